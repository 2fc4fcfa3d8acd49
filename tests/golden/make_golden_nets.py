"""Golden vectors for the conv-GAN rows from the REFERENCE's own networks.py, run on CPU in
the build container.  python tests/golden/make_golden_nets.py  ->  tests/golden/nets_ref.npz"""
import contextlib
import importlib
import io
import os

import numpy as np
import torch

from make_golden import import_reference

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    import_reference()
    R = importlib.import_module("models.networks")
    out = {}

    def quiet(f, *a, **k):
        with contextlib.redirect_stdout(io.StringIO()):
            return f(*a, **k)

    def dump_net(tag, net, x, outs_fn):
        """forward, then backward of sum(out_i * r_i) w.r.t. input and every parameter"""
        x = x.clone().requires_grad_(True)
        outs = outs_fn(net(x))
        g = torch.Generator().manual_seed(99)
        loss = 0
        for i, o in enumerate(outs):
            r = torch.randn(o.shape, generator=g)
            out[f"{tag}/out{i}"] = o.detach().numpy()
            out[f"{tag}/r{i}"] = r.numpy()
            loss = loss + (o * r).sum()
        loss.backward()
        out[f"{tag}/x"] = x.detach().numpy()
        out[f"{tag}/gx"] = x.grad.numpy()
        for k, v in net.state_dict().items():
            out[f"{tag}/w/{k}"] = v.numpy()
        for k, p in net.named_parameters():
            out[f"{tag}/g/{k}"] = p.grad.numpy()
        print(tag, [tuple(o.shape) for o in outs], "params", sum(p.numel() for p in net.parameters()))

    torch.manual_seed(1234)
    # G3L2-shaped LocalEnhancer in miniature: define_G(2, 2, ngf, 'local', n_down_global, n_blocks_global, 1, n_blocks_local)
    G = quiet(R.define_G, 2, 2, 4, 'local', 2, 2, 1, 2, 'instance', [])
    dump_net("G_local", G, torch.rand(2, 2, 32, 16), lambda o: [o])
    # BASELINE config 2 in miniature: GlobalGenerator(1, 1, ngf, n_down, n_blocks)
    Gg = quiet(R.define_G, 1, 1, 8, 'global', 2, 3, 1, 3, 'instance', [])
    dump_net("G_global", Gg, torch.rand(2, 1, 16, 24), lambda o: [o])
    # 2-scale discriminator with intermediate features (define_D(4, ndf, 3, 'instance', False, 2, True))
    D = quiet(R.define_D, 4, 8, 3, 'instance', False, 2, True, [])
    dump_net("D_feat", D, torch.rand(2, 4, 32, 16), lambda o: [t for scale in o for t in scale])
    # time-domain discriminator flavour: no intermediate features, odd sizes
    D2 = quiet(R.define_D, 2, 8, 2, 'instance', False, 2, False, [])
    dump_net("D_plain", D2, torch.rand(1, 2, 19, 33), lambda o: [t for scale in o for t in scale])

    # losses: GANLoss (LSGAN) and feature matching (pix2pixHD_model.py:390-398)
    crit = R.GANLoss(use_lsgan=True, tensor=torch.FloatTensor)
    g = torch.Generator().manual_seed(5)
    fake = [[torch.randn(2, c, h, w, generator=g, requires_grad=True) for c, h, w in ((8, 17, 9), (16, 9, 5), (1, 10, 6))]
            for _ in range(2)]
    real = [[torch.randn_like(t) for t in s] for s in fake]
    l_real, l_fake = crit(fake, True), crit(fake, False)
    feat = 0
    for i in range(2):
        for j in range(len(fake[i]) - 1):
            feat = feat + (1.0 / 2) * (4.0 / (3 + 1)) * torch.nn.L1Loss()(fake[i][j], real[i][j].detach()) * 10.0
    (0.5 * l_real + 0.25 * l_fake + feat).backward()
    out["loss/real"], out["loss/fake"], out["loss/feat"] = l_real.detach().numpy(), l_fake.detach().numpy(), feat.detach().numpy()
    for i in range(2):
        for j in range(3):
            out[f"loss/fake{i}{j}"] = fake[i][j].detach().numpy()
            out[f"loss/real{i}{j}"] = real[i][j].numpy()
            out[f"loss/g{i}{j}"] = fake[i][j].grad.numpy()
    np.savez_compressed(os.path.join(HERE, "nets_ref.npz"), **out)
    print("saved", os.path.getsize(os.path.join(HERE, "nets_ref.npz")), "bytes")


if __name__ == "__main__":
    main()
