"""TEST / BASELINE INFRASTRUCTURE -- not part of the product path (only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this).

Runs the REFERENCE's own `Pix2PixHDModel` (vendored unmodified into oracle/_ref by oracle/build_ref.py) on the host
cores: the G3L2-ngf48 recipe of generated/GEN_VCTK_G3L2_48ngf/opt.txt, one optimisation step = the body of
train.py:148-184 (forward, loss_G.backward + Adam, loss_D.backward + Adam), fp32, `--gpu_ids -1`.

transform = 'mdct4' (BASELINE configs 3/4): the model's `_mdct` hook is the reference's own MDCT4
(models/mdct.py:456-513) exactly as tests/golden/make_golden_model.py wires it; 'mdct2': the model as shipped."""
import contextlib
import importlib
import io
import os
import sys
import tempfile
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available():
    return os.path.isfile(os.path.join(REF_DIR, "models", "pix2pixHD_model.py"))


def _import_reference():
    import warnings
    warnings.filterwarnings("ignore", category=SyntaxWarning)        # the reference's docstrings hold "\s" escapes
    for name in ("pysepm", "matplotlib", "matplotlib.pyplot", "turtle"):        # the complete stub set
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["turtle"].forward = lambda *a, **k: None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    # our package mirrors the reference's module names one level down (pix2pixhdaudiosr_b200.models ...): top-level
    # `models`, `options`, `util`, `dct` resolve to oracle/_ref only
    for top in ("models", "options", "util", "dct"):
        m = sys.modules.get(top)
        if m is not None and not getattr(m, "__file__", "").startswith(REF_DIR):
            raise RuntimeError(f"a foreign top-level module `{top}` is already imported")


def g3l2_args(batch, transform="mdct4", time_d=False):
    n_fft, hop = (1024, 512) if transform == "mdct4" else (512, 256)
    a = ['--gpu_ids', '-1', '--name', 'ref', '--netG', 'local', '--ngf', '48', '--n_downsample_global', '4',
         '--n_blocks_global', '3', '--n_blocks_local', '2', '--n_local_enhancers', '1', '--ndf', '64', '--num_D', '2',
         '--n_layers_D', '3', '--input_nc', '2', '--output_nc', '2', '--label_nc', '0', '--no_instance', '--no_vgg_loss',
         '--explicit_encoding', '--mask', '--mask_mode', 'mode2', '--n_fft', str(n_fft), '--hop_length', str(hop),
         '--win_length', str(n_fft), '--lr_sampling_rate', '8000', '--hr_sampling_rate', '48000', '--batchSize', str(batch),
         '--segment_length', str(127 * hop), '--center']
    if time_d:
        a += ['--use_time_D', '--use_match_loss', '--lambda_time', '10']
    return a


class ReferenceTrainer:
    """the reference model + the step of train.py:148-184"""

    def __init__(self, batch=8, transform="mdct4", time_d=False, threads=None):
        import torch
        self.torch = torch
        self.threads = threads or os.cpu_count() or 1
        torch.set_num_threads(self.threads)
        _import_reference()
        tmp = tempfile.mkdtemp()
        argv, sys.argv = sys.argv, ['train.py', '--checkpoints_dir', tmp] + g3l2_args(batch, transform, time_d)
        try:
            TrainOptions = importlib.import_module("options.train_options").TrainOptions
            create_model = importlib.import_module("models.models").create_model
            with contextlib.redirect_stdout(io.StringIO()):
                opt = TrainOptions().parse()
                torch.manual_seed(opt.seed)
                self.model = create_model(opt)
        finally:
            sys.argv = argv
        self.opt = opt
        if transform == "mdct4":
            if time_d:
                raise NotImplementedError("the reference wires its time-domain path to MDCT2 only (SURVEY H1); use transform='mdct2'")
            MDCT4 = importlib.import_module("models.mdct").MDCT4
            ref = MDCT4(n_fft=opt.n_fft, hop_length=opt.hop_length, win_length=opt.win_length, window=self.model.window, device='cpu')

            class F32(torch.nn.Module):
                def __init__(self, m):
                    super().__init__()
                    self.m = m

                def forward(self, x):
                    return self.m(x.double() if x.dtype != torch.float32 or x.dim() == 3 else x).float()

            self.model._mdct = F32(ref)

    def step(self, lr, hr):
        """train.py:148-184 without AMP: returns the loss dict"""
        torch, model = self.torch, self.model
        losses, _ = model(lr, None, hr, None, infer=False)
        losses = [torch.mean(x) if not isinstance(x, int) else x for x in losses]
        d = dict(zip(model.loss_names, losses))
        loss_D = (d['D_fake'] + d['D_real']) * 0.5 + (d.get('D_fake_t', 0) + d.get('D_real_t', 0)) * 0.5
        loss_G = d['G_GAN'] + d.get('G_GAN_Feat', 0) + d.get('G_VGG', 0) + d.get('G_mat', 0) + d.get('G_GAN_t', 0)
        model.optimizer_G.zero_grad()
        loss_G.backward()
        model.optimizer_G.step()
        model.optimizer_D.zero_grad()
        loss_D.backward()
        model.optimizer_D.step()
        return d

    def time_steps(self, lr, hr, steps, warmup=0, budget_s=None):
        """-> list of per-step seconds (at most `steps`, stopping early once budget_s is spent; at least one)"""
        for _ in range(warmup):
            self.step(lr, hr)
        out, t_begin = [], time.perf_counter()
        for _ in range(steps):
            t0 = time.perf_counter()
            self.step(lr, hr)
            out.append(time.perf_counter() - t0)
            if budget_s is not None and time.perf_counter() - t_begin > budget_s:
                break
        return out
