"""Per-layer timing of the convolution kernels at the G3L2-ngf48 / S128 shapes (batch 8): CUDA events around each
launch, algorithmic FLOPs / time.  `python profiles/layer_bench.py [--reps 20] [--only name]`.  Also the target of
the `ncu --set full` captures committed in this directory (first launches = the layers in LAYERS order)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pix2pixhdaudiosr_b200 import ops

LAYERS = [  # name, transposed, Cin, Cout, k, stride, pad, reflect, (N, H, W), output feeds InstanceNorm (fp32 out)
    ("g_res96_3x3", False, 96, 96, 3, 1, 1, True, (8, 256, 64), True),
    ("g_bottleneck1536_3x3", False, 1536, 1536, 3, 1, 1, True, (8, 16, 4), True),
    ("d_256to512_4x4s1", False, 256, 512, 4, 1, 2, False, (8, 65, 17), True),
    ("g_down_96to192_s2", False, 96, 192, 3, 2, 1, False, (8, 256, 64), True),
    ("g_up_192to96_convT", True, 192, 96, 3, 2, 1, False, (8, 128, 32), True),
    ("d_64to128_4x4s2", False, 64, 128, 4, 2, 2, False, (8, 257, 65), True),
    ("g_first_7x7_2to48", False, 2, 48, 7, 1, 3, True, (8, 512, 128), True),
    ("g_last_7x7_48to2", False, 48, 2, 7, 1, 3, True, (8, 512, 128), False),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    torch.manual_seed(0)
    print("%-24s %-6s %9s %9s %8s" % ("layer", "pass", "GFLOP", "ms", "TFLOP/s"))
    for name, tr, Cin, Cout, k, stride, pad, reflect, (N, H, W), norm in LAYERS:
        if args.only and name not in args.only.split(","):
            continue
        x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
        shape = (Cin, Cout, k, k) if tr else (Cout, Cin, k, k)
        perm = (2, 3, 1, 0) if tr else (2, 3, 0, 1)
        inv = [perm.index(i) for i in range(4)]
        w = (torch.randn(shape, device="cuda") * 0.02).permute(*perm).contiguous().permute(*inv)
        w16 = w.permute(*perm).contiguous().bfloat16()
        b = torch.zeros(Cout, device="cuda")
        cfg = ops.ConvCfg(tr, stride, pad, ops.PAD_REFLECT if reflect else ops.PAD_ZERO, 1 if tr else 0)
        odt = torch.float32 if norm else torch.bfloat16
        y = ops.conv_forward(x, w, b, w16, cfg, odt)
        dy = torch.randn(y.shape, device="cuda").bfloat16()
        for what in ("fwd", "dgrad", "wgrad"):
            def run():
                if what == "fwd":
                    ops.conv_forward(x, w, b, w16, cfg, odt)
                else:
                    ops.conv_backward(x, w, w16, dy, cfg, what == "dgrad", what == "wgrad", False)
            run()
            torch.cuda.synchronize()
            torch.cuda._sleep(int(0.004 * args.reps * 1.9e9))   # the host enqueues all reps behind a device spin: the event pairs
            ops.start_conv_timing()                             # then bracket kernels that run back to back (no launch gaps)
            for _ in range(args.reps):
                run()
            rec = ops.stop_conv_timing()
            ms = sum(r[3] for r in rec) / args.reps
            fl = sum(r[2] for r in rec) / args.reps
            real = 2.0 * N * y.shape[1] * y.shape[2] * Cout * Cin * k * k if not tr else 2.0 * N * H * W * Cout * Cin * k * k
            print("%-24s %-6s %9.2f %9.4f %8.1f   (launched %.2f GFLOP incl. channel padding)" %
                  (name, what, real / 1e9, ms, real / ms / 1e9, fl / 1e9))


if __name__ == "__main__":
    main()
