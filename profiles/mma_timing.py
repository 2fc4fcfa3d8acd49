"""where the MMA-issuing warp of conv_window_kernel spends its cycles (instrumented build: P2P_LIB_PATH=scratch/libp2p_timing.so)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pix2pixhdaudiosr_b200 import ops, _lib
lib = _lib.lib()
import ctypes

dbg = torch.zeros(148 * 17, dtype=torch.int64, device="cuda")
lib.p2p_conv_window_debug_buffer(ctypes.c_void_p(dbg.data_ptr()))
os.environ["P2P_CONV_SWAP"] = "0"
CASES = [("res96 fwd", False, 96, 96, 3, 1, 1, True, (8, 256, 64)), ("d256to512 fwd", False, 256, 512, 4, 1, 2, False, (8, 65, 17)),
         ("d64to128_s2 fwd", False, 64, 128, 4, 2, 2, False, (8, 257, 65)), ("1x1 c112 j48", False, 112, 48, 1, 1, 0, False, (8, 256, 128))]
for name, tr, Cin, Cout, k, stride, pad, reflect, (N, H, W) in CASES:
    x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
    w = (torch.randn(Cout, Cin, k, k, device="cuda") * 0.05).permute(2, 3, 0, 1).contiguous().permute(2, 3, 0, 1)
    w16 = w.permute(2, 3, 0, 1).contiguous().bfloat16()
    b = torch.zeros(Cout, device="cuda")
    cfg = ops.ConvCfg(tr, stride, pad, ops.PAD_REFLECT if reflect else ops.PAD_ZERO, 0)
    for _ in range(3):
        dbg.zero_()
        ops.conv_forward(x, w, b, w16, cfg, torch.float32)
    torch.cuda.synchronize()
    ops.start_conv_timing()
    for _ in range(5):
        ops.conv_forward(x, w, b, w16, cfg, torch.float32)
    rec = ops.stop_conv_timing()
    print("    event-timed kernel: %.1f us (mean of 5)" % (1000 * sum(r[3] for r in rec) / len(rec)))
    life = dbg[148 * 8:148 * 9].double()
    raw7 = dbg[:148 * 8].view(148, 8)[:, 7]
    pre_ns, loop_ns = (raw7 >> 32).double(), (raw7 & 0xffffffff).double()
    print("    ns per CTA: lifetime mean %.0f max %.0f | entry->MMA loop start %.0f | MMA loop %.0f mean / %.0f max" % (
        life.mean().item(), life.max().item(), pre_ns.mean().item(), loop_ns.mean().item(), loop_ns.max().item()))
    d = dbg[:148 * 8].view(148, 8).double()
    d2 = dbg[148 * 9:].view(148, 8).double()
    d2 = d2[d[:, 5] > 0]
    d = d[d[:, 5] > 0]
    th, ch, sh, ct, tt = [d2[:, i].mean().item() for i in range(5)]
    print("    overhead buckets per CTA: tile hdr %.0f | chunk hdr %.0f | stage hdr %.0f | chunk tail %.0f | tile tail %.0f" % (th, ch, sh, ct, tt))
    wa, wb, iss, wt, n, tot, tiles = [d[:, i].mean().item() for i in range(7)]
    print("%-16s CTAs %3d tiles/CTA %.1f stages/CTA %.0f | cycles per CTA: total %.0f = wait window %.0f + wait weights %.0f + issue %.0f + wait tmem %.0f + other %.0f | per stage: issue %.0f wait %.0f"
          % (name, d.shape[0], tiles, n, tot, wa, wb, iss, wt, tot - wa - wb - iss - wt, iss / max(n, 1), wb / max(n, 1)))
