"""Timings of the spectral-norm pieces (run on the GPU box): plain Adam on the D arena, b2g_sn_dot_multi, b2g_adam_sn,
the stand-alone power iteration, each alone with the L2 flushed by reading."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import lib as L
from imagegeneration_b200.sndcgan import SNDCGANEngine, _p, _stream

eng = SNDCGANEngine(4, dropout=0.0, spectral_norm=True)
D, opt = eng.disc_model, eng.d_optimizer
A, sn = D.arena, D._sn_args
lib = L.load()
D.sn_power_iteration()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def run(name, fn, reps=7):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    print(f"{name:44s} {ts[len(ts) // 2]:8.1f} us  (min {ts[0]:.1f})", flush=True)


run("adam (D arena, 172 MB)", lambda: opt.apply_arena(A, 0.0))
run("sn_dot_multi", lambda: L.check(lib.b2g_sn_dot_multi(_p(A.w), _p(A.g), 8, D._sn_off, sn[5], sn[6], _p(D._sn_dots), _stream()), "dot"))
D._sn_dots.zero_()
run("adam_sn", lambda: L.check(lib.b2g_adam_sn(_p(A.w), _p(A.g), _p(A.m), _p(A.v), A.numel, 8, D._sn_off, sn[5], sn[6], sn[1], sn[2],
                                               sn[3], sn[4], _p(D._sn_dots), 1e-4, None, 0.9, 0.999, 1e-7, 0.0, _stream()), "adam_sn"))
run("spectral_norm_multi (stand-alone)", D.sn_power_iteration)
run("D.repack", D.repack)
print("sigma", [round(float(s[1]), 4) for s in D.sn_scal])
