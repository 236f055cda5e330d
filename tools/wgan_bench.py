#!/usr/bin/env python
"""WGAN 256x144 training throughput on one B200 (BASELINE.json configs[3]).

An "iteration" is the body of WGAN.train between two generator updates (wasserstein_gan/WGAN.py:296-318):
critic_learn_iterations (5) x [critic.train_on_batch(real, -1); generator.predict(z); critic.train_on_batch(fake, +1)]
followed by one gan_model.train_on_batch(z, -1).  Value = real images consumed per second (5 * batch per iteration).
Also times the new HBM-bound kernels alone (training-mode BatchNorm forward / backward on the largest activation,
RMSprop + clip on the critic arena) against the measured copy bandwidth.

    python tools/wgan_bench.py [--steps K] [--warmup W] [--batch B]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

IMAGE_SIZE = (144, 256, 3)
G_FWD, D_FWD = 7.4507e9, 15.5295e9          # algorithmic FLOP per image, forward


def flop_per_iteration(batch, n_critic=5):
    critic_step = 3 * D_FWD - 0.1274e9          # fwd + dgrad + wgrad, no conv1 input gradient
    gan_step = 3 * G_FWD - 0.0755e9 + 2 * D_FWD  # G fwd+bwd, critic fwd + input gradients only
    return batch * (n_critic * (2 * critic_step + G_FWD) + gan_step)


def _p(t):
    return C.c_void_p(t.data_ptr())


def cpu_baseline(batch=4):
    """The oracle's restatement of the same loop on the host cores (kind "port": TensorFlow cannot run here): one
    critic update on real images, one predict + critic update on fakes, one generator update at `batch` images, scaled
    to an iteration (5 x the critic pair + 1 generator update).  Images/s = real images per second, as in our arm."""
    import time

    import torch
    from oracle import wgan_oracle as W
    cores = torch.get_num_threads()
    orc = W.OracleWGAN(IMAGE_SIZE, seed=62)
    g = torch.Generator().manual_seed(3)
    real = torch.rand((batch, *IMAGE_SIZE), generator=g) * 2 - 1
    z = torch.randn((batch, 128), generator=g)
    t0 = time.perf_counter()
    orc.train_critic_on_batch(real, -1.0)
    orc.train_critic_on_batch(orc.predict(z), 1.0)
    t_pair = time.perf_counter() - t0
    t0 = time.perf_counter()
    orc.train_gan_on_batch(z)
    t_gan = time.perf_counter() - t0
    t_iter = 5 * t_pair + t_gan
    return {"value": 5 * batch / t_iter, "unit": "images/s", "cores": cores, "kind": "port",
            "sample": f"oracle (PyTorch-CPU restatement of wasserstein_gan/WGAN.py:296-318, fp32) at batch {batch}: one "
                      f"critic real+fake pair {t_pair:.2f} s, one generator update {t_gan:.2f} s, scaled to 5 pairs + 1 "
                      f"update; {cores} threads"}


def run(steps=5, warmup=2, batch=64, n_critic=5, with_cpu_baseline=False):
    import torch
    from imagegeneration_b200 import lib as L
    from imagegeneration_b200.wgan import WGANEngine
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0}
    torch.cuda.set_device(0)
    eng = WGANEngine(IMAGE_SIZE, batch, n_critic, seed=62)
    g = torch.Generator(device="cuda").manual_seed(1)
    reals = [torch.rand((batch, *IMAGE_SIZE), generator=g, device="cuda") * 2 - 1 for _ in range(n_critic)]
    zs = [torch.randn((batch, 128), generator=g, device="cuda") for _ in range(n_critic + 1)]
    for _ in range(warmup):
        out = eng.train_iteration(reals, zs[:n_critic], zs[n_critic])
    torch.cuda.synchronize()
    l0 = L.launch_count
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        out = eng.train_iteration(reals, zs[:n_critic], zs[n_critic])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    launches = L.launch_count - l0
    line = {"metric": "WGAN 256x144 train images/sec (real images; 5 critic iterations per generator step)",
            "value": n_critic * batch / (ms * 1e-3), "unit": "images/s", "n_gpus": 1, "steps": steps, "warmup": warmup,
            "ms_per_iteration": ms, "higher_is_better": True, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "WGAN.train iteration at 256x144: 10 critic RMSprop updates (clip 0.01), 5 generator "
                                   "predicts, 1 generator update; training-mode BatchNorm", "batch_per_gpu": batch,
                       "critic_learn_iterations": n_critic},
            "gpu_launches": launches,
            "tflops_whole_iteration": flop_per_iteration(batch, n_critic) / (ms * 1e-3) / 1e12,
            "frac_of_sustained_bf16_peak": flop_per_iteration(batch, n_critic) / (ms * 1e-3) / 1e12
            / float(peaks.get("bf16_tflops_sustained", 1400.0)),
            "losses_last_iteration": [float(x) for x in out]}
    # ---- HBM-bound kernels alone.  L2 is flushed between repetitions by READING a 512 MB buffer: a write flush
    # leaves ~126 MB of dirty lines whose write-back shares HBM with the first ~100 us of the timed kernel and
    # would dominate a 300 MB read-only kernel.
    lib = L.load()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    flush = torch.zeros(128 * 1024 * 1024, dtype=torch.float32, device="cuda")
    flush_sink = torch.zeros((), dtype=torch.float32, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def timed(fn, reps=5):
        ts = []
        for _ in range(reps):
            flush_sink.copy_(flush.sum())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e-3)
        ts.sort()
        return ts[len(ts) // 2]

    P, Cn = batch * 144 * 256, 64
    x = torch.randn((P, Cn), device="cuda").to(torch.bfloat16)
    dy = (torch.randn((P, Cn), device="cuda") * 0.01).to(torch.bfloat16)
    y, dx = torch.empty_like(x), torch.empty_like(x)
    sums, stat, red = (torch.zeros(16 * Cn, device="cuda") for _ in range(3))
    gamma, beta = torch.ones(Cn, device="cuda"), torch.zeros(Cn, device="cuda")
    mm, mv = torch.zeros(Cn, device="cuda"), torch.ones(Cn, device="cuda")
    n = P * Cn
    hb = {}

    def entry(sec, byt, note):
        return {"bound": "hbm", "achieved": byt / sec / 1e9, "peak": peak, "unit": "GB/s", "frac": byt / sec / 1e9 / peak,
                "bytes_per_launch": byt, "us": sec * 1e6, "note": note}

    sec = timed(lambda: L.check(lib.b2g_bn_stats(_p(x), _p(sums), P, Cn, s), "stats"))
    hb["bn_stats"] = entry(sec, 2.0 * n, "critic conv1 output, 64 x 144 x 256 x 64 bf16: one read")
    L.check(lib.b2g_bn_finalize(_p(sums), P, Cn, 1e-3, 0.99, _p(stat), _p(mm), _p(mv), None, s), "fin")
    sec = timed(lambda: L.check(lib.b2g_bn_apply_act(_p(x), _p(stat), _p(gamma), _p(beta), 0.2, _p(y), P, Cn, s), "apply"))
    hb["bn_apply_act"] = entry(sec, 4.0 * n, "read + write")
    sec = timed(lambda: L.check(lib.b2g_bn_bwd(_p(dy), _p(x), _p(stat), _p(gamma), _p(red), _p(dx), None, None, P, Cn, s), "bwd", 2))
    hb["bn_bwd"] = entry(sec, 10.0 * n, "two launches: reduce reads g, x (4 B/elt); apply reads g, x, writes dx (6 B/elt)")
    a_ = eng.critic_model.arena
    sec = timed(lambda: L.check(lib.b2g_rmsprop(_p(a_.w), _p(a_.g), _p(a_.m), a_.numel, 0.0, 0.9, 1e-7, 1.0,
                                                eng.critic_model.clip_n, 0.01, s), "rmsprop"))
    hb["rmsprop_critic_arena"] = entry(sec, 20.0 * a_.numel, "20 B/param (read w, g, rms; write w, rms); 6.2 M params: "
                                       "latency-bound at 123 MB")
    line["roofline_hbm"] = hb
    if with_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline()
    return line


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--cpu-baseline", action="store_true")
    args = ap.parse_args()
    print(json.dumps(run(args.steps, args.warmup, args.batch, with_cpu_baseline=args.cpu_baseline)), flush=True)
