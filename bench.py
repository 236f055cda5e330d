#!/usr/bin/env python
"""bench.py — SNDCGAN 256x144 training throughput (BASELINE.json metric) on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W            our arm (sm_100a kernels)
  python bench.py --impl reference --steps K --warmup W     the reference's CPU path (PyTorch-CPU restatement,
                                                            TensorFlow is not installable here) on the host cores
For N > 1 launch with `python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...`.

A "step" = one SNDCGAN.train_step (1 G update + 2 D updates, sndcgan/SNDCGAN.py:241-269) on a synthetic batch of
64 images per GPU (BASELINE config[1]; weak scaling: N GPUs -> global batch 64*N, = config[2]'s 512 at N=8).
Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

IMAGE_SIZE = (144, 256, 3)
METRIC = "SNDCGAN 256x144 train images/sec"
FLOP_PER_IMAGE = 146.26e9   # one train_step, algorithmic


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in out.strip().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle restatement of the reference's train_step on the host cores
# ----------------------------------------------------------------------------------------------------------
def _oracle_step_factory(batch):
    import torch
    from oracle import sndcgan_oracle as O
    orc = O.OracleSNDCGAN(IMAGE_SIZE, dropout=0.5, seed=62)
    g = torch.Generator().manual_seed(1234)
    real = torch.rand((batch, *IMAGE_SIZE), generator=g) * 2 - 1
    z = torch.rand((batch, 128), generator=g) * 2 - 1
    shapes = [(144, 256, 64), (72, 128, 128), (72, 128, 128), (36, 64, 256), (36, 64, 256), (18, 32, 512), (18, 32, 512)]

    def masks():
        return [(torch.rand((batch, *s), generator=g) >= 0.5).float() for s in shapes]

    def step():
        return orc.train_step(real, z, masks(), masks(), masks())
    return step


REF_BATCH = 16   # BASELINE configs[0]: the reference's own CPU-runnable case; never shrunk


def _host_threads():
    import torch
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm uses every host core it may run on
    try:
        torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    except (AttributeError, OSError):
        torch.set_num_threads(max(1, os.cpu_count() or 1))
    return torch.get_num_threads()


def _timed_oracle_steps(steps, warmup, budget_s):
    """The oracle's train_step at batch 16 fp32 on the host cores.  The first step is always untimed (oneDNN primitive
    creation, first-touch page faults); if `steps` would overrun `budget_s`, the STEP COUNT is reduced, never the
    batch.  Returns (images/s, timed steps, seconds, threads, untimed steps)."""
    cores = _host_threads()
    step = _oracle_step_factory(REF_BATCH)
    t_begin = time.perf_counter()
    step()
    t1 = None
    for _ in range(max(1, warmup - 1)):
        t0 = time.perf_counter()
        step()
        t1 = time.perf_counter() - t0
    untimed = 1 + max(1, warmup - 1)
    left = budget_s - (time.perf_counter() - t_begin)
    n = max(1, min(steps, int(left / t1)))
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    dt = time.perf_counter() - t0
    return REF_BATCH * n / dt, n, dt, cores, untimed


def _cpu_sample_text(n, dt, cores, untimed):
    return (f"{n} oracle train_step(s) (PyTorch-CPU restatement of SNDCGAN.py:241-269, fp32, dropout 0.5) at batch "
            f"{REF_BATCH} (BASELINE configs[0]) after {untimed} untimed warm-up step(s), {dt:.2f} s, {cores} host threads; "
            "TensorFlow itself is not installable here")


def cpu_baseline(budget_s=25.0):
    """`cpu_baseline` leg of our arm's line: the same function and the same fixed batch as `--impl reference`."""
    value, n, dt, cores, untimed = _timed_oracle_steps(3, 2, budget_s)
    return {"value": value, "unit": "images/s", "cores": cores, "kind": "port",
            "sample": _cpu_sample_text(n, dt, cores, untimed), "batch_per_step": REF_BATCH, "steps": n}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    value, n, dt, cores, untimed = _timed_oracle_steps(args.steps, args.warmup, float(os.environ.get("B2G_REF_BUDGET_S", "240")))
    sample = _cpu_sample_text(n, dt, cores, untimed)
    if n != args.steps:
        sample += f" (asked for {args.steps} steps; the count, not the batch, was cut to stay inside the time budget)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": args.gpus,
            "steps": n, "warmup": untimed, "ms_per_step": dt / n * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "SNDCGAN one G+D train_step at 256x144, fp32, CPU (PyTorch restatement of the "
                                   "reference; TensorFlow unavailable)", "batch_per_step": REF_BATCH,
                       "steps_requested": args.steps},
            "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------------
def _hbm_kernel_rooflines(eng, peaks, peak_kind):
    """Achieved HBM GB/s of the memory-bound parameter kernels, each timed alone with CUDA events on torch's current
    stream (the stream they are launched on), L2 flushed between repetitions by a 512 MB fill.
      adam (G arena): 28 B per parameter (read w, g, m, v; write w, m, v)            SURVEY.md section 8(d)
      spectral norm: the stand-alone power iteration (4 B per weight, W read once) on the 8 D matrices and on a
      294 912 x 128 matrix, and the fused optimizer pass that carries it inside a training step"""
    import torch
    from imagegeneration_b200.sndcgan import SNDCGANEngine
    peak = float(peaks.get("hbm_gbs", 6650.0))
    flush = torch.empty(128 * 1024 * 1024, dtype=torch.float32, device="cuda")
    out = {}

    def timed(fn, reps=5):
        # L2 is flushed by READING 512 MB: after a write flush the 126 MB of dirty lines drain into the first ~20 us
        # of the timed kernel, which is most of a 30 us launch (measured in round 1 on the BatchNorm kernels)
        ts = []
        flush.fill_(1.0)
        for _ in range(reps):
            flush.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e-3)
        ts.sort()
        return ts[len(ts) // 2]

    G = eng.gen_model.arena
    it0 = eng.g_optimizer.iterations
    g_saved = G.g.clone()
    G.g.zero_()                       # zero gradients with m = v as left by training: weights barely move
    sec = timed(lambda: eng.g_optimizer.apply_arena(G, 0.0))
    eng.g_optimizer.iterations = it0
    G.g.copy_(g_saved)
    byt = 28.0 * G.numel
    out["adam_generator_arena"] = {"bound": "hbm", "achieved": byt / sec / 1e9, "peak": peak, "unit": "GB/s",
                                   "frac": byt / sec / 1e9 / peak, "bytes_per_launch": byt, "us": sec * 1e6,
                                   "peak_kind": f"{peak_kind} copy bandwidth"}
    try:
        import ctypes as C
        from imagegeneration_b200 import lib as _L
        lib_ = _L.load()
        sn = SNDCGANEngine(4, dropout=0.0, image_size=IMAGE_SIZE, seed=62, spectral_norm=True)
        D = sn.disc_model
        n_w = float(sum(k * c for k, c in zip(D._sn_K, D._sn_C)))
        # (a) stand-alone power iteration over the 8 discriminator matrices: only runs at init / set_weights / restore
        sec = timed(D.sn_power_iteration)
        byt = 4.0 * n_w
        out["spectral_norm_multi"] = {"bound": "hbm", "achieved": byt / sec / 1e9, "peak": peak, "unit": "GB/s",
                                      "frac": byt / sec / 1e9 / peak, "bytes_per_launch": byt, "us": sec * 1e6,
                                      "note": "stand-alone iteration, 8 matrices = 24.6 MB (3.8 us at the copy peak): "
                                              "latency-bound at this size and OFF the training step since round 2 — "
                                              "training updates carry their power iteration inside b2g_adam_sn",
                                      "peak_kind": f"{peak_kind} copy bandwidth"}
        # (b) what spectral norm costs inside a step now: the fused optimizer pass against plain Adam on the same arena
        A = D.arena
        sec_plain = timed(lambda: sn.d_optimizer.apply_arena(A, 0.0))
        D.sn_power_iteration()

        def fused():
            sn.d_optimizer.apply_model(D, 0.0)
        sec_fused = timed(fused)
        out["adam_sn_discriminator_arena"] = {
            "bound": "hbm", "achieved": 28.0 * A.numel / sec_fused / 1e9, "peak": peak, "unit": "GB/s",
            "frac": 28.0 * A.numel / sec_fused / 1e9 / peak, "bytes_per_launch": 28.0 * A.numel,
            "us": sec_fused * 1e6, "us_plain_adam": sec_plain * 1e6, "extra_us_for_spectral_norm": (sec_fused - sec_plain) * 1e6,
            "extra_bytes": 4.0 * n_w + 8.0 * sum(D._sn_K),
            "note": "b2g_sn_dot_multi (<G,W>: +4 B/weight for W; G is re-read by Adam) + b2g_adam_sn (Adam + gradient "
                    "through W/sigma + next power iteration in one pass); two updates per step",
            "peak_kind": f"{peak_kind} copy bandwidth"}
        del sn
        # (c) SURVEY 7.3's large-matrix case: one power iteration on a 294 912 x 128 fp32 matrix (151 MB, the
        # generator's Dense kernel shape), W read exactly once
        K_, C_ = 294912, 128
        Wl = torch.randn(K_, C_, device="cuda") * 0.02
        ul = torch.nn.functional.normalize(torch.randn(C_, device="cuda"), dim=0)
        vr, tr, sc = torch.zeros(K_, device="cuda"), torch.zeros(C_, device="cuda"), torch.zeros(8, device="cuda")
        arr = lambda t: (C.c_void_p * 1)(t.data_ptr())
        Ks, Cs = (C.c_int * 1)(K_), (C.c_int * 1)(C_)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

        def big():
            rc = lib_.b2g_spectral_norm_multi(1, arr(Wl), arr(ul), arr(vr), arr(tr), arr(sc), Ks, Cs, st)
            if rc:
                raise RuntimeError("b2g_spectral_norm_multi failed")
        sec = timed(big)
        byt = 4.0 * K_ * C_
        out["spectral_norm_large_matrix"] = {"bound": "hbm", "achieved": byt / sec / 1e9, "peak": peak, "unit": "GB/s",
                                             "frac": byt / sec / 1e9 / peak, "bytes_per_launch": byt, "us": sec * 1e6,
                                             "shape": [K_, C_], "peak_kind": f"{peak_kind} copy bandwidth"}
    except Exception as e:  # the option is not on the reference path; never fail the bench line for it
        out["spectral_norm_multi"] = {"error": str(e)[:200]}
    return out


def _parity_check(out, batch, world, args):
    """The first warm-up step of THIS process against the committed oracle outputs for the same seeds
    (tests/golden/bench_step0.json, made by tests/golden/make_bench_golden.py from oracle/sndcgan_oracle.py): the timed
    configuration is parity-checked in the run that times it.  North-star tolerance: 2e-2 relative on the losses, 2e-2
    of the logit scale on the logits.  With several ranks only the quantities computed before the first (all-reduced)
    discriminator update are comparable with the single-process oracle: g_loss, d_loss_real, d_logits_real."""
    path = os.path.join(ROOT, "tests", "golden", "bench_step0.json")
    key = f"batch{batch}"
    if args.loss != "bce" or args.spectral_norm or not os.path.exists(path):
        return {"checked": False, "why": "no committed oracle values for this configuration"}
    with open(path) as f:
        gold = json.load(f)
    if key not in gold:
        return {"checked": False, "why": f"no committed oracle values for batch {batch}"}
    gold = gold[key]
    import torch
    losses = [float(x) for x in out[:4]]
    idx = (0, 1, 2, 3) if world == 1 else (0, 2)
    rel = max(abs(losses[i] - gold["losses"][i]) / (abs(gold["losses"][i]) + 1e-12) for i in idx)
    pairs = [(out[4], gold["d_logits_real"])] + ([(out[5], gold["d_logits_fake"])] if world == 1 else [])
    lerr = 0.0
    for a, b in pairs:
        b = torch.tensor(b, dtype=torch.float32)
        lerr = max(lerr, float((a.float().cpu().flatten() - b).abs().max()) / (1.0 + float(b.abs().max())))
    res = {"checked": True, "against": "tests/golden/bench_step0.json (CPU oracle, fp32)", "step": 0,
           "losses": losses, "oracle_losses": gold["losses"], "max_rel_loss_err": rel, "max_logit_err": lerr,
           "tolerance": 2e-2, "ok": bool(rel < 2e-2 and lerr < 2e-2)}
    if not res["ok"] and not args.no_parity_check:
        raise SystemExit(f"bench.py: the timed configuration FAILED its parity check against the oracle: {res}")
    return res


def _teardown(eng, dist):
    """Drops the captured step graph (it references NCCL kernels), then tears the process group down.  A watchdog ends
    the process if the teardown does not return: the measurement is complete and printed by then."""
    import threading
    sys.stdout.flush()
    t = threading.Timer(30.0, lambda: os._exit(0))
    t.daemon = True
    t.start()
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    t.cancel()


def run_b200(args):
    import torch
    import torch.distributed as dist

    from imagegeneration_b200 import lib
    from imagegeneration_b200.sndcgan import DeviceFeeder, SNDCGANEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    pg = None
    if world > 1:
        # data-parallel steps are replayed as one CUDA graph too (all-reduces captured); _teardown releases it
        os.environ.setdefault("B2G_CUDA_GRAPH_DP", "1")
        # NCCL prints its version banner on STDOUT when the first communicator comes up; the contract is ONE JSON line
        # on stdout, so stdout points at stderr until the communicator exists
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            pg = dist.group.WORLD
            warm = torch.zeros(1, device="cuda")
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    B = args.batch
    scaling = "weak"
    if args.global_batch:
        # BASELINE configs[2]: a FIXED global batch (512) split over the ranks = strong scaling
        if args.global_batch % world:
            raise SystemExit(f"--global-batch {args.global_batch} is not divisible by {world} ranks")
        B = args.global_batch // world
        scaling = "strong"
    eng = SNDCGANEngine(B, dropout=0.5, image_size=IMAGE_SIZE, seed=62, spectral_norm=args.spectral_norm,
                        loss=args.loss, process_group=pg, **({"dtype": args.dtype} if args.dtype != "bf16" else {}))
    g = torch.Generator().manual_seed(1234 + rank)
    real_host = [(torch.rand((B, *IMAGE_SIZE), generator=g) * 2 - 1).pin_memory() for _ in range(2)]
    z_host = [(torch.rand((B, 128), generator=g) * 2 - 1).pin_memory() for _ in range(2)]
    real_dev = [t.cuda() for t in real_host]
    z_dev = [t.cuda() for t in z_host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing: `value`
    parity = None
    for i in range(args.warmup):
        o = eng.train_step(real_dev[i & 1], z_dev[i & 1])
        if i == 0 and rank == 0:
            parity = _parity_check(o, B, world, args)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = lib.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        out = eng.train_step(real_dev[i & 1], z_dev[i & 1])
    e1.record()
    barrier()
    launches = lib.launch_count - l0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if sampler else None
    losses = [float(x) for x in out[:4]]
    ms_per_step = ms_total / args.steps
    value = B * world * args.steps / (ms_total * 1e-3)

    # ---- end-to-end timing through the public API with HOST buffers: `e2e`
    feeder = DeviceFeeder(B, IMAGE_SIZE)
    feeder.put(real_host[0], z_host[0])
    for i in range(2):  # warm the copy path
        feeder.put(real_host[(i + 1) & 1], z_host[(i + 1) & 1])
        r, z, slot = feeder.get()
        o = eng.train_step(r, z)
        feeder.release(slot)
        torch.stack(o[:4]).cpu()
    barrier()
    d2h = 4 * 4 + 2 * B * 4
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(args.steps):
        feeder.put(real_host[(i + 1) & 1], z_host[(i + 1) & 1])   # next batch: H2D overlaps this step
        r, z, slot = feeder.get()
        o = eng.train_step(r, z)
        feeder.release(slot)
        _l = torch.stack(o[:4]).cpu()                               # D2H of the step's losses + logits
        _a, _b = o[4].cpu(), o[5].cpu()
    feeder.get()  # drain the last prefetched batch
    t1.record()
    barrier()
    e2e_ms = max_over_ranks(t0.elapsed_time(t1))
    e2e_value = B * world * args.steps / (e2e_ms * 1e-3)

    # ---- roofline of the dominant kernel family: one instrumented step (events around every tensor-core launch)
    lib.profile_sink = []
    eng.train_step(real_dev[0], z_dev[0])
    torch.cuda.synchronize()
    agg = {}
    for name, flops, a, b, _label in lib.profile_sink:
        d = agg.setdefault(name, [0.0, 0.0, 0])
        d[0] += flops
        d[1] += a.elapsed_time(b) * 1e-3
        d[2] += 1
    lib.profile_sink = None
    peaks, peak_kind = _peaks()
    peak_tf = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
    peak_name = "sustained bf16"
    if args.dtype == "tf32":
        # MEASURED_PEAKS.json holds no tf32 figure: measured here the same way the driver measures bf16 (torch.matmul
        # 8192^3, tf32 enabled, back to back for ~2 s after warm-up)
        torch.backends.cuda.matmul.allow_tf32 = True
        a = torch.randn(8192, 8192, device="cuda")
        b = torch.randn(8192, 8192, device="cuda")
        for _ in range(3):
            a @ b
        torch.cuda.synchronize()
        n, t0 = 0, time.perf_counter()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        while time.perf_counter() - t0 < 2.0:
            for _ in range(10):
                a @ b
            n += 10
            torch.cuda.synchronize()
        ev1.record()
        torch.cuda.synchronize()
        peak_tf = 2.0 * 8192 ** 3 * n / (ev0.elapsed_time(ev1) * 1e-3) / 1e12
        peak_kind, peak_name = "measured in this run", "sustained tf32 (torch.matmul 8192^3, allow_tf32)"
        del a, b
    roof = {}
    for name, (fl, sec, cnt) in agg.items():
        roof[name] = {"bound": "tensor", "achieved": fl / sec / 1e12, "peak": peak_tf, "unit": "TFLOP/s",
                      "frac": fl / sec / 1e12 / peak_tf, "traffic": None, "launches_per_step": cnt,
                      "ms_per_step": sec * 1e3, "peak_kind": f"{peak_kind} {peak_name}"}
    # DRAM bytes per launch of the same kernel families from one ncu capture of one step at the round's HEAD
    # (profiles/r2_conv_traffic.json names the build and the command); None when the file is absent or the dtype differs
    try:
        with open(os.path.join(ROOT, "profiles", "r2_conv_traffic.json")) as f:
            traffic = json.load(f)
        for name, v in roof.items():
            if name in traffic and args.dtype == "bf16":
                v["traffic"] = traffic[name]["dram_bytes_per_launch"]
                v["traffic_source"] = "profiles/r2_conv_traffic.json: " + traffic.get("source", "")
    except (OSError, ValueError, KeyError):
        pass
    tc_ms = sum(v["ms_per_step"] for v in roof.values())
    hbm_roof = _hbm_kernel_rooflines(eng, peaks, peak_kind) if rank == 0 else None

    if rank != 0:
        if world > 1:
            _teardown(eng, dist)
        return
    line = {
        "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": "SNDCGAN one G+D train_step at 256x144 (1 G + 2 D Adam updates, 3 D forwards), "
                               + ("bf16 activations / fp32 accumulate + master weights" if args.dtype == "bf16" else
                                  "fp32 activations, kind::tf32 tensor-core products (parity path, BASELINE.md C2')")
                               + ", dropout 0.5",
                   "batch_per_gpu": B, "global_batch": B * world, "parallelism": f"dp{world}",
                   "loss": args.loss, "spectral_norm": bool(args.spectral_norm),
                   "l2": "per-step working set (~6 GB of activations) is much larger than the 126 MB L2"},
        "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": feeder.bytes_per_batch,
                "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roof.get("conv_igemm"),
        "roofline_wgrad": roof.get("conv_wgrad"),
        "roofline_hbm": hbm_roof,
        "step_model": {"tflops_per_gpu_whole_step": FLOP_PER_IMAGE * B / (ms_per_step * 1e-3) / 1e12,
                       "frac_of_peak_whole_step": FLOP_PER_IMAGE * B / (ms_per_step * 1e-3) / 1e12 / peak_tf,
                       "tensor_core_kernels_ms": tc_ms, "other_ms": ms_per_step - tc_ms},
        "losses_last_step": losses,
        "parity_check": parity,
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline()
    print(json.dumps(line), flush=True)
    if world > 1:
        _teardown(eng, dist)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--loss", default="bce", choices=["bce", "hinge"])
    ap.add_argument("--spectral-norm", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true",
                    help="report a failed step-0 parity check in the JSON line instead of aborting")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="fixed GLOBAL batch split over the ranks (strong scaling; BASELINE configs[2] uses 512)")
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "tf32"],
                    help="bf16 (BASELINE configs[1]) or the fp32-storage / kind::tf32 parity path")
    ap.add_argument("--workload", default="sndcgan", choices=["sndcgan", "wgan"],
                    help="sndcgan (default): the BASELINE metric; wgan: BASELINE configs[3] on one GPU (SURVEY 8(f) N1), "
                         "a step = one WGAN.train iteration (5 critic iterations + 1 generator update)")
    args = ap.parse_args()
    if args.workload == "wgan":
        if args.impl != "b200" or args.gpus != 1:
            raise SystemExit("--workload wgan measures our arm on one GPU (tools/wgan_bench.py)")
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import wgan_bench
        print(json.dumps(wgan_bench.run(max(1, args.steps // 4), max(1, args.warmup), args.batch,
                                        with_cpu_baseline=not args.no_cpu_baseline)), flush=True)
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
