"""CPU: host-side pieces of the drop-in API that need no GPU."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_splitmix_and_threshold():
    from imagegeneration_b200.sndcgan import _splitmix64, rate_to_thresh8
    assert _splitmix64(0) == 0xE220A8397B1DCDAF        # splitmix64 reference value for seed 0
    assert rate_to_thresh8(0.5) == 128 and rate_to_thresh8(0.25) == 64


def test_engine_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from imagegeneration_b200.sndcgan import SNDCGANEngine
    with pytest.raises(Exception):
        SNDCGANEngine(2, image_size=(16, 32, 3))          # no CPU fallback: constructing on a CPU-only box fails


def test_trainer_cli_flags_match_the_reference():
    """sndcgan/Trainer.py:12-31: same positional arguments, flags, dests and defaults."""
    from imagegeneration_b200.trainer import build_parser
    p = build_parser()
    a = p.parse_args(["16", "3"])
    assert (a.bSize, a.epochs, a.ckptFreq, a.dirPath, a.data, a.dropout, a.learnRateDisc, a.learnRateGen, a.liveOutput,
            a.continue_) == (16, 3, 5, "training", "dataset", 0.5, 0.0002, 0.0002, "live", False)
    a = p.parse_args(["8", "1", "-cf", "2", "-d", "out", "-x", "imgs", "-r", "0.3", "-ld", "1e-4", "-lg", "3e-4", "-lo", "pv", "-ct"])
    assert (a.ckptFreq, a.dirPath, a.data, a.dropout, a.learnRateDisc, a.learnRateGen, a.liveOutput, a.continue_) == \
        (2, "out", "imgs", 0.3, 1e-4, 3e-4, "pv", True)


def test_bench_reference_arm_prints_one_json_line():
    import json
    env = dict(os.environ, B2G_REF_MAX_BATCH="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "images/s" and d["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] == "port"
