"""Runs the spectral-norm power iteration a few times on the full-size discriminator (for ncu / timing)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagegeneration_b200.sndcgan import Discriminator

D = Discriminator(0.0, (144, 256, 3), spectral_norm=True)
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
ts = []
for _ in range(6):
    flush.fill_(1.0)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); D.sn_power_iteration(); b.record()
    torch.cuda.synchronize()
    ts.append(a.elapsed_time(b) * 1e3)
# back-to-back (L2-hot) launches: per-launch time without the event overhead
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50):
    D.sn_power_iteration()
b.record(); torch.cuda.synchronize()
print("cold us", [round(t, 1) for t in ts], "hot us/launch", round(a.elapsed_time(b) * 1e3 / 50, 2), "sigma", [float(s[1]) for s in D.sn_scal])
