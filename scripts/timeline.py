"""Kernel timeline of one encode + decode step (torch.profiler / CUPTI): start, duration and the idle gap in front of every
kernel -- where the step's time goes BETWEEN kernels.  usage: python scripts/timeline.py [n] > gpurun_out/timeline.txt"""
import sys, os
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "raht-3dgs-codec_b200"))
from raht_b200 import codec, synth
from torch.profiler import profile, ProfilerActivity

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
scene = [torch.from_numpy(a).cuda() for a in synth.synthetic_gaussians(n, 2001)]
pending = []
def one():
    mine = []
    enc = codec.encode_scene(*scene, J=12, step=1.0 / 64, checks=mine)
    r = codec.decode_scene(enc, reuse_tree=False, checks=mine)
    for c in pending:
        c()
    pending[:] = mine
    return r
for _ in range(3):
    one()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        one()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
# keep the middle step: from the second bounds_init to the third
starts = [i for i, e in enumerate(evs) if "bounds_init" in e.name]
lo, hi = (starts[1], starts[2]) if len(starts) >= 3 else (0, len(evs))
prev_end = None
tot_k = tot_gap = 0.0
for e in evs[lo:hi]:
    s, d = e.time_range.start, e.time_range.end - e.time_range.start
    gap = (s - prev_end) if prev_end is not None else 0.0
    prev_end = max(prev_end or 0, e.time_range.end)
    tot_k += d
    tot_gap += max(gap, 0)
    print(f"{s - evs[lo].time_range.start:10.1f} us  dur {d:8.1f}  gap {gap:7.1f}  {e.name[:90]}")
print(f"# kernels {tot_k:.1f} us, gaps {tot_gap:.1f} us, step {tot_k + tot_gap:.1f} us")
