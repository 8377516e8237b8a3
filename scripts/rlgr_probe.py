#!/usr/bin/env python
"""GPU probe: times raht_rlgr_encode / decode alone on the quantised coefficients of the bench scene and prints the
speculation statistics (chunk solves, fixed-point rounds, repairs).  usage: python scripts/rlgr_probe.py [n] [step]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "raht-3dgs-codec_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from raht_b200 import synth as helpers  # noqa: E402
from raht_b200 import codec, ops  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
step = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0 / 64
J = int(os.environ.get("RLGR_PROBE_J", "12"))
dev = "cuda:0"
if n > 8_000_000:                                       # (numpy takes 3.5 s per million Gaussians)
    scene = list(helpers.synthetic_gaussians_device(n, 4001, dev))
else:
    scene = [torch.from_numpy(a).to(dev) for a in helpers.synthetic_gaussians(n, 2001)]
packed, keys, info = codec.voxelize_merge(*scene, J=J)
tree = ops.tree_from_keys(keys)
order = tree.order()
T = ops.forward(tree, packed, torch.float64)
q = ops.quant_reorder(T, order, step)
del T, packed, scene
torch.cuda.empty_cache()
if os.environ.get("RLGR_PROBE_SEL"):                     # e.g. "0:1": only these channels
    a, b = os.environ["RLGR_PROBE_SEL"].split(":")
    q = q[int(a):int(b)].contiguous()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        r = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, r


ms_enc, st = timed(lambda: ops.rlgr_encode(q, 1, want_stats=True))
ms_dec, _ = timed(lambda: ops.rlgr_decode(st, 1))
ms_ser = float("nan")
if not os.environ.get("RLGR_PROBE_NO_SERIAL"):
    ms_ser, _ = timed(lambda: ops.rlgr_decode(st, 1, use_index=False), reps=1)
zeros = (q == 0).float().mean(dim=1)
print(f"n={q.shape[1]} step={step} sweeps={os.environ.get('RAHT_RLGR_SWEEPS', '2')} encode {ms_enc:.3f} ms (incl. host sync) "
      f"decode indexed {ms_dec:.3f} ms serial {ms_ser:.1f} ms payload {st.total_bytes()} index {st.index_bytes()} stats {st.stats} "
      f"zero fraction min/median/max {zeros.min():.3f}/{zeros.median():.3f}/{zeros.max():.3f}")
if os.environ.get("RLGR_PROBE_SPLIT"):
    for name, sel in (("ch0-2", slice(0, 3)), ("ch3-6", slice(3, 7)), ("ch7-9", slice(7, 10)), ("ch14-58", slice(14, 59)), ("ch0", slice(0, 1)),
                      ("ch3", slice(3, 4))):
        qs = q[sel].contiguous()
        ms, st = timed(lambda: ops.rlgr_encode(qs, 1, want_stats=True))
        print(f"  {name}: encode {ms:.3f} ms stats {st.stats}")
