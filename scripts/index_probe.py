"""Time of raht_rlgr_build_index (the serial walk of index-less streams) per density regime, one channel at a time."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "raht-3dgs-codec_b200"))
from raht_b200 import ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
rng = np.random.default_rng(5)
for scale, z in ((4.0, 0.0), (50.0, 0.0), (0.7, 0.0), (4.0, 0.5), (4.0, 0.9), (4.0, 0.99)):
    x = np.rint(rng.laplace(0, scale, size=(1, n))).astype(np.int32)
    x[0, rng.random(n) < z] = 0
    xd = torch.from_numpy(x).cuda()
    st = ops.rlgr_encode(xd, 1, with_index=False)
    ops.rlgr_build_index(st)
    t0, t1, t2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t0.record(); idx = ops.rlgr_build_index(st); t1.record()
    got_serial = ops.rlgr_decode(st, 1, use_index=False, serial=True); t2.record()
    torch.cuda.synchronize()
    st.sync_index = idx
    ok = torch.equal(ops.rlgr_decode(st, 1), xd) and torch.equal(got_serial, xd)
    nb = int(st.nbytes.sum())
    print(f"scale {scale:5.1f} zeros {z:4.2f}: {nb * 8 / n:5.2f} bits/sym  walk {t0.elapsed_time(t1):7.2f} ms ({t0.elapsed_time(t1) * 1e6 / n * 1.9:5.0f} clk/sym)  serial decode {t1.elapsed_time(t2):7.2f} ms  ok={ok}")
