import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "raht-3dgs-codec_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
import helpers
from test_sharded import OracleBackend
from raht_b200 import sharded
for (n, J, B) in [(3000, 8, 3), (24000, 8, 6), (200000, 12, 8)]:
    P_, MC = helpers.unique_sorted_voxels(n, J, 5)
    keys = torch.from_numpy(MC)
    C = np.random.default_rng(1).normal(size=(len(MC), 7)).astype(np.float32)
    ob, db = OracleBackend(), sharded.DeviceBackend()
    ob.J = J
    cut = 3 * J - B
    lo = ob.levels(keys, cut); ld = db.levels(keys.cuda(), cut)
    print(n, J, B, "levels equal", np.array_equal(lo.numpy(), ld.cpu().numpy()), "roots", int((lo == 255).sum()))
    To = ob.forward_local(keys, lo, torch.from_numpy(C), torch.float64)
    Td = db.forward_local(keys.cuda(), ld, torch.from_numpy(C).cuda(), torch.float64)
    bad = np.flatnonzero((To.numpy() != Td.cpu().numpy()).any(1))
    print("  forward equal", len(bad) == 0, "bad rows", bad[:10], "levels of bad", lo.numpy()[bad[:10]])
    Co = ob.inverse_local(keys, lo, To)
    Cd = db.inverse_local(keys.cuda(), ld, Td)
    bad = np.flatnonzero((Co.numpy() != Cd.cpu().numpy()).any(1))
    print("  inverse equal", len(bad) == 0, "bad rows", bad[:10])
