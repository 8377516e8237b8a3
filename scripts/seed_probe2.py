"""RLGR encoder statistics of the 1 M bench scene per generator seed: where the data-dependent time goes."""
import sys, os
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "raht-3dgs-codec_b200"))
from raht_b200 import codec, synth, ops

seeds = [int(x) for x in sys.argv[1:]] or [2001, 2003, 2004]
for seed in seeds:
    scene = [torch.from_numpy(a).cuda() for a in synth.synthetic_gaussians(1_000_000, seed)]
    enc = codec.encode_scene(*scene, J=12, step=1.0 / 64)
    q = ops.rlgr_decode(enc.streams, 1).contiguous()
    st = ops.rlgr_encode(q, 1, want_stats=True)
    print(seed, st.stats)
    zf = (q == 0).float().mean(1).cpu().numpy()
    # per-channel encode time, one channel at a time
    times = []
    for c in range(q.shape[0]):
        qc = q[c:c + 1].contiguous()
        ops.rlgr_encode(qc, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); s1 = ops.rlgr_encode(qc, 1, want_stats=True); e1.record(); torch.cuda.synchronize()
        times.append((round(e0.elapsed_time(e1), 3), c, round(float(zf[c]), 4), s1.stats["repairs"], s1.stats["table_hits"], s1.stats["slowest_chunk"]["rounds"]))
    times.sort(reverse=True)
    print("  slowest channels (ms, channel, zero fraction, repairs, table hits, rounds):", times[:6])
