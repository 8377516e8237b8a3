"""ms per encode + decode step of the 1 M bench scene for several generator seeds (the replica ranks use seed + rank)."""
import sys, os
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "raht-3dgs-codec_b200"))
from raht_b200 import codec, synth

for seed in (2001, 2002, 2003, 2004, 2005, 2008):
    scene = [torch.from_numpy(a).cuda() for a in synth.synthetic_gaussians(1_000_000, seed)]
    def one(tm=None):
        enc = codec.encode_scene(*scene, J=12, step=1.0 / 64, timer=tm)
        codec.decode_scene(enc, reuse_tree=False, timer=tm)
        return enc
    for _ in range(3):
        enc = one()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        one()
    e1.record()
    torch.cuda.synchronize()
    tm = codec.StageTimer(True)
    enc = one(tm)
    torch.cuda.synchronize()
    print(seed, f"{e0.elapsed_time(e1) / 10:.3f} ms", enc.nvox, enc.payload_bytes(), {k: round(v, 3) for k, v in tm.result().items() if v > 0.3})
