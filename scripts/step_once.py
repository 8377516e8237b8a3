#!/usr/bin/env python
"""Two encode + decode steps of the bench scene and nothing else: the command to put under ncu for a launch list
(bench.py adds the end-to-end, container and CPU legs, which only lengthen a profiled run).
usage: python scripts/step_once.py [n] [J] [--fused]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "raht-3dgs-codec_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from raht_b200 import synth as helpers  # noqa: E402
from raht_b200 import codec  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n = int(args[0]) if args else 1_000_000
J = int(args[1]) if len(args) > 1 else 12
fused = "--fused" in sys.argv
scene = [torch.from_numpy(a).cuda() for a in helpers.synthetic_gaussians(n, 2001)]
for _ in range(2):
    enc = codec.encode_scene(*scene, J=J, step=1 / 64, fused=fused)
    coords, rec = codec.decode_scene(enc, fused=fused)
torch.cuda.synchronize()
print("nvox", enc.nvox, "payload", enc.payload_bytes())
