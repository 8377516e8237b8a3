import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "raht-3dgs-codec_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, helpers
from raht_b200 import codec, ops
dev = torch.device("cuda", 0)
n, J, STEP = 1000000, 12, 1 / 64
scene = helpers.synthetic_gaussians(n, 2001)
host = [torch.from_numpy(a).pin_memory() for a in scene]
devt = [h.to(dev) for h in host]
pre = [[torch.empty_like(d) for d in devt] for _ in range(2)]

def timeit(f, k=10):
    f(2); torch.cuda.synchronize()
    t0 = time.perf_counter(); f(k); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / k * 1e3

def resident(k):
    for _ in range(k):
        enc = codec.encode_scene(*devt, J=J, step=STEP)
        codec.decode_scene(enc)
print("resident", timeit(resident))

s_in = torch.cuda.Stream()
def v1(k):
    for i in range(k):
        with torch.cuda.stream(s_in):
            for d, h in zip(pre[i & 1], host): d.copy_(h, non_blocking=True)
        enc = codec.encode_scene(*devt, J=J, step=STEP)
        codec.decode_scene(enc)
print("V1 side H2D into preallocated", timeit(v1))

def v2(k):
    with torch.cuda.stream(s_in):
        for i in range(k):
            for d, h in zip(pre[i & 1], host): d.copy_(h, non_blocking=True)
    for i in range(k):
        enc = codec.encode_scene(*devt, J=J, step=STEP)
        codec.decode_scene(enc)
print("V2 all H2D queued up front", timeit(v2))

big = torch.empty(46 << 20, dtype=torch.uint8, device=dev)
pin = torch.empty(46 << 20, dtype=torch.uint8).pin_memory()
s_out = torch.cuda.Stream()
def v3(k):
    for i in range(k):
        with torch.cuda.stream(s_out):
            pin.copy_(big, non_blocking=True)
        enc = codec.encode_scene(*devt, J=J, step=STEP)
        codec.decode_scene(enc)
print("V3 side D2H 46 MB", timeit(v3))

def mk(result):
    pipe = codec.ScenePipeline(J, STEP, result=result)
    def run(k):
        pipe.submit(host)
        for i in range(k):
            if i + 1 < k: pipe.submit(host)
            pipe.run()
        pipe.finish()
    return run
print("pipe psnr k=30", timeit(mk("psnr"), 30))
