"""Experiment: self-synchronising speculative parse of RLGR streams by bit segments (host model)."""
import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from oracle import oracle_ctypes as orc
L, U0, D0, U1, D1 = 4, 3, 1, 2, 1

def parse(bits, nbits, pos, kP, kRP, stop_bit, max_steps=1 << 30):
    """steps from bit `pos` in state until pos >= stop_bit; returns (pos, kP, kRP, symbols consumed)."""
    nsym = 0
    while pos < stop_bit and pos < nbits:
        k = kP // L; kR = kRP // L
        run = k != 0
        if run:
            m = 0
            while pos < nbits and bits[pos]:
                pos += 1; m += 1 << k; kP += U1; k = kP // L
                if k > 30: return pos, kP, kRP, nsym   # garbage
            pos += 1
            v = 0
            for _ in range(k):
                v = (v << 1) | (bits[pos] if pos < nbits else 0); pos += 1
            m += v
            nsym += m
        p = 0
        while pos < nbits and bits[pos] and p < 32:
            p += 1; pos += 1
        if p >= 32:
            u = 0
            for _ in range(32):
                u = (u << 1) | (bits[pos] if pos < nbits else 0); pos += 1
        else:
            pos += 1
            v = 0
            for _ in range(kR):
                v = (v << 1) | (bits[pos] if pos < nbits else 0); pos += 1
            u = (p << kR) + v
        pp = u >> kR
        if pp: kRP = min(kRP + pp - 1, 128)
        else: kRP = max(kRP - 2, 0)
        if run or u:
            kP = max(kP - D1, 0)
        else:
            kP += U0
        nsym += 1
    return pos, kP, kRP, nsym

def experiment(x, SEG):
    stream = orc.rlgr_encode_channels(x[None, :].astype(np.int32), 1)[0][0]
    bits = np.unpackbits(np.frombuffer(stream, dtype=np.uint8)).tolist()
    nbits = len(bits)
    nseg = (nbits + SEG - 1) // SEG
    # truth
    truth = []
    pos, kP, kRP = 0, 0, 8
    for i in range(nseg):
        pos, kP, kRP, _ = parse(bits, nbits, pos, kP, kRP, (i + 1) * SEG)
        truth.append((pos, kP, kRP))
    start = [(i * SEG, 0, 8) for i in range(nseg)]
    cross = [None] * nseg
    dirty = [True] * nseg
    it = 0
    work = 0
    while any(dirty):
        it += 1
        newc = list(cross)
        for i in range(nseg):
            if dirty[i]:
                r = parse(bits, nbits, *start[i], (i + 1) * SEG)
                newc[i] = r[:3]; work += 1
        nd = [False] * nseg
        for i in range(1, nseg):
            if newc[i - 1] != start[i] and (dirty[i - 1]):
                if newc[i - 1] != start[i]:
                    start[i] = newc[i - 1]; nd[i] = True
        cross = newc; dirty = nd
        good = sum(1 for i in range(nseg) if cross[i] == truth[i])
        print(f"  iter {it}: {sum(nd)} segments to redo, {good}/{nseg} crossings true")
        if it > 60: break
    print(f" nbits {nbits} nseg {nseg} iterations {it} segment-parses {work} ({work / nseg:.2f} per segment)")

rng = np.random.default_rng(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
SEG = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
for scale, z in ((4.0, 0.0), (4.0, 0.5), (4.0, 0.9), (0.7, 0.0), (50.0, 0.0), (4.0, 0.99)):
    x = np.rint(rng.laplace(0, scale, n)).astype(np.int64)
    x[rng.random(n) < z] = 0
    print(f"scale {scale} zero-frac {z}")
    experiment(x, SEG)
