"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the compressed-scene container and of the RLGR sync index.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the product
(raht-3dgs-codec_b200/) never does.

The container has no counterpart in the reference (python/encode_3dgs.py:247 only takes len(buf) of each stream and
the voxel positions travel uncompressed, python/test_voxelize_3dgs.py:284-288): its layout is defined by
raht-3dgs-codec_b200/raht_b200/container.py and restated here independently (numpy + struct + the RLGR checker of
oracle_ctypes.py, which follows python/PyRLGR/src/libs/rlgr/membuf.cpp:340-423).  Parity of this part is therefore
pinned only against this restatement and the committed known-answer container (tests/golden/container_kat.npz):
"parity unpinned" with respect to the reference, by construction.
"""
import struct

import numpy as np

from . import oracle_ctypes as orc

MAGIC, VERSION, N_GEO, DIGIT_BITS = b"R3GC", 1, 3, 21
FLAG_INDEX, FLAG_F32 = 1, 2
FIXED = "<4sHHIIIIIQQ4fdd"
SYNC_SHIFT = 8
SYNC_RANGE = 1 << SYNC_SHIFT


def pad8(x):
    return (int(x) + 7) // 8 * 8


def geometry_symbols(keys):
    """Sorted unique Morton codes -> (3, n) digits of d[0] = key[0], d[i] = key[i] - key[i-1] - 1."""
    k = np.asarray(keys, dtype=np.int64)
    if k.size and (k[0] < 0 or np.any(np.diff(k) <= 0)):
        raise ValueError("keys must be non-negative, sorted and unique")
    d = k.astype(np.uint64).copy()
    if k.size > 1:
        d[1:] = (k[1:] - k[:-1] - 1).astype(np.uint64)
    mask = np.uint64((1 << DIGIT_BITS) - 1)
    return np.stack([((d >> np.uint64(DIGIT_BITS * j)) & mask).astype(np.int32) for j in range(N_GEO)])


def geometry_keys(sym):
    s = np.asarray(sym).astype(np.uint64)
    d = s[0] | (s[1] << np.uint64(DIGIT_BITS)) | (s[2] << np.uint64(2 * DIGIT_BITS))
    if d.size > 1:
        d[1:] += np.uint64(1)
    return np.cumsum(d.astype(np.int64))


def _gr_len(u, kr):                       # membuf.cpp:242-256 (grWrite): p ones, a zero, kr bits -- or the 32-bit escape
    p = u >> kr
    return p + 1 + kr if p < 32 else 64


def rlgr_sync_index(seq, flag_signed=1):
    """(ceil(n / 256), 2) uint32: for every range of 256 symbols the position of the coder (bit offset, k_P, k_RP) before
    the event -- zeros, then one non-zero -- whose non-zero is the first of the range, and that symbol's place in the range
    (include/raht_b200.h, sync_index).  Ranges without a non-zero keep 0xffffffff in both words.  Walks the coder of
    membuf.cpp:351-414 symbol by symbol, counting bits only."""
    n = len(seq)
    idx = np.full(((n + SYNC_RANGE - 1) >> SYNC_SHIFT, 2), 0xffffffff, dtype=np.uint32)
    kP, kRP, m, bits = 0, 8, 0, 0
    ev = (0, 0, 8)
    last = -1
    for i, v in enumerate(seq):
        v = int(v)
        u = (((-v) << 1) - 1 if v < 0 else v << 1) if flag_signed else v
        k, kR = kP >> 2, kRP >> 2
        coded = None
        if k == 0:
            bits += _gr_len(u, kR)
            coded = u
            kP = (max(kP - 1, 0) if u else kP + 3)
            m = 0
        elif u:
            coded = u - 1
            bits += 1 + k + _gr_len(coded, kR)
            kP = max(kP - 1, 0)
            m = 0
        else:
            m += 1
            if m == 1 << k:
                bits += 1
                kP += 2
                m = 0
        if coded is not None:             # adapt k_RP (membuf.cpp:395-407)
            p = coded >> kR
            kRP = min(kRP + p - 1, 128) if p else max(kRP - 2, 0)
        if u:
            r = i >> SYNC_SHIFT
            if r != last:
                b, a, c = ev
                idx[r, 0] = b & 0xffffffff
                idx[r, 1] = ((b >> 32) & 0x7f) | ((a & 0x7f) << 7) | ((c & 0xff) << 14) | ((i & (SYNC_RANGE - 1)) << 22)
                last = r
            ev = (bits, kP, kRP)
    return idx


def _encode_channels(q, flag_signed):
    q = np.ascontiguousarray(q, dtype=np.int32)
    if q.shape[1] == 0:
        return [b"" for _ in range(q.shape[0])]
    streams, _, _, _ = orc.rlgr_encode_channels(q, flag_signed)
    return [bytes(s) for s in streams]


def pack_scene(J, n_in, steps, vmin, voxel_size, width, keys, q, with_index=True, f32=False):
    """Container bytes from the voxel keys (nvox,) and the quantised, reordered coefficients q (D, nvox) int32."""
    keys = np.asarray(keys, dtype=np.int64)
    q = np.asarray(q, dtype=np.int32)
    D, nvox = q.shape
    sym = geometry_symbols(keys)
    geo, attr = _encode_channels(sym, 0), _encode_channels(q, 1)
    entries = ((nvox + SYNC_RANGE - 1) >> SYNC_SHIFT) if with_index else 0
    steps = [float(s) for s in (steps if isinstance(steps, (list, tuple, np.ndarray)) else [steps])]
    var = struct.pack(f"<{len(steps)}d", *steps) + struct.pack(f"<{N_GEO}Q", *[len(s) for s in geo]) + \
        struct.pack(f"<{D}Q", *[len(s) for s in attr]) + struct.pack("<QQ", entries, entries)
    hb = pad8(struct.calcsize(FIXED) + len(var))
    flags = (FLAG_INDEX if with_index else 0) | (FLAG_F32 if f32 else 0)
    head = struct.pack(FIXED, MAGIC, VERSION, flags, hb, J, D, N_GEO, len(steps), n_in, nvox, float(vmin[0]), float(vmin[1]),
                       float(vmin[2]), 0.0, float(voxel_size), float(width)) + var
    out = [head.ljust(hb, b"\0")]
    for s in geo + attr:
        out.append(s.ljust(pad8(len(s)), b"\0"))
    if with_index:
        for c in range(N_GEO):
            out.append(rlgr_sync_index(sym[c], 0).astype("<u4").tobytes())
        for c in range(D):
            out.append(rlgr_sync_index(q[c], 1).astype("<u4").tobytes())
    return b"".join(out)


def unpack_scene(data):
    """-> dict with the header fields, `keys` (nvox,) int64 and `q` (D, nvox) int32 (sequential decode: the index is
    side information)."""
    fixed = struct.calcsize(FIXED)
    magic, version, flags, hb, J, D, n_geo, n_steps, n_in, nvox, vx, vy, vz, _, voxel_size, width = struct.unpack_from(FIXED, data, 0)
    assert magic == MAGIC and version == VERSION and n_geo == N_GEO
    o = fixed
    steps = list(struct.unpack_from(f"<{n_steps}d", data, o)); o += 8 * n_steps
    geo_bytes = struct.unpack_from(f"<{N_GEO}Q", data, o); o += 8 * N_GEO
    attr_bytes = struct.unpack_from(f"<{D}Q", data, o); o += 8 * D
    geo_entries, attr_entries = struct.unpack_from("<QQ", data, o)
    o = hb
    chans = []
    for nb, signed in [(b, 0) for b in geo_bytes] + [(b, 1) for b in attr_bytes]:
        chans.append(np.asarray(orc.rlgr_decode(data[o:o + nb], nvox, signed), dtype=np.int64) if nvox else np.zeros(0, np.int64))
        o += pad8(nb)
    sym = np.stack(chans[:N_GEO]) if nvox else np.zeros((N_GEO, 0), np.int64)
    return dict(J=J, D=D, n_in=n_in, nvox=nvox, vmin=(vx, vy, vz), voxel_size=voxel_size, width=width, steps=steps, flags=flags,
                geo_bytes=list(geo_bytes), attr_bytes=list(attr_bytes), entries=(geo_entries, attr_entries),
                keys=geometry_keys(sym), q=np.stack(chans[N_GEO:]).astype(np.int32) if D else np.zeros((0, nvox), np.int32),
                body_end=o)
