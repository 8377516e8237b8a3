"""Compressed-scene container (raht_b200/container.py) against its CPU restatement
(oracle/container_oracle.py) and the RLGR sync index against a symbol-by-symbol walk of the reference coder.

CPU (not gpu): the restatement round-trips, its geometry digits are exact at the widest Morton codes, and a committed
known-answer container pins the byte layout.
GPU: the container built on the device is byte-identical to the restatement's, decodes to the same scene as the
EncodedScene it came from, and the encoder's sync index equals the walk.
"""
import hashlib
import os

import numpy as np
import pytest

import helpers
from oracle import oracle_ctypes as orc
from oracle import container_oracle as cor

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "container_kat.npz")


def _toy(seed=5, nvox=1500, D=4, J=7):
    rng = np.random.default_rng(seed)
    keys = np.sort(rng.choice(1 << (3 * J), size=nvox, replace=False)).astype(np.int64)
    q = np.zeros((D, nvox), dtype=np.int32)
    q[0] = rng.integers(-40, 40, nvox)                                   # dense
    q[1] = rng.integers(-3, 4, nvox) * (rng.random(nvox) < 0.2)          # sparse
    q[2, :7] = [100000, -100000, 2147483647, -2147483647, 0, 1, -1]      # escapes
    q[3] = rng.integers(0, 2, nvox) * (rng.random(nvox) < 0.02)          # very sparse, one-signed
    return keys, q


def test_geometry_digits_roundtrip_widest_codes():
    keys = np.array([0, 1, 2, (1 << 21), (1 << 42) + 5, (1 << 62), (1 << 63) - 1], dtype=np.int64)
    sym = cor.geometry_symbols(keys)
    assert sym.shape == (3, 7) and sym.min() >= 0 and sym.max() < (1 << 21)
    assert list(sym[:, 0]) == [0, 0, 0] and list(sym[:, 1]) == [0, 0, 0] and list(sym[:, 3]) == [(1 << 21) - 3, 0, 0]
    assert np.array_equal(cor.geometry_keys(sym), keys)
    with pytest.raises(ValueError):
        cor.geometry_symbols(np.array([3, 3], dtype=np.int64))


def test_sync_index_walk_on_known_streams():
    # all zeros: no entry; a lone non-zero: the event starts at the stream start with the initial state (k_P 0, k_RP 8)
    assert np.all(cor.rlgr_sync_index(np.zeros(1000, np.int64)) == 0xffffffff)
    idx = cor.rlgr_sync_index(np.array([0, 0, 0, 5] + [0] * 600 + [1], np.int64))
    assert idx.shape == (3, 2)                                        # ranges of 256 symbols
    assert idx[0, 0] == 0 and idx[0, 1] == (0 << 7 | 8 << 14 | 3 << 22)
    assert np.all(idx[1] == 0xffffffff)                               # nothing but zeros in [256, 512)
    # third range: its first non-zero is symbol 604 (place 92); the event starts right after the code of the 5
    bits_first = len(orc.rlgr_encode(np.array([0, 0, 0, 5], np.int64))) * 8
    assert idx[2, 1] >> 22 == 604 - 512 and 0 < idx[2, 0] <= bits_first


def test_container_restatement_roundtrip():
    keys, q = _toy()
    for with_index in (True, False):
        data = cor.pack_scene(7, 4321, 0.25, (-1.0, 0.5, 2.0), 0.015625, 2.0, keys, q, with_index)
        assert len(data) % 8 == 0
        h = cor.unpack_scene(data)
        assert h["J"] == 7 and h["n_in"] == 4321 and h["nvox"] == len(keys) and h["steps"] == [0.25]
        assert h["vmin"] == (-1.0, 0.5, 2.0) and h["voxel_size"] == 0.015625 and h["width"] == 2.0
        assert np.array_equal(h["keys"], keys) and np.array_equal(h["q"], q)
        n_index = (3 + 4) * ((len(keys) + 255) // 256) * 8 if with_index else 0
        assert h["body_end"] + n_index == len(data)
    # per-channel steps
    data = cor.pack_scene(7, 10, [1.0, 2.0, 4.0, 8.0], (0, 0, 0), 1.0, 128.0, keys, q, False)
    assert cor.unpack_scene(data)["steps"] == [1.0, 2.0, 4.0, 8.0]
    # empty scene
    e = cor.unpack_scene(cor.pack_scene(3, 0, 1.0, (0, 0, 0), 1.0, 8.0, np.zeros(0, np.int64), np.zeros((2, 0), np.int32), True))
    assert e["nvox"] == 0 and e["keys"].size == 0


def test_container_known_answer():
    """tests/golden/container_kat.npz (made by tests/golden/make_golden.py from this restatement): pins the byte layout."""
    g = np.load(GOLDEN)
    data = cor.pack_scene(int(g["J"]), int(g["n_in"]), float(g["step"]), tuple(g["vmin"]), float(g["voxel_size"]), float(g["width"]),
                          g["keys"], g["q"], True)
    assert hashlib.sha256(data).hexdigest() == str(g["sha256"])
    assert data[:int(g["head"].size)] == g["head"].tobytes()


# ---------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_geometry_symbols_and_keys_cuda():
    import torch
    from raht_b200 import container
    rng = np.random.default_rng(11)
    for n, bits in ((1, 20), (2049, 36), (100000, 48), (5000, 63)):
        keys = np.unique(rng.integers(0, (1 << bits) - 1, size=n, dtype=np.int64))
        if bits == 63:
            keys[-1] = (1 << 63) - 1
        kd = torch.from_numpy(keys).cuda()
        sym = container.geometry_symbols(kd)
        assert np.array_equal(sym.cpu().numpy(), cor.geometry_symbols(keys))
        assert torch.equal(container.geometry_keys(sym), kd)
    from raht_b200._lib import RahtError
    with pytest.raises((RahtError, ValueError, RuntimeError)):
        container.geometry_symbols(torch.tensor([5, 5, 9], dtype=torch.int64).cuda())


@pytest.mark.gpu
def test_sync_index_matches_the_coder_walk_cuda():
    import torch
    from raht_b200 import ops
    keys, q = _toy(seed=9, nvox=6000, D=4)
    for signed, x in ((1, q), (0, np.abs(q) % (1 << 21))):
        s = ops.rlgr_encode(torch.from_numpy(np.ascontiguousarray(x, dtype=np.int32)).cuda(), signed)
        got = s.sync_index.cpu().numpy().view(np.uint32).reshape(x.shape[0], -1, 2)
        for c in range(x.shape[0]):
            assert np.array_equal(got[c], cor.rlgr_sync_index(x[c], signed)), f"channel {c} signed {signed}"


@pytest.mark.gpu
@pytest.mark.parametrize("with_index", [True, False])
def test_container_bytes_match_restatement_and_decode_cuda(with_index, tmp_path):
    import torch
    from raht_b200 import codec, container, ops
    J, step, n = 9, 0.05, 30000
    scene = [torch.from_numpy(a).cuda() for a in helpers.synthetic_gaussians(n, 31)]
    enc = codec.encode_scene(*scene, J=J, step=step)
    buf = container.pack(enc, with_index)
    data = buf.cpu().numpy().tobytes()
    q = ops.rlgr_decode(enc.streams, 1).cpu().numpy()                     # (the RLGR path has its own parity tests)
    want = cor.pack_scene(J, n, step, enc.vmin.cpu().tolist(), enc.voxel_size, enc.width, enc.keys.cpu().numpy(), q, with_index)
    assert len(data) == len(want)
    assert data == want
    # restatement reads the device-built container
    h = cor.unpack_scene(data)
    assert np.array_equal(h["keys"], enc.keys.cpu().numpy()) and np.array_equal(h["q"], q)
    # device reader: from the device tensor, from host bytes, from a file
    _, ref = codec.decode_scene(enc)
    path = os.path.join(tmp_path, "scene.r3gc")
    assert container.write_scene(path, enc, with_index) == len(data)
    for src in (buf, data, None):
        dec = container.read_scene(path) if src is None else container.unpack(src)
        assert (dec.J, dec.n_in, dec.nvox, dec.D, dec.step) == (enc.J, enc.n_in, enc.nvox, enc.D, enc.step)
        assert dec.voxel_size == enc.voxel_size and dec.width == enc.width and torch.equal(dec.vmin, enc.vmin)
        assert torch.equal(dec.keys, enc.keys)
        assert (dec.streams.sync_index is not None) == with_index
        coords, rec = codec.decode_scene(dec)
        assert torch.equal(rec, ref)
    # geometry costs less than the raw keys
    geo_bytes = sum(container.parse_header(data)["geo_bytes"])
    assert geo_bytes < enc.nvox * 8 / 2


@pytest.mark.gpu
def test_container_rejects_damage_cuda():
    import torch
    from raht_b200 import codec, container
    scene = [torch.from_numpy(a).cuda() for a in helpers.synthetic_gaussians(2000, 3)]
    data = container.pack(codec.encode_scene(*scene, J=6, step=0.1)).cpu().numpy().tobytes()
    with pytest.raises(ValueError):
        container.unpack(b"XXXX" + data[4:])
    with pytest.raises(ValueError):
        container.unpack(data[:len(data) // 2])
    with pytest.raises(ValueError):
        container.unpack(data[:40])
    # a header that lies about its sizes (advisor, round 1): every field the kernels index with is checked against the buffer
    import struct
    def patched(off, fmt, value):
        b = bytearray(data)
        struct.pack_into(fmt, b, off, value)
        return bytes(b)
    hb = struct.unpack_from("<I", data, 8)[0]
    n_steps = struct.unpack_from("<I", data, 24)[0]
    var = 76 + 8 * n_steps                       # geo_bytes[3], attr_bytes[D], geo_entries, attr_entries follow the steps
    D = struct.unpack_from("<I", data, 16)[0]
    for bad in (patched(12, "<I", 40),                                   # J out of range
                patched(36, "<Q", 1 << 50),                              # nvox absurd
                patched(var + 8 * (3 + D), "<Q", 1),                     # geo_entries too small for nvox: the decoder would read past the index
                patched(var + 8 * (3 + D) + 8, "<Q", 10 ** 9),           # attr_entries too large
                patched(var, "<Q", 1 << 40)):                            # a stream longer than the file
        with pytest.raises(ValueError):
            container.unpack(bad)
