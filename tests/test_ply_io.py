"""Scene I/O in the reference's PLY dialect (raht_b200/ply_io.py, csrc/ply.cu) against the reference's own
writer (python/quality_eval.py:18-117 save_ply) and reader (python/data_util.py:272-382 read_compressed_3dgs_ply).

CPU (not gpu): the header bytes equal the reference writer's, when the reference checkout is present.
GPU: files written by the product are read by a numpy restatement of the reference reader's dtype (and by the reference
reader itself when it is present) with every value intact; files in the reference writer's layout are read by the product."""
import os
import sys
import warnings

import numpy as np
import pytest
import torch

import helpers

REF_PY = "/root/reference/python"


def _load_ref(name):
    """quality_eval / data_util of the reference checkout (build container only), imported under a private name."""
    import importlib.util
    path = os.path.join(REF_PY, name + ".py")
    if not os.path.isfile(path):
        return None
    spec = importlib.util.spec_from_file_location("refpy_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _record_dtype(cd=48):
    """The vertex record of data_util.py:331-340 (62 floats for 48 colour channels)."""
    return np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("nx", "<f4"), ("ny", "<f4"), ("nz", "<f4")] +
                    [(f"c{i}", "<f4") for i in range(cd)] + [("opacity", "<f4")] + [(f"scale_{i}", "<f4") for i in range(3)] +
                    [(f"rot_{i}", "<f4") for i in range(4)])


def _write_reference_layout(path, scene, voxel_size, vmin):
    """A file exactly as quality_eval.save_ply lays it out, written with numpy (for the box without the reference)."""
    from raht_b200 import ply_io
    means, quats, scales, opac, colors = scene
    n, cd = means.shape[0], colors.shape[1]
    rec = np.zeros(n, dtype=_record_dtype(cd))
    rec["x"], rec["y"], rec["z"] = means[:, 0], means[:, 1], means[:, 2]
    for i in range(cd):
        rec[f"c{i}"] = colors[:, i]
    rec["opacity"] = opac
    for i in range(3):
        rec[f"scale_{i}"] = scales[:, i]
    for i in range(4):
        rec[f"rot_{i}"] = quats[:, i]
    with open(path, "wb") as f:
        f.write(ply_io._header(n, cd, voxel_size, torch.from_numpy(vmin)))
        f.write(rec.tobytes())


def test_header_equals_the_reference_writer(tmp_path):
    qe = _load_ref("quality_eval")
    if qe is None:
        pytest.skip("reference checkout absent")
    from raht_b200 import ply_io
    scene = [torch.from_numpy(a) for a in helpers.synthetic_gaussians(5, 3)]
    vmin = torch.tensor([-1.5, 0.25, 3.0])
    p = tmp_path / "ref.ply"
    qe.save_ply(str(p), *scene, voxel_size=0.00390625, vmin=vmin)
    data = p.read_bytes()
    head = ply_io._header(5, 48, 0.00390625, vmin)
    assert data[:len(head)] == head and len(data) == ply_io.file_size(5, 48, 0.00390625, vmin)
    # and the body is the record layout the kernels build
    rec = np.frombuffer(data[len(head):], dtype=_record_dtype(48))
    assert np.array_equal(rec["x"], scene[0][:, 0].numpy()) and np.array_equal(rec["rot_3"], scene[1][:, 3].numpy())
    assert np.array_equal(rec["c47"], scene[4][:, 47].numpy()) and np.array_equal(rec["opacity"], scene[3].numpy())


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 1000, 300_000])
def test_save_and_read_roundtrip_and_reference_compatibility(n, tmp_path):
    from raht_b200 import ply_io
    sc = helpers.synthetic_gaussians(n, 17)
    J = 9
    # voxelised scene: integer positions stored as floats, like test_voxelize_3dgs.py:284-288
    pos = np.floor((sc[0] - sc[0].min(0)) / (np.ptp(sc[0]) / 2 ** J)).clip(0, 2 ** J - 1).astype(np.float32)
    scene_np = (pos, sc[1], sc[2], sc[3], sc[4])
    scene = [torch.from_numpy(a).cuda() for a in scene_np]
    vmin = torch.tensor([-0.75, 0.5, 1.0])
    p = str(tmp_path / "scene.ply")
    size = ply_io.save_ply(p, *scene, voxel_size=0.001953125, vmin=vmin)
    assert size == os.path.getsize(p) == ply_io.file_size(n, 48, 0.001953125, vmin)
    # (i) the reference reader's dtype sees every value
    head = ply_io._header(n, 48, 0.001953125, vmin)
    rec = np.fromfile(p, dtype=_record_dtype(48), offset=len(head))
    assert rec.shape[0] == n and np.array_equal(rec["y"], pos[:, 1]) and np.all(rec["nx"] == 0)
    assert np.array_equal(rec["c5"], sc[4][:, 5]) and np.array_equal(rec["scale_2"], sc[2][:, 2]) and np.array_equal(rec["rot_0"], sc[1][:, 0])
    # (ii) the product reader
    V_int, attrs, vs, vm = ply_io.read_compressed_3dgs_ply(p)
    assert V_int.dtype == torch.int64 and attrs.dtype == torch.float32 and vs == 0.001953125
    assert torch.equal(vm.cpu(), vmin)
    assert np.array_equal(V_int.cpu().numpy(), pos.astype(np.int64))
    want = np.concatenate([sc[1], sc[2], sc[3][:, None], sc[4]], 1)              # quats | scales | opacity | colours (data_util.py:361)
    assert np.array_equal(attrs.cpu().numpy(), want)
    # (iii) the reference's own reader and writer, when the checkout is here
    du, qe = _load_ref("data_util"), _load_ref("quality_eval")
    if du is not None:
        r = du.read_compressed_3dgs_ply(p)
        assert r is not None and np.array_equal(r[0].numpy(), pos.astype(np.int64)) and np.array_equal(r[1].numpy(), want)
        assert r[2] == vs and torch.equal(r[3], vmin)
    p2 = str(tmp_path / "ref_layout.ply")
    if qe is not None and n <= 1000:
        qe.save_ply(p2, *[torch.from_numpy(a) for a in scene_np], voxel_size=0.001953125, vmin=vmin)
    else:
        _write_reference_layout(p2, scene_np, 0.001953125, vmin.numpy())
    V2, a2, vs2, vm2 = ply_io.read_compressed_3dgs_ply(p2)
    assert torch.equal(V2, V_int) and torch.equal(a2, attrs) and vs2 == vs and torch.equal(vm2, vm)
    # error behaviour of the reference reader: None + warning (data_util.py:375-382)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        assert ply_io.read_compressed_3dgs_ply(str(tmp_path / "missing.ply")) is None and w
    with open(p, "r+b") as f:
        f.truncate(size - 5)
    with warnings.catch_warnings(record=True):
        warnings.simplefilter("always")
        assert ply_io.read_compressed_3dgs_ply(p) is None
