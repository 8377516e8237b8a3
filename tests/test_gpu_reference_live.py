"""The REFERENCE ITSELF on the GPU box (-m gpu), next to this repo's CUDA path on the same inputs:

* the reference's own merge kernel (cuda/merge_cluster.cu:2-111), compiled in place for sm_100 by oracle/Makefile
  into oracle/_ref/libmerge_ref_{ieee,default,fast}.so -- pins row (a4) to reference object code;
* the reference's own Python operators (RAHT_param_reorder_fast, RAHT2_optimized, inverse_RAHT_optimized,
  voxelize_pc_batched) run on cuda:0, their production device (encode_3dgs.py:22), from the copy staged under
  oracle/_ref/python, compared through the reference's own comparers crosscheck.compare_raht_param_torch
  (crosscheck.py:200) and crosscheck.compare_RAHT_outputs(rtol=0, atol=0) (crosscheck.py:363);
* what torch-CUDA does with `tensor / python_scalar` (`Coeff / step` at encode_3dgs.py:204 and
  `V0 / voxel_size` at voxelize_pc.py:98): the findings are asserted here and written to
  gpurun_out/reference_live.json (quoted in DESIGN section 2).
Nothing here reads /root/reference at run time."""
import json
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import oracle_ctypes as orc
from oracle import refpy

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FINDINGS = os.path.join(ROOT, "gpurun_out", "reference_live.json")


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    return t if dtype is None else t.to(dtype)


def record(key, value):
    os.makedirs(os.path.dirname(FINDINGS), exist_ok=True)
    data = json.load(open(FINDINGS)) if os.path.exists(FINDINGS) else {}
    data[key] = value
    json.dump(data, open(FINDINGS, "w"), indent=1, sort_keys=True)


# ---------------------------------------------------------------------------------------------------
# a4: the reference merge kernel
# ---------------------------------------------------------------------------------------------------
def _ref_merge(kind, cidx, coff, scene, by_opacity):
    means, quats, scales, opac, colors = scene
    nclu = coff.shape[0] - 1
    cd = colors.shape[1]
    outs = [torch.zeros(nclu, 3, device=DEV), torch.zeros(nclu, 4, device=DEV), torch.zeros(nclu, 3, device=DEV),
            torch.zeros(nclu, device=DEV), torch.zeros(nclu, cd, device=DEV)]
    torch.cuda.synchronize()
    rc = orc.ref_merge_launch(kind, cidx.data_ptr(), coff.data_ptr(), nclu, means.data_ptr(), quats.data_ptr(), scales.data_ptr(),
                              opac.data_ptr(), colors.data_ptr(), cd, by_opacity, [o.data_ptr() for o in outs], 0)
    assert rc == 0, f"reference merge kernel launch failed: cudaError {rc}"
    torch.cuda.synchronize()
    return outs


@pytest.mark.parametrize("n,J,seed", [(300000, 7, 11), (1_000_000, 12, 2001), (400000, 16, 12)])
def test_merge_segments_vs_the_reference_kernel(n, J, seed):
    """raht_merge_segments against merge_weighted_mean_kernel itself.  IEEE build (--fmad=false): bit-exact on every
    voxel, whatever its member count (same summation order, separately rounded products and sums).  Builds with FMA
    contraction (nvcc default) or --use_fast_math (the reference's shipped flags, cuda/setup.py:27-30): bit-exact on the
    means / scales / opacity / colours of single-member voxels with non-zero weight only in the IEEE build; with
    tolerance 1e-5 relative otherwise."""
    if not orc.ref_merge_available("ieee"):
        pytest.skip("oracle/_ref/libmerge_ref_ieee.so not built (make -C oracle ref in the build container)")
    import merge_cluster_cuda as MG
    import voxelize_pc as VX
    sc_np = helpers.synthetic_gaussians(n, seed)
    scene = [cu(a) for a in sc_np]
    _, _, vi, _, info = VX.voxelize_pc_batched(scene[0], J=J, device=DEV)
    cidx = info["sort_idx"].int().contiguous()
    coff = torch.cat([vi, torch.tensor([n], device=DEV)]).int().contiguous()
    members = (coff[1:] - coff[:-1]).long()
    stats = {"n": n, "J": J, "nvox": int(members.numel()), "max_members": int(members.max()), "multi_member_voxels": int((members > 1).sum())}
    for by_opacity in (True, False):
        got = MG.merge_gaussian_clusters_with_indices(*scene, cidx, coff, by_opacity)
        ref = _ref_merge("ieee", cidx, coff, scene, by_opacity)
        for a, b, nm in zip(got, ref, ("means", "quats", "scales", "opacities", "colors")):
            assert torch.equal(a, b), f"{nm} (weight_by_opacity={by_opacity}) differs from the reference kernel (IEEE build)"
        for kind in ("default", "fast"):
            if not orc.ref_merge_available(kind):
                continue
            ref2 = _ref_merge(kind, cidx, coff, scene, by_opacity)
            worst = 0.0
            for a, b, nm in zip(got, ref2, ("means", "quats", "scales", "opacities", "colors")):
                scale = b.abs().max().item() or 1.0
                err = (a - b).abs().max().item() / scale
                worst = max(worst, err)
                assert err <= 1e-5, f"{nm}: {err} vs the reference kernel ({kind} build)"
            stats[f"max_rel_err_{kind}_{'opacity' if by_opacity else 'uniform'}"] = worst
    record(f"merge_vs_reference_kernel_n{n}_J{J}", stats)


# ---------------------------------------------------------------------------------------------------
# the reference's Python operators on cuda:0
# ---------------------------------------------------------------------------------------------------
def test_reference_python_on_cuda_vs_product():
    if not refpy.available():
        pytest.skip("reference python not staged (make -C oracle ref in the build container)")
    ref = refpy.load()
    import RAHT_param as P
    import RAHT as R
    import iRAHT as IR
    from raht_b200 import ops
    assert not os.path.abspath(P.__file__).startswith(os.path.abspath(ref.directory))
    J, D = 12, 59
    Pts, MC = helpers.unique_sorted_voxels(200_000, J, 31, clustered=True)
    N = len(MC)
    rng = np.random.default_rng(31)
    Cm = (rng.normal(size=(N, D)) * np.linspace(0.5, 40, D)[None, :])
    V = cu(Pts, torch.float64)
    origin = torch.zeros(3, dtype=torch.float64, device=DEV)
    # tree: reference on cuda:0 vs the product shim, through the reference's own comparer
    Lr, Fr, Wr, order_r = ref.RAHT_param.RAHT_param_reorder_fast(V, origin, 2 ** J, J)
    Lp, Fp, Wp, order_p = P.RAHT_param_reorder_fast(V, origin, 2 ** J, J)
    assert ref.crosscheck.compare_raht_param_torch(Lr, Fr, Wr, list(Lp), list(Fp), list(Wp), strict_types=True)
    assert torch.equal(order_r, order_p)
    # RAHT_param (the bounds-checking variant, :6-83) gives the same lists
    L2, F2, W2 = ref.RAHT_param.RAHT_param(V, origin, 2 ** J, J)
    Lq, Fq, Wq = P.RAHT_param(V, origin, 2 ** J, J)
    assert ref.crosscheck.compare_raht_param_torch(L2, F2, W2, list(Lq), list(Fq), list(Wq))
    # transform: bit-exact (rtol = atol = 0), forward and inverse
    C = cu(Cm)
    Tr, wr = ref.RAHT.RAHT2_optimized(C, Lr, Fr, Wr)
    Tp, wp = R.RAHT2_optimized(C, Lp, Fp, Wp)
    assert ref.crosscheck.compare_RAHT_outputs(Tr, wr, Tp, wp, rtol=0.0, atol=0.0)
    assert torch.equal(Tr, Tp)
    Cr = ref.iRAHT.inverse_RAHT_optimized(Tr, Lr, Fr, Wr)
    Cp = IR.inverse_RAHT_optimized(Tp, Lp, Fp, Wp)
    assert torch.equal(Cr, Cp)
    # the product fed with the REFERENCE's lists (plain tensors that did not come from this package)
    Tx, _ = R.RAHT2_optimized(C, Lr, Fr, Wr)
    assert torch.equal(Tx, Tr)
    # quantisation, encode_3dgs.py:204: torch.floor(Coeff / step + 0.5) on cuda:0.  SURVEY 7.2 item 3: torch-CUDA
    # evaluates tensor / python_scalar as tensor * (1 / scalar) -- identical for powers of two, 1 ulp off before the
    # floor for other steps.  Product convention: exact division (torch-CPU); ops.quant_reorder(..., division="reciprocal")
    # reproduces torch-CUDA bit for bit.
    found = {}
    for step in (1, 4, 8, 12, 16, 20, 24, 32, 64, 0.05):
        q_cuda = torch.floor(Tr / step + 0.5)
        q_recip = torch.floor(Tr * (1.0 / step) + 0.5)
        q_exact = torch.floor(torch.from_numpy(Tr.cpu().numpy() / step).to(DEV) + 0.5)          # numpy: true division
        q_prod = ops.quant_reorder(Tp, order_p, step)                                            # (D, N), order_RAGFT order
        q_prod_rows = torch.empty(N, D, dtype=torch.float64, device=DEV)
        q_prod_rows[order_p] = q_prod.t().double()
        assert torch.equal(q_prod_rows, q_exact), f"step {step}: product != floor(T / step + 0.5) with exact division"
        q_prod_r = ops.quant_reorder(Tp, order_p, step, division="reciprocal")
        q_prod_r_rows = torch.empty(N, D, dtype=torch.float64, device=DEV)
        q_prod_r_rows[order_p] = q_prod_r.t().double()
        assert torch.equal(q_prod_r_rows, q_recip), f"step {step}: reciprocal mode != floor(T * (1 / step) + 0.5)"
        found[str(step)] = {"torch_cuda_equals_exact_division": bool(torch.equal(q_cuda, q_exact)),
                            "torch_cuda_equals_reciprocal_multiply": bool(torch.equal(q_cuda, q_recip)),
                            "symbols_differing_exact_vs_torch_cuda": int((q_cuda != q_exact).sum()),
                            "symbols": int(q_cuda.numel())}
        assert torch.equal(q_cuda, q_exact) or torch.equal(q_cuda, q_recip), f"step {step}: torch-CUDA division is neither exact nor reciprocal"
        assert (q_cuda - q_exact).abs().max().item() <= 1.0
    record("quantise_division_on_torch_cuda", found)
    # sqrt / division used by the butterflies' weights: torch-CUDA rounds correctly (the fixtures' assumption)
    w0 = cu(rng.integers(1, 10 ** 6, size=300000).astype(np.float64))
    w1 = cu(rng.integers(1, 10 ** 6, size=300000).astype(np.float64))
    a_np = np.sqrt(w0.cpu().numpy() / (w0.cpu().numpy() + w1.cpu().numpy()))
    assert np.array_equal(torch.sqrt(w0 / (w0 + w1)).cpu().numpy(), a_np)


def test_reference_voxelizer_on_cuda_vs_product():
    """voxelize_pc_batched of the reference on cuda:0 against the product shim.  The reference's torch.sort is not
    stable (ties between Gaussians of one voxel come out in an unspecified order), so the permutation is compared as
    a SET per voxel and everything order-independent bit for bit; `V0 / voxel_size` on torch-CUDA is a multiplication
    by the fp32 reciprocal: the number of Gaussians it moves to a neighbouring voxel is recorded."""
    if not refpy.available():
        pytest.skip("reference python not staged")
    ref = refpy.load()
    import voxelize_pc as VX
    n, J = 500_000, 12
    means = helpers.synthetic_gaussians(n, 41)[0]
    PC = cu(means)
    rv, rs, rvi, rd, rinfo = ref.voxelize_pc.voxelize_pc_batched(PC, J=J, device=DEV)
    pv, ps, pvi, pd, pinfo = VX.voxelize_pc_batched(PC, J=J, device=DEV)
    assert rinfo["N"] == pinfo["N"] == n
    assert float(rinfo["voxel_size"]) == float(pinfo["voxel_size"]) and float(rinfo["width"]) == float(pinfo["width"])
    assert torch.equal(torch.as_tensor(rinfo["vmin"]).to(DEV).float().reshape(-1), torch.as_tensor(pinfo["vmin"]).to(DEV).float().reshape(-1))
    # integer voxel coordinates per Gaussian, both ways of dividing
    vs = float(pinfo["voxel_size"])
    V0 = PC - torch.as_tensor(pinfo["vmin"]).to(DEV).float().reshape(1, 3)
    vint_cuda = torch.clamp(torch.floor(V0 / vs), 0, 2 ** J - 1)                                   # what the reference computes on cuda:0
    vint_exact = torch.clamp(torch.floor(torch.from_numpy(V0.cpu().numpy() / np.float32(vs)).to(DEV)), 0, 2 ** J - 1)
    moved = int((vint_cuda != vint_exact).any(dim=1).sum())
    same_voxels = bool(rinfo["Nvox"] == pinfo["Nvox"] and torch.equal(rv[:, :3], pv[:, :3]))
    record("voxelize_division_on_torch_cuda", {"n": n, "J": J, "gaussians_in_another_voxel_exact_vs_torch_cuda": moved,
                                               "nvox_reference_cuda": int(rinfo["Nvox"]), "nvox_product": int(pinfo["Nvox"]),
                                               "voxel_lists_equal": same_voxels})
    if moved == 0:
        assert same_voxels, "same quantisation, different voxel list"
        # per-voxel member sets agree (tie order inside a voxel is unspecified in the reference)
        ridx, pidx = rinfo["sort_idx"].long(), pinfo["sort_idx"].long()
        seg = torch.zeros(n, dtype=torch.long, device=DEV)
        seg[pvi[1:]] = 1
        seg = torch.cumsum(seg, 0)
        key_r = seg * n + ridx
        key_p = seg * n + pidx
        assert torch.equal(torch.sort(key_r).values, torch.sort(key_p).values)
    else:
        # the product follows the exact-division convention (torch-CPU, the survey's oracle); torch-CUDA's reciprocal
        # moves a few boundary Gaussians: every other Gaussian must land in the same voxel
        assert moved <= max(10, n // 1000), moved
