#!/usr/bin/env python
"""bench.py -- Gaussians/s of the full RAHT-3DGS encode + decode (59 channels) on B200.

One "step" = one pass of the hot path over one synthetic scene:
    encode: voxelise (Morton, radix sort, segments) -> merge -> RAHT_param (tree, plan, order) ->
            RAHT (fp64) -> quantise + reorder -> RLGR encode (59 streams)
    decode: RAHT_param from the voxel keys -> RLGR decode -> dequantise + inverse reorder -> iRAHT
`value`  : Gaussians/s with the scene already resident in HBM (CUDA events, max over ranks).
`e2e`    : the same through the public Python API from pinned HOST buffers, H2D of the scene and D2H
           of the bitstreams (and, `with_attributes`, of the decoded attributes) inside the timed region.
`roofline`: the dominant kernel family of the step, algorithmic bytes / measured duration vs the
           measured HBM peak of MEASURED_PEAKS.json.
`parity` : the GPU bitstreams / keys / decoded attributes of the timed scene compared with the CPU checker
           (the cpu_baseline leg, which runs the whole scene anyway) -- the bench checks what it times.
`cpu_baseline` / --impl reference: the reference's CPU path on the host cores (see run_reference).
Sub-records of the N = 1 line: `c3` (BASELINE configs[2]: 6 M Gaussians, depth 16) and `sparse` (C2 at a coarse step).
N > 1 ranks (torchrun): every rank codes its own scene (scene-per-GPU replicas, weak scaling, no data-path collective)
for the headline; the line additionally carries `sharded` (BASELINE configs[3]: ONE 50 M-Gaussian scene over all
ranks by Morton prefix, strong scaling, NCCL) and `c5` (BASELINE configs[4]: 64 scenes x 2 M, depth 14, streamed).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "raht-3dgs-codec_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (n_gaussians, depth, seed, description)          BASELINE.json configs[1] / configs[2]
    "c2": (1_000_000, 12, 2001, "3DGS scene 1M Gaussians, 59 channels, depth 12 (BASELINE configs[1])"),
    "c3": (6_000_000, 16, 3001, "3DGS scene 6M Gaussians, 59 channels, depth 16 (BASELINE configs[2])"),
    "tiny": (50_000, 10, 2001, "smoke-sized scene"),
    # one scene sharded over the ranks by Morton prefix (raht_b200/sharded.py); --num-gaussians overrides the 50 M
    "c4": (50_000_000, 18, 4001, "50M-Gaussian synthetic scene, depth 18, Morton-prefix sharding (BASELINE configs[3])"),
    "c5": (2_000_000, 14, 5001, "64 scenes x 2M Gaussians, depth 14, one scene stream per GPU (BASELINE configs[4])"),
    # the encode_ply.py workload: a voxelised RGB point cloud, 11 quantisation steps per frame (run_c1)
    "c1": (750_000, 10, 1001, "python/encode_ply.py RAHT+RLGR on a synthetic 10-bit voxelised point cloud (~750K voxels, RGB; BASELINE configs[0])"),
}
STEP = 1.0 / 64          # uniform quantisation step applied to all 59 channels (dense regime: ~300 bits per voxel)
SPARSE_STEP = 0.25       # the coarse step of the `sparse` sub-record (most channels mostly zero: RLGR run mode, tables)
D = 59
METRIC = "Gaussians/s full RAHT encode+decode (59 ch)"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------
# CPU paths: used for cpu_baseline / parity and for --impl reference.  The only places that load oracle/.
# ---------------------------------------------------------------------------------------------------
def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_pipeline(scene, J, step, use_ref_rlgr, want_streams=False):
    """Full encode + decode on the host with the CPU restatement of the reference (oracle/raht_oracle.c, OpenMP)."""
    from oracle import oracle_ctypes as orc
    orc.set_num_threads(host_cores())                     # (torchrun exports OMP_NUM_THREADS=1: the CPU arm gets every core anyway)
    means, quats, scales, opac, colors = scene
    n = means.shape[0]
    r = orc.voxelize(means, J, want_delta=False)
    coff = np.concatenate([r["voxel_indices"], [n]]).astype(np.int32)
    mm, mq, ms, mo, mc = orc.merge_clusters(r["sort_idx"].astype(np.int32), coff, means, quats, scales, opac, colors, True)
    packed = np.concatenate([mm, mq, ms, mo[:, None], mc], 1).astype(np.float64)
    keys = r["MC_sorted"][r["voxel_indices"]]
    t = orc.Tree(keys, J)
    T, _ = t.forward(packed)
    order = t.order()
    q = np.ascontiguousarray(orc.quantize(T, step)[order].T.astype(np.int32))
    nvox = q.shape[1]
    if use_ref_rlgr:
        # the reference codes its channels one after the other on one core (encode_3dgs.py:225-247); here every channel
        # gets its own host thread (ctypes releases the GIL), so the CPU arm uses all the cores it can
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=host_cores()) as pool:
            streams = list(pool.map(lambda c: orc.ref_rlgr_encode(q[c].astype(np.int64), 1)[0], range(q.shape[0])))
            qd = np.stack(list(pool.map(lambda c: orc.ref_rlgr_decode(streams[c], nvox, 1)[0], range(q.shape[0])))).astype(np.int32)
    else:
        streams, out, cap, nb = orc.rlgr_encode_channels(q, 1)
        qd = orc.rlgr_decode_channels(out, cap, nb, nvox, q.shape[0], 1)
    # decoder side: tree again from the keys, dequantise, inverse reorder, iRAHT
    t2 = orc.Tree(keys, J)
    inv = np.argsort(t2.order(), kind="stable")
    Tdec = orc.dequantize(qd.T.astype(np.float64), step)[inv]
    Crec = t2.inverse(Tdec)
    if want_streams:
        return nvox, Crec, keys, list(streams)
    return nvox, Crec


def reference_python_pipeline(scene, J, step, stage_s=None):
    """The REFERENCE's own CPU path, in the call order of its drivers, on torch-CPU with every host core:
        test_voxelize_3dgs.py:203-247   voxelize_pc_batched -> cluster arrays -> merge
        encode_3dgs.py:144-276          RAHT_param_reorder_fast -> RAHT2_optimized -> floor(Coeff/step+0.5) ->
                                        index_select(order_RAGFT) -> int32 -> rlgr per channel -> decode -> * step ->
                                        argsort(order) gather -> inverse_RAHT_optimized
    The Python operators are the reference's files (oracle/refpy.py: /root/reference/python, else the copy staged
    under oracle/_ref/python); the entropy coder is the reference's object code (oracle/_ref/librlgr_ref.so).  Two
    deviations, both in the reference's FAVOUR: the merge, which the reference only has as a CUDA kernel, is the C
    restatement with OpenMP; and the channels are coded one per host thread through a ctypes veneer instead of one
    after the other through Python lists (encode_3dgs.py:218-247: 9.9 s wall for 4.5 s inside the coder)."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle_ctypes as orc
    from oracle import refpy
    ref = refpy.load()
    cores = host_cores()
    torch.set_num_threads(cores)
    orc.set_num_threads(cores)
    means, quats, scales, opac, colors = scene
    n = means.shape[0]
    marks = [("start", time.perf_counter())]

    def mark(name):
        marks.append((name, time.perf_counter()))

    with torch.no_grad():
        PCvox, _, voxel_indices, _, info = ref.voxelize_pc.voxelize_pc_batched(torch.from_numpy(means), J=J, device="cpu")
        mark("voxelize")
        cidx = info["sort_idx"].int().numpy()
        coff = torch.cat([voxel_indices, torch.tensor([n], dtype=voxel_indices.dtype)]).int().numpy()
        mm, mq, ms, mo, mc = orc.merge_clusters(cidx, coff, means, quats, scales, opac, colors, True)
        C = torch.from_numpy(np.concatenate([mm, mq, ms, mo[:, None], mc], 1)).to(torch.float64)
        mark("merge")
        V = PCvox[:, :3].to(torch.float64)
        origin = torch.zeros(3, dtype=torch.float64)
        ListC, FlagsC, weightsC, order = ref.RAHT_param.RAHT_param_reorder_fast(V, origin, 2 ** J, J)
        mark("raht_param")
        Coeff, _ = ref.RAHT.RAHT2_optimized(C, ListC, FlagsC, weightsC)
        mark("raht")
        Coeff_enc = torch.floor(Coeff / step + 0.5)
        reordered = Coeff_enc.index_select(0, order)
        cols = np.ascontiguousarray(reordered.to(torch.int32).numpy().T)
        mark("quant_reorder")
        nvox = cols.shape[1]
        with ThreadPoolExecutor(max_workers=cores) as pool:
            streams = list(pool.map(lambda c: orc.ref_rlgr_encode(cols[c].astype(np.int64), 1)[0], range(cols.shape[0])))
            mark("rlgr_encode")
            dec = list(pool.map(lambda c: orc.ref_rlgr_decode(streams[c], nvox, 1)[0], range(cols.shape[0])))
        mark("rlgr_decode")
        coeff_dec = torch.from_numpy(np.stack(dec, axis=1).astype(np.int32)).to(torch.float64) * step
        coeff_dec = coeff_dec[torch.argsort(order), :]
        mark("dequant_unreorder")
        C_rec = ref.iRAHT.inverse_RAHT_optimized(coeff_dec, ListC, FlagsC, weightsC)
        mark("iraht")
    if stage_s is not None:
        for (_, t0), (name, t1) in zip(marks[:-1], marks[1:]):
            stage_s[name] = stage_s.get(name, 0.0) + (t1 - t0)
    return nvox, C_rec, sum(len(b) for b in streams)


# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and clock-event reasons sampled DURING the timed regions.

    NVML in this process (pynvml: two light queries, SM clock and event-reason mask, every 10 ms from a side thread that
    holds no GIL inside the calls); `nvidia-smi -lms` in a side process only when pynvml is missing.  Both are started
    BEFORE the warm-up.  History: nvidia-smi started together with the timed region stalled kernel launches for ~10 ms
    while it initialised NVML; once running, each of its polls (it reads power and a dozen more fields) still put a
    1-2 ms bubble into whichever stage was executing -- one step in five at 50 ms polling."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.rows = []          # nvidia-smi: (time, csv line); nvml: (time, sm MHz, reason mask)
        self.proc = None
        self.nvml = None
        self.handle = None
        self.max_mhz = None
        self.stop_flag = False
        self.t0 = self.t1 = None

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def _nvml_handle(self, nv):
        try:
            import torch
            p = torch.cuda.get_device_properties(self.idx)
            bus = f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
            return nv.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            return nv.nvmlDeviceGetHandleByIndex(self.idx)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.handle = self._nvml_handle(nv)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM))
            nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
            self.nvml = nv
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll_nvml(self):
        nv = self.nvml
        while not self.stop_flag:
            try:
                t = time.time()
                mhz = float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                self.rows.append((t, mhz, mask))
            except Exception:
                pass
            time.sleep(0.01)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self):
        t0, t1 = self.t0 or 0.0, self.t1 or time.time()
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            nv = self.nvml
            inside = [r for r in self.rows if t0 <= r[0] <= t1 + 0.02]
            names = [("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                     ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap),
                     ("hw_power_brake_slowdown", nv.nvmlClocksEventReasonHwPowerBrakeSlowdown)]
            reasons = sorted({nm for _, _, m in inside for nm, bit in names if m & bit})
            sm = [r[1] for r in inside]
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(sm),
                    "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [r for t, r in self.rows if t0 <= t <= t1 + 0.05]
        for r in inside:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2])
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvidia-smi"}




def stage_bytes(nvox, n_in, payload, J):
    """Algorithmic (compulsory) bytes per stage for one step, SURVEY 8(d)."""
    passes = (3 * J + 7) // 8
    return {
        "morton": n_in * 24, "sort": n_in * 24 * passes, "segments": n_in * 8 + nvox * 20,
        "merge": n_in * (12 + 236) + nvox * 236,
        "raht_param": nvox * 21, "raht_param_dec": nvox * 21,
        "raht": nvox * (D * 4 + D * 8 + 13),            # fp32 packed rows in, fp64 coefficients out
        "quant_reorder": nvox * (D * 8 + 8 + D * 4),
        "raht_quant": nvox * (D * 4 + 13 + 4 + D * 4),  # fused: fp32 rows in, int32 symbols out (the fp64 coefficients stay on chip)
        "dequant_iraht": nvox * (D * 4 + 13 + 4 + D * 8),
        "rlgr_encode": nvox * D * 4 + payload,
        "rlgr_decode": payload + nvox * D * 4,
        "dequant_unreorder": nvox * (D * 4 + 8 + D * 8),
        "iraht": nvox * (2 * D * 8 + 13),
        "geometry": nvox * 20,
    }


# kernels of each pipeline stage with their launches per step: DRAM traffic of a stage = sum over its kernels of the
# per-launch dram__bytes_read.sum + dram__bytes_write.sum of the latest `ncu --set full` capture kept under profiles/
STAGE_KERNELS = {
    "rlgr_encode": [("rlgr_count_kernel", 1), ("rlgr_scan_kernel", 1), ("rlgr_scatter_kernel", 1), ("rlgr_speculate_kernel", 2),
                    ("rlgr_table_kernel", 1), ("rlgr_chunk_table_kernel", 1), ("rlgr_resolve_kernel", 1), ("rlgr_table_lanes_kernel", 1),
                    ("rlgr_zero_kernel", 1), ("rlgr_emit_kernel", 1)],
    "raht_param": [("levels_kernel", 1), ("order_count_kernel", 1), ("scan_counts_kernel", 1), ("order_scatter_kernel", 1),
                   ("plan_init_kernel", 1), ("plan_kernel", 1)],
    "rlgr_decode": [("rlgr_decode_indexed_kernel", 1)],
    "merge": [("merge_kernel<float, 2, 0>", 1)],
    "raht": [("transform_stage0_kernel<float, double, 59, 1, 0>", 1), ("transform_upper_kernel<double, 1>", 5)],
    "iraht": [("transform_stage0_kernel<double, double, 59, 0, 0>", 1), ("transform_upper_kernel<double, 0>", 5)],
    "quant_reorder": [("quant_reorder_kernel<double, 2, 0>", 1)],
    "dequant_unreorder": [("dequant_unreorder_kernel<double>", 1)],
}


def stage_traffic(stage):
    """(bytes per step or None, source file) from the newest profiles/*ncu_full*.json (scripts/ncu_full_summary.py);
    the capture's own STAGE_KERNELS table ("_stage_kernels" in the json) wins over the one above when present."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu_full*.json")),
                   key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])      # r01 ... v8 < v11 < r02
    if not files:
        return None, None
    prof = json.load(open(files[-1]))
    table = prof.get("_stage_kernels", STAGE_KERNELS)
    if stage not in table:
        return None, os.path.basename(files[-1])
    total = 0
    for name, launches in table[stage]:
        if name not in prof:
            return None, os.path.basename(files[-1])
        total += launches * (prof[name]["dram_read_bytes"] + prof[name]["dram_write_bytes"])
    return int(total), os.path.basename(files[-1])


def bind_rank_to_cores(local, world):
    """Before anything is pinned: bind this rank to its own share of the host cores next to its GPU (NVML's ideal CPU set
    for the device, intersected with what the process may use, split between the ranks that share it), so that the
    pinned staging buffers are first-touched on that NUMA node and the ranks' host threads do not migrate over each
    other.  Returns a short description for the JSON line.  N = 1: untouched (the CPU baseline wants every core)."""
    if world <= 1:
        return {"bound": False, "cores": host_cores()}
    if os.environ.get("RAHT_BENCH_BIND", "1") == "0":
        return {"bound": False, "cores": host_cores(), "note": "RAHT_BENCH_BIND=0"}
    try:
        allowed = sorted(os.sched_getaffinity(0))
        pools = []
        try:
            import pynvml as nv
            nv.nvmlInit()
            words = (max(allowed) + 64) // 64
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            ids = [int(x) for x in vis.split(",")] if vis and all(x.strip().isdigit() for x in vis.split(",")) else list(range(world))
            for r in range(world):
                h = nv.nvmlDeviceGetHandleByIndex(ids[r] if r < len(ids) else r)
                mask = nv.nvmlDeviceGetCpuAffinity(h, words)
                ideal = {w * 64 + b for w, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1}
                pools.append(tuple(sorted(ideal & set(allowed))) or tuple(allowed))
        except Exception:
            pools = [tuple(allowed)] * world
        mine = pools[local]
        sharers = [r for r in range(world) if pools[r] == mine]
        pos = sharers.index(local)
        per = len(mine) // len(sharers)
        share = list(mine[pos * per:(pos + 1) * per]) if per >= 2 else list(mine)
        os.sched_setaffinity(0, share)
        return {"bound": per >= 2, "cores": len(share), "first_core": share[0], "gpu_ideal_cores": len(mine), "ranks_sharing": len(sharers)}
    except Exception as e:                                 # binding is an optimisation, never a reason to fail the bench
        return {"bound": False, "error": str(e)[:80]}


def run_gpu(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    binding = bind_rank_to_cores(local, world)
    from raht_b200 import codec, ops, synth
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        from raht_b200 import sharded
        sharded.tune_nccl_env()                            # point-to-point channels for the sharded sub-record's exchanges
        dist.init_process_group("nccl", device_id=dev)
    n, J, seed, desc = WORKLOADS[args.workload]
    if args.n:
        n = args.n
    tdtype = torch.float64 if args.dtype == "f64" else torch.float32

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world == 1:
            return [float(x)]
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    clocks = ClockSampler(local)
    clocks.start()
    scene = synth.synthetic_gaussians(n, seed + rank)
    host = [torch.from_numpy(a).pin_memory() for a in scene]
    devt = [h.to(dev) for h in host]
    h2d_bytes = sum(h.numel() * h.element_size() for h in host)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def measure_resident(inputs, J_, step, steps, warmup):
        """W warm-up steps, then K timed steps (CUDA events, barrier + synchronize on both sides) with per-stage events."""
        # The status words and stream lengths of a step are read and checked ONE STEP BEHIND (codec's `checks=`): every step
        # is checked, but the GPU does not wait for the host between a step's encode, its decode and the next step.
        pending = []

        def one(timer=None):
            mine = []
            enc = codec.encode_scene(*inputs, J=J_, step=step, transform_dtype=tdtype, tile_rows=args.tile_rows, timer=timer, fused=args.fused,
                                     checks=mine)
            _, Crec = codec.decode_scene(enc, reuse_tree=False, tile_rows=args.tile_rows, timer=timer, fused=args.fused, checks=mine)
            for c in pending:
                c()
            pending[:] = mine
            return enc, Crec

        def settle():
            for c in pending:
                c()
            pending[:] = []
        for _ in range(warmup):
            enc, Crec = one()
        settle()
        barrier()
        timers = []
        launches0 = ops.lib.raht_launch_count()
        mem0 = torch.cuda.memory_stats(dev)
        barrier()
        e0.record()
        for _ in range(steps):
            tm = codec.StageTimer(True)
            enc, Crec = one(tm)
            timers.append(tm)
        settle()                                                  # (inside the timed region: the last step's checks)
        e1.record()
        barrier()
        ms_ranks = [round(x / steps, 4) for x in gather_ranks(e0.elapsed_time(e1))]
        ms_step = max(ms_ranks)
        launches = (ops.lib.raht_launch_count() - launches0) // max(steps, 1)
        mem1 = torch.cuda.memory_stats(dev)
        stages, stage_max = {}, {}
        for tm in timers:
            for k, v in tm.result().items():
                stages[k] = stages.get(k, 0.0) + v / len(timers)
                stage_max[k] = max(stage_max.get(k, 0.0), v)
        return {"ms_step": ms_step, "ms_ranks": ms_ranks, "launches": int(launches), "stages": stages, "stage_max": stage_max, "enc": enc, "Crec": Crec,
                "cuda_mallocs": int(mem1.get("segment.all.allocated", 0) - mem0.get("segment.all.allocated", 0))}

    def per_stage_table(res, n_in, J_, peak):
        enc = res["enc"]
        sb = stage_bytes(enc.nvox, n_in, enc.payload_bytes(), J_)
        # legs that run side by side (config.overlap): the tree build's own row is only what the other leg did not cover, so
        # it gets no GB/s; the decoder's two legs are rated together in the rlgr_decode row (bytes of both over the time of both)
        dec_side = os.environ.get("RAHT_DEC_SIDE", "1") != "0"
        enc_side = os.environ.get("RAHT_ENC_SIDE", "1") != "0"
        tab = {}
        for k, v in res["stages"].items():
            b = sb.get(k, 0)
            rate = True
            if k == "raht_param_dec" and dec_side or k == "raht_param" and enc_side:
                rate = False
            if k == "rlgr_decode" and dec_side:
                b += sb.get("raht_param_dec", 0)
            if k == "merge" and enc_side:
                b += sb.get("raht_param", 0)
            tab[k] = {"ms": round(v, 4), "max_ms": round(res["stage_max"][k], 4),
                      "GBps": round(b / (v / 1e3) / 1e9, 1) if v > 0 and rate else None,
                      "frac_of_peak": round(b / (v / 1e3) / 1e9 / peak, 4) if v > 0 and rate else None}
            if not rate:
                tab[k]["overlapped"] = True
        return sb, tab

    peak, peak_src = load_peaks()
    clocks.mark_begin()
    main = measure_resident(devt, J, STEP, args.steps, args.warmup)
    ms_step = main["ms_step"]
    value = n * world / (ms_step / 1e3)
    enc = main["enc"]
    nvox, payload, index_bytes = enc.nvox, enc.payload_bytes(), enc.index_bytes()

    # ---- parity of what was just timed (rank 0, N = 1): GPU results of the timed scene vs the CPU checker ------------
    cpu = parity = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle_ctypes as orc
        got_streams = enc.streams.to_bytes()
        got_keys = enc.keys.cpu().numpy()
        got_rec = main["Crec"].cpu().numpy()
        t0 = time.perf_counter()
        nvox_cpu, Crec_cpu, keys_cpu, streams_cpu = cpu_pipeline(scene, J, STEP, False, want_streams=True)
        secs = time.perf_counter() - t0
        cpu = {"value": round(n / secs, 1), "unit": "Gaussians/s", "cores": orc.num_threads(), "kind": "port",
               "sample": f"the whole timed scene ({n} Gaussians), full encode+decode once ({secs:.1f} s), oracle/raht_oracle.c with OpenMP; "
                         "the reference's own Python path is timed by --impl reference"}
        bad = [c for c in range(len(streams_cpu)) if got_streams[c] != streams_cpu[c]]
        parity = {"checked": True, "against": "oracle/raht_oracle.c (CPU restatement pinned to the reference, DESIGN 2) on the timed scene",
                  "nvox_equal": bool(nvox_cpu == nvox), "keys_equal": bool(np.array_equal(got_keys, keys_cpu)),
                  "streams_equal": not bad, "streams_compared": len(streams_cpu), "bytes_compared": int(sum(len(b) for b in streams_cpu)),
                  "decoded_attributes_equal": bool(np.array_equal(got_rec, Crec_cpu))}
        if bad:
            parity["first_differing_channel"] = bad[0]
        del got_streams, got_rec, Crec_cpu, streams_cpu
    main["Crec"] = None

    # ---- end to end from host buffers ------------------------------------------------------------
    # codec.ScenePipeline: every step copies its scene from pinned host memory (H2D on a copy-in stream, prefetched
    # under the previous step's kernels), encodes + decodes it, and copies the bitstreams, sync index, voxel keys and
    # (with_attributes) the decoded attributes back to pinned host memory (copy-out stream, under the next step's
    # kernels).  All of it, including the pipeline's fill and drain, lies inside the timed region.
    def time_e2e(result, hosts, J_, steps):
        pipe = codec.ScenePipeline(J_, STEP, tdtype, args.tile_rows, dev, result=result)

        def run_e2e(k):
            d2h = 0
            pipe.submit(hosts[0])
            for i in range(k):
                if i + 1 < k:
                    pipe.submit(hosts[(i + 1) % len(hosts)])       # H2D of the next step's scene
                d2h = pipe.run()["d2h_bytes"]
            pipe.finish()
            return d2h

        run_e2e(2)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        d2h = run_e2e(steps)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        return ms, d2h, time.perf_counter() - t0

    ms_e2e, d2h, wall_e2e = time_e2e("psnr", [host], J, args.steps)
    e2e_value = n * world / (ms_e2e / args.steps / 1e3)
    ms_attr, d2h_attr, _ = time_e2e("attributes", [host], J, args.steps)
    e2e_attr_value = n * world / (ms_attr / args.steps / 1e3)

    # plain pinned -> device copy, all ranks at once: the host-side ceiling of the e2e number
    probe_src = torch.empty(1 << 29, dtype=torch.uint8).pin_memory()
    probe_dst = torch.empty(1 << 29, dtype=torch.uint8, device=dev)
    probe_dst.copy_(probe_src, non_blocking=True)
    barrier()
    e0.record()
    for _ in range(4):
        probe_dst.copy_(probe_src, non_blocking=True)
    e1.record()
    barrier()
    h2d_probe = gather_ranks(4 * probe_src.numel() / (e0.elapsed_time(e1) / 1e3) / 1e9)
    del probe_src, probe_dst

    # container, outside the metric: bytes and device time of pack / unpack of the last scene
    from raht_b200 import container as cont
    box = cont.pack(enc)
    cont.unpack(box)
    torch.cuda.synchronize()
    c0, c1, c2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    c0.record()
    for _ in range(3):
        box = cont.pack(enc)
    c1.record()
    for _ in range(3):
        back = cont.unpack(box)
    c2.record()
    torch.cuda.synchronize()
    hb = cont.parse_header(box[:76].cpu().numpy().tobytes())["header_bytes"]
    hdr = cont.parse_header(box[:hb].cpu().numpy().tobytes())
    container_info = {"bytes": int(box.numel()), "geometry_bytes": int(sum(hdr["geo_bytes"])),
                      "raw_keys_bytes": int(enc.nvox * 8), "pack_ms": round(c0.elapsed_time(c1) / 3, 4),
                      "unpack_ms": round(c1.elapsed_time(c2) / 3, 4), "keys_roundtrip": bool(torch.equal(back.keys, enc.keys)),
                      "what": "raht_b200.container: header + RLGR-coded Morton deltas + the attribute streams + sync indices in one device buffer"}
    del box, back
    clocks.mark_end()
    clk = clocks.stop()
    clk["window"] = "samples between the start of the device-timed region and the end of the e2e regions"

    sb, per_stage = per_stage_table(main, n, J, peak)
    stages = main["stages"]
    dom = max(stages, key=stages.get) if stages else None
    roofline = None
    if dom:
        achieved = sb.get(dom, 0) / (stages[dom] / 1e3) / 1e9
        traffic, traffic_src = stage_traffic(dom) if args.workload == "c2" and not args.n else (None, None)
        raht_ms = stages.get("raht", 0.0) + stages.get("iraht", 0.0)
        roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                    "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": sb.get(dom, 0), "ms_per_launch": round(stages[dom], 4),
                    "raht_passes": {"ms": round(raht_ms, 4), "GBps": round((sb["raht"] + sb["iraht"]) / (raht_ms / 1e3) / 1e9, 1),
                                    "frac": round((sb["raht"] + sb["iraht"]) / (raht_ms / 1e3) / 1e9 / peak, 4)} if raht_ms > 0 else None,
                    "note": "kernel = the pipeline stage with the largest CUDA-event time in this run (a group of kernels, STAGE_KERNELS); "
                            "per-stage GB/s and fractions are under `stages`; raht_passes = forward + inverse transform together "
                            "(the north-star bar is >= 0.60 on these)"}

    # ---- sub-records ------------------------------------------------------------------------------
    del devt
    main["enc"] = enc = None
    torch.cuda.empty_cache()
    sub = {}
    if world == 1 and args.workload == "c2" and not args.n and not args.no_subrecords:
        # the same scene at a coarse step: most channels mostly zero (RLGR run mode, helper warps, transition tables)
        devt2 = [h.to(dev) for h in host]
        r = measure_resident(devt2, J, SPARSE_STEP, max(args.steps, 3), 3)
        _, st_tab = per_stage_table(r, n, J, peak)
        sub["sparse"] = {"quant_step": SPARSE_STEP, "ms_per_step": round(r["ms_step"], 4), "value": round(n / (r["ms_step"] / 1e3), 1),
                         "unit": "Gaussians/s", "payload_bytes": r["enc"].payload_bytes(),
                         "bits_per_voxel": round(r["enc"].payload_bytes() * 8 / r["enc"].nvox, 3),
                         "rlgr_encode_ms": round(r["stages"].get("rlgr_encode", 0.0), 4), "rlgr_decode_ms": round(r["stages"].get("rlgr_decode", 0.0), 4)}
        # render-quality evaluation (quality_eval.py:373-520): original against decoded Gaussians of the timed configuration
        try:
            sub["render_eval"] = render_record(devt2, codec, J, STEP)
        except Exception as e:                                   # a reported extra, never a reason to lose the line
            sub["render_eval"] = {"error": str(e)[:120]}
        del devt2, r
        torch.cuda.empty_cache()
        # BASELINE configs[2]: 6 M Gaussians, depth 16, fp64 bit-exact mode (the scene tests/test_gpu_scale_parity.py checks)
        n3, J3, seed3, desc3 = WORKLOADS["c3"]
        scene3 = synth.synthetic_gaussians(n3, seed3)
        dev3 = [torch.from_numpy(a).to(dev) for a in scene3]
        del scene3
        r = measure_resident(dev3, J3, STEP, max(args.steps, 3), 3)
        sb3, st_tab = per_stage_table(r, n3, J3, peak)
        enc_stages = ("morton", "sort", "segments", "merge", "raht_param", "raht", "quant_reorder", "rlgr_encode")
        raht_ms = r["stages"].get("raht", 0.0) + r["stages"].get("iraht", 0.0)
        sub["c3"] = {"workload": desc3, "n_gaussians": n3, "nvox": r["enc"].nvox, "depth": J3, "quant_step": STEP,
                     "ms_per_step": round(r["ms_step"], 4), "value": round(n3 / (r["ms_step"] / 1e3), 1), "unit": "Gaussians/s",
                     "encode_ms": round(sum(r["stages"].get(k, 0.0) for k in enc_stages), 4),
                     "gpu_launches": r["launches"], "payload_bytes": r["enc"].payload_bytes(),
                     "raht_passes": {"ms": round(raht_ms, 4), "frac": round((sb3["raht"] + sb3["iraht"]) / (raht_ms / 1e3) / 1e9 / peak, 4)},
                     "stages": st_tab}
        del dev3, r
        torch.cuda.empty_cache()
    if not args.no_subrecords and args.workload == "c2" and not args.n:
        sub["c5"] = c5_record(args, dev, rank, world, barrier, max_over_ranks, tdtype)
        torch.cuda.empty_cache()
        if world > 1:
            sub["sharded"] = sharded_record(args, dev, rank, world, barrier, max_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    line = {
        "metric": METRIC, "value": round(value, 1), "unit": "Gaussians/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
        "ms_per_step_ranks": main["ms_ranks"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": desc, "n_gaussians_per_gpu": n, "nvox": nvox, "depth": J, "channels": D, "quant_step": STEP,
                   "tile_rows": args.tile_rows or ops.DEFAULT_TILE_ROWS, "payload_bytes": payload, "sync_index_bytes": index_bytes,
                   "bits_per_voxel": round(payload * 8 / nvox, 3),
                   "l2": "inputs (236 MB per scene) and every intermediate exceed the 126 MB L2; no explicit flush",
                   "host_checks": "status words and stream lengths of every step are read one step behind (encode_scene / decode_scene checks=), "
                                  "all of them inside the timed region; the one synchronisation left inside a step is the voxel count",
                   "overlap": "the tree build (levels, order, plan) runs on the library's side stream next to the merge (encoder) and next to "
                              "the RLGR decode (decoder), RAHT_ENC_SIDE / RAHT_DEC_SIDE: in `stages`, raht_param / raht_param_dec are the part "
                              "of the build that the other leg does not cover, and merge / rlgr_decode hold both legs (the merge runs as short-lived CTAs "
                              "and the decoder with 9 CTAs per SM so that the tree's kernels find room)",
                   "parallelism": "scene-per-GPU (no collective)" if world > 1 else "single GPU",
                   "decoder": "rebuilds the tree from the voxel keys (no encoder state reused); RLGR decode is chunk-parallel "
                              "through the encoder's sync index (side information next to the reference-format streams)"},
        "clocks": clk,
        "e2e": {"value": round(e2e_value, 1), "unit": "Gaussians/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": int(d2h),
                "wall_s": round(wall_e2e, 3),
                "how": "codec.ScenePipeline(result='psnr'): H2D of step k+1 and D2H of step k-1 overlap the kernels of step k (3 streams); "
                       "read back per step: bitstreams + sizes, sync index, voxel keys, squared error of the reconstruction "
                       "(what the reference driver reads back, encode_3dgs.py:247,298-310)",
                "h2d_GBps_per_rank_achieved": round(h2d_bytes / (ms_e2e / args.steps / 1e3) / 1e9, 2),
                "h2d_probe_GBps_per_rank": [round(x, 1) for x in h2d_probe],
                "h2d_probe": "plain pinned->device copy of 512 MiB x 4 on every rank at the same time: the host-side ceiling of this number",
                "host_binding": binding,
                "with_attributes": {"value": round(e2e_attr_value, 1), "unit": "Gaussians/s", "d2h_bytes_per_step": int(d2h_attr),
                                    "how": "the same plus the decoded attributes (nvox x 59 fp64) copied to pinned host memory"}},
        "gpu_launches": main["launches"],
        "cuda_mallocs_in_timed_region": main["cuda_mallocs"],
        "roofline": roofline,
        "stages": per_stage,
        "cpu_baseline": cpu,
        "parity": parity,
        "container": container_info,
    }
    line.update(sub)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def render_record(scene_dev, codec, J, step):
    """The reference's image-space check (quality_eval.try_render_comparison) on the CUDA splat rasteriser: the 1 M scene
    against its reconstruction from the coded streams (voxel-merged, quantised at `step`, decoded), 3 views of 512 x 512.
    The synthetic scales are shrunk so that a splat covers a few pixels, as in a trained scene."""
    import torch
    from raht_b200 import render_eval
    m, q, s, o, c = scene_dev
    orig = dict(means=m, quats=q, scales=s * 0.25, opacities=o, colors=c)
    coords, A = codec.decode_scene(codec.encode_scene(m, q, s, o, c, J=J, step=step))
    A = A.float()
    dec = dict(means=A[:, 0:3].contiguous(), quats=A[:, 3:7].contiguous(), scales=A[:, 7:10].contiguous() * 0.25,
               opacities=A[:, 10].contiguous(), colors=A[:, 11:].contiguous())
    g = torch.Generator().manual_seed(4)
    render_eval.try_render_comparison(orig, dec, n_views=1, image_size=512, generator=torch.Generator().manual_seed(4))   # warm-up
    res = render_eval.try_render_comparison(orig, dec, n_views=3, image_size=512, generator=g)
    return {"views": 3, "image": "512x512", "quant_step": step, "n_original": int(m.shape[0]), "n_decoded": int(A.shape[0]),
            "psnr_avg_db": round(res["psnr_avg"], 3), "psnr_min_db": round(res["psnr_min"], 3),
            "ms_per_view_original": round(res["original_render_time_ms"] / 3, 3), "ms_per_view_decoded": round(res["merged_render_time_ms"] / 3, 3),
            "how": "raht_b200.render_eval (csrc/splat.cu): project + tile sort + rasterise per view, host-timed like the reference (quality_eval.py:421-447)"}


def c5_record(args, dev, rank, world, barrier, max_over_ranks, tdtype):
    """BASELINE configs[4] (encode_dataset.py:29,104 walks frames one after the other): 64 scenes x 2 M Gaussians, depth 14,
    one scene stream per GPU -- 64 / N scenes per rank through codec.ScenePipeline from pinned host memory (H2D, encode,
    decode, D2H of streams + index + keys + squared error inside the timed region).  Each rank cycles through
    `distinct` different scenes (generated on the device, copied to pinned host memory before the clock starts)."""
    import torch
    from raht_b200 import codec, synth
    n5, J5, seed5, desc5 = WORKLOADS["c5"]
    total_scenes = 64
    per_rank = max(total_scenes // world, 1)
    distinct = min(per_rank, 4)
    hosts = []
    for k in range(distinct):
        sc = synth.synthetic_gaussians_device(n5, seed5 + rank * 64 + k, dev)
        hosts.append([t.cpu().pin_memory() for t in sc])
        del sc
    torch.cuda.empty_cache()
    pipe = codec.ScenePipeline(J5, STEP, tdtype, args.tile_rows, dev, result="psnr")

    def run(k):
        pipe.submit(hosts[0])
        for i in range(k):
            if i + 1 < k:
                pipe.submit(hosts[(i + 1) % distinct])
            pipe.run()
        pipe.finish()

    run(3)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(per_rank)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    h2d = sum(h.numel() * h.element_size() for h in hosts[0])
    return {"workload": desc5, "scenes": per_rank * world, "scenes_per_rank": per_rank, "distinct_scenes_per_rank": distinct,
            "n_gaussians_per_scene": n5, "depth": J5, "quant_step": STEP, "ms_total": round(ms, 3),
            "scenes_per_s": round(per_rank * world / (ms / 1e3), 2), "value": round(per_rank * world * n5 / (ms / 1e3), 1),
            "unit": "Gaussians/s", "h2d_bytes_per_scene": h2d, "h2d_GBps_per_rank": round(per_rank * h2d / (ms / 1e3) / 1e9, 2),
            "how": "end to end from pinned host memory through codec.ScenePipeline(result='psnr'), copies inside the timed region"}


def sharded_record(args, dev, rank, world, barrier, max_over_ranks, n_total=None):
    """BASELINE configs[3]: ONE scene over all ranks (strong scaling).  Every rank starts with a random 1/P sample of the
    scene; the timed step = the whole sharded encode + decode of raht_b200/sharded.py (NCCL exchanges included)."""
    import torch
    import torch.distributed as dist
    from raht_b200 import ops, sharded, synth
    from raht_b200.codec import StageTimer
    n, J, seed, desc = WORKLOADS["c4"]
    n = n_total or args.sharded_n or n
    n_local = n // world
    devt = list(synth.synthetic_gaussians_device(n_local, seed + rank, dev))
    codec_ = sharded.ShardedCodec(sharded.DeviceBackend(args.tile_rows))

    def step(inputs, timer=None):
        enc = codec_.encode(*inputs, J=J, step=STEP, timer=timer)
        rec = codec_.decode(enc, timer=timer)
        return enc, rec

    for _ in range(2):
        enc, rec = step(devt)
    barrier()
    launches0 = ops.lib.raht_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = max(min(args.steps, 3), 2)
    barrier()
    e0.record()
    timers = []
    for _ in range(steps):
        tm = StageTimer(True)
        enc, rec = step(devt, tm)
        timers.append(tm)
    e1.record()
    barrier()
    launches = (ops.lib.raht_launch_count() - launches0) // steps
    stages = {}
    for tm in timers:
        for k, v in tm.result().items():
            stages[k] = round(stages.get(k, 0.0) + v / len(timers), 3)
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / steps
    payload = torch.tensor([float(enc.streams.total_bytes())], dtype=torch.float64, device=dev)
    rows = torch.tensor([float(enc.n_rows[rank])], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(payload)
        dist.all_reduce(rows, op=dist.ReduceOp.MAX)
    nvox = enc.n_total
    limiting = max(stages, key=stages.get) if stages else None
    out = {"workload": desc, "scaling": "strong", "n_gaussians_total": n_local * world, "nvox": nvox, "depth": J, "quant_step": STEP,
           "ms_per_step": round(ms_step, 3), "value": round(n_local * world / (ms_step / 1e3), 1), "unit": "Gaussians/s",
           "gpu_launches_rank0": int(launches), "payload_bytes": int(payload.item()),
           "max_rows_on_a_rank": int(rows.item()), "balance": round(nvox / world / max(rows.item(), 1.0), 3),
           "stages_rank0_ms": stages, "limiting_stage_rank0": limiting,
           "collective_bytes_per_rank": getattr(enc, "collective_bytes", None),
           "parallelism": f"Morton-prefix shards x{world} (raht_b200/sharded.py): bitstream identical to one GPU"}
    del devt, enc, rec
    torch.cuda.empty_cache()
    return out


def run_sharded(args):
    """--workload c4 on its own: the sharded scene as the headline line (strong scaling)."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    bind_rank_to_cores(local, world)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        from raht_b200 import sharded
        sharded.tune_nccl_env()
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    clocks = ClockSampler(local)
    clocks.start()
    clocks.mark_begin()
    rec = sharded_record(args, dev, rank, world, barrier, max_over_ranks, n_total=args.n or None)
    clocks.mark_end()
    time.sleep(0.06)
    clk = clocks.stop()
    if rank == 0:
        line = {"metric": METRIC, "value": rec["value"], "unit": "Gaussians/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": {"workload": rec["workload"], "n_gaussians_total": rec["n_gaussians_total"], "nvox": rec["nvox"],
                                                "depth": rec["depth"], "channels": D, "quant_step": STEP, "parallelism": rec["parallelism"]},
                "clocks": clk, "gpu_launches": rec["gpu_launches_rank0"], "roofline": None, "cpu_baseline": None, "sharded": rec}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def reference_ply_frame(V, rgb, J, steps):
    """The reference's point-cloud frame (python/encode_ply.py:104-214) on torch-CPU with its own Python operators and its
    own coder object code (three channels, one host thread each)."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle_ctypes as orc
    from oracle import refpy
    ref = refpy.load()
    torch.set_num_threads(host_cores())
    with torch.no_grad():
        C = ref.utils.rgb_to_yuv(torch.from_numpy(rgb).to(torch.float64)).contiguous()
        Vt = torch.from_numpy(V.astype(np.float64))
        ListC, FlagsC, weightsC, order = ref.RAHT_param.RAHT_param_reorder_fast(Vt, torch.zeros(3, dtype=torch.float64), 2 ** J, J)
        Coeff, _ = ref.RAHT.RAHT2_optimized(C, ListC, FlagsC, weightsC)
        inv = torch.argsort(order)
        total = 0
        with ThreadPoolExecutor(max_workers=3) as pool:
            for step in steps:
                enc = torch.floor(Coeff / step + 0.5)
                cols = np.ascontiguousarray(enc.index_select(0, order).to(torch.int32).numpy().T)
                streams = list(pool.map(lambda c: orc.ref_rlgr_encode(cols[c].astype(np.int64), 1)[0], range(3)))
                dec = list(pool.map(lambda c: orc.ref_rlgr_decode(streams[c], cols.shape[1], 1)[0], range(3)))
                total += sum(len(b) for b in streams)
                Cd = (torch.from_numpy(np.stack(dec, axis=1).astype(np.int32)).to(torch.float64) * step)[inv, :]
                ref.iRAHT.inverse_RAHT_optimized(Cd, ListC, FlagsC, weightsC)
    return total


def run_c1(args):
    """BASELINE configs[0]: one frame of python/encode_ply.py per step -- RGB -> YUV, RAHT_param + RAHT, then the reference's
    11 quantisation steps, each with quantise / reorder / RLGR encode / RLGR decode / dequantise / iRAHT."""
    import torch
    from raht_b200 import codec, ops, synth
    n_target, J, seed, desc = WORKLOADS["c1"]
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if int(os.environ.get("RANK", "0")) != 0:
        return
    V, rgb = synth.sphere_point_cloud(seed=seed)
    Vd, rgbd = torch.from_numpy(V.astype(np.float64)).to(dev), torch.from_numpy(rgb).to(dev)
    nvox = V.shape[0]
    steps = codec.PLY_COLOR_STEPS
    clocks = ClockSampler(dev.index)
    clocks.start()

    def measure(fused):
        for _ in range(args.warmup):
            res, _ = codec.encode_ply_frame(Vd, rgbd, J, fused=fused)
        torch.cuda.synchronize()
        launches0 = ops.lib.raht_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        timers = []
        e0.record()
        for _ in range(args.steps):
            tm = codec.StageTimer(True)
            res, _ = codec.encode_ply_frame(Vd, rgbd, J, fused=fused, timer=tm)
            timers.append(tm)
        e1.record()
        torch.cuda.synchronize()
        stages = {}
        for tm in timers:
            for k, v in tm.result().items():
                stages[k] = stages.get(k, 0.0) + v / len(timers)
        return e0.elapsed_time(e1) / args.steps, stages, (ops.lib.raht_launch_count() - launches0) // args.steps, res

    clocks.mark_begin()
    ms, stages, launches, res = measure(False)
    ms_fused, stages_fused, _, res_fused = measure(True)
    clocks.mark_end()
    clk = clocks.stop()
    cpu = None
    if not args.no_cpu_baseline:
        from oracle import refpy
        if refpy.available():
            t0 = time.perf_counter()
            total = reference_ply_frame(V, rgb, J, steps)
            secs = time.perf_counter() - t0
            cpu = {"value": round(nvox * len(steps) / secs, 1), "unit": "voxel-steps/s", "cores": host_cores(), "kind": "reference",
                   "sample": f"the whole frame once ({secs:.1f} s): the reference's Python operators on torch-CPU + its RLGR object code",
                   "payload_bytes_reference_cpu": int(total), "payload_bytes_gpu": int(sum(r["bytes"] for r in res)),
                   "note": "the unmodified reference on torch-CPU differs from the bit-exact convention in the last bits of T (torch-CPU's "
                           "vectorised sqrt is not correctly rounded, its matmul order is the BLAS's): the payloads agree to a few bytes, the "
                           "bit-exact comparison is tests/test_gpu_scale_parity.py::test_c1_point_cloud_frame_is_bit_exact_with_the_oracle"}
    peak, peak_src = load_peaks()
    raht_bytes = nvox * (2 * 3 * 8 + 13)
    line = {
        "metric": "voxel-steps/s RAHT+RLGR encode+decode of an RGB point cloud (3 ch, 11 quantisation steps per frame)",
        "value": round(nvox * len(steps) / (ms / 1e3), 1), "unit": "voxel-steps/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "nvox": nvox, "depth": J, "channels": 3, "quant_steps": steps,
                   "rates_bpp": [round(r["bpp"], 4) for r in res], "psnr_y": [round(r["psnr_y"], 3) for r in res]},
        "clocks": clk, "gpu_launches": int(launches),
        "stages": {k: round(v, 4) for k, v in stages.items()},
        "fused_transform_quantisation": {"ms_per_step": round(ms_fused, 4), "stages": {k: round(v, 4) for k, v in stages_fused.items()},
                                         "bytes_equal": [a["bytes"] for a in res] == [b["bytes"] for b in res_fused],
                                         "verdict": "fused pays" if ms_fused < ms else "two passes are faster"},
        "roofline": {"bound": "hbm", "kernel": "raht", "achieved": round(raht_bytes / (stages.get("raht", 1e9) / 1e3) / 1e9, 2), "peak": peak,
                     "unit": "GB/s", "frac": round(raht_bytes / (stages.get("raht", 1e9) / 1e3) / 1e9 / peak, 4), "traffic": None,
                     "peak_source": peak_src},
        "cpu_baseline": cpu,
    }
    print(json.dumps(line))


def run_reference(args):
    """--impl reference: the reference's own CPU path on the box's host cores, on the WHOLE configured scene
    (same_config: the scene, depth, channels and step of the b200 arm), every step one full encode + decode."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle_ctypes as orc
    from oracle import refpy
    from raht_b200 import synth
    if args.workload == "c1":
        from raht_b200.codec import PLY_COLOR_STEPS
        _, J, seed, desc = WORKLOADS["c1"]
        V, rgb = synth.sphere_point_cloud(seed=seed)
        for _ in range(args.warmup):
            reference_ply_frame(V, rgb, J, PLY_COLOR_STEPS)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            reference_ply_frame(V, rgb, J, PLY_COLOR_STEPS)
        secs = (time.perf_counter() - t0) / max(args.steps, 1)
        v = V.shape[0] * len(PLY_COLOR_STEPS) / secs
        print(json.dumps({"impl": "reference", "metric": "voxel-steps/s RAHT+RLGR encode+decode of an RGB point cloud (3 ch, 11 quantisation steps per frame)",
                          "value": round(v, 1), "unit": "voxel-steps/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": round(secs * 1e3, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                          "data": "synthetic", "config": {"workload": desc, "nvox": int(V.shape[0]), "depth": J, "channels": 3, "same_config": True},
                          "cpu_baseline": {"value": round(v, 1), "unit": "voxel-steps/s", "cores": host_cores(), "kind": "reference",
                                           "sample": "the whole frame per step: the reference's Python operators on torch-CPU + its RLGR object code"},
                          "e2e": {"value": round(v, 1), "unit": "voxel-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    n, J, seed, desc = WORKLOADS[args.workload if args.workload in ("c2", "c3", "tiny") else "c2"]
    if args.n:
        n = args.n
    if args.cpu_sample:
        n = min(n, args.cpu_sample)
    scene = synth.synthetic_gaussians(WORKLOADS[args.workload][0] if not args.n else args.n, seed)
    scene = tuple(a[:n] for a in scene)
    use_python = refpy.available() and orc.ref_rlgr_available() and not args.port
    stage_s = {}
    if use_python:
        def one():
            return reference_python_pipeline(scene, J, STEP, stage_s)[0]
    else:
        use_ref = orc.ref_rlgr_available() and not args.port_rlgr
        def one():
            return cpu_pipeline(scene, J, STEP, use_ref)[0]
    for _ in range(args.warmup):
        one()
    stage_s.clear()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        nvox = one()
    secs = (time.perf_counter() - t0) / max(args.steps, 1)
    v = n / secs
    # the OpenMP port as a second, labelled number (one step)
    t0 = time.perf_counter()
    cpu_pipeline(scene, J, STEP, orc.ref_rlgr_available())
    port_secs = time.perf_counter() - t0
    if use_python:
        kind = "reference"
        sample = (f"the whole scene ({n} Gaussians) per step: the reference's Python operators ({os.path.relpath(refpy.directory(), ROOT) if refpy.directory().startswith(ROOT) else refpy.directory()}) "
                  "on torch-CPU in the order of test_voxelize_3dgs.py:203-247 + encode_3dgs.py:144-276, RLGR by the reference coder's object code "
                  "(oracle/_ref/librlgr_ref.so), one channel per host thread; merge = C restatement (the reference has only a CUDA kernel)")
    else:
        kind = "port"
        sample = f"the whole scene ({n} Gaussians) per step: oracle/raht_oracle.c (OpenMP)" + (" + reference RLGR coder oracle/_ref/librlgr_ref.so" if orc.ref_rlgr_available() and not args.port_rlgr else "")
    cpu = {"value": round(v, 1), "unit": "Gaussians/s", "cores": host_cores(), "kind": kind, "sample": sample,
           "stages_s": {k: round(x / max(args.steps, 1), 3) for k, x in stage_s.items()} or None,
           "port": {"value": round(n / port_secs, 1), "unit": "Gaussians/s", "kind": "port",
                    "sample": f"the same scene once through oracle/raht_oracle.c with OpenMP ({port_secs:.1f} s)"}}
    full = WORKLOADS[args.workload][0] if not args.n else args.n
    line = {
        "impl": "reference", "metric": METRIC, "value": round(v, 1), "unit": "Gaussians/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(secs * 1e3, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "n_gaussians_per_gpu": n, "nvox": int(nvox), "depth": J, "channels": D, "quant_step": STEP,
                   "same_config": bool(n == full)},
        "cpu_baseline": cpu,
        "e2e": {"value": round(v, 1), "unit": "Gaussians/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--num-gaussians", dest="n", type=int, default=0, help="override the number of Gaussians")
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"], help="transform arithmetic (f64 = bit-exact mode)")
    ap.add_argument("--tile-rows", type=int, default=0)
    ap.add_argument("--cpu-sample", type=int, default=0, help="reference arm: only the first K Gaussians of the scene (0 = the whole scene)")
    ap.add_argument("--fused", action="store_true", help="transform and quantisation in one pass (raht_forward_quant / raht_dequant_inverse) instead of the reference's two")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-subrecords", action="store_true", help="only the headline workload (no sparse / c3 / c5 / sharded sub-records)")
    ap.add_argument("--sharded-n", type=int, default=0, help="Gaussians of the sharded sub-record's scene (default: 50 M)")
    ap.add_argument("--port", action="store_true", help="reference arm: the OpenMP port instead of the reference's Python operators")
    ap.add_argument("--port-rlgr", action="store_true", help="reference arm (port): the oracle RLGR instead of oracle/_ref")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    args.tile_rows = args.tile_rows or None
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "c4":
        run_sharded(args)
    elif args.workload == "c1":
        run_c1(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
