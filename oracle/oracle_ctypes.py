"""ctypes access to the CPU ORACLE (oracle/raht_oracle.c) and, when built, to the reference's own
RLGR coder (oracle/_ref/librlgr_ref.so).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs -- never by the product
package (raht-3dgs-codec_b200/), which has no CPU path.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
_REF = os.path.join(_HERE, "_ref", "librlgr_ref.so")


def build(quiet=True):
    """Compile the oracle (and the reference RLGR shim when /root/reference is present)."""
    subprocess.run(["make", "-C", _HERE, "all"], check=True, capture_output=quiet)


def _load():
    if not os.path.exists(_LIB):
        build()
    return C.CDLL(_LIB)


lib = _load()
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C")

lib.orc_morton.argtypes = [_i64p, C.c_int64, C.c_int, _i64p]
lib.orc_morton_from_f64.argtypes = [_f64p, C.c_int64, _f64p, C.c_double, C.c_int, _i64p]
lib.orc_morton_from_f64.restype = C.c_int
lib.orc_voxelize.restype = C.c_int64
lib.orc_voxelize.argtypes = [_f32p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, _i64p, _i64p, _i64p, _i64p, _i64p,
                             C.c_void_p, _f32p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
lib.orc_voxel_attr_mean.argtypes = [_f32p, C.c_int64, C.c_int, _i64p, _i64p, C.c_int64, _f32p]
lib.orc_merge_clusters.argtypes = [_i32p, _i32p, C.c_int64, _f32p, _f32p, _f32p, _f32p, _f32p, C.c_int, C.c_int,
                                   _f32p, _f32p, _f32p, _f32p, _f32p]
lib.orc_tree_build.restype = C.c_void_p
lib.orc_tree_build.argtypes = [_i64p, C.c_int64, C.c_int]
lib.orc_tree_free.argtypes = [C.c_void_p]
lib.orc_tree_nlevels.argtypes = [C.c_void_p]
lib.orc_tree_nlists.argtypes = [C.c_void_p]
lib.orc_tree_level_len.restype = C.c_int64
lib.orc_tree_level_len.argtypes = [C.c_void_p, C.c_int]
lib.orc_tree_level_copy.argtypes = [C.c_void_p, C.c_int, _i64p, _u8p, _i64p]
lib.orc_tree_order_len.restype = C.c_int64
lib.orc_tree_order_len.argtypes = [C.c_void_p]
lib.orc_tree_order_copy.argtypes = [C.c_void_p, _i64p]
lib.orc_raht_forward.argtypes = [C.c_void_p, _f64p, C.c_int, _f64p, _f64p]
lib.orc_raht_inverse.argtypes = [C.c_void_p, _f64p, C.c_int, _f64p]
lib.orc_quantize.argtypes = [_f64p, C.c_int64, C.c_double, _f64p]
lib.orc_dequantize.argtypes = [_f64p, C.c_int64, C.c_double, _f64p]
lib.orc_rlgr_encode.restype = C.c_int64
lib.orc_rlgr_encode.argtypes = [_i64p, C.c_int64, C.c_int, _u8p, C.c_int64]
lib.orc_rlgr_decode.argtypes = [_u8p, C.c_int64, C.c_int64, C.c_int, _i64p]
lib.orc_rlgr_encode_channels.argtypes = [_i32p, C.c_int64, C.c_int64, C.c_int, C.c_int, _u8p, C.c_int64, _i64p]
lib.orc_rlgr_decode_channels.argtypes = [_u8p, C.c_int64, _i64p, C.c_int64, C.c_int64, C.c_int, C.c_int, _i32p]
lib.orc_rgb_to_yuv.argtypes = [_f64p, C.c_int64, _f64p]
lib.orc_num_threads.restype = C.c_int
lib.orc_set_num_threads.argtypes = [C.c_int]


def num_threads():
    return lib.orc_num_threads()


def set_num_threads(t):
    lib.orc_set_num_threads(int(t))


def morton(V, J):
    V = np.ascontiguousarray(V, dtype=np.int64)
    out = np.empty(V.shape[0], dtype=np.int64)
    lib.orc_morton(V, V.shape[0], J, out)
    return out


def morton_from_f64(V, minV, width, depth):
    V = np.ascontiguousarray(V, dtype=np.float64)
    out = np.empty(V.shape[0], dtype=np.int64)
    rc = lib.orc_morton_from_f64(V, V.shape[0], np.ascontiguousarray(minV, dtype=np.float64), float(width), depth, out)
    return out, rc


def voxelize(V, J, vmin=None, width=None, want_delta=True):
    """voxelize_pc_batched on positions only.  Returns dict."""
    V = np.ascontiguousarray(V, dtype=np.float32)
    n = V.shape[0]
    Vint = np.empty((n, 3), dtype=np.int64)
    MC = np.empty(n, dtype=np.int64)
    MCs = np.empty(n, dtype=np.int64)
    idx = np.empty(n, dtype=np.int64)
    vi = np.empty(max(n, 1), dtype=np.int64)
    delta = np.empty((n, 3), dtype=np.float32) if want_delta else None
    vmin_out = np.empty(3, dtype=np.float32)
    w_out, vs_out = C.c_double(), C.c_double()
    vmin_arr = np.ascontiguousarray(vmin, dtype=np.float32) if vmin is not None else None
    w_in = C.c_double(width) if width is not None else None
    nvox = lib.orc_voxelize(V, n, J, vmin_arr.ctypes.data if vmin_arr is not None else None,
                            C.addressof(w_in) if w_in is not None else None, Vint, MC, MCs, idx, vi,
                            delta.ctypes.data if delta is not None else None, vmin_out, C.byref(w_out), C.byref(vs_out))
    return {"Vint": Vint, "MC": MC, "MC_sorted": MCs, "sort_idx": idx, "voxel_indices": vi[:nvox].copy(), "Nvox": int(nvox),
            "DeltaV": delta, "vmin": vmin_out, "width": w_out.value, "voxel_size": vs_out.value}


def voxel_attr_mean(Cattr, sort_idx, voxel_indices):
    Cattr = np.ascontiguousarray(Cattr, dtype=np.float32)
    n, d = Cattr.shape
    out = np.empty((len(voxel_indices), d), dtype=np.float32)
    lib.orc_voxel_attr_mean(Cattr, n, d, np.ascontiguousarray(sort_idx, dtype=np.int64),
                            np.ascontiguousarray(voxel_indices, dtype=np.int64), len(voxel_indices), out)
    return out


def merge_clusters(cidx, coff, means, quats, scales, opac, colors, by_opacity=True):
    nclu = len(coff) - 1
    cd = colors.shape[1]
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    mm, mq, ms = np.zeros((nclu, 3), np.float32), np.zeros((nclu, 4), np.float32), np.zeros((nclu, 3), np.float32)
    mo, mc = np.zeros(nclu, np.float32), np.zeros((nclu, cd), np.float32)
    lib.orc_merge_clusters(np.ascontiguousarray(cidx, dtype=np.int32), np.ascontiguousarray(coff, dtype=np.int32), nclu,
                           f(means), f(quats), f(scales), f(opac), f(colors), cd, 1 if by_opacity else 0, mm, mq, ms, mo, mc)
    return mm, mq, ms, mo, mc


class Tree:
    """RAHT_param_reorder_fast on Morton-sorted unique codes (reference control flow)."""

    def __init__(self, MC, depth):
        self.MC = np.ascontiguousarray(MC, dtype=np.int64)
        self.n = self.MC.shape[0]
        self.h = lib.orc_tree_build(self.MC, self.n, depth)

    def __del__(self):
        if getattr(self, "h", None):
            lib.orc_tree_free(self.h)
            self.h = None

    @property
    def nlevels(self):
        return lib.orc_tree_nlevels(self.h)

    @property
    def nlists(self):
        return lib.orc_tree_nlists(self.h)

    def level(self, l):
        L = lib.orc_tree_level_len(self.h, l)
        lst, fl, w = np.empty(L, np.int64), np.empty(L, np.uint8), np.empty(L, np.int64)
        lib.orc_tree_level_copy(self.h, l, lst, fl, w)
        return lst, fl.astype(bool), w

    def order(self):
        L = lib.orc_tree_order_len(self.h)
        if L < 0:
            return None
        o = np.empty(L, np.int64)
        lib.orc_tree_order_copy(self.h, o)
        return o

    def forward(self, Cmat):
        Cmat = np.ascontiguousarray(Cmat, dtype=np.float64)
        T = np.empty_like(Cmat)
        w = np.empty(self.n, dtype=np.float64)
        lib.orc_raht_forward(self.h, Cmat, Cmat.shape[1], T, w)
        return T, w

    def inverse(self, T):
        T = np.ascontiguousarray(T, dtype=np.float64)
        Cm = np.empty_like(T)
        lib.orc_raht_inverse(self.h, T, T.shape[1], Cm)
        return Cm


def quantize(T, step):
    T = np.ascontiguousarray(T, dtype=np.float64)
    q = np.empty_like(T)
    lib.orc_quantize(T.reshape(-1), T.size, float(step), q.reshape(-1))
    return q


def dequantize(q, step):
    q = np.ascontiguousarray(q, dtype=np.float64)
    T = np.empty_like(q)
    lib.orc_dequantize(q.reshape(-1), q.size, float(step), T.reshape(-1))
    return T


def rlgr_encode(seq, flag_signed=1):
    seq = np.ascontiguousarray(seq, dtype=np.int64)
    cap = max(64, seq.size * 9 + 64)
    out = np.zeros(cap, dtype=np.uint8)
    nb = lib.orc_rlgr_encode(seq, seq.size, flag_signed, out, cap)
    assert nb <= cap
    return out[:nb].tobytes()


def rlgr_decode(buf, N, flag_signed=1):
    b = np.frombuffer(bytes(buf) + b"\0" * 8, dtype=np.uint8).copy()
    seq = np.zeros(N, dtype=np.int64)
    lib.orc_rlgr_decode(b, len(buf), N, flag_signed, seq)
    return seq


def rlgr_encode_channels(q, flag_signed=1, cap=None):
    """q: (nch, n) int32 -> (list of bytes per channel)."""
    q = np.ascontiguousarray(q, dtype=np.int32)
    nch, n = q.shape
    cap = cap or (n * 5 + 64)
    out = np.zeros(nch * cap, dtype=np.uint8)
    nb = np.zeros(nch, dtype=np.int64)
    lib.orc_rlgr_encode_channels(q, n, n, nch, flag_signed, out, cap, nb)
    if (nb > cap).any():
        return rlgr_encode_channels(q, flag_signed, int(nb.max()) + 64)
    return [out[c * cap: c * cap + nb[c]].tobytes() for c in range(nch)], out, cap, nb


def rlgr_decode_channels(out, cap, nb, n, nch, flag_signed=1):
    q = np.zeros((nch, n), dtype=np.int32)
    lib.orc_rlgr_decode_channels(out, cap, nb, n, n, nch, flag_signed, q)
    return q


def rgb_to_yuv(rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float64)
    out = np.empty_like(rgb)
    lib.orc_rgb_to_yuv(rgb, rgb.shape[0], out)
    return out


# ---- the reference's own RLGR coder (compiled from /root/reference by oracle/Makefile) ---------------
def ref_rlgr_available():
    return os.path.exists(_REF)


_ref = None


def _ref_lib():
    global _ref
    if _ref is None:
        _ref = C.CDLL(_REF)
        _ref.ref_rlgr_encode.restype = C.c_longlong
        _ref.ref_rlgr_encode.argtypes = [_i64p, C.c_longlong, C.c_int, _u8p, C.c_longlong, C.POINTER(C.c_longlong)]
        _ref.ref_rlgr_decode.argtypes = [_u8p, C.c_longlong, C.c_longlong, C.c_int, _i64p, C.POINTER(C.c_longlong)]
    return _ref


def ref_rlgr_encode(seq, flag_signed=1):
    seq = np.ascontiguousarray(seq, dtype=np.int64)
    cap = max(64, seq.size * 9 + 64)
    out = np.zeros(cap, dtype=np.uint8)
    ns = C.c_longlong()
    nb = _ref_lib().ref_rlgr_encode(seq, seq.size, flag_signed, out, cap, C.byref(ns))
    return out[:nb].tobytes(), ns.value


def ref_rlgr_decode(buf, N, flag_signed=1):
    b = np.frombuffer(bytes(buf) + b"\0" * 8, dtype=np.uint8).copy()
    seq = np.zeros(N, dtype=np.int64)
    ns = C.c_longlong()
    _ref_lib().ref_rlgr_decode(b, len(buf), N, flag_signed, seq, C.byref(ns))
    return seq, ns.value


# ---- per-attribute quantisation steps (python/encode_3dgs_debug.py:326-381), numpy restatement ---------------------------
IMPORTANCE_56 = {"quats": 1.0 / 21.93, "scales": 1.0 / 26.36, "opacity": 1.0 / 42.22, "colors": 1.0 / 38.67}   # :336-341


def per_attribute_steps(Coeff, attr_ranges, importance_weights, total_levels_budget=1024):
    """Coeff (N, D) float64 -> ({name: (step, levels, range)}, Coeff_enc) exactly as the reference loop :348-381."""
    n_channels = Coeff.shape[1]
    total_importance = sum(importance_weights.values())
    per, enc = {}, np.zeros_like(Coeff)
    for name, (a, b) in attr_ranges.items():
        if a >= n_channels:
            continue
        coeff_attr = Coeff[:, a:b]
        range_attr = coeff_attr.max() - coeff_attr.min()
        levels = max(int(total_levels_budget * importance_weights[name] / total_importance), 2)
        step = max(range_attr / max(levels - 1, 1), 1e-6)
        per[name] = (float(step), levels, float(range_attr))
        enc[:, a:b] = np.floor(Coeff[:, a:b] / step + 0.5)
    return per, enc


# ---- the reference's own merge kernel (cuda/merge_cluster.cu, compiled in place for sm_100 by oracle/Makefile) ----------
def ref_merge_path(kind="ieee"):
    """kind: 'ieee' (--fmad=false), 'default' (nvcc defaults: FMA contraction), 'fast' (--use_fast_math, as shipped)."""
    return os.path.join(_HERE, "_ref", f"libmerge_ref_{kind}.so")


def ref_merge_available(kind="ieee"):
    return os.path.exists(ref_merge_path(kind))


_merge_libs = {}


def ref_merge_launch(kind, cidx_ptr, coff_ptr, nclu, means_ptr, quats_ptr, scales_ptr, opac_ptr, colors_ptr, color_dim, by_opacity,
                     out_ptrs, stream_ptr=0):
    """Launch merge_weighted_mean_kernel on DEVICE pointers (ints, e.g. tensor.data_ptr()); the five outputs must be
    zero-filled by the caller (merge_cluster_wrapper.cu:64-69).  Returns the CUDA error code of the launch."""
    lib_ = _merge_libs.get(kind)
    if lib_ is None:
        lib_ = _merge_libs[kind] = C.CDLL(ref_merge_path(kind))
        lib_.ref_merge_launch.restype = C.c_int
        lib_.ref_merge_launch.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_int] + [C.c_void_p] * 6
    return lib_.ref_merge_launch(cidx_ptr, coff_ptr, int(nclu), means_ptr, quats_ptr, scales_ptr, opac_ptr, colors_ptr, int(color_dim),
                                 1 if by_opacity else 0, *out_ptrs, stream_ptr)
