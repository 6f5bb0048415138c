"""ctypes binding of libgsrast.so (include/gsrast.h).

The product path has NO fallback: if the CUDA library is missing or fails to
load, importing this module raises.  Nothing here imports the oracle.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSR_LIB_PATH: development override to A/B another build of the same library
LIB_PATH = os.environ.get("GSR_LIB_PATH") or os.path.join(os.path.dirname(_HERE), "lib", "libgsrast.so")


class gsr_inputs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int32), ("D", ctypes.c_int32), ("M", ctypes.c_int32),
        ("W", ctypes.c_int32), ("H", ctypes.c_int32),
        ("tanfovx", ctypes.c_float), ("tanfovy", ctypes.c_float), ("scale_modifier", ctypes.c_float),
        ("prefiltered", ctypes.c_int32), ("debug", ctypes.c_int32),
        ("bg", ctypes.c_void_p), ("means3D", ctypes.c_void_p), ("sh", ctypes.c_void_p),
        ("colors_precomp", ctypes.c_void_p), ("opacities", ctypes.c_void_p), ("scales", ctypes.c_void_p),
        ("rotations", ctypes.c_void_p), ("cov3D_precomp", ctypes.c_void_p), ("viewmatrix", ctypes.c_void_p),
        ("projmatrix", ctypes.c_void_p), ("campos", ctypes.c_void_p), ("sh_rest", ctypes.c_void_p),
        ("raw_params", ctypes.c_int32),
    ]


class gsr_views(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in (
        "xy_depth_radius", "conic_opacity", "rgb_flags", "final_T", "n_contrib", "ranges", "tile_count",
        "point_list")]


# every symbol include/gsrast.h declares
EXPORTS = (
    "gsr_version", "gsr_compiled_arch", "gsr_last_error", "gsr_scratch_bytes", "gsr_forward_pre",
    "gsr_forward_post", "gsr_backward", "gsr_mark_visible", "gsr_get_views", "gsr_profile_enable", "gsr_profile_read",
    "gsr_sh_color_factors", "gsr_sh_grad_from_factors", "gsr_knn_scratch_bytes", "gsr_knn_dist2",
    "gsr_loss_scratch_bytes", "gsr_l1_ssim_loss", "gsr_l1_ssim_backward",
    "gsr_backward_blend", "gsr_backward_geometry", "gsr_backward_sh_factors",
    "gsr_adam_step", "gsr_forward", "gsr_forward_capacity", "gsr_backward_geometry_split",
    "gsr_activate", "gsr_activate_backward", "gsr_densification_stats", "gsr_launch_count", "gsr_set_option",
    "gsr_sort_scratch_bytes", "gsr_sort_pairs", "gsr_peer_allreduce", "gsr_sh_grad_from_peers", "gsr_peer_push",
    "gsr_compact_scratch_bytes", "gsr_compact_indices", "gsr_compact_indices_noalloc", "gsr_gather_rows",
    "gsr_selftest_expf",
)
STAGES = ("preprocess_fwd", "depth_sort", "bin_count", "bin_scatter", "render_fwd", "render_bwd", "preprocess_bwd")


def load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(
            f"libgsrast.so not found at {path}: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"or `make -C gaussian-transformer_b200/csrc` (there is no CPU fallback)")
    lib = ctypes.CDLL(path)
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64
    szp = ctypes.POINTER(ctypes.c_size_t)
    inp = ctypes.POINTER(gsr_inputs)
    lib.gsr_version.restype = ctypes.c_int
    lib.gsr_compiled_arch.restype = ctypes.c_int
    lib.gsr_last_error.restype = ctypes.c_char_p
    lib.gsr_scratch_bytes.argtypes = [i32, i64, i32, i32, szp, szp, szp, szp]
    lib.gsr_forward_pre.argtypes = [inp, vp, vp, vp, vp, vp]
    lib.gsr_forward_post.argtypes = [inp, vp, vp, vp, i64, vp, vp]
    lib.gsr_forward.argtypes = [inp, vp, vp, i64, vp, vp, vp, vp, vp]
    lib.gsr_forward_capacity.argtypes = [inp, vp, vp, i64, vp, vp, vp, vp, vp]
    lib.gsr_backward.argtypes = [inp, vp, vp, vp, vp, i64, vp] + [vp] * 9 + [vp, vp]
    lib.gsr_backward_blend.argtypes = [inp, vp, vp, vp, i64, vp, vp, vp]
    lib.gsr_backward_geometry.argtypes = [inp, vp, vp] + [vp] * 9 + [vp, vp]
    lib.gsr_backward_geometry_split.argtypes = [inp, vp, vp] + [vp] * 10 + [vp, vp]
    lib.gsr_backward_sh_factors.argtypes = [i32, vp, vp, vp, vp]
    lib.gsr_adam_step.argtypes = [i32, vp, ctypes.c_double, ctypes.c_double, ctypes.c_double, vp]
    lib.gsr_activate.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp]
    lib.gsr_activate_backward.argtypes = [i32] + [vp] * 9 + [vp]
    lib.gsr_densification_stats.argtypes = [i32, vp, vp, vp, vp, vp, vp]
    lib.gsr_mark_visible.argtypes = [i32, vp, vp, vp, vp, vp]
    lib.gsr_loss_scratch_bytes.argtypes = [i32, i32, i32, szp]
    lib.gsr_l1_ssim_loss.argtypes = [i32, i32, i32, vp, vp, ctypes.c_float, vp, vp, vp, vp]
    lib.gsr_l1_ssim_backward.argtypes = [i32, i32, i32, vp, vp, ctypes.c_float, vp, vp, vp, vp]
    lib.gsr_knn_scratch_bytes.argtypes = [i32, szp]
    lib.gsr_knn_dist2.argtypes = [i32, vp, vp, vp, vp]
    lib.gsr_sh_color_factors.argtypes = [i32, vp, vp, vp, vp]
    lib.gsr_sh_grad_from_factors.argtypes = [i32, i32, i32, i32, vp, vp, vp, vp, i32, vp]
    lib.gsr_get_views.argtypes = [i32, i64, i32, i32, vp, vp, vp, ctypes.POINTER(gsr_views)]
    lib.gsr_selftest_expf.argtypes = [ctypes.c_uint32, ctypes.c_uint32, vp, vp]
    lib.gsr_compact_scratch_bytes.argtypes = [i32, szp]
    lib.gsr_compact_indices.argtypes = [i32, vp, vp, i32, vp, vp, vp]
    lib.gsr_compact_indices_noalloc.argtypes = [i32, vp, vp, i32, vp, vp]
    lib.gsr_gather_rows.argtypes = [i32, vp, i64, vp, vp]
    u64p = ctypes.POINTER(ctypes.c_uint64)
    lib.gsr_peer_allreduce.argtypes = [i32, i32, u64p, ctypes.c_uint64, i64, vp]
    lib.gsr_sh_grad_from_peers.argtypes = [i32, i32, i32, i32, vp, u64p, vp, i32, vp]
    lib.gsr_peer_push.argtypes = [i32, i32, u64p, vp, i64, i64, i32, vp]
    lib.gsr_sort_scratch_bytes.argtypes = [i32, szp]
    lib.gsr_sort_pairs.argtypes = [i32, vp, vp, vp, vp, vp, i32, vp]
    lib.gsr_launch_count.argtypes = [ctypes.c_int]
    lib.gsr_launch_count.restype = ctypes.c_int64
    lib.gsr_set_option.argtypes = [ctypes.c_char_p, i32]
    lib.gsr_profile_enable.argtypes = [ctypes.c_int]
    lib.gsr_profile_read.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32)]
    for name in ("gsr_profile_enable", "gsr_profile_read", "gsr_scratch_bytes", "gsr_forward_pre", "gsr_forward_post", "gsr_backward", "gsr_mark_visible",
                 "gsr_get_views"):
        getattr(lib, name).restype = ctypes.c_int
    return lib


lib = load()


def check(rc):
    if rc != 0:
        msg = lib.gsr_last_error()
        raise RuntimeError("gsrast: " + (msg.decode() if msg else f"error {rc}"))


def scratch_bytes(P, R, W, H):
    g, b, i, w = (ctypes.c_size_t() for _ in range(4))
    check(lib.gsr_scratch_bytes(P, R, W, H, ctypes.byref(g), ctypes.byref(b), ctypes.byref(i), ctypes.byref(w)))
    return g.value, b.value, i.value, w.value


def set_option(name, value):
    """Development / test switch of the library (include/gsrast.h gsr_set_option)."""
    check(lib.gsr_set_option(name.encode(), int(value)))


def launch_count(reset=False):
    """Kernels launched by the library in this process since the last reset."""
    return int(lib.gsr_launch_count(int(reset)))


def profile_enable(on=True):
    check(lib.gsr_profile_enable(int(on)))


def profile_read():
    """-> {stage: (total_ms, launches)} accumulated on this thread since the previous read."""
    ms = (ctypes.c_float * len(STAGES))()
    n = (ctypes.c_int32 * len(STAGES))()
    check(lib.gsr_profile_read(ms, n))
    return {s: (float(ms[i]), int(n[i])) for i, s in enumerate(STAGES)}
