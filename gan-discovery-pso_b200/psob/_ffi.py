"""ctypes binding of libpsob.so (C ABI in include/psob.h).

There is no fallback: if the CUDA library is missing or a call fails, a PsobError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpsob.so")

EXPORTS = [
    "psob_version", "psob_last_error", "psob_device_info", "psob_update", "psob_pack_latent",
    "psob_pbest_minloc", "psob_gbest_candidate", "psob_gbest_apply", "psob_score_fitness",
    "psob_softmax_select", "psob_recon_fitness", "psob_rescale_minmax", "psob_pack_image",
    "psob_pairwise_distance_sum", "psob_vq_nearest", "psob_pack_image_stem", "psob_maxpool_nhwc",
    "psob_layer_create", "psob_layer_destroy", "psob_layer_run", "psob_layer_run_residual", "psob_layer_profile",
    "psob_layer_dot_slots", "psob_layer_run_dot", "psob_dot_finish", "psob_fill_columns_bf16",
    "psob_fitness_tail", "psob_gbest_local", "psob_gbest_record", "psob_layer_recon_slots", "psob_layer_run_recon",
    "psob_pack_grid_nchw", "psob_gather_pack_latent", "psob_layer_run_dot_static",
]
STATIC_SCRATCH_BYTES = 8 << 20
VERSION = 200

MODE_IN_TRAINING, MODE_OUT_TRAINING = 0, 1
IMPL_TCGEN05, IMPL_SIMT = 0, 1


class PsobError(RuntimeError):
    pass


class IterState(C.Structure):
    _fields_ = [("iteration", C.c_uint32), ("w", C.c_float)]


class UpdateArgs(C.Structure):
    _fields_ = [
        ("pos", C.c_void_p), ("vel", C.c_void_p), ("pbest_pos", C.c_void_p), ("gbest_pos", C.c_void_p),
        ("improved", C.c_void_p), ("r1", C.c_void_p), ("r2", C.c_void_p), ("z_bf16", C.c_void_p),
        ("dyn", C.c_void_p),
        ("P", C.c_int64), ("D", C.c_int32), ("z_pitch", C.c_int32),
        ("w", C.c_float), ("c_cogn", C.c_float), ("c_soci", C.c_float),
        ("v_max", C.c_float), ("lo", C.c_float), ("hi", C.c_float),
        ("seed", C.c_uint64), ("index_offset", C.c_uint64),
        ("iteration", C.c_uint32), ("reserved", C.c_uint32),
    ]


class FitnessTailArgs(C.Structure):
    _fields_ = [
        ("part", C.c_void_p), ("n_slots", C.c_int32), ("bias", C.c_float), ("act", C.c_int32),
        ("score", C.c_void_p), ("score_stride", C.c_int32),
        ("logits", C.c_void_p), ("n_class", C.c_int32), ("class_idx", C.c_int32),
        ("fit_in", C.c_void_p),
        ("mode", C.c_int32), ("thr", C.c_float), ("eps", C.c_float),
        ("recon_part", C.c_void_p), ("n_recon_slots", C.c_int32), ("n_pix", C.c_int32),
        ("w_ass", C.c_float), ("w_rec", C.c_float), ("eps_rec", C.c_float),
        ("fit", C.c_void_p), ("pbest_val", C.c_void_p), ("improved", C.c_void_p),
        ("P", C.c_int64), ("index_offset", C.c_uint64), ("key", C.c_void_p),
    ]


class LayerDesc(C.Structure):
    _fields_ = [
        ("mma_n", C.c_int32), ("nt", C.c_int32), ("cc", C.c_int32), ("act", C.c_int32),
        ("in_pitch", C.c_int32), ("out_pitch", C.c_int32), ("out32_pitch", C.c_int32),
        ("b_resident", C.c_int32), ("vec8", C.c_int32),
        ("n_steps", C.c_int32), ("n_jobs", C.c_int32), ("n_wrows", C.c_int32), ("n_bias_rows", C.c_int32),
        ("n_chains", C.c_int32),
        ("in_tiled", C.c_int32), ("out_tiled", C.c_int32), ("band_pitch", C.c_int32), ("dot_head", C.c_int32), ("solo", C.c_int32), ("band_width", C.c_int32),
        ("wpack", C.c_void_p), ("steps", C.c_void_p), ("jobs", C.c_void_p), ("chains", C.c_void_p),
        ("col_map", C.c_void_p), ("col_map32", C.c_void_p), ("bias", C.c_void_p),
    ]


_lib = None


def lib():
    """Load libpsob.so (built by `__graft_entry__.build()` / csrc/build.sh)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PsobError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    l = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float
    sig = {
        "psob_version": ([], C.c_int),
        "psob_last_error": ([], C.c_char_p),
        "psob_device_info": ([C.POINTER(C.c_int)] * 3, C.c_int),
        "psob_update": ([C.POINTER(UpdateArgs), vp], C.c_int),
        "psob_pack_latent": ([vp, i64, i32, vp, i32, vp], C.c_int),
        "psob_pbest_minloc": ([vp, vp, vp, i64, u64, vp, vp], C.c_int),
        "psob_gbest_candidate": ([vp, vp, vp, vp, vp, i64, u64, i32, vp, vp], C.c_int),
        "psob_gbest_apply": ([vp, i32, i32, vp, vp, vp, i32, vp, f32, vp], C.c_int),
        "psob_score_fitness": ([vp, i32, i64, i32, f32, f32, vp, vp], C.c_int),
        "psob_softmax_select": ([vp, i32, i32, i64, vp, vp], C.c_int),
        "psob_recon_fitness": ([vp, vp, i32, i64, i32, vp, f32, f32, f32, vp, vp], C.c_int),
        "psob_rescale_minmax": ([vp, i64, i32, vp, vp], C.c_int),
        "psob_pack_image": ([vp, i64, vp, vp], C.c_int),
        "psob_pairwise_distance_sum": ([vp, i64, i32, i32, vp, vp], C.c_int),
        "psob_vq_nearest": ([vp, i64, i32, i32, vp, i32, vp, vp, vp], C.c_int),
        "psob_pack_image_stem": ([vp, i64, i32, vp, vp, vp], C.c_int),
        "psob_maxpool_nhwc": ([vp, i64, i64, i32, i32, i32, i32, i32, i32, vp, i64, i32, vp], C.c_int),
        "psob_layer_create": ([C.POINTER(LayerDesc), C.POINTER(vp)], C.c_int),
        "psob_layer_destroy": ([vp], C.c_int),
        "psob_layer_run": ([vp, vp, i64, i64, vp, vp, i32, vp], C.c_int),
        "psob_layer_run_residual": ([vp, vp, i64, i64, vp, vp, vp, i64, i32, vp], C.c_int),
        "psob_layer_profile": ([vp, vp], C.c_int),
        "psob_layer_dot_slots": ([vp], C.c_int),
        "psob_layer_run_dot": ([vp, vp, i64, i64, vp, vp, vp], C.c_int),
        "psob_dot_finish": ([vp, i32, i64, f32, i32, vp, vp], C.c_int),
        "psob_fill_columns_bf16": ([vp, i64, i64, i32, i32, f32, vp], C.c_int),
        "psob_fitness_tail": ([C.POINTER(FitnessTailArgs), vp], C.c_int),
        "psob_gbest_local": ([vp, vp, vp, vp, vp, i64, u64, i32, vp, vp, vp, i32, vp, f32, vp], C.c_int),
        "psob_gbest_record": ([vp, vp, vp, vp, vp, i64, u64, i32, vp, vp], C.c_int),
        "psob_pack_grid_nchw": ([vp, i64, i32, i32, i32, i32, vp, i64, vp], C.c_int),
        "psob_gather_pack_latent": ([vp, vp, i64, i32, vp, i32, vp], C.c_int),
        "psob_layer_run_dot_static": ([vp, vp, i64, i64, vp, f32, i32, vp, vp], C.c_int),
        "psob_layer_recon_slots": ([vp], C.c_int),
        "psob_layer_run_recon": ([vp, vp, i64, i64, vp, vp, vp, i32, i32, vp, vp], C.c_int),
    }
    for name, (args, res) in sig.items():
        fn = getattr(l, name)
        fn.argtypes = args
        fn.restype = res
    _lib = l
    return l


N_CALLS = 0      # C-ABI calls checked so far (every compute entry point is one kernel launch; bench.py counts launches with it)


def check(rc: int, what: str = "") -> None:
    global N_CALLS
    N_CALLS += 1
    if rc != 0:
        msg = lib().psob_last_error().decode("utf-8", "replace")
        raise PsobError(f"{what or 'libpsob'} failed with code {rc}: {msg}")
