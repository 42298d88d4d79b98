"""ctypes binding of libevoxmo.so (include/evoxmo.h).  No fallback: if the CUDA library is
missing or fails to load, importing the operators raises."""
import ctypes
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
_SO = _HERE / "libevoxmo.so"
_lib = None

c_i64, c_i32, c_vp, c_sz = ctypes.c_int64, ctypes.c_int32, ctypes.c_void_p, ctypes.c_size_t

_SIGNATURES = {
    "evx_version": (ctypes.c_int, []),
    "evx_last_error": (ctypes.c_char_p, []),
    "evx_dominate_relation_u8": (ctypes.c_int, [c_vp, c_i64, c_vp, c_i64, c_i32, c_vp, c_vp]),
    "evx_dominate_relation_bits": (ctypes.c_int, [c_vp, c_i64, c_vp, c_i64, c_i32, c_vp, c_vp]),
    "evx_nds_workspace_bytes": (c_sz, [c_i64, c_i32]),
    "evx_nds_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, c_sz, c_vp]),
    "evx_nds_host_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_vp]),
    "evx_nds_last_num_fronts": (c_i32, []),
    "evx_nds_noalloc_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_vp, c_vp]),
    "evx_rank_from_matrix_workspace_bytes": (c_sz, [c_i64]),
    "evx_rank_from_matrix_u8": (ctypes.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_sz, c_vp]),
    "evx_crowding_workspace_bytes": (c_sz, [c_i64, c_i32]),
    "evx_crowding_distance_f32": (ctypes.c_int, [c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_sz, c_vp]),
    "evx_crowding_distance_sort_f32": (ctypes.c_int, [c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_sz, c_vp]),
    "evx_non_dominate_workspace_bytes": (c_sz, [c_i64, c_i32]),
    "evx_non_dominate_f32": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "evx_select_by_rank_f32": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "evx_spea2_workspace_bytes": (c_sz, [c_i64, c_i32]),
    "evx_spea2_fitness_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, c_sz, c_vp]),
    "evx_spea2_truncation_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_i64, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "evx_nsga3_associate_f32": (ctypes.c_int, [c_vp, c_vp, c_i64, c_i64, c_i32, c_vp, c_vp, c_vp]),
    "evx_nsga3_niche_workspace_bytes": (c_sz, [c_i64, c_i64]),
    "evx_nsga3_niche": (ctypes.c_int, [c_vp, c_i32, c_vp, c_vp, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "evx_eval_dtlz_f32": (ctypes.c_int, [c_i32, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "evx_eval_maf1_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "evx_lsmop_groups": (ctypes.c_int, [c_i32, c_i32, c_vp, c_vp]),
    "evx_eval_lsmop_f32": (ctypes.c_int, [c_i32, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "evx_variation_sbx_pm_f32": (ctypes.c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, ctypes.c_float, ctypes.c_float,
                                                ctypes.c_float, ctypes.c_float, ctypes.c_uint32, ctypes.c_uint32,
                                                c_vp, c_vp]),
    "evx_nds_shard_workspace_bytes": (c_sz, [c_i64, c_i32, c_i32, c_i32]),
    "evx_nds_shard_create": (c_vp, [c_i64, c_i32, c_i32, c_i32, c_vp, c_sz]),
    "evx_nds_shard_destroy": (None, [c_vp]),
    "evx_nds_shard_build": (ctypes.c_int, [c_vp, c_vp, c_vp]),
    "evx_nds_shard_mask_words": (c_i64, [c_vp]),
    "evx_nds_shard_detect": (ctypes.c_int, [c_vp, c_i32, c_vp, c_vp]),
    "evx_nds_shard_apply": (ctypes.c_int, [c_vp, c_i32, c_vp, c_vp, c_vp]),
    "evx_nds_shard_ranks": (ctypes.c_int, [c_vp, c_vp, c_vp]),
    "evx_nds_shard_peer_bytes": (c_sz, [c_vp]),
    "evx_nds_shard_peer_init": (ctypes.c_int, [c_vp, c_vp, c_vp]),
    "evx_nds_shard_peel_fused": (ctypes.c_int, [c_vp, ctypes.POINTER(c_vp), ctypes.POINTER(c_i32), c_vp]),
    "evx_debug_peel_times": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_uint64)]),
    "evx_debug_peel_cta_times": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_uint64), c_i32]),
    "evx_launch_count": (c_i64, []),
    "evx_launch_count_reset": (None, []),
    "evx_profile_enable": (None, [c_i32]),
    "evx_profile_last": (ctypes.c_int, [ctypes.POINTER(ctypes.c_float)] * 3),
}


class EvoxMoError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            raise EvoxMoError(f"{_SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        L = ctypes.CDLL(str(_SO), mode=os.RTLD_GLOBAL if hasattr(os, "RTLD_GLOBAL") else ctypes.DEFAULT_MODE)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError here = header and library out of sync
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise EvoxMoError(lib().evx_last_error().decode(errors="replace") or f"libevoxmo error {rc}")


def exported_symbols():
    return list(_SIGNATURES)
