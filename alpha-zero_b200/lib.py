"""ctypes binding of libaz_b200.so -- the C ABI declared in include/az_b200.h.

Fails loudly when the CUDA library is missing or a call fails; there is no CPU path.
"""
import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("AZ_B200_LIB") or os.path.join(PKG, "libaz_b200.so")   # AZ_B200_LIB: experiment builds (tools/)

AZ_OK = 0
GAME_IDS = {"tictactoe": 0, "connect4": 1, "gomoku": 2}

vp, i32, u64 = C.c_void_p, C.c_int, C.c_uint64
_SIGS = {
    "az_version": (C.c_char_p, []),
    "az_last_error": (C.c_char_p, []),
    "az_forest_create": (i32, [i32, i32, i32, i32, i32, i32, C.POINTER(vp)]),
    "az_forest_destroy": (None, [vp]),
    "az_forest_set_roots_planes": (i32, [vp, vp, vp, vp, vp]),
    "az_forest_set_roots_words": (i32, [vp, vp, vp, vp, vp]),
    "az_forest_set_active": (i32, [vp, vp, vp]),
    "az_forest_set_noise": (i32, [vp, vp, vp]),
    "az_forest_begin_move": (i32, [vp, i32, vp]),
    "az_forest_advance": (i32, [vp, vp]),
    "az_forest_counts": (i32, [vp, C.POINTER(i32), C.POINTER(i32), vp]),
    "az_eval_hash": (i32, [vp, u64, i32, vp]),
    "az_leaves_to_planes": (i32, [vp, vp, i32, vp]),
    "az_leaves_set_logits": (i32, [vp, vp, vp, i32, vp]),
    "az_forest_search_hash": (i32, [vp, i32, u64, i32, vp]),
    "az_forest_set_mode": (i32, [vp, i32]),
    "az_eval_rollout": (i32, [vp, u64, u64, vp]),
    "az_forest_search_rollout": (i32, [vp, i32, u64, vp]),
    "az_forest_root_visits": (i32, [vp, vp, vp, vp]),
    "az_forest_root_children": (i32, [vp, i32, i32, vp, vp, vp, vp, C.POINTER(i32), C.POINTER(C.c_double), vp]),
    "az_forest_stats": (i32, [vp, vp, i32, vp]),
    "az_forest_info": (i32, [vp] + [C.POINTER(i32)] * 5),
    "az_rules_legal": (i32, [i32, i32, i32, i32, vp, vp, vp]),
    "az_rules_winner": (i32, [i32, i32, i32, i32, vp, vp, vp, vp]),
    "az_net_create": (i32, [i32] * 9 + [C.POINTER(vp)]),
    "az_net_destroy": (None, [vp]),
    "az_net_set_param": (i32, [vp, C.c_char_p, vp, C.c_longlong]),
    "az_net_finalize": (i32, [vp]),
    "az_net_eval_leaves": (i32, [vp, vp, vp]),
    "az_net_forward_planes": (i32, [vp, vp, i32, vp, vp, vp, vp]),
    "az_net_launch_count": (C.c_longlong, [vp]),
    "az_net_proj_width": (i32, [vp]),
    "az_net_path_flags": (i32, [vp]),
    "az_forest_search_net": (i32, [vp, vp, i32, vp]),
    "az_net_invalidate_graphs": (i32, [vp]),
    "az_forest_read_leaves": (i32, [vp, i32, vp, vp, vp, vp, vp, vp]),
    "az_net_profile": (i32, [vp, i32]),
    "az_net_profile_read": (i32, [vp, i32, vp, vp, vp]),
    "az_gemm_bf16": (i32, [vp, i32, vp, vp, vp, i32, vp, i32, i32, i32, i32, i32, vp]),
}
_SIGS.update({
    "az_rules_legal_noalloc": (i32, [i32, i32, i32, i32, vp, vp, vp, vp]),
    "az_forest_root_best": (i32, [vp, vp, vp]),
    "az_replay_create": (i32, [i32, i32, i32, C.c_longlong, C.POINTER(vp)]),
    "az_replay_destroy": (None, [vp]),
    "az_replay_size": (C.c_longlong, [vp]),
    "az_replay_entry_bytes": (C.c_longlong, [vp]),
    "az_replay_clear": (i32, [vp]),
    "az_replay_add": (C.c_longlong, [vp, i32, i32, vp, vp, vp, vp, vp, vp]),
    "az_replay_sample": (i32, [vp, i32, u64, u64, vp, vp, vp, vp, vp]),
    "az_replay_read": (i32, [vp, C.c_longlong, i32, vp, vp, vp, vp]),
    "az_d4_source_cell": (i32, [i32, i32, i32, i32]),
    "az_loss_policy_value": (i32, [vp, vp, vp, vp, i32, i32, vp, vp, vp, vp]),
    "az_forest_cache_config": (i32, [vp, i32]),
    "az_forest_cache_clear": (i32, [vp, vp]),
    "az_forest_cache_info": (i32, [vp, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "az_forest_set_evaluator": (i32, [vp, C.c_ulonglong, vp]),
    "az_selfplay_create": (i32, [vp, i32, i32, i32, u64, i32, i32, C.POINTER(vp)]),
    "az_selfplay_destroy": (None, [vp]),
    "az_selfplay_begin": (i32, [vp, i32, vp]),
    "az_selfplay_prepare": (i32, [vp, vp]),
    "az_selfplay_move": (i32, [vp, vp]),
    "az_selfplay_status": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(C.c_longlong), vp]),
    "az_selfplay_run_net": (i32, [vp, vp, i32, i32, vp]),
    "az_selfplay_run_hash": (i32, [vp, i32, u64, i32, i32, vp]),
    "az_selfplay_run_async_net": (i32, [vp, vp, i32, C.c_longlong, i32, vp]),
    "az_selfplay_run_async_hash": (i32, [vp, i32, u64, i32, vp]),
    "az_forest_set_rounds": (i32, [vp, i32]),
    "az_selfplay_buffers": (i32, [vp, C.POINTER(vp), C.POINTER(i32), C.POINTER(i32)]),
    "az_selfplay_read": (i32, [vp] * 10),
    "az_selfplay_read_slots": (i32, [vp, vp, vp, vp]),
    "az_selfplay_launch_count": (C.c_longlong, [vp]),
})
# entry points added by later translation units; bound when present in the header
_OPTIONAL = {}

_lib = None


class AzError(RuntimeError):
    pass


def declared_symbols():
    """Every function name include/az_b200.h declares (used by the CPU export test)."""
    import re
    hdr = os.path.join(os.path.dirname(PKG), "include", "az_b200.h")
    text = open(hdr).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(az_[a-z0-9_]+)\s*\(", text)))


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise AzError(f"{SO_PATH} is missing: run `python __graft_entry__.py build` "
                      "(nvcc, sm_100a). There is no CPU fallback.")
    L = C.CDLL(SO_PATH)
    for table in (_SIGS, _OPTIONAL):
        for name, (res, args) in table.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
    _lib = L
    return L


def check(rc, what=""):
    if rc < 0:
        raise AzError(f"{what} failed ({rc}): {load().az_last_error().decode()}")
    return rc
