"""ctypes binding of librainrast.so (C ABI declared in include/rain_rast.h).

There is deliberately no fallback: if the CUDA library has not been built, importing the product path raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librainrast.so")

_c = ctypes
_vp, _i, _i64, _f, _sz = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_float, _c.c_size_t

# name -> (restype, argtypes); must list every symbol include/rain_rast.h declares (tests/test_abi.py checks this)
SIGNATURES = {
    "rr_last_error": (_c.c_char_p, []),
    "rr_version": (_i, []),
    "rr_launch_count": (_i64, []),
    "rr_tile_cull_enabled": (_i, []),
    "rr_prof_enable": (_i, [_i]),
    "rr_prof_num_stages": (_i, []),
    "rr_prof_stage_name": (_c.c_char_p, [_i]),
    "rr_prof_read": (_i, [_c.POINTER(_c.c_double), _c.POINTER(_i64), _i]),
    "rr_geom_bytes": (_sz, [_i]),
    "rr_image_bytes": (_sz, [_i, _i]),
    "rr_binning_bytes": (_sz, [_i64, _i, _i]),
    "rr_backward_ws_bytes": (_sz, [_i]),
    "rr_geom_acc_offset": (_sz, [_i]),
    "rr_forward_geom": (_i, [_vp, _i, _i, _i, _i, _i,
                             _vp, _vp, _vp, _vp,
                             _vp, _f, _vp, _vp,
                             _vp, _vp, _vp,
                             _f, _f, _f, _i, _i,
                             _vp, _vp, _sz, _vp, _sz, _vp]),
    "rr_forward_geom_views": (_i, [_vp, _i, _i, _i, _i, _i, _i,
                                   _vp, _vp, _vp, _vp, _vp,
                                   _vp, _vp, _vp, _vp, _vp, _f, _vp, _f,
                                   _vp, _vp, _sz, _vp, _sz, _vp, _i, _i]),
    "rr_forward_raster": (_i, [_vp, _i, _i, _i, _i64, _vp,
                               _vp, _sz, _vp, _sz, _vp, _sz,
                               _vp, _vp, _i]),
    "rr_forward_bin": (_i, [_vp, _i, _i, _i, _i64,
                            _vp, _sz, _vp, _sz, _vp, _sz, _i]),
    "rr_forward_render": (_i, [_vp, _i, _i, _i, _i64, _vp,
                               _vp, _sz, _vp, _sz, _vp, _sz,
                               _vp, _vp, _i]),
    "rr_backward": (_i, [_vp, _i, _i, _i, _i, _i, _i64, _vp,
                         _vp, _vp, _vp,
                         _vp, _f, _vp, _vp,
                         _vp, _vp, _vp,
                         _f, _f, _f, _vp,
                         _vp, _sz, _vp, _sz, _vp, _sz,
                         _vp, _vp, _sz,
                         _vp, _vp, _vp, _vp,
                         _vp, _vp, _vp, _vp, _vp,
                         _i, _i]),
    "rr_backward_render": (_i, [_vp, _i, _i, _i, _i64, _vp,
                                _vp, _sz, _vp, _sz, _vp, _sz,
                                _vp, _vp, _sz, _i]),
    "rr_backward_views": (_i, [_vp, _i, _i, _i, _i, _i, _i,
                               _vp, _vp, _vp, _vp, _vp,
                               _vp, _vp, _vp, _vp,
                               _vp, _vp, _vp, _vp, _f, _vp, _f,
                               _vp, _vp, _vp, _vp, _vp, _vp,
                               _vp, _vp, _vp, _i, _i, _i, _i]),
    "rr_mark_visible": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "rr_selftest_exp": (_i, [_vp, _c.c_uint32, _c.c_uint32, _vp]),
    "rr_loss_ws_bytes": (_sz, [_i, _i, _i]),
    "rr_loss_l1_ssim": (_i, [_vp, _i, _i, _i, _vp, _vp, _c.c_float, _vp, _sz, _vp, _vp]),
    "rr_loss_l2": (_i, [_vp, _i64, _vp, _vp, _vp, _vp]),
    "rr_loss_l2_u8": (_i, [_vp, _i64, _vp, _vp, _c.c_float, _vp, _vp]),
    "rr_param_step": (_i, [_vp, _i, _i] + [_vp] * 6 + [_vp, _vp, _vp, _c.POINTER(_c.c_float), _i] + [_vp] * 4),
    "rr_param_step_leaf": (_i, [_vp, _i, _i] + [_vp] * 6 + [_vp, _vp, _vp, _c.POINTER(_c.c_float), _i]),
    "rr_activate": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "rr_densify_stats": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "rr_allreduce_signal_bytes": (_sz, []),
    "rr_allreduce_sum_f32_ranges": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_uint32, _i, _i]),
    "rr_allreduce_sum_f32": (_i, [_vp, _i, _i, _i64, _vp, _vp, _vp, _vp, _c.c_uint32, _i]),
    "rr_dist2_ws_bytes": (_sz, [_i]),
    "rr_dist2": (_i, [_vp, _i, _vp, _vp, _vp, _sz]),
    "rr_sort_ws_bytes": (_sz, [_i64, _i, _i, _i]),
    "rr_sort_pairs_u32": (_i, [_vp, _i64, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "rr_sort_pairs_u64": (_i, [_vp, _i64, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "rr_sort_noalloc_max_pairs": (_sz, []),
    "rr_sort_pairs_u32_noalloc": (_i, [_vp, _i64, _vp, _vp, _vp, _vp, _i, _i]),
    "rr_geom_layout": (_i, [_i, _c.POINTER(_i64)]),
    "rr_binning_layout": (_i, [_i64, _i, _i, _c.POINTER(_i64)]),
    "rr_image_layout": (_i, [_i, _i, _c.POINTER(_i64)]),
    "rr_export_sorted_keys": (_i, [_vp, _i, _i64, _i, _i, _vp, _vp, _vp, _vp]),
}

_lib = None


def load():
    """Load librainrast.so once; raise (never fall back) if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the sm_100a CUDA library has not been built. "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` (or `make -C rain_b200/csrc`). "
            "rain_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    msg = load().rr_last_error()
    return msg.decode() if msg else ""


def stage_times(reset=True):
    """{stage name: (total ms, calls)} accumulated since the last reset by rr_prof_enable(1) runs."""
    lib = load()
    n = lib.rr_prof_num_stages()
    ms = (_c.c_double * n)()
    calls = (_i64 * n)()
    lib.rr_prof_read(ms, calls, 1 if reset else 0)
    return {lib.rr_prof_stage_name(i).decode(): (ms[i], calls[i]) for i in range(n)}


def check(status, what):
    if status != 0:
        raise RuntimeError(f"{what} failed: {last_error()}")
