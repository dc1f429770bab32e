"""ctypes binding of libvfi_b200.so (include/vfi_b200.h).  No CPU fallback: if the library is
missing or no sm_100 GPU is present every product call raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvfi_b200.so")

OK, E_ARG, E_CUDA, E_STATE, E_NOMEM = 0, -1, -2, -3, -4
PRECISION_FP32, PRECISION_BF16 = 0, 1
FLAG_BGR, FLAG_PAD_REFLECT_EDGE = 1, 2
ROUTE_RIFE, ROUTE_VFIMAMBA, ROUTE_SCENE_CHANGE = 0, 1, 2
ROUTE_NAMES = ("rife", "vfimamba", "scene_change")
IFNET_PARAMS = 5664776
NOALLOC_MAX_PAIRS = 4096
PROFILE_CLASSES = ("pad", "head_cnn0", "conv_tcgen05", "conv_direct", "deconv_direct", "gather", "gather_conv0a",
                   "warp_blend", "reduce", "misc")

_vp, _i, _f, _d, _sz, _ll = C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_size_t, C.c_longlong
SIGNATURES = {
    "vfi_create": (_i, [_i, C.POINTER(_vp)]),
    "vfi_destroy": (None, [_vp]),
    "vfi_last_error": (C.c_char_p, [_vp]),
    "vfi_launch_count": (_ll, [_vp]),
    "vfi_fence": (_i, [_vp, _vp]),
    "vfi_device_bytes": (_sz, [_vp]),
    "vfi_set_option": (_i, [_vp, C.c_char_p, _i]),
    "vfi_ifnet_load_weights": (_i, [_vp, _vp, _sz, _vp]),
    "vfi_ifnet_infer": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _f, _i, _vp, _vp, _vp, _vp, _vp]),
    "vfi_ifnet_infer_ex": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _f, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "vfi_moments_u8": (_i, [_vp, _vp, _vp, _i, _sz, _vp, _vp]),
    "vfi_resize_gray_u8": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "vfi_add_weighted_u8": (_i, [_vp, _vp, _vp, _sz, _d, _d, _vp, _vp]),
    "vfi_canny_u8": (_i, [_vp, _vp, _i, _i, _i, _d, _d, _vp, _vp, _vp]),
    "vfi_remap_flow_u8": (_i, [_vp, _vp, _vp, _d, _i, _i, _i, _vp, _vp]),
    "vfi_rgb_to_yuv420_u8": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "vfi_warp_blend": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "vfi_psnr_u8": (_i, [_vp, _vp, _vp, _i, _sz, _vp, _vp, _vp]),
    "vfi_psnr_u8_noalloc": (_i, [_vp, _vp, _vp, _i, _sz, _vp, _vp, _vp]),
    "vfi_ssim_gray_u8": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "vfi_psnr_ssim_u8": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "vfi_flow_stats": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "vfi_laplacian_var": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "vfi_farneback": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _d, _i, _i, _i, _i, _d, _vp, _vp, _vp]),
    "vfi_hud_update": (_i, [_vp, _vp, _i, _i, _vp, _i, C.POINTER(_ll), _d, _vp, _vp, _vp]),
    "vfi_route_decide": (_i, [_d, _d, _d, _d, _d, _d, _d, _d, _d, C.POINTER(_i), C.POINTER(_d),
                              C.POINTER(_i), C.POINTER(_d)]),
    "vfi_profile_enable": (_i, [_vp, _i]),
    "vfi_profile_read": (_i, [_vp, _vp, _vp, _i]),
    "vfi_debug_conv": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _vp, _vp, _vp]),
}

_lib = None


def lib():
    """Load the shared library (once) and attach the prototypes.  Raises if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(make -C csrc).  There is no CPU fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


class VFIError(RuntimeError):
    pass


def check(handle, rc, what):
    if rc != 0:
        msg = lib().vfi_last_error(handle) if handle else b""
        raise VFIError(f"{what} failed ({rc}): {msg.decode(errors='replace') if msg else ''}")


def route_decide(ssim_score, mean, mx, std, lap_var, scene_thr=0.65, particle_thr=0.4, lo=5.0, hi=25.0):
    """adaptive_pipeline.py:242-275 decision tree (host-side C function)."""
    r, c, hp, ps = _i(), _d(), _i(), _d()
    rc = lib().vfi_route_decide(ssim_score, mean, mx, std, lap_var, scene_thr, particle_thr, lo, hi,
                                C.byref(r), C.byref(c), C.byref(hp), C.byref(ps))
    if rc:
        raise VFIError(f"vfi_route_decide failed ({rc})")
    return r.value, c.value, bool(hp.value), ps.value
