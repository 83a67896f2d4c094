"""ctypes loader of libxpano_b200.so (the C-ABI drop-in library).

The library is built in-tree by `__graft_entry__.build()` /
`make -C xpano_b200/csrc`. There is no CPU fallback: if the library is missing
this module raises at import of the first symbol.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libxpano_b200.so")
_lib = None

c_int, c_float, c_size_t, c_void_p = (ctypes.c_int, ctypes.c_float,
                                      ctypes.c_size_t, ctypes.c_void_p)
P = ctypes.POINTER


class XpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"xpano_b200 status {code}: {msg}")
        self.code = code


class ComposeAux(ctypes.Structure):
    _fields_ = [("gain", c_void_p), ("gain_w", c_int), ("gain_h", c_int),
                ("seam_mask", c_void_p), ("seam_w", c_int), ("seam_h", c_int)]


class BlendOpts(ctypes.Structure):
    _fields_ = [("output_bpp", c_int), ("fixed_levels", c_int),
                ("add_levels", c_int), ("dither", c_int), ("reverse", c_int),
                ("device", c_int), ("keep_debug", c_int),
                ("wideblend", c_int)]


class Timing(ctypes.Structure):
    _fields_ = [(k, ctypes.c_double) for k in
                ("images", "seam", "shrink_mask", "copy", "shrink", "laplace",
                 "blend", "collapse", "out", "total", "pyramid_tail")]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class ComposeImage(ctypes.Structure):
    _fields_ = [("data", c_void_p), ("w", c_int), ("h", c_int),
                ("stride", c_size_t), ("K", c_float * 9), ("R", c_float * 9),
                ("aux", ComposeAux)]


class ComposeOpts(ctypes.Structure):
    _fields_ = [("projection", c_int), ("portrait", c_int), ("a", c_float),
                ("b", c_float), ("warp_scale", c_float),
                ("max_pano_mpx", ctypes.c_double),
                ("warped_image_scale", ctypes.c_double),
                ("compose_work_aspect", ctypes.c_double), ("depth", c_int),
                ("blend", BlendOpts)]


class ComposeInfo(ctypes.Structure):
    _fields_ = [("width", c_int), ("height", c_int), ("min_x", c_int),
                ("min_y", c_int), ("bpp", c_int), ("levels", c_int),
                ("resolution_capped", c_int), ("roi", c_int * 4),
                ("warp_scale", c_float), ("warp_ms", ctypes.c_double),
                ("blend_ms", ctypes.c_double), ("h2d_ms", ctypes.c_double),
                ("d2h_ms", ctypes.c_double), ("timing", Timing),
                ("s_src", ctypes.c_double), ("s_w", ctypes.c_double),
                ("s_t", ctypes.c_double), ("c", ctypes.c_double),
                ("stage_bytes", ctypes.c_double * 9),
                ("stage_launches", c_int * 9),
                ("warp_bytes", ctypes.c_double), ("warp_launches", c_int),
                ("canvas_width", c_int), ("canvas_height", c_int),
                ("tile_x0", c_int), ("window_x0", c_int),
                ("window_width", c_int)]


# every symbol include/xpano_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "xp_last_error": (ctypes.c_char_p, []),
    "xp_version": (ctypes.c_char_p, []),
    "xp_device_count": (c_int, []),
    "xp_kernel_launch_count": (ctypes.c_ulonglong, []),
    "xp_trim_memory": (None, []),
    "xp_warper_create": (c_int, [c_int, c_int, c_float, c_float, c_float,
                                 c_int, P(c_void_p)]),
    "xp_warper_destroy": (None, [c_void_p]),
    "xp_warper_get_scale": (c_float, [c_void_p]),
    "xp_warper_set_scale": (c_int, [c_void_p, c_float]),
    "xp_warper_set_stream": (c_int, [c_void_p, c_void_p]),
    "xp_warper_warp_roi": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p,
                                   c_void_p]),
    "xp_warper_warp_point": (c_int, [c_void_p] * 5),
    "xp_warper_warp_point_backward": (c_int, [c_void_p] * 5),
    "xp_warper_build_maps": (c_int, [c_void_p, c_int, c_int, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_size_t,
                                     c_int, c_void_p]),
    "xp_warper_warp": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t,
                               c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                               c_void_p, c_size_t, c_int, c_void_p]),
    "xp_warper_warp_noalloc": (c_int, [c_void_p, c_void_p, c_int, c_int,
                                       c_size_t, c_int, c_int, c_void_p,
                                       c_void_p, c_int, c_int, c_void_p,
                                       c_size_t, c_void_p, c_void_p]),
    "xp_warper_warp_bgra": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t,
                                    c_int, c_void_p, c_void_p, c_void_p, c_int,
                                    c_void_p, c_size_t, c_int, c_void_p]),
    "xp_blend_opts_default": (None, [P(BlendOpts)]),
    "xp_blender_create": (c_int, [P(BlendOpts), P(c_void_p)]),
    "xp_blender_destroy": (None, [c_void_p]),
    "xp_blender_set_stream": (c_int, [c_void_p, c_void_p]),
    "xp_blender_prepare": (c_int, [c_void_p, c_int, c_int, c_int, c_int]),
    "xp_blender_feed": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                c_size_t, c_int, c_int, c_int, c_int, c_int,
                                c_int]),
    "xp_blender_feed_interleaved": (c_int, [c_void_p, c_void_p, c_size_t,
                                            c_int, c_int, c_int, c_int, c_int,
                                            c_int, c_int]),
    "xp_blender_blend": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p, P(Timing)]),
    "xp_blender_fetch": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                 c_size_t, c_int]),
    "xp_blender_fetch_mask_rle": (c_int, [c_void_p, c_void_p, c_size_t,
                                          P(c_size_t), c_void_p, c_int]),
    "xp_blender_fetch_planes": (c_int, [c_void_p, c_void_p, c_int]),
    "xp_blender_stage_bytes": (c_int, [c_void_p, c_void_p, c_void_p]),
    "xp_blender_debug_fetch": (c_int, [c_void_p, ctypes.c_char_p, c_int, c_int,
                                       c_int, c_void_p, c_size_t, c_void_p,
                                       P(c_size_t)]),
    "xp_pyramid_geometry": (c_int, [c_int] * 5 + [c_void_p] * 7),
    "xp_blend_levels": (c_int, [c_int, c_void_p, c_void_p, c_int, c_int]),
    "xp_compose_opts_default": (None, [P(ComposeOpts)]),
    "xp_compose_capped_scale": (c_int, [ctypes.c_double, ctypes.c_double,
                                        c_int, c_int, ctypes.c_double,
                                        P(c_float)]),
    "xp_composer_create": (c_int, [P(ComposeOpts), P(c_void_p)]),
    "xp_composer_destroy": (None, [c_void_p]),
    "xp_composer_set_stream": (c_int, [c_void_p, c_void_p]),
    "xp_composer_set_images": (c_int, [c_void_p, c_int, P(ComposeImage),
                                       c_int]),
    "xp_composer_run": (c_int, [c_void_p, P(ComposeInfo)]),
    "xp_composer_set_cancel_flag": (c_int, [c_void_p, c_void_p]),
    "xp_composer_set_tile": (c_int, [c_void_p, c_int, c_int]),
    "xp_composer_tile_needs": (c_int, [c_void_p, c_int, c_void_p]),
    "xp_composer_run_tile_begin": (c_int, [c_void_p, c_void_p]),
    "xp_composer_run_tile_end": (c_int, [c_void_p, c_void_p,
                                         P(ComposeInfo)]),
    "xp_tile_halo": (c_int, [c_int]),
    "xp_composer_fetch": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                  c_size_t, c_int]),
    "xp_composer_blender": (c_void_p, [c_void_p]),
    "xp_composer_run_begin": (c_int, [c_void_p]),
    "xp_composer_run_end": (c_int, [c_void_p, c_void_p]),
    "xp_find_largest_crop": (c_int, [c_void_p, c_int, c_int, c_size_t, c_int,
                                     c_int, c_void_p, c_void_p]),
    "xp_blender_auto_crop": (c_int, [c_void_p, c_void_p, c_void_p]),
    "xp_blender_inpaint_mask": (c_int, [c_void_p, c_void_p, c_size_t, c_int,
                                        c_void_p]),
    "xp_largest_crop_from_lines": (c_int, [c_void_p, c_int, c_int, c_void_p,
                                           c_void_p]),
    "xp_gain_compensator_create": (c_int, [c_int, c_int, c_int, c_int,
                                           c_void_p]),
    "xp_gain_compensator_destroy": (None, [c_void_p]),
    "xp_gain_compensator_feed": (c_int, [c_void_p, c_int, c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_int]),
    "xp_gain_compensator_plan": (c_int, [c_void_p, c_int, c_void_p, c_void_p,
                                         c_void_p, c_void_p]),
    "xp_gain_compensator_pairs": (c_int, [c_void_p, c_void_p]),
    "xp_gain_compensator_solve": (c_int, [c_void_p, c_int, c_void_p, c_void_p,
                                          c_void_p]),
    "xp_gain_compensator_map_size": (c_int, [c_void_p, c_int, c_void_p]),
    "xp_gain_compensator_get_map": (c_int, [c_void_p, c_int, c_void_p]),
    "xp_gain_compensator_set_map": (c_int, [c_void_p, c_int, c_void_p, c_int,
                                            c_int]),
    "xp_gain_compensator_apply": (c_int, [c_void_p, c_int, c_void_p, c_int,
                                          c_int, c_size_t, c_int, c_int]),
    "xp_gain_compensator_stats": (c_int, [c_void_p, c_void_p]),
    "xp_codec_create": (c_int, [c_int, c_void_p]),
    "xp_codec_destroy": (None, [c_void_p]),
    "xp_codec_jpeg_info": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                   c_void_p]),
    "xp_codec_stats": (c_int, [c_void_p, c_void_p]),
    "xp_codec_jpeg_decode": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                     c_size_t, c_int]),
    "xp_codec_jpeg_decode_batch": (c_int, [c_void_p, c_int, c_void_p,
                                           c_void_p, c_void_p, c_void_p]),
    "xp_codec_jpeg_encode": (c_int, [c_void_p, c_void_p, c_int, c_int,
                                     c_size_t, c_int, c_int, c_int, c_int,
                                     c_int, c_void_p, c_size_t, c_void_p]),
    "xp_blender_encode_jpeg": (c_int, [c_void_p, c_void_p, c_void_p, c_int,
                                       c_int, c_int, c_int, c_void_p,
                                       c_size_t, c_void_p]),
    "xp_multiband_create": (c_int, [c_int, c_int, c_void_p]),
    "xp_multiband_destroy": (None, [c_void_p]),
    "xp_multiband_prepare": (c_int, [c_void_p, c_int, c_int, c_int, c_int]),
    "xp_multiband_feed": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                  c_size_t, c_int, c_int, c_int, c_int,
                                  c_int]),
    "xp_multiband_blend": (c_int, [c_void_p, c_void_p, c_void_p]),
    "xp_multiband_fetch": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p,
                                   c_size_t, c_int]),
}


def load():
    """Load the library; fail loudly when it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with "
                "`python -c 'import __graft_entry__ as g; g.build()'` "
                "(xpano_b200 has no CPU fallback)")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(status):
    if status != 0:
        raise XpError(status, load().xp_last_error().decode())


MEM_HOST, MEM_DEVICE = 0, 1
