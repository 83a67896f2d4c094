"""Host-side mirror of the compose step: Stitcher::ComposePanorama
(xpano/algorithm/stitcher.cc:299-432) with cameras given, executed by the
device-resident compose driver of libxpano_b200.so.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import c_void_p, check
from .blenders import debug_fetch


_STAGES = ("images", "seam", "shrink_mask", "copy", "shrink", "laplace",
           "blend", "collapse", "out")


@dataclass
class ComposeResult:
    pano: np.ndarray | None
    mask: np.ndarray | None
    info: dict = field(default_factory=dict)


class Composer:
    """Upload sources once, run warp -> feed -> blend on the GPU, fetch."""

    def __init__(self, projection=2, warp_scale=1.0, portrait=False, a=2.0,
                 b=1.0, max_pano_mpx=100.0, depth=8, output_bpp=0,
                 fixed_levels=0, add_levels=0, dither=True, reverse=False,
                 device=0, keep_debug=False, warped_image_scale=0.0,
                 compose_work_aspect=1.0):
        """warp_scale: float(warped_image_scale * compose_work_aspect) as
        stitcher.cc:312-313 forms it; or give the two doubles themselves
        (warped_image_scale > 0) so that the size cap rounds as the
        reference's does."""
        self._lib = _lib.load()
        o = _lib.ComposeOpts()
        self._lib.xp_compose_opts_default(ctypes.byref(o))
        o.projection, o.portrait = int(projection), int(bool(portrait))
        o.a, o.b, o.warp_scale = float(a), float(b), float(warp_scale)
        o.max_pano_mpx, o.depth = float(max_pano_mpx), int(depth)
        o.warped_image_scale = float(warped_image_scale)
        o.compose_work_aspect = float(compose_work_aspect)
        o.blend = _lib.BlendOpts(int(output_bpp), int(fixed_levels),
                                 int(add_levels), int(bool(dither)),
                                 int(bool(reverse)), int(device),
                                 int(bool(keep_debug)))
        self._h = c_void_p()
        check(self._lib.xp_composer_create(ctypes.byref(o),
                                           ctypes.byref(self._h)))
        self._keep = []
        self.info = None

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xp_composer_destroy(self._h)
            self._h = c_void_p()

    def set_stream(self, cuda_stream):
        check(self._lib.xp_composer_set_stream(self._h, c_void_p(cuda_stream)))

    def set_images(self, images, Ks, Rs, gains=None, seam_masks=None):
        """images: list of (h, w, 3) host arrays; Ks/Rs: 3x3 float32 each;
        gains / seam_masks: optional per-image low-resolution arrays."""
        n = len(images)
        arr = (_lib.ComposeImage * n)()
        keep = []
        for i, im in enumerate(images):
            im = np.ascontiguousarray(im)
            keep.append(im)
            arr[i].data = im.ctypes.data
            arr[i].h, arr[i].w = im.shape[:2]
            arr[i].stride = 0
            arr[i].K = (ctypes.c_float * 9)(
                *np.asarray(Ks[i], np.float32).reshape(9))
            arr[i].R = (ctypes.c_float * 9)(
                *np.asarray(Rs[i], np.float32).reshape(9))
            if gains is not None and gains[i] is not None:
                g = np.ascontiguousarray(gains[i], np.float32)
                keep.append(g)
                arr[i].aux.gain = g.ctypes.data
                arr[i].aux.gain_h, arr[i].aux.gain_w = g.shape
            if seam_masks is not None and seam_masks[i] is not None:
                m = np.ascontiguousarray(seam_masks[i], np.uint8)
                keep.append(m)
                arr[i].aux.seam_mask = m.ctypes.data
                arr[i].aux.seam_h, arr[i].aux.seam_w = m.shape
        check(self._lib.xp_composer_set_images(self._h, n, arr,
                                               _lib.MEM_HOST))
        self._keep = keep

    def set_images_device(self, ptrs, sizes, Ks, Rs, gains=None,
                          seam_masks=None, strides=None):
        """The same with the source pixels already on the device (e.g.
        decoded there by JpegCodec.decode_device): ptrs = device addresses of
        interleaved BGR images, sizes = [(w, h)], strides in bytes (None =
        packed). The buffers are adopted, not copied: they must stay valid
        until the run has ended. gains / seam_masks are small host arrays."""
        n = len(ptrs)
        arr = (_lib.ComposeImage * n)()
        keep = []
        for i in range(n):
            arr[i].data = int(ptrs[i])
            arr[i].w, arr[i].h = int(sizes[i][0]), int(sizes[i][1])
            arr[i].stride = int(strides[i]) if strides is not None else 0
            arr[i].K = (ctypes.c_float * 9)(
                *np.asarray(Ks[i], np.float32).reshape(9))
            arr[i].R = (ctypes.c_float * 9)(
                *np.asarray(Rs[i], np.float32).reshape(9))
            if gains is not None and gains[i] is not None:
                g = np.ascontiguousarray(gains[i], np.float32)
                keep.append(g)
                arr[i].aux.gain = g.ctypes.data
                arr[i].aux.gain_h, arr[i].aux.gain_w = g.shape
            if seam_masks is not None and seam_masks[i] is not None:
                m = np.ascontiguousarray(seam_masks[i], np.uint8)
                keep.append(m)
                arr[i].aux.seam_mask = m.ctypes.data
                arr[i].aux.seam_h, arr[i].aux.seam_w = m.shape
        check(self._lib.xp_composer_set_images(self._h, n, arr,
                                               _lib.MEM_DEVICE))
        self._keep = keep

    def set_cancel_flag(self, flag):
        """flag: a ctypes.c_int the caller sets non-zero to cancel; polled
        between images like Stitcher::Cancelled() (stitcher.cc:408-412). A
        cancelled run raises XpError with code 6 (XP_CANCELLED)."""
        self._cancel = flag
        check(self._lib.xp_composer_set_cancel_flag(
            self._h, ctypes.byref(flag) if flag is not None else None))

    # ---- canvas tiling of one panorama over several GPUs -------
    def set_tile(self, index, count):
        """This composer computes strip `index` of `count` vertical strips of
        the canvas; call before set_images (a tile uploads only the sources
        whose colours reach its strip)."""
        check(self._lib.xp_composer_set_tile(self._h, int(index), int(count)))
        self._tile = (int(index), int(count))

    def tile_needs(self, n):
        need = (ctypes.c_int * n)()
        check(self._lib.xp_composer_tile_needs(self._h, n, need))
        return [bool(v) for v in need]

    def run_tile_begin(self):
        """-> the 8 partial sums every rank contributes (all-reduce SUM)."""
        part = (ctypes.c_double * 8)()
        check(self._lib.xp_composer_run_tile_begin(self._h, part))
        return [float(v) for v in part]

    def run_tile_end(self, reduced):
        red = (ctypes.c_double * 8)(*[float(v) for v in reduced])
        info = _lib.ComposeInfo()
        check(self._lib.xp_composer_run_tile_end(self._h, red,
                                                 ctypes.byref(info)))
        return self._set_info(info)

    def run_tile(self, reduce=None):
        """One canvas-tiled run; `reduce` sums a list of 8 floats over the
        ranks (default: xpano_b200.sharding.all_reduce_sum)."""
        if reduce is None:
            from . import sharding
            reduce = sharding.all_reduce_sum
        return self.run_tile_end(reduce(self.run_tile_begin()))

    def run(self):
        info = _lib.ComposeInfo()
        check(self._lib.xp_composer_run(self._h, ctypes.byref(info)))
        return self._set_info(info)

    def run_begin(self):
        """First half of run(): returns when the panorama's last kernel is
        queued; another composer may start its own run on the same GPU."""
        check(self._lib.xp_composer_run_begin(self._h))

    def run_end(self):
        """Second half of run(): waits for the panorama; -> info."""
        info = _lib.ComposeInfo()
        check(self._lib.xp_composer_run_end(self._h, ctypes.byref(info)))
        return self._set_info(info)

    def _set_info(self, info):
        self.info = {
            "width": info.width, "height": info.height, "min_x": info.min_x,
            "min_y": info.min_y, "bpp": info.bpp, "levels": info.levels,
            "resolution_capped": bool(info.resolution_capped),
            "roi": tuple(info.roi), "warp_scale": info.warp_scale,
            "warp_ms": info.warp_ms, "blend_ms": info.blend_ms,
            "h2d_ms": info.h2d_ms, "timing": info.timing.as_dict(),
            "s_src": info.s_src, "s_w": info.s_w, "s_t": info.s_t,
            "c": info.c,
            "stage_bytes": dict(zip(_STAGES, info.stage_bytes)),
            "stage_launches": dict(zip(_STAGES, info.stage_launches)),
            "warp_bytes": info.warp_bytes,
            "warp_launches": info.warp_launches,
            "canvas": (info.canvas_width, info.canvas_height),
            "tile_x0": info.tile_x0,
            "window": (info.window_x0, info.window_width)}
        return self.info

    def fetch(self, pano=None, mask=None):
        W, H = self.info["width"], self.info["height"]
        dt = np.uint16 if self.info["bpp"] == 16 else np.uint8
        if pano is None:
            pano = np.empty((H, W, 3), dt)
        if mask is None:
            mask = np.empty((H, W), np.uint8)
        check(self._lib.xp_composer_fetch(self._h, pano.ctypes.data, 0,
                                          mask.ctypes.data, 0, _lib.MEM_HOST))
        return pano, mask

    def fetch_mask_rle(self):
        """The panorama mask as run-length words (utils::Flex layout) instead
        of the dense array: -> (runs, row byte offsets)."""
        from .blenders import fetch_mask_rle
        return fetch_mask_rle(self._lib,
                              c_void_p(self._lib.xp_composer_blender(self._h)),
                              self.info["height"])

    def auto_crop(self):
        """crop::FindLargestCrop on the panorama mask, on the device
        (auto_crop.cc:118-143) -> (x0, y0, x1, y1) or None."""
        from .auto_crop import blender_auto_crop
        return blender_auto_crop(self._lib.xp_composer_blender(self._h))

    def encode_jpeg(self, codec, crop=None, **kw):
        """The panorama as JPEG bytes, encoded on the device by `codec`
        (xpano_b200.JpegCodec): RunExportPipeline's cv::imwrite
        (stitcher_pipeline.cc:94-116) without the raw download."""
        return codec.encode_panorama(self._lib.xp_composer_blender(self._h),
                                     crop, **kw)

    def inpaint_mask(self):
        """~pano_mask and its non-zero count: what RunInpainting hands to
        cv::inpaint (stitcher_pipeline.cc:384-389)."""
        from .auto_crop import blender_inpaint_mask
        return blender_inpaint_mask(self._lib.xp_composer_blender(self._h),
                                    (self.info["width"], self.info["height"]))

    def debug(self, what, image=0, channel=0, level=0, dtype=None):
        h = self._lib.xp_composer_blender(self._h)
        return debug_fetch(self._lib, c_void_p(h), what, image, channel, level,
                           dtype)


def compose_panorama(images, Ks, Rs, gains=None, seam_masks=None, **opts):
    """One-call form: -> ComposeResult(pano, mask, info)."""
    c = Composer(**opts)
    c.set_images(images, Ks, Rs, gains, seam_masks)
    info = c.run()
    pano, mask = c.fetch()
    return ComposeResult(pano, mask, info)
