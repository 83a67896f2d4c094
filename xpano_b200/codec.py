"""Host-side mirror of the JPEG steps either side of the path -- cv::imread of
the sources (xpano/algorithm/image.cc:83) and cv::imwrite of the panorama
(xpano/pipeline/stitcher_pipeline.cc:94-116) -- executed on the device inside
libxpano_b200.so, so that only compressed bytes cross PCIe: baseline streams
are decoded (bit-identical to cv::imread) and encoded (byte-identical to
cv::imwrite, standard or optimised Huffman tables) by the library's own
kernels, progressive streams by nvJPEG. SURVEY.md 8(f2)."""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from ._lib import c_void_p, check

# pipeline::ChromaSubsampling (xpano/pipeline/options.h) -> sampling factors
K444, K422, K420 = 444, 422, 420


class JpegCodec:
    """Decode to / encode from interleaved 8-bit BGR (cv::imread's layout).
    Defaults follow pipeline::CompressionOptions (options.h:45-49): quality
    95, baseline, no optimised tables, 4:2:2."""

    def __init__(self, device=0):
        self._lib = _lib.load()
        self._h = c_void_p()
        check(self._lib.xp_codec_create(int(device), ctypes.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xp_codec_destroy(self._h)
            self._h = c_void_p()

    @staticmethod
    def _bytes(data):
        buf = np.frombuffer(bytes(data), np.uint8) \
            if not isinstance(data, np.ndarray) else \
            np.ascontiguousarray(data, np.uint8)
        return buf

    def info(self, data):
        """-> (width, height, channels) of a JPEG byte string."""
        buf = self._bytes(data)
        wh = (ctypes.c_int * 2)()
        nc = ctypes.c_int()
        check(self._lib.xp_codec_jpeg_info(self._h, buf.ctypes.data, buf.size,
                                           wh, ctypes.byref(nc)))
        return wh[0], wh[1], nc.value

    def stats(self):
        """-> (images decoded by the library's kernels, images decoded by
        nvJPEG) since this codec was created."""
        return self.all_stats()[:2]

    def all_stats(self):
        """-> (decoded by own kernels, decoded by nvJPEG, encoded by own
        kernels, encoded by nvJPEG)."""
        out = (ctypes.c_int * 4)()
        check(self._lib.xp_codec_stats(self._h, out))
        return tuple(out)

    def decode(self, data):
        """-> (h, w, 3) uint8 BGR host array, like cv2.imdecode(IMREAD_COLOR)."""
        buf = self._bytes(data)
        w, h, _ = self.info(buf)
        out = np.empty((h, w, 3), np.uint8)
        check(self._lib.xp_codec_jpeg_decode(self._h, buf.ctypes.data,
                                             buf.size, out.ctypes.data, 0,
                                             _lib.MEM_HOST))
        return out

    def decode_device(self, data, dst_ptr, stride=0):
        """Decode into device memory at dst_ptr (h rows of `stride` bytes,
        0 = packed): what Composer.set_images adopts with device pointers."""
        buf = self._bytes(data)
        check(self._lib.xp_codec_jpeg_decode(self._h, buf.ctypes.data,
                                             buf.size, c_void_p(dst_ptr),
                                             int(stride), _lib.MEM_DEVICE))

    def decode_batch_device(self, datas, dst_ptrs, strides=None):
        """All sources of a panorama in one batch, into device memory."""
        bufs = [self._bytes(d) for d in datas]
        n = len(bufs)
        dp = (ctypes.c_void_p * n)(*[b.ctypes.data for b in bufs])
        ln = (ctypes.c_size_t * n)(*[b.size for b in bufs])
        op = (ctypes.c_void_p * n)(*[int(p) for p in dst_ptrs])
        st = (ctypes.c_size_t * n)(*[int(s) for s in strides]) \
            if strides is not None else None
        check(self._lib.xp_codec_jpeg_decode_batch(self._h, n, dp, ln, op,
                                                   st))

    def encode(self, image, quality=95, subsampling=K422, progressive=False,
               optimize=False):
        """image: (h, w, 3) uint8 BGR host array -> JPEG bytes."""
        img = np.asarray(image)
        if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3:
            raise ValueError("encode: CV_8UC3 image expected")
        if img.strides[2] != 1 or img.strides[1] != 3 or img.strides[0] < 0:
            img = np.ascontiguousarray(img)
        h, w = img.shape[:2]
        args = (img.ctypes.data, w, h, img.strides[0], _lib.MEM_HOST,
                int(quality), int(subsampling), int(bool(progressive)),
                int(bool(optimize)))
        n = ctypes.c_size_t()
        check(self._lib.xp_codec_jpeg_encode(self._h, *args, None, 0,
                                             ctypes.byref(n)))
        out = np.empty(n.value, np.uint8)
        check(self._lib.xp_codec_jpeg_encode(self._h, *args, out.ctypes.data,
                                             out.size, ctypes.byref(n)))
        return out[:n.value].tobytes()

    def encode_panorama(self, blender_handle, crop=None, quality=95,
                        subsampling=K422, progressive=False, optimize=False):
        """JPEG of the panorama a blended blender (or a composer's blender)
        holds on the device, optionally cropped to (x, y, w, h) as
        RunExportPipeline does (stitcher_pipeline.cc:99-102)."""
        rect = (ctypes.c_int * 4)(*[int(v) for v in crop]) if crop else None
        args = (c_void_p(blender_handle), self._h, rect, int(quality),
                int(subsampling), int(bool(progressive)), int(bool(optimize)))
        n = ctypes.c_size_t()
        check(self._lib.xp_blender_encode_jpeg(*args, None, 0,
                                               ctypes.byref(n)))
        out = np.empty(n.value, np.uint8)
        check(self._lib.xp_blender_encode_jpeg(*args, out.ctypes.data,
                                               out.size, ctypes.byref(n)))
        return out[:n.value].tobytes()
