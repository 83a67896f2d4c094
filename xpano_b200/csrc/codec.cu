// JPEG on either side of the path: cv::imread of the sources
// (xpano/algorithm/image.cc:83) and cv::imwrite of the panorama
// (xpano/pipeline/stitcher_pipeline.cc:94-116, parameters of
// CompressionParameters, :40-65: quality, progressive, optimised Huffman
// tables, chroma subsampling), both on the device, so that only compressed
// bytes cross PCIe: a 12-MP source is 3-5 MB instead of 36 MB, a 60-MP
// panorama ~15 MB instead of 182 MB. Decoded pixels are written straight into
// the buffer the composer warps from (XP_MEM_DEVICE), the panorama is encoded
// from the blender's device buffer.
//
// Decoding: baseline / sequential Huffman streams go to the kernels of
// jpeg_sm.cu, whose pixels are libjpeg-turbo's -- cv::imread's -- bit for bit;
// what those do not cover (progressive, ...) goes to nvJPEG, a library,
// linked statically, whose inverse DCT and chroma upsampling are its own
// (tests: PSNR >= 40 dB against cv2.imdecode for those). Encoding: baseline
// streams (standard or optimised Huffman tables) go to the kernels of
// jpeg_enc.cu, whose stream is cv::imwrite's byte for byte; progressive
// streams to nvJPEG.
// The kernel in this file applies the EXIF orientation, as cv::imread does.
#include <nvjpeg.h>

#include <algorithm>
#include <atomic>
#include <cstring>
#include <exception>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "blender.h"
#include "common.h"

namespace xp {

// jpeg_sm.cu: baseline Huffman streams decoded by this repository's kernels
bool sm_jpeg_decode(const uint8_t* data, size_t len, uint8_t* d_out,
                    size_t out_pitch, cudaStream_t s, std::string* why,
                    int stats[2]);

// jpeg_enc.cu: baseline streams with the standard Huffman tables, encoded by
// this repository's kernels, byte-identical to cv::imwrite's
bool sm_jpeg_encode(const uint8_t* d_bgr, int w, int h, size_t pitch,
                    int quality, int subsampling, int optimize,
                    cudaStream_t s, Mailbox* hold, const uint8_t** stream,
                    size_t* len);

static const char* nvjpeg_name(nvjpegStatus_t s) {
  switch (s) {
    case NVJPEG_STATUS_SUCCESS: return "success";
    case NVJPEG_STATUS_NOT_INITIALIZED: return "not initialized";
    case NVJPEG_STATUS_INVALID_PARAMETER: return "invalid parameter";
    case NVJPEG_STATUS_BAD_JPEG: return "bad jpeg";
    case NVJPEG_STATUS_JPEG_NOT_SUPPORTED: return "jpeg not supported";
    case NVJPEG_STATUS_ALLOCATOR_FAILURE: return "allocator failure";
    case NVJPEG_STATUS_EXECUTION_FAILED: return "execution failed";
    case NVJPEG_STATUS_ARCH_MISMATCH: return "arch mismatch";
    case NVJPEG_STATUS_INTERNAL_ERROR: return "internal error";
    case NVJPEG_STATUS_IMPLEMENTATION_NOT_SUPPORTED:
      return "implementation not supported";
    default: return "unknown";
  }
}

// EXIF orientation (tag 0x0112 of IFD0 in the APP1 "Exif" segment), 1..8;
// 1 when absent or malformed. cv::imread applies it (IMREAD_COLOR without
// IMREAD_IGNORE_ORIENTATION), so the decoder here does too.
static int exif_orientation(const uint8_t* d, size_t n) {
  if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return 1;
  size_t p = 2;
  while (p + 4 <= n) {
    if (d[p] != 0xFF) return 1;
    const uint8_t m = d[p + 1];
    if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) {
      p += 2;
      continue;
    }
    if (m == 0xDA || m == 0xD9) return 1;  // image data: no EXIF before it
    const size_t len = ((size_t)d[p + 2] << 8) | d[p + 3];
    if (len < 2 || p + 2 + len > n) return 1;
    if (m == 0xE1 && len >= 16 && memcmp(d + p + 4, "Exif\0\0", 6) == 0) {
      const uint8_t* t = d + p + 10;
      const size_t tn = len - 8;
      if (tn < 8) return 1;
      const bool le = t[0] == 'I' && t[1] == 'I';
      if (!le && !(t[0] == 'M' && t[1] == 'M')) return 1;
      auto u16 = [&](size_t o) -> unsigned {
        return le ? (unsigned)(t[o] | (t[o + 1] << 8))
                  : (unsigned)((t[o] << 8) | t[o + 1]);
      };
      auto u32 = [&](size_t o) -> size_t {
        return le ? ((size_t)t[o] | ((size_t)t[o + 1] << 8) |
                     ((size_t)t[o + 2] << 16) | ((size_t)t[o + 3] << 24))
                  : (((size_t)t[o] << 24) | ((size_t)t[o + 1] << 16) |
                     ((size_t)t[o + 2] << 8) | (size_t)t[o + 3]);
      };
      const size_t ifd = u32(4);
      if (ifd + 2 > tn) return 1;
      const unsigned cnt = u16(ifd);
      for (unsigned k = 0; k < cnt; ++k) {
        const size_t e = ifd + 2 + 12 * (size_t)k;
        if (e + 12 > tn) return 1;
        if (u16(e) == 0x0112) {
          const unsigned v = u16(e + 8);
          return v >= 1 && v <= 8 ? (int)v : 1;
        }
      }
      return 1;
    }
    p += 2 + len;
  }
  return 1;
}

// cv::imread's ExifTransform on interleaved BGR: dst (W' x H') from src
// (w x h); orientations 5-8 transpose (W' = h, H' = w).
__global__ void orient_kernel(const uint8_t* __restrict__ src, size_t spitch,
                              int w, int h, uint8_t* __restrict__ dst,
                              size_t dpitch, int orientation) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y * blockDim.y + threadIdx.y;
  const bool swap = orientation >= 5;
  const int W = swap ? h : w, H = swap ? w : h;
  if (X >= W || Y >= H) return;
  int x, y;
  switch (orientation) {
    case 2: x = w - 1 - X, y = Y; break;          // flip horizontally
    case 3: x = w - 1 - X, y = h - 1 - Y; break;  // rotate 180
    case 4: x = X, y = h - 1 - Y; break;          // flip vertically
    case 5: x = Y, y = X; break;                  // transpose
    case 6: x = Y, y = h - 1 - X; break;          // rotate 90 clockwise
    case 7: x = w - 1 - Y, y = h - 1 - X; break;  // transverse
    case 8: x = w - 1 - Y, y = X; break;          // rotate 90 counter-clockwise
    default: x = X, y = Y; break;
  }
  const uint8_t* s = src + (size_t)y * spitch + 3 * (size_t)x;
  uint8_t* d = dst + (size_t)Y * dpitch + 3 * (size_t)X;
  d[0] = s[0];
  d[1] = s[1];
  d[2] = s[2];
}

#define XP_JPEG(call)                                                     \
  do {                                                                    \
    const nvjpegStatus_t st_ = (call);                                    \
    if (st_ != NVJPEG_STATUS_SUCCESS)                                     \
      fail(st_ == NVJPEG_STATUS_BAD_JPEG ||                               \
                   st_ == NVJPEG_STATUS_JPEG_NOT_SUPPORTED                \
               ? XP_ERR_INVALID                                           \
               : XP_ERR_CUDA,                                             \
           "nvjpeg: %s (%s)", nvjpeg_name(st_), #call);                   \
  } while (0)

struct Codec {
  int device = 0;
  cudaStream_t stream = nullptr;
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  nvjpegEncoderState_t enc = nullptr;
  nvjpegEncoderParams_t par = nullptr;
  int backend = -1;
  // XP_JPEG_SM: 1 (default) = baseline streams go to the SM decoder of
  // jpeg_sm.cu and only what it does not cover to nvJPEG, 0 = nvJPEG for
  // everything, 2 = a stream the SM decoder does not cover is an error
  int sm_mode = 1;
  std::atomic<int> n_sm{0}, n_lib{0};  // images decoded by either
  std::atomic<int> e_sm{0}, e_lib{0};  // images encoded by either
  Mailbox own_box;  // pinned memory holding the stream our encoder produced
  const uint8_t* own_stream = nullptr;
  size_t own_len = 0;

  // -> true when the SM decoder produced the image
  bool decode_sm(const uint8_t* data, size_t len, uint8_t* d_out, size_t pitch,
                 cudaStream_t s) {
    if (!sm_mode) return false;
    std::string why;
    if (sm_jpeg_decode(data, len, d_out, pitch, s, &why, nullptr)) {
      ++n_sm;
      return true;
    }
    if (sm_mode == 2)
      fail(XP_ERR_UNSUPPORTED, "jpeg: not covered by the SM decoder: %s",
           why.c_str());
    return false;
  }

  void init() {
    select_device(device);
    if (handle) return;
    if (const char* e = getenv("XP_JPEG_SM")) sm_mode = atoi(e);
    stream = create_stream(false);
    // Backend of the decoder: the hardware JPEG engine where the device has
    // one, else Huffman decoding on the SMs, else nvJPEG's default (Huffman
    // on one host thread). XP_JPEG_BACKEND=<0..3> forces an nvjpegBackend_t.
    const nvjpegBackend_t order[] = {NVJPEG_BACKEND_HARDWARE,
                                     NVJPEG_BACKEND_GPU_HYBRID,
                                     NVJPEG_BACKEND_DEFAULT};
    nvjpegStatus_t st = NVJPEG_STATUS_NOT_INITIALIZED;
    if (const char* e = getenv("XP_JPEG_BACKEND")) {
      backend = atoi(e);
      st = nvjpegCreateEx((nvjpegBackend_t)backend, nullptr, nullptr, 0,
                          &handle);
    } else {
      for (nvjpegBackend_t b : order) {
        st = nvjpegCreateEx(b, nullptr, nullptr, 0, &handle);
        if (st == NVJPEG_STATUS_SUCCESS) {
          backend = (int)b;
          break;
        }
        handle = nullptr;
      }
    }
    XP_JPEG(st);
    XP_JPEG(nvjpegJpegStateCreate(handle, &state));
    XP_JPEG(nvjpegEncoderStateCreate(handle, &enc, stream));
    XP_JPEG(nvjpegEncoderParamsCreate(handle, &par, stream));
  }

  ~Codec() {
    if (!handle) return;
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != device) cudaSetDevice(device);
    cudaStreamSynchronize(stream);
    for (Lane& l : lanes) {
      if (l.state) nvjpegJpegStateDestroy(l.state);
      if (l.stream) cudaStreamDestroy(l.stream);
    }
    if (par) nvjpegEncoderParamsDestroy(par);
    if (enc) nvjpegEncoderStateDestroy(enc);
    if (state) nvjpegJpegStateDestroy(state);
    nvjpegDestroy(handle);
    cudaStreamDestroy(stream);
    if (cur != device) cudaSetDevice(cur);
  }

  // stored size, channels and EXIF orientation of a stream
  int stored(const uint8_t* data, size_t len, int wh[2], int* channels) {
    init();
    int nc = 0, ws[NVJPEG_MAX_COMPONENT], hs[NVJPEG_MAX_COMPONENT];
    nvjpegChromaSubsampling_t css;
    XP_JPEG(nvjpegGetImageInfo(handle, data, len, &nc, &css, ws, hs));
    wh[0] = ws[0];
    wh[1] = hs[0];
    if (channels) *channels = nc;
    return exif_orientation(data, len);
  }

  // size as cv::imread returns it (EXIF orientation applied)
  void info(const uint8_t* data, size_t len, int wh[2], int* channels) {
    int s[2];
    const int o = stored(data, len, s, channels);
    wh[0] = o >= 5 ? s[1] : s[0];
    wh[1] = o >= 5 ? s[0] : s[1];
  }

  // -> interleaved 8-bit BGR, the layout and orientation cv::imread returns
  void decode(const uint8_t* data, size_t len, void* dst, size_t stride,
              int mem) {
    int sw[2];
    const int o = stored(data, len, sw, nullptr);
    const int W = o >= 5 ? sw[1] : sw[0], H = o >= 5 ? sw[0] : sw[1];
    const size_t row = (size_t)W * 3;
    if (stride == 0) stride = row;
    XP_REQUIRE(stride >= row, "jpeg_decode: stride smaller than a row");
    // decoded as stored; oriented into `fin` (the caller's device buffer, or
    // a staging buffer that is downloaded)
    DevBuf raw, staging;
    uint8_t* fin = static_cast<uint8_t*>(dst);
    size_t fpitch = stride;
    if (mem != XP_MEM_DEVICE) {
      staging.reserve(row * H);
      fin = staging.as<uint8_t>();
      fpitch = row;
    }
    nvjpegImage_t out = {};
    if (o == 1) {
      out.channel[0] = fin;
      out.pitch[0] = fpitch;
    } else {
      raw.reserve((size_t)sw[0] * 3 * sw[1]);
      out.channel[0] = raw.as<unsigned char>();
      out.pitch[0] = (size_t)sw[0] * 3;
    }
    if (!decode_sm(data, len, out.channel[0], out.pitch[0], stream)) {
      XP_JPEG(nvjpegDecode(handle, state, data, len, NVJPEG_OUTPUT_BGRI, &out,
                           stream));
      ++n_lib;
    }
    if (o != 1) {
      dim3 b(32, 8), g(div_up(W, 32), div_up(H, 8));
      orient_kernel<<<g, b, 0, stream>>>(raw.as<uint8_t>(), (size_t)sw[0] * 3,
                                         sw[0], sw[1], fin, fpitch, o);
      XP_LAUNCHED();
    }
    if (mem != XP_MEM_DEVICE)
      copy_out_2d(dst, stride, staging.p, row, row, (size_t)H, mem, stream);
    XP_CUDA(cudaStreamSynchronize(stream));
  }

  // n streams -> n device buffers (interleaved BGR, oriented). Baseline
  // streams are decoded by the kernels of jpeg_sm.cu. For the others the
  // Huffman stage of nvjpegDecode runs on the calling host thread (measured:
  // 37 ms per 12-MP image with every backend this device accepts, the batched
  // GPU-Huffman API included), so the images are decoded by several host
  // threads at once, each with its own decoder state and stream.
  struct Lane {
    nvjpegJpegState_t state = nullptr;
    cudaStream_t stream = nullptr;
  };
  std::vector<Lane> lanes;

  void decode_batch(int n, const uint8_t* const* data, const size_t* lens,
                    void* const* dst, const size_t* strides) {
    init();
    const int T = std::max(1, std::min(n, host_worker_threads(8) + 1));
    while ((int)lanes.size() < T) {
      Lane l;
      XP_JPEG(nvjpegJpegStateCreate(handle, &l.state));
      l.stream = create_stream(false);
      lanes.push_back(l);
    }
    std::vector<DevBuf> raw(n);
    std::vector<int> orient(n), sw(2 * (size_t)n);
    for (int i = 0; i < n; ++i) {
      XP_REQUIRE(data[i] && lens[i] && dst[i], "jpeg_decode_batch: null entry");
      orient[i] = stored(data[i], lens[i], &sw[2 * i], nullptr);
      const int W = orient[i] >= 5 ? sw[2 * i + 1] : sw[2 * i];
      const size_t stride =
          strides && strides[i] ? strides[i] : (size_t)W * 3;
      XP_REQUIRE(stride >= (size_t)W * 3,
                 "jpeg_decode_batch: stride smaller than a row");
      if (orient[i] != 1)
        raw[i].reserve((size_t)sw[2 * i] * 3 * sw[2 * i + 1]);
    }
    std::vector<nvjpegStatus_t> status(n, NVJPEG_STATUS_SUCCESS);
    std::vector<std::exception_ptr> thrown(T);
    std::vector<std::thread> threads;
    for (int t = 0; t < T; ++t)
      threads.emplace_back([&, t] {
        cudaSetDevice(device);
        try {
        for (int i = t; i < n; i += T) {
          const int w = sw[2 * i], h = sw[2 * i + 1];
          const int W = orient[i] >= 5 ? h : w, H = orient[i] >= 5 ? w : h;
          const size_t stride =
              strides && strides[i] ? strides[i] : (size_t)W * 3;
          nvjpegImage_t out = {};
          if (orient[i] == 1) {
            out.channel[0] = static_cast<unsigned char*>(dst[i]);
            out.pitch[0] = stride;
          } else {
            out.channel[0] = raw[i].as<unsigned char>();
            out.pitch[0] = (size_t)w * 3;
          }
          if (!decode_sm(data[i], lens[i], out.channel[0], out.pitch[0],
                         lanes[t].stream)) {
            status[i] = nvjpegDecode(handle, lanes[t].state, data[i], lens[i],
                                     NVJPEG_OUTPUT_BGRI, &out, lanes[t].stream);
            ++n_lib;
          }
          if (status[i] == NVJPEG_STATUS_SUCCESS && orient[i] != 1) {
            dim3 b(32, 8), g(div_up(W, 32), div_up(H, 8));
            orient_kernel<<<g, b, 0, lanes[t].stream>>>(
                raw[i].as<uint8_t>(), (size_t)w * 3, w, h,
                static_cast<uint8_t*>(dst[i]), stride, orient[i]);
          }
        }
        } catch (...) {
          thrown[t] = std::current_exception();
        }
        cudaStreamSynchronize(lanes[t].stream);
      });
    for (std::thread& th : threads) th.join();
    for (int t = 0; t < T; ++t)
      if (thrown[t]) std::rethrow_exception(thrown[t]);
    for (int i = 0; i < n; ++i) XP_JPEG(status[i]);
    XP_CUDA(cudaGetLastError());
  }

  // d_bgr: device, interleaved 8-bit BGR
  void encode_device(const uint8_t* d_bgr, int w, int h, size_t stride,
                     int quality, int subsampling, int progressive,
                     int optimize, uint8_t* out, size_t cap, size_t* len) {
    init();
    XP_REQUIRE(quality >= 1 && quality <= 100, "jpeg_encode: quality 1..100");
    // A size query (out == NULL) encodes and keeps the stream in the encoder
    // state; the call that follows with a buffer only retrieves it.
    if (out && pending && pending_own) {
      XP_REQUIRE(cap >= pending, "jpeg_encode: %zu bytes needed, %zu given",
                 pending, cap);
      copy_bytes(out, own_stream, pending);
      *len = pending;
      pending = 0;
      pending_own = false;
      return;
    }
    if (out && pending) {
      size_t need = pending;
      XP_REQUIRE(cap >= need, "jpeg_encode: %zu bytes needed, %zu given", need,
                 cap);
      XP_JPEG(nvjpegEncodeRetrieveBitstream(handle, enc, out, &need, stream));
      XP_CUDA(cudaStreamSynchronize(stream));
      *len = need;
      pending = 0;
      return;
    }
    pending = 0;
    pending_own = false;
    // Baseline (standard or optimised Huffman tables): our own kernels,
    // whose stream is cv::imwrite's byte for byte. Progressive streams stay
    // with nvJPEG (XP_JPEG_SM=0: everything does).
    if (sm_mode && !progressive &&
        (subsampling == 444 || subsampling == 422 || subsampling == 420) &&
        sm_jpeg_encode(d_bgr, w, h, stride, quality, subsampling, optimize,
                       stream, &own_box, &own_stream, &own_len)) {
      ++e_sm;
      *len = own_len;
      if (!out) {
        pending = own_len;
        pending_own = true;
        return;
      }
      XP_REQUIRE(cap >= own_len, "jpeg_encode: %zu bytes needed, %zu given",
                 own_len, cap);
      copy_bytes(out, own_stream, own_len);
      return;
    }
    if (sm_mode == 2)
      fail(XP_ERR_UNSUPPORTED, "jpeg_encode: not covered by the SM encoder");
    ++e_lib;
    nvjpegChromaSubsampling_t css;
    switch (subsampling) {
      case 444: css = NVJPEG_CSS_444; break;
      case 422: css = NVJPEG_CSS_422; break;
      case 420: css = NVJPEG_CSS_420; break;
      default:
        fail(XP_ERR_INVALID, "jpeg_encode: subsampling must be 444, 422 or 420");
        return;
    }
    XP_JPEG(nvjpegEncoderParamsSetQuality(par, quality, stream));
    XP_JPEG(nvjpegEncoderParamsSetSamplingFactors(par, css, stream));
    XP_JPEG(nvjpegEncoderParamsSetOptimizedHuffman(par, optimize ? 1 : 0,
                                                   stream));
    XP_JPEG(nvjpegEncoderParamsSetEncoding(
        par,
        progressive ? NVJPEG_ENCODING_PROGRESSIVE_DCT_HUFFMAN
                    : NVJPEG_ENCODING_BASELINE_DCT,
        stream));
    nvjpegImage_t src = {};
    src.channel[0] = const_cast<unsigned char*>(d_bgr);
    src.pitch[0] = stride;
    XP_JPEG(nvjpegEncodeImage(handle, enc, par, &src, NVJPEG_INPUT_BGRI, w, h,
                              stream));
    size_t need = 0;
    XP_JPEG(nvjpegEncodeRetrieveBitstream(handle, enc, nullptr, &need, stream));
    XP_CUDA(cudaStreamSynchronize(stream));
    *len = need;
    if (!out) {
      pending = need;
      return;
    }
    XP_REQUIRE(cap >= need, "jpeg_encode: %zu bytes needed, %zu given", need,
               cap);
    XP_JPEG(nvjpegEncodeRetrieveBitstream(handle, enc, out, &need, stream));
    XP_CUDA(cudaStreamSynchronize(stream));
    *len = need;
  }
  size_t pending = 0;  // a stream of this many bytes waits in the encoder
  bool pending_own = false;  // ... in own_stream, not in nvJPEG's state
};

}  // namespace xp

struct xp_codec : xp::Codec {};

extern "C" {

xp_status xp_codec_create(int device, xp_codec** out) {
  return xp::guard([&] {
    XP_REQUIRE(out, "codec_create: null argument");
    *out = nullptr;
    xp_codec* c = new xp_codec();
    c->device = device;
    *out = c;
  });
}

void xp_codec_destroy(xp_codec* c) { delete c; }

xp_status xp_codec_jpeg_info(xp_codec* c, const uint8_t* data, size_t len,
                             int wh[2], int* channels) {
  return xp::guard([&] {
    XP_REQUIRE(c && data && len && wh, "jpeg_info: null argument");
    c->info(data, len, wh, channels);
  });
}

xp_status xp_codec_stats(xp_codec* c, int out[4]) {
  return xp::guard([&] {
    XP_REQUIRE(c && out, "codec_stats: null argument");
    out[0] = c->n_sm.load();
    out[1] = c->n_lib.load();
    out[2] = c->e_sm.load();
    out[3] = c->e_lib.load();
  });
}

xp_status xp_codec_jpeg_decode(xp_codec* c, const uint8_t* data, size_t len,
                               void* dst_bgr, size_t stride, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(c && data && len && dst_bgr, "jpeg_decode: null argument");
    c->decode(data, len, dst_bgr, stride, mem);
  });
}

xp_status xp_codec_jpeg_decode_batch(xp_codec* c, int n,
                                     const uint8_t* const* data,
                                     const size_t* lens,
                                     void* const* dst_bgr_device,
                                     const size_t* strides) {
  return xp::guard([&] {
    XP_REQUIRE(c && n > 0 && data && lens && dst_bgr_device,
               "jpeg_decode_batch: bad argument");
    c->decode_batch(n, data, lens, dst_bgr_device, strides);
  });
}

xp_status xp_codec_jpeg_encode(xp_codec* c, const void* bgr, int w, int h,
                               size_t stride, int mem, int quality,
                               int subsampling, int progressive, int optimize,
                               uint8_t* out, size_t cap, size_t* len) {
  return xp::guard([&] {
    XP_REQUIRE(c && bgr && len && w > 0 && h > 0, "jpeg_encode: bad argument");
    const size_t row = (size_t)w * 3;
    if (stride == 0) stride = row;
    XP_REQUIRE(stride >= row, "jpeg_encode: stride smaller than a row");
    c->init();
    const uint8_t* d = static_cast<const uint8_t*>(bgr);
    xp::DevBuf staging;
    if (mem != XP_MEM_DEVICE) {
      staging.reserve(row * h);
      xp::copy_in_2d(staging.p, row, bgr, stride, row, (size_t)h, mem,
                     c->stream);
      d = staging.as<uint8_t>();
      stride = row;
    }
    c->encode_device(d, w, h, stride, quality, subsampling, progressive,
                     optimize, out, cap, len);
  });
}

xp_status xp_blender_encode_jpeg(xp_blender* b, xp_codec* c,
                                 const int crop_xywh[4], int quality,
                                 int subsampling, int progressive,
                                 int optimize, uint8_t* out, size_t cap,
                                 size_t* len) {
  return xp::guard([&] {
    XP_REQUIRE(b && c && len, "blender_encode_jpeg: null argument");
    xp::Blender* bl = reinterpret_cast<xp::Blender*>(b);
    XP_REQUIRE(bl->blended, "encode_jpeg before blend()");
    XP_REQUIRE(bl->tile.count == 1, "encode_jpeg: not available for a tile");
    XP_REQUIRE(bl->out_bpp == 8, "encode_jpeg: 8-bit panoramas only");
    XP_REQUIRE(c->device == bl->opts.device,
               "encode_jpeg: codec and blender on different devices");
    int x = 0, y = 0, w = bl->W, h = bl->H;
    if (crop_xywh) {
      x = crop_xywh[0], y = crop_xywh[1], w = crop_xywh[2], h = crop_xywh[3];
      XP_REQUIRE(x >= 0 && y >= 0 && w > 0 && h > 0 && x + w <= bl->W &&
                     y + h <= bl->H,
                 "encode_jpeg: crop outside the panorama");
    }
    const size_t stride = (size_t)bl->W * 3;
    c->encode_device(bl->bgr.as<uint8_t>() + (size_t)y * stride + 3 * (size_t)x,
                     w, h, stride, quality, subsampling, progressive, optimize,
                     out, cap, len);
  });
}

}  // extern "C"
