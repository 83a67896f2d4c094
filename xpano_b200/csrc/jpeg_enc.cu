// Baseline JPEG encoded on the SMs (the cv::imwrite side,
// xpano/pipeline/stitcher_pipeline.cc:94-116): the stream equals the one
// cv::imwrite / libjpeg-turbo writes for the same pixels and settings, byte
// for byte (baseline; the Annex K Huffman tables or tables optimised for the
// image, jchuff.c jpeg_gen_optimal_table; 4:4:4 / 4:2:2 / 4:2:0, any quality). Arithmetic in jpeg_enc.h, tables and headers in jpeg_enc_host.h;
// tests/emu/jpeg_emu.cpp compiles the same headers for the host.
//   1. colour conversion and chroma downsampling into component planes with
//      libjpeg's edge expansion (enc_planes_kernel);
//   2. forward DCT + quantisation, one thread per 8x8 block, blocks in scan
//      order, coefficients in zigzag order (enc_dct_kernel; enc_dummy_kernel
//      gives the dummy blocks of partial MCUs their neighbour's DC);
//   3. Huffman coding is two passes around a prefix sum: every block's code
//      length (enc_count_kernel), a scan of the lengths (per CTA, then over
//      the CTA totals), then every block writes its bits at its offset -- whole
//      words stored, the two words it shares with its neighbours OR-ed in
//      atomically (enc_write_kernel);
//   4. byte stuffing (a zero after every FF) is a second, byte-level
//      expansion: count / scan / scatter (stuff_*_kernel).
#include <cstdlib>
#include <string>
#include <vector>

#include "common.h"
#include "jpeg_enc_host.h"

namespace xp {
namespace jpg {

struct EncJob {
  EncFrame f;
  const uint8_t* bgr;
  size_t pitch;
  uint8_t* plane[3];
  int pw[3], ph[3];
  const EncTabs* tabs;
  int16_t* coef;       // [block][64], zigzag
  uint32_t* bits;      // per block: code length, then offset inside its CTA
  uint32_t* cta_bits;  // per CTA of 256 blocks: total, then prefix
  uint32_t* words;     // the unstuffed scan, big-endian words
  uint32_t* total;     // [0] bits of the scan, [1] stuffed bytes
  uint32_t* ff_count;  // per stuffing block: FF bytes, then prefix
  uint32_t* hist;      // [4][257] symbol counts per table (optimised tables)
  uint8_t* stuffed;
  int ctas, st_blocks;
};

__device__ __forceinline__ void load_ycc(const EncJob& j, int x, int y, int* Y,
                                         int* Cb, int* Cr) {
  x = min(x, j.f.width - 1);
  y = min(y, j.f.height - 1);
  const uint8_t* p = j.bgr + (size_t)y * j.pitch + 3 * (size_t)x;
  bgr_to_ycc(p[0], p[1], p[2], Y, Cb, Cr);
}

// One thread per chroma sample: the luma samples of its hs x vs pixels and
// the downsampled chroma -- jcsample.c fullsize / h2v1 (bias 0, 1, 0, 1 ...)
// / h2v2 (bias 1, 2, 1, 2 ...). Pixels beyond the image repeat its last
// column / row; chroma rows below the last downsampled row repeat THAT row.
__global__ void __launch_bounds__(256) enc_planes_kernel(const EncJob j) {
  const int cx = blockIdx.x * blockDim.x + threadIdx.x;
  const int cy = blockIdx.y * blockDim.y + threadIdx.y;
  if (cx >= j.pw[1] || cy >= j.ph[1]) return;
  const int hs = j.f.hs, vs = j.f.vs;
  const bool own_rows = cy <= j.f.comp_h[1] - 1;
  int sb = 0, sr = 0;
  for (int dy = 0; dy < vs; ++dy)
    for (int dx = 0; dx < hs; ++dx) {
      const int x = cx * hs + dx, y = cy * vs + dy;
      int Y, Cb, Cr;
      load_ycc(j, x, y, &Y, &Cb, &Cr);
      if (x < j.pw[0] && y < j.ph[0])
        j.plane[0][(size_t)y * j.pw[0] + x] = (uint8_t)Y;
      sb += Cb;
      sr += Cr;
    }
  if (!own_rows) {
    const int r = j.f.comp_h[1] - 1;
    sb = sr = 0;
    for (int dy = 0; dy < vs; ++dy)
      for (int dx = 0; dx < hs; ++dx) {
        int Y, Cb, Cr;
        load_ycc(j, cx * hs + dx, r * vs + dy, &Y, &Cb, &Cr);
        sb += Cb;
        sr += Cr;
      }
  }
  if (hs == 2 && vs == 1) {
    sb = (sb + (cx & 1)) >> 1;
    sr = (sr + (cx & 1)) >> 1;
  } else if (hs == 2 && vs == 2) {
    sb = (sb + 1 + (cx & 1)) >> 2;
    sr = (sr + 1 + (cx & 1)) >> 2;
  }
  j.plane[1][(size_t)cy * j.pw[1] + cx] = (uint8_t)sb;
  j.plane[2][(size_t)cy * j.pw[2] + cx] = (uint8_t)sr;
}

__global__ void __launch_bounds__(128) enc_dct_kernel(const EncJob j) {
  __shared__ uint16_t q[2][64];
  for (int i = threadIdx.x; i < 128; i += blockDim.x)
    q[i / 64][i % 64] = j.tabs->q[i / 64][i % 64];
  __syncthreads();
  const int blk = blockIdx.x * blockDim.x + threadIdx.x;
  if (blk >= j.f.total_blocks) return;
  const BlockPos p = block_pos(j.f, blk);
  int4* dst = reinterpret_cast<int4*>(j.coef + (size_t)blk * 64);
  if (p.dummy) {
#pragma unroll
    for (int k = 0; k < 8; ++k) dst[k] = make_int4(0, 0, 0, 0);
    return;
  }
  const uint8_t* src =
      j.plane[p.comp] + (size_t)p.by * 8 * j.pw[p.comp] + (size_t)p.bx * 8;
  uint8_t px[64];
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const uint2 v =
        *reinterpret_cast<const uint2*>(src + (size_t)r * j.pw[p.comp]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      px[8 * r + k] = (uint8_t)(v.x >> (8 * k));
      px[8 * r + 4 + k] = (uint8_t)(v.y >> (8 * k));
    }
  }
  __align__(16) int16_t out[64];
  fdct_quantise(px, q[p.comp ? 1 : 0], out);
#pragma unroll
  for (int k = 0; k < 8; ++k) dst[k] = *reinterpret_cast<const int4*>(out + 8 * k);
}

__global__ void __launch_bounds__(256) enc_dummy_kernel(const EncJob j) {
  const int blk = blockIdx.x * blockDim.x + threadIdx.x;
  if (blk >= j.f.total_blocks) return;
  if (!block_pos(j.f, blk).dummy) return;
  j.coef[(size_t)blk * 64] = j.coef[(size_t)dummy_source(j.f, blk) * 64];
}

// the DC the block is predicted from: the component's previous block in scan
// order, 0 for its first one
__device__ __forceinline__ int dc_pred(const EncJob& j, int blk) {
  const int bpm = j.f.bpm, ny = j.f.hs * j.f.vs;
  const int m = blk / bpm, b = blk - m * bpm;
  int prev;
  if (b < ny)
    prev = b > 0 ? blk - 1 : blk - bpm + ny - 1;
  else
    prev = blk - bpm;
  if (b < ny ? (b == 0 && m == 0) : m == 0) return 0;
  return j.coef[(size_t)prev * 64];
}

__device__ __forceinline__ void load_block(const EncJob& j, int blk,
                                           int16_t c[64]) {
  const int4* src = reinterpret_cast<const int4*>(j.coef + (size_t)blk * 64);
#pragma unroll
  for (int k = 0; k < 8; ++k) *reinterpret_cast<int4*>(c + 8 * k) = src[k];
}

// symbol statistics for optimised Huffman tables (jchuff.c htest_one_block)
__global__ void __launch_bounds__(256) enc_hist_kernel(const EncJob j) {
  __shared__ uint32_t sh[4 * 257];
  for (int i = threadIdx.x; i < 4 * 257; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int blk = blockIdx.x * 256 + threadIdx.x;
  if (blk < j.f.total_blocks) {
    __align__(16) int16_t c[64];
    load_block(j, blk, c);
    const int comp = block_pos(j.f, blk).comp;
    block_symbols(c, dc_pred(j, blk), comp ? 2 : 0, comp ? 3 : 1,
                  [&](int tab, int sym, uint32_t, int) {
                    atomicAdd(&sh[tab * 257 + sym], 1u);
                  });
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * 257; i += blockDim.x)
    if (sh[i]) atomicAdd(&j.hist[i], sh[i]);
}

// code length of every block; inside a CTA of 256 blocks an exclusive scan,
// the CTA's total to cta_bits
__global__ void __launch_bounds__(256) enc_count_kernel(const EncJob j) {
  __shared__ EncTabs tabs;
  __shared__ uint32_t s_sum[8];
  {
    const uint32_t* s = reinterpret_cast<const uint32_t*>(j.tabs);
    uint32_t* d = reinterpret_cast<uint32_t*>(&tabs);
    for (int i = threadIdx.x; i < (int)(sizeof(EncTabs) / 4); i += blockDim.x)
      d[i] = s[i];
  }
  __syncthreads();
  const int blk = blockIdx.x * 256 + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t n = 0;
  if (blk < j.f.total_blocks) {
    __align__(16) int16_t c[64];
    load_block(j, blk, c);
    const int comp = block_pos(j.f, blk).comp;
    encode_block(c, dc_pred(j, blk), tabs, comp ? 2 : 0, comp ? 3 : 1,
                 [&](uint32_t, int len) { n += (uint32_t)len; });
  }
  uint32_t incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_sum[warp] = incl;
  __syncthreads();
  uint32_t off = incl - n;
  for (int k = 0; k < warp; ++k) off += s_sum[k];
  if (blk < j.f.total_blocks) j.bits[blk] = off;
  if (threadIdx.x == 255) j.cta_bits[blockIdx.x] = off + n;
}

// exclusive scan of up to 1024 * k values by one CTA; the total to *total
__global__ void __launch_bounds__(1024)
    scan_u32_kernel(uint32_t* v, int n, uint32_t* total) {
  __shared__ uint32_t s[1024];
  const int t = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int lo = min(n, t * per), hi = min(n, lo + per);
  uint32_t sum = 0;
  for (int k = lo; k < hi; ++k) sum += v[k];
  s[t] = sum;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    const uint32_t add = t >= off ? s[t - off] : 0;
    __syncthreads();
    s[t] += add;
    __syncthreads();
  }
  uint32_t run = s[t] - sum;
  for (int k = lo; k < hi; ++k) {
    const uint32_t c = v[k];
    v[k] = run;
    run += c;
  }
  if (t == 1023) *total = s[t];
}

__global__ void __launch_bounds__(256) enc_write_kernel(const EncJob j) {
  __shared__ EncTabs tabs;
  {
    const uint32_t* s = reinterpret_cast<const uint32_t*>(j.tabs);
    uint32_t* d = reinterpret_cast<uint32_t*>(&tabs);
    for (int i = threadIdx.x; i < (int)(sizeof(EncTabs) / 4); i += blockDim.x)
      d[i] = s[i];
  }
  __syncthreads();
  const int blk = blockIdx.x * 256 + threadIdx.x;
  if (blk >= j.f.total_blocks) return;
  __align__(16) int16_t c[64];
  load_block(j, blk, c);
  const int comp = block_pos(j.f, blk).comp;
  const uint32_t p = j.cta_bits[blockIdx.x] + j.bits[blk];
  // the bits gather behind the (p & 31) bits of the first word that belong
  // to the blocks before; every completed word is written, the first one
  // (shared with them) and the last, partial one (shared with the next
  // block) by atomicOr into the zeroed buffer
  uint32_t* w = j.words + (p >> 5);
  uint64_t acc = 0;
  int n = (int)(p & 31);
  bool first = true;
  auto emit = [&](uint32_t bits, int len) {
    acc = (acc << len) | bits;
    n += len;
    if (n >= 32) {
      const uint32_t v = (uint32_t)(acc >> (n - 32));
      n -= 32;
      if (first)
        atomicOr(w, v);
      else
        *w = v;
      first = false;
      ++w;
    }
  };
  encode_block(c, dc_pred(j, blk), tabs, comp ? 2 : 0, comp ? 3 : 1, emit);
  if (blk == j.f.total_blocks - 1 && (n & 7))  // jchuff.c flush_bits: ones
    emit((1u << (8 - (n & 7))) - 1, 8 - (n & 7));
  if (n) atomicOr(w, (uint32_t)(acc << (32 - n)));
}

// ---------------------------------------------------------------------------
// Byte stuffing: byte i of the scan goes to i + (FF bytes before it), a zero
// follows every FF. 16 bytes per thread.
// ---------------------------------------------------------------------------
__device__ __forceinline__ int load16(const EncJob& j, uint32_t i0,
                                      uint32_t nbytes, uint8_t v[16]) {
  const uint4 q = *reinterpret_cast<const uint4*>(
      reinterpret_cast<const uint8_t*>(j.words) + i0);
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
  int ff = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    v[k] = (uint8_t)(w[k >> 2] >> (24 - 8 * (k & 3)));  // big-endian words
    if (i0 + k < nbytes && v[k] == 0xFF) ++ff;
  }
  return ff;
}

__global__ void __launch_bounds__(256) stuff_count_kernel(const EncJob j) {
  __shared__ int s_sum[8];
  const uint32_t nbytes = (j.total[0] + 7) >> 3;
  const uint32_t g = blockIdx.x * 256 + threadIdx.x;
  uint8_t v[16];
  int n = 16 * g < nbytes ? load16(j, 16 * g, nbytes, v) : 0;
#pragma unroll
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += s_sum[k];
    j.ff_count[blockIdx.x] = (uint32_t)t;
  }
}

__global__ void __launch_bounds__(256) stuff_scatter_kernel(const EncJob j) {
  __shared__ int s_sum[8];
  const uint32_t nbytes = (j.total[0] + 7) >> 3;
  const uint32_t g = blockIdx.x * 256 + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t v[16];
  const int n = 16 * g < nbytes ? load16(j, 16 * g, nbytes, v) : 0;
  int incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_sum[warp] = incl;
  __syncthreads();
  uint32_t off = 16 * g + j.ff_count[blockIdx.x] + (uint32_t)(incl - n);
  for (int k = 0; k < warp; ++k) off += (uint32_t)s_sum[k];
#pragma unroll
  for (int k = 0; k < 16; ++k)
    if (16 * g + k < nbytes) {
      j.stuffed[off++] = v[k];
      if (v[k] == 0xFF) j.stuffed[off++] = 0;
    }
  if (g == 0) j.total[1] = nbytes + j.total[2];
}

}  // namespace jpg

// d_bgr (device, interleaved 8-bit BGR) -> the JPEG stream cv::imwrite would
// write (baseline, standard Huffman tables), left in pinned memory of `hold`
// (*stream, *len; valid until hold is reset). False when the settings are
// not covered (nothing queued). Complete at return.
bool sm_jpeg_encode(const uint8_t* d_bgr, int w, int h, size_t pitch,
                    int quality, int subsampling, int optimize,
                    cudaStream_t s, Mailbox* hold, const uint8_t** stream,
                    size_t* len) {
  using namespace jpg;
  EncHost E;
  if (!enc_setup(w, h, quality, subsampling, &E)) return false;
  const EncFrame& f = E.f;
  EncJob j;
  j.f = f;
  j.bgr = d_bgr;
  j.pitch = pitch;
  size_t plane_bytes[3], planes_total = 0;
  for (int c = 0; c < 3; ++c) {
    j.pw[c] = f.blocks_w[c] * 8;
    j.ph[c] = f.blocks_h[c] * 8;
    plane_bytes[c] = ((size_t)j.pw[c] * j.ph[c] + 255) & ~(size_t)255;
    planes_total += plane_bytes[c];
  }
  j.ctas = div_up(f.total_blocks, 256);
  // worst case of a block: 64 coefficients of 16 + 11 bits
  const size_t coef_bytes = (size_t)f.total_blocks * 64 * sizeof(int16_t);
  const size_t max_scan = (size_t)f.total_blocks * 64 * 27 / 8 + 64;
  // The scan is bounded by the raw size in practice; the buffers are sized
  // for 3 bytes per pixel and the run fails loudly beyond that. Bit offsets
  // are 32-bit: larger images go to the library.
  const size_t cap_bytes =
      std::min(max_scan, ((size_t)w * h * 3 + 4096 + 15) & ~(size_t)15);
  if (cap_bytes >= 0x1FFFFFF0u) return false;
  DevBuf planes, coef, meta, scan;
  Mailbox box;  // pinned staging: released after the stream is drained too
  SyncOnExit drain(s);
  planes.reserve(planes_total);
  coef.reserve(coef_bytes);
  j.st_blocks = div_up((int)div_up(cap_bytes, (size_t)16), 256);
  const size_t meta_bytes =
      sizeof(EncTabs) + 16 +
      ((size_t)f.total_blocks + j.ctas + j.st_blocks + 8 + 4 * 257) *
          sizeof(uint32_t);
  meta.reserve(meta_bytes);
  scan.reserve(cap_bytes + 64 + 2 * cap_bytes + 64);
  uint8_t* pp = planes.as<uint8_t>();
  for (int c = 0; c < 3; ++c) {
    j.plane[c] = pp;
    pp += plane_bytes[c];
  }
  j.coef = coef.as<int16_t>();
  j.tabs = meta.as<EncTabs>();
  j.total = reinterpret_cast<uint32_t*>(meta.as<uint8_t>() +
                                        ((sizeof(EncTabs) + 15) & ~(size_t)15));
  j.bits = j.total + 4;
  j.cta_bits = j.bits + f.total_blocks;
  j.ff_count = j.cta_bits + j.ctas;
  j.hist = j.ff_count + j.st_blocks;
  j.words = scan.as<uint32_t>();
  j.stuffed = scan.as<uint8_t>() + cap_bytes + 64;
  box.reset(2 * sizeof(EncTabs) + 4 * 257 * sizeof(uint32_t) + 256);
  box.to_device(meta.p, &E.t, sizeof(EncTabs), s);
  uint32_t* h_total = static_cast<uint32_t*>(box.take(4 * sizeof(uint32_t)));
  XP_CUDA(cudaMemsetAsync(j.words, 0, cap_bytes + 64, s));
  dim3 b(32, 8);
  enc_planes_kernel<<<dim3(div_up(j.pw[1], 32), div_up(j.ph[1], 8)), b, 0,
                      s>>>(j);
  XP_LAUNCHED();
  enc_dct_kernel<<<div_up(f.total_blocks, 128), 128, 0, s>>>(j);
  XP_LAUNCHED();
  enc_dummy_kernel<<<j.ctas, 256, 0, s>>>(j);
  XP_LAUNCHED();
  if (optimize) {  // tables from the image's own symbol statistics
    uint32_t* h_hist =
        static_cast<uint32_t*>(box.take(4 * 257 * sizeof(uint32_t)));
    XP_CUDA(cudaMemsetAsync(j.hist, 0, 4 * 257 * sizeof(uint32_t), s));
    enc_hist_kernel<<<j.ctas, 256, 0, s>>>(j);
    XP_LAUNCHED();
    XP_CUDA(cudaMemcpyAsync(h_hist, j.hist, 4 * 257 * sizeof(uint32_t),
                            cudaMemcpyDeviceToHost, s));
    XP_CUDA(cudaStreamSynchronize(s));
    long hist[4][257];
    for (int k = 0; k < 4; ++k)
      for (int i = 0; i < 257; ++i) hist[k][i] = (long)h_hist[k * 257 + i];
    enc_optimise(hist, &E);
    box.to_device(meta.p, &E.t, sizeof(EncTabs), s);
  }
  enc_count_kernel<<<j.ctas, 256, 0, s>>>(j);
  XP_LAUNCHED();
  scan_u32_kernel<<<1, 1024, 0, s>>>(j.cta_bits, j.ctas, j.total);
  XP_LAUNCHED();
  // the scan's length decides whether it fits: read it before writing
  XP_CUDA(cudaMemcpyAsync(h_total, j.total, sizeof(uint32_t),
                          cudaMemcpyDeviceToHost, s));
  XP_CUDA(cudaStreamSynchronize(s));
  const size_t nbytes = ((size_t)h_total[0] + 7) / 8;
  XP_REQUIRE(nbytes <= cap_bytes, "jpeg_encode: scan of %zu bytes exceeds %zu",
             nbytes, cap_bytes);
  enc_write_kernel<<<j.ctas, 256, 0, s>>>(j);
  XP_LAUNCHED();
  const int st = div_up((int)div_up(nbytes, (size_t)16), 256);
  stuff_count_kernel<<<st, 256, 0, s>>>(j);
  XP_LAUNCHED();
  scan_u32_kernel<<<1, 1024, 0, s>>>(j.ff_count, st, j.total + 2);
  XP_LAUNCHED();
  stuff_scatter_kernel<<<st, 256, 0, s>>>(j);
  XP_LAUNCHED();
  XP_CUDA(cudaMemcpyAsync(h_total, j.total, 4 * sizeof(uint32_t),
                          cudaMemcpyDeviceToHost, s));
  XP_CUDA(cudaStreamSynchronize(s));
  const size_t stuffed = h_total[1];
  XP_REQUIRE(stuffed >= nbytes && stuffed <= 2 * nbytes,
             "jpeg_encode: inconsistent stream length");
  std::vector<uint8_t> head;
  enc_headers(E, &head);
  const size_t total = head.size() + stuffed + 2;
  hold->reset(total + 64);
  uint8_t* dst = static_cast<uint8_t*>(hold->take(total));
  memcpy(dst, head.data(), head.size());
  XP_CUDA(cudaMemcpyAsync(dst + head.size(), j.stuffed, stuffed,
                          cudaMemcpyDeviceToHost, s));
  XP_CUDA(cudaStreamSynchronize(s));
  dst[total - 2] = 0xFF;
  dst[total - 1] = 0xD9;
  *stream = dst;
  *len = total;
  return true;
}

}  // namespace xp
