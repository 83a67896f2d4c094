// Baseline JPEG decoded on the SMs (the cv::imread side,
// xpano/algorithm/image.cc:83). nvJPEG decodes the Huffman scan of a 12-MP
// source on ONE host thread (37 ms, codec.cu); here the scan is destuffed
// (stuffed zeros and restart markers taken out by a count / scan / scatter
// compaction, bytes packed big-endian into words, so that any 32 bits of it
// are two loads and a funnel shift) and cut into subsequences of kSubBytes
// bytes, one thread each:
//   1. guess pass: every thread decodes its subsequence from "start of a
//      block" -- wrong for most, but Huffman streams resynchronise within a
//      few dozen symbols, so most EXIT states are already right;
//   2. synchronisation: a thread re-decodes whenever its predecessor's exit
//      state differs from the start state it last used, inside a block until
//      the block is stable, across blocks by relaunching until a launch
//      changes nothing (then every exit state follows from the one before it
//      and the first one is exact: the states are the sequential decoder's);
//   3. a scan of the blocks each subsequence completed gives every thread its
//      first coefficient block; the write pass decodes once more and stores;
//   4. DC prediction is a segmented prefix sum per component;
//   5. inverse DCT, chroma upsampling and colour conversion restate
//      libjpeg-turbo (jpeg_sm.h), so the pixels equal cv::imread's bit for bit.
// The arithmetic is in jpeg_sm.h / jpeg_parse.h, which tests/emu/jpeg_emu.cpp
// also compiles for the host. (After Weissenberger & Schmidt, "Massively
// parallel Huffman decoding on GPUs", ICPP 2018, and its JPEG follow-up.)
#include <cstdlib>
#include <string>
#include <vector>

#include "common.h"
#include "jpeg_parse.h"

namespace xp {
namespace jpg {

constexpr int kSubBytes = 128;  // XP_JPEG_SUB overrides (experiments)
static int sub_bytes() {
  static const int v = [] {
    const char* e = getenv("XP_JPEG_SUB");
    const int n = e ? atoi(e) : kSubBytes;
    return n >= 8 && n <= 4096 ? n : kSubBytes;
  }();
  return v;
}
constexpr int kThreads = 256;

struct Job {
  Frame f;
  const uint8_t* raw;  // the stream from a 256-byte boundary before the scan
  uint32_t scan_begin, scan_end;  // the scan's bytes in raw[]
  uint32_t* words;       // the destuffed scan: big-endian words, zero padded
  uint32_t* total_bits;  // its length (device)
  int* ds_count;         // kept bytes per destuffing block, then their offsets
  int ds_blocks;
  const SubRec* subs;
  int n_sub;
  const HuffTab* tabs;
  const uint16_t* quant;  // [3][64]
  uint64_t* exit_state;   // state after a subsequence's last symbol
  uint64_t* in_state;     // start state its last decode used
  int* done;              // blocks it completed
  int* base;              // block index of its first symbol: absolute with
                          // bit 31, else relative to its CTA's base_cta
  int* base_has;          // per CTA of the write pass: it holds an interval
  int* base_cta;          //   start; blocks completed (then: its prefix)
  int* dc_has;            // [3][dc_chunks] DC scan: chunk saw a restart,
  int* dc_val;            //                its sum since the last one
  int dc_chunks;
  int16_t* coef;  // [block][64] zigzag order, AC
  int16_t* dc;    // [block] DC difference, then value
  uint8_t* plane[3];
};

__device__ __forceinline__ void load_tabs(const HuffTab* g, HuffTab* sh) {
  const uint32_t* s = reinterpret_cast<const uint32_t*>(g);
  uint32_t* d = reinterpret_cast<uint32_t*>(sh);
  for (int i = threadIdx.x; i < (int)(4 * sizeof(HuffTab) / 4); i += blockDim.x)
    d[i] = s[i];
  __syncthreads();
}

__device__ __forceinline__ State first_state(const SubRec& r) {
  State s;
  s.p = r.start;
  s.b = 0, s.z = 0;
  return s;
}

// ---------------------------------------------------------------------------
// Destuffing, 16 raw bytes per thread. keep16 -> bit k set when byte i0 + k
// stays (jpeg_sm.h: destuff_keeps), v = the bytes.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t keep16(const Job& j, uint32_t i0,
                                           uint8_t v[16]) {
  const uint4 q = *reinterpret_cast<const uint4*>(j.raw + i0);
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int k = 0; k < 16; ++k) v[k] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
  const uint32_t prev = i0 ? j.raw[i0 - 1] : 0;
  const uint32_t next = j.raw[i0 + 16];  // the buffer is padded with zeros
  uint32_t mask = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const uint32_t i = i0 + k;
    const uint32_t c = v[k], p = k ? v[k - 1] : prev,
                   n = k < 15 ? v[k + 1] : next;
    const bool in = i >= j.scan_begin && i < j.scan_end;
    const bool drop =
        (i > j.scan_begin && p == 0xFF && (c == 0 || (c & 0xF8) == 0xD0)) ||
        (c == 0xFF && i + 1 < j.scan_end && (n & 0xF8) == 0xD0);
    if (in && !drop) mask |= 1u << k;
  }
  return mask;
}

__global__ void __launch_bounds__(256) destuff_count_kernel(const Job j) {
  __shared__ int s_sum[8];
  const uint32_t g = blockIdx.x * 256 + threadIdx.x;
  uint8_t v[16];
  int n = 16 * g < j.scan_end ? __popc(keep16(j, 16 * g, v)) : 0;
#pragma unroll
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += s_sum[k];
    j.ds_count[blockIdx.x] = t;
  }
}

// block counts -> exclusive offsets; the total is the scan's length
__global__ void __launch_bounds__(1024) destuff_offsets_kernel(const Job j) {
  __shared__ int s_val[1024];
  const int t = threadIdx.x;
  const int per = (j.ds_blocks + 1023) / 1024;
  const int lo = min(j.ds_blocks, t * per), hi = min(j.ds_blocks, lo + per);
  int sum = 0;
  for (int k = lo; k < hi; ++k) sum += j.ds_count[k];
  s_val[t] = sum;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    const int add = t >= off ? s_val[t - off] : 0;
    __syncthreads();
    s_val[t] += add;
    __syncthreads();
  }
  int run = s_val[t] - sum;
  for (int k = lo; k < hi; ++k) {
    const int c = j.ds_count[k];
    j.ds_count[k] = run;
    run += c;
  }
  if (t == 1023) *j.total_bits = 8u * (uint32_t)s_val[t];
}

__global__ void __launch_bounds__(256) destuff_scatter_kernel(const Job j) {
  __shared__ int s_sum[8];
  const uint32_t g = blockIdx.x * 256 + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t v[16];
  const uint32_t mask = 16 * g < j.scan_end ? keep16(j, 16 * g, v) : 0;
  const int n = __popc(mask);
  int incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_sum[warp] = incl;
  __syncthreads();
  int off = j.ds_count[blockIdx.x] + incl - n;
  for (int k = 0; k < warp; ++k) off += s_sum[k];
  uint8_t* out = reinterpret_cast<uint8_t*>(j.words);
#pragma unroll
  for (int k = 0; k < 16; ++k)
    if (mask >> k & 1) {
      out[off ^ 3] = v[k];  // big-endian inside each word
      ++off;
    }
}

__global__ void __launch_bounds__(kThreads) huff_guess_kernel(const Job j) {
  __shared__ HuffTab tabs[4];
  load_tabs(j.tabs, tabs);
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= j.n_sub) return;
  const SubRec r = j.subs[i];
  State s = first_state(r);
  j.in_state[i] = ~0ull;  // not a state: the first sweep decodes everything
  j.done[i] = decode_subsequence<false>(j.words, min(r.end, *j.total_bits),
                                        j.f, tabs, &s, nullptr, nullptr, 0,
                                        0);
  j.exit_state[i] = pack(s);
}

__global__ void __launch_bounds__(kThreads)
    huff_sync_kernel(const Job j, int* changed) {
  __shared__ HuffTab tabs[4];
  __shared__ uint64_t sh_exit[kThreads];
  load_tabs(j.tabs, tabs);
  const int t = threadIdx.x, i = blockIdx.x * kThreads + t;
  const bool active = i < j.n_sub;
  SubRec r = {0, 0, 0, 1};
  uint64_t cur = 0, last_in = ~0ull;
  uint32_t sub_end = 0;
  int done = 0;
  if (active) {
    r = j.subs[i];
    sub_end = min(r.end, *j.total_bits);
    cur = j.exit_state[i];
    last_in = j.in_state[i];
    done = j.done[i];
  }
  sh_exit[t] = cur;
  bool wrote = false;
  for (int it = 0; it <= kThreads; ++it) {
    __syncthreads();
    uint64_t in = last_in;
    if (active) {
      if (r.is_first)
        in = pack(first_state(r));
      else if (t > 0)
        in = sh_exit[t - 1];
      else
        in = *reinterpret_cast<volatile uint64_t*>(j.exit_state + i - 1);
    }
    __syncthreads();
    bool ch = false;
    if (active && in != last_in) {
      State s = unpack(in);
      if (s.b >= j.f.bpm) s.b = 0;
      done = decode_subsequence<false>(j.words, sub_end, j.f, tabs, &s,
                                       nullptr, nullptr, 0, 0);
      last_in = in;
      const uint64_t o = pack(s);
      if (o != cur) {
        cur = o;
        sh_exit[t] = o;
        ch = wrote = true;
      }
    }
    if (!__syncthreads_or(ch)) break;
  }
  if (active) {
    j.in_state[i] = last_in;
    j.done[i] = done;
    if (wrote) {
      j.exit_state[i] = cur;
      *changed = 1;
    }
  }
}

// Exclusive scan over the 1024 threads of a block of (has, val) pairs under
// combine(a, b) = b.has ? b : {a.has, a.val + b.val}: "the value since the
// last restart".
__device__ __forceinline__ void block_scan_1024(int* has, int* val, int* s_has,
                                                int* s_val) {
  const int t = threadIdx.x;
  int h = *has, v = *val;
  s_has[t] = h;
  s_val[t] = v;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    if (t >= off && !h) {
      v += s_val[t - off];
      h = s_has[t - off];
    }
    __syncthreads();
    s_has[t] = h;
    s_val[t] = v;
    __syncthreads();
  }
  *has = t ? s_has[t - 1] : 0;
  *val = t ? s_val[t - 1] : 0;
}

// base[i] = first block of subsequence i: its restart interval's first block
// plus the blocks completed by the interval's earlier subsequences. Per CTA
// of kThreads subsequences (the write pass' CTAs): a scan inside the CTA,
// then one CTA scans the CTA totals.
__global__ void __launch_bounds__(kThreads) huff_base_kernel(const Job j) {
  __shared__ int s_has[kThreads], s_val[kThreads];
  const int t = threadIdx.x, i = blockIdx.x * kThreads + t;
  int h = 0, v = 0, first = 0, abs0 = 0;
  if (i < j.n_sub) {
    const SubRec r = j.subs[i];
    first = (int)r.is_first;
    abs0 = (int)r.first_mcu * j.f.bpm;
    h = first;
    v = (first ? abs0 : 0) + j.done[i];
  }
  s_has[t] = h;
  s_val[t] = v;
  __syncthreads();
  for (int off = 1; off < kThreads; off <<= 1) {
    if (t >= off && !h) {
      v += s_val[t - off];
      h = s_has[t - off];
    }
    __syncthreads();
    s_has[t] = h;
    s_val[t] = v;
    __syncthreads();
  }
  if (i < j.n_sub) {
    const int eh = t ? s_has[t - 1] : 0, ev = t ? s_val[t - 1] : 0;
    j.base[i] = first ? (abs0 | (int)0x80000000)
                      : (eh ? (ev | (int)0x80000000) : ev);
  }
  if (t == kThreads - 1) {
    j.base_has[blockIdx.x] = h;
    j.base_cta[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(1024) huff_base_prefix_kernel(const Job j,
                                                                int ctas) {
  __shared__ int s_has[1024], s_val[1024];
  const int t = threadIdx.x;
  const int per = (ctas + 1023) / 1024;
  const int lo = min(ctas, t * per), hi = min(ctas, lo + per);
  int has = 0, val = 0;
  for (int k = lo; k < hi; ++k) {
    if (j.base_has[k])
      has = 1, val = j.base_cta[k];
    else
      val += j.base_cta[k];
  }
  block_scan_1024(&has, &val, s_has, s_val);
  int run = val;
  for (int k = lo; k < hi; ++k) {
    const int h = j.base_has[k], v = j.base_cta[k];
    j.base_cta[k] = run;
    run = h ? v : run + v;
  }
}

__global__ void __launch_bounds__(kThreads) huff_write_kernel(const Job j) {
  __shared__ HuffTab tabs[4];
  load_tabs(j.tabs, tabs);
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= j.n_sub) return;
  const SubRec r = j.subs[i];
  State s = r.is_first ? first_state(r) : unpack(j.exit_state[i - 1]);
  if (s.b >= j.f.bpm) s.b = 0;
  const long long per =
      j.f.restart ? j.f.restart : (long long)j.f.mcus_x * j.f.mcus_y;
  long long limit = ((long long)r.first_mcu + per) * j.f.bpm;
  if (limit > j.f.total_blocks) limit = j.f.total_blocks;
  const int rel = j.base[i];
  const int block =
      rel < 0 ? (rel & 0x7fffffff) : j.base_cta[blockIdx.x] + rel;
  decode_subsequence<true>(j.words, min(r.end, *j.total_bits), j.f, tabs, &s,
                           j.coef, j.dc, block, (int)limit);
}

// DC differences -> DC values: per component a prefix sum over its blocks in
// scan order, restarted with every restart interval. Chunks of kDcChunk MCUs:
// sums per chunk, a scan of the chunk sums (one CTA per component), then
// every chunk is rewritten from its prefix.
constexpr int kDcChunk = 16;

__global__ void __launch_bounds__(256) dc_partial_kernel(const Job j) {
  const int c = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= j.dc_chunks) return;
  const int mcus = j.f.mcus_x * j.f.mcus_y;
  const int nb = j.f.hs[c] * j.f.vs[c], first = j.f.blk_first[c];
  const int lo = q * kDcChunk, hi = min(mcus, lo + kDcChunk);
  int has = 0, val = 0;
  for (int m = lo; m < hi; ++m) {
    if (j.f.restart && m % j.f.restart == 0) has = 1, val = 0;
    for (int b = 0; b < nb; ++b)
      val += j.dc[(size_t)m * j.f.bpm + first + b];
  }
  j.dc_has[c * j.dc_chunks + q] = has;
  j.dc_val[c * j.dc_chunks + q] = val;
}

__global__ void __launch_bounds__(1024) dc_prefix_kernel(const Job j) {
  __shared__ int s_has[1024], s_val[1024];
  const int c = blockIdx.x, t = threadIdx.x;
  int* hs = j.dc_has + c * j.dc_chunks;
  int* vs = j.dc_val + c * j.dc_chunks;
  const int per = (j.dc_chunks + 1023) / 1024;
  const int lo = min(j.dc_chunks, t * per), hi = min(j.dc_chunks, lo + per);
  int has = 0, val = 0;
  for (int q = lo; q < hi; ++q) {
    if (hs[q])
      has = 1, val = vs[q];
    else
      val += vs[q];
  }
  block_scan_1024(&has, &val, s_has, s_val);
  int run = val;
  for (int q = lo; q < hi; ++q) {  // chunk sums -> exclusive prefixes
    const int h = hs[q], v = vs[q];
    vs[q] = run;
    run = h ? v : run + v;
  }
}

__global__ void __launch_bounds__(256) dc_apply_kernel(const Job j) {
  const int c = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= j.dc_chunks) return;
  const int mcus = j.f.mcus_x * j.f.mcus_y;
  const int nb = j.f.hs[c] * j.f.vs[c], first = j.f.blk_first[c];
  const int lo = q * kDcChunk, hi = min(mcus, lo + kDcChunk);
  int run = j.dc_val[c * j.dc_chunks + q];
  for (int m = lo; m < hi; ++m) {
    if (j.f.restart && m % j.f.restart == 0) run = 0;
    for (int b = 0; b < nb; ++b) {
      int16_t* dc = j.dc + (size_t)m * j.f.bpm + first + b;
      run += *dc;
      *dc = (int16_t)run;
    }
  }
}

// one thread per 8x8 block: dequantise, inverse DCT, store into its plane
__global__ void __launch_bounds__(128) idct_kernel(const Job j) {
  __shared__ uint16_t q[3][64];
  for (int i = threadIdx.x; i < 3 * 64; i += blockDim.x)
    q[i / 64][i % 64] = j.quant[i];
  __syncthreads();
  const int blk = blockIdx.x * blockDim.x + threadIdx.x;
  if (blk >= j.f.total_blocks) return;
  const int m = blk / j.f.bpm, b = blk - m * j.f.bpm;
  const int c = j.f.blk_comp[b];
  const int mx = m % j.f.mcus_x, my = m / j.f.mcus_x;
  const int bx = mx * j.f.hs[c] + j.f.blk_x[b];
  const int by = my * j.f.vs[c] + j.f.blk_y[b];
  int16_t in[64];
  const int4* src = reinterpret_cast<const int4*>(j.coef + (size_t)blk * 64);
#pragma unroll
  for (int k = 0; k < 8; ++k)
    *reinterpret_cast<int4*>(in + 8 * k) = src[k];
  uint8_t px[64];
  idct_islow(in, j.dc[blk], q[c], px);
  uint8_t* dst = j.plane[c] + (size_t)by * 8 * j.f.plane_w[c] + (size_t)bx * 8;
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    uint2 v;
    v.x = px[8 * r] | (px[8 * r + 1] << 8) | (px[8 * r + 2] << 16) |
          ((uint32_t)px[8 * r + 3] << 24);
    v.y = px[8 * r + 4] | (px[8 * r + 5] << 8) | (px[8 * r + 6] << 16) |
          ((uint32_t)px[8 * r + 7] << 24);
    *reinterpret_cast<uint2*>(dst + (size_t)r * j.f.plane_w[c]) = v;
  }
}

// Four chroma samples at full resolution for pixels 4g .. 4g + 3 of row Y,
// away from the plane's left and right edges (jpeg_sm.h has the general form)
__device__ __forceinline__ void chroma4(const Frame& f, const uint8_t* p,
                                        int pitch, int ch, int g, int Y,
                                        int c[4]) {
  const int i = 2 * g;
  if (f.chroma == 2) {
    const int r = Y >> 1;
    int rn = (Y & 1) ? r + 1 : r - 1;
    rn = rn < 0 ? 0 : rn > ch - 1 ? ch - 1 : rn;
    const uint8_t* a = p + (size_t)r * pitch + i;
    const uint8_t* b = p + (size_t)rn * pitch + i;
    const int s0 = 3 * a[-1] + b[-1], s1 = 3 * a[0] + b[0],
              s2 = 3 * a[1] + b[1], s3 = 3 * a[2] + b[2];
    c[0] = (3 * s1 + s0 + 8) >> 4;
    c[1] = (3 * s1 + s2 + 7) >> 4;
    c[2] = (3 * s2 + s1 + 8) >> 4;
    c[3] = (3 * s2 + s3 + 7) >> 4;
  } else {
    const uint8_t* a = p + (size_t)Y * pitch + i;
    const int s0 = a[-1], s1 = a[0], s2 = a[1], s3 = a[2];
    c[0] = (3 * s1 + s0 + 1) >> 2;
    c[1] = (3 * s1 + s2 + 2) >> 2;
    c[2] = (3 * s2 + s1 + 1) >> 2;
    c[3] = (3 * s2 + s3 + 2) >> 2;
  }
}

// four pixels per thread; `aligned`: rows of `out` start on 4-byte boundaries
__global__ void __launch_bounds__(256)
    color_kernel(const Job j, uint8_t* __restrict__ out, size_t pitch,
                 int aligned) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y * blockDim.y + threadIdx.y;
  const int X0 = 4 * g;
  if (X0 >= j.f.width || Y >= j.f.height) return;
  const uint8_t* pl[3] = {j.plane[0], j.plane[1], j.plane[2]};
  const int pp[3] = {j.f.plane_w[0], j.f.plane_w[1], j.f.plane_w[2]};
  uint8_t* d = out + (size_t)Y * pitch + 3 * (size_t)X0;
  uint8_t px[12];
  const bool whole = X0 + 3 < j.f.width;
  const bool inner = j.f.ncomp == 3 && j.f.chroma && j.f.fancy && g > 0 &&
                     2 * g + 2 <= j.f.comp_w[1] - 1;
  if (whole && (inner || j.f.ncomp == 1 || j.f.chroma == 0)) {
    const uint32_t y4 =
        *reinterpret_cast<const uint32_t*>(pl[0] + (size_t)Y * pp[0] + X0);
    int cb[4], cr[4];
    if (j.f.ncomp == 3) {
      if (j.f.chroma) {
        chroma4(j.f, pl[1], pp[1], j.f.comp_h[1], g, Y, cb);
        chroma4(j.f, pl[2], pp[2], j.f.comp_h[2], g, Y, cr);
      } else {
        const uint32_t b4 =
            *reinterpret_cast<const uint32_t*>(pl[1] + (size_t)Y * pp[1] + X0);
        const uint32_t r4 =
            *reinterpret_cast<const uint32_t*>(pl[2] + (size_t)Y * pp[2] + X0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          cb[k] = (b4 >> (8 * k)) & 255, cr[k] = (r4 >> (8 * k)) & 255;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int y = (y4 >> (8 * k)) & 255;
      if (j.f.ncomp == 1)
        px[3 * k] = px[3 * k + 1] = px[3 * k + 2] = (uint8_t)y;
      else
        ycc_to_bgr(y, cb[k], cr[k], px + 3 * k);
    }
    if (aligned) {
      uint32_t* o = reinterpret_cast<uint32_t*>(d);
#pragma unroll
      for (int k = 0; k < 3; ++k)
        o[k] = px[4 * k] | (px[4 * k + 1] << 8) | (px[4 * k + 2] << 16) |
               ((uint32_t)px[4 * k + 3] << 24);
    } else {
#pragma unroll
      for (int k = 0; k < 12; ++k) d[k] = px[k];
    }
    return;
  }
  for (int k = 0; k < 4 && X0 + k < j.f.width; ++k) {
    pixel_bgr(j.f, pl, pp, X0 + k, Y, px);
    d[3 * k] = px[0];
    d[3 * k + 1] = px[1];
    d[3 * k + 2] = px[2];
  }
}

}  // namespace jpg

// data (host) -> d_out: width x height interleaved 8-bit BGR as stored (EXIF
// orientation is the caller's), queued on `s` and complete at return. False
// when the stream is not covered (nothing was queued). stats, when given,
// receives {subsequences, synchronisation launches}.
bool sm_jpeg_decode(const uint8_t* data, size_t len, uint8_t* d_out,
                    size_t out_pitch, cudaStream_t s, std::string* why,
                    int stats[2]) {
  using namespace jpg;
  Parsed P;
  if (!parse(data, len, sub_bytes(), &P)) {
    if (why) *why = P.why;
    return false;
  }
  const Frame& f = P.f;
  const int ns = (int)P.subs.size();
  XP_REQUIRE(out_pitch >= (size_t)f.width * 3, "jpeg: pitch below a row");
  Job j;
  j.f = f;
  j.n_sub = ns;
  size_t plane_bytes[3] = {0, 0, 0}, planes_total = 0;
  for (int c = 0; c < f.ncomp; ++c) {
    plane_bytes[c] = ((size_t)f.plane_w[c] * f.plane_h[c] + 255) & ~(size_t)255;
    planes_total += plane_bytes[c];
  }
  const size_t coef_bytes = (size_t)f.total_blocks * 64 * sizeof(int16_t);
  // the scan is uploaded with the bytes before it left out
  const size_t skip = P.scan_begin & ~(size_t)255;
  const size_t raw_bytes = P.scan_end - skip;
  j.scan_begin = (uint32_t)(P.scan_begin - skip);
  j.scan_end = (uint32_t)raw_bytes;
  j.ds_blocks = div_up((int)div_up(raw_bytes, (size_t)16), 256);
  const size_t words_bytes = ((raw_bytes + 3) & ~(size_t)3) + 16;
  const int grid = div_up(ns, kThreads);
  const size_t subs_bytes = (size_t)ns * sizeof(SubRec);
  const size_t meta_bytes = subs_bytes + sizeof(P.huff) + sizeof(P.quant);
  const size_t dc_bytes =
      (((size_t)f.total_blocks * sizeof(int16_t)) + 255) & ~(size_t)255;
  DevBuf raw, words, meta, state, coef, planes;
  Mailbox box;  // pinned staging: released after the stream is drained too
  SyncOnExit drain(s);
  raw.reserve(raw_bytes + 64);
  words.reserve(words_bytes);
  meta.reserve(meta_bytes);
  j.dc_chunks = div_up(f.mcus_x * f.mcus_y, kDcChunk);
  state.reserve((size_t)ns * (2 * sizeof(uint64_t) + 2 * sizeof(int)) +
                (size_t)j.dc_chunks * 6 * sizeof(int) +
                ((size_t)j.ds_blocks + 1 + 2 * (size_t)grid) * sizeof(int));
  coef.reserve(dc_bytes + coef_bytes);
  planes.reserve(planes_total);
  j.raw = raw.as<uint8_t>();
  j.words = words.as<uint32_t>();
  j.subs = meta.as<SubRec>();
  j.tabs = reinterpret_cast<const HuffTab*>(meta.as<uint8_t>() + subs_bytes);
  j.quant = reinterpret_cast<const uint16_t*>(meta.as<uint8_t>() + subs_bytes +
                                              sizeof(P.huff));
  j.exit_state = state.as<uint64_t>();
  j.in_state = j.exit_state + ns;
  j.done = reinterpret_cast<int*>(j.in_state + ns);
  j.base = j.done + ns;
  j.dc_has = j.base + ns;
  j.dc_val = j.dc_has + 3 * (size_t)j.dc_chunks;
  j.ds_count = j.dc_val + 3 * (size_t)j.dc_chunks;
  j.total_bits = reinterpret_cast<uint32_t*>(j.ds_count + j.ds_blocks);
  j.base_has = reinterpret_cast<int*>(j.total_bits + 1);
  j.base_cta = j.base_has + grid;
  j.dc = coef.as<int16_t>();
  j.coef = reinterpret_cast<int16_t*>(coef.as<uint8_t>() + dc_bytes);
  uint8_t* pp = planes.as<uint8_t>();
  for (int c = 0; c < 3; ++c) {
    j.plane[c] = c < f.ncomp ? pp : nullptr;
    pp += plane_bytes[c];
  }
  // The stream's bytes and the tables (subsequences, Huffman, quantisation)
  // go through pinned staging memory of our own, two copies: a
  // cudaMemcpyAsync from pageable memory is staged by the driver under a
  // lock that the other decoding threads also wait for. Sizes are rounded up
  // to a power of two, so that the cached pinned blocks of one panorama's
  // sources serve the next one's.
  size_t pinned = 65536;
  while (pinned < raw_bytes + meta_bytes + 256) pinned *= 2;
  box.reset(pinned);
  int* flag = static_cast<int*>(box.take(4 * sizeof(int)));
  void* d_flag = nullptr;
  XP_CUDA(cudaHostGetDevicePointer(&d_flag, flag, 0));
  void* h_raw = box.take(raw_bytes);
  memcpy(h_raw, data + skip, raw_bytes);
  XP_CUDA(cudaMemcpyAsync(raw.p, h_raw, raw_bytes, cudaMemcpyHostToDevice, s));
  uint8_t* h_meta = static_cast<uint8_t*>(box.take(meta_bytes));
  memcpy(h_meta, P.subs.data(), subs_bytes);
  memcpy(h_meta + subs_bytes, P.huff, sizeof(P.huff));
  memcpy(h_meta + subs_bytes + sizeof(P.huff), P.quant, sizeof(P.quant));
  XP_CUDA(cudaMemcpyAsync(meta.p, h_meta, meta_bytes, cudaMemcpyHostToDevice,
                          s));
  XP_CUDA(cudaMemsetAsync(raw.as<uint8_t>() + raw_bytes, 0, 64, s));
  XP_CUDA(cudaMemsetAsync(words.p, 0, words_bytes, s));
  destuff_count_kernel<<<j.ds_blocks, 256, 0, s>>>(j);
  XP_LAUNCHED();
  destuff_offsets_kernel<<<1, 1024, 0, s>>>(j);
  XP_LAUNCHED();
  destuff_scatter_kernel<<<j.ds_blocks, 256, 0, s>>>(j);
  XP_LAUNCHED();
  huff_guess_kernel<<<grid, kThreads, 0, s>>>(j);
  XP_LAUNCHED();
  // Three synchronisation launches and everything after them are queued
  // without waiting: the first one does the work, the second one repairs what
  // started from a neighbouring block's guessed state, the third one almost
  // always changes nothing, which is all that has to be known -- afterwards.
  // If it did change a state, the round is repeated from there.
  int launches = 0;
  for (int round = 0;; ++round) {
    flag[0] = flag[1] = flag[2] = 0;
    for (int k = 0; k < 3; ++k, ++launches) {
      huff_sync_kernel<<<grid, kThreads, 0, s>>>(j,
                                                 static_cast<int*>(d_flag) + k);
      XP_LAUNCHED();
    }
    XP_CUDA(cudaMemsetAsync(coef.p, 0, dc_bytes + coef_bytes, s));
    huff_base_kernel<<<grid, kThreads, 0, s>>>(j);
    XP_LAUNCHED();
    huff_base_prefix_kernel<<<1, 1024, 0, s>>>(j, grid);
    XP_LAUNCHED();
    huff_write_kernel<<<grid, kThreads, 0, s>>>(j);
    XP_LAUNCHED();
    const dim3 dc_grid(div_up(j.dc_chunks, 256), f.ncomp);
    dc_partial_kernel<<<dc_grid, 256, 0, s>>>(j);
    XP_LAUNCHED();
    dc_prefix_kernel<<<f.ncomp, 1024, 0, s>>>(j);
    XP_LAUNCHED();
    dc_apply_kernel<<<dc_grid, 256, 0, s>>>(j);
    XP_LAUNCHED();
    idct_kernel<<<div_up(f.total_blocks, 128), 128, 0, s>>>(j);
    XP_LAUNCHED();
    dim3 b(32, 8), g(div_up(div_up(f.width, 4), 32), div_up(f.height, 8));
    const int aligned =
        out_pitch % 4 == 0 && reinterpret_cast<uintptr_t>(d_out) % 4 == 0;
    color_kernel<<<g, b, 0, s>>>(j, d_out, out_pitch, aligned);
    XP_LAUNCHED();
    XP_CUDA(cudaStreamSynchronize(s));
    if (!flag[2]) break;
    XP_REQUIRE(round <= grid + 2, "jpeg: synchronisation did not converge");
  }
  if (stats) stats[0] = ns, stats[1] = launches;
  return true;
}

}  // namespace xp
