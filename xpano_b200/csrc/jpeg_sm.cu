// Baseline JPEG decoded on the SMs (the cv::imread side,
// xpano/algorithm/image.cc:83). nvJPEG decodes the Huffman scan of a 12-MP
// source on ONE host thread (37 ms, codec.cu); here the scan is cut into
// subsequences of kSubBytes raw bytes, one thread each:
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

__constant__ uint8_t c_natural[64] = XP_JPG_NATURAL;

struct Job {
  Frame f;
  const uint8_t* raw;
  const SubRec* subs;
  int n_sub;
  uint32_t scan_end;
  const HuffTab* tabs;
  const uint16_t* quant;  // [3][64]
  uint64_t* exit_state;   // state after a subsequence's last symbol
  uint64_t* in_state;     // start state its last decode used
  int* done;              // blocks it completed
  int* base;              // block index of its first symbol
  int* dc_has;            // [3][dc_chunks] DC scan: chunk saw a restart,
  int* dc_val;            //                its sum since the last one
  int dc_chunks;
  int16_t* coef;
  uint8_t* plane[3];
};

__device__ __forceinline__ void load_tabs(const HuffTab* g, HuffTab* sh) {
  const uint32_t* s = reinterpret_cast<const uint32_t*>(g);
  uint32_t* d = reinterpret_cast<uint32_t*>(sh);
  for (int i = threadIdx.x; i < (int)(4 * sizeof(HuffTab) / 4); i += blockDim.x)
    d[i] = s[i];
  __syncthreads();
}

__device__ __forceinline__ State first_state(const SubRec& r,
                                             const uint8_t* raw) {
  State s;
  s.cb = r.start;
  // a guess must not begin on a stuffed zero
  if (!r.is_first && raw[r.start] == 0x00 && raw[r.start - 1] == 0xFF)
    s.cb = r.start + 1;
  s.k = 0, s.b = 0, s.z = 0;
  return s;
}

__global__ void __launch_bounds__(kThreads) huff_guess_kernel(const Job j) {
  __shared__ HuffTab tabs[4];
  load_tabs(j.tabs, tabs);
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= j.n_sub) return;
  const SubRec r = j.subs[i];
  State s = first_state(r, j.raw);
  j.in_state[i] = ~0ull;  // not a state: the first sweep decodes everything
  j.done[i] = decode_subsequence<false>(j.raw, r.end, j.scan_end, j.f, tabs,
                                        c_natural, &s, nullptr, 0, 0);
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
  int done = 0;
  if (active) {
    r = j.subs[i];
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
        in = pack(first_state(r, j.raw));
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
      done = decode_subsequence<false>(j.raw, r.end, j.scan_end, j.f, tabs,
                                       c_natural, &s, nullptr, 0, 0);
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
// plus the blocks completed by the interval's earlier subsequences
__global__ void __launch_bounds__(1024) huff_base_kernel(const Job j) {
  __shared__ int s_has[1024], s_val[1024];
  const int t = threadIdx.x;
  const int chunk = (j.n_sub + 1023) / 1024;
  const int lo = min(j.n_sub, t * chunk), hi = min(j.n_sub, lo + chunk);
  int has = 0, val = 0;
  for (int i = lo; i < hi; ++i) {
    const SubRec r = j.subs[i];
    if (r.is_first) has = 1, val = (int)r.first_mcu * j.f.bpm;
    val += j.done[i];
  }
  block_scan_1024(&has, &val, s_has, s_val);
  int run = val;
  for (int i = lo; i < hi; ++i) {
    const SubRec r = j.subs[i];
    if (r.is_first) run = (int)r.first_mcu * j.f.bpm;
    j.base[i] = run;
    run += j.done[i];
  }
}

__global__ void __launch_bounds__(kThreads) huff_write_kernel(const Job j) {
  __shared__ HuffTab tabs[4];
  load_tabs(j.tabs, tabs);
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= j.n_sub) return;
  const SubRec r = j.subs[i];
  State s = r.is_first ? first_state(r, j.raw) : unpack(j.exit_state[i - 1]);
  if (s.b >= j.f.bpm) s.b = 0;
  const long long per =
      j.f.restart ? j.f.restart : (long long)j.f.mcus_x * j.f.mcus_y;
  long long limit = ((long long)r.first_mcu + per) * j.f.bpm;
  if (limit > j.f.total_blocks) limit = j.f.total_blocks;
  decode_subsequence<true>(j.raw, r.end, j.scan_end, j.f, tabs, c_natural, &s,
                           j.coef, j.base[i], (int)limit);
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
      val += j.coef[((size_t)m * j.f.bpm + first + b) * 64];
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
      int16_t* dc = j.coef + ((size_t)m * j.f.bpm + first + b) * 64;
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
  idct_islow(in, q[c], px);
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

__global__ void __launch_bounds__(256)
    color_kernel(const Job j, uint8_t* __restrict__ out, size_t pitch) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y * blockDim.y + threadIdx.y;
  if (X >= j.f.width || Y >= j.f.height) return;
  const uint8_t* pl[3] = {j.plane[0], j.plane[1], j.plane[2]};
  const int pp[3] = {j.f.plane_w[0], j.f.plane_w[1], j.f.plane_w[2]};
  uint8_t bgr[3];
  pixel_bgr(j.f, pl, pp, X, Y, bgr);
  uint8_t* d = out + (size_t)Y * pitch + 3 * (size_t)X;
  d[0] = bgr[0];
  d[1] = bgr[1];
  d[2] = bgr[2];
}

}  // namespace jpg

// Size of a stream this decoder covers; false (and the reason) otherwise.
bool sm_jpeg_probe(const uint8_t* data, size_t len, int wh[2],
                   std::string* why) {
  jpg::Parsed P;
  if (!jpg::parse(data, len, jpg::sub_bytes(), &P)) {
    if (why) *why = P.why;
    return false;
  }
  wh[0] = P.f.width;
  wh[1] = P.f.height;
  return true;
}

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
  j.scan_end = (uint32_t)P.scan_end;
  size_t plane_bytes[3] = {0, 0, 0}, planes_total = 0;
  for (int c = 0; c < f.ncomp; ++c) {
    plane_bytes[c] = ((size_t)f.plane_w[c] * f.plane_h[c] + 255) & ~(size_t)255;
    planes_total += plane_bytes[c];
  }
  const size_t coef_bytes = (size_t)f.total_blocks * 64 * sizeof(int16_t);
  // the scan is uploaded with the bytes before it left out
  const size_t skip = P.scan_begin & ~(size_t)255;
  const size_t raw_bytes = P.scan_end - skip;
  DevBuf raw, subs, tabs, state, coef, planes;
  raw.reserve(raw_bytes + 16);
  subs.reserve((size_t)ns * sizeof(SubRec));
  tabs.reserve(sizeof(P.huff) + sizeof(P.quant));
  j.dc_chunks = div_up(f.mcus_x * f.mcus_y, kDcChunk);
  state.reserve((size_t)ns * (2 * sizeof(uint64_t) + 2 * sizeof(int)) +
                (size_t)j.dc_chunks * 6 * sizeof(int));
  coef.reserve(coef_bytes);
  planes.reserve(planes_total);
  for (SubRec& r : P.subs) {
    r.start -= (uint32_t)skip;
    r.end -= (uint32_t)skip;
  }
  j.scan_end -= (uint32_t)skip;
  j.raw = raw.as<uint8_t>();
  j.subs = subs.as<SubRec>();
  j.tabs = tabs.as<HuffTab>();
  j.quant = reinterpret_cast<const uint16_t*>(tabs.as<uint8_t>() +
                                              sizeof(P.huff));
  j.exit_state = state.as<uint64_t>();
  j.in_state = j.exit_state + ns;
  j.done = reinterpret_cast<int*>(j.in_state + ns);
  j.base = j.done + ns;
  j.dc_has = j.base + ns;
  j.dc_val = j.dc_has + 3 * (size_t)j.dc_chunks;
  j.coef = coef.as<int16_t>();
  uint8_t* pp = planes.as<uint8_t>();
  for (int c = 0; c < 3; ++c) {
    j.plane[c] = c < f.ncomp ? pp : nullptr;
    pp += plane_bytes[c];
  }
  // The stream's bytes and the subsequence table go through pinned staging
  // memory of our own: a cudaMemcpyAsync from pageable memory is staged by
  // the driver under a lock that the other decoding threads also wait for.
  Mailbox box;
  const size_t subs_bytes = (size_t)ns * sizeof(SubRec);
  // (sizes rounded up to a power of two, so that the cached pinned blocks of
  // one panorama's sources serve the next one's)
  size_t pinned = 65536;
  while (pinned < raw_bytes + subs_bytes + sizeof(P.huff) + sizeof(P.quant) + 256)
    pinned *= 2;
  box.reset(pinned);
  int* flag = static_cast<int*>(box.take(2 * sizeof(int)));
  void* d_flag = nullptr;
  XP_CUDA(cudaHostGetDevicePointer(&d_flag, flag, 0));
  void* h_raw = box.take(raw_bytes);
  memcpy(h_raw, data + skip, raw_bytes);
  XP_CUDA(cudaMemcpyAsync(raw.p, h_raw, raw_bytes, cudaMemcpyHostToDevice, s));
  void* h_subs = box.take(subs_bytes);
  memcpy(h_subs, P.subs.data(), subs_bytes);
  XP_CUDA(cudaMemcpyAsync(subs.p, h_subs, subs_bytes, cudaMemcpyHostToDevice,
                          s));
  box.to_device(tabs.p, P.huff, sizeof(P.huff), s);
  box.to_device(tabs.as<uint8_t>() + sizeof(P.huff), P.quant, sizeof(P.quant),
                s);
  const int grid = div_up(ns, kThreads);
  huff_guess_kernel<<<grid, kThreads, 0, s>>>(j);
  XP_LAUNCHED();
  // Two synchronisation launches and everything after them are queued
  // without waiting: the second launch almost always changes nothing (the
  // first one did the work), which is all that has to be known -- afterwards.
  // If it did change a state, the round is repeated from there.
  int launches = 0;
  for (int round = 0;; ++round) {
    flag[0] = flag[1] = 0;
    for (int k = 0; k < 2; ++k, ++launches) {
      huff_sync_kernel<<<grid, kThreads, 0, s>>>(j,
                                                 static_cast<int*>(d_flag) + k);
      XP_LAUNCHED();
    }
    XP_CUDA(cudaMemsetAsync(coef.p, 0, coef_bytes, s));
    huff_base_kernel<<<1, 1024, 0, s>>>(j);
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
    dim3 b(32, 8), g(div_up(f.width, 32), div_up(f.height, 8));
    color_kernel<<<g, b, 0, s>>>(j, d_out, out_pitch);
    XP_LAUNCHED();
    XP_CUDA(cudaStreamSynchronize(s));
    if (!flag[1]) break;
    XP_REQUIRE(round <= grid + 2, "jpeg: synchronisation did not converge");
  }
  if (stats) stats[0] = ns, stats[1] = launches;
  return true;
}

}  // namespace xp
