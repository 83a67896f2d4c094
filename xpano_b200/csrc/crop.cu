// Auto-crop of the panorama mask: crop::FindLargestCrop
// (xpano/algorithm/auto_crop.cc:118-143), the consumer of result_mask_
// (SURVEY.md 8f3).
//
// The reference walks every column of the mask for its longest run of 255
// (FindLongestLineInColumn, :32-63: H*W byte reads, one column at a time, with
// a 60-MB stride) and then grows rectangles from 1 + W/512 seed columns over
// the per-column lines (:65-113, O(W) each). The first part is the data-
// parallel one and runs on the device -- one thread per column, a warp reads
// 32 adjacent bytes of a row, so the mask is streamed once, coalesced; the
// second part works on W (start, end) pairs and stays host arithmetic.
#include <climits>
#include <vector>

#include "blender.h"
#include "common.h"

namespace xp {

constexpr int kAutoCropSamplingDistance = 512;  // xpano/constants.h:89
constexpr int kMaskValueOn = 255;               // xpano/constants.h

// lines[x] = {start, end} of the longest run of kMaskValueOn in column x (the
// first one among equally long runs: auto_crop.cc:51-53 replaces only on a
// strictly longer run), {rows, 0} when the column has none.
__global__ void __launch_bounds__(128)
    column_runs_kernel(const uint8_t* __restrict__ mask, int w, int h,
                       size_t stride, int2* __restrict__ lines) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= w) return;
  int best = 0, best_start = h, best_end = 0, run = 0, start = 0;
  const uint8_t* p = mask + x;
  int y = 0;
  // eight rows in flight per thread
  for (; y + 8 <= h; y += 8) {
    uint8_t v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __ldg(p + (size_t)(y + k) * stride);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (v[k] == kMaskValueOn) {
        if (run == 0) start = y + k;
        ++run;
        if (run > best) best = run, best_start = start, best_end = y + k + 1;
      } else {
        run = 0;
      }
    }
  }
  for (; y < h; ++y) {
    if (__ldg(p + (size_t)y * stride) == kMaskValueOn) {
      if (run == 0) start = y;
      ++run;
      if (run > best) best = run, best_start = start, best_end = y + 1;
    } else {
      run = 0;
    }
  }
  lines[x] = make_int2(best_start, best_end);
}

namespace {
struct Rect {
  int x0, y0, x1, y1;
};
inline long area(const Rect& r) {
  return (long)(r.x1 - r.x0) * (long)(r.y1 - r.y0);
}
inline bool valid(const int2& l) { return l.x < l.y; }

// auto_crop.cc:65-113
bool expand(const std::vector<int2>& lines, int2 invalid, int seed,
            Rect* out) {
  const int n = (int)lines.size();
  if (!valid(lines[seed])) return false;
  Rect cur{seed, lines[seed].x, seed + 1, lines[seed].y};
  Rect largest = cur;
  if (n == 1) {
    *out = largest;
    return true;
  }
  int left = seed - 1;
  int right = seed + n % 2;
  int2 ll = lines[left], rl = lines[right];
  while (valid(ll) || valid(rl)) {
    const Rect lr{left, std::max(ll.x, cur.y0), cur.x1, std::min(ll.y, cur.y1)};
    const Rect rr{cur.x0, std::max(rl.x, cur.y0), right + 1,
                  std::min(rl.y, cur.y1)};
    if (area(lr) > area(rr)) {
      cur = lr;
      ll = left == 0 ? invalid : lines[--left];
    } else {
      cur = rr;
      rl = right == n - 1 ? invalid : lines[++right];
    }
    if (area(cur) >= area(largest)) largest = cur;
  }
  *out = largest;
  return true;
}
}  // namespace

// auto_crop.cc:118-143 on the per-column lines
bool largest_crop_from_lines(const std::vector<int2>& lines, int rows,
                             int out_rect[4]) {
  const int n = (int)lines.size();
  if (n == 0 || rows == 0) return false;
  const int2 invalid = make_int2(rows, 0);
  bool have = false;
  Rect best{0, 0, 0, 0};
  const int num_samples = 1 + n / kAutoCropSamplingDistance;
  for (int i = 0; i < num_samples; ++i) {
    const int start = (int)((long)(i + 1) * n / (num_samples + 1));
    Rect r;
    if (expand(lines, invalid, start, &r) &&
        (!have || area(r) >= area(best))) {
      best = r;
      have = true;
    }
  }
  if (have) {
    out_rect[0] = best.x0;
    out_rect[1] = best.y0;
    out_rect[2] = best.x1;
    out_rect[3] = best.y1;
  }
  return have;
}

// device mask -> rect; `s` is synchronised
bool find_largest_crop_device(const uint8_t* d_mask, int w, int h,
                              size_t stride, cudaStream_t s, int out_rect[4]) {
  if (w <= 0 || h <= 0) return false;
  DevBuf d_lines;
  d_lines.reserve(sizeof(int2) * (size_t)w);
  column_runs_kernel<<<div_up(w, 128), 128, 0, s>>>(d_mask, w, h, stride,
                                                    d_lines.as<int2>());
  XP_LAUNCHED();
  std::vector<int2> lines(w);
  XP_CUDA(cudaMemcpyAsync(lines.data(), d_lines.p, sizeof(int2) * (size_t)w,
                          cudaMemcpyDeviceToHost, s));
  XP_CUDA(cudaStreamSynchronize(s));
  return largest_crop_from_lines(lines, h, out_rect);
}

// inpaint_mask = ~pano_mask and its countNonZero (the inputs of the auto-fill
// step, stitcher_pipeline.cc:384-389)
__global__ void __launch_bounds__(256)
    invert_count_kernel(const uint8_t* __restrict__ mask, size_t n,
                        uint8_t* __restrict__ inv,
                        unsigned long long* __restrict__ count) {
  unsigned long long c = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    const uint8_t v = (uint8_t)~mask[i];
    inv[i] = v;
    c += v != 0;
  }
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(~0u, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}

}  // namespace xp

extern "C" {

xp_status xp_largest_crop_from_lines(const int* lines_start_end, int n_columns,
                                     int rows, int out_rect[4], int* found) {
  return xp::guard([&] {
    XP_REQUIRE(out_rect && found && (lines_start_end || n_columns == 0),
               "largest_crop_from_lines: null argument");
    std::vector<int2> lines(n_columns > 0 ? n_columns : 0);
    for (int i = 0; i < n_columns; ++i)
      lines[i] = make_int2(lines_start_end[2 * i], lines_start_end[2 * i + 1]);
    *found = xp::largest_crop_from_lines(lines, rows, out_rect) ? 1 : 0;
  });
}

xp_status xp_find_largest_crop(const uint8_t* mask, int w, int h,
                               size_t stride, int mem, int device,
                               int out_rect[4], int* found) {
  return xp::guard([&] {
    XP_REQUIRE(out_rect && found, "find_largest_crop: null argument");
    *found = 0;
    if (!mask || w <= 0 || h <= 0) return;  // empty matrix: no crop
    if (stride == 0) stride = (size_t)w;
    xp::select_device(device);
    const uint8_t* d = mask;
    xp::DevBuf staging;
    size_t dstride = stride;
    if (mem == XP_MEM_HOST) {
      staging.reserve((size_t)w * h);
      xp::copy_in_2d(staging.p, (size_t)w, mask, stride, (size_t)w, (size_t)h,
                     mem, nullptr);
      d = staging.as<uint8_t>();
      dstride = (size_t)w;
    }
    *found = xp::find_largest_crop_device(d, w, h, dstride, nullptr, out_rect)
                 ? 1
                 : 0;
    XP_CUDA(cudaStreamSynchronize(nullptr));
  });
}

xp_status xp_blender_inpaint_mask(xp_blender* b, uint8_t* dst_mask,
                                  size_t mask_stride, int mem,
                                  long long* pixels_to_fill) {
  return xp::guard([&] {
    XP_REQUIRE(b && pixels_to_fill, "inpaint_mask: null argument");
    xp::Blender* bl = reinterpret_cast<xp::Blender*>(b);
    XP_REQUIRE(bl->blended, "inpaint_mask before blend()");
    XP_REQUIRE(bl->tile.count == 1, "inpaint_mask: not available for a tile");
    xp::select_device(bl->opts.device);
    cudaStream_t s = bl->stream;
    const size_t n = (size_t)bl->W * bl->H;
    xp::DevBuf inv, cnt;
    inv.reserve(n);
    cnt.reserve(sizeof(unsigned long long));
    XP_CUDA(cudaMemsetAsync(cnt.p, 0, sizeof(unsigned long long), s));
    xp::invert_count_kernel<<<xp::sm_count() * 8, 256, 0, s>>>(
        bl->full_mask.as<uint8_t>(), n, inv.as<uint8_t>(),
        cnt.as<unsigned long long>());
    XP_LAUNCHED();
    unsigned long long h = 0;
    XP_CUDA(cudaMemcpyAsync(&h, cnt.p, sizeof(h), cudaMemcpyDeviceToHost, s));
    if (dst_mask) {
      if (mask_stride == 0) mask_stride = (size_t)bl->W;
      xp::copy_out_2d(dst_mask, mask_stride, inv.p, (size_t)bl->W,
                      (size_t)bl->W, (size_t)bl->H, mem, s);
    }
    XP_CUDA(cudaStreamSynchronize(s));
    *pixels_to_fill = (long long)h;
  });
}

xp_status xp_blender_auto_crop(xp_blender* b, int out_rect[4], int* found) {
  return xp::guard([&] {
    XP_REQUIRE(b && out_rect && found, "auto_crop: null argument");
    xp::Blender* bl = reinterpret_cast<xp::Blender*>(b);
    XP_REQUIRE(bl->blended, "auto_crop before blend()");
    XP_REQUIRE(bl->tile.count == 1, "auto_crop: not available for a tile");
    xp::select_device(bl->opts.device);
    *found = xp::find_largest_crop_device(bl->full_mask.as<uint8_t>(), bl->W,
                                          bl->H, (size_t)bl->W, bl->stream,
                                          out_rect)
                 ? 1
                 : 0;
  });
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Run-length form of the panorama mask: multiblend's Result.full_mask
// (utils::Flex, mb/include/mb/flex.h:22-75, written by Flex::MaskWrite /
// MaskFinalise, src/flex.cpp:78-98, and decoded by xpano's ToMat,
// blenders.cc:38-62): per row a sequence of 32-bit words
// (white << 31) | run_length over maximal runs, rows back to back, plus the
// byte offset of every row (Flex::rows_). 60 MB of dense mask become a few
// hundred kilobytes on the way to the host.
// ---------------------------------------------------------------------------
namespace xp {

// one block per row: number of runs of the row
__global__ void __launch_bounds__(256)
    rle_count_kernel(const uint8_t* __restrict__ mask, int w, size_t stride,
                     int* __restrict__ counts) {
  const uint8_t* row = mask + (size_t)blockIdx.x * stride;
  int n = 0;
  for (int x = threadIdx.x; x < w; x += blockDim.x)
    n += x == 0 || (row[x] != 0) != (row[x - 1] != 0);
  __shared__ int sh[8];
  for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(~0u, n, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += sh[k];
    counts[blockIdx.x] = t;
  }
}

// exclusive scan of the row counts (one block); offs[h] = total
__global__ void __launch_bounds__(1024)
    rle_scan_kernel(const int* __restrict__ counts, int h,
                    int* __restrict__ offs) {
  __shared__ int sh[1024];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < h; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < h ? counts[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
      const int t = threadIdx.x >= d ? sh[threadIdx.x - d] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < h) offs[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry += sh[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) offs[h] = carry;
}

// one block per row: the run starts, in order, as (white << 31) | x
__global__ void __launch_bounds__(256)
    rle_emit_kernel(const uint8_t* __restrict__ mask, int w, size_t stride,
                    const int* __restrict__ offs, uint32_t* __restrict__ runs) {
  const uint8_t* row = mask + (size_t)blockIdx.x * stride;
  uint32_t* out = runs + offs[blockIdx.x];
  __shared__ int warp_sum[8];
  __shared__ int base;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int x0 = 0; x0 < w; x0 += blockDim.x) {
    const int x = x0 + threadIdx.x;
    const bool start =
        x < w && (x == 0 || (row[x] != 0) != (row[x - 1] != 0));
    const unsigned b = __ballot_sync(~0u, start);
    if (lane == 0) warp_sum[wid] = __popc(b);
    __syncthreads();
    int before = base;
    for (int k = 0; k < wid; ++k) before += warp_sum[k];
    if (start)
      out[before + __popc(b & ((1u << lane) - 1u))] =
          (row[x] != 0 ? 0x80000000u : 0u) | (uint32_t)x;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int k = 0; k < 8; ++k) t += warp_sum[k];
      base += t;
    }
    __syncthreads();
  }
}

// start positions -> run lengths (in place, every word reads its successor
// first: one thread per word, two phases through a register)
__global__ void rle_lengths_kernel(uint32_t* __restrict__ runs,
                                   const int* __restrict__ offs, int h, int w,
                                   const uint32_t* __restrict__ starts) {
  const int y = blockIdx.y;
  const int n = offs[y + 1] - offs[y];
  const uint32_t* in = starts + offs[y];
  uint32_t* out = runs + offs[y];
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n;
       k += gridDim.x * blockDim.x) {
    const uint32_t s = in[k];
    const uint32_t next = k + 1 < n ? (in[k + 1] & 0x7fffffffu) : (uint32_t)w;
    out[k] = (s & 0x80000000u) | (next - (s & 0x7fffffffu));
  }
  (void)h;
}

}  // namespace xp

extern "C" {

xp_status xp_blender_fetch_mask_rle(xp_blender* b, uint32_t* runs,
                                    size_t cap_words, size_t* n_words,
                                    int32_t* row_offsets, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(b && n_words, "fetch_mask_rle: null argument");
    xp::Blender* bl = reinterpret_cast<xp::Blender*>(b);
    XP_REQUIRE(bl->blended, "fetch_mask_rle before blend()");
    XP_REQUIRE(bl->tile.count == 1,
               "fetch_mask_rle: not available for a canvas tile");
    xp::select_device(bl->opts.device);
    cudaStream_t s = bl->stream;
    const int W = bl->W, H = bl->H;
    if (!bl->rle_ready) {
      bl->rle_offs.reserve(sizeof(int) * ((size_t)H + 1));
      xp::DevBuf counts;
      counts.reserve(sizeof(int) * (size_t)H);
      xp::rle_count_kernel<<<H, 256, 0, s>>>(bl->full_mask.as<uint8_t>(), W,
                                             (size_t)W, counts.as<int>());
      XP_LAUNCHED();
      xp::rle_scan_kernel<<<1, 1024, 0, s>>>(counts.as<int>(), H,
                                             bl->rle_offs.as<int>());
      XP_LAUNCHED();
      int total = 0;
      XP_CUDA(cudaMemcpyAsync(&total, bl->rle_offs.as<int>() + H, sizeof(int),
                              cudaMemcpyDeviceToHost, s));
      XP_CUDA(cudaStreamSynchronize(s));
      bl->rle_words = (size_t)total;
      xp::DevBuf starts;
      starts.reserve(sizeof(uint32_t) * std::max<size_t>(bl->rle_words, 1));
      bl->rle_runs.reserve(sizeof(uint32_t) *
                           std::max<size_t>(bl->rle_words, 1));
      xp::rle_emit_kernel<<<H, 256, 0, s>>>(bl->full_mask.as<uint8_t>(), W,
                                            (size_t)W, bl->rle_offs.as<int>(),
                                            starts.as<uint32_t>());
      XP_LAUNCHED();
      xp::rle_lengths_kernel<<<dim3(4, H), 128, 0, s>>>(
          bl->rle_runs.as<uint32_t>(), bl->rle_offs.as<int>(), H, W,
          starts.as<uint32_t>());
      XP_LAUNCHED();
      XP_CUDA(cudaStreamSynchronize(s));  // `starts`, `counts` are released
      bl->rle_ready = true;
    }
    *n_words = bl->rle_words;
    if (!runs) return;  // size query
    XP_REQUIRE(cap_words >= bl->rle_words,
               "fetch_mask_rle: buffer too small (%zu < %zu words)", cap_words,
               bl->rle_words);
    const cudaMemcpyKind kind = mem == XP_MEM_DEVICE
                                    ? cudaMemcpyDeviceToDevice
                                    : cudaMemcpyDeviceToHost;
    if (bl->rle_words)
      XP_CUDA(cudaMemcpyAsync(runs, bl->rle_runs.p,
                              sizeof(uint32_t) * bl->rle_words, kind, s));
    if (row_offsets) {
      // Flex::rows_: byte offset of every row's first word
      std::vector<int> offs((size_t)H + 1);
      XP_CUDA(cudaMemcpyAsync(offs.data(), bl->rle_offs.p,
                              sizeof(int) * ((size_t)H + 1),
                              cudaMemcpyDeviceToHost, s));
      XP_CUDA(cudaStreamSynchronize(s));
      XP_REQUIRE(mem == XP_MEM_HOST, "row_offsets are returned to the host");
      for (int y = 0; y < H; ++y) row_offsets[y] = offs[y] * 4;
    }
    XP_CUDA(cudaStreamSynchronize(s));
  });
}

}  // extern "C"
