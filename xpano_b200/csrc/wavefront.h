// Exact parallel evaluation of raster-scan recurrences (multiblend's seam
// distance transforms and its inpaint fill) as a skewed wavefront.
//
// The reference runs these as single-threaded raster loops
// (external/multiblend/src/multiblend.cpp:271-438, 493-815 and
// src/image.cpp:550-812): pixel (q, r) in processing order depends on
// (q-1, r) in its own row and on (q-1, r-1), (q, r-1), (q+1, r-1) in the
// previously processed row. Nothing about the recurrence is changed here;
// only the order of evaluation:
//
//   * a warp owns a band of 32 consecutive rows, one row per lane;
//   * lane j runs 2 pixels behind lane j-1, so at every step the value it needs
//     from the row above was produced by its neighbour lane one step earlier
//     and arrives by a single shuffle; its own-row predecessor is a register;
//   * bands are claimed in order from an atomic counter; lane 0 of a band
//     reads the last row of the previous band from global memory behind a
//     monotone progress counter (one-directional: a band never waits for a
//     later band, so any number of resident warps makes progress);
//   * inputs and outputs move through per-warp shared-memory tiles so global
//     traffic is coalesced along rows although lanes work on 32 different rows.
//
// Total depth is (W + 2H) steps instead of W*H. The code is written against a
// small warp-context interface so the identical source runs on the device and
// under a host SIMT emulator (tests/emu) used for CPU-side verification.
#pragma once

#include <cstddef>
#include <cstdint>

#include "hd.h"

namespace xp {
namespace wf {

constexpr int kLanes = 32;
constexpr int kChunk = 32;          // steps per tile
constexpr int kTilePitch = kChunk + 1;
constexpr int kSkew = 2;            // pixels a lane trails the lane above it

struct Band {
  int job;    // which image (inpaint) / 0 (seams)
  int rho0;   // first row of the band, in processing order
  int W, H;   // extent of the job in processing order
  int dep;    // global index of the band that owns row rho0-1, or -1
};

// Shared-memory bytes one warp needs for problem P.
template <class P>
constexpr size_t tile_bytes() {
  return (sizeof(typename P::In) + sizeof(uint64_t)) * kLanes * kTilePitch;
}

// Runs problem P over all its bands. `counter` is a zero-initialised claim
// counter, `progress[b]` (zero-initialised) counts the positions of the last
// row of band b that are final in global memory.
template <class Ctx, class P>
XP_HD void run(Ctx& ctx, const P& prob, int* counter, int* progress,
               unsigned char* smem_warp) {
  using In = typename P::In;
  using Val = uint64_t;
  const int lane = ctx.lane();
  In* in_tile = reinterpret_cast<In*>(smem_warp);
  Val* out_tile =
      reinterpret_cast<Val*>(smem_warp + sizeof(In) * kLanes * kTilePitch);

  for (;;) {
    int band_idx = 0;
    if (lane == 0) band_idx = ctx.atomic_add(counter, 1);
    band_idx = ctx.bcast(band_idx);
    if (band_idx >= prob.n_bands) break;
    const Band B = prob.band(band_idx);
    const int W = B.W;
    const int rho = B.rho0 + lane;
    const bool row_ok = rho < B.H;
    typename P::Row row_state = prob.row_init(B, rho);
    const int n_chunks = (W + kSkew * (kLanes - 1) + kChunk - 1) / kChunk;

    Val last = 0, w0 = 0, w1 = 0, w2 = 0, pv = 0;
    if (B.dep >= 0) {
      // lane 0 starts with (q=0, rho0-1) in the slot that shifts into w1
      ctx.wait_ge(progress + B.dep, W < 1 ? W : 1);
      if (lane == 0) w2 = prob.load_val(B, B.rho0 - 1, 0);
    }

    for (int c = 0; c < n_chunks; ++c) {
      const int s0 = c * kChunk;
      // ---- stage inputs: row r, positions s0 - 2r + lane ----
      for (int r = 0; r < kLanes; ++r) {
        const int q = s0 - kSkew * r + lane;
        if (B.rho0 + r < B.H && q >= 0 && q < W)
          in_tile[r * kTilePitch + lane] = prob.load_in(B, B.rho0 + r, q);
      }
      if (B.dep >= 0) {
        int need = s0 + kChunk + 1;
        if (need > W) need = W;
        ctx.wait_ge(progress + B.dep, need);
        const int pos = s0 + 1 + lane;
        pv = pos < W ? prob.load_val(B, B.rho0 - 1, pos) : 0;
      }
      ctx.sync();
      // ---- 32 skewed steps ----
      for (int t = 0; t < kChunk; ++t) {
        const int q = s0 + t - kSkew * lane;
        Val up = ctx.shfl_up1(last);
        const Val hv = ctx.shfl_idx(pv, t);
        if (lane == 0) up = hv;
        w0 = w1;
        w1 = w2;
        w2 = up;
        if (row_ok && q >= 0 && q < W) {
          const In in = in_tile[lane * kTilePitch + t];
          last = prob.step(B, row_state, rho, q, in, last, w0, w1, w2);
          out_tile[lane * kTilePitch + t] = last;
        }
      }
      ctx.sync();
      // ---- flush outputs, coalesced per row ----
      for (int r = 0; r < kLanes; ++r) {
        const int q = s0 - kSkew * r + lane;
        if (B.rho0 + r < B.H && q >= 0 && q < W)
          prob.store(B, B.rho0 + r, q, out_tile[r * kTilePitch + lane]);
      }
      ctx.sync();
      if (lane == 0) {
        int done = s0 + kChunk - kSkew * (kLanes - 1);
        if (done < 0) done = 0;
        if (done > W || c == n_chunks - 1) done = W;
        ctx.publish(progress + band_idx, done);
      }
    }
  }
}

}  // namespace wf
}  // namespace xp
