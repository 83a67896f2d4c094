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
//   * inputs move global -> shared with cp.async one 32-step chunk ahead
//     (double-buffered tiles, coalesced along rows although the lanes work on
//     32 different rows); outputs leave through a shared tile the same way.
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
constexpr int kTileElems = kLanes * kTilePitch;
constexpr int kSkew = 2;            // pixels a lane trails the lane above it

struct Band {
  int job;    // which image (inpaint) / 0 (seams)
  int rho0;   // first row of the band, in processing order
  int W, H;   // extent of the job in processing order
  int dep;    // global index of the band that owns row rho0-1, or -1
};

struct NoInput {};  // second input stream of single-stream problems

template <class T>
struct StreamBytes {
  static constexpr size_t value = sizeof(T);
};
template <>
struct StreamBytes<NoInput> {
  static constexpr size_t value = 0;
};

// Shared-memory bytes one warp needs for problem P: two buffers of each input
// tile, one output tile and one row of hand-over values.
template <class P>
constexpr size_t tile_bytes() {
  return (2 * (StreamBytes<typename P::In0>::value +
               StreamBytes<typename P::In1>::value) +
          sizeof(typename P::Val)) *
             kTileElems +
         sizeof(typename P::Val) * kChunk;
}

// Runs problem P over all its bands. `counter` is a zero-initialised claim
// counter, `progress[b]` (zero-initialised) counts the positions of the last
// row of band b that are final in global memory.
template <class Ctx, class P>
XP_HD void run(Ctx& ctx, const P& prob, int* counter, int* progress,
               unsigned char* smem_warp) {
  using In0 = typename P::In0;
  using In1 = typename P::In1;
  using Val = typename P::Val;
  constexpr size_t kB0 = StreamBytes<In0>::value;
  constexpr size_t kB1 = StreamBytes<In1>::value;
  const int lane = ctx.lane();
  // layout: out tile | hand-over row | in0[2] | in1[2], widest type first
  Val* out_tile = reinterpret_cast<Val*>(smem_warp);
  Val* pv_row = out_tile + kTileElems;
  unsigned char* base_a = reinterpret_cast<unsigned char*>(pv_row + kChunk);
  unsigned char* in0_base =
      kB0 >= kB1 ? base_a : base_a + 2 * kB1 * kTileElems;
  unsigned char* in1_base =
      kB0 >= kB1 ? base_a + 2 * kB0 * kTileElems : base_a;

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

    // stage the inputs of chunk c into buffer c & 1: row r, positions
    // c*32 - 2r + lane
    auto prefetch = [&](int c) {
      const int s0 = c * kChunk;
      In0* t0 = reinterpret_cast<In0*>(in0_base + (c & 1) * kB0 * kTileElems);
      In1* t1 = reinterpret_cast<In1*>(in1_base + (c & 1) * kB1 * kTileElems);
      for (int r = 0; r < kLanes; ++r) {
        const int q = s0 - kSkew * r + lane;
        if (B.rho0 + r < B.H && q >= 0 && q < W) {
          ctx.cp_async(t0 + r * kTilePitch + lane,
                       prob.in0_ptr(B, B.rho0 + r, q), (int)kB0);
          if (kB1)
            ctx.cp_async(t1 + r * kTilePitch + lane,
                         prob.in1_ptr(B, B.rho0 + r, q), (int)kB1);
        }
      }
      ctx.cp_commit();
    };

    Val last = 0, w0 = 0, w1 = 0, w2 = 0, pv = 0, pv_next = 0;
    bool have_next = false;
    prefetch(0);
    if (B.dep >= 0) {
      // lane 0 starts with (q=0, rho0-1) in the slot that shifts into w1
      ctx.wait_ge(progress + B.dep, W < 1 ? W : 1);
      if (lane == 0) w2 = prob.load_val(B, B.rho0 - 1, 0);
    }

    for (int c = 0; c < n_chunks; ++c) {
      const int s0 = c * kChunk;
      if (c + 1 < n_chunks) prefetch(c + 1);
      // previous band's last row, positions s0+1 .. s0+32 (one per lane)
      if (B.dep >= 0) {
        if (have_next) {
          pv = pv_next;
        } else {
          int need = s0 + kChunk + 1;
          if (need > W) need = W;
          ctx.wait_ge(progress + B.dep, need);
          const int pos = s0 + 1 + lane;
          pv = pos < W ? prob.load_val(B, B.rho0 - 1, pos) : 0;
        }
        // opportunistically fetch the next chunk's row now if it is ready
        have_next = false;
        if (c + 1 < n_chunks) {
          int need = s0 + 2 * kChunk + 1;
          if (need > W) need = W;
          if (ctx.peek(progress + B.dep) >= need) {
            const int pos = s0 + kChunk + 1 + lane;
            pv_next = pos < W ? prob.load_val(B, B.rho0 - 1, pos) : 0;
            have_next = true;
          }
        }
      }
      pv_row[lane] = pv;
      if (c + 1 < n_chunks)
        ctx.cp_wait_all_but_one();
      else
        ctx.cp_wait_all();
      ctx.sync();
      const In0* t0 =
          reinterpret_cast<const In0*>(in0_base + (c & 1) * kB0 * kTileElems);
      const In1* t1 =
          reinterpret_cast<const In1*>(in1_base + (c & 1) * kB1 * kTileElems);
      // ---- fast path: every pixel of the chunk is determined by its own
      // input alone (a seed / an opaque pixel), so the tile is filled in
      // parallel and the lane state is rebuilt from it ----
      bool trivial = prob.allow_trivial(B);
      if (trivial) {
        for (int r = 0; r < kLanes; ++r) {
          const int q = s0 - kSkew * r + lane;
          if (B.rho0 + r < B.H && q >= 0 && q < W) {
            In1 b = In1();
            if (kB1) b = t1[r * kTilePitch + lane];
            trivial = trivial && prob.trivial(t0[r * kTilePitch + lane], b);
          }
        }
      }
      trivial = ctx.all(trivial);
#ifdef XP_WF_STATS
      if (lane == 0) XP_WF_STATS(trivial);
#endif
      if (trivial) {
        for (int r = 0; r < kLanes; ++r) {
          const int q = s0 - kSkew * r + lane;
          if (B.rho0 + r < B.H && q >= 0 && q < W) {
            In1 b = In1();
            if (kB1) b = t1[r * kTilePitch + lane];
            out_tile[r * kTilePitch + lane] = prob.trivial_value(
                B, row_state, r - lane, q, t0[r * kTilePitch + lane], b);
          }
        }
        ctx.sync();
        // own row: the last position this lane has produced so far
        int t_last = W - 1 - s0 + kSkew * lane;
        if (t_last > kChunk - 1) t_last = kChunk - 1;
        if (row_ok && t_last >= 0 && s0 + t_last - kSkew * lane >= 0)
          last = out_tile[lane * kTilePitch + t_last];
        // row above at positions q_e and q_e + 1 (columns 29, 30 of lane-1)
        if (lane > 0) {
          w1 = out_tile[(lane - 1) * kTilePitch + kChunk - 3];
          w2 = out_tile[(lane - 1) * kTilePitch + kChunk - 2];
        }
        const Val p30 = ctx.shfl_idx(pv, kChunk - 2);
        const Val p31 = ctx.shfl_idx(pv, kChunk - 1);
        if (lane == 0) {
          w1 = p30;
          w2 = p31;
        }
      } else {
        // ---- 32 skewed steps; everything but the chain through `last` and
        // the shuffle is independent of the previous step ----
        const In0* r0 = t0 + lane * kTilePitch;
        const In1* r1 = t1 + lane * kTilePitch;
        Val* ro = out_tile + lane * kTilePitch;
        const int q0 = s0 - kSkew * lane;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int t = 0; t < kChunk; ++t) {
          const int q = q0 + t;
          const bool active = row_ok && q >= 0 && q < W;
          const In0 a = r0[t];
          In1 b = In1();
          if (kB1) b = r1[t];
          const Val hv = pv_row[t];
          Val up = ctx.shfl_up1(last);
          if (lane == 0) up = hv;
          w0 = w1;
          w1 = w2;
          w2 = up;
          const Val nv =
              prob.step(B, row_state, active, rho, q, a, b, last, w0, w1, w2);
          if (active) {
            last = nv;
            ro[t] = nv;
          }
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
        if (done > 0) ctx.publish(progress + band_idx, done);
      }
    }
  }
}

}  // namespace wf
}  // namespace xp
