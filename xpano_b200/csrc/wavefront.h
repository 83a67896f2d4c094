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
//   * a block of three warps owns a band of 32 consecutive rows: the compute
//     warp carries the recurrence, one row per lane; a loader warp stages the
//     inputs and a storer warp drains the outputs, so that only the steps
//     themselves are on the critical path (hand-over through double-buffered
//     shared tiles and named barriers);
//   * lane j runs 3 pixels behind lane j-1, so at every step the value it needs
//     from the row above was produced by its neighbour lane two steps earlier
//     and arrives by a single shuffle that overlaps the step in between (2
//     pixels would do, but put the shuffle latency on the chain of every
//     step); its own-row predecessor is a register;
//   * bands are claimed in order from an atomic counter; lane 0 of a band
//     reads the last row of the previous band from global memory, polling
//     the values themselves: outputs start as a sentinel no recurrence can
//     produce, so a non-sentinel word is final (one-directional: a band never
//     waits for a later band, so any number of resident warps makes progress);
//   * inputs move global -> shared with cp.async one 32-step chunk ahead
//     (double-buffered tiles, coalesced along rows although the lanes work on
//     32 different rows); outputs leave through a shared tile the same way;
//   * a fully parallel classification pass (classify_tile) runs first: pixels
//     whose value follows from their own input alone (a seam seed, an opaque
//     pixel) are written there, every other output becomes the sentinel, and
//     each 32x32 skewed tile that consists of such pixels only is flagged;
//   * nothing propagates through a flagged tile, so every maximal run of
//     unflagged tiles of a band is a work item of its own (segment_count /
//     segment_emit build the list, band-major): its lane state comes from
//     global memory (all of it written by the classification pass) and it
//     depends only on items of the previous band, which have smaller indices.
//     Items are claimed in order from an atomic counter, so whatever an item
//     waits for is running or done.
//
// Total depth is at most (W + 3H) steps instead of W*H. The code is written against a
// small warp-context interface so the identical source runs on the device and
// under a host SIMT emulator (tests/emu) used for CPU-side verification.
#pragma once

#include <cstddef>
#include <cstdint>
#include <type_traits>

#include "hd.h"

// phase timing hooks (development builds only)
#ifndef XP_WF_CLOCK
#define XP_WF_CLOCK(i)
#define XP_WF_CLOCK_BEGIN()
#define XP_WF_CLOCK_END(band, nb, fast, slow)
#endif

namespace xp {
namespace wf {

constexpr int kLanes = 32;
constexpr int kChunk = 32;          // steps per tile
// Row pitch of the shared tiles: a multiple of four 32-bit words, so a lane
// reads / writes its row 16 bytes at a time, and 4 mod 32 words, so that the
// eight lanes of a quarter warp hit distinct 16-byte bank groups (4-byte
// elements; two-way conflicts for 8-byte ones) while the loader and storer
// warps, which walk rows, stay conflict-free.
constexpr int kTilePitch = kChunk + 4;
constexpr int kTileElems = kLanes * kTilePitch;
constexpr int kSkew = 3;            // pixels a lane trails the lane above it

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

XP_HD int chunks_of(int W) {
  return (W + kSkew * (kLanes - 1) + kChunk - 1) / kChunk;
}

// 16-byte shared-memory accesses of a lane's run of tile elements (src / dst
// 16-byte aligned, N * sizeof(T) a multiple of 16)
struct alignas(16) Quad {
  uint32_t w[4];
};
template <class T, int N>
XP_HD void load_group(const T* src, T (&dst)[N]) {
  constexpr int kQuads = (int)(N * sizeof(T) / 16);
  static_assert(kQuads * 16 == N * sizeof(T), "whole quads");
  union {
    Quad q[kQuads];
    T v[N];
  } u;
  for (int i = 0; i < kQuads; ++i)
    u.q[i] = reinterpret_cast<const Quad*>(src)[i];
  for (int i = 0; i < N; ++i) dst[i] = u.v[i];
}
template <class T, int N>
XP_HD void store_group(T* dst, const T (&src)[N]) {
  constexpr int kQuads = (int)(N * sizeof(T) / 16);
  union {
    Quad q[kQuads];
    T v[N];
  } u;
  for (int i = 0; i < N; ++i) u.v[i] = src[i];
  for (int i = 0; i < kQuads; ++i) reinterpret_cast<Quad*>(dst)[i] = u.q[i];
}

XP_HD int popc32(unsigned v) {
#if defined(__CUDA_ARCH__)
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}

constexpr int kWarpsPerBand = 3;  // compute, loader, storer
enum Role { kCompute = 0, kLoader = 1, kStorer = 2 };
// hand-over signals (one waiter, one signaller each), per buffer parity
// input tiles are staged kInBufs deep: one tile load (32-64 cp.async and a
// trip to L2 / HBM) takes longer than the steps of a chunk, so a band that
// waits for nobody (the first one, which every other band trails) would run
// at the loader's pace with two buffers
constexpr int kInBufs = 4;
constexpr int kFlight = 2;  // tile loads in flight behind the newest one
enum Signal {
  kInFull = 0,
  kInFree = kInBufs,
  kOutFull = 2 * kInBufs,
  kOutFree = 2 * kInBufs + 2,
  kSignals = 2 * kInBufs + 4
};

// Shared-memory bytes one band (block) needs for problem P: kInBufs buffers
// of each input tile, two of the output tile, one row of hand-over values.
template <class P>
constexpr size_t tile_bytes() {
  return (kInBufs * (StreamBytes<typename P::In0>::value +
                     StreamBytes<typename P::In1>::value) +
          2 * sizeof(typename P::Val)) *
             kTileElems +
         sizeof(typename P::Val) * kChunk;
}

// tile flags of the classification pass
enum TileKind : uint8_t {
  kTileGeneral = 0,
  kTileTrivial = 1,    // every pixel follows from its own input: skipped
  kTileUncovered = 2,  // P::uncovered() for every pixel: cheaper step variant
};

// A work item of the wavefront: the unflagged chunks [c0, c1) of one band.
struct Item {
  int band, c0, c1;
};

// Upper bound of the number of items of a problem with `n_bands` bands of at
// most `flag_pitch` chunks (runs are separated by at least one flagged chunk).
XP_HD size_t max_items(int n_bands, int flag_pitch) {
  return (size_t)n_bands * (size_t)((flag_pitch + 1) / 2);
}

// Runs of unflagged chunks in one band, called by a whole warp. With
// items == nullptr it only counts; otherwise item k of the band is written
// to items[k] (starts and ends are ranked separately: the k-th end closes
// the k-th start). A lane reads each flag once; its neighbours' flags come by
// shuffle, the flags across a 32-chunk window boundary as carried bits.
template <class Ctx>
XP_HD int segment_band(Ctx& ctx, const uint8_t* band_flags, int n_chunks,
                       int band_idx, Item* items) {
  const int lane = ctx.lane();
  int n_start = 0, n_end = 0;
  // closed = trivial or outside the band; window w + 1 is loaded one ahead
  auto closed_mask = [&](int base) -> unsigned {
    const int c = base + lane;
    return ctx.ballot(c >= n_chunks || band_flags[c] == kTileTrivial);
  };
  unsigned prev_last = 1;  // chunk -1 counts as closed
  constexpr int kBatch = 8;  // windows whose flags are loaded together
  unsigned cur = closed_mask(0);
  for (int b0 = 0; b0 < n_chunks; b0 += 32 * kBatch) {
    bool closed[kBatch];
    for (int w = 0; w < kBatch; ++w) {  // independent loads, one latency
      const int c = b0 + 32 * (w + 1) + lane;
      closed[w] = c >= n_chunks || band_flags[c] == kTileTrivial;
    }
    for (int w = 0; w < kBatch; ++w) {
      const int base = b0 + 32 * w;
      const unsigned next = ctx.ballot(closed[w]);
      if (base >= n_chunks) continue;
      const unsigned open = ~cur;
      const unsigned prev_closed = (cur << 1) | prev_last;
      const unsigned next_closed = (cur >> 1) | (next << 31);
      const unsigned bs = open & prev_closed, be = open & next_closed;
      const unsigned below = (1u << lane) - 1u, me = 1u << lane;
      if (items) {
        const int c = base + lane;
        if (bs & me) {
          Item& it = items[n_start + popc32(bs & below)];
          it.band = band_idx;
          it.c0 = c;
        }
        if (be & me) items[n_end + popc32(be & below)].c1 = c + 1;
      }
      n_start += popc32(bs);
      n_end += popc32(be);
      prev_last = cur >> 31;
      cur = next;
    }
  }
  return n_start;
}

// Classification pass, one warp per (band, chunk) tile: writes the value of
// every self-determined pixel, the sentinel everywhere else, and the tile's
// "nothing to propagate here" flag.
template <class Ctx, class P>
XP_HD void classify_tile(Ctx& ctx, const P& prob, int band_idx, int c,
                         uint8_t* flags, int flag_pitch) {
  using In0 = typename P::In0;
  using In1 = typename P::In1;
  constexpr size_t kB1 = StreamBytes<In1>::value;
  const int lane = ctx.lane();
  const Band B = prob.band(band_idx);
  if (c >= chunks_of(B.W)) return;
  const typename P::Row row = prob.row_init(B, B.rho0 + lane);
  const int s0 = c * kChunk;
  bool all_trivial = true, all_uncovered = P::kHasUncovered;
  for (int r = 0; r < kLanes; ++r) {
    const int q = s0 - kSkew * r + lane;
    if (B.rho0 + r < B.H && q >= 0 && q < B.W) {
      const In0 a = *prob.in0_ptr(B, B.rho0 + r, q);
      In1 b = In1();
      if (kB1) b = *static_cast<const In1*>(prob.in1_ptr(B, B.rho0 + r, q));
      const bool t = prob.trivial(a, b);
      prob.store(B, B.rho0 + r, q,
                 t ? prob.trivial_value(B, row, r - lane, q, a, b)
                   : prob.sentinel());
      all_trivial = all_trivial && t;
      if constexpr (P::kHasUncovered)
        all_uncovered = all_uncovered && prob.uncovered(a, b);
    }
  }
  all_trivial = ctx.all(all_trivial);
  all_uncovered = ctx.all(all_uncovered);
  if (lane == 0)
    flags[(size_t)band_idx * flag_pitch + c] =
        all_trivial && prob.allow_trivial(B) ? kTileTrivial
        : all_uncovered                      ? kTileUncovered
                                             : kTileGeneral;
}

// Runs problem P over all its work items. `counter` is a zero-initialised
// claim counter; `items` / `*n_items` come from segment_band over `flags`,
// the tile flags of classify_tile (flag_pitch per band). Called by all three warps of a block with the same
// arguments.
template <class Ctx, class P>
XP_HD void run(Ctx& ctx, const P& prob, int* counter, const Item* items,
               const int* n_items, const uint8_t* flags, int flag_pitch,
               unsigned char* smem_block) {
  using In0 = typename P::In0;
  using In1 = typename P::In1;
  using Val = typename P::Val;
  constexpr size_t kB0 = StreamBytes<In0>::value;
  constexpr size_t kB1 = StreamBytes<In1>::value;
  const int lane = ctx.lane();
  const int role = ctx.role();
  const Val kSentinel = prob.sentinel();
  // layout: out[2] | hand-over row | in0[kInBufs] | in1[kInBufs], widest
  // type first
  Val* out_base = reinterpret_cast<Val*>(smem_block);
  Val* pv_row = out_base + 2 * kTileElems;
  unsigned char* base_a = reinterpret_cast<unsigned char*>(pv_row + kChunk);
  unsigned char* in0_base =
      kB0 >= kB1 ? base_a : base_a + kInBufs * kB1 * kTileElems;
  unsigned char* in1_base =
      kB0 >= kB1 ? base_a + kInBufs * kB0 * kTileElems : base_a;
  const int total_items = *n_items;

  for (;;) {
    int item_idx = 0;
    if (role == kCompute && lane == 0) item_idx = ctx.atomic_add(counter, 1);
    item_idx = ctx.block_bcast(item_idx);
    if (item_idx >= total_items) break;
    const Item item = items[item_idx];
    const Band B = prob.band(item.band);
    const int W = B.W;

    if (role == kLoader) {
      // ---- inputs of the k-th chunk of the item -> buffer k % kInBufs: row r,
      // positions c*32 - kSkew*r + lane ----
      int k = 0;
      for (int c = item.c0; c < item.c1; ++c) {
        if (k >= kInBufs) ctx.wait(kInFree + (k % kInBufs));
        const int s0 = c * kChunk;
        In0* t0 = reinterpret_cast<In0*>(in0_base +
                                        (k % kInBufs) * kB0 * kTileElems);
        In1* t1 = reinterpret_cast<In1*>(in1_base +
                                        (k % kInBufs) * kB1 * kTileElems);
#if defined(__CUDA_ARCH__)
#pragma unroll 8
#endif
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
        // kFlight tiles stay in flight behind the one just issued
        if (k >= kFlight) {
          ctx.template cp_wait_pending<kFlight>();
          ctx.signal(kInFull + ((k - kFlight) % kInBufs));
        }
        ++k;
      }
      ctx.template cp_wait_pending<0>();
      for (int j = k < kFlight ? 0 : k - kFlight; j < k; ++j)
        ctx.signal(kInFull + (j % kInBufs));
      // the compute warp frees every buffer; absorb the last kInBufs
      for (int j = k < kInBufs ? 0 : k - kInBufs; j < k; ++j)
        ctx.wait(kInFree + (j % kInBufs));
    } else if (role == kStorer) {
      // ---- outputs of the k-th chunk, coalesced per row ----
      int k = 0;
      for (int c = item.c0; c < item.c1; ++c) {
        ctx.wait(kOutFull + (k & 1));
        const int s0 = c * kChunk;
        const Val* out_tile = out_base + (k & 1) * kTileElems;
#if defined(__CUDA_ARCH__)
#pragma unroll 8
#endif
        for (int r = kLanes - 1; r >= 0; --r) {  // hand-over row first
          const int q = s0 - kSkew * r + lane;
          if (B.rho0 + r < B.H && q >= 0 && q < W)
            prob.store(B, B.rho0 + r, q, out_tile[r * kTilePitch + lane]);
        }
        ctx.signal(kOutFree + (k & 1));
        ++k;
      }
    } else {
      // ---- the recurrence ----
      const int rho = B.rho0 + lane;
      const bool row_ok = rho < B.H;
      typename P::Row row_state = prob.row_init(B, rho);
      // previous band's last row at position pos (lane-private), once final
      auto poll = [&](bool want, int pos) -> Val {
        const bool valid = want && pos >= 0 && pos < W;
        const Val* p = prob.val_ptr(B, B.rho0 - 1, valid ? pos : 0);
        Val v = 0;
        for (;;) {
          if (valid) v = ctx.ld_poll(p);
          if (ctx.all(!valid || v != kSentinel)) break;
          ctx.backoff();
        }
        return v;
      };
      auto in_row = [&](int pos) -> bool { return pos >= 0 && pos < W; };
      // one load of the previous band's last row, final or not
      auto peek = [&](int pos) -> Val {
        const bool valid = in_row(pos);
        const Val* p = prob.val_ptr(B, B.rho0 - 1, valid ? pos : 0);
        return valid ? ctx.ld_poll(p) : Val(0);
      };
      Val pv_ahead = 0;
      bool have_ahead = false;
      // `older` is `last` one step back: with a skew of 3 the lane below needs
      // it a full step after it was formed, so the shuffle is off the chain
      Val last = 0, older = 0, w0 = 0, w1 = 0, w2 = 0;
      XP_WF_CLOCK_BEGIN();
#ifdef XP_WF_STATS
      if (lane == 0) XP_WF_STATS(item.c1 - item.c0);
#endif
      {
        // Lane state at the end of chunk c0-1: all of its pixels in the rows
        // of this band were written by the classification pass (the chunk is
        // flagged, or lies before the first column). A row state that evolves
        // along the row (P::Row) is only kept by problems whose first band
        // is never flagged, i.e. starts at chunk 0.
        const int s0 = item.c0 * kChunk;
        const int qe = s0 - 1 - kSkew * lane;
        if (row_ok && qe >= 0 && qe < W) last = prob.load_val(B, rho, qe);
        if (row_ok && qe >= 1 && qe - 1 < W)
          older = prob.load_val(B, rho, qe - 1);
        if (row_ok && lane > 0) {
          if (qe >= 0 && qe < W) w1 = prob.load_val(B, rho - 1, qe);
          if (qe + 1 >= 0 && qe + 1 < W)
            w2 = prob.load_val(B, rho - 1, qe + 1);
        }
        if (B.dep >= 0) {
          // lane 0's row above belongs to the previous band: lanes 0/1
          // fetch positions s0-1 and s0
          const Val pr = poll(lane < 2, s0 - 1 + lane);
          const Val p0 = ctx.shfl_idx(pr, 0), p1 = ctx.shfl_idx(pr, 1);
          if (lane == 0) w1 = p0, w2 = p1;
        }
      }
      int k = 0;
      const uint8_t* band_flags = flags + (size_t)item.band * flag_pitch;
      unsigned kind_word = 0;  // lane l: kind of chunk c0 + (k & ~31) + l
      for (int c = item.c0; c < item.c1; ++c) {
        const int s0 = c * kChunk;
        XP_WF_CLOCK(7);
        bool uncovered = false;
        if constexpr (P::kHasUncovered) {
          if ((k & 31) == 0)
            kind_word = c + lane < item.c1 ? band_flags[c + lane] : 0u;
          uncovered = ctx.shfl_idx(kind_word, k & 31) == kTileUncovered;
        }
        // previous band's last row, positions s0+1 .. s0+32 (one per lane).
        // The load was issued before the steps of the previous chunk (its
        // trip to L2 is as long as half a chunk of steps); only a value that
        // was not final yet is polled for.
        Val pv = 0;
        if (B.dep >= 0) {
          pv = pv_ahead;
          if (!have_ahead ||
              !ctx.all(!in_row(s0 + 1 + lane) || pv != kSentinel))
            pv = poll(true, s0 + 1 + lane);
        }
        XP_WF_CLOCK(1);
        pv_row[lane] = pv;
        ctx.sync();
        have_ahead = false;
        if (B.dep >= 0 && c + 1 < item.c1) {
          pv_ahead = peek(s0 + kChunk + 1 + lane);
          have_ahead = true;
        }
        ctx.wait(kInFull + (k % kInBufs));
        if (k >= 2) ctx.wait(kOutFree + (k & 1));
        XP_WF_CLOCK(2);
        {
          // 32 skewed steps. Chunks that touch no border of the job (first /
          // last column, first row) run the step variant without the border
          // selects. Inputs are read from the shared tiles a group of steps
          // ahead and outputs are written a group at a time: a shared store
          // between two loads would pin every load behind it and put its
          // latency on the chain of every step.
          const In0* r0 = reinterpret_cast<const In0*>(
                              in0_base + (k % kInBufs) * kB0 * kTileElems) +
                          lane * kTilePitch;
          const In1* r1 = reinterpret_cast<const In1*>(
                              in1_base + (k % kInBufs) * kB1 * kTileElems) +
                          lane * kTilePitch;
          Val* ro = out_base + (k & 1) * kTileElems + lane * kTilePitch;
          const int q0 = s0 - kSkew * lane;
          // rows past H (lanes below the job) may run the border-free step
          // too: nothing depends on them and the storer never writes them
          const bool interior_x = s0 - kSkew * (kLanes - 1) >= 2 &&
                                  s0 + kChunk - 1 <= W - 2;
          auto steps = [&](auto interior_tag, auto edge_y, auto uncov_tag) {
            constexpr bool kInterior = decltype(interior_tag)::value;
            constexpr bool kEdgeY = decltype(edge_y)::value;
            constexpr bool kUncovered = decltype(uncov_tag)::value;
            if constexpr (kUncovered) {
              w1 = prob.uncovered_up(w1);
              w2 = prob.uncovered_up(w2);
              last = prob.uncovered_last(last);
            }
            constexpr int kGroup = 8;
            constexpr int kGroups = kChunk / kGroup;
            In0 ga[2][kGroup];
            In1 gb[2][kGroup];
            Val gh[2][kGroup];
            auto fetch = [&](int g, int slot) {
              load_group(r0 + g * kGroup, ga[slot]);
              if constexpr (kB1 > 0) load_group(r1 + g * kGroup, gb[slot]);
              load_group(pv_row + g * kGroup, gh[slot]);
            };
            fetch(0, 0);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int g = 0; g < kGroups; ++g) {
              if (g + 1 < kGroups) fetch(g + 1, (g + 1) & 1);
              Val go[kGroup];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
              for (int i = 0; i < kGroup; ++i) {
                const int t = g * kGroup + i;
                const In0 a = ga[g & 1][i];
                In1 b = In1();
                if (kB1) b = gb[g & 1][i];
                Val up = ctx.shfl_up1(older);
                if (lane == 0) up = gh[g & 1][i];
                if constexpr (kUncovered) up = prob.uncovered_up(up);
                older = last;
                w0 = w1;
                w1 = w2;
                w2 = up;
                if constexpr (kUncovered) {
                  last = prob.template step_uncovered<kEdgeY>(
                      row_state, rho, a, b, last, w0, w1, w2);
                } else if (kInterior) {
                  last = prob.template step<false, kEdgeY>(
                      B, row_state, true, rho, q0 + t, a, b, last, w0, w1, w2);
                } else {
                  const int q = q0 + t;
                  const bool active = row_ok && q >= 0 && q < W;
                  const Val nv = prob.template step<true, true>(
                      B, row_state, active, rho, q, a, b, last, w0, w1, w2);
                  if (active) last = nv;
                }
                go[i] = last;  // inactive positions are never stored
              }
              store_group(ro + g * kGroup, go);
            }
          };
          auto interior_steps = [&](auto edge_y) {
            if constexpr (P::kHasUncovered) {
              if (uncovered) {
                steps(std::true_type(), edge_y, std::true_type());
                return;
              }
            }
            steps(std::true_type(), edge_y, std::false_type());
          };
          if (interior_x && B.rho0 > 0) {
            interior_steps(std::false_type());
          } else if (interior_x) {
            interior_steps(std::true_type());
          } else {
            steps(std::false_type(), std::true_type(), std::false_type());
          }
        }
        XP_WF_CLOCK(4);
        ctx.signal(kOutFull + (k & 1));
        ctx.signal(kInFree + (k % kInBufs));
        ++k;
      }
      // the storer frees every output tile; absorb the last two
      for (int j = k < 2 ? 0 : k - 2; j < k; ++j) ctx.wait(kOutFree + (j & 1));
      XP_WF_CLOCK_END(item.band, prob.n_bands, item.c0, item.c1);
    }
    ctx.block_sync();  // tiles and signals are quiet before the next item
  }
}

}  // namespace wf
}  // namespace xp
