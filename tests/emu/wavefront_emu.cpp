// TEST INFRASTRUCTURE ONLY: host SIMT emulator for the wavefront kernels.
//
// Compiles the product's device-agnostic sources (xpano_b200/csrc/wavefront.h,
// scan_problems.h) with g++ and runs them with one OS thread per lane, 32
// lanes per emulated warp, warp shuffles and __syncwarp implemented with
// barriers, progress counters with atomics. This lets the CPU test suite
// (-m "not gpu") check the exact code the GPU runs -- skew arithmetic, tile
// staging, band hand-over -- against the oracle without a device.
//
// Build: g++ -O2 -std=c++20 -pthread -shared -fPIC (tests/test_emu_wavefront.py)
#include <atomic>
#include <barrier>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

static std::atomic<long> g_fast_chunks{0}, g_slow_chunks{0}, g_uncovered_chunks{0};
// stepped chunks are reported per work item by wf::run; flagged chunks (never
// visited by the wavefront) are counted by run_emulated
#define XP_WF_STATS(stepped) (g_slow_chunks += (stepped))

#include "../../xpano_b200/csrc/scan_problems.h"
#include "../../xpano_b200/csrc/wavefront.h"

using namespace xp;

namespace {

struct EmuWarp {
  std::barrier<> bar{32};
  uint64_t x[32];
};

struct EmuBlock {
  std::barrier<> bar{32 * wf::kWarpsPerBand};
  int slot = 0;
  std::atomic<int> sig[wf::kSignals];
  EmuWarp warps[wf::kWarpsPerBand];
  EmuBlock() {
    for (auto& s : sig) s.store(0);
  }
};

struct EmuCtx {
  EmuBlock* blk;
  EmuWarp* w;
  int role_, lane_;
  int seen[wf::kSignals] = {};
  int lane() const { return lane_; }
  int role() const { return role_; }
  int atomic_add(int* p, int v) {
    return __atomic_fetch_add(p, v, __ATOMIC_ACQ_REL);
  }
  uint64_t exchange(uint64_t v, int src) {
    w->x[lane_] = v;
    w->bar.arrive_and_wait();
    uint64_t r = (src >= 0 && src < 32) ? w->x[src] : v;
    w->bar.arrive_and_wait();
    return r;
  }
  int block_bcast(int v) {
    if (role_ == 0 && lane_ == 0) blk->slot = v;
    blk->bar.arrive_and_wait();
    const int r = blk->slot;
    blk->bar.arrive_and_wait();
    return r;
  }
  void block_sync() { blk->bar.arrive_and_wait(); }
  void signal(int id) {
    w->bar.arrive_and_wait();
    if (lane_ == 0) blk->sig[id].fetch_add(1, std::memory_order_release);
  }
  void wait(int id) {
    while (blk->sig[id].load(std::memory_order_acquire) <= seen[id])
      std::this_thread::yield();
    ++seen[id];
  }
  template <class T>
  T shfl_up1(T v) {
    return (T)exchange((uint64_t)v, lane_ - 1);
  }
  template <class T>
  T shfl_idx(T v, int src) {
    return (T)exchange((uint64_t)v, src & 31);
  }
  void sync() { w->bar.arrive_and_wait(); }
  bool all(bool v) {
    w->x[lane_] = v ? 1 : 0;
    w->bar.arrive_and_wait();
    bool r = true;
    for (int i = 0; i < 32; ++i) r = r && w->x[i] != 0;
    w->bar.arrive_and_wait();
    return r;
  }
  unsigned ballot(bool v) {
    w->x[lane_] = v ? 1 : 0;
    w->bar.arrive_and_wait();
    unsigned r = 0;
    for (int i = 0; i < 32; ++i) r |= w->x[i] ? 1u << i : 0u;
    w->bar.arrive_and_wait();
    return r;
  }
  template <class T>
  T ld_poll(const T* p) {
    return __atomic_load_n(p, __ATOMIC_RELAXED);
  }
  void backoff() { std::this_thread::yield(); }
  void cp_async(void* dst, const void* src, int bytes) {
    memcpy(dst, src, (size_t)bytes);
  }
  void cp_commit() {}
  template <int kPending>
  void cp_wait_pending() {}
};

template <class P>
void run_emulated(const P& prob, int n_blocks, int max_w) {
  if (prob.n_bands <= 0) return;
  if (n_blocks > prob.n_bands) n_blocks = prob.n_bands;
  int counter = 0;
  const int flag_pitch = wf::chunks_of(max_w);
  std::vector<uint8_t> flags((size_t)prob.n_bands * flag_pitch, 0);
  std::vector<std::unique_ptr<EmuBlock>> blocks;
  std::vector<std::vector<unsigned char>> smem(n_blocks);
  for (int i = 0; i < n_blocks; ++i) {
    blocks.emplace_back(new EmuBlock());
    smem[i].resize(wf::tile_bytes<P>());
  }
  // classification pass: warp 0 of block bi takes tiles bi, bi + n_blocks, ...
  {
    std::vector<std::thread> threads;
    for (int bi = 0; bi < n_blocks; ++bi)
      for (int l = 0; l < 32; ++l)
        threads.emplace_back([&, bi, l] {
          EmuCtx ctx{blocks[bi].get(), &blocks[bi]->warps[0], 0, l};
          const size_t tiles = (size_t)prob.n_bands * flag_pitch;
          for (size_t t = bi; t < tiles; t += n_blocks)
            wf::classify_tile(ctx, prob, (int)(t / flag_pitch),
                              (int)(t % flag_pitch), flags.data(), flag_pitch);
        });
    for (auto& t : threads) t.join();
  }
  // work items, as segment_kernel builds them (one emulated warp)
  std::vector<int> first(prob.n_bands + 1, 0);
  std::vector<wf::Item> items(wf::max_items(prob.n_bands, flag_pitch));
  {
    std::vector<std::thread> threads;
    for (int l = 0; l < 32; ++l)
      threads.emplace_back([&, l] {
        EmuCtx ctx{blocks[0].get(), &blocks[0]->warps[0], 0, l};
        int total = 0;
        for (int b = 0; b < prob.n_bands; ++b) {
          const int nc = wf::chunks_of(prob.band(b).W);
          if (l == 0) first[b] = total;
          total += wf::segment_band(ctx, flags.data() + (size_t)b * flag_pitch,
                                    nc, b, items.data() + total);
        }
        if (l == 0) first[prob.n_bands] = total;
      });
    for (auto& t : threads) t.join();
    for (int b = 0; b < prob.n_bands; ++b) {
      const int nc = wf::chunks_of(prob.band(b).W);
      for (int c = 0; c < nc; ++c)
      {
        const uint8_t f = flags[(size_t)b * flag_pitch + c];
        if (f == wf::kTileTrivial) ++g_fast_chunks;
        if (f == wf::kTileUncovered) ++g_uncovered_chunks;
      }
    }
  }
  const int n_items = first[prob.n_bands];
  std::vector<std::thread> threads;
  for (int bi = 0; bi < n_blocks; ++bi)
    for (int r = 0; r < wf::kWarpsPerBand; ++r)
      for (int l = 0; l < 32; ++l)
        threads.emplace_back([&, bi, r, l] {
          EmuCtx ctx{blocks[bi].get(), &blocks[bi]->warps[r], r, l};
          wf::run(ctx, prob, &counter, items.data(), &n_items, flags.data(),
                  flag_pitch, smem[bi].data());
        });
  for (auto& t : threads) t.join();
}

// mirrors cover_kernel in seams.cu
template <class C>
void build_cover(std::vector<C>& cover, int CW, int CH, int n, const int* xpos,
                 const int* ypos, const int* w, const int* h,
                 const uint8_t* const* opaque) {
  cover.assign((size_t)CW * CH, 0);
  for (int i = 0; i < n; ++i)
    for (int ly = 0; ly < h[i]; ++ly)
      for (int lx = 0; lx <= w[i]; ++lx) {
        int cx = xpos[i] + lx, cy = ypos[i] + ly;
        if (cx >= CW) continue;
        C add = 0;
        if (lx < w[i]) {
          const uint8_t* op = opaque[i] + (size_t)ly * w[i];
          if (op[lx]) add |= (C)1 << i;
          if (lx == 0 || (op[lx] != 0) != (op[lx - 1] != 0))
            add |= CoverTraits<C>::rs_bit;
        } else {
          add |= CoverTraits<C>::rs_bit;
        }
        cover[(size_t)cy * CW + cx] |= add;
      }
}

template <class C, int F>
void seams_t(int n, const int* xpos, const int* ypos, const int* w,
             const int* h, const uint8_t* const* opaque, int W, int H,
             int reverse, int32_t* labels, uint64_t* bwd, uint64_t* fwd,
             int n_warps) {
  std::vector<C> cover;
  build_cover(cover, W, H, n, xpos, ypos, w, h, opaque);
  if (F == kCoverDigest)  // mirrors cover_kernel's digest output
    for (C& c : cover)
      c = (C)make_digest((u32)(c & ~CoverTraits<C>::rs_bit),
                         (c & CoverTraits<C>::rs_bit) != 0, reverse);
  std::vector<uint8_t> lab8((size_t)W * H);
  std::vector<u32> b32((size_t)W * H), f32((size_t)W * H);
  const int n_bands = (H + 31) / 32;
  SeamBwd<C, F> pb{n_bands, W, H, cover.data(), b32.data()};
  run_emulated(pb, n_warps, W);
  SeamFwd<C, F> pf{n_bands, W,          H,          reverse,
                   cover.data(), b32.data(), f32.data(), lab8.data()};
  run_emulated(pf, n_warps, W);
  for (size_t i = 0; i < lab8.size(); ++i) {
    labels[i] = lab8[i];
    bwd[i] = seam_unpack(b32[i]);
    fwd[i] = seam_unpack(f32[i]);
  }
}

}  // namespace

extern "C" {

// chunks that took the parallel fast path / the stepped path since last call
void emu_chunk_stats(long* fast, long* slow) {
  *fast = g_fast_chunks.exchange(0);
  *slow = g_slow_chunks.exchange(0);
}

// chunks classified as uncovered (cheaper step variant) since last call
long emu_uncovered_chunks() { return g_uncovered_chunks.exchange(0); }

void emu_seams(int n, const int* xpos, const int* ypos, const int* w,
               const int* h, const uint8_t* const* opaque, int W, int H,
               int reverse, int wide, int32_t* labels, uint64_t* bwd,
               uint64_t* fwd, int n_warps) {
  // wide: 0 = u32 digest words (the product's choice up to 24 images),
  //       1 = u64 bit words, 2 = u32 bit words (25..31 images)
  if (wide == 1)
    seams_t<u64, kCoverBits>(n, xpos, ypos, w, h, opaque, W, H, reverse,
                             labels, bwd, fwd, n_warps);
  else if (wide == 2)
    seams_t<u32, kCoverBits>(n, xpos, ypos, w, h, opaque, W, H, reverse,
                             labels, bwd, fwd, n_warps);
  else
    seams_t<u32, kCoverDigest>(n, xpos, ypos, w, h, opaque, W, H, reverse,
                               labels, bwd, fwd, n_warps);
}

// Inpaints n images at once (the band table interleaves them like the
// product does). data[i]: 4-sample interleaved pixels, modified in place by
// gathering the final source indices; bbox[i*4..] = left, top, w, h.
void emu_inpaint(int n, void* const* data, const int* tiff_w, int bpp,
                 const int* bbox, int n_warps) {
  std::vector<InpaintJob> jobs(n);
  std::vector<std::vector<u64>> fwd(n), bwd(n);
  int max_bands = 0, max_w = 0;
  for (int i = 0; i < n; ++i) {
    const int* b = bbox + 4 * i;
    max_w = std::max(max_w, b[2]);
    fwd[i].assign((size_t)b[2] * b[3], 0);
    bwd[i].assign((size_t)b[2] * b[3], 0);
    jobs[i] = InpaintJob{data[i],       tiff_w[i],     b[0], b[1], b[2],
                         b[3],          fwd[i].data(), bwd[i].data(), 0};
    max_bands = std::max(max_bands, (b[3] + 31) / 32);
  }
  // band-major table; dep = global index of the previous band of the same job
  std::vector<InpaintBandRef> bands;
  std::vector<int> prev(n, -1);
  for (int lb = 0; lb < max_bands; ++lb)
    for (int i = 0; i < n; ++i)
      if (lb < (jobs[i].h + 31) / 32) {
        bands.push_back(InpaintBandRef{i, lb, prev[i]});
        prev[i] = (int)bands.size() - 1;
      }
  const int nb = (int)bands.size();
  if (bpp == 8) {
    InpaintFwd<8> pf{nb, jobs.data(), bands.data()};
    run_emulated(pf, n_warps, max_w);
  } else {
    InpaintFwd<16> pf{nb, jobs.data(), bands.data()};
    run_emulated(pf, n_warps, max_w);
  }
  InpaintBwd pb{nb, jobs.data(), bands.data()};
  run_emulated(pb, n_warps, max_w);
  for (int i = 0; i < n; ++i) {
    const InpaintJob& j = jobs[i];
    const size_t px = bpp == 8 ? 4 : 8;
    std::vector<unsigned char> copy((size_t)j.tiff_w * (j.top + j.h) * px);
    memcpy(copy.data(), j.data, copy.size());
    for (int y = 0; y < j.h; ++y)
      for (int x = 0; x < j.w; ++x) {
        u32 src = (u32)bwd[i][(size_t)y * j.w + x];
        size_t dst = (size_t)(j.top + y) * j.tiff_w + j.left + x;
        memcpy((unsigned char*)data[i] + dst * px, copy.data() + src * px, px);
      }
  }
}

}  // extern "C"
