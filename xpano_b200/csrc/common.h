// Shared host-side helpers of libxpano_b200: error convention, CUDA checks,
// device buffers, launch counting. Not part of the public ABI.
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/xpano_b200.h"

namespace xp {

void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;

struct Error : std::runtime_error {
  xp_status code;
  Error(xp_status c, const std::string& m) : std::runtime_error(m), code(c) {}
};

[[noreturn]] void fail(xp_status code, const char* fmt, ...);

#define XP_CUDA(expr)                                                        \
  do {                                                                       \
    cudaError_t e__ = (expr);                                                \
    if (e__ != cudaSuccess)                                                  \
      ::xp::fail(XP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                \
                 cudaGetErrorString(e__), __FILE__, __LINE__);               \
  } while (0)

#define XP_REQUIRE(cond, ...)                                 \
  do {                                                        \
    if (!(cond)) ::xp::fail(XP_ERR_INVALID, __VA_ARGS__);     \
  } while (0)

// counts a kernel launch and checks the launch error
#define XP_LAUNCHED()                                                       \
  do {                                                                      \
    ::xp::g_launches.fetch_add(1, std::memory_order_relaxed);               \
    XP_CUDA(cudaGetLastError());                                            \
  } while (0)

// Wraps a C-ABI body: exceptions never cross the ABI.
template <class F>
xp_status guard(F&& f) {
  try {
    f();
    return XP_OK;
  } catch (const Error& e) {
    set_error("%s", e.what());
    return e.code;
  } catch (const std::bad_alloc&) {
    set_error("out of host memory");
    return XP_ERR_NOMEM;
  } catch (const std::exception& e) {
    set_error("%s", e.what());
    return XP_ERR_INVALID;
  }
}

// Owning device allocation.
// caching device allocator (common.cu). The pool itself is not stream-ordered:
// an object may queue work on several streams (a blender uses three), and
// every owner synchronises all of them -- the caller's stream included, also
// when that is the legacy default stream -- before its buffers are released,
// on success and on every error path (the destructors do it).
void* pool_alloc(size_t n, size_t* got, int* device);
void pool_free(void* p, size_t bytes, int device);
void pool_trim();

// Small control transfers (bounding boxes, counters, job tables) bypass the
// copy engines: with several panoramas in flight those queues hold bulk PCIe
// uploads/downloads, and a 100-byte copy would wait milliseconds behind them.
// The mailbox is pinned, mapped host memory; a one-block kernel moves the words.
struct Mailbox {
  unsigned char* host = nullptr;
  size_t cap = 0, used = 0;
  Mailbox() = default;
  Mailbox(const Mailbox&) = delete;
  Mailbox& operator=(const Mailbox&) = delete;
  ~Mailbox();
  void reset(size_t need);  // start a new round with room for `need` bytes
  void* take(size_t bytes); // 16-byte aligned slot
  // device <- host bytes (copied into a slot now, moved by a kernel on s)
  void to_device(void* dst, const void* src, size_t bytes, cudaStream_t s);
  // host slot <- device bytes; readable after s has been synchronised
  void* to_host(const void* src, size_t bytes, cudaStream_t s);
};

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int device = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes), device(o.device) {
    o.p = nullptr;
    o.bytes = 0;
  }
  DevBuf& operator=(DevBuf&& o) noexcept {
    if (this != &o) {
      release();
      p = o.p;
      bytes = o.bytes;
      device = o.device;
      o.p = nullptr;
      o.bytes = 0;
    }
    return *this;
  }
  ~DevBuf() { release(); }
  void release() {
    if (p) pool_free(p, bytes, device);
    p = nullptr;
    bytes = 0;
  }
  // grow-only (re)allocation; contents are not preserved. Blocks come from a
  // per-device caching pool: a blender is single-use like the reference's, so
  // without the pool every panorama would pay cudaMalloc/cudaFree again.
  void reserve(size_t n) {
    if (n <= bytes) return;
    release();
    if (n == 0) return;
    p = pool_alloc(n, &bytes, &device);
  }
  template <class T>
  T* as() const {
    return static_cast<T*>(p);
  }
};

// Declared AFTER the DevBufs of a function that queues work on `s`: on every
// way out, an exception included, the stream is drained before the buffers
// go back to the pool (which is not stream-ordered).
struct SyncOnExit {
  cudaStream_t s;
  explicit SyncOnExit(cudaStream_t stream) : s(stream) {}
  SyncOnExit(const SyncOnExit&) = delete;
  SyncOnExit& operator=(const SyncOnExit&) = delete;
  ~SyncOnExit() { cudaStreamSynchronize(s); }
};

// Non-blocking stream of the highest (main chain: seams, pyramid tail) or
// lowest (work that runs beside it) priority the device offers: when both
// have blocks pending, the main chain's are placed first.
cudaStream_t create_stream(bool high_priority);

// The side streams and events of a blender: blenders are single-use (one per
// panorama, like the reference's), and creating two streams and five events
// costs more host time than a whole pyramid level; sets are kept per device
// and handed out again. A set is returned only after its streams were
// synchronised.
struct StreamSet {
  int device = -1;
  cudaStream_t side = nullptr, aux = nullptr;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // SM partition (green contexts, see sm_partition): a stream that runs on
  // the seam partition only, and two that run on the rest of the device.
  // Null when the device is not partitioned.
  cudaStream_t seam = nullptr, part_side = nullptr, part_aux = nullptr;
  int seam_sms = 0, side_sms = 0;
};
StreamSet acquire_stream_set(int device);
void release_stream_set(const StreamSet& s);

// Wait for a stream without holding a core: for the long waits on bulk
// copies (uploads, downloads) that overlap another thread's kernel launches.
// Polls the stream with 30-us sleeps in between (see common.cu).
void sync_blocking(cudaStream_t s);

// Host threads this process should use for short parallel host work, and
// whether its stream waits should yield the core: several ranks of one job
// share the box's cores (torchrun exports LOCAL_WORLD_SIZE), and a rank that
// spins in the driver or wakes eight workers takes them from its peers.
// Overrides: XP_HOST_THREADS, XP_SYNC_BLOCK.
int host_worker_threads(int wanted);
// memcpy split over a few host threads when it is large (a 16-MB stream
// leaving pinned memory: 1.6 ms on one core)
void copy_bytes(void* dst, const void* src, size_t n);
void sync_stream(cudaStream_t s);

// TMA tensor map (cuTensorMapEncodeTiled, reached through
// cudaGetDriverEntryPoint: no link-time dependency on libcuda) over a
// row-major 2-D array of 4- or 8-byte elements: dim0 elements per row, dim1
// rows, `stride_bytes` between rows (a multiple of 16, base 16-byte aligned),
// box0 x box1 elements per copy (box0 * elem_bytes a multiple of 16), no
// swizzle, zero fill outside the array. `out` is 128 bytes (a CUtensorMap).
// Returns false where the driver refuses (the callers keep a plain-load path).
bool make_tensor_map_2d(void* out, int elem_bytes, const void* base,
                        unsigned long long dim0, unsigned long long dim1,
                        unsigned long long stride_bytes, unsigned box0,
                        unsigned box1);

inline int div_up(int a, int b) { return (a + b - 1) / b; }
inline size_t div_up(size_t a, size_t b) { return (a + b - 1) / b; }

void select_device(int device);
// multiprocessors of the current device (cudaDevAttrMultiProcessorCount,
// queried once per device; 148 on B200): persistent grids are sized from it
int sm_count();

// 2-D copy helper between host/device according to `mem` of the *user* side.
void copy_in_2d(void* dev, size_t dev_pitch, const void* user,
                size_t user_pitch, size_t row_bytes, size_t rows, int mem,
                cudaStream_t s);
void copy_out_2d(void* user, size_t user_pitch, const void* dev,
                 size_t dev_pitch, size_t row_bytes, size_t rows, int mem,
                 cudaStream_t s);

}  // namespace xp
