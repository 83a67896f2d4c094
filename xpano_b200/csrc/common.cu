#include "common.h"

#include <map>
#include <mutex>
#include <thread>
#include <algorithm>
#include <cstdlib>
#include <sched.h>
#include <utility>
#include <vector>
#include <cstring>

namespace xp {

static thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void fail(xp_status code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw Error(code, buf);
}

// ---------------------------------------------------------------------------
// caching device allocator
// ---------------------------------------------------------------------------
namespace {
struct Pool {
  std::mutex mu;
  std::multimap<size_t, void*> free_blocks[16];  // per device, by size
};
Pool& pool() {
  static Pool* p = new Pool();  // leaked on purpose: outlives static dtors
  return *p;
}
}  // namespace

// ---------------------------------------------------------------------------
// mailbox
// ---------------------------------------------------------------------------
namespace {
struct MailboxCache {
  std::mutex mu;
  std::vector<std::pair<unsigned char*, size_t>> free_list;
};
MailboxCache& mailbox_cache() {
  static MailboxCache* c = new MailboxCache();  // outlives static destructors
  return *c;
}
__global__ void copy_words_kernel(uint32_t* __restrict__ dst,
                                  const uint32_t* __restrict__ src, size_t n) {
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}
}  // namespace

Mailbox::~Mailbox() {
  if (!host) return;
  MailboxCache& c = mailbox_cache();
  std::lock_guard<std::mutex> lk(c.mu);
  c.free_list.emplace_back(host, cap);
}

void Mailbox::reset(size_t need) {
  need = (need + 4095) & ~(size_t)4095;
  used = 0;
  if (need <= cap) return;
  MailboxCache& c = mailbox_cache();
  {
    std::lock_guard<std::mutex> lk(c.mu);
    if (host) c.free_list.emplace_back(host, cap);
    host = nullptr;
    cap = 0;
    for (size_t i = 0; i < c.free_list.size(); ++i)
      if (c.free_list[i].second >= need) {
        host = c.free_list[i].first;
        cap = c.free_list[i].second;
        c.free_list.erase(c.free_list.begin() + i);
        break;
      }
  }
  if (!host) {
    void* p = nullptr;
    XP_CUDA(cudaHostAlloc(&p, need, cudaHostAllocMapped | cudaHostAllocPortable));
    host = static_cast<unsigned char*>(p);
    cap = need;
  }
}

void* Mailbox::take(size_t bytes) {
  const size_t b = (bytes + 15) & ~(size_t)15;
  XP_REQUIRE(used + b <= cap, "mailbox overflow (%zu + %zu > %zu)", used, b,
             cap);
  void* p = host + used;
  used += b;
  return p;
}

void Mailbox::to_device(void* dst, const void* src, size_t bytes,
                        cudaStream_t s) {
  if (!bytes) return;
  XP_REQUIRE(bytes % 4 == 0, "mailbox transfers are whole words");
  void* slot = take(bytes);
  memcpy(slot, src, bytes);
  void* dslot = nullptr;
  XP_CUDA(cudaHostGetDevicePointer(&dslot, slot, 0));
  copy_words_kernel<<<1, 256, 0, s>>>(static_cast<uint32_t*>(dst),
                                      static_cast<const uint32_t*>(dslot),
                                      bytes / 4);
  XP_LAUNCHED();
}

void* Mailbox::to_host(const void* src, size_t bytes, cudaStream_t s) {
  XP_REQUIRE(bytes % 4 == 0, "mailbox transfers are whole words");
  void* slot = take(bytes);
  void* dslot = nullptr;
  XP_CUDA(cudaHostGetDevicePointer(&dslot, slot, 0));
  copy_words_kernel<<<1, 256, 0, s>>>(static_cast<uint32_t*>(dslot),
                                      static_cast<const uint32_t*>(src),
                                      bytes / 4);
  XP_LAUNCHED();
  return slot;
}

void* pool_alloc(size_t n, size_t* got, int* device) {
  int dev = 0;
  XP_CUDA(cudaGetDevice(&dev));
  *device = dev;
  // round up so near-equal requests of successive panoramas hit the cache
  const size_t gran = n < (1u << 20) ? 4096 : (size_t)1 << 20;
  n = (n + gran - 1) / gran * gran;
  Pool& P = pool();
  if (dev < 16) {
    std::lock_guard<std::mutex> lk(P.mu);
    auto& fb = P.free_blocks[dev];
    auto it = fb.lower_bound(n);
    if (it != fb.end() && it->first <= n + n / 4 + (1 << 20)) {
      void* p = it->second;
      *got = it->first;
      fb.erase(it);
      return p;
    }
  }
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, n);
  if (e == cudaErrorMemoryAllocation) {
    cudaGetLastError();
    pool_trim();
    e = cudaMalloc(&p, n);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    fail(e == cudaErrorMemoryAllocation ? XP_ERR_NOMEM : XP_ERR_CUDA,
         "cudaMalloc(%zu) failed: %s", n, cudaGetErrorString(e));
  }
  *got = n;
  return p;
}

void pool_free(void* p, size_t bytes, int device) {
  if (!p) return;
  Pool& P = pool();
  if (device >= 0 && device < 16) {
    std::lock_guard<std::mutex> lk(P.mu);
    P.free_blocks[device].emplace(bytes, p);
    return;
  }
  cudaFree(p);
}

void pool_trim() {
  Pool& P = pool();
  std::lock_guard<std::mutex> lk(P.mu);
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < 16; ++d) {
    if (P.free_blocks[d].empty()) continue;
    cudaSetDevice(d);
    for (auto& kv : P.free_blocks[d]) cudaFree(kv.second);
    P.free_blocks[d].clear();
  }
  cudaSetDevice(cur);
}

void select_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    fail(XP_ERR_CUDA,
         "no CUDA device (%s); xpano_b200 has no CPU fallback",
         e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
  if (device < 0 || device >= n)
    fail(XP_ERR_INVALID, "device %d out of range (have %d)", device, n);
  XP_CUDA(cudaSetDevice(device));
}

int sm_count() {
  static std::atomic<int> cached[16];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  std::atomic<int>& c = cached[dev & 15];
  int v = c.load(std::memory_order_relaxed);
  if (v == 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) !=
            cudaSuccess || v <= 0)
      v = 148;
    c.store(v, std::memory_order_relaxed);
  }
  return v;
}

void copy_in_2d(void* dev, size_t dev_pitch, const void* user,
                size_t user_pitch, size_t row_bytes, size_t rows, int mem,
                cudaStream_t s) {
  if (rows == 0 || row_bytes == 0) return;
  XP_CUDA(cudaMemcpy2DAsync(
      dev, dev_pitch, user, user_pitch, row_bytes, rows,
      mem == XP_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
      s));
}

void copy_out_2d(void* user, size_t user_pitch, const void* dev,
                 size_t dev_pitch, size_t row_bytes, size_t rows, int mem,
                 cudaStream_t s) {
  if (rows == 0 || row_bytes == 0) return;
  XP_CUDA(cudaMemcpy2DAsync(
      user, user_pitch, dev, dev_pitch, row_bytes, rows,
      mem == XP_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
      s));
}

cudaStream_t create_stream(bool high_priority) {
  int least = 0, greatest = 0;
  XP_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
  cudaStream_t s = nullptr;
  XP_CUDA(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking,
                                       high_priority ? greatest : least));
  return s;
}

namespace {
struct HostShare {
  int threads;  // cap of worker threads per process
  bool block;   // waits yield the core
  HostShare() {
    const unsigned hc = std::thread::hardware_concurrency();
    int cores = hc ? (int)hc : 8;
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0)
      cores = std::max(1, CPU_COUNT(&set));
    int ranks = 1;
    if (const char* e = getenv("LOCAL_WORLD_SIZE")) ranks = std::max(1, atoi(e));
    const int per_rank = std::max(1, cores / ranks);
    threads = per_rank;
    // spinning waits stay the default even when cores are short: blocking
    // ones (interrupt-driven) were measured at several ms per wake-up with
    // eight ranks on one box
    block = false;
    if (const char* e = getenv("XP_HOST_THREADS")) threads = std::max(1, atoi(e));
    if (const char* e = getenv("XP_SYNC_BLOCK")) block = atoi(e) != 0;
  }
};
const HostShare& host_share() {
  static const HostShare h;
  return h;
}
}  // namespace

int host_worker_threads(int wanted) {
  return std::max(0, std::min(wanted, host_share().threads - 1));
}

void sync_stream(cudaStream_t s) {
  if (host_share().block)
    sync_blocking(s);
  else
    XP_CUDA(cudaStreamSynchronize(s));
}

void sync_blocking(cudaStream_t s) {
  // one event per thread and device; blocking-sync events make
  // cudaEventSynchronize yield the core
  struct Ev {
    cudaEvent_t e[16] = {};
  };
  static thread_local Ev evs;
  int dev = 0;
  XP_CUDA(cudaGetDevice(&dev));
  cudaEvent_t& e = evs.e[dev & 15];
  if (!e)
    XP_CUDA(cudaEventCreateWithFlags(
        &e, cudaEventBlockingSync | cudaEventDisableTiming));
  XP_CUDA(cudaEventRecord(e, s));
  XP_CUDA(cudaEventSynchronize(e));
}

}  // namespace xp

extern "C" {

const char* xp_last_error(void) { return xp::g_err; }
const char* xp_version(void) { return "xpano_b200 0.1 (sm_100a)"; }

int xp_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

void xp_trim_memory(void) { xp::pool_trim(); }

unsigned long long xp_kernel_launch_count(void) {
  return xp::g_launches.load(std::memory_order_relaxed);
}

}  // extern "C"
