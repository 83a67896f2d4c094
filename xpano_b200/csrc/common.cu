#include <time.h>
#include "common.h"

#include <cuda.h>

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
    // best fit: a small request must not take the megabyte-sized staging
    // blocks of the JPEG decoder while a small block is free
    size_t best = c.free_list.size();
    for (size_t i = 0; i < c.free_list.size(); ++i)
      if (c.free_list[i].second >= need &&
          (best == c.free_list.size() ||
           c.free_list[i].second < c.free_list[best].second))
        best = i;
    if (best < c.free_list.size()) {
      host = c.free_list[best].first;
      cap = c.free_list[best].second;
      c.free_list.erase(c.free_list.begin() + best);
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

void copy_bytes(void* dst, const void* src, size_t n) {
  const int t = n >= ((size_t)4 << 20) ? std::min(4, host_worker_threads(4)) : 1;
  if (t <= 1) {
    memcpy(dst, src, n);
    return;
  }
  const size_t part = ((n / t) + 4095) & ~(size_t)4095;
  std::vector<std::thread> th;
  for (int k = 1; k < t; ++k) {
    const size_t o = std::min(n, part * k), e = std::min(n, o + part);
    if (e > o)
      th.emplace_back([=] {
        memcpy(static_cast<char*>(dst) + o, static_cast<const char*>(src) + o,
               e - o);
      });
  }
  memcpy(dst, src, std::min(n, part));
  for (std::thread& x : th) x.join();
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

bool make_tensor_map_2d(void* out, int elem_bytes, const void* base,
                        unsigned long long dim0, unsigned long long dim1,
                        unsigned long long stride_bytes, unsigned box0,
                        unsigned box1) {
  typedef CUresult (*EncodeFn)(
      CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
      CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
      CUtensorMapFloatOOBfill);
  static EncodeFn encode = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault,
                                &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      fn = nullptr;
    }
    return reinterpret_cast<EncodeFn>(fn);
  }();
  if (!encode) return false;
  if (((uintptr_t)base & 15) || (stride_bytes & 15) ||
      ((size_t)box0 * elem_bytes & 15) || box0 > 256 || box1 > 256)
    return false;
  const cuuint64_t dims[2] = {dim0, dim1};
  const cuuint64_t strides[1] = {stride_bytes};
  const cuuint32_t box[2] = {box0, box1};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType type = elem_bytes == 8
                                       ? CU_TENSOR_MAP_DATA_TYPE_UINT64
                                       : CU_TENSOR_MAP_DATA_TYPE_UINT32;
  return encode(static_cast<CUtensorMap*>(out), type, 2,
                const_cast<void*>(base), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
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
  const cudaMemcpyKind kind =
      mem == XP_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  // packed on both sides: one linear copy (the 2-D form is submitted row
  // group by row group and holds the driver's lock for longer)
  if (dev_pitch == row_bytes && user_pitch == row_bytes) {
    XP_CUDA(cudaMemcpyAsync(dev, user, row_bytes * rows, kind, s));
    return;
  }
  XP_CUDA(cudaMemcpy2DAsync(dev, dev_pitch, user, user_pitch, row_bytes, rows,
                            kind, s));
}

void copy_out_2d(void* user, size_t user_pitch, const void* dev,
                 size_t dev_pitch, size_t row_bytes, size_t rows, int mem,
                 cudaStream_t s) {
  if (rows == 0 || row_bytes == 0) return;
  const cudaMemcpyKind kind =
      mem == XP_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  if (dev_pitch == row_bytes && user_pitch == row_bytes) {
    XP_CUDA(cudaMemcpyAsync(user, dev, row_bytes * rows, kind, s));
    return;
  }
  XP_CUDA(cudaMemcpy2DAsync(user, user_pitch, dev, dev_pitch, row_bytes, rows,
                            kind, s));
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
struct StreamSetCache {
  std::mutex mu;
  std::vector<StreamSet> free_sets;
};
StreamSetCache& stream_set_cache() {
  static StreamSetCache* c = new StreamSetCache();  // outlives static dtors
  return *c;
}
}  // namespace

// ---------------------------------------------------------------------------
// SM partition. The seam recurrences are latency chains (one compute warp per
// SM, every step waits for the previous one); a streaming kernel that shares
// their SMs takes issue slots from that warp and stretches every step, while
// the streaming kernels themselves do not need all 148 SMs to saturate HBM.
// Two green contexts (CUDA 12.4 driver API: cuDevSmResourceSplitByCount /
// cuGreenCtxCreate) split the device once: `seam_sms` SMs for the seam chain,
// the rest for what runs beside it. Streams created in a green context accept
// ordinary runtime launches and share the primary context's memory.
// XP_SEAM_SMS=<n> sets the split (0 = no partition).
// ---------------------------------------------------------------------------
namespace {
struct SmPartition {
  bool tried = false, ok = false;
  CUgreenCtx ctx_a = nullptr, ctx_b = nullptr;
  int sms_a = 0, sms_b = 0;
  CUresult (*stream_create)(CUstream*, CUgreenCtx, unsigned, int) = nullptr;
};
SmPartition g_part[16];
std::mutex g_part_mu;

template <class F>
bool driver_fn(const char* name, F& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p) {
    cudaGetLastError();
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

SmPartition& sm_partition(int device) {
  std::lock_guard<std::mutex> lk(g_part_mu);
  SmPartition& P = g_part[device & 15];
  if (P.tried) return P;
  P.tried = true;
  int want = 80;
  if (const char* e = getenv("XP_SEAM_SMS")) want = atoi(e);
  if (want <= 0) return P;
  CUresult (*get_res)(CUdevice, CUdevResource*, CUdevResourceType) = nullptr;
  CUresult (*split)(CUdevResource*, unsigned*, const CUdevResource*,
                    CUdevResource*, unsigned, unsigned) = nullptr;
  CUresult (*gen_desc)(CUdevResourceDesc*, CUdevResource*, unsigned) = nullptr;
  CUresult (*ctx_create)(CUgreenCtx*, CUdevResourceDesc, CUdevice, unsigned) =
      nullptr;
  CUresult (*dev_get)(CUdevice*, int) = nullptr;
  if (!driver_fn("cuDeviceGetDevResource", get_res) ||
      !driver_fn("cuDevSmResourceSplitByCount", split) ||
      !driver_fn("cuDevResourceGenerateDesc", gen_desc) ||
      !driver_fn("cuGreenCtxCreate", ctx_create) ||
      !driver_fn("cuGreenCtxStreamCreate", P.stream_create) ||
      !driver_fn("cuDeviceGet", dev_get))
    return P;
  CUdevice dev;
  if (dev_get(&dev, device) != CUDA_SUCCESS) return P;
  CUdevResource all, part_a, part_b;
  if (get_res(dev, &all, CU_DEV_RESOURCE_TYPE_SM) != CUDA_SUCCESS) return P;
  if ((int)all.sm.smCount < want + 16) return P;
  unsigned groups = 1;
  if (split(&part_a, &groups, &all, &part_b, 0, (unsigned)want) !=
          CUDA_SUCCESS ||
      groups != 1 || part_b.sm.smCount < 8)
    return P;
  CUdevResourceDesc da, db;
  if (gen_desc(&da, &part_a, 1) != CUDA_SUCCESS ||
      gen_desc(&db, &part_b, 1) != CUDA_SUCCESS)
    return P;
  if (ctx_create(&P.ctx_a, da, dev, CU_GREEN_CTX_DEFAULT_STREAM) !=
          CUDA_SUCCESS ||
      ctx_create(&P.ctx_b, db, dev, CU_GREEN_CTX_DEFAULT_STREAM) !=
          CUDA_SUCCESS)
    return P;
  P.sms_a = (int)part_a.sm.smCount;
  P.sms_b = (int)part_b.sm.smCount;
  P.ok = true;
  return P;
}
}  // namespace

StreamSet acquire_stream_set(int device) {
  StreamSetCache& c = stream_set_cache();
  {
    std::lock_guard<std::mutex> lk(c.mu);
    for (size_t i = 0; i < c.free_sets.size(); ++i)
      if (c.free_sets[i].device == device) {
        StreamSet s = c.free_sets[i];
        c.free_sets.erase(c.free_sets.begin() + i);
        return s;
      }
  }
  StreamSet s;
  s.device = device;
  s.side = create_stream(false);
  s.aux = create_stream(false);
  for (cudaEvent_t& e : s.ev)
    XP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  SmPartition& P = sm_partition(device);
  if (P.ok) {
    int least = 0, greatest = 0;
    XP_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    CUstream a = nullptr, b1 = nullptr, b2 = nullptr;
    if (P.stream_create(&a, P.ctx_a, CU_STREAM_NON_BLOCKING, greatest) ==
            CUDA_SUCCESS &&
        P.stream_create(&b1, P.ctx_b, CU_STREAM_NON_BLOCKING, least) ==
            CUDA_SUCCESS &&
        P.stream_create(&b2, P.ctx_b, CU_STREAM_NON_BLOCKING, least) ==
            CUDA_SUCCESS) {
      s.seam = a;
      s.part_side = b1;
      s.part_aux = b2;
      s.seam_sms = P.sms_a;
      s.side_sms = P.sms_b;
    }
  }
  return s;
}

void release_stream_set(const StreamSet& s) {
  if (!s.side) return;
  StreamSetCache& c = stream_set_cache();
  std::lock_guard<std::mutex> lk(c.mu);
  c.free_sets.push_back(s);
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
  // Waits that last milliseconds (bulk uploads and downloads) neither spin
  // on a core nor sleep on an interrupt: the first costs a core per waiting
  // thread -- with several panoramas in flight per rank and several ranks
  // per box that is all of them --, the second was measured at several
  // milliseconds per wake-up with eight ranks on one box. The stream is
  // polled with short sleeps in between instead: no core held, wake-up
  // within tens of microseconds. XP_SYNC_EVENT=1 restores the blocking-sync
  // event.
  static const bool use_event = [] {
    const char* e = getenv("XP_SYNC_EVENT");
    return e && atoi(e) != 0;
  }();
  if (!use_event) {
    for (;;) {
      const cudaError_t q = cudaStreamQuery(s);
      if (q == cudaSuccess) return;
      if (q != cudaErrorNotReady) XP_CUDA(q);
      struct timespec ts = {0, 30000};  // 30 us
      nanosleep(&ts, nullptr);
    }
  }
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
