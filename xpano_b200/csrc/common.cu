#include "common.h"

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

unsigned long long xp_kernel_launch_count(void) {
  return xp::g_launches.load(std::memory_order_relaxed);
}

}  // extern "C"
