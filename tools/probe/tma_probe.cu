// Development probe: 2-D TMA load of 32-bit elements, variants selected by
// argv[1], to find which form this driver / GPU accepts.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

__device__ __forceinline__ unsigned smem_addr(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}

template <int BOXW, int BOXH, int VARIANT>
__global__ void probe(const void* tmap, int x, int y, unsigned* out) {
  __shared__ __align__(128) unsigned tile[BOXH * BOXW];
  __shared__ __align__(8) unsigned long long bar;
  const int tid = threadIdx.x;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&bar)), "r"(1) : "memory");
    if (VARIANT == 0)
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    else
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar)),
                 "r"((unsigned)(BOXW * BOXH * 4))
                 : "memory");
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_addr(tile)),
        "l"(tmap), "r"(x), "r"(y), "r"(smem_addr(&bar))
        : "memory");
  }
  unsigned done;
  do {
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(smem_addr(&bar)), "r"(0u)
        : "memory");
  } while (!done);
  for (int i = tid; i < BOXW * BOXH; i += blockDim.x) out[i] = tile[i];
}

#define CK(x)                                                         \
  do {                                                                \
    cudaError_t e = (x);                                              \
    if (e != cudaSuccess) {                                           \
      printf("%s -> %s\n", #x, cudaGetErrorString(e));                \
      return 1;                                                       \
    }                                                                 \
  } while (0)

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                             CUtensorMapFloatOOBfill);

template <int BOXW, int BOXH, int VARIANT>
int run(bool in_global) {
  const int W = 320, H = 240;
  std::vector<unsigned> h(W * H);
  for (int i = 0; i < W * H; ++i) h[i] = i;
  unsigned *d, *out;
  CK(cudaMalloc(&d, W * H * 4));
  CK(cudaMalloc(&out, BOXW * BOXH * 4));
  CK(cudaMemcpy(d, h.data(), W * H * 4, cudaMemcpyHostToDevice));
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  alignas(64) CUtensorMap map;
  const cuuint64_t dims[2] = {W, H};
  const cuuint64_t strides[1] = {W * 4};
  const cuuint32_t box[2] = {BOXW, BOXH};
  const cuuint32_t es[2] = {1, 1};
  CUresult r = ((EncodeFn)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode box %dx%d -> %d\n", BOXW, BOXH, (int)r);
  if (r != CUDA_SUCCESS) return 1;
  void* dmap;
  CK(cudaMalloc(&dmap, 128));
  CK(cudaMemcpy(dmap, &map, 128, cudaMemcpyHostToDevice));
  (void)in_global;
  probe<BOXW, BOXH, VARIANT><<<1, 256>>>(dmap, 8, 5, out);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  std::vector<unsigned> o(BOXW * BOXH);
  CK(cudaMemcpy(o.data(), out, BOXW * BOXH * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int r2 = 0; r2 < BOXH; ++r2)
    for (int c = 0; c < BOXW; ++c)
      if (o[r2 * BOXW + c] != (unsigned)((5 + r2) * W + 8 + c)) ++bad;
  printf("variant %d box %dx%d: %d mismatches, first %u %u %u\n", VARIANT, BOXW, BOXH, bad, o[0], o[1],
         o[BOXW]);
  return bad != 0;
}

int main(int argc, char** argv) {
  const int v = argc > 1 ? atoi(argv[1]) : 0;
  switch (v) {
    case 0: return run<132, 19, 0>(true);
    case 1: return run<128, 16, 0>(true);
    case 2: return run<132, 19, 1>(true);
    case 3: return run<64, 19, 0>(true);
    case 4: return run<32, 8, 0>(true);
    default: return 2;
  }
}
