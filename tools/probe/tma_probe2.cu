// Development probe 2: closer to shrink_batch_kernel -- many blocks, 43 KB of
// static shared memory, descriptors in global memory written by a copy kernel
// from mapped pinned memory on a non-blocking stream, odd pitches.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

__device__ __forceinline__ unsigned smem_addr(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
struct alignas(64) TM { unsigned char b[128]; };

__global__ void copy_words(unsigned* dst, const unsigned* src, size_t n) {
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

constexpr int BW = 132, BH = 19;
__global__ void __launch_bounds__(256)
    probe(const TM* maps, int W, int H, unsigned* out, int* bad) {
  __shared__ __align__(128) unsigned char stage[30096];
  __shared__ float vert[3][8][134];
  __shared__ __align__(8) unsigned long long bar;
  unsigned(*R)[BW] = reinterpret_cast<unsigned(*)[BW]>(stage);
  const int tid = threadIdx.x;
  const int x = blockIdx.x * 64 + XOFF, y = blockIdx.y * 17 + 2;
  if (x + BW > W || y + BH > H) return;
  const TM* tmap = maps + blockIdx.z;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar)),
                 "r"((unsigned)(BW * BH * 4))
                 : "memory");
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_addr(stage)),
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
  int b = 0;
  for (int i = tid; i < BW * BH; i += 256) {
    const int r = i / BW, c = i % BW;
    const unsigned want = blockIdx.z * 1000000u + (y + r) * W + x + c;
    if (R[r][c] != want) ++b;
    vert[0][r & 7][c] = (float)R[r][c];
  }
  if (b) atomicAdd(bad, b);
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0) out[blockIdx.z] = R[0][0];
}

#define CK(x)                                            \
  do {                                                   \
    cudaError_t e = (x);                                 \
    if (e != cudaSuccess) {                              \
      printf("%s -> %s\n", #x, cudaGetErrorString(e));   \
      return 1;                                          \
    }                                                    \
  } while (0)

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                             CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int W = argc > 1 ? atoi(argv[1]) : 324, H = 240, N = 6;
  const int rounds = argc > 2 ? atoi(argv[2]) : 1;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  cudaStream_t s;
  CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  unsigned* img[N];
  TM* hmaps;
  CK(cudaHostAlloc((void**)&hmaps, sizeof(TM) * N, cudaHostAllocMapped));
  TM* dmaps;
  CK(cudaMalloc(&dmaps, sizeof(TM) * N));
  unsigned* out;
  int* bad;
  CK(cudaMalloc(&out, 4 * N));
  CK(cudaMalloc(&bad, 4));
  for (int round = 0; round < rounds; ++round) {
    CK(cudaMemset(bad, 0, 4));
    for (int k = 0; k < N; ++k) {
      std::vector<unsigned> h(W * H);
      for (int i = 0; i < W * H; ++i) h[i] = k * 1000000u + i;
      CK(cudaMalloc(&img[k], W * H * 4));
      CK(cudaMemcpy(img[k], h.data(), W * H * 4, cudaMemcpyHostToDevice));
      const cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H};
      const cuuint64_t strides[1] = {(cuuint64_t)W * 4};
      const cuuint32_t box[2] = {BW, BH};
      const cuuint32_t es[2] = {1, 1};
      CUresult r = ((EncodeFn)fn)((CUtensorMap*)&hmaps[k], CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, img[k], dims,
                                  strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) {
        printf("encode -> %d\n", (int)r);
        return 1;
      }
    }
    void* dslot;
    CK(cudaHostGetDevicePointer(&dslot, hmaps, 0));
    copy_words<<<1, 256, 0, s>>>((unsigned*)dmaps, (const unsigned*)dslot, sizeof(TM) * N / 4);
    probe<<<dim3(4, 14, N), 256, 0, s>>>(dmaps, W, H, out, bad);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s));
    int hb = -1;
    CK(cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost));
    printf("W %d round %d: %d mismatches\n", W, round, hb);
    for (int k = 0; k < N; ++k) CK(cudaFree(img[k]));
  }
  return 0;
}
