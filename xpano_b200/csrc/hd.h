// Host/device function qualifier shared by code that is compiled both by nvcc
// and by the host-only SIMT emulator used in the CPU tests.
#pragma once
#if defined(__CUDACC__)
#define XP_HD __host__ __device__ __forceinline__
#else
#define XP_HD inline
#endif
