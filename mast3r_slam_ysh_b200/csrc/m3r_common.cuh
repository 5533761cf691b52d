// Shared host/device helpers of the sm_100a backend (no torch, no third-party headers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "m3r_backends.h"

namespace m3r {

// ---- error plumbing -------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define M3R_REQUIRE(cond, ...)                 \
  do {                                         \
    if (!(cond)) {                             \
      ::m3r::set_error(__VA_ARGS__);           \
      return M3R_ERR_INVALID_ARGUMENT;         \
    }                                          \
  } while (0)

#define M3R_CUDA(call)                                                                   \
  do {                                                                                   \
    cudaError_t e__ = (call);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      ::m3r::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return (int)e__;                                                                   \
    }                                                                                    \
  } while (0)

#define M3R_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    ::m3r::count_launch();                                                               \
    cudaError_t e__ = cudaPeekAtLastError();                                             \
    if (e__ != cudaSuccess) {                                                            \
      ::m3r::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
      return (int)e__;                                                                   \
    }                                                                                    \
  } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

int sm_count();  // of the CURRENT device (cached per device)

// cudaFuncSetAttribute applies to the current device only, so the "already opted in" value is remembered PER DEVICE
// (a process that runs on cuda:0 and then on cuda:1 must opt in on both).
constexpr int kMaxDevices = 64;
struct DynSmemOptIn {
  size_t set[kMaxDevices] = {};
};
int current_device_slot();  // cudaGetDevice(), clamped to [0, kMaxDevices)
template <typename K>
static inline int ensure_dyn_smem(K kernel, size_t bytes, DynSmemOptIn& cache) {
  if (bytes <= 48 * 1024) return M3R_OK;
  const int d = current_device_slot();
  if (bytes > cache.set[d]) {
    M3R_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    cache.set[d] = bytes;
  }
  return M3R_OK;
}
// keep stream-ordered scratch cached in the default pool across synchronisations (the default threshold returns it to the OS)
int keep_async_pool();

// ---- read-only vector loads -----------------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }
__device__ __forceinline__ float2 ldg2(const float2* p) { return __ldg(p); }

}  // namespace m3r
