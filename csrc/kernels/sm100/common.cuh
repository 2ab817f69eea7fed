// Shared host/device helpers for the sm_100a kernel library (libsmpb200_kernels.so).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

#define SMPK_API extern "C" __attribute__((visibility("default")))

enum SmpkDtype : int { SMPK_F32 = 0, SMPK_F16 = 1, SMPK_BF16 = 2 };

namespace smpk {

// Thread-local last-error text, readable through smpk_last_error().
char* last_error_buf();
int set_error(const char* fmt, ...);

#define SMPK_CHECK_CUDA(expr)                                                                          \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess)                                                                             \
      return smpk::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define SMPK_REQUIRE(cond, ...)                         \
  do {                                                  \
    if (!(cond)) return smpk::set_error(__VA_ARGS__);   \
  } while (0)

inline int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = n > 0 ? n : 148;
  }
  return cached[dev];
}

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) {
  return __half2float(v);
}
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ __half from_f32<__half>(float v) {
  return __float2half_rn(v);
}
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// tanh-form GELU (HF "gelu_new" / Megatron bias_gelu) and exact erf form, with derivatives.
__device__ __forceinline__ float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * x * (1.f + k1 * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * x * (1.f + k1 * x * x);
  float t = tanhf(u);
  float du = k0 * (1.f + 3.f * k1 * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.f + erff(x * 0.7071067811865475f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
  float cdf = 0.5f * (1.f + erff(x * 0.7071067811865475f));
  float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}


// 8-element vector IO with a runtime dtype (uniform branch; memory-bound kernels only).
// `idx` is an element index that must be a multiple of 8 and the base 16-byte aligned.
__device__ __forceinline__ void load8(const void* __restrict__ base, int dtype, size_t idx, float (&o)[8]) {
  if (dtype == SMPK_F32) {
    const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + idx);
    float4 a = p[0], b = p[1];
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
  } else {
    uint4 raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(base) + idx);
    if (dtype == SMPK_BF16) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float2 f = __bfloat1622float2(h[i]);
        o[2 * i] = f.x; o[2 * i + 1] = f.y;
      }
    } else {
      const __half2* h = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float2 f = __half22float2(h[i]);
        o[2 * i] = f.x; o[2 * i + 1] = f.y;
      }
    }
  }
}
__device__ __forceinline__ void store8(void* __restrict__ base, int dtype, size_t idx, const float (&v)[8]) {
  if (dtype == SMPK_F32) {
    float4* p = reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + idx);
    p[0] = make_float4(v[0], v[1], v[2], v[3]);
    p[1] = make_float4(v[4], v[5], v[6], v[7]);
  } else {
    uint4 raw;
    if (dtype == SMPK_BF16) {
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    } else {
      __half2* h = reinterpret_cast<__half2*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    }
    *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(base) + idx) = raw;
  }
}
__device__ __forceinline__ float load1(const void* __restrict__ base, int dtype, size_t idx) {
  if (dtype == SMPK_F32) return reinterpret_cast<const float*>(base)[idx];
  if (dtype == SMPK_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
  return __half2float(reinterpret_cast<const __half*>(base)[idx]);
}
__device__ __forceinline__ void store1(void* __restrict__ base, int dtype, size_t idx, float v) {
  if (dtype == SMPK_F32) reinterpret_cast<float*>(base)[idx] = v;
  else if (dtype == SMPK_BF16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<__half*>(base)[idx] = __float2half_rn(v);
}
__host__ __device__ inline int dtype_size(int dtype) { return dtype == SMPK_F32 ? 4 : 2; }
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Block-wide sum of one float; `red` must hold >= 32 floats of shared memory.  All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  return warp_max(r);
}

// Dispatch on a runtime SmpkDtype for 16/32-bit floating types.
#define SMPK_DISPATCH_FLOAT(dt, T, ...)                         \
  switch (dt) {                                                 \
    case SMPK_F32: {                                            \
      using T = float;                                          \
      __VA_ARGS__;                                              \
    } break;                                                    \
    case SMPK_F16: {                                            \
      using T = __half;                                         \
      __VA_ARGS__;                                              \
    } break;                                                    \
    case SMPK_BF16: {                                           \
      using T = __nv_bfloat16;                                  \
      __VA_ARGS__;                                              \
    } break;                                                    \
    default:                                                    \
      return smpk::set_error("unsupported dtype code %d", dt);  \
  }

}  // namespace smpk
