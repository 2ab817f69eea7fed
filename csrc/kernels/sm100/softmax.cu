// Scaled causal ("upper-triangular") and scaled masked softmax, forward and backward, for sm_100a.
// One block per score row, the row held in registers as 16-byte vectors; no 2048-key limit (rows up
// to 8192 keys stay in registers, longer rows take the multi-pass scalar kernel).
//
// API parity with the reference's smp_torch_cuda_lib entry points scaled_upper_triang_softmax_
// forward/backward and scaled_masked_softmax_forward/backward (reference smdistributed/
// modelparallel/torch/nn/softmax.py:7-12,40-93; constraint 16 < sk <= 2048 at
// torch/nn/transformer.py:101 is lifted here).
#include "common.cuh"

using namespace smpk;

namespace {

constexpr float NEG_BIG = -1e30f;

// mode 0: no mask, 1: causal (col <= row + Sk - Sq), 2: byte mask [B,1,Sq,Sk] (non-zero = masked out),
// 3: causal, and only columns below the next multiple of 256 past the diagonal are written (the triangle-aware
//    GEMMs that consume the result never read further; saves the zero fill of the upper triangle)
template <int NV>
__global__ void softmax_fwd_kernel(const void* __restrict__ x, void* __restrict__ y, int dtype,
                                   const uint8_t* __restrict__ mask, int mode, float scale, int heads, int Sq, int Sk,
                                   long long rows) {
  __shared__ float red[32];
  const int nvec = Sk >> 3;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int q = (int)(row % Sq);
    const long long bh = row / Sq;
    const int limit = (mode == 1 || mode == 3) ? min(Sk, q + 1 + (Sk - Sq)) : Sk;  // columns [0, limit) are visible
    const int wlimit = (mode == 3) ? min(Sk, (limit + 255) & ~255) : Sk;            // columns [0, wlimit) are written
    const uint8_t* mrow = (mode == 2) ? mask + ((bh / heads) * (long long)Sq + q) * Sk : nullptr;
    const size_t base = (size_t)row * Sk;
    float v[NV][8];
    float mx = NEG_BIG;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = threadIdx.x + i * blockDim.x;
      const int c0 = vi * 8;
      if (vi < nvec && c0 < limit) {
        load8(x, dtype, base + c0, v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const bool dead = (c0 + j >= limit) || (mrow != nullptr && mrow[c0 + j] != 0);
          v[i][j] = dead ? NEG_BIG : v[i][j] * scale;
          mx = fmaxf(mx, v[i][j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = NEG_BIG;
      }
    }
    mx = block_max(mx, red);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float e = (v[i][j] <= NEG_BIG * 0.5f) ? 0.f : __expf(v[i][j] - mx);
        v[i][j] = e;
        sum += e;
      }
    sum = block_sum(sum, red);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = threadIdx.x + i * blockDim.x;
      if (vi < nvec && vi * 8 < wlimit) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] *= inv;
        store8(y, dtype, base + (size_t)vi * 8, v[i]);
      }
    }
  }
}

__global__ void softmax_fwd_scalar_kernel(const void* __restrict__ x, void* __restrict__ y, int dtype,
                                          const uint8_t* __restrict__ mask, int mode, float scale, int heads, int Sq,
                                          int Sk, long long rows) {
  __shared__ float red[32];
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int q = (int)(row % Sq);
    const long long bh = row / Sq;
    const int limit = (mode == 1 || mode == 3) ? min(Sk, q + 1 + (Sk - Sq)) : Sk;   // scalar path: mode 3 fills whole rows
    const uint8_t* mrow = (mode == 2) ? mask + ((bh / heads) * (long long)Sq + q) * Sk : nullptr;
    const size_t base = (size_t)row * Sk;
    float mx = NEG_BIG;
    for (int c = threadIdx.x; c < limit; c += blockDim.x)
      if (!(mrow && mrow[c])) mx = fmaxf(mx, load1(x, dtype, base + c) * scale);
    mx = block_max(mx, red);
    float sum = 0.f;
    for (int c = threadIdx.x; c < limit; c += blockDim.x)
      if (!(mrow && mrow[c])) sum += __expf(load1(x, dtype, base + c) * scale - mx);
    sum = block_sum(sum, red);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    for (int c = threadIdx.x; c < Sk; c += blockDim.x) {
      const bool dead = c >= limit || (mrow && mrow[c]);
      store1(y, dtype, base + c, dead ? 0.f : __expf(load1(x, dtype, base + c) * scale - mx) * inv);
    }
  }
}

// dx = scale * y * (dy - sum(dy * y)); masked positions have y == 0 so need no special casing.
template <int NV>
__global__ void softmax_bwd_kernel(const void* __restrict__ dy, const void* __restrict__ y, void* __restrict__ dx,
                                   int dtype, int causal, float scale, int Sq, int Sk, long long rows) {
  __shared__ float red[32];
  const int nvec = Sk >> 3;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int q = (int)(row % Sq);
    const int limit = causal ? min(Sk, q + 1 + (Sk - Sq)) : Sk;
    const int wlimit = (causal == 2) ? min(Sk, (limit + 255) & ~255) : Sk;
    const size_t base = (size_t)row * Sk;
    float yv[NV][8], dv[NV][8];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = threadIdx.x + i * blockDim.x;
      if (vi < nvec && vi * 8 < limit) {
        load8(y, dtype, base + (size_t)vi * 8, yv[i]);
        load8(dy, dtype, base + (size_t)vi * 8, dv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) dot += yv[i][j] * dv[i][j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          yv[i][j] = 0.f;
          dv[i][j] = 0.f;
        }
      }
    }
    dot = block_sum(dot, red);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = threadIdx.x + i * blockDim.x;
      if (vi < nvec && vi * 8 < wlimit) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = scale * yv[i][j] * (dv[i][j] - dot);
        store8(dx, dtype, base + (size_t)vi * 8, o);
      }
    }
  }
}

__global__ void softmax_bwd_scalar_kernel(const void* __restrict__ dy, const void* __restrict__ y,
                                          void* __restrict__ dx, int dtype, float scale, int Sk, long long rows) {
  __shared__ float red[32];
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * Sk;
    float dot = 0.f;
    for (int c = threadIdx.x; c < Sk; c += blockDim.x) dot += load1(y, dtype, base + c) * load1(dy, dtype, base + c);
    dot = block_sum(dot, red);
    for (int c = threadIdx.x; c < Sk; c += blockDim.x)
      store1(dx, dtype, base + c, scale * load1(y, dtype, base + c) * (load1(dy, dtype, base + c) - dot));
  }
}

inline void pick_geometry(int Sk, int& threads, int& nv) {
  const int nvec = Sk >> 3;
  threads = 32;
  while (threads < 256 && (nvec + threads - 1) / threads > 2) threads <<= 1;
  nv = (nvec + threads - 1) / threads;
}

}  // namespace

// x, y: [batch*heads, Sq, Sk] contiguous; mask (mode 2): uint8 [batch, 1, Sq, Sk].
SMPK_API int smpk_softmax_fwd(void* stream, const void* x, void* y, int dtype, const void* mask, int mode, float scale,
                              int batch_heads, int heads, int Sq, int Sk) {
  SMPK_REQUIRE(mode >= 0 && mode <= 3, "softmax_fwd: bad mode %d", mode);
  SMPK_REQUIRE(mode != 2 || (mask != nullptr && heads > 0), "softmax_fwd: mask mode needs mask and heads");
  const long long rows = (long long)batch_heads * Sq;
  if (rows == 0 || Sk == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int threads, nv;
  pick_geometry(Sk, threads, nv);
  const long long cap = (long long)smpk::num_sms() * 16;
  const int grid = (int)(rows < cap ? rows : cap);
  const bool vec = (Sk % 8 == 0) && nv <= 4 && aligned16(x) && aligned16(y);
  const uint8_t* m = reinterpret_cast<const uint8_t*>(mask);
  if (!vec) {
    softmax_fwd_scalar_kernel<<<grid, 256, 0, s>>>(x, y, dtype, m, mode, scale, heads, Sq, Sk, rows);
  } else {
#define LAUNCH(NV) softmax_fwd_kernel<NV><<<grid, threads, 0, s>>>(x, y, dtype, m, mode, scale, heads, Sq, Sk, rows)
    switch (nv) {
      case 1: LAUNCH(1); break;
      case 2: LAUNCH(2); break;
      case 3: LAUNCH(3); break;
      default: LAUNCH(4); break;
    }
#undef LAUNCH
  }
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_softmax_bwd(void* stream, const void* dy, const void* y, void* dx, int dtype, int causal, float scale,
                              int batch_heads, int Sq, int Sk) {
  const long long rows = (long long)batch_heads * Sq;
  if (rows == 0 || Sk == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int threads, nv;
  pick_geometry(Sk, threads, nv);
  const long long cap = (long long)smpk::num_sms() * 16;
  const int grid = (int)(rows < cap ? rows : cap);
  const bool vec = (Sk % 8 == 0) && nv <= 4 && aligned16(dy) && aligned16(y) && aligned16(dx);
  if (!vec) {
    softmax_bwd_scalar_kernel<<<grid, 256, 0, s>>>(dy, y, dx, dtype, scale, Sk, rows);
  } else {
#define LAUNCH(NV) softmax_bwd_kernel<NV><<<grid, threads, 0, s>>>(dy, y, dx, dtype, causal, scale, Sq, Sk, rows)
    switch (nv) {
      case 1: LAUNCH(1); break;
      case 2: LAUNCH(2); break;
      case 3: LAUNCH(3); break;
      default: LAUNCH(4); break;
    }
#undef LAUNCH
  }
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Megatron-compatible helper kept for API parity (reference torch/nn/softmax.py:12).
SMPK_API int smpk_get_batch_per_block(int sq, int sk, int b, int np) {
  (void)sq; (void)b; (void)np;
  int pow2 = 1;
  while (pow2 < sk) pow2 <<= 1;
  const int warp_size = pow2 < 32 ? pow2 : 32;
  const int batches_per_warp = pow2 <= 128 ? 2 : 1;
  const int warps_per_block = 128 / warp_size;
  return warps_per_block * batches_per_warp;
}
