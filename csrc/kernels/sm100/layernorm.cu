// Fused LayerNorm forward / backward for sm_100a (memory-bound: one pass over x, 16-byte vectors,
// rows kept in registers).  Optional fused residual add in the forward (s = x + residual is
// written out as the new residual stream and normalised in the same pass).
//
// Backward is a persistent grid: each block walks rows with a grid stride, produces dx per row and
// keeps its own dgamma / dbeta partial sums in registers; partials go to a [grid, H] fp32 scratch
// and a second kernel folds them.  This replaces the reference's Apex kernels
// forward_affine / forward_affine_mixed_dtypes / backward_affine (reference smdistributed/
// modelparallel/torch/apex/normalization/fused_layer_norm.py:26-78) and, with externally supplied
// statistics, its hidden-sharded variant forward_affine_mean_var / backward_affine_local_sums /
// backward_affine_finish (reference torch/nn/layer_norm.py:52-81).
#include "common.cuh"

using namespace smpk;

namespace {

constexpr int MAX_NV = 4;

template <int NV>
__global__ void ln_fwd_kernel(const void* __restrict__ x, const void* __restrict__ residual,
                              void* __restrict__ sum_out, int x_dtype, const void* __restrict__ gamma,
                              const void* __restrict__ beta, int w_dtype, void* __restrict__ y, int y_dtype,
                              float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int H, float eps) {
  __shared__ float red[32];
  const int nvec = H >> 3;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * H;
    float xv[NV][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        load8(x, x_dtype, base + (size_t)v * 8, xv[i]);
        if (residual != nullptr) {
          float r[8];
          load8(residual, x_dtype, base + (size_t)v * 8, r);
#pragma unroll
          for (int j = 0; j < 8; ++j) xv[i][j] += r[j];
          if (sum_out != nullptr) store8(sum_out, x_dtype, base + (size_t)v * 8, xv[i]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += xv[i][j];
      }
    }
    const float mean = block_sum(s, red) / H;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float d = xv[i][j] - mean;
          q += d * d;
        }
      }
    }
    const float var = block_sum(q, red) / H;
    const float rstd = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        float g[8], b[8], o[8];
        if (gamma != nullptr) load8(gamma, w_dtype, (size_t)v * 8, g);
        if (beta != nullptr) load8(beta, w_dtype, (size_t)v * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float n = (xv[i][j] - mean) * rstd;
          if (gamma != nullptr) n *= g[j];
          if (beta != nullptr) n += b[j];
          o[j] = n;
        }
        store8(y, y_dtype, base + (size_t)v * 8, o);
      }
    }
  }
}

// Scalar fallback (H not a multiple of 8 or unaligned bases): one block per row.
__global__ void ln_fwd_scalar_kernel(const void* __restrict__ x, const void* __restrict__ residual,
                                     void* __restrict__ sum_out, int x_dtype, const void* __restrict__ gamma,
                                     const void* __restrict__ beta, int w_dtype, void* __restrict__ y, int y_dtype,
                                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int H,
                                     float eps) {
  __shared__ float red[32];
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * H;
    float s = 0.f;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float v = load1(x, x_dtype, base + c);
      if (residual != nullptr) {
        v += load1(residual, x_dtype, base + c);
        if (sum_out != nullptr) store1(sum_out, x_dtype, base + c, v);
      }
      s += v;
    }
    const float mean = block_sum(s, red) / H;
    float q = 0.f;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float v = load1(x, x_dtype, base + c);
      if (residual != nullptr) v += load1(residual, x_dtype, base + c);
      float d = v - mean;
      q += d * d;
    }
    const float rstd = rsqrtf(block_sum(q, red) / H + eps);
    if (threadIdx.x == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float v = load1(x, x_dtype, base + c);
      if (residual != nullptr) v += load1(residual, x_dtype, base + c);
      float n = (v - mean) * rstd;
      if (gamma != nullptr) n *= load1(gamma, w_dtype, c);
      if (beta != nullptr) n += load1(beta, w_dtype, c);
      store1(y, y_dtype, base + c, n);
    }
  }
}

template <int NV>
__global__ void ln_bwd_kernel(const void* __restrict__ dy, int dy_dtype, const void* __restrict__ x, int x_dtype,
                              const float* __restrict__ mean, const float* __restrict__ rstd,
                              const void* __restrict__ gamma, int w_dtype, void* __restrict__ dx, int dx_dtype,
                              const void* __restrict__ dx_add, float* __restrict__ scratch /*[2][grid][H]*/, int rows,
                              int H) {
  __shared__ float red[32];
  const int nvec = H >> 3;
  float dg[NV][8], db[NV][8], g[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = threadIdx.x + i * blockDim.x;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dg[i][j] = 0.f;
      db[i][j] = 0.f;
      g[i][j] = 1.f;
    }
    if (gamma != nullptr && v < nvec) load8(gamma, w_dtype, (size_t)v * 8, g[i]);
  }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * H;
    const float mu = mean[row], rs = rstd[row];
    float xh[NV][8], dyv[NV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        load8(x, x_dtype, base + (size_t)v * 8, xh[i]);
        load8(dy, dy_dtype, base + (size_t)v * 8, dyv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xh[i][j] - mu) * rs;
          const float d = dyv[i][j];
          dg[i][j] += d * xh[i][j];
          db[i][j] += d;
          const float dgm = d * g[i][j];
          dyv[i][j] = dgm;
          s1 += dgm;
          s2 += dgm * xh[i][j];
        }
      }
    }
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    const float inv_h = 1.f / H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (dyv[i][j] - (s1 + xh[i][j] * s2) * inv_h);
        if (dx_add != nullptr) {
          float a[8];
          load8(dx_add, dx_dtype, base + (size_t)v * 8, a);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += a[j];
        }
        store8(dx, dx_dtype, base + (size_t)v * 8, o);
      }
    }
  }
  if (scratch != nullptr) {
    float* sg = scratch + (size_t)blockIdx.x * H;
    float* sb = scratch + (size_t)gridDim.x * H + (size_t)blockIdx.x * H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = threadIdx.x + i * blockDim.x;
      if (v < nvec) {
        store8(sg, SMPK_F32, (size_t)v * 8, dg[i]);
        store8(sb, SMPK_F32, (size_t)v * 8, db[i]);
      }
    }
  }
}

__global__ void ln_bwd_scalar_kernel(const void* __restrict__ dy, int dy_dtype, const void* __restrict__ x, int x_dtype,
                                     const float* __restrict__ mean, const float* __restrict__ rstd,
                                     const void* __restrict__ gamma, int w_dtype, void* __restrict__ dx, int dx_dtype,
                                     const void* __restrict__ dx_add, float* __restrict__ scratch, int rows, int H) {
  // One block per grid slot; dgamma/dbeta partials accumulate straight into the (pre-zeroed) scratch rows.
  __shared__ float red[32];
  float* sg = scratch ? scratch + (size_t)blockIdx.x * H : nullptr;
  float* sb = scratch ? scratch + (size_t)gridDim.x * H + (size_t)blockIdx.x * H : nullptr;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * H;
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      const float xh = (load1(x, x_dtype, base + c) - mu) * rs;
      const float d = load1(dy, dy_dtype, base + c);
      const float dgm = d * (gamma ? load1(gamma, w_dtype, c) : 1.f);
      s1 += dgm;
      s2 += dgm * xh;
      if (sg) {
        sg[c] += d * xh;
        sb[c] += d;
      }
    }
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    const float inv_h = 1.f / H;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      const float xh = (load1(x, x_dtype, base + c) - mu) * rs;
      const float dgm = load1(dy, dy_dtype, base + c) * (gamma ? load1(gamma, w_dtype, c) : 1.f);
      float o = rs * (dgm - (s1 + xh * s2) * inv_h);
      if (dx_add != nullptr) o += load1(dx_add, dx_dtype, base + c);
      store1(dx, dx_dtype, base + c, o);
    }
  }
}

// Fold the [2][nparts][H] partials into dgamma / dbeta.  One thread per column, coalesced over columns.
__global__ void ln_bwd_reduce_kernel(const float* __restrict__ scratch, int nparts, int H, void* __restrict__ dgamma,
                                     void* __restrict__ dbeta, int w_dtype, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= H) return;
  float a = 0.f, b = 0.f;
  const float* sg = scratch;
  const float* sb = scratch + (size_t)nparts * H;
  for (int p = 0; p < nparts; ++p) {
    a += sg[(size_t)p * H + c];
    b += sb[(size_t)p * H + c];
  }
  if (dgamma != nullptr) store1(dgamma, w_dtype, c, a + (accumulate ? load1(dgamma, w_dtype, c) : 0.f));
  if (dbeta != nullptr) store1(dbeta, w_dtype, c, b + (accumulate ? load1(dbeta, w_dtype, c) : 0.f));
}

// ---- statistics-supplied variants for hidden-sharded LayerNorm (tensor-parallel, optimize=memory)
// Phase 1: per-row partial sums over the local shard -> stats[row] = {sum, sum of squares}.
__global__ void ln_partial_stats_kernel(const void* __restrict__ x, int x_dtype, float* __restrict__ stats, int rows,
                                        int H) {
  __shared__ float red[32];
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    float s = 0.f, q = 0.f;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float v = load1(x, x_dtype, (size_t)row * H + c);
      s += v;
      q += v * v;
    }
    s = block_sum(s, red);
    q = block_sum(q, red);
    if (threadIdx.x == 0) {
      stats[2 * row] = s;
      stats[2 * row + 1] = q;
    }
  }
}
// Phase 2: normalise the local shard with globally reduced mean / rstd.
__global__ void ln_apply_stats_kernel(const void* __restrict__ x, int x_dtype, const float* __restrict__ mean,
                                      const float* __restrict__ rstd, const void* __restrict__ gamma,
                                      const void* __restrict__ beta, int w_dtype, void* __restrict__ y, int y_dtype,
                                      int rows, int H) {
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const float mu = mean[row], rs = rstd[row];
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float n = (load1(x, x_dtype, (size_t)row * H + c) - mu) * rs;
      if (gamma) n *= load1(gamma, w_dtype, c);
      if (beta) n += load1(beta, w_dtype, c);
      store1(y, y_dtype, (size_t)row * H + c, n);
    }
  }
}
// Backward phase 1: local sums s1 = sum(dy*g), s2 = sum(dy*g*xhat) per row -> sums[row] = {s1, s2}.
__global__ void ln_bwd_local_sums_kernel(const void* __restrict__ dy, int dy_dtype, const void* __restrict__ x,
                                         int x_dtype, const float* __restrict__ mean, const float* __restrict__ rstd,
                                         const void* __restrict__ gamma, int w_dtype, float* __restrict__ sums, int rows,
                                         int H) {
  __shared__ float red[32];
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      const float xh = (load1(x, x_dtype, (size_t)row * H + c) - mu) * rs;
      const float dgm = load1(dy, dy_dtype, (size_t)row * H + c) * (gamma ? load1(gamma, w_dtype, c) : 1.f);
      s1 += dgm;
      s2 += dgm * xh;
    }
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    if (threadIdx.x == 0) {
      sums[2 * row] = s1;
      sums[2 * row + 1] = s2;
    }
  }
}
// Backward phase 2: dx for the local shard from globally reduced sums; H_total is the full width.
__global__ void ln_bwd_finish_kernel(const void* __restrict__ dy, int dy_dtype, const void* __restrict__ x, int x_dtype,
                                     const float* __restrict__ mean, const float* __restrict__ rstd,
                                     const void* __restrict__ gamma, int w_dtype, const float* __restrict__ sums,
                                     void* __restrict__ dx, int dx_dtype, int rows, int H, int H_total) {
  const float inv_h = 1.f / H_total;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const float mu = mean[row], rs = rstd[row];
    const float s1 = sums[2 * row], s2 = sums[2 * row + 1];
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      const float xh = (load1(x, x_dtype, (size_t)row * H + c) - mu) * rs;
      const float dgm = load1(dy, dy_dtype, (size_t)row * H + c) * (gamma ? load1(gamma, w_dtype, c) : 1.f);
      store1(dx, dx_dtype, (size_t)row * H + c, rs * (dgm - (s1 + xh * s2) * inv_h));
    }
  }
}

inline void pick_geometry(int H, int& threads, int& nv) {
  const int nvec = H >> 3;
  threads = 64;
  while (threads < 1024 && (nvec + threads - 1) / threads > 2) threads <<= 1;
  nv = (nvec + threads - 1) / threads;
}

}  // namespace

SMPK_API int smpk_ln_bwd_grid(int rows, int H) {
  (void)H;
  int g = smpk::num_sms() * 2;
  return rows < g ? (rows > 0 ? rows : 1) : g;
}

SMPK_API int smpk_layernorm_fwd(void* stream, const void* x, const void* residual, void* sum_out, int x_dtype,
                                const void* gamma, const void* beta, int w_dtype, void* y, int y_dtype, float* mean,
                                float* rstd, int rows, int H, float eps) {
  SMPK_REQUIRE(rows >= 0 && H > 0, "layernorm_fwd: bad shape rows=%d H=%d", rows, H);
  if (rows == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int threads, nv;
  pick_geometry(H, threads, nv);
  const bool vec = (H % 8 == 0) && nv <= MAX_NV && aligned16(x) && aligned16(y) && (!residual || aligned16(residual)) &&
                   (!sum_out || aligned16(sum_out)) && (!gamma || aligned16(gamma)) && (!beta || aligned16(beta));
  const int grid = rows < smpk::num_sms() * 8 ? rows : smpk::num_sms() * 8;
  if (!vec) {
    ln_fwd_scalar_kernel<<<grid, 256, 0, s>>>(x, residual, sum_out, x_dtype, gamma, beta, w_dtype, y, y_dtype, mean,
                                              rstd, rows, H, eps);
  } else {
#define LAUNCH(NV)                                                                                                 \
  ln_fwd_kernel<NV><<<grid, threads, 0, s>>>(x, residual, sum_out, x_dtype, gamma, beta, w_dtype, y, y_dtype, mean, \
                                             rstd, rows, H, eps)
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

// scratch: fp32 [2 * smpk_ln_bwd_grid(rows,H) * H]; may be null when neither dgamma nor dbeta is wanted.
SMPK_API int smpk_layernorm_bwd(void* stream, const void* dy, int dy_dtype, const void* x, int x_dtype,
                                const float* mean, const float* rstd, const void* gamma, int w_dtype, void* dx,
                                int dx_dtype, const void* dx_add, void* dgamma, void* dbeta, int accumulate_wgrad,
                                float* scratch, int rows, int H) {
  SMPK_REQUIRE(rows >= 0 && H > 0, "layernorm_bwd: bad shape rows=%d H=%d", rows, H);
  if (rows == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool want_w = dgamma != nullptr || dbeta != nullptr;
  SMPK_REQUIRE(!want_w || scratch != nullptr, "layernorm_bwd: scratch required for dgamma/dbeta");
  int threads, nv;
  pick_geometry(H, threads, nv);
  const bool vec = (H % 8 == 0) && nv <= MAX_NV && aligned16(x) && aligned16(dy) && aligned16(dx) &&
                   (!dx_add || aligned16(dx_add)) && (!gamma || aligned16(gamma)) && (!scratch || aligned16(scratch));
  const int grid = smpk_ln_bwd_grid(rows, H);
  if (!vec) {
    if (want_w) SMPK_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float) * 2 * (size_t)grid * H, s));
    ln_bwd_scalar_kernel<<<grid, 256, 0, s>>>(dy, dy_dtype, x, x_dtype, mean, rstd, gamma, w_dtype, dx, dx_dtype,
                                              dx_add, want_w ? scratch : nullptr, rows, H);
  } else {
#define LAUNCH(NV)                                                                                               \
  ln_bwd_kernel<NV><<<grid, threads, 0, s>>>(dy, dy_dtype, x, x_dtype, mean, rstd, gamma, w_dtype, dx, dx_dtype, \
                                             dx_add, want_w ? scratch : nullptr, rows, H)
    switch (nv) {
      case 1: LAUNCH(1); break;
      case 2: LAUNCH(2); break;
      case 3: LAUNCH(3); break;
      default: LAUNCH(4); break;
    }
#undef LAUNCH
  }
  SMPK_CHECK_CUDA(cudaGetLastError());
  if (want_w) {
    ln_bwd_reduce_kernel<<<(H + 127) / 128, 128, 0, s>>>(scratch, grid, H, dgamma, dbeta, w_dtype, accumulate_wgrad);
    SMPK_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}

SMPK_API int smpk_ln_partial_stats(void* stream, const void* x, int x_dtype, float* stats, int rows, int H) {
  if (rows == 0) return 0;
  const int grid = rows < smpk::num_sms() * 8 ? rows : smpk::num_sms() * 8;
  ln_partial_stats_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, x_dtype, stats, rows, H);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
SMPK_API int smpk_ln_apply_stats(void* stream, const void* x, int x_dtype, const float* mean, const float* rstd,
                                 const void* gamma, const void* beta, int w_dtype, void* y, int y_dtype, int rows,
                                 int H) {
  if (rows == 0) return 0;
  const int grid = rows < smpk::num_sms() * 8 ? rows : smpk::num_sms() * 8;
  ln_apply_stats_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, x_dtype, mean, rstd, gamma, beta,
                                                                                   w_dtype, y, y_dtype, rows, H);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
SMPK_API int smpk_ln_bwd_local_sums(void* stream, const void* dy, int dy_dtype, const void* x, int x_dtype,
                                    const float* mean, const float* rstd, const void* gamma, int w_dtype, float* sums,
                                    int rows, int H) {
  if (rows == 0) return 0;
  const int grid = rows < smpk::num_sms() * 8 ? rows : smpk::num_sms() * 8;
  ln_bwd_local_sums_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(dy, dy_dtype, x, x_dtype, mean, rstd,
                                                                                      gamma, w_dtype, sums, rows, H);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
SMPK_API int smpk_ln_bwd_finish(void* stream, const void* dy, int dy_dtype, const void* x, int x_dtype,
                                const float* mean, const float* rstd, const void* gamma, int w_dtype,
                                const float* sums, void* dx, int dx_dtype, int rows, int H, int H_total) {
  if (rows == 0) return 0;
  const int grid = rows < smpk::num_sms() * 8 ? rows : smpk::num_sms() * 8;
  ln_bwd_finish_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      dy, dy_dtype, x, x_dtype, mean, rstd, gamma, w_dtype, sums, dx, dx_dtype, rows, H, H_total);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
