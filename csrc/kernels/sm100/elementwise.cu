// Small memory-bound fused ops for sm_100a: bias+GELU forward/backward (when not folded into a GEMM
// epilogue), column sums (bias gradients), QKV split + rotary embedding (GPT-J interleaved and
// GPT-NeoX half-rotate) with the matching merge/inverse for backward, and head merge.
//
// bias-GELU mirrors the reference's TorchScript functions bias_gelu / bias_gelu_back (reference
// smdistributed/modelparallel/torch/nn/gelu.py:28-47); rotary mirrors the eager-op code at
// torch/nn/transformer.py:114-163,1565-1615.
#include "common.cuh"

using namespace smpk;

namespace {

// y = gelu(x + bias)  (mode 0 tanh, 1 erf);  backward: dx = dy * gelu'(x + bias)
__global__ void bias_gelu_kernel(const void* __restrict__ x, const void* __restrict__ bias, int bias_dtype,
                                 const void* __restrict__ dy, void* __restrict__ out, int dtype, long long total, int N,
                                 int mode, int backward) {
  const long long nvec = total >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float v[8], b[8], g[8];
    load8(x, dtype, (size_t)i * 8, v);
    if (bias != nullptr) {
      load8(bias, bias_dtype, (size_t)((i * 8) % N), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += b[j];
    }
    if (backward) {
      load8(dy, dtype, (size_t)i * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = g[j] * (mode == 0 ? gelu_tanh_grad(v[j]) : gelu_erf_grad(v[j]));
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = mode == 0 ? gelu_tanh(v[j]) : gelu_erf(v[j]);
    }
    store8(out, dtype, (size_t)i * 8, v);
  }
}

__global__ void bias_gelu_scalar_kernel(const void* __restrict__ x, const void* __restrict__ bias, int bias_dtype,
                                        const void* __restrict__ dy, void* __restrict__ out, int dtype, long long total,
                                        int N, int mode, int backward) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    float v = load1(x, dtype, (size_t)i);
    if (bias != nullptr) v += load1(bias, bias_dtype, (size_t)(i % N));
    if (backward)
      v = load1(dy, dtype, (size_t)i) * (mode == 0 ? gelu_tanh_grad(v) : gelu_erf_grad(v));
    else
      v = mode == 0 ? gelu_tanh(v) : gelu_erf(v);
    store1(out, dtype, (size_t)i, v);
  }
}

constexpr int CS_SPLITS = 64;

// partial[split][col] = sum over this split's rows of x[row][col]
__global__ void colsum_partial_kernel(const void* __restrict__ x, int dtype, int rows, int N, long long ld,
                                      float* __restrict__ partial) {
  __shared__ float sm[16][32][9];
  const int vcol = blockIdx.x * 32 + threadIdx.x;  // 8 columns per thread
  const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per, r1 = min(rows, r0 + rows_per);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (vcol * 8 < N) {
    for (int r = r0 + threadIdx.y; r < r1; r += blockDim.y) {
      float v[8];
      load8(x, dtype, (size_t)r * ld + (size_t)vcol * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[threadIdx.y][threadIdx.x][j] = acc[j];
  __syncthreads();
  if (threadIdx.y == 0 && vcol * 8 < N) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = 0.f;
      for (int k = 0; k < 16; ++k) s += sm[k][threadIdx.x][j];
      partial[(size_t)blockIdx.y * N + (size_t)vcol * 8 + j] = s;
    }
  }
}
__global__ void colsum_scalar_partial_kernel(const void* __restrict__ x, int dtype, int rows, int N, long long ld,
                                             float* __restrict__ partial) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per, r1 = min(rows, r0 + rows_per);
  float s = 0.f;
  for (int r = r0; r < r1; ++r) s += load1(x, dtype, (size_t)r * ld + c);
  partial[(size_t)blockIdx.y * N + c] = s;
}
__global__ void colsum_fold_kernel(const float* __restrict__ partial, int nsplit, int N, void* __restrict__ out,
                                   int out_dtype, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  float s = 0.f;
  for (int k = 0; k < nsplit; ++k) s += partial[(size_t)k * N + c];
  if (accumulate) s += load1(out, out_dtype, c);
  store1(out, out_dtype, c, s);
}

// Rotary position embedding fused with the QKV split / head transpose.
//   split   : fused [T = B*S, 3, heads, d]  ->  q,k,v each [B, heads, S, d]   (rotate q,k)
//   merge   : q,k,v grads [B, heads, S, d]  ->  fused grad [T, 3, heads, d]   (inverse-rotate q,k)
// style 0: none, 1: GPT-J interleaved pairs (2i, 2i+1), 2: GPT-NeoX halves (i, i + rot/2).
// cos/sin: fp32 [S_table, rot/2]; position of token s is s + pos_offset.
__global__ void qkv_rotary_kernel(void* __restrict__ fused, void* __restrict__ q, void* __restrict__ k,
                                  void* __restrict__ v, int dtype, const float* __restrict__ cosT,
                                  const float* __restrict__ sinT, int B, int S, int heads, int d, int rot, int style,
                                  int pos_offset, int merge, long long fused_ld) {
  // one thread per (token, which, head, pair-of-elements)
  const int half_d = d >> 1;
  const long long total = (long long)B * S * 3 * heads * half_d;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int pe = (int)(i % half_d);
    long long r = i / half_d;
    const int head = (int)(r % heads);
    r /= heads;
    const int which = (int)(r % 3);
    const long long tok = r / 3;
    const int s = (int)(tok % S);
    const int b = (int)(tok / S);
    // element indices (e0, e1) handled by this thread
    int e0, e1;
    bool rotate = false;
    float c = 1.f, sn = 0.f;
    const int half_rot = rot >> 1;
    if (style == 1) {
      e0 = 2 * pe;
      e1 = e0 + 1;
      if (which < 2 && e0 < rot) {
        rotate = true;
        c = cosT[(size_t)(s + pos_offset) * half_rot + pe];
        sn = sinT[(size_t)(s + pos_offset) * half_rot + pe];
      }
    } else if (style == 2) {
      if (pe < half_rot) {
        e0 = pe;
        e1 = pe + half_rot;
        if (which < 2) {
          rotate = true;
          c = cosT[(size_t)(s + pos_offset) * half_rot + pe];
          sn = sinT[(size_t)(s + pos_offset) * half_rot + pe];
        }
      } else {  // the un-rotated tail [rot, d): pair up consecutive elements
        e0 = rot + 2 * (pe - half_rot);
        e1 = e0 + 1;
      }
    } else {
      e0 = 2 * pe;
      e1 = e0 + 1;
    }
    void* dst3 = which == 0 ? q : (which == 1 ? k : v);
    const size_t f0 = (size_t)tok * fused_ld + ((size_t)which * heads + head) * d;
    const size_t o0 = (((size_t)b * heads + head) * S + s) * d;
    if (!merge) {
      float a = load1(fused, dtype, f0 + e0), bb = load1(fused, dtype, f0 + e1);
      if (rotate) {
        const float ra = a * c - bb * sn, rb = a * sn + bb * c;
        a = ra;
        bb = rb;
      }
      store1(dst3, dtype, o0 + e0, a);
      store1(dst3, dtype, o0 + e1, bb);
    } else {
      float a = load1(dst3, dtype, o0 + e0), bb = load1(dst3, dtype, o0 + e1);
      if (rotate) {  // transpose of the rotation
        const float ra = a * c + bb * sn, rb = -a * sn + bb * c;
        a = ra;
        bb = rb;
      }
      store1(fused, dtype, f0 + e0, a);
      store1(fused, dtype, f0 + e1, bb);
    }
  }
}

// Vectorised form: one thread per 8 contiguous elements (16 bytes) of a (token, q|k|v, head) row.  Rotation
// partners of the interleaved style sit inside the vector; for the half-split style they are fetched with
// scalar loads that hit the lines the neighbouring threads stream.  Needs d % 8 == 0, 16-byte aligned rows.
template <typename T>
__global__ void __launch_bounds__(256) qkv_rotary_vec_kernel(T* __restrict__ fused, T* __restrict__ q, T* __restrict__ k,
                                                             T* __restrict__ v, const float* __restrict__ cosT,
                                                             const float* __restrict__ sinT, int B, int S, int heads,
                                                             int d, int rot, int style, int pos_offset, int merge,
                                                             long long fused_ld) {
  const int dv = d >> 3;
  const int half_rot = rot >> 1;
  const long long total = (long long)B * S * 3 * heads * dv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % dv);
    long long r = i / dv;
    const int head = (int)(r % heads);
    r /= heads;
    const int which = (int)(r % 3);
    const long long tok = r / 3;
    const int s = (int)(tok % S);
    const int b = (int)(tok / S);
    T* dst3 = which == 0 ? q : (which == 1 ? k : v);
    T* frow = fused + (size_t)tok * fused_ld + ((size_t)which * heads + head) * d;
    T* orow = dst3 + (((size_t)b * heads + head) * S + s) * d;
    const T* src = merge ? orow : frow;
    T* dst = merge ? frow : orow;
    const int e0 = j * 8;
    uint4 raw = *reinterpret_cast<const uint4*>(src + e0);
    if (which < 2 && style != 0 && e0 < rot) {
      T* x = reinterpret_cast<T*>(&raw);
      const float* cr = cosT + (size_t)(s + pos_offset) * half_rot;
      const float* sr = sinT + (size_t)(s + pos_offset) * half_rot;
      const float sgn = merge ? -1.f : 1.f;       // backward applies the transposed rotation
      if (style == 1) {
#pragma unroll
        for (int p2 = 0; p2 < 4; ++p2) {
          const int e = e0 + 2 * p2;
          if (e < rot) {
            const float c = cr[e >> 1], sn = sgn * sr[e >> 1];
            const float a = to_f32(x[2 * p2]), bb = to_f32(x[2 * p2 + 1]);
            x[2 * p2] = from_f32<T>(a * c - bb * sn);
            x[2 * p2 + 1] = from_f32<T>(a * sn + bb * c);
          }
        }
      } else {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int e = e0 + u;
          if (e < rot) {
            const bool lo = e < half_rot;
            const int pi = lo ? e : e - half_rot;
            const float c = cr[pi], sn = sgn * sr[pi];
            const float other = to_f32(src[lo ? e + half_rot : e - half_rot]);
            const float me = to_f32(x[u]);
            x[u] = from_f32<T>(lo ? me * c - other * sn : me * c + other * sn);
          }
        }
      }
    }
    *reinterpret_cast<uint4*>(dst + e0) = raw;
  }
}

// [B, heads, S, d] <-> [B*S, heads*d]
__global__ void head_merge_kernel(const void* __restrict__ src, void* __restrict__ dst, int dtype, int B, int S,
                                  int heads, int d, int to_heads_first) {
  const int dv = d >> 3;
  const long long total = (long long)B * S * heads * dv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int ev = (int)(i % dv);
    long long r = i / dv;
    const int head = (int)(r % heads);
    r /= heads;
    const int s = (int)(r % S);
    const int b = (int)(r / S);
    const size_t tok_major = (((size_t)b * S + s) * heads + head) * d + (size_t)ev * 8;
    const size_t head_major = (((size_t)b * heads + head) * S + s) * d + (size_t)ev * 8;
    float v[8];
    if (to_heads_first) {
      load8(src, dtype, tok_major, v);
      store8(dst, dtype, head_major, v);
    } else {
      load8(src, dtype, head_major, v);
      store8(dst, dtype, tok_major, v);
    }
  }
}

inline int ew_grid(long long work_items, int threads) {
  long long g = (work_items + threads - 1) / threads;
  const long long cap = (long long)smpk::num_sms() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

SMPK_API int smpk_bias_gelu(void* stream, const void* x, const void* bias, int bias_dtype, const void* dy, void* out,
                            int dtype, long long rows, int N, int mode, int backward) {
  const long long total = rows * N;
  if (total == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const bool vec = (N % 8 == 0) && aligned16(x) && aligned16(out) && (!bias || aligned16(bias)) && (!dy || aligned16(dy));
  if (vec)
    bias_gelu_kernel<<<ew_grid(total / 8, 256), 256, 0, s>>>(x, bias, bias_dtype, dy, out, dtype, total, N, mode, backward);
  else
    bias_gelu_scalar_kernel<<<ew_grid(total, 256), 256, 0, s>>>(x, bias, bias_dtype, dy, out, dtype, total, N, mode,
                                                                backward);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_colsum_scratch_floats(int N) { return CS_SPLITS * N; }

// out[c] (+)= sum_r x[r][c]; scratch: fp32 [smpk_colsum_scratch_floats(N)]
SMPK_API int smpk_colsum(void* stream, const void* x, int dtype, int rows, int N, long long ld, void* out,
                         int out_dtype, int accumulate, float* scratch) {
  if (N == 0) return 0;
  SMPK_REQUIRE(scratch != nullptr, "colsum: scratch required");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int nsplit = rows < CS_SPLITS ? (rows > 0 ? rows : 1) : CS_SPLITS;
  const bool vec = (N % 8 == 0) && (ld % 8 == 0) && aligned16(x);
  if (vec) {
    dim3 grid((N / 8 + 31) / 32, nsplit), block(32, 16);
    colsum_partial_kernel<<<grid, block, 0, s>>>(x, dtype, rows, N, ld, scratch);
  } else {
    dim3 grid((N + 127) / 128, nsplit);
    colsum_scalar_partial_kernel<<<grid, 128, 0, s>>>(x, dtype, rows, N, ld, scratch);
  }
  SMPK_CHECK_CUDA(cudaGetLastError());
  colsum_fold_kernel<<<(N + 127) / 128, 128, 0, s>>>(scratch, nsplit, N, out, out_dtype, accumulate);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_qkv_rotary(void* stream, void* fused, void* q, void* k, void* v, int dtype, const float* cosT,
                             const float* sinT, int B, int S, int heads, int d, int rot, int style, int pos_offset,
                             int merge, long long fused_ld) {
  SMPK_REQUIRE(d % 2 == 0 && rot % 2 == 0 && rot <= d, "qkv_rotary: d and rot must be even with rot <= d");
  SMPK_REQUIRE(style == 0 || (cosT != nullptr && sinT != nullptr), "qkv_rotary: tables required");
  const long long total = (long long)B * S * 3 * heads * (d / 2);
  if (total == 0) return 0;
  const bool vec = (d % 8 == 0) && (fused_ld % 8 == 0) && dtype != SMPK_F32 && aligned16(fused) && aligned16(q) &&
                   aligned16(k) && aligned16(v);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (vec) {
    const long long nv = total / 4;
    if (dtype == SMPK_BF16)
      qkv_rotary_vec_kernel<__nv_bfloat16><<<ew_grid(nv, 256), 256, 0, st>>>(
          (__nv_bfloat16*)fused, (__nv_bfloat16*)q, (__nv_bfloat16*)k, (__nv_bfloat16*)v, cosT, sinT, B, S, heads, d, rot,
          style, pos_offset, merge, fused_ld);
    else
      qkv_rotary_vec_kernel<__half><<<ew_grid(nv, 256), 256, 0, st>>>((__half*)fused, (__half*)q, (__half*)k, (__half*)v,
                                                                      cosT, sinT, B, S, heads, d, rot, style, pos_offset,
                                                                      merge, fused_ld);
    SMPK_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  qkv_rotary_kernel<<<ew_grid(total, 256), 256, 0, st>>>(
      fused, q, k, v, dtype, cosT, sinT, B, S, heads, d, rot, style, pos_offset, merge, fused_ld);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_head_transpose(void* stream, const void* src, void* dst, int dtype, int B, int S, int heads, int d,
                                 int to_heads_first) {
  SMPK_REQUIRE(d % 8 == 0 && aligned16(src) && aligned16(dst), "head_transpose: d %% 8 == 0 and 16-byte alignment required");
  const long long total = (long long)B * S * heads * (d / 8);
  if (total == 0) return 0;
  head_merge_kernel<<<ew_grid(total, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(src, dst, dtype, B, S,
                                                                                             heads, d, to_heads_first);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Column sums without any caller-supplied or library-allocated scratch.
//
// Who this is for: callers that reach this library through the plain C ABI and have no allocator to
// lean on - a C / Fortran / ctypes host program without torch's workspace cache, a plugin living inside
// another framework that has already claimed the whole device for its own pool (a raw cudaMalloc then
// fails even though that pool has room), or latency-sensitive code that must not trigger the
// device-wide synchronisation cudaMalloc and cudaFree imply.  smpk_colsum() asks such callers for a
// scratch pointer they cannot produce; this variant keeps the scratch inside the library instead.
//
// The scratch is the file-scope device array below: CS_SPLITS x SMPK_COLSUM_STATIC_MAX_N floats
// (64 x 8192 x 4 B = 2 MiB).  Every GPU that loads the library holds those 2 MiB until the process exits.
// Nothing is allocated or freed here, and no allocating routine (cudaMalloc*, cudaFree*, torch) is called.
//
// Set-up: the first call on each device resolves the array's device address with
// cudaGetSymbolAddress (in host code the array's name is only a stand-in, not an address a kernel
// can use).  Make that first call before starting any stream capture on that device.
//
// Sharing: every call on a device uses the same array, which keeps whatever the previous call left in
// it.  Each call overwrites every partial-sum slot it later reads (rows [0, splits) x columns [0, N)),
// so nothing needs to start at zero and stale contents are never observed.  Two calls that overlap on
// the GPU (different streams, or different host threads) race on the array - serialise them.
#define SMPK_COLSUM_STATIC_MAX_N 8192

static __device__ __align__(256) float g_colsum_static_scratch[CS_SPLITS * SMPK_COLSUM_STATIC_MAX_N];

SMPK_API int smpk_colsum_static_max_n() { return SMPK_COLSUM_STATIC_MAX_N; }

SMPK_API int smpk_colsum_static(void* stream, const void* x, int dtype, int rows, int N, long long ld, void* out,
                                int out_dtype, int accumulate) {
  static float* dev_ptr[64] = {nullptr};
  SMPK_REQUIRE(N >= 0 && rows >= 0, "colsum_static: negative shape rows=%d N=%d", rows, N);
  SMPK_REQUIRE(N <= SMPK_COLSUM_STATIC_MAX_N,
               "colsum_static: N=%d exceeds the built-in scratch (max N = %d); use smpk_colsum with your own scratch",
               N, SMPK_COLSUM_STATIC_MAX_N);
  if (N == 0) return 0;
  int dev = 0;
  SMPK_CHECK_CUDA(cudaGetDevice(&dev));
  SMPK_REQUIRE(dev >= 0 && dev < 64, "colsum_static: device index %d out of range", dev);
  if (dev_ptr[dev] == nullptr) {
    float* p = nullptr;
    SMPK_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch));
    SMPK_REQUIRE(p != nullptr, "colsum_static: cudaGetSymbolAddress returned a null address");
    dev_ptr[dev] = p;
  }
  float* scratch = dev_ptr[dev];
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int nsplit = rows < CS_SPLITS ? (rows > 0 ? rows : 1) : CS_SPLITS;
  const bool vec = (N % 8 == 0) && (ld % 8 == 0) && aligned16(x);
  if (vec) {
    dim3 grid((N / 8 + 31) / 32, nsplit), block(32, 16);
    colsum_partial_kernel<<<grid, block, 0, s>>>(x, dtype, rows, N, ld, scratch);
  } else {
    dim3 grid((N + 127) / 128, nsplit);
    colsum_scalar_partial_kernel<<<grid, 128, 0, s>>>(x, dtype, rows, N, ld, scratch);
  }
  SMPK_CHECK_CUDA(cudaGetLastError());
  colsum_fold_kernel<<<(N + 127) / 128, 128, 0, s>>>(scratch, nsplit, N, out, out_dtype, accumulate);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
