// Multi-tensor apply kernels for sm_100a: scale / axpby / l2norm / Adam(W) / SGD / Adagrad /
// LAMB (two stages) / NovoGrad over lists of tensors, with per-list runtime dtypes so the same
// kernels serve fp32 master weights, bf16/fp16 gradients and the fused "update master + write
// low-precision copy" path used by the mixed-precision optimizer.
//
// Same role as the reference's amp_C extension (multi_tensor_scale, _axpby, _l2norm, _adam, _lamb,
// _novograd, _sgd, _adagrad; call sites reference smdistributed/modelparallel/torch/fp16/
// fp16.py:132,159,555, fp16util.py:205, apex/optimizers/fused_adam.py:148-170, fused_lamb.py:77-80,
// fused_novograd.py:85, fused_sgd.py:104, fused_adagrad.py:55).  Tensor addresses travel in the
// kernel argument block (no device-side pointer tables, no allocation); long lists are split over
// several launches.
#include "common.cuh"

using namespace smpk;

namespace {

constexpr int MT_MAX_TENSORS = 48;
constexpr int MT_MAX_BLOCKS = 640;
constexpr int MT_THREADS = 512;

template <int DEPTH>
struct MTMeta {
  void* addr[DEPTH][MT_MAX_TENSORS];
  long long size[MT_MAX_TENSORS];
  int dtype[DEPTH];
  int tensor_base;  // index of addr[.][0] in the caller's full list (for per-tensor side arrays)
  unsigned char block_to_tensor[MT_MAX_BLOCKS];
  int block_to_chunk[MT_MAX_BLOCKS];
};

// ---- element functors: f(vals[DEPTH]) mutates vals; WRITE_MASK bit d => list d is stored back.
struct ScaleOp {
  static constexpr int DEPTH = 2;
  static constexpr unsigned WRITE_MASK = 0b10;
  float scale;
  const float* scale_dev;
  int* flag;
  __device__ __forceinline__ void prepare(int) {
    if (scale_dev) scale *= *scale_dev;
  }
  __device__ __forceinline__ void operator()(float (&v)[2]) const {
    if (!isfinite(v[0])) *flag = 1;
    v[1] = v[0] * scale;
  }
};
struct AxpbyOp {
  static constexpr int DEPTH = 3;
  static constexpr unsigned WRITE_MASK = 0b100;
  float a, b;
  int check;  // -1: both, 0: x, 1: y
  int* flag;
  __device__ __forceinline__ void prepare(int) {}
  __device__ __forceinline__ void operator()(float (&v)[3]) const {
    const bool bad = (check != 1 && !isfinite(v[0])) || (check != 0 && !isfinite(v[1]));
    if (bad) *flag = 1;
    v[2] = a * v[0] + b * v[1];
  }
};
struct AdamArgs {
  float lr, beta1, beta2, eps, wd, bc1, bc2, grad_scale;
  const float* grad_scale_dev;
  int adamw;
};
template <int D>
struct AdamOp {  // lists: g, p, m, v [, p_lowprec]
  static constexpr int DEPTH = D;
  static constexpr unsigned WRITE_MASK = (D == 5) ? 0b11110 : 0b1110;
  AdamArgs a;
  __device__ __forceinline__ void prepare(int) {
    if (a.grad_scale_dev) a.grad_scale *= *a.grad_scale_dev;
  }
  __device__ __forceinline__ void operator()(float (&v)[D]) const {
    float g = v[0] * a.grad_scale, p = v[1], m = v[2], s = v[3];
    if (!a.adamw) g += a.wd * p;
    m = a.beta1 * m + (1.f - a.beta1) * g;
    s = a.beta2 * s + (1.f - a.beta2) * g * g;
    const float mh = m / a.bc1, sh = s / a.bc2;
    float upd = mh / (sqrtf(sh) + a.eps);
    if (a.adamw) upd += a.wd * p;
    p -= a.lr * upd;
    v[1] = p; v[2] = m; v[3] = s;
    if (D == 5) v[D - 1] = p;
  }
};
struct SgdArgs {
  float lr, momentum, dampening, wd, grad_scale;
  int nesterov, first_run, wd_after;
};
template <int D>
struct SgdOp {  // lists: g, p, mom [, p_lowprec]
  static constexpr int DEPTH = D;
  static constexpr unsigned WRITE_MASK = (D == 4) ? 0b1110 : 0b110;
  SgdArgs a;
  __device__ __forceinline__ void prepare(int) {}
  __device__ __forceinline__ void operator()(float (&v)[D]) const {
    float g = v[0] * a.grad_scale, p = v[1], mom = v[2];
    if (a.wd != 0.f && !a.wd_after) g += a.wd * p;
    if (a.momentum != 0.f) {
      mom = a.first_run ? g : mom * a.momentum + (1.f - a.dampening) * g;
      g = a.nesterov ? g + a.momentum * mom : mom;
    }
    if (a.wd != 0.f && a.wd_after) g += a.wd * p;
    p -= a.lr * g;
    v[1] = p; v[2] = mom;
    if (D == 4) v[3] = p;
  }
};
struct AdagradArgs {
  float lr, eps, wd, grad_scale;
  int decoupled;
};
template <int D>
struct AdagradOp {  // lists: g, p, h [, p_lowprec]
  static constexpr int DEPTH = D;
  static constexpr unsigned WRITE_MASK = (D == 4) ? 0b1110 : 0b110;
  AdagradArgs a;
  __device__ __forceinline__ void prepare(int) {}
  __device__ __forceinline__ void operator()(float (&v)[D]) const {
    float g = v[0] * a.grad_scale, p = v[1], h = v[2];
    if (!a.decoupled) g += a.wd * p;
    h += g * g;
    float upd = g / (sqrtf(h) + a.eps);
    if (a.decoupled) upd += a.wd * p;
    p -= a.lr * upd;
    v[1] = p; v[2] = h;
    if (D == 4) v[3] = p;
  }
};
struct LambArgs {
  float beta1, beta2, beta3, eps, wd, bc1, bc2, grad_scale;
  const float* global_grad_norm;  // optional device scalar
  float max_grad_norm;
  int adamw;
};
struct LambStage1Op {  // lists: g, p, m, v, update(out)
  static constexpr int DEPTH = 5;
  static constexpr unsigned WRITE_MASK = 0b11100;
  LambArgs a;
  float clip;
  __device__ __forceinline__ void prepare(int) {
    clip = 1.f;
    if (a.global_grad_norm && a.max_grad_norm > 0.f) {
      const float n = *a.global_grad_norm * a.grad_scale;
      clip = n > a.max_grad_norm ? a.max_grad_norm / n : 1.f;
    }
  }
  __device__ __forceinline__ void operator()(float (&v)[5]) const {
    float g = v[0] * a.grad_scale * clip, p = v[1], m = v[2], s = v[3];
    if (!a.adamw) g += a.wd * p;
    m = a.beta1 * m + a.beta3 * g;
    s = a.beta2 * s + (1.f - a.beta2) * g * g;
    float upd = (m / a.bc1) / (sqrtf(s / a.bc2) + a.eps);
    if (a.adamw) upd += a.wd * p;
    v[2] = m; v[3] = s; v[4] = upd;
  }
};
template <int D>
struct LambStage2Op {  // lists: update, p [, p_lowprec]; per-tensor trust ratio from norms
  static constexpr int DEPTH = D;
  static constexpr unsigned WRITE_MASK = (D == 3) ? 0b110 : 0b10;
  float lr, wd;
  int use_nvlamb;
  const float* p_norms;
  const float* u_norms;
  float ratio;
  __device__ __forceinline__ void prepare(int tensor_global) {
    const float pn = p_norms[tensor_global], un = u_norms[tensor_global];
    ratio = lr;
    if ((use_nvlamb || wd != 0.f) && pn > 0.f && un > 0.f) ratio = lr * (pn / un);
  }
  __device__ __forceinline__ void operator()(float (&v)[D]) const {
    const float p = v[1] - ratio * v[0];
    v[1] = p;
    if (D == 3) v[2] = p;
  }
};
struct NovoGradArgs {
  float lr, beta1, beta3, eps, wd, bc1, bc2, grad_scale;
  int moment_mode;  // 0: L2 regularisation inside the moment, 1: decoupled
};
template <int D>
struct NovoGradOp {  // lists: g, p, m [, p_lowprec]; per-tensor second moment (already a norm) in v_norms
  static constexpr int DEPTH = D;
  static constexpr unsigned WRITE_MASK = (D == 4) ? 0b1110 : 0b110;
  NovoGradArgs a;
  const float* v_norms;
  float denom;
  __device__ __forceinline__ void prepare(int tensor_global) { denom = v_norms[tensor_global] / a.bc2 + a.eps; }
  __device__ __forceinline__ void operator()(float (&v)[D]) const {
    float g = v[0] * a.grad_scale, p = v[1], m = v[2];
    if (a.moment_mode == 0) {
      g = g / denom + a.wd * p;
      m = a.beta1 * m + a.beta3 * g;
      p -= a.lr * (m / a.bc1);
    } else {
      m = a.beta1 * m + a.beta3 * g;
      p -= a.lr * ((m / a.bc1) / denom + a.wd * p);
    }
    v[1] = p; v[2] = m;
    if (D == 4) v[3] = p;
  }
};

template <typename Op>
__global__ void __launch_bounds__(MT_THREADS) mt_apply_kernel(MTMeta<Op::DEPTH> meta, int chunk, Op op,
                                                              const int* __restrict__ noop) {
  if (noop != nullptr && *noop != 0) return;
  constexpr int D = Op::DEPTH;
  const int t = meta.block_to_tensor[blockIdx.x];
  const long long start = (long long)meta.block_to_chunk[blockIdx.x] * chunk;
  long long n = meta.size[t] - start;
  if (n > chunk) n = chunk;
  op.prepare(meta.tensor_base + t);
  char* base[D];
  bool vec = (start % 8 == 0);
#pragma unroll
  for (int d = 0; d < D; ++d) {
    base[d] = reinterpret_cast<char*>(meta.addr[d][t]) + start * dtype_size(meta.dtype[d]);
    vec = vec && ((reinterpret_cast<uintptr_t>(base[d]) & 15) == 0);
  }
  if (vec) {
    const long long nv = n >> 3;
    for (long long i = threadIdx.x; i < nv; i += blockDim.x) {
      float in[D][8];
#pragma unroll
      for (int d = 0; d < D; ++d) {
        // pure-output lists need no load unless the op reads them; loads are cheap relative to clarity
        load8(base[d], meta.dtype[d], (size_t)i * 8, in[d]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v[D];
#pragma unroll
        for (int d = 0; d < D; ++d) v[d] = in[d][j];
        op(v);
#pragma unroll
        for (int d = 0; d < D; ++d) in[d][j] = v[d];
      }
#pragma unroll
      for (int d = 0; d < D; ++d)
        if (Op::WRITE_MASK & (1u << d)) store8(base[d], meta.dtype[d], (size_t)i * 8, in[d]);
    }
    // tail
    for (long long i = (nv << 3) + threadIdx.x; i < n; i += blockDim.x) {
      float v[D];
#pragma unroll
      for (int d = 0; d < D; ++d) v[d] = load1(base[d], meta.dtype[d], (size_t)i);
      op(v);
#pragma unroll
      for (int d = 0; d < D; ++d)
        if (Op::WRITE_MASK & (1u << d)) store1(base[d], meta.dtype[d], (size_t)i, v[d]);
    }
  } else {
    for (long long i = threadIdx.x; i < n; i += blockDim.x) {
      float v[D];
#pragma unroll
      for (int d = 0; d < D; ++d) v[d] = load1(base[d], meta.dtype[d], (size_t)i);
      op(v);
#pragma unroll
      for (int d = 0; d < D; ++d)
        if (Op::WRITE_MASK & (1u << d)) store1(base[d], meta.dtype[d], (size_t)i, v[d]);
    }
  }
}

// Sum of squares (or max-abs) per chunk.  partials[blockIdx.x] accumulates across launches; the
// per-tensor array (optional) is accumulated with atomics.
__global__ void __launch_bounds__(MT_THREADS) mt_l2norm_kernel(MTMeta<1> meta, int chunk, float* __restrict__ partials,
                                                               float* __restrict__ per_tensor_sq, int* flag,
                                                               const int* __restrict__ noop) {
  if (noop != nullptr && *noop != 0) return;
  __shared__ float red[32];
  const int t = meta.block_to_tensor[blockIdx.x];
  const long long start = (long long)meta.block_to_chunk[blockIdx.x] * chunk;
  long long n = meta.size[t] - start;
  if (n > chunk) n = chunk;
  const char* base = reinterpret_cast<const char*>(meta.addr[0][t]) + start * dtype_size(meta.dtype[0]);
  float acc = 0.f;
  const bool vec = (start % 8 == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0);
  long long done = 0;
  if (vec) {
    const long long nv = n >> 3;
    for (long long i = threadIdx.x; i < nv; i += blockDim.x) {
      float v[8];
      load8(base, meta.dtype[0], (size_t)i * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += v[j] * v[j];
    }
    done = nv << 3;
  }
  for (long long i = done + threadIdx.x; i < n; i += blockDim.x) {
    const float v = load1(base, meta.dtype[0], (size_t)i);
    acc += v * v;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    if (!isfinite(acc) && flag) *flag = 1;
    partials[blockIdx.x] += acc;
    if (per_tensor_sq) atomicAdd(&per_tensor_sq[meta.tensor_base + t], acc);
  }
}

__global__ void mt_l2norm_finish_kernel(const float* __restrict__ partials, int nparts, float* __restrict__ out,
                                        float* __restrict__ per_tensor, int ntensors) {
  __shared__ float red[32];
  float a = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) a += partials[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0 && out) out[0] = sqrtf(a);
  if (per_tensor)
    for (int i = threadIdx.x; i < ntensors; i += blockDim.x) per_tensor[i] = sqrtf(per_tensor[i]);
}

template <int DEPTH, typename LaunchFn>
int for_each_launch(int n, void* const* const* lists, const long long* sizes, const int* dtypes, int chunk,
                    LaunchFn&& fn) {
  SMPK_REQUIRE(chunk > 0, "multi_tensor: chunk must be positive");
  MTMeta<DEPTH> meta;
  for (int d = 0; d < DEPTH; ++d) meta.dtype[d] = dtypes[d];
  int nt = 0, nb = 0;
  meta.tensor_base = 0;
  for (int t = 0; t < n; ++t) {
    if (sizes[t] == 0) {
      // keep tensor_base arithmetic exact: an empty tensor still occupies a slot
    }
    if (nt == 0) meta.tensor_base = t;
    for (int d = 0; d < DEPTH; ++d) meta.addr[d][nt] = lists[d][t];
    meta.size[nt] = sizes[t];
    const long long chunks = (sizes[t] + chunk - 1) / chunk;
    ++nt;
    for (long long c = 0; c < chunks; ++c) {
      meta.block_to_tensor[nb] = (unsigned char)(nt - 1);
      meta.block_to_chunk[nb] = (int)c;
      ++nb;
      const bool last_chunk = (c == chunks - 1);
      const bool tensors_full = (nt == MT_MAX_TENSORS && last_chunk);
      const bool blocks_full = (nb == MT_MAX_BLOCKS);
      const bool very_last = (t == n - 1 && last_chunk);
      if (tensors_full || blocks_full || very_last) {
        int rc = fn(meta, nb);
        if (rc) return rc;
        nb = 0;
        if (last_chunk) {
          nt = 0;
        } else {  // carry the partially processed tensor into slot 0 of the next launch
          for (int d = 0; d < DEPTH; ++d) meta.addr[d][0] = meta.addr[d][nt - 1];
          meta.size[0] = meta.size[nt - 1];
          meta.tensor_base = t;
          nt = 1;
        }
      }
    }
    if (chunks == 0 && t == n - 1 && nb > 0) {
      int rc = fn(meta, nb);
      if (rc) return rc;
      nb = 0;
      nt = 0;
    }
    if (chunks == 0 && nt == MT_MAX_TENSORS) {
      if (nb > 0) {
        int rc = fn(meta, nb);
        if (rc) return rc;
      }
      nb = 0;
      nt = 0;
    }
  }
  return 0;
}

template <typename Op>
int run_op(cudaStream_t s, int n, void* const* const* lists, const long long* sizes, const int* dtypes, int chunk,
           const Op& op, const int* noop) {
  return for_each_launch<Op::DEPTH>(n, lists, sizes, dtypes, chunk, [&](const MTMeta<Op::DEPTH>& meta, int nb) -> int {
    mt_apply_kernel<Op><<<nb, MT_THREADS, 0, s>>>(meta, chunk, op, noop);
    SMPK_CHECK_CUDA(cudaGetLastError());
    return 0;
  });
}

}  // namespace

#define MT_COMMON_ARGS void *stream, int n, void *const *const *lists, const long long *sizes, const int *dtypes, int chunk

SMPK_API int smpk_mt_scale(MT_COMMON_ARGS, float scale, const float* scale_dev, int* flag, const int* noop) {
  if (n == 0) return 0;
  ScaleOp op{scale, scale_dev, flag};
  return run_op(reinterpret_cast<cudaStream_t>(stream), n, lists, sizes, dtypes, chunk, op, noop);
}

SMPK_API int smpk_mt_axpby(MT_COMMON_ARGS, float a, float b, int check, int* flag, const int* noop) {
  if (n == 0) return 0;
  AxpbyOp op{a, b, check, flag};
  return run_op(reinterpret_cast<cudaStream_t>(stream), n, lists, sizes, dtypes, chunk, op, noop);
}

// depth 4: g,p,m,v   depth 5: g,p,m,v,p_lowprec
SMPK_API int smpk_mt_adam(MT_COMMON_ARGS, int depth, float lr, float beta1, float beta2, float eps, int step, int adamw,
                          int bias_correction, float wd, float grad_scale, const float* grad_scale_dev,
                          const int* noop) {
  if (n == 0) return 0;
  AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = wd; a.adamw = adamw;
  a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  a.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
  a.grad_scale = grad_scale; a.grad_scale_dev = grad_scale_dev;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (depth == 4) return run_op(s, n, lists, sizes, dtypes, chunk, AdamOp<4>{a}, noop);
  if (depth == 5) return run_op(s, n, lists, sizes, dtypes, chunk, AdamOp<5>{a}, noop);
  return smpk::set_error("mt_adam: depth must be 4 or 5, got %d", depth);
}

SMPK_API int smpk_mt_sgd(MT_COMMON_ARGS, int depth, float lr, float momentum, float dampening, float wd, int nesterov,
                         int first_run, int wd_after, float grad_scale, const int* noop) {
  if (n == 0) return 0;
  SgdArgs a{lr, momentum, dampening, wd, grad_scale, nesterov, first_run, wd_after};
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (depth == 3) return run_op(s, n, lists, sizes, dtypes, chunk, SgdOp<3>{a}, noop);
  if (depth == 4) return run_op(s, n, lists, sizes, dtypes, chunk, SgdOp<4>{a}, noop);
  return smpk::set_error("mt_sgd: depth must be 3 or 4, got %d", depth);
}

SMPK_API int smpk_mt_adagrad(MT_COMMON_ARGS, int depth, float lr, float eps, float wd, int decoupled, float grad_scale,
                             const int* noop) {
  if (n == 0) return 0;
  AdagradArgs a{lr, eps, wd, grad_scale, decoupled};
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (depth == 3) return run_op(s, n, lists, sizes, dtypes, chunk, AdagradOp<3>{a}, noop);
  if (depth == 4) return run_op(s, n, lists, sizes, dtypes, chunk, AdagradOp<4>{a}, noop);
  return smpk::set_error("mt_adagrad: depth must be 3 or 4, got %d", depth);
}

// lists: g, p, m, v, update
SMPK_API int smpk_mt_lamb_stage1(MT_COMMON_ARGS, float beta1, float beta2, float eps, int step, int bias_correction,
                                 float wd, int grad_averaging, int adamw, const float* global_grad_norm,
                                 float max_grad_norm, float grad_scale, const int* noop) {
  if (n == 0) return 0;
  LambArgs a;
  a.beta1 = beta1; a.beta2 = beta2; a.beta3 = grad_averaging ? 1.f - beta1 : 1.f; a.eps = eps; a.wd = wd;
  a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  a.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
  a.grad_scale = grad_scale; a.global_grad_norm = global_grad_norm; a.max_grad_norm = max_grad_norm; a.adamw = adamw;
  LambStage1Op op{a, 1.f};
  return run_op(reinterpret_cast<cudaStream_t>(stream), n, lists, sizes, dtypes, chunk, op, noop);
}

// lists: update, p [, p_lowprec]; p_norms / u_norms: device fp32 [n] per-tensor L2 norms
SMPK_API int smpk_mt_lamb_stage2(MT_COMMON_ARGS, int depth, float lr, float wd, int use_nvlamb, const float* p_norms,
                                 const float* u_norms, const int* noop) {
  if (n == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (depth == 2) return run_op(s, n, lists, sizes, dtypes, chunk, LambStage2Op<2>{lr, wd, use_nvlamb, p_norms, u_norms, 0.f}, noop);
  if (depth == 3) return run_op(s, n, lists, sizes, dtypes, chunk, LambStage2Op<3>{lr, wd, use_nvlamb, p_norms, u_norms, 0.f}, noop);
  return smpk::set_error("mt_lamb_stage2: depth must be 2 or 3, got %d", depth);
}

// lists: g, p, m [, p_lowprec]; v_norms: device fp32 [n] (blended per-tensor grad norms)
SMPK_API int smpk_mt_novograd(MT_COMMON_ARGS, int depth, float lr, float beta1, float beta2, float eps, int step,
                              int bias_correction, float wd, int grad_averaging, int moment_mode, const float* v_norms,
                              float grad_scale, const int* noop) {
  if (n == 0) return 0;
  NovoGradArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta3 = grad_averaging ? 1.f - beta1 : 1.f; a.eps = eps; a.wd = wd;
  a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  a.bc2 = bias_correction ? sqrtf(1.f - powf(beta2, (float)step)) : 1.f;
  a.grad_scale = grad_scale; a.moment_mode = moment_mode;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (depth == 3) return run_op(s, n, lists, sizes, dtypes, chunk, NovoGradOp<3>{a, v_norms, 1.f}, noop);
  if (depth == 4) return run_op(s, n, lists, sizes, dtypes, chunk, NovoGradOp<4>{a, v_norms, 1.f}, noop);
  return smpk::set_error("mt_novograd: depth must be 3 or 4, got %d", depth);
}

SMPK_API int smpk_mt_l2norm_scratch_floats() { return MT_MAX_BLOCKS; }

// out[0] = L2 norm over every tensor; per_tensor (optional, fp32 [n]) gets each tensor's norm.
// scratch: fp32 [smpk_mt_l2norm_scratch_floats()] supplied by the caller.
SMPK_API int smpk_mt_l2norm(MT_COMMON_ARGS, float* out, float* per_tensor, float* scratch, int* flag,
                            const int* noop) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SMPK_REQUIRE(scratch != nullptr && out != nullptr, "mt_l2norm: out and scratch are required");
  SMPK_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float) * MT_MAX_BLOCKS, s));
  if (per_tensor && n > 0) SMPK_CHECK_CUDA(cudaMemsetAsync(per_tensor, 0, sizeof(float) * n, s));
  if (n > 0) {
    int rc = for_each_launch<1>(n, lists, sizes, dtypes, chunk, [&](const MTMeta<1>& meta, int nb) -> int {
      mt_l2norm_kernel<<<nb, MT_THREADS, 0, s>>>(meta, chunk, scratch, per_tensor, flag, noop);
      SMPK_CHECK_CUDA(cudaGetLastError());
      return 0;
    });
    if (rc) return rc;
  }
  mt_l2norm_finish_kernel<<<1, 256, 0, s>>>(scratch, MT_MAX_BLOCKS, out, per_tensor, n);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
