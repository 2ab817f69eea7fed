// Fused (vocab-parallel capable) softmax cross-entropy for sm_100a.
//
// Phase 1 (smpk_ce_stats): one block per token row does an online-softmax sweep over this rank's
// vocabulary shard and emits {row max, sum exp(logit - max), target logit or 0}.
// Between the phases the three [rows] vectors are reduced across the tensor-parallel group
// (MAX, then rescaled SUM, SUM) - a no-op for tp=1.
// Phase 2 (smpk_ce_finish): writes the per-token loss and overwrites the logits in place with
// d loss / d logits, so the [tokens, vocab/tp] activation is never kept twice.
//
// Replaces the eager-op composition in the reference's _VocabParallelCrossEntropy (reference
// smdistributed/modelparallel/torch/nn/cross_entropy.py:28-98: max, masked gather of the target
// logit, sum-exp, log, and the softmax-minus-onehot backward).
#include "common.cuh"

using namespace smpk;

namespace {

__global__ void ce_stats_kernel(const void* __restrict__ logits, int dtype, const long long* __restrict__ targets,
                                long long vocab_start, int rows, int V, long long ld, float* __restrict__ row_max,
                                float* __restrict__ row_sum, float* __restrict__ row_tgt) {
  __shared__ float red[32];
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * ld;
    const bool vec = (V % 8 == 0) && (ld % 8 == 0) &&
                     ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
    float m = -INFINITY, s = 0.f;
    if (vec) {
      for (int v = threadIdx.x; v < (V >> 3); v += blockDim.x) {
        float x[8];
        load8(logits, dtype, base + (size_t)v * 8, x);
        float lm = x[0];
#pragma unroll
        for (int j = 1; j < 8; ++j) lm = fmaxf(lm, x[j]);
        if (lm > m) {
          s *= __expf(m - lm);
          m = lm;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += __expf(x[j] - m);
      }
    } else {
      for (int c = threadIdx.x; c < V; c += blockDim.x) {
        const float x = load1(logits, dtype, base + c);
        if (x > m) {
          s *= __expf(m - x);
          m = x;
        }
        s += __expf(x - m);
      }
    }
    const float gm = block_max(m, red);
    const float contrib = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    const float gs = block_sum(contrib, red);
    if (threadIdx.x == 0) {
      row_max[row] = gm;
      row_sum[row] = gs;
      const long long t = targets[row] - vocab_start;
      row_tgt[row] = (t >= 0 && t < V) ? load1(logits, dtype, base + (size_t)t) : 0.f;
    }
  }
}

__global__ void ce_finish_kernel(void* __restrict__ logits, int dtype, const long long* __restrict__ targets,
                                 long long vocab_start, int rows, int V, long long ld, const float* __restrict__ gmax,
                                 const float* __restrict__ gsum, const float* __restrict__ gtgt, float grad_scale,
                                 long long ignore_index, float* __restrict__ loss, int write_grad) {
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = (size_t)row * ld;
    const float m = gmax[row];
    const float inv = 1.f / gsum[row];
    const long long tgt = targets[row];
    const bool ignored = (tgt == ignore_index);
    if (threadIdx.x == 0 && loss != nullptr) loss[row] = ignored ? 0.f : (__logf(gsum[row]) + m - gtgt[row]);
    if (!write_grad) continue;
    const long long t = tgt - vocab_start;
    const float gsc = ignored ? 0.f : grad_scale;
    const bool vec = (V % 8 == 0) && (ld % 8 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
    if (vec) {
      for (int v = threadIdx.x; v < (V >> 3); v += blockDim.x) {
        float x[8];
        load8(logits, dtype, base + (size_t)v * 8, x);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float p = __expf(x[j] - m) * inv;
          if ((long long)v * 8 + j == t) p -= 1.f;
          x[j] = p * gsc;
        }
        store8(logits, dtype, base + (size_t)v * 8, x);
      }
    } else {
      for (int c = threadIdx.x; c < V; c += blockDim.x) {
        float p = __expf(load1(logits, dtype, base + c) - m) * inv;
        if (c == t) p -= 1.f;
        store1(logits, dtype, base + c, p * gsc);
      }
    }
  }
}

}  // namespace

SMPK_API int smpk_ce_stats(void* stream, const void* logits, int dtype, const long long* targets,
                           long long vocab_start, int rows, int V, long long ld, float* row_max, float* row_sum,
                           float* row_tgt) {
  if (rows == 0) return 0;
  const int cap = smpk::num_sms() * 4;
  ce_stats_kernel<<<rows < cap ? rows : cap, 512, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      logits, dtype, targets, vocab_start, rows, V, ld, row_max, row_sum, row_tgt);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_ce_finish(void* stream, void* logits, int dtype, const long long* targets, long long vocab_start,
                            int rows, int V, long long ld, const float* gmax, const float* gsum, const float* gtgt,
                            float grad_scale, long long ignore_index, float* loss, int write_grad) {
  if (rows == 0) return 0;
  const int cap = smpk::num_sms() * 4;
  ce_finish_kernel<<<rows < cap ? rows : cap, 512, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      logits, dtype, targets, vocab_start, rows, V, ld, gmax, gsum, gtgt, grad_scale, ignore_index, loss, write_grad);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
