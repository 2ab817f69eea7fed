/* C ABI of libsmpb200_kernels.so - the sm_100a kernel library of smdistributed_modelparallel_b200.
 *
 * Conventions
 *  - every entry point returns 0 on success and non-zero on failure; smpk_last_error() then holds a
 *    thread-local description.
 *  - `stream` is a cudaStream_t passed as void*.  Nothing synchronises the device.
 *  - dtype codes: 0 = float32, 1 = float16, 2 = bfloat16.
 *  - no entry point allocates device memory; scratch buffers are supplied by the caller and their
 *    sizes are reported by the *_scratch_floats helpers.
 *
 * Where an entry point replaces a native function of the reference wheel the comment cites the
 * reference call site (paths relative to smdistributed/modelparallel/).
 */
#ifndef SMPB200_KERNELS_H_
#define SMPB200_KERNELS_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* smpk_last_error(void);
int smpk_version(void);
int smpk_device_cc(void);

/* ---- tcgen05 GEMM ------------------------------------------------------------------------------
 * C[b] = epilogue(alpha * A[b] (MxK) * B[b]^T (NxK)),  b in [0, batch).
 * a_mn_major / b_mn_major = 0: operand stored K-major (row = M resp. N index, K contiguous, ld = row
 * stride);  = 1: stored MN-major (row = K index, M resp. N contiguous).  Strides are in elements and
 * must be multiples of 8; inputs are bf16 or fp16, the accumulator is fp32 in tensor memory.
 * Epilogue, in order: + bias[N] ; epi 0: + beta * R ; epi 1/2: aux = value (optional), value =
 * gelu_tanh / gelu_erf(value) ; epi 3/4: value *= gelu_tanh' / gelu_erf'(R).  R may alias C
 * (in-place accumulate).  max_ctas > 0 caps the persistent grid (to leave SMs to a side stream).
 * Replaces cuBLAS calls at torch/nn/transformer.py:1107-1134,1476-1482,1515,1657,1810-1812 and the
 * TorchScript bias-GELU at torch/nn/gelu.py:28-47. */
int smpk_gemm(void* stream, int in_dtype, int out_dtype, int M, int N, int K, int batch, const void* A, int lda,
              long long a_bs, int a_mn_major, const void* B, int ldb, long long b_bs, int b_mn_major, void* C, int ldc,
              long long c_bs, const void* bias, int bias_dtype, const void* R, int ldr, long long r_bs, int r_dtype,
              void* aux, float alpha, float beta, int epi, int max_ctas);

/* ---- GEMM fused with a tensor-parallel collective over NVLink peer memory -------------------------------
 * mode 1 (all-gather -> GEMM): A = [nranks * rows_per_rank, K] is this GPU's gathered buffer; the local shard is
 *   already in its slot, the first comm_ctas CTAs copy the peers' shards out of peer memory block by block while
 *   the remaining CTAs run the GEMM, local rows first.  mode 2 (GEMM -> reduce-scatter): tiles are stored into the
 *   staging slot of the rank owning their rows (peer_stage[d] is [nranks][rows_per_rank][N]), remote first;
 *   smpk_rs_reduce on the owner sums the slots.  Flags are uint32 epochs in peer memory.
 * Replaces NCCL all-gather / all-reduce next to cuBLAS GEMMs at torch/nn/transformer.py:1436,1474,1522,1141 and
 * torch/nn/utils.py:540-571,701-821. */
typedef struct {
  int mode, nranks, rank, rows_per_rank, comm_ctas;
  unsigned int epoch;
  const void* peer_src[8];
  void* local_buf;
  unsigned int* blk_flags;
  const unsigned int* ready_flags;
  unsigned int* peer_done[8];
  unsigned int* counters;
  void* peer_stage[8];
  unsigned int* peer_arrived[8];
} SmpkCommDesc;
int smpk_gemm_comm(void* stream, int in_dtype, int out_dtype, int M, int N, int K, const void* A, int lda,
                   int a_mn_major, const void* B, int ldb, int b_mn_major, void* C, int ldc, const void* bias,
                   int bias_dtype, const void* R, int ldr, int r_dtype, float beta, void* aux, float alpha, int epi,
                   int max_ctas, const SmpkCommDesc* comm);
/* CTA-pair (tcgen05 cta_group::2, 256 x 256 tiles over the two SMs of a TPC) policy of smpk_gemm: 0 never, 1 when the
 * problem has enough pair tiles (default), 2 whenever N > 128 and M > 128.  Returns the previous mode. */
int smpk_gemm_set_pair_mode(int mode);
int smpk_flags_signal(void* stream, void* const* peer_flags, int n, int index, unsigned int value);
int smpk_flags_wait(void* stream, const void* flags, int n, int skip, unsigned int value);
int smpk_p2p_allgather(void* stream, void* const* peer_src, void* local_buf, long long slot_bytes,
                       const void* ready_flags, void* const* peer_done, void* counters, int nranks, int rank,
                       unsigned int epoch, int ctas);
/* Sharded-data-parallel unit kernels (csrc/kernels/sm100/sdp_comm.cu).  A unit's flat buffer is striped over the
 * ranks in fixed pieces: piece g lives on rank g % n at piece index g / n.  smpk_sdp_gather pulls every piece from
 * its owner's shard by peer loads into `full`; smpk_gemm_comm mode 3 does the same from inside the first GEMM that
 * consumes the unit (B == full, rows_per_rank = shard rows of K elements, 8-row pieces).  smpk_sdp_reduce_adam sums
 * the ranks' fp32 gradient buffers for this rank's pieces (reduce-scatter by peer loads), scales, accumulates into
 * grad_shard (mode 0) or applies Adam/AdamW to master/exp_avg/exp_avg_sq and writes the low-precision shard (mode 1).
 * Engine delegated to DeepSpeed in the reference (torch/optimizers/optimizer.py:46-49,471-491). */
int smpk_sdp_gather(void* stream, void* const* peer_shard, void* full, long long piece_bytes, long long pieces_per_rank,
                    const void* ready, unsigned int epoch, int nranks, int rank, int ctas);
int smpk_sdp_reduce_adam(void* stream, void* const* peer_grad, float* grad_shard, float* master, float* exp_avg,
                         float* exp_avg_sq, void* lowp, int lowp_dtype, long long piece_elems, long long pieces_per_rank,
                         const void* ready, void* const* peer_consumed, void* counter, unsigned int epoch, int nranks,
                         int rank, float scale, float beta, int mode, float lr, float beta1, float beta2, float eps,
                         float wd, int step, int adamw, int bias_correction, float grad_scale, int ctas);
int smpk_rs_reduce(void* stream, const void* staging, const void* arrived, void* const* peer_reduced, void* counter,
                   const void* residual, int res_dtype, const void* bias, int bias_dtype, void* out, int out_dtype,
                   long long rows, int N, int nranks, int rank, int dtype, unsigned int epoch);

/* ---- LayerNorm (Apex fused_layer_norm_cuda: torch/apex/normalization/fused_layer_norm.py:26-78) */
int smpk_layernorm_fwd(void* stream, const void* x, const void* residual, void* sum_out, int x_dtype,
                       const void* gamma, const void* beta, int w_dtype, void* y, int y_dtype, float* mean, float* rstd,
                       int rows, int H, float eps);
/* number of partial-sum slots the backward uses; scratch must hold 2 * grid * H floats */
int smpk_ln_bwd_grid(int rows, int H);
int smpk_layernorm_bwd(void* stream, const void* dy, int dy_dtype, const void* x, int x_dtype, const float* mean,
                       const float* rstd, const void* gamma, int w_dtype, void* dx, int dx_dtype, const void* dx_add,
                       void* dgamma, void* dbeta, int accumulate_wgrad, float* scratch, int rows, int H);
/* hidden-sharded LayerNorm phases (smp_torch_cuda_lib forward_affine_mean_var / backward_affine_local_sums /
 * backward_affine_finish: torch/nn/layer_norm.py:52-81) */
int smpk_ln_partial_stats(void* stream, const void* x, int x_dtype, float* stats, int rows, int H);
int smpk_ln_apply_stats(void* stream, const void* x, int x_dtype, const float* mean, const float* rstd,
                        const void* gamma, const void* beta, int w_dtype, void* y, int y_dtype, int rows, int H);
int smpk_ln_bwd_local_sums(void* stream, const void* dy, int dy_dtype, const void* x, int x_dtype, const float* mean,
                           const float* rstd, const void* gamma, int w_dtype, float* sums, int rows, int H);
int smpk_ln_bwd_finish(void* stream, const void* dy, int dy_dtype, const void* x, int x_dtype, const float* mean,
                       const float* rstd, const void* gamma, int w_dtype, const float* sums, void* dx, int dx_dtype,
                       int rows, int H, int H_total);

/* ---- softmax (smp_torch_cuda_lib scaled_*_softmax_*: torch/nn/softmax.py:7-12,40-93)
 * mode 0: plain, 1: causal, 2: uint8 mask [batch,1,Sq,Sk] (non-zero = masked). */
int smpk_softmax_fwd(void* stream, const void* x, void* y, int dtype, const void* mask, int mode, float scale,
                     int batch_heads, int heads, int Sq, int Sk);
int smpk_softmax_bwd(void* stream, const void* dy, const void* y, void* dx, int dtype, int causal, float scale,
                     int batch_heads, int Sq, int Sk);
int smpk_get_batch_per_block(int sq, int sk, int b, int np);

/* ---- vocab-parallel cross entropy (torch/nn/cross_entropy.py:28-98) */
int smpk_ce_stats(void* stream, const void* logits, int dtype, const long long* targets, long long vocab_start,
                  int rows, int V, long long ld, float* row_max, float* row_sum, float* row_tgt);
int smpk_ce_finish(void* stream, void* logits, int dtype, const long long* targets, long long vocab_start, int rows,
                   int V, long long ld, const float* gmax, const float* gsum, const float* gtgt, float grad_scale,
                   long long ignore_index, float* loss, int write_grad);

/* ---- elementwise / layout (torch/nn/gelu.py:28-47; torch/nn/transformer.py:114-163,1565-1615) */
int smpk_bias_gelu(void* stream, const void* x, const void* bias, int bias_dtype, const void* dy, void* out, int dtype,
                   long long rows, int N, int mode, int backward);
int smpk_colsum_scratch_floats(int N);
int smpk_colsum(void* stream, const void* x, int dtype, int rows, int N, long long ld, void* out, int out_dtype,
                int accumulate, float* scratch);
/* Column sums with the scratch held inside the library (a 2 MiB static device array per GPU) for callers
 * that cannot allocate: bare C-ABI hosts with no workspace cache, plugins inside a process whose device
 * memory is fully owned by another pool, code that must avoid cudaMalloc's device-wide synchronisation.
 * Same result as smpk_colsum.  N must be <= smpk_colsum_static_max_n() (8192), else an error is returned.
 * The first call on a device resolves the array's address (make it before any stream capture); calls that
 * overlap on one GPU race on the shared array.  No counterpart in the reference. */
int smpk_colsum_static_max_n(void);
int smpk_colsum_static(void* stream, const void* x, int dtype, int rows, int N, long long ld, void* out,
                       int out_dtype, int accumulate);
int smpk_qkv_rotary(void* stream, void* fused, void* q, void* k, void* v, int dtype, const float* cosT,
                    const float* sinT, int B, int S, int heads, int d, int rot, int style, int pos_offset, int merge,
                    long long fused_ld);
int smpk_head_transpose(void* stream, const void* src, void* dst, int dtype, int B, int S, int heads, int d,
                        int to_heads_first);

/* ---- fused attention (replaces the reference's baddbmm -> mask -> softmax -> dropout -> matmul composition,
 * torch/nn/transformer.py:1617-1657, forward and backward; no Sq x Sk tensor is written to memory)
 * q, k, v, o, d_o, dq, dk, dv: 16-bit, head size d a multiple of 32 in [32, 256], addressed by element strides
 * (batch, head, sequence; the head dimension is contiguous), so both [B, H, S, d] and [B, S, H, d] work.
 * lse [B*H*Sq] fp32: per-row log-sum-exp in the log2 domain (+inf for rows with no visible key), written by the
 * forward and read by the backward.  delta [B*H*Sq] fp32: scratch the backward fills.  scale multiplies q k^T.
 * causal: key j visible to query i iff j <= i + (Sk - Sq); window > 0 additionally requires j > i + (Sk - Sq) - window.
 * mask: optional bytes, non-zero = excluded, addressed by (m_sb, m_sh, m_sq) element strides (0 broadcasts) with
 * contiguous keys.  p_drop > 0: dropout decided by Philox4x32-10 over (seed, offset, batch*head, query, key / 4) - the
 * backward regenerates the forward's mask from the same seed / offset. */
int smpk_attn_fwd(void* stream, int dtype, int B, int H, int Sq, int Sk, int d, const void* q, long long q_sb,
                  long long q_sh, long long q_ss, const void* k, long long k_sb, long long k_sh, long long k_ss,
                  const void* v, long long v_sb, long long v_sh, long long v_ss, void* o, long long o_sb,
                  long long o_sh, long long o_ss, float* lse, float scale, int causal, int window, const void* mask,
                  long long m_sb, long long m_sh, long long m_sq, float p_drop, unsigned long long seed,
                  unsigned long long offset);
int smpk_attn_bwd(void* stream, int dtype, int B, int H, int Sq, int Sk, int d, const void* q, long long q_sb,
                  long long q_sh, long long q_ss, const void* k, long long k_sb, long long k_sh, long long k_ss,
                  const void* v, long long v_sb, long long v_sh, long long v_ss, const void* o, long long o_sb,
                  long long o_sh, long long o_ss, const void* d_o, long long g_sb, long long g_sh, long long g_ss,
                  const float* lse, float* delta, void* dq, long long dq_sb, long long dq_sh, long long dq_ss, void* dk,
                  long long dk_sb, long long dk_sh, long long dk_ss, void* dv, long long dv_sb, long long dv_sh,
                  long long dv_ss, float scale, int causal, int window, const void* mask, long long m_sb,
                  long long m_sh, long long m_sq, float p_drop, unsigned long long seed, unsigned long long offset);

/* ---- multi-tensor apply (amp_C: torch/fp16/fp16.py:132,159,555; torch/apex/optimizers/*.py)
 * lists: array of `depth` pointers, each to an array of n device addresses; sizes: n element counts;
 * dtypes: one code per list; chunk: elements per thread block. noop (device int, optional): skip
 * the whole launch when non-zero.  flag (device int): set to 1 when a non-finite input is seen. */
#define SMPK_MT_ARGS void *stream, int n, void *const *const *lists, const long long *sizes, const int *dtypes, int chunk
int smpk_mt_scale(SMPK_MT_ARGS, float scale, const float* scale_dev, int* flag, const int* noop);
int smpk_mt_axpby(SMPK_MT_ARGS, float a, float b, int check, int* flag, const int* noop);
int smpk_mt_adam(SMPK_MT_ARGS, int depth, float lr, float beta1, float beta2, float eps, int step, int adamw,
                 int bias_correction, float wd, float grad_scale, const float* grad_scale_dev, const int* noop);
int smpk_mt_sgd(SMPK_MT_ARGS, int depth, float lr, float momentum, float dampening, float wd, int nesterov,
                int first_run, int wd_after, float grad_scale, const int* noop);
int smpk_mt_adagrad(SMPK_MT_ARGS, int depth, float lr, float eps, float wd, int decoupled, float grad_scale,
                    const int* noop);
int smpk_mt_lamb_stage1(SMPK_MT_ARGS, float beta1, float beta2, float eps, int step, int bias_correction, float wd,
                        int grad_averaging, int adamw, const float* global_grad_norm, float max_grad_norm,
                        float grad_scale, const int* noop);
int smpk_mt_lamb_stage2(SMPK_MT_ARGS, int depth, float lr, float wd, int use_nvlamb, const float* p_norms,
                        const float* u_norms, const int* noop);
int smpk_mt_novograd(SMPK_MT_ARGS, int depth, float lr, float beta1, float beta2, float eps, int step,
                     int bias_correction, float wd, int grad_averaging, int moment_mode, const float* v_norms,
                     float grad_scale, const int* noop);
int smpk_mt_l2norm_scratch_floats(void);
int smpk_mt_l2norm(SMPK_MT_ARGS, float* out, float* per_tensor, float* scratch, int* flag, const int* noop);

#ifdef __cplusplus
}
#endif
#endif /* SMPB200_KERNELS_H_ */
