// Sharded-data-parallel (ZeRO-3 style) communication kernels over NVLink peer memory.
//
// Layout: a unit's flat buffer is cut into fixed-size pieces; piece g belongs to rank g % n and is stored
// at piece index g / n of that rank's shard ("striped" sharding).  Any contiguous range of the unit is
// therefore spread evenly over all ranks, so every reader pulls from every peer at once instead of all
// readers converging on the one or two ranks that own the weight being multiplied.
//
//   smpk_sdp_gather      parameter all-gather by peer loads (prefetch / backward re-gather); the forward
//                        gather is normally fused into the first consuming GEMM (COMM_AGB in gemm_tcgen05.cu)
//   smpk_sdp_reduce_adam gradient reduce-scatter by peer loads of every rank's fp32 unit gradient buffer,
//                        fused with the scale, the accumulation over microbatches and (optionally) the
//                        sharded Adam update + 16-bit parameter write
//
// The reference delegates this engine to a patched DeepSpeed (reference torch/optimizers/optimizer.py:
// 46-49,471-491; backend/ds_config_defaults.json:6-27), which all-gathers / reduce-scatters
// with NCCL and updates with a separate optimizer kernel.
#include "common.cuh"

using namespace smpk;

namespace {

constexpr int MAX_RANKS = 8;

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void spin_ge(const uint32_t* p, uint32_t v) {
  while ((int32_t)(ld_acquire_sys(p) - v) < 0) __nanosleep(64);
}

struct GatherArgs {
  const void* peer_shard[MAX_RANKS];
  void* full;
  long long piece_bytes, pieces_per_rank;
  const uint32_t* ready;  // [n]: peer j's shard is final for `epoch`
  uint32_t epoch;
  int nranks, rank;
};

__global__ void __launch_bounds__(256) sdp_gather_kernel(GatherArgs a) {
  const int lane = threadIdx.x & 31;
  const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long wstride = (long long)gridDim.x * (blockDim.x >> 5);
  const long long total = a.pieces_per_rank * a.nranks;
  uint32_t waited = 1u << a.rank;
  for (long long pc = wid; pc < total; pc += wstride) {
    const long long q = pc / a.nranks;
    const int src = (a.rank + (int)(pc % a.nranks)) % a.nranks;
    if (!((waited >> src) & 1u)) {
      if (lane == 0) spin_ge(&a.ready[src], a.epoch);
      __syncwarp();
      waited |= 1u << src;
    }
    const uint4* from = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(a.peer_shard[src]) + q * a.piece_bytes);
    uint4* to = reinterpret_cast<uint4*>(reinterpret_cast<char*>(a.full) + (q * a.nranks + src) * a.piece_bytes);
    const int nvec = (int)(a.piece_bytes >> 4);
    constexpr int U = 8;
    for (int i = lane; i < nvec; i += U * 32) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * 32 < nvec) v[u] = from[i + u * 32];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * 32 < nvec) to[i + u * 32] = v[u];
    }
  }
}

struct ReduceArgs {
  const float* peer_grad[MAX_RANKS];  // every rank's full fp32 gradient buffer of the unit (this slot)
  float* grad_shard;                  // [per] local accumulator over microbatches
  float* master;                      // [per] fp32 parameters (Adam only)
  float* exp_avg;
  float* exp_avg_sq;
  void* lowp;                         // [per] 16-bit parameter shard
  int lowp_dtype;
  long long piece_elems, pieces_per_rank;
  const uint32_t* ready;              // [n] local: peer j's gradient slot is complete for `epoch`
  uint32_t* peer_consumed[MAX_RANKS]; // peer j's consumed[n]; [rank] = epoch once I have read its slot
  uint32_t* counter;
  uint32_t epoch;
  int nranks, rank;
  float scale, beta;                  // g = scale * (sum over ranks + beta * grad_shard)
  int mode;                           // 0: write grad_shard; 1: Adam update (grad_shard is read, not written)
  float lr, beta1, beta2, eps, wd, bc1, bc2, grad_scale;
  int adamw;
};

// One thread owns RU float4 vectors per trip (strided by the block so every load instruction is coalesced) and
// pulls them from every rank in turn: RU x 16 B x 256 threads x CTAs stay in flight per source, which is what
// it takes to cover NVLink latency (a single vector per thread leaves the links ~85 % idle).
constexpr int RU = 4;

__device__ __forceinline__ float4 ld_nc_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}

__global__ void __launch_bounds__(256) sdp_reduce_kernel(ReduceArgs a) {
  if (threadIdx.x < a.nranks && threadIdx.x != a.rank) spin_ge(&a.ready[threadIdx.x], a.epoch);
  __syncthreads();
  const long long per = a.piece_elems * a.pieces_per_rank;
  const long long nvec = per >> 2;
  const long long pvec = a.piece_elems >> 2;
  const long long trip = (long long)blockDim.x * RU;
  for (long long base = (long long)blockIdx.x * trip; base < nvec; base += (long long)gridDim.x * trip) {
    long long idx[RU], goff[RU];
    float4 acc[RU];
#pragma unroll
    for (int u = 0; u < RU; ++u) {
      idx[u] = base + (long long)u * blockDim.x + threadIdx.x;
      const long long i = idx[u] < nvec ? idx[u] : 0;
      const long long q = i / pvec;
      goff[u] = (q * a.nranks + a.rank) * pvec + (i - q * pvec);
      acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // optimizer state of this trip, requested before the (slower) peer reads so it is there when they return
    float4 pre_p[RU], pre_m[RU], pre_s[RU];
    if (a.mode == 1) {
#pragma unroll
      for (int u = 0; u < RU; ++u) {
        const long long i = idx[u] < nvec ? idx[u] : 0;
        pre_p[u] = reinterpret_cast<const float4*>(a.master)[i];
        pre_m[u] = reinterpret_cast<const float4*>(a.exp_avg)[i];
        pre_s[u] = reinterpret_cast<const float4*>(a.exp_avg_sq)[i];
      }
    }
    // fixed rank order: the sum is bitwise identical on every run.  Ranks are read two at a time so that the
    // loads of both (2 x RU x 16 B per thread) are in flight before either is consumed.
#pragma unroll
    for (int j = 0; j < MAX_RANKS; j += 2) {
      if (j < a.nranks) {
        const bool two = j + 1 < a.nranks;
        float4 v0[RU], v1[RU];
        const float4* s0 = reinterpret_cast<const float4*>(a.peer_grad[j]);
        const float4* s1 = reinterpret_cast<const float4*>(a.peer_grad[two ? j + 1 : j]);
#pragma unroll
        for (int u = 0; u < RU; ++u) v0[u] = ld_nc_f4(s0 + goff[u]);
        if (two) {
#pragma unroll
          for (int u = 0; u < RU; ++u) v1[u] = ld_nc_f4(s1 + goff[u]);
        }
#pragma unroll
        for (int u = 0; u < RU; ++u) { acc[u].x += v0[u].x; acc[u].y += v0[u].y; acc[u].z += v0[u].z; acc[u].w += v0[u].w; }
        if (two) {
#pragma unroll
          for (int u = 0; u < RU; ++u) { acc[u].x += v1[u].x; acc[u].y += v1[u].y; acc[u].z += v1[u].z; acc[u].w += v1[u].w; }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < RU; ++u) {
      const long long i = idx[u];
      if (i >= nvec) continue;
      float g[4] = {acc[u].x, acc[u].y, acc[u].z, acc[u].w};
      if (a.beta != 0.f) {
        const float4 old = reinterpret_cast<const float4*>(a.grad_shard)[i];
        g[0] += a.beta * old.x; g[1] += a.beta * old.y; g[2] += a.beta * old.z; g[3] += a.beta * old.w;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) g[k] *= a.scale;
      if (a.mode == 0) {
        reinterpret_cast<float4*>(a.grad_shard)[i] = make_float4(g[0], g[1], g[2], g[3]);
        continue;
      }
      const float4 p4 = pre_p[u], m4 = pre_m[u], s4 = pre_s[u];
      float p[4] = {p4.x, p4.y, p4.z, p4.w}, m[4] = {m4.x, m4.y, m4.z, m4.w}, sq[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float gg = g[k] * a.grad_scale;
        if (!a.adamw) gg += a.wd * p[k];
        m[k] = a.beta1 * m[k] + (1.f - a.beta1) * gg;
        sq[k] = a.beta2 * sq[k] + (1.f - a.beta2) * gg * gg;
        float upd = (m[k] / a.bc1) / (sqrtf(sq[k] / a.bc2) + a.eps);
        if (a.adamw) upd += a.wd * p[k];
        p[k] -= a.lr * upd;
      }
      reinterpret_cast<float4*>(a.master)[i] = make_float4(p[0], p[1], p[2], p[3]);
      reinterpret_cast<float4*>(a.exp_avg)[i] = make_float4(m[0], m[1], m[2], m[3]);
      reinterpret_cast<float4*>(a.exp_avg_sq)[i] = make_float4(sq[0], sq[1], sq[2], sq[3]);
      if (a.lowp != nullptr) {
        if (a.lowp_dtype == SMPK_BF16) {
          __nv_bfloat162 lo = __floats2bfloat162_rn(p[0], p[1]), hi = __floats2bfloat162_rn(p[2], p[3]);
          uint2 w;
          w.x = *reinterpret_cast<uint32_t*>(&lo);
          w.y = *reinterpret_cast<uint32_t*>(&hi);
          reinterpret_cast<uint2*>(a.lowp)[i] = w;
        } else if (a.lowp_dtype == SMPK_F16) {
          __half2 lo = __floats2half2_rn(p[0], p[1]), hi = __floats2half2_rn(p[2], p[3]);
          uint2 w;
          w.x = *reinterpret_cast<uint32_t*>(&lo);
          w.y = *reinterpret_cast<uint32_t*>(&hi);
          reinterpret_cast<uint2*>(a.lowp)[i] = w;
        } else {
          reinterpret_cast<float4*>(a.lowp)[i] = make_float4(p[0], p[1], p[2], p[3]);
        }
      }
    }
  }
  // all peers' slots have been read by this rank once every CTA is through
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t old = atomicAdd(a.counter, 1u);
    if (old == gridDim.x - 1) {
      *a.counter = 0;
      __threadfence_system();
      for (int j = 0; j < a.nranks; ++j) st_release_sys(&a.peer_consumed[j][a.rank], a.epoch);
    }
  }
}

}  // namespace

SMPK_API int smpk_sdp_gather(void* stream, void* const* peer_shard, void* full, long long piece_bytes,
                             long long pieces_per_rank, const void* ready, unsigned int epoch, int nranks, int rank,
                             int ctas) {
  SMPK_REQUIRE(nranks >= 1 && nranks <= MAX_RANKS && piece_bytes > 0 && piece_bytes % 16 == 0,
               "sdp_gather: bad arguments (nranks=%d piece_bytes=%lld)", nranks, piece_bytes);
  GatherArgs a;
  for (int i = 0; i < MAX_RANKS; ++i) a.peer_shard[i] = i < nranks ? peer_shard[i] : nullptr;
  a.full = full; a.piece_bytes = piece_bytes; a.pieces_per_rank = pieces_per_rank;
  a.ready = reinterpret_cast<const uint32_t*>(ready); a.epoch = epoch; a.nranks = nranks; a.rank = rank;
  if (ctas < 1) ctas = 96;
  const long long need = (pieces_per_rank * nranks + 7) / 8;
  if (ctas > need) ctas = (int)(need < 1 ? 1 : need);
  sdp_gather_kernel<<<ctas, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_sdp_reduce_adam(void* stream, void* const* peer_grad, float* grad_shard, float* master, float* exp_avg,
                                  float* exp_avg_sq, void* lowp, int lowp_dtype, long long piece_elems,
                                  long long pieces_per_rank, const void* ready, void* const* peer_consumed,
                                  void* counter, unsigned int epoch, int nranks, int rank, float scale, float beta,
                                  int mode, float lr, float beta1, float beta2, float eps, float wd, int step,
                                  int adamw, int bias_correction, float grad_scale, int ctas) {
  SMPK_REQUIRE(nranks >= 1 && nranks <= MAX_RANKS && piece_elems > 0 && piece_elems % 4 == 0,
               "sdp_reduce_adam: bad arguments (nranks=%d piece_elems=%lld)", nranks, piece_elems);
  SMPK_REQUIRE(mode == 0 || (master && exp_avg && exp_avg_sq), "sdp_reduce_adam: Adam mode needs master/exp_avg/exp_avg_sq");
  ReduceArgs a;
  for (int i = 0; i < MAX_RANKS; ++i) {
    a.peer_grad[i] = i < nranks ? reinterpret_cast<const float*>(peer_grad[i]) : nullptr;
    a.peer_consumed[i] = i < nranks ? reinterpret_cast<uint32_t*>(peer_consumed[i]) : nullptr;
  }
  a.grad_shard = grad_shard; a.master = master; a.exp_avg = exp_avg; a.exp_avg_sq = exp_avg_sq; a.lowp = lowp;
  a.lowp_dtype = lowp_dtype; a.piece_elems = piece_elems; a.pieces_per_rank = pieces_per_rank;
  a.ready = reinterpret_cast<const uint32_t*>(ready); a.counter = reinterpret_cast<uint32_t*>(counter);
  a.epoch = epoch; a.nranks = nranks; a.rank = rank; a.scale = scale; a.beta = beta; a.mode = mode;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = wd; a.adamw = adamw; a.grad_scale = grad_scale;
  a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  a.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
  const long long nvec = piece_elems * pieces_per_rank / 4;
  if (ctas < 1) ctas = 2 * smpk::num_sms();
  long long g = (nvec + 256 * RU - 1) / (256 * RU);
  if (g < 1) g = 1;
  if (ctas > g) ctas = (int)g;
  sdp_reduce_kernel<<<ctas, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
