// Flag signalling, peer all-gather and the reduce half of GEMM->reduce-scatter for tensor-parallel
// groups on NVLink peer memory.  Companion of the COMM_AG / COMM_RS modes of gemm_tcgen05.cu.
//
// The reference performs these collectives with NCCL through torch.distributed on the compute stream
// (reference smdistributed/modelparallel/torch/nn/utils.py:701-821), never overlapped with the GEMMs
// next to them; here they are device code reading / writing the peers' memory
// directly, sequenced by epoch-valued flags in peer memory.
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

struct PtrTable {
  uint32_t* p[MAX_RANKS];
};

// flags[index] = value on every listed peer (after everything earlier on this stream is visible)
__global__ void flags_signal_kernel(PtrTable peers, int n, int index, uint32_t value) {
  const int j = threadIdx.x;
  if (j < n && peers.p[j] != nullptr) {
    __threadfence_system();
    st_release_sys(peers.p[j] + index, value);
  }
}

// wait until flags[i] >= value for every i in [0, n) except `skip`
__global__ void flags_wait_kernel(const uint32_t* flags, int n, int skip, uint32_t value) {
  const int j = threadIdx.x;
  if (j < n && j != skip) spin_ge(flags + j, value);
}

struct GatherArgs {
  const void* peer_src[MAX_RANKS];
  void* local_buf;
  long long slot_bytes;
  const uint32_t* ready_flags;
  uint32_t* peer_done[MAX_RANKS];
  uint32_t* counters;
  int nranks, rank;
  uint32_t epoch;
};

// Stand-alone all-gather by peer loads (same copy loop as the copy CTAs of the fused GEMM).
__global__ void __launch_bounds__(512) p2p_allgather_kernel(GatherArgs a) {
  for (int ci = 1; ci < a.nranks; ++ci) {
    const int src = (a.rank + ci) % a.nranks;
    if (threadIdx.x == 0) spin_ge(&a.ready_flags[src], a.epoch);
    __syncthreads();
    const long long off = (long long)src * a.slot_bytes;
    const uint4* from = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(a.peer_src[src]) + off);
    uint4* to = reinterpret_cast<uint4*>(reinterpret_cast<char*>(a.local_buf) + off);
    const long long nvec = a.slot_bytes >> 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += 4 * stride) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i + u * stride < nvec) v[u] = from[i + u * stride];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i + u * stride < nvec) to[i + u * stride] = v[u];
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t old = atomicAdd(&a.counters[src], 1u);
      if (old == gridDim.x - 1) {
        a.counters[src] = 0;
        __threadfence_system();
        st_release_sys(&a.peer_done[src][a.rank], a.epoch);
      }
    }
  }
}

struct ReduceArgs {
  const void* staging;          // [nranks][rows][N] in `dtype`
  const uint32_t* arrived;      // [nranks] flags raised by the producers
  uint32_t* peer_reduced[MAX_RANKS];   // each producer's reduced[nranks]; I set [rank] when done reading
  uint32_t* counter;
  const void* residual;         // optional [rows][N]
  const void* bias;             // optional [N]
  void* out;                    // [rows][N]
  long long rows;
  int N, nranks, rank, dtype, out_dtype, res_dtype, bias_dtype;
  uint32_t epoch;
};

// out = sum_src staging[src] (+ bias) (+ residual), summed in rank order (deterministic).
__global__ void __launch_bounds__(256) rs_reduce_kernel(ReduceArgs a) {
  if (threadIdx.x < a.nranks) spin_ge(&a.arrived[threadIdx.x], a.epoch);
  __syncthreads();
  const long long total = a.rows * a.N;
  const long long nvec = total >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float acc[8], v[8];
    load8(a.staging, a.dtype, (size_t)i * 8, acc);
    for (int s = 1; s < a.nranks; ++s) {
      load8(a.staging, a.dtype, (size_t)s * total + (size_t)i * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    if (a.bias != nullptr) {
      load8(a.bias, a.bias_dtype, (size_t)((i * 8) % a.N), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    if (a.residual != nullptr) {
      load8(a.residual, a.res_dtype, (size_t)i * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    store8(a.out, a.out_dtype, (size_t)i * 8, acc);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t old = atomicAdd(a.counter, 1u);
    if (old == gridDim.x - 1) {
      *a.counter = 0;
      __threadfence_system();
      for (int s = 0; s < a.nranks; ++s)
        if (a.peer_reduced[s] != nullptr) st_release_sys(&a.peer_reduced[s][a.rank], a.epoch);
    }
  }
}

}  // namespace

SMPK_API int smpk_flags_signal(void* stream, void* const* peer_flags, int n, int index, unsigned int value) {
  SMPK_REQUIRE(n >= 1 && n <= MAX_RANKS, "flags_signal: n=%d out of range", n);
  PtrTable t;
  for (int i = 0; i < MAX_RANKS; ++i) t.p[i] = i < n ? reinterpret_cast<uint32_t*>(peer_flags[i]) : nullptr;
  flags_signal_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(t, n, index, value);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_flags_wait(void* stream, const void* flags, int n, int skip, unsigned int value) {
  SMPK_REQUIRE(n >= 1 && n <= 32, "flags_wait: n=%d out of range", n);
  flags_wait_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(reinterpret_cast<const uint32_t*>(flags), n, skip,
                                                                          value);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_p2p_allgather(void* stream, void* const* peer_src, void* local_buf, long long slot_bytes,
                                const void* ready_flags, void* const* peer_done, void* counters, int nranks, int rank,
                                unsigned int epoch, int ctas) {
  SMPK_REQUIRE(nranks >= 2 && nranks <= MAX_RANKS && slot_bytes % 16 == 0, "p2p_allgather: bad arguments");
  GatherArgs a;
  for (int i = 0; i < MAX_RANKS; ++i) {
    a.peer_src[i] = i < nranks ? peer_src[i] : nullptr;
    a.peer_done[i] = i < nranks ? reinterpret_cast<uint32_t*>(peer_done[i]) : nullptr;
  }
  a.local_buf = local_buf; a.slot_bytes = slot_bytes;
  a.ready_flags = reinterpret_cast<const uint32_t*>(ready_flags);
  a.counters = reinterpret_cast<uint32_t*>(counters);
  a.nranks = nranks; a.rank = rank; a.epoch = epoch;
  if (ctas < 1) ctas = 32;
  p2p_allgather_kernel<<<ctas, 512, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

SMPK_API int smpk_rs_reduce(void* stream, const void* staging, const void* arrived, void* const* peer_reduced,
                            void* counter, const void* residual, int res_dtype, const void* bias, int bias_dtype,
                            void* out, int out_dtype, long long rows, int N, int nranks, int rank, int dtype,
                            unsigned int epoch) {
  SMPK_REQUIRE(nranks >= 2 && nranks <= MAX_RANKS && N % 8 == 0, "rs_reduce: bad arguments (N %% 8 != 0?)");
  ReduceArgs a;
  a.staging = staging; a.arrived = reinterpret_cast<const uint32_t*>(arrived);
  for (int i = 0; i < MAX_RANKS; ++i) a.peer_reduced[i] = i < nranks ? reinterpret_cast<uint32_t*>(peer_reduced[i]) : nullptr;
  a.counter = reinterpret_cast<uint32_t*>(counter);
  a.residual = residual; a.bias = bias; a.out = out; a.rows = rows; a.N = N; a.nranks = nranks; a.rank = rank;
  a.dtype = dtype; a.out_dtype = out_dtype; a.res_dtype = res_dtype; a.bias_dtype = bias_dtype; a.epoch = epoch;
  const long long nvec = rows * N / 8;
  long long g = (nvec + 255) / 256;
  const long long cap = (long long)smpk::num_sms() * 8;
  const int grid = (int)(g < 1 ? 1 : (g > cap ? cap : g));
  rs_reduce_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
