// Persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   C[b] = epilogue( alpha * A[b] x B[b]^T )        A: [M,K]   B: [N,K]   C: [M,N] row-major
//
// Each operand may be stored K-major (reduction dim contiguous) or MN-major (M resp. N
// contiguous), so one kernel serves the three GEMMs of a linear layer without transposes:
//   forward  y  = x  W^T   (A K-major,  B K-major)
//   dgrad    dx = dy W     (A K-major,  B MN-major)
//   wgrad    dW = dy^T x   (A MN-major, B MN-major; fp32 accumulate into the grad buffer)
// and the two batched GEMMs of attention (QK^T, PV).
//
// Structure: warp 0 = TMA producer, warp 1 = MMA issuer (single thread, tcgen05.mma with the
// accumulator in TMEM), warp 2 = TMEM allocator, warps 4-7 = epilogue (tcgen05.ld -> registers ->
// fused bias / GELU / dGELU / residual / cast -> global).  The accumulator is double-buffered in
// TMEM (2 x BLOCK_N columns) so the epilogue of tile i overlaps the main loop of tile i+1.
//
// Replaces the reference's cuBLAS calls (reference smdistributed/modelparallel/torch/nn/
// transformer.py:1107-1134,1476-1482,1515,1657,1810-1812; nn/linear.py:62) and its TorchScript
// bias-GELU (nn/gelu.py:28-47), which becomes a GEMM epilogue here.
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

using namespace smp100;

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 x 2 B = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 256;
constexpr int EPI_WARP0 = 4;  // epilogue warps 4..7 (warp_idx % 4 selects the TMEM lane quarter)
constexpr int COPY_ROWS = 8;   // rows per all-gather copy piece
constexpr int GROUP_M = 8;    // L2-friendly raster: 8 m-blocks share one sweep over n

enum EpiMode : int { EPI_NONE = 0, EPI_GELU_TANH = 1, EPI_GELU_ERF = 2, EPI_DGELU_TANH = 3, EPI_DGELU_ERF = 4 };

// Communication fused into the GEMM (tensor-parallel groups of up to 8 GPUs on one NVSwitch domain).
//   COMM_AG: the A operand is the token all-gather of every rank's shard.  The first `comm_ctas` CTAs
//            of the grid copy the peers' shards out of peer memory (P2P loads over NVLink) into this
//            GPU's gathered buffer, 128-row block by block, publishing a flag per block; the GEMM CTAs
//            start with the local shard and consume remote blocks as their flags appear.
//   COMM_RS: the output is reduce-scattered over tokens.  The epilogue stores every tile straight into
//            the staging slot of the rank that owns those rows (P2P stores), remote destinations first,
//            and raises that rank's arrival flag once all tiles for it are written; a small reduce
//            kernel on the owner sums the slots.
//   COMM_AGB: the B operand (a weight) is the first tensor of a sharded-data-parallel unit whose flat buffer is
//            striped over the ranks in 8-row pieces (piece g lives on rank g % n at index g / n).  The copy
//            warps gather the WHOLE unit into `local_buf` (the weight first), counting pieces per BN-row
//            block; the TMA producer waits for the n-block it is about to load.  This is the ZeRO-3
//            parameter all-gather fused into the first GEMM that consumes the unit.
enum CommMode : int { COMM_NONE = 0, COMM_AG = 1, COMM_RS = 2, COMM_AGB = 3 };
constexpr int MAX_RANKS = 8;

struct CommParams {
  int mode, nranks, rank, rows_per_rank, comm_ctas;
  uint32_t epoch;
  // all-gather of A
  const void* peer_src[MAX_RANKS];     // peer j's gathered buffer (same layout as mine); I read its slot j
  void* local_buf;                     // my gathered buffer [nranks * rows_per_rank, K]
  long long row_bytes;                 // K * 2
  long long total_rows;                // COMM_AGB: rows of the whole unit (nranks * rows_per_rank)
  uint32_t* blk_flags;                 // local [M / 128]: block landed (== epoch)
  const uint32_t* ready_flags;         // local [nranks]: peer j's slot is final for this epoch (written by j)
  uint32_t* peer_done[MAX_RANKS];      // peer j's done[nranks]; I set [rank] = epoch after copying its slot
  uint32_t* counters;                  // local [nranks] scratch counters, zero between uses
  // reduce-scatter of C
  void* peer_stage[MAX_RANKS];         // rank d's staging [nranks][rows_per_rank][N]; I write slot `rank`
  uint32_t* peer_arrived[MAX_RANKS];   // rank d's arrived[nranks]; I set [rank] = epoch when my tiles landed
};

struct GemmParams {
  CommParams comm;
  int M, N, K, batch;
  int num_m, num_n, num_k;
  int a_mn_major, b_mn_major;
  // output
  void* C;
  int ldc;
  long long c_bs;
  int out_dtype;
  // optional pre-activation output (EPI_GELU_*)
  void* aux;
  // optional residual / accumulate input or dGELU pre-activation input
  const void* R;
  int ldr;
  long long r_bs;
  int r_dtype;
  // optional bias over N
  const void* bias;
  int bias_dtype;
  float alpha, beta;
  int epi;
  // causal structure (square attention problems, M == N or M == K): 0 none;
  //   1: only output tiles that touch the lower triangle (n0 < tile row end) are computed - QK^T, dO V^T;
  //   2: A is lower triangular [M, K]: the reduction stops at the tile's last row - P V, dS K;
  //   3: A^T is lower triangular (A given MN-major [K, M]): the reduction starts at the tile's first row - P^T dO, dS^T Q
  int tri;
  uint32_t idesc;
};

// Rows covered by the tile that m_blk belongs to (a CTA pair covers two 128-row blocks).
template <bool CTA2>
__device__ __forceinline__ void tile_rows(int m_blk, int& row0, int& row1) {
  const int first = CTA2 ? (m_blk & ~1) : m_blk;
  row0 = first * BLOCK_M;
  row1 = row0 + (CTA2 ? 2 : 1) * BLOCK_M;
}
template <bool CTA2>
__device__ __forceinline__ bool tile_skipped(const GemmParams& p, int m_blk, int n0) {
  if (p.tri != 1) return false;
  int r0, r1;
  tile_rows<CTA2>(m_blk, r0, r1);
  return n0 >= r1;
}
template <bool CTA2>
__device__ __forceinline__ void tile_k_range(const GemmParams& p, int m_blk, int& kb0, int& kb1) {
  kb0 = 0;
  kb1 = p.num_k;
  if (p.tri < 2) return;
  int r0, r1;
  tile_rows<CTA2>(m_blk, r0, r1);
  if (p.tri == 2) kb1 = min(p.num_k, (r1 + BLOCK_K - 1) / BLOCK_K);
  else kb0 = min(p.num_k - 1, r0 / BLOCK_K);
}

// CTA2: the tile is computed by a CTA pair (cta_group::2): 256 x BN output, each CTA holds its 128 rows of A and
// half of the B rows per stage, so a stage is 32 KB instead of 48 KB and six stages fit.
template <int BN, bool CTA2 = false>
struct Cfg {
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_ROWS = CTA2 ? BN / 2 : BN;
  static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = CTA2 ? 6 : ((BN == 256) ? 4 : (BN == 128 ? 6 : 8));
  static constexpr int TMEM_COLS = 2 * BN;  // double-buffered accumulator (power of two >= 32)
  static constexpr int EPI_STAGE_BYTES = 4 * 32 * 256;  // per epilogue warp: 32 rows x 256 B, for coalesced stores
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ void tile_coords(int t, const GemmParams& p, int num_m, int& b, int& m_blk, int& n_blk) {
  const int per_batch = num_m * p.num_n;
  b = t / per_batch;
  const int r = t - b * per_batch;
  const int group_size = GROUP_M * p.num_n;
  const int group_id = r / group_size;
  const int first_m = group_id * GROUP_M;
  const int gsm = min(num_m - first_m, GROUP_M);
  const int in_group = r - group_id * group_size;
  m_blk = first_m + in_group % gsm;
  n_blk = in_group / gsm;
}

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void spin_until_ge(const uint32_t* p, uint32_t v, bool sys) {
  while ((int32_t)((sys ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - v) < 0) __nanosleep(64);
}

// Order m-blocks so communication overlaps compute: AG starts with the local shard and walks the ring;
// RS finishes with the local shard (remote tiles leave first).
__device__ __forceinline__ int permute_m_block(int m_blk, const GemmParams& p, int blocks_per_tile = 1) {
  const CommParams& c = p.comm;
  if (c.mode != COMM_AG && c.mode != COMM_RS) return m_blk;
  const int mpr = c.rows_per_rank / (BLOCK_M * blocks_per_tile);   // tiles (single blocks or CTA-pair tiles) per rank
  const int ci = m_blk / mpr, within = m_blk - ci * mpr;
  const int chunk = (c.mode == COMM_AG) ? (c.rank + ci) % c.nranks : (c.rank + 1 + ci) % c.nranks;
  return chunk * mpr + within;
}

template <typename T>
__device__ __forceinline__ void load32(const T* __restrict__ ptr, float (&o)[32], bool vec, int nvalid) {
  if (vec) {
    constexpr int PER = 16 / sizeof(T);
    constexpr int NV = 32 / PER;
    const uint4* p4 = reinterpret_cast<const uint4*>(ptr);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      uint4 raw = p4[i];
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int j = 0; j < PER; ++j) o[i * PER + j] = smpk::to_f32<T>(e[j]);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = (i < nvalid) ? smpk::to_f32<T>(ptr[i]) : 0.f;
  }
}

template <typename T>
__device__ __forceinline__ void store32(T* __restrict__ ptr, const float (&v)[32], bool vec, int nvalid) {
  if (vec) {
    constexpr int PER = 16 / sizeof(T);
    constexpr int NV = 32 / PER;
    uint4* p4 = reinterpret_cast<uint4*>(ptr);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      uint4 raw;
      T* e = reinterpret_cast<T*>(&raw);
#pragma unroll
      for (int j = 0; j < PER; ++j) e[j] = smpk::from_f32<T>(v[i * PER + j]);
      p4[i] = raw;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < nvalid) ptr[i] = smpk::from_f32<T>(v[i]);
  }
}

__device__ __forceinline__ void load32_dyn(const void* base, int dtype, long long off, float (&o)[32], bool vec,
                                           int nvalid) {
  if (dtype == SMPK_F32)
    load32<float>(reinterpret_cast<const float*>(base) + off, o, vec, nvalid);
  else if (dtype == SMPK_BF16)
    load32<__nv_bfloat16>(reinterpret_cast<const __nv_bfloat16*>(base) + off, o, vec, nvalid);
  else
    load32<__half>(reinterpret_cast<const __half*>(base) + off, o, vec, nvalid);
}
__device__ __forceinline__ void store32_dyn(void* base, int dtype, long long off, const float (&v)[32], bool vec,
                                            int nvalid) {
  if (dtype == SMPK_F32)
    store32<float>(reinterpret_cast<float*>(base) + off, v, vec, nvalid);
  else if (dtype == SMPK_BF16)
    store32<__nv_bfloat16>(reinterpret_cast<__nv_bfloat16*>(base) + off, v, vec, nvalid);
  else
    store32<__half>(reinterpret_cast<__half*>(base) + off, v, vec, nvalid);
}

template <int BN, bool CTA2>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                        const GemmParams p) {
  using C = Cfg<BN, CTA2>;
  // CTA pair: rank 0 (leader) issues the MMAs for both; each CTA loads, and later stores, its own 128 rows
  const int cta_rank = CTA2 ? (int)cluster_ctarank() : 0;
  const bool leader = cta_rank == 0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_stage = smem + C::STAGES * C::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_stage + C::EPI_STAGE_BYTES);
  uint64_t* full_bar = bars;                   // [STAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + C::STAGES;      // [STAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * C::STAGES;  // [2]       MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + 2;        // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], CTA2 ? 8 : 4);  // one arrive per epilogue warp (of both CTAs of a pair)
    }
    fence_mbar_init();
    if (CTA2) fence_mbar_init_cluster();
  }
  if (warp == 2) {
    if (CTA2) {
      tmem_alloc_2sm<C::TMEM_COLS>(tmem_slot);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc<C::TMEM_COLS>(tmem_slot);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CTA2) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled across CTAs
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_m_eff = CTA2 ? (p.num_m + 1) / 2 : p.num_m;  // pair tiles span two 128-row blocks
  const int total_tiles = p.batch * num_m_eff * p.num_n;
  const int tile0 = CTA2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tstride = CTA2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (warp == 0) {
    // ===================================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = tile0; t < total_tiles; t += tstride) {
        int b, m_blk, n_blk;
        tile_coords(t, p, num_m_eff, b, m_blk, n_blk);
        m_blk = CTA2 ? permute_m_block(m_blk, p, 2) * 2 + cta_rank : permute_m_block(m_blk, p);
        const int m0 = m_blk * BLOCK_M, n0 = n_blk * BN;
        const int nb0 = n0 + cta_rank * C::B_ROWS;  // this CTA's share of the B rows (all of them without pairing)
        if (tile_skipped<CTA2>(p, m_blk, n0)) continue;
        int kb0, kb1;
        tile_k_range<CTA2>(p, m_blk, kb0, kb1);
        if (p.comm.mode == COMM_AG && m0 / p.comm.rows_per_rank != p.comm.rank) {
          // rows of a remote shard: wait until a copy CTA has landed this 128-row block, then make the
          // generic-proxy writes visible to the TMA (async proxy) reads that follow
          spin_until_ge(&p.comm.blk_flags[m_blk], (uint32_t)(BLOCK_M / COPY_ROWS), false);
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        if (p.comm.mode == COMM_AGB) {
          // weight rows of this n-block are being gathered by the copy warps (all CTAs); pieces are 8 rows
          const long long hi = min((long long)n0 + BN, p.comm.total_rows);
          spin_until_ge(&p.comm.blk_flags[n_blk], (uint32_t)((hi - n0) / COPY_ROWS), false);
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          uint8_t* sb = sa + C::A_BYTES;
          if (leader) mbar_expect_tx(&full_bar[stage], (CTA2 ? 2 : 1) * C::STAGE_BYTES);
          const int k0 = kb * BLOCK_K;
          if (!p.a_mn_major) {
            if (CTA2) tma_load_3d_2sm(sa, &tma_a, &full_bar[stage], k0, m0, b);
            else tma_load_3d(sa, &tma_a, &full_bar[stage], k0, m0, b);
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_M / 64; ++j) {
              if (CTA2) tma_load_3d_2sm(sa + j * (BLOCK_K * 128), &tma_a, &full_bar[stage], m0 + j * 64, k0, b);
              else tma_load_3d(sa + j * (BLOCK_K * 128), &tma_a, &full_bar[stage], m0 + j * 64, k0, b);
            }
          }
          if (!p.b_mn_major) {
            if (CTA2) tma_load_3d_2sm(sb, &tma_b, &full_bar[stage], k0, nb0, b);
            else tma_load_3d(sb, &tma_b, &full_bar[stage], k0, nb0, b);
          } else {
#pragma unroll
            for (int j = 0; j < C::B_ROWS / 64; ++j) {
              if (CTA2) tma_load_3d_2sm(sb + j * (BLOCK_K * 128), &tma_b, &full_bar[stage], nb0 + j * 64, k0, b);
              else tma_load_3d(sb + j * (BLOCK_K * 128), &tma_b, &full_bar[stage], nb0 + j * 64, k0, b);
            }
          }
          if (++stage == C::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer (pair: leader CTA only)
    if (lane == 0 && leader) {
      // K-major: 8-row groups 1024 B apart.  MN-major: K rows 128 B apart inside a 64-wide atom,
      // 8-row groups 1024 B apart, atoms BLOCK_K*128 B apart.
      const uint64_t hi_a = p.a_mn_major ? make_smem_desc_hi(BLOCK_K * 128, 1024) : make_smem_desc_hi(16, 1024);
      const uint64_t hi_b = p.b_mn_major ? make_smem_desc_hi(BLOCK_K * 128, 1024) : make_smem_desc_hi(16, 1024);
      // per-UMMA_K advance of the start-address field (16-byte units)
      const uint32_t adv_a = p.a_mn_major ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      const uint32_t adv_b = p.b_mn_major ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = tile0; t < total_tiles; t += tstride) {
        int kb0 = 0, kb1 = p.num_k;
        if (p.tri != 0) {
          int b, m_blk, n_blk;
          tile_coords(t, p, num_m_eff, b, m_blk, n_blk);
          m_blk = CTA2 ? m_blk * 2 : m_blk;
          if (tile_skipped<CTA2>(p, m_blk, n_blk * BN)) continue;
          tile_k_range<CTA2>(p, m_blk, kb0, kb1);
        }
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          uint8_t* sb = sa + C::A_BYTES;
          const uint64_t da = make_smem_desc(sa, hi_a);
          const uint64_t db = make_smem_desc(sb, hi_b);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            if (CTA2) umma_f16_2sm(tmem_d, da + uint64_t(k * adv_a), db + uint64_t(k * adv_b), p.idesc, (kb != kb0) || k != 0);
            else umma_f16(tmem_d, da + uint64_t(k * adv_a), db + uint64_t(k * adv_b), p.idesc, (kb != kb0) || k != 0);
          }
          // frees the smem slot (in both CTAs of a pair) once these MMAs retire
          if (CTA2) umma_commit_2sm(&empty_bar[stage]);
          else umma_commit(&empty_bar[stage]);
          if (++stage == C::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        // accumulator complete -> epilogue warps (of both CTAs)
        if (CTA2) umma_commit_2sm(&tfull_bar[acc]);
        else umma_commit(&tfull_bar[acc]);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp < EPI_WARP0) {
    // ===================================================================== all-gather copy warps (2, 3)
    // Every CTA's two spare warps pull 8-row pieces of the peers' shards over NVLink (16-byte peer loads,
    // 16 in flight per lane) into the local gathered buffer, ring order, and count each piece into its
    // 128-row block's arrival counter; the TMA producer of whichever CTA needs that block waits on it.
    if (p.comm.mode == COMM_AG) {
      const CommParams& c = p.comm;
      const long long piece_bytes = (long long)COPY_ROWS * c.row_bytes;
      const int pieces_per_rank = c.rows_per_rank / COPY_ROWS;
      const int total_pieces = (c.nranks - 1) * pieces_per_rank;
      const int wid = (int)blockIdx.x * 2 + (warp - 2), wstride = (int)gridDim.x * 2;
      int waited_src = -1;
      for (int pc = wid; pc < total_pieces; pc += wstride) {
        const int ci = pc / pieces_per_rank + 1, within = pc % pieces_per_rank;
        const int src = (c.rank + ci) % c.nranks;
        if (src != waited_src) {
          if (lane == 0) spin_until_ge(&c.ready_flags[src], c.epoch, true);
          __syncwarp();
          waited_src = src;
        }
        const long long off = ((long long)src * pieces_per_rank + within) * piece_bytes;
        const uint4* from = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(c.peer_src[src]) + off);
        uint4* to = reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.local_buf) + off);
        const int nvec = (int)(piece_bytes >> 4);
        constexpr int U = 16;
        for (int i = lane; i < nvec; i += U * 32) {
          uint4 v[U];
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (i + u * 32 < nvec) v[u] = from[i + u * 32];
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (i + u * 32 < nvec) to[i + u * 32] = v[u];
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicAdd(&c.blk_flags[(src * pieces_per_rank + within) / (BLOCK_M / COPY_ROWS)], 1u);
      }
      // when every copy warp of the grid is through, release the peers' shards for their next write
      __syncwarp();
      if (lane == 0) {
        const uint32_t old = atomicAdd(&c.counters[0], 1u);
        if (old == (uint32_t)wstride - 1u) {
          c.counters[0] = 0;
          __threadfence_system();
          for (int j = 0; j < c.nranks; ++j)
            if (j != c.rank) st_release_sys(&c.peer_done[j][c.rank], c.epoch);
        }
      }
    }
    if (p.comm.mode == COMM_AGB) {
      const CommParams& c = p.comm;
      const long long piece_bytes = (long long)COPY_ROWS * c.row_bytes;
      const long long total_pieces = c.total_rows / COPY_ROWS;
      const int wid = (int)blockIdx.x * 2 + (warp - 2), wstride = (int)gridDim.x * 2;
      uint32_t waited = 1u << c.rank;
      for (long long pc = wid; pc < total_pieces; pc += wstride) {
        const long long q = pc / c.nranks;
        const int src = (c.rank + (int)(pc % c.nranks)) % c.nranks;
        if (!((waited >> src) & 1u)) {
          if (lane == 0) spin_until_ge(&c.ready_flags[src], c.epoch, true);
          __syncwarp();
          waited |= 1u << src;
        }
        const long long g = q * c.nranks + src;
        const uint4* from = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(c.peer_src[src]) + q * piece_bytes);
        uint4* to = reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.local_buf) + g * piece_bytes);
        const int nvec = (int)(piece_bytes >> 4);
        constexpr int U = 16;
        for (int i = lane; i < nvec; i += U * 32) {
          uint4 v[U];
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (i + u * 32 < nvec) v[u] = from[i + u * 32];
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (i + u * 32 < nvec) to[i + u * 32] = v[u];
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicAdd(&c.blk_flags[(g * COPY_ROWS) / BN], 1u);
      }
    }
  } else {
    // ===================================================================== epilogue
    const int ew = warp - EPI_WARP0;  // == warp % 4: TMEM lanes [32*ew, 32*ew+32)
    int acc = 0;
    uint32_t acc_phase = 0;
    const int out_es = (p.out_dtype == SMPK_F32) ? 4 : 2;
    const bool c_vec_ok = ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0) && ((p.ldc * out_es) % 16 == 0) &&
                          ((p.c_bs * out_es) % 16 == 0) &&
                          (p.aux == nullptr || (reinterpret_cast<uintptr_t>(p.aux) & 15) == 0);
    const int r_es = (p.r_dtype == SMPK_F32) ? 4 : 2;
    const bool r_vec_ok = p.R == nullptr || (((reinterpret_cast<uintptr_t>(p.R) & 15) == 0) &&
                                             ((p.ldr * r_es) % 16 == 0) && ((p.r_bs * r_es) % 16 == 0));
    const int b_es = (p.bias_dtype == SMPK_F32) ? 4 : 2;
    const bool bias_vec_ok = p.bias == nullptr || ((reinterpret_cast<uintptr_t>(p.bias) & 15) == 0);
    (void)b_es;
    for (int t = tile0; t < total_tiles; t += tstride) {
      int b, m_blk, n_blk;
      tile_coords(t, p, num_m_eff, b, m_blk, n_blk);
      m_blk = CTA2 ? permute_m_block(m_blk, p, 2) * 2 + cta_rank : permute_m_block(m_blk, p);
      const int row = m_blk * BLOCK_M + ew * 32 + lane;
      const int n0 = n_blk * BN;
      if (tile_skipped<CTA2>(p, m_blk, n0)) continue;
      // reduce-scatter: this tile's rows belong to rank `dest`; write into its staging slot for this rank
      int dest = 0;
      char* c_base = reinterpret_cast<char*>(p.C);
      long long row_in_c = row;
      if (p.comm.mode == COMM_RS) {
        dest = (m_blk * BLOCK_M) / p.comm.rows_per_rank;
        c_base = reinterpret_cast<char*>(p.comm.peer_stage[dest]);
        row_in_c = (long long)p.comm.rank * p.comm.rows_per_rank + (row - (long long)dest * p.comm.rows_per_rank);
      }
      // whole tile in range and a 16-bit output: stage through shared memory for coalesced row stores
      // (always for reduce-scatter; for plain GEMMs only when the reduction is short - QK^T-like, store-bound
      // tiles - because the extra shared-memory pass costs ~3 % on compute-bound ones)
      const bool staged = (p.comm.mode == COMM_RS || (p.K <= 512 && p.aux == nullptr)) && out_es == 2 && c_vec_ok &&
                          (m_blk * BLOCK_M + BLOCK_M <= p.M) && (n0 + BN <= p.N);
      uint8_t* wstage = epi_stage + ew * (32 * 256);
      const long long row_in_c0 = row_in_c - lane;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(ew * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t raw[32];
        tmem_ld_32x32(taddr + c * 32, raw);
        tmem_ld_wait();
        const int col0 = n0 + c * 32;
        const int nvalid = min(32, p.N - col0);
        if (row < p.M && nvalid > 0) {
          const bool full = nvalid == 32;
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]) * p.alpha;
          if (p.bias != nullptr) {
            float bv[32];
            load32_dyn(p.bias, p.bias_dtype, col0, bv, full && bias_vec_ok, nvalid);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += bv[i];
          }
          const long long coff = (long long)b * p.c_bs + row_in_c * p.ldc + col0;
          if (p.R != nullptr) {
            float rv[32];
            const long long roff = (long long)b * p.r_bs + (long long)row * p.ldr + col0;
            load32_dyn(p.R, p.r_dtype, roff, rv, full && r_vec_ok, nvalid);
            if (p.epi == EPI_DGELU_TANH) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] *= smpk::gelu_tanh_grad(rv[i]);
            } else if (p.epi == EPI_DGELU_ERF) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] *= smpk::gelu_erf_grad(rv[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] += p.beta * rv[i];
            }
          }
          if (p.epi == EPI_GELU_TANH || p.epi == EPI_GELU_ERF) {
            if (p.aux != nullptr) store32_dyn(p.aux, p.out_dtype, coff, v, full && c_vec_ok, nvalid);
            if (p.epi == EPI_GELU_TANH) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = smpk::gelu_tanh(v[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = smpk::gelu_erf(v[i]);
            }
          }
          if (staged) {
            // 32 fp32 -> 32 x 16 bit = 4 x 16 B, written to this warp's staging rows with an XOR swizzle
            uint4 pk[4];
            if (p.out_dtype == SMPK_BF16) {
              __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
              for (int i = 0; i < 16; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            } else {
              __half2* h = reinterpret_cast<__half2*>(pk);
#pragma unroll
              for (int i = 0; i < 16; ++i) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int u = ((c & 3) * 4 + j) ^ (lane & 15);
              *reinterpret_cast<uint4*>(wstage + lane * 256 + u * 16) = pk[j];
            }
          } else {
            store32_dyn(c_base, p.out_dtype, coff, v, full && c_vec_ok, nvalid);
          }
        }
        if (staged && ((c & 3) == 3 || c == BN / 32 - 1)) {
          // flush up to 128 columns: each row's bytes leave as one contiguous run (coalesced / NVLink friendly)
          constexpr int PASS_COLS = BN < 128 ? BN : 128;
          constexpr int UPR = PASS_COLS * 2 / 16;            // 16-byte units per row
          constexpr int RPI = 32 / UPR;                      // rows per store instruction
          __syncwarp();
          const int pass_col0 = n0 + (c / 4) * 128;
          const int u = lane % UPR;
#pragma unroll
          for (int ps = 0; ps < 32 / RPI; ++ps) {
            const int rr = ps * RPI + lane / UPR;
            const uint4 val = *reinterpret_cast<const uint4*>(wstage + rr * 256 + ((u ^ (rr & 15)) * 16));
            const long long off = (long long)b * p.c_bs + (row_in_c0 + rr) * p.ldc + pass_col0;
            *reinterpret_cast<uint4*>(c_base + off * 2 + u * 16) = val;
          }
          __syncwarp();
        }
      }
      if (p.comm.mode == COMM_RS) {
        // publish: once every epilogue warp of every tile bound for `dest` has stored, raise its flag
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
          const int mpr = p.comm.rows_per_rank / BLOCK_M;
          const uint32_t expect = (uint32_t)(mpr * p.num_n * 4);
          const uint32_t old = atomicAdd(&p.comm.counters[dest], 1u);
          if (old == expect - 1u) {
            p.comm.counters[dest] = 0;
            __threadfence_system();
            st_release_sys(&p.comm.peer_arrived[dest][p.comm.rank], p.comm.epoch);
          }
        }
      }
      // all TMEM reads of this accumulator stage are done -> hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CTA2) mbar_arrive_leader(&tempty_bar[acc]);
        else mbar_arrive(&tempty_bar[acc]);
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CTA2) cluster_sync_all();  // the peer may still be reading this CTA's shared memory / TMEM until here
  if (warp == 2) {
    tc_fence_after();
    if (CTA2) tmem_dealloc_2sm<C::TMEM_COLS>(tmem_base);
    else tmem_dealloc<C::TMEM_COLS>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

// Operand tensor map.  K-major: dims {K, rows, batch}, box {64, box_rows, 1}.
// MN-major: dims {rows(MN), K, batch}, box {64, BLOCK_K, 1}.
int make_operand_map(CUtensorMap* map, int in_dtype, const void* ptr, int rows, int K, int batch, int ld,
                     long long bs, bool mn_major, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return smpk::set_error("cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[3];
  cuuint64_t strides[2];
  cuuint32_t box[3];
  cuuint32_t estr[3] = {1, 1, 1};
  if (!mn_major) {
    dims[0] = (cuuint64_t)K;
    dims[1] = (cuuint64_t)rows;
    box[0] = BLOCK_K;
    box[1] = (cuuint32_t)box_rows;
  } else {
    dims[0] = (cuuint64_t)rows;
    dims[1] = (cuuint64_t)K;
    box[0] = 64;
    box[1] = BLOCK_K;
  }
  dims[2] = (cuuint64_t)batch;
  box[2] = 1;
  strides[0] = (cuuint64_t)ld * 2;
  strides[1] = (cuuint64_t)(batch > 1 ? bs : (long long)dims[1] * ld) * 2;
  if (strides[1] == 0) strides[1] = strides[0];
  CUtensorMapDataType dt = in_dtype == SMPK_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUresult r = enc(map, dt, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return smpk::set_error("cuTensorMapEncodeTiled failed (%d): rows=%d K=%d batch=%d ld=%d bs=%lld mn_major=%d", (int)r,
                           rows, K, batch, ld, bs, (int)mn_major);
  return 0;
}

template <int BN, bool CTA2>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, GemmParams& p, cudaStream_t stream, int max_ctas) {
  using C = Cfg<BN, CTA2>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    SMPK_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, CTA2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  p.num_n = (p.N + BN - 1) / BN;
  p.idesc = make_idesc_f16(/*fmt filled by caller in idesc low bits*/ p.idesc, p.a_mn_major, p.b_mn_major,
                           CTA2 ? 2 * BLOCK_M : BLOCK_M, BN);
  const int num_m_eff = CTA2 ? (p.num_m + 1) / 2 : p.num_m;
  const int total = p.batch * num_m_eff * p.num_n;
  int grid = smpk::num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (CTA2) grid &= ~1;                          // whole pairs
  const int per_tile = CTA2 ? 2 : 1;
  if (p.comm.mode != COMM_AGB && per_tile * total < grid) grid = per_tile * total;
  // (gathered-weight mode: the copy warps of EVERY resident CTA gather the unit, so the grid stays full)
  if (p.comm.mode == COMM_AG) {
    // arrival counters of the row blocks start at zero for this launch
    SMPK_CHECK_CUDA(cudaMemsetAsync(p.comm.blk_flags, 0, sizeof(uint32_t) * (size_t)p.num_m, stream));
  }
  if (p.comm.mode == COMM_AGB) {
    const size_t nblk = (size_t)((p.comm.total_rows + BN - 1) / BN);
    SMPK_CHECK_CUDA(cudaMemsetAsync(p.comm.blk_flags, 0, sizeof(uint32_t) * nblk, stream));
  }
  if (CTA2) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    SMPK_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<BN, CTA2>, ta, tb, p));
    return 0;
  }
  gemm_tcgen05_kernel<BN, CTA2><<<grid, NUM_THREADS, C::SMEM_BYTES, stream>>>(ta, tb, p);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Pairing (cta_group::2): 0 = never, 1 = when the problem is large enough (default), 2 = whenever it applies
// (tests).  SMPK_GEMM_2CTA sets the initial value; smpk_gemm_set_pair_mode changes it at run time.
int g_pair_mode = -1;
int pair_mode() {
  if (g_pair_mode < 0) {
    const char* e = getenv("SMPK_GEMM_2CTA");
    g_pair_mode = (e == nullptr) ? 1 : (e[0] == '0' ? 0 : (e[0] == '2' ? 2 : 1));
  }
  return g_pair_mode;
}

}  // namespace

// Host-visible mirror of CommParams (include/smpb200_kernels.h: SmpkCommDesc).
struct SmpkCommDesc {
  int mode, nranks, rank, rows_per_rank, comm_ctas;
  unsigned int epoch;
  const void* peer_src[MAX_RANKS];
  void* local_buf;
  unsigned int* blk_flags;
  const unsigned int* ready_flags;
  unsigned int* peer_done[MAX_RANKS];
  unsigned int* counters;
  void* peer_stage[MAX_RANKS];
  unsigned int* peer_arrived[MAX_RANKS];
};

static int gemm_impl(void* stream, int in_dtype, int out_dtype, int M, int N, int K, int batch, const void* A,
                     int lda, long long a_bs, int a_mn_major, const void* B, int ldb, long long b_bs,
                     int b_mn_major, void* C, int ldc, long long c_bs, const void* bias, int bias_dtype,
                     const void* R, int ldr, long long r_bs, int r_dtype, void* aux, float alpha, float beta,
                     int epi, int max_ctas, const SmpkCommDesc* comm);

SMPK_API int smpk_gemm_set_pair_mode(int mode) {
  const int old = pair_mode();
  if (mode >= 0 && mode <= 2) g_pair_mode = mode;
  return old;
}

// See include/smpb200_kernels.h for the contract.
SMPK_API int smpk_gemm(void* stream, int in_dtype, int out_dtype, int M, int N, int K, int batch, const void* A,
                       int lda, long long a_bs, int a_mn_major, const void* B, int ldb, long long b_bs,
                       int b_mn_major, void* C, int ldc, long long c_bs, const void* bias, int bias_dtype,
                       const void* R, int ldr, long long r_bs, int r_dtype, void* aux, float alpha, float beta,
                       int epi, int max_ctas) {
  return gemm_impl(stream, in_dtype, out_dtype, M, N, K, batch, A, lda, a_bs, a_mn_major, B, ldb, b_bs, b_mn_major, C,
                   ldc, c_bs, bias, bias_dtype, R, ldr, r_bs, r_dtype, aux, alpha, beta, epi, max_ctas, nullptr);
}

// GEMM fused with a tensor-parallel collective over NVLink peer memory (see CommMode above).
SMPK_API int smpk_gemm_comm(void* stream, int in_dtype, int out_dtype, int M, int N, int K, const void* A, int lda,
                            int a_mn_major, const void* B, int ldb, int b_mn_major, void* C, int ldc,
                            const void* bias, int bias_dtype, const void* R, int ldr, int r_dtype, float beta,
                            void* aux, float alpha, int epi, int max_ctas, const SmpkCommDesc* comm) {
  SMPK_REQUIRE(comm != nullptr && (comm->mode == COMM_AG || comm->mode == COMM_RS || comm->mode == COMM_AGB),
               "gemm_comm: bad comm descriptor");
  SMPK_REQUIRE(comm->nranks >= 2 && comm->nranks <= MAX_RANKS, "gemm_comm: nranks %d out of range", comm->nranks);
  if (comm->mode == COMM_AGB) {
    SMPK_REQUIRE(!b_mn_major && ldb == K && B == comm->local_buf,
                 "gemm_comm: gathered-weight mode needs a dense K-major B at the start of the unit buffer");
    SMPK_REQUIRE(comm->rows_per_rank % COPY_ROWS == 0 && (long long)comm->rows_per_rank * comm->nranks >= N,
                 "gemm_comm: unit of %d x %d rows does not cover the %d weight rows", comm->nranks,
                 comm->rows_per_rank, N);
  } else {
    SMPK_REQUIRE(comm->rows_per_rank % BLOCK_M == 0 && comm->rows_per_rank * comm->nranks == M,
                 "gemm_comm: M=%d must be nranks x rows_per_rank with rows_per_rank %% 128 == 0 (got %d x %d)", M,
                 comm->nranks, comm->rows_per_rank);
  }
  SMPK_REQUIRE(comm->mode != COMM_RS || aux == nullptr, "gemm_comm: aux output is not supported with reduce-scatter");
  SMPK_REQUIRE(comm->mode != COMM_AG || !a_mn_major, "gemm_comm: all-gather mode needs a K-major A operand");
  return gemm_impl(stream, in_dtype, out_dtype, M, N, K, 1, A, lda, 0, a_mn_major, B, ldb, 0, b_mn_major, C, ldc, 0, bias,
                   bias_dtype, R, ldr, 0, r_dtype, aux, alpha, beta, epi, max_ctas, comm);
}

static int gemm_impl(void* stream, int in_dtype, int out_dtype, int M, int N, int K, int batch, const void* A,
                     int lda, long long a_bs, int a_mn_major, const void* B, int ldb, long long b_bs,
                     int b_mn_major, void* C, int ldc, long long c_bs, const void* bias, int bias_dtype,
                     const void* R, int ldr, long long r_bs, int r_dtype, void* aux, float alpha, float beta,
                     int epi, int max_ctas, const SmpkCommDesc* comm) {
  SMPK_REQUIRE(in_dtype == SMPK_BF16 || in_dtype == SMPK_F16, "smpk_gemm: inputs must be bf16 or fp16");
  SMPK_REQUIRE(M > 0 && N > 0 && K > 0 && batch > 0, "smpk_gemm: empty problem M=%d N=%d K=%d batch=%d", M, N, K, batch);
  SMPK_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && (a_bs % 8) == 0 && (b_bs % 8) == 0,
               "smpk_gemm: TMA needs 16-byte aligned strides (lda=%d ldb=%d a_bs=%lld b_bs=%lld)", lda, ldb, a_bs, b_bs);
  SMPK_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "smpk_gemm: operand pointers must be 16-byte aligned");
  const int tri = (epi >> 8) & 3;       // causal structure rides in bits 8-9 of the epilogue argument
  epi &= 0xff;
  SMPK_REQUIRE(epi >= 0 && epi <= 4, "smpk_gemm: bad epilogue mode %d", epi);
  SMPK_REQUIRE(tri == 0 || (comm == nullptr && (tri == 1 ? M == N : M == K) && (tri != 3 || a_mn_major)),
               "smpk_gemm: causal structure %d needs a square problem (M=%d N=%d K=%d) and no fused collective", tri, M,
               N, K);
  SMPK_REQUIRE(!(epi >= 3) || R != nullptr, "smpk_gemm: dGELU epilogue needs the pre-activation in R");

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K; p.batch = batch;
  p.num_m = (M + BLOCK_M - 1) / BLOCK_M;
  p.num_k = (K + BLOCK_K - 1) / BLOCK_K;
  p.a_mn_major = a_mn_major ? 1 : 0;
  p.b_mn_major = b_mn_major ? 1 : 0;
  p.C = C; p.ldc = ldc; p.c_bs = c_bs; p.out_dtype = out_dtype;
  p.aux = aux;
  p.R = R; p.ldr = ldr; p.r_bs = r_bs; p.r_dtype = r_dtype;
  p.bias = bias; p.bias_dtype = bias_dtype;
  p.alpha = alpha; p.beta = beta; p.epi = epi; p.tri = tri;
  p.idesc = (in_dtype == SMPK_BF16) ? 1u : 0u;  // format code, expanded in launch<>
  if (comm != nullptr) {
    CommParams& c = p.comm;
    c.mode = comm->mode; c.nranks = comm->nranks; c.rank = comm->rank; c.rows_per_rank = comm->rows_per_rank;
    c.comm_ctas = comm->comm_ctas; c.epoch = comm->epoch;
    c.local_buf = comm->local_buf; c.row_bytes = (long long)K * 2;
    c.total_rows = (long long)comm->rows_per_rank * comm->nranks;
    c.blk_flags = comm->blk_flags; c.ready_flags = comm->ready_flags; c.counters = comm->counters;
    for (int i = 0; i < MAX_RANKS; ++i) {
      c.peer_src[i] = comm->peer_src[i]; c.peer_done[i] = comm->peer_done[i];
      c.peer_stage[i] = comm->peer_stage[i]; c.peer_arrived[i] = comm->peer_arrived[i];
    }
    if (c.mode == COMM_RS) p.C = c.peer_stage[c.rank];
  }

  const int BN = (N > 128) ? 256 : (N > 64 ? 128 : 64);
  // CTA pairs for plain GEMMs with enough 256 x 256 tiles to occupy every pair; fused-collective modes and small
  // problems stay on single CTAs
  const long long pair_tiles = (long long)batch * ((p.num_m + 1) / 2) * ((N + 255) / 256);
  const int pm = pair_mode();
  // (token all-gather / reduce-scatter modes order tiles by owning rank, so a pair tile must not straddle two ranks)
  const bool comm_ok = comm == nullptr || comm->mode == COMM_AGB || comm->rows_per_rank % (2 * BLOCK_M) == 0;
  const bool pair = BN == 256 && comm_ok && M > BLOCK_M && pm > 0 && (pm == 2 || pair_tiles >= smpk::num_sms() / 4);
  CUtensorMap ta, tb;
  int rc = make_operand_map(&ta, in_dtype, A, M, K, batch, lda, a_bs, p.a_mn_major, BLOCK_M);
  if (rc) return rc;
  rc = make_operand_map(&tb, in_dtype, B, N, K, batch, ldb, b_bs, p.b_mn_major, pair ? BN / 2 : BN);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (pair) return launch<256, true>(ta, tb, p, s, max_ctas);
  if (BN == 256) return launch<256, false>(ta, tb, p, s, max_ctas);
  if (BN == 128) return launch<128, false>(ta, tb, p, s, max_ctas);
  return launch<64, false>(ta, tb, p, s, max_ctas);
}
