// Flash-style attention for sm_100a: softmax(scale q k^T + mask) v without ever writing the [Sq, Sk] score or
// probability matrices to global memory, forward and backward.
//
// Forward (attn_fwd_kernel), one CTA per 128-query tile of one (batch, head):
//   warp 0      TMA producer for Q (once) and the K tiles        (cp.async.bulk.tensor, 128-byte swizzle)
//   warp 3      TMA producer for the V tiles
//   warp 1      MMA issuer: S_j = Q K_j^T (SS mode, accumulator in TMEM, double buffered), O += P_j V_j with P_j read
//               from TENSOR MEMORY (TS mode) - the probabilities never pass through shared or global memory
//   warp 2      TMEM allocator
//   warps 4-7   online softmax: tcgen05.ld a row of S per thread, running max / sum in the log2 domain, lazy
//               rescale of the O accumulator (only when the row maximum grew by more than 2^8), Philox dropout,
//               byte / causal / local-window masking, P packed to 16 bit and written back over S with tcgen05.st;
//               finally O / l -> global and the log-sum-exp row statistics for the backward pass.
//
// Backward (attn_bwd_kernel) keeps one 128-row tile of two operands resident and streams 64-row tiles of the other
// side; both score-shaped products are recomputed on the tensor cores into TMEM, the elementwise stage turns them
// into P / dS (16 bit, TMEM) and the output products consume those straight from TMEM:
//   key-stationary   (transposed):  S^T = K Q_i^T, dP^T = V dO_i^T  ->  dV += P^T dO_i,  dK += dS^T Q_i
//   query-stationary              :  S = Q K_j^T,  dP = dO V_j^T    ->  dQ += dS K_j
// For head_dim 256 the two key-stationary outputs do not fit in tensor memory together (2 x 256 accumulator columns
// plus the score tiles), so they run as two launches.
//
// Replaces the reference's baddbmm -> fused softmax -> dropout -> matmul composition and its autograd backward
// (reference smdistributed/modelparallel/torch/nn/transformer.py:1617-1657,1769-1835) and the Megatron-style
// softmax kernels on that path (torch/nn/softmax.py:40-93), including their 2048-key limit (transformer.py:101).
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

using namespace smp100;

namespace {

constexpr int BM = 128;          // rows of the resident tile (= TMEM lanes)
constexpr int NUM_THREADS = 256;
constexpr int SM_WARP0 = 4;      // softmax / elementwise warps 4..7 (warp % 4 = TMEM lane quarter)
constexpr float RESCALE_THRESHOLD = 8.0f;   // log2 units: the stale maximum may lag by up to 2^8

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
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

// Operand [B, H, S, d] with element strides (sb, sh, ss, 1): tensor map dims {d, S, H, B}, box {64, rows, 1, 1}.
// Columns beyond d and rows beyond S are zero-filled by the TMA unit.
int make_map4(CUtensorMap* map, int dtype, const void* ptr, int B, int H, int S, int d, long long sb, long long sh,
              long long ss, int box_rows) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return smpk::set_error("cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)S, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)ss * 2, (cuuint64_t)sh * 2, (cuuint64_t)sb * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  if (H == 1) strides[1] = strides[0] * (cuuint64_t)S;   // unused dimension still needs a legal stride
  if (B == 1) strides[2] = strides[1] * (cuuint64_t)H;
  for (int i = 0; i < 3; ++i)
    if (strides[i] == 0 || (strides[i] & 15)) return smpk::set_error("attention: operand stride %d (%llu B) is not a non-zero multiple of 16 bytes", i, (unsigned long long)strides[i]);
  CUtensorMapDataType dt = dtype == SMPK_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUresult r = enc(map, dt, 4, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return smpk::set_error("attention: cuTensorMapEncodeTiled failed (%d): B=%d H=%d S=%d d=%d strides %lld %lld %lld", (int)r,
                           B, H, S, d, sb, sh, ss);
  return 0;
}

// ------------------------------------------------------------------------------------------------ Philox4x32-10
// Counter-based generator: the dropout mask of element (bh, q, k) is a pure function of (seed, offset, bh, q, k), so
// the backward kernels and an activation-checkpoint recomputation regenerate exactly the mask of the forward pass.
__device__ __forceinline__ uint4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
struct DropoutParams {
  uint32_t seed_lo, seed_hi, off_lo, off_hi;
  uint32_t threshold;     // drop when the random word < threshold (= p * 2^32)
  float inv_keep;         // 1 / (1 - p)
  int enabled;
};
// random words for keys 4*kg .. 4*kg+3 of query row q in (batch*head) bh
__device__ __forceinline__ uint4 dropout_words(const DropoutParams& dp, int bh, int q, int kg) {
  return philox4x32((uint32_t)kg, (uint32_t)q, (uint32_t)bh ^ dp.off_lo, dp.off_hi, dp.seed_lo, dp.seed_hi);
}

struct MaskParams {
  const uint8_t* mask;          // non-zero = excluded; [B, (H), Sq|1, Sk] with element strides, last stride 1
  long long sb, sh, sq;
  int causal;                   // key k visible to query q iff k <= q + (Sk - Sq)
  int window;                   // > 0: additionally k > q + (Sk - Sq) - window (local attention)
};

struct FwdParams {
  int B, H, Sq, Sk, d;
  float scale_log2;             // softmax scale * log2(e)
  MaskParams m;
  DropoutParams drop;
  void* O;
  long long o_sb, o_sh, o_ss;
  int out_dtype;
  float* lse;                   // [B*H, Sq], log2 domain: m + log2(l); +inf for rows with no visible key
  uint32_t idesc_s, idesc_o;
};

template <int NCH, int BN, int KST, int VST>
struct FwdCfg {
  static constexpr int Q_BYTES = NCH * BM * 128;
  static constexpr int KV_BYTES = NCH * BN * 128;
  static constexpr int SMEM_BYTES = Q_BYTES + (KST + VST) * KV_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

__device__ __forceinline__ bool key_visible(const MaskParams& m, int q, int k, int Sq, int Sk) {
  if (k >= Sk) return false;
  const int off = Sk - Sq;
  if (m.causal && k > q + off) return false;
  if (m.window > 0 && k <= q + off - m.window) return false;
  return true;
}

// Tile range [j0, j1) of BN-key tiles that hold a visible key for query rows [q0, q0 + rows).
__device__ __forceinline__ void kv_tile_range(const MaskParams& m, int q0, int rows, int Sq, int Sk, int BN, int& j0, int& j1) {
  const int off = Sk - Sq;
  int hi = Sk - 1;                                   // last visible key of the tile's last row
  if (m.causal) hi = min(hi, min(q0 + rows, Sq) - 1 + off);
  int lo = 0;                                        // first visible key of the tile's first row
  if (m.window > 0) lo = max(0, q0 + off - m.window + 1);
  j0 = lo / BN;
  j1 = hi < 0 ? 0 : hi / BN + 1;
  if (j1 < j0) j1 = j0;
}

template <int NCH, int BN, int KST, int VST>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    attn_fwd_kernel(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
                    const __grid_constant__ CUtensorMap tma_v, const FwdParams p) {
  using C = FwdCfg<NCH, BN, KST, VST>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + C::Q_BYTES;
  uint8_t* sV = sK + KST * C::KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + VST * C::KV_BYTES);
  uint64_t* q_full = bars;                 // [1]
  uint64_t* k_full = bars + 1;             // [KST]
  uint64_t* k_empty = k_full + KST;        // [KST]
  uint64_t* v_full = k_empty + KST;        // [VST]
  uint64_t* v_empty = v_full + VST;        // [VST]
  uint64_t* s_full = v_empty + VST;        // [2]  MMA -> softmax: S_j complete
  uint64_t* p_full = s_full + 2;           // [2]  softmax -> MMA: P_j written (4 warp arrivals)
  uint64_t* o_bar = p_full + 2;            // [1]  MMA -> softmax: O += P_j V_j retired
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int qt = (int)gridDim.x - 1 - (int)blockIdx.x;   // heaviest (causal) tiles first
  const int q0 = qt * BM;
  int j0, j1;
  kv_tile_range(p.m, q0, BM, p.Sq, p.Sk, BN, j0, j1);
  const int ntiles = j1 - j0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_q);
    tma_prefetch_desc(&tma_k);
    tma_prefetch_desc(&tma_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < KST; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < VST; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4); }
    mbar_init(o_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<512>(tmem_slot);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_o = tmem_base + 2 * BN;      // O accumulator: columns [2*BN, 2*BN + d)

  if (warp == 0) {
    // ================================================================= Q (once) and K tiles
    if (lane == 0 && ntiles > 0) {
      mbar_expect_tx(q_full, C::Q_BYTES);
#pragma unroll
      for (int c = 0; c < NCH; ++c) tma_load_4d(sQ + c * (BM * 128), &tma_q, q_full, c * 64, q0, h, b);
      int st = 0;
      uint32_t ph = 0;
      for (int j = j0; j < j1; ++j) {
        mbar_wait(&k_empty[st], ph ^ 1);
        mbar_expect_tx(&k_full[st], C::KV_BYTES);
#pragma unroll
        for (int c = 0; c < NCH; ++c)
          tma_load_4d(sK + st * C::KV_BYTES + c * (BN * 128), &tma_k, &k_full[st], c * 64, j * BN, h, b);
        if (++st == KST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 3) {
    // ================================================================= V tiles
    if (lane == 0) {
      int st = 0;
      uint32_t ph = 0;
      for (int j = j0; j < j1; ++j) {
        mbar_wait(&v_empty[st], ph ^ 1);
        mbar_expect_tx(&v_full[st], C::KV_BYTES);
#pragma unroll
        for (int c = 0; c < NCH; ++c)
          tma_load_4d(sV + st * C::KV_BYTES + c * (BN * 128), &tma_v, &v_full[st], c * 64, j * BN, h, b);
        if (++st == VST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0 && ntiles > 0) {
      const uint64_t hi_k = make_smem_desc_hi(16, 1024);              // K-major operands (Q, K)
      const uint64_t hi_v = make_smem_desc_hi(BN * 128, 1024);        // V as the N-major B operand of P V
      const int ksteps_s = p.d / 16;                                  // reduction over head_dim
      int kst = 0, vst = 0;
      uint32_t kph = 0, vph = 0;
      auto issue_s = [&](int t) {   // S_t = Q K_t^T into score buffer t & 1
        mbar_wait(&k_full[kst], kph);
        tc_fence_after();
        const uint32_t tmem_s = tmem_base + (t & 1) * BN;
        for (int ks = 0; ks < ksteps_s; ++ks) {
          const int c = ks >> 2, within = (ks & 3) * 32;
          const uint64_t da = make_smem_desc(sQ + c * (BM * 128) + within, hi_k);
          const uint64_t db = make_smem_desc(sK + kst * C::KV_BYTES + c * (BN * 128) + within, hi_k);
          umma_f16(tmem_s, da, db, p.idesc_s, ks != 0);
        }
        umma_commit(&k_empty[kst]);
        umma_commit(&s_full[t & 1]);
        if (++kst == KST) { kst = 0; kph ^= 1; }
      };
      mbar_wait(q_full, 0);
      issue_s(0);
      for (int t = 0; t < ntiles; ++t) {
        if (t + 1 < ntiles) issue_s(t + 1);
        mbar_wait(&p_full[t & 1], (t >> 1) & 1);
        mbar_wait(&v_full[vst], vph);
        tc_fence_after();
        const uint32_t tmem_p = tmem_base + (t & 1) * BN;             // P_t packed over the first BN/2 columns of S_t
#pragma unroll
        for (int ks = 0; ks < BN / 16; ++ks) {
          const uint64_t db = make_smem_desc(sV + vst * C::KV_BYTES + ks * (16 * 128), hi_v);
          umma_f16_ts(tmem_o, tmem_p + ks * 8, db, p.idesc_o, (t != 0) || ks != 0);
        }
        umma_commit(&v_empty[vst]);
        umma_commit(o_bar);
        if (++vst == VST) { vst = 0; vph ^= 1; }
      }
    }
  } else if (warp >= SM_WARP0) {
    // ================================================================= online softmax + epilogue
    const int ew = warp - SM_WARP0;
    const int row = ew * 32 + lane;
    const int q = q0 + row;
    const uint32_t lane_addr = uint32_t(ew * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f;
    const uint8_t* mrow = p.m.mask == nullptr ? nullptr
                                              : p.m.mask + (long long)b * p.m.sb + (long long)h * p.m.sh + (long long)min(q, p.Sq - 1) * p.m.sq;
    const bool mask_vec = mrow != nullptr && ((reinterpret_cast<uintptr_t>(mrow) & 15) == 0);
    for (int t = 0; t < ntiles; ++t) {
      const int kv0 = (j0 + t) * BN;
      mbar_wait(&s_full[t & 1], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_s = tmem_base + lane_addr + (t & 1) * BN;
      float s[BN];
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t raw[32];
        tmem_ld_32x32(tmem_s + c * 32, raw);
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(raw[i]);
      }
      tmem_ld_wait();
      const int off = p.Sk - p.Sq;
      const bool need_mask = (kv0 + BN > p.Sk) || (p.m.causal && kv0 + BN - 1 > q0 + off) ||
                             (p.m.window > 0 && kv0 <= q0 + BM - 1 + off - p.m.window) || mrow != nullptr;
      if (need_mask) {
        if (mrow != nullptr) {
          if (mask_vec && kv0 + BN <= p.Sk) {
#pragma unroll
            for (int c = 0; c < BN / 16; ++c) {
              const uint4 mv = *reinterpret_cast<const uint4*>(mrow + kv0 + c * 16);
              const uint8_t* mb = reinterpret_cast<const uint8_t*>(&mv);
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (mb[i]) s[c * 16 + i] = -INFINITY;
            }
          } else {
#pragma unroll
            for (int i = 0; i < BN; ++i)
              if (kv0 + i < p.Sk && mrow[kv0 + i]) s[i] = -INFINITY;
          }
        }
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (!key_visible(p.m, q, kv0 + i, p.Sq, p.Sk)) s[i] = -INFINITY;
      }
      float mx0 = s[0], mx1 = s[1], mx2 = s[2], mx3 = s[3];
#pragma unroll
      for (int i = 4; i < BN; i += 4) {
        mx0 = fmaxf(mx0, s[i]); mx1 = fmaxf(mx1, s[i + 1]); mx2 = fmaxf(mx2, s[i + 2]); mx3 = fmaxf(mx3, s[i + 3]);
      }
      const float m_tile = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;
      // lazy rescale: keep the stale maximum unless it would let exp2 grow beyond 2^RESCALE_THRESHOLD
      const bool grow = m_tile > m_run + RESCALE_THRESHOLD || (m_run == -INFINITY && m_tile > -INFINITY);
      float alpha = 1.f;
      if (grow) {
        alpha = (m_run == -INFINITY) ? 0.f : fast_exp2(m_run - m_tile);
        m_run = m_tile;
        l_run *= alpha;
      }
      if (t > 0 && __any_sync(0xffffffffu, grow)) {
        // O accumulates P V products computed against the old maximum: wait for the last one, scale this row
        mbar_wait(o_bar, (t - 1) & 1);
        tc_fence_after();
        for (int c = 0; c < p.d / 32; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32(tmem_o + lane_addr + c * 32, raw);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) raw[i] = __float_as_uint(__uint_as_float(raw[i]) * alpha);
          tmem_st_32x32(tmem_o + lane_addr + c * 32, raw);
        }
      }
      const float m_use = (m_run == -INFINITY) ? 0.f : m_run;
      float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
      for (int c = 0; c < BN / 64; ++c) {
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          float pv[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) pv[e] = fast_exp2(fmaf(s[c * 64 + i + e], p.scale_log2, -m_use));
          sum0 += pv[0] + pv[2];
          sum1 += pv[1] + pv[3];
          if (p.drop.enabled) {
            const uint4 w = dropout_words(p.drop, bh, q, (kv0 + c * 64 + i) >> 2);   // 4 consecutive keys per call
            pv[0] = (w.x < p.drop.threshold) ? 0.f : pv[0] * p.drop.inv_keep;
            pv[1] = (w.y < p.drop.threshold) ? 0.f : pv[1] * p.drop.inv_keep;
            pv[2] = (w.z < p.drop.threshold) ? 0.f : pv[2] * p.drop.inv_keep;
            pv[3] = (w.w < p.drop.threshold) ? 0.f : pv[3] * p.drop.inv_keep;
          }
          if (p.out_dtype == SMPK_BF16) {
            __nv_bfloat162 h0 = __floats2bfloat162_rn(pv[0], pv[1]), h1 = __floats2bfloat162_rn(pv[2], pv[3]);
            pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h0);
            pk[(i >> 1) + 1] = *reinterpret_cast<uint32_t*>(&h1);
          } else {
            __half2 h0 = __floats2half2_rn(pv[0], pv[1]), h1 = __floats2half2_rn(pv[2], pv[3]);
            pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h0);
            pk[(i >> 1) + 1] = *reinterpret_cast<uint32_t*>(&h1);
          }
        }
        tmem_st_32x32(tmem_s + c * 32, pk);
      }
      l_run += sum0 + sum1;
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t & 1]);
    }
    // ---- epilogue: O / l -> global, log-sum-exp
    const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
    if (ntiles > 0) {
      mbar_wait(o_bar, (ntiles - 1) & 1);
      tc_fence_after();
    }
    char* orow = reinterpret_cast<char*>(p.O) + ((long long)b * p.o_sb + (long long)h * p.o_sh + (long long)q * p.o_ss) * 2;
    const bool o_vec = (reinterpret_cast<uintptr_t>(orow) & 15) == 0;
    for (int c = 0; c < p.d / 32; ++c) {
      uint32_t raw[32];
      if (ntiles > 0) {
        tmem_ld_32x32(tmem_o + lane_addr + c * 32, raw);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) raw[i] = 0u;
      }
      if (q < p.Sq) {
        uint4 pk[4];
        if (p.out_dtype == SMPK_BF16) {
          __nv_bfloat162* hv = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
          for (int i = 0; i < 16; ++i) hv[i] = __floats2bfloat162_rn(__uint_as_float(raw[2 * i]) * inv_l, __uint_as_float(raw[2 * i + 1]) * inv_l);
        } else {
          __half2* hv = reinterpret_cast<__half2*>(pk);
#pragma unroll
          for (int i = 0; i < 16; ++i) hv[i] = __floats2half2_rn(__uint_as_float(raw[2 * i]) * inv_l, __uint_as_float(raw[2 * i + 1]) * inv_l);
        }
        if (o_vec) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(orow + c * 64 + j * 16) = pk[j];
        } else {
          const uint16_t* e = reinterpret_cast<const uint16_t*>(pk);
#pragma unroll
          for (int i = 0; i < 32; ++i) *reinterpret_cast<uint16_t*>(orow + c * 64 + i * 2) = e[i];
        }
      }
    }
    if (q < p.Sq && p.lse != nullptr) p.lse[(long long)bh * p.Sq + q] = l_run > 0.f ? m_run + log2f(l_run) : INFINITY;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

template <int NCH, int BN, int KST, int VST>
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const FwdParams& p, cudaStream_t stream) {
  using C = FwdCfg<NCH, BN, KST, VST>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    SMPK_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<NCH, BN, KST, VST>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  dim3 grid((unsigned)((p.Sq + BM - 1) / BM), (unsigned)(p.B * p.H));
  attn_fwd_kernel<NCH, BN, KST, VST><<<grid, NUM_THREADS, C::SMEM_BYTES, stream>>>(tq, tk, tv, p);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

// See include/smpb200_kernels.h for the contract.
SMPK_API int smpk_attn_fwd(void* stream, int dtype, int B, int H, int Sq, int Sk, int d, const void* q, long long q_sb,
                           long long q_sh, long long q_ss, const void* k, long long k_sb, long long k_sh, long long k_ss,
                           const void* v, long long v_sb, long long v_sh, long long v_ss, void* o, long long o_sb,
                           long long o_sh, long long o_ss, float* lse, float scale, int causal, int window,
                           const void* mask, long long m_sb, long long m_sh, long long m_sq, float p_drop,
                           unsigned long long seed, unsigned long long offset) {
  SMPK_REQUIRE(dtype == SMPK_BF16 || dtype == SMPK_F16, "attn_fwd: inputs must be bf16 or fp16");
  SMPK_REQUIRE(B > 0 && H > 0 && Sq > 0 && Sk > 0, "attn_fwd: empty problem");
  SMPK_REQUIRE(d % 32 == 0 && d >= 32 && d <= 256, "attn_fwd: head_dim %d must be a multiple of 32 in [32, 256]", d);
  SMPK_REQUIRE(scale > 0.f, "attn_fwd: scale must be positive");
  SMPK_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "attn_fwd: dropout probability %f out of range", p_drop);
  FwdParams p;
  memset(&p, 0, sizeof(p));
  p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.d = d;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.m.mask = reinterpret_cast<const uint8_t*>(mask); p.m.sb = m_sb; p.m.sh = m_sh; p.m.sq = m_sq;
  p.m.causal = causal; p.m.window = window;
  p.drop.enabled = p_drop > 0.f;
  p.drop.seed_lo = (uint32_t)seed; p.drop.seed_hi = (uint32_t)(seed >> 32);
  p.drop.off_lo = (uint32_t)offset; p.drop.off_hi = (uint32_t)(offset >> 32);
  p.drop.threshold = (uint32_t)fmin(4294967295.0, (double)p_drop * 4294967296.0);
  p.drop.inv_keep = 1.f / (1.f - p_drop);
  p.O = o; p.o_sb = o_sb; p.o_sh = o_sh; p.o_ss = o_ss; p.out_dtype = dtype;
  p.lse = lse;
  const uint32_t fmt = dtype == SMPK_BF16 ? 1u : 0u;
  const int nch = (d + 63) / 64;
  const int BN = 128;
  p.idesc_s = make_idesc_f16(fmt, 0, 0, BM, BN);
  p.idesc_o = make_idesc_f16(fmt, 0, 1, BM, d);
  CUtensorMap tq, tk, tv;
  int rc = make_map4(&tq, dtype, q, B, H, Sq, d, q_sb, q_sh, q_ss, BM);
  if (rc) return rc;
  rc = make_map4(&tk, dtype, k, B, H, Sk, d, k_sb, k_sh, k_ss, BN);
  if (rc) return rc;
  rc = make_map4(&tv, dtype, v, B, H, Sk, d, v_sb, v_sh, v_ss, BN);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (nch) {
    case 1: return launch_fwd<1, 128, 3, 3>(tq, tk, tv, p, s);
    case 2: return launch_fwd<2, 128, 2, 2>(tq, tk, tv, p, s);
    case 3: return launch_fwd<3, 128, 2, 1>(tq, tk, tv, p, s);
    default: return launch_fwd<4, 128, 1, 1>(tq, tk, tv, p, s);
  }
}
