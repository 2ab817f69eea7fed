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

// GENERAL = false: no byte mask and no dropout (causal / local-window / length masks only) - the straight-line fast path.
template <int NCH, int BN, int KST, int VST, bool GENERAL>
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
      int kst = 0, vst = 0;
      uint32_t kph = 0, vph = 0;
      auto issue_s = [&](int t) {   // S_t = Q K_t^T into score buffer t & 1
        mbar_wait(&k_full[kst], kph);
        tc_fence_after();
        const uint32_t tmem_s = tmem_base + (t & 1) * BN;
        // reduction over the padded head_dim (TMA zero-fills the columns beyond d): a compile-time MMA count
#pragma unroll
        for (int ks = 0; ks < NCH * 4; ++ks)
          umma_f16(tmem_s, make_smem_desc(sQ + (ks >> 2) * (BM * 128) + (ks & 3) * 32, hi_k),
                   make_smem_desc(sK + kst * C::KV_BYTES + (ks >> 2) * (BN * 128) + (ks & 3) * 32, hi_k), p.idesc_s, ks != 0);
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
        for (int ks = 0; ks < BN / 16; ++ks)
          umma_f16_ts(tmem_o, tmem_p + ks * 8, make_smem_desc(sV + vst * C::KV_BYTES + ks * (16 * 128), hi_v), p.idesc_o,
                      (t != 0) || ks != 0);
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
    const uint8_t* mrow = (!GENERAL || p.m.mask == nullptr) ? nullptr
                                              : p.m.mask + (long long)b * p.m.sb + (long long)h * p.m.sh + (long long)min(q, p.Sq - 1) * p.m.sq;
    const bool mask_vec = mrow != nullptr && ((reinterpret_cast<uintptr_t>(mrow) & 15) == 0);
    const int off = p.Sk - p.Sq;
    // keys visible to this row form one interval [k_lo, k_hi] (length, causal and window limits)
    const int k_hi = p.m.causal ? min(p.Sk - 1, q + off) : p.Sk - 1;
    const int k_lo = p.m.window > 0 ? max(0, q + off - p.m.window + 1) : 0;
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
      // the whole tile is visible to every row unless it crosses a length / causal / window boundary
      const bool need_range = (kv0 + BN > p.Sk) || (p.m.causal && kv0 + BN - 1 > q0 + off) ||
                              (p.m.window > 0 && kv0 <= q0 + BM - 1 + off - p.m.window);
      if (GENERAL && mrow != nullptr) {
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
      if (need_range) {
        const int lo = k_lo - kv0;
        const unsigned span = (unsigned)(k_hi - k_lo);      // k_hi < k_lo (no visible key) wraps to a huge span only
        const bool none = k_hi < k_lo;                      // when negative, so test it separately
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (none || (unsigned)(i - lo) > span) s[i] = -INFINITY;
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
          if (GENERAL && p.drop.enabled) {
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

template <int NCH, int BN, int KST, int VST, bool GENERAL>
int launch_fwd_g(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const FwdParams& p, cudaStream_t stream) {
  using C = FwdCfg<NCH, BN, KST, VST>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    SMPK_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<NCH, BN, KST, VST, GENERAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  dim3 grid((unsigned)((p.Sq + BM - 1) / BM), (unsigned)(p.B * p.H));
  attn_fwd_kernel<NCH, BN, KST, VST, GENERAL><<<grid, NUM_THREADS, C::SMEM_BYTES, stream>>>(tq, tk, tv, p);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
template <int NCH, int BN, int KST, int VST>
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const FwdParams& p, cudaStream_t stream) {
  if (p.m.mask != nullptr || p.drop.enabled) return launch_fwd_g<NCH, BN, KST, VST, true>(tq, tk, tv, p, stream);
  return launch_fwd_g<NCH, BN, KST, VST, false>(tq, tk, tv, p, stream);
}

// ================================================================================================ backward
// delta[bh, q] = sum_d dO[q, d] * O[q, d]   (the softmax-Jacobian row term), one warp per row.
struct DeltaParams {
  const void *O, *dO;
  long long o_sb, o_sh, o_ss, g_sb, g_sh, g_ss;
  float* delta;
  int B, H, Sq, d, dtype;
};
__global__ void __launch_bounds__(256) attn_delta_kernel(const DeltaParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rowid = (long long)blockIdx.x * 8 + warp;
  if (rowid >= (long long)p.B * p.H * p.Sq) return;
  const int q = (int)(rowid % p.Sq);
  const int bh = (int)(rowid / p.Sq), b = bh / p.H, h = bh % p.H;
  const uint16_t* o = reinterpret_cast<const uint16_t*>(p.O) + (long long)b * p.o_sb + (long long)h * p.o_sh + (long long)q * p.o_ss;
  const uint16_t* g = reinterpret_cast<const uint16_t*>(p.dO) + (long long)b * p.g_sb + (long long)h * p.g_sh + (long long)q * p.g_ss;
  float acc = 0.f;
  const bool vec = ((reinterpret_cast<uintptr_t>(o) | reinterpret_cast<uintptr_t>(g)) & 15) == 0;
  if (vec) {
    for (int c = lane * 8; c < p.d; c += 256) {
      float a[8], bq[8];
      smpk::load8(o, p.dtype, c, a);
      smpk::load8(g, p.dtype, c, bq);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc += a[i] * bq[i];
    }
  } else {
    for (int c = lane; c < p.d; c += 32) acc += smpk::load1(o, p.dtype, c) * smpk::load1(g, p.dtype, c);
  }
  acc = smpk::warp_sum(acc);
  if (lane == 0) p.delta[rowid] = acc;
}

struct BwdParams {
  int B, H, Sq, Sk, d;
  float scale, scale_log2;
  MaskParams m;
  DropoutParams drop;
  const float* lse;      // [B*H, Sq] log2 domain (from the forward kernel)
  const float* delta;    // [B*H, Sq]
  // outputs (16 bit, strided [B, H, S, d]); out_p = dV (key-stationary), out_ds = dK (key-stationary) or dQ
  void* out_p;
  long long p_sb, p_sh, p_ss;
  void* out_ds;
  long long s_sb, s_sh, s_ss;
  int dtype;
  uint32_t idesc_t, idesc_o;
};

template <int NCH, int BS, int Y1ST, int Y2ST, bool NEED_X2>
struct BwdCfg {
  static constexpr int X_BYTES = NCH * BM * 128;             // one resident operand
  static constexpr int Y_BYTES = NCH * BS * 128;             // one streamed operand tile
  static constexpr int STAT_BYTES = 2 * 2 * BS * 4;          // lse / delta of the streamed queries, double buffered
  static constexpr int SMEM_BYTES = (NEED_X2 ? 2 : 1) * X_BYTES + (Y1ST + Y2ST) * Y_BYTES + STAT_BYTES + 1024 + 256;
};

constexpr int BWD_THREADS = 384;   // warps 0-3: producers / MMA / TMEM, warps 4-11: two elementwise warpgroups

template <int N>
__device__ __forceinline__ void tmem_ld_n(uint32_t taddr, uint32_t (&r)[N]) {
  if constexpr (N == 32) tmem_ld_32x32(taddr, r);
  else tmem_ld_32x16(taddr, r);
}
template <int N>
__device__ __forceinline__ void tmem_st_n(uint32_t taddr, const uint32_t (&r)[N]) {
  if constexpr (N == 16) tmem_st_32x16(taddr, r);
  else tmem_st_32x8(taddr, r);
}
__device__ __forceinline__ uint32_t pack2(float a, float b, int dtype) {
  if (dtype == SMPK_BF16) {
    __nv_bfloat162 hv = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&hv);
  }
  __half2 hv = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&hv);
}

// TR  = true : resident rows are KEYS   (X1 = K, X2 = V; streamed Y1 = Q_i, Y2 = dO_i), T = S^T / dP^T
// TR  = false: resident rows are QUERIES (X1 = Q, X2 = dO; streamed Y1 = K_j, Y2 = V_j), T = S / dP
// OUT_P : accumulate P^T dO_i  (dV)           - only with TR
// OUT_DS: accumulate dS^T Q_i (dK) with TR, dS K_j (dQ) without
// GENERAL = false: no byte mask and no dropout - the straight-line elementwise stage.
// Y1ST / Y2ST: ring depth of the two streamed operands.  The first one is read by a score product and by an output
// product of the same tile, so it needs two stages for the score products to run one tile ahead; the second one
// is read only by the second score product when dV is not accumulated here and then gets by with one stage.
template <int NCH, int BS, int Y1ST, int Y2ST, bool TR, bool OUT_P, bool OUT_DS, bool GENERAL>
__global__ void __launch_bounds__(BWD_THREADS, 1)
    attn_bwd_kernel(const __grid_constant__ CUtensorMap tma_x1, const __grid_constant__ CUtensorMap tma_x2,
                    const __grid_constant__ CUtensorMap tma_y1, const __grid_constant__ CUtensorMap tma_y2,
                    const BwdParams p) {
  constexpr bool NEED_T2 = OUT_DS;              // dP is only needed for dS
  constexpr bool NEED_X2 = NEED_T2;
  constexpr bool NEED_Y2 = NEED_T2 || OUT_P;
  using C = BwdCfg<NCH, BS, Y1ST, Y2ST, NEED_X2>;
  static_assert(Y1ST >= 2 && (Y2ST >= 2 || (!OUT_P && Y2ST == 1)), "ring depths");
  constexpr int HC = BS / 2;                    // columns of a score tile handled by one elementwise warpgroup
  constexpr int DPAD = NCH * 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sX1 = smem;
  uint8_t* sX2 = sX1 + C::X_BYTES;
  uint8_t* sY1 = sX1 + (NEED_X2 ? 2 : 1) * C::X_BYTES;
  uint8_t* sY2 = sY1 + Y1ST * C::Y_BYTES;
  float* sStat = reinterpret_cast<float*>(sY2 + Y2ST * C::Y_BYTES);     // [2][2][BS]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sStat) + C::STAT_BYTES);
  uint64_t* x_full = bars;
  uint64_t* y1_full = bars + 1;
  uint64_t* y1_empty = y1_full + Y1ST;
  uint64_t* y2_full = y1_empty + Y1ST;
  uint64_t* y2_empty = y2_full + Y2ST;
  uint64_t* t_full = y2_empty + Y2ST;      // [2] MMA -> elementwise: score tiles of buffer b complete
  uint64_t* pd_full = t_full + 2;          // [2] elementwise -> MMA: P / dS of buffer b written (8 warp arrivals)
  uint64_t* acc_done = pd_full + 2;        // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int off = p.Sk - p.Sq;
  const int rt = TR ? (int)blockIdx.x : (int)gridDim.x - 1 - (int)blockIdx.x;   // resident tile
  const int r0 = rt * BM;
  // streamed tile range [i0, i1)
  int i0, i1;
  if (!TR) {
    kv_tile_range(p.m, r0, BM, p.Sq, p.Sk, BS, i0, i1);
  } else {
    int lo = 0, hi = p.Sq - 1;                     // queries that can see a key of [r0, r0 + BM)
    if (p.m.causal) lo = max(0, r0 - off);
    if (p.m.window > 0) hi = min(hi, min(r0 + BM, p.Sk) - 1 - off + p.m.window - 1);
    i0 = lo / BS;
    i1 = hi < 0 ? 0 : hi / BS + 1;
    if (i1 < i0) i1 = i0;
  }
  const int ntiles = i1 - i0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_x1);
    tma_prefetch_desc(&tma_y1);
    if (NEED_X2) tma_prefetch_desc(&tma_x2);
    if (NEED_Y2) tma_prefetch_desc(&tma_y2);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(x_full, 1);
    for (int i = 0; i < Y1ST; ++i) { mbar_init(&y1_full[i], 1); mbar_init(&y1_empty[i], 1); }
    for (int i = 0; i < Y2ST; ++i) { mbar_init(&y2_full[i], 1); mbar_init(&y2_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&t_full[i], 1); mbar_init(&pd_full[i], 8); }
    mbar_init(acc_done, 1);
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
  const uint32_t tmem_acc0 = tmem_base + 4 * BS;               // dV (OUT_P) or, without OUT_P, the dS product
  const uint32_t tmem_acc1 = tmem_acc0 + DPAD;                 // dK when both outputs are accumulated

  if (warp == 0) {
    // ================================================================= resident operands, then Y1 tiles
    if (lane == 0 && ntiles > 0) {
      mbar_expect_tx(x_full, (NEED_X2 ? 2 : 1) * C::X_BYTES);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        tma_load_4d(sX1 + c * (BM * 128), &tma_x1, x_full, c * 64, r0, h, b);
        if (NEED_X2) tma_load_4d(sX2 + c * (BM * 128), &tma_x2, x_full, c * 64, r0, h, b);
      }
      int st = 0;
      uint32_t ph = 0;
      for (int i = i0; i < i1; ++i) {
        mbar_wait(&y1_empty[st], ph ^ 1);
        mbar_expect_tx(&y1_full[st], C::Y_BYTES);
#pragma unroll
        for (int c = 0; c < NCH; ++c)
          tma_load_4d(sY1 + st * C::Y_BYTES + c * (BS * 128), &tma_y1, &y1_full[st], c * 64, i * BS, h, b);
        if (++st == Y1ST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 3) {
    // ================================================================= Y2 tiles
    if (lane == 0 && NEED_Y2) {
      int st = 0;
      uint32_t ph = 0;
      for (int i = i0; i < i1; ++i) {
        mbar_wait(&y2_empty[st], ph ^ 1);
        mbar_expect_tx(&y2_full[st], C::Y_BYTES);
#pragma unroll
        for (int c = 0; c < NCH; ++c)
          tma_load_4d(sY2 + st * C::Y_BYTES + c * (BS * 128), &tma_y2, &y2_full[st], c * 64, i * BS, h, b);
        if (++st == Y2ST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0 && ntiles > 0) {
      const uint64_t hi_k = make_smem_desc_hi(16, 1024);            // K-major operands of the score products
      const uint64_t hi_n = make_smem_desc_hi(BS * 128, 1024);      // streamed tile as the N-major B of an output product
      int t1 = 0, t2 = 0, o1 = 0, o2 = 0;          // ring positions: score products (t*) and output products (o*)
      uint32_t ph1 = 0, ph2 = 0, oph2 = 0;
      auto issue_t = [&](int t) {
        const uint32_t tm = tmem_base + (t & 1) * (2 * BS);
        mbar_wait(&y1_full[t1], ph1);
        tc_fence_after();
        // the reduction runs over the padded head_dim (NCH * 64 columns; TMA zero-fills the padding), so the number
        // of MMAs is a compile-time constant and the loop unrolls without predication
#pragma unroll
        for (int ks = 0; ks < NCH * 4; ++ks)
          umma_f16(tm, make_smem_desc(sX1 + (ks >> 2) * (BM * 128) + (ks & 3) * 32, hi_k),
                   make_smem_desc(sY1 + t1 * C::Y_BYTES + (ks >> 2) * (BS * 128) + (ks & 3) * 32, hi_k), p.idesc_t, ks != 0);
        if (NEED_T2) {
          mbar_wait(&y2_full[t2], ph2);
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < NCH * 4; ++ks)
            umma_f16(tm + BS, make_smem_desc(sX2 + (ks >> 2) * (BM * 128) + (ks & 3) * 32, hi_k),
                     make_smem_desc(sY2 + t2 * C::Y_BYTES + (ks >> 2) * (BS * 128) + (ks & 3) * 32, hi_k), p.idesc_t, ks != 0);
          if (!OUT_P) umma_commit(&y2_empty[t2]);     // the second streamed operand is not read again
          if (++t2 == Y2ST) { t2 = 0; ph2 ^= 1; }
        }
        umma_commit(&t_full[t & 1]);
        if (++t1 == Y1ST) { t1 = 0; ph1 ^= 1; }
      };
      mbar_wait(x_full, 0);
      issue_t(0);
      for (int t = 0; t < ntiles; ++t) {
        if (t + 1 < ntiles) issue_t(t + 1);
        mbar_wait(&pd_full[t & 1], (t >> 1) & 1);
        tc_fence_after();
        const uint32_t tm = tmem_base + (t & 1) * (2 * BS);
        if (OUT_P) {
          // (the streamed dO tile was already waited for by issue_t when NEED_T2; otherwise wait here)
          if (!NEED_T2) {
            mbar_wait(&y2_full[o2], oph2);
            tc_fence_after();
          }
#pragma unroll
          for (int ks = 0; ks < BS / 16; ++ks)
            umma_f16_ts(tmem_acc0, tm + (ks / (HC / 16)) * HC + (ks % (HC / 16)) * 8,
                        make_smem_desc(sY2 + o2 * C::Y_BYTES + ks * (16 * 128), hi_n), p.idesc_o, (t != 0) || ks != 0);
          umma_commit(&y2_empty[o2]);
          if (++o2 == Y2ST) { o2 = 0; oph2 ^= 1; }
        }
        if (OUT_DS) {
          const uint32_t acc = OUT_P ? tmem_acc1 : tmem_acc0;
#pragma unroll
          for (int ks = 0; ks < BS / 16; ++ks)
            umma_f16_ts(acc, tm + BS + (ks / (HC / 16)) * HC + (ks % (HC / 16)) * 8,
                        make_smem_desc(sY1 + o1 * C::Y_BYTES + ks * (16 * 128), hi_n), p.idesc_o, (t != 0) || ks != 0);
        }
        umma_commit(&y1_empty[o1]);
        if (++o1 == Y1ST) o1 = 0;
      }
      umma_commit(acc_done);
    }
  } else if (warp >= SM_WARP0) {
    // ================================================================= elementwise: scores -> P, dS
    const int ew = warp & 3;                       // TMEM lane quarter
    const int g = (warp - SM_WARP0) >> 2;          // column half
    const int row = ew * 32 + lane;
    const int rglob = r0 + row;                    // key index (TR) or query index
    const uint32_t lane_addr = uint32_t(ew * 32) << 16;
    const int et = threadIdx.x - SM_WARP0 * 32;    // 0..255
    float my_lse = 0.f, my_delta = 0.f;
    if (!TR) {
      const int qq = min(rglob, p.Sq - 1);
      my_lse = p.lse[(long long)bh * p.Sq + qq];
      my_delta = p.delta[(long long)bh * p.Sq + qq];
    }
    // streamed indices visible to this row form one interval [v_lo, v_hi]: keys of a query row, or (transposed)
    // queries of a key row
    int v_lo, v_hi;
    if (!TR) {
      v_hi = p.m.causal ? min(p.Sk - 1, rglob + off) : p.Sk - 1;
      v_lo = p.m.window > 0 ? max(0, rglob + off - p.m.window + 1) : 0;
      if (rglob >= p.Sq) v_hi = -1;
    } else {
      v_lo = p.m.causal ? max(0, rglob - off) : 0;
      v_hi = p.m.window > 0 ? min(p.Sq - 1, rglob - off + p.m.window - 1) : p.Sq - 1;
      if (rglob >= p.Sk) v_hi = -1;
    }
    const bool v_none = v_hi < v_lo;
    const unsigned v_span = (unsigned)(v_hi - v_lo);
    const uint8_t* mbase = (!GENERAL || p.m.mask == nullptr) ? nullptr : p.m.mask + (long long)b * p.m.sb + (long long)h * p.m.sh;
    for (int t = 0; t < ntiles; ++t) {
      const int c0 = (i0 + t) * BS;                // first streamed index of the tile (query if TR, key otherwise)
      float* stat = sStat + (t & 1) * (2 * BS);
      if (TR) {
        // row statistics of the streamed queries, shared by all threads of the tile
        if (et < 2 * BS) {
          const int qq = min(c0 + (et % BS), p.Sq - 1);
          stat[et] = (et < BS ? p.lse : p.delta)[(long long)bh * p.Sq + qq];
        }
        named_bar_sync(1, 256);
      }
      mbar_wait(&t_full[t & 1], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t tm = tmem_base + lane_addr + (t & 1) * (2 * BS);
      uint32_t sraw[HC], draw[HC];
      tmem_ld_n<HC>(tm + g * HC, sraw);
      if (NEED_T2) tmem_ld_n<HC>(tm + BS + g * HC, draw);
      tmem_ld_wait();
      uint32_t ppk[HC / 2], dpk[HC / 2];
      const int lo = v_lo - (c0 + g * HC);        // tile-relative start of the visible interval
#pragma unroll
      for (int i = 0; i < HC; i += 4) {
        float pv[4], dv[4], lse4[4], dl4[4];
        if (TR) {
          const float4 a = *reinterpret_cast<const float4*>(stat + g * HC + i);
          const float4 d4 = *reinterpret_cast<const float4*>(stat + BS + g * HC + i);
          lse4[0] = a.x; lse4[1] = a.y; lse4[2] = a.z; lse4[3] = a.w;
          dl4[0] = d4.x; dl4[1] = d4.y; dl4[2] = d4.z; dl4[3] = d4.w;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) { lse4[e] = my_lse; dl4[e] = my_delta; }
        }
        uint4 w = make_uint4(0, 0, 0, 0);
        if (GENERAL && p.drop.enabled && !TR) w = dropout_words(p.drop, bh, rglob, (c0 + g * HC + i) >> 2);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const bool vis = !v_none && (unsigned)(i + e - lo) <= v_span;
          const float arg = fmaf(__uint_as_float(sraw[i + e]), p.scale_log2, -lse4[e]);
          float de = NEED_T2 ? __uint_as_float(draw[i + e]) : 0.f;
          bool dead = !vis;
          if (GENERAL) {
            const int cc = c0 + g * HC + i + e;
            const int qi = TR ? cc : rglob, ki = TR ? rglob : cc;
            if (mbase != nullptr && vis) dead = mbase[(long long)min(qi, p.Sq - 1) * p.m.sq + ki] != 0;
          }
          const float pe = dead ? 0.f : fast_exp2(arg);
          float pd = pe;                           // probability as used by the P V product (after dropout)
          if (GENERAL && p.drop.enabled) {
            uint32_t word;
            if (TR) {
              const int cc = c0 + g * HC + i + e;
              const uint4 wq = dropout_words(p.drop, bh, cc, rglob >> 2);
              word = (rglob & 3) == 0 ? wq.x : ((rglob & 3) == 1 ? wq.y : ((rglob & 3) == 2 ? wq.z : wq.w));
            } else {
              word = e == 0 ? w.x : (e == 1 ? w.y : (e == 2 ? w.z : w.w));
            }
            const bool keep = word >= p.drop.threshold;
            pd = keep ? pe * p.drop.inv_keep : 0.f;
            de = keep ? de * p.drop.inv_keep : 0.f;
          }
          pv[e] = pd;
          dv[e] = pe * (de - dl4[e]);
        }
        ppk[i >> 1] = pack2(pv[0], pv[1], p.dtype);
        ppk[(i >> 1) + 1] = pack2(pv[2], pv[3], p.dtype);
        dpk[i >> 1] = pack2(dv[0], dv[1], p.dtype);
        dpk[(i >> 1) + 1] = pack2(dv[2], dv[3], p.dtype);
      }
      // (each warpgroup overwrites only columns it has just read itself: P / dS of half g live at the start of that half)
      if (OUT_P) tmem_st_n<HC / 2>(tm + g * HC, ppk);
      if (OUT_DS) tmem_st_n<HC / 2>(tm + BS + g * HC, dpk);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&pd_full[t & 1]);
    }
    // ---- epilogue: accumulators -> global (the two warpgroups interleave 32-column chunks)
    if (ntiles > 0) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
    }
    const int S_out = TR ? p.Sk : p.Sq;
    auto store_acc = [&](uint32_t tmem_acc, void* out, long long sb, long long sh, long long ss, float mul) {
      char* orow = reinterpret_cast<char*>(out) + ((long long)b * sb + (long long)h * sh + (long long)rglob * ss) * 2;
      const bool vec = (reinterpret_cast<uintptr_t>(orow) & 15) == 0;
      for (int c = g; c < p.d / 32; c += 2) {
        uint32_t raw[32];
        if (ntiles > 0) {
          tmem_ld_32x32(tmem_acc + lane_addr + c * 32, raw);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) raw[i] = 0u;
        }
        if (rglob < S_out) {
          uint4 pk[4];
          uint32_t* pw = reinterpret_cast<uint32_t*>(pk);
#pragma unroll
          for (int i = 0; i < 16; ++i) pw[i] = pack2(__uint_as_float(raw[2 * i]) * mul, __uint_as_float(raw[2 * i + 1]) * mul, p.dtype);
          if (vec) {
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(orow + c * 64 + j * 16) = pk[j];
          } else {
            const uint16_t* e = reinterpret_cast<const uint16_t*>(pk);
#pragma unroll
            for (int i = 0; i < 32; ++i) *reinterpret_cast<uint16_t*>(orow + c * 64 + i * 2) = e[i];
          }
        }
      }
    };
    if (OUT_P) store_acc(tmem_acc0, p.out_p, p.p_sb, p.p_sh, p.p_ss, 1.f);
    if (OUT_DS) store_acc(OUT_P ? tmem_acc1 : tmem_acc0, p.out_ds, p.s_sb, p.s_sh, p.s_ss, p.scale);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

template <int NCH, int BS, int Y1ST, int Y2ST, bool TR, bool OUT_P, bool OUT_DS, bool GENERAL>
int launch_bwd_g(const CUtensorMap& x1, const CUtensorMap& x2, const CUtensorMap& y1, const CUtensorMap& y2,
                 const BwdParams& p, cudaStream_t stream) {
  using C = BwdCfg<NCH, BS, Y1ST, Y2ST, OUT_DS>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  auto kern = attn_bwd_kernel<NCH, BS, Y1ST, Y2ST, TR, OUT_P, OUT_DS, GENERAL>;
  if (!attr_set[dev & 63]) {
    SMPK_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int S_res = TR ? p.Sk : p.Sq;
  dim3 grid((unsigned)((S_res + BM - 1) / BM), (unsigned)(p.B * p.H));
  kern<<<grid, BWD_THREADS, C::SMEM_BYTES, stream>>>(x1, x2, y1, y2, p);
  SMPK_CHECK_CUDA(cudaGetLastError());
  return 0;
}
template <int NCH, int BS, int Y1ST, int Y2ST, bool TR, bool OUT_P, bool OUT_DS>
int launch_bwd(const CUtensorMap& x1, const CUtensorMap& x2, const CUtensorMap& y1, const CUtensorMap& y2,
               const BwdParams& p, cudaStream_t stream) {
  if (p.m.mask != nullptr || p.drop.enabled)
    return launch_bwd_g<NCH, BS, Y1ST, Y2ST, TR, OUT_P, OUT_DS, true>(x1, x2, y1, y2, p, stream);
  return launch_bwd_g<NCH, BS, Y1ST, Y2ST, TR, OUT_P, OUT_DS, false>(x1, x2, y1, y2, p, stream);
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
  const int BN = nch == 4 ? 64 : 128;      // head_dim 256: 64-key tiles so that K and V are both double buffered
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
    default: return launch_fwd<4, 64, 2, 2>(tq, tk, tv, p, s);
  }
}

// Backward of smpk_attn_fwd.  `delta` is caller-provided scratch of B*H*Sq floats.  dq / dk / dv may be strided
// [B, H, S, d] views like the inputs.
SMPK_API int smpk_attn_bwd(void* stream, int dtype, int B, int H, int Sq, int Sk, int d, const void* q, long long q_sb,
                           long long q_sh, long long q_ss, const void* k, long long k_sb, long long k_sh, long long k_ss,
                           const void* v, long long v_sb, long long v_sh, long long v_ss, const void* o, long long o_sb,
                           long long o_sh, long long o_ss, const void* d_o, long long g_sb, long long g_sh,
                           long long g_ss, const float* lse, float* delta, void* dq, long long dq_sb, long long dq_sh,
                           long long dq_ss, void* dk, long long dk_sb, long long dk_sh, long long dk_ss, void* dv,
                           long long dv_sb, long long dv_sh, long long dv_ss, float scale, int causal, int window,
                           const void* mask, long long m_sb, long long m_sh, long long m_sq, float p_drop,
                           unsigned long long seed, unsigned long long offset) {
  SMPK_REQUIRE(dtype == SMPK_BF16 || dtype == SMPK_F16, "attn_bwd: inputs must be bf16 or fp16");
  SMPK_REQUIRE(B > 0 && H > 0 && Sq > 0 && Sk > 0, "attn_bwd: empty problem");
  SMPK_REQUIRE(d % 32 == 0 && d >= 32 && d <= 256, "attn_bwd: head_dim %d must be a multiple of 32 in [32, 256]", d);
  SMPK_REQUIRE(lse != nullptr && delta != nullptr, "attn_bwd: lse and delta scratch are required");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  {
    DeltaParams dp;
    dp.O = o; dp.dO = d_o; dp.o_sb = o_sb; dp.o_sh = o_sh; dp.o_ss = o_ss; dp.g_sb = g_sb; dp.g_sh = g_sh; dp.g_ss = g_ss;
    dp.delta = delta; dp.B = B; dp.H = H; dp.Sq = Sq; dp.d = d; dp.dtype = dtype;
    const long long rows = (long long)B * H * Sq;
    attn_delta_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(dp);
    SMPK_CHECK_CUDA(cudaGetLastError());
  }
  BwdParams p;
  memset(&p, 0, sizeof(p));
  p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.d = d;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  p.m.mask = reinterpret_cast<const uint8_t*>(mask); p.m.sb = m_sb; p.m.sh = m_sh; p.m.sq = m_sq;
  p.m.causal = causal; p.m.window = window;
  p.drop.enabled = p_drop > 0.f;
  p.drop.seed_lo = (uint32_t)seed; p.drop.seed_hi = (uint32_t)(seed >> 32);
  p.drop.off_lo = (uint32_t)offset; p.drop.off_hi = (uint32_t)(offset >> 32);
  p.drop.threshold = (uint32_t)fmin(4294967295.0, (double)p_drop * 4294967296.0);
  p.drop.inv_keep = 1.f / (1.f - p_drop);
  p.lse = lse; p.delta = delta; p.dtype = dtype;
  const uint32_t fmt = dtype == SMPK_BF16 ? 1u : 0u;
  const int nch = (d + 63) / 64;
  const int BS = 64;                           // rows of a streamed tile
  CUtensorMap tq_res, tdo_res, tk_res, tv_res, tq_str, tdo_str, tk_str, tv_str;
  int rc;
  if ((rc = make_map4(&tq_res, dtype, q, B, H, Sq, d, q_sb, q_sh, q_ss, BM))) return rc;
  if ((rc = make_map4(&tdo_res, dtype, d_o, B, H, Sq, d, g_sb, g_sh, g_ss, BM))) return rc;
  if ((rc = make_map4(&tk_res, dtype, k, B, H, Sk, d, k_sb, k_sh, k_ss, BM))) return rc;
  if ((rc = make_map4(&tv_res, dtype, v, B, H, Sk, d, v_sb, v_sh, v_ss, BM))) return rc;
  if ((rc = make_map4(&tk_str, dtype, k, B, H, Sk, d, k_sb, k_sh, k_ss, BS))) return rc;
  if ((rc = make_map4(&tv_str, dtype, v, B, H, Sk, d, v_sb, v_sh, v_ss, BS))) return rc;
  if ((rc = make_map4(&tq_str, dtype, q, B, H, Sq, d, q_sb, q_sh, q_ss, BS))) return rc;
  if ((rc = make_map4(&tdo_str, dtype, d_o, B, H, Sq, d, g_sb, g_sh, g_ss, BS))) return rc;
  p.idesc_t = make_idesc_f16(fmt, 0, 0, BM, BS);
  p.idesc_o = make_idesc_f16(fmt, 0, 1, BM, d);
  // ---- query-stationary pass: dQ
  p.out_p = nullptr;
  p.out_ds = dq; p.s_sb = dq_sb; p.s_sh = dq_sh; p.s_ss = dq_ss;
  switch (nch) {
    case 1: rc = launch_bwd<1, 64, 3, 3, false, false, true>(tq_res, tdo_res, tk_str, tv_str, p, s); break;
    case 2: rc = launch_bwd<2, 64, 2, 2, false, false, true>(tq_res, tdo_res, tk_str, tv_str, p, s); break;
    case 3: rc = launch_bwd<3, 64, 2, 2, false, false, true>(tq_res, tdo_res, tk_str, tv_str, p, s); break;
    default: rc = launch_bwd<4, 64, 2, 1, false, false, true>(tq_res, tdo_res, tk_str, tv_str, p, s); break;
  }
  if (rc) return rc;
  // ---- key-stationary pass(es): dK, dV
  p.out_p = dv; p.p_sb = dv_sb; p.p_sh = dv_sh; p.p_ss = dv_ss;
  p.out_ds = dk; p.s_sb = dk_sb; p.s_sh = dk_sh; p.s_ss = dk_ss;
  switch (nch) {
    case 1: return launch_bwd<1, 64, 3, 3, true, true, true>(tk_res, tv_res, tq_str, tdo_str, p, s);
    case 2: return launch_bwd<2, 64, 2, 2, true, true, true>(tk_res, tv_res, tq_str, tdo_str, p, s);
    case 3: {
      // 192 padded columns: two accumulators (384) + score tiles (256) exceed tensor memory - split like head_dim 256
      rc = launch_bwd<3, 64, 2, 2, true, false, true>(tk_res, tv_res, tq_str, tdo_str, p, s);
      if (rc) return rc;
      return launch_bwd<3, 64, 2, 2, true, true, false>(tk_res, tv_res, tq_str, tdo_str, p, s);
    }
    default: {
      rc = launch_bwd<4, 64, 2, 1, true, false, true>(tk_res, tv_res, tq_str, tdo_str, p, s);      // dK
      if (rc) return rc;
      return launch_bwd<4, 64, 2, 2, true, true, false>(tk_res, tv_res, tq_str, tdo_str, p, s);    // dV
    }
  }
}
