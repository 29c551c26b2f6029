// Flash attention forward/backward for sm_100a on tcgen05 + TMEM + TMA (head_dim 128, GQA, causal).
//
// the reference reaches attention through transformers'
// F.scaled_dot_product_attention (flash-attn2-style mma.sync or cuDNN chosen by torch); here both
// matmuls of every block run as tcgen05.mma with fp32 accumulators in TMEM, the softmax runs with
// one thread per query row straight out of TMEM (no shuffles), and P goes back to the tensor core
// through a 128B-swizzled smem tile.
//
// Forward: CTA = 128 query rows of one (batch, q-head); KV blocks of 64. Two CTAs per SM overlap
// each other's MMA and softmax phases. O accumulates in TMEM and is only rescaled (TMEM
// read-modify-write) when a row max grows by more than 2^8 ("lazy rescale").
//
// Backward: CTA = 128 KV rows of one (batch, kv-head), looping over the q-heads of the GQA group
// and over 64-row query blocks. dK and dV accumulate in TMEM over the whole loop; S^T, dP^T and
// dQ^T are per-iteration TMEM tiles; dQ is reduced with fp32 atomics into a scratch buffer.
#include "common.cuh"

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, long long inner, long long outer,
                      long long row_stride, int box_inner, int box_outer);

static constexpr int FA_D = 128;
static constexpr float kLog2e = 1.4426950408889634f;
static constexpr float kLn2 = 0.6931471805599453f;

__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
      "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]),
      "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]),
      "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}

// write 64 bf16 values (one 128 B row) of a K-major SWIZZLE_128B tile: 16 B chunk c lands at c ^ (row & 7)
__device__ __forceinline__ void store_row_sw128(uint8_t* tile, int row, const float (&f)[64]) {
  uint8_t* rp = tile + row * 128;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    uint4 o;
    o.x = f2_to_bf2(f[c * 8 + 0], f[c * 8 + 1]);
    o.y = f2_to_bf2(f[c * 8 + 2], f[c * 8 + 3]);
    o.z = f2_to_bf2(f[c * 8 + 4], f[c * 8 + 5]);
    o.w = f2_to_bf2(f[c * 8 + 6], f[c * 8 + 7]);
    *reinterpret_cast<uint4*>(rp + ((c ^ (row & 7)) << 4)) = o;
  }
}

struct FaFwdParams {
  __nv_bfloat16* o;
  float* lse;
  int B, S, Hq, Hkv;
  long long o_stride;
  float scale_log2;
  int causal;
};

static constexpr int FWD_BM = 128, FWD_BN = 64;
static constexpr int FWD_Q_BYTES = FWD_BM * FA_D * 2;   // 32 KB (2 chunks of 128 x 64)
static constexpr int FWD_KV_BYTES = FWD_BN * FA_D * 2;  // 16 KB (2 chunks of 64 x 64)
static constexpr int FWD_P_BYTES = FWD_BM * FWD_BN * 2;  // 16 KB
static constexpr int FWD_SMEM = FWD_Q_BYTES + 2 * FWD_KV_BYTES + FWD_P_BYTES + 1024 + 128;

__global__ void __launch_bounds__(192, 2)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const FaFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + FWD_Q_BYTES;
  uint8_t* sV = sK + FWD_KV_BYTES;
  uint8_t* sP = sV + FWD_KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + FWD_P_BYTES);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;
  uint64_t* k_empty = bars + 2;
  uint64_t* v_full = bars + 3;
  uint64_t* v_empty = bars + 4;
  uint64_t* s_full = bars + 5;
  uint64_t* p_ready = bars + 6;
  uint64_t* o_full = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qb = gridDim.x - 1 - blockIdx.x;  // heavy (late) causal blocks first
  const int hq = blockIdx.y, b = blockIdx.z;
  const int hkv = hq / (p.Hq / p.Hkv);
  const int q0 = qb * FWD_BM;
  const int kv_end = p.causal ? min(p.S, q0 + FWD_BM) : p.S;
  const int nblk = (kv_end + FWD_BN - 1) / FWD_BN;
  const int row_base = b * p.S;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_q);
    prefetch_tensormap(&tmap_k);
    prefetch_tensormap(&tmap_v);
    mbar_init(q_full, 1); mbar_init(k_full, 1); mbar_init(k_empty, 1); mbar_init(v_full, 1);
    mbar_init(v_empty, 1); mbar_init(s_full, 1); mbar_init(p_ready, 128); mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 64;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, FWD_Q_BYTES);
      tma_load_2d(sQ, &tmap_q, q_full, hq * FA_D, row_base + q0);
      tma_load_2d(sQ + FWD_Q_BYTES / 2, &tmap_q, q_full, hq * FA_D + 64, row_base + q0);
      for (int j = 0; j < nblk; ++j) {
        const uint32_t par = (j & 1) ^ 1;
        const int kvrow = row_base + j * FWD_BN;
        mbar_wait(k_empty, par);
        mbar_arrive_expect_tx(k_full, FWD_KV_BYTES);
        tma_load_2d(sK, &tmap_k, k_full, hkv * FA_D, kvrow);
        tma_load_2d(sK + FWD_KV_BYTES / 2, &tmap_k, k_full, hkv * FA_D + 64, kvrow);
        mbar_wait(v_empty, par);
        mbar_arrive_expect_tx(v_full, FWD_KV_BYTES);
        tma_load_2d(sV, &tmap_v, v_full, hkv * FA_D, kvrow);
        tma_load_2d(sV + FWD_KV_BYTES / 2, &tmap_v, v_full, hkv * FA_D + 64, kvrow);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, FWD_BN, 0, 0);
      constexpr uint32_t idesc_o = make_idesc_bf16(128, FA_D, 0, 1);
      const uint32_t aQ = smem_u32(sQ), aK = smem_u32(sK), aV = smem_u32(sV), aP = smem_u32(sP);
      mbar_wait(q_full, 0);
      for (int j = 0; j < nblk; ++j) {
        const uint32_t par = j & 1;
        mbar_wait(k_full, par);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < FA_D / 16; ++k) {
          const uint64_t da = make_umma_desc(aQ + (k >> 2) * (FWD_Q_BYTES / 2) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_umma_desc(aK + (k >> 2) * (FWD_KV_BYTES / 2) + (k & 3) * 32, 16, 1024);
          umma_bf16_ss(tS, da, db, idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(k_empty);
        umma_commit(s_full);
        mbar_wait(p_ready, par);
        mbar_wait(v_full, par);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < FWD_BN / 16; ++k) {
          const uint64_t da = make_umma_desc(aP + k * 32, 16, 1024);
          // V tile: kv rows x 64-wide d chunks -> MN-major B (N = d): LBO = chunk stride, SBO = 8 kv rows
          const uint64_t db = make_umma_desc(aV + k * 16 * 128, FWD_KV_BYTES / 2, 1024);
          umma_bf16_ss(tO, da, db, idesc_o, (j > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(v_empty);
        if (j == nblk - 1) umma_commit(o_full);
      }
    }
  } else {
    // ---------------- softmax / correction / epilogue: one thread per query row ----------------
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const int q_idx = q0 + r;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    float m_used = -INFINITY, l = 0.f;
    for (int j = 0; j < nblk; ++j) {
      const int kv0 = j * FWD_BN;
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      float s[64];
      {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(tS + lane_addr, v0);
        tmem_ld_32x32b_x32(tS + lane_addr + 32, v1);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          s[i] = __uint_as_float(v0[i]) * p.scale_log2;
          s[32 + i] = __uint_as_float(v1[i]) * p.scale_log2;
        }
      }
      const bool need_mask = (p.causal && (kv0 + FWD_BN - 1 > q0)) || (kv0 + FWD_BN > p.S);
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const int kv = kv0 + i;
          if (kv >= p.S || (p.causal && kv > q_idx)) s[i] = -INFINITY;
        }
      }
      float mx = s[0];
#pragma unroll
      for (int i = 1; i < 64; ++i) mx = fmaxf(mx, s[i]);
      const float m_new = fmaxf(m_used, mx);
      float alpha = 1.f;
      const bool need = (m_new - m_used) > 8.f;  // also true for the first finite max (m_used = -inf)
      if (need) {
        alpha = exp2f(m_used - m_new);  // 0 when m_used == -inf
        m_used = m_new;
        l *= alpha;
      }
      const float m_eff = (m_used == -INFINITY) ? 0.f : m_used;
      float rs = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        s[i] = exp2f(s[i] - m_eff);
        rs += s[i];
      }
      l += rs;
      store_row_sw128(sP, r, s);
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        // PV_{j-1} has retired (it was issued before S_j, whose completion we waited for)
#pragma unroll 1
        for (int c = 0; c < FA_D; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(tO + lane_addr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
          tmem_st_32x32b_x32(tO + lane_addr + c, v);
        }
        tmem_st_wait();
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    mbar_wait(o_full, 0);
    tc_fence_after();
    const float inv_l = (l > 0.f) ? 1.f / l : 0.f;
    const bool ok = q_idx < p.S;
    __nv_bfloat16* orow = p.o + (long long)(row_base + q_idx) * p.o_stride + hq * FA_D;
#pragma unroll 1
    for (int c = 0; c < FA_D; c += 32) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(tO + lane_addr + c, v);
      tmem_ld_wait();
      if (ok) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 o;
          o.x = f2_to_bf2(__uint_as_float(v[i]) * inv_l, __uint_as_float(v[i + 1]) * inv_l);
          o.y = f2_to_bf2(__uint_as_float(v[i + 2]) * inv_l, __uint_as_float(v[i + 3]) * inv_l);
          o.z = f2_to_bf2(__uint_as_float(v[i + 4]) * inv_l, __uint_as_float(v[i + 5]) * inv_l);
          o.w = f2_to_bf2(__uint_as_float(v[i + 6]) * inv_l, __uint_as_float(v[i + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c + i) = o;
        }
      }
    }
    if (ok) p.lse[((long long)b * p.Hq + hq) * p.S + q_idx] = m_used * kLn2 + __logf(l);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------ backward
// delta[b,h,s] = sum_d dO * O   (one warp per (token, head))
__global__ void __launch_bounds__(256)
flash_bwd_prep_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ out,
                      float* __restrict__ delta, int B, int S, int Hq, long long do_stride, long long o_stride) {
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long total = (long long)B * S * Hq;
  if (gw >= total) return;
  const int h = (int)(gw % Hq);
  const long long tok = gw / Hq;
  const int b = (int)(tok / S), s = (int)(tok % S);
  const uint2 a = *reinterpret_cast<const uint2*>(dout + tok * do_stride + h * FA_D + lane * 4);
  const uint2 c = *reinterpret_cast<const uint2*>(out + tok * o_stride + h * FA_D + lane * 4);
  float2 a0 = bf2_to_f2(a.x), a1 = bf2_to_f2(a.y), c0 = bf2_to_f2(c.x), c1 = bf2_to_f2(c.y);
  float v = a0.x * c0.x + a0.y * c0.y + a1.x * c1.x + a1.y * c1.y;
  v = warp_sum(v);
  if (lane == 0) delta[((long long)b * Hq + h) * S + s] = v;
}

struct FaBwdParams {
  const float* lse;
  const float* delta;
  float* dq_acc;  // [B*S, Hq*D] fp32, zero-initialised
  __nv_bfloat16* dk;
  __nv_bfloat16* dv;
  int B, S, Hq, Hkv;
  long long dkv_stride;
  float scale, scale_log2;
  int causal;
};

static constexpr int BWD_BN = 128;  // kv rows per CTA
static constexpr int BWD_BM = 64;   // q rows per iteration
static constexpr int BWD_KV_BYTES = BWD_BN * FA_D * 2;   // 32 KB each for K and V
static constexpr int BWD_QDO_BYTES = BWD_BM * FA_D * 2;  // 16 KB each for Q_i and dO_i
static constexpr int BWD_STAGE_BYTES = 2 * BWD_QDO_BYTES;
static constexpr int BWD_PT_BYTES = BWD_BN * BWD_BM * 2;  // 16 KB
static constexpr int BWD_SMEM =
    2 * BWD_KV_BYTES + 2 * BWD_STAGE_BYTES + 2 * BWD_PT_BYTES + 2 * 2 * BWD_BM * 4 + 1024 + 128;

__global__ void __launch_bounds__(192, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_do,
                 const FaBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sK = smem;
  uint8_t* sV = sK + BWD_KV_BYTES;
  uint8_t* sQdO = sV + BWD_KV_BYTES;           // [2 stages][Q_i 16 KB | dO_i 16 KB]
  uint8_t* sPt = sQdO + 2 * BWD_STAGE_BYTES;   // P^T   [128 kv x 64 q]
  uint8_t* sdSt = sPt + BWD_PT_BYTES;          // dS^T  [128 kv x 64 q]
  float* sLse = reinterpret_cast<float*>(sdSt + BWD_PT_BYTES);  // [2][64]
  float* sDelta = sLse + 2 * BWD_BM;                             // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + 2 * BWD_BM);
  uint64_t* kv_full = bars + 0;
  uint64_t* qdo_full = bars + 1;   // [2]
  uint64_t* qdo_empty = bars + 3;  // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* pds_ready = bars + 6;
  uint64_t* dq_full = bars + 7;
  uint64_t* dkv_full = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kvb = blockIdx.x, hkv = blockIdx.y, b = blockIdx.z;
  const int G = p.Hq / p.Hkv;
  const int kv0 = kvb * BWD_BN;
  const int row_base = b * p.S;
  const int nq = (p.S + BWD_BM - 1) / BWD_BM;
  const int i_begin = p.causal ? (kv0 / BWD_BM) : 0;
  const int per_head = nq - i_begin;
  const int n_iter = per_head * G;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_q); prefetch_tensormap(&tmap_k);
    prefetch_tensormap(&tmap_v); prefetch_tensormap(&tmap_do);
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 1); }
    mbar_init(s_full, 1); mbar_init(pds_ready, 128); mbar_init(dq_full, 1); mbar_init(dkv_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tdK = tmem_base, tdV = tmem_base + 128, tSt = tmem_base + 256, tdPt = tmem_base + 320,
                 tdQt = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(kv_full, 2 * BWD_KV_BYTES);
      tma_load_2d(sK, &tmap_k, kv_full, hkv * FA_D, row_base + kv0);
      tma_load_2d(sK + BWD_KV_BYTES / 2, &tmap_k, kv_full, hkv * FA_D + 64, row_base + kv0);
      tma_load_2d(sV, &tmap_v, kv_full, hkv * FA_D, row_base + kv0);
      tma_load_2d(sV + BWD_KV_BYTES / 2, &tmap_v, kv_full, hkv * FA_D + 64, row_base + kv0);
      for (int it = 0; it < n_iter; ++it) {
        const int st = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        const int hq = hkv * G + it / per_head;
        const int qrow = row_base + (i_begin + it % per_head) * BWD_BM;
        mbar_wait(&qdo_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&qdo_full[st], BWD_STAGE_BYTES);
        uint8_t* sq = sQdO + st * BWD_STAGE_BYTES;
        uint8_t* sd = sq + BWD_QDO_BYTES;
        tma_load_2d(sq, &tmap_q, &qdo_full[st], hq * FA_D, qrow);
        tma_load_2d(sq + BWD_QDO_BYTES / 2, &tmap_q, &qdo_full[st], hq * FA_D + 64, qrow);
        tma_load_2d(sd, &tmap_do, &qdo_full[st], hq * FA_D, qrow);
        tma_load_2d(sd + BWD_QDO_BYTES / 2, &tmap_do, &qdo_full[st], hq * FA_D + 64, qrow);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_st = make_idesc_bf16(128, BWD_BM, 0, 0);  // K Q^T / V dO^T  (N = 64)
      constexpr uint32_t idesc_dkv = make_idesc_bf16(128, FA_D, 0, 1);   // P^T dO / dS^T Q  (N = 128)
      constexpr uint32_t idesc_dq = make_idesc_bf16(128, BWD_BM, 1, 1);  // K^T dS^T        (N = 64)
      const uint32_t aK = smem_u32(sK), aV = smem_u32(sV), aPt = smem_u32(sPt), adSt = smem_u32(sdSt);
      mbar_wait(kv_full, 0);
      for (int it = 0; it < n_iter; ++it) {
        const int st = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        const uint32_t aQ = smem_u32(sQdO + st * BWD_STAGE_BYTES);
        const uint32_t adO = aQ + BWD_QDO_BYTES;
        mbar_wait(&qdo_full[st], ph);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < FA_D / 16; ++k) {  // S^T = K Q^T
          const uint64_t da = make_umma_desc(aK + (k >> 2) * (BWD_KV_BYTES / 2) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_umma_desc(aQ + (k >> 2) * (BWD_QDO_BYTES / 2) + (k & 3) * 32, 16, 1024);
          umma_bf16_ss(tSt, da, db, idesc_st, k > 0 ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < FA_D / 16; ++k) {  // dP^T = V dO^T
          const uint64_t da = make_umma_desc(aV + (k >> 2) * (BWD_KV_BYTES / 2) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_umma_desc(adO + (k >> 2) * (BWD_QDO_BYTES / 2) + (k & 3) * 32, 16, 1024);
          umma_bf16_ss(tdPt, da, db, idesc_st, k > 0 ? 1u : 0u);
        }
        umma_commit(s_full);
        mbar_wait(pds_ready, it & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BWD_BM / 16; ++k) {  // dV += P^T dO ; dK += dS^T Q   (reduction over q rows)
          const uint64_t dap = make_umma_desc(aPt + k * 32, 16, 1024);
          const uint64_t dbo = make_umma_desc(adO + k * 16 * 128, BWD_QDO_BYTES / 2, 1024);
          umma_bf16_ss(tdV, dap, dbo, idesc_dkv, (it > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < BWD_BM / 16; ++k) {
          const uint64_t das = make_umma_desc(adSt + k * 32, 16, 1024);
          const uint64_t dbq = make_umma_desc(aQ + k * 16 * 128, BWD_QDO_BYTES / 2, 1024);
          umma_bf16_ss(tdK, das, dbq, idesc_dkv, (it > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < BWD_BN / 16; ++k) {  // dQ^T[d, q] = sum_kv K[kv, d] dS^T[kv, q]
          const uint64_t da = make_umma_desc(aK + k * 16 * 128, BWD_KV_BYTES / 2, 1024);
          const uint64_t db = make_umma_desc(adSt + k * 16 * 128, BWD_PT_BYTES, 1024);
          umma_bf16_ss(tdQt, da, db, idesc_dq, k > 0 ? 1u : 0u);
        }
        umma_commit(&qdo_empty[st]);
        umma_commit(dq_full);
      }
      umma_commit(dkv_full);
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;  // kv row inside the block, and d index for the dQ^T readout
    const int tid = threadIdx.x - 64;  // 0..127
    const int kv_idx = kv0 + r;
    const bool kv_ok = kv_idx < p.S;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    for (int it = 0; it < n_iter; ++it) {
      const int st = it & 1;
      const int hq = hkv * G + it / per_head;
      const int q0 = (i_begin + it % per_head) * BWD_BM;
      // stage lse / delta of this q block (OOB rows get lse = +inf -> p = 0)
      {
        const int qi = q0 + (tid & 63);
        const long long idx = ((long long)b * p.Hq + hq) * p.S + qi;
        if (tid < 64) sLse[st * BWD_BM + tid] = (qi < p.S) ? p.lse[idx] * kLog2e : INFINITY;
        else sDelta[st * BWD_BM + (tid - 64)] = (qi < p.S) ? p.delta[idx] : 0.f;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      float pr[64];
      {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(tSt + lane_addr, v0);
        tmem_ld_32x32b_x32(tSt + lane_addr + 32, v1);
        tmem_ld_wait();
        const bool diag = p.causal && (q0 < kv0 + BWD_BN);  // some (q, kv) pairs of this tile are masked
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float a = exp2f(__uint_as_float(v0[i]) * p.scale_log2 - sLse[st * BWD_BM + i]);
          float c = exp2f(__uint_as_float(v1[i]) * p.scale_log2 - sLse[st * BWD_BM + 32 + i]);
          if (!kv_ok || (diag && (q0 + i < kv_idx))) a = 0.f;
          if (!kv_ok || (diag && (q0 + 32 + i < kv_idx))) c = 0.f;
          pr[i] = a;
          pr[32 + i] = c;
        }
      }
      store_row_sw128(sPt, r, pr);
      {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(tdPt + lane_addr, v0);
        tmem_ld_32x32b_x32(tdPt + lane_addr + 32, v1);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          pr[i] = pr[i] * (__uint_as_float(v0[i]) - sDelta[st * BWD_BM + i]);
          pr[32 + i] = pr[32 + i] * (__uint_as_float(v1[i]) - sDelta[st * BWD_BM + 32 + i]);
        }
      }
      store_row_sw128(sdSt, r, pr);
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(pds_ready);
      // dQ^T readout: thread r owns feature d = r, columns are the 64 q rows of this block
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      float* dq_base = p.dq_acc + (long long)(row_base + q0) * (p.Hq * FA_D) + hq * FA_D + r;
#pragma unroll 1
      for (int c = 0; c < BWD_BM; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tdQt + lane_addr + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (q0 + c + i < p.S) atomicAdd(dq_base + (long long)(c + i) * (p.Hq * FA_D), __uint_as_float(v[i]));
      }
      tc_fence_before();
      asm volatile("bar.sync 1, 128;" ::: "memory");  // lse/delta stage and dQ^T tile are free again
    }
    mbar_wait(dkv_full, 0);
    tc_fence_after();
    __nv_bfloat16* dkrow = p.dk + (long long)(row_base + kv_idx) * p.dkv_stride + hkv * FA_D;
    __nv_bfloat16* dvrow = p.dv + (long long)(row_base + kv_idx) * p.dkv_stride + hkv * FA_D;
#pragma unroll 1
    for (int c = 0; c < FA_D; c += 32) {
      uint32_t vk[32], vv[32];
      tmem_ld_32x32b_x32(tdK + lane_addr + c, vk);
      tmem_ld_32x32b_x32(tdV + lane_addr + c, vv);
      tmem_ld_wait();
      if (kv_ok) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 ok, ov;
          ok.x = f2_to_bf2(__uint_as_float(vk[i]) * p.scale, __uint_as_float(vk[i + 1]) * p.scale);
          ok.y = f2_to_bf2(__uint_as_float(vk[i + 2]) * p.scale, __uint_as_float(vk[i + 3]) * p.scale);
          ok.z = f2_to_bf2(__uint_as_float(vk[i + 4]) * p.scale, __uint_as_float(vk[i + 5]) * p.scale);
          ok.w = f2_to_bf2(__uint_as_float(vk[i + 6]) * p.scale, __uint_as_float(vk[i + 7]) * p.scale);
          ov.x = f2_to_bf2(__uint_as_float(vv[i]), __uint_as_float(vv[i + 1]));
          ov.y = f2_to_bf2(__uint_as_float(vv[i + 2]), __uint_as_float(vv[i + 3]));
          ov.z = f2_to_bf2(__uint_as_float(vv[i + 4]), __uint_as_float(vv[i + 5]));
          ov.w = f2_to_bf2(__uint_as_float(vv[i + 6]), __uint_as_float(vv[i + 7]));
          *reinterpret_cast<uint4*>(dkrow + c + i) = ok;
          *reinterpret_cast<uint4*>(dvrow + c + i) = ov;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// dq (bf16, strided) = dq_acc (fp32, dense [T, width]) * scale
__global__ void __launch_bounds__(256)
dq_finalize_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dst, long long T, int width,
                   long long dst_stride, float scale) {
  const int nvec = width >> 2;
  const long long total = T * nvec;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / nvec;
    const int i = (int)(idx - t * nvec);
    const float4 v = *reinterpret_cast<const float4*>(acc + t * width + i * 4);
    uint2 o;
    o.x = f2_to_bf2(v.x * scale, v.y * scale);
    o.y = f2_to_bf2(v.z * scale, v.w * scale);
    *reinterpret_cast<uint2*>(dst + t * dst_stride + i * 4) = o;
  }
}

int num_sms();

}  // namespace b200

using namespace b200;

B200_API int b200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, void* lse, int B, int S,
                                 int Hq, int Hkv, int D, long long q_stride, long long k_stride,
                                 long long v_stride, long long o_stride, float scale, int causal, void* stream) {
  if (D != FA_D || Hq % Hkv != 0 || (o_stride % 8)) return B200_ERR_UNSUPPORTED;
  if (B * S == 0) return 0;
  CUtensorMap tq, tk, tv;
  int rc;
  const long long T = (long long)B * S;
  if ((rc = make_tmap_bf16_2d(&tq, q, (long long)Hq * D, T, q_stride, 64, FWD_BM))) return rc;
  if ((rc = make_tmap_bf16_2d(&tk, k, (long long)Hkv * D, T, k_stride, 64, FWD_BN))) return rc;
  if ((rc = make_tmap_bf16_2d(&tv, v, (long long)Hkv * D, T, v_stride, 64, FWD_BN))) return rc;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(flash_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    attr = true;
  }
  FaFwdParams p;
  p.o = (__nv_bfloat16*)o; p.lse = (float*)lse; p.B = B; p.S = S; p.Hq = Hq; p.Hkv = Hkv;
  p.o_stride = o_stride; p.scale_log2 = scale * kLog2e; p.causal = causal;
  dim3 grid((S + FWD_BM - 1) / FWD_BM, Hq, B);
  flash_fwd_kernel<<<grid, 192, FWD_SMEM, (cudaStream_t)stream>>>(tq, tk, tv, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_flash_attn_bwd_prep(const void* dout, const void* out, void* delta, int B, int S, int Hq,
                                      long long do_stride, long long o_stride, void* stream) {
  const long long warps = (long long)B * S * Hq;
  if (warps == 0) return 0;
  flash_bwd_prep_kernel<<<ceil_div(warps * 32, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dout, (const __nv_bfloat16*)out, (float*)delta, B, S, Hq, do_stride, o_stride);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// dq_acc: fp32 [B*S, Hq*D] zero-initialised scratch (unscaled dQ lands there); dk/dv are bf16 outputs
B200_API int b200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* dout, const void* lse,
                                 const void* delta, void* dq_acc, void* dk, void* dv, int B, int S, int Hq,
                                 int Hkv, int D, long long q_stride, long long k_stride, long long v_stride,
                                 long long do_stride, long long dkv_stride, float scale, int causal,
                                 void* stream) {
  if (D != FA_D || Hq % Hkv != 0 || (dkv_stride % 8)) return B200_ERR_UNSUPPORTED;
  if (B * S == 0) return 0;
  CUtensorMap tq, tk, tv, tdo;
  int rc;
  const long long T = (long long)B * S;
  if ((rc = make_tmap_bf16_2d(&tq, q, (long long)Hq * D, T, q_stride, 64, BWD_BM))) return rc;
  if ((rc = make_tmap_bf16_2d(&tk, k, (long long)Hkv * D, T, k_stride, 64, BWD_BN))) return rc;
  if ((rc = make_tmap_bf16_2d(&tv, v, (long long)Hkv * D, T, v_stride, 64, BWD_BN))) return rc;
  if ((rc = make_tmap_bf16_2d(&tdo, dout, (long long)Hq * D, T, do_stride, 64, BWD_BM))) return rc;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(flash_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM));
    attr = true;
  }
  FaBwdParams p;
  p.lse = (const float*)lse; p.delta = (const float*)delta; p.dq_acc = (float*)dq_acc;
  p.dk = (__nv_bfloat16*)dk; p.dv = (__nv_bfloat16*)dv; p.B = B; p.S = S; p.Hq = Hq; p.Hkv = Hkv;
  p.dkv_stride = dkv_stride; p.scale = scale; p.scale_log2 = scale * kLog2e; p.causal = causal;
  dim3 grid((S + BWD_BN - 1) / BWD_BN, Hkv, B);
  flash_bwd_kernel<<<grid, 192, BWD_SMEM, (cudaStream_t)stream>>>(tq, tk, tv, tdo, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_dq_finalize(const void* acc, void* dst, long long T, int width, long long dst_stride,
                              float scale, void* stream) {
  if (width % 4) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  long long work = T * (width / 4);
  long long g = (work + 255) / 256;
  long long cap = (long long)num_sms() * 16;
  if (g > cap) g = cap;
  dq_finalize_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>((const float*)acc, (__nv_bfloat16*)dst, T,
                                                                    width, dst_stride, scale);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
