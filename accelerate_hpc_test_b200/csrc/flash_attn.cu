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
#include <stdlib.h>

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
  const uint32_t rp = smem_u32(tile) + row * 128;
#pragma unroll
  for (int c = 0; c < 8; ++c)
    sts_v4(rp + ((c ^ (row & 7)) << 4), f2_to_bf2(f[c * 8 + 0], f[c * 8 + 1]), f2_to_bf2(f[c * 8 + 2], f[c * 8 + 3]),
           f2_to_bf2(f[c * 8 + 4], f[c * 8 + 5]), f2_to_bf2(f[c * 8 + 6], f[c * 8 + 7]));
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
// Q + 2-stage K ring + 2-stage V ring + P = 112 KB; two CTAs per SM must fit in 228 KB together with the 1 KB the
// driver reserves per CTA, so there is no 1 KB alignment slack: the kernel checks the (in practice 1024-aligned)
// dynamic shared memory base instead.
static constexpr int FWD_SMEM = FWD_Q_BYTES + 4 * FWD_KV_BYTES + FWD_P_BYTES + 128;

// One CTA = one 128-row Q tile of one head; two CTAs per SM.  Per 64-row KV block j:
//   tensor:  S_j = Q K_j^T  (TMEM stage j&1, issued TWO blocks ahead),  O += P_j V_j
//   softmax: 128 threads, one query row each: online softmax with lazy rescale, P_j -> smem (bf16)
// S is double buffered in TMEM and K / V each have a 2-stage TMA ring, so S_{j+1} is already waiting when the
// softmax warps finish block j and the TMA latency of block j+2 is hidden behind a whole iteration.
__global__ void __launch_bounds__(192, 2)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const FaFwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + FWD_Q_BYTES;      // [2 stages]
  uint8_t* sV = sK + 2 * FWD_KV_BYTES;  // [2 stages]
  uint8_t* sP = sV + 2 * FWD_KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + FWD_P_BYTES);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* k_empty = bars + 3;   // [2]
  uint64_t* v_full = bars + 5;    // [2]
  uint64_t* v_empty = bars + 7;   // [2]
  uint64_t* s_full = bars + 9;    // [2]
  uint64_t* p_ready = bars + 11;  // softmax (128) -> MMA
  uint64_t* pv_done = bars + 12;  // PV_j retired: P buffer free, O stable
  uint64_t* o_full = bars + 13;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid ordered q-block-major with the heavy (late) causal blocks first: the hardware scheduler hands out CTAs
  // in index order, so this is longest-processing-time-first across ALL (head, batch) pairs, not per pair
  const int nqb = (p.S + FWD_BM - 1) / FWD_BM;
  const int qb = nqb - 1 - (int)(blockIdx.x / (unsigned)(p.Hq * p.B));
  const int hb = (int)(blockIdx.x % (unsigned)(p.Hq * p.B));
  const int hq = hb % p.Hq, b = hb / p.Hq;
  const int hkv = hq / (p.Hq / p.Hkv);
  const int q0 = qb * FWD_BM;
  const int kv_end = p.causal ? min(p.S, q0 + FWD_BM) : p.S;
  const int nblk = (kv_end + FWD_BN - 1) / FWD_BN;
  const int row_base = b * p.S;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) {
      printf("b200 flash_fwd: dynamic shared memory base is not 1024-byte aligned\n");
      __trap();
    }
    prefetch_tensormap(&tmap_q);
    prefetch_tensormap(&tmap_k);
    prefetch_tensormap(&tmap_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
    }
    mbar_init(p_ready, 128); mbar_init(pv_done, 1); mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 128;  // S stage s at tS + 64 s

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, FWD_Q_BYTES);
      tma_load_2d(sQ, &tmap_q, q_full, hq * FA_D, row_base + q0);
      tma_load_2d(sQ + FWD_Q_BYTES / 2, &tmap_q, q_full, hq * FA_D + 64, row_base + q0);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        const uint32_t par = ((j >> 1) & 1) ^ 1;
        const int kvrow = row_base + j * FWD_BN;
        uint8_t* k_dst = sK + st * FWD_KV_BYTES;
        uint8_t* v_dst = sV + st * FWD_KV_BYTES;
        mbar_wait(&k_empty[st], par);
        mbar_arrive_expect_tx(&k_full[st], FWD_KV_BYTES);
        tma_load_2d(k_dst, &tmap_k, &k_full[st], hkv * FA_D, kvrow);
        tma_load_2d(k_dst + FWD_KV_BYTES / 2, &tmap_k, &k_full[st], hkv * FA_D + 64, kvrow);
        mbar_wait(&v_empty[st], par);
        mbar_arrive_expect_tx(&v_full[st], FWD_KV_BYTES);
        tma_load_2d(v_dst, &tmap_v, &v_full[st], hkv * FA_D, kvrow);
        tma_load_2d(v_dst + FWD_KV_BYTES / 2, &tmap_v, &v_full[st], hkv * FA_D + 64, kvrow);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, FWD_BN, 0, 0);
      constexpr uint32_t idesc_o = make_idesc_bf16(128, FA_D, 0, 1);
      const uint32_t aQ = smem_u32(sQ), aP = smem_u32(sP);
      auto issue_S = [&](int j) {
        const int st = j & 1;
        const uint32_t aK = smem_u32(sK + st * FWD_KV_BYTES);
        mbar_wait(&k_full[st], (j >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < FA_D / 16; ++k) {
          const uint64_t da = make_umma_desc(aQ + (k >> 2) * (FWD_Q_BYTES / 2) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_umma_desc(aK + (k >> 2) * (FWD_KV_BYTES / 2) + (k & 3) * 32, 16, 1024);
          umma_bf16_ss(tS + st * 64, da, db, idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(&k_empty[st]);
        umma_commit(&s_full[st]);
      };
      mbar_wait(q_full, 0);
      issue_S(0);
      if (nblk > 1) issue_S(1);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        const uint32_t aV = smem_u32(sV + st * FWD_KV_BYTES);
        mbar_wait(p_ready, j & 1);
        mbar_wait(&v_full[st], (j >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < FWD_BN / 16; ++k) {
          const uint64_t da = make_umma_desc(aP + k * 32, 16, 1024);
          // V tile: kv rows x 64-wide d chunks -> MN-major B (N = d): LBO = chunk stride, SBO = 8 kv rows
          const uint64_t db = make_umma_desc(aV + k * 16 * 128, FWD_KV_BYTES / 2, 1024);
          umma_bf16_ss(tO, da, db, idesc_o, (j > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&v_empty[st]);
        umma_commit(pv_done);
        if (j == nblk - 1) umma_commit(o_full);
        if (j + 2 < nblk) issue_S(j + 2);  // the softmax warps finished reading S_j before p_ready(j)
      }
    }
  } else {
    // ---------------- softmax / correction / epilogue: one thread per query row ----------------
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const int q_idx = q0 + r;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    float m_used = -INFINITY, l = 0.f;  // running max (in scaled log2 units) and row sum
    const float sl2 = p.scale_log2;
    for (int j = 0; j < nblk; ++j) {
      const int kv0 = j * FWD_BN;
      const int st = j & 1;
      mbar_wait(&s_full[st], (j >> 1) & 1);
      tc_fence_after();
      float s[64];
      {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(tS + st * 64 + lane_addr, v0);
        tmem_ld_32x32b_x32(tS + st * 64 + lane_addr + 32, v1);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          s[i] = __uint_as_float(v0[i]);
          s[32 + i] = __uint_as_float(v1[i]);
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
      float mx0 = fmaxf(s[0], s[1]), mx1 = fmaxf(s[2], s[3]);
#pragma unroll
      for (int i = 4; i < 64; i += 4) {
        mx0 = fmaxf(mx0, fmaxf(s[i], s[i + 1]));
        mx1 = fmaxf(mx1, fmaxf(s[i + 2], s[i + 3]));
      }
      const float m_new = fmaxf(m_used, fmaxf(mx0, mx1) * sl2);  // scale > 0: max commutes with it
      float alpha = 1.f;
      const bool need = (m_new - m_used) > 8.f;  // also true for the first finite max (m_used = -inf)
      if (need) {
        alpha = exp2f(m_used - m_new);  // 0 when m_used == -inf
        m_used = m_new;
        l *= alpha;
      }
      const float neg_m = (m_used == -INFINITY) ? 0.f : -m_used;
      float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; i += 2) {
        s[i] = ex2_approx(fmaf(s[i], sl2, neg_m));
        s[i + 1] = ex2_approx(fmaf(s[i + 1], sl2, neg_m));
        rs0 += s[i];
        rs1 += s[i + 1];
      }
      l += rs0 + rs1;
      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1);  // PV_{j-1} retired: the P buffer is free and O is stable
        tc_fence_after();
      }
      store_row_sw128(sP, r, s);
      if (j > 0 && __any_sync(0xffffffffu, need)) {
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
  float* dq_acc;  // [B, Hq, S_pad, D] fp32, zero-initialised; S_pad = S rounded up to 64
  __nv_bfloat16* dk;
  __nv_bfloat16* dv;
  int B, S, S_pad, Hq, Hkv;
  long long dkv_stride;
  float scale, scale_log2;
  int causal;
  int dbg;
};

// One CTA owns a 128-row K/V block of one kv head and walks over the q blocks (64 rows) of its GQA group.
// Per q block:   A:  S^T = K Q^T, dP^T = V dO^T                         (tensor cores -> TMEM stage it&1)
//                    P^T = exp2(S^T*scale - lse), dS^T = P^T (dP^T - delta)   (8 softmax warps -> smem stage it&1)
//                B:  dQ^T = K^T dS^T (into the S^T columns of the stage), dV += P^T dO, dK += dS^T Q
//                    dQ tile: TMEM -> smem -> ONE bulk reduce-add into the fp32 accumulator  (4 readout warps)
// A(it+1) is issued before B(it), the TMEM S/dP stage is double buffered, Q/dO sit in a 3-deep TMA ring, and the dQ
// readout has its own warps, so the tensor cores run A(it+1) while the softmax warps work on `it`, and B(it) while
// the readout warps drain dQ(it-1).  P^T / dS^T have a single smem buffer: the softmax warps compute both in
// registers and only wait for B(it-1) to retire before the (short) stores.  dQ goes to global memory as a 32 KB cp.reduce.async.bulk (TMA engine) instead of 8192 scalar atomics
// per q block, which is what bounded the first version of this kernel (L2 atomic throughput).
static constexpr int BWD_BN = 128;  // kv rows per CTA
static constexpr int BWD_BM = 64;   // q rows per iteration
static constexpr int BWD_KV_BYTES = BWD_BN * FA_D * 2;   // 32 KB each for K and V
static constexpr int BWD_QDO_BYTES = BWD_BM * FA_D * 2;  // 16 KB each for Q_i and dO_i
static constexpr int BWD_STAGE_BYTES = 2 * BWD_QDO_BYTES;
static constexpr int BWD_PT_BYTES = BWD_BN * BWD_BM * 2;  // 16 KB
static constexpr int BWD_DQ_BYTES = BWD_BM * FA_D * 4;    // 32 KB fp32 staging tile [64 q][128 d]
static constexpr int BWD_THREADS = 512;
static constexpr int BWD_QSTAGES = 3;  // Q/dO ring: TMA latency spans more than one iteration
static constexpr int BWD_SMEM = 2 * BWD_KV_BYTES + BWD_QSTAGES * BWD_STAGE_BYTES + 2 * BWD_PT_BYTES + BWD_DQ_BYTES +
                                2 * 2 * BWD_BM * 4 + 256 + 1024;

// 32 bf16 values (half of a 128 B row) of a K-major SWIZZLE_128B tile: chunks 4*half .. 4*half+3
__device__ __forceinline__ void store_half_row_sw128(uint8_t* tile, int row, int half, const float (&f)[32]) {
  uint8_t* rp = tile + row * 128;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint4 o;
    o.x = f2_to_bf2(f[c * 8 + 0], f[c * 8 + 1]);
    o.y = f2_to_bf2(f[c * 8 + 2], f[c * 8 + 3]);
    o.z = f2_to_bf2(f[c * 8 + 4], f[c * 8 + 5]);
    o.w = f2_to_bf2(f[c * 8 + 6], f[c * 8 + 7]);
    *reinterpret_cast<uint4*>(rp + ((((half << 2) + c) ^ (row & 7)) << 4)) = o;
  }
}

__global__ void __launch_bounds__(BWD_THREADS, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_do,
                 const FaBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sK = smem;
  uint8_t* sV = sK + BWD_KV_BYTES;
  uint8_t* sQdO = sV + BWD_KV_BYTES;                      // [3 stages][Q_i 16 KB | dO_i 16 KB]
  uint8_t* sPt = sQdO + BWD_QSTAGES * BWD_STAGE_BYTES;    // P^T   [128 kv x 64 q]
  uint8_t* sdSt = sPt + BWD_PT_BYTES;                     // dS^T  [128 kv x 64 q]
  float* sDQ = reinterpret_cast<float*>(sdSt + BWD_PT_BYTES);  // [64 q][128 d] fp32
  float* sLse = sDQ + BWD_BM * FA_D;            // [2][64]
  float* sDelta = sLse + 2 * BWD_BM;            // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + 2 * BWD_BM);
  uint64_t* kv_full = bars + 0;
  uint64_t* qdo_full = bars + 1;    // [3] TMA -> MMA
  uint64_t* qdo_free = bars + 4;    // [3] B(it) retired: Q/dO stage it%3 may be refilled   MMA -> TMA
  uint64_t* s_full = bars + 7;      // [2] A(it) retired            MMA -> softmax
  uint64_t* pds_ready = bars + 9;   // [2] P^T / dS^T written       softmax (256) -> MMA
  uint64_t* dq_full = bars + 11;    // [2] dQ^T(it) retired         MMA -> readout
  uint64_t* dq_done = bars + 13;    // [2] dQ^T(it) is in registers: TMEM stage free       readout (128) -> MMA
  uint64_t* pds_free = bars + 15;   //     B(it) retired: the P^T / dS^T buffer may be overwritten  MMA -> softmax
  uint64_t* dkv_full = bars + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid, kv-block-major: kv block 0 sees every q block (most work) -> longest CTAs are scheduled first
  const int kvb = (int)(blockIdx.x / (unsigned)(p.Hkv * p.B));
  const int hb = (int)(blockIdx.x % (unsigned)(p.Hkv * p.B));
  const int hkv = hb % p.Hkv, b = hb / p.Hkv;
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
    for (int i = 0; i < BWD_QSTAGES; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_free[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&pds_ready[i], 256);
      mbar_init(&dq_full[i], 1); mbar_init(&dq_done[i], 128);
    }
    mbar_init(pds_free, 1);
    mbar_init(dkv_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tdK = tmem_base, tdV = tmem_base + 128;
  // stage s: S^T (later dQ^T) at 256 + 128 s, dP^T at 320 + 128 s

  if (warp == 0) {
    // ================================================= TMA producer
    if (elect_one()) {
      mbar_arrive_expect_tx(kv_full, 2 * BWD_KV_BYTES);
      tma_load_2d(sK, &tmap_k, kv_full, hkv * FA_D, row_base + kv0);
      tma_load_2d(sK + BWD_KV_BYTES / 2, &tmap_k, kv_full, hkv * FA_D + 64, row_base + kv0);
      tma_load_2d(sV, &tmap_v, kv_full, hkv * FA_D, row_base + kv0);
      tma_load_2d(sV + BWD_KV_BYTES / 2, &tmap_v, kv_full, hkv * FA_D + 64, row_base + kv0);
      int st = 0;
      uint32_t ph = 0;
      for (int it = 0; it < n_iter; ++it) {
        const int hq = hkv * G + it / per_head;
        const int qrow = row_base + (i_begin + it % per_head) * BWD_BM;
        mbar_wait(&qdo_free[st], ph ^ 1);  // B(it-3) has consumed this stage
        mbar_arrive_expect_tx(&qdo_full[st], BWD_STAGE_BYTES);
        uint8_t* sq = sQdO + st * BWD_STAGE_BYTES;
        uint8_t* sd = sq + BWD_QDO_BYTES;
        tma_load_2d(sq, &tmap_q, &qdo_full[st], hq * FA_D, qrow);
        tma_load_2d(sq + BWD_QDO_BYTES / 2, &tmap_q, &qdo_full[st], hq * FA_D + 64, qrow);
        tma_load_2d(sd, &tmap_do, &qdo_full[st], hq * FA_D, qrow);
        tma_load_2d(sd + BWD_QDO_BYTES / 2, &tmap_do, &qdo_full[st], hq * FA_D + 64, qrow);
        if (++st == BWD_QSTAGES) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================================================= MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc_st = make_idesc_bf16(128, BWD_BM, 0, 0);  // K Q^T / V dO^T  (N = 64)
      constexpr uint32_t idesc_dkv = make_idesc_bf16(128, FA_D, 0, 1);   // P^T dO / dS^T Q  (N = 128)
      constexpr uint32_t idesc_dq = make_idesc_bf16(128, BWD_BM, 1, 1);  // K^T dS^T        (N = 64)
      const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);
      auto issue_A = [&](int it) {
        const int st = it & 1, qs = it % BWD_QSTAGES;
        const uint32_t aQ = smem_u32(sQdO + qs * BWD_STAGE_BYTES);
        const uint32_t adO = aQ + BWD_QDO_BYTES;
        const uint32_t tSt = tmem_base + 256 + st * 128, tdPt = tSt + 64;
        mbar_wait(&qdo_full[qs], (it / BWD_QSTAGES) & 1);
        mbar_wait(&dq_done[st], ((it >> 1) & 1) ^ 1);  // dQ^T(it-2) left this TMEM stage
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
        umma_commit(&s_full[st]);
      };
      mbar_wait(kv_full, 0);
      if (n_iter > 0) issue_A(0);
      for (int it = 0; it < n_iter; ++it) {
        const int st = it & 1, qs = it % BWD_QSTAGES;
        if (it + 1 < n_iter) issue_A(it + 1);
        const uint32_t aQ = smem_u32(sQdO + qs * BWD_STAGE_BYTES);
        const uint32_t adO = aQ + BWD_QDO_BYTES;
        const uint32_t aPt = smem_u32(sPt), adSt = smem_u32(sdSt);
        const uint32_t tdQt = tmem_base + 256 + st * 128;
        mbar_wait(&pds_ready[st], (it >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BWD_BN / 16; ++k) {  // dQ^T[d, q] = sum_kv K[kv, d] dS^T[kv, q]  (first: its readout overlaps dV/dK)
          const uint64_t da = make_umma_desc(aK + k * 16 * 128, BWD_KV_BYTES / 2, 1024);
          const uint64_t db = make_umma_desc(adSt + k * 16 * 128, BWD_PT_BYTES, 1024);
          umma_bf16_ss(tdQt, da, db, idesc_dq, k > 0 ? 1u : 0u);
        }
        umma_commit(&dq_full[st]);
#pragma unroll
        for (int k = 0; k < BWD_BM / 16; ++k) {  // dV += P^T dO   (reduction over the q rows)
          const uint64_t dap = make_umma_desc(aPt + k * 32, 16, 1024);
          const uint64_t dbo = make_umma_desc(adO + k * 16 * 128, BWD_QDO_BYTES / 2, 1024);
          umma_bf16_ss(tdV, dap, dbo, idesc_dkv, (it > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < BWD_BM / 16; ++k) {  // dK += dS^T Q
          const uint64_t das = make_umma_desc(adSt + k * 32, 16, 1024);
          const uint64_t dbq = make_umma_desc(aQ + k * 16 * 128, BWD_QDO_BYTES / 2, 1024);
          umma_bf16_ss(tdK, das, dbq, idesc_dkv, (it > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&qdo_free[qs]);
        umma_commit(pds_free);
      }
      umma_commit(dkv_full);
    }
  } else if (warp >= 4 && warp < 12) {
    // ================================================= softmax / dS warps: 128 kv rows x two 32-column halves
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int r = quad * 32 + lane;     // kv row inside the block
    const int tid = threadIdx.x - 128;  // 0..255
    const int kv_idx = kv0 + r;
    const bool kv_ok = kv_idx < p.S;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    // lse / delta of a q block are staged through smem; each of the first 128 threads prefetches its value one
    // iteration ahead so the global-load latency is off the critical path (rows past S: lse = +inf -> p = 0)
    auto fetch_stat = [&](int it) -> float {
      if (tid >= 128 || it >= n_iter) return 0.f;
      const int hq = hkv * G + it / per_head;
      const int qi = (i_begin + it % per_head) * BWD_BM + (tid & 63);
      const long long idx = ((long long)b * p.Hq + hq) * p.S + qi;
      if (tid < 64) return (qi < p.S) ? p.lse[idx] * kLog2e : INFINITY;
      return (qi < p.S) ? p.delta[idx] : 0.f;
    };
    float stat = fetch_stat(0);
    const float sl2 = p.scale_log2;
    for (int it = 0; it < n_iter; ++it) {
      const int st = it & 1;
      const uint32_t ph = (it >> 1) & 1;
      const int q0 = (i_begin + it % per_head) * BWD_BM;
      if (tid < 64) sts_f32(smem_u32(sLse + st * BWD_BM + tid), stat);
      else if (tid < 128) sts_f32(smem_u32(sDelta + st * BWD_BM + (tid - 64)), stat);
      stat = fetch_stat(it + 1);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mbar_wait(&s_full[st], ph);
      tc_fence_after();
      const uint32_t tSt = tmem_base + 256 + st * 128 + half * 32, tdPt = tSt + 64;
      const uint32_t lse_a = smem_u32(sLse + st * BWD_BM + half * 32), del_a = smem_u32(sDelta + st * BWD_BM + half * 32);
      const int qc = q0 + half * 32;
      uint32_t pp[16], dd[16];  // P^T and dS^T of this thread's 32 q columns, packed bf16 pairs
      {
        uint32_t v[32], w[32];
        tmem_ld_32x32b_x32(tSt + lane_addr, v);
        tmem_ld_32x32b_x32(tdPt + lane_addr, w);
        tmem_ld_wait();
        // only tiles on the causal diagonal / past the end of the sequence need the per-element mask
        const bool masked = (p.causal && (q0 < kv0 + BWD_BN)) || (kv0 + BWD_BN > p.S);
        if (!masked) {
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 l4 = lds_f4(lse_a + i * 4), d4 = lds_f4(del_a + i * 4);
            const float a0 = ex2_approx(fmaf(__uint_as_float(v[i]), sl2, -l4.x));
            const float a1 = ex2_approx(fmaf(__uint_as_float(v[i + 1]), sl2, -l4.y));
            const float a2 = ex2_approx(fmaf(__uint_as_float(v[i + 2]), sl2, -l4.z));
            const float a3 = ex2_approx(fmaf(__uint_as_float(v[i + 3]), sl2, -l4.w));
            pp[i >> 1] = f2_to_bf2(a0, a1);
            pp[(i >> 1) + 1] = f2_to_bf2(a2, a3);
            dd[i >> 1] = f2_to_bf2(a0 * (__uint_as_float(w[i]) - d4.x), a1 * (__uint_as_float(w[i + 1]) - d4.y));
            dd[(i >> 1) + 1] = f2_to_bf2(a2 * (__uint_as_float(w[i + 2]) - d4.z), a3 * (__uint_as_float(w[i + 3]) - d4.w));
          }
        } else {
          const bool diag = p.causal && (q0 < kv0 + BWD_BN);
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 l4 = lds_f4(lse_a + i * 4), d4 = lds_f4(del_a + i * 4);
            const float lv[4] = {l4.x, l4.y, l4.z, l4.w}, dv4[4] = {d4.x, d4.y, d4.z, d4.w};
            float a[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              a[j] = ex2_approx(fmaf(__uint_as_float(v[i + j]), sl2, -lv[j]));
              if (!kv_ok || (diag && (qc + i + j < kv_idx))) a[j] = 0.f;
            }
            pp[i >> 1] = f2_to_bf2(a[0], a[1]);
            pp[(i >> 1) + 1] = f2_to_bf2(a[2], a[3]);
            dd[i >> 1] = f2_to_bf2(a[0] * (__uint_as_float(w[i]) - dv4[0]), a[1] * (__uint_as_float(w[i + 1]) - dv4[1]));
            dd[(i >> 1) + 1] = f2_to_bf2(a[2] * (__uint_as_float(w[i + 2]) - dv4[2]), a[3] * (__uint_as_float(w[i + 3]) - dv4[3]));
          }
        }
      }
      mbar_wait(pds_free, (it & 1) ^ 1);  // B(it-1) no longer reads the P^T / dS^T buffer
      {
        // half a 128-byte row of the K-major SWIZZLE_128B tiles: 16-byte chunks 4*half .. 4*half+3, chunk c at c ^ (r & 7)
        const uint32_t prow = smem_u32(sPt) + r * 128, drow = smem_u32(sdSt) + r * 128;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t o = (uint32_t)((((half << 2) + c) ^ (r & 7)) << 4);
          sts_v4(prow + o, pp[4 * c], pp[4 * c + 1], pp[4 * c + 2], pp[4 * c + 3]);
          sts_v4(drow + o, dd[4 * c], dd[4 * c + 1], dd[4 * c + 2], dd[4 * c + 3]);
        }
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(&pds_ready[st]);
    }
    // ---- dK / dV epilogue: this warp group's 64 of the 128 feature columns
    mbar_wait(dkv_full, 0);
    tc_fence_after();
    __nv_bfloat16* dkrow = p.dk + (long long)(row_base + kv_idx) * p.dkv_stride + hkv * FA_D;
    __nv_bfloat16* dvrow = p.dv + (long long)(row_base + kv_idx) * p.dkv_stride + hkv * FA_D;
#pragma unroll 1
    for (int c = half * 64; c < half * 64 + 64; c += 32) {
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
  } else if (warp >= 12) {
    // ================================================= dQ readout warps: TMEM -> smem tile -> bulk reduce-add
    const int quad = warp & 3;
    const int d = quad * 32 + lane;  // feature index = TMEM lane of dQ^T
    const int tid = threadIdx.x - 384;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    for (int it = 0; it < n_iter; ++it) {
      const int st = it & 1;
      const int hq = hkv * G + it / per_head;
      const int q0 = (i_begin + it % per_head) * BWD_BM;
      mbar_wait(&dq_full[st], (it >> 1) & 1);
      tc_fence_after();
      uint32_t v0[32], v1[32];
      const uint32_t tdQt = tmem_base + 256 + st * 128;
      tmem_ld_32x32b_x32(tdQt + lane_addr, v0);
      tmem_ld_32x32b_x32(tdQt + lane_addr + 32, v1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&dq_done[st]);
      if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // previous tile left the staging buffer
      asm volatile("bar.sync 2, 128;" ::: "memory");
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        sDQ[i * FA_D + d] = __uint_as_float(v0[i]);
        sDQ[(32 + i) * FA_D + d] = __uint_as_float(v1[i]);
      }
      fence_proxy_async();
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (tid == 0 && !(p.dbg & 1)) {
        float* dst = p.dq_acc + (((long long)b * p.Hq + hq) * p.S_pad + q0) * FA_D;
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst),
                     "r"(smem_u32(sDQ)), "n"(BWD_DQ_BYTES)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// dq[(b*S + s), h*D + d] (bf16, strided rows) = dq_acc[b, h, s, d] (fp32, S padded to S_pad) * scale
__global__ void __launch_bounds__(256)
dq_finalize_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dst, int B, int S, int S_pad, int Hq,
                   long long dst_stride, float scale) {
  constexpr int nvec = FA_D / 4;
  const long long total = (long long)B * S * Hq * nvec;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(idx % nvec);
    long long t = idx / nvec;
    const int h = (int)(t % Hq);
    t /= Hq;  // token index b*S + s
    const int bb = (int)(t / S), ss = (int)(t - (long long)bb * S);
    const float4 v = *reinterpret_cast<const float4*>(acc + (((long long)bb * Hq + h) * S_pad + ss) * FA_D + i * 4);
    uint2 o;
    o.x = f2_to_bf2(v.x * scale, v.y * scale);
    o.y = f2_to_bf2(v.z * scale, v.w * scale);
    *reinterpret_cast<uint2*>(dst + t * dst_stride + h * FA_D + i * 4) = o;
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
  const unsigned grid = (unsigned)((S + FWD_BM - 1) / FWD_BM) * Hq * B;
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
  p.S_pad = (S + BWD_BM - 1) / BWD_BM * BWD_BM;
  { const char* e = getenv("B200_FA_DBG"); p.dbg = e ? atoi(e) : 0; }
  p.dkv_stride = dkv_stride; p.scale = scale; p.scale_log2 = scale * kLog2e; p.causal = causal;
  const unsigned grid = (unsigned)((S + BWD_BN - 1) / BWD_BN) * Hkv * B;
  flash_bwd_kernel<<<grid, BWD_THREADS, BWD_SMEM, (cudaStream_t)stream>>>(tq, tk, tv, tdo, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_dq_finalize(const void* acc, void* dst, int B, int S, int Hq, long long dst_stride, float scale,
                              void* stream) {
  if ((long long)B * S == 0) return 0;
  const int S_pad = (S + BWD_BM - 1) / BWD_BM * BWD_BM;
  long long work = (long long)B * S * Hq * (FA_D / 4);
  long long g = (work + 255) / 256;
  long long cap = (long long)num_sms() * 16;
  if (g > cap) g = cap;
  dq_finalize_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>((const float*)acc, (__nv_bfloat16*)dst, B, S, S_pad, Hq,
                                                                    dst_stride, scale);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
