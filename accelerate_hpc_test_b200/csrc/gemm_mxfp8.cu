// MXFP8 (OCP microscaling: e4m3 elements, one UE8M0 scale per 32 elements along K) GEMM and quantisers for sm_100a.
//
//   D[M,N] (+)= sum_k (A_q[m,k] * 2^(sa[m,k/32]-127)) * (B_q[n,k] * 2^(sb[n,k/32]-127))   (+ bias[n])
//
// The scales are applied by the tensor core itself: tcgen05.mma.kind::mxf8f6f4.block_scale reads them from TMEM, where
// tcgen05.cp puts them from shared memory each k-block.  This is the K3 row of SURVEY.md §2.6 — the reference reaches
// it through TransformerEngine's MXFP8BlockScaling recipe + cuBLASLt (reference src/accelerate/utils/transformer_engine.py
// :157-177); here it is one warp-specialised persistent kernel:
//   tile 128 x 128 x 128, 6 TMA stages (A 16 KB + B 16 KB + SFA 512 B + SFB 512 B), 2 TMEM accumulator stages,
//   w0 TMA producer, w1 MMA issuer (UTCCP scale copies + 4 MMAs of K = 32 per stage), w2 TMEM allocator, w4-7 epilogue.
// Scale-factor storage (global and shared): 512-byte blocks, one per (128 rows, 128 k): byte (r % 32) * 16 + (r / 32) * 4 + g
// holds the scale of row r, k-group g — the layout tcgen05.cp.32x128b.warpx4 turns into "lane r % 32, column r / 32".
#include "cluster2.cuh"
#include <cuda_fp8.h>

namespace b200 {

int make_tmap_u8_2d(CUtensorMap* out, const void* base, long long inner, long long outer, long long row_stride,
                    int box_inner, int box_outer);
int make_tmap_sf_blocks(CUtensorMap* out, const void* base, long long nblocks);
int num_sms();

namespace mx {

static constexpr int BM = 128, BN = 128, BK = 128;
static constexpr int STAGES = 6;
static constexpr int A_BYTES = BM * BK;  // 16 KB
static constexpr int B_BYTES = BN * BK;  // 16 KB
static constexpr int SF_BYTES = 512;
static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + 2 * SF_BYTES;
static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 + 256;
static constexpr int TMEM_SF_BASE = 2 * BN;  // columns 256.. : 8 per stage (SFA 4 + SFB 4)
static constexpr int GROUP_M = 8;

struct Params {
  void* C;
  const void* bias;
  const uint8_t* sfa;  // [ceil(M/128)][K/128][512]
  const uint8_t* sfb;  // [ceil(N/128)][K/128][512]
  int M, N, K;
  long long ldc;
  int out_fp32, accumulate;
  int num_m_blocks, num_n_blocks;
};

// instruction descriptor of kind::mxf8f6f4.block_scale: e4m3 x e4m3, UE8M0 scales, both K-major
__device__ __forceinline__ uint32_t idesc_mx(int M, int N, uint32_t sf_id) {
  return (sf_id << 4)                  // b_sf_id: which of the 4 bytes of the scale word this MMA uses
         | (0u << 7) | (0u << 10)      // a_format = b_format = E4M3
         | ((uint32_t)(N >> 3) << 17)  // n_dim
         | (1u << 23)                  // scale_format = E8M0
         | ((uint32_t)(M >> 4) << 24)  // m_dim
         | (sf_id << 29);              // a_sf_id
}
// shared memory descriptor of one 32 x 16 B scale block (no swizzle: 8-row core matrices 128 B apart)
__device__ __forceinline__ uint64_t sf_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(16 >> 4) << 16;   // leading byte offset (one 16 B chunk per row: unused)
  d |= (uint64_t)(128 >> 4) << 32;  // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ void utccp_32x128b_warpx4(uint32_t tmem_dst, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(tmem_dst), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_mxf8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc,
                                          uint32_t tsfa, uint32_t tsfb) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(tsfa), "r"(tsfb)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int per_group = GROUP_M * num_n;
  const int group = tile / per_group;
  const int first_m = group * GROUP_M;
  const int gm = min(GROUP_M, num_m - first_m);
  const int r = tile - group * per_group;
  m_blk = first_m + (r % gm);
  n_blk = r / gm;
}

__global__ void __launch_bounds__(256, 1)
gemm_mxfp8_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + STAGES * A_BYTES;
  uint8_t* smem_sf = smem_b + STAGES * B_BYTES;  // [STAGES][SFA 512 | SFB 512]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_sf + STAGES * 2 * SF_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_k_blocks = p.K / BK;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
        const uint8_t* sfa = p.sfa + (long long)m_blk * num_k_blocks * SF_BYTES;
        const uint8_t* sfb = p.sfb + (long long)n_blk * num_k_blocks * SF_BYTES;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
          tma_load_2d(smem_a + stage * A_BYTES, &tmap_a, &full_bar[stage], kb * BK, m_blk * BM);
          tma_load_2d(smem_b + stage * B_BYTES, &tmap_b, &full_bar[stage], kb * BK, n_blk * BN);
          bulk_g2s(smem_sf + stage * 2 * SF_BYTES, sfa + (long long)kb * SF_BYTES, SF_BYTES, &full_bar[stage]);
          bulk_g2s(smem_sf + stage * 2 * SF_BYTES + SF_BYTES, sfb + (long long)kb * SF_BYTES, SF_BYTES, &full_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * A_BYTES);
          const uint32_t sb = smem_u32(smem_b + stage * B_BYTES);
          const uint32_t ssf = smem_u32(smem_sf + stage * 2 * SF_BYTES);
          const uint32_t tsfa = tmem_base + TMEM_SF_BASE + stage * 8, tsfb = tsfa + 4;
          utccp_32x128b_warpx4(tsfa, sf_desc(ssf));
          utccp_32x128b_warpx4(tsfb, sf_desc(ssf + SF_BYTES));
#pragma unroll
          for (int k = 0; k < BK / 32; ++k) {
            const uint64_t da = make_umma_desc(sa + k * 32, 16, 1024);
            const uint64_t db = make_umma_desc(sb + k * 32, 16, 1024);
            umma_mxf8(d_tmem, da, db, idesc_mx(BM, BN, (uint32_t)k), (kb > 0 || k > 0) ? 1u : 0u, tsfa, tsfb);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == num_k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    const int e = warp - 4;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
      const int row = m_blk * BM + e * 32 + lane;
      const int n0 = n_blk * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + acc * BN;
      const bool row_ok = row < p.M;
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(taddr + c, v);
        tmem_ld_wait();
        const int col0 = n0 + c;
        if (row_ok && col0 < p.N) {
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
          if (p.bias != nullptr) {
            const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(p.bias);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col0 + i < p.N) f[i] += __bfloat162float(__ldg(bias + col0 + i));
          }
          if (p.out_fp32) {
            float* crow = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col0;
            for (int i = 0; i < 32; ++i)
              if (col0 + i < p.N) crow[i] = p.accumulate ? crow[i] + f[i] : f[i];
          } else {
            __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col0;
            if (col0 + 32 <= p.N && (p.ldc % 8 == 0)) {
#pragma unroll
              for (int i = 0; i < 32; i += 8) {
                if (p.accumulate) {
                  uint4 old = *reinterpret_cast<uint4*>(crow + i);
                  float2 a0 = bf2_to_f2(old.x), a1 = bf2_to_f2(old.y), a2 = bf2_to_f2(old.z), a3 = bf2_to_f2(old.w);
                  f[i] += a0.x; f[i + 1] += a0.y; f[i + 2] += a1.x; f[i + 3] += a1.y;
                  f[i + 4] += a2.x; f[i + 5] += a2.y; f[i + 6] += a3.x; f[i + 7] += a3.y;
                }
                uint4 o;
                o.x = f2_to_bf2(f[i], f[i + 1]); o.y = f2_to_bf2(f[i + 2], f[i + 3]);
                o.z = f2_to_bf2(f[i + 4], f[i + 5]); o.w = f2_to_bf2(f[i + 6], f[i + 7]);
                *reinterpret_cast<uint4*>(crow + i) = o;
              }
            } else {
              for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) {
                  float o = f[i];
                  if (p.accumulate) o += __bfloat162float(crow[i]);
                  crow[i] = __float2bfloat16(o);
                }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ quantisers
// shared exponent of a 32-element group (OCP MX): floor(log2(amax)) - emax(e4m3 = 8), stored biased by 127
__device__ __forceinline__ uint32_t mx_biased_exp(float amax) {
  const uint32_t bits = __float_as_uint(amax);
  int e = (int)((bits >> 23) & 0xFF) - 8;  // biased exponent of amax minus emax
  if (amax == 0.f) e = 0;
  e = e < 0 ? 0 : (e > 254 ? 254 : e);
  return (uint32_t)e;
}
__device__ __forceinline__ float mx_inv_scale(uint32_t biased) {  // 2^-(biased - 127)
  return __uint_as_float((uint32_t)(254 - biased) << 23);
}
__device__ __forceinline__ uint8_t f_to_e4m3(float v) {
  return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
}
__device__ __forceinline__ long long sf_offset(long long r, long long g, long long kblocks) {
  return ((r >> 7) * kblocks + (g >> 2)) * 512 + (r & 31) * 16 + ((r >> 5) & 3) * 4 + (g & 3);
}

// x [R, C] bf16 (row stride ldx) -> q [R, C] e4m3 (row stride ldq), scales along C.
// One thread per 32-element group: a warp reads 2 KB of one row contiguously (64 B per lane), every lane owns the
// scale of its group, and each quad of lanes merges its four scale bytes into one 32-bit store.
__global__ void __launch_bounds__(256)
quantize_mx_rowwise_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int R, int C, uint8_t* __restrict__ q,
                           long long ldq, uint8_t* __restrict__ sf) {
  const int groups = C >> 5;  // multiple of 4
  const long long total = (long long)R * groups;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx / groups), g = (int)(idx - (long long)r * groups);
    const uint4* src = reinterpret_cast<const uint4*>(x + (long long)r * ldx + g * 32);
    float v[32];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint4 t = ld_nc_v4(src + i);
      const float2 a = bf2_to_f2(t.x), b = bf2_to_f2(t.y), c = bf2_to_f2(t.z), d = bf2_to_f2(t.w);
      v[i * 8 + 0] = a.x; v[i * 8 + 1] = a.y; v[i * 8 + 2] = b.x; v[i * 8 + 3] = b.y;
      v[i * 8 + 4] = c.x; v[i * 8 + 5] = c.y; v[i * 8 + 6] = d.x; v[i * 8 + 7] = d.y;
    }
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) amax = fmaxf(amax, fabsf(v[i]));
    const uint32_t be = mx_biased_exp(amax);
    const float inv = mx_inv_scale(be);
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      w[i] = (uint32_t)f_to_e4m3(v[4 * i] * inv) | ((uint32_t)f_to_e4m3(v[4 * i + 1] * inv) << 8) |
             ((uint32_t)f_to_e4m3(v[4 * i + 2] * inv) << 16) | ((uint32_t)f_to_e4m3(v[4 * i + 3] * inv) << 24);
    uint4* dst = reinterpret_cast<uint4*>(q + (long long)r * ldq + g * 32);
    dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
    dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
    // the four groups of one 128-column block sit in four consecutive lanes (groups % 4 == 0, 256 % 4 == 0)
    uint32_t sfw = be << (8 * (g & 3));
    sfw |= __shfl_xor_sync(0xffffffffu, sfw, 1);
    sfw |= __shfl_xor_sync(0xffffffffu, sfw, 2);
    if ((g & 3) == 0) *reinterpret_cast<uint32_t*>(sf + sf_offset(r, g, groups >> 2)) = sfw;
  }
}

// x [R, C] bf16 -> qt [C, R] e4m3 (row stride ldqt), scales along R (groups of 32 consecutive x-rows).
// CTA tile: 128 x-rows x 64 x-columns staged in shared memory with 16-byte loads; thread = (column, 32-row group).
__global__ void __launch_bounds__(256)
quantize_mx_colwise_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int R, int C, uint8_t* __restrict__ qt,
                           long long ldqt, uint8_t* __restrict__ sft) {
  __shared__ __align__(16) __nv_bfloat16 tile[128][64 + 8];
  const int r0 = blockIdx.y * 128, c0 = blockIdx.x * 64;
  const bool vec = (ldx % 8 == 0) && (c0 + 64 <= C);
  if (vec) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int i = threadIdx.x + it * 256;  // 1024 16-byte pieces
      const int rr = i >> 3, cc = (i & 7) * 8;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (r0 + rr < R) v = ld_nc_v4(reinterpret_cast<const uint4*>(x + (long long)(r0 + rr) * ldx + c0 + cc));
      *reinterpret_cast<uint4*>(&tile[rr][cc]) = v;
    }
  } else {
    for (int i = threadIdx.x; i < 128 * 64; i += 256) {
      const int rr = i >> 6, cc = i & 63;
      tile[rr][cc] = (r0 + rr < R && c0 + cc < C) ? x[(long long)(r0 + rr) * ldx + c0 + cc] : __float2bfloat16(0.f);
    }
  }
  __syncthreads();
  const int c = threadIdx.x & 63, g = threadIdx.x >> 6;  // 4 groups of 32 rows
  if (c0 + c >= C) return;
  float v[32];
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    v[i] = __bfloat162float(tile[g * 32 + i][c]);
    amax = fmaxf(amax, fabsf(v[i]));
  }
  const uint32_t be = mx_biased_exp(amax);
  const float inv = mx_inv_scale(be);
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    w[i] = (uint32_t)f_to_e4m3(v[4 * i] * inv) | ((uint32_t)f_to_e4m3(v[4 * i + 1] * inv) << 8) |
           ((uint32_t)f_to_e4m3(v[4 * i + 2] * inv) << 16) | ((uint32_t)f_to_e4m3(v[4 * i + 3] * inv) << 24);
  uint8_t* dst = qt + (long long)(c0 + c) * ldqt + r0 + g * 32;
  reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
  reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4], w[5], w[6], w[7]);
  sft[sf_offset(c0 + c, (r0 >> 5) + g, R >> 7)] = (uint8_t)be;
}

}  // namespace mx

// ------------------------------------------------------------------------------------------------ SM-pair version
// Cluster of two CTAs, 256 x 256 x 128 tile: each CTA stages its 128 rows of A, half of B, its own 128 rows of A-scales
// and the B-scales of all 256 columns; the leader issues tcgen05.cp.cta_group::2 (scales smem -> TMEM in both CTAs) and
// tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale (M = 256).  512 TMEM columns cannot hold two 256-column
// accumulators plus the scale columns, so there is ONE accumulator stage; eight epilogue warps drain it into
// registers (bf16-packed), hand TMEM back, and only then do the global stores, which overlap the next tile's MMAs.
namespace mx2 {

static constexpr int BM = 128, BN = 256, BNH = 128, BK = 128;
static constexpr int STAGES = 5;
static constexpr int A_BYTES = BM * BK, B_BYTES = BNH * BK, SFA_BYTES = 512, SFB_BYTES = 1024;
static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + SFA_BYTES + SFB_BYTES;  // 34,304
static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 + 256;
static constexpr int THREADS = 384;  // w0 TMA, w1 MMA, w2 TMEM alloc, w3 idle, w4-11 epilogue
static constexpr int GROUP_M = 4;

struct Params {
  void* C;
  const void* bias;
  int M, N, K;
  long long ldc;
  int out_fp32, accumulate;
  int num_m_tiles, num_n_tiles;
};

__device__ __forceinline__ void utccp_2sm(uint32_t tmem_dst, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], %1;" ::"r"(tmem_dst), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_mxf8_2sm(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc,
                                              uint32_t tsfa, uint32_t tsfb) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(tsfa), "r"(tsfb)
      : "memory");
}
__device__ __forceinline__ void tile_coords2(int tile, int num_m, int num_n, int& m_t, int& n_t) {
  const int per_group = GROUP_M * num_n;
  const int group = tile / per_group;
  const int first_m = group * GROUP_M;
  const int gm = min(GROUP_M, num_m - first_m);
  const int r = tile - group * per_group;
  m_t = first_m + (r % gm);
  n_t = r / gm;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gemm2_mxfp8_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                   const __grid_constant__ CUtensorMap tmap_sfa, const __grid_constant__ CUtensorMap tmap_sfb, const Params p) {
  using namespace c2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + STAGES * A_BYTES;
  uint8_t* smem_sf = smem_b + STAGES * B_BYTES;  // [STAGES][SFA 512 | SFB 1024]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_sf + STAGES * (SFA_BYTES + SFB_BYTES));
  uint64_t* full_bar = bars;                 // [STAGES] leader's are waited on
  uint64_t* empty_bar = bars + STAGES;       // [STAGES] per CTA (multicast commit)
  uint64_t* tmem_full = bars + 2 * STAGES;   // per CTA (multicast commit)
  uint64_t* tmem_empty = bars + 2 * STAGES + 1;  // leader's: 16 arrivals (8 epilogue warps x 2 CTAs)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int num_k_blocks = p.K / BK;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a); prefetch_tensormap(&tmap_b);
    prefetch_tensormap(&tmap_sfa); prefetch_tensormap(&tmap_sfb);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 16);
    fence_barrier_init();
  }
  cluster_sync_all();
  if (warp == 2) tmem_alloc_2sm<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tsfa = tmem_base + 256, tsfb = tmem_base + 260;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int m_t, n_t;
        tile_coords2(tile, p.num_m_tiles, p.num_n_tiles, m_t, n_t);
        const int m_blk = m_t * 2 + (int)cta;  // my 128 rows of A (and their scales)
        const int n_blk = n_t * 2 + (int)cta;  // my 128 rows of B
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          const uint32_t lbar = smem_u32(&full_bar[stage]) & PEER_MASK;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          uint8_t* ssf = smem_sf + stage * (SFA_BYTES + SFB_BYTES);
          tma_load_2d_2sm(smem_a + stage * A_BYTES, &tmap_a, lbar, kb * BK, m_blk * BM);
          tma_load_2d_2sm(smem_b + stage * B_BYTES, &tmap_b, lbar, kb * BK, n_blk * BNH);
          tma_load_2d_2sm(ssf, &tmap_sfa, lbar, 0, m_blk * num_k_blocks + kb);
          tma_load_2d_2sm(ssf + SFA_BYTES, &tmap_sfb, lbar, 0, (n_t * 2) * num_k_blocks + kb);
          tma_load_2d_2sm(ssf + SFA_BYTES + 512, &tmap_sfb, lbar, 0, (n_t * 2 + 1) * num_k_blocks + kb);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && elect_one()) {
      int stage = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(tmem_empty, acc_phase ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * A_BYTES);
          const uint32_t sb = smem_u32(smem_b + stage * B_BYTES);
          const uint32_t ssf = smem_u32(smem_sf + stage * (SFA_BYTES + SFB_BYTES));
          // the tensor pipe executes these in issue order: the copies overwrite the scale columns only after the
          // previous k-block's MMAs have read them
          utccp_2sm(tsfa, mx::sf_desc(ssf));
          utccp_2sm(tsfb, mx::sf_desc(ssf + SFA_BYTES));
          utccp_2sm(tsfb + 4, mx::sf_desc(ssf + SFA_BYTES + 512));
#pragma unroll
          for (int k = 0; k < BK / 32; ++k) {
            const uint64_t da = make_umma_desc(sa + k * 32, 16, 1024);
            const uint64_t db = make_umma_desc(sb + k * 32, 16, 1024);
            umma_mxf8_2sm(tmem_base, da, db, mx::idesc_mx(256, BN, (uint32_t)k), (kb > 0 || k > 0) ? 1u : 0u, tsfa, tsfb);
          }
          umma_commit_2sm(&empty_bar[stage]);
          if (kb == num_k_blocks - 1) umma_commit_2sm(tmem_full);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ---- epilogue: 8 warps; warp pair (quadrant q) splits the 256 columns in two halves
    const int q = warp & 3, half = (warp - 4) >> 2;
    uint32_t acc_phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int m_t, n_t;
      tile_coords2(tile, p.num_m_tiles, p.num_n_tiles, m_t, n_t);
      const int row = m_t * 256 + (int)cta * BM + q * 32 + lane;
      const int n0 = n_t * BN + half * 128;
      mbar_wait(tmem_full, acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * 128;
      if (!p.out_fp32) {
        const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(p.bias);
#pragma unroll 1
        for (int rnd = 0; rnd < 2; ++rnd) {  // 64 columns per round, held as 32 packed bf16 pairs
          uint32_t packed[32];
#pragma unroll
          for (int c = 0; c < 64; c += 32) {
            uint32_t v[32];
            tmem_ld_32x32b_x32(taddr + rnd * 64 + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i += 2) packed[(c + i) >> 1] = f2_to_bf2(__uint_as_float(v[i]), __uint_as_float(v[i + 1]));
          }
          if (rnd == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(smem_u32(tmem_empty) & PEER_MASK);  // TMEM is free: the next tile may start
          }
          const int nc = n0 + rnd * 64;
          if (row < p.M && nc < p.N) {
            __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + nc;
            const bool fast = (nc + 64 <= p.N) && (p.ldc % 8 == 0) && bias == nullptr && !p.accumulate;
            if (fast) {
#pragma unroll
              for (int i = 0; i < 32; i += 4)
                *reinterpret_cast<uint4*>(crow + 2 * i) = make_uint4(packed[i], packed[i + 1], packed[i + 2], packed[i + 3]);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) {  // compile-time indices keep `packed` in registers
                const int col = nc + 2 * i;
                if (col < p.N) {
                  float2 t = bf2_to_f2(packed[i]);
                  if (bias) {
                    t.x += __bfloat162float(__ldg(bias + col));
                    if (col + 1 < p.N) t.y += __bfloat162float(__ldg(bias + col + 1));
                  }
                  if (p.accumulate) {
                    t.x += __bfloat162float(crow[2 * i]);
                    if (col + 1 < p.N) t.y += __bfloat162float(crow[2 * i + 1]);
                  }
                  crow[2 * i] = __float2bfloat16(t.x);
                  if (col + 1 < p.N) crow[2 * i + 1] = __float2bfloat16(t.y);
                }
              }
            }
          }
        }
      } else {
        // fp32 output: stream straight from TMEM (keeps full precision), then release
#pragma unroll 1
        for (int c = 0; c < 128; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
          if (row < p.M) {
            float* crow = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + n0 + c;
            for (int i = 0; i < 32; ++i)
              if (n0 + c + i < p.N) {
                float o = __uint_as_float(v[i]);
                if (p.bias) o += __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(p.bias) + n0 + c + i));
                crow[i] = p.accumulate ? crow[i] + o : o;
              }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(smem_u32(tmem_empty) & PEER_MASK);
      }
      acc_phase ^= 1;
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm<512>(tmem_base);
  }
}

}  // namespace mx2
}  // namespace b200

using namespace b200;

// A_q [M,K] e4m3 (row stride lda), B_q [N,K] e4m3 (row stride ldb), sfa / sfb: swizzled UE8M0 blocks (see file header).
// K % 128 == 0; M, N arbitrary (scale buffers are padded to whole 128-row blocks by the quantisers' callers).
B200_API int b200_gemm_mxfp8(const void* A, const void* B, const void* sfa, const void* sfb, void* C, const void* bias, int M,
                             int N, int K, long long lda, long long ldb, long long ldc, int out_fp32, int accumulate,
                             void* stream) {
  using namespace b200::mx;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((K % BK) || (lda % 16) || (ldb % 16) || !sfa || !sfb) return B200_ERR_BAD_ARG;
  CUtensorMap ta, tb;
  int rc = make_tmap_u8_2d(&ta, A, K, M, lda, BK, BM);
  if (rc) return rc;
  rc = make_tmap_u8_2d(&tb, B, K, N, ldb, BK, BN);
  if (rc) return rc;
  Params p;
  p.C = C; p.bias = bias; p.sfa = (const uint8_t*)sfa; p.sfb = (const uint8_t*)sfb; p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.num_m_blocks = (M + BM - 1) / BM; p.num_n_blocks = (N + BN - 1) / BN;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(gemm_mxfp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr = true;
  }
  int ctas = p.num_m_blocks * p.num_n_blocks;
  if (ctas > num_sms()) ctas = num_sms();
  gemm_mxfp8_kernel<<<ctas, 256, SMEM, (cudaStream_t)stream>>>(ta, tb, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// bytes of the scale buffer for a [rows, k] operand (k % 128 == 0)
B200_API long long b200_mx_sf_bytes(long long rows, long long k) { return ((rows + 127) / 128) * (k / 128) * 512; }

// q [R,C] = mx(x) with scales along C (C % 128 == 0); sf must hold b200_mx_sf_bytes(R, C) bytes, zero-initialised
B200_API int b200_quantize_mx_rowwise(const void* x, long long ldx, int R, int C, void* q, long long ldq, void* sf, void* stream) {
  using namespace b200::mx;
  if (R <= 0 || C <= 0) return 0;
  if ((C % 128) || (ldx % 8) || (ldq % 16)) return B200_ERR_BAD_ARG;
  const long long work = (long long)R * (C / 32);
  long long blocks = (work + 255) / 256;
  const long long cap = (long long)num_sms() * 32;
  if (blocks > cap) blocks = cap;
  quantize_mx_rowwise_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, ldx, R, C, (uint8_t*)q, ldq, (uint8_t*)sf);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// qt [C,R] = mx(x^T) with scales along R (R % 128 == 0); sft must hold b200_mx_sf_bytes(C, R) bytes, zero-initialised
B200_API int b200_quantize_mx_colwise(const void* x, long long ldx, int R, int C, void* qt, long long ldqt, void* sft,
                                      void* stream) {
  using namespace b200::mx;
  if (R <= 0 || C <= 0) return 0;
  if ((R % 128) || (ldx % 2) || (ldqt % 16)) return B200_ERR_BAD_ARG;
  dim3 grid((C + 63) / 64, R / 128);
  quantize_mx_colwise_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ldx, R, C, (uint8_t*)qt, ldqt,
                                                                     (uint8_t*)sft);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// Same contract as b200_gemm_mxfp8 on an SM pair (256 x 256 tiles).  The scale buffers must cover whole 256-row
// tiles: b200_mx_sf_bytes(round_up(rows, 256), k) bytes, zero-initialised.
B200_API int b200_gemm2_mxfp8(const void* A, const void* B, const void* sfa, const void* sfb, void* C, const void* bias, int M,
                              int N, int K, long long lda, long long ldb, long long ldc, int out_fp32, int accumulate,
                              void* stream) {
  using namespace b200::mx2;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((K % BK) || (lda % 16) || (ldb % 16) || !sfa || !sfb) return B200_ERR_BAD_ARG;
  CUtensorMap ta, tb, tsa, tsb;
  int rc;
  if ((rc = make_tmap_u8_2d(&ta, A, K, M, lda, BK, BM))) return rc;
  if ((rc = make_tmap_u8_2d(&tb, B, K, N, ldb, BK, BNH))) return rc;
  Params p;
  p.C = C; p.bias = bias; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.num_m_tiles = (M + 255) / 256; p.num_n_tiles = (N + BN - 1) / BN;
  const long long kblocks = K / BK;
  if ((rc = make_tmap_sf_blocks(&tsa, sfa, (long long)p.num_m_tiles * 2 * kblocks))) return rc;
  if ((rc = make_tmap_sf_blocks(&tsb, sfb, (long long)p.num_n_tiles * 2 * kblocks))) return rc;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(gemm2_mxfp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr = true;
  }
  int clusters = p.num_m_tiles * p.num_n_tiles;
  if (clusters > num_sms() / 2) clusters = num_sms() / 2;
  gemm2_mxfp8_kernel<<<clusters * 2, THREADS, SMEM, (cudaStream_t)stream>>>(ta, tb, tsa, tsb, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
