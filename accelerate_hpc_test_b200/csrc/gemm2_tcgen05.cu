// 2-CTA ("cta_group::2") persistent bf16 GEMM for sm_100a: a cluster of two CTAs (one SM pair) computes a
// 256 x 256 tile.  Each CTA stages its own 128 rows of A and *half* of the B tile (128 of the 256 n-rows), the
// leader issues tcgen05.mma.cta_group::2 (M = 256) reading B from both CTAs' shared memory, and each CTA keeps its
// 128 x 256 fp32 half of the accumulator in its own TMEM.  Compared with the 1-CTA kernel (gemm_tcgen05.cu) the
// per-CTA operand traffic drops from 48 KB to 32 KB per k-block, which buys 6 pipeline stages instead of 4 and
// halves the B reads from L2.  Same operand layouts and epilogue options as b200_gemm_bf16.
#include "common.cuh"
#include <string.h>

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, long long inner, long long outer, long long row_stride,
                      int box_inner, int box_outer);
int make_tmap_u8_2d(CUtensorMap* out, const void* base, long long inner, long long outer, long long row_stride,
                    int box_inner, int box_outer);
int num_sms();

namespace g2 {

static constexpr int BM = 128;   // rows per CTA (256 per cluster)
static constexpr int BN = 256;   // columns per cluster tile
static constexpr int BNH = 128;  // B rows staged per CTA
static constexpr int BK = 64;
static constexpr int STAGES = 6;
static constexpr int A_BYTES = BM * BK * 2;   // 16 KB
static constexpr int B_BYTES = BNH * BK * 2;  // 16 KB
static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
static constexpr int CHUNK = 64 * BK * 2;  // one MN-major 64 x 64 box
static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 + 256;
static constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address
static constexpr int GROUP_M = 4;                   // in units of 256-row cluster tiles

struct Params {
  void* C;
  const void* bias;
  int M, N, K;
  long long ldc;
  int out_fp32, accumulate;
  int num_m_tiles, num_n_tiles;
  const float* deq_a;  // fp8 only: per-tensor dequantisation factors (device scalars), result *= deq_a * deq_b
  const float* deq_b;
  // ---- GEMM -> reduce-scatter (rs_world > 0): row r of D belongs to rank r / rs_rows; the epilogue stores this rank's
  // partial straight into slot rs_rank of the owner's [rs_world][rs_rows][N] bf16 buffer (byte offset rs_off of its
  // symmetric buffer) over NVLink.
  uint8_t* rs_peer[8];
  long long rs_off;
  int rs_rows, rs_rank, rs_world;
  // ---- GEMM -> Ulysses all-to-all (ux_world > 0, uses the same TMA-store epilogue): D = X W^T is the fused projection of
  // this rank's S_loc tokens for ALL heads.  Column block `d` of every segment (q | k | v, or one segment) belongs to rank d,
  // which keeps the whole sequence for its heads: rows land at (b * S + ux_rank * S_loc + s) of d's [B*S, part] buffer.
  int ux_world, ux_rank, ux_S_loc, ux_nseg;
  int ux_start[3], ux_width[3], ux_dst[3];  // per segment: first column, columns per rank, first destination column
  // ---- all-gather -> GEMM (FUSE_AG): A = [ag_world * ag_rows, K] is assembled in ag_full from the shards every rank
  // keeps at byte offset ag_off + rank * shard_bytes of its symmetric buffer.  Warp 3 of every CTA pulls the peers'
  // shards with cp.async.bulk while the tensor cores already work on rows that are present; ag_ready[p] counts the
  // bytes of shard p that landed (reaches ag_expect when complete).
  //
  // The same gather warp also serves the FSDP weight gather (ag_mode 1-3): every rank keeps its dim-0 piece of a weight
  // W[rows * world, cols] at byte offset ag_off of its symmetric buffer (ag_src_stride = 0) and the pieces are assembled in
  // ag_full.  ag_mode says what the tensor cores must wait for:
  //   0  A rows (activations, above)            1  B rows indexed by n  (forward:  D = X W^T, W is this GEMM's B operand)
  //   2  B rows indexed by k (dgrad: D = dY W)  3  nothing: the buffer belongs to a LATER kernel of the stream ("carry")
  uint8_t* ag_peer[8];
  long long ag_off;
  long long ag_src_stride;   // byte distance between consecutive ranks' pieces inside one source buffer (0: same offset everywhere)
  long long ag_shard_bytes;  // bytes of one rank's piece
  uint8_t* ag_full;
  unsigned long long* ag_ready;
  unsigned long long ag_expect;
  int ag_rows, ag_rank, ag_world, ag_mode;
};

// GEMM -> reduce-scatter epilogue: every epilogue warp stages 32 rows x 64 columns (4 KB, 128-byte swizzled) and sends
// them to the owner's slot with ONE 2-D TMA store — LSU stores over NVLink are an order of magnitude slower.
struct RsMaps {
  CUtensorMap m[8];  // m[d]: rank d's [world * rows, N] bf16 slot buffer
};
static constexpr int RS_STAGING = 4 * 4096;

static constexpr int AGG_CHUNK = 8192;
static constexpr int AGG_STAGES = 3;
static constexpr int AGG_SMEM = AGG_CHUNK * AGG_STAGES + 64;
static constexpr uint32_t kSpinLimit2 = 1u << 27;

__device__ __forceinline__ void bulk_g2s_g2(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g_g2(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
// with the fused all-gather, visit the 256-row tiles of my own shard first, then the shards in arrival order
__device__ __forceinline__ int ag_rotate2(int m_t, const Params& p) {
  // start at the tile that holds the first row of my shard (a tile may span several small shards, or a shard several tiles)
  if (p.ag_mode != 0) return m_t;
  const int start = (int)(((long long)p.ag_rank * p.ag_rows) / 256);
  return (m_t + start) % p.num_m_tiles;
}
// weight gather, forward: begin with the n-tiles made of my own rows of W
__device__ __forceinline__ int ag_rotate_n(int n_t, const Params& p) {
  if (p.ag_mode != 1) return n_t;
  const int start = (int)(((long long)p.ag_rank * p.ag_rows) / BN);
  return (n_t + start) % p.num_n_tiles;
}
// bounded wait until the pieces that hold rows [r0, r1] of the gathered matrix have landed
__device__ __forceinline__ void ag_wait_rows(const Params& p, int r0, int r1) {
  const int s_lo = r0 / p.ag_rows;
  const int s_hi = min(r1 / p.ag_rows, p.ag_world - 1);
  for (int sh = s_lo; sh <= s_hi; ++sh) {
    if (sh == p.ag_rank && p.ag_mode == 0) continue;
    uint32_t spins = 0;
    while (*reinterpret_cast<volatile unsigned long long*>(p.ag_ready + sh) < p.ag_expect) {
      __nanosleep(32);
      if (++spins > kSpinLimit2) {
        printf("b200 gemm2: all-gather piece %d never arrived (block %d, mode %d)\n", sh, blockIdx.x, p.ag_mode);
        __trap();
      }
    }
  }
  __threadfence();
  asm volatile("fence.proxy.async;" ::: "memory");
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA load whose completion is signalled on the LEADER CTA's mbarrier
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, uint32_t leader_bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(acc)
      : "memory");
}
// arrive (once all earlier MMAs of this thread retired) on the same barrier in BOTH CTAs of the pair
__device__ __forceinline__ void umma_e4m3_2sm(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_result) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}

// groups of GROUP_M n-tiles, m fastest: the order used when B arrives piece by piece (rows of B = n)
__device__ __forceinline__ void tile_coords_nmajor(int tile, int num_m, int num_n, int& m_t, int& n_t) {
  const int per_group = GROUP_M * num_m;
  const int group = tile / per_group;
  const int first_n = group * GROUP_M;
  const int gn = min(GROUP_M, num_n - first_n);
  const int r = tile - group * per_group;
  n_t = first_n + (r % gn);
  m_t = r / gn;
}
__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_t, int& n_t) {
  const int per_group = GROUP_M * num_n;
  const int group = tile / per_group;
  const int first_m = group * GROUP_M;
  const int gm = min(GROUP_M, num_m - first_m);
  const int r = tile - group * per_group;
  m_t = first_m + (r % gm);
  n_t = r / gm;
}

// FP8 = 1: operands are e4m3 bytes, K-major only; a k-block is still 128 bytes per row (128 elements), the MMA is
// kind::f8f6f4 with K = 32 per instruction, and the epilogue applies the two per-tensor dequantisation factors.
template <int A_MN, int B_MN, int FP8 = 0, int FUSE_AG = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
             const __grid_constant__ RsMaps rs_maps, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full_bar = bars;                    // [STAGES]  (only the leader's are waited on)
  uint64_t* empty_bar = bars + STAGES;          // [STAGES]  per CTA, signalled by the multicast commit
  uint64_t* tmem_full = bars + 2 * STAGES;      // [2]       per CTA, multicast commit
  uint64_t* tmem_empty = bars + 2 * STAGES + 2; // [2]       leader's: 8 arrivals (4 epilogue warps x 2 CTAs)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  uint64_t* ag_bars = bars + 2 * STAGES + 6;                        // [AGG_STAGES] (FUSE_AG)
  uint8_t* ag_ring = smem + STAGES * STAGE_BYTES + 256;   // after the 256-byte barrier block (FUSE_AG)
  uint8_t* rs_stage = smem + STAGES * STAGE_BYTES + 1024;  // 4 x 4 KB, 1024-aligned (reduce-scatter epilogue)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  constexpr int BKE = FP8 ? 128 : BK;  // elements per k-block
  const int num_k_blocks = (p.K + BKE - 1) / BKE;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8);
    }
    if (FUSE_AG)
      for (int i = 0; i < AGG_STAGES; ++i) mbar_init(&ag_bars[i], 1);
    fence_barrier_init();
  }
  cluster_sync_all();  // barriers of both CTAs exist before anybody signals across the pair
  if (warp == 2) tmem_alloc_2sm<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ TMA producer (one per CTA) ================================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int m_t, n_t;
        if (FUSE_AG && p.ag_mode == 1) tile_coords_nmajor(tile, p.num_m_tiles, p.num_n_tiles, m_t, n_t);
        else tile_coords(tile, p.num_m_tiles, p.num_n_tiles, m_t, n_t);
        if (FUSE_AG) { m_t = ag_rotate2(m_t, p); n_t = ag_rotate_n(n_t, p); }
        const int m0 = m_t * 256 + (int)cta * BM;
        const int n0 = n_t * BN + (int)cta * BNH;
        // rows [m0, m0 + BM) of A (mode 0) / rows [n0, n0 + BNH) of B (mode 1) must have arrived
        if (FUSE_AG && p.ag_world > 1 && p.ag_mode == 0 && m0 < p.M) ag_wait_rows(p, m0, m0 + BM - 1);
        if (FUSE_AG && p.ag_mode == 1 && n0 < p.N) ag_wait_rows(p, n0, min(n0 + BNH, p.N) - 1);
        // mode 2: the reduction runs over the rows of the gathered matrix; every tile starts in my own piece
        const int kb_start = (FUSE_AG && p.ag_mode == 2) ? (int)(((long long)p.ag_rank * p.ag_rows) / BKE) % num_k_blocks : 0;
        int k_piece = -1;
        for (int kq = 0; kq < num_k_blocks; ++kq) {
          int kb = kq + kb_start;
          if (kb >= num_k_blocks) kb -= num_k_blocks;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          const uint32_t lbar = smem_u32(&full_bar[stage]) & PEER_MASK;  // the leader's barrier
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          uint8_t* sa = smem_a + stage * A_BYTES;
          uint8_t* sb = smem_b + stage * B_BYTES;
          const int k0 = kb * BKE;
          if (FUSE_AG && p.ag_mode == 2) {
            const int k1 = min(k0 + BKE, p.K) - 1;
            if (k1 / p.ag_rows != k_piece || k0 / p.ag_rows != k_piece) {
              ag_wait_rows(p, k0, k1);
              k_piece = k1 / p.ag_rows;
            }
          }
          if (A_MN == 0) {
            tma_load_2d_2sm(sa, &tmap_a, lbar, k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d_2sm(sa + c * CHUNK, &tmap_a, lbar, m0 + c * 64, k0);
          }
          if (B_MN == 0) {
            tma_load_2d_2sm(sb, &tmap_b, lbar, k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BNH / 64; ++c) tma_load_2d_2sm(sb + c * CHUNK, &tmap_b, lbar, n0 + c * 64, k0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (leader CTA only) ================================
    if (leader && elect_one()) {
      constexpr uint32_t idesc = FP8 ? make_idesc_e4m3(256, BN) : make_idesc_bf16(256, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * A_BYTES);
          const uint32_t sb = smem_u32(smem_b + stage * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN == 0 ? make_umma_desc(sa + k * 32, 16, 1024) : make_umma_desc(sa + k * 16 * 128, CHUNK, 1024);
            const uint64_t db = B_MN == 0 ? make_umma_desc(sb + k * 32, 16, 1024) : make_umma_desc(sb + k * 16 * 128, CHUNK, 1024);
            if (FP8) umma_e4m3_2sm(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            else umma_bf16_2sm(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_2sm(&empty_bar[stage]);
          if (kb == num_k_blocks - 1) umma_commit_2sm(&tmem_full[acc]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (FUSE_AG && warp == 3) {
    // ================================ all-gather warp (every CTA moves its share of the chunks) ================================
    if ((p.ag_world > 1 || p.ag_mode != 0) && elect_one()) {
      const long long shard_bytes = p.ag_shard_bytes;
      const long long cps = (shard_bytes + AGG_CHUNK - 1) / AGG_CHUNK;
      // weight gather (modes 1-3): slot 0 is my own piece (a local copy out of my shard), then the peers; activation gather
      // (mode 0): the caller already placed my rows
      const int first_slot = p.ag_mode == 0 ? 1 : 0;
      const long long total = cps * (p.ag_world - first_slot);
      const long long n = (total > blockIdx.x) ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
      auto locate = [&](long long k, const uint8_t*& src, uint8_t*& dst, uint32_t& bytes, int& peer) {
        const long long g = blockIdx.x + k * (long long)gridDim.x;
        const int slot = (int)(g / cps);
        const long long c = g - (long long)slot * cps;
        peer = (p.ag_rank + first_slot + slot) % p.ag_world;
        const long long o = c * AGG_CHUNK;
        const long long left = shard_bytes - o;
        bytes = (uint32_t)(left < AGG_CHUNK ? left : AGG_CHUNK);
        src = p.ag_peer[peer] + p.ag_off + (long long)peer * p.ag_src_stride + o;
        dst = p.ag_full + (long long)peer * shard_bytes + o;
      };
      auto issue_load = [&](long long k) {
        const uint8_t* src; uint8_t* dst; uint32_t bytes; int peer;
        locate(k, src, dst, bytes, peer);
        const int st = (int)(k % AGG_STAGES);
        mbar_arrive_expect_tx(&ag_bars[st], bytes);
        bulk_g2s_g2(ag_ring + st * AGG_CHUNK, src, bytes, &ag_bars[st]);
      };
      auto publish = [&](long long k) {  // chunk k is in local memory: account for its bytes
        const uint8_t* s2; uint8_t* d2; uint32_t b2; int p2;
        locate(k, s2, d2, b2, p2);
        asm volatile("fence.proxy.async;" ::: "memory");
        __threadfence();
        atomicAdd(p.ag_ready + p2, (unsigned long long)b2);
      };
      for (long long k = 0; k < 2 && k < n; ++k) issue_load(k);
      for (long long k = 0; k < n; ++k) {
        const int st = (int)(k % AGG_STAGES);
        mbar_wait(&ag_bars[st], (uint32_t)((k / AGG_STAGES) & 1));
        const uint8_t* src; uint8_t* dst; uint32_t bytes; int peer;
        locate(k, src, dst, bytes, peer);
        bulk_s2g_g2(dst, ag_ring + st * AGG_CHUNK, bytes);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (k >= 1) {
          asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");  // store k-1 done: publish it, its slot may refill
          publish(k - 1);
        }
        if (k + 2 < n) issue_load(k + 2);
      }
      if (n > 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        publish(n - 1);
      }
    }
  } else if (warp >= 4) {
    // ================================ epilogue (both CTAs, own 128 rows) ================================
    const int e = warp - 4;
    int acc = 0;
    uint32_t acc_phase = 0;
    float alpha = 1.f;
    if (FP8) alpha = __ldg(p.deq_a) * __ldg(p.deq_b);
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int m_t, n_t;
      if (FUSE_AG && p.ag_mode == 1) tile_coords_nmajor(tile, p.num_m_tiles, p.num_n_tiles, m_t, n_t);
      else tile_coords(tile, p.num_m_tiles, p.num_n_tiles, m_t, n_t);
      if (FUSE_AG) { m_t = ag_rotate2(m_t, p); n_t = ag_rotate_n(n_t, p); }
      const int row = m_t * 256 + (int)cta * BM + e * 32 + lane;
      const int n0 = n_t * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + acc * BN;
      const bool row_ok = row < p.M;
      const bool vec_ok = (p.ldc % 8 == 0);
      if ((p.rs_world > 0 && (p.rs_rows % 32) == 0) || p.ux_world > 0) {
        // ---- reduce-scatter / Ulysses scatter: TMEM -> swizzled smem tile -> one TMA store per 32 x 64 block to its owner
        const int row0 = row - lane;  // first of this warp's 32 rows: all owned by the same rank
        int owner = 0, dst_row = 0;
        if (p.ux_world > 0) {
          const int b = row0 / p.ux_S_loc;
          dst_row = b * p.ux_S_loc * p.ux_world + p.ux_rank * p.ux_S_loc + (row0 - b * p.ux_S_loc);
        } else {
          owner = row0 / p.rs_rows;
          dst_row = p.rs_rank * p.rs_rows + (row0 - owner * p.rs_rows);
        }
        uint8_t* stg = rs_stage + e * 4096;
#pragma unroll 1
        for (int c = 0; c < BN; c += 64) {
          uint32_t v0[32], v1[32];
          tmem_ld_32x32b_x32(taddr + c, v0);
          tmem_ld_32x32b_x32(taddr + c + 32, v1);
          tmem_ld_wait();
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // previous block left the staging tile
          __syncwarp();
          uint8_t* rp = stg + lane * 128;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 o;
            o.x = f2_to_bf2(__uint_as_float(v0[8 * j]), __uint_as_float(v0[8 * j + 1]));
            o.y = f2_to_bf2(__uint_as_float(v0[8 * j + 2]), __uint_as_float(v0[8 * j + 3]));
            o.z = f2_to_bf2(__uint_as_float(v0[8 * j + 4]), __uint_as_float(v0[8 * j + 5]));
            o.w = f2_to_bf2(__uint_as_float(v0[8 * j + 6]), __uint_as_float(v0[8 * j + 7]));
            *reinterpret_cast<uint4*>(rp + ((j ^ (lane & 7)) << 4)) = o;
            o.x = f2_to_bf2(__uint_as_float(v1[8 * j]), __uint_as_float(v1[8 * j + 1]));
            o.y = f2_to_bf2(__uint_as_float(v1[8 * j + 2]), __uint_as_float(v1[8 * j + 3]));
            o.z = f2_to_bf2(__uint_as_float(v1[8 * j + 4]), __uint_as_float(v1[8 * j + 5]));
            o.w = f2_to_bf2(__uint_as_float(v1[8 * j + 6]), __uint_as_float(v1[8 * j + 7]));
            *reinterpret_cast<uint4*>(rp + (((j + 4) ^ (lane & 7)) << 4)) = o;
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0 && row0 < p.M && n0 + c < p.N) {
            int dst_col = n0 + c;
            if (p.ux_world > 0) {  // 64-column blocks never straddle a segment or a rank (widths are multiples of 64)
              int sgm = 0;
              while (sgm + 1 < p.ux_nseg && dst_col >= p.ux_start[sgm + 1]) ++sgm;
              const int rel = dst_col - p.ux_start[sgm];
              owner = rel / p.ux_width[sgm];
              dst_col = p.ux_dst[sgm] + (rel - owner * p.ux_width[sgm]);
            }
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                             reinterpret_cast<uint64_t>(&rs_maps.m[owner])),
                         "r"(smem_u32(stg)), "r"(dst_col), "r"(dst_row)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // data is on its way before TMEM / smem are recycled
        __syncwarp();
      } else
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
          if (FP8) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] *= alpha;
          }
          if (p.bias != nullptr) {
            const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(p.bias);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col0 + i < p.N) f[i] += __bfloat162float(__ldg(bias + col0 + i));
          }
          if (p.out_fp32) {
            float* crow = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col0;
            if (col0 + 32 <= p.N && (p.ldc % 4 == 0)) {
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                float4 o = make_float4(f[i], f[i + 1], f[i + 2], f[i + 3]);
                if (p.accumulate) {
                  float4 old = *reinterpret_cast<float4*>(crow + i);
                  o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                }
                *reinterpret_cast<float4*>(crow + i) = o;
              }
            } else {
              for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) crow[i] = p.accumulate ? crow[i] + f[i] : f[i];
            }
          } else {
            __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col0;
            if (p.rs_world > 0) {
              // scatter: this row lives on rank `owner`; my partial goes to my slot of its buffer
              const int owner = row / p.rs_rows;
              const long long lrow = (long long)p.rs_rank * p.rs_rows + (row - owner * p.rs_rows);
              crow = reinterpret_cast<__nv_bfloat16*>(p.rs_peer[owner] + p.rs_off) + lrow * p.N + col0;
            }
            if (col0 + 32 <= p.N && vec_ok) {
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
      if (lane == 0) mbar_arrive_cluster(smem_u32(&tmem_empty[acc]) & PEER_MASK);  // the leader's barrier
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  cluster_sync_all();  // nobody frees TMEM / exits while the partner may still signal or read
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm<512>(tmem_base);
  }
}

static RsMaps g_no_rs_maps;  // zero-filled: launches without the reduce-scatter epilogue never touch it

template <int A_MN, int B_MN, int FP8 = 0, int FUSE_AG = 0>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, cudaStream_t st,
                  const RsMaps& rs = g_no_rs_maps) {
  static bool attr_set = false;
  constexpr int smem_max = SMEM + (FUSE_AG ? AGG_SMEM : RS_STAGING + 768);
  // plain GEMMs stay at 198 KB so that a comm CTA (< 25 KB ring) still fits next to them on every SM
  const int smem_bytes = SMEM + (FUSE_AG ? AGG_SMEM : ((p.rs_world > 0 || p.ux_world > 0) ? RS_STAGING + 768 : 0));
  if (!attr_set) {
    B200_CUDA(cudaFuncSetAttribute(gemm2_kernel<A_MN, B_MN, FP8, FUSE_AG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    attr_set = true;
  }
  int clusters = p.num_m_tiles * p.num_n_tiles;
  const int max_clusters = num_sms() / 2;
  if (clusters > max_clusters) clusters = max_clusters;
  gemm2_kernel<A_MN, B_MN, FP8, FUSE_AG><<<clusters * 2, 256, smem_bytes, st>>>(ta, tb, rs, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

}  // namespace g2
}  // namespace b200

// Same contract as b200_gemm_bf16 (see include/b200_ops.h); picked by the Python layer for plain GEMMs.
B200_API int b200_gemm2_bf16(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, long long lda,
                             long long ldb, long long ldc, int a_mn_major, int b_mn_major, int out_fp32, int accumulate,
                             void* stream) {
  using namespace b200;
  using namespace b200::g2;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((lda % 8) || (ldb % 8) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return B200_ERR_BAD_ARG;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BK, BM);
  else rc = make_tmap_bf16_2d(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BNH);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  Params p = {};
  p.C = C; p.bias = bias; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = (N + BN - 1) / BN;
  p.deq_a = p.deq_b = nullptr;
  p.rs_world = 0; p.ag_world = 0; p.ux_world = 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a_mn_major && !b_mn_major) return launch<0, 0>(ta, tb, p, st);
  if (!a_mn_major && b_mn_major) return launch<0, 1>(ta, tb, p, st);
  if (a_mn_major && !b_mn_major) return launch<1, 0>(ta, tb, p, st);
  return launch<1, 1>(ta, tb, p, st);
}

// e4m3 x e4m3 -> bf16/fp32 with per-tensor scaling: D[M,N] (+)= deq_a * deq_b * sum_k A(m,k) B(n,k) (+ bias[n]).
// A is [M,K] bytes (row stride lda), B is [N,K] bytes (row stride ldb): both K-major, strides multiples of 16.
// deq_a / deq_b are device fp32 scalars (written by b200_quantize_e4m3).
B200_API int b200_gemm2_e4m3(const void* A, const void* B, void* C, const void* bias, const void* deq_a, const void* deq_b, int M,
                             int N, int K, long long lda, long long ldb, long long ldc, int out_fp32, int accumulate,
                             void* stream) {
  using namespace b200;
  using namespace b200::g2;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((lda % 16) || (ldb % 16) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) || !deq_a || !deq_b)
    return B200_ERR_BAD_ARG;
  CUtensorMap ta, tb;
  int rc = make_tmap_u8_2d(&ta, A, K, M, lda, 128, BM);
  if (rc) return rc;
  rc = make_tmap_u8_2d(&tb, B, K, N, ldb, 128, BNH);
  if (rc) return rc;
  Params p = {};
  p.C = C; p.bias = bias; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = (N + BN - 1) / BN;
  p.deq_a = (const float*)deq_a;
  p.deq_b = (const float*)deq_b;
  p.rs_world = 0; p.ag_world = 0; p.ux_world = 0;
  return launch<0, 0, 1>(ta, tb, p, reinterpret_cast<cudaStream_t>(stream));
}

// all-gather -> GEMM and GEMM -> reduce-scatter on the SM-pair kernel; same contracts as the b200_gemm_bf16_* versions
B200_API int b200_gemm2_bf16_allgather_a(const void* const* peer_ptrs, int rank, int world, long long a_off, void* a_full,
                                         int rows, const void* B, void* C, int N, int K, long long ldb, long long ldc,
                                         int b_mn_major, void* ready, unsigned long long epoch, void* stream) {
  using namespace b200;
  using namespace b200::g2;
  if (world < 1 || world > 8 || rows <= 0 || (K % 8) || (ldb % 8)) return B200_ERR_BAD_ARG;
  const int M = rows * world;
  CUtensorMap ta, tb;
  int rc;
  if ((rc = make_tmap_bf16_2d(&ta, a_full, K, M, K, BK, BM))) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BNH);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  Params p = {};
  p.C = C; p.bias = nullptr; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.out_fp32 = 0; p.accumulate = 0;
  p.num_m_tiles = (M + 255) / 256; p.num_n_tiles = (N + BN - 1) / BN;
  p.deq_a = p.deq_b = nullptr;
  p.rs_world = 0; p.ux_world = 0;
  for (int i = 0; i < 8; ++i) p.ag_peer[i] = i < world ? (uint8_t*)peer_ptrs[i] : nullptr;
  p.ag_off = a_off; p.ag_full = (uint8_t*)a_full; p.ag_ready = (unsigned long long*)ready;
  p.ag_expect = epoch * (unsigned long long)rows * K * 2; p.ag_rows = rows; p.ag_rank = rank; p.ag_world = world;
  p.ag_mode = 0; p.ag_shard_bytes = (long long)rows * K * 2; p.ag_src_stride = p.ag_shard_bytes;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!b_mn_major) return launch<0, 0, 0, 1>(ta, tb, p, st);
  return launch<0, 1, 0, 1>(ta, tb, p, st);
}

// GEMM with a weight-gather job riding on it (FSDP2 parameter all-gather).  Same GEMM contract as
// b200_gemm2_bf16.  Every rank keeps its dim-0 piece (piece_bytes) of a weight at byte offset src_off of its symmetric buffer;
// warp 3 of every CTA pulls the peers' pieces into gathered + q * piece_bytes (q = owning rank; this rank's own piece is copied
// out of its shard the same way, first) while the tensor cores run.  mode 1: the weight is this GEMM's B operand, stored [N, K] (forward) - n-tiles wait
// for the piece that holds their rows and the tile order starts in this rank's piece; mode 2: the weight is B stored [K, N]
// (b_mn_major, dgrad) - k-blocks wait for their piece, every tile starts its reduction in this rank's piece; mode 3: the
// gathered buffer is for a later kernel of the same stream, nothing waits.  piece_rows = rows of the weight per rank (modes 1, 2).
// ready: 8 device counters (bytes landed per piece, monotonic across calls); expect = value that means "complete" for this call.
B200_API int b200_gemm2_bf16_gather(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, long long lda,
                                    long long ldb, long long ldc, int a_mn_major, int b_mn_major, int out_fp32, int accumulate,
                                    const void* const* peer_ptrs, int rank, int world, long long src_off, void* gathered,
                                    long long piece_bytes, int piece_rows, int mode, void* ready, unsigned long long expect,
                                    void* stream) {
  using namespace b200;
  using namespace b200::g2;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if (world < 1 || world > 8 || mode < 1 || mode > 3 || piece_bytes <= 0 || (piece_bytes % 16) || (src_off % 16)) return B200_ERR_BAD_ARG;
  if ((mode == 1 && (a_mn_major || b_mn_major)) || (mode == 2 && (a_mn_major || !b_mn_major))) return B200_ERR_BAD_ARG;
  if (mode != 3 && piece_rows <= 0) return B200_ERR_BAD_ARG;
  if ((lda % 8) || (ldb % 8) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
      (reinterpret_cast<uintptr_t>(gathered) & 15))
    return B200_ERR_BAD_ARG;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BK, BM);
  else rc = make_tmap_bf16_2d(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BNH);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  Params p = {};
  p.C = C; p.bias = bias; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = (N + BN - 1) / BN;
  for (int i = 0; i < 8; ++i) p.ag_peer[i] = i < world ? (uint8_t*)peer_ptrs[i] : nullptr;
  p.ag_off = src_off; p.ag_src_stride = 0; p.ag_shard_bytes = piece_bytes; p.ag_full = (uint8_t*)gathered;
  p.ag_ready = (unsigned long long*)ready; p.ag_expect = expect; p.ag_rows = piece_rows > 0 ? piece_rows : 1;
  p.ag_rank = rank; p.ag_world = world; p.ag_mode = mode;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a_mn_major && !b_mn_major) return launch<0, 0, 0, 1>(ta, tb, p, st);
  if (!a_mn_major && b_mn_major) return launch<0, 1, 0, 1>(ta, tb, p, st);
  if (a_mn_major && b_mn_major) return launch<1, 1, 0, 1>(ta, tb, p, st);
  return B200_ERR_UNSUPPORTED;
}

B200_API int b200_gemm2_bf16_scatter(const void* A, const void* B, const void* const* peer_ptrs, int rank, int world,
                                     long long out_off, int rows, int N, int K, long long lda, long long ldb, int a_mn_major,
                                     int b_mn_major, void* stream) {
  using namespace b200;
  using namespace b200::g2;
  if (world < 1 || world > 8 || rows <= 0 || (N % 8) || (lda % 8) || (ldb % 8)) return B200_ERR_BAD_ARG;
  const int M = rows * world;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BK, BM);
  else rc = make_tmap_bf16_2d(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BNH);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  Params p = {};
  p.C = nullptr; p.bias = nullptr; p.M = M; p.N = N; p.K = K; p.ldc = N; p.out_fp32 = 0; p.accumulate = 0;
  p.num_m_tiles = (M + 255) / 256; p.num_n_tiles = (N + BN - 1) / BN;
  p.deq_a = p.deq_b = nullptr;
  p.ag_world = 0; p.ux_world = 0;
  for (int i = 0; i < 8; ++i) p.rs_peer[i] = i < world ? (uint8_t*)peer_ptrs[i] : nullptr;
  p.rs_off = out_off; p.rs_rows = rows; p.rs_rank = rank; p.rs_world = world;
  RsMaps maps;
  memset(&maps, 0, sizeof(maps));
  if (rows % 32 == 0) {
    for (int d = 0; d < world; ++d)
      if ((rc = make_tmap_bf16_2d(&maps.m[d], (const uint8_t*)peer_ptrs[d] + out_off, N, (long long)world * rows, N, 64, 32))) return rc;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a_mn_major && !b_mn_major) return launch<0, 0>(ta, tb, p, st, maps);
  if (!a_mn_major && b_mn_major) return launch<0, 1>(ta, tb, p, st, maps);
  if (a_mn_major && !b_mn_major) return launch<1, 0>(ta, tb, p, st, maps);
  return launch<1, 1>(ta, tb, p, st, maps);
}

// GEMM -> Ulysses sequence->head all-to-all in one kernel: D = A W^T for this rank's B*S_loc tokens; the epilogue sends every
// 32 x 64 block straight to the rank that owns those heads (2-D TMA store over NVLink).  widths[i] = full width (columns of D)
// of segment i (q | k | v for the fused QKV projection, a single segment for a plain matrix), each a multiple of 64 * world.
// Rank d's destination buffer is [B * S_loc * world, sum(widths) / world] bf16 at byte offset out_off of its workspace.
// The caller runs a device barrier afterwards (all ranks' blocks have landed) before reading its buffer.
B200_API int b200_gemm2_bf16_ulysses_scatter(const void* A, const void* W, const void* const* peer_ptrs, int rank, int world,
                                             long long out_off, int B, int S_loc, const int* widths, int nseg, int K, long long lda,
                                             long long ldb, int b_mn_major, void* stream) {
  using namespace b200;
  using namespace b200::g2;
  if (world < 1 || world > 8 || nseg < 1 || nseg > 3 || (S_loc % 32) || (lda % 8) || (ldb % 8)) return B200_ERR_BAD_ARG;
  const int M = B * S_loc;
  int N = 0;
  for (int i = 0; i < nseg; ++i) {
    if (widths[i] <= 0 || widths[i] % (64 * world)) return B200_ERR_BAD_ARG;
    N += widths[i];
  }
  const int part = N / world;
  CUtensorMap ta, tb;
  int rc;
  if ((rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BK, BM))) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, W, K, N, ldb, BK, BNH);
  else rc = make_tmap_bf16_2d(&tb, W, N, K, ldb, 64, BK);
  if (rc) return rc;
  Params p = {};
  p.C = nullptr; p.bias = nullptr; p.M = M; p.N = N; p.K = K; p.ldc = N; p.out_fp32 = 0; p.accumulate = 0;
  p.num_m_tiles = (M + 255) / 256; p.num_n_tiles = (N + BN - 1) / BN;
  p.deq_a = p.deq_b = nullptr;
  p.ag_world = 0; p.rs_world = 0;
  p.ux_world = world; p.ux_rank = rank; p.ux_S_loc = S_loc; p.ux_nseg = nseg;
  int start = 0, dst = 0;
  for (int i = 0; i < 3; ++i) {
    p.ux_start[i] = start; p.ux_width[i] = i < nseg ? widths[i] / world : 1; p.ux_dst[i] = dst;
    if (i < nseg) { start += widths[i]; dst += widths[i] / world; }
  }
  RsMaps maps;
  memset(&maps, 0, sizeof(maps));
  for (int d = 0; d < world; ++d)
    if ((rc = make_tmap_bf16_2d(&maps.m[d], (const uint8_t*)peer_ptrs[d] + out_off, part, (long long)M * world, part, 64, 32))) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!b_mn_major) return launch<0, 0>(ta, tb, p, st, maps);
  return launch<0, 1>(ta, tb, p, st, maps);
}
