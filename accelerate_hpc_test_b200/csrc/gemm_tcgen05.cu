// Persistent warp-specialised bf16 GEMM for sm_100a: TMA -> smem ring -> tcgen05.mma -> TMEM
// (double-buffered accumulator) -> tcgen05.ld epilogue.
//
//   D[M,N] (+)= sum_k A(m,k) * B(n,k)   (+ bias[n])
//
// Each operand may be stored K-major (row = m or n, K contiguous) or MN-major (row = k, m or n
// contiguous), which covers the three training GEMMs without any transposes in memory:
//   fwd   Y  = X  W^T : A = X  [M,K]  K-major,   B = W [N,K]   K-major
//   dgrad dX = dY W   : A = dY [M,N'] K-major,   B = W [N',K'] MN-major (N' is the reduction)
//   wgrad dW = dY^T X : A = dY [T,N'] MN-major,  B = X [T,K']  MN-major (T is the reduction)
// This is the K1 row of SURVEY.md §2.6: the reference gets these from cuBLASLt through
// nn.Linear (reference src/accelerate/accelerator.py:1797-1802 only wraps autocast around it).
//
// Tile: 128 x 256 x 64, 4 smem stages (48 KB each), 2 TMEM accumulator stages (2 x 256 cols).
// Warp roles (256 threads): w0 TMA producer, w1 MMA issuer, w2 TMEM allocator, w4-7 epilogue.
#include "common.cuh"

namespace b200 {

static constexpr int BM = 128;
static constexpr int BN = 256;
static constexpr int BK = 64;
static constexpr int STAGES = 4;
static constexpr int UMMA_K = 16;
static constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB
static constexpr int B_STAGE_BYTES = BN * BK * 2;  // 32 KB
static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
static constexpr int CHUNK_BYTES = 64 * BK * 2;  // one 64(MN) x 64(K) MN-major TMA box: 8 KB
static constexpr int GEMM_SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
static constexpr int TMEM_COLS = 512;
static constexpr int GROUP_M = 8;

struct GemmParams {
  void* C;
  const void* bias;  // bf16 [N] or null
  int M, N, K;
  long long ldc;
  int out_fp32;
  int accumulate;
  int num_m_blocks, num_n_blocks;
  // ---- GEMM -> reduce-scatter: rows of D are scattered to the rank that owns them (rs_world > 0).
  // Row r belongs to rank r / rs_rows; this rank's partial goes to slot rs_rank of the owner's buffer
  // [rs_world][rs_rows][N] (bf16) at byte offset rs_off of its symmetric buffer.
  uint8_t* rs_peer[8];
  long long rs_off;
  int rs_rows, rs_rank, rs_world;
  // ---- all-gather -> GEMM: A is [ag_world * ag_rows, K] assembled in `ag_full` (local) from the
  // ag_rows-row shards every rank keeps at byte offset ag_off + rank * shard_bytes of its symmetric
  // buffer.  Warp 3 of every CTA pulls the peers' shards with cp.async.bulk while the tensor cores
  // already work on the rows that are present; ag_ready[p] counts the bytes of shard p that landed.
  uint8_t* ag_peer[8];
  long long ag_off;
  uint8_t* ag_full;
  unsigned long long* ag_ready;
  unsigned long long ag_expect;  // value ag_ready[p] reaches when shard p is complete (epoch * shard_bytes)
  int ag_rows, ag_rank, ag_world;
};

static constexpr int AGG_CHUNK = 8192;
static constexpr int AGG_STAGES = 3;
static constexpr int AGG_SMEM = AGG_CHUNK * AGG_STAGES + 64;

__device__ __forceinline__ void bulk_g2s_gemm(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g_gemm(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes)
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

// with the fused all-gather, visit the m-blocks of my own shard first, then the shards in arrival order
__device__ __forceinline__ int ag_rotate(int m_blk, const GemmParams& p) {
  const int start = (int)(((long long)p.ag_rank * p.ag_rows) / BM);  // block holding the first row of my shard
  return (m_blk + start) % p.num_m_blocks;
}

template <int A_MN, int B_MN, int FUSE_AG>
__global__ void __launch_bounds__(256, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024 B alignment
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full_bar = bars;                  // [STAGES]
  uint64_t* empty_bar = bars + STAGES;        // [STAGES]
  uint64_t* tmem_full = bars + 2 * STAGES;    // [2]
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  uint64_t* ag_bars = bars + 2 * STAGES + 6;                                   // [AGG_STAGES]
  uint8_t* ag_ring = smem + STAGES * STAGE_BYTES + 256;                        // only with FUSE_AG

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_k_blocks = (p.K + BK - 1) / BK;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;

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
      mbar_init(&tmem_empty[i], 4);  // one arrive per epilogue warp
    }
    if (FUSE_AG) {
      for (int i = 0; i < AGG_STAGES; ++i) mbar_init(&ag_bars[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc<TMEM_COLS>(tmem_base_smem);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
        if (FUSE_AG) m_blk = ag_rotate(m_blk, p);
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        if (FUSE_AG) {
          // rows [m0, m0 + BM) must have arrived: wait for the shard(s) they belong to
          const int s_lo = m0 / p.ag_rows;
          const int s_hi = min((m0 + BM - 1) / p.ag_rows, p.ag_world - 1);
          for (int sh = s_lo; sh <= s_hi; ++sh) {
            if (sh == p.ag_rank) continue;
            uint32_t spins = 0;
            while (*reinterpret_cast<volatile unsigned long long*>(p.ag_ready + sh) < p.ag_expect) {
              __nanosleep(32);
              if (++spins > kSpinLimit) {
                printf("b200 gemm: all-gather shard %d never arrived (block %d)\n", sh, blockIdx.x);
                __trap();
              }
            }
          }
          __threadfence();
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
          uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
          uint8_t* sb = smem_b + stage * B_STAGE_BYTES;
          const int k0 = kb * BK;
          if (A_MN == 0) {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);  // box {64 k, 128 m}
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c)
              tma_load_2d(sa + c * CHUNK_BYTES, &tmap_a, &full_bar[stage], m0 + c * 64, k0);
          }
          if (B_MN == 0) {
            tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);  // box {64 k, 256 n}
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)
              tma_load_2d(sb + c * CHUNK_BYTES, &tmap_b, &full_bar[stage], n0 + c * 64, k0);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
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
          const uint32_t sa = smem_u32(smem_a + stage * A_STAGE_BYTES);
          const uint32_t sb = smem_u32(smem_b + stage * B_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // K-major: advance 32 B inside the 128 B swizzle row; MN-major: 16 k-rows = 2048 B
            const uint64_t da = A_MN == 0 ? make_umma_desc(sa + k * UMMA_K * 2, 16, 1024)
                                          : make_umma_desc(sa + k * UMMA_K * 128, CHUNK_BYTES, 1024);
            const uint64_t db = B_MN == 0 ? make_umma_desc(sb + k * UMMA_K * 2, 16, 1024)
                                          : make_umma_desc(sb + k * UMMA_K * 128, CHUNK_BYTES, 1024);
            umma_bf16_ss(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // smem slot reusable once these MMAs retire
          if (kb == num_k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (FUSE_AG && warp == 3) {
    // ================================ all-gather warp ================================
    if (p.ag_world > 1 && elect_one()) {
      const long long shard_bytes = (long long)p.ag_rows * p.K * 2;
      const long long cps = (shard_bytes + AGG_CHUNK - 1) / AGG_CHUNK;
      const long long total = cps * (p.ag_world - 1);
      const long long n = (total > blockIdx.x) ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
      auto locate = [&](long long k, const uint8_t*& src, uint8_t*& dst, uint32_t& bytes, int& peer) {
        const long long g = blockIdx.x + k * (long long)gridDim.x;
        const int slot = (int)(g / cps);
        const long long c = g - (long long)slot * cps;
        peer = (p.ag_rank + 1 + slot) % p.ag_world;
        const long long o = c * AGG_CHUNK;
        const long long left = shard_bytes - o;
        bytes = (uint32_t)(left < AGG_CHUNK ? left : AGG_CHUNK);
        src = p.ag_peer[peer] + p.ag_off + (long long)peer * shard_bytes + o;
        dst = p.ag_full + (long long)peer * shard_bytes + o;
      };
      // loads run 2 chunks ahead; a chunk is accounted for once its store to local memory is complete
      auto issue_load = [&](long long k) {
        const uint8_t* src; uint8_t* dst; uint32_t bytes; int peer;
        locate(k, src, dst, bytes, peer);
        const int st = (int)(k % AGG_STAGES);
        mbar_arrive_expect_tx(&ag_bars[st], bytes);
        bulk_g2s_gemm(ag_ring + st * AGG_CHUNK, src, bytes, &ag_bars[st]);
      };
      for (long long k = 0; k < 2 && k < n; ++k) issue_load(k);
      for (long long k = 0; k < n; ++k) {
        const int st = (int)(k % AGG_STAGES);
        mbar_wait(&ag_bars[st], (uint32_t)((k / AGG_STAGES) & 1));
        const uint8_t* src; uint8_t* dst; uint32_t bytes; int peer;
        locate(k, src, dst, bytes, peer);
        bulk_s2g_gemm(dst, ag_ring + st * AGG_CHUNK, bytes);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (k >= 1) {
          // store k-1 has completed: publish its bytes, and its ring slot may be refilled
          asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");
          const uint8_t* s2; uint8_t* d2; uint32_t b2; int p2;
          locate(k - 1, s2, d2, b2, p2);
          asm volatile("fence.proxy.async;" ::: "memory");
          __threadfence();
          atomicAdd(p.ag_ready + p2, (unsigned long long)b2);
        }
        if (k + 2 < n) issue_load(k + 2);
      }
      if (n > 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        const uint8_t* s2; uint8_t* d2; uint32_t b2; int p2;
        locate(n - 1, s2, d2, b2, p2);
        asm volatile("fence.proxy.async;" ::: "memory");
        __threadfence();
        atomicAdd(p.ag_ready + p2, (unsigned long long)b2);
      }
    }
  } else if (warp >= 4) {
    // ================================ epilogue ================================
    const int e = warp - 4;  // == warp % 4 -> TMEM lane quadrant
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
      if (FUSE_AG) m_blk = ag_rotate(m_blk, p);
      const int row = m_blk * BM + e * 32 + lane;
      const int n0 = n_blk * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(e * 32) << 16) + acc * BN;
      const bool row_ok = row < p.M;
      const bool vec_ok = (p.ldc % 8 == 0);
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
                  float2 a0 = bf2_to_f2(old.x), a1 = bf2_to_f2(old.y), a2 = bf2_to_f2(old.z),
                         a3 = bf2_to_f2(old.w);
                  f[i] += a0.x; f[i + 1] += a0.y; f[i + 2] += a1.x; f[i + 3] += a1.y;
                  f[i + 4] += a2.x; f[i + 5] += a2.y; f[i + 6] += a3.x; f[i + 7] += a3.y;
                }
                uint4 o;
                o.x = f2_to_bf2(f[i], f[i + 1]);
                o.y = f2_to_bf2(f[i + 2], f[i + 3]);
                o.z = f2_to_bf2(f[i + 4], f[i + 5]);
                o.w = f2_to_bf2(f[i + 6], f[i + 7]);
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
      // hand the accumulator stage back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// --------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(ptr);
  return fn;
}

// 2-D bf16 tensor map: `inner` contiguous elements, `outer` rows of `row_stride` elements
int make_tmap_bf16_2d(CUtensorMap* out, const void* base, long long inner, long long outer,
                      long long row_stride, int box_inner, int box_outer) {
  // cuTensorMapEncodeTiled is a driver call: make sure this host thread (e.g. an autograd worker)
  // has the primary context bound, which runtime calls do lazily and driver calls do not.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(0);
    ctx_bound = true;
  }
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_DRIVER;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)row_stride * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "b200: cuTensorMapEncodeTiled failed (CUresult %d): base=%p inner=%lld outer=%lld stride=%lld box=%dx%d\n",
            (int)r, base, inner, outer, row_stride, box_inner, box_outer);
    return B200_ERR_DRIVER;
  }
  return 0;
}

// same for a row-major byte matrix (fp8 operands): 128-byte swizzled boxes of `box_inner` bytes
int make_tmap_u8_2d(CUtensorMap* out, const void* base, long long inner, long long outer, long long row_stride,
                    int box_inner, int box_outer) {
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(0);
    ctx_bound = true;
  }
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_DRIVER;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)row_stride};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "b200: cuTensorMapEncodeTiled(u8) failed (CUresult %d): base=%p inner=%lld outer=%lld stride=%lld box=%dx%d\n",
            (int)r, base, inner, outer, row_stride, box_inner, box_outer);
    return B200_ERR_DRIVER;
  }
  return 0;
}

// scale-factor blocks of the MXFP8 GEMM: `nblocks` rows of 512 bytes, loaded one row at a time, no swizzle
int make_tmap_sf_blocks(CUtensorMap* out, const void* base, long long nblocks) {
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(0);
    ctx_bound = true;
  }
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return B200_ERR_DRIVER;
  cuuint64_t dims[2] = {128, (cuuint64_t)nblocks};  // 128 x uint32 = 512 bytes per row
  cuuint64_t strides[1] = {512};
  cuuint32_t box[2] = {128, 1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "b200: cuTensorMapEncodeTiled(sf) failed (CUresult %d): base=%p nblocks=%lld\n", (int)r, base, nblocks);
    return B200_ERR_DRIVER;
  }
  return 0;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) cudaDeviceGetAttribute(&cached[dev], cudaDevAttrMultiProcessorCount, dev);
  return cached[dev];
}

template <int A_MN, int B_MN, int FUSE_AG = 0>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st,
                       int max_ctas) {
  static bool attr_set = false;
  constexpr int smem_bytes = GEMM_SMEM + (FUSE_AG ? AGG_SMEM : 0);
  if (!attr_set) {
    B200_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<A_MN, B_MN, FUSE_AG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   smem_bytes));
    attr_set = true;
  }
  int grid = p.num_m_blocks * p.num_n_blocks;
  int sms = num_sms();
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
  if (grid > sms) grid = sms;
  gemm_bf16_kernel<A_MN, B_MN, FUSE_AG><<<grid, 256, smem_bytes, st>>>(ta, tb, p);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

}  // namespace b200

// D[M,N] (+)= A x B^T (+ bias).  a_mn_major == 0: A is [M,K] with row stride lda; == 1: A is [K,M].
// b_mn_major == 0: B is [N,K] with row stride ldb; == 1: B is [K,N].  All bf16; D bf16 or fp32.
// max_ctas > 0 limits the persistent grid (leave SMs free for a concurrent comm kernel).
B200_API int b200_gemm_bf16(const void* A, const void* B, void* C, const void* bias, int M, int N, int K,
                            long long lda, long long ldb, long long ldc, int a_mn_major, int b_mn_major,
                            int out_fp32, int accumulate, int max_ctas, void* stream) {
  using namespace b200;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((lda % 8) || (ldb % 8) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return B200_ERR_BAD_ARG;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major)
    rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BK, BM);
  else
    rc = make_tmap_bf16_2d(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major)
    rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BN);
  else
    rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  GemmParams p;
  p.C = C;
  p.bias = bias;
  p.M = M;
  p.N = N;
  p.K = K;
  p.ldc = ldc;
  p.out_fp32 = out_fp32;
  p.accumulate = accumulate;
  p.num_m_blocks = (M + BM - 1) / BM;
  p.num_n_blocks = (N + BN - 1) / BN;
  p.rs_world = 0;
  p.ag_world = 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a_mn_major && !b_mn_major) return launch_gemm<0, 0>(ta, tb, p, st, max_ctas);
  if (!a_mn_major && b_mn_major) return launch_gemm<0, 1>(ta, tb, p, st, max_ctas);
  if (a_mn_major && !b_mn_major) return launch_gemm<1, 0>(ta, tb, p, st, max_ctas);
  return launch_gemm<1, 1>(ta, tb, p, st, max_ctas);
}


// ---- tensor-parallel fused GEMMs --------------------------------------

// All-gather -> GEMM in ONE kernel.  Every rank keeps its [rows, K] bf16 shard of A at byte offset
// a_off + rank * rows*K*2 of its symmetric buffer (peer_ptrs: host array of the mapped bases) and has already written
// its own shard there; `a_full` is the local [world*rows, K] assembly buffer whose own slot is also filled.
// `ready` is a local array of `world` zero-initialised uint64 counters, `epoch` the 1-based use count of them.
// D[world*rows, N] = A_full x B^T with B as in b200_gemm_bf16.
B200_API int b200_gemm_bf16_allgather_a(const void* const* peer_ptrs, int rank, int world, long long a_off, void* a_full,
                                        int rows, const void* B, void* C, int N, int K, long long ldb, long long ldc,
                                        int b_mn_major, void* ready, unsigned long long epoch, void* stream) {
  using namespace b200;
  if (world < 1 || world > 8 || rows <= 0 || (K % 8) || (ldb % 8)) return B200_ERR_BAD_ARG;
  const int M = rows * world;
  CUtensorMap ta, tb;
  int rc;
  if ((rc = make_tmap_bf16_2d(&ta, a_full, K, M, K, BK, BM))) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BN);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  GemmParams p;
  p.C = C; p.bias = nullptr; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.out_fp32 = 0; p.accumulate = 0;
  p.num_m_blocks = (M + BM - 1) / BM; p.num_n_blocks = (N + BN - 1) / BN;
  p.rs_world = 0;
  for (int i = 0; i < 8; ++i) p.ag_peer[i] = i < world ? (uint8_t*)peer_ptrs[i] : nullptr;
  p.ag_off = a_off; p.ag_full = (uint8_t*)a_full; p.ag_ready = (unsigned long long*)ready;
  p.ag_expect = epoch * (unsigned long long)rows * K * 2; p.ag_rows = rows; p.ag_rank = rank; p.ag_world = world;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!b_mn_major) return launch_gemm<0, 0, 1>(ta, tb, p, st, 0);
  return launch_gemm<0, 1, 1>(ta, tb, p, st, 0);
}

// GEMM -> reduce-scatter, first half: D = A x B^T is never written locally; the epilogue stores row r of this rank's
// partial straight into slot `rank` of the owner's [world][rows][N] bf16 buffer (byte offset out_off of its symmetric
// buffer) over NVLink.  After a cross-rank barrier, b200_reduce_slots folds the slots on the owner.
B200_API int b200_gemm_bf16_scatter(const void* A, const void* B, const void* const* peer_ptrs, int rank, int world,
                                    long long out_off, int rows, int N, int K, long long lda, long long ldb, int a_mn_major,
                                    int b_mn_major, void* stream) {
  using namespace b200;
  if (world < 1 || world > 8 || rows <= 0 || (N % 8) || (lda % 8) || (ldb % 8)) return B200_ERR_BAD_ARG;
  const int M = rows * world;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BK, BM);
  else rc = make_tmap_bf16_2d(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BK, BN);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  GemmParams p;
  p.C = nullptr; p.bias = nullptr; p.M = M; p.N = N; p.K = K; p.ldc = N; p.out_fp32 = 0; p.accumulate = 0;
  p.num_m_blocks = (M + BM - 1) / BM; p.num_n_blocks = (N + BN - 1) / BN;
  p.ag_world = 0;
  for (int i = 0; i < 8; ++i) p.rs_peer[i] = i < world ? (uint8_t*)peer_ptrs[i] : nullptr;
  p.rs_off = out_off; p.rs_rows = rows; p.rs_rank = rank; p.rs_world = world;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a_mn_major && !b_mn_major) return launch_gemm<0, 0>(ta, tb, p, st, 0);
  if (!a_mn_major && b_mn_major) return launch_gemm<0, 1>(ta, tb, p, st, 0);
  if (a_mn_major && !b_mn_major) return launch_gemm<1, 0>(ta, tb, p, st, 0);
  return launch_gemm<1, 1>(ta, tb, p, st, 0);
}

namespace b200 {
// out[r, :] = sum_s slots[s][r, :]   (bf16 in, fp32 accumulate, bf16 out)
__global__ void __launch_bounds__(256)
reduce_slots_kernel(const __nv_bfloat16* __restrict__ slots, __nv_bfloat16* __restrict__ out, int world, long long per_slot_vec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < per_slot_vec;
       i += (long long)gridDim.x * blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < world; ++s) {
      const uint4 v = ld_nc_v4(reinterpret_cast<const uint4*>(slots) + (long long)s * per_slot_vec + i);
      float2 t;
      t = bf2_to_f2(v.x); acc[0] += t.x; acc[1] += t.y;
      t = bf2_to_f2(v.y); acc[2] += t.x; acc[3] += t.y;
      t = bf2_to_f2(v.z); acc[4] += t.x; acc[5] += t.y;
      t = bf2_to_f2(v.w); acc[6] += t.x; acc[7] += t.y;
    }
    uint4 o;
    o.x = f2_to_bf2(acc[0], acc[1]); o.y = f2_to_bf2(acc[2], acc[3]);
    o.z = f2_to_bf2(acc[4], acc[5]); o.w = f2_to_bf2(acc[6], acc[7]);
    st_v4(reinterpret_cast<uint4*>(out) + i, o);
  }
}
}  // namespace b200

B200_API int b200_reduce_slots(const void* slots, void* out, int world, long long elems_per_slot, void* stream) {
  if (elems_per_slot % 8) return B200_ERR_BAD_ARG;
  if (elems_per_slot == 0) return 0;
  long long nv = elems_per_slot / 8;
  long long g = (nv + 255) / 256;
  long long cap = (long long)b200::num_sms() * 8;
  if (g > cap) g = cap;
  b200::reduce_slots_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)slots, (__nv_bfloat16*)out, world, nv);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
