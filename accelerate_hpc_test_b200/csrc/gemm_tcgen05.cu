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
};

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int per_group = GROUP_M * num_n;
  const int group = tile / per_group;
  const int first_m = group * GROUP_M;
  const int gm = min(GROUP_M, num_m - first_m);
  const int r = tile - group * per_group;
  m_blk = first_m + (r % gm);
  n_blk = r / gm;
}

template <int A_MN, int B_MN>
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
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
        const int m0 = m_blk * BM, n0 = n_blk * BN;
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
    if (lane == 0) {
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
  } else if (warp >= 4) {
    // ================================ epilogue ================================
    const int e = warp - 4;  // == warp % 4 -> TMEM lane quadrant
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

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) cudaDeviceGetAttribute(&cached[dev], cudaDevAttrMultiProcessorCount, dev);
  return cached[dev];
}

template <int A_MN, int B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st,
                       int max_ctas) {
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   GEMM_SMEM));
    attr_set = true;
  }
  int grid = p.num_m_blocks * p.num_n_blocks;
  int sms = num_sms();
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
  if (grid > sms) grid = sms;
  gemm_bf16_kernel<A_MN, B_MN><<<grid, 256, GEMM_SMEM, st>>>(ta, tb, p);
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
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a_mn_major && !b_mn_major) return launch_gemm<0, 0>(ta, tb, p, st, max_ctas);
  if (!a_mn_major && b_mn_major) return launch_gemm<0, 1>(ta, tb, p, st, max_ctas);
  if (a_mn_major && !b_mn_major) return launch_gemm<1, 0>(ta, tb, p, st, max_ctas);
  return launch_gemm<1, 1>(ta, tb, p, st, max_ctas);
}
