// FP8 (e4m3) tensor quantisation with just-in-time per-tensor scaling, feeding the kind::f8f6f4 GEMM
// (gemm2_tcgen05.cu).  Counterpart of the reference's fp8 back-ends (reference src/accelerate/utils/transformer_engine.py
// convert_model / utils/ao.py convert_model_to_fp8_ao: they swap nn.Linear for TE / torchao modules); here the cast is
// two bandwidth-bound kernels:
//   1. b200_amax:           amax[0] = max |x|           (atomicMax on the bit pattern of a non-negative float)
//   2. b200_quantize_e4m3:  q = sat(x * 448/amax), optionally also q^T (for the dgrad GEMM, which needs the weight
//                           K-major along the *output* features), and deq[0] = amax/448 for the GEMM epilogue.
#include "common.cuh"
#include <cuda_fp8.h>

namespace b200 {

static constexpr float E4M3_MAX = 448.f;

__global__ void __launch_bounds__(256) amax_kernel(const __nv_bfloat16* __restrict__ x, long long n, float* __restrict__ amax) {
  float m = 0.f;
  const long long nv = n >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    const uint4 v = ld_nc_v4(reinterpret_cast<const uint4*>(x) + i);
    float2 t;
    t = bf2_to_f2(v.x); m = fmaxf(m, fmaxf(fabsf(t.x), fabsf(t.y)));
    t = bf2_to_f2(v.y); m = fmaxf(m, fmaxf(fabsf(t.x), fabsf(t.y)));
    t = bf2_to_f2(v.z); m = fmaxf(m, fmaxf(fabsf(t.x), fabsf(t.y)));
    t = bf2_to_f2(v.w); m = fmaxf(m, fmaxf(fabsf(t.x), fabsf(t.y)));
  }
  for (long long i = (nv << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    m = fmaxf(m, fabsf(__bfloat162float(x[i])));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  __shared__ float wm[8];
  if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 1; i < 8; ++i) m = fmaxf(m, wm[i]);
    if (m == m) atomicMax(reinterpret_cast<unsigned int*>(amax), __float_as_uint(m));  // NaN never wins
    else atomicMax(reinterpret_cast<unsigned int*>(amax), 0x7f800000u);                // but is reported as +inf
  }
}

__device__ __forceinline__ uint8_t to_e4m3(float v) {
  return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
}

// 64 x 64 tiles; q written row-major, qt (if any) written through a transposed smem tile so both stores coalesce
__global__ void __launch_bounds__(256)
quantize_e4m3_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int R, int C, const float* __restrict__ amax,
                     uint8_t* __restrict__ q, long long ldq, uint8_t* __restrict__ qt, long long ldqt, float* __restrict__ deq) {
  __shared__ uint8_t tile[64][64 + 4];
  const float a = fmaxf(*amax, 1e-12f);
  const float scale = (a < 3.0e38f) ? E4M3_MAX / a : 1.f;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) *deq = 1.f / scale;
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  // each thread converts 16 elements: row = tid / 4 (+0..63), 16 consecutive columns
  const int tr = threadIdx.x >> 2, tc = (threadIdx.x & 3) * 16;
  const int r = r0 + tr;
  uint8_t out[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) out[i] = 0;
  if (r < R) {
    const __nv_bfloat16* src = x + (long long)r * ldx + c0 + tc;
    if (c0 + tc + 16 <= C && (ldx % 8 == 0)) {
      const uint4 v0 = ld_nc_v4(reinterpret_cast<const uint4*>(src));
      const uint4 v1 = ld_nc_v4(reinterpret_cast<const uint4*>(src) + 1);
      const uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 t = bf2_to_f2(w[i]);
        out[2 * i] = to_e4m3(t.x * scale);
        out[2 * i + 1] = to_e4m3(t.y * scale);
      }
    } else {
      for (int i = 0; i < 16; ++i)
        if (c0 + tc + i < C) out[i] = to_e4m3(__bfloat162float(src[i]) * scale);
    }
    uint8_t* dst = q + (long long)r * ldq + c0 + tc;
    if (c0 + tc + 16 <= C && (ldq % 16 == 0)) {
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(out);
    } else {
      for (int i = 0; i < 16; ++i)
        if (c0 + tc + i < C) dst[i] = out[i];
    }
  }
  if (qt == nullptr) return;
#pragma unroll
  for (int i = 0; i < 16; ++i) tile[tc + i][tr] = out[i];  // tile[col][row]
  __syncthreads();
  const int c = c0 + tr;  // output row of qt = input column
  if (c < C) {
    uint8_t* dst = qt + (long long)c * ldqt + r0 + tc;
    if (r0 + tc + 16 <= R && (ldqt % 16 == 0)) {
      uint4 v;
      uint8_t* pv = reinterpret_cast<uint8_t*>(&v);
#pragma unroll
      for (int i = 0; i < 16; ++i) pv[i] = tile[tr][tc + i];
      *reinterpret_cast<uint4*>(dst) = v;
    } else {
      for (int i = 0; i < 16; ++i)
        if (r0 + tc + i < R) dst[i] = tile[tr][tc + i];
    }
  }
}

}  // namespace b200

using namespace b200;

// amax[0] = max(amax[0], max|x|) over n bf16 elements; the caller zeroes amax first (b200_memset_async or a fill)
B200_API int b200_amax_bf16(const void* x, long long n, void* amax, void* stream) {
  if (n <= 0) return 0;
  if (reinterpret_cast<uintptr_t>(x) & 15) return B200_ERR_BAD_ARG;
  long long blocks = (n / 8 + 255) / 256;
  const int cap = 148 * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  amax_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, n, (float*)amax);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// q[R,C] = e4m3(x * 448/amax), qt[C,R] = q^T when qt != NULL, deq[0] = amax/448
B200_API int b200_quantize_e4m3(const void* x, long long ldx, int R, int C, const void* amax, void* q, long long ldq, void* qt,
                                long long ldqt, void* deq, void* stream) {
  if (R <= 0 || C <= 0) return 0;
  dim3 grid((C + 63) / 64, (R + 63) / 64);
  quantize_e4m3_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ldx, R, C, (const float*)amax, (uint8_t*)q, ldq,
                                                              (uint8_t*)qt, ldqt, (float*)deq);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
