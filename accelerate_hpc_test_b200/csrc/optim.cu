// Fused AdamW (fp32 master weights + moments, bf16/fp32 grads, bf16 shadow params out),
// squared-gradient-norm reduction and the clip coefficient, all stream-ordered with no host sync.
// the reference calls torch.optim.AdamW (multi-tensor ATen) and
// torch.nn.utils.clip_grad_norm_ (reference src/accelerate/accelerator.py:2918-2979).
#include <cstdlib>
#include "common.cuh"

namespace b200 {
int num_sms();

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay;
  float bias_c1, bias_c2;  // 1 - beta1^t, 1 - beta2^t
  float grad_scale;        // multiplies every grad (1/GA, 1/loss_scale, ...)
};

template <typename GradT>
__device__ __forceinline__ float load_grad(const GradT* g, long long i);
template <>
__device__ __forceinline__ float load_grad<float>(const float* g, long long i) { return g[i]; }
template <>
__device__ __forceinline__ float load_grad<__nv_bfloat16>(const __nv_bfloat16* g, long long i) {
  return __bfloat162float(g[i]);
}

// one AdamW element update (decoupled weight decay, bias-corrected)
__device__ __forceinline__ void adamw_one(float& p, float& mi, float& vi, float g, const AdamArgs& a) {
  mi = mi * a.beta1 + (1.f - a.beta1) * g;
  vi = vi * a.beta2 + (1.f - a.beta2) * g * g;
  p *= (1.f - a.lr * a.weight_decay);
  p -= (a.lr / a.bias_c1) * (mi / (sqrtf(vi / a.bias_c2) + a.eps));
}

__device__ __forceinline__ void load_grad4(const float* g, long long i, float (&o)[4]) {
  const float4 t = __ldcs(reinterpret_cast<const float4*>(g + i));
  o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
}
__device__ __forceinline__ void load_grad4(const __nv_bfloat16* g, long long i, float (&o)[4]) {
  const uint2 t = __ldcs(reinterpret_cast<const uint2*>(g + i));
  const float2 a = bf2_to_f2(t.x), b = bf2_to_f2(t.y);
  o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}

// AdamW over elements [begin, end) of one tensor, shared by `nthreads` threads (this thread is `tid`).
// The optimizer state is touched exactly once per step, so everything goes through 16-byte streaming
// loads / stores (28 bytes of traffic per element: the kernel is pure HBM bandwidth).
template <typename GradT>
__device__ __forceinline__ void adamw_span(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                                           const GradT* __restrict__ grad, __nv_bfloat16* __restrict__ shadow,
                                           long long begin, long long end, long long tid, long long nthreads,
                                           const AdamArgs& a, float gs) {
  const uintptr_t mis = reinterpret_cast<uintptr_t>(master + begin) | reinterpret_cast<uintptr_t>(m + begin) |
                        reinterpret_cast<uintptr_t>(v + begin);
  const bool vec = ((mis & 15) == 0) && ((reinterpret_cast<uintptr_t>(grad + begin) & (sizeof(GradT) * 4 - 1)) == 0) &&
                   (shadow == nullptr || (reinterpret_cast<uintptr_t>(shadow + begin) & 7) == 0);
  long long i = begin;
  if (vec) {
    const long long nvec = (end - begin) >> 2;
    for (long long k = tid; k < nvec; k += nthreads) {
      const long long j = begin + (k << 2);
      float4 p4 = __ldcs(reinterpret_cast<const float4*>(master + j));
      float4 m4 = __ldcs(reinterpret_cast<const float4*>(m + j));
      float4 v4 = __ldcs(reinterpret_cast<const float4*>(v + j));
      float g[4];
      load_grad4(grad, j, g);
      adamw_one(p4.x, m4.x, v4.x, g[0] * gs, a);
      adamw_one(p4.y, m4.y, v4.y, g[1] * gs, a);
      adamw_one(p4.z, m4.z, v4.z, g[2] * gs, a);
      adamw_one(p4.w, m4.w, v4.w, g[3] * gs, a);
      __stcs(reinterpret_cast<float4*>(master + j), p4);
      __stcs(reinterpret_cast<float4*>(m + j), m4);
      __stcs(reinterpret_cast<float4*>(v + j), v4);
      if (shadow) {
        uint2 o;
        o.x = f2_to_bf2(p4.x, p4.y);
        o.y = f2_to_bf2(p4.z, p4.w);
        *reinterpret_cast<uint2*>(shadow + j) = o;  // the bf16 parameter is read again soon: normal store
      }
    }
    i = begin + (nvec << 2);
  }
  for (long long j = i + tid; j < end; j += nthreads) {
    float p = master[j], mi = m[j], vi = v[j];
    adamw_one(p, mi, vi, load_grad<GradT>(grad, j) * gs, a);
    master[j] = p; m[j] = mi; v[j] = vi;
    if (shadow) shadow[j] = __float2bfloat16(p);
  }
}

// flat AdamW over n elements.  clip_coef (device scalar, may be null) multiplies grads;
// found_inf (device scalar, may be null) != 0 skips the whole update (GradScaler semantics).
template <typename GradT>
__global__ void __launch_bounds__(256)
adamw_flat_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                  const GradT* __restrict__ grad, __nv_bfloat16* __restrict__ shadow, long long n,
                  AdamArgs a, const float* __restrict__ clip_coef, const float* __restrict__ found_inf) {
  if (found_inf && *found_inf != 0.f) return;
  const float gs = a.grad_scale * (clip_coef ? *clip_coef : 1.f);
  adamw_span<GradT>(master, m, v, grad, shadow, 0, n, (long long)blockIdx.x * blockDim.x + threadIdx.x,
                    (long long)gridDim.x * blockDim.x, a, gs);
}

// multi-tensor variant: tensor table in device memory, one CTA per (tensor, 64K-element chunk)
struct MTEntry {
  float* master;
  float* m;
  float* v;
  const void* grad;
  __nv_bfloat16* shadow;
  long long n;
  long long chunk_begin;  // first global chunk index of this tensor
};
static constexpr int MT_CHUNK = 65536;

template <typename GradT>
__global__ void __launch_bounds__(256)
adamw_mt_kernel(const MTEntry* __restrict__ table, int ntensors, AdamArgs a,
                const float* __restrict__ clip_coef, const float* __restrict__ found_inf) {
  if (found_inf && *found_inf != 0.f) return;
  // binary search the tensor owning this chunk
  const long long chunk = blockIdx.x;
  int lo = 0, hi = ntensors - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (table[mid].chunk_begin <= chunk) lo = mid; else hi = mid - 1;
  }
  const MTEntry e = table[lo];
  const long long begin = (chunk - e.chunk_begin) * MT_CHUNK;
  const long long end = min(begin + (long long)MT_CHUNK, e.n);
  const float gs = a.grad_scale * (clip_coef ? *clip_coef : 1.f);
  const GradT* grad = reinterpret_cast<const GradT*>(e.grad);
  adamw_span<GradT>(e.master, e.m, e.v, grad, e.shadow, begin, end, threadIdx.x, blockDim.x, a, gs);
}

// out[0] += sum(g^2) ; also flags non-finite values in out[1]
template <typename GradT>
__global__ void __launch_bounds__(256)
sumsq_kernel(const GradT* __restrict__ g, long long n, float* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.f;
  bool bad = false;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  long long i0 = 0;
  if ((reinterpret_cast<uintptr_t>(g) & (sizeof(GradT) * 4 - 1)) == 0) {
    const long long nvec = n >> 2;
    for (long long k = tid; k < nvec; k += nth) {
      float x[4];
      load_grad4(g, k << 2, x);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s += x[j] * x[j];
        bad |= !isfinite(x[j]);
      }
    }
    i0 = nvec << 2;
  }
  for (long long i = i0 + tid; i < n; i += nth) {
    const float x = load_grad<GradT>(g, i);
    s += x * x;
    bad |= !isfinite(x);
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) atomicAdd(out, s);
  if (bad) out[1] = 1.f;
}

// norm_out[0] = sqrt(sumsq) ; coef_out[0] = min(1, max_norm / (norm + 1e-6))
__global__ void clip_coef_kernel(const float* __restrict__ sumsq, float max_norm, float* __restrict__ norm_out,
                                 float* __restrict__ coef_out) {
  const float norm = sqrtf(sumsq[0]);
  norm_out[0] = norm;
  const float c = max_norm / (norm + 1e-6f);
  coef_out[0] = c < 1.f ? c : 1.f;
}

}  // namespace b200

using namespace b200;

static inline int flat_grid(long long n) {
  long long g = (n / 4 + 255) / 256;  // 4 elements per thread per iteration
  long long cap = (long long)num_sms() * 8;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

// The update usually runs on a side stream beside the backward GEMMs (optimizer in backward): how many CTAs per SM it may take
// is a trade between its own bandwidth and the SM slots / issue cycles it takes from them.  ACCELERATE_B200_ADAMW_CTAS_PER_SM.
static inline int adamw_grid(long long n) {
  static int per_sm = 0;
  if (per_sm == 0) {
    const char* e = getenv("ACCELERATE_B200_ADAMW_CTAS_PER_SM");
    per_sm = e ? atoi(e) : 4;  // 1 GPU, 8B model: 212.5 / 209.9 / 215.3 / 230.9 ms per step with 8 / 4 / 2 / 1
    if (per_sm < 1 || per_sm > 8) per_sm = 8;
  }
  long long g = (n / 4 + 255) / 256;
  long long cap = (long long)num_sms() * per_sm;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

B200_API int b200_adamw_flat(void* master, void* m, void* v, const void* grad, void* shadow_bf16,
                             long long n, int grad_is_bf16, float lr, float beta1, float beta2, float eps,
                             float weight_decay, int step, float grad_scale, const void* clip_coef,
                             const void* found_inf, void* stream) {
  if (n == 0) return 0;
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step),
             1.f - powf(beta2, (float)step), grad_scale};
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_is_bf16)
    adamw_flat_kernel<__nv_bfloat16><<<adamw_grid(n), 256, 0, st>>>(
        (float*)master, (float*)m, (float*)v, (const __nv_bfloat16*)grad, (__nv_bfloat16*)shadow_bf16, n, a,
        (const float*)clip_coef, (const float*)found_inf);
  else
    adamw_flat_kernel<float><<<adamw_grid(n), 256, 0, st>>>((float*)master, (float*)m, (float*)v,
                                                           (const float*)grad, (__nv_bfloat16*)shadow_bf16,
                                                           n, a, (const float*)clip_coef,
                                                           (const float*)found_inf);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// table: device array of ntensors MTEntry records (7 x 8 bytes each, see MTEntry); total_chunks =
// sum(ceil(n_i / 65536)).  The Python side builds the table once per optimizer and re-uses it.
B200_API int b200_adamw_multi(const void* table, int ntensors, long long total_chunks, int grad_is_bf16,
                              float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                              float grad_scale, const void* clip_coef, const void* found_inf, void* stream) {
  if (ntensors == 0 || total_chunks == 0) return 0;
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step),
             1.f - powf(beta2, (float)step), grad_scale};
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_is_bf16)
    adamw_mt_kernel<__nv_bfloat16><<<(unsigned)total_chunks, 256, 0, st>>>(
        (const MTEntry*)table, ntensors, a, (const float*)clip_coef, (const float*)found_inf);
  else
    adamw_mt_kernel<float><<<(unsigned)total_chunks, 256, 0, st>>>((const MTEntry*)table, ntensors, a,
                                                                   (const float*)clip_coef,
                                                                   (const float*)found_inf);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// out (2 floats, caller-zeroed before the first call of a step): out[0] += sum g^2, out[1] = inf flag
B200_API int b200_sumsq(const void* g, long long n, int is_bf16, void* out2, void* stream) {
  if (n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    sumsq_kernel<__nv_bfloat16><<<flat_grid(n), 256, 0, st>>>((const __nv_bfloat16*)g, n, (float*)out2);
  else
    sumsq_kernel<float><<<flat_grid(n), 256, 0, st>>>((const float*)g, n, (float*)out2);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_clip_coef(const void* sumsq, float max_norm, void* norm_out, void* coef_out, void* stream) {
  clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((const float*)sumsq, max_norm, (float*)norm_out,
                                                      (float*)coef_out);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
