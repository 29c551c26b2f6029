// Cross-entropy over a chunk of bf16 logits, forward and backward in one launch:
//   loss_row = logsumexp(z) - z[label];   z <- (softmax(z) - onehot(label)) * (scale * inv_count)
// The logits chunk is produced by the lm_head GEMM and immediately overwritten with its own
// gradient, so full [tokens, vocab] logits (2.1 GB bf16 at 8192 tokens, twice that after the
// reference's fp32 upcast in convert_outputs_to_fp32, reference src/accelerate/accelerator.py:
// 1797-1802) never exist. SURVEY.md §2.6 row K8 / §5.7.
#include "common.cuh"

namespace b200 {

// one CTA per row. Pass 1: online max/sum (single read); pass 2: re-read (L2-resident) and write grad.
__global__ void __launch_bounds__(1024)
ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ labels,
                  float* __restrict__ loss_rows, float* __restrict__ lse_rows, long long ld, int V,
                  const float* __restrict__ inv_count, float scale, long long ignore_index,
                  int write_grad) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  __nv_bfloat16* z = logits + row * ld;
  const long long label = labels[row];
  const bool valid = (label != ignore_index) && label >= 0 && label < V;
  const int nvec = V >> 3;       // whole 16-byte vectors of the row (the row pitch `ld` is a multiple of 8 elements)
  const int tail0 = nvec << 3;   // the last V % 8 elements (odd vocabularies, e.g. 50257) are handled one per thread
  const uint4* zv = reinterpret_cast<const uint4*>(z);

  float mx = -INFINITY, sum = 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 u = ld_v4(zv + i);
    const float2 a = bf2_to_f2(u.x), b = bf2_to_f2(u.y), c = bf2_to_f2(u.z), d = bf2_to_f2(u.w);
    const float lm = fmaxf(fmaxf(fmaxf(a.x, a.y), fmaxf(b.x, b.y)), fmaxf(fmaxf(c.x, c.y), fmaxf(d.x, d.y)));
    if (lm > mx) {
      sum *= __expf(mx - lm);
      mx = lm;
    }
    sum += __expf(a.x - mx) + __expf(a.y - mx) + __expf(b.x - mx) + __expf(b.y - mx) + __expf(c.x - mx) +
           __expf(c.y - mx) + __expf(d.x - mx) + __expf(d.y - mx);
  }
  if ((int)threadIdx.x < V - tail0) {
    const float a = __bfloat162float(z[tail0 + threadIdx.x]);
    if (a > mx) {
      sum *= __expf(mx - a);
      mx = a;
    }
    sum += __expf(a - mx);
  }
  const float gmx = block_max(mx, red);
  sum *= (mx == -INFINITY) ? 0.f : __expf(mx - gmx);
  const float gsum = block_sum(sum, red);
  const float lse = gmx + __logf(gsum);
  if (threadIdx.x == 0) {
    const float zl = valid ? __bfloat162float(z[label]) : 0.f;
    loss_rows[row] = valid ? (lse - zl) : 0.f;
    if (lse_rows) lse_rows[row] = lse;
  }
  if (!write_grad) return;
  __syncthreads();  // z[label] read above must precede the overwrite below
  const float g = valid ? scale * (inv_count ? *inv_count : 1.f) : 0.f;
  uint4* zo = reinterpret_cast<uint4*>(z);
  const int lvec = valid ? (int)(label >> 3) : -1;
  const int lsub = (int)(label & 7);
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 u = ld_v4(zv + i);
    float f[8];
    float2 t;
    t = bf2_to_f2(u.x); f[0] = t.x; f[1] = t.y;
    t = bf2_to_f2(u.y); f[2] = t.x; f[3] = t.y;
    t = bf2_to_f2(u.z); f[4] = t.x; f[5] = t.y;
    t = bf2_to_f2(u.w); f[6] = t.x; f[7] = t.y;
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = __expf(f[k] - lse) * g;
    if (i == lvec) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k == lsub) f[k] -= g;
    }
    uint4 o;
    o.x = f2_to_bf2(f[0], f[1]); o.y = f2_to_bf2(f[2], f[3]);
    o.z = f2_to_bf2(f[4], f[5]); o.w = f2_to_bf2(f[6], f[7]);
    st_v4(zo + i, o);
  }
  if ((int)threadIdx.x < V - tail0) {
    const int idx = tail0 + threadIdx.x;
    float f = __expf(__bfloat162float(z[idx]) - lse) * g;
    if (valid && idx == label) f -= g;
    z[idx] = __float2bfloat16(f);
  }
}

// counts labels != ignore_index and writes 1/max(count,1) (mean reduction denominator)
__global__ void __launch_bounds__(1024)
count_valid_kernel(const long long* __restrict__ labels, long long n, long long ignore_index,
                   float* __restrict__ inv_count, float* __restrict__ count_out) {
  __shared__ float red[32];
  float c = 0.f;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) c += (labels[i] != ignore_index) ? 1.f : 0.f;
  c = block_sum(c, red);
  if (threadIdx.x == 0) {
    if (count_out) count_out[0] = c;
    inv_count[0] = 1.f / fmaxf(c, 1.f);
  }
}

// out[0] (+)= sum(rows) * (*inv_count) * scale  (single CTA, deterministic)
__global__ void __launch_bounds__(1024)
loss_reduce_kernel(const float* __restrict__ rows, long long n, const float* __restrict__ inv_count,
                   float scale, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float s = 0.f;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) s += rows[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) {
    const float r = s * (inv_count ? *inv_count : 1.f) * scale;
    out[0] = accumulate ? out[0] + r : r;
  }
}

}  // namespace b200

using namespace b200;

B200_API int b200_ce_fwd_bwd(void* logits, const void* labels, void* loss_rows, void* lse_rows,
                             long long T, long long ld, int V, const void* inv_count, float scale,
                             long long ignore_index, int write_grad, void* stream) {
  if (ld % 8 || V > ld) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  ce_fwd_bwd_kernel<<<(unsigned)T, 1024, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)logits, (const long long*)labels, (float*)loss_rows, (float*)lse_rows, ld, V,
      (const float*)inv_count, scale, ignore_index, write_grad);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_count_valid(const void* labels, long long n, long long ignore_index, void* inv_count,
                              void* count_out, void* stream) {
  count_valid_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((const long long*)labels, n, ignore_index,
                                                           (float*)inv_count, (float*)count_out);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_loss_reduce(const void* rows, long long n, const void* inv_count, float scale, void* out,
                              int accumulate, void* stream) {
  loss_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((const float*)rows, n, (const float*)inv_count,
                                                           scale, (float*)out, accumulate);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
