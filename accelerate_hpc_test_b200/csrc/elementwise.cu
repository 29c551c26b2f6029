// Bandwidth-bound fused ops of the transformer block, hand-written for sm_100a:
// RMSNorm (+residual add) fwd/bwd, RoPE (in place on the fused QKV buffer), SwiGLU fwd/bwd,
// embedding gather / scatter-add, casts and scaled copies.
// in the reference these run as chains of eager ATen
// kernels inside transformers' LlamaRMSNorm / rotate_half / silu*up; here each is one pass.
#include "common.cuh"

namespace b200 {

// ------------------------------------------------------------------------------- RMSNorm fwd
// s = x (+ res); y = s * rsqrt(mean(s^2) + eps) * w.  One CTA per row; H % 8 == 0.
__global__ void __launch_bounds__(256)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                   const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y,
                   __nv_bfloat16* __restrict__ sum_out, float* __restrict__ rstd_out, int H, float eps) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * H);
  const uint4* rr = res ? reinterpret_cast<const uint4*>(res + row * H) : nullptr;
  uint4* sr = sum_out ? reinterpret_cast<uint4*>(sum_out + row * H) : nullptr;
  const int nvec = H >> 3;
  // cache up to 4 vectors per thread in registers (H <= 8192 with 256 threads)
  uint4 cache[4];
  float ss = 0.f;
  int it = 0;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x, ++it) {
    uint4 v = ld_v4(xr + i);
    float2 a = bf2_to_f2(v.x), b = bf2_to_f2(v.y), c = bf2_to_f2(v.z), d = bf2_to_f2(v.w);
    if (rr) {
      uint4 r = ld_v4(rr + i);
      float2 ra = bf2_to_f2(r.x), rb = bf2_to_f2(r.y), rc = bf2_to_f2(r.z), rd = bf2_to_f2(r.w);
      a.x += ra.x; a.y += ra.y; b.x += rb.x; b.y += rb.y;
      c.x += rc.x; c.y += rc.y; d.x += rd.x; d.y += rd.y;
      v.x = f2_to_bf2(a.x, a.y); v.y = f2_to_bf2(b.x, b.y);
      v.z = f2_to_bf2(c.x, c.y); v.w = f2_to_bf2(d.x, d.y);
      // the normalised value must see the same rounding the saved sum has
      a = bf2_to_f2(v.x); b = bf2_to_f2(v.y); c = bf2_to_f2(v.z); d = bf2_to_f2(v.w);
    }
    if (sr) st_v4(sr + i, v);
    if (it < 4) cache[it] = v;
    ss += a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y + c.x * c.x + c.y * c.y + d.x * d.x + d.y * d.y;
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / (float)H + eps);
  if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
  const uint4* wr = reinterpret_cast<const uint4*>(w);
  uint4* yr = reinterpret_cast<uint4*>(y + row * H);
  it = 0;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x, ++it) {
    uint4 v;
    if (it < 4) v = cache[it];
    else if (sr) v = ld_v4(sr + i);
    else {
      v = ld_v4(xr + i);  // (res == null whenever sum_out == null is enforced by the host wrapper)
    }
    uint4 g = __ldg(wr + i);
    float2 a = bf2_to_f2(v.x), b = bf2_to_f2(v.y), c = bf2_to_f2(v.z), d = bf2_to_f2(v.w);
    float2 ga = bf2_to_f2(g.x), gb = bf2_to_f2(g.y), gc = bf2_to_f2(g.z), gd = bf2_to_f2(g.w);
    uint4 o;
    o.x = f2_to_bf2(a.x * rstd * ga.x, a.y * rstd * ga.y);
    o.y = f2_to_bf2(b.x * rstd * gb.x, b.y * rstd * gb.y);
    o.z = f2_to_bf2(c.x * rstd * gc.x, c.y * rstd * gc.y);
    o.w = f2_to_bf2(d.x * rstd * gd.x, d.y * rstd * gd.y);
    st_v4(yr + i, o);
  }
}

// ------------------------------------------------------------------------------- RMSNorm bwd
// dx = rstd * (dy*w - xhat * mean(dy*w*xhat)) (+ dres);  dw_partial[b] = sum_rows dy * xhat.
// Persistent over rows: CTA b handles rows b, b+grid, ... and keeps its dw partial in registers;
// partials land in `dw_partial` [grid, H] fp32 (scratch) and are folded by rmsnorm_dw_reduce.
template <int VPT>  // uint4 vectors per thread; H == VPT * 8 * blockDim.x exactly covered or less
__global__ void __launch_bounds__(256)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd,
                   const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                   float* __restrict__ dw_partial, int T, int H) {
  __shared__ float red[32];
  const int nvec = H >> 3;
  float dwacc[VPT][8];
#pragma unroll
  for (int j = 0; j < VPT; ++j)
#pragma unroll
    for (int k = 0; k < 8; ++k) dwacc[j][k] = 0.f;
  float wreg[VPT][8];
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < nvec) {
      uint4 g = __ldg(reinterpret_cast<const uint4*>(w) + i);
      float2 a = bf2_to_f2(g.x), b = bf2_to_f2(g.y), c = bf2_to_f2(g.z), d = bf2_to_f2(g.w);
      wreg[j][0] = a.x; wreg[j][1] = a.y; wreg[j][2] = b.x; wreg[j][3] = b.y;
      wreg[j][4] = c.x; wreg[j][5] = c.y; wreg[j][6] = d.x; wreg[j][7] = d.y;
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) wreg[j][k] = 0.f;
    }
  }
  for (long long row = blockIdx.x; row < T; row += gridDim.x) {
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + row * H);
    const uint4* xr = reinterpret_cast<const uint4*>(x + row * H);
    const float rs = rstd[row];
    float dyv[VPT][8], xh[VPT][8];
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int i = threadIdx.x + j * blockDim.x;
      if (i < nvec) {
        uint4 a = ld_nc_v4(dyr + i), b = ld_nc_v4(xr + i);
        float2 t;
        t = bf2_to_f2(a.x); dyv[j][0] = t.x; dyv[j][1] = t.y;
        t = bf2_to_f2(a.y); dyv[j][2] = t.x; dyv[j][3] = t.y;
        t = bf2_to_f2(a.z); dyv[j][4] = t.x; dyv[j][5] = t.y;
        t = bf2_to_f2(a.w); dyv[j][6] = t.x; dyv[j][7] = t.y;
        t = bf2_to_f2(b.x); xh[j][0] = t.x * rs; xh[j][1] = t.y * rs;
        t = bf2_to_f2(b.y); xh[j][2] = t.x * rs; xh[j][3] = t.y * rs;
        t = bf2_to_f2(b.z); xh[j][4] = t.x * rs; xh[j][5] = t.y * rs;
        t = bf2_to_f2(b.w); xh[j][6] = t.x * rs; xh[j][7] = t.y * rs;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          dot += dyv[j][k] * wreg[j][k] * xh[j][k];
          dwacc[j][k] += dyv[j][k] * xh[j][k];
        }
      }
    }
    dot = block_sum(dot, red);
    const float mean = dot / (float)H;
    uint4* dxr = reinterpret_cast<uint4*>(dx + row * H);
    const uint4* drr = dres ? reinterpret_cast<const uint4*>(dres + row * H) : nullptr;
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int i = threadIdx.x + j * blockDim.x;
      if (i < nvec) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rs * (dyv[j][k] * wreg[j][k] - xh[j][k] * mean);
        if (drr) {
          uint4 r = ld_nc_v4(drr + i);
          float2 t;
          t = bf2_to_f2(r.x); o[0] += t.x; o[1] += t.y;
          t = bf2_to_f2(r.y); o[2] += t.x; o[3] += t.y;
          t = bf2_to_f2(r.z); o[4] += t.x; o[5] += t.y;
          t = bf2_to_f2(r.w); o[6] += t.x; o[7] += t.y;
        }
        uint4 ov;
        ov.x = f2_to_bf2(o[0], o[1]); ov.y = f2_to_bf2(o[2], o[3]);
        ov.z = f2_to_bf2(o[4], o[5]); ov.w = f2_to_bf2(o[6], o[7]);
        st_v4(dxr + i, ov);
      }
    }
  }
  float* part = dw_partial + (long long)blockIdx.x * H;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < nvec) {
      float4* dst = reinterpret_cast<float4*>(part + i * 8);
      dst[0] = make_float4(dwacc[j][0], dwacc[j][1], dwacc[j][2], dwacc[j][3]);
      dst[1] = make_float4(dwacc[j][4], dwacc[j][5], dwacc[j][6], dwacc[j][7]);
    }
  }
}

// dw[h] (+)= sum_b partial[b,h]
__global__ void rmsnorm_dw_reduce_kernel(const float* __restrict__ partial, __nv_bfloat16* __restrict__ dw,
                                         int nparts, int H, int accumulate) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= H) return;
  float s = 0.f;
  for (int b = 0; b < nparts; ++b) s += partial[(long long)b * H + h];
  if (accumulate) s += __bfloat162float(dw[h]);
  dw[h] = __float2bfloat16(s);
}

// ------------------------------------------------------------------------------- RoPE
// In place on rows of `buf` ([T, row_stride]); rotates heads 0..n_heads-1 of width D laid out
// back to back (the Q then K slots of the fused QKV projection output).  HF "rotate_half"
// convention: pairs (i, i + D/2).  sign = +1 forward, -1 backward (inverse rotation).
__global__ void __launch_bounds__(256)
rope_kernel(__nv_bfloat16* __restrict__ buf, const float* __restrict__ cos_tab,
            const float* __restrict__ sin_tab, const int* __restrict__ pos, long long row_stride,
            int n_heads, int D, float sign, long long total_pairs2) {
  // each thread handles 2 adjacent pairs (bf16x2 at i and at i + D/2)
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_pairs2) return;
  const int half = D >> 1;
  const int per_head = half >> 1;
  const int per_row = n_heads * per_head;
  const long long t = idx / per_row;
  const int r = (int)(idx - t * per_row);
  const int h = r / per_head;
  const int i = (r - h * per_head) * 2;
  const int p = pos[t];
  const float2 c = *reinterpret_cast<const float2*>(cos_tab + (long long)p * half + i);
  float2 s = *reinterpret_cast<const float2*>(sin_tab + (long long)p * half + i);
  s.x *= sign; s.y *= sign;
  __nv_bfloat16* base = buf + t * row_stride + (long long)h * D;
  uint32_t* lo = reinterpret_cast<uint32_t*>(base + i);
  uint32_t* hi = reinterpret_cast<uint32_t*>(base + i + half);
  const float2 a = bf2_to_f2(*lo), b = bf2_to_f2(*hi);
  *lo = f2_to_bf2(a.x * c.x - b.x * s.x, a.y * c.y - b.y * s.y);
  *hi = f2_to_bf2(b.x * c.x + a.x * s.x, b.y * c.y + a.y * s.y);
}

// ------------------------------------------------------------------------------- SwiGLU
// gu = [gate | up] of shape [T, 2I] (fused gate/up projection output); act = silu(gate) * up
__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act, long long T,
                  int I) {
  const int nvec = I >> 3;
  const long long total = T * nvec;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / nvec;
    const int i = (int)(idx - t * nvec);
    const uint4 g = ld_nc_v4(reinterpret_cast<const uint4*>(gu + t * 2 * I) + i);
    const uint4 u = ld_nc_v4(reinterpret_cast<const uint4*>(gu + t * 2 * I + I) + i);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w};
    uint32_t ow[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float2 a = bf2_to_f2(gw[k]), b = bf2_to_f2(uw[k]);
      float s0 = a.x / (1.f + __expf(-a.x)), s1 = a.y / (1.f + __expf(-a.y));
      ow[k] = f2_to_bf2(s0 * b.x, s1 * b.y);
    }
    st_v4(reinterpret_cast<uint4*>(act + t * I) + i, make_uint4(ow[0], ow[1], ow[2], ow[3]));
  }
}

// dgu = [dact * up * dsilu(gate) | dact * silu(gate)]
__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ dact,
                  __nv_bfloat16* __restrict__ dgu, long long T, int I) {
  const int nvec = I >> 3;
  const long long total = T * nvec;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / nvec;
    const int i = (int)(idx - t * nvec);
    const uint4 g = ld_nc_v4(reinterpret_cast<const uint4*>(gu + t * 2 * I) + i);
    const uint4 u = ld_nc_v4(reinterpret_cast<const uint4*>(gu + t * 2 * I + I) + i);
    const uint4 d = ld_nc_v4(reinterpret_cast<const uint4*>(dact + t * I) + i);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w}, dw[4] = {d.x, d.y, d.z, d.w};
    uint32_t og[4], ou[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float2 a = bf2_to_f2(gw[k]), b = bf2_to_f2(uw[k]), c = bf2_to_f2(dw[k]);
      float sg0 = 1.f / (1.f + __expf(-a.x)), sg1 = 1.f / (1.f + __expf(-a.y));
      float si0 = a.x * sg0, si1 = a.y * sg1;
      float ds0 = sg0 * (1.f + a.x * (1.f - sg0)), ds1 = sg1 * (1.f + a.y * (1.f - sg1));
      og[k] = f2_to_bf2(c.x * b.x * ds0, c.y * b.y * ds1);
      ou[k] = f2_to_bf2(c.x * si0, c.y * si1);
    }
    st_v4(reinterpret_cast<uint4*>(dgu + t * 2 * I) + i, make_uint4(og[0], og[1], og[2], og[3]));
    st_v4(reinterpret_cast<uint4*>(dgu + t * 2 * I + I) + i, make_uint4(ou[0], ou[1], ou[2], ou[3]));
  }
}

// ------------------------------------------------------------------------------- embedding
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ table,
                     __nv_bfloat16* __restrict__ out, long long T, int H, long long vocab) {
  const int nvec = H >> 3;
  const long long total = T * nvec;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / nvec;
    const int i = (int)(idx - t * nvec);
    long long id = ids[t];
    uint4 v = make_uint4(0, 0, 0, 0);
    if (id >= 0 && id < vocab) v = ld_v4(reinterpret_cast<const uint4*>(table + id * H) + i);
    st_v4(reinterpret_cast<uint4*>(out + t * H) + i, v);
  }
}

// dtable[ids[t]] += dout[t]  (bf16x2 atomics; dtable pre-zeroed by the caller unless accumulating)
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ dout,
                     __nv_bfloat16* __restrict__ dtable, long long T, int H, long long vocab) {
  const int npair = H >> 1;
  const long long total = T * npair;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / npair;
    const int i = (int)(idx - t * npair);
    const long long id = ids[t];
    if (id < 0 || id >= vocab) continue;
    const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(dout + t * H + 2 * i);
    atomicAdd(reinterpret_cast<__nv_bfloat162*>(dtable + id * H + 2 * i), v);
  }
}

// ------------------------------------------------------------------------------- casts
// out = (OutT)(in * scale) ; covers fp32<->bf16 casts and the grad / GA division (K11)
template <typename InT, typename OutT>
__global__ void __launch_bounds__(256)
scale_cast_kernel(const InT* __restrict__ in, OutT* __restrict__ out, long long n, float scale) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = static_cast<OutT>(static_cast<float>(in[i]) * scale);
}

// y += a (bf16), vectorised; used for residual-gradient accumulation
__global__ void __launch_bounds__(256)
add_inplace_bf16_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ a, long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    uint4 u = ld_v4(reinterpret_cast<const uint4*>(y) + i), v = ld_nc_v4(reinterpret_cast<const uint4*>(a) + i);
    float2 a0 = bf2_to_f2(u.x), a1 = bf2_to_f2(u.y), a2 = bf2_to_f2(u.z), a3 = bf2_to_f2(u.w);
    float2 b0 = bf2_to_f2(v.x), b1 = bf2_to_f2(v.y), b2 = bf2_to_f2(v.z), b3 = bf2_to_f2(v.w);
    uint4 o;
    o.x = f2_to_bf2(a0.x + b0.x, a0.y + b0.y); o.y = f2_to_bf2(a1.x + b1.x, a1.y + b1.y);
    o.z = f2_to_bf2(a2.x + b2.x, a2.y + b2.y); o.w = f2_to_bf2(a3.x + b3.x, a3.y + b3.y);
    st_v4(reinterpret_cast<uint4*>(y) + i, o);
  }
}

// out (+)= in * (*scale): the scale lives on the device (an upstream gradient, a clip coefficient), so no host sync
__global__ void __launch_bounds__(256)
scale_by_scalar_bf16_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, long long nvec,
                            const float* __restrict__ scale, int accumulate) {
  const float s = *scale;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const uint4 v = ld_nc_v4(reinterpret_cast<const uint4*>(in) + i);
    float2 b0 = bf2_to_f2(v.x), b1 = bf2_to_f2(v.y), b2 = bf2_to_f2(v.z), b3 = bf2_to_f2(v.w);
    b0.x *= s; b0.y *= s; b1.x *= s; b1.y *= s; b2.x *= s; b2.y *= s; b3.x *= s; b3.y *= s;
    if (accumulate) {
      const uint4 u = ld_v4(reinterpret_cast<const uint4*>(out) + i);
      const float2 a0 = bf2_to_f2(u.x), a1 = bf2_to_f2(u.y), a2 = bf2_to_f2(u.z), a3 = bf2_to_f2(u.w);
      b0.x += a0.x; b0.y += a0.y; b1.x += a1.x; b1.y += a1.y; b2.x += a2.x; b2.y += a2.y; b3.x += a3.x; b3.y += a3.y;
    }
    uint4 o;
    o.x = f2_to_bf2(b0.x, b0.y); o.y = f2_to_bf2(b1.x, b1.y); o.z = f2_to_bf2(b2.x, b2.y); o.w = f2_to_bf2(b3.x, b3.y);
    st_v4(reinterpret_cast<uint4*>(out) + i, o);
  }
}

int num_sms();

}  // namespace b200

using namespace b200;

B200_API int b200_rmsnorm_fwd(const void* x, const void* res, const void* w, void* y, void* sum_out,
                              void* rstd, long long T, int H, float eps, void* stream) {
  if (H % 8) return B200_ERR_BAD_ARG;
  if (res && !sum_out) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  rmsnorm_fwd_kernel<<<(unsigned)T, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)res, (const __nv_bfloat16*)w, (__nv_bfloat16*)y,
      (__nv_bfloat16*)sum_out, (float*)rstd, H, eps);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// number of fp32 partial rows the backward needs as scratch: [b200_rmsnorm_bwd_parts(T), H]
B200_API int b200_rmsnorm_bwd_parts(long long T) {
  long long g = (long long)num_sms() * 4;
  return (int)(T < g ? T : g);
}

B200_API int b200_rmsnorm_bwd(const void* dy, const void* x, const void* w, const void* rstd,
                              const void* dres, void* dx, void* dw, void* dw_partial_scratch,
                              long long T, int H, int accumulate_dw, void* stream) {
  if (H % 8 || H > 8 * 256 * 8) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  const int parts = b200_rmsnorm_bwd_parts(T);
  const int nvec = H / 8;
  const int vpt = (nvec + 255) / 256;
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH_RB(V)                                                                                   \
  rmsnorm_bwd_kernel<V><<<parts, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,       \
                                               (const __nv_bfloat16*)w, (const float*)rstd,             \
                                               (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx,          \
                                               (float*)dw_partial_scratch, (int)T, H)
  if (vpt <= 1) LAUNCH_RB(1);
  else if (vpt <= 2) LAUNCH_RB(2);
  else if (vpt <= 4) LAUNCH_RB(4);
  else LAUNCH_RB(8);
#undef LAUNCH_RB
  b200_count_launch();
  B200_CHECK_LAUNCH();
  rmsnorm_dw_reduce_kernel<<<ceil_div(H, 256), 256, 0, st>>>((const float*)dw_partial_scratch,
                                                             (__nv_bfloat16*)dw, parts, H, accumulate_dw);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_rope_inplace(void* buf, const void* cos_tab, const void* sin_tab, const void* pos,
                               long long T, long long row_stride, int n_heads, int D, int backward,
                               void* stream) {
  if (D % 4) return B200_ERR_BAD_ARG;
  const long long total = T * n_heads * (D / 4);
  if (total == 0) return 0;
  rope_kernel<<<ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)buf, (const float*)cos_tab, (const float*)sin_tab, (const int*)pos, row_stride,
      n_heads, D, backward ? -1.f : 1.f, total);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

static inline int ew_grid(long long work_items) {
  long long g = (work_items + 255) / 256;
  long long cap = (long long)num_sms() * 16;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

B200_API int b200_swiglu_fwd(const void* gu, void* act, long long T, int I, void* stream) {
  if (I % 8) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  swiglu_fwd_kernel<<<ew_grid(T * (I / 8)), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)gu,
                                                                           (__nv_bfloat16*)act, T, I);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_swiglu_bwd(const void* gu, const void* dact, void* dgu, long long T, int I, void* stream) {
  if (I % 8) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  swiglu_bwd_kernel<<<ew_grid(T * (I / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)gu, (const __nv_bfloat16*)dact, (__nv_bfloat16*)dgu, T, I);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_embedding_fwd(const void* ids, const void* table, void* out, long long T, int H,
                                long long vocab, void* stream) {
  if (H % 8) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  embedding_fwd_kernel<<<ew_grid(T * (H / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)ids, (const __nv_bfloat16*)table, (__nv_bfloat16*)out, T, H, vocab);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_embedding_bwd(const void* ids, const void* dout, void* dtable, long long T, int H,
                                long long vocab, void* stream) {
  if (H % 2) return B200_ERR_BAD_ARG;
  if (T == 0) return 0;
  embedding_bwd_kernel<<<ew_grid(T * (H / 2)), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)ids, (const __nv_bfloat16*)dout, (__nv_bfloat16*)dtable, T, H, vocab);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// kind: 0 = fp32->bf16, 1 = bf16->fp32, 2 = fp32->fp32, 3 = bf16->bf16
B200_API int b200_scale_cast(const void* in, void* out, long long n, float scale, int kind, void* stream) {
  if (n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int g = ew_grid(n);
  switch (kind) {
    case 0: scale_cast_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)in, (__nv_bfloat16*)out, n, scale); break;
    case 1: scale_cast_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)in, (float*)out, n, scale); break;
    case 2: scale_cast_kernel<float, float><<<g, 256, 0, st>>>((const float*)in, (float*)out, n, scale); break;
    case 3: scale_cast_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, scale); break;
    default: return B200_ERR_BAD_ARG;
  }
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_scale_by_scalar_bf16(const void* in, void* out, long long n, const void* scale, int accumulate, void* stream) {
  if (n % 8 || scale == nullptr || (reinterpret_cast<uintptr_t>(in) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) return B200_ERR_BAD_ARG;
  if (n == 0) return 0;
  scale_by_scalar_bf16_kernel<<<ew_grid(n / 8), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, n / 8,
                                                                               (const float*)scale, accumulate);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_add_inplace_bf16(void* y, const void* a, long long n, void* stream) {
  if (n % 8) return B200_ERR_BAD_ARG;
  if (n == 0) return 0;
  add_inplace_bf16_kernel<<<ew_grid(n / 8), 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)y,
                                                                           (const __nv_bfloat16*)a, n / 8);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// RMSNorm backward for callers that must not allocate.
//
// b200_rmsnorm_bwd() needs a [parts, H] fp32 scratch for the per-CTA partial sums of dW, and its
// callers normally get that from an allocator.  Some callers cannot: a host whose device heap is
// entirely owned by someone else's allocator (a fixed memory budget handed out at start-up, or a
// framework caching allocator that must not be re-entered from the place this is called from, e.g.
// inside one of its own callbacks), where cudaMalloc / cudaFree / a tensor allocation would fail,
// deadlock, or break the budget.  For them the scratch lives in the library image itself: a
// file-scope __device__ array of fixed size.  Every GPU that loads libb200.so holds those
// NOALLOC_MAX_PARTS * NOALLOC_MAX_H * 4 bytes (8 MiB) until the process exits.
//
// Limits: H <= 8192 (larger H returns B200_ERR_TOO_LARGE); at most 256 partial rows are used, so
// very long inputs run with fewer CTAs than b200_rmsnorm_bwd would pick.
// Set-up: the first call on each device resolves the device address of the array with
// cudaGetSymbolAddress (the array's name is only a host-side handle, never a pointer a kernel can
// use) and caches it; make that first call before any stream capture begins on that device.
// Sharing: all calls use the same array and it keeps whatever the previous call left in it.  Nothing
// has to start at zero: each launch overwrites rows [0, parts) x [0, H) before its second kernel reads
// them, and reads nothing else.  Two calls whose kernels overlap on the GPU (different streams with
// no ordering between them) race on the array: serialise them.
// No cudaMalloc, cudaFree or allocating library routine is called here.
static constexpr int NOALLOC_MAX_PARTS = 256;
static constexpr int NOALLOC_MAX_H = 8192;
static __device__ __align__(256) float g_rmsnorm_bwd_noalloc_scratch[(size_t)NOALLOC_MAX_PARTS * NOALLOC_MAX_H];

static int rmsnorm_noalloc_scratch(float** out) {
  // written without a lock: two host threads racing on the first call of a device both store the same address
  static float* cached[64] = {nullptr};
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return B200_ERR_UNSUPPORTED;
  if (cached[dev] == nullptr) {
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_rmsnorm_bwd_noalloc_scratch));
    cached[dev] = static_cast<float*>(p);
  }
  *out = cached[dev];
  return 0;
}

B200_API int b200_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const void* rstd, const void* dres,
                                      void* dx, void* dw, long long T, int H, int accumulate_dw, void* stream) {
  if (H % 8) return B200_ERR_BAD_ARG;
  if (H > NOALLOC_MAX_H) return B200_ERR_TOO_LARGE;
  if (T == 0) return 0;
  float* scratch = nullptr;
  int rc = rmsnorm_noalloc_scratch(&scratch);
  if (rc) return rc;
  long long want = (long long)num_sms() * 4;
  if (want > NOALLOC_MAX_PARTS) want = NOALLOC_MAX_PARTS;
  const int parts = (int)(T < want ? T : want);
  const int nvec = H / 8;
  const int vpt = (nvec + 255) / 256;
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH_RB_NA(V)                                                                                \
  rmsnorm_bwd_kernel<V><<<parts, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,       \
                                               (const __nv_bfloat16*)w, (const float*)rstd,             \
                                               (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, \
                                               (int)T, H)
  if (vpt <= 1) LAUNCH_RB_NA(1);
  else if (vpt <= 2) LAUNCH_RB_NA(2);
  else LAUNCH_RB_NA(4);
#undef LAUNCH_RB_NA
  b200_count_launch();
  B200_CHECK_LAUNCH();
  rmsnorm_dw_reduce_kernel<<<ceil_div(H, 256), 256, 0, st>>>(scratch, (__nv_bfloat16*)dw, parts, H, accumulate_dw);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
