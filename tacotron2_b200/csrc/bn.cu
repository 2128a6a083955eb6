// Train-mode BatchNorm1d + activation + dropout around the tensor-core convolutions, forward and backward
// (model/encoder.py:31-46, model/postnet.py:8-47 under model.train(): biased batch statistics over ALL (batch, time)
// positions including padding -- SURVEY appendix 13 -- running statistics updated with momentum 0.1 and the unbiased
// variance, exactly nn.BatchNorm1d).  Channels-last fp32 [rows][C]; per-channel sums are accumulated in double in a fixed
// order (deterministic).
#include "train.cuh"

#define BN_NS 64  // row splits of the per-channel reductions

namespace {

__device__ __forceinline__ float act_fwd(float a, int act) {
  return act == ACT_RELU ? fmaxf(a, 0.f) : (act == ACT_TANH ? tanhf(a) : a);
}

// partial[split][c] = (sum z, sum z^2) over this split's rows
__global__ void __launch_bounds__(256) bn_stats_partial_kernel(const float* __restrict__ z, long long rows, int C, double2* __restrict__ partial) {
  __shared__ double2 red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  double s = 0.0, q = 0.0;
  if (c < C)
    for (long long r = blockIdx.y * 8 + threadIdx.y; r < rows; r += (long long)gridDim.y * 8) {
      const double v = (double)z[r * C + c];
      s += v; q += v * v;
    }
  red[threadIdx.y][threadIdx.x] = make_double2(s, q);
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    for (int i = 1; i < 8; ++i) { s += red[i][threadIdx.x].x; q += red[i][threadIdx.x].y; }
    partial[(size_t)blockIdx.y * C + c] = make_double2(s, q);
  }
}

__global__ void bn_stats_final_kernel(const double2* __restrict__ partial, int nsplit, long long rows, int C, float* mean, float* rstd,
                                      float* running_mean, float* running_var, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int i = 0; i < nsplit; ++i) { s += partial[(size_t)i * C + c].x; q += partial[(size_t)i * C + c].y; }
  const double m = s / (double)rows;
  double var = q / (double)rows - m * m;
  if (var < 0.0) var = 0.0;
  mean[c] = (float)m;
  rstd[c] = (float)(1.0 / sqrt(var + 1e-5));
  if (running_mean) {  // nn.BatchNorm1d: running = (1 - momentum) * running + momentum * batch (unbiased variance)
    const double unb = rows > 1 ? var * (double)rows / (double)(rows - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unb;
  }
}

__global__ void __launch_bounds__(256) bn_apply_kernel(BnLayer L, long long rows, float* __restrict__ out_f32,
                                                       __nv_bfloat16* __restrict__ out_b16) {
  const int C = L.C;
  const long long n = rows * C;
  const float dscale = 1.0f / (1.0f - L.p);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long r = i / C;
    const float a = (L.z[i] - L.mean[c]) * L.rstd[c] * L.gamma[c] + L.beta[c];
    float y = act_fwd(a, L.act);
    if (L.dropout) {
      const bool k = L.keep ? (L.keep[r * L.keep_ld + c] != 0) : rng_keep(L.seed, L.rng_base + (unsigned long long)(r * L.keep_ld + c), L.p);
      y = k ? y * dscale : 0.f;
    }
    if (out_f32) out_f32[i] = y;
    if (out_b16) {
      __nv_bfloat16 hi, lo;
      tc::split_bf16(y, hi, lo);
      out_b16[i] = hi;
      out_b16[n + i] = lo;
    }
  }
}

// d_a = dy * keep/(1-p) * act'(a); partial sums (sum d_a, sum d_a * zhat) per channel
__device__ __forceinline__ float bn_da(const BnLayer& L, const float* __restrict__ dy, long long r, int c, long long i, float& zhat) {
  zhat = (L.z[i] - L.mean[c]) * L.rstd[c];
  const float a = zhat * L.gamma[c] + L.beta[c];
  float d = dy[i];
  if (L.dropout) {
    const bool k = L.keep ? (L.keep[r * L.keep_ld + c] != 0) : rng_keep(L.seed, L.rng_base + (unsigned long long)(r * L.keep_ld + c), L.p);
    d = k ? d * (1.0f / (1.0f - L.p)) : 0.f;
  }
  if (L.act == ACT_RELU) d = a > 0.f ? d : 0.f;
  else if (L.act == ACT_TANH) { const float t = tanhf(a); d *= 1.f - t * t; }
  return d;
}

__global__ void __launch_bounds__(256) bn_bwd_partial_kernel(BnLayer L, const float* __restrict__ dy, long long rows, double2* __restrict__ partial) {
  __shared__ double2 red[8][33];
  const int C = L.C;
  const int c = blockIdx.x * 32 + threadIdx.x;
  double s = 0.0, q = 0.0;
  if (c < C)
    for (long long r = blockIdx.y * 8 + threadIdx.y; r < rows; r += (long long)gridDim.y * 8) {
      float zh;
      const float d = bn_da(L, dy, r, c, r * C + c, zh);
      s += (double)d; q += (double)d * (double)zh;
    }
  red[threadIdx.y][threadIdx.x] = make_double2(s, q);
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    for (int i = 1; i < 8; ++i) { s += red[i][threadIdx.x].x; q += red[i][threadIdx.x].y; }
    partial[(size_t)blockIdx.y * C + c] = make_double2(s, q);
  }
}

__global__ void bn_bwd_final_kernel(const double2* __restrict__ partial, int nsplit, long long rows, int C, float* s1n, float* s2n,
                                    float* dgamma, float* dbeta) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int i = 0; i < nsplit; ++i) { s += partial[(size_t)i * C + c].x; q += partial[(size_t)i * C + c].y; }
  s1n[c] = (float)(s / (double)rows);
  s2n[c] = (float)(q / (double)rows);
  if (dbeta) dbeta[c] = (float)s;
  if (dgamma) dgamma[c] = (float)q;
}

// dz = gamma * rstd * (d_a - mean(d_a) - zhat * mean(d_a * zhat))  -> bf16 (hi,lo) channels-last [2][rows][C]
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(BnLayer L, const float* __restrict__ dy, long long rows, const float* __restrict__ s1n,
                                                           const float* __restrict__ s2n, __nv_bfloat16* __restrict__ dz_b16,
                                                           float* __restrict__ dz_f32) {
  const int C = L.C;
  const long long n = rows * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long r = i / C;
    float zh;
    const float d = bn_da(L, dy, r, c, i, zh);
    const float dz = L.gamma[c] * L.rstd[c] * (d - s1n[c] - zh * s2n[c]);
    if (dz_f32) dz_f32[i] = dz;
    if (dz_b16) {
      __nv_bfloat16 hi, lo;
      tc::split_bf16(dz, hi, lo);
      dz_b16[i] = hi;
      dz_b16[n + i] = lo;
    }
  }
}

// ---- four channels per thread (C % 4 == 0, 16-byte aligned tensors): 16-byte loads and stores, one row / channel split per four
// elements instead of a 64-bit division per element.  Same arithmetic and the same summation order as the scalar kernels above.
struct Bn4 { float4 mean, rstd, gamma, beta; };
__device__ __forceinline__ Bn4 bn4_params(const BnLayer& L, int c) {
  Bn4 p;
  p.mean = __ldg(reinterpret_cast<const float4*>(L.mean + c)); p.rstd = __ldg(reinterpret_cast<const float4*>(L.rstd + c));
  p.gamma = __ldg(reinterpret_cast<const float4*>(L.gamma + c)); p.beta = __ldg(reinterpret_cast<const float4*>(L.beta + c));
  return p;
}
__device__ __forceinline__ void bn4_keep(const BnLayer& L, long long r, int c, bool (&k)[4]) {
  if (L.keep) {
    const uchar4 m = *reinterpret_cast<const uchar4*>(L.keep + r * L.keep_ld + c);
    k[0] = m.x != 0; k[1] = m.y != 0; k[2] = m.z != 0; k[3] = m.w != 0;
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) k[e] = rng_keep(L.seed, L.rng_base + (unsigned long long)(r * L.keep_ld + c + e), L.p);
  }
}
__device__ __forceinline__ void st_split4(__nv_bfloat16* hi_p, __nv_bfloat16* lo_p, const float (&v)[4]) {
  __align__(8) __nv_bfloat16 hi[4], lo[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) tc::split_bf16(v[e], hi[e], lo[e]);
  *reinterpret_cast<uint2*>(hi_p) = *reinterpret_cast<const uint2*>(hi);
  *reinterpret_cast<uint2*>(lo_p) = *reinterpret_cast<const uint2*>(lo);
}

__global__ void __launch_bounds__(256) bn_apply4_kernel(BnLayer L, long long rows, float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_b16) {
  const int C4 = L.C >> 2;
  const long long n4 = rows * C4, n = rows * L.C;
  const float dscale = 1.0f / (1.0f - L.p);
  for (long long i4 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i4 < n4; i4 += (long long)gridDim.x * blockDim.x) {
    const long long r = i4 / C4;
    const int c = (int)(i4 - r * C4) * 4;
    const Bn4 p = bn4_params(L, c);
    const float4 z = *reinterpret_cast<const float4*>(L.z + i4 * 4);
    float y[4] = {act_fwd((z.x - p.mean.x) * p.rstd.x * p.gamma.x + p.beta.x, L.act), act_fwd((z.y - p.mean.y) * p.rstd.y * p.gamma.y + p.beta.y, L.act),
                  act_fwd((z.z - p.mean.z) * p.rstd.z * p.gamma.z + p.beta.z, L.act), act_fwd((z.w - p.mean.w) * p.rstd.w * p.gamma.w + p.beta.w, L.act)};
    if (L.dropout) {
      bool k[4];
      bn4_keep(L, r, c, k);
#pragma unroll
      for (int e = 0; e < 4; ++e) y[e] = k[e] ? y[e] * dscale : 0.f;
    }
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + i4 * 4) = make_float4(y[0], y[1], y[2], y[3]);
    if (out_b16) st_split4(out_b16 + i4 * 4, out_b16 + n + i4 * 4, y);
  }
}

// d_a of four channels of one row (bn_da), zhat alongside
__device__ __forceinline__ void bn_da4(const BnLayer& L, const float* __restrict__ dy, long long r, int c, float (&d)[4], float (&zh)[4]) {
  const Bn4 p = bn4_params(L, c);
  const long long i = r * L.C + c;
  const float4 z = *reinterpret_cast<const float4*>(L.z + i), g = *reinterpret_cast<const float4*>(dy + i);
  zh[0] = (z.x - p.mean.x) * p.rstd.x; zh[1] = (z.y - p.mean.y) * p.rstd.y; zh[2] = (z.z - p.mean.z) * p.rstd.z; zh[3] = (z.w - p.mean.w) * p.rstd.w;
  const float a[4] = {zh[0] * p.gamma.x + p.beta.x, zh[1] * p.gamma.y + p.beta.y, zh[2] * p.gamma.z + p.beta.z, zh[3] * p.gamma.w + p.beta.w};
  d[0] = g.x; d[1] = g.y; d[2] = g.z; d[3] = g.w;
  if (L.dropout) {
    bool k[4];
    bn4_keep(L, r, c, k);
#pragma unroll
    for (int e = 0; e < 4; ++e) d[e] = k[e] ? d[e] * (1.0f / (1.0f - L.p)) : 0.f;
  }
  if (L.act == ACT_RELU) {
#pragma unroll
    for (int e = 0; e < 4; ++e) d[e] = a[e] > 0.f ? d[e] : 0.f;
  } else if (L.act == ACT_TANH) {
#pragma unroll
    for (int e = 0; e < 4; ++e) { const float t = tanhf(a[e]); d[e] *= 1.f - t * t; }
  }
}

__global__ void __launch_bounds__(256) bn_bwd_partial4_kernel(BnLayer L, const float* __restrict__ dy, long long rows, double2* __restrict__ partial) {
  __shared__ double2 red[8][4][33];
  const int C = L.C;
  const int c = blockIdx.x * 128 + threadIdx.x * 4;
  double s[4] = {0.0, 0.0, 0.0, 0.0}, q[4] = {0.0, 0.0, 0.0, 0.0};
  if (c < C)
    for (long long r = blockIdx.y * 8 + threadIdx.y; r < rows; r += (long long)gridDim.y * 8) {
      float d[4], zh[4];
      bn_da4(L, dy, r, c, d, zh);
#pragma unroll
      for (int e = 0; e < 4; ++e) { s[e] += (double)d[e]; q[e] += (double)d[e] * (double)zh[e]; }
    }
#pragma unroll
  for (int e = 0; e < 4; ++e) red[threadIdx.y][e][threadIdx.x] = make_double2(s[e], q[e]);
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      for (int i = 1; i < 8; ++i) { s[e] += red[i][e][threadIdx.x].x; q[e] += red[i][e][threadIdx.x].y; }
      partial[(size_t)blockIdx.y * C + c + e] = make_double2(s[e], q[e]);
    }
  }
}

__global__ void __launch_bounds__(256) bn_bwd_apply4_kernel(BnLayer L, const float* __restrict__ dy, long long rows, const float* __restrict__ s1n,
                                                            const float* __restrict__ s2n, __nv_bfloat16* __restrict__ dz_b16,
                                                            float* __restrict__ dz_f32) {
  const int C4 = L.C >> 2;
  const long long n4 = rows * C4, n = rows * L.C;
  for (long long i4 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i4 < n4; i4 += (long long)gridDim.x * blockDim.x) {
    const long long r = i4 / C4;
    const int c = (int)(i4 - r * C4) * 4;
    float d[4], zh[4];
    bn_da4(L, dy, r, c, d, zh);
    const float4 ga = __ldg(reinterpret_cast<const float4*>(L.gamma + c)), rs = __ldg(reinterpret_cast<const float4*>(L.rstd + c));
    const float4 m1 = __ldg(reinterpret_cast<const float4*>(s1n + c)), m2 = __ldg(reinterpret_cast<const float4*>(s2n + c));
    const float dz[4] = {ga.x * rs.x * (d[0] - m1.x - zh[0] * m2.x), ga.y * rs.y * (d[1] - m1.y - zh[1] * m2.y),
                         ga.z * rs.z * (d[2] - m1.z - zh[2] * m2.z), ga.w * rs.w * (d[3] - m1.w - zh[3] * m2.w)};
    if (dz_f32) *reinterpret_cast<float4*>(dz_f32 + i4 * 4) = make_float4(dz[0], dz[1], dz[2], dz[3]);
    if (dz_b16) st_split4(dz_b16 + i4 * 4, dz_b16 + n + i4 * 4, dz);
  }
}

__host__ inline bool bn_vec4(const BnLayer& L, const void* a, const void* b, const void* c2) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  return L.C % 4 == 0 && al(L.z) && al(L.mean) && al(L.rstd) && al(L.gamma) && al(L.beta) && al(a) && al(b) && al(c2) &&
         (!L.keep || ((L.keep_ld & 3) == 0 && (reinterpret_cast<uintptr_t>(L.keep) & 3) == 0)) && (L.keep_ld & 3) == 0 && !getenv("T2B200_BN_SCALAR");
}

inline int gs(long long n) { long long g = (n + 255) / 256; return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g)); }

}  // namespace

long long bn_scratch_bytes(int C) { return (long long)BN_NS * C * sizeof(double2) + 2ll * C * sizeof(float) + 512; }

int bn_stats_launch(const float* z, long long rows, int C, float* mean, float* rstd, float* running_mean, float* running_var,
                    void* scratch, cudaStream_t st) {
  double2* partial = reinterpret_cast<double2*>(scratch);
  dim3 grid((C + 31) / 32, BN_NS), block(32, 8);
  bn_stats_partial_kernel<<<grid, block, 0, st>>>(z, rows, C, partial);
  bn_stats_final_kernel<<<(C + 127) / 128, 128, 0, st>>>(partial, BN_NS, rows, C, mean, rstd, running_mean, running_var, 0.1f);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int bn_apply_launch(const BnLayer& L, long long rows, float* out_f32, __nv_bfloat16* out_b16, cudaStream_t st) {
  if (bn_vec4(L, out_f32, out_b16, nullptr)) bn_apply4_kernel<<<gs(rows * L.C / 4), 256, 0, st>>>(L, rows, out_f32, out_b16);
  else bn_apply_kernel<<<gs(rows * L.C), 256, 0, st>>>(L, rows, out_f32, out_b16);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int bn_bwd_launch(const BnLayer& L, const float* dy, long long rows, float* dgamma, float* dbeta, __nv_bfloat16* dz_b16, float* dz_f32,
                  void* scratch, cudaStream_t st) {
  double2* partial = reinterpret_cast<double2*>(scratch);
  float* s1n = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + (size_t)BN_NS * L.C * sizeof(double2));
  float* s2n = s1n + L.C;
  dim3 grid((L.C + 31) / 32, BN_NS), block(32, 8);
  const bool v4 = bn_vec4(L, dy, dz_b16, dz_f32);
  if (v4) bn_bwd_partial4_kernel<<<dim3((L.C + 127) / 128, BN_NS), block, 0, st>>>(L, dy, rows, partial);
  else bn_bwd_partial_kernel<<<grid, block, 0, st>>>(L, dy, rows, partial);
  bn_bwd_final_kernel<<<(L.C + 127) / 128, 128, 0, st>>>(partial, BN_NS, rows, L.C, s1n, s2n, dgamma, dbeta);
  if (v4) bn_bwd_apply4_kernel<<<gs(rows * L.C / 4), 256, 0, st>>>(L, dy, rows, s1n, s2n, dz_b16, dz_f32);
  else bn_bwd_apply_kernel<<<gs(rows * L.C), 256, 0, st>>>(L, dy, rows, s1n, s2n, dz_b16, dz_f32);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
