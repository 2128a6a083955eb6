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
  bn_apply_kernel<<<gs(rows * L.C), 256, 0, st>>>(L, rows, out_f32, out_b16);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int bn_bwd_launch(const BnLayer& L, const float* dy, long long rows, float* dgamma, float* dbeta, __nv_bfloat16* dz_b16, float* dz_f32,
                  void* scratch, cudaStream_t st) {
  double2* partial = reinterpret_cast<double2*>(scratch);
  float* s1n = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + (size_t)BN_NS * L.C * sizeof(double2));
  float* s2n = s1n + L.C;
  dim3 grid((L.C + 31) / 32, BN_NS), block(32, 8);
  bn_bwd_partial_kernel<<<grid, block, 0, st>>>(L, dy, rows, partial);
  bn_bwd_final_kernel<<<(L.C + 127) / 128, 128, 0, st>>>(partial, BN_NS, rows, L.C, s1n, s2n, dgamma, dbeta);
  bn_bwd_apply_kernel<<<gs(rows * L.C), 256, 0, st>>>(L, dy, rows, s1n, s2n, dz_b16, dz_f32);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
