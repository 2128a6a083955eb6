// Training-step shell: global-norm gradient clipping + Adam with L2 weight decay in three launches over a
// list of parameter tensors -- torch.nn.utils.clip_grad_norm_(params, max_norm) followed by torch.optim.Adam(lr, betas, eps,
// weight_decay).step() as the reference configures them (model/tts_model.py:78-91: Adam(lr 1e-3, weight_decay 1e-6);
// run/train.py:239-240: gradient_clip_val 1.0).  No host synchronisation: the clip coefficient stays on the device.
//
//   1  sumsq_kernel   one CTA per 2048-element chunk: sum of g^2 (fp32 per thread, double across the CTA) -> partial[chunk]
//   2  norm_kernel    one CTA adds the partials in a fixed order (deterministic) -> norm, clip = min(1, max_norm / (norm + 1e-6))
//   3  adam_kernel    per chunk: g' = clip*g + wd*p;  m = b1 m + (1-b1) g';  v = b2 v + (1-b2) g'^2;
//                     p -= (lr / (1-b1^t)) * m / (sqrt(v) / sqrt(1-b2^t) + eps)          (torch/optim/adam.py, single-tensor path)
//
// HBM-bound: 5 reads + 3 writes of 4 bytes per parameter (g twice).
#include "../../include/t2b200.h"
#include "common.cuh"

#include <cstdint>
#include <cstring>
#include <vector>

#define OP_CHUNK 2048
#define OP_THREADS 256

namespace {

struct OpTensor { float* p; float* g; float* m; float* v; long long numel; };
struct OpChunk { int tensor; int start_div;  /* start element / OP_CHUNK */ };

__global__ void __launch_bounds__(OP_THREADS) sumsq_kernel(const OpTensor* __restrict__ tens, const OpChunk* __restrict__ chunks,
                                                           double* __restrict__ partial) {
  const OpChunk c = chunks[blockIdx.x];
  const OpTensor t = tens[c.tensor];
  const long long base = (long long)c.start_div * OP_CHUNK;
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < OP_CHUNK / OP_THREADS; ++i) {
    const long long e = base + (long long)i * OP_THREADS + threadIdx.x;
    if (e < t.numel) { const float g = t.g[e]; s = fmaf(g, g, s); }
  }
  __shared__ double red[OP_THREADS / 32];
  double d = (double)s;
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = d;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int i = 0; i < OP_THREADS / 32; ++i) tot += red[i];
    partial[blockIdx.x] = tot;
  }
}

__global__ void __launch_bounds__(OP_THREADS) norm_kernel(const double* __restrict__ partial, int n, float max_norm, float* __restrict__ scal,
                                                          float* __restrict__ norm_out) {
  __shared__ double red[OP_THREADS];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += OP_THREADS) s += partial[i];   // fixed assignment and order: deterministic
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = OP_THREADS / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float norm = (float)sqrt(red[0]);
    const float coef = max_norm > 0.f ? fminf(1.0f, max_norm / (norm + 1e-6f)) : 1.0f;   // torch clip_grad_norm_: clamp(max_norm / (norm + 1e-6), max 1)
    scal[0] = norm;
    scal[1] = coef;
    if (norm_out) *norm_out = norm;
  }
}

__global__ void __launch_bounds__(OP_THREADS) adam_kernel(const OpTensor* __restrict__ tens, const OpChunk* __restrict__ chunks,
                                                          const float* __restrict__ scal, float step_size, float beta1, float beta2,
                                                          float inv_sqrt_bc2, float eps, float wd) {
  const OpChunk c = chunks[blockIdx.x];
  const OpTensor t = tens[c.tensor];
  const long long base = (long long)c.start_div * OP_CHUNK;
  const float clip = scal[1];
  // a non-finite gradient norm (overflow, or a NaN-poisoned gradient after an aborted kernel) skips the update, as the GradScaler of
  // the reference's 16-mixed Trainer does (run/train.py:239); the caller sees it in grad_norm_out
  if (!isfinite(scal[0])) return;
#pragma unroll
  for (int i = 0; i < OP_CHUNK / OP_THREADS; ++i) {
    const long long e = base + (long long)i * OP_THREADS + threadIdx.x;
    if (e < t.numel) {
      const float p = t.p[e];
      const float g = fmaf(wd, p, t.g[e] * clip);
      const float m = fmaf(beta1, t.m[e], (1.0f - beta1) * g);
      const float v = fmaf(beta2, t.v[e], (1.0f - beta2) * g * g);
      t.m[e] = m;
      t.v[e] = v;
      t.p[e] = p - step_size * (m / (sqrtf(v) * inv_sqrt_bc2 + eps));
    }
  }
}

inline long long align256(long long x) { return (x + 255) & ~255ll; }

}  // namespace

// Workspace layout: [header 256 B: signature of the cached chunk table][tensor table][chunk table][partials (double)][scalars]
extern "C" long long t2_clip_adam_workspace_bytes(int n_tensors, long long total_chunks) {
  if (n_tensors <= 0 || total_chunks <= 0) return -1;
  return 256 + align256((long long)n_tensors * sizeof(OpTensor)) + align256(total_chunks * (long long)sizeof(OpChunk)) +
         align256(total_chunks * (long long)sizeof(double)) + 256;
}

extern "C" long long t2_clip_adam_chunks(int n_tensors, const long long* numel) {
  if (n_tensors <= 0 || !numel) return -1;
  long long n = 0;
  for (int i = 0; i < n_tensors; ++i) n += (numel[i] + OP_CHUNK - 1) / OP_CHUNK;
  return n;
}

extern "C" int t2_clip_adam_step(int n_tensors, float* const* params, float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                                 const long long* numel, float lr, float beta1, float beta2, float eps, float weight_decay, float max_norm,
                                 long long step, int table_cached, float* grad_norm_out, void* ws, long long ws_bytes, void* stream) {
  T2_REQUIRE(n_tensors > 0 && params && grads && exp_avg && exp_avg_sq && numel && ws && step >= 1, T2_ERR_ARG,
             "t2_clip_adam_step: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const long long nchunks = t2_clip_adam_chunks(n_tensors, numel);
  T2_REQUIRE(nchunks > 0 && nchunks < (1ll << 31), T2_ERR_ARG, "t2_clip_adam_step: empty or oversized tensor list");
  T2_REQUIRE(ws_bytes >= t2_clip_adam_workspace_bytes(n_tensors, nchunks), T2_ERR_WORKSPACE, "t2_clip_adam_step: workspace too small");
  char* base = (char*)ws + 256;
  OpTensor* d_t = (OpTensor*)base;                      base += align256((long long)n_tensors * sizeof(OpTensor));
  OpChunk* d_c = (OpChunk*)base;                        base += align256(nchunks * (long long)sizeof(OpChunk));
  double* d_part = (double*)base;                       base += align256(nchunks * (long long)sizeof(double));
  float* d_scal = (float*)base;
  // pointer table: every step (gradient buffers move); chunk table: only when the caller says the tensor list changed
  std::vector<OpTensor> ht((size_t)n_tensors);
  for (int i = 0; i < n_tensors; ++i) {
    T2_REQUIRE(params[i] && grads[i] && exp_avg[i] && exp_avg_sq[i] && numel[i] > 0, T2_ERR_ARG, "t2_clip_adam_step: null tensor %d", i);
    ht[(size_t)i] = OpTensor{params[i], grads[i], exp_avg[i], exp_avg_sq[i], numel[i]};
  }
  T2_CHECK_CUDA(cudaMemcpyAsync(d_t, ht.data(), ht.size() * sizeof(OpTensor), cudaMemcpyHostToDevice, st));
  if (!table_cached) {
    std::vector<OpChunk> hc((size_t)nchunks);
    size_t k = 0;
    for (int i = 0; i < n_tensors; ++i)
      for (long long s = 0; s < numel[i]; s += OP_CHUNK) hc[k++] = OpChunk{i, (int)(s / OP_CHUNK)};
    T2_CHECK_CUDA(cudaMemcpyAsync(d_c, hc.data(), hc.size() * sizeof(OpChunk), cudaMemcpyHostToDevice, st));
    T2_CHECK_CUDA(cudaStreamSynchronize(st));   // hc / ht are host temporaries; only on the (rare) table upload
  }
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  sumsq_kernel<<<(unsigned)nchunks, OP_THREADS, 0, st>>>(d_t, d_c, d_part);
  norm_kernel<<<1, OP_THREADS, 0, st>>>(d_part, (int)nchunks, max_norm, d_scal, grad_norm_out);
  adam_kernel<<<(unsigned)nchunks, OP_THREADS, 0, st>>>(d_t, d_c, d_scal, (float)(lr / bc1), beta1, beta2, (float)(1.0 / sqrt(bc2)), eps,
                                                         weight_decay);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
