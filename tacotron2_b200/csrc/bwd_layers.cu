// Small training kernels: loss (model/tts_model.py:197-201), output-mask backward (model/tacotron2.py:333-345),
// column sums, conditioning / embedding backward, BiLSTM train forward + backward (model/encoder.py:59-65).
#include "train.cuh"

namespace {

inline int gs(long long n) { long long g = (n + 255) / 256; return (int)(g < 1 ? 1 : (g > 148 * 8 ? 148 * 8 : g)); }

__device__ __forceinline__ double block_sum_d(double v) {
  __shared__ double red[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double r = 0.0;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) r += red[i];
  return r;
}

// loss = BCEWithLogits(gates, gate_target) + MSE(mels, target) + MSE(mels_post, target), plain means over the padded
// tensors (model/tts_model.py:197-201); also writes the gradients w.r.t. the three (masked) model outputs.
__global__ void __launch_bounds__(256) loss_partial_kernel(const float* __restrict__ mels, const float* __restrict__ post,
                                                           const float* __restrict__ gates, const float* __restrict__ tgt,
                                                           const float* __restrict__ gtgt, long long n_mel, long long n_gate,
                                                           float* __restrict__ d_mels, float* __restrict__ d_post, float* __restrict__ d_gates,
                                                           double* __restrict__ partial) {
  double acc = 0.0;
  const float im = 1.0f / (float)n_mel, ig = 1.0f / (float)n_gate;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_mel; i += (long long)gridDim.x * blockDim.x) {
    const float y = tgt[i], a = mels[i] - y, b = post[i] - y;
    acc += ((double)a * a + (double)b * b) * (double)im;
    if (d_mels) d_mels[i] = 2.f * a * im;
    if (d_post) d_post[i] = 2.f * b * im;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_gate; i += (long long)gridDim.x * blockDim.x) {
    const float x = gates[i], y = gtgt[i];
    const float l = fmaxf(x, 0.f) - x * y + log1pf(expf(-fabsf(x)));
    acc += (double)l * (double)ig;
    if (d_gates) d_gates[i] = (1.f / (1.f + expf(-x)) - y) * ig;
  }
  const double s = block_sum_d(acc);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}
__global__ void loss_final_kernel(const double* partial, int n, float* loss) {
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += partial[i];
  *loss = (float)s;
}

// backward of the residual + output masking (model/tacotron2.py:333-345)
__global__ void finalize_bwd_kernel(const float* __restrict__ d_mels, const float* __restrict__ d_post, const float* __restrict__ d_gates,
                                    const long long* __restrict__ lengths, int B, int T, int M, float* __restrict__ d_mels_raw,
                                    float* __restrict__ d_postout, float* __restrict__ d_gates_raw) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)B * T * M; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % M);
    const long long p = i / M;
    const int b = (int)(p / T), t = (int)(p % T);
    const bool valid = t < lengths[b];
    const float dp = (valid && d_post) ? d_post[i] : 0.f;
    d_postout[i] = dp;
    d_mels_raw[i] = ((valid && d_mels) ? d_mels[i] : 0.f) + dp;
    if (c == 0) d_gates_raw[p] = (valid && d_gates) ? d_gates[p] : 0.f;
  }
}

__global__ void __launch_bounds__(256) colsum_partial_kernel(const float* __restrict__ x, long long rows, int C, long long ld,
                                                             double* __restrict__ partial) {
  __shared__ double red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  double s = 0.0;
  if (c < C)
    for (long long r = blockIdx.y * 8 + threadIdx.y; r < rows; r += (long long)gridDim.y * 8) s += (double)x[r * ld + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    for (int i = 1; i < 8; ++i) s += red[i][threadIdx.x];
    partial[(size_t)blockIdx.y * C + c] = s;
  }
}
__global__ void colsum_final_kernel(const double* partial, int nsplit, int C, float* out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0.0;
  for (int i = 0; i < nsplit; ++i) s += partial[(size_t)i * C + c];
  out[c] = (float)s;
}

}  // namespace

int colsum_launch(const float* x, long long rows, int C, long long ld, float* out, void* scratch, cudaStream_t st) {
  dim3 grid((C + 31) / 32, 64), block(32, 8);
  colsum_partial_kernel<<<grid, block, 0, st>>>(x, rows, C, ld, reinterpret_cast<double*>(scratch));
  colsum_final_kernel<<<(C + 127) / 128, 128, 0, st>>>(reinterpret_cast<double*>(scratch), 64, C, out);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int loss_launch(const float* mels, const float* post, const float* gates, const float* tgt, const float* gtgt, long long n_mel,
                long long n_gate, float* loss, float* d_mels, float* d_post, float* d_gates, void* scratch, cudaStream_t st) {
  const int nb = 148 * 2;
  loss_partial_kernel<<<nb, 256, 0, st>>>(mels, post, gates, tgt, gtgt, n_mel, n_gate, d_mels, d_post, d_gates,
                                          reinterpret_cast<double*>(scratch));
  loss_final_kernel<<<1, 1, 0, st>>>(reinterpret_cast<double*>(scratch), nb, loss);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int finalize_bwd_launch(const float* d_mels, const float* d_post, const float* d_gates, const long long* lengths, int B, int T, int M,
                        float* d_mels_raw, float* d_postout, float* d_gates_raw, cudaStream_t st) {
  finalize_bwd_kernel<<<gs((long long)B * T * M), 256, 0, st>>>(d_mels, d_post, d_gates, lengths, B, T, M, d_mels_raw, d_postout, d_gates_raw);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
