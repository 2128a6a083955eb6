// fp32 CUDA-core implementations of the once-per-batch layers (exact-parity engine):
//   Encoder  model/encoder.py:54-67   embedding, 3x conv1d k5 + BN + ReLU (+dropout), packed BiLSTM
//   Postnet  model/postnet.py:8-52    5x conv1d k5 + BN (+tanh) (+dropout)
//   glue     model/tacotron2.py:201-212 (conditioning), :229 (att_encoder), :332-345 (residual + masks)
#include "layers.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// Channels-last conv1d as an implicit GEMM: M = B*T positions, N = Cout, K = k*Cin.
// 64x64 output tile, K step 16, 256 threads, 4x4 outputs per thread.
#define CV_BM 64
#define CV_BN 64
#define CV_BK 16

__global__ void __launch_bounds__(256) conv1d_cl_kernel(ConvArgs a) {
  __shared__ float As[2][CV_BK][CV_BM + 4];
  __shared__ float Bs[2][CV_BK][CV_BN];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * CV_BM, n0 = blockIdx.y * CV_BN;
  const int Mtot = a.B * a.T;
  const int pad = (a.k - 1) / 2;
  // loader roles
  const int a_pos = tid >> 2, a_c4 = (tid & 3) * 4;        // A: 64 positions x 16 channels
  const int b_k = tid >> 4, b_n4 = (tid & 15) * 4;         // B: 16 k x 64 couts
  const int m_a = m0 + a_pos;
  const int ab = m_a / a.T, at = m_a % a.T;
  const int kchunks = a.k * (a.Cin / CV_BK);
  const int cin_chunks = a.Cin / CV_BK;
  float4 areg, breg;
  auto load = [&](int kc) {
    const int j = kc / cin_chunks, ci0 = (kc % cin_chunks) * CV_BK;
    const int ts = at + j - pad;
    areg = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m_a < Mtot && ts >= 0 && ts < a.T)
      areg = __ldg(reinterpret_cast<const float4*>(a.in + (size_t)ab * a.in_bstride + (size_t)ts * a.Cin + ci0 + a_c4));
    breg = make_float4(0.f, 0.f, 0.f, 0.f);
    const int n = n0 + b_n4;
    const float* wrow = a.Wr + ((size_t)j * a.Cin + ci0 + b_k) * a.Cout + n;
    if (n + 3 < a.Cout) breg = __ldg(reinterpret_cast<const float4*>(wrow));
    else {
      if (n + 0 < a.Cout) breg.x = __ldg(wrow + 0);
      if (n + 1 < a.Cout) breg.y = __ldg(wrow + 1);
      if (n + 2 < a.Cout) breg.z = __ldg(wrow + 2);
    }
  };
  auto store = [&](int buf) {
    As[buf][a_c4 + 0][a_pos] = areg.x;
    As[buf][a_c4 + 1][a_pos] = areg.y;
    As[buf][a_c4 + 2][a_pos] = areg.z;
    As[buf][a_c4 + 3][a_pos] = areg.w;
    *reinterpret_cast<float4*>(&Bs[buf][b_k][b_n4]) = breg;
  };
  const int ty = tid >> 4, tx = tid & 15;  // thread tile rows ty*4.., cols tx*4..
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  load(0);
  store(0);
  __syncthreads();
  for (int kc = 0; kc < kchunks; ++kc) {
    const int buf = kc & 1;
    if (kc + 1 < kchunks) load(kc + 1);
#pragma unroll
    for (int kk = 0; kk < CV_BK; ++kk) {
      const float4 av = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float am[4] = {av.x, av.y, av.z, av.w};
      const float bm[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(am[i], bm[j], acc[i][j]);
    }
    if (kc + 1 < kchunks) store(buf ^ 1);
    __syncthreads();
  }
  // epilogue: bias / folded BN, activation, dropout
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= Mtot) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= a.Cout) continue;
      float v = acc[i][j];
      if (a.bn_scale) v = fmaf(v, __ldg(a.bn_scale + n), __ldg(a.bn_shift + n));
      else if (a.bias) v += __ldg(a.bias + n);
      if (a.act == ACT_RELU) v = fmaxf(v, 0.f);
      else if (a.act == ACT_TANH) v = tanhf(v);
      if (a.keep) v = a.keep[(size_t)m * a.keep_ld + n] ? v * a.drop_scale : 0.f;
      a.out[(size_t)m * a.Cout + n] = v;
    }
  }
}

__global__ void embed_kernel(const long long* idx, const float* table, float* out, int n_pos, int E, int n_rows) {
  const int e4 = E / 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pos * e4; i += gridDim.x * blockDim.x) {
    const int p = i / e4, c = i % e4;
    long long r = idx[p];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r >= 0 && r < n_rows) v = __ldg(reinterpret_cast<const float4*>(table + (size_t)r * E) + c);
    reinterpret_cast<float4*>(out + (size_t)p * E)[c] = v;
  }
}

// ------------------------------------------------------------------------------------------
// Packed BiLSTM recurrence (model/encoder.py:59-65).  One CTA per (direction, utterance); thread j
// owns hidden unit j.  xproj already holds x W_ih^T + b_ih + b_hh.  The reverse direction starts at
// the utterance's own last character; positions >= len are written as zeros.
__global__ void __launch_bounds__(256) bilstm_kernel(const float* __restrict__ xproj, const float* __restrict__ WhhT,
                                                     const long long* __restrict__ lens, float* __restrict__ out, int B,
                                                     int L, int Hh) {
  extern __shared__ float hs[];  // [Hh]
  const int dir = blockIdx.x & 1, b = blockIdx.x >> 1;
  const int n = (int)min((long long)L, max(0ll, lens[b]));
  const float* W = WhhT + (size_t)dir * Hh * 4 * Hh;  // [k][4Hh]
  for (int j = threadIdx.x; j < Hh; j += blockDim.x) hs[j] = 0.f;
  for (int t = n + (int)threadIdx.x / Hh; t < L; t += blockDim.x / Hh) {  // zero the padded tail
    for (int j = threadIdx.x % Hh; j < Hh; j += Hh) out[((size_t)b * L + t) * 2 * Hh + dir * Hh + j] = 0.f;
  }
  float c = 0.f;
  __syncthreads();
  for (int s = 0; s < n; ++s) {
    const int t = dir ? (n - 1 - s) : s;
    const int j = threadIdx.x;
    float hnew = 0.f;
    if (j < Hh) {
      const float* xp = xproj + (((size_t)b * L + t) * 2 + dir) * 4 * Hh;
      float g0 = xp[j], g1 = xp[Hh + j], g2 = xp[2 * Hh + j], g3 = xp[3 * Hh + j];
#pragma unroll 4
      for (int k = 0; k < Hh; ++k) {
        const float hk = hs[k];
        const float* wr = W + (size_t)k * 4 * Hh;
        g0 = fmaf(__ldg(wr + j), hk, g0);
        g1 = fmaf(__ldg(wr + Hh + j), hk, g1);
        g2 = fmaf(__ldg(wr + 2 * Hh + j), hk, g2);
        g3 = fmaf(__ldg(wr + 3 * Hh + j), hk, g3);
      }
      c = sigmoidf_(g1) * c + sigmoidf_(g0) * tanhf(g2);
      hnew = sigmoidf_(g3) * tanhf(c);
      out[((size_t)b * L + t) * 2 * Hh + dir * Hh + j] = hnew;
    }
    __syncthreads();
    if (j < Hh) hs[j] = hnew;
    __syncthreads();
  }
}

// tanh(enc + speaker_emb) over ALL positions, then concat the description vector (model/tacotron2.py:201-212)
__global__ void condition_kernel(const float* enc, const float* spk_table, const long long* spk_id, const float* desc_vec,
                                 int D, float* out, int B, int L, int E) {
  const int EF = E + D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)B * L * EF;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % EF);
    const long long p = i / EF;
    const int b = (int)(p / L);
    float v;
    if (c < E) {
      v = enc[p * E + c];
      if (spk_table) v = tanhf(v + spk_table[(size_t)spk_id[b] * E + c]);
    } else {
      v = desc_vec[(size_t)b * D + (c - E)];
    }
    out[i] = v;
  }
}

// Tanh(Linear(description)) (model/tacotron2.py:99-103); one warp per output
__global__ void desc_linear_kernel(const float* desc, const float* W, const float* bias, float* out, int B, int Din, int Dout) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (gw >= B * Dout) return;
  const int b = gw / Dout, o = gw % Dout;
  float acc = 0.f;
  for (int k = lane; k < Din; k += 32) acc = fmaf(W[(size_t)o * Din + k], desc[(size_t)b * Din + k], acc);
  acc = warp_sum(acc);
  if (lane == 0) out[gw] = tanhf(acc + bias[o]);
}

// mels_post = mels + postnet(mels); masked_fill by length (model/tacotron2.py:333-345)
__global__ void finalize_kernel(float* mels, float* post, float* gates, const long long* lengths, int B, int T, int Tstride, int M) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)B * T * M;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % M);
    const long long p = i / M;
    const int b = (int)(p / T), t = (int)(p % T);
    const size_t mi = ((size_t)b * Tstride + t) * M + c;
    const bool masked = t >= lengths[b];
    const float m = mels[mi];
    post[i] = masked ? 0.f : m + post[i];
    if (masked) mels[mi] = 0.f;
    if (c == 0 && masked) gates[(size_t)b * Tstride + t] = -1000.0f;
  }
}

__global__ void bn_fold_kernel(const float* w, const float* b, const float* mean, const float* var, const float* conv_bias,
                               float* scale, float* shift, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C) return;
  const float sc = w[i] * rsqrtf(var[i] + 1e-5f);
  scale[i] = sc;
  shift[i] = b[i] + ((conv_bias ? conv_bias[i] : 0.f) - mean[i]) * sc;
}

__global__ void repack_conv_kernel(const float* W, float* Wr, int Cout, int Cin, int k) {
  const long long n = (long long)Cout * Cin * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    const int ci = (int)((i / Cout) % Cin);
    const int j = (int)(i / ((long long)Cout * Cin));
    Wr[i] = W[((size_t)co * Cin + ci) * k + j];
  }
}

__global__ void transpose_kernel(const float* in, float* out, int R, int C) {
  __shared__ float tile[32][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = blockIdx.y * 32 + i;
    if (r < R && c < C) tile[i][threadIdx.x] = in[(size_t)r * C + c];
  }
  __syncthreads();
  const int r2 = blockIdx.y * 32 + threadIdx.x;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c2 = blockIdx.x * 32 + i;
    if (r2 < R && c2 < C) out[(size_t)c2 * R + r2] = tile[threadIdx.x][i];
  }
}

inline int gs(long long n, int bs) { long long g = (n + bs - 1) / bs; return (int)(g < 148 * 16 ? (g < 1 ? 1 : g) : 148 * 16); }

}  // namespace

int conv1d_cl_launch(const ConvArgs& a, cudaStream_t st) {
  T2_REQUIRE(a.Cin % CV_BK == 0 && a.Cin % 4 == 0 && (a.k & 1), T2_ERR_UNSUPPORTED, "conv1d: Cin=%d k=%d unsupported", a.Cin, a.k);
  T2_REQUIRE((a.in_bstride % 4) == 0, T2_ERR_ARG, "conv1d: batch stride must be a multiple of 4 floats");
  if (a.B * a.T == 0) return T2_OK;
  dim3 grid((a.B * a.T + CV_BM - 1) / CV_BM, (a.Cout + CV_BN - 1) / CV_BN);
  conv1d_cl_kernel<<<grid, 256, 0, st>>>(a);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int embed_launch(const long long* idx, const float* table, float* out, int n_pos, int E, int n_rows, cudaStream_t st) {
  embed_kernel<<<gs((long long)n_pos * E / 4, 256), 256, 0, st>>>(idx, table, out, n_pos, E, n_rows);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int bilstm_launch(const float* xproj, const float* WhhT, const long long* lens, float* out, int B, int L, int Hh, cudaStream_t st) {
  T2_REQUIRE(Hh <= 256 && 256 % Hh == 0, T2_ERR_UNSUPPORTED, "bilstm: hidden %d unsupported (must divide 256)", Hh);
  bilstm_kernel<<<2 * B, 256, Hh * sizeof(float), st>>>(xproj, WhhT, lens, out, B, L, Hh);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int condition_launch(const float* enc, const float* spk_table, const long long* spk_id, const float* desc_vec, int D,
                     float* out, int B, int L, int E, cudaStream_t st) {
  condition_kernel<<<gs((long long)B * L * (E + D), 256), 256, 0, st>>>(enc, spk_table, spk_id, desc_vec, D, out, B, L, E);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int desc_linear_launch(const float* desc, const float* W, const float* bias, float* out, int B, int Din, int Dout, cudaStream_t st) {
  desc_linear_kernel<<<(B * Dout * 32 + 255) / 256, 256, 0, st>>>(desc, W, bias, out, B, Din, Dout);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int finalize_launch(float* mels, float* post, float* gates, const long long* lengths, int B, int T, int Tstride, int M, cudaStream_t st) {
  if (B * T == 0) return T2_OK;
  finalize_kernel<<<gs((long long)B * T * M, 256), 256, 0, st>>>(mels, post, gates, lengths, B, T, Tstride, M);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int bn_fold_launch(const float* w, const float* b, const float* mean, const float* var, const float* conv_bias, float* scale,
                   float* shift, int C, cudaStream_t st) {
  bn_fold_kernel<<<(C + 127) / 128, 128, 0, st>>>(w, b, mean, var, conv_bias, scale, shift, C);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int repack_conv_launch(const float* W, float* Wr, int Cout, int Cin, int k, cudaStream_t st) {
  repack_conv_kernel<<<gs((long long)Cout * Cin * k, 256), 256, 0, st>>>(W, Wr, Cout, Cin, k);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int transpose_launch(const float* in, float* out, int R, int C, cudaStream_t st) {
  dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
  transpose_kernel<<<grid, block, 0, st>>>(in, out, R, C);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
