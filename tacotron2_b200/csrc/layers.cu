// Once-per-batch layers that are not GEMM-shaped (the convolutions live in convtc.cu):
//   packed BiLSTM recurrence  model/encoder.py:59-65
//   glue                      model/tacotron2.py:201-212 (conditioning), :332-345 (residual + output masks)
#include "layers.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// Packed BiLSTM recurrence (model/encoder.py:59-65).  One cluster of 8 CTAs per (direction, group of 8 utterances):
// each CTA keeps the W_hh rows of Hh/8 hidden units (all four gates) resident in shared memory for the whole
// sequence, computes their gates for the 8 utterances, and broadcasts the new hidden values to the 8 CTAs through
// distributed shared memory; one cluster barrier per time step.  xproj already holds x W_ih^T + b_ih + b_hh.
// The reverse direction starts at each utterance's own last character; positions >= len stay zero (pre-zeroed output).
#define BL_NB 8
#define BL_CL 8
__global__ void __launch_bounds__(256, 1) bilstm_cluster_kernel(const float* __restrict__ xproj, const float* __restrict__ Whh,
                                                                const long long* __restrict__ lens, float* __restrict__ out,
                                                                int B, int L, int Hh, float* __restrict__ sav) {
  extern __shared__ __align__(16) float bsm[];
  const int UPC = Hh / BL_CL, R = 4 * UPC, ldw = Hh + 1;
  float* Ws = bsm;                              // [R][Hh+1]
  float* hb = Ws + (size_t)R * ldw;             // [2][Hh][BL_NB]
  hb += (4 - ((hb - bsm) & 3)) & 3;
  float* gs = hb + 2 * Hh * BL_NB;              // [R][BL_NB+1]
  __shared__ int len_s[BL_NB];
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int cid = blockIdx.x / BL_CL, dir = cid & 1, grp = cid >> 1;
  const int tid = threadIdx.x;
  const float* W = Whh + (size_t)dir * 4 * Hh * Hh;  // [4Hh][Hh] of this direction
  for (int i = tid; i < R * Hh; i += blockDim.x) {
    const int lr = i / Hh, k = i - lr * Hh;
    const int g = lr / UPC, ul = lr - g * UPC;
    Ws[lr * ldw + k] = W[(size_t)(g * Hh + rank * UPC + ul) * Hh + k];
  }
  for (int i = tid; i < 2 * Hh * BL_NB; i += blockDim.x) hb[i] = 0.f;
  if (tid < BL_NB) {
    const int b = grp * BL_NB + tid;
    len_s[tid] = b < B ? (int)min((long long)L, max(0ll, lens[b])) : 0;
  }
  __syncthreads();
  int maxlen = 0;
  for (int i = 0; i < BL_NB; ++i) maxlen = max(maxlen, len_s[i]);
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const int r = tid % R, half = tid / R;          // gate-row role (needs 2R <= blockDim)
  const int gq = r / UPC, uq = r - gq * UPC;
  const int pu = tid % UPC, pb = tid / UPC;       // pointwise role: (unit, utterance), needs UPC * BL_NB <= blockDim
  float c = 0.f, hprev = 0.f;
  const uint32_t hb_addr = (uint32_t)__cvta_generic_to_shared(hb);
  for (int s = 0; s < maxlen; ++s) {
    const int cur = s & 1, nxt = cur ^ 1;
    if (half < 2) {
      float acc[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int bl = half * 4 + j, n = len_s[bl];
        float x = 0.f;
        if (s < n) {
          const int t = dir ? (n - 1 - s) : s;
          x = __ldg(xproj + (((size_t)(grp * BL_NB + bl) * L + t) * 2 + dir) * 4 * Hh + gq * Hh + rank * UPC + uq);
        }
        acc[j] = x;
      }
      const float* wr = Ws + r * ldw;
      const float* hc = hb + (size_t)cur * Hh * BL_NB + half * 4;
#pragma unroll 8
      for (int k = 0; k < Hh; ++k) {
        const float w = wr[k];
        const float4 h4 = *reinterpret_cast<const float4*>(hc + k * BL_NB);
        acc[0] = fmaf(w, h4.x, acc[0]); acc[1] = fmaf(w, h4.y, acc[1]);
        acc[2] = fmaf(w, h4.z, acc[2]); acc[3] = fmaf(w, h4.w, acc[3]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) gs[r * (BL_NB + 1) + half * 4 + j] = acc[j];
    }
    __syncthreads();
    if (pb < BL_NB) {
      const int n = len_s[pb];
      float h = hprev;
      if (s < n) {
        const float gi = gs[(0 * UPC + pu) * (BL_NB + 1) + pb], gf = gs[(1 * UPC + pu) * (BL_NB + 1) + pb];
        const float gg = gs[(2 * UPC + pu) * (BL_NB + 1) + pb], go = gs[(3 * UPC + pu) * (BL_NB + 1) + pb];
        const float si = sigmoidf_(gi), sf = sigmoidf_(gf), tg = tanhf(gg), so = sigmoidf_(go);
        c = sf * c + si * tg;
        h = so * tanhf(c);
        const int t = dir ? (n - 1 - s) : s;
        if (sav) {  // training tape: [B][L][2][5][Hh] = i, f, g, o (post-activation), c
          float* sv = sav + (((size_t)(grp * BL_NB + pb) * L + t) * 2 + dir) * 5 * Hh + rank * UPC + pu;
          sv[0] = si; sv[Hh] = sf; sv[2 * Hh] = tg; sv[3 * Hh] = so; sv[4 * Hh] = c;
        }
        out[((size_t)(grp * BL_NB + pb) * L + t) * 2 * Hh + dir * Hh + rank * UPC + pu] = h;
      }
      hprev = h;
      const uint32_t off = hb_addr + (uint32_t)(((size_t)nxt * Hh * BL_NB + (size_t)(rank * UPC + pu) * BL_NB + pb) * 4);
#pragma unroll
      for (uint32_t pr = 0; pr < BL_CL; ++pr) {
        uint32_t ra;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(off), "r"(pr));
        asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(h) : "memory");
      }
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
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

int bilstm_launch(const float* xproj, const float* Whh, const long long* lens, float* out, int B, int L, int Hh, cudaStream_t st,
                  float* sav) {
  const int UPC = Hh / BL_CL;
  T2_REQUIRE(Hh % BL_CL == 0 && 8 * UPC <= 256 && UPC * BL_NB <= 256 && Hh % 4 == 0, T2_ERR_UNSUPPORTED,
             "bilstm: hidden size %d unsupported (needs Hh %% 8 == 0 and Hh <= 256)", Hh);
  T2_CHECK_CUDA(cudaMemsetAsync(out, 0, (size_t)B * L * 2 * Hh * sizeof(float), st));
  const size_t smem = ((size_t)4 * UPC * (Hh + 1) + 4 + 2 * (size_t)Hh * BL_NB + (size_t)4 * UPC * (BL_NB + 1)) * sizeof(float);
  T2_CHECK_CUDA(cudaFuncSetAttribute(bilstm_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(2 * ((B + BL_NB - 1) / BL_NB) * BL_CL));
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = BL_CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  T2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, bilstm_cluster_kernel, xproj, Whh, lens, out, B, L, Hh, sav));
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

int transpose_launch(const float* in, float* out, int R, int C, cudaStream_t st) {
  dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
  transpose_kernel<<<grid, block, 0, st>>>(in, out, R, C);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
