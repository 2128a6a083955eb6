// Once-per-batch layers that are not GEMM-shaped (the convolutions live in convtc.cu):
//   packed BiLSTM recurrence  model/encoder.py:59-65
//   glue                      model/tacotron2.py:201-212 (conditioning), :332-345 (residual + output masks)
#include <cstdlib>
#include <cstring>
#include "layers.cuh"
#include "tc.cuh"

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

// ------------------------------------------------------------------------------------------
// The same recurrence with the gate contraction on tcgen05 (Hh = 256): one cluster of 8 CTAs per (direction, group of 16 utterances).
// CTA `rank` owns 32 hidden units: A = its 128 gate rows of W_hh as bf16 (hi,lo) K-major sw128 tiles, resident in shared memory
// (128 KB, packed once per launch); B = h_{t-1} of the 16 utterances as bf16 (hi,lo), K-major in 32-element K chunks with 64-byte
// swizzle -- chunk c is exactly what cluster rank c produces, a contiguous 2 KB block, so the all-gather of the new hidden state is ONE
// bulk copy shared::cta -> shared::cluster per peer, counted on the receiver's transaction barrier (no cluster barrier in the loop).
// D[128 gate rows x 16 utterances] in tensor memory, three bf16 passes (fp32-parity arithmetic as everywhere else).  The CUDA-core
// form above spends 13 us per character (two shared-memory loads per four FMAs, two cluster barriers, eight remote 4-byte stores per
// thread); this one is bound by the exchange latency.
#define BT_NB 16
#define BT_A_BYTES (4 * 2 * 16384)            // [4 K blocks][hi, lo][128 rows x 128 B]
#define BT_B_BUF (8 * 2048)                   // [8 source ranks][hi, lo][16 rows x 64 B]
__host__ __device__ inline size_t bilstm_tc_smem() { return 1024 + BT_A_BYTES + 2 * BT_B_BUF + 2 * 2048 + 128 * 17 * 4 + 256; }
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t saddr) {   // K-major, 64-byte swizzle: rows 64 B apart, 8-row groups 512 B apart
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;                            // layout type SWIZZLE_64B
  return d;
}
__device__ __forceinline__ uint32_t sw64_off(int r, int k) {   // element (row r < 16, k < 32) of a [16][32] bf16 block
  return (uint32_t)(r * 64 + ((((k >> 3) & 3) ^ ((r >> 1) & 3)) << 4) + ((k & 7) << 1));
}
__global__ void __launch_bounds__(256, 1) bilstm_tc_kernel(const float* __restrict__ xproj, const float* __restrict__ Whh,
                                                           const long long* __restrict__ lens, float* __restrict__ out, int B, int L,
                                                           float* __restrict__ sav) {
  constexpr int Hh = 256, UPC = 32;
  extern __shared__ __align__(1024) uint8_t bsm_raw[];
  uint8_t* base = bsm_raw + ((1024u - (tc::smem_u32(bsm_raw) & 1023u)) & 1023u);
  uint8_t* At = base;                                     // A tiles
  uint8_t* Bt = At + BT_A_BYTES;                          // [2 buffers] B blocks
  uint8_t* stg = Bt + 2 * BT_B_BUF;                       // [2] outgoing block of this CTA (same layout as one source block)
  float* gs = reinterpret_cast<float*>(stg + 2 * 2048);   // [128][17] gate pre-activations (W_hh h part)
  uint64_t* bars = reinterpret_cast<uint64_t*>(gs + 128 * 17);   // full[2], mma_done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
  int* len_s = reinterpret_cast<int*>(tmem_slot + 2);     // [16]
  const uint32_t rank = tc::cluster_ctarank();
  const int cid = blockIdx.x / BL_CL, dir = cid & 1, grp = cid >> 1;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  {  // pack this CTA's rows of W_hh: local row lr = gate * 32 + unit
    const float* W = Whh + (size_t)dir * 4 * Hh * Hh;
#pragma unroll 1
    for (int i = tid; i < 128 * Hh; i += 256) {
      const int lr = i >> 8, k = i & 255, g = lr >> 5, ul = lr & 31;
      __nv_bfloat16 hi, lo;
      tc::split_bf16(__ldg(W + (size_t)(g * Hh + (int)rank * UPC + ul) * Hh + k), hi, lo);
      uint8_t* t = At + (size_t)(k >> 6) * 32768 + tc::sw128_off(lr, k & 63);
      *reinterpret_cast<__nv_bfloat16*>(t) = hi;
      *reinterpret_cast<__nv_bfloat16*>(t + 16384) = lo;
    }
  }
  if (tid < BT_NB) {
    const int b = grp * BT_NB + tid;
    len_s[tid] = b < B ? (int)min((long long)L, max(0ll, lens[b])) : 0;
  }
  if (tid == 0) {
    tc::mbar_init(bars + 0, 1); tc::mbar_init(bars + 1, 1); tc::mbar_init(bars + 2, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc<32>(tmem_slot);
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  int maxlen = 0;
  for (int i = 0; i < BT_NB; ++i) maxlen = max(maxlen, len_s[i]);
  tc::cluster_sync();   // barrier inits of the whole cluster are visible before any remote copy

  // pointwise role: unit pu, utterances pb and pb + 8
  const int pu = lane, pb = warp;
  float cst[2] = {0.f, 0.f}, hprev[2] = {0.f, 0.f};
  float xg[2][4];   // x W_ih^T + biases of the current step for the two (unit, utterance) pairs, requested one step ahead
  auto load_x = [&](int s, float (&x)[2][4]) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int bl = pb + 8 * j, n = len_s[bl];
#pragma unroll
      for (int g = 0; g < 4; ++g) x[j][g] = 0.f;
      if (s < n) {
        const int t = dir ? (n - 1 - s) : s;
        const float* xp = xproj + (((size_t)(grp * BT_NB + bl) * L + t) * 2 + dir) * 4 * Hh + (int)rank * UPC + pu;
#pragma unroll
        for (int g = 0; g < 4; ++g) x[j][g] = __ldg(xp + g * Hh);
      }
    }
  };
  load_x(0, xg);
  uint32_t fph[2] = {0u, 0u}, dph = 0u;
  bool ok = true;
#pragma unroll 1
  for (int s = 0; s < maxlen; ++s) {
    const int cur = s & 1, nxt = cur ^ 1;
    float xn[2][4];
    if (s + 1 < maxlen) load_x(s + 1, xn);
    if (s > 0) {
      if (tid == 0) {   // W_hh h_{s-1}: 8 K chunks of 32 (one per source rank), two K16 steps each, three passes
        if (!tc::mbar_wait_cluster(bars + cur, fph[cur])) ok = false;
        tc::tc_fence_after();
        const uint32_t idesc = tc::idesc_bf16_f32(128, BT_NB);
        const uint32_t ab = tc::smem_u32(At), bb = tc::smem_u32(Bt + (size_t)cur * BT_B_BUF);
        uint32_t acc = 0u;
#pragma unroll 1
        for (int c = 0; c < 8; ++c) {
          const uint64_t ahi = tc::smem_desc_sw128(ab + (uint32_t)(c >> 1) * 32768u), alo = tc::smem_desc_sw128(ab + (uint32_t)(c >> 1) * 32768u + 16384u);
          const uint64_t bhi = smem_desc_sw64(bb + (uint32_t)c * 2048u), blo = smem_desc_sw64(bb + (uint32_t)c * 2048u + 1024u);
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const uint64_t aadv = (uint64_t)((((c & 1) * 2 + j) * 32) >> 4), badv = (uint64_t)((j * 32) >> 4);
            tc::umma_bf16(tmem, alo + aadv, bhi + badv, idesc, acc);
            tc::umma_bf16(tmem, ahi + aadv, blo + badv, idesc, 1u);
            tc::umma_bf16(tmem, ahi + aadv, bhi + badv, idesc, 1u);
            acc = 1u;
          }
        }
        tc::umma_commit(bars + 2);
      }
      fph[cur] ^= 1u;
      if (!tc::mbar_wait(bars + 2, dph)) ok = false;
      dph ^= 1u;
      tc::tc_fence_after();
      if (warp < 4) {   // accumulator row = TMEM lane -> gs[row][utterance]
        float v[16];
        tc::tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16), v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) gs[(32 * warp + lane) * 17 + i] = v[i];
      }
      tc::tc_fence_before();
    }
    __syncthreads();
    uint8_t* sb = stg + (size_t)nxt * 2048;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int bl = pb + 8 * j, n = len_s[bl];
      float h = hprev[j];
      if (s < n) {
        float pre[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) pre[g] = xg[j][g] + (s > 0 ? gs[(g * UPC + pu) * 17 + bl] : 0.f);
        const float si = sigmoidf_(pre[0]), sf = sigmoidf_(pre[1]), tg = tanhf(pre[2]), so = sigmoidf_(pre[3]);
        cst[j] = sf * cst[j] + si * tg;
        h = so * tanhf(cst[j]);
        const int t = dir ? (n - 1 - s) : s;
        const size_t row = (size_t)(grp * BT_NB + bl) * L + t;
        if (sav) {  // training tape: [B][L][2][5][Hh] = i, f, g, o (post-activation), c
          float* sv = sav + (row * 2 + dir) * 5 * Hh + (int)rank * UPC + pu;
          sv[0] = si; sv[Hh] = sf; sv[2 * Hh] = tg; sv[3 * Hh] = so; sv[4 * Hh] = cst[j];
        }
        out[row * 2 * Hh + dir * Hh + (int)rank * UPC + pu] = h;
      }
      hprev[j] = h;
      __nv_bfloat16 hi, lo;
      tc::split_bf16(h, hi, lo);
      const uint32_t off = sw64_off(bl, pu);
      *reinterpret_cast<__nv_bfloat16*>(sb + off) = hi;
      *reinterpret_cast<__nv_bfloat16*>(sb + 1024 + off) = lo;
    }
    if (s + 1 < maxlen) {   // all-gather h_s: this CTA's block -> chunk `rank` of every CTA's next B buffer
      tc::fence_proxy_async();
      __syncthreads();
      if (tid == 0) tc::mbar_arrive_expect_tx(bars + nxt, 8u * 2048u);
      if (tid < 8) {
        const uint32_t dst = tc::mapa(tc::smem_u32(Bt + (size_t)nxt * BT_B_BUF + (size_t)rank * 2048), (uint32_t)tid);
        const uint32_t rb = tc::mapa(tc::smem_u32(bars + nxt), (uint32_t)tid);
        asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "r"(tc::smem_u32(sb)), "r"(2048u), "r"(rb) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the block sent two steps ago (same staging buffer next time) is out
      }
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int g = 0; g < 4; ++g) xg[j][g] = xn[j][g];
    }
  }
  if (tid < 8) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  tc::tc_fence_before();
  __syncthreads();
  tc::cluster_sync();   // no CTA exits while a peer may still copy into its shared memory
  if (warp == 0) tc::tmem_dealloc<32>(tmem);
  (void)ok;
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
  if (Hh == 256 && !(getenv("T2B200_BILSTM") && !strcmp(getenv("T2B200_BILSTM"), "cuda"))) {   // tensor-core form (the reference's size)
    const size_t smem = bilstm_tc_smem();
    T2_CHECK_CUDA(cudaFuncSetAttribute(bilstm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * ((B + BT_NB - 1) / BT_NB) * BL_CL));
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = BL_CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    T2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, bilstm_tc_kernel, xproj, Whh, lens, out, B, L, sav));
    return T2_OK;
  }
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
