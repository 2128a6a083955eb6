// Persistent autoregressive decoder kernel (one cooperative launch for the whole utterance
// batch; grid = one CTA per SM).  Per frame it runs, separated by grid barriers:
//   P1 attention LSTMCell     (model/decoder.py:70-75)
//   P2 query projection       (model/attention.py:52)
//   P3 location-sensitive attention: conv31 -> dense -> tanh energies -> masked softmax
//      -> context, cumulative-weight update (model/attention.py:54-67, model/decoder.py:90)
//   P4 decoder LSTMCell       (model/decoder.py:94-101)
//   P5 mel / gate projection + stop logic (model/decoder.py:103-109, model/tacotron2.py:311-322)
//   P6/P7 prenet layers on the emitted frame (model/tacotron2.py:85-92,324-325)
// All arithmetic is fp32 FMA in this revision (exact-parity engine).
#include "decoder.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// Grid barrier: monotonic arrival counter.  bar[0] = counter, bar[1] = abort flag.
// Cooperative launch guarantees co-residency; the spin is bounded so a bug cannot hang the GPU.
struct GridBar {
  unsigned* ctr;
  unsigned* abort_flag;
  unsigned nblocks;
  unsigned epoch;
  __device__ bool sync() {  // returns false if the kernel must abort
    __syncthreads();
    __shared__ int ok_s;
    if (threadIdx.x == 0) {
      ++epoch;
      __threadfence();
      atomicAdd(ctr, 1u);
      const unsigned target = epoch * nblocks;
      long long spins = 0;
      int ok = 1;
      while (ld_acquire_u32(ctr) < target) {
        if (++spins > (1ll << 26)) {
          if (ld_acquire_u32(abort_flag) != 0u || spins > (1ll << 28)) { atomicExch(abort_flag, 1u); ok = 0; break; }
        }
      }
      if (ld_acquire_u32(abort_flag) != 0u) ok = 0;
      __threadfence();
      ok_s = ok;
    }
    __syncthreads();
    return ok_s != 0;
  }
};

// ------------------------------------------------------------------------------------------
// Small dense phases: out[b][r] = sum_k W[r][k] * x[b][k], W row-major [rows][ldw], x row-major
// [B][ldx] (mutable -> ld.global.cg).  One warp owns a 4x4 (rows x batch) tile, lanes split K in
// float4 steps, butterfly reduction, then lane (i*4+j) applies the epilogue to output (r0+i, b0+j).
template <typename Epi>
__device__ __forceinline__ void warp_gemm_rm(const float* __restrict__ W, int ldw, int rows, const float* x, int ldx,
                                             int B, int K, Epi epi) {
  const int lane = threadIdx.x & 31;
  const int gwarp = blockIdx.x * (DEC_THREADS / 32) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (DEC_THREADS / 32);
  const int rt = (rows + 3) >> 2, bt = (B + 3) >> 2;
  const int K4 = K >> 2;
  for (int task = gwarp; task < rt * bt; task += nwarps) {
    const int r0 = (task / bt) * 4, b0 = (task % bt) * 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const float* wp[4];
    const float* xp[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      wp[i] = W + (size_t)min(r0 + i, rows - 1) * ldw;
      xp[i] = x + (size_t)min(b0 + i, B - 1) * ldx;
    }
    for (int k4 = lane; k4 < K4; k4 += 32) {
      float4 wv[4], xv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        wv[i] = __ldg(reinterpret_cast<const float4*>(wp[i]) + k4);
        xv[i] = __ldcg(reinterpret_cast<const float4*>(xp[i]) + k4);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          acc[i][j] = fmaf(wv[i].x, xv[j].x, acc[i][j]);
          acc[i][j] = fmaf(wv[i].y, xv[j].y, acc[i][j]);
          acc[i][j] = fmaf(wv[i].z, xv[j].z, acc[i][j]);
          acc[i][j] = fmaf(wv[i].w, xv[j].w, acc[i][j]);
        }
    }
    float mine = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float s = warp_sum(acc[i][j]);
        if (lane == i * 4 + j) mine = s;
      }
    if (lane < 16) {
      const int r = r0 + (lane >> 2), b = b0 + (lane & 3);
      if (r < rows && b < B) epi(r, b, mine);
    }
  }
}

// ------------------------------------------------------------------------------------------
// LSTM gate tile: CTA `tile` owns 8 hidden units x 4 gates = 32 gate rows for all DEC_BP batch
// slots.  gates[row][b] = sum_k Wp[tile][k][row] * x[k][b]; x is k-major and comes from up to
// three row ranges (segments) so the prenet part can live in its own buffer.
// 256 threads = 4 k-groups x (2 row-halves x 32 lanes); thread tile 4 rows x 8 batch.
#define LT_KC 64
#define LT_STAGES 3
#define LT_STAGE_FLOATS (LT_KC * 32 + LT_KC * DEC_BP)

struct XSeg { const float* ptr; int rows; };  // consecutive k-major row ranges forming x

__device__ __forceinline__ const float* xrow(const XSeg* seg, int k) {
  if (k < seg[0].rows) return seg[0].ptr + (size_t)k * DEC_BP;
  k -= seg[0].rows;
  return seg[1].ptr + (size_t)k * DEC_BP;
}

__device__ __forceinline__ void lt_fill(float* stage, const float* Wp_tile, const XSeg* seg, int k0) {
  float* Ws = stage;                  // [KC][32]
  float* xs = stage + LT_KC * 32;     // [KC][BP]
  const int tid = threadIdx.x;
  // W: KC*32 floats contiguous in the packed layout
  const float* wsrc = Wp_tile + (size_t)k0 * 32;
#pragma unroll
  for (int i = 0; i < (LT_KC * 32 / 4) / DEC_THREADS; ++i) {
    int idx = tid + i * DEC_THREADS;
    cp_async16(Ws + idx * 4, wsrc + idx * 4);
  }
#pragma unroll
  for (int i = 0; i < (LT_KC * DEC_BP / 4) / DEC_THREADS; ++i) {
    int idx = tid + i * DEC_THREADS;  // float4 index
    int k = idx / (DEC_BP / 4), c = idx % (DEC_BP / 4);
    cp_async16(xs + k * DEC_BP + c * 4, xrow(seg, k0 + k) + c * 4);
  }
}

// Result: red[4][32][BP] partial sums in shared memory (caller reduces over the 4 k-groups).
__device__ void lstm_tile_gemm(const float* Wp_tile, const XSeg* seg, int KP, float* smem) {
  float* red = smem;  // aliases the stages after the main loop (32 KB <= 3 stages)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kg = warp >> 1, wr = warp & 1;
  const int tr = lane >> 3, tb = lane & 7;
  const int row0 = wr * 16 + tr * 4, col0 = tb * 4;  // batch columns {col0..col0+3} and {32+col0..}
  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  const int nchunk = KP / LT_KC;
#pragma unroll
  for (int s = 0; s < LT_STAGES - 1; ++s) {
    if (s < nchunk) lt_fill(smem + s * LT_STAGE_FLOATS, Wp_tile, seg, s * LT_KC);
    cp_async_commit();
  }
  for (int c = 0; c < nchunk; ++c) {
    cp_async_wait<LT_STAGES - 2>();
    __syncthreads();
    {
      int nc = c + LT_STAGES - 1;
      if (nc < nchunk) lt_fill(smem + (nc % LT_STAGES) * LT_STAGE_FLOATS, Wp_tile, seg, nc * LT_KC);
      cp_async_commit();
    }
    const float* Ws = smem + (c % LT_STAGES) * LT_STAGE_FLOATS;
    const float* xs = Ws + LT_KC * 32;
#pragma unroll 4
    for (int kk = 0; kk < LT_KC / 4; ++kk) {
      const int k = kg * (LT_KC / 4) + kk;
      const float4 w4 = *reinterpret_cast<const float4*>(Ws + k * 32 + row0);
      const float4 xa = *reinterpret_cast<const float4*>(xs + k * DEC_BP + col0);
      const float4 xb = *reinterpret_cast<const float4*>(xs + k * DEC_BP + 32 + col0);
      const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
      const float xv[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(wv[i], xv[j], acc[i][j]);
    }
  }
  cp_async_wait<0>();
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float* dst = red + ((size_t)kg * 32 + row0 + i) * DEC_BP + col0;
    *reinterpret_cast<float4*>(dst) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    *reinterpret_cast<float4*>(dst + 32) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------
// LSTM pointwise on the reduced gate tile.  Thread handles (unit_in_tile, b) pairs.
struct HDest {
  float* km0; int km0_row;   // k-major destination [row][BP] (may be null)
  float* km1; int km1_row;
  float* rm;  int rm_ld, rm_col;  // row-major destination [b][ld]
  float* rm2; int rm2_ld, rm2_col;
};

__device__ void lstm_pointwise(const float* red, const float* __restrict__ bias, int H, int tile, float* cbuf, int B,
                               const HDest& hd, const unsigned char* keep, bool hidden_dropout, unsigned long long seed,
                               unsigned long long rng_base) {
  for (int idx = threadIdx.x; idx < 8 * DEC_BP; idx += DEC_THREADS) {
    const int ul = idx / DEC_BP, b = idx % DEC_BP;
    const int u = tile * 8 + ul;
    float pre[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int row = g * 8 + ul;
      float s = red[(0 * 32 + row) * DEC_BP + b] + red[(1 * 32 + row) * DEC_BP + b];
      s += red[(2 * 32 + row) * DEC_BP + b] + red[(3 * 32 + row) * DEC_BP + b];
      pre[g] = s + __ldg(bias + g * H + u);
    }
    float h = 0.f, c = 0.f;
    if (b < B) {
      const float c_old = cbuf[(size_t)u * DEC_BP + b];
      c = sigmoidf_(pre[1]) * c_old + sigmoidf_(pre[0]) * tanhf(pre[2]);
      h = sigmoidf_(pre[3]) * tanhf(c);
      if (hidden_dropout) {  // nn.Dropout(0.1), train only (model/decoder.py:75,101)
        bool k = keep ? (keep[(size_t)b * H + u] != 0) : rng_keep(seed, rng_base + (unsigned long long)b * H + u, 0.1f);
        h = k ? h * (1.0f / 0.9f) : 0.f;
      }
    }
    cbuf[(size_t)u * DEC_BP + b] = c;
    if (hd.km0) hd.km0[(size_t)(hd.km0_row + u) * DEC_BP + b] = h;
    if (hd.km1) hd.km1[(size_t)(hd.km1_row + u) * DEC_BP + b] = h;
    if (b < B) {
      if (hd.rm) hd.rm[(size_t)b * hd.rm_ld + hd.rm_col + u] = h;
      if (hd.rm2) hd.rm2[(size_t)b * hd.rm2_ld + hd.rm2_col + u] = h;
    }
  }
}

// ------------------------------------------------------------------------------------------
// Attention for one (batch element, column split).  Shared-memory plan (floats):
//   att_w: conv[32*2*31] | dense[A*32] | v[A]          (loaded once per launch)
//   att_s: q[A] | wp[L+30] | wc[L+30] | e[L] | red[64] | cred[4*colsmax]
struct AttSmem {
  float *conv, *dense, *v, *q, *wp, *wc, *e, *red, *cred;
};

__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int i = 1; i < DEC_THREADS / 32; ++i) r = is_max ? fmaxf(r, red[i]) : r + red[i];
  return r;
}

__device__ void attention_task(const DecDims& d, const DecState& s, const DecIO& io, const AttSmem& sm, int b, int split,
                               int nsplit, int cur, int t_out) {
  const int tid = threadIdx.x, L = d.L, A = d.A, EF = d.EF;
  const int nxt = cur ^ 1;
  const float* wprev = s.w_prev + ((size_t)cur * DEC_BP + b) * L;
  const float* wcum = s.w_cum + ((size_t)cur * DEC_BP + b) * L;
  for (int i = tid; i < A; i += DEC_THREADS) sm.q[i] = ld_cg(s.q + (size_t)b * A + i);
  for (int i = tid; i < L + 30; i += DEC_THREADS) {
    const int l = i - 15;
    const bool in = (l >= 0 && l < L);
    sm.wp[i] = in ? ld_cg(wprev + l) : 0.f;
    sm.wc[i] = in ? ld_cg(wcum + l) : 0.f;
  }
  __syncthreads();
  const unsigned char* mk = io.mask + (size_t)b * L;
  // energies
  for (int l = tid; l < L; l += DEC_THREADS) {
    float e = -INFINITY;
    if (mk[l] == 0) {
      float f[DEC_LOC_F];
#pragma unroll
      for (int i = 0; i < DEC_LOC_F; ++i) f[i] = 0.f;
      for (int j = 0; j < DEC_LOC_K; ++j) {
        const float a0 = sm.wp[l + j], a1 = sm.wc[l + j];
#pragma unroll
        for (int i = 0; i < DEC_LOC_F; ++i) {
          f[i] = fmaf(sm.conv[(i * 2 + 0) * DEC_LOC_K + j], a0, f[i]);
          f[i] = fmaf(sm.conv[(i * 2 + 1) * DEC_LOC_K + j], a1, f[i]);
        }
      }
      const float* pm = io.pmem + ((size_t)b * L + l) * A;
      e = 0.f;
      for (int a = 0; a < A; a += 4) {
        const float4 p4 = __ldg(reinterpret_cast<const float4*>(pm + a));
        const float pv[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
        for (int aa = 0; aa < 4; ++aa) {
          const float4* dr = reinterpret_cast<const float4*>(sm.dense + (a + aa) * DEC_LOC_F);
          float loc = 0.f;
#pragma unroll
          for (int i = 0; i < DEC_LOC_F / 4; ++i) {
            const float4 w4 = dr[i];
            loc = fmaf(w4.x, f[4 * i + 0], loc);
            loc = fmaf(w4.y, f[4 * i + 1], loc);
            loc = fmaf(w4.z, f[4 * i + 2], loc);
            loc = fmaf(w4.w, f[4 * i + 3], loc);
          }
          e = fmaf(sm.v[a + aa], tanhf(sm.q[a + aa] + loc + pv[aa]), e);
        }
      }
    }
    sm.e[l] = e;
  }
  __syncthreads();
  // masked softmax over L (model/attention.py:63-65)
  float mx = -INFINITY;
  for (int l = tid; l < L; l += DEC_THREADS) mx = fmaxf(mx, sm.e[l]);
  mx = block_reduce(mx, sm.red, true);
  float sum = 0.f;
  for (int l = tid; l < L; l += DEC_THREADS) {
    const float p = expf(sm.e[l] - mx);
    sm.e[l] = p;
    sum += p;
  }
  sum = block_reduce(sum, sm.red + 32, false);
  const float inv = 1.0f / sum;
  for (int l = tid; l < L; l += DEC_THREADS) {
    const float w = sm.e[l] * inv;
    sm.e[l] = w;
    if (split == 0) {
      s.w_prev[((size_t)nxt * DEC_BP + b) * L + l] = w;
      s.w_cum[((size_t)nxt * DEC_BP + b) * L + l] = sm.wc[l + 15] + w;  // model/decoder.py:90
      io.aligns[((size_t)b * io.Tmax + t_out) * L + l] = w;
    }
  }
  __syncthreads();
  // context columns [c0, c1) of this split (model/attention.py:66-67)
  const int cols_per = ((EF + nsplit - 1) / nsplit + 3) & ~3;
  const int c0 = split * cols_per, c1 = min(EF, c0 + cols_per);
  const int nq = (c1 - c0 + 3) >> 2;  // float4 columns
  const float* mem = io.memory + (size_t)b * L * EF;
  const int lg = tid >> 6, cq0 = tid & 63;
  for (int cq = cq0; cq < nq; cq += 64) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int l = lg; l < L; l += 4) {
      const float w = sm.e[l];
      if (w != 0.f) {
        const float4 m4 = __ldg(reinterpret_cast<const float4*>(mem + (size_t)l * EF + c0) + cq);
        acc.x = fmaf(w, m4.x, acc.x); acc.y = fmaf(w, m4.y, acc.y);
        acc.z = fmaf(w, m4.z, acc.z); acc.w = fmaf(w, m4.w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(sm.cred + ((size_t)lg * cols_per) + cq * 4) = acc;
  }
  __syncthreads();
  for (int c = tid; c < c1 - c0; c += DEC_THREADS) {
    const float v = sm.cred[c] + sm.cred[cols_per + c] + sm.cred[2 * cols_per + c] + sm.cred[3 * cols_per + c];
    const int col = c0 + c;
    s.xatt[nxt][(size_t)col * DEC_BP + b] = v;             // next frame's attention-LSTM input
    s.xdec[cur][(size_t)(d.H + col) * DEC_BP + b] = v;     // this frame's decoder-LSTM input
    s.projin[(size_t)b * d.KprojP + d.R + col] = v;        // this frame's projection input
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DEC_THREADS, 1)
decoder_persistent_kernel(DecDims d, DecWeights w, DecState s, DecIO io) {
  extern __shared__ __align__(16) float smem[];
  float* gemm_smem = smem;                                  // LT_STAGES * LT_STAGE_FLOATS
  AttSmem sm;
  float* p = smem + LT_STAGES * LT_STAGE_FLOATS;
  sm.conv = p;  p += DEC_LOC_F * 2 * DEC_LOC_K;
  sm.dense = p; p += d.A * DEC_LOC_F;
  sm.v = p;     p += d.A;
  sm.q = p;     p += d.A;
  sm.wp = p;    p += d.L + 30 + 2;
  sm.wc = p;    p += d.L + 30 + 2;
  sm.e = p;     p += d.L + 4;
  sm.red = p;   p += 64;
  p = smem + (((p - smem) + 3) & ~3);
  sm.cred = p;

  const int tid = threadIdx.x;
  for (int i = tid; i < DEC_LOC_F * 2 * DEC_LOC_K; i += DEC_THREADS) sm.conv[i] = w.loc_conv[i];
  for (int i = tid; i < d.A * DEC_LOC_F; i += DEC_THREADS) sm.dense[i] = w.loc_dense[i];
  for (int i = tid; i < d.A; i += DEC_THREADS) sm.v[i] = w.v[i];
  __syncthreads();

  GridBar bar{s.barrier, s.barrier + 1, gridDim.x, 0u};
  long long prof_last = clock64();
  auto prof_mark = [&](int slot) {
    if (blockIdx.x == 0 && threadIdx.x == 0 && s.prof) { const long long now = clock64(); s.prof[slot] += (unsigned long long)(now - prof_last); prof_last = now; }
  };
  const int B = d.B;
  const int n_att_tiles = d.H / 8, n_dec_tiles = d.R / 8;
  const int nsplit = max(1, min((int)gridDim.x / B, d.EF / 64));
  const float pscale = 1.0f / (1.0f - d.p_prenet);
  int frames = 0;
  bool alive = true;

  for (int t = 0; t < io.n_steps && alive; ++t) {
    const int cur = t & 1, nxt = cur ^ 1;
    const int tg = io.t0 + t;  // global frame index (mask / output row)

    // ---- P1: attention LSTMCell -------------------------------------------------------
    for (int tile = blockIdx.x; tile < n_att_tiles; tile += gridDim.x) {
      XSeg seg[2];
      seg[0].ptr = io.teacher_forcing ? io.tf_prenet + (size_t)tg * d.P * DEC_BP : s.xpre[cur];
      seg[0].rows = d.P;
      seg[1].ptr = s.xatt[cur];
      seg[1].rows = d.KattP - d.P;
      lstm_tile_gemm(w.Watt_p + (size_t)tile * d.KattP * 32, seg, d.KattP, gemm_smem);
      HDest hd{s.xatt[nxt], d.EF, s.xdec[cur], 0, s.att_h_rm, d.H, 0, nullptr, 0, 0};
      lstm_pointwise(gemm_smem, w.batt, d.H, tile, s.att_c, B, hd,
                     io.att_keep ? io.att_keep + (size_t)tg * B * d.H : nullptr, io.hidden_dropout != 0, io.seed,
                     (1ull << 40) + (unsigned long long)tg * B * d.H);
      __syncthreads();
    }
    if (!(alive = bar.sync())) break;
    prof_mark(0);

    // ---- P2: query projection (model/attention.py:52) ---------------------------------
    warp_gemm_rm(w.Wq, d.H, d.A, s.att_h_rm, d.H, B, d.H,
                 [&](int r, int b, float v) { s.q[(size_t)b * d.A + r] = v; });
    if (!(alive = bar.sync())) break;
    prof_mark(1);

    // ---- P3: location-sensitive attention ---------------------------------------------
    for (int task = blockIdx.x; task < B * nsplit; task += gridDim.x)
      attention_task(d, s, io, sm, task / nsplit, task % nsplit, nsplit, cur, tg);
    if (!(alive = bar.sync())) break;
    prof_mark(2);

    // ---- P4: decoder LSTMCell ----------------------------------------------------------
    for (int tile = blockIdx.x; tile < n_dec_tiles; tile += gridDim.x) {
      XSeg seg[2];
      seg[0].ptr = s.xdec[cur];
      seg[0].rows = d.KdecP;
      seg[1].ptr = s.xdec[cur];
      seg[1].rows = 0;
      lstm_tile_gemm(w.Wdec_p + (size_t)tile * d.KdecP * 32, seg, d.KdecP, gemm_smem);
      HDest hd{s.xdec[nxt], d.H + d.EF + d.C, nullptr, 0, s.projin, d.KprojP, 0,
               io.projin_all ? io.projin_all + (size_t)tg * B * d.KprojP : nullptr, d.KprojP, 0};
      lstm_pointwise(gemm_smem, w.bdec, d.R, tile, s.rnn_c, B, hd,
                     io.rnn_keep ? io.rnn_keep + (size_t)tg * B * d.R : nullptr, io.hidden_dropout != 0, io.seed,
                     (2ull << 40) + (unsigned long long)tg * B * d.R);
      __syncthreads();
    }
    if (!(alive = bar.sync())) break;
    prof_mark(3);

    // ---- P5: mel / gate projection + stop bookkeeping ----------------------------------
    warp_gemm_rm(w.Wproj, d.KprojP, d.M + 1, s.projin, d.KprojP, B, d.KprojP, [&](int r, int b, float v) {
      v += __ldg(w.bproj + r);
      if (r < d.M) {
        io.mels[((size_t)b * io.Tmax + tg) * d.M + r] = v;
        s.mel_cur[(size_t)b * d.M + r] = v;
      } else {
        io.gates[(size_t)b * io.Tmax + tg] = v;
        if (io.stop_enabled) {  // model/tacotron2.py:319-320
          if (v < 0.0f) s.done[b] = 1;
          if (v >= 0.0f) s.lengths[b] += 1;
        }
      }
    });
    // TF mode also stashes ctx / controls of projin for the backward pass
    if (io.projin_all) {
      for (int i = blockIdx.x * DEC_THREADS + tid; i < B * (d.KprojP - d.R); i += gridDim.x * DEC_THREADS) {
        const int b = i / (d.KprojP - d.R), c = d.R + i % (d.KprojP - d.R);
        io.projin_all[((size_t)tg * B + b) * d.KprojP + c] = ld_cg(s.projin + (size_t)b * d.KprojP + c);
      }
    }
    frames = t + 1;
    if (!(alive = bar.sync())) break;
    prof_mark(4);
    if (io.teacher_forcing) continue;

    if (io.stop_enabled) {  // model/tacotron2.py:321-322: break once every item has fired
      int mine = 1;
      for (int b = tid; b < B; b += DEC_THREADS) mine &= (ld_acquire_u32((const unsigned*)s.done + b) != 0u);
      if (__syncthreads_and(mine)) break;
    }
    if (t + 1 >= io.n_steps && io.n_steps > 1) break;  // last frame: its prenet output is never used

    // ---- P6 / P7: prenet on the emitted frame (model/tacotron2.py:324-325) -------------
    const unsigned char* k1 = io.prenet_keep ? io.prenet_keep + ((size_t)(tg + 1) * 2 + 0) * B * d.P : nullptr;
    const unsigned char* k2 = io.prenet_keep ? io.prenet_keep + ((size_t)(tg + 1) * 2 + 1) * B * d.P : nullptr;
    warp_gemm_rm(w.W1, d.M, d.P, s.mel_cur, d.M, B, d.M, [&](int r, int b, float v) {
      const bool keep = k1 ? (k1[(size_t)b * d.P + r] != 0)
                           : rng_keep(io.seed, ((unsigned long long)(tg + 1) * 2 + 0) * B * d.P + (size_t)b * d.P + r, d.p_prenet);
      s.p1[(size_t)b * d.P + r] = (v > 0.f && keep) ? v * pscale : 0.f;
    });
    if (!(alive = bar.sync())) break;
    prof_mark(5);
    warp_gemm_rm(w.W2, d.P, d.P, s.p1, d.P, B, d.P, [&](int r, int b, float v) {
      const bool keep = k2 ? (k2[(size_t)b * d.P + r] != 0)
                           : rng_keep(io.seed, ((unsigned long long)(tg + 1) * 2 + 1) * B * d.P + (size_t)b * d.P + r, d.p_prenet);
      s.xpre[nxt][(size_t)r * DEC_BP + b] = (v > 0.f && keep) ? v * pscale : 0.f;
    });
    if (!(alive = bar.sync())) break;
    prof_mark(6);
  }
  if (blockIdx.x == 0 && tid == 0) *s.n_frames = alive ? frames : -1;
}

}  // namespace

size_t dec_smem_bytes(const DecDims& d) {
  size_t f = (size_t)LT_STAGES * LT_STAGE_FLOATS + DEC_LOC_F * 2 * DEC_LOC_K + (size_t)d.A * DEC_LOC_F + 2 * d.A +
             2 * (d.L + 32) + d.L + 4 + 64 + 8 + 4 * (size_t)((d.EF + 3) & ~3);
  return f * sizeof(float);
}

int dec_launch(const DecDims& d, const DecWeights& w, const DecState& s, const DecIO& io, int num_sms, cudaStream_t stream) {
  T2_REQUIRE(d.B >= 1 && d.B <= DEC_BP, T2_ERR_ARG, "decoder: batch %d not in [1,%d] per launch", d.B, DEC_BP);
  T2_REQUIRE(d.H % 8 == 0 && d.R % 8 == 0 && d.EF % 4 == 0 && d.M % 4 == 0 && d.P % 4 == 0 && d.A <= DEC_MAX_A && d.A % 4 == 0,
             T2_ERR_UNSUPPORTED, "decoder: unsupported dims H=%d R=%d EF=%d M=%d P=%d A=%d", d.H, d.R, d.EF, d.M, d.P, d.A);
  T2_REQUIRE(d.KattP % LT_KC == 0 && d.KdecP % LT_KC == 0 && d.P % LT_KC == 0, T2_ERR_UNSUPPORTED, "decoder: K padding");
  const size_t smem = dec_smem_bytes(d);
  T2_REQUIRE(smem <= 227 * 1024, T2_ERR_UNSUPPORTED, "decoder: L=%d needs %zu B shared memory", d.L, smem);
  T2_CHECK_CUDA(cudaFuncSetAttribute(decoder_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  T2_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, decoder_persistent_kernel, DEC_THREADS, smem));
  T2_REQUIRE(per_sm >= 1, T2_ERR_UNSUPPORTED, "decoder: kernel does not fit on an SM");
  T2_CHECK_CUDA(cudaMemsetAsync(s.barrier, 0, 2 * sizeof(unsigned), stream));
  DecDims dd = d; DecWeights ww = w; DecState ss = s; DecIO ii = io;
  void* args[] = {&dd, &ww, &ss, &ii};
  T2_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)decoder_persistent_kernel, dim3(num_sms), dim3(DEC_THREADS), args, smem, stream));
  return T2_OK;
}
