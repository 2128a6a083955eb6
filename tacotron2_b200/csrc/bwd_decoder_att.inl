// Attention backward for one frame (included by bwd_decoder.cu).  CTA = (utterance b, part `part` of S; the S parts of one
// utterance sit in the same cluster).  Forward being differentiated (model/attention.py:52-67, model/decoder.py:78-90):
//   f = conv1d([w_{t-1}; cum_{t-1}]) (2->32, k 31);  s = tanh(q + f Wd^T + pm);  e = s . v (masked);  w = softmax(e);
//   ctx = w^T memory;  cum_t = cum_{t-1} + w.
namespace {

struct AttAcc {           // per-thread accumulators that live for the whole utterance
  float wd[16];           // d location_dense: (a-quad = lane, f-quad = warp) 4x4 tile
  float wc[8];            // d location_conv:  (f-quad = tid&7, cj = tid>>3 and +32)
  float dv[4];            // d v: lane's 4 attention dims (partial over the positions this warp handled)
};

// scratch (floats): q[A] | wp[Lh4] | wc[Lh4] | fs[Lc][36] | dctx[EF] | ws[Lc4] | gw[Lc4] | de[Lc4] | red[64] | xch[8] |
//                   dqw[8][A] | ds[Lc][A+4] | dfe[Lc+30][36] | gsm[2][31][Lc+30]
__device__ __forceinline__ size_t att_bwd_scratch_floats(int L, int S, int A, int EF) {
  const int Lc = (L + S - 1) / S, Lh4 = (Lc + 30 + 3) & ~3, Lc4 = (Lc + 3) & ~3;
  return (size_t)A + 2 * Lh4 + (size_t)Lc * 36 + EF + 3 * Lc4 + 64 + 8 + 8 * (size_t)A + (size_t)Lc * (A + 4) + (size_t)(Lc + 30) * 36 + (size_t)(Lc + 30) * 64;
}

#define ABW_MARK(slot) do { if (blockIdx.x == 0 && threadIdx.x == 0 && prof) { const long long now_ = clock64(); prof[slot] += (unsigned long long)(now_ - *prof_last); *prof_last = now_; } } while (0)
__device__ void attention_bwd_phase(const D2Dims& d, const DBTape& tp, const DBWork& wk, const BSmem& sm, float* scratch, bool active,
                                    int b, int part, int S, int t, int T, bool have_att_next, int KattT, int KdecT, AttAcc& acc,
                                    unsigned long long* prof, long long* prof_last) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int L = d.L, A = d.A, EF = d.EF;
  const int Lc = (L + S - 1) / S, l0 = min(L, part * Lc), l1 = min(L, l0 + Lc), nl = l1 - l0;
  const int Lh4 = (Lc + 30 + 3) & ~3, Lc4 = (Lc + 3) & ~3;
  float* q_s = scratch;
  float* wp = q_s + A;
  float* wc = wp + Lh4;
  float* fs = wc + Lh4;
  float* dctx = fs + (size_t)Lc * 36;
  float* w_s = dctx + EF;
  float* gw = w_s + Lc4;
  float* de = gw + Lc4;
  float* red = de + Lc4;
  float* xch = red + 64;
  float* dqw = xch + 8;
  float* ds = dqw + 8 * A;
  float* dfe = ds + (size_t)Lc * (A + 4);
  float* gsm = dfe + (size_t)(Lc + 30) * 36;
  const int ldd = A + 4;
  const uint32_t my_rank = tc::cluster_ctarank();
  const uint32_t base_rank = my_rank - (uint32_t)part;
  const float* xloc = tp.xloc ? tp.xloc + ((size_t)t * 128 + blockIdx.x) * Lc * A : nullptr;   // this CTA's slice of frame t

  if (active) {
    // ---- stage: q_t, wcat of this frame (w_{t-1}, cum_{t-1}) with halo, total d ctx_t, w_t slice
    const float* wprev = tp.w_hist + ((size_t)t * D2_BP + b) * L;
    const float* wcum = tp.wcum_hist + ((size_t)t * D2_BP + b) * L;
    const float* wcur = tp.w_hist + ((size_t)(t + 1) * D2_BP + b) * L;
#pragma unroll 1
    for (int i = tid; i < A; i += D2_THREADS) q_s[i] = tp.q[((size_t)t * D2_BP + b) * A + i];
#pragma unroll 1
    for (int i = tid; i < nl + 30; i += D2_THREADS) {
      const int l = l0 + i - 15;
      const bool in = (l >= 0 && l < L);
      wp[i] = in ? wprev[l] : 0.f;
      wc[i] = in ? wcum[l] : 0.f;
    }
#pragma unroll 1
    for (int c = tid; c < EF; c += D2_THREADS) {
      float v = ld_cg(wk.dprojin + ((size_t)t * D2_BP + b) * d.KprojP + d.R + c) + ld_cg(wk.dctxT_dec + (size_t)b * EF + c);
      if (have_att_next) v += ld_cg(wk.dctxT_att + (size_t)b * EF + c);
      dctx[c] = v;
      if (part == 0) wk.dctx_tot[((size_t)t * D2_BP + b) * EF + c] = v;
    }
#pragma unroll 1
    for (int i = tid; i < nl; i += D2_THREADS) w_s[i] = wcur[l0 + i];
    __syncthreads();

    // ---- location conv features of this slice: from the forward's tape when it has them, else recomputed (as the forward)
    if (tp.floc) {
      const float4* fsrc = reinterpret_cast<const float4*>(tp.floc + ((size_t)t * 128 + blockIdx.x) * Lc * D2_LOC_F);
#pragma unroll 1
      for (int i = tid; i < nl * (D2_LOC_F / 4); i += D2_THREADS)
        *reinterpret_cast<float4*>(fs + (size_t)(i >> 3) * 36 + (i & 7) * 4) = __ldcs(fsrc + i);
    } else {
      const int nchunk = (nl + 31) >> 5;
#pragma unroll 1
      for (int task = warp; task < nchunk * 4; task += D2_THREADS / 32) {
        const int li = (task >> 2) * 32 + lane, fg = task & 3;
        float a8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a8[i] = 0.f;
        const int lw = min(li, max(nl - 1, 0));
#pragma unroll 1
        for (int j = 0; j < D2_LOC_K; ++j) {
          const float a0 = wp[lw + j], a1 = wc[lw + j];
          const float4* w0 = reinterpret_cast<const float4*>(sm.conv_t + (0 * D2_LOC_K + j) * D2_LOC_F + fg * 8);
          const float4* w1 = reinterpret_cast<const float4*>(sm.conv_t + (1 * D2_LOC_K + j) * D2_LOC_F + fg * 8);
          const float4 p0 = w0[0], p1 = w0[1], r0 = w1[0], r1 = w1[1];
          a8[0] = fmaf(p0.x, a0, a8[0]); a8[1] = fmaf(p0.y, a0, a8[1]); a8[2] = fmaf(p0.z, a0, a8[2]); a8[3] = fmaf(p0.w, a0, a8[3]);
          a8[4] = fmaf(p1.x, a0, a8[4]); a8[5] = fmaf(p1.y, a0, a8[5]); a8[6] = fmaf(p1.z, a0, a8[6]); a8[7] = fmaf(p1.w, a0, a8[7]);
          a8[0] = fmaf(r0.x, a1, a8[0]); a8[1] = fmaf(r0.y, a1, a8[1]); a8[2] = fmaf(r0.z, a1, a8[2]); a8[3] = fmaf(r0.w, a1, a8[3]);
          a8[4] = fmaf(r1.x, a1, a8[4]); a8[5] = fmaf(r1.y, a1, a8[5]); a8[6] = fmaf(r1.z, a1, a8[6]); a8[7] = fmaf(r1.w, a1, a8[7]);
        }
        if (li < nl) {
          float4* dst = reinterpret_cast<float4*>(fs + (size_t)li * 36 + fg * 8);
          dst[0] = make_float4(a8[0], a8[1], a8[2], a8[3]);
          dst[1] = make_float4(a8[4], a8[5], a8[6], a8[7]);
        }
      }
    }

    ABW_MARK(3);
    // ---- g_w[l] = d ctx . memory[l] + d conv ch0 (from frame t+1) + g_cum;  g_cum += d conv ch1
    // warp per position, two positions in flight per warp (8 independent 16-byte loads per lane); the per-position scalars
    // (g_cum, d conv) are staged by all threads first so that no lane-0 load chain sits on the critical path
    {
      const float* mem = tp.memory + (size_t)b * L * EF;
      float* gc = wk.gcum + (size_t)b * L;
#pragma unroll 1
      for (int i = tid; i < nl; i += D2_THREADS) {
        const int l = l0 + i;
        float g = gc[l], s0 = 0.f;
        if (t + 1 < T) { g += wk.dconv[((size_t)1 * D2_BP + b) * L + l]; s0 = wk.dconv[((size_t)0 * D2_BP + b) * L + l]; }
        gc[l] = g;             // d/d cum_t (cum_t feeds cum_{t+1} and conv channel 1 of frame t+1)
        gw[i] = s0 + g;        // w_t also enters cum_t; the context term is added below
      }
      __syncthreads();
      const int nq = EF / 4;
#pragma unroll 1
      for (int li = 2 * warp; li < nl; li += 2 * (D2_THREADS / 32)) {
        const int li1 = min(li + 1, nl - 1);
        const float4* mp0 = reinterpret_cast<const float4*>(mem + (size_t)(l0 + li) * EF);
        const float4* mp1 = reinterpret_cast<const float4*>(mem + (size_t)(l0 + li1) * EF);
        float s0 = 0.f, s1 = 0.f;
#pragma unroll 1
        for (int c4 = lane; c4 < nq; c4 += 128) {
          float4 m0[4], m1[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int cc = min(c4 + 32 * u, nq - 1);
            m0[u] = __ldg(mp0 + cc); m1[u] = __ldg(mp1 + cc);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (c4 + 32 * u >= nq) continue;
            const float4 d4 = *reinterpret_cast<const float4*>(dctx + (c4 + 32 * u) * 4);
            s0 = fmaf(m0[u].x, d4.x, s0); s0 = fmaf(m0[u].y, d4.y, s0); s0 = fmaf(m0[u].z, d4.z, s0); s0 = fmaf(m0[u].w, d4.w, s0);
            s1 = fmaf(m1[u].x, d4.x, s1); s1 = fmaf(m1[u].y, d4.y, s1); s1 = fmaf(m1[u].z, d4.z, s1); s1 = fmaf(m1[u].w, d4.w, s1);
          }
        }
        s0 = warp_sum(s0); s1 = warp_sum(s1);
        if (lane == 0) { gw[li] += s0; if (li + 1 < nl) gw[li + 1] += s1; }
      }
    }
    __syncthreads();
    ABW_MARK(4);
    // ---- softmax backward: de = w * (g_w - sum_l w g_w); the dot needs all parts
    float pd = 0.f;
#pragma unroll 1
    for (int i = tid; i < nl; i += D2_THREADS) pd = fmaf(w_s[i], gw[i], pd);
    pd = block_reduce_sum(pd, red);
    if (tid == 0) xch[0] = pd;
  }
  tc::cluster_sync();
  if (active) {
    float tot = 0.f;
    {
      const uint32_t xa = tc::smem_u32(xch);
      for (int pr = 0; pr < S; ++pr) {
        float v;
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(tc::mapa(xa, base_rank + (uint32_t)pr)) : "memory");
        tot += v;
      }
    }
#pragma unroll 1
    for (int i = tid; i < nl; i += D2_THREADS) de[i] = w_s[i] * (gw[i] - tot);   // masked positions have w = 0 -> de = 0
    __syncthreads();

    // ---- recompute s = tanh(q + f Wd^T + pm); ds = de * v * (1 - s^2); accumulate dv, dq, d pm   (warp = 4 positions, lane = 4 dims)
    {
      const int a4 = lane * 4;
      const bool lane_on = a4 < A;
      const int ac = lane_on ? a4 : 0;
      const float4 q4 = *reinterpret_cast<const float4*>(q_s + ac);
      const float4 v4 = *reinterpret_cast<const float4*>(sm.v + ac);
      float dq4[4] = {0.f, 0.f, 0.f, 0.f};
      float dvl[4] = {acc.dv[0], acc.dv[1], acc.dv[2], acc.dv[3]};   // register copies (acc may alias the scratch for the compiler)
      const int ntile = (nl + 3) >> 2;
#pragma unroll 1
      for (int tl = warp; tl < ntile; tl += D2_THREADS / 32) {
        const int lt = tl * 4;
        float4 pm4[4];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          const int lp = min(lt + l, nl - 1);
          pm4[l] = xloc ? __ldcs(reinterpret_cast<const float4*>(xloc + (size_t)lp * A + ac))   // taped f Wd^T + pm (read once)
                        : __ldg(reinterpret_cast<const float4*>(tp.pmem + ((size_t)b * L + l0 + lp) * A + ac));
        }
        float ac4[4][4];
#pragma unroll
        for (int l = 0; l < 4; ++l)
#pragma unroll
          for (int a = 0; a < 4; ++a) ac4[l][a] = 0.f;
#pragma unroll 2
        for (int f4 = 0; f4 < (xloc ? 0 : D2_LOC_F / 4); ++f4) {
          float4 fv[4];
#pragma unroll
          for (int l = 0; l < 4; ++l) fv[l] = *reinterpret_cast<const float4*>(fs + (size_t)min(lt + l, nl - 1) * 36 + f4 * 4);
#pragma unroll
          for (int ff = 0; ff < 4; ++ff) {
            const float4 wv = *reinterpret_cast<const float4*>(sm.dense + (f4 * 4 + ff) * A + ac);
#pragma unroll
            for (int l = 0; l < 4; ++l) {
              const float fl = ff == 0 ? fv[l].x : (ff == 1 ? fv[l].y : (ff == 2 ? fv[l].z : fv[l].w));
              ac4[l][0] = fmaf(fl, wv.x, ac4[l][0]); ac4[l][1] = fmaf(fl, wv.y, ac4[l][1]);
              ac4[l][2] = fmaf(fl, wv.z, ac4[l][2]); ac4[l][3] = fmaf(fl, wv.w, ac4[l][3]);
            }
          }
        }
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          if (lt + l >= nl || !lane_on) continue;
          const float del = de[lt + l];
          const float s0 = tanh_fast(q4.x + ac4[l][0] + pm4[l].x), s1 = tanh_fast(q4.y + ac4[l][1] + pm4[l].y);
          const float s2 = tanh_fast(q4.z + ac4[l][2] + pm4[l].z), s3 = tanh_fast(q4.w + ac4[l][3] + pm4[l].w);
          dvl[0] = fmaf(del, s0, dvl[0]); dvl[1] = fmaf(del, s1, dvl[1]);
          dvl[2] = fmaf(del, s2, dvl[2]); dvl[3] = fmaf(del, s3, dvl[3]);
          const float4 d4 = make_float4(del * v4.x * (1.f - s0 * s0), del * v4.y * (1.f - s1 * s1), del * v4.z * (1.f - s2 * s2),
                                        del * v4.w * (1.f - s3 * s3));
          dq4[0] += d4.x; dq4[1] += d4.y; dq4[2] += d4.z; dq4[3] += d4.w;
          *reinterpret_cast<float4*>(ds + (size_t)(lt + l) * ldd + ac) = d4;
          float4* pmg = reinterpret_cast<float4*>(wk.dpm + ((size_t)b * L + l0 + lt + l) * A + ac);   // this thread owns (l, a) in every frame
          float4 o = *pmg;
          o.x += d4.x; o.y += d4.y; o.z += d4.z; o.w += d4.w;
          *pmg = o;
        }
      }
      acc.dv[0] = dvl[0]; acc.dv[1] = dvl[1]; acc.dv[2] = dvl[2]; acc.dv[3] = dvl[3];
      if (lane_on) *reinterpret_cast<float4*>(dqw + warp * A + ac) = make_float4(dq4[0], dq4[1], dq4[2], dq4[3]);
    }
    __syncthreads();
#pragma unroll 1
    for (int a = tid; a < A; a += D2_THREADS) {
      float s = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < D2_THREADS / 32; ++w8) s += dqw[w8 * A + a];
      wk.dq[(((size_t)t * S + part) * A + a) * D2_BP + b] = s;   // [t][part][a][b]: utterance-minor, the layout phase P stages
    }

    ABW_MARK(5);
    // ---- d f[l][f] = sum_a ds[l][a] Wd[a][f]   (warp = 16 positions; lane -> position quad lane>>3, filter quad lane&7)
    {
      const int fq = (lane & 7) * 4, lq = lane >> 3;
      const int ngrp = (nl + 15) >> 4;
#pragma unroll 1
      for (int g = warp; g < ngrp; g += D2_THREADS / 32) {
        const int lb = g * 16 + lq * 4;
        float o[4][4];
#pragma unroll
        for (int l = 0; l < 4; ++l)
#pragma unroll
          for (int f = 0; f < 4; ++f) o[l][f] = 0.f;
        // four attention dims per pass: 4 + 4 conflict-free 16-byte shared loads feed 64 FMAs (ds rows are 16-byte aligned:
        // ldd = A + 4; the weights come from the a-major copy of location_dense, row stride 36)
#pragma unroll 2
        for (int a = 0; a < A; a += 4) {
          float4 w4[4], x4[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) w4[i] = *reinterpret_cast<const float4*>(sm.dense_af + (size_t)(a + i) * 36 + fq);
#pragma unroll
          for (int l = 0; l < 4; ++l) x4[l] = *reinterpret_cast<const float4*>(ds + (size_t)min(lb + l, nl - 1) * ldd + a);
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            o[l][0] = fmaf(x4[l].x, w4[0].x, o[l][0]); o[l][1] = fmaf(x4[l].x, w4[0].y, o[l][1]);
            o[l][2] = fmaf(x4[l].x, w4[0].z, o[l][2]); o[l][3] = fmaf(x4[l].x, w4[0].w, o[l][3]);
            o[l][0] = fmaf(x4[l].y, w4[1].x, o[l][0]); o[l][1] = fmaf(x4[l].y, w4[1].y, o[l][1]);
            o[l][2] = fmaf(x4[l].y, w4[1].z, o[l][2]); o[l][3] = fmaf(x4[l].y, w4[1].w, o[l][3]);
            o[l][0] = fmaf(x4[l].z, w4[2].x, o[l][0]); o[l][1] = fmaf(x4[l].z, w4[2].y, o[l][1]);
            o[l][2] = fmaf(x4[l].z, w4[2].z, o[l][2]); o[l][3] = fmaf(x4[l].z, w4[2].w, o[l][3]);
            o[l][0] = fmaf(x4[l].w, w4[3].x, o[l][0]); o[l][1] = fmaf(x4[l].w, w4[3].y, o[l][1]);
            o[l][2] = fmaf(x4[l].w, w4[3].z, o[l][2]); o[l][3] = fmaf(x4[l].w, w4[3].w, o[l][3]);
          }
        }
#pragma unroll
        for (int l = 0; l < 4; ++l)
          if (lb + l < nl) *reinterpret_cast<float4*>(dfe + (size_t)(15 + lb + l) * 36 + fq) = make_float4(o[l][0], o[l][1], o[l][2], o[l][3]);
      }
    }

    // ---- d location_dense[a][f] += sum_l ds[l][a] f[l][f]   (thread tile: a-quad = lane, f-quad = warp)
    {
      const int a4 = lane * 4, f4 = warp * 4;
      if (a4 < A) {
        float wd[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) wd[i] = acc.wd[i];
#pragma unroll 2
        for (int l = 0; l < nl; ++l) {
          const float4 dv4 = *reinterpret_cast<const float4*>(ds + (size_t)l * ldd + a4);
          const float4 fv = *reinterpret_cast<const float4*>(fs + (size_t)l * 36 + f4);
          wd[0] = fmaf(dv4.x, fv.x, wd[0]);  wd[1] = fmaf(dv4.x, fv.y, wd[1]);
          wd[2] = fmaf(dv4.x, fv.z, wd[2]);  wd[3] = fmaf(dv4.x, fv.w, wd[3]);
          wd[4] = fmaf(dv4.y, fv.x, wd[4]);  wd[5] = fmaf(dv4.y, fv.y, wd[5]);
          wd[6] = fmaf(dv4.y, fv.z, wd[6]);  wd[7] = fmaf(dv4.y, fv.w, wd[7]);
          wd[8] = fmaf(dv4.z, fv.x, wd[8]);  wd[9] = fmaf(dv4.z, fv.y, wd[9]);
          wd[10] = fmaf(dv4.z, fv.z, wd[10]); wd[11] = fmaf(dv4.z, fv.w, wd[11]);
          wd[12] = fmaf(dv4.w, fv.x, wd[12]); wd[13] = fmaf(dv4.w, fv.y, wd[13]);
          wd[14] = fmaf(dv4.w, fv.z, wd[14]); wd[15] = fmaf(dv4.w, fv.w, wd[15]);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) acc.wd[i] = wd[i];
      }
    }
  }
  ABW_MARK(6);
  tc::cluster_sync();   // every part's d f slice is in its shared memory
  ABW_MARK(11);
  if (active) {
    // ---- halo rows of d f from the neighbouring parts (15 positions each side)
#pragma unroll 1
    for (int i = tid; i < 30 * 8; i += D2_THREADS) {
      const int e = i >> 3, f4 = (i & 7) * 4;
      const int row = e < 15 ? e : nl + e;                 // rows [0,15) and [15+nl, 30+nl)
      const int l = l0 - 15 + row;                          // global position of that row
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (l >= 0 && l < L) {
        const int pr = l / Lc, li = l - pr * Lc;
        const uint32_t ra = tc::mapa(tc::smem_u32(dfe + (size_t)(15 + li) * 36 + f4), base_rank + (uint32_t)pr);
        v = tc::ld_dsmem4(ra);
      }
      *reinterpret_cast<float4*>(dfe + (size_t)row * 36 + f4) = v;
    }
    __syncthreads();
    ABW_MARK(12);
    // ---- d location_conv[c][j][f] += sum_{l in slice} d f[l][f] * wcat[c][l + j - 15]
    {
      const int f4 = (tid & 7) * 4, cj0 = tid >> 3;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int cj = cj0 + 32 * h;
        if (cj >= 2 * D2_LOC_K) continue;
        const int c = cj / D2_LOC_K, j = cj - c * D2_LOC_K;
        const float* wsrc = (c ? wc : wp) + j;
        float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
#pragma unroll 4
        for (int l = 0; l < nl; ++l) {
          const float x = wsrc[l];
          const float4 fv = *reinterpret_cast<const float4*>(dfe + (size_t)(15 + l) * 36 + f4);
          o0 = fmaf(x, fv.x, o0); o1 = fmaf(x, fv.y, o1); o2 = fmaf(x, fv.z, o2); o3 = fmaf(x, fv.w, o3);
        }
        acc.wc[4 * h + 0] += o0; acc.wc[4 * h + 1] += o1; acc.wc[4 * h + 2] += o2; acc.wc[4 * h + 3] += o3;
      }
    }
    ABW_MARK(13);
    // ---- d wcat[c][l'] = sum_j sum_f d f[l' - j + 15][f] Wc[c][j][f]  for the slice, in two steps so that every d f row is
    // read from shared memory once instead of 31 times (the direct form was shared-memory-bandwidth bound):
    //   G[m][c][j] = d f[m] . Wc[c][j]   (thread = (row m incl. halo, channel c): row in registers, weights broadcast)
    //   d wcat[c][l'] = sum_j G[l' - j + 30][c][j]
    {
      const int nrow = nl + 30, c = tid >> 7, m = tid & 127, MR = Lc + 30;
      __syncthreads();   // gsm aliases nothing live, but every warp must be done with the d f halo writes above
#pragma unroll 1
      for (int m0 = 0; m0 < nrow; m0 += 128) {
        const int mm = m0 + m;
        if (mm < nrow) {
          float4 x[8];
#pragma unroll
          for (int f4 = 0; f4 < 8; ++f4) x[f4] = *reinterpret_cast<const float4*>(dfe + (size_t)mm * 36 + f4 * 4);
          float* gout = gsm + (size_t)c * D2_LOC_K * MR + mm;   // [c][j][m]: lanes = consecutive rows -> conflict-free
#pragma unroll 1
          for (int j = 0; j < D2_LOC_K; ++j) {
            const float4* wr = reinterpret_cast<const float4*>(sm.conv_t + (c * D2_LOC_K + j) * D2_LOC_F);
            float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
#pragma unroll
            for (int f4 = 0; f4 < 8; ++f4) {
              const float4 w = wr[f4];
              o0 = fmaf(x[f4].x, w.x, o0); o1 = fmaf(x[f4].y, w.y, o1); o2 = fmaf(x[f4].z, w.z, o2); o3 = fmaf(x[f4].w, w.w, o3);
            }
            gout[(size_t)j * MR] = (o0 + o1) + (o2 + o3);
          }
        }
      }
      __syncthreads();
#pragma unroll 1
      for (int i = tid; i < 2 * nl; i += D2_THREADS) {
        const int cc = i / nl, li = i - cc * nl;
        float o = 0.f;
#pragma unroll 1
        for (int j = 0; j < D2_LOC_K; ++j) o += gsm[((size_t)cc * D2_LOC_K + j) * MR + (li - j + 30)];
        wk.dconv[((size_t)cc * D2_BP + b) * L + l0 + li] = o;
      }
    }
  }
  ABW_MARK(7);
  // No cluster barrier here: the caller's grid barrier follows, and the scratch (the peers read this CTA's d f rows from it) is not
  // written again before every CTA has passed that barrier.
}

}  // namespace
