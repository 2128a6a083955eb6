// Attention backward for one frame with its contractions on tcgen05 (included by bwd_decoder.cu; the CUDA-core version is
// bwd_decoder_att.inl).  Taken when the forward taped the pre-tanh argument (tp.xloc), att_dim == 128 and a slice has <= 128 positions.
// Forward being differentiated (model/attention.py:52-67, model/decoder.py:78-90):
//   f = conv1d([w_{t-1}; cum_{t-1}]) (2->32, k 31);  s = tanh(q + f Wd^T + pm);  e = s . v (masked);  w = softmax(e);  ctx = w^T memory.
// With ds[l][a] = d e[l] v[a] (1 - s[l][a]^2) (fp32, CUDA cores) everything downstream of ds is linear, and the two small layers fold
// into the 62-tap filter bank G[(c,j)][a] = sum_f Wd[a][f] Wc[f][c][j] the forward also uses (decoder_ar.cu):
//   d wcat[c][l']  = sum_j H[l' - j + 15][(c,j)],   H = ds G^T                       (per frame: M = positions, N = 64, K = 128)
//   d Wd[a][f]     = sum_{c,j} Y[a][(c,j)] Wc[f][c][j],  d Wc[f][c][j] = sum_a Wd[a][f] Y[a][(c,j)]
//   Y[a][(c,j)]    = sum_{t,l} ds_t[l][a] wcat_t[c][l + j - 15]                       (per frame: M = 128 dims, N = 64, K = positions;
//                                                                                      accumulated in tensor memory for the whole utterance)
//   d pm[l][a]     = sum_t ds_t[l][a]                                                 (identity contraction, accumulated in tensor memory)
// so neither the conv features f nor d f are formed.  ds is written once per frame as bf16 (hi,lo) K-major sw128 tiles [position][dim];
// H and d pm read them K-major, Y reads the same bytes MN-major (tc::smem_desc_sw128_mn).  Three bf16 passes (hi*hi + lo*hi + hi*lo).
namespace {

__device__ __forceinline__ uint32_t abw_pack2(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

struct AttTcState {
  uint32_t phase;   // parity of the two per-frame MMA barriers
  bool issued;      // this CTA has accumulated at least one frame into Y / d pm
};

// scratch bytes: ds [2 hl][2 kt][128][128 B] | toeplitz [2 hl][2 lt][64][128 B] | floats
__device__ __forceinline__ size_t att_tc_scratch_bytes(int Lc, int A, int EF) {
  const int Lh4 = (Lc + 30 + 3) & ~3, Lc4 = (Lc + 3) & ~3;
  return 65536 + 32768 + 4 * ((size_t)A + 2 * Lh4 + EF + 3 * Lc4 + 64 + 8 + 8 * (size_t)A + 2 * Lh4);
}

__device__ void attention_bwd_phase_tc(const D2Dims& d, const DBTape& tp, const DBWork& wk, const BSmem& sm, uint8_t* scratch, bool active,
                                       int b, int part, int S, int t, int T, bool have_att_next, uint32_t tmem_g, uint32_t tmem_p,
                                       float* dv_acc, AttTcState& stt, bool& fail, unsigned long long* prof, long long* prof_last) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int L = d.L, A = d.A, EF = d.EF;   // A == 128 here
  const int Lc = (L + S - 1) / S, l0 = min(L, part * Lc), l1 = min(L, l0 + Lc), nl = l1 - l0;
  const int Lh4 = (Lc + 30 + 3) & ~3, Lc4 = (Lc + 3) & ~3;
  const int Kl = (nl + 15) & ~15;           // positions rounded up to the MMA K step
  uint8_t* ds_t = scratch;
  uint8_t* toep = scratch + 65536;
  float* q_s = reinterpret_cast<float*>(toep + 32768);
  float* wp = q_s + A;
  float* wc = wp + Lh4;
  float* dctx = wc + Lh4;
  float* w_s = dctx + EF;
  float* gw = w_s + Lc4;
  float* de = gw + Lc4;
  float* red = de + Lc4;
  float* xch = red + 64;
  float* dqw = xch + 8;
  float* outp = dqw + 8 * A;                // [2][Lh4]: this slice's contributions to d wcat at positions l0 - 15 + x
  float* gsm = sm.part;                     // [2][31][128] staging of H for the diagonal sums (the split-K buffer is idle in this phase)
  const uint32_t my_rank = tc::cluster_ctarank();
  const uint32_t base_rank = my_rank - (uint32_t)part;
  const bool has = active && nl > 0;
  const float* xloc = tp.xloc + ((size_t)t * 128 + blockIdx.x) * Lc * A;
  const int a4 = lane * 4;
  const int ntile = (nl + 3) >> 2;
  uint64_t* bar_h = sm.bars + 2 * D2_STAGES + 1;
  uint64_t* bar_all = sm.bars + 2 * D2_STAGES + 2;

  float4 pre[3][4];   // taped pre-tanh arguments of this thread's first three position quads: requested behind the staging loads, used after the context step
  if (active) {
    // ---- stage: q_t, wcat of this frame (w_{t-1}, cum_{t-1}) with halo, total d ctx_t, w_t slice
    const float* wprev = tp.w_hist + ((size_t)t * D2_BP + b) * L;
    const float* wcum = tp.wcum_hist + ((size_t)t * D2_BP + b) * L;
    const float* wcur = tp.w_hist + ((size_t)(t + 1) * D2_BP + b) * L;
    // Every load of the staging step is issued before the first store (the loops it replaces paid one L2 round trip each: A == 128,
    // Lh4 <= 160 and nl <= 128 give one element per thread, d ctx two; wider encoder outputs take the trailing loop): the step costs one round
    // trip (~1.5 us under this load; storing the weight rows first and building the Toeplitz operand under the other loads was no faster).
    const float qv = tid < A ? tp.q[((size_t)t * D2_BP + b) * A + tid] : 0.f;
    const int lw = l0 + tid - 15;
    const bool inw = tid < nl + 30 && lw >= 0 && lw < L;
    const float wpv = inw ? wprev[lw] : 0.f, wcv = inw ? wcum[lw] : 0.f;
    float dcv[2] = {0.f, 0.f};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int c = tid + h * D2_THREADS;
      if (c < EF) {
        dcv[h] = ld_cg(wk.dprojin + ((size_t)t * D2_BP + b) * d.KprojP + d.R + c) + ld_cg(wk.dctxT_dec + (size_t)b * EF + c);
        if (have_att_next) dcv[h] += ld_cg(wk.dctxT_att + (size_t)b * EF + c);
      }
    }
    float* gc = wk.gcum + (size_t)b * L;
    float wsv = 0.f, g = 0.f, s0 = 0.f;
    if (tid < nl) {
      const int l = l0 + tid;
      wsv = wcur[l];
      g = gc[l];
      if (t + 1 < T) { g += wk.dconv[((size_t)1 * D2_BP + b) * L + l]; s0 = wk.dconv[((size_t)0 * D2_BP + b) * L + l]; }
    }
    if (has) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int lt = (warp + 8 * i) * 4;
#pragma unroll
        for (int l = 0; l < 4; ++l)
          pre[i][l] = (warp + 8 * i < ntile) ? __ldcs(reinterpret_cast<const float4*>(xloc + (size_t)min(lt + l, nl - 1) * A + a4))
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    if (tid < Lh4) { wp[tid] = wpv; wc[tid] = wcv; }
    if (tid < A) q_s[tid] = qv;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int c = tid + h * D2_THREADS;
      if (c < EF) {
        dctx[c] = dcv[h];
        if (part == 0) wk.dctx_tot[((size_t)t * D2_BP + b) * EF + c] = dcv[h];
      }
    }
    if (tid < nl) {
      w_s[tid] = wsv;
      gc[l0 + tid] = g;        // d/d cum_t (cum_t feeds cum_{t+1} and conv channel 1 of frame t+1)
      gw[tid] = s0 + g;        // w_t also enters cum_t; the context term is added below
    }
#pragma unroll 1
    for (int c = tid + 2 * D2_THREADS; c < EF; c += D2_THREADS) {
      float v = ld_cg(wk.dprojin + ((size_t)t * D2_BP + b) * d.KprojP + d.R + c) + ld_cg(wk.dctxT_dec + (size_t)b * EF + c);
      if (have_att_next) v += ld_cg(wk.dctxT_att + (size_t)b * EF + c);
      dctx[c] = v;
      if (part == 0) wk.dctx_tot[((size_t)t * D2_BP + b) * EF + c] = v;
    }
    __syncthreads();
    ABW_MARK(12);
    // ---- Toeplitz operand of Y: row (c,j) = c*32 + j holds wcat[c][l0 + li + j - 15] over li (K-major in li); rows 31 / 63 and li >= nl are zero
#pragma unroll 1
    for (int task = tid; task < 64 * (Kl >> 3); task += D2_THREADS) {
      const int cj = task & 63, q8 = task >> 6, c = cj >> 5, j = cj & 31;
      const float* src = (c ? wc : wp) + q8 * 8 + j;
      __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float x = (j < D2_LOC_K && q8 * 8 + e < nl) ? src[e] : 0.f;
        tc::split_bf16(x, hi[e], lo[e]);
      }
      const uint32_t off = (uint32_t)((q8 >> 3) * 8192) + tc::sw128_off(cj, (q8 & 7) * 8);
      *reinterpret_cast<uint4*>(toep + off) = *reinterpret_cast<const uint4*>(hi);
      *reinterpret_cast<uint4*>(toep + 16384 + off) = *reinterpret_cast<const uint4*>(lo);
    }
    ABW_MARK(3);
    // ---- g_w[l] += d ctx . memory[l]: warp per position, four positions in flight per warp (16 independent 16-byte loads per lane: the
    // phase is a chain of L2 round trips, one per group of positions -- two per group took 6-7 trips, 3.0 us)
    {
      const float* mem = tp.memory + (size_t)b * L * EF;
      const int nq = EF / 4;
#pragma unroll 1
      for (int li = 4 * warp; li < nl; li += 4 * (D2_THREADS / 32)) {
        const float4* mp[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) mp[r] = reinterpret_cast<const float4*>(mem + (size_t)(l0 + min(li + r, nl - 1)) * EF);
        float sacc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
        for (int c4 = lane; c4 < nq; c4 += 128) {
          float4 m[4][4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int cc = min(c4 + 32 * u, nq - 1);
#pragma unroll
            for (int r = 0; r < 4; ++r) m[r][u] = __ldg(mp[r] + cc);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (c4 + 32 * u >= nq) continue;
            const float4 d4 = *reinterpret_cast<const float4*>(dctx + (c4 + 32 * u) * 4);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              sacc[r] = fmaf(m[r][u].x, d4.x, sacc[r]); sacc[r] = fmaf(m[r][u].y, d4.y, sacc[r]);
              sacc[r] = fmaf(m[r][u].z, d4.z, sacc[r]); sacc[r] = fmaf(m[r][u].w, d4.w, sacc[r]);
            }
          }
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) sacc[r] = warp_sum(sacc[r]);
        if (lane == 0) {
#pragma unroll
          for (int r = 0; r < 4; ++r) if (li + r < nl) gw[li + r] += sacc[r];
        }
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
  // between the arrival and the wait of the cluster barrier (the peers' dot products): everything of the next step that does not need
  // the total -- tanh of the taped arguments in place, the zero rows of the ds tiles
  tc::cluster_arrive();
  if (has) {
    const float4 q4 = *reinterpret_cast<const float4*>(q_s + a4);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        pre[i][l].x = tanh_fast(q4.x + pre[i][l].x); pre[i][l].y = tanh_fast(q4.y + pre[i][l].y);
        pre[i][l].z = tanh_fast(q4.z + pre[i][l].z); pre[i][l].w = tanh_fast(q4.w + pre[i][l].w);
      }
  }
  if (active) {
    // rows [nl, Kl) of the ds tiles are part of the Y contraction: zero (the region is the TMA ring of the GEMM phase)
#pragma unroll 1
    for (int i = tid; i < (Kl - nl) * 4 * 8; i += D2_THREADS) {
      const int r = nl + (i >> 5), tile = (i >> 3) & 3, ch = i & 7;
      *reinterpret_cast<uint4*>(ds_t + tile * 16384 + ((r >> 3) << 10) + ((r & 7) << 7) + (ch << 4)) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  tc::cluster_wait();
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

    // ---- s = tanh(q + taped argument); ds = de * v * (1 - s^2) -> bf16 (hi,lo) operand tiles; dv, dq   (warp = 4 positions, lane = 4 dims)
    if (has) {
      const float4 q4 = *reinterpret_cast<const float4*>(q_s + a4);
      const float4 v4 = *reinterpret_cast<const float4*>(sm.v + a4);
      float dq4[4] = {0.f, 0.f, 0.f, 0.f};
      float dvl[4] = {dv_acc[0], dv_acc[1], dv_acc[2], dv_acc[3]};
      const uint32_t colb = (uint32_t)((lane >> 4) * 16384), sub = (uint32_t)((lane & 1) * 8), ch = (uint32_t)((lane & 15) >> 1);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int tl = warp + 8 * i;
        if (tl >= ntile) continue;
        const int lt = tl * 4;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          if (lt + l >= nl) continue;
          float s0, s1, s2, s3;
          if (i < 3) {
            const float4 sv = pre[i < 3 ? i : 0][l];   // tanh applied above
            s0 = sv.x; s1 = sv.y; s2 = sv.z; s3 = sv.w;
          } else {
            const float4 pm = __ldcs(reinterpret_cast<const float4*>(xloc + (size_t)(lt + l) * A + a4));
            s0 = tanh_fast(q4.x + pm.x); s1 = tanh_fast(q4.y + pm.y); s2 = tanh_fast(q4.z + pm.z); s3 = tanh_fast(q4.w + pm.w);
          }
          const float del = de[lt + l];
          dvl[0] = fmaf(del, s0, dvl[0]); dvl[1] = fmaf(del, s1, dvl[1]); dvl[2] = fmaf(del, s2, dvl[2]); dvl[3] = fmaf(del, s3, dvl[3]);
          const float d0 = del * v4.x * (1.f - s0 * s0), d1 = del * v4.y * (1.f - s1 * s1), d2 = del * v4.z * (1.f - s2 * s2),
                      d3 = del * v4.w * (1.f - s3 * s3);
          dq4[0] += d0; dq4[1] += d1; dq4[2] += d2; dq4[3] += d3;
          __nv_bfloat16 h0, h1, h2, h3, e0, e1, e2, e3;
          tc::split_bf16(d0, h0, e0); tc::split_bf16(d1, h1, e1); tc::split_bf16(d2, h2, e2); tc::split_bf16(d3, h3, e3);
          const int r = lt + l;
          const uint32_t off = colb + (uint32_t)(((r >> 3) << 10) + ((r & 7) << 7)) + ((ch ^ (uint32_t)(r & 7)) << 4) + sub;
          *reinterpret_cast<uint2*>(ds_t + off) = make_uint2(abw_pack2(h0, h1), abw_pack2(h2, h3));
          *reinterpret_cast<uint2*>(ds_t + 32768 + off) = make_uint2(abw_pack2(e0, e1), abw_pack2(e2, e3));
        }
      }
      dv_acc[0] = dvl[0]; dv_acc[1] = dvl[1]; dv_acc[2] = dvl[2]; dv_acc[3] = dvl[3];
      *reinterpret_cast<float4*>(dqw + warp * A + a4) = make_float4(dq4[0], dq4[1], dq4[2], dq4[3]);
    } else {
      *reinterpret_cast<float4*>(dqw + warp * A + a4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    tc::fence_proxy_async();   // the operand tiles were written with generic stores; the MMAs read them through the async proxy
    __syncthreads();
    ABW_MARK(5);
    // ---- the three contractions (one thread issues); H is read back below, Y and d pm stay in tensor memory until the kernel ends
    if (tid == 0 && has) {
      tc::tc_fence_after();
      const uint32_t dsb = tc::smem_u32(ds_t), gb = tc::smem_u32(sm.gfT), tb = tc::smem_u32(toep);
      const uint32_t i64 = tc::idesc_bf16_f32(128, 64), i16 = tc::idesc_bf16_f32(128, 16), iy = tc::idesc_bf16_f32_amn(128, 64);
      const uint64_t idn = tc::smem_desc_sw128(tc::smem_u32(sm.ident));
      const uint32_t first = stt.issued ? 1u : 0u;
#pragma unroll
      for (int k = 0; k < 8; ++k) {      // H[l][(c,j)] = sum_a ds[l][a] G[(c,j)][a]
        const uint32_t kt = (uint32_t)(k >> 2);
        const uint64_t adv = (uint64_t)((k & 3) * 2);
        const uint64_t ahi = tc::smem_desc_sw128(dsb + kt * 16384) + adv, alo = tc::smem_desc_sw128(dsb + 32768 + kt * 16384) + adv;
        const uint64_t ghi = tc::smem_desc_sw128(gb + kt * 8192) + adv, glo = tc::smem_desc_sw128(gb + 16384 + kt * 8192) + adv;
        tc::umma_bf16(tmem_g, alo, ghi, i64, k ? 1u : 0u);
        tc::umma_bf16(tmem_g, ahi, glo, i64, 1u);
        tc::umma_bf16(tmem_g, ahi, ghi, i64, 1u);
      }
      tc::umma_commit(bar_h);
#pragma unroll
      for (int k = 0; k < 8; ++k) {      // d pm[l][16k..16k+15] += ds[l][16k..16k+15] (identity block; hi + lo is the fp32 value to 2^-17)
        const uint32_t kt = (uint32_t)(k >> 2);
        const uint64_t adv = (uint64_t)((k & 3) * 2);
        const uint64_t ahi = tc::smem_desc_sw128(dsb + kt * 16384) + adv, alo = tc::smem_desc_sw128(dsb + 32768 + kt * 16384) + adv;
        tc::umma_bf16(tmem_p + 16u * k, ahi, idn, i16, first);
        tc::umma_bf16(tmem_p + 16u * k, alo, idn, i16, 1u);
      }
#pragma unroll 1
      for (int s16 = 0; s16 < (Kl >> 4); ++s16) {   // Y[a][(c,j)] += sum_l ds[l][a] wcat[c][l + j - 15]: ds read MN-major
        const uint64_t ahi = tc::smem_desc_sw128_mn(dsb + (uint32_t)s16 * 2048, 16384, 1024);
        const uint64_t alo = tc::smem_desc_sw128_mn(dsb + 32768 + (uint32_t)s16 * 2048, 16384, 1024);
        const uint32_t lt = (uint32_t)(s16 >> 2);
        const uint64_t adv = (uint64_t)((s16 & 3) * 2);
        const uint64_t thi = tc::smem_desc_sw128(tb + lt * 8192) + adv, tlo = tc::smem_desc_sw128(tb + 16384 + lt * 8192) + adv;
        tc::umma_bf16(tmem_p + 128u, alo, thi, iy, s16 ? 1u : first);
        tc::umma_bf16(tmem_p + 128u, ahi, tlo, iy, 1u);
        tc::umma_bf16(tmem_p + 128u, ahi, thi, iy, 1u);
      }
      tc::umma_commit(bar_all);
    }
    if (has) stt.issued = true;
    // d query of this slice: the eight warps' sums, left in row 0 (each thread reads and writes its own column); part 0 adds the parts
    // after the cluster barrier below
#pragma unroll 1
    for (int a = tid; a < A; a += D2_THREADS) {
      float s = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < D2_THREADS / 32; ++w8) s += dqw[w8 * A + a];
      dqw[a] = s;
    }
    // ---- H -> shared memory, [c][j][position] (lanes = consecutive positions: conflict-free)
    if (has) {
      if (!tc::mbar_wait(bar_h, stt.phase)) fail = true;
      tc::tc_fence_after();
      const int lane0 = 32 * (warp & 3), c = warp >> 2, pz = lane0 + lane;
      float v[32];
      tc::tmem_ld32(tmem_g + ((uint32_t)lane0 << 16) + 32u * c, v);
      tc::tmem_ld_wait();
      if (pz < nl) {
#pragma unroll
        for (int j = 0; j < D2_LOC_K; ++j) gsm[(c * D2_LOC_K + j) * 128 + pz] = v[j];
      }
      tc::tc_fence_before();
    }
    __syncthreads();
    ABW_MARK(6);
    // ---- this slice's contributions to d wcat[c][l0 - 15 + x] = sum_j H[x - j][(c,j)]
#pragma unroll 1
    for (int i = tid; i < 2 * Lh4; i += D2_THREADS) {
      const int c = i / Lh4, x = i - c * Lh4;
      float o = 0.f;
      if (x < nl + 30) {
        const int j0 = max(0, x - nl + 1), j1 = min(D2_LOC_K - 1, x);
#pragma unroll 1
        for (int j = j0; j <= j1; ++j) o += gsm[(c * D2_LOC_K + j) * 128 + x - j];
      }
      outp[i] = o;
    }
  }
  tc::cluster_sync();   // every part's contributions are in its shared memory
  ABW_MARK(11);
  if (active) {
#pragma unroll 1
    for (int i = tid; i < 2 * nl; i += D2_THREADS) {
      const int c = i / nl, li = i - c * nl, l = l0 + li;
      float o = 0.f;
      for (int pr = 0; pr < S; ++pr) {
        const int pl0 = min(L, pr * Lc), pnl = min(L, pl0 + Lc) - pl0, x = l - pl0 + 15;
        if (pnl <= 0 || x < 0 || x >= pnl + 30) continue;
        float v;
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(tc::mapa(tc::smem_u32(outp + c * Lh4 + x), base_rank + (uint32_t)pr)) : "memory");
        o += v;
      }
      wk.dconv[((size_t)c * D2_BP + b) * L + l] = o;
    }
    // d query of the utterance = sum over the parts, [t][a][b] utterance-minor (the layout phase P reads); the upper threads, beside
    // the d wcat rows above
    if (part == 0) {
#pragma unroll 1
      for (int a = D2_THREADS - 1 - tid; a < A; a += D2_THREADS) {
        float s = dqw[a];
        for (int pr = 1; pr < S; ++pr) {
          float v;
          asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(tc::mapa(tc::smem_u32(dqw + a), base_rank + (uint32_t)pr)) : "memory");
          s += v;
        }
        wk.dq[((size_t)t * S * A + a) * D2_BP + b] = s;
      }
    }
    if (has) {   // all MMAs of this frame have retired: the operand tiles may be overwritten (the next phase streams through them)
      if (!tc::mbar_wait(bar_all, stt.phase)) fail = true;
      stt.phase ^= 1u;
    }
  }
  ABW_MARK(7);
  // No cluster barrier here: the caller's grid barrier follows, and the scratch (the peers read this CTA's `outp` from it) is not
  // written again before every CTA has passed that barrier.
}

}  // namespace
