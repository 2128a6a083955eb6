// Persistent kernel + launch (included by bwd_decoder.cu).
namespace {

struct DBParams {
  D2Dims d;
  int T, KattT, KdecT, S;
  int wpol_att, wpol_dec;     // L2 eviction priority of the transposed-weight tiles of the two cells (1 evict-first, 0 default, 2 evict-last)
  int tc_att;                 // 1: attention backward contractions on tcgen05 (bwd_decoder_att_tc.inl), 0: CUDA cores (bwd_decoder_att.inl)
  DBWeights w;
  DBTape tp;
  DBWork wk;
  unsigned long long* prof;   // [16] or null: cycles of CTA 0 (13 G until the accumulator is complete, 0 G epilogue + barrier, 12 / 3..7 / 11 inside A,
                              // 1 A tail + barrier, 8..10 inside P, 2 P barrier)
};

// one 128-row tile of an input-gradient GEMM: dx[k0 + r][b] = sum_p WT[k0 + r][p] dg[b][p]
// Rows [c0, c1) of the product are the context block of the cell input: the attention backward reads them per utterance, so a tile that
// holds any of them is reduced with lanes over rows and those rows go to dxT [BP][c1 - c0] (utterance-major, 128-byte stores) instead of
// the k-major dx (whose column read costs one 32-byte sector per element: 1.2 us of the attention phase).
template <class Side, class Mark>
__device__ __forceinline__ bool dx_tile(tg::Pipe& pipe, const BSmem& sm, const CUtensorMap* wmap, const CUtensorMap* dgmap, int ktile,
                                        int npb, uint32_t rank, float* dx /*[K][BP]*/, float* dxT, int c0, int c1, Side side, int wpol, Mark mark) {
  tg::XSrc xs{dgmap, 0, dgmap, 0, npb};
  // the transposed weights of the decoder cell stream through L2 evict-first: the per-frame working set of the attention backward stays resident
  const bool ok = tg::gemm(pipe, wmap, ktile * TG_M, npb, xs, (int)rank, npb, D2_CLUSTER, 0u, side, wpol);
  mark();
  tg::accum_to_smem(pipe, 0u, sm.part);
  tc::cluster_sync();
  // CTA `rank` reduces rows [32*rank, 32*rank+32) of the tile over the four partials
  if (ktile * TG_M < c1 && (ktile + 1) * TG_M > c0) {
    // the sums of the context rows go back into this CTA's own rows of its partial (only this CTA reads them, each element by the thread
    // that rewrites it), then leave transposed from local shared memory.  (Reading the peers' partials with lanes over rows instead
    // cost 4.6 us per frame: four times the DSMEM packets.)
#pragma unroll 1
    for (int i = threadIdx.x; i < 32 * (TG_N / 2); i += D2_THREADS) {
      const int row = 32 * (int)rank + i / (TG_N / 2), col = 2 * (i % (TG_N / 2));
      const float2 s = tg::reduce4(sm.part, row, col);
      const int k = ktile * TG_M + row;
      if (k >= c0 && k < c1) *reinterpret_cast<float2*>(sm.part + row * TG_PART_LD + col) = s;
      else *reinterpret_cast<float2*>(dx + (size_t)k * D2_BP + col) = s;
    }
    __syncthreads();
#pragma unroll 1
    for (int i = threadIdx.x; i < 32 * TG_N; i += D2_THREADS) {
      const int row = 32 * (int)rank + (i & 31), col = i >> 5;
      const int k = ktile * TG_M + row;
      if (k >= c0 && k < c1) dxT[(size_t)col * (c1 - c0) + (k - c0)] = sm.part[row * TG_PART_LD + col];
    }
  } else {
#pragma unroll 1
    for (int i = threadIdx.x; i < 32 * (TG_N / 2); i += D2_THREADS) {
      const int row = 32 * (int)rank + i / (TG_N / 2), col = 2 * (i % (TG_N / 2));
      const float2 s = tg::reduce4(sm.part, row, col);
      const int k = ktile * TG_M + row;
      *reinterpret_cast<float2*>(dx + (size_t)k * D2_BP + col) = s;
    }
  }
  // (the caller separates two tiles of one cluster with a cluster barrier -- the peers must be done reading this CTA's partial
  //  before it is overwritten; after the last tile the grid barrier that follows covers it)
  return ok;
}

__global__ void __launch_bounds__(D2_THREADS, 1)
decoder_bwd_kernel(const __grid_constant__ DBMaps maps, const DBParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  const D2Dims& d = p.d;
  const BSmem sm = carve(base, d);
  float* scratch = reinterpret_cast<float*>(base);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t rank = tc::cluster_ctarank();
  const int cl = blockIdx.x / D2_CLUSTER;
  const int unit0 = cl * 32 + (int)rank * 8;
  const int pbase = cl * 128 + (int)rank * 32;
  const int B = d.B, H = d.H, T = p.T;
  const DBWork& wk = p.wk;
  const DBTape& tp = p.tp;

  if (tid == 0) {
    tc::prefetch_tmap(&maps.wt_att); tc::prefetch_tmap(&maps.wt_dec); tc::prefetch_tmap(&maps.dg_att); tc::prefetch_tmap(&maps.dg_dec);
  }
  if (p.tc_att) {
    // folded filter bank G[(c,j)][a] = sum_f Wd[a][f] Wc[c][j][f] (row c*32 + j; rows 31 and 63 zero) as bf16 (hi,lo) operand tiles, and
    // the 16 x 16 identity block of the d pm accumulation
#pragma unroll 1
    for (int i = tid; i < 64 * d.A; i += D2_THREADS) {
      const int cj = i / d.A, a_ = i - cj * d.A, c = cj >> 5, j = cj & 31;
      float g = 0.f;
      if (j < D2_LOC_K)
#pragma unroll 8
        for (int f = 0; f < D2_LOC_F; ++f) g = fmaf(__ldg(p.w.loc_dense + f * d.A + a_), __ldg(p.w.loc_conv_t + (c * D2_LOC_K + j) * D2_LOC_F + f), g);
      __nv_bfloat16 hi, lo;
      tc::split_bf16(g, hi, lo);
      const uint32_t off = (uint32_t)((a_ >> 6) * 8192) + tc::sw128_off(cj, a_ & 63);
      *reinterpret_cast<__nv_bfloat16*>(sm.gfT + off) = hi;
      *reinterpret_cast<__nv_bfloat16*>(sm.gfT + 16384 + off) = lo;
    }
#pragma unroll 1
    for (int i = tid; i < 16 * 64; i += D2_THREADS) {
      const int r = i >> 6, k = i & 63;
      *reinterpret_cast<__nv_bfloat16*>(sm.ident + tc::sw128_off(r, k)) = __float2bfloat16_rn(r == k ? 1.f : 0.f);
    }
#pragma unroll 1
    for (int i = tid; i < d.A * 8; i += D2_THREADS) sm.wq[i] = __ldg(p.w.Wq + (size_t)(i >> 3) * d.H + unit0 + (i & 7));
    tc::fence_proxy_async();
  } else {
#pragma unroll 1
    for (int i = tid; i < 2 * D2_LOC_K * D2_LOC_F; i += D2_THREADS) sm.conv_t[i] = p.w.loc_conv_t[i];
#pragma unroll 1
    for (int i = tid; i < d.A * D2_LOC_F; i += D2_THREADS) {
      const float v = p.w.loc_dense[i];           // [f][a]
      sm.dense[i] = v;
      sm.dense_af[(i % d.A) * 36 + i / d.A] = v;  // [a][f]
    }
  }
#pragma unroll 1
  for (int i = tid; i < d.A; i += D2_THREADS) sm.v[i] = p.w.v[i];
#pragma unroll 1
  for (int i = tid; i < 8 * D2_BP; i += D2_THREADS) { sm.dc_att[i] = 0.f; sm.dc_dec[i] = 0.f; }
  tg::Pipe pipe;
  // 512 tensor-memory columns: [0,256) GEMM accumulators (H of the attention phase aliases them), [256,384) d pm, [384,448) Y
  tg::pipe_init<512>(pipe, sm.bars, sm.stages, D2_STAGES, d.npass, sm.tmem_slot);
  if (tid == 0) {
    tc::mbar_init(sm.bars + 2 * D2_STAGES + 1, 1);
    tc::mbar_init(sm.bars + 2 * D2_STAGES + 2, 1);
    tc::fence_barrier_init();
  }
  tc::tc_fence_before();
  tc::cluster_sync();
  tc::tc_fence_after();
  const uint32_t tmem_p = pipe.tmem + 256u;   // d pm [0,128) and Y [128,192) of the tensor-core attention path: live for the whole kernel
  AttTcState stt{0u, false};

  GridBar bar{wk.barrier, wk.barrier + 1, gridDim.x, 0u};
  long long prof_last = clock64();
  auto prof_mark = [&](int slot) {
    if (blockIdx.x == 0 && tid == 0 && p.prof) { const long long now = clock64(); p.prof[slot] += (unsigned long long)(now - prof_last); prof_last = now; }
  };
  const int S = p.S;
  const int att_b = blockIdx.x / S, att_part = blockIdx.x % S;
  const bool att_active = att_b < B;
  const int npb = 4 * H / TG_BK;
  // the prenet block of the attention-cell input (first P rows of dx_att) is off the recurrent chain: its gradient is one
  // bulk GEMM after the loop, so the in-loop GEMM skips those tiles (vanilla: 20 + 12 = 32 tiles = one per cluster)
  const int att_tile0 = d.P / TG_M;
  const int ntile_dec = p.KdecT / TG_M, ntile_att = p.KattT / TG_M - att_tile0;
  const int nclusters = gridDim.x / D2_CLUSTER;
  AttAcc acc;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc.wd[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc.wc[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) acc.dv[i] = 0.f;
  float bias_att[4] = {0.f, 0.f, 0.f, 0.f}, bias_dec[4] = {0.f, 0.f, 0.f, 0.f};
  float* stg = scratch;                       // [64][33] staging of the pointwise phase: attention cell
  float* stg2 = scratch + 64 * 33 + 16;       // decoder cell
  float* qred = stg2 + 64 * 33 + 16;          // [8 warps][8 units][64] partial d att_h of the query layer
  float* wq_s = qred + 64 * D2_BP;            // [A][8] query weight columns when they are not resident (CUDA-core attention path)
  bool alive = true, fail = false;

  // Every first touch of the forward's tape is a DRAM miss (written once by the forward kernel).  The slices an iteration reads are pulled
  // into L2 ahead of use -- NOT beside the input-gradient GEMMs, whose weight stream owns the HBM bandwidth of that phase (prefetching
  // there moved 11 MB per frame into the one bandwidth-bound phase), but from the latency-bound phase before: the pointwise phase of
  // frame t+1 requests what the attention phase of frame t reads, the attention phase what the pointwise phase after it reads.
  auto pf = [&](const void* q, size_t bytes) {
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(q);
    const char* c = reinterpret_cast<const char*>(a0 & ~(uintptr_t)127);
    bytes += (size_t)(a0 & 127);
    for (size_t o = (size_t)tid * 128; o < bytes; o += (size_t)D2_THREADS * 128) tc::prefetch_l2(c + o);
  };
  auto prefetch_for_attention = [&](int t) {   // location term, conv features, query, weight rows with halo, the projection's share of d ctx
    if (t < 0 || !att_active) return;
    const int Lc = (d.L + S - 1) / S, l0 = min(d.L, att_part * Lc), nl = max(0, min(d.L, l0 + Lc) - l0);
    if (tp.xloc) pf(tp.xloc + ((size_t)t * 128 + blockIdx.x) * Lc * d.A, (size_t)nl * d.A * 4);
    if (tp.floc) pf(tp.floc + ((size_t)t * 128 + blockIdx.x) * Lc * D2_LOC_F, (size_t)nl * D2_LOC_F * 4);
    const int lo = max(0, l0 - 15), hi = min(d.L, l0 + nl + 15);
    if (hi > lo) {
      pf(tp.w_hist + ((size_t)t * D2_BP + att_b) * d.L + lo, (size_t)(hi - lo) * 4);
      pf(tp.wcum_hist + ((size_t)t * D2_BP + att_b) * d.L + lo, (size_t)(hi - lo) * 4);
      pf(tp.w_hist + ((size_t)(t + 1) * D2_BP + att_b) * d.L + lo, (size_t)(hi - lo) * 4);
    }
    pf(tp.q + ((size_t)t * D2_BP + att_b) * d.A, (size_t)d.A * 4);
    pf(wk.dprojin + ((size_t)t * D2_BP + att_b) * d.KprojP + d.R, (size_t)d.EF * 4);
  };
  auto prefetch_for_pointwise = [&](int t) {   // gates and cell states of this CTA's 8 units (attention cell t, decoder cell t-1), d rnn_h columns
#pragma unroll 1
    for (int g = 0; g < 5; ++g) {
      pf(tp.sav_att + (((size_t)t * 5 + g) * H + unit0) * D2_BP, (size_t)8 * D2_BP * 4);
      if (t > 0) pf(tp.sav_dec + (((size_t)(t - 1) * 5 + g) * H + unit0) * D2_BP, (size_t)8 * D2_BP * 4);
    }
    if (t > 0) pf(tp.sav_att + (((size_t)(t - 1) * 5 + 4) * H + unit0) * D2_BP, (size_t)8 * D2_BP * 4);
    if (t > 1) pf(tp.sav_dec + (((size_t)(t - 2) * 5 + 4) * H + unit0) * D2_BP, (size_t)8 * D2_BP * 4);
    if (t > 0 && tid < D2_BP) tc::prefetch_l2(wk.dprojin + ((size_t)(t - 1) * D2_BP + tid) * d.KprojP + unit0);
  };
  auto args_dec = [&](int t, bool have_next) {   // decoder LSTMCell of frame t
    PwArgs a{};
    a.sav = tp.sav_dec + (size_t)t * 5 * H * D2_BP;
    a.cprev = t > 0 ? tp.sav_dec + ((size_t)(t - 1) * 5 + 4) * H * D2_BP : nullptr;
    a.dh0 = have_next ? wk.dx_dec : nullptr; a.row0 = d.H + d.EF + d.C;     // d rnn_h_t from the decoder cell of frame t+1
    a.dh_rm = wk.dprojin + (size_t)t * D2_BP * d.KprojP; a.ld_rm = d.KprojP; a.col_rm = 0;   // from the mel / gate projection
    a.keep = tp.rnn_keep ? tp.rnn_keep + (size_t)t * B * H : nullptr;
    a.hidden_dropout = tp.hidden_dropout != 0; a.seed = tp.seed; a.rng_base = (2ull << 40) + (unsigned long long)t * B * H;
    a.dg = wk.dg_dec; a.dgT = wk.dgT_dec; a.ldT = (long long)T * D2_BP; a.tcol = (long long)t * D2_BP;
    return a;
  };
  auto args_att = [&](int t, bool have_next) {   // attention LSTMCell of frame t (its h also feeds the query and the decoder cell)
    PwArgs a{};
    a.sav = tp.sav_att + (size_t)t * 5 * H * D2_BP;
    a.cprev = t > 0 ? tp.sav_att + ((size_t)(t - 1) * 5 + 4) * H * D2_BP : nullptr;
    a.dh0 = wk.dx_dec; a.row0 = 0;                                         // att_h_t is the first input block of the decoder cell
    a.dh1 = have_next ? wk.dx_att : nullptr; a.row1 = d.P + d.EF;          // recurrent input of the attention cell of frame t+1
    a.keep = tp.att_keep ? tp.att_keep + (size_t)t * B * H : nullptr;
    a.hidden_dropout = tp.hidden_dropout != 0; a.seed = tp.seed; a.rng_base = (1ull << 40) + (unsigned long long)t * B * H;
    a.dg = wk.dg_att; a.dgT = wk.dgT_att; a.ldT = (long long)T * D2_BP; a.tcol = (long long)t * D2_BP;
    return a;
  };
  auto pw_dec = [&](int t, bool have_next) {
    const PwArgs a = args_dec(t, have_next);
    const PwIn in = lstm_pointwise_load(a, H, B, unit0);
    lstm_pointwise_bwd(a, in, H, B, unit0, pbase, sm.dc_dec, stg, nullptr, bias_dec);
  };
  // Phase P: attention cell of frame t and decoder cell of frame t-1.  Everything the phase reads is final when it starts (dx of phase
  // G, d query of phase A, the tape), so ALL loads -- both cells' tape / gradient rows and the d query -- are issued before the first
  // use: one L2 round trip for the phase instead of one per step.  d att_h through the query layer is split over the attention dims:
  // warp w multiplies dims [16w, 16w+16) for all 8 units of the CTA straight from registers (every d query element is read once per CTA,
  // not once per warp), the eight partials meet in shared memory.  Both cells share the two block barriers.
  auto pw_phase = [&](int t, bool have_next) {
    const PwArgs aa = args_att(t, have_next);
    const PwIn ia = lstm_pointwise_load(aa, H, B, unit0);
    PwArgs ad{};
    PwIn id{};
    if (t > 0) { ad = args_dec(t - 1, true); id = lstm_pointwise_load(ad, H, B, unit0); }
    prefetch_for_attention(t - 1);   // behind this phase's own loads
    const float* wq = sm.wq;
    if (!p.tc_att) {   // CUDA-core attention path: the resident slot holds its filter banks, the columns are restaged per frame
#pragma unroll 1
      for (int i = tid; i < d.A * 8; i += D2_THREADS) wq_s[i] = __ldg(p.w.Wq + (size_t)(i >> 3) * H + unit0 + (i & 7));
      __syncthreads();
      wq = wq_s;
    }
    {
      const int SP = p.tc_att ? 1 : S;             // the tensor-core attention path sums the parts itself (DSMEM), the other leaves S blocks
      const size_t pstride = (size_t)D2_BP * d.A;  // floats between parts; columns b >= B are stale, never used
      const float* src = wk.dq + (size_t)t * S * pstride + 2 * lane;
      float acc0[8], acc1[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) { acc0[u] = 0.f; acc1[u] = 0.f; }
#pragma unroll 1
      for (int ab = 16 * warp; ab < d.A; ab += 16 * (D2_THREADS / 32)) {
        float2 q[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) q[k] = __ldcg(reinterpret_cast<const float2*>(src + (size_t)(ab + k) * D2_BP));
        for (int pr = 1; pr < SP; ++pr) {
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const float2 w2 = __ldcg(reinterpret_cast<const float2*>(src + pr * pstride + (size_t)(ab + k) * D2_BP));
            q[k].x += w2.x; q[k].y += w2.y;
          }
        }
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float4 wa = *reinterpret_cast<const float4*>(wq + (ab + k) * 8), wb = *reinterpret_cast<const float4*>(wq + (ab + k) * 8 + 4);
          acc0[0] = fmaf(wa.x, q[k].x, acc0[0]); acc1[0] = fmaf(wa.x, q[k].y, acc1[0]);
          acc0[1] = fmaf(wa.y, q[k].x, acc0[1]); acc1[1] = fmaf(wa.y, q[k].y, acc1[1]);
          acc0[2] = fmaf(wa.z, q[k].x, acc0[2]); acc1[2] = fmaf(wa.z, q[k].y, acc1[2]);
          acc0[3] = fmaf(wa.w, q[k].x, acc0[3]); acc1[3] = fmaf(wa.w, q[k].y, acc1[3]);
          acc0[4] = fmaf(wb.x, q[k].x, acc0[4]); acc1[4] = fmaf(wb.x, q[k].y, acc1[4]);
          acc0[5] = fmaf(wb.y, q[k].x, acc0[5]); acc1[5] = fmaf(wb.y, q[k].y, acc1[5]);
          acc0[6] = fmaf(wb.z, q[k].x, acc0[6]); acc1[6] = fmaf(wb.z, q[k].y, acc1[6]);
          acc0[7] = fmaf(wb.w, q[k].x, acc0[7]); acc1[7] = fmaf(wb.w, q[k].y, acc1[7]);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) *reinterpret_cast<float2*>(qred + (warp * 8 + u) * D2_BP + 2 * lane) = make_float2(acc0[u], acc1[u]);
    }
    __syncthreads();
    prof_mark(8);
    float ex[2] = {0.f, 0.f};
#pragma unroll
    for (int w8 = 0; w8 < D2_THREADS / 32; ++w8) {
      const float2 r = *reinterpret_cast<const float2*>(qred + (w8 * 8 + warp) * D2_BP + 2 * lane);
      ex[0] += r.x; ex[1] += r.y;
    }
    if (2 * lane >= B) ex[0] = 0.f;        // utterance slots >= B hold stale scratch
    if (2 * lane + 1 >= B) ex[1] = 0.f;
    prof_mark(9);
    lstm_pointwise_compute(aa, ia, H, B, unit0, sm.dc_att, stg, ex, bias_att);
    if (t > 0) lstm_pointwise_compute(ad, id, H, B, unit0, sm.dc_dec, stg2, nullptr, bias_dec);
    __syncthreads();
    lstm_pointwise_pack(aa, H, pbase, stg);
    if (t > 0) lstm_pointwise_pack(ad, H, pbase, stg2);
    prof_mark(10);
    // (no block barrier here: the grid barrier that follows synchronises the block before the ring -- where the staging lies -- is refilled)
  };

  // prologue: decoder cell of the last frame
  pw_dec(T - 1, false);
  prefetch_for_attention(T - 1);
  alive = bar.sync(fail, true);
  prof_mark(2);

#pragma unroll 1
  for (int t = T - 1; t >= 0 && alive; --t) {
    // ---- G: dx_dec(t) and dx_att(t+1)
    const bool have_att_next = t + 1 <= T - 1;
    {
      const int n_dec = t >= 0 ? ntile_dec : 0, n_att = have_att_next ? ntile_att : 0;
#pragma unroll 1
      for (int tile = cl; tile < n_dec + n_att; tile += nclusters) {
        bool ok;
        if (tile != cl) tc::cluster_sync();   // second tile of this cluster in the same frame: the partial buffer is about to be rewritten
        auto no_side = []() {};
        if (tile < n_dec) ok = dx_tile(pipe, sm, &maps.wt_dec, &maps.dg_dec, tile, npb, rank, wk.dx_dec, wk.dctxT_dec, d.H, d.H + d.EF, no_side, p.wpol_dec, [&]() { prof_mark(13); });
        else ok = dx_tile(pipe, sm, &maps.wt_att, &maps.dg_att, tile - n_dec + att_tile0, npb, rank, wk.dx_att, wk.dctxT_att, d.P, d.P + d.EF, no_side, p.wpol_att, [&]() { prof_mark(13); });
        if (!ok) fail = true;
      }
    }
    if (!(alive = bar.sync(fail, false))) break;
    prof_mark(0);
    // ---- A: attention backward of frame t
    prefetch_for_pointwise(t);
    if (p.tc_att)
      attention_bwd_phase_tc(d, tp, wk, sm, base, att_active, att_b, att_part, S, t, T, have_att_next, pipe.tmem, tmem_p, acc.dv, stt, fail,
                             p.prof, &prof_last);
    else
      attention_bwd_phase(d, tp, wk, sm, scratch, att_active, att_b, att_part, S, t, T, have_att_next, p.KattT, p.KdecT, acc, p.prof, &prof_last);
    if (!(alive = bar.sync(fail, false))) break;
    prof_mark(1);
    // ---- P: attention cell of frame t, decoder cell of frame t-1
    pw_phase(t, have_att_next);
    if (!(alive = bar.sync(fail, true))) break;
    prof_mark(2);
  }

  // ---- per-CTA partials of the small attention weight gradients, bias gradients
  {
    float* sp = wk.small_partial + (size_t)blockIdx.x * (p.tc_att ? d.A * 64 + d.A : d.A * 32 + 2 * D2_LOC_K * D2_LOC_F + d.A);
    const int a4 = lane * 4, f4 = warp * 4;
    float* spc = sp + d.A * 32;
    if (p.tc_att) {
      // Y[a][(c,j)] (lanes = attention dims) -> this CTA's partial; d pm (lanes = positions) -> its rows of the gradient
      const bool has = att_active && stt.issued;
      const int lane0 = 32 * (warp & 3), half = warp >> 2, row = lane0 + lane;
      float v[32];
      tc::tc_fence_after();
      if (has) { tc::tmem_ld32(tmem_p + ((uint32_t)lane0 << 16) + 128u + 32u * half, v); tc::tmem_ld_wait(); }
      else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0.f;
      }
      if (row < d.A) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(sp + (size_t)row * 64 + 32 * half + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
      if (has) {
        const int Lc = (d.L + S - 1) / S, l0 = min(d.L, att_part * Lc), nl = min(d.L, l0 + Lc) - l0;
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          tc::tmem_ld32(tmem_p + ((uint32_t)lane0 << 16) + 64u * half + 32u * cc, v);
          tc::tmem_ld_wait();
          if (row < nl) {
            float4* dst = reinterpret_cast<float4*>(wk.dpm + ((size_t)att_b * d.L + l0 + row) * d.A + 64 * half + 32 * cc);
#pragma unroll
            for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          }
        }
      }
      tc::tc_fence_before();
    } else {
      if (a4 < d.A)
#pragma unroll
        for (int i = 0; i < 16; ++i) sp[(a4 + (i >> 2)) * 32 + f4 + (i & 3)] = acc.wd[i];            // [a][f] = location_dense.weight layout
      const int ff = (tid & 7) * 4, cj0 = tid >> 3;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int cj = cj0 + 32 * h;
        if (cj >= 2 * D2_LOC_K) continue;
        const int c = cj / D2_LOC_K, j = cj - c * D2_LOC_K;
#pragma unroll
        for (int i = 0; i < 4; ++i) spc[((ff + i) * 2 + c) * D2_LOC_K + j] = acc.wc[4 * h + i];       // [f][c][j] = location_conv.weight layout
      }
    }
    __syncthreads();
    float* red = scratch;   // [8][A]
    if (a4 < d.A) *reinterpret_cast<float4*>(red + warp * d.A + a4) = make_float4(acc.dv[0], acc.dv[1], acc.dv[2], acc.dv[3]);
    __syncthreads();
    float* spv = p.tc_att ? sp + d.A * 64 : spc + 2 * D2_LOC_K * D2_LOC_F;
#pragma unroll 1
    for (int a_ = tid; a_ < d.A; a_ += D2_THREADS) {
      float s = 0.f;
      for (int w8 = 0; w8 < D2_THREADS / 32; ++w8) s += red[w8 * d.A + a_];
      spv[a_] = s;
    }
    // bias gradients: this warp's unit, sum over the 64 utterances
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const float sa = warp_sum(bias_att[g]), sd = warp_sum(bias_dec[g]);
      if (lane == 0) {
        wk.dbias[g * H + unit0 + warp] = sa;
        wk.dbias[4 * H + g * H + unit0 + warp] = sd;
      }
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    wk.barrier[2] = alive ? 0u : 1u;
    if (!alive && wk.status) atomicExch(wk.status, 1u);
  }
  tc::tc_fence_before();
  tc::cluster_sync();
  if ((tid >> 5) == 1) {
    tc::tmem_dealloc<512>(pipe.tmem);
  }
}

}  // namespace

int db_parts(const D2Dims& d) {
  const int grid = (d.H / 32) * D2_CLUSTER;
  return (d.B * 4 <= grid) ? 4 : ((d.B * 2 <= grid) ? 2 : 1);
}
// The tensor-core attention backward needs the taped pre-tanh argument, att_dim == 128 (the M of the Y contraction) and slices of at
// most 128 positions (TMEM lanes).  T2B200_BWD_ATT=cuda forces the CUDA-core version (A/B runs, tests of both).
bool db_tc_att(const D2Dims& d, bool have_xloc) {
  const char* e = getenv("T2B200_BWD_ATT");
  if (e && e[0] == 'c') return false;
  const int S = db_parts(d), Lc = (d.L + S - 1) / S;
  return have_xloc && d.A == 128 && Lc <= 128 && d.EF % 4 == 0;
}
// per-CTA partial of the small attention gradients: tensor-core path Y[A][64] | dv[A]; CUDA-core path dWd[A][32] | dWc[32][2][31] | dv[A]
size_t db_small_floats(const D2Dims& d, bool tc_att) {
  return tc_att ? (size_t)d.A * 64 + d.A : (size_t)d.A * 32 + 2 * D2_LOC_K * D2_LOC_F + d.A;
}

static size_t att_bwd_scratch_floats_host(int L, int S, int A, int EF) {
  const int Lc = (L + S - 1) / S, Lh4 = (Lc + 30 + 3) & ~3, Lc4 = (Lc + 3) & ~3;
  return (size_t)A + 2 * Lh4 + (size_t)Lc * 36 + EF + 3 * Lc4 + 64 + 8 + 8 * (size_t)A + (size_t)Lc * (A + 4) + (size_t)(Lc + 30) * 36 + (size_t)(Lc + 30) * 64;
}

static size_t att_tc_scratch_bytes_host(int L, int S, int A, int EF) {
  const int Lc = (L + S - 1) / S, Lh4 = (Lc + 30 + 3) & ~3, Lc4 = (Lc + 3) & ~3;
  return 65536 + 32768 + 4 * ((size_t)A + 2 * Lh4 + EF + 3 * Lc4 + 64 + 8 + 8 * (size_t)A + 2 * Lh4);
}

int db_launch(const D2Dims& d, int T, int KattT, int KdecT, const DBWeights& w, const DBTape& tp, const DBWork& wk, const DBMaps& maps,
              unsigned long long* prof, cudaStream_t st) {
  T2_REQUIRE(d.B >= 1 && d.B <= D2_BP && T >= 1, T2_ERR_ARG, "decoder backward: batch %d not in [1,%d] per launch", d.B, D2_BP);
  T2_REQUIRE(d.H == d.R && d.H % 32 == 0 && d.P % TG_M == 0 && d.EF % 4 == 0 && d.A % 32 == 0 && d.A <= D2_MAX_A, T2_ERR_UNSUPPORTED,
             "decoder backward: unsupported dims H=%d R=%d EF=%d P=%d A=%d", d.H, d.R, d.EF, d.P, d.A);
  const int grid = (d.H / 32) * D2_CLUSTER;
  T2_REQUIRE(grid == 128, T2_ERR_UNSUPPORTED, "decoder backward: laid out for 128 CTAs (att_rnn_dim = 1024), got %d", grid);
  const int S = db_parts(d);
  const size_t ring = (size_t)D2_STAGES * TG_STAGE_BYTES;
  const bool tc_att = db_tc_att(d, tp.xloc != nullptr);
  T2_REQUIRE((tc_att ? att_tc_scratch_bytes_host(d.L, S, d.A, d.EF) : att_bwd_scratch_floats_host(d.L, S, d.A, d.EF) * 4) <= ring,
             T2_ERR_UNSUPPORTED, "decoder backward: L=%d too long for the attention scratch", d.L);
  T2_REQUIRE(!tc_att || 2 * D2_LOC_K * 128 * 4 <= TG_M * TG_PART_LD * 4, T2_ERR_UNSUPPORTED, "decoder backward: H staging");
  T2_REQUIRE((size_t)(2 * (64 * 33 + 16) + 64 * D2_BP + d.A * 8) * 4 <= ring, T2_ERR_UNSUPPORTED, "decoder backward: pointwise staging");
  const size_t smem = bsmem_bytes(d);
  T2_REQUIRE(smem <= 227 * 1024, T2_ERR_UNSUPPORTED, "decoder backward: %zu B shared memory", smem);
  T2_CHECK_CUDA(cudaFuncSetAttribute(decoder_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute at[1];
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(D2_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = D2_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int nclusters = 0;
  T2_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&nclusters, decoder_bwd_kernel, &cfg));
  T2_REQUIRE(nclusters * D2_CLUSTER >= grid, T2_ERR_UNSUPPORTED, "decoder backward: only %d clusters co-resident, %d needed", nclusters,
             grid / D2_CLUSTER);
  T2_CHECK_CUDA(cudaMemsetAsync(wk.barrier, 0, 4 * sizeof(unsigned), st));
  // the attention cell's transposed weights (25 MB) with default L2 priority, the decoder cell's (46 MB) evict-first: measured at C3
  // (att, dec) = (first, first) 43.2 ms per backward, (default, first) 42.8, (last, first) 42.8, (first, default) 43.0, (default, default) 45.5
  const int wpol_att = getenv("T2B200_BWD_WPOL_ATT") ? atoi(getenv("T2B200_BWD_WPOL_ATT")) : 0;
  const int wpol_dec = getenv("T2B200_BWD_WPOL_DEC") ? atoi(getenv("T2B200_BWD_WPOL_DEC")) : 1;
  DBParams p{d, T, KattT, KdecT, S, wpol_att, wpol_dec, tc_att ? 1 : 0, w, tp, wk, prof};
  T2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, decoder_bwd_kernel, maps, p));
  return T2_OK;
}
