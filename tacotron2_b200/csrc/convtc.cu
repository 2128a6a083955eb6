// tcgen05 implicit-GEMM conv1d (see convtc.cuh).  Persistent: one CTA per SM walks the (batch item, 128-position tile, N tile) list;
// warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..9 = epilogue (TMEM -> registers -> BN/activation -> global).  Two accumulators in
// tensor memory: the epilogue of tile i runs while the producer and the issuer are already on tile i + 1 (the one-tile-per-CTA form
// left the tensor pipe idle during every prologue and epilogue: 44 % active on the postnet layers, profiles/r02_v1_tc_kernels_summary.txt).
#include "convtc.cuh"
#include "layers.cuh"

#define CT_STAGES 3
#define CT_A_TILE (128 * 64 * 2)
#define CT_STAGE_BYTES (4 * CT_A_TILE)  // A hi|lo (32 KB) + B hi|lo (up to 32 KB)
#define CT_TMEM_COLS 256                // two accumulators of up to 128 columns
#define CT_THREADS 320

namespace {

// What one 16-channel chunk of the epilogue adds from global memory, requested ahead of its use.
struct CtPre { float4 r[4], ac[4]; uint4 kp; };

__device__ __forceinline__ void ct_load_pre(const ConvTcArgs& a, int b, int t, int n, bool keep16, CtPre& p) {
  if (t >= a.T) return;
  const size_t pos = (size_t)b * a.T + t;
  if (a.residual) {
    const float4* src = reinterpret_cast<const float4*>(a.residual + (size_t)b * a.res_bstride + (size_t)t * a.Cout + n);
#pragma unroll
    for (int i = 0; i < 4; ++i) p.r[i] = __ldcs(src + i);   // read once: keep it from evicting the per-channel tables and the weights
  }
  if (a.accum) {
    const float4* src = reinterpret_cast<const float4*>(a.accum + pos * a.Cout + n);
#pragma unroll
    for (int i = 0; i < 4; ++i) p.ac[i] = __ldcs(src + i);   // (may alias out_f32: each element is read, then written, by this thread only)
  }
  if (a.keep) {
    const unsigned char* kp = a.keep + pos * a.keep_ld + n;
    if (keep16) p.kp = __ldg(reinterpret_cast<const uint4*>(kp));
    else {
      unsigned char* dstb = reinterpret_cast<unsigned char*>(&p.kp);
#pragma unroll
      for (int i = 0; i < 16; ++i) dstb[i] = __ldg(kp + i);
    }
  }
}

// bias / folded BatchNorm, activation, dropout, residual, receptive-field sum, then the fp32 and bf16 (hi,lo) stores of one chunk
// tab: the per-channel shift (and scale) of all Cout columns in shared memory ([Cout] shift, then [Cout] scale), or null: from global
// memory.  (ncu on the narrow HiFi-GAN layers: 27 % of all stall samples sat on the sixteen bias loads of a chunk -- the activations
// streaming through L1 kept evicting the 128-byte bias vector, one L2 round trip per chunk.)
#define CT_TAB_MAX 2048
__device__ __forceinline__ void ct_fill_tab(const ConvTcArgs& a, float* tab) {
  for (int n = threadIdx.x; n < a.Cout; n += blockDim.x) {
    const int ci = a.chan_mod > 0 ? n % a.chan_mod : n;
    tab[n] = a.shift ? __ldg(a.shift + ci) : 0.f;
    tab[a.Cout + n] = a.scale ? __ldg(a.scale + ci) : 1.f;
  }
}

__device__ __forceinline__ void ct_finish_chunk(const ConvTcArgs& a, float (&v)[16], const CtPre& cur, size_t pos, int n, long long plane,
                                                const float* tab) {
  const float* rr = reinterpret_cast<const float*>(cur.r);
  const float* aa = reinterpret_cast<const float*>(cur.ac);
  const unsigned char* kk = reinterpret_cast<const unsigned char*>(&cur.kp);
  float sh[16], sc[16];
  if (tab) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      *reinterpret_cast<float4*>(sh + 4 * i) = *reinterpret_cast<const float4*>(tab + n + 4 * i);
      if (a.scale) *reinterpret_cast<float4*>(sc + 4 * i) = *reinterpret_cast<const float4*>(tab + a.Cout + n + 4 * i);
    }
  }
  // every option is tested once per chunk, not once per element: the executed path is short straight-line code (with the tests inside
  // the element loop the narrow layers spent a third of their stall samples on instruction fetch and branch resolution)
  if (tab) {
    if (a.scale) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(v[i], sc[i], sh[i]);
    } else if (a.shift) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] += sh[i];
    }
  } else if (a.shift) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int ci = a.chan_mod > 0 ? (n + i) % a.chan_mod : n + i;
      v[i] = a.scale ? fmaf(v[i], __ldg(a.scale + ci), __ldg(a.shift + ci)) : v[i] + __ldg(a.shift + ci);
    }
  }
  if (a.act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (a.act == ACT_TANH) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = tanhf(v[i]);
  } else if (a.act == ACT_GELU) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = 0.5f * v[i] * (1.0f + erff(v[i] * 0.70710678118654752f));
  }
  if (a.keep) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = kk[i] ? v[i] * a.drop_scale : 0.f;
  }
  if (a.residual) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += rr[i];
  }
  if (a.out_scale != 0.f) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] *= a.out_scale;
  }
  if (a.accum) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += aa[i];
  }
  if (a.out_c0 && n == 0) a.out_c0[pos] = v[0];
  if (a.out_f32) {
    float4* dst = reinterpret_cast<float4*>(a.out_f32 + pos * a.Cout + n);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  if (a.out_b16) {
    __align__(16) __nv_bfloat16 hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) tc::split_bf16((a.b16_slope != 0.f && v[i] < 0.f) ? v[i] * a.b16_slope : v[i], hi[i], lo[i]);
    uint4* dh = reinterpret_cast<uint4*>(a.out_b16 + pos * a.Cout + n);
    uint4* dl = reinterpret_cast<uint4*>(a.out_b16 + plane + pos * a.Cout + n);
    dh[0] = reinterpret_cast<const uint4*>(hi)[0]; dh[1] = reinterpret_cast<const uint4*>(hi)[1];
    dl[0] = reinterpret_cast<const uint4*>(lo)[0]; dl[1] = reinterpret_cast<const uint4*>(lo)[1];
  }
}

__global__ void __launch_bounds__(CT_THREADS, 1) conv_tc_kernel(const __grid_constant__ ConvTcArgs a, int ntiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(base + CT_STAGES * CT_STAGE_BYTES);
  uint64_t* empty = full + CT_STAGES;
  uint64_t* acc_full = empty + CT_STAGES;    // [2]
  uint64_t* acc_empty = acc_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  volatile int* fail_s = reinterpret_cast<volatile int*>(tmem_slot + 1);
  float* tab = a.Cout <= CT_TAB_MAX ? reinterpret_cast<float*>(base + CT_STAGES * CT_STAGE_BYTES + 128) : nullptr;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_t = (a.T + 127) >> 7;
  if (tab) ct_fill_tab(a, tab);
  const int dil = a.dil > 1 ? a.dil : 1;
  const int cblocks = (a.Cin + 63) >> 6, nkb = a.k * cblocks, pad = ((a.k - 1) >> 1) * dil;
  const uint32_t b_tile = (uint32_t)a.NT * 128u;  // bytes of one bf16 weight tile (NT rows x 128 B)

  if (tid == 0) {
    tc::prefetch_tmap(&a.in_map);
    tc::prefetch_tmap(&a.w_map);
    for (int i = 0; i < CT_STAGES; ++i) { tc::mbar_init(full + i, 1); tc::mbar_init(empty + i, 1); }
    for (int i = 0; i < 2; ++i) { tc::mbar_init(acc_full + i, 1); tc::mbar_init(acc_empty + i, 8); }
    *fail_s = 0;
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<CT_TMEM_COLS>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // tile id -> (position tile, N tile): consecutive ids share the activation tile
  const int ntn = a.Cout / a.NT;
  if (warp == 0) {
    if (lane == 0) {
      asm volatile("fence.proxy.async;" ::: "memory");
      int st = 0; uint32_t ph = 0;
      const uint32_t tx = (uint32_t)(a.npass == 3 ? 2 : 1) * (CT_A_TILE + b_tile);
      for (int tile = blockIdx.x; tile < ntiles && !*fail_s; tile += gridDim.x) {
        const int bt = tile / ntn, n0 = (tile - bt * ntn) * a.NT;
        const int b = bt / tiles_t, t0 = (bt - b * tiles_t) * 128;
        for (int kb = 0; kb < nkb; ++kb) {
          const int j = kb / cblocks, c0 = (kb - j * cblocks) * 64;
          if (!tc::mbar_wait(empty + st, ph ^ 1)) { *fail_s = 1; break; }
          uint8_t* sp = base + (size_t)st * CT_STAGE_BYTES;
          tc::mbar_arrive_expect_tx(full + st, tx);
          tc::tma_load_4d(sp, &a.in_map, full + st, c0, t0 + j * dil - pad, b, 0);    // zero fill outside [0,T) = conv padding
          tc::tma_load_4d(sp + 2 * CT_A_TILE, &a.w_map, full + st, c0, j, n0, 0);
          if (++st == CT_STAGES) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int st = 0; uint32_t ph = 0;
      const uint32_t idesc = tc::idesc_bf16_f32(128, a.NT);
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles && !*fail_s; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        if (!tc::mbar_wait(acc_empty + buf, (uint32_t)((it >> 1) & 1) ^ 1u)) { *fail_s = 1; break; }   // the epilogue has drained this accumulator
        tc::tc_fence_after();
        const uint32_t dcol = tmem + (uint32_t)buf * 128u;
        uint32_t acc = 0;
        for (int kb = 0; kb < nkb; ++kb) {
          if (!tc::mbar_wait(full + st, ph)) { *fail_s = 1; break; }
          tc::tc_fence_after();
          const uint32_t sb = tc::smem_u32(base + (size_t)st * CT_STAGE_BYTES);
          const uint64_t ahi = tc::smem_desc_sw128(sb), alo = tc::smem_desc_sw128(sb + CT_A_TILE);
          const uint64_t bhi = tc::smem_desc_sw128(sb + 2 * CT_A_TILE), blo = tc::smem_desc_sw128(sb + 2 * CT_A_TILE + b_tile);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adv = (uint64_t)(k * 2);
            if (a.npass == 3) {
              tc::umma_bf16(dcol, alo + adv, bhi + adv, idesc, acc);
              tc::umma_bf16(dcol, ahi + adv, blo + adv, idesc, 1u);
              tc::umma_bf16(dcol, ahi + adv, bhi + adv, idesc, 1u);
            } else {
              tc::umma_bf16(dcol, ahi + adv, bhi + adv, idesc, acc);
            }
            acc = 1u;
          }
          tc::umma_commit(empty + st);
          if (++st == CT_STAGES) { st = 0; ph ^= 1; }
        }
        tc::umma_commit(acc_full + buf);
      }
    }
  } else {
    // ---- epilogue warps 2..9: thread = output position (TMEM lane quarter = warp % 4), 16-channel chunks alternate between the two groups.
    // What a chunk adds from global memory (residual, running sum, dropout mask: first-touch DRAM reads in the narrow HiFi-GAN stages)
    // is requested one chunk AHEAD -- the next chunk of this tile or the first of the next tile -- so its latency runs under the
    // accumulator wait, the tensor-memory read and the stores of the current chunk (it used to sit between them: 5.8 us per tile).
    const int q = warp & 3, grp = (warp - 2) >> 2, row = 32 * q + lane;
    const long long plane = (long long)a.B * a.T * a.Cout;
    const int nch = a.NT / 16;
    const bool keep16 = (a.keep_ld & 15) == 0 && (reinterpret_cast<uintptr_t>(a.keep) & 15) == 0;
    auto load_pre = [&](int tile, int c, CtPre& p) {
      const int bt = tile / ntn, n = (tile - bt * ntn) * a.NT + c * 16;
      const int b = bt / tiles_t, t = (bt - b * tiles_t) * 128 + row;
      ct_load_pre(a, b, t, n, keep16, p);
    };
    CtPre cur, nxt;
    if ((int)blockIdx.x < ntiles && grp < nch) load_pre(blockIdx.x, grp, cur);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      const int bt = tile / ntn, n0 = (tile - bt * ntn) * a.NT;
      const int b = bt / tiles_t, t0 = (bt - b * tiles_t) * 128;
      const bool got = tc::mbar_wait(acc_full + buf, (uint32_t)((it >> 1) & 1));
      if (!got) *fail_s = 1;
      tc::tc_fence_after();
      const int t = t0 + row;
      const bool valid = got && t < a.T && !*fail_s;
      const size_t pos = (size_t)b * a.T + min(t, a.T - 1);
#pragma unroll 1
      for (int c = grp; c < nch; c += 2) {
        {
          int ntile = tile, nc = c + 2;
          if (nc >= nch) { ntile = tile + gridDim.x; nc = grp; }
          if (ntile < ntiles) load_pre(ntile, nc, nxt);
        }
        float v[16];
        tc::tmem_ld16(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * 128 + c * 16), v);
        tc::tmem_ld_wait();
        if (valid) ct_finish_chunk(a, v, cur, pos, n0 + c * 16, plane, tab);
        cur = nxt;
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(acc_empty + buf);   // this warp has read its lanes of the accumulator
      if (!got) break;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<CT_TMEM_COLS>(tmem);
}

// ---------------------------------------------------------------------------------------------------------------------------------
// Narrow layers (HiFi-GAN stages of 64 / 32 channels, conv_post).  conv_tc_kernel fetches the [128 positions x 64 channels] activation
// tile once per TAP: k = 11 reads every position eleven times from L2, and at 32-64 channels that stream -- not the tensor pipe, not HBM
// -- is what the layer costs (0.37 us per K-block on every SM).  Here a CTA loads 256 positions plus the halo ((k-1)*dil <= 64 rows) ONCE
// into a K-major SWIZZLE_128B buffer of 320 rows and tap j multiplies the view that starts j*dil rows further down: the shared-memory
// descriptor's start address moves by whole 128-byte rows, the swizzle is a function of the address bits, so the shifted view is
// read exactly as TMA wrote it.  Weights (one [NT x 64] tile per tap) stream through a small ring and serve both 128-position halves.
#define CH_ROWS 320
#define CH_PLANE (CH_ROWS * 128)          // one plane (hi or lo) of the haloed tile
#define CH_ABUF (2 * CH_PLANE)
#define CH_WSTAGE 16384                   // one tap's weights: hi | lo, NT <= 64 rows of 128 B
#define CH_WSTAGES 3
#define CH_G 2                            // 128-position sub-tiles per tile
#ifndef CH_NG
#define CH_NG 2                           // epilogue warp groups (4 warps each: one per tensor-memory lane quarter); 3 and 4 measured slower
                                          // (20.3 / 21.7 ms against 19.2 for the HiFi-GAN generator: 128 / 96 registers per thread spill)
#endif
#define CH_THREADS (64 + CH_NG * 128)

__global__ void __launch_bounds__(CH_THREADS, 1) conv_halo_kernel(const __grid_constant__ ConvTcArgs a, int ntiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* wring = base + 2 * CH_ABUF;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(wring + CH_WSTAGES * CH_WSTAGE);   // [2]
  uint64_t* a_empty = a_full + 2;            // [2]
  uint64_t* w_full = a_empty + 2;            // [CH_WSTAGES]
  uint64_t* w_empty = w_full + CH_WSTAGES;
  uint64_t* acc_full = w_empty + CH_WSTAGES; // [2]
  uint64_t* acc_empty = acc_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  volatile int* fail_s = reinterpret_cast<volatile int*>(tmem_slot + 1);
  float* tab = reinterpret_cast<float*>(wring + CH_WSTAGES * CH_WSTAGE + 256);   // [2][Cout <= 64]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_t = (a.T + 128 * CH_G - 1) / (128 * CH_G);
  ct_fill_tab(a, tab);
  const int dil = a.dil > 1 ? a.dil : 1, pad = ((a.k - 1) >> 1) * dil;
  const int planes = a.npass == 3 ? 2 : 1;
  const uint32_t b_tile = (uint32_t)a.NT * 128u;

  if (tid == 0) {
    tc::prefetch_tmap(&a.in_map_h);
    tc::prefetch_tmap(&a.in_map_t);
    tc::prefetch_tmap(&a.w_map);
    for (int i = 0; i < 2; ++i) { tc::mbar_init(a_full + i, 1); tc::mbar_init(a_empty + i, 1); tc::mbar_init(acc_full + i, 1); tc::mbar_init(acc_empty + i, 4 * CH_NG); }
    for (int i = 0; i < CH_WSTAGES; ++i) { tc::mbar_init(w_full + i, 1); tc::mbar_init(w_empty + i, 1); }
    *fail_s = 0;
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<CT_TMEM_COLS>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      asm volatile("fence.proxy.async;" ::: "memory");
      int ws = 0; uint32_t wph = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles && !*fail_s; tile += gridDim.x, ++it) {
        const int b = tile / tiles_t, t0 = (tile - b * tiles_t) * 128 * CH_G;
        const int ab = it & 1;
        if (!tc::mbar_wait(a_empty + ab, (uint32_t)((it >> 1) & 1) ^ 1u)) { *fail_s = 1; break; }
        uint8_t* ap = base + (size_t)ab * CH_ABUF;
        tc::mbar_arrive_expect_tx(a_full + ab, (uint32_t)planes * (uint32_t)CH_PLANE);
        for (int pl = 0; pl < planes; ++pl) {   // rows t0 - pad ... of the input; outside [0, T) TMA fills zeros = the conv padding
          tc::tma_load_4d(ap + pl * CH_PLANE, &a.in_map_h, a_full + ab, 0, t0 - pad, b, pl);
          tc::tma_load_4d(ap + pl * CH_PLANE + 16384, &a.in_map_h, a_full + ab, 0, t0 - pad + 128, b, pl);
          tc::tma_load_4d(ap + pl * CH_PLANE + 32768, &a.in_map_t, a_full + ab, 0, t0 - pad + 256, b, pl);
        }
        for (int j = 0; j < a.k; ++j) {
          if (!tc::mbar_wait(w_empty + ws, wph ^ 1)) { *fail_s = 1; break; }
          tc::mbar_arrive_expect_tx(w_full + ws, (uint32_t)planes * b_tile);
          tc::tma_load_4d(wring + (size_t)ws * CH_WSTAGE, &a.w_map, w_full + ws, 0, j, 0, 0);
          if (++ws == CH_WSTAGES) { ws = 0; wph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int ws = 0; uint32_t wph = 0;
      const uint32_t idesc = tc::idesc_bf16_f32(128, a.NT);
      const int ksteps = (a.Cin + 15) >> 4;   // channels beyond Cin are zero fill: no MMA for them
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles && !*fail_s; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        const uint32_t par = (uint32_t)((it >> 1) & 1);
        if (!tc::mbar_wait(acc_empty + buf, par ^ 1u)) { *fail_s = 1; break; }   // the epilogue has drained this accumulator
        if (!tc::mbar_wait(a_full + buf, par)) { *fail_s = 1; break; }
        tc::tc_fence_after();
        const uint32_t ab = tc::smem_u32(base + (size_t)buf * CH_ABUF);
        for (int j = 0; j < a.k; ++j) {
          if (!tc::mbar_wait(w_full + ws, wph)) { *fail_s = 1; break; }
          tc::tc_fence_after();
          const uint32_t wb = tc::smem_u32(wring + (size_t)ws * CH_WSTAGE);
          const uint64_t bhi = tc::smem_desc_sw128(wb), blo = tc::smem_desc_sw128(wb + b_tile);
#pragma unroll 1
          for (int g = 0; g < CH_G; ++g) {
            const uint32_t arow = ab + (uint32_t)(g * 128 + j * dil) * 128u;
            const uint64_t ahi = tc::smem_desc_sw128(arow), alo = tc::smem_desc_sw128(arow + CH_PLANE);
            const uint32_t dcol = tmem + (uint32_t)(buf * 128 + g * a.NT);
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t adv = (uint64_t)(k * 2);
              const uint32_t acc = (j || k) ? 1u : 0u;
              if (a.npass == 3) {
                tc::umma_bf16(dcol, alo + adv, bhi + adv, idesc, acc);
                tc::umma_bf16(dcol, ahi + adv, blo + adv, idesc, 1u);
                tc::umma_bf16(dcol, ahi + adv, bhi + adv, idesc, 1u);
              } else {
                tc::umma_bf16(dcol, ahi + adv, bhi + adv, idesc, acc);
              }
            }
          }
          tc::umma_commit(w_empty + ws);
          if (++ws == CH_WSTAGES) { ws = 0; wph ^= 1; }
        }
        tc::umma_commit(acc_full + buf);
        tc::umma_commit(a_empty + buf);
      }
    }
  } else {
    // ---- epilogue warps 2.. : CH_NG groups of four warps (thread = output position of a 128-position sub-tile, tensor-memory lane quarter
    // = warp % 4).  The items of a tile -- (sub-tile g, 16-channel chunk c), e = g * nch + c -- go round-robin to the groups; the global
    // reads of a warp's next item are requested before its current one is finished.  The epilogue, not the tensor pipe, bounds these
    // layers (8 bytes of output and ~20 instructions per element).
    const int q = warp & 3, grp = (warp - 2) >> 2, row = 32 * q + lane;
    const long long plane = (long long)a.B * a.T * a.Cout;
    const int nch = a.NT / 16, nitem = CH_G * nch;
    const bool keep16 = (a.keep_ld & 15) == 0 && (reinterpret_cast<uintptr_t>(a.keep) & 15) == 0;
    auto load_pre = [&](int tile, int e, CtPre& p) {
      const int g = e / nch, c = e - g * nch;
      const int b = tile / tiles_t, t = (tile - b * tiles_t) * 128 * CH_G + g * 128 + row;
      ct_load_pre(a, b, t, c * 16, keep16, p);
    };
    CtPre cur, nx1;
    int ptile = blockIdx.x, pe = grp;   // the item whose reads are requested next
    auto advance = [&]() {
      pe += CH_NG;
      if (pe >= nitem) { pe = grp; ptile += gridDim.x; }
    };
    if (grp < nitem) {
      if (ptile < ntiles) load_pre(ptile, pe, cur);
      advance();
    }
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      const int b = tile / tiles_t, t0 = (tile - b * tiles_t) * 128 * CH_G;
      const bool got = tc::mbar_wait(acc_full + buf, (uint32_t)((it >> 1) & 1));
      if (!got) *fail_s = 1;
      tc::tc_fence_after();
      const bool live = got && !*fail_s;
#pragma unroll 1
      for (int e = grp; e < nitem; e += CH_NG) {
        const int g = e / nch, c = e - g * nch;
        const int t = t0 + g * 128 + row;
        const size_t pos = (size_t)b * a.T + min(t, a.T - 1);
        if (ptile < ntiles) load_pre(ptile, pe, nx1);
        advance();
        float v[16];
        tc::tmem_ld16(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * 128 + g * a.NT + c * 16), v);
        tc::tmem_ld_wait();
        if (live && t < a.T) ct_finish_chunk(a, v, cur, pos, c * 16, plane, tab);
        cur = nx1;
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(acc_empty + buf);   // this warp has read its lanes of the accumulators
      if (!got) break;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<CT_TMEM_COLS>(tmem);
}

__global__ void split_cl_kernel(const float* src, long long src_bstride, __nv_bfloat16* dst, int B, int T, int C) {
  const long long n = (long long)B * T * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int b = (int)(p / T), t = (int)(p % T);
    __nv_bfloat16 hi, lo;
    tc::split_bf16(src[(size_t)b * src_bstride + (size_t)t * C + c], hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}

__global__ void pack_conv_w_kernel(const float* W, __nv_bfloat16* dst, int Cout, int Cin, int k) {
  const long long n = (long long)Cout * Cin * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int j = (int)((i / Cin) % k);
    const int co = (int)(i / ((long long)Cin * k));
    __nv_bfloat16 hi, lo;
    tc::split_bf16(W[((size_t)co * Cin + ci) * k + j], hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}

__global__ void embed_split_kernel(const long long* idx, const float* table, __nv_bfloat16* dst, int n_pos, int E, int n_rows) {
  const long long n = (long long)n_pos * E;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = idx[i / E];
    const float v = (r >= 0 && r < n_rows) ? table[(size_t)r * E + i % E] : 0.f;
    __nv_bfloat16 hi, lo;
    tc::split_bf16(v, hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}

// four channels per thread (C % 4 == 0, 16-byte aligned): 16-byte loads, 8-byte stores, one division per four elements
__global__ void split_cl4_kernel(const float* src, long long src_bstride, __nv_bfloat16* dst, int B, int T, int C) {
  const int C4 = C >> 2;
  const long long n4 = (long long)B * T * C4, n = n4 * 4;
  for (long long i4 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i4 < n4; i4 += (long long)gridDim.x * blockDim.x) {
    const long long p = i4 / C4;
    const int c = (int)(i4 - p * C4) * 4;
    const int b = (int)(p / T), t = (int)(p - (long long)b * T);
    const float4 v = *reinterpret_cast<const float4*>(src + (size_t)b * src_bstride + (size_t)t * C + c);
    __align__(8) __nv_bfloat16 hi[4], lo[4];
    tc::split_bf16(v.x, hi[0], lo[0]); tc::split_bf16(v.y, hi[1], lo[1]); tc::split_bf16(v.z, hi[2], lo[2]); tc::split_bf16(v.w, hi[3], lo[3]);
    *reinterpret_cast<uint2*>(dst + i4 * 4) = *reinterpret_cast<const uint2*>(hi);
    *reinterpret_cast<uint2*>(dst + n + i4 * 4) = *reinterpret_cast<const uint2*>(lo);
  }
}

inline int gs(long long n) { long long g = (n + 255) / 256; return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g)); }

}  // namespace

int convtc_launch(const ConvTcArgs& a, cudaStream_t st) {
  T2_REQUIRE(a.NT % 16 == 0 && a.NT >= 16 && a.NT <= 128 && a.Cout % a.NT == 0 && (a.k & 1) && a.Cin % 8 == 0 && a.Cout % 8 == 0,
             T2_ERR_UNSUPPORTED, "convtc: unsupported shape Cin=%d Cout=%d k=%d NT=%d", a.Cin, a.Cout, a.k, a.NT);
  if (a.B * a.T == 0) return T2_OK;
  const size_t smem = 1024 + (size_t)CT_STAGES * CT_STAGE_BYTES + 128 + 2 * CT_TAB_MAX * sizeof(float);
  static bool attr_set = false;
  static int num_sms = 0;
  if (!attr_set) {
    T2_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0;
    T2_CHECK_CUDA(cudaGetDevice(&dev));
    T2_CHECK_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    attr_set = true;
  }
  if (a.halo) {
    T2_REQUIRE(convtc_halo_ok(a.Cin, a.Cout, a.NT, a.k, a.dil), T2_ERR_ARG, "convtc: layer does not qualify for the halo kernel");
    const size_t hsmem = 1024 + 2 * (size_t)CH_ABUF + (size_t)CH_WSTAGES * CH_WSTAGE + 256 + 2 * 64 * sizeof(float);
    static bool hattr_set = false;
    if (!hattr_set) {
      T2_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));
      hattr_set = true;
    }
    const int htiles = a.B * ((a.T + 128 * CH_G - 1) / (128 * CH_G));
    conv_halo_kernel<<<std::min(htiles, num_sms), CH_THREADS, hsmem, st>>>(a, htiles);
    T2_CHECK_CUDA(cudaGetLastError());
    return T2_OK;
  }
  const int ntiles = a.B * ((a.T + 127) / 128) * (a.Cout / a.NT);
  conv_tc_kernel<<<std::min(ntiles, num_sms), CT_THREADS, smem, st>>>(a, ntiles);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

bool convtc_halo_ok(int Cin, int Cout, int NT, int k, int dil) {
  const char* e = getenv("T2B200_CONV_HALO");
  if (e && e[0] == '0') return false;
  const int d = dil > 1 ? dil : 1;
  return Cin <= 64 && Cin % 8 == 0 && Cout == NT && NT <= 64 && NT % 16 == 0 && (k & 1) && k >= 3 && (k - 1) * d <= CH_ROWS - 128 * CH_G;
}

int convtc_make_halo_maps(ConvTcArgs& a, const __nv_bfloat16* x, int B, int T, int Cin) {
  unsigned long long dims[4] = {(unsigned long long)Cin, (unsigned long long)T, (unsigned long long)B, 2ull};
  unsigned long long st[3] = {(unsigned long long)Cin * 2, (unsigned long long)T * Cin * 2, (unsigned long long)B * T * Cin * 2};
  unsigned boxh[4] = {64u, 128u, 1u, 1u}, boxt[4] = {64u, 64u, 1u, 1u};
  const int rc = tc_make_tmap_bf16(&a.in_map_h, x, 4, dims, st, boxh);
  if (rc != T2_OK) return rc;
  return tc_make_tmap_bf16(&a.in_map_t, x, 4, dims, st, boxt);
}

int split_cl_launch(const float* src, long long src_bstride, __nv_bfloat16* dst, int B, int T, int C, cudaStream_t st) {
  if (B * T == 0) return T2_OK;
  if (C % 4 == 0 && src_bstride % 4 == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0 &&
      ((long long)B * T * C) % 8 == 0)
    split_cl4_kernel<<<gs((long long)B * T * C / 4), 256, 0, st>>>(src, src_bstride, dst, B, T, C);
  else
    split_cl_kernel<<<gs((long long)B * T * C), 256, 0, st>>>(src, src_bstride, dst, B, T, C);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int pack_conv_w_launch(const float* W, __nv_bfloat16* dst, int Cout, int Cin, int k, cudaStream_t st) {
  pack_conv_w_kernel<<<gs((long long)Cout * Cin * k), 256, 0, st>>>(W, dst, Cout, Cin, k);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int embed_split_launch(const long long* idx, const float* table, __nv_bfloat16* dst, int n_pos, int E, int n_rows, cudaStream_t st) {
  embed_split_kernel<<<gs((long long)n_pos * E), 256, 0, st>>>(idx, table, dst, n_pos, E, n_rows);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
