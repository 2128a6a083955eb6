// Device helpers shared by the persistent decoder kernels (decoder.cu: teacher-forced / training / single-step schedule;
// decoder_ar.cu: pipelined autoregressive schedule): CUDA-core dense layers of the output stage, the LSTM pointwise tail with its
// split-K reduce over the cluster, the query-projection partials.  Included inside each translation unit's anonymous namespace
// user; D2_SYNC() is the CTA-wide barrier of the calling kernel's 256 compute threads (default __syncthreads()).
#pragma once
#include "decoder.cuh"

#ifndef D2_SYNC
#define D2_SYNC() __syncthreads()
#endif
#ifndef D2_LDW
#define D2_LDW(p) __ldg(p)   // weight-row load of the dense layers (decoder_ar.cu: without L1 allocation)
#endif

namespace {

// ------------------------------------------------------------------------------------------
// Small dense layers of the output phase on CUDA cores: y_j[r] = sum_k W[r][k] x_j[k] for the two utterances (j = 0,1) a
// cluster owns.  One warp per block of consecutive rows [rb, re), four rows in flight, lanes stride the row in 16-byte loads
// (coalesced, weights straight from L2), fixed-order warp reduction; lane 0 leaves the sums in y[j*ys + r - ybase] (shared
// memory).  One out-of-line copy serves all three layers (instruction-cache footprint).
// XG: the two input rows are read straight from global memory (L2-coherent loads, they were written by other CTAs before the
// last grid barrier) in the same batch as the weight loads, instead of being staged in shared memory first (one L2 round trip
// and a barrier less on the dependent chain).
template <bool XG>
__device__ __noinline__ void rows_dot2(const float* __restrict__ W, int ldw, int K4, int rb, int re, const float* x0, const float* x1, float* y,
                                       int ybase, int ys) {
  const int lane = threadIdx.x & 31;
#pragma unroll 1
  for (int r = rb; r < re; r += 4) {
    float a0[4] = {0.f, 0.f, 0.f, 0.f}, a1[4] = {0.f, 0.f, 0.f, 0.f};
    const float4* w0 = reinterpret_cast<const float4*>(W + (size_t)min(r + 0, re - 1) * ldw);
    const float4* w1 = reinterpret_cast<const float4*>(W + (size_t)min(r + 1, re - 1) * ldw);
    const float4* w2 = reinterpret_cast<const float4*>(W + (size_t)min(r + 2, re - 1) * ldw);
    const float4* w3 = reinterpret_cast<const float4*>(W + (size_t)min(r + 3, re - 1) * ldw);
#pragma unroll 4
    for (int k = lane; k < K4; k += 32) {
      const float4 xa = XG ? __ldcg(reinterpret_cast<const float4*>(x0) + k) : *reinterpret_cast<const float4*>(x0 + 4 * k);
      const float4 xb = XG ? __ldcg(reinterpret_cast<const float4*>(x1) + k) : *reinterpret_cast<const float4*>(x1 + 4 * k);
      const float4 wv[4] = {D2_LDW(w0 + k), D2_LDW(w1 + k), D2_LDW(w2 + k), D2_LDW(w3 + k)};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a0[i] = fmaf(wv[i].x, xa.x, a0[i]); a0[i] = fmaf(wv[i].y, xa.y, a0[i]); a0[i] = fmaf(wv[i].z, xa.z, a0[i]); a0[i] = fmaf(wv[i].w, xa.w, a0[i]);
        a1[i] = fmaf(wv[i].x, xb.x, a1[i]); a1[i] = fmaf(wv[i].y, xb.y, a1[i]); a1[i] = fmaf(wv[i].z, xb.z, a1[i]); a1[i] = fmaf(wv[i].w, xb.w, a1[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float y0 = warp_sum(a0[i]), y1 = warp_sum(a1[i]);
      if (lane == 0 && r + i < re) { y[r + i - ybase] = y0; y[ys + r + i - ybase] = y1; }
    }
  }
  __syncwarp();
}

// Short-row variant (K4 <= 32 * NK float4 per row): NR rows and all their loads in flight at once, so a layer costs one or two
// L2 round trips instead of one per group of four rows.
template <int NR, int NK, bool WS>   // WS: the weight rows are in shared memory
__device__ __noinline__ void rows_short(const float* __restrict__ W, int ldw, int K4, int rb, int re, const float* x0, const float* x1, float* y,
                                        int ybase, int ys) {
  const int lane = threadIdx.x & 31;
  float4 xa[NK], xb[NK];
#pragma unroll
  for (int k = 0; k < NK; ++k) {
    const int kk = lane + 32 * k;
    xa[k] = kk < K4 ? *reinterpret_cast<const float4*>(x0 + 4 * kk) : make_float4(0.f, 0.f, 0.f, 0.f);
    xb[k] = kk < K4 ? *reinterpret_cast<const float4*>(x1 + 4 * kk) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll 1
  for (int r = rb; r < re; r += NR) {
    float4 wv[NR][NK];
#pragma unroll
    for (int i = 0; i < NR; ++i)
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        const int kk = lane + 32 * k;
        const float4* wp = reinterpret_cast<const float4*>(W + (ptrdiff_t)min(r + i, re - 1) * ldw) + kk;
        wv[i][k] = kk < K4 ? (WS ? *wp : D2_LDW(wp)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        a0 = fmaf(wv[i][k].x, xa[k].x, a0); a0 = fmaf(wv[i][k].y, xa[k].y, a0); a0 = fmaf(wv[i][k].z, xa[k].z, a0); a0 = fmaf(wv[i][k].w, xa[k].w, a0);
        a1 = fmaf(wv[i][k].x, xb[k].x, a1); a1 = fmaf(wv[i][k].y, xb[k].y, a1); a1 = fmaf(wv[i][k].z, xb[k].z, a1); a1 = fmaf(wv[i][k].w, xb[k].w, a1);
      }
      const float y0 = warp_sum(a0), y1 = warp_sum(a1);
      if (lane == 0 && r + i < re) { y[r + i - ybase] = y0; y[ys + r + i - ybase] = y1; }
    }
  }
  __syncwarp();
}

// value -> the same word of all CTAs of the cluster, each copy counted on the receiver's transaction barrier `bar`
__device__ __forceinline__ void st_async_all(float* p, float v, uint64_t* bar) {
  const uint32_t a = tc::smem_u32(p), b = tc::smem_u32(bar);
#pragma unroll
  for (uint32_t r = 0; r < D2_CLUSTER; ++r) tc::st_async_f32(tc::mapa(a, r), v, tc::mapa(b, r));
}

// ------------------------------------------------------------------------------------------
// LSTM phase tail: split-K reduce over the cluster + gate nonlinearities for this CTA's 8 units.
// Tile rows are packed so that rows [32*rank, 32*rank+32) = gates (i,f,g,o) x 8 units of CTA `rank`.
struct HOut {
  __nv_bfloat16* x0; int ld0, col0;   // bf16 (hi,lo) destination: element (b, col0+u) of [2][BP][ld0]; lo plane at +plane0
  long long plane0;
  __nv_bfloat16* x1; int ld1, col1;
  long long plane1;
  float* rm; int rm_ld, rm_col;       // fp32 row-major destination [b][rm_ld]
};

__device__ __forceinline__ void store_split(__nv_bfloat16* base, long long plane, int ld, int b, int col, float h) {
  __nv_bfloat16 hi, lo;
  tc::split_bf16(h, hi, lo);
  base[(size_t)b * ld + col] = hi;
  base[plane + (size_t)b * ld + col] = lo;
}

__device__ __forceinline__ uint32_t pack2_bf16(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

// One plane (hi or lo) of 8 consecutive hidden values -> one 16-byte store when the destination allows it.
__device__ __forceinline__ void store_plane8(__nv_bfloat16* dst, const float* h8, int lo_plane, bool vec) {
  __nv_bfloat16 o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    __nv_bfloat16 hi, lo;
    tc::split_bf16(h8[i], hi, lo);
    o[i] = lo_plane ? lo : hi;
  }
  if (vec) {
    *reinterpret_cast<uint4*>(dst) = make_uint4(pack2_bf16(o[0], o[1]), pack2_bf16(o[2], o[3]), pack2_bf16(o[4], o[5]), pack2_bf16(o[6], o[7]));
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = o[i];
  }
}

// Pointwise part: thread (unit ul, utterance pair bp).  The new hidden values are staged in h_s[8][64] and leave the CTA as
// one 16-byte store per (utterance, destination, plane) instead of 2-byte scattered stores.
__device__ __forceinline__ void lstm_tail(const float* part, uint32_t rank, int unit0, const float* __restrict__ bias, int H,
                                          float* cst, int B, const HOut& ho, const unsigned char* keep, bool hidden_dropout,
                                          unsigned long long seed, unsigned long long rng_base, float* sav, float* h_s) {
  const int ul = threadIdx.x >> 5, bp = threadIdx.x & 31;
  const int u = unit0 + ul;
  float2 pre[4];
  {  // all sixteen remote loads of the split-K exchange in flight before the first sum (one DSMEM round trip, not four)
    float2 pv[4][4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint32_t a = tc::smem_u32(part + (size_t)(32 * (int)rank + g * 8 + ul) * TG_PART_LD + 2 * bp);
#pragma unroll
      for (uint32_t r = 0; r < 4; ++r) pv[g][r] = tc::ld_dsmem2(tc::mapa(a, r));
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) {   // same order as tg::reduce4: ranks 0..3
      const float bv = __ldg(bias + g * H + u);
      pre[g].x = ((pv[g][0].x + pv[g][1].x) + pv[g][2].x) + pv[g][3].x + bv;
      pre[g].y = ((pv[g][0].y + pv[g][1].y) + pv[g][2].y) + pv[g][3].y + bv;
    }
  }
  float hv[2], sv[5][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int b = 2 * bp + j;
    const float pi = j ? pre[0].y : pre[0].x, pf = j ? pre[1].y : pre[1].x;
    const float pg = j ? pre[2].y : pre[2].x, po = j ? pre[3].y : pre[3].x;
    const float si = sigmoid_fast(pi), sf = sigmoid_fast(pf), tg = tanh_fast(pg), so = sigmoid_fast(po);
    const float c = sf * cst[ul * D2_BP + b] + si * tg;
    float h = so * tanh_fast(c);
    if (b < B) cst[ul * D2_BP + b] = c;
    sv[0][j] = si; sv[1][j] = sf; sv[2][j] = tg; sv[3][j] = so; sv[4][j] = c;
    if (hidden_dropout && b < B) {  // nn.Dropout(0.1), train only (model/decoder.py:75,101)
      const bool k = keep ? (keep[(size_t)b * H + u] != 0) : rng_keep(seed, rng_base + (unsigned long long)b * H + u, 0.1f);
      h = k ? h * (1.0f / 0.9f) : 0.f;
    }
    hv[j] = b < B ? h : 0.f;
  }
  if (sav) {  // training tape: [5][H][BP]
    const size_t o = (size_t)u * D2_BP + 2 * bp, gs = (size_t)H * D2_BP;
    if (2 * bp + 1 < B) {
#pragma unroll
      for (int g = 0; g < 5; ++g) __stcs(reinterpret_cast<float2*>(sav + g * gs + o), make_float2(sv[g][0], sv[g][1]));   // write-once tape
    } else if (2 * bp < B) {
#pragma unroll
      for (int g = 0; g < 5; ++g) sav[g * gs + o] = sv[g][0];
    }
  }
  *reinterpret_cast<float2*>(h_s + ul * D2_BP + 2 * bp) = make_float2(hv[0], hv[1]);
  D2_SYNC();
  const int b = threadIdx.x & 63, w = threadIdx.x >> 6;
  if (b < B) {
    float h8[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) h8[i] = h_s[i * D2_BP + b];
    __nv_bfloat16* x = w < 2 ? ho.x0 : ho.x1;
    if (x) {
      const long long plane = w < 2 ? ho.plane0 : ho.plane1;
      const int ld = w < 2 ? ho.ld0 : ho.ld1, col = (w < 2 ? ho.col0 : ho.col1) + unit0;
      store_plane8(x + (w & 1) * plane + (size_t)b * ld + col, h8, w & 1, ((ld | col) & 7) == 0);
    }
    if (ho.rm && (ho.x1 ? w < 2 : w >= 2)) {
      float* dst = ho.rm + (size_t)b * ho.rm_ld + ho.rm_col + unit0 + (w & 1) * 4;
      const float* src = h8 + (w & 1) * 4;
      if (((ho.rm_ld | (ho.rm_col + unit0)) & 3) == 0) *reinterpret_cast<float4*>(dst) = make_float4(src[0], src[1], src[2], src[3]);
      else { dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; dst[3] = src[3]; }
    }
  }
}

// Query projection (model/attention.py:52) folded into the attention-LSTM phase: this CTA's 8 hidden units contribute
// qp[b][a] = sum_u Wq[a][unit0+u] h[u][b]; the attention phase adds the 128 per-CTA partials in a fixed order.
// Thread = (4 attention dims, 8 utterances); h_s is the staging buffer lstm_tail just filled.
__device__ __forceinline__ void query_partial(const float* h_s, const float* wq_s, int A, int B, float* qp, size_t bstride) {
  const int a4 = (threadIdx.x & 31) * 4, b0 = (threadIdx.x >> 5) * 8;
  if (a4 >= A) return;
  float acc[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[j][0] = 0.f; acc[j][1] = 0.f; acc[j][2] = 0.f; acc[j][3] = 0.f; }
#pragma unroll 2
  for (int uu = 0; uu < 8; ++uu) {
    const float4 w4 = *reinterpret_cast<const float4*>(wq_s + uu * A + a4);
    const float4 ha = *reinterpret_cast<const float4*>(h_s + uu * D2_BP + b0);
    const float4 hb = *reinterpret_cast<const float4*>(h_s + uu * D2_BP + b0 + 4);
    const float hj[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc[j][0] = fmaf(hj[j], w4.x, acc[j][0]); acc[j][1] = fmaf(hj[j], w4.y, acc[j][1]);
      acc[j][2] = fmaf(hj[j], w4.z, acc[j][2]); acc[j][3] = fmaf(hj[j], w4.w, acc[j][3]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (b0 + j < B) *reinterpret_cast<float4*>(qp + (size_t)(b0 + j) * bstride + a4) = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
}

}  // namespace
