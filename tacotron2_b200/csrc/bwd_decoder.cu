// Persistent backward kernel of the teacher-forced decoder loop (BPTT over model/tacotron2.py:276-317 with
// model/decoder.py:53-119 and model/attention.py:34-69).  Same grid as the forward: 32 clusters x 4 CTAs, one CTA per SM.
// Per frame t (descending) three phases separated by grid barriers:
//   G  input-gradient GEMMs on tcgen05: dx_dec(t) = W_dec^T dgates_dec(t) and dx_att(t+1) = W_att^T dgates_att(t+1)
//      (transposed tile-major weight packs, split-K over the cluster, DSMEM reduce) -- both in the same phase
//   A  attention backward for frame t: context / softmax / energies / location features, recomputing the forward's
//      tanh arguments; accumulates the small attention weight gradients in registers for the whole utterance
//   P  pointwise LSTMCell backward: attention cell of frame t (needs dq from A) and decoder cell of frame t-1
// Everything that is not on the recurrent dependency chain is left to bulk tensor-core GEMMs after the loop (weight
// gradients from the saved dgates^T, memory gradient from the saved d ctx, prenet backward).
#include "train.cuh"

namespace {

struct BSmem {
  uint8_t* stages;    // TMA ring / phase scratch
  float* part;        // [128][68] split-K partial
  float* dc_att;      // [8][64] carried d/d c of this CTA's units
  float* dc_dec;
  float* conv_t;      // [2][31][32]
  float* dense;       // [32][A]   (f-major, as the forward)
  float* dense_af;    // [A][36]   (a-major copy for d f = ds * Wd: 16-byte rows over f)
  float* v;           // [A]
  uint8_t* gfT;       // tensor-core attention path (bwd_decoder_att_tc.inl), in place of conv_t / dense / dense_af:
                      // folded filter bank G[(c,j)][a] as bf16 (hi,lo) K-major sw128 tiles [2 hl][2 kt][64 rows][128 B]
  uint8_t* ident;     // 16 x 16 identity as one K-major sw128 tile (2 KB)
  float* wq;          // [A][8] query-weight columns of this CTA's 8 units, resident (tensor-core attention path: behind `ident`)
  uint64_t* bars;     // ring full[3] empty[3], accumulator, spare, attention H done, attention all done
  uint32_t* tmem_slot;   // [0] GEMM accumulators (256 columns), [1] d pm + Y accumulators of the tensor-core attention path (256 columns)
};

__device__ __forceinline__ BSmem carve(uint8_t* base, const D2Dims& d) {
  BSmem s;
  s.stages = base;
  uint8_t* p = base + D2_STAGES * TG_STAGE_BYTES;
  s.part = reinterpret_cast<float*>(p);    p += TG_M * TG_PART_LD * 4;
  s.dc_att = reinterpret_cast<float*>(p);  p += 8 * D2_BP * 4;
  s.dc_dec = reinterpret_cast<float*>(p);  p += 8 * D2_BP * 4;
  s.gfT = p; s.ident = p + 32768;          // (offset 182 KB from the 1024-aligned base: tile aligned)
  s.wq = reinterpret_cast<float*>(p + 32768 + 2048);
  s.conv_t = reinterpret_cast<float*>(p);  p += 2 * D2_LOC_K * D2_LOC_F * 4;
  s.dense = reinterpret_cast<float*>(p);   p += d.A * D2_LOC_F * 4;
  s.dense_af = reinterpret_cast<float*>(p); p += d.A * 36 * 4;
  s.v = reinterpret_cast<float*>(p);       p += d.A * 4;
  s.bars = reinterpret_cast<uint64_t*>(p); p += (2 * D2_STAGES + 4) * 8;
  s.tmem_slot = reinterpret_cast<uint32_t*>(p);
  return s;
}

size_t bsmem_bytes(const D2Dims& d) {
  return 1024 + (size_t)D2_STAGES * TG_STAGE_BYTES + TG_M * TG_PART_LD * 4 + 2 * 8 * D2_BP * 4 + 2 * D2_LOC_K * D2_LOC_F * 4 +
         (size_t)d.A * D2_LOC_F * 4 + (size_t)d.A * 36 * 4 + d.A * 4 + (2 * D2_STAGES + 4) * 8 + 64;
}

__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int i = 1; i < D2_THREADS / 32; ++i) r += red[i];
  return r;
}

// ------------------------------------------------------------------------------------------------------------------
// Pointwise LSTMCell backward for this CTA's 8 units x 64 utterances (thread = (unit = warp, utterance pair = lane)).
// dh_extra: optional per-thread additive term (query backward).  Writes dgates to the packed operand buffer (through
// `stg`, 16-byte stores) and to the K-major weight-gradient operand; carries d c in shared memory; accumulates bias sums.
struct PwArgs {
  const float* sav;        // [5][H][BP] of this frame
  const float* cprev;      // [H][BP] c of the previous frame or null (zero state)
  const float* dh0;        // [..][BP] k-major rows: dh0[(row0 + u)][b]  (may be null)
  const float* dh1;        // second k-major source (may be null)
  int row0, row1;
  const float* dh_rm;      // row-major [BP][ld_rm] source, column col_rm + u (may be null)
  int ld_rm, col_rm;
  const unsigned char* keep;  // [B][H] of this frame or null
  bool hidden_dropout;
  unsigned long long seed, rng_base;
  __nv_bfloat16* dg;       // [2 hl][BP][4H] packed row order
  __nv_bfloat16* dgT;      // [2 hl][4H][ldT] reference row order, column tcol + b
  long long ldT; long long tcol;
};

// The inputs of one thread (tape and gradient rows of both its utterances), loaded ahead of use: the pointwise phase issues the loads of
// both cells and of the query gradient together, so the phase pays one L2 round trip instead of one per step.
struct PwIn {
  float2 si2, sf2, tg2, so2, c2, cp2, d02, d12;
  float rm[2];
  bool keep[2];
};

__device__ __forceinline__ PwIn lstm_pointwise_load(const PwArgs& a, int H, int B, int unit0) {
  const int ul = threadIdx.x >> 5, bp = threadIdx.x & 31;
  const int u = unit0 + ul;
  const size_t gs = (size_t)H * D2_BP;
  // 8-byte loads; the tape is read once -> streaming (evict-first)
  const size_t o2 = (size_t)u * D2_BP + 2 * bp;
  const float2 zero2 = make_float2(0.f, 0.f);
  const bool any = 2 * bp < B;
  PwIn in;
  in.si2 = any ? __ldcs(reinterpret_cast<const float2*>(a.sav + o2)) : zero2;
  in.sf2 = any ? __ldcs(reinterpret_cast<const float2*>(a.sav + gs + o2)) : zero2;
  in.tg2 = any ? __ldcs(reinterpret_cast<const float2*>(a.sav + 2 * gs + o2)) : zero2;
  in.so2 = any ? __ldcs(reinterpret_cast<const float2*>(a.sav + 3 * gs + o2)) : zero2;
  in.c2 = any ? __ldcs(reinterpret_cast<const float2*>(a.sav + 4 * gs + o2)) : zero2;
  in.cp2 = (any && a.cprev) ? __ldcg(reinterpret_cast<const float2*>(a.cprev + o2)) : zero2;   // read again as c of the next iteration
  in.d02 = (any && a.dh0) ? __ldcg(reinterpret_cast<const float2*>(a.dh0 + (size_t)(a.row0 + u) * D2_BP + 2 * bp)) : zero2;
  in.d12 = (any && a.dh1) ? __ldcg(reinterpret_cast<const float2*>(a.dh1 + (size_t)(a.row1 + u) * D2_BP + 2 * bp)) : zero2;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int b = 2 * bp + j;
    in.rm[j] = (b < B && a.dh_rm) ? ld_cg(a.dh_rm + (size_t)b * a.ld_rm + a.col_rm + u) : 0.f;
    in.keep[j] = (b < B && a.hidden_dropout && a.keep) ? (a.keep[(size_t)b * H + u] != 0) : true;
  }
  return in;
}

// compute half: d gates of this thread's (unit, utterance pair) -> weight-gradient operand (global) and the staging tile `stg` [64][33]
__device__ __forceinline__ void lstm_pointwise_compute(const PwArgs& a, const PwIn& in, int H, int B, int unit0, float* dcs, float* stg,
                                                       const float* dh_extra2, float* bias_acc /*[4]*/) {
  const int ul = threadIdx.x >> 5, bp = threadIdx.x & 31;
  const int u = unit0 + ul;
  float dgv[4][2];
  const float2 si2 = in.si2, sf2 = in.sf2, tg2 = in.tg2, so2 = in.so2, c2 = in.c2, cp2 = in.cp2, d02 = in.d02, d12 = in.d12;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int b = 2 * bp + j;
    float di = 0.f, df = 0.f, dg_ = 0.f, do_ = 0.f;
    if (b < B) {
      const float si = j ? si2.y : si2.x, sf = j ? sf2.y : sf2.x, tg = j ? tg2.y : tg2.x, so = j ? so2.y : so2.x, c = j ? c2.y : c2.x;
      const float cp = j ? cp2.y : cp2.x;
      float dh = dh_extra2 ? dh_extra2[j] : 0.f;
      dh += j ? d02.y : d02.x;
      dh += j ? d12.y : d12.x;
      dh += in.rm[j];
      if (a.hidden_dropout) {  // nn.Dropout(0.1) on h (model/decoder.py:75,101): same draw as the forward
        const bool k = a.keep ? in.keep[j] : rng_keep(a.seed, a.rng_base + (unsigned long long)b * H + u, 0.1f);
        dh = k ? dh * (1.0f / 0.9f) : 0.f;
      }
      const float tc_ = tanh_fast(c);
      do_ = dh * tc_ * so * (1.f - so);
      const float dc = dh * so * (1.f - tc_ * tc_) + dcs[ul * D2_BP + b];
      di = dc * tg * si * (1.f - si);
      df = dc * cp * sf * (1.f - sf);
      dg_ = dc * si * (1.f - tg * tg);
      dcs[ul * D2_BP + b] = dc * sf;
    }
    dgv[0][j] = di; dgv[1][j] = df; dgv[2][j] = dg_; dgv[3][j] = do_;
  }
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    bias_acc[g] += dgv[g][0] + dgv[g][1];
    // K-major weight-gradient operand: row g*H + u, columns tcol + 2bp, +1  (4-byte store, 128 B per warp)
    __nv_bfloat16 h0, l0, h1, l1;
    tc::split_bf16(dgv[g][0], h0, l0);
    tc::split_bf16(dgv[g][1], h1, l1);
    const size_t ro = (size_t)(g * H + u) * a.ldT + a.tcol + 2 * bp;
    // consumed only by the bulk weight-gradient GEMMs after the kernel: streaming stores
    const __nv_bfloat162 hv = __halves2bfloat162(h0, h1), lv = __halves2bfloat162(l0, l1);
    __stcs(reinterpret_cast<unsigned*>(a.dgT + ro), *reinterpret_cast<const unsigned*>(&hv));
    __stcs(reinterpret_cast<unsigned*>(a.dgT + (size_t)4 * H * a.ldT + ro), *reinterpret_cast<const unsigned*>(&lv));
    stg[(2 * bp) * 33 + g * 8 + ul] = dgv[g][0];
    stg[(2 * bp + 1) * 33 + g * 8 + ul] = dgv[g][1];
  }
}

// pack half (after a block barrier): the staged d gates as the packed GEMM operand, rows [pbase, pbase+32) of every utterance, both
// planes: 64 x 2 x 4 chunks of 8 bf16
__device__ __forceinline__ void lstm_pointwise_pack(const PwArgs& a, int H, int pbase, const float* stg) {
#pragma unroll 1
  for (int i = threadIdx.x; i < D2_BP * 2 * 4; i += D2_THREADS) {
    const int ch = i & 3, hl = (i >> 2) & 1, b = i >> 3;
    __align__(16) __nv_bfloat16 v8[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      __nv_bfloat16 hi, lo;
      tc::split_bf16(stg[b * 33 + ch * 8 + e], hi, lo);
      v8[e] = hl ? lo : hi;
    }
    *reinterpret_cast<uint4*>(a.dg + ((size_t)hl * D2_BP + b) * 4 * H + pbase + ch * 8) = *reinterpret_cast<const uint4*>(v8);
  }
}

__device__ __forceinline__ void lstm_pointwise_bwd(const PwArgs& a, const PwIn& in, int H, int B, int unit0, int pbase, float* dcs, float* stg,
                                                   const float* dh_extra2, float* bias_acc /*[4]*/) {
  lstm_pointwise_compute(a, in, H, B, unit0, dcs, stg, dh_extra2, bias_acc);
  __syncthreads();
  lstm_pointwise_pack(a, H, pbase, stg);
  __syncthreads();
}

}  // namespace

#include "bwd_decoder_att.inl"
#include "bwd_decoder_att_tc.inl"
#include "bwd_decoder_main.inl"
