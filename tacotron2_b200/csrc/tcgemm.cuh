// Cluster split-K GEMM on tcgen05: D[128 rows x 64 batch] (fp32, TMEM) = sum over this CTA's K-blocks of
// W[128 x 64k] * x[64 x 64k]^T with operands streamed by TMA into a 128B-swizzled shared-memory ring.
// fp32-parity mode runs three bf16 passes per K-step on (hi,lo) splits: Whi*xhi + Whi*xlo + Wlo*xhi.
// Used by the decoder's LSTM gate contractions (model/decoder.py:74,100) and by the debug GEMM entry.
#pragma once
#include "tc.cuh"

#define TG_BK 64
#define TG_M 128
#define TG_N 64
#define TG_W_TILE (TG_M * TG_BK * 2)  // bytes of one bf16 W tile (16 KB)
#define TG_X_TILE (TG_N * TG_BK * 2)  // bytes of one bf16 x tile (8 KB)
#define TG_STAGE_BYTES (2 * TG_W_TILE + 2 * TG_X_TILE)  // hi+lo of both operands (48 KB)
#define TG_TMEM_COLS 256   // two accumulators of 128 columns: [Whi*xhi + Wlo*xhi | Whi*xlo] each
#define TG_PART_LD 68  // floats per row of the split-K partial buffer (64 + 4 pad: conflict-free float4 rows)

#ifdef __CUDACC__
namespace tg {

struct Pipe {
  uint64_t* full;    // [nstages]
  uint64_t* empty;   // [nstages]
  uint64_t* accum;   // [1]
  uint8_t* stages;   // nstages * TG_STAGE_BYTES, 1024-B aligned
  int nstages;
  int npass;         // 3 = fp32-parity (hi/lo split), 1 = plain bf16
  uint32_t tmem;     // TMEM base (lane 0, column 0 of the allocation)
  // ring state; each role thread advances its own copy
  int p_stage; uint32_t p_phase;
  int c_stage; uint32_t c_phase;
  uint32_t a_phase;
};

struct XSrc {  // where the x (activation) K-blocks come from: blocks [0,split) from map0, the rest from map1
  const CUtensorMap* map0; int z0;
  const CUtensorMap* map1; int z1;
  int split;
};

template <int COLS = TG_TMEM_COLS>   // tensor-memory columns to allocate (a CTA allocates once: the permit is relinquished right after)
__device__ __forceinline__ void pipe_init(Pipe& p, uint64_t* bars, uint8_t* stages, int nstages, int npass, uint32_t* tmem_slot) {
  p.full = bars; p.empty = bars + nstages; p.accum = bars + 2 * nstages;
  p.stages = stages; p.nstages = nstages; p.npass = npass;
  p.p_stage = p.c_stage = 0; p.p_phase = p.c_phase = p.a_phase = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < nstages; ++i) { tc::mbar_init(p.full + i, 1); tc::mbar_init(p.empty + i, 1); }
    tc::mbar_init(p.accum, 1);
    tc::fence_barrier_init();
  }
  if ((threadIdx.x >> 5) == 1) tc::tmem_alloc<COLS>(tmem_slot);  // warp 1 owns alloc + dealloc
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  p.tmem = *tmem_slot;
}

// Runs the K-blocks kb = kb_begin, kb_begin+kb_stride, ... < kb_end of one 128-row tile into TMEM columns
// [d_col, d_col+64) (bf16 mode) or [d_col, d_col+128) (fp32-parity mode, two half sums).  All threads of the CTA must call it; returns false on a pipeline timeout.
// w_nkb > 0: W is stored tile-major ([tile][kb][128][64], every TMA box one contiguous 16 KB chunk) with w_nkb K-blocks per
// tile; w_nkb == 0: W is a plain row-major [rows][K] matrix.
// `side` runs on warps 2..7 (192 threads, named barrier 1 is theirs) while warps 0/1 stream and issue the GEMM: work that does not
// depend on this GEMM's result and needs no part of the stage ring.
struct NoSide { __device__ __forceinline__ void operator()() const {} };
// w_policy: L2 eviction priority of the W tiles: 0 = default, 1 = evict-first (pure stream), 2 = evict-last (keep resident)
// acc_init: 0 = the first MMA overwrites the accumulator columns, 1 = continue a sum an earlier call left there (a contraction
// issued in two parts because some K-blocks of x become available later than the others)
// prefetch_w_kb >= 0: after its own K-blocks the producer also claims the next ring stage and starts the W box of that K-block
// (the x box of it is not ready yet); the follow-up call passes first_w_issued = true and only adds the x box of its first
// K-block to that stage.  The follow-up call MUST be made (the stage's transaction count is otherwise never met).
template <class Side>
__device__ __forceinline__ bool gemm(Pipe& p, const CUtensorMap* wmap, int w_row0, int w_nkb, const XSrc& xs, int kb_begin, int kb_end,
                                     int kb_stride, uint32_t d_col, Side side, int w_policy = 0, uint32_t acc_init = 0u,
                                     int prefetch_w_kb = -1, bool first_w_issued = false) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  bool ok = true;
  const uint32_t stage_tx = (uint32_t)(p.npass == 3 ? TG_STAGE_BYTES : TG_W_TILE + TG_X_TILE);
  if (warp == 0 && lane == 0) {
    // ---------------- TMA producer
    asm volatile("fence.proxy.async;" ::: "memory");  // operands written by generic stores in an earlier phase
    // The weight stream (tens of MB per frame, each byte used once per frame by one SM) must not push the small per-frame
    // working set (attention memory, processed memory, partial sums) out of L2: evict-first.
    const uint64_t wpol = w_policy == 2 ? tc::l2_policy_evict_last() : tc::l2_policy_evict_first();
    auto load_w = [&](uint8_t* st, int kb) {
      const int wc0 = w_nkb ? 0 : kb * TG_BK, wc1 = w_nkb ? ((w_row0 / TG_M) * w_nkb + kb) * TG_M : w_row0;
      if (w_policy) tc::tma_load_3d_hint(st, wmap, p.full + p.p_stage, wc0, wc1, 0, wpol);   // [hi|lo] 32 KB (hi only when box z = 1)
      else tc::tma_load_3d(st, wmap, p.full + p.p_stage, wc0, wc1, 0);
    };
    for (int kb = kb_begin; kb < kb_end; kb += kb_stride) {
      uint8_t* st = p.stages + (size_t)p.p_stage * TG_STAGE_BYTES;
      if (!(first_w_issued && kb == kb_begin)) {
        if (!tc::mbar_wait(p.empty + p.p_stage, p.p_phase ^ 1)) { ok = false; break; }
        tc::mbar_arrive_expect_tx(p.full + p.p_stage, stage_tx);
        load_w(st, kb);
      }
      const bool first = kb < xs.split;
      const CUtensorMap* xm = first ? xs.map0 : xs.map1;
      const int xk = (first ? kb : kb - xs.split) * TG_BK, xz = first ? xs.z0 : xs.z1;
      tc::tma_load_3d(st + 2 * TG_W_TILE, xm, p.full + p.p_stage, xk, 0, xz);      // [hi|lo] 16 KB
      if (++p.p_stage == p.nstages) { p.p_stage = 0; p.p_phase ^= 1; }
    }
    if (ok && prefetch_w_kb >= 0) {
      if (!tc::mbar_wait(p.empty + p.p_stage, p.p_phase ^ 1)) ok = false;
      else {
        tc::mbar_arrive_expect_tx(p.full + p.p_stage, stage_tx);
        load_w(p.stages + (size_t)p.p_stage * TG_STAGE_BYTES, prefetch_w_kb);
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ---------------- MMA issuer
    const uint32_t idesc = tc::idesc_bf16_f32(TG_M, TG_N), idesc2 = tc::idesc_bf16_f32(TG_M, 2 * TG_N);
    uint32_t acc = acc_init;
    for (int kb = kb_begin; kb < kb_end; kb += kb_stride) {
      if (!tc::mbar_wait(p.full + p.c_stage, p.c_phase)) { ok = false; break; }
      tc::tc_fence_after();
      const uint32_t sbase = tc::smem_u32(p.stages + (size_t)p.c_stage * TG_STAGE_BYTES);
      const uint64_t whi = tc::smem_desc_sw128(sbase), wlo = tc::smem_desc_sw128(sbase + TG_W_TILE);
      const uint64_t xhi = tc::smem_desc_sw128(sbase + 2 * TG_W_TILE), xlo = tc::smem_desc_sw128(sbase + 2 * TG_W_TILE + TG_X_TILE);
#pragma unroll
      for (int k = 0; k < TG_BK / 16; ++k) {
        const uint64_t adv = (uint64_t)(k * 32 >> 4);  // 16 bf16 = 32 B along K inside the 128-B swizzle atom
        if (p.npass == 3) {
          // The x stage holds the hi tile and the lo tile back to back (64 + 64 rows of one K-major swizzled tile): ONE N = 128
          // MMA gives Whi*xhi in columns [0,64) and Whi*xlo in columns [64,128) and reads the Whi tile from shared memory once
          // instead of twice; Wlo*xhi then adds into columns [0,64).  The epilogue sums the two halves (accum_to_smem).
          tc::umma_bf16(p.tmem + d_col, whi + adv, xhi + adv, idesc2, acc);
          tc::umma_bf16(p.tmem + d_col, wlo + adv, xhi + adv, idesc, 1u);
        } else {
          tc::umma_bf16(p.tmem + d_col, whi + adv, xhi + adv, idesc, acc);
        }
        acc = 1u;
      }
      tc::umma_commit(p.empty + p.c_stage);  // frees the smem slot when these MMAs retire
      if (++p.c_stage == p.nstages) { p.c_stage = 0; p.c_phase ^= 1; }
    }
    tc::umma_commit(p.accum);  // accumulator complete
  } else if (warp >= 2) {
    side();
  }
  __syncwarp();
  // everyone waits for the accumulator
  if (!tc::mbar_wait(p.accum, p.a_phase)) ok = false;
  p.a_phase ^= 1;
  tc::tc_fence_after();
  return __syncthreads_and(ok ? 1 : 0) != 0;
}

__device__ __forceinline__ bool gemm(Pipe& p, const CUtensorMap* wmap, int w_row0, int w_nkb, const XSrc& xs, int kb_begin, int kb_end,
                                     int kb_stride, uint32_t d_col, int w_policy = 0) {
  return gemm(p, wmap, w_row0, w_nkb, xs, kb_begin, kb_end, kb_stride, d_col, NoSide{}, w_policy);
}

// TMEM accumulator columns [d_col, d_col+64) -> this CTA's partial buffer part[128][TG_PART_LD] (shared memory).
// 8 warps: warp w reads lanes 32*(w%4).. and columns 32*(w/4)..
__device__ __forceinline__ void accum_to_smem(const Pipe& p, uint32_t d_col, float* part) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < 8) {
    const int lane0 = 32 * (warp & 3), col0 = 32 * (warp >> 2);
    float v[32];
    tc::tmem_ld32(p.tmem + ((uint32_t)lane0 << 16) + d_col + col0, v);
    if (p.npass == 3) {   // fp32-parity mode: columns [64,128) hold the Whi*xlo part of the sum
      float u[32];
      tc::tmem_ld32(p.tmem + ((uint32_t)lane0 << 16) + d_col + TG_N + col0, u);
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += u[i];
    }
    tc::tmem_ld_wait();
    float* dst = part + (size_t)(lane0 + lane) * TG_PART_LD + col0;
#pragma unroll
    for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(dst + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  tc::tc_fence_before();
}

// Sum of the 4 cluster CTAs' partials at (row, col..col+1); fixed order -> deterministic.
__device__ __forceinline__ float2 reduce4(const float* part, int row, int col) {
  const uint32_t a = tc::smem_u32(part + (size_t)row * TG_PART_LD + col);
  float2 s = tc::ld_dsmem2(tc::mapa(a, 0));
#pragma unroll
  for (uint32_t r = 1; r < 4; ++r) {
    const float2 t = tc::ld_dsmem2(tc::mapa(a, r));
    s.x += t.x; s.y += t.y;
  }
  return s;
}

}  // namespace tg
#endif

// Host-side helpers (tcgemm.cu)
int tg_split_bf16_launch(const float* src, __nv_bfloat16* dst_hi_lo /*[2][rows][ld_dst]*/, int rows, int cols, int ld_src, int ld_dst,
                         long long plane_stride, cudaStream_t st);
int tg_debug_gemm(const float* W, const float* x, float* out, int rows, int K, int npass, void* ws, long long ws_bytes, cudaStream_t st);
