// Pipelined autoregressive decoder kernel (eval-mode synthesis: t2_decode_ar / t2_decode_ar_chunk).
//
// Same arithmetic, tile ownership and reduction orders as decoder_persistent_kernel (decoder.cu) -- 32 clusters x 4 CTAs, one CTA
// per SM, CTA `rank` of cluster `tile` owns 8 hidden units of both LSTM cells and every fourth K-block of their gate contractions --
// but a different SCHEDULE.  The round-1 profile showed a latency chain (DRAM 20 %, tensor pipe 7 %, 77 % of issue slots without an
// eligible warp): HBM and the tensor pipe idled during the attention phase, the output stage and the four grid barriers of a frame,
// and the CUDA cores idled while the two weight streams ran.  Here
//   * two dedicated role warps (TMA producer, MMA issuer) run the weight stream CONTINUOUSLY and in dataflow order: a K-block is
//     issued as soon as the activations it multiplies exist (its "gate" = the grid barrier that publishes them).  Of the 17 K-blocks a
//     CTA contracts per frame only 3 sit on the dependent chain (the prenet K-block of the attention LSTM, the two context K-blocks
//     of the decoder LSTM); the other 14 (att_h / rnn_h / old-context columns) stream behind the attention phase and the output stage;
//   * eight worker warps run the dependent chain: LSTM tails, attention, output stage, grid barriers;
//   * the location term of the attention energies (location conv k=31 + location dense, model/attention.py:54-56) is ONE tcgen05
//     contraction per frame: the two linear layers are folded into a 62-tap filter bank G[a][c][j] at commit, the A operand is the
//     Toeplitz matrix of (previous, cumulative) weights written straight into TENSOR MEMORY by the threads that own the rows, the
//     result stays in TMEM next to a resident copy of this CTA's processed-memory slice, and the energy step reads both with
//     tcgen05.ld.  No shared or global memory holds a location term any more (round 1: 6.5 MB per frame through L2 and 7.3 us of
//     CUDA-core work beside the decoder GEMM), and the stage ring is never used as scratch, which is what lets the stream run on.
#include "decoder.cuh"
#define AR_LDW(p) __ldg(p)   // (ld.global.nc.L1::no_allocate made the context sum 2x slower: the 28 KB L1 does serve these loads)
#define D2_LDW(p) AR_LDW(p)
#define D2_SYNC() asm volatile("bar.sync 1, 256;" ::: "memory")
#include "dec_shared.cuh"

#define AR_THREADS 320     // warps 0..7 workers, warp 8 TMA producer, warp 9 MMA issuer
#define AR_STAGES 2        // two 48 KB ring stages: a third bought nothing (measured), its shared memory holds projection rows instead
#define AR_COL_ATT 0u      // TMEM columns: attention-LSTM accumulator [128 gate rows x 64 utterances]
#define AR_COL_DEC 64u     // decoder-LSTM accumulator
#define AR_COL_LOC 128u    // location term [128 positions x 128 attention dims]
#define AR_COL_PM 256u     // processed memory of this CTA's positions, resident for the whole launch
#define AR_COL_THI 384u    // Toeplitz operand, hi plane: 128 positions x 64 taps bf16 = 32 columns
#define AR_COL_TLO 416u    // lo plane
#define AR_COL_WP 448u     // 8 rows of the mel / gate projection weights (rnn_h half): lane = K / 4 index mod 128, 2 x 32 columns
#define AR_WRES_ROWS 5     // rows of the same weights resident in shared memory (20 KB; the staging buffers of the exchange take 25.5 KB)
#define AR_PROJ_ROWS 21    // most projection rows per CTA on the fast path: the rest (beyond shared and tensor memory) comes from L2
#define AR_WTM_ROWS 8      // rows resident in tensor memory
#define AR_MAX_ENT 40

namespace {

struct KbEnt { unsigned char gemm, kb, gate, flags; };   // flags: 1 first block of its accumulator, 2 last, 4 belongs to frame t+1
enum { G_X = 0, G_Q = 1, G_C = 2, G_R = 3 };

struct ArSmem {
  uint8_t* stages;
  float* part;       // split-K partials | attention scratch | output-stage scratch (phases never overlap, see the hazards note)
  uint8_t* gfilt;    // [2 hi/lo][128 a][64 taps] bf16, 128-B swizzled K-major: B operand of the location contraction
  float* wres;       // [AR_WRES_ROWS][H] first rows of this CTA's slice of the projection weights (rnn_h half), resident
  float* stg;        // [3][32][TG_PART_LD] outgoing partial rows of the split-K exchange, one block per peer
  float* c_att;
  float* c_dec;
  float* wq;
  float* h_s;
  float* v;
  float* octx;
  float* wcum;       // [15 | L | 17] cumulative attention weights of this CTA's utterance (full row, zero halo)
  float* wrow;       // [15 | L | 17] energies -> attention weights of the current frame (full row, zero halo)
  KbEnt* prog;       // per-frame K-block program of this CTA
  uint64_t* bars;
  uint32_t* tmem_slot;
  volatile unsigned* ctl;   // [0] gate count, [1] quit, [2] role failure, [3] entries in prog
};

__host__ __device__ inline size_t ar_smem_bytes(const D2Dims& d) {
  return 1024 + (size_t)AR_STAGES * TG_STAGE_BYTES + (size_t)AR_WRES_ROWS * d.H * 4 + 3 * 32 * TG_PART_LD * 4 + TG_M * TG_PART_LD * 4 + 2 * 128 * 64 * 2 + 2 * 8 * D2_BP * 4 + 8 * (size_t)d.A * 4 + 8 * D2_BP * 4 +
         (size_t)d.A * 4 + 2 * 64 * 4 + 2 * (size_t)(d.L + 32) * 4 + AR_MAX_ENT * sizeof(KbEnt) + 24 * 8 + 80;
}

__device__ __forceinline__ ArSmem ar_carve(uint8_t* base, const D2Dims& d) {
  ArSmem s;
  s.stages = base;
  uint8_t* p = base + (size_t)AR_STAGES * TG_STAGE_BYTES;
  s.gfilt = p;                             p += 2 * 128 * 64 * 2;    // keeps the 1024-B alignment of the ring
  s.wres = reinterpret_cast<float*>(p);    p += (size_t)AR_WRES_ROWS * d.H * 4;
  s.stg = reinterpret_cast<float*>(p);     p += 3 * 32 * TG_PART_LD * 4;
  s.part = reinterpret_cast<float*>(p);    p += TG_M * TG_PART_LD * 4;
  s.c_att = reinterpret_cast<float*>(p);   p += 8 * D2_BP * 4;
  s.c_dec = reinterpret_cast<float*>(p);   p += 8 * D2_BP * 4;
  s.wq = reinterpret_cast<float*>(p);      p += 8 * d.A * 4;
  s.h_s = reinterpret_cast<float*>(p);     p += 8 * D2_BP * 4;
  s.v = reinterpret_cast<float*>(p);       p += d.A * 4;
  s.octx = reinterpret_cast<float*>(p);    p += 2 * 64 * 4;
  s.wcum = reinterpret_cast<float*>(p);    p += (size_t)(d.L + 32) * 4;
  s.wrow = reinterpret_cast<float*>(p);    p += (size_t)(d.L + 32) * 4;
  s.prog = reinterpret_cast<KbEnt*>(p);    p += AR_MAX_ENT * sizeof(KbEnt);
  p = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 7) & ~(uintptr_t)7);
  s.bars = reinterpret_cast<uint64_t*>(p); p += 24 * 8;
  s.tmem_slot = reinterpret_cast<uint32_t*>(p); p += 16;
  s.ctl = reinterpret_cast<volatile unsigned*>(p);
  return s;
}
// barrier slots
enum { B_FULL = 0, B_EMPTY = AR_STAGES, B_ACC_ATT = 2 * AR_STAGES, B_ACC_DEC, B_LOC, B_RV, B_E, B_MEL, B_P1, B_W1, B_G, B_FIN };

__device__ __forceinline__ void wsync() { D2_SYNC(); }

__device__ __forceinline__ unsigned ld_acquire_cta(const volatile unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(tc::smem_u32(const_cast<const unsigned*>(p))) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_cta(volatile unsigned* p, unsigned v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(tc::smem_u32(const_cast<unsigned*>(p))), "r"(v) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {   // transaction count only, no arrival
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T  (A: lane = row, one 32-bit column = two consecutive bf16 along K)
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]),
        "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------------------------ lean math
// __expf / __fdividef carry range fix-up code (FSETP + two FMUL per call); the arguments here never need it.
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_ftz(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float ar_sigmoid(float x) { return rcp_ftz(1.0f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float ar_tanh(float x) {
  const float t = ex2_ftz(-2.8853900817779268f * fabsf(x));
  return copysignf((1.0f - t) * rcp_ftz(1.0f + t), x);
}

// K-block classes by the data the columns hold (decoder.cuh: x_att = [prenet | ctx | att_h], x_dec = [att_h | ctx | controls | rnn_h])
__host__ __device__ inline int att_class(const D2Dims& d, int kb) {   // 0 prenet_t (gate X), 1 ctx_{t-1} (gate C of t-1), 2 att_h_{t-1} (gate Q of t-1)
  const int c0 = kb * TG_BK;
  if (c0 < d.P) return 0;
  if (c0 < d.P + d.EF) return 1;
  return 2;
}
__host__ __device__ inline int dec_class(const D2Dims& d, int kb) {   // 0 att_h_t (gate Q), 1 ctx_t (gate C), 2 rnn_h_{t-1} / controls / padding (gate R of t-1)
  const int c0 = kb * TG_BK, c1 = c0 + TG_BK;
  if (c0 < d.H + d.EF && c1 > d.H) return 1;
  if (c0 < d.H) return 0;
  return 2;
}

// The K-block program of CTA `rank`, in issue order (see the file comment).  Returns the entry count.
__host__ __device__ inline int build_program(const D2Dims& d, int rank, KbEnt* out, bool tf) {
  // att.H blocks streamed behind dec.A in the attention window instead of the decoder-LSTM tail window.  Measured at C3 / C2: teacher
  // forcing 0 -> 19.9 ms per 800-frame forward, 1 -> 19.4, 2 -> 19.3, 3 / 4 -> 19.5; autoregressive 0 -> 30.3 us per frame, 1 -> 30.4, 2 -> 31.0
  const int early_h = 2, early_ar = 0;
  const int nka = d.KattP / TG_BK, nkd = d.KdecP / TG_BK;
  int n = 0;
  auto emit = [&](int gemm, int cls, int gate, int next, int skip = 0, int take = 1 << 20) {   // blocks [skip, skip + take) of the class
    int i = 0;
    for (int kb = rank; kb < (gemm ? nkd : nka); kb += D2_CLUSTER)
      if ((gemm ? dec_class(d, kb) : att_class(d, kb)) == cls) {
        if (i >= skip && i < skip + take) {
          if (n < AR_MAX_ENT) out[n] = KbEnt{(unsigned char)gemm, (unsigned char)kb, (unsigned char)gate, (unsigned char)(next ? 4 : 0)};
          ++n;
        }
        ++i;
      }
  };
  // Order = which barrier window a background block streams in.  The windows after Q (query sum, context sum: 270 KB of CUDA-core
  // loads per CTA) and after X share the SM's L2 ingress with whatever streams there, so only the blocks the decoder cell needs before
  // the context blocks close its accumulator (dec.A) follow Q; att.H / att.C of the next frame follow the on-chain context blocks
  // (decoder-LSTM tail window), dec.R of the next frame follows R (the output stage reads its projection weights from chip memory).
  // Measured at C2 (profiles/r02_v14_ar_ab.log): this order 32.7 us per frame, round 2's first order (att.H right after dec.A) 32.8,
  // dec.R behind the next att.P instead (after X) 33.4.
  if (tf) {
    // Teacher forcing: the prenet block has no gate (bulk prenet before the launch) and there is no output stage after R for the rnn_h
    // blocks to hide behind -- in the order below they sat in front of the next frame's prenet block and the workers waited 5 us per
    // frame for the attention cell's accumulator.  Here the prenet block of frame t+1 closes that accumulator in the decoder-LSTM
    // tail window of frame t, and the rnn_h blocks stream behind the attention-LSTM tail of frame t+1.
    emit(1, 0, G_Q, 0);   // dec.A(t)
    int first_att = n;
    emit(0, 2, G_Q, 1, 0, early_h);        // att.H(t+1), first blocks: behind dec.A in the attention window
    const bool none_early = n == first_att;
    emit(1, 1, G_C, 0);   // dec.C(t)      [on the chain]
    const int last_dec = n - 1;
    if (none_early) first_att = n;
    emit(0, 2, G_Q, 1, early_h);           // att.H(t+1), the rest: with att.C and att.P in the decoder-LSTM tail window
    emit(0, 1, G_C, 1);   // att.C(t+1)
    emit(0, 0, G_X, 1);   // att.P(t+1)
    const int last_att = n - 1;
    const int first_dec = n;
    emit(1, 2, G_R, 1);   // dec.R(t+1)
    if (n <= AR_MAX_ENT && last_dec >= 0 && first_att < n && last_att > first_att && first_dec < n) {
      out[last_att].flags |= 2; out[first_att].flags |= 1; out[last_dec].flags |= 2; out[first_dec].flags |= 1;
    }
    return n;
  }
  emit(0, 0, G_X, 0);   // att.P(t)      [on the chain]
  const int last_att = n - 1;
  emit(1, 0, G_Q, 0);   // dec.A(t)
  int first_att = n;
  emit(0, 2, G_Q, 1, 0, early_ar);   // att.H(t+1), first blocks
  const bool none_early = n == first_att;
  emit(1, 1, G_C, 0);   // dec.C(t)      [on the chain]
  const int last_dec = n - 1;
  if (none_early) first_att = n;
  emit(0, 2, G_Q, 1, early_ar);      // att.H(t+1), the rest
  emit(0, 1, G_C, 1);   // att.C(t+1)
  const int first_dec = n;
  emit(1, 2, G_R, 1);   // dec.R(t+1)
  if (n <= AR_MAX_ENT && last_att >= 0 && first_att < n && last_dec > last_att && first_dec < n) {
    out[last_att].flags |= 2; out[first_att].flags |= 1; out[last_dec].flags |= 2; out[first_dec].flags |= 1;
  }
  return n;
}

// ------------------------------------------------------------------------------------------ grid barrier (workers only)
// One 64-bit word: the low half counts arrivals (release add by thread 0 of every CTA, acquire poll; bit 31 = abort), the high half
// counts the utterances whose stop gate has fired -- it is added with the arrival at the X barrier and read with the poll, so the
// stop check (model/tacotron2.py:321-322) costs no L2 round trip of its own.  Measured alternatives (profiles/r02_v2_ar_barrier*.log):
// one arrival word per CTA polled by a warp +1.5 us per barrier, one add per cluster after a shared-memory count +0.3 us.
struct ArGridBar {
  unsigned* ctr;          // the 64-bit word (8-byte aligned)
  unsigned* abort_flag;   // host-visible copy of the abort state
  unsigned nblocks, epoch;
  int* ok_s;              // shared word
  unsigned poll_sleep;    // ns between polls of the barrier word (fewer readers on its L2 line while the arrivals land)
  __device__ void arrive(bool fail, volatile unsigned* ctl) {
    asm volatile("fence.proxy.async.global;" ::: "memory");   // operands written with generic stores are read by TMA after the barrier
    if (fail) ctl[2] = 1u;
    wsync();
    ++epoch;
    if (threadIdx.x == 0) {
      if (ctl[2] != 0u) { atomicExch(abort_flag, 1u); atomicOr(ctr, 0x80000000u); }
      const unsigned fired = ctl[9];   // stop gates that fired in this CTA since its last arrival (output stage)
      ctl[9] = 0u;
      asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(ctr), "l"(1ull | ((unsigned long long)fired << 32)) : "memory");
    }
  }
  // returns false when the grid aborts; on success thread 0 opens the producer's next gate (ctl[0]) and leaves the done count in ctl[10]
  __device__ bool wait(volatile unsigned* ctl) {
    if (threadIdx.x == 0) {
      const unsigned target = epoch * nblocks;
      unsigned long long vv;
      unsigned v;
      long long spins = 0;
      for (;;) {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(vv) : "l"(ctr) : "memory");
        v = (unsigned)vv;
        if ((v & 0x7fffffffu) >= target || (v & 0x80000000u)) break;
        if (poll_sleep) __nanosleep(poll_sleep);
        if (++spins > (1ll << 24)) { atomicExch(abort_flag, 1u); atomicOr(ctr, 0x80000000u); v = 0x80000000u; break; }   // a peer died: never hang the GPU
      }
      const int ok = (v & 0x80000000u) == 0u;
      *ok_s = ok;
      ctl[10] = (unsigned)(vv >> 32);   // utterances done, as every CTA sees it at this barrier
      if (ok) st_release_cta(ctl, ctl[0] + 1u);
    }
    wsync();
    return *ok_s != 0;
  }
  __device__ bool sync(bool fail, volatile unsigned* ctl) { arrive(fail, ctl); return wait(ctl); }
};

// Split-K exchange of an LSTM gate accumulator, push mode.  Rows [32r, 32r+32) of the 128-row tile belong to cluster rank r, and
// they are exactly the TMEM lanes worker warps r and r+4 can read: every thread sends its 32 partial sums (one row, 32 utterances)
// straight to the owner's receive buffer recv[source rank][row][utterance] with st.async, whose bytes count on the owner's
// transaction barrier.  One trip through the SM-to-SM network, no staging store, no rendezvous, no remote loads.
__device__ __forceinline__ void st_async_v4(uint32_t cluster_addr, float4 v, uint32_t cluster_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(cluster_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void ar_push_partials(uint32_t tmem, uint32_t col, float* recv, float* stg, uint64_t* bar, uint32_t my_rank) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int owner = warp & 3, col0 = 32 * (warp >> 2);
  float v[32];
  // Staged form: the 32 rows for peer r leave as ONE bulk copy shared::cta -> shared::cluster (about 20 B/clk; the per-thread st.async
  // form moved 6 B/clk), the CTA's own rows go straight into its receive block.  Block layout [32 rows][TG_PART_LD] on both sides.
  constexpr uint32_t BLK = 32 * TG_PART_LD * 4;
  if (threadIdx.x == 0) tc::mbar_arrive_expect_tx(bar, 3u * BLK);
  tc::tmem_ld32(tmem + ((uint32_t)(32 * owner) << 16) + col + col0, v);
  tc::tmem_ld_wait();
  tc::tc_fence_before();
  float* blk = (uint32_t)owner == my_rank ? recv + (size_t)my_rank * 32 * TG_PART_LD : stg + (size_t)(((uint32_t)owner - my_rank - 1u) & 3u) * 32 * TG_PART_LD;
  float* dst = blk + (size_t)lane * TG_PART_LD + col0;
#pragma unroll
  for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(dst + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the bulk copy reads the block through the async proxy
  wsync();
  if (threadIdx.x < 3) {
    const uint32_t peer = (my_rank + 1u + threadIdx.x) & 3u;
    const uint32_t src = tc::smem_u32(stg + (size_t)threadIdx.x * 32 * TG_PART_LD);
    const uint32_t dstc = tc::mapa(tc::smem_u32(recv + (size_t)my_rank * 32 * TG_PART_LD), peer), rb = tc::mapa(tc::smem_u32(bar), peer);
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dstc), "r"(src), "r"(BLK), "r"(rb) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the block may be overwritten by the next exchange
  }
}

// LSTM pointwise tail on the received partials (same summation order as lstm_tail of dec_shared.cuh: ranks 0..3, then the bias).
// Thread = (unit ul, utterance pair bp).
__device__ __forceinline__ void ar_lstm_tail(const float* recv, int unit0, const float* __restrict__ bias, int H, float* cst, int B, const HOut& ho,
                                             float* h_s, const unsigned char* keep = nullptr, bool hidden_dropout = false,
                                             unsigned long long seed = 0ull, unsigned long long rng_base = 0ull, float* sav = nullptr) {
  const int ul = threadIdx.x >> 5, bp = threadIdx.x & 31;
  const int u = unit0 + ul;
  float2 pre[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const float* rp = recv + (size_t)(g * 8 + ul) * TG_PART_LD + 2 * bp;
    const float2 p0 = *reinterpret_cast<const float2*>(rp), p1 = *reinterpret_cast<const float2*>(rp + 32 * TG_PART_LD);
    const float2 p2 = *reinterpret_cast<const float2*>(rp + 64 * TG_PART_LD), p3 = *reinterpret_cast<const float2*>(rp + 96 * TG_PART_LD);
    const float bv = __ldg(bias + g * H + u);
    pre[g].x = ((p0.x + p1.x) + p2.x) + p3.x + bv;
    pre[g].y = ((p0.y + p1.y) + p2.y) + p3.y + bv;
  }
  float hv[2], sv[5][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int b = 2 * bp + j;
    const float pi = j ? pre[0].y : pre[0].x, pf = j ? pre[1].y : pre[1].x;
    const float pg = j ? pre[2].y : pre[2].x, po = j ? pre[3].y : pre[3].x;
    const float si = ar_sigmoid(pi), sf = ar_sigmoid(pf), tg = ar_tanh(pg), so = ar_sigmoid(po);
    const float c = sf * cst[ul * D2_BP + b] + si * tg;
    float h = so * ar_tanh(c);
    if (b < B) cst[ul * D2_BP + b] = c;
    sv[0][j] = si; sv[1][j] = sf; sv[2][j] = tg; sv[3][j] = so; sv[4][j] = c;
    if (hidden_dropout && b < B) {  // nn.Dropout(0.1), train only (model/decoder.py:75,101)
      const bool k = keep ? (keep[(size_t)b * H + u] != 0) : rng_keep(seed, rng_base + (unsigned long long)b * H + u, 0.1f);
      h = k ? h * (1.0f / 0.9f) : 0.f;
    }
    hv[j] = b < B ? h : 0.f;
  }
  if (sav) {  // training tape: [5][H][BP] (i, f, g, o post-activation, c), write-once
    const size_t o = (size_t)u * D2_BP + 2 * bp, gs = (size_t)H * D2_BP;
    if (2 * bp + 1 < B) {
#pragma unroll
      for (int g = 0; g < 5; ++g) __stcs(reinterpret_cast<float2*>(sav + g * gs + o), make_float2(sv[g][0], sv[g][1]));
    } else if (2 * bp < B) {
#pragma unroll
      for (int g = 0; g < 5; ++g) sav[g * gs + o] = sv[g][0];
    }
  }
  *reinterpret_cast<float2*>(h_s + ul * D2_BP + 2 * bp) = make_float2(hv[0], hv[1]);
  wsync();
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

// Sums over the 32 lanes of C values per lane by value halving: at each step a lane keeps one half of its values and receives the
// partner's partial sums of that half (C - 1 shuffles in total instead of 5 C).  Afterwards lane l holds the total of v[l * C / 32 ...]
// -- for C = 32: value l in v[0]; for C = 16: value (l >> 1) in v[0] (lane pairs hold the same total).  Fixed order: deterministic.
template <int C>
__device__ __forceinline__ void xreduce(float* v) {
  const int lane = threadIdx.x & 31;
  int m = 16;
#pragma unroll
  for (int c = C; c > 1; c >>= 1, m >>= 1) {
    const bool upper = (lane & m) != 0;
#pragma unroll
    for (int i = 0; i < c / 2; ++i) {
      const float send = upper ? v[i] : v[i + c / 2], keep = upper ? v[i + c / 2] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
    }
  }
#pragma unroll
  for (; m > 0; m >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], m);
}

// Short dense layer for the two utterances of a cluster, 8 rows per warp and pass (as rows_short<8, NK, WS> of dec_shared.cuh), with
// the sixteen sums of a pass reduced by value halving.
template <int NK, bool WS>
__device__ __forceinline__ void ar_rows8_body(const float* __restrict__ W, int ldw, int K4, int rb, int re, const float* x0, const float* x1, float* y,
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
  for (int r = rb; r < re; r += 8) {
    float4 wv[8][NK];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        const int kk = lane + 32 * k;
        const float4* wp = reinterpret_cast<const float4*>(W + (ptrdiff_t)min(r + i, re - 1) * ldw) + kk;
        wv[i][k] = kk < K4 ? (WS ? *wp : AR_LDW(wp)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    __syncwarp();   // scheduling fence: every load of the pass is issued before the first dependent FMA (ptxas otherwise interleaves
                    // them to save registers, which turns one L2 round trip into several)
    float v[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        a0 = fmaf(wv[i][k].x, xa[k].x, a0); a0 = fmaf(wv[i][k].y, xa[k].y, a0); a0 = fmaf(wv[i][k].z, xa[k].z, a0); a0 = fmaf(wv[i][k].w, xa[k].w, a0);
        a1 = fmaf(wv[i][k].x, xb[k].x, a1); a1 = fmaf(wv[i][k].y, xb[k].y, a1); a1 = fmaf(wv[i][k].z, xb[k].z, a1); a1 = fmaf(wv[i][k].w, xb[k].w, a1);
      }
      v[i] = a0; v[8 + i] = a1;
    }
    xreduce<16>(v);
    const int q = lane >> 1, i = q & 7, j = q >> 3;   // lane pair q holds sum q: row r + (q & 7) of utterance q >> 3
    if ((lane & 1) == 0 && r + i < re) y[j * ys + r + i - ybase] = v[0];
  }
  __syncwarp();
}
template <int NK, bool WS>
__device__ __noinline__ void ar_rows8(const float* __restrict__ W, int ldw, int K4, int rb, int re, const float* x0, const float* x1, float* y,
                                      int ybase, int ys) {
  ar_rows8_body<NK, WS>(W, ldw, K4, rb, re, x0, x1, y, ybase, ys);
}

// ------------------------------------------------------------------------------------------ attention
// Scratch (floats, in `part`): q[A] | qh[8A] | ep[2][128] | red[64] | cred[4*cols]; the weight row has its own buffer (sm.wrow)
struct ArAtt { float *q_s, *qh, *ep, *red, *cred, *wrow; };
__device__ __forceinline__ ArAtt ar_att_scratch(const ArSmem& sm, int A) {
  ArAtt a;
  a.q_s = sm.part;
  a.qh = a.q_s + A;
  a.ep = a.qh + 8 * A;
  a.red = a.ep + 256;
  a.cred = a.red + 64;
  a.wrow = sm.wrow;
  return a;
}
__host__ __device__ inline size_t ar_att_scratch_floats(int L, int A, int EF, int S) {
  return 9 * (size_t)A + 256 + 64 + 4 * (size_t)((((EF + S - 1) / S) + 3) & ~3);
}

// Toeplitz rows of this CTA's positions -> tensor memory, then the location contraction (issued by thread 0).
//   T[p][k]      = w   [l0 + p + k - 15]   k = 0..30
//   T[p][32 + k] = wcum[l0 + p + k - 15]   k = 0..30        (taps 31 and 63 are zero, and so are those columns of G)
// Worker warp w owns TMEM lanes 32*(w%4)..+31 (= positions) and writes plane w/4 (0 = hi, 1 = lo).
// The work is cut in two so that each half hides behind a grid barrier of its own (between arrive and wait): `which` 0 = the taps of
// the previous weights (columns 0..15 of a plane, K steps 0-1 of the contraction), 1 = the taps of the cumulative weights (columns
// 16..31, K steps 2-3, then the commit that publishes the location term); 2 = both (prologue).
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void toeplitz_half(const float* src, bool on, int plane, uint32_t taddr) {
  uint32_t r[16];
  if (on && plane == 0) {   // hi plane: one packed conversion per pair of taps (the plane is uniform over the warp)
#pragma unroll
    for (int cc = 0; cc < 16; ++cc) {
      const float x0 = src[2 * cc], x1 = (2 * cc + 1 < D2_LOC_K) ? src[2 * cc + 1] : 0.f;
      const __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
      r[cc] = *reinterpret_cast<const uint32_t*>(&h);
    }
  } else if (on) {          // lo plane: x - bf16(x), rounded again
#pragma unroll
    for (int cc = 0; cc < 16; ++cc) {
      const float x0 = src[2 * cc], x1 = (2 * cc + 1 < D2_LOC_K) ? src[2 * cc + 1] : 0.f;
      const __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
      const uint32_t hb = *reinterpret_cast<const uint32_t*>(&h);
      const __nv_bfloat162 l = __floats2bfloat162_rn(x0 - __uint_as_float(hb << 16), x1 - __uint_as_float(hb & 0xffff0000u));
      r[cc] = *reinterpret_cast<const uint32_t*>(&l);
    }
  } else {
#pragma unroll
    for (int cc = 0; cc < 16; ++cc) r[cc] = 0u;
  }
  tmem_st16(taddr, r);
}
__device__ __forceinline__ void build_toeplitz_and_issue(const ArSmem& sm, const float* wrow, uint32_t tmem, int l0, int nl, int npass,
                                                         uint64_t* loc_bar, int which) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = 32 * (warp & 3) + lane, plane = warp >> 2;
  const bool on = p < nl && (plane == 0 || npass == 3);
  const uint32_t taddr = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + (plane ? AR_COL_TLO : AR_COL_THI);
  // wrow[15 + l] holds position l: tap k reads wrow[l0 + p + k]
  if (which != 1) toeplitz_half(wrow + l0 + p, on, plane, taddr);
  if (which != 0) toeplitz_half(sm.wcum + l0 + p, on, plane, taddr + 16u);
  tmem_st_wait();
  tc::tc_fence_before();
  wsync();
  if (threadIdx.x == 0) {
    tc::tc_fence_after();
    const uint32_t idesc = tc::idesc_bf16_f32(128, 128);
    const uint64_t ghi = tc::smem_desc_sw128(tc::smem_u32(sm.gfilt)), glo = tc::smem_desc_sw128(tc::smem_u32(sm.gfilt) + 128 * 64 * 2);
    const int k0 = which == 1 ? 2 : 0, k1 = which == 0 ? 2 : 4;
#pragma unroll 1
    for (int k = k0; k < k1; ++k) {
      const uint64_t adv = (uint64_t)(k * 32 >> 4);
      umma_bf16_ts(tmem + AR_COL_LOC, tmem + AR_COL_THI + 8u * k, ghi + adv, idesc, k ? 1u : 0u);
      if (npass == 3) {
        umma_bf16_ts(tmem + AR_COL_LOC, tmem + AR_COL_TLO + 8u * k, ghi + adv, idesc, 1u);
        umma_bf16_ts(tmem + AR_COL_LOC, tmem + AR_COL_THI + 8u * k, glo + adv, idesc, 1u);
      }
    }
    if (which != 0) tc::umma_commit(loc_bar);
  }
}

// e + v . tanh(z) for four arguments: tanh(z) = sign(z) (1 - t) / (1 + t), t = exp(-2|z|) in (0, 1]; the four denominators (each in
// (1, 2]) share ONE reciprocal of their product -- 5 MUFU operations instead of 8 (the energy step is MUFU-bound)
__device__ __forceinline__ float tanh4_dot(float4 v, float z0, float z1, float z2, float z3, float e) {
  const float k = -2.8853900817779268f;   // -2 log2(e)
  const float t0 = ex2_ftz(k * fabsf(z0)), t1 = ex2_ftz(k * fabsf(z1)), t2 = ex2_ftz(k * fabsf(z2)), t3 = ex2_ftz(k * fabsf(z3));
  const float a0 = 1.0f + t0, a1 = 1.0f + t1, a2 = 1.0f + t2, a3 = 1.0f + t3;
  const float p01 = a0 * a1, p23 = a2 * a3;
  const float r = rcp_ftz(p01 * p23);
  const float r01 = r * p23, r23 = r * p01;   // 1 / (a0 a1), 1 / (a2 a3)
  e = fmaf(v.x, copysignf((1.0f - t0) * (r01 * a1), z0), e);
  e = fmaf(v.y, copysignf((1.0f - t1) * (r01 * a0), z1), e);
  e = fmaf(v.z, copysignf((1.0f - t2) * (r23 * a3), z2), e);
  e = fmaf(v.w, copysignf((1.0f - t3) * (r23 * a2), z3), e);
  return e;
}

struct ArAttState {
  int b, part, S, l0, nl, Leff;
  bool active, masked;     // masked: the position this thread owns in the energy step is padding
};

// Attention of frame t for this CTA's utterance slice.  Returns false on a wait timeout.  The grid barrier that publishes the context
// is ARRIVED at inside (after the context stores) and the next frame's Toeplitz operand is built while it completes.
template <bool TF>
__device__ bool ar_attention(const D2Dims& d, const D2State& s, const D2IO& io, const ArSmem& sm, const ArAttState& at, uint32_t tmem,
                             int cur, int nxt, int t_out, uint32_t fphase, ArGridBar& bar, bool& alive, long long* prof_last, bool split_toeplitz) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int L = d.L, A = d.A, EF = d.EF, S = at.S;
  const ArAtt as = ar_att_scratch(sm, A);
  bool ok = true;
  auto mark = [&](int slot) {
    if (blockIdx.x == 0 && tid == 0 && s.prof) { const long long now = clock64(); s.prof[slot] += (unsigned long long)(now - *prof_last); *prof_last = now; }
  };
  if (at.active) {
    if (tid == 0) tc::mbar_arrive_expect_tx(sm.bars + B_E, (uint32_t)(L * 4));   // the S energy slices of this frame
    {  // processed query = sum of the 128 per-CTA partials in a fixed order (as decoder.cu)
      const int a4 = lane * 4;
      if (a4 < A) {
        const float* qp = s.qpart + ((size_t)at.b * 128 + warp * 16) * A + a4;
        float4 v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __ldcg(reinterpret_cast<const float4*>(qp + (size_t)i * A));
        float4 acc = v[0];
#pragma unroll
        for (int i = 1; i < 16; ++i) { acc.x += v[i].x; acc.y += v[i].y; acc.z += v[i].z; acc.w += v[i].w; }
        *reinterpret_cast<float4*>(as.qh + warp * A + a4) = acc;
      }
    }
    if (tid < 32) { as.wrow[tid < 15 ? tid : L + tid] = 0.f; }   // halo: 15 zeros in front, 17 behind
    wsync();
    if (tid < A) {
      float qv = as.qh[tid];
#pragma unroll
      for (int g = 1; g < 8; ++g) qv += as.qh[g * A + tid];
      as.q_s[tid] = qv;
      if (TF && at.part == 0) s.q[((size_t)t_out * D2_BP + at.b) * A + tid] = qv;   // training tape
    }
    if (!tc::mbar_wait(sm.bars + B_LOC, fphase)) ok = false;   // location term of this frame is in TMEM
    tc::tc_fence_after();
    wsync();
    mark(10);
    {  // energies: thread = (position = TMEM lane, half of the attention dims); location term + processed memory from tensor memory
      const int lane0 = 32 * (warp & 3), half = warp >> 2, p = lane0 + lane;
      float e = 0.f;
#pragma unroll 1
      for (int c32 = 0; c32 < 2; ++c32) {
        const int a0 = 64 * half + 32 * c32;
        float x[32], pm[32];
        tc::tmem_ld32(tmem + ((uint32_t)lane0 << 16) + AR_COL_LOC + a0, x);
        tc::tmem_ld32(tmem + ((uint32_t)lane0 << 16) + AR_COL_PM + a0, pm);
        tc::tmem_ld_wait();
        if (TF && s.accpm && a0 < A && p < at.nl) {   // training tape: location term + processed memory of this frame (the
          float4* dst = reinterpret_cast<float4*>(s.accpm + (((size_t)t_out * 128 + blockIdx.x) * ((L + S - 1) / S) + p) * A + a0);   // backward's pre-tanh argument - q)
#pragma unroll
          for (int i = 0; i < 8; ++i) __stcs(dst + i, make_float4(x[4 * i] + pm[4 * i], x[4 * i + 1] + pm[4 * i + 1], x[4 * i + 2] + pm[4 * i + 2], x[4 * i + 3] + pm[4 * i + 3]));
        }
        if (a0 < A) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {   // q and v as 16-byte broadcast loads
            const float4 q4 = *reinterpret_cast<const float4*>(as.q_s + a0 + 4 * i), v4 = *reinterpret_cast<const float4*>(sm.v + a0 + 4 * i);
            e = tanh4_dot(v4, q4.x + x[4 * i] + pm[4 * i], q4.y + x[4 * i + 1] + pm[4 * i + 1], q4.z + x[4 * i + 2] + pm[4 * i + 2],
                          q4.w + x[4 * i + 3] + pm[4 * i + 3], e);
          }
        }
      }
      as.ep[half * 128 + p] = e;
    }
    tc::tc_fence_before();
    wsync();
    if (tid < at.nl) {   // masked_fill (model/attention.py:63) and the exchange: every part of the utterance gets the whole energy row
      const float e = at.masked ? -INFINITY : (as.ep[tid] + as.ep[128 + tid]);
      const uint32_t my_rank = tc::cluster_ctarank(), base_rank = my_rank - (uint32_t)at.part;
      const uint32_t dst = tc::smem_u32(as.wrow + 15 + at.l0 + tid), eb = tc::smem_u32(sm.bars + B_E);
      for (int r = 0; r < S; ++r) tc::st_async_f32(tc::mapa(dst, base_rank + (uint32_t)r), e, tc::mapa(eb, base_rank + (uint32_t)r));
    }
    mark(11);
    // context operand (model/attention.py:66-67): the memory rows do not depend on the weights -- the first 8 rows of this thread's
    // column quad are requested now and land during the exchange and the softmax
    const int cols_per = ((EF + S - 1) / S + 3) & ~3;
    const int c0 = at.part * cols_per, c1 = min(EF, c0 + cols_per);
    const int nq = max(0, (c1 - c0 + 3) >> 2);
    const int lg = tid >> 6, cq0 = tid & 63;
    const size_t rs = (size_t)EF / 4;
    const float4* mp0 = reinterpret_cast<const float4*>(io.memory + (size_t)at.b * L * EF + c0);
    const int lclamp = max(at.Leff - 1, 0);
    float4 pf[8];
    if (cq0 < nq) {
#pragma unroll
      for (int u = 0; u < 8; ++u) pf[u] = AR_LDW(mp0 + cq0 + (size_t)min(lg + 4 * u, lclamp) * rs);
    }
    if (!tc::mbar_wait_cluster(sm.bars + B_E, fphase)) ok = false;
    mark(12);
    // softmax over the full row (model/attention.py:65); every thread keeps to its own positions through the three passes
    float* wr = as.wrow + 15;
    float mx = -INFINITY;
#pragma unroll 1
    for (int l = tid; l < L; l += 256) mx = fmaxf(mx, wr[l]);
    mx = warp_max(mx);
    if (lane == 0) as.red[warp] = mx;
    wsync();
    mx = as.red[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, as.red[i]);
    float sum = 0.f;
#pragma unroll 1
    for (int l = tid; l < L; l += 256) {
      const float pe = expf(wr[l] - mx);
      wr[l] = pe;
      sum += pe;
    }
    sum = warp_sum(sum);
    if (lane == 0) as.red[32 + warp] = sum;
    wsync();
    sum = as.red[32];
#pragma unroll
    for (int i = 1; i < 8; ++i) sum += as.red[32 + i];
    const float inv = 1.0f / sum;
    const int l1 = at.l0 + at.nl;
#pragma unroll 1
    for (int l = tid; l < L; l += 256) {
      const float w = wr[l] * inv;
      wr[l] = w;
      sm.wcum[15 + l] += w;                                                   // model/decoder.py:90 (every part keeps the whole row)
      if (l >= at.l0 && l < l1) {
        io.aligns[((size_t)at.b * io.Tmax + t_out) * L + l] = w;
        if (TF) {   // training tape: weights / cumulative weights after this frame (slot t+1)
          s.w_prev[((size_t)nxt * D2_BP + at.b) * L + l] = w;
          s.w_cum[((size_t)nxt * D2_BP + at.b) * L + l] = sm.wcum[15 + l];
        }
      }
    }
    wsync();
    mark(13);
    // context columns of this part: thread = (row group lg of 4, column quad); rows lg, lg + 4, ...: the 8 requested above, then batches of 16 loads; rows
    // beyond the last unmasked position carry weight 0 exactly and are skipped (clamped address, zero weight inside a batch)
#pragma unroll 1
    for (int cq = cq0; cq < nq; cq += 64) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* mp = mp0 + cq;
      if (cq != cq0) {
#pragma unroll
        for (int u = 0; u < 8; ++u) pf[u] = AR_LDW(mp + (size_t)min(lg + 4 * u, lclamp) * rs);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const float w = (lg + 4 * u < at.Leff) ? wr[lg + 4 * u] : 0.f;
        acc.x = fmaf(w, pf[u].x, acc.x); acc.y = fmaf(w, pf[u].y, acc.y);
        acc.z = fmaf(w, pf[u].z, acc.z); acc.w = fmaf(w, pf[u].w, acc.w);
      }
#pragma unroll 1
      for (int l = lg + 32; l < at.Leff; l += 64) {
        float4 m[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) m[u] = AR_LDW(mp + (size_t)min(l + 4 * u, lclamp) * rs);
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          const float w = (l + 4 * u < at.Leff) ? wr[l + 4 * u] : 0.f;
          acc.x = fmaf(w, m[u].x, acc.x); acc.y = fmaf(w, m[u].y, acc.y);
          acc.z = fmaf(w, m[u].z, acc.z); acc.w = fmaf(w, m[u].w, acc.w);
        }
      }
      *reinterpret_cast<float4*>(as.cred + (size_t)lg * cols_per + cq * 4) = acc;
    }
    wsync();
    const long long pa = (long long)D2_BP * d.KattP, pd = (long long)D2_BP * d.KdecP;
    __nv_bfloat16* xa = s.xatt + (size_t)nxt * 2 * pa;   // next frame's attention-LSTM input
    __nv_bfloat16* xd = s.xdec + (size_t)cur * 2 * pd;   // this frame's decoder-LSTM input
#pragma unroll 1
    for (int c = tid; c < c1 - c0; c += 256) {
      const float v = as.cred[c] + as.cred[cols_per + c] + as.cred[2 * cols_per + c] + as.cred[3 * cols_per + c];
      const int col = c0 + c;
      store_split(xa, pa, d.KattP, at.b, d.P + col, v);
      store_split(xd, pd, d.KdecP, at.b, d.H + col, v);
      if (!TF) s.projin[(size_t)at.b * d.KprojP + d.R + col] = v;
    }
  }
  mark(14);
  bar.arrive(!ok, sm.ctl);
  // while the barrier completes: the Toeplitz operand of the NEXT frame's location term (depends on this frame's weights only)
  if (at.active) build_toeplitz_and_issue(sm, sm.wrow, tmem, at.l0, at.nl, d.npass, sm.bars + B_LOC, split_toeplitz ? 0 : 2);
  alive = bar.wait(sm.ctl);
  return ok;
}

// The rnn_h half of the mel / gate projection reads its weights from chip memory when the CTA's rows fit: AR_WRES_ROWS rows in shared
// memory, AR_WTM_ROWS in tensor memory, at most one more from L2 (vanilla: 21 rows per CTA; lane = one float4 of K, K = 1024).
__host__ __device__ inline bool ar_proj_fast(const D2Dims& d) {
  const int rows_per = (d.M + 1 + D2_CLUSTER - 1) / D2_CLUSTER, rq = d.P / D2_CLUSTER;
  const int w1_off = 2 * ((d.M + 4) & ~3) + 2 * d.P + 2 * rq + 6 * (rq > 64 ? rq : 64);
  const bool w1_smem = (d.M >> 2) <= 32 && (w1_off & 3) == 0 && w1_off + rq * d.M <= TG_M * TG_PART_LD;
  return d.H == 1024 && rows_per <= AR_PROJ_ROWS && w1_off + (w1_smem ? rq * d.M : 0) + 8 * 2 * 24 <= TG_M * TG_PART_LD;
}

// ------------------------------------------------------------------------------------------ output stage (all 8 worker warps)
// As output_side / proj_ctx_side of decoder.cu, on eight warps and on the dependent chain (nothing streams on these warps).
__device__ void ar_proj_ctx(const D2Dims& d, const D2Weights& w, const D2State& s, const ArSmem& sm, uint32_t rank, int cluster) {
  const int pw = threadIdx.x >> 5, NW = 8;
  const int B = d.B, M = d.M, KP = d.KprojP, Kc = KP - d.H;
  const int b0 = 2 * cluster;
  if (b0 >= B) return;
  const float* x0 = s.projin + (size_t)b0 * KP + d.H;
  const float* x1 = s.projin + (size_t)min(b0 + 1, B - 1) * KP + d.H;
  const int rows_per = (M + 1 + D2_CLUSTER - 1) / D2_CLUSTER, r0 = (int)rank * rows_per, nr = max(0, min(rows_per, M + 1 - r0));
  const int rpw = (nr + NW - 1) / NW, rb = r0 + pw * rpw, re = min(r0 + nr, rb + rpw);
  rows_dot2<true>(w.Wproj + d.H, KP, Kc >> 2, rb, re, x0, x1, sm.octx, r0, 64);

}

__device__ bool ar_output(const D2Dims& d, const D2Weights& w, const D2State& s, const D2IO& io, const ArSmem& sm, uint32_t rank, int cluster,
                          int tg_, int nxt, uint32_t& cphase, long long* prof_last, uint32_t tmem, bool proj_fast) {
  const int tid = threadIdx.x, lane = tid & 31, pw = tid >> 5, NW = 8;
  const int B = d.B, M = d.M, P = d.P, KP = d.KprojP, MP = (M + 1 + 3) & ~3;
  const int rq = P / D2_CLUSTER;
  float* mel_s = sm.part;
  float* p1_s = mel_s + 2 * MP;
  float* out_s = p1_s + 2 * P;
  const int YS = max(64, rq);
  float* y_s = out_s + 2 * rq;
  float* w1_s = y_s + 6 * YS;
  const int w1_off = 2 * MP + 2 * P + 2 * rq + 6 * YS;
  const bool w1_smem = (M >> 2) <= 32 && (w1_off & 3) == 0 && w1_off + rq * M <= TG_M * TG_PART_LD;
  uint64_t* mel_bar = sm.bars + B_MEL;
  uint64_t* p1_bar = sm.bars + B_P1;
  uint64_t* wbar = sm.bars + B_W1;
  const int b0 = 2 * cluster;
  if (b0 >= B) return true;         // uniform over the cluster
  bool ok = true;
  const int j = lane & 1, b = b0 + j;
  auto mark = [&](int slot) {
    if (blockIdx.x == 0 && tid == 0 && s.prof) { const long long now = clock64(); s.prof[slot] += (unsigned long long)(now - *prof_last); *prof_last = now; }
  };
  if (tid == 0) {
    const uint32_t nu = (uint32_t)min(2, B - b0);
    tc::mbar_arrive_expect_tx(mel_bar, nu * (uint32_t)M * 4u);
    tc::mbar_arrive_expect_tx(p1_bar, nu * (uint32_t)P * 4u);
    if (w1_smem) {
      tc::mbar_arrive_expect_tx(wbar, (uint32_t)(rq * M * 4));
      tc::bulk_load_1d(w1_s, w.W1 + (size_t)rank * rq * M, (uint32_t)(rq * M * 4), wbar);
    }
  }
  const int H4 = d.H >> 2;
  {
    const int rows_per = (M + 1 + D2_CLUSTER - 1) / D2_CLUSTER, r0 = (int)rank * rows_per, nr = max(0, min(rows_per, M + 1 - r0));
    const int rpw = (nr + NW - 1) / NW, rb = r0 + pw * rpw, re = min(r0 + nr, rb + rpw);
    float* pr = mel_s + w1_off + (w1_smem ? rq * M : 0);   // [8 warps][2][24] partial sums of the fast path
    if (proj_fast) {
      // K split over the threads (this thread: float4 index k of every row), weights from shared memory (first AR_WRES_ROWS rows), tensor memory (next 8)
      // and L2 (the rest), all requested at once: one L2 round trip on the chain
      const int k = tid;
      float tw[4 * AR_WTM_ROWS];
      tc::tmem_ld32(tmem + ((uint32_t)(32 * (pw & 3)) << 16) + AR_COL_WP + 32u * (uint32_t)(pw >> 2), tw);
      const float4 xa = __ldcg(reinterpret_cast<const float4*>(s.projin + (size_t)b0 * KP) + k);
      const float4 xb = __ldcg(reinterpret_cast<const float4*>(s.projin + (size_t)min(b0 + 1, B - 1) * KP) + k);
      constexpr int NL2 = AR_PROJ_ROWS - AR_WRES_ROWS - AR_WTM_ROWS;
      float4 wl[NL2];
#pragma unroll
      for (int i = 0; i < NL2; ++i)
        wl[i] = AR_LDW(reinterpret_cast<const float4*>(w.Wproj + (size_t)(r0 + min(AR_WRES_ROWS + AR_WTM_ROWS + i, nr - 1)) * KP) + k);
      tc::tmem_ld_wait();
      const float4* ws = reinterpret_cast<const float4*>(sm.wres) + k;
#pragma unroll
      for (int c = 0; c < 24; c += 8) {
        float v[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int ri = c + i;
          float4 wv;
          if (ri < AR_WRES_ROWS) wv = ws[ri * (1024 / 4)];
          else if (ri < AR_WRES_ROWS + AR_WTM_ROWS) wv = make_float4(tw[4 * (ri - AR_WRES_ROWS)], tw[4 * (ri - AR_WRES_ROWS) + 1], tw[4 * (ri - AR_WRES_ROWS) + 2], tw[4 * (ri - AR_WRES_ROWS) + 3]);
          else if (ri < AR_PROJ_ROWS) wv = wl[ri - AR_WRES_ROWS - AR_WTM_ROWS < 0 ? 0 : ri - AR_WRES_ROWS - AR_WTM_ROWS];
          else wv = make_float4(0.f, 0.f, 0.f, 0.f);
          v[i] = fmaf(wv.w, xa.w, fmaf(wv.z, xa.z, fmaf(wv.y, xa.y, wv.x * xa.x)));
          v[8 + i] = fmaf(wv.w, xb.w, fmaf(wv.z, xb.z, fmaf(wv.y, xb.y, wv.x * xb.x)));
        }
        xreduce<16>(v);
        const int q = lane >> 1, i = c + (q & 7);
        if ((lane & 1) == 0 && i < nr) pr[(pw * 2 + (q >> 3)) * 24 + i] = v[0];
      }
      tc::tc_fence_before();
      wsync();
    } else {
      rows_dot2<true>(w.Wproj, KP, H4, rb, re, s.projin + (size_t)b0 * KP, s.projin + (size_t)min(b0 + 1, B - 1) * KP, y_s, r0, YS);
    }
#pragma unroll 1
    for (int r = proj_fast ? r0 + (tid >> 1) : rb + (lane >> 1); r < (proj_fast ? min(r0 + nr, r0 + (tid >> 1) + 1) : re) && b < B; r += 16) {
      float v;
      if (proj_fast) {
        v = pr[j * 24 + r - r0];
#pragma unroll
        for (int wi = 1; wi < 8; ++wi) v += pr[(wi * 2 + j) * 24 + r - r0];
      } else {
        v = y_s[j * YS + r - r0];
      }
      v += sm.octx[j * 64 + r - r0] + __ldg(w.bproj + r);
      if (r < M) {
        io.mels[((size_t)b * io.Tmax + tg_) * M + r] = v;
        st_async_all(mel_s + j * MP + r, v, mel_bar);
      } else {
        io.gates[(size_t)b * io.Tmax + tg_] = v;
        if (io.stop_enabled) {  // model/tacotron2.py:319-320; the running values live in shared memory (this thread is their only writer)
          if (v < 0.0f && sm.ctl[11 + j] == 0u) {
            sm.ctl[11 + j] = 1u;
            s.done[b] = 1;
            atomicAdd(const_cast<unsigned*>(sm.ctl + 9), 1u);   // travels with this CTA's arrival at the X barrier
          }
          if (v >= 0.0f) {
            const unsigned n = sm.ctl[13 + j] + 1u;
            sm.ctl[13 + j] = n;
            s.lengths[b] = (long long)n;
          }
        }
      }
    }
  }
  mark(4);
  const float pscale = 1.0f / (1.0f - d.p_prenet);
  const int pr0 = (int)rank * rq;
  const int rpw = (rq + NW - 1) / NW, rb = pr0 + pw * rpw, re = min(pr0 + rq, rb + rpw);
  float* y1 = y_s + 2 * YS;
  float* y2 = y_s + 4 * YS;
  // prenet layer 2 reads 64 KB of weights per CTA from L2; they depend on nothing, so when the shape is the usual one (8 rows per warp,
  // 256 columns) this lane's sixteen float4 are requested before the wait for the all-gather of layer 1's output
  const bool w2_pre = w1_smem && rq == 64 && P == 256;
  if (!tc::mbar_wait_cluster(mel_bar, cphase)) ok = false;
  if (w2_pre) {
    if (!tc::mbar_wait(wbar, cphase)) ok = false;
    ar_rows8_body<1, true>(w1_s - (ptrdiff_t)pr0 * M, M, M >> 2, rb, re, mel_s, mel_s + MP, y1, pr0, YS);
  } else if (w1_smem) {
    if (!tc::mbar_wait(wbar, cphase)) ok = false;
    ar_rows8<1, true>(w1_s - (ptrdiff_t)pr0 * M, M, M >> 2, rb, re, mel_s, mel_s + MP, y1, pr0, YS);
  } else if ((M >> 2) <= 32) ar_rows8<1, false>(w.W1, M, M >> 2, rb, re, mel_s, mel_s + MP, y1, pr0, YS);
  else rows_dot2<false>(w.W1, M, M >> 2, rb, re, mel_s, mel_s + MP, y1, pr0, YS);
#pragma unroll 1
  for (int r = rb + (lane >> 1); r < re && b < B; r += 16) {
    const size_t kidx = (((size_t)(tg_ + 1) * 2 + 0) * B + b) * P + r;
    const bool keep = io.prenet_keep ? (io.prenet_keep[kidx] != 0) : rng_keep(io.seed, kidx, d.p_prenet);
    const float v = y1[j * YS + r - pr0];
    st_async_all(p1_s + j * P + r, (v > 0.f && keep) ? v * pscale : 0.f, p1_bar);
  }
  mark(5);
  const uint32_t p1_phase = cphase;
  cphase ^= 1u;
  const long long pa = (long long)D2_BP * d.KattP;
  __nv_bfloat16* xa = s.xatt + (size_t)nxt * 2 * pa;
  const bool vec = (rq & 7) == 0 && (d.KattP & 7) == 0;
  if (w2_pre) {
    float4 w2v[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4* wp = reinterpret_cast<const float4*>(w.W2 + (size_t)(rb + i) * P) + lane;
      w2v[i][0] = AR_LDW(wp); w2v[i][1] = AR_LDW(wp + 32);
    }
    if (!tc::mbar_wait_cluster(p1_bar, p1_phase)) ok = false;
    float4 p1a[2], p1b[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      p1a[k] = *reinterpret_cast<const float4*>(p1_s + 4 * (lane + 32 * k));
      p1b[k] = *reinterpret_cast<const float4*>(p1_s + P + 4 * (lane + 32 * k));
    }
    float v[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        a0 = fmaf(w2v[i][k].x, p1a[k].x, a0); a0 = fmaf(w2v[i][k].y, p1a[k].y, a0); a0 = fmaf(w2v[i][k].z, p1a[k].z, a0); a0 = fmaf(w2v[i][k].w, p1a[k].w, a0);
        a1 = fmaf(w2v[i][k].x, p1b[k].x, a1); a1 = fmaf(w2v[i][k].y, p1b[k].y, a1); a1 = fmaf(w2v[i][k].z, p1b[k].z, a1); a1 = fmaf(w2v[i][k].w, p1b[k].w, a1);
      }
      v[i] = a0; v[8 + i] = a1;
    }
    xreduce<16>(v);
    const int q = lane >> 1;
    if ((lane & 1) == 0) y2[(q >> 3) * YS + rb + (q & 7) - pr0] = v[0];
    __syncwarp();
  } else {
    if (!tc::mbar_wait_cluster(p1_bar, p1_phase)) ok = false;
    if ((P >> 2) <= 64) ar_rows8<2, false>(w.W2, P, P >> 2, rb, re, p1_s, p1_s + P, y2, pr0, YS);
    else rows_dot2<false>(w.W2, P, P >> 2, rb, re, p1_s, p1_s + P, y2, pr0, YS);
  }
#pragma unroll 1
  for (int r = rb + (lane >> 1); r < re; r += 16) {
    float v = 0.f;
    if (b < B) {
      const size_t kidx = (((size_t)(tg_ + 1) * 2 + 1) * B + b) * P + r;
      const bool keep = io.prenet_keep ? (io.prenet_keep[kidx] != 0) : rng_keep(io.seed, kidx, d.p_prenet);
      v = y2[j * YS + r - pr0];
      v = (v > 0.f && keep) ? v * pscale : 0.f;
      if (!vec) store_split(xa, pa, d.KattP, b, r, v);
    }
    if (vec) out_s[j * rq + r - pr0] = v;
  }
  if (vec) {
    wsync();
#pragma unroll 1
    for (int i = tid; i < 4 * (rq >> 3); i += 256) {
      const int pl = i & 1, jj = (i >> 1) & 1, c8 = i >> 2;
      if (b0 + jj < B) store_plane8(xa + pl * pa + (size_t)(b0 + jj) * d.KattP + pr0 + 8 * c8, out_s + jj * rq + 8 * c8, pl, true);
    }
  }
  mark(6);
  return ok;
}

// ------------------------------------------------------------------------------------------ the kernel
// Profiling slots (CTA 0, thread 0; SM-clock cycles summed over the launch):
//   8 wait for the attention-LSTM accumulator (prenet K-block on the chain), 0 its tail, 1 query partials + Q barrier,
//   10 query sum + location-term wait, 11 energies + exchange issue, 12 exchange wait, 13 softmax, 14 context + stores,
//   2 C barrier with the Toeplitz build beside it, 15 context half of the projection, 9 wait for the decoder-LSTM accumulator,
//   3 its tail + R barrier, 4/5/6 projection / prenet 1 / prenet 2, 7 X barrier + stop check
template <bool TF>   // TF: teacher-forced training forward (history mode: per-frame tape slots), else autoregressive eval
__global__ void __launch_bounds__(AR_THREADS, 1)
decoder_ar_kernel(const __grid_constant__ D2Maps maps, const __grid_constant__ CUtensorMap gmap, D2Dims d, D2Weights w, D2State s, D2IO io) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  const ArSmem sm = ar_carve(base, d);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool worker = warp < 8;
  const uint32_t rank = tc::cluster_ctarank();
  const int tile = blockIdx.x / D2_CLUSTER;
  const int unit0 = tile * 32 + (int)rank * 8;
  const int B = d.B, L = d.L, A = d.A;
  const int S = (B * 4 <= (int)gridDim.x) ? 4 : ((B * 2 <= (int)gridDim.x) ? 2 : 1);
  int& ok_word = *reinterpret_cast<int*>(const_cast<unsigned*>(sm.ctl + 6));     // (dynamic shared memory: the kernel has no static allocation)

  // ---- one-time setup
  if (tid == 0) {
    tc::prefetch_tmap(&maps.w_att); tc::prefetch_tmap(&maps.w_dec); tc::prefetch_tmap(&maps.x_att); tc::prefetch_tmap(&maps.x_dec);
    tc::prefetch_tmap(&gmap);
    for (int i = 0; i < AR_STAGES; ++i) { tc::mbar_init(sm.bars + B_FULL + i, 1); tc::mbar_init(sm.bars + B_EMPTY + i, 1); }
    tc::mbar_init(sm.bars + B_ACC_ATT, 1); tc::mbar_init(sm.bars + B_ACC_DEC, 1); tc::mbar_init(sm.bars + B_LOC, 1);
    tc::mbar_init(sm.bars + B_RV, 1); tc::mbar_init(sm.bars + B_E, 1); tc::mbar_init(sm.bars + B_MEL, 1);
    tc::mbar_init(sm.bars + B_P1, 1); tc::mbar_init(sm.bars + B_W1, 1); tc::mbar_init(sm.bars + B_G, 1); tc::mbar_init(sm.bars + B_FIN, 1);
    tc::fence_barrier_init();
    sm.ctl[0] = TF ? 0u : 1u;   // gate count; autoregressive: the (virtual) barrier that publishes the prenet columns of
                                                // frame 0 counts as passed.  Teacher forcing has three barriers per frame (no X)
    sm.ctl[1] = 0u; sm.ctl[2] = 0u; sm.ctl[4] = 0u; sm.ctl[5] = 0u;
    for (int i = 8; i < 15; ++i) sm.ctl[i] = 0u;   // 8 cluster counter (barrier mode 2), 9 fired stop gates, 10 done count, 11-12 fired flags, 13-14 lengths
    sm.ctl[3] = (unsigned)build_program(d, (int)rank, sm.prog, TF);
  }
  if (warp == 9) tc::tmem_alloc<512>(sm.tmem_slot);   // warp 9 owns alloc + dealloc
  if (worker) {
#pragma unroll 1
    for (int i = tid; i < A; i += 256) sm.v[i] = w.v[i];
#pragma unroll 1
    for (int i = tid; i < 8 * A; i += 256) sm.wq[i] = w.Wq[(size_t)(i % A) * d.H + unit0 + i / A];
#pragma unroll 1
    for (int i = tid; i < 8 * D2_BP; i += 256) {
      sm.c_att[i] = s.att_c[(size_t)(unit0 + i / D2_BP) * D2_BP + i % D2_BP];
      sm.c_dec[i] = s.rnn_c[(size_t)(unit0 + i / D2_BP) * D2_BP + i % D2_BP];
    }
#pragma unroll 1
    for (int i = tid; i < L + 32; i += 256) sm.wcum[i] = 0.f;
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *sm.tmem_slot;
  if (tid == 0) {   // the folded location filter bank G (hi | lo), 128-B swizzled by TMA
    tc::mbar_arrive_expect_tx(sm.bars + B_G, 2u * 128u * 64u * 2u);
    tc::tma_load_3d(sm.gfilt, &gmap, sm.bars + B_G, 0, 0, 0);
  }
  tc::cluster_sync();   // barrier inits of the whole cluster are visible before any remote arrive

  const int nprog = (int)sm.ctl[3];
  constexpr int nst = AR_STAGES;
  const int nkb_att = d.KattP / TG_BK, nkb_dec = d.KdecP / TG_BK;
  const uint32_t stage_w = (uint32_t)(d.npass == 3 ? 2 * TG_W_TILE : TG_W_TILE), stage_x = (uint32_t)(d.npass == 3 ? 2 * TG_X_TILE : TG_X_TILE);

  if (warp == 8) {
    // =============================================================== TMA producer
    if (lane == 0) {
      const uint64_t wpol = tc::l2_policy_evict_first();
      // sequence position -> (frame, entry): frame -1 (prologue) issues only the entries that belong to the following frame
      int w_f = -1, w_e = 0, x_f = -1, x_e = 0;
      long long w_n = 0, x_n = 0;
      auto skip = [&](int f, int e) {   // entry e is not issued in frame f
        const bool next = (sm.prog[e].flags & 4) != 0;
        return (f < 0 && !next) || (f == io.n_steps - 1 && next);
      };
      auto advance = [&](int& f, int& e) {
        do { if (++e == nprog) { e = 0; ++f; } } while (f < io.n_steps && skip(f, e));
      };
      if (skip(w_f, w_e)) advance(w_f, w_e);
      if (skip(x_f, x_e)) advance(x_f, x_e);
      int w_stage = 0, x_stage = 0;
      uint32_t w_phase = 0;
      long long idle = 0;
      unsigned fenced_gate = 0;
      while (x_f < io.n_steps) {
        bool progressed = false;
        if (w_f < io.n_steps && w_n - x_n < nst && tc::mbar_try_wait(sm.bars + B_EMPTY + w_stage, w_phase ^ 1)) {
          const KbEnt en = sm.prog[w_e];
          uint8_t* st = sm.stages + (size_t)w_stage * TG_STAGE_BYTES;
          mbar_expect_tx(sm.bars + B_FULL + w_stage, stage_w);
          const int nkb = en.gemm ? nkb_dec : nkb_att;
          tc::tma_load_3d_hint(st, en.gemm ? &maps.w_dec : &maps.w_att, sm.bars + B_FULL + w_stage, 0, (tile * nkb + en.kb) * TG_M, 0, wpol);
          if (++w_stage == nst) { w_stage = 0; w_phase ^= 1; }
          ++w_n;
          advance(w_f, w_e);
          progressed = true;
        }
        if (x_n < w_n) {
          const KbEnt en = sm.prog[x_e];
          // teacher forcing: the prenet columns of every frame exist before the launch (bulk prenet), three barriers per frame (Q, C, R)
          const int need = TF ? (en.gate == G_X ? -1 : 3 * x_f + (int)en.gate) : 4 * x_f + (int)en.gate + 1;
          const unsigned have = ld_acquire_cta(sm.ctl);
          if ((int)have >= need) {
            if (have != fenced_gate) { asm volatile("fence.proxy.async;" ::: "memory"); fenced_gate = have; }   // peers' generic stores -> TMA reads
            const int gf = x_f + ((en.flags & 4) ? 1 : 0);
            uint8_t* st = sm.stages + (size_t)x_stage * TG_STAGE_BYTES;
            tc::mbar_arrive_expect_tx(sm.bars + B_FULL + x_stage, stage_x);
            const int slot = TF ? gf : (gf & 1);   // history mode (training tape): one slot per frame instead of two parities
            const bool pre = TF && !en.gemm && en.gate == G_X;   // prenet K-block: from the bulk prenet output
            tc::tma_load_3d(st + 2 * TG_W_TILE, pre ? &maps.tf_pre : (en.gemm ? &maps.x_dec : &maps.x_att), sm.bars + B_FULL + x_stage, en.kb * TG_BK, 0,
                            (pre ? gf : slot) * 2);
            if (++x_stage == nst) x_stage = 0;
            ++x_n;
            advance(x_f, x_e);
            progressed = true;
          }
        }
        if (progressed) { idle = 0; continue; }
        __nanosleep(32);   // nothing to issue: leave the issue slots of this scheduler to the worker warps
        if (sm.ctl[1] != 0u) break;                              // workers stopped (stop gate, abort): no more gates will open
        if (++idle > (1ll << 24)) { sm.ctl[2] = 1u; break; }     // never hang the GPU
      }
      // weight tiles already in flight for K-blocks that will never run (early stop): complete their stages with the (stale) x
      // tile so that no bulk copy is outstanding when the CTA exits; the MMA issuer drains them
      while (x_n < w_n) {
        const KbEnt en = sm.prog[x_e];
        uint8_t* st = sm.stages + (size_t)x_stage * TG_STAGE_BYTES;
        tc::mbar_arrive_expect_tx(sm.bars + B_FULL + x_stage, stage_x);
        tc::tma_load_3d(st + 2 * TG_W_TILE, en.gemm ? &maps.x_dec : &maps.x_att, sm.bars + B_FULL + x_stage, en.kb * TG_BK, 0, 0);
        if (++x_stage == nst) x_stage = 0;
        ++x_n;
        advance(x_f, x_e);
      }
      st_release_cta(sm.ctl + 4, (unsigned)x_n);   // total K-blocks issued
      st_release_cta(sm.ctl + 5, 1u);              // producer done
    }
    __syncwarp();
  } else if (warp == 9) {
    // =============================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = tc::idesc_bf16_f32(TG_M, TG_N);
      int f = -1, e = 0;
      auto skip = [&](int ff, int ee) {
        const bool next = (sm.prog[ee].flags & 4) != 0;
        return (ff < 0 && !next) || (ff == io.n_steps - 1 && next);
      };
      auto advance = [&]() { do { if (++e == nprog) { e = 0; ++f; } } while (f < io.n_steps && skip(f, e)); };
      if (skip(f, e)) advance();
      int stage = 0;
      uint32_t phase = 0;
      long long n = 0;
      bool out = false;
      while (f < io.n_steps && !out) {
        long long spins = 0;
        while (!tc::mbar_try_wait(sm.bars + B_FULL + stage, phase)) {
          if (ld_acquire_cta(sm.ctl + 5) != 0u && (long long)ld_acquire_cta(sm.ctl + 4) == n) { out = true; break; }   // producer ended: nothing more comes
          if (++spins > (1ll << 26)) { sm.ctl[2] = 1u; out = true; break; }
        }
        if (out) break;
        tc::tc_fence_after();
        const KbEnt en = sm.prog[e];
        const uint32_t sbase = tc::smem_u32(sm.stages + (size_t)stage * TG_STAGE_BYTES);
        const uint64_t whi = tc::smem_desc_sw128(sbase), wlo = tc::smem_desc_sw128(sbase + TG_W_TILE);
        const uint64_t xhi = tc::smem_desc_sw128(sbase + 2 * TG_W_TILE), xlo = tc::smem_desc_sw128(sbase + 2 * TG_W_TILE + TG_X_TILE);
        const uint32_t dcol = tmem + (en.gemm ? AR_COL_DEC : AR_COL_ATT);
        uint32_t acc = (en.flags & 1) ? 0u : 1u;
#pragma unroll
        for (int k = 0; k < TG_BK / 16; ++k) {
          const uint64_t adv = (uint64_t)(k * 32 >> 4);
          tc::umma_bf16(dcol, whi + adv, xhi + adv, idesc, acc);
          if (d.npass == 3) {
            tc::umma_bf16(dcol, whi + adv, xlo + adv, idesc, 1u);
            tc::umma_bf16(dcol, wlo + adv, xhi + adv, idesc, 1u);
          }
          acc = 1u;
        }
        tc::umma_commit(sm.bars + B_EMPTY + stage);
        if (en.flags & 2) tc::umma_commit(sm.bars + (en.gemm ? B_ACC_DEC : B_ACC_ATT));
        if (++stage == nst) { stage = 0; phase ^= 1; }
        ++n;
        advance();
      }
      tc::umma_commit(sm.bars + B_FIN);   // every MMA of this thread has retired when this fires
      tc::mbar_wait(sm.bars + B_FIN, 0);
    }
    __syncwarp();
  } else {
    // =============================================================== workers
    ArGridBar bar{s.barrier, s.barrier + 2, gridDim.x, 0u, &ok_word, 50u};   // measured at C2: no pause 30.6 us per frame, 50 ns 30.3, 150 ns 30.5, 300 ns 31.0
    constexpr bool split_toeplitz = true;   // (whole build behind the C barrier: C barrier 3.0 instead of 2.3 us)
    long long prof_last = clock64();
    auto prof_mark = [&](int slot) {
      if (blockIdx.x == 0 && tid == 0 && s.prof) { const long long now = clock64(); s.prof[slot] += (unsigned long long)(now - prof_last); prof_last = now; }
    };
    ArAttState at{};
    at.S = S; at.b = blockIdx.x / S; at.part = blockIdx.x % S; at.active = at.b < B;
    {
      const int Lc = (L + S - 1) / S;
      at.l0 = at.part * Lc;
      at.nl = max(0, min(L, at.l0 + Lc) - at.l0);
    }
    at.masked = false; at.Leff = 0;
    if (tid == 0 && !tc::mbar_wait(sm.bars + B_G, 0)) sm.ctl[2] = 1u;   // the filter bank has landed (every CTA: no copy may be in flight at exit)
    wsync();
    if (at.active) {
      const unsigned char* mk = io.mask + (size_t)at.b * L;
      if (tid < at.nl) at.masked = mk[at.l0 + tid] != 0;
      int le = 0;
#pragma unroll 1
      for (int l = tid; l < L; l += 256) if (mk[l] == 0) le = max(le, l + 1);
      le = (int)warp_max((float)le);
      if (lane == 0) sm.octx[warp] = (float)le;   // (not `part`: a peer's first partial sums may already be arriving there)
      wsync();
      for (int i = 0; i < 8; ++i) at.Leff = max(at.Leff, (int)sm.octx[i]);
      wsync();
      // processed memory of this CTA's positions -> tensor memory (lane = position, column = attention dim), once
      const int lane0 = 32 * (warp & 3), half = warp >> 2, p = lane0 + lane;
#pragma unroll 1
      for (int c32 = 0; c32 < 2; ++c32) {
        const int a0 = 64 * half + 32 * c32;
        uint32_t r[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = 0u;
        if (p < at.nl && a0 < A) {
          const float4* src = reinterpret_cast<const float4*>(io.pmem + ((size_t)at.b * L + at.l0 + p) * A + a0);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 v4 = __ldg(src + i);
            r[4 * i] = __float_as_uint(v4.x); r[4 * i + 1] = __float_as_uint(v4.y); r[4 * i + 2] = __float_as_uint(v4.z); r[4 * i + 3] = __float_as_uint(v4.w);
          }
        }
        tmem_st32(tmem + ((uint32_t)lane0 << 16) + AR_COL_PM + a0, r);
      }
      tmem_st_wait();
      // frame 0: previous and cumulative weights are zero (model/tacotron2.py:126-153) -> the location term is zero; run the
      // same path so that the accumulator columns are defined
      float* wrow = sm.wrow;
#pragma unroll 1
      for (int i = tid; i < L + 32; i += 256) wrow[i] = 0.f;
      wsync();
      build_toeplitz_and_issue(sm, wrow, tmem, at.l0, at.nl, d.npass, sm.bars + B_LOC, 2);
    }
    const long long pa = (long long)D2_BP * d.KattP, pd = (long long)D2_BP * d.KdecP;
    bool alive = true, fail = false, stop = false;
    constexpr bool has_out = !TF;
    const bool proj_fast = has_out && ar_proj_fast(d) && 2 * tile < B;
    if (proj_fast) {   // this CTA's rows of the projection weights (rnn_h half) -> shared memory and tensor memory, once per launch
      const int rows_per = (d.M + 1 + D2_CLUSTER - 1) / D2_CLUSTER, r0 = (int)rank * rows_per, nr = max(0, min(rows_per, d.M + 1 - r0));
      const float4* wp = reinterpret_cast<const float4*>(w.Wproj);
      const size_t ld4 = (size_t)(d.KprojP >> 2);
#pragma unroll 1
      for (int i = tid; i < min(AR_WRES_ROWS, nr) * 256; i += 256)
        reinterpret_cast<float4*>(sm.wres)[i] = __ldg(wp + (size_t)(r0 + (i >> 8)) * ld4 + (i & 255));
      uint32_t r[32];
#pragma unroll
      for (int i = 0; i < AR_WTM_ROWS; ++i) {
        const int row = AR_WRES_ROWS + i;
        const float4 v4 = row < nr ? __ldg(wp + (size_t)(r0 + row) * ld4 + tid) : make_float4(0.f, 0.f, 0.f, 0.f);
        r[4 * i] = __float_as_uint(v4.x); r[4 * i + 1] = __float_as_uint(v4.y); r[4 * i + 2] = __float_as_uint(v4.z); r[4 * i + 3] = __float_as_uint(v4.w);
      }
      tmem_st32(tmem + ((uint32_t)(32 * (warp & 3)) << 16) + AR_COL_WP + 32u * (uint32_t)(warp >> 2), r);
      tmem_st_wait();
      tc::tc_fence_before();
      wsync();
    }
    uint32_t rv_phase = 0, cphase = 0;
    int frames = 0;
#pragma unroll 1
    for (int t = 0; t < io.n_steps && alive && !stop; ++t) {
      const int tg_ = io.t0 + t;
      const int cur = TF ? tg_ : (t & 1), nxt = TF ? tg_ + 1 : (cur ^ 1);   // history mode: one slot per frame (training tape)
      const unsigned char* att_keep = (TF && io.att_keep) ? io.att_keep + (size_t)tg_ * B * d.H : nullptr;
      const unsigned char* rnn_keep = (TF && io.rnn_keep) ? io.rnn_keep + (size_t)tg_ * B * d.H : nullptr;
      const unsigned long long rb_att = (1ull << 40) + (unsigned long long)tg_ * B * d.H, rb_dec = (2ull << 40) + (unsigned long long)tg_ * B * d.H;
      float* sav_att = (TF && s.sav_att) ? s.sav_att + (size_t)tg_ * 5 * d.H * D2_BP : nullptr;
      float* sav_dec = (TF && s.sav_dec) ? s.sav_dec + (size_t)tg_ * 5 * d.H * D2_BP : nullptr;
      const bool hdrop = TF && io.hidden_dropout != 0;
      // ================= attention LSTMCell (model/decoder.py:70-75): only the prenet K-block of the contraction is still to come
      if (!tc::mbar_wait(sm.bars + B_ACC_ATT, (uint32_t)(t & 1))) fail = true;
      tc::tc_fence_after();
      prof_mark(8);
      {
        const HOut ho{s.xatt + (size_t)nxt * 2 * pa, d.KattP, d.P + d.EF, pa, s.xdec + (size_t)cur * 2 * pd, d.KdecP, 0, pd, s.att_h_rm, d.H, 0};
        ar_push_partials(tmem, AR_COL_ATT, sm.part, sm.stg, sm.bars + B_RV, rank);
        if (!tc::mbar_wait_cluster(sm.bars + B_RV, rv_phase)) fail = true;
        rv_phase ^= 1u;
        ar_lstm_tail(sm.part, unit0, w.batt, d.H, sm.c_att, B, ho, sm.h_s, att_keep, hdrop, io.seed, rb_att, sav_att);
      }
      prof_mark(0);
      query_partial(sm.h_s, sm.wq, A, B, s.qpart + (size_t)blockIdx.x * A, (size_t)128 * A);
      if (!(alive = bar.sync(fail, sm.ctl))) break;
      prof_mark(1);
      // ================= location-sensitive attention (model/attention.py:54-67, model/decoder.py:90) + C barrier
      if (!ar_attention<TF>(d, s, io, sm, at, tmem, cur, nxt, tg_, (uint32_t)(t & 1), bar, alive, &prof_last, split_toeplitz)) fail = true;
      if (!alive) break;
      prof_mark(2);
      // ================= decoder LSTMCell (model/decoder.py:94-101): the context K-blocks are on their way; meanwhile the context
      // half of this frame's mel / gate projection
      if (has_out) ar_proj_ctx(d, w, s, sm, rank, tile);
      prof_mark(15);
      if (!tc::mbar_wait(sm.bars + B_ACC_DEC, (uint32_t)(t & 1))) fail = true;
      tc::tc_fence_after();
      prof_mark(9);
      {
        const HOut ho{s.xdec + (size_t)nxt * 2 * pd, d.KdecP, d.H + d.EF + d.C, pd, nullptr, 0, 0, 0, TF ? nullptr : s.projin, d.KprojP, 0};
        ar_push_partials(tmem, AR_COL_DEC, sm.part, sm.stg, sm.bars + B_RV, rank);
        if (!tc::mbar_wait_cluster(sm.bars + B_RV, rv_phase)) fail = true;
        rv_phase ^= 1u;
        ar_lstm_tail(sm.part, unit0, w.bdec, d.H, sm.c_dec, B, ho, sm.h_s, rnn_keep, hdrop, io.seed, rb_dec, sav_dec);
      }
      bar.arrive(fail, sm.ctl);
      // behind the R barrier (training forward) or the X barrier (autoregressive: the output stage reads tensor memory right after R):
      // the second half of the next frame's Toeplitz operand (cumulative weights) and the commit
      if (TF && split_toeplitz && at.active) build_toeplitz_and_issue(sm, sm.wrow, tmem, at.l0, at.nl, d.npass, sm.bars + B_LOC, 1);
      if (!(alive = bar.wait(sm.ctl))) break;
      prof_mark(3);
      // ================= mel / gate projection, stop bookkeeping, prenet on the emitted frame (model/decoder.py:103-109,
      // model/tacotron2.py:311-325)
      frames = t + 1;
      if (!has_out) continue;   // teacher-forced training: the projection is one bulk GEMM after the kernel, the prenet ran before it
      if (!ar_output(d, w, s, io, sm, rank, tile, tg_, nxt, cphase, &prof_last, tmem, proj_fast)) fail = true;
      if (t + 1 < io.n_steps) {
        bar.arrive(fail, sm.ctl);
        if (split_toeplitz && at.active) build_toeplitz_and_issue(sm, sm.wrow, tmem, at.l0, at.nl, d.npass, sm.bars + B_LOC, 1);
        if (!(alive = bar.wait(sm.ctl))) break;
        // model/tacotron2.py:321-322: leave once every utterance has fired its stop gate (the count arrived with the barrier word)
        if (io.stop_enabled && (int)sm.ctl[10] >= B && t + 1 >= io.min_frames) stop = true;
        prof_mark(7);
      }
    }
    if (fail) sm.ctl[2] = 1u;
    wsync();
    if (tid == 0) {
      st_release_cta(sm.ctl + 1, 1u);   // role warps: no more gates
      if (blockIdx.x == 0) {
        const bool good = alive && sm.ctl[2] == 0u;
        *s.n_frames = good ? frames : -1;
        if (!good && s.status) atomicExch(s.status, 1u);
      }
    }
#pragma unroll 1
    for (int i = tid; i < 8 * D2_BP; i += 256) {
      s.att_c[(size_t)(unit0 + i / D2_BP) * D2_BP + i % D2_BP] = sm.c_att[i];
      s.rnn_c[(size_t)(unit0 + i / D2_BP) * D2_BP + i % D2_BP] = sm.c_dec[i];
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::cluster_sync();   // no CTA exits while a peer may still write into its shared memory
  if (warp == 9) tc::tmem_dealloc<512>(tmem);
}

}  // namespace

bool d2ar_supported(const D2Dims& d) {
  if (getenv("T2B200_AR_KERNEL") && !strcmp(getenv("T2B200_AR_KERNEL"), "old")) return false;
  const int grid = (d.H / 32) * D2_CLUSTER;
  if (grid != 128 || d.H != d.R || d.B < 1 || d.B > D2_BP) return false;
  const int S = (d.B * 4 <= grid) ? 4 : ((d.B * 2 <= grid) ? 2 : 1);
  const int Lc = (d.L + S - 1) / S;
  if (Lc > 128 || d.A > 128 || d.A % 32 != 0 || d.P % (TG_BK * D2_CLUSTER) != 0 || d.EF % 4 != 0 || d.M % 4 != 0) return false;
  if (d.H % (TG_BK * D2_CLUSTER) != 0 || d.KattP / TG_BK > 255 || d.KdecP / TG_BK > 255) return false;
  KbEnt tmp[AR_MAX_ENT];
  for (int r = 0; r < 2 * D2_CLUSTER; ++r) {   // both orders (autoregressive, teacher forcing)
    const int n = build_program(d, r % D2_CLUSTER, tmp, r >= D2_CLUSTER);
    if (n > AR_MAX_ENT) return false;
    int la = 0, fa = 0, ld = 0, fd = 0;
    for (int i = 0; i < n; ++i) {
      if (tmp[i].flags & 1) (tmp[i].gemm ? fd : fa)++;
      if (tmp[i].flags & 2) (tmp[i].gemm ? ld : la)++;
    }
    if (la != 1 || fa != 1 || ld != 1 || fd != 1) return false;   // every rank needs a prenet, an att_h, a ctx and an rnn_h K-block
  }
  if (ar_att_scratch_floats(d.L, d.A, d.EF, S) * 4 > (size_t)TG_M * TG_PART_LD * 4) return false;
  if (((size_t)2 * (d.M + 4) + 2 * d.P + 2 * (d.P / D2_CLUSTER) + 6 * 64) * 4 > (size_t)TG_M * TG_PART_LD * 4 || d.P / D2_CLUSTER > 64 ||
      (d.M + 1 + D2_CLUSTER - 1) / D2_CLUSTER > 64)
    return false;
  return ar_smem_bytes(d) <= 227 * 1024 - 1024;   // (the 1024-aligned extern array costs up to 1 KB of the 227 KB opt-in limit)
}

int d2ar_launch(const D2Dims& d, const D2Weights& w, const D2State& s, const D2IO& io, const D2Maps& maps, const CUtensorMap& gmap,
                cudaStream_t st) {
  T2_REQUIRE(d2ar_supported(d), T2_ERR_UNSUPPORTED, "decoder (pipelined AR): unsupported shape B=%d L=%d", d.B, d.L);
  T2_REQUIRE((!io.teacher_forcing && !io.hist && !io.hidden_dropout) || (io.teacher_forcing && io.hist), T2_ERR_ARG,
             "decoder (pipelined): autoregressive eval mode or the teacher-forced training forward (history mode) only");
  const size_t smem = ar_smem_bytes(d);
  auto kern = io.teacher_forcing ? decoder_ar_kernel<true> : decoder_ar_kernel<false>;
  T2_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute at[1];
  cfg.gridDim = dim3(128);
  cfg.blockDim = dim3(AR_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = D2_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int nclusters = 0;
  T2_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg));
  T2_REQUIRE(nclusters * D2_CLUSTER >= 128, T2_ERR_UNSUPPORTED, "decoder (pipelined AR): only %d clusters of %d can be co-resident", nclusters,
             D2_CLUSTER);
  T2_CHECK_CUDA(cudaMemsetAsync(s.barrier, 0, 4 * sizeof(unsigned), st));   // [0..1] 64-bit barrier word, [2] abort flag
  T2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, maps, gmap, d, w, s, io));
  return T2_OK;
}
