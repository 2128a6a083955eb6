// C ABI of the training path: train-mode forward entry points that record a tape, the matching backward entry
// points, the loss, and the gradient-destination registry (include/t2b200.h, "training" section).
#include "ctx.cuh"
#include "train.cuh"
#include "gemmnt.cuh"

#include <algorithm>
#include <vector>

extern "C" int t2_ctx_set_grad(t2_ctx* c, const char* key, float* ptr, long long numel) {
  T2_REQUIRE(c && key, T2_ERR_ARG, "t2_ctx_set_grad: null argument");
  if (!ptr) { c->grads.erase(key); return T2_OK; }
  c->grads[key] = GradSlot{ptr, numel};
  return T2_OK;
}

#define NEED_GRAD(var, key, n)                                                                       \
  float* var = find_grad(c, key, (long long)(n));                                                    \
  T2_REQUIRE(var, T2_ERR_STATE, "gradient destination for %s not registered or wrong size (t2_ctx_set_grad)", std::string(key).c_str())
#define NEED_PARAM(var, key, n)                                                                      \
  const Param* var = find(c, key);                                                                   \
  T2_REQUIRE(var && var->numel == (long long)(n), T2_ERR_STATE, "parameter %s missing or wrong size", std::string(key).c_str())

namespace {

inline int gsz(long long n) { return (int)std::max<long long>(1, std::min<long long>((n + 255) / 256, 148 * 8)); }
inline int round8(int x) { return (x + 7) & ~7; }

// channels-last bf16 (hi,lo) [2][B][T][C] -> K-major [2][C][B][Tp], optionally time-shifted: dst[c][b][t] = x[b][t+shift][c]
int to_kmajor(const __nv_bfloat16* x, __nv_bfloat16* xT, int B, int T, int Tp, int C, cudaStream_t st, int shift = 0,
              long long dst_plane_rows = 0) {
  (void)dst_plane_rows;
  return transpose_bf16_launch(x, C, (long long)T * C, (long long)B * T * C, xT, T, C, Tp, (long long)B * Tp, Tp, B, shift, (long long)C * B * Tp, st);
}

// conv weight gradient dW[co][ci][j] = sum_{b,t} dz[b][t][co] x[b][t+j-pad][ci] on the tensor cores: ONE launch for all k taps.  Both
// operands stay channels-last ([2][B][T][C] bf16 hi,lo: dz as the BatchNorm backward leaves it, x as the forward taped it) and are read
// MN-major, so time is an outer TMA coordinate: the tap shift is a coordinate offset with out-of-range rows zero-filled (the conv
// padding) -- no K-major transposes, no k shifted copies (they were 37 launches and 3.1 ms of a training step).
int conv_wgrad(t2_ctx* c, const __nv_bfloat16* dz, const __nv_bfloat16* x, int B, int T, int Cout, int Cin, int k, float* dW, float* partial,
               long long partial_floats, cudaStream_t st) {
  GemmNtArgs a{};
  a.M = Cout; a.N = Cin; a.NT = Cin % 128 == 0 ? 128 : ((Cin + 15) & ~15);
  T2_REQUIRE(a.NT <= 128, T2_ERR_UNSUPPORTED, "conv_wgrad: Cin=%d", Cin);
  a.Kin = T; a.Kout = B; a.nz = k; a.shift0 = -((k - 1) / 2); a.shift_step = 1; a.a_zrow = 0; a.b_zrow = 0; a.mn = 3;
  a.splitk = 0; a.npass = c->npass;
  const long long need = gemmnt_partial_floats(a, c->num_sms);
  T2_REQUIRE(need <= partial_floats, T2_ERR_WORKSPACE, "conv_wgrad: partial buffer too small");
  a.partial = partial;
  RC(gemmnt_make_map_cl(&a.a_map, dz, T, B, Cout, (long long)B * T * Cout, c->npass));
  RC(gemmnt_make_map_cl(&a.b_map, x, T, B, Cin, (long long)B * T * Cin, c->npass));
  GemmNtOut o{dW, (long long)Cin * k, (long long)k, 1ll, 1.0f, 0};
  return gemmnt_launch(a, o, st);
}
long long conv_wgrad_partial_floats(t2_ctx* c, int B, int T, int Cout, int Cin, int k) {
  GemmNtArgs a{};
  a.M = Cout; a.N = Cin; a.NT = Cin % 128 == 0 ? 128 : ((Cin + 15) & ~15);
  a.Kin = T; a.Kout = B; a.nz = k; a.splitk = 0;
  return gemmnt_partial_floats(a, c->num_sms);
}

int conv_tc_res(t2_ctx* c, const __nv_bfloat16* in, const CUtensorMap& wmap, int B, int T, int Cin, int Cout, int k, int NT,
                const float* residual, long long res_bstride, float* out_f32, cudaStream_t st) {
  ConvTcArgs a{};
  RC(make_in_map(&a.in_map, in, B, T, Cin, c->npass));
  a.w_map = wmap;
  a.B = B; a.T = T; a.Cin = Cin; a.Cout = Cout; a.k = k; a.NT = NT; a.npass = c->npass; a.act = ACT_NONE;
  a.drop_scale = 1.f; a.out_f32 = out_f32; a.residual = residual; a.res_bstride = res_bstride;
  return convtc_launch(a, st);
}

// ------------------------------------------------------------------------------------------ postnet
struct PostTape { __nv_bfloat16* x[5]; float* z[5]; float* mean[5]; float* rstd[5]; };
PostTape plan_post_tape(const t2_ctx* c, int B, int T, Arena& ar) {
  PostTape t{};
  const int chans[6] = {c->M, c->PD, c->PD, c->PD, c->PD, c->M};
  for (int i = 0; i < 5; ++i) {
    t.x[i] = ar.take<__nv_bfloat16>((size_t)2 * B * T * chans[i]);
    t.z[i] = ar.take<float>((size_t)B * T * chans[i + 1]);
    t.mean[i] = ar.take<float>(chans[i + 1]);
    t.rstd[i] = ar.take<float>(chans[i + 1]);
  }
  return t;
}

BnLayer post_layer(const t2_ctx* c, const PostTape& tp, int i, int B, int T, const unsigned char* keep, unsigned long long seed,
                   const float* gamma, const float* beta) {
  const int chans[6] = {c->M, c->PD, c->PD, c->PD, c->PD, c->M};
  BnLayer L{};
  L.z = tp.z[i]; L.mean = tp.mean[i]; L.rstd = tp.rstd[i]; L.gamma = gamma; L.beta = beta;
  L.C = chans[i + 1]; L.act = i < 4 ? ACT_TANH : ACT_NONE; L.dropout = 1;
  L.keep = keep ? keep + (size_t)i * B * T * c->PD : nullptr; L.keep_ld = c->PD; L.p = c->cfg.dropout;
  L.seed = seed; L.rng_base = (0x50ull + i) << 40;
  return L;
}

}  // namespace

extern "C" long long t2_postnet_tape_bytes(const t2_ctx* c, int B, int T) {
  if (!c || B <= 0 || T <= 0) return -1;
  Arena ar(nullptr, 0);
  plan_post_tape(c, B, T, ar);
  return ar.off + 1024;
}

// Postnet.forward under model.train() (model/postnet.py:8-52): batch statistics, running-stat update, dropout.
extern "C" int t2_postnet_train_fwd(t2_ctx* c, const float* mels, int t_stride, int B, int T, const unsigned char* keep,
                                    unsigned long long seed, float* out, void* tape, long long tape_bytes, void* ws, long long ws_bytes,
                                    void* stream) {
  CHECK_CTX("t2_postnet_train_fwd");
  T2_REQUIRE(mels && out && tape && ws && B > 0 && T > 0 && t_stride >= T, T2_ERR_ARG, "t2_postnet_train_fwd: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  Arena ta(tape, tape_bytes);
  PostTape tp = plan_post_tape(c, B, T, ta);
  T2_REQUIRE(!ta.overflow, T2_ERR_WORKSPACE, "t2_postnet_train_fwd: tape too small");
  Arena wa(ws, ws_bytes);
  void* scratch = wa.take<char>(bn_scratch_bytes(c->PD));
  T2_REQUIRE(!wa.overflow, T2_ERR_WORKSPACE, "t2_postnet_train_fwd: workspace too small");
  const int chans[6] = {c->M, c->PD, c->PD, c->PD, c->PD, c->M};
  RC(split_cl_launch(mels, (long long)t_stride * c->M, tp.x[0], B, T, c->M, st));
  for (int i = 0; i < 5; ++i) {
    const std::string bn = "postnet.postnet." + std::to_string(4 * i + 1);
    NEED_PARAM(bw, bn + ".weight", chans[i + 1]);
    NEED_PARAM(bb, bn + ".bias", chans[i + 1]);
    NEED_PARAM(bm, bn + ".running_mean", chans[i + 1]);
    NEED_PARAM(bv, bn + ".running_var", chans[i + 1]);
    const int Cout = chans[i + 1], NT = Cout % 128 == 0 ? 128 : Cout;
    RC(conv_tc(c, tp.x[i], c->post_wmap[i], B, T, chans[i], Cout, 5, NT, ACT_NONE, nullptr, nullptr, nullptr, 0, tp.z[i], nullptr, st));
    RC(bn_stats_launch(tp.z[i], (long long)B * T, Cout, tp.mean[i], tp.rstd[i], const_cast<float*>(bm->ptr), const_cast<float*>(bv->ptr),
                       scratch, st));
    BnLayer L = post_layer(c, tp, i, B, T, keep, seed, bw->ptr, bb->ptr);
    RC(bn_apply_launch(L, (long long)B * T, i == 4 ? out : nullptr, i == 4 ? nullptr : tp.x[i + 1], st));
  }
  return T2_OK;
}

static int postnet_bwd_impl(t2_ctx* c, int B, int T, const unsigned char* keep, unsigned long long seed, const float* d_out,
                            const float* d_mels_in, const PostTape& tp, float* d_mels, Arena& wa, cudaStream_t st) {
  const int PD = c->PD, M = c->M;
  const int chans[6] = {M, PD, PD, PD, PD, M};
  void* scratch = wa.take<char>(bn_scratch_bytes(PD));
  float* dyA = wa.take<float>((size_t)B * T * PD);
  float* dyB = wa.take<float>((size_t)B * T * PD);
  __nv_bfloat16* dz = wa.take<__nv_bfloat16>((size_t)2 * B * T * PD);
  long long pf = 0;
  for (int i = 0; i < 5; ++i) pf = std::max(pf, conv_wgrad_partial_floats(c, B, T, chans[i + 1], chans[i], 5));
  float* partial = wa.take<float>(pf);
  if (!wa.base) return T2_OK;
  T2_REQUIRE(!wa.overflow, T2_ERR_WORKSPACE, "t2_postnet_bwd: workspace too small");
  const float* dy = d_out;
  float* bufs[2] = {dyA, dyB};
  for (int i = 4; i >= 0; --i) {
    const std::string cv = "postnet.postnet." + std::to_string(4 * i), bn = "postnet.postnet." + std::to_string(4 * i + 1);
    const int Cout = chans[i + 1], Cin = chans[i];
    NEED_PARAM(bw, bn + ".weight", Cout);
    NEED_PARAM(bb, bn + ".bias", Cout);
    NEED_GRAD(gW, cv + ".weight", (size_t)Cout * Cin * 5);
    NEED_GRAD(gg, bn + ".weight", Cout);
    NEED_GRAD(gb, bn + ".bias", Cout);
    BnLayer L = post_layer(c, tp, i, B, T, keep, seed, bw->ptr, bb->ptr);
    RC(bn_bwd_launch(L, dy, (long long)B * T, gg, gb, dz, nullptr, scratch, st));
    RC(conv_wgrad(c, dz, tp.x[i], B, T, Cout, Cin, 5, gW, partial, pf, st));
    float* dst = i == 0 ? d_mels : bufs[i & 1];
    RC(conv_tc_res(c, dz, c->post_wtmap[i], B, T, Cout, Cin, 5, Cin % 128 == 0 ? 128 : Cin, i == 0 ? d_mels_in : nullptr,
                   (long long)T * Cin, dst, st));
    dy = dst;
  }
  return T2_OK;
}

// Backward of Postnet.forward.  d_out (B,T,M) = gradient w.r.t. the postnet output; d_mels_in (B,T,M) or NULL is added to the
// input gradient (the residual / direct decoder-mel gradient); d_mels (B,T,M) receives the total.  Parameter gradients go to
// the destinations registered with t2_ctx_set_grad.
extern "C" int t2_postnet_bwd(t2_ctx* c, int B, int T, const unsigned char* keep, unsigned long long seed, const float* d_out,
                              const float* d_mels_in, const void* tape, long long tape_bytes, float* d_mels, void* ws, long long ws_bytes,
                              void* stream) {
  CHECK_CTX("t2_postnet_bwd");
  T2_REQUIRE(d_out && tape && d_mels && ws && B > 0 && T > 0, T2_ERR_ARG, "t2_postnet_bwd: bad argument");
  Arena ta(const_cast<void*>(tape), tape_bytes);
  PostTape tp = plan_post_tape(c, B, T, ta);
  T2_REQUIRE(!ta.overflow, T2_ERR_WORKSPACE, "t2_postnet_bwd: tape too small");
  Arena wa(ws, ws_bytes);
  return postnet_bwd_impl(c, B, T, keep, seed, d_out, d_mels_in, tp, d_mels, wa, (cudaStream_t)stream);
}

long long postnet_bwd_ws_bytes(t2_ctx* c, int B, int T) {
  Arena wa(nullptr, 0);
  PostTape tp{};
  postnet_bwd_impl(c, B, T, nullptr, 0, nullptr, nullptr, tp, nullptr, wa, nullptr);
  return wa.off;
}

// Loss of training_step (model/tts_model.py:197-201) and its gradients w.r.t. the model outputs.
// mels/post/mel_target (B,T,M), gates/gate_target (B,T); loss (1) device; d_* may be NULL.  ws >= 4 KB.
extern "C" int t2_loss_fwd_bwd(t2_ctx* c, const float* mels, const float* post, const float* gates, const float* mel_target,
                               const float* gate_target, int B, int T, float* loss, float* d_mels, float* d_post, float* d_gates,
                               void* ws, long long ws_bytes, void* stream) {
  T2_REQUIRE(c && mels && post && gates && mel_target && gate_target && loss && ws && ws_bytes >= 4096, T2_ERR_ARG,
             "t2_loss_fwd_bwd: bad argument");
  return loss_launch(mels, post, gates, mel_target, gate_target, (long long)B * T * c->M, (long long)B * T, loss, d_mels, d_post, d_gates,
                     ws, (cudaStream_t)stream);
}

// Backward of t2_finalize (model/tacotron2.py:333-345): gradients w.r.t. the masked outputs -> gradient w.r.t. the raw decoder
// mels through the residual (d_mels_raw), w.r.t. the postnet output (d_postnet_out) and the raw gate logits.
extern "C" int t2_finalize_bwd(t2_ctx* c, const float* d_mels, const float* d_post, const float* d_gates, const long long* lengths, int B,
                               int T, float* d_mels_raw, float* d_postnet_out, float* d_gates_raw, void* stream) {
  T2_REQUIRE(c && lengths && d_mels_raw && d_postnet_out && d_gates_raw && B > 0 && T > 0, T2_ERR_ARG, "t2_finalize_bwd: bad argument");
  return finalize_bwd_launch(d_mels, d_post, d_gates, lengths, B, T, c->M, d_mels_raw, d_postnet_out, d_gates_raw, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ decoder (teacher forced) backward
namespace {

struct Opnd { const __nv_bfloat16* base; int Kin, kin_ld, Kout, rows; long long plane; };

#define BULK_PARTIAL_FLOATS (24ll << 20)   // 96 MB of split-K slabs shared by all bulk weight-gradient GEMMs

// C[z][m][n] = sum_k A[m + z*a_zrow][k] B[n + z*b_zrow][k + shiftB]  (see gemmnt.cuh)
int bulk_gemm(t2_ctx* c, const Opnd& A, const Opnd& Bo, int M, int N, int shiftB, const GemmNtOut& o, float* partial, cudaStream_t st,
              int nz = 1, int a_zrow = 0, int b_zrow = 0, bool direct = false) {
  GemmNtArgs a{};
  a.M = M; a.N = N; a.NT = (N >= 128 || N % 16) ? (N >= 128 ? 128 : ((N + 15) & ~15)) : N;
  a.Kin = A.Kin; a.Kout = A.Kout; a.nz = nz; a.shift0 = shiftB; a.shift_step = 0; a.a_zrow = a_zrow; a.b_zrow = b_zrow;
  a.npass = c->npass;
  T2_REQUIRE(shiftB % 8 == 0, T2_ERR_ARG, "bulk_gemm: operand shift must be a multiple of 8 elements");
  if (direct) { a.splitk = 1; a.direct_out = o.out; a.direct_ldm = o.ldm; }
  else {
    a.splitk = 0;
    const long long need = gemmnt_partial_floats(a, c->num_sms);
    T2_REQUIRE(need <= BULK_PARTIAL_FLOATS, T2_ERR_WORKSPACE, "bulk_gemm: split-K slab of %lld floats exceeds the reserved %lld", need,
               (long long)BULK_PARTIAL_FLOATS);
    a.partial = partial;
  }
  RC(gemmnt_make_map(&a.a_map, A.base, A.Kin, A.kin_ld, A.Kout, A.rows, A.plane, 128, c->npass));
  RC(gemmnt_make_map(&a.b_map, Bo.base, Bo.Kin, Bo.kin_ld, Bo.Kout, Bo.rows, Bo.plane, a.NT, c->npass));
  return gemmnt_launch(a, o, st);
}

// The same product with the B operand read MN-major from the decoder's operand rows as the forward kernel left them,
// x [slot][plane][64 utterances][KP]: C[m][n] = sum_{t < T, b} A[m][t*64 + b] x[slot0 + t][b][col0 + n].  No K-major copy of x is made
// (the three transposes of the LSTM inputs were 0.65 ms of the training step).  col0 and KP must be multiples of 8 elements.
struct XRows { const __nv_bfloat16* x; int KP, col0, slot0; };
int bulk_gemm_xrows(t2_ctx* c, const Opnd& A, const XRows& X, int T, int M, int N, const GemmNtOut& o, float* partial, cudaStream_t st) {
  GemmNtArgs a{};
  a.M = M; a.N = N; a.NT = (N >= 128 || N % 16) ? (N >= 128 ? 128 : ((N + 15) & ~15)) : N;
  a.Kin = A.Kin; a.Kout = A.Kout; a.nz = 1; a.npass = c->npass; a.mn = 2; a.KinB = D2_BP;
  T2_REQUIRE(A.Kout == 1 && A.Kin == T * D2_BP && X.col0 % 8 == 0 && X.KP % 8 == 0, T2_ERR_ARG, "bulk_gemm_xrows: operand shapes");
  a.splitk = 0;
  const long long need = gemmnt_partial_floats(a, c->num_sms);
  T2_REQUIRE(need <= BULK_PARTIAL_FLOATS, T2_ERR_WORKSPACE, "bulk_gemm_xrows: split-K slab of %lld floats exceeds the reserved %lld", need,
             (long long)BULK_PARTIAL_FLOATS);
  a.partial = partial;
  RC(gemmnt_make_map(&a.a_map, A.base, A.Kin, A.kin_ld, A.Kout, A.rows, A.plane, 128, c->npass));
  const long long slot = 2ll * D2_BP * X.KP;
  RC(gemmnt_make_map_cl_strided(&a.b_map, X.x + (size_t)X.slot0 * slot + X.col0, D2_BP, T, N, X.KP, slot, (long long)D2_BP * X.KP, c->npass));
  return gemmnt_launch(a, o, st);
}

// tmp[m][t*64 + b] (+ bias) -> mels[b][t][m], gates[b][t]; one CTA per frame, transposed through shared memory
__global__ void proj_finish_kernel(const float* __restrict__ tmp, const float* __restrict__ bias, int B, int T, int M, float* __restrict__ mels,
                                   float* __restrict__ gates) {
  extern __shared__ float tile[];   // [M+1][65]
  const int t = blockIdx.x;
  const long long N = (long long)T * D2_BP;
  for (int i = threadIdx.x; i < (M + 1) * D2_BP; i += blockDim.x) {
    const int m = i / D2_BP, b = i % D2_BP;
    tile[m * 65 + b] = tmp[(size_t)m * N + (size_t)t * D2_BP + b] + bias[m];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < B * M; i += blockDim.x) {
    const int b = i / M, m = i % M;
    mels[((size_t)b * T + t) * M + m] = tile[m * 65 + b];
  }
  for (int b = threadIdx.x; b < B; b += blockDim.x) gates[(size_t)b * T + t] = tile[M * 65 + b];
}

}  // namespace

int decode_proj_bulk(t2_ctx* c, const __nv_bfloat16* xdec, int B, int T, float* mels, float* gates, float* tmp, cudaStream_t st) {
  const int M = c->M, KdecP = c->dd.KdecP;
  const long long slot = 2ll * D2_BP * KdecP, plane = (long long)D2_BP * KdecP;
  GemmNtArgs a{};
  a.M = M + 1; a.N = T * D2_BP; a.NT = 64; a.nz = 1; a.splitk = 1; a.npass = c->npass;
  a.b_row_step = 2 * D2_BP;   // rows of one slot: 64 hi rows then 64 lo rows
  a.direct_out = tmp; a.direct_ldm = (long long)T * D2_BP;
  // segment 1: rnn_h_t = columns [projA_col0, KdecP) of slot t+1;  segment 2: [ctx_t | controls] = columns [H, H+KprojB) of slot t
  a.Kin = c->KprojA; a.Kout = 1;
  RC(gemmnt_make_map(&a.a_map, c->WprojA_b, c->KprojA, c->KprojA, 1, M + 1, (long long)(M + 1) * c->KprojA, 128, c->npass));
  RC(gemmnt_make_map(&a.b_map, xdec + slot + c->projA_col0, c->KprojA, KdecP, 1, T * 2 * D2_BP, plane, 64, c->npass));
  a.Kin2 = c->KprojB;
  RC(gemmnt_make_map(&a.a_map2, c->WprojB_b, c->KprojB, c->KprojB, 1, M + 1, (long long)(M + 1) * c->KprojB, 128, c->npass));
  RC(gemmnt_make_map(&a.b_map2, xdec + c->H, c->KprojB, KdecP, 1, T * 2 * D2_BP, plane, 64, c->npass));
  RC(gemmnt_launch(a, GemmNtOut{}, st));
  proj_finish_kernel<<<T, 256, (size_t)(M + 1) * 65 * sizeof(float), st>>>(tmp, c->d2w.bproj, B, T, M, mels, gates);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

namespace {

int decode_bwd_impl(t2_ctx* c, const float* memory, const float* pmem, const unsigned char* mask, const float* mel_target, int B, int L,
                    int T, const unsigned char* att_keep, const unsigned char* rnn_keep, unsigned long long seed, const float* aligns,
                    const float* d_mels, const float* d_gates, const DecTape& tp, float* d_memory, float* d_pmem, Arena& wa,
                    cudaStream_t st) {
  const D2Dims d = dims2(c, B, L);
  const int H = d.H, R = d.R, P = d.P, A = d.A, M = d.M, EF = d.EF, C = d.C;
  const int S = db_parts(d);
  const long long TB = (long long)T * D2_BP, TB1 = (long long)(T + 1) * D2_BP;
  const int Tp = round8(T);
  const int Nn = T * B, Np = round8(Nn);
  const int nsmall = (int)std::max(db_small_floats(d, true), db_small_floats(d, false));
  // ---- workspace
  void* scratch = wa.take<char>(bn_scratch_bytes(std::max(M, 128)));
  __nv_bfloat16* dP = wa.take<__nv_bfloat16>((size_t)2 * TB * PROJ_LD);
  __nv_bfloat16* dPT = wa.take<__nv_bfloat16>((size_t)2 * PROJ_LD * TB);
  float* dprojin = wa.take<float>((size_t)TB * d.KprojP);
  __nv_bfloat16* dg_att = wa.take<__nv_bfloat16>((size_t)2 * D2_BP * 4 * H);
  __nv_bfloat16* dg_dec = wa.take<__nv_bfloat16>((size_t)2 * D2_BP * 4 * R);
  __nv_bfloat16* dgT_att = wa.take<__nv_bfloat16>((size_t)2 * 4 * H * TB);
  __nv_bfloat16* dgT_dec = wa.take<__nv_bfloat16>((size_t)2 * 4 * R * TB);
  const int64_t zoff0 = wa.off;
  float* dx_att = wa.take<float>((size_t)c->KattT * D2_BP);
  float* dx_dec = wa.take<float>((size_t)c->KdecT * D2_BP);
  float* dctxT_dec = wa.take<float>((size_t)D2_BP * EF);
  float* dctxT_att = wa.take<float>((size_t)D2_BP * EF);
  float* dconv = wa.take<float>((size_t)2 * D2_BP * L);
  float* gcum = wa.take<float>((size_t)D2_BP * L);
  unsigned* barrier = wa.take<unsigned>(4);
  float* dbias = wa.take<float>((size_t)8 * H);
  const int64_t zoff1 = wa.off;
  float* dpre = wa.take<float>((size_t)TB * P);
  float* dctx_tot = wa.take<float>((size_t)TB * EF);
  float* dq = wa.take<float>((size_t)T * S * D2_BP * A);
  float* small_partial = wa.take<float>((size_t)128 * nsmall);
  float* small_sum = wa.take<float>(nsmall);
  __nv_bfloat16* xT_att = wa.take<__nv_bfloat16>((size_t)2 * d.KattP * TB1);
  __nv_bfloat16* xT_dec = wa.take<__nv_bfloat16>((size_t)2 * d.KdecP * TB1);
  __nv_bfloat16* dqT = wa.take<__nv_bfloat16>((size_t)2 * A * TB);
  __nv_bfloat16* alT = wa.take<__nv_bfloat16>((size_t)2 * B * L * Tp);
  __nv_bfloat16* dcT = wa.take<__nv_bfloat16>((size_t)2 * B * EF * Tp);
  float* d2 = wa.take<float>((size_t)Nn * P);
  float* dp1 = wa.take<float>((size_t)Nn * P);
  __nv_bfloat16* d2b = wa.take<__nv_bfloat16>((size_t)2 * Nn * P);
  __nv_bfloat16* d2T = wa.take<__nv_bfloat16>((size_t)2 * P * Np);
  __nv_bfloat16* p1T = wa.take<__nv_bfloat16>((size_t)2 * P * Np);
  __nv_bfloat16* xT = wa.take<__nv_bfloat16>((size_t)2 * M * Np);
  float* wtmp = wa.take<float>((size_t)(M + 1) * d.KprojP);
  __nv_bfloat16* dgH = wa.take<__nv_bfloat16>((size_t)2 * TB * 4 * H);
  float* partial = wa.take<float>(BULK_PARTIAL_FLOATS);
  if (!wa.base) return T2_OK;
  T2_REQUIRE(!wa.overflow, T2_ERR_WORKSPACE, "t2_decode_tf_bwd: workspace too small");

  NEED_PARAM(wq, "decoder.attention.query_layer.weight", (size_t)A * H);
  NEED_PARAM(wv, "decoder.attention.v.weight", A);
  NEED_GRAD(g_att_ih, "decoder.att_rnn.weight_ih", (size_t)4 * H * (P + EF));
  NEED_GRAD(g_att_hh, "decoder.att_rnn.weight_hh", (size_t)4 * H * H);
  NEED_GRAD(g_att_bi, "decoder.att_rnn.bias_ih", 4 * H);
  NEED_GRAD(g_att_bh, "decoder.att_rnn.bias_hh", 4 * H);
  NEED_GRAD(g_dec_ih, "decoder.lstm.weight_ih", (size_t)4 * R * (H + EF + C));
  NEED_GRAD(g_dec_hh, "decoder.lstm.weight_hh", (size_t)4 * R * R);
  NEED_GRAD(g_dec_bi, "decoder.lstm.bias_ih", 4 * R);
  NEED_GRAD(g_dec_bh, "decoder.lstm.bias_hh", 4 * R);
  NEED_GRAD(g_wq, "decoder.attention.query_layer.weight", (size_t)A * H);
  NEED_GRAD(g_v, "decoder.attention.v.weight", A);
  NEED_GRAD(g_lconv, "decoder.attention.location_conv.weight", 32 * 2 * 31);
  NEED_GRAD(g_ldense, "decoder.attention.location_dense.weight", (size_t)A * 32);
  NEED_GRAD(g_wmel, "decoder.mel_out.weight", (size_t)M * (R + EF + C));
  NEED_GRAD(g_bmel, "decoder.mel_out.bias", M);
  NEED_GRAD(g_wgate, "decoder.gate.weight", R + EF);
  NEED_GRAD(g_bgate, "decoder.gate.bias", 1);
  NEED_GRAD(g_w1, "prenet.0.weight", (size_t)P * M);
  NEED_GRAD(g_w2, "prenet.3.weight", (size_t)P * P);

  // ---- projection backward (bulk): d[rnn_h | ctx | controls] for every frame; bias gradients
  RC(build_dproj_launch(d_mels, d_gates, B, T, M, PROJ_LD, dP, dPT, st));
  {
    Opnd Aop{dP, M + 1, PROJ_LD, 1, (int)TB, TB * PROJ_LD};
    Opnd Bop{c->WprojT_b, M + 1, PROJ_LD, 1, d.KprojP, (long long)d.KprojP * PROJ_LD};
    RC(bulk_gemm(c, Aop, Bop, (int)TB, d.KprojP, 0, GemmNtOut{dprojin, d.KprojP, 1, 0, 1.f, 0}, nullptr, st, 1, 0, 0, true));
  }
  RC(colsum_launch(d_mels, (long long)B * T, M, M, g_bmel, scratch, st));
  RC(colsum_launch(d_gates, (long long)B * T, 1, 1, g_bgate, scratch, st));

  // ---- the sequential part
  T2_CHECK_CUDA(cudaMemsetAsync(wa.base + zoff0, 0, (size_t)(zoff1 - zoff0), st));
  T2_CHECK_CUDA(cudaMemsetAsync(d_pmem, 0, (size_t)B * L * A * sizeof(float), st));
  DBWeights w{wq->ptr, wv->ptr, c->loc_conv_t, c->loc_dense_t};
  DBTape t{};
  t.sav_att = tp.sav_att; t.sav_dec = tp.sav_dec; t.q = tp.q; t.w_hist = tp.w_hist; t.wcum_hist = tp.wcum_hist;
  // same decision as the forward that wrote the tape: the pipelined forward (decoder_ar.cu) tapes the location term only
  if (d2ar_supported(d)) { t.xloc = tp.xloc; t.floc = nullptr; }
  else if (d2_precompute_fits(d)) { t.xloc = tp.xloc; t.floc = tp.floc; }
  t.memory = memory; t.pmem = pmem; t.mask = mask; t.att_keep = att_keep; t.rnn_keep = rnn_keep; t.seed = seed; t.hidden_dropout = 1;
  DBWork k{};
  k.dprojin = dprojin; k.dg_att = dg_att; k.dg_dec = dg_dec; k.dgT_att = dgT_att; k.dgT_dec = dgT_dec; k.dx_att = dx_att; k.dx_dec = dx_dec; k.dctxT_dec = dctxT_dec; k.dctxT_att = dctxT_att;
  k.dpre = dpre; k.dctx_tot = dctx_tot; k.dq = dq; k.dpm = d_pmem; k.dconv = dconv; k.gcum = gcum; k.small_partial = small_partial;
  k.dbias = dbias; k.barrier = barrier; k.status = c->status_dev;
  DBMaps maps;
  maps.wt_att = c->map_wt_att; maps.wt_dec = c->map_wt_dec;
  {
    const unsigned zb = c->npass == 3 ? 2u : 1u;
    unsigned box[3] = {TG_BK, TG_N, zb};
    unsigned long long dims[3] = {4ull * H, (unsigned long long)D2_BP, 2ull};
    unsigned long long sb[2] = {4ull * H * 2, 4ull * H * D2_BP * 2};
    RC(tc_make_tmap_bf16(&maps.dg_att, dg_att, 3, dims, sb, box));
    RC(tc_make_tmap_bf16(&maps.dg_dec, dg_dec, 3, dims, sb, box));
  }
  T2_CHECK_CUDA(cudaMemsetAsync(c->prof_buf, 0, 16 * sizeof(unsigned long long), st));
  timing_begin(c, 1, st);
  RC(db_launch(d, T, c->KattT, c->KdecT, w, t, k, maps, c->prof_buf, st));
  timing_end(c, 1, st);

  // ---- small attention weights, biases
  if (db_tc_att(d, t.xloc != nullptr)) {
    const int ns = (int)db_small_floats(d, true);
    RC(partial_sum_launch(small_partial, 128, ns, small_sum, st));
    RC(att_small_finish_launch(small_sum, c->loc_conv_t, c->loc_dense_t, A, g_ldense, g_lconv, st));
    RC(copy2_launch(small_sum + A * 64, g_v, nullptr, A, st));
  } else {
    const int ns = (int)db_small_floats(d, false);
    RC(partial_sum_launch(small_partial, 128, ns, small_sum, st));
    RC(copy2_launch(small_sum, g_ldense, nullptr, A * 32, st));
    RC(copy2_launch(small_sum + A * 32, g_lconv, nullptr, 2 * 31 * 32, st));
    RC(copy2_launch(small_sum + A * 32 + 2 * 31 * 32, g_v, nullptr, A, st));
  }
  RC(copy2_launch(dbias, g_att_bi, g_att_bh, 4 * H, st));
  RC(copy2_launch(dbias + 4 * H, g_dec_bi, g_dec_bh, 4 * R, st));

  // The weight gradients that contract d gates / d query / d projection with the per-frame operand rows read those rows MN-major as the
  // forward left them ([slot][plane][utterance][k]) when every column window starts on a 16-byte boundary; otherwise (controls of odd
  // width) they go through K-major copies xT[k][slot*64 + b].
  const bool x_mn = P % 8 == 0 && (P + EF) % 8 == 0 && H % 8 == 0 && (H + EF + C) % 8 == 0 && d.KattP % 8 == 0 && d.KdecP % 8 == 0 &&
                    !getenv("T2B200_WGRAD_KMAJOR");
  const Opnd Ga{dgT_att, (int)TB, (int)TB, 1, 4 * H, 4ll * H * TB}, Gd{dgT_dec, (int)TB, (int)TB, 1, 4 * R, 4ll * R * TB};
  const Opnd Q{dqT, (int)TB, (int)TB, 1, A, (long long)A * TB}, Pm{dPT, (int)TB, (int)TB, 1, M + 1, (long long)PROJ_LD * TB};
  RC(dq_sumT_launch(dq, T, S, db_tc_att(d, t.xloc != nullptr) ? 1 : S, A, B, dqT, st));
  if (x_mn) {
    const XRows Xpre{tp.tfp, P, 0, 0};   // teacher forcing: the prenet block of the attention-cell input lives in tfp
    auto xa = [&](int col0) { return XRows{tp.xatt, d.KattP, col0, 0}; };
    auto xd = [&](int col0, int slot0 = 0) { return XRows{tp.xdec, d.KdecP, col0, slot0}; };
    // LSTMCell weight gradients  dW[r][k] = sum_{t,b} dgates[t,b][r] x[t,b][k]
    RC(bulk_gemm_xrows(c, Ga, Xpre, T, 4 * H, P, GemmNtOut{g_att_ih, P + EF, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm_xrows(c, Ga, xa(P), T, 4 * H, EF, GemmNtOut{g_att_ih + P, P + EF, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm_xrows(c, Ga, xa(P + EF), T, 4 * H, H, GemmNtOut{g_att_hh, H, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm_xrows(c, Gd, xd(0), T, 4 * R, H + EF + C, GemmNtOut{g_dec_ih, H + EF + C, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm_xrows(c, Gd, xd(H + EF + C), T, 4 * R, R, GemmNtOut{g_dec_hh, R, 1, 0, 1.f, 0}, partial, st));
    // query layer: dWq[a][h] = sum dq[t,b][a] att_h_t[b][h]   (att_h_t = first block of the decoder-cell input of frame t)
    RC(bulk_gemm_xrows(c, Q, xd(0), T, A, H, GemmNtOut{g_wq, H, 1, 0, 1.f, 0}, partial, st));
    // mel / gate projection weights: input [rnn_h_t | ctx_t | controls]; rnn_h_t sits in slot t+1 of the decoder-cell input
    RC(bulk_gemm_xrows(c, Pm, xd(H + EF + C, 1), T, M + 1, R, GemmNtOut{wtmp, d.KprojP, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm_xrows(c, Pm, xd(H), T, M + 1, EF + C, GemmNtOut{wtmp + R, d.KprojP, 1, 0, 1.f, 0}, partial, st));
  } else {
    RC(transpose_bf16_launch(tp.xatt, d.KattP, (long long)2 * D2_BP * d.KattP, (long long)D2_BP * d.KattP, xT_att, D2_BP, d.KattP, D2_BP, TB1,
                             D2_BP, T + 1, 0, (long long)d.KattP * TB1, st));
    RC(transpose_bf16_launch(tp.tfp, P, (long long)2 * D2_BP * P, (long long)D2_BP * P, xT_att, D2_BP, P, D2_BP, TB1, D2_BP, T + 1, 0,
                             (long long)d.KattP * TB1, st));   // teacher forcing: the prenet block of the input lives in tfp
    RC(transpose_bf16_launch(tp.xdec, d.KdecP, (long long)2 * D2_BP * d.KdecP, (long long)D2_BP * d.KdecP, xT_dec, D2_BP, d.KdecP, D2_BP, TB1,
                             D2_BP, T + 1, 0, (long long)d.KdecP * TB1, st));
    auto xrows = [&](const __nv_bfloat16* xTb, int KP, int row0, int rows, bool all_slots = false) {
      return Opnd{xTb + (size_t)row0 * TB1, (int)(all_slots ? TB1 : TB), (int)TB1, 1, rows, (long long)KP * TB1};
    };
    RC(bulk_gemm(c, Ga, xrows(xT_att, d.KattP, 0, P + EF), 4 * H, P + EF, 0, GemmNtOut{g_att_ih, P + EF, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm(c, Ga, xrows(xT_att, d.KattP, P + EF, H), 4 * H, H, 0, GemmNtOut{g_att_hh, H, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm(c, Gd, xrows(xT_dec, d.KdecP, 0, H + EF + C), 4 * R, H + EF + C, 0, GemmNtOut{g_dec_ih, H + EF + C, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm(c, Gd, xrows(xT_dec, d.KdecP, H + EF + C, R), 4 * R, R, 0, GemmNtOut{g_dec_hh, R, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm(c, Q, xrows(xT_dec, d.KdecP, 0, H), A, H, 0, GemmNtOut{g_wq, H, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm(c, Pm, xrows(xT_dec, d.KdecP, H + EF + C, R, true), M + 1, R, D2_BP, GemmNtOut{wtmp, d.KprojP, 1, 0, 1.f, 0}, partial, st));
    RC(bulk_gemm(c, Pm, xrows(xT_dec, d.KdecP, H, EF + C), M + 1, EF + C, 0, GemmNtOut{wtmp + R, d.KprojP, 1, 0, 1.f, 0}, partial, st));
  }
  T2_CHECK_CUDA(cudaMemcpy2DAsync(g_wmel, (size_t)(R + EF + C) * 4, wtmp, (size_t)d.KprojP * 4, (size_t)(R + EF + C) * 4, M,
                                  cudaMemcpyDeviceToDevice, st));
  T2_CHECK_CUDA(cudaMemcpyAsync(g_wgate, wtmp + (size_t)M * d.KprojP, (size_t)(R + EF) * 4, cudaMemcpyDeviceToDevice, st));
  {  // attention memory: d memory[b][l][c] = sum_t w_t[b][l] dctx_t[b][c]   (one GEMM per utterance, batched over z = b)
    RC(transpose_split_launch(aligns, L, (long long)T * L, alT, T, L, Tp, Tp, (long long)L * Tp, B, (long long)B * L * Tp, st));
    RC(transpose_split_launch(dctx_tot, (long long)D2_BP * EF, EF, dcT, T, EF, Tp, Tp, (long long)EF * Tp, B, (long long)B * EF * Tp, st));
    Opnd Aw{alT, T, Tp, 1, B * L, (long long)B * L * Tp}, Dc{dcT, T, Tp, 1, B * EF, (long long)B * EF * Tp};
    RC(bulk_gemm(c, Aw, Dc, L, EF, 0, GemmNtOut{d_memory, EF, 1, (long long)L * EF, 1.f, 0}, partial, st, B, L, EF));
  }
  {  // d prenet output [t][b][k] = sum_r dgates_att[t,b][r] W_ih[r][k], k < P: off the recurrent chain, one bulk GEMM
    // A = the d gates history dgT_att [4H][T*64] read MN-major (rows of the product = its contiguous index): no [T*64][4H] copy
    if (!getenv("T2B200_WGRAD_KMAJOR")) {
      GemmNtArgs a{};
      a.M = (int)TB; a.N = P; a.NT = P >= 128 ? 128 : ((P + 15) & ~15);
      a.Kin = 4 * H; a.Kout = 1; a.nz = 1; a.npass = c->npass; a.mn = 1; a.splitk = 1; a.direct_out = dpre; a.direct_ldm = P;
      RC(gemmnt_make_map_cl(&a.a_map, dgT_att, 4 * H, 1, (int)TB, 4ll * H * TB, c->npass));
      RC(gemmnt_make_map(&a.b_map, c->WpreT_b, 4 * H, 4 * H, 1, P, (long long)P * 4 * H, a.NT, c->npass));
      RC(gemmnt_launch(a, GemmNtOut{dpre, P, 1, 0, 1.f, 0}, st));
    } else {
      RC(transpose_bf16_launch(dgT_att, TB, 0, 4ll * H * TB, dgH, 4 * H, (int)TB, 4 * H, 4 * H, 0, 1, 0, TB * 4 * H, st));
      Opnd Gh{dgH, 4 * H, 4 * H, 1, (int)TB, TB * 4 * H}, Wp{c->WpreT_b, 4 * H, 4 * H, 1, P, (long long)P * 4 * H};
      RC(bulk_gemm(c, Gh, Wp, (int)TB, P, 0, GemmNtOut{dpre, P, 1, 0, 1.f, 0}, nullptr, st, 1, 0, 0, true));
    }
  }
  {  // prenet backward (model/tacotron2.py:85-92), rows n = t*B + b, t < T (the T+1-th prenet frame is never used, :317)
    const float scale = 1.0f / (1.0f - c->cfg.dropout);
    RC(prenet_d2_launch(dpre, tp.tfp, B, T, P, scale, d2, st));
    RC(split_cl_launch(d2, (long long)Nn * P, d2b, 1, Nn, P, st));
    RC(transpose_split_launch(d2, P, 0, d2T, Nn, P, Np, Np, 0, 1, (long long)P * Np, st));
    RC(transpose_split_launch(tp.p1, P, 0, p1T, Nn, P, Np, Np, 0, 1, (long long)P * Np, st));
    Opnd D2T{d2T, Nn, Np, 1, P, (long long)P * Np}, P1T{p1T, Nn, Np, 1, P, (long long)P * Np};
    RC(bulk_gemm(c, D2T, P1T, P, P, 0, GemmNtOut{g_w2, P, 1, 0, 1.f, 0}, partial, st));
    Opnd D2{d2b, P, P, 1, Nn, (long long)Nn * P}, W2T{c->W2T_b, P, P, 1, P, (long long)P * P};
    RC(bulk_gemm(c, D2, W2T, Nn, P, 0, GemmNtOut{dp1, P, 1, 0, 1.f, 0}, nullptr, st, 1, 0, 0, true));
    RC(prenet_d1_launch(dp1, tp.p1, (long long)Nn * P, scale, d2, st));   // d2 buffer reused for d1
    RC(transpose_split_launch(d2, P, 0, d2T, Nn, P, Np, Np, 0, 1, (long long)P * Np, st));
    RC(prenet_xT_launch(mel_target, B, T, M, Np, xT, st));
    Opnd XT{xT, Nn, Np, 1, M, (long long)M * Np};
    RC(bulk_gemm(c, D2T, XT, P, M, 0, GemmNtOut{g_w1, M, 1, 0, 1.f, 0}, partial, st));
  }
  // an aborted backward kernel must not pass for a result: the gradient norm downstream turns NaN (and the optimizer skips the step)
  RC(poison_on_abort_launch(c->status_dev, d_memory, (long long)B * L * EF, st));
  RC(poison_on_abort_launch(c->status_dev, g_dec_hh, (long long)4 * R * R, st));
  return T2_OK;
}

}  // namespace

// Backward of the teacher-forced frame loop.  d_mels (B,T,M) / d_gates (B,T): gradients w.r.t. the raw decoder outputs;
// alignments (B,T,L) as returned by the forward.  Outputs: d_memory (B,L,EF), d_processed_memory (B,L,A); decoder.* and
// prenet.* parameter gradients go to their registered destinations.
extern "C" int t2_decode_tf_bwd(t2_ctx* c, const float* memory, const float* processed_memory, const unsigned char* mask,
                                const float* mel_target, int B, int L, int T, const unsigned char* att_keep, const unsigned char* rnn_keep,
                                unsigned long long seed, const float* alignments, const float* d_mels, const float* d_gates, const void* tape,
                                long long tape_bytes, float* d_memory, float* d_processed_memory, void* ws, long long ws_bytes,
                                void* stream) {
  CHECK_CTX("t2_decode_tf_bwd");
  T2_REQUIRE(memory && processed_memory && mask && mel_target && alignments && d_mels && d_gates && tape && d_memory && d_processed_memory && ws,
             T2_ERR_ARG, "t2_decode_tf_bwd: null argument");
  T2_REQUIRE(B > 0 && B <= D2_BP && L > 0 && T > 0, T2_ERR_ARG, "t2_decode_tf_bwd: bad sizes");
  Arena ta(const_cast<void*>(tape), tape_bytes);
  DecTape tp = plan_dec_tape(dims2(c, B, L), B, L, T, ta);
  T2_REQUIRE(!ta.overflow, T2_ERR_WORKSPACE, "t2_decode_tf_bwd: tape too small");
  Arena wa(ws, ws_bytes);
  return decode_bwd_impl(c, memory, processed_memory, mask, mel_target, B, L, T, att_keep, rnn_keep, seed, alignments, d_mels, d_gates, tp,
                         d_memory, d_processed_memory, wa, (cudaStream_t)stream);
}

long long decode_bwd_ws_bytes(t2_ctx* c, int B, int L, int T) {
  Arena wa(nullptr, 0);
  DecTape tp{};
  decode_bwd_impl(c, nullptr, nullptr, nullptr, nullptr, std::min(B, D2_BP), L, T, nullptr, nullptr, 0, nullptr, nullptr, nullptr, tp, nullptr,
                  nullptr, wa, nullptr);
  return wa.off;
}

// ------------------------------------------------------------------------------------------ encoder (train fwd / bwd)
namespace {

struct EncTape { __nv_bfloat16* x[4]; float* z[3]; float* mean[3]; float* rstd[3]; float* lstm_sav; };
EncTape plan_enc_tape(const t2_ctx* c, int B, int L, Arena& ar) {
  EncTape t{};
  const int E = c->E, Hh = E / 2;
  for (int i = 0; i < 4; ++i) t.x[i] = ar.take<__nv_bfloat16>((size_t)2 * B * L * E);
  for (int i = 0; i < 3; ++i) {
    t.z[i] = ar.take<float>((size_t)B * L * E);
    t.mean[i] = ar.take<float>(E);
    t.rstd[i] = ar.take<float>(E);
  }
  t.lstm_sav = ar.take<float>((size_t)B * L * 2 * 5 * Hh);
  return t;
}
BnLayer enc_layer(const t2_ctx* c, const EncTape& tp, int i, int B, int L, const unsigned char* keep, unsigned long long seed,
                  const float* gamma, const float* beta) {
  BnLayer Ly{};
  Ly.z = tp.z[i]; Ly.mean = tp.mean[i]; Ly.rstd = tp.rstd[i]; Ly.gamma = gamma; Ly.beta = beta;
  Ly.C = c->E; Ly.act = ACT_RELU; Ly.dropout = 1;
  Ly.keep = keep ? keep + (size_t)i * B * L * c->E : nullptr; Ly.keep_ld = c->E; Ly.p = c->cfg.dropout;
  Ly.seed = seed; Ly.rng_base = (0x60ull + i) << 40;
  return Ly;
}

int encoder_bwd_impl(t2_ctx* c, const long long* chars, const long long* lens, int B, int L, const unsigned char* keep,
                     unsigned long long seed, const float* encoded, const float* d_encoded, const EncTape& tp, Arena& wa, cudaStream_t st) {
  const int E = c->E, Hh = E / 2, KS = c->KS, Lp = round8(L), G8 = 8 * Hh;
  void* scratch = wa.take<char>(bn_scratch_bytes(G8));
  float* dgates = wa.take<float>((size_t)B * L * G8);
  __nv_bfloat16* dgb = wa.take<__nv_bfloat16>((size_t)2 * B * L * G8);
  __nv_bfloat16* dgT = wa.take<__nv_bfloat16>((size_t)2 * G8 * B * Lp);
  __nv_bfloat16* hT = wa.take<__nv_bfloat16>((size_t)2 * 2 * Hh * B * Lp);     // per direction: shifted h_{prev}, K-major
  __nv_bfloat16* xT = wa.take<__nv_bfloat16>((size_t)2 * E * B * Lp);
  float* wtmp = wa.take<float>((size_t)G8 * E);
  float* btmp = wa.take<float>(G8);
  float* dyA = wa.take<float>((size_t)B * L * E);
  float* dyB = wa.take<float>((size_t)B * L * E);
  float* dzf = wa.take<float>((size_t)B * L * E);
  __nv_bfloat16* dz = wa.take<__nv_bfloat16>((size_t)2 * B * L * E);
  __nv_bfloat16* dzT = wa.take<__nv_bfloat16>((size_t)2 * E * B * Lp);
  float* partial = wa.take<float>(BULK_PARTIAL_FLOATS);
  if (!wa.base) return T2_OK;
  T2_REQUIRE(!wa.overflow, T2_ERR_WORKSPACE, "t2_encoder_bwd: workspace too small");

  // ---- packed BiLSTM (model/encoder.py:59-65)
  RC(bilstm_bwd_launch(tp.lstm_sav, c->lstm_Whh, lens, d_encoded, dgates, B, L, Hh, st));
  RC(colsum_launch(dgates, (long long)B * L, G8, G8, btmp, scratch, st));
  RC(transpose_split_launch(dgates, G8, (long long)L * G8, dgT, L, G8, Lp, (long long)B * Lp, Lp, B, (long long)G8 * B * Lp, st));
  RC(to_kmajor(tp.x[3], xT, B, L, Lp, E, st));
  {  // W_ih of both directions in one GEMM: rows [dir*4Hh, +4Hh)
    Opnd G{dgT, L, Lp, B, G8, (long long)G8 * B * Lp}, X{xT, L, Lp, B, E, (long long)E * B * Lp};
    RC(bulk_gemm(c, G, X, G8, E, 0, GemmNtOut{wtmp, E, 1, 0, 1.f, 0}, partial, st));
  }
  for (int dir = 0; dir < 2; ++dir) {
    const std::string sfx = dir ? "_reverse" : "";
    NEED_GRAD(g_ih, "encoder.lstm.weight_ih_l0" + sfx, (size_t)4 * Hh * E);
    NEED_GRAD(g_hh, "encoder.lstm.weight_hh_l0" + sfx, (size_t)4 * Hh * Hh);
    NEED_GRAD(g_bi, "encoder.lstm.bias_ih_l0" + sfx, 4 * Hh);
    NEED_GRAD(g_bh, "encoder.lstm.bias_hh_l0" + sfx, 4 * Hh);
    T2_CHECK_CUDA(cudaMemcpyAsync(g_ih, wtmp + (size_t)dir * 4 * Hh * E, (size_t)4 * Hh * E * 4, cudaMemcpyDeviceToDevice, st));
    RC(copy2_launch(btmp + dir * 4 * Hh, g_bi, g_bh, 4 * Hh, st));
    // recurrent weights: h_{prev}[b][t] = out[b][t-1] (forward) / out[b][t+1] (reverse), zero at the sequence ends
    __nv_bfloat16* hTd = hT + (size_t)dir * 2 * Hh * B * Lp;
    RC(transpose_split_launch(encoded + dir * Hh, 2 * Hh, (long long)L * 2 * Hh, hTd, L, Hh, Lp, (long long)B * Lp, Lp, B,
                              (long long)Hh * B * Lp, st, dir ? 1 : -1));
    Opnd G{dgT + (size_t)dir * 4 * Hh * B * Lp, L, Lp, B, 4 * Hh, (long long)G8 * B * Lp}, Hp{hTd, L, Lp, B, Hh, (long long)Hh * B * Lp};
    RC(bulk_gemm(c, G, Hp, 4 * Hh, Hh, 0, GemmNtOut{g_hh, Hh, 1, 0, 1.f, 0}, partial, st));
  }
  // d (conv stack output) = dgates W_ih  (k = 1 "convolution" with the transposed pack)
  RC(split_cl_launch(dgates, (long long)L * G8, dgb, B, L, G8, st));
  RC(conv_tc_res(c, dgb, c->lstm_wtmap, B, L, G8, E, 1, 128, nullptr, 0, dyA, st));

  // ---- 3 x [Conv1d k5 + BatchNorm + ReLU + Dropout] (model/encoder.py:31-46)
  const float* dy = dyA;
  float* bufs[2] = {dyB, dyA};
  for (int i = 2; i >= 0; --i) {
    const std::string cv = "encoder.convolutions." + std::to_string(4 * i), bn = "encoder.convolutions." + std::to_string(4 * i + 1);
    NEED_PARAM(bw, bn + ".weight", E);
    NEED_PARAM(bb, bn + ".bias", E);
    NEED_GRAD(gW, cv + ".weight", (size_t)E * E * KS);
    NEED_GRAD(gB, cv + ".bias", E);
    NEED_GRAD(gg, bn + ".weight", E);
    NEED_GRAD(gb, bn + ".bias", E);
    BnLayer Ly = enc_layer(c, tp, i, B, L, keep, seed, bw->ptr, bb->ptr);
    RC(bn_bwd_launch(Ly, dy, (long long)B * L, gg, gb, dz, dzf, scratch, st));
    RC(colsum_launch(dzf, (long long)B * L, E, E, gB, scratch, st));   // conv bias (mathematically 0 through BatchNorm; kept as computed)
    RC(conv_wgrad(c, dz, tp.x[i], B, L, E, E, KS, gW, partial, BULK_PARTIAL_FLOATS, st));
    float* dst = bufs[i & 1];
    RC(conv_tc_res(c, dz, c->enc_wtmap[i], B, L, E, E, KS, 128, nullptr, 0, dst, st));
    dy = dst;
  }
  NEED_GRAD(gE, "encoder.embedding.weight", (size_t)(c->cfg.num_chars + 1) * E);
  RC(embedding_grad_launch(dy, chars, (long long)B * L, E, c->cfg.num_chars + 1, gE, st));
  return T2_OK;
}

}  // namespace

extern "C" long long t2_encoder_tape_bytes(const t2_ctx* c, int B, int L) {
  if (!c || B <= 0 || L <= 0) return -1;
  Arena ar(nullptr, 0);
  plan_enc_tape(c, B, L, ar);
  return ar.off + 1024;
}

// Encoder.forward under train() (model/encoder.py:54-67): BatchNorm batch statistics over all (b, l) incl. padding,
// running statistics updated in place, dropout after each conv block.  keep (3,B,L,E) or NULL.
extern "C" int t2_encoder_train_fwd(t2_ctx* c, const long long* chars_idx, const long long* lens, int B, int L, const unsigned char* keep,
                                    unsigned long long seed, float* encoded, void* tape, long long tape_bytes, void* ws, long long ws_bytes,
                                    void* stream) {
  CHECK_CTX("t2_encoder_train_fwd");
  T2_REQUIRE(chars_idx && lens && encoded && tape && ws && B > 0 && L > 0, T2_ERR_ARG, "t2_encoder_train_fwd: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const int E = c->E, Hh = E / 2;
  Arena ta(tape, tape_bytes);
  EncTape tp = plan_enc_tape(c, B, L, ta);
  T2_REQUIRE(!ta.overflow, T2_ERR_WORKSPACE, "t2_encoder_train_fwd: tape too small");
  Arena wa(ws, ws_bytes);
  void* scratch = wa.take<char>(bn_scratch_bytes(E));
  float* xproj = wa.take<float>((size_t)B * L * 8 * Hh);
  T2_REQUIRE(!wa.overflow, T2_ERR_WORKSPACE, "t2_encoder_train_fwd: workspace too small");
  NEED_PARAM(emb, "encoder.embedding.weight", (size_t)(c->cfg.num_chars + 1) * E);
  RC(embed_split_launch(chars_idx, emb->ptr, tp.x[0], B * L, E, c->cfg.num_chars + 1, st));
  for (int i = 0; i < 3; ++i) {
    const std::string cv = "encoder.convolutions." + std::to_string(4 * i), bn = "encoder.convolutions." + std::to_string(4 * i + 1);
    NEED_PARAM(cb, cv + ".bias", E);
    NEED_PARAM(bw, bn + ".weight", E);
    NEED_PARAM(bb, bn + ".bias", E);
    NEED_PARAM(bm, bn + ".running_mean", E);
    NEED_PARAM(bv, bn + ".running_var", E);
    RC(conv_tc(c, tp.x[i], c->enc_wmap[i], B, L, E, E, c->KS, 128, ACT_NONE, nullptr, cb->ptr, nullptr, 0, tp.z[i], nullptr, st));
    RC(bn_stats_launch(tp.z[i], (long long)B * L, E, tp.mean[i], tp.rstd[i], const_cast<float*>(bm->ptr), const_cast<float*>(bv->ptr), scratch, st));
    BnLayer Ly = enc_layer(c, tp, i, B, L, keep, seed, bw->ptr, bb->ptr);
    RC(bn_apply_launch(Ly, (long long)B * L, nullptr, tp.x[i + 1], st));
  }
  RC(conv_tc(c, tp.x[3], c->lstm_wmap, B, L, E, 8 * Hh, 1, 128, ACT_NONE, nullptr, c->lstm_bias, nullptr, 0, xproj, nullptr, st));
  RC(bilstm_launch(xproj, c->lstm_Whh, lens, encoded, B, L, Hh, st, tp.lstm_sav));
  return T2_OK;
}

// Backward of Encoder.forward: d_encoded (B,L,E) -> encoder.* parameter gradients.  `encoded` is the forward's output.
extern "C" int t2_encoder_bwd(t2_ctx* c, const long long* chars_idx, const long long* lens, int B, int L, const unsigned char* keep,
                              unsigned long long seed, const float* encoded, const float* d_encoded, const void* tape, long long tape_bytes,
                              void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_encoder_bwd");
  T2_REQUIRE(chars_idx && lens && encoded && d_encoded && tape && ws && B > 0 && L > 0, T2_ERR_ARG, "t2_encoder_bwd: bad argument");
  Arena ta(const_cast<void*>(tape), tape_bytes);
  EncTape tp = plan_enc_tape(c, B, L, ta);
  T2_REQUIRE(!ta.overflow, T2_ERR_WORKSPACE, "t2_encoder_bwd: tape too small");
  Arena wa(ws, ws_bytes);
  return encoder_bwd_impl(c, chars_idx, lens, B, L, keep, seed, encoded, d_encoded, tp, wa, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ conditioning + att_encoder backward
static int memory_bwd_impl(t2_ctx* c, const float* memory, const long long* spk, const float* desc, const float* d_memory, const float* d_pmem,
                           int B, int L, float* d_encoded, Arena& wa, cudaStream_t st) {
  const int E = c->E, EF = c->EF, A = c->A, D = EF - E, Lp = round8(L);
  __nv_bfloat16* dpb = wa.take<__nv_bfloat16>((size_t)2 * B * L * A);
  __nv_bfloat16* dpT = wa.take<__nv_bfloat16>((size_t)2 * A * B * Lp);
  __nv_bfloat16* memT = wa.take<__nv_bfloat16>((size_t)2 * EF * B * Lp);
  float* dmem = wa.take<float>((size_t)B * L * EF);
  float* dpre = wa.take<float>((size_t)B * 128);
  float* partial = wa.take<float>(BULK_PARTIAL_FLOATS);
  if (!wa.base) return T2_OK;
  T2_REQUIRE(!wa.overflow, T2_ERR_WORKSPACE, "t2_memory_bwd: workspace too small");
  NEED_GRAD(g_ae, "att_encoder.weight", (size_t)A * EF);
  // d memory (total) = d_memory + d_pmem W_ae ; d W_ae[a][c] = sum_{b,l} d_pmem[b,l][a] memory[b,l][c]   (model/tacotron2.py:229)
  RC(split_cl_launch(d_pmem, (long long)L * A, dpb, B, L, A, st));
  RC(conv_tc_res(c, dpb, c->attenc_wtmap, B, L, A, EF, 1, 128, d_memory, (long long)L * EF, dmem, st));
  RC(to_kmajor(dpb, dpT, B, L, Lp, A, st));
  RC(transpose_split_launch(memory, EF, (long long)L * EF, memT, L, EF, Lp, (long long)B * Lp, Lp, B, (long long)EF * B * Lp, st));
  {
    Opnd Pm{dpT, L, Lp, B, A, (long long)A * B * Lp}, Mm{memT, L, Lp, B, EF, (long long)EF * B * Lp};
    RC(bulk_gemm(c, Pm, Mm, A, EF, 0, GemmNtOut{g_ae, EF, 1, 0, 1.f, 0}, partial, st));
  }
  // conditioning (model/tacotron2.py:201-212)
  const bool speaker = c->cfg.num_speakers > 0;
  RC(condition_bwd_launch(dmem, memory, speaker ? 1 : 0, d_encoded, (long long)B * L, E, EF, st));
  if (speaker) {
    T2_REQUIRE(spk, T2_ERR_ARG, "speaker_id tensor required when speaker tokens are active!");
    NEED_GRAD(g_spk, "speaker_embedding.weight", (size_t)c->cfg.num_speakers * E);
    RC(speaker_grad_launch(d_encoded, spk, B, L, E, c->cfg.num_speakers, g_spk, partial /* free again: the GEMM above has reduced it */, st));
  }
  if (D > 0) {
    T2_REQUIRE(desc, T2_ERR_ARG, "description tensor required when description tokens are active!");
    NEED_GRAD(g_dw, "description_embeddings_linear.0.weight", (size_t)D * c->cfg.description_dim);
    NEED_GRAD(g_db, "description_embeddings_linear.0.bias", D);
    RC(desc_bwd_launch(dmem, memory, desc, B, L, E, D, c->cfg.description_dim, dpre, g_dw, g_db, st));
  }
  return T2_OK;
}

// Backward of t2_memory_fwd: d_memory (B,L,EF), d_processed_memory (B,L,A) -> d_encoded (B,L,E); writes att_encoder.weight,
// speaker_embedding.weight and description_embeddings_linear.0.* gradients.  `memory` is the forward's output.
extern "C" int t2_memory_bwd(t2_ctx* c, const float* memory, const long long* speaker_id, const float* description, const float* d_memory,
                             const float* d_processed_memory, int B, int L, float* d_encoded, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_memory_bwd");
  T2_REQUIRE(memory && d_memory && d_processed_memory && d_encoded && ws && B > 0 && L > 0, T2_ERR_ARG, "t2_memory_bwd: bad argument");
  Arena wa(ws, ws_bytes);
  return memory_bwd_impl(c, memory, speaker_id, description, d_memory, d_processed_memory, B, L, d_encoded, wa, (cudaStream_t)stream);
}

extern "C" long long t2_train_workspace_bytes(const t2_ctx* cc, int B, int L, int T) {
  t2_ctx* c = const_cast<t2_ctx*>(cc);
  if (!c || B <= 0 || L <= 0 || T <= 0) return -1;
  long long best = t2_workspace_bytes(cc, B, L, T) + (long long)(c->M + 1) * T * D2_BP * 4 + 256;   // + bulk-projection staging of the train forward
  best = std::max(best, postnet_bwd_ws_bytes(c, B, T) + bn_scratch_bytes(c->PD));
  best = std::max(best, decode_bwd_ws_bytes(c, B, L, T));
  { Arena wa(nullptr, 0); EncTape tp{}; encoder_bwd_impl(c, nullptr, nullptr, B, L, nullptr, 0, nullptr, nullptr, tp, wa, nullptr); best = std::max<long long>(best, wa.off); }
  { Arena wa(nullptr, 0); memory_bwd_impl(c, nullptr, nullptr, nullptr, nullptr, nullptr, B, L, nullptr, wa, nullptr); best = std::max<long long>(best, wa.off); }
  best = std::max<long long>(best, bn_scratch_bytes(c->E) + (long long)B * L * 8 * (c->E / 2) * 4 + 4096);
  return best + 4096;
}


