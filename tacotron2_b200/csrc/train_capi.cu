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
  return transpose_bf16_launch(x, C, (long long)T * C, (long long)B * T * C, xT, T, C, Tp, (long long)B * Tp, Tp, B, shift, st);
}

// conv weight gradient dW[co][ci][j] = sum_{b,t} dz[b][t][co] x[b][t+j-pad][ci] on the tensor cores.
// dzT [2][Cout][B][Tp] K-major in time; x channels-last [2][B][T][Cin]; xTk = scratch for the k time-shifted K-major copies
// of x, [k][2][Cin][B][Tp] (tap shifts cannot be TMA coordinates: 16-byte alignment of the innermost dimension).
int conv_wgrad(t2_ctx* c, const __nv_bfloat16* dzT, const __nv_bfloat16* x, __nv_bfloat16* xTk, int B, int T, int Tp, int Cout, int Cin,
               int k, float* dW, float* partial, long long partial_floats, cudaStream_t st) {
  GemmNtArgs a{};
  a.M = Cout; a.N = Cin; a.NT = Cin % 128 == 0 ? 128 : ((Cin + 15) & ~15);
  T2_REQUIRE(a.NT <= 128, T2_ERR_UNSUPPORTED, "conv_wgrad: Cin=%d", Cin);
  a.Kin = T; a.Kout = B; a.nz = 1; a.shift0 = 0; a.shift_step = 0; a.a_zrow = 0; a.b_zrow = 0;
  a.splitk = 0; a.npass = c->npass;
  const long long need = gemmnt_partial_floats(a, c->num_sms);
  T2_REQUIRE(need <= partial_floats, T2_ERR_WORKSPACE, "conv_wgrad: partial buffer too small");
  a.partial = partial;
  RC(gemmnt_make_map(&a.a_map, dzT, T, Tp, B, Cout, (long long)Cout * B * Tp, 128, c->npass));
  const int pad = (k - 1) / 2;
  const size_t copy = (size_t)2 * Cin * B * Tp;
  for (int j = 0; j < k; ++j) RC(to_kmajor(x, xTk + j * copy, B, T, Tp, Cin, st, j - pad));
  for (int j = 0; j < k; ++j) {
    RC(gemmnt_make_map(&a.b_map, xTk + j * copy, T, Tp, B, Cin, (long long)Cin * B * Tp, a.NT, c->npass));
    GemmNtOut o{dW + j, (long long)Cin * k, (long long)k, 0ll, 1.0f, 0};
    RC(gemmnt_launch(a, o, st));
  }
  return T2_OK;
}
long long conv_wgrad_partial_floats(t2_ctx* c, int B, int T, int Cout, int Cin, int k) {
  GemmNtArgs a{};
  a.M = Cout; a.N = Cin; a.NT = Cin % 128 == 0 ? 128 : ((Cin + 15) & ~15);
  a.Kin = T; a.Kout = B; a.nz = 1; a.splitk = 0;
  (void)k;
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
  const int PD = c->PD, M = c->M, Tp = round8(T);
  const int chans[6] = {M, PD, PD, PD, PD, M};
  void* scratch = wa.take<char>(bn_scratch_bytes(PD));
  float* dyA = wa.take<float>((size_t)B * T * PD);
  float* dyB = wa.take<float>((size_t)B * T * PD);
  __nv_bfloat16* dz = wa.take<__nv_bfloat16>((size_t)2 * B * T * PD);
  __nv_bfloat16* dzT = wa.take<__nv_bfloat16>((size_t)2 * PD * B * Tp);
  __nv_bfloat16* xT = wa.take<__nv_bfloat16>((size_t)5 * 2 * PD * B * Tp);
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
    RC(to_kmajor(dz, dzT, B, T, Tp, Cout, st));
    RC(conv_wgrad(c, dzT, tp.x[i], xT, B, T, Tp, Cout, Cin, 5, gW, partial, pf, st));
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

extern "C" long long t2_train_workspace_bytes(const t2_ctx* cc, int B, int L, int T) {
  t2_ctx* c = const_cast<t2_ctx*>(cc);
  if (!c || B <= 0 || L <= 0 || T <= 0) return -1;
  long long best = t2_workspace_bytes(cc, B, L, T);
  best = std::max(best, postnet_bwd_ws_bytes(c, B, T) + bn_scratch_bytes(c->PD));
  return best + 4096;
}
