// Context object shared by the C-ABI translation units (capi.cu, train_capi.cu).
#pragma once
#include "../../include/t2b200.h"
#include "common.cuh"
#include "decoder.cuh"
#include "convtc.cuh"
#include "layers.cuh"

#include <map>
#include <string>

struct Param { const float* ptr; long long numel; };
struct GradSlot { float* ptr; long long numel; };

struct t2_ctx {
  t2_model_config cfg;
  int device, num_sms;
  int EF, C, H, R, P, A, M, E, PD, KS;
  struct { int KattP, KdecP, KprojP; } dd;  // padded contraction lengths of the packed decoder operands
  std::map<std::string, Param> params;
  std::map<std::string, GradSlot> grads;   // destination of each parameter gradient (t2_ctx_set_grad)
  bool committed;
  // packed device memory (one slab)
  char* slab;
  size_t slab_bytes;
  float *enc_scale[3], *enc_shift[3];      // folded BatchNorm (+conv bias) of the encoder convs
  float *post_scale[5], *post_shift[5];
  float *lstm_bias, *lstm_Whh;  // BiLSTM: [1][E][2*4Hh], [2*4Hh], [2][4Hh][Hh]
  unsigned long long* prof_buf;            // [16] device, decoder phase cycles
  unsigned* status_dev;                    // sticky: set by any persistent kernel that aborted (barrier / pipeline timeout); t2_ctx_status reads + clears
  int timing_on;                           // t2_debug_set_timing: CUDA events around the persistent decoder kernels
  cudaEvent_t tev[2][2];                   // [0 = decoder forward, 1 = decoder backward][begin, end]
  bool tev_valid[2];
  // tensor-core decoder (decoder2)
  __nv_bfloat16 *Watt_b, *Wdec_b;          // [2 hi/lo][4H permuted][KattP] / [2][4R][KdecP]
  float* loc_conv_t;                       // [2][31][32]
  float* loc_dense_t;                      // [32][A]
  D2Weights d2w;
  CUtensorMap map_w_att, map_w_dec;
  // location conv (k = 31, 2 -> 32) and location dense (32 -> A) folded into one 62-tap filter bank per attention dim:
  // locG_b[2 hi/lo][128 a][64]: column c*32 + j = sum_f dense[a][f] conv[f][c][j], columns 31 / 63 and rows >= A zero
  __nv_bfloat16* locG_b;
  CUtensorMap map_locG;
  int npass;
  // tensor-core conv operands (bf16 hi/lo [2][Cout][k][Cin]) and their TMA descriptors
  __nv_bfloat16 *enc_Wb[3], *post_Wb[5], *lstm_Wb, *attenc_Wb;
  CUtensorMap enc_wmap[3], post_wmap[5], lstm_wmap, attenc_wmap;
  // training: transposed (+tap-flipped) conv packs for the input gradients, [2][Cin][k][Cout]
  __nv_bfloat16 *enc_WTb[3], *post_WTb[5], *lstm_WTb, *attenc_WTb;
  CUtensorMap enc_wtmap[3], post_wtmap[5], lstm_wtmap, attenc_wtmap;
  // training: transposed LSTMCell packs (tile-major [ktile][pblock][128][64], packed gate-row order) for the in-loop
  // input-gradient GEMM; projection / prenet transposes for the bulk input gradients
  __nv_bfloat16 *WTatt_b, *WTdec_b;
  CUtensorMap map_wt_att, map_wt_dec;
  int KattT, KdecT;                        // K padded to 128
  __nv_bfloat16 *WprojT_b;                 // [2][KprojP][PROJ_LD]  (row k, contraction r = mel rows + gate row)
  // bulk mel/gate projection of the teacher-forced training forward, laid out over the decoder-cell input rows of the tape:
  // WprojA_b[2][M+1][KprojA] multiplies columns [projA_col0, KdecP) of slot t+1 (rnn_h_t), WprojB_b[2][M+1][KprojB] multiplies
  // columns [H, H+KprojB) of slot t (ctx_t | controls); columns outside the segment carry zero weights.
  __nv_bfloat16 *WprojA_b, *WprojB_b;
  int KprojA, KprojB, projA_col0;
  __nv_bfloat16 *W2T_b;                    // [2][P][P]  prenet.3.weight transposed
  __nv_bfloat16 *WpreT_b;                  // [2][P][4H] att_rnn.weight_ih[:, :P] transposed (bulk prenet-output gradient)
};
#define PROJ_LD 96


inline const Param* find(const t2_ctx* c, const std::string& k) {
  auto it = c->params.find(k);
  return it == c->params.end() ? nullptr : &it->second;
}
inline float* find_grad(const t2_ctx* c, const std::string& k, long long numel) {
  auto it = c->grads.find(k);
  return (it == c->grads.end() || it->second.numel != numel) ? nullptr : it->second.ptr;
}

#define CHECK_CTX(name)                                                                        \
  T2_REQUIRE(c, T2_ERR_ARG, name ": null context");                                            \
  T2_REQUIRE(c->committed, T2_ERR_STATE, name ": weights not committed (call t2_ctx_commit)")
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

// implicit-GEMM conv on the tensor cores (capi.cu); scale/shift/keep/act as in ConvTcArgs
int conv_tc(t2_ctx* c, const __nv_bfloat16* in, const CUtensorMap& wmap, int B, int T, int Cin, int Cout, int k, int NT, int act,
            const float* scale, const float* shift, const unsigned char* keep, int keep_ld, float* out_f32,
            __nv_bfloat16* out_b16, cudaStream_t st);
int make_w_map(CUtensorMap* m, const __nv_bfloat16* w, int Cout, int Cin, int k, int NT, int npass);
int make_in_map(CUtensorMap* m, const __nv_bfloat16* x, int B, int T, int Cin, int npass);

// ---- teacher-forced decoder tape (history-mode buffers written by the forward, read by the backward)
struct DecTape {
  __nv_bfloat16* tfp;    // [T+1][2][BP][P]      prenet outputs (bf16 hi,lo)
  __nv_bfloat16* xatt;   // [T+1][2][BP][KattP]  slot t = attention-LSTM input of frame t: [- | ctx_{t-1} | att_h_{t-1}]
  __nv_bfloat16* xdec;   // [T+1][2][BP][KdecP]  slot t = decoder-LSTM input of frame t: [att_h_t | ctx_t | controls | rnn_h_{t-1}]
  float* p1;             // [(T+1)*B][P]         prenet layer-1 output (after dropout), row n = t*B + b
  float* sav_att;        // [T][5][H][BP]
  float* sav_dec;        // [T][5][R][BP]
  float* q;              // [T][BP][A]
  float* w_hist;         // [T+1][BP][L]
  float* wcum_hist;      // [T+1][BP][L]
  float* xloc;           // [T][128 CTAs][ceil(L/S)][A]  location term + processed memory of every frame (pre-tanh energy argument - q)
  float* floc;           // [T][128 CTAs][ceil(L/S)][32] location conv features of every frame
  char* zero_begin; size_t zero_bytes;   // region that must be zeroed before the forward
};
inline DecTape plan_dec_tape(const D2Dims& d, int B, int L, int T, Arena& ar) {
  DecTape t{};
  const int64_t off0 = ar.off;
  t.tfp = ar.take<__nv_bfloat16>((size_t)(T + 1) * 2 * D2_BP * d.P);
  t.xatt = ar.take<__nv_bfloat16>((size_t)(T + 1) * 2 * D2_BP * d.KattP);
  t.xdec = ar.take<__nv_bfloat16>((size_t)(T + 1) * 2 * D2_BP * d.KdecP);
  t.w_hist = ar.take<float>((size_t)(T + 1) * D2_BP * L);
  t.wcum_hist = ar.take<float>((size_t)(T + 1) * D2_BP * L);
  t.zero_begin = ar.base ? ar.base + off0 : nullptr;
  t.zero_bytes = (size_t)(ar.off - off0);
  t.p1 = ar.take<float>((size_t)(T + 1) * B * d.P);
  {
    const int S = d2_att_parts(d.B), Lc = (L + S - 1) / S;
    t.xloc = ar.take<float>((size_t)T * 128 * Lc * d.A);
    t.floc = ar.take<float>((size_t)T * 128 * Lc * D2_LOC_F);
  }
  t.sav_att = ar.take<float>((size_t)T * 5 * d.H * D2_BP);
  t.sav_dec = ar.take<float>((size_t)T * 5 * d.R * D2_BP);
  t.q = ar.take<float>((size_t)T * D2_BP * d.A);
  return t;
}
D2Dims dims2(const t2_ctx* c, int B, int L);
// After a persistent kernel: if the context's sticky status word is set, the first floats of `out` become NaN, so that an abort
// can never pass for a result (the loss / gradient norm downstream turns NaN) even where no host code looks at the status.
int poison_on_abort_launch(const unsigned* status, float* out, long long n, cudaStream_t st);
inline void timing_begin(t2_ctx* c, int which, cudaStream_t st) { if (c->timing_on) { cudaEventRecord(c->tev[which][0], st); } }
inline void timing_end(t2_ctx* c, int which, cudaStream_t st) { if (c->timing_on) { cudaEventRecord(c->tev[which][1], st); c->tev_valid[which] = true; } }
