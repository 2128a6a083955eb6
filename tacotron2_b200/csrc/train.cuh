// Training path (teacher-forced forward with a tape + backward): shared declarations.
// The reference obtains the backward pass from autograd over model/tacotron2.py:155-347 (training_step,
// model/tts_model.py:165-202); here it is explicit kernels.
#pragma once
#include "tc.cuh"
#include "layers.cuh"
#include "decoder.cuh"

// One conv -> BatchNorm -> activation -> dropout layer in train mode (bn.cu)
struct BnLayer {
  const float* z;        // [rows][C] pre-BatchNorm conv output (incl. conv bias)
  const float* mean;     // [C] batch mean
  const float* rstd;     // [C] 1/sqrt(biased var + eps)
  const float* gamma;    // [C]
  const float* beta;     // [C]
  int C, act, dropout;
  const unsigned char* keep;  // [rows][keep_ld] or null (counter RNG)
  int keep_ld;
  float p;
  unsigned long long seed, rng_base;
};

long long bn_scratch_bytes(int C);
int bn_stats_launch(const float* z, long long rows, int C, float* mean, float* rstd, float* running_mean, float* running_var,
                    void* scratch, cudaStream_t st);
int bn_apply_launch(const BnLayer& L, long long rows, float* out_f32, __nv_bfloat16* out_b16, cudaStream_t st);
// dy [rows][C] -> dgamma, dbeta (may be null), dz as bf16 (hi,lo) [2][rows][C] and/or fp32
int bn_bwd_launch(const BnLayer& L, const float* dy, long long rows, float* dgamma, float* dbeta, __nv_bfloat16* dz_b16, float* dz_f32,
                  void* scratch, cudaStream_t st);

// ---- decoder backward (bwd_decoder.cu) ----------------------------------------------------------------------------
struct DBWeights {
  const float* Wq;          // [A][H]
  const float* v;           // [A]
  const float* loc_conv_t;  // [2][31][32]
  const float* loc_dense;   // [32][A]
};

struct DBTape {             // written by the teacher-forced forward in history mode
  const float* sav_att;     // [T][5][H][BP]  i,f,g,o (post-activation), c
  const float* sav_dec;     // [T][5][R][BP]
  const float* q;           // [T][BP][A]
  const float* xloc;        // [T][128][ceil(L/S)][A]  taped location term + processed memory, or null (recompute)
  const float* floc;        // [T][128][ceil(L/S)][32] taped location conv features, or null (recompute)
  const float* w_hist;      // [T+1][BP][L]  slot t+1 = w_t, slot 0 = zeros
  const float* wcum_hist;   // [T+1][BP][L]  slot t+1 = cum_t
  const float* memory;      // [B][L][EF]
  const float* pmem;        // [B][L][A]
  const unsigned char* mask;      // [B][L]
  const unsigned char* att_keep;  // [T][B][H] or null
  const unsigned char* rnn_keep;  // [T][B][R] or null
  unsigned long long seed;
  int hidden_dropout;
};

struct DBWork {
  const float* dprojin;     // [T][BP][KprojP]  d(loss)/d[rnn_h | ctx | controls] from the mel/gate projection (bulk GEMM)
  __nv_bfloat16* dg_att;    // [2 parity][2 hi/lo][BP][4H]   packed row order (operand of the in-loop input-gradient GEMM)
  __nv_bfloat16* dg_dec;    // [2][2][BP][4R]
  __nv_bfloat16* dgT_att;   // [2 hi/lo][4H][T*BP]   reference row order (operand of the weight-gradient GEMM)
  __nv_bfloat16* dgT_dec;   // [2][4R][T*BP]
  float* dx_att;            // [2 parity][KattT][BP]  output of the att input-gradient GEMM (k-major)
  float* dx_dec;            // [2 parity][KdecT][BP]
  float* dctxT_dec;         // [BP][EF] the context rows of dx_dec / dx_att utterance-major: the attention backward reads one utterance's
  float* dctxT_att;         //          row (the k-major layout costs one 32-byte sector per element there)
  float* dpre;              // [T][BP][P]     d/d prenet output
  float* dctx_tot;          // [T][BP][EF]    total d/d ctx_t (for the bulk memory gradient)
  float* dq;                // [T][S][A][BP]  per-part partial d/d query (utterance-minor)
  float* dpm;               // [B][L][A]      accumulated d/d processed memory
  float* dconv;             // [2 parity][2 ch][BP][L]   d/d (w_{t-1}, cum_{t-1}) from the location conv
  float* gcum;              // [BP][L]        running d/d cum
  float* small_partial;     // [grid][A*32 + 2*31*32 + A]  per-CTA partials of location_dense / location_conv / v gradients
  float* dbias;             // [2][4H]   att, dec bias gradients
  unsigned* barrier;        // [4]
  unsigned* status;         // context's sticky abort word or null
};

struct DBMaps {
  CUtensorMap wt_att, wt_dec;   // transposed packed weights: tile-major [ktile][pblock][128][64] x 2 planes
  CUtensorMap dg_att, dg_dec;   // [4H][BP][4 = parity*2 + hl]
};

int db_launch(const D2Dims& d, int T, int KattT, int KdecT, const DBWeights& w, const DBTape& tp, const DBWork& wk, const DBMaps& maps,
              unsigned long long* prof, cudaStream_t st);
// Bulk mel/gate projection of the teacher-forced training forward (model/decoder.py:103-109 for all T frames at once, off the
// recurrence): reads rnn_h_t / ctx_t / controls from the decoder-cell operand rows of the tape.  tmp: (M+1) * T*64 floats.
struct t2_ctx;
int decode_proj_bulk(t2_ctx* c, const __nv_bfloat16* xdec_tape, int B, int T, float* mels, float* gates, float* tmp, cudaStream_t st);
int db_parts(const D2Dims& d);        // S: CTAs per utterance in the attention backward
bool db_tc_att(const D2Dims& d, bool have_xloc);   // attention backward contractions on tcgen05 (else CUDA cores)
size_t db_small_floats(const D2Dims& d, bool tc_att);

// ---- BiLSTM backward (bwd_layers.cu) ------------------------------------------------------------------------------
// sav [B][L][2][5][Hh] (i,f,g,o,c) from the train-mode forward; d_out [B][L][2Hh]; dgates out [B][L][2][4Hh] (zeros at t >= len)
int bilstm_bwd_launch(const float* sav, const float* Whh, const long long* lens, const float* d_out, float* dgates, int B, int L, int Hh,
                      cudaStream_t st);

// ---- small pointwise / reduction kernels (bwd_layers.cu) ----------------------------------------------------------
int colsum_launch(const float* x, long long rows, int C, long long ld, float* out, void* scratch /*bn_scratch_bytes(C)*/, cudaStream_t st);
int loss_launch(const float* mels, const float* post, const float* gates, const float* tgt, const float* gtgt, long long n_mel,
                long long n_gate, float* loss, float* d_mels, float* d_post, float* d_gates, void* scratch /*>= 4 KB*/, cudaStream_t st);
int finalize_bwd_launch(const float* d_mels, const float* d_post, const float* d_gates, const long long* lengths, int B, int T, int M,
                        float* d_mels_raw, float* d_postout, float* d_gates_raw, cudaStream_t st);
int build_dproj_launch(const float* d_mels, const float* d_gates, int B, int T, int M, int ld, __nv_bfloat16* dP, __nv_bfloat16* dPT,
                       cudaStream_t st);
int dq_sumT_launch(const float* dq, int T, int S, int SP, int A, int B, __nv_bfloat16* out, cudaStream_t st);
int partial_sum_launch(const float* partial, int nparts, int n, float* out, cudaStream_t st);
int copy2_launch(const float* src, float* a, float* b, int n, cudaStream_t st);
int att_small_finish_launch(const float* Y, const float* conv_t, const float* dense_t, int A, float* g_dense, float* g_conv, cudaStream_t st);
int prenet_d2_launch(const float* dpre, const __nv_bfloat16* tfp, int B, int T, int P, float scale, float* d2, cudaStream_t st);
int prenet_d1_launch(const float* dp1, const float* p1, long long n, float scale, float* d1, cudaStream_t st);
int prenet_xT_launch(const float* mel_target, int B, int T, int M, int Np, __nv_bfloat16* xT, cudaStream_t st);
int condition_bwd_launch(const float* dmem, const float* mem, int speaker, float* d_enc, long long rows, int E, int EF, cudaStream_t st);
int speaker_grad_launch(const float* d_enc, const long long* spk, int B, int L, int E, int nspk, float* out, float* scratch /*[B][E] or null*/,
                        cudaStream_t st);
int desc_bwd_launch(const float* dmem, const float* mem, const float* desc, int B, int L, int E, int D, int Din, float* dpre, float* dW,
                    float* db, cudaStream_t st);
int embedding_grad_launch(const float* dx, const long long* chars, long long npos, int E, int nrows, float* out, cudaStream_t st);
int zero_launch(float* p, long long n, cudaStream_t st);
