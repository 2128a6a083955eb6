// Channels-last conv1d (odd k, "same" padding) as an implicit GEMM on tcgen05:
//   out[(b,t)][n] = sum_{j<k} sum_c in[b][t+j-pad][c] * W[n][c][j]
// A = activations [128 positions x 64 channels] per (tap, channel block), fetched by TMA from a bf16 (hi,lo)
// [2][B][T][C] tensor with out-of-bounds zero fill providing the padding; B = weights [N x 64] from a bf16 (hi,lo)
// [2][Cout][k][Cin] pack; D = [128 x N] fp32 in TMEM.  fp32-parity mode issues three bf16 passes per K-step.
// Used for Encoder convs (model/encoder.py:31-46), Postnet convs (model/postnet.py:8-47), and the 1x1 cases
// att_encoder (model/tacotron2.py:229) and the BiLSTM input projection (model/encoder.py:49-52).
#pragma once
#include "tc.cuh"

struct ConvTcArgs {
  CUtensorMap in_map;   // dims {Cin, T, B, 2}, box {64, 128, 1, npass==3 ? 2 : 1}
  CUtensorMap w_map;    // dims {Cin, k, Cout, 2}, box {64, 1, NT, ...}
  int B, T, Cin, Cout, k, NT;   // NT = N tile (multiple of 16, <= 128)
  int npass, act;
  const float* scale;   // [Cout] folded BatchNorm scale or null
  const float* shift;   // [Cout] folded shift / bias or null
  const unsigned char* keep;  // dropout keep-mask [B][T][keep_ld] or null
  int keep_ld;
  float drop_scale;
  float* out_f32;             // [B][T][Cout] or null
  __nv_bfloat16* out_b16;     // [2][B][T][Cout] (hi plane, lo plane) or null
  const float* residual;      // optional [B][res_tstride][Cout] added before the store (postnet residual), or null
  long long res_bstride;
  // --- extensions used by the HiFi-GAN generator (hifigan.cu); zero-initialised = the plain conv above
  int dil;                    // dilation (0 or 1 = none): tap j reads position t + (j - (k-1)/2) * dil
  int chan_mod;               // > 0: scale / shift are indexed by (channel % chan_mod) (transposed conv as u phase groups of Cout)
  float out_scale;            // 0 = 1: value is multiplied by this after the residual add ...
  const float* accum;         // ... and this [B][T][Cout] tensor is added (may alias out_f32): multi-receptive-field sum
  float b16_slope;            // != 0: the bf16 (hi,lo) copy holds leaky_relu(value, slope) -- the next conv's input -- out_f32 the value
  float* out_c0;              // optional [B][T]: channel 0 only (conv_post: one real output channel in a 16-wide tile)
  // --- narrow layers (Cin <= 64, Cout == NT <= 64, k >= 3): conv_halo_kernel loads each 256-position activation tile ONCE with its halo
  // and gives every tap a shifted view of it, instead of reloading the tile per tap.  halo = 1 asks for it; the two extra maps are
  // single-plane views of the input (convtc_make_halo_maps): box {64, 128, 1, 1} and {64, 64, 1, 1}.
  int halo;
  CUtensorMap in_map_h, in_map_t;
};

// Whether the layer qualifies for conv_halo_kernel (T2B200_CONV_HALO=0 disables it), and the two maps it needs.
bool convtc_halo_ok(int Cin, int Cout, int NT, int k, int dil);
int convtc_make_halo_maps(ConvTcArgs& a, const __nv_bfloat16* x, int B, int T, int Cin);

int convtc_launch(const ConvTcArgs& a, cudaStream_t st);
// fp32 [B][T(stride)][C] -> bf16 (hi,lo) [2][B][T][C]
int split_cl_launch(const float* src, long long src_bstride, __nv_bfloat16* dst, int B, int T, int C, cudaStream_t st);
// conv weight [Cout][Cin][k] fp32 -> bf16 (hi,lo) [2][Cout][k][Cin]
int pack_conv_w_launch(const float* W, __nv_bfloat16* dst, int Cout, int Cin, int k, cudaStream_t st);
// embedding lookup straight to bf16 (hi,lo) [2][n_pos][E]
int embed_split_launch(const long long* idx, const float* table, __nv_bfloat16* dst, int n_pos, int E, int n_rows, cudaStream_t st);
