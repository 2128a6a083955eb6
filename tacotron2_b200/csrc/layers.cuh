// Encoder / postnet / conditioning layers (channels-last activations [B][T][C]).
#pragma once
#include "common.cuh"

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2 };

struct ConvArgs {
  const float* in;      // [B][*][Cin] channels-last; element (b,t,c) at in + b*in_bstride + t*Cin + c
  long long in_bstride;
  const float* Wr;      // [k][Cin][Cout] repacked weight
  const float* bias;    // [Cout] or null
  const float* bn_scale;  // [Cout] or null: y = conv*scale + shift (eval BatchNorm folded with bias)
  const float* bn_shift;
  const unsigned char* keep;  // dropout keep-mask [B][T][keep_ld] or null
  int keep_ld;
  float drop_scale;     // 1/(1-p)
  float* out;           // [B][T][Cout] dense
  int B, T, Cin, Cout, k, act;
};

int conv1d_cl_launch(const ConvArgs& a, cudaStream_t st);
int embed_launch(const long long* idx, const float* table, float* out, int n_pos, int E, int n_rows, cudaStream_t st);
int bilstm_launch(const float* xproj /*[B][L][2][4Hh]*/, const float* Whh /*[2][4Hh][Hh] (reference layout)*/, const long long* lens,
                  float* out /*[B][L][2Hh]*/, int B, int L, int Hh, cudaStream_t st);
int condition_launch(const float* enc /*[B][L][E]*/, const float* spk_table, const long long* spk_id,
                     const float* desc_vec /*[B][D] already tanh(linear)*/, int D, float* out /*[B][L][E+D]*/, int B, int L,
                     int E, cudaStream_t st);
int desc_linear_launch(const float* desc /*[B][Din]*/, const float* W /*[128][Din]*/, const float* bias, float* out, int B,
                       int Din, int Dout, cudaStream_t st);
int finalize_launch(float* mels, float* post /*in: postnet out, out: mels_post*/, float* gates, const long long* lengths,
                    int B, int T, int Tstride, int M, cudaStream_t st);
int bn_fold_launch(const float* w, const float* b, const float* mean, const float* var, const float* conv_bias, float* scale,
                   float* shift, int C, cudaStream_t st);
int repack_conv_launch(const float* W /*[Cout][Cin][k]*/, float* Wr /*[k][Cin][Cout]*/, int Cout, int Cin, int k, cudaStream_t st);
int transpose_launch(const float* in /*[R][C]*/, float* out /*[C][R]*/, int R, int C, cudaStream_t st);
