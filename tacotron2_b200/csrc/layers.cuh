// Encoder / postnet / conditioning layers (channels-last activations [B][T][C]).
#pragma once
#include "common.cuh"

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2, ACT_GELU = 3 };   // GELU: exact erf form (BERT "gelu")

int bilstm_launch(const float* xproj /*[B][L][2][4Hh]*/, const float* Whh /*[2][4Hh][Hh] (reference layout)*/, const long long* lens,
                  float* out /*[B][L][2Hh]*/, int B, int L, int Hh, cudaStream_t st,
                  float* sav = nullptr /*training tape [B][L][2][5][Hh] or null*/);
int condition_launch(const float* enc /*[B][L][E]*/, const float* spk_table, const long long* spk_id,
                     const float* desc_vec /*[B][D] already tanh(linear)*/, int D, float* out /*[B][L][E+D]*/, int B, int L,
                     int E, cudaStream_t st);
int desc_linear_launch(const float* desc /*[B][Din]*/, const float* W /*[128][Din]*/, const float* bias, float* out, int B,
                       int Din, int Dout, cudaStream_t st);
int finalize_launch(float* mels, float* post /*in: postnet out, out: mels_post*/, float* gates, const long long* lengths,
                    int B, int T, int Tstride, int M, cudaStream_t st);
int bn_fold_launch(const float* w, const float* b, const float* mean, const float* var, const float* conv_bias, float* scale,
                   float* shift, int C, cudaStream_t st);
int transpose_launch(const float* in /*[R][C]*/, float* out /*[C][R]*/, int R, int C, cudaStream_t st);
