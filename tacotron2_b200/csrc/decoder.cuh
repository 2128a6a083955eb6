// Persistent autoregressive decoder: parameter blocks shared by host and device.
// Replaces the Python frame loop of the reference (model/tacotron2.py:276-325) and the
// per-frame Decoder.forward (model/decoder.py:53-119) + Attention.forward
// (model/attention.py:34-69) + prenet (model/tacotron2.py:85-92).
#pragma once
#include "common.cuh"

#define DEC_BP 64          // batch slots per launch (batch padded to this in k-major buffers)
#define DEC_THREADS 256
#define DEC_LOC_F 32       // attention_location_n_filters  (model/decoder.py:35, hard-coded)
#define DEC_LOC_K 31       // attention_location_kernel_size (model/decoder.py:36, hard-coded)
#define DEC_MAX_A 128      // att_dim upper bound for the smem-resident attention weights

struct DecDims {
  int B, L, EF, H, R, P, A, M, C;
  int Katt, KattP;    // P+EF+H, padded to 64
  int Kdec, KdecP;    // H+EF+C+R, padded to 64
  int Kproj, KprojP;  // R+EF+C, padded to 4
  float p_prenet;     // AlwaysDropout p (config dropout)
};

struct DecWeights {
  const float* Watt_p;  // [H/8][KattP][32]  tile-packed, k-major; tile row = gate*8 + unit_in_tile
  const float* batt;    // [4H] b_ih + b_hh (reference gate order i,f,g,o)
  const float* Wdec_p;  // [R/8][KdecP][32]
  const float* bdec;    // [4R]
  const float* Wq;      // [A][H]        decoder.attention.query_layer.weight
  const float* v;       // [A]           decoder.attention.v.weight
  const float* loc_conv;   // [32][2][31] decoder.attention.location_conv.weight
  const float* loc_dense;  // [A][32]     decoder.attention.location_dense.weight
  const float* Wproj;   // [M+1][KprojP] rows 0..M-1 = mel_out.weight, row M = gate.weight (controls columns 0)
  const float* bproj;   // [M+1]
  const float* W1;      // [P][M]  prenet.0.weight
  const float* W2;      // [P][P]  prenet.3.weight
};

// Mutable state, all in the caller's workspace.  k-major buffers are [K][DEC_BP].
struct DecState {
  float* xpre[2];    // [P][BP]      prenet output feeding the attention LSTM (AR mode), ping-pong by step parity
  float* xatt[2];    // [KattP-P][BP] rows: ctx (EF), att_h (H), zero pad
  float* xdec[2];    // [KdecP][BP]  rows: att_h (H), ctx (EF), controls (C), rnn_h (R), zero pad
  float* att_c;      // [H][BP]
  float* rnn_c;      // [R][BP]
  float* att_h_rm;   // [BP][H]      row-major copy for the query projection
  float* projin;     // [BP][KprojP] row-major [rnn_h | ctx | controls | pad]
  float* q;          // [BP][A]
  float* w_prev;     // [BP][L]
  float* w_cum;      // [BP][L]
  float* mel_cur;    // [BP][M]
  float* p1;         // [BP][P]
  int* done;         // [BP]
  long long* lengths;  // [BP]
  int* n_frames;     // [1]
  unsigned* barrier; // [2]  monotonic arrival counter, abort flag
  unsigned long long* prof;  // [16] per-phase SM-clock cycles seen by CTA 0 (debug / DESIGN.md breakdown)
};

struct DecIO {
  const float* memory;        // [B][L][EF]
  const float* pmem;          // [B][L][A]
  const unsigned char* mask;  // [B][L] 1 = padded (encoded_mask, model/tacotron2.py:223-226)
  const unsigned char* prenet_keep;  // [T+1][2][B][P] or NULL (in-kernel RNG)
  const unsigned char* att_keep;     // [T][B][H] train-mode p=0.1 keep mask, or NULL
  const unsigned char* rnn_keep;     // [T][B][R]
  const float* tf_prenet;     // teacher forcing: [T+1][P][BP] k-major bulk prenet output, or NULL (AR)
  float* mels;                // [B][Tmax][M]
  float* gates;               // [B][Tmax]
  float* aligns;              // [B][Tmax][L]
  float* projin_all;          // TF mode: [T][B][KprojP] saved projection inputs (bulk projection afterwards), or NULL
  int Tmax;                   // rows allocated in mels/gates/aligns
  int n_steps;                // frames to run (<= Tmax)
  int t0;                     // index of the first frame (mask/out offset; single-step API uses t0 = 0)
  int teacher_forcing;
  int hidden_dropout;         // 1 = train mode: apply att_keep / rnn_keep (or RNG) with p = 0.1
  int stop_enabled;           // AR: evaluate the stop gate (model/tacotron2.py:319-322)
  unsigned long long seed;
};

int dec_launch(const DecDims& d, const DecWeights& w, const DecState& s, const DecIO& io, int num_sms, cudaStream_t stream);
size_t dec_smem_bytes(const DecDims& d);
