// Persistent autoregressive decoder: parameter blocks shared by host and device.
// Replaces the Python frame loop of the reference (model/tacotron2.py:276-325), the per-frame
// Decoder.forward (model/decoder.py:53-119), Attention.forward (model/attention.py:34-69) and the
// prenet (model/tacotron2.py:85-92).
#pragma once
#include "tcgemm.cuh"

#define D2_BP 64           // batch slots per launch (UMMA N)
#define D2_THREADS 256
#define D2_CLUSTER 4       // CTAs per cluster = split-K factor of the gate GEMMs
#define D2_STAGES 3
#define D2_LOC_F 32        // attention_location_n_filters  (model/decoder.py:35, hard-coded)
#define D2_LOC_K 31        // attention_location_kernel_size (model/decoder.py:36, hard-coded)
#define D2_MAX_A 128

struct D2Dims {
  int B, L, EF, H, R, P, A, M, C;
  int KattP;   // P+EF+H padded to 64: columns [prenet | ctx | att_h | 0]
  int KdecP;   // H+EF+C+R padded to 64: columns [att_h | ctx | controls | rnn_h | 0]
  int KprojP;  // R+EF+C padded to 4:   columns [rnn_h | ctx | controls | 0]
  int npass;   // 3 = fp32-parity (bf16 hi/lo split), 1 = bf16
  float p_prenet;
};

struct D2Weights {
  const float* batt;       // [4H] b_ih + b_hh, reference row order (i,f,g,o blocks)
  const float* bdec;       // [4R]
  const float* Wq;         // [A][H]
  const float* v;          // [A]
  const float* loc_conv_t; // [2][31][32]  location_conv.weight transposed to (channel, tap, filter)
  const float* loc_dense;  // [32][A]  location_dense.weight transposed
  const float* Wproj;      // [M+1][KprojP]
  const float* bproj;      // [M+1]
  const float* W1;         // [P][M]
  const float* W2;         // [P][P]
};

struct D2State {
  __nv_bfloat16* xatt;   // [2 parity][2 hi/lo][BP][KattP]
  __nv_bfloat16* xdec;   // [2][2][BP][KdecP]
  float* att_c;          // [H][BP]  (loaded at kernel start, stored at kernel end)
  float* rnn_c;          // [R][BP]
  float* att_h_rm;       // [BP][H]
  float* projin;         // [BP][KprojP]
  float* q;              // history mode only: [T][BP][A] processed query of every frame (training tape)
  float* accpm;          // [128 CTAs][ceil(L/S)][A] location term + processed memory of each CTA's attention slice (precomputed
                         // during the attention-LSTM GEMM of the same frame, consumed by the attention phase)
  float* floc;           // history mode only: [T][128 CTAs][ceil(L/S)][32] location conv features (training tape), else null;
                         // accpm then also holds every frame: [T][128 CTAs][ceil(L/S)][A]
  float* qpart;          // [BP][128 CTAs][A] per-CTA partial sums of the query projection (each CTA owns 8 units of att_h)
  float* w_prev;         // [2][BP][L]
  float* w_cum;          // [2][BP][L]
  float* mel_cur;        // [BP][M]
  float* p1;             // [BP][P]
  int* done;             // [BP]
  long long* lengths;    // [BP]
  int* n_frames;         // [1]
  unsigned* barrier;     // [2]
  unsigned* bar_flags;   // [256] per-CTA arrival epochs of the pipelined AR kernel's grid barrier
  unsigned long long* prof;  // [16] or null
  unsigned* status;          // context's sticky abort word or null
  // history mode (training tape): xatt/xdec/w_prev/w_cum hold T+1 slots and q holds T slots instead of two parities
  float* sav_att;        // [T][5][H][BP] i,f,g,o (post-activation), c of the attention LSTMCell, or null
  float* sav_dec;        // [T][5][R][BP]
};

struct D2IO {
  const float* memory;        // [B][L][EF]
  const float* pmem;          // [B][L][A]
  const unsigned char* mask;  // [B][L]
  const unsigned char* prenet_keep;  // [T+1][2][B][P] or null
  const unsigned char* att_keep;     // [T][B][H] or null
  const unsigned char* rnn_keep;     // [T][B][R] or null
  float* mels;   // [B][Tmax][M]
  float* gates;  // [B][Tmax]
  float* aligns; // [B][Tmax][L]
  int Tmax, n_steps, t0;
  int teacher_forcing, hidden_dropout, stop_enabled;
  int min_frames;     // the stop condition is honoured only once this many frames have been emitted (chunks of a larger batch)
  int debug_barrier;  // experiment bit mask (env T2B200_DEBUG_BARRIER): read by profiling builds of the decoder kernels only
  int precompute;     // set by d2_launch: shared memory allows the location conv + dense to run beside the attention-LSTM GEMM
  int w_policy_att, w_policy_dec;  // set by d2_launch: L2 eviction priority of the LSTM weight tiles (0 default, 1 evict-first, 2 evict-last)
  int hist;           // 1: state buffers are indexed by frame (slot t / t+1) instead of parity -- the training tape
  unsigned long long seed;
};

struct D2Maps {  // TMA descriptors (bf16, 128-B swizzle, box 64 x rows x {1|2})
  CUtensorMap w_att, w_dec;   // [K][4H permuted rows][2]
  CUtensorMap x_att;          // [KattP][BP][4 = parity*2 + hl]
  CUtensorMap x_att_rest;     // same buffer starting at column P (teacher forcing: prenet part comes from tf_pre)
  CUtensorMap x_dec;          // [KdecP][BP][4]
  CUtensorMap tf_pre;         // [P][BP][(T+1)*2]
};

// Pipelined autoregressive schedule (decoder_ar.cu): eval-mode synthesis when the shape fits (else decoder_persistent_kernel)
bool d2ar_supported(const D2Dims& d);
int d2ar_launch(const D2Dims& d, const D2Weights& w, const D2State& s, const D2IO& io, const D2Maps& maps, const CUtensorMap& gmap,
                cudaStream_t stream);
int d2_launch(const D2Dims& d, const D2Weights& w, const D2State& s, const D2IO& io, const D2Maps& maps, cudaStream_t stream);
int d2_max_clusters(int* out);
size_t d2_smem_bytes(const D2Dims& d);
int d2_att_parts(int B);   // S: CTAs per utterance in the attention phase
bool d2_precompute_fits(const D2Dims& d);   // the location term is computed beside the attention-LSTM GEMM (and, in training, taped)
