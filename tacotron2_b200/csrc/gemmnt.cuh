// Bulk "NT" GEMM on tcgen05 for the training path: C[z][m][n] = sum_k A[m + z*a_zrow][k] * B[n + z*b_zrow][k(+shift_z)]
// with BOTH operands K-major bf16 (hi,lo) planes fetched by TMA.  The contraction index is two-level (kin, kout), e.g.
// (time, utterance): per kout the kin range is walked in 64-element blocks and the B operand's kin coordinate can be
// shifted per z (conv taps, the h_{t-1} shift of a recurrent weight gradient); out-of-range elements are zero-filled by
// TMA, which is exactly the conv zero padding / the zero initial state.  Split-K partials go to fp32 slabs that a second
// kernel reduces in a fixed order (deterministic) into an arbitrarily strided destination.
// Used for: LSTMCell / BiLSTM / conv / linear weight gradients and the bulk input gradients of the teacher-forced
// backward pass (the reference gets these from autograd over model/tacotron2.py:155-347).
#pragma once
#include "tc.cuh"

struct GemmNtArgs {
  CUtensorMap a_map;   // dims {Kin, Kout, Mrows, 2}, box {64, 1, 128, npass==3 ? 2 : 1}
  CUtensorMap b_map;   // dims {Kin, Kout, Nrows, 2}, box {64, 1, NT, ...}
  int M, N, NT;        // logical output size; tiles = ceil(M/128) x ceil(N/NT)
  int Kin, Kout;
  int nz, shift0, shift_step;   // B kin shift for batch z: shift0 + z*shift_step (must be a multiple of 8 elements)
  int a_zrow, b_zrow;           // operand row offsets per z
  int splitk, npass;
  float* partial;               // [splitk][nz][Mt*128][Nt*NT]
  float* direct_out;            // splitk == 1 && nz == 1 only: write C[m][n] straight to direct_out[m*direct_ldm + n] (no partial slab)
  long long direct_ldm;
  // optional second contraction segment (own operand maps, Kout = 1, no shift): C += A2 * B2^T over Kin2 -- lets one launch
  // contract over two column windows that live in different places (e.g. different time slots of a tape)
  CUtensorMap a_map2, b_map2;
  int Kin2;
  int b_row_step;               // B-operand row coordinate of N-tile nt is nt*b_row_step (0: nt*NT) -- rows interleaved with other data
  // mn = 1: both operands are read MN-major from channels-last tensors [2][Kout][Kin][rows] (maps from gemmnt_make_map_cl: dims
  // {rows, Kin, Kout, 2}, box {64, 64, 1, planes}): the contraction index kin is then an OUTER TMA coordinate, so the per-z shift may be
  // any integer (conv taps need no shifted copies) and neither operand needs a K-major transpose.  No second segment, no z row offsets.
  // mn is a bit mask: 1 = A MN-major, 2 = B MN-major, 3 = both (the conv weight gradients).  With one MN-major operand the other keeps the
  // K-major form above and the two operands may split the contraction index differently: KinB (0 = Kin) is the B operand's inner extent,
  // K-block kb reads B at (kin, kout) = ((kb mod KinB/64)*64 + shift, kb div KinB/64).  E.g. the LSTM weight gradients: A = d gates
  // K-major over (frame, utterance) as one run of T*64, B = the operand rows [frame][plane][utterance][k] as they lie: KinB = 64, kout = frame.
  int mn;
  int KinB;
};

struct GemmNtOut {
  float* out;
  long long ldm, ldn, ldz;   // element strides of the destination
  float alpha;
  int accumulate;            // 1: out += alpha * sum, 0: out = alpha * sum
};

// Chooses splitk (stored into a.splitk) unless already > 0; returns partial-buffer floats needed.
long long gemmnt_partial_floats(GemmNtArgs& a, int num_sms);
int gemmnt_launch(const GemmNtArgs& a, const GemmNtOut& o, cudaStream_t st);

// Operand maps for a K-major bf16 (hi,lo) matrix stored as [2][rows][Kout][kin_ld] (kin contiguous, kin_ld % 8 == 0).
int gemmnt_make_map(CUtensorMap* m, const __nv_bfloat16* base, int Kin, int kin_ld, int Kout, int rows, long long plane_elems,
                    int box_rows, int npass);

// Operand map for a channels-last bf16 (hi,lo) tensor [2][Kout][Kin][rows] (rows contiguous, rows % 8 == 0), read MN-major (a.mn = 1).
int gemmnt_make_map_cl(CUtensorMap* m, const __nv_bfloat16* base, int Kin, int Kout, int rows, long long plane_elems, int npass);
// The same with explicit element strides of the kin / kout / plane coordinates (operand rows that are interleaved with other data).
int gemmnt_make_map_cl_strided(CUtensorMap* m, const __nv_bfloat16* base, int Kin, int Kout, int rows, long long kin_elems, long long kout_elems,
                               long long plane_elems, int npass);

// Batched transposes into K-major operands.  Batch bb reads src + bb*src_bstride ([R][C], row stride ld) and writes
// dst[plane][c][bb*dst_boff + r] for r < R_ld (zeros for r >= R); dst row stride dst_ld, plane stride dst_plane elements.
int transpose_split_launch(const float* src, long long ld, long long src_bstride, __nv_bfloat16* dst, int R, int C, int R_ld,
                           long long dst_ld, long long dst_boff, int nbatch, long long dst_plane, cudaStream_t st, int row_shift = 0);
// row_shift: dst column r holds src row r + row_shift (zeros outside [0,R)) -- a TMA box may only start at a 16-byte
// aligned element of the innermost dimension, so conv-tap / recurrence shifts are materialised as shifted copies.
int transpose_bf16_launch(const __nv_bfloat16* src, long long src_ld, long long src_bstride, long long src_plane, __nv_bfloat16* dst,
                          int R, int C, int R_ld, long long dst_ld, long long dst_boff, int nbatch, int row_shift, long long dst_plane,
                          cudaStream_t st);
