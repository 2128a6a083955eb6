// Leading / trailing silence trim of a waveform batch on the GPU: what
//   librosa.effects.trim(wav, top_db=trim_top_db, frame_length=trim_frame_length)          datasets/tts_dataset.py:194-200
// computes -- frame RMS over centred, zero-padded frames (hop 512 by default), in dB relative to the utterance's loudest frame
// (amin 1e-5), frames above -top_db are "non-silent", the kept interval is [first * hop, min(len, (last + 1) * hop)).
// librosa is an external dependency that is not vendored with the reference: parity is pinned to the CPU restatement in
// oracle/trim_oracle.py ("parity unpinned vs librosa").
// HBM-bound: every sample is read frame_length / hop (= 4) times through L1/L2, once from HBM.
#include "../../include/t2b200.h"
#include "common.cuh"
#include <algorithm>

namespace {

// one warp per frame: mean square over [f*hop - fl/2, f*hop + fl/2) with zeros outside [0, len)
__global__ void __launch_bounds__(256) trim_rms_kernel(const float* __restrict__ wav, const long long* __restrict__ wav_len, long long stride, int fl,
                                                       int hop, float* __restrict__ rms, long long max_frames) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const long long len = wav_len[b], nf = 1 + len / hop;
  const float* x = wav + (size_t)b * stride;
  for (long long f = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); f < nf; f += (long long)gridDim.x * 8) {
    const long long s0 = f * hop - fl / 2;
    float acc = 0.f;
    for (int i = lane; i < fl; i += 32) {
      const long long s = s0 + i;
      const float v = (s >= 0 && s < len) ? __ldg(x + s) : 0.f;
      acc = fmaf(v, v, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) rms[(size_t)b * max_frames + f] = sqrtf(acc / (float)fl);
  }
}

__global__ void __launch_bounds__(256) trim_bounds_kernel(const float* __restrict__ rms, const long long* __restrict__ wav_len, int hop, float top_db,
                                                          long long max_frames, long long* __restrict__ start, long long* __restrict__ end) {
  __shared__ float red[8];
  __shared__ long long lo_s[8], hi_s[8];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long len = wav_len[b], nf = 1 + len / hop;
  const float* r = rms + (size_t)b * max_frames;
  float mx = 0.f;
  for (long long f = tid; f < nf; f += 256) mx = fmaxf(mx, r[f]);
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
  const float amin = 1e-5f;
  const float ref_db = 20.0f * log10f(fmaxf(amin, mx));   // amplitude_to_db(ref=np.max, top_db=None)
  long long lo = nf, hi = -1;
  for (long long f = tid; f < nf; f += 256) {
    const float db = 20.0f * log10f(fmaxf(amin, r[f])) - ref_db;
    if (db > -top_db) { lo = f < lo ? f : lo; hi = f > hi ? f : hi; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const long long l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
    lo = l2 < lo ? l2 : lo; hi = h2 > hi ? h2 : hi;
  }
  if (lane == 0) { lo_s[warp] = lo; hi_s[warp] = hi; }
  __syncthreads();
  if (tid == 0) {
    for (int i = 1; i < 8; ++i) { lo = lo_s[i] < lo ? lo_s[i] : lo; hi = hi_s[i] > hi ? hi_s[i] : hi; }
    lo = lo_s[0] < lo ? lo_s[0] : lo; hi = hi_s[0] > hi ? hi_s[0] : hi;
    if (hi < 0) { start[b] = 0; end[b] = 0; }              // nothing above the threshold: librosa returns an empty signal
    else {
      const long long e = (hi + 1) * hop;
      start[b] = lo * hop;
      end[b] = e < len ? e : len;
    }
  }
}

}  // namespace

extern "C" long long t2_trim_workspace_bytes(int B, long long max_len, int hop_length) {
  if (B <= 0 || max_len < 0 || hop_length <= 0) return -1;
  return (long long)B * (1 + max_len / hop_length) * (long long)sizeof(float) + 256;
}

extern "C" int t2_trim(const float* wav, const long long* wav_len, int B, long long wav_stride, long long max_len, float top_db, int frame_length,
                       int hop_length, long long* start, long long* end, void* ws, long long ws_bytes, void* stream) {
  T2_REQUIRE(wav && wav_len && start && end && ws && B > 0 && frame_length > 0 && hop_length > 0 && max_len <= wav_stride, T2_ERR_ARG,
             "t2_trim: bad argument");
  const long long max_frames = 1 + max_len / hop_length;
  T2_REQUIRE(ws_bytes >= t2_trim_workspace_bytes(B, max_len, hop_length), T2_ERR_WORKSPACE, "t2_trim: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  float* rms = (float*)ws;
  const unsigned gx = (unsigned)std::min<long long>((max_frames + 7) / 8, 148 * 4);
  trim_rms_kernel<<<dim3(gx, (unsigned)B), 256, 0, st>>>(wav, wav_len, wav_stride, frame_length, hop_length, rms, max_frames);
  trim_bounds_kernel<<<B, 256, 0, st>>>(rms, wav_len, hop_length, top_db, max_frames, start, end);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
