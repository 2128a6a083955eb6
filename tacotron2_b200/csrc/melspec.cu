// Batched log-mel extraction: the arithmetic of torchaudio MelSpectrogram(n_fft 1024, win 1024, hop 256,
// center=True, pad_mode="reflect", periodic Hann, power 1, slaney mel scale + norm) followed by log(clamp(min=1e-5)) and a
// transpose to (T, n_mels) -- the parameter statement of the reference (datasets/prosody_dataset.py:39-50,66-67; call site
// datasets/tts_dataset.py:166-168,204; inverse run/say.py:161-171).
//
// One CTA = 256 threads = MS_FPC consecutive frames of one utterance.  The samples the frames share (75 % overlap) are
// staged once in shared memory with the reflect padding applied; two real frames are packed into one 1024-point complex
// radix-4 Stockham FFT (5 passes, twiddles from a shared-memory table), unpacked to |X_a|, |X_b| for bins 0..512, and the
// triangular mel filters are applied by one THREAD per (frame, mel) over the filter's [lo,hi) bin range, weights from a compact
// shared-memory copy of the non-zero part of the filterbank (the kernel is instruction-issue bound: a warp per filter spent 45
// instructions, most of them the warp reduction, on 6..40 multiply-adds).
#include "../../include/t2b200.h"
#include "common.cuh"

#define MS_NFFT 1024
#define MS_HOP 256
#define MS_BINS 513
#define MS_FPC 8          // frames per CTA (4 packed FFTs)
#define MS_THREADS 256
#ifndef MS_NB
#define MS_NB 4           // frame blocks per CTA: the twiddle table and the compact filterbank are built once for 32 frames (2, 3, 4, 6
#endif                    // measured within 8 %: 4 is the best at 64 x 800 frames)
#define MS_FBC 1024       // capacity of the compact filterbank (sum of the filters' bin ranges; slaney 80 x 513: ~760)
#define MS_MAXMEL 128

namespace {

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

__global__ void __launch_bounds__(MS_THREADS) melspec_kernel(const float* __restrict__ wav, const long long* __restrict__ wav_len,
                                                             long long wav_stride, const float* __restrict__ window,
                                                             const float* __restrict__ fb /*[n_mels][513]*/,
                                                             const int* __restrict__ fb_range /*[n_mels][2]*/, int n_mels,
                                                             float* __restrict__ out, int t_stride, long long* __restrict__ n_frames_out,
                                                             float log_floor) {
  __shared__ float2 bufA[MS_NFFT];
  __shared__ float2 bufB[MS_NFFT];
  __shared__ float2 tw[3 * MS_NFFT / 4];                          // exp(-2 pi i m / 1024), m < 768 (largest index used: 3 * 255)
  __shared__ float samples[MS_NFFT + (MS_FPC - 1) * MS_HOP];
  __shared__ float mag[2][MS_BINS + 3];
  __shared__ float fbc[MS_FBC];            // non-zero ranges of the filters, back to back
  __shared__ int fb_off[MS_MAXMEL + 1];    // start of filter m in fbc
  __shared__ int fb_lo[MS_MAXMEL];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.y;
  const long long S = wav_len[b];
  const int T = S > 0 ? (int)(1 + S / MS_HOP) : 0;
  if (blockIdx.x == 0 && tid == 0 && n_frames_out) n_frames_out[b] = T;
  float* outb = out + (size_t)b * t_stride * n_mels;
  const int fblk0 = blockIdx.x * MS_NB * MS_FPC;
  if (fblk0 >= T) {  // pad_sequence zero padding of the batch (datasets/tts_dataloader.py:28)
    for (int i = tid; i < MS_NB * MS_FPC * n_mels; i += MS_THREADS) {
      const int f = fblk0 + i / n_mels;
      if (f < t_stride) outb[(size_t)f * n_mels + i % n_mels] = 0.f;
    }
    return;
  }
  {  // exp(-2 pi i m / 1024): one accurate sincos per thread (m < 256), the next two quadrants by rotation (x -i, x -1)
    float sn, cs;
    sincospif(-2.0f * (float)tid / (float)MS_NFFT, &sn, &cs);
    tw[tid] = make_float2(cs, sn);
    tw[tid + 256] = make_float2(sn, -cs);
    tw[tid + 512] = make_float2(-cs, -sn);
  }
  // compact filterbank: filter m occupies fbc[fb_off[m] .. fb_off[m+1]) = fb[m][lo_m .. hi_m)
  if (warp == 0) {
    int run = 0;
    for (int m0 = 0; m0 < n_mels; m0 += 32) {
      const int m = m0 + lane;
      const int lo = m < n_mels ? fb_range[2 * m] : 0, n = m < n_mels ? max(0, fb_range[2 * m + 1] - lo) : 0;
      int inc = n;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
      }
      if (m < n_mels) { fb_off[m] = run + inc - n; fb_lo[m] = lo; }
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) fb_off[n_mels] = run;
  }
  __syncthreads();
  const bool fb_compact = fb_off[n_mels] <= MS_FBC;
  if (fb_compact) {
    for (int m = warp; m < n_mels; m += MS_THREADS / 32) {
      const int o = fb_off[m], n = fb_off[m + 1] - o, lo = fb_lo[m];
      for (int i = lane; i < n; i += 32) fbc[o + i] = __ldg(fb + (size_t)m * MS_BINS + lo + i);
    }
  }
  const float* x = wav + (size_t)b * wav_stride;
#pragma unroll 1
  for (int blk = 0; blk < MS_NB; ++blk) {
  const int f0 = fblk0 + blk * MS_FPC;
  if (f0 >= t_stride) break;
  if (f0 >= T) {   // beyond this utterance: zero rows
    for (int i = tid; i < MS_FPC * n_mels; i += MS_THREADS) {
      const int f = f0 + i / n_mels;
      if (f < t_stride) outb[(size_t)f * n_mels + i % n_mels] = 0.f;
    }
    continue;
  }
  const long long base = (long long)f0 * MS_HOP - MS_NFFT / 2;
  for (int i = tid; i < MS_NFFT + (MS_FPC - 1) * MS_HOP; i += MS_THREADS) {
    long long n = base + i;
    if (n < 0) n = -n;                         // reflect (no edge repeat), torch.stft center=True
    if (n >= S) n = 2 * (S - 1) - n;
    samples[i] = (n >= 0 && n < S) ? x[n] : 0.f;
  }
  __syncthreads();

  for (int pair = 0; pair < MS_FPC / 2; ++pair) {
    const int fa = f0 + 2 * pair;
    if (fa >= T && fa >= t_stride) break;
    // pack frame a (real) and frame b (imag)
    for (int i = tid; i < MS_NFFT; i += MS_THREADS) {
      const float w = __ldg(window + i);   // 4 KB, L1-resident
      bufA[i] = make_float2(samples[(2 * pair) * MS_HOP + i] * w, samples[(2 * pair + 1) * MS_HOP + i] * w);
    }
    __syncthreads();
    float2* src = bufA;
    float2* dst = bufB;
#pragma unroll
    for (int pass = 0; pass < 5; ++pass) {
      const int Ns = 1 << (2 * pass);
      const int j = tid, k = j & (Ns - 1);
      const int tstep = k * (MS_NFFT / (Ns * 4));  // twiddle index for r = 1
      float2 v0 = src[j], v1 = src[j + 256], v2 = src[j + 512], v3 = src[j + 768];
      if (pass > 0) {
        v1 = cmul(v1, tw[tstep]);
        v2 = cmul(v2, tw[2 * tstep]);
        v3 = cmul(v3, tw[3 * tstep]);
      }
      const float2 a = make_float2(v0.x + v2.x, v0.y + v2.y), bb = make_float2(v0.x - v2.x, v0.y - v2.y);
      const float2 c = make_float2(v1.x + v3.x, v1.y + v3.y), d = make_float2(v1.x - v3.x, v1.y - v3.y);
      const float2 dr = make_float2(d.y, -d.x);  // -i * d
      const int o = (j / Ns) * Ns * 4 + k;
      dst[o] = make_float2(a.x + c.x, a.y + c.y);
      dst[o + Ns] = make_float2(bb.x + dr.x, bb.y + dr.y);
      dst[o + 2 * Ns] = make_float2(a.x - c.x, a.y - c.y);
      dst[o + 3 * Ns] = make_float2(bb.x - dr.x, bb.y - dr.y);
      __syncthreads();
      float2* t = src; src = dst; dst = t;
    }
    // src holds Z = FFT(a + i b) in natural order: X_a[k] = (Z[k] + conj Z[N-k]) / 2, X_b[k] = (Z[k] - conj Z[N-k]) / (2i)
    for (int k = tid; k < MS_BINS; k += MS_THREADS) {
      const float2 z = src[k], zc = src[(MS_NFFT - k) & (MS_NFFT - 1)];
      const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);
      const float br = 0.5f * (z.y + zc.y), bi = -0.5f * (z.x - zc.x);
      mag[0][k] = sqrtf(ar * ar + ai * ai);   // power = 1: magnitude
      mag[1][k] = sqrtf(br * br + bi * bi);
    }
    __syncthreads();
    if (fb_compact) {   // thread = (frame of the pair, mel): consecutive threads write consecutive mels of one output row
      for (int o = tid; o < 2 * n_mels; o += MS_THREADS) {
        const int which = o / n_mels, m = o - which * n_mels;
        const int f = fa + which;
        const int off = fb_off[m], n = fb_off[m + 1] - off;
        const float* mg = mag[which] + fb_lo[m];
        float acc = 0.f;
        for (int i = 0; i < n; ++i) acc = fmaf(mg[i], fbc[off + i], acc);
        if (f < t_stride) outb[(size_t)f * n_mels + m] = f < T ? logf(fmaxf(acc, log_floor)) : 0.f;
      }
    } else {            // a filterbank denser than the compact copy holds: one warp per (frame, mel), weights from global memory
      for (int o = warp; o < 2 * n_mels; o += MS_THREADS / 32) {
        const int which = o / n_mels, m = o - which * n_mels;
        const int f = fa + which;
        const int lo = fb_range[2 * m], hi = fb_range[2 * m + 1];
        float acc = 0.f;
        for (int k = lo + lane; k < hi; k += 32) acc = fmaf(mag[which][k], __ldg(fb + (size_t)m * MS_BINS + k), acc);
        acc = warp_sum(acc);
        if (lane == 0 && f < t_stride) outb[(size_t)f * n_mels + m] = f < T ? logf(fmaxf(acc, log_floor)) : 0.f;
      }
    }
    __syncthreads();
  }
  }  // frame blocks
}

}  // namespace

extern "C" int t2_melspec(const float* wav, const long long* wav_len, int B, long long wav_stride, const float* window,
                          const float* fb, const int* fb_range, int n_fft, int hop, int n_mels, float* out, int t_stride,
                          long long* n_frames, void* stream) {
  T2_REQUIRE(wav && wav_len && window && fb && fb_range && out, T2_ERR_ARG, "t2_melspec: null argument");
  T2_REQUIRE(n_fft == MS_NFFT && hop == MS_HOP, T2_ERR_UNSUPPORTED,
             "t2_melspec: specialised for n_fft=1024, hop=256 (datasets/prosody_dataset.py:39-50), got %d/%d", n_fft, hop);
  T2_REQUIRE(B > 0 && n_mels > 0 && n_mels <= MS_MAXMEL && t_stride > 0 && wav_stride > 0, T2_ERR_ARG, "t2_melspec: bad sizes (n_mels <= %d)", MS_MAXMEL);
  dim3 grid((unsigned)((t_stride + MS_NB * MS_FPC - 1) / (MS_NB * MS_FPC)), (unsigned)B);
  melspec_kernel<<<grid, MS_THREADS, 0, (cudaStream_t)stream>>>(wav, wav_len, wav_stride, window, fb, fb_range, n_mels, out, t_stride,
                                                                n_frames, 1e-5f);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
