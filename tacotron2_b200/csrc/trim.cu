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

// ---------------------------------------------------------------------------------------------------------------------------------
// t2_trim_static: t2_trim for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.  The frame-RMS scratch that t2_trim takes as a
// caller workspace lives in the file-scope device array below instead: T2_TRIM_STATIC_MAX_FRAMES floats = 1 MiB, held by every GPU
// that loads the library until the process exits.  Limits and sharing:
//   * B * (1 + max_len / hop_length) <= T2_TRIM_STATIC_MAX_FRAMES (262144: e.g. 64 utterances of 2.09 M samples = 95 s at 22.05 kHz
//     with librosa's hop of 512); larger requests are rejected with T2_ERR_WORKSPACE before any launch;
//   * every call on a device shares the one array, which keeps the previous call's contents.  Nothing in it has to start at zero:
//     trim_rms_kernel writes frames [0, 1 + len/hop) of every utterance and trim_bounds_kernel reads exactly those, so there is no
//     clearing pass;
//   * calls that overlap on the GPU (two streams, or two host threads on one device) race on the array: serialise them, or use
//     t2_trim with one workspace per stream.
// In host code `g_trim_static_rms` is only a host-side stand-in for the symbol, not a device address: the address handed to the
// kernels comes from cudaGetSymbolAddress, looked up once per device with that device current (first call on the device) and cached.
// The call allocates nothing: no cudaMalloc / cudaFree, no allocating library routine.
static_assert(T2_TRIM_STATIC_MAX_FRAMES == 262144, "the limit is stated in include/t2b200.h, DESIGN.md and INTEGRATION.md");
static __device__ __align__(256) float g_trim_static_rms[T2_TRIM_STATIC_MAX_FRAMES];

namespace {
constexpr int TRIM_STATIC_MAX_DEVICES = 64;
float* g_trim_static_ptr[TRIM_STATIC_MAX_DEVICES];   // per-device address of g_trim_static_rms (null = not looked up yet)

// address of the scratch array on the CURRENT device (set-up on the first call per device; racing first calls store the same value)
int trim_static_scratch(float** out) {
  int dev = -1;
  T2_CHECK_CUDA(cudaGetDevice(&dev));
  T2_REQUIRE(dev >= 0 && dev < TRIM_STATIC_MAX_DEVICES, T2_ERR_UNSUPPORTED, "t2_trim_static: device ordinal %d beyond %d", dev,
             TRIM_STATIC_MAX_DEVICES);
  float* p = __atomic_load_n(&g_trim_static_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    T2_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_trim_static_rms));
    T2_REQUIRE(p != nullptr, T2_ERR_CUDA, "t2_trim_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    __atomic_store_n(&g_trim_static_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return T2_OK;
}
}  // namespace

extern "C" int t2_trim_static(const float* wav, const long long* wav_len, int B, long long wav_stride, long long max_len, float top_db,
                              int frame_length, int hop_length, long long* start, long long* end, void* stream) {
  T2_REQUIRE(wav && wav_len && start && end && B > 0 && frame_length > 0 && hop_length > 0 && max_len >= 0 && max_len <= wav_stride,
             T2_ERR_ARG, "t2_trim_static: bad argument");
  const long long max_frames = 1 + max_len / hop_length;
  T2_REQUIRE(max_frames <= T2_TRIM_STATIC_MAX_FRAMES && (long long)B * max_frames <= T2_TRIM_STATIC_MAX_FRAMES, T2_ERR_WORKSPACE,
             "t2_trim_static: B=%d x %lld frames (max_len %lld, hop %d) = %lld frames exceeds the built-in scratch of %lld frames; "
             "use t2_trim with a caller workspace",
             B, max_frames, max_len, hop_length, (long long)B * max_frames, (long long)T2_TRIM_STATIC_MAX_FRAMES);
  float* rms = nullptr;
  const int rc = trim_static_scratch(&rms);
  if (rc != T2_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned gx = (unsigned)std::min<long long>((max_frames + 7) / 8, 148 * 4);
  trim_rms_kernel<<<dim3(gx, (unsigned)B), 256, 0, st>>>(wav, wav_len, wav_stride, frame_length, hop_length, rms, max_frames);
  trim_bounds_kernel<<<B, 256, 0, st>>>(rms, wav_len, hop_length, top_db, max_frames, start, end);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
