// HiFi-GAN generator: the vocoder the reference runs right after the acoustic model
// (model/hifi_gan.py:154-216 Generator.forward, ResBlock1 :20-112, ResBlock2 :115-151; call sites run/say.py:153-157, run/test.py:170).
//   mel (B, 80, T) -> conv_pre k7 -> 4 x [ leaky_relu(0.1) -> ConvTranspose1d(stride u) -> mean of the multi-receptive-field
//   resblocks ] -> leaky_relu(0.01) -> conv_post k7 -> tanh -> wav (B, 1, T * prod(u))
// Everything is channels-last and runs on the implicit-GEMM tensor-core conv of convtc.cu (tcgen05, TMA out-of-bounds zero fill =
// padding, three bf16 (hi,lo) passes in fp32-parity mode):
//   * dilated convs: the TMA time coordinate of tap j is t + (j - (k-1)/2) * dilation;
//   * ConvTranspose1d(Cin -> Cout, k, stride u, padding (k-u)/2) = ONE conv with u * Cout output channels (phase r = output position mod
//     u is the channel group): out[u s + r][co] = sum_delta in[s + delta] . W[:, co, r + p - u delta], so the output tensor
//     [B][S][u * Cout] IS the upsampled channels-last tensor [B][S u][Cout]; taps outside [0, k) are zero weights;
//   * every leaky_relu is applied by the PRODUCER's epilogue to the bf16 copy the next conv reads (the fp32 copy keeps the value for the
//     residual), the residual add, the 1/num_kernels scaling and the running sum over the resblocks are in the last conv's epilogue.
// Weights are the reference's state dict after Generator.remove_weight_norm() (run/say.py:88): conv_pre.{weight,bias}, ups.i.*,
// resblocks.n.convs1.m.*, resblocks.n.convs2.m.* (ResBlock2: resblocks.n.convs.m.*), conv_post.*.
#include "../../include/t2b200.h"
#include "common.cuh"
#include "convtc.cuh"
#include "layers.cuh"

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

int make_w_map(CUtensorMap* m, const __nv_bfloat16* w, int Cout, int Cin, int k, int NT, int npass);   // capi.cu
int make_in_map(CUtensorMap* m, const __nv_bfloat16* x, int B, int T, int Cin, int npass);

namespace {

struct HParam { const float* ptr; long long numel; };
struct HConv {   // one packed conv
  __nv_bfloat16* w;    // [2][Cout'][k'][Cin]
  const float* bias;   // [Cout] (registered pointer) or null
  CUtensorMap wmap;
  int Cin, Cout, k, dil, NT, chan_mod;   // Cout = packed output channels (u * Cout for a transposed conv, 16 for conv_post)
};

// ConvTranspose1d weight (Cin, Cout, k) -> conv pack [2][u*Cout][kp][Cin]: row n = r*Cout + co, tap jp <-> delta = jp - (kp-1)/2
__global__ void pack_convT_w_kernel(const float* W, __nv_bfloat16* dst, int Cin, int Cout, int k, int u, int p, int kp) {
  const long long n = (long long)u * Cout * kp * Cin;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int jp = (int)((i / Cin) % kp);
    const int row = (int)(i / ((long long)Cin * kp));
    const int r = row / Cout, co = row % Cout;
    const int j = r + p - u * (jp - (kp - 1) / 2);
    const float v = (j >= 0 && j < k) ? W[((size_t)ci * Cout + co) * k + j] : 0.f;
    __nv_bfloat16 hi, lo;
    tc::split_bf16(v, hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}
// conv weight (Cout, Cin, k) with Cout padded to CoutP rows of zeros -> [2][CoutP][k][Cin]
__global__ void pack_conv_w_pad_kernel(const float* W, __nv_bfloat16* dst, int Cout, int CoutP, int Cin, int k) {
  const long long n = (long long)CoutP * Cin * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int j = (int)((i / Cin) % k);
    const int co = (int)(i / ((long long)Cin * k));
    __nv_bfloat16 hi, lo;
    tc::split_bf16(co < Cout ? W[((size_t)co * Cin + ci) * k + j] : 0.f, hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}
// mel (B, M, T) fp32 -> channels-last bf16 (hi,lo) [2][B][T][M]
__global__ void mel_to_cl_kernel(const float* mel, __nv_bfloat16* dst, int B, int M, int T) {
  const long long n = (long long)B * T * M;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % M);
    const long long p = i / M;
    const int b = (int)(p / T), t = (int)(p % T);
    __nv_bfloat16 hi, lo;
    tc::split_bf16(mel[((size_t)b * M + c) * T + t], hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}
inline int gsz(long long n) { return (int)std::min<long long>((n + 255) / 256, 148 * 16); }

}  // namespace

struct t2_hifigan {
  t2_hifigan_config cfg;
  int npass;
  std::map<std::string, HParam> params;
  bool committed;
  char* slab;
  size_t slab_bytes;
  HConv pre, post;
  std::vector<HConv> ups;
  std::vector<HConv> rb;     // resblock convs in execution order
  int n_up, n_kern, n_dil;   // n_dil: dilations per resblock (3 for "1", 2 for "2")
};

static const HParam* hfind(const t2_hifigan* g, const std::string& k, long long numel) {
  auto it = g->params.find(k);
  return (it == g->params.end() || it->second.numel != numel) ? nullptr : &it->second;
}

extern "C" int t2_hifigan_create(const t2_hifigan_config* cfg, t2_hifigan** out) {
  T2_REQUIRE(cfg && out, T2_ERR_ARG, "t2_hifigan_create: null argument");
  T2_REQUIRE(cfg->num_upsamples >= 1 && cfg->num_upsamples <= 8 && cfg->num_kernels >= 1 && cfg->num_kernels <= 8, T2_ERR_ARG,
             "t2_hifigan_create: bad layer counts");
  T2_REQUIRE(cfg->resblock == 1 || cfg->resblock == 2, T2_ERR_ARG, "t2_hifigan_create: resblock must be 1 or 2");
  T2_REQUIRE(cfg->num_mels % 8 == 0 && cfg->upsample_initial_channel % (16 << cfg->num_upsamples) == 0, T2_ERR_UNSUPPORTED,
             "t2_hifigan_create: channels must stay multiples of 16 (initial %d, %d upsamples)", cfg->upsample_initial_channel, cfg->num_upsamples);
  int dev = 0;
  cudaDeviceProp prop;
  T2_CHECK_CUDA(cudaGetDevice(&dev));
  T2_CHECK_CUDA(cudaGetDeviceProperties(&prop, dev));
  T2_REQUIRE(prop.major == 10, T2_ERR_UNSUPPORTED, "t2_hifigan_create: device sm_%d%d; libt2b200 is built for sm_100a only", prop.major, prop.minor);
  for (int i = 0; i < cfg->num_upsamples; ++i) {
    const int k = cfg->upsample_kernel_sizes[i], u = cfg->upsample_rates[i];
    T2_REQUIRE(u >= 1 && k >= u && (k - u) % 2 == 0, T2_ERR_UNSUPPORTED, "t2_hifigan_create: upsample %d: kernel %d, rate %d", i, k, u);
  }
  for (int j = 0; j < cfg->num_kernels; ++j) T2_REQUIRE(cfg->resblock_kernel_sizes[j] % 2 == 1, T2_ERR_UNSUPPORTED, "resblock kernel sizes must be odd");
  t2_hifigan* g = new t2_hifigan();
  g->cfg = *cfg;
  g->npass = cfg->precision == 1 ? 1 : 3;
  g->committed = false;
  g->slab = nullptr;
  g->slab_bytes = 0;
  g->n_up = cfg->num_upsamples;
  g->n_kern = cfg->num_kernels;
  g->n_dil = cfg->resblock == 1 ? 3 : 2;
  *out = g;
  return T2_OK;
}

extern "C" int t2_hifigan_destroy(t2_hifigan* g) {
  if (!g) return T2_OK;
  if (g->slab) cudaFree(g->slab);
  delete g;
  return T2_OK;
}

extern "C" int t2_hifigan_set_param(t2_hifigan* g, const char* key, const float* ptr, long long numel) {
  T2_REQUIRE(g && key && ptr, T2_ERR_ARG, "t2_hifigan_set_param: null argument");
  g->params[key] = HParam{ptr, numel};
  g->committed = false;
  return T2_OK;
}

static int nt_for(int C) { return C % 128 == 0 ? 128 : (C <= 128 ? C : (C % 64 == 0 ? 64 : 16)); }

static int hifigan_commit_impl(t2_hifigan* g, Arena& ar, cudaStream_t st) {
  const bool sizing = ar.base == nullptr;
  const t2_hifigan_config& c = g->cfg;
  const int M = c.num_mels, C0 = c.upsample_initial_channel;
  auto fail_key = [&](const std::string& k) { t2_set_error("t2_hifigan_commit: parameter %s missing or wrong size", k.c_str()); return T2_ERR_STATE; };
  auto conv = [&](HConv& h, const std::string& name, int Cin, int Cout, int k, int dil) -> int {
    h.Cin = Cin; h.Cout = Cout; h.k = k; h.dil = dil; h.NT = nt_for(Cout); h.chan_mod = 0; h.bias = nullptr;
    h.w = ar.take<__nv_bfloat16>((size_t)2 * Cout * k * Cin);
    if (sizing) return T2_OK;
    const HParam* w = hfind(g, name + ".weight", (long long)Cout * Cin * k);
    const HParam* b = hfind(g, name + ".bias", Cout);
    if (!w) return fail_key(name + ".weight");
    if (!b) return fail_key(name + ".bias");
    h.bias = b->ptr;
    RC(pack_conv_w_launch(w->ptr, h.w, Cout, Cin, k, st));
    return make_w_map(&h.wmap, h.w, Cout, Cin, k, h.NT, g->npass);
  };
  RC(conv(g->pre, "conv_pre", M, C0, 7, 1));
  g->ups.assign(g->n_up, HConv{});
  g->rb.clear();
  int ch = C0;
  for (int i = 0; i < g->n_up; ++i) {
    const int k = c.upsample_kernel_sizes[i], u = c.upsample_rates[i], p = (k - u) / 2, Cin = ch, Cout = ch / 2;
    // taps delta with j = r + p - u*delta in [0,k) for some phase r in [0,u)
    const int dmax = (u - 1 + p) / u, dmin_neg = (k - 1 - p + u - 1) / u;   // |ceil((p - k + 1) / u)|
    const int kp = 2 * std::max(dmax, dmin_neg) + 1;
    HConv& h = g->ups[i];
    h.Cin = Cin; h.Cout = u * Cout; h.k = kp; h.dil = 1; h.NT = nt_for(u * Cout); h.chan_mod = Cout; h.bias = nullptr;
    T2_REQUIRE((u * Cout) % h.NT == 0 && h.NT % 16 == 0, T2_ERR_UNSUPPORTED, "t2_hifigan_commit: upsample %d: %d packed output channels", i, u * Cout);
    h.w = ar.take<__nv_bfloat16>((size_t)2 * u * Cout * kp * Cin);
    if (!sizing) {
      const std::string name = "ups." + std::to_string(i);
      const HParam* w = hfind(g, name + ".weight", (long long)Cin * Cout * k);
      const HParam* b = hfind(g, name + ".bias", Cout);
      if (!w) return fail_key(name + ".weight");
      if (!b) return fail_key(name + ".bias");
      h.bias = b->ptr;
      pack_convT_w_kernel<<<gsz((long long)u * Cout * kp * Cin), 256, 0, st>>>(w->ptr, h.w, Cin, Cout, k, u, p, kp);
      T2_CHECK_CUDA(cudaGetLastError());
      RC(make_w_map(&h.wmap, h.w, u * Cout, Cin, kp, h.NT, g->npass));
    }
    ch = Cout;
    for (int j = 0; j < g->n_kern; ++j) {
      const std::string rbn = "resblocks." + std::to_string(i * g->n_kern + j);
      const int kk = c.resblock_kernel_sizes[j];
      for (int m = 0; m < g->n_dil; ++m) {
        const int dil = c.resblock_dilation_sizes[j][m];
        T2_REQUIRE(dil >= 1, T2_ERR_ARG, "t2_hifigan_commit: dilation");
        if (c.resblock == 1) {
          HConv a1{}, a2{};
          RC(conv(a1, rbn + ".convs1." + std::to_string(m), ch, ch, kk, dil));
          RC(conv(a2, rbn + ".convs2." + std::to_string(m), ch, ch, kk, 1));
          g->rb.push_back(a1); g->rb.push_back(a2);
        } else {
          HConv a1{};
          RC(conv(a1, rbn + ".convs." + std::to_string(m), ch, ch, kk, dil));
          g->rb.push_back(a1);
        }
      }
    }
  }
  {  // conv_post: one real output channel in a 16-row pack
    HConv& h = g->post;
    h.Cin = ch; h.Cout = 16; h.k = 7; h.dil = 1; h.NT = 16; h.chan_mod = 0; h.bias = nullptr;
    h.w = ar.take<__nv_bfloat16>((size_t)2 * 16 * 7 * ch);
    float* bias16 = ar.take<float>(16);
    if (!sizing) {
      const HParam* w = hfind(g, "conv_post.weight", (long long)ch * 7);
      const HParam* b = hfind(g, "conv_post.bias", 1);
      if (!w) return fail_key("conv_post.weight");
      if (!b) return fail_key("conv_post.bias");
      pack_conv_w_pad_kernel<<<gsz(16ll * 7 * ch), 256, 0, st>>>(w->ptr, h.w, 1, 16, ch, 7);
      T2_CHECK_CUDA(cudaMemsetAsync(bias16, 0, 16 * sizeof(float), st));
      T2_CHECK_CUDA(cudaMemcpyAsync(bias16, b->ptr, sizeof(float), cudaMemcpyDeviceToDevice, st));
      h.bias = bias16;
      T2_CHECK_CUDA(cudaGetLastError());
      RC(make_w_map(&h.wmap, h.w, 16, ch, 7, 16, g->npass));
    }
  }
  if (!sizing) {
    T2_REQUIRE(!ar.overflow, T2_ERR_STATE, "t2_hifigan_commit: slab overflow");
    g->committed = true;
  }
  return T2_OK;
}

extern "C" int t2_hifigan_commit(t2_hifigan* g, void* stream) {
  T2_REQUIRE(g, T2_ERR_ARG, "t2_hifigan_commit: null context");
  if (!g->slab) {
    Arena sizing(nullptr, 0);
    RC(hifigan_commit_impl(g, sizing, nullptr));
    g->slab_bytes = (size_t)sizing.off;
    T2_CHECK_CUDA(cudaMalloc((void**)&g->slab, g->slab_bytes));
  }
  Arena ar(g->slab, (int64_t)g->slab_bytes);
  return hifigan_commit_impl(g, ar, (cudaStream_t)stream);
}

static int hconv(const t2_hifigan* g, const HConv& h, const __nv_bfloat16* in, int B, int T, int act, float* out_f32, __nv_bfloat16* out_b16,
                 float b16_slope, const float* residual, float out_scale, const float* accum, float* out_c0, cudaStream_t st) {
  ConvTcArgs a{};
  RC(make_in_map(&a.in_map, in, B, T, h.Cin, g->npass));
  a.w_map = h.wmap;
  a.B = B; a.T = T; a.Cin = h.Cin; a.Cout = h.Cout; a.k = h.k; a.NT = h.NT; a.npass = g->npass; a.act = act;
  a.scale = nullptr; a.shift = h.bias; a.keep = nullptr; a.keep_ld = 0; a.drop_scale = 1.f;
  a.out_f32 = out_f32; a.out_b16 = out_b16; a.residual = residual; a.res_bstride = (long long)T * h.Cout;
  a.dil = h.dil; a.chan_mod = h.chan_mod; a.out_scale = out_scale; a.accum = accum; a.b16_slope = b16_slope; a.out_c0 = out_c0;
  if (convtc_halo_ok(h.Cin, h.Cout, h.NT, h.k, h.dil)) {
    a.halo = 1;
    RC(convtc_make_halo_maps(a, in, B, T, h.Cin));
  }
  return convtc_launch(a, st);
}

static int hifigan_fwd_impl(t2_hifigan* g, const float* mel, int B, int T, float* wav, Arena& ar, cudaStream_t st) {
  const t2_hifigan_config& c = g->cfg;
  const int M = c.num_mels, C0 = c.upsample_initial_channel;
  long long nmax = (long long)T * C0, len = T;
  {
    int ch = C0;
    for (int i = 0; i < g->n_up; ++i) { len *= c.upsample_rates[i]; ch /= 2; nmax = std::max(nmax, len * ch); }
  }
  const long long Tw = len;
  nmax *= B;
  __nv_bfloat16* melb = ar.take<__nv_bfloat16>((size_t)2 * B * T * M);
  float* X = ar.take<float>(nmax);
  float* Y = ar.take<float>(nmax);
  float* XS = ar.take<float>(nmax);
  __nv_bfloat16* XA = ar.take<__nv_bfloat16>(2 * nmax);
  __nv_bfloat16* YA = ar.take<__nv_bfloat16>(2 * nmax);
  __nv_bfloat16* T1 = ar.take<__nv_bfloat16>(2 * nmax);
  __nv_bfloat16* XSA = ar.take<__nv_bfloat16>(2 * nmax);
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_hifigan_fwd: workspace too small");
  const float S = 0.1f;   // LRELU_SLOPE (model/hifi_gan.py:7)
  mel_to_cl_kernel<<<gsz((long long)B * T * M), 256, 0, st>>>(mel, melb, B, M, T);
  T2_CHECK_CUDA(cudaGetLastError());
  // conv_pre; its consumer is leaky_relu -> ups[0]
  RC(hconv(g, g->pre, melb, B, T, ACT_NONE, nullptr, XSA, S, nullptr, 0.f, nullptr, nullptr, st));
  int Ti = T, ch = C0;
  size_t ri = 0;
  const float inv_k = 1.0f / (float)g->n_kern;
  for (int i = 0; i < g->n_up; ++i) {
    const int u = c.upsample_rates[i];
    // ConvTranspose1d as one conv with u * Cout channels at the input rate: the output [B][Ti][u*Cout] is [B][Ti*u][Cout]
    RC(hconv(g, g->ups[i], XSA, B, Ti, ACT_NONE, X, XA, S, nullptr, 0.f, nullptr, nullptr, st));
    Ti *= u; ch /= 2;
    const bool last_stage = i + 1 == g->n_up;
    const float next_slope = last_stage ? 0.01f : S;   // F.leaky_relu(x) before conv_post uses the default slope (model/hifi_gan.py:205)
    for (int j = 0; j < g->n_kern; ++j) {
      const float* cur_x = X;
      const __nv_bfloat16* cur_xa = XA;
      for (int m = 0; m < g->n_dil; ++m) {
        const bool last = m + 1 == g->n_dil;
        const __nv_bfloat16* c2_in = cur_xa;
        if (c.resblock == 1) {   // xt = c1(lrelu(x)); xt = c2(lrelu(xt)); x = xt + x   (model/hifi_gan.py:96-103)
          RC(hconv(g, g->rb[ri++], cur_xa, B, Ti, ACT_NONE, nullptr, T1, S, nullptr, 0.f, nullptr, nullptr, st));
          c2_in = T1;
        }
        const HConv& h2 = g->rb[ri++];
        if (!last) {
          RC(hconv(g, h2, c2_in, B, Ti, ACT_NONE, Y, YA, S, cur_x, 0.f, nullptr, nullptr, st));
          cur_x = Y; cur_xa = YA;
        } else {   // resblock output -> running mean over the kernels (model/hifi_gan.py:197-203)
          const bool fin = j + 1 == g->n_kern;
          RC(hconv(g, h2, c2_in, B, Ti, ACT_NONE, XS, fin ? XSA : nullptr, fin ? next_slope : 0.f, cur_x, inv_k, j ? XS : nullptr, nullptr, st));
        }
      }
    }
  }
  // conv_post + tanh; channel 0 of the 16-wide tile is the waveform
  RC(hconv(g, g->post, XSA, B, Ti, ACT_TANH, nullptr, nullptr, 0.f, nullptr, 0.f, nullptr, wav, st));
  (void)Tw;
  return T2_OK;
}

extern "C" long long t2_hifigan_workspace_bytes(const t2_hifigan* gc, int B, int T) {
  t2_hifigan* g = const_cast<t2_hifigan*>(gc);
  if (!g || B <= 0 || T <= 0) return -1;
  Arena a(nullptr, 0);
  hifigan_fwd_impl(g, nullptr, B, T, nullptr, a, nullptr);
  return a.off + 1024;
}

extern "C" long long t2_hifigan_out_len(const t2_hifigan* g, int T) {
  if (!g || T <= 0) return -1;
  long long len = T;
  for (int i = 0; i < g->n_up; ++i) len *= g->cfg.upsample_rates[i];
  return len;
}

extern "C" int t2_hifigan_fwd(t2_hifigan* g, const float* mel, int B, int T, float* wav, void* ws, long long ws_bytes, void* stream) {
  T2_REQUIRE(g && mel && wav && ws && B > 0 && T > 0, T2_ERR_ARG, "t2_hifigan_fwd: bad argument");
  T2_REQUIRE(g->committed, T2_ERR_STATE, "t2_hifigan_fwd: weights not committed (call t2_hifigan_commit)");
  Arena ar(ws, ws_bytes);
  return hifigan_fwd_impl(g, mel, B, T, wav, ar, (cudaStream_t)stream);
}
