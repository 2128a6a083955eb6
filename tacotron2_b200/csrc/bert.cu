// Description-embedding front end: BertModel(...).pooler_output as run/say.py:93-106 computes it for a
// description string (google-bert/bert-base-uncased: 12 layers, hidden 768, 12 heads, intermediate 3072), batched.
//   embeddings (word + position + token type) -> LayerNorm -> 12 x [ self-attention -> dense + residual + LayerNorm ->
//   dense + GELU -> dense + residual + LayerNorm ] -> pooler: tanh(dense(hidden[:, 0]))
// Every dense layer is the k = 1 case of the tcgen05 implicit-GEMM conv (convtc.cu; bias, GELU / tanh and the residual add in its
// epilogue; three bf16 (hi,lo) passes in fp32-parity mode); LayerNorm, the embedding sum and the per-head attention (sequences of a
// few dozen tokens) are small CUDA-core kernels.  Parameters: the Hugging Face BertModel state-dict keys.  The tokenizer (WordPiece
// over the bert-base-uncased vocabulary file) is host string processing outside this library: the entry point takes token ids.
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

struct BParam { const float* ptr; long long numel; };
struct BLin { __nv_bfloat16* w; const float* bias; CUtensorMap wmap; int Cin, Cout; };

__device__ __forceinline__ float block_sum256(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  for (int i = 0; i < 8; ++i) r += red[i];
  return r;
}

// row = one token: x = (in_a + in_b) (either may be an embedding gather), y = LayerNorm(x) * gamma + beta -> fp32 + bf16 (hi,lo)
struct LnArgs {
  const float* a; const float* b;            // [N][H] each (b may be null)
  const long long* ids; const long long* type_ids; const float* word; const float* pos; const float* type; int S; int vocab; int max_pos; int type_vocab;
  const float* gamma; const float* beta; float eps; int N, H;
  float* out_f32; __nv_bfloat16* out_b16;    // [N][H], [2][N][H]
};
__global__ void __launch_bounds__(256) bert_ln_kernel(LnArgs p) {
  __shared__ float red[8];
  extern __shared__ float row[];
  const int n = blockIdx.x, H = p.H;
  float s = 0.f;
  for (int i = threadIdx.x; i < H; i += 256) {
    float v;
    if (p.ids) {   // embeddings: word[id] + position[s] + token_type[type]   (BertEmbeddings.forward)
      const long long id = p.ids[n], ty = p.type_ids ? p.type_ids[n] : 0;
      const int pos = n % p.S;
      v = ((id >= 0 && id < p.vocab) ? p.word[(size_t)id * H + i] : 0.f) + (pos < p.max_pos ? p.pos[(size_t)pos * H + i] : 0.f) +
          ((ty >= 0 && ty < p.type_vocab) ? p.type[(size_t)ty * H + i] : 0.f);
    } else {
      v = p.a[(size_t)n * H + i] + (p.b ? p.b[(size_t)n * H + i] : 0.f);
    }
    row[i] = v;
    s += v;
  }
  const float mean = block_sum256(s, red) / (float)H;
  float q = 0.f;
  for (int i = threadIdx.x; i < H; i += 256) { const float dlt = row[i] - mean; q += dlt * dlt; }
  const float var = block_sum256(q, red) / (float)H;   // biased variance, as torch.nn.LayerNorm
  const float rstd = rsqrtf(var + p.eps);
  const long long plane = (long long)p.N * H;
  for (int i = threadIdx.x; i < H; i += 256) {
    const float y = (row[i] - mean) * rstd * p.gamma[i] + p.beta[i];
    if (p.out_f32) p.out_f32[(size_t)n * H + i] = y;
    if (p.out_b16) {
      __nv_bfloat16 hi, lo;
      tc::split_bf16(y, hi, lo);
      p.out_b16[(size_t)n * H + i] = hi;
      p.out_b16[plane + (size_t)n * H + i] = lo;
    }
  }
}

// Self-attention of one (sequence, head): scores = Q K^T / sqrt(dh) + (1 - mask) * (-inf), softmax, context = P V.
// qkv [N][3H] fp32 (query | key | value blocks of H); one warp per query row; K and V of the head in shared memory.
__global__ void __launch_bounds__(256) bert_attention_kernel(const float* __restrict__ qkv, const float* __restrict__ mask, int S, int H, int heads,
                                                             __nv_bfloat16* __restrict__ ctx_b16, long long plane) {
  extern __shared__ float sh[];
  const int dh = H / heads, ld = dh + 1;
  float* Ks = sh;                  // [S][dh+1]
  float* Vs = Ks + (size_t)S * ld; // [S][dh+1]
  float* Ps = Vs + (size_t)S * ld; // [8 warps][S]
  const int b = blockIdx.x / heads, hd = blockIdx.x % heads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* base = qkv + (size_t)b * S * 3 * H;
  for (int i = threadIdx.x; i < S * dh; i += 256) {
    const int j = i / dh, e = i % dh;
    Ks[j * ld + e] = base[(size_t)j * 3 * H + H + hd * dh + e];
    Vs[j * ld + e] = base[(size_t)j * 3 * H + 2 * H + hd * dh + e];
  }
  __syncthreads();
  const float scale = rsqrtf((float)dh);
  float* P = Ps + (size_t)warp * S;
  for (int qi = warp; qi < S; qi += 8) {
    const float* q = base + (size_t)qi * 3 * H + hd * dh;
    float mx = -INFINITY;
    for (int j = lane; j < S; j += 32) {
      float acc = 0.f;
      for (int e = 0; e < dh; ++e) acc = fmaf(__ldg(q + e), Ks[j * ld + e], acc);
      acc = acc * scale + ((mask && mask[(size_t)b * S + j] == 0.f) ? -INFINITY : 0.f);
      P[j] = acc;
      mx = fmaxf(mx, acc);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < S; j += 32) { const float e_ = expf(P[j] - mx); P[j] = e_; sum += e_; }
    sum = warp_sum(sum);
    __syncwarp();
    const float inv = 1.0f / sum;
    for (int e = lane; e < dh; e += 32) {
      float acc = 0.f;
      for (int j = 0; j < S; ++j) acc = fmaf(P[j], Vs[j * ld + e], acc);
      acc *= inv;
      __nv_bfloat16 hi, lo;
      tc::split_bf16(acc, hi, lo);
      const size_t o = ((size_t)b * S + qi) * H + hd * dh + e;
      ctx_b16[o] = hi;
      ctx_b16[plane + o] = lo;
    }
    __syncwarp();
  }
}

// hidden[:, 0] of every sequence -> bf16 (hi,lo) [2][B][H]
__global__ void bert_cls_kernel(const float* x, __nv_bfloat16* dst, int B, int S, int H) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * H; i += gridDim.x * blockDim.x) {
    const int b = i / H, e = i % H;
    __nv_bfloat16 hi, lo;
    tc::split_bf16(x[(size_t)b * S * H + e], hi, lo);
    dst[i] = hi;
    dst[(size_t)B * H + i] = lo;
  }
}

}  // namespace

struct t2_bert {
  t2_bert_config cfg;
  int npass;
  std::map<std::string, BParam> params;
  bool committed;
  char* slab;
  size_t slab_bytes;
  std::vector<BLin> qkv, ao, ff1, ff2;
  std::vector<float*> qkv_bias;
  BLin pool;
};

static const BParam* bfind(const t2_bert* g, const std::string& k, long long numel) {
  auto it = g->params.find(k);
  return (it == g->params.end() || it->second.numel != numel) ? nullptr : &it->second;
}

extern "C" int t2_bert_create(const t2_bert_config* cfg, t2_bert** out) {
  T2_REQUIRE(cfg && out, T2_ERR_ARG, "t2_bert_create: null argument");
  T2_REQUIRE(cfg->hidden_size % 128 == 0 && cfg->intermediate_size % 128 == 0 && cfg->num_heads > 0 && cfg->hidden_size % cfg->num_heads == 0 &&
                 cfg->num_layers > 0, T2_ERR_UNSUPPORTED, "t2_bert_create: hidden %d / intermediate %d / heads %d", cfg->hidden_size,
             cfg->intermediate_size, cfg->num_heads);
  int dev = 0;
  cudaDeviceProp prop;
  T2_CHECK_CUDA(cudaGetDevice(&dev));
  T2_CHECK_CUDA(cudaGetDeviceProperties(&prop, dev));
  T2_REQUIRE(prop.major == 10, T2_ERR_UNSUPPORTED, "t2_bert_create: device sm_%d%d; libt2b200 is built for sm_100a only", prop.major, prop.minor);
  t2_bert* g = new t2_bert();
  g->cfg = *cfg;
  g->npass = cfg->precision == 1 ? 1 : 3;
  g->committed = false;
  g->slab = nullptr;
  g->slab_bytes = 0;
  *out = g;
  return T2_OK;
}
extern "C" int t2_bert_destroy(t2_bert* g) {
  if (!g) return T2_OK;
  if (g->slab) cudaFree(g->slab);
  delete g;
  return T2_OK;
}
extern "C" int t2_bert_set_param(t2_bert* g, const char* key, const float* ptr, long long numel) {
  T2_REQUIRE(g && key && ptr, T2_ERR_ARG, "t2_bert_set_param: null argument");
  g->params[key] = BParam{ptr, numel};
  g->committed = false;
  return T2_OK;
}

static int bert_commit_impl(t2_bert* g, Arena& ar, cudaStream_t st) {
  const bool sizing = ar.base == nullptr;
  const t2_bert_config& c = g->cfg;
  const int H = c.hidden_size, I = c.intermediate_size, NL = c.num_layers;
  auto fail_key = [&](const std::string& k) { t2_set_error("t2_bert_commit: parameter %s missing or wrong size", k.c_str()); return T2_ERR_STATE; };
  auto need = [&](const std::string& k, long long n, const float** out) -> int {
    const BParam* p = bfind(g, k, n);
    if (!p) return fail_key(k);
    *out = p->ptr;
    return T2_OK;
  };
  auto lin = [&](BLin& l, const float* w, const float* bias, int Cout, int Cin) -> int {
    l.Cin = Cin; l.Cout = Cout; l.bias = bias;
    if (sizing) return T2_OK;
    RC(pack_conv_w_launch(w, l.w, Cout, Cin, 1, st));
    return make_w_map(&l.wmap, l.w, Cout, Cin, 1, 128, g->npass);
  };
  float* stage = ar.take<float>((size_t)3 * H * H);   // q | k | v weights concatenated before packing
  g->qkv.assign(NL, BLin{}); g->ao.assign(NL, BLin{}); g->ff1.assign(NL, BLin{}); g->ff2.assign(NL, BLin{});
  g->qkv_bias.assign(NL, nullptr);
  for (int i = 0; i < NL; ++i) {
    const std::string p = "encoder.layer." + std::to_string(i) + ".";
    g->qkv[i].w = ar.take<__nv_bfloat16>((size_t)2 * 3 * H * H);
    g->qkv_bias[i] = ar.take<float>(3 * H);
    g->ao[i].w = ar.take<__nv_bfloat16>((size_t)2 * H * H);
    g->ff1[i].w = ar.take<__nv_bfloat16>((size_t)2 * I * H);
    g->ff2[i].w = ar.take<__nv_bfloat16>((size_t)2 * H * I);
    if (sizing) continue;
    const char* names[3] = {"query", "key", "value"};
    for (int j = 0; j < 3; ++j) {
      const float *w = nullptr, *b = nullptr;
      RC(need(p + "attention.self." + names[j] + ".weight", (long long)H * H, &w));
      RC(need(p + "attention.self." + names[j] + ".bias", H, &b));
      T2_CHECK_CUDA(cudaMemcpyAsync(stage + (size_t)j * H * H, w, (size_t)H * H * sizeof(float), cudaMemcpyDeviceToDevice, st));
      T2_CHECK_CUDA(cudaMemcpyAsync(g->qkv_bias[i] + j * H, b, H * sizeof(float), cudaMemcpyDeviceToDevice, st));
    }
    RC(lin(g->qkv[i], stage, g->qkv_bias[i], 3 * H, H));
    const float *w = nullptr, *b = nullptr;
    RC(need(p + "attention.output.dense.weight", (long long)H * H, &w)); RC(need(p + "attention.output.dense.bias", H, &b));
    RC(lin(g->ao[i], w, b, H, H));
    RC(need(p + "intermediate.dense.weight", (long long)I * H, &w)); RC(need(p + "intermediate.dense.bias", I, &b));
    RC(lin(g->ff1[i], w, b, I, H));
    RC(need(p + "output.dense.weight", (long long)H * I, &w)); RC(need(p + "output.dense.bias", H, &b));
    RC(lin(g->ff2[i], w, b, H, I));
  }
  g->pool.w = ar.take<__nv_bfloat16>((size_t)2 * H * H);
  if (!sizing) {
    const float *w = nullptr, *b = nullptr;
    RC(need("pooler.dense.weight", (long long)H * H, &w)); RC(need("pooler.dense.bias", H, &b));
    RC(lin(g->pool, w, b, H, H));
    T2_REQUIRE(!ar.overflow, T2_ERR_STATE, "t2_bert_commit: slab overflow");
    g->committed = true;
  }
  return T2_OK;
}

extern "C" int t2_bert_commit(t2_bert* g, void* stream) {
  T2_REQUIRE(g, T2_ERR_ARG, "t2_bert_commit: null context");
  if (!g->slab) {
    Arena sizing(nullptr, 0);
    RC(bert_commit_impl(g, sizing, nullptr));
    g->slab_bytes = (size_t)sizing.off;
    T2_CHECK_CUDA(cudaMalloc((void**)&g->slab, g->slab_bytes));
  }
  Arena ar(g->slab, (int64_t)g->slab_bytes);
  return bert_commit_impl(g, ar, (cudaStream_t)stream);
}

static int blin(const t2_bert* g, const BLin& l, const __nv_bfloat16* in, int N, int act, float* out_f32, __nv_bfloat16* out_b16, const float* residual,
                cudaStream_t st) {
  ConvTcArgs a{};
  RC(make_in_map(&a.in_map, in, 1, N, l.Cin, g->npass));
  a.w_map = l.wmap;
  a.B = 1; a.T = N; a.Cin = l.Cin; a.Cout = l.Cout; a.k = 1; a.NT = 128; a.npass = g->npass; a.act = act;
  a.shift = l.bias; a.drop_scale = 1.f; a.out_f32 = out_f32; a.out_b16 = out_b16; a.residual = residual; a.res_bstride = 0;
  return convtc_launch(a, st);
}

static int bert_fwd_impl(t2_bert* g, const long long* ids, const long long* type_ids, const float* mask, int B, int S, float* pooled, float* hidden_out,
                         Arena& ar, cudaStream_t st) {
  const t2_bert_config& c = g->cfg;
  const int H = c.hidden_size, I = c.intermediate_size, N = B * S;
  float* X = ar.take<float>((size_t)N * H);
  float* Y = ar.take<float>((size_t)N * H);
  float* QKV = ar.take<float>((size_t)N * 3 * H);
  __nv_bfloat16* Xb = ar.take<__nv_bfloat16>((size_t)2 * N * H);
  __nv_bfloat16* Cb = ar.take<__nv_bfloat16>((size_t)2 * N * H);
  __nv_bfloat16* Hb = ar.take<__nv_bfloat16>((size_t)2 * N * I);
  __nv_bfloat16* cls = ar.take<__nv_bfloat16>((size_t)2 * B * H);
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_bert_pooler_fwd: workspace too small");
  const float *word, *pos, *type, *eg, *eb;
  {
    const BParam* p;
    T2_REQUIRE((p = bfind(g, "embeddings.word_embeddings.weight", (long long)c.vocab_size * H)), T2_ERR_STATE, "embeddings.word_embeddings.weight missing"); word = p->ptr;
    T2_REQUIRE((p = bfind(g, "embeddings.position_embeddings.weight", (long long)c.max_position * H)), T2_ERR_STATE, "embeddings.position_embeddings.weight missing"); pos = p->ptr;
    T2_REQUIRE((p = bfind(g, "embeddings.token_type_embeddings.weight", (long long)c.type_vocab_size * H)), T2_ERR_STATE, "embeddings.token_type_embeddings.weight missing"); type = p->ptr;
    T2_REQUIRE((p = bfind(g, "embeddings.LayerNorm.weight", H)), T2_ERR_STATE, "embeddings.LayerNorm.weight missing"); eg = p->ptr;
    T2_REQUIRE((p = bfind(g, "embeddings.LayerNorm.bias", H)), T2_ERR_STATE, "embeddings.LayerNorm.bias missing"); eb = p->ptr;
  }
  const size_t ln_smem = (size_t)H * sizeof(float);
  LnArgs e{};
  e.ids = ids; e.type_ids = type_ids; e.word = word; e.pos = pos; e.type = type; e.S = S; e.vocab = c.vocab_size; e.max_pos = c.max_position;
  e.type_vocab = c.type_vocab_size; e.gamma = eg; e.beta = eb; e.eps = c.layer_norm_eps; e.N = N; e.H = H; e.out_f32 = X; e.out_b16 = Xb;
  bert_ln_kernel<<<N, 256, ln_smem, st>>>(e);
  const int dh = H / c.num_heads;
  const size_t att_smem = ((size_t)2 * S * (dh + 1) + 8 * (size_t)S) * sizeof(float);
  T2_REQUIRE(att_smem <= 220 * 1024, T2_ERR_UNSUPPORTED, "t2_bert_pooler_fwd: sequence length %d too long for the attention kernel", S);
  T2_CHECK_CUDA(cudaFuncSetAttribute(bert_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_smem));
  for (int i = 0; i < c.num_layers; ++i) {
    const std::string p = "encoder.layer." + std::to_string(i) + ".";
    const BParam *g1 = bfind(g, p + "attention.output.LayerNorm.weight", H), *b1 = bfind(g, p + "attention.output.LayerNorm.bias", H);
    const BParam *g2 = bfind(g, p + "output.LayerNorm.weight", H), *b2 = bfind(g, p + "output.LayerNorm.bias", H);
    T2_REQUIRE(g1 && b1 && g2 && b2, T2_ERR_STATE, "t2_bert_pooler_fwd: LayerNorm parameters of layer %d missing", i);
    RC(blin(g, g->qkv[i], Xb, N, ACT_NONE, QKV, nullptr, nullptr, st));                          // BertSelfAttention: query | key | value
    bert_attention_kernel<<<B * c.num_heads, 256, att_smem, st>>>(QKV, mask, S, H, c.num_heads, Cb, (long long)N * H);
    RC(blin(g, g->ao[i], Cb, N, ACT_NONE, Y, nullptr, X, st));                                   // BertSelfOutput: dense + residual ...
    LnArgs l1{}; l1.a = Y; l1.gamma = g1->ptr; l1.beta = b1->ptr; l1.eps = c.layer_norm_eps; l1.N = N; l1.H = H; l1.out_f32 = X; l1.out_b16 = Xb;
    bert_ln_kernel<<<N, 256, ln_smem, st>>>(l1);                                                 // ... + LayerNorm
    RC(blin(g, g->ff1[i], Xb, N, ACT_GELU, nullptr, Hb, nullptr, st));                           // BertIntermediate: dense + GELU
    RC(blin(g, g->ff2[i], Hb, N, ACT_NONE, Y, nullptr, X, st));                                  // BertOutput: dense + residual + LayerNorm
    LnArgs l2{}; l2.a = Y; l2.gamma = g2->ptr; l2.beta = b2->ptr; l2.eps = c.layer_norm_eps; l2.N = N; l2.H = H; l2.out_f32 = X; l2.out_b16 = Xb;
    bert_ln_kernel<<<N, 256, ln_smem, st>>>(l2);
  }
  T2_CHECK_CUDA(cudaGetLastError());
  if (hidden_out) T2_CHECK_CUDA(cudaMemcpyAsync(hidden_out, X, (size_t)N * H * sizeof(float), cudaMemcpyDeviceToDevice, st));
  bert_cls_kernel<<<std::max(1, std::min((B * H + 255) / 256, 148 * 4)), 256, 0, st>>>(X, cls, B, S, H);
  RC(blin(g, g->pool, cls, B, ACT_TANH, pooled, nullptr, nullptr, st));                          // BertPooler: tanh(dense(hidden[:, 0]))
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

extern "C" long long t2_bert_workspace_bytes(const t2_bert* gc, int B, int S) {
  t2_bert* g = const_cast<t2_bert*>(gc);
  if (!g || B <= 0 || S <= 0) return -1;
  Arena a(nullptr, 0);
  bert_fwd_impl(g, nullptr, nullptr, nullptr, B, S, nullptr, nullptr, a, nullptr);
  return a.off + 1024;
}

extern "C" int t2_bert_pooler_fwd(t2_bert* g, const long long* input_ids, const long long* token_type_ids, const float* attention_mask, int B, int S,
                                  float* pooler_output, float* last_hidden_state, void* ws, long long ws_bytes, void* stream) {
  T2_REQUIRE(g && input_ids && pooler_output && ws && B > 0 && S > 0, T2_ERR_ARG, "t2_bert_pooler_fwd: bad argument");
  T2_REQUIRE(g->committed, T2_ERR_STATE, "t2_bert_pooler_fwd: weights not committed (call t2_bert_commit)");
  T2_REQUIRE(S <= g->cfg.max_position, T2_ERR_ARG, "t2_bert_pooler_fwd: %d tokens exceed max_position_embeddings %d", S, g->cfg.max_position);
  Arena ar(ws, ws_bytes);
  return bert_fwd_impl(g, input_ids, token_type_ids, attention_mask, B, S, pooler_output, last_hidden_state, ar, (cudaStream_t)stream);
}
