// C ABI of libt2b200 (see include/t2b200.h): context, weight packing, op orchestration.
#include "../../include/t2b200.h"
#include "common.cuh"
#include "decoder.cuh"
#include "convtc.cuh"
#include "layers.cuh"
#include "ctx.cuh"

#include <cstdarg>
#include <vector>

static thread_local char g_err[1024] = "";
void t2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* t2_last_error(void) { return g_err; }
extern "C" int t2_abi_version(void) { return 1; }

namespace {

__global__ void poison_on_abort_kernel(const unsigned* status, float* out, long long n) {
  if (*status == 0u) return;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) out[i] = __int_as_float(0x7fc00000);
}

__global__ void add_vec_kernel(const float* a, const float* b, float* o, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = a[i] + b[i];
}

__global__ void pack_proj_kernel(const float* Wmel, const float* Wgate, const float* bmel, const float* bgate, float* Wproj,
                                 float* bproj, int M, int Kmel, int Kgate, int KP) {
  const int n = (M + 1) * KP;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int r = i / KP, k = i % KP;
    float v = 0.f;
    if (r < M) { if (k < Kmel) v = Wmel[(size_t)r * Kmel + k]; }
    else if (k < Kgate) v = Wgate[k];
    Wproj[i] = v;
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= M; i += gridDim.x * blockDim.x) bproj[i] = i < M ? bmel[i] : bgate[0];
}

// row-major (B,dim) -> k-major [dim][BP] rows at row offset, batch slots >= B zeroed
__global__ void rm_to_km_kernel(const float* src, int ld, float* dst, int row0, int dim, int B) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < dim * D2_BP; i += gridDim.x * blockDim.x) {
    const int r = i / D2_BP, b = i % D2_BP;
    dst[(size_t)(row0 + r) * D2_BP + b] = (b < B && src) ? src[(size_t)b * ld + r] : 0.f;
  }
}
__global__ void km_to_rm_kernel(const float* src, int row0, float* dst, int ld, int dim, int B) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < dim * B; i += gridDim.x * blockDim.x) {
    const int b = i / dim, r = i % dim;
    dst[(size_t)b * ld + r] = src[(size_t)(row0 + r) * D2_BP + b];
  }
}
__global__ void copy2d_kernel(const float* src, int lds, float* dst, int ldd, int rows, int cols) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += gridDim.x * blockDim.x) {
    const int r = i / cols, c = i % cols;
    dst[(size_t)r * ldd + c] = src[(size_t)r * lds + c];
  }
}

// prenet (model/tacotron2.py:85-92) on N independent rows, 8 rows per CTA pass.
// mode 0: x (N,M) -> out (N,P), keep1/keep2 (N,P).
// mode 2 (teacher forcing, model/tacotron2.py:255-257): row n = t*B + b, t in [0,T]; input is
//   mel_target[b][t-1] (zeros for t = 0, the F.pad row); keep = prenet_keep (T+1,2,B,P);
//   output is bf16 (hi,lo) [T+1][2][D2_BP][P], the layout the decoder's TMA descriptor reads.
struct PrenetArgs {
  const float* x; const float* W1; const float* W2;
  const unsigned char* k1; const unsigned char* k2;
  unsigned long long seed; float p; int N, M, P, mode, B, T; float* out;
  __nv_bfloat16* out_b;  // mode 2: bf16 (hi,lo) [T+1][2][D2_BP][P] for the tensor-core decoder
  float* p1_out;         // optional [N][P]: layer-1 output after dropout (training tape)
};
// 32 rows per CTA, both layers as shared-memory tiled products: thread (ty = warp, tx = lane) accumulates rows 4 ty .. 4 ty + 3 for the
// outputs tx + 32 j; the inputs sit transposed in shared memory (one broadcast LDS.128 per k for the four rows), the weights pass
// through a [32 k][P] tile (filled with coalesced 128-byte row pieces, read conflict-free).  Products are added in ascending k as before.
#define PN_ROWS 32
#define PN_KC 32
__host__ __device__ inline size_t prenet_smem_bytes(int M, int P) { return ((size_t)M * PN_ROWS + (size_t)P * PN_ROWS + (size_t)PN_KC * (P + 1)) * sizeof(float); }
__global__ void __launch_bounds__(256) prenet_rows_kernel(PrenetArgs a) {
  extern __shared__ float sh[];
  const int M = a.M, P = a.P;
  float* xT = sh;                       // [M][PN_ROWS]
  float* hT = xT + (size_t)M * PN_ROWS; // [P][PN_ROWS]   layer-1 output, transposed
  float* Ws = hT + (size_t)P * PN_ROWS; // [PN_KC][P + 1]
  const float scale = 1.0f / (1.0f - a.p);
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int NJ = P / 32;                // outputs per thread (P % 32 == 0, P <= 256: checked by the callers' shapes)
  for (int n0 = blockIdx.x * PN_ROWS; n0 < a.N; n0 += gridDim.x * PN_ROWS) {
    __syncthreads();
    for (int i = threadIdx.x; i < PN_ROWS * M; i += blockDim.x) {
      const int r = i / M, cidx = i - r * M, n = n0 + r;
      float v = 0.f;
      if (n < a.N) {
        if (a.mode == 0) v = a.x[(size_t)n * M + cidx];
        else { const int t = n / a.B, bb = n % a.B; if (t > 0) v = a.x[((size_t)bb * a.T + (t - 1)) * M + cidx]; }
      }
      xT[cidx * PN_ROWS + r] = v;
    }
    for (int layer = 0; layer < 2; ++layer) {
      const float* W = layer ? a.W2 : a.W1;
      const int K = layer ? P : M;
      const float* inT = layer ? hT : xT;
      float acc[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) { acc[j][0] = 0.f; acc[j][1] = 0.f; acc[j][2] = 0.f; acc[j][3] = 0.f; }
      for (int k0 = 0; k0 < K; k0 += PN_KC) {
        __syncthreads();   // the previous tile (and, first time, the inputs) are done with / in place
        for (int o = ty; o < P; o += 8) Ws[tx * (P + 1) + o] = (k0 + tx < K) ? __ldg(W + (size_t)o * K + k0 + tx) : 0.f;
        __syncthreads();
        const int kn = min(PN_KC, K - k0);
#pragma unroll 4
        for (int kk = 0; kk < kn; ++kk) {
          const float4 x4 = *reinterpret_cast<const float4*>(inT + (size_t)(k0 + kk) * PN_ROWS + 4 * ty);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j < NJ) {
              const float w = Ws[kk * (P + 1) + tx + 32 * j];
              acc[j][0] = fmaf(w, x4.x, acc[j][0]); acc[j][1] = fmaf(w, x4.y, acc[j][1]);
              acc[j][2] = fmaf(w, x4.z, acc[j][2]); acc[j][3] = fmaf(w, x4.w, acc[j][3]);
            }
          }
        }
      }
      const unsigned char* km = layer ? a.k2 : a.k1;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (j >= NJ) continue;
        const int o = tx + 32 * j;
        float v4[4];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
          const int n = n0 + 4 * ty + rr;
          float v = 0.f;
          if (n < a.N) {
            size_t kidx; unsigned long long ridx;
            if (a.mode == 0) { kidx = (size_t)n * P + o; ridx = ((unsigned long long)n * 2 + layer) * P + o; }
            else {
              const int t = n / a.B, bb = n % a.B;
              kidx = (((size_t)t * 2 + layer) * a.B + bb) * P + o;
              ridx = kidx;
            }
            const bool keep = km ? (km[kidx] != 0) : rng_keep(a.seed, ridx, a.p);
            v = (acc[j][rr] > 0.f && keep) ? acc[j][rr] * scale : 0.f;
            if (layer == 0) { if (a.p1_out) a.p1_out[(size_t)n * P + o] = v; }
            else if (a.mode == 0) a.out[(size_t)n * P + o] = v;
            else {
              const int t = n / a.B, bb = n % a.B;
              __nv_bfloat16 hi, lo;
              tc::split_bf16(v, hi, lo);
              a.out_b[(((size_t)t * 2 + 0) * D2_BP + bb) * P + o] = hi;
              a.out_b[(((size_t)t * 2 + 1) * D2_BP + bb) * P + o] = lo;
            }
          }
          v4[rr] = v;
        }
        if (layer == 0) *reinterpret_cast<float4*>(hT + (size_t)o * PN_ROWS + 4 * ty) = make_float4(v4[0], v4[1], v4[2], v4[3]);
      }
    }
  }
}
static int prenet_rows_launch(const PrenetArgs& pa, cudaStream_t st) {
  T2_REQUIRE(pa.P % 32 == 0 && pa.P <= 256, T2_ERR_UNSUPPORTED, "prenet: prenet_dim %d (multiple of 32, at most 256)", pa.P);
  const size_t smem = prenet_smem_bytes(pa.M, pa.P);
  T2_CHECK_CUDA(cudaFuncSetAttribute(prenet_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  prenet_rows_kernel<<<std::min((pa.N + PN_ROWS - 1) / PN_ROWS, 148 * 16), 256, smem, st>>>(pa);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

// bf16 (hi,lo) planes of [W_ih | W_hh] with rows permuted so that each 128-row tile holds 32 hidden units and
// rows [32j, 32j+32) of a tile are gates (i,f,g,o) x 8 units: packed row p = tile*128 + j*32 + g*8 + ul.
__global__ void pack_lstm_bf16_kernel(const float* Wih, const float* Whh, __nv_bfloat16* out, int H, int Kin, int KP) {
  const long long n = 4ll * H * KP;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    // destination is tile-major: [tile][kb][128 rows][64 k] so every TMA box is one contiguous 16 KB chunk
    const int kk = (int)(i % 64);
    const int rr = (int)((i / 64) % 128);
    const int nkb = KP / 64;
    const int kb = (int)((i / (64 * 128)) % nkb);
    const int tile = (int)(i / (64ll * 128 * nkb));
    const int k = kb * 64 + kk;
    const int j = (rr >> 5) & 3, g = (rr >> 3) & 3, ul = rr & 7;
    const int row = g * H + tile * 32 + j * 8 + ul;
    float v = 0.f;
    if (k < Kin) v = Wih[(size_t)row * Kin + k];
    else if (k < Kin + H) v = Whh[(size_t)row * H + (k - Kin)];
    __nv_bfloat16 hi, lo;
    tc::split_bf16(v, hi, lo);
    out[i] = hi;
    out[n + i] = lo;
  }
}
// transposed pack for the in-loop input-gradient GEMM dx[k][b] = sum_p W[row(p)][k] dg[b][p]:
// tile-major [ktile][pblock][128 k][64 p], p = packed gate-row index (same permutation as pack_lstm_bf16_kernel)
__global__ void pack_lstm_T_bf16_kernel(const float* Wih, const float* Whh, __nv_bfloat16* out, int H, int Kin, int KT) {
  const long long n = 4ll * H * KT;
  const int npb = 4 * H / 64;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int pp = (int)(i % 64);
    const int kk = (int)((i / 64) % 128);
    const int pb = (int)((i / (64 * 128)) % npb);
    const int kt = (int)(i / (64ll * 128 * npb));
    const int k = kt * 128 + kk, p = pb * 64 + pp;
    const int tile = p >> 7, rr = p & 127;
    const int j = (rr >> 5) & 3, g = (rr >> 3) & 3, ul = rr & 7;
    const int row = g * H + tile * 32 + j * 8 + ul;
    float v = 0.f;
    if (k < Kin) v = Wih[(size_t)row * Kin + k];
    else if (k < Kin + H) v = Whh[(size_t)row * H + (k - Kin)];
    __nv_bfloat16 hi, lo;
    tc::split_bf16(v, hi, lo);
    out[i] = hi;
    out[n + i] = lo;
  }
}
// dst[2][rows_out][ld]: dst[k][r] = r < rows_in ? W[r][k] : 0   (fp32 [rows_in][ldw] -> transposed bf16 hi/lo)
__global__ void pack_T_kernel(const float* W, int rows_in, int ldw, __nv_bfloat16* dst, int rows_out, int ld) {
  const long long n = (long long)rows_out * ld;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i % ld), k = (int)(i / ld);
    __nv_bfloat16 hi, lo;
    tc::split_bf16(r < rows_in ? W[(size_t)r * ldw + k] : 0.f, hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}
// dst[2][rows][K]: dst[m][j] = W[m][src_off + j - lead] for 0 <= j - lead < src_n, else 0  (column window of an fp32 matrix -> bf16 hi/lo)
__global__ void pack_cols_kernel(const float* W, int rows, int ldw, int src_off, int src_n, int lead, __nv_bfloat16* dst, int K) {
  const long long n = (long long)rows * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % K) - lead, m = (int)(i / K);
    __nv_bfloat16 hi, lo;
    tc::split_bf16((j >= 0 && j < src_n) ? W[(size_t)m * ldw + src_off + j] : 0.f, hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}
// conv weight [Cout][Cin][k] -> bf16 (hi,lo) [2][Cin][k][Cout] with the taps flipped (input-gradient conv)
__global__ void pack_conv_wT_kernel(const float* W, __nv_bfloat16* dst, int Cout, int Cin, int k) {
  const long long n = (long long)Cout * Cin * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    const int jj = (int)((i / Cout) % k);
    const int ci = (int)(i / ((long long)Cout * k));
    __nv_bfloat16 hi, lo;
    tc::split_bf16(W[((size_t)co * Cin + ci) * k + (k - 1 - jj)], hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}
__global__ void pack_loc_filter_kernel(const float* conv /*[32][2][31]*/, const float* dense /*[A][32]*/, int A, __nv_bfloat16* out /*[2][128][64]*/) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 128 * 64) return;
  const int a = i / 64, k = i % 64, c = k / 32, j = k % 32;
  float g = 0.f;
  if (a < A && j < 31)
    for (int f = 0; f < 32; ++f) g = fmaf(dense[a * 32 + f], conv[(f * 2 + c) * 31 + j], g);
  __nv_bfloat16 hi, lo;
  tc::split_bf16(g, hi, lo);
  out[i] = hi;
  out[128 * 64 + i] = lo;
}
__global__ void conv_t_kernel(const float* w /*[32][2][31]*/, float* out /*[2][31][32]*/) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 32 * 2 * 31) return;
  const int f = i % 32, j = (i / 32) % 31, c = i / (32 * 31);
  out[i] = w[(f * 2 + c) * 31 + j];
}
// controls (B,C) -> bf16 (hi,lo) columns [col0, col0+C) of every slot of x[nslots][2][BP][ld]
__global__ void controls_slots_kernel(const float* src, int C, __nv_bfloat16* dst, long long plane, int ld, int col0, int B, int nslots) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)nslots * B * C; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C), b = (int)((i / C) % B), sl = (int)(i / ((long long)C * B));
    __nv_bfloat16 hi, lo;
    tc::split_bf16(src[(size_t)b * C + c], hi, lo);
    __nv_bfloat16* x = dst + (size_t)sl * 2 * plane;
    x[(size_t)b * ld + col0 + c] = hi;
    x[plane + (size_t)b * ld + col0 + c] = lo;
  }
}
// fp32 row-major (B,dim) -> bf16 (hi,lo) columns [col0, col0+dim) of x[2][BP][ld]
__global__ void rm_to_xsplit_kernel(const float* src, int lds, __nv_bfloat16* dst, long long plane, int ld, int col0, int dim, int B) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < dim * B; i += gridDim.x * blockDim.x) {
    const int b = i / dim, c = i % dim;
    __nv_bfloat16 hi, lo;
    tc::split_bf16(src[(size_t)b * lds + c], hi, lo);
    dst[(size_t)b * ld + col0 + c] = hi;
    dst[plane + (size_t)b * ld + col0 + c] = lo;
  }
}

inline int gsz(long long n) { return (int)std::min<long long>((n + 255) / 256, 148 * 8); }

}  // namespace

int poison_on_abort_launch(const unsigned* status, float* out, long long n, cudaStream_t st) {
  if (!status || !out || n <= 0) return T2_OK;
  poison_on_abort_kernel<<<1, 256, 0, st>>>(status, out, std::min<long long>(n, 4096));
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int make_w_map(CUtensorMap* m, const __nv_bfloat16* w, int Cout, int Cin, int k, int NT, int npass) {
  unsigned long long dims[4] = {(unsigned long long)Cin, (unsigned long long)k, (unsigned long long)Cout, 2ull};
  unsigned long long st[3] = {(unsigned long long)Cin * 2, (unsigned long long)k * Cin * 2, (unsigned long long)Cout * k * Cin * 2};
  unsigned box[4] = {64u, 1u, (unsigned)NT, npass == 3 ? 2u : 1u};
  return tc_make_tmap_bf16(m, w, 4, dims, st, box);
}
int make_in_map(CUtensorMap* m, const __nv_bfloat16* x, int B, int T, int Cin, int npass) {
  unsigned long long dims[4] = {(unsigned long long)Cin, (unsigned long long)T, (unsigned long long)B, 2ull};
  unsigned long long st[3] = {(unsigned long long)Cin * 2, (unsigned long long)T * Cin * 2, (unsigned long long)B * T * Cin * 2};
  unsigned box[4] = {64u, 128u, 1u, npass == 3 ? 2u : 1u};
  return tc_make_tmap_bf16(m, x, 4, dims, st, box);
}

// ------------------------------------------------------------------------------------------ context
extern "C" int t2_ctx_create(const t2_model_config* cfg, t2_ctx** out) {
  T2_REQUIRE(cfg && out, T2_ERR_ARG, "t2_ctx_create: null argument");
  int dev = 0, count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  T2_REQUIRE(e == cudaSuccess && count > 0, T2_ERR_CUDA, "t2_ctx_create: no CUDA device (%s); libt2b200 has no CPU fallback",
             cudaGetErrorString(e));
  T2_CHECK_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  T2_CHECK_CUDA(cudaGetDeviceProperties(&prop, dev));
  // the binary carries sm_100a SASS only (tcgen05 / TMEM): any other architecture, newer ones included, cannot run it
  T2_REQUIRE(prop.major == 10, T2_ERR_UNSUPPORTED, "t2_ctx_create: device sm_%d%d; libt2b200 is built for sm_100a only", prop.major,
             prop.minor);
  t2_ctx* c = new t2_ctx();
  c->cfg = *cfg;
  c->device = dev;
  c->num_sms = prop.multiProcessorCount;
  c->E = cfg->encoded_dim;
  c->EF = cfg->encoded_dim + (cfg->description_dim > 0 ? 128 : 0);
  c->C = cfg->controls_dim;
  c->H = cfg->att_rnn_dim;
  c->R = cfg->rnn_hidden_dim;
  c->P = cfg->prenet_dim;
  c->A = cfg->att_dim;
  c->M = cfg->num_mels;
  c->PD = cfg->postnet_dim;
  c->KS = cfg->encoder_kernel_size;
  c->dd.KattP = (int)round_up64(c->P + c->EF + c->H, 64);
  c->dd.KdecP = (int)round_up64(c->H + c->EF + c->C + c->R, 64);
  c->dd.KprojP = (int)round_up64(c->R + c->EF + c->C, 4);
  c->npass = cfg->precision == 1 ? 1 : 3;
  c->committed = false;
  c->slab = nullptr;
  c->slab_bytes = 0;
  c->prof_buf = nullptr;
  T2_CHECK_CUDA(cudaMalloc((void**)&c->prof_buf, 16 * sizeof(unsigned long long)));
  c->status_dev = nullptr;
  T2_CHECK_CUDA(cudaMalloc((void**)&c->status_dev, 4 * sizeof(unsigned)));
  T2_CHECK_CUDA(cudaMemset(c->status_dev, 0, 4 * sizeof(unsigned)));
  c->timing_on = 0;
  c->tev_valid[0] = c->tev_valid[1] = false;
  for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) T2_CHECK_CUDA(cudaEventCreate(&c->tev[i][j]));
  *out = c;
  return T2_OK;
}

extern "C" int t2_ctx_destroy(t2_ctx* c) {
  if (!c) return T2_OK;
  if (c->slab) cudaFree(c->slab);
  if (c->prof_buf) cudaFree(c->prof_buf);
  if (c->status_dev) cudaFree(c->status_dev);
  for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) cudaEventDestroy(c->tev[i][j]);
  delete c;
  return T2_OK;
}

extern "C" int t2_ctx_set_param(t2_ctx* c, const char* key, const float* ptr, long long numel) {
  T2_REQUIRE(c && key && ptr, T2_ERR_ARG, "t2_ctx_set_param: null argument");
  c->params[key] = Param{ptr, numel};
  c->committed = false;
  return T2_OK;
}

#define NEED(var, key, n)                                                                                   \
  const Param* var = find(c, key);                                                                          \
  T2_REQUIRE(var && var->numel == (long long)(n), T2_ERR_STATE, "t2_ctx_commit: parameter %s missing or wrong size (want %lld)", \
             std::string(key).c_str(), (long long)(n))

static int commit_impl(t2_ctx* c, Arena& ar, cudaStream_t st) {
  const bool sizing = ar.base == nullptr;
  const auto& d = c->dd;
  const int H = c->H, R = c->R, P = c->P, A = c->A, M = c->M, E = c->E, EF = c->EF, C = c->C, PD = c->PD, KS = c->KS;
  const int Hh = E / 2;
  float* batt = ar.take<float>(4 * H);
  float* bdec = ar.take<float>(4 * R);
  float* Wproj = ar.take<float>((size_t)(M + 1) * d.KprojP);
  float* bproj = ar.take<float>(M + 1);
  int chans[6] = {M, PD, PD, PD, PD, M};
  for (int i = 0; i < 3; ++i) {
    c->enc_scale[i] = ar.take<float>(E);
    c->enc_shift[i] = ar.take<float>(E);
  }
  for (int i = 0; i < 5; ++i) {
    c->post_scale[i] = ar.take<float>(chans[i + 1]);
    c->post_shift[i] = ar.take<float>(chans[i + 1]);
  }
  c->lstm_bias = ar.take<float>(8 * Hh);
  c->lstm_Whh = ar.take<float>((size_t)2 * Hh * 4 * Hh);
  float* tmp = ar.take<float>((size_t)8 * Hh * E);  // staging for the concatenated BiLSTM W_ih
  c->Watt_b = ar.take<__nv_bfloat16>((size_t)2 * 4 * H * d.KattP);
  c->Wdec_b = ar.take<__nv_bfloat16>((size_t)2 * 4 * R * d.KdecP);
  c->loc_conv_t = ar.take<float>(2 * 31 * 32);
  c->loc_dense_t = ar.take<float>((size_t)32 * A);
  c->locG_b = ar.take<__nv_bfloat16>((size_t)2 * 128 * 64);
  for (int i = 0; i < 3; ++i) c->enc_Wb[i] = ar.take<__nv_bfloat16>((size_t)2 * E * KS * E);
  for (int i = 0; i < 5; ++i) c->post_Wb[i] = ar.take<__nv_bfloat16>((size_t)2 * chans[i + 1] * 5 * chans[i]);
  c->lstm_Wb = ar.take<__nv_bfloat16>((size_t)2 * 8 * Hh * E);
  c->attenc_Wb = ar.take<__nv_bfloat16>((size_t)2 * A * EF);
  for (int i = 0; i < 3; ++i) c->enc_WTb[i] = ar.take<__nv_bfloat16>((size_t)2 * E * KS * E);
  for (int i = 0; i < 5; ++i) c->post_WTb[i] = ar.take<__nv_bfloat16>((size_t)2 * chans[i + 1] * 5 * chans[i]);
  c->lstm_WTb = ar.take<__nv_bfloat16>((size_t)2 * 8 * Hh * E);
  c->attenc_WTb = ar.take<__nv_bfloat16>((size_t)2 * A * EF);
  c->KattT = (int)round_up64(P + EF + H, 128);
  c->KdecT = (int)round_up64(H + EF + C + R, 128);
  c->WTatt_b = ar.take<__nv_bfloat16>((size_t)2 * 4 * H * c->KattT);
  c->WTdec_b = ar.take<__nv_bfloat16>((size_t)2 * 4 * R * c->KdecT);
  c->WprojT_b = ar.take<__nv_bfloat16>((size_t)2 * d.KprojP * PROJ_LD);
  c->projA_col0 = (H + EF + C) & ~7;
  c->KprojA = d.KdecP - c->projA_col0;
  c->KprojB = (EF + C + 7) & ~7;
  c->WprojA_b = ar.take<__nv_bfloat16>((size_t)2 * (M + 1) * c->KprojA);
  c->WprojB_b = ar.take<__nv_bfloat16>((size_t)2 * (M + 1) * c->KprojB);
  c->W2T_b = ar.take<__nv_bfloat16>((size_t)2 * P * P);
  c->WpreT_b = ar.take<__nv_bfloat16>((size_t)2 * P * 4 * H);
  if (sizing) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_STATE, "t2_ctx_commit: slab overflow");

  NEED(att_ih, "decoder.att_rnn.weight_ih", (size_t)4 * H * (P + EF));
  NEED(att_hh, "decoder.att_rnn.weight_hh", (size_t)4 * H * H);
  NEED(att_bi, "decoder.att_rnn.bias_ih", 4 * H);
  NEED(att_bh, "decoder.att_rnn.bias_hh", 4 * H);
  NEED(dec_ih, "decoder.lstm.weight_ih", (size_t)4 * R * (H + EF + C));
  NEED(dec_hh, "decoder.lstm.weight_hh", (size_t)4 * R * R);
  NEED(dec_bi, "decoder.lstm.bias_ih", 4 * R);
  NEED(dec_bh, "decoder.lstm.bias_hh", 4 * R);
  NEED(wq, "decoder.attention.query_layer.weight", (size_t)A * H);
  NEED(wv, "decoder.attention.v.weight", A);
  NEED(lconv, "decoder.attention.location_conv.weight", 32 * 2 * 31);
  NEED(ldense, "decoder.attention.location_dense.weight", (size_t)A * 32);
  NEED(wmel, "decoder.mel_out.weight", (size_t)M * (R + EF + C));
  NEED(bmel, "decoder.mel_out.bias", M);
  NEED(wgate, "decoder.gate.weight", R + EF);
  NEED(bgate, "decoder.gate.bias", 1);
  NEED(w1, "prenet.0.weight", (size_t)P * M);
  NEED(w2, "prenet.3.weight", (size_t)P * P);
  NEED(wae, "att_encoder.weight", (size_t)A * EF);
  T2_REQUIRE(H == R, T2_ERR_UNSUPPORTED, "att_rnn_dim != rnn_hidden_dim is not supported by the packed layout");

  add_vec_kernel<<<(4 * H + 255) / 256, 256, 0, st>>>(att_bi->ptr, att_bh->ptr, batt, 4 * H);
  add_vec_kernel<<<(4 * R + 255) / 256, 256, 0, st>>>(dec_bi->ptr, dec_bh->ptr, bdec, 4 * R);
  pack_proj_kernel<<<gsz((long long)(M + 1) * d.KprojP), 256, 0, st>>>(wmel->ptr, wgate->ptr, bmel->ptr, bgate->ptr, Wproj, bproj, M,
                                                                      R + EF + C, R + EF, d.KprojP);
  T2_CHECK_CUDA(cudaGetLastError());
  // tensor-core decoder operands
  pack_lstm_bf16_kernel<<<gsz(4ll * H * d.KattP), 256, 0, st>>>(att_ih->ptr, att_hh->ptr, c->Watt_b, H, P + EF, d.KattP);
  pack_lstm_bf16_kernel<<<gsz(4ll * R * d.KdecP), 256, 0, st>>>(dec_ih->ptr, dec_hh->ptr, c->Wdec_b, R, H + EF + C, d.KdecP);
  conv_t_kernel<<<(32 * 2 * 31 + 255) / 256, 256, 0, st>>>(lconv->ptr, c->loc_conv_t);
  if (A <= 128) {
    pack_loc_filter_kernel<<<(128 * 64 + 255) / 256, 256, 0, st>>>(lconv->ptr, ldense->ptr, A, c->locG_b);
    unsigned long long gd[3] = {64ull, 128ull, 2ull};
    unsigned long long gs[2] = {128ull, 128ull * 128ull};
    unsigned gb[3] = {64u, 128u, 2u};
    int rcg;
    if ((rcg = tc_make_tmap_bf16(&c->map_locG, c->locG_b, 3, gd, gs, gb))) return rcg;
  }
  T2_CHECK_CUDA(cudaGetLastError());
  { int rc3; if ((rc3 = transpose_launch(ldense->ptr, c->loc_dense_t, A, 32, st))) return rc3; }
  c->d2w = D2Weights{batt, bdec, wq->ptr, wv->ptr, c->loc_conv_t, c->loc_dense_t, Wproj, bproj, w1->ptr, w2->ptr};
  // training operands
  T2_REQUIRE(M + 1 <= PROJ_LD, T2_ERR_UNSUPPORTED, "num_mels %d too large for the projection-gradient pack", M);
  pack_lstm_T_bf16_kernel<<<gsz(4ll * H * c->KattT), 256, 0, st>>>(att_ih->ptr, att_hh->ptr, c->WTatt_b, H, P + EF, c->KattT);
  pack_lstm_T_bf16_kernel<<<gsz(4ll * R * c->KdecT), 256, 0, st>>>(dec_ih->ptr, dec_hh->ptr, c->WTdec_b, R, H + EF + C, c->KdecT);
  pack_T_kernel<<<gsz((long long)d.KprojP * PROJ_LD), 256, 0, st>>>(Wproj, M + 1, d.KprojP, c->WprojT_b, d.KprojP, PROJ_LD);
  pack_cols_kernel<<<gsz((long long)(M + 1) * c->KprojA), 256, 0, st>>>(Wproj, M + 1, d.KprojP, 0, R, H + EF + C - c->projA_col0, c->WprojA_b,
                                                                       c->KprojA);
  pack_cols_kernel<<<gsz((long long)(M + 1) * c->KprojB), 256, 0, st>>>(Wproj, M + 1, d.KprojP, R, EF + C, 0, c->WprojB_b, c->KprojB);
  pack_T_kernel<<<gsz((long long)P * P), 256, 0, st>>>(w2->ptr, P, P, c->W2T_b, P, P);
  pack_T_kernel<<<gsz((long long)P * 4 * H), 256, 0, st>>>(att_ih->ptr, 4 * H, P + EF, c->WpreT_b, P, 4 * H);
  T2_CHECK_CUDA(cudaGetLastError());
  {
    const unsigned zb = c->npass == 3 ? 2u : 1u;
    unsigned box[3] = {TG_BK, TG_M, zb};
    int rc2;
    unsigned long long dims[3] = {64ull, (unsigned long long)c->KattT * (4 * H / 64), 2ull};
    unsigned long long strides[2] = {128ull, 4ull * H * c->KattT * 2};
    if ((rc2 = tc_make_tmap_bf16(&c->map_wt_att, c->WTatt_b, 3, dims, strides, box))) return rc2;
    unsigned long long dims2[3] = {64ull, (unsigned long long)c->KdecT * (4 * R / 64), 2ull};
    unsigned long long strides2[2] = {128ull, 4ull * R * c->KdecT * 2};
    if ((rc2 = tc_make_tmap_bf16(&c->map_wt_dec, c->WTdec_b, 3, dims2, strides2, box))) return rc2;
  }
  {
    const unsigned zb = c->npass == 3 ? 2u : 1u;
    unsigned long long dims[3] = {64ull, 4ull * H * (d.KattP / 64), 2ull};
    unsigned long long strides[2] = {128ull, 4ull * H * d.KattP * 2};
    unsigned box[3] = {TG_BK, TG_M, zb};
    int rc2;
    if ((rc2 = tc_make_tmap_bf16(&c->map_w_att, c->Watt_b, 3, dims, strides, box))) return rc2;
    unsigned long long dims2[3] = {64ull, 4ull * R * (d.KdecP / 64), 2ull};
    unsigned long long strides2[2] = {128ull, 4ull * R * d.KdecP * 2};
    if ((rc2 = tc_make_tmap_bf16(&c->map_w_dec, c->Wdec_b, 3, dims2, strides2, box))) return rc2;
  }

  int rc;
  for (int i = 0; i < 3; ++i) {
    const std::string cv = "encoder.convolutions." + std::to_string(4 * i), bn = "encoder.convolutions." + std::to_string(4 * i + 1);
    NEED(cw, cv + ".weight", (size_t)E * E * KS);
    NEED(cb, cv + ".bias", E);
    NEED(bw, bn + ".weight", E);
    NEED(bb, bn + ".bias", E);
    NEED(bm, bn + ".running_mean", E);
    NEED(bv, bn + ".running_var", E);
    if ((rc = pack_conv_w_launch(cw->ptr, c->enc_Wb[i], E, E, KS, st))) return rc;
    if ((rc = make_w_map(&c->enc_wmap[i], c->enc_Wb[i], E, E, KS, 128, c->npass))) return rc;
    if ((rc = bn_fold_launch(bw->ptr, bb->ptr, bm->ptr, bv->ptr, cb->ptr, c->enc_scale[i], c->enc_shift[i], E, st))) return rc;
    pack_conv_wT_kernel<<<gsz((long long)E * E * KS), 256, 0, st>>>(cw->ptr, c->enc_WTb[i], E, E, KS);
    if ((rc = make_w_map(&c->enc_wtmap[i], c->enc_WTb[i], E, E, KS, 128, c->npass))) return rc;
  }
  for (int i = 0; i < 5; ++i) {
    const std::string cv = "postnet.postnet." + std::to_string(4 * i), bn = "postnet.postnet." + std::to_string(4 * i + 1);
    NEED(cw, cv + ".weight", (size_t)chans[i + 1] * chans[i] * 5);
    NEED(bw, bn + ".weight", chans[i + 1]);
    NEED(bb, bn + ".bias", chans[i + 1]);
    NEED(bm, bn + ".running_mean", chans[i + 1]);
    NEED(bv, bn + ".running_var", chans[i + 1]);
    if ((rc = pack_conv_w_launch(cw->ptr, c->post_Wb[i], chans[i + 1], chans[i], 5, st))) return rc;
    if ((rc = make_w_map(&c->post_wmap[i], c->post_Wb[i], chans[i + 1], chans[i], 5, chans[i + 1] % 128 == 0 ? 128 : chans[i + 1], c->npass))) return rc;
    if ((rc = bn_fold_launch(bw->ptr, bb->ptr, bm->ptr, bv->ptr, nullptr, c->post_scale[i], c->post_shift[i], chans[i + 1], st))) return rc;
    pack_conv_wT_kernel<<<gsz((long long)chans[i + 1] * chans[i] * 5), 256, 0, st>>>(cw->ptr, c->post_WTb[i], chans[i + 1], chans[i], 5);
    if ((rc = make_w_map(&c->post_wtmap[i], c->post_WTb[i], chans[i], chans[i + 1], 5, chans[i] % 128 == 0 ? 128 : chans[i], c->npass))) return rc;
  }
  // BiLSTM: input projection for both directions as one k=1 "conv": Wr [E][8Hh]; recurrent W_hh^T per direction
  for (int dir = 0; dir < 2; ++dir) {
    const std::string sfx = dir ? "_reverse" : "";
    NEED(wih, "encoder.lstm.weight_ih_l0" + sfx, (size_t)4 * Hh * E);
    NEED(whh, "encoder.lstm.weight_hh_l0" + sfx, (size_t)4 * Hh * Hh);
    NEED(bih, "encoder.lstm.bias_ih_l0" + sfx, 4 * Hh);
    NEED(bhh, "encoder.lstm.bias_hh_l0" + sfx, 4 * Hh);
    T2_CHECK_CUDA(cudaMemcpyAsync(tmp + (size_t)dir * 4 * Hh * E, wih->ptr, (size_t)4 * Hh * E * sizeof(float),
                                  cudaMemcpyDeviceToDevice, st));
    T2_CHECK_CUDA(cudaMemcpyAsync(c->lstm_Whh + (size_t)dir * 4 * Hh * Hh, whh->ptr, (size_t)4 * Hh * Hh * sizeof(float),
                                  cudaMemcpyDeviceToDevice, st));
    add_vec_kernel<<<(4 * Hh + 255) / 256, 256, 0, st>>>(bih->ptr, bhh->ptr, c->lstm_bias + dir * 4 * Hh, 4 * Hh);
  }
  if ((rc = pack_conv_w_launch(tmp, c->lstm_Wb, 8 * Hh, E, 1, st))) return rc;   // tmp = [8Hh][E] = conv weight with k = 1
  if ((rc = make_w_map(&c->lstm_wmap, c->lstm_Wb, 8 * Hh, E, 1, 128, c->npass))) return rc;
  if ((rc = pack_conv_w_launch(wae->ptr, c->attenc_Wb, A, EF, 1, st))) return rc;
  if ((rc = make_w_map(&c->attenc_wmap, c->attenc_Wb, A, EF, 1, A, c->npass))) return rc;
  pack_conv_wT_kernel<<<gsz((long long)8 * Hh * E), 256, 0, st>>>(tmp, c->lstm_WTb, 8 * Hh, E, 1);
  if ((rc = make_w_map(&c->lstm_wtmap, c->lstm_WTb, E, 8 * Hh, 1, 128, c->npass))) return rc;
  pack_conv_wT_kernel<<<gsz((long long)A * EF), 256, 0, st>>>(wae->ptr, c->attenc_WTb, A, EF, 1);
  if ((rc = make_w_map(&c->attenc_wtmap, c->attenc_WTb, EF, A, 1, 128, c->npass))) return rc;
  T2_CHECK_CUDA(cudaGetLastError());
  c->committed = true;
  return T2_OK;
}

extern "C" int t2_ctx_commit(t2_ctx* c, void* stream) {
  T2_REQUIRE(c, T2_ERR_ARG, "t2_ctx_commit: null context");
  if (!c->slab) {
    Arena sizing(nullptr, 0);
    commit_impl(c, sizing, nullptr);
    c->slab_bytes = (size_t)sizing.off;
    T2_CHECK_CUDA(cudaMalloc((void**)&c->slab, c->slab_bytes));
  }
  Arena ar(c->slab, (int64_t)c->slab_bytes);
  return commit_impl(c, ar, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ ops
int conv_tc(t2_ctx* c, const __nv_bfloat16* in, const CUtensorMap& wmap, int B, int T, int Cin, int Cout, int k, int NT, int act,
                   const float* scale, const float* shift, const unsigned char* keep, int keep_ld, float* out_f32,
                   __nv_bfloat16* out_b16, cudaStream_t st) {
  ConvTcArgs a{};
  RC(make_in_map(&a.in_map, in, B, T, Cin, c->npass));
  a.w_map = wmap;
  a.B = B; a.T = T; a.Cin = Cin; a.Cout = Cout; a.k = k; a.NT = NT; a.npass = c->npass; a.act = act;
  a.scale = scale; a.shift = shift; a.keep = keep; a.keep_ld = keep_ld; a.drop_scale = 1.0f / (1.0f - c->cfg.dropout);
  a.out_f32 = out_f32; a.out_b16 = out_b16; a.residual = nullptr; a.res_bstride = 0;
  return convtc_launch(a, st);
}

static int encoder_impl(t2_ctx* c, const long long* chars, const long long* lens, int B, int L, int train,
                        const unsigned char* keep, float* encoded, Arena& ar, cudaStream_t st) {
  const int E = c->E, Hh = E / 2;
  __nv_bfloat16* x0 = ar.take<__nv_bfloat16>((size_t)2 * B * L * E);
  __nv_bfloat16* x1 = ar.take<__nv_bfloat16>((size_t)2 * B * L * E);
  float* xproj = ar.take<float>((size_t)B * L * 8 * Hh);
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_encoder_fwd: workspace too small");
  T2_REQUIRE(!train, T2_ERR_UNSUPPORTED, "t2_encoder_fwd: train-mode BatchNorm statistics are not implemented in this revision");
  RC(embed_split_launch(chars, find(c, "encoder.embedding.weight")->ptr, x0, B * L, E, c->cfg.num_chars + 1, st));
  __nv_bfloat16* cur = x0;
  __nv_bfloat16* nxt = x1;
  for (int i = 0; i < 3; ++i) {  // Conv1d k5 + BatchNorm (folded) + ReLU (+ Dropout)  model/encoder.py:31-46
    RC(conv_tc(c, cur, c->enc_wmap[i], B, L, E, E, c->KS, 128, ACT_RELU, c->enc_scale[i], c->enc_shift[i],
               keep ? keep + (size_t)i * B * L * E : nullptr, E, nullptr, nxt, st));
    std::swap(cur, nxt);
  }
  // BiLSTM input projection for both directions (k = 1), then the packed recurrence  model/encoder.py:59-65
  RC(conv_tc(c, cur, c->lstm_wmap, B, L, E, 8 * Hh, 1, 128, ACT_NONE, nullptr, c->lstm_bias, nullptr, 0, xproj, nullptr, st));
  RC(bilstm_launch(xproj, c->lstm_Whh, lens, encoded, B, L, Hh, st));
  return T2_OK;
}

extern "C" int t2_encoder_fwd(t2_ctx* c, const long long* chars_idx, const long long* lens, int B, int L, int train,
                              const unsigned char* keep, float* encoded, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_encoder_fwd");
  T2_REQUIRE(chars_idx && lens && encoded && ws && B > 0 && L > 0, T2_ERR_ARG, "t2_encoder_fwd: bad argument");
  T2_REQUIRE(find(c, "encoder.embedding.weight"), T2_ERR_STATE, "t2_encoder_fwd: encoder.embedding.weight not set");
  Arena ar(ws, ws_bytes);
  return encoder_impl(c, chars_idx, lens, B, L, train, keep, encoded, ar, (cudaStream_t)stream);
}

static int memory_impl(t2_ctx* c, const float* encoded, const long long* spk, const float* desc, int B, int L, float* memory,
                       float* pmem, Arena& ar, cudaStream_t st) {
  const int E = c->E, EF = c->EF, D = EF - E;
  float* dvec = ar.take<float>((size_t)B * 128);
  __nv_bfloat16* membf = ar.take<__nv_bfloat16>((size_t)2 * B * L * (c->EF));
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_memory_fwd: workspace too small");
  const float* spk_table = nullptr;
  if (c->cfg.num_speakers > 0) {
    T2_REQUIRE(spk, T2_ERR_ARG, "speaker_id tensor required when speaker tokens are active!");  // model/tacotron2.py:175-177
    const Param* p = find(c, "speaker_embedding.weight");
    T2_REQUIRE(p && p->numel == (long long)c->cfg.num_speakers * E, T2_ERR_STATE, "speaker_embedding.weight missing");
    spk_table = p->ptr;
  }
  if (D > 0) {
    T2_REQUIRE(desc, T2_ERR_ARG, "description tensor required when description tokens are active!");  // :179-181
    const Param* w = find(c, "description_embeddings_linear.0.weight");
    const Param* b = find(c, "description_embeddings_linear.0.bias");
    T2_REQUIRE(w && b && w->numel == 128ll * c->cfg.description_dim, T2_ERR_STATE, "description_embeddings_linear missing");
    RC(desc_linear_launch(desc, w->ptr, b->ptr, dvec, B, c->cfg.description_dim, 128, st));
  }
  if (spk_table || D > 0) RC(condition_launch(encoded, spk_table, spk, dvec, D, memory, B, L, E, st));
  else T2_CHECK_CUDA(cudaMemcpyAsync(memory, encoded, (size_t)B * L * E * sizeof(float), cudaMemcpyDeviceToDevice, st));
  RC(split_cl_launch(memory, (long long)L * EF, membf, B, L, EF, st));
  RC(conv_tc(c, membf, c->attenc_wmap, B, L, EF, c->A, 1, c->A, ACT_NONE, nullptr, nullptr, nullptr, 0, pmem, nullptr, st));  // model/tacotron2.py:229
  return T2_OK;
}

extern "C" int t2_memory_fwd(t2_ctx* c, const float* encoded, const long long* speaker_id, const float* description, int B, int L,
                             float* memory, float* processed_memory, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_memory_fwd");
  T2_REQUIRE(encoded && memory && processed_memory && ws && B > 0 && L > 0, T2_ERR_ARG, "t2_memory_fwd: bad argument");
  Arena ar(ws, ws_bytes);
  return memory_impl(c, encoded, speaker_id, description, B, L, memory, processed_memory, ar, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ decoder
struct State2Plan { D2State s; char* zero_begin; size_t zero_bytes; };
static State2Plan plan_state2(const D2Dims& d, int L, Arena& ar) {
  State2Plan pl{};
  const int64_t off0 = ar.off;
  D2State& s = pl.s;
  s.xatt = ar.take<__nv_bfloat16>((size_t)4 * D2_BP * d.KattP);
  s.xdec = ar.take<__nv_bfloat16>((size_t)4 * D2_BP * d.KdecP);
  s.att_c = ar.take<float>((size_t)d.H * D2_BP);
  s.rnn_c = ar.take<float>((size_t)d.R * D2_BP);
  s.att_h_rm = ar.take<float>((size_t)D2_BP * d.H);
  s.projin = ar.take<float>((size_t)D2_BP * d.KprojP);
  s.q = nullptr;
  s.w_prev = ar.take<float>((size_t)2 * D2_BP * L);
  s.w_cum = ar.take<float>((size_t)2 * D2_BP * L);
  s.mel_cur = ar.take<float>((size_t)D2_BP * d.M);
  s.p1 = ar.take<float>((size_t)D2_BP * d.P);
  s.done = ar.take<int>(D2_BP);
  s.lengths = ar.take<long long>(D2_BP);
  s.n_frames = ar.take<int>(4);
  s.barrier = ar.take<unsigned>(4);
  s.bar_flags = ar.take<unsigned>(256);   // pipelined AR kernel: one arrival word per CTA
  pl.zero_begin = ar.base ? ar.base + off0 : nullptr;
  pl.zero_bytes = (size_t)(ar.off - off0);
  s.qpart = ar.take<float>((size_t)D2_BP * 128 * d.A);   // fully rewritten every frame: not part of the zeroed region
  {
    const int S = d2_att_parts(d.B);
    s.accpm = ar.take<float>((size_t)128 * ((L + S - 1) / S) * d.A);
  }
  return pl;
}

int decode_proj_bulk(t2_ctx* c, const __nv_bfloat16* xdec_tape, int B, int T, float* mels, float* gates, float* tmp, cudaStream_t st);  // train_capi.cu

D2Dims dims2(const t2_ctx* c, int B, int L) {
  D2Dims d{};
  d.B = B; d.L = L; d.EF = c->EF; d.H = c->H; d.R = c->R; d.P = c->P; d.A = c->A; d.M = c->M; d.C = c->C;
  d.KattP = c->dd.KattP; d.KdecP = c->dd.KdecP; d.KprojP = c->dd.KprojP;
  d.npass = c->npass; d.p_prenet = c->cfg.dropout;
  return d;
}

// activation tensor maps: x[4 = parity*2 + hl][BP][K] bf16, box {64, 64, 1|2}
static int make_xmaps(const t2_ctx* c, const D2Dims& d, const D2State& s, const __nv_bfloat16* tfp, int T1, D2Maps* m, int nslots = 2) {
  const unsigned zb = d.npass == 3 ? 2u : 1u;
  unsigned box[3] = {TG_BK, TG_N, zb};
  int rc;
  {
    unsigned long long dims[3] = {(unsigned long long)d.KattP, (unsigned long long)D2_BP, 2ull * nslots};
    unsigned long long st[2] = {(unsigned long long)d.KattP * 2, (unsigned long long)D2_BP * d.KattP * 2};
    if ((rc = tc_make_tmap_bf16(&m->x_att, s.xatt, 3, dims, st, box))) return rc;
    unsigned long long dims_r[3] = {(unsigned long long)(d.KattP - d.P), (unsigned long long)D2_BP, 2ull * nslots};
    if ((rc = tc_make_tmap_bf16(&m->x_att_rest, s.xatt + d.P, 3, dims_r, st, box))) return rc;
  }
  {
    unsigned long long dims[3] = {(unsigned long long)d.KdecP, (unsigned long long)D2_BP, 2ull * nslots};
    unsigned long long st[2] = {(unsigned long long)d.KdecP * 2, (unsigned long long)D2_BP * d.KdecP * 2};
    if ((rc = tc_make_tmap_bf16(&m->x_dec, s.xdec, 3, dims, st, box))) return rc;
  }
  if (tfp) {
    unsigned long long dims[3] = {(unsigned long long)d.P, (unsigned long long)D2_BP, (unsigned long long)T1 * 2};
    unsigned long long st[2] = {(unsigned long long)d.P * 2, (unsigned long long)D2_BP * d.P * 2};
    if ((rc = tc_make_tmap_bf16(&m->tf_pre, tfp, 3, dims, st, box))) return rc;
  } else {
    m->tf_pre = m->x_att;
  }
  m->w_att = c->map_w_att;
  m->w_dec = c->map_w_dec;
  return T2_OK;
}

static int fill_controls2(const D2Dims& d, D2State& s, const float* controls, int B, cudaStream_t st, int nslots = 2) {
  if (d.C == 0) return T2_OK;
  T2_REQUIRE(controls, T2_ERR_ARG, "Controls are enabled, but no control vector was passed to the model!");  // model/tacotron2.py:183-185
  const long long pd = (long long)D2_BP * d.KdecP;
  controls_slots_kernel<<<gsz((long long)nslots * B * d.C), 256, 0, st>>>(controls, d.C, s.xdec, pd, d.KdecP, d.H + d.EF, B, nslots);
  copy2d_kernel<<<gsz(B * d.C), 256, 0, st>>>(controls, d.C, s.projin + d.R + d.EF, d.KprojP, B, d.C);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

static int decode2_impl(t2_ctx* c, bool tf, const float* memory, const float* pmem, const unsigned char* mask, const float* controls,
                        const float* mel_target, int B, int L, int T, const unsigned char* prenet_keep, int train,
                        const unsigned char* att_keep, const unsigned char* rnn_keep, unsigned long long seed, float* mels,
                        float* gates, float* aligns, long long* lengths, int* n_frames, Arena& ar, cudaStream_t st,
                        const DecTape* tape, int min_frames = 0) {
  D2Dims d = dims2(c, B, L);
  State2Plan pl = plan_state2(d, L, ar);
  __nv_bfloat16* tfp = (tf && !tape) ? ar.take<__nv_bfloat16>((size_t)(T + 1) * 2 * D2_BP * d.P) : nullptr;
  float* proj_tmp = tape ? ar.take<float>((size_t)(d.M + 1) * T * D2_BP) : nullptr;
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_decode: workspace too small");
  T2_CHECK_CUDA(cudaMemsetAsync(pl.zero_begin, 0, pl.zero_bytes, st));
  D2State s = pl.s;
  s.prof = c->prof_buf;
  s.status = c->status_dev;
  T2_CHECK_CUDA(cudaMemsetAsync(c->prof_buf, 0, 16 * sizeof(unsigned long long), st));
  const int nslots = tape ? T + 1 : 2;
  if (tape) {  // training: every per-frame buffer keeps its whole history
    T2_CHECK_CUDA(cudaMemsetAsync(tape->zero_begin, 0, tape->zero_bytes, st));
    tfp = tape->tfp;
    s.xatt = tape->xatt; s.xdec = tape->xdec; s.q = tape->q; s.w_prev = tape->w_hist; s.w_cum = tape->wcum_hist;
    s.sav_att = tape->sav_att; s.sav_dec = tape->sav_dec;
    s.accpm = tape->xloc; s.floc = tape->floc;
  }
  RC(fill_controls2(d, s, controls, B, st, nslots));
  D2IO io{};
  io.memory = memory; io.pmem = pmem; io.mask = mask; io.prenet_keep = prenet_keep;
  io.att_keep = att_keep; io.rnn_keep = rnn_keep; io.mels = mels; io.gates = gates; io.aligns = aligns;
  io.Tmax = T; io.n_steps = T; io.t0 = 0; io.seed = seed; io.hidden_dropout = train ? 1 : 0;
  io.debug_barrier = getenv("T2B200_DEBUG_BARRIER") ? atoi(getenv("T2B200_DEBUG_BARRIER")) : 0;   // experiment switches (bit mask)
  io.hist = tape ? 1 : 0;
  if (tf) {
    if (!tape) T2_CHECK_CUDA(cudaMemsetAsync(tfp, 0, (size_t)(T + 1) * 2 * D2_BP * d.P * sizeof(__nv_bfloat16), st));
    PrenetArgs pa{mel_target, c->d2w.W1, c->d2w.W2, prenet_keep, prenet_keep, seed, d.p_prenet, (T + 1) * B, d.M, d.P, 2, B, T, nullptr, tfp, tape ? tape->p1 : nullptr};
    RC(prenet_rows_launch(pa, st));
    io.teacher_forcing = 1; io.stop_enabled = 0;
  } else {
    io.teacher_forcing = 0; io.stop_enabled = 1; io.min_frames = min_frames;
    T2_CHECK_CUDA(cudaMemsetAsync(lengths, 0, (size_t)B * sizeof(long long), st));
    s.lengths = lengths;
    s.n_frames = n_frames;
  }
  D2Maps maps;
  RC(make_xmaps(c, d, s, tfp, T + 1, &maps, nslots));
  timing_begin(c, 0, st);
  // pipelined schedule: autoregressive synthesis and the teacher-forced training forward (history mode)
  if (((!tf && !tape) || (tf && tape)) && d2ar_supported(d)) {
    if (tape) s.floc = nullptr;   // the pipelined kernel tapes the location term (xloc) but not the conv features: the backward recomputes them
    RC(d2ar_launch(d, c->d2w, s, io, maps, c->map_locG, st));
  }
  else RC(d2_launch(d, c->d2w, s, io, maps, st));
  timing_end(c, 0, st);
  if (tape) RC(decode_proj_bulk(c, tape->xdec, B, T, mels, gates, proj_tmp, st));
  RC(poison_on_abort_launch(c->status_dev, mels, (long long)B * T * d.M, st));   // an aborted launch must not pass for a result
  return T2_OK;
}

static int step2_impl(t2_ctx* c, const float* prev, const float* att_h, const float* att_c, const float* ctx, const float* w,
                      const float* w_cum, const float* rnn_h, const float* rnn_c, const float* memory, const float* pmem,
                      const unsigned char* mask, const float* controls, int B, int L, float* mel_out, float* gate_out,
                      float* att_h_out, float* att_c_out, float* ctx_out, float* w_out, float* w_cum_out, float* rnn_h_out,
                      float* rnn_c_out, Arena& ar, cudaStream_t st) {
  D2Dims d = dims2(c, B, L);
  State2Plan pl = plan_state2(d, L, ar);
  __nv_bfloat16* tfp = ar.take<__nv_bfloat16>((size_t)2 * D2_BP * d.P);
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_decoder_step: workspace too small");
  T2_CHECK_CUDA(cudaMemsetAsync(pl.zero_begin, 0, pl.zero_bytes, st));
  T2_CHECK_CUDA(cudaMemsetAsync(tfp, 0, (size_t)2 * D2_BP * d.P * sizeof(__nv_bfloat16), st));
  D2State s = pl.s;
  s.status = c->status_dev;
  RC(fill_controls2(d, s, controls, B, st));
  const long long pa = (long long)D2_BP * d.KattP, pd = (long long)D2_BP * d.KdecP, pp = (long long)D2_BP * d.P;
  rm_to_xsplit_kernel<<<gsz(B * d.P), 256, 0, st>>>(prev, d.P, tfp, pp, d.P, 0, d.P, B);
  rm_to_xsplit_kernel<<<gsz(B * d.EF), 256, 0, st>>>(ctx, d.EF, s.xatt, pa, d.KattP, d.P, d.EF, B);
  rm_to_xsplit_kernel<<<gsz(B * d.H), 256, 0, st>>>(att_h, d.H, s.xatt, pa, d.KattP, d.P + d.EF, d.H, B);
  rm_to_xsplit_kernel<<<gsz(B * d.R), 256, 0, st>>>(rnn_h, d.R, s.xdec, pd, d.KdecP, d.H + d.EF + d.C, d.R, B);
  rm_to_km_kernel<<<gsz(d.H * D2_BP), 256, 0, st>>>(att_c, d.H, s.att_c, 0, d.H, B);
  rm_to_km_kernel<<<gsz(d.R * D2_BP), 256, 0, st>>>(rnn_c, d.R, s.rnn_c, 0, d.R, B);
  copy2d_kernel<<<gsz(B * L), 256, 0, st>>>(w, L, s.w_prev, L, B, L);
  copy2d_kernel<<<gsz(B * L), 256, 0, st>>>(w_cum, L, s.w_cum, L, B, L);
  T2_CHECK_CUDA(cudaGetLastError());
  D2IO io{};
  io.memory = memory; io.pmem = pmem; io.mask = mask; io.teacher_forcing = 1; io.stop_enabled = 0;
  io.mels = mel_out; io.gates = gate_out; io.aligns = w_out; io.Tmax = 1; io.n_steps = 1; io.t0 = 0;
  D2Maps maps;
  RC(make_xmaps(c, d, s, tfp, 1, &maps));
  RC(d2_launch(d, c->d2w, s, io, maps, st));
  copy2d_kernel<<<gsz(B * d.H), 256, 0, st>>>(s.att_h_rm, d.H, att_h_out, d.H, B, d.H);
  km_to_rm_kernel<<<gsz(B * d.H), 256, 0, st>>>(s.att_c, 0, att_c_out, d.H, d.H, B);
  copy2d_kernel<<<gsz(B * d.EF), 256, 0, st>>>(s.projin + d.R, d.KprojP, ctx_out, d.EF, B, d.EF);
  copy2d_kernel<<<gsz(B * L), 256, 0, st>>>(s.w_cum + (size_t)D2_BP * L, L, w_cum_out, L, B, L);
  copy2d_kernel<<<gsz(B * d.R), 256, 0, st>>>(s.projin, d.KprojP, rnn_h_out, d.R, B, d.R);
  km_to_rm_kernel<<<gsz(B * d.R), 256, 0, st>>>(s.rnn_c, 0, rnn_c_out, d.R, d.R, B);
  T2_CHECK_CUDA(cudaGetLastError());
  RC(poison_on_abort_launch(c->status_dev, mel_out, (long long)B * d.M, st));
  return T2_OK;
}

extern "C" int t2_decode_ar(t2_ctx* c, const float* memory, const float* processed_memory, const unsigned char* mask,
                            const float* controls, int B, int L, int max_len, const unsigned char* prenet_keep,
                            unsigned long long seed, float* mels, float* gates, float* alignments, long long* lengths,
                            int* n_frames, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_decode_ar");
  T2_REQUIRE(memory && processed_memory && mask && mels && gates && alignments && lengths && n_frames && ws, T2_ERR_ARG,
             "t2_decode_ar: null argument");
  T2_REQUIRE(B > 0 && L > 0 && max_len > 0, T2_ERR_ARG, "t2_decode_ar: bad sizes B=%d L=%d max_len=%d", B, L, max_len);
  Arena ar(ws, ws_bytes);
  return decode2_impl(c, false, memory, processed_memory, mask, controls, nullptr, B, L, max_len, prenet_keep,
                                                 0, nullptr, nullptr, seed, mels, gates, alignments, lengths, n_frames, ar,
                                                 (cudaStream_t)stream, nullptr);
}

extern "C" int t2_decode_ar_chunk(t2_ctx* c, const float* memory, const float* processed_memory, const unsigned char* mask,
                                  const float* controls, int B, int L, int max_len, int min_frames, const unsigned char* prenet_keep,
                                  unsigned long long seed, float* mels, float* gates, float* alignments, long long* lengths,
                                  int* n_frames, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_decode_ar_chunk");
  T2_REQUIRE(memory && processed_memory && mask && mels && gates && alignments && lengths && n_frames && ws, T2_ERR_ARG,
             "t2_decode_ar_chunk: null argument");
  T2_REQUIRE(B > 0 && L > 0 && max_len > 0 && min_frames >= 0 && min_frames <= max_len, T2_ERR_ARG,
             "t2_decode_ar_chunk: bad sizes B=%d L=%d max_len=%d min_frames=%d", B, L, max_len, min_frames);
  Arena ar(ws, ws_bytes);
  return decode2_impl(c, false, memory, processed_memory, mask, controls, nullptr, B, L, max_len, prenet_keep, 0, nullptr, nullptr, seed,
                      mels, gates, alignments, lengths, n_frames, ar, (cudaStream_t)stream, nullptr, min_frames);
}

extern "C" int t2_decode_tf(t2_ctx* c, const float* memory, const float* processed_memory, const unsigned char* mask,
                            const float* controls, const float* mel_target, int B, int L, int T, const unsigned char* prenet_keep,
                            int train, const unsigned char* att_keep, const unsigned char* rnn_keep, unsigned long long seed,
                            float* mels, float* gates, float* alignments, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_decode_tf");
  T2_REQUIRE(memory && processed_memory && mask && mel_target && mels && gates && alignments && ws, T2_ERR_ARG,
             "Ground-truth Mel spectrogram is required for teacher forcing");  // model/tacotron2.py:168-170
  T2_REQUIRE(B > 0 && L > 0 && T > 0, T2_ERR_ARG, "t2_decode_tf: bad sizes");
  Arena ar(ws, ws_bytes);
  return decode2_impl(c, true, memory, processed_memory, mask, controls, mel_target, B, L, T, prenet_keep,
                                                 train, att_keep, rnn_keep, seed, mels, gates, alignments, nullptr, nullptr, ar,
                                                 (cudaStream_t)stream, nullptr);
}

extern "C" long long t2_decode_tape_bytes(const t2_ctx* c, int B, int L, int T) {
  if (!c || B <= 0 || L <= 0 || T <= 0) return -1;
  Arena ar(nullptr, 0);
  plan_dec_tape(dims2(c, B, L), B, L, T, ar);
  return ar.off + 1024;
}

// Teacher-forced frame loop under train(): same kernel in history mode (every per-frame buffer, the LSTM gate activations,
// queries and attention weights are kept on the tape for t2_decode_tf_bwd); hidden-state dropouts p = 0.1 are active.
extern "C" int t2_decode_tf_train_fwd(t2_ctx* c, const float* memory, const float* processed_memory, const unsigned char* mask,
                                      const float* controls, const float* mel_target, int B, int L, int T,
                                      const unsigned char* prenet_keep, const unsigned char* att_keep, const unsigned char* rnn_keep,
                                      unsigned long long seed, float* mels, float* gates, float* alignments, void* tape,
                                      long long tape_bytes, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_decode_tf_train_fwd");
  T2_REQUIRE(memory && processed_memory && mask && mel_target && mels && gates && alignments && ws && tape, T2_ERR_ARG,
             "Ground-truth Mel spectrogram is required for teacher forcing");
  T2_REQUIRE(B > 0 && L > 0 && T > 0, T2_ERR_ARG, "t2_decode_tf_train_fwd: bad sizes");
  Arena ta(tape, tape_bytes);
  DecTape tp = plan_dec_tape(dims2(c, B, L), B, L, T, ta);
  T2_REQUIRE(!ta.overflow, T2_ERR_WORKSPACE, "t2_decode_tf_train_fwd: tape too small");
  Arena ar(ws, ws_bytes);
  return decode2_impl(c, true, memory, processed_memory, mask, controls, mel_target, B, L, T, prenet_keep, 1, att_keep, rnn_keep, seed,
                      mels, gates, alignments, nullptr, nullptr, ar, (cudaStream_t)stream, &tp);
}

extern "C" int t2_decoder_step(t2_ctx* c, const float* prev_mel_prenet, const float* att_h, const float* att_c, const float* att_ctx,
                               const float* w, const float* w_cum, const float* rnn_h, const float* rnn_c, const float* memory,
                               const float* processed_memory, const unsigned char* mask, const float* controls, int B, int L,
                               float* mel_out, float* gate_out, float* att_h_out, float* att_c_out, float* att_ctx_out,
                               float* w_out, float* w_cum_out, float* rnn_h_out, float* rnn_c_out, void* ws, long long ws_bytes,
                               void* stream) {
  CHECK_CTX("t2_decoder_step");
  T2_REQUIRE(prev_mel_prenet && att_h && att_c && att_ctx && w && w_cum && rnn_h && rnn_c && memory && processed_memory && mask &&
                 mel_out && gate_out && att_h_out && att_c_out && att_ctx_out && w_out && w_cum_out && rnn_h_out && rnn_c_out && ws,
             T2_ERR_ARG, "t2_decoder_step: null argument");
  Arena ar(ws, ws_bytes);
  return step2_impl(c, prev_mel_prenet, att_h, att_c, att_ctx, w, w_cum, rnn_h, rnn_c, memory, processed_memory,
                                             mask, controls, B, L, mel_out, gate_out, att_h_out, att_c_out, att_ctx_out, w_out,
                                             w_cum_out, rnn_h_out, rnn_c_out, ar, (cudaStream_t)stream);
}

extern "C" int t2_prenet_fwd(t2_ctx* c, const float* x, const unsigned char* keep1, const unsigned char* keep2,
                             unsigned long long seed, int N, float* out, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_prenet_fwd");
  T2_REQUIRE(x && out && N > 0, T2_ERR_ARG, "t2_prenet_fwd: bad argument");
  (void)ws; (void)ws_bytes;
  PrenetArgs pa{x, c->d2w.W1, c->d2w.W2, keep1, keep2, seed, c->cfg.dropout, N, c->M, c->P, 0, 0, 0, out, nullptr, nullptr};
  return prenet_rows_launch(pa, (cudaStream_t)stream);
}

static int postnet_impl(t2_ctx* c, const float* mels, int t_stride, int B, int T, int train, const unsigned char* keep, float* out,
                        Arena& ar, cudaStream_t st) {
  const int M = c->M, PD = c->PD;
  __nv_bfloat16* xin = ar.take<__nv_bfloat16>((size_t)2 * B * T * M);
  __nv_bfloat16* a0 = ar.take<__nv_bfloat16>((size_t)2 * B * T * PD);
  __nv_bfloat16* a1 = ar.take<__nv_bfloat16>((size_t)2 * B * T * PD);
  if (!ar.base) return T2_OK;
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_postnet_fwd: workspace too small");
  T2_REQUIRE(!train, T2_ERR_UNSUPPORTED, "t2_postnet_fwd: train-mode BatchNorm statistics are not implemented in this revision");
  const int chans[6] = {M, PD, PD, PD, PD, M};
  RC(split_cl_launch(mels, (long long)t_stride * M, xin, B, T, M, st));
  const __nv_bfloat16* in = xin;
  __nv_bfloat16* bufs[2] = {a0, a1};
  for (int i = 0; i < 5; ++i) {  // Conv1d k5 'same' + BatchNorm (folded) (+ Tanh) (+ Dropout)  model/postnet.py:8-47
    const int Cout = chans[i + 1], NT = Cout % 128 == 0 ? 128 : Cout;
    RC(conv_tc(c, in, c->post_wmap[i], B, T, chans[i], Cout, 5, NT, i < 4 ? ACT_TANH : ACT_NONE, c->post_scale[i], c->post_shift[i],
               keep ? keep + (size_t)i * B * T * PD : nullptr, PD, i == 4 ? out : nullptr, i == 4 ? nullptr : bufs[i & 1], st));
    in = bufs[i & 1];
  }
  return T2_OK;
}

extern "C" int t2_postnet_fwd(t2_ctx* c, const float* mels, int t_stride, int B, int T, int train, const unsigned char* keep,
                              float* out, void* ws, long long ws_bytes, void* stream) {
  CHECK_CTX("t2_postnet_fwd");
  T2_REQUIRE(mels && out && ws && B > 0 && T > 0 && t_stride >= T, T2_ERR_ARG, "t2_postnet_fwd: bad argument");
  Arena ar(ws, ws_bytes);
  return postnet_impl(c, mels, t_stride, B, T, train, keep, out, ar, (cudaStream_t)stream);
}

extern "C" int t2_finalize(t2_ctx* c, float* mels, float* post, float* gates, const long long* lengths, int B, int T, int t_stride,
                           void* stream) {
  T2_REQUIRE(c && mels && post && gates && lengths, T2_ERR_ARG, "t2_finalize: null argument");
  return finalize_launch(mels, post, gates, lengths, B, T, t_stride, c->M, (cudaStream_t)stream);
}

extern "C" int t2_ctx_status(t2_ctx* c, int* status_out) {
  T2_REQUIRE(c && status_out, T2_ERR_ARG, "t2_ctx_status: null argument");
  unsigned v = 0;
  T2_CHECK_CUDA(cudaMemcpy(&v, c->status_dev, sizeof(unsigned), cudaMemcpyDeviceToHost));   // synchronises with the device
  if (v) T2_CHECK_CUDA(cudaMemset(c->status_dev, 0, sizeof(unsigned)));
  *status_out = v ? T2_ERR_DEVICE_ABORT : 0;
  if (v) t2_set_error("a persistent decoder kernel aborted (grid barrier or TMA pipeline timeout); its outputs were NaN-filled");
  return T2_OK;
}

extern "C" int t2_debug_set_timing(t2_ctx* c, int on) {
  T2_REQUIRE(c, T2_ERR_ARG, "t2_debug_set_timing: null context");
  c->timing_on = on ? 1 : 0;
  c->tev_valid[0] = c->tev_valid[1] = false;
  return T2_OK;
}

extern "C" int t2_debug_kernel_ms(t2_ctx* c, int which, float* ms_out) {
  T2_REQUIRE(c && ms_out && (which == 0 || which == 1), T2_ERR_ARG, "t2_debug_kernel_ms: bad argument");
  T2_REQUIRE(c->tev_valid[which], T2_ERR_STATE, "t2_debug_kernel_ms: no timed launch of kernel %d (call t2_debug_set_timing first)", which);
  T2_CHECK_CUDA(cudaEventSynchronize(c->tev[which][1]));
  T2_CHECK_CUDA(cudaEventElapsedTime(ms_out, c->tev[which][0], c->tev[which][1]));
  return T2_OK;
}

extern "C" int t2_debug_phase_cycles(t2_ctx* c, unsigned long long* host_out16) {
  T2_REQUIRE(c && host_out16, T2_ERR_ARG, "t2_debug_phase_cycles: null argument");
  T2_CHECK_CUDA(cudaMemcpy(host_out16, c->prof_buf, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return T2_OK;
}

extern "C" long long t2_workspace_bytes(const t2_ctx* cc, int B, int L, int T) {
  t2_ctx* c = const_cast<t2_ctx*>(cc);
  if (!c || B <= 0 || L <= 0 || T < 0) return -1;
  long long best = 0;
  { Arena a(nullptr, 0); encoder_impl(c, nullptr, nullptr, B, L, 0, nullptr, nullptr, a, nullptr); best = std::max<long long>(best, a.off); }
  { Arena a(nullptr, 0); memory_impl(c, nullptr, nullptr, nullptr, B, L, nullptr, nullptr, a, nullptr); best = std::max<long long>(best, a.off); }
  { Arena a(nullptr, 0); decode2_impl(c, true, nullptr, nullptr, nullptr, nullptr, nullptr, B, L, std::max(T, 1), nullptr, 0, nullptr, nullptr, 0,
                                     nullptr, nullptr, nullptr, nullptr, nullptr, a, nullptr, nullptr); best = std::max<long long>(best, a.off); }
  { Arena a(nullptr, 0); postnet_impl(c, nullptr, std::max(T, 1), B, std::max(T, 1), 0, nullptr, nullptr, a, nullptr); best = std::max<long long>(best, a.off); }
  return best + 1024;
}
