// Host helpers for the tcgen05 split-K GEMM + a self-contained debug kernel (t2_debug_tcgemm) that exercises
// exactly the device path the decoder uses: TMA -> swizzled smem ring -> tcgen05.mma -> TMEM -> DSMEM reduce.
#include "tcgemm.cuh"

#include <cudaTypedefs.h>

namespace {

__global__ void split_bf16_kernel(const float* src, __nv_bfloat16* dst, int rows, int cols, int ld_src, int ld_dst, long long plane) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)rows * ld_dst; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / ld_dst), c = (int)(i % ld_dst);
    const float v = c < cols ? src[(size_t)r * ld_src + c] : 0.f;
    __nv_bfloat16 hi, lo;
    tc::split_bf16(v, hi, lo);
    dst[i] = hi;
    dst[plane + i] = lo;
  }
}

// grid = (rows/128) clusters of 4 CTAs; CTA rank r handles K-blocks r, r+4, ...
__global__ void __launch_bounds__(256, 1)
tg_debug_kernel(const __grid_constant__ CUtensorMap wmap, const __grid_constant__ CUtensorMap xmap, float* out, int nkb, int npass,
                unsigned* err) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int nstages = 3;
  uint8_t* stages = smem;
  float* part = reinterpret_cast<float*>(smem + nstages * TG_STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(part + TG_M * TG_PART_LD);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * nstages + 1);
  const uint32_t rank = tc::cluster_ctarank();
  const int tile = blockIdx.x / 4;
  if (threadIdx.x == 0) { tc::prefetch_tmap(&wmap); tc::prefetch_tmap(&xmap); }
  tg::Pipe p;
  tg::pipe_init(p, bars, stages, nstages, npass, tmem_slot);
  tc::cluster_sync();
  tg::XSrc xs{&xmap, 0, &xmap, 0, nkb};
  const bool ok = tg::gemm(p, &wmap, tile * TG_M, 0, xs, (int)rank, nkb, 4, 0);
  if (!ok && threadIdx.x == 0) atomicExch(err, 1u);
  tg::accum_to_smem(p, 0, part);
  tc::cluster_sync();
  // CTA `rank` reduces rows [32*rank, 32*rank+32) of the tile over the four partials
  for (int i = threadIdx.x; i < 32 * (TG_N / 2); i += blockDim.x) {
    const int row = 32 * rank + i / (TG_N / 2), col = 2 * (i % (TG_N / 2));
    const float2 s = tg::reduce4(part, row, col);
    *reinterpret_cast<float2*>(out + ((size_t)tile * TG_M + row) * TG_N + col) = s;
  }
  tc::cluster_sync();  // nobody leaves while a peer may still read its shared memory
  if ((threadIdx.x >> 5) == 1) tc::tmem_dealloc<TG_TMEM_COLS>(p.tmem);
}

PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

}  // namespace

int tc_make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const unsigned long long* dims,
                      const unsigned long long* strides_bytes, const unsigned* box) {
  auto enc = get_encode();
  T2_REQUIRE(enc, T2_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gd[5], gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  T2_REQUIRE(r == CUDA_SUCCESS, T2_ERR_CUDA, "cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
  return T2_OK;
}

int tg_split_bf16_launch(const float* src, __nv_bfloat16* dst, int rows, int cols, int ld_src, int ld_dst, long long plane,
                         cudaStream_t st) {
  const long long n = (long long)rows * ld_dst;
  split_bf16_kernel<<<(int)std::min<long long>((n + 255) / 256, 148 * 8), 256, 0, st>>>(src, dst, rows, cols, ld_src, ld_dst, plane);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int tg_debug_gemm(const float* W, const float* x, float* out, int rows, int K, int npass, void* ws, long long ws_bytes, cudaStream_t st) {
  T2_REQUIRE(rows % TG_M == 0 && K % (4 * TG_BK) == 0 && (npass == 1 || npass == 3), T2_ERR_ARG,
             "t2_debug_tcgemm: rows %% 128, K %% 256, npass in {1,3}");
  Arena ar(ws, ws_bytes);
  __nv_bfloat16* Wb = ar.take<__nv_bfloat16>((size_t)2 * rows * K);
  __nv_bfloat16* xb = ar.take<__nv_bfloat16>((size_t)2 * TG_N * K);
  unsigned* err = ar.take<unsigned>(4);
  T2_REQUIRE(!ar.overflow, T2_ERR_WORKSPACE, "t2_debug_tcgemm: workspace too small");
  int rc;
  if ((rc = tg_split_bf16_launch(W, Wb, rows, K, K, K, (long long)rows * K, st))) return rc;
  if ((rc = tg_split_bf16_launch(x, xb, TG_N, K, K, K, (long long)TG_N * K, st))) return rc;
  T2_CHECK_CUDA(cudaMemsetAsync(err, 0, sizeof(unsigned), st));
  CUtensorMap wmap, xmap;
  const unsigned zbox = npass == 3 ? 2u : 1u;
  {
    unsigned long long dims[3] = {(unsigned long long)K, (unsigned long long)rows, 2ull};
    unsigned long long strides[2] = {(unsigned long long)K * 2, (unsigned long long)rows * K * 2};
    unsigned box[3] = {TG_BK, TG_M, zbox};
    if ((rc = tc_make_tmap_bf16(&wmap, Wb, 3, dims, strides, box))) return rc;
  }
  {
    unsigned long long dims[3] = {(unsigned long long)K, (unsigned long long)TG_N, 2ull};
    unsigned long long strides[2] = {(unsigned long long)K * 2, (unsigned long long)TG_N * K * 2};
    unsigned box[3] = {TG_BK, TG_N, zbox};
    if ((rc = tc_make_tmap_bf16(&xmap, xb, 3, dims, strides, box))) return rc;
  }
  const size_t smem = 1024 + 3 * TG_STAGE_BYTES + TG_M * TG_PART_LD * sizeof(float) + 256;
  T2_CHECK_CUDA(cudaFuncSetAttribute(tg_debug_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((rows / TG_M) * 4);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 4; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  T2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, tg_debug_kernel, wmap, xmap, out, K / TG_BK, npass, err));
  unsigned herr = 0;  // debug entry: synchronous
  T2_CHECK_CUDA(cudaMemcpyAsync(&herr, err, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  T2_CHECK_CUDA(cudaStreamSynchronize(st));
  T2_REQUIRE(herr == 0, T2_ERR_DEVICE_ABORT, "t2_debug_tcgemm: pipeline timeout inside the kernel");
  return T2_OK;
}

extern "C" int t2_debug_tcgemm(const float* W, const float* x, float* out, int rows, int K, int npass, void* ws, long long ws_bytes,
                               void* stream) {
  T2_REQUIRE(W && x && out && ws, T2_ERR_ARG, "t2_debug_tcgemm: null argument");
  return tg_debug_gemm(W, x, out, rows, K, npass, ws, ws_bytes, (cudaStream_t)stream);
}
