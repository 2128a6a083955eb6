// tcgen05 implicit-GEMM conv1d (see convtc.cuh).  Persistent: one CTA per SM walks the (batch item, 128-position tile, N tile) list;
// warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..9 = epilogue (TMEM -> registers -> BN/activation -> global).  Two accumulators in
// tensor memory: the epilogue of tile i runs while the producer and the issuer are already on tile i + 1 (the one-tile-per-CTA form
// left the tensor pipe idle during every prologue and epilogue: 44 % active on the postnet layers, profiles/r02_v1_tc_kernels_summary.txt).
#include "convtc.cuh"
#include "layers.cuh"

#define CT_STAGES 3
#define CT_A_TILE (128 * 64 * 2)
#define CT_STAGE_BYTES (4 * CT_A_TILE)  // A hi|lo (32 KB) + B hi|lo (up to 32 KB)
#define CT_TMEM_COLS 256                // two accumulators of up to 128 columns
#define CT_THREADS 320

namespace {

__global__ void __launch_bounds__(CT_THREADS, 1) conv_tc_kernel(const __grid_constant__ ConvTcArgs a, int ntiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(base + CT_STAGES * CT_STAGE_BYTES);
  uint64_t* empty = full + CT_STAGES;
  uint64_t* acc_full = empty + CT_STAGES;    // [2]
  uint64_t* acc_empty = acc_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  volatile int* fail_s = reinterpret_cast<volatile int*>(tmem_slot + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_t = (a.T + 127) >> 7, tiles_bt = a.B * tiles_t;
  const int dil = a.dil > 1 ? a.dil : 1;
  const int cblocks = (a.Cin + 63) >> 6, nkb = a.k * cblocks, pad = ((a.k - 1) >> 1) * dil;
  const uint32_t b_tile = (uint32_t)a.NT * 128u;  // bytes of one bf16 weight tile (NT rows x 128 B)

  if (tid == 0) {
    tc::prefetch_tmap(&a.in_map);
    tc::prefetch_tmap(&a.w_map);
    for (int i = 0; i < CT_STAGES; ++i) { tc::mbar_init(full + i, 1); tc::mbar_init(empty + i, 1); }
    for (int i = 0; i < 2; ++i) { tc::mbar_init(acc_full + i, 1); tc::mbar_init(acc_empty + i, 8); }
    *fail_s = 0;
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<CT_TMEM_COLS>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // tile id -> (position tile, N tile): consecutive ids share the activation tile
  const int ntn = a.Cout / a.NT;
  if (warp == 0) {
    if (lane == 0) {
      asm volatile("fence.proxy.async;" ::: "memory");
      int st = 0; uint32_t ph = 0;
      const uint32_t tx = (uint32_t)(a.npass == 3 ? 2 : 1) * (CT_A_TILE + b_tile);
      for (int tile = blockIdx.x; tile < ntiles && !*fail_s; tile += gridDim.x) {
        const int bt = tile / ntn, n0 = (tile - bt * ntn) * a.NT;
        const int b = bt / tiles_t, t0 = (bt - b * tiles_t) * 128;
        for (int kb = 0; kb < nkb; ++kb) {
          const int j = kb / cblocks, c0 = (kb - j * cblocks) * 64;
          if (!tc::mbar_wait(empty + st, ph ^ 1)) { *fail_s = 1; break; }
          uint8_t* sp = base + (size_t)st * CT_STAGE_BYTES;
          tc::mbar_arrive_expect_tx(full + st, tx);
          tc::tma_load_4d(sp, &a.in_map, full + st, c0, t0 + j * dil - pad, b, 0);    // zero fill outside [0,T) = conv padding
          tc::tma_load_4d(sp + 2 * CT_A_TILE, &a.w_map, full + st, c0, j, n0, 0);
          if (++st == CT_STAGES) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int st = 0; uint32_t ph = 0;
      const uint32_t idesc = tc::idesc_bf16_f32(128, a.NT);
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles && !*fail_s; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        if (!tc::mbar_wait(acc_empty + buf, (uint32_t)((it >> 1) & 1) ^ 1u)) { *fail_s = 1; break; }   // the epilogue has drained this accumulator
        tc::tc_fence_after();
        const uint32_t dcol = tmem + (uint32_t)buf * 128u;
        uint32_t acc = 0;
        for (int kb = 0; kb < nkb; ++kb) {
          if (!tc::mbar_wait(full + st, ph)) { *fail_s = 1; break; }
          tc::tc_fence_after();
          const uint32_t sb = tc::smem_u32(base + (size_t)st * CT_STAGE_BYTES);
          const uint64_t ahi = tc::smem_desc_sw128(sb), alo = tc::smem_desc_sw128(sb + CT_A_TILE);
          const uint64_t bhi = tc::smem_desc_sw128(sb + 2 * CT_A_TILE), blo = tc::smem_desc_sw128(sb + 2 * CT_A_TILE + b_tile);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adv = (uint64_t)(k * 2);
            if (a.npass == 3) {
              tc::umma_bf16(dcol, alo + adv, bhi + adv, idesc, acc);
              tc::umma_bf16(dcol, ahi + adv, blo + adv, idesc, 1u);
              tc::umma_bf16(dcol, ahi + adv, bhi + adv, idesc, 1u);
            } else {
              tc::umma_bf16(dcol, ahi + adv, bhi + adv, idesc, acc);
            }
            acc = 1u;
          }
          tc::umma_commit(empty + st);
          if (++st == CT_STAGES) { st = 0; ph ^= 1; }
        }
        tc::umma_commit(acc_full + buf);
      }
    }
  } else {
    // ---- epilogue warps 2..9: thread = output position (TMEM lane quarter = warp % 4), 16-channel chunks alternate between the two groups.
    // What a chunk adds from global memory (residual, running sum, dropout mask: first-touch DRAM reads in the narrow HiFi-GAN stages)
    // is requested one chunk AHEAD -- the next chunk of this tile or the first of the next tile -- so its latency runs under the
    // accumulator wait, the tensor-memory read and the stores of the current chunk (it used to sit between them: 5.8 us per tile).
    const int q = warp & 3, grp = (warp - 2) >> 2, row = 32 * q + lane;
    const long long plane = (long long)a.B * a.T * a.Cout;
    const int nch = a.NT / 16;
    struct Pre { float4 r[4], ac[4]; uint4 kp; };
    const bool keep16 = (a.keep_ld & 15) == 0 && (reinterpret_cast<uintptr_t>(a.keep) & 15) == 0;
    auto load_pre = [&](int tile, int c, Pre& p) {
      const int bt = tile / ntn, n = (tile - bt * ntn) * a.NT + c * 16;
      const int b = bt / tiles_t, t = (bt - b * tiles_t) * 128 + row;
      if (t >= a.T) return;
      const size_t pos = (size_t)b * a.T + t;
      if (a.residual) {
        const float4* src = reinterpret_cast<const float4*>(a.residual + (size_t)b * a.res_bstride + (size_t)t * a.Cout + n);
#pragma unroll
        for (int i = 0; i < 4; ++i) p.r[i] = src[i];
      }
      if (a.accum) {
        const float4* src = reinterpret_cast<const float4*>(a.accum + pos * a.Cout + n);
#pragma unroll
        for (int i = 0; i < 4; ++i) p.ac[i] = src[i];   // (may alias out_f32: each element is read, then written, by this thread only)
      }
      if (a.keep) {
        const unsigned char* kp = a.keep + pos * a.keep_ld + n;
        if (keep16) p.kp = __ldg(reinterpret_cast<const uint4*>(kp));
        else {
          unsigned char* dstb = reinterpret_cast<unsigned char*>(&p.kp);
#pragma unroll
          for (int i = 0; i < 16; ++i) dstb[i] = __ldg(kp + i);
        }
      }
    };
    Pre cur, nxt;
    if ((int)blockIdx.x < ntiles && grp < nch) load_pre(blockIdx.x, grp, cur);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      const int bt = tile / ntn, n0 = (tile - bt * ntn) * a.NT;
      const int b = bt / tiles_t, t0 = (bt - b * tiles_t) * 128;
      const bool got = tc::mbar_wait(acc_full + buf, (uint32_t)((it >> 1) & 1));
      if (!got) *fail_s = 1;
      tc::tc_fence_after();
      const int t = t0 + row;
      const bool valid = got && t < a.T && !*fail_s;
      const size_t pos = (size_t)b * a.T + min(t, a.T - 1);
#pragma unroll 1
      for (int c = grp; c < nch; c += 2) {
        {
          int ntile = tile, nc = c + 2;
          if (nc >= nch) { ntile = tile + gridDim.x; nc = grp; }
          if (ntile < ntiles) load_pre(ntile, nc, nxt);
        }
        float v[16];
        tc::tmem_ld16(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * 128 + c * 16), v);
        tc::tmem_ld_wait();
        if (valid) {
          const int n = n0 + c * 16;
          const float* rr = reinterpret_cast<const float*>(cur.r);
          const float* aa = reinterpret_cast<const float*>(cur.ac);
          const unsigned char* kk = reinterpret_cast<const unsigned char*>(&cur.kp);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x = v[i];
            const int ci = a.chan_mod > 0 ? (n + i) % a.chan_mod : n + i;
            if (a.scale) x = fmaf(x, __ldg(a.scale + ci), __ldg(a.shift + ci));
            else if (a.shift) x += __ldg(a.shift + ci);
            if (a.act == ACT_RELU) x = fmaxf(x, 0.f);
            else if (a.act == ACT_TANH) x = tanhf(x);
            else if (a.act == ACT_GELU) x = 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
            if (a.keep) x = kk[i] ? x * a.drop_scale : 0.f;
            if (a.residual) x += rr[i];
            if (a.out_scale != 0.f) x *= a.out_scale;
            if (a.accum) x += aa[i];
            v[i] = x;
          }
          if (a.out_c0 && n == 0) a.out_c0[pos] = v[0];
          if (a.out_f32) {
            float4* dst = reinterpret_cast<float4*>(a.out_f32 + pos * a.Cout + n);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          }
          if (a.out_b16) {
            __align__(16) __nv_bfloat16 hi[16], lo[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) tc::split_bf16((a.b16_slope != 0.f && v[i] < 0.f) ? v[i] * a.b16_slope : v[i], hi[i], lo[i]);
            uint4* dh = reinterpret_cast<uint4*>(a.out_b16 + pos * a.Cout + n);
            uint4* dl = reinterpret_cast<uint4*>(a.out_b16 + plane + pos * a.Cout + n);
            dh[0] = reinterpret_cast<const uint4*>(hi)[0]; dh[1] = reinterpret_cast<const uint4*>(hi)[1];
            dl[0] = reinterpret_cast<const uint4*>(lo)[0]; dl[1] = reinterpret_cast<const uint4*>(lo)[1];
          }
        }
        cur = nxt;
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(acc_empty + buf);   // this warp has read its lanes of the accumulator
      if (!got) break;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<CT_TMEM_COLS>(tmem);
}

__global__ void split_cl_kernel(const float* src, long long src_bstride, __nv_bfloat16* dst, int B, int T, int C) {
  const long long n = (long long)B * T * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int b = (int)(p / T), t = (int)(p % T);
    __nv_bfloat16 hi, lo;
    tc::split_bf16(src[(size_t)b * src_bstride + (size_t)t * C + c], hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}

__global__ void pack_conv_w_kernel(const float* W, __nv_bfloat16* dst, int Cout, int Cin, int k) {
  const long long n = (long long)Cout * Cin * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int j = (int)((i / Cin) % k);
    const int co = (int)(i / ((long long)Cin * k));
    __nv_bfloat16 hi, lo;
    tc::split_bf16(W[((size_t)co * Cin + ci) * k + j], hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}

__global__ void embed_split_kernel(const long long* idx, const float* table, __nv_bfloat16* dst, int n_pos, int E, int n_rows) {
  const long long n = (long long)n_pos * E;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = idx[i / E];
    const float v = (r >= 0 && r < n_rows) ? table[(size_t)r * E + i % E] : 0.f;
    __nv_bfloat16 hi, lo;
    tc::split_bf16(v, hi, lo);
    dst[i] = hi;
    dst[n + i] = lo;
  }
}

inline int gs(long long n) { long long g = (n + 255) / 256; return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g)); }

}  // namespace

int convtc_launch(const ConvTcArgs& a, cudaStream_t st) {
  T2_REQUIRE(a.NT % 16 == 0 && a.NT >= 16 && a.NT <= 128 && a.Cout % a.NT == 0 && (a.k & 1) && a.Cin % 8 == 0 && a.Cout % 8 == 0,
             T2_ERR_UNSUPPORTED, "convtc: unsupported shape Cin=%d Cout=%d k=%d NT=%d", a.Cin, a.Cout, a.k, a.NT);
  if (a.B * a.T == 0) return T2_OK;
  const size_t smem = 1024 + (size_t)CT_STAGES * CT_STAGE_BYTES + 128;
  static bool attr_set = false;
  static int num_sms = 0;
  if (!attr_set) {
    T2_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0;
    T2_CHECK_CUDA(cudaGetDevice(&dev));
    T2_CHECK_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    attr_set = true;
  }
  const int ntiles = a.B * ((a.T + 127) / 128) * (a.Cout / a.NT);
  conv_tc_kernel<<<std::min(ntiles, num_sms), CT_THREADS, smem, st>>>(a, ntiles);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int split_cl_launch(const float* src, long long src_bstride, __nv_bfloat16* dst, int B, int T, int C, cudaStream_t st) {
  if (B * T == 0) return T2_OK;
  split_cl_kernel<<<gs((long long)B * T * C), 256, 0, st>>>(src, src_bstride, dst, B, T, C);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int pack_conv_w_launch(const float* W, __nv_bfloat16* dst, int Cout, int Cin, int k, cudaStream_t st) {
  pack_conv_w_kernel<<<gs((long long)Cout * Cin * k), 256, 0, st>>>(W, dst, Cout, Cin, k);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int embed_split_launch(const long long* idx, const float* table, __nv_bfloat16* dst, int n_pos, int E, int n_rows, cudaStream_t st) {
  embed_split_kernel<<<gs((long long)n_pos * E), 256, 0, st>>>(idx, table, dst, n_pos, E, n_rows);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
