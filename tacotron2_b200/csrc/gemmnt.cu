// Bulk NT GEMM on tcgen05 (see gemmnt.cuh).  One CTA per (128 x NT output tile, batch z, K split):
// warp 0 = TMA producer, warp 1 = MMA issuer, all 8 warps = epilogue (TMEM -> registers -> fp32 partial slab).
#include "gemmnt.cuh"

#define GN_STAGES 3
#define GN_A_TILE (128 * 64 * 2)
#define GN_STAGE_BYTES (4 * GN_A_TILE)
#define GN_TMEM_COLS 128

namespace {

__global__ void __launch_bounds__(256, 1) gemm_nt_kernel(const __grid_constant__ GemmNtArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(base + GN_STAGES * GN_STAGE_BYTES);
  uint64_t* empty = full + GN_STAGES;
  uint64_t* accum = empty + GN_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Nt = (a.N + a.NT - 1) / a.NT, Mt = (a.M + 127) >> 7;
  const int mt = blockIdx.x / Nt, nt = blockIdx.x - mt * Nt;
  const int z = blockIdx.y, sp = blockIdx.z;
  const int m0 = mt * 128, n0 = nt * a.NT;
  const int kin_blocks = (a.Kin + 63) >> 6;
  const long long nkb_seg1 = (long long)a.Kout * kin_blocks;
  const long long nkb_all = nkb_seg1 + ((a.Kin2 + 63) >> 6);
  const int kb0 = (int)(nkb_all * sp / a.splitk), kb1 = (int)(nkb_all * (sp + 1) / a.splitk);
  const int brow0 = nt * (a.b_row_step ? a.b_row_step : a.NT);
  const int shift = a.shift0 + z * a.shift_step;
  const uint32_t b_tile = (uint32_t)a.NT * 128u;

  if (tid == 0) {
    tc::prefetch_tmap(&a.a_map);
    tc::prefetch_tmap(&a.b_map);
    for (int i = 0; i < GN_STAGES; ++i) { tc::mbar_init(full + i, 1); tc::mbar_init(empty + i, 1); }
    tc::mbar_init(accum, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<GN_TMEM_COLS>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  bool ok = true;

  if (warp == 0 && lane == 0) {
    asm volatile("fence.proxy.async;" ::: "memory");
    int st = 0; uint32_t ph = 0;
    const bool amn = (a.mn & 1) != 0, bmn = (a.mn & 2) != 0;
    const int kinb_blocks = ((a.KinB ? a.KinB : a.Kin) + 63) >> 6;   // the B operand's own two-level split of the contraction index
    const uint32_t tx = (uint32_t)(a.npass == 3 ? 2 : 1) * (GN_A_TILE + (bmn ? (uint32_t)((a.NT + 63) / 64) * 8192u : b_tile));
    for (int kb = kb0; kb < kb1; ++kb) {
      if (!tc::mbar_wait(empty + st, ph ^ 1)) { ok = false; break; }
      uint8_t* sp_ = base + (size_t)st * GN_STAGE_BYTES;
      tc::mbar_arrive_expect_tx(full + st, tx);
      if (a.mn) {   // an MN-major operand arrives as [64 kin rows][64 channels] tiles, hi then lo plane per 64-channel block
        const int ko = kb / kin_blocks, ki = (kb - ko * kin_blocks) * 64;
        const int kob = kb / kinb_blocks, kib = (kb - kob * kinb_blocks) * 64;
        const uint32_t blk = (a.npass == 3 ? 2u : 1u) * 8192u;
        if (amn) {
          tc::tma_load_4d(sp_, &a.a_map, full + st, m0, ki, ko, 0);
          tc::tma_load_4d(sp_ + blk, &a.a_map, full + st, m0 + 64, ki, ko, 0);
        } else {
          tc::tma_load_4d(sp_, &a.a_map, full + st, ki, ko, m0 + z * a.a_zrow, 0);
        }
        if (bmn) {
          for (int nb = 0; nb * 64 < a.NT; ++nb)
            tc::tma_load_4d(sp_ + 2 * GN_A_TILE + nb * blk, &a.b_map, full + st, n0 + nb * 64, kib + shift, kob, 0);
        } else {
          tc::tma_load_4d(sp_ + 2 * GN_A_TILE, &a.b_map, full + st, kib + shift, kob, brow0 + z * a.b_zrow, 0);
        }
      } else if (kb < nkb_seg1) {
        const int ko = kb / kin_blocks, ki = (kb - ko * kin_blocks) * 64;
        tc::tma_load_4d(sp_, &a.a_map, full + st, ki, ko, m0 + z * a.a_zrow, 0);
        tc::tma_load_4d(sp_ + 2 * GN_A_TILE, &a.b_map, full + st, ki + shift, ko, brow0 + z * a.b_zrow, 0);
      } else {
        const int ki = (int)(kb - nkb_seg1) * 64;
        tc::tma_load_4d(sp_, &a.a_map2, full + st, ki, 0, m0 + z * a.a_zrow, 0);
        tc::tma_load_4d(sp_ + 2 * GN_A_TILE, &a.b_map2, full + st, ki, 0, brow0 + z * a.b_zrow, 0);
      }
      if (++st == GN_STAGES) { st = 0; ph ^= 1; }
    }
  } else if (warp == 1 && lane == 0) {
    int st = 0; uint32_t ph = 0, acc = 0;
    const bool amn = (a.mn & 1) != 0, bmn = (a.mn & 2) != 0;
    const uint32_t idesc = tc::idesc_bf16_f32(128, a.NT) | (amn ? (1u << 15) : 0u) | (bmn ? (1u << 16) : 0u);   // bits 15 / 16: A / B MN-major
    for (int kb = kb0; kb < kb1; ++kb) {
      if (!tc::mbar_wait(full + st, ph)) { ok = false; break; }
      tc::tc_fence_after();
      const uint32_t sb = tc::smem_u32(base + (size_t)st * GN_STAGE_BYTES);
      uint64_t ahi, alo, bhi, blo;
      // MN-major: 64-channel blocks lie blk bytes apart (hi | lo per block), 8 kin rows = one 1024-byte swizzle atom
      const uint32_t blk = (a.npass == 3 ? 2u : 1u) * 8192u;
      if (amn) { ahi = tc::smem_desc_sw128_mn(sb, blk, 1024); alo = tc::smem_desc_sw128_mn(sb + 8192, blk, 1024); }
      else { ahi = tc::smem_desc_sw128(sb); alo = tc::smem_desc_sw128(sb + GN_A_TILE); }
      if (bmn) { bhi = tc::smem_desc_sw128_mn(sb + 2 * GN_A_TILE, blk, 1024); blo = tc::smem_desc_sw128_mn(sb + 2 * GN_A_TILE + 8192, blk, 1024); }
      else { bhi = tc::smem_desc_sw128(sb + 2 * GN_A_TILE); blo = tc::smem_desc_sw128(sb + 2 * GN_A_TILE + b_tile); }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        // 16 kin: two atoms (2048 B) MN-major, 32 B K-major
        const uint64_t adva = amn ? (uint64_t)(k * 128) : (uint64_t)(k * 2), advb = bmn ? (uint64_t)(k * 128) : (uint64_t)(k * 2);
        if (a.npass == 3) {
          tc::umma_bf16(tmem, alo + adva, bhi + advb, idesc, acc);
          tc::umma_bf16(tmem, ahi + adva, blo + advb, idesc, 1u);
          tc::umma_bf16(tmem, ahi + adva, bhi + advb, idesc, 1u);
        } else {
          tc::umma_bf16(tmem, ahi + adva, bhi + advb, idesc, acc);
        }
        acc = 1u;
      }
      tc::umma_commit(empty + st);
      if (++st == GN_STAGES) { st = 0; ph ^= 1; }
    }
    tc::umma_commit(accum);
  }
  __syncwarp();
  if (!tc::mbar_wait(accum, 0)) ok = false;
  tc::tc_fence_after();
  ok = __syncthreads_and(ok ? 1 : 0) != 0;

  // epilogue: thread = output row (TMEM lane); 16-column chunks alternate between the two warp groups
  const int q = warp & 3, row = 32 * q + lane;
  const size_t ldp = (size_t)Nt * a.NT;
  float* dst_row = a.partial + (((size_t)sp * a.nz + z) * Mt * 128 + m0 + row) * ldp + n0;
  const bool empty_k = kb1 <= kb0;   // nothing accumulated: TMEM is undefined, write zeros
#pragma unroll 1
  for (int c = warp >> 2; c < a.NT / 16; c += 2) {
    float v[16];
    tc::tmem_ld16(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(c * 16), v);
    tc::tmem_ld_wait();
    if (a.direct_out) {
      const int m = m0 + row;
      if (m < a.M) {
        float* drow = a.direct_out + (size_t)m * a.direct_ldm;
        const int nb = n0 + c * 16;
        if (nb + 16 <= a.N && (a.direct_ldm & 3) == 0 && (reinterpret_cast<uintptr_t>(a.direct_out) & 15) == 0) {
          // 16-byte stores: the scalar form wrote 4 bytes per 32-byte sector per instruction (the 51200 x 1536 projection gradient took
          // 0.67 ms for 315 MB)
          const bool live = ok && !empty_k;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            reinterpret_cast<float4*>(drow + nb)[i] = live ? make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int n = nb + i;
            if (n < a.N) drow[n] = (ok && !empty_k) ? v[i] : 0.f;
          }
        }
      }
      continue;
    }
    float4* dst = reinterpret_cast<float4*>(dst_row + c * 16);
#pragma unroll
    for (int i = 0; i < 4; ++i)
      dst[i] = (ok && !empty_k) ? make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<GN_TMEM_COLS>(tmem);
}

__global__ void gemm_nt_reduce_kernel(const float* __restrict__ partial, int splitk, int nz, int Mp, int Np, int M, int N, GemmNtOut o) {
  const long long total = (long long)nz * M * N;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i % N);
    const int m = (int)((i / N) % M);
    const int z = (int)(i / ((long long)N * M));
    float s = 0.f;
    for (int sp = 0; sp < splitk; ++sp) s += partial[(((size_t)sp * nz + z) * Mp + m) * Np + n];
    float* d = o.out + (size_t)m * o.ldm + (size_t)n * o.ldn + (size_t)z * o.ldz;
    *d = o.accumulate ? *d + o.alpha * s : o.alpha * s;
  }
}

// fp32 [nb][R][C] -> bf16 (hi,lo) [2][C][dst_ld] with batch bb at column offset bb*dst_boff (transposed); 32x32 tiles
__global__ void transpose_split_kernel(const float* __restrict__ src, long long ld, long long src_bstride,
                                       __nv_bfloat16* __restrict__ dst, int R, int C, int R_ld, long long dst_ld, long long dst_boff,
                                       long long plane, int row_shift) {
  __shared__ float tile[32][33];
  const int bb = blockIdx.z;
  const float* s = src + (size_t)bb * src_bstride;
  __nv_bfloat16* d = dst + (size_t)bb * dst_boff;
  const int c = blockIdx.x * 32 + threadIdx.x;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = blockIdx.y * 32 + i + row_shift;
    tile[i][threadIdx.x] = (r >= 0 && r < R && c < C) ? s[(size_t)r * ld + c] : 0.f;
  }
  __syncthreads();
  const int r2 = blockIdx.y * 32 + threadIdx.x;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c2 = blockIdx.x * 32 + i;
    if (r2 < R_ld && c2 < C) {
      __nv_bfloat16 hi, lo;
      tc::split_bf16(tile[threadIdx.x][i], hi, lo);
      d[(size_t)c2 * dst_ld + r2] = hi;
      d[plane + (size_t)c2 * dst_ld + r2] = lo;
    }
  }
}

__global__ void transpose_bf16_kernel(const __nv_bfloat16* __restrict__ src, long long src_ld, long long src_bstride, long long src_plane,
                                      __nv_bfloat16* __restrict__ dst, int R, int C, int R_ld, long long dst_ld, long long dst_boff,
                                      int row_shift, long long plane) {
  __shared__ __nv_bfloat16 tile[32][34];
  const int pl = blockIdx.z & 1, bb = blockIdx.z >> 1;
  const __nv_bfloat16* s = src + (size_t)pl * src_plane + (size_t)bb * src_bstride;
  __nv_bfloat16* d = dst + (size_t)pl * plane + (size_t)bb * dst_boff;
  const int c = blockIdx.x * 32 + threadIdx.x;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = blockIdx.y * 32 + i + row_shift;   // dst row r' holds src row r' + row_shift (zero outside [0,R))
    tile[i][threadIdx.x] = (r >= 0 && r < R && c < C) ? s[(size_t)r * src_ld + c] : __float2bfloat16(0.f);
  }
  __syncthreads();
  const int r2 = blockIdx.y * 32 + threadIdx.x;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c2 = blockIdx.x * 32 + i;
    if (r2 < R_ld && c2 < C) d[(size_t)c2 * dst_ld + r2] = tile[threadIdx.x][i];
  }
}

// the same, four columns per thread (unit column stride, N / strides multiples of 4, 16-byte aligned destination)
__global__ void gemm_nt_reduce4_kernel(const float* __restrict__ partial, int splitk, int nz, int Mp, int Np, int M, int N, GemmNtOut o) {
  const int N4 = N >> 2;
  const long long total = (long long)nz * M * N4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long mz = i / N4;
    const int n = (int)(i - mz * N4) * 4;
    const int m = (int)(mz % M), z = (int)(mz / M);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int sp = 0; sp < splitk; ++sp) {
      const float4 v = *reinterpret_cast<const float4*>(partial + (((size_t)sp * nz + z) * Mp + m) * Np + n);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    float4* d = reinterpret_cast<float4*>(o.out + (size_t)m * o.ldm + n + (size_t)z * o.ldz);
    if (o.accumulate) { const float4 p = *d; *d = make_float4(p.x + o.alpha * s.x, p.y + o.alpha * s.y, p.z + o.alpha * s.z, p.w + o.alpha * s.w); }
    else *d = make_float4(o.alpha * s.x, o.alpha * s.y, o.alpha * s.z, o.alpha * s.w);
  }
}

// 64x64 tiles, 4-byte (bf16x2) global accesses on both sides: 128 B per warp row instead of 64 B.  Needs even C, src_ld, dst_ld,
// dst_boff, R_ld and 4-byte aligned bases (the launch falls back to the 32x32 kernel otherwise).
__global__ void __launch_bounds__(256) transpose_bf16x2_kernel(const __nv_bfloat16* __restrict__ src, long long src_ld, long long src_bstride,
                                                               long long src_plane, __nv_bfloat16* __restrict__ dst, int R, int C, int R_ld,
                                                               long long dst_ld, long long dst_boff, int row_shift, long long plane) {
  __shared__ __align__(4) __nv_bfloat16 tile[64][66];
  const int pl = blockIdx.z & 1, bb = blockIdx.z >> 1;
  const __nv_bfloat16* s = src + (size_t)pl * src_plane + (size_t)bb * src_bstride;
  __nv_bfloat16* d = dst + (size_t)pl * plane + (size_t)bb * dst_boff;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 64 + 2 * tx;
  const __nv_bfloat162 zero2 = __floats2bfloat162_rn(0.f, 0.f);
#pragma unroll
  for (int i = ty; i < 64; i += 8) {
    const int r = blockIdx.y * 64 + i + row_shift;   // dst row r' holds src row r' + row_shift (zero outside [0,R))
    *reinterpret_cast<__nv_bfloat162*>(&tile[i][2 * tx]) =
        (r >= 0 && r < R && c < C) ? *reinterpret_cast<const __nv_bfloat162*>(s + (size_t)r * src_ld + c) : zero2;
  }
  __syncthreads();
  const int r2 = blockIdx.y * 64 + 2 * tx;
#pragma unroll
  for (int i = ty; i < 64; i += 8) {
    const int c2 = blockIdx.x * 64 + i;
    if (r2 < R_ld && c2 < C)
      *reinterpret_cast<__nv_bfloat162*>(d + (size_t)c2 * dst_ld + r2) = __halves2bfloat162(tile[2 * tx][i], tile[2 * tx + 1][i]);
  }
}

}  // namespace

long long gemmnt_partial_floats(GemmNtArgs& a, int num_sms) {
  const int Nt = (a.N + a.NT - 1) / a.NT, Mt = (a.M + 127) / 128;
  const long long nkb = (long long)a.Kout * ((a.Kin + 63) / 64) + (a.Kin2 + 63) / 64;
  if (a.splitk <= 0) {
    // K split: fill whole waves of one-CTA-per-SM launches (448 tiles on 148 SMs are 3.03 waves -- a quarter of the time on four CTAs),
    // at least 8 K-blocks per CTA, slabs within the callers' reserved buffer (24 M floats); a small penalty per split for the slab traffic
    const long long tiles = (long long)Mt * Nt * a.nz, per_split = (long long)a.nz * Mt * 128 * Nt * a.NT;
    long long best = 1;
    double best_score = -1.0;
    for (long long sk = 1; sk <= 64; ++sk) {
      if (sk > 1 && (nkb / sk < 8 || sk * per_split > (24ll << 20))) break;
      const long long ctas = tiles * sk, waves = (ctas + num_sms - 1) / num_sms;
      const double score = (double)ctas / (double)(waves * num_sms) - 0.004 * (double)sk;
      if (score > best_score + 1e-9) { best_score = score; best = sk; }
    }
    a.splitk = (int)best;
  }
  return (long long)a.splitk * a.nz * Mt * 128 * Nt * a.NT;
}

int gemmnt_launch(const GemmNtArgs& a, const GemmNtOut& o, cudaStream_t st) {
  T2_REQUIRE(a.NT % 16 == 0 && a.NT >= 16 && a.NT <= 128 && a.splitk >= 1 && a.nz >= 1 && (a.partial || a.direct_out), T2_ERR_ARG,
             "gemmnt: bad arguments NT=%d splitk=%d nz=%d", a.NT, a.splitk, a.nz);
  T2_REQUIRE(!a.direct_out || (a.splitk == 1 && a.nz == 1), T2_ERR_ARG, "gemmnt: direct output needs splitk == 1 and nz == 1");
  T2_REQUIRE(!a.mn || (a.Kin2 == 0 && a.b_row_step == 0 && (!(a.mn & 1) || a.a_zrow == 0) && (!(a.mn & 2) || a.b_zrow == 0)), T2_ERR_ARG,
             "gemmnt: an MN-major operand takes one segment and no row offsets");
  T2_REQUIRE(a.mn >= 0 && a.mn <= 3 && (a.KinB == 0 || a.mn), T2_ERR_ARG, "gemmnt: mn = %d / KinB = %d", a.mn, a.KinB);
  if (a.M <= 0 || a.N <= 0) return T2_OK;
  const int Nt = (a.N + a.NT - 1) / a.NT, Mt = (a.M + 127) / 128;
  const size_t smem = 1024 + (size_t)GN_STAGES * GN_STAGE_BYTES + 128;
  static bool attr_set = false;
  if (!attr_set) {
    T2_CHECK_CUDA(cudaFuncSetAttribute(gemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  dim3 grid((unsigned)(Mt * Nt), (unsigned)a.nz, (unsigned)a.splitk);
  gemm_nt_kernel<<<grid, 256, smem, st>>>(a);
  T2_CHECK_CUDA(cudaGetLastError());
  if (a.direct_out) return T2_OK;
  const long long total = (long long)a.nz * a.M * a.N;
  if (o.ldn == 1 && a.N % 4 == 0 && o.ldm % 4 == 0 && o.ldz % 4 == 0 && (reinterpret_cast<uintptr_t>(o.out) & 15) == 0) {
    const int g4 = (int)std::min<long long>((total / 4 + 255) / 256, 148 * 8);
    gemm_nt_reduce4_kernel<<<g4, 256, 0, st>>>(a.partial, a.splitk, a.nz, Mt * 128, Nt * a.NT, a.M, a.N, o);
    T2_CHECK_CUDA(cudaGetLastError());
    return T2_OK;
  }
  const int g = (int)std::min<long long>((total + 255) / 256, 148 * 8);
  gemm_nt_reduce_kernel<<<g, 256, 0, st>>>(a.partial, a.splitk, a.nz, Mt * 128, Nt * a.NT, a.M, a.N, o);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int gemmnt_make_map(CUtensorMap* m, const __nv_bfloat16* base, int Kin, int kin_ld, int Kout, int rows, long long plane_elems,
                    int box_rows, int npass) {
  T2_REQUIRE(kin_ld % 8 == 0 && ((uintptr_t)base & 15) == 0, T2_ERR_ARG, "gemmnt: operand stride %d / base not 16-byte aligned", kin_ld);
  unsigned long long dims[4] = {(unsigned long long)Kin, (unsigned long long)Kout, (unsigned long long)rows, 2ull};
  unsigned long long st[3] = {(unsigned long long)kin_ld * 2, (unsigned long long)Kout * kin_ld * 2, (unsigned long long)plane_elems * 2};
  unsigned box[4] = {64u, 1u, (unsigned)box_rows, npass == 3 ? 2u : 1u};
  return tc_make_tmap_bf16(m, base, 4, dims, st, box);
}

int gemmnt_make_map_cl_strided(CUtensorMap* m, const __nv_bfloat16* base, int Kin, int Kout, int rows, long long kin_elems, long long kout_elems,
                               long long plane_elems, int npass) {
  T2_REQUIRE(kin_elems % 8 == 0 && kout_elems % 8 == 0 && plane_elems % 8 == 0 && ((uintptr_t)base & 15) == 0, T2_ERR_ARG,
             "gemmnt: MN-major operand strides %lld / %lld / %lld or base not 16-byte aligned", kin_elems, kout_elems, plane_elems);
  unsigned long long dims[4] = {(unsigned long long)rows, (unsigned long long)Kin, (unsigned long long)Kout, 2ull};
  unsigned long long st[3] = {(unsigned long long)kin_elems * 2, (unsigned long long)kout_elems * 2, (unsigned long long)plane_elems * 2};
  unsigned box[4] = {64u, 64u, 1u, npass == 3 ? 2u : 1u};
  return tc_make_tmap_bf16(m, base, 4, dims, st, box);
}

int gemmnt_make_map_cl(CUtensorMap* m, const __nv_bfloat16* base, int Kin, int Kout, int rows, long long plane_elems, int npass) {
  T2_REQUIRE(rows % 8 == 0 && ((uintptr_t)base & 15) == 0, T2_ERR_ARG, "gemmnt: channels-last operand with %d channels / base not 16-byte aligned", rows);
  unsigned long long dims[4] = {(unsigned long long)rows, (unsigned long long)Kin, (unsigned long long)Kout, 2ull};
  unsigned long long st[3] = {(unsigned long long)rows * 2, (unsigned long long)Kin * rows * 2, (unsigned long long)plane_elems * 2};
  unsigned box[4] = {64u, 64u, 1u, npass == 3 ? 2u : 1u};
  return tc_make_tmap_bf16(m, base, 4, dims, st, box);
}

int transpose_split_launch(const float* src, long long ld, long long src_bstride, __nv_bfloat16* dst, int R, int C, int R_ld,
                           long long dst_ld, long long dst_boff, int nbatch, long long dst_plane, cudaStream_t st, int row_shift) {
  if (R_ld <= 0 || C <= 0 || nbatch <= 0) return T2_OK;
  dim3 grid((C + 31) / 32, (R_ld + 31) / 32, nbatch), block(32, 8);
  transpose_split_kernel<<<grid, block, 0, st>>>(src, ld, src_bstride, dst, R, C, R_ld, dst_ld, dst_boff, dst_plane, row_shift);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}

int transpose_bf16_launch(const __nv_bfloat16* src, long long src_ld, long long src_bstride, long long src_plane, __nv_bfloat16* dst,
                          int R, int C, int R_ld, long long dst_ld, long long dst_boff, int nbatch, int row_shift, long long dst_plane,
                          cudaStream_t st) {
  if (R_ld <= 0 || C <= 0 || nbatch <= 0) return T2_OK;
  const bool even = ((C | R_ld | src_ld | dst_ld | dst_boff | src_bstride | src_plane | dst_plane) & 1) == 0 &&
                    (((uintptr_t)src | (uintptr_t)dst) & 3) == 0;
  if (even) {
    dim3 grid2((C + 63) / 64, (R_ld + 63) / 64, 2 * nbatch), block2(32, 8);
    transpose_bf16x2_kernel<<<grid2, block2, 0, st>>>(src, src_ld, src_bstride, src_plane, dst, R, C, R_ld, dst_ld, dst_boff, row_shift,
                                                      dst_plane);
    T2_CHECK_CUDA(cudaGetLastError());
    return T2_OK;
  }
  dim3 grid((C + 31) / 32, (R_ld + 31) / 32, 2 * nbatch), block(32, 8);
  transpose_bf16_kernel<<<grid, block, 0, st>>>(src, src_ld, src_bstride, src_plane, dst, R, C, R_ld, dst_ld, dst_boff, row_shift, dst_plane);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
