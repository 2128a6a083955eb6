// Shared device/host helpers for libt2b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

#define T2_OK 0
#define T2_ERR_ARG -1
#define T2_ERR_CUDA -2
#define T2_ERR_STATE -3
#define T2_ERR_WORKSPACE -4
#define T2_ERR_UNSUPPORTED -5
#define T2_ERR_DEVICE_ABORT -6

void t2_set_error(const char* fmt, ...);

#define T2_CHECK_CUDA(expr)                                                                   \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      t2_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));     \
      return T2_ERR_CUDA;                                                                     \
    }                                                                                         \
  } while (0)

#define T2_REQUIRE(cond, code, ...)  \
  do {                               \
    if (!(cond)) {                   \
      t2_set_error(__VA_ARGS__);     \
      return (code);                 \
    }                                \
  } while (0)

static inline int64_t round_up64(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// Bump allocator over the caller-provided workspace.
struct Arena {
  char* base;
  int64_t size, off;
  bool overflow;
  Arena(void* p, int64_t n) : base((char*)p), size(n), off(0), overflow(false) {}
  template <typename T>
  T* take(int64_t count) {
    int64_t bytes = round_up64(count * (int64_t)sizeof(T), 256);
    if (base == nullptr) { off += bytes; return nullptr; }  // sizing pass
    if (off + bytes > size) { overflow = true; return nullptr; }
    T* r = (T*)(base + off);
    off += bytes;
    return r;
  }
};

#ifdef __CUDACC__
__device__ __forceinline__ float ld_cg(const float* p) { return __ldcg(p); }
__device__ __forceinline__ float4 ld_cg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Software grid barrier of the persistent kernels (all CTAs co-resident: one CTA per SM, grid <= #SMs).  One counter: every CTA
// adds 1 with release semantics and polls it with acquire loads until it reaches epoch * nblocks; bit 31 of the same word is the
// abort flag, so leaving the barrier costs no second round trip to L2.
// `fail` = this CTA hit a pipeline timeout: it raises the abort bit before arriving so every CTA leaves together.
// `wrote_x`: this phase stored GEMM operands with ordinary (generic-proxy) stores; they must be ordered before the TMA
// (async-proxy) reads of a later phase, hence the proxy fence on the writer side (the TMA producer thread issues the matching
// fence before its first load).
struct GridBar {
  unsigned* ctr;
  unsigned* abort_flag;   // host-visible copy of the abort state
  unsigned nblocks;
  unsigned epoch;
  __device__ bool sync(bool fail, bool wrote_x) {
    // the GEMM operands the CTA wrote with generic stores are read by TMA (async proxy) after this barrier; they live in global
    // memory, so the fence is restricted to that state space (ncu: the unrestricted form was 4.8% of all stall samples)
    if (wrote_x) asm volatile("fence.proxy.async.global;" ::: "memory");
    fail = __syncthreads_or(fail ? 1 : 0) != 0;   // a pipeline timeout seen by any role thread aborts the grid
    __shared__ int ok_s;
    if (threadIdx.x == 0) {
      ++epoch;
      if (fail) { atomicExch(abort_flag, 1u); atomicOr(ctr, 0x80000000u); }
      // release: cumulative over the CTA's writes that the bar.sync above ordered before this thread
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
      const unsigned target = epoch * nblocks;
      unsigned v;
      long long spins = 0;
      while ((((v = ld_acquire_u32(ctr)) & 0x7fffffffu) < target) && !(v & 0x80000000u)) {
        if (++spins > (1ll << 24)) {   // ~10 s: a peer died; never hang the GPU
          atomicExch(abort_flag, 1u);
          atomicOr(ctr, 0x80000000u);
          v = 0x80000000u;
          break;
        }
      }
      ok_s = (v & 0x80000000u) == 0u;
    }
    __syncthreads();
    return ok_s != 0;
  }
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// exp2-based forms (MUFU.EX2): absolute error ~1e-7..1e-6, far inside the parity tolerances, a fraction of the code
// size and latency of the libm versions (the decoder executes them on the critical path of every frame).
__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) {
  const float t = __expf(-2.0f * fabsf(x));
  return copysignf(__fdividef(1.0f - t, 1.0f + t), x);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Counter-based RNG for in-kernel dropout when no mask is injected (masks == NULL).
// Not bit-compatible with torch's Philox stream (the reference draws from the global
// RNG, model/modules.py:10-12); statistically equivalent Bernoulli(1-p) keep decisions.
__device__ __forceinline__ uint32_t mix32(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + idx * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (uint32_t)(z >> 32);
}
__device__ __forceinline__ bool rng_keep(uint64_t seed, uint64_t idx, float p) {
  return (mix32(seed, idx) >> 8) * (1.0f / 16777216.0f) >= p;
}
#endif
