// Batch assembly on the GPU: the reference's `_collate` (datasets/tts_dataloader.py:8-35) pads every
// data key of the batch with nn.utils.rnn.pad_sequence(batch_first=True) on DataLoader worker CPUs; the gate targets come from
// datasets/tts_dataset.py:211-214 (ones, last valid frame 0).  Here the items of a field arrive concatenated (ragged rows + row
// offsets) and ONE launch writes all padded fields and the gate targets.  Pure byte movement: HBM-bound, every output byte is
// written once, every valid input byte read once, 16-byte accesses where the row size allows.
#include "../../include/t2b200.h"
#include "common.cuh"

#define CL_MAX_FIELDS 8
#define CL_THREADS 256

namespace {

struct CollateField {
  const unsigned char* src;   // items' rows, concatenated
  const long long* off;       // [B+1] row offsets of the items
  unsigned char* out;         // [B][max_rows][row_bytes]
  long long row_units;        // row size in copy units
  long long max_rows;
  long long units;            // B * max_rows * row_units
  int unit_bytes;             // 16 or 4
};

struct CollateArgs {
  CollateField f[CL_MAX_FIELDS];
  int n_fields, B;
  const long long* gate_len;  // [B] or null
  float* gate;                // [B][gate_T]
  long long gate_T;
};

template <class V>
__device__ __forceinline__ void pad_field(const CollateField& fd, long long i0, long long stride) {
  const long long per_item = fd.max_rows * fd.row_units;
  for (long long i = i0; i < fd.units; i += stride) {
    const long long b = i / per_item, r = (i - b * per_item) / fd.row_units, c = i - b * per_item - r * fd.row_units;
    const long long o0 = __ldg(fd.off + b), n = __ldg(fd.off + b + 1) - o0;
    V v{};
    if (r < n) v = __ldg(reinterpret_cast<const V*>(fd.src) + (o0 + r) * fd.row_units + c);
    reinterpret_cast<V*>(fd.out)[i] = v;
  }
}

__global__ void __launch_bounds__(CL_THREADS) collate_kernel(const __grid_constant__ CollateArgs a) {
  const long long i0 = blockIdx.x * (long long)CL_THREADS + threadIdx.x, stride = (long long)gridDim.x * CL_THREADS;
  for (int k = 0; k < a.n_fields; ++k) {
    if (a.f[k].unit_bytes == 16) pad_field<uint4>(a.f[k], i0, stride);
    else pad_field<unsigned>(a.f[k], i0, stride);
  }
  if (a.gate) {  // ones, 0 at the last valid frame (datasets/tts_dataset.py:213-214), 0 beyond (pad_sequence)
    for (long long i = i0; i < a.B * a.gate_T; i += stride) {
      const long long b = i / a.gate_T, t = i - b * a.gate_T;
      a.gate[i] = t < __ldg(a.gate_len + b) - 1 ? 1.0f : 0.0f;
    }
  }
}

}  // namespace

extern "C" int t2_collate(int n_fields, const void* const* src, const long long* const* off, const long long* row_bytes,
                          const long long* max_rows, void* const* out, int B, const long long* gate_len, float* gate,
                          long long gate_T, void* stream) {
  T2_REQUIRE(n_fields >= 0 && n_fields <= CL_MAX_FIELDS && B >= 1, T2_ERR_ARG, "t2_collate: n_fields %d (max %d), B %d", n_fields,
             CL_MAX_FIELDS, B);
  T2_REQUIRE(n_fields == 0 || (src && off && row_bytes && max_rows && out), T2_ERR_ARG, "t2_collate: null table");
  T2_REQUIRE((gate == nullptr) == (gate_len == nullptr) && gate_T >= 0, T2_ERR_ARG, "t2_collate: gate and gate_len come together");
  CollateArgs a{};
  a.n_fields = n_fields;
  a.B = B;
  a.gate_len = gate_len;
  a.gate = gate;
  a.gate_T = gate_T;
  long long work = gate ? (long long)B * gate_T : 0;
  for (int k = 0; k < n_fields; ++k) {
    T2_REQUIRE(off[k] && out[k] && row_bytes[k] > 0 && row_bytes[k] % 4 == 0 && max_rows[k] >= 0, T2_ERR_ARG,
               "t2_collate: field %d (row_bytes %lld must be a positive multiple of 4)", k, row_bytes[k]);
    T2_REQUIRE(src[k] || max_rows[k] == 0, T2_ERR_ARG, "t2_collate: field %d has no source", k);
    const bool v16 = row_bytes[k] % 16 == 0 && ((uintptr_t)src[k] % 16 == 0) && ((uintptr_t)out[k] % 16 == 0);
    CollateField& f = a.f[k];
    f.src = static_cast<const unsigned char*>(src[k]);
    f.off = off[k];
    f.out = static_cast<unsigned char*>(out[k]);
    f.unit_bytes = v16 ? 16 : 4;
    f.row_units = row_bytes[k] / f.unit_bytes;
    f.max_rows = max_rows[k];
    f.units = (long long)B * f.max_rows * f.row_units;
    work = std::max(work, f.units);
  }
  if (work == 0) return T2_OK;
  const int grid = (int)std::min<long long>((work + CL_THREADS - 1) / CL_THREADS, 148 * 8);   // a multiple of the SM count when large
  collate_kernel<<<grid, CL_THREADS, 0, (cudaStream_t)stream>>>(a);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
