// Attention.forward as a stand-alone call (model/attention.py:34-69): the location-sensitive attention of ONE decoder step from
// explicit tensors.  Inside Tacotron2.forward / Decoder.forward the same arithmetic is fused into the persistent decoder kernel
// (decoder.cu); this entry point exists so that the reference's module boundary is callable on its own (C ABI t2_attention_step).
//   q      = query_layer(h)                                   (:52)
//   loc    = location_dense(location_conv(wcat)^T)            (:54-56)   conv k = 31, pad 15, 2 -> 32 channels, no bias
//   e      = v . tanh(q + loc + processed_memory)             (:58-62)
//   e[mask] = -inf ; w = softmax(e) ; ctx = w @ memory        (:63-67)
// One CTA per utterance, fp32 FMA arithmetic throughout (no tensor-core passes: this is 2.7 MFLOP per utterance).
#include "ctx.cuh"

namespace {

constexpr int AS_THREADS = 256;

__global__ void __launch_bounds__(AS_THREADS)
attention_step_kernel(const float* __restrict__ h, const float* __restrict__ memory, const float* __restrict__ pmem,
                      const float* __restrict__ wcat, const unsigned char* __restrict__ mask, const float* __restrict__ Wq,
                      const float* __restrict__ v, const float* __restrict__ conv_t /*[2][31][32]*/,
                      const float* __restrict__ dense_t /*[32][A]*/, int L, int H, int A, int EF, float* __restrict__ ctx,
                      float* __restrict__ w_out) {
  extern __shared__ float sh[];
  float* q_s = sh;                    // [A]
  float* wc_s = q_s + A;              // [2][L + 30] channels with a 15-sample zero halo
  float* e_s = wc_s + 2 * (L + 30);   // [L]
  float* red = e_s + L;               // [32]
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = AS_THREADS / 32;
  // processed query: one warp per output row, lanes stride the 1024-wide dot product in 16-byte loads
  for (int a = warp; a < A; a += NW) {
    const float4* wr = reinterpret_cast<const float4*>(Wq + (size_t)a * H);
    const float4* hr = reinterpret_cast<const float4*>(h + (size_t)b * H);
    float acc = 0.f;
    for (int k = lane; k < H / 4; k += 32) {
      const float4 w4 = __ldg(wr + k), h4 = __ldg(hr + k);
      acc = fmaf(w4.x, h4.x, acc); acc = fmaf(w4.y, h4.y, acc); acc = fmaf(w4.z, h4.z, acc); acc = fmaf(w4.w, h4.w, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) q_s[a] = acc;
  }
  for (int i = tid; i < 2 * (L + 30); i += AS_THREADS) {
    const int c = i / (L + 30), l = i % (L + 30) - 15;
    wc_s[i] = (l >= 0 && l < L) ? wcat[((size_t)b * 2 + c) * L + l] : 0.f;
  }
  __syncthreads();
  // energies: one warp per position.  lane = conv filter for the location conv, then lane = 4 attention dims for the dense layer
  for (int l = warp; l < L; l += NW) {
    float f = 0.f;
    for (int c = 0; c < 2; ++c) {
      const float* x = wc_s + c * (L + 30) + l;
      const float* wt = conv_t + (size_t)c * 31 * 32 + lane;
#pragma unroll 1
      for (int j = 0; j < 31; ++j) f = fmaf(__ldg(wt + j * 32), x[j], f);
    }
    float e = 0.f;
    for (int it = 0; it < (A + 127) / 128; ++it) {   // uniform trip count: the shuffles need all 32 lanes
      const int a0 = 4 * lane + 128 * it;
      const bool on = a0 < A;
      const int ac = on ? a0 : 0;
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int ff = 0; ff < 32; ++ff) {
        const float fv = __shfl_sync(0xffffffffu, f, ff);
        const float4 d4 = __ldg(reinterpret_cast<const float4*>(dense_t + (size_t)ff * A + ac));
        acc[0] = fmaf(d4.x, fv, acc[0]); acc[1] = fmaf(d4.y, fv, acc[1]); acc[2] = fmaf(d4.z, fv, acc[2]); acc[3] = fmaf(d4.w, fv, acc[3]);
      }
      const float4 p4 = __ldg(reinterpret_cast<const float4*>(pmem + ((size_t)b * L + l) * A + ac));
      float t = v[ac + 0] * tanhf(q_s[ac + 0] + acc[0] + p4.x);
      t = fmaf(v[ac + 1], tanhf(q_s[ac + 1] + acc[1] + p4.y), t);
      t = fmaf(v[ac + 2], tanhf(q_s[ac + 2] + acc[2] + p4.z), t);
      t = fmaf(v[ac + 3], tanhf(q_s[ac + 3] + acc[3] + p4.w), t);
      if (on) e += t;
    }
    e = warp_sum(e);
    if (lane == 0) e_s[l] = mask[(size_t)b * L + l] ? -INFINITY : e;
  }
  __syncthreads();
  float mx = -INFINITY;
  for (int l = tid; l < L; l += AS_THREADS) mx = fmaxf(mx, e_s[l]);
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < NW; ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int l = tid; l < L; l += AS_THREADS) { const float p = expf(e_s[l] - mx); e_s[l] = p; sum += p; }
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = red[0];
  for (int i = 1; i < NW; ++i) sum += red[i];
  const float inv = 1.0f / sum;
  for (int l = tid; l < L; l += AS_THREADS) { const float w = e_s[l] * inv; e_s[l] = w; w_out[(size_t)b * L + l] = w; }
  __syncthreads();
  for (int c = tid; c < EF; c += AS_THREADS) {   // context: thread = column, coalesced rows of `memory`
    const float* m = memory + (size_t)b * L * EF + c;
    float acc = 0.f;
    for (int l = 0; l < L; ++l) acc = fmaf(e_s[l], __ldg(m + (size_t)l * EF), acc);
    ctx[(size_t)b * EF + c] = acc;
  }
}

}  // namespace

extern "C" int t2_attention_step(t2_ctx* c, const float* attention_hidden_state, const float* memory, const float* processed_memory,
                                 const float* attention_weights_cat, const unsigned char* mask, int B, int L, float* context,
                                 float* weights, void* stream) {
  CHECK_CTX("t2_attention_step");
  T2_REQUIRE(attention_hidden_state && memory && processed_memory && attention_weights_cat && mask && context && weights, T2_ERR_ARG,
             "t2_attention_step: null argument");
  T2_REQUIRE(B > 0 && L > 0, T2_ERR_ARG, "t2_attention_step: bad sizes B=%d L=%d", B, L);
  T2_REQUIRE(c->H % 4 == 0 && c->A % 4 == 0, T2_ERR_UNSUPPORTED, "t2_attention_step: att_rnn_dim / att_dim must be multiples of 4");
  const size_t smem = ((size_t)c->A + 2 * (L + 30) + L + 32) * sizeof(float);
  T2_REQUIRE(smem <= 200 * 1024, T2_ERR_UNSUPPORTED, "t2_attention_step: L=%d too long", L);
  if (smem > 48 * 1024) T2_CHECK_CUDA(cudaFuncSetAttribute(attention_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  attention_step_kernel<<<B, AS_THREADS, smem, (cudaStream_t)stream>>>(attention_hidden_state, memory, processed_memory, attention_weights_cat,
                                                                        mask, c->d2w.Wq, c->d2w.v, c->loc_conv_t, c->loc_dense_t, L, c->H, c->A,
                                                                        c->EF, context, weights);
  T2_CHECK_CUDA(cudaGetLastError());
  return T2_OK;
}
