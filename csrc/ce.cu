// K17 (part): cross-entropy over one slab of logits, forward and gradient in a single pass pair.
// The lm_head GEMM (csrc/gemm.cu) writes an [R, V] slab — bf16 in training (the precision the reference's
// logits have under autocast, /root/reference/train.py:211; halves the slab traffic and lets the gradient
// overwrite the logits in place), fp32 for loss-only scoring — and this kernel turns it into
//   loss[r]      = logsumexp(logits[r, :]) - logits[r, target[r]]
//   dlogits[r,:] = (softmax(logits[r, :]) - onehot(target[r])) * scale        (bf16, in place of a
//                  separate softmax / scatter / scale / cast sequence of ATen kernels)
// which then feeds the dgrad and wgrad GEMMs.  Replaces ATen's log_softmax + nll_loss forward and
// backward as reached from F.cross_entropy at /root/reference/model.py:43-46.
// One CTA per row; the row (V = 50304 floats = 196 KB) is read twice (second read hits L2).
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kCeThreads = 512;

// exp(m - nm) with the empty-partial case (m = -inf, possibly nm = -inf too) mapped to 0, not NaN
__device__ __forceinline__ float scale_to(float m, float nm) { return m == -INFINITY ? 0.f : fast_exp(m - nm); }

// 8 consecutive logits of a row as fp32, from either slab precision
template <bool BF16_IN>
__device__ __forceinline__ void ld_logits8(const void* row, int i8, float* f) {
  if (BF16_IN) unpack8(ld8(reinterpret_cast<const bf16*>(row) + 8 * i8), f);
  else ld8f(reinterpret_cast<const float*>(row) + 8 * i8, f);
}
template <bool BF16_IN>
__device__ __forceinline__ float ld_logit(const void* row, long long i) {
  return BF16_IN ? __bfloat162float(reinterpret_cast<const bf16*>(row)[i]) : reinterpret_cast<const float*>(row)[i];
}

// dlogits may alias logits (bf16 slab): element i is read and then written by the same thread.
template <bool BF16_IN>
__global__ void __launch_bounds__(kCeThreads)
ce_slab_kernel(const void* logits, long long ld, const long long* __restrict__ targets,
               float* __restrict__ loss, bf16* dlogits, long long ldd, int V, float scale) {
  __shared__ float s_m[kCeThreads / 32], s_s[kCeThreads / 32];
  __shared__ float s_max, s_lse;
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  const int row = blockIdx.x, tid = threadIdx.x;
  const void* lr = reinterpret_cast<const char*>(logits) + (size_t)row * ld * (BF16_IN ? 2 : 4);
  const int nv8 = V >> 3;
  // pass 1: online max / sum-exp
  float m = -INFINITY, s = 0.f;
  for (int i = tid; i < nv8; i += kCeThreads) {
    float v[8];
    ld_logits8<BF16_IN>(lr, i, v);
    float vm = v[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) vm = fmaxf(vm, v[j]);
    if (vm > m) { s *= fast_exp(m - vm); m = vm; }
#pragma unroll
    for (int j = 0; j < 8; ++j) s += fast_exp(v[j] - m);
  }
  for (int i = (nv8 << 3) + tid; i < V; i += kCeThreads) {
    const float v = ld_logit<BF16_IN>(lr, i);
    if (v > m) { s *= fast_exp(m - v); m = v; }
    s += fast_exp(v - m);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    const float nm = fmaxf(m, om);
    s = s * scale_to(m, nm) + os * scale_to(om, nm);
    m = nm;
  }
  if ((tid & 31) == 0) { s_m[tid >> 5] = m; s_s[tid >> 5] = s; }
  __syncthreads();
  if (tid == 0) {
    float M = s_m[0], S = s_s[0];
    for (int w = 1; w < kCeThreads / 32; ++w) {
      const float nm = fmaxf(M, s_m[w]);
      S = S * scale_to(M, nm) + s_s[w] * scale_to(s_m[w], nm);
      M = nm;
    }
    const float lse = M + __logf(S);
    s_max = M;
    s_lse = lse;
    const long long t = targets[row];
    loss[row] = (t >= 0 && t < V) ? lse - ld_logit<BF16_IN>(lr, t) : 0.f;
  }
  __syncthreads();
  if (dlogits == nullptr) return;
  // pass 2: gradient
  const float lse = s_lse;
  const long long t = targets[row];
  bf16* dr = dlogits + (size_t)row * ldd;
  for (int i = tid; i < nv8; i += kCeThreads) {
    float f[8];
    ld_logits8<BF16_IN>(lr, i, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float p = fast_exp(f[j] - lse);
      if (8 * i + j == t) p -= 1.f;
      f[j] = p * scale;
    }
    st8(dr + 8 * i, pack8(f));
  }
  for (int i = (nv8 << 3) + tid; i < V; i += kCeThreads) {
    float p = fast_exp(ld_logit<BF16_IN>(lr, i) - lse);
    if (i == t) p -= 1.f;
    dr[i] = __float2bfloat16(p * scale);
  }
}

}  // namespace mb
using namespace mb;

// logits: [R, V] slab, fp32 (logits_bf16 = 0) or bf16 (1); dlogits_bf16 may be the bf16 logits slab itself (in place)
extern "C" int mb_ce_slab(const void* logits, int logits_bf16, long long ld, const long long* targets, float* loss,
                          void* dlogits_bf16, long long ldd, int R, int V, float scale, cudaStream_t stream) {
  MB_REQUIRE(ld % 8 == 0 && (dlogits_bf16 == nullptr || ldd % 8 == 0), "ce_slab: row strides must keep 16-byte alignment");
  MB_REQUIRE(dlogits_bf16 != logits || logits_bf16, "ce_slab: the gradient can only overwrite a bf16 logits slab");
  if (R == 0) return 0;
  if (logits_bf16)
    MB_CUDA(launch_pdl(ce_slab_kernel<true>, dim3(R), dim3(kCeThreads), 0, stream, logits, ld, targets, loss, (bf16*)dlogits_bf16, ldd, V, scale));
  else
    MB_CUDA(launch_pdl(ce_slab_kernel<false>, dim3(R), dim3(kCeThreads), 0, stream, logits, ld, targets, loss, (bf16*)dlogits_bf16, ldd, V, scale));
  return check_launch("ce_slab");
}
