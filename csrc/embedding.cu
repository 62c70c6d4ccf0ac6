// K18: embedding gather (forward) and scatter-add (backward) on the fp32 residual stream.
// Forward writes the fp32 row straight into the residual stream buffer (the first block's
// add+RMSNorm then needs no separate add).  Backward accumulates into the *same* fp32 gradient
// slab that the tied lm_head wgrad wrote (tied grad is complete
// last), using 16-byte vector reductions (red.global.add.v4.f32, sm_90+).
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

__global__ void embedding_fwd_kernel(const long long* __restrict__ ids, const float* __restrict__ w,
                                     float* __restrict__ out, int M, int D, int V) {
  const int warps_per_block = blockDim.x >> 5;
  const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  if (row >= M) return;
  long long id = ids[row];
  if (id < 0 || id >= V) id = 0;
  const float4* src = reinterpret_cast<const float4*>(w + (size_t)id * D);
  float4* dst = reinterpret_cast<float4*>(out + (size_t)row * D);
  for (int i = threadIdx.x & 31; i < (D >> 2); i += 32) dst[i] = __ldg(src + i);
}

__global__ void embedding_bwd_kernel(const long long* __restrict__ ids, const float* __restrict__ dout,
                                     float* __restrict__ dw, int M, int D, int V) {
  const int warps_per_block = blockDim.x >> 5;
  const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  if (row >= M) return;
  long long id = ids[row];
  if (id < 0 || id >= V) return;
  const float4* src = reinterpret_cast<const float4*>(dout + (size_t)row * D);
  float4* dst = reinterpret_cast<float4*>(dw + (size_t)id * D);
  for (int i = threadIdx.x & 31; i < (D >> 2); i += 32) atomicAdd(dst + i, src[i]);
}

// fp32 -> bf16 cast of a flat buffer (used for the bf16 compute copy of parameters when no
// optimizer-maintained shadow exists, and by tests).
__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long n) {
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  const long long stride = (long long)gridDim.x * blockDim.x * 8;
  for (; i + 8 <= n; i += stride) {
    float f[8];
    ld8f(in + i, f);
    st8(out + i, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long k = n & ~7ll; k < n; ++k) out[k] = __float2bfloat16(in[k]);
}

}  // namespace mb
using namespace mb;

extern "C" int mb_embedding_fwd(const long long* ids, const float* w, float* out, int M, int D, int V,
                                cudaStream_t stream) {
  MB_REQUIRE(D % 4 == 0, "embedding_fwd: D=%d must be a multiple of 4", D);
  if (M == 0) return 0;
  embedding_fwd_kernel<<<cdiv(M, 8), 256, 0, stream>>>(ids, w, out, M, D, V);
  return check_launch("embedding_fwd");
}

extern "C" int mb_embedding_bwd(const long long* ids, const float* dout, float* dw, int M, int D, int V,
                                cudaStream_t stream) {
  MB_REQUIRE(D % 4 == 0, "embedding_bwd: D=%d must be a multiple of 4", D);
  if (M == 0) return 0;
  embedding_bwd_kernel<<<cdiv(M, 8), 256, 0, stream>>>(ids, dout, dw, M, D, V);
  return check_launch("embedding_bwd");
}

extern "C" int mb_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t stream) {
  if (n == 0) return 0;
  MB_REQUIRE(((uintptr_t)in % 16 == 0) && ((uintptr_t)out % 16 == 0), "cast_f32_bf16: pointers must be 16-byte aligned");
  long long blocks = (n / 8 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  cast_f32_bf16_kernel<<<(int)blocks, 256, 0, stream>>>(in, (bf16*)out, n);
  return check_launch("cast_f32_bf16");
}
