// Micro-benchmark: cycles per tcgen05.mma (M = 128, K = 16, bf16, operands in shared memory) as a function of N and of
// how many independent accumulator chains the issuing thread interleaves.  One CTA per SM, one issuing warp.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I csrc -I include bench/mma_chain_bench.cu -o /tmp/mma_chain_bench
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace mb::ptx;

template <int N, int CHAINS, bool B_MN>
__global__ void __launch_bounds__(128, 1) k(long long* out, int reps) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 0) {
    const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
    constexpr uint32_t idesc = make_idesc_bf16(128, N, MAJOR_K, B_MN ? MAJOR_MN : MAJOR_K);
    long long t0 = 0, t1 = 0;
    for (int it = 0; it < 2; ++it) {   // first pass warms up
      t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint64_t da = make_smem_desc(a_addr + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(b_addr + kk * 2048, 128 * 128, 1024)
                                     : make_smem_desc(b_addr + (kk >> 2) * (N * 128) + (kk & 3) * 32, 16, 1024);
            mma_bf16_ss(tmem + (kk % CHAINS) * N, da, db, idesc, (r > 0 || kk >= CHAINS) ? 1u : 0u);
          }
        }
        __syncwarp();
      }
      if (elect_one()) mma_commit(&bar);
      __syncwarp();
      mbar_wait_warp(&bar, it & 1);
      tc_fence_after();
      t1 = clock64();
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int N, int CHAINS, bool B_MN>
void run(const char* name) {
  long long* d;
  cudaMalloc(&d, 8);
  const int reps = 64;
  auto kern = k<N, CHAINS, B_MN>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  kern<<<148, 128, 100 * 1024>>>(d, reps);
  long long h = 0;
  cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("%-28s N=%3d chains=%d : %6.1f cycles per MMA (floor %d)  %s\n", name, N, CHAINS, (double)h / (reps * 8), 128 * N / 256,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run<64, 1, false>("K-major B");  run<64, 2, false>("K-major B");  run<64, 4, false>("K-major B");
  run<128, 1, false>("K-major B"); run<128, 2, false>("K-major B");
  run<256, 1, false>("K-major B"); run<256, 2, false>("K-major B");
  run<64, 1, true>("MN-major B");  run<64, 2, true>("MN-major B");
  run<128, 1, true>("MN-major B"); run<128, 2, true>("MN-major B");
  return 0;
}
