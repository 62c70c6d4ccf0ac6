// Building blocks shared by the tcgen05 SSD kernels (ssd_fwd_tc.cu / ssd_bwd_tc.cu).
//
// Operand tiles are written to shared memory by the CTA's threads (they have to be: almost every
// SSD operand is scaled by a per-token, per-head decay factor on its way in) in exactly the
// physical pattern a TMA SWIZZLE_128B load would produce, so the same shared-memory matrix
// descriptors as the GEMM apply:
//   a tile is a sequence of 64-column "blocks", each [rows][64 bf16] = rows x 128 bytes, with the
//   16-byte chunk index XOR-ed with (row % 8).
//   * read as a K-major operand  : a block is a 64-wide K slice;  rows = M (or N) index
//   * read as an MN-major operand: a block is a 64-wide MN atom;  rows = K index
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace mb {
namespace ssd {
using namespace ptx;

constexpr int kL = 256;  // chunk length
constexpr int kP = 64;   // head dim
constexpr int kN = 128;  // state dim

struct NoScale {
  __device__ __forceinline__ float operator()(int /*row*/, int /*blk*/) const { return 1.f; }
};

// Cooperative global -> swizzled smem copy of a [rows x 64*NBLK] bf16 tile (row stride ld).
// Rows >= valid_rows are zero-filled.
//   SCALED = false : cp.async (LDGSTS) 16-byte copies — asynchronous, no register staging, any
//                    number in flight; the caller must cp_async_wait_all() (or wait_group) before
//                    the tile is consumed.  Lets the next head's tiles stream in behind the math.
//   SCALED = true  : `scale(row, blk)` multiplies every element of that (row, 64-column block);
//                    loads are issued 8 deep per thread before the first is consumed.
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g, bool pred) {
  const int sz = pred ? 16 : 0;  // src-size 0 => destination is zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int NBLK, bool SCALED, class ScaleFn>
__device__ __forceinline__ void load_tile(uint8_t* sdst, const bf16* __restrict__ g, long long ld, int rows,
                                          int valid_rows, ScaleFn scale, int tid, int nthreads) {
  const int total = rows * NBLK * 8;
  if (!SCALED) {
    const uint32_t sbase = smem_u32(sdst);
    for (int idx = tid; idx < total; idx += nthreads) {
      const int c = idx & 7;
      const int blk = (idx >> 3) % NBLK;
      const int row = idx / (8 * NBLK);
      const bool ok = row < valid_rows;
      const bf16* src = g + (size_t)(ok ? row : 0) * ld + blk * 64 + c * 8;
      cp_async16(sbase + blk * rows * 128 + row * 128 + ((c ^ (row & 7)) << 4), src, ok);
    }
    return;
  }
  constexpr int U = 8;
  for (int base = tid; base < total; base += nthreads * U) {
    bf16x8 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads;
      const int c = idx & 7;
      const int blk = (idx >> 3) % NBLK;
      const int row = idx / (8 * NBLK);
      if (idx < total && row < valid_rows) v[u] = ld8(g + (size_t)row * ld + blk * 64 + c * 8);
      else v[u].v[0] = v[u].v[1] = v[u].v[2] = v[u].v[3] = __floats2bfloat162_rn(0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = base + u * nthreads;
      if (idx >= total) break;
      const int c = idx & 7;
      const int blk = (idx >> 3) % NBLK;
      const int row = idx / (8 * NBLK);
      float f[8];
      unpack8(v[u], f);
      const float sc = row < valid_rows ? scale(row, blk) : 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] *= sc;
      *reinterpret_cast<bf16x8*>(sdst + (size_t)blk * rows * 128 + row * 128 + ((c ^ (row & 7)) << 4)) = pack8(f);
    }
  }
}

// K-major operand: block kb (64 wide), k-step ks (16 wide) of a tile with `rows` rows.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_addr, int rows, int kb, int ks) {
  return make_smem_desc(tile_addr + kb * rows * 128 + ks * 32, 16, 1024);
}
// MN-major operand: k-step k16 (16 K-rows) of a tile with `rows` K-rows; MN atoms are rows*128 apart.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_addr, int rows, int k16) {
  return make_smem_desc(tile_addr + k16 * 2048, rows * 128, 1024);
}

__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
  const uint32_t a = smem_u32(p);
  return p + ((1024u - (a & 1023u)) & 1023u);
}

}  // namespace ssd
}  // namespace mb
