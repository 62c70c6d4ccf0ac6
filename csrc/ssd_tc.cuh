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
// Rows >= valid_rows are zero-filled.  `scale(row, blk)` multiplies every element of that
// (row, 64-column block); pass NoScale for a plain copy.
template <int NBLK, bool SCALED, class ScaleFn>
__device__ __forceinline__ void load_tile(uint8_t* sdst, const bf16* __restrict__ g, long long ld, int rows,
                                          int valid_rows, ScaleFn scale, int tid, int nthreads) {
  const int total = rows * NBLK * 8;
  for (int idx = tid; idx < total; idx += nthreads) {
    const int c = idx & 7;
    const int blk = (idx >> 3) % NBLK;
    const int row = idx / (8 * NBLK);
    bf16x8 v;
    if (row < valid_rows) {
      v = ld8(g + (size_t)row * ld + blk * 64 + c * 8);
      if (SCALED) {
        float f[8];
        unpack8(v, f);
        const float s = scale(row, blk);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] *= s;
        v = pack8(f);
      }
    } else {
      v.v[0] = v.v[1] = v.v[2] = v.v[3] = __floats2bfloat162_rn(0.f, 0.f);
    }
    *reinterpret_cast<bf16x8*>(sdst + (size_t)blk * rows * 128 + row * 128 + ((c ^ (row & 7)) << 4)) = v;
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
