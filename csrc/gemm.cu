// K16: bf16 GEMM on tcgen05 / TMEM / TMA for sm_100a.
//
// Replaces the cuBLAS GEMMs behind nn.Linear (in_proj / out_proj / x_proj / dt_proj / lm_head;
// SURVEY.md section 2.3 row K16) in all three layouts autograd needs:
//   forward  y  = x  W^T   : A K-major [M,K],   B K-major  [N,K]
//   dgrad    dx = dy W     : A K-major [M,N'],  B MN-major [N',K]   (W read as stored)
//   wgrad    dW += dy^T x  : A MN-major [M',N], B MN-major [M',K]   (activations read as stored)
// so no operand is ever transposed in memory.  wgrad accumulates in fp32 straight into the
// gradient arena (vector reductions), which is also what makes split-K free.
//
// Structure (one CTA per SM, persistent over output tiles, 6 warps):
//   warp 0    TMA producer   cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx-count
//   warp 1    MMA issuer     tcgen05.mma.kind::f16, fp32 accum in TMEM: 256 x BN x 16 across a CTA pair
//                            (cta_group::2, the default for BN = 256) or 128 x BN x 16 per CTA (cta_group::1);
//                            tcgen05.commit releases smem stages and publishes the accumulator
//   warps 2-5 epilogue       tcgen05.ld (32 lanes x 32 cols) -> registers -> global
// Two TMEM accumulators (2 x BN columns) let the epilogue of tile i overlap the mainloop of
// tile i+1.  Out-of-bounds rows / columns / K are zero-filled by TMA and predicated on store,
// so ragged shapes (N = 3352, K = 48, ...) need no padding copies.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "ptx.cuh"
#include "../include/mamba_b200.h"

namespace mb {
using namespace ptx;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kGemmThreads = 192;

// Tile rasterisation.  With few N tiles (dgrad: N = d_model) consecutive CTAs take all N tiles of
// one M block, so the large A operand (activations / gradients, > L2) is streamed from HBM once;
// with many N tiles consecutive CTAs share one N tile (the weight tile) and A (which then fits in
// L2) is re-read from L2.
__device__ __forceinline__ void decode_tile(int tile, int tiles_m, int tiles_n, int& m_blk, int& n_blk, int& split) {
  const int per_split = tiles_m * tiles_n;
  split = tile / per_split;
  const int t = tile - split * per_split;
  if (tiles_n <= 8) { m_blk = t / tiles_n; n_blk = t - m_blk * tiles_n; }
  else { n_blk = t / tiles_m; m_blk = t - n_blk * tiles_m; }
}

enum { EPI_BF16 = 0, EPI_F32 = 1, EPI_F32_ATOMIC = 2 };

// Columns the MMAs of N tile `n_blk` have to cover: the tile width, or the remaining columns rounded up to 32 (each
// CTA of a pair then holds a multiple of 16 B rows, i.e. whole 8-row swizzle groups).
__device__ __forceinline__ int tile_n_eff(int Ng, int n_blk, int BN) {
  const int rem = Ng - n_blk * BN;
  return rem >= BN ? BN : ((rem + 31) & ~31);
}

enum { MODE_SINGLE = 0, MODE_PAIR = 1 };

template <int BN, int MODE = MODE_SINGLE>
struct GemmCfg {
  // MODE_PAIR: each CTA of the pair keeps only its half of the B tile, which buys two more stages
  static constexpr int STAGES = MODE == MODE_PAIR ? 6 : (BN == 256 ? 4 : 6);
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = (MODE == MODE_PAIR ? BN / 2 : BN) * BK * 2;
  static constexpr int STAGING = 4 * 2 * 4096;  // 4 epilogue warps x 2 buffers x [32 rows][128 B]
  static constexpr int SMEM = STAGES * (A_BYTES + B_BYTES) + STAGING + 1024 /*align slack*/ + 256 /*barriers*/;
};

// MODE_PAIR: launched as clusters of 2 CTAs that work on two vertically adjacent output tiles (same N block)
// and form one `cta_group::2` MMA: a 256 x BN tile whose A rows and TMEM
// accumulator rows are split across the two SMs and whose B tile is split in halves of BN/2 rows, one per
// SM.  Only the even CTA issues tcgen05.mma / tcgen05.commit; both run a TMA producer (whose transaction
// bytes are all counted on the even CTA's full barrier) and an epilogue over their own TMEM.  Per k-block
// each SM now reads 16 KB (A) + 16 KB (half of B) from L2 AND from its own shared memory, instead of
// 16 + 32 KB: the tensor pipe was waiting on the shared-memory operand feed (ncu: 64 % active).
template <int BN, uint32_t A_MAJOR, uint32_t B_MAJOR, int EPI, bool TMA_STORE, int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmD, void* __restrict__ Dout, const float* __restrict__ bias, int Mg, int Ng, int Kg,
                 long long ldd, int splitk, int kb_per_split, float alpha) {
  using Cfg = GemmCfg<BN, MODE>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr bool PAIR = MODE == MODE_PAIR;   // cluster of 2 CTAs, cta_group::2 MMA
  constexpr int CL = PAIR ? 2 : 1;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint8_t* sStage = smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sStage + Cfg::STAGING);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m_real = (Mg + BM - 1) / BM, tiles_n = (Ng + BN - 1) / BN;
  // with PAIR the scheduler hands out pairs of M blocks to a cluster; CTA `crank` takes block 2*pm + crank
  const int tiles_m = (tiles_m_real + CL - 1) / CL;
  const int num_tiles = tiles_m * tiles_n * splitk;
  const int kblocks = (Kg + BK - 1) / BK;
  const int crank = PAIR ? (int)cluster_ctarank() : 0;
  const int worker = (int)blockIdx.x / CL;
  const int nworkers = (int)gridDim.x / CL;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmA);
    prefetch_tensormap(&tmB);
    if (TMA_STORE) prefetch_tensormap(&tmD);
    // pair: the issuing CTA's commit arrives once on the empty barrier of both CTAs
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    // pair: the issuing CTA waits for the epilogue warps of both CTAs before it reuses an accumulator
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], PAIR ? 8 : 4); }
    fence_mbar_init();
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc2(tmem_slot, 2 * BN); tmem_relinquish2(); }
    else { tmem_alloc(tmem_slot, 2 * BN); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel's tail (common.cuh: PDL)
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    // ================================ TMA producer =====================================
    // all 32 lanes run the loops and the waits; one elected lane issues the TMA loads of a stage (see
    // "warp-uniform single-lane issue" in ptx.cuh)
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = worker; tile < num_tiles; tile += nworkers) {
        int m_blk, n_blk, split;
        decode_tile(tile, tiles_m, tiles_n, m_blk, n_blk, split);
        if (PAIR) m_blk = CL * m_blk + crank;
        const int kb0 = split * kb_per_split;
        const int kb1 = min(kb0 + kb_per_split, kblocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait_warp(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = sA + stage * Cfg::A_BYTES;
          uint8_t* b_dst = sB + stage * Cfg::B_BYTES;
          if (elect_one()) {
            if (PAIR) {
              // both CTAs' loads complete on the even CTA's barrier, which expects the bytes of both
              if (crank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * (Cfg::A_BYTES + Cfg::B_BYTES));
              if (A_MAJOR == MAJOR_K) {
                tma_load_2d_pair(a_dst, &tmA, &full_bar[stage], kb * BK, m_blk * BM);
              } else {
#pragma unroll
                for (int j = 0; j < BM / 64; ++j)
                  tma_load_2d_pair(a_dst + j * (BK * 128), &tmA, &full_bar[stage], m_blk * BM + j * 64, kb * BK);
              }
              if (B_MAJOR == MAJOR_K) {  // tmB here is the half-height (BN/2 rows) box
                // the MMA of a narrow last N tile (see n_eff in the MMA issuer) takes its second half of the B rows
                // from the odd CTA, so that CTA fetches from n0 + n_eff / 2 instead of n0 + BN / 2
                tma_load_2d_pair(b_dst, &tmB, &full_bar[stage], kb * BK, n_blk * BN + crank * (tile_n_eff(Ng, n_blk, BN) / 2));
              } else {
#pragma unroll
                for (int j = 0; j < BN / 128; ++j)
                  tma_load_2d_pair(b_dst + j * (BK * 128), &tmB, &full_bar[stage],
                                   n_blk * BN + (crank * (BN / 128) + j) * 64, kb * BK);
              }
            } else {
              mbar_arrive_expect_tx(&full_bar[stage], Cfg::A_BYTES + Cfg::B_BYTES);
              if (A_MAJOR == MAJOR_K) {
                tma_load_2d(a_dst, &tmA, &full_bar[stage], kb * BK, m_blk * BM);
              } else {
#pragma unroll
                for (int j = 0; j < BM / 64; ++j)
                  tma_load_2d(a_dst + j * (BK * 128), &tmA, &full_bar[stage], m_blk * BM + j * 64, kb * BK);
              }
              if (B_MAJOR == MAJOR_K) {
                tma_load_2d(b_dst, &tmB, &full_bar[stage], kb * BK, n_blk * BN);
              } else {
#pragma unroll
                for (int j = 0; j < BN / 64; ++j)
                  tma_load_2d(b_dst + j * (BK * 128), &tmB, &full_bar[stage], n_blk * BN + j * 64, kb * BK);
              }
            }
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ========================================
    if (!PAIR || crank == 0) {   // whole warp; the MMAs / commits of a k-block are issued by one elected lane
      constexpr uint32_t idesc_full = make_idesc_bf16(PAIR ? 2 * BM : BM, BN, A_MAJOR, B_MAJOR);
      // K-major: +32 B per UMMA_K inside the swizzle row; MN-major: +16 k-rows * 128 B
      constexpr uint32_t a_kstep = A_MAJOR == MAJOR_K ? 32 : 16 * 128;
      constexpr uint32_t b_kstep = B_MAJOR == MAJOR_K ? 32 : 16 * 128;
      constexpr uint32_t a_lbo = A_MAJOR == MAJOR_K ? 16 : BK * 128;  // K-major: field value 1 (ignored by HW)
      constexpr uint32_t b_lbo = B_MAJOR == MAJOR_K ? 16 : BK * 128;
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = worker; tile < num_tiles; tile += nworkers, ++local) {
        const int split = tile / (tiles_m * tiles_n);
        const int kb0 = split * kb_per_split;
        const int kb1 = min(kb0 + kb_per_split, kblocks);
        const int acc = local & 1;
        const uint32_t acc_phase = (local >> 1) & 1;
        // A ragged last N tile (in_proj: 3352 = 13 x 256 + 24) runs its MMAs at N = 32 instead of 256: the MMA time
        // is proportional to N and the rest of the tile would only multiply zero-filled rows.
        uint32_t idesc = idesc_full;
        if (PAIR && B_MAJOR == MAJOR_K) {
          int m_unused, n_blk_t, s_unused;
          decode_tile(tile, tiles_m, tiles_n, m_unused, n_blk_t, s_unused);
          idesc = make_idesc_bf16(2 * BM, (uint32_t)tile_n_eff(Ng, n_blk_t, BN), A_MAJOR, B_MAJOR);
        }
#ifdef MB_GEMM_TIMING
        const long long tt0 = clock64();
#endif
        mbar_wait_warp(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
#ifdef MB_GEMM_TIMING
        const long long tt1 = clock64();
        long long twait_full = 0;
#endif
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
#ifdef MB_GEMM_TIMING
          const long long tf0 = clock64();
#endif
          mbar_wait_warp(&full_bar[stage], phase);
          tc_fence_after();
#ifdef MB_GEMM_TIMING
          twait_full += clock64() - tf0;
#endif
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::A_BYTES);
          const uint32_t b_addr = smem_u32(sB + stage * Cfg::B_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = make_smem_desc(a_addr + k * a_kstep, a_lbo, 1024);
              const uint64_t db = make_smem_desc(b_addr + k * b_kstep, b_lbo, 1024);
              if (PAIR) mma_bf16_ss_pair(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
              else mma_bf16_ss(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
            // frees this smem stage when the MMAs above retire (with PAIR in both CTAs: the MMA read both stages)
            if (PAIR) mma_commit_pair_mc(&empty_bar[stage], 0b11);
            else mma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        // accumulator complete -> epilogue (of both CTAs with PAIR)
        if (elect_one()) { if (PAIR) mma_commit_pair_mc(&tfull_bar[acc], 0b11); else mma_commit(&tfull_bar[acc]); }
        __syncwarp();
#ifdef MB_GEMM_TIMING
        if (blockIdx.x == 4 && local < 30 && lane == 0)
          printf("G tile %d start %lld wait_acc %lld wait_full %lld issue_total %lld\n", local, tt0, tt1 - tt0, twait_full, clock64() - tt0);
#endif
      }
    }
  } else {
    // ================================ epilogue ==========================================
    const int lane_grp = warp & 3;  // TMEM lanes [32*lane_grp, +32) are the ones this warp may read
    int local = 0;
    int st_iter = 0;
    for (int tile = worker; tile < num_tiles; tile += nworkers, ++local) {
      int m_blk, n_blk, split_unused;
      decode_tile(tile, tiles_m, tiles_n, m_blk, n_blk, split_unused);
      if (PAIR) m_blk = CL * m_blk + crank;
      const int acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const int row = m_blk * BM + lane_grp * 32 + lane;
      const int n0 = n_blk * BN;
      const uint32_t t_addr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + acc * BN;
      if (TMA_STORE) {
        // registers -> 128B-swizzled staging tile [32 rows][128 B] -> one TMA store per tile slab.
        // Global writes are full 128-byte row segments; ragged M / N edges are clipped by TMA.
        uint8_t* my_stage = sStage + (warp - 2) * 8192;
        constexpr int COLS = EPI == EPI_BF16 ? 64 : 32;  // columns per 128-byte staged row
#pragma unroll 1
        for (int c = 0; c < BN / COLS; ++c) {
          const int col0 = n0 + c * COLS;
          if (col0 >= Ng) break;  // warp-uniform
          uint8_t* buf = my_stage + (st_iter & 1) * 4096;
          if (st_iter >= 2) {  // the store issued two slabs ago must have finished reading this buffer
            if (elect_one()) tma_store_wait_read<1>();
            __syncwarp();
          }
#pragma unroll
          for (int hseg = 0; hseg < COLS / 32; ++hseg) {
            uint32_t r[32];
            tmem_ld32(t_addr + c * COLS + hseg * 32, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * alpha;
            if (bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (col0 + hseg * 32 + j < Ng) v[j] += bias[col0 + hseg * 32 + j];
            }
            if (EPI == EPI_BF16) {
#pragma unroll
              for (int q = 0; q < 4; ++q)
                *reinterpret_cast<bf16x8*>(buf + lane * 128 + (((hseg * 4 + q) ^ (lane & 7)) << 4)) = pack8(v + 8 * q);
            } else {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                *reinterpret_cast<float4*>(buf + lane * 128 + ((q ^ (lane & 7)) << 4)) =
                    make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {   // always the same lane (elect.sync is deterministic): bulk groups are per thread
            tma_store_2d(&tmD, buf, col0, m_blk * BM + lane_grp * 32);
            tma_store_commit();
          }
          ++st_iter;
        }
      } else {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= Ng) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld32(t_addr + c * 32, r);
        tmem_ld_wait();
        if (row < Mg) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * alpha;
          if (bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j) if (col0 + j < Ng) v[j] += bias[col0 + j];
          }
          if (EPI == EPI_BF16) {
            bf16* dst = reinterpret_cast<bf16*>(Dout) + (size_t)row * ldd + col0;
            if (col0 + 32 <= Ng && (ldd & 7) == 0) {
#pragma unroll
              for (int j = 0; j < 4; ++j) st8(dst + j * 8, pack8(v + j * 8));
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (col0 + j < Ng) dst[j] = __float2bfloat16(v[j]);
            }
          } else if (EPI == EPI_F32) {
            float* dst = reinterpret_cast<float*>(Dout) + (size_t)row * ldd + col0;
            if (col0 + 32 <= Ng && (ldd & 3) == 0) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(dst + j * 4) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (col0 + j < Ng) dst[j] = v[j];
            }
          } else {
            float* dst = reinterpret_cast<float*>(Dout) + (size_t)row * ldd + col0;
            if (col0 + 32 <= Ng && (ldd & 3) == 0) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                atomicAdd(reinterpret_cast<float4*>(dst + j * 4),
                          make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (col0 + j < Ng) atomicAdd(dst + j, v[j]);
            }
          }
        }
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR && crank != 0) mbar_arrive_cluster(&tempty_bar[acc], 0);  // the issuing (even) CTA owns the wait
        else mbar_arrive(&tempty_bar[acc]);
      }
    }
    if (TMA_STORE && elect_one()) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // no CTA leaves while its peer may still signal / read from it
  if (warp == 1) { if (PAIR) tmem_dealloc2(tmem_base, 2 * BN); else tmem_dealloc(tmem_base, 2 * BN); }
}

// ---------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  // The encode call is a driver-API entry point: it needs the primary context bound to the calling thread.
  // Runtime calls bind it lazily, but an autograd worker thread may reach this before making any.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { cudaFree(nullptr); ctx_bound = true; }
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// ---- tensor-map cache -----------------------------------------------------------------------------------------
// A CUtensorMap is a pure function of (address, extents, strides, box, element type, swizzle): encoding the same
// key twice gives the same 128 bytes, and nothing about it depends on what the memory currently holds or on who
// owns the allocation.  Training reuses the same arenas and (under the caching allocator / graph pools) the same
// activation addresses every step, so every launcher here asks this small per-thread table first: a GEMM call
// drops from four driver encode calls to four hash probes.  Direct-mapped, 4096 entries; a collision just re-encodes.
struct TmapKey {
  const void* ptr;
  unsigned long long d0, d1, d2, s0, s1;
  unsigned int b0, b1, b2, kind;   // kind: element type | rank << 4 | swizzle << 8
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && s0 == o.s0 && s1 == o.s1 && b0 == o.b0 && b1 == o.b1 &&
           b2 == o.b2 && kind == o.kind;
  }
};
struct TmapSlot { TmapKey key; CUtensorMap map; bool used; };
constexpr int kTmapSlots = 4096;
static thread_local TmapSlot* g_tmap_cache = nullptr;

static int encode_cached(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* ptr, const cuuint64_t* dims,
                         const cuuint64_t* strides, const cuuint32_t* box, CUtensorMapSwizzle swz) {
  TmapKey k = {};
  k.ptr = ptr;
  k.d0 = dims[0]; k.d1 = dims[1]; k.d2 = rank > 2 ? dims[2] : 0;
  k.s0 = strides[0]; k.s1 = rank > 2 ? strides[1] : 0;
  k.b0 = box[0]; k.b1 = box[1]; k.b2 = rank > 2 ? box[2] : 0;
  k.kind = (unsigned)dt | ((unsigned)rank << 4) | ((unsigned)swz << 8);
  unsigned long long h = (unsigned long long)(uintptr_t)ptr * 0x9E3779B97F4A7C15ull;
  h ^= (k.d0 * 0x100000001B3ull) ^ (k.d1 << 17) ^ (k.s0 << 7) ^ (k.s1 << 29) ^ ((unsigned long long)k.b1 << 40) ^ ((unsigned long long)k.b0 << 48) ^ k.kind;
  h ^= h >> 29;
  if (g_tmap_cache == nullptr) g_tmap_cache = static_cast<TmapSlot*>(calloc(kTmapSlots, sizeof(TmapSlot)));
  TmapSlot* slot = g_tmap_cache != nullptr ? &g_tmap_cache[h & (kTmapSlots - 1)] : nullptr;
  if (slot != nullptr && slot->used && slot->key == k) { *out = slot->map; return 0; }
  EncodeTiledFn enc = get_encode_fn();
  MB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, dt, (cuuint32_t)rank, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (rank %d, dims %llu x %llu x %llu, stride %llu, box %u x %u)",
             (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1], rank > 2 ? (unsigned long long)dims[2] : 0ull,
             (unsigned long long)strides[0], box[0], box[1]);
  if (slot != nullptr) { slot->key = k; slot->map = *out; slot->used = true; }
  return 0;
}

// 2-D fp32 row-major output map with a {32, 32} box (128-byte rows), 128B swizzle.
int make_tmap_f32_2d(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld) {
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {32, 32};
  return encode_cached(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ptr, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

// 2-D bf16 row-major [rows, cols] (row stride ld elements) with a {64, box_rows} box, 128B swizzle.
int make_tmap_bf16_2d(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows) {
  MB_REQUIRE(((uintptr_t)ptr & 15) == 0 && (ld * 2) % 16 == 0,
             "TMA operand must be 16-byte aligned with a row stride that is a multiple of 8 elements (ptr=%p ld=%lld)", ptr, ld);
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  return encode_cached(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

// 3-D bf16 [d2][d1][cols] view of a row-major matrix whose rows are (d2, d1) pairs (row stride ld elements),
// box {box_cols, box_rows, 1}; box_cols 64 -> 128B swizzle, 32 -> 64B swizzle.  Rows past d1 inside a box read
// as zeros and are dropped on stores (ragged last chunk of a sequence).
int make_tmap_bf16_3d(CUtensorMap* map, const void* ptr, long long d2, long long d1, long long cols, long long ld,
                      int box_rows, int box_cols) {
  MB_REQUIRE(((uintptr_t)ptr & 15) == 0 && (ld * 2) % 16 == 0,
             "TMA operand must be 16-byte aligned with a row stride that is a multiple of 8 elements (ptr=%p ld=%lld)", ptr, ld);
  MB_REQUIRE(box_cols == 64 || box_cols == 32, "make_tmap_bf16_3d: box_cols must be 64 or 32");
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)d1, (cuuint64_t)d2};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)d1 * (cuuint64_t)ld * 2};
  cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
  return encode_cached(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, ptr, dims, strides, box,
                       box_cols == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
}

template <int BN, uint32_t AM, uint32_t BMJ, int EPI, bool TS, int MODE>
static int launch_gemm_mode(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, void* D,
                            const float* bias, int M, int N, int K, long long ldd, int splitk, int kb_per_split,
                            float alpha, cudaStream_t stream) {
  constexpr int CL = MODE == MODE_PAIR ? 2 : 1;
  auto kern = gemm_bf16_kernel<BN, AM, BMJ, EPI, TS, MODE>;
  static bool attr_set = false;
  if (!attr_set) {
    MB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<BN, MODE>::SMEM));
    attr_set = true;
  }
  const int tiles_m = (cdiv(M, BM) + CL - 1) / CL;
  const int tiles = tiles_m * cdiv(N, BN) * splitk;       // groups of CL vertically adjacent tiles
  const int grid = CL * (tiles < num_sms() / CL ? tiles : num_sms() / CL);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = GemmCfg<BN, MODE>::SMEM;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_for(stream) ? 2 : 1;
  MB_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, td, D, bias, M, N, K, ldd, splitk, kb_per_split, alpha));
  return check_launch("gemm_bf16");
}

// No TMA load in this file is a multicast load: the operands may be any device-accessible pointer the caller
// passes through the C ABI, and TMA multicast loads here read only this GPU's memory, since on B200 a TMA
// multicast load from another GPU's memory hung the GPU.  CTA pairs share the B tile through the cta_group::2
// MMA (each CTA loads its own half) instead.
static bool use_pair() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("MAMBA_B200_GEMM_PAIR"); v = (e == nullptr || e[0] != '0') ? 1 : 0; }  // default on
  return v == 1;
}

template <int BN, uint32_t AM, uint32_t BMJ, int EPI, bool TS>
static int launch_gemm_ts(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tb_half, const CUtensorMap& td,
                          void* D, const float* bias, int M, int N, int K, long long ldd, int splitk, int kb_per_split,
                          float alpha, cudaStream_t stream) {
  if (BN == 256 && use_pair() && cdiv(M, BM) >= 2)
    return launch_gemm_mode<256, AM, BMJ, EPI, TS, MODE_PAIR>(ta, tb_half, td, D, bias, M, N, K, ldd, splitk, kb_per_split, alpha, stream);
  return launch_gemm_mode<BN, AM, BMJ, EPI, TS, MODE_SINGLE>(ta, tb, td, D, bias, M, N, K, ldd, splitk, kb_per_split, alpha, stream);
}

template <int BN, uint32_t AM, uint32_t BMJ, int EPI>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tb_half, void* D,
                       const float* bias, int M, int N, int K, long long ldd, int splitk, int kb_per_split,
                       float alpha, cudaStream_t stream) {
  // stores go through TMA whenever the output satisfies its alignment rules; the additive
  // (wgrad) epilogue uses vector reductions straight from registers
  const bool ts = EPI != EPI_F32_ATOMIC && (EPI == EPI_BF16 ? (ldd % 8 == 0) : (ldd % 4 == 0));
  CUtensorMap td;
  if (ts) {
    if (EPI == EPI_BF16) { if (make_tmap_bf16_2d(&td, D, M, N, ldd, 32)) return 1; }
    else { if (make_tmap_f32_2d(&td, D, M, N, ldd)) return 1; }
    return launch_gemm_ts<BN, AM, BMJ, EPI, (EPI != EPI_F32_ATOMIC)>(ta, tb, tb_half, td, D, bias, M, N, K, ldd, splitk, kb_per_split, alpha, stream);
  }
  td = ta;
  return launch_gemm_ts<BN, AM, BMJ, EPI, false>(ta, tb, tb_half, td, D, bias, M, N, K, ldd, splitk, kb_per_split, alpha, stream);
}

template <int BN, int EPI>
static int dispatch_major(int a_major, int b_major, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tbh,
                          void* D, const float* bias, int M, int N, int K, long long ldd, int splitk, int kbps,
                          float alpha, cudaStream_t s) {
  if (a_major == 0 && b_major == 0) return launch_gemm<BN, MAJOR_K, MAJOR_K, EPI>(ta, tb, tbh, D, bias, M, N, K, ldd, splitk, kbps, alpha, s);
  if (a_major == 0 && b_major == 1) return launch_gemm<BN, MAJOR_K, MAJOR_MN, EPI>(ta, tb, tbh, D, bias, M, N, K, ldd, splitk, kbps, alpha, s);
  if (a_major == 1 && b_major == 1) return launch_gemm<BN, MAJOR_MN, MAJOR_MN, EPI>(ta, tb, tbh, D, bias, M, N, K, ldd, splitk, kbps, alpha, s);
  return launch_gemm<BN, MAJOR_MN, MAJOR_K, EPI>(ta, tb, tbh, D, bias, M, N, K, ldd, splitk, kbps, alpha, s);
}

}  // namespace mb
using namespace mb;

// D[M,N] (=|+=) alpha * op(A) op(B) (+ bias[n]).
//   a_major 0: A is [M,K] row-major (lda);  1: A is [K,M] row-major (lda)
//   b_major 0: B is [N,K] row-major (ldb);  1: B is [K,N] row-major (ldb)
//   d_mode  0: bf16 store, 1: fp32 store, 2: fp32 accumulate (atomic; enables split-K)
extern "C" int mb_gemm_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb,
                            int b_major, void* D, long long ldd, int d_mode, const float* bias, int M,
                            int N, int K, float alpha, cudaStream_t stream) {
  MB_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm: negative shape");
  MB_REQUIRE(d_mode >= 0 && d_mode <= 2, "gemm: bad d_mode %d", d_mode);
  if (M == 0 || N == 0) return 0;
  MB_REQUIRE(K > 0, "gemm: K must be positive");
  MB_REQUIRE(((uintptr_t)D & 15) == 0, "gemm: output must be 16-byte aligned");
  const int BN = N > 128 ? 256 : 128;
  CUtensorMap ta, tb, tbh;
  if (a_major == 0) { if (make_tmap_bf16_2d(&ta, A, M, K, lda, BM)) return 1; }
  else { if (make_tmap_bf16_2d(&ta, A, K, M, lda, BK)) return 1; }
  if (b_major == 0) { if (make_tmap_bf16_2d(&tb, B, N, K, ldb, BN)) return 1; }
  else { if (make_tmap_bf16_2d(&tb, B, K, N, ldb, BK)) return 1; }
  // half-height box of the K-major B tile for the CTA-pair path (MN-major boxes are per 64-wide atom already)
  if (b_major == 0) { if (make_tmap_bf16_2d(&tbh, B, N, K, ldb, BN / 2)) return 1; }
  else tbh = tb;
  // split-K only where the epilogue is additive.  The persistent grid runs whole waves of work units
  // (tile pairs x splits when two CTAs share a tile column), so pick the split count that minimises
  // waves x (k-blocks per split + the cost of one more additive epilogue), not just "enough units".
  const int kblocks = cdiv(K, BK);
  int splitk = 1;
  if (d_mode == 2) {
    const bool paired = cdiv(M, BM) >= 2 && BN == 256 && use_pair();
    const int cl = paired ? 2 : 1;
    const long long units = (long long)((cdiv(M, BM) + cl - 1) / cl) * cdiv(N, BN);
    const int workers = num_sms() / cl;
    int max_split = kblocks / 8;  // keep >= 8 k-blocks (512 deep) per split
    if (max_split < 1) max_split = 1;
    if (max_split > 64) max_split = 64;
    long long best = -1;
    for (int sk = 1; sk <= max_split; ++sk) {
      const int per = cdiv(kblocks, sk);
      const int real = cdiv(kblocks, per);
      const long long waves = (units * real + workers - 1) / workers;
      const long long cost = waves * (per + 6);
      if (best < 0 || cost < best) { best = cost; splitk = real; }
    }
  }
  int kbps = cdiv(kblocks, splitk);
  splitk = cdiv(kblocks, kbps);
#define GO(BN_, EPI_) return dispatch_major<BN_, EPI_>(a_major, b_major, ta, tb, tbh, D, bias, M, N, K, ldd, splitk, kbps, alpha, stream)
  if (BN == 256) {
    if (d_mode == 0) GO(256, EPI_BF16);
    if (d_mode == 1) GO(256, EPI_F32);
    GO(256, EPI_F32_ATOMIC);
  } else {
    if (d_mode == 0) GO(128, EPI_BF16);
    if (d_mode == 1) GO(128, EPI_F32);
    GO(128, EPI_F32_ATOMIC);
  }
#undef GO
}
