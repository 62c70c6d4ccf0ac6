// SSD (Mamba-2) tensor-core kernels, part 1:
//   ssd_outer_kernel  K9 chunk states  S[p,n] = sum_t x[t,p] * (exp(a_last - a_t) dt_t) * B[t,n]
//                     and (DSTATE) the backward dS_prev[p,n] = sum_t dy[t,p] * exp(a_t) * C[t,n]
//   ssd_scan_kernel   K8 + K11 fused: CB = C B^T in TMEM, per head mask/decay/dt scaling in
//                     registers -> bf16 operand in smem -> (CB o L) X and C S_prev^T on tcgen05,
//                     epilogue adds D x.  The same kernel in BWD mode computes
//                     dx = dt (M^T dY) + w (B dS^T) + D dY and the two d(dt) partials.
// They replace mamba_ssm's Triton _chunk_state_fwd, _bmm_chunk_fwd, _chunk_scan_fwd,
// _chunk_scan_bwd_dstates and _chunk_scan_chunk_state_bwd_dx.
// CB never touches HBM.  Fixed tile geometry: chunk 256, headdim 64, d_state 128.
#include "ssd_tc.cuh"
#include "../include/mamba_b200.h"

namespace mb {
namespace ssd {

constexpr int kThreads = 256;

// ------------------------------------------------------------------------------------------
// outer-product kernel: out[(hh,p), n] = sum_t arow[t, (hh,p)] * scale_hh[t] * bop[t, n]
// ------------------------------------------------------------------------------------------
constexpr int kOuterHalf = 128;  // K rows staged per pass: 2 passes keep smem at 64 KB -> 3 CTAs / SM
constexpr int kOuterSmem = 2 * (kOuterHalf * 2 * 128) + 2 * kL * 4 + 64 + 1024;
constexpr int kSeqSmem = kOuterSmem + 2 * 2 * 2 * kL * 4;   // + the a | dt stages of ssd_states_seq_kernel

template <bool DSTATE>
__global__ void __launch_bounds__(kThreads, 3)
ssd_outer_kernel(const bf16* __restrict__ arow, long long lda, const __grid_constant__ CUtensorMap tmBop,
                 const float* __restrict__ dtv, const float* __restrict__ av, const __grid_constant__ CUtensorMap tmOut,
                 int T, int H, int G, int nc, int hpc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sA = smem;                                // [128 k-rows][2 atoms]  MN-major A
  uint8_t* sB = smem + kOuterHalf * 2 * 128;         // [128 k-rows][2 atoms]  MN-major B
  float* s_scale = reinterpret_cast<float*>(smem + 2 * kOuterHalf * 2 * 128);  // [2][256]
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_scale + 2 * kL);
  uint64_t* ld_bar = bar + 1;   // TMA transaction barrier of the B / C operand tile
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bc = blockIdx.x;
  const int b = bc / nc, c = bc % nc;
  const int h0 = blockIdx.y * hpc;
  const int g = h0 / (H / G);
  const int valid = min(kL, T - c * kL);

  if (tid == 0) { mbar_init(bar, 1); mbar_init(ld_bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(tmem_slot, 128); tmem_relinquish(); }
  for (int idx = tid; idx < hpc * kL; idx += kThreads) {
    const int hh = idx / kL, t = idx % kL;
    const size_t base = ((size_t)bc * H + h0 + hh) * kL;
    const float a = av[base + t];
    s_scale[hh * kL + t] = DSTATE ? fast_exp(a) : fast_exp(av[base + kL - 1] - a) * dtv[base + t];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const size_t row0 = (size_t)b * T + (size_t)c * kL;

  for (int half = 0; half < kL / kOuterHalf; ++half) {
    const int k0 = half * kOuterHalf;
    const int v = max(0, min(kOuterHalf, valid - k0));
    auto scale = [&](int row, int blk) { return s_scale[blk * kL + k0 + row]; };
    if (warp == 0 && elect_one()) {  // B / C rows of this half by TMA (two 64-column atoms); rows past the sequence end read as zeros
      mbar_arrive_expect_tx(ld_bar, kOuterHalf * 2 * 128);
      for (int j2 = 0; j2 < 2; ++j2)
        tma_load_3d(sB + j2 * kOuterHalf * 128, &tmBop, ld_bar, g * kN + j2 * 64, c * kL + k0, b);
    }
    if (hpc == 2) load_tile<2, true>(sA, arow + (row0 + k0) * lda + h0 * kP, lda, kOuterHalf, v, scale, tid, kThreads);
    else load_tile<1, true>(sA, arow + (row0 + k0) * lda + h0 * kP, lda, kOuterHalf, v, scale, tid, kThreads);
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {  // converged warp, one elected lane issues (ptx.cuh: warp-uniform single-lane issue)
      mbar_wait_warp(ld_bar, half & 1);
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc_bf16(128, kN, MAJOR_MN, MAJOR_MN);
      const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < kOuterHalf / 16; ++k)
          mma_bf16_ss(tmem, desc_mnmajor(a_addr, kOuterHalf, k), desc_mnmajor(b_addr, kOuterHalf, k), idesc,
                      (half > 0 || k > 0) ? 1u : 0u);
        mma_commit(bar);
      }
      __syncwarp();
    }
    mbar_wait(bar, half & 1);  // operands consumed: smem may be refilled / accumulator readable
    tc_fence_after();
  }

  // Epilogue: a thread owns one accumulator row, so direct stores would hit 32 different lines with 16 bytes
  // each.  Every warp instead stages its [32 rows][32 fp32] blocks in the (now idle) operand buffers,
  // 128B-swizzled, and one lane hands each block to a TMA store.
  const int lg = warp & 3, colhalf = warp >> 2;
  if (lg * 32 < hpc * kP) {
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      const int col = colhalf * 64 + ch * 32;
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + col, v);
      tmem_ld_wait();
      uint8_t* stg = smem + warp * 8192 + ch * 4096;
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<uint4*>(stg + lane * 128 + ((q ^ (lane & 7)) << 4)) =
            make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      if (elect_one()) {   // the same lane every time (elect.sync is deterministic for a given mask)
        tma_store_2d(&tmOut, stg, col, (bc * H + h0) * kP + lg * 32);
        tma_store_commit();
      }
    }
    if (elect_one()) tma_store_wait_read<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

// ------------------------------------------------------------------------------------------
// K9 + K10 in one kernel: chunk states AND the inter-chunk recurrence.
//
// One CTA owns a (batch, head pair) and walks the chunks of the sequence in order (backward: in reverse order),
// so the recurrence  S_c = exp(a_last_c) S_{c-1} + X_c^T (w_c o B_c)  is carried in registers (64 fp32 per thread
// in the TMEM lane layout of the accumulator) and only what the scan kernels consume ever reaches HBM: the bf16
// state ENTERING each chunk (forward) / the bf16 gradient of each chunk's own state (backward).  The separate
// state-passing kernels and their fp32 [B, nc, H, P, N] round trip (write 100 MB, read 100 MB per layer and
// direction at B = 32, T = 1024) disappear.  Also native here: an initial state / its gradient and the final state
// (mamba_chunk_scan_combined's initial_states / return_final_states).
//   forward : sprev[c] = S_{c-1} ;  S_c = d_c S_{c-1} + states_c                         (S_{-1} = init or 0)
//   backward: ds[c]    = g_c     ;  da_last[c] += d_c <g_c, sprev[c]> ;  g_{c-1} = d_c g_c + dstates_c   (g_{nc-1} = dfinal or 0)
// Parallelism is B x H/2 CTAs: used when that fills the GPU, otherwise the per-chunk kernels above run.
// ------------------------------------------------------------------------------------------
template <bool DSTATE>
__global__ void __launch_bounds__(kThreads, 2)
ssd_states_seq_kernel(const bf16* __restrict__ arow, long long lda, const __grid_constant__ CUtensorMap tmBop,
                      const float* __restrict__ dtv, const float* __restrict__ av,
                      const __grid_constant__ CUtensorMap tmOut,          // bf16 [B*nc*H*P][N], box {64, 32}
                      const bf16* __restrict__ sprev, float* __restrict__ da_last,   // backward only
                      const float* __restrict__ init, float* __restrict__ last_out,  // fp32 [B, H, P, N] or null
                      int T, int H, int G, int nc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sA = smem;                                // [128 k-rows][2 atoms]  MN-major A
  uint8_t* sB = smem + kOuterHalf * 2 * 128;         // [128 k-rows][2 atoms]  MN-major B
  float* s_scale = reinterpret_cast<float*>(smem + 2 * kOuterHalf * 2 * 128);  // [2][256]
  float* s_raw = s_scale + 2 * kL;   // [2 stages][2 heads][a(256) | dt(256)]: a | dt of the NEXT chunk, by cp.async
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_raw + 2 * 2 * 2 * kL);
  uint64_t* ld_bar = bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x;
  const int h0 = blockIdx.y * 2;
  const int g = h0 / (H / G);
  const int lg = warp & 3, colhalf = warp >> 2;
  const int row = lg * 32 + lane;              // accumulator row = (head of the pair, p)
  const int hh = row >> 6, pp = row & 63;

  if (tid == 0) { mbar_init(bar, 1); mbar_init(ld_bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(tmem_slot, 128); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();

  float run[64];   // this thread's 64 columns (colhalf * 64 ..) of the carried state / gradient
  {
    const float* src = init != nullptr ? init + (((size_t)b * H + h0 + hh) * kP + pp) * kN + colhalf * 64 : nullptr;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      const float4 v = src != nullptr ? *reinterpret_cast<const float4*>(src + 4 * q) : make_float4(0.f, 0.f, 0.f, 0.f);
      run[4 * q] = v.x; run[4 * q + 1] = v.y; run[4 * q + 2] = v.z; run[4 * q + 3] = v.w;
    }
  }
  uint32_t mma_phase = 0, ld_phase = 0;
  // a | dt of a chunk (2 heads x 256 tokens x 2 values = 4 KB) are fetched one chunk ahead with cp.async: their
  // global latency was 14 % of this kernel's stall samples when each chunk loaded them on the spot.
  auto fetch_adt = [&](int c_, int stage) {
    const int bc_ = b * nc + c_;
    for (int idx = tid; idx < 2 * kL; idx += kThreads) {
      const int h2 = idx / kL, t = idx % kL;
      const size_t base = ((size_t)bc_ * H + h0 + h2) * kL + t;
      const uint32_t dst = smem_u32(s_raw + (stage * 2 + h2) * 2 * kL + t);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(av + base) : "memory");
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + kL * 4), "l"(dtv + base) : "memory");
    }
    cp_async_commit();
  };
  fetch_adt(DSTATE ? nc - 1 : 0, 0);
  for (int ci = 0; ci < nc; ++ci) {
    const int c = DSTATE ? nc - 1 - ci : ci;
    const int bc = b * nc + c;
    const int valid = min(kL, T - c * kL);
    if (ci + 1 < nc) { fetch_adt(DSTATE ? c - 1 : c + 1, (ci + 1) & 1); cp_async_wait<1>(); }
    else cp_async_wait<0>();
    __syncthreads();   // this chunk's a | dt (copied by all threads) are visible
    const float* ra = s_raw + (ci & 1) * 2 * 2 * kL;
    for (int idx = tid; idx < 2 * kL; idx += kThreads) {
      const int h2 = idx / kL, t = idx % kL;
      const float a = ra[h2 * 2 * kL + t];
      s_scale[h2 * kL + t] = DSTATE ? fast_exp(a) : fast_exp(ra[h2 * 2 * kL + kL - 1] - a) * ra[h2 * 2 * kL + kL + t];
    }
    const float decay = fast_exp(ra[hh * 2 * kL + kL - 1]);
    __syncthreads();   // scale vector visible; the previous chunk's TMA stores have finished reading the operand buffers
    const size_t row0 = (size_t)b * T + (size_t)c * kL;
    for (int half = 0; half < kL / kOuterHalf; ++half) {
      const int k0 = half * kOuterHalf;
      const int v = max(0, min(kOuterHalf, valid - k0));
      auto scale = [&](int r_, int blk) { return s_scale[blk * kL + k0 + r_]; };
      if (warp == 0 && elect_one()) {
        mbar_arrive_expect_tx(ld_bar, kOuterHalf * 2 * 128);
        for (int j2 = 0; j2 < 2; ++j2)
          tma_load_3d(sB + j2 * kOuterHalf * 128, &tmBop, ld_bar, g * kN + j2 * 64, c * kL + k0, b);
      }
      load_tile<2, true>(sA, arow + (row0 + k0) * lda + h0 * kP, lda, kOuterHalf, v, scale, tid, kThreads);
      fence_proxy_async_smem();
      __syncthreads();
      if (warp == 0) {
        mbar_wait_warp(ld_bar, ld_phase);
        tc_fence_after();
        constexpr uint32_t idesc = make_idesc_bf16(128, kN, MAJOR_MN, MAJOR_MN);
        const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kOuterHalf / 16; ++k)
            mma_bf16_ss(tmem, desc_mnmajor(a_addr, kOuterHalf, k), desc_mnmajor(b_addr, kOuterHalf, k), idesc,
                        (half > 0 || k > 0) ? 1u : 0u);
          mma_commit(bar);
        }
        __syncwarp();
      }
      ld_phase ^= 1;
      mbar_wait(bar, mma_phase);  // operands consumed: smem may be refilled / accumulator readable
      mma_phase ^= 1;
      tc_fence_after();
    }
    // ---- epilogue of the chunk: emit the carried value (bf16), then advance the recurrence ----
    uint8_t* stg = smem + warp * 4096;   // [32 rows][128 B] staging in the idle operand buffers (128B swizzle)
    float part = 0.f;
    if (DSTATE) {  // <g_c, S_{c-1}> for d a_last: this thread's 64 columns of its row, 128 contiguous bytes
      const bf16* sp = sprev + (((size_t)bc * H + h0 + hh) * kP + pp) * kN + colhalf * 64;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float f[8];
        unpack8(ld8(sp + 8 * q), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) part += run[8 * q + e] * f[e];
      }
    }
#pragma unroll
    for (int q = 0; q < 8; ++q)
      *reinterpret_cast<bf16x8*>(stg + lane * 128 + ((q ^ (lane & 7)) << 4)) = pack8(run + 8 * q);
    fence_proxy_async_smem();
    __syncwarp();
    if (elect_one()) {
      tma_store_2d(&tmOut, stg, colhalf * 64, (bc * H + h0) * kP + lg * 32);
      tma_store_commit();
    }
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + colhalf * 64 + ch * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) run[ch * 32 + j] = fmaf(decay, run[ch * 32 + j], __uint_as_float(v[j]));
    }
    if (DSTATE) {
      part = warp_sum(part);   // all 32 rows of a warp belong to one head
      if (lane == 0) atomicAdd(da_last + (size_t)bc * H + h0 + hh, decay * part);
    }
    if (elect_one()) tma_store_wait_read<0>();   // staging (= operand buffer) free before the next chunk's loads
    tc_fence_before();
  }
  if (last_out != nullptr) {   // forward: final state; backward: gradient of the initial state
    float* dst = last_out + (((size_t)b * H + h0 + hh) * kP + pp) * kN + colhalf * 64;
#pragma unroll
    for (int q = 0; q < 16; ++q)
      *reinterpret_cast<float4*>(dst + 4 * q) = make_float4(run[4 * q], run[4 * q + 1], run[4 * q + 2], run[4 * q + 3]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

// ------------------------------------------------------------------------------------------
// scan kernel (FWD: y ; BWD: dx, ddt partials) — warp-specialised, software-pipelined over heads
//
//   warps 0-7   "scale"     read the CB tile from TMEM, apply decay / mask / dt, write the bf16
//                           operand G one 128-column half-tile at a time into a 2-slot smem ring
//   warps 8-15  "epilogue"  read the finished accumulators of the previous head from TMEM, add the
//                           skip terms, store y / dx
//   warp 16     "mma"       one lane issues tcgen05.mma: C.S^T (per head) and G.X (per half-tile);
//                           tcgen05.commit frees G slots and publishes accumulators
//   warp 17     "producer"  one lane issues the TMA loads: row / column operands, then the X / state
//                           tiles of head+2 as soon as the epilogue warps hold their X rows of head
// Accumulators are double-buffered in TMEM (CB 256 cols + 2 x (64 + 64)), per-head tiles in smem, so
// head h+1 is being scaled while head h is multiplied and head h-1 is written out.
// ------------------------------------------------------------------------------------------
#ifndef MB_TBLK
#define MB_TBLK 5
#endif
#ifndef MB_TBWD
#define MB_TBWD 0
#endif
#ifdef MB_SSD_TIMING
__device__ long long g_tlog[32][64];   // [warp][event]: store-only marks (no atomics -> ~tens of cycles each)
__device__ int g_tmeta[32][64];
#endif
constexpr int kScanWarpsScale = 8;   // measured: 16 scale warps (one 32-column chunk each) is not faster
constexpr int kScanWarpsProd = 1;    // one lane of it issues every TMA load (operands, per-head X / state tiles)
constexpr int kScanWarpsEpi = 8;     // two per 32-row TMEM lane group, 32 head-dim columns each
constexpr int kScanThreads = (kScanWarpsScale + kScanWarpsEpi + 1 + kScanWarpsProd) * 32;
constexpr int kScanMaxHeads = 6;     // heads per CTA sharing one C.B^T (amortises the prologue)
constexpr int kScanSmem = 128 * 2 * 128     /* sR  row operand  [128][2 kblk]               32 KB */
                        + 2 * (128 * 128 * 2) /* sG ring: 2 x [128][128] bf16 (aliases sC)     64 KB */
                        + 2 * kL * 128      /* sV  X_h / dY_h   [256][1 atom]   x2 buffers   64 KB */
                        + 2 * kP * 2 * 128  /* sS  S_prev_h / dS_h [64][2 kblk] x2           32 KB */
                        + kScanMaxHeads * 2 * kL * 4 /* a | dt of every head of the group, loaded once    16 KB */
                        + kScanWarpsScale * 64 * 4 /* per-scale-warp column-factor scratch                2 KB */
                        + kScanWarpsEpi * 2048 /* per-epilogue-warp TMA-store staging [32][64 B]      16 KB */
                        + 256 + 1024;

template <bool BWD>
__global__ void __launch_bounds__(kScanThreads, 1)
ssd_scan_kernel(const __grid_constant__ CUtensorMap tmRow, const __grid_constant__ CUtensorMap tmCol,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmS,
                const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmOut2,
                const float* __restrict__ dtv, const float* __restrict__ av, const float* __restrict__ Dp,
 float* __restrict__ ddt1, float* __restrict__ ddt2, float* __restrict__ dD, int T,
                int H, int G, int nc, int hgs) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sR = smem;
  uint8_t* sG = sR + 128 * 2 * 128;   // ring of two 32 KB half-tiles; holds the column operand first
  uint8_t* sC = sG;
  uint8_t* sV0 = sG + 2 * 32768;
  uint8_t* sS0 = sV0 + 2 * kL * 128;
  float* s_ad = reinterpret_cast<float*>(sS0 + 2 * kP * 2 * 128);  // [hgs <= kScanMaxHeads][a(256) | dt(256)]
  uint8_t* s_stage = reinterpret_cast<uint8_t*>(s_ad + kScanMaxHeads * 2 * kL);  // [epilogue warp][32 rows][64 B], 64B swizzle
  float* s_wscr = reinterpret_cast<float*>(s_stage + kScanWarpsEpi * 2048);      // [scale warp][2][32] column factors
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_wscr + kScanWarpsScale * 64);
  uint64_t* cb_bar = bars;            // 1
  uint64_t* g_full = bars + 1;        // [2]  one arrival per scale warp
  uint64_t* g_empty = bars + 3;       // [2]  tcgen05.commit
  uint64_t* acc_full = bars + 5;      // [2]  tcgen05.commit
  uint64_t* acc_empty = bars + 7;     // [2]  one arrival per epilogue warp
  uint64_t* tiles_full = bars + 9;    // [2]  TMA transaction barrier (expect_tx by the producer lane)
  uint64_t* tiles_empty = bars + 11;  // [2]  one arrival per epilogue warp once its X row is in registers
  uint64_t* ld_bar = bars + 13;       // 1    TMA transaction barrier for the row / column operands
  uint64_t* x_full = bars + 14;       // [kScanWarpsEpi]  backward: per-warp TMA load of its 32 x 32 tile of x
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14 + kScanWarpsEpi);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // Long CTAs first: the row half with two column halves of work (forward: the lower rows; backward: the upper
  // ones) takes ~1.6x as long as the other, so launching all of them before the short ones leaves only short
  // CTAs in the ragged last wave.
  const int half_ctas = (int)gridDim.x >> 1;
  const int first = (int)blockIdx.x < half_ctas ? 1 : 0;
  const int rh = BWD ? 1 - first : first;
  const int bc = (int)blockIdx.x - (first ? 0 : half_ctas);
  const int b = bc / nc, c = bc % nc;
  const int h_first = blockIdx.y * hgs;
  const int g = h_first / (H / G);
  const int valid = min(kL, T - c * kL);
  const int r0 = rh * 128;
  const int c0 = BWD ? rh * 128 : 0;
  const int ncols = BWD ? (2 - rh) * 128 : (rh + 1) * 128;
  const int nhalves = ncols >> 7;

  // per-head tiles by TMA (one lane): V tile [ncols][64] as 128-row boxes, state tile [64][128] as two
  // 64-column boxes; both land 128B-swizzled exactly as the UMMA descriptors expect them
  auto issue_head = [&](int hh) {
    const int h = h_first + hh;
    const int bsel = hh & 1;
    uint64_t* bar = &tiles_full[bsel];
    mbar_arrive_expect_tx(bar, (uint32_t)(ncols * 128 + kP * 2 * 128));
    for (int rb = 0; rb < nhalves; ++rb)
      tma_load_3d(sV0 + bsel * kL * 128 + rb * 128 * 128, &tmV, bar, h * kP, c * kL + c0 + rb * 128, b);
    for (int kb = 0; kb < 2; ++kb)
      tma_load_2d(sS0 + bsel * kP * 2 * 128 + kb * kP * 128, &tmS, bar, kb * 64, (bc * H + h) * kP);
  };

  // The producer lane initialises the barriers and, without waiting for anyone, starts the TMA loads of the
  // row / column operands and of the tiles of heads 0 and 1: they fly while the other warps allocate TMEM
  // and fetch a | dt (everything else waits for the __syncthreads below before touching a barrier).
  if (warp == kScanWarpsScale + kScanWarpsEpi + 1 && elect_one()) {
    mbar_init(cb_bar, 1);
    mbar_init(ld_bar, 1);
    for (int i = 0; i < kScanWarpsEpi; ++i) mbar_init(&x_full[i], 1);
    for (int i = 0; i < 2; ++i) {
      // one arrival per WARP (after __syncwarp): hundreds of per-thread arrivals on one mbarrier word serialise
      mbar_init(&g_full[i], kScanWarpsScale); mbar_init(&g_empty[i], 1); mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kScanWarpsEpi); mbar_init(&tiles_full[i], 1); mbar_init(&tiles_empty[i], kScanWarpsEpi);
    }
    fence_mbar_init();
    pdl_wait();   // the loads below read the previous kernel's output (common.cuh: programmatic dependent launch)
    mbar_arrive_expect_tx(ld_bar, (uint32_t)((128 + ncols) * 2 * 128));
    for (int kb = 0; kb < 2; ++kb) {
      tma_load_3d(sR + kb * 128 * 128, &tmRow, ld_bar, g * kN + kb * 64, c * kL + r0, b);
      for (int rb = 0; rb < nhalves; ++rb)
        tma_load_3d(sC + kb * ncols * 128 + rb * 128 * 128, &tmCol, ld_bar, g * kN + kb * 64, c * kL + c0 + rb * 128, b);
    }
    issue_head(0);
    if (hgs > 1) issue_head(1);
  }
  if (warp == kScanWarpsScale + kScanWarpsEpi) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();

  // ---- prologue: a | dt of every head of the group (2 KB each, cp.async by everyone) ----
  for (int i = tid; i < hgs * kL; i += kScanThreads) {
    const int hh = i / kL, t = i % kL;
    const size_t sb = ((size_t)bc * H + h_first + hh) * kL;
    const uint32_t ad = smem_u32(s_ad + hh * 2 * kL);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ad + t * 4), "l"(av + sb + t) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ad + (kL + t) * 4), "l"(dtv + sb + t) : "memory");
  }
  cp_async_commit();
  cp_async_wait<0>();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tCB = tmem;
  if (warp == kScanWarpsScale + kScanWarpsEpi) {
    mbar_wait_warp(ld_bar, 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, ncols, MAJOR_K, MAJOR_K);
    const uint32_t a_addr = smem_u32(sR), b_addr = smem_u32(sC);
    if (elect_one()) {
#pragma unroll
      for (int k = 0; k < kN / 16; ++k)
        mma_bf16_ss(tCB, desc_kmajor(a_addr, 128, k >> 2, k & 3), desc_kmajor(b_addr, ncols, k >> 2, k & 3), idesc,
                    k > 0 ? 1u : 0u);
      mma_commit(cb_bar);
    }
    __syncwarp();
  }
  mbar_wait(cb_bar, 0);  // CB complete: the column operand is dead, its smem becomes the G ring
  tc_fence_after();

#ifdef MB_SSD_TIMING
  const bool tprint = blockIdx.x == MB_TBLK && blockIdx.y == 0 && (BWD == (MB_TBWD != 0));
  const long long t_start = clock64();
  int tidx = 0;
#define TMARK(role, hh, what) if (tprint && lane == 0 && tidx < 64) { g_tlog[warp][tidx] = clock64() - t_start; g_tmeta[warp][tidx] = ((hh) << 8) | (what[0]); ++tidx; }
#else
#define TMARK(role, hh, what)
#endif
  if (warp < kScanWarpsScale) {
    // =================================== scale warps ===================================
    const int lg = warp & 3, sub = warp >> 2;      // 32-row lane group, column slice of the half-tile
    const int r = lg * 32 + lane, ri = r0 + r;
    const uint32_t lane_sel = (uint32_t)(lg * 32) << 16;
    const int warp_row_min = r0 + lg * 32, warp_row_max = warp_row_min + 31;
    float* wscr = s_wscr + warp * 64;
    int unit = 0;
    for (int hh = 0; hh < hgs; ++hh) {
      TMARK("scale", hh, "tiles_ready");
      const float* s_a = s_ad + hh * 2 * kL;
      const float* s_dt = s_a + kL;
      const float a_i = s_a[ri];
      const float a_ref = s_a[BWD ? warp_row_max : warp_row_min];
      const float rowf = fast_exp(fminf(BWD ? a_ref - a_i : a_i - a_ref, 0.f));
      const bool diag_ok = fabsf(s_a[warp_row_min] - s_a[warp_row_max]) < 60.f;  // warp-uniform
      for (int half = 0; half < nhalves; ++half, ++unit) {
        const int slot = unit & 1, use = unit >> 1;
        if (use > 0) mbar_wait(&g_empty[slot], (use - 1) & 1);
        TMARK("scale", hh, "slot_free");
        constexpr int kChunks = 128 / (32 * (kScanWarpsScale / 4));  // 32-column chunks per warp per half-tile
#pragma unroll
        for (int cq = 0; cq < kChunks; ++cq) {
          const int c32 = sub * kChunks + cq;              // 32-column chunk index (0..3) inside the half-tile
          uint8_t* gdst = sG + slot * 32768 + (size_t)(c32 >> 1) * (128 * 128) + r * 128;
          const int kq = c32 & 1;                          // which 32-column half of the 64-column k-block
          const int kcol = half * 128 + c32 * 32;          // column inside the CB tile
          const int cj0 = c0 + kcol;
          const bool skip = BWD ? (cj0 + 31 < warp_row_min) : (cj0 > warp_row_max);  // warp-uniform
          // strictly off-diagonal chunk: every (i, j) of the 32 x 32 block is unmasked and
          // a_i <= a_ref <= a_j (fwd; mirrored in bwd), so exp(a_i - a_j) factors into a per-row and a
          // per-column term that are both <= 1 -- 2 multiplies per element instead of an exp + mask
          const bool offdiag = BWD ? (cj0 > warp_row_max) : (cj0 + 31 < warp_row_min);  // warp-uniform
          // The chunk ON the diagonal factors the same way around the same a_ref (the column factor is then
          // >= 1 instead of <= 1) as long as the decay across its 32 rows stays far from the fp32 range
          // (diag_ok, checked per head); only the causal mask is added.  Otherwise: the exact path below.
          const bool fast = offdiag || (!skip && diag_ok);
          uint32_t pk[16];  // 32 bf16 results, packed as soon as they are produced
          if (fast) {
            float* scr = wscr + cq * 32;
            {
              const float aj = s_a[cj0 + lane];
              float cf = fast_exp(fminf(BWD ? aj - a_ref : a_ref - aj, offdiag ? 0.f : 64.f));
              if (!BWD) cf *= s_dt[cj0 + lane];
              scr[lane] = cf;
            }
            uint32_t v[32];
            tmem_ld32(tCB + lane_sel + kcol, v);
            __syncwarp();
            tmem_ld_wait();
            const float2 rf2 = f2(rowf, rowf);
            if (offdiag) {
#pragma unroll
              for (int q = 0; q < 8; ++q) {  // column pairs: FMUL2 x 2 + one pack per pair
                const float4 c4 = *reinterpret_cast<const float4*>(scr + 4 * q);
                const float2 lo2 = mul2(mul2(f2(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1])), rf2), f2(c4.x, c4.y));
                const float2 hi2 = mul2(mul2(f2(__uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])), rf2), f2(c4.z, c4.w));
                const bf162 lo = __floats2bfloat162_rn(lo2.x, lo2.y), hi = __floats2bfloat162_rn(hi2.x, hi2.y);
                pk[2 * q] = *reinterpret_cast<const uint32_t*>(&lo);
                pk[2 * q + 1] = *reinterpret_cast<const uint32_t*>(&hi);
              }
            } else {
#pragma unroll
              for (int q = 0; q < 8; ++q) {  // same, plus the causal mask: column t of the chunk vs row `lane`
                const float4 c4 = *reinterpret_cast<const float4*>(scr + 4 * q);
                float2 lo2 = mul2(mul2(f2(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1])), rf2), f2(c4.x, c4.y));
                float2 hi2 = mul2(mul2(f2(__uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])), rf2), f2(c4.z, c4.w));
                lo2.x = (BWD ? (4 * q >= lane) : (4 * q <= lane)) ? lo2.x : 0.f;
                lo2.y = (BWD ? (4 * q + 1 >= lane) : (4 * q + 1 <= lane)) ? lo2.y : 0.f;
                hi2.x = (BWD ? (4 * q + 2 >= lane) : (4 * q + 2 <= lane)) ? hi2.x : 0.f;
                hi2.y = (BWD ? (4 * q + 3 >= lane) : (4 * q + 3 <= lane)) ? hi2.y : 0.f;
                const bf162 lo = __floats2bfloat162_rn(lo2.x, lo2.y), hi = __floats2bfloat162_rn(hi2.x, hi2.y);
                pk[2 * q] = *reinterpret_cast<const uint32_t*>(&lo);
                pk[2 * q + 1] = *reinterpret_cast<const uint32_t*>(&hi);
              }
            }
          } else if (!skip) {
            uint32_t v[32];
            tmem_ld32(tCB + lane_sel + kcol, v);
            tmem_ld_wait();
            // branch-free (the causal mask is a bit mask) and in groups of 4 columns whose a_j / dt_j
            // come from one 128-bit broadcast load each: keeps ~16 independent exp chains in flight
            // without holding 4 x 32 fp32 values live
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 a4 = *reinterpret_cast<const float4*>(s_a + cj0 + 4 * q);
              float4 d4 = make_float4(1.f, 1.f, 1.f, 1.f);
              if (!BWD) d4 = *reinterpret_cast<const float4*>(s_dt + cj0 + 4 * q);
              const float aj[4] = {a4.x, a4.y, a4.z, a4.w}, dj[4] = {d4.x, d4.y, d4.z, d4.w};
              float o4[4];
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const int cj = cj0 + 4 * q + t;
                const float e = fast_exp(fminf(BWD ? aj[t] - a_i : a_i - aj[t], 0.f));
                const float val = __uint_as_float(v[4 * q + t]) * e * dj[t];
                const unsigned keep = (BWD ? (cj >= ri) : (cj <= ri)) ? 0xffffffffu : 0u;
                o4[t] = __uint_as_float(__float_as_uint(val) & keep);
              }
              const bf162 lo = __floats2bfloat162_rn(o4[0], o4[1]), hi = __floats2bfloat162_rn(o4[2], o4[3]);
              pk[2 * q] = *reinterpret_cast<const uint32_t*>(&lo);
              pk[2 * q + 1] = *reinterpret_cast<const uint32_t*>(&hi);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[j] = 0u;
          }
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<uint4*>(gdst + (((kq * 4 + q) ^ (r & 7)) << 4)) =
                make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&g_full[slot]);
        TMARK("scale", hh, "half_done");
      }
    }
  } else if (warp < kScanWarpsScale + kScanWarpsEpi) {
    // =================================== epilogue warps ===================================
    const int lg = warp & 3;
    const int pc0 = ((warp - kScanWarpsScale) >> 2) * 32;   // this warp's 32 of the 64 head-dim columns
    const int r = lg * 32 + lane, ri = r0 + r;
    const uint32_t lane_sel = (uint32_t)(lg * 32) << 16;
    const int ew = warp - kScanWarpsScale;
    uint8_t* stg = s_stage + ew * 2048;            // this warp's [32 rows][64 B] staging tile (64B swizzle)
    const int trow = c * kL + r0 + lg * 32;
    for (int hh = 0; hh < hgs; ++hh) {
      const int h = h_first + hh;
      const int acc = hh & 1;
      if (BWD && elect_one()) {  // backward also needs x[rows, h, cols]: TMA it into the staging tile while waiting
        tma_store_wait_read<0>();
        mbar_arrive_expect_tx(&x_full[ew], 2048);
        tma_load_3d(stg, &tmOut2, &x_full[ew], h * kP + pc0, trow, b);
      }
      TMARK("epi", hh, "wait_acc");
      mbar_wait(&acc_full[acc], (hh >> 1) & 1);
      tc_fence_after();
      TMARK("epi", hh, "acc_ready");
      const float* s_a = s_ad + hh * 2 * kL;
      const float* s_dt = s_a + kL;
      const uint8_t* sV = sV0 + acc * kL * 128;
      const int vrow = ri - c0;
      // The accumulators being complete means the tensor core is done with this head's X / state tiles.
      // Take this thread's X row (128 B) into registers; once all four epilogue warps have theirs the
      // slot is free and the producer warps stream the tiles of head hh+2 in underneath this epilogue.
      mbar_wait(&tiles_full[acc], (hh >> 1) & 1);  // long complete; makes the TMA-written tile visible to this thread
      uint4 xrow[4];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        xrow[q] = *reinterpret_cast<const uint4*>(sV + vrow * 128 + ((((pc0 >> 3) + q) ^ (vrow & 7)) << 4));
      __syncwarp();
      if (lane == 0) mbar_arrive(&tiles_empty[acc]);
      const uint32_t tOff = tmem + 256 + acc * 128, tDiag = tOff + 64;
      const size_t sbase = ((size_t)bc * H + h) * kL;
      const float a_i = s_a[ri];
      const float Dh = Dp[h];
      float s1 = 0.f, s2 = 0.f, sd = 0.f;
      const float dt_i = s_dt[ri];
      const float el = fast_exp(s_a[kL - 1] - a_i);
      const float ea = fast_exp(a_i);
      if (BWD) mbar_wait(&x_full[ew], hh & 1);
      uint4 pk_y[4], pk_off[4];  // this row's 32 output columns (and, forward, the y_off ones), packed bf16
#pragma unroll
      for (int ph = 0; ph < 2; ++ph) {  // 16 head-dim columns at a time
        const int p0 = pc0 + ph * 16;
        uint32_t vo[16], vd[16];
        tmem_ld16(tOff + lane_sel + p0, vo);
        tmem_ld16(tDiag + lane_sel + p0, vd);
        tmem_ld_wait();
        float xv[16];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          bf16x8 t8;
          memcpy(&t8, &xrow[2 * ph + q], 16);
          unpack8(t8, xv + 8 * q);
        }
        float o[16];
        if (!BWD) {
#pragma unroll
          for (int j = 0; j < 16; ++j) o[j] = ea * __uint_as_float(vo[j]);
#pragma unroll
          for (int q = 0; q < 2; ++q) {  // y_off = exp(a_t) C_t S_prev^T alone is kept for backward (see ssd_bwd_finalize)
            const bf16x8 t8 = pack8(o + 8 * q);
            memcpy(&pk_off[2 * ph + q], &t8, 16);
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) o[j] += __uint_as_float(vd[j]) + Dh * xv[j];
        } else {
          {  // rows past the end of the sequence read as zeros (TMA fill), so no guard is needed
#pragma unroll
            for (int q = 0; q < 2; ++q) {
              float xx[8];
              bf16x8 t8;
              memcpy(&t8, stg + lane * 64 + (((2 * ph + q) ^ ((lane >> 1) & 3)) << 4), 16);
              unpack8(t8, xx);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                s1 += xx[j] * __uint_as_float(vd[8 * q + j]);
                s2 += xx[j] * __uint_as_float(vo[8 * q + j]);
                sd += xx[j] * xv[8 * q + j];  // dD_h = sum dy . x
              }
            }
          }
          const float w = el * dt_i;
#pragma unroll
          for (int j = 0; j < 16; ++j) o[j] = dt_i * __uint_as_float(vd[j]) + w * __uint_as_float(vo[j]) + Dh * xv[j];
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const bf16x8 t8 = pack8(o + 8 * q);
          memcpy(&pk_y[2 * ph + q], &t8, 16);
        }
      }
      // the accumulators are in registers: hand the TMEM buffer back before the stores
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
      // Row-per-thread global stores would touch 32 different lines with 16 bytes each.  Instead the warp
      // lays its 32 x 32 tile out in shared memory (64B-swizzled rows) and one lane issues a TMA store;
      // the tensor map is [B][T][H*P], so rows past the end of the sequence are dropped by the hardware.
      auto stage_and_store = [&](const uint4* pk, const CUtensorMap* map) {
        if (elect_one()) tma_store_wait_read<0>();  // the previous store has finished reading the staging tile
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(stg + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4)) = pk[q];
        fence_proxy_async_smem();
        __syncwarp();
        if (elect_one()) {
          tma_store_3d(map, stg, h * kP + pc0, trow, b);
          tma_store_commit();
        }
      };
      if (!BWD) stage_and_store(pk_off, &tmOut2);
      stage_and_store(pk_y, &tmOut);
      if (BWD) {
        if (ri < valid) {  // two threads (one per column half) own a (b, c, h, row): buffers are pre-zeroed
          atomicAdd(ddt1 + sbase + ri, s1);
          atomicAdd(ddt2 + sbase + ri, el * s2);
        }
        sd = warp_sum(sd);
        if (lane == 0) atomicAdd(dD + h, sd);
      }
      TMARK("epi", hh, "epi_done");
    }
    if (elect_one()) tma_store_wait_read<0>();  // staging must outlive the last store's read
  } else if (warp == kScanWarpsScale + kScanWarpsEpi) {
    // =================================== MMA issuer ===================================
    // all 32 lanes run the loops and the waits (converged), one elected lane issues each group of MMAs: the
    // descriptors then live in uniform registers and the 8 MMAs of a group go out back to back
    {
    const uint32_t r_addr = smem_u32(sR);
    constexpr uint32_t idesc_off = make_idesc_bf16(128, kP, MAJOR_K, MAJOR_K);
    constexpr uint32_t idesc_diag = make_idesc_bf16(128, kP, MAJOR_K, MAJOR_MN);
    int unit = 0;
    for (int hh = 0; hh < hgs; ++hh) {
      const int acc = hh & 1;
      if (hh >= 2) mbar_wait_warp(&acc_empty[acc], ((hh >> 1) - 1) & 1);
      mbar_wait_warp(&tiles_full[acc], (hh >> 1) & 1);
      tc_fence_after();
      TMARK("mma", hh, "ready");
      const uint32_t tOff = tmem + 256 + acc * 128, tDiag = tOff + 64;
      const uint32_t s_addr = smem_u32(sS0 + acc * kP * 2 * 128), v_addr = smem_u32(sV0 + acc * kL * 128);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < kN / 16; ++k)  // (C | B)[128 x 128] . (S_prev | dS)[64 x 128]^T
          mma_bf16_ss(tOff, desc_kmajor(r_addr, 128, k >> 2, k & 3), desc_kmajor(s_addr, kP, k >> 2, k & 3), idesc_off,
                      k > 0 ? 1u : 0u);
      }
      __syncwarp();
      for (int half = 0; half < nhalves; ++half, ++unit) {
        const int slot = unit & 1;
        mbar_wait_warp(&g_full[slot], (unit >> 1) & 1);
        tc_fence_after();
        TMARK("mma", hh, "g_full");
        const uint32_t g_addr = smem_u32(sG + slot * 32768);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)  // G half-tile [128 x 128] . V[128 x 64]
            mma_bf16_ss(tDiag, desc_kmajor(g_addr, 128, k >> 2, k & 3), desc_mnmajor(v_addr, ncols, half * 8 + k),
                        idesc_diag, (half > 0 || k > 0) ? 1u : 0u);
          mma_commit(&g_empty[slot]);
        }
        __syncwarp();
      }
      if (elect_one()) mma_commit(&acc_full[acc]);
      __syncwarp();
    }
    }
  } else {
    // =================================== TMA producer ===================================
    for (int hh = 2; hh < hgs; ++hh) {
      mbar_wait_warp(&tiles_empty[hh & 1], ((hh >> 1) - 1) & 1);
      if (elect_one()) issue_head(hh);
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kScanWarpsScale + kScanWarpsEpi) tmem_dealloc(tmem, 512);
#ifdef MB_SSD_TIMING
  if (tprint && lane == 0) {
    for (int i = 0; i < tidx; ++i) printf("T %d %d %c %lld\n", warp, (g_tmeta[warp][i] >> 8) & 255, (char)(g_tmeta[warp][i] & 255), g_tlog[warp][i]);
  }
#endif
}

// Heads per CTA: more heads amortise the C.B^T prologue (~ one head's worth of time), fewer heads
// quantise better into waves of one CTA per SM.  Pick the divisor that minimises the modelled time.
static int pick_head_group(int heads_per_group, int ctas_per_head, int heads) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int best = 1;
  long long best_cost = -1;
  for (int s = 1; s <= kScanMaxHeads; ++s) {
    if (heads_per_group % s) continue;
    const long long ctas = (long long)ctas_per_head * (heads / s);
    const long long cost = ((ctas + sms - 1) / sms) * (s + 1);
    if (best_cost < 0 || cost <= best_cost) { best = s; best_cost = cost; }
  }
  return best;
}

}  // namespace ssd
// tensor-map builders shared with the GEMM (gemm.cu)
int make_tmap_f32_2d(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld);
int make_tmap_bf16_2d(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows);
int make_tmap_bf16_3d(CUtensorMap* map, const void* ptr, long long d2, long long d1, long long cols, long long ld, int box_rows, int box_cols);
}  // namespace mb
using namespace mb;
using namespace mb::ssd;

#define SSD_SHAPE_CHECK(name)                                                                         \
  MB_REQUIRE(L == kL && P == kP && N == kN, name ": tensor-core SSD kernels need chunk=256, headdim=64, " \
             "d_state=128 (got %d, %d, %d)", L, P, N);                                                \
  MB_REQUIRE(H % G == 0, name ": nheads %% ngroups != 0")

// states[b,c,h,p,n] = sum_t x[t,h,p] * exp(a_last - a_t) * dt_t * Bm[t,g,n]          (dstate = 0)
// dstates[b,c,h,p,n] = sum_t dy[t,h,p] * exp(a_t) * Cm[t,g,n]                        (dstate = 1)
extern "C" int mb_ssd_chunk_state(const void* arow, long long lda, const void* bop, long long ldb,
                                  const float* dt, const float* a, float* out, int dstate, int B, int T,
                                  int H, int G, int P, int N, int L, cudaStream_t stream) {
  SSD_SHAPE_CHECK("ssd_chunk_state");
  MB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "ssd_chunk_state: row strides must be multiples of 8");
  const int nc = cdiv(T, L);
  const int hpc = ((H / G) % 2 == 0) ? 2 : 1;
  dim3 grid(B * nc, H / hpc);
  CUtensorMap tmOut, tmBop;
  if (make_tmap_f32_2d(&tmOut, out, (long long)B * nc * H * P, N, N) ||
      make_tmap_bf16_3d(&tmBop, bop, B, T, (long long)G * N, ldb, kOuterHalf, 64)) return 1;
  if (dstate) {
    static bool once = false;
    if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_outer_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kOuterSmem)); once = true; }
    ssd_outer_kernel<true><<<grid, kThreads, kOuterSmem, stream>>>((const bf16*)arow, lda, tmBop, dt, a, tmOut, T, H, G, nc, hpc);
  } else {
    static bool once = false;
    if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_outer_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kOuterSmem)); once = true; }
    ssd_outer_kernel<false><<<grid, kThreads, kOuterSmem, stream>>>((const bf16*)arow, lda, tmBop, dt, a, tmOut, T, H, G, nc, hpc);
  }
  return check_launch("ssd_chunk_state");
}

// Whether the sequential chunk-state + state-passing kernel is the better choice for this shape: it needs an even
// number of heads per group and enough (batch, head pair) CTAs to occupy the GPU (two per SM).
extern "C" int mb_ssd_states_seq_ok(int B, int H, int G) {
  return ((H / G) % 2 == 0 && (long long)B * (H / 2) >= (long long)num_sms() * 3 / 2) ? 1 : 0;
}

// Forward (dstate = 0): sprev[b,c] = state entering chunk c (bf16), optional init / final fp32 [B,H,P,N].
// Backward (dstate = 1): arow = dy, bop = Cm; out = ds[b,c] (bf16 gradient of chunk c's own state),
//   da_last[b,c,h] += exp(a_last) <ds, sprev> (pre-zeroed by the caller), `init` = gradient of the final state
//   (or null), `last_out` = gradient of the initial state (or null).
extern "C" int mb_ssd_states_seq(const void* arow, long long lda, const void* bop, long long ldb, const float* dt,
                                 const float* a, void* out_bf16, const void* sprev_bf16, float* da_last,
                                 const float* init, float* last_out, int dstate, int B, int T, int H, int G, int P,
                                 int N, int L, cudaStream_t stream) {
  SSD_SHAPE_CHECK("ssd_states_seq");
  MB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "ssd_states_seq: row strides must be multiples of 8");
  MB_REQUIRE((H / G) % 2 == 0, "ssd_states_seq: needs an even number of heads per group");
  MB_REQUIRE(!dstate || (sprev_bf16 != nullptr && da_last != nullptr), "ssd_states_seq: backward needs sprev and da_last");
  const int nc = cdiv(T, L);
  dim3 grid(B, H / 2);
  CUtensorMap tmOut, tmBop;
  if (make_tmap_bf16_2d(&tmOut, out_bf16, (long long)B * nc * H * P, N, N, 32) ||
      make_tmap_bf16_3d(&tmBop, bop, B, T, (long long)G * N, ldb, kOuterHalf, 64)) return 1;
  if (dstate) {
    static bool once = false;
    if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_states_seq_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSeqSmem)); once = true; }
    MB_CUDA(launch_pdl(ssd_states_seq_kernel<true>, grid, dim3(kThreads), kSeqSmem, stream, (const bf16*)arow, lda, tmBop, dt, a,
                       tmOut, (const bf16*)sprev_bf16, da_last, init, last_out, T, H, G, nc));
  } else {
    static bool once = false;
    if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_states_seq_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSeqSmem)); once = true; }
    MB_CUDA(launch_pdl(ssd_states_seq_kernel<false>, grid, dim3(kThreads), kSeqSmem, stream, (const bf16*)arow, lda, tmBop, dt, a,
                       tmOut, (const bf16*)nullptr, (float*)nullptr, init, last_out, T, H, G, nc));
  }
  return check_launch("ssd_states_seq");
}

// y[t,h,:] = exp(a_t) C_t S_prev^T + sum_{s<=t} (C_t.B_s) exp(a_t - a_s) dt_s x_s + D_h x_t ;  y_off = first term only
extern "C" int mb_ssd_chunk_scan_fwd(const void* Cm, const void* Bm, long long ldbc, const void* x, long long ldx,
                                     const void* sprev_bf16, const float* dt, const float* a, const float* D,
                                     void* y, void* y_off, long long ldy, int B, int T, int H, int G, int P, int N, int L,
                                     cudaStream_t stream) {
  SSD_SHAPE_CHECK("ssd_chunk_scan_fwd");
  const int nc = cdiv(T, L);
  const int hgs = pick_head_group(H / G, B * nc * 2, H);  // <= kScanMaxHeads (smem holds a | dt of the group)
  static bool once = false;
  if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kScanSmem)); once = true; }
  dim3 grid(B * nc * 2, H / hgs);
  CUtensorMap tmRow, tmCol, tmV, tmS;
  if (make_tmap_bf16_3d(&tmRow, Cm, B, T, (long long)G * N, ldbc, 128, 64) || make_tmap_bf16_3d(&tmCol, Bm, B, T, (long long)G * N, ldbc, 128, 64) ||
      make_tmap_bf16_3d(&tmV, x, B, T, (long long)H * P, ldx, 128, 64) ||
      make_tmap_bf16_2d(&tmS, sprev_bf16, (long long)B * nc * H * P, N, N, 64)) return 1;
  CUtensorMap tmOut, tmOut2;
  if (make_tmap_bf16_3d(&tmOut, y, B, T, (long long)H * P, ldy, 32, 32) ||
      make_tmap_bf16_3d(&tmOut2, y_off, B, T, (long long)H * P, ldy, 32, 32)) return 1;
  MB_CUDA(launch_pdl(ssd_scan_kernel<false>, grid, dim3(kScanThreads), kScanSmem, stream, tmRow, tmCol, tmV, tmS, tmOut, tmOut2,
                     dt, a, D, (float*)nullptr, (float*)nullptr, (float*)nullptr, T, H, G, nc, hgs));
  return check_launch("ssd_chunk_scan_fwd");
}

// dx[s,h,:] = dt_s sum_{t>=s} (B_s.C_t) exp(a_t - a_s) dy_t + exp(a_last - a_s) dt_s B_s dS^T + D_h dy_s
// ddt1[s] += sum_p x (M-part / dt_s),  ddt2[s] += exp(a_last - a_s) sum_p x (B dS^T),  dD[h] += sum dy.x
// (dD, ddt1 and ddt2 are pre-zeroed by the caller: each gets two commutative fp32 atomic adds per element)
extern "C" int mb_ssd_chunk_scan_bwd_dx(const void* Bm, const void* Cm, long long ldbc, const void* dy, long long lddy,
                                        const void* ds_bf16, const float* dt, const float* a, const float* D,
                                        void* dx, long long lddx, const void* x, long long ldx, float* ddt1,
                                        float* ddt2, float* dD, int B, int T, int H, int G, int P, int N, int L,
                                        cudaStream_t stream) {
  SSD_SHAPE_CHECK("ssd_chunk_scan_bwd_dx");
  const int nc = cdiv(T, L);
  const int hgs = pick_head_group(H / G, B * nc * 2, H);
  static bool once = false;
  if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kScanSmem)); once = true; }
  dim3 grid(B * nc * 2, H / hgs);
  CUtensorMap tmRow, tmCol, tmV, tmS;
  if (make_tmap_bf16_3d(&tmRow, Bm, B, T, (long long)G * N, ldbc, 128, 64) || make_tmap_bf16_3d(&tmCol, Cm, B, T, (long long)G * N, ldbc, 128, 64) ||
      make_tmap_bf16_3d(&tmV, dy, B, T, (long long)H * P, lddy, 128, 64) ||
      make_tmap_bf16_2d(&tmS, ds_bf16, (long long)B * nc * H * P, N, N, 64)) return 1;
  CUtensorMap tmOut, tmX;  // the kernel's second output map is unused in backward: it carries x instead
  if (make_tmap_bf16_3d(&tmOut, dx, B, T, (long long)H * P, lddx, 32, 32) ||
      make_tmap_bf16_3d(&tmX, x, B, T, (long long)H * P, ldx, 32, 32)) return 1;
  MB_CUDA(launch_pdl(ssd_scan_kernel<true>, grid, dim3(kScanThreads), kScanSmem, stream, tmRow, tmCol, tmV, tmS, tmOut, tmX, dt, a,
                     D, ddt1, ddt2, dD, T, H, G, nc, hgs));
  return check_launch("ssd_chunk_scan_bwd_dx");
}
