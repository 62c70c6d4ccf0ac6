// SSD (Mamba-2) tensor-core kernels, part 2 (backward only):
//   ssd_dcb_kernel    dCB[i,j] = sum_h (dY_h X_h^T)[i,j] * exp(a^h_i - a^h_j) * dt^h_j * [j <= i]
//                     per (batch, chunk, group); the per-head products live in TMEM, the decay
//                     scaling and the sum over heads happen in registers.
//   ssd_dbdc_kernel   dC = dCB . B + sum_h (e^{a_i} dY_h) S_prev_h
//                     dB = dCB^T . C + sum_h (w_j X_h) dS_h
//                     as ONE K-concatenated tcgen05 accumulation per 128-row block: the sum over
//                     heads is the K dimension of the MMA chain, so no cross-head reduction pass.
// They replace mamba_ssm's Triton _chunk_scan_bwd_dcb, _bmm_chunk_bwd, _chunk_scan_bwd_dC and
// _chunk_state_bwd_db (SURVEY.md section 2.3, K8/K9/K11/K12 backward variants).
#include "ssd_tc.cuh"
#include "../include/mamba_b200.h"

namespace mb {
namespace ssd {

constexpr int kThreadsB = 256;

constexpr int kDcbSmem = 2 * 2 * (128 * 128) + 2 * 2 * kL * 4 + 2 * 4 * 128 * 4 + 64 + 1024;  // double-buffered per-head tiles + column factors

__global__ void __launch_bounds__(kThreadsB, 2)
ssd_dcb_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
               const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmB,
               const float* __restrict__ dtv, const float* __restrict__ av, bf16* __restrict__ dcb,
               float* __restrict__ rowsum_m, int T, int H, int G, int nc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  float* s_ad = reinterpret_cast<float*>(smem + 2 * 2 * 128 * 128);  // [2][a(256) | dt(256)]
  float* s_cf = s_ad + 2 * 2 * kL;                                   // [4 lane groups][128]  exp(a_ref - a_j)
  float* s_cd = s_cf + 4 * 128;                                      // [4 lane groups][128]  exp(a_ref - a_j) dt_j
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_cd + 4 * 128);
  uint64_t* tile_bar = bar + 1;   // [2] TMA transaction barriers of the double-buffered per-head tiles
  uint64_t* ld_bar = bar + 3;     // C / B tiles of the prologue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // block 1 (rows 128.., columns 0..127) lies entirely below the diagonal and costs ~1.6x a diagonal block: all of
  // them are launched first so that the second, partial wave of CTAs consists of short ones only
  const int nbc = (int)gridDim.x / 3;
  const int blk = (int)blockIdx.x < nbc ? 1 : (((int)blockIdx.x - nbc) & 1) * 2;
  const int bc = (int)blockIdx.x < nbc ? (int)blockIdx.x : ((int)blockIdx.x - nbc) >> 1;
  const int b = bc / nc, c = bc % nc;
  const int g = blockIdx.y;
  const int hpg = H / G;
  const int ri0 = blk == 0 ? 0 : 128;
  const int cj0 = blk == 2 ? 128 : 0;
  const int valid = min(kL, T - c * kL);

  if (warp == 0 && elect_one()) {
    mbar_init(bar, 1); mbar_init(&tile_bar[0], 1); mbar_init(&tile_bar[1], 1); mbar_init(ld_bar, 1);
    fence_mbar_init();
    pdl_wait();   // the loads below read the previous kernel's output
    // C / B tiles for the CB block: two 64-column boxes each, straight away (they borrow the per-head buffers)
    mbar_arrive_expect_tx(ld_bar, 2 * 2 * 128 * 128);
    for (int kb = 0; kb < 2; ++kb) {
      tma_load_3d(smem + kb * 128 * 128, &tmC, ld_bar, g * kN + kb * 64, c * kL + ri0, b);
      tma_load_3d(smem + 2 * 128 * 128 + kb * 128 * 128, &tmB, ld_bar, g * kN + kb * 64, c * kL + cj0, b);
    }
  }
  if (warp == 0) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tCB = tmem + 128;
  uint32_t phase = 0;
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();

  // CB block (rows ri0.., columns cj0..) once per CTA: needed for the row sums of dM o M below.
  // The C / B tiles borrow the per-head double buffer (2 x 32 KB) before the head loop starts.
  {
    uint8_t* sCt = smem;
    uint8_t* sBt = smem + 2 * 128 * 128;
    if (warp == 0) {  // converged warp, one elected lane issues (ptx.cuh: warp-uniform single-lane issue)
      mbar_wait_warp(ld_bar, 0);
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc_bf16(128, 128, MAJOR_K, MAJOR_K);
      const uint32_t a_addr = smem_u32(sCt), b_addr = smem_u32(sBt);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < kN / 16; ++k)
          mma_bf16_ss(tCB, desc_kmajor(a_addr, 128, k >> 2, k & 3), desc_kmajor(b_addr, 128, k >> 2, k & 3), idesc,
                      k > 0 ? 1u : 0u);
        mma_commit(bar);
      }
      __syncwarp();
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    tc_fence_after();
    __syncthreads();  // everyone is past the wait before the tiles are overwritten by the head prefetch
  }

  const int lg = warp & 3, half = warp >> 2;
  const int r = lg * 32 + lane;
  const int ri = ri0 + r;
  const uint32_t lane_sel = (uint32_t)(lg * 32) << 16;
  const int warp_row_min = ri0 + lg * 32, warp_row_max = warp_row_min + 31;
  float2 acc2[32];  // dCB accumulators, column pairs (FFMA2 operands need aligned register pairs)
#pragma unroll
  for (int j = 0; j < 32; ++j) acc2[j] = f2(0.f, 0.f);

  auto prefetch_head = [&](int hh) {
    const int h = g * hpg + hh;
    const int bsel = hh & 1;
    const size_t sb = ((size_t)bc * H + h) * kL;
    const uint32_t ad = smem_u32(s_ad + bsel * 2 * kL);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ad + tid * 4), "l"(av + sb + tid) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ad + (kL + tid) * 4), "l"(dtv + sb + tid) : "memory");
    if (warp == 0 && elect_one()) {  // dY rows / X rows of this head by TMA: one 128 x 64 box each, 128B-swizzled
      uint8_t* t0 = smem + bsel * 2 * 128 * 128;
      mbar_arrive_expect_tx(&tile_bar[bsel], 2 * 128 * 128);
      tma_load_3d(t0, &tmDY, &tile_bar[bsel], h * kP, c * kL + ri0, b);
      tma_load_3d(t0 + 128 * 128, &tmX, &tile_bar[bsel], h * kP, c * kL + cj0, b);
    }
    cp_async_commit();
  };
  prefetch_head(0);
  for (int hh = 0; hh < hpg; ++hh) {
    const int h = g * hpg + hh;
    uint8_t* sDY = smem + (hh & 1) * 2 * 128 * 128;
    uint8_t* sX = sDY + 128 * 128;
    const float* s_a = s_ad + (hh & 1) * 2 * kL;
    const float* s_dt = s_a + kL;
    if (hh + 1 < hpg) { prefetch_head(hh + 1); cp_async_wait<1>(); }
    else cp_async_wait<0>();
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      mbar_wait_warp(&tile_bar[hh & 1], (hh >> 1) & 1);
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc_bf16(128, 128, MAJOR_K, MAJOR_K);
      const uint32_t a_addr = smem_u32(sDY), b_addr = smem_u32(sX);
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          mma_bf16_ss(tmem, desc_kmajor(a_addr, 128, 0, ks), desc_kmajor(b_addr, 128, 0, ks), idesc, ks > 0 ? 1u : 0u);
        mma_commit(bar);
      }
      __syncwarp();
    }
    {  // per (lane group, column) decay factors for the blocks strictly below the diagonal, where
       // exp(a_i - a_j) = exp(a_i - a_ref) exp(a_ref - a_j) with both factors <= 1 (a_ref = first row of the group)
      const int lgq = tid >> 6, col = (tid & 63) * 2;
      const float aref = s_a[ri0 + lgq * 32];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float cf = fast_exp(fminf(aref - s_a[cj0 + col + u], 0.f));
        s_cf[lgq * 128 + col + u] = cf;
        s_cd[lgq * 128 + col + u] = cf * s_dt[cj0 + col + u];
      }
    }
    __syncthreads();
    mbar_wait(bar, phase);
    phase ^= 1;
    tc_fence_after();
    const float a_i = s_a[ri];
    const float rowf = fast_exp(fminf(a_i - s_a[warp_row_min], 0.f));
    float2 rs2 = f2(0.f, 0.f);  // the same sum, even / odd columns of the blocks below the diagonal
    float rsum = 0.f;  // sum_j (dM o M)[ri, j] over this thread's columns, with M rounded exactly as
                       // the dx kernel rounds it (see ssd_bwd_finalize: makes row - column sums cancel)
#pragma unroll
    for (int ch = 0; ch < 4; ++ch) {  // 16 columns at a time: keeps (P, CB) staging at 32 registers
      const int k0 = half * 64 + ch * 16;
      if (cj0 + k0 > warp_row_max) continue;  // warp-uniform: block entirely above the diagonal
      uint32_t v[16], cbv[16];
      tmem_ld16(tmem + lane_sel + k0, v);
      tmem_ld16(tCB + lane_sel + k0, cbv);
      tmem_ld_wait();
      if (cj0 + k0 + 15 < warp_row_min) {  // warp-uniform: strictly below the diagonal, no mask, no exp
        // column pairs through the packed fp32 pipe (FMUL2 / FFMA2): ~4 issued instructions per element
        const float2 rf2 = f2(rowf, rowf);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 f4 = *reinterpret_cast<const float4*>(s_cf + lg * 128 + k0 + 4 * q);
          const float4 d4 = *reinterpret_cast<const float4*>(s_cd + lg * 128 + k0 + 4 * q);
          const float4 t4 = *reinterpret_cast<const float4*>(s_dt + cj0 + k0 + 4 * q);
          const float2 cf[2] = {f2(f4.x, f4.y), f2(f4.z, f4.w)}, cd[2] = {f2(d4.x, d4.y), f2(d4.z, d4.w)},
                       dj[2] = {f2(t4.x, t4.y), f2(t4.z, t4.w)};
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const int e0 = 4 * q + 2 * t;
            const float2 p = f2(__uint_as_float(v[e0]), __uint_as_float(v[e0 + 1]));
            const float2 cb = f2(__uint_as_float(cbv[e0]), __uint_as_float(cbv[e0 + 1]));
            acc2[(ch * 16 + e0) >> 1] = fma2(p, mul2(rf2, cd[t]), acc2[(ch * 16 + e0) >> 1]);
            const float2 m = mul2(cb, mul2(rf2, cf[t]));
            const float2 mr = __bfloat1622float2(__floats2bfloat162_rn(m.x, m.y));
            rs2 = fma2(mul2(p, dj[t]), mr, rs2);
          }
        }
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) {  // branch-free, 128-bit broadcast loads of a_j / dt_j
          const float4 a4 = *reinterpret_cast<const float4*>(s_a + cj0 + k0 + 4 * q);
          const float4 d4 = *reinterpret_cast<const float4*>(s_dt + cj0 + k0 + 4 * q);
          const float aj[4] = {a4.x, a4.y, a4.z, a4.w}, dj[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int cj = cj0 + k0 + 4 * q + t;
            const float e = fast_exp(fminf(a_i - aj[t], 0.f));
            const float pd = __uint_as_float(v[4 * q + t]) * dj[t];   // P * dt_j
            const unsigned keep = (cj <= ri) ? 0xffffffffu : 0u;
            {
              float2& a2 = acc2[(ch * 16 + 4 * q + t) >> 1];
              const float add = __uint_as_float(__float_as_uint(pd * e) & keep);
              if (t & 1) a2.y += add; else a2.x += add;
            }
            // M'[ri,cj] = bf16(CB * e), the operand value the dx kernel multiplies with
            const float mr = __bfloat162float(__float2bfloat16(__uint_as_float(cbv[4 * q + t]) * e));
            rsum += __uint_as_float(__float_as_uint(pd * mr) & keep);
          }
        }
      }
    }
    rsum += rs2.x + rs2.y;
    if (ri < valid) atomicAdd(rowsum_m + ((size_t)bc * H + h) * kL + ri, rsum);
    tc_fence_before();
    __syncthreads();
  }
  bf16* dst = dcb + (((size_t)bc * G + g) * kL + ri) * kL + cj0 + half * 64;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    bf16x8 o;
#pragma unroll
    for (int e = 0; e < 4; ++e) o.v[e] = __floats2bfloat162_rn(acc2[4 * q + e].x, acc2[4 * q + e].y);
    st8(dst + 8 * q, o);
  }
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------
constexpr int kDbdcStages = 3;
constexpr int kDbdcStage = 128 * 128 + 64 * 2 * 128 + 2048;  // A tile 16 KB + B tile 16 KB + a | dt | a_last of the rows
constexpr int kDbdcSmem = kDbdcStages * kDbdcStage + 64 + 1024;

// Software pipeline: every operand travels by cp.async into a 3-stage ring, two stages ahead of the MMA.
// The activation tiles need a per-row factor (e^{a_i} or w_j); each thread scales, in place, exactly the
// 16-byte chunks its own cp.async brought in, so that pass needs no barrier of its own.
template <bool IS_DB>
__global__ void __launch_bounds__(kThreadsB, 2)
ssd_dbdc_kernel(const __grid_constant__ CUtensorMap tmG,    // dCB   [B*nc*G*256][256], box 64 x 64
                const __grid_constant__ CUtensorMap tmBC,   // dC: Bm ; dB: Cm   [B][T][G*N], box 64 x 64
                const __grid_constant__ CUtensorMap tmAct,  // dC: dy ; dB: x    [B][T][H*P], box 64 x 64
                const __grid_constant__ CUtensorMap tmSt,   // dC: S_prev bf16 ; dB: dS bf16  [B*nc*H*P][N], box 64 x 64
                const float* __restrict__ dtv, const float* __restrict__ av,
                bf16* __restrict__ outp, long long ldo,             // dC / dB column base of group 0 in dxBC
                int T, int H, int G, int nc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kDbdcStages * kDbdcStage);   // MMAs of a stage retired
  uint64_t* full = bars + kDbdcStages;                                             // TMA bytes of a stage landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(full + kDbdcStages);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rb = blockIdx.x & 1;
  const int bc = blockIdx.x >> 1;
  const int b = bc / nc, c = bc % nc;
  const int g = blockIdx.y;
  const int hpg = H / G;
  const int m0 = rb * 128;  // output row block (i for dC, j for dB)
  const int valid = min(kL, T - c * kL);
  const size_t row_base = (size_t)b * T + (size_t)c * kL;
  const int g_row0 = (bc * G + g) * kL;   // first row of this (b, c, g) block in the dCB matrix

  if (warp == 0 && elect_one()) {
    for (int i = 0; i < kDbdcStages; ++i) { mbar_init(&bars[i], 1); mbar_init(&full[i], 1); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(tmem_slot, 128); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();

  // part 1: K = j in [0, (rb+1)*128) for dC ; K = i in [rb*128, 256) for dB   (64 per stage)
  const int k_begin = IS_DB ? m0 : 0;
  const int k_end = IS_DB ? kL : m0 + 128;
  const int n_part1 = (k_end - k_begin) / 64;
  const int n_stages = n_part1 + hpg;
  const int valid_m = max(0, min(128, valid - m0));

  // every tile travels by TMA (one lane issues 4 boxes of 64 x 64 per operand pair and stage); only the 1 KB of
  // per-row scale factors of the activation stages is copied by the threads themselves
  auto issue = [&](int s) {
    uint8_t* sA = smem + (s % kDbdcStages) * kDbdcStage;
    uint8_t* sB = sA + 128 * 128;
    uint64_t* fb = &full[s % kDbdcStages];
    if (s < n_part1) {
      if (warp == 0 && elect_one()) {
        const int kk = k_begin + s * 64;
        mbar_arrive_expect_tx(fb, 2 * 128 * 128);
        for (int j2 = 0; j2 < 2; ++j2) {
          if (!IS_DB)  // A = G[m0 + r, kk .. kk+63]  (K-major, 128 rows as two 64-row boxes)
            tma_load_2d(sA + j2 * 64 * 128, &tmG, fb, kk, g_row0 + m0 + j2 * 64);
          else         // A = G[kk + krow, m0 .. m0+127]  (MN-major: 64 k-rows x 2 atoms of 64 columns)
            tma_load_2d(sA + j2 * 64 * 128, &tmG, fb, m0 + j2 * 64, g_row0 + kk);
          // B = (Bm | Cm)[kk + krow, n]  (MN-major, 64 k-rows x 2 atoms); rows past the sequence end read as zeros
          tma_load_3d(sB + j2 * 64 * 128, &tmBC, fb, g * kN + j2 * 64, c * kL + kk, b);
        }
      }
    } else {
      const int h = g * hpg + (s - n_part1);
      const size_t sbase = ((size_t)bc * H + h) * kL;
      if (warp == 0 && elect_one()) {
        mbar_arrive_expect_tx(fb, 2 * 128 * 128);
        for (int j2 = 0; j2 < 2; ++j2) {
          tma_load_3d(sA + j2 * 64 * 128, &tmAct, fb, h * kP, c * kL + m0 + j2 * 64, b);   // [128 rows][64] K-major
          tma_load_2d(sB + j2 * 64 * 128, &tmSt, fb, j2 * 64, (bc * H + h) * kP);          // [64 p][128 n] MN-major
        }
      }
      const uint32_t sc = smem_u32(sB + 64 * 2 * 128);
      if (tid < 128) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sc + tid * 4), "l"(av + sbase + m0 + tid) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sc + 512 + tid * 4), "l"(dtv + sbase + m0 + tid) : "memory");
      } else if (tid == 128) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sc + 1024), "l"(av + sbase + kL - 1) : "memory");
      }
    }
  };

  issue(0);
  cp_async_commit();
  if (n_stages > 1) issue(1);
  cp_async_commit();
  for (int s = 0; s < n_stages; ++s) {
    const int buf = s % kDbdcStages;
    if (s + 2 < n_stages) {
      // the ring slot of stage s+2 was last read by the MMAs of stage s-1
      if (s >= 1) mbar_wait(&bars[(s + 2) % kDbdcStages], ((s - 1) / kDbdcStages) & 1);
      issue(s + 2);
    }
    cp_async_commit();   // (possibly empty) keeps one group per stage
    cp_async_wait<2>();  // this thread's scale-vector copies of stage s have landed
    mbar_wait(&full[buf], (s / kDbdcStages) & 1);  // ... and so have the stage's TMA tiles
    uint8_t* sA = smem + buf * kDbdcStage;
    uint8_t* sB = sA + 128 * 128;
    const bool a_mn = IS_DB && s < n_part1;
    if (s >= n_part1) {
      // the scale vectors were copied by other threads: one barrier for them, then in-place scaling, 4 chunks each
      __syncthreads();
      const float* s_a = reinterpret_cast<const float*>(sB + 64 * 2 * 128);
      const float a_last = s_a[256];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int idx = tid + k * kThreadsB;
        const int cc = idx & 7, row = idx >> 3;
        if (row < valid_m) {
          const float f = IS_DB ? fast_exp(a_last - s_a[row]) * s_a[128 + row] : fast_exp(s_a[row]);
          uint8_t* ptr = sA + row * 128 + ((cc ^ (row & 7)) << 4);
          bf16x8 t8;
          memcpy(&t8, ptr, 16);
          float v[8];
          unpack8(t8, v);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] *= f;
          t8 = pack8(v);
          memcpy(ptr, &t8, 16);
        }
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      tc_fence_after();
      const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
      const uint32_t idesc = make_idesc_bf16(128, kN, a_mn ? MAJOR_MN : MAJOR_K, MAJOR_MN);
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t da = a_mn ? desc_mnmajor(a_addr, 64, ks) : desc_kmajor(a_addr, 128, 0, ks);
          mma_bf16_ss(tmem, da, desc_mnmajor(b_addr, 64, ks), idesc, (s > 0 || ks > 0) ? 1u : 0u);
        }
        mma_commit(&bars[buf]);
      }
      __syncwarp();
    }
  }
  {
    const int s_last = n_stages - 1;
    mbar_wait(&bars[s_last % kDbdcStages], (s_last / kDbdcStages) & 1);  // MMAs retire in order: the last commit covers all
    tc_fence_after();
  }
  const int lg = warp & 3, half = warp >> 2;
  const int r = lg * 32 + lane;
  const int mi = m0 + r;
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int col = half * 64 + ch * 32;
    uint32_t v[32];
    tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + col, v);
    tmem_ld_wait();
    if (mi < valid) {
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
      bf16* dst = outp + (row_base + mi) * ldo + g * kN + col;
#pragma unroll
      for (int q = 0; q < 4; ++q) st8(dst + 8 * q, pack8(f + 8 * q));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

}  // namespace ssd
int make_tmap_bf16_2d(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows);
int make_tmap_bf16_3d(CUtensorMap* map, const void* ptr, long long d2, long long d1, long long cols, long long ld, int box_rows, int box_cols);
}  // namespace mb
using namespace mb;
using namespace mb::ssd;

// rowsum_m[b,c,h,i] += sum_j (dM o M)[i,j]  (pre-zeroed by the caller), the intra-chunk part of d a_i
extern "C" int mb_ssd_bwd_dcb(const void* dy, long long lddy, const void* x, long long ldx, const void* Cm,
                              const void* Bm, long long ldbc, const float* dt, const float* a, void* dcb,
                              float* rowsum_m, int B, int T, int H, int G, int P, int N, int L,
                              cudaStream_t stream) {
  MB_REQUIRE(L == kL && P == kP && N == kN, "ssd_bwd_dcb: needs chunk=256, headdim=64, d_state=128");
  const int nc = cdiv(T, L);
  static bool once = false;
  if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_dcb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDcbSmem)); once = true; }
  dim3 grid(B * nc * 3, G);
  CUtensorMap tmDY, tmX, tmC, tmB;
  if (make_tmap_bf16_3d(&tmDY, dy, B, T, (long long)H * P, lddy, 128, 64) || make_tmap_bf16_3d(&tmX, x, B, T, (long long)H * P, ldx, 128, 64) ||
      make_tmap_bf16_3d(&tmC, Cm, B, T, (long long)G * N, ldbc, 128, 64) || make_tmap_bf16_3d(&tmB, Bm, B, T, (long long)G * N, ldbc, 128, 64))
    return 1;
  MB_CUDA(launch_pdl(ssd_dcb_kernel, grid, dim3(kThreadsB), kDcbSmem, stream, tmDY, tmX, tmC, tmB, dt, a, (bf16*)dcb, rowsum_m,
                     T, H, G, nc));
  return check_launch("ssd_bwd_dcb");
}

extern "C" int mb_ssd_bwd_dbdc(const void* dcb, const void* Bm, const void* Cm, long long ldbc, const void* dy,
                               long long lddy, const void* x, long long ldx, const void* sprev_bf16,
                               const void* ds_bf16, const float* dt, const float* a, void* dB, void* dC,
                               long long ldo, int B, int T, int H, int G, int P, int N, int L,
                               cudaStream_t stream) {
  MB_REQUIRE(L == kL && P == kP && N == kN, "ssd_bwd_dbdc: needs chunk=256, headdim=64, d_state=128");
  const int nc = cdiv(T, L);
  static bool once = false;
  if (!once) {
    MB_CUDA(cudaFuncSetAttribute(ssd_dbdc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDbdcSmem));
    MB_CUDA(cudaFuncSetAttribute(ssd_dbdc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDbdcSmem));
    once = true;
  }
  dim3 grid(B * nc * 2, G);
  CUtensorMap tmG, tmBm, tmCm, tmDy, tmX, tmSp, tmDs;
  if (make_tmap_bf16_2d(&tmG, dcb, (long long)B * nc * G * L, L, L, 64) ||
      make_tmap_bf16_3d(&tmBm, Bm, B, T, (long long)G * N, ldbc, 64, 64) || make_tmap_bf16_3d(&tmCm, Cm, B, T, (long long)G * N, ldbc, 64, 64) ||
      make_tmap_bf16_3d(&tmDy, dy, B, T, (long long)H * P, lddy, 64, 64) || make_tmap_bf16_3d(&tmX, x, B, T, (long long)H * P, ldx, 64, 64) ||
      make_tmap_bf16_2d(&tmSp, sprev_bf16, (long long)B * nc * H * P, N, N, 64) ||
      make_tmap_bf16_2d(&tmDs, ds_bf16, (long long)B * nc * H * P, N, N, 64))
    return 1;
  // dB or dC may be null: the two kernels are independent, a caller can put them on different streams
  if (dC != nullptr) {
    MB_CUDA(launch_pdl(ssd_dbdc_kernel<false>, grid, dim3(kThreadsB), kDbdcSmem, stream, tmG, tmBm, tmDy, tmSp, dt, a,
                       (bf16*)dC, ldo, T, H, G, nc));
    if (check_launch("ssd_bwd_dbdc/dC")) return 1;
  }
  if (dB != nullptr) {
    MB_CUDA(launch_pdl(ssd_dbdc_kernel<true>, grid, dim3(kThreadsB), kDbdcSmem, stream, tmG, tmCm, tmX, tmDs, dt, a,
                       (bf16*)dB, ldo, T, H, G, nc));
    if (check_launch("ssd_bwd_dbdc/dB")) return 1;
  }
  return 0;
}
