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
constexpr int kOuterSmem = 2 * (kL * 2 * 128) + 2 * kL * 4 + 64 + 1024;

template <bool DSTATE>
__global__ void __launch_bounds__(kThreads, 1)
ssd_outer_kernel(const bf16* __restrict__ arow, long long lda, const bf16* __restrict__ bop, long long ldb,
                 const float* __restrict__ dtv, const float* __restrict__ av, float* __restrict__ out, int T,
                 int H, int G, int nc, int hpc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sA = smem;                       // [256 k-rows][2 atoms]  MN-major A
  uint8_t* sB = smem + kL * 2 * 128;        // [256 k-rows][2 atoms]  MN-major B
  float* s_scale = reinterpret_cast<float*>(smem + 2 * kL * 2 * 128);  // [2][256]
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_scale + 2 * kL);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bc = blockIdx.x;
  const int b = bc / nc, c = bc % nc;
  const int h0 = blockIdx.y * hpc;
  const int g = h0 / (H / G);
  const int valid = min(kL, T - c * kL);

  if (tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(tmem_slot, 128); tmem_relinquish(); }
  for (int idx = tid; idx < hpc * kL; idx += kThreads) {
    const int hh = idx / kL, t = idx % kL;
    const size_t base = ((size_t)bc * H + h0 + hh) * kL;
    const float a = av[base + t];
    s_scale[hh * kL + t] = DSTATE ? __expf(a) : __expf(av[base + kL - 1] - a) * dtv[base + t];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const size_t row0 = (size_t)b * T + (size_t)c * kL;
  auto scale = [&](int row, int blk) { return s_scale[blk * kL + row]; };
  load_tile<2, false>(sB, bop + row0 * ldb + g * kN, ldb, kL, valid, NoScale(), tid, kThreads);  // async
  if (hpc == 2) load_tile<2, true>(sA, arow + row0 * lda + h0 * kP, lda, kL, valid, scale, tid, kThreads);
  else load_tile<1, true>(sA, arow + row0 * lda + h0 * kP, lda, kL, valid, scale, tid, kThreads);
  cp_async_commit();
  cp_async_wait<0>();
  fence_proxy_async_smem();
  __syncthreads();

  if (tid == 0) {
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, kN, MAJOR_MN, MAJOR_MN);
    const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
#pragma unroll
    for (int k = 0; k < kL / 16; ++k)
      mma_bf16_ss(tmem, desc_mnmajor(a_addr, kL, k), desc_mnmajor(b_addr, kL, k), idesc, k > 0 ? 1u : 0u);
    mma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();

  const int lg = warp & 3, half = warp >> 2;
  const int r = lg * 32 + lane;
  const int hh = r / kP, p = r % kP;
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    const int col = half * 64 + ch * 32;
    uint32_t v[32];
    tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + col, v);
    tmem_ld_wait();
    if (hh < hpc) {
      float* dst = out + (((size_t)bc * H + h0 + hh) * kP + p) * kN + col;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(dst + 4 * j) =
            make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                        __uint_as_float(v[4 * j + 3]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

// ------------------------------------------------------------------------------------------
// scan kernel (FWD: y ; BWD: dx, ddt partials)
// ------------------------------------------------------------------------------------------
constexpr int kScanSmem = 128 * 2 * 128   /* sR  row operand  [128][2 kblk]            32 KB */
                        + kL * 2 * 128    /* sC  col operand  [256][2 kblk]; later G   64 KB */
                        + 2 * kL * 128    /* sV  X_h / dY_h   [256][1 atom]  x2 buffers 64 KB */
                        + 2 * kP * 2 * 128/* sS  S_prev_h / dS_h [64][2 kblk] x2       32 KB */
                        + 2 * 2 * kL * 4 + 64 + 1024;

template <bool BWD>
__global__ void __launch_bounds__(kThreads, 1)
ssd_scan_kernel(const bf16* __restrict__ rowop, const bf16* __restrict__ colop, long long ldbc,
                const bf16* __restrict__ vsrc, long long ldv, const bf16* __restrict__ ssrc,
                const float* __restrict__ dtv, const float* __restrict__ av, const float* __restrict__ Dp,
                bf16* __restrict__ outp, long long ldo, bf16* __restrict__ out2, const bf16* __restrict__ xsrc,
                long long ldx, float* __restrict__ ddt1, float* __restrict__ ddt2, float* __restrict__ dD, int T,
                int H, int G, int nc, int hgs) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sR = smem;
  uint8_t* sC = sR + 128 * 2 * 128;
  uint8_t* sG = sC;  // aliases the column operand once CB is in TMEM
  uint8_t* sV0 = sC + kL * 2 * 128;          // two buffers: the next head's tiles stream in (cp.async)
  uint8_t* sS0 = sV0 + 2 * kL * 128;         // while the current head is being scaled / multiplied
  float* s_ad = reinterpret_cast<float*>(sS0 + 2 * kP * 2 * 128);  // [2 buffers][a(256) | dt(256)]
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_ad + 2 * 2 * kL);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rh = blockIdx.x & 1;
  const int bc = blockIdx.x >> 1;
  const int b = bc / nc, c = bc % nc;
  const int h_first = blockIdx.y * hgs;
  const int g = h_first / (H / G);
  const int valid = min(kL, T - c * kL);        // valid tokens in this chunk
  const int r0 = rh * 128;                      // first row of this CTA
  const int c0 = BWD ? rh * 128 : 0;            // first column
  const int ncols = BWD ? (2 - rh) * 128 : (rh + 1) * 128;
  const size_t row_base = (size_t)b * T + (size_t)c * kL;

  if (tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tCB = tmem, tOff = tmem + 256, tDiag = tmem + 320;
  uint32_t phase = 0;

  // ---- CB (or its transpose) for this row half: [128 x ncols], K = d_state ----
  load_tile<2, false>(sR, rowop + (row_base + r0) * ldbc + g * kN, ldbc, 128, max(0, min(128, valid - r0)),
                      NoScale(), tid, kThreads);
  load_tile<2, false>(sC, colop + (row_base + c0) * ldbc + g * kN, ldbc, ncols, max(0, min(ncols, valid - c0)),
                      NoScale(), tid, kThreads);
  cp_async_commit();
  auto prefetch_head = [&](int hh) {
    const int h = h_first + hh;
    const int bsel = hh & 1;
    const size_t sb = ((size_t)bc * H + h) * kL;
    const uint32_t ad = smem_u32(s_ad + bsel * 2 * kL);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ad + tid * 4), "l"(av + sb + tid) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ad + (kL + tid) * 4), "l"(dtv + sb + tid) : "memory");
    load_tile<1, false>(sV0 + bsel * kL * 128, vsrc + (row_base + c0) * ldv + h * kP, ldv, ncols,
                        max(0, min(ncols, valid - c0)), NoScale(), tid, kThreads);
    load_tile<2, false>(sS0 + bsel * kP * 2 * 128, ssrc + ((size_t)bc * H + h) * kP * kN, kN, kP, kP, NoScale(), tid,
                        kThreads);
    cp_async_commit();
  };
  prefetch_head(0);
  cp_async_wait<1>();  // row / column operands have landed (head 0's tiles may still be in flight)
  fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, ncols, MAJOR_K, MAJOR_K);
    const uint32_t a_addr = smem_u32(sR), b_addr = smem_u32(sC);
#pragma unroll
    for (int k = 0; k < kN / 16; ++k)
      mma_bf16_ss(tCB, desc_kmajor(a_addr, 128, k >> 2, k & 3), desc_kmajor(b_addr, ncols, k >> 2, k & 3), idesc,
                  k > 0 ? 1u : 0u);
    mma_commit(bar);
  }
  mbar_wait(bar, phase);
  phase ^= 1;
  tc_fence_after();

  const int lg = warp & 3, half = warp >> 2;
  const int r = lg * 32 + lane;       // local row
  const int ri = r0 + r;              // row inside the chunk
  const uint32_t lane_sel = (uint32_t)(lg * 32) << 16;
  const int warp_row_min = r0 + lg * 32, warp_row_max = warp_row_min + 31;

  for (int hh = 0; hh < hgs; ++hh) {
    const int h = h_first + hh;
    const size_t sbase = ((size_t)bc * H + h) * kL;
    uint8_t* sV = sV0 + (hh & 1) * kL * 128;
    uint8_t* sS = sS0 + (hh & 1) * kP * 2 * 128;
    const float* s_a = s_ad + (hh & 1) * 2 * kL;
    const float* s_dt = s_a + kL;
    if (hh + 1 < hgs) { prefetch_head(hh + 1); cp_async_wait<1>(); }
    else cp_async_wait<0>();
    __syncthreads();

    // ---- G = CB o decay o mask (o dt_j in FWD) -> bf16 K-major operand ----
    const float a_i = s_a[ri];
    const int cols_per_thread = ncols >> 1;
    for (int k0 = half * cols_per_thread; k0 < (half + 1) * cols_per_thread; k0 += 32) {
      const int cj0 = c0 + k0;
      const bool skip = BWD ? (cj0 + 31 < warp_row_min) : (cj0 > warp_row_max);  // warp-uniform
      float f[32];
      if (!skip) {
        uint32_t v[32];
        tmem_ld32(tCB + lane_sel + k0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int cj = cj0 + j;
          float val;
          if (BWD) val = (cj >= ri) ? __uint_as_float(v[j]) * __expf(fminf(s_a[cj] - a_i, 0.f)) : 0.f;
          else val = (cj <= ri) ? __uint_as_float(v[j]) * __expf(fminf(a_i - s_a[cj], 0.f)) * s_dt[cj] : 0.f;
          f[j] = val;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = 0.f;
      }
      uint8_t* gblk = sG + (size_t)(k0 >> 6) * (128 * 128) + r * 128;
      const int ch0 = (k0 & 63) >> 3;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<bf16x8*>(gblk + (((ch0 + q) ^ (r & 7)) << 4)) = pack8(f + 8 * q);
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();

    if (tid == 0) {
      tc_fence_after();
      const uint32_t r_addr = smem_u32(sR), s_addr = smem_u32(sS), g_addr = smem_u32(sG), v_addr = smem_u32(sV);
      constexpr uint32_t idesc_off = make_idesc_bf16(128, kP, MAJOR_K, MAJOR_K);
      constexpr uint32_t idesc_diag = make_idesc_bf16(128, kP, MAJOR_K, MAJOR_MN);
#pragma unroll
      for (int k = 0; k < kN / 16; ++k)  // (C | B)[128 x 128] . (S_prev | dS)[64 x 128]^T
        mma_bf16_ss(tOff, desc_kmajor(r_addr, 128, k >> 2, k & 3), desc_kmajor(s_addr, kP, k >> 2, k & 3),
                    idesc_off, k > 0 ? 1u : 0u);
      for (int k = 0; k < ncols / 16; ++k)  // G[128 x ncols] . V[ncols x 64]
        mma_bf16_ss(tDiag, desc_kmajor(g_addr, 128, k >> 2, k & 3), desc_mnmajor(v_addr, ncols, k), idesc_diag,
                    k > 0 ? 1u : 0u);
      mma_commit(bar);
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    tc_fence_after();

    // ---- epilogue: this thread owns row ri, 32 of the 64 head-dim columns ----
    {
      const int p0 = half * 32;
      uint32_t vo[32], vd[32];
      tmem_ld32(tOff + lane_sel + p0, vo);
      tmem_ld32(tDiag + lane_sel + p0, vd);
      tmem_ld_wait();
      const float Dh = Dp[h];
      const int vrow = ri - c0;  // row of this token inside sV
      float xv[32];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        unpack8(*reinterpret_cast<const bf16x8*>(sV + vrow * 128 + ((((p0 >> 3) + q) ^ (vrow & 7)) << 4)), xv + 8 * q);
      float o[32];
      if (!BWD) {
        const float ea = __expf(a_i);
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = ea * __uint_as_float(vo[j]) + __uint_as_float(vd[j]);
        // the state-space part alone (without the D x skip) is kept for backward: d a_t needs
        // sum_p dy (y - D x), which would cancel catastrophically if formed from the rounded y
        if (ri < valid) {
          bf16* dst2 = out2 + (row_base + ri) * ldo + h * kP + p0;
#pragma unroll
          for (int q = 0; q < 4; ++q) st8(dst2 + 8 * q, pack8(o + 8 * q));
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] += Dh * xv[j];
      } else {
        const float dt_i = s_dt[ri];
        const float el = __expf(s_a[kL - 1] - a_i);
        const float w = el * dt_i;
        float s1 = 0.f, s2 = 0.f, sd = 0.f;
        if (ri < valid) {
          const bf16* xr = xsrc + (row_base + ri) * ldx + h * kP + p0;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float xx[8];
            unpack8(ld8(xr + 8 * q), xx);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              s1 += xx[j] * __uint_as_float(vd[8 * q + j]);
              s2 += xx[j] * __uint_as_float(vo[8 * q + j]);
              sd += xx[j] * xv[8 * q + j];  // dD_h = sum dy . x
            }
          }
          atomicAdd(ddt1 + sbase + ri, s1);
          atomicAdd(ddt2 + sbase + ri, el * s2);
        }
        sd = warp_sum(sd);
        if (lane == 0) atomicAdd(dD + h, sd);
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = dt_i * __uint_as_float(vd[j]) + w * __uint_as_float(vo[j]) + Dh * xv[j];
      }
      if (ri < valid) {
        bf16* dst = outp + (row_base + ri) * ldo + h * kP + p0;
#pragma unroll
        for (int q = 0; q < 4; ++q) st8(dst + 8 * q, pack8(o + 8 * q));
      }
    }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem, 512);
}

static int pick_head_group(int heads_per_group) {
  for (int s = 6; s >= 1; --s)
    if (heads_per_group % s == 0) return s;
  return 1;
}

}  // namespace ssd
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
  if (dstate) {
    static bool once = false;
    if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_outer_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kOuterSmem)); once = true; }
    ssd_outer_kernel<true><<<grid, kThreads, kOuterSmem, stream>>>((const bf16*)arow, lda, (const bf16*)bop, ldb, dt, a, out, T, H, G, nc, hpc);
  } else {
    static bool once = false;
    if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_outer_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kOuterSmem)); once = true; }
    ssd_outer_kernel<false><<<grid, kThreads, kOuterSmem, stream>>>((const bf16*)arow, lda, (const bf16*)bop, ldb, dt, a, out, T, H, G, nc, hpc);
  }
  return check_launch("ssd_chunk_state");
}

// y[t,h,:] = exp(a_t) C_t S_prev^T + sum_{s<=t} (C_t.B_s) exp(a_t - a_s) dt_s x_s + D_h x_t ;  y_ssm = y - D_h x_t
extern "C" int mb_ssd_chunk_scan_fwd(const void* Cm, const void* Bm, long long ldbc, const void* x, long long ldx,
                                     const void* sprev_bf16, const float* dt, const float* a, const float* D,
                                     void* y, void* y_ssm, long long ldy, int B, int T, int H, int G, int P, int N, int L,
                                     cudaStream_t stream) {
  SSD_SHAPE_CHECK("ssd_chunk_scan_fwd");
  const int nc = cdiv(T, L);
  const int hgs = pick_head_group(H / G);
  static bool once = false;
  if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kScanSmem)); once = true; }
  dim3 grid(B * nc * 2, H / hgs);
  ssd_scan_kernel<false><<<grid, kThreads, kScanSmem, stream>>>((const bf16*)Cm, (const bf16*)Bm, ldbc, (const bf16*)x, ldx,
                                                                (const bf16*)sprev_bf16, dt, a, D, (bf16*)y, ldy, (bf16*)y_ssm,
                                                                nullptr, 0, nullptr, nullptr, nullptr, T, H, G, nc, hgs);
  return check_launch("ssd_chunk_scan_fwd");
}

// dx[s,h,:] = dt_s sum_{t>=s} (B_s.C_t) exp(a_t - a_s) dy_t + exp(a_last - a_s) dt_s B_s dS^T + D_h dy_s
// ddt1[s] += sum_p x (M-part / dt_s),  ddt2[s] += exp(a_last - a_s) sum_p x (B dS^T),  dD[h] += sum dy.x
// (ddt1, ddt2, dD are pre-zeroed by the caller)
extern "C" int mb_ssd_chunk_scan_bwd_dx(const void* Bm, const void* Cm, long long ldbc, const void* dy, long long lddy,
                                        const void* ds_bf16, const float* dt, const float* a, const float* D,
                                        void* dx, long long lddx, const void* x, long long ldx, float* ddt1,
                                        float* ddt2, float* dD, int B, int T, int H, int G, int P, int N, int L,
                                        cudaStream_t stream) {
  SSD_SHAPE_CHECK("ssd_chunk_scan_bwd_dx");
  const int nc = cdiv(T, L);
  const int hgs = pick_head_group(H / G);
  static bool once = false;
  if (!once) { MB_CUDA(cudaFuncSetAttribute(ssd_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kScanSmem)); once = true; }
  dim3 grid(B * nc * 2, H / hgs);
  ssd_scan_kernel<true><<<grid, kThreads, kScanSmem, stream>>>((const bf16*)Bm, (const bf16*)Cm, ldbc, (const bf16*)dy, lddy,
                                                               (const bf16*)ds_bf16, dt, a, D, (bf16*)dx, lddx, nullptr,
                                                               (const bf16*)x, ldx, ddt1, ddt2, dD, T, H, G, nc, hgs);
  return check_launch("ssd_chunk_scan_bwd_dx");
}
