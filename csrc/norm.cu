// K14 / K15: fused residual-add + RMSNorm and gated RMSNorm, forward and backward.
//
// Replaces the Triton layer-norm kernels the reference reaches through mamba_ssm
// (SURVEY.md section 2.3 rows K14, K15).  Memory-bound: one warp owns one row (or one
// (row, group) for the gated norm), the whole row lives in registers between the statistic
// and the scale pass, all global traffic is 16-byte vectors.  Backward kernels are
// grid-persistent so the weight gradient is reduced in registers across rows; per-block
// partials go to a scratch buffer and a second kernel folds them into the fp32 grad arena.
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kWarpsPerBlock = 4;

// ------------------------------------------------------------------------------------------
// add + rmsnorm forward.  x: bf16 or fp32 [M, D]; residual: fp32 [M, D] or null.
// ------------------------------------------------------------------------------------------
template <int NV, bool X_FP32>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
add_rmsnorm_fwd_kernel(const void* __restrict__ xin, const float* __restrict__ residual,
                       const float* __restrict__ weight, bf16* __restrict__ y,
                       float* __restrict__ res_out, float* __restrict__ rstd_out, int M, int D,
                       float eps) {
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= M) return;
  const int nvec = D >> 3;
  float v[NV][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = lane + 32 * i;
    if (vec < nvec) {
      const size_t off = (size_t)row * D + vec * 8;
      if (X_FP32) ld8f(reinterpret_cast<const float*>(xin) + off, v[i]);
      else unpack8(ld8(reinterpret_cast<const bf16*>(xin) + off), v[i]);
      if (residual != nullptr) {
        float r[8];
        ld8f(residual + off, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] += r[j];
      }
      if (res_out != nullptr) st8f(res_out + off, v[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += v[i][j] * v[i][j];
    }
  }
  ss = warp_sum(ss);
  const float rstd = rsqrtf(ss / (float)D + eps);
  if (lane == 0 && rstd_out != nullptr) rstd_out[row] = rstd;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = lane + 32 * i;
    if (vec < nvec) {
      float w[8], o[8];
      ld8f(weight + vec * 8, w);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = v[i][j] * rstd * w[j];
      st8(y + (size_t)row * D + vec * 8, pack8(o));
    }
  }
}

// ------------------------------------------------------------------------------------------
// add + rmsnorm backward.
//   r     fp32 [M, D]  the normalised input saved by forward (x + residual)
//   dy    bf16 [M, D]
//   dres_in  fp32 [M, D] or null : gradient arriving on the residual stream from later layers
//   dres_out fp32 [M, D]         : total gradient wrt r (may alias dres_in)
//   dx_bf16  bf16 [M, D] or null : same values in bf16 (operand of the previous out_proj dgrad)
//   dw_partial fp32 [gridDim.x, D] scratch
// ------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
add_rmsnorm_bwd_kernel(const float* __restrict__ r, const bf16* __restrict__ dy,
                       const float* __restrict__ rstd_in, const float* __restrict__ weight,
                       const float* dres_in, float* dres_out, bf16* __restrict__ dx_bf16,
                       float* __restrict__ dw_partial, int M, int D) {
  __shared__ float red[kWarpsPerBlock][32 * 8 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nvec = D >> 3;
  float w[NV][8], dw[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = lane + 32 * i;
#pragma unroll
    for (int j = 0; j < 8; ++j) { dw[i][j] = 0.f; w[i][j] = 0.f; }
    if (vec < nvec) ld8f(weight + vec * 8, w[i]);
  }
  for (int row = blockIdx.x * kWarpsPerBlock + warp; row < M; row += gridDim.x * kWarpsPerBlock) {
    const float rstd = rstd_in[row];
    float xh[NV][8], wdy[NV][8];
    float c = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      if (vec < nvec) {
        const size_t off = (size_t)row * D + vec * 8;
        float g[8];
        ld8f(r + off, xh[i]);
        unpack8(ld8(dy + off), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] *= rstd;
          wdy[i][j] = g[j] * w[i][j];
          dw[i][j] += g[j] * xh[i][j];
          c += xh[i][j] * wdy[i][j];
        }
      }
    }
    c = warp_sum(c) / (float)D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      if (vec < nvec) {
        const size_t off = (size_t)row * D + vec * 8;
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (wdy[i][j] - xh[i][j] * c) * rstd;
        if (dres_in != nullptr) {
          float a[8];
          ld8f(dres_in + off, a);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += a[j];
        }
        st8f(dres_out + off, o);
        if (dx_bf16 != nullptr) st8(dx_bf16 + off, pack8(o));
      }
    }
  }
  // block-level reduction of dw, one 256-column slab at a time
#pragma unroll
  for (int i = 0; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = dw[i][j];
    __syncthreads();
    for (int k = threadIdx.x; k < 256; k += blockDim.x) {
      const int col = i * 256 + k;
      if (col < D) {
        float s = 0.f;
#pragma unroll
        for (int ww = 0; ww < kWarpsPerBlock; ++ww) s += red[ww][k];
        dw_partial[(size_t)blockIdx.x * D + col] = s;
      }
    }
    __syncthreads();
  }
}

// dw[col] (+)= sum_p partial[p][col].   Block = 32 columns x 32 row-lanes; rows strided by 32.
__global__ void colsum_partials_kernel(const float* __restrict__ partial, float* __restrict__ dw,
                                       int P, int D, int accumulate) {
  __shared__ float red[32][33];
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + tx;
  float s = 0.f;
  if (col < D)
    for (int p = ty; p < P; p += 32) s += partial[(size_t)p * D + col];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && col < D) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) t += red[i][tx];
    dw[col] = accumulate ? dw[col] + t : t;
  }
}

// ------------------------------------------------------------------------------------------
// gated rmsnorm: out = rmsnorm_group(x * silu(z)) * w.   rows = M * G, row length gs = D / G.
// ------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
gated_rmsnorm_fwd_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ z, int ldz,
                         const float* __restrict__ weight, bf16* __restrict__ out, int ldo,
                         float* __restrict__ rstd_out, int M, int G, int gs, float eps) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= M * G) return;
  const int m = row / G, g = row - m * G;
  const int nvec = gs >> 3;
  float2 v[NV][4];   // column pairs: FMUL2 / FFMA2
  float2 ss2 = f2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = lane + 32 * i;
    if (vec < nvec) {
      const int col = g * gs + vec * 8;
      const bf16x8 px = ld8(x + (size_t)m * ldx + col), pz = ld8(z + (size_t)m * ldz + col);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[i][j] = mul2(__bfloat1622float2(px.v[j]), silu2(__bfloat1622float2(pz.v[j])));
        ss2 = fma2(v[i][j], v[i][j], ss2);
      }
    }
  }
  const float ss = warp_sum(ss2.x + ss2.y);
  const float rstd = rsqrtf(ss / (float)gs + eps);
  if (lane == 0) rstd_out[row] = rstd;
  const float2 rs2 = f2(rstd, rstd);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int vec = lane + 32 * i;
    if (vec < nvec) {
      const int col = g * gs + vec * 8;
      const float4 w0 = *reinterpret_cast<const float4*>(weight + col), w1 = *reinterpret_cast<const float4*>(weight + col + 4);
      const float2 w[4] = {f2(w0.x, w0.y), f2(w0.z, w0.w), f2(w1.x, w1.y), f2(w1.z, w1.w)};
      bf16x8 o;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 r = mul2(mul2(v[i][j], rs2), w[j]);
        o.v[j] = __floats2bfloat162_rn(r.x, r.y);
      }
      st8(out + (size_t)m * ldo + col, o);
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
gated_rmsnorm_bwd_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ z, int ldz,
                         const float* __restrict__ weight, const float* __restrict__ rstd_in,
                         const bf16* __restrict__ dout, int ldo, bf16* __restrict__ dx, int lddx,
                         bf16* __restrict__ dz, int lddz, float* __restrict__ dw_partial, int M,
                         int G, int gs) {
  __shared__ float red[kWarpsPerBlock][32 * 8 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nvec = gs >> 3;
  const int D = G * gs;
  // dw accumulators are per group-column; a warp visits rows of a single group index only when
  // G == 1.  For G > 1 the column base changes per row, so dw is accumulated through atomics
  // into the partial slab instead of registers (G > 1 is not used by the benchmark configs).
  float dw[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dw[i][j] = 0.f;
  float* my_partial = dw_partial + (size_t)blockIdx.x * D;
  for (int row = blockIdx.x * kWarpsPerBlock + warp; row < M * G; row += gridDim.x * kWarpsPerBlock) {
    const int m = row / G, g = row - m * G;
    const float rstd = rstd_in[row];
    // inputs stay packed (bf16) in registers between the two passes; the fp32 intermediates
    // are recomputed in pass 2 — keeps the kernel spill-free at row length 1536 (NV = 6)
    bf16x8 px[NV], pz[NV], pg[NV];
    float c = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      if (vec < nvec) {
        const int col = g * gs + vec * 8;
        float xv[8], zz[8], go[8], w[8];
        px[i] = ld8(x + (size_t)m * ldx + col);
        pz[i] = ld8(z + (size_t)m * ldz + col);
        pg[i] = ld8(dout + (size_t)m * ldo + col);
        unpack8(px[i], xv); unpack8(pz[i], zz); unpack8(pg[i], go);
        ld8f(weight + col, w);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = xv[j] * silu_f(zz[j]) * rstd;
          c += xh * go[j] * w[j];
          if (G == 1) dw[i][j] += go[j] * xh;
          else atomicAdd(my_partial + col + j, go[j] * xh);
        }
      }
    }
    c = warp_sum(c) / (float)gs;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      if (vec < nvec) {
        const int col = g * gs + vec * 8;
        float xv[8], zz[8], go[8], w[8], ox[8], oz[8];
        unpack8(px[i], xv); unpack8(pz[i], zz); unpack8(pg[i], go);
        ld8f(weight + col, w);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float sg, dsg;
          silu_both(zz[j], sg, dsg);
          const float xh = xv[j] * sg * rstd;
          const float dxs = (go[j] * w[j] - xh * c) * rstd;
          ox[j] = dxs * sg;
          oz[j] = dxs * xv[j] * dsg;
        }
        st8(dx + (size_t)m * lddx + col, pack8(ox));
        st8(dz + (size_t)m * lddz + col, pack8(oz));
      }
    }
  }
  if (G == 1) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = dw[i][j];
      __syncthreads();
      for (int k = threadIdx.x; k < 256; k += blockDim.x) {
        const int col = i * 256 + k;
        if (col < D) {
          float s = 0.f;
#pragma unroll
          for (int ww = 0; ww < kWarpsPerBlock; ++ww) s += red[ww][k];
          my_partial[col] = s;
        }
      }
      __syncthreads();
    }
  }
}


// Gated RMSNorm backward, G == 1 fast path: persistent warps, one row at a time, two passes over the row.
// The second pass re-reads the row (9 KB per warp, L1-resident) instead of parking 72 registers of packed
// inputs; that leaves room for the 8 x NV weight-gradient accumulators a lane owns (its columns are the
// same for every row), so dw costs one FFMA per element here and no second sweep over x / z / dout.
// Per-block partials go to dw_partial[blockIdx][D]; colsum_partials_kernel folds them.
template <int NV>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
gated_rmsnorm_bwd_rows_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ z, int ldz,
                              const float* __restrict__ weight, const float* __restrict__ rstd_in,
                              const bf16* __restrict__ dout, int ldo, bf16* __restrict__ dx, int lddx,
                              bf16* __restrict__ dz, int lddz, float* __restrict__ dw_partial, int M, int D) {
  __shared__ float red[kWarpsPerBlock][32 * 8 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nvec = D >> 3;
  // column pairs in float2 throughout: FMUL2 / FFMA2 halve the issue slots of the arithmetic
  float2 dw[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dw[i][j] = f2(0.f, 0.f);
  for (int row = blockIdx.x * kWarpsPerBlock + warp; row < M; row += gridDim.x * kWarpsPerBlock) {
    const float rstd = rstd_in[row];
    float2 c2 = f2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      if (vec < nvec) {
        const int col = vec * 8;
        const bf16x8 px = ld8(x + (size_t)row * ldx + col), pz = ld8(z + (size_t)row * ldz + col),
                     pg = ld8(dout + (size_t)row * ldo + col);
        const float4 w0 = *reinterpret_cast<const float4*>(weight + col), w1 = *reinterpret_cast<const float4*>(weight + col + 4);
        const float2 w[4] = {f2(w0.x, w0.y), f2(w0.z, w0.w), f2(w1.x, w1.y), f2(w1.z, w1.w)};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 xv = __bfloat1622float2(px.v[j]), zz = __bfloat1622float2(pz.v[j]), go = __bfloat1622float2(pg.v[j]);
          c2 = fma2(mul2(xv, silu2(zz)), mul2(go, w[j]), c2);
        }
      }
    }
    const float c = warp_sum(c2.x + c2.y) * rstd / (float)D;  // mean(xhat * w * dout)
    const float2 rs2 = f2(rstd, rstd), nc2 = f2(-c, -c);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      if (vec < nvec) {
        const int col = vec * 8;
        const bf16x8 px = ld8(x + (size_t)row * ldx + col), pz = ld8(z + (size_t)row * ldz + col),
                     pg = ld8(dout + (size_t)row * ldo + col);
        const float4 w0 = *reinterpret_cast<const float4*>(weight + col), w1 = *reinterpret_cast<const float4*>(weight + col + 4);
        const float2 w[4] = {f2(w0.x, w0.y), f2(w0.z, w0.w), f2(w1.x, w1.y), f2(w1.z, w1.w)};
        bf16x8 ox, oz;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 xv = __bfloat1622float2(px.v[j]), zz = __bfloat1622float2(pz.v[j]), go = __bfloat1622float2(pg.v[j]);
          const float2 sig = sigmoid2(zz);
          const float2 sg = mul2(zz, sig);                                                   // silu(z)
          const float2 dsg = mul2(sig, add2(f2(1.f, 1.f), fma2(f2(-1.f, -1.f), sg, zz)));   // silu'(z)
          const float2 xh = mul2(mul2(xv, sg), rs2);
          const float2 dxs = mul2(fma2(xh, nc2, mul2(go, w[j])), rs2);                       // (go*w - xh*c) * rstd
          const float2 o1 = mul2(dxs, sg), o2 = mul2(mul2(dxs, xv), dsg);
          ox.v[j] = __floats2bfloat162_rn(o1.x, o1.y);
          oz.v[j] = __floats2bfloat162_rn(o2.x, o2.y);
          dw[i][j] = fma2(go, xh, dw[i][j]);
        }
        st8(dx + (size_t)row * lddx + col, ox);
        st8(dz + (size_t)row * lddz + col, oz);
      }
    }
  }
  float* my_partial = dw_partial + (size_t)blockIdx.x * D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) { red[warp][lane * 8 + 2 * j] = dw[i][j].x; red[warp][lane * 8 + 2 * j + 1] = dw[i][j].y; }
    __syncthreads();
    for (int k = threadIdx.x; k < 256; k += blockDim.x) {
      const int col = i * 256 + k;
      if (col < D) {
        float sum = 0.f;
#pragma unroll
        for (int ww = 0; ww < kWarpsPerBlock; ++ww) sum += red[ww][k];
        my_partial[col] = sum;
      }
    }
    __syncthreads();
  }
}

// Same backward, rows staged through shared memory by cp.async, one row ahead.  The kernel above pays the global
// latency twice per row (statistic pass, then the re-read for the scale pass) with nothing else in flight for that
// warp — ncu: 8 long-scoreboard stall cycles per issued instruction, 48 % of the DRAM rate.  Here every lane copies
// its own 16-byte chunks of x, z and dout of the NEXT row into its warp's second stage before it starts on the
// current one, both passes read shared memory, and the weight vector sits in shared memory once per block.
// A lane only ever reads back the chunks it copied itself, so cp.async.wait_group is the only synchronisation.
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
// Gated RMSNorm forward (G == 1) with the same one-row-ahead cp.async staging: the one-row-per-warp kernel above has
// nothing in flight for a warp between its loads landing and its stores leaving (ncu: 54 % of the DRAM rate at 39 %
// occupancy); here a persistent warp always has the next row of x and z on the way while it runs the SiLU / reduce
// / scale of the current one.
template <int NV, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
gated_rmsnorm_fwd_staged_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ z, int ldz,
                                const float* __restrict__ weight, bf16* __restrict__ out, int ldo,
                                float* __restrict__ rstd_out, int M, int D, float eps) {
  extern __shared__ __align__(16) uint8_t st_smem[];
  constexpr int ROWB = NV * 32 * 16;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* my = st_smem + (size_t)warp * (2 * 2 * ROWB);
  float* s_w = reinterpret_cast<float*>(st_smem + (size_t)WARPS * 2 * 2 * ROWB);
  for (int i = threadIdx.x; i < D; i += WARPS * 32) s_w[i] = weight[i];   // a parameter: nothing the previous kernel wrote
  __syncthreads();
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  auto issue = [&](int row, int stage) {
    uint8_t* base = my + stage * 2 * ROWB;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (lane + 32 * i) * 8;
      cp_async_16(base + (lane + 32 * i) * 16, x + (size_t)row * ldx + col);
      cp_async_16(base + ROWB + (lane + 32 * i) * 16, z + (size_t)row * ldz + col);
    }
  };
  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + warp;
  if (row < M) issue(row, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int stage = 0; row < M; row += stride, stage ^= 1) {
    if (row + stride < M) issue(row + stride, stage ^ 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    const uint8_t* base = my + stage * 2 * ROWB;
    float2 v[NV][4];
    float2 ss2 = f2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      bf16x8 px, pz;
      memcpy(&px, base + vec * 16, 16); memcpy(&pz, base + ROWB + vec * 16, 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[i][j] = mul2(__bfloat1622float2(px.v[j]), silu2(__bfloat1622float2(pz.v[j])));
        ss2 = fma2(v[i][j], v[i][j], ss2);
      }
    }
    const float ss = warp_sum(ss2.x + ss2.y);
    const float rstd = rsqrtf(ss / (float)D + eps);
    if (lane == 0) rstd_out[row] = rstd;
    const float2 rs2 = f2(rstd, rstd);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (lane + 32 * i) * 8;
      const float4 w0 = *reinterpret_cast<const float4*>(s_w + col), w1 = *reinterpret_cast<const float4*>(s_w + col + 4);
      const float2 w[4] = {f2(w0.x, w0.y), f2(w0.z, w0.w), f2(w1.x, w1.y), f2(w1.z, w1.w)};
      bf16x8 o;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 r = mul2(mul2(v[i][j], rs2), w[j]);
        o.v[j] = __floats2bfloat162_rn(r.x, r.y);
      }
      st8(out + (size_t)row * ldo + col, o);
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int NV, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
gated_rmsnorm_bwd_rows_staged_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ z, int ldz,
                                     const float* __restrict__ weight, const float* __restrict__ rstd_in,
                                     const bf16* __restrict__ dout, int ldo, bf16* __restrict__ dx, int lddx,
                                     bf16* __restrict__ dz, int lddz, float* __restrict__ dw_partial, int M, int D,
                                     int dw_atomic) {   // 1: dw_partial is the gradient itself, added to with atomics
  extern __shared__ __align__(16) uint8_t st_smem[];
  constexpr int ROWB = NV * 32 * 16;                 // bytes of one staged row of one operand (D = NV * 256 exactly)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* my = st_smem + (size_t)warp * (2 * 3 * ROWB);
  float* s_w = reinterpret_cast<float*>(st_smem + (size_t)WARPS * 2 * 3 * ROWB);
  float* red = s_w + D;                              // [WARPS][256 + 8] for the final dw fold
  for (int i = threadIdx.x; i < D; i += WARPS * 32) s_w[i] = weight[i];   // a parameter: nothing the previous kernel wrote
  __syncthreads();
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  auto issue = [&](int row, int stage) {
    uint8_t* base = my + stage * 3 * ROWB;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (lane + 32 * i) * 8;
      cp_async_16(base + (lane + 32 * i) * 16, x + (size_t)row * ldx + col);
      cp_async_16(base + ROWB + (lane + 32 * i) * 16, z + (size_t)row * ldz + col);
      cp_async_16(base + 2 * ROWB + (lane + 32 * i) * 16, dout + (size_t)row * ldo + col);
    }
  };
  float2 dw[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dw[i][j] = f2(0.f, 0.f);
  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + warp;
  if (row < M) issue(row, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int stage = 0; row < M; row += stride, stage ^= 1) {
    if (row + stride < M) issue(row + stride, stage ^ 1);
    asm volatile("cp.async.commit_group;" ::: "memory");   // one group per iteration, possibly empty
    asm volatile("cp.async.wait_group 1;" ::: "memory");   // this lane's chunks of the current row have landed
    const uint8_t* base = my + stage * 3 * ROWB;
    const float rstd = rstd_in[row];
    float2 c2 = f2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      bf16x8 px, pz, pg;
      memcpy(&px, base + vec * 16, 16); memcpy(&pz, base + ROWB + vec * 16, 16); memcpy(&pg, base + 2 * ROWB + vec * 16, 16);
      const float4 w0 = *reinterpret_cast<const float4*>(s_w + vec * 8), w1 = *reinterpret_cast<const float4*>(s_w + vec * 8 + 4);
      const float2 w[4] = {f2(w0.x, w0.y), f2(w0.z, w0.w), f2(w1.x, w1.y), f2(w1.z, w1.w)};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 xv = __bfloat1622float2(px.v[j]), zz = __bfloat1622float2(pz.v[j]), go = __bfloat1622float2(pg.v[j]);
        c2 = fma2(mul2(xv, silu2(zz)), mul2(go, w[j]), c2);
      }
    }
    const float c = warp_sum(c2.x + c2.y) * rstd / (float)D;  // mean(xhat * w * dout)
    const float2 rs2 = f2(rstd, rstd), nc2 = f2(-c, -c);
    // the second pass re-reads shared memory on purpose: without this fence the compiler keeps all 3 x NV vectors of
    // the first pass live in registers (> 255 with the dw accumulators) and spills
    asm volatile("" ::: "memory");
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      const int col = vec * 8;
      bf16x8 px, pz, pg;
      memcpy(&px, base + vec * 16, 16); memcpy(&pz, base + ROWB + vec * 16, 16); memcpy(&pg, base + 2 * ROWB + vec * 16, 16);
      const float4 w0 = *reinterpret_cast<const float4*>(s_w + col), w1 = *reinterpret_cast<const float4*>(s_w + col + 4);
      const float2 w[4] = {f2(w0.x, w0.y), f2(w0.z, w0.w), f2(w1.x, w1.y), f2(w1.z, w1.w)};
      bf16x8 ox, oz;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 xv = __bfloat1622float2(px.v[j]), zz = __bfloat1622float2(pz.v[j]), go = __bfloat1622float2(pg.v[j]);
        const float2 sig = sigmoid2(zz);
        const float2 sg = mul2(zz, sig);                                                   // silu(z)
        const float2 dsg = mul2(sig, add2(f2(1.f, 1.f), fma2(f2(-1.f, -1.f), sg, zz)));   // silu'(z)
        const float2 xh = mul2(mul2(xv, sg), rs2);
        const float2 dxs = mul2(fma2(xh, nc2, mul2(go, w[j])), rs2);                       // (go*w - xh*c) * rstd
        const float2 o1 = mul2(dxs, sg), o2 = mul2(mul2(dxs, xv), dsg);
        ox.v[j] = __floats2bfloat162_rn(o1.x, o1.y);
        oz.v[j] = __floats2bfloat162_rn(o2.x, o2.y);
        dw[i][j] = fma2(go, xh, dw[i][j]);
      }
      st8(dx + (size_t)row * lddx + col, ox);
      st8(dz + (size_t)row * lddz + col, oz);
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  float* my_partial = dw_partial + (size_t)blockIdx.x * D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) { red[warp * 264 + lane * 8 + 2 * j] = dw[i][j].x; red[warp * 264 + lane * 8 + 2 * j + 1] = dw[i][j].y; }
    __syncthreads();
    for (int k = threadIdx.x; k < 256; k += WARPS * 32) {
      float sum = 0.f;
#pragma unroll
      for (int ww = 0; ww < WARPS; ++ww) sum += red[ww * 264 + k];
      if (dw_atomic) atomicAdd(dw_partial + i * 256 + k, sum);   // 600 blocks x D adds spread over D addresses
      else my_partial[i * 256 + k] = sum;
    }
  }
}
template <int NV, int WARPS>
constexpr int staged_smem_bytes(int D) { return WARPS * 2 * 3 * NV * 32 * 16 + D * 4 + WARPS * 264 * 4; }

// add + RMSNorm backward with the same cp.async row pipeline as the gated kernel above: r (fp32), dy (bf16) and the
// incoming residual gradient (fp32) of the NEXT row are copied into the warp's second stage while the current row is
// processed, so neither the first loads nor the late dres_in loads of the plain kernel (issued only after the row
// reduction) expose their latency.  Weight vector in shared memory; dw accumulated in registers across rows.
template <int NV, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
add_rmsnorm_bwd_staged_kernel(const float* __restrict__ r, const bf16* __restrict__ dy, const float* __restrict__ rstd_in,
                              const float* __restrict__ weight, const float* dres_in, float* dres_out,
                              bf16* __restrict__ dx_bf16, float* __restrict__ dw_partial, int M, int D, int dw_atomic) {
  extern __shared__ __align__(16) uint8_t st_smem[];
  constexpr int RB = NV * 32 * 32;     // bytes of one staged fp32 row (D = NV * 256)
  constexpr int GB = NV * 32 * 16;     // bytes of one staged bf16 row
  constexpr int STAGE = 2 * RB + GB;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* my = st_smem + (size_t)warp * 2 * STAGE;
  float* s_w = reinterpret_cast<float*>(st_smem + (size_t)WARPS * 2 * STAGE);
  float* red = s_w + D;
  for (int i = threadIdx.x; i < D; i += WARPS * 32) s_w[i] = weight[i];   // a parameter: nothing the previous kernel wrote
  __syncthreads();
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  const bool has_res = dres_in != nullptr;
  auto issue = [&](int row, int stage) {
    uint8_t* base = my + stage * STAGE;
    const size_t off = (size_t)row * D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      cp_async_16(base + vec * 32, r + off + vec * 8);
      cp_async_16(base + vec * 32 + 16, r + off + vec * 8 + 4);
      cp_async_16(base + 2 * RB + vec * 16, dy + off + vec * 8);
      if (has_res) {
        cp_async_16(base + RB + vec * 32, dres_in + off + vec * 8);
        cp_async_16(base + RB + vec * 32 + 16, dres_in + off + vec * 8 + 4);
      }
    }
  };
  float dw[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dw[i][j] = 0.f;
  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + warp;
  if (row < M) issue(row, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int stage = 0; row < M; row += stride, stage ^= 1) {
    if (row + stride < M) issue(row + stride, stage ^ 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    const uint8_t* base = my + stage * STAGE;
    const float rstd = rstd_in[row];
    float xh[NV][8], wdy[NV][8];
    float c = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      float g[8], w[8];
      ld8f(reinterpret_cast<const float*>(base + vec * 32), xh[i]);
      bf16x8 pg;
      memcpy(&pg, base + 2 * RB + vec * 16, 16);
      unpack8(pg, g);
      ld8f(s_w + vec * 8, w);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        xh[i][j] *= rstd;
        wdy[i][j] = g[j] * w[j];
        dw[i][j] += g[j] * xh[i][j];
        c += xh[i][j] * wdy[i][j];
      }
    }
    c = warp_sum(c) / (float)D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vec = lane + 32 * i;
      const size_t off = (size_t)row * D + vec * 8;
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (wdy[i][j] - xh[i][j] * c) * rstd;
      if (has_res) {
        float a[8];
        ld8f(reinterpret_cast<const float*>(base + RB + vec * 32), a);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += a[j];
      }
      st8f(dres_out + off, o);
      if (dx_bf16 != nullptr) st8(dx_bf16 + off, pack8(o));
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  float* my_partial = dw_partial + (size_t)blockIdx.x * D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp * 264 + lane * 8 + j] = dw[i][j];
    __syncthreads();
    for (int k = threadIdx.x; k < 256; k += WARPS * 32) {
      float sum = 0.f;
#pragma unroll
      for (int ww = 0; ww < WARPS; ++ww) sum += red[ww * 264 + k];
      if (dw_atomic) atomicAdd(dw_partial + i * 256 + k, sum);   // 600 blocks x D adds spread over D addresses
      else my_partial[i * 256 + k] = sum;
    }
  }
}
template <int NV, int WARPS>
constexpr int add_staged_smem_bytes(int D) { return WARPS * 2 * (2 * NV * 32 * 32 + NV * 32 * 16) + D * 4 + WARPS * 264 * 4; }

static int bwd_grid(int rows) {
  int g = cdiv(rows, kWarpsPerBlock);
  int cap = num_sms() * 8;
  return g < cap ? g : cap;
}

}  // namespace mb

using namespace mb;

#define DISPATCH_NV(nv, ...)                                   \
  switch (nv) {                                                \
    case 1: { constexpr int NV = 1; __VA_ARGS__; } break;      \
    case 2: { constexpr int NV = 2; __VA_ARGS__; } break;      \
    case 3: { constexpr int NV = 3; __VA_ARGS__; } break;      \
    case 4: { constexpr int NV = 4; __VA_ARGS__; } break;      \
    case 5: case 6: { constexpr int NV = 6; __VA_ARGS__; } break;   \
    case 7: case 8: { constexpr int NV = 8; __VA_ARGS__; } break;   \
    case 9: case 10: { constexpr int NV = 10; __VA_ARGS__; } break; \
    case 11: case 12: { constexpr int NV = 12; __VA_ARGS__; } break; \
    case 13: case 14: case 15: case 16: { constexpr int NV = 16; __VA_ARGS__; } break; \
    case 17: case 18: case 19: case 20: { constexpr int NV = 20; __VA_ARGS__; } break; \
    case 21: case 22: case 23: case 24: { constexpr int NV = 24; __VA_ARGS__; } break; \
    default: { constexpr int NV = 32; __VA_ARGS__; } break;    \
  }

extern "C" int mb_add_rmsnorm_fwd(const void* x, int x_is_fp32, const float* residual,
                                  const float* weight, void* y, float* res_out, float* rstd, int M,
                                  int D, float eps, cudaStream_t stream) {
  MB_REQUIRE(D % 8 == 0 && D <= 8192, "add_rmsnorm_fwd: D=%d must be a multiple of 8 and <= 8192", D);
  if (M == 0) return 0;
  const int nv = cdiv(D, 256);
  dim3 grid(cdiv(M, kWarpsPerBlock)), block(kWarpsPerBlock * 32);
  DISPATCH_NV(nv, {
    if (x_is_fp32)
      MB_CUDA(launch_pdl(add_rmsnorm_fwd_kernel<NV, true>, grid, block, 0, stream, x, residual, weight, (bf16*)y, res_out, rstd, M, D, eps));
    else
      MB_CUDA(launch_pdl(add_rmsnorm_fwd_kernel<NV, false>, grid, block, 0, stream, x, residual, weight, (bf16*)y, res_out, rstd, M, D, eps));
  });
  return check_launch("add_rmsnorm_fwd");
}

extern "C" long long mb_norm_bwd_scratch_bytes(int rows, int D) {
  return (long long)bwd_grid(rows) * D * sizeof(float);
}

extern "C" int mb_add_rmsnorm_bwd(const float* r, const void* dy, const float* rstd,
                                  const float* weight, const float* dres_in, float* dres_out,
                                  void* dx_bf16, float* dw, int dw_accumulate, float* scratch, int M,
                                  int D, cudaStream_t stream) {
  MB_REQUIRE(D % 8 == 0 && D <= 8192, "add_rmsnorm_bwd: D=%d must be a multiple of 8 and <= 8192", D);
  if (M == 0) return 0;
  const int nv = cdiv(D, 256);
  int grid = bwd_grid(M);
  if (D % 256 == 0 && (nv == 3 || nv == 6 || nv == 10)) {   // 280M / 780M / 2.7B widths: cp.async row pipeline
#define LAUNCH_ADD_STAGED(NV_, W_)                                                                                     \
    {                                                                                                                  \
      auto kern = add_rmsnorm_bwd_staged_kernel<NV_, W_>;                                                              \
      const int smem = add_staged_smem_bytes<NV_, W_>(D);                                                              \
      static int per_sm = 0;                                                                                           \
      if (per_sm == 0) {                                                                                               \
        MB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                        \
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W_ * 32, smem));                          \
        if (per_sm < 1) per_sm = 1;                                                                                    \
      }                                                                                                                \
      int g2 = cdiv(M, W_);                                                                                            \
      if (g2 > num_sms() * per_sm) g2 = num_sms() * per_sm;                                                            \
      if (g2 > grid) g2 = grid;   /* never more partial rows than the caller's scratch holds */                        \
      grid = g2;                                                                                                       \
      MB_CUDA(launch_pdl(kern, dim3(grid), dim3(W_ * 32), smem, stream, r, (const bf16*)dy, rstd, weight, dres_in,    \
                         dres_out, (bf16*)dx_bf16, dw_accumulate ? dw : scratch, M, D, dw_accumulate ? 1 : 0));        \
    }
    if (nv == 3) LAUNCH_ADD_STAGED(3, 4) else if (nv == 6) LAUNCH_ADD_STAGED(6, 2) else LAUNCH_ADD_STAGED(10, 2)
#undef LAUNCH_ADD_STAGED
    if (check_launch("add_rmsnorm_bwd/staged")) return 1;
    // accumulating into an existing gradient (training): the blocks added their column sums with atomics, no second launch
    if (dw_accumulate) return 0;
    MB_CUDA(launch_pdl(colsum_partials_kernel, dim3(cdiv(D, 32)), dim3(1024), 0, stream, scratch, dw, grid, D, dw_accumulate));
    return check_launch("add_rmsnorm_bwd/colsum");
  }
  DISPATCH_NV(nv, {
    add_rmsnorm_bwd_kernel<NV><<<grid, kWarpsPerBlock * 32, 0, stream>>>(
        r, (const bf16*)dy, rstd, weight, dres_in, dres_out, (bf16*)dx_bf16, scratch, M, D);
  });
  if (check_launch("add_rmsnorm_bwd")) return 1;
  MB_CUDA(launch_pdl(colsum_partials_kernel, dim3(cdiv(D, 32)), dim3(1024), 0, stream, scratch, dw, grid, D, dw_accumulate));
  return check_launch("add_rmsnorm_bwd/colsum");
}

extern "C" int mb_gated_rmsnorm_fwd(const void* x, int ldx, const void* z, int ldz,
                                    const float* weight, void* out, int ldo, float* rstd, int M,
                                    int D, int group_size, float eps, cudaStream_t stream) {
  MB_REQUIRE(D % group_size == 0 && group_size % 8 == 0 && group_size <= 8192,
             "gated_rmsnorm_fwd: bad D=%d group_size=%d", D, group_size);
  MB_REQUIRE(ldx % 8 == 0 && ldz % 8 == 0 && ldo % 8 == 0, "gated_rmsnorm_fwd: row strides must be multiples of 8");
  if (M == 0) return 0;
  const int G = D / group_size, nv = cdiv(group_size, 256);
  if (G == 1 && D % 256 == 0 && (nv == 6 || nv == 12)) {   // d_inner 1536 / 3072: persistent warps, rows staged one ahead
#define LAUNCH_GFWD_STAGED(NV_, W_)                                                                                    \
    {                                                                                                                  \
      auto kern = gated_rmsnorm_fwd_staged_kernel<NV_, W_>;                                                            \
      const int smem = W_ * 2 * 2 * NV_ * 32 * 16 + D * 4;                                                             \
      static int per_sm = 0;                                                                                           \
      if (per_sm == 0) {                                                                                               \
        MB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                        \
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W_ * 32, smem));                          \
        if (per_sm < 1) per_sm = 1;                                                                                    \
      }                                                                                                                \
      int g2 = cdiv(M, W_);                                                                                            \
      if (g2 > num_sms() * per_sm) g2 = num_sms() * per_sm;                                                            \
      MB_CUDA(launch_pdl(kern, dim3(g2), dim3(W_ * 32), smem, stream, (const bf16*)x, ldx, (const bf16*)z, ldz, weight, \
                         (bf16*)out, ldo, rstd, M, D, eps));                                                           \
    }
    if (nv == 6) LAUNCH_GFWD_STAGED(6, 4) else LAUNCH_GFWD_STAGED(12, 2)
#undef LAUNCH_GFWD_STAGED
    return check_launch("gated_rmsnorm_fwd/staged");
  }
  dim3 grid(cdiv((long long)M * G, kWarpsPerBlock)), block(kWarpsPerBlock * 32);
  DISPATCH_NV(nv, {
    gated_rmsnorm_fwd_kernel<NV><<<grid, block, 0, stream>>>((const bf16*)x, ldx, (const bf16*)z, ldz, weight,
                                                             (bf16*)out, ldo, rstd, M, G, group_size, eps);
  });
  return check_launch("gated_rmsnorm_fwd");
}

extern "C" int mb_gated_rmsnorm_bwd(const void* x, int ldx, const void* z, int ldz,
                                    const float* weight, const float* rstd, const void* dout, int ldo,
                                    void* dx, int lddx, void* dz, int lddz, float* dw,
                                    int dw_accumulate, float* scratch, int M, int D, int group_size,
                                    cudaStream_t stream) {
  MB_REQUIRE(D % group_size == 0 && group_size % 8 == 0 && group_size <= 8192,
             "gated_rmsnorm_bwd: bad D=%d group_size=%d", D, group_size);
  if (M == 0) return 0;
  const int G = D / group_size, nv = cdiv(group_size, 256);
  if (G == 1) {
    int grid1 = bwd_grid(M);
    // rows staged by cp.async one ahead (D a whole number of 256-column vectors, shared memory permitting)
    if (D % 256 == 0 && (nv == 6 || nv == 12) && ldx % 8 == 0 && ldz % 8 == 0 && ldo % 8 == 0) {
#define LAUNCH_STAGED(NV_, W_)                                                                                        \
      {                                                                                                               \
        auto kern = gated_rmsnorm_bwd_rows_staged_kernel<NV_, W_>;                                                     \
        const int smem = staged_smem_bytes<NV_, W_>(D);                                                                \
        static int per_sm = 0;                                                                                         \
        if (per_sm == 0) {                                                                                             \
          MB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                      \
          MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W_ * 32, smem));                        \
          if (per_sm < 1) per_sm = 1;                                                                                  \
        }                                                                                                              \
        grid1 = cdiv(M, W_);                                                                                           \
        const int cap = bwd_grid(M);                                                                                   \
        if (grid1 > num_sms() * per_sm) grid1 = num_sms() * per_sm;                                                    \
        if (grid1 > cap) grid1 = cap;                                                                                  \
        MB_CUDA(launch_pdl(kern, dim3(grid1), dim3(W_ * 32), smem, stream, (const bf16*)x, ldx, (const bf16*)z, ldz,  \
                           weight, rstd, (const bf16*)dout, ldo, (bf16*)dx, lddx, (bf16*)dz, lddz,                     \
                           dw_accumulate ? dw : scratch, M, D, dw_accumulate ? 1 : 0));                                \
      }
      if (nv == 6) LAUNCH_STAGED(6, 4) else LAUNCH_STAGED(12, 2)
#undef LAUNCH_STAGED
      if (check_launch("gated_rmsnorm_bwd/rows_staged")) return 1;
      if (dw_accumulate) return 0;   // column sums already added to dw with atomics
      MB_CUDA(launch_pdl(colsum_partials_kernel, dim3(cdiv(D, 32)), dim3(1024), 0, stream, scratch, dw, grid1, D, dw_accumulate));
      return check_launch("gated_rmsnorm_bwd/colsum");
    }
    DISPATCH_NV(nv, {
      // persistent: exactly one wave of resident blocks (never more than the scratch rows bwd_grid allows)
      static int per_sm = 0;
      if (per_sm == 0)
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gated_rmsnorm_bwd_rows_kernel<NV>, kWarpsPerBlock * 32, 0));
      if (per_sm > 0 && grid1 > num_sms() * per_sm) grid1 = num_sms() * per_sm;
      gated_rmsnorm_bwd_rows_kernel<NV><<<grid1, kWarpsPerBlock * 32, 0, stream>>>(
          (const bf16*)x, ldx, (const bf16*)z, ldz, weight, rstd, (const bf16*)dout, ldo, (bf16*)dx, lddx, (bf16*)dz, lddz,
          scratch, M, D);
    });
    if (check_launch("gated_rmsnorm_bwd/rows")) return 1;
    MB_CUDA(launch_pdl(colsum_partials_kernel, dim3(cdiv(D, 32)), dim3(1024), 0, stream, scratch, dw, grid1, D, dw_accumulate));
    return check_launch("gated_rmsnorm_bwd/colsum");
  }
  const int grid = bwd_grid(M * G);
  if (G > 1) MB_CUDA(cudaMemsetAsync(scratch, 0, (size_t)grid * D * sizeof(float), stream));
  DISPATCH_NV(nv, {
    gated_rmsnorm_bwd_kernel<NV><<<grid, kWarpsPerBlock * 32, 0, stream>>>(
        (const bf16*)x, ldx, (const bf16*)z, ldz, weight, rstd, (const bf16*)dout, ldo, (bf16*)dx, lddx,
        (bf16*)dz, lddz, scratch, M, G, group_size);
  });
  if (check_launch("gated_rmsnorm_bwd")) return 1;
  MB_CUDA(launch_pdl(colsum_partials_kernel, dim3(cdiv(D, 32)), dim3(1024), 0, stream, scratch, dw, grid, D, dw_accumulate));
  return check_launch("gated_rmsnorm_bwd/colsum");
}

// ------------------------------------------------------------------------------------------
// mb_add_rmsnorm_bwd_noalloc — the same backward as mb_add_rmsnorm_bwd for callers that cannot
// hand us a scratch buffer because they cannot allocate one.
//
// Who that is: a host that gave the whole device to its own memory pool at start-up and has no
// spare bytes to carve out later, or a latency-critical loop that must never call cudaMalloc /
// cudaFree (both synchronise the device) and has no allocator of ours (or torch's) to ask.
// For them the per-block weight-gradient partials live in a fixed array that is part of this
// library's image: every GPU that loads libmamba_b200.so holds these 1 MiB until the process
// exits, whether or not this entry point is ever called.
//
// Limits (checked; a violation returns an error and launches nothing):
//   D % 8 == 0 and D <= kNoAllocMaxD (1024);  M > 0 fits in an int.
// The partial rows are capped at kNoAllocMaxRows (256) blocks, so large M runs on a smaller grid
// than the scratch-taking entry point — same result, somewhat lower bandwidth.
//
// Sharing: there is ONE array per device.  Every call overwrites the rows it uses (each launched
// block writes all D columns of its own row and the reduction reads exactly those rows), so
// nothing needs to start at zero and nothing is cleared; stale rows from an earlier, larger call
// are never read.  Two calls whose kernels overlap on the same GPU (different streams) race on
// the array — serialise them.
//
// Set-up: `g_noalloc_scratch` below is, in host code, only a handle for the symbol, not a device
// address.  The first call on each device (with that device current, which is the device `stream`
// must belong to) resolves the address with cudaGetSymbolAddress and caches it; make that first
// call outside any stream capture.  No cudaMalloc, cudaFree or allocating library call is made.
// ------------------------------------------------------------------------------------------
namespace mb {
constexpr int kNoAllocMaxD = 1024;
constexpr int kNoAllocMaxRows = 256;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocMaxRows * kNoAllocMaxD];
static float* g_noalloc_ptr[64] = {nullptr};  // resolved device address of the array, per device
}  // namespace mb

extern "C" int mb_add_rmsnorm_bwd_noalloc(const float* r, const void* dy, const float* rstd,
                                          const float* weight, const float* dres_in, float* dres_out,
                                          void* dx_bf16, float* dw, int dw_accumulate, int M, int D,
                                          cudaStream_t stream) {
  MB_REQUIRE(D > 0 && D % 8 == 0 && D <= kNoAllocMaxD,
             "add_rmsnorm_bwd_noalloc: D=%d outside the built-in scratch limit (multiple of 8, <= %d); "
             "use mb_add_rmsnorm_bwd with caller scratch", D, kNoAllocMaxD);
  MB_REQUIRE(M >= 0, "add_rmsnorm_bwd_noalloc: negative M");
  if (M == 0) return 0;
  int dev = 0;
  MB_CUDA(cudaGetDevice(&dev));
  MB_REQUIRE(dev >= 0 && dev < 64, "add_rmsnorm_bwd_noalloc: device index %d not supported", dev);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    MB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // per device: the current one
    MB_REQUIRE(p != nullptr, "add_rmsnorm_bwd_noalloc: could not resolve the scratch symbol");
    g_noalloc_ptr[dev] = static_cast<float*>(p);
  }
  float* scratch = g_noalloc_ptr[dev];
  const int nv = cdiv(D, 256);
  int grid = cdiv(M, kWarpsPerBlock);
  if (grid > kNoAllocMaxRows) grid = kNoAllocMaxRows;
  DISPATCH_NV(nv, {
    add_rmsnorm_bwd_kernel<NV><<<grid, kWarpsPerBlock * 32, 0, stream>>>(
        r, (const bf16*)dy, rstd, weight, dres_in, dres_out, (bf16*)dx_bf16, scratch, M, D);
  });
  if (check_launch("add_rmsnorm_bwd_noalloc")) return 1;
  MB_CUDA(launch_pdl(colsum_partials_kernel, dim3(cdiv(D, 32)), dim3(1024), 0, stream, scratch, dw, grid, D, dw_accumulate));
  return check_launch("add_rmsnorm_bwd_noalloc/colsum");
}
