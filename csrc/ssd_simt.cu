// SSD (Mamba-2) support kernels that are not GEMM-shaped: K7 (dt softplus + chunk cumsum), K10
// (inter-chunk state passing, forward and reverse) and the backward "finalize" pass that turns
// the per-token partials produced by the tcgen05 kernels into d(dt_raw), dA_log, d(dt_bias).
// Replaces the Triton kernels of mamba_ssm/ops/triton/ssd_chunk_state.py::_chunk_cumsum_* and
// ssd_state_passing.py.  All fp32.
//
// Layouts:  dt, a (= cumsum of dt*A inside the chunk):  [B, nc, H, L] fp32
//           states / dstates:                           [B, nc, H, P, N] fp32
//           S_prev / dS operand copies:                 [B, nc, H, P, N] bf16
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

// one block = one (b, c, group of kPrepHeads heads); thread t = token c*L + t (L == blockDim.x == 256 or 128 or 64).
// Eight heads per block: the eight dt loads of a thread are independent and issue together, each of the block's
// warps scans one head, and the grid is 3x the (b, c) count -- one block per (b, c) walking all heads in turn
// left 128 blocks serialising 24 dependent load + softplus rounds (15 us for 5 MB).
constexpr int kPrepHeads = 8;
__global__ void ssd_prep_kernel(const bf16* __restrict__ dt_raw, long long ld_dt,
                                const float* __restrict__ dt_bias, const float* __restrict__ A_log,
                                float* __restrict__ dt_out, float* __restrict__ a_out, int T, int H, int L,
                                int nc, float dt_min, float dt_max) {
  extern __shared__ float sm[];  // [kPrepHeads][L] dA, then scanned in place
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  const int b = blockIdx.x / nc, c = blockIdx.x % nc;
  const int h0 = blockIdx.y * kPrepHeads;
  const int t = threadIdx.x;
  const int tok = min(c * L + t, T - 1);
  const bool valid = c * L + t < T;
  const bf16* src = dt_raw + ((size_t)b * T + tok) * ld_dt;
  float* dt_o = dt_out + ((size_t)(b * nc + c) * H) * L;
  float* a_o = a_out + ((size_t)(b * nc + c) * H) * L;
  float raw[kPrepHeads];
#pragma unroll
  for (int i = 0; i < kPrepHeads; ++i) raw[i] = __bfloat162float(src[min(h0 + i, H - 1)]);
#pragma unroll
  for (int i = 0; i < kPrepHeads; ++i) {
    const int h = h0 + i;
    if (h < H) {
      float dt = 0.f;
      if (valid) dt = fminf(fmaxf(softplus_f(raw[i] + dt_bias[h]), dt_min), dt_max);  // dt_limit
      dt_o[(size_t)h * L + t] = dt;
      sm[i * L + t] = dt * (-fast_exp(A_log[h]));
    }
  }
  __syncthreads();
  // inclusive scan over t for each head: warp w handles heads w, w+nw, ...
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int per = L / 32;  // consecutive elements per lane
  for (int i = warp; i < kPrepHeads && h0 + i < H; i += nw) {
    const float* row = sm + i * L;
    float run = 0.f;
    float loc[8];
    for (int k = 0; k < per; ++k) { run += row[lane * per + k]; loc[k] = run; }
    float incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const float excl = incl - run;
    for (int k = 0; k < per; ++k) a_o[(size_t)(h0 + i) * L + lane * per + k] = loc[k] + excl;
  }
}

// Forward recurrence over chunks.  One thread = 4 consecutive elements of the (P*N) state.
__global__ void ssd_state_passing_fwd_kernel(const float* __restrict__ states, const float* __restrict__ a,
                                             bf16* __restrict__ sprev_bf16, float* __restrict__ final_state,
                                             int nc, int H, int PN, int L) {
  const int bh = blockIdx.y;  // b * H + h
  const int b = bh / H, h = bh % H;
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (e >= PN) return;
  float4 prev = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int c = 0; c < nc; ++c) {
    const size_t base = ((size_t)(b * nc + c) * H + h) * PN + e;
    bf162 lo = __floats2bfloat162_rn(prev.x, prev.y), hi = __floats2bfloat162_rn(prev.z, prev.w);
    *reinterpret_cast<bf162*>(sprev_bf16 + base) = lo;
    *reinterpret_cast<bf162*>(sprev_bf16 + base + 2) = hi;
    const float decay = fast_exp(a[((size_t)(b * nc + c) * H + h) * L + (L - 1)]);
    const float4 s = *reinterpret_cast<const float4*>(states + base);
    prev.x = decay * prev.x + s.x; prev.y = decay * prev.y + s.y;
    prev.z = decay * prev.z + s.z; prev.w = decay * prev.w + s.w;
  }
  if (final_state != nullptr) *reinterpret_cast<float4*>(final_state + (size_t)bh * PN + e) = prev;
}

// Reverse recurrence.  dstates_direct[c] = gradient reaching the state that ENTERS chunk c through
// the C.S_prev term.  Outputs: ds_bf16[c] = total gradient of chunk c's own contribution S_c
// (== total gradient of the state entering chunk c+1), and da_last[b,c,h] += sum(g * decay * S_prev).
__global__ void ssd_state_passing_bwd_kernel(const float* __restrict__ dstates_direct,
                                             const bf16* __restrict__ sprev_bf16, const float* __restrict__ a,
                                             bf16* __restrict__ ds_bf16, float* __restrict__ da_last, int nc,
                                             int H, int PN, int L) {
  const int bh = blockIdx.y;
  const int b = bh / H, h = bh % H;
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const bool active = e < PN;
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);  // gradient of the state entering chunk c+1
  __shared__ float red[32];
  for (int c = nc - 1; c >= 0; --c) {
    const size_t base = ((size_t)(b * nc + c) * H + h) * PN + e;
    const float decay = fast_exp(a[((size_t)(b * nc + c) * H + h) * L + (L - 1)]);
    float part = 0.f;
    if (active) {
      *reinterpret_cast<bf162*>(ds_bf16 + base) = __floats2bfloat162_rn(g.x, g.y);
      *reinterpret_cast<bf162*>(ds_bf16 + base + 2) = __floats2bfloat162_rn(g.z, g.w);
      const float2 p0 = __bfloat1622float2(*reinterpret_cast<const bf162*>(sprev_bf16 + base));
      const float2 p1 = __bfloat1622float2(*reinterpret_cast<const bf162*>(sprev_bf16 + base + 2));
      part = decay * (g.x * p0.x + g.y * p0.y + g.z * p1.x + g.w * p1.y);
      const float4 d = *reinterpret_cast<const float4*>(dstates_direct + base);
      g.x = d.x + decay * g.x; g.y = d.y + decay * g.y; g.z = d.z + decay * g.z; g.w = d.w + decay * g.w;
    }
    part = warp_sum(part);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x < 32) {
      float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
      v = warp_sum(v);
      if (threadIdx.x == 0 && v != 0.f) atomicAdd(da_last + (size_t)(b * nc + c) * H + h, v);
    }
    __syncthreads();
  }
}

// Backward finalize.  One block = one (b, c, head), thread t = token:
//   rowpart_t = rowsum_m[t] + sum_p dY[t,p] * y_off[t,p]
//     rowsum_m = row sums of dM o M from the dCB kernel, built from the SAME bf16-rounded M the dx kernel
//     uses for the column sums (dt_t * ddt1_t), so the suffix sum of (row - column) sums — which equals the
//     rectangle sum_{i >= t > j} (dM o M)[i,j] — cancels exactly instead of amplifying operand rounding;
//     y_off = exp(a_t) C_t S_prev^T is the inter-chunk part, stored by forward.
//   da_t      = rowpart_t - dt_t * (ddt1_t + ddt2_t)  (+ da_last at t = L-1, + sum_s dt_s ddt2_s at L-1)
//   ddA_t     = suffix-sum_{s >= t} da_s ;  ddt_t = ddt1_t + ddt2_t + A_h * ddA_t
//   draw_t    = ddt_t * sigmoid(raw_t + bias_h) -> bf16 into the dt columns of d(zxbcdt)
//   dA_log_h += A_h * sum_t dt_t * ddA_t ;  ddt_bias_h += sum_t draw_t
__global__ void ssd_bwd_finalize_kernel(const bf16* __restrict__ dy, long long lddy, const bf16* __restrict__ y,
                                        long long ldy, const bf16* __restrict__ dt_raw, long long ld_dt,
                                        const float* __restrict__ dt, const float* __restrict__ ddt1,
                                        const float* __restrict__ ddt2, const float* __restrict__ rowsum_m,
                                        const float* __restrict__ da_last,
                                        const float* __restrict__ dt_bias, const float* __restrict__ A_log,
                                        bf16* __restrict__ draw_out,
                                        long long ld_draw, float* __restrict__ dA_log, float* __restrict__ ddt_bias,
                                        int T, int H, int P, int L, int nc, float dt_min, float dt_max) {
  __shared__ float wsum[8];
  __shared__ float red3[2][8];
  __shared__ float s_rowpart[256];
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  // heads are the fastest grid index: blocks resident together read neighbouring 128-byte pieces of the same dY /
  // y_off rows (and write neighbouring bf16 d(dt) columns), not the same piece of 128 different chunks
  const int b = blockIdx.y / nc, c = blockIdx.y % nc;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nw = blockDim.x >> 5;
  const int tok = c * L + t;
  const bool valid = tok < T;
  const size_t row = (size_t)b * T + tok;
  {
    const int h = blockIdx.x;
    const size_t sidx = ((size_t)(b * nc + c) * H + h) * L + t;
    const float A = -fast_exp(A_log[h]);
    const float dtv = dt[sidx];
    // sum_p dY[t,p] * y_off[t,p]: 8 lanes share one token row (16 bytes each = one full 128-byte line per
    // row for P = 64), so every warp load is 4 complete lines instead of 32 partial ones
    {
      const int per_row = P >> 3;                   // 16-byte chunks per row (8 for P = 64)
      const int rows_per_pass = blockDim.x / per_row;
      const int chunk = t % per_row, r0 = t / per_row;
      for (int rr = r0; rr < L; rr += rows_per_pass) {
        const int tk = c * L + rr;
        float acc = 0.f;
        if (tk < T) {
          const size_t ro = (size_t)b * T + tk;
          float a8[8], b8[8];
          unpack8(ld8(dy + ro * lddy + h * P + chunk * 8), a8);
          unpack8(ld8(y + ro * ldy + h * P + chunk * 8), b8);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc += a8[j] * b8[j];
        }
        for (int o = per_row >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (chunk == 0) s_rowpart[rr] = acc;
      }
    }
    __syncthreads();
    float rowpart = s_rowpart[t];
    const float d1 = ddt1[sidx], d2 = ddt2[sidx];
    rowpart += rowsum_m[sidx];
    float da = rowpart - dtv * (d1 + d2);
    // a_last receives: state-passing term and sum_s dt_s * ddt2_s  (w_s = exp(a_last - a_s) dt_s)
    float q = dtv * d2;
    q = warp_sum(q);
    if (lane == 0) wsum[warp] = q;
    __syncthreads();
    if (t == L - 1) {
      float tot = 0.f;
      for (int w = 0; w < nw; ++w) tot += wsum[w];
      da += tot + da_last[(size_t)(b * nc + c) * H + h];
    }
    __syncthreads();
    // suffix sum over t (inclusive): reverse inclusive scan
    float incl = da;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float v = __shfl_down_sync(0xffffffffu, incl, o);
      if (lane + o < 32) incl += v;
    }
    if (lane == 0) wsum[warp] = incl;  // total of this warp
    __syncthreads();
    float after = 0.f;
    for (int w = warp + 1; w < nw; ++w) after += wsum[w];
    const float ddA = incl + after;
    const float ddt = d1 + d2 + A * ddA;
    float draw = 0.f;
    if (valid) {
      const float raw = __bfloat162float(dt_raw[row * ld_dt + h]) + dt_bias[h];
      const float sp = softplus_f(raw);
      draw = (sp >= dt_min && sp <= dt_max) ? ddt * sigmoid_f(raw) : 0.f;   // no gradient through a clamped dt
      draw_out[row * ld_draw + h] = __float2bfloat16(draw);
    }
    float r0 = warp_sum(dtv * ddA), r1 = warp_sum(draw);
    __syncthreads();
    if (lane == 0) { red3[0][warp] = r0; red3[1][warp] = r1; }
    __syncthreads();
    if (t == 0) {
      float s0 = 0.f, s1 = 0.f;
      for (int w = 0; w < nw; ++w) { s0 += red3[0][w]; s1 += red3[1][w]; }
      atomicAdd(dA_log + h, A * s0);
      atomicAdd(ddt_bias + h, s1);
    }
    __syncthreads();
  }
}


// K23 (Mamba-2 decode): one token of the recurrence for every (batch, head).
//   state[p][n] = state[p][n] * exp(dt * A_h) + dt * x[p] * B[n]
//   y[p]        = sum_n state[p][n] * C[n] + D_h * x[p]        (* silu(z[p]) when z is given)
// One block per (b, h); a warp walks rows p, each lane owning N/32 consecutive state columns, so the fp32
// state (the only real traffic: P*N*4 bytes read + written) moves in whole 128-byte lines.
__global__ void __launch_bounds__(256)
ssd_state_update_kernel(float* __restrict__ state, const bf16* __restrict__ x, long long ldx,
                        const bf16* __restrict__ dt_raw, long long lddt, const float* __restrict__ A,
                        const bf16* __restrict__ Bm, const bf16* __restrict__ Cm, long long ldbc,
                        const float* __restrict__ D, const bf16* __restrict__ z, long long ldz,
                        const float* __restrict__ dt_bias, bf16* __restrict__ y, long long ldy, int H, int G,
                        int P, int N, int softplus) {
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int g = h / (H / G);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float dt = __bfloat162float(dt_raw[(size_t)b * lddt + h]) + (dt_bias != nullptr ? dt_bias[h] : 0.f);
  if (softplus) dt = softplus_f(dt);
  const float dA = fast_exp(dt * A[h]);
  const int per = N / 32;  // 4 for d_state 128
  float bv[8], cv[8];
  for (int i = 0; i < per; ++i) {
    bv[i] = __bfloat162float(Bm[(size_t)b * ldbc + g * N + lane * per + i]) * dt;
    cv[i] = __bfloat162float(Cm[(size_t)b * ldbc + g * N + lane * per + i]);
  }
  float* st = state + ((size_t)b * H + h) * P * N;
  const float Dh = D != nullptr ? D[h] : 0.f;
  for (int p = warp; p < P; p += nw) {
    const float xv = __bfloat162float(x[(size_t)b * ldx + h * P + p]);
    float acc = 0.f;
    float* row = st + (size_t)p * N + lane * per;
    for (int i = 0; i < per; ++i) {
      const float sv = row[i] * dA + bv[i] * xv;
      row[i] = sv;
      acc += sv * cv[i];
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      float o = acc + Dh * xv;
      if (z != nullptr) o *= silu_f(__bfloat162float(z[(size_t)b * ldz + h * P + p]));
      y[(size_t)b * ldy + h * P + p] = __float2bfloat16(o);
    }
  }
}

}  // namespace mb
using namespace mb;

extern "C" int mb_ssd_prep(const void* dt_raw, long long ld_dt, const float* dt_bias, const float* A_log,
                           float* dt_out, float* a_out, int B, int T, int H, int L, float dt_min, float dt_max,
                           cudaStream_t stream) {
  MB_REQUIRE(L == 64 || L == 128 || L == 256, "ssd_prep: chunk size %d not in {64,128,256}", L);
  const int nc = cdiv(T, L);
  const size_t smem = (size_t)kPrepHeads * L * sizeof(float);
  MB_CUDA(launch_pdl(ssd_prep_kernel, dim3(B * nc, cdiv(H, kPrepHeads)), dim3(L), smem, stream, (const bf16*)dt_raw, ld_dt,
                     dt_bias, A_log, dt_out, a_out, T, H, L, nc, dt_min, dt_max));
  return check_launch("ssd_prep");
}

extern "C" int mb_ssd_state_passing_fwd(const float* states, const float* a, void* sprev_bf16,
                                        float* final_state, int B, int nc, int H, int PN, int L,
                                        cudaStream_t stream) {
  MB_REQUIRE(PN % 4 == 0, "ssd_state_passing_fwd: P*N must be a multiple of 4");
  dim3 grid(cdiv(PN / 4, 256), B * H);
  ssd_state_passing_fwd_kernel<<<grid, 256, 0, stream>>>(states, a, (bf16*)sprev_bf16, final_state, nc, H, PN, L);
  return check_launch("ssd_state_passing_fwd");
}

extern "C" int mb_ssd_state_passing_bwd(const float* dstates_direct, const void* sprev_bf16, const float* a,
                                        void* ds_bf16, float* da_last, int B, int nc, int H, int PN, int L,
                                        cudaStream_t stream) {
  MB_REQUIRE(PN % 4 == 0, "ssd_state_passing_bwd: P*N must be a multiple of 4");
  MB_CUDA(cudaMemsetAsync(da_last, 0, (size_t)B * nc * H * sizeof(float), stream));
  dim3 grid(cdiv(PN / 4, 256), B * H);
  ssd_state_passing_bwd_kernel<<<grid, 256, 0, stream>>>(dstates_direct, (const bf16*)sprev_bf16, a, (bf16*)ds_bf16,
                                                         da_last, nc, H, PN, L);
  return check_launch("ssd_state_passing_bwd");
}

extern "C" int mb_ssd_bwd_finalize(const void* dy, long long lddy, const void* y_off, long long ldy,
                                   const void* dt_raw, long long ld_dt, const float* dt, const float* ddt1,
                                   const float* ddt2, const float* rowsum_m, const float* da_last,
                                   const float* dt_bias,
                                   const float* A_log, void* draw_out, long long ld_draw, float* dA_log,
                                   float* ddt_bias, int B, int T, int H, int P, int L, float dt_min, float dt_max,
                                   cudaStream_t stream) {
  MB_REQUIRE(L == 64 || L == 128 || L == 256, "ssd_bwd_finalize: chunk size %d not in {64,128,256}", L);
  MB_REQUIRE(P == 8 || P == 16 || P == 32 || P == 64 || P == 128 || P == 256,
             "ssd_bwd_finalize: headdim %d must be a power of two in [8, 256] (lanes of one row reduce by shuffles)", P);
  const int nc = cdiv(T, L);
  MB_REQUIRE((long long)B * nc <= 65535, "ssd_bwd_finalize: B * chunks = %lld exceeds the grid's y extent", (long long)B * nc);
  MB_CUDA(launch_pdl(ssd_bwd_finalize_kernel, dim3(H, B * nc), dim3(L), 0, stream, (const bf16*)dy, lddy, (const bf16*)y_off,
                     ldy, (const bf16*)dt_raw, ld_dt, dt, ddt1, ddt2, rowsum_m, da_last, dt_bias, A_log, (bf16*)draw_out,
                     ld_draw, dA_log, ddt_bias, T, H, P, L, nc, dt_min, dt_max));
  return check_launch("ssd_bwd_finalize");
}

extern "C" int mb_ssd_state_update(float* state, const void* x, long long ldx, const void* dt_raw, long long lddt,
                                   const float* A, const void* Bm, const void* Cm, long long ldbc, const float* D,
                                   const void* z, long long ldz, const float* dt_bias, void* y, long long ldy,
                                   int B, int H, int G, int P, int N, int softplus, cudaStream_t stream) {
  MB_REQUIRE(N % 32 == 0 && N <= 256, "ssd_state_update: d_state %d must be a multiple of 32 and <= 256", N);
  MB_REQUIRE(H % G == 0, "ssd_state_update: nheads %% ngroups != 0");
  if (B * H == 0) return 0;
  ssd_state_update_kernel<<<B * H, 256, 0, stream>>>(state, (const bf16*)x, ldx, (const bf16*)dt_raw, lddt, A,
                                                     (const bf16*)Bm, (const bf16*)Cm, ldbc, D, (const bf16*)z, ldz,
                                                     dt_bias, (bf16*)y, ldy, H, G, P, N, softplus);
  return check_launch("ssd_state_update");
}
