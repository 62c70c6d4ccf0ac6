// K4 / K5 / K23: Mamba-1 selective scan (forward, backward) and the single-token state update.
//
// Replaces mamba_ssm's csrc/selective_scan CUDA kernels and the Triton selective_state_update
// (SURVEY.md section 2.3 rows K4, K5, K23).  This is the literal-compatibility path: the
// reference's `MambaConfig(d_model=768, vocab_size=50304)` builds Mamba-1 mixers.
//
//   h_t[d,n] = exp(dt_t[d] A[d,n]) h_{t-1}[d,n] + dt_t[d] B_t[n] u_t[d]
//   y_t[d]   = (sum_n C_t[n] h_t[d,n] + D[d] u_t[d]) * silu(z_t[d]),   dt = softplus(delta + bias)
//
// Layout is channel-last ([B, T, Dn] activations, [B, T, N] for B/C, all possibly strided column
// slices), N = 16.  No tensor-core content: a first-order recurrence.  Parallelisation: time is cut
// into tiles of 32 tokens; FOUR lanes own one channel of one (batch, tile), each lane carrying 4 of the
// 16 states in registers (4 independent recurrences per thread: the FMA latency of one hides behind the
// other three, and a sum over n is 3 in-thread adds + 2 shuffles instead of 4 shuffle rounds).  A block is
// 32 channels x 4 lanes.  Tile-entry states come from a carry pass:
//   fwd:  pass 1 per tile (decay product, local end state) -> sequential carry over tiles ->
//         pass 2 per tile produces y (and keeps the tile-entry states for backward)
//   bwd:  tiles in parallel using the stored entry states (the state history of a tile is rebuilt into
//         registers 8 steps at a time); the gradient flowing between tiles (dh at the tile boundary) is
//         resolved with the same two-pass scheme in reverse.
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kSN = 16;        // d_state
constexpr int kSQ = 4;         // states per lane
constexpr int kSTile = 32;     // tokens per tile
constexpr int kSChan = 32;     // channels per block (32 x 4 lanes = 128 threads)
constexpr int kSThreads = kSChan * (kSN / kSQ);
constexpr float kLog2e = 1.4426950408889634f;

struct SScanArgs {
  const bf16* u; long long ldu;
  const bf16* delta; long long ldd;
  const float* A;               // [Dn, N]
  const bf16* Bm; long long ldb;
  const bf16* Cm; long long ldc;
  const float* D;               // [Dn] or null
  const bf16* z; long long ldz; // or null
  const float* delta_bias;      // [Dn] or null
  int Bsz, T, Dn, ntiles, softplus;
};

// sum over a half-warp (single-token update kernel: one lane per state)
__device__ __forceinline__ float half_warp_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// sum over the 4 lanes of a channel (lanes 4c .. 4c+3 of a warp)
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}
// sum over the 8 channels of a warp for a fixed state quad (lanes q, q+4, ..., q+28)
__device__ __forceinline__ float chan8_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  return v;
}

// Per-token inputs of one (batch, tile, 32-channel block), staged in shared memory so that the 32-step
// recurrences below never wait on global memory inside the loop (dt = 0 / u = 0 pad the ragged end: the
// state then just carries through).
struct SScanTile {
  float dt[kSTile][kSChan], u[kSTile][kSChan];
  float B[kSTile][kSN], C[kSTile][kSN];
};
__device__ __forceinline__ void sscan_stage(const SScanArgs& a, SScanTile& s, int b, int t0, int nt, int d0,
                                            bool want_u, bool want_b, bool want_c) {
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += kSThreads) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    float dt = 0.f, uu = 0.f;
    if (i < nt && dch < a.Dn) {
      const size_t row = (size_t)b * a.T + t0 + i;
      dt = __bfloat162float(a.delta[row * a.ldd + dch]) + (a.delta_bias ? a.delta_bias[dch] : 0.f);
      if (a.softplus) dt = softplus_f(dt);
      if (want_u) uu = __bfloat162float(a.u[row * a.ldu + dch]);
    }
    s.dt[i][cch] = dt; s.u[i][cch] = uu;
  }
  for (int idx = threadIdx.x; idx < kSTile * kSN; idx += kSThreads) {
    const int i = idx / kSN, nn = idx % kSN;
    float bv = 0.f, cv = 0.f;
    if (i < nt) {
      const size_t row = (size_t)b * a.T + t0 + i;
      if (want_b) bv = __bfloat162float(a.Bm[row * a.ldb + nn]);
      if (want_c) cv = __bfloat162float(a.Cm[row * a.ldc + nn]);
    }
    s.B[i][nn] = bv; s.C[i][nn] = cv;
  }
}

struct Lane {
  int q, dl, d;       // state quad, channel inside the block, channel (clamped to 0 when out of range)
  bool active;
  float2 An2[2];      // A[d, 4q + {0,1}] and A[d, 4q + {2,3}], times log2(e): exp(dt A) = ex2(dt * An2)
};
// exp2 of both halves of a pair (one MUFU each; everything around it is packed fp32: FMUL2 / FFMA2)
__device__ __forceinline__ float2 ex2_pair(float2 x) { return f2(ex2_approx(x.x), ex2_approx(x.y)); }
__device__ __forceinline__ Lane make_lane(const SScanArgs& a, int d0) {
  Lane l;
  l.q = threadIdx.x & 3;
  l.dl = threadIdx.x >> 2;
  const int dr = d0 + l.dl;
  l.active = dr < a.Dn;
  l.d = l.active ? dr : 0;
  const float4 A4 = *reinterpret_cast<const float4*>(a.A + (size_t)l.d * kSN + l.q * kSQ);
  l.An2[0] = f2(A4.x * kLog2e, A4.y * kLog2e);
  l.An2[1] = f2(A4.z * kLog2e, A4.w * kLog2e);
  return l;
}

// pass 1: per (b, tile, d, n): P = log of prod_t exp(dt_t A) = A sum_t dt_t, S = local end state from zero
__global__ void __launch_bounds__(kSThreads)
sscan_local_kernel(SScanArgs a, float* __restrict__ P, float* __restrict__ S) {
  __shared__ SScanTile st;
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, nt = min(t0 + kSTile, a.T) - t0;
  const int d0 = blockIdx.x * kSChan;
  sscan_stage(a, st, b, t0, nt, d0, true, true, false);
  __syncthreads();
  const Lane l = make_lane(a, d0);
  float2 h[2] = {f2(0.f, 0.f), f2(0.f, 0.f)};
  float sum_dt = 0.f;
#pragma unroll 8
  for (int i = 0; i < kSTile; ++i) {
    const float dt = st.dt[i][l.dl];
    const float dtu = dt * st.u[i][l.dl];
    const float4 b4 = *reinterpret_cast<const float4*>(&st.B[i][l.q * kSQ]);
    const float2 bb[2] = {f2(b4.x, b4.y), f2(b4.z, b4.w)};
    sum_dt += dt;
#pragma unroll
    for (int p = 0; p < 2; ++p) h[p] = fma2(ex2_pair(mul2(f2(dt, dt), l.An2[p])), h[p], mul2(f2(dtu, dtu), bb[p]));
  }
  if (!l.active) return;
  const size_t o = (((size_t)b * a.ntiles + tile) * a.Dn + l.d) * kSN + l.q * kSQ;
  const float k = sum_dt * 0.6931471805599453f;   // P stays in natural-log space (carry kernels use exp)
  *reinterpret_cast<float4*>(P + o) = make_float4(k * l.An2[0].x, k * l.An2[0].y, k * l.An2[1].x, k * l.An2[1].y);
  *reinterpret_cast<float4*>(S + o) = make_float4(h[0].x, h[0].y, h[1].x, h[1].y);
}

// carry: hin[tile] = state entering the tile  (forward order), one thread per (b, d, n)
__global__ void sscan_carry_fwd_kernel(const float* __restrict__ P, const float* __restrict__ S, float* __restrict__ hin,
                                       int Bsz, int ntiles, int DnN) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Bsz * DnN) return;
  const int b = i / DnN, e = i % DnN;
  float h = 0.f;
  for (int c = 0; c < ntiles; ++c) {
    const size_t o = ((size_t)b * ntiles + c) * DnN + e;
    hin[o] = h;
    h = fast_exp(P[o]) * h + S[o];
  }
}

// pass 2: output.  Four lanes per channel, 4 states each.
__global__ void __launch_bounds__(kSThreads)
sscan_fwd_kernel(SScanArgs a, const float* __restrict__ hin, bf16* __restrict__ y, long long ldy) {
  __shared__ SScanTile st;
  __shared__ float s_g[kSTile][kSChan];   // silu(z) on the way in, y on the way out
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, nt = min(t0 + kSTile, a.T) - t0;
  const int d0 = blockIdx.x * kSChan;
  sscan_stage(a, st, b, t0, nt, d0, true, true, true);
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += kSThreads) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    float gz = 1.f;
    if (a.z && i < nt && dch < a.Dn) gz = silu_f(__bfloat162float(a.z[((size_t)b * a.T + t0 + i) * a.ldz + dch]));
    s_g[i][cch] = gz;
  }
  __syncthreads();
  const Lane l = make_lane(a, d0);
  const float Dd = a.D ? a.D[l.d] : 0.f;
  float2 h[2] = {f2(0.f, 0.f), f2(0.f, 0.f)};
  if (l.active) {
    const float4 h4 = *reinterpret_cast<const float4*>(hin + (((size_t)b * a.ntiles + tile) * a.Dn + l.d) * kSN + l.q * kSQ);
    h[0] = f2(h4.x, h4.y); h[1] = f2(h4.z, h4.w);
  }
#pragma unroll 8
  for (int i = 0; i < kSTile; ++i) {
    const float dt = st.dt[i][l.dl], uu = st.u[i][l.dl];
    const float dtu = dt * uu;
    const float4 b4 = *reinterpret_cast<const float4*>(&st.B[i][l.q * kSQ]);
    const float4 c4 = *reinterpret_cast<const float4*>(&st.C[i][l.q * kSQ]);
    const float2 bb[2] = {f2(b4.x, b4.y), f2(b4.z, b4.w)}, cc[2] = {f2(c4.x, c4.y), f2(c4.z, c4.w)};
    float2 yp = f2(0.f, 0.f);
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      h[p] = fma2(ex2_pair(mul2(f2(dt, dt), l.An2[p])), h[p], mul2(f2(dtu, dtu), bb[p]));
      yp = fma2(cc[p], h[p], yp);
    }
    const float yy = quad_sum(yp.x + yp.y);
    if (l.q == 0) s_g[i][l.dl] = (yy + Dd * uu) * s_g[i][l.dl];   // only this lane reads / writes the slot
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < nt * kSChan; idx += kSThreads) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    if (dch < a.Dn) y[((size_t)b * a.T + t0 + i) * ldy + dch] = __float2bfloat16(s_g[i][cch]);
  }
}

// Backward, pass 1: per tile, with dh entering from the right assumed zero, compute the local
// gradient that leaves the tile on the left: G = dh_{t0-1 <- tile} and the decay product (log).
// Needs only C, dy, z, dt: dh_{t} = C_t dyy_t + exp(dt_{t+1} A) dh_{t+1}.
__global__ void __launch_bounds__(kSThreads)
sscan_bwd_local_kernel(SScanArgs a, const bf16* __restrict__ dy, long long lddy, float* __restrict__ Gl) {
  __shared__ SScanTile st;                 // u slot holds dyy = dy * silu(z)
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, nt = min(t0 + kSTile, a.T) - t0;
  const int d0 = blockIdx.x * kSChan;
  sscan_stage(a, st, b, t0, nt, d0, false, false, true);
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += kSThreads) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    float dyy = 0.f;
    if (i < nt && dch < a.Dn) {
      const size_t row = (size_t)b * a.T + t0 + i;
      dyy = __bfloat162float(dy[row * lddy + dch]);
      // y = (ssm + D u) * silu(z): the part flowing into the SSM output is dy * silu(z)
      if (a.z) dyy *= silu_f(__bfloat162float(a.z[row * a.ldz + dch]));
    }
    st.u[i][cch] = dyy;
  }
  __syncthreads();
  const Lane l = make_lane(a, d0);
  // gradient w.r.t. the state *entering* the tile, from the tile's own outputs only
  float2 g[2] = {f2(0.f, 0.f), f2(0.f, 0.f)};  // dL/dh_t accumulated right-to-left (padding steps: dt = 0, dyy = 0 leave it unchanged)
#pragma unroll 8
  for (int i = kSTile - 1; i >= 0; --i) {
    const float dt = st.dt[i][l.dl], dyy = st.u[i][l.dl];
    const float4 c4 = *reinterpret_cast<const float4*>(&st.C[i][l.q * kSQ]);
    const float2 cc[2] = {f2(c4.x, c4.y), f2(c4.z, c4.w)};
#pragma unroll
    for (int p = 0; p < 2; ++p)   // now dL/dh_{t-1} through h_t
      g[p] = mul2(fma2(cc[p], f2(dyy, dyy), g[p]), ex2_pair(mul2(f2(dt, dt), l.An2[p])));
  }
  if (!l.active) return;
  *reinterpret_cast<float4*>(Gl + (((size_t)b * a.ntiles + tile) * a.Dn + l.d) * kSN + l.q * kSQ) =
      make_float4(g[0].x, g[0].y, g[1].x, g[1].y);
}

// reverse carry: gin[tile] = dL/dh at the tile's right boundary (gradient w.r.t. the state leaving
// the tile), given each tile's local left-going gradient and decay product.
__global__ void sscan_carry_bwd_kernel(const float* __restrict__ P, const float* __restrict__ Gl, float* __restrict__ gin,
                                       int Bsz, int ntiles, int DnN) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Bsz * DnN) return;
  const int b = i / DnN, e = i % DnN;
  float g = 0.f;  // gradient w.r.t. the state leaving the last tile = 0
  for (int c = ntiles - 1; c >= 0; --c) {
    const size_t o = ((size_t)b * ntiles + c) * DnN + e;
    gin[o] = g;
    g = Gl[o] + fast_exp(P[o]) * g;  // gradient w.r.t. the state entering tile c
  }
}

// Backward, pass 2: everything.  Rebuilds the tile's states from hin (8 steps at a time, in registers), then
// walks right-to-left starting from gin.
//
// All per-token inputs of the tile are staged in shared memory first and all per-token outputs leave through
// shared memory afterwards (they overwrite the input slots they no longer need), so the recurrences touch no
// global memory.  The arithmetic runs on state PAIRS (FFMA2 / FMUL2: two fp32 lanes per issued instruction).
// dB[t, n] / dC[t, n] are sums over all channels: every thread drops its 4 + 4 terms of a step into a shared
// [4 steps][32 channels][16 states] buffer (one 16-byte store each), and after each block of 4 steps the CTA sums
// the 32 channels with conflict-free loads — 18 instructions per thread and step instead of the 48 a shuffle
// reduction over the warp's 8 channels costs — and issues one global atomic per (token, state).
constexpr int kBwdBlk = 4;   // steps per rebuilt block (4 keeps the term buffers at 17 KB -> 5 blocks per SM)
constexpr int kVStep = kSChan * kSN + 16;   // floats per step in the dB / dC term buffers: +16 keeps the two steps a warp sums in different banks
constexpr int kBwdSmem = (6 * kSTile * kSChan + 2 * kSTile * kSN + 2 * kBwdBlk * kVStep) * 4;

__global__ void __launch_bounds__(kSThreads)
sscan_bwd_kernel(SScanArgs a, const bf16* __restrict__ dy, long long lddy, const float* __restrict__ hin,
                 const float* __restrict__ gin, bf16* __restrict__ du, long long lddu, bf16* __restrict__ ddelta,
                 long long lddd, float* __restrict__ dA, float* __restrict__ dBf, float* __restrict__ dCf,
                 float* __restrict__ dD, bf16* __restrict__ dz, long long lddz, float* __restrict__ dbias) {
  extern __shared__ __align__(16) float sm[];
  // inputs; three of them are overwritten by outputs once consumed: s_sig -> d(delta), s_u -> du, s_dsz -> dz
  float (*s_dt)[kSChan] = reinterpret_cast<float (*)[kSChan]>(sm);
  float (*s_u)[kSChan] = s_dt + kSTile;
  float (*s_dy)[kSChan] = s_u + kSTile;
  float (*s_sig)[kSChan] = s_dy + kSTile;
  float (*s_sz)[kSChan] = s_sig + kSTile;
  float (*s_dsz)[kSChan] = s_sz + kSTile;
  float (*s_B)[kSN] = reinterpret_cast<float (*)[kSN]>(s_dsz + kSTile);
  float (*s_C)[kSN] = s_B + kSTile;
  float* s_vb = reinterpret_cast<float*>(s_C + kSTile);    // [4 steps][32 channels][16 states] per-channel terms
  float* s_vc = s_vb + kBwdBlk * kVStep;
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, t1 = min(t0 + kSTile, a.T);
  const int nt = t1 - t0;
  const int d0 = blockIdx.x * kSChan;
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += kSThreads) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    float dt = 0.f, uu = 0.f, dyv = 0.f, sig = 0.f, sz = 1.f, dsz = 0.f;
    if (i < nt && dch < a.Dn) {
      const size_t row = (size_t)b * a.T + t0 + i;
      const float raw = __bfloat162float(a.delta[row * a.ldd + dch]) + (a.delta_bias ? a.delta_bias[dch] : 0.f);
      dt = a.softplus ? softplus_f(raw) : raw;
      sig = a.softplus ? sigmoid_f(raw) : 1.f;
      uu = __bfloat162float(a.u[row * a.ldu + dch]);
      dyv = __bfloat162float(dy[row * lddy + dch]);
      if (a.z) silu_both(__bfloat162float(a.z[row * a.ldz + dch]), sz, dsz);
    }
    s_dt[i][cch] = dt; s_u[i][cch] = uu; s_dy[i][cch] = dyv; s_sig[i][cch] = sig; s_sz[i][cch] = sz; s_dsz[i][cch] = dsz;
  }
  for (int idx = threadIdx.x; idx < kSTile * kSN; idx += kSThreads) {
    const int i = idx / kSN, nn = idx % kSN;
    float bv = 0.f, cv = 0.f;
    if (i < nt) {
      const size_t row = (size_t)b * a.T + t0 + i;
      bv = __bfloat162float(a.Bm[row * a.ldb + nn]);
      cv = __bfloat162float(a.Cm[row * a.ldc + nn]);
    }
    s_B[i][nn] = bv; s_C[i][nn] = cv;
  }
  __syncthreads();

  const Lane l = make_lane(a, d0);     // inactive channels run along on channel 0 with zero inputs
  const float Dd = a.D ? a.D[l.d] : 0.f;
  const size_t o = (((size_t)b * a.ntiles + tile) * a.Dn + l.d) * kSN + l.q * kSQ;
  const float2 An[2] = {mul2(l.An2[0], f2(0.6931471805599453f, 0.6931471805599453f)),
                        mul2(l.An2[1], f2(0.6931471805599453f, 0.6931471805599453f))};   // A in natural units
  // State history: keeping all 32 x 4 states of a thread live costs > 255 registers.  Instead a first sweep keeps
  // the state entering each block of 4 steps, and the backward walk rebuilds one block's states at a time
  // (5 x 4 values) right before it consumes them: a few more instructions per state, half the registers.
  constexpr int kBlk = kBwdBlk, kNBlk = kSTile / kBlk;
  float2 ck[kNBlk][2];   // indexed by the (rolled) block loops: lives in local memory, 16 values, touched once per block
  {
    const float4 h4 = l.active ? *reinterpret_cast<const float4*>(hin + o) : make_float4(0.f, 0.f, 0.f, 0.f);
    float2 h[2] = {f2(h4.x, h4.y), f2(h4.z, h4.w)};
    // (the block loops are deliberately NOT unrolled: unrolled, the compiler recognises the rebuild below as the same
    // expressions as this sweep and keeps all 32 x 4 states live instead of recomputing them)
#pragma unroll 1
    for (int blk = 0; blk < kNBlk; ++blk) {
      ck[blk][0] = h[0]; ck[blk][1] = h[1];
      if (blk == kNBlk - 1) break;   // the last block's states are rebuilt below anyway
#pragma unroll
      for (int s2 = 0; s2 < kBlk; ++s2) {
        const int i = blk * kBlk + s2;
        const float dt = s_dt[i][l.dl];   // 0 past the end of the tile: h carries through unchanged
        const float dtu = dt * s_u[i][l.dl];
        const float4 b4 = *reinterpret_cast<const float4*>(&s_B[i][l.q * kSQ]);
        const float2 bb[2] = {f2(b4.x, b4.y), f2(b4.z, b4.w)};
#pragma unroll
        for (int p = 0; p < 2; ++p) h[p] = fma2(ex2_pair(mul2(f2(dt, dt), l.An2[p])), h[p], mul2(f2(dtu, dtu), bb[p]));
      }
    }
  }
  float2 g[2] = {f2(0.f, 0.f), f2(0.f, 0.f)};  // dL/dh_{t1-1} coming from the right
  if (l.active) {
    const float4 g4 = *reinterpret_cast<const float4*>(gin + o);
    g[0] = f2(g4.x, g4.y); g[1] = f2(g4.z, g4.w);
  }
  float2 dA_acc[2] = {f2(0.f, 0.f), f2(0.f, 0.f)};
  float dD_acc = 0.f, db_acc = 0.f;
#pragma unroll 1
  for (int blk = kNBlk - 1; blk >= 0; --blk) {
    float2 hist[kBlk + 1][2];
    hist[0][0] = ck[blk][0]; hist[0][1] = ck[blk][1];
#pragma unroll
    for (int s2 = 0; s2 < kBlk; ++s2) {
      const int i = blk * kBlk + s2;
      const float dt = s_dt[i][l.dl];
      const float dtu = dt * s_u[i][l.dl];
      const float4 b4 = *reinterpret_cast<const float4*>(&s_B[i][l.q * kSQ]);
      const float2 bb[2] = {f2(b4.x, b4.y), f2(b4.z, b4.w)};
#pragma unroll
      for (int p = 0; p < 2; ++p)
        hist[s2 + 1][p] = fma2(ex2_pair(mul2(f2(dt, dt), l.An2[p])), hist[s2][p], mul2(f2(dtu, dtu), bb[p]));
    }
#pragma unroll
    for (int s2 = kBlk - 1; s2 >= 0; --s2) {
      const int i = blk * kBlk + s2;
      // steps past the end of the tile have dt = u = dy = 0 (staged so): they leave g unchanged and contribute zeros
      const float dt = s_dt[i][l.dl], uu = s_u[i][l.dl];
      const float dyv = s_dy[i][l.dl], sz = s_sz[i][l.dl];
      const float4 b4 = *reinterpret_cast<const float4*>(&s_B[i][l.q * kSQ]);
      const float4 c4 = *reinterpret_cast<const float4*>(&s_C[i][l.q * kSQ]);
      const float2 bb[2] = {f2(b4.x, b4.y), f2(b4.z, b4.w)}, cc[2] = {f2(c4.x, c4.y), f2(c4.z, c4.w)};
      const float dyy = dyv * sz;                         // gradient into (ssm + D u)
      const float2 dyy2 = f2(dyy, dyy), uu2 = f2(uu, uu), dt2 = f2(dt, dt);
      float2 ssm_p = f2(0.f, 0.f), ddt_p = f2(0.f, 0.f), du_p = f2(0.f, 0.f);
      float2 vb[2], vc[2];
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const float2 h_t = hist[s2 + 1][p], h_prev = hist[s2][p];
        ssm_p = fma2(cc[p], h_t, ssm_p);
        g[p] = fma2(cc[p], dyy2, g[p]);                   // dL/dh_t (total)
        const float2 dec = ex2_pair(mul2(dt2, l.An2[p]));
        const float2 dh = mul2(dec, h_prev);
        ddt_p = fma2(g[p], fma2(An[p], dh, mul2(bb[p], uu2)), ddt_p);
        du_p = fma2(g[p], bb[p], du_p);
        const float2 gdt = mul2(g[p], dt2);
        dA_acc[p] = fma2(gdt, dh, dA_acc[p]);
        vb[p] = mul2(gdt, uu2);                           // dB[t, n] terms of this channel
        vc[p] = mul2(dyy2, h_t);                          // dC[t, n] terms of this channel
        g[p] = mul2(g[p], dec);                           // becomes dL/dh_{t-1} through h_t
      }
      const float ssm = quad_sum(ssm_p.x + ssm_p.y), ddt = quad_sum(ddt_p.x + ddt_p.y), duu = quad_sum(du_p.x + du_p.y) * dt;
      *reinterpret_cast<float4*>(s_vb + s2 * kVStep + l.dl * kSN + l.q * kSQ) = make_float4(vb[0].x, vb[0].y, vb[1].x, vb[1].y);
      *reinterpret_cast<float4*>(s_vc + s2 * kVStep + l.dl * kSN + l.q * kSQ) = make_float4(vc[0].x, vc[0].y, vc[1].x, vc[1].y);
      if (l.q == 0) {
        const float draw = ddt * s_sig[i][l.dl];
        s_sig[i][l.dl] = draw;                            // d(delta) out
        db_acc += draw;
        s_u[i][l.dl] = duu + Dd * dyy;                    // du out (s_u of this step is dead for every lane of the warp)
        dD_acc += dyy * uu;
        s_dsz[i][l.dl] = dyv * (ssm + Dd * uu) * s_dsz[i][l.dl];   // dz out
      }
    }
    // sum the 32 channels of this block of steps: thread -> (step, state), consecutive lanes = consecutive states
    __syncthreads();
    {
      // 128 threads = 4 steps x 16 states x {dB, dC}: each sums its 32 channel terms and adds them to the global
      // fp32 [B*T, N] accumulator (one atomic per (token, state) and block of 32 channels)
      const int which = threadIdx.x >> 6, s2 = (threadIdx.x >> 4) & 3, nn = threadIdx.x & 15;
      const float* pv = (which ? s_vc : s_vb) + s2 * kVStep + nn;
      float sv = 0.f;
#pragma unroll 8
      for (int c = 0; c < kSChan; ++c) sv += pv[c * kSN];
      const int i = blk * kBlk + s2;
      if (i < nt) atomicAdd((which ? dCf : dBf) + ((size_t)b * a.T + t0 + i) * kSN + nn, sv);
    }
    __syncthreads();
  }
  if (l.active) {
    float* dAp = dA + (size_t)l.d * kSN + l.q * kSQ;
    atomicAdd(dAp + 0, dA_acc[0].x); atomicAdd(dAp + 1, dA_acc[0].y);
    atomicAdd(dAp + 2, dA_acc[1].x); atomicAdd(dAp + 3, dA_acc[1].y);
    if (l.q == 0) {
      if (dD) atomicAdd(dD + l.d, dD_acc);
      if (dbias) atomicAdd(dbias + l.d, db_acc);
    }
  }
  for (int idx = threadIdx.x; idx < nt * kSChan; idx += kSThreads) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    if (dch < a.Dn) {
      const size_t row = (size_t)b * a.T + t0 + i;
      ddelta[row * lddd + dch] = __float2bfloat16(s_sig[i][cch]);
      du[row * lddu + dch] = __float2bfloat16(s_u[i][cch]);
      if (a.z) dz[row * lddz + dch] = __float2bfloat16(s_dsz[i][cch]);
    }
  }
}

// K23: single-token update.  state [B, Dn, N] fp32 in place; one half-warp per (b, d).
__global__ void sscan_update_kernel(float* __restrict__ state, const bf16* __restrict__ x, long long ldx,
                                    const bf16* __restrict__ delta, long long ldd, const float* __restrict__ A,
                                    const bf16* __restrict__ Bm, long long ldb, const bf16* __restrict__ Cm, long long ldc,
                                    const float* __restrict__ D, const bf16* __restrict__ z, long long ldz,
                                    const float* __restrict__ bias, bf16* __restrict__ y, long long ldy, int Bsz, int Dn,
                                    int softplus) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = gid & (kSN - 1), bd = gid >> 4;
  if (bd >= Bsz * Dn) return;
  const int b = bd / Dn, d = bd % Dn;
  float dt = __bfloat162float(delta[(size_t)b * ldd + d]) + (bias ? bias[d] : 0.f);
  if (softplus) dt = softplus_f(dt);
  const float uu = __bfloat162float(x[(size_t)b * ldx + d]);
  float h = state[(size_t)bd * kSN + n];
  h = fast_exp(dt * A[d * kSN + n]) * h + dt * __bfloat162float(Bm[(size_t)b * ldb + n]) * uu;
  state[(size_t)bd * kSN + n] = h;
  float yy = half_warp_sum(__bfloat162float(Cm[(size_t)b * ldc + n]) * h);
  if (n == 0) {
    if (D) yy += D[d] * uu;
    if (z) yy *= silu_f(__bfloat162float(z[(size_t)b * ldz + d]));
    y[(size_t)b * ldy + d] = __float2bfloat16(yy);
  }
}

static SScanArgs make_args(const void* u, long long ldu, const void* delta, long long ldd, const float* A,
                           const void* Bm, long long ldb, const void* Cm, long long ldc, const float* D, const void* z,
                           long long ldz, const float* bias, int Bsz, int T, int Dn, int softplus) {
  SScanArgs a;
  a.u = (const bf16*)u; a.ldu = ldu; a.delta = (const bf16*)delta; a.ldd = ldd; a.A = A;
  a.Bm = (const bf16*)Bm; a.ldb = ldb; a.Cm = (const bf16*)Cm; a.ldc = ldc; a.D = D;
  a.z = (const bf16*)z; a.ldz = ldz; a.delta_bias = bias;
  a.Bsz = Bsz; a.T = T; a.Dn = Dn; a.ntiles = cdiv(T, kSTile); a.softplus = softplus;
  return a;
}

}  // namespace mb
using namespace mb;

// scratch layout (fp32): P | S | hin, each [B, ntiles, Dn, N]
extern "C" long long mb_sscan_scratch_floats(int B, int T, int Dn, int N) {
  return 3ll * B * cdiv(T, kSTile) * Dn * N;
}

extern "C" int mb_sscan_fwd(const void* u, long long ldu, const void* delta, long long ldd, const float* A,
                            const void* Bm, long long ldb, const void* Cm, long long ldc, const float* D,
                            const void* z, long long ldz, const float* delta_bias, void* y, long long ldy,
                            float* scratch, int B, int T, int Dn, int N, int softplus, cudaStream_t stream) {
  MB_REQUIRE(N == kSN, "sscan: d_state must be %d (got %d)", kSN, N);
  if (B * T * Dn == 0) return 0;
  SScanArgs a = make_args(u, ldu, delta, ldd, A, Bm, ldb, Cm, ldc, D, z, ldz, delta_bias, B, T, Dn, softplus);
  const size_t one = (size_t)B * a.ntiles * Dn * N;
  float *P = scratch, *S = scratch + one, *hin = scratch + 2 * one;
  dim3 grid(cdiv(Dn, kSChan), a.ntiles, B);
  sscan_local_kernel<<<grid, kSThreads, 0, stream>>>(a, P, S);
  if (check_launch("sscan_fwd/local")) return 1;
  sscan_carry_fwd_kernel<<<cdiv((long long)B * Dn * N, 256), 256, 0, stream>>>(P, S, hin, B, a.ntiles, Dn * N);
  if (check_launch("sscan_fwd/carry")) return 1;
  sscan_fwd_kernel<<<grid, kSThreads, 0, stream>>>(a, hin, (bf16*)y, ldy);
  return check_launch("sscan_fwd");
}

// scratch: the forward's P | S | hin must be passed back unchanged; `scratch2` (2 x [B,ntiles,Dn,N]) is
// workspace.  dA [Dn,N], dB/dC fp32 [B*T, N], dD [Dn], dbias [Dn] are ACCUMULATED into (pre-zero them).
extern "C" int mb_sscan_bwd(const void* u, long long ldu, const void* delta, long long ldd, const float* A,
                            const void* Bm, long long ldb, const void* Cm, long long ldc, const float* D,
                            const void* z, long long ldz, const float* delta_bias, const void* dy, long long lddy,
                            const float* scratch, float* scratch2, void* du, long long lddu, void* ddelta,
                            long long lddd, float* dA, float* dB, float* dC, float* dD, void* dz, long long lddz,
                            float* dbias, int B, int T, int Dn, int N, int softplus, cudaStream_t stream) {
  MB_REQUIRE(N == kSN, "sscan: d_state must be %d (got %d)", kSN, N);
  if (B * T * Dn == 0) return 0;
  SScanArgs a = make_args(u, ldu, delta, ldd, A, Bm, ldb, Cm, ldc, D, z, ldz, delta_bias, B, T, Dn, softplus);
  const size_t one = (size_t)B * a.ntiles * Dn * N;
  const float *P = scratch, *hin = scratch + 2 * one;
  float *Gl = scratch2, *gin = scratch2 + one;
  dim3 grid(cdiv(Dn, kSChan), a.ntiles, B);
  sscan_bwd_local_kernel<<<grid, kSThreads, 0, stream>>>(a, (const bf16*)dy, lddy, Gl);
  if (check_launch("sscan_bwd/local")) return 1;
  sscan_carry_bwd_kernel<<<cdiv((long long)B * Dn * N, 256), 256, 0, stream>>>(P, Gl, gin, B, a.ntiles, Dn * N);
  if (check_launch("sscan_bwd/carry")) return 1;
  static bool attr_set = false;
  if (!attr_set) { MB_CUDA(cudaFuncSetAttribute(sscan_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem)); attr_set = true; }
  sscan_bwd_kernel<<<grid, kSThreads, kBwdSmem, stream>>>(a, (const bf16*)dy, lddy, hin, gin, (bf16*)du, lddu, (bf16*)ddelta,
                                                      lddd, dA, dB, dC, dD, (bf16*)dz, lddz, dbias);
  return check_launch("sscan_bwd");
}

extern "C" int mb_sscan_update(float* state, const void* x, long long ldx, const void* delta, long long ldd,
                               const float* A, const void* Bm, long long ldb, const void* Cm, long long ldc,
                               const float* D, const void* z, long long ldz, const float* delta_bias, void* y,
                               long long ldy, int B, int Dn, int N, int softplus, cudaStream_t stream) {
  MB_REQUIRE(N == kSN, "sscan_update: d_state must be %d (got %d)", kSN, N);
  if (B * Dn == 0) return 0;
  const long long threads = (long long)B * Dn * kSN;
  sscan_update_kernel<<<cdiv(threads, 256), 256, 0, stream>>>(state, (const bf16*)x, ldx, (const bf16*)delta, ldd, A,
                                                             (const bf16*)Bm, ldb, (const bf16*)Cm, ldc, D, (const bf16*)z,
                                                             ldz, delta_bias, (bf16*)y, ldy, B, Dn, softplus);
  return check_launch("sscan_update");
}
