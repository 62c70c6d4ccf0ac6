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
// into tiles of 64 tokens; one half-warp owns one channel of one (batch, tile) with one lane per
// state n, so the state history of a tile lives in registers (needed by backward) and the sum
// over n is four shuffles.  Tile-entry states come from a carry pass:
//   fwd:  pass 1 per tile (decay product, local end state) -> sequential carry over tiles ->
//         pass 2 per tile produces y (and keeps the tile-entry states for backward)
//   bwd:  tiles in parallel using the stored entry states; the gradient flowing between tiles
//         (dh at the tile boundary) is resolved with the same two-pass scheme in reverse.
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kSN = 16;        // d_state
constexpr int kSTile = 64;     // tokens per tile
constexpr int kSChan = 8;      // channels per block (8 half-warps = 128 threads)

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

__device__ __forceinline__ float half_warp_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Per-token inputs of one (batch, tile, 8-channel block), staged in shared memory so that the 64-step
// recurrences below never wait on global memory inside the loop (dt = 0 / u = 0 pad the ragged end: the
// state then just carries through).
struct SScanTile {
  float dt[kSTile][kSChan], u[kSTile][kSChan];
  float B[kSTile][kSN], C[kSTile][kSN];
};
__device__ __forceinline__ void sscan_stage(const SScanArgs& a, SScanTile& s, int b, int t0, int nt, int d0,
                                            bool want_u, bool want_b, bool want_c) {
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += blockDim.x) {
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
  for (int idx = threadIdx.x; idx < kSTile * kSN; idx += blockDim.x) {
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

// pass 1: per (b, tile, d, n): P = prod_t exp(dt_t A), S = local end state from zero
__global__ void __launch_bounds__(kSChan * kSN)
sscan_local_kernel(SScanArgs a, float* __restrict__ P, float* __restrict__ S) {
  __shared__ SScanTile st;
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, nt = min(t0 + kSTile, a.T) - t0;
  const int d0 = blockIdx.x * kSChan;
  sscan_stage(a, st, b, t0, nt, d0, true, true, false);
  __syncthreads();
  const int n = threadIdx.x & (kSN - 1), dl = threadIdx.x >> 4;
  const int d = d0 + dl;
  if (d >= a.Dn) return;
  const float An = a.A[d * kSN + n];
  float h = 0.f, lp = 0.f;
#pragma unroll 8
  for (int i = 0; i < kSTile; ++i) {
    const float dt = st.dt[i][dl];
    const float la = dt * An;
    lp += la;
    h = fast_exp(la) * h + dt * st.B[i][n] * st.u[i][dl];
  }
  const size_t o = (((size_t)b * a.ntiles + tile) * a.Dn + d) * kSN + n;
  P[o] = lp;  // log of the decay product (kept in log space: no underflow surprises)
  S[o] = h;
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

// pass 2: output.  One half-warp per channel, lane = n.
__global__ void __launch_bounds__(kSChan * kSN)
sscan_fwd_kernel(SScanArgs a, const float* __restrict__ hin, bf16* __restrict__ y, long long ldy) {
  __shared__ SScanTile st;
  __shared__ float s_g[kSTile][kSChan];   // silu(z) on the way in, y on the way out
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, nt = min(t0 + kSTile, a.T) - t0;
  const int d0 = blockIdx.x * kSChan;
  sscan_stage(a, st, b, t0, nt, d0, true, true, true);
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += blockDim.x) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    float gz = 1.f;
    if (a.z && i < nt && dch < a.Dn) gz = silu_f(__bfloat162float(a.z[((size_t)b * a.T + t0 + i) * a.ldz + dch]));
    s_g[i][cch] = gz;
  }
  __syncthreads();
  const int n = threadIdx.x & (kSN - 1), dl = threadIdx.x >> 4;
  const int d_real = d0 + dl;
  const bool active = d_real < a.Dn;
  const int d = active ? d_real : 0;
  const float An = a.A[d * kSN + n];
  const float Dd = a.D ? a.D[d] : 0.f;
  float h = active ? hin[(((size_t)b * a.ntiles + tile) * a.Dn + d) * kSN + n] : 0.f;
#pragma unroll 8
  for (int i = 0; i < kSTile; ++i) {
    const float dt = st.dt[i][dl], uu = st.u[i][dl];
    h = fast_exp(dt * An) * h + dt * st.B[i][n] * uu;
    const float yy = half_warp_sum(st.C[i][n] * h);
    if (n == 0) s_g[i][dl] = (yy + Dd * uu) * s_g[i][dl];   // only this lane reads / writes the slot
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < nt * kSChan; idx += blockDim.x) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    if (dch < a.Dn) y[((size_t)b * a.T + t0 + i) * ldy + dch] = __float2bfloat16(s_g[i][cch]);
  }
}

// Backward, pass 1: per tile, with dh entering from the right assumed zero, compute the local
// gradient that leaves the tile on the left: G = dh_{t0-1 <- tile} and the decay product (log).
// Needs only C, dy, z, dt: dh_{t} = C_t dyy_t + exp(dt_{t+1} A) dh_{t+1}.
__global__ void __launch_bounds__(kSChan * kSN)
sscan_bwd_local_kernel(SScanArgs a, const bf16* __restrict__ dy, long long lddy, const float* __restrict__ hin,
                       float* __restrict__ Gl) {
  __shared__ SScanTile st;                 // u slot holds dyy = dy * silu(z)
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, nt = min(t0 + kSTile, a.T) - t0;
  const int d0 = blockIdx.x * kSChan;
  sscan_stage(a, st, b, t0, nt, d0, false, false, true);
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += blockDim.x) {
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
  const int n = threadIdx.x & (kSN - 1), dl = threadIdx.x >> 4;
  const int d = d0 + dl;
  if (d >= a.Dn) return;
  const float An = a.A[d * kSN + n];
  // gradient w.r.t. the state *entering* the tile, from the tile's own outputs only
  float g = 0.f;  // dL/dh_t accumulated right-to-left (padding steps: dt = 0, dyy = 0 leave it unchanged)
#pragma unroll 8
  for (int i = kSTile - 1; i >= 0; --i)
    g = (g + st.C[i][n] * st.u[i][dl]) * fast_exp(st.dt[i][dl] * An);  // after this line g = dL/dh_{t-1} through h_t
  (void)hin;
  Gl[(((size_t)b * a.ntiles + tile) * a.Dn + d) * kSN + n] = g;
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

// Backward, pass 2: everything.  Recomputes the tile's states from hin (kept in registers), then
// walks right-to-left starting from gin.
//
// All per-token inputs of the tile are staged in shared memory first and all per-token outputs leave through
// shared memory afterwards, so the two 64-step recurrences touch no global memory: with loads, stores and
// atomics inside the loop every step paid a full global round trip (5.4 ms per call, 70 % of a Mamba-1 step).
// dB[t, n] / dC[t, n] are sums over all channels: the block sums its 8 channels (warp shuffle + shared-memory
// atomics) and issues one global atomic per (token, state).
__global__ void __launch_bounds__(kSChan * kSN)
sscan_bwd_kernel(SScanArgs a, const bf16* __restrict__ dy, long long lddy, const float* __restrict__ hin,
                 const float* __restrict__ gin, bf16* __restrict__ du, long long lddu, bf16* __restrict__ ddelta,
                 long long lddd, float* __restrict__ dA, float* __restrict__ dBf, float* __restrict__ dCf,
                 float* __restrict__ dD, bf16* __restrict__ dz, long long lddz, float* __restrict__ dbias) {
  __shared__ float s_dt[kSTile][kSChan], s_u[kSTile][kSChan], s_dy[kSTile][kSChan], s_sig[kSTile][kSChan];
  __shared__ float s_sz[kSTile][kSChan], s_dsz[kSTile][kSChan];
  __shared__ float s_o_dt[kSTile][kSChan], s_o_du[kSTile][kSChan], s_o_dz[kSTile][kSChan];
  __shared__ float s_B[kSTile][kSN], s_C[kSTile][kSN], s_dB[kSTile][kSN], s_dC[kSTile][kSN];
  const int tile = blockIdx.y, b = blockIdx.z;
  const int t0 = tile * kSTile, t1 = min(t0 + kSTile, a.T);
  const int nt = t1 - t0;
  const int d0 = blockIdx.x * kSChan;
  for (int idx = threadIdx.x; idx < kSTile * kSChan; idx += blockDim.x) {
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
  for (int idx = threadIdx.x; idx < kSTile * kSN; idx += blockDim.x) {
    const int i = idx / kSN, nn = idx % kSN;
    float bv = 0.f, cv = 0.f;
    if (i < nt) {
      const size_t row = (size_t)b * a.T + t0 + i;
      bv = __bfloat162float(a.Bm[row * a.ldb + nn]);
      cv = __bfloat162float(a.Cm[row * a.ldc + nn]);
    }
    s_B[i][nn] = bv; s_C[i][nn] = cv; s_dB[i][nn] = 0.f; s_dC[i][nn] = 0.f;
  }
  __syncthreads();

  const int n = threadIdx.x & (kSN - 1), dl = threadIdx.x >> 4;
  const int d_real = d0 + dl;
  const bool active = d_real < a.Dn;     // inactive half-warps run along on channel 0 with zero inputs
  const int d = active ? d_real : 0;
  const float An = a.A[d * kSN + n];
  const float Dd = a.D ? a.D[d] : 0.f;
  const size_t o = (((size_t)b * a.ntiles + tile) * a.Dn + d) * kSN + n;
  float hist[kSTile + 1];
  hist[0] = active ? hin[o] : 0.f;
#pragma unroll
  for (int i = 0; i < kSTile; ++i) {
    const float dt = s_dt[i][dl];   // 0 past the end of the tile: h carries through unchanged
    hist[i + 1] = fast_exp(dt * An) * hist[i] + dt * s_B[i][n] * s_u[i][dl];
  }
  float g = active ? gin[o] : 0.f;  // dL/dh_{t1-1} coming from the right
  float dA_acc = 0.f, dD_acc = 0.f, db_acc = 0.f;
#pragma unroll
  for (int i = kSTile - 1; i >= 0; --i) {
    if (i < nt) {   // uniform over the block
      const float dt = s_dt[i][dl], uu = s_u[i][dl], Bn = s_B[i][n], Cn = s_C[i][n];
      const float dyv = s_dy[i][dl], sz = s_sz[i][dl];
      const float dyy = dyv * sz;                         // gradient into (ssm + D u)
      const float h_t = hist[i + 1], h_prev = hist[i];
      const float ssm = half_warp_sum(Cn * h_t);
      g += Cn * dyy;                                      // dL/dh_t (total)
      const float dec = fast_exp(dt * An);
      const float ddt = half_warp_sum(g * (An * dec * h_prev + Bn * uu));
      const float duu = half_warp_sum(g * dt * Bn);
      dA_acc += g * dt * dec * h_prev;
      float vb = g * dt * uu, vc = dyy * h_t;             // summed over channels d (inactive ones contribute 0)
      vb += __shfl_xor_sync(0xffffffffu, vb, 16);
      vc += __shfl_xor_sync(0xffffffffu, vc, 16);
      if ((threadIdx.x & 16) == 0) { atomicAdd(&s_dB[i][n], vb); atomicAdd(&s_dC[i][n], vc); }
      if (n == 0) {
        const float draw = ddt * s_sig[i][dl];
        s_o_dt[i][dl] = draw;
        db_acc += draw;
        s_o_du[i][dl] = duu + Dd * dyy;
        dD_acc += dyy * uu;
        s_o_dz[i][dl] = dyv * (ssm + Dd * uu) * s_dsz[i][dl];
      }
      g *= dec;                                           // becomes dL/dh_{t-1} through h_t
    }
  }
  if (active) {
    atomicAdd(dA + d * kSN + n, dA_acc);
    if (n == 0) {
      if (dD) atomicAdd(dD + d, dD_acc);
      if (dbias) atomicAdd(dbias + d, db_acc);
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < nt * kSChan; idx += blockDim.x) {
    const int i = idx / kSChan, cch = idx % kSChan, dch = d0 + cch;
    if (dch < a.Dn) {
      const size_t row = (size_t)b * a.T + t0 + i;
      ddelta[row * lddd + dch] = __float2bfloat16(s_o_dt[i][cch]);
      du[row * lddu + dch] = __float2bfloat16(s_o_du[i][cch]);
      if (a.z) dz[row * lddz + dch] = __float2bfloat16(s_o_dz[i][cch]);
    }
  }
  for (int idx = threadIdx.x; idx < nt * kSN; idx += blockDim.x) {
    const size_t row = (size_t)b * a.T + t0 + idx / kSN;
    atomicAdd(dBf + row * kSN + (idx % kSN), (&s_dB[0][0])[idx]);
    atomicAdd(dCf + row * kSN + (idx % kSN), (&s_dC[0][0])[idx]);
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
  sscan_local_kernel<<<grid, kSChan * kSN, 0, stream>>>(a, P, S);
  if (check_launch("sscan_fwd/local")) return 1;
  sscan_carry_fwd_kernel<<<cdiv((long long)B * Dn * N, 256), 256, 0, stream>>>(P, S, hin, B, a.ntiles, Dn * N);
  if (check_launch("sscan_fwd/carry")) return 1;
  sscan_fwd_kernel<<<grid, kSChan * kSN, 0, stream>>>(a, hin, (bf16*)y, ldy);
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
  sscan_bwd_local_kernel<<<grid, kSChan * kSN, 0, stream>>>(a, (const bf16*)dy, lddy, hin, Gl);
  if (check_launch("sscan_bwd/local")) return 1;
  sscan_carry_bwd_kernel<<<cdiv((long long)B * Dn * N, 256), 256, 0, stream>>>(P, Gl, gin, B, a.ntiles, Dn * N);
  if (check_launch("sscan_bwd/carry")) return 1;
  sscan_bwd_kernel<<<grid, kSChan * kSN, 0, stream>>>(a, (const bf16*)dy, lddy, hin, gin, (bf16*)du, lddu, (bf16*)ddelta,
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
