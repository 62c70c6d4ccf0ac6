// K1 / K2 / K3: depthwise causal conv1d (+bias +SiLU), channel-last, forward / backward / update.
//
// Replaces causal-conv1d's CUDA kernels.  B200-first
// differences: input and output are channel-last [B, T, C] *strided views* — the kernel reads
// the xBC columns of the in_proj GEMM output in place and the backward writes dx straight into
// the xBC columns of d(zxbcdt) — so no transpose / split / cat copies exist around it.
// Memory-bound; each thread owns 8 (forward) or 4 (backward) channels and walks a tile of the flat [B*T] token
// axis with the sliding window in registers; rows are prefetched with cp.async into a thread-private ring.
#include <type_traits>
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kConvThreads = 128;
constexpr int kConvTile = 64;  // smallest token tile per block (the launcher picks the actual one, see conv_tile)

// Forward.  The kernel is a pure stream (read x once, write y once), so what matters is bytes in flight:
// a thread prefetches its own 16-byte column of the next kFwdDepth batches of U tokens with cp.async into a
// private slice of shared memory (ring of kFwdRing rows).  Nobody else touches that slice, so the only
// synchronisation is the thread's own cp.async.wait_group -- no barriers, no registers parked on loads.
constexpr int kFwdU = 4;                          // tokens per batch (one cp.async group)
constexpr int kFwdDepth = 3;                      // batches in flight (6 measured the same; 3 leaves room for 4 blocks per SM)
constexpr int kFwdMinBlocks = 4;
constexpr int kFwdRing = (kFwdDepth + 1) * kFwdU; // rows in the ring (the batch being read + those in flight)
constexpr int kFwdSmem = kFwdRing * kConvThreads * 16;

__device__ __forceinline__ void cp_async16_zfill(uint32_t saddr, const void* g, bool pred) {
  const int n = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(g), "r"(n) : "memory");
}

template <int W, bool SILU>
__global__ void __launch_bounds__(kConvThreads, kFwdMinBlocks)
conv1d_fwd_kernel(const bf16* __restrict__ x, long long ldx, const float* __restrict__ w,
                  const float* __restrict__ bias, bf16* __restrict__ y, long long ldy, int T, int C, int N,
                  int tile, int ntiles, int ngroups) {
  extern __shared__ __align__(16) uint8_t conv_smem[];
  // work item = (token tile of the flat [B*T] axis, group of 32 column vectors), one per warp: see conv_flat_tiles
  const int item = blockIdx.x * (kConvThreads / 32) + (threadIdx.x >> 5);
  if (item >= ntiles * ngroups) return;
  const int tl = item / ngroups, cg = item - tl * ngroups;
  const int c0 = (cg * 32 + (threadIdx.x & 31)) * 8;
  if (c0 >= C) return;
  constexpr int U = kFwdU;
  float2 wt[4][W], bs[4];  // channel pairs: the taps run as FFMA2 (two fp32 lanes per issued instruction)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
#pragma unroll
    for (int k = 0; k < W; ++k) wt[j][k] = f2(w[(c0 + 2 * j) * W + k], w[(c0 + 2 * j + 1) * W + k]);
    bs[j] = bias != nullptr ? f2(bias[c0 + 2 * j], bias[c0 + 2 * j + 1]) : f2(0.f, 0.f);
  }
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  uint8_t* my = conv_smem + threadIdx.x * 16;                 // row r of the ring lives at my + r * (threads * 16)
  const uint32_t my_s = (uint32_t)__cvta_generic_to_shared(my);
  constexpr int kRowBytes = kConvThreads * 16;
  // a flat tile that crosses a sequence boundary is walked as one segment per sequence (the window restarts)
  const int flat_end = min((tl + 1) * tile, N);
  for (int flat = tl * tile; flat < flat_end;) {
  const int b = flat / T;
  const int t0 = flat - b * T;
  const int t1 = min(T, t0 + (flat_end - flat));
  flat += t1 - t0;
  const bf16* xb = x + ((size_t)b * T) * ldx + c0;
  bf16* yb = y + ((size_t)b * T) * ldy + c0;
  // batch i covers tokens t0 + i*U .. +U-1 and ring rows (i % (kFwdDepth+1)) * U ..; rows outside [0, T) are zero-filled
  auto issue = [&](int i) {
    const int tb = t0 + i * U, slot = (i % (kFwdDepth + 1)) * U;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int t = tb + u;
      cp_async16_zfill(my_s + (slot + u) * kRowBytes, xb + (size_t)min(t, T - 1) * ldx, t < T);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  bf16x8 halo[W - 1];   // x[tb - (W-1) + k], packed
#pragma unroll
  for (int k = 0; k < W - 1; ++k) {
    const int t = t0 - (W - 1) + k;
    halo[k] = ld8(xb + (size_t)max(t, 0) * ldx);
    if (t < 0) halo[k].v[0] = halo[k].v[1] = halo[k].v[2] = halo[k].v[3] = __floats2bfloat162_rn(0.f, 0.f);
  }
  const int nb = (t1 - t0 + U - 1) / U;
#pragma unroll
  for (int i = 0; i < kFwdDepth; ++i) {
    if (i < nb) issue(i);
    else asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int i = 0; i < nb; ++i) {
    asm volatile("cp.async.wait_group %0;" ::"n"(kFwdDepth - 1) : "memory");  // batch i has landed
    const int tb = t0 + i * U, slot = (i % (kFwdDepth + 1)) * U;
    bf16x8 raw[U + W - 1];
#pragma unroll
    for (int k = 0; k < W - 1; ++k) raw[k] = halo[k];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      uint4 t4 = *reinterpret_cast<const uint4*>(my + (slot + u) * kRowBytes);
      memcpy(&raw[W - 1 + u], &t4, 16);
    }
    // refill: batch i + depth goes into the slot freed by batch i - 1 (its rows are already in registers / halo)
    if (i + kFwdDepth < nb) issue(i + kFwdDepth);
    else asm volatile("cp.async.commit_group;" ::: "memory");
#pragma unroll
    for (int k = 0; k < W - 1; ++k) halo[k] = raw[U + k];
    float2 acc[U][4];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[u][j] = bs[j];
#pragma unroll
    for (int r = 0; r < U + W - 1; ++r) {  // each input row feeds up to W outputs
      float2 xr[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) xr[j] = __bfloat1622float2(raw[r].v[j]);
#pragma unroll
      for (int k = 0; k < W; ++k) {
        const int u = r - k;  // out[u] += w[k] * x[u + k]
        if (u >= 0 && u < U) {
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[u][j] = fma2(wt[j][k], xr[j], acc[u][j]);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (tb + u < t1) {
        bf16x8 o;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 v = SILU ? silu2(acc[u][j]) : acc[u][j];
          o.v[j] = __floats2bfloat162_rn(v.x, v.y);
        }
        st8(yb + (size_t)(tb + u) * ldy, o);
      }
    }
  }
  }  // segments
}

// Backward.  Each thread owns 4 channels (one 8-byte vector) and walks its token tile in register
// batches of U tokens with everything statically indexed: the g = dy * silu'(pre) values and the x
// rows a batch shares with the next one are carried in registers, so every x / dy row is loaded
// once and no window is ever shifted.  dw / dbias: one fp32 atomic per (warp tile, channel, tap).
struct __align__(8) bf16x4 { bf162 v[2]; };
__device__ __forceinline__ bf16x4 ld4(const bf16* p) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  bf16x4 r;
  memcpy(&r, &t, 8);
  return r;
}
__device__ __forceinline__ void st4(bf16* p, const float* f) {
  bf16x4 r;
  r.v[0] = __floats2bfloat162_rn(f[0], f[1]);
  r.v[1] = __floats2bfloat162_rn(f[2], f[3]);
  uint2 t;
  memcpy(&t, &r, 8);
  *reinterpret_cast<uint2*>(p) = t;
}
__device__ __forceinline__ void unpack4(const bf16x4& p, float* f) {
  const float2 a = __bfloat1622float2(p.v[0]), b = __bfloat1622float2(p.v[1]);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}

// Backward prefetch ring: like the forward kernel, a thread copies its own column of the next kBwdDepth batches
// (x rows and dy rows) into a private slice of shared memory with cp.async, so the loads of later batches are in
// flight while the ~300 instructions of the current batch issue; nothing is parked in registers.
constexpr int kBwdU = 4;
constexpr int kBwdDepth = 3;
constexpr int kBwdStages = kBwdDepth + 1;
constexpr int kBwdMinBlocks = 3;   // 164 registers, no spills; capping at 128 (4 blocks) spills and measured 12 % slower
template <int NP> constexpr int conv_bwd_smem() { return kBwdStages * 2 * kBwdU * kConvThreads * 4 * NP; }
template <int BYTES>
__device__ __forceinline__ void cp_async_small(uint32_t saddr, const void* g) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(saddr), "l"(g), "n"(BYTES) : "memory");
}

template <bool E> struct ConvEdgeTag { static constexpr bool value = E; };

// NP = channel pairs per thread: 1 (4-byte accesses, ~70 registers, high occupancy) or 2 (8-byte accesses).
template <int W, bool SILU, int NP>
__global__ void __launch_bounds__(kConvThreads, kBwdMinBlocks)
conv1d_bwd_kernel(const bf16* __restrict__ x, long long ldx, const float* __restrict__ w,
                  const float* __restrict__ bias, const bf16* __restrict__ dy, long long lddy,
                  bf16* __restrict__ dx, long long lddx, float* __restrict__ dw,
                  float* __restrict__ dbias, int T, int C, int N, int tile, int ntiles, int ngroups) {
  constexpr int H = W - 1;  // halo
  constexpr int U = 4;      // tokens per register batch
  const int item = blockIdx.x * (kConvThreads / 32) + (threadIdx.x >> 5);   // (flat token tile, 32-thread column group)
  if (item >= ntiles * ngroups) return;
  const int tl = item / ngroups, cg = item - tl * ngroups;
  const int c0 = (cg * 32 + (threadIdx.x & 31)) * (2 * NP);
  if (c0 >= C) return;
  // all arithmetic on channel PAIRS (float2): FFMA2 / FMUL2 / FADD2 halve the issue slots of this
  // instruction-bound kernel (~40 instructions per element before)
  float2 wt[NP][W], bs[NP], dwa[NP][W], dba[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) {
#pragma unroll
    for (int k = 0; k < W; ++k) {
      wt[j][k] = f2(w[(c0 + 2 * j) * W + k], w[(c0 + 2 * j + 1) * W + k]);
      dwa[j][k] = f2(0.f, 0.f);
    }
    bs[j] = bias != nullptr ? f2(bias[c0 + 2 * j], bias[c0 + 2 * j + 1]) : f2(0.f, 0.f);
    dba[j] = f2(0.f, 0.f);
  }
  pdl_wait();      // common.cuh: programmatic dependent launch
  pdl_trigger();
  const int flat_end = min((tl + 1) * tile, N);
  for (int flat = tl * tile; flat < flat_end;) {   // one segment per sequence the flat tile touches
  const int b = flat / T;
  const int t0 = flat - b * T;
  const int t1 = min(T, t0 + (flat_end - flat));
  flat += t1 - t0;
  const bf16* xb = x + ((size_t)b * T) * ldx + c0;
  const bf16* gb = dy + ((size_t)b * T) * lddy + c0;
  bf16* dxb = dx + ((size_t)b * T) * lddx + c0;
  // Branch-free: out-of-range rows read a clamped (valid) row and are zeroed by a multiply, so the
  // compiler can issue a batch's loads back to back (a conditional load ends the basic block and
  // each load would then pay its full latency in turn — ncu: 67 % long-scoreboard stalls).
  // `edge` (a compile-time tag): batches that cannot touch a sequence / tile boundary skip the clamps and masks
  auto load_x = [&](int t, float2* f, auto edge) {
    constexpr bool E = decltype(edge)::value;
    const int tc = E ? min(max(t, 0), T - 1) : t;
    const float m = (!E || (t >= 0 && t < T)) ? 1.f : 0.f;
    if (NP == 2) {
      const bf16x4 p = ld4(xb + (size_t)tc * ldx);
      f[0] = mul2(__bfloat1622float2(p.v[0]), f2(m, m));
      f[NP - 1] = mul2(__bfloat1622float2(p.v[1]), f2(m, m));
    } else {
      const bf162 p = *reinterpret_cast<const bf162*>(xb + (size_t)tc * ldx);
      f[0] = mul2(__bfloat1622float2(p), f2(m, m));
    }
  };
  // g[t] = dy[t] * silu'(pre[t]) from the W x-rows ending at t; also this token's dw / db terms
  auto make_g = [&](int t, const float2 (*xw)[NP], const float2* d2, float2* g, auto edge) {  // xw[k] = x[t - H + k]
    constexpr bool E = decltype(edge)::value;
    const float live = (!E || t < T) ? 1.f : 0.f, own = (!E || t < t1) ? 1.f : 0.f;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      float2 gg = E ? mul2(d2[j], f2(live, live)) : d2[j];
      if (SILU) {
        float2 a = bs[j];
#pragma unroll
        for (int k = 0; k < W; ++k) a = fma2(wt[j][k], xw[k][j], a);
        float2 sv, dsv;
        silu_both2(a, sv, dsv);
        gg = mul2(gg, dsv);
      }
      g[j] = gg;
      const float2 go = E ? mul2(gg, f2(own, own)) : gg;  // only the tile that owns token t accumulates its dw / db terms
      dba[j] = add2(dba[j], go);
#pragma unroll
      for (int k = 0; k < W; ++k) dwa[j][k] = fma2(go, xw[k][j], dwa[j][k]);
    }
  };
  auto load_dy = [&](int t, float2* f, auto edge) {
    const int tc = decltype(edge)::value ? min(t, T - 1) : t;
    if (NP == 2) {
      const bf16x4 p = ld4(gb + (size_t)tc * lddy);
      f[0] = __bfloat1622float2(p.v[0]);
      f[NP - 1] = __bfloat1622float2(p.v[1]);
    } else {
      f[0] = __bfloat1622float2(*reinterpret_cast<const bf162*>(gb + (size_t)tc * lddy));
    }
  };
  ConvEdgeTag<true> edge_t;
  ConvEdgeTag<false> inner_t;

  // Rows live in two rings of 8 slots, x[t] and g[t] in slot (t - t0) & 7.  The loop alternates two unrolled
  // 4-token phases, so every slot index is a compile-time constant and no row is ever moved between registers
  // (a shifted window cost 46 MOVs per 16 elements).
  static_assert(U == kBwdU && H + U <= 8, "H carried rows + U new rows must fit the 8-slot ring");
  extern __shared__ __align__(16) uint8_t conv_bwd_smem_buf[];
  constexpr int VB = 4 * NP;                       // bytes per thread per row
  constexpr int kRowBytes = kConvThreads * VB;
  uint8_t* my = conv_bwd_smem_buf + threadIdx.x * VB;
  const uint32_t my_s = (uint32_t)__cvta_generic_to_shared(my);
  // batch i brings tokens t0 + i*U + H .. + U-1 (clamped to the sequence; masked when consumed) into stage i % kBwdStages
  auto issue = [&](int i) {
    const int tn = t0 + i * U + H, st = (i % kBwdStages) * 2 * U;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int t = min(tn + u, T - 1);
      cp_async_small<VB>(my_s + (st + u) * kRowBytes, xb + (size_t)t * ldx);
      cp_async_small<VB>(my_s + (st + U + u) * kRowBytes, gb + (size_t)t * lddy);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int nb = (t1 - t0 + U - 1) / U;
#pragma unroll
  for (int i = 0; i < kBwdDepth; ++i) {
    if (i < nb) issue(i);
    else asm volatile("cp.async.commit_group;" ::: "memory");
  }
  float2 xr[8][NP], gr[8][NP];
  {
    float2 xp[2 * H][NP];  // x[t0-H .. t0+H)
#pragma unroll
    for (int i = 0; i < 2 * H; ++i) load_x(t0 - H + i, xp[i], edge_t);
#pragma unroll
    for (int i = 0; i < H; ++i) {
      float2 xw[W][NP];
#pragma unroll
      for (int k = 0; k < W; ++k)
#pragma unroll
        for (int jj = 0; jj < NP; ++jj) xw[k][jj] = xp[i + k][jj];
      float2 d2[NP];
      load_dy(t0 + i, d2, edge_t);
      make_g(t0 + i, xw, d2, gr[i], edge_t);
#pragma unroll
      for (int jj = 0; jj < NP; ++jj) xr[i][jj] = xp[H + i][jj];
    }
  }
  auto staged_row = [&](int row, float2* f) {
    if (NP == 2) {
      bf16x4 v;
      const uint2 t2 = *reinterpret_cast<const uint2*>(my + row * kRowBytes);
      memcpy(&v, &t2, 8);
      f[0] = __bfloat1622float2(v.v[0]);
      f[NP - 1] = __bfloat1622float2(v.v[1]);
    } else {
      f[0] = __bfloat1622float2(*reinterpret_cast<const bf162*>(my + row * kRowBytes));
    }
  };
  // one batch: tokens tb+H .. tb+H+U-1 arrive, dx[tb .. tb+U) leaves.  P = (tb - t0) & 7, 0 or 4.
  auto step = [&](int i, auto phase) {
    constexpr int P = decltype(phase)::value;
    const int tb = t0 + i * U, st = (i % kBwdStages) * 2 * U;
    asm volatile("cp.async.wait_group %0;" ::"n"(kBwdDepth - 1) : "memory");   // batch i has landed
    float2 dyb[U][NP];
#pragma unroll
    for (int u = 0; u < U; ++u) { staged_row(st + u, xr[(P + H + u) & 7]); staged_row(st + U + u, dyb[u]); }
    // refill the stage batch i-1 was read from
    if (i + kBwdDepth < nb) issue(i + kBwdDepth);
    else asm volatile("cp.async.commit_group;" ::: "memory");
    auto batch = [&](auto edge) {
      constexpr bool E = decltype(edge)::value;
      if (E) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const float m = tb + H + u < T ? 1.f : 0.f;
#pragma unroll
          for (int jj = 0; jj < NP; ++jj) xr[(P + H + u) & 7][jj] = mul2(xr[(P + H + u) & 7][jj], f2(m, m));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float2 xw[W][NP];
#pragma unroll
        for (int k = 0; k < W; ++k)
#pragma unroll
          for (int jj = 0; jj < NP; ++jj) xw[k][jj] = xr[(P + u + k) & 7][jj];  // x[(tb+H+u) - H + k] = x[tb + u + k]
        make_g(tb + H + u, xw, dyb[u], gr[(P + H + u) & 7], edge);
      }
      // dx[tb+u] = sum_k w[k] * g[tb+u + H - k]
      bf16* op = dxb + (size_t)tb * lddx;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!E || tb + u < t1) {
          float2 o[NP];
#pragma unroll
          for (int jj = 0; jj < NP; ++jj) {
            float2 a = mul2(wt[jj][0], gr[(P + u + H) & 7][jj]);
#pragma unroll
            for (int k = 1; k < W; ++k) a = fma2(wt[jj][k], gr[(P + u + H - k) & 7][jj], a);
            o[jj] = a;
          }
          if (NP == 2) {
            const float of[4] = {o[0].x, o[0].y, o[NP - 1].x, o[NP - 1].y};
            st4(op, of);
          } else {
            *reinterpret_cast<bf162*>(op) = __floats2bfloat162_rn(o[0].x, o[0].y);
          }
        }
        op += lddx;
      }
    };
    if (tb + H + U <= t1) batch(inner_t);   // every token of the batch is inside this tile (hence inside the sequence)
    else batch(edge_t);
  };
  for (int i = 0; i < nb; i += 2) {
    step(i, std::integral_constant<int, 0>{});
    if (i + 1 < nb) step(i + 1, std::integral_constant<int, U>{});
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  }  // segments
#pragma unroll
  for (int jj = 0; jj < NP; ++jj) {
#pragma unroll
    for (int k = 0; k < W; ++k) {
      atomicAdd(dw + (c0 + 2 * jj) * W + k, dwa[jj][k].x);
      atomicAdd(dw + (c0 + 2 * jj + 1) * W + k, dwa[jj][k].y);
    }
    if (dbias != nullptr) { atomicAdd(dbias + c0 + 2 * jj, dba[jj].x); atomicAdd(dbias + c0 + 2 * jj + 1, dba[jj].y); }
  }
}

// Single-token update with a rolling state [B, C, W] (K3; decode path).
template <bool SILU>
__global__ void conv1d_update_kernel(const bf16* __restrict__ x, long long ldx, float* __restrict__ state,
                                     const float* __restrict__ w, const float* __restrict__ bias,
                                     bf16* __restrict__ y, long long ldy, int B, int C, int W) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * C) return;
  const int b = idx / C, c = idx - b * C;
  float* st = state + (size_t)idx * W;
  float a = bias != nullptr ? bias[c] : 0.f;
  for (int k = 0; k < W - 1; ++k) { const float v = st[k + 1]; st[k] = v; a += w[c * W + k] * v; }
  const float xv = __bfloat162float(x[(size_t)b * ldx + c]);
  st[W - 1] = xv;
  a += w[c * W + W - 1] * xv;
  y[(size_t)b * ldy + c] = __float2bfloat16(SILU ? silu_f(a) : a);
}

}  // namespace mb

using namespace mb;

constexpr int kBwdNP = 2;  // channel pairs per backward thread (1 = more occupancy, measured no faster: issue-bound)

// Work decomposition shared by forward and backward.  The unit is a warp: 32 column vectors x one tile of the flat
// [B*T] token axis.  Tiles are as long as possible while the warps still fill whole waves of resident warps --
// per-sequence tiles and 128-thread column blocks left the machine 75 % occupied at C = 1792 (7 or 14 column
// groups never divide a power-of-two block count), worth 20 % of the kernel time.
static void conv_flat_tiles(long long N, int ngroups, int warps_per_sm, int* tile, int* ntiles) {
  const long long slots = (long long)num_sms() * warps_per_sm;
  int best_tile = kConvTile;
  double best = 1e30;
  for (int waves = 1; waves <= 64; ++waves) {
    long long nt = waves * slots / ngroups;
    if (nt < 1) nt = 1;
    long long L = cdiv(N, nt);
    if (L < kConvTile) L = kConvTile;
    nt = cdiv(N, L);
    const double cost = (double)cdiv(nt * ngroups, slots) * (double)(L + 8);   // waves x (tokens + per-tile overhead)
    if (cost < best) { best = cost; best_tile = (int)L; }
    if (L == kConvTile) break;
  }
  *tile = best_tile;
  *ntiles = (int)cdiv(N, (long long)best_tile);
}

// Host-only view of the planner (no kernel launch): lets the work decomposition be checked without a GPU.
extern "C" int mb_conv1d_plan(long long tokens, int column_groups, int warps_per_sm, int* tile, int* ntiles) {
  MB_REQUIRE(tokens > 0 && column_groups > 0 && warps_per_sm > 0, "conv1d_plan: arguments must be positive");
  conv_flat_tiles(tokens, column_groups, warps_per_sm, tile, ntiles);
  return 0;
}

#define DISPATCH_W(W_, SILU_, ...)                                                     \
  if (W_ == 4 && SILU_) { constexpr int W = 4; constexpr bool S = true; __VA_ARGS__; }  \
  else if (W_ == 4) { constexpr int W = 4; constexpr bool S = false; __VA_ARGS__; }    \
  else if (W_ == 3 && SILU_) { constexpr int W = 3; constexpr bool S = true; __VA_ARGS__; } \
  else if (W_ == 3) { constexpr int W = 3; constexpr bool S = false; __VA_ARGS__; }    \
  else if (W_ == 2 && SILU_) { constexpr int W = 2; constexpr bool S = true; __VA_ARGS__; } \
  else { constexpr int W = 2; constexpr bool S = false; __VA_ARGS__; }

extern "C" int mb_conv1d_fwd(const void* x, long long ldx, const float* w, const float* bias,
                             void* y, long long ldy, int B, int T, int C, int width, int silu,
                             cudaStream_t stream) {
  MB_REQUIRE(width >= 2 && width <= 4, "conv1d_fwd: width %d not in [2,4]", width);
  MB_REQUIRE(C % 8 == 0 && ldx % 8 == 0 && ldy % 8 == 0, "conv1d_fwd: C and row strides must be multiples of 8");
  if (B * T == 0) return 0;
  MB_REQUIRE((long long)B * T < (1ll << 31), "conv1d_fwd: B*T must fit in 31 bits");
  DISPATCH_W(width, silu, {
    static int per_sm = 0;
    if (per_sm == 0) {
      MB_CUDA(cudaFuncSetAttribute(conv1d_fwd_kernel<W, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
      MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, conv1d_fwd_kernel<W, S>, kConvThreads, kFwdSmem));
      if (per_sm < 1) per_sm = 1;
    }
    int tile, ntiles;
    const int ngroups = cdiv(C / 8, 32);
    conv_flat_tiles((long long)B * T, ngroups, per_sm * (kConvThreads / 32), &tile, &ntiles);
    MB_CUDA(launch_pdl(conv1d_fwd_kernel<W, S>, dim3(cdiv(ntiles * ngroups, kConvThreads / 32)), dim3(kConvThreads), kFwdSmem,
                       stream, (const bf16*)x, ldx, w, bias, (bf16*)y, ldy, T, C, B * T, tile, ntiles, ngroups));
  });
  return check_launch("conv1d_fwd");
}

extern "C" int mb_conv1d_bwd(const void* x, long long ldx, const float* w, const float* bias,
                             const void* dy, long long lddy, void* dx, long long lddx, float* dw,
                             float* dbias, int B, int T, int C, int width, int silu,
                             cudaStream_t stream) {
  MB_REQUIRE(width >= 2 && width <= 4, "conv1d_bwd: width %d not in [2,4]", width);
  MB_REQUIRE(C % 8 == 0 && ldx % 8 == 0 && lddy % 8 == 0 && lddx % 8 == 0,
             "conv1d_bwd: C and row strides must be multiples of 8");
  if (B * T == 0) return 0;
  MB_REQUIRE((long long)B * T < (1ll << 31), "conv1d_bwd: B*T must fit in 31 bits");
  DISPATCH_W(width, silu, {
    static int per_sm = 0;
    if (per_sm == 0) {
      MB_CUDA(cudaFuncSetAttribute(conv1d_bwd_kernel<W, S, kBwdNP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   conv_bwd_smem<kBwdNP>()));
      MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, conv1d_bwd_kernel<W, S, kBwdNP>, kConvThreads,
                                                            conv_bwd_smem<kBwdNP>()));
      if (per_sm < 1) per_sm = 1;
    }
    int tile, ntiles;
    const int ngroups = cdiv(C / (2 * kBwdNP), 32);
    conv_flat_tiles((long long)B * T, ngroups, per_sm * (kConvThreads / 32), &tile, &ntiles);
    MB_CUDA(launch_pdl(conv1d_bwd_kernel<W, S, kBwdNP>, dim3(cdiv(ntiles * ngroups, kConvThreads / 32)), dim3(kConvThreads),
                       conv_bwd_smem<kBwdNP>(), stream, (const bf16*)x, ldx, w, bias, (const bf16*)dy, lddy, (bf16*)dx, lddx,
                       dw, dbias, T, C, B * T, tile, ntiles, ngroups));
  });
  return check_launch("conv1d_bwd");
}

extern "C" int mb_conv1d_update(const void* x, long long ldx, float* state, const float* w,
                                const float* bias, void* y, long long ldy, int B, int C, int width,
                                int silu, cudaStream_t stream) {
  if (B * C == 0) return 0;
  const int n = B * C;
  if (silu) conv1d_update_kernel<true><<<cdiv(n, 256), 256, 0, stream>>>((const bf16*)x, ldx, state, w, bias, (bf16*)y, ldy, B, C, width);
  else conv1d_update_kernel<false><<<cdiv(n, 256), 256, 0, stream>>>((const bf16*)x, ldx, state, w, bias, (bf16*)y, ldy, B, C, width);
  return check_launch("conv1d_update");
}
