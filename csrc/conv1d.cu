// K1 / K2 / K3: depthwise causal conv1d (+bias +SiLU), channel-last, forward / backward / update.
//
// Replaces causal-conv1d's CUDA kernels.  B200-first
// differences: input and output are channel-last [B, T, C] *strided views* — the kernel reads
// the xBC columns of the in_proj GEMM output in place and the backward writes dx straight into
// the xBC columns of d(zxbcdt) — so no transpose / split / cat copies exist around it.
// Memory-bound; each thread owns 8 channels (one 16-byte vector) and walks a token tile with the
// sliding window in registers, so every global access is a coalesced 16-byte vector.
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kConvThreads = 128;
constexpr int kConvTile = 64;  // tokens per block

template <int W, bool SILU>
__global__ void __launch_bounds__(kConvThreads, 4)
conv1d_fwd_kernel(const bf16* __restrict__ x, long long ldx, const float* __restrict__ w,
                  const float* __restrict__ bias, bf16* __restrict__ y, long long ldy, int T, int C) {
  const int c0 = (blockIdx.x * kConvThreads + threadIdx.x) * 8;
  if (c0 >= C) return;
  const int b = blockIdx.z;
  const int t0 = blockIdx.y * kConvTile;
  const int t1 = min(t0 + kConvTile, T);
  float wt[8][W], bs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int k = 0; k < W; ++k) wt[j][k] = w[(c0 + j) * W + k];
    bs[j] = bias != nullptr ? bias[c0 + j] : 0.f;
  }
  const bf16* xb = x + ((size_t)b * T) * ldx + c0;
  bf16* yb = y + ((size_t)b * T) * ldy + c0;
  constexpr int U = 8;  // tokens per register batch: U loads in flight, static indexing, no window shifts
  bf16x8 halo[W - 1];   // x[tb - (W-1) + k], packed
#pragma unroll
  for (int k = 0; k < W - 1; ++k) {
    const int t = t0 - (W - 1) + k;
    if (t >= 0) halo[k] = ld8(xb + (size_t)t * ldx);
    else halo[k].v[0] = halo[k].v[1] = halo[k].v[2] = halo[k].v[3] = __floats2bfloat162_rn(0.f, 0.f);
  }
  for (int tb = t0; tb < t1; tb += U) {
    bf16x8 raw[U + W - 1];
#pragma unroll
    for (int k = 0; k < W - 1; ++k) raw[k] = halo[k];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (tb + u < t1) raw[W - 1 + u] = ld8(xb + (size_t)(tb + u) * ldx);
      else raw[W - 1 + u] = raw[W - 2 + u];
    }
#pragma unroll
    for (int k = 0; k < W - 1; ++k) halo[k] = raw[U + k];
    float acc[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[u][j] = bs[j];
#pragma unroll
    for (int r = 0; r < U + W - 1; ++r) {  // each input row feeds up to W outputs
      float xr[8];
      unpack8(raw[r], xr);
#pragma unroll
      for (int k = 0; k < W; ++k) {
        const int u = r - k;  // out[u] += w[k] * x[u + k]
        if (u >= 0 && u < U) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[u][j] += wt[j][k] * xr[j];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (tb + u < t1) {
        if (SILU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[u][j] = silu_f(acc[u][j]);
        }
        st8(yb + (size_t)(tb + u) * ldy, pack8(acc[u]));
      }
    }
  }
}

// Backward.  dw / dbias are accumulated with one fp32 atomic per (block, channel, tap).
template <int W, bool SILU>
__global__ void __launch_bounds__(kConvThreads)
conv1d_bwd_kernel(const bf16* __restrict__ x, long long ldx, const float* __restrict__ w,
                  const float* __restrict__ bias, const bf16* __restrict__ dy, long long lddy,
                  bf16* __restrict__ dx, long long lddx, float* __restrict__ dw,
                  float* __restrict__ dbias, int T, int C) {
  const int c0 = (blockIdx.x * kConvThreads + threadIdx.x) * 8;
  if (c0 >= C) return;
  const int b = blockIdx.z;
  const int t0 = blockIdx.y * kConvTile;
  const int t1 = min(t0 + kConvTile, T);
  float wt[8][W], bs[8], dwa[8][W], dba[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int k = 0; k < W; ++k) { wt[j][k] = w[(c0 + j) * W + k]; dwa[j][k] = 0.f; }
    bs[j] = bias != nullptr ? bias[c0 + j] : 0.f;
    dba[j] = 0.f;
  }
  const bf16* xb = x + ((size_t)b * T) * ldx + c0;
  const bf16* gb = dy + ((size_t)b * T) * lddy + c0;
  bf16* dxb = dx + ((size_t)b * T) * lddx + c0;
  float win[W][8];  // x window ending at t
  float gw[W][8];   // gw[k] = g[t - (W-1) + k]
#pragma unroll
  for (int k = 0; k < W; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) { win[k][j] = 0.f; gw[k][j] = 0.f; }
#pragma unroll
  for (int k = 0; k < W - 1; ++k) {
    const int t = t0 - (W - 1) + k;
    if (t >= 0) unpack8(ld8(xb + (size_t)t * ldx), win[k + 1]);
  }
  const int tend = t1 + (W - 1);  // g of the W-1 tokens after the tile feeds dx of the tile's tail
  constexpr int U = 8;
  for (int tb = t0; tb < tend; tb += U) {
    bf16x8 rx[U], rg[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (tb + u < tend && tb + u < T) {
        rx[u] = ld8(xb + (size_t)(tb + u) * ldx);
        rg[u] = ld8(gb + (size_t)(tb + u) * lddy);
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int t = tb + u;
      if (t >= tend) break;
#pragma unroll
      for (int k = 0; k < W - 1; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) { win[k][j] = win[k + 1][j]; gw[k][j] = gw[k + 1][j]; }
      if (t < T) {
        float g[8];
        unpack8(rx[u], win[W - 1]);
        unpack8(rg[u], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (SILU) {
            float a = bs[j];
#pragma unroll
            for (int k = 0; k < W; ++k) a += wt[j][k] * win[k][j];
            g[j] *= dsilu_f(a);
          }
          gw[W - 1][j] = g[j];
          if (t < t1) {  // this tile owns the weight-gradient contribution of token t
            dba[j] += g[j];
#pragma unroll
            for (int k = 0; k < W; ++k) dwa[j][k] += g[j] * win[k][j];
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) { gw[W - 1][j] = 0.f; win[W - 1][j] = 0.f; }
      }
      const int s_ = t - (W - 1);  // dx[s] = sum_k w[k] * g[s + (W-1) - k] = sum_k w[k] * gw[W-1-k]
      if (s_ >= t0 && s_ < t1) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float a = 0.f;
#pragma unroll
          for (int k = 0; k < W; ++k) a += wt[j][k] * gw[W - 1 - k][j];
          o[j] = a;
        }
        st8(dxb + (size_t)s_ * lddx, pack8(o));
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int k = 0; k < W; ++k) atomicAdd(dw + (c0 + j) * W + k, dwa[j][k]);
    if (dbias != nullptr) atomicAdd(dbias + c0 + j, dba[j]);
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
  dim3 grid(cdiv(C / 8, kConvThreads), cdiv(T, kConvTile), B);
  DISPATCH_W(width, silu, {
    conv1d_fwd_kernel<W, S><<<grid, kConvThreads, 0, stream>>>((const bf16*)x, ldx, w, bias, (bf16*)y, ldy, T, C);
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
  dim3 grid(cdiv(C / 8, kConvThreads), cdiv(T, kConvTile), B);
  DISPATCH_W(width, silu, {
    conv1d_bwd_kernel<W, S><<<grid, kConvThreads, 0, stream>>>((const bf16*)x, ldx, w, bias, (const bf16*)dy, lddy,
                                                               (bf16*)dx, lddx, dw, dbias, T, C);
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
