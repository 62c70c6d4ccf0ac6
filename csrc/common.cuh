// Shared device/host helpers for the sm_100a kernels (no external dependencies).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace mb {

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

// ---- error reporting through the C ABI ------------------------------------------------------
void set_error(const char* fmt, ...);          // defined in runtime.cu
int check_launch(const char* what);            // cudaGetLastError -> set_error; returns 0/err

#define MB_REQUIRE(cond, ...)                                                       \
  do {                                                                              \
    if (!(cond)) { mb::set_error(__VA_ARGS__); return 1; }                          \
  } while (0)

#define MB_CUDA(call)                                                               \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      mb::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return 1;                                                                     \
    }                                                                               \
  } while (0)

// ---- programmatic dependent launch ----------------------------------------------------------------
// Every hot kernel is launched with the programmatic-stream-serialization attribute: its blocks may become resident
// (shared-memory carve-up, barrier init, TMEM allocation, weight staging) while the tail of the previous kernel is
// still running, and `pdl_wait()` -- executed before the first read of anything an earlier kernel produced and
// before any global write -- blocks until that kernel has completed and flushed.  `pdl_trigger()` tells the
// hardware this block no longer minds the NEXT kernel's blocks being scheduled.  Both are no-ops for a kernel
// launched the ordinary way.  ~1600 launches per micro-step at 2-3 us of launch gap each is what this recovers.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Measured on the 280M step: eager launches -0.8 % device time; replayed from a CUDA graph (whose launch gaps are
// already ~1 us) +0.7 %, so launches that are being captured do not carry the attribute.
bool pdl_enabled();                       // runtime.cu: MAMBA_B200_PDL=0 turns the attribute off everywhere
inline bool pdl_for(cudaStream_t stream) {
  if (!pdl_enabled()) return false;
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  return cudaStreamIsCapturing(stream, &st) == cudaSuccess && st == cudaStreamCaptureStatusNone;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_for(stream) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---- small math -------------------------------------------------------------------------------
// sigmoid with one MUFU.EX2 + one MUFU.RCP.  (__frcp_rn / operator/ expand to an IEEE-rounded
// multi-instruction sequence; rcp.approx is a single MUFU op, 1 ulp-ish, plenty for an activation.)
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// exp through one FMUL + one MUFU.EX2.  (__expf without -use_fast_math wraps the same MUFU in a range
// fix-up — FSETP + two extra FMULs per call — to keep denormal results; here they flush to zero, which is
// what every caller wants: decays and sigmoids whose results are rounded to bf16 or multiplied away.)
__device__ __forceinline__ float fast_exp(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));
  return r;
}
__device__ __forceinline__ float sigmoid_f(float x) { return rcp_approx(1.f + fast_exp(-x)); }
__device__ __forceinline__ float silu_f(float x) { return x * sigmoid_f(x); }
// d/dx silu(x) = s + x*s*(1-s), s = sigmoid(x)
__device__ __forceinline__ float dsilu_f(float x) {
  const float s = sigmoid_f(x);
  return s * (1.f + x * (1.f - s));
}
// silu and its derivative from a single sigmoid evaluation
__device__ __forceinline__ void silu_both(float x, float& f, float& df) {
  const float s = sigmoid_f(x);
  f = x * s;
  df = s * (1.f + x * (1.f - s));
}
// ---- packed fp32 (sm_100: FFMA2 / FMUL2 / FADD2 do two fp32 lanes per issued instruction) -----------
// The fp32 pipe does the same number of flops either way; what the pairs buy is issue slots, which is
// what the elementwise kernels (conv, norms, SSD scaling) run out of first at their occupancy.
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float2 sigmoid2(float2 x) {
  const float2 t = mul2(x, f2(-1.4426950408889634f, -1.4426950408889634f));
  const float2 d = add2(f2(ex2_approx(t.x), ex2_approx(t.y)), f2(1.f, 1.f));
  return f2(rcp_approx(d.x), rcp_approx(d.y));
}
__device__ __forceinline__ float2 silu2(float2 x) { return mul2(x, sigmoid2(x)); }
// silu and its derivative s + x*s*(1-s) = s * (1 + x - x*s) from one sigmoid
__device__ __forceinline__ void silu_both2(float2 x, float2& f, float2& df) {
  const float2 s = sigmoid2(x);
  f = mul2(x, s);
  df = mul2(s, add2(f2(1.f, 1.f), fma2(f2(-1.f, -1.f), f, x)));  // 1 + (x - x*s)
}
__device__ __forceinline__ float softplus_f(float x) { return x > 20.f ? x : log1pf(__expf(x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- 16-byte vectors of 8 bf16 ----------------------------------------------------------------
struct __align__(16) bf16x8 { bf162 v[4]; };

__device__ __forceinline__ void unpack8(const bf16x8& p, float* f) {
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(p.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ bf16x8 pack8(const float* f) {
  bf16x8 p;
#pragma unroll
  for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return p;
}
// Forced 128-bit accesses: a plain struct copy of bf16x8 is scalarised by nvcc into four 32-bit
// LDG/STG (seen in SASS), which quarters the bytes in flight per instruction.
__device__ __forceinline__ bf16x8 ld8(const bf16* p) {
  const uint4 t = *reinterpret_cast<const uint4*>(p);
  bf16x8 r;
  memcpy(&r, &t, 16);
  return r;
}
__device__ __forceinline__ void st8(bf16* p, const bf16x8& v) {
  uint4 t;
  memcpy(&t, &v, 16);
  *reinterpret_cast<uint4*>(p) = t;
}
__device__ __forceinline__ void ld8f(const float* p, float* f) {
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void st8f(float* p, const float* f) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
}

inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
int num_sms();  // cached cudaDevAttrMultiProcessorCount of the current device (runtime.cu)

}  // namespace mb
