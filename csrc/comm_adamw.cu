// K19 + K20 + K21: gradient synchronisation fused with clipping, AdamW and the bf16 parameter
// refresh, over NVLink peer memory — no NCCL call on this path.
//
// Replaces, from the reference's step (/root/reference/train.py:86,210-227, model.py:141-148):
// DistributedDataParallel's bucketed ncclAllReduce + copy-out, clip_grad_norm_ (foreach norm +
// scale) and torch.optim.AdamW(fused=True), i.e. ~600 launches and ~9 GB of HBM traffic per step,
// by two kernels:
//
//  A  mb_comm_reduce_scatter   one launch per gradient bucket, as soon as the bucket's last wgrad
//     of the last micro-step is enqueued (side stream; overlaps the rest of backward).
//     Every rank owns 1/world of the bucket.  It publishes "my bucket is final" to all peers
//     (release store into their control block), waits for theirs (acquire loads), then pulls the
//     owned shard from every peer's gradient arena with 16-byte P2P loads over NVLink, sums in a
//     fixed rank order, scales by 1/world, writes the averaged shard back in place and
//     accumulates its sum of squares.
//
//  B  mb_comm_adamw_allgather  once per step.  Exchanges the per-rank squared-norm partials
//     (packed {seq, value} 64-bit words, summed in rank order so every rank gets the same bits),
//     computes the clip coefficient min(1, max_norm / (norm + 1e-6)) exactly as clip_grad_norm_,
//     runs decoupled-weight-decay Adam on the owned shard only (optimizer state is sharded,
//     ZeRO-1), writes the fp32 master in place and stores the new bf16 parameters into EVERY
//     rank's bf16 arena (all-gather by P2P stores).  A final flag exchange guarantees that when
//     the kernel retires all peers' shards have landed locally, so the next forward can start.
//
// All cross-GPU waits are bounded: a flag that never arrives prints rank / phase / bucket and
// traps instead of hanging.  world == 1 runs the same code with itself
// as the only peer (fused clip + AdamW + bf16 cast).
#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kMaxWorld = 16;
constexpr int kMaxBuckets = 512;
constexpr int kCommThreads = 256;

// control block (one per rank, peer-writable), in 64-bit words
//   [0, kMaxBuckets*kMaxWorld)            bucket-final flags   flagA[bucket][src]
//   [+0, +kMaxWorld)                      norm partials        {seq:32 | float bits:32}
//   [+kMaxWorld, +2*kMaxWorld)            phase-B done flags
constexpr int kCtrlFlagA = 0;
constexpr int kCtrlNorm = kMaxBuckets * kMaxWorld;
constexpr int kCtrlDone = kCtrlNorm + kMaxWorld;
constexpr int kCtrlWords = kCtrlDone + kMaxWorld;

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;  // relaxed system-scope load: never served from a stale non-coherent line
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// wait until *p (local control word) satisfies pred; bounded.
template <class Pred>
__device__ __forceinline__ unsigned long long spin_wait(const unsigned long long* p, Pred pred, int rank, int phase,
                                                        int bucket, int src) {
  unsigned long long v = ld_acquire_sys(p);
  unsigned long long spins = 0;
  while (!pred(v)) {
    __nanosleep(64);
    v = ld_acquire_sys(p);
    if (++spins > (1ull << 26)) {  // ~ several seconds
      printf("mamba_b200 comm timeout: rank %d phase %d bucket %d waiting for rank %d (word=%llx)\n", rank, phase,
             bucket, src, v);
      __trap();
    }
  }
  return v;
}

struct PeerTable {
  float* grad[kMaxWorld];
  bf16* pbf16[kMaxWorld];
  unsigned long long* ctrl[kMaxWorld];
  float* master[kMaxWorld];  // fp32 master arenas (phase B only)
};

// ------------------------------------------------------------------------------------------ A
__global__ void __launch_bounds__(kCommThreads)
comm_reduce_scatter_kernel(PeerTable peers, int world, int rank, int bucket, long long start, long long shard_len,
                           unsigned long long seq, float* __restrict__ norm_acc) {
  __shared__ float red[kCommThreads / 32];
  const int tid = threadIdx.x;
  if (world > 1) {
    if (blockIdx.x == 0 && tid < world) {
      __threadfence_system();  // this rank's gradients (written by earlier kernels) before the flag
      st_release_sys(peers.ctrl[tid] + kCtrlFlagA + bucket * kMaxWorld + rank, seq);
    }
    if (tid < world)
      spin_wait(peers.ctrl[rank] + kCtrlFlagA + bucket * kMaxWorld + tid,
                [seq](unsigned long long v) { return v >= seq; }, rank, 0, bucket, tid);
    __syncthreads();
  }
  const long long off0 = start + (long long)rank * shard_len;
  const float inv = 1.f / (float)world;
  float ss = 0.f;
  float* mine = peers.grad[rank] + off0;
  const long long nvec = shard_len >> 2;
  for (long long i = (long long)blockIdx.x * kCommThreads + tid; i < nvec; i += (long long)gridDim.x * kCommThreads) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int p = 0; p < world; ++p) {  // fixed order: every rank reduces a shard exactly once -> reproducible
      const float4 v = (p == rank) ? *reinterpret_cast<const float4*>(mine + 4 * i)
                                   : ld_peer_f4(peers.grad[p] + off0 + 4 * i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
    ss += acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
    *reinterpret_cast<float4*>(mine + 4 * i) = acc;
  }
  ss = warp_sum(ss);
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int w = 0; w < kCommThreads / 32; ++w) t += red[w];
    atomicAdd(norm_acc, t);
  }
}

// ------------------------------------------------------------------------------------------ B
struct AdamArgs {
  float lr[4], wd[4];
  float beta1, beta2, eps, bc1, bc2, max_norm;
};

__global__ void __launch_bounds__(kCommThreads)
comm_adamw_allgather_kernel(PeerTable peers, int world, int rank, unsigned long long seq,
                            const long long* __restrict__ items, int n_items, float* __restrict__ m_state, float* __restrict__ v_state, AdamArgs a,
                            float* __restrict__ norm_acc, float* __restrict__ norm_out,
                            unsigned int* __restrict__ block_counter) {
  __shared__ float s_clip;
  __shared__ bool s_skip;
  const int tid = threadIdx.x;
  // ---- global gradient norm: exchange per-rank partials ----
  if (blockIdx.x == 0 && tid < world) {
    const float mine = *reinterpret_cast<volatile float*>(norm_acc);
    const unsigned long long word = (seq << 32) | (unsigned long long)__float_as_uint(mine);
    __threadfence_system();
    st_release_sys(peers.ctrl[tid] + kCtrlNorm + rank, word);
  }
  if (tid == 0) {
    float total = 0.f;
    for (int p = 0; p < world; ++p) {
      const unsigned long long w = spin_wait(peers.ctrl[rank] + kCtrlNorm + p,
                                             [seq](unsigned long long v) { return (v >> 32) == (seq & 0xffffffffull); },
                                             rank, 1, -1, p);
      total += __uint_as_float((unsigned int)(w & 0xffffffffull));
    }
    const float norm = sqrtf(total);
    float clip = 1.f;
    if (a.max_norm > 0.f) clip = fminf(1.f, a.max_norm / (norm + 1e-6f));
    // NaN/Inf guard: a non-finite global norm means some gradient is NaN/Inf; every rank sees the
    // same total, so all of them skip this update (parameters, m and v stay as they are; 0 * NaN would not).
    s_skip = !isfinite(norm);
    s_clip = clip;
    if (blockIdx.x == 0) *norm_out = norm;
  }
  __syncthreads();
  const float clip = s_clip;
  const float inv_bc1 = 1.f / a.bc1, inv_sqrt_bc2 = rsqrtf(a.bc2);
  float* grad = peers.grad[rank];
  float* master = peers.master[rank];

  for (int it = blockIdx.x; it < n_items && !s_skip; it += gridDim.x) {
    const long long arena_off = items[4 * it + 0];
    const long long local_off = items[4 * it + 1];
    const int count = (int)items[4 * it + 2];
    const int grp = (int)(items[4 * it + 3] & 0xff);
    // parameters that kernels consume in fp32 (norm weights, conv, A_log, dt_bias, D, embedding):
    // the owner also stores the fp32 master into every peer's master arena
    const bool gather_fp32 = ((items[4 * it + 3] >> 8) & 1) != 0 && world > 1;
    const float lr = a.lr[grp], wd = a.wd[grp];
    const float decay = 1.f - lr * wd;
    for (int i = tid * 4; i < count; i += kCommThreads * 4) {
      const float4 g4 = *reinterpret_cast<const float4*>(grad + arena_off + i);
      float4 p4 = *reinterpret_cast<const float4*>(master + arena_off + i);
      float4 m4 = *reinterpret_cast<const float4*>(m_state + local_off + i);
      float4 v4 = *reinterpret_cast<const float4*>(v_state + local_off + i);
      float g[4] = {g4.x * clip, g4.y * clip, g4.z * clip, g4.w * clip};
      float p[4] = {p4.x, p4.y, p4.z, p4.w}, m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        p[j] *= decay;
        m[j] = a.beta1 * m[j] + (1.f - a.beta1) * g[j];
        v[j] = a.beta2 * v[j] + (1.f - a.beta2) * g[j] * g[j];
        const float denom = sqrtf(v[j]) * inv_sqrt_bc2 + a.eps;
        p[j] -= lr * inv_bc1 * m[j] / denom;
      }
      *reinterpret_cast<float4*>(master + arena_off + i) = make_float4(p[0], p[1], p[2], p[3]);
      *reinterpret_cast<float4*>(m_state + local_off + i) = make_float4(m[0], m[1], m[2], m[3]);
      *reinterpret_cast<float4*>(v_state + local_off + i) = make_float4(v[0], v[1], v[2], v[3]);
      const bf162 lo = __floats2bfloat162_rn(p[0], p[1]), hi = __floats2bfloat162_rn(p[2], p[3]);
      uint2 packed;
      packed.x = *reinterpret_cast<const unsigned int*>(&lo);
      packed.y = *reinterpret_cast<const unsigned int*>(&hi);
      for (int q = 0; q < world; ++q)  // all-gather: the owner stores its shard into every rank's arena
        *reinterpret_cast<uint2*>(peers.pbf16[q] + arena_off + i) = packed;
      if (gather_fp32) {
        for (int q = 0; q < world; ++q)
          if (q != rank) *reinterpret_cast<float4*>(peers.master[q] + arena_off + i) = make_float4(p[0], p[1], p[2], p[3]);
      }
    }
  }

  // ---- completion: my shard has been stored everywhere; wait until everyone else's has landed here ----
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (tid == 0) last = (atomicAdd(block_counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    if (tid < world) {
      __threadfence_system();
      st_release_sys(peers.ctrl[tid] + kCtrlDone + rank, seq);
    }
    if (tid < world)
      spin_wait(peers.ctrl[rank] + kCtrlDone + tid, [seq](unsigned long long v) { return v >= seq; }, rank, 2, -1, tid);
    if (tid == 0) {
      *block_counter = 0;
      *norm_acc = 0.f;
    }
  }
}

}  // namespace mb
using namespace mb;

static int fill_peers(PeerTable& t, const long long* grad_ptrs, const long long* bf16_ptrs, const long long* ctrl_ptrs,
                      const long long* master_ptrs, int world) {
  MB_REQUIRE(world >= 1 && world <= kMaxWorld, "comm: world size %d not in [1, %d]", world, kMaxWorld);
  for (int i = 0; i < kMaxWorld; ++i) {
    const int j = i < world ? i : 0;
    t.grad[i] = reinterpret_cast<float*>(grad_ptrs[j]);
    t.pbf16[i] = reinterpret_cast<bf16*>(bf16_ptrs[j]);
    t.ctrl[i] = reinterpret_cast<unsigned long long*>(ctrl_ptrs[j]);
    t.master[i] = master_ptrs != nullptr ? reinterpret_cast<float*>(master_ptrs[j]) : nullptr;
  }
  return 0;
}

extern "C" long long mb_comm_ctrl_words(void) { return kCtrlWords; }

// grad_ptrs / bf16_ptrs / ctrl_ptrs: HOST arrays of `world` device addresses (peer-mapped arenas).
extern "C" int mb_comm_reduce_scatter(const long long* grad_ptrs, const long long* bf16_ptrs,
                                      const long long* ctrl_ptrs, int world, int rank, int bucket,
                                      long long start, long long shard_len, long long seq, float* norm_acc,
                                      int num_blocks, cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, grad_ptrs, bf16_ptrs, ctrl_ptrs, nullptr, world)) return 1;
  MB_REQUIRE(bucket >= 0 && bucket < kMaxBuckets, "comm: bucket index %d >= %d", bucket, kMaxBuckets);
  MB_REQUIRE(shard_len % 4 == 0 && start % 4 == 0, "comm: shard must be 16-byte aligned");
  if (num_blocks <= 0) num_blocks = 32;
  comm_reduce_scatter_kernel<<<num_blocks, kCommThreads, 0, stream>>>(t, world, rank, bucket, start, shard_len,
                                                                      (unsigned long long)seq, norm_acc);
  return check_launch("comm_reduce_scatter");
}

// items: device int64 [n_items][4] = {arena offset, optimizer-state offset, count (multiple of 4),
//        group | (fp32_gather << 8)};  master_ptrs: HOST array of the peer-mapped fp32 master arenas
// hyper: host float[14] = lr[4], wd[4], beta1, beta2, eps, bias_correction1, bias_correction2, max_norm
extern "C" int mb_comm_adamw_allgather(const long long* grad_ptrs, const long long* bf16_ptrs,
                                       const long long* ctrl_ptrs, const long long* master_ptrs, int world,
                                       int rank, long long seq, const long long* items, int n_items, float* m_state,
                                       float* v_state, const float* hyper, float* norm_acc, float* norm_out,
                                       unsigned int* block_counter, int num_blocks, cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, grad_ptrs, bf16_ptrs, ctrl_ptrs, master_ptrs, world)) return 1;
  AdamArgs a;
  for (int i = 0; i < 4; ++i) { a.lr[i] = hyper[i]; a.wd[i] = hyper[4 + i]; }
  a.beta1 = hyper[8]; a.beta2 = hyper[9]; a.eps = hyper[10]; a.bc1 = hyper[11]; a.bc2 = hyper[12]; a.max_norm = hyper[13];
  if (num_blocks <= 0) num_blocks = 2 * num_sms();
  comm_adamw_allgather_kernel<<<num_blocks, kCommThreads, 0, stream>>>(t, world, rank, (unsigned long long)seq, items,
                                                                       n_items, m_state, v_state, a, norm_acc,
                                                                       norm_out, block_counter);
  return check_launch("comm_adamw_allgather");
}
