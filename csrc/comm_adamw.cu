// K19 + K20 + K21: gradient synchronisation fused with clipping, AdamW, the loss mean and the bf16
// parameter refresh, over NVLink peer / multicast memory — no NCCL call on this path.
//
// Replaces, from the reference's step (/root/reference/train.py:86,210-227, model.py:141-148):
// DistributedDataParallel's bucketed ncclAllReduce + copy-out, the scalar loss all-reduce,
// clip_grad_norm_ (foreach norm + scale) and torch.optim.AdamW(fused=True), i.e. ~600 launches and
// ~9 GB of HBM traffic per step, by two kernels:
//
//  A  mb_comm_reduce_scatter   one launch per gradient bucket, as soon as the bucket's last wgrad
//     of the last micro-step is enqueued (side stream; overlaps the rest of backward).
//     Every rank owns 1/world of the bucket.  It publishes "my bucket is final" to all peers
//     (release store into their control block), waits for theirs (acquire loads), then reduces the
//     owned shard over all ranks, scales by 1/world, writes the averaged shard back in place and
//     accumulates its sum of squares.  Three algorithms, chosen per bucket by the host:
//       NVLS     multimem.ld_reduce.add.v4.f32 on the multicast mapping of the gradient arena: the
//                NVSwitch sums the world copies, each rank receives its shard once (1/world of the
//                bytes of the pull);
//       P2P      16-byte loads from every peer's arena, summed in rank order (no multicast object,
//                or world == 2 where the switch saves nothing);
//       ONESHOT  the P2P body on a single CTA for buckets below 256 KiB: one flag round, no grid to
//                schedule behind the persistent GEMMs of backward.
//
//  B  mb_comm_adamw_allgather  once per step.  Exchanges the per-rank squared-norm partials and
//     the per-rank loss (packed {seq, value} 64-bit words, summed in rank order so every rank gets
//     the same bits), computes the clip coefficient min(1, max_norm / (norm + 1e-6)) exactly as
//     clip_grad_norm_, runs decoupled-weight-decay Adam on the owned shard only (optimizer state
//     is sharded, ZeRO-1), writes the fp32 master in place and stores the new bf16 parameters into
//     EVERY rank's bf16 arena: one multimem.st per 16 bytes through the multicast mapping, or one
//     16-byte P2P store per peer.  A final flag exchange guarantees that when the kernel retires
//     all peers' shards have landed locally, so the next forward can start.
//
// The step sequence number lives in device memory (control word kCtrlSeq, advanced by kernel B),
// so neither kernel has a per-step launch argument and both can sit in a captured CUDA graph.
// All cross-GPU waits are bounded by a wall-clock budget (mb_comm_set_timeout_ms): a flag that
// never arrives prints rank / phase / bucket and traps instead of hanging.
// world == 1 runs the same code with itself as the only peer (fused clip + AdamW + bf16 cast).
#include <stdlib.h>

#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {

constexpr int kMaxWorld = 16;
constexpr int kMaxBuckets = 512;
constexpr int kCommThreads = 256;

// control block (one per rank, peer-writable), in 64-bit words
//   flagA[bucket][src]   bucket-final flags (step sequence numbers)
//   norm[src], loss[src] per-rank partials     {seq:32 | float bits:32}
//   done[src]            phase-B done flags
//   bar[src]             arrivals of the stand-alone barrier kernel
//   scal[src]            operands of the stand-alone scalar-mean kernel {seq:32 | float bits:32}
//   seq[0..2]            LOCAL counters: completed steps, barriers, scalar means
constexpr int kCtrlFlagA = 0;
constexpr int kCtrlNorm = kMaxBuckets * kMaxWorld;
constexpr int kCtrlLoss = kCtrlNorm + kMaxWorld;
constexpr int kCtrlDone = kCtrlLoss + kMaxWorld;
constexpr int kCtrlBar = kCtrlDone + kMaxWorld;
constexpr int kCtrlScal = kCtrlBar + kMaxWorld;
constexpr int kCtrlSeq = kCtrlScal + kMaxWorld;
constexpr int kCtrlWords = kCtrlSeq + 8;

enum { ALGO_P2P = 0, ALGO_NVLS = 1, ALGO_ONESHOT = 2 };

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
// NVLS: one load on the multicast address returns the sum of every rank's copy (reduced in the switch)
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
// NVLS: one store on the multicast address lands in every rank's copy
__device__ __forceinline__ void multimem_st_f4(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// wait until *p (local control word) satisfies pred; bounded by timeout_ns of wall clock.
template <class Pred>
__device__ __forceinline__ unsigned long long spin_wait(const unsigned long long* p, Pred pred, unsigned long long timeout_ns,
                                                        int rank, int phase, int bucket, int src) {
  unsigned long long v = ld_acquire_sys(p);
  if (pred(v)) return v;
  const unsigned long long t0 = globaltimer_ns();
  unsigned int spins = 0;
  while (!pred(v)) {
    __nanosleep(64);
    v = ld_acquire_sys(p);
    if ((++spins & 1023u) == 0 && globaltimer_ns() - t0 > timeout_ns) {
      printf("mamba_b200 comm timeout: rank %d phase %d bucket %d waited %llu ms for rank %d (word=%llx)\n", rank, phase,
             bucket, timeout_ns / 1000000ull, src, v);
      __trap();
    }
  }
  return v;
}

struct PeerTable {
  float* grad[kMaxWorld];
  bf16* pbf16[kMaxWorld];
  unsigned long long* ctrl[kMaxWorld];
  float* master[kMaxWorld];  // fp32 master arenas (phase B / gather only)
  float* mc_grad;            // multicast mappings of the three arenas (nullptr: no NVLS)
  bf16* mc_bf16;
  float* mc_master;
};

// ------------------------------------------------------------------------------------------ A
// W: world size known at compile time (all peer loads of a vector are issued before the adds), 0: generic.
template <int W>
__device__ __forceinline__ float4 pull_sum(const PeerTable& peers, int world, int rank, const float* mine, long long off, long long i) {
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (W > 0) {
    float4 v[W > 0 ? W : 1];
#pragma unroll
    for (int p = 0; p < W; ++p)
      v[p] = (p == rank) ? *reinterpret_cast<const float4*>(mine + 4 * i) : ld_peer_f4(peers.grad[p] + off + 4 * i);
#pragma unroll
    for (int p = 0; p < W; ++p) { acc.x += v[p].x; acc.y += v[p].y; acc.z += v[p].z; acc.w += v[p].w; }
  } else {
    for (int p = 0; p < world; ++p) {  // fixed order: every rank reduces a shard exactly once -> reproducible
      const float4 v = (p == rank) ? *reinterpret_cast<const float4*>(mine + 4 * i) : ld_peer_f4(peers.grad[p] + off + 4 * i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  return acc;
}

template <int W, bool NVLS>
__global__ void __launch_bounds__(kCommThreads)
comm_reduce_scatter_kernel(PeerTable peers, int world, int rank, int bucket, long long start, long long shard_len,
                           float* __restrict__ norm_acc, unsigned long long timeout_ns) {
  __shared__ float red[kCommThreads / 32];
  const int tid = threadIdx.x;
  const unsigned long long seq = peers.ctrl[rank][kCtrlSeq] + 1;  // this step (advanced by kernel B)
  if (world > 1) {
    if (blockIdx.x == 0 && tid < world) {
      __threadfence_system();  // this rank's gradients (written by earlier kernels) before the flag
      st_release_sys(peers.ctrl[tid] + kCtrlFlagA + bucket * kMaxWorld + rank, seq);
    }
    if (tid < world)
      spin_wait(peers.ctrl[rank] + kCtrlFlagA + bucket * kMaxWorld + tid,
                [seq](unsigned long long v) { return v >= seq; }, timeout_ns, rank, 0, bucket, tid);
    __syncthreads();
  }
  const long long off0 = start + (long long)rank * shard_len;
  const float inv = 1.f / (float)world;
  float ss = 0.f;
  float* mine = peers.grad[rank] + off0;
  const long long nvec = shard_len >> 2;
  const long long stride = (long long)gridDim.x * kCommThreads;
  constexpr int U = 4;  // independent vectors in flight per thread
  for (long long i0 = (long long)blockIdx.x * kCommThreads + tid; i0 < nvec; i0 += U * stride) {
    float4 acc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < nvec) acc[u] = NVLS ? multimem_ld_reduce_f4(peers.mc_grad + off0 + 4 * i) : pull_sum<W>(peers, world, rank, mine, off0, i);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < nvec) {
        float4 a = acc[u];
        a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
        ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
        *reinterpret_cast<float4*>(mine + 4 * i) = a;
      }
    }
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
comm_adamw_allgather_kernel(PeerTable peers, int world, int rank, const long long* __restrict__ items, int n_items,
                            float* __restrict__ m_state, float* __restrict__ v_state, AdamArgs a,
                            float* __restrict__ norm_acc, float* __restrict__ norm_out, const float* __restrict__ loss_in,
                            float* __restrict__ loss_out, unsigned int* __restrict__ block_counter,
                            unsigned long long timeout_ns) {
  __shared__ float s_clip;
  __shared__ bool s_skip;
  const int tid = threadIdx.x;
  unsigned long long* ctrl = peers.ctrl[rank];
  const unsigned long long seq = ctrl[kCtrlSeq] + 1;
  // ---- global gradient norm and loss mean: exchange per-rank partials ----
  if (blockIdx.x == 0 && tid < world) {
    const float mine = *reinterpret_cast<volatile float*>(norm_acc);
    const float myloss = loss_in != nullptr ? *reinterpret_cast<const volatile float*>(loss_in) : 0.f;
    __threadfence_system();
    st_release_sys(peers.ctrl[tid] + kCtrlLoss + rank, (seq << 32) | (unsigned long long)__float_as_uint(myloss));
    st_release_sys(peers.ctrl[tid] + kCtrlNorm + rank, (seq << 32) | (unsigned long long)__float_as_uint(mine));
  }
  if (tid == 0) {
    float total = 0.f;
    const auto same_seq = [seq](unsigned long long v) { return (v >> 32) == (seq & 0xffffffffull); };
    for (int p = 0; p < world; ++p) {
      const unsigned long long w = spin_wait(ctrl + kCtrlNorm + p, same_seq, timeout_ns, rank, 1, -1, p);
      total += __uint_as_float((unsigned int)(w & 0xffffffffull));
    }
    if (blockIdx.x == 0 && loss_out != nullptr) {  // K21: the loss mean rides in the same exchange
      float lsum = 0.f;
      for (int p = 0; p < world; ++p) {
        const unsigned long long w = spin_wait(ctrl + kCtrlLoss + p, same_seq, timeout_ns, rank, 1, -2, p);
        lsum += __uint_as_float((unsigned int)(w & 0xffffffffull));
      }
      *loss_out = lsum / (float)world;
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
  const bool nvls = peers.mc_bf16 != nullptr && world > 1;

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
    for (int i = tid * 8; i < count; i += kCommThreads * 8) {
      float g[8], p[8], m[8], v[8];
      ld8f(grad + arena_off + i, g);
      ld8f(master + arena_off + i, p);
      ld8f(m_state + local_off + i, m);
      ld8f(v_state + local_off + i, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gj = g[j] * clip;
        p[j] *= decay;
        m[j] = a.beta1 * m[j] + (1.f - a.beta1) * gj;
        v[j] = a.beta2 * v[j] + (1.f - a.beta2) * gj * gj;
        const float denom = sqrtf(v[j]) * inv_sqrt_bc2 + a.eps;
        p[j] -= lr * inv_bc1 * m[j] / denom;
      }
      st8f(master + arena_off + i, p);
      st8f(m_state + local_off + i, m);
      st8f(v_state + local_off + i, v);
      const bf16x8 packed = pack8(p);
      float4 pk;
      memcpy(&pk, &packed, 16);
      if (nvls) {  // all-gather: one multicast store reaches every rank's arena (this one included)
        multimem_st_f4(peers.mc_bf16 + arena_off + i, pk);
        if (gather_fp32) {
          multimem_st_f4(peers.mc_master + arena_off + i, make_float4(p[0], p[1], p[2], p[3]));
          multimem_st_f4(peers.mc_master + arena_off + i + 4, make_float4(p[4], p[5], p[6], p[7]));
        }
      } else {     // all-gather: the owner stores its shard into every rank's arena
        for (int q = 0; q < world; ++q) *reinterpret_cast<float4*>(peers.pbf16[q] + arena_off + i) = pk;
        if (gather_fp32) {
          for (int q = 0; q < world; ++q)
            if (q != rank) st8f(peers.master[q] + arena_off + i, p);
        }
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
      spin_wait(ctrl + kCtrlDone + tid, [seq](unsigned long long v) { return v >= seq; }, timeout_ns, rank, 2, -1, tid);
    __syncthreads();
    if (tid == 0) {
      *block_counter = 0;
      *norm_acc = 0.f;
      ctrl[kCtrlSeq] = seq;  // the step is complete on every rank: next step's kernels use seq + 1
    }
  }
}

// ------------------------------------------------------------------- small stand-alone collectives
// Barrier over all ranks (one warp): used around checkpoint gathers, never on the step path.
__global__ void comm_barrier_kernel(PeerTable peers, int world, int rank, unsigned long long timeout_ns) {
  const int tid = threadIdx.x;
  unsigned long long* ctrl = peers.ctrl[rank];
  const unsigned long long bseq = ctrl[kCtrlSeq + 1] + 1;
  __syncwarp();
  if (tid < world) {
    __threadfence_system();
    st_release_sys(peers.ctrl[tid] + kCtrlBar + rank, bseq);
    spin_wait(ctrl + kCtrlBar + tid, [bseq](unsigned long long v) { return v >= bseq; }, timeout_ns, rank, 3, -1, tid);
  }
  __syncwarp();
  if (tid == 0) ctrl[kCtrlSeq + 1] = bseq;
}

// Mean over ranks of n <= kMaxWorld... of ONE fp32 scalar (the validation loss, /root/reference/train.py:147).
__global__ void comm_scalar_mean_kernel(PeerTable peers, int world, int rank, const float* __restrict__ in,
                                        float* __restrict__ out, unsigned long long timeout_ns) {
  const int tid = threadIdx.x;
  unsigned long long* ctrl = peers.ctrl[rank];
  const unsigned long long sseq = ctrl[kCtrlSeq + 2] + 1;
  __syncwarp();
  if (tid < world) {
    const float mine = *in;
    __threadfence_system();
    st_release_sys(peers.ctrl[tid] + kCtrlScal + rank, (sseq << 32) | (unsigned long long)__float_as_uint(mine));
  }
  __syncwarp();
  if (tid == 0) {
    float s = 0.f;
    for (int p = 0; p < world; ++p) {  // rank order: every rank computes the same bits
      const unsigned long long w = spin_wait(ctrl + kCtrlScal + p,
                                             [sseq](unsigned long long v) { return (v >> 32) == (sseq & 0xffffffffull); },
                                             timeout_ns, rank, 4, -1, p);
      s += __uint_as_float((unsigned int)(w & 0xffffffffull));
    }
    *out = s / (float)world;
    ctrl[kCtrlSeq + 2] = sseq;
  }
}

// Checkpoint path: copy every peer's owned shard [start + q*shard_len, +shard_len) of one bucket from ITS master
// arena into this rank's (the owner holds the current fp32 values after a fused step).  Callers bracket the
// launches of a whole gather with mb_comm_barrier on both sides.
__global__ void __launch_bounds__(kCommThreads)
comm_gather_master_kernel(PeerTable peers, int world, int rank, long long start, long long shard_len) {
  const long long nvec = shard_len >> 2;
  for (int q = 0; q < world; ++q) {
    if (q == rank) continue;
    const float* src = peers.master[q] + start + (long long)q * shard_len;
    float* dst = peers.master[rank] + start + (long long)q * shard_len;
    for (long long i = (long long)blockIdx.x * kCommThreads + threadIdx.x; i < nvec; i += (long long)gridDim.x * kCommThreads)
      *reinterpret_cast<float4*>(dst + 4 * i) = ld_peer_f4(src + 4 * i);
  }
}

__global__ void timestamp_kernel(unsigned long long* out) { *out = globaltimer_ns(); }

}  // namespace mb
using namespace mb;

static unsigned long long g_timeout_ns = 0;
static unsigned long long comm_timeout_ns() {
  if (g_timeout_ns == 0) {
    const char* e = getenv("MAMBA_B200_COMM_TIMEOUT_MS");
    long long ms = e != nullptr ? atoll(e) : 0;
    if (ms <= 0) ms = 60000;  // a minute: long enough for a rank that is writing a checkpoint or compiling
    g_timeout_ns = (unsigned long long)ms * 1000000ull;
  }
  return g_timeout_ns;
}

static int fill_peers(PeerTable& t, const long long* grad_ptrs, const long long* bf16_ptrs, const long long* ctrl_ptrs,
                      const long long* master_ptrs, int world) {
  MB_REQUIRE(world >= 1 && world <= kMaxWorld, "comm: world size %d not in [1, %d]", world, kMaxWorld);
  for (int i = 0; i < kMaxWorld; ++i) {
    const int j = i < world ? i : 0;
    t.grad[i] = grad_ptrs != nullptr ? reinterpret_cast<float*>(grad_ptrs[j]) : nullptr;
    t.pbf16[i] = bf16_ptrs != nullptr ? reinterpret_cast<bf16*>(bf16_ptrs[j]) : nullptr;
    t.ctrl[i] = reinterpret_cast<unsigned long long*>(ctrl_ptrs[j]);
    t.master[i] = master_ptrs != nullptr ? reinterpret_cast<float*>(master_ptrs[j]) : nullptr;
  }
  t.mc_grad = nullptr; t.mc_bf16 = nullptr; t.mc_master = nullptr;
  return 0;
}

extern "C" long long mb_comm_ctrl_words(void) { return kCtrlWords; }

extern "C" int mb_comm_set_timeout_ms(long long ms) {
  MB_REQUIRE(ms > 0, "comm: timeout must be positive");
  g_timeout_ns = (unsigned long long)ms * 1000000ull;
  return 0;
}

template <bool NVLS>
static void launch_rs(int world, int blocks, cudaStream_t stream, const PeerTable& t, int rank, int bucket, long long start,
                      long long shard_len, float* norm_acc) {
  const unsigned long long to = comm_timeout_ns();
#define RS(W_) comm_reduce_scatter_kernel<W_, NVLS><<<blocks, kCommThreads, 0, stream>>>(t, world, rank, bucket, start, shard_len, norm_acc, to)
  if (NVLS) RS(0);
  else if (world == 2) RS(2);
  else if (world == 4) RS(4);
  else if (world == 8) RS(8);
  else RS(0);
#undef RS
}

// grad_ptrs / ctrl_ptrs: HOST arrays of `world` device addresses (peer-mapped arenas); mc_grad: multicast
// address of the gradient arena (0: none).  algo: 0 P2P pull, 1 NVLS (multimem.ld_reduce), 2 one-shot (single CTA).
extern "C" int mb_comm_reduce_scatter(const long long* grad_ptrs, const long long* ctrl_ptrs, long long mc_grad,
                                      int world, int rank, int bucket, long long start, long long shard_len,
                                      float* norm_acc, int algo, int num_blocks, cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, grad_ptrs, nullptr, ctrl_ptrs, nullptr, world)) return 1;
  MB_REQUIRE(bucket >= 0 && bucket < kMaxBuckets, "comm: bucket index %d >= %d", bucket, kMaxBuckets);
  MB_REQUIRE(shard_len % 4 == 0 && start % 4 == 0, "comm: shard must be 16-byte aligned");
  MB_REQUIRE(algo >= ALGO_P2P && algo <= ALGO_ONESHOT, "comm: unknown algorithm %d", algo);
  MB_REQUIRE(algo != ALGO_NVLS || (mc_grad != 0 && world > 1), "comm: NVLS reduce-scatter needs a multicast mapping");
  t.mc_grad = reinterpret_cast<float*>(mc_grad);
  if (num_blocks <= 0) num_blocks = 32;
  if (algo == ALGO_ONESHOT) num_blocks = 1;
  if (algo == ALGO_NVLS) launch_rs<true>(world, num_blocks, stream, t, rank, bucket, start, shard_len, norm_acc);
  else launch_rs<false>(world, num_blocks, stream, t, rank, bucket, start, shard_len, norm_acc);
  return check_launch("comm_reduce_scatter");
}

// items: device int64 [n_items][4] = {arena offset, optimizer-state offset, count (multiple of 8),
//        group | (fp32_gather << 8)};  master_ptrs: HOST array of the peer-mapped fp32 master arenas
// mc_bf16 / mc_master: multicast addresses of the bf16 and master arenas (0: all-gather by P2P stores)
// hyper: host float[14] = lr[4], wd[4], beta1, beta2, eps, bias_correction1, bias_correction2, max_norm
// loss_in / loss_out: optional device scalars; loss_out receives the mean of loss_in over ranks (K21)
extern "C" int mb_comm_adamw_allgather(const long long* grad_ptrs, const long long* bf16_ptrs,
                                       const long long* ctrl_ptrs, const long long* master_ptrs, long long mc_bf16,
                                       long long mc_master, int world, int rank, const long long* items, int n_items,
                                       float* m_state, float* v_state, const float* hyper, float* norm_acc,
                                       float* norm_out, const float* loss_in, float* loss_out,
                                       unsigned int* block_counter, int num_blocks, cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, grad_ptrs, bf16_ptrs, ctrl_ptrs, master_ptrs, world)) return 1;
  if (mc_bf16 != 0 && mc_master != 0) {
    t.mc_bf16 = reinterpret_cast<bf16*>(mc_bf16);
    t.mc_master = reinterpret_cast<float*>(mc_master);
  }
  AdamArgs a;
  for (int i = 0; i < 4; ++i) { a.lr[i] = hyper[i]; a.wd[i] = hyper[4 + i]; }
  a.beta1 = hyper[8]; a.beta2 = hyper[9]; a.eps = hyper[10]; a.bc1 = hyper[11]; a.bc2 = hyper[12]; a.max_norm = hyper[13];
  if (num_blocks <= 0) num_blocks = 2 * num_sms();
  comm_adamw_allgather_kernel<<<num_blocks, kCommThreads, 0, stream>>>(t, world, rank, items, n_items, m_state, v_state, a,
                                                                       norm_acc, norm_out, loss_in, loss_out,
                                                                       block_counter, comm_timeout_ns());
  return check_launch("comm_adamw_allgather");
}

extern "C" int mb_comm_barrier(const long long* ctrl_ptrs, int world, int rank, cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, nullptr, nullptr, ctrl_ptrs, nullptr, world)) return 1;
  comm_barrier_kernel<<<1, 32, 0, stream>>>(t, world, rank, comm_timeout_ns());
  return check_launch("comm_barrier");
}

extern "C" int mb_comm_scalar_mean(const long long* ctrl_ptrs, int world, int rank, const float* in, float* out,
                                   cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, nullptr, nullptr, ctrl_ptrs, nullptr, world)) return 1;
  comm_scalar_mean_kernel<<<1, 32, 0, stream>>>(t, world, rank, in, out, comm_timeout_ns());
  return check_launch("comm_scalar_mean");
}

extern "C" int mb_comm_gather_master(const long long* master_ptrs, const long long* ctrl_ptrs, int world, int rank,
                                     long long start, long long shard_len, int num_blocks, cudaStream_t stream) {
  PeerTable t;
  if (fill_peers(t, nullptr, nullptr, ctrl_ptrs, master_ptrs, world)) return 1;
  MB_REQUIRE(shard_len % 4 == 0 && start % 4 == 0, "comm: shard must be 16-byte aligned");
  if (world == 1) return 0;
  if (num_blocks <= 0) num_blocks = 64;
  comm_gather_master_kernel<<<num_blocks, kCommThreads, 0, stream>>>(t, world, rank, start, shard_len);
  return check_launch("comm_gather_master");
}

// Writes the GPU's nanosecond wall clock to *out when the stream reaches this point (works inside a captured graph,
// where CUDA-event timing does not): used for the "exposed reduce-scatter" figure.
extern "C" int mb_timestamp(unsigned long long* out, cudaStream_t stream) {
  timestamp_kernel<<<1, 1, 0, stream>>>(out);
  return check_launch("timestamp");
}
