/* C ABI of libmamba_b200.so — hand-written sm_100a kernels for Mamba training on B200.
 *
 * Conventions
 *   - every launcher returns 0 on success, non-zero on error; mb_last_error() gives the text
 *     (thread-local).  Nothing here allocates device memory unless stated; all work is enqueued
 *     on `stream` and is asynchronous.
 *   - "bf16" pointers are void* to __nv_bfloat16 data; row strides (ld*) are in elements.
 *   - activations are channel-last; "strided" operands may be column slices of a wider matrix.
 * Each block cites the component of the reference stack it replaces (SURVEY.md section 2.3;
 * the reference itself, /root/reference, has no native code — these kernels stand in for the
 * mamba_ssm / causal-conv1d / cuBLAS / NCCL pieces it reaches through its imports).
 */
#ifndef MAMBA_B200_H_
#define MAMBA_B200_H_

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------ */
const char* mb_last_error(void);
int mb_version(void);
int mb_num_sms(void);
unsigned long long mb_launch_count(void); /* kernel launches issued by this library so far */

/* ---- K14: fused residual-add + RMSNorm  (mamba_ssm layer_norm_fn as called from Block,
 *      reached via /root/reference/model.py:35-41) ------------------------------------------ */
int mb_add_rmsnorm_fwd(const void* x, int x_is_fp32, const float* residual, const float* weight,
                       void* y_bf16, float* res_out, float* rstd, int M, int D, float eps,
                       cudaStream_t stream);
long long mb_norm_bwd_scratch_bytes(int rows, int D);
int mb_add_rmsnorm_bwd(const float* r, const void* dy_bf16, const float* rstd, const float* weight,
                       const float* dres_in, float* dres_out, void* dx_bf16, float* dw,
                       int dw_accumulate, float* scratch, int M, int D, cudaStream_t stream);
/* Same backward, no scratch argument and no allocation of any kind: the weight-gradient partials
 * go to a 1 MiB array built into the library image (held by every GPU that loads the library).
 * For hosts that cannot allocate — the device is owned by their own pool, or the call sits in a
 * loop that may not hit cudaMalloc/cudaFree.  Limits: D % 8 == 0, D <= 1024 (else an error is
 * returned).  One array per device, shared by all calls: overlapping calls on one GPU race.  The
 * first call on a device resolves the array's address; issue it outside stream capture.
 * (No counterpart in the reference: it has no native code and no such entry point.) */
int mb_add_rmsnorm_bwd_noalloc(const float* r, const void* dy_bf16, const float* rstd,
                               const float* weight, const float* dres_in, float* dres_out,
                               void* dx_bf16, float* dw, int dw_accumulate, int M, int D,
                               cudaStream_t stream);

/* ---- K15: gated RMSNorm  y = rmsnorm(x * silu(z)) * w  (mamba_ssm RMSNormGated) ----------- */
int mb_gated_rmsnorm_fwd(const void* x, int ldx, const void* z, int ldz, const float* weight,
                         void* out, int ldo, float* rstd, int M, int D, int group_size, float eps,
                         cudaStream_t stream);
int mb_gated_rmsnorm_bwd(const void* x, int ldx, const void* z, int ldz, const float* weight,
                         const float* rstd, const void* dout, int ldo, void* dx, int lddx, void* dz,
                         int lddz, float* dw, int dw_accumulate, float* scratch, int M, int D,
                         int group_size, cudaStream_t stream);

/* ---- K1-K3: causal depthwise conv1d (+bias, +SiLU), channel-last (causal-conv1d) ----------- */
int mb_conv1d_fwd(const void* x, long long ldx, const float* w, const float* bias, void* y,
                  long long ldy, int B, int T, int C, int width, int silu, cudaStream_t stream);
int mb_conv1d_bwd(const void* x, long long ldx, const float* w, const float* bias, const void* dy,
                  long long lddy, void* dx, long long lddx, float* dw, float* dbias, int B, int T,
                  int C, int width, int silu, cudaStream_t stream);
int mb_conv1d_update(const void* x, long long ldx, float* state, const float* w, const float* bias,
                     void* y, long long ldy, int B, int C, int width, int silu, cudaStream_t stream);

/* Work decomposition of the two conv kernels (host only, no launch): tokens = B*T on the flat token axis,
 * column_groups = groups of 32 threads across the channels, warps_per_sm = resident warps per SM.  Returns the
 * tile length (>= 64 tokens) and the number of tiles; work items = ntiles * column_groups, one per warp. */
int mb_conv1d_plan(long long tokens, int column_groups, int warps_per_sm, int* tile, int* ntiles);

/* ---- K18: embedding gather / scatter-add (ATen embedding fwd/bwd) ------------------------- */
int mb_embedding_fwd(const long long* ids, const float* w, float* out, int M, int D, int V,
                     cudaStream_t stream);
int mb_embedding_bwd(const long long* ids, const float* dout, float* dw, int M, int D, int V,
                     cudaStream_t stream);
int mb_cast_f32_bf16(const float* in, void* out_bf16, long long n, cudaStream_t stream);

/* ---- K17: cross-entropy over a slab of lm_head logits: per-row loss and, in the same launch,
 *      (softmax - onehot) * scale in bf16 for the dgrad / wgrad GEMMs (ATen log_softmax + nll_loss
 *      fwd/bwd behind F.cross_entropy, /root/reference/model.py:43-46).  The slab is bf16 in training
 *      (the gradient then overwrites it in place) or fp32 for loss-only scoring -------------------- */
int mb_ce_slab(const void* logits, int logits_bf16, long long ld, const long long* targets, float* loss,
               void* dlogits_bf16, long long ldd, int R, int V, float scale, cudaStream_t stream);

/* ---- K16: bf16 GEMM on tcgen05/TMEM/TMA (cuBLAS behind nn.Linear) ------------------------- */
int mb_gemm_bf16(const void* A, long long lda, int a_major, const void* B, long long ldb, int b_major,
                 void* D, long long ldd, int d_mode, const float* bias, int M, int N, int K,
                 float alpha, cudaStream_t stream);

/* ---- K7-K12: Mamba-2 SSD chunked scan (mamba_ssm Triton ssd_* kernels, forward + backward) ----
 * Layouts: dt, a = cumsum(dt*A) per chunk: fp32 [B, nc, H, L]; states: fp32 [B, nc, H, P, N];
 * bf16 state operand copies: [B, nc, H, P, N]; x | B | C are column slices of the conv output. */
int mb_ssd_prep(const void* dt_raw, long long ld_dt, const float* dt_bias, const float* A_log,
                float* dt_out, float* a_out, int B, int T, int H, int L, float dt_min, float dt_max,
                cudaStream_t stream);
int mb_ssd_chunk_state(const void* arow, long long lda, const void* bop, long long ldb, const float* dt,
                       const float* a, float* out, int dstate, int B, int T, int H, int G, int P, int N,
                       int L, cudaStream_t stream);
/* K9 + K10 fused: chunk states and the inter-chunk recurrence in one kernel (one CTA per batch x head pair walks
 * the chunks in order and carries the state in registers), with native initial / final states and their
 * gradients (mamba_chunk_scan_combined's initial_states / return_final_states).  dstate = 0: out = bf16 state
 * entering each chunk; dstate = 1: arow = dy, bop = Cm, out = bf16 gradient of each chunk's own state,
 * da_last += exp(a_last) <ds, sprev> (pre-zeroed), init = gradient of the final state, last_out = gradient of the
 * initial state.  mb_ssd_states_seq_ok says whether the shape has enough (batch, head pair) parallelism for it. */
int mb_ssd_states_seq_ok(int B, int H, int G);
int mb_ssd_states_seq(const void* arow, long long lda, const void* bop, long long ldb, const float* dt,
                      const float* a, void* out_bf16, const void* sprev_bf16, float* da_last, const float* init,
                      float* last_out, int dstate, int B, int T, int H, int G, int P, int N, int L,
                      cudaStream_t stream);
int mb_ssd_state_passing_fwd(const float* states, const float* a, void* sprev_bf16, float* final_state,
                             int B, int nc, int H, int PN, int L, cudaStream_t stream);
int mb_ssd_state_passing_bwd(const float* dstates_direct, const void* sprev_bf16, const float* a,
                             void* ds_bf16, float* da_last, int B, int nc, int H, int PN, int L,
                             cudaStream_t stream);
int mb_ssd_chunk_scan_fwd(const void* Cm, const void* Bm, long long ldbc, const void* x, long long ldx,
                          const void* sprev_bf16, const float* dt, const float* a, const float* D, void* y,
                          void* y_off, long long ldy, int B, int T, int H, int G, int P, int N, int L,
                          cudaStream_t stream);
int mb_ssd_chunk_scan_bwd_dx(const void* Bm, const void* Cm, long long ldbc, const void* dy, long long lddy,
                             const void* ds_bf16, const float* dt, const float* a, const float* D, void* dx,
                             long long lddx, const void* x, long long ldx, float* ddt1, float* ddt2,
                             float* dD, int B, int T, int H, int G, int P, int N, int L, cudaStream_t stream);
int mb_ssd_bwd_dcb(const void* dy, long long lddy, const void* x, long long ldx, const void* Cm,
                   const void* Bm, long long ldbc, const float* dt, const float* a, void* dcb,
                   float* rowsum_m, int B, int T, int H, int G, int P, int N, int L, cudaStream_t stream);
/* dB or dC may be NULL: that kernel is skipped (the two are independent and may run on different streams). */
int mb_ssd_bwd_dbdc(const void* dcb, const void* Bm, const void* Cm, long long ldbc, const void* dy,
                    long long lddy, const void* x, long long ldx, const void* sprev_bf16,
                    const void* ds_bf16, const float* dt, const float* a, void* dB, void* dC, long long ldo,
                    int B, int T, int H, int G, int P, int N, int L, cudaStream_t stream);
int mb_ssd_bwd_finalize(const void* dy, long long lddy, const void* y_off, long long ldy,
                        const void* dt_raw, long long ld_dt, const float* dt, const float* ddt1,
                        const float* ddt2, const float* rowsum_m, const float* da_last,
                        const float* dt_bias, const float* A_log,
                        void* draw_out, long long ld_draw, float* dA_log, float* ddt_bias, int B, int T, int H,
                        int P, int L, float dt_min, float dt_max, cudaStream_t stream);

/* ---- K4 / K5 / K23: Mamba-1 selective scan fwd / bwd / single-token update (mamba_ssm
 * csrc/selective_scan + Triton selective_state_update), channel-last, d_state = 16.
 * fwd scratch: mb_sscan_scratch_floats() floats, must be handed back unchanged to bwd; bwd scratch2:
 * two thirds of that.  dA, dB, dC (fp32 [B*T, N]), dD, dbias are accumulated into: pre-zero them. */
long long mb_sscan_scratch_floats(int B, int T, int Dn, int N);
int mb_sscan_fwd(const void* u, long long ldu, const void* delta, long long ldd, const float* A,
                 const void* Bm, long long ldb, const void* Cm, long long ldc, const float* D,
                 const void* z, long long ldz, const float* delta_bias, void* y, long long ldy,
                 float* scratch, int B, int T, int Dn, int N, int softplus, cudaStream_t stream);
int mb_sscan_bwd(const void* u, long long ldu, const void* delta, long long ldd, const float* A,
                 const void* Bm, long long ldb, const void* Cm, long long ldc, const float* D,
                 const void* z, long long ldz, const float* delta_bias, const void* dy, long long lddy,
                 const float* scratch, float* scratch2, void* du, long long lddu, void* ddelta,
                 long long lddd, float* dA, float* dB, float* dC, float* dD, void* dz, long long lddz,
                 float* dbias, int B, int T, int Dn, int N, int softplus, cudaStream_t stream);
/* K23 (Mamba-2 decode, mamba_ssm selective_state_update with heads): in-place fp32 state [B,H,P,N] step for
 * one token; x, z, y [B, H*P] bf16 rows, dt_raw [B, H] bf16, Bm / Cm [B, G*N] bf16 rows (stride ldbc). */
int mb_ssd_state_update(float* state, const void* x, long long ldx, const void* dt_raw, long long lddt,
                        const float* A, const void* Bm, const void* Cm, long long ldbc, const float* D,
                        const void* z, long long ldz, const float* dt_bias, void* y, long long ldy,
                        int B, int H, int G, int P, int N, int softplus, cudaStream_t stream);
int mb_sscan_update(float* state, const void* x, long long ldx, const void* delta, long long ldd,
                    const float* A, const void* Bm, long long ldb, const void* Cm, long long ldc,
                    const float* D, const void* z, long long ldz, const float* delta_bias, void* y,
                    long long ldy, int B, int Dn, int N, int softplus, cudaStream_t stream);

/* ---- K19-K21: gradient reduce-scatter + global-norm clip + loss mean + sharded AdamW + bf16
 * parameter all-gather over NVLink peer / NVLS multicast memory (replaces DDP's NCCL all-reduce,
 * the scalar loss all-reduce, clip_grad_norm_ and torch.optim.AdamW of
 * /root/reference/train.py:86,219-227).  *_ptrs are HOST arrays of `world` device addresses of
 * the peer-mapped gradient / bf16-parameter / control / fp32-master arenas; mc_* are the
 * multicast addresses of the same arenas (0 when the system has no multicast object: the P2P
 * forms are used).  The step sequence number lives in the control block, so no launch argument
 * changes from step to step.
 *   algo (reduce-scatter, chosen per bucket): 0 P2P pull, 1 NVLS multimem.ld_reduce,
 *   2 one-shot single-CTA pull for buckets under 256 KiB. */
long long mb_comm_ctrl_words(void);
int mb_comm_set_timeout_ms(long long ms); /* wall-clock budget of every cross-GPU wait (default 60 s) */
int mb_comm_reduce_scatter(const long long* grad_ptrs, const long long* ctrl_ptrs, long long mc_grad,
                           int world, int rank, int bucket, long long start, long long shard_len,
                           float* norm_acc, int algo, int num_blocks, cudaStream_t stream);
int mb_comm_adamw_allgather(const long long* grad_ptrs, const long long* bf16_ptrs,
                            const long long* ctrl_ptrs, const long long* master_ptrs, long long mc_bf16,
                            long long mc_master, int world, int rank, const long long* items,
                            int n_items, float* m_state, float* v_state, const float* hyper,
                            float* norm_acc, float* norm_out, const float* loss_in, float* loss_out,
                            unsigned int* block_counter, int num_blocks, cudaStream_t stream);
/* Stand-alone pieces of the same control-block protocol (not on the step path):
 *   barrier over all ranks; mean of one fp32 scalar over ranks (validation loss,
 *   /root/reference/train.py:147); pull of every peer's owned fp32 master shard of one bucket
 *   (checkpointing a ZeRO-1 sharded model, /root/reference/train.py:152-163). */
int mb_comm_barrier(const long long* ctrl_ptrs, int world, int rank, cudaStream_t stream);
int mb_comm_scalar_mean(const long long* ctrl_ptrs, int world, int rank, const float* in, float* out,
                        cudaStream_t stream);
int mb_comm_gather_master(const long long* master_ptrs, const long long* ctrl_ptrs, int world, int rank,
                          long long start, long long shard_len, int num_blocks, cudaStream_t stream);
/* GPU wall clock (ns) written to *out when the stream gets there; usable inside graph capture. */
int mb_timestamp(unsigned long long* out, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MAMBA_B200_H_ */
