/* C ABI of libdpfs_b200.so — the sm_100a kernel library behind distributed_pytorch_from_scratch_b200.
 *
 * Conventions
 *   - Every entry point enqueues work on `stream` and returns immediately; the return value is 0 on success,
 *     a positive cudaError_t, or a negative library error (bad alignment / unsupported shape).
 *   - Pointers are device pointers unless stated otherwise. bf16 tensors are `__nv_bfloat16` (passed as void*).
 *   - No entry point allocates device memory unless its comment says so; scratch is passed in by the caller.
 *   - "ref:" comments cite the file:line in ldh127/distributed_pytorch_from_scratch whose behaviour the kernel
 *     provides on B200.
 */
#pragma once

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- GEMM (tcgen05 / TMEM / TMA) -------------------------------------------------------------------------
 * out[M,N] = A[M,K] * B[N,K]^T (+ bias[N]) (+ residual[M,N]),  bf16 inputs, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A is stored [M][lda] (K contiguous).   1: A is stored [K][lda] (M contiguous).
 *   b_mn_major = 0: B is stored [N][ldb] (K contiguous).   1: B is stored [K][ldb] (N contiguous).
 *   out_fp32   = 0: out is bf16 [M][ldo];                  1: out is fp32 [M][ldo].
 *   bias (fp32, nullable), residual (fp32 [M][ldr], nullable, only with splits == 1).
 *   splits > 1 : split-K over `splits` CTAs groups; `workspace` must hold splits*M*N floats.
 *   bn         : N tile (128 or 256; 0 = 128).
 * Requirements: N % 8 == 0, lda/ldb % 8 == 0, 16-byte aligned bases.
 * ref: F.linear in models/layers.py:49,93 and its autograd dgrad/wgrad (cuBLAS in the reference). */
int dpfs_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                   void* out, long long ldo, int out_fp32, const float* bias, const float* residual,
                   long long ldr, int M, int N, int K, int splits, float* workspace, int bn,
                   cudaStream_t stream);

/* Same contract on CTA pairs: a 2-CTA cluster computes 256x256 tiles with tcgen05.mma.cta_group::2; each CTA stages
 * its 128 rows of A and half of the B tile, i.e. 2/3 of the operand traffic of the 128x256 1-SM kernel. */
int dpfs_gemm2sm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                      void* out, long long ldo, int out_fp32, const float* bias, const float* residual,
                      long long ldr, int M, int N, int K, int splits, float* workspace, cudaStream_t stream);

/* Fused gate|up projection + SwiGLU, forward (2-SM kernel with the activation in the epilogue): gu[M][2F] = A W^T + b
 * (bf16; gate columns [0,F), up columns [F,2F): the backward reads it) and act[M][F] = silu(gate) * up, in ONE pass.
 * W: bf16 [2F][ldb], gate rows then up rows (the two weights back to back), K-major; bias fp32 [2F] (nullable).
 * F % 128 == 0.  ref: FFN.forward, models/model.py:95 (two cuBLAS GEMMs + two ATen kernels in the reference). */
int dpfs_gemm2sm_swiglu_bf16(const void* A, long long lda, const void* B, long long ldb, void* gu, void* act,
                             const float* bias, int M, int F, int K, cudaStream_t stream);

/* Fused down_proj dgrad + SwiGLU backward (2-SM kernel, SwiGLU' in the epilogue): dgu[M][2F] = swiglu'(dy W, gu) in
 * bf16 and bias_grad[2F] = column sums of dgu (nullable).  dy: bf16 [M][ldy]; W: the down_proj weight as stored,
 * bf16 [K = d][ldw] with F contiguous; gu: the saved gate|up activation [M][2F].  F % 256 == 0.  scratch: fp32
 * [dpfs_gemm2sm_swiglu_bwd_rows(M)][2F].  dact = dy W is never written to memory.
 * ref: autograd of FFN.forward, models/model.py:95 (a cuBLAS dgrad GEMM + ~6 ATen kernels in the reference). */
int dpfs_gemm2sm_swiglu_bwd_bf16(const void* dy, long long ldy, const void* W, long long ldw, const void* gu, void* dgu,
                                 float* bias_grad, float* scratch, int M, int F, int K, cudaStream_t stream);
int dpfs_gemm2sm_swiglu_bwd_rows(int M);

/* Fused GEMM -> all-reduce over NVLink (one kernel): every rank computes its partial out = A_r B_r^T (bf16) into
 * its symmetric buffer `out` ([M][N], row stride N) and the ranks sum the tiles through the NVSwitch with
 * multimem.ld_reduce / multimem.st on `mc_out` (the multicast alias of `out`).  flag_ptrs: HOST array of the n ranks'
 * flag-buffer addresses (dpfs_comm_flag_words() uint32 each, zero-initialised, peer-mapped).  On return of the kernel
 * every rank's `out` holds the full sum.  ref: RowParallelLinear.forward GEMM + Reduce (models/layers.py:49-51,
 * models/comm_ops.py:39) and the Copy.backward all-reduce after the dgrad GEMM (models/comm_ops.py:59). */
int dpfs_gemm_allreduce_bf16(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, void* out,
                             void* mc_out /* NVLS reduce path; NULL = P2P reduce path */,
                             const unsigned long long* data_ptrs /* HOST array: every rank's symmetric base */,
                             long long out_offset /* byte offset of `out` inside the symmetric buffer */, int M,
                             int N, int K, const unsigned long long* flag_ptrs, int rank, int n, int bn,
                             cudaStream_t stream);

/* Fused GEMM -> reduce-scatter over rows (one kernel; the sequence-parallel form of the same call sites, SURVEY
 * K-c1 "rowpar_gemm_reducescatter"): partial products go to the symmetric scratch `partial` ([M][N] bf16) and rank r
 * receives rows [r*M/n, (r+1)*M/n) of the sum in its PRIVATE rs_out ([M/n][N] bf16).  Row-blocks are computed in
 * owner-interleaved rounds so every owner's reduction overlaps the GEMM from the start.  mc_partial != NULL: the owner
 * pulls switch-reduced rows with multimem.ld_reduce (NVLS); NULL: it sums the peers' copies in rank order (P2P).
 * Requires M % n == 0 and M / n >= 128.  No counterpart in the reference (it has no sequence parallelism; its
 * Reduce is always a full all-reduce, models/comm_ops.py:39). */
int dpfs_gemm_reduce_scatter_bf16(const void* A, long long lda, const void* B, long long ldb, int b_mn_major,
                                  void* partial, void* mc_partial, const unsigned long long* data_ptrs,
                                  long long partial_offset, void* rs_out, int M, int N, int K,
                                  const unsigned long long* flag_ptrs, int rank, int n, int bn, cudaStream_t stream);

/* ---- Sequence-parallel hand-offs: reduce-scatter -> row op -> all-gather in ONE kernel (csrc/sp_fused.cu) ------------
 * mc_partial: MULTICAST alias of the [n * rows_per_rank][D] bf16 partial products every rank wrote at the same
 * symmetric offset (the output of its row-parallel GEMM).  Rank r owns rows [r * rows_per_rank, (r+1) * rows_per_rank).
 * One warp per row: multimem.ld_reduce pulls the switch-reduced row, the row op runs in registers, multimem.st pushes
 * the bf16 result row into every rank's copy of the landing zone (mc_h / mc_dx: MULTICAST alias of [n * rpr][D] bf16).
 * Two rank barriers (partials complete / results landed) per call.  D % 8 == 0, D <= 1024.
 *
 * forward:  xout_l = x_l + sum_r partial_r + bias (fp32, [rpr][D]; x_l / bias nullable), rstd_l[rpr],
 *           h = rmsnorm(xout_l) * scale -> every rank.   mc_h == NULL: residual add only (no norm, no broadcast).
 * ref: Reduce after the row-parallel GEMM + bias + residual + the next RMSNorm (models/comm_ops.py:39,
 * models/layers.py:51-54, models/model.py:119-120); the reference has no sequence parallelism. */
int dpfs_rs_add_rmsnorm_ag(const void* mc_partial, const float* x_l, const float* bias, const float* scale,
                           float* xout_l, float* rstd_l, void* mc_h, int rows_per_rank, int D, float eps,
                           const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
/* backward: dh = sum_r partial_r (this rank's rows of a column-parallel dgrad); dx = dres_l + rmsnorm'(dh) with the
 * saved x_l / rstd_l; dx -> dx_l (fp32 private, nullable) and bf16 -> every rank (mc_dx, nullable).
 * dscale_out[D] (+)= sum over this rank's rows of dh * xhat, dbias_out[D] (+)= column sums of dx over this rank's rows
 * (both nullable; partial sums over LOCAL rows: the caller all-reduces them once per step).  scratch: fp32
 * [dpfs_sp_blocks(rows_per_rank, D)][2 * D].
 * ref: Copy.backward all-reduce after the dgrad GEMM (models/comm_ops.py:59) + RMSNorm backward. */
int dpfs_rs_rmsnorm_bwd_ag(const void* mc_partial, const float* x_l, const float* rstd_l, const float* scale,
                           const float* dres_l, float* dx_l, void* mc_dx, float* dscale_out, float* dbias_out,
                           int accumulate, float* scratch, int rows_per_rank, int D,
                           const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
int dpfs_sp_blocks(int rows_per_rank, int D);
/* The same hand-offs with the row-parallel GEMM in front of them in the SAME kernel (csrc/gemm_handoff.cu): the 1-SM
 * tcgen05 GEMM writes bf16 partial tiles into `partial` (this rank's symmetric buffer, multicast alias mc_partial) and
 * flags them to the owners of their rows; comm warps of the same CTAs pull / row-op / push 16-row slices as soon as all
 * n ranks have flagged their tiles.  No opening barrier (tile flags carry a launch epoch kept in device memory, so the
 * call is CUDA-graph replayable), one closing grid-wide rank barrier.  A: bf16 [M][lda] (K-major); B: [N][ldb] or, with
 * b_mn_major, [K][ldb]; N = model width, N % 8 == 0, N <= 1024, M % n == 0, M / n >= 128.  Other arguments as above;
 * the backward's scratch holds dpfs_gemm_handoff_scratch_rows() x 2N floats.  No counterpart in the reference beyond
 * the call sites cited above (it runs cuBLAS, then NCCL, then ATen kernels). */
int dpfs_gemm_handoff_fwd(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, void* partial,
                          const void* mc_partial, int M, int N, int K, const float* x_l, const float* bias,
                          const float* scale, float* xout_l, float* rstd_l, void* mc_h, float eps,
                          const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
int dpfs_gemm_handoff_bwd(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, void* partial,
                          const void* mc_partial, int M, int N, int K, const float* x_l, const float* rstd_l,
                          const float* scale, const float* dres_l, float* dx_l, void* mc_dx, float* dscale_out,
                          float* dbias_out, float* scratch, const unsigned long long* flag_ptrs, int rank, int n,
                          cudaStream_t stream);
int dpfs_gemm_handoff_scratch_rows(void);
/* host-side view of that kernel's tile order and comm work items, for the CPU tests */
int dpfs_debug_gh_schedule(int M, int N, int n, int rank, int* order, int* item_lo, int* item_hi, int* blk_lo,
                           int* blk_hi, int cap);
/* Tuning hook of the two stand-alone kernels above (tools/bench_handoff.py): grid_cap = most blocks to launch (0 = as many as are
 * resident at once), bar_mode = 0 one rank barrier per block (n remote atomics each), 1 grid-wide barrier with a
 * single n-way exchange per rank (default), 2 no barrier (timing of the data phase only; results are undefined). */
int dpfs_sp_debug_config(int grid_cap, int bar_mode);
/* `reps` rank barriers in each of `blocks` blocks of `threads` threads: the cost of the barrier itself. */
int dpfs_comm_barrier_ex(const unsigned long long* flag_ptrs, int rank, int n, int blocks, int threads, int reps,
                         cudaStream_t stream);


/* Host-only (no GPU): the work items (heavy-first list positions) CTA `bid` of `grid` persistent attention CTAs
 * processes, in order; snake != 0 = the reversed-odd-rounds order of the opt-in kernels.  Returns the count. */
int dpfs_debug_attn_items(int n_items, int grid, int bid, int snake, int* out, int cap);

/* Host-only (no GPU): the fused reduce-scatter's schedule for M rows over n ranks -- order[i] = row-block computed
 * i-th, owner_lo/hi[b] = ranks signalled when block b completes (arrays of ceil(M/128) ints).  For the CPU tests. */
int dpfs_debug_rs_schedule(int M, int n, int* order, int* owner_lo, int* owner_hi);

/* ---- Stand-alone NVLS collectives over symmetric memory ----------------------------------------------------------
 * In-place two-shot all-reduce of the `nbytes` at multicast address mc_ptr (every rank passes the same offset),
 * all-gather of nbytes_per_rank from local `src` into slot `rank` of mc_dst on every rank, and a rank-wide barrier.
 * ref: dist.all_reduce / dist.all_gather in models/comm_ops.py:26,39,59,74 (NCCL in the reference). */
int dpfs_comm_flag_words(void);
int dpfs_allreduce_nvls(void* mc_ptr, long long nbytes, int is_bf16, const unsigned long long* flag_ptrs, int rank,
                        int n, cudaStream_t stream);
/* Same all-reduce over plain peer loads / stores (fp32 accumulation in rank order; n in {2,4,8}). */
int dpfs_allreduce_p2p(const unsigned long long* data_ptrs, long long offset, long long nbytes, int is_bf16,
                       const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
/* Sequence-parallel halves of the all-reduce.  reduce-scatter: out (local bf16, nbytes_total / n bytes) = sum over
 * ranks of this rank's block of rows of the symmetric partial matrix at `offset`.  all-gather: every rank's
 * nbytes_per_rank block lands in slot `rank` of every rank's symmetric buffer at dst_offset.  No counterpart in the
 * reference (it has no sequence parallelism; SURVEY §2.3 lists it as a required extension). */
int dpfs_reduce_scatter_p2p(const unsigned long long* data_ptrs, long long offset, long long nbytes_total, void* out,
                            const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
int dpfs_reduce_scatter_nvls(const void* mc_ptr, long long nbytes_total, void* out,
                             const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
int dpfs_allgather_p2p(const void* src, const unsigned long long* data_ptrs, long long dst_offset,
                       long long nbytes_per_rank, const unsigned long long* flag_ptrs, int rank, int n,
                       cudaStream_t stream);
int dpfs_allgather_nvls(const void* src, void* mc_dst, long long nbytes_per_rank,
                        const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);
/* experimental data-path variants of the all-reduce (tools/gpu_ar_sweep.py); mode 3/4 are correct all-reduces */
int dpfs_allreduce_exp(void* mc_base, const unsigned long long* data_ptrs, long long offset, long long nbytes,
                       int mode, int blocks, const unsigned long long* flag_ptrs, int rank, int n,
                       cudaStream_t stream);
int dpfs_comm_barrier(const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream);

/* number of kernels launched by this library so far (bench.py's gpu_launches) */
unsigned long long dpfs_launch_count(void);

/* ---- Fused causal flash attention (tcgen05 / TMEM / TMA) ---------------------------------------------------
 * qkv: bf16 [B*T][3*H*HD] packed q|k|v (RoPE already applied); out: bf16 [B*T][H*HD]; lse: fp32 [B][H][T].
 * HD must be 64 or 128.  ref: Attention.forward, models/model.py:68-78 (matmul + triu mask + softmax + matmul). */
int dpfs_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int HD, float scale,
                  cudaStream_t stream);
/* Backward: writes dqkv bf16 [B*T][3*H*HD].  Scratch supplied by the caller: delta fp32 [B][H][T] and
 * dq_acc fp32 [B*T][H*HD] (zeroed here, accumulated with red.global.add.v4.f32 across kv tiles). */
int dpfs_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                  float* dq_acc, void* dqkv, int B, int T, int H, int HD, float scale, cudaStream_t stream);

/* ---- Normalisation ------------------------------------------------------------------------------------------
 * One warp per row, fp32 statistics, D % 4 == 0 and D <= 2048.  *_bf16 flags select the element type (0 = fp32).
 * ref: RMSNorm, models/layers.py:145-155.  (LayerNorm has no counterpart in the reference model; it is provided
 * because the project brief names it.) */
int dpfs_norm_bwd_blocks(int M); /* number of CTAs the backward kernels launch == rows of scratch they need */
int dpfs_rmsnorm_fwd(const void* x, int x_bf16, const float* scale, void* y, int y_bf16, float* rstd /*nullable*/,
                     int M, int D, float eps, cudaStream_t stream);
/* x_out = x + y + bias (fp32); h = rmsnorm(x_out) * scale (bf16, nullable together with scale/rstd).  y is the
 * all-reduced bf16 output of a row-parallel GEMM.  ref: models/layers.py:51-54 + models/model.py:119-120. */
int dpfs_add_rmsnorm_fwd(const float* x, const void* y_bf16, const float* bias /*nullable*/,
                         const float* scale /*nullable*/, float* x_out, void* h_bf16 /*nullable*/,
                         float* rstd /*nullable*/, int M, int D, float eps, cudaStream_t stream);
/* dx = dres + d(rmsnorm)(dy); writes dx (fp32, nullable) and/or dx_bf16 (nullable); dscale[D] (+= if accumulate).
 * scratch: dpfs_norm_bwd_blocks(M) * D floats. */
int dpfs_rmsnorm_bwd(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* rstd, const float* scale,
                     const float* dres /*nullable*/, float* dx, void* dx_bf16, float* dscale, int dscale_accumulate,
                     float* scratch, int M, int D, cudaStream_t stream);
/* Same, plus dbias[D] = column sums of the produced dx (overwritten): dx is the output gradient of the row-parallel
 * linear in front of the norm (wo / down_proj, ref models/layers.py:51-54), so its bias gradient comes out of this
 * pass.  scratch: dpfs_norm_bwd_blocks(M) * 2 * D floats. */
int dpfs_rmsnorm_bwd_bias(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* rstd,
                          const float* scale, const float* dres /*nullable*/, float* dx, void* dx_bf16, float* dscale,
                          int dscale_accumulate, float* dbias, float* scratch, int M, int D, cudaStream_t stream);
/* ---- "push" all-gather variants (sequence parallelism; no counterpart in the reference, which has no SP) --------
 * The bf16 activation a kernel produces is written through the NVLS MULTICAST mapping of the all-gather landing zone
 * (`*_mc` = multicast address of THIS rank's block), so it lands in every rank's copy as it is produced; the caller
 * follows with dpfs_comm_barrier.  Replaces "write local block + all-gather kernel".  fp32 outputs stay local. */
int dpfs_rmsnorm_fwd_mc(const float* x, const float* scale, void* y_mc, float* rstd /*nullable*/, int M, int D,
                        float eps, cudaStream_t stream);
int dpfs_add_rmsnorm_fwd_mc(const float* x, const void* y_bf16, const float* bias /*nullable*/, const float* scale,
                            float* x_out, void* h_mc, float* rstd /*nullable*/, int M, int D, float eps,
                            cudaStream_t stream);
/* dy bf16, x fp32.  scratch: dpfs_norm_bwd_blocks(M) * D floats. */
int dpfs_rmsnorm_bwd_mc(const void* dy_bf16, const float* x, const float* rstd, const float* scale,
                        const float* dres /*nullable*/, float* dx /*nullable*/, void* dx_mc, float* dscale,
                        int dscale_accumulate, float* scratch, int M, int D, cudaStream_t stream);
int dpfs_cast_f32_to_bf16_mc(const float* x, void* y_mc, long long n, cudaStream_t stream);
int dpfs_layernorm_fwd(const void* x, int x_bf16, const float* gamma, const float* beta /*nullable*/, void* y,
                       int y_bf16, float* mean, float* rstd, int M, int D, float eps, cudaStream_t stream);
/* scratch: (dpfs_norm_bwd_blocks(M) + 1) * 2 * D floats. */
int dpfs_layernorm_bwd(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* mean, const float* rstd,
                       const float* gamma, float* dx, float* dgamma, float* dbeta /*nullable*/, float* scratch,
                       int M, int D, cudaStream_t stream);

/* ---- Activations ----------------------------------------------------------------------------------------------
 * gu: bf16 [M][2F] = gate | up;  act = silu(gate) * up.  ref: FFN.forward, models/model.py:95. */
int dpfs_swiglu_fwd(const void* gu, void* act, int M, int F, cudaStream_t stream);
int dpfs_swiglu_bwd(const void* dact, const void* gu, void* dgu, int M, int F, cudaStream_t stream);
/* swiglu backward that also produces the gate|up bias gradient (column sums of dgu) from the same pass.
 * scratch: dpfs_colsum_scratch_floats(M, 2F) floats.  ref: autograd of models/model.py:95 + models/layers.py:96. */
int dpfs_swiglu_bwd_bias(const void* dact, const void* gu, void* dgu, float* bias_grad, int accumulate,
                         float* scratch, int M, int F, cudaStream_t stream);
/* scratch floats dpfs_swiglu_bwd_bias needs for an [M][2F] gate|up activation (strip height depends on F) */
int dpfs_swiglu_bwd_scratch_floats(int M, int F);
/* tanh-GELU on bf16, n % 8 == 0 (no counterpart in the reference model; named by the project brief). */
int dpfs_gelu_fwd(const void* x, void* y, long long n, cudaStream_t stream);
int dpfs_gelu_bwd(const void* dy, const void* x, void* dx, long long n, cudaStream_t stream);

/* ---- RoPE -----------------------------------------------------------------------------------------------------
 * In-place rotate_half RoPE on the first `nseg` segments of a packed bf16 [M][ld] buffer; each segment has H heads
 * of hd; cos/sin: fp32 [maxlen][hd/2]; pos: int64 [M]; inverse != 0 applies the transpose rotation (backward).
 * ref: rotate_half / apply_rotary_pos_emb / get_cos_sin, models/model.py:17-46. */
int dpfs_rope_inplace(void* buf, long long ld, const long long* pos, const float* cos_t, const float* sin_t, int M,
                      int nseg, int seg_stride, int H, int hd, int inverse, cudaStream_t stream);

/* ---- Column sum (bias gradients) ------------------------------------------------------------------------------
 * out[N] (+)= sum over rows of bf16 x[M][ld].  Two stages through `scratch` (dpfs_colsum_scratch_floats(M, N)
 * floats), summed in a fixed order.  ref: the bias-grad reductions autograd runs for models/layers.py:54,96. */
int dpfs_colsum_scratch_floats(int M, int N);
int dpfs_colsum_bf16(const void* x, long long ld, float* out, int accumulate, float* scratch, int M, int N,
                     cudaStream_t stream);
int dpfs_cast_f32_to_bf16(const float* x, void* y, long long n, cudaStream_t stream);
/* y = bf16(x) for fp32 x [M][N] plus colsum[N] (+)= column sums of y, one pass (the residual-stream gradient feeds a
 * GEMM as bf16 and is also the gradient of the replicated row-parallel bias).  scratch: ceil(M / 16) * N floats. */
int dpfs_cast_colsum(const float* x, void* y, float* colsum, int accumulate, float* scratch, int M, int N,
                     cudaStream_t stream);

/* Column sum for callers that cannot allocate device memory and have no scratch buffer to lend: a host process whose
 * own allocator already owns the whole GPU (so a cudaMalloc may fail, and would stall the device), or a plain C /
 * embedded host with no device allocator at all.  Same result as dpfs_colsum_bf16, but the partial sums go through a
 * fixed array that is part of the library image, so nothing is allocated (no cudaMalloc / cudaFree / allocating
 * library call).  This entry point has no counterpart in the reference.
 *   - Limit: ceil(M / 64) * N <= DPFS_COLSUM_STATIC_MAX_FLOATS; larger problems return -25 and launch nothing.
 *   - x must be 16-byte aligned (-26 otherwise: the kernel reads 16-byte vectors) and M <= 65535 * 64 rows (-27).
 *   - Cost: the array (8 MiB) is resident on every GPU that loads the library until the process exits.
 *   - The first call on a device resolves the array's device address; make that first call before any stream
 *     capture on that device.
 *   - Every call shares the array and it keeps the previous call's contents (nothing needs clearing: each slot is
 *     written before it is read).  Calls that can overlap on the GPU race on it; keep them ordered. */
#define DPFS_COLSUM_STATIC_MAX_FLOATS (2 * 1024 * 1024)
int dpfs_colsum_bf16_static(const void* x, long long ld, float* out, int accumulate, int M, int N,
                            cudaStream_t stream);

/* ---- Vocab-parallel embedding -----------------------------------------------------------------------------------
 * out[m] = weight[ids[m] - vocab_start] for owned ids, 0 otherwise (ids are never modified).
 * ref: ParallelVocabularyEmbedding.forward, models/layers.py:134-141. */
int dpfs_embedding_fwd(const long long* ids, const float* weight, void* out, int out_bf16, int M, int D,
                       long long vocab_start, long long vocab_end, cudaStream_t stream);
/* Vocab-parallel lookup + its collective in one kernel (SURVEY K-c3): the rank that owns a token's row writes it (bf16)
 * through the multicast address out_mc of the [M][D] output, the others write nothing; follow with dpfs_comm_barrier.
 * ref: ParallelVocabularyEmbedding.forward + Reduce (models/layers.py:134-141). */
int dpfs_embedding_fwd_mc(const long long* ids, const float* weight, void* out_mc, int M, int D, long long vocab_start,
                          long long vocab_end, cudaStream_t stream);
int dpfs_embedding_bwd(const long long* ids, const float* dout, float* dweight, int zero_first,
                       long long local_vocab, int M, int D, long long vocab_start, long long vocab_end,
                       cudaStream_t stream);
/* same, dout is bf16 [M][D] */
int dpfs_embedding_bwd_bf16(const long long* ids, const void* dout_bf16, float* dweight, int zero_first,
                            long long local_vocab, int M, int D, long long vocab_start, long long vocab_end,
                            cudaStream_t stream);

/* ---- Vocab-parallel softmax cross-entropy ------------------------------------------------------------------------
 * Phase 1 (dpfs_ce_stats): stats[m] = {max, sum exp(x - max), target logit if owned} over the local shard
 * [M][Vl]; loss_acc[2] is reset and loss_acc[1] receives the number of non-ignored targets.  The ranks then
 * exchange their stats into all_stats [nranks][M][3] (tiny all-gather).  dpfs_ce_loss: loss_acc[0] += sum of row
 * losses (mean loss = loss_acc[0] / loss_acc[1]); row_loss nullable.  dpfs_ce_grad: dlogits (bf16 [M][ldd]) =
 * (softmax - onehot) * upstream[0] / n_valid, where `upstream` is a device scalar (nullable = 1).
 * ref: F.cross_entropy on gathered logits, train.py:101-104 + Gather in models/layers.py:99. */
int dpfs_ce_stats(const void* logits, int logits_bf16, long long ld, const long long* targets, float* stats,
                  float* loss_acc, int M, int Vl, long long vocab_start, long long ignore_index,
                  cudaStream_t stream);
int dpfs_ce_loss(const long long* targets, const float* all_stats, int nranks, float* loss_acc, float* row_loss,
                 int M, long long ignore_index, cudaStream_t stream);
int dpfs_ce_grad(const void* logits, int logits_bf16, long long ld, const long long* targets,
                 const float* all_stats, int nranks, void* dlogits, long long ldd, const float* loss_acc,
                 const float* upstream, int M, int Vl, long long vocab_start, long long ignore_index,
                 cudaStream_t stream);

/* ---- Fused Adam / AdamW over a flat arena ------------------------------------------------------------------------
 * hyper (device): {lr, beta1, beta2, eps, bias_correction1, bias_correction2, decoupled weight decay}.
 * Also refreshes the bf16 shadow of the parameters (nullable).  n % 4 == 0.
 * ref: optim.Adam + OneCycleLR (which cycles lr and beta1 every step), train.py:83-84,106-109. */
int dpfs_adam_step(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n, const float* hyper,
                   cudaStream_t stream);

#ifdef __cplusplus
}
#endif
