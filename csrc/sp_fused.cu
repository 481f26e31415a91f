// Sequence-parallel "reduce -> row op -> broadcast" kernels: ONE kernel per residual-stream hand-off.
//
// With Megatron sequence parallelism every row-parallel GEMM (wo, down_proj, the dgrads of q|k|v and gate|up) is
// followed by   reduce-scatter -> (residual add + RMSNorm | RMSNorm backward) on this rank's rows -> all-gather
// before the next column-parallel GEMM can start.  As three kernels that is two collectives with four rank barriers,
// the reduce-scatter saturating only the GPU's NVLink egress and the all-gather only its ingress.  Here the rank's rows
// are processed by one kernel, a warp per row:
//     multimem.ld_reduce  (the NVSwitch returns the fp32-accumulated sum of the n partial rows)
//  -> the row op in registers
//  -> multimem.st         (the NVSwitch writes the bf16 result row into every rank's landing zone)
// so the pull of row i+1 and the push of row i use both link directions at once, the intermediate [M/n, D] tensors
// never touch HBM, and there are two barriers (partials complete / results landed) instead of four.
//
// The reference has no sequence parallelism; the call sites these serve are its Reduce after the row-parallel GEMM
// (models/comm_ops.py:39 <- models/layers.py:51) + residual add + next RMSNorm (models/model.py:119-120), and the
// Copy.backward all-reduce (models/comm_ops.py:59 <- models/layers.py:91) + RMSNorm backward.
#include "comm.cuh"
#include "dpfs_b200.h"
#include "sp_rowops.cuh"

#include <stdlib.h>

namespace dpfs {

int num_sms();

// ---- rank barrier of a whole grid ---------------------------------------------------------------------------------
// sync_ranks (comm.cuh) pairs block b of every rank: n remote atomics per block, ~1000 per barrier at 125 blocks.
// Here the blocks of ONE grid first meet on a local counter; the last one to arrive does a single n-way exchange with
// its peers and then releases the others -- 8 remote atomics per rank per barrier.  Needs every block of the grid
// resident at once (sp_grid caps the grid at the occupancy limit).  bar_mode: 0 = per-block barriers, 1 = this one,
// 2 = none (timing experiments only: the result is garbage).
__device__ unsigned int g_grid_count = 0;
__device__ unsigned int g_grid_gen = 0;

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// ALLFENCE = false: one system-scope fence per block, issued by thread 0 after the block barrier (the pattern of
// cooperative-groups grid sync: the barrier orders every thread's stores before thread 0's cumulative fence).
// ALLFENCE = true: every thread fences its own stores first (kept for A/B timing; 64 K system fences per barrier).
template <bool ALLFENCE>
__device__ __forceinline__ void sync_ranks_grid(const CommCtx& c) {
  __shared__ int s_last;
  if (ALLFENCE) __threadfence_system();
  __syncthreads();
  unsigned int gen = 0;
  if (threadIdx.x == 0) {
    if (!ALLFENCE) __threadfence_system();  // this block's stores (plain and multimem) are performed system-wide
    gen = ld_acquire_gpu(&g_grid_gen);      // read before arriving: the release of THIS barrier flips it
    unsigned int prev;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(prev) : "l"(&g_grid_count) : "memory");
    s_last = prev == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x < c.n) {
      const int peer = threadIdx.x;
      put_signal(c.flags[peer] + c.rank);   // slot 0 of the block-barrier region
      wait_signal(c.flags[c.rank] + peer);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      g_grid_count = 0;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&g_grid_gen), "r"(gen + 1) : "memory");
    }
  } else if (threadIdx.x == 0) {
    const uint64_t t0 = globaltimer_ns();
    while (ld_acquire_gpu(&g_grid_gen) == gen) {
      if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void sp_barrier(const CommCtx& c, int bar_mode) {
  if (bar_mode == 0) sync_ranks(c, blockIdx.x);
  else if (bar_mode == 1) sync_ranks_grid<false>(c);
  else if (bar_mode == 3) sync_ranks_grid<true>(c);
}

// Block shape: T threads = T / 32 warps, each owns whole rows (512 threads for D <= 512, 256 for the wider rows, whose
// per-lane accumulators need more than 128 registers).

// A lane owns the 16-byte chunks (8 bf16 columns) j * 32 + lane, j < NC, of a row: D <= NC * 256.
// R rows per warp are in flight together (the loads of all R rows are issued before the first use).

// ------------------------------------------------------------------------------------------------ forward
// x_out = x + sum_r partial_r + bias   (fp32, this rank's rows);   h = rmsnorm(x_out) * scale -> every rank (bf16).
//   mc_part : multicast alias of the partial matrix [n * rpr][D] bf16 (row r0 + i of it is this rank's row i)
//   x_l     : fp32 [rpr][D] residual or nullptr (treated as zero: the embedding hand-off)
//   mc_h    : multicast alias of the landing zone [n * rpr][D] bf16, or nullptr (no norm wanted)
template <int NC, int R, int T>
__global__ void __launch_bounds__(T) rs_add_rmsnorm_ag_kernel(
    CommCtx c, const uint8_t* __restrict__ mc_part, const float* __restrict__ x_l, const float* __restrict__ bias,
    const float* __restrict__ scale, float* __restrict__ xout_l, float* __restrict__ rstd_l, uint8_t* mc_h, int rpr,
    int D, float eps, int bar_mode) {
  sp_barrier(c, bar_mode);  // every rank's partial matrix is complete; every rank is done with the landing zone
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int wpb = T / 32;
  const int nchunk = D >> 3;
  float sc[NC][8], bs[NC][8];
  sp_load_cols<NC>(scale, nchunk, lane, sc);
  sp_load_cols<NC>(bias, nchunk, lane, bs);
  const long long r0 = static_cast<long long>(c.rank) * rpr;
  const long long rowb = static_cast<long long>(D) * 2;
  for (int row = (blockIdx.x * wpb + warp) * R; row < rpr; row += gridDim.x * wpb * R)
    sp_fwd_rows<NC, R>(mc_part + (r0 + row) * rowb, mc_h ? mc_h + (r0 + row) * rowb : nullptr, x_l, xout_l, rstd_l, row,
                       rpr, D, eps, lane, sc, bs);
  sp_barrier(c, bar_mode);  // every rank's h rows have landed everywhere; nobody still reads the partials
}

// ------------------------------------------------------------------------------------------------ backward
// dh = sum_r partial_r (this rank's rows);  dx = dres + rstd * (g - xhat * mean(g * xhat)),  g = dh * scale,
// xhat = x * rstd;  dx -> fp32 private rows (optional) and bf16 to every rank (optional).
// Per-block partial sums over this rank's rows go to scratch[blockIdx.x][2 * D]: [0, D) = dscale = sum dh * xhat,
// [D, 2D) = column sums of dx (the bias gradient of the row-parallel linear whose output gradient dx is).
template <int NC, int R, int T>
__global__ void __launch_bounds__(T) rs_rmsnorm_bwd_ag_kernel(
    CommCtx c, const uint8_t* __restrict__ mc_part, const float* __restrict__ x_l, const float* __restrict__ rstd_l,
    const float* __restrict__ scale, const float* __restrict__ dres_l, float* __restrict__ dx_l, uint8_t* mc_dx,
    float* __restrict__ scratch, int rpr, int D, int bar_mode) {
  __shared__ float red[T / 32][32 * 8];
  sp_barrier(c, bar_mode);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int wpb = T / 32;
  const int nchunk = D >> 3;
  float sc[NC][8], ds[NC][8], cs[NC][8];
  sp_load_cols<NC>(scale, nchunk, lane, sc);
  sp_load_cols<NC>(nullptr, nchunk, lane, ds);
  sp_load_cols<NC>(nullptr, nchunk, lane, cs);
  const long long r0 = static_cast<long long>(c.rank) * rpr;
  const long long rowb = static_cast<long long>(D) * 2;
  for (int row = (blockIdx.x * wpb + warp) * R; row < rpr; row += gridDim.x * wpb * R)
    sp_bwd_rows<NC, R>(mc_part + (r0 + row) * rowb, mc_dx ? mc_dx + (r0 + row) * rowb : nullptr, x_l, rstd_l, dres_l,
                       dx_l, row, rpr, D, lane, sc, ds, cs);
  // block-level reduction of the two partial-sum rows: warps -> smem -> warp 0 sums in a fixed order
  sp_bwd_block_reduce<NC, wpb>(red, scratch + static_cast<long long>(blockIdx.x) * 2 * D, D, warp, lane, ds, cs,
                               [] { __syncthreads(); });
  sp_barrier(c, bar_mode);
}

// out0[c] (+)= sum_b scratch[b][c], out1[c] (+)= sum_b scratch[b][D + c]: fixed order, deterministic.
__global__ void __launch_bounds__(256) sp_reduce_partials_kernel(const float* __restrict__ scratch, int nblocks, int D,
                                                                 float* __restrict__ out0, float* __restrict__ out1,
                                                                 int accumulate) {
  __shared__ float sm[8][33];
  const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;  // column of the [2 * D] partial row
  float s = 0.f;
  if (c < 2 * D)
    for (int b = grp; b < nblocks; b += 8) s += scratch[static_cast<long long>(b) * 2 * D + c];
  sm[grp][lane] = s;
  __syncthreads();
  if (grp == 0 && c < 2 * D) {
    float t = 0.f;
#pragma unroll
    for (int g = 0; g < 8; ++g) t += sm[g][lane];
    float* o = c < D ? (out0 ? out0 + c : nullptr) : (out1 ? out1 + (c - D) : nullptr);
    if (o) *o = accumulate ? *o + t : t;
  }
}

static CommCtx sp_ctx(const unsigned long long* flag_ptrs, int rank, int n) {
  CommCtx c;
  for (int i = 0; i < kMaxRanks; ++i) c.flags[i] = i < n ? reinterpret_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
  c.rank = rank;
  c.n = n;
  return c;
}

// Debug / tuning knobs (tools/bench_handoff.py): grid cap (0 = automatic) and barrier flavour.
static int g_sp_grid_cap = 0;
static int g_sp_bar_mode = 1;
static void sp_env_init() {   // DPFS_SP_GRID_CAP / DPFS_SP_BAR_MODE, read once per process
  static const bool done = [] {
    if (const char* e = getenv("DPFS_SP_GRID_CAP")) g_sp_grid_cap = atoi(e);
    if (const char* e = getenv("DPFS_SP_BAR_MODE")) g_sp_bar_mode = atoi(e);
    return true;
  }();
  (void)done;
}

// Grid: about eight passes of rows per block.  Measured on 8 x B200 (profiles/handoff_microbench_tp8_r2.md): with
// [32000, 512] the kernel takes 118 us with one pass on 125 blocks, 101 us on 37 blocks and 95 us on 16 -- the data
// phase is bound by the NVLS pull / push rate whatever the grid, and every extra block adds to the grid-wide barrier
// (and takes an SM away from the weight-gradient GEMMs that run beside it).  Never more blocks than can be resident
// together (`resident` = SMs x blocks per SM from the occupancy calculator): the grid barrier spins until all arrived.
static int sp_grid(int rpr, int R, int T, int resident) {
  const int rows_per_block = (T / 32) * R;
  int b = (rpr + 8 * rows_per_block - 1) / (8 * rows_per_block);
  if (b < 16) b = 16;
  const int one_pass = (rpr + rows_per_block - 1) / rows_per_block;
  if (b > one_pass) b = one_pass;
  int cap = resident < kMaxBlocks ? resident : kMaxBlocks;
  if (g_sp_grid_cap > 0 && g_sp_grid_cap < cap) cap = g_sp_grid_cap;
  if (b > cap) b = cap;
  return b < 1 ? 1 : b;
}

template <typename K>
static int resident_blocks(K kfn, int threads) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  return per_sm * num_sms();
}

}  // namespace dpfs

using namespace dpfs;

// upper bound of the backward kernel's grid (sizes its [blocks][2 * D] partial-sum scratch)
extern "C" int dpfs_sp_blocks(int rows_per_rank, int D) {
  const int rows_per_block = D <= 512 ? 32 : 8;
  const int b = (rows_per_rank + rows_per_block - 1) / rows_per_block;
  return b < kMaxBlocks ? (b < 1 ? 1 : b) : kMaxBlocks;
}

// tuning hook for the micro-benchmarks: grid_cap (0 = automatic), bar_mode (0 per-block, 1 grid-wide, 2 none)
extern "C" int dpfs_sp_debug_config(int grid_cap, int bar_mode) {
  sp_env_init();
  g_sp_grid_cap = grid_cap;
  g_sp_bar_mode = bar_mode;
  return 0;
}

extern "C" int dpfs_rs_add_rmsnorm_ag(const void* mc_partial, const float* x_l, const float* bias, const float* scale,
                                      float* xout_l, float* rstd_l, void* mc_h, int rows_per_rank, int D, float eps,
                                      const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (D % 8 != 0 || D > 1024 || rows_per_rank <= 0) return -60;
  if (!mc_partial || !xout_l || (mc_h && (!scale || !rstd_l))) return -61;
  if ((reinterpret_cast<uintptr_t>(mc_partial) & 15) || (reinterpret_cast<uintptr_t>(mc_h) & 15)) return -51;
  sp_env_init();
  CommCtx c = sp_ctx(flag_ptrs, rank, n);
  const uint8_t* part = static_cast<const uint8_t*>(mc_partial);
  uint8_t* h = static_cast<uint8_t*>(mc_h);
#define DPFS_SP_FWD(NC, R, T)                                                                               \
  {                                                                                                         \
    auto kfn = rs_add_rmsnorm_ag_kernel<NC, R, T>;                                                          \
    static const int resident = resident_blocks(kfn, T);                                                    \
    kfn<<<(DPFS_COUNT(), sp_grid(rows_per_rank, R, T, resident)), T, 0, stream>>>(                          \
        c, part, x_l, bias, scale, xout_l, rstd_l, h, rows_per_rank, D, eps, g_sp_bar_mode);                \
  }
  if (D <= 256) DPFS_SP_FWD(1, 2, 512)
  else if (D <= 512) DPFS_SP_FWD(2, 2, 512)
  else DPFS_SP_FWD(4, 1, 256)
#undef DPFS_SP_FWD
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_rs_rmsnorm_bwd_ag(const void* mc_partial, const float* x_l, const float* rstd_l, const float* scale,
                                      const float* dres_l, float* dx_l, void* mc_dx, float* dscale_out,
                                      float* dbias_out, int accumulate, float* scratch, int rows_per_rank, int D,
                                      const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (D % 8 != 0 || D > 1024 || rows_per_rank <= 0) return -60;
  if (!mc_partial || !x_l || !rstd_l || !scale || !scratch) return -61;
  if ((reinterpret_cast<uintptr_t>(mc_partial) & 15) || (reinterpret_cast<uintptr_t>(mc_dx) & 15)) return -51;
  sp_env_init();
  CommCtx c = sp_ctx(flag_ptrs, rank, n);
  const uint8_t* part = static_cast<const uint8_t*>(mc_partial);
  uint8_t* dx = static_cast<uint8_t*>(mc_dx);
  int grid = 0;
#define DPFS_SP_BWD(NC, R, T)                                                                      \
  {                                                                                                \
    auto kfn = rs_rmsnorm_bwd_ag_kernel<NC, R, T>;                                                 \
    static const int resident = resident_blocks(kfn, T);                                           \
    grid = sp_grid(rows_per_rank, R, T, resident);                                                 \
    kfn<<<(DPFS_COUNT(), grid), T, 0, stream>>>(c, part, x_l, rstd_l, scale, dres_l, dx_l, dx,     \
                                                scratch, rows_per_rank, D, g_sp_bar_mode);         \
  }
  if (D <= 256) { DPFS_SP_BWD(1, 2, 512); }
  else if (D <= 512) { DPFS_SP_BWD(2, 2, 512); }
  else { DPFS_SP_BWD(4, 1, 256); }
#undef DPFS_SP_BWD
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  sp_reduce_partials_kernel<<<(DPFS_COUNT(), (2 * D + 31) / 32), 256, 0, stream>>>(scratch, grid, D, dscale_out,
                                                                                    dbias_out, accumulate);
  return static_cast<int>(cudaGetLastError());
}
