// Row-parallel GEMM -> reduce-scatter -> row op -> all-gather in ONE kernel (sequence-parallel hand-off, fused form).
//
//   partial_r = A_r W_r^T            tcgen05 GEMM (TMA -> smem ring -> tcgen05.mma -> TMEM -> tcgen05.ld epilogue), bf16
//                                    tiles into this rank's symmetric buffer, one "tile ready" flag per tile to the
//                                    rank(s) that own its rows
//   rows of rank r:                  comm warps of the SAME CTAs wait until a 16-row slice's tiles are flagged by all n
//       y   = sum_r partial_r        ranks, pull the switch-reduced rows (multimem.ld_reduce),
//       out = rowop(y, ...)          run residual + bias + RMSNorm (forward) / RMSNorm backward in registers and
//       out -> every rank            push the bf16 result rows into every rank's landing zone (multimem.st)
//
// so the reduction of the first row-blocks runs under the GEMM of the later ones, there is no opening rank barrier (the
// tile flags of all ranks gate every row) and the tile flags are plain release stores: they carry a launch EPOCH instead
// of being consumed, so several CTAs may wait on the same tile and nothing has to be reset.  The epoch is a per-GPU
// device counter advanced by the kernel itself (CUDA-graph replays re-use the same launch arguments); all ranks launch
// the same sequence of these kernels, so their counters agree.  One closing rank barrier (grid-wide, one n-way exchange)
// makes the pushed rows visible everywhere before the consumer GEMM starts.
//
// Serves the same call sites as sp_fused.cu plus the GEMM in front of them: RowParallelLinear.forward + Reduce + residual
// + next RMSNorm (ref models/layers.py:49-54, models/comm_ops.py:39, models/model.py:119-120) and the column-parallel
// dgrad + Copy.backward all-reduce + RMSNorm backward (models/comm_ops.py:59 <- models/layers.py:91).
#include "comm.cuh"
#include "dpfs_b200.h"
#include "sp_rowops.cuh"

#include <cudaTypedefs.h>
#include <stdlib.h>

namespace dpfs {

int make_tmap_bf16_2d(CUtensorMap* tm, const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                      uint32_t box_inner, uint32_t box_outer);
int num_sms();

namespace gh {

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int kStages = 4;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kGemmThreads = 192;  // warp 0 TMA, warp 1 MMA + TMEM, warps 2-5 epilogue
constexpr int kCommWarps = 8;      // warps 6-13: reduce / row op / broadcast
constexpr int kThreads = kGemmThreads + kCommWarps * 32;
constexpr int kRedBytes = kCommWarps * 32 * 8 * 4;  // backward: block-level partial-sum reduction
constexpr int kSmem = kStages * kStageBytes + 256 + kRedBytes + 1024;
static_assert(kSmem <= 232448, "exceeds the 227 KB shared-memory limit");
constexpr int kEpochBase = (kMaxBlocks + kMaxTiles) * kMaxRanks;  // third region of the flag buffer

__device__ unsigned int g_epoch = 0;       // launches of this kernel family completed on this GPU
__device__ unsigned int g_bar_count = 0;   // grid barrier: arrivals
__device__ unsigned int g_bar_gen = 0;     //               generation

struct Args {
  int M, N, K;                 // N = D, the model width: a comm warp owns whole rows
  __nv_bfloat16* out;          // partial products [M][N] bf16, this rank's symmetric buffer
  const uint8_t* mc_part;      // multicast alias of `out`
  CommCtx comm;
  int rpr;                     // rows per rank (M / n)
  float eps;
  // forward row op
  const float* x_l;            // fp32 [rpr][N] residual (nullable)
  const float* bias;           // fp32 [N] (nullable)
  const float* scale;          // fp32 [N]
  float* xout_l;               // fp32 [rpr][N]
  float* rstd_l;               // fp32 [rpr] (forward: written, backward: read)
  uint8_t* mc_dst;             // multicast alias of the landing zone [M][N] bf16 (h / dx)
  // backward row op
  const float* dres_l;         // fp32 [rpr][N] (nullable)
  float* dx_l;                 // fp32 [rpr][N] (nullable)
  float* scratch;              // fp32 [gridDim.x][2 * N] partial sums
  int comm_ctas;               // CTAs whose comm warps work (the others only run the GEMM), see launch()
};

// same owner-interleaved tile order as the fused reduce-scatter (gemm_tcgen05.cu): the j-th row-block of owner 0, of
// owner 1, ... so that every owner's comm warps are fed from the start
__host__ __device__ __forceinline__ int first_blk(int o, int rpr) { return (o * rpr + BM - 1) / BM; }
__host__ __device__ __forceinline__ int block_of(int wb, int n, int rpr) {
  int cmin = 0x7fffffff;
  for (int o = 0; o < n; ++o) {
    const int cnt = first_blk(o + 1, rpr) - first_blk(o, rpr);
    cmin = cnt < cmin ? cnt : cmin;
  }
  if (wb < cmin * n) return first_blk(wb % n, rpr) + wb / n;
  int idx = wb - cmin * n;
  for (int o = 0; o < n; ++o) {
    if (first_blk(o + 1, rpr) - first_blk(o, rpr) > cmin) {
      if (idx == 0) return first_blk(o, rpr) + cmin;
      --idx;
    }
  }
  return 0;
}

__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// OP = 0: residual + bias + RMSNorm (forward), OP = 1: RMSNorm backward.  NC: 16-byte chunks per lane (N <= NC * 256).
template <bool B_MN, int OP, int NC>
__global__ void __launch_bounds__(kThreads, 1)
gemm_handoff_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Args args) {
  constexpr int R = NC <= 2 ? 2 : 1;           // rows per comm warp per item
  constexpr int IR = kCommWarps * R;           // rows per comm work item
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  int* s_last = reinterpret_cast<int*>(tmem_slot + 1);
  float(*red)[32 * 8] = reinterpret_cast<float(*)[32 * 8]>(smem + kStages * kStageBytes + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = args.comm.n, rank = args.comm.rank, rpr = args.rpr;
  const int tiles_m = (args.M + BM - 1) / BM;
  const int tiles_n = (args.N + BN - 1) / BN;
  const int num_work = tiles_m * tiles_n;
  const int kb_total = (args.K + BK - 1) / BK;
  const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(&g_epoch) + 1u;
  auto tile_of = [&](int w) { return block_of(w / tiles_n, n, rpr) * tiles_n + w % tiles_n; };

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const int t = tile_of(w);
        const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN;
        for (int kb = 0; kb < kb_total; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], kStageBytes);
          uint8_t* a_dst = sA + stage * kABytes;
          uint8_t* b_dst = sB + stage * kBBytes;
          tma_load_2d(a_dst, &tmA, &full_bar[stage], kb * BK, m0);
          if constexpr (!B_MN) {
            tma_load_2d(b_dst, &tmB, &full_bar[stage], kb * BK, n0);
          } else {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b)
              tma_load_2d(b_dst + b * (BK * 128), &tmB, &full_bar[stage], n0 + 64 * b, kb * BK);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, false, B_MN);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < kb_total; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * kABytes);
          const uint32_t b_base = smem_u32(sB + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(a_base + k * 32, 0, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_base + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(b_base + k * 32, 0, 1024);
            umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[as]);
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp < 6) {
    // ------------------------------------------------------------------ epilogue: TMEM -> bf16 partial tile -> flag
    const int q = warp & 3;
    int as = 0;
    uint32_t aphase = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
      const int t = tile_of(w);
      const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN;
      const int row = m0 + q * 32 + lane;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + as * BN + (static_cast<uint32_t>(q * 32) << 16);
      __nv_bfloat16* orow = args.out + static_cast<long long>(row) * args.N + n0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(t_addr + c * 32, r);
        tmem_ld_wait();
        const int ncols_valid = args.N - (n0 + c * 32);
        if (row < args.M && ncols_valid > 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j * 8 + 8 <= ncols_valid) {
              uint4 o;
              o.x = pack_bf16(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
              o.y = pack_bf16(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
              o.z = pack_bf16(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
              o.w = pack_bf16(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
              *reinterpret_cast<uint4*>(orow + c * 32 + 8 * j) = o;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
      // all 128 epilogue threads have stored their rows of the partial tile -> tell the rank(s) that own these rows
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 64) {
        __threadfence_system();
        const int o_lo = m0 / rpr;
        const int o_hi = (min(m0 + BM, args.M) - 1) / rpr;
        st_release_sys(args.comm.flags[o_lo] + kEpochBase + t * kMaxRanks + rank, epoch);
        if (o_hi != o_lo) st_release_sys(args.comm.flags[o_hi] + kEpochBase + t * kMaxRanks + rank, epoch);
      }
    }
  } else {
    // ------------------------------------------------------------------ comm warps: this rank's rows, IR at a time
    const int ct = threadIdx.x - kGemmThreads;
    const int cw = ct >> 5;
    const int nchunk = args.N >> 3;
    const long long rowb = static_cast<long long>(args.N) * 2;
    const long long g_row0 = static_cast<long long>(rank) * rpr;
    float sc[NC][8], c1[NC][8], c2[NC][8];   // forward: scale, bias, -   backward: scale, dscale sums, dx column sums
    sp_load_cols<NC>(args.scale, nchunk, lane, sc);
    sp_load_cols<NC>(OP == 0 ? args.bias : nullptr, nchunk, lane, c1);
    sp_load_cols<NC>(nullptr, nchunk, lane, c2);
    const int n_items = (rpr + IR - 1) / IR;
    const int comm_ctas = min(args.comm_ctas, static_cast<int>(gridDim.x));
    for (int it = blockIdx.x < comm_ctas ? blockIdx.x : n_items; it < n_items; it += comm_ctas) {
      const int r_lo = it * IR, r_hi = min(rpr, r_lo + IR);                 // local rows of this item
      const int mb_lo = static_cast<int>((g_row0 + r_lo) / BM), mb_hi = static_cast<int>((g_row0 + r_hi - 1) / BM);
      const int per_blk = tiles_n * n;
      const int cnt = (mb_hi - mb_lo + 1) * per_blk;                        // <= 2 * 4 * 8 tile flags
      if (ct < cnt) {
        const int mb = mb_lo + ct / per_blk, rem = ct % per_blk;
        const unsigned int* f =
            args.comm.flags[rank] + kEpochBase + (mb * tiles_n + rem / n) * kMaxRanks + rem % n;
        const uint64_t t0 = globaltimer_ns();
        while (static_cast<int>(ld_acquire_sys(f) - epoch) < 0) {
          if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
        }
      }
      asm volatile("bar.sync 2, 256;" ::: "memory");
      const int row = r_lo + cw * R;
      if (row < r_hi) {
        const uint8_t* src = args.mc_part + (g_row0 + row) * rowb;
        uint8_t* dst = args.mc_dst ? args.mc_dst + (g_row0 + row) * rowb : nullptr;
        if constexpr (OP == 0)
          sp_fwd_rows<NC, R>(src, dst, args.x_l, args.xout_l, args.rstd_l, row, r_hi, args.N, args.eps, lane, sc, c1);
        else
          sp_bwd_rows<NC, R>(src, dst, args.x_l, args.rstd_l, args.dres_l, args.dx_l, row, r_hi, args.N, lane, sc, c1,
                             c2);
      }
    }
    if constexpr (OP == 1)
      sp_bwd_block_reduce<NC, kCommWarps>(red, args.scratch + static_cast<long long>(blockIdx.x) * 2 * args.N, args.N,
                                          cw, lane, c1, c2, [] { asm volatile("bar.sync 2, 256;" ::: "memory"); });
  }

  // ---- closing rank barrier: every rank's pushed rows have landed everywhere, nobody still reads the partials ----
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int gen = ld_acquire_gpu(&g_bar_gen);
    unsigned int prev;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(prev) : "l"(&g_bar_count) : "memory");
    *s_last = prev == gridDim.x - 1;
    if (!*s_last) {
      const uint64_t t0 = globaltimer_ns();
      while (ld_acquire_gpu(&g_bar_gen) == gen) {
        if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
      }
    }
  }
  __syncthreads();
  if (*s_last) {   // the last block of this rank to arrive: one n-way exchange, then release the others
    if (threadIdx.x < n) {
      const int peer = threadIdx.x;
      put_signal(args.comm.flags[peer] + rank);
      wait_signal(args.comm.flags[rank] + peer);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      g_bar_count = 0;
      g_epoch = epoch;   // this launch is complete: the next one (same stream) waits on epoch + 1
      __threadfence();
      const unsigned int gen = *reinterpret_cast<volatile unsigned int*>(&g_bar_gen);
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&g_bar_gen), "r"(gen + 1) : "memory");
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// out0[c] (+)= sum_b scratch[b][c], out1[c] (+)= sum_b scratch[b][D + c]  (fixed order)
__global__ void __launch_bounds__(256) reduce_partials2_kernel(const float* __restrict__ scratch, int nblocks, int D,
                                                               float* __restrict__ out0, float* __restrict__ out1) {
  __shared__ float sm[8][33];
  const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
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
    if (o) *o = t;
  }
}

// How many CTAs lend their comm warps to the pull / row op / push.  The NVLS data phase runs at the same rate with 16
// blocks of pulls in flight as with 148 (profiles/handoff_microbench_tp8_r2.md) and every extra puller only adds
// contention, so the rows are dealt to the first `comm_ctas` CTAs: about eight 16-row items each, at least 16 CTAs
// (DPFS_GH_COMM_CTAS overrides).
static int comm_ctas_for(int rpr, int rows_per_item) {
  static const int env = [] {
    const char* e = getenv("DPFS_GH_COMM_CTAS");
    return e ? atoi(e) : 0;
  }();
  if (env > 0) return env;
  const int items = (rpr + rows_per_item - 1) / rows_per_item;
  int c = (items + 7) / 8;
  return c < 16 ? 16 : c;
}

template <bool B_MN, int OP, int NC>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const Args& args_in, int grid, cudaStream_t stream) {
  Args args = args_in;
  args.comm_ctas = comm_ctas_for(args.rpr, kCommWarps * (NC <= 2 ? 2 : 1));
  auto kfn = gemm_handoff_kernel<B_MN, OP, NC>;
  DPFS_ENSURE_DYN_SMEM(kfn, kSmem);
  kfn<<<(DPFS_COUNT(), grid), kThreads, kSmem, stream>>>(tmA, tmB, args);
  return static_cast<int>(cudaGetLastError());
}

template <int OP>
static int dispatch(bool b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB, const Args& args, int grid,
                    cudaStream_t stream) {
  const int nc = (args.N + 255) / 256;
  if (nc <= 1) return b_mn ? launch<true, OP, 1>(tmA, tmB, args, grid, stream) : launch<false, OP, 1>(tmA, tmB, args, grid, stream);
  if (nc <= 2) return b_mn ? launch<true, OP, 2>(tmA, tmB, args, grid, stream) : launch<false, OP, 2>(tmA, tmB, args, grid, stream);
  return b_mn ? launch<true, OP, 4>(tmA, tmB, args, grid, stream) : launch<false, OP, 4>(tmA, tmB, args, grid, stream);
}

static int grid_for(int M, int N) {
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  int g = num_sms();   // one CTA per SM (200 KB of shared memory each): all resident, as the grid barrier requires
  return tiles < g ? tiles : g;
}

static int prepare(const void* A, long long lda, const void* B, long long ldb, int b_mn, void* partial,
                   const void* mc_partial, int M, int N, int K, const unsigned long long* flag_ptrs, int rank, int n,
                   CUtensorMap* tmA, CUtensorMap* tmB, Args* a) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (M <= 0 || K <= 0 || N % 8 != 0 || N > 1024 || M % n != 0 || (lda % 8) != 0 || (ldb % 8) != 0) return -60;
  if (!partial || !mc_partial) return -61;
  if (M / n < BM) return -53;   // a 128-row block may straddle at most two owners
  if (((M + BM - 1) / BM) * ((N + BN - 1) / BN) > kMaxTiles) return -52;
  int rc = make_tmap_bf16_2d(tmA, A, K, M, lda, 64, BM);
  if (rc) return rc;
  rc = !b_mn ? make_tmap_bf16_2d(tmB, B, K, N, ldb, 64, BN) : make_tmap_bf16_2d(tmB, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  *a = Args{};
  a->M = M;
  a->N = N;
  a->K = K;
  a->out = static_cast<__nv_bfloat16*>(partial);
  a->mc_part = static_cast<const uint8_t*>(mc_partial);
  for (int i = 0; i < kMaxRanks; ++i) a->comm.flags[i] = i < n ? reinterpret_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
  a->comm.rank = rank;
  a->comm.n = n;
  a->rpr = M / n;
  return 0;
}

}  // namespace gh
}  // namespace dpfs

using namespace dpfs;

// rows of the backward kernel's partial-sum scratch ([rows][2 * N] floats)
extern "C" int dpfs_gemm_handoff_scratch_rows(void) { return num_sms(); }

extern "C" int dpfs_gemm_handoff_fwd(const void* A, long long lda, const void* B, long long ldb, int b_mn_major,
                                     void* partial, const void* mc_partial, int M, int N, int K, const float* x_l,
                                     const float* bias, const float* scale, float* xout_l, float* rstd_l, void* mc_h,
                                     float eps, const unsigned long long* flag_ptrs, int rank, int n,
                                     cudaStream_t stream) {
  CUtensorMap tmA, tmB;
  gh::Args a;
  int rc = gh::prepare(A, lda, B, ldb, b_mn_major, partial, mc_partial, M, N, K, flag_ptrs, rank, n, &tmA, &tmB, &a);
  if (rc) return rc;
  if (!xout_l || (mc_h && (!scale || !rstd_l))) return -61;
  a.eps = eps;
  a.x_l = x_l;
  a.bias = bias;
  a.scale = scale;
  a.xout_l = xout_l;
  a.rstd_l = rstd_l;
  a.mc_dst = static_cast<uint8_t*>(mc_h);
  return gh::dispatch<0>(b_mn_major != 0, tmA, tmB, a, gh::grid_for(M, N), stream);
}

extern "C" int dpfs_gemm_handoff_bwd(const void* A, long long lda, const void* B, long long ldb, int b_mn_major,
                                     void* partial, const void* mc_partial, int M, int N, int K, const float* x_l,
                                     const float* rstd_l, const float* scale, const float* dres_l, float* dx_l,
                                     void* mc_dx, float* dscale_out, float* dbias_out, float* scratch,
                                     const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  CUtensorMap tmA, tmB;
  gh::Args a;
  int rc = gh::prepare(A, lda, B, ldb, b_mn_major, partial, mc_partial, M, N, K, flag_ptrs, rank, n, &tmA, &tmB, &a);
  if (rc) return rc;
  if (!x_l || !rstd_l || !scale || !scratch) return -61;
  a.x_l = x_l;
  a.rstd_l = const_cast<float*>(rstd_l);
  a.scale = scale;
  a.dres_l = dres_l;
  a.dx_l = dx_l;
  a.mc_dst = static_cast<uint8_t*>(mc_dx);
  a.scratch = scratch;
  const int grid = gh::grid_for(M, N);
  rc = gh::dispatch<1>(b_mn_major != 0, tmA, tmB, a, grid, stream);
  if (rc) return rc;
  gh::reduce_partials2_kernel<<<(DPFS_COUNT(), (2 * N + 31) / 32), 256, 0, stream>>>(scratch, grid, N, dscale_out,
                                                                                     dbias_out);
  return static_cast<int>(cudaGetLastError());
}

// Host-side view of the kernel's schedule for the CPU tests (same functions the kernel runs): order[i] = row-block
// computed i-th by every rank; for rank `rank`, item k (16 rows for N <= 512, 8 above) covers local rows
// [item_lo[k], item_hi[k]) and waits for the tiles of row-blocks [blk_lo[k], blk_hi[k]].  Returns the item count.
extern "C" int dpfs_debug_gh_schedule(int M, int N, int n, int rank, int* order, int* item_lo, int* item_hi,
                                      int* blk_lo, int* blk_hi, int cap) {
  using namespace dpfs::gh;
  if (n < 1 || M <= 0 || M % n != 0 || M / n < BM || rank < 0 || rank >= n) return -53;
  const int rpr = M / n, tiles_m = (M + BM - 1) / BM;
  for (int i = 0; i < tiles_m; ++i) order[i] = block_of(i, n, rpr);
  const int IR = kCommWarps * ((N + 255) / 256 <= 2 ? 2 : 1);
  const int n_items = (rpr + IR - 1) / IR;
  const long long g0 = static_cast<long long>(rank) * rpr;
  for (int k = 0; k < n_items && k < cap; ++k) {
    item_lo[k] = k * IR;
    item_hi[k] = (k + 1) * IR < rpr ? (k + 1) * IR : rpr;
    blk_lo[k] = static_cast<int>((g0 + item_lo[k]) / BM);
    blk_hi[k] = static_cast<int>((g0 + item_hi[k] - 1) / BM);
  }
  return n_items;
}
