// Persistent, warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma with fp32
// accumulators in TMEM (double buffered) -> tcgen05.ld epilogue (bias / residual / dtype cast) -> global.
//
//   D[M,N] = A[M,K] * B[N,K]^T  (+ bias[N]) (+ residual[M,N])
//
// Both operands may be K-major (reduction dim contiguous) or MN-major (output dim contiguous), which gives the
// three shapes a linear layer needs without ever transposing in memory:
//   forward  y  = x  W^T : A = x  (K-major),  B = W  (K-major)
//   dgrad    dx = dy W   : A = dy (K-major),  B = W  (MN-major: W is [N_red, K_out] row-major)
//   wgrad    dW = dy^T x : A = dy (MN-major), B = x  (MN-major), reduction over tokens, split-K over the SMs
//
// Replaces the cuBLAS calls behind F.linear in the reference (models/layers.py:49,93 and their autograd).
#include "comm.cuh"
#include "dpfs_b200.h"

#include <cudaTypedefs.h>
#include <stdio.h>
#include <stdlib.h>

namespace dpfs {

unsigned long long g_launches = 0;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kGemmThreads = 192;  // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue
constexpr int kCommThreads = 256;  // FUSE_AR only: warps 6-13 reduce finished tiles over NVLink while the GEMM runs

struct GemmArgs {
  int M, N, K;
  int splits;        // split-K factor
  int kb_per_split;  // K blocks (of BK) per split
  void* out;         // OutT [splits][M][ldo]
  long long ldo;
  long long split_stride;  // elements between split slabs
  const float* bias;       // [N] or nullptr
  const float* residual;   // fp32 [M][ldr] or nullptr
  long long ldr;
  // fused all-reduce (FUSE_AR kernels only): `out` is this rank's symmetric buffer, `mc_out` its multicast alias
  uint8_t* mc_out;   // multicast alias of out (NVLS reduce path) or nullptr (P2P reduce path)
  PeerPtrs peers;    // every rank's `out` (P2P reduce path)
  int debug;         // bring-up only: bit0 skip tile signals, bit1 skip tile waits, bit2 skip the reduce traffic
  CommCtx comm;
  // fused reduce-scatter (FUSE_AR kernels with rs_out != nullptr): rank r keeps rows [r * rs_rows, (r + 1) * rs_rows)
  // of the sum in its private rs_out (bf16 [rs_rows][N]) instead of every rank receiving everything
  __nv_bfloat16* rs_out;
  int rs_rows;
};

// ---- tile order / ownership of the fused reduce-scatter -------------------------------------------------------------
// A row-block (BM rows) has a *primary* owner: the rank whose row range holds the block's first row.  Blocks are
// computed in rounds -- the j-th block of owner 0, of owner 1, ... -- by every rank in the same order, so every
// owner's comm warps get finished tiles at an even rate from the start (with the natural order owner 0 would be served
// first and owner n-1 would reduce its whole share after the GEMM).  A block that straddles two row ranges is flagged
// to both owners.
__host__ __device__ __forceinline__ int rs_first_blk(int o, int rs_rows) { return (o * rs_rows + BM - 1) / BM; }
__host__ __device__ __forceinline__ int rs_block_of(int wb, int n, int rs_rows) {  // wb-th block in compute order
  int cmin = 0x7fffffff;
  for (int o = 0; o < n; ++o) {
    const int cnt = rs_first_blk(o + 1, rs_rows) - rs_first_blk(o, rs_rows);
    cmin = cnt < cmin ? cnt : cmin;
  }
  if (wb < cmin * n) return rs_first_blk(wb % n, rs_rows) + wb / n;
  int idx = wb - cmin * n;  // the owners that have one more block than the others, in rank order
  for (int o = 0; o < n; ++o) {
    if (rs_first_blk(o + 1, rs_rows) - rs_first_blk(o, rs_rows) > cmin) {
      if (idx == 0) return rs_first_blk(o, rs_rows) + cmin;
      --idx;
    }
  }
  return 0;  // unreachable
}

template <int BN>
struct GemmSmem {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kBarrierBytes = 256;
  static constexpr int kTotal = kStages * kStageBytes + kBarrierBytes + 1024;  // + alignment slack
};

template <typename OutT>
__device__ __forceinline__ void store_chunk32(OutT* dst, const float (&v)[32], int ncols_valid);

template <>
__device__ __forceinline__ void store_chunk32<float>(float* dst, const float (&v)[32], int ncols_valid) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (j * 4 + 4 <= ncols_valid) {
      float4 o = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      *reinterpret_cast<float4*>(dst + 4 * j) = o;
    }
  }
}
template <>
__device__ __forceinline__ void store_chunk32<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[32],
                                                             int ncols_valid) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (j * 8 + 8 <= ncols_valid) {
      uint4 o;
      o.x = pack_bf16(v[8 * j + 0], v[8 * j + 1]);
      o.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
      o.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]);
      o.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
      *reinterpret_cast<uint4*>(dst + 8 * j) = o;
    }
  }
}

// FUSE_AR: GEMM -> all-reduce in ONE kernel.  Each CTA writes the bf16 partial tiles it computes into this rank's
// symmetric buffer and raises a per-tile flag on the tile's owner rank (tile row-block m_blk belongs to rank
// m_blk % n).  When its GEMM tiles are done, the CTA turns to the tiles its rank owns: as soon as all n partials of a
// tile are flagged, it pulls their sum through the NVSwitch (multimem.ld_reduce) and broadcasts the result into every
// rank's buffer (multimem.st).  A closing rank-wide barrier makes the full result visible before the kernel exits.
// FUSE_RS (implies FUSE_AR): the reduce-scatter form -- rank r keeps only its rows of the sum (args.rs_out).  A separate
// instantiation so that the all-reduce kernels stay exactly the binaries that were validated on hardware.
template <int BN, bool A_MN, bool B_MN, typename OutT, bool FUSE_AR = false, bool FUSE_RS = false>
__global__ void __launch_bounds__(FUSE_AR ? kGemmThreads + kCommThreads : kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmArgs args) {
  using S = GemmSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + S::kStages * S::kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kStages * S::kStageBytes);
  uint64_t* empty_bar = full_bar + S::kStages;
  uint64_t* tmem_full = empty_bar + S::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int tiles_m = (args.M + BM - 1) / BM;
  const int tiles_n = (args.N + BN - 1) / BN;
  const int tiles_mn = tiles_m * tiles_n;
  const int num_work = tiles_mn * args.splits;
  const int kb_total = (args.K + BK - 1) / BK;
  // work index -> tile id: identity, except for the fused reduce-scatter (see rs_block_of)
  auto tile_of = [&](int wt) -> int {
    if constexpr (FUSE_RS) return rs_block_of(wt / tiles_n, args.comm.n, args.rs_rows) * tiles_n + wt % tiles_n;
    return wt;
  };

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < S::kStages; ++i) {
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
        const int split = w / tiles_mn;
        const int t = tile_of(w - split * tiles_mn);
        const int m0 = (t / tiles_n) * BM;
        const int n0 = (t % tiles_n) * BN;
        const int kb0 = split * args.kb_per_split;
        const int kb1 = min(kb_total, kb0 + args.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], S::kStageBytes);
          uint8_t* a_dst = sA + stage * S::kABytes;
          uint8_t* b_dst = sB + stage * S::kBBytes;
          if constexpr (!A_MN) {
            tma_load_2d(a_dst, &tmA, &full_bar[stage], kb * BK, m0);
          } else {
#pragma unroll
            for (int b = 0; b < BM / 64; ++b)
              tma_load_2d(a_dst + b * (BK * 128), &tmA, &full_bar[stage], m0 + 64 * b, kb * BK);
          }
          if constexpr (!B_MN) {
            tma_load_2d(b_dst, &tmB, &full_bar[stage], kb * BK, n0);
          } else {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b)
              tma_load_2d(b_dst + b * (BK * 128), &tmB, &full_bar[stage], n0 + 64 * b, kb * BK);
          }
          if (++stage == S::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (single thread)
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const int split = w / tiles_mn;
        const int kb0 = split * args.kb_per_split;
        const int kb1 = min(kb_total, kb0 + args.kb_per_split);
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * S::kABytes);
          const uint32_t b_base = smem_u32(sB + stage * S::kBBytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: one 128B swizzle atom spans BK; step 16 elements (32 B) inside the atom; 8-row groups are
            // 1024 B apart.  MN-major: rows are k, 128 B each; 64-wide MN atoms are BK*128 B apart (LBO), 8-row k
            // groups are 1024 B apart (SBO); step 16 k-rows = 2048 B.
            const uint64_t adesc = A_MN ? make_smem_desc_sw128(a_base + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(a_base + k * 32, 0, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_base + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(b_base + k * 32, 0, 1024);
            umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
          if (++stage == S::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[as]);  // accumulator complete -> epilogue
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp < 6) {
    // ------------------------------------------------------------------ epilogue warps (TMEM -> regs -> global)
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    int as = 0;
    uint32_t aphase = 0;
    OutT* out = reinterpret_cast<OutT*>(args.out);
    for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
      const int split = w / tiles_mn;
      const int t = tile_of(w - split * tiles_mn);
      const int m0 = (t / tiles_n) * BM;
      const int n0 = (t % tiles_n) * BN;
      const int row = m0 + q * 32 + lane;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + as * BN + (static_cast<uint32_t>(q * 32) << 16);
      OutT* orow = out + split * args.split_stride + static_cast<long long>(row) * args.ldo + n0;
      const float* rrow = args.residual ? args.residual + static_cast<long long>(row) * args.ldr + n0 : nullptr;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(t_addr + c * 32, r);
        tmem_ld_wait();
        const int ncols_valid = args.N - (n0 + c * 32);
        if (row < args.M && ncols_valid > 0) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          if (args.bias) {  // 8 x LDG.128 (broadcast across the warp) instead of 32 scalar loads
            const float4* b4 = reinterpret_cast<const float4*>(args.bias + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j * 4 + 4 <= ncols_valid) {
                const float4 bv = __ldg(b4 + j);
                v[4 * j] += bv.x;
                v[4 * j + 1] += bv.y;
                v[4 * j + 2] += bv.z;
                v[4 * j + 3] += bv.w;
              }
            }
          }
          if (rrow) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j * 4 + 4 <= ncols_valid) {
                const float4 rv = *reinterpret_cast<const float4*>(rrow + c * 32 + 4 * j);
                v[4 * j] += rv.x;
                v[4 * j + 1] += rv.y;
                v[4 * j + 2] += rv.z;
                v[4 * j + 3] += rv.w;
              }
            }
          }
          store_chunk32<OutT>(orow + c * 32, v, ncols_valid);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
      if constexpr (FUSE_AR) {
        // all 128 epilogue threads have stored their rows of the partial tile -> tell the owner rank
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64 && !(args.debug & 1)) {
          const int m_blk = t / tiles_n;
          __threadfence_system();
          if constexpr (FUSE_RS) {
            const int o_lo = m0 / args.rs_rows;
            const int o_hi = (min(m0 + BM, args.M) - 1) / args.rs_rows;
            put_signal(args.comm.flags[o_lo] + (kMaxBlocks + t) * kMaxRanks + args.comm.rank);
            if (o_hi != o_lo) put_signal(args.comm.flags[o_hi] + (kMaxBlocks + t) * kMaxRanks + args.comm.rank);
          } else {
            const int owner = m_blk % args.comm.n;
            put_signal(args.comm.flags[owner] + (kMaxBlocks + t) * kMaxRanks + args.comm.rank);
          }
        }
      }
    }
  } else {
    if constexpr (FUSE_AR) {
      // ---------------------------------------------------------------- comm warps: reduce the tiles this rank owns
      // Tile row-block m_blk belongs to rank m_blk % n; owned tiles are dealt round-robin to the CTAs.  As soon as the
      // n partial copies of a tile are flagged, sum them over NVLink (fp32, rank order) and write the bf16 result into
      // every rank's buffer — while warps 0-5 of this CTA are still computing other tiles.
      const int ct = threadIdx.x - kGemmThreads;
      const int n = args.comm.n, rank = args.comm.rank;
      if constexpr (FUSE_RS) {
        // ---- reduce-scatter: the blocks that intersect this rank's rows, primary blocks in compute order first, the
        // leading straddler (primary owner rank - 1, computed in that owner's last round) last
        const int rpr = args.rs_rows;
        const int row_lo = rank * rpr, row_hi = row_lo + rpr;
        const int b_first = rs_first_blk(rank, rpr);          // first primary block
        const int b_lo = row_lo / BM, b_hi = (row_hi - 1) / BM;
        const int n_primary = b_hi - b_first + 1;
        const int n_blk_owned = b_hi - b_lo + 1;               // n_primary (+ 1 with a leading straddler)
        uint8_t* dst_base = reinterpret_cast<uint8_t*>(args.rs_out);
        for (int k = blockIdx.x; k < n_blk_owned * tiles_n; k += gridDim.x) {
          const int i = k / tiles_n, n_blk = k % tiles_n;
          const int m_blk = i < n_primary ? b_first + i : b_lo;
          const int t = m_blk * tiles_n + n_blk;
          if (ct < n && !(args.debug & 2)) wait_signal(args.comm.flags[rank] + (kMaxBlocks + t) * kMaxRanks + ct);
          asm volatile("bar.sync 2, 256;" ::: "memory");
          if (args.debug & 4) continue;
          const int m0 = m_blk * BM, n0 = n_blk * BN;
          const int r0 = max(m0, row_lo), r1 = min(min(m0 + BM, row_hi), args.M);
          const int cpr = min(BN, args.N - n0) / 8;  // 16-byte chunks per row
          const long long off0 = (static_cast<long long>(r0) * args.ldo + n0) * 2;
          const long long ldb = args.ldo * 2;
          uint8_t* dst = dst_base + (static_cast<long long>(r0 - row_lo) * args.N + n0) * 2;
          const long long dst_ld = static_cast<long long>(args.N) * 2;
          if (args.mc_out) {
            // args.mc_out is the multicast alias of this rank's `out`: offsets are relative to the buffer start
            reduce_region_to_local_nvls<4>(args.mc_out, off0, ldb, dst, dst_ld, r1 - r0, cpr, ct, kCommThreads);
          } else if (n == 2) {
            reduce_region_to_local_p2p<2, 8>(args.peers, off0, ldb, dst, dst_ld, r1 - r0, cpr, ct, kCommThreads);
          } else if (n == 4) {
            reduce_region_to_local_p2p<4, 8>(args.peers, off0, ldb, dst, dst_ld, r1 - r0, cpr, ct, kCommThreads);
          } else {
            reduce_region_to_local_p2p<8, 8>(args.peers, off0, ldb, dst, dst_ld, r1 - r0, cpr, ct, kCommThreads);
          }
        }
      } else {
      const int owned_rows = (tiles_m - rank + n - 1) / n;  // row-blocks m_blk = rank, rank + n, ...
      const int n_owned = owned_rows * tiles_n;
      for (int k = blockIdx.x; k < n_owned; k += gridDim.x) {
        {
          const int m_blk = rank + n * (k / tiles_n);
          const int n_blk = k % tiles_n;
          const int t = m_blk * tiles_n + n_blk;
          if (ct < n && !(args.debug & 2)) wait_signal(args.comm.flags[rank] + (kMaxBlocks + t) * kMaxRanks + ct);
          asm volatile("bar.sync 2, 256;" ::: "memory");
          if (args.debug & 4) continue;
          const int m0 = m_blk * BM, n0 = n_blk * BN;
          const int rows = min(BM, args.M - m0);
          const int cpr = min(BN, args.N - n0) / 8;  // 16-byte chunks per row
          const long long off0 = (static_cast<long long>(m0) * args.ldo + n0) * 2;
          const long long ldb = args.ldo * 2;
          if (args.mc_out) {   // NVLS data path: the switch sums, the owner re-broadcasts
            reduce_region_nvls<4>(args.mc_out, off0, ldb, rows, cpr, ct, kCommThreads);
          } else if (n == 2) {
            reduce_region_p2p<2, 8>(args.peers, off0, ldb, rows, cpr, ct, kCommThreads);
          } else if (n == 4) {
            reduce_region_p2p<4, 8>(args.peers, off0, ldb, rows, cpr, ct, kCommThreads);
          } else {
            reduce_region_p2p<8, 8>(args.peers, off0, ldb, rows, cpr, ct, kCommThreads);
          }
        }
      }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (FUSE_AR) {
    if (!(args.debug & 8)) sync_ranks(args.comm, blockIdx.x);
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// Sum split-K partial slabs (fp32) and write the final result; optional bias and dtype cast.
template <typename OutT>
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, OutT* __restrict__ out, long long ldo,
                                     const float* __restrict__ bias, int M, int N, int splits) {
  const long long total4 = static_cast<long long>(M) * N / 4;
  const long long slab = static_cast<long long>(M) * N;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 acc = *reinterpret_cast<const float4*>(ws + i * 4);
    for (int s = 1; s < splits; ++s) {
      const float4 p = *reinterpret_cast<const float4*>(ws + s * slab + i * 4);
      acc.x += p.x;
      acc.y += p.y;
      acc.z += p.z;
      acc.w += p.w;
    }
    const long long e = i * 4;
    const int r = static_cast<int>(e / N);
    const int c = static_cast<int>(e % N);
    if (bias) {
      acc.x += bias[c];
      acc.y += bias[c + 1];
      acc.z += bias[c + 2];
      acc.w += bias[c + 3];
    }
    OutT* o = out + static_cast<long long>(r) * ldo + c;
    if constexpr (sizeof(OutT) == 4) {
      *reinterpret_cast<float4*>(o) = acc;
    } else {
      uint2 pk;
      pk.x = pack_bf16(acc.x, acc.y);
      pk.y = pack_bf16(acc.z, acc.w);
      *reinterpret_cast<uint2*>(o) = pk;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    cudaDriverEntryPointQueryResult q;
    void* p = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

// 2-D bf16 tensor map over a row-major [outer][ld] array whose logical inner extent is `inner`;
// box = 64 (inner, 128 B, 128B swizzle) x box_outer.
int make_tmap_bf16_2d(CUtensorMap* tm, const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                      uint32_t box_inner, uint32_t box_outer) {
  auto fn = get_encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <int BN, bool A_MN, bool B_MN, typename OutT>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmArgs& args, cudaStream_t stream) {
  using S = GemmSmem<BN>;
  auto kfn = gemm_bf16_kernel<BN, A_MN, B_MN, OutT>;
  DPFS_ENSURE_DYN_SMEM(kfn, S::kTotal);
  const int tiles = ((args.M + BM - 1) / BM) * ((args.N + BN - 1) / BN) * args.splits;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  kfn<<<(DPFS_COUNT(), grid), kGemmThreads, S::kTotal, stream>>>(tmA, tmB, args);
  return static_cast<int>(cudaGetLastError());
}

template <int BN, bool B_MN, bool RS = false>
static int launch_gemm_ar(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmArgs& args, cudaStream_t stream) {
  using S = GemmSmem<BN>;
  auto kfn = gemm_bf16_kernel<BN, false, B_MN, __nv_bfloat16, true, RS>;
  DPFS_ENSURE_DYN_SMEM(kfn, S::kTotal);
  const int tiles = ((args.M + BM - 1) / BM) * ((args.N + BN - 1) / BN);
  if (tiles > kMaxTiles) return -52;
  int grid = tiles < num_sms() ? tiles : num_sms();
  if (grid > kMaxBlocks) grid = kMaxBlocks;
  kfn<<<(DPFS_COUNT(), grid), kGemmThreads + kCommThreads, S::kTotal, stream>>>(tmA, tmB, args);
  return static_cast<int>(cudaGetLastError());
}

template <int BN, typename OutT>
static int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB,
                          const GemmArgs& args, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm<BN, false, false, OutT>(tmA, tmB, args, stream);
  if (!a_mn && b_mn) return launch_gemm<BN, false, true, OutT>(tmA, tmB, args, stream);
  if (a_mn && b_mn) return launch_gemm<BN, true, true, OutT>(tmA, tmB, args, stream);
  return launch_gemm<BN, true, false, OutT>(tmA, tmB, args, stream);
}

}  // namespace dpfs

using namespace dpfs;

extern "C" int dpfs_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb,
                              int b_mn_major, void* out, long long ldo, int out_fp32, const float* bias,
                              const float* residual, long long ldr, int M, int N, int K, int splits,
                              float* workspace, int bn, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((N % 8) != 0 || (lda % 8) != 0 || (ldb % 8) != 0 || (ldo % (out_fp32 ? 4 : 8)) != 0) return -10;
  if (splits < 1) splits = 1;
  const int kb_total = (K + BK - 1) / BK;
  if (splits > kb_total) splits = kb_total;
  int kb_per_split = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per_split - 1) / kb_per_split;
  if (splits > 1 && workspace == nullptr) return -11;
  if (bn != 256) bn = 128;

  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn_major)
    rc = make_tmap_bf16_2d(&tmA, A, K, M, lda, 64, BM);
  else
    rc = make_tmap_bf16_2d(&tmA, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major)
    rc = make_tmap_bf16_2d(&tmB, B, K, N, ldb, 64, bn);
  else
    rc = make_tmap_bf16_2d(&tmB, B, N, K, ldb, 64, BK);
  if (rc) return rc;

  GemmArgs args;
  args.mc_out = nullptr;
  args.rs_out = nullptr;
  args.rs_rows = 0;
  args.debug = 0;
  for (int i = 0; i < kMaxRanks; ++i) args.peers.p[i] = nullptr;
  args.comm.rank = 0;
  args.comm.n = 1;
  for (int i = 0; i < kMaxRanks; ++i) args.comm.flags[i] = nullptr;
  args.M = M;
  args.N = N;
  args.K = K;
  args.splits = splits;
  args.kb_per_split = kb_per_split;
  args.bias = bias;
  args.residual = residual;
  args.ldr = ldr;
  const bool direct = (splits == 1);
  if (direct) {
    args.out = out;
    args.ldo = ldo;
    args.split_stride = 0;
  } else {
    args.out = workspace;
    args.ldo = N;
    args.split_stride = static_cast<long long>(M) * N;
    args.bias = nullptr;
    args.residual = nullptr;
  }
  const bool kernel_fp32 = direct ? (out_fp32 != 0) : true;
  if (bn == 256) {
    rc = kernel_fp32 ? dispatch_major<256, float>(a_mn_major, b_mn_major, tmA, tmB, args, stream)
                     : dispatch_major<256, __nv_bfloat16>(a_mn_major, b_mn_major, tmA, tmB, args, stream);
  } else {
    rc = kernel_fp32 ? dispatch_major<128, float>(a_mn_major, b_mn_major, tmA, tmB, args, stream)
                     : dispatch_major<128, __nv_bfloat16>(a_mn_major, b_mn_major, tmA, tmB, args, stream);
  }
  if (rc) return rc;
  if (!direct) {
    if (residual) return -12;  // residual is only fused on the direct path
    const long long total4 = static_cast<long long>(M) * N / 4;
    int blocks = static_cast<int>((total4 + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    if (out_fp32)
      splitk_reduce_kernel<float><<<(DPFS_COUNT(), blocks), 256, 0, stream>>>(workspace, reinterpret_cast<float*>(out), ldo, bias,
                                                             M, N, splits);
    else
      splitk_reduce_kernel<__nv_bfloat16><<<(DPFS_COUNT(), blocks), 256, 0, stream>>>(
          workspace, reinterpret_cast<__nv_bfloat16*>(out), ldo, bias, M, N, splits);
    rc = static_cast<int>(cudaGetLastError());
  }
  return rc;
}

extern "C" unsigned long long dpfs_launch_count(void) { return dpfs::g_launches; }

// Fused GEMM -> all-reduce: out (bf16 [M][N], row stride N) must be this rank's symmetric buffer and mc_out the
// multicast alias of the same offset; on return every rank's `out` holds sum_r A_r B_r^T.
static int gemm_fused_collective(const void* A, long long lda, const void* B, long long ldb, int b_mn_major,
                                 void* out, void* mc_out, const unsigned long long* data_ptrs, long long out_offset,
                                 void* rs_out, int M, int N, int K, const unsigned long long* flag_ptrs, int rank,
                                 int n, int bn, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if (n < 2 || n > kMaxRanks) return -50;
  if (n != 2 && n != 4 && n != 8) return -50;  // the in-kernel reduce is templated on the rank count
  if (rs_out && (M % n != 0 || M / n < BM)) return -53;  // a row-block may straddle at most two owners
  if ((N % 8) != 0 || (lda % 8) != 0 || (ldb % 8) != 0) return -10;
  if (bn != 256) bn = 128;
  CUtensorMap tmA, tmB;
  int rc = make_tmap_bf16_2d(&tmA, A, K, M, lda, 64, BM);
  if (rc) return rc;
  if (!b_mn_major)
    rc = make_tmap_bf16_2d(&tmB, B, K, N, ldb, 64, bn);
  else
    rc = make_tmap_bf16_2d(&tmB, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  GemmArgs args;
  args.M = M;
  args.N = N;
  args.K = K;
  args.splits = 1;
  args.kb_per_split = (K + BK - 1) / BK;
  args.out = out;
  args.ldo = N;
  args.split_stride = 0;
  args.bias = nullptr;
  args.residual = nullptr;
  args.ldr = 0;
  args.mc_out = static_cast<uint8_t*>(mc_out);
  args.rs_out = static_cast<__nv_bfloat16*>(rs_out);
  args.rs_rows = rs_out ? M / n : 0;
  {
    static const int fuse_debug = [] {   // read once per process, not once per launch
      const char* e = getenv("DPFS_FUSE_DEBUG");
      return e ? atoi(e) : 0;
    }();
    args.debug = fuse_debug;
  }
  for (int i = 0; i < kMaxRanks; ++i)
    args.peers.p[i] = (i < n && data_ptrs) ? reinterpret_cast<uint8_t*>(data_ptrs[i]) + out_offset : nullptr;
  for (int i = 0; i < kMaxRanks; ++i) args.comm.flags[i] = i < n ? reinterpret_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
  args.comm.rank = rank;
  args.comm.n = n;
  if (rs_out) {
    if (bn == 256)
      return b_mn_major ? launch_gemm_ar<256, true, true>(tmA, tmB, args, stream)
                        : launch_gemm_ar<256, false, true>(tmA, tmB, args, stream);
    return b_mn_major ? launch_gemm_ar<128, true, true>(tmA, tmB, args, stream)
                      : launch_gemm_ar<128, false, true>(tmA, tmB, args, stream);
  }
  if (bn == 256)
    return b_mn_major ? launch_gemm_ar<256, true>(tmA, tmB, args, stream)
                      : launch_gemm_ar<256, false>(tmA, tmB, args, stream);
  return b_mn_major ? launch_gemm_ar<128, true>(tmA, tmB, args, stream)
                    : launch_gemm_ar<128, false>(tmA, tmB, args, stream);
}

extern "C" int dpfs_gemm_allreduce_bf16(const void* A, long long lda, const void* B, long long ldb, int b_mn_major,
                                        void* out, void* mc_out, const unsigned long long* data_ptrs,
                                        long long out_offset, int M, int N, int K,
                                        const unsigned long long* flag_ptrs, int rank, int n, int bn,
                                        cudaStream_t stream) {
  return gemm_fused_collective(A, lda, B, ldb, b_mn_major, out, mc_out, data_ptrs, out_offset, nullptr, M, N, K,
                               flag_ptrs, rank, n, bn, stream);
}

// Fused GEMM -> reduce-scatter over rows (sequence-parallel form of the row-parallel linear): `partial` (bf16 [M][N])
// is this rank's symmetric scratch for the partial products, `rs_out` (bf16 [M / n][N], private memory) receives rows
// [rank * M / n, (rank + 1) * M / n) of sum_r A_r B_r^T.  mc_partial != nullptr selects the NVLS data path
// (multimem.ld_reduce through the switch), otherwise the partials are pulled from the peers (data_ptrs) and summed in
// rank order.  Requires M % n == 0 and M / n >= 128.
extern "C" int dpfs_gemm_reduce_scatter_bf16(const void* A, long long lda, const void* B, long long ldb,
                                             int b_mn_major, void* partial, void* mc_partial,
                                             const unsigned long long* data_ptrs, long long partial_offset,
                                             void* rs_out, int M, int N, int K,
                                             const unsigned long long* flag_ptrs, int rank, int n, int bn,
                                             cudaStream_t stream) {
  if (!rs_out) return -54;
  return gemm_fused_collective(A, lda, B, ldb, b_mn_major, partial, mc_partial, data_ptrs, partial_offset, rs_out, M,
                               N, K, flag_ptrs, rank, n, bn, stream);
}

// Host-side view of the fused reduce-scatter's schedule, for the CPU tests: order[i] = row-block computed i-th by every
// rank (i < ceil(M / 128)); owner_lo / owner_hi[b] = the rank(s) whose comm warps are signalled when block b is done.
extern "C" int dpfs_debug_rs_schedule(int M, int n, int* order, int* owner_lo, int* owner_hi) {
  if (n < 1 || M <= 0 || M % n != 0 || M / n < BM) return -53;
  const int rpr = M / n, tiles_m = (M + BM - 1) / BM;
  for (int i = 0; i < tiles_m; ++i) order[i] = rs_block_of(i, n, rpr);
  for (int b = 0; b < tiles_m; ++b) {
    const int m0 = b * BM, m1 = (m0 + BM < M ? m0 + BM : M) - 1;
    owner_lo[b] = m0 / rpr;
    owner_hi[b] = m1 / rpr;
  }
  return 0;
}
