// Stand-alone NVLink / NVSwitch collectives over symmetric memory (the fused GEMM + all-reduce lives in
// gemm_tcgen05.cu and shares comm.cuh).  These replace the NCCL calls behind the reference's Reduce / Copy / Gather
// ops (models/comm_ops.py:26,39,59,74) on the hot path; NCCL stays as the bootstrap and the correctness oracle.
//
// all-reduce (two-shot through the switch): rank r owns slice r of the buffer; for its slice it issues
// multimem.ld_reduce (the NVSwitch returns the sum over all ranks) and multimem.st (the switch writes the result into
// every rank's copy).  Link traffic per GPU ~ S (1 + 1/n) instead of 2S(n-1)/n for a ring.
#include "comm.cuh"
#include "dpfs_b200.h"

#include <stdlib.h>

namespace dpfs {

int num_sms();

template <bool BF16>
__global__ void __launch_bounds__(512) allreduce_nvls_kernel(CommCtx c, uint8_t* mc_base, long long units) {
  sync_ranks(c, blockIdx.x);  // every rank's input is complete and visible
  const long long per = (units + c.n - 1) / c.n;
  const long long u0 = per * c.rank;
  const long long u1 = min(units, u0 + per);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long ub = u0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; ub < u1; ub += stride * 8) {
    uint4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const long long u = ub + k * stride;
      if (u < u1) v[k] = BF16 ? multimem_ld_reduce_bf16x8(mc_base + u * 16) : multimem_ld_reduce_f32x4(mc_base + u * 16);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const long long u = ub + k * stride;
      if (u < u1) multimem_st_16B(mc_base + u * 16, v[k]);
    }
  }
  sync_ranks(c, blockIdx.x);  // every rank's slice has landed everywhere
}


// P2P two-shot all-reduce: rank r sums slice r of every rank's buffer (plain NVLink loads, fp32 accumulation in rank
// order) and writes the result into every rank's buffer (plain NVLink stores).
template <int NR, bool BF16>
__global__ void __launch_bounds__(512) allreduce_p2p_kernel(CommCtx c, PeerPtrs peers, long long units) {
  sync_ranks(c, blockIdx.x);
  const long long per = (units + NR - 1) / NR;
  const long long u0 = per * c.rank;
  const long long cnt = max(0ll, min(units, u0 + per) - u0);
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  if constexpr (BF16) {
    // treat the slice as one long row of 16-byte chunks, processed in pieces that fit an int
    const long long piece = 1ll << 24;
    for (long long p0 = 0; p0 < cnt; p0 += piece)
      reduce_region_p2p<NR>(peers, (u0 + p0) * 16, 0, 1, static_cast<int>(min(piece, cnt - p0)), tid,
                            static_cast<int>(nthreads));
  } else {
    reduce_region_p2p_f32<NR>(peers, u0 * 16, cnt, tid, nthreads);
  }
  sync_ranks(c, blockIdx.x);
}


// P2P reduce-scatter over rows (sequence parallelism): the n ranks hold partial [rows_total][row_units*16 B] bf16
// matrices at the same symmetric offset; rank r sums ITS block of rows over all ranks (fp32, rank order) and writes the
// bf16 result to a private local buffer.  This is the first half of the two-shot all-reduce; the all-gather half is
// allgather_p2p_kernel.  ref: the reference has no reduce-scatter (sequence parallel is an extension).
template <int NR>
__global__ void __launch_bounds__(512) reduce_scatter_p2p_kernel(CommCtx c, PeerPtrs peers, uint4* __restrict__ out,
                                                                 long long units_per_rank) {
  sync_ranks(c, blockIdx.x);  // every rank's partial matrix is complete
  constexpr int KC = 16 / NR;
  const long long src0 = static_cast<long long>(c.rank) * units_per_rank;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long base = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; base < units_per_rank;
       base += nthreads * KC) {
    uint4 v[KC][NR];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const long long u = base + k * nthreads;
      if (u < units_per_rank) {
#pragma unroll
        for (int r = 0; r < NR; ++r) v[k][r] = ld_peer_16B(peers.p[r] + (src0 + u) * 16);
      }
    }
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const long long u = base + k * nthreads;
      if (u < units_per_rank) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          const uint32_t w[4] = {v[k][r].x, v[k][r].y, v[k][r].z, v[k][r].w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = unpack_bf16(w[i]);
            acc[2 * i] += f.x;
            acc[2 * i + 1] += f.y;
          }
        }
        out[u] = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]),
                            pack_bf16(acc[6], acc[7]));
      }
    }
  }
  sync_ranks(c, blockIdx.x);  // nobody overwrites its partial matrix while a peer still reads it
}


// NVLS reduce-scatter: rank r pulls ITS block of rows already summed by the NVSwitch (multimem.ld_reduce) into private
// memory — each GPU receives only S/n bytes.
__global__ void __launch_bounds__(512) reduce_scatter_nvls_kernel(CommCtx c, const uint8_t* mc_base,
                                                                  uint4* __restrict__ out, long long units_per_rank) {
  sync_ranks(c, blockIdx.x);
  const uint8_t* src = mc_base + static_cast<long long>(c.rank) * units_per_rank * 16;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long ub = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; ub < units_per_rank;
       ub += stride * 8) {
    uint4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const long long u = ub + k * stride;
      if (u < units_per_rank) v[k] = multimem_ld_reduce_bf16x8(src + u * 16);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const long long u = ub + k * stride;
      if (u < units_per_rank) out[u] = v[k];
    }
  }
  sync_ranks(c, blockIdx.x);
}

// all-gather: rank r broadcasts its `units` 16-byte units into slot r of every rank's buffer with multimem.st
__global__ void __launch_bounds__(512) allgather_nvls_kernel(CommCtx c, const uint4* __restrict__ src, uint8_t* mc_base,
                                                             long long units) {
  sync_ranks(c, blockIdx.x);  // the destination buffers are free on every rank
  uint8_t* dst = mc_base + static_cast<long long>(c.rank) * units * 16;
  for (long long u = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; u < units;
       u += static_cast<long long>(gridDim.x) * blockDim.x)
    multimem_st_16B(dst + u * 16, src[u]);
  sync_ranks(c, blockIdx.x);
}


// Experimental variants used by tools/gpu_ar_sweep.py to find the fastest data path on a given box:
//   mode 1: multimem.ld_reduce only   mode 2: multimem.st only
//   mode 3: P2P loads from every peer + fp32 add on the SM + multimem.st      mode 4: P2P loads + P2P stores
__device__ __forceinline__ uint4 add_bf16x8(uint4 a, uint4 b) {
  uint4 o;
  const uint32_t* pa = &a.x;
  const uint32_t* pb = &b.x;
  uint32_t* po = &o.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = unpack_bf16(pa[i]), y = unpack_bf16(pb[i]);
    po[i] = pack_bf16(x.x + y.x, x.y + y.y);
  }
  return o;
}
__global__ void __launch_bounds__(512) allreduce_exp_kernel(CommCtx c, PeerPtrs peers, uint8_t* mc_base, long long off,
                                                            long long units, int mode) {
  const bool nosync = (mode & 16) != 0;   // timing experiments on concurrent streams: no barriers (flags are per block)
  mode &= 15;
  if (!nosync) sync_ranks(c, blockIdx.x);
  const long long per = (units + c.n - 1) / c.n;
  const long long u0 = per * c.rank;
  const long long u1 = min(units, u0 + per);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long ub = u0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; ub < u1; ub += stride * 4) {
    uint4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long u = ub + k * stride;
      if (u >= u1) continue;
      if (mode == 1) {
        v[k] = multimem_ld_reduce_bf16x8(mc_base + off + u * 16);
      } else if (mode == 2) {
        v[k] = make_uint4(0, 0, 0, 0);
      } else {
        // fp32 accumulation over the ranks in rank order (deterministic)
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int r = 0; r < c.n; ++r) {
          const uint4 t = ld_peer_16B(peers.p[r] + off + u * 16);
          const uint32_t* pt = &t.x;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = unpack_bf16(pt[i]);
            acc[2 * i] += f.x;
            acc[2 * i + 1] += f.y;
          }
        }
        v[k] = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]),
                          pack_bf16(acc[6], acc[7]));
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long u = ub + k * stride;
      if (u >= u1) continue;
      if (mode == 1) {
        if (v[k].x == 0x12345678u) st_peer_16B(peers.p[c.rank] + off + u * 16, v[k]);  // keep the load alive
      } else if (mode == 2 || mode == 3) {
        multimem_st_16B(mc_base + off + u * 16, v[k]);
      } else {
        for (int r = 0; r < c.n; ++r) st_peer_16B(peers.p[r] + off + u * 16, v[k]);
      }
    }
  }
  if (!nosync) sync_ranks(c, blockIdx.x);
}

// all-gather over plain peer stores (used when the buffers have no multicast mapping)
__global__ void __launch_bounds__(512) allgather_p2p_kernel(CommCtx c, PeerPtrs peers, const uint4* __restrict__ src,
                                                            long long units) {
  sync_ranks(c, blockIdx.x);
  const long long dst_off = static_cast<long long>(c.rank) * units * 16;
  for (long long u = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; u < units;
       u += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint4 v = src[u];
    for (int r = 0; r < c.n; ++r) st_peer_16B(peers.p[r] + dst_off + u * 16, v);
  }
  sync_ranks(c, blockIdx.x);
}

// rank-wide barrier only (used by tests and to order non-kernel accesses)
__global__ void barrier_kernel(CommCtx c) { sync_ranks(c, blockIdx.x); }
// `reps` back-to-back barriers in every block (micro-benchmark of the barrier itself)
__global__ void barrier_reps_kernel(CommCtx c, int reps) {
  for (int i = 0; i < reps; ++i) sync_ranks(c, blockIdx.x);
}

static CommCtx make_ctx(const unsigned long long* flag_ptrs, int rank, int n) {
  CommCtx c;
  for (int i = 0; i < kMaxRanks; ++i) c.flags[i] = i < n ? reinterpret_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
  c.rank = rank;
  c.n = n;
  return c;
}

}  // namespace dpfs

using namespace dpfs;

extern "C" int dpfs_comm_flag_words(void) { return kFlagWords; }

static PeerPtrs make_peers(const unsigned long long* data_ptrs, long long offset, int n) {
  PeerPtrs pp;
  for (int i = 0; i < kMaxRanks; ++i) pp.p[i] = i < n ? reinterpret_cast<uint8_t*>(data_ptrs[i]) + offset : nullptr;
  return pp;
}

extern "C" int dpfs_allreduce_nvls(void* mc_ptr, long long nbytes, int is_bf16, const unsigned long long* flag_ptrs,
                                   int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (nbytes % 16 != 0 || (reinterpret_cast<uintptr_t>(mc_ptr) & 15)) return -51;
  if (nbytes == 0) return 0;
  const long long units = nbytes / 16;
  long long blocks = (units / n + 511) / 512;
  if (blocks > 64) blocks = 64;
  if (blocks < 1) blocks = 1;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  if (is_bf16)
    allreduce_nvls_kernel<true><<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
        c, static_cast<uint8_t*>(mc_ptr), units);
  else
    allreduce_nvls_kernel<false><<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
        c, static_cast<uint8_t*>(mc_ptr), units);
  return static_cast<int>(cudaGetLastError());
}

template <int NR>
static int launch_ar_p2p(const CommCtx& c, const PeerPtrs& pp, long long units, int is_bf16, cudaStream_t stream) {
  long long blocks = (units / NR + 511) / 512;
  const long long cap = 2ll * num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (is_bf16)
    allreduce_p2p_kernel<NR, true><<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(c, pp, units);
  else
    allreduce_p2p_kernel<NR, false><<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(c, pp, units);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_allreduce_p2p(const unsigned long long* data_ptrs, long long offset, long long nbytes, int is_bf16,
                                  const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n != 2 && n != 4 && n != 8) return -50;
  if (nbytes % 16 != 0 || offset % 16 != 0) return -51;
  if (nbytes == 0) return 0;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  PeerPtrs pp = make_peers(data_ptrs, offset, n);
  const long long units = nbytes / 16;
  if (n == 2) return launch_ar_p2p<2>(c, pp, units, is_bf16, stream);
  if (n == 4) return launch_ar_p2p<4>(c, pp, units, is_bf16, stream);
  return launch_ar_p2p<8>(c, pp, units, is_bf16, stream);
}

extern "C" int dpfs_allgather_nvls(const void* src, void* mc_dst, long long nbytes_per_rank,
                                   const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (nbytes_per_rank % 16 != 0 || (reinterpret_cast<uintptr_t>(mc_dst) & 15) ||
      (reinterpret_cast<uintptr_t>(src) & 15))
    return -51;
  if (nbytes_per_rank == 0) return 0;
  const long long units = nbytes_per_rank / 16;
  long long blocks = (units + 511) / 512;
  if (blocks > 64) blocks = 64;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  allgather_nvls_kernel<<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
      c, static_cast<const uint4*>(src), static_cast<uint8_t*>(mc_dst), units);
  return static_cast<int>(cudaGetLastError());
}


template <int NR>
static int launch_rs_p2p(const CommCtx& c, const PeerPtrs& pp, void* out, long long units, cudaStream_t stream) {
  long long blocks = (units + 511) / 512;
  const long long cap = 2ll * num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  reduce_scatter_p2p_kernel<NR><<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
      c, pp, static_cast<uint4*>(out), units);
  return static_cast<int>(cudaGetLastError());
}

// out (local, bf16) = sum over ranks of rows [rank*rows/n, (rank+1)*rows/n) of the symmetric partial matrix at `offset`
extern "C" int dpfs_reduce_scatter_p2p(const unsigned long long* data_ptrs, long long offset, long long nbytes_total,
                                       void* out, const unsigned long long* flag_ptrs, int rank, int n,
                                       cudaStream_t stream) {
  if (n != 2 && n != 4 && n != 8) return -50;
  if (nbytes_total % (16ll * n) != 0 || offset % 16 != 0 || (reinterpret_cast<uintptr_t>(out) & 15)) return -51;
  if (nbytes_total == 0) return 0;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  PeerPtrs pp = make_peers(data_ptrs, offset, n);
  const long long units = nbytes_total / 16 / n;
  if (n == 2) return launch_rs_p2p<2>(c, pp, out, units, stream);
  if (n == 4) return launch_rs_p2p<4>(c, pp, out, units, stream);
  return launch_rs_p2p<8>(c, pp, out, units, stream);
}


extern "C" int dpfs_reduce_scatter_nvls(const void* mc_ptr, long long nbytes_total, void* out,
                                        const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (nbytes_total % (16ll * n) != 0 || (reinterpret_cast<uintptr_t>(mc_ptr) & 15) ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return -51;
  if (nbytes_total == 0) return 0;
  const long long units = nbytes_total / 16 / n;
  long long blocks = (units + 511) / 512;
  if (blocks > 64) blocks = 64;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  reduce_scatter_nvls_kernel<<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
      c, static_cast<const uint8_t*>(mc_ptr), static_cast<uint4*>(out), units);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_allgather_p2p(const void* src, const unsigned long long* data_ptrs, long long dst_offset,
                                  long long nbytes_per_rank, const unsigned long long* flag_ptrs, int rank, int n,
                                  cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (nbytes_per_rank % 16 != 0 || dst_offset % 16 != 0 || (reinterpret_cast<uintptr_t>(src) & 15)) return -51;
  if (nbytes_per_rank == 0) return 0;
  const long long units = nbytes_per_rank / 16;
  long long blocks = (units + 511) / 512;
  if (blocks > 2ll * num_sms()) blocks = 2ll * num_sms();
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  PeerPtrs pp = make_peers(data_ptrs, dst_offset, n);
  allgather_p2p_kernel<<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
      c, pp, static_cast<const uint4*>(src), units);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_comm_barrier(const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  barrier_kernel<<<(DPFS_COUNT(), 1), 32, 0, stream>>>(c);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_comm_barrier_ex(const unsigned long long* flag_ptrs, int rank, int n, int blocks, int threads,
                                    int reps, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks || blocks < 1 || blocks > kMaxBlocks || threads < 32 || threads > 1024) return -50;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  barrier_reps_kernel<<<(DPFS_COUNT(), blocks), threads, 0, stream>>>(c, reps);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_allreduce_exp(void* mc_base, const unsigned long long* data_ptrs, long long offset,
                                  long long nbytes, int mode, int blocks, const unsigned long long* flag_ptrs,
                                  int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks || nbytes % 16) return -50;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  PeerPtrs pp;
  for (int i = 0; i < kMaxRanks; ++i) pp.p[i] = i < n ? reinterpret_cast<uint8_t*>(data_ptrs[i]) : nullptr;
  allreduce_exp_kernel<<<(DPFS_COUNT(), blocks), 512, 0, stream>>>(c, pp, static_cast<uint8_t*>(mc_base), offset,
                                                                  nbytes / 16, mode);
  return static_cast<int>(cudaGetLastError());
}
