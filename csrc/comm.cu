// Stand-alone NVLink / NVSwitch collectives over symmetric memory (the fused GEMM + all-reduce lives in
// gemm_tcgen05.cu and shares comm.cuh).  These replace the NCCL calls behind the reference's Reduce / Copy / Gather
// ops (models/comm_ops.py:26,39,59,74) on the hot path; NCCL stays as the bootstrap and the correctness oracle.
//
// all-reduce (two-shot through the switch): rank r owns slice r of the buffer; for its slice it issues
// multimem.ld_reduce (the NVSwitch returns the sum over all ranks) and multimem.st (the switch writes the result into
// every rank's copy).  Link traffic per GPU ~ S (1 + 1/n) instead of 2S(n-1)/n for a ring.
#include "comm.cuh"
#include "dpfs_b200.h"

namespace dpfs {

int num_sms();

template <bool BF16>
__global__ void __launch_bounds__(512) allreduce_nvls_kernel(CommCtx c, uint8_t* mc_base, long long units) {
  sync_ranks(c, blockIdx.x);  // every rank's input is complete and visible
  const long long per = (units + c.n - 1) / c.n;
  const long long u0 = per * c.rank;
  const long long u1 = min(units, u0 + per);
  for (long long u = u0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; u < u1;
       u += static_cast<long long>(gridDim.x) * blockDim.x) {
    uint8_t* p = mc_base + u * 16;
    const uint4 v = BF16 ? multimem_ld_reduce_bf16x8(p) : multimem_ld_reduce_f32x4(p);
    multimem_st_16B(p, v);
  }
  sync_ranks(c, blockIdx.x);  // every rank's slice has landed everywhere
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

// rank-wide barrier only (used by tests and to order non-kernel accesses)
__global__ void barrier_kernel(CommCtx c) { sync_ranks(c, blockIdx.x); }

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

extern "C" int dpfs_allgather_nvls(const void* src, void* mc_dst, long long nbytes_per_rank,
                                   const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  if (nbytes_per_rank % 16 != 0 || (reinterpret_cast<uintptr_t>(mc_dst) & 15) ||
      (reinterpret_cast<uintptr_t>(src) & 15))
    return -51;
  if (nbytes_per_rank == 0) return 0;
  const long long units = nbytes_per_rank / 16;
  long long blocks = (units + 511) / 512;
  if (blocks > 32) blocks = 32;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  allgather_nvls_kernel<<<(DPFS_COUNT(), static_cast<int>(blocks)), 512, 0, stream>>>(
      c, static_cast<const uint4*>(src), static_cast<uint8_t*>(mc_dst), units);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_comm_barrier(const unsigned long long* flag_ptrs, int rank, int n, cudaStream_t stream) {
  if (n < 2 || n > kMaxRanks) return -50;
  CommCtx c = make_ctx(flag_ptrs, rank, n);
  barrier_kernel<<<(DPFS_COUNT(), 1), 32, 0, stream>>>(c);
  return static_cast<int>(cudaGetLastError());
}
