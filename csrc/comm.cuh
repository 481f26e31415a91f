// Device-side building blocks of the in-kernel NVLink collectives: system-scope signal flags in peer memory,
// rank-wide block barriers, and NVLS multicast loads/stores (multimem.*) through the NVSwitch.
//
// Symmetric-memory layout (allocated and exchanged by parallel/symm_comm.py):
//   data   : one buffer per rank, mapped in every peer (P2P) and through one multicast address (NVLS)
//   flags  : uint32 per rank; region 0 = [kMaxBlocks][kMaxRanks] block-barrier semaphores,
//            region 1 = [kMaxTiles][kMaxRanks] "partial tile ready" semaphores of the fused GEMM + all-reduce
// Every flag is a binary semaphore between one sender and one receiver (sender flips 0->1 with release semantics,
// receiver flips 1->0 with acquire semantics), so there are no epochs to keep and CUDA-graph replays stay correct.
#pragma once

#include "common.cuh"

namespace dpfs {

constexpr int kMaxRanks = 8;
constexpr int kMaxBlocks = 1024;
constexpr int kMaxTiles = 16384;
constexpr int kFlagWords = (kMaxBlocks + kMaxTiles) * kMaxRanks;

struct CommCtx {
  uint32_t* flags[kMaxRanks];  // flags[r] = rank r's flag buffer (peer-mapped)
  int rank;
  int n;
};

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t v;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v)::"memory");
  return v;
}
__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val)
               : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val)
               : "memory");
  return old;
}
// A peer that never arrives must not hang the box: after ~20 s of spinning the kernel traps.
constexpr uint64_t kSpinTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;
__device__ __forceinline__ void put_signal(uint32_t* addr) {
  const uint64_t t0 = globaltimer_ns();
  while (cas_release_sys(addr, 0u, 1u) != 0u) {
    if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
  }
}
__device__ __forceinline__ void wait_signal(uint32_t* addr) {
  const uint64_t t0 = globaltimer_ns();
  while (cas_acquire_sys(addr, 1u, 0u) != 1u) {
    if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
  }
}

// Barrier between the CTAs with the same `slot` on every rank.  All threads of the CTA must call it.  Writes made by
// any thread of the CTA before the call are visible to every thread of the matching CTAs after it.
__device__ __forceinline__ void sync_ranks(const CommCtx& c, int slot) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < c.n) {
    const int peer = threadIdx.x;
    put_signal(c.flags[peer] + slot * kMaxRanks + c.rank);
    wait_signal(c.flags[c.rank] + slot * kMaxRanks + peer);
  }
  __syncthreads();
}

// NVLS: one load returns the sum over all ranks' copies (reduced inside the switch, fp32 accumulation for bf16).
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc_ptr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc_ptr)
               : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_f32x4(const void* mc_ptr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc_ptr)
               : "memory");
  return v;
}
// NVLS: one store lands in every rank's copy.
__device__ __forceinline__ void multimem_st_16B(void* mc_ptr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_ptr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

}  // namespace dpfs
