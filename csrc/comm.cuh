// Device-side building blocks of the in-kernel NVLink collectives: system-scope signal flags in peer memory,
// rank-wide block barriers, and NVLS multicast loads/stores (multimem.*) through the NVSwitch.
//
// Symmetric-memory layout (allocated and exchanged by parallel/symm_comm.py):
//   data   : one buffer per rank, mapped in every peer (P2P) and through one multicast address (NVLS)
//   flags  : uint32 per rank; region 0 = [kMaxBlocks][kMaxRanks] block-barrier semaphores,
//            region 1 = [kMaxTiles][kMaxRanks] "partial tile ready" semaphores of the fused GEMM + all-reduce
//            region 2 = [kMaxTiles][kMaxRanks] "partial tile ready" EPOCH words of the fused GEMM + hand-off
//                       (gemm_handoff.cu: written with plain release stores, never reset)
// Every flag is a binary semaphore between one sender and one receiver (sender flips 0->1 with release semantics,
// receiver flips 1->0 with acquire semantics), so there are no epochs to keep and CUDA-graph replays stay correct.
#pragma once

#include "common.cuh"

namespace dpfs {

constexpr int kMaxRanks = 8;
constexpr int kMaxBlocks = 1024;
constexpr int kMaxTiles = 16384;
constexpr int kFlagWords = (kMaxBlocks + 2 * kMaxTiles) * kMaxRanks;

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

// ---------------------------------------------------------------------------------------------------------------
// P2P data path: plain loads / stores on peer-mapped pointers.  Measured on 2xB200 (profiles/comm_paths_r1.md) this
// moves ~1.5-2x the bytes per SM-thread of the multimem path, so it is the default for the reduce phase.
struct PeerPtrs {
  uint8_t* p[kMaxRanks];
};
__device__ __forceinline__ uint4 ld_peer_16B(const void* p) {
  uint4 v;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_16B(void* p, uint4 v) {
  asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// Sum the NR ranks' copies of a [rows][cpr x 16 B] bf16 region (row pitch ld_bytes, starting at byte offset off0) in
// fp32, in rank order (bitwise identical on every rank and run to run), and write the bf16 result into every rank's
// copy.  16 / NR independent 16-byte loads per rank are in flight per thread.
template <int NR, int INFLIGHT = 16>
__device__ __forceinline__ void reduce_region_p2p(const PeerPtrs& peers, long long off0, long long ld_bytes, int rows,
                                                  int cpr, int tid, int nthreads) {
  constexpr int KC = INFLIGHT / NR;
  const int total = rows * cpr;
  for (int base = tid; base < total; base += nthreads * KC) {
    uint4 v[KC][NR];
    long long off[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const int idx = base + k * nthreads;
      const int r_ = idx / cpr, cc = idx - r_ * cpr;
      off[k] = off0 + r_ * ld_bytes + cc * 16;
      if (idx < total) {
#pragma unroll
        for (int r = 0; r < NR; ++r) v[k][r] = ld_peer_16B(peers.p[r] + off[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      if (base + k * nthreads < total) {
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
        const uint4 o = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]),
                                   pack_bf16(acc[6], acc[7]));
#pragma unroll
        for (int r = 0; r < NR; ++r) st_peer_16B(peers.p[r] + off[k], o);
      }
    }
  }
}

// Reduce-scatter flavour: sum the NR ranks' copies of rows [0, rows) x (cpr x 16 B) of a bf16 region (row pitch
// ld_bytes, byte offset off0 inside every rank's buffer) in fp32, rank order, and write the bf16 result ONLY to this
// rank's private output `dst` (row pitch dst_ld_bytes).
template <int NR, int INFLIGHT = 8>
__device__ __forceinline__ void reduce_region_to_local_p2p(const PeerPtrs& peers, long long off0, long long ld_bytes,
                                                           uint8_t* dst, long long dst_ld_bytes, int rows, int cpr,
                                                           int tid, int nthreads) {
  constexpr int KC = INFLIGHT / NR > 0 ? INFLIGHT / NR : 1;
  const int total = rows * cpr;
  for (int base = tid; base < total; base += nthreads * KC) {
    uint4 v[KC][NR];
    long long doff[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const int idx = base + k * nthreads;
      const int r_ = idx / cpr, cc = idx - r_ * cpr;
      doff[k] = r_ * dst_ld_bytes + cc * 16;
      if (idx < total) {
        const long long off = off0 + r_ * ld_bytes + cc * 16;
#pragma unroll
        for (int r = 0; r < NR; ++r) v[k][r] = ld_peer_16B(peers.p[r] + off);
      }
    }
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      if (base + k * nthreads < total) {
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
        *reinterpret_cast<uint4*>(dst + doff[k]) =
            make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]),
                       pack_bf16(acc[6], acc[7]));
      }
    }
  }
}
// Same through the switch: one multimem.ld_reduce returns the fp32-accumulated sum of all ranks' copies.
template <int INFLIGHT = 4>
__device__ __forceinline__ void reduce_region_to_local_nvls(const uint8_t* mc_base, long long off0, long long ld_bytes,
                                                            uint8_t* dst, long long dst_ld_bytes, int rows, int cpr,
                                                            int tid, int nthreads) {
  const int total = rows * cpr;
  for (int base = tid; base < total; base += nthreads * INFLIGHT) {
    uint4 v[INFLIGHT];
    long long doff[INFLIGHT];
#pragma unroll
    for (int k = 0; k < INFLIGHT; ++k) {
      const int idx = base + k * nthreads;
      const int r_ = idx / cpr, cc = idx - r_ * cpr;
      doff[k] = r_ * dst_ld_bytes + cc * 16;
      if (idx < total) v[k] = multimem_ld_reduce_bf16x8(mc_base + off0 + r_ * ld_bytes + cc * 16);
    }
#pragma unroll
    for (int k = 0; k < INFLIGHT; ++k)
      if (base + k * nthreads < total) *reinterpret_cast<uint4*>(dst + doff[k]) = v[k];
  }
}

// All-reduce flavour through the switch, in place: pull the switch-reduced chunk and broadcast it back into every
// rank's copy.  Every chunk is read by exactly one thread of one rank (its owner) before that thread overwrites it.
template <int INFLIGHT = 4>
__device__ __forceinline__ void reduce_region_nvls(uint8_t* mc_base, long long off0, long long ld_bytes, int rows,
                                                   int cpr, int tid, int nthreads) {
  const int total = rows * cpr;
  for (int base = tid; base < total; base += nthreads * INFLIGHT) {
    uint4 v[INFLIGHT];
    long long off[INFLIGHT];
#pragma unroll
    for (int k = 0; k < INFLIGHT; ++k) {
      const int idx = base + k * nthreads;
      const int r_ = idx / cpr, cc = idx - r_ * cpr;
      off[k] = off0 + r_ * ld_bytes + cc * 16;
      if (idx < total) v[k] = multimem_ld_reduce_bf16x8(mc_base + off[k]);
    }
#pragma unroll
    for (int k = 0; k < INFLIGHT; ++k)
      if (base + k * nthreads < total) multimem_st_16B(mc_base + off[k], v[k]);
  }
}

// fp32 variant (4 floats per 16-byte unit)
template <int NR>
__device__ __forceinline__ void reduce_region_p2p_f32(const PeerPtrs& peers, long long off0, long long units, int tid,
                                                      long long nthreads) {
  constexpr int KC = 16 / NR;
  for (long long base = tid; base < units; base += nthreads * KC) {
    uint4 v[KC][NR];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const long long u = base + k * nthreads;
      if (u < units) {
#pragma unroll
        for (int r = 0; r < NR; ++r) v[k][r] = ld_peer_16B(peers.p[r] + off0 + u * 16);
      }
    }
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const long long u = base + k * nthreads;
      if (u < units) {
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          a0 += __uint_as_float(v[k][r].x);
          a1 += __uint_as_float(v[k][r].y);
          a2 += __uint_as_float(v[k][r].z);
          a3 += __uint_as_float(v[k][r].w);
        }
        const uint4 o = make_uint4(__float_as_uint(a0), __float_as_uint(a1), __float_as_uint(a2), __float_as_uint(a3));
#pragma unroll
        for (int r = 0; r < NR; ++r) st_peer_16B(peers.p[r] + off0 + u * 16, o);
      }
    }
  }
}

}  // namespace dpfs
