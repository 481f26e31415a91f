// 2-SM (cta_group::2) variant of the bf16 GEMM: a thread-block cluster of two CTAs on one TPC computes a 256x256 output
// tile with ONE stream of tcgen05.mma.cta_group::2 instructions issued by the leader CTA.  Each CTA stages only its own
// 128 rows of A and HALF of the B tile (128 of the 256 N rows), so the L2->SMEM operand traffic per FLOP is 2/3 of the
// 1-SM 128x256 kernel — the 1-SM kernel is bound by exactly that traffic on the transformer shapes
// (profiles/step_breakdown_tp1_r1.md).  Accumulators: 128 lanes x 256 fp32 columns per CTA, double buffered (512 TMEM
// columns), so the epilogue of tile i still overlaps the main loop of tile i+1.
//
// Pipeline (per CTA): warp 0 = TMA producer (both CTAs load; transaction bytes of BOTH land on the leader's full
// barrier), warp 1 = MMA issuer (leader CTA only) + TMEM owner (both CTAs allocate with cta_group::2), warps 2-5 =
// epilogue.  tcgen05.commit multicasts "slot free" / "accumulator ready" to the same barrier offsets in both CTAs; the
// follower's epilogue warps release the accumulator with a remote mbarrier arrive on the leader.
//
// Same operand-major options and epilogues as gemm_tcgen05.cu (ref: F.linear in models/layers.py:49,93).
#include "common.cuh"
#include "dpfs_b200.h"

#include <cudaTypedefs.h>

namespace dpfs {

int make_tmap_bf16_2d(CUtensorMap* tm, const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                      uint32_t box_inner, uint32_t box_outer);
int num_sms();

namespace sm2 {

constexpr int BM = 128;   // rows per CTA (256 per cluster)
constexpr int BNH = 128;  // B rows staged per CTA (half of the 256-wide tile)
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int kABytes = BM * BK * 2;
constexpr int kBBytes = BNH * BK * 2;
constexpr int kStageBytes = kABytes + kBBytes;
// TMA-store staging: kOutBufs slabs of 32 rows x 128 B per epilogue warp (a ring of 4 measured no faster).
constexpr int kOutBufs = 2;
// Two shapes of the same kernel.  EW = 4: four epilogue warps, six operand stages -- for long K, where the epilogue
// hides behind the next tile's main loop.  EW = 8: two warps per TMEM lane quarter (each takes half of the tile's
// columns), five stages -- for short K (<= 1024), where a 256x256 tile's main loop (2.2 us at K = 512) is shorter
// than one warp set's epilogue and the epilogue paces the kernel.  (10 warps put 3 on one SM sub-partition, so the
// compiler caps that shape at 168 registers: 16384 / 3 per warp.)
// SWIGLU flavour (gate|up projection with the activation in the epilogue, see gemm2sm_kernel): three output slabs per
// warp (gate, up, act) instead of a ring of two, paid for with one operand stage (K = d is short there anyway).
// EPI: 0 plain epilogue, 1 SwiGLU forward (three slabs per warp), 2 SwiGLU backward (four slabs per warp: the gate and
// up halves of two 64-column groups; three operand stages -- that kernel is bound by its 524 MB of epilogue traffic).
template <int EW, int EPI = 0>
struct Cfg {
  static constexpr int kStages = EPI == 1 ? 4 : EPI == 2 ? 3 : (EW == 8 ? 5 : 6);
  static constexpr int kThreads = 64 + EW * 32;
  static constexpr int kBufs = EPI == 1 ? 3 : EPI == 2 ? 4 : kOutBufs;
  static constexpr int kOutStageBytes = EW * kBufs * 4096;
  static constexpr int kSmem = kStages * kStageBytes + kOutStageBytes + 256 + 1024;
  static_assert(kSmem <= 232448, "exceeds the 227 KB shared-memory limit");
};
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared-window address -> leader CTA's copy

struct Args {
  int M, N, K;
  int splits, kb_per_split;
  void* out;
  long long ldo, split_stride;
  const float* bias;
  const float* residual;
  long long ldr;
  int F;  // SwiGLU flavours: hidden units (forward: N = 2 F, gate rows [0, F), up rows [F, 2F) of B; backward: N = F)
  const __nv_bfloat16* gu;  // SwiGLU backward: the saved gate|up activation [M][2F]
  float* colsum;            // SwiGLU backward: per-(32-row group) column sums of d(gate|up), [ceil(M/256)*8][2F]
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma2_bf16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      :
      : "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all previously issued MMAs completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma2_commit_mc(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(static_cast<uint16_t>(3))
      : "memory");
}
// TMA tile load whose transaction bytes are accounted on the LEADER CTA's mbarrier
__device__ __forceinline__ void tma2_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];"
      :
      : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerMask) : "memory");
}

template <typename OutT>
__device__ __forceinline__ void store32(OutT* dst, const float (&v)[32], int ncols_valid);
template <>
__device__ __forceinline__ void store32<float>(float* dst, const float (&v)[32], int ncols_valid) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j * 4 + 4 <= ncols_valid)
      *reinterpret_cast<float4*>(dst + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
template <>
__device__ __forceinline__ void store32<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[32], int ncols_valid) {
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

// SWIGLU (forward of the fused gate|up projection, EW = 8, K-major operands, bf16 out): the B tile of a CTA pair is
// "128 gate rows | the 128 up rows of the SAME hidden units" (CTA 0 loads rows [128 j, +128), CTA 1 rows [F + 128 j,
// +128) of the concatenated weight), so an accumulator row holds g in TMEM columns [0, 128) and u in [128, 256) and an
// epilogue thread has both of a hidden unit in registers: it emits gate|up (+ bias, bf16: the backward needs it) AND
// act = silu(g) * u (through tmAct) in one pass -- the 2 x M x F re-read of the stand-alone SwiGLU kernel disappears.
// Parameters, checkpoints, wgrad and dgrad are untouched.  ref: models/model.py:95.
//
// EPI = 2, SwiGLU BACKWARD in the epilogue of the down_proj dgrad (A = dy K-major, B = W_down MN-major, N = F): the
// accumulator is dact = dy W_down; an epilogue thread (one row) reads the saved gate / up values of its 32 hidden units,
// turns dact into d(gate) and d(up) in registers and emits both halves of d(gate|up) [M][2F] (bf16, through tmOut) plus
// the column sums of each 32-row slab (the gate|up bias gradient, finished by a fixed-order reduction).  dact is
// never written: the stand-alone SwiGLU-backward kernel's 131 MB write + 131 MB read per layer disappear.
template <bool A_MN, bool B_MN, typename OutT, int EW, int EPI = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Cfg<EW, EPI>::kThreads, 1)
gemm2sm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmOut, const Args args,
               const __grid_constant__ CUtensorMap tmAct) {
  constexpr bool SWIGLU = EPI == 1;
  constexpr bool SWIGLU_BWD = EPI == 2;
  static_assert(!SWIGLU || (EW == 8 && !A_MN && !B_MN && sizeof(OutT) == 2), "SWIGLU: forward bf16 shape only");
  static_assert(!SWIGLU_BWD || (EW == 8 && !A_MN && B_MN && sizeof(OutT) == 2), "SWIGLU_BWD: dgrad bf16 shape only");
  constexpr int kStages = Cfg<EW, EPI>::kStages;
  constexpr int kOutStageBytes = Cfg<EW, EPI>::kOutStageBytes;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kABytes;
  uint8_t* sOut = smem + kStages * kStageBytes;                                    // TMA-store staging
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sOut + kOutStageBytes);         // used on the leader only
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;  // used on the leader only
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  const int tiles_m = (args.M + 2 * BM - 1) / (2 * BM);
  const int tiles_n = SWIGLU ? args.F / BNH : (args.N + BN - 1) / BN;
  const int tiles_mn = tiles_m * tiles_n;
  const int num_work = tiles_mn * args.splits;
  const int kb_total = (args.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmOut);
    if constexpr (SWIGLU || SWIGLU_BWD) prefetch_tmap(&tmAct);
    if constexpr (SWIGLU_BWD) {   // one "gate|up slabs landed" barrier per epilogue warp, behind the pipeline barriers
      uint64_t* gu_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(full_bar) + 128);
      for (int i = 0; i < EW; ++i) mbar_init(&gu_bar[i], 1);
    }
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 2 * EW);  // EW epilogue warps x 2 CTAs
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc2(tmem_slot, 2 * BN);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync();  // barriers of both CTAs initialised, TMEM allocated, before any cross-CTA signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = cluster_id; w < num_work; w += n_clusters) {
        const int split = w / tiles_mn;
        const int t = w - split * tiles_mn;
        const int m0 = (t / tiles_n) * (2 * BM) + cta * BM;   // this CTA's 128 rows of the 256-row tile
        // this CTA's half of the B tile (SWIGLU: the gate rows of hidden units [128 j, +128) / their up rows)
        const int n0 = SWIGLU ? (t % tiles_n) * BNH + static_cast<int>(cta) * args.F : (t % tiles_n) * BN + cta * BNH;
        const int kb0 = split * args.kb_per_split;
        const int kb1 = min(kb_total, kb0 + args.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);  // bytes of both CTAs
          uint8_t* a_dst = sA + stage * kABytes;
          uint8_t* b_dst = sB + stage * kBBytes;
          if constexpr (!A_MN) {
            tma2_load_2d(a_dst, &tmA, &full_bar[stage], kb * BK, m0);
          } else {
#pragma unroll
            for (int b = 0; b < BM / 64; ++b)
              tma2_load_2d(a_dst + b * (BK * 128), &tmA, &full_bar[stage], m0 + 64 * b, kb * BK);
          }
          if constexpr (!B_MN) {
            tma2_load_2d(b_dst, &tmB, &full_bar[stage], kb * BK, n0);
          } else {
#pragma unroll
            for (int b = 0; b < BNH / 64; ++b)
              tma2_load_2d(b_dst + b * (BK * 128), &tmB, &full_bar[stage], n0 + 64 * b, kb * BK);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int w = cluster_id; w < num_work; w += n_clusters) {
        const int split = w / tiles_mn;
        const int kb0 = split * args.kb_per_split;
        const int kb1 = min(kb_total, kb0 + args.kb_per_split);
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * kABytes);
          const uint32_t b_base = smem_u32(sB + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t adesc = A_MN ? make_smem_desc_sw128(a_base + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(a_base + k * 32, 0, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_base + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(b_base + k * 32, 0, 1024);
            umma2_bf16_ss(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma2_commit_mc(&empty_bar[stage]);  // frees this slot in both CTAs
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma2_commit_mc(&tmem_full[as]);  // accumulators of both CTAs complete
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else {
    // Epilogue: TMEM -> registers (bias / residual) -> 128B-swizzled smem slab -> TMA store.  Every thread owns one
    // output ROW, so direct global stores would touch 32 different rows per instruction (half-filled 32 B sectors);
    // staging 32 rows x 128 B per warp and letting the TMA unit write them gives full-line, coalesced writes, clips the
    // M / N edges for free, and takes the store instructions off the epilogue's critical path.
    const int q = warp & 3;   // the hardware ties a warp to TMEM lanes [32 * (warp % 4), +32)
    if constexpr (SWIGLU_BWD) {
      // warps 2-5: columns [0, 128) of the tile, warps 6-9: [128, 256) -- two groups of 64 hidden units each.  Four
      // slabs per warp: (gate, up) x (group 0, group 1), 32 rows x 64 bf16 columns each.
      const int half = (warp - 2) >> 2;
      uint8_t* slabs = sOut + (warp - 2) * 4 * 4096;
      uint64_t* gu_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(full_bar) + 128) + (warp - 2);
      uint32_t gphase = 0;
      const int F = args.F;
      int as = 0;
      uint32_t aphase = 0;
      for (int w = cluster_id; w < num_work; w += n_clusters) {
        const int t = w;   // splits == 1
        const int m0 = (t / tiles_n) * (2 * BM) + cta * BM;
        const int n0 = (t % tiles_n) * BN;
        // The saved gate / up values of this warp's 32 rows x 2 x 64 hidden units are fetched by TMA straight into the
        // four output slabs (same 128B-swizzled layout as the store), under the tile's main loop; each thread then
        // turns its own slab row into d(gate) / d(up) in place.  Rows past M are zero-filled by the load.
        if (lane == 0) {
          tma_store_wait_read<0>();   // the stores of the previous tile have read the slabs out
          mbar_expect_tx(gu_bar, 4 * 4096);
#pragma unroll
          for (int grp = 0; grp < 2; ++grp) {
            const int c0 = n0 + half * 128 + grp * 64;
            tma_load_3d(slabs + (2 * grp) * 4096, &tmAct, gu_bar, c0, m0 + q * 32, 0);
            tma_load_3d(slabs + (2 * grp + 1) * 4096, &tmAct, gu_bar, F + c0, m0 + q * 32, 0);
          }
        }
        __syncwarp();
        mbar_wait(&tmem_full[as], aphase);
        tc_fence_after();
        mbar_wait(gu_bar, gphase);
        gphase ^= 1;
        const uint32_t t_addr = tmem_base + as * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
        for (int grp = 0; grp < 2; ++grp) {
          const int c0 = n0 + half * 128 + grp * 64;   // first hidden unit of this group of 64
          uint8_t* row_g = slabs + (2 * grp) * 4096 + lane * 128;
          uint8_t* row_u = row_g + 4096;
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            uint32_t r[32];
            tmem_ld32(t_addr + half * 128 + grp * 64 + cc * 32, r);
            uint4 gq[4], uq[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int idx = ((cc * 4 + i) ^ (lane & 7)) << 4;   // 128B swizzle of the slab row
              gq[i] = *reinterpret_cast<const uint4*>(row_g + idx);
              uq[i] = *reinterpret_cast<const uint4*>(row_u + idx);
            }
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t gw[4] = {gq[i].x, gq[i].y, gq[i].z, gq[i].w}, uw[4] = {uq[i].x, uq[i].y, uq[i].z, uq[i].w};
              uint32_t og[4], ou[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float2 g = unpack_bf16(gw[k]), u = unpack_bf16(uw[k]);
                // dact goes through bf16 like the stand-alone path (GEMM output rounded, then the SwiGLU backward)
                const float2 d = unpack_bf16(pack_bf16(__uint_as_float(r[8 * i + 2 * k]),
                                                       __uint_as_float(r[8 * i + 2 * k + 1])));
                float t0, t1;
                asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(0.5f * g.x));
                asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(0.5f * g.y));
                const float s0 = fmaf(0.5f, t0, 0.5f), s1 = fmaf(0.5f, t1, 0.5f);
                og[k] = pack_bf16(d.x * u.x * s0 * (1.f + g.x * (1.f - s0)), d.y * u.y * s1 * (1.f + g.y * (1.f - s1)));
                ou[k] = pack_bf16(d.x * g.x * s0, d.y * g.y * s1);
              }
              const int idx = ((cc * 4 + i) ^ (lane & 7)) << 4;
              *reinterpret_cast<uint4*>(row_g + idx) = make_uint4(og[0], og[1], og[2], og[3]);
              *reinterpret_cast<uint4*>(row_u + idx) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
            }
          }
          __syncwarp();
          // column sums of the two finished slabs (what is stored, bf16-rounded): lane = (8-column unit, 8-row group)
          {
            const int unit = lane & 7, rg = lane >> 3;
            float* dst = args.colsum +
                         (static_cast<long long>(t / tiles_n) * 8 + cta * 4 + q) * 2 * F + c0 + unit * 8;
#pragma unroll
            for (int which = 0; which < 2; ++which) {
              const uint8_t* slab = slabs + (2 * grp + which) * 4096;
              float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
              for (int rr = 0; rr < 8; ++rr) {
                const int srow = rg * 8 + rr;
                const uint4 v = *reinterpret_cast<const uint4*>(slab + srow * 128 + ((unit ^ (srow & 7)) << 4));
                const uint32_t vw[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float2 f = unpack_bf16(vw[k]);
                  acc[2 * k] += f.x;
                  acc[2 * k + 1] += f.y;
                }
              }
#pragma unroll
              for (int e = 0; e < 8; ++e) {   // fixed order: row groups 0+1, 2+3, then the pairs
                acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 8);
                acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
              }
              if (rg == 0) {
                *reinterpret_cast<float4*>(dst + which * F) = make_float4(acc[0], acc[1], acc[2], acc[3]);
                *reinterpret_cast<float4*>(dst + which * F + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
              }
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_3d(&tmOut, slabs + (2 * grp) * 4096, c0, m0 + q * 32, 0);
            tma_store_3d(&tmOut, slabs + (2 * grp + 1) * 4096, F + c0, m0 + q * 32, 0);
            tma_store_commit();
          }
          __syncwarp();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(&tmem_empty[as]);
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
      if (lane == 0) tma_store_wait<0>();
    } else if constexpr (SWIGLU) {
      // warps 2-5: hidden-unit chunks 0, 1 of the tile (gate columns [0, 64) + up columns [128, 192));
      // warps 6-9: chunks 2, 3.  Three slabs per warp: gate, up, act (32 rows x 64 bf16 columns each).
      const int half = (warp - 2) >> 2;
      uint8_t* slab_g = sOut + (warp - 2) * 3 * 4096;
      uint8_t* slab_u = slab_g + 4096;
      uint8_t* slab_a = slab_u + 4096;
      int as = 0;
      uint32_t aphase = 0;
      for (int w = cluster_id; w < num_work; w += n_clusters) {
        const int t = w;   // splits == 1
        const int m0 = (t / tiles_n) * (2 * BM) + cta * BM;
        const int j0 = (t % tiles_n) * BNH;   // first hidden unit of the tile
        if (lane == 0) tma_store_wait_read<0>();   // the three slabs of the previous tile have been read out
        __syncwarp();
        mbar_wait(&tmem_full[as], aphase);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + as * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          const int c = half * 2 + cc;   // chunk of 32 hidden units inside the tile
          uint32_t rg[32], ru[32];
          tmem_ld32(t_addr + c * 32, rg);
          tmem_ld32(t_addr + BNH + c * 32, ru);
          float4 bg[8], bu[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            bg[i] = args.bias ? __ldg(reinterpret_cast<const float4*>(args.bias + j0 + c * 32) + i)
                              : make_float4(0.f, 0.f, 0.f, 0.f);
            bu[i] = args.bias ? __ldg(reinterpret_cast<const float4*>(args.bias + args.F + j0 + c * 32) + i)
                              : make_float4(0.f, 0.f, 0.f, 0.f);
          }
          tmem_ld_wait();
          uint8_t* row_g = slab_g + lane * 128;
          uint8_t* row_u = slab_u + lane * 128;
          uint8_t* row_a = slab_a + lane * 128;
#pragma unroll
          for (int i = 0; i < 4; ++i) {   // 8 hidden units at a time -> one 16-byte unit of each slab row
            const float bgv[8] = {bg[2 * i].x, bg[2 * i].y, bg[2 * i].z, bg[2 * i].w,
                                  bg[2 * i + 1].x, bg[2 * i + 1].y, bg[2 * i + 1].z, bg[2 * i + 1].w};
            const float buv[8] = {bu[2 * i].x, bu[2 * i].y, bu[2 * i].z, bu[2 * i].w,
                                  bu[2 * i + 1].x, bu[2 * i + 1].y, bu[2 * i + 1].z, bu[2 * i + 1].w};
            uint32_t pg[4], pu[4], pa[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float g0 = __uint_as_float(rg[8 * i + 2 * k]) + bgv[2 * k];
              const float g1 = __uint_as_float(rg[8 * i + 2 * k + 1]) + bgv[2 * k + 1];
              const float u0 = __uint_as_float(ru[8 * i + 2 * k]) + buv[2 * k];
              const float u1 = __uint_as_float(ru[8 * i + 2 * k + 1]) + buv[2 * k + 1];
              pg[k] = pack_bf16(g0, g1);
              pu[k] = pack_bf16(u0, u1);
              // the activation is computed from the ROUNDED gate / up, i.e. from exactly what the backward will read
              const float2 gr = unpack_bf16(pg[k]), ur = unpack_bf16(pu[k]);
              float t0, t1;
              asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(0.5f * gr.x));
              asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(0.5f * gr.y));
              pa[k] = pack_bf16(gr.x * fmaf(0.5f, t0, 0.5f) * ur.x, gr.y * fmaf(0.5f, t1, 0.5f) * ur.y);
            }
            const int idx = ((cc * 4 + i) ^ (lane & 7)) << 4;   // 128B swizzle of the slab row
            *reinterpret_cast<uint4*>(row_g + idx) = make_uint4(pg[0], pg[1], pg[2], pg[3]);
            *reinterpret_cast<uint4*>(row_u + idx) = make_uint4(pu[0], pu[1], pu[2], pu[3]);
            *reinterpret_cast<uint4*>(row_a + idx) = make_uint4(pa[0], pa[1], pa[2], pa[3]);
          }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive_leader(&tmem_empty[as]);
          const int col = j0 + half * 64, row = m0 + q * 32;
          tma_store_3d(&tmOut, slab_g, col, row, 0);
          tma_store_3d(&tmOut, slab_u, args.F + col, row, 0);
          tma_store_3d(&tmAct, slab_a, col, row, 0);
          tma_store_commit();
        }
        __syncwarp();
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
      if (lane == 0) tma_store_wait<0>();
    } else {
    // EW = 8: warps 2-5 take columns [0, 128) of the tile, warps 6-9 take [128, 256)
    const int c_begin = (EW == 8) ? ((warp - 2) >> 2) * (BN / 64) : 0;
    const int c_end = (EW == 8) ? c_begin + BN / 64 : BN / 32;
    uint8_t* stage = sOut + (warp - 2) * kOutBufs * 4096;
    int sbuf = 0;
    int as = 0;
    uint32_t aphase = 0;
    constexpr bool kBf16 = sizeof(OutT) == 2;
    // The residual is read by row (every thread its own 128-byte segments), which leaves few bytes in flight per SM:
    // its lines are pulled into L2 one whole tile ahead (prefetch, no registers), and the register loads run one
    // 32-column chunk ahead of the TMEM read they are added to.
    auto residual_row = [&](int w) -> const float* {
      if (!args.residual || w >= num_work) return nullptr;
      const int t = w % tiles_mn;
      const int row = (t / tiles_n) * (2 * BM) + cta * BM + q * 32 + lane;
      return row < args.M ? args.residual + static_cast<long long>(row) * args.ldr + (t % tiles_n) * BN : nullptr;
    };
    auto prefetch_row = [&](int w) {
      const float* p = residual_row(w);
      if (!p) return;
      const int n0 = (w % tiles_mn % tiles_n) * BN;
#pragma unroll
      for (int j = 0; j < BN / 32; ++j)
        if (n0 + j * 32 < args.N) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + j * 32));
    };
    prefetch_row(cluster_id);
    for (int w = cluster_id; w < num_work; w += n_clusters) {
      const int split = w / tiles_mn;
      const int t = w - split * tiles_mn;
      const int m0 = (t / tiles_n) * (2 * BM) + cta * BM;
      const int n0 = (t % tiles_n) * BN;
      const int row = m0 + q * 32 + lane;
      const float* rrow = (args.residual && row < args.M)
                              ? args.residual + static_cast<long long>(row) * args.ldr + n0
                              : nullptr;
      float4 rnext[8], bnext[8];
      auto load_res = [&](int c) {
        const int valid = args.N - (n0 + c * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          rnext[j] = (rrow && c < c_end && j * 4 + 4 <= valid)
                         ? *reinterpret_cast<const float4*>(rrow + c * 32 + 4 * j)
                         : make_float4(0.f, 0.f, 0.f, 0.f);
      };
      // the bias segment of a chunk is a first touch (an L2 round trip) for every tile: fetch it one chunk ahead too
      auto load_bias = [&](int c) {
        const int valid = args.N - (n0 + c * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          bnext[j] = (args.bias && c < c_end && j * 4 + 4 <= valid)
                         ? __ldg(reinterpret_cast<const float4*>(args.bias + n0 + c * 32) + j)
                         : make_float4(0.f, 0.f, 0.f, 0.f);
      };
      load_res(c_begin);
      load_bias(c_begin);
      prefetch_row(w + n_clusters);
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + as * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        uint32_t r[32];
        tmem_ld32(t_addr + c * 32, r);
        float4 bcur[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) bcur[j] = bnext[j];
        if (args.bias) load_bias(c + 1);
        tmem_ld_wait();
        const int ncols_valid = args.N - (n0 + c * 32);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if (ncols_valid > 0) {
          if (args.bias) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {   // (zeros where the column group is out of range)
              v[4 * j] += bcur[j].x;
              v[4 * j + 1] += bcur[j].y;
              v[4 * j + 2] += bcur[j].z;
              v[4 * j + 3] += bcur[j].w;
            }
          }
          if (rrow) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {   // (zeros where the column group is out of range)
              v[4 * j] += rnext[j].x;
              v[4 * j + 1] += rnext[j].y;
              v[4 * j + 2] += rnext[j].z;
              v[4 * j + 3] += rnext[j].w;
            }
          }
        }
        if (rrow) load_res(c + 1);   // (registers are tight: the residual runs the rest of this chunk ahead, not a full one)
        uint8_t* slab = stage + sbuf * 4096 + lane * 128;
        if constexpr (kBf16) {
          // two 32-column chunks share one 64-column (128 B) slab row
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int idx = (c & 1) * 4 + j;
            uint4 o;
            o.x = pack_bf16(v[8 * j + 0], v[8 * j + 1]);
            o.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
            o.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]);
            o.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
            *reinterpret_cast<uint4*>(slab + ((idx ^ (lane & 7)) << 4)) = o;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(slab + ((j ^ (lane & 7)) << 4)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
        if (!kBf16 || (c & 1)) {
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            const int col0 = n0 + (kBf16 ? (c - 1) : c) * 32;
            if (col0 < args.N) tma_store_3d(&tmOut, stage + sbuf * 4096, col0, m0 + q * 32, split);
            tma_store_commit();
            tma_store_wait_read<kOutBufs - 1>();  // the slab written next has been read out by its previous store
          }
          __syncwarp();
          sbuf = (sbuf + 1) % kOutBufs;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(&tmem_empty[as]);
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
    if (lane == 0) tma_store_wait<0>();  // all output tiles have landed before the CTA exits
    }
  }

  tc_fence_before();
  cluster_sync();  // no CTA frees TMEM / exits while its peer may still touch its barriers or shared memory
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, 2 * BN);
  }
}

static PFN_cuTensorMapEncodeTiled_v12000 encode_fn2() {
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

// 3-D store map over out[splits][M][ldo] (N valid columns): box = 128 bytes of a row x 32 rows x 1 slab, 128B swizzle
static int make_out_tmap(CUtensorMap* tm, void* out, bool fp32, int M, int N, long long ldo, int splits,
                         long long split_stride) {
  auto fn = encode_fn2();
  if (!fn) return -1;
  const cuuint64_t es = fp32 ? 4 : 2;
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(N), static_cast<cuuint64_t>(M), static_cast<cuuint64_t>(splits)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ldo) * es,
                           (splits > 1 ? static_cast<cuuint64_t>(split_stride) : static_cast<cuuint64_t>(M) * ldo) * es};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(128 / es), 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, out, dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

template <bool A_MN, bool B_MN, typename OutT>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const Args& args, cudaStream_t stream) {
  CUtensorMap tmOut;
  {
    const int rc = make_out_tmap(&tmOut, args.out, sizeof(OutT) == 4, args.M, args.N, args.ldo, args.splits,
                                 args.split_stride);
    if (rc) return rc;
  }
  const int work = ((args.M + 2 * BM - 1) / (2 * BM)) * ((args.N + BN - 1) / BN) * args.splits;
  int clusters = num_sms() / 2;
  if (work < clusters) clusters = work;
  // short reductions per tile are epilogue-paced: use the 8-epilogue-warp shape (DPFS_GEMM_EPI8=0/1 forces a shape)
  static int force = -2;
  if (force == -2) {
    const char* e = getenv("DPFS_GEMM_EPI8");
    force = e ? atoi(e) : -1;
  }
  const int k_per_tile = args.kb_per_split * BK < args.K ? args.kb_per_split * BK : args.K;
  const bool epi8 = force >= 0 ? force != 0 : k_per_tile <= 1024;
  if (epi8) {
    auto kfn = gemm2sm_kernel<A_MN, B_MN, OutT, 8>;
    DPFS_ENSURE_DYN_SMEM(kfn, Cfg<8>::kSmem);
    kfn<<<(DPFS_COUNT(), 2 * clusters), Cfg<8>::kThreads, Cfg<8>::kSmem, stream>>>(tmA, tmB, tmOut, args, tmOut);
  } else {
    auto kfn = gemm2sm_kernel<A_MN, B_MN, OutT, 4>;
    DPFS_ENSURE_DYN_SMEM(kfn, Cfg<4>::kSmem);
    kfn<<<(DPFS_COUNT(), 2 * clusters), Cfg<4>::kThreads, Cfg<4>::kSmem, stream>>>(tmA, tmB, tmOut, args, tmOut);
  }
  return static_cast<int>(cudaGetLastError());
}

template <typename OutT>
static int dispatch(bool a_mn, bool b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB, const Args& args,
                    cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch<false, false, OutT>(tmA, tmB, args, stream);
  if (!a_mn && b_mn) return launch<false, true, OutT>(tmA, tmB, args, stream);
  if (a_mn && b_mn) return launch<true, true, OutT>(tmA, tmB, args, stream);
  return launch<true, false, OutT>(tmA, tmB, args, stream);
}

// fp32 split-K partial sums -> final output (same contract as the 1-SM kernel's reducer)
template <typename OutT>
__global__ void splitk_reduce2_kernel(const float* __restrict__ ws, OutT* __restrict__ out, long long ldo,
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
    const int r = static_cast<int>(e / N), c = static_cast<int>(e % N);
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

}  // namespace sm2
}  // namespace dpfs

using namespace dpfs;

// Same contract as dpfs_gemm_bf16 (without the N-tile choice): 256x256 tiles on CTA pairs.
extern "C" int dpfs_gemm2sm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb,
                                 int b_mn_major, void* out, long long ldo, int out_fp32, const float* bias,
                                 const float* residual, long long ldr, int M, int N, int K, int splits,
                                 float* workspace, cudaStream_t stream) {
  using namespace dpfs::sm2;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((N % 8) != 0 || (lda % 8) != 0 || (ldb % 8) != 0 || (ldo % (out_fp32 ? 4 : 8)) != 0) return -10;
  if (splits < 1) splits = 1;
  const int kb_total = (K + BK - 1) / BK;
  if (splits > kb_total) splits = kb_total;
  const int kb_per_split = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per_split - 1) / kb_per_split;
  if (splits > 1 && workspace == nullptr) return -11;
  CUtensorMap tmA, tmB;
  int rc = !a_mn_major ? make_tmap_bf16_2d(&tmA, A, K, M, lda, 64, BM) : make_tmap_bf16_2d(&tmA, A, M, K, lda, 64, BK);
  if (rc) return rc;
  rc = !b_mn_major ? make_tmap_bf16_2d(&tmB, B, K, N, ldb, 64, BNH) : make_tmap_bf16_2d(&tmB, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  Args args;
  args.M = M;
  args.N = N;
  args.K = K;
  args.splits = splits;
  args.kb_per_split = kb_per_split;
  args.bias = bias;
  args.residual = residual;
  args.ldr = ldr;
  const bool direct = splits == 1;
  if (direct) {
    args.out = out;
    args.ldo = ldo;
    args.split_stride = 0;
  } else {
    if (residual) return -12;
    args.out = workspace;
    args.ldo = N;
    args.split_stride = static_cast<long long>(M) * N;
    args.bias = nullptr;
    args.residual = nullptr;
  }
  const bool kernel_fp32 = direct ? (out_fp32 != 0) : true;
  rc = kernel_fp32 ? dispatch<float>(a_mn_major, b_mn_major, tmA, tmB, args, stream)
                   : dispatch<__nv_bfloat16>(a_mn_major, b_mn_major, tmA, tmB, args, stream);
  if (rc) return rc;
  if (!direct) {
    const long long total4 = static_cast<long long>(M) * N / 4;
    int blocks = static_cast<int>((total4 + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    if (out_fp32)
      splitk_reduce2_kernel<float><<<(DPFS_COUNT(), blocks), 256, 0, stream>>>(
          workspace, reinterpret_cast<float*>(out), ldo, bias, M, N, splits);
    else
      splitk_reduce2_kernel<__nv_bfloat16><<<(DPFS_COUNT(), blocks), 256, 0, stream>>>(
          workspace, reinterpret_cast<__nv_bfloat16*>(out), ldo, bias, M, N, splits);
    rc = static_cast<int>(cudaGetLastError());
  }
  return rc;
}

// Fused gate|up projection + SwiGLU (forward): gu[M][2F] = x W^T + b (bf16), act[M][F] = silu(gu[:, :F]) * gu[:, F:].
// W is the concatenated [2F][ldb] weight (gate rows, then up rows), K-major; F % 128 == 0.
extern "C" int dpfs_gemm2sm_swiglu_bf16(const void* A, long long lda, const void* B, long long ldb, void* gu, void* act,
                                        const float* bias, int M, int F, int K, cudaStream_t stream) {
  using namespace dpfs::sm2;
  if (M <= 0 || F <= 0 || K <= 0) return 0;
  if (F % BNH != 0 || (lda % 8) != 0 || (ldb % 8) != 0) return -10;
  CUtensorMap tmA, tmB, tmOut, tmAct;
  int rc = make_tmap_bf16_2d(&tmA, A, K, M, lda, 64, BM);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&tmB, B, K, 2 * F, ldb, 64, BNH);
  if (rc) return rc;
  rc = make_out_tmap(&tmOut, gu, false, M, 2 * F, 2LL * F, 1, 0);
  if (rc) return rc;
  rc = make_out_tmap(&tmAct, act, false, M, F, F, 1, 0);
  if (rc) return rc;
  Args args;
  args.M = M;
  args.N = 2 * F;
  args.K = K;
  args.splits = 1;
  args.kb_per_split = (K + BK - 1) / BK;
  args.out = gu;
  args.ldo = 2LL * F;
  args.split_stride = 0;
  args.bias = bias;
  args.residual = nullptr;
  args.ldr = 0;
  args.F = F;
  args.gu = nullptr;
  args.colsum = nullptr;
  const int work = ((M + 2 * BM - 1) / (2 * BM)) * (F / BNH);
  int clusters = num_sms() / 2;
  if (work < clusters) clusters = work;
  auto kfn = gemm2sm_kernel<false, false, __nv_bfloat16, 8, 1>;
  DPFS_ENSURE_DYN_SMEM(kfn, (Cfg<8, 1>::kSmem));
  kfn<<<(DPFS_COUNT(), 2 * clusters), Cfg<8, 1>::kThreads, Cfg<8, 1>::kSmem, stream>>>(tmA, tmB, tmOut, args,
                                                                                              tmAct);
  return static_cast<int>(cudaGetLastError());
}

namespace dpfs { int reduce_partials_rows(const float* partial, float* out, int nrows, int D, cudaStream_t stream); }

// rows of the column-sum scratch of dpfs_gemm2sm_swiglu_bwd_bf16 ([rows][2F] floats)
extern "C" int dpfs_gemm2sm_swiglu_bwd_rows(int M) { return ((M + 255) / 256) * 8; }

// Fused down_proj dgrad + SwiGLU backward: dgu[M][2F] = swiglu'(dy W_down, gu) (bf16) and bias_grad[2F] = column sums
// of dgu.  dy: bf16 [M][ldy] (K = d contiguous), W: bf16 [d][ldw] (the down_proj weight as stored: [odim = d][F]),
// gu: the saved gate|up activation [M][2F].  F % 256 == 0.  scratch: dpfs_gemm2sm_swiglu_bwd_rows(M) x 2F floats.
extern "C" int dpfs_gemm2sm_swiglu_bwd_bf16(const void* dy, long long ldy, const void* W, long long ldw, const void* gu,
                                            void* dgu, float* bias_grad, float* scratch, int M, int F, int K,
                                            cudaStream_t stream) {
  using namespace dpfs::sm2;
  if (M <= 0 || F <= 0 || K <= 0) return 0;
  if (F % BN != 0 || (ldy % 8) != 0 || (ldw % 8) != 0 || !gu || !scratch) return -10;
  CUtensorMap tmA, tmB, tmOut;
  int rc = make_tmap_bf16_2d(&tmA, dy, K, M, ldy, 64, BM);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&tmB, W, F, K, ldw, 64, BK);   // MN-major B: [K][N] with N contiguous
  if (rc) return rc;
  rc = make_out_tmap(&tmOut, dgu, false, M, 2 * F, 2LL * F, 1, 0);
  if (rc) return rc;
  CUtensorMap tmGu;   // same geometry over the saved gate|up activation: its slabs are loaded, not stored
  rc = make_out_tmap(&tmGu, const_cast<void*>(gu), false, M, 2 * F, 2LL * F, 1, 0);
  if (rc) return rc;
  Args args;
  args.M = M;
  args.N = F;
  args.K = K;
  args.splits = 1;
  args.kb_per_split = (K + BK - 1) / BK;
  args.out = dgu;
  args.ldo = 2LL * F;
  args.split_stride = 0;
  args.bias = nullptr;
  args.residual = nullptr;
  args.ldr = 0;
  args.F = F;
  args.gu = static_cast<const __nv_bfloat16*>(gu);
  args.colsum = scratch;
  const int work = ((M + 2 * BM - 1) / (2 * BM)) * (F / BN);
  int clusters = num_sms() / 2;
  if (work < clusters) clusters = work;
  auto kfn = gemm2sm_kernel<false, true, __nv_bfloat16, 8, 2>;
  DPFS_ENSURE_DYN_SMEM(kfn, (Cfg<8, 2>::kSmem));
  kfn<<<(DPFS_COUNT(), 2 * clusters), Cfg<8, 2>::kThreads, Cfg<8, 2>::kSmem, stream>>>(tmA, tmB, tmOut, args, tmGu);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  if (!bias_grad) return 0;
  return reduce_partials_rows(scratch, bias_grad, dpfs_gemm2sm_swiglu_bwd_rows(M), 2 * F, stream);
}
