// Causal flash attention forward / backward on tcgen05 tensor cores (sm_100a).
//
// Replaces the materialised attention of the reference (models/model.py:68-78: QK^T / sqrt(hd), a per-call
// fp32 [b,1,t,t] triu mask, masked_fill(-10000), softmax, PV, plus two einops head rearranges): scores and
// probabilities never leave the SM (S and dP live in TMEM, P and dS in shared memory), the causal mask is implicit
// (tiles above the diagonal are skipped), and the head split / merge is folded into the TMA tensor maps over the
// packed [b*t, 3*Dl] QKV buffer — no transposes, no [b,H,t,t] tensors.
//
// Layouts
//   qkv : bf16 [b*t][3*Dl]   q | k | v segments, each H heads of hd (RoPE already applied to q and k)
//   out : bf16 [b*t][Dl]
//   lse : fp32 [b][H][t]     log-sum-exp of the scaled scores (natural log)
//   dqkv: bf16 [b*t][3*Dl]   (backward; dq is accumulated in fp32 [b*t][Dl] across kv tiles and converted after)
//
// Forward CTA  = one (batch, head, 128-row q tile); loops over kv tiles up to the diagonal.
// Backward CTA = one (batch, head, 128-row kv tile); loops over q tiles from the diagonal down.
// Warp roles (192 threads): warp 0 TMA producer, warp 1 tcgen05.mma issuer (+ TMEM owner), warps 2-5 softmax /
// gradient math, one thread per tile row (TMEM lane).
#include "common.cuh"
#include "dpfs_b200.h"

#include <cudaTypedefs.h>
#include <stdlib.h>

#include <type_traits>

namespace dpfs {

int num_sms();

constexpr int kAttnFwdThreads = 320;  // warp 0 TMA, warp 1 MMA, warps 2-9: two softmax warps per SM sub-partition
constexpr int kAttnBwdThreads = 320;  // warp 0 TMA, warp 1 MMA, warps 2-9: two groups of 4 gradient-math warps
constexpr int kTile = 128;  // q rows and kv rows per tile
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---- work-item order of the persistent kernels ------------------------------------------------------------------------
// Items are sorted heavy-first.  Round-robin (CTA b takes items b, b + G, b + 2G, ...) always hands CTA 0 the heaviest
// item of every round; with few rounds (TP = 8: 256 items on 148 CTAs) the makespan is 12 tile-steps where 8 would do.
// The "snake" order reverses every other round -- CTA b takes position G-1-b of the odd rounds -- so the CTAs that got
// the heaviest items of one round get the lightest of the next (LPT-like).  Only the last round can be partial, and a
// CTA whose position there is past the end has no later item either, so "w >= n_items" still ends the loop.
__host__ __device__ __forceinline__ int snake_item(int k, int bid, int G) {
  return k * G + ((k & 1) ? (G - 1 - bid) : bid);
}
// k-th item of this CTA given the previous one (SNAKE = false: the plain round-robin of the validated kernels)
template <bool SNAKE>
__device__ __forceinline__ int item_w(int k, int w_prev) {
  if constexpr (SNAKE) return snake_item(k, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x));
  else return w_prev + static_cast<int>(gridDim.x);
}

// the item sequence of one CTA as a loop variable: SNAKE = false is literally "w = blockIdx.x; w += gridDim.x"
template <bool SNAKE>
struct ItemWalk {
  int w, k;
  __device__ __forceinline__ ItemWalk() : w(static_cast<int>(blockIdx.x)), k(0) {}
  __device__ __forceinline__ void next() {
    if constexpr (SNAKE) {
      ++k;
      w = snake_item(k, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x));
    } else {
      w += gridDim.x;
    }
  }
  __device__ __forceinline__ int peek() const {   // the item after the current one (may be >= n_items)
    if constexpr (SNAKE) return snake_item(k + 1, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x));
    else return w + gridDim.x;
  }
};

// K-major SW128 operand [rows][cols] made of 64-column atoms (each atom = rows x 128 B, atoms `atom_bytes` apart):
// descriptor for the 16-wide K slice kk.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int kk, uint32_t atom_bytes) {
  return make_smem_desc_sw128(base + (kk >> 2) * atom_bytes + (kk & 3) * 32, 0, 1024);
}
// MN-major SW128 operand stored as [k rows][MN cols] in 64-column atoms: descriptor for k rows [16*kk, 16*kk+16).
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int kk, uint32_t atom_bytes) {
  return make_smem_desc_sw128(base + kk * 2048, atom_bytes, 1024);
}

// write 32 consecutive bf16 (columns [c0, c0+32) of row r) into a K-major SW128 [128][128] tile (two 64-col atoms)
__device__ __forceinline__ void st_tile_row32(uint8_t* tile, int r, int c0, const uint32_t (&packed)[16]) {
  uint8_t* atom = tile + (c0 >> 6) * (kTile * 128) + r * 128;
  const int cc0 = (c0 & 63) >> 3;  // first 16-byte chunk inside the atom row
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int phys = (cc0 + q) ^ (r & 7);
    *reinterpret_cast<uint4*>(atom + (phys << 4)) =
        make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
  }
}

// ================================================================================================= forward
template <int HD>
struct FwdCfg {
  static constexpr int kStages = 2;
  static constexpr int kQBytes = kTile * HD * 2;
  static constexpr int kKVBytes = kTile * HD * 2;  // one K or V tile
  static constexpr int kPBytes = kTile * kTile * 2;
  static constexpr int kXchgBytes = 2 * 2 * kTile * 4;  // row-max / row-sum exchange between the two column halves
  static constexpr int kSmem = kQBytes + kStages * 2 * kKVBytes + kPBytes + 256 + kXchgBytes + 1024;
  // One CTA per SM; S is double buffered in TMEM [0,128),[128,256) and O sits at [256,256+HD), so QK^T of tile j+1
  // overlaps the softmax of tile j inside the CTA.  (Measured: two CTAs/SM with a single S buffer were 30% SLOWER for
  // hd = 64 — 193 vs 149 us on the flagship shape; the single-buffer path is kept for experiments.)
  static constexpr int kSBufs = 2;
  static constexpr int kTmemCols = 512;
  static constexpr int kOCol = 256;
  static constexpr int kCtasPerSm = 1;
};

template <int HD>
__global__ void __launch_bounds__(kAttnFwdThreads, FwdCfg<HD>::kCtasPerSm)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, __nv_bfloat16* __restrict__ out,
                float* __restrict__ lse, int T, int H, int Dl, float scale_log2) {
  using C = FwdCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + C::kQBytes;                    // [stage]
  uint8_t* sV = sK + C::kStages * C::kKVBytes;      // [stage]
  uint8_t* sP = sV + C::kStages * C::kKVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + C::kPBytes);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;                       // [2]
  uint64_t* v_full = bars + 3;                       // [2]
  uint64_t* kv_empty = bars + 5;                     // [2]
  uint64_t* s_full = bars + 7;                       // [2] (S is double buffered in TMEM)
  uint64_t* p_full = bars + 9;
  uint64_t* o_done = bars + 10;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);
  float* s_xchg = reinterpret_cast<float*>(bars + 12);  // [2 parities][2 halves][128 rows] row-max / row-sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (T + kTile - 1) / kTile;
  const int qt = nq - 1 - blockIdx.x;  // heavy tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = qt * kTile;
  const int n_kv = qt + 1;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(&s_full[0], 1);
    mbar_init(&s_full[1], 1);
    mbar_init(p_full, 8);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + C::kOCol;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, C::kQBytes);
#pragma unroll
      for (int a = 0; a < HD / 64; ++a) tma_load_3d(sQ + a * (kTile * 128), &tm_qkv, q_full, h * HD + a * 64, q0, b);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&k_full[st], C::kKVBytes);
#pragma unroll
        for (int a = 0; a < HD / 64; ++a)
          tma_load_3d(sK + st * C::kKVBytes + a * (kTile * 128), &tm_qkv, &k_full[st], Dl + h * HD + a * 64,
                      j * kTile, b);
        mbar_expect_tx(&v_full[st], C::kKVBytes);
#pragma unroll
        for (int a = 0; a < HD / 64; ++a)
          tma_load_3d(sV + st * C::kKVBytes + a * (kTile * 128), &tm_qkv, &v_full[st], 2 * Dl + h * HD + a * 64,
                      j * kTile, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(kTile, HD, false, true);
      const uint32_t q_base = smem_u32(sQ), p_base = smem_u32(sP);
      mbar_wait(q_full, 0);
      // S(0); afterwards S(j+1) is issued before PV(j) so that QK^T overlaps the softmax of tile j
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk)
        umma_bf16_ss(tS, desc_kmajor(q_base, kk, kTile * 128), desc_kmajor(smem_u32(sK), kk, kTile * 128), idesc_s,
                     kk > 0);
      umma_commit(&s_full[0]);
      auto issue_s = [&](int jn) {
        const int sn = jn & 1;
        mbar_wait(&k_full[sn], (jn >> 1) & 1);
        tc_fence_after();
        const uint32_t k_base = smem_u32(sK + sn * C::kKVBytes);
        const uint32_t s_buf = tS + (jn % C::kSBufs) * 128;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk)
          umma_bf16_ss(s_buf, desc_kmajor(q_base, kk, kTile * 128), desc_kmajor(k_base, kk, kTile * 128), idesc_s,
                       kk > 0);
        umma_commit(&s_full[jn % C::kSBufs]);
      };
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        // double-buffered S: QK^T of tile j+1 goes out before PV(j) (its buffer was drained before p_full(j-1));
        // single-buffered S: it has to wait until the softmax of tile j has read S, i.e. until p_full(j)
        if (C::kSBufs == 2 && j + 1 < n_kv) issue_s(j + 1);
        mbar_wait(p_full, j & 1);
        mbar_wait(&v_full[st], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t v_base = smem_u32(sV + st * C::kKVBytes);
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)
          umma_bf16_ss(tO, desc_kmajor(p_base, kk, kTile * 128), desc_mnmajor(v_base, kk, kTile * 128), idesc_o,
                       (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit(o_done);
        umma_commit(&kv_empty[st]);
        if (C::kSBufs == 1 && j + 1 < n_kv) issue_s(j + 1);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax warps: TWO threads per q row
    // Warp group `half` (warps 2-5 / 6-9) owns kv columns [64*half, 64*half+64) of every row; inside a group warp q
    // owns TMEM lanes (rows) [32q, 32q+32).  Two warps per SM sub-partition double the issue rate of the exp / FMA
    // chains (the forward kernel is bound by exactly that).  The two halves of a row agree on the running max through
    // a double-buffered smem slot + one named barrier per tile; each keeps a partial row sum, combined at the end.
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    constexpr int kOChunks = HD / 64;  // 32-column chunks of O per half
    float m_run = -INFINITY, l_run = 0.f;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(&s_full[j % C::kSBufs], (j / C::kSBufs) & 1);
      tc_fence_after();
      float s[64];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t tmp[32];
        tmem_ld32(tS + (j % C::kSBufs) * 128 + lane_off + (2 * half + c) * 32, tmp);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(tmp[i]) * scale_log2;
      }
      if (j == qt) {  // diagonal tile: mask kv > q
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (half * 64 + i > r) s[i] = -INFINITY;
      }
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // 4 independent chains instead of one of 64
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        m4[0] = fmaxf(m4[0], s[i]);
        m4[1] = fmaxf(m4[1], s[i + 1]);
        m4[2] = fmaxf(m4[2], s[i + 2]);
        m4[3] = fmaxf(m4[3], s[i + 3]);
      }
      const float mloc = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      float* xb = s_xchg + (j & 1) * 256;
      xb[half * 128 + r] = mloc;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const float mcand = fmaxf(mloc, xb[(half ^ 1) * 128 + r]);   // column 0 is never masked: finite
      // Lazy rescaling: the reference max only moves when the new row max exceeds it by more than 8 (log2 domain), so
      // P stays below 2^8 and the TMEM round trip that rescales O is skipped on almost every tile.  O and l are
      // accumulated against the same reference, so the final O / l and the LSE are exact.
      const float mx = (mcand > m_run + 8.f) ? mcand : m_run;
      const float alpha = fast_exp2(m_run - mx);
      m_run = mx;
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t packed[32];
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        const float p0 = fast_exp2(s[i] - mx), p1 = fast_exp2(s[i + 1] - mx);
        const float p2 = fast_exp2(s[i + 2] - mx), p3 = fast_exp2(s[i + 3] - mx);
        s4[0] += p0;
        s4[1] += p1;
        s4[2] += p2;
        s4[3] += p3;
        packed[i >> 1] = pack_bf16(p0, p1);
        packed[(i >> 1) + 1] = pack_bf16(p2, p3);
      }
      const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
      l_run = l_run * alpha + sum;
      if (j > 0) {
        // previous PV must be complete before O is rescaled and P is overwritten
        mbar_wait(o_done, (j - 1) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
          for (int c = half * kOChunks; c < (half + 1) * kOChunks; ++c) {
            uint32_t o[32];
            tmem_ld32(tO + lane_off + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(tO + lane_off + c * 32, o);
          }
          tmem_st_wait();
        }
      }
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t chunk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) chunk[i] = packed[c * 16 + i];
        st_tile_row32(sP, r, half * 64 + c * 32, chunk);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    // epilogue: combine the two partial row sums, O / l -> bf16, lse
    float* xb = s_xchg + (n_kv & 1) * 256;
    xb[half * 128 + r] = l_run;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const float l_tot = l_run + xb[(half ^ 1) * 128 + r];
    mbar_wait(o_done, (n_kv - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.f / l_tot;
    const bool valid = (q0 + r) < T;
    __nv_bfloat16* orow = out + (static_cast<long long>(b) * T + q0 + r) * Dl + h * HD;
#pragma unroll
    for (int c = half * kOChunks; c < (half + 1) * kOChunks; ++c) {
      uint32_t o[32];
      tmem_ld32(tO + lane_off + c * 32, o);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int v8 = 0; v8 < 4; ++v8) {
          uint4 w;
          w.x = pack_bf16(__uint_as_float(o[8 * v8 + 0]) * inv_l, __uint_as_float(o[8 * v8 + 1]) * inv_l);
          w.y = pack_bf16(__uint_as_float(o[8 * v8 + 2]) * inv_l, __uint_as_float(o[8 * v8 + 3]) * inv_l);
          w.z = pack_bf16(__uint_as_float(o[8 * v8 + 4]) * inv_l, __uint_as_float(o[8 * v8 + 5]) * inv_l);
          w.w = pack_bf16(__uint_as_float(o[8 * v8 + 6]) * inv_l, __uint_as_float(o[8 * v8 + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + v8 * 8) = w;
        }
      }
    }
    if (half == 0 && valid && lse) lse[(static_cast<long long>(b) * H + h) * T + q0 + r] = (m_run + log2f(l_tot)) * kLn2;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ================================================================================================= forward, persistent
// Same math as attn_fwd_kernel, but ONE CTA per SM walks a heavy-first list of (q tile, head, batch) work items.  The
// per-CTA fixed costs of the one-item-per-CTA kernel (TMEM allocation, barrier setup, the cold Q / K / V loads and the
// pipeline fill and drain — ~14 CTAs ran back to back on every SM with nothing overlapping between them) are paid once;
// the loader and the MMA warp run ahead into the next item while the softmax warps finish the current one.
template <int HD>
struct FwdPCfg {
  static constexpr int kStages = 2;
  static constexpr int kQBufs = (HD == 64) ? 2 : 1;  // hd = 128: 2 x 32 KB of Q would not fit next to K / V / P
  static constexpr int kQBytes = kTile * HD * 2;
  static constexpr int kKVBytes = kTile * HD * 2;
  static constexpr int kPBytes = kTile * kTile * 2;
  static constexpr int kXchgBytes = 2 * 2 * kTile * 4;
  static constexpr int kSmem = kQBufs * kQBytes + kStages * 2 * kKVBytes + kPBytes + 256 + kXchgBytes + 1024;
  static constexpr int kTmemCols = 512;  // S0 [0,128), S1 [128,256), O [256, 256+HD)
  static constexpr int kOCol = 256;
};

template <int HD>
__global__ void __launch_bounds__(kAttnFwdThreads, 1)
attn_fwd_persistent_kernel(const __grid_constant__ CUtensorMap tm_qkv, __nv_bfloat16* __restrict__ out,
                           float* __restrict__ lse, int B, int T, int H, int Dl, float scale_log2) {
  using C = FwdPCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                 // [kQBufs]
  uint8_t* sK = sQ + C::kQBufs * C::kQBytes;         // [stage]
  uint8_t* sV = sK + C::kStages * C::kKVBytes;       // [stage]
  uint8_t* sP = sV + C::kStages * C::kKVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + C::kPBytes);
  uint64_t* q_full = bars;        // [2]
  uint64_t* q_empty = bars + 2;   // [2]
  uint64_t* k_full = bars + 4;    // [2]
  uint64_t* v_full = bars + 6;    // [2]
  uint64_t* kv_empty = bars + 8;  // [2]
  uint64_t* s_full = bars + 10;   // [2]
  uint64_t* p_full = bars + 12;
  uint64_t* o_done = bars + 13;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
  float* s_xchg = reinterpret_cast<float*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (T + kTile - 1) / kTile;
  const int HB = H * B;
  const int n_items = nq * HB;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&s_full[i], 1);
    }
    mbar_init(p_full, 8);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + C::kOCol;

  // work item w (heavy q tiles first): qt = nq-1 - w / HB, h = (w % HB) % H, b = (w % HB) / H
  if (warp == 0) {
    if (lane == 0) {
      int g = 0, it = 0;
      for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
        const int qt = nq - 1 - w / HB, rem = w % HB, h = rem % H, b = rem / H;
        const int qb = it % C::kQBufs;
        mbar_wait(&q_empty[qb], ((it / C::kQBufs) & 1) ^ 1);
        mbar_expect_tx(&q_full[qb], C::kQBytes);
#pragma unroll
        for (int a = 0; a < HD / 64; ++a)
          tma_load_3d(sQ + qb * C::kQBytes + a * (kTile * 128), &tm_qkv, &q_full[qb], h * HD + a * 64, qt * kTile, b);
        for (int j = 0; j <= qt; ++j, ++g) {
          const int st = g & 1;
          mbar_wait(&kv_empty[st], ((g >> 1) & 1) ^ 1);
          mbar_expect_tx(&k_full[st], C::kKVBytes);
#pragma unroll
          for (int a = 0; a < HD / 64; ++a)
            tma_load_3d(sK + st * C::kKVBytes + a * (kTile * 128), &tm_qkv, &k_full[st], Dl + h * HD + a * 64,
                        j * kTile, b);
          mbar_expect_tx(&v_full[st], C::kKVBytes);
#pragma unroll
          for (int a = 0; a < HD / 64; ++a)
            tma_load_3d(sV + st * C::kKVBytes + a * (kTile * 128), &tm_qkv, &v_full[st], 2 * Dl + h * HD + a * 64,
                        j * kTile, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(kTile, HD, false, true);
      const uint32_t p_base = smem_u32(sP);
      // S(tile g) = Q(item) K(g)^T into S buffer g & 1; commits "Q buffer free" after the last S of an item
      auto issue_s = [&](int g, int it, bool first_of_item, bool last_of_item) {
        const int qb = it % C::kQBufs;
        if (first_of_item) mbar_wait(&q_full[qb], (it / C::kQBufs) & 1);
        const int sn = g & 1;
        mbar_wait(&k_full[sn], (g >> 1) & 1);
        tc_fence_after();
        const uint32_t q_base = smem_u32(sQ + qb * C::kQBytes);
        const uint32_t k_base = smem_u32(sK + sn * C::kKVBytes);
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk)
          umma_bf16_ss(tS + sn * 128, desc_kmajor(q_base, kk, kTile * 128), desc_kmajor(k_base, kk, kTile * 128),
                       idesc_s, kk > 0);
        umma_commit(&s_full[sn]);
        if (last_of_item) umma_commit(&q_empty[qb]);
      };
      int g = 0, it = 0;
      for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
        const int n_kv = nq - w / HB;  // qt + 1
        if (it == 0) issue_s(0, 0, true, n_kv == 1);
        for (int j = 0; j < n_kv; ++j, ++g) {
          // look one tile ahead (possibly into the next work item) so QK^T overlaps this tile's softmax
          if (j + 1 < n_kv) {
            issue_s(g + 1, it, false, j + 2 == n_kv);
          } else if (w + static_cast<int>(gridDim.x) < n_items) {
            const int n_kv_next = nq - (w + static_cast<int>(gridDim.x)) / HB;
            issue_s(g + 1, it + 1, true, n_kv_next == 1);
          }
          const int st = g & 1;
          mbar_wait(p_full, g & 1);
          mbar_wait(&v_full[st], (g >> 1) & 1);
          tc_fence_after();
          const uint32_t v_base = smem_u32(sV + st * C::kKVBytes);
#pragma unroll
          for (int kk = 0; kk < kTile / 16; ++kk)
            umma_bf16_ss(tO, desc_kmajor(p_base, kk, kTile * 128), desc_mnmajor(v_base, kk, kTile * 128), idesc_o,
                         (j > 0 || kk > 0) ? 1u : 0u);
          umma_commit(o_done);
          umma_commit(&kv_empty[st]);
        }
      }
    }
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    constexpr int kOChunks = HD / 64;
    int g = 0, xc = 0;
    for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
      const int qt = nq - 1 - w / HB, rem = w % HB, h = rem % H, b = rem / H;
      const int q0 = qt * kTile;
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = 0; j <= qt; ++j, ++g) {
        mbar_wait(&s_full[g & 1], (g >> 1) & 1);
        tc_fence_after();
        float s[64];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t tmp[32];
          tmem_ld32(tS + (g & 1) * 128 + lane_off + (2 * half + c) * 32, tmp);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(tmp[i]) * scale_log2;
        }
        if (j == qt) {
#pragma unroll
          for (int i = 0; i < 64; ++i)
            if (half * 64 + i > r) s[i] = -INFINITY;
        }
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          m4[0] = fmaxf(m4[0], s[i]);
          m4[1] = fmaxf(m4[1], s[i + 1]);
          m4[2] = fmaxf(m4[2], s[i + 2]);
          m4[3] = fmaxf(m4[3], s[i + 3]);
        }
        const float mloc = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        float* xb = s_xchg + (xc & 1) * 256;
        ++xc;
        xb[half * 128 + r] = mloc;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float mcand = fmaxf(mloc, xb[(half ^ 1) * 128 + r]);
        const float mx = (mcand > m_run + 8.f) ? mcand : m_run;  // lazy rescaling (see attn_fwd_kernel)
        const float alpha = fast_exp2(m_run - mx);
        m_run = mx;
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
        uint32_t packed[32];
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          const float p0 = fast_exp2(s[i] - mx), p1 = fast_exp2(s[i + 1] - mx);
          const float p2 = fast_exp2(s[i + 2] - mx), p3 = fast_exp2(s[i + 3] - mx);
          s4[0] += p0;
          s4[1] += p1;
          s4[2] += p2;
          s4[3] += p3;
          packed[i >> 1] = pack_bf16(p0, p1);
          packed[(i >> 1) + 1] = pack_bf16(p2, p3);
        }
        l_run = l_run * alpha + ((s4[0] + s4[1]) + (s4[2] + s4[3]));
        if (j > 0) {
          mbar_wait(o_done, (g - 1) & 1);  // PV of the previous tile: O may be rescaled, P may be overwritten
          tc_fence_after();
          if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
            for (int c = half * kOChunks; c < (half + 1) * kOChunks; ++c) {
              uint32_t o[32];
              tmem_ld32(tO + lane_off + c * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st32(tO + lane_off + c * 32, o);
            }
            tmem_st_wait();
          }
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t chunk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) chunk[i] = packed[c * 16 + i];
          st_tile_row32(sP, r, half * 64 + c * 32, chunk);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
      }
      // item epilogue (the previous item's P / O are free again once this returns: the next p_full needs us)
      float* xb = s_xchg + (xc & 1) * 256;
      ++xc;
      xb[half * 128 + r] = l_run;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const float l_tot = l_run + xb[(half ^ 1) * 128 + r];
      mbar_wait(o_done, (g - 1) & 1);
      tc_fence_after();
      const float inv_l = 1.f / l_tot;
      const bool valid = (q0 + r) < T;
      __nv_bfloat16* orow = out + (static_cast<long long>(b) * T + q0 + r) * Dl + h * HD;
#pragma unroll
      for (int c = half * kOChunks; c < (half + 1) * kOChunks; ++c) {
        uint32_t o[32];
        tmem_ld32(tO + lane_off + c * 32, o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int v8 = 0; v8 < 4; ++v8) {
            uint4 wv;
            wv.x = pack_bf16(__uint_as_float(o[8 * v8 + 0]) * inv_l, __uint_as_float(o[8 * v8 + 1]) * inv_l);
            wv.y = pack_bf16(__uint_as_float(o[8 * v8 + 2]) * inv_l, __uint_as_float(o[8 * v8 + 3]) * inv_l);
            wv.z = pack_bf16(__uint_as_float(o[8 * v8 + 4]) * inv_l, __uint_as_float(o[8 * v8 + 5]) * inv_l);
            wv.w = pack_bf16(__uint_as_float(o[8 * v8 + 6]) * inv_l, __uint_as_float(o[8 * v8 + 7]) * inv_l);
            *reinterpret_cast<uint4*>(orow + c * 32 + v8 * 8) = wv;
          }
        }
      }
      if (half == 0 && valid && lse)
        lse[(static_cast<long long>(b) * H + h) * T + q0 + r] = (m_run + log2f(l_tot)) * kLn2;
      tc_fence_before();   // the O reads above are ordered before the next item's first PV (via p_full)
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ================================================================================================= forward, ping-pong
// EXPERIMENTAL (DPFS_ATTN_FWD_PP=1, hd = 64 only).  The persistent kernel above is paced by the softmax warps' own
// dependent chain (TMEM load -> row max -> exchange -> exp2 -> pack -> smem -> fence -> signal): the tensor pipe is 13 %
// busy and the MUFU pipe would allow ~4x the rate.  Here one CTA works on TWO q tiles of the same (batch, head) -- tile
// A = 2p, tile B = 2p + 1 -- with one softmax warp group each (one thread per row, no cross-thread exchange at all).
// Every SM sub-partition then hosts one A warp and one B warp that are in different phases of the chain, and the MMA
// thread alternates [PV_A(j), QK_A(j+1)], [PV_B(j), QK_B(j+1)], so a group's MMA latency hides behind the other
// group's exponentials.  K / V tiles are loaded once for both q tiles (3-stage ring), Q is double buffered across
// work items.  TMEM: S_A [0,128) S_B [128,256) O_A [256,320) O_B [320,384); S is single buffered per tile and is
// released by the p_full signal (the group has read it twice -- max pass, exp pass -- before it signals).
struct FwdPPCfg {
  static constexpr int HD = 64;
  static constexpr int kKvStages = 3;
  static constexpr int kQBytes = kTile * HD * 2;      // one q tile
  static constexpr int kKVBytes = kTile * HD * 2;     // one K or V tile
  static constexpr int kPBytes = kTile * kTile * 2;   // one P tile
  static constexpr int kSmem = 2 * 2 * kQBytes + kKvStages * 2 * kKVBytes + 2 * kPBytes + 256 + 1024;
  static constexpr int kTmemCols = 512;
  static_assert(kSmem <= 232448, "ping-pong forward does not fit in shared memory");
};

__global__ void __launch_bounds__(kAttnFwdThreads, 1)
attn_fwd_pp_kernel(const __grid_constant__ CUtensorMap tm_qkv, __nv_bfloat16* __restrict__ out,
                   float* __restrict__ lse, int B, int T, int H, int Dl, float scale_log2) {
  using C = FwdPPCfg;
  constexpr int HD = C::HD;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                      // [q buffer][tile A | tile B]
  uint8_t* sK = sQ + 2 * 2 * C::kQBytes;                   // [stage]
  uint8_t* sV = sK + C::kKvStages * C::kKVBytes;           // [stage]
  uint8_t* sP = sV + C::kKvStages * C::kKVBytes;           // [tile A | tile B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * C::kPBytes);
  uint64_t* q_full = bars;         // [2]  per Q buffer
  uint64_t* q_empty = bars + 2;    // [2]
  uint64_t* k_full = bars + 4;     // [3]
  uint64_t* v_full = bars + 7;     // [3]
  uint64_t* kv_empty = bars + 10;  // [3]
  uint64_t* s_full = bars + 13;    // [2]  per tile (A, B)
  uint64_t* p_full = bars + 15;    // [2]
  uint64_t* o_done = bars + 17;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (T + kTile - 1) / kTile;
  const int n_pairs = (nq + 1) / 2;
  const int HB = H * B;
  const int n_items = n_pairs * HB;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);   // the four warps of a group
      mbar_init(&o_done[i], 1);
    }
    for (int i = 0; i < C::kKvStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work item w (heavy pairs first): pair = n_pairs-1 - w / HB; q tiles 2*pair (A) and 2*pair+1 (B, if it exists);
  // kv tiles 0 .. n_kv-1 with n_kv = (B exists ? 2*pair+2 : 2*pair+1); A takes part in tiles j <= 2*pair.
  if (warp == 0) {
    if (lane == 0) {
      int g = 0, it = 0;
      for (int k_ = 0, w = blockIdx.x; w < n_items; ++k_, w = item_w<true>(k_, w), ++it) {
        const int pair = n_pairs - 1 - w / HB, rem = w % HB, h = rem % H, b = rem / H;
        const bool has_b = 2 * pair + 1 < nq;
        const int n_kv = has_b ? 2 * pair + 2 : 2 * pair + 1;
        const int qb = it & 1;
        mbar_wait(&q_empty[qb], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&q_full[qb], (has_b ? 2 : 1) * C::kQBytes);
        tma_load_3d(sQ + (qb * 2 + 0) * C::kQBytes, &tm_qkv, &q_full[qb], h * HD, (2 * pair) * kTile, b);
        if (has_b)
          tma_load_3d(sQ + (qb * 2 + 1) * C::kQBytes, &tm_qkv, &q_full[qb], h * HD, (2 * pair + 1) * kTile, b);
        for (int j = 0; j < n_kv; ++j, ++g) {
          const int st = g % C::kKvStages;
          mbar_wait(&kv_empty[st], ((g / C::kKvStages) & 1) ^ 1);
          mbar_expect_tx(&k_full[st], C::kKVBytes);
          tma_load_3d(sK + st * C::kKVBytes, &tm_qkv, &k_full[st], Dl + h * HD, j * kTile, b);
          mbar_expect_tx(&v_full[st], C::kKVBytes);
          tma_load_3d(sV + st * C::kKVBytes, &tm_qkv, &v_full[st], 2 * Dl + h * HD, j * kTile, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(kTile, HD, false, true);
      int g = 0, it = 0;
      int cnt_a = 0, cnt_b = 0;   // steps issued so far per tile slot (phase of s_full / p_full / o_done)
      for (int k_ = 0, w = blockIdx.x; w < n_items; ++k_, w = item_w<true>(k_, w), ++it) {
        const int pair = n_pairs - 1 - w / HB;
        const bool has_b = 2 * pair + 1 < nq;
        const int n_kv = has_b ? 2 * pair + 2 : 2 * pair + 1;
        const int n_of_a = 2 * pair + 1, n_of_b = has_b ? 2 * pair + 2 : 0;   // kv tiles each q tile takes part in
        const int qb = it & 1;
        const int g0 = g;   // ring position of kv tile 0 of this item
        // S_x(j) = Q_x K(j)^T
        auto issue_qk = [&](int x, int j) {
          const int st = (g0 + j) % C::kKvStages;
          mbar_wait(&k_full[st], ((g0 + j) / C::kKvStages) & 1);
          tc_fence_after();
          const uint32_t q_base = smem_u32(sQ + (qb * 2 + x) * C::kQBytes);
          const uint32_t k_base = smem_u32(sK + st * C::kKVBytes);
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk)
            umma_bf16_ss(tmem_base + x * 128, desc_kmajor(q_base, kk, kTile * 128),
                         desc_kmajor(k_base, kk, kTile * 128), idesc_s, kk > 0);
          umma_commit(&s_full[x]);
        };
        mbar_wait(&q_full[qb], (it >> 1) & 1);
        issue_qk(0, 0);
        if (has_b) issue_qk(1, 0);
        if (n_kv == 1) umma_commit(&q_empty[qb]);   // (single-tile item: no further QK)
        for (int j = 0; j < n_kv; ++j) {
          const int st = (g0 + j) % C::kKvStages;
          bool v_ready = false;
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            const int n_of_x = x == 0 ? n_of_a : n_of_b;
            int& cnt_x = x == 0 ? cnt_a : cnt_b;
            if (j >= n_of_x) continue;
            mbar_wait(&p_full[x], cnt_x & 1);       // P_x(j) is in smem, S_x has been read out
            if (!v_ready) {
              mbar_wait(&v_full[st], ((g0 + j) / C::kKvStages) & 1);
              v_ready = true;
            }
            tc_fence_after();
            const uint32_t p_base = smem_u32(sP + x * C::kPBytes);
            const uint32_t v_base = smem_u32(sV + st * C::kKVBytes);
#pragma unroll
            for (int kk = 0; kk < kTile / 16; ++kk)
              umma_bf16_ss(tmem_base + 256 + x * HD, desc_kmajor(p_base, kk, kTile * 128),
                           desc_mnmajor(v_base, kk, kTile * 128), idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
            umma_commit(&o_done[x]);
            ++cnt_x;
            if (j + 1 < n_of_x) {
              issue_qk(x, j + 1);
              // the item's last QK is tile B's (or A's when there is no B) for kv tile n_kv - 1
              if (j + 2 == n_kv && (x == 1 || !has_b)) umma_commit(&q_empty[qb]);
            }
          }
          umma_commit(&kv_empty[st]);   // K(j) / V(j) have no reader left once the MMAs issued so far retire
        }
        g += n_kv;
      }
    }
  } else {
    // softmax groups: warps 2-5 -> tile A, warps 6-9 -> tile B; thread = one q row, all 128 kv columns
    const int x = (warp - 2) >> 2;
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t tS = tmem_base + x * 128, tO = tmem_base + 256 + x * HD;
    uint8_t* sPx = sP + x * C::kPBytes;
    int cnt = 0;   // steps done by this group (phase of s_full / p_full / o_done of slot x)
    for (int k_ = 0, w = blockIdx.x; w < n_items; ++k_, w = item_w<true>(k_, w)) {
      const int pair = n_pairs - 1 - w / HB, rem = w % HB, h = rem % H, b = rem / H;
      const int qt = 2 * pair + x;
      if (qt >= nq) continue;   // odd number of q tiles: the last pair has no tile B
      const int q0 = qt * kTile;
      float m_run = -INFINITY, l_run = 0.f;
      // One kv tile.  DIAG (the causal diagonal tile, the last one of the item) is a compile-time flag so that the
      // 127 other tiles of a long row do not pay per-element mask predicates.
      auto kv_step = [&](int j, auto diag_tag) {
        constexpr bool DIAG = decltype(diag_tag)::value;
        mbar_wait(&s_full[x], cnt & 1);
        tc_fence_after();
        // pass 1: row max over the 128 columns (two TMEM loads in flight)
        float mloc = -INFINITY;
#pragma unroll 1
        for (int c2 = 0; c2 < 2; ++c2) {
          uint32_t t0[32], t1[32];
          tmem_ld32(tS + lane_off + (2 * c2) * 32, t0);
          tmem_ld32(tS + lane_off + (2 * c2 + 1) * 32, t1);
          tmem_ld_wait();
          if constexpr (DIAG) {   // causal mask, also hides kv columns past T
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if ((2 * c2) * 32 + i > r) t0[i] = 0xff800000u;       // -inf
              if ((2 * c2 + 1) * 32 + i > r) t1[i] = 0xff800000u;
            }
          }
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
          for (int i = 0; i < 32; ++i)
            m4[i & 3] = fmaxf(m4[i & 3], fmaxf(__uint_as_float(t0[i]), __uint_as_float(t1[i])));
          mloc = fmaxf(mloc, fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])));
        }
        mloc *= scale_log2;   // scale > 0: max commutes with the scaling
        const float mx = (mloc > m_run + 8.f) ? mloc : m_run;   // lazy rescaling (see attn_fwd_kernel)
        const float alpha = fast_exp2(m_run - mx);
        m_run = mx;
        if (j > 0) {
          mbar_wait(&o_done[x], (cnt - 1) & 1);   // PV_x(j-1): O may be rescaled, P may be overwritten
          tc_fence_after();
          if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
            for (int c = 0; c < HD / 32; ++c) {
              uint32_t o[32];
              tmem_ld32(tO + lane_off + c * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st32(tO + lane_off + c * 32, o);
            }
            tmem_st_wait();
          }
        }
        // pass 2: p = exp2(s * scale - m) -> bf16 P tile in smem, row sum
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
        auto emit_chunk = [&](int c, uint32_t (&tv)[32]) {
          if constexpr (DIAG) {   // masked scores become -inf: exp2 gives an exact 0
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c * 32 + i > r) tv[i] = 0xff800000u;
          }
          uint32_t packed[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float p0 = fast_exp2(__uint_as_float(tv[i]) * scale_log2 - mx);
            const float p1 = fast_exp2(__uint_as_float(tv[i + 1]) * scale_log2 - mx);
            s4[i & 3] += p0;
            s4[(i + 1) & 3] += p1;
            packed[i >> 1] = pack_bf16(p0, p1);
          }
          st_tile_row32(sPx, r, c * 32, packed);
        };
#pragma unroll 1
        for (int c2 = 0; c2 < 2; ++c2) {   // two TMEM loads in flight, like pass 1
          uint32_t ta[32], tb[32];
          tmem_ld32(tS + lane_off + (2 * c2) * 32, ta);
          tmem_ld32(tS + lane_off + (2 * c2 + 1) * 32, tb);
          tmem_ld_wait();
          emit_chunk(2 * c2, ta);
          emit_chunk(2 * c2 + 1, tb);
        }
        l_run = l_run * alpha + ((s4[0] + s4[1]) + (s4[2] + s4[3]));
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[x]);
      };
      for (int j = 0; j < qt; ++j, ++cnt) kv_step(j, std::false_type{});
      kv_step(qt, std::true_type{});
      ++cnt;
      // item epilogue for this group's tile
      mbar_wait(&o_done[x], (cnt - 1) & 1);
      tc_fence_after();
      const float inv_l = 1.f / l_run;
      const bool valid = (q0 + r) < T;
      __nv_bfloat16* orow = out + (static_cast<long long>(b) * T + q0 + r) * Dl + h * HD;
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t o[32];
        tmem_ld32(tO + lane_off + c * 32, o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int v8 = 0; v8 < 4; ++v8) {
            uint4 wv;
            wv.x = pack_bf16(__uint_as_float(o[8 * v8 + 0]) * inv_l, __uint_as_float(o[8 * v8 + 1]) * inv_l);
            wv.y = pack_bf16(__uint_as_float(o[8 * v8 + 2]) * inv_l, __uint_as_float(o[8 * v8 + 3]) * inv_l);
            wv.z = pack_bf16(__uint_as_float(o[8 * v8 + 4]) * inv_l, __uint_as_float(o[8 * v8 + 5]) * inv_l);
            wv.w = pack_bf16(__uint_as_float(o[8 * v8 + 6]) * inv_l, __uint_as_float(o[8 * v8 + 7]) * inv_l);
            *reinterpret_cast<uint4*>(orow + c * 32 + v8 * 8) = wv;
          }
        }
      }
      if (valid && lse) lse[(static_cast<long long>(b) * H + h) * T + q0 + r] = (m_run + log2f(l_run)) * kLn2;
      tc_fence_before();   // the O reads above are ordered before the next item's first PV_x (via p_full)
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ================================================================================================= backward
// delta[b][h][i] = sum_c dO[i][c] * O[i][c]   (one warp per (row, head))
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                  float* __restrict__ delta, int B, int T, int H, int HD) {
  // one 16-byte vector (8 bf16) of O and dO per thread; the HD/8 consecutive lanes of a head reduce by shuffle
  const int Dl = H * HD;
  const int vec = Dl >> 3, grp = HD >> 3;  // grp = 8 (hd 64) or 16 (hd 128): divides the warp
  const long long total = static_cast<long long>(B) * T * vec;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long rounds = (total + stride - 1) / stride;
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  for (long long it = 0; it < rounds; ++it, i += stride) {
    float acc = 0.f;
    const bool live = i < total;
    if (live) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(o) + i);
      const uint4 g = __ldg(reinterpret_cast<const uint4*>(d_o) + i);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 x = unpack_bf16(aw[q]), y = unpack_bf16(gw[q]);
        acc += x.x * y.x + x.y * y.y;
      }
    }
    for (int off = grp >> 1; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (live && (i % grp) == 0) {
      const long long row = i / vec;
      const int h = static_cast<int>(i % vec) / grp;
      const long long bb = row / T, tt = row % T;
      delta[(bb * H + h) * T + tt] = acc;
    }
  }
}
__global__ void attn_dq_convert_kernel(const float* __restrict__ dq, __nv_bfloat16* __restrict__ dqkv, long long M,
                                       int Dl) {
  const int vec = Dl / 8;
  const long long total = M * vec;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long row = i / vec;
    const int c = static_cast<int>(i % vec) * 8;
    const float4 v0 = __ldcs(reinterpret_cast<const float4*>(dq + row * Dl + c));
    const float4 v1 = __ldcs(reinterpret_cast<const float4*>(dq + row * Dl + c + 4));
    uint4 o;
    o.x = pack_bf16(v0.x, v0.y);
    o.y = pack_bf16(v0.z, v0.w);
    o.z = pack_bf16(v1.x, v1.y);
    o.w = pack_bf16(v1.z, v1.w);
    *reinterpret_cast<uint4*>(dqkv + row * 3 * Dl + c) = o;
  }
}

template <int HD>
struct BwdCfg {
  static constexpr int kStages = (HD == 64) ? 2 : 1;
  static constexpr int kTileBytes = kTile * HD * 2;  // K, V, Q or dO tile
  static constexpr int kPBytes = kTile * kTile * 2;
  static constexpr int kSmem = 2 * kTileBytes + kStages * 2 * kTileBytes + 2 * kPBytes + 256 + 1024;
  // persistent variant: K / V double-buffered across work items when they fit (hd = 64: 192 KB)
  static constexpr int kKvBufs = (HD == 64) ? 2 : 1;
  static constexpr int kDqSlabBytes = 8 * 4096;  // one 32x32 fp32 block per math warp
  static constexpr int kSmemPersistent = kSmem + (kKvBufs - 1) * 2 * kTileBytes + kDqSlabBytes;
  static_assert(kSmemPersistent <= 232448, "persistent backward does not fit in shared memory");
  // S [0,128), dP [128,256), dV [256,256+HD), dK [256+HD,256+2HD); dQ has its own columns [384,448) when hd = 64
  // (the next tile's S / dP MMAs then overlap the dQ drain) and aliases S when hd = 128 (all 512 columns are taken)
  static constexpr int kTmemCols = 512;
  static constexpr bool kDqAliasesS = (HD != 64);
  static constexpr int kDqCol = kDqAliasesS ? 0 : 384;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

template <int HD>
__global__ void __launch_bounds__(kAttnBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_acc,
                __nv_bfloat16* __restrict__ dqkv, int T, int H, int Dl, float scale, float scale_log2) {
  using C = BwdCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + C::kTileBytes;
  uint8_t* sQ = sV + C::kTileBytes;                  // [stage]
  uint8_t* sDO = sQ + C::kStages * C::kTileBytes;    // [stage]
  uint8_t* sP = sDO + C::kStages * C::kTileBytes;
  uint8_t* sDS = sP + C::kPBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + C::kPBytes);
  uint64_t* kv_full = bars;
  uint64_t* qdo_full = bars + 1;   // [2]
  uint64_t* qdo_empty = bars + 3;  // [2]
  uint64_t* sdp_full = bars + 5;
  uint64_t* pds_full = bars + 6;
  uint64_t* dq_full = bars + 7;
  uint64_t* s_free = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (T + kTile - 1) / kTile;
  const int jt = blockIdx.x;  // kv tile; tile 0 has the most q tiles -> scheduled first
  const int h = blockIdx.y, b = blockIdx.z;
  const int kv0 = jt * kTile;
  const int n_it = nq - jt;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    prefetch_tmap(&tm_do);
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(pds_full, 8);
    mbar_init(dq_full, 1);
    mbar_init(s_free, 8);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDV = tmem_base + 256, tDK = tmem_base + 256 + HD;
  const uint32_t tDQ = tmem_base + C::kDqCol;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * C::kTileBytes);
#pragma unroll
      for (int a = 0; a < HD / 64; ++a) {
        tma_load_3d(sK + a * (kTile * 128), &tm_qkv, kv_full, Dl + h * HD + a * 64, kv0, b);
        tma_load_3d(sV + a * (kTile * 128), &tm_qkv, kv_full, 2 * Dl + h * HD + a * 64, kv0, b);
      }
      for (int it = 0; it < n_it; ++it) {
        const int st = it % C::kStages;
        const int q0 = (jt + it) * kTile;
        mbar_wait(&qdo_empty[st], ((it / C::kStages) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[st], 2 * C::kTileBytes);
#pragma unroll
        for (int a = 0; a < HD / 64; ++a) {
          tma_load_3d(sQ + st * C::kTileBytes + a * (kTile * 128), &tm_qkv, &qdo_full[st], h * HD + a * 64, q0, b);
          tma_load_3d(sDO + st * C::kTileBytes + a * (kTile * 128), &tm_do, &qdo_full[st], h * HD + a * 64, q0, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);   // S, dP
      constexpr uint32_t idesc_kv = make_idesc_bf16(kTile, HD, true, true);       // dV, dK
      constexpr uint32_t idesc_q = make_idesc_bf16(kTile, HD, false, true);       // dQ
      const uint32_t k_base = smem_u32(sK), v_base = smem_u32(sV), p_base = smem_u32(sP), ds_base = smem_u32(sDS);
      constexpr uint32_t kAtom = kTile * 128;
      mbar_wait(kv_full, 0);
      for (int it = 0; it < n_it; ++it) {
        const int st = it % C::kStages;
        const uint32_t q_base = smem_u32(sQ + st * C::kTileBytes), do_base = smem_u32(sDO + st * C::kTileBytes);
        mbar_wait(&qdo_full[st], (it / C::kStages) & 1);
        if (C::kDqAliasesS && it > 0) mbar_wait(s_free, (it - 1) & 1);   // dQ(it-1) drained out of the S columns
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk)
          umma_bf16_ss(tS, desc_kmajor(q_base, kk, kAtom), desc_kmajor(k_base, kk, kAtom), idesc_s, kk > 0);
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk)
          umma_bf16_ss(tDP, desc_kmajor(do_base, kk, kAtom), desc_kmajor(v_base, kk, kAtom), idesc_s, kk > 0);
        umma_commit(sdp_full);
        mbar_wait(pds_full, it & 1);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)  // dV += P^T dO   (reduction over the 128 q rows)
          umma_bf16_ss(tDV, desc_mnmajor(p_base, kk, kAtom), desc_mnmajor(do_base, kk, kAtom), idesc_kv,
                       (it > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)  // dK += dS^T Q
          umma_bf16_ss(tDK, desc_mnmajor(ds_base, kk, kAtom), desc_mnmajor(q_base, kk, kAtom), idesc_kv,
                       (it > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&qdo_empty[st]);
        if (!C::kDqAliasesS && it > 0) {
          mbar_wait(s_free, (it - 1) & 1);   // previous dQ tile drained out of its own columns
          tc_fence_after();
        }
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)  // dQ = dS K   (reduction over the 128 kv rows)
          umma_bf16_ss(tDQ, desc_kmajor(ds_base, kk, kAtom), desc_mnmajor(k_base, kk, kAtom), idesc_q, kk > 0);
        umma_commit(dq_full);
      }
    }
  } else {
    // 8 warps: warp group `half` (0/1) owns kv columns [64*half, 64*half+64) of every row; inside a group warp q owns
    // TMEM lanes (rows) [32q, 32q+32).  Two warps per SM sub-partition hide each other's exp / TMEM latency.
    const int q = warp & 3;            // the hardware ties a warp to TMEM lanes [32*(warp%4), +32)
    const int half = (warp - 2) >> 2;  // warps 2-5 -> 0, warps 6-9 -> 1 (each group covers all four lane quarters)
    const int r = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    constexpr int kOutChunks = HD / 64;  // 32-column chunks of dQ / dK / dV per warp group
    for (int it = 0; it < n_it; ++it) {
      const int q0 = (jt + it) * kTile;
      const bool row_valid = (q0 + r) < T;
      const long long stat_idx = (static_cast<long long>(b) * H + h) * T + q0 + r;
      const float lse2 = row_valid ? lse[stat_idx] * kLog2e : 0.f;
      const float dlt = row_valid ? delta[stat_idx] : 0.f;
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 2 * half; c < 2 * half + 2; ++c) {
        uint32_t sv[32], dv[32];
        tmem_ld32(tS + lane_off + c * 32, sv);
        tmem_ld32(tDP + lane_off + c * 32, dv);
        tmem_ld_wait();
        uint32_t pp[16], dd[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = fast_exp2(__uint_as_float(sv[i]) * scale_log2 - lse2);
          float p1 = fast_exp2(__uint_as_float(sv[i + 1]) * scale_log2 - lse2);
          const int col = c * 32 + i;
          if (!row_valid || (it == 0 && col > r)) p0 = 0.f;
          if (!row_valid || (it == 0 && col + 1 > r)) p1 = 0.f;
          const float d0 = p0 * (__uint_as_float(dv[i]) - dlt) * scale;
          const float d1 = p1 * (__uint_as_float(dv[i + 1]) - dlt) * scale;
          pp[i >> 1] = pack_bf16(p0, p1);
          dd[i >> 1] = pack_bf16(d0, d1);
        }
        st_tile_row32(sP, r, c * 32, pp);
        st_tile_row32(sDS, r, c * 32, dd);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
      // dQ tile -> fp32 atomics
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      float* dq_row = dq_acc + (static_cast<long long>(b) * T + q0 + r) * Dl + h * HD;
#pragma unroll 1
      for (int c = half * kOutChunks; c < (half + 1) * kOutChunks; ++c) {
        uint32_t v[32];
        tmem_ld32(tDQ + lane_off + c * 32, v);
        tmem_ld_wait();
        if (row_valid) {
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            red_add_v4(dq_row + c * 32 + i, __uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                       __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_free);
    }
    // all MMAs (incl. the last dV / dK updates) completed before the last dq_full fired
    const bool kv_valid = (kv0 + r) < T;
    __nv_bfloat16* drow = dqkv + (static_cast<long long>(b) * T + kv0 + r) * 3 * Dl + h * HD;
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {  // 0: dK -> segment 1, 1: dV -> segment 2
      const uint32_t tsrc = which == 0 ? tDK : tDV;
      __nv_bfloat16* dst = drow + (which + 1) * Dl;
#pragma unroll 1
      for (int c = half * kOutChunks; c < (half + 1) * kOutChunks; ++c) {
        uint32_t v[32];
        tmem_ld32(tsrc + lane_off + c * 32, v);
        tmem_ld_wait();
        if (kv_valid) {
#pragma unroll
          for (int v8 = 0; v8 < 4; ++v8) {
            uint4 w;
            w.x = pack_bf16(__uint_as_float(v[8 * v8 + 0]), __uint_as_float(v[8 * v8 + 1]));
            w.y = pack_bf16(__uint_as_float(v[8 * v8 + 2]), __uint_as_float(v[8 * v8 + 3]));
            w.z = pack_bf16(__uint_as_float(v[8 * v8 + 4]), __uint_as_float(v[8 * v8 + 5]));
            w.w = pack_bf16(__uint_as_float(v[8 * v8 + 6]), __uint_as_float(v[8 * v8 + 7]));
            *reinterpret_cast<uint4*>(dst + c * 32 + v8 * 8) = w;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// V2 (opt-in, DPFS_ATTN_BWD_V2=1): the per-element causal / row-validity predicates of the P, dS computation are paid
// only on the diagonal tile (compile-time flag per step) and rows past T get lse = +inf (p = exp2(-inf) = 0) instead
// of a select per element -- about a quarter fewer instructions per element for the math warps.
template <int HD, bool V2 = false>
__global__ void __launch_bounds__(kAttnBwdThreads, 1)
attn_bwd_persistent_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                const __grid_constant__ CUtensorMap tm_dq, const __grid_constant__ CUtensorMap tm_dkv,
                const float* __restrict__ lse, const float* __restrict__ delta,
                __nv_bfloat16* __restrict__ dqkv, int B, int T, int H, int Dl, float scale, float scale_log2) {
  using C = BwdCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sKV = smem;                               // [kv buffer]{K, V}
  uint8_t* sQ = sKV + C::kKvBufs * 2 * C::kTileBytes;  // [stage]
  uint8_t* sDO = sQ + C::kStages * C::kTileBytes;    // [stage]
  uint8_t* sP = sDO + C::kStages * C::kTileBytes;
  uint8_t* sDS = sP + C::kPBytes;
  uint8_t* sDQ = sDS + C::kPBytes;                   // [math warp] 32 rows x 32 fp32, 128B-swizzled (TMA reduce source)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDQ + C::kDqSlabBytes);
  uint64_t* kv_full = bars;        // [2]
  uint64_t* qdo_full = bars + 2;   // [2]
  uint64_t* qdo_empty = bars + 4;  // [2]
  uint64_t* sdp_full = bars + 6;
  uint64_t* pds_full = bars + 7;
  uint64_t* dq_full = bars + 8;
  uint64_t* s_free = bars + 9;
  uint64_t* kv_empty = bars + 10;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
  uint64_t* sdp_read = bars + 13;  // kLook only: the math warps have read S / dP of the current step out of TMEM
  // V2, hd = 64 (private dQ columns, two K/V buffers): S / dP of step g+1 are issued as soon as the math warps have
  // READ S / dP of step g, i.e. while they are still computing P / dS(g) -- instead of queueing behind dQ / dV / dK(g).
  // The math warps then find S / dP(g+1) complete when they come back from the dQ drain.  Because S / dP(g+1) no longer
  // orders dV / dK(g) before the next step's math, the math warps wait for the Q/dO stage release of step g (committed
  // after dV / dK(g)) before they overwrite P / dS.
  constexpr bool kLook = V2 && !C::kDqAliasesS && C::kKvBufs == 2;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (T + kTile - 1) / kTile;
  const int HB = H * B;
  const int n_items = nq * HB;  // work item w: kv tile jt = w / HB (tile 0 has the most q tiles -> first), h, b

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    prefetch_tmap(&tm_do);
    prefetch_tmap(&tm_dq);
    prefetch_tmap(&tm_dkv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(pds_full, 8);
    mbar_init(dq_full, 1);
    mbar_init(s_free, 8);
    if constexpr (kLook) mbar_init(sdp_read, 8);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDV = tmem_base + 256, tDK = tmem_base + 256 + HD;
  const uint32_t tDQ = tmem_base + C::kDqCol;

  if (warp == 0) {
    if (lane == 0) {
      int gs = 0, item = 0;
      for (ItemWalk<V2> iw_; iw_.w < n_items; iw_.next(), ++item) {
      const int w = iw_.w;
        const int jt = w / HB, rem = w % HB, h = rem % H, b = rem / H;
        const int kv0 = jt * kTile, n_it = nq - jt;
        const int kb = item % C::kKvBufs;
        uint8_t* sK = sKV + kb * 2 * C::kTileBytes;
        uint8_t* sV = sK + C::kTileBytes;
        mbar_wait(&kv_empty[kb], ((item / C::kKvBufs) & 1) ^ 1);   // the MMAs that read this K / V buffer are done
        mbar_expect_tx(&kv_full[kb], 2 * C::kTileBytes);
#pragma unroll
        for (int a = 0; a < HD / 64; ++a) {
          tma_load_3d(sK + a * (kTile * 128), &tm_qkv, &kv_full[kb], Dl + h * HD + a * 64, kv0, b);
          tma_load_3d(sV + a * (kTile * 128), &tm_qkv, &kv_full[kb], 2 * Dl + h * HD + a * 64, kv0, b);
        }
        for (int it = 0; it < n_it; ++it, ++gs) {
          const int st = gs % C::kStages;
          const int q0 = (jt + it) * kTile;
          mbar_wait(&qdo_empty[st], ((gs / C::kStages) & 1) ^ 1);
          mbar_expect_tx(&qdo_full[st], 2 * C::kTileBytes);
#pragma unroll
          for (int a = 0; a < HD / 64; ++a) {
            tma_load_3d(sQ + st * C::kTileBytes + a * (kTile * 128), &tm_qkv, &qdo_full[st], h * HD + a * 64, q0, b);
            tma_load_3d(sDO + st * C::kTileBytes + a * (kTile * 128), &tm_do, &qdo_full[st], h * HD + a * 64, q0, b);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, false, false);   // S, dP
      constexpr uint32_t idesc_kv = make_idesc_bf16(kTile, HD, true, true);       // dV, dK
      constexpr uint32_t idesc_q = make_idesc_bf16(kTile, HD, false, true);       // dQ
      const uint32_t p_base = smem_u32(sP), ds_base = smem_u32(sDS);
      constexpr uint32_t kAtom = kTile * 128;
      if constexpr (kLook) {
        int gs = 0, item = 0;
        auto kbase = [&](int it_) { return smem_u32(sKV + (it_ % C::kKvBufs) * 2 * C::kTileBytes); };
        // S(g) = Q K^T and dP(g) = dO V^T for global step g, K / V of the item it belongs to
        auto issue_sdp = [&](int g, uint32_t k_base, uint32_t v_base) {
          const int st = g % C::kStages;
          const uint32_t q_base = smem_u32(sQ + st * C::kTileBytes), do_base = smem_u32(sDO + st * C::kTileBytes);
          mbar_wait(&qdo_full[st], (g / C::kStages) & 1);
          tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk)
            umma_bf16_ss(tS, desc_kmajor(q_base, kk, kAtom), desc_kmajor(k_base, kk, kAtom), idesc_s, kk > 0);
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk)
            umma_bf16_ss(tDP, desc_kmajor(do_base, kk, kAtom), desc_kmajor(v_base, kk, kAtom), idesc_s, kk > 0);
          umma_commit(sdp_full);
        };
        ItemWalk<V2> iw_;
        if (iw_.w < n_items) {
          mbar_wait(&kv_full[0], 0);
          issue_sdp(0, kbase(0), kbase(0) + C::kTileBytes);
        }
        for (; iw_.w < n_items; iw_.next(), ++item) {
          const int w = iw_.w;
          const int n_it = nq - w / HB;
          const int kb = item % C::kKvBufs;
          const uint32_t k_base = kbase(item), v_base = k_base + C::kTileBytes;
          for (int it = 0; it < n_it; ++it, ++gs) {
            const int st = gs % C::kStages;
            const uint32_t q_base = smem_u32(sQ + st * C::kTileBytes), do_base = smem_u32(sDO + st * C::kTileBytes);
            // look ahead: S / dP of the next step (of this item, or the first step of this CTA's next item)
            bool has_next = true;
            uint32_t nk = k_base;
            if (it + 1 == n_it) {
              has_next = iw_.peek() < n_items;
              if (has_next) {
                mbar_wait(&kv_full[(item + 1) % C::kKvBufs], ((item + 1) / C::kKvBufs) & 1);
                nk = kbase(item + 1);
              }
            }
            if (has_next) {
              mbar_wait(sdp_read, gs & 1);   // S / dP(gs) are in the math warps' registers
              tc_fence_after();
              issue_sdp(gs + 1, nk, nk + C::kTileBytes);
            }
            mbar_wait(pds_full, gs & 1);
            tc_fence_after();
            if (gs > 0) {
              mbar_wait(s_free, (gs - 1) & 1);   // previous dQ tile drained out of its own columns
              tc_fence_after();
            }
#pragma unroll
            for (int kk = 0; kk < kTile / 16; ++kk)  // dQ = dS K   (reduction over the 128 kv rows)
              umma_bf16_ss(tDQ, desc_kmajor(ds_base, kk, kAtom), desc_mnmajor(k_base, kk, kAtom), idesc_q, kk > 0);
            umma_commit(dq_full);
#pragma unroll
            for (int kk = 0; kk < kTile / 16; ++kk)  // dV += P^T dO   (reduction over the 128 q rows)
              umma_bf16_ss(tDV, desc_mnmajor(p_base, kk, kAtom), desc_mnmajor(do_base, kk, kAtom), idesc_kv,
                           (it > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
            for (int kk = 0; kk < kTile / 16; ++kk)  // dK += dS^T Q
              umma_bf16_ss(tDK, desc_mnmajor(ds_base, kk, kAtom), desc_mnmajor(q_base, kk, kAtom), idesc_kv,
                           (it > 0 || kk > 0) ? 1u : 0u);
            umma_commit(&qdo_empty[st]);   // also tells the math warps that P / dS(gs) have been consumed
          }
          umma_commit(&kv_empty[kb]);   // K / V of this item are no longer read
        }
      } else {
      int gs = 0, item = 0;
      for (ItemWalk<V2> iw_; iw_.w < n_items; iw_.next(), ++item) {
      const int w = iw_.w;
        const int n_it = nq - w / HB;
        const int kb = item % C::kKvBufs;
        const uint32_t k_base = smem_u32(sKV + kb * 2 * C::kTileBytes), v_base = k_base + C::kTileBytes;
        mbar_wait(&kv_full[kb], (item / C::kKvBufs) & 1);
        for (int it = 0; it < n_it; ++it, ++gs) {
          const int st = gs % C::kStages;
          const uint32_t q_base = smem_u32(sQ + st * C::kTileBytes), do_base = smem_u32(sDO + st * C::kTileBytes);
          mbar_wait(&qdo_full[st], (gs / C::kStages) & 1);
          if (C::kDqAliasesS && gs > 0) mbar_wait(s_free, (gs - 1) & 1);   // dQ(gs-1) drained out of the S columns
          tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk)
            umma_bf16_ss(tS, desc_kmajor(q_base, kk, kAtom), desc_kmajor(k_base, kk, kAtom), idesc_s, kk > 0);
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk)
            umma_bf16_ss(tDP, desc_kmajor(do_base, kk, kAtom), desc_kmajor(v_base, kk, kAtom), idesc_s, kk > 0);
          umma_commit(sdp_full);
          mbar_wait(pds_full, gs & 1);
          tc_fence_after();
          if (!C::kDqAliasesS && gs > 0) {
            mbar_wait(s_free, (gs - 1) & 1);   // previous dQ tile drained out of its own columns
            tc_fence_after();
          }
          // dQ first: the math warps drain it (smem slab + TMA reduce) while dV / dK accumulate
#pragma unroll
          for (int kk = 0; kk < kTile / 16; ++kk)  // dQ = dS K   (reduction over the 128 kv rows)
            umma_bf16_ss(tDQ, desc_kmajor(ds_base, kk, kAtom), desc_mnmajor(k_base, kk, kAtom), idesc_q, kk > 0);
          umma_commit(dq_full);
#pragma unroll
          for (int kk = 0; kk < kTile / 16; ++kk)  // dV += P^T dO   (reduction over the 128 q rows)
            umma_bf16_ss(tDV, desc_mnmajor(p_base, kk, kAtom), desc_mnmajor(do_base, kk, kAtom), idesc_kv,
                         (it > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < kTile / 16; ++kk)  // dK += dS^T Q
            umma_bf16_ss(tDK, desc_mnmajor(ds_base, kk, kAtom), desc_mnmajor(q_base, kk, kAtom), idesc_kv,
                         (it > 0 || kk > 0) ? 1u : 0u);
          umma_commit(&qdo_empty[st]);
        }
        umma_commit(&kv_empty[kb]);   // K / V of this item are no longer read
      }
      }
    }
  } else {
    // 8 warps: warp group `half` (0/1) owns kv columns [64*half, 64*half+64) of every row; inside a group warp q owns
    // TMEM lanes (rows) [32q, 32q+32).  Two warps per SM sub-partition hide each other's exp / TMEM latency.
    const int q = warp & 3;            // the hardware ties a warp to TMEM lanes [32*(warp%4), +32)
    const int half = (warp - 2) >> 2;  // warps 2-5 -> 0, warps 6-9 -> 1 (each group covers all four lane quarters)
    const int r = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    constexpr int kOutChunks = HD / 64;  // 32-column chunks of dQ / dK / dV per warp group
    // row statistics of step (w, it); loaded one step ahead so the global-load latency hides behind the dQ drain
    auto load_stats = [&](int w, int it, float& l2, float& dl) {
      const int jt_ = w / HB, rem_ = w % HB;
      const int row = (jt_ + it) * kTile + r;
      l2 = V2 ? INFINITY : 0.f;
      dl = 0.f;
      if (w < n_items && row < T) {
        const long long idx = (static_cast<long long>(rem_ / H) * H + rem_ % H) * T + row;
        l2 = lse[idx] * kLog2e;
        dl = delta[idx];
      }
    };
    int gs = 0, item = 0;
    float lse2_nx, dlt_nx;
    load_stats(blockIdx.x, 0, lse2_nx, dlt_nx);
    for (ItemWalk<V2> iw_; iw_.w < n_items; iw_.next(), ++item) {
      const int w = iw_.w;
    const int jt = w / HB, rem = w % HB, h = rem % H, b = rem / H;
    const int kv0 = jt * kTile, n_it = nq - jt;
    for (int it = 0; it < n_it; ++it, ++gs) {
      const int q0 = (jt + it) * kTile;
      const bool row_valid = (q0 + r) < T;
      const float lse2 = lse2_nx, dlt = dlt_nx;
      mbar_wait(sdp_full, gs & 1);
      tc_fence_after();
      if constexpr (V2) {
        auto chunks = [&](auto diag_tag) {
          constexpr bool DIAG = decltype(diag_tag)::value;
#pragma unroll 1
          for (int c = 2 * half; c < 2 * half + 2; ++c) {
            uint32_t sv[32], dv[32];
            tmem_ld32(tS + lane_off + c * 32, sv);
            tmem_ld32(tDP + lane_off + c * 32, dv);
            tmem_ld_wait();
            if constexpr (kLook) {
              if (c == 2 * half + 1) {   // last TMEM read of this step: S / dP(gs+1) may overwrite the columns
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(sdp_read);
              } else if (gs > 0) {       // before the first P / dS store: dV / dK(gs-1) have consumed the tiles
                mbar_wait(&qdo_empty[(gs - 1) % C::kStages], ((gs - 1) / C::kStages) & 1);
              }
            }
            if constexpr (DIAG) {   // causal mask: a masked score becomes -inf, so p and dS come out as exact zeros
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (c * 32 + i > r) sv[i] = 0xff800000u;
            }
            uint32_t pp[16], dd[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const float p0 = fast_exp2(__uint_as_float(sv[i]) * scale_log2 - lse2);
              const float p1 = fast_exp2(__uint_as_float(sv[i + 1]) * scale_log2 - lse2);
              const float d0 = p0 * (__uint_as_float(dv[i]) - dlt) * scale;
              const float d1 = p1 * (__uint_as_float(dv[i + 1]) - dlt) * scale;
              pp[i >> 1] = pack_bf16(p0, p1);
              dd[i >> 1] = pack_bf16(d0, d1);
            }
            st_tile_row32(sP, r, c * 32, pp);
            st_tile_row32(sDS, r, c * 32, dd);
          }
        };
        if (it == 0) chunks(std::true_type{}); else chunks(std::false_type{});
      } else {
#pragma unroll 1
      for (int c = 2 * half; c < 2 * half + 2; ++c) {
        uint32_t sv[32], dv[32];
        tmem_ld32(tS + lane_off + c * 32, sv);
        tmem_ld32(tDP + lane_off + c * 32, dv);
        tmem_ld_wait();
        uint32_t pp[16], dd[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = fast_exp2(__uint_as_float(sv[i]) * scale_log2 - lse2);
          float p1 = fast_exp2(__uint_as_float(sv[i + 1]) * scale_log2 - lse2);
          const int col = c * 32 + i;
          if (!row_valid || (it == 0 && col > r)) p0 = 0.f;
          if (!row_valid || (it == 0 && col + 1 > r)) p1 = 0.f;
          const float d0 = p0 * (__uint_as_float(dv[i]) - dlt) * scale;
          const float d1 = p1 * (__uint_as_float(dv[i + 1]) - dlt) * scale;
          pp[i >> 1] = pack_bf16(p0, p1);
          dd[i >> 1] = pack_bf16(d0, d1);
        }
        st_tile_row32(sP, r, c * 32, pp);
        st_tile_row32(sDS, r, c * 32, dd);
      }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
      if (it + 1 < n_it) load_stats(w, it + 1, lse2_nx, dlt_nx);
      else load_stats(iw_.peek(), 0, lse2_nx, dlt_nx);
      // dQ tile -> this warp's smem slab -> one TMA reduce-add per 32x32 fp32 block (rows past T are clipped by the
      // tensor map); the L2 does the accumulation in 128-byte lines while the warp goes on with the next tile
      mbar_wait(dq_full, gs & 1);
      tc_fence_after();
      uint8_t* slab = sDQ + (warp - 2) * 4096;
#pragma unroll 1
      for (int c = half * kOutChunks; c < (half + 1) * kOutChunks; ++c) {
        uint32_t v[32];
        tmem_ld32(tDQ + lane_off + c * 32, v);
        tmem_ld_wait();
        if (lane == 0) tma_store_wait_read<0>();   // the previous reduce has finished reading the slab
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(slab + lane * 128 + ((j ^ (lane & 7)) << 4)) =
              make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_reduce_add_3d(&tm_dq, slab, h * HD + c * 32, q0 + q * 32, b);
          tma_store_commit();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_free);
    }
    // dV / dK of this item are complete once every MMA of the item has retired (the K / V release barrier)
    mbar_wait(&kv_empty[item % C::kKvBufs], (item / C::kKvBufs) & 1);
    tc_fence_after();
    uint8_t* slab = sDQ + (warp - 2) * 4096;
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {  // 0: dK -> segment 1, 1: dV -> segment 2
      const uint32_t tsrc = which == 0 ? tDK : tDV;
#pragma unroll 1
      for (int c = half * kOutChunks; c < (half + 1) * kOutChunks; ++c) {
        uint32_t v[32];
        tmem_ld32(tsrc + lane_off + c * 32, v);
        tmem_ld_wait();
        if (lane == 0) tma_store_wait_read<0>();   // the slab's previous contents have been read out
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j) {   // 32 rows x 64 B, 64B-swizzled: 16-byte chunk j of row l at j ^ ((l >> 1) & 3)
          uint4 o;
          o.x = pack_bf16(__uint_as_float(v[8 * j + 0]), __uint_as_float(v[8 * j + 1]));
          o.y = pack_bf16(__uint_as_float(v[8 * j + 2]), __uint_as_float(v[8 * j + 3]));
          o.z = pack_bf16(__uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]));
          o.w = pack_bf16(__uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
          *reinterpret_cast<uint4*>(slab + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) = o;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {   // rows past T are clipped by the tensor map
          tma_store_3d(&tm_dkv, slab, (which + 1) * Dl + h * HD + c * 32, kv0 + q * 32, b);
          tma_store_commit();
        }
      }
    }
    tc_fence_before();   // dK / dV reads are ordered before the next item's first MMAs (via pds_full)
    }
    if (lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------- host
static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
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

// 3-D bf16 map over [B][T][ld] (row-major rows of `cols` valid columns): box = 64 cols x 128 rows x 1 batch.
static int make_tmap_rows3d(CUtensorMap* tm, const void* ptr, int cols, int T, int B, long long ld) {
  auto fn = encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(T), static_cast<cuuint64_t>(B)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 2, static_cast<cuuint64_t>(ld) * 2 * T};
  cuuint32_t box[3] = {64, kTile, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

// 3-D fp32 map over the dQ accumulator [B][T][Dl]: box = 32 cols (128 B, swizzled) x 32 rows x 1 batch.
static int make_tmap_dq(CUtensorMap* tm, const void* ptr, int Dl, int T, int B) {
  auto fn = encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(Dl), static_cast<cuuint64_t>(T), static_cast<cuuint64_t>(B)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(Dl) * 4, static_cast<cuuint64_t>(Dl) * 4 * T};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

// 3-D bf16 map over dqkv [B][T][3*Dl] for the dK / dV stores: box = 32 cols (64 B, swizzled) x 32 rows x 1 batch.
static int make_tmap_dkv(CUtensorMap* tm, const void* ptr, int cols, int T, int B) {
  auto fn = encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(T), static_cast<cuuint64_t>(B)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(cols) * 2, static_cast<cuuint64_t>(cols) * 2 * T};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

template <int HD>
static int launch_fwd(const CUtensorMap& tm, void* out, float* lse, int B, int T, int H, float scale,
                      cudaStream_t stream) {
  using C = FwdCfg<HD>;
  auto kfn = attn_fwd_kernel<HD>;
  DPFS_ENSURE_DYN_SMEM(kfn, C::kSmem);
  dim3 grid((T + kTile - 1) / kTile, H, B);
  kfn<<<(DPFS_COUNT(), grid), kAttnFwdThreads, C::kSmem, stream>>>(tm, reinterpret_cast<__nv_bfloat16*>(out), lse, T, H, H * HD,
                                               scale * kLog2e);
  return static_cast<int>(cudaGetLastError());
}

template <int HD>
static int launch_fwd_persistent(const CUtensorMap& tm, void* out, float* lse, int B, int T, int H, float scale,
                                 cudaStream_t stream) {
  using C = FwdPCfg<HD>;
  auto kfn = attn_fwd_persistent_kernel<HD>;
  DPFS_ENSURE_DYN_SMEM(kfn, C::kSmem);
  const int items = ((T + kTile - 1) / kTile) * H * B;
  const int grid = items < num_sms() ? items : num_sms();
  kfn<<<(DPFS_COUNT(), grid), kAttnFwdThreads, C::kSmem, stream>>>(tm, reinterpret_cast<__nv_bfloat16*>(out), lse, B, T,
                                                                  H, H * HD, scale * kLog2e);
  return static_cast<int>(cudaGetLastError());
}

static int launch_fwd_pp(const CUtensorMap& tm, void* out, float* lse, int B, int T, int H, float scale,
                         cudaStream_t stream) {
  using C = FwdPPCfg;
  auto kfn = attn_fwd_pp_kernel;
  DPFS_ENSURE_DYN_SMEM(kfn, C::kSmem);
  const int nq = (T + kTile - 1) / kTile;
  const int items = ((nq + 1) / 2) * H * B;
  const int grid = items < num_sms() ? items : num_sms();
  kfn<<<(DPFS_COUNT(), grid), kAttnFwdThreads, C::kSmem, stream>>>(tm, reinterpret_cast<__nv_bfloat16*>(out), lse, B, T,
                                                                  H, H * C::HD, scale * kLog2e);
  return static_cast<int>(cudaGetLastError());
}

template <int HD>
static int launch_bwd(const CUtensorMap& tm_qkv, const CUtensorMap& tm_do, const float* lse, const float* delta,
                      float* dq_acc, void* dqkv, int B, int T, int H, float scale, cudaStream_t stream) {
  using C = BwdCfg<HD>;
  auto kfn = attn_bwd_kernel<HD>;
  DPFS_ENSURE_DYN_SMEM(kfn, C::kSmem);
  dim3 grid((T + kTile - 1) / kTile, H, B);
  kfn<<<(DPFS_COUNT(), grid), kAttnBwdThreads, C::kSmem, stream>>>(tm_qkv, tm_do, lse, delta, dq_acc,
                                               reinterpret_cast<__nv_bfloat16*>(dqkv), T, H, H * HD, scale,
                                               scale * kLog2e);
  return static_cast<int>(cudaGetLastError());
}

template <int HD, bool V2 = false>
static int launch_bwd_persistent(const CUtensorMap& tm_qkv, const CUtensorMap& tm_do, const CUtensorMap& tm_dq,
                                 const CUtensorMap& tm_dkv,                                 const float* lse, const float* delta, void* dqkv, int B, int T, int H, float scale,
                                 cudaStream_t stream) {
  using C = BwdCfg<HD>;
  auto kfn = attn_bwd_persistent_kernel<HD, V2>;
  DPFS_ENSURE_DYN_SMEM(kfn, C::kSmemPersistent);
  const int items = ((T + kTile - 1) / kTile) * H * B;
  const int grid = items < num_sms() ? items : num_sms();
  kfn<<<(DPFS_COUNT(), grid), kAttnBwdThreads, C::kSmemPersistent, stream>>>(tm_qkv, tm_do, tm_dq, tm_dkv, lse, delta,
                                                                  reinterpret_cast<__nv_bfloat16*>(dqkv), B, T, H,
                                                                  H * HD, scale, scale * kLog2e);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace dpfs

using namespace dpfs;

extern "C" int dpfs_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int HD, float scale,
                             cudaStream_t stream) {
  if (HD != 64 && HD != 128) return -40;
  if (B <= 0 || T <= 0 || H <= 0) return 0;
  CUtensorMap tm;
  int rc = make_tmap_rows3d(&tm, qkv, 3 * H * HD, T, B, 3LL * H * HD);
  if (rc) return rc;
  static int persistent = -1;
  if (persistent < 0) {
    const char* e = getenv("DPFS_ATTN_PERSISTENT");
    persistent = e ? atoi(e) : 1;
  }
  // two-q-tile ping-pong kernel (hd = 64), see attn_fwd_pp_kernel: 128.5 -> 102.6 us on b32 h8 t1000 (B200,
  // profiles/attn_variants_r2.md), default since it passed the fp32 oracle on hardware; DPFS_ATTN_FWD_PP=0 turns it off
  static int pingpong = -1;
  if (pingpong < 0) {
    const char* e = getenv("DPFS_ATTN_FWD_PP");
    pingpong = e ? atoi(e) : 1;
  }
  if (pingpong && HD == 64) return launch_fwd_pp(tm, out, lse, B, T, H, scale, stream);
  if (persistent)
    return HD == 64 ? launch_fwd_persistent<64>(tm, out, lse, B, T, H, scale, stream)
                    : launch_fwd_persistent<128>(tm, out, lse, B, T, H, scale, stream);
  return HD == 64 ? launch_fwd<64>(tm, out, lse, B, T, H, scale, stream)
                  : launch_fwd<128>(tm, out, lse, B, T, H, scale, stream);
}

// dq_acc: fp32 [B*T][H*HD] scratch (zeroed here); delta: fp32 [B][H][T] scratch.
extern "C" int dpfs_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                             float* dq_acc, void* dqkv, int B, int T, int H, int HD, float scale,
                             cudaStream_t stream) {
  if (HD != 64 && HD != 128) return -40;
  if (B <= 0 || T <= 0 || H <= 0) return 0;
  const int Dl = H * HD;
  const long long M = static_cast<long long>(B) * T;
  CUtensorMap tm_qkv, tm_do;
  int rc = make_tmap_rows3d(&tm_qkv, qkv, 3 * Dl, T, B, 3LL * Dl);
  if (rc) return rc;
  rc = make_tmap_rows3d(&tm_do, d_out, Dl, T, B, Dl);
  if (rc) return rc;
  cudaMemsetAsync(dq_acc, 0, sizeof(float) * M * Dl, stream);
  {
    const long long vecs = M * (Dl / 8);
    long long blocks = (vecs + 255) / 256;
    if (blocks > num_sms() * 16LL) blocks = num_sms() * 16LL;
    attn_delta_kernel<<<(DPFS_COUNT(), static_cast<int>(blocks)), 256, 0, stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(out), reinterpret_cast<const __nv_bfloat16*>(d_out), delta, B, T, H,
        HD);
  }
  static int persistent = -1;
  if (persistent < 0) {
    const char* e = getenv("DPFS_ATTN_PERSISTENT");
    persistent = e ? atoi(e) : 1;
  }
  if (persistent) {
    CUtensorMap tm_dq;
    rc = make_tmap_dq(&tm_dq, dq_acc, Dl, T, B);
    if (rc) return rc;
    CUtensorMap tm_dkv;
    rc = make_tmap_dkv(&tm_dkv, dqkv, 3 * Dl, T, B);
    if (rc) return rc;
    // V2 math path (mask predicates only on the diagonal step): measured on B200 (profiles/attn_variants_r2.md)
    // 4 % faster at hd = 128 (717 -> 687 us, b8 h16 t2048) but 4 % slower at hd = 64, where its S / dP look-ahead
    // does not pay (250 -> 260 us, b32 h8 t1000).  Default: V2 for hd = 128 only; DPFS_ATTN_BWD_V2=0/1 forces a choice.
    static int v2_env = -2;
    if (v2_env == -2) {
      const char* e = getenv("DPFS_ATTN_BWD_V2");
      v2_env = e ? atoi(e) : -1;
    }
    const bool v2 = v2_env >= 0 ? v2_env != 0 : HD == 128;
    if (v2)
      rc = HD == 64 ? launch_bwd_persistent<64, true>(tm_qkv, tm_do, tm_dq, tm_dkv, lse, delta, dqkv, B, T, H, scale,
                                                      stream)
                    : launch_bwd_persistent<128, true>(tm_qkv, tm_do, tm_dq, tm_dkv, lse, delta, dqkv, B, T, H, scale,
                                                       stream);
    else
      rc = HD == 64
               ? launch_bwd_persistent<64>(tm_qkv, tm_do, tm_dq, tm_dkv, lse, delta, dqkv, B, T, H, scale, stream)
               : launch_bwd_persistent<128>(tm_qkv, tm_do, tm_dq, tm_dkv, lse, delta, dqkv, B, T, H, scale, stream);
  } else
    rc = HD == 64 ? launch_bwd<64>(tm_qkv, tm_do, lse, delta, dq_acc, dqkv, B, T, H, scale, stream)
                  : launch_bwd<128>(tm_qkv, tm_do, lse, delta, dq_acc, dqkv, B, T, H, scale, stream);
  if (rc) return rc;
  {
    const long long total = M * (Dl / 8);
    long long blocks = (total + 255) / 256;
    if (blocks > num_sms() * 16LL) blocks = num_sms() * 16LL;
    attn_dq_convert_kernel<<<(DPFS_COUNT(), static_cast<int>(blocks)), 256, 0, stream>>>(
        dq_acc, reinterpret_cast<__nv_bfloat16*>(dqkv), M, Dl);
  }
  return static_cast<int>(cudaGetLastError());
}

// Host-side view of the persistent kernels' work distribution, for the CPU tests: the items CTA `bid` of `grid`
// processes, in order (snake != 0: the reversed-odd-rounds order of the opt-in kernels).  Returns the count.
extern "C" int dpfs_debug_attn_items(int n_items, int grid, int bid, int snake, int* out, int cap) {
  int n = 0;
  for (int k = 0, w = bid; w < n_items; ++k, w = snake ? dpfs::snake_item(k, bid, grid) : w + grid) {
    if (n < cap) out[n] = w;
    ++n;
  }
  return n;
}
