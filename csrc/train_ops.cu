// Training-step kernels that are not plain elementwise maps: vocab-parallel embedding lookup (+ scatter-add
// backward), vocab-parallel softmax cross-entropy (two-phase, stats exchanged between ranks), fused flat Adam/AdamW.
#include "common.cuh"
#include "dpfs_b200.h"

namespace dpfs {

int num_sms();

// --------------------------------------------------------------------------------------------- embedding
// out[m][:] = weight[ids[m] - vocab_start][:] if vocab_start <= ids[m] < vocab_end else 0
// (masked lookup of reference models/layers.py:134-140, without mutating ids — reference defect D3).
template <typename OT>
__global__ void embedding_fwd_kernel(const long long* __restrict__ ids, const float* __restrict__ weight,
                                     OT* __restrict__ out, int M, int D, long long vocab_start,
                                     long long vocab_end) {
  const int vec = D / 4;
  const long long total = static_cast<long long>(M) * vec;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(i / vec);
    const int c = static_cast<int>(i % vec) * 4;
    const long long id = ids[row];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (id >= vocab_start && id < vocab_end)
      v = *reinterpret_cast<const float4*>(weight + (id - vocab_start) * D + c);
    if constexpr (sizeof(OT) == 4) {
      *reinterpret_cast<float4*>(out + static_cast<long long>(row) * D + c) = v;
    } else {
      uint2 o;
      o.x = pack_bf16(v.x, v.y);
      o.y = pack_bf16(v.z, v.w);
      *reinterpret_cast<uint2*>(out + static_cast<long long>(row) * D + c) = o;
    }
  }
}

// Vocab-parallel lookup fused with its collective (SURVEY K-c3 "vocab_embed_allreduce"): exactly one rank owns each
// token's row, so the all-reduce of the masked lookups degenerates into a scatter.  The owner writes the bf16 row
// through the NVLS multicast mapping of the output (it lands in every rank's copy); the other ranks write nothing and
// nothing is summed.  The caller follows with a rank barrier.  out_mc: multicast address of the [M][D] bf16 output.
__global__ void embedding_fwd_mc_kernel(const long long* __restrict__ ids, const float* __restrict__ weight,
                                        __nv_bfloat16* __restrict__ out_mc, int M, int D, long long vocab_start,
                                        long long vocab_end) {
  const int vec = D / 4;
  const long long total = static_cast<long long>(M) * vec;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(i / vec);
    const int c = static_cast<int>(i % vec) * 4;
    const long long id = ids[row];
    if (id < vocab_start || id >= vocab_end) continue;   // another rank owns this token
    const float4 v = *reinterpret_cast<const float4*>(weight + (id - vocab_start) * D + c);
    asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(out_mc +
                                                                                  static_cast<long long>(row) * D + c),
                 "r"(pack_bf16(v.x, v.y)), "r"(pack_bf16(v.z, v.w))
                 : "memory");
  }
}

// dweight[ids[m] - vocab_start][:] += dout[m][:] for owned ids (dweight must be zeroed by the caller).
template <typename GT>
__global__ void embedding_bwd_kernel(const long long* __restrict__ ids, const GT* __restrict__ dout,
                                     float* __restrict__ dweight, int M, int D, long long vocab_start,
                                     long long vocab_end) {
  const int vec = D / 4;
  const long long total = static_cast<long long>(M) * vec;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(i / vec);
    const int c = static_cast<int>(i % vec) * 4;
    const long long id = ids[row];
    if (id < vocab_start || id >= vocab_end) continue;
    float4 g;
    if constexpr (sizeof(GT) == 4) {
      g = *reinterpret_cast<const float4*>(dout + static_cast<long long>(row) * D + c);
    } else {
      const uint2 u = *reinterpret_cast<const uint2*>(dout + static_cast<long long>(row) * D + c);
      const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y);
      g = make_float4(a.x, a.y, b.x, b.y);
    }
    float* w = dweight + (id - vocab_start) * D + c;
    atomicAdd(w + 0, g.x);
    atomicAdd(w + 1, g.y);
    atomicAdd(w + 2, g.z);
    atomicAdd(w + 3, g.w);
  }
}

// --------------------------------------------------------------------------------------------- cross-entropy
// Vocab-parallel softmax cross-entropy in two phases so the full [M, V] logits never exist on one rank
// (replaces all_gather + F.cross_entropy on replicated logits, reference train.py:101-104 / models/layers.py:99).
//
// phase 1 (local): stats[m] = {max_j x[m][j], sum_j exp(x[m][j] - max), x[m][target - vocab_start] or 0}
// exchange       : all-gather the [M][3] stats of the n ranks (tiny)
// phase 2 (local): combine the n stats per row -> loss_m; dlogits = (softmax - onehot) / n_valid (bf16);
//                  accumulates sum(loss_m) and n_valid into `loss_acc` {loss_sum, n_valid}.
constexpr int kCeThreads = 256;

template <typename LT>
__global__ void __launch_bounds__(kCeThreads) ce_stats_kernel(const LT* __restrict__ logits, long long ld,
                                                              const long long* __restrict__ targets,
                                                              float* __restrict__ stats, int M, int Vl,
                                                              long long vocab_start, long long ignore_index) {
  __shared__ float sred[kCeThreads / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int row = blockIdx.x; row < M; row += gridDim.x) {
    const LT* x = logits + static_cast<long long>(row) * ld;
    float mx = -INFINITY;
    for (int j = threadIdx.x; j < Vl; j += kCeThreads) mx = fmaxf(mx, static_cast<float>(x[j]));
    mx = warp_max(mx);
    if (lane == 0) sred[warp] = mx;
    __syncthreads();
    mx = sred[0];
#pragma unroll
    for (int w = 1; w < kCeThreads / 32; ++w) mx = fmaxf(mx, sred[w]);
    __syncthreads();
    float s = 0.f;
    for (int j = threadIdx.x; j < Vl; j += kCeThreads) s += __expf(static_cast<float>(x[j]) - mx);
    s = warp_sum(s);
    if (lane == 0) sred[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kCeThreads / 32; ++w) tot += sred[w];
      const long long t = targets[row];
      float tl = 0.f;
      if (t != ignore_index && t >= vocab_start && t < vocab_start + Vl) tl = static_cast<float>(x[t - vocab_start]);
      stats[row * 3 + 0] = mx;
      stats[row * 3 + 1] = tot;
      stats[row * 3 + 2] = tl;
    }
    __syncthreads();
  }
}

// n_valid = #targets != ignore_index (same on every rank) -> counts[0]
__global__ void ce_count_kernel(const long long* __restrict__ targets, int M, long long ignore_index,
                                float* __restrict__ loss_acc) {
  __shared__ int sred[8];
  int c = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < M; i += gridDim.x * blockDim.x)
    c += targets[i] != ignore_index;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < blockDim.x / 32; ++w) t += sred[w];
    if (t) atomicAdd(loss_acc + 1, static_cast<float>(t));
  }
}

// combine the per-rank stats of every row: loss_acc[0] += sum of row losses (valid rows); optional per-row loss
__global__ void ce_loss_kernel(const long long* __restrict__ targets, const float* __restrict__ all_stats, int nranks,
                               float* __restrict__ loss_acc, float* __restrict__ row_loss, int M,
                               long long ignore_index) {
  __shared__ float sred[8];
  float acc = 0.f;
  for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < M; row += gridDim.x * blockDim.x) {
    float l = 0.f;
    if (targets[row] != ignore_index) {
      float gmax = -INFINITY;
      for (int r = 0; r < nranks; ++r) gmax = fmaxf(gmax, all_stats[(static_cast<long long>(r) * M + row) * 3]);
      float gsum = 0.f, gt = 0.f;
      for (int r = 0; r < nranks; ++r) {
        const float* st = all_stats + (static_cast<long long>(r) * M + row) * 3;
        gsum += st[1] * __expf(st[0] - gmax);
        gt += st[2];
      }
      l = logf(gsum) + gmax - gt;
    }
    if (row_loss) row_loss[row] = l;
    acc += l;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < blockDim.x / 32; ++w) t += sred[w];
    if (t != 0.f) atomicAdd(loss_acc, t);
  }
}

// dlogits[m][j] = (softmax_j - [j == target]) * upstream / n_valid  (bf16), zero rows for ignored targets
template <typename LT>
__global__ void __launch_bounds__(kCeThreads) ce_grad_kernel(const LT* __restrict__ logits, long long ld,
                                                             const long long* __restrict__ targets,
                                                             const float* __restrict__ all_stats, int nranks,
                                                             __nv_bfloat16* __restrict__ dlogits, long long ldd,
                                                             const float* __restrict__ loss_acc,
                                                             const float* __restrict__ upstream, int M, int Vl,
                                                             long long vocab_start, long long ignore_index) {
  __shared__ float sh[2];
  const float n_valid = loss_acc[1];
  const float g = (n_valid > 0.f ? 1.f / n_valid : 0.f) * (upstream ? upstream[0] : 1.f);
  for (int row = blockIdx.x; row < M; row += gridDim.x) {
    const long long t = targets[row];
    if (threadIdx.x == 0) {
      float gmax = -INFINITY;
      for (int r = 0; r < nranks; ++r) gmax = fmaxf(gmax, all_stats[(static_cast<long long>(r) * M + row) * 3]);
      float gsum = 0.f;
      for (int r = 0; r < nranks; ++r) {
        const float* st = all_stats + (static_cast<long long>(r) * M + row) * 3;
        gsum += st[1] * __expf(st[0] - gmax);
      }
      sh[0] = gmax;
      sh[1] = gsum;
    }
    __syncthreads();
    const float gmax = sh[0];
    const float inv_sum = 1.f / sh[1];
    const LT* x = logits + static_cast<long long>(row) * ld;
    __nv_bfloat16* d = dlogits + static_cast<long long>(row) * ldd;
    const bool valid = t != ignore_index;
    const long long tl = t - vocab_start;
    const float sc = valid ? g : 0.f;
    for (int j = threadIdx.x; j < Vl; j += kCeThreads) {
      float p = __expf(static_cast<float>(x[j]) - gmax) * inv_sum;
      if (j == tl) p -= 1.f;
      d[j] = __float2bfloat16(p * sc);
    }
    __syncthreads();
  }
}

// --------------------------------------------------------------------------------------------- Adam / AdamW
// One launch over a flat fp32 parameter arena (torch.optim.Adam semantics, reference train.py:83,106-109).
// hyper (device, fp32): {lr, beta1, beta2, eps, bias_correction1, bias_correction2, weight_decay(decoupled)}
// so lr and beta1 — which OneCycleLR changes every step — are runtime values and the launch is graph-capturable.
// Also refreshes the bf16 shadow copy of the parameters that the GEMMs read.
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, long long n4,
                            const float* __restrict__ hyper) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], bc1 = hyper[4], bc2 = hyper[5],
              wd = hyper[6];
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 pv = *reinterpret_cast<float4*>(p + i * 4);
    const float4 gv = *reinterpret_cast<const float4*>(g + i * 4);
    float4 mv = *reinterpret_cast<float4*>(m + i * 4);
    float4 vv = *reinterpret_cast<float4*>(v + i * 4);
    float pa[4] = {pv.x, pv.y, pv.z, pv.w};
    const float ga[4] = {gv.x, gv.y, gv.z, gv.w};
    float ma[4] = {mv.x, mv.y, mv.z, mv.w};
    float va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (wd != 0.f) pa[j] *= (1.f - lr * wd);
      ma[j] = b1 * ma[j] + (1.f - b1) * ga[j];
      va[j] = b2 * va[j] + (1.f - b2) * ga[j] * ga[j];
      const float denom = sqrtf(va[j]) * inv_sqrt_bc2 + eps;
      pa[j] -= step_size * (ma[j] / denom);
    }
    *reinterpret_cast<float4*>(p + i * 4) = make_float4(pa[0], pa[1], pa[2], pa[3]);
    *reinterpret_cast<float4*>(m + i * 4) = make_float4(ma[0], ma[1], ma[2], ma[3]);
    *reinterpret_cast<float4*>(v + i * 4) = make_float4(va[0], va[1], va[2], va[3]);
    if (shadow) {
      uint2 o;
      o.x = pack_bf16(pa[0], pa[1]);
      o.y = pack_bf16(pa[2], pa[3]);
      *reinterpret_cast<uint2*>(shadow + i * 4) = o;
    }
  }
}

static inline int ew_grid2(long long work_items, int threads) {
  long long b = (work_items + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace dpfs

using namespace dpfs;

extern "C" int dpfs_embedding_fwd(const long long* ids, const float* weight, void* out, int out_bf16, int M, int D,
                                  long long vocab_start, long long vocab_end, cudaStream_t stream) {
  if (D % 4 != 0) return -30;
  if (M <= 0) return 0;
  const long long total = static_cast<long long>(M) * (D / 4);
  if (out_bf16)
    embedding_fwd_kernel<__nv_bfloat16><<<(DPFS_COUNT(), ew_grid2(total, 256)), 256, 0, stream>>>(ids, weight, (__nv_bfloat16*)out, M,
                                                                                 D, vocab_start, vocab_end);
  else
    embedding_fwd_kernel<float><<<(DPFS_COUNT(), ew_grid2(total, 256)), 256, 0, stream>>>(ids, weight, (float*)out, M, D, vocab_start,
                                                                         vocab_end);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_embedding_fwd_mc(const long long* ids, const float* weight, void* out_mc, int M, int D,
                                     long long vocab_start, long long vocab_end, cudaStream_t stream) {
  if (D % 4 != 0) return -30;
  if (M <= 0) return 0;
  const long long total = static_cast<long long>(M) * (D / 4);
  embedding_fwd_mc_kernel<<<(DPFS_COUNT(), ew_grid2(total, 256)), 256, 0, stream>>>(
      ids, weight, reinterpret_cast<__nv_bfloat16*>(out_mc), M, D, vocab_start, vocab_end);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_embedding_bwd(const long long* ids, const float* dout, float* dweight, int zero_first,
                                  long long local_vocab, int M, int D, long long vocab_start, long long vocab_end,
                                  cudaStream_t stream) {
  if (D % 4 != 0) return -30;
  if (zero_first) cudaMemsetAsync(dweight, 0, sizeof(float) * local_vocab * D, stream);
  if (M <= 0) return 0;
  const long long total = static_cast<long long>(M) * (D / 4);
  embedding_bwd_kernel<float><<<(DPFS_COUNT(), ew_grid2(total, 256)), 256, 0, stream>>>(ids, dout, dweight, M, D,
                                                                                        vocab_start, vocab_end);
  return static_cast<int>(cudaGetLastError());
}

// same with a bf16 output gradient (the sequence-parallel path all-gathers the residual-stream gradient in bf16)
extern "C" int dpfs_embedding_bwd_bf16(const long long* ids, const void* dout_bf16, float* dweight, int zero_first,
                                       long long local_vocab, int M, int D, long long vocab_start,
                                       long long vocab_end, cudaStream_t stream) {
  if (D % 4 != 0) return -30;
  if (zero_first) cudaMemsetAsync(dweight, 0, sizeof(float) * local_vocab * D, stream);
  if (M <= 0) return 0;
  const long long total = static_cast<long long>(M) * (D / 4);
  embedding_bwd_kernel<__nv_bfloat16><<<(DPFS_COUNT(), ew_grid2(total, 256)), 256, 0, stream>>>(
      ids, static_cast<const __nv_bfloat16*>(dout_bf16), dweight, M, D, vocab_start, vocab_end);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_ce_stats(const void* logits, int logits_bf16, long long ld, const long long* targets,
                             float* stats, float* loss_acc, int M, int Vl, long long vocab_start,
                             long long ignore_index, cudaStream_t stream) {
  if (M <= 0) return 0;
  cudaMemsetAsync(loss_acc, 0, 2 * sizeof(float), stream);
  ce_count_kernel<<<(DPFS_COUNT(), (M + 255) / 256 > 64 ? 64 : (M + 255) / 256), 256, 0, stream>>>(targets, M, ignore_index,
                                                                                  loss_acc);
  const int grid = M < num_sms() * 8 ? M : num_sms() * 8;
  if (logits_bf16)
    ce_stats_kernel<__nv_bfloat16><<<(DPFS_COUNT(), grid), kCeThreads, 0, stream>>>((const __nv_bfloat16*)logits, ld, targets, stats,
                                                                   M, Vl, vocab_start, ignore_index);
  else
    ce_stats_kernel<float><<<(DPFS_COUNT(), grid), kCeThreads, 0, stream>>>((const float*)logits, ld, targets, stats, M, Vl,
                                                            vocab_start, ignore_index);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_ce_loss(const long long* targets, const float* all_stats, int nranks, float* loss_acc,
                            float* row_loss, int M, long long ignore_index, cudaStream_t stream) {
  if (M <= 0) return 0;
  int grid = (M + 255) / 256;
  if (grid > num_sms() * 4) grid = num_sms() * 4;
  ce_loss_kernel<<<(DPFS_COUNT(), grid), 256, 0, stream>>>(targets, all_stats, nranks, loss_acc, row_loss, M, ignore_index);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_ce_grad(const void* logits, int logits_bf16, long long ld, const long long* targets,
                            const float* all_stats, int nranks, void* dlogits, long long ldd, const float* loss_acc,
                            const float* upstream, int M, int Vl, long long vocab_start, long long ignore_index,
                            cudaStream_t stream) {
  if (M <= 0) return 0;
  const int grid = M < num_sms() * 8 ? M : num_sms() * 8;
  if (logits_bf16)
    ce_grad_kernel<__nv_bfloat16><<<(DPFS_COUNT(), grid), kCeThreads, 0, stream>>>(
        (const __nv_bfloat16*)logits, ld, targets, all_stats, nranks, (__nv_bfloat16*)dlogits, ldd, loss_acc,
        upstream, M, Vl, vocab_start, ignore_index);
  else
    ce_grad_kernel<float><<<(DPFS_COUNT(), grid), kCeThreads, 0, stream>>>((const float*)logits, ld, targets, all_stats, nranks,
                                                           (__nv_bfloat16*)dlogits, ldd, loss_acc, upstream, M, Vl,
                                                           vocab_start, ignore_index);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_adam_step(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n,
                              const float* hyper, cudaStream_t stream) {
  if (n % 4 != 0) return -31;
  if (n <= 0) return 0;
  adam_kernel<<<(DPFS_COUNT(), ew_grid2(n / 4, 256)), 256, 0, stream>>>(p, g, m, v, (__nv_bfloat16*)shadow_bf16, n / 4, hyper);
  return static_cast<int>(cudaGetLastError());
}
