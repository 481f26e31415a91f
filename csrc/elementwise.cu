// Fused memory-bound kernels of the training step: RMSNorm / LayerNorm fwd+bwd, SwiGLU / GELU fwd+bwd, RoPE,
// column sums (bias grads), dtype casts.  Each replaces a chain of ATen elementwise kernels in the reference
// (models/layers.py:151-155, models/model.py:17-31,95) with a single pass over HBM.
#include "common.cuh"
#include "dpfs_b200.h"

#include <stdio.h>

#include <mutex>

namespace dpfs {

// --------------------------------------------------------------------------------------------- helpers
struct alignas(16) F4 {
  float x, y, z, w;
};

template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&v)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x;
  v[1] = t.y;
  v[2] = t.z;
  v[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[4]) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  const float2 a = unpack_bf16(t.x), b = unpack_bf16(t.y);
  v[0] = a.x;
  v[1] = a.y;
  v[2] = b.x;
  v[3] = b.y;
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const float (&v)[4]);
template <>
__device__ __forceinline__ void store4<float>(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[4]) {
  uint2 t;
  t.x = pack_bf16(v[0], v[1]);
  t.y = pack_bf16(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = t;
}

// A bf16 element addressed through an NVLS MULTICAST mapping: one store lands in every rank's copy of the symmetric
// buffer (the NVSwitch replicates it).  Kernels instantiated with this output type all-gather what they produce as
// they produce it ("push" all-gather: the separate all-gather kernel and its read of the local block disappear).
struct McBf16 {
  __nv_bfloat16 v;
};
static_assert(sizeof(McBf16) == 2, "McBf16 must index like bf16");
template <>
__device__ __forceinline__ void store4<McBf16>(McBf16* p, const float (&v)[4]) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(p), "r"(pack_bf16(v[0], v[1])),
               "r"(pack_bf16(v[2], v[3]))
               : "memory");
}

constexpr int kNormThreads = 256;  // 8 warps, one row per warp at a time
constexpr int kMaxNV = 16;         // float4 per lane -> D <= 2048 per warp-row

// --------------------------------------------------------------------------------------------- RMSNorm fwd
// y = scale * x * rsqrt(mean(x^2) + eps); statistics in fp32 (reference models/layers.py:151-155).
template <typename XT, typename YT, int NV>
__global__ void __launch_bounds__(kNormThreads) rmsnorm_fwd_kernel(const XT* __restrict__ x,
                                                                   const float* __restrict__ scale,
                                                                   YT* __restrict__ y, float* __restrict__ rstd_out,
                                                                   int M, int D, float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wpb = kNormThreads / 32;
  float sc[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * 4;
    if (c < D) load4<float>(scale + c, sc[i]);
  }
  for (int row = blockIdx.x * wpb + warp; row < M; row += gridDim.x * wpb) {
    const XT* xr = x + static_cast<long long>(row) * D;
    float v[NV][4];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        load4<XT>(xr + c, v[i]);
        ss += v[i][0] * v[i][0] + v[i][1] * v[i][1] + v[i][2] * v[i][2] + v[i][3] * v[i][3];
      }
    }
    ss = warp_sum(ss);
    const float rstd = rsqrtf(ss / D + eps);
    if (lane == 0 && rstd_out) rstd_out[row] = rstd;
    YT* yr = y + static_cast<long long>(row) * D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = v[i][j] * rstd * sc[i][j];
        store4<YT>(yr + c, o);
      }
    }
  }
}


// --------------------------------------------------------------------------------------------- residual + RMSNorm
// x_out = x + y + bias (fp32 residual stream), h = rmsnorm(x_out) * scale (bf16).  `y` is the all-reduced bf16 output
// of a row-parallel GEMM (reference models/layers.py:51-54 followed by models/model.py:119-120): the residual add,
// the replicated bias and the next layer's norm become one pass.  h / scale may be null (residual add only).
template <int NV, typename HT = __nv_bfloat16>
__global__ void __launch_bounds__(kNormThreads) add_rmsnorm_fwd_kernel(
    const float* __restrict__ x, const __nv_bfloat16* __restrict__ y, const float* __restrict__ bias,
    const float* __restrict__ scale, float* __restrict__ x_out, HT* __restrict__ h,
    float* __restrict__ rstd_out, int M, int D, float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wpb = kNormThreads / 32;
  float sc[NV][4], bs[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) sc[i][j] = bs[i][j] = 0.f;
    if (c < D) {
      if (scale) load4<float>(scale + c, sc[i]);
      if (bias) load4<float>(bias + c, bs[i]);
    }
  }
  for (int row = blockIdx.x * wpb + warp; row < M; row += gridDim.x * wpb) {
    const long long off = static_cast<long long>(row) * D;
    float v[NV][4];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float a[4], b[4];
        load4<float>(x + off + c, a);
        load4<__nv_bfloat16>(y + off + c, b);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[i][j] = a[j] + b[j] + bs[i][j];
          ss += v[i][j] * v[i][j];
        }
        store4<float>(x_out + off + c, v[i]);
      }
    }
    if (h) {
      ss = warp_sum(ss);
      const float rstd = rsqrtf(ss / D + eps);
      if (lane == 0 && rstd_out) rstd_out[row] = rstd;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = (i * 32 + lane) * 4;
        if (c < D) {
          float o[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) o[j] = v[i][j] * rstd * sc[i][j];
          store4<HT>(h + off + c, o);
        }
      }
    }
  }
}

// --------------------------------------------------------------------------------------------- RMSNorm bwd
// dx = dres + rstd * (g - xhat * mean(g * xhat)),  g = dy * scale, xhat = x * rstd;  dscale = sum_rows dy * xhat.
// Each block writes one row of per-block dscale partials into `scratch` [gridDim.x][D]; a second kernel sums them
// in a fixed order (deterministic).
// CS: also emit per-block column sums of the produced dx (the bias gradient of the row-parallel linear whose output
// this residual-stream gradient is) as a second row of partials: scratch is then [gridDim.x][2 * D].
template <typename XT, typename GT, int NV, bool CS = false, typename BT = __nv_bfloat16>
__global__ void __launch_bounds__(kNormThreads) rmsnorm_bwd_kernel(
    const GT* __restrict__ dy, const XT* __restrict__ x, const float* __restrict__ rstd_in,
    const float* __restrict__ scale, const float* __restrict__ dres, float* __restrict__ dx,
    BT* __restrict__ dx_bf16, float* __restrict__ scratch, int M, int D) {
  __shared__ float red[kNormThreads / 32][32 * 4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wpb = kNormThreads / 32;
  const int ldp = CS ? 2 * D : D;  // row pitch of the partials
  float sc[NV][4], ds[NV][4], cs[CS ? NV : 1][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ds[i][j] = 0.f;
      if (CS) cs[i][j] = 0.f;
    }
    if (c < D) load4<float>(scale + c, sc[i]);
  }
  for (int row = blockIdx.x * wpb + warp; row < M; row += gridDim.x * wpb) {
    const long long off = static_cast<long long>(row) * D;
    const float rstd = rstd_in[row];
    float xh[NV][4], g[NV][4], rs[NV][4];
    float dot = 0.f;
    // every global load of the row (x, dy and the residual-stream gradient) is issued before the first use, so the
    // row costs ONE memory round trip instead of two
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) rs[i][j] = 0.f;
      if (c < D) {
        load4<XT>(x + off + c, xh[i]);
        load4<GT>(dy + off + c, g[i]);
        if (dres) load4<float>(dres + off + c, rs[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float xv = xh[i][j] * rstd, dv = g[i][j];
          xh[i][j] = xv;
          g[i][j] = dv * sc[i][j];
          ds[i][j] += dv * xv;
          dot += g[i][j] * xv;
        }
      }
    }
    dot = warp_sum(dot) / D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = rstd * (g[i][j] - xh[i][j] * dot) + rs[i][j];
        if (dx) store4<float>(dx + off + c, o);
        if (dx_bf16) store4<BT>(dx_bf16 + off + c, o);
        if (CS) {
#pragma unroll
          for (int j = 0; j < 4; ++j) cs[i][j] += o[j];
        }
      }
    }
  }
  // block-level reduction of the dscale partials: warps -> smem -> warp 0 sums in fixed order
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) red[warp][lane * 4 + j] = ds[i][j];
    __syncthreads();
    if (warp == 0) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        for (int w = 0; w < wpb; ++w)
#pragma unroll
          for (int j = 0; j < 4; ++j) o[j] += red[w][lane * 4 + j];
        store4<float>(scratch + static_cast<long long>(blockIdx.x) * ldp + c, o);
      }
    }
  }
  if (CS) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 4; ++j) red[warp][lane * 4 + j] = cs[i][j];
      __syncthreads();
      if (warp == 0) {
        const int c = (i * 32 + lane) * 4;
        if (c < D) {
          float o[4] = {0.f, 0.f, 0.f, 0.f};
          for (int w = 0; w < wpb; ++w)
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] += red[w][lane * 4 + j];
          store4<float>(scratch + static_cast<long long>(blockIdx.x) * ldp + D + c, o);
        }
      }
    }
  }
}

// out[c] = sum_b partial[b][c].  Block = 32 columns x 32 row-groups: each warp sums a strided subset of the rows of
// its 32 columns (coalesced 128-byte reads, 8 independent loads in flight: the kernel is pure L2 latency), the 32
// partial sums meet in shared memory in a fixed order.  Columns [split, D) (if any) go to out1[c - split] and are
// never accumulated.
constexpr int kReduceGroups = 32;
__global__ void __launch_bounds__(kReduceGroups * 32) reduce_partials_kernel(const float* __restrict__ partial,
                                                                             float* __restrict__ out, int nblocks,
                                                                             int D, int accumulate,
                                                                             float* __restrict__ out1 = nullptr,
                                                                             int split = 0x7fffffff) {
  __shared__ float sm[kReduceGroups][33];
  const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < D) {
    constexpr int G = kReduceGroups;
    const float* p = partial + c;
    int b = grp;
    for (; b + 7 * G < nblocks; b += 8 * G) {
      float a[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) a[u] = p[static_cast<long long>(b + u * G) * D];
      s += ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
    }
    for (; b < nblocks; b += G) s += p[static_cast<long long>(b) * D];
  }
  sm[grp][lane] = s;
  __syncthreads();
  if (grp == 0 && c < D) {
    float t = 0.f;
#pragma unroll
    for (int g = 0; g < kReduceGroups; ++g) t += sm[g][lane];
    if (c >= split) out1[c - split] = t;
    else out[c] = accumulate ? out[c] + t : t;
  }
}

// --------------------------------------------------------------------------------------------- LayerNorm
// y = (x - mean) * rstd * gamma + beta.  (The reference only has RMSNorm; LayerNorm is provided because the
// north-star names it.)  One warp per row, fp32 statistics.
template <typename XT, typename YT, int NV>
__global__ void __launch_bounds__(kNormThreads) layernorm_fwd_kernel(const XT* __restrict__ x,
                                                                     const float* __restrict__ gamma,
                                                                     const float* __restrict__ beta,
                                                                     YT* __restrict__ y, float* __restrict__ mean_out,
                                                                     float* __restrict__ rstd_out, int M, int D,
                                                                     float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wpb = kNormThreads / 32;
  for (int row = blockIdx.x * wpb + warp; row < M; row += gridDim.x * wpb) {
    const XT* xr = x + static_cast<long long>(row) * D;
    float v[NV][4];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        load4<XT>(xr + c, v[i]);
        s += v[i][0] + v[i][1] + v[i][2] + v[i][3];
      }
    }
    const float mean = warp_sum(s) / D;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float d = v[i][j] - mean;
          ss += d * d;
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(ss) / D + eps);
    if (lane == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    YT* yr = y + static_cast<long long>(row) * D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float gm[4], bt[4] = {0.f, 0.f, 0.f, 0.f}, o[4];
        load4<float>(gamma + c, gm);
        if (beta) load4<float>(beta + c, bt);
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = (v[i][j] - mean) * rstd * gm[j] + bt[j];
        store4<YT>(yr + c, o);
      }
    }
  }
}

// dx = rstd * (g - mean(g) - xhat * mean(g*xhat)), g = dy*gamma; dgamma = sum dy*xhat; dbeta = sum dy.
// scratch: [gridDim.x][2*D] (dgamma partials then dbeta partials).
template <typename XT, typename GT, int NV>
__global__ void __launch_bounds__(kNormThreads) layernorm_bwd_kernel(
    const GT* __restrict__ dy, const XT* __restrict__ x, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, const float* __restrict__ gamma, float* __restrict__ dx,
    float* __restrict__ scratch, int M, int D) {
  __shared__ float red[kNormThreads / 32][32 * 4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wpb = kNormThreads / 32;
  float dg[NV][4], db[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dg[i][j] = db[i][j] = 0.f;
  for (int row = blockIdx.x * wpb + warp; row < M; row += gridDim.x * wpb) {
    const long long off = static_cast<long long>(row) * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[NV][4], g[NV][4];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float xv[4], dv[4], gm[4];
        load4<XT>(x + off + c, xv);
        load4<GT>(dy + off + c, dv);
        load4<float>(gamma + c, gm);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          xh[i][j] = (xv[j] - mean) * rstd;
          g[i][j] = dv[j] * gm[j];
          dg[i][j] += dv[j] * xh[i][j];
          db[i][j] += dv[j];
          s1 += g[i][j];
          s2 += g[i][j] * xh[i][j];
        }
      }
    }
    s1 = warp_sum(s1) / D;
    s2 = warp_sum(s2) / D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * 4;
      if (c < D) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = rstd * (g[i][j] - s1 - xh[i][j] * s2);
        store4<float>(dx + off + c, o);
      }
    }
  }
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 4; ++j) red[warp][lane * 4 + j] = pass == 0 ? dg[i][j] : db[i][j];
      __syncthreads();
      if (warp == 0) {
        const int c = (i * 32 + lane) * 4;
        if (c < D) {
          float o[4] = {0.f, 0.f, 0.f, 0.f};
          for (int w = 0; w < wpb; ++w)
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] += red[w][lane * 4 + j];
          store4<float>(scratch + static_cast<long long>(blockIdx.x) * 2 * D + pass * D + c, o);
        }
      }
    }
  }
}

// --------------------------------------------------------------------------------------------- SwiGLU / GELU
// gu: [M][2F] bf16 with the gate in columns [0,F) and the up projection in [F,2F) (concatenated weights);
// act = silu(gate) * up   (reference models/model.py:95).
// sigmoid through the two SFU approximations (ex2, rcp: ~1 ulp each, far below bf16 resolution) instead of an
// IEEE division: the elementwise SwiGLU kernels were instruction-bound on the division's Newton steps.
__device__ __forceinline__ float fast_sigmoid(float x) {
  // sigmoid(x) = 0.5 * tanh(x / 2) + 0.5: ONE SFU op (tanh.approx, rel. error 2^-11) instead of ex2 + rcp -- the
  // elementwise SwiGLU kernels sit on the SFU pipe (16 ops / clk / SM) otherwise
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  return fmaf(0.5f, t, 0.5f);
}

// Row-blocked: blockIdx.y walks rows, threads cover the F / 8 column groups -- no 64-bit division per element, and two
// rows per thread keep 4 x 16 B loads in flight.
__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu,
                                                         __nv_bfloat16* __restrict__ act, int M, int F) {
  const int groups = F >> 3;
  const int cg = blockIdx.x * blockDim.x + threadIdx.x;
  if (cg >= groups) return;
  const int c = cg * 8;
  for (int row = blockIdx.y * 2; row < M; row += gridDim.y * 2) {
    uint4 gv[2], uv[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (row + r < M) {
        const __nv_bfloat16* p = gu + static_cast<long long>(row + r) * 2 * F + c;
        gv[r] = *reinterpret_cast<const uint4*>(p);
        uv[r] = *reinterpret_cast<const uint4*>(p + F);
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (row + r < M) {
        const uint32_t gw[4] = {gv[r].x, gv[r].y, gv[r].z, gv[r].w}, uw[4] = {uv[r].x, uv[r].y, uv[r].z, uv[r].w};
        uint32_t ow[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 g = unpack_bf16(gw[j]), u = unpack_bf16(uw[j]);
          ow[j] = pack_bf16(g.x * fast_sigmoid(g.x) * u.x, g.y * fast_sigmoid(g.y) * u.y);
        }
        *reinterpret_cast<uint4*>(act + static_cast<long long>(row + r) * F + c) =
            make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
  }
}

// dgu[:, :F] = dact * up * silu'(gate);  dgu[:, F:] = dact * silu(gate)
__global__ void swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* __restrict__ gu,
                                  __nv_bfloat16* __restrict__ dgu, int M, int F) {
  const long long total = static_cast<long long>(M) * (F / 8);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(i / (F / 8));
    const int c = static_cast<int>(i % (F / 8)) * 8;
    const long long go = static_cast<long long>(row) * 2 * F + c;
    const uint4 gv = *reinterpret_cast<const uint4*>(gu + go);
    const uint4 uv = *reinterpret_cast<const uint4*>(gu + go + F);
    const uint4 dv = *reinterpret_cast<const uint4*>(dact + static_cast<long long>(row) * F + c);
    const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w}, uw[4] = {uv.x, uv.y, uv.z, uv.w},
                   dw[4] = {dv.x, dv.y, dv.z, dv.w};
    uint32_t og[4], ou[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 g = unpack_bf16(gw[j]), u = unpack_bf16(uw[j]), d = unpack_bf16(dw[j]);
      const float s0 = fast_sigmoid(g.x), s1 = fast_sigmoid(g.y);
      const float dg0 = d.x * u.x * s0 * (1.f + g.x * (1.f - s0));
      const float dg1 = d.y * u.y * s1 * (1.f + g.y * (1.f - s1));
      og[j] = pack_bf16(dg0, dg1);
      ou[j] = pack_bf16(d.x * g.x * s0, d.y * g.y * s1);
    }
    *reinterpret_cast<uint4*>(dgu + go) = make_uint4(og[0], og[1], og[2], og[3]);
    *reinterpret_cast<uint4*>(dgu + go + F) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
  }
}

// tanh-approximation GELU (GPT-2 flavour), elementwise bf16.  Provided because the north-star names GELU; the
// reference model itself only uses SiLU gating.
__device__ __forceinline__ float gelu_f(float x) {
  const float k = 0.7978845608028654f, a = 0.044715f;
  return 0.5f * x * (1.f + tanhf(k * (x + a * x * x * x)));
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  const float k = 0.7978845608028654f, a = 0.044715f;
  const float t = tanhf(k * (x + a * x * x * x));
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k * (1.f + 3.f * a * x * x);
}
__global__ void gelu_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n8) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint4 xv = *reinterpret_cast<const uint4*>(x + i * 8);
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    uint32_t ow[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 a = unpack_bf16(xw[j]);
      ow[j] = pack_bf16(gelu_f(a.x), gelu_f(a.y));
    }
    *reinterpret_cast<uint4*>(y + i * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  }
}
__global__ void gelu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                __nv_bfloat16* __restrict__ dx, long long n8) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint4 xv = *reinterpret_cast<const uint4*>(x + i * 8);
    const uint4 dv = *reinterpret_cast<const uint4*>(dy + i * 8);
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w}, dw[4] = {dv.x, dv.y, dv.z, dv.w};
    uint32_t ow[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 a = unpack_bf16(xw[j]), d = unpack_bf16(dw[j]);
      ow[j] = pack_bf16(d.x * gelu_grad_f(a.x), d.y * gelu_grad_f(a.y));
    }
    *reinterpret_cast<uint4*>(dx + i * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  }
}

// --------------------------------------------------------------------------------------------- RoPE
// In-place rotary embedding ("rotate_half" convention, reference models/model.py:17-31) on the q and k segments of
// a packed [M][ld] bf16 buffer.  Segment s (0..nseg-1) starts at column s*seg_stride and holds H heads of hd.
// cos/sin tables: fp32 [maxlen][hd/2], indexed by pos[m].  sign = +1 forward, -1 backward (inverse rotation).
__global__ void rope_kernel(__nv_bfloat16* __restrict__ buf, long long ld, const long long* __restrict__ pos,
                            const float* __restrict__ cos_t, const float* __restrict__ sin_t, int M, int nseg,
                            int seg_stride, int H, int hd, float sign) {
  const int half = hd / 2;
  const int vec_per_head = half / 8;
  const long long per_row = static_cast<long long>(nseg) * H * vec_per_head;
  const long long total = static_cast<long long>(M) * per_row;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(i / per_row);
    int r = static_cast<int>(i % per_row);
    const int seg = r / (H * vec_per_head);
    r -= seg * H * vec_per_head;
    const int h = r / vec_per_head;
    const int j0 = (r % vec_per_head) * 8;
    const long long p = pos[row];
    __nv_bfloat16* base = buf + static_cast<long long>(row) * ld + seg * seg_stride + h * hd + j0;
    const uint4 a = *reinterpret_cast<const uint4*>(base);
    const uint4 b = *reinterpret_cast<const uint4*>(base + half);
    const float4 c0 = *reinterpret_cast<const float4*>(cos_t + p * half + j0);
    const float4 c1 = *reinterpret_cast<const float4*>(cos_t + p * half + j0 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(sin_t + p * half + j0);
    const float4 s1 = *reinterpret_cast<const float4*>(sin_t + p * half + j0 + 4);
    const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
    uint32_t oa[4], ob[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 x1 = unpack_bf16(aw[j]), x2 = unpack_bf16(bw[j]);
      const float sA = sign * sn[2 * j], sB = sign * sn[2 * j + 1];
      oa[j] = pack_bf16(x1.x * cs[2 * j] - x2.x * sA, x1.y * cs[2 * j + 1] - x2.y * sB);
      ob[j] = pack_bf16(x2.x * cs[2 * j] + x1.x * sA, x2.y * cs[2 * j + 1] + x1.y * sB);
    }
    *reinterpret_cast<uint4*>(base) = make_uint4(oa[0], oa[1], oa[2], oa[3]);
    *reinterpret_cast<uint4*>(base + half) = make_uint4(ob[0], ob[1], ob[2], ob[3]);
  }
}

// --------------------------------------------------------------------------------------------- column sum
// out[c] = sum_r x[r][c] for bf16 x [M][ld] (bias gradients).  Stage 1 writes per-block partials to scratch
// [nblocks][N]; stage 2 (reduce_partials_kernel) sums them in a fixed order.
constexpr int kColsumRowsPerBlock = 64;
__global__ void colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, long long ld, float* __restrict__ scratch,
                                      int M, int N) {
  // block = (N/8 column groups covered by threadIdx.x .. ) x row strip blockIdx.y
  const int cg = blockIdx.x * blockDim.x + threadIdx.x;  // group of 8 columns
  if (cg * 8 >= N) return;
  const int r0 = blockIdx.y * kColsumRowsPerBlock;
  const int r1 = min(M, r0 + kColsumRowsPerBlock);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int r = r0; r < r1; ++r) {
    const uint4 v = *reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * ld + cg * 8);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16(w[j]);
      acc[2 * j] += f.x;
      acc[2 * j + 1] += f.y;
    }
  }
  float* o = scratch + static_cast<long long>(blockIdx.y) * N + cg * 8;
  *reinterpret_cast<float4*>(o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4*>(o + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}


// --------------------------------------------------------------------------------- fused producers + bias gradients
// The bias gradient of a linear layer is the column sum of its output gradient.  Instead of re-reading that gradient in
// a separate colsum pass, the kernels that PRODUCE it keep per-thread column partials (each thread owns a fixed group
// of 8 columns over a strip of 64 rows), write them to scratch [strips][ncols] and the usual fixed-order second stage
// finishes the sum.

// dgu = swiglu'(dact, gu) and partial column sums of dgu (= gradient of the gate|up biases)
constexpr int kSwigluRowLanes = 4;
__global__ void __launch_bounds__(256) swiglu_bwd_colsum_kernel(const __nv_bfloat16* __restrict__ dact,
                                                                const __nv_bfloat16* __restrict__ gu,
                                                                __nv_bfloat16* __restrict__ dgu,
                                                                float* __restrict__ scratch, int M, int F,
                                                                int strip_rows) {
  // block = blockDim.x column groups x 4 row lanes: lane ty walks rows r0 + ty, r0 + ty + 4, ... of the strip, the four
  // partial sums of a column meet in shared memory in a fixed order (4x the threads in flight of one walker per column)
  __shared__ float red[kSwigluRowLanes][64][17];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int cg = blockIdx.x * blockDim.x + tx;
  const bool active = cg * 8 < F;
  const int r0 = blockIdx.y * strip_rows;
  const int r1 = min(M, r0 + strip_rows);
  float sg[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, su[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (active) {
#pragma unroll 4
    for (int r = r0 + ty; r < r1; r += kSwigluRowLanes) {
      const long long go = static_cast<long long>(r) * 2 * F + cg * 8;
      const uint4 gv = *reinterpret_cast<const uint4*>(gu + go);
      const uint4 uv = *reinterpret_cast<const uint4*>(gu + go + F);
      const uint4 dv = *reinterpret_cast<const uint4*>(dact + static_cast<long long>(r) * F + cg * 8);
      const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w}, uw[4] = {uv.x, uv.y, uv.z, uv.w},
                     dw[4] = {dv.x, dv.y, dv.z, dv.w};
      uint32_t og[4], ou[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 g = unpack_bf16(gw[j]), u = unpack_bf16(uw[j]), d = unpack_bf16(dw[j]);
        const float s0 = fast_sigmoid(g.x), s1 = fast_sigmoid(g.y);
        const float dg0 = d.x * u.x * s0 * (1.f + g.x * (1.f - s0));
        const float dg1 = d.y * u.y * s1 * (1.f + g.y * (1.f - s1));
        const float du0 = d.x * g.x * s0, du1 = d.y * g.y * s1;
        og[j] = pack_bf16(dg0, dg1);
        ou[j] = pack_bf16(du0, du1);
        // sum what is actually stored (bf16-rounded), like a separate colsum pass over dgu would
        const float2 rg = unpack_bf16(og[j]), ru = unpack_bf16(ou[j]);
        sg[2 * j] += rg.x;
        sg[2 * j + 1] += rg.y;
        su[2 * j] += ru.x;
        su[2 * j + 1] += ru.y;
      }
      *reinterpret_cast<uint4*>(dgu + go) = make_uint4(og[0], og[1], og[2], og[3]);
      *reinterpret_cast<uint4*>(dgu + go + F) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    red[ty][tx][e] = sg[e];
    red[ty][tx][8 + e] = su[e];
  }
  __syncthreads();
  if (ty == 0 && active) {
    float o[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) o[e] = ((red[0][tx][e] + red[1][tx][e]) + red[2][tx][e]) + red[3][tx][e];
    float* dst = scratch + static_cast<long long>(blockIdx.y) * 2 * F + cg * 8;
    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(o[4], o[5], o[6], o[7]);
    *reinterpret_cast<float4*>(dst + F) = make_float4(o[8], o[9], o[10], o[11]);
    *reinterpret_cast<float4*>(dst + F + 4) = make_float4(o[12], o[13], o[14], o[15]);
  }
}

// y = bf16(x) and partial column sums of y
constexpr int kCastColsumRows = 16;  // short strips: N is small here (the model width), parallelism comes from M
__global__ void __launch_bounds__(128) cast_colsum_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                          float* __restrict__ scratch, int M, int N) {
  const int cg = blockIdx.x * blockDim.x + threadIdx.x;
  if (cg * 8 >= N) return;
  const int r0 = blockIdx.y * kCastColsumRows;
  const int r1 = min(M, r0 + kCastColsumRows);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int r = r0; r < r1; ++r) {
    const long long o = static_cast<long long>(r) * N + cg * 8;
    const float4 a = *reinterpret_cast<const float4*>(x + o);
    const float4 b = *reinterpret_cast<const float4*>(x + o + 4);
    uint4 p;
    p.x = pack_bf16(a.x, a.y);
    p.y = pack_bf16(a.z, a.w);
    p.z = pack_bf16(b.x, b.y);
    p.w = pack_bf16(b.z, b.w);
    *reinterpret_cast<uint4*>(y + o) = p;
    const uint32_t w[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16(w[j]);
      acc[2 * j] += f.x;
      acc[2 * j + 1] += f.y;
    }
  }
  float* o = scratch + static_cast<long long>(blockIdx.y) * N + cg * 8;
  *reinterpret_cast<float4*>(o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4*>(o + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

// --------------------------------------------------------------------------------------------- casts
__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n4) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(x + i * 4);
    uint2 o;
    o.x = pack_bf16(v.x, v.y);
    o.y = pack_bf16(v.z, v.w);
    *reinterpret_cast<uint2*>(y + i * 4) = o;
  }
}

// cast + push all-gather: y_mc is the multicast address of this rank's block in the landing zone
__global__ void cast_f32_to_bf16_mc_kernel(const float* __restrict__ x, McBf16* __restrict__ y_mc, long long n4) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(x + i * 4);
    const float o[4] = {v.x, v.y, v.z, v.w};
    store4<McBf16>(y_mc + i * 4, o);
  }
}

int num_sms();

static inline int ew_grid(long long work_items, int threads) {
  long long b = (work_items + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

static inline int norm_grid(int M) {
  int b = (M + 7) / 8;
  const int cap = num_sms() * 8;   // as many resident warps as registers allow: these kernels live on bytes in flight
  return b < cap ? (b < 1 ? 1 : b) : cap;
}

}  // namespace dpfs

namespace dpfs {
// out[c] = sum over nrows of partial[r][c] (fixed order); for other translation units that produce partial-sum rows
int reduce_partials_rows(const float* partial, float* out, int nrows, int D, cudaStream_t stream) {
  reduce_partials_kernel<<<(DPFS_COUNT(), (D + 31) / 32), kReduceGroups * 32, 0, stream>>>(partial, out, nrows, D, 0);
  return static_cast<int>(cudaGetLastError());
}
}  // namespace dpfs

using namespace dpfs;

#define DISPATCH_NV(D, ...)                   \
  do {                                        \
    const int nv_ = ((D) + 127) / 128;        \
    if (nv_ <= 1) {                           \
      constexpr int NV = 1;                   \
      __VA_ARGS__;                            \
    } else if (nv_ <= 2) {                    \
      constexpr int NV = 2;                   \
      __VA_ARGS__;                            \
    } else if (nv_ <= 4) {                    \
      constexpr int NV = 4;                   \
      __VA_ARGS__;                            \
    } else if (nv_ <= 8) {                    \
      constexpr int NV = 8;                   \
      __VA_ARGS__;                            \
    } else {                                  \
      constexpr int NV = 16;                  \
      __VA_ARGS__;                            \
    }                                         \
  } while (0)

extern "C" int dpfs_norm_bwd_blocks(int M) { return norm_grid(M); }

extern "C" int dpfs_rmsnorm_fwd(const void* x, int x_bf16, const float* scale, void* y, int y_bf16, float* rstd,
                                int M, int D, float eps, cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    if (x_bf16 && y_bf16)
      rmsnorm_fwd_kernel<__nv_bfloat16, __nv_bfloat16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)x, scale, (__nv_bfloat16*)y, rstd, M, D, eps);
    else if (x_bf16 && !y_bf16)
      rmsnorm_fwd_kernel<__nv_bfloat16, float, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)x, scale, (float*)y, rstd, M, D, eps);
    else if (!x_bf16 && y_bf16)
      rmsnorm_fwd_kernel<float, __nv_bfloat16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const float*)x, scale, (__nv_bfloat16*)y, rstd, M, D, eps);
    else
      rmsnorm_fwd_kernel<float, float, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>((const float*)x, scale, (float*)y,
                                                                              rstd, M, D, eps);
  });
  return static_cast<int>(cudaGetLastError());
}


extern "C" int dpfs_add_rmsnorm_fwd(const float* x, const void* y_bf16, const float* bias, const float* scale,
                                    float* x_out, void* h_bf16, float* rstd, int M, int D, float eps,
                                    cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    add_rmsnorm_fwd_kernel<NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
        x, (const __nv_bfloat16*)y_bf16, bias, scale, x_out, (__nv_bfloat16*)h_bf16, rstd, M, D, eps);
  });
  return static_cast<int>(cudaGetLastError());
}

template <bool CS>
static int rmsnorm_bwd_launch(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* rstd,
                              const float* scale, const float* dres, float* dx, void* dx_bf16, float* dscale,
                              int dscale_accumulate, float* dbias, float* scratch, int M, int D,
                              cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    if (dy_bf16 && !x_bf16)
      rmsnorm_bwd_kernel<float, __nv_bfloat16, NV, CS><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)dy, (const float*)x, rstd, scale, dres, dx, (__nv_bfloat16*)dx_bf16, scratch, M, D);
    else if (dy_bf16 && x_bf16)
      rmsnorm_bwd_kernel<__nv_bfloat16, __nv_bfloat16, NV, CS><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, rstd, scale, dres, dx, (__nv_bfloat16*)dx_bf16, scratch,
          M, D);
    else if (!dy_bf16 && !x_bf16)
      rmsnorm_bwd_kernel<float, float, NV, CS><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const float*)dy, (const float*)x, rstd, scale, dres, dx, (__nv_bfloat16*)dx_bf16, scratch, M, D);
    else
      rmsnorm_bwd_kernel<__nv_bfloat16, float, NV, CS><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const float*)dy, (const __nv_bfloat16*)x, rstd, scale, dres, dx, (__nv_bfloat16*)dx_bf16, scratch, M, D);
  });
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  if (CS)
    reduce_partials_kernel<<<(DPFS_COUNT(), (2 * D + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, dscale, grid, 2 * D,
                                                                                 dscale_accumulate, dbias, D);
  else
    reduce_partials_kernel<<<(DPFS_COUNT(), (D + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, dscale, grid, D,
                                                                             dscale_accumulate);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_rmsnorm_bwd(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* rstd,
                                const float* scale, const float* dres, float* dx, void* dx_bf16, float* dscale,
                                int dscale_accumulate, float* scratch, int M, int D, cudaStream_t stream) {
  return rmsnorm_bwd_launch<false>(dy, dy_bf16, x, x_bf16, rstd, scale, dres, dx, dx_bf16, dscale, dscale_accumulate,
                                   nullptr, scratch, M, D, stream);
}

// Same, plus dbias[c] = sum_rows dx[r][c] (overwritten): the residual-stream gradient this kernel produces is the
// output gradient of the row-parallel linear in front of it, so that layer's bias gradient needs no pass of its own.
// scratch: dpfs_norm_bwd_blocks(M) * 2 * D floats.
extern "C" int dpfs_rmsnorm_bwd_bias(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* rstd,
                                     const float* scale, const float* dres, float* dx, void* dx_bf16, float* dscale,
                                     int dscale_accumulate, float* dbias, float* scratch, int M, int D,
                                     cudaStream_t stream) {
  if (!dbias) return -21;
  return rmsnorm_bwd_launch<true>(dy, dy_bf16, x, x_bf16, rstd, scale, dres, dx, dx_bf16, dscale, dscale_accumulate,
                                  dbias, scratch, M, D, stream);
}

extern "C" int dpfs_layernorm_fwd(const void* x, int x_bf16, const float* gamma, const float* beta, void* y,
                                  int y_bf16, float* mean, float* rstd, int M, int D, float eps,
                                  cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    if (x_bf16 && y_bf16)
      layernorm_fwd_kernel<__nv_bfloat16, __nv_bfloat16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)x, gamma, beta, (__nv_bfloat16*)y, mean, rstd, M, D, eps);
    else if (!x_bf16 && y_bf16)
      layernorm_fwd_kernel<float, __nv_bfloat16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const float*)x, gamma, beta, (__nv_bfloat16*)y, mean, rstd, M, D, eps);
    else if (!x_bf16 && !y_bf16)
      layernorm_fwd_kernel<float, float, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>((const float*)x, gamma, beta,
                                                                                (float*)y, mean, rstd, M, D, eps);
    else
      layernorm_fwd_kernel<__nv_bfloat16, float, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)x, gamma, beta, (float*)y, mean, rstd, M, D, eps);
  });
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_layernorm_bwd(const void* dy, int dy_bf16, const void* x, int x_bf16, const float* mean,
                                  const float* rstd, const float* gamma, float* dx, float* dgamma, float* dbeta,
                                  float* scratch, int M, int D, cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    if (dy_bf16 && x_bf16)
      layernorm_bwd_kernel<__nv_bfloat16, __nv_bfloat16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, mean, rstd, gamma, dx, scratch, M, D);
    else if (dy_bf16 && !x_bf16)
      layernorm_bwd_kernel<float, __nv_bfloat16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const __nv_bfloat16*)dy, (const float*)x, mean, rstd, gamma, dx, scratch, M, D);
    else if (!dy_bf16 && !x_bf16)
      layernorm_bwd_kernel<float, float, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>((const float*)dy, (const float*)x,
                                                                                mean, rstd, gamma, dx, scratch, M, D);
    else
      layernorm_bwd_kernel<__nv_bfloat16, float, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
          (const float*)dy, (const __nv_bfloat16*)x, mean, rstd, gamma, dx, scratch, M, D);
  });
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  // partial rows are [dgamma(D) | dbeta(D)]: reduce as one 2D-wide vector into a strided pair of outputs
  reduce_partials_kernel<<<(DPFS_COUNT(), (2 * D + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, scratch + (long long)grid * 2 * D, grid,
                                                                  2 * D, 0);
  e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  cudaMemcpyAsync(dgamma, scratch + (long long)grid * 2 * D, D * sizeof(float), cudaMemcpyDeviceToDevice, stream);
  if (dbeta)
    cudaMemcpyAsync(dbeta, scratch + (long long)grid * 2 * D + D, D * sizeof(float), cudaMemcpyDeviceToDevice,
                    stream);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_swiglu_fwd(const void* gu, void* act, int M, int F, cudaStream_t stream) {
  if (F % 8 != 0) return -21;
  if (M <= 0) return 0;
  const int groups = F / 8;
  const int threads = groups >= 256 ? 256 : ((groups + 31) / 32) * 32;
  const int gx = (groups + threads - 1) / threads;
  long long gy = (static_cast<long long>(M) + 1) / 2;
  const long long cap = static_cast<long long>(num_sms()) * 8 * 256 / (static_cast<long long>(threads) * gx);
  if (gy > cap) gy = cap < 1 ? 1 : cap;
  if (gy > 65535) gy = 65535;
  swiglu_fwd_kernel<<<(DPFS_COUNT(), dim3(gx, static_cast<unsigned>(gy))), threads, 0, stream>>>(
      (const __nv_bfloat16*)gu, (__nv_bfloat16*)act, M, F);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_swiglu_bwd(const void* dact, const void* gu, void* dgu, int M, int F, cudaStream_t stream) {
  if (F % 8 != 0) return -21;
  if (M <= 0) return 0;
  const long long total = static_cast<long long>(M) * (F / 8);
  swiglu_bwd_kernel<<<(DPFS_COUNT(), ew_grid(total, 256)), 256, 0, stream>>>((const __nv_bfloat16*)dact, (const __nv_bfloat16*)gu,
                                                             (__nv_bfloat16*)dgu, M, F);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_gelu_fwd(const void* x, void* y, long long n, cudaStream_t stream) {
  if (n % 8 != 0) return -21;
  if (n <= 0) return 0;
  gelu_fwd_kernel<<<(DPFS_COUNT(), ew_grid(n / 8, 256)), 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n / 8);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_gelu_bwd(const void* dy, const void* x, void* dx, long long n, cudaStream_t stream) {
  if (n % 8 != 0) return -21;
  if (n <= 0) return 0;
  gelu_bwd_kernel<<<(DPFS_COUNT(), ew_grid(n / 8, 256)), 256, 0, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                                           (__nv_bfloat16*)dx, n / 8);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_rope_inplace(void* buf, long long ld, const long long* pos, const float* cos_t,
                                 const float* sin_t, int M, int nseg, int seg_stride, int H, int hd, int inverse,
                                 cudaStream_t stream) {
  if (hd % 16 != 0 || ld % 8 != 0 || seg_stride % 8 != 0) return -22;
  if (M <= 0) return 0;
  const long long total = static_cast<long long>(M) * nseg * H * (hd / 16);
  rope_kernel<<<(DPFS_COUNT(), ew_grid(total, 256)), 256, 0, stream>>>((__nv_bfloat16*)buf, ld, pos, cos_t, sin_t, M, nseg,
                                                       seg_stride, H, hd, inverse ? -1.f : 1.f);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_colsum_scratch_floats(int M, int N) {
  return ((M + kColsumRowsPerBlock - 1) / kColsumRowsPerBlock) * N;
}

extern "C" int dpfs_colsum_bf16(const void* x, long long ld, float* out, int accumulate, float* scratch, int M, int N,
                                cudaStream_t stream) {
  if (N % 8 != 0 || ld % 8 != 0) return -23;
  if (reinterpret_cast<uintptr_t>(x) & 15) return -26;
  if (M <= 0) return 0;
  const int strips = (M + kColsumRowsPerBlock - 1) / kColsumRowsPerBlock;
  if (strips > 65535) return -27;
  dim3 grid((N / 8 + 127) / 128, strips);
  colsum_partial_kernel<<<(DPFS_COUNT(), grid), 128, 0, stream>>>((const __nv_bfloat16*)x, ld, scratch, M, N);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  reduce_partials_kernel<<<(DPFS_COUNT(), (N + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, out, strips, N, accumulate);
  return static_cast<int>(cudaGetLastError());
}


// dgu = swiglu backward; bias_grad[2F] (+)= column sums of dgu.  scratch: dpfs_colsum_scratch_floats(M, 2F) floats.
// Strip height of the fused SwiGLU backward: a thread walks one 8-column group down a strip, so a narrow F (high TP
// degree) has few threads per strip -- shorter strips keep the number of threads in flight up (F = 256: 32 threads per
// strip, 16-row strips -> 2000 blocks for M = 32000 instead of 500 three-quarters-empty ones).
static inline int swiglu_strip_rows(int F) {
  const int groups = F / 8;
  if (groups >= 64) return kColsumRowsPerBlock;
  const int r = kColsumRowsPerBlock * groups / 64;
  return r < 16 ? 16 : r;
}
extern "C" int dpfs_swiglu_bwd_scratch_floats(int M, int F) {
  const int rows = swiglu_strip_rows(F);
  return ((M + rows - 1) / rows) * 2 * F;
}
extern "C" int dpfs_swiglu_bwd_bias(const void* dact, const void* gu, void* dgu, float* bias_grad, int accumulate,
                                    float* scratch, int M, int F, cudaStream_t stream) {
  if (F % 8 != 0) return -21;
  if (M <= 0) return 0;
  const int rows = swiglu_strip_rows(F);
  const int strips = (M + rows - 1) / rows;
  if (strips > 65535) return -27;
  const int groups = F / 8;
  const int tx = groups >= 64 ? 64 : ((groups + 31) / 32) * 32;   // <= 64: the smem reduction is sized for it
  dim3 grid((groups + tx - 1) / tx, strips), block(tx, kSwigluRowLanes);
  swiglu_bwd_colsum_kernel<<<(DPFS_COUNT(), grid), block, 0, stream>>>((const __nv_bfloat16*)dact,
                                                                      (const __nv_bfloat16*)gu, (__nv_bfloat16*)dgu,
                                                                      scratch, M, F, rows);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  reduce_partials_kernel<<<(DPFS_COUNT(), (2 * F + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, bias_grad, strips, 2 * F,
                                                                               accumulate);
  return static_cast<int>(cudaGetLastError());
}

// y = bf16(x) for fp32 x [M][N] (contiguous); colsum[N] (+)= column sums of y.  scratch: ceil(M / 16) * N floats.
extern "C" int dpfs_cast_colsum(const float* x, void* y, float* colsum, int accumulate, float* scratch, int M, int N,
                                cudaStream_t stream) {
  if (N % 8 != 0) return -23;
  if (M <= 0) return 0;
  const int strips = (M + kCastColsumRows - 1) / kCastColsumRows;
  const int tpb = N / 8 < 128 ? ((N / 8 + 31) / 32) * 32 : 128;
  dim3 grid((N / 8 + tpb - 1) / tpb, strips);
  cast_colsum_kernel<<<(DPFS_COUNT(), grid), tpb, 0, stream>>>(x, (__nv_bfloat16*)y, scratch, M, N);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  reduce_partials_kernel<<<(DPFS_COUNT(), (N + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, colsum, strips, N, accumulate);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_cast_f32_to_bf16(const float* x, void* y, long long n, cudaStream_t stream) {
  if (n % 4 != 0) return -24;
  if (n <= 0) return 0;
  cast_f32_to_bf16_kernel<<<(DPFS_COUNT(), ew_grid(n / 4, 256)), 256, 0, stream>>>(x, (__nv_bfloat16*)y, n / 4);
  return static_cast<int>(cudaGetLastError());
}

// ------------------------------------------------------------------------------------- push all-gather variants
// Sequence parallelism needs the bf16 activations these kernels produce (h = norm(x), the bf16 copy of a
// residual-stream gradient) on EVERY rank.  Instead of writing the local block and running an all-gather kernel over
// it, the *_mc entry points write through the NVLS multicast mapping of the all-gather landing zone: `*_mc` arguments
// are multicast addresses of THIS rank's block (mc_base + zone offset + rank * block bytes).  The caller runs
// dpfs_comm_barrier afterwards; every rank then holds all n blocks.  fp32 outputs (x_out, dx, rstd, dscale) stay local.
extern "C" int dpfs_rmsnorm_fwd_mc(const float* x, const float* scale, void* y_mc, float* rstd, int M, int D, float eps,
                                   cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    rmsnorm_fwd_kernel<float, McBf16, NV><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
        x, scale, reinterpret_cast<McBf16*>(y_mc), rstd, M, D, eps);
  });
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_add_rmsnorm_fwd_mc(const float* x, const void* y_bf16, const float* bias, const float* scale,
                                       float* x_out, void* h_mc, float* rstd, int M, int D, float eps,
                                       cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (!h_mc || !scale) return -21;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    add_rmsnorm_fwd_kernel<NV, McBf16><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
        x, (const __nv_bfloat16*)y_bf16, bias, scale, x_out, reinterpret_cast<McBf16*>(h_mc), rstd, M, D, eps);
  });
  return static_cast<int>(cudaGetLastError());
}

// dy bf16, x fp32 (the sequence-parallel layer's combination); dx (fp32, local, nullable) and dx_mc (pushed bf16 copy)
extern "C" int dpfs_rmsnorm_bwd_mc(const void* dy_bf16, const float* x, const float* rstd, const float* scale,
                                   const float* dres, float* dx, void* dx_mc, float* dscale, int dscale_accumulate,
                                   float* scratch, int M, int D, cudaStream_t stream) {
  if (D % 4 != 0 || D > 128 * kMaxNV) return -20;
  if (!dx_mc) return -21;
  if (M <= 0) return 0;
  const int grid = norm_grid(M);
  DISPATCH_NV(D, {
    rmsnorm_bwd_kernel<float, __nv_bfloat16, NV, false, McBf16><<<(DPFS_COUNT(), grid), kNormThreads, 0, stream>>>(
        (const __nv_bfloat16*)dy_bf16, x, rstd, scale, dres, dx, reinterpret_cast<McBf16*>(dx_mc), scratch, M, D);
  });
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  reduce_partials_kernel<<<(DPFS_COUNT(), (D + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, dscale, grid, D,
                                                                                       dscale_accumulate);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int dpfs_cast_f32_to_bf16_mc(const float* x, void* y_mc, long long n, cudaStream_t stream) {
  if (n % 4 != 0) return -24;
  if (n <= 0) return 0;
  cast_f32_to_bf16_mc_kernel<<<(DPFS_COUNT(), ew_grid(n / 4, 256)), 256, 0, stream>>>(
      x, reinterpret_cast<McBf16*>(y_mc), n / 4);
  return static_cast<int>(cudaGetLastError());
}

// ------------------------------------------------------------------------------------- allocation-free column sum
// Some hosts of this library cannot give it scratch memory: a serving process whose own allocator already owns the
// whole GPU (any cudaMalloc may fail, and it synchronises the device), or a plain C / embedded caller that has no
// device allocator at all and only holds the input and output buffers it was handed.  For them the two-stage column
// sum is also offered with its partial-sum scratch living in a fixed array inside the library image.
//
// The array costs DPFS_COLSUM_STATIC_MAX_FLOATS * 4 bytes (8 MiB) on EVERY GPU that loads libdpfs_b200.so, for the
// life of the process, whether or not this entry point is ever called.
static __device__ __align__(256) float g_colsum_static_scratch[DPFS_COLSUM_STATIC_MAX_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
float* g_static_scratch_ptr[kMaxDevices] = {nullptr};
std::mutex g_static_scratch_mutex;

// In host code `g_colsum_static_scratch` is only a stand-in for the symbol, NOT a device address, so it must never be
// handed to a kernel.  Resolve the real address once per device (with that device current) and cache it.
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= kMaxDevices) return -26;
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_colsum_static_scratch);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (p == nullptr) return -27;
    g_static_scratch_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}
}  // namespace

// Same result as dpfs_colsum_bf16 (same kernels, same summation order) without any caller-provided scratch and
// without allocating: no cudaMalloc / cudaFree / allocating library call happens here.
//   * Size limit: ceil(M / 64) * N <= DPFS_COLSUM_STATIC_MAX_FLOATS, otherwise -25 is returned and nothing runs.
//   * The first call on each device looks up the scratch address (cudaGetSymbolAddress); make that first call
//     before starting any stream capture on that device.
//   * All calls share the one array and it keeps whatever the previous call left in it.  Nothing has to be cleared:
//     stage 1 overwrites every slot that stage 2 reads.  For the same reason two calls that may run on the GPU at the
//     same time (different streams, no ordering between them) race on the array — serialise them.
extern "C" int dpfs_colsum_bf16_static(const void* x, long long ld, float* out, int accumulate, int M, int N,
                                       cudaStream_t stream) {
  if (N % 8 != 0 || ld % 8 != 0) return -23;
  if (reinterpret_cast<uintptr_t>(x) & 15) return -26;   // the kernel reads 16-byte vectors
  if (M <= 0) return 0;
  const long long strips = (static_cast<long long>(M) + kColsumRowsPerBlock - 1) / kColsumRowsPerBlock;
  if (strips > 65535) return -27;                        // strips ride in gridDim.y
  if (strips * N > static_cast<long long>(DPFS_COLSUM_STATIC_MAX_FLOATS)) {
    fprintf(stderr,
            "dpfs_colsum_bf16_static: M=%d N=%d needs %lld scratch floats but the built-in scratch holds %d; "
            "use dpfs_colsum_bf16 with caller-provided scratch\n",
            M, N, strips * N, DPFS_COLSUM_STATIC_MAX_FLOATS);
    return -25;
  }
  float* scratch = nullptr;
  const int rc = static_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  dim3 grid((N / 8 + 127) / 128, static_cast<unsigned>(strips));
  colsum_partial_kernel<<<(DPFS_COUNT(), grid), 128, 0, stream>>>((const __nv_bfloat16*)x, ld, scratch, M, N);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  reduce_partials_kernel<<<(DPFS_COUNT(), (N + 31) / 32), kReduceGroups * 32, 0, stream>>>(scratch, out, static_cast<int>(strips),
                                                                             N, accumulate);
  return static_cast<int>(cudaGetLastError());
}
