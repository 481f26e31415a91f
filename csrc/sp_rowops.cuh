// Row operations of the sequence-parallel hand-offs, shared by the stand-alone kernels (sp_fused.cu) and by the comm
// warps of the fused GEMM + hand-off kernel (gemm_handoff.cu).  One warp owns R consecutive rows at a time; a lane owns
// the 16-byte chunks (8 bf16 columns) j * 32 + lane, j < NC, of a row (D <= NC * 256).  Every load of the R rows is
// issued before the first use, so the rows cost one NVLink round trip, not R.
#pragma once

#include "comm.cuh"

namespace dpfs {

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = unpack_bf16(w[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
__device__ __forceinline__ void ld8f(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
}
__device__ __forceinline__ void st8f(float* p, const float (&f)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
}

// per-lane column constants: 8 floats per owned chunk (zeros where the chunk is out of range or the pointer is null)
template <int NC>
__device__ __forceinline__ void sp_load_cols(const float* p, int nchunk, int lane, float (&v)[NC][8]) {
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    const int ch = j * 32 + lane;
#pragma unroll
    for (int e = 0; e < 8; ++e) v[j][e] = 0.f;
    if (p && ch < nchunk) ld8f(p + ch * 8, v[j]);
  }
}

// ---- forward: x_out = x + sum_r partial_r + bias (fp32, private rows); h = rmsnorm(x_out) * scale -> every rank.
//   mc_part_row0 : multicast address of the partial row that corresponds to local row `row`
//   mc_h_row0    : multicast address of the same row in the landing zone (nullptr: residual add only)
//   rows [row, row + R) of the rank, valid while < row_end
template <int NC, int R>
__device__ __forceinline__ void sp_fwd_rows(const uint8_t* mc_part_row0, uint8_t* mc_h_row0, const float* x_l,
                                            float* xout_l, float* rstd_l, int row, int row_end, int D, float eps,
                                            int lane, const float (&sc)[NC][8], const float (&bs)[NC][8]) {
  const int nchunk = D >> 3;
  const long long rowb = static_cast<long long>(D) * 2;
  uint4 p[R][NC];
  float v[R][NC][8];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (row + r < row_end) {
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const int ch = j * 32 + lane;
        if (ch < nchunk) p[r][j] = multimem_ld_reduce_bf16x8(mc_part_row0 + r * rowb + ch * 16);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (row + r < row_end) {
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const int ch = j * 32 + lane;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[r][j][e] = 0.f;
        if (ch < nchunk && x_l) ld8f(x_l + static_cast<long long>(row + r) * D + ch * 8, v[r][j]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (row + r >= row_end) continue;
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      const int ch = j * 32 + lane;
      if (ch < nchunk) {
        float y[8];
        unpack8(p[r][j], y);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          v[r][j][e] += y[e] + bs[j][e];
          ss += v[r][j][e] * v[r][j][e];
        }
        st8f(xout_l + static_cast<long long>(row + r) * D + ch * 8, v[r][j]);
      }
    }
    if (mc_h_row0) {
      ss = warp_sum(ss);
      const float rstd = rsqrtf(ss / D + eps);
      if (lane == 0) rstd_l[row + r] = rstd;
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const int ch = j * 32 + lane;
        if (ch < nchunk) {
          float o[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = v[r][j][e] * rstd * sc[j][e];
          multimem_st_16B(mc_h_row0 + r * rowb + ch * 16, pack8(o));
        }
      }
    }
  }
}

// ---- backward: dh = sum_r partial_r; dx = dres + rstd * (g - xhat * mean(g * xhat)), g = dh * scale, xhat = x * rstd;
// dx -> fp32 private rows (dx_l, nullable) and bf16 -> every rank (mc_dx_row0, nullable).  ds += dh * xhat and
// cs += dx accumulate this lane's column partial sums over the rows it has seen.
template <int NC, int R>
__device__ __forceinline__ void sp_bwd_rows(const uint8_t* mc_part_row0, uint8_t* mc_dx_row0, const float* x_l,
                                            const float* rstd_l, const float* dres_l, float* dx_l, int row,
                                            int row_end, int D, int lane, const float (&sc)[NC][8],
                                            float (&ds)[NC][8], float (&cs)[NC][8]) {
  const int nchunk = D >> 3;
  const long long rowb = static_cast<long long>(D) * 2;
  uint4 p[R][NC];
  float xh[R][NC][8], rs[R][NC][8];
  float rstd[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    rstd[r] = 0.f;
    if (row + r < row_end) {
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const int ch = j * 32 + lane;
        if (ch < nchunk) p[r][j] = multimem_ld_reduce_bf16x8(mc_part_row0 + r * rowb + ch * 16);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (row + r < row_end) {
      rstd[r] = rstd_l[row + r];
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const int ch = j * 32 + lane;
#pragma unroll
        for (int e = 0; e < 8; ++e) xh[r][j][e] = rs[r][j][e] = 0.f;
        if (ch < nchunk) {
          const long long off = static_cast<long long>(row + r) * D + ch * 8;
          ld8f(x_l + off, xh[r][j]);
          if (dres_l) ld8f(dres_l + off, rs[r][j]);
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (row + r >= row_end) continue;
    float g[NC][8];
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      const int ch = j * 32 + lane;
      if (ch < nchunk) {
        float dh[8];
        unpack8(p[r][j], dh);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xv = xh[r][j][e] * rstd[r];
          xh[r][j][e] = xv;
          g[j][e] = dh[e] * sc[j][e];
          ds[j][e] += dh[e] * xv;
          dot += g[j][e] * xv;
        }
      }
    }
    dot = warp_sum(dot) / D;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      const int ch = j * 32 + lane;
      if (ch < nchunk) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          o[e] = rstd[r] * (g[j][e] - xh[r][j][e] * dot) + rs[r][j][e];
          cs[j][e] += o[e];
        }
        if (dx_l) st8f(dx_l + static_cast<long long>(row + r) * D + ch * 8, o);
        if (mc_dx_row0) multimem_st_16B(mc_dx_row0 + r * rowb + ch * 16, pack8(o));
      }
    }
  }
}

// Block-level reduction of the backward partial sums: every warp's per-lane ds / cs -> red (smem, [WARPS][32 * 8]) ->
// warp `w0` of the group sums them in a fixed order into srow[0, D) (ds) and srow[D, 2D) (cs).  `sync()` must be a
// barrier over exactly the WARPS participating warps; gw = this warp's index within the group.
template <int NC, int WARPS, typename SyncFn>
__device__ __forceinline__ void sp_bwd_block_reduce(float (*red)[32 * 8], float* srow, int D, int gw, int lane,
                                                    const float (&ds)[NC][8], const float (&cs)[NC][8], SyncFn sync) {
  const int nchunk = D >> 3;
#pragma unroll
  for (int which = 0; which < 2; ++which) {
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      sync();
#pragma unroll
      for (int e = 0; e < 8; ++e) red[gw][lane * 8 + e] = which ? cs[j][e] : ds[j][e];
      sync();
      if (gw == 0) {
        const int ch = j * 32 + lane;
        if (ch < nchunk) {
          float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          for (int w = 0; w < WARPS; ++w)
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] += red[w][lane * 8 + e];
          st8f(srow + which * D + ch * 8, o);
        }
      }
    }
  }
}

}  // namespace dpfs
