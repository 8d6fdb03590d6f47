// LayerNorm forward and input-gradient backward over bf16 rows.  One warp per PAIR of rows: the
// rows are read once with 16-byte loads into registers (kept packed), reduced with warp shuffles
// (no shared memory, no block barrier), and written once.  Two rows per warp because one was
// latency-bound: 32 warps x 1.5 KB = 48 KB in flight per SM is ~3.3 TB/s by Little's law.  HBM-bound: fwd 4*d bytes/row, bwd 8*d (+2*d
// with the fused residual-gradient add).
#include "common.cuh"
#include "../../include/b200ft.h"

#ifndef LN_BWD_PREFETCH_RES
#define LN_BWD_PREFETCH_RES 1
#endif
#ifndef LN_BWD_BLOCKS
#define LN_BWD_BLOCKS 2
#endif
#ifndef LN_FWD_ROWS
#define LN_FWD_ROWS 2
#endif
#ifndef LN_FWD_BLOCKS
#define LN_FWD_BLOCKS 3
#endif

namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kRowsPerWarp = 2;
constexpr int kFwdRows = LN_FWD_ROWS;   // rows per warp in the forward kernel
constexpr int kMaxVecPerLane = 32;   // d <= 32 lanes * 32 vectors * 8 = 8192

// Rows are kept as the raw 16-byte bf16 vectors they were loaded as (4 registers per vector
// instead of 8) and unpacked on the fly in each pass; all loads of both rows are issued before
// the first reduction.
template <int VPL>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, VPL <= 4 ? LN_FWD_BLOCKS : 1)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean, float* __restrict__ rstd, int64_t rows, int d, float eps) {
  grid_dep_sync();
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5)) * kFwdRows;
  if (row0 >= rows) return;
  const int nvec = d >> 3;
  uint4 q[kFwdRows][VPL];
#pragma unroll
  for (int r = 0; r < kFwdRows; ++r) {
    const bool live = row0 + r < rows;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int vi = lane + i * 32;
      q[r][i] = (live && vi < nvec) ? ld_stream16(x + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
    }
  }
  float mu[kFwdRows], rs[kFwdRows];
#pragma unroll
  for (int r = 0; r < kFwdRows; ++r) {
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      float f[8];
      unpack8(q[r][i], f);  // vectors beyond the row are zero and add nothing
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += f[j];
    }
    mu[r] = warp_sum(sum) / (float)d;
  }
#pragma unroll
  for (int r = 0; r < kFwdRows; ++r) {
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      if (lane + i * 32 < nvec) {
        float f[8];
        unpack8(q[r][i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float c = f[j] - mu[r]; sq += c * c; }
      }
    }
    rs[r] = rsqrtf(warp_sum(sq) / (float)d + eps);
    if (lane == 0 && row0 + r < rows) { mean[row0 + r] = mu[r]; rstd[row0 + r] = rs[r]; }
  }
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + (vi << 3)));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int r = 0; r < kFwdRows; ++r) {
        if (row0 + r < rows) {
          float f[8], o[8];
          unpack8(q[r][i], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = (f[j] - mu[r]) * rs[r] * gg[j] + bb[j];
          st_stream16(y + (row0 + r) * d + (vi << 3), pack8(o));
        }
      }
    }
  }
}

// dx = dres + rstd * (g*dy - mean(g*dy) - xhat * mean(g*dy*xhat))
template <int VPL>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, VPL <= 4 ? LN_BWD_BLOCKS : 1)
layernorm_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                        const float* __restrict__ gamma, const float* __restrict__ mean,
                        const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres,
                        __nv_bfloat16* __restrict__ dx, int64_t rows, int d) {
  grid_dep_sync();
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5)) * kRowsPerWarp;
  if (row0 >= rows) return;
  const int nvec = d >> 3;
  uint4 qdy[kRowsPerWarp][VPL], qx[kRowsPerWarp][VPL];
#if LN_BWD_PREFETCH_RES
  uint4 qr[kRowsPerWarp][VPL];
#endif
  float mu[kRowsPerWarp], rs[kRowsPerWarp];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    const bool live = row0 + r < rows;
    mu[r] = live ? mean[row0 + r] : 0.f;
    rs[r] = live ? rstd[row0 + r] : 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int vi = lane + i * 32;
      const bool ok = live && vi < nvec;
      qdy[r][i] = ok ? ld_stream16(dy + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
      qx[r][i] = ok ? ld_stream16(x + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
#if LN_BWD_PREFETCH_RES
      qr[r][i] = (ok && dres) ? ld_stream16(dres + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
#endif
    }
  }
  float s1[kRowsPerWarp], s2[kRowsPerWarp];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) { s1[r] = 0.f; s2[r] = 0.f; }
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
      for (int r = 0; r < kRowsPerWarp; ++r) {
        float a[8], b[8];
        unpack8(qdy[r][i], a);
        unpack8(qx[r][i], b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gdy = a[j] * gg[j];
          s1[r] += gdy;
          s2[r] += gdy * ((b[j] - mu[r]) * rs[r]);
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    s1[r] = warp_sum(s1[r]) / (float)d;
    s2[r] = warp_sum(s2[r]) / (float)d;
  }
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
      for (int r = 0; r < kRowsPerWarp; ++r) {
        if (row0 + r < rows) {
          float a[8], b[8], o[8];
          unpack8(qdy[r][i], a);
          unpack8(qx[r][i], b);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = rs[r] * (a[j] * gg[j] - s1[r] - (b[j] - mu[r]) * rs[r] * s2[r]);
          if (dres) {
            float rr[8];
#if LN_BWD_PREFETCH_RES
            unpack8(qr[r][i], rr);
#else
            unpack8(ld_stream16(dres + (row0 + r) * d + (vi << 3)), rr);
#endif
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] += rr[j];
          }
          st_stream16(dx + (row0 + r) * d + (vi << 3), pack8(o));
        }
      }
    }
  }
}

inline int vpl_for(int d) { return ((d >> 3) + 31) / 32; }

}  // namespace

#define LN_DISPATCH(VPL_RT, CALL)                       \
  if ((VPL_RT) <= 1) { CALL(1); }                       \
  else if ((VPL_RT) <= 2) { CALL(2); }                  \
  else if ((VPL_RT) <= 3) { CALL(3); }                  \
  else if ((VPL_RT) <= 4) { CALL(4); }                  \
  else if ((VPL_RT) <= 8) { CALL(8); }                  \
  else if ((VPL_RT) <= 16) { CALL(16); }                \
  else { CALL(32); }

extern "C" int b200ft_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y,
                                    float* mean, float* rstd, int64_t rows, int d, float eps,
                                    void* stream) {
  B200FT_CHECK_ARG(x && gamma && beta && y && mean && rstd, "layernorm_fwd: null buffer");
  B200FT_CHECK_ARG(rows > 0 && d > 0 && d % 8 == 0 && d <= 8 * 32 * kMaxVecPerLane,
                   "layernorm_fwd: rows=%lld d=%d (d must be a multiple of 8, <= 8192)", (long long)rows, d);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows_per_block = (int64_t)kWarpsPerBlock * kFwdRows;
  unsigned blocks = (unsigned)((rows + rows_per_block - 1) / rows_per_block);
  int vpl = vpl_for(d);
#define CALL(V)                                                                                     \
  launch_k(layernorm_fwd_kernel<V>, blocks, kWarpsPerBlock * 32, 0, st,                              \
           (const __nv_bfloat16*)x, gamma, beta, (__nv_bfloat16*)y, mean, rstd, rows, d, eps)
  LN_DISPATCH(vpl, CALL)
#undef CALL
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_layernorm_bwd_dx(const void* dy, const void* x, const float* gamma,
                                       const float* mean, const float* rstd, const void* dres,
                                       void* dx, int64_t rows, int d, void* stream) {
  B200FT_CHECK_ARG(dy && x && gamma && mean && rstd && dx, "layernorm_bwd_dx: null buffer");
  B200FT_CHECK_ARG(rows > 0 && d > 0 && d % 8 == 0 && d <= 8 * 32 * kMaxVecPerLane,
                   "layernorm_bwd_dx: rows=%lld d=%d (d must be a multiple of 8, <= 8192)", (long long)rows, d);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows_per_block = (int64_t)kWarpsPerBlock * kRowsPerWarp;
  unsigned blocks = (unsigned)((rows + rows_per_block - 1) / rows_per_block);
  int vpl = vpl_for(d);
#define CALL(V)                                                                                     \
  launch_k(layernorm_bwd_dx_kernel<V>, blocks, kWarpsPerBlock * 32, 0, st,                           \
           (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, mean, rstd, (const __nv_bfloat16*)dres, \
           (__nv_bfloat16*)dx, rows, d)
  LN_DISPATCH(vpl, CALL)
#undef CALL
  B200FT_LAUNCH_CHECK();
  return 0;
}
