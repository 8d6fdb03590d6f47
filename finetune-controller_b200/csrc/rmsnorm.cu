// RMSNorm forward and input-gradient backward over bf16 rows (Llama family):
//   y = x * rstd * g,  rstd = rsqrt(mean(x^2) + eps)
//   dx = dres + rstd * (g*dy - xhat * mean(g*dy*xhat)),  xhat = x * rstd
// Same shape as layernorm.cu: one warp per pair of rows (one row once a row no longer fits the
// register budget twice: d >= 2048), rows kept as packed 16-byte vectors in registers,
// warp-shuffle reductions, no shared memory.  HBM-bound: fwd 4*d bytes/row, bwd 6*d
// (+2*d with the fused residual-gradient add).
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kWarpsPerBlock = 8;

template <int VPL, int kRowsPerWarp>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                   __nv_bfloat16* __restrict__ y, float* __restrict__ rstd, int64_t rows, int d, float eps) {
  grid_dep_sync();
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5)) * kRowsPerWarp;
  if (row0 >= rows) return;
  const int nvec = d >> 3;
  uint4 q[kRowsPerWarp][VPL];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    const bool live = row0 + r < rows;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int vi = lane + i * 32;
      q[r][i] = (live && vi < nvec) ? ld_stream16(x + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
    }
  }
  float rs[kRowsPerWarp];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      float f[8];
      unpack8(q[r][i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) sq += f[j] * f[j];
    }
    rs[r] = rsqrtf(warp_sum(sq) / (float)d + eps);
    if (lane == 0 && row0 + r < rows) rstd[row0 + r] = rs[r];
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
          float f[8], o[8];
          unpack8(q[r][i], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = f[j] * rs[r] * gg[j];
          st_stream16(y + (row0 + r) * d + (vi << 3), pack8(o));
        }
      }
    }
  }
}

template <int VPL, int kRowsPerWarp>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
rmsnorm_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                      const float* __restrict__ gamma, const float* __restrict__ rstd,
                      const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx, int64_t rows, int d) {
  grid_dep_sync();
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5)) * kRowsPerWarp;
  if (row0 >= rows) return;
  const int nvec = d >> 3;
  uint4 qdy[kRowsPerWarp][VPL], qx[kRowsPerWarp][VPL];
  float rs[kRowsPerWarp], s2[kRowsPerWarp];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    const bool live = row0 + r < rows;
    rs[r] = live ? rstd[row0 + r] : 0.f;
    s2[r] = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int vi = lane + i * 32;
      const bool ok = live && vi < nvec;
      qdy[r][i] = ok ? ld_stream16(dy + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
      qx[r][i] = ok ? ld_stream16(x + (row0 + r) * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
    }
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
        float a[8], b[8];
        unpack8(qdy[r][i], a);
        unpack8(qx[r][i], b);
#pragma unroll
        for (int j = 0; j < 8; ++j) s2[r] += a[j] * gg[j] * (b[j] * rs[r]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) s2[r] = warp_sum(s2[r]) / (float)d;
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
          for (int j = 0; j < 8; ++j) o[j] = rs[r] * (a[j] * gg[j] - b[j] * rs[r] * s2[r]);
          if (dres) {
            float rr[8];
            unpack8(ld_stream16(dres + (row0 + r) * d + (vi << 3)), rr);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] += rr[j];
          }
          st_stream16(dx + (row0 + r) * d + (vi << 3), pack8(o));
        }
      }
    }
  }
}

// ---- wide rows (d > 1024): four warps per row, so that a row's registers are spread over 128
// threads (d = 4096: 4 vectors per lane per array) and 16 CTAs = 16 rows per SM stay in flight.
constexpr int kWideThreads = 128;

__device__ __forceinline__ float block4_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  const float t = red[0] + red[1] + red[2] + red[3];
  __syncthreads();  // red is reused by the caller's next reduction
  return t;
}

template <int VPL>
__global__ void __launch_bounds__(kWideThreads)
rmsnorm_fwd_wide_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                        __nv_bfloat16* __restrict__ y, float* __restrict__ rstd, int d, float eps) {
  grid_dep_sync();
  __shared__ float red[4];
  const int64_t row = blockIdx.x;
  const int nvec = d >> 3;
  uint4 q[VPL];
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = threadIdx.x + i * kWideThreads;
    q[i] = vi < nvec ? ld_stream16(x + row * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
  }
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    float f[8];
    unpack8(q[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) sq += f[j] * f[j];
  }
  const float rs = rsqrtf(block4_sum(sq, red) / (float)d + eps);
  if (threadIdx.x == 0) rstd[row] = rs;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = threadIdx.x + i * kWideThreads;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      float f[8], o[8];
      unpack8(q[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = f[j] * rs * gg[j];
      st_stream16(y + row * d + (vi << 3), pack8(o));
    }
  }
}

template <int VPL>
__global__ void __launch_bounds__(kWideThreads)
rmsnorm_bwd_dx_wide_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                           const float* __restrict__ gamma, const float* __restrict__ rstd,
                           const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx, int d) {
  grid_dep_sync();
  __shared__ float red[4];
  const int64_t row = blockIdx.x;
  const int nvec = d >> 3;
  const float rs = rstd[row];
  uint4 qdy[VPL], qx[VPL], qr[VPL];
#pragma unroll
  for (int i = 0; i < VPL; ++i) {  // every load of the row is issued before the reduction
    const int vi = threadIdx.x + i * kWideThreads;
    const bool ok = vi < nvec;
    qdy[i] = ok ? ld_stream16(dy + row * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
    qx[i] = ok ? ld_stream16(x + row * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
    qr[i] = (ok && dres) ? ld_stream16(dres + row * d + (vi << 3)) : make_uint4(0u, 0u, 0u, 0u);
  }
  float s2 = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = threadIdx.x + i * kWideThreads;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      float a[8], b[8];
      unpack8(qdy[i], a);
      unpack8(qx[i], b);
#pragma unroll
      for (int j = 0; j < 8; ++j) s2 += a[j] * gg[j] * (b[j] * rs);
    }
  }
  s2 = block4_sum(s2, red) / (float)d;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = threadIdx.x + i * kWideThreads;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      float a[8], b[8], rr[8], o[8];
      unpack8(qdy[i], a);
      unpack8(qx[i], b);
      unpack8(qr[i], rr);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rs * (a[j] * gg[j] - b[j] * rs * s2) + rr[j];
      st_stream16(dx + row * d + (vi << 3), pack8(o));
    }
  }
}

inline int vpl_for(int d) { return ((d >> 3) + 31) / 32; }

}  // namespace

// (vectors per lane, rows per warp)
#define RMS_DISPATCH(VPL_RT, CALL)                      \
  if ((VPL_RT) <= 1) { CALL(1, 2); }                    \
  else if ((VPL_RT) <= 2) { CALL(2, 2); }               \
  else { CALL(4, 2); }

extern "C" int b200ft_rmsnorm_fwd(const void* x, const float* gamma, void* y, float* rstd, int64_t rows, int d,
                                  float eps, void* stream) {
  B200FT_CHECK_ARG(x && gamma && y && rstd, "rmsnorm_fwd: null buffer");
  B200FT_CHECK_ARG(rows > 0 && d > 0 && d % 8 == 0 && d <= 4096,
                   "rmsnorm_fwd: rows=%lld d=%d (d must be a multiple of 8, <= 4096)", (long long)rows, d);
  cudaStream_t st = (cudaStream_t)stream;
  if (d > 1024) {  // wide rows: one 128-thread CTA per row
    const int wv = ((d >> 3) + kWideThreads - 1) / kWideThreads;
    const __nv_bfloat16* xi = (const __nv_bfloat16*)x;
    __nv_bfloat16* yo = (__nv_bfloat16*)y;
    if (wv <= 2) launch_k(rmsnorm_fwd_wide_kernel<2>, (unsigned)rows, kWideThreads, 0, st, xi, gamma, yo, rstd, d, eps);
    else launch_k(rmsnorm_fwd_wide_kernel<4>, (unsigned)rows, kWideThreads, 0, st, xi, gamma, yo, rstd, d, eps);
    B200FT_LAUNCH_CHECK();
    return 0;
  }
  int vpl = vpl_for(d);
#define CALL(V, R)                                                                                  \
  launch_k(rmsnorm_fwd_kernel<V, R>, (unsigned)((rows + kWarpsPerBlock * R - 1) / (kWarpsPerBlock * R)), \
                             kWarpsPerBlock * 32, 0, st, (const __nv_bfloat16*)x, gamma,           \
                                                           (__nv_bfloat16*)y, rstd, rows, d, eps)
  RMS_DISPATCH(vpl, CALL)
#undef CALL
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_rmsnorm_bwd_dx(const void* dy, const void* x, const float* gamma, const float* rstd,
                                     const void* dres, void* dx, int64_t rows, int d, void* stream) {
  B200FT_CHECK_ARG(dy && x && gamma && rstd && dx, "rmsnorm_bwd_dx: null buffer");
  B200FT_CHECK_ARG(rows > 0 && d > 0 && d % 8 == 0 && d <= 4096,
                   "rmsnorm_bwd_dx: rows=%lld d=%d (d must be a multiple of 8, <= 4096)", (long long)rows, d);
  cudaStream_t st = (cudaStream_t)stream;
  if (d > 1024) {
    const int wv = ((d >> 3) + kWideThreads - 1) / kWideThreads;
    const __nv_bfloat16 *dyi = (const __nv_bfloat16*)dy, *xi = (const __nv_bfloat16*)x, *ri = (const __nv_bfloat16*)dres;
    __nv_bfloat16* o = (__nv_bfloat16*)dx;
    if (wv <= 2) launch_k(rmsnorm_bwd_dx_wide_kernel<2>, (unsigned)rows, kWideThreads, 0, st, dyi, xi, gamma, rstd, ri, o, d);
    else launch_k(rmsnorm_bwd_dx_wide_kernel<4>, (unsigned)rows, kWideThreads, 0, st, dyi, xi, gamma, rstd, ri, o, d);
    B200FT_LAUNCH_CHECK();
    return 0;
  }
  int vpl = vpl_for(d);
#define CALL(V, R)                                                                                     \
  launch_k(rmsnorm_bwd_dx_kernel<V, R>, (unsigned)((rows + kWarpsPerBlock * R - 1) / (kWarpsPerBlock * R)), \
                                kWarpsPerBlock * 32, 0, st,                                           \
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, rstd, (const __nv_bfloat16*)dres,       \
      (__nv_bfloat16*)dx, rows, d)
  RMS_DISPATCH(vpl, CALL)
#undef CALL
  B200FT_LAUNCH_CHECK();
  return 0;
}
