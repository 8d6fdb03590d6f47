// LayerNorm forward and input-gradient backward over bf16 rows.  One warp per row: the row
// is read once with 16-byte loads into registers, reduced with warp shuffles (no shared
// memory, no block barrier), and written once.  HBM-bound: fwd 4*d bytes/row, bwd 8*d (+2*d
// with the fused residual-gradient add).
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kMaxVecPerLane = 32;  // d <= 32 lanes * 32 vectors * 8 = 8192

template <int VPL>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean, float* __restrict__ rstd, int64_t rows, int d, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = d >> 3;
  const __nv_bfloat16* xr = x + row * d;
  float v[VPL][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
      unpack8(ld_stream16(xr + (vi << 3)), v[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[i][j];
    }
  }
  sum = warp_sum(sum);
  const float mu = sum / (float)d;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { float c = v[i][j] - mu; sq += c * c; }
    }
  }
  sq = warp_sum(sq);
  const float rs = rsqrtf(sq / (float)d + eps);
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
  __nv_bfloat16* yr = y + row * d;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + (vi << 3)));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mu) * rs * gg[j] + bb[j];
      st_stream16(yr + (vi << 3), pack8(o));
    }
  }
}

// dx = dres + rstd * (g*dy - mean(g*dy) - xhat * mean(g*dy*xhat))
template <int VPL>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
layernorm_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                        const float* __restrict__ gamma, const float* __restrict__ mean,
                        const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres,
                        __nv_bfloat16* __restrict__ dx, int64_t rows, int d) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = d >> 3;
  const float mu = mean[row], rs = rstd[row];
  float gdy[VPL][8], xh[VPL][8];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
      float a[8], b[8];
      unpack8(ld_stream16(dy + row * d + (vi << 3)), a);
      unpack8(ld_stream16(x + row * d + (vi << 3)), b);
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3)));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + (vi << 3) + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        gdy[i][j] = a[j] * gg[j];
        xh[i][j] = (b[j] - mu) * rs;
        s1 += gdy[i][j];
        s2 += gdy[i][j] * xh[i][j];
      }
    }
  }
  s1 = warp_sum(s1) / (float)d;
  s2 = warp_sum(s2) / (float)d;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    int vi = lane + i * 32;
    if (vi < nvec) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rs * (gdy[i][j] - s1 - xh[i][j] * s2);
      if (dres) {
        float r[8];
        unpack8(ld_stream16(dres + row * d + (vi << 3)), r);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += r[j];
      }
      st_stream16(dx + row * d + (vi << 3), pack8(o));
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
  unsigned blocks = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
  int vpl = vpl_for(d);
#define CALL(V)                                                                                     \
  layernorm_fwd_kernel<V><<<blocks, kWarpsPerBlock * 32, 0, st>>>(                                   \
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
  unsigned blocks = (unsigned)((rows + kWarpsPerBlock - 1) / kWarpsPerBlock);
  int vpl = vpl_for(d);
#define CALL(V)                                                                                     \
  layernorm_bwd_dx_kernel<V><<<blocks, kWarpsPerBlock * 32, 0, st>>>(                                \
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, mean, rstd, (const __nv_bfloat16*)dres, \
      (__nv_bfloat16*)dx, rows, d)
  LN_DISPATCH(vpl, CALL)
#undef CALL
  B200FT_LAUNCH_CHECK();
  return 0;
}
