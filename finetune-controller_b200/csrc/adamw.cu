// Fused AdamW + gradient scale, and the global grad-norm reduction that feeds its clip factor.
// HBM-bound: 28 B/element (read g2 p4 m4 v4, write p4 m4 v4 shadow2).  Every access is a
// 16-byte vector; the grid is a multiple of the SM count and strides over the flat buffer.
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

struct AdamScalars {
  float lr, beta1, beta2, eps, wd_mul, step_size, inv_bc2_sqrt, grad_scale;
};

// One element, in the operation order of torch.optim.AdamW's single-tensor path
// (mul_ by 1-lr*wd; lerp_; mul_/addcmul_; sqrt/div/add eps; addcdiv_).
__device__ __forceinline__ void adam1(float& p, float& m, float& v, float g, const AdamScalars& s) {
  p = p * s.wd_mul;
  m = m + (g - m) * (1.f - s.beta1);
  v = v * s.beta2 + (1.f - s.beta2) * g * g;
  float denom = sqrtf(v) * s.inv_bc2_sqrt + s.eps;
  p = p - s.step_size * (m / denom);
}

template <bool G_BF16>
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
             const void* __restrict__ g, __nv_bfloat16* __restrict__ shadow, int64_t n,
             AdamScalars s, const float* __restrict__ grad_scale_dev) {
  grid_dep_sync();
  if (grad_scale_dev) s.grad_scale *= __ldg(grad_scale_dev);
  const int64_t nvec = n >> 3;  // 8 elements per thread per iteration
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const int64_t e = i << 3;
    float4 p0 = *reinterpret_cast<const float4*>(p + e), p1 = *reinterpret_cast<const float4*>(p + e + 4);
    float4 m0 = *reinterpret_cast<const float4*>(m + e), m1 = *reinterpret_cast<const float4*>(m + e + 4);
    float4 v0 = *reinterpret_cast<const float4*>(v + e), v1 = *reinterpret_cast<const float4*>(v + e + 4);
    float gg[8];
    if (G_BF16) {
      uint4 q = ld_stream16(reinterpret_cast<const __nv_bfloat16*>(g) + e);
      unpack8(q, gg);
    } else {
      const float* gf = reinterpret_cast<const float*>(g) + e;
      float4 g0 = *reinterpret_cast<const float4*>(gf), g1 = *reinterpret_cast<const float4*>(gf + 4);
      gg[0] = g0.x; gg[1] = g0.y; gg[2] = g0.z; gg[3] = g0.w;
      gg[4] = g1.x; gg[5] = g1.y; gg[6] = g1.z; gg[7] = g1.w;
    }
    float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) adam1(pp[j], mm[j], vv[j], gg[j] * s.grad_scale, s);
    *reinterpret_cast<float4*>(p + e) = make_float4(pp[0], pp[1], pp[2], pp[3]);
    *reinterpret_cast<float4*>(p + e + 4) = make_float4(pp[4], pp[5], pp[6], pp[7]);
    *reinterpret_cast<float4*>(m + e) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    *reinterpret_cast<float4*>(m + e + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
    *reinterpret_cast<float4*>(v + e) = make_float4(vv[0], vv[1], vv[2], vv[3]);
    *reinterpret_cast<float4*>(v + e + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
    if (shadow) *reinterpret_cast<uint4*>(shadow + e) = pack8(pp);
  }
  // ragged tail (< 8 elements): one thread
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (int64_t e = nvec << 3; e < n; ++e) {
      float gv = G_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g)[e])
                        : reinterpret_cast<const float*>(g)[e];
      float pe = p[e], me = m[e], ve = v[e];
      adam1(pe, me, ve, gv * s.grad_scale, s);
      p[e] = pe; m[e] = me; v[e] = ve;
      if (shadow) shadow[e] = __float2bfloat16_rn(pe);
    }
  }
}

// ---- sum of squares: per-block partials into the workspace, last block folds them ----
template <bool G_BF16>
__global__ void __launch_bounds__(256)
sumsq_kernel(const void* __restrict__ g, int64_t n, float* __restrict__ partials,
             unsigned int* __restrict__ ticket, float* __restrict__ out) {
  grid_dep_sync();
  float acc = 0.f;
  const int64_t nvec = n >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float f[8];
    if (G_BF16) {
      unpack8(ld_stream16(reinterpret_cast<const __nv_bfloat16*>(g) + (i << 3)), f);
    } else {
      const float* gf = reinterpret_cast<const float*>(g) + (i << 3);
      float4 a = *reinterpret_cast<const float4*>(gf), b = *reinterpret_cast<const float4*>(gf + 4);
      f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int64_t e = nvec << 3; e < n; ++e) {
      float gv = G_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g)[e])
                        : reinterpret_cast<const float*>(g)[e];
      acc += gv * gv;
    }
  __shared__ float red[8];
  __shared__ bool last;
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    partials[blockIdx.x] = t;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {  // deterministic fold of the per-block partials, fixed order
    float t = 0.f;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) t += __ldcg(partials + i);
    t = warp_sum(t);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int w = 0; w < 8; ++w) s += red[w];
      out[0] = s;
      *ticket = 0;  // leave the workspace ready for the next call
    }
  }
}

__global__ void clip_coef_kernel(const float* norm_sq, float max_norm, float pre_scale, float* coef) {
  grid_dep_sync();
  float nrm = sqrtf(norm_sq[0]) * pre_scale;
  coef[0] = fminf(1.f, max_norm / (nrm + 1e-6f));
}

constexpr int kSumsqBlocks = B200FT_NUM_SMS * 4;

}  // namespace

extern "C" int b200ft_adamw_step(float* p, float* m, float* v, const void* g, int g_is_bf16,
                                 void* shadow_bf16, int64_t n, float lr, float beta1, float beta2,
                                 float eps, float weight_decay, int step, float grad_scale,
                                 const float* grad_scale_dev, void* stream) {
  return b200ft_adamw_step_limited(p, m, v, g, g_is_bf16, shadow_bf16, n, lr, beta1, beta2, eps, weight_decay, step,
                                   grad_scale, grad_scale_dev, 0, stream);
}

extern "C" int b200ft_adamw_step_limited(float* p, float* m, float* v, const void* g, int g_is_bf16,
                                         void* shadow_bf16, int64_t n, float lr, float beta1, float beta2,
                                         float eps, float weight_decay, int step, float grad_scale,
                                         const float* grad_scale_dev, int max_blocks, void* stream) {
  B200FT_CHECK_ARG(max_blocks >= 0, "adamw_step: max_blocks=%d", max_blocks);
  B200FT_CHECK_ARG(p && m && v && g, "adamw_step: null buffer");
  B200FT_CHECK_ARG(n >= 0 && step >= 1, "adamw_step: n=%lld step=%d", (long long)n, step);
  B200FT_CHECK_ARG(((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)g | (uintptr_t)shadow_bf16) % 16 == 0,
                   "adamw_step: buffers must be 16-byte aligned");
  if (n == 0) return 0;
  AdamScalars s;
  s.lr = lr; s.beta1 = beta1; s.beta2 = beta2; s.eps = eps;
  s.wd_mul = 1.f - lr * weight_decay;
  // bias corrections in double on the host, as torch computes them with python floats
  double bc1 = 1.0 - pow((double)beta1, (double)step);
  double bc2 = 1.0 - pow((double)beta2, (double)step);
  s.step_size = (float)((double)lr / bc1);
  s.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
  s.grad_scale = grad_scale;
  int64_t nvec = n >> 3;
  int64_t want = (nvec + 255) / 256;
  int blocks = (int)(want < 1 ? 1 : (want > B200FT_NUM_SMS * 8 ? B200FT_NUM_SMS * 8 : want));
  if (max_blocks > 0 && blocks > max_blocks) blocks = max_blocks;
  cudaStream_t st = (cudaStream_t)stream;
  if (g_is_bf16)
    launch_k(adamw_kernel<true>, blocks, 256, 0, st, p, m, v, g, (__nv_bfloat16*)shadow_bf16, n, s, grad_scale_dev);
  else
    launch_k(adamw_kernel<false>, blocks, 256, 0, st, p, m, v, g, (__nv_bfloat16*)shadow_bf16, n, s, grad_scale_dev);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" size_t b200ft_gradnorm_sq_workspace_bytes(int64_t n) {
  (void)n;
  return (size_t)(kSumsqBlocks + 64) * sizeof(float);  // partials + ticket (zeroed by caller once)
}

extern "C" int b200ft_gradnorm_sq(const void* g, int g_is_bf16, int64_t n, float* out_sq,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  B200FT_CHECK_ARG(g && out_sq && workspace, "gradnorm_sq: null buffer");
  if (workspace_bytes < b200ft_gradnorm_sq_workspace_bytes(n)) {
    b200ft_set_error("gradnorm_sq: workspace %zu < %zu", workspace_bytes, b200ft_gradnorm_sq_workspace_bytes(n));
    return B200FT_ERR_WORKSPACE;
  }
  float* partials = (float*)workspace;
  unsigned int* ticket = (unsigned int*)(partials + kSumsqBlocks);
  cudaStream_t st = (cudaStream_t)stream;
  B200FT_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st));
  int64_t want = ((n >> 3) + 255) / 256;
  int blocks = (int)(want < 1 ? 1 : (want > kSumsqBlocks ? kSumsqBlocks : want));
  if (g_is_bf16) launch_k(sumsq_kernel<true>, blocks, 256, 0, st, g, n, partials, ticket, out_sq);
  else launch_k(sumsq_kernel<false>, blocks, 256, 0, st, g, n, partials, ticket, out_sq);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_clip_coef(const float* norm_sq, float max_norm, float pre_scale, float* coef, void* stream) {
  B200FT_CHECK_ARG(norm_sq && coef && max_norm > 0.f, "clip_coef: bad arguments");
  launch_k(clip_coef_kernel, 1, 1, 0, (cudaStream_t)stream, norm_sq, max_norm, pre_scale, coef);
  B200FT_LAUNCH_CHECK();
  return 0;
}
