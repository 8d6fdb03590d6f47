// Adapter-gradient contraction over the token dimension:
//   out[i, c] += alpha * sum_m S[m, i] * T[m, c]        i < R (<= 16), c < cols
// used for dB = (s*X*A)^T * dY  (S = U, T = dY) and dA^T = dU^T * X  (S = dU, T = X).
// Both operands have the reduction index as their slow dimension, so the wide operand T is
// streamed once from HBM in row order (coalesced 4-byte lanes -> 128 B per warp request) and
// the skinny operand S is staged in shared memory and broadcast.  HBM-bound on T:
// algorithmic bytes = 2*M*cols (+ 2*M*R).
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kRowsPerBlock = 128;
constexpr int kThreads = 256;

template <int R>
__global__ void __launch_bounds__(kThreads)
lora_grad_kernel(const __nv_bfloat16* __restrict__ S, int64_t ld_s, const __nv_bfloat16* __restrict__ T,
                 int64_t ld_t, int cols, int64_t M, float* __restrict__ out, int64_t ld_out, float alpha) {
  __shared__ __align__(16) float sS[kRowsPerBlock][R];
  const int64_t m0 = (int64_t)blockIdx.y * kRowsPerBlock;
  const int nrows = (int)((M - m0) < kRowsPerBlock ? (M - m0) : kRowsPerBlock);
  for (int idx = threadIdx.x; idx < kRowsPerBlock * R; idx += kThreads) {
    int rr = idx / R, ii = idx % R;
    sS[rr][ii] = (rr < nrows) ? __bfloat162float(S[(m0 + rr) * ld_s + ii]) : 0.f;
  }
  __syncthreads();
  const int c = (blockIdx.x * kThreads + threadIdx.x) * 2;
  if (c >= cols) return;  // cols is even (checked on the host)
  float acc0[R], acc1[R];
#pragma unroll
  for (int i = 0; i < R; ++i) { acc0[i] = 0.f; acc1[i] = 0.f; }
  const __nv_bfloat16* tp = T + m0 * ld_t + c;
#pragma unroll 4
  for (int rr = 0; rr < nrows; ++rr) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(tp + (int64_t)rr * ld_t));
    const float t0 = bf16lo(w), t1 = bf16hi(w);
#pragma unroll
    for (int i = 0; i < R; i += 4) {
      const float4 s4 = *reinterpret_cast<const float4*>(&sS[rr][i]);
      acc0[i] += s4.x * t0; acc1[i] += s4.x * t1;
      acc0[i + 1] += s4.y * t0; acc1[i + 1] += s4.y * t1;
      acc0[i + 2] += s4.z * t0; acc1[i + 2] += s4.z * t1;
      acc0[i + 3] += s4.w * t0; acc1[i + 3] += s4.w * t1;
    }
  }
#pragma unroll
  for (int i = 0; i < R; ++i) {
    atomicAdd(out + (int64_t)i * ld_out + c, alpha * acc0[i]);
    atomicAdd(out + (int64_t)i * ld_out + c + 1, alpha * acc1[i]);
  }
}

}  // namespace

extern "C" int b200ft_lora_grad(const void* s, int64_t ld_s, int r, const void* t, int64_t ld_t,
                                int cols, int64_t m, float* out, int64_t ld_out, float alpha,
                                void* stream) {
  B200FT_CHECK_ARG(s && t && out, "lora_grad: null buffer");
  B200FT_CHECK_ARG(r > 0 && r <= 16 && r % 4 == 0 && ld_s >= r, "lora_grad: r=%d ld_s=%lld (r in {4,8,12,16})", r,
                   (long long)ld_s);
  B200FT_CHECK_ARG(cols > 0 && cols % 2 == 0 && ld_t >= cols && ld_t % 2 == 0 && ld_out >= cols && m > 0,
                   "lora_grad: cols=%d ld_t=%lld ld_out=%lld m=%lld", cols, (long long)ld_t,
                   (long long)ld_out, (long long)m);
  B200FT_CHECK_ARG((uintptr_t)t % 4 == 0, "lora_grad: T must be 4-byte aligned");
  dim3 grid((cols / 2 + kThreads - 1) / kThreads, (unsigned)((m + kRowsPerBlock - 1) / kRowsPerBlock));
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16* S = (const __nv_bfloat16*)s;
  const __nv_bfloat16* T = (const __nv_bfloat16*)t;
  if (r <= 8 && r % 8 == 0)
    lora_grad_kernel<8><<<grid, kThreads, 0, st>>>(S, ld_s, T, ld_t, cols, m, out, ld_out, alpha);
  else if (r == 4)
    lora_grad_kernel<4><<<grid, kThreads, 0, st>>>(S, ld_s, T, ld_t, cols, m, out, ld_out, alpha);
  else if (r == 12)
    lora_grad_kernel<12><<<grid, kThreads, 0, st>>>(S, ld_s, T, ld_t, cols, m, out, ld_out, alpha);
  else
    lora_grad_kernel<16><<<grid, kThreads, 0, st>>>(S, ld_s, T, ld_t, cols, m, out, ld_out, alpha);
  B200FT_LAUNCH_CHECK();
  return 0;
}
