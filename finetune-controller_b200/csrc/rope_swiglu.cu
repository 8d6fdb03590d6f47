// Llama-family elementwise pieces of the step, all HBM-bound with 16-byte accesses:
//   rotary position embedding applied in place to the q and k heads of the fused qkv buffer
//   (forward, and its transpose for backward), and SwiGLU forward / backward.
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

// One thread per (row, head, 8-wide slice of the first half of the head): rotates the pairs
// (i, i + dh/2) — HF's rotate_half convention.  sign = +1 forward, -1 backward (inverse rotation).
__global__ void __launch_bounds__(256)
rope_kernel(__nv_bfloat16* __restrict__ x, int64_t ld, int64_t rows, int seq, int n_heads, int dh,
            const float* __restrict__ cos_t, const float* __restrict__ sin_t, float sign) {
  grid_dep_sync();
  // flat index in 32 bits (the host checks rows * heads * vectors < 2^31): 64-bit divisions cost ~100 instructions each
  const int half = dh >> 1;
  const int vph = half >> 3;  // 8-wide vectors per half head
  const unsigned per_row = (unsigned)(n_heads * vph);
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (unsigned)rows * per_row) return;
  const unsigned urow = idx / per_row, rem = idx - urow * per_row;
  const int64_t row = urow;
  const int head = (int)(rem / (unsigned)vph), vi = (int)rem - head * vph;
  const int pos = (int)(urow % (unsigned)seq);
  __nv_bfloat16* p1 = x + row * ld + head * dh + (vi << 3);
  __nv_bfloat16* p2 = p1 + half;
  float a[8], b[8];
  unpack8(*reinterpret_cast<const uint4*>(p1), a);
  unpack8(*reinterpret_cast<const uint4*>(p2), b);
  const float* cp = cos_t + (int64_t)pos * half + (vi << 3);
  const float* sp = sin_t + (int64_t)pos * half + (vi << 3);
  const float4 c0 = __ldg(reinterpret_cast<const float4*>(cp)), c1 = __ldg(reinterpret_cast<const float4*>(cp + 4));
  const float4 s0 = __ldg(reinterpret_cast<const float4*>(sp)), s1 = __ldg(reinterpret_cast<const float4*>(sp + 4));
  const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
  const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  float o1[8], o2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float sj = s[j] * sign;
    o1[j] = a[j] * c[j] - b[j] * sj;
    o2[j] = b[j] * c[j] + a[j] * sj;
  }
  *reinterpret_cast<uint4*>(p1) = pack8(o1);
  *reinterpret_cast<uint4*>(p2) = pack8(o2);
}

__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

// gu [rows, 2*ff]: gate in columns [0, ff), up in [ff, 2*ff).  act = silu(gate) * up.
__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act, int64_t rows, int ff) {
  grid_dep_sync();
  const int vpr = ff >> 3;
  const int64_t total = rows * vpr;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // (row, vector) advance with the grid stride: one division per thread, none per element
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step_rows = stride / vpr;
  const int step_v = (int)(stride - step_rows * vpr);
  int64_t row = i0 / vpr;
  int v = (int)(i0 - row * vpr);
  for (int64_t i = i0; i < total; i += stride, row += step_rows, v += step_v) {
    if (v >= vpr) { v -= vpr; ++row; }
    float g[8], u[8], o[8];
    unpack8(ld_stream16(gu + row * 2 * ff + (v << 3)), g);
    unpack8(ld_stream16(gu + row * 2 * ff + ff + (v << 3)), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = g[j] * sigmoid_f(g[j]) * u[j];
    st_stream16(act + row * ff + (v << 3), pack8(o));
  }
}

// d gate = dact * up * sig(g) * (1 + g * (1 - sig(g))),  d up = dact * silu(g)
__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* __restrict__ gu,
                  __nv_bfloat16* __restrict__ dgu, __nv_bfloat16* __restrict__ act, int64_t rows, int ff) {
  grid_dep_sync();
  const int vpr = ff >> 3;
  const int64_t total = rows * vpr;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // (row, vector) advance with the grid stride: one division per thread, none per element
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step_rows = stride / vpr;
  const int step_v = (int)(stride - step_rows * vpr);
  int64_t row = i0 / vpr;
  int v = (int)(i0 - row * vpr);
  for (int64_t i = i0; i < total; i += stride, row += step_rows, v += step_v) {
    if (v >= vpr) { v -= vpr; ++row; }
    float g[8], u[8], d[8], dg[8], du[8], a[8];
    unpack8(ld_stream16(gu + row * 2 * ff + (v << 3)), g);
    unpack8(ld_stream16(gu + row * 2 * ff + ff + (v << 3)), u);
    unpack8(ld_stream16(dact + row * ff + (v << 3)), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = sigmoid_f(g[j]);
      dg[j] = d[j] * u[j] * sg * (1.f + g[j] * (1.f - sg));
      du[j] = d[j] * g[j] * sg;
      a[j] = g[j] * sg * u[j];
    }
    st_stream16(dgu + row * 2 * ff + (v << 3), pack8(dg));
    st_stream16(dgu + row * 2 * ff + ff + (v << 3), pack8(du));
    if (act) st_stream16(act + row * ff + (v << 3), pack8(a));  // the forward product, recomputed for dW_down
  }
}

}  // namespace

extern "C" int b200ft_rope_inplace(void* x, int64_t ld, int64_t rows, int seq, int n_heads, int d_head,
                                   const float* cos_table, const float* sin_table, int backward, void* stream) {
  B200FT_CHECK_ARG(x && cos_table && sin_table, "rope: null buffer");
  B200FT_CHECK_ARG(rows > 0 && seq > 0 && n_heads > 0 && d_head > 0 && d_head % 16 == 0 && ld % 8 == 0 &&
                       ld >= (int64_t)n_heads * d_head && (uintptr_t)x % 16 == 0,
                   "rope: rows=%lld seq=%d heads=%d d_head=%d ld=%lld (d_head %% 16 == 0, ld %% 8 == 0)",
                   (long long)rows, seq, n_heads, d_head, (long long)ld);
  const int64_t total = rows * n_heads * (d_head / 16);
  B200FT_CHECK_ARG(total < (1ll << 31), "rope: rows * heads * d_head / 16 = %lld does not fit 31 bits", (long long)total);
  launch_k(rope_kernel, (unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream,
      (__nv_bfloat16*)x, ld, rows, seq, n_heads, d_head, cos_table, sin_table, backward ? -1.f : 1.f);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_swiglu_fwd(const void* gate_up, void* act, int64_t rows, int d_ff, void* stream) {
  B200FT_CHECK_ARG(gate_up && act && rows > 0 && d_ff > 0 && d_ff % 8 == 0, "swiglu_fwd: rows=%lld d_ff=%d",
                   (long long)rows, d_ff);
  const int64_t total = rows * (d_ff / 8);
  int64_t want = (total + 255) / 256;
  int blocks = (int)(want > B200FT_NUM_SMS * 16 ? B200FT_NUM_SMS * 16 : want);
  launch_k(swiglu_fwd_kernel, blocks, 256, 0, (cudaStream_t)stream, (const __nv_bfloat16*)gate_up, (__nv_bfloat16*)act, rows, d_ff);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_swiglu_bwd(const void* dact, const void* gate_up, void* dgate_up, int64_t rows, int d_ff,
                                 void* stream) {
  return b200ft_swiglu_bwd_act(dact, gate_up, dgate_up, nullptr, rows, d_ff, stream);
}

extern "C" int b200ft_swiglu_bwd_act(const void* dact, const void* gate_up, void* dgate_up, void* act, int64_t rows,
                                     int d_ff, void* stream) {
  B200FT_CHECK_ARG(dact && gate_up && dgate_up && rows > 0 && d_ff > 0 && d_ff % 8 == 0,
                   "swiglu_bwd: rows=%lld d_ff=%d", (long long)rows, d_ff);
  const int64_t total = rows * (d_ff / 8);
  int64_t want = (total + 255) / 256;
  int blocks = (int)(want > B200FT_NUM_SMS * 16 ? B200FT_NUM_SMS * 16 : want);
  launch_k(swiglu_bwd_kernel, blocks, 256, 0, (cudaStream_t)stream, (const __nv_bfloat16*)dact, (const __nv_bfloat16*)gate_up,
                                                               (__nv_bfloat16*)dgate_up, (__nv_bfloat16*)act, rows, d_ff);
  B200FT_LAUNCH_CHECK();
  return 0;
}
