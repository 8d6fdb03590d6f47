// Shared device/host helpers for libb200ft (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#ifndef B200FT_NUM_SMS
#define B200FT_NUM_SMS 148  // B200: 2 dies x 74 SMs; grids below are sized in multiples of this
#endif

// ----------------------------------------------------------------------------------
// error plumbing: nothing throws across the C boundary; the last message is kept
// per host thread and handed out by b200ft_last_error_string().
// ----------------------------------------------------------------------------------
void b200ft_set_error(const char* fmt, ...);

#define B200FT_ERR_ARG (-1)
#define B200FT_ERR_UNSUPPORTED (-2)
#define B200FT_ERR_WORKSPACE (-3)

#define B200FT_CHECK_ARG(cond, ...)              \
  do {                                           \
    if (!(cond)) {                               \
      b200ft_set_error(__VA_ARGS__);             \
      return B200FT_ERR_ARG;                     \
    }                                            \
  } while (0)

#define B200FT_CUDA(call)                                                            \
  do {                                                                               \
    cudaError_t e__ = (call);                                                        \
    if (e__ != cudaSuccess) {                                                        \
      b200ft_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),      \
                       __FILE__, __LINE__);                                          \
      return (int)e__;                                                               \
    }                                                                                \
  } while (0)

#define B200FT_LAUNCH_CHECK()                                                        \
  do {                                                                               \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess) {                                                        \
      b200ft_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__),  \
                       __FILE__, __LINE__);                                          \
      return (int)e__;                                                               \
    }                                                                                \
  } while (0)

// ----------------------------------------------------------------------------------
// kernel launch: one path for every kernel of the library
// ----------------------------------------------------------------------------------
#ifdef __CUDACC__
bool b200ft_pdl_enabled();  // abi.cu: B200FT_PDL=0 in the environment turns programmatic dependent launch off

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                            Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = b200ft_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, args...);
}
#endif

// ----------------------------------------------------------------------------------
// device helpers
// ----------------------------------------------------------------------------------
#ifdef __CUDACC__

// Programmatic dependent launch.  Every kernel of the library is launched with the programmatic-stream-serialization
// attribute (launch_k below): its CTAs may be scheduled while the previous kernel of the stream is still draining, so
// that block dispatch and the prologue (barrier init, TMEM allocation, descriptor prefetch) run under the predecessor's
// tail instead of after it.  grid_dep_sync() is the point up to which that is safe: it returns once the
// predecessor grid has completed and its writes are visible, and then lets the NEXT kernel's CTAs be scheduled.
// Nothing before it may read or write global memory.  Without the launch attribute both instructions are no-ops.
__device__ __forceinline__ void grid_dep_sync() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// 16-byte streaming load / store (data touched once: keep it out of L1)
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
// plain (coherent) 16-byte load: for data another kernel phase of the SAME launch wrote
__device__ __forceinline__ uint4 ld_cg16(const void* p) {
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);  // .x = lo (low 16 bits)
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void unpack8(const uint4& q, float* f) {
  f[0] = bf16lo(q.x); f[1] = bf16hi(q.x); f[2] = bf16lo(q.y); f[3] = bf16hi(q.y);
  f[4] = bf16lo(q.z); f[5] = bf16hi(q.z); f[6] = bf16lo(q.w); f[7] = bf16hi(q.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 q;
  q.x = pack_bf16x2(f[0], f[1]); q.y = pack_bf16x2(f[2], f[3]);
  q.z = pack_bf16x2(f[4], f[5]); q.w = pack_bf16x2(f[6], f[7]);
  return q;
}

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// GPT-2 "gelu_new" and its derivative (tanh form)
__device__ __forceinline__ float gelu_new_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float t = tanh_approx(k0 * (x + k1 * x * x * x));
  return 0.5f * x * (1.f + t);
}
__device__ __forceinline__ float gelu_new_grad_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float x2 = x * x;
  float t = tanh_approx(k0 * (x + k1 * x * x2));
  float dt = (1.f - t * t) * k0 * (1.f + 3.f * k1 * x2);
  return 0.5f * (1.f + t) + 0.5f * x * dt;
}

// ----------------------------------------------------------------------------------
// packed fp32 pairs (sm_100 FFMA2 / FMUL2 / FADD2): one instruction per two elements.  A pair
// lives in a 64-bit register; pack/unpack are register-pair moves the assembler elides.
// ----------------------------------------------------------------------------------
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t pk2(float lo, float hi) {
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(f32x2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2_t fma2(f32x2_t a, f32x2_t b, f32x2_t c) {
  f32x2_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2_t mul2(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t add2(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t sub2(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t tanh2(f32x2_t u) {
  float a, b;
  upk2(u, a, b);
  return pk2(tanh_approx(a), tanh_approx(b));
}
// a bf16x2 word <-> a packed fp32 pair (.lo = low 16 bits)
__device__ __forceinline__ f32x2_t bf16x2_to_f32x2(uint32_t w) { return pk2(bf16lo(w), bf16hi(w)); }
__device__ __forceinline__ uint32_t f32x2_to_bf16x2(f32x2_t v) {
  float lo, hi;
  upk2(v, lo, hi);
  return pack_bf16x2(lo, hi);
}
// gelu_new and its derivative on a pair; same formulas as the scalar versions above
__device__ __forceinline__ f32x2_t gelu_new2(f32x2_t x) {
  const f32x2_t k0 = pk2(0.7978845608028654f, 0.7978845608028654f);
  const f32x2_t k0k1 = pk2(0.7978845608028654f * 0.044715f, 0.7978845608028654f * 0.044715f);
  const f32x2_t half = pk2(0.5f, 0.5f);
  const f32x2_t x2 = mul2(x, x);
  const f32x2_t t = tanh2(mul2(x, fma2(x2, k0k1, k0)));
  const f32x2_t hx = mul2(x, half);
  return fma2(hx, t, hx);
}
__device__ __forceinline__ f32x2_t gelu_new_grad2(f32x2_t x) {
  const f32x2_t k0 = pk2(0.7978845608028654f, 0.7978845608028654f);
  const f32x2_t k0k1 = pk2(0.7978845608028654f * 0.044715f, 0.7978845608028654f * 0.044715f);
  const f32x2_t k0k1x3 = pk2(3.f * 0.7978845608028654f * 0.044715f, 3.f * 0.7978845608028654f * 0.044715f);
  const f32x2_t half = pk2(0.5f, 0.5f), one = pk2(1.f, 1.f);
  const f32x2_t x2 = mul2(x, x);
  const f32x2_t t = tanh2(mul2(x, fma2(x2, k0k1, k0)));
  const f32x2_t sech2 = sub2(one, mul2(t, t));
  const f32x2_t w = mul2(mul2(x, half), fma2(x2, k0k1x3, k0));   // 0.5 x k0 (1 + 3 k1 x^2)
  return fma2(w, sech2, fma2(half, t, half));
}

#endif  // __CUDACC__
