// tcgen05 / TMA / mbarrier plumbing shared by the attention kernels (attention.cu, attention_bwd.cu): thin PTX
// wrappers, the SW128 shared-memory matrix descriptor, the kind::f16 instruction descriptor, and the host-side
// tensor-map helpers.  (The GEMM keeps its own copies: its cta_group::2 forms and descriptor caching differ.)
#pragma once
#include <cuda.h>
#include <mutex>
#include "common.cuh"
#include "../../include/b200ft.h"

namespace b200ft_attn {

constexpr long long kWaitLimit = 4000000000ll;   // clock64 ticks (~2 s) before a barrier wait traps

static __device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
static __device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kWaitLimit) __trap();  // deadlock guard
  }
}
static __device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
static __device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
static __device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
static __device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
static __device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
static __device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
static __device__ __forceinline__ float ex2(float x) {
#ifdef B200FT_EXP_NO_EX2   // bottleneck experiment (tools/attn_exp.py): no MUFU work, wrong numbers
  return x * 0.001f;
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}
static __device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
static __device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// smem matrix descriptor (SW128): start>>4 | LBO>>4 << 16 || SBO>>4 | version 1 | swizzle 128B
static __device__ __forceinline__ uint64_t sdesc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  const uint32_t lo = ((addr & 0x3FFFFu) >> 4) | (((lbo >> 4) & 0x3FFFu) << 16);
  const uint32_t hi = ((sbo >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// kind::f16 instruction descriptor: fp32 accumulate, bf16 x bf16, M = 128
static __device__ __forceinline__ uint32_t idesc(int n, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(n >> 3) << 17) | (8u << 24);
}

// 2^x for a pair on the FMA pipe instead of the special-function unit (16 exponentials per clock and SM is the
// forward kernel's longest resource at d_head 64): x = n + f with n = round(x) taken from the mantissa of x + 1.5 * 2^23,
// 2^f on [-0.5, 0.5] by a degree-3 polynomial (relative error 1e-4: P is rounded to bf16, 4e-3, right after), 2^n added
// into the exponent field.  x is clamped at -126: masked scores (-inf) and underflow give 2^-126, i.e. 0 in bf16.
static __device__ __forceinline__ f32x2_t exp2_fma2(f32x2_t x) {
  float x0, x1;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x0), "=f"(x1) : "l"(x));
  x0 = fmaxf(x0, -126.f); x1 = fmaxf(x1, -126.f);
  f32x2_t xc, t, f, q;
  asm("mov.b64 %0, {%1, %2};" : "=l"(xc) : "f"(x0), "f"(x1));
  const float mg = 12582912.f;
  f32x2_t magic, c0, c1, c2, c3;
  asm("mov.b64 %0, {%1, %1};" : "=l"(magic) : "f"(mg));
  asm("mov.b64 %0, {%1, %1};" : "=l"(c0) : "f"(0.99992454f));
  asm("mov.b64 %0, {%1, %1};" : "=l"(c1) : "f"(0.69313675f));
  asm("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(0.24263948f));
  asm("mov.b64 %0, {%1, %1};" : "=l"(c3) : "f"(0.05583828f));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(xc), "l"(magic));
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(f) : "l"(t), "l"(magic));      // n as a float
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(f) : "l"(xc), "l"(f));         // f = x - n
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(q) : "l"(c3), "l"(f), "l"(c2));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(q) : "l"(q), "l"(f), "l"(c1));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(q) : "l"(q), "l"(f), "l"(c0));
  uint32_t q0, q1, t0, t1;
  asm("mov.b64 {%0, %1}, %2;" : "=r"(q0), "=r"(q1) : "l"(q));
  asm("mov.b64 {%0, %1}, %2;" : "=r"(t0), "=r"(t1) : "l"(t));
  q0 += t0 << 23; q1 += t1 << 23;                                       // the low mantissa bits of t hold n (two's complement)
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(q0), "r"(q1));
  return r;
}
static __device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* w) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]),
        "r"(w[8]), "r"(w[9]), "r"(w[10]), "r"(w[11]), "r"(w[12]), "r"(w[13]), "r"(w[14]), "r"(w[15]) : "memory");
}
static __device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* w) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(taddr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
}
static __device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem descriptor]
static __device__ __forceinline__ void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

// ---- host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)f;
  });
  return fn;
}

// [batch][seq][columns] view of a [batch*seq, ld] bf16 buffer, boxes of 64 columns x 128 rows, 128B swizzle;
// rows past the end of a sequence are zero-filled
static inline int make_map(EncodeTiledFn enc, CUtensorMap* tm, const void* base, int64_t ld, int seq, int batch) {
  cuuint64_t dims[3] = {(cuuint64_t)ld, (cuuint64_t)seq, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)ld * 2 * (cuuint64_t)seq};
  cuuint32_t box[3] = {64, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { b200ft_set_error("attention: cuTensorMapEncodeTiled failed (%d)", (int)r); return B200FT_ERR_ARG; }
  return 0;
}

// d_head 128 backward (attention_bwd_d128.cu): dK|dV kernel + dQ kernel; `delta` has been computed by the caller
struct Bwd128Args {
  int batch, seq, n_head, n_kv_head;
  int q_off, k_off, v_off;
  float scale;
  const float* lse2;     // [batch*n_head, seq], log2 domain incl. the softmax scale
  const float* delta;    // [batch*n_head, seq]
  __nv_bfloat16* dqkv; int64_t ld_d;
  const __nv_bfloat16* qkv; int64_t ld_qkv;     // the same buffers the tensor maps describe: the dQ kernel reads the
  const __nv_bfloat16* dout; int64_t ld_dout;   // rows of its resident Q / dO tile directly (A operands kept in TMEM)
};
int launch_bwd_d128(const CUtensorMap& tm_qkv, const CUtensorMap& tm_do, const Bwd128Args& a, cudaStream_t st);

}  // namespace b200ft_attn
