// Adapter-gradient contraction over the token dimension:
//   out[i, c] += alpha * sum_m S[m, i] * T[m, c]        i < r (8, 16, or 32 = two stacked adapters), c < cols
// used for dB = (s*X*A)^T * dY  (S = U, T = dY) and dA^T = dU^T * X  (S = dU, T = X).
//
// Both operands have the reduction index (tokens) as their slow dimension.  That is exactly
// the MN-major x MN-major case of the tcgen05 GEMM: D[c, i] = sum_m T[m, c] * S[m, i] with
// A = T (M' = cols), B = S (N' = r).  The output is tiny (cols x r) and the reduction long
// (tokens), so the work is split over K across the SMs and the partial tiles are reduced with
// fp32 atomics straight into out[i, c] (the transposed-accumulate epilogue).  HBM-bound on
// T: algorithmic bytes = 2*tokens*(cols + r).
#include "common.cuh"
#include "../../include/b200ft.h"

extern "C" int b200ft_lora_grad(const void* s, int64_t ld_s, int r, const void* t, int64_t ld_t,
                                int cols, int64_t m, float* out, int64_t ld_out, float alpha,
                                void* stream) {
  B200FT_CHECK_ARG(s && t && out, "lora_grad: null buffer");
  B200FT_CHECK_ARG((r == 8 || r == 16 || r == 32) && ld_s >= r && ld_s % 8 == 0, "lora_grad: r=%d ld_s=%lld (r in {8,16,32}, ld_s %% 8 == 0)",
                   r, (long long)ld_s);
  B200FT_CHECK_ARG(cols > 0 && ld_t >= cols && ld_t % 8 == 0 && ld_out >= cols && m > 0 && m < (1ll << 31),
                   "lora_grad: cols=%d ld_t=%lld ld_out=%lld m=%lld", cols, (long long)ld_t,
                   (long long)ld_out, (long long)m);
  b200ft_gemm_args g;
  memset(&g, 0, sizeof(g));
  g.m = cols; g.n = r; g.k1 = (int32_t)m;
  g.a1 = t; g.lda1 = ld_t;
  g.b1 = s; g.ldb1 = ld_s;
  g.layout = B200FT_A_MN_MAJOR | B200FT_B_MN_MAJOR;
  g.epilogue = B200FT_EPI_ATOMIC_F32_T;
  g.alpha = alpha;
  g.c = out; g.ldc = ld_out;
  return b200ft_gemm_bf16(&g, stream);
}
