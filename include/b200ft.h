/*
 * libb200ft — C ABI of the B200-native fine-tuning step.
 *
 * What this boundary replaces.  The reference (acceleratedscience/finetune-controller) has
 * no FFI and no compute code: its model plugin names a container image and a shell command
 * (reference app/models/base/finetuning.py:51-108, app/models/examples/mnist.py:41-99) and
 * app/jobs/kubeflow/PyTorchJobDeployer.py:93-119 runs that command in the "pytorch"
 * container.  The arithmetic the command performs lives in an external image that is absent
 * from the reference tree, so there is no reference symbol to bind one-to-one.  Each entry
 * point below therefore cites the op of the training step it implements (the row of
 * SURVEY.md §8(a2), restated on CPU in oracle/gpt2_lora_ref.py) instead of a reference
 * function; INTEGRATION.md shows the ctypes binding the payload uses.
 *
 * Conventions (all entry points):
 *   - return 0 on success; a negative B200FT_ERR_* for rejected arguments; a positive
 *     cudaError_t for a CUDA failure.  b200ft_last_error_string() describes the last failure
 *     on the calling host thread.  Nothing throws across the boundary.
 *   - every pointer is a DEVICE pointer unless its name ends in _host.  The caller owns all
 *     memory; the library allocates none on the device.  Scratch is passed in as
 *     (workspace, workspace_bytes) sized by the matching *_workspace_bytes().
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued and the call returns
 *     without synchronising.
 *   - bf16 tensors are row-major with an explicit leading dimension in ELEMENTS; bases are
 *     16-byte aligned and leading dimensions are multiples of 8 elements.
 *   - there is no CPU fallback: on a machine without an sm_100 GPU the calls fail.
 */
#ifndef B200FT_H_
#define B200FT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200FT_ERR_ARG (-1)
#define B200FT_ERR_UNSUPPORTED (-2)
#define B200FT_ERR_WORKSPACE (-3)

/* library identity / diagnostics */
const char* b200ft_version(void);
const char* b200ft_last_error_string(void);

/* ---------------------------------------------------------------------------------------
 * Projection / LoRA GEMM on tcgen05 tensor cores (SURVEY.md §8 a2.1, a2.2, a2.3, a2.5;
 * oracle/gpt2_lora_ref.py:_linear and its autograd transpose).
 *
 *   acc[m,n] = sum_s sum_{k<k_s} A_s[m,k]*B_s[n,k],  s = 1..4                 (fp32 in TMEM)
 *   v        = alpha*acc + bias[n]
 *   epilogue NONE     : C = v
 *            RESIDUAL : C = v + AUX[m,n]
 *            GELU     : C = v, C2 = gelu_new(v)
 *            DGELU    : C = v * gelu_new'(AUX[m,n])
 *            ATOMIC_F32_T : C is fp32 [n, ldc]; C[n_i, m_i] += alpha*acc[m_i, n_i] (atomic), so
 *                       the reduction can be split over K slabs (split_k) across CTAs; no bias.
 *                       Used for the adapter-gradient contractions over tokens.
 *
 * Extra K segments are how a LoRA branch is fused: with A2 = s*(X*A_lora) and B2 = B_lora^T the
 * rank-r product lands in the same TMEM accumulator as the frozen projection; backward of the
 * fused qkv projection uses four (dq, dk, dv and the adapter term).  Operands are bf16.
 * "K-major" = the reduction index is contiguous (A: [M, lda], B: [N, ldb]); "MN-major" = the M
 * (resp. N) index is contiguous (A: [K, lda], B: [K, ldb]).  Segments 2..4 are always K-major;
 * a NULL a_i ends the list.
 * ------------------------------------------------------------------------------------- */
enum {
  B200FT_EPI_NONE = 0,
  B200FT_EPI_RESIDUAL = 1,
  B200FT_EPI_GELU = 2,
  B200FT_EPI_DGELU = 3,
  B200FT_EPI_ATOMIC_F32_T = 4
};
enum { B200FT_A_MN_MAJOR = 1, B200FT_B_MN_MAJOR = 2 };

typedef struct b200ft_gemm_args {
  int32_t m, n;
  const void* a1; int64_t lda1;
  const void* b1; int64_t ldb1;
  int32_t k1;
  const void* a2; int64_t lda2;   /* NULL when there is no second segment */
  const void* b2; int64_t ldb2;
  int32_t k2;
  const void* a3; int64_t lda3;
  const void* b3; int64_t ldb3;
  int32_t k3;
  const void* a4; int64_t lda4;
  const void* b4; int64_t ldb4;
  int32_t k4;
  int32_t layout;                 /* B200FT_A_MN_MAJOR | B200FT_B_MN_MAJOR, segment 1 only */
  int32_t epilogue;               /* B200FT_EPI_* */
  float alpha;
  const float* bias;              /* [n] fp32 or NULL */
  const void* aux; int64_t ld_aux;/* bf16 [m, ld_aux] for RESIDUAL / DGELU */
  void* c; int64_t ldc;           /* bf16 [m, ldc] */
  void* c2; int64_t ldc2;         /* bf16 [m, ldc2], GELU only */
  int32_t block_n;                /* 0 = choose; else multiple of 16 in [16,256] */
  int32_t max_ctas;               /* 0 = one per SM */
  int32_t split_k;                /* ATOMIC_F32_T only: 0 = choose, else number of K splits */
  int32_t cta_group;              /* 0 = choose; 1 = one CTA per 128 x block_n tile; 2 = CTA pair per
                                     256 x block_n tile (K-major or MN-major operands; non-atomic epilogues only) */
} b200ft_gemm_args;

int b200ft_gemm_bf16(const b200ft_gemm_args* args_host, void* stream);

/* ---------------------------------------------------------------------------------------
 * Adapter-gradient contraction over tokens (dB = s*(XA)^T*dY,
 * dA = X^T*(s*dY*B^T)):   out[i, c] += alpha * sum_m S[m, i] * T[m, c],  i < r, c < cols.
 * S, T bf16; out fp32 [r, ld_out], accumulated into (zero it first).  r = 8, 16 or 32 (two stacked adapters); ld_s, ld_t
 * multiples of 8.  Runs on the tcgen05 GEMM (both operands MN-major, split-K over tokens,
 * ATOMIC_F32_T epilogue): the sum order across splits is not fixed, results are reproducible
 * to fp32 rounding, not bitwise.
 * ------------------------------------------------------------------------------------- */
int b200ft_lora_grad(const void* s, int64_t ld_s, int r, const void* t, int64_t ld_t, int cols,
                     int64_t m, float* out, int64_t ld_out, float alpha, void* stream);

/* ---------------------------------------------------------------------------------------
 * Fused softmax-cross-entropy forward+backward (SURVEY.md §8 a2.4;
 * oracle/gpt2_lora_ref.py:softmax_xent_ref).  logits bf16 [rows, ld], the first `vocab`
 * columns valid; overwritten in place with d(loss)/d(logits) * grad_scale (padding columns
 * and ignored rows get 0).  loss_out[0] = mean over rows whose target != ignore_index.
 * A target outside [0, vocab) other than ignore_index is treated like an ignored one (no loss, zero
 * gradient row, not counted in the mean) and counted in 32-bit scratch word 2, which the host may read
 * after the call has completed (word 0: number of valid targets).
 * Scratch: b200ft_softmax_xent_workspace_bytes(rows) bytes.
 * ------------------------------------------------------------------------------------- */
size_t b200ft_softmax_xent_workspace_bytes(int64_t rows);
int b200ft_softmax_xent_fwd_bwd(void* logits, int64_t rows, int vocab, int64_t ld,
                                const int64_t* targets, int64_t ignore_index, float grad_scale,
                                float* loss_out, void* workspace, size_t workspace_bytes,
                                void* stream);

/* Allocation-free form of b200ft_softmax_xent_fwd_bwd (no counterpart in the reference; it
 * exists for this library's own callers).  Same arguments minus the workspace: the scratch
 * is a fixed array inside the library's device image, so a caller that cannot allocate device
 * memory at this point (its pool owns all of it, or it runs inside another framework's
 * allocator) can still compute the loss.  Limits and rules:
 *   - rows <= b200ft_softmax_xent_noalloc_max_rows() (65536); more is refused with
 *     B200FT_ERR_WORKSPACE and a message, nothing is launched;
 *   - the array (about 256 KB) is resident on every GPU that loads the library until exit;
 *   - the first call made with a given device current resolves the array's device address
 *     (cudaGetSymbolAddress); make that first call outside any stream capture;
 *   - all calls on a device share the one array: two calls that may overlap on the GPU
 *     (different streams, no dependency) race and give wrong losses — serialise them;
 *   - the call itself allocates and frees nothing, on the host or the device. */
int64_t b200ft_softmax_xent_noalloc_max_rows(void);
int b200ft_softmax_xent_fwd_bwd_noalloc(void* logits, int64_t rows, int vocab, int64_t ld,
                                        const int64_t* targets, int64_t ignore_index, float grad_scale,
                                        float* loss_out, void* stream);

/* HF causal-LM target shift (oracle/gpt2_lora_ref.py:causal_lm_loss):
 * targets[b, t] = labels[b, t+1] for t < seq-1, ignore_index at t = seq-1. */
int b200ft_shift_targets(const int64_t* labels, int64_t* targets, int64_t batch, int seq,
                         int64_t ignore_index, void* stream);

/* ---------------------------------------------------------------------------------------
 * Fused AdamW + gradient scale (oracle/gpt2_lora_ref.py:adamw_ref, the
 * formula of torch.optim.AdamW).  Flat buffers of n elements: p, m, v fp32; g bf16 or fp32;
 * optional bf16 shadow of the updated p.  g is multiplied by grad_scale and, when
 * grad_scale_dev != NULL, also by *grad_scale_dev (a device scalar, e.g. a clip factor).
 * 28 bytes of HBM traffic per element with a bf16 g and a shadow.
 * ------------------------------------------------------------------------------------- */
int b200ft_adamw_step(float* p, float* m, float* v, const void* g, int g_is_bf16, void* shadow_bf16,
                      int64_t n, float lr, float beta1, float beta2, float eps,
                      float weight_decay, int step, float grad_scale,
                      const float* grad_scale_dev, void* stream);
/* Same update on at most max_blocks thread blocks of 256 threads (0 = no limit: up to 8 per SM).  For a launch
 * that runs BESIDE other kernels: the full-parameter step updates a layer on a second stream while backward
 * still computes the layers below it, and one block per SM (148) leaves the GEMM's CTAs room to stay resident. */
int b200ft_adamw_step_limited(float* p, float* m, float* v, const void* g, int g_is_bf16, void* shadow_bf16,
                              int64_t n, float lr, float beta1, float beta2, float eps, float weight_decay,
                              int step, float grad_scale, const float* grad_scale_dev, int max_blocks,
                              void* stream);

/* Sum of squares of a flat gradient buffer -> out_sq[0] (fp32), and the clip factor
 * coef[0] = min(1, max_norm / (pre_scale*sqrt(norm_sq) + 1e-6))  (torch clip_grad_norm_). */
size_t b200ft_gradnorm_sq_workspace_bytes(int64_t n);
int b200ft_gradnorm_sq(const void* g, int g_is_bf16, int64_t n, float* out_sq, void* workspace,
                       size_t workspace_bytes, void* stream);
int b200ft_clip_coef(const float* norm_sq, float max_norm, float pre_scale, float* coef, void* stream);

/* ---------------------------------------------------------------------------------------
 * LayerNorm (F.layer_norm in oracle forward()).  x, y, dy, dx, dres bf16
 * [rows, d]; gamma/beta fp32 [d]; mean/rstd fp32 [rows].  bwd_dx: dx = dres + dLN/dx
 * (dres may be NULL): the residual-stream gradient add is fused.
 * d must be a multiple of 8 and <= 8192.
 * ------------------------------------------------------------------------------------- */
int b200ft_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean,
                         float* rstd, int64_t rows, int d, float eps, void* stream);
int b200ft_layernorm_bwd_dx(const void* dy, const void* x, const float* gamma, const float* mean,
                            const float* rstd, const void* dres, void* dx, int64_t rows, int d,
                            void* stream);

/* ---------------------------------------------------------------------------------------
 * Llama-family pieces (BASELINE.json configs[2..4]; oracle/llama_lora_ref.py).
 * RMSNorm: y = x * rstd * gamma, rstd = rsqrt(mean(x^2) + eps) saved per row; bwd_dx fuses the
 * residual-gradient add like the LayerNorm one.  d %% 8 == 0, d <= 4096.
 * RoPE: rotates, in place, the first n_heads * d_head columns of each row of a bf16 matrix with
 * pitch ld (the q and k heads of a fused qkv buffer) by the angle table cos/sin [seq, d_head/2]
 * (fp32), pairs (i, i + d_head/2) — HF's rotate_half; position = row %% seq; backward != 0 applies
 * the inverse rotation (the gradient map).
 * SwiGLU: gate_up [rows, 2*d_ff] (gate columns first) -> act = silu(gate) * up, and its backward.
 * ------------------------------------------------------------------------------------- */
int b200ft_rmsnorm_fwd(const void* x, const float* gamma, void* y, float* rstd, int64_t rows, int d, float eps,
                       void* stream);
int b200ft_rmsnorm_bwd_dx(const void* dy, const void* x, const float* gamma, const float* rstd, const void* dres,
                          void* dx, int64_t rows, int d, void* stream);
int b200ft_rope_inplace(void* x, int64_t ld, int64_t rows, int seq, int n_heads, int d_head,
                        const float* cos_table, const float* sin_table, int backward, void* stream);
int b200ft_swiglu_fwd(const void* gate_up, void* act, int64_t rows, int d_ff, void* stream);
int b200ft_swiglu_bwd(const void* dact, const void* gate_up, void* dgate_up, int64_t rows, int d_ff, void* stream);
/* same, and also rewrites act = silu(gate) * up [rows, d_ff] from gate_up (may be NULL): the full-parameter
 * step does not keep the forward product and needs it again for the down-projection's weight gradient */
int b200ft_swiglu_bwd_act(const void* dact, const void* gate_up, void* dgate_up, void* act, int64_t rows,
                          int d_ff, void* stream);

/* ---------------------------------------------------------------------------------------
 * Causal flash-attention forward on tcgen05 (oracle causal_attention()).
 * q, k, v are read in place from the fused projection buffer qkv [batch*seq, ld] (bf16): head h of
 * q starts at column q_off + h*d_head, kv head g at k_off + g*d_head / v_off + g*d_head, with
 * g = h / (n_head / n_kv_head) (grouped-query attention).  out [batch*seq, ld_out] receives head h
 * at column h*d_head; lse2 [batch*n_head, seq] receives max + log2(sum) of the scaled scores in
 * the log2 domain (what the backward needs).  d_head = 64 or 128; other sizes return
 * B200FT_ERR_UNSUPPORTED.  seq may be ragged (not a multiple of the 128-row tile).
 * ------------------------------------------------------------------------------------- */
int b200ft_attn_fwd(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, void* out, int64_t ld_out,
                    float* lse2, int batch, int seq, int n_head, int n_kv_head, int d_head, float scale,
                    void* stream);

/* Causal flash-attention backward on tcgen05, the counterpart of b200ft_attn_fwd (same layout
 * conventions; oracle: autograd of causal_attention()).  out / dout [batch*seq, ld_*] carry head h at
 * column h*d_head; lse2 is the forward's output.  dqkv [batch*seq, ld_dqkv] receives dq of head h at
 * q_off + h*d_head and dk / dv of kv head g at k_off / v_off + g*d_head (summed over the q heads of the
 * group), i.e. the column layout of the fused projection output itself; every element of those
 * column ranges is written.  d_head must be 64 or 128; other sizes return B200FT_ERR_UNSUPPORTED.
 * workspace: 256-byte aligned scratch of b200ft_attn_bwd_workspace_bytes() bytes (row sums of dout*out).
 * d_head 64, flags = 0: one fused kernel per (kv tile, batch, kv head) computes dK, dV in TMEM and adds
 * its dQ partials, already scaled and rounded to bf16, into the (pre-zeroed) q columns of dqkv with TMA
 * reduce-add: one bf16 rounding per visible kv tile, summation order varying run to run in the last bits
 * of dq.  flags = B200FT_ATTN_BWD_DETERMINISTIC: two kernels (dK|dV, then dQ accumulated in fp32 in TMEM),
 * no atomics, bit-reproducible, ~1.3x the time.  d_head 128 always runs its two-kernel form
 * (bit-reproducible). */
#define B200FT_ATTN_BWD_DETERMINISTIC 1
size_t b200ft_attn_bwd_workspace_bytes(int batch, int seq, int n_head, int d_head);
int b200ft_attn_bwd(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, const void* out,
                    int64_t ld_out, const void* dout, int64_t ld_dout, const float* lse2, void* workspace,
                    size_t workspace_bytes, void* dqkv, int64_t ld_dqkv, int batch, int seq, int n_head,
                    int n_kv_head, int d_head, float scale, int flags, void* stream);

/* Token + position embedding (oracle forward(): F.embedding(ids,wte)+F.embedding(pos,wpe)).
 * wte bf16 [vocab, d], wpe bf16 [ctx, d], x bf16 [rows, d], position = row % seq.  wpe may be NULL
 * (Llama: rotary positions, no learned table). */
int b200ft_embed_fwd(const int64_t* ids, const void* wte, const void* wpe, void* x, int64_t rows,
                     int seq, int d, int vocab, void* stream);

/* ---------------------------------------------------------------------------------------
 * Full-parameter fine-tune only (BASELINE.json configs[4]; SURVEY.md §8 a2.5 "full-FT adds dW",
 * a2.9).  The weight gradients themselves are b200ft_gemm_bf16 calls with MN-major operands;
 * these two cover what is not a GEMM.  Oracle: autograd of oracle/llama_lora_ref.py rmsnorm() /
 * F.embedding under FullOracleTrainer.
 *   rmsnorm_bwd_dgamma: dgamma[c] += sum_r dy[r,c] * x[r,c] * rstd[r]   (fp32 atomics: zero it first)
 *   embed_bwd:          dwte[ids[r], :] += dx[r, :]   (no atomics: the first row of every id sums all rows
 *                       of that id in fp32 and adds the sum to the bf16 table row once — one rounding however
 *                       often a token repeats, bit-reproducible; ids clamped to [0, vocab) like b200ft_embed_fwd;
 *                       d a multiple of 8)
 * --------------------------------------------------------------------------------------- */
int b200ft_rmsnorm_bwd_dgamma(const void* dy, const void* x, const float* rstd, float* dgamma,
                              int64_t rows, int d, void* stream);
int b200ft_embed_bwd(const int64_t* ids, const void* dx, void* dwte, int64_t rows, int d, int vocab,
                     void* stream);

/* out[c, r] = in[r, c] for bf16 matrices (weight re-layout at load; adapter shadows). */
int b200ft_transpose_bf16(const void* in, int64_t ld_in, void* out, int64_t ld_out, int rows,
                          int cols, void* stream);

/* fp32 -> bf16 cast of a flat buffer (parameter upload). */
int b200ft_cast_f32_to_bf16(const float* in, void* out, int64_t n, void* stream);

/* Stream-ordered zero fill of `bytes` bytes (gradient accumulators that the split-K and atomic kernels add into:
 * the step's "zero_grad").  A DMA memset, no kernel launch.  No counterpart in the reference (no compute there). */
int b200ft_zero(void* ptr, size_t bytes, void* stream);

/* Programmatic dependent launch, process-wide (default on; B200FT_PDL=0 in the environment forces it off).  With it on,
 * the CTAs of a kernel are dispatched while the previous kernel of the stream drains and wait, resident, for it to
 * finish.  Turn it off around work that shares the GPU with a kernel on ANOTHER stream which needs SMs of its own — the
 * data-parallel step does so for the backward pass whose GEMMs leave SMs to an overlapped all-reduce: resident, waiting
 * CTAs would take exactly those SMs (measured, 2 x B200, Llama-3-8B full-parameter step: 259.8 -> 263.3 ms).
 * Returns the previous setting.  No counterpart in the reference (no compute there). */
int b200ft_set_dependent_launch(int enabled);

#ifdef __cplusplus
}
#endif
#endif /* B200FT_H_ */
