// Fused softmax-cross-entropy forward + backward over bf16 logits, in place.
//
// One CTA per row.  Sweep 1 streams the row from HBM with 16-byte loads keeping an online
// (max, sum-of-exp2) pair per thread, folded with warp shuffles and one shared-memory hop.
// Sweep 2 re-reads the row (it is still in the 126 MB L2: 148 SMs x resident CTAs x 100 KB)
// and overwrites it with (softmax - onehot) * grad_scale / n_valid.  Algorithmic HBM traffic is
// therefore one read + one write of the row: 2*rows*ld*2 bytes, plus 12 bytes per row.
//
// Scratch (the caller's workspace): [0] number of valid targets, [1] completion ticket,
// [16 .. 16+rows) per-row losses.  The last CTA to finish folds the row losses in a fixed
// order, so the loss is bit-reproducible from run to run.
#include <mutex>
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__global__ void __launch_bounds__(256)
count_valid_kernel(const int64_t* __restrict__ targets, int64_t rows, int64_t ignore_index,
                   unsigned int* __restrict__ ws) {
  // single CTA: rows is at most a few 10^5, this is ~1 us and keeps the count exact
  unsigned int c = 0;
  for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) c += (targets[i] != ignore_index);
  __shared__ unsigned int red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = 0;
    for (int w = 0; w < 8; ++w) t += red[w];
    ws[0] = t;
    ws[1] = 0;  // ticket
  }
}

struct MS { float m, s; };
__device__ __forceinline__ MS ms_merge(MS a, MS b) {
  float M = fmaxf(a.m, b.m);
  MS r;
  r.m = M;
  // guard (-inf) - (-inf): a fully masked partial contributes 0
  float ea = (a.m == -INFINITY) ? 0.f : exp2f(a.m - M);
  float eb = (b.m == -INFINITY) ? 0.f : exp2f(b.m - M);
  r.s = a.s * ea + b.s * eb;
  return r;
}

__global__ void __launch_bounds__(kThreads)
softmax_xent_kernel(__nv_bfloat16* __restrict__ logits, int64_t rows, int vocab, int64_t ld,
                    const int64_t* __restrict__ targets, int64_t ignore_index, float grad_scale,
                    float* __restrict__ loss_out, unsigned int* __restrict__ ws) {
  const int64_t row = blockIdx.x;
  __nv_bfloat16* z = logits + row * ld;
  const int nvec = (int)(ld >> 3);  // the whole padded row is rewritten
  const int64_t tgt = targets[row];
  const bool valid = (tgt != ignore_index);
  float* row_loss = reinterpret_cast<float*>(ws + 16);
  __shared__ MS red[kThreads / 32];
  __shared__ float s_lse2;
  __shared__ bool s_last;

  if (valid) {
    // ---- sweep 1: online max / sum in the log2 domain, kUnroll independent 16-byte loads in
    // flight per thread (the running max is revisited once per 8*kUnroll elements) ----
    MS acc{-INFINITY, 0.f};
    for (int i0 = threadIdx.x; i0 < nvec; i0 += kThreads * kUnroll) {
      uint4 q[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        if (i < nvec) q[u] = ld_cg16(z + ((int64_t)i << 3));
      }
      float f[kUnroll][8];
      float vmax = -INFINITY;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        const int c0 = i << 3;
        if (i < nvec) {
          unpack8(q[u], f[u]);
          if (c0 + 8 <= vocab) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { f[u][j] *= kLog2e; vmax = fmaxf(vmax, f[u][j]); }
          } else {  // the vector straddling (or beyond) the end of the vocabulary
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              f[u][j] = (c0 + j < vocab) ? f[u][j] * kLog2e : -INFINITY;
              vmax = fmaxf(vmax, f[u][j]);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[u][j] = -INFINITY;
        }
      }
      if (vmax > acc.m) {  // rare after the first few iterations
        acc.s *= ex2_approx(acc.m - vmax);  // ex2(-inf) = 0 on the first hit
        acc.m = vmax;
      }
      if (acc.m != -INFINITY) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc.s += ex2_approx(f[u][j] - acc.m);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      MS other;
      other.m = __shfl_xor_sync(0xffffffffu, acc.m, o);
      other.s = __shfl_xor_sync(0xffffffffu, acc.s, o);
      acc = ms_merge(acc, other);
    }
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      MS t = red[0];
      for (int w = 1; w < kThreads / 32; ++w) t = ms_merge(t, red[w]);
      float lse2 = t.m + log2f(t.s);
      s_lse2 = lse2;
      const int64_t tsafe = tgt < 0 ? 0 : (tgt >= vocab ? vocab - 1 : tgt);  // never read out of the row
      float zt = __bfloat162float(z[tsafe]) * kLog2e;
      row_loss[row] = (lse2 - zt) * kLn2;
    }
    __syncthreads();
  } else if (threadIdx.x == 0) {
    row_loss[row] = 0.f;
  }

  // ---- sweep 2: write the gradient in place (row re-read from L2) ----
  const unsigned int n_valid = ws[0];
  const float scale = valid ? grad_scale / (float)(n_valid > 0 ? n_valid : 1) : 0.f;
  const float lse2 = valid ? s_lse2 : 0.f;
  const int tcol = valid ? (int)tgt : -1;
  if (valid) {
    for (int i0 = threadIdx.x; i0 < nvec; i0 += kThreads * kUnroll) {
      uint4 q[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        if (i < nvec) q[u] = ld_cg16(z + ((int64_t)i << 3));
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        if (i < nvec) {
          const int c0 = i << 3;
          float f[8];
          unpack8(q[u], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = ex2_approx(f[j] * kLog2e - lse2) * scale;
          if (tcol >= c0 && tcol < c0 + 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (c0 + j == tcol) f[j] -= scale;
          }
          if (c0 + 8 > vocab) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (c0 + j >= vocab) f[j] = 0.f;
          }
          st_stream16(z + ((int64_t)i << 3), pack8(f));
        }
      }
    }
  } else {
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < nvec; i += kThreads) st_stream16(z + ((int64_t)i << 3), zero);
  }

  // ---- last CTA folds the per-row losses ----
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(ws + 1, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    float t = 0.f;
    for (int64_t i = threadIdx.x; i < rows; i += kThreads) t += __ldcg(row_loss + i);
    t = warp_sum(t);
    __shared__ float fr[kThreads / 32];
    if ((threadIdx.x & 31) == 0) fr[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      float sum = 0.f;
      for (int w = 0; w < kThreads / 32; ++w) sum += fr[w];
      loss_out[0] = sum / (float)(n_valid > 0 ? n_valid : 1);
    }
  }
}

__global__ void shift_targets_kernel(const int64_t* __restrict__ labels, int64_t* __restrict__ targets,
                                     int64_t total, int seq, int64_t ignore_index) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int t = (int)(i % seq);
  targets[i] = (t == seq - 1) ? ignore_index : labels[i + 1];
}

}  // namespace

extern "C" size_t b200ft_softmax_xent_workspace_bytes(int64_t rows) {
  return (size_t)(rows + 16) * sizeof(float);
}

extern "C" int b200ft_softmax_xent_fwd_bwd(void* logits, int64_t rows, int vocab, int64_t ld,
                                           const int64_t* targets, int64_t ignore_index,
                                           float grad_scale, float* loss_out, void* workspace,
                                           size_t workspace_bytes, void* stream) {
  B200FT_CHECK_ARG(logits && targets && loss_out && workspace, "softmax_xent: null buffer");
  B200FT_CHECK_ARG(rows > 0 && rows < (1ll << 31) && vocab > 0 && ld >= vocab && ld % 8 == 0,
                   "softmax_xent: rows=%lld vocab=%d ld=%lld (ld must be a multiple of 8 and >= vocab)",
                   (long long)rows, vocab, (long long)ld);
  B200FT_CHECK_ARG((uintptr_t)logits % 16 == 0, "softmax_xent: logits must be 16-byte aligned");
  if (workspace_bytes < b200ft_softmax_xent_workspace_bytes(rows)) {
    b200ft_set_error("softmax_xent: workspace %zu < %zu", workspace_bytes,
                     b200ft_softmax_xent_workspace_bytes(rows));
    return B200FT_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* ws = (unsigned int*)workspace;
  count_valid_kernel<<<1, 256, 0, st>>>(targets, rows, ignore_index, ws);
  B200FT_LAUNCH_CHECK();
  softmax_xent_kernel<<<(unsigned)rows, kThreads, 0, st>>>((__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                                           ignore_index, grad_scale, loss_out, ws);
  B200FT_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant.  Some callers of this library are not in a position to hand over a
// workspace: a host that has given every byte of device memory to its own pool and may not
// call cudaMalloc again, or a plug-in running inside another framework's allocator where a
// fresh allocation would fragment or stall it.  For them the scratch of the cross-entropy
// kernel (valid-target count, completion ticket, one loss per row) lives in this file-scope
// device array instead.  It costs (kNoAllocMaxRows + 16) * 4 bytes on every GPU that loads the
// library, until the process exits.
//
// The array's name is only a host-side handle; the kernels need its device address, which
// cudaGetSymbolAddress() yields for the device that is current at the time.  That look-up is
// done once per device, on the first call made with that device current.
// ---------------------------------------------------------------------------------------
namespace {
constexpr int64_t kNoAllocMaxRows = 65536;
static __device__ __align__(256) unsigned int g_xent_static_scratch[kNoAllocMaxRows + 16];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
unsigned int* g_scratch_addr[kMaxDevices] = {nullptr};

// Device address of the static scratch on the CURRENT device (resolved once per device).
int static_scratch_for_current_device(unsigned int** out) {
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  B200FT_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "softmax_xent_noalloc: device ordinal %d unsupported", dev);
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (g_scratch_addr[dev] == nullptr) {
    unsigned int* ptr = nullptr;
    B200FT_CUDA(cudaGetSymbolAddress((void**)&ptr, g_xent_static_scratch));
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return 0;
}
}  // namespace

extern "C" int64_t b200ft_softmax_xent_noalloc_max_rows(void) { return kNoAllocMaxRows; }

extern "C" int b200ft_softmax_xent_fwd_bwd_noalloc(void* logits, int64_t rows, int vocab, int64_t ld,
                                                   const int64_t* targets, int64_t ignore_index,
                                                   float grad_scale, float* loss_out, void* stream) {
  B200FT_CHECK_ARG(logits && targets && loss_out, "softmax_xent_noalloc: null buffer");
  B200FT_CHECK_ARG(rows > 0 && vocab > 0 && ld >= vocab && ld % 8 == 0,
                   "softmax_xent_noalloc: rows=%lld vocab=%d ld=%lld (ld must be a multiple of 8 and >= vocab)",
                   (long long)rows, vocab, (long long)ld);
  if (rows > kNoAllocMaxRows) {
    b200ft_set_error("softmax_xent_noalloc: rows=%lld exceeds the built-in scratch (max %lld rows); split the "
                     "batch or use b200ft_softmax_xent_fwd_bwd with a caller workspace",
                     (long long)rows, (long long)kNoAllocMaxRows);
    return B200FT_ERR_WORKSPACE;
  }
  B200FT_CHECK_ARG((uintptr_t)logits % 16 == 0, "softmax_xent_noalloc: logits must be 16-byte aligned");
  unsigned int* ws = nullptr;
  int rc = static_scratch_for_current_device(&ws);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // count_valid_kernel rewrites the count and zeroes the ticket; every row's loss slot is
  // rewritten by its CTA, so nothing of the previous call's contents is read.
  count_valid_kernel<<<1, 256, 0, st>>>(targets, rows, ignore_index, ws);
  B200FT_LAUNCH_CHECK();
  softmax_xent_kernel<<<(unsigned)rows, kThreads, 0, st>>>((__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                                           ignore_index, grad_scale, loss_out, ws);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_shift_targets(const int64_t* labels, int64_t* targets, int64_t batch, int seq,
                                    int64_t ignore_index, void* stream) {
  B200FT_CHECK_ARG(labels && targets && batch > 0 && seq > 0, "shift_targets: bad arguments");
  int64_t total = batch * seq;
  shift_targets_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      labels, targets, total, seq, ignore_index);
  B200FT_LAUNCH_CHECK();
  return 0;
}
