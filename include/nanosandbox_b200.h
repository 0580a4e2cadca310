/*
 * nanosandbox_b200 -- C-ABI of the B200-native nanoGPT data-parallel training step.
 *
 * Boundary: the reference (fxcawley/nanoSandbox) has no FFI of its own; its k8s
 * jobs run nanoGPT's `train.py` (/root/reference/README.md:7-8, 55-66;
 * /root/reference/notebooks/colab_nanoGPT_companion.ipynb:70-79, 107-116), whose arithmetic is
 * stock PyTorch.  Each entry point below names the torch op (site-packages/torch/<file>:<line>,
 * torch 2.11) it replaces on that path -- that is what a maintainer's model.py/train.py would
 * bind instead (see INTEGRATION.md for the ctypes stubs).
 *
 * Conventions: plain pointers and sizes only, all pointers are DEVICE pointers unless a name
 * says host; the caller owns every allocation; every launcher enqueues on `stream` and returns
 * immediately; return value 0 = ok, anything else = failure with text in nsb_last_error().
 * Tensors are row-major.  "lowp" = bfloat16.  No entry point allocates or frees device memory.
 */
#ifndef NANOSANDBOX_B200_H
#define NANOSANDBOX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* nsb_stream_t; /* == cudaStream_t */

enum { NSB_F32 = 0, NSB_BF16 = 1 };

/* library / diagnostics ------------------------------------------------------------------- */
int nsb_version(void);
const char* nsb_last_error(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
unsigned long long nsb_launch_count(void);

/* K1: token + position embedding.  Replaces nn.Embedding x2 + add in GPT.forward
 * (torch/nn/functional.py:2437 F.embedding).  x[b,t,:] = wte[idx[b,t],:] + wpe[t,:] (bit-exact). */
int nsb_embedding_fwd(const int64_t* idx, const float* wte, const float* wpe, float* x,
                      int B, int T, int C, int V, nsb_stream_t stream);
/* K1': dwte[idx[b,t],:] += dx[b,t,:] (fp32 atomics, on top of the lm_head wgrad already there:
 * tied weight); dwpe[t,:] = (accumulate ? dwpe[t,:] : 0) + sum_b dx[b,t,:].
 * Replaces embedding_dense_backward (autograd of F.embedding). */
int nsb_embedding_bwd(const int64_t* idx, const float* dx, float* dwte, float* dwpe,
                      int B, int T, int C, int V, int accumulate_wpe, nsb_stream_t stream);

/* K2: LayerNorm over the last dim, eps inside the sqrt, bias optional (NULL).
 * Replaces F.layer_norm (torch/nn/functional.py:2914).  y is f32 or bf16 (y_dtype); mean/rstd
 * [M] are saved for the backward. */
int nsb_layernorm_fwd(const float* x, const float* w, const float* b, void* y, int y_dtype,
                      float* mean, float* rstd, int M, int C, float eps, nsb_stream_t stream);
/* K2': dx = (dres_in ? dres_in : 0) + LN'(dy); optional bf16 copy of dx (dx_lowp, may be NULL);
 * dw (+)= sum_rows dy*xhat, db (+)= sum_rows dy (db may be NULL).  Two-stage column reduction
 * through `scratch` (>= nsb_layernorm_bwd_scratch_bytes(M, C) bytes of device memory). */
size_t nsb_layernorm_bwd_scratch_bytes(int M, int C);
int nsb_layernorm_bwd(const void* dy, int dy_dtype, const float* x, const float* w,
                      const float* mean, const float* rstd, const float* dres_in, float* dx,
                      void* dx_lowp, float* dw, float* db, int accumulate, float* scratch,
                      size_t scratch_bytes, int M, int C, nsb_stream_t stream);

/* LayerNorm backward without caller scratch.  No counterpart in the reference: it exists for
 * hosts that can pass inputs and outputs but cannot allocate device memory themselves -- an
 * operator plugged into another runtime that only hands over its tensors, a caller inside a stream
 * capture, a deployment with a fixed memory plan.  Same arguments and results as
 * nsb_layernorm_bwd minus scratch/scratch_bytes: the partial sums go to a 1.16 MiB array built
 * into the library (one copy per GPU that loads it, held until the process exits), so it serves
 * rows of at most 1024 columns (any number of rows) and fails with a message beyond that.  It
 * calls no cudaMalloc/cudaFree.  The first call on each device looks up the array's device address;
 * make that call before starting any stream capture.  The address is that of the CURRENT device's
 * copy: `stream` must belong to the current device (checked; the call fails otherwise).  All calls on a device share the one array
 * (nothing in it needs clearing between calls), therefore two calls that may run on the GPU at the
 * same time -- different streams, no ordering between them -- race on it: keep them ordered. */
int nsb_layernorm_bwd_noalloc(const void* dy, int dy_dtype, const float* x, const float* w,
                              const float* mean, const float* rstd, const float* dres_in, float* dx,
                              void* dx_lowp, float* dw, float* db, int accumulate, int M, int C,
                              nsb_stream_t stream);

/* K9: softmax cross-entropy with ignore_index=-1, mean over non-ignored rows.
 * Replaces F.cross_entropy (torch/nn/functional.py:3421).  logits [M, ld] (ld >= V), dtype f32
 * or bf16.  Writes lse[M], row_loss[M] and stats[0]=mean loss, stats[1]=#valid rows. */
int nsb_cross_entropy_fwd(const void* logits, int dtype, const int64_t* targets, float* lse,
                          float* row_loss, float* stats, int M, int V, int ld,
                          nsb_stream_t stream);
/* K9': in place, logits <- (softmax(logits) - onehot(target)) * gscale / n_valid, 0 for ignored
 * rows; gscale is a device scalar (NULL = 1).  Columns V..ld-1 are set to 0. */
int nsb_cross_entropy_bwd(void* logits, int dtype, const int64_t* targets, const float* lse,
                          const float* stats, const float* gscale, int M, int V, int ld,
                          nsb_stream_t stream);

/* K9+K9': both passes in one kernel for bf16 logits whose row (ld * 2 bytes) fits in shared memory:
 * the row is read from HBM once.  Same lse / row_loss / stats as nsb_cross_entropy_fwd; logits <-
 * softmax(logits) - onehot(target), 0 for ignored rows and for columns V..ld-1, NOT scaled: the caller
 * multiplies by gscale / stats[1] downstream (the alpha of the two GEMMs that read it).  Fails with a
 * message for fp32 logits, rows that are not 16-byte aligned, or rows that do not fit. */
int nsb_cross_entropy_fwd_bwd(void* logits, int dtype, const int64_t* targets, float* lse,
                              float* row_loss, float* stats, int M, int V, int ld,
                              nsb_stream_t stream);

/* K12: global L2 norm of a flat fp32 gradient arena and the clip coefficient
 * min(1, max_norm / (norm + 1e-6)).  Replaces clip_grad_norm_ (torch/nn/utils/clip_grad.py:
 * 96-106, 165-174).  out[0] = total norm, out[1] = coefficient (applied inside nsb_adamw_step). */
size_t nsb_grad_norm_scratch_bytes(size_t n);
int nsb_grad_clip_coef(const float* g, size_t n, float max_norm, float* scratch,
                       size_t scratch_bytes, float* out, nsb_stream_t stream);

/* K13: ONE fused AdamW pass over the flat arenas (decoupled weight decay; arithmetic order of
 * torch/optim/adam.py:417-419, 457, 476, 531-547).  n is a multiple of 1024; chunk_decay[i] != 0
 * means elements [1024 i, 1024 i + 1024) belong to a weight-decayed tensor.  decay_factor =
 * 1 - lr*wd, step_size = lr / (1 - beta1^t), bc2_sqrt = sqrt(1 - beta2^t) are computed by the
 * caller in double.  gscale (device scalar, may be NULL) multiplies every gradient first (the
 * clip coefficient).  p_lowp (may be NULL) receives a bf16 copy of the updated parameters. */
int nsb_adamw_step(float* p, const float* g, float* m, float* v, void* p_lowp,
                   const uint8_t* chunk_decay, size_t n, float beta1, float beta2, float eps,
                   float decay_factor, float step_size, float bc2_sqrt, const float* gscale,
                   nsb_stream_t stream);

/* casts / scaling (weight shadow copies; DDP bucket pack = cast * 1/world, unpack) */
int nsb_cast_f32_to_bf16(const float* in, void* out, size_t n, float scale, nsb_stream_t stream);
int nsb_cast_bf16_to_f32(const void* in, float* out, size_t n, nsb_stream_t stream);
int nsb_scale_f32(float* x, size_t n, float scale, nsb_stream_t stream);
/* Stand-alone exact-erf GELU on bf16 activations (nn.GELU(), torch/nn/modules/activation.py:777-808):
 * a = gelu(h), and in place da <- da * gelu'(h).  The same math is also available fused into the
 * GEMM epilogues (NSB_EPI_GELU / NSB_EPI_DGELU); on B200 the split form is faster for K = 768. */
int nsb_gelu_fwd_bf16(const void* h, void* a, size_t n, nsb_stream_t stream);
int nsb_gelu_bwd_bf16(void* da, const void* h, size_t n, nsb_stream_t stream);
/* db[n] (+)= sum_m dy[m, n]  (bias gradients; autograd of F.linear's bias add) */
int nsb_colsum(const void* dy, int dtype, int ld, float* db, int accumulate, int M, int N,
               nsb_stream_t stream);

/* K3,5,6,7,8,10: C[M,N] = alpha * sum_k A[m,k] B[n,k] with a fused epilogue.
 * Replaces F.linear (torch/nn/functional.py:2329) forward and its two autograd GEMMs.
 *   a_kmajor = 1: A is stored [M, lda] (k contiguous);  0: A is stored [K, lda] (m contiguous)
 *   b_kmajor = 1: B is stored [N, ldb] (k contiguous);  0: B is stored [K, ldb] (n contiguous)
 *   in_dtype = NSB_BF16 -> tcgen05/TMEM/TMA kernel (fp32 accumulate); NSB_F32 -> exact-fp32
 *   SIMT kernel (the 1e-5 parity mode).  Operand strides must be multiples of 16 bytes for bf16. */
enum {
  NSB_EPI_STORE = 0,    /* C = acc (+bias[n])                       C: c_dtype            */
  NSB_EPI_GELU = 1,     /* C = acc (+bias); C2 = gelu(C)            exact erf GELU        */
  NSB_EPI_RESIDUAL = 2, /* C(f32) = aux(f32) + round_lowp(acc(+bias)) residual stream      */
  NSB_EPI_ACCUM = 3,    /* C(f32) = (accumulate ? C : 0) + acc      weight gradients      */
  NSB_EPI_DGELU = 4     /* C = acc * gelu'(aux)   aux: pre-activation, same dtype/ld as C */
};
typedef struct {
  int M, N, K;
  const void* A; int lda; int a_kmajor;
  const void* B; int ldb; int b_kmajor;
  int in_dtype;
  int epilogue;
  void* C; int ldc; int c_dtype;
  void* C2;             /* second output (NSB_EPI_GELU), same ld/dtype as C */
  const void* aux;      /* residual (f32) or pre-activation, ld = ldc       */
  const float* bias;    /* [N] or NULL                                      */
  const float* alpha;   /* device scalar or NULL (=1)                        */
  int accumulate;       /* NSB_EPI_ACCUM: beta = 1                           */
} nsb_gemm_args;
int nsb_gemm(const nsb_gemm_args* args, nsb_stream_t stream);

/* K4: causal self-attention, softmax(q k^T / sqrt(D)) v.  Replaces
 * F.scaled_dot_product_attention(is_causal=True) (torch/nn/functional.py:5948) and the
 * split/view/transpose/contiguous around it: qkv is c_attn's output [B, T, 3, H, D] read in
 * place, y is [B, T, H, D] (= [B, T, C]) ready for c_proj.  lse [B, H, T] is saved for backward. */
int nsb_attention_fwd(const void* qkv, void* y, float* lse, int dtype, int B, int T, int H,
                      int D, nsb_stream_t stream);
/* K4': dqkv [B, T, 3, H, D] from dy.  delta [B, H, T] is caller-provided scratch.  dq_acc is an
 * fp32 [B*T, H*D] accumulator for the tensor-core (tcgen05) path, bf16 with D = 64: the key-block
 * CTAs reduce their dQ contributions into it; it must be all zeros on entry and is zero again on
 * return.  NULL selects the two-pass kernels that need no accumulator. */
int nsb_attention_bwd(const void* qkv, const void* y, const void* dy, const float* lse,
                      void* dqkv, float* delta, float* dq_acc, int dtype, int B, int T, int H, int D,
                      nsb_stream_t stream);

/* ---- dropout (config.dropout > 0: nanoGPT's shakespeare_char recipe trains with 0.2) -----------------------
 * Masks are not stored: an element is kept when its word of Philox4x32-10(key = seed, counter = (element
 * index / 4, stream_id)) is >= floor(p * 2^32), so forward and backward regenerate the same mask from the same
 * (seed, stream_id).  Kept values are scaled by 1 / (1 - p). */

/* nn.Dropout (torch/nn/functional.py: dropout) on a flat tensor of n elements, optionally fused with the
 * residual add that follows it in Block.forward: y = (residual ? residual : 0) + dropout(x).  x / y: f32 or
 * bf16 (y must be f32 when a residual is given); in place (y == x) is allowed.  Used for transformer.drop,
 * attn.resid_dropout and mlp.dropout, and with the same (seed, stream_id) on the incoming gradient in backward. */
int nsb_dropout(const void* x, int x_dtype, const float* residual, void* y, int y_dtype, size_t n, float p,
                unsigned long long seed, unsigned long long stream_id, nsb_stream_t stream);

/* K4 with F.scaled_dot_product_attention's dropout_p: the mask index of probability (b, h, q, k) is
 * ((b H + h) T + q) T + k.  The softmax normaliser is that of the undropped probabilities, as in torch.
 * mask_scratch: B*H*T*T bytes of caller-owned device memory (contents need not survive the call), or NULL.
 * bf16 with D = 64 and a scratch buffer: one element-wise pass writes the keep bytes there and the mma.sync flash
 * kernels read them; otherwise (fp32, other head dims, no scratch) the exact SIMT kernels draw the bits themselves. */
int nsb_attention_fwd_dropout(const void* qkv, void* y, float* lse, int dtype, int B, int T, int H, int D,
                              float p, unsigned long long seed, unsigned long long stream_id,
                              void* mask_scratch, size_t mask_scratch_bytes, nsb_stream_t stream);
int nsb_attention_bwd_dropout(const void* qkv, const void* y, const void* dy, const float* lse, void* dqkv,
                              float* delta, int dtype, int B, int T, int H, int D, float p,
                              unsigned long long seed, unsigned long long stream_id, void* mask_scratch,
                              size_t mask_scratch_bytes, nsb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NANOSANDBOX_B200_H */
