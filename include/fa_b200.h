/* fa_b200.h -- C ABI of libflash_attention.so (B200 / sm_100a).
 *
 * This is the drop-in boundary.  The reference loads a plugin of the same file name,
 * build/libflash_attention.so, with ctypes (flash_attention_ipu/__init__.py:4-20) and
 * the plugin's static initialisers register the popart op
 * ("ai.graphcore","FlashAttentionQKVPacked",1) and its grad
 * (flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:516-517, :618-625, :675-676).
 * Here the same library instead exports plain C entry points that the Python side
 * binds with ctypes and wraps as a torch.library op.
 *
 * Conventions
 *  - every function returns 0 on success, a negative fa_b200_status otherwise;
 *    fa_b200_last_error() gives a thread-local message for the last failure.
 *  - the library never allocates or frees device memory (its own state on the device is two
 *    static arrays: the work counters of the persistent forward and the scratch of
 *    fa_b200_bwd_noalloc), never synchronises the device (fa_b200_bwd_sm_clock, a
 *    measurement hook, excepted) and only enqueues work on the stream it is given
 *    (cudaStream_t as void*).  The tcgen05 kernels are launched with programmatic stream
 *    serialization among themselves (FA_B200_PDL=0 in the environment: plain launches).
 *  - the caller makes the tensors' device current before the call.
 *  - tensors are 4-D (B, H, L, D) with D contiguous; strides are in ELEMENTS.
 *    The packed (3,G,L,D) layout of the reference API is B=1, H=G,
 *    q/k/v = base + {0,1,2}*G*L*D (zero copy).
 */
#ifndef FA_B200_H_
#define FA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  FA_B200_OK = 0,
  FA_B200_ERR_INVALID = -1,     /* bad argument (shape, stride, alignment, null pointer) */
  FA_B200_ERR_UNSUPPORTED = -2, /* dtype / head-dim combination with no kernel           */
  FA_B200_ERR_CUDA = -3,        /* a CUDA runtime / driver call failed                   */
  FA_B200_ERR_WORKSPACE = -4    /* workspace too small                                   */
} fa_b200_status;

/* element types; matches FLOAT16 / FLOAT of the reference op definition
 * (cpp/flash_attention_qkv_packed.cpp:609) plus bf16, which the IPU lacks. */
typedef enum { FA_B200_F16 = 0, FA_B200_BF16 = 1, FA_B200_F32 = 2 } fa_b200_dtype;

/* one (B,H,L,D) operand: device pointer + element strides (stride of D is 1). */
typedef struct {
  void *ptr;
  int64_t stride_b, stride_h, stride_l;
} fa_b200_tensor;

/* problem description shared by forward and backward. */
typedef struct {
  int32_t B, H;          /* G = B*H independent attention groups                          */
  int32_t Lq, Lk, D;     /* query rows, key/value rows, head dim                          */
  int32_t dtype;         /* fa_b200_dtype of q,k,v,o,do,dq,dk,dv                          */
  int32_t causal;        /* 1: key c visible to query r iff kv_pos0+c <= q_pos0+r         */
  float softmax_scale;   /* S = scale * q.k ; the packed reference API uses 1.0           */
  int32_t block_q_hint;  /* L/num_chunks_q  of the reference API: a tile-size HINT only   */
  int32_t block_kv_hint; /* L/num_chunks_kv (cpp/flash_attention_qkv_packed.cpp:194-195)  */
  int32_t q_pos0, kv_pos0; /* global sequence position of q row 0 / kv row 0 (ring steps) */
  int32_t H_kv;          /* K/V heads: 0 or H = multi-head; a divisor of H = grouped-query /  */
                         /* multi-query attention (query head h reads K/V head h/(H/H_kv));   */
                         /* k, v, dk, dv are then (B, H_kv, Lk, D).  The reference refuses    */
                         /* this case (_impl/scaled_dot_product_attention.py:33-37).          */
  const int32_t *kv_lens; /* NULL, or a DEVICE array of B*H ints: group g = b*H+h only has its first  */
                         /* kv_lens[g] (<= Lk) keys, the rest is padding (the key-padding form of    */
                         /* attn_mask, which the reference refuses: _impl/scaled_dot_product_        */
                         /* attention.py:21-24).  All heads of a batch element must agree.  KV tiles  */
                         /* made of padding are skipped; dK = dV = 0 there.  Tensor path only.        */
  float dropout_p;       /* dropout on the attention probabilities (the reference refuses dropout_p > 0: */
                         /* _impl/scaled_dot_product_attention.py:25-28): 0 = none; else each (query,    */
                         /* key) pair is dropped with probability round(256 p) / 256 and the survivors   */
                         /* are scaled by its complement's inverse.  Tensor path only; kv_pos0 % 4 == 0. */
  uint64_t dropout_seed; /* the mask is a pure function of (dropout_seed, b*H+h, q_pos0 + row,           */
                         /* kv_pos0 + key) -- csrc/sm100_ptx.cuh: drop_hash -- so the backward must be   */
                         /* given the forward's seed; restated by oracle/attention_oracle.py.            */
  const void *bias;      /* NULL, or a DEVICE tensor added to the scaled scores before the softmax (the   */
                         /* additive form of attn_mask, which the reference refuses:                      */
                         /* _impl/scaled_dot_product_attention.py:21-24): S' = scale*q.k + bias[b][h][q][k], */
                         /* elements of `dtype`, -inf = masked.  Element strides below; 0 broadcasts the   */
                         /* batch / head / query dim; the key dim has unit stride.  Tensor path only; Lk % 8 == 0, */
                         /* 16-byte aligned pointer and row stride; no gradient is produced for the bias.   */
  int64_t bias_stride_b, bias_stride_h, bias_stride_l;
} fa_b200_problem;

/* Forward: replaces flashAttentionQKVPackedWithStash
 * (cpp/flash_attention_qkv_packed.cpp:175-334; declared .hpp:15-24).
 * o: (B,H,Lq,D) in `dtype`; lse: fp32 contiguous [B*H][Lq], natural-log logsumexp of
 * the scaled, masked scores (the `logSumExp` stash, .hpp:10-13).  Rows with no visible
 * key get o = 0, lse = -inf. */
int fa_b200_fwd(const fa_b200_problem *p, const fa_b200_tensor *q, const fa_b200_tensor *k,
                const fa_b200_tensor *v, const fa_b200_tensor *o, float *lse, void *stream);

/* Backward: replaces flashAttentionQKVPackedGrad
 * (cpp/flash_attention_qkv_packed.cpp:348-512; declared .hpp:31-38), except that O and
 * LSE saved by the forward are consumed instead of re-running the forward (:357-359).
 * workspace: device scratch of at least fa_b200_bwd_workspace_bytes(p) bytes,
 * 256-byte aligned; contents need not be initialised. */
int fa_b200_bwd(const fa_b200_problem *p, const fa_b200_tensor *q, const fa_b200_tensor *k,
                const fa_b200_tensor *v, const fa_b200_tensor *o, const fa_b200_tensor *dout,
                const float *lse, const fa_b200_tensor *dq, const fa_b200_tensor *dk,
                const fa_b200_tensor *dv, void *workspace, size_t workspace_bytes,
                void *stream);

size_t fa_b200_bwd_workspace_bytes(const fa_b200_problem *p);

/* Backward of one block of a larger (sequence-sharded) problem, dQ accumulated in place: as
 * fa_b200_bwd, except that `dq_acc` is the caller's fp32 (B,H,Lq,D) accumulator (element strides,
 * unit stride on D) and every partial dQ tile is reduce-added into it WITHOUT softmax_scale
 * (dq_acc += dS K; multiply by softmax_scale once, after the last block).  It is neither zeroed
 * nor converted, so the ring's hops need no separate add pass and dQ is rounded to 16 bits once.
 * Cross-device form of the `dQ +=` inside the reference's loop
 * (cpp/flash_attention_qkv_packed.cpp:470).  Tensor path only (16-bit dtype, D in {64, 128}). */
int fa_b200_bwd_dq_accum(const fa_b200_problem *p, const fa_b200_tensor *q, const fa_b200_tensor *k,
                         const fa_b200_tensor *v, const fa_b200_tensor *o, const fa_b200_tensor *dout,
                         const float *lse, const fa_b200_tensor *dq_acc, const fa_b200_tensor *dk,
                         const fa_b200_tensor *dv, void *workspace, size_t workspace_bytes, void *stream);

/* Backward without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same computation and results as fa_b200_bwd, but the
 * workspace is a static __device__ array inside the library, FA_B200_NOALLOC_SCRATCH_BYTES (32 MiB)
 * on every GPU that loads the library, held until the process exits.  Problems with
 * fa_b200_bwd_workspace_bytes(p) above that limit are refused with FA_B200_ERR_WORKSPACE (tensor
 * path: about B*H*Lq*(4*D + 24) bytes, e.g. B*H = 8, Lq = 4096, D = 128 needs 17 MiB).  The call
 * itself makes no cudaMalloc / cudaFree and uses no allocating library routine.  ALL calls on a
 * device share the one array, which keeps the previous call's contents: what has to start at zero
 * (dQ accumulator, work counter, semaphores) is cleared on `stream` at the start of every call, and
 * calls that overlap on the GPU (two streams, or two host threads on one device) RACE -- order them. */
#define FA_B200_NOALLOC_SCRATCH_BYTES (32u << 20)
int fa_b200_bwd_noalloc(const fa_b200_problem *p, const fa_b200_tensor *q, const fa_b200_tensor *k,
                        const fa_b200_tensor *v, const fa_b200_tensor *o, const fa_b200_tensor *dout,
                        const float *lse, const fa_b200_tensor *dq, const fa_b200_tensor *dk,
                        const fa_b200_tensor *dv, void *stream);

/* Measurement hook (no reference counterpart; the reference's C++ benchmark reads
 * poplar::cycleCount the same way, tests/cpp/test_flash_attention.cpp:236-239): SM cycles and
 * nanoseconds that CTA 0 of the last tensor-family backward kernel launched with `workspace`
 * spent on its SM -- cycles / ns = the effective SM clock under load.  Synchronises on the copy;
 * call after the stream the backward ran on has been synchronised. */
int fa_b200_bwd_sm_clock(const fa_b200_problem *p, const void *workspace, uint64_t *cycles, uint64_t *ns);

/* LSE merge of two partial results over disjoint key sets (ring-attention hop):
 *   lse = logaddexp(lse_a, lse_b);  o = o_a*exp(lse_a-lse) + o_b*exp(lse_b-lse)
 * written in place into (o_a, lse_a).  rows = B*H*Lq, o_* contiguous [rows][D].
 * No reference counterpart (the reference is single-device). */
int fa_b200_merge(void *o_a, float *lse_a, const void *o_b, const float *lse_b, int64_t rows,
                  int32_t D, int32_t dtype, void *stream);

/* Ring accumulator form of the merge: (o_acc, lse_acc) are fp32 and persist across all hops of
 * a ring (no per-hop rounding); the hop's partial (o_b in dtype_b, lse_b) has n_outer groups of
 * n_inner rows and partial row (g, r) is folded into accumulator row
 * g*acc_outer_stride + acc_row0 + r.  lse_acc = -inf marks an empty accumulator row. */
int fa_b200_merge_acc(float *o_acc, float *lse_acc, const void *o_b, const float *lse_b,
                      int64_t n_outer, int64_t n_inner, int64_t acc_outer_stride, int64_t acc_row0,
                      int32_t D, int32_t dtype_b, void *stream);

/* Deterministic gradients.  dQ tiles are accumulated across KV tiles by fp32 reduce-adds; by default
 * in arrival order (run-to-run differences ~1e-7 relative).  With on != 0 every later tensor-family
 * backward adds the partials of a Q tile in a fixed (descending KV tile) order behind per-tile
 * semaphores: bit-identical results from run to run, at a throughput cost.  Process-wide; returns
 * the previous setting.  (dK, dV and the forward are deterministic either way.) */
int fa_b200_set_deterministic(int on);

/* The dropout mask of a problem, evaluated on the HOST with the very functions the kernels compile
 * (csrc/sm100_ptx.cuh: drop_hash): keep[c] = 1 iff key k_first + c of query row q_row of group
 * (b*H + h) survives; *scale (optional) = the factor applied to the survivors.  No reference
 * counterpart (the reference refuses dropout); for debugging and to pin the CPU oracle's restatement
 * of the hash without a GPU (tests/test_oracle.py).  No device work. */
int fa_b200_dropout_keep_host(const fa_b200_problem *p, int32_t group, int32_t q_row, int32_t k_first, int32_t n,
                              uint8_t *keep, float *scale);

/* Which kernel family a problem dispatches to: 1 = tcgen05/TMA sm_100a kernels,
 * 0 = generic SIMT kernels (fp32 inputs / head dims the tensor path does not cover),
 * negative = unsupported. */
int fa_b200_kernel_family(const fa_b200_problem *p);

/* Number of kernel launches the library has enqueued since load (all threads). */
int64_t fa_b200_launch_count(void);

/* Tensor-map (TMA descriptor) cache of the shim: descriptors are keyed by (pointer, dtype, shape,
 * strides, box) and reused across calls; counters since load (all threads). */
void fa_b200_tmap_cache_stats(int64_t *hits, int64_t *misses);

const char *fa_b200_last_error(void);
int fa_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* FA_B200_H_ */
