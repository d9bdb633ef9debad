/* C ABI of libhstd_b200.so — the sm_100a kernel library of
 * huggingface_sagemaker_tensorflow_distributed_b200.
 *
 * Every entry point takes raw device pointers plus a `stream` (a cudaStream_t passed as
 * void*), launches asynchronously on that stream and returns 0 on success or a negative
 * error code; `hstd_last_error()` returns a human-readable message for the calling thread.
 * Unless a comment says otherwise, tensors are row-major bf16 and no entry point allocates
 * device memory (workspaces are passed by the caller).
 *
 * The reference (philschmid/huggingface_sagemaker_tensorflow_distributed) has no native
 * code; each entry point cites the reference line whose *implied* library call it replaces
 * (SURVEY.md §2.4, K1..K13).
 */
#ifndef HSTD_B200_H_
#define HSTD_B200_H_

#include <stdint.h>

#if defined(HSTD_BUILDING_LIB)
#define HSTD_API __attribute__((visibility("default")))
#else
#define HSTD_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

const char* hstd_last_error(void);
int hstd_version(void);

/* ------------------------------------------------------------------ GEMM (K2,K4,K5,K6 + attention GEMMs)
 * tcgen05/TMEM GEMM fed by TMA.  Replaces the cuBLAS calls behind model.fit()
 * (reference scripts/train.py:145).  For every batch index z = (z0, z1):
 *     acc[m,n] = sum_k A_z(m,k) * B_z(n,k)
 *     C_z[m,n] = epilogue(alpha * acc[m,n] + bias[n])
 * Operand X (A or B) in K-major form is read at  X + z0*bs0 + z1*bs1 + row*ld + k   (row = m or n),
 * in MN-major form ("transposed") at            X + z0*bs0 + z1*bs1 + k*ld   + row.
 * All strides are in elements; ld, bs0, bs1 must be multiples of 8 and pointers 16-byte aligned.
 */
enum {
  HSTD_EPI_NONE = 0,     /* C = v                                                        */
  HSTD_EPI_GELU = 1,     /* aux (if non-null) = v ; C = gelu_erf(v)                       */
  HSTD_EPI_ADD = 2,      /* C = v + aux        (aux: bf16 [M,N] read with ldaux)          */
  HSTD_EPI_GELU_BWD = 3, /* C = v * gelu_erf'(aux)  (aux: pre-activation, ldaux)          */
};

typedef struct hstd_gemm_args {
  const void* A;
  const void* B;
  void* C;
  const void* bias; /* bf16 [N] or NULL */
  void* aux;        /* see epilogue */
  int64_t lda, a_bs0, a_bs1;
  int64_t ldb, b_bs0, b_bs1;
  int64_t ldc, c_bs0, c_bs1;
  int64_t ldaux; /* aux uses the same batch strides as C */
  int32_t M, N, K;
  int32_t nb0, nb1; /* batch extents (>=1) */
  int32_t a_mn_major, b_mn_major;
  int32_t epilogue;
  float alpha;
  int32_t block_n; /* 0 = auto, else 64/128/192/256 */
  int32_t max_ctas; /* 0 = one persistent CTA per SM */
  int32_t cluster_m; /* 0 = auto (2-CTA clusters multicasting B when M > 128), 1 = no clusters */
} hstd_gemm_args;

int hstd_gemm_bf16(const hstd_gemm_args* args, void* stream);

/* Split-K variant for weight gradients (K = tokens is long, M x N is small):
 * dW[M,N] = sum_k A(m,k) B(n,k) with `splits` partial sums written as fp32 to
 * `workspace` ([splits, M, N] floats, caller-allocated) and reduced to bf16 C by a second
 * kernel.  Only alpha and EPI_NONE are honoured.  No batch dimensions. */
int hstd_gemm_bf16_splitk(const hstd_gemm_args* args, int splits, void* workspace, int64_t workspace_bytes,
                          void* stream);
int64_t hstd_gemm_splitk_workspace_bytes(int M, int N, int splits);

/* ------------------------------------------------------------------ per-step device state
 * Lives in device memory so a captured CUDA graph can be replayed every step: the prologue
 * kernel advances the step counter, evaluates the LR schedule and the Keras-Adam bias
 * correction (reference scripts/train.py:112-113: Adam(lr * size), Keras eps placement) and
 * re-keys the dropout RNG.  `metrics` is {loss_sum, correct, count} of the last step and
 * `metrics_acc` their running sums since the last hstd_metrics_reset. */
typedef struct hstd_step_state {
  int64_t step;  /* optimizer steps taken, including the current one after the prologue */
  uint64_t seed; /* dropout RNG key */
  float lr;      /* schedule value for this step */
  float lr_t;    /* lr * sqrt(1 - b2^t) / (1 - b1^t) */
  float decay;   /* 1 - lr * weight_decay */
  float beta1, beta2, eps;
  float metrics[4];
  double metrics_acc[4];
} hstd_step_state;

enum { HSTD_SCHED_CONSTANT = 0, HSTD_SCHED_LINEAR = 1 };

int hstd_step_state_init(hstd_step_state* dev_state, uint64_t seed, int64_t step, void* stream);
int hstd_step_prologue(hstd_step_state* dev_state, float base_lr, int schedule, int64_t total_steps,
                       int64_t warmup_steps, float beta1, float beta2, float eps, float weight_decay, void* stream);
int hstd_metrics_reset(hstd_step_state* dev_state, void* stream);

/* Keras-Adam(W) on a flat parameter arena (K11; reference scripts/train.py:113).  grads are
 * bf16, master/m/v fp32, the refreshed bf16 weights are written to `w_bf16`.  `grad_scale`
 * multiplies the gradient first (1/world_size when the sum was formed elsewhere). */
int hstd_adamw_flat(const hstd_step_state* st, const void* grads_bf16, float* master, float* m, float* v,
                    void* w_bf16, int64_t n, float grad_scale, void* stream);

/* ------------------------------------------------------------------ LayerNorm family (K1, K4, K6)
 * Row-wise kernels, hidden size d must be a multiple of 256 and <= 1024.
 *   forward :  pre = dropout_in(a) + resid ;  y = dropout_out(LN(pre) * gamma + beta)
 *   embedding forward: a is gathered as word[ids] + pos[pos_ids] (+ type0)
 * `pre`, `mean`, `rstd` are saved for backward; dropout masks are regenerated from
 * (state->seed, state->step, site) and never stored.  state==NULL or p==0 disables dropout. */
typedef struct hstd_ln_args {
  const void* a;      /* bf16 [M,d] (NULL for the embedding variant) */
  const void* resid;  /* bf16 [M,d] or NULL */
  const void* gamma;  /* bf16 [d] */
  const void* beta;   /* bf16 [d] */
  void* pre;          /* bf16 [M,d] out (may alias a) */
  void* y;            /* bf16 [M,d] out */
  float* mean;        /* [M] out */
  float* rstd;        /* [M] out */
  const int32_t* ids; /* embedding variant: [M] token ids */
  const int32_t* pos_ids;
  const void* word;   /* bf16 [V,d] */
  const void* pos;    /* bf16 [P,d] */
  const void* type0;  /* bf16 [d] or NULL */
  const hstd_step_state* state;
  int32_t M, d;
  float eps, p_in, p_out;
  uint32_t site_in, site_out;
} hstd_ln_args;
int hstd_ln_fwd(const hstd_ln_args* args, void* stream);

/* backward: given dy, recompute xhat from (pre, mean, rstd):
 *   dpre = LN'(dropout_out'(dy)) ; da = dropout_in'(dpre)  (da may be NULL when p_in == 0)
 *   dgamma, dbeta, dbias (= column sums of da, or of dpre when da==NULL) reduced over rows.
 * One kernel; the column sums go through `partials`, an fp32 accumulator of hstd_ln_bwd_partials_floats(d)
 * floats that must be ZERO before the first call and is left zeroed by every call (calls that share it must
 * not overlap on the GPU). */
typedef struct hstd_ln_bwd_args {
  const void* dy; const void* pre; const float* mean; const float* rstd; const void* gamma;
  void* dpre; void* da;
  void* dgamma; void* dbeta; void* dbias; /* bf16 [d] each; dbias may be NULL */
  float* partials;
  const hstd_step_state* state;
  int32_t M, d;
  float p_in, p_out;
  uint32_t site_in, site_out;
} hstd_ln_bwd_args;
int hstd_ln_bwd(const hstd_ln_bwd_args* args, void* stream);
int64_t hstd_ln_bwd_partials_floats(int d);

/* Column sums of a bf16 [M,N] matrix into bf16 [N] (bias gradients).  `partials`: scratch
 * of hstd_colsum_partials_floats(N) floats.  `x` must be 16-byte aligned and `ld` (elements)
 * a multiple of 8: rows are read with 16-byte vector loads; anything else is refused. */
int hstd_colsum(const void* x, int64_t ld, int M, int N, void* out_bf16, float* partials, void* stream);
int64_t hstd_colsum_partials_floats(int N);

/* Column sums for callers that can neither pass a workspace nor tolerate a device allocation
 * (no counterpart in the reference, which has no native code at all): a plain-C user with no
 * device allocator to carve scratch from, or a process whose GPU memory is wholly owned by
 * another pool, where a cudaMalloc from inside a library could fail or stall every stream with
 * its implicit synchronisation.  Same result as hstd_colsum, but the partial sums are kept in a
 * 512 KiB array that is built into the library and resident on every GPU that loads it, so:
 *   - N is limited to hstd_colsum_noalloc_max_n() (4096); larger N is refused with an error
 *     and should go through hstd_colsum;
 *   - the call makes no cudaMalloc/cudaFree and uses no allocating library routine;
 *   - `x` must be 16-byte aligned and `ld` a multiple of 8 elements, as for hstd_colsum
 *     (refused with an error otherwise, never a fault);
 *   - the first call on each device looks up the array's device address (with that device
 *     current); make that first call before starting any stream capture.  The address is
 *     cached for the life of the process: it does not survive a cudaDeviceReset(), after
 *     which this entry point must not be used again in that process;
 *   - all calls on a device share the one array: two calls whose kernels can run at the same
 *     time (different streams, no dependency between them) race on it — order them.  The
 *     array keeps the previous call's partial sums; nothing needs clearing, every cell that is
 *     read is first rewritten by the same call. */
int hstd_colsum_noalloc(const void* x, int64_t ld, int M, int N, void* out_bf16, void* stream);
int hstd_colsum_noalloc_max_n(void);

/* ------------------------------------------------------------------ attention softmax (K3)
 * scores: bf16 [B,H,Sq,Sk] (already scaled), key mask int32 [B,Sk] (1 = attend).
 * forward : P = softmax(scores + mask) written in place; Pd = dropout(P) written to `pd`
 *           (pd may be NULL when p == 0: use P).
 * backward: dpd (= dO·Vᵀ) is overwritten with dS = P ∘ (dP − rowsum(dP ∘ P)), dP = dropout'(dpd). */
int hstd_softmax_fwd(void* scores_p, void* pd, const int32_t* key_mask, int B, int H, int Sq, int Sk, float p_drop,
                     uint32_t site, const hstd_step_state* st, void* stream);
int hstd_softmax_bwd(const void* p, void* dpd_ds, int B, int H, int Sq, int Sk, float p_drop, uint32_t site,
                     const hstd_step_state* st, void* stream);

/* ------------------------------------------------------------------ fused attention (K3)
 * Flash-style attention on tcgen05: ctx = dropout(softmax(Q Kᵀ * scale + key mask)) V per
 * (batch, head), reading Q/K/V in place from the fused projection output qkv [B*S, 3*H*64]
 * (Q | K | V column blocks) and writing ctx [B*S, H*64].  head_dim is 64, S in {128..512}.
 * The [B,H,S,S] probabilities never touch memory; forward saves only lse [B,H,S] (fp32).
 * Backward consumes qkv, ctx, dctx and lse and writes dqkv [B*S, 3*H*64];
 * `workspace` = hstd_attn_bwd_workspace_bytes(B,H,S) bytes of caller scratch.
 * Dropout uses the same Philox stream as hstd_softmax_fwd (same site/seed/step => same mask). */
int hstd_attn_fwd(const void* qkv, void* ctx, float* lse, const int32_t* key_mask, int B, int H, int S, float scale,
                  float p_drop, uint32_t site, const hstd_step_state* st, void* stream);
int hstd_attn_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse, const int32_t* key_mask,
                  void* dqkv, void* workspace, int B, int H, int S, float scale, float p_drop, uint32_t site,
                  const hstd_step_state* st, void* stream);
int64_t hstd_attn_bwd_workspace_bytes(int B, int H, int S);
/* Debug aids (tools/attn_trace.py, tools/gemm_trace.py): a device buffer of >= 64 / >= 512 int64 that CTA 0 of
 * the following attention / GEMM launches fills with clock64() stamps of its phases; NULL switches it off. */
void hstd_attn_set_trace(void* device_buffer);
void hstd_gemm_set_trace(void* device_buffer);

/* ------------------------------------------------------------------ embeddings (K1)
 * position ids: mode 0 = arange(S); mode 1 = RoBERTa (cumsum(ids != pad) * (ids != pad) + pad). */
int hstd_position_ids(const int32_t* ids, int32_t* pos_ids, int B, int S, int mode, int pad_id, void* stream);
/* scatter-add de [M,d] into word_grad[ids] and pos_grad[pos_ids] (bf16 tables, zeroed by the caller). */
int hstd_embed_bwd(const void* de, const int32_t* ids, const int32_t* pos_ids, void* word_grad, void* pos_grad,
                   int M, int d, void* stream);

/* ------------------------------------------------------------------ classification head + loss (K7, K8)
 * x: bf16 [B*S, d] final hidden states; pooled = row b*S of x.  kind: 0 DistilBERT
 * (Linear-ReLU-drop-Linear), 1 BERT (Linear-tanh-drop-Linear), 2 RoBERTa
 * (drop-Linear-tanh-drop-Linear).  labels int32 [B], -1 = ignore.  Loss is the mean over valid
 * rows of sparse softmax cross-entropy from logits (reference scripts/train.py:118-119).
 * forward writes logits [B,C] fp32, dlogits (d loss / d logits, already divided by the valid
 * count), h [B,d] fp32 (post-activation; with `train` != 0 already multiplied by the dropout
 * mask), sets state->metrics = {loss_sum, correct, count} and adds them to state->metrics_acc.
 * `train` != 0 enables dropout and, when dz is given, leaves mask * act'(h) in dz [B,d] fp32.
 * backward (train only; must follow a forward with train != 0 on the same h / dz buffers, which
 * carry the dropout mask so it is not regenerated) writes dW1,db1,dW2,db2 (bf16), dz, and the CLS
 * rows of dx (bf16 [B*S,d], other rows are zeroed here).  d <= 1024. */
typedef struct hstd_head_args {
  const void* x; const void* w1; const void* b1; const void* w2; const void* b2;
  const int32_t* labels;
  float* logits; float* dlogits; float* h; float* dz; /* fp32 [B,C], [B,C], [B,d], [B,d] */
  void* dw1; void* db1; void* dw2; void* db2; void* dx;
  hstd_step_state* state;
  int32_t B, S, d, C, kind, train;
  float p_drop;
  uint32_t site0, site1;
} hstd_head_args;
int hstd_head_fwd(const hstd_head_args* args, void* stream);
int hstd_head_bwd(const hstd_head_args* args, void* stream);

/* ------------------------------------------------------------------ fused gradient all-reduce (K9, K10, K11)
 * One launch per gradient bucket [start, start+numel) of the flat arenas.  Replaces
 * hvd.DistributedOptimizer's ncclAllReduce + 1/size postscale and Keras' per-variable Adam
 * apply (reference scripts/train.py:113-114) with one NVLink peer-memory kernel:
 * reduce (peer loads or in-switch multimem.ld_reduce) -> x 1/world -> bf16->fp32 -> Keras-AdamW on
 * the owned fp32 master/m/v -> bf16 weights stored to every rank (peer stores / multimem.st).
 * g_peers / w_peers / flag_peers: the same arena as mapped from every rank (index = rank;
 * entry [rank] is the local pointer).  flag area: hstd_allreduce_flag_words(num_buckets)
 * zero-initialised uint32 per rank; epoch: hstd_allreduce_max_blocks() zeroed uint32 per bucket
 * (local); error: one zeroed uint32 (local) that becomes non-zero if a peer never arrived.
 * master/m/v point at this bucket's *owned* range: numel/world elements (rank-th slice) for
 * TWO_SHOT and NVLS, numel elements (replicated) for ONE_SHOT and LOCAL.
 * update == 0 turns the kernel into a plain averaging all-reduce of the gradient arena
 * (ONE_SHOT writes its result to the weight arena instead: it cannot be in place). */
enum { HSTD_AR_LOCAL = 0, HSTD_AR_TWO_SHOT = 1, HSTD_AR_ONE_SHOT = 2, HSTD_AR_NVLS = 3 };
typedef struct hstd_allreduce_args {
  void* g_peers[8];
  void* w_peers[8];
  void* flag_peers[8];
  void* g_mc; /* multicast mapping of the gradient / weight arenas (NVLS only) */
  void* w_mc;
  void* epoch;
  void* error;
  float* master; float* m; float* v;
  const hstd_step_state* state;
  int64_t start, numel; /* elements; numel % (8*world) == 0 */
  int32_t world, rank, bucket_index, mode, update, blocks;
  /* row-sparse buckets (see below); all zero for an ordinary bucket */
  void* row_stamp;     /* uint32 per bucket row: == state->step when the row is live in this step */
  void* ids_peers[8];  /* hstd_sparse_rows_mark: this step's int32 token ids of every rank */
  int32_t row_elems;   /* elements per row (the hidden size) */
  int32_t rows;        /* HSTD_AR_ROWS_* filter of this launch */
  int32_t n_ids;       /* ids per rank */
  int32_t dense_rows;  /* leading rows of the bucket that are live in every step */
  int32_t vocab_rows;  /* rows addressable by an id, following the dense ones */
  int32_t tail;        /* non-zero: nothing else runs on the GPU any more (last buckets of a step): use big blocks */
  int64_t row_elem0;   /* bucket-relative element offset of `start` when a launch covers part of a row-sparse bucket */
  int32_t phase;       /* sharded optimizer form: 0 reduce + update in one launch; 1 reduce only (the averaged gradient
                        * of this rank's slice is written back, bf16, into its own gradient arena); 2 update only
                        * (AdamW from those averages + weight broadcast; no barrier before it) */
  int32_t end_barrier; /* phase 2: finish with the cross-rank barrier (the last launch of a step must) */
} hstd_allreduce_args;
int hstd_fused_allreduce(const hstd_allreduce_args* args, void* stream);

/* Row-sparse form for the embedding bucket (Horovod ships IndexedSlices gradients by allgather unless
 * sparse_as_dense is set, which the reference does not do — scripts/train.py:114; Keras' Adam then
 * still decays m / v and steps EVERY row).  The bucket is a stack of rows of `row_elems` elements:
 * `dense_rows` always-live rows (LayerNorm, position, token-type tables) followed by `vocab_rows`
 * word-embedding rows.  Per step:
 *   1. hstd_sparse_rows_mark, as soon as the batch is on the device: flag barrier, every rank reads
 *      every rank's ids (peer loads), stamps the rows somebody will produce a gradient for with the
 *      step number and clears exactly those rows of its own gradient arena.
 *   2. hstd_fused_allreduce(rows = HSTD_AR_ROWS_IDLE), any time after 1: rows nobody touches have a
 *      zero gradient everywhere, so the owner applies the Adam step (m, v decay; w moves by its
 *      history) and broadcasts the new bf16 row with no reduction and no barrier — off the critical
 *      path, under the forward pass (no rank reads an idle row during the step).
 *   3. hstd_fused_allreduce(rows = HSTD_AR_ROWS_LIVE) after backward: reduce + AdamW + broadcast of
 *      the live rows only.  Together 2 and 3 equal the dense update exactly. */
enum { HSTD_AR_ROWS_ALL = 0, HSTD_AR_ROWS_LIVE = 1, HSTD_AR_ROWS_IDLE = 2,
       HSTD_AR_ROWS_BOTH = 3 /* phase 2 only: live rows take the reduced gradient, idle rows zero */ };
int hstd_sparse_rows_mark(const hstd_allreduce_args* args, void* stream);

/* Single-GPU emulation of a `world`-rank launch for tests and profiling on a box with fewer GPUs
 * than ranks: one cooperative grid runs every rank's blocks (the same per-block code as the real
 * kernels, flag barriers included) over `world` arenas that all live on the current device.
 * `per_rank[r]` is what rank r would pass to the real entry point (two-shot / one-shot only). */
int hstd_fused_allreduce_emulated(const hstd_allreduce_args* per_rank, int world, void* stream);
int hstd_sparse_rows_mark_emulated(const hstd_allreduce_args* per_rank, int world, void* stream);
int hstd_allreduce_max_blocks(void);
int64_t hstd_allreduce_flag_words(int num_buckets);

/* CUDA IPC plumbing for the symmetric arenas (the C++ side of what MPI + NCCL bootstrap do for
 * the reference, launch.py:22 / scripts/train.py:24): allocate a zeroed device arena and export
 * its 64-byte handle; open a peer's handle; close / free. */
int hstd_ipc_alloc(int64_t bytes, void** ptr, void* handle64);
int hstd_ipc_open(const void* handle64, void** ptr);
int hstd_ipc_close(void* ptr);
int hstd_ipc_free(void* ptr);

/* ------------------------------------------------------------------ host runtime: tokenizer (N6)
 * Native WordPiece tokenizer: the C++ side of AutoTokenizer.from_pretrained(...)(text,
 * truncation=True, padding=...) (reference scripts/train.py:69,75-77,89-92).  `vocab_path` is a
 * BERT vocab.txt (one token per line; [PAD]/[UNK]/[CLS]/[SEP] are looked up by name).
 * hstd_tok_encode_batch writes dense row-major int32 [n, max_len] ids and masks (rows truncated
 * to max_len, padded with [PAD]); lengths[r] (optional) is the untruncated token count with the
 * specials.  num_threads <= 0 picks a default.  None of these touch the GPU. */
typedef struct hstd_tokenizer hstd_tokenizer;
hstd_tokenizer* hstd_tok_create(const char* vocab_path, int do_lower_case, int strip_accents);
void hstd_tok_destroy(hstd_tokenizer* tok);
int hstd_tok_vocab_size(const hstd_tokenizer* tok);
int hstd_tok_token_to_id(const hstd_tokenizer* tok, const char* token); /* -1 if absent */
const char* hstd_tok_id_to_token(const hstd_tokenizer* tok, int id);    /* NULL if out of range */
int hstd_tok_special_ids(const hstd_tokenizer* tok, int32_t* out4);     /* pad, unk, cls, sep */
int hstd_tok_encode_batch(const hstd_tokenizer* tok, const char* const* texts, int n, int max_len,
                          int add_special_tokens, int32_t* ids, int32_t* mask, int32_t* lengths, int num_threads);

/* Byte-level BPE (RoBERTa / GPT-2 vocab.json + merges.txt): GPT-2 pre-tokenisation, byte -> printable
 * code point mapping, lowest-rank-first merges, <s> ... </s> framing.  Same output contract as
 * hstd_tok_encode_batch.  special ids: pad, unk, bos, eos. */
typedef struct hstd_bpe hstd_bpe;
hstd_bpe* hstd_bpe_create(const char* vocab_json_path, const char* merges_path);
void hstd_bpe_destroy(hstd_bpe* tok);
int hstd_bpe_vocab_size(const hstd_bpe* tok);
int hstd_bpe_token_to_id(const hstd_bpe* tok, const char* token);
const char* hstd_bpe_id_to_token(const hstd_bpe* tok, int id);
int hstd_bpe_special_ids(const hstd_bpe* tok, int32_t* out4);
int hstd_bpe_encode_batch(const hstd_bpe* tok, const char* const* texts, int n, int max_len,
                          int add_special_tokens, int32_t* ids, int32_t* mask, int32_t* lengths, int num_threads);

/* ------------------------------------------------------------------ host runtime: fusion-bucket manager (N1)
 * What Horovod's C++ core does with its tensor-fusion buffer (reached from reference
 * scripts/train.py:114): decide which gradients travel together and notice when a group is
 * complete.  Parameters are given in arena (= backward-readiness) order by their element offsets;
 * a bucket is a contiguous range cut at parameter boundaries once it would exceed bucket_elems.
 *   hstd_bucket_plan: fills starts/ends (capacity max_buckets) and param_bucket[n]; returns the
 *     number of buckets (negative on error).
 *   tracker: mark parameters whose gradients now exist; buckets are released strictly in arena
 *     order, a bucket only after all its non-exempt parameters are marked. */
int hstd_bucket_plan(const int64_t* offsets, const int64_t* ends, int n, int64_t bucket_elems, int64_t* starts,
                     int64_t* bucket_ends, int32_t* param_bucket, int max_buckets);
typedef struct hstd_bucket_tracker hstd_bucket_tracker;
hstd_bucket_tracker* hstd_bucket_tracker_create(const int32_t* param_bucket, const uint8_t* exempt, int n, int num_buckets);
void hstd_bucket_tracker_destroy(hstd_bucket_tracker* t);
void hstd_bucket_tracker_reset(hstd_bucket_tracker* t);
/* returns how many bucket indices were written to ready_out (capacity = num_buckets) */
int hstd_bucket_tracker_mark(hstd_bucket_tracker* t, const int32_t* params, int k, int32_t* ready_out);
int hstd_bucket_tracker_done(const hstd_bucket_tracker* t); /* number of buckets released so far */

/* ------------------------------------------------------------------ host runtime: prefetching batch loader (C11)
 * tf.data's from_tensor_slices(...).batch(bs) runs in TF's C++ runtime (reference
 * scripts/train.py:84-86,98-100).  Equivalent here: a worker thread gathers the rows
 * indices[b*bs .. (b+1)*bs) of the int32 [N,S] id / mask tables and the [N] labels into a ring
 * of caller-owned (pinned) slots ahead of the consumer.  A short final batch is filled with
 * pad_id / mask 0 (first position 1) / ignore_label.  next() blocks until the next batch is
 * staged and returns its slot (or -1 after the last batch); release() hands the slot back. */
typedef struct hstd_loader hstd_loader;
hstd_loader* hstd_loader_create(const int32_t* ids, const int32_t* mask, const int32_t* labels, int64_t n_rows,
                                int seq_len, const int64_t* indices, int64_t n_indices, int batch, int num_batches,
                                int num_slots, int32_t* const* slot_ids, int32_t* const* slot_mask,
                                int32_t* const* slot_labels, int32_t pad_id, int32_t ignore_label);
void hstd_loader_destroy(hstd_loader* l);
int hstd_loader_start_epoch(hstd_loader* l);
int hstd_loader_next(hstd_loader* l);
int hstd_loader_release(hstd_loader* l, int slot);

#ifdef __cplusplus
}
#endif
#endif /* HSTD_B200_H_ */
