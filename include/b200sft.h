/*
 * b200sft.h -- C ABI of libb200sft.so: the B200 (sm_100a) data-parallel hot path of the SFT step
 * of thesteve0/llm-fine-tune-distributed.
 *
 * The reference (a Python script, training.py) has no native code; the arithmetic of its step
 * runs in third-party packages.  Each entry point below names the reference call site it replaces
 * (paths under the reference checkout) and the third-party function that carries the arithmetic
 * ("SP/" = site-packages of torch 2.11 / transformers 5.5).
 *
 * Conventions
 *   - every pointer is a CUDA device pointer on the current device unless marked "host";
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - no entry point synchronises the host or allocates device memory: scratch is supplied by the
 *     caller, sized by the matching *_workspace_bytes() query (b200sft_flce_fwd_noalloc carries
 *     its own fixed scratch instead);
 *   - return value 0 = success, negative = error (b200sft_last_error() gives the text);
 *   - bf16 tensors are row-major and 16-byte aligned, with hidden size H and vocabulary V
 *     multiples of 8 (TMA needs 16-byte row pitches).
 */
#ifndef B200SFT_H_
#define B200SFT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Flags for the fused LM-head + cross-entropy calls. */
#define B200SFT_FLAG_ROUND_LOGITS_BF16 1u /* round each logit to bf16 before the softmax, as the      \
                                             reference does between lm_head (bf16 out) and            \
                                             logits.float() -- SP/transformers/loss/loss_utils.py:55 */

/* How the summed loss is normalised (SP/transformers/loss/loss_utils.py:28-42). */
#define B200SFT_DENOM_NONE 0   /* loss = scale * sum                                  */
#define B200SFT_DENOM_DEVICE 1 /* loss = scale * sum / *denom  (num_items_in_batch)    */
#define B200SFT_DENOM_VALID 2  /* loss = scale * sum / (#rows with label != ignore): the
                                  reduction="mean" branch taken when num_items_in_batch is None */

const char* b200sft_last_error(void);
int b200sft_version(void);
/* Process-wide tile configuration of the tcgen05 GEMM engine: 1 = one CTA per 128x256 tile
 * (tcgen05.mma.cta_group::1), 2 = two-CTA clusters on 256x256 tiles (cta_group::2, operands split
 * across the SM pair; the default).  Results are identical; no reference counterpart. */
int b200sft_set_cta_group(int cta_group);

/* ---------------------------------------------------------------------------------------------
 * Label front-end.  Replaces the shift in ForCausalLMLoss (SP/transformers/loss/loss_utils.py:57-60:
 * pad right with ignore_index, drop the first column, flatten) and the count in
 * Trainer._get_num_items_in_batch (SP/transformers/trainer.py:2136: labels.ne(-100).sum(), on the
 * UNSHIFTED labels), both reached from trainer.train() at training.py:300.
 *   labels   [B,T] int64
 *   shifted  [B*T] int32  out: shifted[b,t] = labels[b,t+1], last column = ignore_index
 *   counts   [2]   int64  out: counts[0] = #(labels != ignore) unshifted, counts[1] = same, shifted
 * Bit-exact integer work.
 */
int b200sft_label_prep(const int64_t* labels, int B, int T, int64_t ignore_index, int32_t* shifted,
                       int64_t* counts, void* stream);

/* The whole label front-end in one launch: what the collators do to the labels on the host, the
 * shift, both counts and the live-row compaction list that b200sft_gather_rows / flce consume.
 * The unshifted label of token i is
 *   (labels ? labels : input_ids)[i]          labels = input_ids copy: data_collator.py:1428
 *   -> ignore_index where attention_mask[i] == 0     padding (labels padded with -100)
 *   -> ignore_index where completion_mask[i] == 0    prompt tokens under completion-only loss
 *                                                    [TRL collator, not on disk]
 *   -> ignore_index where position_ids[i] == 0       first token of a packed sequence, the separator
 *                                                    of DataCollatorWithFlattening
 *                                                    (SP/transformers/data/data_collator.py:1426-1428)
 * followed by the shift and counts of b200sft_label_prep.  Every mask pointer may be NULL (= no such
 * mask); all are [B,T] int64.
 *   labels_out [B,T] int64 out or NULL: the unshifted labels (what inputs["labels"] would hold)
 *   shifted    [B*T] int32 out;  counts [2] int64 out (as b200sft_label_prep)
 *   row_index, labels_c, pos [B*T] int32, n_live int32 scalar: outputs as b200sft_compact_rows, or
 *   all four NULL to skip the compaction.  B*T < 2^31.  Bit-exact integer work; reached from
 *   training.py:300 (collator -> compute_loss).
 */
int b200sft_label_frontend(const int64_t* input_ids, const int64_t* labels, const int64_t* attention_mask,
                           const int64_t* completion_mask, const int64_t* position_ids, int B, int T,
                           int64_t ignore_index, int64_t* labels_out, int32_t* shifted, int64_t* counts,
                           int32_t* row_index, int32_t* labels_c, int32_t* pos, int32_t* n_live,
                           void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused LM-head + masked cross-entropy, forward.  Replaces, for training.py:300's step,
 *   logits = lm_head(hidden)            SP/transformers/models/smollm3/modeling_smollm3.py:504
 *   logits.float(); cross_entropy(sum)  SP/transformers/loss/loss_utils.py:55,35-36
 *   loss / num_items_in_batch           SP/transformers/loss/loss_utils.py:37-41
 * without materialising the [N,V] logits: a TMA-fed tcgen05 GEMM whose epilogue keeps an online
 * max / sum-exp per row.
 *   x        [N,H] bf16   final hidden states (already shifted-aligned with `labels`)
 *   w        [V,H] bf16   lm_head.weight (tied to embed_tokens.weight)
 *   labels   [N]   int32  shifted labels; ignore_index rows contribute nothing.  Any other label
 *            outside [0,V) makes the loss NaN (the reference trips nll_loss's device assert there;
 *            a kernel cannot raise, and the host never reads the labels).
 *   denom    int64 scalar (device) used when denom_mode == B200SFT_DENOM_DEVICE, else may be NULL
 *   n_rows_dev  int32 scalar (device) or NULL: number of live rows (<= N) when x/labels are the
 *            compacted tensors from b200sft_compact_rows; N is then the capacity.  NULL = all N rows.
 *   loss     fp32 scalar out
 *   lse      [N] fp32 out: log-sum-exp of every (live) row (saved for backward)
 *   n_valid  int64 scalar out (may be NULL): #rows with label != ignore_index
 */
size_t b200sft_flce_fwd_workspace_bytes(int N, int H, int V);
int b200sft_flce_fwd(const void* x, const void* w, const int32_t* labels, int N, int H, int V,
                     int ignore_index, float scale, int denom_mode, const int64_t* denom, uint32_t flags,
                     const int32_t* n_rows_dev, float* loss, float* lse, int64_t* n_valid, void* workspace,
                     size_t workspace_bytes, void* stream);

/* Same forward, additionally producing the two logits-based training metrics TRL's SFTTrainer logs
 * ([TRL, not on disk]: mean_token_accuracy and entropy, computed there from outputs.logits) without
 * logits: n_correct = #valid rows whose target logit is the row maximum (argmax == label; exact
 * ties count as correct), entropy_sum = sum over valid rows of the softmax entropy
 * lse - E_p[z].  One extra fused multiply-add per logit in the epilogue.
 */
size_t b200sft_flce_fwd_stats_workspace_bytes(int N, int H, int V);
int b200sft_flce_fwd_stats(const void* x, const void* w, const int32_t* labels, int N, int H, int V,
                           int ignore_index, float scale, int denom_mode, const int64_t* denom,
                           uint32_t flags, const int32_t* n_rows_dev, float* loss, float* lse,
                           int64_t* n_valid, int64_t* n_correct, float* entropy_sum, void* workspace,
                           size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Valid-row compaction.  Rows whose label is ignore_index (padding, prompt tokens, the first token
 * of every packed sample -- SP/transformers/data/data_collator.py:1426-1428) contribute nothing to
 * the loss or to dW and get a zero dX row (torch: ignore_index in F.cross_entropy), yet the
 * reference still pays the full [N,V] GEMMs for them.  b200sft_compact_rows gathers the live rows
 * so that flce_fwd / flce_bwd run on them only; b200sft_scatter_rows expands dX back.  The live
 * count never leaves the device.
 *   x [N,H] bf16, labels [N] int32                                  in
 *   xc [N,H] bf16        live rows first (stable order), zero-filled up to the next 128-row boundary
 *   labels_c [N] int32   their labels, ignore_index beyond the live count
 *   row_index [N] int32  row_index[j] = original row of compact row j
 *   pos [N] int32        compact position of each original row, -1 if ignored
 *   n_live int32 scalar  live row count
 */
int b200sft_compact_rows(const void* x, const int32_t* labels, int N, int H, int ignore_index, void* xc,
                         int32_t* labels_c, int32_t* row_index, int32_t* pos, int32_t* n_live, void* stream);
/*   The list half alone (row_index, labels_c, pos, n_live as above) from shifted int32 labels. */
int b200sft_compact_labels(const int32_t* labels, int N, int ignore_index, int32_t* row_index, int32_t* labels_c,
                           int32_t* pos, int32_t* n_live, void* stream);
/*   The gather half alone, for a compaction list made by b200sft_label_frontend:
 *   xc[j,:] = x[row_index[j],:] for j < *n_live, zero rows up to the next 128-row boundary. */
int b200sft_gather_rows(const void* x, const int32_t* row_index, const int32_t* n_live, int N, int H, void* xc,
                        void* stream);
/*   dx[r,:] = dxc[pos[r],:] when pos[r] >= 0, else 0 (every row of dx is written). */
int b200sft_scatter_rows(const void* dxc, const int32_t* pos, int N, int H, void* dx, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Final RMSNorm as the producer of the head's X operand.  Replaces, for the rows the loss scores,
 *   hidden_states = self.norm(hidden_states)   SP/transformers/models/smollm3/modeling_smollm3.py:438
 *   SmolLM3RMSNorm.forward                      modeling_smollm3.py:271-276
 * (fp32 statistics; the normalised value rounded to bf16; bf16 multiply by the weight) and its
 * autograd backward, fused with the row gather / scatter of the compaction above.
 *   h [N,H] bf16 pre-norm hidden states;  gamma [H] bf16 norm weight;  eps = variance_epsilon
 *   row_index, n_live: compaction list (b200sft_label_frontend) or both NULL = all N rows in place
 *   xc [N,H] bf16 out: normalised live rows first, zero rows up to the next 128-row boundary
 *   rstd [N] fp32 out: rsqrt(mean(h^2) + eps) per (compact) row, saved for the backward
 */
int b200sft_rmsnorm_gather(const void* h, const void* gamma, float eps, const int32_t* row_index,
                           const int32_t* n_live, int N, int H, void* xc, float* rstd, void* stream);
/*   dyc [N,H] bf16: gradient w.r.t. the normalised (compact) rows, i.e. dX of b200sft_flce_bwd
 *   pos, row_index, n_live: the compaction maps, or all NULL
 *   dh [N,H] bf16 out (or NULL): gradient w.r.t. h, zero rows where pos < 0
 *   dgamma [H] bf16 out (or NULL): gradient w.r.t. gamma (fixed summation order); needs
 *   b200sft_rmsnorm_bwd_workspace_bytes(H) bytes of workspace */
size_t b200sft_rmsnorm_bwd_workspace_bytes(int H);
int b200sft_rmsnorm_scatter_bwd(const void* dyc, const void* h, const void* gamma, const float* rstd,
                                const int32_t* pos, const int32_t* row_index, const int32_t* n_live, int N,
                                int H, void* dh, void* dgamma, void* workspace, size_t workspace_bytes,
                                void* stream);

/* ---------------------------------------------------------------------------------------------
 * Forward, for callers that cannot allocate.  Has no counterpart in the reference.
 *
 * b200sft_flce_fwd needs scratch that somebody must allocate.  Some users of this ABI cannot do
 * that: a ctypes / cgo-style stub that is handed tensors by another framework and owns no device
 * allocator; a loss or evaluation callback that runs while PyTorch's caching allocator already
 * holds nearly all of HBM, so a raw cudaMalloc fails; code on the training loop's critical path,
 * where cudaMalloc / cudaFree would synchronise the whole device and stall the communication
 * stream that is overlapping backward.  This variant takes the same inputs and outputs, computes
 * exactly what b200sft_flce_fwd computes with the same kernels, and uses a fixed scratch array
 * that is part of the library image instead.  It calls no cudaMalloc, cudaFree or any allocating
 * routine.
 *
 * Limit: the call needs b200sft_flce_fwd_workspace_bytes(N,H,V) bytes and is refused with an
 * error if that exceeds B200SFT_NOALLOC_SCRATCH_BYTES (40 MiB -- enough for 8,192 rows against
 * the 128,256-token vocabulary, or 16,384 rows against 64k).  Larger batches: call it per row
 * block, or use b200sft_flce_fwd.  Every GPU that loads the library keeps those 40 MiB reserved
 * until the process exits.
 *
 * The first call on each device looks up the array's device address (with that device current);
 * make that first call before starting any stream capture.  All calls on a device share the one
 * array, which still holds the previous call's contents: the call itself clears the part that
 * must start at zero, but two calls that overlap on the GPU (different streams, no dependency)
 * race on it -- serialise them.  The 40 MiB are reserved when the CUDA runtime loads the library's
 * device image on that device, which happens at the latest during this first call: make it while
 * that much HBM is still free, not after another allocator has taken the rest.
 */
#define B200SFT_NOALLOC_SCRATCH_BYTES (40u << 20)
int b200sft_flce_fwd_noalloc(const void* x, const void* w, const int32_t* labels, int N, int H, int V,
                             int ignore_index, float scale, int denom_mode, const int64_t* denom,
                             uint32_t flags, float* loss, float* lse, int64_t* n_valid, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused LM-head + masked cross-entropy, backward.  Replaces autograd's backward of the three
 * calls above (nll_loss_backward, log_softmax_backward, the fp32->bf16 cast of dlogits and the two
 * cuBLAS GEMMs dX = dlogits.W, dW = dlogits^T.X).  dlogits = (softmax - onehot) * grad_out * scale
 * / denom on valid rows, 0 on ignored rows, rounded to bf16 exactly where the reference rounds;
 * it only ever exists as one [N, chunk_cols] bf16 slab of vocabulary columns at a time, never as
 * an [N,V] tensor; dW slabs are final after one pass, dX accumulates in fp32 across slabs.
 *   lse       [N] fp32 from the forward call
 *   grad_out  fp32 scalar (device): upstream gradient of `loss`
 *   dx        [N,H] bf16 out, or NULL to skip the dX GEMM
 *   dw        [V,H] bf16 out (overwritten), or NULL to skip the dW GEMM (frozen head)
 *   chunk_cols  vocabulary columns of dlogits held in scratch at a time (rounded up to a multiple
 *               of 256; 0 = library default: about 1 GiB of scratch, split into equal slabs)
 *   n_rows_dev  as in b200sft_flce_fwd (compacted inputs; dx is then compact too)
 */
size_t b200sft_flce_bwd_workspace_bytes(int N, int H, int V, int chunk_cols);
/* Number of kernels one b200sft_flce_bwd call launches for this shape (per vocabulary slab: dlogits,
 * dX, dW GEMMs; plus the split-K combine) -- bookkeeping for benchmarks; no reference counterpart. */
int b200sft_flce_bwd_launch_count(int N, int H, int V, int chunk_cols, int need_dx, int need_dw);
int b200sft_flce_bwd(const void* x, const void* w, const int32_t* labels, const float* lse,
                     const float* grad_out, int N, int H, int V, int ignore_index, float scale,
                     int denom_mode, const int64_t* denom, uint32_t flags, int chunk_cols,
                     const int32_t* n_rows_dev, void* dx, void* dw, void* workspace, size_t workspace_bytes,
                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * Plain bf16 GEMM on the same tcgen05 engine: C[M,N] = A * B^T, fp32 accumulate, bf16 out.
 * Exported so the parity tests can pin each operand-layout path of the engine in isolation.
 *   a_mn_major = 0: A is [M,K] row-major (K contiguous); 1: A is stored [K,M] (M contiguous)
 *   b_mn_major = 0: B is [N,K] row-major;                1: B is stored [K,N]
 * lda/ldb/ldc are row pitches in elements.  No reference counterpart (cuBLAS does this there).
 */
int b200sft_gemm_bf16(const void* a, const void* b, void* c, int M, int N, int K, int lda, int ldb, int ldc,
                      int a_mn_major, int b_mn_major, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Gradient-bucket scale / cast kernels for the DDP communication hook.  Replace
 *   tensor.div_(world_size)                      SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-32
 *   buffer.to(bf16).div_(world_size); copy back  .../default_hooks.py:57-93 (bf16_compress_hook)
 * (DDP is configured at training.py:249-256).  128-bit vectorised, one contiguous run per block, HBM-bound.
 * Pass scale = 1/world rounded to fp32: on CUDA, ATen's div_ by a host scalar is itself a multiply by
 * that fp32 reciprocal (BinaryDivTrueKernel.cu: "a * reciprocal(b)"), so the result has the reference's
 * bits for every world size, not only powers of two (tests: worlds 2..8 against tensor.div_ on the GPU).
 * scale_cast rounds to bf16 first, scales in fp32 and rounds again, as buffer.to(bf16).div_(world) does.
 */
int b200sft_scale_bf16_inplace(void* buf, size_t n, float scale, void* stream);
int b200sft_scale_f32_inplace(float* buf, size_t n, float scale, void* stream);
int b200sft_scale_cast_f32_to_bf16(const float* src, void* dst, size_t n, float scale, void* stream);
int b200sft_cast_bf16_to_f32(const void* src, float* dst, size_t n, float scale, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Bucket allreduce over NVLink 5 / NVSwitch with the scale fused in: one pass that reduces a bf16
 * bucket across ranks (fp32 accumulate), multiplies by `scale` (1/world) and leaves the result in
 * every rank's copy.  Replaces the div_ + ncclAllReduce pair of default_hooks.py:18-32 for buckets
 * held in symmetric memory (torch.distributed._symmetric_memory).
 *   multimem: `multicast_ptr` is the NVLS multicast address of the bucket (multimem.ld_reduce /
 *             multimem.st: the reduction and the broadcast happen inside the switch);
 *   p2p:      `peer_ptrs_host` is a HOST array of `world` device pointers, the bucket's address on
 *             each rank (own rank included), for boxes without multicast support.
 * Precision: p2p returns bf16(fp32 sum * scale) exactly; the in-switch reduction of the multimem path hands back a
 * bf16 sum that is not rounded to nearest (measured on B200: up to 1.92 * 2^-8 relative off the exact sum, i.e. within
 * one bf16 spacing) -- the same order as the bf16 ring sum of the reference's NCCL allreduce.
 * Each rank processes slice `rank` of `world` equal slices; n (elements) must be a multiple of
 * 8*world.  The kernels never wait on another rank: the caller must order them across ranks with a
 * symmetric-memory barrier before (all copies written) and after (all results written) the call.
 */
int b200sft_allreduce_scale_bf16_multimem(void* multicast_ptr, size_t n, int rank, int world, float scale,
                                          void* stream);
int b200sft_allreduce_scale_bf16_p2p(const uint64_t* peer_ptrs_host, size_t n, int rank, int world, float scale,
                                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * Tied-embedding gradient exchanged as rows.  lm_head.weight is embed_tokens.weight
 * (SP/transformers/models/smollm3/modeling_smollm3.py:367,447) and is trained (training.py:136-141); the
 * embedding's part of its gradient -- autograd's embedding_dense_backward of `self.embed_tokens(input_ids)`
 * (modeling_smollm3.py:396), the last op of backward() -- is what keeps DDP from reducing that parameter's
 * 521 MiB bucket before backward() ends (training.py:249-256; default_hooks.py:18-32).  This call replaces that
 * op AND the cross-rank averaging of its result: every rank publishes its token ids and the gradient rows dY
 * w.r.t. the embedding output (symmetric memory, or plain device memory at world 1), then each rank adds
 *   grad[v,:] = bf16(grad[v,:] + scale * sum_{r, t: ids_r[t] == v} dY_r[t,:])          (fp32 sum)
 * for the touched rows only, in ONE fixed order (rank, then position) on every rank, so replicas stay bit-identical.
 * The dense head part of the gradient is then free to go through the bucket hook at the START of backward.
 *   ids_ptrs_host, dy_ptrs_host, n_ptrs_host   HOST arrays of `world` device pointers: rank r's int32 ids [cap]
 *            (negative id = skip: padding position, or nn.Embedding's padding_idx), its bf16 dY [cap,H] and an
 *            int32 device scalar with the number of valid entries (<= cap); peers' pointers are read with plain
 *            loads over NVLink -- the caller orders the ranks (barrier before: all published; after: all done)
 *   grad     [V,H] bf16 in/out (the tied weight's gradient after DDP has finalised the backward)
 *   scale    1/world for DDP's average (the loss already carries compute_loss's x num_processes)
 * H a multiple of 8, at most 8192; world <= 16.
 */
size_t b200sft_embed_grad_workspace_bytes(int world, int cap, int V, int H);
int b200sft_embed_grad_rows(const uint64_t* ids_ptrs_host, const uint64_t* dy_ptrs_host, const uint64_t* n_ptrs_host,
                            int world, int cap, int V, int H, float scale, void* grad, void* workspace,
                            size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Step tail on bf16 parameters: global gradient-norm clip + AdamW as two multi-tensor kernels.
 * Replaces accelerator.clip_grad_norm_(params, max_grad_norm=1.0) (training.py:264 ->
 * SP/transformers/trainer.py:2489-2499 -> torch.nn.utils.clip_grad_norm_) and torch.optim.AdamW.step on
 * the trainable tensors.  The pointer arrays are HOST arrays of n_tensors device pointers.
 *   grad_norm : total_norm (fp32 device scalar) = sqrt(sum g^2) over all tensors, fixed summation order
 *   adamw     : g = bf16(grad * min(1, max_norm / (total_norm + 1e-6)))  (max_norm <= 0: no clipping),
 *               then torch's AdamW update in fp32 with one bf16 rounding per stored value; `step`
 *               counts from 1.  Gradients are not modified.
 */
size_t b200sft_grad_norm_workspace_bytes(const size_t* numels, int n_tensors);
int b200sft_grad_norm_bf16(void* const* grads, const size_t* numels, int n_tensors, float* total_norm,
                           void* workspace, size_t workspace_bytes, void* stream);
int b200sft_adamw_clip_bf16(void* const* params, void* const* grads, void* const* exp_avgs,
                            void* const* exp_avg_sqs, const size_t* numels, int n_tensors, float lr, float beta1,
                            float beta2, float eps, float weight_decay, int step, float max_norm,
                            const float* total_norm, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SFT_H_ */
