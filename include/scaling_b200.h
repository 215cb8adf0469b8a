/* C ABI of the scaling_b200 native library (scaling_b200/_kernels.so, sm_100a).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on and returns 0 on success, a CUDA
 * error code (> 0) or a negative argument error.  None of them allocates device memory: scratch buffers, where a
 * kernel needs one, are passed in by the caller.  dtype codes: 0 = bf16, 1 = fp32, 2 = fp16.
 *
 * File:line citations name the reference (marcobellagente93/scaling, paths relative to src/scaling/) behaviour each
 * entry point replaces.
 */
#pragma once
#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- tcgen05 GEMM family (csrc/gemm_sm100.cu) — replaces cuBLAS behind F.linear:
 *      core/nn/linear/column_parallel_linear.py:151, row_parallel_linear.py:158, core/nn/mlp.py:157-161 ---- */
/* D[M,N] (+)= A·Bᵀ.  a_layout/b_layout: 0 = K-major ([M,K] / [N,K] row-major), 1 = MN-major ([K,M] / [K,N]).
 * out_dtype: 0 bf16, 1 fp32.  epilogue: 0 store, 1 accumulate into D.  bias: optional bf16 [N].  max_ctas: 0 = all SMs. */
int sb_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb,
                 int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas, cudaStream_t stream);
/* H[M,F] = silu(X·W1ᵀ) ⊙ (X·W2ᵀ), W = [W1; W2] as one [2F,K] buffer; pre (optional) keeps (gate | up) [M,2F]. */
int sb_gemm_swiglu_bf16(const void* X, const void* W, void* H, void* pre, int M, int F, int K, int ldx, int ldw,
                        int ldh, int ldpre, int max_ctas, cudaStream_t stream);

/* 2-CTA (tcgen05 cta_group::2, 256x256 tiles) variants with the same contracts; the Python layer selects them for
 * M >= 256 and N >= 128.  The last, partial wave of tiles is split along K when each part keeps >= 112 k-blocks
 * (fp32 partials meet in the caller-owned workspace registered with sb_gemm2_set_split_workspace; without one the
 * tail is not split; disable with SCALING_B200_GEMM_SPLITK=0).
 * Reference counterpart: cuBLAS GEMM behind F.linear (core/nn/linear/column_parallel_linear.py:151). */
int sb_gemm2_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb,
                  int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas, cudaStream_t st);
int sb_gemm2_swiglu_bf16(const void* X, const void* W, void* H, void* pre, int M, int F, int K, int ldx, int ldw,
                         int ldh, int ldpre, int max_ctas, cudaStream_t st);
/* fp16 operands on the same 2-CTA kernel, for the three GEMMs of a linear layer: fprop (a_layout 0, b_layout 0), dgrad
 * (0, 1), wgrad (1, 1).  out_dtype 0 = fp16, 1 = fp32 (wgrad only); epilogue 1 accumulates into an fp32 D (wgrad only);
 * bias is fp16.  Returns -10 for any other combination.  Reference counterpart: the fp16 + dynamic-loss-scaler
 * training grid (tests/core/test_training/test_training.py:149-235), cuBLAS behind F.linear. */
int sb_gemm2_f16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb, int ldd,
                 int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas, cudaStream_t st);
/* Split-K tail workspace of the 2-CTA GEMMs: size for the current device, and registration (once per device, with that
 * device current) of a caller-owned 16-byte aligned buffer of at least that size whose trailing counter words are
 * zero.  GEMMs that split must then be issued on one stream at a time per device.  NULL unregisters. */
long long sb_gemm2_split_workspace_bytes(void);
int sb_gemm2_set_split_workspace(void* base, unsigned long long bytes);

/* Fused tensor-parallel GEMM + all-reduce over NVLink peer memory: D = sum over ranks of op(A)·op(B) (+ bias), bf16,
 * bit-identical on every rank.  data_ptrs[r] / flag_ptrs[r]: rank r's symmetric workspace regions of this call's
 * parity, as mapped into this process (host arrays of tp_size addresses); epoch: per-workspace call counter, equal on
 * all ranks; slot_bytes >= sb_gemm2_allreduce_slot_bytes(M, N); flag_slot >= slot_bytes / 65536 u32 flags per sender.
 * a_layout must be 0 (K-major).  Reference counterpart: GEMM followed by torch.distributed.all_reduce
 * (core/nn/linear/row_parallel_linear.py:150-160; backward of copy_to_tensor_model_parallel_region, utils.py:20-60). */
int sb_gemm2_allreduce_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda,
                            int ldb, int ldd, int a_layout, int b_layout, const unsigned long long* data_ptrs,
                            const unsigned long long* flag_ptrs, int tp_rank, int tp_size, unsigned int epoch,
                            unsigned long long slot_bytes, unsigned int flag_slot, int max_ctas, cudaStream_t st);
long long sb_gemm2_allreduce_slot_bytes(int M, int N);
/* Same launch, reduce-scatter along the sequence (Megatron sequence parallelism): rows of the product are [batch, seq],
 * rank r receives its slice seq/tp of every batch item, D = [M / tp_size, N]; needs (seq / tp_size) % 256 == 0.
 * Reference counterpart: F.linear then reduce_scatter (core/nn/linear/utils.py:279-300). */
int sb_gemm2_reduce_scatter_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda,
                                 int ldb, int ldd, int a_layout, int b_layout, const unsigned long long* data_ptrs,
                                 const unsigned long long* flag_ptrs, int tp_rank, int tp_size, unsigned int epoch,
                                 unsigned long long slot_bytes, unsigned int flag_slot, int seq, int max_ctas,
                                 cudaStream_t st);
/* LoRA-fused linear: D (+)= A·Bᵀ + A2·B2ᵀ in one accumulator; the low-rank pair (A2 [M,K2]; B2 [N,K2] or,
 * with b_layout 1, [K2,N]) rides as extra k-blocks behind the main K range.  K2 % 8 == 0.  The _allreduce_ form sums the
 * result over the TP group in the same launch.  Reference: extra F.linear calls + add (core/nn/lora.py:103-112) and an
 * all-gather of the attention output for the dense projection (core/nn/attention/attention.py:690-701). */
int sb_gemm2_lowrank_bf16(const void* A, const void* B, const void* A2, const void* B2, void* D, const void* bias, int M,
                          int N, int K, int K2, int lda, int ldb, int lda2, int ldb2, int ldd, int b_layout, int out_dtype,
                          int epilogue, int max_ctas, cudaStream_t st);
int sb_gemm2_allreduce_lowrank_bf16(const void* A, const void* B, const void* A2, const void* B2, void* D,
                                    const void* bias, int M, int N, int K, int K2, int lda, int ldb, int lda2, int ldb2,
                                    int ldd, int b_layout, const unsigned long long* data_ptrs,
                                    const unsigned long long* flag_ptrs, int tp_rank, int tp_size, unsigned int epoch,
                                    unsigned long long slot_bytes, unsigned int flag_slot, int max_ctas, cudaStream_t st);
/* LM-head GEMM with the cross-entropy statistics in its epilogue: logits [M,N] bf16 and stats
 * [M, sb_gemm2_xent_tiles(N)] float4 = (max, Σexp(x − max), unused, arg-max column as int bits) per 256-column tile.
 * Reference: F.linear + all-gather + fp32 copy + cross_entropy (lm_head_tied.py:41-43, model.py:56-76). */
long long sb_gemm2_xent_tiles(int N);
int sb_gemm2_xent_stats_bf16(const void* X, const void* W, void* logits, float* stats, int M, int N, int K, int ldx,
                             int ldw, int ldl, int max_ctas, cudaStream_t st);
/* all-gather -> GEMM: arrival gate for the NEXT sb_gemm2_* launch on the current device.  gate[tm] counts
 * the row chunks of A's 256-row tile tm that have arrived from the TP peers (sb_seq_gather_push_counted); the TMA
 * producer starts a tile when gate[tm] >= target.  One-shot.  Reference: all_gather_into_tensor before F.linear
 * (core/nn/linear/utils.py:305, core/nn/norm/rms_norm.py:57-61). */
int sb_gemm2_set_gate(const unsigned int* gate, unsigned int target);
/* Optional phase profile of the fused collective kernels: caller-owned buffer of grid x 8 u64 %globaltimer stamps for
 * the current device (NULL = off).  No reference counterpart. */
int sb_gemm2_set_profile_buffer(unsigned long long* buf);

/* ---- peer-memory runtime and collectives issued by this library's own kernels (csrc/runtime.cu, elementwise.cu,
 *      optim.cu) — SURVEY §8 C1 / E1 / O2 / G2.  All pointer arrays are host arrays of peer-mapped addresses in rank
 *      order (own rank included). ---- */
/* Device-side barrier over <= 8 GPUs: flag_ptrs[r] = rank r's u32[world] flag array; epoch grows by 1 per barrier. */
int sb_device_barrier(const unsigned long long* flag_ptrs, int rank, int world, unsigned int epoch, cudaStream_t st);
/* Data-parallel gradient all-reduce without ZeRO: this rank sums ITS chunk of every arena and stores the sum into every
 * arena (reference core/optimizer/allreduce.py:16-22). */
int sb_dp_allreduce(const unsigned long long* peer_chunks, int n_peers, long long n, int grad_dtype, cudaStream_t st);
/* Vocab-parallel embedding reduced by placement: the rank owning a token's row stores it into every rank's output (seq ==
 * 0) or into the output of the rank owning the token's sequence position (seq > 0: sequence-parallel scatter); no dense
 * all-reduce (reference core/nn/linear/vocab_parallel_embedding.py:122-140). */
int sb_embedding_fwd_peers(const long long* ids, const void* W, const unsigned long long* out_ptrs, int tp, int tokens,
                           int hidden, long long vocab_start, long long vocab_end, int seq, cudaStream_t st);
/* Sequence-parallel all-gather / reduce-scatter by peer stores (reference core/nn/linear/utils.py:279-305); follow each
 * with sb_device_barrier, except the _counted gather whose consumer is a gated GEMM (sb_gemm2_set_gate). */
int sb_seq_gather_push(const void* src, const unsigned long long* out_ptrs, int tp, int rank, int batch, int s_local,
                       int hidden, cudaStream_t st);
int sb_seq_gather_chunks_per_tile(void);
int sb_seq_gather_push_counted(const void* src, const unsigned long long* out_ptrs, const unsigned long long* gate_ptrs,
                               int tp, int rank, int batch, int s_local, int hidden, cudaStream_t st);
int sb_seq_scatter_push(const void* src, const unsigned long long* out_ptrs, int tp, int rank, int batch, int s_local,
                        int hidden, cudaStream_t st);
int sb_sum_slots(const void* slots, void* out, int tp, long long n, cudaStream_t st);
/* Cross-entropy from the LM-head epilogue statistics: merge the tiles, pick the target logit, exchange 16 bytes per
 * token across the TP group (peer stores), then — after a barrier — loss / lse / arg-max per token. */
int sb_xent_combine(const float* partial, int tiles, const void* logits, long long ld, const long long* targets,
                    int tokens, int vocab_local, long long vocab_start, const unsigned long long* out_ptrs, int tp,
                    int rank, cudaStream_t st);
int sb_xent_finalize(const float* all, int tp, int tokens, float* loss, float* lse, long long* argmax, cudaStream_t st);

/* Fused ZeRO-1 step over data-parallel peer memory (csrc/optim.cu): reduce-scatter as peer loads, all-gather as peer
 * stores inside AdamW.  Reference counterpart: all-reduce + slice, AdamW, all_gather (parameter_group.py:346-481). */
int sb_zero_reduce(const unsigned long long* peer_grads, int n_peers, void* own, long long n, int grad_dtype,
                   cudaStream_t st);
int sb_adamw_zero(float* p32, float* m, float* v, void* g, const unsigned long long* peer_outs, int n_peers, long long n,
                  int grad_dtype, int out_dtype, float lr, float beta1, float beta2, float eps, float weight_decay,
                  int step, float inv_loss_scale, float clip, const float* norm_sq, const int* skip, int zero_grad,
                  cudaStream_t st);
/* Fused q|k|v split + rotary over up to three head segments (csrc/elementwise.cu). */
int sb_qkv_rope(const unsigned long long* ins, const unsigned long long* outs, const long long* in_pitch,
                const long long* out_pitch, const int* heads, const int* rotate, int n_seg, const float* cos_t,
                const float* sin_t, const int* pos, int tokens, int head_dim, int rot_dim, int seq, int mode,
                int backward, cudaStream_t st);

/* Flash attention for packed (varlen) sequences, head_dim 128, bf16, GQA; tcgen05 + TMEM + TMA.
 * q/out: [T, heads, 128] (row pitch q_pitch elements), k/v: [T, kv_heads, 128]; seg_start/seg_end: per token, the
 * first token / one past the last token of its sequence; lse: [heads, T] fp32.
 * Backward is ONE launch: dK/dV blocks and dQ blocks share the grid (heaviest first).
 * Reference counterpart: flash_attn_varlen_func (core/nn/attention/attention.py:245-258). */
int sb_attention_fwd(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                     const int* seg_end, int T, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                     long long v_pitch, float scale, int causal, cudaStream_t st);
/* Generalised key side: k, v hold Tk rows (KV-cache decoding, Tk != T needs causal == 0) and with causal == 0 the arrays
 * are per-row KEY bounds "row t attends to keys [seg_start[t], seg_end[t])", non-decreasing in t: causal, right-aligned
 * causal, sliding windows (reference attention.py:242-257, 571-592, 599-667) are all just bounds.
 * dropout_p > 0 drops attention probabilities with a stateless hash of (dropout_seed, head_offset + head, q row, key):
 * forward and backward regenerate the same mask, nothing is stored; head_offset is the first global head of a
 * tensor-parallel shard so every shard draws its own stream. */
int sb_attention_fwd_ex(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                        const int* seg_end, int T, int Tk, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                        long long v_pitch, float scale, int causal, float dropout_p, unsigned long long dropout_seed,
                        int head_offset, cudaStream_t st);
int sb_attention_bwd_pitch(int T);
long long sb_attention_bwd_scratch_floats(int T, int heads);
int sb_attention_bwd(const void* dout, const void* q, const void* k, const void* v, const void* out, const float* lse,
                     float* scratch, void* dq, void* dk, void* dv, const int* seg_start, const int* seg_end, int T,
                     int heads, int kv_heads, long long q_pitch, long long k_pitch, long long v_pitch, float scale,
                     int causal, cudaStream_t st);
/* As sb_attention_bwd plus key_qend (int32 [T] or NULL): one past the last query that sees key t (sliding windows: the
 * dK/dV blocks stop there instead of at the end of the sequence). */
int sb_attention_bwd_ex(const void* dout, const void* q, const void* k, const void* v, const void* out, const float* lse,
                        float* scratch, void* dq, void* dk, void* dv, const int* seg_start, const int* seg_end,
                        const int* key_qend, int T, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                        long long v_pitch, float scale, int causal, float dropout_p, unsigned long long dropout_seed,
                        int head_offset, cudaStream_t st);

/* sb_attention_fwd_ex / sb_attention_bwd_ex with fp16 tensors (q, k, v, out, dout, dq, dk, dv and the probabilities
 * handed back to the tensor core are fp16; statistics stay fp32). */
int sb_attention_fwd_f16(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                         const int* seg_end, int T, int Tk, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                         long long v_pitch, float scale, int causal, float dropout_p, unsigned long long dropout_seed,
                         int head_offset, cudaStream_t st);
int sb_attention_bwd_f16(const void* dout, const void* q, const void* k, const void* v, const void* out, const float* lse,
                         float* scratch, void* dq, void* dk, void* dv, const int* seg_start, const int* seg_end,
                         const int* key_qend, int T, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                         long long v_pitch, float scale, int causal, float dropout_p, unsigned long long dropout_seed,
                         int head_offset, cudaStream_t st);

/* Bind `device`'s primary context to the calling thread (tensor-map encoding is a driver call and needs one; PyTorch's
 * autograd worker threads set their device lazily). */
int sb_bind_device(int device);

/* ---- norms (csrc/norm.cu) — core/nn/norm/rms_norm.py:45-53, core/nn/norm/layernorm.py:72-79 ---- */
int sb_norm_fwd(const void* x, const void* res, const void* w, const void* b, void* y, void* sum_out, float* mean,
                float* rstd, int rows, int cols, float eps, int rms, int dtype, cudaStream_t st);
long long sb_norm_bwd_scratch_floats(int rows, int cols, int rms);
int sb_norm_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd, const void* dres,
                void* dx, void* dw, void* db, int grad_fp32, int accumulate, float* scratch, int rows, int cols,
                int rms, int dtype, cudaStream_t st);

/* No counterpart in the reference.  sb_norm_bwd for callers that cannot allocate device memory (stream capture,
 * memory fully committed elsewhere, foreign allocator, plain C hosts): the fp32 partial buffer is a static array inside
 * the library (4.625 MiB per GPU that loads it; the current device must own the stream) instead of a caller argument.  Limits: cols <= 8192 (RMSNorm) /
 * 4096 (LayerNorm), else status -3.  The first call on each device resolves the array's device address — issue it
 * outside stream capture.  Calls share the array: do not overlap two of them on the GPU. */
int sb_norm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                        const void* dres, void* dx, void* dw, void* db, int grad_fp32, int accumulate, int rows,
                        int cols, int rms, int dtype, cudaStream_t st);
int sb_norm_bwd_noalloc_max_cols(int rms);
const char* sb_norm_bwd_noalloc_strerror(int code);

/* ---- element-wise (csrc/elementwise.cu) — core/nn/mlp.py:157-160, core/nn/rotary.py:68-139,
 *      transformer/model/layers/layer.py:211-216, core/nn/linear/vocab_parallel_embedding.py:119-147 ---- */
int sb_swiglu_fwd(const void* pre, void* h, int rows, int F, cudaStream_t st);
int sb_swiglu_bwd(const void* pre, const void* dh, void* dpre, void* h_out, int rows, int F, cudaStream_t st);
int sb_act_fwd(const void* x, void* y, long long n, int act, cudaStream_t st);
int sb_act_bwd(const void* x, const void* dy, void* dx, long long n, int act, cudaStream_t st);
int sb_rope(void* x, const float* cos_t, const float* sin_t, const int* pos, int tokens, int heads, int head_dim,
            int rot_dim, long long token_stride, int seq, int mode, int backward, cudaStream_t st);
int sb_bias_dropout_add(const void* x, const void* bias, const void* residual, void* out, long long n, int cols,
                        float p, unsigned long long seed, unsigned long long offset, cudaStream_t st);
int sb_dropout_bwd(const void* dout, void* dx, long long n, float p, unsigned long long seed,
                   unsigned long long offset, cudaStream_t st);
int sb_embedding_fwd(const long long* ids, const void* W, void* out, int tokens, int hidden, long long vocab_start,
                     long long vocab_end, cudaStream_t st);
int sb_embedding_bwd(const long long* ids, const void* dout, void* dW, int grad_fp32, int tokens, int hidden,
                     long long vocab_start, long long vocab_end, cudaStream_t st);
/* Deterministic form: sorted_ids = stable sort of the token ids, perm = the token each sorted id came from; one block per
 * run of equal ids sums the run in token order and adds it to dW without atomics. */
int sb_embedding_bwd_sorted(const long long* sorted_ids, const long long* perm, const void* dout, void* dW, int grad_fp32,
                            int tokens, int hidden, long long vocab_start, long long vocab_end, cudaStream_t st);

/* ---- vocab-parallel cross-entropy (csrc/xent.cu) — transformer/model/model.py:56-76,
 *      transformer/model/layers/lm_head_tied.py:41-43 ---- */
int sb_xent_stats(const void* logits, const long long* targets, float* stats, int tokens, int vocab_local,
                  long long ld, long long vocab_start, int dtype, cudaStream_t st);
int sb_xent_grad(void* logits, const long long* targets, const float* lse, const float* coeff, const float* gscale,
                 int tokens, int vocab_local, long long ld, long long vocab_start, int dtype, cudaStream_t st);

/* ---- optimizer (csrc/optim.cu) — core/optimizer/optimizer.py:107-208, parameter_group.py:346-535,
 *      loss_scaler.py:23-41 ---- */
int sb_grad_sumsq_scratch_floats(void);
int sb_grad_sumsq(const void* g, long long n, int dtype, float* scratch, float* out, int accumulate, int* nonfinite,
                  cudaStream_t st);
int sb_adamw(float* p32, float* m, float* v, void* g, void* out, long long n, int grad_dtype, int out_dtype, float lr,
             float beta1, float beta2, float eps, float weight_decay, int step, float inv_loss_scale, float clip,
             const float* norm_sq, const int* skip, int zero_grad, cudaStream_t st);
int sb_cast_from_fp32(const float* src, void* dst, long long n, int out_dtype, cudaStream_t st);

/* ---- data (csrc/blend_index.cpp) — core/data/blended_dataset.py:316-329 (Rust wheel blended_dataset_loop),
 *      transformer/data/text_dataset.py:371-385 ---- */
long long sb_blend_index(const long long* counts, int n_datasets, long long* out, long long capacity);
long long sb_gather_window_i32(const int* tokens, const long long* doc_offsets, const long long* pieces, int n_pieces,
                               int* out, long long capacity);

#ifdef __cplusplus
}
#endif
