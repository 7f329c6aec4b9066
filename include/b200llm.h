/* libb200llm - C ABI of the hand-written sm_100a kernels.
 *
 * Conventions: every function returns 0 on success, non-zero on error (message via
 * b200_last_error()); pointers are device pointers unless named h_*; `st` is the CUDA stream the
 * work is enqueued on; nothing synchronises the device and - unless a function says otherwise -
 * nothing allocates device memory, so the calls are capturable in CUDA graphs after a first
 * eager call (which sets function attributes and fills the TMA-descriptor cache).
 * bf16 tensors are passed as void*.  Matrices are row-major with an explicit leading dimension
 * in elements.  Citations "ref:" point at the reference code each entry point replaces.
 *
 * The Python side parses THIS FILE to build its ctypes signatures (ops/lib.py): keep one
 * declaration per statement, plain C types only.
 */
#ifndef B200LLM_H
#define B200LLM_H

#include <cuda_runtime.h>

#define B200_ABI_VERSION 1

#ifdef __cplusplus
extern "C" {
#endif

const char* b200_last_error(void);
int b200_abi_version(void);

/* ---- utility ------------------------------------------------------------------------------ */
int b200_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t st);
/* out[N] (+)= column sums of bf16 a[M,N] (bias gradients; ref: autograd of nn.Linear bias, mlp.py:24-26) */
int b200_colsum_bf16(const void* a, float* out, int M, int N, int ld, int accumulate, cudaStream_t st);
int b200_l2_flush(void* buf, long long bytes, cudaStream_t st);
/* zero-fill as a memset node (capturable; no kernel) */
int b200_memset_zero(void* ptr, long long bytes, cudaStream_t st);

/* ---- embedding (ref: src/models/transformer.py:51-54) -------------------------------------- */
/* x[b*T+t, :] = wte[idx[b,t], :] + wpe[pos0 + t, :]   (fp32) */
int b200_embed_fwd(const long long* idx, const float* wte, const float* wpe, float* x, int B, int T, int d, int V, int pos0, const int* pos_dev, cudaStream_t st);
/* dwte[idx[m], :] += dx[m, :] (vector RED; caller zeroes dwte on the first micro-step);
 * dwpe[t, :] (+)= sum_b dx[b, t, :] */
int b200_embed_bwd(const long long* idx, const float* dx, float* dwte, float* dwpe, int B, int T, int d, int V, int accumulate, cudaStream_t st);

/* Deterministic variant: `order` = stable argsort of idx (device, int64 [B*T]); runs of equal tokens are summed in
 * position order by one warp and the row is written once - no floating-point atomics, bit-identical run to run. */
int b200_embed_bwd_sorted(const long long* idx, const long long* order, const float* dx, float* dwte, float* dwpe, int B, int T, int d, int V, int accumulate, cudaStream_t st);

/* ---- fused residual-add + LayerNorm (ref: src/models/transformer_block.py:44,46; transformer.py:70) */
int b200_ln_supported(int d);
/* r = x (+ branch, then stored to x_out, which may alias x); out = LN(r) * gamma + beta.
 * out_bf16 / out_f32 / mean / rstd may each be null. */
int b200_ln_fwd(const float* x, const void* branch, float* x_out, const float* gamma, const float* beta, void* out_bf16, float* out_f32, float* mean, float* rstd, int M, int d, float eps, cudaStream_t st);
int b200_ln_bwd_nparts(int M, int d);
/* dx_out = (dx_in or 0) + LN-backward(dy); dx_bf16 = bf16(dx_out) (nullable); dgamma/dbeta (+)=;
 * colsum (nullable) (+)= column sums of dx_in.  `partial` is scratch of
 * b200_ln_bwd_nparts(M,d) * (colsum ? 3 : 2) * d floats.  dx_in may alias dx_out. */
int b200_ln_bwd(const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma, const float* dx_in, float* dx_out, void* dx_bf16, float* partial, float* dgamma, float* dbeta, float* colsum, int M, int d, int accumulate, cudaStream_t st);
/* Two-step form of the same backward: `_partial` runs the main kernel only and leaves the per-CTA partial sums in
 * `partial` (same size as above; with_colsum selects the 3-accumulator layout); `_finalize2` then reduces the partials
 * of TWO LayerNorms of the same (M, d) in one launch (pass colsum_x iff that LayerNorm ran with_colsum). */
int b200_ln_bwd_partial(const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma, const float* dx_in, float* dx_out, void* dx_bf16, float* partial, int with_colsum, int M, int d, cudaStream_t st);
int b200_ln_bwd_finalize2(const float* partial_a, float* dgamma_a, float* dbeta_a, float* colsum_a, const float* partial_b, float* dgamma_b, float* dbeta_b, float* colsum_b, int M, int d, int accumulate, cudaStream_t st);

/* Same operation for callers that cannot allocate device memory (a host that gave the whole GPU to its own
 * arena; a loop that must not meet cudaMalloc's device-wide synchronisation).  This entry point has no
 * counterpart in the reference.  The partial-sum scratch is a fixed array inside the library image
 * (b200_ln_bwd_static_capacity() floats = 6.94 MiB held by every GPU that loads the library); requests that
 * need more are refused.  The first call on a device resolves the array's address (make it before any stream
 * capture); calls that can overlap on one GPU race on the shared array; nothing needs zeroing between calls. */
long long b200_ln_bwd_static_capacity(void);
int b200_ln_bwd_static(const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma, const float* dx_in, float* dx_out, void* dx_bf16, float* dgamma, float* dbeta, float* colsum, int M, int d, int accumulate, cudaStream_t st);

/* ---- tcgen05 GEMM (ref: every nn.Linear fwd/bwd: attention.py:47-55, mlp.py:53,66, transformer.py:71)
 * C[M,N] = alpha * opA * opB^T.   a_mn = 0: A is [M,K] (K contiguous); 1: A is [K,M].
 *                                  b_mn = 0: B is [N,K] (K contiguous); 1: B is [K,N].
 * out_f32 = 0: C bf16 via TMA store, epilogue = (+bias[N]) (ReLU) (* (mask[M,N] > 0)), and if
 *              `stats` != null: float2 (max, sum exp) of the bf16 outputs per row, n-tile (width
 *              b200_gemm_pick_bn() or bn_override) and epilogue warp group g (g < 2 for tiles of 128 or 256 columns,
 *              1 group for 64) -> stats[(n_tile * groups + g) * M + row]; consumers merge any number of entries.
 *              `relu_bits_out` / `relu_bits`: one 64-bit word per row and 64 columns, [M][ceil(N/64)]; the GEMM
 *              that applies the ReLU writes bit j = (output column 64 w + j > 0), the matching dgrad GEMM zeroes
 *              the outputs whose bit is clear (cheaper than re-reading the activations as `mask`).
 * out_f32 = 1: C fp32, f32_mode 0 store / 1 accumulate / 2 atomic add (enables split-K;
 *              split_k = 0 picks automatically, C must then hold the running sum). */
int b200_gemm_pick_bn(int M, int N, int K, int out_f32);
int b200_gemm(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb, long long ldc, int a_mn, int b_mn, int out_f32, int f32_mode, float alpha, const float* bias, int relu, const void* mask, int ld_mask, void* stats, void* relu_bits_out, const void* relu_bits, int bn_override, int split_k, cudaStream_t st);

/* Same contract on CTA pairs (tcgen05 cta_group::2, cluster of 2): tiles of 256 x bn (bn = 256 | 128), each CTA
 * stages half of the operands, 6-8 pipeline stages.  The default for large problems. */
int b200_gemm_2cta(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb, long long ldc, int a_mn, int b_mn, int out_f32, int f32_mode, float alpha, const float* bias, int relu, const void* mask, int ld_mask, void* stats, void* relu_bits_out, const void* relu_bits, int bn, int split_k, cudaStream_t st);

/* bf16-output form of b200_gemm_2cta that also accumulates colsum[N] (fp32) += column sums of the output - the bias
 * gradient of the Linear whose output gradient this dgrad GEMM produces (ref: autograd of nn.Linear bias, mlp.py:53),
 * summed from the staged smem tile instead of by a second pass over the output.  Zero colsum for a fresh sum. */
int b200_gemm_2cta_colsum(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb, long long ldc, int a_mn, int b_mn, float alpha, const float* bias, int relu, void* relu_bits_out, const void* relu_bits, int bn, float* colsum, cudaStream_t st);

/* All-gather fused into a GEMM (sharded data parallelism; absent in the reference - SURVEY N6/T11).
 * C[M,N] = A[M,K] * opB^T (+bias) (ReLU -> relu_bits_out | masked by relu_bits), bf16 out; Bw points into a LOCAL
 * window of gathered layer weights (b_mn = 0: [N,K]; 1: [K,N]).  The launch carries njobs (0..2) gather jobs: extra
 * warps copy a layer chunk by chunk from h_srcJ[r] (rank r's chunk in its peer-mapped bf16 shard; own chunk first)
 * into dstJ + r * chunkJ (totalJ elements in all) and raise flagsJ[r] = *gen_dev; countersJ: [world] zeroed ints.
 * wait_flags != null (K-major B only): the B operand's own window is being filled by job 0 of this very launch -
 * tiles wait per chunk (wait_chunk elements per rank, B starts at element w_offset of the window), own rows first. */
int b200_gemm_ag(const void* A, const void* Bw, void* C, int M, int N, int K, long long lda, long long ldb, long long ldc, int b_mn, const float* bias, int relu, void* relu_bits_out, const void* relu_bits, const int* wait_flags, long long wait_chunk, long long w_offset, int world, int rank, const int* gen_dev, int* err, int njobs, void* const* h_src0, void* dst0, long long chunk0, long long total0, int* flags0, void* counters0, void* const* h_src1, void* dst1, long long chunk1, long long total1, int* flags1, void* counters1, cudaStream_t st);

/* Weight-streaming GEMM for decode-sized batches (M <= 128): out[M,N] = act(x[M,K] @ W[N,K]^T + bias), operands
 * swapped so the weight tile is the 128-row MMA operand; K is split over the CTAs of a thread-block cluster
 * (split_k = 0: chosen automatically, else 1/2/4/8) whose partial tiles are combined through distributed shared
 * memory - no scratch, no atomics. */
int b200_gemm_skinny_splits(int M, int N, int K);
int b200_gemm_skinny(const void* x, const void* W, void* out, int M, int N, int K, long long ldx, long long ldw, long long ldc, const float* bias, int relu, int split_k, cudaStream_t st);

/* ---- fused AdamW (ref: scripts/train_transformer.py:92,126 torch.optim.AdamW defaults) -------
 * hyper (device, 8 floats) = 1 - lr*weight_decay, beta2, 1 - beta1, 1 - beta2, lr / bias_correction1,
 * sqrt(bias_correction2), eps, grad_scale (derive them in double on the host: ops.adamw_hyper).
 * p_bf16 (nullable) receives the bf16 shadow weights. */
int b200_adamw(float* p, const float* g, float* m, float* v, void* p_bf16, const float* hyper, long long n, cudaStream_t st);

/* ---- softmax cross-entropy on bf16 logits (ref: src/models/transformer.py:74-77) ------------
 * finalize: combines the GEMM-epilogue statistics stats[n_tiles][M] into lse[M], reads the
 * target logit, writes loss_rows[M] = lse - logit[target] and loss_sum[0] += sum_scale * sum of rows. */
int b200_ce_finalize(const void* stats, int n_tiles, const void* logits, long long ld, const long long* targets, float* lse, float* loss_rows, float* loss_sum, int M, int V, float sum_scale, cudaStream_t st);
/* in place: logits[m, v] = (exp(logits - lse[m]) - [v == target[m]]) * scale  (bf16), and
 * dbias[V] (+)= column sums of the result (ref: lm_head bias gradient) */
int b200_ce_grad(void* logits, long long ld, const float* lse, const long long* targets, float* dbias, int M, int V, float scale, int accumulate, cudaStream_t st);

/* ---- causal flash attention on packed qkv (ref: src/models/attention.py:45-58,95) -----------
 * qkv [B*T, 3*H*hs] bf16 (q | k | v thirds, head h at columns h*hs); out [B*T, H*hs] bf16
 * head-major (the reference's torch.cat of heads); lse [B, H, T] fp32. hs in {64, 128}. */
int b200_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale, cudaStream_t st);
/* same contract on tcgen05/TMEM/TMA (S and O accumulate in tensor memory; the default forward) */
int b200_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale, cudaStream_t st);
/* Same, with the number of query tiles per CTA chosen by the caller: 1 (two CTAs per SM at hs = 64) or 2 (persistent
 * CTAs, two softmax groups half a tile apart); 0 = the default of b200_attn_fwd_tc (1 for hs = 64, 2 for hs = 128). */
int b200_attn_fwd_tc_tiles(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale, int tiles, cudaStream_t st);
/* dqkv [B*T, 3*H*hs] bf16; delta [B, H, T] fp32 scratch; dq_acc [B*T, H*hs] fp32 scratch */
int b200_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, float* dq_acc, void* dqkv, int B, int T, int H, int hs, float scale, cudaStream_t st);

/* tcgen05 backward, two passes of one kernel (key-stationary: dK, dV; query-stationary: dQ), all
 * accumulators in TMEM, no atomics; delta [B, H, T] fp32 is the only scratch. */
int b200_attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, void* dqkv, int B, int T, int H, int hs, float scale, cudaStream_t st);

/* ---- generation (ref: src/models/transformer.py:96-114) ------------------------------------
 * pos_dev (nullable, device int): when given it overrides pos0 / len (= *pos_dev + 1) / adds to the
 * sampler step, so that one captured CUDA graph can be replayed for every decode position.
 * KV cache layout [B, H, Tmax, hs] bf16 for k and v.  append: copies the k/v thirds of
 * qkv [B*t, 3*H*hs] into the caches at positions pos0..pos0+t-1. */
int b200_kv_append(const void* qkv, void* kcache, void* vcache, int B, int t, int H, int hs, int Tmax, int pos0, const int* pos_dev, cudaStream_t st);
/* single-position decode attention: q = q-third of qkv [B, 3*H*hs], attends to cache[0..len).
 * append != 0: first write this position's k / v thirds into the caches at index len-1 (fused kv_append). */
int b200_attn_decode(const void* qkv, void* kcache, void* vcache, void* out, int B, int H, int hs, int Tmax, int len, float scale, const int* pos_dev, int append, cudaStream_t st);
/* Split-KV form (same contract): grid (H, B, splits), each CTA takes 1/splits of the cached keys with coalesced row
 * loads (8 key + 8 value rows in flight per warp); the last CTA of a (b, h) pair to arrive merges the partials.
 * scratch: B*H*splits*(hs+2) floats; tickets: B*H zeroed unsigned ints (left zero); splits == 1 needs neither. */
int b200_attn_decode_split(const void* qkv, void* kcache, void* vcache, void* out, float* scratch, void* tickets, int B, int H, int hs, int Tmax, int len, float scale, const int* pos_dev, int append, int splits, cudaStream_t st);
/* next[b] ~ softmax(logits[b, :] / temperature) restricted to the top_k largest (top_k <= 0: all);
 * Philox-free counter RNG: (seed, step, b) -> uniform.  logits bf16 [B, ld]. */
int b200_sample(const void* logits, long long ld, long long* next, int B, int V, float temperature, int top_k, unsigned long long seed, unsigned long long step, const int* pos_dev, cudaStream_t st);

int b200_incr_int(int* p, int delta, cudaStream_t st);
/* seq[b, col] = next[b]  (col = *pos_dev + 1 when pos_dev is given) */
int b200_scatter_tokens(const long long* next, long long* seq, int B, long long ld, int col, const int* pos_dev, cudaStream_t st);
/* out[b, :] = in[b*T + T-1, :] : the last-position rows, for the last-token-only LM head */
int b200_gather_last_rows(const void* in, void* out, int B, int T, int row_bytes, cudaStream_t st);

/* ---- data parallel over NVLink (ref: torch DDP + AdamW, scripts/train_transformer.py:122-123,92) ----
 * Symmetric memory bootstrap: cudaMalloc'd buffers exported / imported as 64-byte CUDA IPC handles.
 * b200_ipc_alloc is the one allocating call of this library (set-up time, never per step). */
int b200_ipc_alloc(void** ptr, long long bytes);
int b200_ipc_free(void* ptr);
int b200_ipc_export(void* ptr, void* h_handle_out);
int b200_ipc_import(const void* h_handle, void** ptr_out);
int b200_ipc_close(void* ptr);
/* Symmetric heap on the virtual-memory-management API + NVSwitch multicast objects (csrc/comm/vmm.cu).
 * query: out[5] = {VMM supported, POSIX-fd handles supported, multicast supported, allocation granularity,
 * multicast granularity for `world` devices}.  create: cuMemCreate on the current device (+ an fd of THIS process
 * to hand to the peers over a unix socket); import: fd -> handle; map: reserve + map + read/write access for the
 * current device (local memory, a peer's memory and a multicast object alike).  Multicast: one rank creates the
 * object, every rank adds its device, then - once ALL have - binds its local allocation and maps the object:
 * stores to that mapping (multimem.st) land in every rank's copy, multimem.ld_reduce returns the switch-reduced sum. */
int b200_vmm_query(long long* out, int world);
int b200_vmm_create(long long bytes, unsigned long long* handle_out, int* fd_out);
int b200_vmm_import(int fd, unsigned long long* handle_out);
int b200_vmm_map(unsigned long long handle, long long bytes, void** ptr_out);
int b200_vmm_unmap(void* ptr, long long bytes);
int b200_vmm_release(unsigned long long handle);
int b200_vmm_close_fd(int fd);
int b200_mc_create(long long bytes, int world, unsigned long long* handle_out, int* fd_out);
int b200_mc_add_device(unsigned long long mc_handle);
int b200_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, long long bytes);
/* Fused per-bucket gradient reduce-scatter (P2P loads) + scale + AdamW on the owned 1/W slice
 * [start, start+n) + all-gather of the updated weights (P2P stores of bf16, and fp32 where flagged).
 * h_grad/h_wb/h_flat/h_sig: host arrays of `world` peer-mapped device pointers (own rank included);
 * sig: int[2][world][num_buckets] per rank (ready | done), zero-initialised; *epoch_dev: 1, 2, 3, ...
 * (a device int the caller increments once per step, so a captured graph can be replayed);
 * m, v: this rank's moments for the slice; h_slot_end / h_slot_fp32: element end offsets and
 * "consumers read fp32" flags of the parameter slots inside the bucket.
 * mc_grad / mc_wb / mc_flat (all three or none): multicast mappings of the same buffers; when given, the pull is a
 * multimem.ld_reduce (summed inside the NVSwitch) and each push one multimem.st replicated by the switch. */
int b200_dp_reduce_adamw(void* const* h_grad, void* const* h_wb, void* const* h_flat, void* const* h_sig, int world, int rank, int bucket, int num_buckets, const int* epoch_dev, long long start, long long n, float* m, float* v, const float* hyper, const long long* h_slot_end, const int* h_slot_fp32, int n_slots, void* cta_counter, int* err, int ctas, void* mc_grad, void* mc_wb, void* mc_flat, cudaStream_t st);
/* Sharded mode (csrc/comm/zero.cu): same fused reduce + AdamW, but nothing full-size is kept - h_grad[r] is rank r's
 * gradient buffer OF THIS BUCKET (a rolling window or the head / embedding buffer; mc_grad its multicast mapping or
 * null), the owned slice [lo, lo+n) of the bucket is reduced (multimem.ld_reduce or W peer loads), master / m / v /
 * shard16 are this rank's slices (fp32 master, moments, bf16 shard the peers' all-gather reads), and parameters
 * consumed in fp32 are also stored into every rank's small replicated store (h_small / mc_small) at
 * bucket offset + h_slot_small_off[k] (slots whose value is -(1<<62) are not small).  gacc (nullable): this rank's
 * fp32 accumulator for the slice (gradient accumulation): accumulate_only != 0 adds the reduced slice to it and updates
 * nothing; otherwise the accumulator is added to the reduced slice, cleared, and the update runs. */
int b200_zero_reduce_adamw(void* const* h_grad, void* mc_grad, void* const* h_small, void* mc_small, void* const* h_sig, int world, int rank, int bucket, int num_buckets, const int* epoch_dev, long long lo, long long n, float* master, float* m, float* v, void* shard16, const float* hyper, const long long* h_slot_end, const long long* h_slot_small_off, int n_slots, void* cta_counter, int* err, int ctas, float* gacc, int accumulate_only, cudaStream_t st);
/* blocks the stream until every rank's done flag of buckets [bucket_lo, bucket_hi) has reached *epoch_dev - lag */
int b200_zero_wait(const int* sig, int world, int num_buckets, int bucket_lo, int bucket_hi, const int* epoch_dev, int lag, int* err, cudaStream_t st);
/* x[b*T+t, :] = wte[idx[b,t], :] + wpe[t, :] with the fp32 tables sharded: h_shard[r] = rank r's master slice of the
 * embedding bucket (chunk elements per rank), wte_off / wpe_off = element offsets of the two tables in the bucket */
int b200_embed_fwd_sharded(const long long* idx, void* const* h_shard, int world, long long chunk, long long wte_off, long long wpe_off, float* x, int B, int T, int d, int V, cudaStream_t st);
/* blocks the stream until every owner's done flag of every bucket has reached `epoch` */
int b200_dp_wait_done(const int* sig, int world, int num_buckets, const int* epoch_dev, int* err, cudaStream_t st);

/* ---- vocab-parallel LM head + cross-entropy (ref: replicated lm_head + F.cross_entropy, transformer.py:71-77)
 * Every rank owns Vs consecutive vocabulary rows.  All h_* arguments are host arrays of `world` peer-mapped device
 * pointers; h_sig: signal pads int[4][world]; `slot` selects the hand-shake; *epoch_dev is the device step counter.
 * gather_rows : dst_all[src*M + i, :] = hf[src][i, :], tgt_all[src*M + i] = tgt[src][i] - tgt_shift
 * ce_finalize : for this rank's M rows, combine stats (float2 [nt_local][world*M]) of all slices, fetch the target
 *               logit from its owner's logits [world*M, ldl], loss_sum += sum_scale * sum, push lse to every rank
 * reduce_rows : out[i, :] = scale * sum_src partial[src][rank*M + i, :]   (bf16) */
int b200_vp_barrier(void* const* h_sig, int world, int rank, int slot, const int* epoch_dev, int* err, cudaStream_t st);
int b200_vp_gather_rows(void* const* h_hf, void* const* h_tgt, void* dst_all, long long* tgt_all, int M, int d, int world, int rank, long long tgt_shift, void* const* h_sig, int slot, const int* epoch_dev, int* err, cudaStream_t st);
int b200_vp_ce_finalize(void* const* h_stats, void* const* h_logits, void* const* h_lse_all, const long long* targets, float* loss_sum, float* loss_rows, int M, int world, int rank, int V, int Vs, int nt_local, long long ldl, float sum_scale, void* const* h_sig, int slot, const int* epoch_dev, int* err, cudaStream_t st);
int b200_vp_reduce_rows(void* const* h_partial, void* out, int M, int d, int world, int rank, float scale, void* const* h_sig, int slot, const int* epoch_dev, int* err, cudaStream_t st);

#ifdef __cplusplus
}
#endif
#endif
