/* C ABI of libpdb200_kernels.so - hand-written sm_100a kernels of the paddle_b200 framework.
 *
 * Conventions: every entry point returns 0 on success, a cudaError_t value or -1 on failure (message via
 * pdb200_last_error(), thread-local).  `stream` is a cudaStream_t passed as void*.  dtype codes: 0 = float32,
 * 1 = bfloat16, 2 = float16.  Pointers are device pointers unless noted.  No entry point synchronises the
 * device.  This header is also parsed by paddle_b200/lib/__init__.py to build the ctypes signatures, so keep one
 * declaration per statement and only the scalar / pointer types already used here.
 */
#ifndef PDB200_KERNELS_H_
#define PDB200_KERNELS_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define PDB_API

PDB_API const char* pdb200_last_error();
PDB_API int pdb200_abi_version();
PDB_API int pdb200_sm_count();

/* Local-memory guard (no counterpart in the reference).  Rule of this library: every TMA tensor map and every
 * tensor-core operand is this GPU's own memory; data of another GPU is first landed in local memory.  The
 * symmetric-memory layer registers each address range that maps another GPU's memory (peer segments, NVLS windows);
 * the GEMM / attention launchers refuse operands that overlap a registered range, and additionally verify with
 * cudaPointerGetAttributes that a base address belongs to the current device when they first build a map for it. */
PDB_API int pdb200_register_foreign_range(uint64_t base, uint64_t bytes);
PDB_API int pdb200_unregister_foreign_range(uint64_t base);
PDB_API int pdb200_check_local(const void* p, int64_t bytes);

/* K18/K19 fused RMSNorm / LayerNorm (reference: phi/kernels/gpu/rms_norm_kernel.cu:420-520, rms_norm_funcs.h:507-942,
 * layer_norm_kernel.cu).  residual (optional): y = norm(x + residual), residual_out = x + residual. */
PDB_API int pdb200_norm_fwd(const void* x, const void* residual, const void* w, const void* b, void* y, void* residual_out, float* mean, float* rstd, int64_t rows, int64_t cols, float eps, int x_dtype, int w_is_f32, int is_layernorm, void* stream);
PDB_API int64_t pdb200_norm_bwd_scratch_bytes(int64_t rows, int64_t cols, int is_layernorm);
PDB_API int pdb200_norm_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd, void* dx, void* dw, void* db, void* scratch, int64_t scratch_bytes, int64_t rows, int64_t cols, int x_dtype, int w_is_f32, int is_layernorm, void* stream);

/* K21 SwiGLU (reference: phi/kernels/gpu/swiglu_kernel.cu, swiglu_grad_kernel.cu); ld* are row strides in elements */
PDB_API int pdb200_swiglu_fwd(const void* x, const void* y, void* out, int64_t rows, int64_t H, int64_t ldx, int64_t ldy, int dtype, void* stream);
PDB_API int pdb200_swiglu_bwd(const void* x, const void* y, const void* dout, void* dx, void* dy, int64_t rows, int64_t H, int64_t ldx, int64_t ldy, int64_t lddx, int64_t lddy, int dtype, void* stream);

/* K20 fused rotary embedding (reference: phi/kernels/fusion/gpu/fused_rope_kernel.cu, fused_rope_utils.h:235-426).
 * rotate_half=1: halves rotate together (Paddle use_neox_rotary_style=False); 0: adjacent pairs (=True).
 * sign=+1 forward, -1 backward. sin/cos: fp32 [positions, d] or both null (computed from base). */
PDB_API int pdb200_rope(const void* q, const void* k, const void* v, void* oq, void* ok, void* ov, int hq, int hk, int hv, const float* sin_t, const float* cos_t, const int64_t* pos_ids, int64_t batch, int64_t seq, int d, int rotate_half, int time_major, float base, float sign, int dtype, void* stream);

/* K20b fused QKV split + rotary (split=1) / merge + inverse rotary (split=0, sign=-1): one pass instead of three
 * strided copies + rope (no counterpart op in the reference, which runs split/transpose kernels around fused_rope). */
PDB_API int pdb200_qkv_rope(const void* packed, void* q, void* k, void* v, int hq, int hk, const float* sin_t, const float* cos_t, const int64_t* pos_ids, int64_t batch, int64_t seq, int d, int rotate_half, float base, float sign, int split, int dtype, void* stream);

/* K22 fused bias + activation (reference: phi/kernels/fusion/gpu/fused_bias_act_kernel.cu:30-545).
 * act: 0 gelu, 1 relu, 2 swiglu, 3 geglu, 4 silu, 5 identity; GLU variants read 2*cols_out input columns. */
PDB_API int pdb200_bias_act(const void* x, const void* bias, void* out, int64_t rows, int64_t cols_out, int act, int dtype, void* stream);

/* K12/K15 multi-tensor optimizer kernels (reference: phi/kernels/gpu/adamw_kernel.cu:35-372, fused_adam_kernel.cu,
 * squared_l2_norm_kernel.cu, amp_kernel.cu).  table: MTDesc[ntensors] (64 B each, see csrc/kernels/optim.cu);
 * chunk_map: int32 pairs (tensor index, chunk index) - one CTA per pair, pdb200_mt_chunk_elems() elements each. */
PDB_API int pdb200_mt_chunk_elems();
PDB_API int pdb200_adamw_multi(const void* table, const void* chunk_map, int n_chunks, float lr, float beta1, float beta2, float eps, float weight_decay, float beta1_pow, float beta2_pow, const float* beta_pows_dev, const float* grad_scale, const float* found_inf, void* stream);
PDB_API int pdb200_momentum_multi(const void* table, const void* chunk_map, int n_chunks, float lr, float mu, float l2_decay, int nesterov, const float* grad_scale, const float* found_inf, void* stream);
PDB_API int pdb200_l2norm_multi(const void* table, const void* chunk_map, int n_chunks, float* scratch, int64_t scratch_floats, float* out, int accumulate, void* stream);
PDB_API int pdb200_unscale_check_multi(const void* table, const void* chunk_map, int n_chunks, const float* inv_scale, float* found_inf, void* stream);
/* No counterpart in the reference.  Same result as pdb200_l2norm_multi, but the per-chunk partial sums live in a
 * fixed 4 MiB array inside the library, for callers that cannot allocate device memory at the call site (a C host
 * without a tensor allocator; a step loop that must not let cudaMalloc synchronise the device).  Allocates nothing.
 * n_chunks above pdb200_l2norm_multi_static_max_chunks() is refused.  The first call on each device looks up the
 * array's address: make it before any stream capture.  The array is shared per device: calls that may overlap
 * (unordered streams / threads) race - keep them ordered on one stream. */
PDB_API int pdb200_l2norm_multi_static_max_chunks();
PDB_API int pdb200_l2norm_multi_static(const void* table, const void* chunk_map, int n_chunks, float* out, int accumulate, void* stream);

/* A2 tensor check (reference: phi/kernels/gpu/check_numerics_kernel.cu): stats = device int64[3] {nan, inf, zero counts},
 * values = device float[3] {max, min, sum} over the finite elements.  Nothing is read back by the library. */
PDB_API int pdb200_check_numerics(const void* x, int64_t numel, int dtype, int64_t* stats, float* values, void* stream);

/* K8 softmax cross-entropy, local / per-rank part (reference: phi/kernels/gpu/cross_entropy_kernel.cu,
 * fluid/operators/collective/c_softmax_with_cross_entropy_op.cu:47-114,133-340). stats: [rows,3] = max,sumexp,target */
PDB_API int pdb200_ce_fwd(const void* logits, const int64_t* labels, float* stats, float* loss, float* lse, int64_t rows, int64_t V, int64_t ld, int64_t vocab_start, int64_t ignore_index, int dtype, void* stream);
PDB_API int pdb200_ce_combine(const float* all_stats, const int64_t* labels, float* loss, float* lse, int64_t rows, int nranks, int64_t ignore_index, void* stream);
PDB_API int pdb200_ce_bwd(const void* logits, const int64_t* labels, const float* lse, const float* dloss, float dloss_scalar, void* dlogits, int64_t rows, int64_t V, int64_t ld, int64_t vocab_start, int64_t ignore_index, int dtype, void* stream);

/* K11 MoE routing (reference: phi/kernels/gpu/{number_count,assign_pos,limit_by_capacity,prune_gate_by_capacity,
 * random_routing}_kernel.cu) */
PDB_API int pdb200_number_count(const int64_t* ids, int64_t n, int64_t* out, int n_bins, void* stream);
PDB_API int pdb200_assign_pos(const int64_t* ids, int64_t n, const int64_t* cum_count, int n_bins, int64_t* out, void* stream);
PDB_API int pdb200_limit_by_capacity(const int64_t* expert_count, const int64_t* capacity, int64_t* out, int n_expert, int n_worker, void* stream);
PDB_API int pdb200_prune_gate_by_capacity(const int64_t* gate, int64_t* new_gate, const int64_t* expert_count, int* remaining_i32, int n_bins, int64_t n, void* stream);
PDB_API int pdb200_random_routing(int64_t* topk_idx, const void* topk_val, const void* prob, int64_t n, int dtype, void* stream);
PDB_API int pdb200_topk_softmax(const void* logits, int64_t tokens, int E, int k, int renorm, float* weights, int64_t* idx, int64_t* counts, int dtype, void* stream);

/* K23/K13 tcgen05 + TMA bf16 GEMM: C[M,N] (=|+=) act(op(A) op(B) + bias), fp32 accumulation in TMEM (reference:
 * cuBLASLt behind phi/kernels/funcs/blas/blaslt_impl.cu.h, fusion/gpu/fused_gemm_epilogue_kernel.cu:63-125,
 * fused_linear_param_grad_add_kernel.cu:35-289).  A: [M,K] row-major, or [K,M] when trans_a.  B: [K,N] row-major, or
 * [N,K] when trans_b.  out_dtype 0 = fp32 (beta = 1 accumulates into C), 1 = bf16.  act: 0 none, 1 gelu, 2 relu. */
PDB_API int pdb200_gemm_bf16(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K, int trans_a, int trans_b, int out_dtype, float beta, int act, void* stream);
/* Same contract, 256x256 output tiles computed by CTA pairs (tcgen05.mma.cta_group::2, cluster of 2). */
PDB_API int pdb200_gemm_bf16_pair(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K, int trans_a, int trans_b, int out_dtype, float beta, int act, void* stream);

/* Same contract (beta must be 0 or 1): 256x512 blocks per CTA pair - two tcgen05 accumulators sharing A - with a
 * swizzled-smem + TMA-store epilogue; fp32 C += is a TMA reduce-add (csrc/kernels/gemm_tc3.cu). */
PDB_API int pdb200_gemm_bf16_macro(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K, int trans_a, int trans_b, int out_dtype, float beta, int act, void* stream);
PDB_API int pdb200_gemm_set_band(int n);
/* K6 GEMM -> reduce-scatter, compute half: row blocks of block_rows rows are finished one after the other starting at pair-row
 * m_rot, and every finished piece is counted into done_counters[row / block_rows] (device uint32, +ceil(N/256)*8 per
 * 256 rows), so a stream can wait for a block (pdb200_nvl_wait_counter) and ship it while the rest is computed. */
PDB_API int pdb200_gemm_bf16_counted(const void* a, const void* b, void* c, int64_t M, int64_t N, int64_t K, int trans_b, void* done_counters, int block_rows, int m_rot, void* stream);
/* K5 all-gather -> GEMM, compute half: A is the gathered activation in LOCAL memory; rows [i*gate_rows, (i+1)*gate_rows)
 * are loaded only after the local uint32 gate_flags[i] has reached gate_epoch (set by the peer whose chunk landed, see
 * pdb200_nvl_signal); row blocks are visited from pair-row m_rot on (this rank's own rows first).  gate_rows % 256 == 0. */
PDB_API int pdb200_gemm_bf16_gated(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K, int trans_b, int act, const void* gate_flags, int gate_rows, int64_t gate_epoch, int m_rot, void* stream);
/* K9/K10 grouped GEMMs for the local experts of an MoE layer, scheduled on the device from int64 row counts (reference:
 * per-expert loop after blocking count reads, incubate/distributed/models/moe/moe_layer.py:441-470).  Rows are
 * expert-major in slabs of rows_cap rows.  forward/dgrad: C[r,:] = A[r,:] . B[e] (B stacked [E][K][N], or [E][N][K] when
 * trans_b).  wgrad: dW[e] (=|+=) A_e^T . G_e over the rows of expert e (out_dtype 0 = fp32, beta 1 accumulates). */
PDB_API int pdb200_gemm_grouped(const void* a, const void* b, void* c, const int64_t* counts, int E, int64_t rows_cap, int64_t N, int64_t K, int trans_b, int act, void* stream);
PDB_API int pdb200_gemm_grouped_wgrad(const void* a, const void* g, void* dw, const int64_t* counts, int E, int64_t rows_cap, int64_t Kin, int64_t N, int out_dtype, float beta, void* stream);
/* Tuning knob: raster band height (in 256-row pair tiles) of the CTA-pair kernel; 0 restores the default. */
PDB_API int pdb200_gemm_set_group_m(int n);

/* Fused collectives over NVLink symmetric memory (K1, K2+K3, K8; csrc/kernels/nvl_collectives.cu).  peer_ptrs: host
 * array of nranks device addresses = every rank's segment as mapped in this process (pdb200_symm_peer_ptr), mc_ptr:
 * NVLS (multimem) mapping of the segment or 0 (P2P fallback).  *_off: byte offsets inside the segment, identical on all ranks.
 * pad_off: 32-byte signal pad used by the in-stream barriers; each call consumes epochs `epoch` and `epoch + 1`.
 * Reference counterparts: reducer.cc:1093-1127 (K1), tensor_fusion_helper.py:559-645 + adamw_kernel.cu (K2+K3),
 * c_softmax_with_cross_entropy_op.cu:211-314 (K8). */
/* Tests on one GPU: emulation mode skips every cross-rank barrier, so the entry points can be called once per
 * emulated rank (segments of one GPU's memory as peer_ptrs, mc_ptr = 0), one after the other on one stream. */
PDB_API int pdb200_nvl_set_emulation(int on);
/* Optional trace: a mapped pinned host word (from pdb200_wd_progress_ptr of the runtime library) that every stand-alone
 * barrier of the fused kernels overwrites with the epoch it just passed; 0 switches it off. */
PDB_API int pdb200_nvl_set_progress(uint64_t host_mapped_ptr);
/* flag word (flag_off + 4*index) of every rank's segment <- value, system-scope release; enqueue it behind the copies it
 * announces. */
PDB_API int pdb200_nvl_signal(const uint64_t* peer_ptrs, int nranks, int rank, int64_t flag_off, int index, int64_t value, void* stream);
PDB_API int pdb200_nvl_signal_one(const uint64_t* peer_ptrs, int nranks, int rank, int target, int64_t flag_off, int index, int64_t value, void* stream);
/* Stream-ordered waits: until *counter (local, device-scope atomics of a running kernel) >= target; until `count` local
 * flag words (peers' signals) have all reached epoch (skip_index < 0: none skipped). */
PDB_API int pdb200_nvl_wait_counter(const void* counter, int64_t target, void* stream);
PDB_API int pdb200_nvl_wait_flags(const void* flags, int count, int64_t epoch, int skip_index, void* stream);
PDB_API int pdb200_nvl_barrier(const uint64_t* peer_ptrs, int nranks, int rank, int64_t pad_off, int64_t epoch, void* stream);
PDB_API int pdb200_nvl_allreduce(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t data_off, int64_t numel, int dtype, float scale, int64_t pad_off, int64_t epoch, int max_ctas, void* stream);
PDB_API int pdb200_nvl_rs_adamw_ag(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t grad_off, int grad_dtype, int64_t param_off, int64_t shard_start, int64_t shard_numel, float* master, float* m, float* v, float lr, float beta1, float beta2, float eps, float weight_decay, const float* beta_pows, const float* grad_scale, const float* found_inf, const float* local_grad, int64_t pad_off, int64_t epoch, void* stream);
PDB_API int pdb200_nvl_rs_grid(int64_t shard_numel);
/* Cap the grid of the reduce-scatter kernels (0 = default 4 CTAs per SM); affects pdb200_nvl_rs_grid too. */
PDB_API int pdb200_nvl_set_rs_max_ctas(int n);
PDB_API int pdb200_nvl_reduce_scatter(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t grad_off, int grad_dtype, int64_t shard_start, int64_t shard_numel, float* out, float* partial_sq, int64_t pad_off, int64_t epoch, void* stream);
/* K3 stand-alone: parameter all-gather as a push (every rank stores its shard into slot `rank` of all ranks' buffers: one NVLS multimem.st per 16 bytes, or a loop of peer stores). */
PDB_API int pdb200_nvl_allgather(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t src_off, int64_t dst_off, int64_t shard_bytes, int64_t pad_off, int64_t epoch, int max_ctas, void* stream);
PDB_API int pdb200_nvl_allgather_from(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, const void* src, int64_t dst_off, int64_t shard_bytes, int64_t pad_off, int64_t epoch, int max_ctas, void* stream);
/* ZeRO-3 unit transport (K3 / K2 pass 1 for GroupSharded stage 3; reference: group_sharded_stage3.py:741-803, 849-1118
 * per-parameter all_gather / all_reduce + slice).  One launch per unit: `segs` is a HOST array of nseg (<= 16) int64
 * triples and the cross-GPU barriers run inside the kernel.  all-gather triple = {src_off, dst_off, shard_bytes};
 * reduce-scatter triple = {stage_off, shard_numel, local fp32 out pointer}: out (+)= scale * sum_r stage_r[shard]. */
PDB_API int pdb200_nvl_allgather_multi(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, const int64_t* segs, int nseg, int64_t pad_off, int64_t epoch, int max_ctas, void* stream);
PDB_API int pdb200_nvl_reduce_scatter_multi(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, const int64_t* segs, int nseg, int grad_dtype, float scale, int accumulate, int64_t pad_off, int64_t epoch, int max_ctas, void* stream);
/* Local half of the copy-engine reduce-scatter: out (+)= scale * sum over the nranks receive slots; segs = HOST triples
 * {recv pointer (local), shard_numel, fp32 out pointer}. */
PDB_API int pdb200_reduce_slots_multi(const int64_t* segs, int nseg, int nranks, int grad_dtype, float scale, int accumulate, int max_ctas, void* stream);
PDB_API int pdb200_reduce_slots_multi2(const int64_t* segs, int nseg, int nranks, int grad_dtype, float scale, int accumulate, int out_dtype, int max_ctas, void* stream);
/* Sum reduce-scatter of a symmetric bf16 buffer into a local bf16 shard (tail of GEMM -> reduce-scatter; NVLS in-switch add). */
PDB_API int pdb200_nvl_rs_sum_bf16(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t src_off, int64_t shard_start, int64_t shard_numel, void* out, int64_t pad_off, int64_t epoch, void* stream);
PDB_API int pdb200_nvl_ce_combine(const uint64_t* peer_ptrs, int nranks, int rank, int64_t stats_off, const int64_t* labels, float* loss, float* lse, int64_t rows, int64_t ignore_index, int64_t pad_off, int64_t epoch, void* stream);
/* K7 vocab-parallel embedding forward as a gather from the owning rank's table shard (replaces masked lookup + all-reduce). */
PDB_API int pdb200_nvl_vocab_embed(const uint64_t* peer_ptrs, int nranks, int rank, int64_t table_off, const int64_t* ids, void* out, int64_t n_tokens, int hidden, int64_t vocab_per_rank, int dtype, int64_t pad_off, int64_t epoch, void* stream);

/* K16 flash-attention forward on tcgen05 (reference: phi/kernels/gpu/flash_attn_kernel.cu:36-170, 305-447 ->
 * FlashAttention-2 sm80 kernels).  bf16 or fp16, d in {64,128}, GQA, causal aligned to the last key.
 * Dense: q [b,sq,hq,d], k/v [b,sk,hk,d], any sq/sk >= 1; o like q; lse [b,hq,sq] fp32 (natural log of the softmax
 * denominator of the scaled scores; -inf for a row that sees no key, whose output is 0).
 * Packed (varlen): q [total_q,hq,d], k/v [total_k,hk,d], cu_seqlens_* int32 [b+1] ON THE DEVICE (nothing is read
 * back), lse [hq,total_q].
 * Dropout: score (b*hq+h, q, k) is kept iff an 8-bit hash of (seed, offset, b*hq+h, q, k) is below round((1-p)*256);
 * the backward regenerates the same mask from the same (p_drop, seed, offset). */
PDB_API int pdb200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale, int causal, int dtype, void* stream);
PDB_API int pdb200_flash_attn_fwd_dropout(const void* q, const void* k, const void* v, void* o, float* lse, int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale, int causal, float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream);
PDB_API int pdb200_flash_attn_varlen_fwd(const void* q, const void* k, const void* v, void* o, float* lse, const int* cu_seqlens_q, const int* cu_seqlens_k, int64_t b, int64_t total_q, int64_t total_k, int64_t max_seqlen_q, int64_t max_seqlen_k, int64_t hq, int64_t hk, int64_t d, float scale, int causal, float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream);
/* K17 backward: Delta pre-pass + dK/dV kernel (transposed products, GQA heads looped in-CTA) + dQ kernel; deterministic,
 * no atomics.  workspace: pdb200_flash_attn_bwd_workspace_bytes() bytes of device memory (16-byte aligned); for the
 * packed layout call it with b = 1, sq = total_q. */
PDB_API int64_t pdb200_flash_attn_bwd_workspace_bytes(int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d);
PDB_API int pdb200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse, void* dq, void* dk, void* dv, void* workspace, int64_t workspace_bytes, int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale, int causal, int dtype, void* stream);
PDB_API int pdb200_flash_attn_bwd_dropout(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse, void* dq, void* dk, void* dv, void* workspace, int64_t workspace_bytes, int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale, int causal, float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream);
PDB_API int pdb200_flash_attn_varlen_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse, void* dq, void* dk, void* dv, void* workspace, int64_t workspace_bytes, const int* cu_seqlens_q, const int* cu_seqlens_k, int64_t b, int64_t total_q, int64_t total_k, int64_t max_seqlen_q, int64_t max_seqlen_k, int64_t hq, int64_t hk, int64_t d, float scale, int causal, float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream);

/* K9 / K10 MoE dispatch / combine over NVLink symmetric memory with device-side routing (csrc/kernels/moe_nvl.cu). */
PDB_API int pdb200_moe_dispatch(const uint64_t* peer_ptrs, int nranks, int rank, int64_t counts_off, int64_t slab_off, int64_t slab_rows, const void* src, const int64_t* pos, int div, const int64_t* lec, int n_expert_local, int64_t n_slots, int row_bytes, int32_t* route_rank, int32_t* route_row, int64_t* recv_counts, int32_t* overflow, int64_t pad_off, int64_t epoch, void* stream);
PDB_API int pdb200_moe_pull(const uint64_t* peer_ptrs, int nranks, int rank, int64_t slab_off, const int32_t* route_rank, const int32_t* route_row, const int64_t* inv, int64_t n_entries, void* out, int row_bytes, int sum_over_k, int dtype, int64_t pad_off, int64_t epoch, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PDB200_KERNELS_H_ */
