/*
 * b200_kv.h — C ABI of the B200 (sm_100a) paged-KV hot path.
 *
 * This is the drop-in boundary underneath the reference's Python surface
 * (Baozhi888/distributed-gpu-inference).  The reference has no native code, so every entry point below cites
 * the Python routine whose work it takes over; citations are relative to the reference checkout.
 *
 * Conventions (all entry points)
 *   - extern "C", plain pointers and sizes; no torch / C++ types cross the boundary.
 *   - every data pointer is a DEVICE pointer owned by the caller and kept alive until the stream has run.
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*); nothing here synchronises, allocates
 *     or frees device memory, so calls may be captured into a CUDA graph (b200_paged_decode_attn_static has a
 *     one-time set-up step of its own; see its comment).
 *   - return value: 0 on success, otherwise a negative B200_E* code (argument errors) or a positive
 *     cudaError_t; b200_last_error() returns a thread-local description of the last failure.
 *   - `dtype` uses the reference's wire ids (common/serialization.py:20-27):
 *         0 = float16, 1 = float32, 2 = bfloat16.
 *   - pool layout is the reference's (worker/distributed/kv_cache.py:130-144):
 *         k_pool, v_pool : [max_blocks, Hkv, block_size, D], contiguous.
 */
#ifndef B200_KV_H_
#define B200_KV_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 1

#define B200_DTYPE_F16 0
#define B200_DTYPE_F32 1
#define B200_DTYPE_BF16 2

#define B200_EINVAL (-1)   /* bad argument (null pointer, negative size, unsupported dtype) */
#define B200_ESHAPE (-2)   /* shape not supported by any kernel variant */
#define B200_ENOSPC (-3)   /* workspace too small / size beyond a fixed limit */

/* b200_paged_decode_attn flags */
#define B200_ATTN_FORCE_GENERIC 1u /* skip the TMA + tensor-core variant (used by tests to cross-check) */
#define B200_ATTN_ROUND_SCORES 2u  /* round q.k and scaled scores to the storage dtype before softmax, the way
                                      the reference's eager HF attention does (bf16/f16 only) */
#define B200_ATTN_WS_COUNTERS_ZERO 4u /* shared-prefix entry points, only when the partials are merged inside the kernel
                                      by the last CTA to arrive (B200_PREFIX_SEPARATE_COMBINE=0; the default is a
                                      dependent combine launch, which needs no counters): the first 4 * B * Hq bytes of
                                      the workspace (one arrival counter per (sequence, q head)) are known to be zero:
                                      the workspace was zero-filled once, is
                                      used by nothing but these entry points with the same B and Hq, and no earlier call
                                      on it was aborted.  Every call leaves the counters zero again.  Without the flag
                                      the library clears them itself (one cudaMemsetAsync on the stream per call). */

int b200_abi_version(void);
const char* b200_last_error(void);

/*
 * Paged KV append.  Replaces the `torch.cat((past_k, k), dim=2)` growth of the dense cache inside the HF layer
 * that ModelShard.forward calls (worker/distributed/model_shard.py:205-222) and fills the blocks that
 * PagedKVCache.allocate_block hands out (worker/distributed/kv_cache.py:153-194).
 *
 *   k_new, v_new : [n_tok, Hkv, D]; row n starts at k_new + n * k_row_stride (elements), heads contiguous
 *   slot_mapping : int64[n_tok]; slot = block_table[t / block_size] * block_size + t % block_size;
 *                  a negative slot skips the token (padding)
 * Element address in a pool: ((slot / bs * Hkv + h) * bs + slot % bs) * D + d.
 */
int b200_kv_append(const void* k_new, const void* v_new, int64_t k_row_stride, int64_t v_row_stride,
                   void* k_pool, void* v_pool, const int64_t* slot_mapping, int n_tok,
                   int Hkv, int D, int block_size, int64_t max_blocks, int dtype, void* stream);

/*
 * Paged GQA decode attention, one query token per sequence.  Replaces repeat_kv + QK^T + softmax + PV over the
 * dense cache in the HF layer invoked from ModelShard.forward (worker/distributed/model_shard.py:205-222).
 *
 *   out           : [B, Hq, D] contiguous
 *   q             : [B, Hq, D]; row b starts at q + b * q_row_stride (elements)
 *   block_tables  : int32[B, table_stride]; entry j of row b is the pool slot of the j-th block of sequence b
 *   context_lens  : int32[B]; tokens to attend to (including the one just appended); 0 gives a zero row
 *   max_context_len : host-side upper bound of context_lens, used only to size the split-K partition
 *   num_splits    : partitions of the context per (sequence, kv head); 0 = let the library choose
 *                   (b200_paged_decode_attn_plan); > 1 needs `workspace`
 *   workspace     : >= b200_paged_decode_attn_workspace(B, Hq, D, num_splits) bytes of device scratch
 * kv head j serves q heads j*g .. j*g+g-1 with g = Hq / Hkv (HF repeat_kv order).
 */
int b200_paged_decode_attn(void* out, const void* q, int64_t q_row_stride,
                           const void* k_pool, const void* v_pool,
                           const int32_t* block_tables, int table_stride, const int32_t* context_lens,
                           int B, int Hq, int Hkv, int D, int block_size, int64_t max_blocks,
                           float scale, int max_context_len, int num_splits,
                           void* workspace, size_t workspace_bytes,
                           int dtype, unsigned flags, void* stream);
/* Bytes of scratch for `num_splits` partitions (0 for num_splits <= 1): B * Hq * num_splits * (D + 2) floats of
 * partials (un-normalised outputs, running max and sum) behind a header of 4 * B * Hq bytes rounded up to 256 — the
 * arrival counters of the shared-prefix entry points (B200_ATTN_WS_COUNTERS_ZERO above); the plain entry point
 * ignores the header. */
size_t b200_paged_decode_attn_workspace(int B, int Hq, int D, int num_splits);
int b200_paged_decode_attn_plan(int B, int Hq, int Hkv, int D, int block_size, int max_context_len, int dtype);

/*
 * One decode step of a layer's attention in ONE launch: RoPE on q and on the new token's k, the append of that k and
 * v into the pools, and the paged attention over the result — b200_rope_kv_append followed by
 * b200_paged_decode_attn, fused (same rounding points, bit-identical output and pool contents).  q is read
 * un-rotated and left untouched; context_lens[b] already counts the new token, whose slot is the last one of the
 * sequence (block_tables[b][(ctx-1)/16], row (ctx-1)%16).  Replaces apply_rotary_pos_emb + torch.cat + eager
 * attention of the HF layer called from ModelShard.forward (worker/distributed/model_shard.py:209-215).
 * Shapes: f16/bf16, block_size 16, head_dim 64/128, group <= 8, 16-byte aligned rows; anything else returns
 * B200_ESHAPE and the two separate entry points apply.
 */
int b200_rope_append_decode_attn(void* out, const void* q, int64_t q_row_stride,
                                 const void* k_new, int64_t k_row_stride, const void* v_new, int64_t v_row_stride,
                                 const void* cos_table, const void* sin_table,
                                 void* k_pool, void* v_pool,
                                 const int32_t* block_tables, int table_stride, const int32_t* context_lens,
                                 int B, int Hq, int Hkv, int D, int block_size, int64_t max_blocks,
                                 float scale, int max_context_len, int num_splits,
                                 void* workspace, size_t workspace_bytes,
                                 int dtype, unsigned flags, void* stream);

/*
 * Shared-prefix ("cascade") variant of b200_paged_decode_attn: when block_tables[b][0:prefix_blocks] are the same
 * slots for every sequence (a shared system prompt — the blocks the reference's allocator would hold with
 * CacheBlock.ref_count > 1, worker/distributed/kv_cache.py:63-72, README "RadixAttention" prefix sharing), those
 * blocks are streamed from HBM once for the whole batch instead of once per sequence, then each sequence's own
 * blocks are processed as usual and the partials are merged.  Requirements: every context_lens[b] >=
 * prefix_blocks * block_size; f16/bf16, block_size 16, head_dim 64/128, group <= 8 (else B200_ESHAPE — use the
 * plain entry point, which gives the same result).  A workspace is always needed:
 * b200_paged_decode_attn_workspace(B, Hq, D, prefix_splits + suffix_splits); 0 for a split count = library's choice
 * (b200_paged_decode_attn_prefix_plan).  prefix_blocks = 0 forwards to b200_paged_decode_attn.
 */
int b200_paged_decode_attn_prefix(void* out, const void* q, int64_t q_row_stride,
                                  const void* k_pool, const void* v_pool,
                                  const int32_t* block_tables, int table_stride, const int32_t* context_lens,
                                  int B, int Hq, int Hkv, int D, int block_size, int64_t max_blocks,
                                  float scale, int max_context_len, int prefix_blocks,
                                  int prefix_splits, int suffix_splits,
                                  void* workspace, size_t workspace_bytes,
                                  int dtype, unsigned flags, void* stream);
int b200_paged_decode_attn_prefix_plan(int B, int Hq, int Hkv, int D, int block_size, int prefix_blocks,
                                       int max_context_len, int dtype, int* prefix_splits, int* suffix_splits);

/*
 * b200_paged_decode_attn_prefix with the decode step's RoPE and KV append inside the same launch (what
 * b200_rope_append_decode_attn is to b200_paged_decode_attn): q arrives un-rotated; the new token's k row is rotated
 * and written, with its v row, into the sequence's last block before that block is attended to.  Every sequence must
 * have at least one token after the shared prefix (context_lens[b] > prefix_blocks * block_size): shared blocks are
 * never written.  Output and pool contents are bit-identical to b200_rope_kv_append followed by
 * b200_paged_decode_attn_prefix.
 */
int b200_rope_append_decode_attn_prefix(void* out, const void* q, int64_t q_row_stride, const void* k_new,
                                        int64_t k_row_stride, const void* v_new, int64_t v_row_stride,
                                        const void* cos_table, const void* sin_table, void* k_pool, void* v_pool,
                                        const int32_t* block_tables, int table_stride, const int32_t* context_lens,
                                        int B, int Hq, int Hkv, int D, int block_size, int64_t max_blocks, float scale,
                                        int max_context_len, int prefix_blocks, int prefix_splits, int suffix_splits,
                                        void* workspace, size_t workspace_bytes, int dtype, unsigned flags,
                                        void* stream);

/*
 * Paged decode attention for callers that cannot allocate device memory.  No counterpart in the reference: it
 * exists for hosts that are handed pointers and nothing else — a plugin inside another runtime's serving loop
 * where cudaMalloc is off limits because it stalls the whole device, or a binding without an allocator behind it
 * (cgo / JNI / bare ctypes with no torch underneath).
 *
 * Same arguments and result as b200_paged_decode_attn, minus the workspace: the split-K partials go to an array
 * of B200_STATIC_SCRATCH_BYTES that is part of the library image (each GPU that loads the library holds it until
 * the process exits).  The call itself never calls cudaMalloc / cudaFree or any allocating routine.
 *   - Sizes whose partials do not fit (b200_paged_decode_attn_workspace(B, Hq, D, num_splits) >
 *     B200_STATIC_SCRATCH_BYTES) are rejected with B200_ENOSPC; pass a smaller num_splits (1 needs no scratch).
 *   - The first call on each device looks up the array's device address (with that device current).  Make that
 *     first call before starting any stream capture on the device.
 *   - All calls on a device share the one array, which keeps the previous call's partials.  Nothing in it needs
 *     to be zero beforehand (every call rewrites all it reads), but two calls that overlap on the GPU — different
 *     streams, or different host threads feeding different streams — race on it.  Keep them on one stream, or
 *     order them with events.
 */
#define B200_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
int b200_paged_decode_attn_static(void* out, const void* q, int64_t q_row_stride,
                                  const void* k_pool, const void* v_pool,
                                  const int32_t* block_tables, int table_stride, const int32_t* context_lens,
                                  int B, int Hq, int Hkv, int D, int block_size, int64_t max_blocks,
                                  float scale, int max_context_len, int num_splits,
                                  int dtype, unsigned flags, void* stream);

/*
 * Wire-format pack / unpack for hand-offs that leave the box (no NVLink): bfloat16 pools <-> a float16 packed
 * buffer [n_layers, 2, n_blocks, Hkv, block_size, D], i.e. exactly the bytes TensorSerializer.serialize emits for a
 * bf16 block (common/serialization.py:73-76: bf16 is cast to fp16 before tobytes(); :143-155 widens it back), in
 * the per-layer keys/values order of KVCacheLayer (proto/inference.proto:110-127).  Gather + cast in one pass on
 * the GPU instead of per-block .to(float16).cpu().  Lossy outside the fp16 range exactly like the reference.
 */
int b200_kv_pack_layers_wire(void* packed_f16, const void* const* k_pools, const void* const* v_pools, int n_layers,
                             const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                             int64_t max_blocks, void* stream);
int b200_kv_unpack_layers_wire(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed_f16,
                               const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                               int64_t max_blocks, void* stream);

/*
 * Block-to-block move between two sets of pools, all layers in one launch: block src_table[j] of source layer l
 * is copied to block dst_table[j] of destination layer l.  The destination pool pointers may be peer memory —
 * another GPU's pools mapped over NVLink — in which case the prefill GPU writes a sequence's KV straight into the
 * blocks the decode GPU allocated, with no packed staging buffer and no unpack.  This is the fused form of the
 * hand-off of InferenceServicer.TransferKVCache (worker/distributed/grpc_server.py:190-235); with local pointers
 * it is a block copy (copy-on-write of shared blocks, CacheBlock.add_ref — kv_cache.py:67-68).
 */
int b200_kv_move_blocks_layers(void* const* dst_k_pools, void* const* dst_v_pools,
                               const void* const* src_k_pools, const void* const* src_v_pools, int n_layers,
                               const int32_t* src_table, const int32_t* dst_table, int n_blocks,
                               int Hkv, int D, int block_size, int64_t src_max_blocks, int64_t dst_max_blocks,
                               int dtype, void* stream);

/*
 * Fused element-wise pieces of the decoder layer invoked from ModelShard.forward
 * (worker/distributed/model_shard.py:209-215; arithmetic of the HF Llama layer).  Each reproduces the rounding
 * points of the eager torch recipe in the storage dtype; see csrc/b200_layer_ops.cu.
 *
 * b200_add_rmsnorm : h = x + delta (delta may be NULL), resid_out = h (may be NULL), out = weight * rmsnorm(h).
 *                    x: [rows, hidden] with row stride x_stride, delta likewise with d_stride; out/resid_out
 *                    contiguous.  Replaces the residual add + LlamaRMSNorm.
 * b200_rope_kv_append : rotate-half RoPE on q (in place) and k with cos/sin tables [n_tok, D] (storage dtype),
 *                    then scatter the rotated k and v into the pools through slot_mapping (as b200_kv_append).
 *                    Replaces apply_rotary_pos_emb + the dense cache growth.
 * b200_silu_mul    : out[r, c] = silu(gate[r, c]) * up[r, c]; gate/up row strides free, out contiguous.
 *                    Replaces act_fn(gate_proj(x)) * up_proj(x) of LlamaMLP.
 */
int b200_add_rmsnorm(void* out, void* resid_out, const void* x, const void* delta, const void* weight, float eps,
                     int rows, int hidden, int64_t x_stride, int64_t d_stride, int dtype, void* stream);
int b200_rope_kv_append(void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride,
                        const void* v, int64_t v_row_stride, const void* cos_table, const void* sin_table,
                        void* k_pool, void* v_pool, const int64_t* slot_mapping, int n_tok,
                        int Hq, int Hkv, int D, int block_size, int64_t max_blocks, int dtype, void* stream);
/* Prompt-pass form of b200_rope_kv_append: k is rotated IN PLACE as well (the causal attention among the new
 * tokens needs the rotated k densely), besides being scattered into the pool.  Same arithmetic, same arguments. */
int b200_rope_kv_append_prompt(void* q, int64_t q_row_stride, void* k, int64_t k_row_stride,
                               const void* v, int64_t v_row_stride, const void* cos_table, const void* sin_table,
                               void* k_pool, void* v_pool, const int64_t* slot_mapping, int n_tok,
                               int Hq, int Hkv, int D, int block_size, int64_t max_blocks, int dtype, void* stream);
int b200_silu_mul(void* out, const void* gate, const void* up, int64_t rows, int cols,
                  int64_t gate_row_stride, int64_t up_row_stride, int dtype, void* stream);
/* Greedy token: out[r] = argmax over x[r, 0..cols) — largest value, smallest index among equals, NaN largest, i.e.
 * torch.argmax(logits, dim=-1) of the sampling step (the reference: worker/engines/speculative.py:449).  x rows are
 * row_stride elements apart; out is int64[rows]. */
int b200_argmax_rows(int64_t* out, const void* x, int64_t rows, int64_t cols, int64_t row_stride, int dtype,
                     void* stream);
/* Greedy token when the vocabulary projection is split by rows of lm_head across ranks (ModelShard.get_logits,
 * worker/distributed/model_shard.py:230-246, computed as N column slices of the logits): keys[r] encodes the best
 * (value, col_offset + column) of this rank's slice x[r, 0..cols) so that the element-wise signed MAX of the N ranks'
 * key vectors — one int64 all-reduce — is the key of torch.argmax over the whole row (largest value, smallest
 * global column among equals, NaN largest).  b200_argmax_keys_decode turns reduced keys into token ids.
 * col_offset + cols must stay below 2^32. */
int b200_argmax_rows_keys(int64_t* keys, const void* x, int64_t rows, int64_t cols, int64_t row_stride,
                          int64_t col_offset, int dtype, void* stream);
int b200_argmax_keys_decode(int64_t* out, const int64_t* keys, int64_t rows, void* stream);

/*
 * Gather to dense: the first n_tokens tokens of B sequences, read through their block-table rows, written as
 *   k_out, v_out : [B, Hkv, n_tokens, D]
 * in ONE pass.  Replaces what a dense-cache layer does when a prompt continues after cached tokens — the
 * torch.cat((past_k, k), dim=2) of the HF layer ModelShard.forward calls (worker/distributed/model_shard.py:209-215) —
 * for paged pools: the prompt pass appends its new tokens first (b200_rope_kv_append_prompt), then gathers prefix and
 * new tokens together for the dense causal attention.  Bit-exact copy.
 *   block_tables : int32[B, max_blocks_per_seq], the first ceil(n_tokens / block_size) entries of each row are read
 */
int b200_kv_gather_dense(void* k_out, void* v_out, const void* k_pool, const void* v_pool,
                         const int32_t* block_tables, int max_blocks_per_seq, int B, int n_tokens, int Hkv, int D,
                         int block_size, int64_t max_blocks, int dtype, void* stream);

/*
 * KV block pack / unpack for the prefill -> decode hand-off.  Replaces the per-block, per-layer
 * TensorSerializer.serialize -> wire -> deserialize -> block.keys.copy_ round trip of
 * InferenceServicer.TransferKVCache (worker/distributed/grpc_server.py:190-235) and
 * DistributedKVCacheManager._promote_to_gpu (worker/distributed/kv_cache.py:447-462).
 *
 *   packed      : [2, n_blocks, Hkv, block_size, D] — all K blocks in table order, then all V blocks
 *   block_table : int32[n_blocks] pool slots (source slots for pack, destination slots for unpack)
 * Equivalent to key_pool.index_select(0, table) / key_pool.index_copy_(0, table, ...), bit-exact.
 */
int b200_kv_pack(void* packed, const void* k_pool, const void* v_pool, const int32_t* block_table,
                 int n_blocks, int Hkv, int D, int block_size, int64_t max_blocks, int dtype, void* stream);
int b200_kv_unpack(void* k_pool, void* v_pool, const void* packed, const int32_t* block_table,
                   int n_blocks, int Hkv, int D, int block_size, int64_t max_blocks, int dtype, void* stream);

/*
 * Same for all layers of a shard in one launch.  `k_pools` / `v_pools` are DEVICE arrays of n_layers pool
 * pointers; packed is [n_layers, 2, n_blocks, Hkv, block_size, D]; one block table serves every layer
 * (KVCachePool.allocate_sequence yields the same slots in every layer — kv_cache.py:283-313).
 */
int b200_kv_pack_layers(void* packed, const void* const* k_pools, const void* const* v_pools, int n_layers,
                        const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                        int64_t max_blocks, int dtype, void* stream);
int b200_kv_unpack_layers(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed,
                          const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                          int64_t max_blocks, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_KV_H_ */
