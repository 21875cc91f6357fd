// Internal interface between the C ABI (b200_abi.cu) and the attention kernels.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

namespace b200 {

struct AttnArgs {
  void* out;                    // [B, Hq, D]
  const void* q;                // [B, Hq, D], row stride q_row_stride elements
  int64_t q_row_stride;
  const void* k_pool;           // [max_blocks, Hkv, block_size, D]
  const void* v_pool;
  const int32_t* block_tables;  // [B, table_stride]
  int table_stride;
  const int32_t* context_lens;  // [B]
  int B, Hq, Hkv, D, block_size;
  int64_t max_blocks;
  float scale;
  int blocks_per_split, num_splits;
  float* ws_acc;                // [B, Hq, num_splits, D] un-normalised partial outputs
  float* ws_ml;                 // [B, Hq, num_splits, 2] running max and sum
  int round_scores;
  int* ws_counters;             // [B * Hq] arrival counters in front of the partials (shared-prefix entry points)
  bool ws_counters_zero;        // the caller vouches that they read zero (B200_ATTN_WS_COUNTERS_ZERO)
  // fused RoPE + KV append (b200_rope_append_decode_attn); all null / zero for the plain entry points
  const void* k_new;   // [B, Hkv, D] un-rotated k of the new token, row stride k_new_stride
  int64_t k_new_stride;
  const void* v_new;
  int64_t v_new_stride;
  const void* cos_t;   // [B, D]
  const void* sin_t;
};

// shape-generic CUDA-core variant (b200_attn_generic.cu)
int attn_generic(const AttnArgs& a, int dtype, cudaStream_t s);
// TMA + tensor-core variant (b200_attn_tma.cu); attn_tma_supported says whether it covers the shape
bool attn_tma_supported(const AttnArgs& a, int dtype);
int attn_tma(const AttnArgs& a, int dtype, cudaStream_t s);
// shared-prefix cascade (b200_attn_prefix.cu): prefix blocks once for all sequences, then per-sequence suffixes
bool attn_prefix_supported(const AttnArgs& a, int dtype);
int attn_prefix(const AttnArgs& a, int dtype, int prefix_blocks, int prefix_splits, int suffix_splits, cudaStream_t s);
void attn_prefix_plan(int B, int Hq, int Hkv, int block_size, int prefix_blocks, int max_context_len,
                      int* prefix_splits, int* suffix_splits);
// merge of split-K partials
int attn_combine(void* out, const float* ws_acc, const float* ws_ml, int rows, int D, int num_splits, int dtype,
                 cudaStream_t stream);

// kv ops (b200_kv_ops.cu)
int kv_append(const void* k_new, const void* v_new, int64_t k_row_stride, int64_t v_row_stride, void* k_pool,
              void* v_pool, const int64_t* slot_mapping, int n_tok, int Hkv, int D, int block_size,
              int64_t max_blocks, int dtype, cudaStream_t stream);
int kv_gather_dense(void* k_out, void* v_out, const void* k_pool, const void* v_pool, const int32_t* tables,
                    int table_stride, int B, int n_tokens, int Hkv, int D, int bs, int64_t max_blocks, int dtype,
                    cudaStream_t s);
int kv_pack(void* packed, const void* k_pool, const void* v_pool, const int32_t* table, int n_blocks, int Hkv, int D,
            int bs, int64_t max_blocks, int dtype, cudaStream_t s);
int kv_unpack(void* k_pool, void* v_pool, const void* packed, const int32_t* table, int n_blocks, int Hkv, int D,
              int bs, int64_t max_blocks, int dtype, cudaStream_t s);
int kv_pack_layers(void* packed, const void* const* k_pools, const void* const* v_pools, int n_layers,
                   const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, int dtype,
                   cudaStream_t s);
int kv_unpack_layers(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed,
                     const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, int dtype,
                     cudaStream_t s);

int kv_move_blocks_layers(void* const* dst_k_pools, void* const* dst_v_pools, const void* const* src_k_pools,
                          const void* const* src_v_pools, int n_layers, const int32_t* src_table,
                          const int32_t* dst_table, int n_blocks, int Hkv, int D, int bs, int64_t src_max_blocks,
                          int64_t dst_max_blocks, int dtype, cudaStream_t s);
int kv_pack_layers_wire(void* packed_f16, const void* const* k_pools, const void* const* v_pools, int n_layers,
                        const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, cudaStream_t s);
int kv_unpack_layers_wire(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed_f16,
                          const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks,
                          cudaStream_t s);

// fused layer pieces (b200_layer_ops.cu)
int add_rmsnorm(void* out, void* resid_out, const void* x, const void* delta, const void* weight, float eps, int rows,
                int hidden, int64_t x_stride, int64_t d_stride, int dtype, cudaStream_t s);
int rope_kv_append(void* q, int64_t q_stride, const void* k, int64_t k_stride, void* k_rot, const void* v,
                   int64_t v_stride,
                   const void* cos_t, const void* sin_t, void* k_pool, void* v_pool, const int64_t* slot_mapping,
                   int n_tok, int Hq, int Hkv, int D, int block_size, int64_t max_blocks, int dtype, cudaStream_t s);
int argmax_rows(int64_t* out, const void* x, int64_t rows, int64_t cols, int64_t row_stride, int dtype, cudaStream_t s);
int argmax_rows_keys(int64_t* keys, const void* x, int64_t rows, int64_t cols, int64_t row_stride, int64_t col_offset,
                     int dtype, cudaStream_t s);
int argmax_keys_decode(int64_t* out, const int64_t* keys, int64_t rows, cudaStream_t s);
int silu_mul(void* out, const void* gate, const void* up, int64_t rows, int cols, int64_t g_stride, int64_t u_stride,
             int dtype, cudaStream_t s);

}  // namespace b200
