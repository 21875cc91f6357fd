// extern "C" surface of libb200_kv.so — argument checking, split-K planning and dispatch to the kernels.
// See include/b200_kv.h for the contract of every entry point.
#include <algorithm>

#include "b200_attn.cuh"
#include "b200_common.cuh"

namespace b200 {

char* last_error_buf() {
  static thread_local char buf[kErrBufLen] = {0};
  return buf;
}

namespace {

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// How many partitions of the context each (sequence, kv head) pair gets.  Measured on B200 (tools/kbench.py,
// profiles/r01_kbench.md): a second launch (the combine) and the partials cost 10-25 % on these 50-170 us
// kernels, so one partition wins whenever the pairs alone put a CTA on most SMs — 512 pairs (config 2): 98 % of
// the copy peak with 1 split, 83 % with 3; 128 pairs (config 4, micro-batch 16): 86 % with 1, 76 % with 2.  Below
// that the context is cut to roughly 3.5 CTAs per SM (64 pairs, config 5: 93-95 % for 4..12 splits), keeping every
// partition at least 256 tokens long so the per-CTA pipeline ramp stays amortised.
int plan_splits(int B, int Hkv, int block_size, int max_context_len) {
  const int n_blocks = std::max(1, ceil_div(std::max(1, max_context_len), block_size));
  const int pairs = std::max(1, B * Hkv);
  const int min_blocks = std::max(1, 256 / block_size);
  const int max_by_len = std::max(1, n_blocks / min_blocks);
  if (pairs >= 120) {
    // One CTA per pair is right when the pairs put (at most) one CTA on every SM, or so many that the block scheduler
    // balances them by itself.  In between every CTA is resident from the start, some SMs hold two and the others one,
    // and the launch takes as long as the doubly loaded SMs: 160 pairs (batch 20, 8 kv heads, context 4k) ran at 0.58
    // of the copy peak, 192 at 0.68, 320 at 0.82.  Cutting the context so that there are roughly 700 CTAs (a bit over
    // two waves of the 296 slots) lets the scheduler even them out, and with the combine launched as a programmatic
    // dependent the second launch costs little: 0.80-0.83 / 0.88 / 0.94 for those three, 256 pairs 0.90 -> 0.92,
    // 384 pairs 0.94 -> 0.97 (tools/kbench.py --cases grid:BxCTX, profiles/r02_kbench_midsize_grids.txt).
    if (pairs <= kNumSMs || pairs >= 500) return 1;
    const int s = std::min(6, std::max(2, ceil_div(700, pairs)));
    return std::max(1, std::min(s, max_by_len));
  }
  int splits = std::max(1, ceil_div(kNumSMs * 7 / 2, pairs));
  // One or two long sequences (pairs <= 16): exactly one wave of CTAs (2 per SM) beats 1.75 waves of shorter ones —
  // batch 1 at 16k / 32k / 64k / 128k context: 37 splits 25.7 / 32.5 / 52.7 / 92.6 us vs 26.7 / 35.7 / 56.6 / 94.5 us;
  // batch 2 at 32k: 18 splits 54.3 vs 55.3 us.  Not so from 32 pairs up (batch 4 at 8k: 9 splits 35.3 us, 17: 33.2 us).
  if (pairs <= 16) splits = std::min(splits, std::max(1, 2 * kNumSMs / pairs));
  splits = std::min(splits, max_by_len);
  return std::min(splits, 128);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_abi_version(void) { return B200_ABI_VERSION; }

const char* b200_last_error(void) { return last_error_buf(); }

int b200_kv_append(const void* k_new, const void* v_new, int64_t k_row_stride, int64_t v_row_stride, void* k_pool,
                   void* v_pool, const int64_t* slot_mapping, int n_tok, int Hkv, int D, int block_size,
                   int64_t max_blocks, int dtype, void* stream) {
  return kv_append(k_new, v_new, k_row_stride, v_row_stride, k_pool, v_pool, slot_mapping, n_tok, Hkv, D, block_size,
                   max_blocks, dtype, static_cast<cudaStream_t>(stream));
}

namespace {
// bytes in front of the partials that hold one arrival counter per (sequence, q head) row
size_t counter_bytes(int B, int Hq) { return (static_cast<size_t>(B) * Hq * sizeof(int) + 255) / 256 * 256; }
}  // namespace

size_t b200_paged_decode_attn_workspace(int B, int Hq, int D, int num_splits) {
  if (B <= 0 || Hq <= 0 || D <= 0 || num_splits <= 1) return 0;
  return counter_bytes(B, Hq) + static_cast<size_t>(B) * Hq * num_splits * (static_cast<size_t>(D) + 2) * sizeof(float);
}

int b200_paged_decode_attn_plan(int B, int Hq, int Hkv, int D, int block_size, int max_context_len, int dtype) {
  (void)Hq;
  (void)D;
  (void)dtype;
  if (B <= 0 || Hkv <= 0 || block_size <= 0) return 1;
  return plan_splits(B, Hkv, block_size, max_context_len);
}

int b200_paged_decode_attn(void* out, const void* q, int64_t q_row_stride, const void* k_pool, const void* v_pool,
                           const int32_t* block_tables, int table_stride, const int32_t* context_lens, int B, int Hq,
                           int Hkv, int D, int block_size, int64_t max_blocks, float scale, int max_context_len,
                           int num_splits, void* workspace, size_t workspace_bytes, int dtype, unsigned flags,
                           void* stream) {
  if (dtype_size(dtype) == 0) return fail(B200_EINVAL, "b200_paged_decode_attn: unsupported dtype %d", dtype);
  if (B < 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || block_size <= 0 || max_blocks <= 0 || table_stride < 0)
    return fail(B200_EINVAL, "b200_paged_decode_attn: bad sizes B=%d Hq=%d Hkv=%d D=%d block_size=%d", B, Hq, Hkv, D,
                block_size);
  if (Hq % Hkv != 0) return fail(B200_ESHAPE, "b200_paged_decode_attn: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
  if (B == 0) return 0;
  if (!out || !q || !k_pool || !v_pool || !block_tables || !context_lens)
    return fail(B200_EINVAL, "b200_paged_decode_attn: null pointer");
  if (q_row_stride < static_cast<int64_t>(Hq) * D)
    return fail(B200_EINVAL, "b200_paged_decode_attn: q_row_stride smaller than Hq*D");
  if (max_context_len < 0) return fail(B200_EINVAL, "b200_paged_decode_attn: negative max_context_len");
  if (num_splits < 0) return fail(B200_EINVAL, "b200_paged_decode_attn: negative num_splits");
  if (num_splits == 0) num_splits = workspace ? plan_splits(B, Hkv, block_size, max_context_len) : 1;
  const int n_blocks = std::max(1, ceil_div(std::max(1, max_context_len), block_size));
  num_splits = std::min(num_splits, n_blocks);

  AttnArgs a{};
  a.out = out;
  a.q = q;
  a.q_row_stride = q_row_stride;
  a.k_pool = k_pool;
  a.v_pool = v_pool;
  a.block_tables = block_tables;
  a.table_stride = table_stride;
  a.context_lens = context_lens;
  a.B = B;
  a.Hq = Hq;
  a.Hkv = Hkv;
  a.D = D;
  a.block_size = block_size;
  a.max_blocks = max_blocks;
  a.scale = scale;
  a.num_splits = num_splits;
  a.blocks_per_split = ceil_div(n_blocks, num_splits);
  a.round_scores = (flags & B200_ATTN_ROUND_SCORES) ? 1 : 0;
  if (num_splits > 1) {
    const size_t need = b200_paged_decode_attn_workspace(B, Hq, D, num_splits);
    if (!workspace || workspace_bytes < need)
      return fail(B200_ENOSPC, "b200_paged_decode_attn: workspace of %zu bytes needed for %d splits, got %zu", need,
                  num_splits, workspace_bytes);
    a.ws_acc = static_cast<float*>(workspace);
    a.ws_ml = a.ws_acc + static_cast<size_t>(B) * Hq * num_splits * D;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc;
  if (!(flags & B200_ATTN_FORCE_GENERIC) && attn_tma_supported(a, dtype))
    rc = attn_tma(a, dtype, s);
  else
    rc = attn_generic(a, dtype, s);
  if (rc != 0) return rc;
  if (num_splits > 1) return attn_combine(out, a.ws_acc, a.ws_ml, B * Hq, D, num_splits, dtype, s);
  return 0;
}

int b200_rope_append_decode_attn(void* out, const void* q, int64_t q_row_stride, const void* k_new,
                                 int64_t k_row_stride, const void* v_new, int64_t v_row_stride, const void* cos_table,
                                 const void* sin_table, void* k_pool, void* v_pool, const int32_t* block_tables,
                                 int table_stride, const int32_t* context_lens, int B, int Hq, int Hkv, int D,
                                 int block_size, int64_t max_blocks, float scale, int max_context_len, int num_splits,
                                 void* workspace, size_t workspace_bytes, int dtype, unsigned flags, void* stream) {
  const char* name = "b200_rope_append_decode_attn";
  if (dtype_size(dtype) == 0) return fail(B200_EINVAL, "%s: unsupported dtype %d", name, dtype);
  if (B < 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || block_size <= 0 || max_blocks <= 0 || Hq % Hkv != 0 || num_splits < 0)
    return fail(B200_EINVAL, "%s: bad sizes", name);
  if (B == 0) return 0;
  if (!out || !q || !k_new || !v_new || !cos_table || !sin_table || !k_pool || !v_pool || !block_tables || !context_lens)
    return fail(B200_EINVAL, "%s: null pointer", name);
  if (q_row_stride < static_cast<int64_t>(Hq) * D || k_row_stride < static_cast<int64_t>(Hkv) * D ||
      v_row_stride < static_cast<int64_t>(Hkv) * D)
    return fail(B200_EINVAL, "%s: row stride smaller than heads * D", name);
  if (num_splits == 0) num_splits = workspace ? plan_splits(B, Hkv, block_size, max_context_len) : 1;
  const int n_blocks = std::max(1, ceil_div(std::max(1, max_context_len), block_size));
  num_splits = std::min(num_splits, n_blocks);
  AttnArgs a{};
  a.out = out;
  a.q = q;
  a.q_row_stride = q_row_stride;
  a.k_pool = k_pool;
  a.v_pool = v_pool;
  a.block_tables = block_tables;
  a.table_stride = table_stride;
  a.context_lens = context_lens;
  a.B = B;
  a.Hq = Hq;
  a.Hkv = Hkv;
  a.D = D;
  a.block_size = block_size;
  a.max_blocks = max_blocks;
  a.scale = scale;
  a.num_splits = num_splits;
  a.blocks_per_split = ceil_div(n_blocks, num_splits);
  a.round_scores = (flags & B200_ATTN_ROUND_SCORES) ? 1 : 0;
  a.k_new = k_new;
  a.k_new_stride = k_row_stride;
  a.v_new = v_new;
  a.v_new_stride = v_row_stride;
  a.cos_t = cos_table;
  a.sin_t = sin_table;
  const int es = dtype_size(dtype);
  if (!attn_tma_supported(a, dtype) || (k_row_stride * es) % 16 || (v_row_stride * es) % 16 || !aligned16(k_new) ||
      !aligned16(v_new) || !aligned16(cos_table) || !aligned16(sin_table))
    return fail(B200_ESHAPE, "%s: needs f16/bf16, block_size 16, head_dim 64/128, group <= 8, 16-byte aligned rows; "
                "use b200_rope_kv_append + b200_paged_decode_attn for other shapes", name);
  if (num_splits > 1) {
    const size_t need = b200_paged_decode_attn_workspace(B, Hq, D, num_splits);
    if (!workspace || workspace_bytes < need)
      return fail(B200_ENOSPC, "%s: workspace of %zu bytes needed for %d splits, got %zu", name, need, num_splits,
                  workspace_bytes);
    a.ws_acc = static_cast<float*>(workspace);
    a.ws_ml = a.ws_acc + static_cast<size_t>(B) * Hq * num_splits * D;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int rc = attn_tma(a, dtype, s);
  if (rc != 0) return rc;
  if (num_splits > 1) return attn_combine(out, a.ws_acc, a.ws_ml, B * Hq, D, num_splits, dtype, s);
  return 0;
}

int b200_paged_decode_attn_prefix_plan(int B, int Hq, int Hkv, int D, int block_size, int prefix_blocks,
                                       int max_context_len, int dtype, int* prefix_splits, int* suffix_splits) {
  (void)D;
  (void)dtype;
  if (!prefix_splits || !suffix_splits) return fail(B200_EINVAL, "b200_paged_decode_attn_prefix_plan: null pointer");
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || block_size <= 0 || prefix_blocks < 0)
    return fail(B200_EINVAL, "b200_paged_decode_attn_prefix_plan: bad sizes");
  attn_prefix_plan(B, Hq, Hkv, block_size, prefix_blocks, max_context_len, prefix_splits, suffix_splits);
  return 0;
}

namespace {
struct NewToken {  // operands of the fused decode step; all null = attention only
  const void* k_new = nullptr;
  int64_t k_stride = 0;
  const void* v_new = nullptr;
  int64_t v_stride = 0;
  const void* cos_t = nullptr;
  const void* sin_t = nullptr;
};

int attn_prefix_entry(const char* name, const NewToken& nt, void* out, const void* q, int64_t q_row_stride,
                      const void* k_pool, const void* v_pool, const int32_t* block_tables, int table_stride,
                      const int32_t* context_lens, int B, int Hq, int Hkv, int D, int block_size, int64_t max_blocks,
                      float scale, int max_context_len, int prefix_blocks, int prefix_splits, int suffix_splits,
                      void* workspace, size_t workspace_bytes, int dtype, unsigned flags, void* stream) {
  if (dtype_size(dtype) == 0) return fail(B200_EINVAL, "%s: unsupported dtype %d", name, dtype);
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || block_size <= 0 || max_blocks <= 0 || Hq % Hkv != 0)
    return fail(B200_EINVAL, "%s: bad sizes", name);
  if (!out || !q || !k_pool || !v_pool || !block_tables || !context_lens || !workspace)
    return fail(B200_EINVAL, "%s: null pointer (a workspace is always needed)", name);
  if (prefix_blocks > table_stride || static_cast<int64_t>(prefix_blocks) * block_size > max_context_len)
    return fail(B200_EINVAL, "%s: prefix of %d blocks exceeds the table / max_context_len", name, prefix_blocks);
  int ps = prefix_splits, ss = suffix_splits;
  if (ps == 0 || ss == 0) {
    int aps, ass;
    attn_prefix_plan(B, Hq, Hkv, block_size, prefix_blocks, max_context_len, &aps, &ass);
    if (ps == 0) ps = aps;
    if (ss == 0) ss = ass;
  }
  ps = std::min(ps, prefix_blocks);
  const int suffix_blocks = std::max(1, ceil_div(std::max(1, max_context_len - prefix_blocks * block_size), block_size));
  ss = std::min(ss, suffix_blocks);
  const int total = ps + ss;
  const size_t need = b200_paged_decode_attn_workspace(B, Hq, D, total);
  if (workspace_bytes < need)
    return fail(B200_ENOSPC, "%s: workspace of %zu bytes needed for %d + %d splits, got %zu", name, need, ps, ss,
                workspace_bytes);
  AttnArgs a{};
  a.out = out;
  a.q = q;
  a.q_row_stride = q_row_stride;
  a.k_pool = k_pool;
  a.v_pool = v_pool;
  a.block_tables = block_tables;
  a.table_stride = table_stride;
  a.context_lens = context_lens;
  a.B = B;
  a.Hq = Hq;
  a.Hkv = Hkv;
  a.D = D;
  a.block_size = block_size;
  a.max_blocks = max_blocks;
  a.scale = scale;
  a.num_splits = ss;
  a.blocks_per_split = ceil_div(suffix_blocks, ss);
  a.round_scores = (flags & B200_ATTN_ROUND_SCORES) ? 1 : 0;
  a.ws_counters = static_cast<int*>(workspace);
  a.ws_counters_zero = (flags & B200_ATTN_WS_COUNTERS_ZERO) != 0;
  a.ws_acc = reinterpret_cast<float*>(static_cast<unsigned char*>(workspace) + counter_bytes(B, Hq));
  a.ws_ml = a.ws_acc + static_cast<size_t>(B) * Hq * total * D;
  a.k_new = nt.k_new;
  a.k_new_stride = nt.k_stride;
  a.v_new = nt.v_new;
  a.v_new_stride = nt.v_stride;
  a.cos_t = nt.cos_t;
  a.sin_t = nt.sin_t;
  if (!attn_prefix_supported(a, dtype))
    return fail(B200_ESHAPE, "%s: needs f16/bf16, block_size 16, head_dim 64/128, group <= 8", name);
  return attn_prefix(a, dtype, prefix_blocks, ps, ss, static_cast<cudaStream_t>(stream));
}
}  // namespace

int b200_paged_decode_attn_prefix(void* out, const void* q, int64_t q_row_stride, const void* k_pool,
                                  const void* v_pool, const int32_t* block_tables, int table_stride,
                                  const int32_t* context_lens, int B, int Hq, int Hkv, int D, int block_size,
                                  int64_t max_blocks, float scale, int max_context_len, int prefix_blocks,
                                  int prefix_splits, int suffix_splits, void* workspace, size_t workspace_bytes,
                                  int dtype, unsigned flags, void* stream) {
  const char* name = "b200_paged_decode_attn_prefix";
  if (prefix_blocks < 0 || prefix_splits < 0 || suffix_splits < 0) return fail(B200_EINVAL, "%s: negative count", name);
  if (prefix_blocks == 0)
    return b200_paged_decode_attn(out, q, q_row_stride, k_pool, v_pool, block_tables, table_stride, context_lens, B, Hq,
                                  Hkv, D, block_size, max_blocks, scale, max_context_len, suffix_splits, workspace,
                                  workspace_bytes, dtype, flags, stream);
  return attn_prefix_entry(name, NewToken{}, out, q, q_row_stride, k_pool, v_pool, block_tables, table_stride,
                           context_lens, B, Hq, Hkv, D, block_size, max_blocks, scale, max_context_len, prefix_blocks,
                           prefix_splits, suffix_splits, workspace, workspace_bytes, dtype, flags, stream);
}

int b200_rope_append_decode_attn_prefix(void* out, const void* q, int64_t q_row_stride, const void* k_new,
                                        int64_t k_row_stride, const void* v_new, int64_t v_row_stride,
                                        const void* cos_table, const void* sin_table, void* k_pool, void* v_pool,
                                        const int32_t* block_tables, int table_stride, const int32_t* context_lens, int B,
                                        int Hq, int Hkv, int D, int block_size, int64_t max_blocks, float scale,
                                        int max_context_len, int prefix_blocks, int prefix_splits, int suffix_splits,
                                        void* workspace, size_t workspace_bytes, int dtype, unsigned flags, void* stream) {
  const char* name = "b200_rope_append_decode_attn_prefix";
  if (prefix_blocks < 0 || prefix_splits < 0 || suffix_splits < 0) return fail(B200_EINVAL, "%s: negative count", name);
  if (prefix_blocks == 0)
    return b200_rope_append_decode_attn(out, q, q_row_stride, k_new, k_row_stride, v_new, v_row_stride, cos_table,
                                        sin_table, k_pool, v_pool, block_tables, table_stride, context_lens, B, Hq, Hkv, D,
                                        block_size, max_blocks, scale, max_context_len, suffix_splits, workspace,
                                        workspace_bytes, dtype, flags, stream);
  if (!k_new || !v_new || !cos_table || !sin_table) return fail(B200_EINVAL, "%s: null pointer", name);
  if (Hkv > 0 && D > 0 && (k_row_stride < static_cast<int64_t>(Hkv) * D || v_row_stride < static_cast<int64_t>(Hkv) * D))
    return fail(B200_EINVAL, "%s: row stride smaller than heads * D", name);
  NewToken nt;
  nt.k_new = k_new;
  nt.k_stride = k_row_stride;
  nt.v_new = v_new;
  nt.v_stride = v_row_stride;
  nt.cos_t = cos_table;
  nt.sin_t = sin_table;
  return attn_prefix_entry(name, nt, out, q, q_row_stride, k_pool, v_pool, block_tables, table_stride, context_lens, B,
                           Hq, Hkv, D, block_size, max_blocks, scale, max_context_len, prefix_blocks, prefix_splits,
                           suffix_splits, workspace, workspace_bytes, dtype, flags, stream);
}

int b200_kv_gather_dense(void* k_out, void* v_out, const void* k_pool, const void* v_pool,
                         const int32_t* block_tables, int max_blocks_per_seq, int B, int n_tokens, int Hkv, int D,
                         int block_size, int64_t max_blocks, int dtype, void* stream) {
  return kv_gather_dense(k_out, v_out, k_pool, v_pool, block_tables, max_blocks_per_seq, B, n_tokens, Hkv, D,
                         block_size, max_blocks, dtype, static_cast<cudaStream_t>(stream));
}

int b200_kv_pack(void* packed, const void* k_pool, const void* v_pool, const int32_t* block_table, int n_blocks,
                 int Hkv, int D, int block_size, int64_t max_blocks, int dtype, void* stream) {
  return kv_pack(packed, k_pool, v_pool, block_table, n_blocks, Hkv, D, block_size, max_blocks, dtype,
                 static_cast<cudaStream_t>(stream));
}

int b200_kv_unpack(void* k_pool, void* v_pool, const void* packed, const int32_t* block_table, int n_blocks, int Hkv,
                   int D, int block_size, int64_t max_blocks, int dtype, void* stream) {
  return kv_unpack(k_pool, v_pool, packed, block_table, n_blocks, Hkv, D, block_size, max_blocks, dtype,
                   static_cast<cudaStream_t>(stream));
}

int b200_kv_pack_layers(void* packed, const void* const* k_pools, const void* const* v_pools, int n_layers,
                        const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size, int64_t max_blocks,
                        int dtype, void* stream) {
  return kv_pack_layers(packed, k_pools, v_pools, n_layers, block_table, n_blocks, Hkv, D, block_size, max_blocks,
                        dtype, static_cast<cudaStream_t>(stream));
}

int b200_kv_unpack_layers(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed,
                          const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                          int64_t max_blocks, int dtype, void* stream) {
  return kv_unpack_layers(k_pools, v_pools, n_layers, packed, block_table, n_blocks, Hkv, D, block_size, max_blocks,
                          dtype, static_cast<cudaStream_t>(stream));
}

int b200_kv_move_blocks_layers(void* const* dst_k_pools, void* const* dst_v_pools, const void* const* src_k_pools,
                               const void* const* src_v_pools, int n_layers, const int32_t* src_table,
                               const int32_t* dst_table, int n_blocks, int Hkv, int D, int block_size,
                               int64_t src_max_blocks, int64_t dst_max_blocks, int dtype, void* stream) {
  return kv_move_blocks_layers(dst_k_pools, dst_v_pools, src_k_pools, src_v_pools, n_layers, src_table, dst_table,
                               n_blocks, Hkv, D, block_size, src_max_blocks, dst_max_blocks, dtype,
                               static_cast<cudaStream_t>(stream));
}

int b200_kv_pack_layers_wire(void* packed_f16, const void* const* k_pools, const void* const* v_pools, int n_layers,
                             const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                             int64_t max_blocks, void* stream) {
  return kv_pack_layers_wire(packed_f16, k_pools, v_pools, n_layers, block_table, n_blocks, Hkv, D, block_size,
                             max_blocks, static_cast<cudaStream_t>(stream));
}

int b200_kv_unpack_layers_wire(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed_f16,
                               const int32_t* block_table, int n_blocks, int Hkv, int D, int block_size,
                               int64_t max_blocks, void* stream) {
  return kv_unpack_layers_wire(k_pools, v_pools, n_layers, packed_f16, block_table, n_blocks, Hkv, D, block_size,
                               max_blocks, static_cast<cudaStream_t>(stream));
}

int b200_add_rmsnorm(void* out, void* resid_out, const void* x, const void* delta, const void* weight, float eps,
                     int rows, int hidden, int64_t x_stride, int64_t d_stride, int dtype, void* stream) {
  return add_rmsnorm(out, resid_out, x, delta, weight, eps, rows, hidden, x_stride, d_stride, dtype,
                     static_cast<cudaStream_t>(stream));
}

int b200_rope_kv_append(void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride, const void* v,
                        int64_t v_row_stride, const void* cos_table, const void* sin_table, void* k_pool, void* v_pool,
                        const int64_t* slot_mapping, int n_tok, int Hq, int Hkv, int D, int block_size,
                        int64_t max_blocks, int dtype, void* stream) {
  return rope_kv_append(q, q_row_stride, k, k_row_stride, nullptr, v, v_row_stride, cos_table, sin_table, k_pool,
                        v_pool, slot_mapping, n_tok, Hq, Hkv, D, block_size, max_blocks, dtype,
                        static_cast<cudaStream_t>(stream));
}

int b200_rope_kv_append_prompt(void* q, int64_t q_row_stride, void* k, int64_t k_row_stride, const void* v,
                               int64_t v_row_stride, const void* cos_table, const void* sin_table, void* k_pool,
                               void* v_pool, const int64_t* slot_mapping, int n_tok, int Hq, int Hkv, int D,
                               int block_size, int64_t max_blocks, int dtype, void* stream) {
  return rope_kv_append(q, q_row_stride, k, k_row_stride, k, v, v_row_stride, cos_table, sin_table, k_pool, v_pool,
                        slot_mapping, n_tok, Hq, Hkv, D, block_size, max_blocks, dtype,
                        static_cast<cudaStream_t>(stream));
}

int b200_argmax_rows(int64_t* out, const void* x, int64_t rows, int64_t cols, int64_t row_stride, int dtype,
                     void* stream) {
  return argmax_rows(out, x, rows, cols, row_stride, dtype, static_cast<cudaStream_t>(stream));
}

int b200_argmax_rows_keys(int64_t* keys, const void* x, int64_t rows, int64_t cols, int64_t row_stride,
                          int64_t col_offset, int dtype, void* stream) {
  return argmax_rows_keys(keys, x, rows, cols, row_stride, col_offset, dtype, static_cast<cudaStream_t>(stream));
}

int b200_argmax_keys_decode(int64_t* out, const int64_t* keys, int64_t rows, void* stream) {
  return argmax_keys_decode(out, keys, rows, static_cast<cudaStream_t>(stream));
}

int b200_silu_mul(void* out, const void* gate, const void* up, int64_t rows, int cols, int64_t gate_row_stride,
                  int64_t up_row_stride, int dtype, void* stream) {
  return silu_mul(out, gate, up, rows, cols, gate_row_stride, up_row_stride, dtype, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
