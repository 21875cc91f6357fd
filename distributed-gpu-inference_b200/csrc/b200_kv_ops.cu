// Paged KV append (scatter through a slot mapping) and KV block pack / unpack (gather / scatter of whole
// blocks) for sm_100a.  All three are pure byte movers bounded by HBM bandwidth: every access is a 16-byte
// vector, consecutive threads touch consecutive vectors of one contiguous run (a token-head row of D elements
// for append, a whole [Hkv, block_size, D] block for pack/unpack), and each thread keeps several independent
// loads in flight before its first store.
#include <algorithm>

#include "b200_attn.cuh"
#include "b200_common.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------------------------
// append
// ---------------------------------------------------------------------------------------------------------
// A CTA takes TOKS consecutive tokens.  The first TOKS threads turn the tokens' slots into pool row offsets once (the
// only divisions of the kernel; the flat-index version of round 1 spent two 64-bit divisions per 16-byte vector and
// was issue-bound at 62 % of the copy peak, profiles/r02_ncu_append.md); after that thread x owns 16-byte vector x of
// the K|V row pair of every token of the CTA — TOKS independent loads in flight before the first store, consecutive
// threads on consecutive vectors of one contiguous head row.
template <int TOKS>
__global__ void __launch_bounds__(256)
kv_append_vec16_kernel(const char* __restrict__ k_new, const char* __restrict__ v_new,
                       int64_t k_row_bytes, int64_t v_row_bytes,
                       char* __restrict__ k_pool, char* __restrict__ v_pool,
                       const int64_t* __restrict__ slot_mapping,
                       int n_tok, int Hkv, int block_size, int64_t max_blocks,
                       int vec_per_head /* D * elem_size / 16 */) {
  __shared__ int64_t row0_s[TOKS];  // (block * Hkv * block_size + offset in block) of the token, -1 = skip
  const int tok0 = blockIdx.x * TOKS;
  if (threadIdx.x < TOKS) {
    const int tok = tok0 + threadIdx.x;
    int64_t row0 = -1;
    if (tok < n_tok) {
      const int64_t slot = slot_mapping[tok];
      if (slot >= 0) {
        const int64_t blk = slot / block_size;
        if (blk < max_blocks) row0 = blk * Hkv * block_size + (slot - blk * block_size);
      }
    }
    row0_s[threadIdx.x] = row0;
  }
  __syncthreads();
  const int vec_per_tok = Hkv * vec_per_head;  // vectors of one K (or V) row
  const int head_bytes = vec_per_head * 16;
  for (int x = threadIdx.x; x < 2 * vec_per_tok; x += blockDim.x) {
    const bool is_v = x >= vec_per_tok;
    const int r = is_v ? x - vec_per_tok : x;
    const int h = r / vec_per_head;
    const int col = (r - h * vec_per_head) * 16;  // byte offset inside the head row
    const char* src = (is_v ? v_new : k_new) + h * head_bytes + col;
    const int64_t src_row = is_v ? v_row_bytes : k_row_bytes;
    char* pool = (is_v ? v_pool : k_pool) + col;
    int4 val[TOKS];
#pragma unroll
    for (int u = 0; u < TOKS; ++u)
      if (row0_s[u] >= 0) val[u] = ld_stream16(src + (tok0 + u) * src_row);
#pragma unroll
    for (int u = 0; u < TOKS; ++u)
      if (row0_s[u] >= 0)
        *reinterpret_cast<int4*>(pool + (row0_s[u] + static_cast<int64_t>(h) * block_size) * head_bytes) = val[u];
  }
}

// Fallback for rows that are not 16-byte aligned (odd head_dim, fp32 with D % 4 != 0, unaligned views).
template <typename T>
__global__ void kv_append_scalar_kernel(const T* __restrict__ k_new, const T* __restrict__ v_new,
                                        int64_t k_row_stride, int64_t v_row_stride,
                                        T* __restrict__ k_pool, T* __restrict__ v_pool,
                                        const int64_t* __restrict__ slot_mapping,
                                        int n_tok, int Hkv, int D, int block_size, int64_t max_blocks) {
  const int64_t per_tok = static_cast<int64_t>(Hkv) * D;
  const int64_t total = 2 * static_cast<int64_t>(n_tok) * per_tok;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const bool is_v = i >= n_tok * per_tok;
    const int64_t j = is_v ? i - n_tok * per_tok : i;
    const int64_t tok = j / per_tok;
    const int64_t rem = j - tok * per_tok;
    const int64_t h = rem / D, d = rem - h * D;
    const int64_t slot = slot_mapping[tok];
    if (slot < 0) continue;
    const int64_t blk = slot / block_size, off = slot - blk * block_size;
    if (blk >= max_blocks) continue;
    const int64_t dst = ((blk * Hkv + h) * block_size + off) * D + d;
    if (is_v)
      v_pool[dst] = v_new[tok * v_row_stride + h * D + d];
    else
      k_pool[dst] = k_new[tok * k_row_stride + h * D + d];
  }
}

int kv_append(const void* k_new, const void* v_new, int64_t k_row_stride, int64_t v_row_stride, void* k_pool,
              void* v_pool, const int64_t* slot_mapping, int n_tok, int Hkv, int D, int block_size,
              int64_t max_blocks, int dtype, cudaStream_t stream) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "b200_kv_append: unsupported dtype %d", dtype);
  if (n_tok < 0 || Hkv <= 0 || D <= 0 || block_size <= 0 || max_blocks <= 0)
    return fail(B200_EINVAL, "b200_kv_append: bad sizes n_tok=%d Hkv=%d D=%d block_size=%d max_blocks=%lld", n_tok,
                Hkv, D, block_size, static_cast<long long>(max_blocks));
  if (n_tok == 0) return 0;
  if (!k_new || !v_new || !k_pool || !v_pool || !slot_mapping)
    return fail(B200_EINVAL, "b200_kv_append: null pointer");
  if (k_row_stride < static_cast<int64_t>(Hkv) * D || v_row_stride < static_cast<int64_t>(Hkv) * D)
    return fail(B200_EINVAL, "b200_kv_append: row stride smaller than Hkv*D");

  const bool vec_ok = (D * es) % 16 == 0 && (k_row_stride * es) % 16 == 0 && (v_row_stride * es) % 16 == 0 &&
                      aligned16(k_new) && aligned16(v_new) && aligned16(k_pool) && aligned16(v_pool);
  if (vec_ok) {
    constexpr int TOKS = 8;
    const int vec_per_head = D * es / 16;
    const int threads = std::min(256, std::max(32, (2 * Hkv * vec_per_head + 31) / 32 * 32));
    const int grid = (n_tok + TOKS - 1) / TOKS;
    kv_append_vec16_kernel<TOKS><<<static_cast<unsigned>(grid), threads, 0, stream>>>(
        static_cast<const char*>(k_new), static_cast<const char*>(v_new), k_row_stride * es, v_row_stride * es,
        static_cast<char*>(k_pool), static_cast<char*>(v_pool), slot_mapping, n_tok, Hkv, block_size, max_blocks,
        vec_per_head);
    return check_launch("b200_kv_append");
  }
  const int64_t total = 2ll * n_tok * Hkv * D;
  const int threads = 256;
  const int grid = static_cast<int>(std::min<int64_t>((total + threads - 1) / threads, 148 * 16));
#define LAUNCH_SCALAR(T)                                                                                        \
  kv_append_scalar_kernel<T><<<grid, threads, 0, stream>>>(                                                     \
      static_cast<const T*>(k_new), static_cast<const T*>(v_new), k_row_stride, v_row_stride,                   \
      static_cast<T*>(k_pool), static_cast<T*>(v_pool), slot_mapping, n_tok, Hkv, D, block_size, max_blocks)
  if (es == 2)
    LAUNCH_SCALAR(uint16_t);
  else
    LAUNCH_SCALAR(uint32_t);
#undef LAUNCH_SCALAR
  return check_launch("b200_kv_append(scalar)");
}

// ---------------------------------------------------------------------------------------------------------
// pack / unpack
// ---------------------------------------------------------------------------------------------------------
struct PoolSet {
  const void* const* k_pools;  // device array [n_layers] or nullptr
  const void* const* v_pools;
  const void* k_single;  // used when the arrays are null (single-layer entry points)
  const void* v_single;
};

// Work item = (layer, K|V, table entry); each item is one contiguous run of `vec_per_block` 16-byte vectors.
// A CTA of 256 threads walks chunks of 256*UNROLL vectors; the grid is flat over all chunks so the hardware
// scheduler balances it (grid sizes here are thousands of CTAs, i.e. many waves of 148 SMs).
template <bool PACK, int UNROLL>
__global__ void __launch_bounds__(256)
kv_block_copy_kernel(char* __restrict__ packed, PoolSet pools, const int32_t* __restrict__ block_table,
                     int n_layers, int n_blocks, int64_t max_blocks, int vec_per_block, int chunks_per_block) {
  const int64_t chunk = blockIdx.x;
  const int64_t item = chunk / chunks_per_block;  // (layer * 2 + kv) * n_blocks + j
  const int c = static_cast<int>(chunk - item * chunks_per_block);
  const int j = static_cast<int>(item % n_blocks);
  const int lk = static_cast<int>(item / n_blocks);
  const int layer = lk >> 1;
  const bool is_v = lk & 1;
  const int32_t slot = block_table[j];
  if (slot < 0 || slot >= max_blocks) return;

  const void* pool_base;
  if (pools.k_pools != nullptr)
    pool_base = is_v ? pools.v_pools[layer] : pools.k_pools[layer];
  else
    pool_base = is_v ? pools.v_single : pools.k_single;
  const int64_t block_bytes = static_cast<int64_t>(vec_per_block) * 16;
  char* pool_ptr = const_cast<char*>(static_cast<const char*>(pool_base)) + slot * block_bytes;
  char* pack_ptr = packed + item * block_bytes;
  const char* src = PACK ? pool_ptr : pack_ptr;
  char* dst = PACK ? pack_ptr : pool_ptr;

  const int v0 = c * (256 * UNROLL) + threadIdx.x;
  int4 val[UNROLL];
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    const int v = v0 + u * 256;
    if (v < vec_per_block) val[u] = ld_stream16(src + static_cast<int64_t>(v) * 16);
  }
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    const int v = v0 + u * 256;
    if (v < vec_per_block) st_stream16(dst + static_cast<int64_t>(v) * 16, val[u]);
  }
}

template <bool PACK>
int kv_block_copy(void* packed, PoolSet pools, int n_layers, const int32_t* block_table, int n_blocks, int Hkv, int D,
                  int block_size, int64_t max_blocks, int dtype, cudaStream_t stream, const char* name) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "%s: unsupported dtype %d", name, dtype);
  if (n_blocks < 0 || n_layers <= 0 || Hkv <= 0 || D <= 0 || block_size <= 0 || max_blocks <= 0)
    return fail(B200_EINVAL, "%s: bad sizes", name);
  if (n_blocks == 0) return 0;
  if (!packed || !block_table) return fail(B200_EINVAL, "%s: null pointer", name);
  if (pools.k_pools == nullptr && (!pools.k_single || !pools.v_single))
    return fail(B200_EINVAL, "%s: null pool pointer", name);
  const int64_t block_bytes = static_cast<int64_t>(Hkv) * block_size * D * es;
  if (block_bytes % 16 != 0 || !aligned16(packed) ||
      (pools.k_pools == nullptr && (!aligned16(pools.k_single) || !aligned16(pools.v_single))))
    return fail(B200_ESHAPE, "%s: block of %lld bytes / pointers must be 16-byte aligned", name,
                static_cast<long long>(block_bytes));
  if (block_bytes / 16 > INT32_MAX) return fail(B200_ESHAPE, "%s: block too large", name);

  constexpr int UNROLL = 8;  // 8 x 16 B in flight per thread, 32 KiB per CTA pass
  const int vec_per_block = static_cast<int>(block_bytes / 16);
  const int chunks_per_block = (vec_per_block + 256 * UNROLL - 1) / (256 * UNROLL);
  const int64_t grid = 2ll * n_layers * n_blocks * chunks_per_block;
  if (grid > INT32_MAX) return fail(B200_ESHAPE, "%s: too many chunks", name);
  kv_block_copy_kernel<PACK, UNROLL><<<static_cast<unsigned>(grid), 256, 0, stream>>>(
      static_cast<char*>(packed), pools, block_table, n_layers, n_blocks, max_blocks, vec_per_block, chunks_per_block);
  return check_launch(name);
}

// Block-to-block copy between two sets of pools: block src_table[j] of every source layer goes to block
// dst_table[j] of the matching destination layer.  The destination pointers may be PEER memory (another GPU's
// pools mapped over NVLink): the prefill rank then pushes a sequence straight into the decode rank's blocks with
// 16-byte stores over NVSwitch and no staging buffer on either side.
template <int UNROLL>
__global__ void __launch_bounds__(256)
kv_block_move_kernel(PoolSet dst, PoolSet src, const int32_t* __restrict__ src_table,
                     const int32_t* __restrict__ dst_table, int n_blocks, int64_t src_max_blocks,
                     int64_t dst_max_blocks, int vec_per_block, int chunks_per_block) {
  const int64_t chunk = blockIdx.x;
  const int64_t item = chunk / chunks_per_block;
  const int c = static_cast<int>(chunk - item * chunks_per_block);
  const int j = static_cast<int>(item % n_blocks);
  const int lk = static_cast<int>(item / n_blocks);
  const int layer = lk >> 1;
  const bool is_v = lk & 1;
  const int32_t s_slot = src_table[j], d_slot = dst_table[j];
  if (s_slot < 0 || s_slot >= src_max_blocks || d_slot < 0 || d_slot >= dst_max_blocks) return;
  const int64_t block_bytes = static_cast<int64_t>(vec_per_block) * 16;
  const char* sp = static_cast<const char*>(is_v ? src.v_pools[layer] : src.k_pools[layer]) + s_slot * block_bytes;
  char* dp = const_cast<char*>(static_cast<const char*>(is_v ? dst.v_pools[layer] : dst.k_pools[layer])) + d_slot * block_bytes;
  const int v0 = c * (256 * UNROLL) + threadIdx.x;
  int4 val[UNROLL];
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    const int v = v0 + u * 256;
    if (v < vec_per_block) val[u] = ld_stream16(sp + static_cast<int64_t>(v) * 16);
  }
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    const int v = v0 + u * 256;
    if (v < vec_per_block) st_stream16(dp + static_cast<int64_t>(v) * 16, val[u]);
  }
}

int kv_move_blocks_layers(void* const* dst_k_pools, void* const* dst_v_pools, const void* const* src_k_pools,
                          const void* const* src_v_pools, int n_layers, const int32_t* src_table,
                          const int32_t* dst_table, int n_blocks, int Hkv, int D, int bs, int64_t src_max_blocks,
                          int64_t dst_max_blocks, int dtype, cudaStream_t s) {
  const char* name = "b200_kv_move_blocks_layers";
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "%s: unsupported dtype %d", name, dtype);
  if (n_blocks < 0 || n_layers <= 0 || Hkv <= 0 || D <= 0 || bs <= 0 || src_max_blocks <= 0 || dst_max_blocks <= 0)
    return fail(B200_EINVAL, "%s: bad sizes", name);
  if (n_blocks == 0) return 0;
  if (!dst_k_pools || !dst_v_pools || !src_k_pools || !src_v_pools || !src_table || !dst_table)
    return fail(B200_EINVAL, "%s: null pointer", name);
  const int64_t block_bytes = static_cast<int64_t>(Hkv) * bs * D * es;
  if (block_bytes % 16 != 0 || block_bytes / 16 > INT32_MAX) return fail(B200_ESHAPE, "%s: block size", name);
  constexpr int UNROLL = 8;
  const int vec_per_block = static_cast<int>(block_bytes / 16);
  const int chunks_per_block = (vec_per_block + 256 * UNROLL - 1) / (256 * UNROLL);
  const int64_t grid = 2ll * n_layers * n_blocks * chunks_per_block;
  if (grid > INT32_MAX) return fail(B200_ESHAPE, "%s: too many chunks", name);
  PoolSet d{const_cast<const void* const*>(dst_k_pools), const_cast<const void* const*>(dst_v_pools), nullptr, nullptr};
  PoolSet sr{src_k_pools, src_v_pools, nullptr, nullptr};
  kv_block_move_kernel<UNROLL><<<static_cast<unsigned>(grid), 256, 0, s>>>(d, sr, src_table, dst_table, n_blocks,
                                                                         src_max_blocks, dst_max_blocks,
                                                                         vec_per_block, chunks_per_block);
  return check_launch(name);
}

// Wire-format pack / unpack: as pack / unpack, but the packed side is float16 while the pools are bfloat16 — the
// byte stream the reference's host serializer produces for a bf16 tensor (common/serialization.py:73-76 casts
// bf16 to fp16 before tobytes(); :143-155 widens it back on receipt).  One pass over the data instead of
// pack -> .to(float16) (-> D2H): each thread converts one 16-byte vector of 8 elements.  Same rounding as torch's
// cast: bf16 -> fp32 is exact, fp32 -> fp16 rounds to nearest even, overflow gives inf, fp16 -> bf16 rounds to
// nearest even.
template <bool PACK, int UNROLL>
__global__ void __launch_bounds__(256)
kv_block_wire_kernel(__half* __restrict__ packed, PoolSet pools, const int32_t* __restrict__ block_table, int n_blocks,
                     int64_t max_blocks, int vec_per_block, int chunks_per_block) {
  const int64_t chunk = blockIdx.x;
  const int64_t item = chunk / chunks_per_block;
  const int c = static_cast<int>(chunk - item * chunks_per_block);
  const int j = static_cast<int>(item % n_blocks);
  const int lk = static_cast<int>(item / n_blocks);
  const int layer = lk >> 1;
  const bool is_v = lk & 1;
  const int32_t slot = block_table[j];
  if (slot < 0 || slot >= max_blocks) return;
  const void* pool_base = is_v ? pools.v_pools[layer] : pools.k_pools[layer];
  const int64_t block_elems = static_cast<int64_t>(vec_per_block) * 8;
  __nv_bfloat16* pool_ptr = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(pool_base)) + slot * block_elems;
  __half* pack_ptr = packed + item * block_elems;
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    const int v = c * (256 * UNROLL) + threadIdx.x + u * 256;
    if (v >= vec_per_block) continue;
    union { int4 raw; __nv_bfloat16 b[8]; __half h[8]; } in, out;
    if (PACK) {
      in.raw = ld_stream16(pool_ptr + static_cast<int64_t>(v) * 8);
#pragma unroll
      for (int e = 0; e < 8; ++e) out.h[e] = __float2half_rn(__bfloat162float(in.b[e]));
      st_stream16(pack_ptr + static_cast<int64_t>(v) * 8, out.raw);
    } else {
      in.raw = ld_stream16(pack_ptr + static_cast<int64_t>(v) * 8);
#pragma unroll
      for (int e = 0; e < 8; ++e) out.b[e] = __float2bfloat16_rn(__half2float(in.h[e]));
      st_stream16(pool_ptr + static_cast<int64_t>(v) * 8, out.raw);
    }
  }
}

template <bool PACK>
int kv_wire(void* packed_f16, const void* const* k_pools, const void* const* v_pools, int n_layers,
            const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, cudaStream_t s,
            const char* name) {
  if (n_blocks < 0 || n_layers <= 0 || Hkv <= 0 || D <= 0 || bs <= 0 || max_blocks <= 0)
    return fail(B200_EINVAL, "%s: bad sizes", name);
  if (n_blocks == 0) return 0;
  if (!packed_f16 || !k_pools || !v_pools || !table) return fail(B200_EINVAL, "%s: null pointer", name);
  const int64_t block_elems = static_cast<int64_t>(Hkv) * bs * D;
  if (block_elems % 8 != 0 || !aligned16(packed_f16)) return fail(B200_ESHAPE, "%s: block size / alignment", name);
  constexpr int UNROLL = 8;
  const int vec_per_block = static_cast<int>(block_elems / 8);
  const int chunks_per_block = (vec_per_block + 256 * UNROLL - 1) / (256 * UNROLL);
  const int64_t grid = 2ll * n_layers * n_blocks * chunks_per_block;
  if (grid > INT32_MAX) return fail(B200_ESHAPE, "%s: too many chunks", name);
  PoolSet p{k_pools, v_pools, nullptr, nullptr};
  kv_block_wire_kernel<PACK, UNROLL><<<static_cast<unsigned>(grid), 256, 0, s>>>(
      static_cast<__half*>(packed_f16), p, table, n_blocks, max_blocks, vec_per_block, chunks_per_block);
  return check_launch(name);
}

int kv_pack_layers_wire(void* packed_f16, const void* const* k_pools, const void* const* v_pools, int n_layers,
                        const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, cudaStream_t s) {
  return kv_wire<true>(packed_f16, k_pools, v_pools, n_layers, table, n_blocks, Hkv, D, bs, max_blocks, s,
                       "b200_kv_pack_layers_wire");
}
int kv_unpack_layers_wire(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed_f16,
                          const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks,
                          cudaStream_t s) {
  return kv_wire<false>(const_cast<void*>(packed_f16), const_cast<const void* const*>(k_pools),
                        const_cast<const void* const*>(v_pools), n_layers, table, n_blocks, Hkv, D, bs, max_blocks, s,
                        "b200_kv_unpack_layers_wire");
}

// ---------------------------------------------------------------------------------------------------------
// gather to dense: the first n_tokens tokens of every sequence, read through its block-table row, laid out
// [B, Hkv, n_tokens, D] — what a prompt pass that continues a cached prefix hands to the dense attention.
// Work item = (sequence, table entry, K|V): Hkv runs of `rows x D` elements, each contiguous on both sides.
// ---------------------------------------------------------------------------------------------------------
template <int HU>
__global__ void __launch_bounds__(256)
kv_gather_dense_kernel(char* __restrict__ k_out, char* __restrict__ v_out, const char* __restrict__ k_pool,
                       const char* __restrict__ v_pool, const int32_t* __restrict__ tables, int table_stride,
                       int n_blk, int n_tokens, int Hkv, int block_size, int row_bytes, int64_t max_blocks) {
  const int item = blockIdx.x;  // (b * n_blk + j) * 2 + kv
  const bool is_v = item & 1;
  const int bj = item >> 1;
  const int j = bj % n_blk, b = bj / n_blk;
  const int32_t slot = tables[static_cast<int64_t>(b) * table_stride + j];
  if (slot < 0 || slot >= max_blocks) return;
  const int rows = min(block_size, n_tokens - j * block_size);
  const int vec_valid = rows * (row_bytes / 16);
  const int64_t head_bytes = static_cast<int64_t>(block_size) * row_bytes;  // one head of one block in the pool
  const int64_t seq_bytes = static_cast<int64_t>(n_tokens) * row_bytes;      // one head of one sequence in the output
  const char* src = (is_v ? v_pool : k_pool) + static_cast<int64_t>(slot) * Hkv * head_bytes;
  char* dst = (is_v ? v_out : k_out) + static_cast<int64_t>(b) * Hkv * seq_bytes +
              static_cast<int64_t>(j) * block_size * row_bytes;
  for (int h0 = 0; h0 < Hkv; h0 += HU) {
    for (int x = threadIdx.x; x < vec_valid; x += 256) {
      int4 val[HU];
#pragma unroll
      for (int u = 0; u < HU; ++u)
        if (h0 + u < Hkv) val[u] = ld_stream16(src + (h0 + u) * head_bytes + static_cast<int64_t>(x) * 16);
#pragma unroll
      for (int u = 0; u < HU; ++u)
        if (h0 + u < Hkv) st_stream16(dst + (h0 + u) * seq_bytes + static_cast<int64_t>(x) * 16, val[u]);
    }
  }
}

int kv_gather_dense(void* k_out, void* v_out, const void* k_pool, const void* v_pool, const int32_t* tables,
                    int table_stride, int B, int n_tokens, int Hkv, int D, int bs, int64_t max_blocks, int dtype,
                    cudaStream_t s) {
  const char* name = "b200_kv_gather_dense";
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "%s: unsupported dtype %d", name, dtype);
  if (B < 0 || n_tokens < 0 || Hkv <= 0 || D <= 0 || bs <= 0 || max_blocks <= 0)
    return fail(B200_EINVAL, "%s: bad sizes", name);
  if (B == 0 || n_tokens == 0) return 0;
  if (!k_out || !v_out || !k_pool || !v_pool || !tables) return fail(B200_EINVAL, "%s: null pointer", name);
  const int n_blk = (n_tokens + bs - 1) / bs;
  if (table_stride < n_blk) return fail(B200_ESHAPE, "%s: %d tokens need %d table entries per row, rows hold %d", name,
                                        n_tokens, n_blk, table_stride);
  const int64_t row_bytes = static_cast<int64_t>(D) * es;
  if (row_bytes % 16 != 0 || row_bytes > INT32_MAX || !aligned16(k_out) || !aligned16(v_out) || !aligned16(k_pool) ||
      !aligned16(v_pool))
    return fail(B200_ESHAPE, "%s: rows of %lld bytes / pointers must be 16-byte aligned", name,
                static_cast<long long>(row_bytes));
  const int64_t grid = 2ll * B * n_blk;
  if (grid > INT32_MAX) return fail(B200_ESHAPE, "%s: too many blocks", name);
  kv_gather_dense_kernel<8><<<static_cast<unsigned>(grid), 256, 0, s>>>(
      static_cast<char*>(k_out), static_cast<char*>(v_out), static_cast<const char*>(k_pool),
      static_cast<const char*>(v_pool), tables, table_stride, n_blk, n_tokens, Hkv, bs, static_cast<int>(row_bytes),
      max_blocks);
  return check_launch(name);
}

int kv_pack(void* packed, const void* k_pool, const void* v_pool, const int32_t* table, int n_blocks, int Hkv, int D,
            int bs, int64_t max_blocks, int dtype, cudaStream_t s) {
  PoolSet p{nullptr, nullptr, k_pool, v_pool};
  return kv_block_copy<true>(packed, p, 1, table, n_blocks, Hkv, D, bs, max_blocks, dtype, s, "b200_kv_pack");
}
int kv_unpack(void* k_pool, void* v_pool, const void* packed, const int32_t* table, int n_blocks, int Hkv, int D,
              int bs, int64_t max_blocks, int dtype, cudaStream_t s) {
  PoolSet p{nullptr, nullptr, k_pool, v_pool};
  return kv_block_copy<false>(const_cast<void*>(packed), p, 1, table, n_blocks, Hkv, D, bs, max_blocks, dtype, s,
                              "b200_kv_unpack");
}
int kv_pack_layers(void* packed, const void* const* k_pools, const void* const* v_pools, int n_layers,
                   const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, int dtype,
                   cudaStream_t s) {
  if (!k_pools || !v_pools) return fail(B200_EINVAL, "b200_kv_pack_layers: null pool array");
  PoolSet p{k_pools, v_pools, nullptr, nullptr};
  return kv_block_copy<true>(packed, p, n_layers, table, n_blocks, Hkv, D, bs, max_blocks, dtype, s,
                             "b200_kv_pack_layers");
}
int kv_unpack_layers(void* const* k_pools, void* const* v_pools, int n_layers, const void* packed,
                     const int32_t* table, int n_blocks, int Hkv, int D, int bs, int64_t max_blocks, int dtype,
                     cudaStream_t s) {
  if (!k_pools || !v_pools) return fail(B200_EINVAL, "b200_kv_unpack_layers: null pool array");
  PoolSet p{const_cast<const void* const*>(k_pools), const_cast<const void* const*>(v_pools), nullptr, nullptr};
  return kv_block_copy<false>(const_cast<void*>(packed), p, n_layers, table, n_blocks, Hkv, D, bs, max_blocks, dtype,
                              s, "b200_kv_unpack_layers");
}

}  // namespace b200
