// Paged GQA decode attention for sm_100a — TMA-staged KV tiles + tensor-core inner products.
//
// Shapes covered: f16 / bf16 storage, block_size 16, head_dim 64 or 128, group size (Hq / Hkv) <= 8.
// Everything else goes to b200_attn_generic.cu.
//
// Work decomposition
//   grid = (num_splits, Hkv, B): one CTA per (context partition, kv head, sequence); 4 warps per CTA.
//   A KV block of the reference layout is, per kv head, one contiguous [16 tokens, D] tile for K and one for V
//   (4 KiB each at D = 128).  Warp w of a CTA owns tiles w, w+4, ... of the CTA's block range and runs its own
//   R-stage ring: lane 0 issues the TMA copies (cp.async.bulk.tensor.2d, 128-byte swizzle) of a K and a V tile
//   into the ring and arms the stage's mbarrier with the byte count; the whole warp waits on that mbarrier,
//   consumes the stage, and lane 0 refills it with the tile R steps ahead.  No block-wide barrier is taken
//   until the final merge, so a warp never waits for another warp's loads, and with 2 CTAs per SM up to
//   2 * 4 * R * 8 KiB of KV is in flight per SM.
//
// Math per 16-token tile (transposed formulation so that the 8-wide N dimension of mma.m16n8k16 holds the
// q heads of the group and nothing is padded along M):
//   S^T[tok, g]  = K[tok, :] . Q[g, :]        A = K tile (ldmatrix), B = Q (registers, loaded once)
//   online softmax over tok per g: running max via warp shuffles (xor 4, 8, 16), exp2 with folded scale
//   O^T[d, g]   += V^T[d, tok] P^T[tok, g]    A = V tile (ldmatrix.trans), B = P (movmatrix.trans of S^T frags)
// Partials (m, l, O) of the 4 warps are merged through shared memory; with num_splits > 1 the CTA stores its
// partial to the workspace for attn_combine_kernel.
#include <type_traits>
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "b200_attn.cuh"
#include "b200_common.cuh"
#include "b200_tma_common.cuh"

namespace b200 {

namespace tma {

namespace {
// ---- host side: tensor maps -------------------------------------------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess ||
        qr != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t rows;
  int D;
  int dtype;
  bool operator==(const MapKey& o) const { return ptr == o.ptr && rows == o.rows && D == o.D && dtype == o.dtype; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (std::hash<uint64_t>()(k.rows) * 1315423911u) ^ (k.D * 2654435761u) ^
           static_cast<size_t>(k.dtype);
  }
};

}  // namespace

// Pools are long-lived (one pair per layer), so descriptors are cached by (pointer, rows, D, dtype).
int get_tensor_map(const void* pool, uint64_t rows, int D, int dtype, CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  const MapKey key{pool, rows, D, dtype};
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(B200_EINVAL, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[2] = {static_cast<cuuint64_t>(D), rows};
  const cuuint64_t strides[1] = {static_cast<cuuint64_t>(D) * 2};
  const cuuint32_t box[2] = {64, static_cast<cuuint32_t>(kTok)};
  const cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, dtype == B200_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                   const_cast<void*>(pool), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B200_EINVAL, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  {
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 8192) cache.clear();
    cache.emplace(key, m);
  }
  *out = m;
  return 0;
}


bool tensor_maps_available() { return encode_fn() != nullptr; }

}  // namespace tma

using namespace tma;

namespace {

// WARPITEM: launches of many SHORT items (a large batch at a short context).  Each warp of a CTA takes a whole
// (partition, kv head, sequence) item — grid.x = ceil(items / 4), every block of the item walked by that one warp,
// result filed by that warp — so the per-CTA fixed cost (barrier set-up, table read, first TMA round trip, the
// cross-warp merge through shared memory) is paid once for four items and no block-wide barrier is taken at all.
template <typename T, int D, int STAGES, bool FUSED, bool WARPITEM = false>
__global__ void __launch_bounds__(kWarps * 32, (STAGES <= 3 ? 2 : 1))
paged_attn_tma_kernel(const __grid_constant__ CUtensorMap k_map, const __grid_constant__ CUtensorMap v_map,
                      T* __restrict__ out, const T* __restrict__ q, int64_t q_row_stride,
                      const int32_t* __restrict__ block_tables, int table_stride,
                      const int32_t* __restrict__ context_lens, int Hq, int Hkv, int G, float scale,
                      int blocks_per_split, int num_splits, float* __restrict__ ws_acc, float* __restrict__ ws_ml,
                      int round_scores, const FusedAppend<T> fa, int n_items) {
  using S = Smem<D, STAGES>;
  constexpr int KSTEPS = D / 16;   // mma k-steps over head_dim for S^T
  constexpr int MTILES = D / 16;   // mma m-tiles over head_dim for O^T
  constexpr int BOXES = D / 64;    // TMA boxes per tile

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad = lane >> 2, tq = lane & 3;
  constexpr int WSTRIDE = WARPITEM ? 1 : kWarps;  // distance between the blocks one warp takes
  const int wofs = WARPITEM ? 0 : warp;           // its first block within the CTA's (or its own) run
  int split, kvh, b;
  if constexpr (WARPITEM) {
    const int item = blockIdx.x * kWarps + warp;
    split = item % num_splits;
    kvh = (item / num_splits) % Hkv;
    b = item / (num_splits * Hkv);
  } else {
    split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  }
  // split-K: the combine grid is a programmatic dependent of this one (attn_combine) and may be scheduled as soon as
  // every CTA here has started; a no-op when nothing depends on this launch
  asm volatile("griddepcontrol.launch_dependents;");
  if constexpr (WARPITEM) {
    if (blockIdx.x * kWarps + warp >= n_items) return;  // the last CTA may carry fewer than four items
  }

  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t ring = smem_u32(smem) + warp * STAGES * S::kStageBytes;
  const uint32_t bars = smem_u32(smem + S::kRingBytes) + warp * STAGES * 8;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(bars + s * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();

  // ---- this CTA's slice of the sequence ----
  const int ctx = min(context_lens[b], table_stride * kTok);  // never walk past the end of the block-table row
  const int n_blocks = (ctx + kTok - 1) / kTok;
  const int blk_begin = min(n_blocks, split * blocks_per_split);
  const int blk_end = (split == num_splits - 1) ? n_blocks : min(n_blocks, (split + 1) * blocks_per_split);
  const int n_cta = blk_end - blk_begin;
  const int n_mine = n_cta > wofs ? (n_cta - wofs + WSTRIDE - 1) / WSTRIDE : 0;  // tiles warp, warp+4, ... (or all)
  const int32_t* table = block_tables + static_cast<int64_t>(b) * table_stride + blk_begin + wofs;

  // Block-table window: lane L holds the pool slot of this warp's tile (win * 32 + L).
  int win_base = 0;
  int32_t win_slot = (lane < n_mine) ? table[lane * WSTRIDE] : 0;

  auto issue = [&](int j) {  // all lanes call; lane 0 issues tile j of this warp into stage j % STAGES
    if (j - win_base >= 32) {
      win_base += 32;
      win_slot = (win_base + lane < n_mine) ? table[(win_base + lane) * WSTRIDE] : 0;
    }
    const int32_t slot = __shfl_sync(0xffffffffu, win_slot, j - win_base);
    if (lane == 0) {
      const int st = j % STAGES;
      const uint32_t bar = bars + st * 8;
      const uint32_t dst = ring + st * S::kStageBytes;
      const int row = (slot * Hkv + kvh) * kTok;
      mbar_expect_tx(bar, S::kStageBytes);
#pragma unroll
      for (int bx = 0; bx < BOXES; ++bx) {
        tma_load_2d(dst + bx * (kTok * 128), &k_map, bx * 64, row, bar);
        tma_load_2d(dst + S::kTileBytes + bx * (kTok * 128), &v_map, bx * 64, row, bar);
      }
    }
  };

  const int n_pro = min(STAGES, n_mine);
  for (int j = 0; j < n_pro; ++j) issue(j);

  // ---- Q as the B operand of S^T = K Q^T:  b0 = Q[g = quad][16 kk + 2 tq .. +1], b1 = same at d + 8 ----
  uint32_t qf[KSTEPS][2];
  {
    const T* qrow = q + static_cast<int64_t>(b) * q_row_stride + static_cast<int64_t>(kvh * G + quad) * D;
#pragma unroll
    for (int kk = 0; kk < KSTEPS; ++kk) {
      qf[kk][0] = quad < G ? *reinterpret_cast<const uint32_t*>(qrow + kk * 16 + 2 * tq) : 0u;
      qf[kk][1] = quad < G ? *reinterpret_cast<const uint32_t*>(qrow + kk * 16 + 8 + 2 * tq) : 0u;
    }
  }

  if constexpr (FUSED) {
    // RoPE on q in registers: the rotate-half partner of d is d + D/2, i.e. k-step kk + KSTEPS/2 at the same
    // register slot of the same lane, so the rotation is lane-local.
    const T* cs = fa.cos_t + static_cast<int64_t>(b) * D;
    const T* sn = fa.sin_t + static_cast<int64_t>(b) * D;
#pragma unroll
    for (int kk = 0; kk < KSTEPS / 2; ++kk) {
#pragma unroll
      for (int j = 0; j < 2; ++j) rope_q_regs<T, D>(qf[kk][j], qf[kk + KSTEPS / 2][j], cs, sn, kk * 16 + 8 * j + 2 * tq);
    }
  }
  // the warp that owns the sequence's last KV tile (the one the new token lands in) does the append
  const int last_blk = n_blocks - 1;
  const bool owns_last = ctx > 0 && last_blk >= blk_begin && last_blk < blk_end && ((last_blk - blk_begin) % WSTRIDE) == wofs;
  const bool append_here = FUSED && owns_last;
  // The new token's k / v / cos / sin rows and its pool slot are fetched NOW, while the first KV tiles are in
  // flight, so that the patch of the last tile (end of this warp's loop) does not wait on global memory.
  AppendRows ap{};
  if constexpr (FUSED) {
    if (append_here)
      ap = append_fetch<T, D>(fa, b, kvh, lane, block_tables[static_cast<int64_t>(b) * table_stride + last_blk]);
  }

  // per-lane state for the two q heads g = 2 tq, 2 tq + 1: running max (scaled log2 domain), partial sum, O^T
  const float sl2 = scale * 1.4426950408889634f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  float o[MTILES][4];
#pragma unroll
  for (int mt = 0; mt < MTILES; ++mt) o[mt][0] = o[mt][1] = o[mt][2] = o[mt][3] = 0.f;

  // ldmatrix lane roles (see the file header): matrix index and row inside it
  const int lm_mat = lane >> 3, lm_row = lane & 7;
  const int k_row = (lm_mat & 1) * 8 + lm_row;   // K: matrices 0/2 -> tok 0-7, 1/3 -> tok 8-15
  const int k_chunk = lm_mat >> 1;               //    matrices 0/1 -> chunk 2kk, 2/3 -> chunk 2kk+1
  const int v_row = (lm_mat >> 1) * 8 + lm_row;  // V: matrices 0/1 -> tok 0-7, 2/3 -> tok 8-15
  const int v_chunk = lm_mat & 1;                //    matrices 0/2 -> chunk 2mt, 1/3 -> chunk 2mt+1

  // One KV tile of this warp.  `patch` (compile-time) = this is the sequence's last, partly filled tile: the fused
  // variant appends the new token into it, and its V rows past the context are cleared; kept out of the steady-state
  // loop.
  auto tile = [&](const int i, auto patch) {
    const int st = i % STAGES;
    const uint32_t k_tile = ring + st * S::kStageBytes;
    const uint32_t v_tile = k_tile + S::kTileBytes;
    mbar_wait(bars + st * 8, (i / STAGES) & 1);

    if constexpr (decltype(patch)::value) {
      if constexpr (FUSED) append_patch<T, D>(ap, fa, Hkv, kvh, (ctx - 1) & (kTok - 1), lane, k_tile, v_tile);
      zero_v_tail<D>(v_tile, ((ctx - 1) & (kTok - 1)) + 1, lane);
    }

    // S^T = K Q^T
    float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int kk = 0; kk < KSTEPS; ++kk) {
      uint32_t a[4];
      ldmatrix_x4(a, k_tile + tile_off(k_row, 2 * kk + k_chunk));
      Mma<T>::run(s, a, qf[kk][0], qf[kk][1]);
    }

    // online softmax.  s[0], s[1]: token `quad`; s[2], s[3]: token `quad + 8`; columns g = 2 tq, 2 tq + 1
    const int tok0 = (blk_begin + wofs + i * WSTRIDE) * kTok + quad;
    const bool ok_lo = tok0 < ctx, ok_hi = tok0 + 8 < ctx;
    float x[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float v = s[e];
      if (round_scores) {
        v = Mma<T>::round(Mma<T>::round(v) * scale) * 1.4426950408889634f;
      } else {
        v *= sl2;
      }
      x[e] = ((e < 2) ? ok_lo : ok_hi) ? v : -INFINITY;
    }
    float corr[2];
    bool rescale = false;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float mx = fmaxf(x[c], x[c + 2]);
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
      const float m_new = fmaxf(m_run[c], mx);  // finite: every tile holds at least one valid token
      rescale |= m_new != m_run[c];
      corr[c] = fast_exp2(m_run[c] - m_new);    // 2^-inf = 0 on the first tile
      m_run[c] = m_new;
      x[c] = fast_exp2(x[c] - m_new);
      x[c + 2] = fast_exp2(x[c + 2] - m_new);
      l_run[c] = l_run[c] * corr[c] + x[c] + x[c + 2];
    }
    // P^T as the B operand: transpose the two 8x8 (tok x g) fragments into (g x tok) register layout
    const uint32_t pb0 = movmatrix_trans(Mma<T>::pack(x[0], x[1]));
    const uint32_t pb1 = movmatrix_trans(Mma<T>::pack(x[2], x[3]));

    // O^T = O^T * corr + V^T P^T; once the running maxima have settled every corr is exactly 1 and the multiplies
    // are skipped for the whole warp
    if (__any_sync(0xffffffffu, rescale)) {
#pragma unroll
      for (int mt = 0; mt < MTILES; ++mt) {
        o[mt][0] *= corr[0];
        o[mt][1] *= corr[1];
        o[mt][2] *= corr[0];
        o[mt][3] *= corr[1];
      }
    }
#pragma unroll
    for (int mt = 0; mt < MTILES; ++mt) {
      uint32_t a[4];
      ldmatrix_x4_trans(a, v_tile + tile_off(v_row, 2 * mt + v_chunk));
      Mma<T>::run(o[mt], a, pb0, pb1);
    }

    __syncwarp();  // every lane's reads of this stage are done before lane 0 hands it back to TMA
    if (i + STAGES < n_mine) issue(i + STAGES);
  };
  const bool tail_tile = owns_last && (FUSED || (ctx & (kTok - 1)) != 0);
  const int n_plain = tail_tile ? n_mine - 1 : n_mine;
  for (int i = 0; i < n_plain; ++i) tile(i, std::false_type{});
  if (tail_tile) tile(n_mine - 1, std::true_type{});

  // ---- merge the 4 warps (buffers alias the ring; all TMA traffic of this CTA has landed and been consumed) ----
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    l_run[c] += __shfl_xor_sync(0xffffffffu, l_run[c], 4);
    l_run[c] += __shfl_xor_sync(0xffffffffu, l_run[c], 8);
    l_run[c] += __shfl_xor_sync(0xffffffffu, l_run[c], 16);
  }
  if constexpr (WARPITEM) {
    // this warp's ring is idle (every tile it asked for has been consumed): stage the result there so that lanes
    // write whole rows, and file it without any block-wide barrier
    __syncwarp();
    float* o_w = reinterpret_cast<float*>(smem + warp * STAGES * S::kStageBytes);  // [8][D]
    float* m_w = o_w + 8 * D;                                                      // [8]
    float* l_w = m_w + 8;                                                          // [8]
#pragma unroll
    for (int mt = 0; mt < MTILES; ++mt)
#pragma unroll
      for (int e = 0; e < 4; ++e) o_w[(2 * tq + (e & 1)) * D + 16 * mt + quad + ((e >> 1) << 3)] = o[mt][e];
    if (quad == 0) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        m_w[2 * tq + c] = m_run[c];
        l_w[2 * tq + c] = l_run[c];
      }
    }
    __syncwarp();
    for (int idx = lane; idx < G * D; idx += 32) {
      const int g = idx / D, d = idx - g * D;
      const float acc = o_w[g * D + d], L = l_w[g];
      const int hq = kvh * G + g;
      if (num_splits == 1) {
        out[(static_cast<int64_t>(b) * Hq + hq) * D + d] = Cvt<T>::from_f32(L > 0.f ? acc / L : 0.f);
      } else {
        const int64_t pidx = (static_cast<int64_t>(b) * Hq + hq) * num_splits + split;
        ws_acc[pidx * D + d] = acc;
        if (d == 0) {
          ws_ml[pidx * 2 + 0] = m_w[g] * 0.6931471805599453f;  // natural-log domain, as attn_combine_kernel expects
          ws_ml[pidx * 2 + 1] = L;
        }
      }
    }
    return;
  }
  __syncthreads();
  float* o_s = reinterpret_cast<float*>(smem);        // [W][8][D]
  float* m_s = o_s + kWarps * 8 * D;                  // [W][8]
  float* l_s = m_s + kWarps * 8;                      // [W][8]
#pragma unroll
  for (int mt = 0; mt < MTILES; ++mt) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int g = 2 * tq + (e & 1);
      const int d = 16 * mt + quad + ((e >> 1) << 3);
      o_s[(warp * 8 + g) * D + d] = o[mt][e];
    }
  }
  if (quad == 0) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      m_s[warp * 8 + 2 * tq + c] = m_run[c];
      l_s[warp * 8 + 2 * tq + c] = l_run[c];
    }
  }
  __syncthreads();
  const float inv_l2 = 0.6931471805599453f;  // ln 2: back from the log2 domain for the workspace
  for (int idx = threadIdx.x; idx < G * D; idx += kWarps * 32) {
    const int g = idx / D, d = idx - g * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) M = fmaxf(M, m_s[w * 8 + g]);
    float acc = 0.f, L = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float mw = m_s[w * 8 + g];
      const float wgt = (mw == -INFINITY) ? 0.f : fast_exp2(mw - M);
      acc = fmaf(wgt, o_s[(w * 8 + g) * D + d], acc);
      L = fmaf(wgt, l_s[w * 8 + g], L);
    }
    const int hq = kvh * G + g;
    if (num_splits == 1) {
      out[(static_cast<int64_t>(b) * Hq + hq) * D + d] = Cvt<T>::from_f32(L > 0.f ? acc / L : 0.f);
    } else {
      const int64_t pidx = (static_cast<int64_t>(b) * Hq + hq) * num_splits + split;
      ws_acc[pidx * D + d] = acc;
      if (d == 0) {
        ws_ml[pidx * 2 + 0] = M * inv_l2;  // natural-log domain, as attn_combine_kernel expects
        ws_ml[pidx * 2 + 1] = L;
      }
    }
  }
}

template <typename T, int D, int STAGES, bool FUSED, bool WARPITEM = false>
int launch_tma(const AttnArgs& a, int dtype, cudaStream_t stream) {
  using S = Smem<D, STAGES>;
  const uint64_t rows = static_cast<uint64_t>(a.max_blocks) * a.Hkv * kTok;
  CUtensorMap km, vm;
  int rc = get_tensor_map(a.k_pool, rows, D, dtype, &km);
  if (rc) return rc;
  rc = get_tensor_map(a.v_pool, rows, D, dtype, &vm);
  if (rc) return rc;
  auto kern = paged_attn_tma_kernel<T, D, STAGES, FUSED, WARPITEM>;
  FusedAppend<T> fa{};
  if (FUSED) {
    fa.k_new = static_cast<const T*>(a.k_new);
    fa.k_stride = a.k_new_stride;
    fa.v_new = static_cast<const T*>(a.v_new);
    fa.v_stride = a.v_new_stride;
    fa.cos_t = static_cast<const T*>(a.cos_t);
    fa.sin_t = static_cast<const T*>(a.sin_t);
    fa.k_pool = static_cast<T*>(const_cast<void*>(a.k_pool));
    fa.v_pool = static_cast<T*>(const_cast<void*>(a.v_pool));
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal);
  if (e != cudaSuccess) return fail(static_cast<int>(e), "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  const int n_items = a.num_splits * a.Hkv * a.B;
  const dim3 grid = WARPITEM ? dim3((n_items + kWarps - 1) / kWarps) : dim3(a.num_splits, a.Hkv, a.B);
  kern<<<grid, kWarps * 32, S::kTotal, stream>>>(km, vm, static_cast<T*>(a.out), static_cast<const T*>(a.q),
                                                  a.q_row_stride, a.block_tables, a.table_stride, a.context_lens, a.Hq,
                                                  a.Hkv, a.Hq / a.Hkv, a.scale, a.blocks_per_split, a.num_splits,
                                                  a.ws_acc, a.ws_ml, a.round_scores, fa, n_items);
  return check_launch("b200_paged_decode_attn(tma)");
}

}  // namespace

bool attn_tma_supported(const AttnArgs& a, int dtype) {
  if (dtype != B200_DTYPE_BF16 && dtype != B200_DTYPE_F16) return false;
  if (a.block_size != kTok) return false;
  if (a.D != 64 && a.D != 128) return false;
  if (a.Hq / a.Hkv > 8) return false;
  if (!aligned16(a.k_pool) || !aligned16(a.v_pool)) return false;
  if ((reinterpret_cast<uintptr_t>(a.q) & 3u) || (a.q_row_stride & 1)) return false;
  if (static_cast<uint64_t>(a.max_blocks) * a.Hkv * kTok >= (1ull << 31)) return false;
  return tensor_maps_available();
}

template <int STAGES, bool FUSED, bool WARPITEM = false>
int attn_tma_stages(const AttnArgs& a, int dtype, cudaStream_t s) {
  if (dtype == B200_DTYPE_BF16)
    return a.D == 128 ? launch_tma<__nv_bfloat16, 128, STAGES, FUSED, WARPITEM>(a, dtype, s)
                      : launch_tma<__nv_bfloat16, 64, STAGES, FUSED, WARPITEM>(a, dtype, s);
  return a.D == 128 ? launch_tma<__half, 128, STAGES, FUSED, WARPITEM>(a, dtype, s)
                    : launch_tma<__half, 64, STAGES, FUSED, WARPITEM>(a, dtype, s);
}

// Ring depth: 3 stages x 2 CTAs per SM when the grid has at least two CTAs for every SM; 6 stages x 1 CTA per SM
// when it has not (then a second resident CTA never materialises and the deeper ring keeps the same number of
// bytes in flight per SM).  B200_ATTN_STAGES=3|6 overrides the choice (used by tools/kbench.py sweeps).
int attn_tma(const AttnArgs& a, int dtype, cudaStream_t s) {
  static const int forced = [] {
    const char* e = getenv("B200_ATTN_STAGES");
    return e ? atoi(e) : 0;
  }();
  static const bool warp_items = [] {  // B200_ATTN_WARP_ITEMS=0: always one CTA per item (A/B measurements)
    const char* e = getenv("B200_ATTN_WARP_ITEMS");
    return !(e && atoi(e) == 0);
  }();
  const long ctas = static_cast<long>(a.num_splits) * a.Hkv * a.B;
  const int stages = forced ? forced : (ctas >= 2L * kNumSMs ? 3 : 6);
  // many short items (more than two per SM, at most 32 blocks = 512 tokens each): one item per warp, see the kernel
  if (warp_items && !forced && ctas > 2L * kNumSMs && a.blocks_per_split <= 32)
    return a.k_new != nullptr ? attn_tma_stages<3, true, true>(a, dtype, s) : attn_tma_stages<3, false, true>(a, dtype, s);
  if (a.k_new != nullptr)  // fused RoPE + append + attention (3-stage ring only)
    return attn_tma_stages<3, true>(a, dtype, s);
  return stages >= 6 ? attn_tma_stages<6, false>(a, dtype, s) : attn_tma_stages<3, false>(a, dtype, s);
}

}  // namespace b200
