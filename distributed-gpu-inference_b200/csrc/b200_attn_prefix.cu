// Shared-prefix ("cascade") paged decode attention for sm_100a.
//
// When every sequence of a batch starts with the same blocks (a shared system prompt / radix prefix: the blocks
// whose CacheBlock.ref_count > 1, reference worker/distributed/kv_cache.py:63-72), the plain kernel streams those
// blocks once PER SEQUENCE; they come out of L2 after the first reader, but the launch is still paced by L2 and the
// unique HBM bytes are a fraction of what moves.  Here the prefix is processed ONCE for all sequences:
//
//   launch 1 (this file)   ONE grid with two kinds of CTAs (measured: as separate launches the two phases cost
//                          46 + 37 us back to back; fused they overlap and share the machine):
//       prefix CTAs  (prefix_splits x Hkv x q-groups): a run of prefix blocks of one kv head against NT*8 query
//                    rows at a time — rows are (sequence, q head of the group) pairs, so one K/V tile staged by
//                    TMA feeds several sequences' queries through the same MMAs (S^T = K Q^T with N = 8 NT);
//       suffix CTAs  (suffix_splits x Hkv x B): a sequence's own blocks after the prefix, one kv head, its group
//                    of q heads — the plain kernel's work item.
//                    Every CTA files a partial (m, l, O) in the split-K workspace.
//   launch 2               attn_combine_kernel merges the partials of every (sequence, q head) row; it is a
//                          programmatic dependent of launch 1, so it is resident by the time the last CTA finishes.
//   (in-kernel merge)      alternative kept behind B200_PREFIX_SEPARATE_COMBINE=0: every CTA counts its arrival per row
//                          in a counter in front of the workspace and the CTA that completes a row merges it — no
//                          second launch, but measured slower than the dependent launch at every shape (attn_prefix()).
//
// The caller guarantees: block_tables[b][0:prefix_blocks] are identical for all b, and every context_lens[b] >=
// prefix_blocks * 16 (the prefix blocks are full for everybody, so launch 1 needs no masking).
#include <algorithm>
#include <cstdlib>

#include "b200_attn.cuh"
#include "b200_common.cuh"
#include "b200_tma_common.cuh"

namespace b200 {

using namespace tma;

namespace {

struct CascadeParams {
  const void* q;
  int64_t q_row_stride;
  const int32_t* block_tables;
  int table_stride;
  const int32_t* context_lens;
  int prefix_blocks, B, Hq, Hkv, G;
  float scale;
  int prefix_splits, prefix_bps;   // prefix role: splits and blocks per split
  int suffix_splits, suffix_bps;   // suffix role
  int n_prefix_ctas;               // CTAs [0, n_prefix_ctas) run the prefix role
  int suffix_warp_items;           // != 0: a suffix CTA's four warps each take one (split, kv head, sequence) item
  int n_suffix_items;              // suffix_splits * Hkv * B
  float* ws_acc;
  float* ws_ml;
  int* counters;                   // [B * Hq] arrival counters, zero before and after the launch; null: separate combine
  void* out;                       // [B, Hq, D] (written by the merging CTAs)
  // fused decode step (FUSED kernels only): un-rotated q, the new token's rows, RoPE tables, writable pools
  const void* k_new;
  int64_t k_new_stride;
  const void* v_new;
  int64_t v_new_stride;
  const void* cos_t;
  const void* sin_t;
  void* k_pool;
  void* v_pool;
};

// One CTA's work in either role.  SUFFIX = false: prefix role, columns are (sequence, head) pairs of q-group `zi`.
// SUFFIX = true (NT = 1): columns are the q heads of sequence `zi`; tiles past the context are masked.
// FUSED: q arrives un-rotated and is rotated in registers (both roles); the suffix warp that owns a sequence's last
// tile appends the new token's rotated k and its v, exactly as paged_attn_tma_kernel<.., FUSED> does.
constexpr int kMaxMergeSplits = 256;  // as attn_combine_kernel

// WARPITEM (suffix role only): the calling WARP owns the whole work item — it walks every block of the run itself and
// files the partial on its own, so a CTA carries four items and nothing is merged across warps (short runs: the
// per-CTA fixed cost, not the data, is what a launch of many short items pays for).
template <typename T, int D, int STAGES, int NT, bool SUFFIX, bool FUSED, bool ROUND, bool WARPITEM = false>
__device__ __forceinline__ void cascade_role(const CUtensorMap& k_map, const CUtensorMap& v_map, const CascadeParams& p,
                                             int split, int kvh, int zi) {
  using S = Smem<D, STAGES>;
  constexpr int KSTEPS = D / 16;
  constexpr int MTILES = D / 16;
  constexpr int BOXES = D / 64;
  constexpr int COLS = 8 * NT;  // query rows handled by this CTA

  const T* __restrict__ q = static_cast<const T*>(p.q);
  const int64_t q_row_stride = p.q_row_stride;
  const int B = p.B, Hq = p.Hq, Hkv = p.Hkv, G = p.G;
  const float scale = p.scale;
  float* ws_acc = p.ws_acc;
  float* ws_ml = p.ws_ml;
  const int splits_total = p.prefix_splits + p.suffix_splits;
  const int qgroup = zi;
  static_assert(!WARPITEM || SUFFIX, "one warp per item is a suffix-role variant");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad = lane >> 2, tq = lane & 3;
  constexpr int WSTRIDE = WARPITEM ? 1 : kWarps;  // distance between the blocks one warp takes
  const int wofs = WARPITEM ? 0 : warp;           // its first block within the run

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

  int blk_begin, blk_end, ctx = 0x7fffffff;
  const int32_t* table_row;
  if constexpr (SUFFIX) {
    ctx = min(p.context_lens[zi], p.table_stride * kTok);  // never walk past the end of the block-table row
    const int n_blocks = (ctx + kTok - 1) / kTok;
    blk_begin = min(n_blocks, p.prefix_blocks + split * p.suffix_bps);
    blk_end = (split == p.suffix_splits - 1) ? n_blocks : min(n_blocks, p.prefix_blocks + (split + 1) * p.suffix_bps);
    table_row = p.block_tables + static_cast<int64_t>(zi) * p.table_stride;
  } else {
    blk_begin = min(p.prefix_blocks, split * p.prefix_bps);
    blk_end = min(p.prefix_blocks, (split + 1) * p.prefix_bps);
    table_row = p.block_tables;  // the prefix entries are the same in every row
  }
  const int n_cta = max(0, blk_end - blk_begin);
  const int n_mine = n_cta > wofs ? (n_cta - wofs + WSTRIDE - 1) / WSTRIDE : 0;
  const int32_t* table = table_row + blk_begin + wofs;

  int win_base = 0;
  int32_t win_slot = (lane < n_mine) ? table[lane * WSTRIDE] : 0;
  auto issue = [&](int j) {
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

  // Q rows of this CTA as B operands: column c = nt * 8 + quad  <->  row r = qgroup * COLS + c  <->  (b, g)
  uint32_t qf[KSTEPS][NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    int bq, g;
    bool ok;
    if constexpr (SUFFIX) {
      bq = zi;
      g = quad;
      ok = quad < G;
    } else {
      const int r = qgroup * COLS + nt * 8 + quad;
      bq = r / G;
      g = r - bq * G;
      ok = bq < B;
    }
    const T* qrow = q + static_cast<int64_t>(ok ? bq : 0) * q_row_stride + static_cast<int64_t>(kvh * G + g) * D;
#pragma unroll
    for (int kk = 0; kk < KSTEPS; ++kk) {
      qf[kk][nt][0] = ok ? *reinterpret_cast<const uint32_t*>(qrow + kk * 16 + 2 * tq) : 0u;
      qf[kk][nt][1] = ok ? *reinterpret_cast<const uint32_t*>(qrow + kk * 16 + 8 + 2 * tq) : 0u;
    }
    if constexpr (FUSED) {
      if (ok) {
        const T* cs = static_cast<const T*>(p.cos_t) + static_cast<int64_t>(bq) * D;
        const T* sn = static_cast<const T*>(p.sin_t) + static_cast<int64_t>(bq) * D;
#pragma unroll
        for (int kk = 0; kk < KSTEPS / 2; ++kk) {
#pragma unroll
          for (int j = 0; j < 2; ++j)
            rope_q_regs<T, D>(qf[kk][nt][j], qf[kk + KSTEPS / 2][nt][j], cs, sn, kk * 16 + 8 * j + 2 * tq);
        }
      }
    }
  }

  // suffix role: the warp that owns the sequence's last KV tile appends the new token into it
  bool append_here = false, owns_last = false;
  AppendRows ap{};
  FusedAppend<T> fa{};
  if constexpr (SUFFIX) {
    const int last_blk = (ctx + kTok - 1) / kTok - 1;
    owns_last = ctx > 0 && last_blk >= blk_begin && last_blk < blk_end && ((last_blk - blk_begin) % WSTRIDE) == wofs;
  }
  if constexpr (FUSED && SUFFIX) {
    const int last_blk = (ctx + kTok - 1) / kTok - 1;
    append_here = owns_last;
    fa.k_new = static_cast<const T*>(p.k_new);
    fa.k_stride = p.k_new_stride;
    fa.v_new = static_cast<const T*>(p.v_new);
    fa.v_stride = p.v_new_stride;
    fa.cos_t = static_cast<const T*>(p.cos_t);
    fa.sin_t = static_cast<const T*>(p.sin_t);
    fa.k_pool = static_cast<T*>(p.k_pool);
    fa.v_pool = static_cast<T*>(p.v_pool);
    if (append_here) ap = append_fetch<T, D>(fa, zi, kvh, lane, table_row[last_blk]);
  }

  const float sl2 = scale * 1.4426950408889634f;
  float m_run[NT][2], l_run[NT][2], o[MTILES][NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    m_run[nt][0] = m_run[nt][1] = -INFINITY;
    l_run[nt][0] = l_run[nt][1] = 0.f;
#pragma unroll
    for (int mt = 0; mt < MTILES; ++mt) o[mt][nt][0] = o[mt][nt][1] = o[mt][nt][2] = o[mt][nt][3] = 0.f;
  }

  const int lm_mat = lane >> 3, lm_row = lane & 7;
  const int k_row = (lm_mat & 1) * 8 + lm_row, k_chunk = lm_mat >> 1;
  const int v_row = (lm_mat >> 1) * 8 + lm_row, v_chunk = lm_mat & 1;

  for (int i = 0; i < n_mine; ++i) {
    const int st = i % STAGES;
    const uint32_t k_tile = ring + st * S::kStageBytes;
    const uint32_t v_tile = k_tile + S::kTileBytes;
    mbar_wait(bars + st * 8, (i / STAGES) & 1);
    if constexpr (SUFFIX) {
      if (owns_last && i == n_mine - 1) {  // the sequence's last, partly filled tile
        if constexpr (FUSED) append_patch<T, D>(ap, fa, Hkv, kvh, (ctx - 1) & (kTok - 1), lane, k_tile, v_tile);
        zero_v_tail<D>(v_tile, ((ctx - 1) & (kTok - 1)) + 1, lane);  // rows past the context: 0 x NaN must stay 0
      }
    }

    // S^T = K Q^T.  Two accumulators per column group (even / odd k-steps) halve the dependent-MMA chain.
    float s[NT][4], s2[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      s2[nt][0] = s2[nt][1] = s2[nt][2] = s2[nt][3] = 0.f;
    }
#pragma unroll
    for (int kk = 0; kk < KSTEPS; kk += 2) {
      uint32_t a[4], a2[4];
      ldmatrix_x4(a, k_tile + tile_off(k_row, 2 * kk + k_chunk));
      ldmatrix_x4(a2, k_tile + tile_off(k_row, 2 * kk + 2 + k_chunk));
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        Mma<T>::run(s[nt], a, qf[kk][nt][0], qf[kk][nt][1]);
        Mma<T>::run(s2[nt], a2, qf[kk + 1][nt][0], qf[kk + 1][nt][1]);
      }
    }

    float corr[NT][2];
    uint32_t pb[NT][2];
    bool rescale = false;
    const int tok0 = (blk_begin + wofs + i * WSTRIDE) * kTok + quad;
    const bool ok_lo = !SUFFIX || tok0 < ctx, ok_hi = !SUFFIX || tok0 + 8 < ctx;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      float x[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float sv = s[nt][e] + s2[nt][e];
        const float v = ROUND ? Mma<T>::round(Mma<T>::round(sv) * scale) * 1.4426950408889634f : sv * sl2;
        x[e] = ((e < 2) ? ok_lo : ok_hi) ? v : -INFINITY;
      }
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float mx = fmaxf(x[c], x[c + 2]);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
        const float m_new = fmaxf(m_run[nt][c], mx);  // finite: every tile holds at least one valid token
        rescale |= m_new != m_run[nt][c];
        corr[nt][c] = fast_exp2(m_run[nt][c] - m_new);  // 2^-inf = 0 on the first tile
        m_run[nt][c] = m_new;
        x[c] = fast_exp2(x[c] - m_new);
        x[c + 2] = fast_exp2(x[c + 2] - m_new);
        l_run[nt][c] = l_run[nt][c] * corr[nt][c] + x[c] + x[c + 2];
      }
      pb[nt][0] = movmatrix_trans(Mma<T>::pack(x[0], x[1]));
      pb[nt][1] = movmatrix_trans(Mma<T>::pack(x[2], x[3]));
    }

    // O^T = O^T * corr + V^T P^T.  Once the running maxima have settled (after the first few tiles of a run) every
    // corr is exactly 1 and the 4 * NT * MTILES multiplies are skipped for the whole warp.
    if (__any_sync(0xffffffffu, rescale)) {
#pragma unroll
      for (int mt = 0; mt < MTILES; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          o[mt][nt][0] *= corr[nt][0];
          o[mt][nt][1] *= corr[nt][1];
          o[mt][nt][2] *= corr[nt][0];
          o[mt][nt][3] *= corr[nt][1];
        }
    }
#pragma unroll
    for (int mt = 0; mt < MTILES; ++mt) {
      uint32_t a[4];
      ldmatrix_x4_trans(a, v_tile + tile_off(v_row, 2 * mt + v_chunk));
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) Mma<T>::run(o[mt][nt], a, pb[nt][0], pb[nt][1]);
    }
    __syncwarp();
    if (i + STAGES < n_mine) issue(i + STAGES);
  }

  // merge the 4 warps, then file one partial per (sequence, q head) of this CTA's columns
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      l_run[nt][c] += __shfl_xor_sync(0xffffffffu, l_run[nt][c], 4);
      l_run[nt][c] += __shfl_xor_sync(0xffffffffu, l_run[nt][c], 8);
      l_run[nt][c] += __shfl_xor_sync(0xffffffffu, l_run[nt][c], 16);
    }
  if constexpr (WARPITEM) {
    // The warp's own ring is idle now (every tile it asked for has been consumed): the partial is staged there, so
    // that lanes can write whole rows, and filed by this warp alone — no block-wide barrier anywhere on this path.
    __syncwarp();
    float* o_w = reinterpret_cast<float*>(smem + warp * STAGES * S::kStageBytes);  // [COLS][D]
    float* m_w = o_w + COLS * D;                                                   // [COLS]
    float* l_w = m_w + COLS;                                                       // [COLS]
#pragma unroll
    for (int mt = 0; mt < MTILES; ++mt)
#pragma unroll
      for (int e = 0; e < 4; ++e) o_w[(2 * tq + (e & 1)) * D + 16 * mt + quad + ((e >> 1) << 3)] = o[mt][0][e];
    if (quad == 0) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        m_w[2 * tq + c] = m_run[0][c];
        l_w[2 * tq + c] = l_run[0][c];
      }
    }
    __syncwarp();
    for (int idx = lane; idx < G * D; idx += 32) {
      const int c = idx / D, d = idx - c * D;
      const int64_t pidx = (static_cast<int64_t>(zi) * Hq + kvh * G + c) * splits_total + p.prefix_splits + split;
      ws_acc[pidx * D + d] = o_w[c * D + d];
      if (d == 0) {
        ws_ml[pidx * 2 + 0] = m_w[c] * 0.6931471805599453f;  // natural-log domain, as attn_combine_kernel expects
        ws_ml[pidx * 2 + 1] = l_w[c];
      }
    }
    return;
  }
  __syncthreads();
  float* o_s = reinterpret_cast<float*>(smem);   // [W][COLS][D]
  float* m_s = o_s + kWarps * COLS * D;          // [W][COLS]
  float* l_s = m_s + kWarps * COLS;              // [W][COLS]
#pragma unroll
  for (int mt = 0; mt < MTILES; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = nt * 8 + 2 * tq + (e & 1);
        const int d = 16 * mt + quad + ((e >> 1) << 3);
        o_s[(warp * COLS + c) * D + d] = o[mt][nt][e];
      }
  if (quad == 0) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        m_s[warp * COLS + nt * 8 + 2 * tq + c] = m_run[nt][c];
        l_s[warp * COLS + nt * 8 + 2 * tq + c] = l_run[nt][c];
      }
  }
  __syncthreads();
  const float ln2 = 0.6931471805599453f;
  auto row_of = [&](int c, int& bq, int& g) -> bool {  // column c of this CTA -> (sequence, head of the group)
    if constexpr (SUFFIX) {
      bq = zi;
      g = c;
      return c < G;
    } else {
      const int r = qgroup * COLS + c;
      bq = r / G;
      g = r - bq * G;
      return bq < B;
    }
  };
  for (int idx = threadIdx.x; idx < COLS * D; idx += kWarps * 32) {
    const int c = idx / D, d = idx - c * D;
    int bq, g;
    if (!row_of(c, bq, g)) continue;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) M = fmaxf(M, m_s[w * COLS + c]);
    float acc = 0.f, L = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float mw = m_s[w * COLS + c];
      const float wgt = (mw == -INFINITY) ? 0.f : fast_exp2(mw - M);
      acc = fmaf(wgt, o_s[(w * COLS + c) * D + d], acc);
      L = fmaf(wgt, l_s[w * COLS + c], L);
    }
    const int64_t pidx = (static_cast<int64_t>(bq) * Hq + kvh * G + g) * splits_total + (SUFFIX ? p.prefix_splits : 0) + split;
    ws_acc[pidx * D + d] = acc;
    if (d == 0) {
      ws_ml[pidx * 2 + 0] = M * ln2;  // natural-log domain, as attn_combine_kernel expects
      ws_ml[pidx * 2 + 1] = L;
    }
  }
  if (p.counters == nullptr) return;  // two-launch form: attn_combine_kernel merges

  // ---- merge by the last-arriving CTA of each row ------------------------------------------------------------------
  // One warp per row (the CTA's rows are dealt round-robin to its 4 warps): lanes stride over the partials for the
  // weights, then every lane owns D / 32 consecutive output elements and streams the partial rows with independent
  // 8- or 16-byte loads, so a row costs about two global round trips.
  constexpr int VPL = D / 32;  // output elements per lane: 4 (float4) or 2 (float2)
  __shared__ int last_s[COLS];
  __shared__ float w_s[kWarps][kMaxMergeSplits];
  __threadfence();  // this CTA's partials are visible device-wide before its arrival is counted
  __syncthreads();
  if (threadIdx.x < COLS) {
    int bq, g, last = 0;
    if (row_of(threadIdx.x, bq, g)) {
      int* cnt = p.counters + static_cast<int64_t>(bq) * Hq + kvh * G + g;
      last = atomicAdd(cnt, 1) == splits_total - 1;
      if (last) *cnt = 0;  // everybody has arrived: leave the counter as the next launch expects it
    }
    last_s[threadIdx.x] = last;
  }
  __syncthreads();
  T* __restrict__ out = static_cast<T*>(p.out);
  float* wv = w_s[warp];
  for (int c = warp; c < COLS; c += kWarps) {
    if (!last_s[c]) continue;  // warp-uniform
    int bq, g;
    row_of(c, bq, g);
    const int64_t row = static_cast<int64_t>(bq) * Hq + kvh * G + g;
    __threadfence();
    const float2* ml = reinterpret_cast<const float2*>(ws_ml) + row * splits_total;
    float M = -INFINITY;
    for (int sp = lane; sp < splits_total; sp += 32) M = fmaxf(M, __ldcg(ml + sp).x);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, off));
    float L = 0.f;
    for (int sp = lane; sp < splits_total; sp += 32) {
      const float2 v = __ldcg(ml + sp);
      const float w = (v.x == -INFINITY) ? 0.f : __expf(v.x - M);
      wv[sp] = w;
      L = fmaf(w, v.y, L);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) L += __shfl_xor_sync(0xffffffffu, L, off);
    const float inv = L > 0.f ? 1.f / L : 0.f;
    __syncwarp();
    const float* acc = ws_acc + row * splits_total * D + lane * VPL;
    float o[VPL];
#pragma unroll
    for (int e = 0; e < VPL; ++e) o[e] = 0.f;
#pragma unroll 8
    for (int sp = 0; sp < splits_total; ++sp) {
      const float w = wv[sp];
      if constexpr (VPL == 4) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(acc + static_cast<int64_t>(sp) * D));
        o[0] = fmaf(w, v.x, o[0]);
        o[1] = fmaf(w, v.y, o[1]);
        o[2] = fmaf(w, v.z, o[2]);
        o[3] = fmaf(w, v.w, o[3]);
      } else {
        const float2 v = __ldcg(reinterpret_cast<const float2*>(acc + static_cast<int64_t>(sp) * D));
        o[0] = fmaf(w, v.x, o[0]);
        o[1] = fmaf(w, v.y, o[1]);
      }
    }
    T* orow = out + row * D + lane * VPL;
#pragma unroll
    for (int e = 0; e < VPL; ++e) orow[e] = Cvt<T>::from_f32(o[e] * inv);
    __syncwarp();  // wv is reused by this warp's next row
  }
}

template <typename T, int D, int STAGES, int NT, bool FUSED, bool ROUND>
__global__ void __launch_bounds__(kWarps * 32, (STAGES <= 2 ? 3 : 2))
paged_attn_cascade_kernel(const __grid_constant__ CUtensorMap k_map, const __grid_constant__ CUtensorMap v_map,
                          const __grid_constant__ CascadeParams p) {
  // lets the combine grid (a programmatic dependent, see attn_combine) be scheduled once every CTA of this grid has
  // started; it still waits for this grid to complete before it reads a partial
  asm volatile("griddepcontrol.launch_dependents;");
  const int id = blockIdx.x;
  if (id < p.n_prefix_ctas) {  // heavier CTAs first
    const int split = id % p.prefix_splits;
    const int rest = id / p.prefix_splits;
    cascade_role<T, D, STAGES, NT, false, FUSED, ROUND>(k_map, v_map, p, split, rest % p.Hkv, rest / p.Hkv);
  } else {
    const int j = id - p.n_prefix_ctas;
    if (p.suffix_warp_items) {  // short runs: one item per warp
      const int item = j * kWarps + (threadIdx.x >> 5);
      if (item >= p.n_suffix_items) return;
      const int split = item % p.suffix_splits;
      const int rest = item / p.suffix_splits;
      cascade_role<T, D, STAGES, 1, true, FUSED, ROUND, true>(k_map, v_map, p, split, rest % p.Hkv, rest / p.Hkv);
      return;
    }
    const int split = j % p.suffix_splits;
    const int rest = j / p.suffix_splits;
    cascade_role<T, D, STAGES, 1, true, FUSED, ROUND>(k_map, v_map, p, split, rest % p.Hkv, rest / p.Hkv);
  }
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

constexpr int kPrefixNT = 2;       // 16 query rows ((sequence, head) pairs) per prefix CTA
constexpr int kPrefixStages = 3;   // ring depth at 2 CTAs per SM (2 stages x 3 CTAs measured the same: 63.2 vs 63.5 us on config 5)

bool warp_items_allowed() {  // B200_PREFIX_WARP_ITEMS=0 keeps one CTA per suffix item (for A/B measurements)
  static const bool v = [] {
    const char* e = getenv("B200_PREFIX_WARP_ITEMS");
    return !(e && atoi(e) == 0);
  }();
  return v;
}

int separate_combine() {  // -1: library's choice, 0 / 1: forced by B200_PREFIX_SEPARATE_COMBINE
  static const int v = [] {
    const char* e = getenv("B200_PREFIX_SEPARATE_COMBINE");
    return e ? (atoi(e) != 0 ? 1 : 0) : -1;
  }();
  return v;
}

template <typename T, int D, bool FUSED, bool ROUND>
int launch_cascade_impl(const AttnArgs& a, int dtype, int prefix_blocks, int prefix_splits, int suffix_splits,
                        int* counters, cudaStream_t stream) {
  constexpr int NT = kPrefixNT;
  using S = Smem<D, kPrefixStages>;
  const uint64_t rows = static_cast<uint64_t>(a.max_blocks) * a.Hkv * kTok;
  CUtensorMap km, vm;
  int rc = get_tensor_map(a.k_pool, rows, D, dtype, &km);
  if (rc) return rc;
  rc = get_tensor_map(a.v_pool, rows, D, dtype, &vm);
  if (rc) return rc;
  auto kern = paged_attn_cascade_kernel<T, D, kPrefixStages, NT, FUSED, ROUND>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal);
  if (e != cudaSuccess) return fail(static_cast<int>(e), "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  CascadeParams p{};
  p.q = a.q;
  p.q_row_stride = a.q_row_stride;
  p.block_tables = a.block_tables;
  p.table_stride = a.table_stride;
  p.context_lens = a.context_lens;
  p.prefix_blocks = prefix_blocks;
  p.B = a.B;
  p.Hq = a.Hq;
  p.Hkv = a.Hkv;
  p.G = a.Hq / a.Hkv;
  p.scale = a.scale;
  p.prefix_splits = prefix_splits;
  p.prefix_bps = ceil_div(prefix_blocks, prefix_splits);
  p.suffix_splits = suffix_splits;
  p.suffix_bps = a.blocks_per_split;
  const int qgroups = ceil_div(a.B * p.G, 8 * NT);
  p.n_prefix_ctas = prefix_splits * a.Hkv * qgroups;
  p.ws_acc = a.ws_acc;
  p.ws_ml = a.ws_ml;
  p.counters = counters;
  p.out = a.out;
  if (FUSED) {
    p.k_new = a.k_new;
    p.k_new_stride = a.k_new_stride;
    p.v_new = a.v_new;
    p.v_new_stride = a.v_new_stride;
    p.cos_t = a.cos_t;
    p.sin_t = a.sin_t;
    p.k_pool = const_cast<void*>(a.k_pool);
    p.v_pool = const_cast<void*>(a.v_pool);
  }
  // Short runs after the prefix (<= 32 blocks per piece) in a launch of many items: one item per WARP, four per CTA —
  // the fixed cost of a CTA (barriers, table read, first TMA round trip, cross-warp merge, partial write) is paid a
  // quarter as often.  Needs the combine launch (no block-wide barrier on that path, hence no arrival counting).
  p.n_suffix_items = suffix_splits * a.Hkv * a.B;
  p.suffix_warp_items = warp_items_allowed() && counters == nullptr && p.n_suffix_items > 2 * kNumSMs &&
                        p.suffix_bps <= 32;
  const int suffix_ctas = p.suffix_warp_items ? ceil_div(p.n_suffix_items, kWarps) : p.n_suffix_items;
  const int grid = p.n_prefix_ctas + suffix_ctas;
  kern<<<grid, kWarps * 32, S::kTotal, stream>>>(km, vm, p);
  return check_launch("b200_paged_decode_attn_prefix(cascade)");
}

template <typename T, int D>
int launch_cascade(const AttnArgs& a, int dtype, int prefix_blocks, int prefix_splits, int suffix_splits, int* counters,
                   cudaStream_t stream) {
  const bool fused = a.k_new != nullptr, round = a.round_scores != 0;
#define B200_CASCADE_GO(F, R) \
  launch_cascade_impl<T, D, F, R>(a, dtype, prefix_blocks, prefix_splits, suffix_splits, counters, stream)
  if (fused) return round ? B200_CASCADE_GO(true, true) : B200_CASCADE_GO(true, false);
  return round ? B200_CASCADE_GO(false, true) : B200_CASCADE_GO(false, false);
#undef B200_CASCADE_GO
}

}  // namespace

bool attn_prefix_supported(const AttnArgs& a, int dtype) { return attn_tma_supported(a, dtype); }

// Split counts for the two roles.  Measured on B200 at config 5 (tools/kbench.py --cases cfg5c, sweep in
// profiles/r01_kbench_cascade_splits.txt): about ONE prefix CTA per SM is best (8-10 prefix splits: 64 us; 19: 78 us;
// 38: 87 us; 4: 94 us) — the suffix CTAs of the same launch fill the second CTA slot of every SM — with at least
// 256 tokens per split; the suffix role follows the plain kernel's rule on the context left after the prefix.
void attn_prefix_plan(int B, int Hq, int Hkv, int block_size, int prefix_blocks, int max_context_len,
                      int* prefix_splits, int* suffix_splits) {
  const int G = std::max(1, Hq / std::max(1, Hkv));
  const int qgroups = std::max(1, ceil_div(B * G, 8 * kPrefixNT));
  const int min_blocks = std::max(1, 256 / block_size);
  int ps = ceil_div(kNumSMs, std::max(1, Hkv * qgroups));
  // a grid that nearly fills the machine unsplit (>= 100 prefix CTAs) is split in two only for runs of >= 1024 tokens
  // per piece (profiles/r02_prefix_split_sweep.jsonl: 64 sequences, 512-token prompt: 24.9 us unsplit vs 27.5 us;
  // 2048-token prompt: 41.8 us split in two vs 46.2 us)
  const int min_prefix_blocks = Hkv * qgroups >= 100 ? std::max(min_blocks, 1024 / block_size) : min_blocks;
  ps = std::max(1, std::min(ps, std::max(1, prefix_blocks / min_prefix_blocks)));
  *prefix_splits = std::min(ps, 96);
  const int suffix_tokens = std::max(0, max_context_len - prefix_blocks * block_size);
  const int suffix_blocks = std::max(1, ceil_div(std::max(1, suffix_tokens), block_size));
  const int pairs = std::max(1, B * Hkv);
  int ss = pairs >= 120 ? 1 : ceil_div(kNumSMs * 7 / 2, pairs);
  ss = std::max(1, std::min(ss, std::max(1, suffix_blocks / min_blocks)));
  *suffix_splits = std::min(ss, 96);
}

int attn_prefix(const AttnArgs& a, int dtype, int prefix_blocks, int prefix_splits, int suffix_splits, cudaStream_t s) {
  const int total = prefix_splits + suffix_splits;
  if (total > kMaxMergeSplits)
    return fail(B200_ESHAPE, "b200_paged_decode_attn_prefix: more than %d splits cannot be merged", kMaxMergeSplits);
  // Arrival counters in front of the partials.  They must read zero when the kernel starts; the merging CTAs put
  // them back to zero, so a caller that owns the workspace exclusively can promise that (B200_ATTN_WS_COUNTERS_ZERO)
  // and save the memset node.
  // Who merges: a separate combine launch (attn_combine_kernel, a programmatic dependent of this grid, so its launch
  // latency hides behind the attention CTAs).  Merging inside the kernel by the last-arriving CTA of each row was
  // the better choice for short prefix runs before the combine became a dependent launch; measured again afterwards
  // on B200 (tools/prefix_merge_probe.py, profiles/r02_prefix_merge_probe.jsonl, counters vouched for): the separate
  // launch wins at every shape tried — 8 x (512 + 64): 17.5 vs 24.6 us, 64 x (2048 + 176): 41.0 vs 43.7 us,
  // config 5: 72.3 vs 80.4 us — because the rows of a CTA are merged by whichever CTA finishes last, at the very end
  // of the launch, with nothing left to overlap.  B200_PREFIX_SEPARATE_COMBINE=0 still selects the in-kernel merge.
  const int force = separate_combine();
  const bool in_kernel = force == 0;
  int* counters = in_kernel ? a.ws_counters : nullptr;
  if (counters != nullptr && !a.ws_counters_zero) {
    cudaError_t e = cudaMemsetAsync(counters, 0, sizeof(int) * static_cast<size_t>(a.B) * a.Hq, s);
    if (e != cudaSuccess) return fail(static_cast<int>(e), "cudaMemsetAsync(counters): %s", cudaGetErrorString(e));
  }
  int rc;
  if (dtype == B200_DTYPE_BF16)
    rc = a.D == 128 ? launch_cascade<__nv_bfloat16, 128>(a, dtype, prefix_blocks, prefix_splits, suffix_splits, counters, s)
                    : launch_cascade<__nv_bfloat16, 64>(a, dtype, prefix_blocks, prefix_splits, suffix_splits, counters, s);
  else
    rc = a.D == 128 ? launch_cascade<__half, 128>(a, dtype, prefix_blocks, prefix_splits, suffix_splits, counters, s)
                    : launch_cascade<__half, 64>(a, dtype, prefix_blocks, prefix_splits, suffix_splits, counters, s);
  if (rc || counters != nullptr) return rc;
  return attn_combine(a.out, a.ws_acc, a.ws_ml, a.B * a.Hq, a.D, total, dtype, s);
}

}  // namespace b200
