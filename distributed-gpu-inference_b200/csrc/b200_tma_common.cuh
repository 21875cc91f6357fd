// PTX wrappers, fragment helpers and the tensor-map cache shared by the TMA + tensor-core attention kernels
// (b200_attn_tma.cu, b200_attn_prefix.cu).
#pragma once
#include <cuda.h>

#include "b200_common.cuh"

namespace b200 {
namespace tma {

constexpr int kWarps = 4;
constexpr int kTok = 16;  // tokens per tile == block_size

// ---- PTX wrappers ---------------------------------------------------------------------------------------
static __device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

static __device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
static __device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
static __device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
static __device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
static __device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
static __device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

// 2^x in one MUFU instruction (exp2f() wraps the same instruction in range scaling for denormal results, four more
// instructions per call).  ex2.approx(-inf) = +0, which the online softmax relies on.
static __device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T>
struct Mma;
template <>
struct Mma<__nv_bfloat16> {
  static __device__ __forceinline__ void run(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  static __device__ __forceinline__ float round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
};
template <>
struct Mma<__half> {
  static __device__ __forceinline__ void run(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  static __device__ __forceinline__ float round(float x) { return __half2float(__float2half_rn(x)); }
};

// Byte offset of (row, 16-byte chunk) inside a K or V tile staged by TMA with 128-byte swizzle.
// The tile is stored as D/64 boxes of [16 rows][128 B]; inside a box chunk c of row r sits at c ^ (r & 7).
static __device__ __forceinline__ uint32_t tile_off(int row, int chunk) {
  return static_cast<uint32_t>((chunk >> 3) * (kTok * 128) + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}

template <int D, int STAGES>
struct Smem {
  static constexpr int kTileBytes = kTok * D * 2;       // one K or V tile
  static constexpr int kStageBytes = 2 * kTileBytes;    // K then V
  static constexpr int kRingBytes = kWarps * STAGES * kStageBytes;
  static constexpr int kBarBytes = kWarps * STAGES * 8;
  static constexpr int kTotal = kRingBytes + kBarBytes + 1024 /* alignment slack */;
  // the merge buffers alias the ring after the main loop: [W][8][D] floats + 2 * [W][8]
  static_assert(kWarps * 8 * (D + 2) * 4 <= kRingBytes, "merge buffers must fit in the ring");
};


// ---- fused decode step: RoPE + append inside the attention launch ----------------------------------------------
// Operands of the fused variants: this step's un-rotated q, the new token's k and v rows and the RoPE tables.  The
// kernel rotates q in registers, and the one warp that owns a sequence's last KV tile rotates the new k row, writes
// it and the v row into the pool AND into its staged tile — so RoPE, the KV append and the attention of a decode step
// are one launch.
template <typename T>
struct FusedAppend {
  const T* k_new;       // [B, Hkv, D], row stride k_stride
  int64_t k_stride;
  const T* v_new;
  int64_t v_stride;
  const T* cos_t;       // [B, D]
  const T* sin_t;
  T* k_pool;            // writable views of the pools behind k_map / v_map
  T* v_pool;
};

// rotate-half RoPE on one element pair with torch's rounding points (see b200_layer_ops.cu)
template <typename T>
static __device__ __forceinline__ void rope_pair(T x_lo, T x_hi, T c_lo, T s_lo, T c_hi, T s_hi, T& o_lo, T& o_hi) {
  const float xl = Cvt<T>::to_f32(x_lo), xh = Cvt<T>::to_f32(x_hi);
  const T a = Cvt<T>::from_f32(__fmul_rn(xl, Cvt<T>::to_f32(c_lo)));
  const T b = Cvt<T>::from_f32(__fmul_rn(-xh, Cvt<T>::to_f32(s_lo)));
  o_lo = Cvt<T>::from_f32(__fadd_rn(Cvt<T>::to_f32(a), Cvt<T>::to_f32(b)));
  const T c = Cvt<T>::from_f32(__fmul_rn(xh, Cvt<T>::to_f32(c_hi)));
  const T d = Cvt<T>::from_f32(__fmul_rn(xl, Cvt<T>::to_f32(s_hi)));
  o_hi = Cvt<T>::from_f32(__fadd_rn(Cvt<T>::to_f32(c), Cvt<T>::to_f32(d)));
}

// RoPE on one query row held as MMA B-operand fragments: `lo` is the register of k-step kk (elements d, d + 1 with
// d = 16 kk + 8 j + 2 tq), `hi` the same register of k-step kk + D/32 — the rotate-half partners d + D/2 sit in the
// same lane, so the rotation needs no shuffles.  cs / sn: this sequence's [D] cos / sin rows.
template <typename T, int D>
static __device__ __forceinline__ void rope_q_regs(uint32_t& lo_reg, uint32_t& hi_reg, const T* cs, const T* sn, int d) {
  union { uint32_t u; T e[2]; } lo, hi, c_lo, s_lo, c_hi, s_hi, o_lo, o_hi;
  lo.u = lo_reg;
  hi.u = hi_reg;
  c_lo.u = *reinterpret_cast<const uint32_t*>(cs + d);
  s_lo.u = *reinterpret_cast<const uint32_t*>(sn + d);
  c_hi.u = *reinterpret_cast<const uint32_t*>(cs + D / 2 + d);
  s_hi.u = *reinterpret_cast<const uint32_t*>(sn + D / 2 + d);
  rope_pair<T>(lo.e[0], hi.e[0], c_lo.e[0], s_lo.e[0], c_hi.e[0], s_hi.e[0], o_lo.e[0], o_hi.e[0]);
  rope_pair<T>(lo.e[1], hi.e[1], c_lo.e[1], s_lo.e[1], c_hi.e[1], s_hi.e[1], o_lo.e[1], o_hi.e[1]);
  lo_reg = o_lo.u;
  hi_reg = o_hi.u;
}

// The new token's rows, fetched by the appending warp while its first KV tiles are in flight: lanes 0..D/8-1 hold
// one 16-byte chunk of k (x), its rotate-half partner (p) and the cos / sin chunks; lanes 16.. hold a chunk of v (x).
struct AppendRows {
  int4 x, p, c, s;
  int64_t slot;
};

template <typename T, int D>
static __device__ __forceinline__ AppendRows append_fetch(const FusedAppend<T>& fa, int b, int kvh, int lane, int64_t slot) {
  constexpr int CH = D / 8;  // 16-byte chunks per row
  AppendRows r;
  r.x = r.p = r.c = r.s = make_int4(0, 0, 0, 0);
  r.slot = slot;
  if (lane < CH) {
    const int c = lane, pc = (c + CH / 2) % CH;
    const T* kr = fa.k_new + static_cast<int64_t>(b) * fa.k_stride + static_cast<int64_t>(kvh) * D;
    r.x = *reinterpret_cast<const int4*>(kr + 8 * c);
    r.p = *reinterpret_cast<const int4*>(kr + 8 * pc);
    r.c = *reinterpret_cast<const int4*>(fa.cos_t + static_cast<int64_t>(b) * D + 8 * c);
    r.s = *reinterpret_cast<const int4*>(fa.sin_t + static_cast<int64_t>(b) * D + 8 * c);
  } else if (lane >= 16 && lane < 16 + CH) {
    const T* vr = fa.v_new + static_cast<int64_t>(b) * fa.v_stride + static_cast<int64_t>(kvh) * D;
    r.x = *reinterpret_cast<const int4*>(vr + 8 * (lane - 16));
  }
  return r;
}

// Row `row` of the staged tile is the new token: rotate k, then write k and v to the pool (for later steps) and into
// the staged tile (for this step).  Call after the tile's TMA copy has completed, so nothing is written concurrently
// with the async proxy; ends with __syncwarp().
template <typename T, int D>
static __device__ __forceinline__ void append_patch(const AppendRows& pre, const FusedAppend<T>& fa, int Hkv, int kvh,
                                                    int row, int lane, uint32_t k_tile, uint32_t v_tile) {
  constexpr int CH = D / 8;
  const int64_t row_elems = ((pre.slot * Hkv + kvh) * kTok + row) * static_cast<int64_t>(D);
  if (lane < CH) {
    const int c = lane;
    union { int4 raw; T e[8]; } x, xp, cv, sv, o;
    x.raw = pre.x;
    xp.raw = pre.p;
    cv.raw = pre.c;
    sv.raw = pre.s;
    const bool first_half = c < CH / 2;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float xf = Cvt<T>::to_f32(x.e[e]);
      const float pf = first_half ? -Cvt<T>::to_f32(xp.e[e]) : Cvt<T>::to_f32(xp.e[e]);
      const T a = Cvt<T>::from_f32(__fmul_rn(xf, Cvt<T>::to_f32(cv.e[e])));
      const T bb = Cvt<T>::from_f32(__fmul_rn(pf, Cvt<T>::to_f32(sv.e[e])));
      o.e[e] = Cvt<T>::from_f32(__fadd_rn(Cvt<T>::to_f32(a), Cvt<T>::to_f32(bb)));
    }
    *reinterpret_cast<int4*>(fa.k_pool + row_elems + 8 * c) = o.raw;
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(k_tile + tile_off(row, c)), "r"(o.raw.x), "r"(o.raw.y),
                 "r"(o.raw.z), "r"(o.raw.w)
                 : "memory");
  } else if (lane >= 16 && lane < 16 + CH) {
    const int c = lane - 16;
    const int4 raw = pre.x;
    *reinterpret_cast<int4*>(fa.v_pool + row_elems + 8 * c) = raw;
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(v_tile + tile_off(row, c)), "r"(raw.x), "r"(raw.y),
                 "r"(raw.z), "r"(raw.w)
                 : "memory");
  }
  __syncwarp();
}


// Rows `first_row`..15 of a staged V tile <- 0.  The last tile of a sequence is staged whole, but its rows past the
// context length belong to nobody: their probabilities are exactly 0, and 0 x (whatever the pool holds there, possibly
// NaN / Inf in a pool that was never zero-filled) must stay 0 in the P V product.  Ends with __syncwarp().
template <int D>
static __device__ __forceinline__ void zero_v_tail(uint32_t v_tile, int first_row, int lane) {
  constexpr int CH = D / 8;  // 16-byte chunks per row
  for (int idx = first_row * CH + lane; idx < kTok * CH; idx += 32) {
    const int row = idx / CH, c = idx - row * CH;
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(v_tile + tile_off(row, c)), "r"(0) : "memory");
  }
  __syncwarp();
}

// 2D tensor map over a pool viewed as [max_blocks * Hkv * 16 rows, D] with a 64 x 16 box and 128-byte swizzle;
// cached per (pointer, rows, D, dtype).  Defined in b200_attn_tma.cu.
int get_tensor_map(const void* pool, uint64_t rows, int D, int dtype, CUtensorMap* out);
bool tensor_maps_available();

}  // namespace tma
}  // namespace b200
