// Fused element-wise pieces of the decoder layer that ModelShard.forward calls once per owned layer
// (reference: worker/distributed/model_shard.py:205-222 → the HF Llama layer): residual-add + RMSNorm,
// RoPE + paged KV append, SiLU·up.  In torch each of these is 5-10 tiny launches; at decode (64..128 rows) the
// step is launch-bound around the GEMMs, so each becomes ONE launch here.
//
// Rounding contract: every kernel reproduces the rounding points of the eager torch recipe in the storage
// dtype T (each torch op computes in fp32 and rounds its result to T):
//   add+norm : h = T(x + d);  y = T(w * T(float(h) * rsqrt(mean(float(h)^2) + eps)))
//   rope     : out = T( T(x * cos) + T(rot(x) * sin) ),  rot(x) = cat(-x[D/2:], x[:D/2])
//   silu·up  : out = T( T(silu(float(g))) * u ),          silu(x) = x / (1 + exp(-x))
// Products of two 16-bit floats are exact in fp32 and the fp32 adds are done with explicit round-to-nearest
// intrinsics (no FMA contraction), so RoPE and SiLU·up are bit-identical to torch; the norm differs from torch
// only through the summation order of the fp32 mean (≤ 1 ulp of fp32 before the rounding to T).
#include <algorithm>
#include <type_traits>

#include "b200_attn.cuh"
#include "b200_common.cuh"

namespace b200 {

namespace {

template <typename T>
struct Vec16 {
  static constexpr int N = 16 / sizeof(T);
  union {
    int4 raw;
    T e[16 / sizeof(T)];
  };
};

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : 0.f;
  t = warp_sum(t);
  return t;  // every thread holds the total
}

// ---- residual add + RMSNorm: one CTA per row ------------------------------------------------------------
constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 8;  // vectors cached in registers per thread: hidden <= 256 * 8 * (16 / sizeof(T))

// NV = 16-byte vectors cached per thread (ceil(hidden / VN / 256)): sizing the register cache to the row keeps the
// kernel at ~40 registers for hidden = 4096 instead of ~90, i.e. 6 instead of 2 CTAs per SM when thousands of rows
// (a prompt pass) are in flight.  Arithmetic and its order do not depend on NV.
template <typename T, int NV>
__global__ void __launch_bounds__(kNormThreads)
add_rmsnorm_kernel(T* __restrict__ out, T* __restrict__ resid_out, const T* __restrict__ x, const T* __restrict__ delta,
                   const T* __restrict__ weight, float eps, int hidden, int64_t x_stride, int64_t d_stride) {
  constexpr int VN = Vec16<T>::N;
  __shared__ float red[32];
  const int row = blockIdx.x;
  const T* xr = x + row * x_stride;
  const T* dr = delta ? delta + row * d_stride : nullptr;
  const int nvec = hidden / VN;
  Vec16<T> cache[NV], wcache[NV];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {  // weights first: their latency hides behind the row loads + reduction
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) wcache[i].raw = *reinterpret_cast<const int4*>(weight + v * VN);
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      cache[i].raw = *reinterpret_cast<const int4*>(xr + v * VN);
      if (dr) {
        Vec16<T> dv;
        dv.raw = *reinterpret_cast<const int4*>(dr + v * VN);
#pragma unroll
        for (int e = 0; e < VN; ++e)
          cache[i].e[e] = Cvt<T>::from_f32(__fadd_rn(Cvt<T>::to_f32(cache[i].e[e]), Cvt<T>::to_f32(dv.e[e])));
        if (resid_out) *reinterpret_cast<int4*>(resid_out + static_cast<int64_t>(row) * hidden + v * VN) = cache[i].raw;
      }
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const float f = Cvt<T>::to_f32(cache[i].e[e]);
        ss = fmaf(f, f, ss);
      }
    }
  }
  const float total = block_sum(ss, red);
  const float inv = rsqrtf(total * (1.0f / static_cast<float>(hidden)) + eps);  // torch's mean multiplies by 1/N
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      Vec16<T> ov;
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const T n = Cvt<T>::from_f32(__fmul_rn(Cvt<T>::to_f32(cache[i].e[e]), inv));
        ov.e[e] = Cvt<T>::from_f32(__fmul_rn(Cvt<T>::to_f32(wcache[i].e[e]), Cvt<T>::to_f32(n)));
      }
      *reinterpret_cast<int4*>(out + static_cast<int64_t>(row) * hidden + v * VN) = ov.raw;
    }
  }
}

// ---- RoPE on q (in place) and k, then K/V scatter into the pools -----------------------------------------
// Grid (groups of kRopeToks tokens, chunks of 128 work items); a thread handles one 16-byte vector of the first half of
// a head together with its partner vector in the second half (rotate-half pairs d and d + D/2), for every token of
// the CTA's group.  The tokens' slots are turned into pool row offsets once per CTA (shared memory), so the per-thread
// work is loads, the rotation and stores; a CTA per token (round 1) spent its time on launch overhead and on two
// 64-bit divisions per thread (0.69 of the copy peak on a 16k-token prompt chunk).
constexpr int kRopeToks = 4;

template <typename T>
__global__ void __launch_bounds__(128)
rope_append_kernel(T* __restrict__ q, int64_t q_stride, const T* k, int64_t k_stride, T* k_rot,
                   const T* __restrict__ v, int64_t v_stride, const T* __restrict__ cos_t, const T* __restrict__ sin_t,
                   T* __restrict__ k_pool, T* __restrict__ v_pool, const int64_t* __restrict__ slot_mapping, int n_tok,
                   int Hq, int Hkv, int D, int block_size, int64_t max_blocks) {
  constexpr int VN = Vec16<T>::N;
  __shared__ int64_t row0_s[kRopeToks];  // (block * Hkv * block_size + offset in block) of the token, -1 = no store
  const int tok0 = blockIdx.x * kRopeToks;
  if (threadIdx.x < kRopeToks) {
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
  const int per_head = D / 2 / VN;  // work items per head (each covers both halves)
  const int n_items = (Hq + 2 * Hkv) * per_head;
  const int it = blockIdx.y * blockDim.x + threadIdx.x;  // one (head, vector pair) per thread
  if (it >= n_items) return;
  const int head = it / per_head;
  const int d0 = (it - head * per_head) * VN;
  const int n_here = min(kRopeToks, n_tok - tok0);
  if (head < Hq + Hkv) {
    const bool is_q = head < Hq;
    const int hh = is_q ? head : head - Hq;
    Vec16<T> lo[kRopeToks], hi[kRopeToks], c_lo[kRopeToks], c_hi[kRopeToks], s_lo[kRopeToks], s_hi[kRopeToks];
#pragma unroll
    for (int u = 0; u < kRopeToks; ++u) {
      if (u >= n_here) break;
      const int64_t tok = tok0 + u;
      const T* src = is_q ? q + tok * q_stride + static_cast<int64_t>(hh) * D : k + tok * k_stride + static_cast<int64_t>(hh) * D;
      const T* cs = cos_t + tok * D;
      const T* sn = sin_t + tok * D;
      lo[u].raw = *reinterpret_cast<const int4*>(src + d0);
      hi[u].raw = *reinterpret_cast<const int4*>(src + D / 2 + d0);
      c_lo[u].raw = *reinterpret_cast<const int4*>(cs + d0);
      c_hi[u].raw = *reinterpret_cast<const int4*>(cs + D / 2 + d0);
      s_lo[u].raw = *reinterpret_cast<const int4*>(sn + d0);
      s_hi[u].raw = *reinterpret_cast<const int4*>(sn + D / 2 + d0);
    }
#pragma unroll
    for (int u = 0; u < kRopeToks; ++u) {
      if (u >= n_here) break;
      const int64_t tok = tok0 + u;
      Vec16<T> o_lo, o_hi;
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const float xl = Cvt<T>::to_f32(lo[u].e[e]), xh = Cvt<T>::to_f32(hi[u].e[e]);
        // first half: x_lo * cos - x_hi * sin ; second half: x_hi * cos + x_lo * sin
        const T a = Cvt<T>::from_f32(__fmul_rn(xl, Cvt<T>::to_f32(c_lo[u].e[e])));
        const T b = Cvt<T>::from_f32(__fmul_rn(-xh, Cvt<T>::to_f32(s_lo[u].e[e])));
        o_lo.e[e] = Cvt<T>::from_f32(__fadd_rn(Cvt<T>::to_f32(a), Cvt<T>::to_f32(b)));
        const T c = Cvt<T>::from_f32(__fmul_rn(xh, Cvt<T>::to_f32(c_hi[u].e[e])));
        const T d = Cvt<T>::from_f32(__fmul_rn(xl, Cvt<T>::to_f32(s_hi[u].e[e])));
        o_hi.e[e] = Cvt<T>::from_f32(__fadd_rn(Cvt<T>::to_f32(c), Cvt<T>::to_f32(d)));
      }
      if (is_q) {
        T* dst = q + tok * q_stride + static_cast<int64_t>(hh) * D;
        *reinterpret_cast<int4*>(dst + d0) = o_lo.raw;
        *reinterpret_cast<int4*>(dst + D / 2 + d0) = o_hi.raw;
      } else {
        if (k_rot != nullptr) {  // prompt pass: the rotated k is also needed densely (may alias k: same thread, read first)
          T* dense = k_rot + tok * k_stride + static_cast<int64_t>(hh) * D;
          *reinterpret_cast<int4*>(dense + d0) = o_lo.raw;
          *reinterpret_cast<int4*>(dense + D / 2 + d0) = o_hi.raw;
        }
        if (row0_s[u] >= 0) {
          T* dst = k_pool + (row0_s[u] + static_cast<int64_t>(hh) * block_size) * D;
          *reinterpret_cast<int4*>(dst + d0) = o_lo.raw;
          *reinterpret_cast<int4*>(dst + D / 2 + d0) = o_hi.raw;
        }
      }
    }
  } else {
    const int h = head - Hq - Hkv;
    int4 a[kRopeToks], b2[kRopeToks];
#pragma unroll
    for (int u = 0; u < kRopeToks; ++u) {
      if (u >= n_here || row0_s[u] < 0) continue;
      const T* src = v + (tok0 + u) * v_stride + static_cast<int64_t>(h) * D;
      a[u] = *reinterpret_cast<const int4*>(src + d0);
      b2[u] = *reinterpret_cast<const int4*>(src + D / 2 + d0);
    }
#pragma unroll
    for (int u = 0; u < kRopeToks; ++u) {
      if (u >= n_here || row0_s[u] < 0) continue;
      T* dst = v_pool + (row0_s[u] + static_cast<int64_t>(h) * block_size) * D;
      *reinterpret_cast<int4*>(dst + d0) = a[u];
      *reinterpret_cast<int4*>(dst + D / 2 + d0) = b2[u];
    }
  }
}

// ---- SiLU(gate) * up -------------------------------------------------------------------------------------
// Grid (column chunks, rows): a thread takes one 16-byte vector of one row, rows beyond gridDim.y are walked with a
// stride — no per-element index division (the flat-index version spent ~40 % of its instructions on 64-bit div/mod).
template <typename T>
__global__ void __launch_bounds__(256)
silu_mul_kernel(T* __restrict__ out, const T* __restrict__ gate, const T* __restrict__ up, int64_t rows, int cols,
                int64_t g_stride, int64_t u_stride) {
  constexpr int VN = Vec16<T>::N;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * VN;
  if (c >= cols) return;
  for (int64_t r = blockIdx.y; r < rows; r += gridDim.y) {
    Vec16<T> g, u, o;
    g.raw = *reinterpret_cast<const int4*>(gate + r * g_stride + c);
    u.raw = *reinterpret_cast<const int4*>(up + r * u_stride + c);
#pragma unroll
    for (int e = 0; e < VN; ++e) {
      const float x = Cvt<T>::to_f32(g.e[e]);
      const T s = Cvt<T>::from_f32(x / (1.0f + expf(-x)));
      o.e[e] = Cvt<T>::from_f32(__fmul_rn(Cvt<T>::to_f32(s), Cvt<T>::to_f32(u.e[e])));
    }
    *reinterpret_cast<int4*>(out + r * cols + c) = o.raw;
  }
}

// ---- greedy token: row-wise argmax of the logits --------------------------------------------------------------
// One CTA per row, 16-byte loads.  Same answer as torch.argmax: the largest value, the smallest index among equal
// values, NaN counts as larger than everything (first NaN wins).
struct Best {
  float v;
  int i;
};
__device__ __forceinline__ bool better(float v, int i, const Best& b) {
  const bool vn = v != v, bn = b.v != b.v;
  if (bn) return vn && i < b.i;
  return vn || v > b.v || (v == b.v && i < b.i);
}
constexpr int kArgmaxThreads = 1024;

// Vocabulary-parallel form: every rank reduces its own column slice of the logits and the N partial answers are
// merged by ONE max-reduction over int64 keys.  key = (order-preserving image of the value) << 32 | ~(global column),
// top bit flipped so that signed comparison orders keys like (value, -column): the maximum key is the largest value
// and, among equal values, the smallest column — the same winner torch.argmax picks on the whole row.
__device__ __forceinline__ int64_t argmax_key(float v, int64_t col) {
  uint32_t ord;
  if (v != v) {
    ord = 0xffffffffu;  // NaN beats everything
  } else {
    if (v == 0.f) v = 0.f;  // -0 and +0 compare equal in torch.argmax
    const uint32_t b = __float_as_uint(v);
    ord = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  }
  const uint64_t key = (static_cast<uint64_t>(ord) << 32) | (0xffffffffu - static_cast<uint32_t>(col));
  return static_cast<int64_t>(key ^ 0x8000000000000000ull);
}

__global__ void argmax_keys_decode_kernel(int64_t* __restrict__ out, const int64_t* __restrict__ keys, int rows) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rows) {
    const uint64_t key = static_cast<uint64_t>(keys[r]) ^ 0x8000000000000000ull;
    out[r] = static_cast<int64_t>(0xffffffffu - static_cast<uint32_t>(key & 0xffffffffu));
  }
}

template <typename T>
__global__ void __launch_bounds__(kArgmaxThreads)
argmax_rows_kernel(int64_t* __restrict__ out, const T* __restrict__ x, int cols, int64_t row_stride, bool vec_ok,
                   int64_t col_offset, bool as_key) {
  constexpr int VN = Vec16<T>::N;
  const T* row = x + static_cast<int64_t>(blockIdx.x) * row_stride;
  Best best{-INFINITY, 0x7fffffff};
  int done = 0;
  if (vec_ok) {
    const int nvec = cols / VN;
    for (int v = threadIdx.x; v < nvec; v += kArgmaxThreads) {
      Vec16<T> a;
      a.raw = *reinterpret_cast<const int4*>(row + static_cast<int64_t>(v) * VN);
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const float f = Cvt<T>::to_f32(a.e[e]);
        if (better(f, v * VN + e, best)) best = Best{f, v * VN + e};
      }
    }
    done = nvec * VN;
  }
  for (int c = done + threadIdx.x; c < cols; c += kArgmaxThreads) {
    const float f = Cvt<T>::to_f32(row[c]);
    if (better(f, c, best)) best = Best{f, c};
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best.v, off);
    const int oi = __shfl_xor_sync(0xffffffffu, best.i, off);
    if (better(ov, oi, best)) best = Best{ov, oi};
  }
  __shared__ float sv[kArgmaxThreads / 32];
  __shared__ int si[kArgmaxThreads / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    sv[warp] = best.v;
    si[warp] = best.i;
  }
  __syncthreads();
  if (warp == 0) {
    best = Best{sv[lane], si[lane]};
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best.v, off);
      const int oi = __shfl_xor_sync(0xffffffffu, best.i, off);
      if (better(ov, oi, best)) best = Best{ov, oi};
    }
    if (lane == 0) out[blockIdx.x] = as_key ? argmax_key(best.v, col_offset + best.i) : col_offset + best.i;
  }
}

template <typename F>
int by_dtype(int dtype, const char* name, F&& f) {
  switch (dtype) {
    case B200_DTYPE_F16:
      return f(static_cast<__half*>(nullptr));
    case B200_DTYPE_BF16:
      return f(static_cast<__nv_bfloat16*>(nullptr));
    case B200_DTYPE_F32:
      return f(static_cast<float*>(nullptr));
    default:
      return fail(B200_EINVAL, "%s: unsupported dtype %d", name, dtype);
  }
}

}  // namespace

int add_rmsnorm(void* out, void* resid_out, const void* x, const void* delta, const void* weight, float eps, int rows,
                int hidden, int64_t x_stride, int64_t d_stride, int dtype, cudaStream_t s) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "b200_add_rmsnorm: unsupported dtype %d", dtype);
  if (rows < 0 || hidden <= 0) return fail(B200_EINVAL, "b200_add_rmsnorm: bad sizes");
  if (rows == 0) return 0;
  if (!out || !x || !weight) return fail(B200_EINVAL, "b200_add_rmsnorm: null pointer");
  const int vn = 16 / es;
  if (hidden % vn != 0 || hidden / vn > kNormThreads * kNormMaxVec || (x_stride * es) % 16 != 0 ||
      (delta && (d_stride * es) % 16 != 0) || !aligned16(out) || !aligned16(x) || !aligned16(weight) ||
      (delta && !aligned16(delta)) || (resid_out && !aligned16(resid_out)))
    return fail(B200_ESHAPE, "b200_add_rmsnorm: hidden=%d / alignment not supported", hidden);
  return by_dtype(dtype, "b200_add_rmsnorm", [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    const int nv = (hidden / (16 / static_cast<int>(sizeof(T))) + kNormThreads - 1) / kNormThreads;
    auto launch = [&](auto kern) {
      kern<<<rows, kNormThreads, 0, s>>>(static_cast<T*>(out), static_cast<T*>(resid_out), static_cast<const T*>(x),
                                         static_cast<const T*>(delta), static_cast<const T*>(weight), eps, hidden, x_stride,
                                         d_stride);
    };
    if (nv <= 1) launch(add_rmsnorm_kernel<T, 1>);
    else if (nv <= 2) launch(add_rmsnorm_kernel<T, 2>);
    else if (nv <= 4) launch(add_rmsnorm_kernel<T, 4>);
    else launch(add_rmsnorm_kernel<T, kNormMaxVec>);
    return check_launch("b200_add_rmsnorm");
  });
}

int rope_kv_append(void* q, int64_t q_stride, const void* k, int64_t k_stride, void* k_rot, const void* v,
                   int64_t v_stride, const void* cos_t, const void* sin_t, void* k_pool, void* v_pool, const int64_t* slot_mapping,
                   int n_tok, int Hq, int Hkv, int D, int block_size, int64_t max_blocks, int dtype, cudaStream_t s) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "b200_rope_kv_append: unsupported dtype %d", dtype);
  if (n_tok < 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || block_size <= 0 || max_blocks <= 0)
    return fail(B200_EINVAL, "b200_rope_kv_append: bad sizes");
  if (n_tok == 0) return 0;
  if (!q || !k || !v || !cos_t || !sin_t || !k_pool || !v_pool || !slot_mapping)
    return fail(B200_EINVAL, "b200_rope_kv_append: null pointer");
  const int vn = 16 / es;
  if ((D / 2) % vn != 0 || (q_stride * es) % 16 || (k_stride * es) % 16 || (v_stride * es) % 16 || !aligned16(q) ||
      !aligned16(k) || !aligned16(v) || !aligned16(cos_t) || !aligned16(sin_t) || !aligned16(k_pool) || !aligned16(v_pool))
    return fail(B200_ESHAPE, "b200_rope_kv_append: head_dim %d / alignment not supported", D);
  return by_dtype(dtype, "b200_rope_kv_append", [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    const int n_items = (Hq + 2 * Hkv) * ((D / 2) / (16 / static_cast<int>(sizeof(T))));
    const int threads = 128;
    dim3 grid((n_tok + kRopeToks - 1) / kRopeToks, (n_items + threads - 1) / threads);
    rope_append_kernel<T><<<grid, threads, 0, s>>>(static_cast<T*>(q), q_stride, static_cast<const T*>(k), k_stride,
                                                static_cast<T*>(k_rot), static_cast<const T*>(v), v_stride, static_cast<const T*>(cos_t),
                                                static_cast<const T*>(sin_t), static_cast<T*>(k_pool),
                                                static_cast<T*>(v_pool), slot_mapping, n_tok, Hq, Hkv, D, block_size,
                                                max_blocks);
    return check_launch("b200_rope_kv_append");
  });
}

int silu_mul(void* out, const void* gate, const void* up, int64_t rows, int cols, int64_t g_stride, int64_t u_stride,
             int dtype, cudaStream_t s) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "b200_silu_mul: unsupported dtype %d", dtype);
  if (rows < 0 || cols <= 0) return fail(B200_EINVAL, "b200_silu_mul: bad sizes");
  if (rows == 0) return 0;
  if (!out || !gate || !up) return fail(B200_EINVAL, "b200_silu_mul: null pointer");
  const int vn = 16 / es;
  if (cols % vn || (g_stride * es) % 16 || (u_stride * es) % 16 || !aligned16(out) || !aligned16(gate) || !aligned16(up))
    return fail(B200_ESHAPE, "b200_silu_mul: cols=%d / alignment not supported", cols);
  const dim3 grid(static_cast<unsigned>((cols / vn + 255) / 256), static_cast<unsigned>(std::min<int64_t>(rows, 32768)));
  return by_dtype(dtype, "b200_silu_mul", [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    silu_mul_kernel<T><<<grid, 256, 0, s>>>(static_cast<T*>(out), static_cast<const T*>(gate),
                                            static_cast<const T*>(up), rows, cols, g_stride, u_stride);
    return check_launch("b200_silu_mul");
  });
}

int argmax_rows(int64_t* out, const void* x, int64_t rows, int64_t cols, int64_t row_stride, int dtype, cudaStream_t s) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "b200_argmax_rows: unsupported dtype %d", dtype);
  if (rows < 0 || cols <= 0 || cols > 0x7fffffff || rows > 0x7fffffff || row_stride < cols)
    return fail(B200_EINVAL, "b200_argmax_rows: bad sizes");
  if (rows == 0) return 0;
  if (!out || !x) return fail(B200_EINVAL, "b200_argmax_rows: null pointer");
  const bool vec_ok = aligned16(x) && (row_stride * es) % 16 == 0;
  return by_dtype(dtype, "b200_argmax_rows", [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    argmax_rows_kernel<T><<<static_cast<unsigned>(rows), kArgmaxThreads, 0, s>>>(
        out, static_cast<const T*>(x), static_cast<int>(cols), row_stride, vec_ok, 0, false);
    return check_launch("b200_argmax_rows");
  });
}

int argmax_rows_keys(int64_t* keys, const void* x, int64_t rows, int64_t cols, int64_t row_stride, int64_t col_offset,
                     int dtype, cudaStream_t s) {
  const int es = dtype_size(dtype);
  if (es == 0) return fail(B200_EINVAL, "b200_argmax_rows_keys: unsupported dtype %d", dtype);
  if (rows < 0 || cols <= 0 || cols > 0x7fffffff || rows > 0x7fffffff || row_stride < cols || col_offset < 0 ||
      col_offset + cols > 0xffffffffll)
    return fail(B200_EINVAL, "b200_argmax_rows_keys: bad sizes");
  if (rows == 0) return 0;
  if (!keys || !x) return fail(B200_EINVAL, "b200_argmax_rows_keys: null pointer");
  const bool vec_ok = aligned16(x) && (row_stride * es) % 16 == 0;
  return by_dtype(dtype, "b200_argmax_rows_keys", [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    argmax_rows_kernel<T><<<static_cast<unsigned>(rows), kArgmaxThreads, 0, s>>>(
        keys, static_cast<const T*>(x), static_cast<int>(cols), row_stride, vec_ok, col_offset, true);
    return check_launch("b200_argmax_rows_keys");
  });
}

int argmax_keys_decode(int64_t* out, const int64_t* keys, int64_t rows, cudaStream_t s) {
  if (rows < 0 || rows > 0x7fffffff) return fail(B200_EINVAL, "b200_argmax_keys_decode: bad sizes");
  if (rows == 0) return 0;
  if (!out || !keys) return fail(B200_EINVAL, "b200_argmax_keys_decode: null pointer");
  argmax_keys_decode_kernel<<<static_cast<unsigned>((rows + 127) / 128), 128, 0, s>>>(out, keys, static_cast<int>(rows));
  return check_launch("b200_argmax_keys_decode");
}

}  // namespace b200
