// Paged GQA decode attention — shape-generic variant and the split-K combine kernel.
//
// This variant takes any storage dtype (f16 / bf16 / f32), any block size, head_dim in {32, 64, 128, 256} and
// group sizes up to 16.  It is the numerically plain formulation (fp32 everywhere after the load) and exists
// for the shapes the TMA + tensor-core variant (b200_attn_tma.cu) does not cover, and as its on-device
// cross-check.  One CTA per (split, kv head, sequence); each of its 4 warps walks 32 tokens at a time:
//   scores : lane = token, the lane reads its K row with 16-byte loads, q comes from shared memory (broadcast)
//   softmax: running max / sum per q head, warp-shuffle reductions
//   PV     : lane = head_dim slice (d = lane + 32 i), probabilities broadcast from shared memory
// Partials (m, l, acc) of the warps are merged in shared memory; with num_splits > 1 the CTA writes its partial
// to the workspace and `attn_combine_kernel` merges the splits.
#include <algorithm>
#include <cmath>

#include "b200_attn.cuh"
#include "b200_common.cuh"

namespace b200 {

constexpr int kGenWarps = 4;

template <typename T, int GP, int DPL>
__global__ void __launch_bounds__(kGenWarps * 32)
paged_attn_generic_kernel(T* __restrict__ out, const T* __restrict__ q, int64_t q_row_stride,
                          const T* __restrict__ k_pool, const T* __restrict__ v_pool,
                          const int32_t* __restrict__ block_tables, int table_stride,
                          const int32_t* __restrict__ context_lens, int Hq, int Hkv, int G, int block_size,
                          int64_t max_blocks, float scale, int blocks_per_split, int num_splits, float* __restrict__ ws_acc,
                          float* __restrict__ ws_ml, int round_scores) {
  constexpr int D = DPL * 32;
  constexpr int VEC = 16 / sizeof(T);
  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* q_s = reinterpret_cast<float*>(smem_raw);                 // [GP][D]
  float* p_s = q_s + GP * D;                                       // [W][GP][32]
  float* acc_s = p_s + kGenWarps * GP * 32;                        // [W][GP][D]
  float* m_s = acc_s + kGenWarps * GP * D;                         // [W][GP]
  float* l_s = m_s + kGenWarps * GP;                               // [W][GP]
  int64_t* row_s = reinterpret_cast<int64_t*>(l_s + kGenWarps * GP);  // [W][32]

  const int ctx = min(context_lens[b], table_stride * block_size);  // never walk past the end of the table row
  const int t_begin = min(ctx, split * blocks_per_split * block_size);
  const int t_end = (split == num_splits - 1) ? ctx : min(ctx, (split + 1) * blocks_per_split * block_size);

  for (int i = threadIdx.x; i < GP * D; i += blockDim.x) {
    const int g = i / D, d = i - g * D;
    q_s[i] = g < G ? Cvt<T>::to_f32(q[b * q_row_stride + static_cast<int64_t>(kvh * G + g) * D + d]) : 0.f;
  }
  __syncthreads();

  float m[GP], l[GP], acc[GP][DPL];
#pragma unroll
  for (int g = 0; g < GP; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int i = 0; i < DPL; ++i) acc[g][i] = 0.f;
  }

  const int32_t* table = block_tables + static_cast<int64_t>(b) * table_stride;
  float* my_p = p_s + warp * GP * 32;
  int64_t* my_row = row_s + warp * 32;

  for (int base = t_begin + warp * 32; base < t_end; base += kGenWarps * 32) {
    const int t = base + lane;
    const bool valid = t < t_end;
    float s[GP];
#pragma unroll
    for (int g = 0; g < GP; ++g) s[g] = 0.f;
    int64_t row = 0;
    int64_t slot = 0;
    if (valid) slot = table[t / block_size];
    // a slot outside the pool (corrupt block table) is treated like a masked token instead of being dereferenced
    const bool in_pool = slot >= 0 && slot < max_blocks;
    if (valid && in_pool) {
      row = ((slot * Hkv + kvh) * block_size + (t % block_size)) * D;
      const T* kp = k_pool + row;
#pragma unroll 2
      for (int d0 = 0; d0 < D; d0 += VEC) {
        const int4 raw = *reinterpret_cast<const int4*>(kp + d0);
        const T* kv = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          const float kf = Cvt<T>::to_f32(kv[e]);
#pragma unroll
          for (int g = 0; g < GP; ++g) s[g] = fmaf(q_s[g * D + d0 + e], kf, s[g]);
        }
      }
    }
    my_row[lane] = row;
#pragma unroll
    for (int g = 0; g < GP; ++g) {
      float sv = s[g];
      if (round_scores) {
        sv = Cvt<T>::to_f32(Cvt<T>::from_f32(sv));
        sv = Cvt<T>::to_f32(Cvt<T>::from_f32(sv * scale));
      } else {
        sv *= scale;
      }
      sv = (valid && in_pool) ? sv : -INFINITY;
      const float m_new = fmaxf(m[g], warp_max(sv));
      const float p = (valid && in_pool && m_new > -INFINITY) ? __expf(sv - m_new) : 0.f;
      const float corr = (m[g] == -INFINITY) ? 0.f : __expf(m[g] - m_new);
      l[g] = l[g] * corr + warp_sum(p);
      m[g] = m_new;
#pragma unroll
      for (int i = 0; i < DPL; ++i) acc[g][i] *= corr;
      my_p[g * 32 + lane] = p;
    }
    __syncwarp();
    const int cnt = min(32, t_end - base);
    for (int j = 0; j < cnt; ++j) {
      const T* vp = v_pool + my_row[j];
      float vf[DPL];
#pragma unroll
      for (int i = 0; i < DPL; ++i) vf[i] = Cvt<T>::to_f32(vp[lane + 32 * i]);
#pragma unroll
      for (int g = 0; g < GP; ++g) {
        const float p = my_p[g * 32 + j];
#pragma unroll
        for (int i = 0; i < DPL; ++i) acc[g][i] = fmaf(p, vf[i], acc[g][i]);
      }
    }
    __syncwarp();
  }

  // merge the warps
#pragma unroll
  for (int g = 0; g < GP; ++g) {
    if (lane == 0) {
      m_s[warp * GP + g] = m[g];
      l_s[warp * GP + g] = l[g];
    }
#pragma unroll
    for (int i = 0; i < DPL; ++i) acc_s[(warp * GP + g) * D + lane + 32 * i] = acc[g][i];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
    const int g = i / D, d = i - g * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kGenWarps; ++w) M = fmaxf(M, m_s[w * GP + g]);
    float o = 0.f, L = 0.f;
#pragma unroll
    for (int w = 0; w < kGenWarps; ++w) {
      const float mw = m_s[w * GP + g];
      const float wgt = (mw == -INFINITY) ? 0.f : __expf(mw - M);
      o = fmaf(wgt, acc_s[(w * GP + g) * D + d], o);
      L = fmaf(wgt, l_s[w * GP + g], L);
    }
    const int hq = kvh * G + g;
    if (num_splits == 1) {
      out[(static_cast<int64_t>(b) * Hq + hq) * D + d] = Cvt<T>::from_f32(L > 0.f ? o / L : 0.f);
    } else {
      const int64_t pidx = (static_cast<int64_t>(b) * Hq + hq) * num_splits + split;
      ws_acc[pidx * D + d] = o;
      if (d == 0) {
        ws_ml[pidx * 2 + 0] = M;
        ws_ml[pidx * 2 + 1] = L;
      }
    }
  }
}

// Merge split-K partials: one CTA per (sequence, q head) row.  The weights exp(m_s - M) are computed once per row
// into shared memory; then thread d streams acc[s][d] over all splits with four independent accumulators, so a row
// of S partials is S coalesced 4*D-byte reads in flight rather than a serial chain (the first version walked the
// splits with one warp per row and cost 13-26 us at 9-28 splits; profiles/r01_cascade_launches.csv).
constexpr int kCombineThreads = 128;
constexpr int kCombineMaxSplits = 256;

template <typename T>
__global__ void __launch_bounds__(kCombineThreads)
attn_combine_kernel(T* __restrict__ out, const float* __restrict__ ws_acc, const float* __restrict__ ws_ml,
                    int D, int num_splits) {
  __shared__ float w_s[kCombineMaxSplits];
  __shared__ float red[2][kCombineThreads / 32];
  // Programmatic dependent launch: this grid may be scheduled while the attention kernel that produces the partials
  // is still running (its CTAs signal as they start); it waits here until that grid has completed and its writes are
  // visible.  Launched without the attribute the instruction does nothing.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* ml = ws_ml + static_cast<int64_t>(row) * num_splits * 2;

  float m_loc = -INFINITY;
  for (int s = tid; s < num_splits; s += kCombineThreads) m_loc = fmaxf(m_loc, ml[2 * s]);
  m_loc = warp_max(m_loc);
  if (lane == 0) red[0][warp] = m_loc;
  __syncthreads();
  float M = red[0][0];
#pragma unroll
  for (int w = 1; w < kCombineThreads / 32; ++w) M = fmaxf(M, red[0][w]);

  float l_loc = 0.f;
  for (int s = tid; s < num_splits; s += kCombineThreads) {
    const float ms = ml[2 * s];
    const float w = (ms == -INFINITY) ? 0.f : __expf(ms - M);
    w_s[s] = w;
    l_loc = fmaf(w, ml[2 * s + 1], l_loc);
  }
  l_loc = warp_sum(l_loc);
  if (lane == 0) red[1][warp] = l_loc;
  __syncthreads();
  float L = 0.f;
#pragma unroll
  for (int w = 0; w < kCombineThreads / 32; ++w) L += red[1][w];
  const float inv = L > 0.f ? 1.f / L : 0.f;

  const float* acc = ws_acc + static_cast<int64_t>(row) * num_splits * D;
  for (int d = tid; d < D; d += kCombineThreads) {
    float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
    int s = 0;
    for (; s + 4 <= num_splits; s += 4) {
      o0 = fmaf(w_s[s], acc[static_cast<int64_t>(s) * D + d], o0);
      o1 = fmaf(w_s[s + 1], acc[static_cast<int64_t>(s + 1) * D + d], o1);
      o2 = fmaf(w_s[s + 2], acc[static_cast<int64_t>(s + 2) * D + d], o2);
      o3 = fmaf(w_s[s + 3], acc[static_cast<int64_t>(s + 3) * D + d], o3);
    }
    for (; s < num_splits; ++s) o0 = fmaf(w_s[s], acc[static_cast<int64_t>(s) * D + d], o0);
    out[static_cast<int64_t>(row) * D + d] = Cvt<T>::from_f32(((o0 + o1) + (o2 + o3)) * inv);
  }
}

template <typename T>
int launch_combine(void* out, const float* ws_acc, const float* ws_ml, int rows, int D, int num_splits,
                   cudaStream_t stream) {
  if (num_splits > kCombineMaxSplits)
    return fail(B200_ESHAPE, "b200_paged_decode_attn: more than %d splits cannot be merged", kCombineMaxSplits);
  // launched as a programmatic dependent of the attention kernel just before it on the stream: the launch latency of
  // this small grid overlaps that kernel's tail instead of following it (config 5 in the step: 68.7 -> 66.5 us per
  // layer; 9-way split-K on the unshared tables 162.6 -> 158.4 us).  Fetching the first partial rows ahead of the
  // weights was tried as well and measured slower (66.0 -> 67.5 us, 30.8 -> 34.9 us for one 32k sequence).
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(rows));
  cfg.blockDim = dim3(kCombineThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  T* out_t = static_cast<T*>(out);
  cudaError_t e = cudaLaunchKernelEx(&cfg, attn_combine_kernel<T>, out_t, ws_acc, ws_ml, D, num_splits);
  if (e != cudaSuccess) return fail(static_cast<int>(e), "b200_paged_decode_attn(combine): %s", cudaGetErrorString(e));
  return check_launch("b200_paged_decode_attn(combine)");
}

int attn_combine(void* out, const float* ws_acc, const float* ws_ml, int rows, int D, int num_splits, int dtype,
                 cudaStream_t stream) {
  switch (dtype) {
    case B200_DTYPE_F16:
      return launch_combine<__half>(out, ws_acc, ws_ml, rows, D, num_splits, stream);
    case B200_DTYPE_BF16:
      return launch_combine<__nv_bfloat16>(out, ws_acc, ws_ml, rows, D, num_splits, stream);
    default:
      return launch_combine<float>(out, ws_acc, ws_ml, rows, D, num_splits, stream);
  }
}


template <typename T, int GP, int DPL>
int launch_generic(const AttnArgs& a, cudaStream_t stream) {
  constexpr int D = DPL * 32;
  const size_t smem = sizeof(float) * (GP * D + kGenWarps * GP * 32 + kGenWarps * GP * D + 2 * kGenWarps * GP) +
                      sizeof(int64_t) * kGenWarps * 32;
  auto kern = paged_attn_generic_kernel<T, GP, DPL>;
  if (smem > 48 * 1024) {  // opt in per launch: the attribute is per device and costs no stream work
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return fail(static_cast<int>(e), "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  }
  dim3 grid(a.num_splits, a.Hkv, a.B);
  kern<<<grid, kGenWarps * 32, smem, stream>>>(
      static_cast<T*>(a.out), static_cast<const T*>(a.q), a.q_row_stride, static_cast<const T*>(a.k_pool),
      static_cast<const T*>(a.v_pool), a.block_tables, a.table_stride, a.context_lens, a.Hq, a.Hkv, a.Hq / a.Hkv,
      a.block_size, a.max_blocks, a.scale, a.blocks_per_split, a.num_splits, a.ws_acc, a.ws_ml, a.round_scores);
  return check_launch("b200_paged_decode_attn(generic)");
}

template <typename T, int GP>
int dispatch_d(const AttnArgs& a, cudaStream_t s) {
  switch (a.D) {
    case 32:
      return launch_generic<T, GP, 1>(a, s);
    case 64:
      return launch_generic<T, GP, 2>(a, s);
    case 128:
      return launch_generic<T, GP, 4>(a, s);
    case 256:
      if constexpr (GP <= 8) return launch_generic<T, GP, 8>(a, s);
      [[fallthrough]];
    default:
      return fail(B200_ESHAPE, "b200_paged_decode_attn: head_dim %d with group %d not supported", a.D, a.Hq / a.Hkv);
  }
}

template <typename T>
int dispatch_g(const AttnArgs& a, cudaStream_t s) {
  const int G = a.Hq / a.Hkv;
  if (G == 1) return dispatch_d<T, 1>(a, s);
  if (G == 2) return dispatch_d<T, 2>(a, s);
  if (G <= 4) return dispatch_d<T, 4>(a, s);
  if (G <= 8) return dispatch_d<T, 8>(a, s);
  if (G <= 16) return dispatch_d<T, 16>(a, s);
  return fail(B200_ESHAPE, "b200_paged_decode_attn: group size %d > 16 not supported", G);
}

int attn_generic(const AttnArgs& a, int dtype, cudaStream_t s) {
  switch (dtype) {
    case B200_DTYPE_F16:
      return dispatch_g<__half>(a, s);
    case B200_DTYPE_BF16:
      return dispatch_g<__nv_bfloat16>(a, s);
    case B200_DTYPE_F32:
      return dispatch_g<float>(a, s);
    default:
      return fail(B200_EINVAL, "b200_paged_decode_attn: unsupported dtype %d", dtype);
  }
}

}  // namespace b200
