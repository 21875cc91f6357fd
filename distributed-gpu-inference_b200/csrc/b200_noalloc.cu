// b200_paged_decode_attn_static — paged decode attention for callers that cannot allocate device memory.
//
// The regular entry point needs a caller-provided workspace whenever the context is cut into more than one
// partition (split-K).  Some hosts have no way to obtain one: a plugin running inside somebody else's serving
// loop where every cudaMalloc is forbidden because it synchronises the device, or a thin FFI binding (cgo / JNI /
// ctypes without torch) that only ever sees pointers handed to it.  For those callers the split-K partials live
// in a fixed array that is part of this library's image.
#include <mutex>

#include "b200_common.cuh"

// Scratch for the split-K partials.  Every GPU that loads the library holds these bytes until the process exits.
// Nothing in it has to start at zero: a call writes every partial it later reads, so stale contents of an
// earlier call are harmless.  The array is shared by all calls on a device — see the entry point's comment.
static __device__ __align__(256) unsigned char g_static_scratch[B200_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
void* g_scratch_addr[kMaxDevices] = {nullptr};  // device address of g_static_scratch, per device ordinal

// In host code `g_static_scratch` is only a handle for the symbol, not an address a kernel can use; the real
// device address differs per GPU and is looked up once per device, with that device current.
int scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return b200::fail(static_cast<int>(e), "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return b200::fail(B200_EINVAL, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return b200::fail(static_cast<int>(e), "cudaGetSymbolAddress(static scratch): %s", cudaGetErrorString(e));
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace

extern "C" int b200_paged_decode_attn_static(void* out, const void* q, int64_t q_row_stride, const void* k_pool,
                                             const void* v_pool, const int32_t* block_tables, int table_stride,
                                             const int32_t* context_lens, int B, int Hq, int Hkv, int D,
                                             int block_size, int64_t max_blocks, float scale, int max_context_len,
                                             int num_splits, int dtype, unsigned flags, void* stream) {
  if (num_splits < 0) return b200::fail(B200_EINVAL, "b200_paged_decode_attn_static: negative num_splits");
  if (num_splits == 0) num_splits = b200_paged_decode_attn_plan(B, Hq, Hkv, D, block_size, max_context_len, dtype);
  if (block_size > 0) {  // the regular entry point never uses more partitions than the context has blocks
    const int n_blocks = (max_context_len > 1 ? max_context_len - 1 : 0) / block_size + 1;
    if (num_splits > n_blocks) num_splits = n_blocks;
  }
  const size_t need = b200_paged_decode_attn_workspace(B, Hq, D, num_splits);
  if (need > static_cast<size_t>(B200_STATIC_SCRATCH_BYTES))
    return b200::fail(B200_ENOSPC,
                      "b200_paged_decode_attn_static: B=%d Hq=%d D=%d with %d splits needs %zu bytes of scratch, the "
                      "built-in array holds %d; use fewer splits or b200_paged_decode_attn with your own workspace",
                      B, Hq, D, num_splits, need, B200_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (need > 0) {
    const int rc = scratch_for_current_device(&scratch);
    if (rc != 0) return rc;
  }
  return b200_paged_decode_attn(out, q, q_row_stride, k_pool, v_pool, block_tables, table_stride, context_lens, B, Hq,
                                Hkv, D, block_size, max_blocks, scale, max_context_len, num_splits, scratch,
                                need > 0 ? static_cast<size_t>(B200_STATIC_SCRATCH_BYTES) : 0, dtype, flags, stream);
}
