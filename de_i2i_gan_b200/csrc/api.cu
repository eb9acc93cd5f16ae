// Library-wide plumbing of the C ABI: error reporting, tuning knobs, launch counter, TMA descriptor encoding.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "common.cuh"
#include "../../include/dei2i.h"

namespace dei2i {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return DEI2I_ERR_CUDA;
}

int g_tune_conv_stages = 0;
int g_tune_conv_bn = 0;
int g_tune_wgrad_bn = 0;
int g_tune_wgrad_waves = 0;
int g_tune_tma_ms = 0;
int g_tune_mma_issue = -1;   // conv_tma_kernel MMA issue loop: -1 = auto by N tile, 1 = warp-uniform with an elected lane, 0 = single thread
int g_tune_tma_pair = 3;     // CTA pairs (tcgen05 cta_group::2), conv_tma.cu: 0 = never, 1 = N = 256 tiles, 2 = also large N = 128 layers, 3 = and N = 64
int g_tune_wgrad_pair = 0;   // 1 = weight gradients with >= 256 output channels on CTA pairs (cta_group::2), conv_tma.cu; measured: no gain (off)
int g_tune_wgrad_ring = 1;   // 1 = kh x 1 weight gradients on the row-ring kernel where applicable (conv_tma.cu)
int g_tune_conv_ring = 1;    // 1 = kh x 1 convolutions on the row-ring kernel where applicable (conv_tma.cu)
int g_tune_wgrad_swap = 1;   // 1 = operand-swapped wgrad where applicable (conv_igemm.cu), 0 = always the ordinary operand roles
int g_tune_f32_tf32 = 0;   // FP32-parity path: 1 = forward / dgrad of TMA-able geometries on the kind::tf32 tensor-core kernel
int g_tune_conv_path = 0;  // 0 = auto (TMA im2col kernels where applicable), 1 = always the cp.async gather kernels
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// Which kernel variants the calling thread's most recent convolution entry point enqueued ("conv_tma<128,4,2>",
// "wgrad_tma<1,4>+wgrad_reduce", ...): lets the parity tests assert that a geometry really ran on the variant it is meant to
// cover (dei2i_last_kernel).  Host-side bookkeeping only.
static thread_local char g_trace[256] = "";
void trace_begin() { g_trace[0] = 0; }
void trace_kernel(const char* fmt, ...) {
  size_t n = strlen(g_trace);
  if (n + 2 >= sizeof(g_trace)) return;
  if (n) g_trace[n++] = '+';
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_trace + n, sizeof(g_trace) - n, fmt, ap);
  va_end(ap);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The tensor-map encoders are DRIVER entry points: unlike runtime calls they do not bind the primary context to the calling
// thread, and fail with CUDA_ERROR_INVALID_CONTEXT when they are the first CUDA call a thread makes (e.g. a convolution
// backward as the first node an autograd worker thread runs).  One runtime call per thread binds it.
static void bind_primary_context() {
  static thread_local bool bound = false;
  if (!bound) {
    cudaFree(nullptr);
    bound = true;
  }
}

// resolved once, thread-safely (the autograd worker thread and the main thread both encode descriptors)
static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

static int encode_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_pitch_bytes,
                          uint32_t box_inner, uint32_t box_outer, int swizzle128, CUtensorMapDataType dt);
int encode_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_pitch_bytes,
                        uint32_t box_inner, uint32_t box_outer, int swizzle128) {
  return encode_tmap_2d(out, base, inner, outer, row_pitch_bytes, box_inner, box_outer, swizzle128, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
}
int encode_tmap_2d_f32(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_pitch_bytes,
                       uint32_t box_inner, uint32_t box_outer, int swizzle128) {
  return encode_tmap_2d(out, base, inner, outer, row_pitch_bytes, box_inner, box_outer, swizzle128, CU_TENSOR_MAP_DATA_TYPE_FLOAT32);
}
static int encode_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_pitch_bytes,
                          uint32_t box_inner, uint32_t box_outer, int swizzle128, CUtensorMapDataType dt) {
  bind_primary_context();
  EncodeTiledFn fn = get_encode();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is not available (no CUDA driver / no sm_100 device)");
    return DEI2I_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (row_pitch_bytes & 15) != 0) {
    set_error("TMA operand must be 16-byte aligned (base %p, pitch %llu)", base, (unsigned long long)row_pitch_bytes);
    return DEI2I_ERR_INVALID;
  }
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_pitch_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (inner %llu outer %llu pitch %llu box %ux%u)", (int)r,
              (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)row_pitch_bytes, box_inner, box_outer);
    return DEI2I_ERR_CUDA;
  }
  return 0;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_tmap_im2col(CUtensorMap* out, const void* base, uint64_t C, uint64_t W, uint64_t H, uint64_t N, int lower_w,
                              int upper_w, int lower_h, int upper_h, uint32_t stride, uint32_t channels_per_pixel,
                              uint32_t pixels_per_column, CUtensorMapDataType dt, uint64_t esz);
int encode_tmap_im2col_bf16(CUtensorMap* out, const void* base, uint64_t C, uint64_t W, uint64_t H, uint64_t N, int lower_w,
                            int upper_w, int lower_h, int upper_h, uint32_t stride, uint32_t channels_per_pixel,
                            uint32_t pixels_per_column) {
  return encode_tmap_im2col(out, base, C, W, H, N, lower_w, upper_w, lower_h, upper_h, stride, channels_per_pixel, pixels_per_column,
                            CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2);
}
int encode_tmap_im2col_f32(CUtensorMap* out, const void* base, uint64_t C, uint64_t W, uint64_t H, uint64_t N, int lower_w, int upper_w,
                           int lower_h, int upper_h, uint32_t stride, uint32_t channels_per_pixel, uint32_t pixels_per_column) {
  return encode_tmap_im2col(out, base, C, W, H, N, lower_w, upper_w, lower_h, upper_h, stride, channels_per_pixel, pixels_per_column,
                            CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4);
}
static int encode_tmap_im2col(CUtensorMap* out, const void* base, uint64_t C, uint64_t W, uint64_t H, uint64_t N, int lower_w,
                              int upper_w, int lower_h, int upper_h, uint32_t stride, uint32_t channels_per_pixel,
                              uint32_t pixels_per_column, CUtensorMapDataType dt, uint64_t esz) {
  bind_primary_context();
  static EncodeIm2colFn fn = nullptr;
  static int driver_version = 0;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeIm2colFn)p;
    cudaDriverGetVersion(&driver_version);
  });
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeIm2col is not available (no CUDA driver / no sm_100 device)");
    return DEI2I_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (C * esz) % 16 != 0) {
    set_error("im2col TMA operand must be 16-byte aligned (base %p, C %llu)", base, (unsigned long long)C);
    return DEI2I_ERR_INVALID;
  }
  cuuint64_t dims[4] = {C, W, H, N};
  cuuint64_t strides[3] = {C * esz, W * C * esz, H * W * C * esz};
  int lo[2] = {lower_w, lower_h};
  int up[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, stride, stride, 1};
  CUresult r = fn(out, dt, 4, const_cast<void*>(base), dims, strides, lo, up, channels_per_pixel,
                  pixels_per_column, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeIm2col failed with CUresult %d (C %llu W %llu H %llu N %llu lower %d upper %d stride %u)", (int)r,
              (unsigned long long)C, (unsigned long long)W, (unsigned long long)H, (unsigned long long)N, lower_h, upper_h, stride);
    return DEI2I_ERR_CUDA;
  }
  // Known driver issue (drivers <= 13.1; same workaround as CUTLASS make_im2col_tma_copy_desc): for tensors smaller than
  // 128 KiB a descriptor bit must be cleared or the load faults.
  if (driver_version <= 13010 && C * W * H * N * esz < 131072)
    reinterpret_cast<uint64_t*>(out)[1] &= ~(1llu << 21);
  return 0;
}

}  // namespace dei2i

extern "C" int dei2i_version(void) { return 100; }
extern "C" const char* dei2i_last_error(void) { return dei2i::g_err; }
extern "C" long long dei2i_launch_count(void) { return dei2i::g_launches.load(); }
extern "C" const char* dei2i_last_kernel(void) { return dei2i::g_trace; }
extern "C" int dei2i_set_tuning(const char* key, int value) {
  int* slot = nullptr;
  if (!strcmp(key, "conv_stages")) slot = &dei2i::g_tune_conv_stages;
  else if (!strcmp(key, "conv_bn")) slot = &dei2i::g_tune_conv_bn;
  else if (!strcmp(key, "wgrad_bn")) slot = &dei2i::g_tune_wgrad_bn;
  else if (!strcmp(key, "wgrad_waves")) slot = &dei2i::g_tune_wgrad_waves;
  else if (!strcmp(key, "conv_path")) slot = &dei2i::g_tune_conv_path;
  else if (!strcmp(key, "f32_tf32")) slot = &dei2i::g_tune_f32_tf32;
  else if (!strcmp(key, "tma_ms")) slot = &dei2i::g_tune_tma_ms;
  else if (!strcmp(key, "wgrad_swap")) slot = &dei2i::g_tune_wgrad_swap;
  else if (!strcmp(key, "conv_ring")) slot = &dei2i::g_tune_conv_ring;
  else if (!strcmp(key, "mma_issue")) slot = &dei2i::g_tune_mma_issue;
  else if (!strcmp(key, "tma_pair")) slot = &dei2i::g_tune_tma_pair;
  else if (!strcmp(key, "wgrad_ring")) slot = &dei2i::g_tune_wgrad_ring;
  else if (!strcmp(key, "wgrad_pair")) slot = &dei2i::g_tune_wgrad_pair;
  if (slot == nullptr) return -1;
  int old = *slot;
  *slot = value;
  return old;
}
