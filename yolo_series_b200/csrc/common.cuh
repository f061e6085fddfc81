// Shared host/device helpers for libyolo_b200.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/yolo_b200.h"

namespace yb {

void set_error(const char* fmt, ...);

#define YB_CHECK_ARG(cond, ...)                     \
  do {                                              \
    if (!(cond)) {                                  \
      ::yb::set_error(__VA_ARGS__);                 \
      return YB_ERR_INVALID;                        \
    }                                               \
  } while (0)

#define YB_CUDA(expr)                                                                        \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      ::yb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return YB_ERR_CUDA;                                                                    \
    }                                                                                        \
  } while (0)

#define YB_LAUNCH_CHECK()                                                                    \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess) {                                                                 \
      ::yb::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return YB_ERR_CUDA;                                                                    \
    }                                                                                        \
  } while (0)

inline int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

inline size_t dtype_size(int dt) { return dt == YB_F32 ? 4 : dt == YB_BF16 ? 2 : 1; }

inline bool view_ok(const yb_tensor& t) {
  return t.data && t.N > 0 && t.H > 0 && t.W > 0 && t.C > 0 && t.c_stride >= t.C;
}

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- dtype traits for the elementwise / stencil kernels
template <typename T> struct Elem;
template <> struct Elem<float> {
  static __device__ __forceinline__ float load(const float* p) { return *p; }
  static __device__ __forceinline__ void store(float* p, float v) { *p = v; }
};
template <> struct Elem<__nv_bfloat16> {
  static __device__ __forceinline__ float load(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

// Launch with the programmatic-stream-serialization attribute (the kernel must call griddepcontrol.wait before
// touching anything an earlier kernel writes).  Captured into CUDA graphs as a programmatic dependency edge.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3((unsigned)block, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  static const bool off = getenv("YB_NO_PDL") != nullptr;   // debugging aid: plain stream-ordered launches
  cfg.attrs = at;
  cfg.numAttrs = off ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Same, as clusters of two CTAs (CTA pairs for tcgen05 cta_group::2); `grid` must be even.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_pairs(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3((unsigned)block, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  static const bool off = getenv("YB_NO_PDL") != nullptr;
  cfg.attrs = at;
  cfg.numAttrs = off ? 1 : 2;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// internal launchers shared by the C ABI wrappers and the engine
int conv_fp32_launch(const yb_conv_desc& d, cudaStream_t s);

struct ConvTcPlan;  // opaque: tensor maps + launch geometry for one bf16 conv
int conv_tc_plan_create(const yb_conv_desc& d, ConvTcPlan** out);
int conv_tc_launch(const ConvTcPlan* p, cudaStream_t s);
void conv_tc_plan_destroy(ConvTcPlan* p);

struct Conv3Plan;   // 3x3 stride-1 slab kernel (conv3_tc.cu); plan_create returns YB_ERR_UNSUPPORTED to decline
int conv3_tc_plan_create(const yb_conv_desc& d, Conv3Plan** out);
int conv3_tc_launch(const Conv3Plan* p, cudaStream_t s);
void conv3_tc_plan_destroy(Conv3Plan* p);

const char* conv3_tc_kernel_name(const Conv3Plan* p);   // "k_conv3_tc" or "k_conv3w_tc" (the wide-slab variant it may hold)
struct Conv3wPlan;   // 3x3 stride-1 wide-slab kernel (conv3w_tc.cu): one activation slab per K chunk; held by Conv3Plan
int conv3w_plan_create(const yb_conv_desc& d, int force, Conv3wPlan** out);
int conv3w_launch(const Conv3wPlan* p, cudaStream_t s);
void conv3w_plan_destroy(Conv3wPlan* p);
struct ConvS2pPlan;  // 3x3 stride-2 pair-slab kernel for 32-channel inputs (conv3s2p_tc.cu)
int conv3s2p_plan_create(const yb_conv_desc& d, ConvS2pPlan** out);
int conv3s2p_launch(const ConvS2pPlan* p, cudaStream_t s);
void conv3s2p_plan_destroy(ConvS2pPlan* p);

// tcgen05 first layer for 3-channel images (stem_tc.cu); YB_ERR_UNSUPPORTED -> caller uses the mma.sync stem
int stem_tc_launch(const void* img, int img_u8, int B, int H, int W, const void* weight, const float* bias,
                   int cin_pad, int stride, int act, float slope, const yb_tensor* y, cudaStream_t st);

// tcgen05 first layer of the P6 cfgs: ReOrg + 12 -> 64 | 80 (stemr_tc.cu); YB_ERR_UNSUPPORTED -> mma.sync stem
int stemr_tc_launch(const void* img, int img_u8, int B, int H, int W, const void* weight, const float* bias,
                    int cin_pad, int stride, int act, float slope, const yb_tensor* y, cudaStream_t st);

}  // namespace yb
