// HBM-bound data-movement / stencil kernels of the YOLOv7 trunk, NHWC with channel-slice views.
//   MP  MaxPool2d(2,2)                reference models/common.py:30-36
//   SP  MaxPool2d(k,1,k/2)            reference models/common.py:39-45
//   SPPCSPC pools k=5,9,13 fused      reference models/common.py:271,278   (13 = 5o5o5, 9 = 5o5: exact for max)
//   nn.Upsample(None,2,'nearest')     reference cfg/training/yolov7.yaml:78
//   Concat fallback copy / Shortcut   reference models/common.py:61-62, 85-86
//   input staging (+ReOrg)            reference detect.py:100-104, models/common.py:52-53
// All accesses are 128-bit vectors over the channel dimension (8 bf16 / 4 fp32 per thread).
#include "common.cuh"

namespace yb {

// ---- 16-byte vector helpers ---------------------------------------------------------------
template <typename T> struct Vec16;
template <> struct Vec16<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ uint4 vmax(uint4 a, uint4 b) {
    float4 x = *reinterpret_cast<float4*>(&a), y = *reinterpret_cast<float4*>(&b);
    float4 r = make_float4(fmaxf(x.x, y.x), fmaxf(x.y, y.y), fmaxf(x.z, y.z), fmaxf(x.w, y.w));
    return *reinterpret_cast<uint4*>(&r);
  }
  static __device__ __forceinline__ uint4 vadd(uint4 a, uint4 b) {
    float4 x = *reinterpret_cast<float4*>(&a), y = *reinterpret_cast<float4*>(&b);
    float4 r = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    return *reinterpret_cast<uint4*>(&r);
  }
  static __device__ __forceinline__ uint4 neg_inf() {
    float4 r = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    return *reinterpret_cast<uint4*>(&r);
  }
};
template <> struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ uint4 vmax(uint4 a, uint4 b) {
    uint4 r;
    const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* y = reinterpret_cast<const __nv_bfloat162*>(&b);
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = __hmax2(x[i], y[i]);
    return r;
  }
  static __device__ __forceinline__ uint4 vadd(uint4 a, uint4 b) {
    // fp32 add then one rounding to bf16 (what a bf16 torch add does)
    uint4 r;
    const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* y = reinterpret_cast<const __nv_bfloat162*>(&b);
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 fx = __bfloat1622float2(x[i]), fy = __bfloat1622float2(y[i]);
      o[i] = __floats2bfloat162_rn(fx.x + fy.x, fx.y + fy.y);
    }
    return r;
  }
  static __device__ __forceinline__ uint4 neg_inf() {
    const uint32_t v = 0xFF80FF80u;  // bf16 -inf pair
    return make_uint4(v, v, v, v);
  }
};

template <typename T>
__device__ __forceinline__ const uint4* vptr(const yb_tensor& t, int n, int h, int w, int cv) {
  return reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(t.data) +
                                        (((size_t)n * t.H + h) * t.W + w) * t.c_stride) + cv;
}
template <typename T>
__device__ __forceinline__ uint4* vptr_w(const yb_tensor& t, int n, int h, int w, int cv) {
  return reinterpret_cast<uint4*>(reinterpret_cast<T*>(t.data) + (((size_t)n * t.H + h) * t.W + w) * t.c_stride) + cv;
}

static bool vec_ok(const yb_tensor& t) {
  const int n = t.dtype == YB_F32 ? 4 : 8;
  return (t.C % n == 0) && (t.c_stride % n == 0) && ((uintptr_t)t.data % 16 == 0);
}

static inline int grid_for(size_t total, int block) {
  size_t g = (total + block - 1) / block;
  size_t cap = (size_t)num_sms() * 32;
  return (int)(g < cap ? (g ? g : 1) : cap);
}

// ---- MP 2x2 s2 -----------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_maxpool2(yb_tensor x, yb_tensor y) {
  using V = Vec16<T>;
  const int cvn = y.C / V::N;
  const size_t total = (size_t)y.N * y.H * y.W * cvn;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % cvn);
    size_t p = i / cvn;
    int w = (int)(p % y.W); p /= y.W;
    int h = (int)(p % y.H);
    int n = (int)(p / y.H);
    uint4 a = __ldg(vptr<T>(x, n, 2 * h, 2 * w, cv));
    uint4 b = __ldg(vptr<T>(x, n, 2 * h, 2 * w + 1, cv));
    uint4 c = __ldg(vptr<T>(x, n, 2 * h + 1, 2 * w, cv));
    uint4 d = __ldg(vptr<T>(x, n, 2 * h + 1, 2 * w + 1, cv));
    *vptr_w<T>(y, n, h, w, cv) = V::vmax(V::vmax(a, b), V::vmax(c, d));
  }
}

// ---- generic stride-1 k x k max (SP) ---------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_pool_s1(yb_tensor x, yb_tensor y, int k) {
  using V = Vec16<T>;
  const int cvn = y.C / V::N, r = k / 2;
  const size_t total = (size_t)y.N * y.H * y.W * cvn;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % cvn);
    size_t p = i / cvn;
    int w = (int)(p % y.W); p /= y.W;
    int h = (int)(p % y.H);
    int n = (int)(p / y.H);
    uint4 m = V::neg_inf();
    for (int dy = -r; dy <= r; ++dy) {
      int hh = h + dy;
      if (hh < 0 || hh >= x.H) continue;
      for (int dx = -r; dx <= r; ++dx) {
        int ww = w + dx;
        if (ww < 0 || ww >= x.W) continue;
        m = V::vmax(m, __ldg(vptr<T>(x, n, hh, ww, cv)));
      }
    }
    *vptr_w<T>(y, n, h, w, cv) = m;
  }
}

// ---- fused SPP: one CTA = one image x (VPP x 16)-byte channel group; plane staged in shared memory.
// pass(src)->dst is a separable 5-tap max (rows then columns); m5 = pass(x), m9 = pass(m5), m13 = pass(m9).
// VPP = 2 (32-byte sectors) keeps a CTA's plane small (many CTAs per SM, many CTAs even at batch 1: the kernel is
// seven dependent phases, i.e. latency bound per CTA); pixel -> (h, w) by multiplication (magic_w = 2^32 / W + 1).
template <typename T, int VPP>
__global__ void __launch_bounds__(256) k_spp3(yb_tensor x, yb_tensor y5, yb_tensor y9, yb_tensor y13, unsigned magic_w) {
  using V = Vec16<T>;
  extern __shared__ __align__(16) uint4 smem[];
  const int H = x.H, W = x.W, HW = H * W;
  uint4* A = smem;
  uint4* Bf = smem + (size_t)HW * VPP;
  const int groups = ceil_div(x.C / V::N, VPP);
  const int n = blockIdx.x / groups, g = blockIdx.x % groups;
  const int cv0 = g * VPP;
  const int nv = min(VPP, x.C / V::N - cv0);
  const int items = HW * VPP;
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int v = i % VPP, p = i / VPP;
    const int h = magic_w ? (int)__umulhi((unsigned)p, magic_w) : p, w = p - h * W;
    if (v < nv) A[i] = __ldg(vptr<T>(x, n, h, w, cv0 + v));
  }
  __syncthreads();
  yb_tensor outs[3] = {y5, y9, y13};
#pragma unroll 1
  for (int pass = 0; pass < 3; ++pass) {
    for (int i = threadIdx.x; i < items; i += blockDim.x) {  // horizontal: A -> Bf
      const int v = i % VPP, p = i / VPP;
      const int h = magic_w ? (int)__umulhi((unsigned)p, magic_w) : p, w = p - h * W;
      if (v >= nv) continue;
      uint4 m = A[i];
#pragma unroll
      for (int d = 1; d <= 2; ++d) {
        if (w - d >= 0) m = V::vmax(m, A[i - d * VPP]);
        if (w + d < W) m = V::vmax(m, A[i + d * VPP]);
      }
      Bf[i] = m;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < items; i += blockDim.x) {  // vertical: Bf -> A (+ global)
      const int v = i % VPP, p = i / VPP;
      const int h = magic_w ? (int)__umulhi((unsigned)p, magic_w) : p, w = p - h * W;
      if (v >= nv) continue;
      uint4 m = Bf[i];
#pragma unroll
      for (int d = 1; d <= 2; ++d) {
        if (h - d >= 0) m = V::vmax(m, Bf[i - d * W * VPP]);
        if (h + d < H) m = V::vmax(m, Bf[i + d * W * VPP]);
      }
      A[i] = m;
      *vptr_w<T>(outs[pass], n, h, w, cv0 + v) = m;
    }
    __syncthreads();
  }
}

// ---- 2x nearest upsample / copy / add --------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_upsample2(yb_tensor x, yb_tensor y) {
  using V = Vec16<T>;
  const int cvn = y.C / V::N;
  const size_t total = (size_t)y.N * y.H * y.W * cvn;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % cvn);
    size_t p = i / cvn;
    int w = (int)(p % y.W); p /= y.W;
    int h = (int)(p % y.H);
    int n = (int)(p / y.H);
    *vptr_w<T>(y, n, h, w, cv) = __ldg(vptr<T>(x, n, h >> 1, w >> 1, cv));
  }
}

template <typename T, bool ADD>
__global__ void __launch_bounds__(256) k_copy_add(yb_tensor a, yb_tensor b, yb_tensor y) {
  using V = Vec16<T>;
  const int cvn = y.C / V::N;
  const size_t total = (size_t)y.N * y.H * y.W * cvn;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % cvn);
    size_t p = i / cvn;
    int w = (int)(p % y.W); p /= y.W;
    int h = (int)(p % y.H);
    int n = (int)(p / y.H);
    uint4 v = __ldg(vptr<T>(a, n, h, w, cv));
    if (ADD) v = V::vadd(v, __ldg(vptr<T>(b, n, h, w, cv)));
    *vptr_w<T>(y, n, h, w, cv) = v;
  }
}

// ---- input staging: NCHW (u8 | f32) -> NHWC (bf16 | f32), zero channel padding, optional ReOrg
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) k_stage(const TI* __restrict__ img, int B, int C, int H, int W, int reorg,
                                               yb_tensor y) {
  const size_t total = (size_t)y.N * y.H * y.W;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int w = (int)(i % y.W);
    size_t p = i / y.W;
    int h = (int)(p % y.H);
    int n = (int)(p / y.H);
    TO* o = reinterpret_cast<TO*>(y.data) + i * y.c_stride;
    const int live = reorg ? 4 * C : C;
    for (int c = 0; c < y.C; ++c) {
      float v = 0.f;
      if (c < live) {
        int ch = c, hh = h, ww = w;
        if (reorg) {
          int blk = c / C;
          ch = c - blk * C;
          hh = 2 * h + (blk & 1);
          ww = 2 * w + (blk >> 1);
        }
        TI raw = img[(((size_t)n * C + ch) * H + hh) * W + ww];
        if (sizeof(TI) == 1) v = __fdiv_rn((float)raw, 255.0f);  // img /= 255.0 (detect.py:102)
        else v = (float)raw;
      }
      Elem<TO>::store(o + c, v);
    }
  }
}

// ---- host wrappers -----------------------------------------------------------------------------
#define YB_DISPATCH_T(dt, ...)                                  \
  if ((dt) == YB_F32) { using T = float; __VA_ARGS__; }         \
  else { using T = __nv_bfloat16; __VA_ARGS__; }

static int check_pair(const yb_tensor* x, const yb_tensor* y, const char* what) {
  YB_CHECK_ARG(x && y && view_ok(*x) && view_ok(*y), "%s: bad tensor view", what);
  YB_CHECK_ARG(x->dtype == y->dtype && (x->dtype == YB_F32 || x->dtype == YB_BF16), "%s: dtype mismatch", what);
  YB_CHECK_ARG(x->C == y->C && x->N == y->N, "%s: N/C mismatch", what);
  YB_CHECK_ARG(vec_ok(*x) && vec_ok(*y), "%s: channels/pointers must be 16-byte aligned", what);
  return YB_OK;
}

}  // namespace yb

using namespace yb;

extern "C" int yb_maxpool2x2(const yb_tensor* x, const yb_tensor* y, void* stream) {
  int rc = check_pair(x, y, "yb_maxpool2x2");
  if (rc) return rc;
  YB_CHECK_ARG(y->H == x->H / 2 && y->W == x->W / 2, "yb_maxpool2x2: output must be floor(H/2) x floor(W/2)");
  size_t total = (size_t)y->N * y->H * y->W * (y->C / (y->dtype == YB_F32 ? 4 : 8));
  YB_DISPATCH_T(x->dtype, (k_maxpool2<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*x, *y)));
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_maxpool_s1(const yb_tensor* x, const yb_tensor* y, int k, void* stream) {
  int rc = check_pair(x, y, "yb_maxpool_s1");
  if (rc) return rc;
  YB_CHECK_ARG(k % 2 == 1 && k >= 1 && k <= 31, "yb_maxpool_s1: k must be odd");
  YB_CHECK_ARG(y->H == x->H && y->W == x->W, "yb_maxpool_s1: shape mismatch");
  size_t total = (size_t)y->N * y->H * y->W * (y->C / (y->dtype == YB_F32 ? 4 : 8));
  YB_DISPATCH_T(x->dtype, (k_pool_s1<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*x, *y, k)));
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_spp3(const yb_tensor* x, const yb_tensor* y5, const yb_tensor* y9, const yb_tensor* y13,
                       void* stream) {
  int rc = check_pair(x, y5, "yb_spp3");
  if (!rc) rc = check_pair(x, y9, "yb_spp3");
  if (!rc) rc = check_pair(x, y13, "yb_spp3");
  if (rc) return rc;
  YB_CHECK_ARG(y5->H == x->H && y5->W == x->W && y9->H == x->H && y9->W == x->W && y13->H == x->H && y13->W == x->W,
               "yb_spp3: shape mismatch");
  constexpr int kSppVpp = 2;
  const size_t smem = (size_t)x->H * x->W * kSppVpp * 16 * 2;
  if (smem > 200 * 1024 || (long long)x->H * x->W * x->W >= 0x100000000LL) {  // plane does not fit: three generic passes
    if ((rc = yb_maxpool_s1(x, y5, 5, stream))) return rc;
    if ((rc = yb_maxpool_s1(x, y9, 9, stream))) return rc;
    return yb_maxpool_s1(x, y13, 13, stream);
  }
  const int vn = x->dtype == YB_F32 ? 4 : 8;
  const int groups = ceil_div(x->C / vn, kSppVpp);
  const unsigned magic_w = x->W > 1 ? (unsigned)(0x100000000ULL / (unsigned long long)x->W) + 1u : 0u;
  YB_DISPATCH_T(x->dtype, {
    if (smem > 48 * 1024)
      YB_CUDA(cudaFuncSetAttribute(k_spp3<T, kSppVpp>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_spp3<T, kSppVpp><<<x->N * groups, 256, smem, (cudaStream_t)stream>>>(*x, *y5, *y9, *y13, magic_w);
  });
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_upsample2x(const yb_tensor* x, const yb_tensor* y, void* stream) {
  int rc = check_pair(x, y, "yb_upsample2x");
  if (rc) return rc;
  YB_CHECK_ARG(y->H == 2 * x->H && y->W == 2 * x->W, "yb_upsample2x: output must be 2H x 2W");
  size_t total = (size_t)y->N * y->H * y->W * (y->C / (y->dtype == YB_F32 ? 4 : 8));
  YB_DISPATCH_T(x->dtype, (k_upsample2<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*x, *y)));
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_copy(const yb_tensor* x, const yb_tensor* y, void* stream) {
  int rc = check_pair(x, y, "yb_copy");
  if (rc) return rc;
  YB_CHECK_ARG(y->H == x->H && y->W == x->W, "yb_copy: shape mismatch");
  size_t total = (size_t)y->N * y->H * y->W * (y->C / (y->dtype == YB_F32 ? 4 : 8));
  YB_DISPATCH_T(x->dtype,
                (k_copy_add<T, false><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*x, *x, *y)));
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_add(const yb_tensor* a, const yb_tensor* b, const yb_tensor* y, void* stream) {
  int rc = check_pair(a, y, "yb_add");
  if (!rc) rc = check_pair(b, y, "yb_add");
  if (rc) return rc;
  YB_CHECK_ARG(y->H == a->H && y->W == a->W && b->H == a->H && b->W == a->W, "yb_add: shape mismatch");
  size_t total = (size_t)y->N * y->H * y->W * (y->C / (y->dtype == YB_F32 ? 4 : 8));
  YB_DISPATCH_T(a->dtype, (k_copy_add<T, true><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*a, *b, *y)));
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_stage_input(const void* img, int img_dtype, int B, int C, int H, int W, int reorg,
                              const yb_tensor* y, void* stream) {
  YB_CHECK_ARG(img && y && view_ok(*y), "yb_stage_input: bad arguments");
  YB_CHECK_ARG(img_dtype == YB_U8 || img_dtype == YB_F32, "yb_stage_input: image must be uint8 or float32");
  YB_CHECK_ARG(y->dtype == YB_F32 || y->dtype == YB_BF16, "yb_stage_input: output must be fp32/bf16");
  if (reorg) {
    YB_CHECK_ARG(H % 2 == 0 && W % 2 == 0 && y->H == H / 2 && y->W == W / 2 && y->C >= 4 * C,
                 "yb_stage_input: ReOrg output must be (B,H/2,W/2,>=4C)");
  } else {
    YB_CHECK_ARG(y->H == H && y->W == W && y->C >= C, "yb_stage_input: output must be (B,H,W,>=C)");
  }
  YB_CHECK_ARG(y->N == B, "yb_stage_input: batch mismatch");
  size_t total = (size_t)y->N * y->H * y->W;
  int grid = grid_for(total, 256);
  cudaStream_t s = (cudaStream_t)stream;
  if (img_dtype == YB_U8) {
    if (y->dtype == YB_F32) k_stage<uint8_t, float><<<grid, 256, 0, s>>>((const uint8_t*)img, B, C, H, W, reorg, *y);
    else k_stage<uint8_t, __nv_bfloat16><<<grid, 256, 0, s>>>((const uint8_t*)img, B, C, H, W, reorg, *y);
  } else {
    if (y->dtype == YB_F32) k_stage<float, float><<<grid, 256, 0, s>>>((const float*)img, B, C, H, W, reorg, *y);
    else k_stage<float, __nv_bfloat16><<<grid, 256, 0, s>>>((const float*)img, B, C, H, W, reorg, *y);
  }
  YB_LAUNCH_CHECK();
  return YB_OK;
}
