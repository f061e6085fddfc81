// Test-time augmentation input side: scale_img (reference utils/torch_utils.py:247-257) fused with the flip of
// models/yolo.py:588 -- out = F.pad(F.interpolate(flip(x), (h, w), 'bilinear', align_corners=False), value).
// The bilinear arithmetic follows ATen's CPU kernel step by step (pinned against torch 2.11 by
// tests/test_oracle_tta.py): scale = in/out in fp32, src = fma(scale, dst + 0.5, -0.5) clamped at 0, lambda1 = src - i0,
// lambda0 = 1 - lambda1, value = fma(top, ly0, bot*ly1) with top = fma(a, lx0, b*lx1).  HBM-bound: reads each source
// pixel ~once (L1/L2 serve the 2x2 footprints), writes fp32 NCHW once.
#include "common.cuh"

namespace yb {
namespace {

struct ScaleImgParams {
  const void* src;
  float* dst;
  int B, C, H, W;        // source
  int h, w;              // resized extent inside the destination
  int dh, dw;            // destination extent (>= h, w)
  int flip;              // 0 none, 2 up-down, 3 left-right
  float scale_y, scale_x, pad;
};

__device__ __forceinline__ void src_index(float scale, int d, int in_size, int& i0, int& i1, float& l0, float& l1) {
  float s = __fmaf_rn(scale, __fadd_rn((float)d, 0.5f), -0.5f);
  s = s < 0.0f ? 0.0f : s;
  i0 = min((int)floorf(s), in_size - 1);
  l1 = fminf(fmaxf(__fsub_rn(s, (float)i0), 0.0f), 1.0f);
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l0 = __fsub_rn(1.0f, l1);
}

template <typename T> __device__ __forceinline__ float px(const T* p);
template <> __device__ __forceinline__ float px<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float px<unsigned char>(const unsigned char* p) {
  return __fdiv_rn((float)__ldg(p), 255.0f);   // img /= 255.0 (detect.py:103)
}

template <typename T>
__global__ void __launch_bounds__(256) k_scale_img(ScaleImgParams P) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  if (x >= P.dw) return;
  const bool inside = x < P.w && y < P.h;
  int x0 = 0, x1 = 0, y0 = 0, y1 = 0;
  float lx0 = 0, lx1 = 0, ly0 = 0, ly1 = 0;
  if (inside) {
    src_index(P.scale_x, x, P.W, x0, x1, lx0, lx1);
    src_index(P.scale_y, y, P.H, y0, y1, ly0, ly1);
    if (P.flip == 3) { x0 = P.W - 1 - x0; x1 = P.W - 1 - x1; }
    if (P.flip == 2) { y0 = P.H - 1 - y0; y1 = P.H - 1 - y1; }
  }
  const T* src = static_cast<const T*>(P.src);
  const size_t plane = (size_t)P.H * P.W, dplane = (size_t)P.dh * P.dw;
  for (int bc = blockIdx.z; bc < P.B * P.C; bc += gridDim.z) {
    float v = P.pad;
    if (inside) {
      const T* s = src + bc * plane;
      const float a = px<T>(s + (size_t)y0 * P.W + x0), b = px<T>(s + (size_t)y0 * P.W + x1);
      const float c = px<T>(s + (size_t)y1 * P.W + x0), d = px<T>(s + (size_t)y1 * P.W + x1);
      const float top = __fmaf_rn(a, lx0, __fmul_rn(b, lx1));
      const float bot = __fmaf_rn(c, lx0, __fmul_rn(d, lx1));
      v = __fmaf_rn(top, ly0, __fmul_rn(bot, ly1));
    }
    P.dst[bc * dplane + (size_t)y * P.dw + x] = v;
  }
}

}  // namespace
}  // namespace yb

extern "C" int yb_scale_img(const void* src, int src_dtype, int B, int C, int H, int W, int flip, int h, int w,
                            float* dst, int dst_h, int dst_w, float pad_value, void* stream) {
  using namespace yb;
  YB_CHECK_ARG(src && dst && B > 0 && C > 0 && H > 0 && W > 0, "yb_scale_img: bad source");
  YB_CHECK_ARG(h > 0 && w > 0 && dst_h >= h && dst_w >= w, "yb_scale_img: resized %dx%d does not fit %dx%d", h, w, dst_h, dst_w);
  YB_CHECK_ARG(flip == 0 || flip == 2 || flip == 3, "yb_scale_img: flip must be 0, 2 (up-down) or 3 (left-right)");
  YB_CHECK_ARG(src_dtype == YB_F32 || src_dtype == YB_U8, "yb_scale_img: source must be fp32 or uint8");
  ScaleImgParams P;
  P.src = src; P.dst = dst; P.B = B; P.C = C; P.H = H; P.W = W; P.h = h; P.w = w; P.dh = dst_h; P.dw = dst_w;
  P.flip = flip; P.pad = pad_value;
  P.scale_y = (float)H / (float)h;   // area_pixel_compute_scale: input_size / output_size, fp32
  P.scale_x = (float)W / (float)w;
  const int planes = B * C;
  dim3 grid(ceil_div(dst_w, 256), dst_h, planes < 8 ? planes : 8), block(256);
  if (src_dtype == YB_F32) k_scale_img<float><<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(P);
  else k_scale_img<unsigned char><<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(P);
  YB_LAUNCH_CHECK();
  return YB_OK;
}
