// Pre-/post-processing either side of the hot path.
//
// k_letterbox: reference utils/datasets.py:1277-1307 (letterbox) fused with datasets.py:198-199 (BGR->RGB,
// HWC->CHW): one pass from the decoded uint8 HWC frame to the planar uint8 batch the first-layer kernel reads.
// The resampling restates cv2.resize(INTER_LINEAR) on uint8 bit for bit (oracle/preprocess_ref.py, pinned against
// the installed OpenCV): per axis  f = float((d + 0.5) * scale - 0.5), s = floor(f), weights
// round_half_even((1 - f) * 2048), round_half_even(f * 2048); horizontal taps collapse onto the border pixel,
// vertical taps clamp their row; horizontal pass in int32, vertical pass
// (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2.  An exact 2x reduction is a 2x2 box average.
//
// k_scale_coords: reference utils/general.py:340-361 (+ detect.py:183 .round()) on the packed NMS output.
#include "common.cuh"

namespace yb {

namespace {

constexpr int kLbMaxImgs = 32;          // images per launch (descriptors travel as kernel parameters)
constexpr int kLbBX = 32, kLbBY = 8;    // block: 32 threads x 4 pixels wide, 8 rows
constexpr int kLbPx = 4;

enum { LB_COPY = 0, LB_LINEAR = 1, LB_AREA2 = 2 };

struct LbImg {
  const unsigned char* src;
  long long pitch;
  int src_h, src_w, new_w, new_h, top, left;
  double scale_x, scale_y;
  int mode;
};

struct LbParams {
  LbImg img[kLbMaxImgs];
  unsigned char* dst;
  int dst_h, dst_w;
  int swap_rb;
  int vec_ok;
  unsigned char border[4];
};

// Source index and 11-bit weights of one destination index (OpenCV's xofs/ialpha, yofs/ibeta tables).
__device__ __forceinline__ void axis_coeff(int d, double scale, int src_n, bool collapse, int& s, int& w0, int& w1) {
  float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
  const float fl = floorf(f);
  s = (int)fl;
  f = __fsub_rn(f, fl);
  if (collapse) {
    if (s < 0) { f = 0.f; s = 0; }
    if (s >= src_n - 1) { f = 0.f; s = src_n - 1; }
  }
  w0 = __float2int_rn(__fmul_rn(__fsub_rn(1.0f, f), 2048.0f));
  w1 = __float2int_rn(__fmul_rn(f, 2048.0f));
}

template <int LAYOUT>
__global__ void __launch_bounds__(kLbBX * kLbBY) k_letterbox(const __grid_constant__ LbParams P) {
  const LbImg& I = P.img[blockIdx.z];
  const int x0 = (blockIdx.x * kLbBX + threadIdx.x) * kLbPx;
  const int y = blockIdx.y * kLbBY + threadIdx.y;
  if (x0 >= P.dst_w || y >= P.dst_h) return;

  unsigned char px[kLbPx][3];   // [pixel][source channel]
#pragma unroll
  for (int i = 0; i < kLbPx; ++i) { px[i][0] = P.border[0]; px[i][1] = P.border[1]; px[i][2] = P.border[2]; }

  const int ry = y - I.top;
  if (ry >= 0 && ry < I.new_h && x0 + kLbPx > I.left && x0 < I.left + I.new_w) {
    if (I.mode == LB_LINEAR) {
      int sy, b0, b1;
      axis_coeff(ry, I.scale_y, I.src_h, false, sy, b0, b1);
      const int y0 = min(max(sy, 0), I.src_h - 1), y1 = min(max(sy + 1, 0), I.src_h - 1);
      const unsigned char* r0 = I.src + (long long)y0 * I.pitch;
      const unsigned char* r1 = I.src + (long long)y1 * I.pitch;
#pragma unroll
      for (int i = 0; i < kLbPx; ++i) {
        const int rx = x0 + i - I.left;
        if (rx < 0 || rx >= I.new_w) continue;
        int sx, a0, a1;
        axis_coeff(rx, I.scale_x, I.src_w, true, sx, a0, a1);
        const int o0 = sx * 3, o1 = min(sx + 1, I.src_w - 1) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int h0 = (int)__ldg(r0 + o0 + c) * a0 + (int)__ldg(r0 + o1 + c) * a1;
          const int h1 = (int)__ldg(r1 + o0 + c) * a0 + (int)__ldg(r1 + o1 + c) * a1;
          const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
          px[i][c] = (unsigned char)min(max(v, 0), 255);
        }
      }
    } else if (I.mode == LB_AREA2) {
      const unsigned char* r0 = I.src + (long long)(2 * ry) * I.pitch;
      const unsigned char* r1 = r0 + I.pitch;
#pragma unroll
      for (int i = 0; i < kLbPx; ++i) {
        const int rx = x0 + i - I.left;
        if (rx < 0 || rx >= I.new_w) continue;
        const int o = rx * 6;
#pragma unroll
        for (int c = 0; c < 3; ++c)
          px[i][c] = (unsigned char)(((int)__ldg(r0 + o + c) + (int)__ldg(r0 + o + 3 + c) + (int)__ldg(r1 + o + c) +
                                      (int)__ldg(r1 + o + 3 + c) + 2) >> 2);
      }
    } else {
      const unsigned char* r0 = I.src + (long long)ry * I.pitch;
#pragma unroll
      for (int i = 0; i < kLbPx; ++i) {
        const int rx = x0 + i - I.left;
        if (rx < 0 || rx >= I.new_w) continue;
#pragma unroll
        for (int c = 0; c < 3; ++c) px[i][c] = __ldg(r0 + rx * 3 + c);
      }
    }
  }

  const size_t plane = (size_t)P.dst_h * P.dst_w;
  if (LAYOUT == YB_LAYOUT_CHW) {
    unsigned char* d = P.dst + (size_t)blockIdx.z * 3 * plane + (size_t)y * P.dst_w + x0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int cs = P.swap_rb ? 2 - c : c;
      if (P.vec_ok) {
        *reinterpret_cast<uchar4*>(d + c * plane) = make_uchar4(px[0][cs], px[1][cs], px[2][cs], px[3][cs]);
      } else {
#pragma unroll
        for (int i = 0; i < kLbPx; ++i)
          if (x0 + i < P.dst_w) d[c * plane + i] = px[i][cs];
      }
    }
  } else {
    unsigned char* d = P.dst + ((size_t)blockIdx.z * plane + (size_t)y * P.dst_w + x0) * 3;
    unsigned char o[12];
#pragma unroll
    for (int i = 0; i < kLbPx; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c) o[i * 3 + c] = px[i][P.swap_rb ? 2 - c : c];
    if (P.vec_ok) {
#pragma unroll
      for (int q = 0; q < 3; ++q)
        *reinterpret_cast<uchar4*>(d + 4 * q) = make_uchar4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < kLbPx; ++i)
        if (x0 + i < P.dst_w) { d[i * 3] = o[i * 3]; d[i * 3 + 1] = o[i * 3 + 1]; d[i * 3 + 2] = o[i * 3 + 2]; }
    }
  }
}

__global__ void __launch_bounds__(256) k_scale_coords(float* dets, int B, int max_rows, int row_stride,
                                                      const int* counts, const float* table, int do_round) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * max_rows) return;
  const int b = i / max_rows, r = i - b * max_rows;
  if (counts && r >= counts[b]) return;
  const float* t = table + 5 * b;
  const float gain = t[0], px = t[1], py = t[2], w0 = t[3], h0 = t[4];
  float* d = dets + (size_t)i * row_stride;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const bool isx = (k & 1) == 0;
    float v = __fdiv_rn(__fsub_rn(d[k], isx ? px : py), gain);
    if (v == v) v = fminf(fmaxf(v, 0.0f), isx ? w0 : h0);       // clamp_(0, limit); NaN stays NaN like torch.clamp
    d[k] = do_round ? rintf(v) : v;
  }
}

}  // namespace

}  // namespace yb

extern "C" int yb_letterbox(const yb_letterbox_src* srcs, int B, void* dst, int dst_h, int dst_w, int layout,
                            int swap_rb, const uint8_t border[3], void* stream) {
  using namespace yb;
  YB_CHECK_ARG(srcs && dst && B > 0 && dst_h > 0 && dst_w > 0 && border, "yb_letterbox: bad arguments");
  YB_CHECK_ARG(layout == YB_LAYOUT_HWC || layout == YB_LAYOUT_CHW, "yb_letterbox: layout %d", layout);
  const size_t img_bytes = (size_t)dst_h * dst_w * 3;
  for (int base = 0; base < B; base += kLbMaxImgs) {
    const int nb = B - base < kLbMaxImgs ? B - base : kLbMaxImgs;
    LbParams P;
    memset(&P, 0, sizeof(P));
    for (int i = 0; i < nb; ++i) {
      const yb_letterbox_src& s = srcs[base + i];
      YB_CHECK_ARG(s.src && s.src_h > 0 && s.src_w > 0 && s.new_w > 0 && s.new_h > 0 && s.src_pitch >= 3ll * s.src_w,
                   "yb_letterbox: image %d: bad source description", base + i);
      YB_CHECK_ARG(s.top >= 0 && s.left >= 0 && s.top + s.new_h <= dst_h && s.left + s.new_w <= dst_w,
                   "yb_letterbox: image %d: %dx%d at (%d,%d) does not fit %dx%d", base + i, s.new_w, s.new_h, s.left,
                   s.top, dst_w, dst_h);
      LbImg& I = P.img[i];
      I.src = static_cast<const unsigned char*>(s.src);
      I.pitch = s.src_pitch;
      I.src_h = s.src_h; I.src_w = s.src_w; I.new_w = s.new_w; I.new_h = s.new_h; I.top = s.top; I.left = s.left;
      // cv::resize: inv_scale = dsize / ssize, scale = 1 / inv_scale (both double)
      I.scale_x = 1.0 / ((double)s.new_w / (double)s.src_w);
      I.scale_y = 1.0 / ((double)s.new_h / (double)s.src_h);
      I.mode = (s.new_w == s.src_w && s.new_h == s.src_h) ? LB_COPY
               : (s.src_w == 2 * s.new_w && s.src_h == 2 * s.new_h) ? LB_AREA2 : LB_LINEAR;
    }
    P.dst = static_cast<unsigned char*>(dst) + (size_t)base * img_bytes;
    P.dst_h = dst_h; P.dst_w = dst_w; P.swap_rb = swap_rb ? 1 : 0;
    P.vec_ok = (dst_w % 4 == 0) && ((reinterpret_cast<uintptr_t>(P.dst) & 3u) == 0);
    P.border[0] = border[0]; P.border[1] = border[1]; P.border[2] = border[2];
    dim3 grid(ceil_div(dst_w, kLbBX * kLbPx), ceil_div(dst_h, kLbBY), nb), block(kLbBX, kLbBY);
    if (layout == YB_LAYOUT_CHW) k_letterbox<YB_LAYOUT_CHW><<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(P);
    else k_letterbox<YB_LAYOUT_HWC><<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(P);
    YB_LAUNCH_CHECK();
  }
  return YB_OK;
}

extern "C" int yb_scale_coords(float* dets, int B, int max_rows, int row_stride, const int32_t* counts,
                               const float* table, int do_round, void* stream) {
  using namespace yb;
  YB_CHECK_ARG(dets && table && B > 0 && max_rows >= 0 && row_stride >= 4, "yb_scale_coords: bad arguments");
  if (max_rows == 0) return YB_OK;
  const int total = B * max_rows;
  k_scale_coords<<<ceil_div(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dets, B, max_rows, row_stride,
                                                                                    counts, table, do_round);
  YB_LAUNCH_CHECK();
  return YB_OK;
}
