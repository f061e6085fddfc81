// Detect-head decode: IDetect.fuseforward inference branch (reference models/yolo.py:145-162;
// exact arithmetic order in SURVEY.md A.5).  Reads the NHWC logits written by the 1x1 head conv
// once and writes pred (B,A,no) fp32 once (+ optionally the raw (B,na,ny,nx,no) tensors).
//   p = sigmoid(t); xy = (p*2 - 0.5 + g) * stride; wh = (p*2)^2 * anchor; obj/cls = p
// Every fp32 step is an explicit round-to-nearest intrinsic so nvcc cannot contract it.
#include "common.cuh"

namespace yb {

__device__ __forceinline__ float sigmoid_rn(float t) {
  return __frcp_rn(__fadd_rn(1.0f, expf(-t)));   // == __fdiv_rn(1, x): both correctly rounded
}

__device__ __forceinline__ float decode_elem(float p, int o, int gx, int gy, float stride, float aw, float ah) {
  if (o >= 4) return p;
  float p2 = __fmul_rn(p, 2.0f);
  if (o < 2) {
    float g = (float)(o == 0 ? gx : gy);
    return __fmul_rn(__fadd_rn(__fsub_rn(p2, 0.5f), g), stride);
  }
  return __fmul_rn(__fmul_rn(p2, p2), o == 2 ? aw : ah);
}

struct DecodeLevel {
  const void* logits;
  int dtype, ny, nx, c_stride;
  float stride;
  float anchor_wh[2 * YB_MAX_ANCHORS];
  int row_off;  // first pred row of this level within an image
  float* raw;   // nullable
  // TTA pass (yolo.py:591-595): xywh /= scale, then y = img_h - y (flip 2) or x = img_w - x (flip 3)
  float scale, img_h, img_w;
  int flip;
};

// Block = one run of kDecCells cells of one (image, grid row); thread = channel c of the cell (c = a*no + o, strided
// by the block size when na*no > 256).  Loads are coalesced over the channels of a cell, stores are na runs of `no`
// contiguous floats per cell; no integer division anywhere (a is found by comparison, na <= 8).  The kDecCells loads
// of a thread are issued before any arithmetic.
constexpr int kDecCells = 8;

template <typename T>
__global__ void __launch_bounds__(256) k_decode(DecodeLevel L, int B, int A, int na, int no, float* __restrict__ pred) {
  const int C = na * no;
  const int y = blockIdx.x % L.ny, b = blockIdx.x / L.ny;
  const int x0 = blockIdx.y * kDecCells;
  const int ncell = min(kDecCells, L.nx - x0);
  const T* in = reinterpret_cast<const T*>(L.logits) + ((size_t)(b * L.ny + y) * L.nx + x0) * L.c_stride;
  const size_t plane = (size_t)L.ny * L.nx;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    int a = 0;
#pragma unroll
    for (int q = 1; q < YB_MAX_ANCHORS; ++q) a += (q < na && c >= q * no) ? 1 : 0;
    const int o = c - a * no;
    const float aw = L.anchor_wh[2 * a], ah = L.anchor_wh[2 * a + 1];
    float t[kDecCells];
#pragma unroll
    for (int i = 0; i < kDecCells; ++i) t[i] = i < ncell ? Elem<T>::load(in + (size_t)i * L.c_stride + c) : 0.0f;
    const size_t cell0 = ((size_t)a * L.ny + y) * L.nx + x0;
    float* prow = pred + ((size_t)b * A + L.row_off + cell0) * no + o;
    float* rrow = L.raw ? L.raw + ((size_t)b * na * plane + cell0) * no + o : nullptr;
#pragma unroll
    for (int i = 0; i < kDecCells; ++i) {
      if (i < ncell) {
        float v = decode_elem(sigmoid_rn(t[i]), o, x0 + i, y, L.stride, aw, ah);
        if (o < 4) {
          if (L.scale != 1.0f) v = __fdiv_rn(v, L.scale);
          if (o == 1 && L.flip == 2) v = __fsub_rn(L.img_h, v);
          if (o == 0 && L.flip == 3) v = __fsub_rn(L.img_w, v);
        }
        prow[(size_t)i * no] = v;
        if (rrow) rrow[(size_t)i * no] = t[i];
      }
    }
  }
}

}  // namespace yb

using namespace yb;

static int decode_impl(const yb_head_level* levels, int nl, int na, int no, float* pred, float* const* raw,
                       const yb_tta_pass* pass, void* stream) {
  YB_CHECK_ARG(levels && pred && nl > 0 && nl <= YB_MAX_LEVELS, "yb_decode: bad levels");
  YB_CHECK_ARG(na > 0 && na <= YB_MAX_ANCHORS && no > 5 - 1, "yb_decode: bad na/no");
  int A = 0;
  const int B = levels[0].logits.N;
  for (int l = 0; l < nl; ++l) {
    const yb_tensor& t = levels[l].logits;
    YB_CHECK_ARG(view_ok(t) && t.N == B && t.C >= na * no, "yb_decode: level %d logits view invalid", l);
    YB_CHECK_ARG(t.dtype == YB_F32 || t.dtype == YB_BF16, "yb_decode: logits must be fp32/bf16");
    A += na * t.H * t.W;
  }
  int off = 0, rows = A;
  if (pass) {
    YB_CHECK_ARG(pass->row_off >= 0 && pass->rows_total >= pass->row_off + A, "yb_decode_tta: rows [%d, %d) exceed %d",
                 pass->row_off, pass->row_off + A, pass->rows_total);
    YB_CHECK_ARG(pass->scale > 0.0f && (pass->flip == 0 || pass->flip == 2 || pass->flip == 3), "yb_decode_tta: bad scale/flip");
    off = pass->row_off;
    rows = pass->rows_total;
  }
  for (int l = 0; l < nl; ++l) {
    const yb_tensor& t = levels[l].logits;
    DecodeLevel L;
    L.logits = t.data; L.dtype = t.dtype; L.ny = t.H; L.nx = t.W; L.c_stride = t.c_stride;
    L.stride = levels[l].stride;
    memcpy(L.anchor_wh, levels[l].anchor_wh, sizeof(L.anchor_wh));
    L.row_off = off;
    L.raw = raw ? raw[l] : nullptr;
    L.scale = pass ? pass->scale : 1.0f;
    L.flip = pass ? pass->flip : 0;
    L.img_h = pass ? pass->img_h : 0.0f;
    L.img_w = pass ? pass->img_w : 0.0f;
    YB_CHECK_ARG(ceil_div(t.W, kDecCells) <= 65535, "yb_decode: level %d is too wide", l);
    dim3 grid((unsigned)(B * t.H), (unsigned)ceil_div(t.W, kDecCells));
    if (t.dtype == YB_F32) k_decode<float><<<grid, 256, 0, (cudaStream_t)stream>>>(L, B, rows, na, no, pred);
    else k_decode<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(L, B, rows, na, no, pred);
    YB_LAUNCH_CHECK();
    off += na * t.H * t.W;
  }
  return YB_OK;
}

extern "C" int yb_decode(const yb_head_level* levels, int nl, int na, int no, float* pred, float* const* raw,
                         void* stream) {
  return decode_impl(levels, nl, na, no, pred, raw, nullptr, stream);
}

extern "C" int yb_decode_tta(const yb_head_level* levels, int nl, int na, int no, float* pred, const yb_tta_pass* pass,
                             void* stream) {
  YB_CHECK_ARG(pass, "yb_decode_tta: pass is NULL");
  return decode_impl(levels, nl, na, no, pred, nullptr, pass, stream);
}
