// Fused input staging + first convolution ("stem") for the bf16 path.
//
// The first conv of every yolov7 cfg reads the raw image: 3 channels (or 12 after ReOrg), K = 27 / 108.
// As an implicit GEMM on tcgen05 it is pure overhead (K padded to 144+, nine tiny TMA chunks per tile,
// and a 16-channel NHWC staging buffer written and re-read 9x), yet the layer is bound by writing its
// 640x640x32 bf16 output.  This kernel therefore reads the NCHW uint8 / fp32 image directly
// (detect.py:100-104: x/255), applies ReOrg on the fly (models/common.py:52-53), and does the 3x3 conv
// + bias + SiLU/LeakyReLU (models/common.py:110-111) with warp-level bf16 MMAs (fp32 accumulate):
//   * a CTA owns 128 consecutive output pixels of one output row; the 3-row input patch is staged in
//     shared memory as bf16 [dy][w][CP] (CP = 8 or 16 channels per pixel, zero padded);
//   * for a fixed filter row dy the im2col matrix A(p, dx*CP + c) = patch[dy][s*p + dx][c] is an
//     ordinary strided matrix (row stride s*CP) over the patch, so no im2col copy is ever built;
//   * weights are re-packed once per CTA into W_dy[cout][dx*CP + c] (zeros elsewhere).
// 0.7 % of the network FLOPs: the legacy mma.sync path is the right tool here; everything else is tcgen05.
#include <mma.h>

#include "common.cuh"

namespace yb {
using namespace nvcuda;

struct StemParams {
  const void* img;       // NCHW planar, uint8 (scaled by 1/255) or fp32
  int img_u8;
  int B, C, H, W;        // raw image dims (before ReOrg)
  int reorg;
  int Hi, Wi, Cin;       // conv input dims (after ReOrg): Cin = C or 4C
  int Ho, Wo, Cout, stride;
  const __nv_bfloat16* w;  // [Cout][9][cin_pad]
  int cin_pad;
  const float* bias;
  int act; float slope;
  __nv_bfloat16* y; int y_cs;   // NHWC output view
  int CP, KD;            // channels per pixel in the patch (8/16), K per filter row (32/48)
  int tiles_w;           // ceil(Wo / 128)
};

__device__ __forceinline__ float stem_act(float v, int act, float slope) {
  if (act == YB_ACT_SILU) {
    const float h = 0.5f * v;
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
  }
  if (act == YB_ACT_LEAKY) return v > 0.f ? v : v * slope;
  return v;
}

constexpr int kStemWarps = 4;
constexpr int kStemTile = 16 * kStemWarps;  // output pixels per CTA tile; small CTAs, many per SM (latency hiding)

// NT = Cout/16, CP = channels per patch pixel (8|16), S = stride.  FAST3: plain 3-channel image with W % 4 == 0
// (4 pixels per 32-bit / 128-bit load, uint8 -> bf16 through an exact 256-entry table of bf16(v/255)).
template <int NT, int CP, int S, bool FAST3>
__global__ void __launch_bounds__(32 * kStemWarps) k_stem(const StemParams P) {
  constexpr int COUT = NT * 16, KD = (3 * CP + 15) / 16 * 16;
  constexpr int PATCH_W = kStemTile * S + 8;               // pixels per patch row (+ slack for the K overrun)
  constexpr int CPITCH = COUT + 4;
  constexpr int NEED_W = kStemTile * S + 2;
  constexpr bool BREG = (NT * (KD / 16) * 3 <= 12);        // keep all weight fragments in registers
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __nv_bfloat16* wsm = reinterpret_cast<__nv_bfloat16*>(smem);                    // [3][COUT][KD]
  __nv_bfloat16* patch = wsm + 3 * COUT * KD;                                     // [3][PATCH_W][CP]
  float* cst = reinterpret_cast<float*>(patch + 3 * PATCH_W * CP);                // [warps][16][CPITCH]
  float* bias_s = cst + kStemWarps * 16 * CPITCH;                                 // [COUT]
  __nv_bfloat16* lut = reinterpret_cast<__nv_bfloat16*>(bias_s + COUT);           // [256] bf16(v / 255)
  float* mine = cst + warp * 16 * CPITCH;
  for (int i = threadIdx.x; i < 3 * COUT * KD; i += blockDim.x) {
    const int k = i % KD, co = (i / KD) % COUT, dy = i / (KD * COUT);
    const int dx = k / CP, c = k - dx * CP;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (dx < 3 && c < P.Cin) v = P.w[((size_t)co * 9 + dy * 3 + dx) * P.cin_pad + c];
    wsm[i] = v;
  }
  for (int i = threadIdx.x; i < 3 * PATCH_W * CP; i += blockDim.x) patch[i] = __float2bfloat16_rn(0.f);
  for (int i = threadIdx.x; i < COUT; i += blockDim.x) bias_s[i] = P.bias[i];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = __float2bfloat16_rn(__fdiv_rn((float)i, 255.0f));
  __syncthreads();
  wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> breg[BREG ? 3 * (KD / 16) * NT : 1];
  if (BREG) {
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
      for (int kk = 0; kk < KD / 16; ++kk)
#pragma unroll
        for (int j = 0; j < NT; ++j)
          wmma::load_matrix_sync(breg[BREG ? (dy * (KD / 16) + kk) * NT + j : 0], wsm + (dy * COUT + j * 16) * KD + kk * 16, KD);
  }
  const long long total = (long long)P.B * P.Ho * P.tiles_w;
  for (long long t = blockIdx.x; t < total; t += gridDim.x) {
    const int tw = (int)(t % P.tiles_w);
    const long long r = t / P.tiles_w;
    const int ho = (int)(r % P.Ho), n = (int)(r / P.Ho);
    const int wo0 = tw * kStemTile;
    const int wi0 = wo0 * S - 1, hi0 = ho * S - 1;
    if (FAST3) {
      // warp dy (< 3) stages image row hi0 + dy: lane l owns the aligned 4-pixel group l of the window
      if (warp < 3) {
        const int hi = hi0 + warp;
        constexpr int NGRP = (NEED_W + 3) / 4 + 1;
        for (int gl = lane; gl < NGRP; gl += 32) {
          const int grp = (wi0 < 0 ? -1 : wi0 / 4) + gl;     // floor(wi0 / 4) + gl   (wi0 >= -1)
          const bool ok = hi >= 0 && hi < P.Hi && grp >= 0 && grp * 4 < P.Wi;
          float v[3][4];
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const size_t off = (((size_t)n * 3 + c) * P.H + (ok ? hi : 0)) * P.W + (ok ? grp * 4 : 0);
            if (P.img_u8) {
              const uint32_t wv = ok ? __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(P.img) + off)) : 0u;
#pragma unroll
              for (int e = 0; e < 4; ++e) v[c][e] = ok ? __bfloat162float(lut[(wv >> (8 * e)) & 0xff]) : 0.f;
            } else {
              const float4 fv = ok ? __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(P.img) + off))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
              v[c][0] = fv.x; v[c][1] = fv.y; v[c][2] = fv.z; v[c][3] = fv.w;
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int pw = grp * 4 + e - wi0;               // position inside the patch row
            if (pw >= 0 && pw < NEED_W) {
              __nv_bfloat162 lo = __floats2bfloat162_rn(v[0][e], v[1][e]);
              __nv_bfloat162 hi2 = __floats2bfloat162_rn(v[2][e], 0.f);
              uint2 o;
              o.x = *reinterpret_cast<uint32_t*>(&lo);
              o.y = *reinterpret_cast<uint32_t*>(&hi2);
              *reinterpret_cast<uint2*>(patch + ((size_t)warp * PATCH_W + pw) * CP) = o;
            }
          }
        }
      }
    } else {
      const int rows = 3 * P.Cin;
      for (int rr = warp; rr < rows; rr += kStemWarps) {
        const int dy = rr / P.Cin, c = rr - dy * P.Cin;
        const int hi = hi0 + dy;
        int ch = c, radd = 0, cadd = 0, mul = 1;
        if (P.reorg) {  // channel block order (even,even),(odd,even),(even,odd),(odd,odd) in (row, col)
          const int blk = c / P.C;
          ch = c - blk * P.C; radd = blk & 1; cadd = blk >> 1; mul = 2;
        }
        const bool hok = hi >= 0 && hi < P.Hi;
        const size_t rowoff = (((size_t)n * P.C + ch) * P.H + (hok ? mul * hi + radd : 0)) * P.W;
        __nv_bfloat16* prow = patch + (size_t)dy * PATCH_W * CP + c;
        for (int pw = lane; pw < NEED_W; pw += 32) {
          const int wi = wi0 + pw;
          __nv_bfloat16 v = __float2bfloat16_rn(0.f);
          if (hok && wi >= 0 && wi < P.Wi) {
            const size_t off = rowoff + (size_t)(mul * wi + cadd);
            v = P.img_u8 ? lut[__ldg(reinterpret_cast<const uint8_t*>(P.img) + off)]
                         : __float2bfloat16_rn(__ldg(reinterpret_cast<const float*>(P.img) + off));
          }
          prow[pw * CP] = v;
        }
      }
    }
    __syncthreads();
    // ---- each warp: 16 output pixels x COUT
    const int p0 = warp * 16;
    if (wo0 + p0 < P.Wo) {
      wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[NT];
#pragma unroll
      for (int j = 0; j < NT; ++j) wmma::fill_fragment(acc[j], 0.f);
#pragma unroll
      for (int dy = 0; dy < 3; ++dy) {
        const __nv_bfloat16* abase = patch + (dy * PATCH_W + p0 * S) * CP;
#pragma unroll
        for (int kk = 0; kk < KD / 16; ++kk) {
          wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> a;
          wmma::load_matrix_sync(a, abase + kk * 16, S * CP);
#pragma unroll
          for (int j = 0; j < NT; ++j) {
            if (BREG) {
              wmma::mma_sync(acc[j], a, breg[BREG ? (dy * (KD / 16) + kk) * NT + j : 0], acc[j]);
            } else {
              wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> b;
              wmma::load_matrix_sync(b, wsm + (dy * COUT + j * 16) * KD + kk * 16, KD);
              wmma::mma_sync(acc[j], a, b, acc[j]);
            }
          }
        }
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) wmma::store_matrix_sync(mine + j * 16, acc[j], CPITCH, wmma::mem_row_major);
      __syncwarp();
      // ---- bias + activation + bf16 pack: 8 channels (16 bytes) per lane-iteration, pixels are contiguous in y
      constexpr int VPP = COUT / 8;
      __nv_bfloat16* yrow = P.y + (((size_t)n * P.Ho + ho) * P.Wo + wo0 + p0) * P.y_cs;
      const int act = P.act;
      const float slope = P.slope;
#pragma unroll
      for (int i0 = 0; i0 < 16 * VPP; i0 += 32) {
        const int i = i0 + lane;
        const int p = i / VPP, q = i - p * VPP;
        if (i < 16 * VPP && wo0 + p0 + p < P.Wo) {
          const float4 a0 = *reinterpret_cast<const float4*>(mine + p * CPITCH + 8 * q);
          const float4 a1 = *reinterpret_cast<const float4*>(mine + p * CPITCH + 8 * q + 4);
          const float4 b0 = *reinterpret_cast<const float4*>(bias_s + 8 * q);
          const float4 b1 = *reinterpret_cast<const float4*>(bias_s + 8 * q + 4);
          uint4 o;
          __nv_bfloat162 h;
          h = __floats2bfloat162_rn(stem_act(a0.x + b0.x, act, slope), stem_act(a0.y + b0.y, act, slope));
          o.x = *reinterpret_cast<uint32_t*>(&h);
          h = __floats2bfloat162_rn(stem_act(a0.z + b0.z, act, slope), stem_act(a0.w + b0.w, act, slope));
          o.y = *reinterpret_cast<uint32_t*>(&h);
          h = __floats2bfloat162_rn(stem_act(a1.x + b1.x, act, slope), stem_act(a1.y + b1.y, act, slope));
          o.z = *reinterpret_cast<uint32_t*>(&h);
          h = __floats2bfloat162_rn(stem_act(a1.z + b1.z, act, slope), stem_act(a1.w + b1.w, act, slope));
          o.w = *reinterpret_cast<uint32_t*>(&h);
          *reinterpret_cast<uint4*>(yrow + (size_t)p * P.y_cs + 8 * q) = o;
        }
      }
    }
    __syncthreads();
  }
}

template <int NT, int CP, int S, bool FAST3>
static int launch_stem(const StemParams& P, cudaStream_t st) {
  constexpr int COUT = NT * 16, KD = (3 * CP + 15) / 16 * 16, PATCH_W = kStemTile * S + 8;
  const size_t smem = (size_t)3 * COUT * KD * 2 + (size_t)3 * PATCH_W * CP * 2 + (size_t)kStemWarps * 16 * (COUT + 4) * 4 +
                      (size_t)COUT * 4 + 512;
  YB_CUDA(cudaFuncSetAttribute(k_stem<NT, CP, S, FAST3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long total = (long long)P.B * P.Ho * P.tiles_w;
  int per_sm = (int)((200 * 1024) / (smem + 1024));
  if (per_sm > 8) per_sm = 8;
  if (per_sm < 1) per_sm = 1;
  long long grid = (long long)num_sms() * per_sm;
  if (grid > total) grid = total;
  k_stem<NT, CP, S, FAST3><<<(int)grid, 32 * kStemWarps, smem, st>>>(P);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

}  // namespace yb

using namespace yb;

extern "C" int yb_stem_conv(const void* img, int img_dtype, int B, int C, int H, int W, int reorg,
                            const void* weight, const float* bias, int cin_pad, int stride, int act,
                            float leaky_slope, const yb_tensor* y, void* stream) {
  YB_CHECK_ARG(img && weight && bias && y && view_ok(*y), "yb_stem_conv: bad arguments");
  YB_CHECK_ARG(img_dtype == YB_U8 || img_dtype == YB_F32, "yb_stem_conv: image must be uint8 or float32");
  YB_CHECK_ARG(y->dtype == YB_BF16, "yb_stem_conv: bf16 output only (fp32 mode uses yb_stage_input + yb_conv2d)");
  YB_CHECK_ARG(stride == 1 || stride == 2, "yb_stem_conv: stride must be 1 or 2");
  StemParams P;
  P.img = img; P.img_u8 = img_dtype == YB_U8; P.B = B; P.C = C; P.H = H; P.W = W; P.reorg = reorg ? 1 : 0;
  P.Hi = reorg ? H / 2 : H; P.Wi = reorg ? W / 2 : W; P.Cin = reorg ? 4 * C : C;
  P.stride = stride;
  P.Ho = (P.Hi + 2 - 3) / stride + 1; P.Wo = (P.Wi + 2 - 3) / stride + 1;
  P.Cout = y->C;
  YB_CHECK_ARG(P.Cin <= 16 && P.Cin <= cin_pad, "yb_stem_conv: at most 16 input channels");
  YB_CHECK_ARG(P.Cout % 16 == 0 && P.Cout <= 96, "yb_stem_conv: Cout must be a multiple of 16, <= 96");
  YB_CHECK_ARG(y->N == B && y->H == P.Ho && y->W == P.Wo && y->c_stride % 8 == 0 && (uintptr_t)y->data % 16 == 0,
               "yb_stem_conv: output dims / alignment mismatch");
  YB_CHECK_ARG(!reorg || (H % 2 == 0 && W % 2 == 0), "yb_stem_conv: ReOrg needs even H, W");
  P.w = (const __nv_bfloat16*)weight; P.cin_pad = cin_pad; P.bias = bias; P.act = act; P.slope = leaky_slope;
  P.y = (__nv_bfloat16*)y->data; P.y_cs = y->c_stride;
  P.CP = P.Cin <= 8 ? 8 : 16;
  P.KD = (3 * P.CP + 15) / 16 * 16;
  P.tiles_w = (P.Wo + kStemTile - 1) / kStemTile;
  cudaStream_t st = (cudaStream_t)stream;
  if (!reorg && C == 3) {   // plain RGB first layer: tcgen05 kernel (stem_tc.cu)
    const int rc = stem_tc_launch(img, P.img_u8, B, H, W, weight, bias, cin_pad, stride, act, leaky_slope, y, st);
    if (rc != YB_ERR_UNSUPPORTED) return rc;
  }
  static const bool no_tc = getenv("YB_STEM_MMA_SYNC") != nullptr;   // tuning / comparison aid: force the mma.sync stem
  if (reorg && C == 3 && !no_tc) {   // ReOrg first layer of the P6 cfgs: tcgen05 kernel (stemr_tc.cu)
    const int rc = stemr_tc_launch(img, P.img_u8, B, H, W, weight, bias, cin_pad, stride, act, leaky_slope, y, st);
    if (rc != YB_ERR_UNSUPPORTED) return rc;
  }
  const int nt = P.Cout / 16;
  if (P.CP == 8 && stride == 1 && nt == 2) return launch_stem<2, 8, 1, false>(P, st);
  if (P.CP == 8 && stride == 2 && nt == 2) return launch_stem<2, 8, 2, false>(P, st);
  if (P.CP == 16 && stride == 1 && nt == 4) return launch_stem<4, 16, 1, false>(P, st);
  if (P.CP == 16 && stride == 1 && nt == 5) return launch_stem<5, 16, 1, false>(P, st);
  set_error("yb_stem_conv: no specialisation for Cin=%d Cout=%d stride=%d (supported: 3->32 s1|s2, 12->64, 12->80)", P.Cin,
            P.Cout, stride);
  return YB_ERR_UNSUPPORTED;
}
