// fp32 "reference mode" convolution: NHWC implicit GEMM on the CUDA cores with fp32 FMA accumulate.
// Same operator contract as the tcgen05 kernel (conv_tc.cu): act(conv2d(x, W', b')) of
// Conv.fuseforward (reference models/common.py:110-111), k in {1,3}, stride in {1,2}, pad k/2,
// optional fused 2x nearest upsample and Shortcut residual.  Used for the north-star's fp32 parity
// mode (<= 1e-2 px / 1e-3 score vs the reference CPU path) and as the on-device checker of the bf16
// tensor-core path.  Tile: 64 pixels x 64 output channels per CTA, 4x4 outputs per thread, K step 16.
#include "common.cuh"

namespace yb {

struct ConvF32Params {
  const float* x; float* y; const float* w; const float* bias; const float* res;
  int N, H, W, Cin, x_cs;           // input view
  int Ho, Wo, Cout, y_cs;           // conv output extent (before upsample), y channel stride
  int yH, yW;                       // y spatial dims (= Ho*ups)
  int res_cs;
  int cin_pad, taps, k, stride, pad, act, ups;
  float slope;
};

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
  if (act == YB_ACT_SILU) return __fdividef(v, 1.0f + __expf(-v));
  if (act == YB_ACT_LEAKY) return v > 0.f ? v : v * slope;
  return v;
}
__device__ __forceinline__ float apply_act_precise(float v, int act, float slope) {
  if (act == YB_ACT_SILU) return v / (1.0f + expf(-v));
  if (act == YB_ACT_LEAKY) return v > 0.f ? v : v * slope;
  return v;
}

__global__ void __launch_bounds__(256) k_conv_fp32(const ConvF32Params P) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long M = (long long)P.N * P.Ho * P.Wo;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;

  // loader role: pixel lp (0..63), cin quad lq (0..3)
  const int lp = tid >> 2, lq = tid & 3;
  long long lm = m0 + lp;
  const bool lvalid = lm < M;
  int ln = 0, lho = 0, lwo = 0;
  if (lvalid) {
    lwo = (int)(lm % P.Wo);
    long long t = lm / P.Wo;
    lho = (int)(t % P.Ho);
    ln = (int)(t / P.Ho);
  }
  const int lcout = n0 + lp;  // weight row loaded by this thread
  const bool wvalid = lcout < P.Cout;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int tap = 0; tap < P.taps; ++tap) {
    const int dy = tap / P.k, dx = tap - dy * P.k;
    const int hi = lho * P.stride + dy - P.pad, wi = lwo * P.stride + dx - P.pad;
    const bool pvalid = lvalid && hi >= 0 && hi < P.H && wi >= 0 && wi < P.W;
    const float* xrow = P.x + (((size_t)ln * P.H + (pvalid ? hi : 0)) * P.W + (pvalid ? wi : 0)) * P.x_cs;
    const float* wrow = P.w + ((size_t)(wvalid ? lcout : 0) * P.taps + tap) * P.cin_pad;
    for (int c0 = 0; c0 < P.Cin; c0 += BK) {
      const int c = c0 + lq * 4;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
      if (pvalid && c < P.Cin) a = __ldg(reinterpret_cast<const float4*>(xrow + c));
      if (wvalid && c < P.Cin) b = __ldg(reinterpret_cast<const float4*>(wrow + c));
      __syncthreads();
      As[lq * 4 + 0][lp] = a.x; As[lq * 4 + 1][lp] = a.y; As[lq * 4 + 2][lp] = a.z; As[lq * 4 + 3][lp] = a.w;
      Bs[lq * 4 + 0][lp] = b.x; Bs[lq * 4 + 1][lp] = b.y; Bs[lq * 4 + 2][lp] = b.z; Bs[lq * 4 + 3][lp] = b.w;
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        const float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
        const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
        const float aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
      }
    }
  }
  // epilogue
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
    const int wo = (int)(m % P.Wo);
    const long long t = m / P.Wo;
    const int ho = (int)(t % P.Ho), n = (int)(t / P.Ho);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= P.Cout) continue;
      float v = apply_act_precise(acc[i][j] + P.bias[co], P.act, P.slope);
      if (P.res) v += P.res[(((size_t)n * P.Ho + ho) * P.Wo + wo) * P.res_cs + co];
      for (int uy = 0; uy < P.ups; ++uy)
        for (int ux = 0; ux < P.ups; ++ux)
          P.y[(((size_t)n * P.yH + ho * P.ups + uy) * P.yW + wo * P.ups + ux) * P.y_cs + co] = v;
    }
  }
}

int conv_fp32_launch(const yb_conv_desc& d, cudaStream_t s) {
  const yb_tensor& x = d.x;
  const yb_tensor& y = d.y;
  YB_CHECK_ARG(x.dtype == YB_F32 && y.dtype == YB_F32, "conv(fp32): x and y must be fp32");
  YB_CHECK_ARG(x.c_stride % 4 == 0 && d.cin_pad % 4 == 0 && (uintptr_t)x.data % 16 == 0 && (uintptr_t)d.weight % 16 == 0,
               "conv(fp32): input channels must be padded to a multiple of 4 and 16-byte aligned");
  ConvF32Params P;
  P.x = (const float*)x.data; P.y = (float*)y.data; P.w = (const float*)d.weight; P.bias = d.bias;
  P.res = d.res.data ? (const float*)d.res.data : nullptr;
  P.res_cs = d.res.data ? d.res.c_stride : 0;
  P.N = x.N; P.H = x.H; P.W = x.W; P.x_cs = x.c_stride;
  P.Cin = (x.C + 3) / 4 * 4;
  if (P.Cin > d.cin_pad) P.Cin = d.cin_pad;
  P.k = d.k; P.stride = d.stride; P.pad = d.k / 2; P.taps = d.k * d.k; P.cin_pad = d.cin_pad;
  P.Ho = (x.H + 2 * P.pad - d.k) / d.stride + 1;
  P.Wo = (x.W + 2 * P.pad - d.k) / d.stride + 1;
  P.ups = d.upsample; P.yH = y.H; P.yW = y.W; P.Cout = y.C; P.y_cs = y.c_stride;
  P.act = d.act; P.slope = d.leaky_slope;
  const long long M = (long long)P.N * P.Ho * P.Wo;
  dim3 grid((unsigned)((M + 63) / 64), (unsigned)((P.Cout + 63) / 64));
  k_conv_fp32<<<grid, 256, 0, s>>>(P);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

}  // namespace yb
