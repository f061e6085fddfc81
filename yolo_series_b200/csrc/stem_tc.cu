// First layer (3 -> 32 channels, 3x3, stride 1|2) on tcgen05, fused with the input staging.
//
// Path: reference models/yolo.py Model.forward_once -> models/common.py Conv.forward on the raw image
// (utils/datasets.py hands uint8 NCHW, detect.py divides by 255).  K = 27 is too small for a TMA-fed im2col
// (8-byte taps), so four builder warps assemble the [128 px][32 k] bf16 A tile in shared memory themselves
// (64-byte rows in the SW64 UMMA layout), one elected lane issues two 128x32x16 MMAs per tile against a
// CTA-resident weight tile, and four epilogue warps take the accumulator out of TMEM (bias + activation)
// and store 64 contiguous bytes per pixel.  Every image byte is read from global memory once per tile,
// converted once (bf16(v/255) == bf16(v * (1/255)) for all 256 byte values, checked exhaustively on the host).
#include "common.cuh"
#include "tc_common.cuh"

namespace yb {

namespace {

constexpr int kStBW = 64, kStBH = 2;            // output tile: 64 x 2 pixels = 128 MMA rows
constexpr int kStStages = 4;                    // A-tile ring
constexpr int kStAcc = 4;                       // TMEM accumulators (32 columns each)
constexpr int kStThreads = 32 * 9;              // warp 0 MMA, 1-4 builders, 5-8 epilogue
constexpr int kStCout = 32, kStK = 32;

struct StemTcParams {
  const void* img;
  int B, H, W;                  // 3-channel planar image
  int Ho, Wo;
  const __nv_bfloat16* w;       // [32][9][cin_pad]
  int cin_pad;
  const float* bias;
  int act; float slope;
  __nv_bfloat16* y; int y_cs;
  int tiles_w, tiles_h, total_tiles;
};

template <bool U8> struct ImgElem;
template <> struct ImgElem<true> {
  typedef unsigned char T;
  static __device__ __forceinline__ __nv_bfloat16 cvt(unsigned char v) { return __float2bfloat16_rn((float)v * (1.0f / 255.0f)); }
};
template <> struct ImgElem<false> {
  typedef float T;
  static __device__ __forceinline__ __nv_bfloat16 cvt(float v) { return __float2bfloat16_rn(v); }
};

template <int S, bool U8, int ACT>
__global__ void __launch_bounds__(kStThreads, 1) k_stem_tc(const StemTcParams P) {
  constexpr int PH = (kStBH - 1) * S + 3, PW = (kStBW - 1) * S + 3;   // input patch of one tile
  constexpr int PWP = PW + 1;                                         // row pitch (elements)
  constexpr int PATCH = 3 * PH * PW;
  constexpr int PER_T = (PATCH + 127) / 128;
  typedef typename ImgElem<U8>::T in_t;

  __shared__ __align__(1024) unsigned char a_tiles[kStStages][128 * 64];
  __shared__ __align__(1024) unsigned char b_tile[kStCout * 64];
  __shared__ __align__(16) __nv_bfloat16 patch[2][3 * PH * PWP];
  __shared__ __align__(8) uint64_t bars[2 * kStStages + 2 * kStAcc];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full0 = smem_u32(bars), bar_empty0 = bar_full0 + 8u * kStStages;
  const uint32_t bar_tfull0 = bar_empty0 + 8u * kStStages, bar_tempty0 = bar_tfull0 + 8u * kStAcc;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStStages; ++s) { mbar_init(bar_full0 + 8u * s, 128); mbar_init(bar_empty0 + 8u * s, 1); }
    for (int i = 0; i < kStAcc; ++i) { mbar_init(bar_tfull0 + 8u * i, 1); mbar_init(bar_tempty0 + 8u * i, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"((uint32_t)(kStAcc * kStCout)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // resident weight tile: B[n][k], k = (c*3 + dy)*3 + dx, 64-byte rows, SW64 (16-byte chunk ^= (row >> 1) & 3)
  for (int i = threadIdx.x; i < kStCout * kStK; i += kStThreads) {
    const int n = i >> 5, k = i & 31;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (k < 27) {
      const int c = k / 9, tap = k - c * 9;
      v = P.w[((size_t)n * 9 + tap) * P.cin_pad + c];
    }
    const uint32_t off = (uint32_t)n * 64u + ((((uint32_t)k >> 3) ^ (((uint32_t)n >> 1) & 3u)) << 4) + ((uint32_t)k & 7u) * 2u;
    *reinterpret_cast<__nv_bfloat16*>(b_tile + off) = v;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_slot;
  const int tiles_per_img = P.tiles_w * P.tiles_h;

  if (warp == 0) {
    // ================================ MMA issuer ==================================
    const uint32_t hi32 = (uint32_t)(make_smem_desc(0, 512u, 4u) >> 32);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kStCout >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t b_lo = (smem_u32(b_tile) >> 4) & 0x3FFF;
    uint32_t s = 0, ph = 0, acc = 0, aph = 1;
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      mbar_wait(bar_tempty0 + 8u * acc, aph);
      mbar_wait(bar_full0 + 8u * s, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a_lo = (smem_u32(a_tiles[s]) >> 4) & 0x3FFF;
      if (elect_one()) {
        umma_bf16_lh(tmem_base + acc * kStCout, a_lo, b_lo, hi32, idesc, 0u);
        umma_bf16_lh(tmem_base + acc * kStCout, a_lo + 2u, b_lo + 2u, hi32, idesc, 1u);
        umma_commit(bar_empty0 + 8u * s);
        umma_commit(bar_tfull0 + 8u * acc);
      }
      __syncwarp();
      if (++s == kStStages) { s = 0; ph ^= 1u; }
      if (++acc == kStAcc) { acc = 0; aph ^= 1u; }
    }
  } else if (warp <= 4) {
    // ================================ A-tile builders =============================
    const int bt = threadIdx.x - 32;                 // 0..127 = MMA row = output pixel of the tile
    const int ty = bt / kStBW, tx = bt - ty * kStBW;
    const in_t* img = reinterpret_cast<const in_t*>(P.img);
    const size_t plane = (size_t)P.H * P.W;
    in_t regs[PER_T];
    auto fetch = [&](int t) {
      int m = t;
      const int b = m / tiles_per_img;
      m -= b * tiles_per_img;
      const int th = m / P.tiles_w, tw = m - th * P.tiles_w;
      const int y0 = th * kStBH * S - 1, x0 = tw * kStBW * S - 1;
      const in_t* base = img + (size_t)b * 3 * plane;
#pragma unroll
      for (int i = 0; i < PER_T; ++i) {
        const int e = bt + 128 * i;
        const int c = e / (PH * PW), rem = e - c * (PH * PW);
        const int r = rem / PW, x = rem - r * PW;
        const int yy = y0 + r, xx = x0 + x;
        in_t v = (in_t)0;
        if (e < PATCH && yy >= 0 && yy < P.H && xx >= 0 && xx < P.W) v = __ldg(base + (size_t)c * plane + (size_t)yy * P.W + xx);
        regs[i] = v;
      }
    };
    uint32_t s = 0, ph = 1;
    int it = 0;
    if ((int)blockIdx.x < P.total_tiles) fetch(blockIdx.x);
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x, ++it) {
      __nv_bfloat16* pb = patch[it & 1];
#pragma unroll
      for (int i = 0; i < PER_T; ++i) {
        const int e = bt + 128 * i;
        const int c = e / (PH * PW), rem = e - c * (PH * PW);
        const int r = rem / PW, x = rem - r * PW;
        if (e < PATCH) pb[(c * PH + r) * PWP + x] = ImgElem<U8>::cvt(regs[i]);
      }
      if (t + (int)gridDim.x < P.total_tiles) fetch(t + gridDim.x);   // next tile's bytes fly during this build
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(bar_empty0 + 8u * s, ph);
      // row bt: k = (c*3 + dy)*3 + dx  <-  patch[c][ty*S + dy][tx*S + dx]
      __nv_bfloat16 kv[32];
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
          const __nv_bfloat16* src = pb + (c * PH + ty * S + dy) * PWP + tx * S;
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) kv[(c * 3 + dy) * 3 + dx] = src[dx];
        }
#pragma unroll
      for (int k = 27; k < 32; ++k) kv[k] = __float2bfloat16_rn(0.f);
      const uint32_t row = smem_u32(a_tiles[s]) + (uint32_t)bt * 64u;
      const uint32_t xr = ((uint32_t)bt >> 1) & 3u;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t p[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          __nv_bfloat162 hh = __halves2bfloat162(kv[8 * j + 2 * e], kv[8 * j + 2 * e + 1]);
          p[e] = *reinterpret_cast<uint32_t*>(&hh);
        }
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((((uint32_t)j) ^ xr) << 4)), "r"(p[0]),
                     "r"(p[1]), "r"(p[2]), "r"(p[3]) : "memory");
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
      mbar_arrive(bar_full0 + 8u * s);
      if (++s == kStStages) { s = 0; ph ^= 1u; }
    }
  } else {
    // ================================ epilogue ====================================
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;
    const int ty = row / kStBW, tx = row - ty * kStBW;
    float bias_r[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) bias_r[i] = P.bias[i];
    uint32_t acc = 0, aph = 0;
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      int m = t;
      const int b = m / tiles_per_img;
      m -= b * tiles_per_img;
      const int th = m / P.tiles_w, tw = m - th * P.tiles_w;
      const int oy = th * kStBH + ty, ox = tw * kStBW + tx;
      mbar_wait(bar_tfull0 + 8u * acc, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * kStCout, v);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty0 + 8u * acc);    // accumulator is in registers: hand it back
      if (++acc == kStAcc) { acc = 0; aph ^= 1u; }
      if (oy < P.Ho && ox < P.Wo) {
        __nv_bfloat16* dst = P.y + (((size_t)b * P.Ho + oy) * P.Wo + ox) * P.y_cs;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t p[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i0 = 8 * j + 2 * e;
            float f0 = __uint_as_float(v[i0]) + bias_r[i0], f1 = __uint_as_float(v[i0 + 1]) + bias_r[i0 + 1];
            if (ACT == YB_ACT_SILU) { f0 = silu_fast(f0); f1 = silu_fast(f1); }
            else if (ACT == YB_ACT_LEAKY) { f0 = f0 > 0.f ? f0 : f0 * P.slope; f1 = f1 > 0.f ? f1 : f1 * P.slope; }
            __nv_bfloat162 hh = __floats2bfloat162_rn(f0, f1);
            p[e] = *reinterpret_cast<uint32_t*>(&hh);
          }
          *reinterpret_cast<uint4*>(dst + 8 * j) = make_uint4(p[0], p[1], p[2], p[3]);
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"((uint32_t)(kStAcc * kStCout)) : "memory");
  }
}

template <int S, bool U8>
int launch_act(const StemTcParams& P, int grid, cudaStream_t st) {
  if (P.act == YB_ACT_SILU) k_stem_tc<S, U8, YB_ACT_SILU><<<grid, kStThreads, 0, st>>>(P);
  else if (P.act == YB_ACT_LEAKY) k_stem_tc<S, U8, YB_ACT_LEAKY><<<grid, kStThreads, 0, st>>>(P);
  else k_stem_tc<S, U8, YB_ACT_NONE><<<grid, kStThreads, 0, st>>>(P);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

}  // namespace

// 3-channel, 32-filter first layer; anything else returns YB_ERR_UNSUPPORTED so that the caller falls back to
// the mma.sync stem (ReOrg / 12-channel inputs, other widths).
int stem_tc_launch(const void* img, int img_u8, int B, int H, int W, const void* weight, const float* bias,
                   int cin_pad, int stride, int act, float slope, const yb_tensor* y, cudaStream_t st) {
  if (y->C != kStCout || y->dtype != YB_BF16 || (stride != 1 && stride != 2)) return YB_ERR_UNSUPPORTED;
  StemTcParams P;
  P.img = img; P.B = B; P.H = H; P.W = W;
  P.Ho = (H + 2 - 3) / stride + 1; P.Wo = (W + 2 - 3) / stride + 1;
  P.w = (const __nv_bfloat16*)weight; P.cin_pad = cin_pad; P.bias = bias; P.act = act; P.slope = slope;
  P.y = (__nv_bfloat16*)y->data; P.y_cs = y->c_stride;
  P.tiles_w = ceil_div(P.Wo, kStBW); P.tiles_h = ceil_div(P.Ho, kStBH);
  const long long total = (long long)P.tiles_w * P.tiles_h * B;
  P.total_tiles = (int)total;
  const int grid = (int)(total < num_sms() ? total : num_sms());
  if (stride == 1) return img_u8 ? launch_act<1, true>(P, grid, st) : launch_act<1, false>(P, grid, st);
  return img_u8 ? launch_act<2, true>(P, grid, st) : launch_act<2, false>(P, grid, st);
}

}  // namespace yb
