// First layer of the P6 cfgs on tcgen05: ReOrg (models/common.py:48-53) + Conv 12 -> 64 | 80, 3x3, stride 1
// (cfg/training/yolov7-w6.yaml / yolov7-e6e.yaml node 1, models/common.py:110 Conv.fuseforward), fused with the input
// staging (uint8 NCHW / 255, detect.py:102-103) -- WITHOUT an im2col.
//
// After ReOrg a pixel has 12 channels; padded to 16 it is exactly ONE K = 16 step of a bf16 tcgen05.mma.  The builder
// warps only convert the image into a shared-memory patch of such pixels, stored as two 16-byte planes per pixel
// (channels 0-7 | channels 8-11, 1, 1, 0, 0); the filter window is addressed by the UMMA shared-memory DESCRIPTORS:
//   * a SWIZZLE_NONE K-major operand consists of core matrices of 8 rows x 16 bytes (rows 16 bytes apart); the next 8
//     rows lie SBO bytes further, the second 16-byte half of K lies LBO bytes further;
//   * MMA row m = output pixel ox0 + m of one output row, so 8 consecutive rows = 8 consecutive patch pixels of one
//     plane (SBO = 128) and the two K halves are the two planes (LBO = plane pitch);
//   * tap (dy, dx) is the same descriptor with its start address moved by dy patch rows and dx pixels: nine
//     128 x N x 16 MMAs per 128 output pixels, no data movement between taps.
// The bias rides on the two constant-one channels of the CENTRE tap (split into bf16 hi + lo; padding pixels are all
// zero, and the centre tap is never a padding pixel), SiLU's factor 1/2 is folded into the weights (accumulator = x/2,
// silu(x) = h * tanh(h) + h).  The mma.sync kernel this replaces (stem.cu) took 1.14 ms per 8 images at 1280 x 1280
// (24 % of the yolov7-w6 step) against an HBM floor of 0.07 ms.
#include "common.cuh"
#include "tc_common.cuh"

namespace yb {

namespace {

constexpr int kSrBW = 128;                       // output pixels per tile (one output row segment)
constexpr int kSrSlots = 2;                      // patch buffers / accumulators in flight
constexpr int kSrThreads = 32 * 9;               // warp 0 MMA, 1-4 builders, 5-8 epilogue
constexpr int kSrQuads = kSrBW / 4 + 2;          // aligned 4-pixel quads per patch row: pixels ox0 - 4 ... ox0 + 131
constexpr int kSrPix = kSrQuads * 4;             // 136 patch pixels per row
constexpr int kSrPlane = kSrPix * 16;            // bytes of one plane of one patch row
constexpr int kSrRow = 2 * kSrPlane;             // [plane][pixel][16 B]
constexpr int kSrPatch = 3 * kSrRow;             // three input rows under one output row

struct StemRParams {
  const void* img;              // NCHW planar 3-channel image BEFORE ReOrg (uint8 or fp32)
  int B, H, W;                  // raw image dims
  int Hi, Wi;                   // after ReOrg: H / 2, W / 2 (stride 1, pad 1 -> output is Hi x Wi)
  const __nv_bfloat16* w;       // [N][9][cin_pad], channel order = ReOrg's cat order
  int cin_pad;
  const float* bias;
  float slope;
  __nv_bfloat16* y; int y_cs;
  int tiles_w, total_tiles;
};

// 8 horizontally adjacent image values of one plane and row (= 4 reorganised pixels x {even, odd} column)
template <bool U8> struct Oct;
template <> struct Oct<true> {
  typedef uint2 T;
  static __device__ __forceinline__ T zero() { return make_uint2(0u, 0u); }
  static __device__ __forceinline__ T load(const void* img, long long e) {
    return __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const unsigned char*>(img) + e));
  }
  static __device__ __forceinline__ void to_f(T v, float (&f)[8]) {   // bf16(b / 255) == bf16(b * (1/255)) for every byte b
    const float k = 1.0f / 255.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[i] = (float)((v.x >> (8 * i)) & 0xffu) * k;
      f[4 + i] = (float)((v.y >> (8 * i)) & 0xffu) * k;
    }
  }
};
template <> struct Oct<false> {
  struct T { float4 a, b; };
  static __device__ __forceinline__ T zero() { T t; t.a = t.b = make_float4(0.f, 0.f, 0.f, 0.f); return t; }
  static __device__ __forceinline__ T load(const void* img, long long e) {
    T t;
    t.a = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(img) + e));
    t.b = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(img) + e + 4));
    return t;
  }
  static __device__ __forceinline__ void to_f(T v, float (&f)[8]) {
    f[0] = v.a.x; f[1] = v.a.y; f[2] = v.a.z; f[3] = v.a.w; f[4] = v.b.x; f[5] = v.b.y; f[6] = v.b.z; f[7] = v.b.w;
  }
};

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// SWIZZLE_NONE K-major descriptor (version 1): start | LBO (K halves) | SBO (8-row groups), all in 16-byte units
__device__ __forceinline__ uint64_t desc_none(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct RowWalk {   // (image, output row, tile column) of a persistent CTA's tiles without per-tile divisions
  int tw, oy, b, s_tw, s_oy, s_b;
  __device__ __forceinline__ void init(int t, int step, int tiles_w, int rows) {
    const int per_img = tiles_w * rows;
    b = t / per_img; t -= b * per_img; oy = t / tiles_w; tw = t - oy * tiles_w;
    s_b = step / per_img; step -= s_b * per_img; s_oy = step / tiles_w; s_tw = step - s_oy * tiles_w;
  }
  __device__ __forceinline__ void next(int tiles_w, int rows) {
    tw += s_tw; oy += s_oy; b += s_b;
    if (tw >= tiles_w) { tw -= tiles_w; ++oy; }
    if (oy >= rows) { oy -= rows; ++b; }
  }
};

template <int N> struct SrGeo {
  static constexpr int kTmemCols = kSrSlots * N <= 128 ? 128 : 256;
  static constexpr int kPitch = N * 2 + 16;       // staging row: +16 B keeps thread-per-pixel writes conflict-free
  static constexpr size_t kSmem = 128 + (size_t)kSrSlots * kSrPatch + 9 * N * 32 + 4 * 32 * kPitch + 8 * 4 * kSrSlots + 16;
};

template <int N, bool U8, int ACT>
__global__ void __launch_bounds__(kSrThreads, 2) k_stemr_tc(const StemRParams P) {
  constexpr int PITCH = SrGeo<N>::kPitch;
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* sm = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);
  unsigned char* patch = sm;                                          // [slots][3 rows][2 planes][136 px][16 B]
  unsigned char* b_tile = patch + kSrSlots * kSrPatch;                // [9 taps][2 K halves][N/8 groups][8 rows][16 B]
  unsigned char* out_stage = b_tile + 9 * N * 32;                     // [4 warps][32 px][PITCH]
  uint64_t* bars = reinterpret_cast<uint64_t*>(out_stage + 4 * 32 * PITCH);
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(bars + 4 * kSrSlots);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full0 = smem_u32(bars), bar_empty0 = bar_full0 + 8u * kSrSlots;
  const uint32_t bar_tfull0 = bar_empty0 + 8u * kSrSlots, bar_tempty0 = bar_tfull0 + 8u * kSrSlots;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kSrSlots; ++s) {
      mbar_init(bar_full0 + 8u * s, 128); mbar_init(bar_empty0 + 8u * s, 1);
      mbar_init(bar_tfull0 + 8u * s, 1); mbar_init(bar_tempty0 + 8u * s, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"((uint32_t)SrGeo<N>::kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // Resident weights, one N x 16 tile per tap: k < 12 the filter taps of the 12 reorganised channels, k = 12 / 13 the
  // bias (bf16 hi / lo) on the centre tap -- the patch holds 1.0 there --, k = 14, 15 zero.
  for (int i = threadIdx.x; i < 9 * 2 * N * 8; i += kSrThreads) {
    const int c = i & 7, n = (i >> 3) % N, j = (i / (8 * N)) & 1, tap = i / (16 * N);
    const int k = j * 8 + c;
    const float sc = ACT == YB_ACT_SILU ? 0.5f : 1.0f;
    float v = 0.f;
    if (k < 12) v = __bfloat162float(P.w[((size_t)n * 9 + tap) * P.cin_pad + k]);
    else if (tap == 4 && k == 12) v = __bfloat162float(__float2bfloat16_rn(P.bias[n]));
    else if (tap == 4 && k == 13) v = P.bias[n] - __bfloat162float(__float2bfloat16_rn(P.bias[n]));
    *reinterpret_cast<__nv_bfloat16*>(b_tile + tap * N * 32 + j * N * 16 + (n >> 3) * 128 + (n & 7) * 16 + c * 2) =
        __float2bfloat16_rn(v * sc);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_slot;
  pdl_launch_dependents();
  pdl_wait();   // the image (and the output buffer's previous readers) belong to earlier work in the stream

  if (warp == 0) {
    // ================================ MMA issuer ==================================
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t a_hi = (uint32_t)(desc_none(0, kSrPlane, 128u) >> 32), a_mid = ((uint32_t)kSrPlane >> 4) << 16;
    const uint32_t b_hi = (uint32_t)(desc_none(0, N * 16, 128u) >> 32), b_mid = ((uint32_t)(N * 16) >> 4) << 16;
    const uint32_t b0 = (smem_u32(b_tile) >> 4) & 0x3FFF;
    uint32_t s = 0, ph = 0;
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      mbar_wait(bar_tempty0 + 8u * s, ph ^ 1u);      // the epilogue has taken this slot's accumulator
      mbar_wait(bar_full0 + 8u * s, ph);             // the builders have filled this slot's patch
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t pbase = smem_u32(patch) + s * kSrPatch;
      const uint32_t tmem_d = tmem_base + s * N;
      if (elect_one()) {
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            // output pixel m of the tile reads patch pixel m + dx + 3 (the patch starts 4 pixels left of the tile)
            const uint32_t start = pbase + (uint32_t)dy * kSrRow + (uint32_t)(dx + 3) * 16u;
            const uint32_t a_lo = ((start >> 4) & 0x3FFF) | a_mid;
            const uint32_t b_lo = ((b0 + (uint32_t)(dy * 3 + dx) * (N * 2)) & 0x3FFF) | b_mid;
            asm volatile(
                "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
                "setp.ne.b32 p, %6, 0;\n\t"
                "mov.b64 da, {%1, %3};\n\t"
                "mov.b64 db, {%2, %4};\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
                ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(a_hi), "r"(b_hi), "r"(idesc), "r"((uint32_t)(dy | dx)) : "memory");
          }
        }
        umma_commit(bar_empty0 + 8u * s);
        umma_commit(bar_tfull0 + 8u * s);
      }
      __syncwarp();
      if (++s == kSrSlots) { s = 0; ph ^= 1u; }
    }
  } else if (warp <= 4) {
    // ================================ patch builders ==============================
    const int bt = threadIdx.x - 32;                 // 0..127; 3 rows x 34 quads = 102 of them carry a quad
    const int q_row = bt / kSrQuads, q_quad = bt - q_row * kSrQuads;
    const bool active = q_row < 3;
    const long long plane = (long long)P.H * P.W;
    RowWalk wk;
    wk.init(blockIdx.x, gridDim.x, P.tiles_w, P.Hi);
    uint32_t s = 0, ph = 1;
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      const int yp = wk.oy - 1 + q_row, x0 = wk.tw * kSrBW - 4 + 4 * q_quad;
      const long long base = (long long)wk.b * 3 * plane + (long long)(2 * yp) * P.W + 2 * x0;
      wk.next(P.tiles_w, P.Hi);
      const bool in = active && yp >= 0 && yp < P.Hi && x0 >= 0 && x0 < P.Wi;   // Wi % 4 == 0: whole quads
      float f[3][2][8];
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          typename Oct<U8>::T o = in ? Oct<U8>::load(P.img, base + c * plane + (long long)r * P.W) : Oct<U8>::zero();
          Oct<U8>::to_f(o, f[c][r]);
        }
      mbar_wait(bar_empty0 + 8u * s, ph);            // the MMAs that read this patch buffer have retired
      if (active) {
        unsigned char* rowp = patch + s * kSrPatch + q_row * kSrRow + (4 * q_quad) * 16;
        const uint32_t ones = in ? 0x3F803F80u : 0u;                             // halo pixels are all-zero (conv padding)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int ev = 2 * e, od = 2 * e + 1;
          // ReOrg channel order: (row even, col even) c0-2 | (row odd, col even) | (row even, col odd) | (row odd, col odd)
          uint4 p0, p1;
          p0.x = pack_bf16(f[0][0][ev], f[1][0][ev]);
          p0.y = pack_bf16(f[2][0][ev], f[0][1][ev]);
          p0.z = pack_bf16(f[1][1][ev], f[2][1][ev]);
          p0.w = pack_bf16(f[0][0][od], f[1][0][od]);
          p1.x = pack_bf16(f[2][0][od], f[0][1][od]);
          p1.y = pack_bf16(f[1][1][od], f[2][1][od]);
          p1.z = ones;                                                           // channels 12, 13 = 1.0 (bias hi, lo)
          p1.w = 0u;
          *reinterpret_cast<uint4*>(rowp + e * 16) = p0;
          *reinterpret_cast<uint4*>(rowp + kSrPlane + e * 16) = p1;
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the MMA
      mbar_arrive(bar_full0 + 8u * s);
      if (++s == kSrSlots) { s = 0; ph ^= 1u; }
    }
  } else {
    // ================================ epilogue ====================================
    constexpr int CH = N / 8;                        // 16-byte chunks per output pixel
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    unsigned char* stg = out_stage + (warp - 5) * 32 * PITCH;
    RowWalk wk;
    wk.init(blockIdx.x, gridDim.x, P.tiles_w, P.Hi);
    uint32_t s = 0, ph = 0;
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      const int b = wk.b, oy = wk.oy, ox0 = wk.tw * kSrBW + q * 32;
      wk.next(P.tiles_w, P.Hi);
      mbar_wait(bar_tfull0 + 8u * s, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int c16 = 0; c16 < N / 16; ++c16) {
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + s * N + c16 * 16, v);
        if (c16 == N / 16 - 1) {                     // the whole accumulator row is in registers: hand the slot back
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_tempty0 + 8u * s);
        }
        uint32_t p[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float f0 = __uint_as_float(v[2 * e]), f1 = __uint_as_float(v[2 * e + 1]);   // bias already accumulated
          if (ACT == YB_ACT_SILU) {
            float t0, t1;
            asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(f0));
            asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(f1));
            f0 = fmaf(f0, t0, f0); f1 = fmaf(f1, t1, f1);
          } else if (ACT == YB_ACT_LEAKY) {
            f0 = f0 > 0.f ? f0 : f0 * P.slope; f1 = f1 > 0.f ? f1 : f1 * P.slope;
          }
          p[e] = pack_bf16(f0, f1);
        }
        *reinterpret_cast<uint4*>(stg + lane * PITCH + (2 * c16) * 16) = make_uint4(p[0], p[1], p[2], p[3]);
        *reinterpret_cast<uint4*>(stg + lane * PITCH + (2 * c16 + 1) * 16) = make_uint4(p[4], p[5], p[6], p[7]);
      }
      __syncwarp();
      // 32 pixels x CH chunks leave as runs of consecutive 16-byte chunks: a store instruction covers 512 contiguous
      // bytes (when the output is not a concat slice) instead of 32 pixel-strided pieces
      __nv_bfloat16* dst0 = P.y + (((size_t)b * P.Hi + oy) * P.Wi + ox0) * P.y_cs;
#pragma unroll
      for (int i = 0; i < CH; ++i) {
        const int id = i * 32 + lane, px = id / CH, ch = id - px * CH;
        const uint4 val = *reinterpret_cast<const uint4*>(stg + px * PITCH + ch * 16);
        if (ox0 + px < P.Wi) *reinterpret_cast<uint4*>(dst0 + (size_t)px * P.y_cs + ch * 8) = val;
      }
      __syncwarp();                                  // the staging buffer is rewritten by the next tile
      if (++s == kSrSlots) { s = 0; ph ^= 1u; }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"((uint32_t)SrGeo<N>::kTmemCols) : "memory");
  }
}

template <int N, bool U8, int ACT>
int launch_one(const StemRParams& P, int grid, cudaStream_t st) {
  constexpr size_t smem = SrGeo<N>::kSmem;
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_stemr_tc<N, U8, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  YB_CUDA(launch_pdl(k_stemr_tc<N, U8, ACT>, grid, kSrThreads, smem, st, P));
  return YB_OK;
}

template <int N>
int launch_n(const StemRParams& P, int img_u8, int act, int grid, cudaStream_t st) {
#define YB_SR(U8_)                                                                         \
  (act == YB_ACT_SILU ? launch_one<N, U8_, YB_ACT_SILU>(P, grid, st)                        \
   : act == YB_ACT_LEAKY ? launch_one<N, U8_, YB_ACT_LEAKY>(P, grid, st)                    \
                         : launch_one<N, U8_, YB_ACT_NONE>(P, grid, st))
  return img_u8 ? YB_SR(true) : YB_SR(false);
#undef YB_SR
}

}  // namespace

// ReOrg + 12 -> 64 | 80 first layer (stride 1); anything else returns YB_ERR_UNSUPPORTED so that the caller falls back
// to the mma.sync stem (stem.cu).
int stemr_tc_launch(const void* img, int img_u8, int B, int H, int W, const void* weight, const float* bias, int cin_pad,
                    int stride, int act, float slope, const yb_tensor* y, cudaStream_t st) {
  if ((y->C != 64 && y->C != 80) || y->dtype != YB_BF16 || stride != 1 || cin_pad < 12) return YB_ERR_UNSUPPORTED;
  // builders load aligned groups of 8 image columns (= 4 reorganised pixels) from two image rows
  if (W % 8 != 0 || H % 2 != 0 || (uintptr_t)img % 16 != 0) return YB_ERR_UNSUPPORTED;
  StemRParams P;
  P.img = img; P.B = B; P.H = H; P.W = W; P.Hi = H / 2; P.Wi = W / 2;
  P.w = (const __nv_bfloat16*)weight; P.cin_pad = cin_pad; P.bias = bias; P.slope = slope;
  P.y = (__nv_bfloat16*)y->data; P.y_cs = y->c_stride;
  P.tiles_w = ceil_div(P.Wi, kSrBW);
  const long long total = (long long)P.tiles_w * P.Hi * B;
  if (total > 0x7fffffffLL) return YB_ERR_UNSUPPORTED;
  P.total_tiles = (int)total;
  const int cap = 2 * num_sms();   // two CTAs per SM (registers; TMEM 2 x 256 columns at N = 80)
  const int grid = (int)(total < cap ? total : cap);
  return y->C == 64 ? launch_n<64>(P, img_u8, act, grid, st) : launch_n<80>(P, img_u8, act, grid, st);
}

}  // namespace yb
