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

// (image, tile row, tile column) of a persistent CTA's current tile, advanced by a fixed tile-count step without
// divisions (the decode of a linear tile index costs two integer divisions per warp per tile otherwise).
struct TileWalk {
  int tw, th, b;
  int s_tw, s_th, s_b;
  __device__ __forceinline__ void init(int t, int step, int tiles_w, int tiles_h) {
    const int per_img = tiles_w * tiles_h;
    b = t / per_img; t -= b * per_img; th = t / tiles_w; tw = t - th * tiles_w;
    s_b = step / per_img; step -= s_b * per_img; s_th = step / tiles_w; s_tw = step - s_th * tiles_w;
  }
  __device__ __forceinline__ void next(int tiles_w, int tiles_h) {
    tw += s_tw; th += s_th; b += s_b;
    if (tw >= tiles_w) { tw -= tiles_w; ++th; }
    if (th >= tiles_h) { th -= tiles_h; ++b; }
  }
};

// Four horizontally adjacent pixels of one image row: one aligned 32-bit (uint8) / 128-bit (fp32) load.
template <bool U8> struct Quad;
template <> struct Quad<true> {
  typedef uint32_t T;
  static constexpr int kPrefetch = 4;           // tiles in flight per builder thread (registers: 1 word each)
  static __device__ __forceinline__ T zero() { return 0u; }
  static __device__ __forceinline__ T load(const void* img, size_t elem) {
    return __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const unsigned char*>(img) + elem));
  }
  static __device__ __forceinline__ uint2 to_bf16(T v) {   // bf16(b / 255) == bf16(b * (1/255)) for every byte b
    const float k = 1.0f / 255.0f;
    __nv_bfloat162 lo = __floats2bfloat162_rn((float)(v & 0xffu) * k, (float)((v >> 8) & 0xffu) * k);
    __nv_bfloat162 hi = __floats2bfloat162_rn((float)((v >> 16) & 0xffu) * k, (float)(v >> 24) * k);
    return make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
};
template <> struct Quad<false> {
  typedef float4 T;
  static constexpr int kPrefetch = 2;
  static __device__ __forceinline__ T zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
  static __device__ __forceinline__ T load(const void* img, size_t elem) {
    return __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(img) + elem));
  }
  static __device__ __forceinline__ uint2 to_bf16(T v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    return make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
};

template <int S, bool U8, int ACT>
__global__ void __launch_bounds__(kStThreads, 3) k_stem_tc(const StemTcParams P) {
  // input patch of one tile, in 4-pixel quads starting at x = x0*S - 4 (keeps every load aligned)
  constexpr int PH = (kStBH - 1) * S + 3;
  constexpr int PQ = ((kStBW - 1) * S + 3 + 3 + 3) / 4;     // quads per patch row: pixels x0*S-4 .. x0*S+(BW-1)*S+1
  constexpr int PWP = PQ * 4;                               // patch row pitch (elements)
  constexpr int NQ = 3 * PH * PQ;                           // quads per tile
  constexpr int PER_T = (NQ + 127) / 128;
  constexpr int D = Quad<U8>::kPrefetch;
  typedef typename Quad<U8>::T quad_t;

  constexpr int PATCH_ELEMS = 3 * PH * PWP;
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* sm = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char (*a_tiles)[128 * 64] = reinterpret_cast<unsigned char (*)[128 * 64]>(sm);             // [stages]
  unsigned char* b_tile = sm + kStStages * 8192;                                                       // 2 KB
  unsigned char* out_stage = b_tile + kStCout * 64;                                                    // [4 warps][2][32 px][64 B]
  __nv_bfloat16 (*patch)[PATCH_ELEMS] = reinterpret_cast<__nv_bfloat16 (*)[PATCH_ELEMS]>(out_stage + 4 * 2 * 2048);
  uint64_t* bars = reinterpret_cast<uint64_t*>(out_stage + 4 * 2 * 2048 + ((2 * PATCH_ELEMS * 2 + 15) / 16) * 16);
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(bars + 2 * kStStages + 2 * kStAcc);

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
  // Resident weight tile B[n][k], 64-byte rows, SW64 (16-byte chunk ^= (row >> 1) & 3):
  //   k = (c*3 + dy)*3 + dx < 27 : filter taps;  k = 27, 28 : bias split into bf16 hi + lo (A holds 1.0 there);
  //   SiLU: everything scaled by 0.5 (exact), so the accumulator is h = x/2 and silu(x) = h*tanh(h) + h.
  for (int i = threadIdx.x; i < kStCout * kStK; i += kStThreads) {
    const int n = i >> 5, k = i & 31;
    const float sc = ACT == YB_ACT_SILU ? 0.5f : 1.0f;
    float v = 0.f;
    if (k < 27) {
      const int c = k / 9, tap = k - c * 9;
      v = __bfloat162float(P.w[((size_t)n * 9 + tap) * P.cin_pad + c]);
    } else if (k == 27) {
      v = __bfloat162float(__float2bfloat16_rn(P.bias[n]));
    } else if (k == 28) {
      v = P.bias[n] - __bfloat162float(__float2bfloat16_rn(P.bias[n]));
    }
    const uint32_t off = (uint32_t)n * 64u + ((((uint32_t)k >> 3) ^ (((uint32_t)n >> 1) & 3u)) << 4) + ((uint32_t)k & 7u) * 2u;
    *reinterpret_cast<__nv_bfloat16*>(b_tile + off) = __float2bfloat16_rn(v * sc);
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
    const size_t plane = (size_t)P.H * P.W;
    // tile-invariant description of the quads this thread fetches
    int q_r[PER_T], q_x[PER_T], q_dst[PER_T];
    size_t q_off[PER_T];
#pragma unroll
    for (int i = 0; i < PER_T; ++i) {
      const int w = bt + 128 * i;
      const int prow = w / PQ, pq = w - prow * PQ;
      const int c = prow / PH, r = prow - c * PH;
      q_r[i] = w < NQ ? r : (1 << 28);               // out-of-range rows never pass the bounds test
      q_x[i] = 4 * pq;
      q_dst[i] = prow * PWP + 4 * pq;
      q_off[i] = (size_t)c * plane + (size_t)r * P.W + 4 * pq;
    }
    quad_t regs[D][PER_T];
    TileWalk fw;                                     // tile the next fetch() is for
    fw.init(blockIdx.x, gridDim.x, P.tiles_w, P.tiles_h);
    auto fetch = [&](quad_t (&dst)[PER_T]) {
      const int y0 = fw.th * kStBH * S - 1, x0 = fw.tw * kStBW * S - 4;
      const long long base = (long long)fw.b * 3 * (long long)plane + (long long)y0 * P.W + x0;
#pragma unroll
      for (int i = 0; i < PER_T; ++i) {
        const int yy = y0 + q_r[i], xx = x0 + q_x[i];
        dst[i] = Quad<U8>::zero();
        if (fw.b < P.B && yy >= 0 && yy < P.H && xx >= 0 && xx < P.W)
          dst[i] = Quad<U8>::load(P.img, (size_t)(base + (long long)q_off[i]));
      }
      fw.next(P.tiles_w, P.tiles_h);
    };
#pragma unroll
    for (int d = 0; d < D; ++d) fetch(regs[d]);
    uint32_t s = 0, ph = 1;
    const uint32_t one2 = 0x3F803F80u;               // bf16 {1.0, 1.0}: the bias columns k = 27, 28
    bool more = true;
    for (int t0 = blockIdx.x; more; t0 += D * gridDim.x) {
#pragma unroll
      for (int d = 0; d < D; ++d) {
        const int t = t0 + d * gridDim.x;
        if (t >= P.total_tiles) { more = false; break; }
        __nv_bfloat16* pb = patch[d & 1];
#pragma unroll
        for (int i = 0; i < PER_T; ++i)
          if (q_r[i] < (1 << 28)) *reinterpret_cast<uint2*>(pb + q_dst[i]) = Quad<U8>::to_bf16(regs[d][i]);
        fetch(regs[d]);                              // D tiles ahead: hides the global-load latency
        asm volatile("bar.sync 1, 128;" ::: "memory");
        mbar_wait(bar_empty0 + 8u * s, ph);
        // row bt: k = (c*3 + dy)*3 + dx  <-  patch[c][ty*S + dy][tx*S + dx + 3]
        __nv_bfloat16 kv[28];
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            const __nv_bfloat16* src = pb + (c * PH + ty * S + dy) * PWP + tx * S + 3;
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) kv[(c * 3 + dy) * 3 + dx] = src[dx];
          }
        kv[27] = __float2bfloat16_rn(1.0f);
        const uint32_t row = smem_u32(a_tiles[s]) + (uint32_t)bt * 64u;
        const uint32_t xr = ((uint32_t)bt >> 1) & 3u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t p[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k0 = 8 * j + 2 * e;
            if (k0 + 1 < 28) {
              __nv_bfloat162 hh = __halves2bfloat162(kv[k0], kv[k0 + 1]);
              p[e] = *reinterpret_cast<uint32_t*>(&hh);
            } else {
              p[e] = k0 == 28 ? (one2 & 0xffffu) : 0u;   // k = 28 -> 1.0, k >= 29 -> 0
            }
          }
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((((uint32_t)j) ^ xr) << 4)), "r"(p[0]),
                       "r"(p[1]), "r"(p[2]), "r"(p[3]) : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
        mbar_arrive(bar_full0 + 8u * s);
        if (++s == kStStages) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ================================ epilogue ====================================
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;
    const int ty = row / kStBW, tx = row - ty * kStBW;
    uint32_t acc = 0, aph = 0, obuf = 0;
    TileWalk tw_;
    tw_.init(blockIdx.x, gridDim.x, P.tiles_w, P.tiles_h);
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      const int b = tw_.b;
      const int oy = tw_.th * kStBH + ty, ox = tw_.tw * kStBW + tx;
      tw_.next(P.tiles_w, P.tiles_h);
      mbar_wait(bar_tfull0 + 8u * acc, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * kStCout, v);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty0 + 8u * acc);    // accumulator is in registers: hand it back
      if (++acc == kStAcc) { acc = 0; aph ^= 1u; }
      uint32_t p[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        float f0 = __uint_as_float(v[2 * e]), f1 = __uint_as_float(v[2 * e + 1]);   // bias already accumulated
        if (ACT == YB_ACT_SILU) {
          float t0, t1;
          asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(f0));
          asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(f1));
          f0 = fmaf(f0, t0, f0); f1 = fmaf(f1, t1, f1);
        } else if (ACT == YB_ACT_LEAKY) {
          f0 = f0 > 0.f ? f0 : f0 * P.slope; f1 = f1 > 0.f ? f1 : f1 * P.slope;
        }
        __nv_bfloat162 hh = __floats2bfloat162_rn(f0, f1);
        p[e] = *reinterpret_cast<uint32_t*>(&hh);
      }
      // Transpose through shared memory (64-byte rows, 16-byte chunks XOR-swizzled by (row >> 1) & 3: conflict-free
      // stores and loads) so that each store instruction writes eight pixels x 64 contiguous bytes instead of 32
      // half-filled sectors.
      unsigned char* stg = out_stage + ((warp - 5) * 2 + (int)obuf) * 2048;
      const uint32_t wx = ((uint32_t)lane >> 1) & 3u;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ wx) << 4)) = make_uint4(p[4 * j], p[4 * j + 1], p[4 * j + 2], p[4 * j + 3]);
      __syncwarp();
      const int ox0 = ox - lane;                       // first pixel of this warp's 32-pixel run (same output row)
      __nv_bfloat16* dst0 = P.y + (((size_t)b * P.Ho + oy) * P.Wo + ox0) * P.y_cs + (lane & 3) * 8;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int pix = (lane >> 2) + 8 * i;
        const uint32_t rx = ((uint32_t)pix >> 1) & 3u;
        const uint4 val = *reinterpret_cast<const uint4*>(stg + pix * 64 + ((((uint32_t)lane & 3u) ^ rx) << 4));
        if (oy < P.Ho && ox0 + pix < P.Wo) *reinterpret_cast<uint4*>(dst0 + (size_t)pix * P.y_cs) = val;
      }
      obuf ^= 1u;                                      // next tile stages into the other buffer
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

template <int S>
constexpr size_t stem_smem_bytes() {
  constexpr int PH = (kStBH - 1) * S + 3, PWP = (((kStBW - 1) * S + 9) / 4) * 4;
  return 1024 + kStStages * 8192 + kStCout * 64 + 4 * 2 * 2048 + ((2 * 3 * PH * PWP * 2 + 15) / 16) * 16 +
         8 * (2 * kStStages + 2 * kStAcc) + 16;
}

template <int S, bool U8, int ACT>
int launch_one(const StemTcParams& P, int grid, cudaStream_t st) {
  constexpr size_t smem = stem_smem_bytes<S>();
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_stem_tc<S, U8, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  YB_CUDA(launch_pdl(k_stem_tc<S, U8, ACT>, grid, kStThreads, smem, st, P));
  return YB_OK;
}

template <int S, bool U8>
int launch_act(const StemTcParams& P, int grid, cudaStream_t st) {
  if (P.act == YB_ACT_SILU) return launch_one<S, U8, YB_ACT_SILU>(P, grid, st);
  if (P.act == YB_ACT_LEAKY) return launch_one<S, U8, YB_ACT_LEAKY>(P, grid, st);
  return launch_one<S, U8, YB_ACT_NONE>(P, grid, st);
}

}  // namespace

// 3-channel, 32-filter first layer; anything else returns YB_ERR_UNSUPPORTED so that the caller falls back to
// the mma.sync stem (ReOrg / 12-channel inputs, other widths).
int stem_tc_launch(const void* img, int img_u8, int B, int H, int W, const void* weight, const float* bias,
                   int cin_pad, int stride, int act, float slope, const yb_tensor* y, cudaStream_t st) {
  if (y->C != kStCout || y->dtype != YB_BF16 || (stride != 1 && stride != 2)) return YB_ERR_UNSUPPORTED;
  if (W % 4 != 0 || (uintptr_t)img % 16 != 0) return YB_ERR_UNSUPPORTED;   // the builders load aligned 4-pixel quads
  StemTcParams P;
  P.img = img; P.B = B; P.H = H; P.W = W;
  P.Ho = (H + 2 - 3) / stride + 1; P.Wo = (W + 2 - 3) / stride + 1;
  P.w = (const __nv_bfloat16*)weight; P.cin_pad = cin_pad; P.bias = bias; P.act = act; P.slope = slope;
  P.y = (__nv_bfloat16*)y->data; P.y_cs = y->c_stride;
  P.tiles_w = ceil_div(P.Wo, kStBW); P.tiles_h = ceil_div(P.Ho, kStBH);
  const long long total = (long long)P.tiles_w * P.tiles_h * B;
  P.total_tiles = (int)total;
  const int cap = 3 * num_sms();   // three CTAs per SM: the builders' load -> convert -> build chain is latency-bound
  const int grid = (int)(total < cap ? total : cap);
  if (stride == 1) return img_u8 ? launch_act<1, true>(P, grid, st) : launch_act<1, false>(P, grid, st);
  return img_u8 ? launch_act<2, true>(P, grid, st) : launch_act<2, false>(P, grid, st);
}

}  // namespace yb
