// bf16 implicit-GEMM convolution on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM),
// operands staged by TMA.  Replaces Conv.fuseforward / fused RepConv / head 1x1 convs of the
// reference (models/common.py:110-111, 498-500; models/yolo.py:145) after host-side folding.
//
// GEMM view:  D[pixel][cout] = sum_{tap, cin} X[pixel shifted by tap][cin] * W[cout][tap][cin]
//   M tile = 128 output pixels = a BW x BH spatial box of one image (or 128 consecutive pixels of
//            the flattened N*H*W axis for 1x1/s1 convs), loaded per filter tap with a 4-D TMA box
//            whose negative / overflowing coordinates are zero-filled by the hardware (= conv padding);
//            stride-2 convs use the tensor map's elementStrides.
//   N tile = BN output channels (32/64/128/256), weights [Cout][taps][Cin] loaded with a 3-D TMA box.
//   K loop = taps x ceil(Cin/BK) chunks, BK = 64/32/16 channels (SWIZZLE_128B/64B/32B, K-major).
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (+TMEM alloc), warps 2..5 = epilogue
// (tcgen05.ld -> +bias -> SiLU/LeakyReLU -> bf16/fp32 -> swizzled smem -> TMA store).  The store
// tensor map addresses a channel slice of the consumer's concat buffer (Concat fused away) and, for
// convs feeding nn.Upsample, four maps with doubled strides write the 2x2 replicas.
#include "tc_common.cuh"

namespace yb {

struct ConvTcArgs {
  CUtensorMap tm_a;
  CUtensorMap tm_b;
  CUtensorMap tm_y[4];
  const float* bias;
  int tiles_n, tiles_w, tiles_h;   // tile grid; m tiles = tiles_w * tiles_h * images
  int BW, BH, BN, BK;              // spatial box, N tile, K chunk (channels)
  int taps, ksz, stride, pad;
  int kchunks;                     // ceil(Cin / BK)
  int stages;
  int act; float slope;
  int n_store_maps;                // 1 or 4 (upsample)
  int out_f32;
  int flat;                        // 1: M axis is the flattened pixel list
  int sw_mode;                     // UMMA layout type: 2 = SW128, 4 = SW64, 6 = SW32
  int tmem_cols;
  int bias_len;
  int total_tiles;
  uint32_t idesc;
};

// Persistent kernel: one CTA per SM walks tiles t = blockIdx.x, +gridDim.x, ...  The TMA producer runs
// ahead across tile boundaries (smem ring of `stages` K-chunks), accumulators are double-buffered in
// TMEM so the MMA of tile i+1 overlaps the epilogue of tile i, and the epilogue owns a dedicated
// swizzled staging buffer for the TMA store.
__global__ void __launch_bounds__(kTcThreads, 1) k_conv_tc(const __grid_constant__ ConvTcArgs G) {
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  unsigned char* tiles_ptr = smem_dyn + (tiles - raw);
  const uint32_t a_bytes = 128u * G.BK * 2u;
  const uint32_t b_bytes = (uint32_t)G.BN * G.BK * 2u;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t out_es = G.out_f32 ? 4u : 2u;
  const uint32_t pipe_bytes = stage_bytes * G.stages;
  const uint32_t staging = tiles + pipe_bytes;                       // 1024-aligned (stage_bytes % 1024 == 0)
  const uint32_t staging_bytes = ((uint32_t)G.BN * out_es + 127u) / 128u * 16384u;   // 128 rows x 128 B per group
  unsigned char* aux = tiles_ptr + pipe_bytes + staging_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(aux);  // full[S], empty[S], tmem_full[2], tmem_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux + 8 * (2 * G.stages + 4));
  float* bias_s = reinterpret_cast<float*>(aux + 8 * (2 * G.stages + 4) + 16);
  const uint32_t bar_full0 = smem_u32(bars), bar_empty0 = bar_full0 + 8u * G.stages;
  const uint32_t bar_tfull0 = bar_full0 + 16u * G.stages, bar_tempty0 = bar_tfull0 + 16u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&G.tm_a);
    tma_prefetch_desc(&G.tm_b);
    for (int i = 0; i < G.n_store_maps; ++i) tma_prefetch_desc(&G.tm_y[i]);
    for (int s = 0; s < G.stages; ++s) { mbar_init(bar_full0 + 8u * s, 1); mbar_init(bar_empty0 + 8u * s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull0 + 8u * i, 1); mbar_init(bar_tempty0 + 8u * i, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)G.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < kMaxBias; i += kTcThreads) bias_s[i] = (i < G.bias_len) ? G.bias[i] : 0.f;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const int total_chunks = G.taps * G.kchunks;
  const int tiles_m_per_img = G.tiles_w * G.tiles_h;

  if (warp == 0) {
    // ================================ TMA producer ================================
    // Warp-uniform control flow, one elected lane issues the copies (coordinates live in uniform registers).
    const uint32_t tx = (uint32_t)(G.BW * G.BH) * G.BK * 2u + b_bytes;
    uint32_t s = 0, ph = 1;  // ring slot + parity to wait for on its empty barrier
    for (int t = blockIdx.x; t < G.total_tiles; t += gridDim.x) {
      const int n_tile = t % G.tiles_n;
      int m_tile = t / G.tiles_n;
      const int img = m_tile / tiles_m_per_img;
      m_tile -= img * tiles_m_per_img;
      const int th = m_tile / G.tiles_w, tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * G.BW, h0 = th * G.BH;
      int dy = 0, dx = 0;
      for (int tap = 0; tap < G.taps; ++tap) {
        const int cw = w0 * G.stride + dx - G.pad;
        const int chh = G.flat ? 0 : h0 * G.stride + dy - G.pad;
        for (int kk = 0; kk < G.kchunks; ++kk) {
          mbar_wait(bar_empty0 + 8u * s, ph);
          if (elect_one()) {
            const uint32_t full = bar_full0 + 8u * s;
            mbar_expect_tx(full, tx);
            const uint32_t sa = tiles + s * stage_bytes;
            tma_load_4d(&G.tm_a, full, sa, kk * G.BK, cw, chh, img);
            tma_load_3d(&G.tm_b, full, sa + a_bytes, kk * G.BK, tap, n0);
          }
          __syncwarp();
          if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
        }
        if (++dx == G.ksz) { dx = 0; ++dy; }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    // Whole warp runs the (uniform) control flow; one elected lane issues tcgen05.mma / commit.
    const uint32_t sbo = 8u * (uint32_t)G.BK * 2u;
    const int nk = G.BK / 16;
    const uint32_t hi32 = (uint32_t)(make_smem_desc(0, sbo, (uint32_t)G.sw_mode) >> 32);
    const uint32_t boff = a_bytes >> 4;
    uint32_t s = 0, ph = 0;
    int it = 0;
    for (int t = blockIdx.x; t < G.total_tiles; t += gridDim.x, ++it) {
      const uint32_t acc = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      mbar_wait(bar_tempty0 + 8u * acc, aph ^ 1u);       // epilogue has drained this accumulator
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_d = tmem_base + acc * (uint32_t)G.BN;
      for (int c = 0; c < total_chunks; ++c) {
        mbar_wait(bar_full0 + 8u * s, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t da = ((tiles + s * stage_bytes) >> 4) & 0x3FFF;
        if (elect_one()) {
          const uint32_t acc0 = (uint32_t)(c != 0);
          if (nk == 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16_lh(tmem_d, da + 2u * k, da + boff + 2u * k, hi32, G.idesc, k ? 1u : acc0);
          } else if (nk == 2) {
#pragma unroll
            for (int k = 0; k < 2; ++k) umma_bf16_lh(tmem_d, da + 2u * k, da + boff + 2u * k, hi32, G.idesc, k ? 1u : acc0);
          } else {
            umma_bf16_lh(tmem_d, da, da + boff, hi32, G.idesc, acc0);
          }
          umma_commit(bar_empty0 + 8u * s);   // frees the smem stage once these MMAs retire
        }
        __syncwarp();
        if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
      }
      if (elect_one()) umma_commit(bar_tfull0 + 8u * acc);
      __syncwarp();
    }
  } else {
    // ================================ epilogue ====================================
    const int ew = warp - 2;
    const int q = warp & 3;                 // TMEM lane quarter this warp may read
    const int half = ew >> 2;               // which half of the 32-column chunks this warp owns
    const int row = q * 32 + lane;
    const uint32_t xr = (uint32_t)(row & 7);
    const int nchunks = G.BN / 32;
    const int c_begin = (kEpiWarps == 8) ? (half * ((nchunks + 1) / 2)) : 0;
    const int c_end = (kEpiWarps == 8) ? (half == 0 ? (nchunks + 1) / 2 : nchunks) : nchunks;
    const bool store_thread = (warp == 2 && lane == 0);
    int it = 0;
    for (int t = blockIdx.x; t < G.total_tiles; t += gridDim.x, ++it) {
      const uint32_t acc = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      const int n_tile = t % G.tiles_n;
      int m_tile = t / G.tiles_n;
      const int img = m_tile / tiles_m_per_img;
      m_tile -= img * tiles_m_per_img;
      const int th = m_tile / G.tiles_w, tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * G.BW, h0 = th * G.BH;
      mbar_wait(bar_tfull0 + 8u * acc, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // staging free again
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (uint32_t)G.BN;
      epi_chunks_dispatch(G.act, taddr, c_begin, c_end, bias_s + n0, G.slope, staging, row, xr, G.out_f32 != 0);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> visible to TMA
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
      if (store_thread) {
        mbar_arrive(bar_tempty0 + 8u * acc);   // every epilogue warp has finished reading this accumulator
        const int per_group = G.out_f32 ? 32 : 64;
        const int ngroups = (G.BN + per_group - 1) / per_group;
        for (int mp = 0; mp < G.n_store_maps; ++mp)
          for (int g = 0; g < ngroups; ++g)
            tma_store_4d(&G.tm_y[mp], staging + (uint32_t)g * 16384u, n0 + g * per_group, w0, G.flat ? 0 : h0, img);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (store_thread) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G.tmem_cols)
                 : "memory");
  }
}

struct ConvTcPlan {
  ConvTcArgs args;
  int grid;
  size_t smem;
};

// Tile-N heuristic: estimated cycles = rounds of the persistent grid x per-tile cost, where a tile costs
// max(MMA issue, operand bytes / ~42 B/clk/SM of L2->smem TMA throughput, epilogue) -- see DESIGN.md.
static int pick_tile_n(int cout, long long m_tiles, int chunks, int bk, bool out_f32) {
  if (cout <= 32) return 32;
  const int cands[3] = {64, 128, 256};
  int best = 64;
  double best_cost = 1e30;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (out_f32 && bn > 128) continue;  // fp32 staging tile would not leave room for the pipeline
    const long long tiles = m_tiles * ((cout + bn - 1) / bn);
    const double rounds = (double)((tiles + num_sms() - 1) / num_sms());
    const double mma = chunks * (bk / 16) * (bn / 2.0);
    const double load = chunks * (128.0 * bk * 2 + (double)bn * bk * 2) / 42.0;
    const double epi = bn * 10.0 + 600.0;
    double tile = mma > load ? mma : load;
    if (epi > tile) tile = epi;
    const double cost = rounds * (tile + 400.0);
    if (cost < best_cost * 0.999) { best_cost = cost; best = bn; }
  }
  return best;
}

int conv_tc_plan_create(const yb_conv_desc& d, ConvTcPlan** out) {
  const yb_tensor& x = d.x;
  const yb_tensor& y = d.y;
  YB_CHECK_ARG(x.dtype == YB_BF16, "conv(tc): x must be bf16");
  YB_CHECK_ARG(y.dtype == YB_BF16 || y.dtype == YB_F32, "conv(tc): y must be bf16 or fp32");
  YB_CHECK_ARG(d.res.data == nullptr, "conv(tc): fused residual not supported; use yb_add");
  YB_CHECK_ARG(x.c_stride % 8 == 0 && d.cin_pad % 8 == 0 && (uintptr_t)x.data % 16 == 0 && (uintptr_t)d.weight % 16 == 0,
               "conv(tc): input channel stride / cin_pad must be multiples of 8 and pointers 16-byte aligned");
  const int y_es = y.dtype == YB_F32 ? 4 : 2;
  YB_CHECK_ARG(((size_t)y.c_stride * y_es) % 16 == 0 && (uintptr_t)y.data % 16 == 0,
               "conv(tc): output channel stride must be a multiple of 16 bytes");
  const int k = d.k, s = d.stride, pad = k / 2, ups = d.upsample;
  const int Ho = (x.H + 2 * pad - k) / s + 1, Wo = (x.W + 2 * pad - k) / s + 1;
  const int Cin = x.C, Cout = y.C;
  ConvTcPlan* P = new ConvTcPlan();
  ConvTcArgs& A = P->args;
  memset(&A, 0, sizeof(A));
  A.bias = d.bias; A.bias_len = d.cout_pad;
  A.ksz = k; A.taps = k * k; A.stride = s; A.pad = pad;
  A.act = d.act; A.slope = d.leaky_slope;
  A.out_f32 = y.dtype == YB_F32;
  A.BK = Cin >= 64 ? 64 : (Cin >= 32 ? 32 : 16);
  if (A.BK > d.cin_pad) A.BK = d.cin_pad >= 32 ? 32 : 16;
  YB_CHECK_ARG(d.cin_pad >= 16, "conv(tc): cin_pad must be >= 16");
  A.kchunks = (Cin + A.BK - 1) / A.BK;
  A.sw_mode = A.BK == 64 ? 2 : (A.BK == 32 ? 4 : 6);
  const CUtensorMapSwizzle sw_in = A.BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                                              : (A.BK == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  A.flat = (k == 1 && s == 1 && ups == 1) ? 1 : 0;
  int images = x.N;
  if (A.flat) {
    const long long M = (long long)x.N * x.H * x.W;
    A.BW = 128; A.BH = 1;
    A.tiles_w = (int)((M + 127) / 128); A.tiles_h = 1;
    images = 1;
  } else {
    int best_bw = 1, best_cost = 1 << 30;
    for (int bw = 1; bw <= 128 && bw <= Wo; ++bw) {
      int bh = 128 / bw;
      if (bh > Ho) bh = Ho;
      if (bh * s > 256 || bw * s > 256) continue;
      int cost = ((Wo + bw - 1) / bw) * ((Ho + bh - 1) / bh);
      if (cost < best_cost || (cost == best_cost && bw > best_bw)) { best_cost = cost; best_bw = bw; }
    }
    A.BW = best_bw;
    A.BH = 128 / best_bw < Ho ? 128 / best_bw : Ho;
    A.tiles_w = (Wo + A.BW - 1) / A.BW;
    A.tiles_h = (Ho + A.BH - 1) / A.BH;
  }
  A.BN = d.tile_n ? d.tile_n : pick_tile_n(Cout, (long long)A.tiles_w * A.tiles_h * images, A.taps * A.kchunks, A.BK,
                                          y.dtype == YB_F32);
  YB_CHECK_ARG(A.BN == 32 || A.BN == 64 || A.BN == 128 || A.BN == 256, "conv(tc): tile_n must be 32/64/128/256");
  A.tiles_n = (Cout + A.BN - 1) / A.BN;
  A.n_store_maps = ups == 2 ? 4 : 1;
  A.tmem_cols = 2 * A.BN < 32 ? 32 : 2 * A.BN;   // double-buffered accumulators (64/128/256/512 columns)
  A.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(A.BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  // stages / shared memory
  const size_t stage_bytes = (size_t)128 * A.BK * 2 + (size_t)A.BN * A.BK * 2;
  const size_t stage_out = ((size_t)A.BN * y_es + 127) / 128 * 16384;
  const size_t aux_bytes = 8 * (2 * 12 + 4) + 16 + (size_t)kMaxBias * 4 + 64;
  const size_t budget = 226 * 1024 - 1024 - stage_out - aux_bytes;
  int stages = d.stages ? d.stages : (int)(budget / stage_bytes);
  if (stages > 12) stages = 12;
  if (stages < 2) stages = 2;
  A.stages = stages;
  P->smem = 1024 + stage_bytes * stages + stage_out + aux_bytes;
  YB_CHECK_ARG(d.cout_pad <= kMaxBias, "conv(tc): cout_pad > %d unsupported", kMaxBias);
  if (P->smem > 227 * 1024) {
    set_error("conv(tc): shared memory budget exceeded (%zu bytes)", P->smem);
    delete P;
    return YB_ERR_INVALID;
  }
  int rc;
  const uint32_t one4[4] = {1, 1, 1, 1};
  // ---- A: activations
  {
    uint64_t dims[4], str[3];
    uint32_t box[4], es[4] = {1, (uint32_t)s, (uint32_t)s, 1};
    const uint64_t cs = (uint64_t)x.c_stride * 2;
    if (A.flat) {
      const uint64_t M = (uint64_t)x.N * x.H * x.W;
      dims[0] = Cin; dims[1] = M; dims[2] = 1; dims[3] = 1;
      str[0] = cs; str[1] = cs * M; str[2] = cs * M;
      box[0] = A.BK; box[1] = 128; box[2] = 1; box[3] = 1;
      es[1] = es[2] = 1;
    } else {
      dims[0] = Cin; dims[1] = x.W; dims[2] = x.H; dims[3] = x.N;
      str[0] = cs; str[1] = cs * x.W; str[2] = cs * x.W * x.H;
      box[0] = A.BK; box[1] = A.BW * s; box[2] = A.BH * s; box[3] = 1;
    }
    rc = encode(&A.tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data, dims, str, box, es, sw_in, "activations");
    if (rc) { delete P; return rc; }
  }
  // ---- B: weights [cout_pad][taps][cin_pad]
  {
    uint64_t dims[3] = {(uint64_t)d.cin_pad, (uint64_t)A.taps, (uint64_t)d.cout_pad};
    uint64_t str[2] = {(uint64_t)d.cin_pad * 2, (uint64_t)d.cin_pad * 2 * A.taps};
    uint32_t box[3] = {(uint32_t)A.BK, 1, (uint32_t)A.BN};
    rc = encode(&A.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, d.weight, dims, str, box, one4, sw_in, "weights");
    if (rc) { delete P; return rc; }
  }
  // ---- Y: outputs (swizzle-128B staging, 128-byte inner box)
  {
    const CUtensorMapDataType dt = A.out_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const uint32_t inner = A.out_f32 ? 32 : 64;
    const uint64_t cs = (uint64_t)y.c_stride * y_es;
    for (int mp = 0; mp < A.n_store_maps; ++mp) {
      uint64_t dims[4], str[3];
      uint32_t box[4] = {inner, (uint32_t)A.BW, (uint32_t)A.BH, 1};
      const char* base = (const char*)y.data;
      if (A.flat) {
        const uint64_t M = (uint64_t)x.N * x.H * x.W;
        dims[0] = Cout; dims[1] = M; dims[2] = 1; dims[3] = 1;
        str[0] = cs; str[1] = cs * M; str[2] = cs * M;
        box[1] = 128; box[2] = 1;
      } else if (ups == 2) {
        const int uy = mp >> 1, ux = mp & 1;
        base += ((size_t)uy * y.W + ux) * cs;
        dims[0] = Cout; dims[1] = Wo; dims[2] = Ho; dims[3] = x.N;
        str[0] = 2 * cs; str[1] = 2 * cs * y.W; str[2] = cs * y.W * y.H;
      } else {
        dims[0] = Cout; dims[1] = Wo; dims[2] = Ho; dims[3] = x.N;
        str[0] = cs; str[1] = cs * y.W; str[2] = cs * y.W * y.H;
      }
      rc = encode(&A.tm_y[mp], dt, 4, base, dims, str, box, one4, CU_TENSOR_MAP_SWIZZLE_128B, "output");
      if (rc) { delete P; return rc; }
    }
  }
  YB_CHECK_ARG(y.H == Ho * ups && y.W == Wo * ups && y.N == x.N, "conv(tc): output spatial dims mismatch");
  const long long grid = (long long)A.tiles_n * A.tiles_w * A.tiles_h * images;
  YB_CHECK_ARG(grid < 0x7fffffffLL, "conv(tc): grid too large");
  A.total_tiles = (int)grid;
  P->grid = (int)(grid < num_sms() ? grid : num_sms());
  *out = P;
  return YB_OK;
}

int conv_tc_launch(const ConvTcPlan* P, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_conv_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  k_conv_tc<<<P->grid, kTcThreads, P->smem, s>>>(P->args);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

void conv_tc_plan_destroy(ConvTcPlan* p) { delete p; }

}  // namespace yb

using namespace yb;

extern "C" int yb_conv2d(const yb_conv_desc* d, void* stream) {
  YB_CHECK_ARG(d && view_ok(d->x) && view_ok(d->y) && d->weight && d->bias, "yb_conv2d: bad descriptor");
  YB_CHECK_ARG((d->k == 1 || d->k == 3) && (d->stride == 1 || d->stride == 2), "yb_conv2d: k in {1,3}, stride in {1,2}");
  YB_CHECK_ARG(d->upsample == 1 || d->upsample == 2, "yb_conv2d: upsample must be 1 or 2");
  YB_CHECK_ARG(d->cin_pad >= d->x.C && d->cout_pad >= d->y.C, "yb_conv2d: padded weight extents too small");
  const int pad = d->k / 2;
  const int Ho = (d->x.H + 2 * pad - d->k) / d->stride + 1, Wo = (d->x.W + 2 * pad - d->k) / d->stride + 1;
  YB_CHECK_ARG(d->y.H == Ho * d->upsample && d->y.W == Wo * d->upsample && d->y.N == d->x.N,
               "yb_conv2d: output dims must be (%d,%d)", Ho * d->upsample, Wo * d->upsample);
  if (d->x.dtype == YB_F32) return conv_fp32_launch(*d, (cudaStream_t)stream);
  yb_conv_desc dd = *d;
  if (dd.tile_n >= 0) {   // negative tile_n forces the generic kernel (tests / tuning)
    Conv3Plan* p3 = nullptr;
    int rc3 = conv3_tc_plan_create(dd, &p3);
    if (rc3 == YB_OK) {
      rc3 = conv3_tc_launch(p3, (cudaStream_t)stream);
      conv3_tc_plan_destroy(p3);
      return rc3;
    }
    if (rc3 != YB_ERR_UNSUPPORTED) return rc3;
  } else {
    dd.tile_n = dd.tile_n == -1 ? 0 : -dd.tile_n;
  }
  ConvTcPlan* p = nullptr;
  int rc = conv_tc_plan_create(dd, &p);
  if (rc) return rc;
  rc = conv_tc_launch(p, (cudaStream_t)stream);
  conv_tc_plan_destroy(p);
  return rc;
}
