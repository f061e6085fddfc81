// 3x3 stride-1 convolution on tcgen05, ONE activation slab per K chunk ("wide slab" kernel, round 2).
//
// OPT-IN variant (YB_C3_WIDE=1 | 2, see conv3_tc_plan_create): correct and bit-identical to k_conv3_tc, but not faster
// in the network -- kept for the operand-addressing scheme below, which removes two thirds of the slab traffic.
//
// k_conv3_tc (conv3_tc.cu) loads one slab per (K chunk, filter COLUMN): inside its BW x BH tiles a shift by one pixel
// would wrap from the end of one image row into the next, so the three columns are three TMA loads of nearly the same
// pixels.  A probe build that loaded only the dx = 0 slab (wrong results, same MMAs) ran the batch-32 step 7.5 % faster,
// the isolated layers 1-8 %.  This kernel delivers the 4 % on the isolated 64 -> 64 layers; its fixed 8 x 16 sub-tiles
// waste rows on maps whose height is not a multiple of 16 / 32, it has no CTA-pair mode (k_conv3_tc with pairs is at
// 1.5-1.65 PFLOP/s for Cin >= 128), and in the network it measured 0.7-2.6 % slower (profiles/README.md).
//
// Here a sub-tile is 8 pixels wide and 16 rows high, and the slab keeps the halo INSIDE its rows: box {BK channels,
// 10 pixels, NSUB*16 + 2 rows} at (w0 - 1, h0 - 1), rows of 10 x 128 B.  The 128 rows of an MMA are 16 groups of 8
// consecutive pixels -- exactly what a K-major SWIZZLE_128B operand is made of (8-row core groups) -- and the stride
// between groups (SBO) is the slab pitch, 10 x 128 B = 1280 B, instead of 1024 B: group g is output row g of the
// sub-tile, and a filter tap (dy, dx) is the descriptor start address advanced by (dy * 10 + dx) x 128 B.  No junk rows,
// no second copy of any pixel; the swizzle is a function of the absolute shared-memory address both for the TMA write
// and for the UMMA read, so a start address inside a 1024-byte atom reads what TMA put there (the property
// k_conv3s2p_tc relies on for its +128 B tap).
//
// Operand rings: slabs (2 x 44 KB at NSUB = 2) and weight stages (3 filter rows of one column = 3 tiles of BN x 64
// channels) are separate, because a slab lives for three weight stages.  Per K chunk: 1 slab + 9 weight tiles for
// NSUB * 9 * 4 MMAs.  Epilogue, TMEM layout (NSUB sub-accumulators, double buffered) and store as in k_conv3_tc.
// Reference op: Conv.fuseforward / fused RepConv, models/common.py:110-111, 498-500.
#include "tc_common.cuh"

namespace yb {

constexpr int kWsBW = 8, kWsBH = 16, kWsPitch = kWsBW + 2;   // sub-tile 8 x 16 pixels, slab rows of 10 pixels
constexpr int kWsSlabs = 2;

struct Conv3wArgs {
  CUtensorMap tm_a;
  CUtensorMap tm_b;
  CUtensorMap tm_y;
  const float* bias;
  int tiles_n, tiles_w, tiles_h;   // tile grid per image (tile = 8 x NSUB*16 pixels)
  FastDiv fd_n, fd_mimg, fd_w;
  int BN, nsub, kchunks, wstages;
  int act; float slope;
  int tmem_cols, bias_len, total_tiles;
  uint32_t a_bytes, a_tx;          // slab bytes rounded up to 1024 / bytes TMA delivers
  uint32_t idesc;
  unsigned long long* dbg;
};

template <int NSUB>
__device__ __forceinline__ void issue_wide_stage(uint32_t tmem_d0, uint32_t bn, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                                 uint32_t b_hi, uint32_t bstep, uint32_t idesc, uint32_t acc0) {
  // NSUB sub-tiles x 3 filter rows x 4 K steps of 16 channels; rows of 128 B => one K step = 32 B = 2 address units
#pragma unroll
  for (int j = 0; j < NSUB; ++j) {
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
      const uint32_t a = a_lo + (uint32_t)(((j * kWsBH + dy) * kWsPitch * 128) >> 4);
      const uint32_t b = b_lo + (uint32_t)dy * bstep;
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_bf16_lh2(tmem_d0 + (uint32_t)j * bn, a + (uint32_t)(2 * k), a_hi, b + (uint32_t)(2 * k), b_hi, idesc,
                      (dy | k) ? 1u : acc0);
    }
  }
}

__global__ void __launch_bounds__(kTcThreads, 1) k_conv3w_tc(const __grid_constant__ Conv3wArgs G) {
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  YB_CTS(0, threadIdx.x == 0);
  if (warp == 2) warm_param_lines(G);
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  unsigned char* tiles_ptr = smem_dyn + (tiles - raw);
  const uint32_t b_bytes = (uint32_t)G.BN * 128u;                 // one weight tile: BN rows of 64 channels
  const uint32_t wstage_bytes = 3u * b_bytes;
  const uint32_t slab0 = tiles;
  const uint32_t wring = tiles + (uint32_t)kWsSlabs * G.a_bytes;
  const uint32_t pipe_bytes = (uint32_t)kWsSlabs * G.a_bytes + wstage_bytes * (uint32_t)G.wstages;
  const uint32_t staging = tiles + pipe_bytes;
  const uint32_t staging_bytes = (uint32_t)((G.BN + 63) / 64) * 16384u;
  unsigned char* aux = tiles_ptr + pipe_bytes + staging_bytes;
  // barriers: slab_full[2], slab_empty[2], w_full[S], w_empty[S], tmem_full[2], tmem_empty[2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(aux);
  const int nbars = 2 * kWsSlabs + 2 * G.wstages + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux + 8 * nbars);
  float* bias_s = reinterpret_cast<float*>(aux + 8 * nbars + 16);
  const uint32_t bar_sfull0 = smem_u32(bars), bar_sempty0 = bar_sfull0 + 8u * kWsSlabs;
  const uint32_t bar_wfull0 = bar_sempty0 + 8u * kWsSlabs, bar_wempty0 = bar_wfull0 + 8u * G.wstages;
  const uint32_t bar_tfull0 = bar_wempty0 + 8u * G.wstages, bar_tempty0 = bar_tfull0 + 16u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&G.tm_a);
    tma_prefetch_desc(&G.tm_b);
    tma_prefetch_desc(&G.tm_y);
    for (int s = 0; s < kWsSlabs; ++s) { mbar_init(bar_sfull0 + 8u * s, 1); mbar_init(bar_sempty0 + 8u * s, 1); }
    for (int s = 0; s < G.wstages; ++s) { mbar_init(bar_wfull0 + 8u * s, 1); mbar_init(bar_wempty0 + 8u * s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull0 + 8u * i, 1); mbar_init(bar_tempty0 + 8u * i, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)G.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {
    const int nb = min(kMaxBias, (G.bias_len + 255) & ~255);
    const float bsc = epi_bias_scale(G.act);
    for (int i = threadIdx.x; i < nb; i += kTcThreads) bias_s[i] = (i < G.bias_len) ? bsc * G.bias[i] : 0.f;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const int walker = (int)blockIdx.x, walkers = (int)gridDim.x;
  pdl_launch_dependents();
  const int tiles_m_per_img = G.tiles_w * G.tiles_h;

  if (warp == 0) {
    // ================================ TMA producer ================================
    uint32_t sl = 0, sph = 1, ws = 0, wph = 1;    // ring slots + the parity to wait for on their empty barriers
    bool first = true;
    for (int t = walker; t < G.total_tiles; t += walkers) {
      const int m_all = fd_div(t, G.fd_n);
      const int n_tile = t - m_all * G.tiles_n;
      const int img = fd_div(m_all, G.fd_mimg);
      const int m_tile = m_all - img * tiles_m_per_img;
      const int th = fd_div(m_tile, G.fd_w), tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * kWsBW, h0 = th * G.nsub * kWsBH;
      if (first) {
        asm volatile("" ::"r"(n0), "r"(w0), "r"(h0), "r"(img));     // coordinates first, then the wait
        pdl_wait();                // activations written by the previous kernel are visible from here on
        YB_CTS(2, lane == 0);
        first = false;
      }
      for (int kk = 0; kk < G.kchunks; ++kk) {
        mbar_wait(bar_sempty0 + 8u * sl, sph);
        if (elect_one()) {
          const uint32_t full = bar_sfull0 + 8u * sl;
          mbar_expect_tx(full, G.a_tx);
          tma_load_4d(&G.tm_a, full, slab0 + sl * G.a_bytes, kk * 64, w0 - 1, h0 - 1, img);
        }
        __syncwarp();
        if (++sl == (uint32_t)kWsSlabs) { sl = 0; sph ^= 1u; }
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          mbar_wait(bar_wempty0 + 8u * ws, wph);
          if (elect_one()) {
            const uint32_t full = bar_wfull0 + 8u * ws;
            mbar_expect_tx(full, wstage_bytes);
            const uint32_t sb = wring + ws * wstage_bytes;
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) tma_load_3d(&G.tm_b, full, sb + (uint32_t)dy * b_bytes, kk * 64, dy * 3 + dx, n0);
          }
          __syncwarp();
          if (++ws == (uint32_t)G.wstages) { ws = 0; wph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const uint32_t a_hi = (uint32_t)(make_smem_desc(0, (uint32_t)kWsPitch * 128u, 2u) >> 32);   // SBO = slab pitch
    const uint32_t b_hi = (uint32_t)(make_smem_desc(0, 1024u, 2u) >> 32);
    const uint32_t bstep = b_bytes >> 4;
    uint32_t sl = 0, sph = 0, ws = 0, wph = 0;
    int it = 0;
    for (int t = walker; t < G.total_tiles; t += walkers, ++it) {
      const uint32_t par = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      mbar_wait(bar_tempty0 + 8u * par, aph ^ 1u);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_d0 = tmem_base + (par * (uint32_t)G.nsub) * (uint32_t)G.BN;
      for (int kk = 0; kk < G.kchunks; ++kk) {
        mbar_wait(bar_sfull0 + 8u * sl, sph);
        const uint32_t a_base = ((slab0 + sl * G.a_bytes) >> 4) & 0x3FFF;
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          mbar_wait(bar_wfull0 + 8u * ws, wph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t b_base = ((wring + ws * wstage_bytes) >> 4) & 0x3FFF;
          if (elect_one()) {
            const uint32_t acc0 = (uint32_t)((kk | dx) != 0);
            if (G.nsub == 2) issue_wide_stage<2>(tmem_d0, (uint32_t)G.BN, a_base + (uint32_t)(dx * 8), a_hi, b_base, b_hi, bstep, G.idesc, acc0);
            else             issue_wide_stage<1>(tmem_d0, (uint32_t)G.BN, a_base + (uint32_t)(dx * 8), a_hi, b_base, b_hi, bstep, G.idesc, acc0);
            umma_commit(bar_wempty0 + 8u * ws);
            if (dx == 2) umma_commit(bar_sempty0 + 8u * sl);     // the slab has fed its three columns
          }
          __syncwarp();
          if (++ws == (uint32_t)G.wstages) { ws = 0; wph ^= 1u; }
        }
        if (++sl == (uint32_t)kWsSlabs) { sl = 0; sph ^= 1u; }
      }
      if (elect_one()) umma_commit(bar_tfull0 + 8u * par);
      __syncwarp();
    }
  } else {
    // ================================ epilogue (as k_conv3_tc) ====================================
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int row = q * 32 + lane;
    const uint32_t xr = (uint32_t)(row & 7);
    const int nchunks = G.BN / 32;
    const int c_begin = half * ((nchunks + 1) / 2);
    const int c_end = half == 0 ? (nchunks + 1) / 2 : nchunks;
    const bool store_thread = (warp == 2 && lane == 0);
    int it = 0;
    for (int t = walker; t < G.total_tiles; t += walkers, ++it) {
      const uint32_t par = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      const int m_all = fd_div(t, G.fd_n);
      const int n_tile = t - m_all * G.tiles_n;
      const int img = fd_div(m_all, G.fd_mimg);
      const int m_tile = m_all - img * tiles_m_per_img;
      const int th = fd_div(m_tile, G.fd_w), tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * kWsBW, h0 = th * G.nsub * kWsBH;
      mbar_wait(bar_tfull0 + 8u * par, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      for (int j = 0; j < G.nsub; ++j) {
        if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // staging free again
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (par * (uint32_t)G.nsub + (uint32_t)j) * (uint32_t)G.BN;
        epi_chunks_dispatch(G.act, taddr, c_begin, c_end, bias_s + n0, G.slope, staging, row, xr, false);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
        if (store_thread) {
          if (j == G.nsub - 1) mbar_arrive(bar_tempty0 + 8u * par);   // all sub-accumulators have been read
          const int ngroups = (G.BN + 63) / 64;
          for (int g = 0; g < ngroups; ++g)
            tma_store_4d(&G.tm_y, staging + (uint32_t)g * 16384u, n0 + g * 64, w0, h0 + j * kWsBH, img);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  YB_CTS(9, threadIdx.x == 64);
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G.tmem_cols)
                 : "memory");
  }
}

struct Conv3wPlan {
  Conv3wArgs args;
  int grid;
  size_t smem;
};

// Declines (YB_ERR_UNSUPPORTED) what the 16-row / 8-column tiles or the 64-channel chunks do not fit well; the caller then
// takes k_conv3_tc.  force: 1 = take every shape the kernel can run (tests, sweeps), 0 = the measured rule.
int conv3w_plan_create(const yb_conv_desc& d, int force, Conv3wPlan** out) {
  const yb_tensor& x = d.x;
  const yb_tensor& y = d.y;
  if (d.k != 3 || d.stride != 1 || d.upsample != 1 || x.dtype != YB_BF16 || y.dtype != YB_BF16 || d.res.data || d.pool.data ||
      d.pair > 0)
    return YB_ERR_UNSUPPORTED;
  const int Cin = x.C, Cout = y.C, H = x.H, W = x.W;
  if (Cin < 64 || d.cin_pad < 64 || Cout <= 32 || d.cout_pad > kMaxBias) return YB_ERR_UNSUPPORTED;
  if (x.c_stride % 8 || d.cin_pad % 8 || (uintptr_t)x.data % 16 || (uintptr_t)d.weight % 16 || y.c_stride % 8 ||
      (uintptr_t)y.data % 16)
    return YB_ERR_UNSUPPORTED;
  const int bn = d.tile_n > 0 ? d.tile_n : (Cout <= 64 ? 64 : 128);
  if (bn != 64 && bn != 128) return YB_ERR_UNSUPPORTED;
  const int tiles_w = (W + kWsBW - 1) / kWsBW;
  const int tiles_n = (Cout + bn - 1) / bn;
  // sub-tiles per CTA tile: two stacked ones share every weight tile; one when that would leave SMs without work or
  // waste more than a sixth of the rows
  int nsub = 0;
  double best = 1e30;
  const char* ns_env = getenv("YB_C3W_NSUB");      // tuning aid
  for (int ns = 2; ns >= 1; --ns) {
    if (ns_env && atoi(ns_env) != ns) continue;
    const int th = (H + ns * kWsBH - 1) / (ns * kWsBH);
    // rows / columns of work per row / column of image: small maps (20 x 20: 24 x 32) belong to the other kernel
    const double eff = ((double)W / (tiles_w * kWsBW)) * ((double)H / (th * ns * kWsBH));
    if (!force && eff < 0.99) continue;
    const long long tiles = (long long)tiles_w * th * x.N * tiles_n;
    const long long grid = tiles < num_sms() ? tiles : num_sms();
    const double rounds = (double)((tiles + grid - 1) / grid);
    const double mma = ns * 9 * 4 * (bn >= 128 ? 64.0 : 48.0);
    const double load = ((double)(ns * kWsBH + 2) * kWsPitch + 9.0 * bn) * 128.0 / 42.0;
    const double t = rounds * ((Cin + 63) / 64) * (mma > load ? mma : load) + ns * (bn * 10.0 + 600.0) + 3000.0;
    if (t < best * 0.97) { best = t; nsub = ns; }
  }
  if (!nsub) return YB_ERR_UNSUPPORTED;
  const int tiles_h = (H + nsub * kWsBH - 1) / (nsub * kWsBH);
  Conv3wPlan* P = new Conv3wPlan();
  Conv3wArgs& A = P->args;
  memset(&A, 0, sizeof(A));
  A.bias = d.bias; A.bias_len = d.cout_pad; A.act = d.act; A.slope = d.leaky_slope;
  A.BN = bn; A.nsub = nsub; A.kchunks = (Cin + 63) / 64;
  A.tiles_w = tiles_w; A.tiles_h = tiles_h; A.tiles_n = tiles_n;
  A.fd_n = make_fastdiv(tiles_n); A.fd_mimg = make_fastdiv(tiles_w * tiles_h); A.fd_w = make_fastdiv(tiles_w);
  A.tmem_cols = 2 * nsub * bn < 32 ? 32 : 2 * nsub * bn;        // 128 / 256 / 512 columns: a power of two
  A.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint32_t slab_rows = (uint32_t)(nsub * kWsBH + 2) * kWsPitch;
  A.a_tx = slab_rows * 128u;
  // the last tap (dy = dx = 2) of the last sub-tile starts at row ((nsub-1)*16 + 2)*10 + 2; its last group of 8 rows ends
  // 15 * 10 + 8 rows further = exactly at the end of the slab: no operand read leaves the buffer
  A.a_bytes = (uint32_t)align_up((size_t)slab_rows * 128, 1024);
  const size_t wstage = 3 * (size_t)bn * 128;
  const size_t stage_out = (size_t)((bn + 63) / 64) * 16384;
  const size_t bias_bytes = (size_t)((d.cout_pad + 255) & ~255) * 4;
  int ws = d.stages ? d.stages : 8;
  size_t smem = 0;
  for (; ws >= 2; --ws) {
    smem = 1024 + (size_t)kWsSlabs * A.a_bytes + wstage * ws + stage_out + 8 * (2 * kWsSlabs + 2 * ws + 4) + 16 + bias_bytes + 64;
    if (smem <= 227 * 1024) break;
  }
  if (ws < 2) { delete P; return YB_ERR_UNSUPPORTED; }
  A.wstages = ws;
  P->smem = smem;
  const uint32_t one4[4] = {1, 1, 1, 1};
  int rc;
  {
    const uint64_t cs = (uint64_t)x.c_stride * 2;
    uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)x.N};
    uint64_t str[3] = {cs, cs * W, cs * W * H};
    uint32_t box[4] = {64, (uint32_t)kWsPitch, (uint32_t)(nsub * kWsBH + 2), 1};
    rc = encode(&A.tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data, dims, str, box, one4, CU_TENSOR_MAP_SWIZZLE_128B,
                "conv3w activations");
    if (rc) { delete P; return rc; }
  }
  {
    uint64_t dims[3] = {(uint64_t)d.cin_pad, 9, (uint64_t)d.cout_pad};
    uint64_t str[2] = {(uint64_t)d.cin_pad * 2, (uint64_t)d.cin_pad * 2 * 9};
    uint32_t box[3] = {64, 1, (uint32_t)bn};
    rc = encode(&A.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, d.weight, dims, str, box, one4, CU_TENSOR_MAP_SWIZZLE_128B,
                "conv3w weights");
    if (rc) { delete P; return rc; }
  }
  {
    const uint64_t cs = (uint64_t)y.c_stride * 2;
    uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)W, (uint64_t)H, (uint64_t)x.N};
    uint64_t str[3] = {cs, cs * W, cs * W * H};
    uint32_t box[4] = {64, (uint32_t)kWsBW, (uint32_t)kWsBH, 1};
    rc = encode(&A.tm_y, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, y.data, dims, str, box, one4, CU_TENSOR_MAP_SWIZZLE_128B,
                "conv3w output");
    if (rc) { delete P; return rc; }
  }
  const long long units = (long long)tiles_w * tiles_h * x.N * tiles_n;
  if (units >= 0x3fffffffLL) { delete P; return YB_ERR_UNSUPPORTED; }
  A.total_tiles = (int)units;
  P->grid = (int)(units < num_sms() ? units : num_sms());
  A.dbg = conv_dbg_slot();
  *out = P;
  return YB_OK;
}

int conv3w_launch(const Conv3wPlan* P, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_conv3w_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  YB_CUDA(launch_pdl(k_conv3w_tc, P->grid, kTcThreads, P->smem, s, P->args));
  return YB_OK;
}

void conv3w_plan_destroy(Conv3wPlan* p) { delete p; }

}  // namespace yb
