// 3x3 stride-1 convolution on tcgen05 with operand reuse in shared memory ("slab" kernel).
//
// The generic kernel (conv_tc.cu) loads one 128-pixel activation tile per filter tap: 9 loads of the
// same pixels, and on B200 the 3x3 layers of yolov7 are bound by L2->smem TMA throughput, not by the
// tensor pipe (profiles/round1_*).  This kernel cuts the operand bytes per MAC ~2.2x:
//   * a CTA owns a SUPER-tile of 2 stacked 128-pixel sub-tiles (BW x 2*BH output pixels, BW % 8 == 0);
//   * per K chunk and per filter COLUMN dx it loads ONE activation slab: box {BK ch, BW, 2*BH+2 rows}
//     at (w0+dx-1, h0-1) -- hardware zero-fill gives the padding.  Because BW is a multiple of the
//     8-row swizzle atom, the operand of filter row dy for sub-tile j is the same slab advanced by
//     (j*BH + dy)*BW rows: only the UMMA descriptor start address changes, no data moves;
//   * the three weight tiles (dy = 0..2) of that column are loaded once and feed both sub-tiles.
// Bytes per (dx, chunk) stage: A (2BH+2)*BW*128 B (36 KB at 16x8) + B 3*BN*128 B, for 2*3*BK/16 MMAs.
// TMEM: 2 sub-accumulators x 2 (double buffering) x BN columns (BN <= 128).
// Same epilogue as conv_tc.cu (bias + SiLU/LeakyReLU -> bf16 -> swizzled staging -> TMA store into the
// concat slice).  Reference op: Conv.fuseforward / fused RepConv, models/common.py:110-111, 498-500.
#include "tc_common.cuh"

namespace yb {

struct Conv3Args {
  CUtensorMap tm_a;
  CUtensorMap tm_b;
  CUtensorMap tm_y;
  const float* bias;
  int tiles_n, tiles_w, tiles_h;   // super-tile grid per image
  FastDiv fd_n, fd_mimg, fd_w;     // divisions by tiles_n, tiles_w * tiles_h, tiles_w (tc_common.cuh)
  int BW, BH, BN, BK;
  int nsub;                        // sub-tiles per CTA tile: 2 (super-tile, weights shared) or 1 (grids short of CTAs)
  int kchunks, stages;
  int act; float slope;
  int sw_mode, tmem_cols, bias_len, total_tiles;   // total_tiles = work units: tiles, or (pair of M tiles, N tile)
  int pair;                        // 1: CTA pairs (cta_group::2): M = 256 per MMA, each CTA holds half of every weight tile
  uint32_t a_bytes;                // slab bytes, rounded up to 1024
  uint32_t a_tx;                   // bytes the slab TMA actually delivers
  unsigned long long* dbg;         // tuning aid: globaltimer stamps, 16 per CTA (tc_common.cuh YB_CTS)
  uint32_t idesc;
};

// NSUB * 3 * NK MMAs of one pipeline stage: NSUB sub-tiles x 3 filter rows x NK 16-wide K steps.
template <int NK, int NSUB, bool PAIR>
__device__ __forceinline__ void issue_stage(uint32_t tmem_d0, uint32_t bn, uint32_t dlo, uint32_t hi32,
                                            const uint32_t (&aoff)[2][3], const uint32_t (&boff)[3],
                                            uint32_t idesc, uint32_t acc0) {
#pragma unroll
  for (int j = 0; j < NSUB; ++j) {
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        if (PAIR) umma_bf16_lh_2sm(tmem_d0 + (uint32_t)j * bn, dlo + aoff[j][dy] + (uint32_t)(2 * k),
                                   dlo + boff[dy] + (uint32_t)(2 * k), hi32, idesc, (dy | k) ? 1u : acc0);
        else umma_bf16_lh(tmem_d0 + (uint32_t)j * bn, dlo + aoff[j][dy] + (uint32_t)(2 * k),
                          dlo + boff[dy] + (uint32_t)(2 * k), hi32, idesc, (dy | k) ? 1u : acc0);
      }
    }
  }
}

// PAIR = true (clusters of two CTAs): the pair takes two consecutive M tiles of one N tile; every MMA is one
// tcgen05.mma.cta_group::2 of M = 256 whose B operand is split between the two CTAs (each loads and reads only BN / 2
// rows of every weight tile) -- for N <= 64, where the single-CTA MMA is bound by shared-memory operand reads
// (A 4 KB + B 2 KB per 32 tensor cycles), that takes the B share off each SM.  Protocol: tc_common.cuh "CTA pairs".
template <bool PAIR>
__global__ void __launch_bounds__(kTcThreads, 1) k_conv3_tc(const __grid_constant__ Conv3Args G) {
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = rank == 0u;
  YB_CTS(0, threadIdx.x == 0);
  if (warp == 2) warm_param_lines(G);     // tc_common.cuh: parameter lines are cold per launch
  const uint32_t bn_cta = PAIR ? (uint32_t)G.BN / 2u : (uint32_t)G.BN;
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  unsigned char* tiles_ptr = smem_dyn + (tiles - raw);
  const uint32_t row_bytes = (uint32_t)G.BK * 2u;
  const uint32_t b_bytes = bn_cta * row_bytes;
  const uint32_t stage_bytes = G.a_bytes + 3u * b_bytes;
  const uint32_t pipe_bytes = stage_bytes * G.stages;
  const uint32_t staging = tiles + pipe_bytes;
  const uint32_t staging_bytes = (uint32_t)((G.BN + 63) / 64) * 16384u;   // 128 rows x 128 B per 64-channel group
  unsigned char* aux = tiles_ptr + pipe_bytes + staging_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(aux);  // full[S], empty[S], tmem_full[2], tmem_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux + 8 * (2 * G.stages + 4));
  float* bias_s = reinterpret_cast<float*>(aux + 8 * (2 * G.stages + 4) + 16);
  const uint32_t bar_full0 = smem_u32(bars), bar_empty0 = bar_full0 + 8u * G.stages;
  const uint32_t bar_tfull0 = bar_full0 + 16u * G.stages, bar_tempty0 = bar_tfull0 + 16u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&G.tm_a);
    tma_prefetch_desc(&G.tm_b);
    tma_prefetch_desc(&G.tm_y);
    for (int s = 0; s < G.stages; ++s) { mbar_init(bar_full0 + 8u * s, 1); mbar_init(bar_empty0 + 8u * s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull0 + 8u * i, 1); mbar_init(bar_tempty0 + 8u * i, PAIR ? 2 : 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"((uint32_t)G.tmem_cols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"((uint32_t)G.tmem_cols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  {   // bias of the channels this launch can touch (the last N tile may reach past Cout: zeros there)
    const int nb = min(kMaxBias, (G.bias_len + 255) & ~255);
    const float bsc = epi_bias_scale(G.act);
    for (int i = threadIdx.x; i < nb; i += kTcThreads) bias_s[i] = (i < G.bias_len) ? bsc * G.bias[i] : 0.f;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (PAIR) cluster_sync_all();     // the peer's barriers must exist before anything signals them
  else __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const int walker = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int walkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  pdl_launch_dependents();   // set-up done: the next kernel's CTAs may take SMs as ours retire
  // griddepcontrol.wait: the producer warp alone, after the arithmetic of its first tile (see k_conv_tc)
  const int tiles_m_per_img = G.tiles_w * G.tiles_h;
  const int stages_per_tile = 3 * G.kchunks;

  if (warp == 0) {
    // ================================ TMA producer ================================
    // Warp-uniform control flow, one elected lane issues the copies (coordinates live in uniform registers).
    const uint32_t tx = (PAIR ? 2u : 1u) * (G.a_tx + 3u * b_bytes);   // the leader's barrier collects both CTAs' bytes
    uint32_t s = 0, ph = 1;  // ring slot + parity to wait for on its empty barrier
    bool first = true;
    for (int t = walker; t < G.total_tiles; t += walkers) {
      const int t_m = fd_div(t, G.fd_n);
      const int n_tile = t - t_m * G.tiles_n;
      int m_tile = PAIR ? 2 * t_m + (int)rank : t_m;   // a tile past the end loads zeros and stores nothing (TMA clips)
      const int img = fd_div(m_tile, G.fd_mimg);
      m_tile -= img * tiles_m_per_img;
      const int th = fd_div(m_tile, G.fd_w), tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * G.BW, h0 = th * G.nsub * G.BH;
      if (first) {
        asm volatile("" ::"r"(n0), "r"(w0), "r"(h0), "r"(img));     // coordinates first, then the wait
        pdl_wait();                // activations written by the previous kernel are visible from here on
        YB_CTS(2, lane == 0);
        first = false;
      }
      for (int kk = 0; kk < G.kchunks; ++kk) {
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          mbar_wait(bar_empty0 + 8u * s, ph);
          if (elect_one()) {
            const uint32_t full = bar_full0 + 8u * s;
            if (leader) mbar_expect_tx(full, tx);
            const uint32_t sa = tiles + s * stage_bytes;
            if (PAIR) {
              tma_load_4d_2sm(&G.tm_a, full, sa, kk * G.BK, w0 + dx - 1, h0 - 1, img);
#pragma unroll
              for (int dy = 0; dy < 3; ++dy)
                tma_load_3d_2sm(&G.tm_b, full, sa + G.a_bytes + (uint32_t)dy * b_bytes, kk * G.BK, dy * 3 + dx,
                                n0 + (int)(rank * bn_cta));
            } else {
              tma_load_4d(&G.tm_a, full, sa, kk * G.BK, w0 + dx - 1, h0 - 1, img);
#pragma unroll
              for (int dy = 0; dy < 3; ++dy)
                tma_load_3d(&G.tm_b, full, sa + G.a_bytes + (uint32_t)dy * b_bytes, kk * G.BK, dy * 3 + dx, n0);
            }
          }
          __syncwarp();
          if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (pair mode: the leader CTA only) ==================================
    // Whole warp runs the (uniform) control flow; one elected lane issues tcgen05.mma / commit.
    if (leader) {
    const uint32_t sbo = 8u * row_bytes;
    const int nk = G.BK / 16;
    const uint32_t sub_rows = (uint32_t)(G.BH * G.BW);   // rows between the two sub-tiles
    const uint32_t hi32 = (uint32_t)(make_smem_desc(0, sbo, (uint32_t)G.sw_mode) >> 32);
    uint32_t aoff[2][3], boff[3];                        // descriptor address-field offsets (bytes >> 4)
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
      boff[dy] = (G.a_bytes + (uint32_t)dy * b_bytes) >> 4;
#pragma unroll
      for (int j = 0; j < 2; ++j) aoff[j][dy] = (((uint32_t)j * sub_rows + (uint32_t)(dy * G.BW)) * row_bytes) >> 4;
    }
    uint32_t s = 0, ph = 0;
    int it = 0;
    for (int t = walker; t < G.total_tiles; t += walkers, ++it) {
      const uint32_t par = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      if (PAIR) mbar_wait_cluster(bar_tempty0 + 8u * par, aph ^ 1u);   // both CTAs' epilogues have drained it
      else mbar_wait(bar_tempty0 + 8u * par, aph ^ 1u);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_d0 = tmem_base + (par * 2u) * (uint32_t)G.BN;
      uint32_t idesc = G.idesc;      // last N tile of Cout = 192, 320, ...: only the columns that exist (see k_conv_tc)
      if (!PAIR) {
        const int n_tile_m = t - fd_div(t, G.fd_n) * G.tiles_n;
        const int n_valid = min(G.BN, (G.bias_len - n_tile_m * G.BN + 15) & ~15);
        idesc = (idesc & ~(0x3Fu << 17)) | ((uint32_t)(n_valid >> 3) << 17);
      }
      for (int c = 0; c < stages_per_tile; ++c) {
        mbar_wait(bar_full0 + 8u * s, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t dlo = ((tiles + s * stage_bytes) >> 4) & 0x3FFF;
        if (elect_one()) {
          const uint32_t acc0 = (uint32_t)(c != 0);
          if (G.nsub == 2) {
            if (nk == 4) issue_stage<4, 2, PAIR>(tmem_d0, (uint32_t)G.BN, dlo, hi32, aoff, boff, idesc, acc0);
            else         issue_stage<2, 2, PAIR>(tmem_d0, (uint32_t)G.BN, dlo, hi32, aoff, boff, idesc, acc0);
          } else {
            if (nk == 4) issue_stage<4, 1, PAIR>(tmem_d0, (uint32_t)G.BN, dlo, hi32, aoff, boff, idesc, acc0);
            else         issue_stage<2, 1, PAIR>(tmem_d0, (uint32_t)G.BN, dlo, hi32, aoff, boff, idesc, acc0);
          }
          if (PAIR) umma_commit_2sm(bar_empty0 + 8u * s); else umma_commit(bar_empty0 + 8u * s);
        }
        __syncwarp();
        if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
      }
      if (elect_one()) { if (PAIR) umma_commit_2sm(bar_tfull0 + 8u * par); else umma_commit(bar_tfull0 + 8u * par); }
      __syncwarp();
    }
    }
  } else {
    // ================================ epilogue ====================================
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int row = q * 32 + lane;
    const uint32_t xr = (uint32_t)(row & 7);
    const int nchunks = G.BN / 32;
    const bool store_thread = (warp == 2 && lane == 0);
    int it = 0;
    for (int t = walker; t < G.total_tiles; t += walkers, ++it) {
      const uint32_t par = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      const int t_m = fd_div(t, G.fd_n);
      const int n_tile = t - t_m * G.tiles_n;
      int m_tile = PAIR ? 2 * t_m + (int)rank : t_m;   // a tile past the end loads zeros and stores nothing (TMA clips)
      const int img = fd_div(m_tile, G.fd_mimg);
      m_tile -= img * tiles_m_per_img;
      const int th = fd_div(m_tile, G.fd_w), tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * G.BW, h0 = th * G.nsub * G.BH;
      // only the 32-column chunks that hold real output channels (the last N tile of Cout = 192, 320, ...)
      const int valid_cols = min(G.BN, G.bias_len - n0);
      const int nch = min(nchunks, (valid_cols + 31) >> 5);
      const int c_begin = half * ((nch + 1) / 2);
      const int c_end = half == 0 ? (nch + 1) / 2 : nch;
      mbar_wait(bar_tfull0 + 8u * par, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      for (int j = 0; j < G.nsub; ++j) {
        if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // staging free again
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (par * 2u + (uint32_t)j) * (uint32_t)G.BN;
        epi_chunks_dispatch(G.act, taddr, c_begin, c_end, bias_s + n0, G.slope, staging, row, xr, false);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
        if (store_thread) {
          if (j == G.nsub - 1) {   // all sub-accumulators have been read: hand them back to the (leader's) MMA warp
            if (PAIR && !leader) mbar_arrive_remote(bar_tempty0 + 8u * par, 0u);
            else mbar_arrive(bar_tempty0 + 8u * par);
          }
          const int ngroups = (valid_cols + 63) / 64;
          for (int g = 0; g < ngroups; ++g)
            tma_store_4d(&G.tm_y, staging + (uint32_t)g * 16384u, n0 + g * 64, w0, h0 + j * G.BH, img);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // smem may go once it has been read; the writes complete with the grid
  }
  YB_CTS(9, threadIdx.x == 64);     // an epilogue thread, after its last store has left shared memory
  if (PAIR) {
    // neither CTA may leave (or free TMEM) while the other can still signal its barriers or read its operands
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    cluster_sync_all();
    if (warp == 1) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G.tmem_cols)
                   : "memory");
    }
    return;
  }
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G.tmem_cols)
                 : "memory");
  }
}

struct Conv3Plan {
  Conv3Args args;
  int grid;
  size_t smem;
  Conv3wPlan* wide = nullptr;      // set: the layer runs on the wide-slab kernel (conv3w_tc.cu) instead
  ~Conv3Plan() { if (wide) conv3w_plan_destroy(wide); }
};

const char* conv3_tc_kernel_name(const Conv3Plan* p) { return p->wide ? "k_conv3w_tc" : "k_conv3_tc"; }

// Returns YB_ERR_UNSUPPORTED (without touching the error string semantics of real failures) when the shape
// is better served by the generic kernel; the caller then falls back to conv_tc.
int conv3_tc_plan_create(const yb_conv_desc& d, Conv3Plan** out) {
  const yb_tensor& x = d.x;
  const yb_tensor& y = d.y;
  if (d.k != 3 || d.stride != 1 || d.upsample != 1 || x.dtype != YB_BF16 || y.dtype != YB_BF16 || d.res.data || d.pool.data)
    return YB_ERR_UNSUPPORTED;
  const int Cin = x.C, Cout = y.C, H = x.H, W = x.W;
  if (Cin < 32 || d.cin_pad < 32) return YB_ERR_UNSUPPORTED;
  YB_CHECK_ARG(x.c_stride % 8 == 0 && d.cin_pad % 8 == 0 && (uintptr_t)x.data % 16 == 0 && (uintptr_t)d.weight % 16 == 0 &&
                   y.c_stride % 8 == 0 && (uintptr_t)y.data % 16 == 0, "conv3(tc): alignment");
  YB_CHECK_ARG(d.cout_pad <= kMaxBias, "conv3(tc): cout_pad > %d unsupported", kMaxBias);
  Conv3Plan* P = new Conv3Plan();
  Conv3Args& A = P->args;
  memset(&A, 0, sizeof(A));
  {   // The wide-slab kernel (conv3w_tc.cu: one activation slab per K chunk instead of three) is OPT-IN: YB_C3_WIDE = 1
      // takes every shape it can run, 2 only maps it tiles without waste.  Measured (profiles/README.md, round 2): 4 %
      // faster on the isolated 64 -> 64 layers, 7-9 % slower where this kernel runs as CTA pairs (Cin >= 128), and in
      // the network 0.7 % (rule 2) to 2.6 % SLOWER at batch 32 -- so the default stays off.
    const char* wide_env = getenv("YB_C3_WIDE");          // read per plan (tests switch it between plans)
    const int wm = wide_env ? atoi(wide_env) : 0;
    if (wm != 0) {
      const int rcw = conv3w_plan_create(d, wm == 1 ? 1 : 0, &P->wide);
      if (rcw == YB_OK) { *out = P; return YB_OK; }
      if (rcw != YB_ERR_UNSUPPORTED) { delete P; return rcw; }
    }
  }
  A.bias = d.bias; A.bias_len = d.cout_pad; A.act = d.act; A.slope = d.leaky_slope;
  A.BK = Cin >= 64 ? 64 : 32;
  A.kchunks = (Cin + A.BK - 1) / A.BK;
  A.sw_mode = A.BK == 64 ? 2 : 4;
  const CUtensorMapSwizzle sw_in = A.BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  // ---- tile shape: spatial sub-tile BW x BH (BW multiple of 8, BH = 128 / BW), NSUB stacked sub-tiles per CTA tile,
  // N tile.  The super-tile (NSUB = 2) shares every weight tile between two sub-tiles, N = 128 halves the activation
  // traffic of N = 64 -- but at small maps / small batches such tiles are fewer than the SMs (512 -> 512 at 20 x 20,
  // batch 8: 52 CTAs of 19 us each), so the shape is chosen by an estimate of the whole launch:
  //   rounds of the persistent grid x stages x max(MMA cycles, operand bytes / 42 B/clk of L2 -> smem) + fill / drain.
  const int bw_step = A.BK == 64 ? 8 : 16;   // sub-tile / filter-row offsets stay 1024-byte aligned
  const uint32_t row_bytes = A.BK * 2;
  static const char* nsub_env = getenv("YB_C3_NSUB");       // tuning aids
  const int nsub_force = nsub_env ? atoi(nsub_env) : 0;
  double best_time = 1e30;
  for (int nsub = 2; nsub >= 1; --nsub) {
    if (nsub_force && nsub != nsub_force) continue;
    long long best_cost = -1;
    int cbw = 0, cbh = 0;
    for (int bw = bw_step; bw <= 128; bw += bw_step) {
      const int bh = 128 / bw;
      if (bh < 1 || nsub * bh + 2 > 256) continue;
      const long long nt = (long long)((W + bw - 1) / bw) * ((H + nsub * bh - 1) / (nsub * bh));
      const long long cost = nt * ((nsub * bh + 2) * bw + 96);   // slab rows (+ weight rows equivalent) per tile
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; cbw = bw; cbh = bh; }
    }
    const long long m_tiles = (long long)((W + cbw - 1) / cbw) * ((H + nsub * cbh - 1) / (nsub * cbh)) * x.N;
    const int bn_cands[3] = {128, 64, 32};
    for (int bi = 0; bi < 3; ++bi) {
      const int bn = bn_cands[bi];
      if (d.tile_n ? bn != d.tile_n : ((bn == 32) != (Cout <= 32) || (bn == 128 && Cout <= 64))) continue;
      const int tiles_n = (Cout + bn - 1) / bn;
      const long long tiles = m_tiles * tiles_n;
      const long long grid = tiles < num_sms() ? tiles : num_sms();
      const double rounds = (double)((tiles + grid - 1) / grid);
      const double mma = nsub * 3 * (A.BK / 16) * (bn >= 128 ? 64.0 : bn == 64 ? 48.0 : 40.0);   // N <= 64: smem-operand bound
      const double load = ((double)(nsub * cbh + 2) * cbw + 3.0 * bn) * row_bytes / 42.0;
      const double t = rounds * 3 * A.kchunks * (mma > load ? mma : load) + nsub * (bn * 10.0 + 600.0) + 3000.0;
      if (t < best_time * 0.97) {   // prefer the larger shape unless the smaller one clearly wins
        best_time = t; A.nsub = nsub; A.BW = cbw; A.BH = cbh; A.BN = bn;
      }
    }
  }
  YB_CHECK_ARG(A.nsub && (A.BN == 32 || A.BN == 64 || A.BN == 128), "conv3(tc): tile_n must be 32/64/128");
  A.tiles_w = (W + A.BW - 1) / A.BW;
  A.tiles_h = (H + A.nsub * A.BH - 1) / (A.nsub * A.BH);
  A.tiles_n = (Cout + A.BN - 1) / A.BN;
  A.fd_n = make_fastdiv(A.tiles_n); A.fd_mimg = make_fastdiv(A.tiles_w * A.tiles_h); A.fd_w = make_fastdiv(A.tiles_w);
  // CTA pairs, by measurement (scripts/pair3_sweep_specs.txt, profiles/round2_pair3_sweep.txt, batch 32): 0.81-0.95x where
  // the K loop is long (Cin >= 128: the weight tiles are half of a stage's L2 -> SM bytes and the pair halves them),
  // 1.12-1.21x (slower) for the 64 -> 64 layers, whose three stages per tile do not amortise the cross-CTA handshakes --
  // the smem-operand bound of N = 64 is NOT what pairs fix.  d.pair / YB_C3_PAIR force (1) or forbid (-1).
  {
    static const char* pair_env = getenv("YB_C3_PAIR");
    int pm = d.pair ? d.pair : (pair_env ? atoi(pair_env) : 0);
    const long long tiles_all = (long long)A.tiles_w * A.tiles_h * x.N * A.tiles_n;
    if (pm == 0) pm = (Cin >= 128 && tiles_all >= num_sms()) ? 1 : -1;
    A.pair = pm > 0 ? 1 : 0;
  }
  const int bn_cta = A.pair ? A.BN / 2 : A.BN;
  A.tmem_cols = 4 * A.BN < 32 ? 32 : 4 * A.BN;
  A.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(A.BN >> 3) << 17) | ((uint32_t)((A.pair ? 256 : 128) >> 4) << 24);
  A.a_tx = (uint32_t)(A.nsub * A.BH + 2) * A.BW * row_bytes;
  // the last sub-tile's MMA (filter row 2) reads 128 rows from row ((NSUB-1)*BH + 2)*BW: keep those reads inside the stage
  uint32_t need_rows = (uint32_t)(((A.nsub - 1) * A.BH + 2) * A.BW + 128);
  uint32_t slab_rows = (uint32_t)(A.nsub * A.BH + 2) * A.BW;
  if (need_rows > slab_rows) slab_rows = need_rows;
  A.a_bytes = (uint32_t)align_up((size_t)slab_rows * row_bytes, 1024);
  const size_t stage_bytes = (size_t)A.a_bytes + 3 * (size_t)bn_cta * row_bytes;
  const size_t stage_out = (size_t)((A.BN + 63) / 64) * 16384;
  const size_t aux_bytes = 8 * (2 * 12 + 4) + 16 + (size_t)kMaxBias * 4 + 64;
  const size_t budget = 226 * 1024 - 1024 - stage_out - aux_bytes;
  int stages = d.stages ? d.stages : (int)(budget / stage_bytes);
  if (stages > 12) stages = 12;
  if (stages < 2) { delete P; return YB_ERR_UNSUPPORTED; }
  A.stages = stages;
  P->smem = 1024 + stage_bytes * stages + stage_out + aux_bytes;
  const uint32_t one4[4] = {1, 1, 1, 1};
  int rc;
  {
    const uint64_t cs = (uint64_t)x.c_stride * 2;
    uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)x.N};
    uint64_t str[3] = {cs, cs * W, cs * W * H};
    uint32_t box[4] = {(uint32_t)A.BK, (uint32_t)A.BW, (uint32_t)(A.nsub * A.BH + 2), 1};
    rc = encode(&A.tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data, dims, str, box, one4, sw_in, "conv3 activations");
    if (rc) { delete P; return rc; }
  }
  {
    uint64_t dims[3] = {(uint64_t)d.cin_pad, 9, (uint64_t)d.cout_pad};
    uint64_t str[2] = {(uint64_t)d.cin_pad * 2, (uint64_t)d.cin_pad * 2 * 9};
    uint32_t box[3] = {(uint32_t)A.BK, 1, (uint32_t)bn_cta};     // pair mode: each CTA loads half of the N tile
    rc = encode(&A.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, d.weight, dims, str, box, one4, sw_in, "conv3 weights");
    if (rc) { delete P; return rc; }
  }
  {
    const uint64_t cs = (uint64_t)y.c_stride * 2;
    uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)W, (uint64_t)H, (uint64_t)x.N};
    uint64_t str[3] = {cs, cs * W, cs * W * H};
    uint32_t box[4] = {64, (uint32_t)A.BW, (uint32_t)A.BH, 1};
    rc = encode(&A.tm_y, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, y.data, dims, str, box, one4, CU_TENSOR_MAP_SWIZZLE_128B,
                "conv3 output");
    if (rc) { delete P; return rc; }
  }
  const long long m_total = (long long)A.tiles_w * A.tiles_h * x.N;
  const long long units = (long long)A.tiles_n * (A.pair ? (m_total + 1) / 2 : m_total);
  A.total_tiles = (int)units;
  const int walkers_max = A.pair ? num_sms() / 2 : num_sms();
  const int walkers = (int)(units < walkers_max ? units : walkers_max);
  P->grid = A.pair ? 2 * walkers : walkers;
  A.dbg = conv_dbg_slot();
  *out = P;
  return YB_OK;
}

int conv3_tc_launch(const Conv3Plan* P, cudaStream_t s) {
  if (P->wide) return conv3w_launch(P->wide, s);
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_conv3_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    YB_CUDA(cudaFuncSetAttribute(k_conv3_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  if (P->args.pair) YB_CUDA(launch_pdl_pairs(k_conv3_tc<true>, P->grid, kTcThreads, P->smem, s, P->args));
  else YB_CUDA(launch_pdl(k_conv3_tc<false>, P->grid, kTcThreads, P->smem, s, P->args));
  return YB_OK;
}

void conv3_tc_plan_destroy(Conv3Plan* p) { delete p; }

}  // namespace yb
