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
#include <type_traits>

#include "tc_common.cuh"

namespace yb {

unsigned long long* conv_dbg_slot() {
  static long long next_slot = 0;
  const char* ts = getenv("YB_CONV_TS");
  if (!ts) return nullptr;
  auto* base = reinterpret_cast<unsigned long long*>(strtoull(ts, nullptr, 0));
  const char* st = getenv("YB_CONV_TS_STRIDE");
  const long long stride = st ? strtoll(st, nullptr, 0) : 0;
  return base + (stride > 0 ? stride * next_slot++ : 0);
}

struct ConvTcArgs {
  CUtensorMap tm_a;
  CUtensorMap tm_b;
  CUtensorMap tm_y[4];
  const float* bias;
  int tiles_n, tiles_w, tiles_h;   // tile grid; m tiles = tiles_w * tiles_h * images
  FastDiv fd_n, fd_mimg, fd_w;     // divisions by tiles_n, tiles_w * tiles_h, tiles_w
  int BW, BH, BN, BK;              // spatial box, N tile, K chunk (channels)
  int taps, kw, stride_h, stride_w, pad_h, pad_w;   // taps = kh * kw, tap = dy * kw + dx
  int epi_ring, epi_cpr;           // output staging ring (in rounds) and 32-column chunks per warp per round
  int resident;                    // 1: all weight tiles of the CTA's N tile stay in shared memory
  int kchunks;                     // ceil(Cin / BK)
  int stages;
  int act; float slope;
  int n_store_maps;                // 1 or 4 (upsample)
  int out_f32;
  int flat;                        // 1: M axis is the flattened pixel list
  int sw_mode;                     // UMMA layout type: 2 = SW128, 4 = SW64, 6 = SW32
  int tmem_cols;
  int bias_len;
  int total_tiles;                 // work units: (m, n) tiles, or (pair of m tiles, n) in pair mode
  int pair;                        // 1: CTA pairs (clusters of 2, tcgen05 cta_group::2, M = 256 per MMA)
  int pool;                        // 1: also write the 2x2 max-pool of every (even-sided, even-origin) output tile
  __nv_bfloat16* pool_ptr;         // pooled output (N, pool_h, pool_w, .) with pool_cs elements per pixel
  int pool_h, pool_w, pool_cs, pool_c, tiles_imgs;
  const __nv_bfloat16* res_ptr;    // residual added after the activation (NULL = none): (N, out_h, out_w, .) bf16
  int res_cs, res_c, out_h, out_w; // elements per residual pixel, its channel count rounded up to 8; output extent
  long long out_pixels;            // N * out_h * out_w (flat mode)
  uint32_t idesc;
  unsigned long long* dbg;         // tuning aid (env YB_CONV_TS = device address): globaltimer stamps, 16 per CTA
};


// Persistent kernel: one CTA per SM walks tiles t = blockIdx.x, +gridDim.x, ...  The TMA producer runs
// ahead across tile boundaries (smem ring of `stages` K-chunks), accumulators are double-buffered in
// TMEM so the MMA of tile i+1 overlaps the epilogue of tile i, and the epilogue owns a dedicated
// swizzled staging buffer for the TMA store.
//
// PAIR = true (launched as clusters of two CTAs): the two CTAs of a pair take two consecutive M tiles of the same N
// tile and run them as ONE M = 256 tcgen05.mma.cta_group::2 -- each CTA loads its own A tile and only HALF of the
// weight tile (rows n0 + rank * BN/2 ...), so the weight bytes through L2 -> SM and through shared memory are halved
// (tc_common.cuh "CTA pairs").  Everything below is per CTA unless it says "leader".
template <bool PAIR>
__global__ void __launch_bounds__(kTcThreads, 1) k_conv_tc(const __grid_constant__ ConvTcArgs G) {
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = rank == 0u;
  YB_CTS(0, threadIdx.x == 0);
  if (warp == 2) warm_param_lines(G);
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  unsigned char* tiles_ptr = smem_dyn + (tiles - raw);
  const uint32_t a_bytes = 128u * G.BK * 2u;
  const uint32_t bn_cta = PAIR ? (uint32_t)G.BN / 2u : (uint32_t)G.BN;      // weight rows this CTA holds
  const uint32_t b_bytes = bn_cta * G.BK * 2u;
  // resident mode: [all taps x chunks weight tiles][A ring]; streaming mode: ring of (A, B) stages
  const uint32_t res_bytes = G.resident ? b_bytes * (uint32_t)(G.taps * G.kchunks) : 0u;
  const uint32_t stage_bytes = G.resident ? a_bytes : a_bytes + b_bytes;
  const uint32_t ring = tiles + res_bytes;
  const uint32_t pipe_bytes = res_bytes + stage_bytes * G.stages;
  const uint32_t staging = tiles + pipe_bytes;                       // 1024-aligned (all pieces are multiples of 1024)
  const uint32_t round_bytes = (G.out_f32 ? 32768u : 16384u) * (uint32_t)G.epi_cpr;
  const uint32_t staging_bytes = round_bytes * (uint32_t)G.epi_ring;
  const uint32_t pool_groups = (uint32_t)(G.BN + 63) / 64u;
  unsigned char* aux = tiles_ptr + pipe_bytes + staging_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(aux);  // full[S], empty[S], tmem_full[2], tmem_empty[2], resident
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux + 8 * (2 * G.stages + 6));
  float* bias_s = reinterpret_cast<float*>(aux + 8 * (2 * G.stages + 6) + 16);
  const uint32_t bar_full0 = smem_u32(bars), bar_empty0 = bar_full0 + 8u * G.stages;
  const uint32_t bar_tfull0 = bar_full0 + 16u * G.stages, bar_tempty0 = bar_tfull0 + 16u;
  const uint32_t bar_res = bar_tempty0 + 16u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&G.tm_a);
    tma_prefetch_desc(&G.tm_b);
    for (int i = 0; i < G.n_store_maps; ++i) tma_prefetch_desc(&G.tm_y[i]);
    for (int s = 0; s < G.stages; ++s) { mbar_init(bar_full0 + 8u * s, 1); mbar_init(bar_empty0 + 8u * s, 1); }
    // tmem_empty: one arrival per CTA of the pair (the leader's MMA warp waits for both epilogues)
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull0 + 8u * i, 1); mbar_init(bar_tempty0 + 8u * i, PAIR ? 2 : 1); }
    mbar_init(bar_res, 1);
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
  const int total_chunks = G.taps * G.kchunks;
  const int tiles_m_per_img = G.tiles_w * G.tiles_h;
  // Tile walk.  Streaming: t = blockIdx.x + i * grid over (m, n) pairs, n fastest.  Resident: the CTA keeps one
  // N tile (blockIdx.x % tiles_n) and strides over the M tiles.
  // In pair mode the work unit is (pair of M tiles, N tile) and the walker is the pair: unit u covers M tiles
  // 2u and 2u + 1 (this CTA: 2u + rank; a tile past the end loads zeros and stores nothing -- TMA clips both).
  const int walker = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int walkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int res_n = walker % G.tiles_n;
  const int t_first = G.resident ? walker / G.tiles_n : walker;
  const int t_step = G.resident ? walkers / G.tiles_n : walkers;
  const int t_end = G.resident ? G.total_tiles / G.tiles_n : G.total_tiles;
  const int b_row0 = (int)(rank * bn_cta);            // first weight row of this CTA inside the N tile
  if (G.resident && warp == 0) {
    // weights are constants: fetch them while the previous kernel is still draining
    if (elect_one()) {
      if (leader) mbar_expect_tx(bar_res, PAIR ? 2u * res_bytes : res_bytes);
      for (int tap = 0; tap < G.taps; ++tap)
        for (int kk = 0; kk < G.kchunks; ++kk) {
          const uint32_t dst = tiles + (uint32_t)(tap * G.kchunks + kk) * b_bytes;
          if (PAIR) tma_load_3d_2sm(&G.tm_b, bar_res, dst, kk * G.BK, tap, res_n * G.BN + b_row0);
          else tma_load_3d(&G.tm_b, bar_res, dst, kk * G.BK, tap, res_n * G.BN);
        }
    }
    __syncwarp();
  }
  YB_CTS(1, threadIdx.x == 0);
  pdl_launch_dependents();   // set-up done: the next kernel's CTAs may take SMs as ours retire
  // griddepcontrol.wait is taken by the producer warp only, AFTER it has worked out its first tile (that arithmetic
  // then overlaps the previous kernel's tail instead of following it); everything the other warps touch that an
  // earlier kernel wrote or still reads -- accumulators, the output, the residual -- is ordered behind the
  // producer's loads through the mbarrier chain (the epilogue takes the wait itself before residual loads).

  if (warp == 0) {
    // ================================ TMA producer ================================
    // Warp-uniform control flow, one elected lane issues the copies (coordinates live in uniform registers).
    const uint32_t tx1 = (uint32_t)(G.BW * G.BH) * G.BK * 2u + (G.resident ? 0u : b_bytes);
    const uint32_t tx = PAIR ? 2u * tx1 : tx1;      // the leader's barrier collects the bytes of both CTAs
    uint32_t s = 0, ph = 1;  // ring slot + parity to wait for on its empty barrier
    bool first = true;
    for (int t = t_first; t < t_end; t += t_step) {
      const int t_m = G.resident ? t : fd_div(t, G.fd_n);
      const int n_tile = G.resident ? res_n : t - t_m * G.tiles_n;
      int m_tile = PAIR ? 2 * t_m + (int)rank : t_m;
      const int img = fd_div(m_tile, G.fd_mimg);
      m_tile -= img * tiles_m_per_img;
      const int th = fd_div(m_tile, G.fd_w), tw = m_tile - th * G.tiles_w;
      const int n0 = n_tile * G.BN, w0 = tw * G.BW, h0 = th * G.BH;
      int dy = 0, dx = 0;
      for (int tap = 0; tap < G.taps; ++tap) {
        const int cw = w0 * G.stride_w + dx - G.pad_w;
        const int chh = G.flat ? 0 : h0 * G.stride_h + dy - G.pad_h;
        for (int kk = 0; kk < G.kchunks; ++kk) {
          if (first) {
            asm volatile("" ::"r"(cw), "r"(chh), "r"(img), "r"(n0));     // coordinates first, then the wait
            pdl_wait();                 // activations written by the previous kernel are visible from here on
            YB_CTS(2, lane == 0);
            first = false;
          }
          mbar_wait(bar_empty0 + 8u * s, ph);
          if (elect_one()) {
            const uint32_t full = bar_full0 + 8u * s;
            if (leader) mbar_expect_tx(full, tx);
            const uint32_t sa = ring + s * stage_bytes;
            if (PAIR) {
              tma_load_4d_2sm(&G.tm_a, full, sa, kk * G.BK, cw, chh, img);
              if (!G.resident) tma_load_3d_2sm(&G.tm_b, full, sa + a_bytes, kk * G.BK, tap, n0 + b_row0);
            } else {
              tma_load_4d(&G.tm_a, full, sa, kk * G.BK, cw, chh, img);
              if (!G.resident) tma_load_3d(&G.tm_b, full, sa + a_bytes, kk * G.BK, tap, n0);
            }
          }
          __syncwarp();
          YB_CTS(3, lane == 0 && t == t_first && tap == 0 && kk == 0);
          if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
        }
        if (++dx == G.kw) { dx = 0; ++dy; }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (pair mode: the leader CTA only) ==================================
    // Whole warp runs the (uniform) control flow; one elected lane issues tcgen05.mma / commit.
    if (leader) {
    const uint32_t sbo = 8u * (uint32_t)G.BK * 2u;
    const int nk = G.BK / 16;
    const uint32_t hi32 = (uint32_t)(make_smem_desc(0, sbo, (uint32_t)G.sw_mode) >> 32);
    const uint32_t boff = a_bytes >> 4;
    const uint32_t bres = (tiles >> 4) & 0x3FFF, bstep = b_bytes >> 4;
    if (G.resident) mbar_wait(bar_res, 0);
    uint32_t s = 0, ph = 0;
    int it = 0;
    for (int t = t_first; t < t_end; t += t_step, ++it) {
      const uint32_t acc = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      if (PAIR) mbar_wait_cluster(bar_tempty0 + 8u * acc, aph ^ 1u);   // both CTAs' epilogues have drained it
      else mbar_wait(bar_tempty0 + 8u * acc, aph ^ 1u);  // epilogue has drained this accumulator
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_d = tmem_base + acc * (uint32_t)G.BN;
      // N of this tile's MMAs: the last N tile of a Cout that is not a multiple of BN (320 = 256 + 64 ...) only computes
      // the columns that exist, rounded up to the instruction's granularity of 16 (single-CTA form; a pair splits the
      // weight rows between its CTAs by BN / 2, so it keeps the full N)
      uint32_t idesc = G.idesc;
      if (!PAIR) {
        const int n_tile_m = G.resident ? res_n : t - fd_div(t, G.fd_n) * G.tiles_n;
        const int n_valid = min(G.BN, (G.bias_len - n_tile_m * G.BN + 15) & ~15);
        idesc = (idesc & ~(0x3Fu << 17)) | ((uint32_t)(n_valid >> 3) << 17);
      }
      for (int c = 0; c < total_chunks; ++c) {
        mbar_wait(bar_full0 + 8u * s, ph);
        YB_CTS(4, lane == 0 && it == 0 && c == 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t da = ((ring + s * stage_bytes) >> 4) & 0x3FFF;
        const uint32_t db = G.resident ? bres + (uint32_t)c * bstep : da + boff;
        if (elect_one()) {
          const uint32_t acc0 = (uint32_t)(c != 0);
          if (PAIR) {
            if (nk == 4) {
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_bf16_lh_2sm(tmem_d, da + 2u * k, db + 2u * k, hi32, G.idesc, k ? 1u : acc0);
            } else if (nk == 2) {
#pragma unroll
              for (int k = 0; k < 2; ++k) umma_bf16_lh_2sm(tmem_d, da + 2u * k, db + 2u * k, hi32, G.idesc, k ? 1u : acc0);
            } else {
              umma_bf16_lh_2sm(tmem_d, da, db, hi32, G.idesc, acc0);
            }
            umma_commit_2sm(bar_empty0 + 8u * s);   // frees this stage in BOTH CTAs once these MMAs retire
          } else {
          if (nk == 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16_lh(tmem_d, da + 2u * k, db + 2u * k, hi32, idesc, k ? 1u : acc0);
          } else if (nk == 2) {
#pragma unroll
            for (int k = 0; k < 2; ++k) umma_bf16_lh(tmem_d, da + 2u * k, db + 2u * k, hi32, idesc, k ? 1u : acc0);
          } else {
            umma_bf16_lh(tmem_d, da, db, hi32, idesc, acc0);
          }
          umma_commit(bar_empty0 + 8u * s);   // frees the smem stage once these MMAs retire
          }
        }
        __syncwarp();
        if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
      }
      if (elect_one()) { if (PAIR) umma_commit_2sm(bar_tfull0 + 8u * acc); else umma_commit(bar_tfull0 + 8u * acc); }
      __syncwarp();
      YB_CTS(5, lane == 0 && it == 0);
    }
    }
  } else {
    // ================================ epilogue ====================================
    const int ew = warp - 2;
    const int q = warp & 3;                 // TMEM lane quarter this warp may read
    const int half = ew >> 2;               // which half of the 32-column chunks this warp owns
    const int row = q * 32 + lane;
    const uint32_t xr = (uint32_t)(row & 7);
    const int nchunks = G.BN / 32;
    // Two ways to spread a tile over the eight epilogue warps:
    //  alt (two staging slots fit): the two groups of four warps (each covers the four TMEM lane quarters) take
    //      alternate tiles -- group g owns accumulator g and slot g -- so two tiles' tcgen05.ld -> math -> store
    //      chains are in flight; for short-K layers that chain's latency is the critical path.
    //  single slot: all eight warps share every tile, the two warps of a lane quarter taking the even / odd
    //      32-column chunks.
    const bool alt = G.epi_ring >= 2;
    const int grp = alt ? half : 0;
    const int nthr = alt ? 128 : 256;
    const bool store_thread = alt ? ((ew & 3) == 0 && lane == 0) : (warp == 2 && lane == 0);
    const uint32_t tile_stage = (G.out_f32 ? 32768u : 16384u) * (uint32_t)G.epi_cpr;
    const uint32_t sbase = staging + (uint32_t)grp * tile_stage;
    const int c_first = alt ? 0 : half, c_step = alt ? 1 : 2;
    const int r_h = row / G.BW, r_w = row - r_h * G.BW;     // this thread's pixel inside a 2-D tile
    if (G.res_ptr) pdl_wait();     // the residual is read with plain loads, ahead of the accumulator barrier
    // The activation is a compile-time constant inside the loop (straight-line element code).
    auto run = [&](auto act_c) {
      constexpr int ACT = decltype(act_c)::value;
      int it = grp;
      for (int t = t_first + grp * t_step; t < t_end; t += (alt ? 2 : 1) * t_step, it += alt ? 2 : 1) {
        const uint32_t acc = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
        const int t_m = G.resident ? t : fd_div(t, G.fd_n);
        const int n_tile = G.resident ? res_n : t - t_m * G.tiles_n;
        int m_tile = PAIR ? 2 * t_m + (int)rank : t_m;
        const int img = fd_div(m_tile, G.fd_mimg);
        m_tile -= img * tiles_m_per_img;
        const int th = fd_div(m_tile, G.fd_w), tw = m_tile - th * G.tiles_w;
        const int n0 = n_tile * G.BN, w0 = tw * G.BW, h0 = th * G.BH;
        mbar_wait(bar_tfull0 + 8u * acc, aph);
        YB_CTS(6, store_thread && it == 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // the slot's previous stores (issued a whole tile ago by this store thread) must have read it
        if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(nthr) : "memory");
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (uint32_t)G.BN;
        // residual: this thread's pixel of the other branch (flat mode: pixel index; 2-D tiles: row / BW, row % BW)
        const __nv_bfloat16* rpix = nullptr;
        if (G.res_ptr) {
          if (G.flat) {
            const long long pix = (long long)w0 + row;
            if (img == 0 && pix < G.out_pixels) rpix = G.res_ptr + pix * G.res_cs;
          } else {
            const int hh = h0 + r_h, ww = w0 + r_w;
            if (img < G.tiles_imgs && hh < G.out_h && ww < G.out_w)
              rpix = G.res_ptr + (((long long)img * G.out_h + hh) * G.out_w + ww) * G.res_cs;
          }
        }
        // 32-column chunks that hold real output channels: the last N tile of a Cout that is not a multiple of the tile
        // (160, 320, 480 ... in the P6 cfgs) skips the TMEM reads, the activation math and the staging of the rest
        const int valid_cols = min(G.BN, G.bias_len - n0);
        const int nch = min(nchunks, (valid_cols + 31) >> 5);
        for (int cc = c_first; cc < nch; cc += c_step) {
          uint32_t v[32];
          if (G.res_ptr) {   // loads issued before the TMEM wait so that their latency overlaps it
            uint4 res[4];
            const int ch0 = n0 + cc * 32;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              res[j] = (rpix && ch0 + 8 * j < G.res_c) ? __ldg(reinterpret_cast<const uint4*>(rpix + ch0 + 8 * j))
                                                       : make_uint4(0u, 0u, 0u, 0u);
            tmem_ld32(taddr + (uint32_t)(cc * 32), v);
            epi_chunk_bf16_res<ACT>(v, bias_s + ch0, G.slope, res, sbase + (uint32_t)(cc >> 1) * 16384u + (uint32_t)row * 128u,
                                    (uint32_t)(cc & 1), xr);
            continue;
          }
          tmem_ld32(taddr + (uint32_t)(cc * 32), v);
          const float* bs = bias_s + n0 + cc * 32;
          if (G.out_f32) epi_chunk_f32<ACT>(v, bs, G.slope, sbase + (uint32_t)cc * 16384u + (uint32_t)row * 128u, xr);
          else epi_chunk_bf16<ACT>(v, bs, G.slope, sbase + (uint32_t)(cc >> 1) * 16384u + (uint32_t)row * 128u,
                                   (uint32_t)(cc & 1), xr);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> visible to TMA
        asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(nthr) : "memory");
        if (store_thread) {
          // every warp of the tile has read this accumulator: hand it back to the (leader's) MMA warp
          if (PAIR && !leader) mbar_arrive_remote(bar_tempty0 + 8u * acc, 0u);
          else mbar_arrive(bar_tempty0 + 8u * acc);
          const int per_group = G.out_f32 ? 32 : 64;
          const int ngroups = (valid_cols + per_group - 1) / per_group;
          for (int mp = 0; mp < G.n_store_maps; ++mp)
            for (int g = 0; g < ngroups; ++g)
              tma_store_4d(&G.tm_y[mp], sbase + (uint32_t)g * 16384u, n0 + g * per_group, w0, G.flat ? 0 : h0, img);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          YB_CTS(7, it == 0);
        }
        if (G.pool) {
          // MaxPool2d(2, 2) of the tile just staged (rows = pixels py * BW + px, 128-byte rows, SW128): thread ->
          // (pooled pixel q, 16-byte chunk j, 64-channel group g) takes the max of its four pixels and stores the 16
          // bytes straight to global memory (8 lanes = one 128-byte run).  No extra shared memory: the TMA store of
          // the full tile reads the same staging concurrently, which is fine (both only read); the next tile's writes
          // to this slot come after the bar.sync at its start.
          const int tig = alt ? (int)threadIdx.x - 64 - grp * 128 : (int)threadIdx.x - 64;
          const int half_w = G.BW >> 1;
          for (int item = tig; item < (int)pool_groups * 256; item += nthr) {
            const int g = item >> 8, q = (item >> 3) & 31, j = item & 7;
            const int qy = q / half_w, qx = q - qy * half_w;
            const int py = (h0 >> 1) + qy, px = (w0 >> 1) + qx, ch = n0 + g * 64 + j * 8;
            if (py >= G.pool_h || px >= G.pool_w || ch >= G.pool_c || img >= G.tiles_imgs) continue;
            const uint32_t r00 = (uint32_t)(2 * qy * G.BW + 2 * qx);
            const uint32_t gb = sbase + (uint32_t)g * 16384u;
            uint4 v[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t r = r00 + (uint32_t)(e & 1) + (uint32_t)(e >> 1) * (uint32_t)G.BW;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[e].x), "=r"(v[e].y), "=r"(v[e].z), "=r"(v[e].w)
                           : "r"(gb + r * 128u + ((((uint32_t)j) ^ (r & 7u)) << 4)));
            }
            uint4 o;
            uint32_t* ow = &o.x;
            const uint32_t* a0 = &v[0].x; const uint32_t* a1 = &v[1].x; const uint32_t* a2 = &v[2].x; const uint32_t* a3 = &v[3].x;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
              const __nv_bfloat162 m01 = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(a0 + w), *reinterpret_cast<const __nv_bfloat162*>(a1 + w));
              const __nv_bfloat162 m23 = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(a2 + w), *reinterpret_cast<const __nv_bfloat162*>(a3 + w));
              const __nv_bfloat162 m = __hmax2(m01, m23);
              ow[w] = *reinterpret_cast<const uint32_t*>(&m);
            }
            *reinterpret_cast<uint4*>(G.pool_ptr + (((size_t)img * G.pool_h + py) * G.pool_w + px) * G.pool_cs + ch) = o;
          }
        }
      }
    };
    if (G.act == YB_ACT_SILU) run(std::integral_constant<int, YB_ACT_SILU>{});
    else if (G.act == YB_ACT_LEAKY) run(std::integral_constant<int, YB_ACT_LEAKY>{});
    else run(std::integral_constant<int, YB_ACT_NONE>{});
    if ((ew & 3) == 0 && lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // smem may go once it has been read; the writes complete with the grid
    YB_CTS(8, warp == 2 && lane == 0);
    YB_CTS(9, warp == 2 && lane == 0);     // (pair mode returns before the stamp at the end)
  }
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

struct ConvTcPlan {
  ConvTcArgs args;
  int grid;
  size_t smem;
  void* w_repack = nullptr;   // pair-view weights owned by the plan (stride-2, 32-channel inputs)
  ~ConvTcPlan() { if (w_repack) cudaFree(w_repack); }
};

constexpr size_t kAuxBytes = 8 * (2 * 12 + 6) + 16 + (size_t)kMaxBias * 4 + 64;

// one round per tile: each epilogue warp converts half of the tile's 32-column chunks
static int epi_cpr_for(int bn, int y_es) { (void)y_es; return (bn / 32 + 1) / 2; }

// `bn_cta` = weight rows a CTA holds (BN, or BN / 2 in pair mode); the output staging is always the CTA's own 128 x BN.
static size_t tile_smem(int bn, int bk, int y_es, bool resident, int chunks, int stages, int ring = 1, int bn_cta = 0) {
  if (!bn_cta) bn_cta = bn;
  const size_t a = (size_t)128 * bk * 2, b = (size_t)bn_cta * bk * 2;
  const size_t out = (size_t)ring * (y_es == 4 ? 32768 : 16384) * epi_cpr_for(bn, y_es);
  return 1024 + (resident ? b * chunks + a * stages : (a + b) * stages) + out + kAuxBytes;
}

// Tile-N / weight-residency / CTA-pair heuristic: estimated cycles = rounds of the persistent grid x per-tile cost,
// where a tile costs max(MMA issue, operand bytes / ~42 B/clk/SM of L2->smem throughput, epilogue).  With resident
// weights the per-tile operand bytes are the activations only and the weights are fetched once per CTA.  In pair mode
// (cta_group::2) every CTA loads half of the weight tile: same MMA time per tile, fewer operand bytes, but the walkers
// are pairs (half as many) each covering two M tiles, and a pair waits for the slower of its two CTAs.
// pair_mode: 0 = choose, 1 = force, -1 = never.
static void pick_tile(int cout, long long m_tiles, int chunks, int bk, int y_es, int pair_mode, int* bn_out, int* resident_out,
                      int* pair_out) {
  const int cands[4] = {32, 64, 128, 256};
  static const char* force_env = getenv("YB_RESIDENT");   // tuning aid: 0 = never, 1 = whenever it fits
  const int force = force_env ? atoi(force_env) : -1;
  static const char* pair_env = getenv("YB_PAIR");        // tuning aid: overrides a pair_mode of 0
  if (pair_mode == 0 && pair_env) pair_mode = atoi(pair_env);
  if (pair_mode == 0) {
    // Measured on the yolov7 shapes at batch 32 (scripts/pair_sweep_specs.txt, profiles/round2_pair_sweep.txt): pairs win
    // 5-12 % where the K loop is long (1x1 with Cin >= 1024, stride-2 3x3 with Cin >= 128) and lose 5-30 % on short-K
    // layers, where the extra cluster handshakes are not amortised; with fewer tiles than SMs they only pay off for
    // very long K.
    const long long k_total = (long long)chunks * bk;
    // (profiles/round2_tile_sweep.txt: with fewer M tiles than SMs pairs still win 3-7 % once K >= 2048 or Cout spans
    // two or more 256-wide tiles; a single N tile at 20 x 20 loses)
    pair_mode = (k_total >= 1024 && (k_total >= 2048 || m_tiles >= num_sms() || cout > 256)) ? 1 : -1;
  }
  int best = 64, best_res = 0, best_pair = 0;
  double best_cost = 1e30;
  for (int pr = 0; pr < 2; ++pr) {
    if ((pr && pair_mode < 0) || (!pr && pair_mode > 0)) continue;
    for (int i = 0; i < 4; ++i) {
      const int bn = cands[i];
      if (bn == 32 && cout > 32) continue;
      if (bn > 32 && cout <= 32) continue;
      if (y_es == 4 && bn > 128) continue;  // fp32 staging tile would not leave room for the pipeline
      const int bn_cta = pr ? bn / 2 : bn;
      const int tiles_n = (cout + bn - 1) / bn;
      const int walkers_max = pr ? num_sms() / 2 : num_sms();
      const long long units_m = pr ? (m_tiles + 1) / 2 : m_tiles;
      for (int res = 0; res < 2; ++res) {
        if (force == 0 && res) continue;
        if (res && tiles_n != 1) continue;   // measured: splitting Cout to make the weights fit re-reads A and loses
        if (res && (tile_smem(bn, bk, y_es, true, chunks, 4, 1, bn_cta) > 227 * 1024 || tiles_n > walkers_max)) continue;
        if (!res && tile_smem(bn, bk, y_es, false, chunks, 2, 1, bn_cta) > 227 * 1024) continue;
        const long long units = units_m * tiles_n;
        int walkers = (int)(units < walkers_max ? units : walkers_max);
        if (res) walkers = walkers / tiles_n * tiles_n;
        if (walkers < 1) continue;
        const double rounds = (double)((units + walkers - 1) / walkers);
        const double mma = chunks * (bk / 16) * (bn / 2.0);
        const double a_b = 128.0 * bk * 2, b_b = (double)bn_cta * bk * 2;
        const double load = chunks * (a_b + (res ? 0.0 : b_b)) / 42.0;
        const double epi = ((double)cout / tiles_n < bn ? (double)cout / tiles_n : bn) * 10.0 + 600.0;   // chunks past Cout are skipped
        double tile = mma > load ? mma : load;
        if (epi > tile) tile = epi;
        // pairs: cluster launch / cross-CTA handshakes cost a little per tile and at start-up
        const double cost = rounds * (tile + (pr ? 500.0 : 400.0)) + (res ? chunks * b_b / 42.0 : 0.0) + (pr ? 1500.0 : 0.0);
        if (force == 1 && res) { const double c1 = cost * 1e-6; if (c1 < best_cost) { best_cost = c1; best = bn; best_res = 1; best_pair = pr; } continue; }
        if (cost < best_cost * 0.999) { best_cost = cost; best = bn; best_res = res; best_pair = pr; }
      }
    }
  }
  *bn_out = best; *resident_out = best_res; *pair_out = best_pair;
}

// [Cout_pad][9][cin_pad] (Cin = 32) -> [Cout_pad][3 dy][2 pair taps][64]: a stride-2 3x3 conv over 32-channel pixels
// read as a (3 x 2, stride 2 x 1) conv over 64-channel pixel PAIRS: pair tap 0 = (zeros | dx 0) of pair ox - 1,
// pair tap 1 = (dx 1 | dx 2) of pair ox.  128-byte TMA rows instead of 64-byte ones, 6 taps instead of 9.
__global__ void k_repack_pair(const __nv_bfloat16* w, int cin_pad, int cout_pad, __nv_bfloat16* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cout_pad * 6 * 64) return;
  const int ch = i & 63, pt = (i >> 6) % 2, dy = (i >> 7) % 3, co = i / 384;
  const int c = ch & 31, hi = ch >> 5;
  const int dx = pt == 0 ? (hi ? 0 : -1) : (hi ? 2 : 1);
  out[i] = dx < 0 ? __float2bfloat16_rn(0.f) : w[((size_t)co * 9 + dy * 3 + dx) * cin_pad + c];
}

int conv_tc_plan_create(const yb_conv_desc& d, ConvTcPlan** out) {
  const yb_tensor& x = d.x;
  const yb_tensor& y = d.y;
  YB_CHECK_ARG(x.dtype == YB_BF16, "conv(tc): x must be bf16");
  YB_CHECK_ARG(y.dtype == YB_BF16 || y.dtype == YB_F32, "conv(tc): y must be bf16 or fp32");
  YB_CHECK_ARG(x.c_stride % 8 == 0 && d.cin_pad % 8 == 0 && (uintptr_t)x.data % 16 == 0 && (uintptr_t)d.weight % 16 == 0,
               "conv(tc): input channel stride / cin_pad must be multiples of 8 and pointers 16-byte aligned");
  const int y_es = y.dtype == YB_F32 ? 4 : 2;
  YB_CHECK_ARG(((size_t)y.c_stride * y_es) % 16 == 0 && (uintptr_t)y.data % 16 == 0,
               "conv(tc): output channel stride must be a multiple of 16 bytes");
  const int k = d.k, s = d.stride, pad = k / 2, ups = d.upsample;
  const int Ho = (x.H + 2 * pad - k) / s + 1, Wo = (x.W + 2 * pad - k) / s + 1;
  const int Cout = y.C;
  ConvTcPlan* P = new ConvTcPlan();
  ConvTcArgs& A = P->args;
  memset(&A, 0, sizeof(A));
  A.bias = d.bias; A.bias_len = d.cout_pad;
  // pair view (see k_repack_pair): needs pixel pairs to be contiguous in memory
  const bool pair_view = k == 3 && s == 2 && x.C == 32 && x.c_stride == 32 && x.W % 2 == 0 && ups == 1 && d.tile_n >= 0;
  int Cin = x.C, cin_pad = d.cin_pad, xW = x.W, x_cs = x.c_stride;
  int kh = k, kw = k, sh = s, sw = s, ph = pad, pw = pad;
  const void* weight = d.weight;
  if (pair_view) {
    Cin = 64; cin_pad = 64; xW = x.W / 2; x_cs = 64;
    kh = 3; kw = 2; sh = 2; sw = 1; ph = 1; pw = 1;
    const size_t n = (size_t)d.cout_pad * 6 * 64;
    if (cudaMalloc(&P->w_repack, n * 2) != cudaSuccess) { delete P; set_error("conv(tc): cudaMalloc failed"); return YB_ERR_CUDA; }
    k_repack_pair<<<(int)((n + 255) / 256), 256>>>((const __nv_bfloat16*)d.weight, d.cin_pad, d.cout_pad,
                                                   (__nv_bfloat16*)P->w_repack);
    if (cudaDeviceSynchronize() != cudaSuccess) { delete P; set_error("conv(tc): weight repack failed"); return YB_ERR_CUDA; }
    weight = P->w_repack;
  }
  A.kw = kw; A.taps = kh * kw; A.stride_h = sh; A.stride_w = sw; A.pad_h = ph; A.pad_w = pw;
  A.act = d.act; A.slope = d.leaky_slope;
  A.out_f32 = y.dtype == YB_F32;
  A.BK = Cin >= 64 ? 64 : (Cin >= 32 ? 32 : 16);
  if (A.BK > cin_pad) A.BK = cin_pad >= 32 ? 32 : 16;
  YB_CHECK_ARG(cin_pad >= 16, "conv(tc): cin_pad must be >= 16");
  A.kchunks = (Cin + A.BK - 1) / A.BK;
  A.sw_mode = A.BK == 64 ? 2 : (A.BK == 32 ? 4 : 6);
  const CUtensorMapSwizzle sw_in = A.BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                                              : (A.BK == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  const bool pool = d.pool.data != nullptr;
  if (pool) {
    YB_CHECK_ARG(ups == 1 && y.dtype == YB_BF16 && d.pool.dtype == YB_BF16 && d.pool.C == y.C && d.pool.N == y.N &&
                     d.pool.H == Ho / 2 && d.pool.W == Wo / 2 && d.pool.c_stride % 8 == 0 && (uintptr_t)d.pool.data % 16 == 0,
                 "conv(tc): fused max-pool output must be bf16 (N, H/2, W/2, C), 16-byte aligned, without upsample");
    YB_CHECK_ARG(d.pool.H > 0 && d.pool.W > 0, "conv(tc): fused max-pool needs an output of at least 2 x 2");
  }
  if (d.res.data) {
    YB_CHECK_ARG(ups == 1 && y.dtype == YB_BF16 && d.res.dtype == YB_BF16 && d.res.N == y.N && d.res.H == Ho && d.res.W == Wo &&
                     d.res.C == y.C && d.res.c_stride % 8 == 0 && (uintptr_t)d.res.data % 16 == 0,
                 "conv(tc): residual must be a bf16 (N, Ho, Wo, Cout) tensor, 16-byte aligned, without upsample");
    A.res_ptr = (const __nv_bfloat16*)d.res.data; A.res_cs = d.res.c_stride;
    A.res_c = (Cout + 7) / 8 * 8 <= d.res.c_stride ? (Cout + 7) / 8 * 8 : Cout / 8 * 8;
  }
  A.out_h = Ho; A.out_w = Wo; A.out_pixels = (long long)x.N * Ho * Wo; A.tiles_imgs = x.N;
  A.pool = pool ? 1 : 0;
  A.flat = (k == 1 && s == 1 && ups == 1 && !pool) ? 1 : 0;
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
      if (pool && ((bw & 1) || (bh & 1) || bw * bh != 128)) continue;   // 2x2 blocks must not straddle tiles
      if (bh > Ho && !pool) bh = Ho;
      if (bh * sh > 256 || bw * sw > 256) continue;
      int cost = ((Wo + bw - 1) / bw) * ((Ho + bh - 1) / bh);
      if (cost < best_cost || (cost == best_cost && bw > best_bw)) { best_cost = cost; best_bw = bw; }
    }
    static const char* bw_env = getenv("YB_TILE_BW");   // tuning aid
    if (bw_env && atoi(bw_env) > 0 && atoi(bw_env) <= 128) best_bw = atoi(bw_env);
    A.BW = best_bw;
    A.BH = (pool || 128 / best_bw < Ho) ? 128 / best_bw : Ho;
    A.tiles_w = (Wo + A.BW - 1) / A.BW;
    A.tiles_h = (Ho + A.BH - 1) / A.BH;
  }
  const int total_chunks = A.taps * A.kchunks;
  const long long m_tiles = (long long)A.tiles_w * A.tiles_h * images;
  // the 2x2 upsample store walks four tensor maps per tile; pairs are not worth it there (tiny layers)
  const int pair_mode = ups == 2 ? -1 : d.pair;
  if (d.tile_n) {
    A.BN = d.tile_n < 0 ? -d.tile_n : d.tile_n;
    A.resident = 0;
    A.pair = pair_mode > 0 ? 1 : 0;
    if (A.BN == 1) pick_tile(Cout, m_tiles, total_chunks, A.BK, y_es, pair_mode, &A.BN, &A.resident, &A.pair);
  } else {
    pick_tile(Cout, m_tiles, total_chunks, A.BK, y_es, pair_mode, &A.BN, &A.resident, &A.pair);
  }
  YB_CHECK_ARG(A.BN == 32 || A.BN == 64 || A.BN == 128 || A.BN == 256, "conv(tc): tile_n must be 32/64/128/256");
  const int bn_cta = A.pair ? A.BN / 2 : A.BN;
  A.tiles_n = (Cout + A.BN - 1) / A.BN;
  YB_CHECK_ARG(m_tiles < 0x3fffffffLL, "conv(tc): grid too large");
  A.fd_n = make_fastdiv(A.tiles_n); A.fd_mimg = make_fastdiv(A.tiles_w * A.tiles_h); A.fd_w = make_fastdiv(A.tiles_w);
  A.n_store_maps = ups == 2 ? 4 : 1;
  A.tmem_cols = 2 * A.BN < 32 ? 32 : 2 * A.BN;   // double-buffered accumulators (64/128/256/512 columns)
  A.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(A.BN >> 3) << 17) | ((uint32_t)((A.pair ? 256 : 128) >> 4) << 24);
  // stages / shared memory
  const size_t a_stage = (size_t)128 * A.BK * 2, b_tile = (size_t)bn_cta * A.BK * 2;
  const size_t stage_bytes = A.resident ? a_stage : a_stage + b_tile;
  // output staging ring: deep enough to overlap the TMA store of one round with the next round's math, as long
  // as the operand pipeline keeps at least 4 stages
  A.epi_cpr = epi_cpr_for(A.BN, y_es);
  // A deeper ring overlaps a tile's TMA store with the next tile's epilogue math: worth its shared memory for
  // short-K layers (the epilogue is the critical path), not when it would thin out the operand pipeline.
  int ring = 2;
  size_t fixed = 0;
  for (; ring >= 1; --ring) {
    fixed = tile_smem(A.BN, A.BK, y_es, A.resident != 0, total_chunks, 0, ring, bn_cta);
    if (ring == 1) break;
    if (fixed >= 226 * 1024) continue;
    const int st = (int)((226 * 1024 - fixed) / stage_bytes);
    if (st >= 8 || (total_chunks <= 6 && st >= 4)) break;
  }
  A.epi_ring = ring;
  int stages = d.stages ? d.stages : (int)((226 * 1024 - fixed) / stage_bytes);
  if (stages > 12) stages = 12;
  if (stages < 2) stages = 2;
  A.stages = stages;
  P->smem = fixed + stage_bytes * stages;
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
    uint32_t box[4], es[4] = {1, (uint32_t)sw, (uint32_t)sh, 1};
    const uint64_t cs = (uint64_t)x_cs * 2;
    if (A.flat) {
      const uint64_t M = (uint64_t)x.N * x.H * x.W;
      dims[0] = Cin; dims[1] = M; dims[2] = 1; dims[3] = 1;
      str[0] = cs; str[1] = cs * M; str[2] = cs * M;
      box[0] = A.BK; box[1] = 128; box[2] = 1; box[3] = 1;
      es[1] = es[2] = 1;
    } else {
      dims[0] = Cin; dims[1] = xW; dims[2] = x.H; dims[3] = x.N;
      str[0] = cs; str[1] = cs * xW; str[2] = cs * xW * x.H;
      box[0] = A.BK; box[1] = A.BW * sw; box[2] = A.BH * sh; box[3] = 1;
    }
    rc = encode(&A.tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data, dims, str, box, es, sw_in, "activations");
    if (rc) { delete P; return rc; }
  }
  // ---- B: weights [cout_pad][taps][cin_pad]
  {
    uint64_t dims[3] = {(uint64_t)cin_pad, (uint64_t)A.taps, (uint64_t)d.cout_pad};
    uint64_t str[2] = {(uint64_t)cin_pad * 2, (uint64_t)cin_pad * 2 * A.taps};
    uint32_t box[3] = {(uint32_t)A.BK, 1, (uint32_t)bn_cta};     // pair mode: each CTA loads half of the N tile
    rc = encode(&A.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, weight, dims, str, box, one4, sw_in, "weights");
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
  if (pool) {
    A.pool_ptr = (__nv_bfloat16*)d.pool.data; A.pool_h = d.pool.H; A.pool_w = d.pool.W; A.pool_cs = d.pool.c_stride;
    A.pool_c = (Cout + 7) / 8 * 8 <= d.pool.c_stride ? (Cout + 7) / 8 * 8 : Cout / 8 * 8;   // whole 8-channel chunks
    A.tiles_imgs = x.N;
  }
  YB_CHECK_ARG(y.H == Ho * ups && y.W == Wo * ups && y.N == x.N, "conv(tc): output spatial dims mismatch");
  const long long units = (long long)A.tiles_n * (A.pair ? (m_tiles + 1) / 2 : m_tiles);
  YB_CHECK_ARG(units < 0x3fffffffLL, "conv(tc): grid too large");
  A.total_tiles = (int)units;
  const int walkers_max = A.pair ? num_sms() / 2 : num_sms();
  int walkers = (int)(units < walkers_max ? units : walkers_max);
  if (A.resident) walkers = walkers / A.tiles_n * A.tiles_n;
  YB_CHECK_ARG(walkers >= 1, "conv(tc): no walker fits the resident layout");
  P->grid = A.pair ? 2 * walkers : walkers;
  A.dbg = conv_dbg_slot();
  *out = P;
  return YB_OK;
}

int conv_tc_launch(const ConvTcPlan* P, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_conv_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    YB_CUDA(cudaFuncSetAttribute(k_conv_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  if (P->args.pair) YB_CUDA(launch_pdl_pairs(k_conv_tc<true>, P->grid, kTcThreads, P->smem, s, P->args));
  else YB_CUDA(launch_pdl(k_conv_tc<false>, P->grid, kTcThreads, P->smem, s, P->args));
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
  YB_CHECK_ARG(d->x.dtype != YB_F32 || (d->pool.data == nullptr && d->res.data == nullptr),
               "yb_conv2d: the fused max-pool output / residual need the bf16 path");
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
    ConvS2pPlan* ps = nullptr;
    rc3 = conv3s2p_plan_create(dd, &ps);
    if (rc3 == YB_OK) {
      rc3 = conv3s2p_launch(ps, (cudaStream_t)stream);
      conv3s2p_plan_destroy(ps);   // cudaFree of the repacked weights waits for the kernel
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

// ---- reusable single-conv plans (tensor maps encoded once; yb_conv2d re-encodes them on every call)
struct yb_conv_plan {
  yb::ConvTcPlan* tc = nullptr;
  yb::Conv3Plan* c3 = nullptr;
  yb::ConvS2pPlan* s2p = nullptr;
};

extern "C" int yb_conv_plan_create(const yb_conv_desc* d, yb_conv_plan** out) {
  YB_CHECK_ARG(d && out && view_ok(d->x) && view_ok(d->y) && d->weight && d->bias, "yb_conv_plan_create: bad descriptor");
  YB_CHECK_ARG(d->x.dtype == YB_BF16, "yb_conv_plan_create: bf16 activations only");
  YB_CHECK_ARG((d->k == 1 || d->k == 3) && (d->stride == 1 || d->stride == 2), "yb_conv_plan_create: k in {1,3}, stride in {1,2}");
  yb_conv_desc dd = *d;
  yb_conv_plan* p = new yb_conv_plan();
  int rc = YB_ERR_UNSUPPORTED;
  if (dd.tile_n >= 0) rc = conv3_tc_plan_create(dd, &p->c3);
  if (rc == YB_ERR_UNSUPPORTED && dd.tile_n >= 0) rc = conv3s2p_plan_create(dd, &p->s2p);
  else dd.tile_n = dd.tile_n == -1 ? 0 : -dd.tile_n;
  if (rc == YB_ERR_UNSUPPORTED) rc = conv_tc_plan_create(dd, &p->tc);
  if (rc != YB_OK) { delete p; return rc; }
  *out = p;
  return YB_OK;
}

extern "C" int yb_conv_plan_launch(const yb_conv_plan* p, void* stream) {
  YB_CHECK_ARG(p && (p->tc || p->c3 || p->s2p), "yb_conv_plan_launch: null plan");
  if (p->s2p) return conv3s2p_launch(p->s2p, (cudaStream_t)stream);
  return p->c3 ? conv3_tc_launch(p->c3, (cudaStream_t)stream) : conv_tc_launch(p->tc, (cudaStream_t)stream);
}

extern "C" void yb_conv_plan_destroy(yb_conv_plan* p) {
  if (!p) return;
  if (p->tc) conv_tc_plan_destroy(p->tc);
  if (p->c3) conv3_tc_plan_destroy(p->c3);
  if (p->s2p) conv3s2p_plan_destroy(p->s2p);
  delete p;
}
