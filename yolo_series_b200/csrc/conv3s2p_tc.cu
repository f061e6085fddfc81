// 3x3 stride-2 convolution over 32-channel inputs (the second layer of yolov7 / yolov7-tiny: Conv(32, 64, 3, 2),
// cfg/deploy/yolov7.yaml node 1) as a "pair slab" kernel.
//
// The input is read as 64-channel pixel PAIRS (x = 2p, 2p+1 are adjacent in NHWC memory), so the conv becomes a
// (3 x 2, stride 2 x 1) conv over pairs: pair tap 0 = (zeros | dx 0) of pair ox - 1, pair tap 1 = (dx 1 | dx 2) of
// pair ox (weights repacked once per plan, see k_repack_pair in conv_tc.cu).  Per tile of BH x BW outputs ONE slab
// of the input is loaded -- even rows 2*oy (filter row 1) and odd rows 2*oy - 1 (filter rows 0 and 2), each
// BWS = BW + 1 pairs wide -- and all six taps read it through UMMA descriptor offsets:
//   filter row  -> whole slab rows (multiples of BWS*128 B, 1024-byte aligned),
//   pair tap 1  -> + one 128-byte row.  The SW128 XOR is a function of the absolute shared-memory address, so a
//                  descriptor that starts one row into a swizzle atom simply reads the rows one further down
//                  (verified on B200: base_offset must stay 0).
// An MMA covers BH image rows of BWS slab columns; the last column of each row is a junk output (its pair tap 1
// wraps into the next row) and is never stored.  The six weight tiles stay resident; every input byte crosses
// L2 -> SM once per tile instead of six times, which is what bounded the per-tap kernel on this layer.
#include <type_traits>

#include "tc_common.cuh"

namespace yb {

__global__ void k_repack_pair(const __nv_bfloat16* w, int cin_pad, int cout_pad, __nv_bfloat16* out);   // conv_tc.cu

struct ConvS2pArgs {
  CUtensorMap tm_e, tm_o;          // pair view of the input, rows strided by 2: even-row / odd-row boxes
  CUtensorMap tm_b;                // [cout_pad][6][64] pair-tap weights
  CUtensorMap tm_y;                // output, box {64, BW, 1, 1}
  const float* bias;
  int tiles_w, tiles_h, total_tiles;
  int BWS, BH, BN;                 // slab width (pairs), output rows per tile, N tile (= padded Cout: 32 | 64)
  int Ho;
  int stages, epi_ring;
  int act; float slope;
  int bias_len, tmem_cols;
  uint32_t a_bytes, a_tx;          // stage bytes (1024-aligned, incl. the one-row overrun) / bytes TMA delivers
  unsigned long long* dbg;         // tuning aid: globaltimer stamps, 16 per CTA (tc_common.cuh YB_CTS)
  uint32_t idesc;
};

__global__ void __launch_bounds__(kTcThreads, 1) k_conv3s2p_tc(const __grid_constant__ ConvS2pArgs G) {
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  YB_CTS(0, threadIdx.x == 0);
  if (warp == 2) warm_param_lines(G);     // tc_common.cuh: parameter lines are cold per launch
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  unsigned char* tiles_ptr = smem_dyn + (tiles - raw);
  const uint32_t b_bytes = (uint32_t)G.BN * 128u;
  const uint32_t res_bytes = 6u * b_bytes;
  const uint32_t ring = tiles + res_bytes;
  const uint32_t pipe_bytes = res_bytes + G.a_bytes * (uint32_t)G.stages;
  const uint32_t staging = tiles + pipe_bytes;
  const uint32_t staging_bytes = 16384u * (uint32_t)G.epi_ring;
  unsigned char* aux = tiles_ptr + pipe_bytes + staging_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(aux);  // full[S], empty[S], tmem_full[2], tmem_empty[2], resident
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux + 8 * (2 * G.stages + 6));
  float* bias_s = reinterpret_cast<float*>(aux + 8 * (2 * G.stages + 6) + 16);
  const uint32_t bar_full0 = smem_u32(bars), bar_empty0 = bar_full0 + 8u * G.stages;
  const uint32_t bar_tfull0 = bar_full0 + 16u * G.stages, bar_tempty0 = bar_tfull0 + 16u;
  const uint32_t bar_res = bar_tempty0 + 16u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&G.tm_e);
    tma_prefetch_desc(&G.tm_o);
    tma_prefetch_desc(&G.tm_b);
    tma_prefetch_desc(&G.tm_y);
    for (int s = 0; s < G.stages; ++s) { mbar_init(bar_full0 + 8u * s, 1); mbar_init(bar_empty0 + 8u * s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull0 + 8u * i, 1); mbar_init(bar_tempty0 + 8u * i, 1); }
    mbar_init(bar_res, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)G.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {   // bias of the channels this launch can touch (the last N tile may reach past Cout: zeros there)
    const int nb = min(kMaxBias, (G.bias_len + 255) & ~255);
    const float bsc = epi_bias_scale(G.act);
    for (int i = threadIdx.x; i < nb; i += kTcThreads) bias_s[i] = (i < G.bias_len) ? bsc * G.bias[i] : 0.f;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) {   // weights are constants: fetch them while the previous kernel is still draining
    if (elect_one()) {
      mbar_expect_tx(bar_res, res_bytes);
      for (int tap = 0; tap < 6; ++tap) tma_load_3d(&G.tm_b, bar_res, tiles + (uint32_t)tap * b_bytes, 0, tap, 0);
    }
    __syncwarp();
  }
  pdl_launch_dependents();
  pdl_wait();
  YB_CTS(2, threadIdx.x == 0);
  const int tiles_per_img = G.tiles_w * G.tiles_h;
  const int BW = G.BWS - 1;
  const uint32_t e_bytes = (uint32_t)(G.BH * G.BWS) * 128u;     // even-row part of the slab; odd rows follow

  if (warp == 0) {
    // ================================ TMA producer ================================
    uint32_t s = 0, ph = 1;
    for (int t = blockIdx.x; t < G.total_tiles; t += gridDim.x) {
      const int img = t / tiles_per_img;
      const int m = t - img * tiles_per_img;
      const int th = m / G.tiles_w, tw = m - th * G.tiles_w;
      const int w0 = tw * BW, h0 = th * G.BH;
      mbar_wait(bar_empty0 + 8u * s, ph);
      if (elect_one()) {
        const uint32_t full = bar_full0 + 8u * s;
        mbar_expect_tx(full, G.a_tx);
        const uint32_t sa = ring + s * G.a_bytes;
        tma_load_4d(&G.tm_e, full, sa, 0, w0 - 1, 2 * h0, img);               // rows 2*(h0+i)
        tma_load_4d(&G.tm_o, full, sa + e_bytes, 0, w0 - 1, 2 * h0 - 1, img);  // rows 2*(h0+i) - 1, i = 0..BH
      }
      __syncwarp();
      if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    const uint32_t hi32 = (uint32_t)(make_smem_desc(0, 1024u, 2u) >> 32);   // SW128, 8-row groups 1024 B apart
    const uint32_t bres = (tiles >> 4) & 0x3FFF, bstep = b_bytes >> 4;
    // slab row offsets of the three filter rows (in 16-byte units): dy 1 -> even rows, dy 0 / 2 -> odd rows i / i+1
    uint32_t roff[3];
    roff[1] = 0u;
    roff[0] = e_bytes >> 4;
    roff[2] = (e_bytes + (uint32_t)G.BWS * 128u) >> 4;
    mbar_wait(bar_res, 0);
    uint32_t s = 0, ph = 0;
    int it = 0;
    for (int t = blockIdx.x; t < G.total_tiles; t += gridDim.x, ++it) {
      const uint32_t acc = (uint32_t)it & 1u, aph = ((uint32_t)it >> 1) & 1u;
      mbar_wait(bar_tempty0 + 8u * acc, aph ^ 1u);
      mbar_wait(bar_full0 + 8u * s, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_d = tmem_base + acc * (uint32_t)G.BN;
      const uint32_t da = ((ring + s * G.a_bytes) >> 4) & 0x3FFF;
      if (elect_one()) {
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
          for (int pt = 0; pt < 2; ++pt) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_lh(tmem_d, da + roff[dy] + 8u * (uint32_t)pt + 2u * (uint32_t)k,
                           bres + (uint32_t)(dy * 2 + pt) * bstep + 2u * (uint32_t)k, hi32, G.idesc,
                           (dy | pt | k) ? 1u : 0u);
          }
        }
        umma_commit(bar_empty0 + 8u * s);
        umma_commit(bar_tfull0 + 8u * acc);
      }
      __syncwarp();
      if (++s == (uint32_t)G.stages) { s = 0; ph ^= 1u; }
    }
  } else {
    // ================================ epilogue ====================================
    // Two groups of four warps (each group covers the four TMEM lane quarters) take alternate tiles, so two
    // tiles' tcgen05.ld -> math -> store chains are in flight: with one MMA stage per tile the chain latency,
    // not its throughput, is what limits a single group.  Group g always owns accumulator g and staging slot g.
    const int ew = warp - 2;
    const int q = warp & 3;
    const int grp = ew >> 2;
    const int row = q * 32 + lane;
    const uint32_t xr = (uint32_t)(row & 7);
    const int nchunks = G.BN / 32;
    const bool store_thread = ((ew & 3) == 0 && lane == 0);
    const uint32_t sbase = staging + (uint32_t)grp * 16384u;
    auto run = [&](auto act_c) {
      constexpr int ACT = decltype(act_c)::value;
      uint32_t aph = 0;
      for (int t = blockIdx.x + grp * (int)gridDim.x; t < G.total_tiles; t += 2 * (int)gridDim.x, aph ^= 1u) {
        const int img = t / tiles_per_img;
        const int m = t - img * tiles_per_img;
        const int th = m / G.tiles_w, tw = m - th * G.tiles_w;
        const int w0 = tw * BW, h0 = th * G.BH;
        mbar_wait(bar_tfull0 + 8u * (uint32_t)grp, aph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // this group's previous stores (two tiles ago) have long been issued; make sure they have read the slot
        if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        for (int cc = 0; cc < nchunks; ++cc) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)grp * (uint32_t)G.BN + (uint32_t)(cc * 32), v);
          epi_chunk_bf16<ACT>(v, bias_s + cc * 32, G.slope, sbase + (uint32_t)row * 128u, (uint32_t)cc, xr);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        if (store_thread) {
          mbar_arrive(bar_tempty0 + 8u * (uint32_t)grp);
          for (int i = 0; i < G.BH; ++i)   // one box per image row: BW valid pixels, the junk column is skipped
            if (h0 + i < G.Ho) tma_store_4d(&G.tm_y, sbase + (uint32_t)(i * G.BWS) * 128u, 0, w0, h0 + i, img);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    };
    if (G.act == YB_ACT_SILU) run(std::integral_constant<int, YB_ACT_SILU>{});
    else if (G.act == YB_ACT_LEAKY) run(std::integral_constant<int, YB_ACT_LEAKY>{});
    else run(std::integral_constant<int, YB_ACT_NONE>{});
    if (store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // smem may go once it has been read; the writes complete with the grid
  }
  YB_CTS(9, threadIdx.x == 64);
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G.tmem_cols)
                 : "memory");
  }
}

struct ConvS2pPlan {
  ConvS2pArgs args;
  int grid;
  size_t smem;
  void* w_repack = nullptr;
  ~ConvS2pPlan() { if (w_repack) cudaFree(w_repack); }
};

// YB_ERR_UNSUPPORTED = shape not handled here (caller falls back to the generic kernel).
int conv3s2p_plan_create(const yb_conv_desc& d, ConvS2pPlan** out) {
  if (d.pool.data || d.res.data) return YB_ERR_UNSUPPORTED;   // fused max-pool / residual live in the generic kernel
  const yb_tensor& x = d.x;
  const yb_tensor& y = d.y;
  if (d.k != 3 || d.stride != 2 || d.upsample != 1 || x.dtype != YB_BF16 || y.dtype != YB_BF16 || d.res.data)
    return YB_ERR_UNSUPPORTED;
  if (x.C != 32 || x.c_stride != 32 || x.W % 2 || x.H % 2 || y.C > 64 || d.cin_pad < 32 || d.tile_n < 0)
    return YB_ERR_UNSUPPORTED;
  if ((uintptr_t)x.data % 16 || (uintptr_t)y.data % 16 || y.c_stride % 8 || (uintptr_t)d.weight % 16 || d.cout_pad > 64)
    return YB_ERR_UNSUPPORTED;
  const int Ho = x.H / 2, Wo = x.W / 2, Cout = y.C;
  ConvS2pPlan* P = new ConvS2pPlan();
  ConvS2pArgs& A = P->args;
  memset(&A, 0, sizeof(A));
  A.bias = d.bias; A.bias_len = d.cout_pad; A.act = d.act; A.slope = d.leaky_slope; A.Ho = Ho;
  A.BN = Cout <= 32 ? 32 : 64;
  A.tmem_cols = 2 * A.BN < 32 ? 32 : 2 * A.BN;
  A.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(A.BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  // slab width: BWS in {16, 32, 64} pairs (BW = BWS - 1 valid outputs per row), BH = 128 / BWS rows
  long long best = -1;
  for (int bws = 16; bws <= 64; bws *= 2) {
    const int bh = 128 / bws;
    const long long nt = (long long)((Wo + bws - 2) / (bws - 1)) * ((Ho + bh - 1) / bh);
    const long long cost = nt * ((2 * bh + 1) * bws + 128);   // slab rows + output rows per tile
    if (best < 0 || cost < best) { best = cost; A.BWS = bws; A.BH = bh; }
  }
  const int BW = A.BWS - 1;
  A.tiles_w = (Wo + BW - 1) / BW;
  A.tiles_h = (Ho + A.BH - 1) / A.BH;
  A.a_tx = (uint32_t)((2 * A.BH + 1) * A.BWS) * 128u;
  A.a_bytes = (uint32_t)align_up((size_t)A.a_tx + 128, 1024);   // + the one-row overrun of pair tap 1
  const size_t aux_bytes = 8 * (2 * 12 + 6) + 16 + (size_t)kMaxBias * 4 + 64;
  const size_t fixed1 = 1024 + 6 * (size_t)A.BN * 128 + aux_bytes;
  A.epi_ring = 2;
  int stages = (int)((226 * 1024 - fixed1 - 2 * 16384) / A.a_bytes);
  if (stages > 8) stages = 8;
  if (stages < 2) { delete P; return YB_ERR_UNSUPPORTED; }
  A.stages = stages;
  P->smem = fixed1 + 2 * 16384 + (size_t)A.a_bytes * stages;
  // ---- weights -> pair taps
  {
    const size_t n = (size_t)d.cout_pad * 6 * 64;
    if (cudaMalloc(&P->w_repack, n * 2) != cudaSuccess) { delete P; set_error("conv3s2p: cudaMalloc failed"); return YB_ERR_CUDA; }
    k_repack_pair<<<(int)((n + 255) / 256), 256>>>((const __nv_bfloat16*)d.weight, d.cin_pad, d.cout_pad,
                                                   (__nv_bfloat16*)P->w_repack);
    if (cudaDeviceSynchronize() != cudaSuccess) { delete P; set_error("conv3s2p: weight repack failed"); return YB_ERR_CUDA; }
  }
  const uint32_t one4[4] = {1, 1, 1, 1};
  int rc;
  {
    const uint64_t cs = 128;   // one pixel pair = 64 bf16
    uint64_t dims[4] = {64, (uint64_t)(x.W / 2), (uint64_t)x.H, (uint64_t)x.N};
    uint64_t str[3] = {cs, cs * (x.W / 2), cs * (x.W / 2) * x.H};
    uint32_t es[4] = {1, 1, 2, 1};
    uint32_t box_e[4] = {64, (uint32_t)A.BWS, (uint32_t)(2 * A.BH), 1};
    uint32_t box_o[4] = {64, (uint32_t)A.BWS, (uint32_t)(2 * (A.BH + 1)), 1};
    rc = encode(&A.tm_e, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data, dims, str, box_e, es, CU_TENSOR_MAP_SWIZZLE_128B,
                "conv3s2p even rows");
    if (!rc) rc = encode(&A.tm_o, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data, dims, str, box_o, es,
                         CU_TENSOR_MAP_SWIZZLE_128B, "conv3s2p odd rows");
    if (rc) { delete P; return rc; }
  }
  {
    uint64_t dims[3] = {64, 6, (uint64_t)d.cout_pad};
    uint64_t str[2] = {128, 128 * 6};
    uint32_t box[3] = {64, 1, (uint32_t)A.BN};
    rc = encode(&A.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, P->w_repack, dims, str, box, one4,
                CU_TENSOR_MAP_SWIZZLE_128B, "conv3s2p weights");
    if (rc) { delete P; return rc; }
  }
  {
    const uint64_t cs = (uint64_t)y.c_stride * 2;
    uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)x.N};
    uint64_t str[3] = {cs, cs * Wo, cs * Wo * Ho};
    uint32_t box[4] = {64, (uint32_t)BW, 1, 1};
    rc = encode(&A.tm_y, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, y.data, dims, str, box, one4, CU_TENSOR_MAP_SWIZZLE_128B,
                "conv3s2p output");
    if (rc) { delete P; return rc; }
  }
  if (y.H != Ho || y.W != Wo || y.N != x.N) { delete P; set_error("conv3s2p: output dims mismatch"); return YB_ERR_INVALID; }
  const long long total = (long long)A.tiles_w * A.tiles_h * x.N;
  A.total_tiles = (int)total;
  P->grid = (int)(total < num_sms() ? total : num_sms());
  A.dbg = conv_dbg_slot();
  *out = P;
  return YB_OK;
}

int conv3s2p_launch(const ConvS2pPlan* P, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    YB_CUDA(cudaFuncSetAttribute(k_conv3s2p_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  YB_CUDA(launch_pdl(k_conv3s2p_tc, P->grid, kTcThreads, P->smem, s, P->args));
  return YB_OK;
}

void conv3s2p_plan_destroy(ConvS2pPlan* p) { delete p; }

}  // namespace yb
