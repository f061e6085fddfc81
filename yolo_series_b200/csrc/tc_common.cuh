// Shared device-side PTX wrappers (TMA, mbarrier, tcgen05, TMEM) and the host-side tensor-map encoder
// for the tcgen05 convolution kernels (conv_tc.cu, conv3_tc.cu).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace yb {

// ------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* tm, uint32_t bar, uint32_t dst, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_3d(const CUtensorMap* tm, uint32_t bar, uint32_t dst, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(tm), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
// stamp `slot` of this CTA's 16-entry row from one thread when the debug buffer is set (scripts/conv_timeline.py,
// scripts/conv_chain_timeline.py; the buffer holds 16 x grid entries)
#define YB_CTS(slot, cond)                                                                                  \
  do {                                                                                                      \
    if (G.dbg && (cond)) {                                                                                  \
      unsigned long long t_;                                                                                \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                \
      G.dbg[blockIdx.x * 16 + slot] = t_;                                                                   \
    }                                                                                                       \
  } while (0)
// Debug buffer of a newly created plan (host side; NULL without the env).  YB_CONV_TS = device address; with
// YB_CONV_TS_STRIDE = entries per plan every successfully created plan takes the next slot (creation order = the
// engine's op order), which gives a timeline of a whole graph replay (scripts/step_timeline.py).
unsigned long long* conv_dbg_slot();

// Division of a tile index by a launch constant: n / d = (n * mul) >> shift for 0 <= n <= 2^30 with mul = ceil(2^k / d),
// k = 31 + floor(log2 d) (the error term n * (mul * d - 2^k) stays below 2^k).  A runtime 32-bit division is ~30
// dependent instructions; the role warps decompose a tile index with three or four of them per tile, which a lone warp
// executes in about a microsecond (scripts/conv_chain_timeline.py) -- on the critical path of the first tile of every
// launch and of the same order as a short-K tile's period.
struct FastDiv {
  uint32_t mul, shift;
};
inline FastDiv make_fastdiv(int d) {
  FastDiv f;
  int l = 0;
  while ((2ll << l) <= (long long)d) ++l;
  const int k = 31 + l;
  f.mul = (uint32_t)(((1ull << k) + (unsigned long long)d - 1ull) / (unsigned long long)d);
  f.shift = (uint32_t)k;
  return f;
}
__device__ __forceinline__ int fd_div(int n, FastDiv f) {
  return (int)(((unsigned long long)(uint32_t)n * f.mul) >> f.shift);
}

// Kernel parameters live in constant bank 0 and every launch gets a fresh copy, so an SM's first read of each 64-byte
// line is a miss of ~0.4-0.5 us; the role loops read their fields one after another, which serialised three or four such
// misses between griddepcontrol.wait and the first TMA issue (scripts/conv_chain_timeline.py: 1.1-1.9 us).  One warp
// reads a word of every line up front -- independent loads with constant offsets (LDC), so the misses overlap and they
// happen while the previous kernel is still draining.
template <class T>
__device__ __forceinline__ void warm_param_lines(const T& G) {
  constexpr int kLines = (int)((sizeof(T) + 63) / 64);
  const uint32_t* p = reinterpret_cast<const uint32_t*>(&G);
  uint32_t acc = 0;
#pragma unroll
  for (int l = 0; l < kLines; ++l) acc ^= p[l * 16];
  // a use ptxas cannot drop (an empty asm leaves the loads dead at the PTX level and they are removed)
  asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %0, 0xD1CEB00C;\n\t@p nanosleep.u32 1;\n\t}" ::"r"(acc));
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// Same, with the descriptors split into their 32-bit halves: only the low word (smem address field)
// changes between MMAs of one stage, so the issuing lane does one 32-bit add per operand.
__device__ __forceinline__ void umma_bf16_lh(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// A and B with different upper descriptor words (different group strides: conv3w_tc.cu)
__device__ __forceinline__ void umma_bf16_lh2(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                              uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
//   [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout type
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout_type & 7) << 61;
  return d;
}

__device__ __forceinline__ float act_fast(float v, int act, float slope) {
  if (act == YB_ACT_SILU) return __fdividef(v, 1.0f + __expf(-v));
  if (act == YB_ACT_LEAKY) return v > 0.f ? v : v * slope;
  return v;
}

constexpr int kEpiWarps = 8;                       // two warps per TMEM lane quarter, each takes half the columns
constexpr int kTcThreads = 64 + 32 * kEpiWarps;    // warp 0 producer, warp 1 MMA, warps 2.. epilogue
constexpr int kMaxBias = 2048;

// SiLU with ONE MUFU op per element: x*sigmoid(x) = 0.5x*tanh(0.5x) + 0.5x (tanh.approx rel. err ~2^-11 << bf16 ulp)
__device__ __forceinline__ float silu_fast(float v) {
  const float h = 0.5f * v;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}
__device__ __forceinline__ float act_epi(float v, int act, float slope) {
  if (act == YB_ACT_SILU) return silu_fast(v);
  if (act == YB_ACT_LEAKY) return v > 0.f ? v : v * slope;
  return v;
}
// factor the kernels apply to the bias when they copy it into shared memory (see epi_bias_act)
__device__ __forceinline__ float epi_bias_scale(int act) { return act == YB_ACT_SILU ? 0.5f : 1.0f; }
// ---- epilogue: one 32-column TMEM chunk of one accumulator row -> bias + activation -> swizzled staging.
// ACT is a template parameter so the 32 elements form one straight-line block (a per-element runtime
// switch serialises the LDS -> FADD -> MUFU chains); the bias is read with 8 LDS.128.
template <int ACT>
__device__ __forceinline__ void epi_bias_act(const uint32_t (&v)[32], const float* bs, float slope, float (&f)[32]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 b = reinterpret_cast<const float4*>(bs)[j];
    const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (ACT == YB_ACT_SILU) {
        // the kernels keep bias / 2 in shared memory for SiLU layers (epi_bias_scale): h = (acc + bias) / 2 is ONE FFMA,
        // bit-identical to the add followed by the halving (a power-of-two scale commutes with the rounding)
        const float h = fmaf(__uint_as_float(v[4 * j + e]), 0.5f, bb[e]);
        float t;
        asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
        f[4 * j + e] = fmaf(h, t, h);
      } else {
        const float x = __uint_as_float(v[4 * j + e]) + bb[e];
        f[4 * j + e] = ACT == YB_ACT_LEAKY ? (x > 0.f ? x : x * slope) : x;
      }
    }
  }
}
// bf16 output: the chunk is half (sub = 0/1) of a 128-byte staging row (64 channels), SW128 pattern.
template <int ACT>
__device__ __forceinline__ void epi_chunk_bf16(const uint32_t (&v)[32], const float* bs, float slope, uint32_t row_base,
                                               uint32_t sub, uint32_t xr) {
  float f[32];
  epi_bias_act<ACT>(v, bs, slope, f);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t p[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      __nv_bfloat162 hh = __floats2bfloat162_rn(f[8 * j + 2 * e], f[8 * j + 2 * e + 1]);
      p[e] = *reinterpret_cast<uint32_t*>(&hh);
    }
    const uint32_t addr = row_base + (((sub * 4u + (uint32_t)j) ^ xr) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(p[0]), "r"(p[1]), "r"(p[2]), "r"(p[3])
                 : "memory");
  }
}
// Same with a residual (Shortcut, models/common.py:85-86: y = act(conv) + res): `res` holds this pixel's 32 bf16 values of
// the other branch (zeros where it has none); the sum is formed in fp32 and rounded to bf16 once.
template <int ACT>
__device__ __forceinline__ void epi_chunk_bf16_res(const uint32_t (&v)[32], const float* bs, float slope, const uint4 (&res)[4],
                                                   uint32_t row_base, uint32_t sub, uint32_t xr) {
  float f[32];
  epi_bias_act<ACT>(v, bs, slope, f);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t rw[4] = {res[j].x, res[j].y, res[j].z, res[j].w};
    uint32_t p[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 r2 = *reinterpret_cast<const __nv_bfloat162*>(&rw[e]);
      __nv_bfloat162 hh = __floats2bfloat162_rn(f[8 * j + 2 * e] + __low2float(r2), f[8 * j + 2 * e + 1] + __high2float(r2));
      p[e] = *reinterpret_cast<uint32_t*>(&hh);
    }
    const uint32_t addr = row_base + (((sub * 4u + (uint32_t)j) ^ xr) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(p[0]), "r"(p[1]), "r"(p[2]), "r"(p[3])
                 : "memory");
  }
}
// fp32 output: the chunk is a whole 128-byte staging row (32 channels).
template <int ACT>
__device__ __forceinline__ void epi_chunk_f32(const uint32_t (&v)[32], const float* bs, float slope, uint32_t row_base,
                                              uint32_t xr) {
  float f[32];
  epi_bias_act<ACT>(v, bs, slope, f);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t addr = row_base + ((((uint32_t)j) ^ xr) << 4);
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(f[4 * j]), "f"(f[4 * j + 1]),
                 "f"(f[4 * j + 2]), "f"(f[4 * j + 3]) : "memory");
  }
}
// All chunks [c_begin, c_end) of one accumulator for this warp.
template <int ACT>
__device__ __forceinline__ void epi_chunks(uint32_t taddr, int c_begin, int c_end, const float* bias_n0, float slope,
                                           uint32_t staging, int row, uint32_t xr, bool out_f32) {
  for (int cc = c_begin; cc < c_end; ++cc) {
    uint32_t v[32];
    tmem_ld32(taddr + (uint32_t)(cc * 32), v);
    const float* bs = bias_n0 + cc * 32;
    if (out_f32) epi_chunk_f32<ACT>(v, bs, slope, staging + (uint32_t)cc * 16384u + (uint32_t)row * 128u, xr);
    else epi_chunk_bf16<ACT>(v, bs, slope, staging + (uint32_t)(cc >> 1) * 16384u + (uint32_t)row * 128u,
                             (uint32_t)(cc & 1), xr);
  }
}
__device__ __forceinline__ void epi_chunks_dispatch(int act, uint32_t taddr, int c_begin, int c_end, const float* bias_n0,
                                                    float slope, uint32_t staging, int row, uint32_t xr, bool out_f32) {
  if (act == YB_ACT_SILU) epi_chunks<YB_ACT_SILU>(taddr, c_begin, c_end, bias_n0, slope, staging, row, xr, out_f32);
  else if (act == YB_ACT_LEAKY) epi_chunks<YB_ACT_LEAKY>(taddr, c_begin, c_end, bias_n0, slope, staging, row, xr, out_f32);
  else epi_chunks<YB_ACT_NONE>(taddr, c_begin, c_end, bias_n0, slope, staging, row, xr, out_f32);
}

// Programmatic dependent launch: a kernel launched with launch_pdl() may start while its predecessor in the
// stream is still draining.  Everything before pdl_wait() must only touch memory no earlier kernel writes
// (weights, bias, tensor maps, shared memory / TMEM set-up); pdl_wait() returns once the predecessor has
// completed and its writes are visible.  pdl_launch_dependents() lets the successor's CTAs be scheduled as
// soon as SM resources free up.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// One elected lane of a fully converged warp.  The surrounding control flow stays warp-uniform, so
// descriptor / address arithmetic is done once on the uniform datapath instead of per MMA via R2UR.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}


// ------------------------------------------------------------------------------ CTA pairs (cta_group::2)
// Two CTAs of a 2-CTA cluster (the two SMs of a TPC) run ONE tcgen05.mma of M = 256: each CTA supplies its own 128 rows
// of A and HALF of the B tile (N/2 rows) from its own shared memory -- same offsets in both CTAs -- and receives its
// 128 rows of D in its own TMEM.  Per CTA the B traffic from L2 and the B reads from shared memory are halved.
// Protocol (the leader = cluster rank 0 issues the MMAs):
//   * both CTAs issue their TMA loads with the .cta_group::2 form, completing on the LEADER's full barrier (the
//     barrier address with the peer bit cleared); the leader arms it with the bytes of both CTAs;
//   * tcgen05.commit.cta_group::2 with a multicast mask releases the smem stage / publishes the accumulator on the
//     barriers at the same offset in BOTH CTAs;
//   * the peer's epilogue hands its accumulator half back with a remote arrive on the leader's tmem_empty barrier.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's even CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// wait with cluster-scope acquire: the arrival may come from the peer CTA
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAITC_LOOP:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAITC_DONE;\n\t"
      "bra WAITC_LOOP;\n\t"
      "WAITC_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// arrive on the barrier at the same offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar), "r"(rank));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void tma_load_4d_2sm(const CUtensorMap* tm, uint32_t bar, uint32_t dst, int c0, int c1, int c2,
                                                int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_3d_2sm(const CUtensorMap* tm, uint32_t bar, uint32_t dst, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void umma_bf16_lh_2sm(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi,
                                                 uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once these MMAs retire) on the barrier at this offset in both CTAs of the pair
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}

// ------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

inline int encode(CUtensorMap* tm, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* estr, CUtensorMapSwizzle sw,
                  const char* what) {
  EncodeTiledFn fn = get_encode();
  if (!fn) {
    set_error("conv(tc): cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return YB_ERR_CUDA;
  }
  cuuint64_t d[5], st[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = estr[i]; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  CUresult r = fn(tm, dt, (cuuint32_t)rank, const_cast<void*>(base), d, st, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("conv(tc): cuTensorMapEncodeTiled(%s) failed with %d (dims %llu,%llu,%llu,%llu box %u,%u,%u,%u)", what,
              (int)r, (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)(rank > 2 ? dims[2] : 0),
              (unsigned long long)(rank > 3 ? dims[3] : 0), box[0], box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return YB_ERR_CUDA;
  }
  return YB_OK;
}


}  // namespace yb
