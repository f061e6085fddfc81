// Batched non_max_suppression for B200.
// Restates reference utils/general.py:628-720 (candidate filter, conf = obj*cls, xywh->xyxy,
// best-class / multi-label, class whitelist, top-max_nms, 4096-px class offset, max_det) and the
// greedy suppression of torchvision.ops.nms (call site general.py:704): stable descending sort,
// iou = inter / (area_i + area_j - inter) in fp32, suppress iff (double)iou > iou_thres, only kept
// boxes suppress.  Bit-exact by construction: every fp32 step is an explicit *_rn intrinsic.
//
// Pipeline (no host sync, no per-image launches):
//   k_filter    one warp per prediction row; reads the row once, emits 64-bit sort keys
//               (~score_bits << 32 | row_or_row*nc+cls) by warp-aggregated atomics
//   k_sort_nms  one CTA per image: LSD radix sort of the keys (stable => ties resolved by original
//               order, as torch's stable sort does), then chunked greedy NMS against the kept set
//               held in shared memory, stopping as soon as max_det boxes are kept.
#include "common.cuh"

namespace yb {

constexpr int kMaxClassWords = 32;  // class whitelist bitmask: nc <= 1024
constexpr float kMaxWH = 4096.0f;   // general.py:640 max_wh

struct HeadLevelDev {  // fused decode source (yb_decode_nms)
  const void* logits;
  int dtype, ny, nx, c_stride;
  float stride;
  float anchor_wh[2 * YB_MAX_ANCHORS];
  int row_off;
  unsigned magic_cells, magic_nx;   // floor(2^32 / d) + 1 for d = ny*nx and d = nx
  unsigned magic_cs;                // same for d = c_stride (k_filter_cells staging)
  int blk_off;                      // k_filter_cells: first block (within an image) of this level
};

struct FilterParams {
  // source A: decoded predictions (B, A, no) fp32
  const float* pred;
  // source B: raw head logits (fused decode)
  HeadLevelDev lv[YB_MAX_LEVELS];
  int nl, na;
  int B, A, no, nc;
  float conf_thres;
  int multi_label, use_class_mask;
  uint32_t class_mask[kMaxClassWords];
  int cap;
  int rows_per_block, blocks_per_img;   // launch geometry: grid = B * blocks_per_img
  int cells_per_block;                  // k_filter_cells
  // outputs
  int* count;            // [B]
  uint64_t* keys;        // [B][cap]
  uint16_t* cls_of_row;  // [B][A]      (best-class mode)
  float* boxes;          // [B][A][4]   xywh of candidate rows (fused mode only)
  int* overflow;         // nullable
};

// __frcp_rn(x) and __fdiv_rn(1, x) are both the correctly rounded 1/x, i.e. the same float; the former is cheaper.
__device__ __forceinline__ float sigmoid_rn_(float t) { return __frcp_rn(__fadd_rn(1.0f, expf(-t))); }

__device__ __forceinline__ bool class_allowed(const FilterParams& P, int c) {
  return !P.use_class_mask || ((P.class_mask[c >> 5] >> (c & 31)) & 1u);
}

// One warp per row; lane l holds channels l, l+32, l+64, ... of the row.  A block owns a contiguous run of rows
// of ONE image and collects its candidate keys in shared memory, so the per-image counter sees one atomicAdd per
// batch instead of one per candidate row (25 200 same-address atomics per image serialise in L2 otherwise).
// Slot order inside an image is arbitrary: every key is distinct (it carries the row index) and k_sort_nms
// sorts them, so the result does not depend on it.
constexpr int kFilterBuf = 2048;   // keys buffered per block between flushes

// NV = number of 32-channel groups a row spans (compile-time so that the per-group code is not predicated 8 ways).
template <bool FUSED, typename TL, int NV>
__global__ void __launch_bounds__(256) k_filter(const FilterParams P) {
  __shared__ uint64_t kbuf[kFilterBuf];
  __shared__ int s_n, s_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.x / P.blocks_per_img, blk = blockIdx.x - b * P.blocks_per_img;
  const int row_begin = blk * P.rows_per_block;
  const int row_end = min(P.A, row_begin + P.rows_per_block);
  const int batch_rows = P.multi_label ? kFilterBuf / P.nc : kFilterBuf;   // worst case fills the buffer exactly
  constexpr int MAXV = NV;  // no <= 32 * NV channels
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  for (int batch0 = row_begin; batch0 < row_end; batch0 += batch_rows) {
    const int batch1 = min(row_end, batch0 + batch_rows);
    for (int row = batch0 + warp; row < batch1; row += 8) {
      // ---- locate the row
      const float* src = nullptr;
      const TL* lsrc = nullptr;
      float stride = 0.f, aw = 0.f, ah = 0.f;
      int gx = 0, gy = 0;
      if (!FUSED) {
        src = P.pred + ((size_t)b * P.A + row) * P.no;
      } else {
        int l = 0;
        while (l + 1 < P.nl && row >= P.lv[l + 1].row_off) ++l;
        const HeadLevelDev& L = P.lv[l];
        const int cell = row - L.row_off;
        // cell / (ny*nx) and rem / nx by multiplication (exact for n*d < 2^32; d == 1 has no 32-bit magic)
        const int a = L.magic_cells ? (int)__umulhi((unsigned)cell, L.magic_cells) : cell;
        const int rem = cell - a * L.ny * L.nx;
        gy = L.magic_nx ? (int)__umulhi((unsigned)rem, L.magic_nx) : rem; gx = rem - gy * L.nx;
        lsrc = reinterpret_cast<const TL*>(L.logits) + (((size_t)b * L.ny + gy) * L.nx + gx) * L.c_stride + a * P.no;
        stride = L.stride; aw = L.anchor_wh[2 * a]; ah = L.anchor_wh[2 * a + 1];
      }
      // ---- all channels of the row up front (independent loads), then the arithmetic
      float raw[MAXV];
#pragma unroll
      for (int it = 0; it < MAXV; ++it) {
        const int ch = it * 32 + lane;
        raw[it] = 0.f;
        if (ch < P.no) raw[it] = FUSED ? Elem<TL>::load(lsrc + ch) : __ldg(src + ch);
      }
      const float v0 = FUSED ? sigmoid_rn_(raw[0]) : raw[0];
      const float obj = __shfl_sync(0xffffffffu, v0, 4);
      if (!(obj > P.conf_thres)) continue;  // general.py:637,653
      // xywh (decoded when fused)
      float bx = __shfl_sync(0xffffffffu, v0, 0), by = __shfl_sync(0xffffffffu, v0, 1);
      float bw = __shfl_sync(0xffffffffu, v0, 2), bh = __shfl_sync(0xffffffffu, v0, 3);
      if (FUSED) {
        bx = __fmul_rn(__fadd_rn(__fsub_rn(__fmul_rn(bx, 2.f), 0.5f), (float)gx), stride);
        by = __fmul_rn(__fadd_rn(__fsub_rn(__fmul_rn(by, 2.f), 0.5f), (float)gy), stride);
        float w2 = __fmul_rn(bw, 2.f), h2 = __fmul_rn(bh, 2.f);
        bw = __fmul_rn(__fmul_rn(w2, w2), aw);
        bh = __fmul_rn(__fmul_rn(h2, h2), ah);
      }
      // ---- class confidences: conf_c = cls_c * obj (general.py:673); nc == 1: conf = obj (:670)
      float best = -INFINITY;
      int best_c = 0x7fffffff;
      int n_emit_total = 0;
#pragma unroll
      for (int it = 0; it < MAXV; ++it) {
        const int ch = it * 32 + lane;
        const float v = it == 0 ? v0 : (FUSED ? sigmoid_rn_(raw[it]) : raw[it]);
        const int c = ch - 5;
        const bool is_cls = (c >= 0) && (ch < P.no);
        const float conf = (P.nc == 1) ? obj : __fmul_rn(v, obj);
        if (!P.multi_label) {
          if (is_cls && (conf > best)) { best = conf; best_c = c; }  // strict > keeps the first index in-lane
        } else {
          const bool emit = is_cls && (conf > P.conf_thres) && class_allowed(P, c);
          const unsigned m = __ballot_sync(0xffffffffu, emit);
          if (m) {
            int base = 0;
            const int cnt = __popc(m);
            if (lane == 0) base = atomicAdd(&s_n, cnt);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (emit) {
              const uint32_t idx = (uint32_t)row * (uint32_t)P.nc + (uint32_t)c;
              kbuf[base + __popc(m & ((1u << lane) - 1u))] = ((uint64_t)(~__float_as_uint(conf)) << 32) | idx;
            }
            n_emit_total += cnt;
          }
        }
      }
      if (!P.multi_label) {
        // warp argmax with first-index tie-break (torch.max returns the first maximal index)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          float ob = __shfl_xor_sync(0xffffffffu, best, o);
          int oc = __shfl_xor_sync(0xffffffffu, best_c, o);
          if (ob > best || (ob == best && oc < best_c)) { best = ob; best_c = oc; }
        }
        if (lane == 0 && best > P.conf_thres && class_allowed(P, best_c)) {  // general.py:684,688
          kbuf[atomicAdd(&s_n, 1)] = ((uint64_t)(~__float_as_uint(best)) << 32) | (uint32_t)row;
          P.cls_of_row[(size_t)b * P.A + row] = (uint16_t)best_c;
          n_emit_total = 1;
        }
        n_emit_total = __shfl_sync(0xffffffffu, n_emit_total, 0);
      }
      if (FUSED && n_emit_total > 0 && lane == 0) {
        *reinterpret_cast<float4*>(P.boxes + ((size_t)b * P.A + row) * 4) = make_float4(bx, by, bw, bh);
      }
    }
    // ---- flush the batch: one global atomic, coalesced key stores
    __syncthreads();
    if (threadIdx.x == 0) s_base = s_n ? atomicAdd(P.count + b, s_n) : 0;
    __syncthreads();
    const int n = s_n, base = s_base;
    for (int i = threadIdx.x; i < n; i += 256) {
      const int slot = base + i;
      if (slot < P.cap) P.keys[(size_t)b * P.cap + slot] = kbuf[i];
      else if (P.overflow) *P.overflow = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// k_filter_best: best-class mode (multi_label == False, general.py:682-684), one THREAD per row.
// The warps first stage a tile of rows into shared memory with coalesced loads (a row is `no` contiguous
// values, rows are c_stride / no apart), then every thread walks its own row (row pitch odd -> no bank
// conflicts).  In fused mode the class sigmoid is only evaluated where it can matter:
//   conf_c = fl(sigmoid(l_c) * obj) and the reference takes the FIRST index of the maximum.  With M the largest
//   class logit, every class with l_c < M - kLogitMargin has sigmoid(l_c) <= sigmoid(M - kLogitMargin) =: s_lo
//   (sigmoid_rn_ is monotone up to a few ulps, covered by stepping s_lo up eight ulps), so if fl(s_lo' * obj) is
//   strictly below conf(M) none of them can reach the maximum and only the classes within the margin are
//   evaluated -- exactly (classes equal to M reuse conf(M)).  When the test fails (saturated logits) every
//   class is evaluated.
// ------------------------------------------------------------------------------------------------
constexpr int kFbThreads = 128;
constexpr int kFbInflight = 8;       // rows whose loads a warp has in flight while staging
constexpr float kLogitMargin = 0.015625f;

// One row of the best-class filter: r = the row's `no` values (logits when FUSED, decoded predictions otherwise).
// Returns true and the sort key when the row is a candidate; writes its class (and decoded box when FUSED).
template <bool FUSED>
__device__ __forceinline__ bool fb_process_row(const FilterParams& P, const float* r, int b, int row, int gx, int gy,
                                               float stride, float aw, float ah, unsigned long long& key) {
  bool emit = false;
  const float obj = FUSED ? sigmoid_rn_(r[4]) : r[4];
  if (obj > P.conf_thres) {  // general.py:637,653
    float best = -INFINITY;
    int best_c = 0;
    if (P.nc == 1) {          // general.py:670: conf = obj
      best = obj;
    } else if (!FUSED) {
      for (int c = 0; c < P.nc; ++c) {
        const float conf = __fmul_rn(r[5 + c], obj);
        if (conf > best) { best = conf; best_c = c; }   // strict >: first index of the maximum
      }
    } else {
      float M = -INFINITY;
#pragma unroll 8
      for (int c = 0; c < P.nc; ++c) M = fmaxf(M, r[5 + c]);
      const float conf_M = __fmul_rn(sigmoid_rn_(M), obj);
      const float cut = M - kLogitMargin;
      float s_lo = sigmoid_rn_(cut);
      s_lo = __uint_as_float(__float_as_uint(s_lo) + 8u);            // eight ulps up (s_lo >= 0)
      const bool margin_ok = __fmul_rn(s_lo, obj) < conf_M;          // NaN logits fail the test -> full scan
      // classes below the gate cannot reach the maximum: four of them are dismissed with one branch
      const float gate = margin_ok ? cut : -INFINITY;
      int c = 0;
#pragma unroll 2
      for (; c + 4 <= P.nc; c += 4) {
        const float v0 = r[5 + c], v1 = r[6 + c], v2 = r[7 + c], v3 = r[8 + c];
        if (!(v0 < gate) || !(v1 < gate) || !(v2 < gate) || !(v3 < gate)) {
          const float v[4] = {v0, v1, v2, v3};
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (!(v[u] < gate)) {
              const float conf = v[u] == M ? conf_M : __fmul_rn(sigmoid_rn_(v[u]), obj);   // same logit, same product
              if (conf > best) { best = conf; best_c = c + u; }
            }
          }
        }
      }
      for (; c < P.nc; ++c) {
        const float lc = r[5 + c];
        if (lc < gate) continue;
        const float conf = lc == M ? conf_M : __fmul_rn(sigmoid_rn_(lc), obj);
        if (conf > best) { best = conf; best_c = c; }
      }
    }
    if (best > P.conf_thres && class_allowed(P, best_c)) {  // general.py:684,688
      emit = true;
      key = ((unsigned long long)(~__float_as_uint(best)) << 32) | (uint32_t)row;
      P.cls_of_row[(size_t)b * P.A + row] = (uint16_t)best_c;
      if (FUSED) {
        const float sx = sigmoid_rn_(r[0]), sy = sigmoid_rn_(r[1]), sw = sigmoid_rn_(r[2]), sh = sigmoid_rn_(r[3]);
        const float bx = __fmul_rn(__fadd_rn(__fsub_rn(__fmul_rn(sx, 2.f), 0.5f), (float)gx), stride);
        const float by = __fmul_rn(__fadd_rn(__fsub_rn(__fmul_rn(sy, 2.f), 0.5f), (float)gy), stride);
        const float w2 = __fmul_rn(sw, 2.f), h2 = __fmul_rn(sh, 2.f);
        *reinterpret_cast<float4*>(P.boxes + ((size_t)b * P.A + row) * 4) =
            make_float4(bx, by, __fmul_rn(__fmul_rn(w2, w2), aw), __fmul_rn(__fmul_rn(h2, h2), ah));
      }
    }
  }
  return emit;
}

template <bool FUSED, typename TL, int NV>
__global__ void __launch_bounds__(kFbThreads) k_filter_best(const FilterParams P) {
  extern __shared__ __align__(16) unsigned char fb_smem[];
  float* tile = reinterpret_cast<float*>(fb_smem);             // [kFbThreads][pitch]
  __shared__ uint64_t kbuf[kFbThreads];
  __shared__ int s_n, s_cnt, s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int pitch = P.no | 1;
  const int b = blockIdx.x / P.blocks_per_img, blk = blockIdx.x - b * P.blocks_per_img;
  const int row_begin = blk * P.rows_per_block;
  const int row_end = min(P.A, row_begin + P.rows_per_block);
  if (tid == 0) s_n = 0;       // re-armed by thread 0 after every flush (never while another thread may read it)
  for (int t0 = row_begin; t0 < row_end; t0 += kFbThreads) {
    const int t1 = min(row_end, t0 + kFbThreads);
    // ---- stage rows [t0, t1): warp w owns rows t0 + w + 4k; eight rows' loads are issued before their stores
    constexpr int kWarpsFb = kFbThreads / 32;
    for (int row0 = t0 + warp; row0 < t1; row0 += kFbInflight * kWarpsFb) {
      float v[kFbInflight][NV];
#pragma unroll
      for (int u = 0; u < kFbInflight; ++u) {
        const int row = row0 + u * kWarpsFb;
        const bool live = row < t1;
        const float* src = nullptr;
        const TL* lsrc = nullptr;
        if (live) {
          if (!FUSED) {
            src = P.pred + ((size_t)b * P.A + row) * P.no;
          } else {
            int l = 0;
            while (l + 1 < P.nl && row >= P.lv[l + 1].row_off) ++l;
            const HeadLevelDev& L = P.lv[l];
            const int cell = row - L.row_off;
            const int a = L.magic_cells ? (int)__umulhi((unsigned)cell, L.magic_cells) : cell;
            const int rem = cell - a * L.ny * L.nx;     // = gy * nx + gx
            lsrc = reinterpret_cast<const TL*>(L.logits) + ((size_t)b * L.ny * L.nx + rem) * L.c_stride + a * P.no;
          }
        }
#pragma unroll
        for (int it = 0; it < NV; ++it) {
          const int ch = it * 32 + lane;
          v[u][it] = 0.f;
          if (live && ch < P.no) v[u][it] = FUSED ? Elem<TL>::load(lsrc + ch) : __ldg(src + ch);
        }
      }
#pragma unroll
      for (int u = 0; u < kFbInflight; ++u) {
        const int row = row0 + u * kWarpsFb;
        if (row < t1) {
          float* dstp = tile + (row - t0) * pitch;
#pragma unroll
          for (int it = 0; it < NV; ++it) {
            const int ch = it * 32 + lane;
            if (ch < P.no) dstp[ch] = v[u][it];
          }
        }
      }
    }
    __syncthreads();
    // ---- one thread per row
    const int row = t0 + tid;
    bool emit = false;
    unsigned long long key = 0ull;
    if (row < t1) {
      const float* r = tile + tid * pitch;
      float stride = 0.f, aw = 0.f, ah = 0.f;
      int gx = 0, gy = 0;
      if (FUSED) {
        int l = 0;
        while (l + 1 < P.nl && row >= P.lv[l + 1].row_off) ++l;
        const HeadLevelDev& L = P.lv[l];
        const int cell = row - L.row_off;
        const int a = L.magic_cells ? (int)__umulhi((unsigned)cell, L.magic_cells) : cell;
        const int rem = cell - a * L.ny * L.nx;
        gy = L.magic_nx ? (int)__umulhi((unsigned)rem, L.magic_nx) : rem;
        gx = rem - gy * L.nx;
        stride = L.stride; aw = L.anchor_wh[2 * a]; ah = L.anchor_wh[2 * a + 1];
      }
      emit = fb_process_row<FUSED>(P, r, b, row, gx, gy, stride, aw, ah, key);
    }
    const unsigned mk = __ballot_sync(0xffffffffu, emit);
    int pos = 0;
    if (lane == 0 && mk) pos = atomicAdd(&s_n, __popc(mk));
    pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(mk & ((1u << lane) - 1u));
    if (emit) kbuf[pos] = key;
    __syncthreads();
    if (tid == 0) {
      s_cnt = s_n;
      s_base = s_cnt ? atomicAdd(P.count + b, s_cnt) : 0;
      s_n = 0;
    }
    __syncthreads();
    if (tid < s_cnt) {
      const int slot = s_base + tid;
      if (slot < P.cap) P.keys[(size_t)b * P.cap + slot] = kbuf[tid];
      else if (P.overflow) *P.overflow = 1;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// k_filter_bulk: best-class filter over decoded predictions (the stand-alone yb_nms, e.g. the 100 000-row stress
// config) as a bulk-copy pipeline.  (B, A, no) fp32 is dense, so a tile of kBkRows rows is ONE contiguous span of
// kBkRows*no*4 bytes: thread 0 streams tiles into an S-stage shared-memory ring with cp.async.bulk (completion on
// an mbarrier per stage), S-1 tiles are always in flight per SM while the 256 threads walk one row each of the
// current tile (dense pitch `no` must be odd -> conflict-free).  The class scan keeps four independent running
// maxima (classes c mod 4) merged by (larger conf, then smaller index) = the first index of the maximum.
// Needs pred 16-byte aligned, A*no % 4 == 0 (every tile start / size is then a multiple of 16 bytes) and odd no;
// anything else takes k_filter_best.
// ------------------------------------------------------------------------------------------------
constexpr int kBkRows = 256;

__device__ __forceinline__ uint32_t nms_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bk_mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bk_issue(uint32_t bar, uint32_t dst, const void* src, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

__global__ void __launch_bounds__(kBkRows) k_filter_bulk(const FilterParams P, int stages, int tile_floats) {
  extern __shared__ __align__(128) unsigned char bk_smem[];
  float* ring = reinterpret_cast<float*>(bk_smem);             // [stages][tile_floats]
  __shared__ __align__(8) uint64_t full[8];
  __shared__ uint64_t kbuf[kBkRows];
  __shared__ int s_n, s_cnt, s_base;
  const int tid = threadIdx.x, lane = tid & 31;
  const int b = blockIdx.x / P.blocks_per_img, blk = blockIdx.x - b * P.blocks_per_img;
  const int row_begin = blk * P.rows_per_block;
  const int row_end = min(P.A, row_begin + P.rows_per_block);
  const int ntiles = (row_end - row_begin + kBkRows - 1) / kBkRows;
  const float* img = P.pred + (size_t)b * P.A * P.no;
  if (tid == 0) {
    for (int i = 0; i < stages; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(nms_smem_u32(&full[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    s_n = 0;
    for (int i = 0; i < stages && i < ntiles; ++i) {
      const int t0 = row_begin + i * kBkRows, rows = min(kBkRows, row_end - t0);
      bk_issue(nms_smem_u32(&full[i]), nms_smem_u32(ring + (size_t)i * tile_floats), img + (size_t)t0 * P.no,
               (uint32_t)rows * P.no * 4u);
    }
  }
  __syncthreads();
  const int nc = P.nc, no = P.no;
  for (int i = 0; i < ntiles; ++i) {
    const int st = i % stages;
    const int t0 = row_begin + i * kBkRows, t1 = min(row_end, t0 + kBkRows);
    bk_mbar_wait(nms_smem_u32(&full[st]), (uint32_t)(i / stages) & 1u);
    const int row = t0 + tid;
    bool emit = false;
    unsigned long long key = 0ull;
    if (row < t1) {
      const float* r = ring + (size_t)st * tile_floats + (size_t)tid * no;
      const float obj = r[4];
      if (obj > P.conf_thres) {                 // general.py:637,653
        float bv[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        int bi[4] = {0, 1, 2, 3};
        const float* cl = r + 5;
        int c = 0;
        for (; c + 4 <= nc; c += 4) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const float conf = __fmul_rn(cl[c + u], obj);   // general.py:673
            if (conf > bv[u]) { bv[u] = conf; bi[u] = c + u; }
          }
        }
        for (int u = 0; c + u < nc; ++u) {
          const float conf = __fmul_rn(cl[c + u], obj);
          if (conf > bv[u]) { bv[u] = conf; bi[u] = c + u; }
        }
        float best = bv[0];
        int best_c = bi[0];
#pragma unroll
        for (int u = 1; u < 4; ++u) {
          if (bv[u] > best || (bv[u] == best && bi[u] < best_c)) { best = bv[u]; best_c = bi[u]; }
        }
        if (best == -INFINITY) best_c = 0;      // no class beat -inf (all NaN): argmax of the reference's scan = 0
        if (best > P.conf_thres && class_allowed(P, best_c)) {   // general.py:684,688
          emit = true;
          key = ((unsigned long long)(~__float_as_uint(best)) << 32) | (uint32_t)row;
          P.cls_of_row[(size_t)b * P.A + row] = (uint16_t)best_c;
        }
      }
    }
    __syncthreads();                              // every thread is done reading stage st
    if (tid == 0 && i + stages < ntiles) {
      const int n0 = row_begin + (i + stages) * kBkRows, rows = min(kBkRows, row_end - n0);
      bk_issue(nms_smem_u32(&full[st]), nms_smem_u32(ring + (size_t)st * tile_floats), img + (size_t)n0 * no,
               (uint32_t)rows * no * 4u);
    }
    const unsigned mk = __ballot_sync(0xffffffffu, emit);
    int pos = 0;
    if (lane == 0 && mk) pos = atomicAdd(&s_n, __popc(mk));
    pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(mk & ((1u << lane) - 1u));
    if (emit) kbuf[pos] = key;
    __syncthreads();
    if (tid == 0) {                               // s_n is re-armed here: nobody touches it until the next ballot
      s_cnt = s_n;
      s_base = s_cnt ? atomicAdd(P.count + b, s_cnt) : 0;
      s_n = 0;
    }
    __syncthreads();
    if (tid < s_cnt) {
      const int slot = s_base + tid;
      if (slot < P.cap) P.keys[(size_t)b * P.cap + slot] = kbuf[tid];
      else if (P.overflow) *P.overflow = 1;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// k_filter_cells: fused best-class filter reading the head maps cell-major.  A head level stores, per grid cell,
// the na*no logits of all its anchors contiguously (c_stride values per cell), so a tile of consecutive cells is
// ONE contiguous span: it is staged with 16-byte coalesced loads (cell pitch c_stride + 1 in shared memory keeps
// the thread-per-row walk conflict-free) and thread (a, cell) handles prediction row row_off + a*ny*nx + cell.
// Block -> (image, level, cell range) through lv[].blk_off.
// ------------------------------------------------------------------------------------------------
template <typename TL> struct Vec4Load;
template <> struct Vec4Load<float> {
  static __device__ __forceinline__ float4 load(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
};
template <> struct Vec4Load<__nv_bfloat16> {
  static __device__ __forceinline__ float4 load(const __nv_bfloat16* p) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), c = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
    return make_float4(__low2float(a), __high2float(a), __low2float(c), __high2float(c));
  }
};

// Multi-label mode (ML, general.py:679-681: every (row, class) with conf > conf_thres is a candidate) keeps the
// thread-per-row walk.  A class can only be a candidate if fl(sigmoid(l_c) * obj) > conf_thres, so per row a logit gate
// g is derived from logit(conf_thres / obj) and PROVEN safe before use: with s_hi = sigmoid_rn_(g) stepped up eight ulps
// (sigmoid_rn_ is monotone up to a few ulps) the test fl(s_hi * obj) <= conf_thres implies that no logit below g can
// pass; classes below the gate are dismissed four per branch without evaluating their sigmoid, the others are
// evaluated exactly as the reference does.  If the proof fails the gate is lowered once and then dropped (full scan).
// Rows of a warp emit different numbers of keys: the 32 lanes walk the classes in lock step and compact each class's
// keys with a ballot into a per-warp buffer that is flushed with one global atomic per ~200 keys.
constexpr int kMlBuf = 256;          // keys buffered per warp between flushes (a class step adds at most 4 * 32)

template <typename TL, bool ML>
__global__ void __launch_bounds__(kFbThreads) k_filter_cells(const FilterParams P) {
  extern __shared__ __align__(16) unsigned char fb_smem[];
  float* tile = reinterpret_cast<float*>(fb_smem);             // [cells_per_tile][c_stride + 1]
  __shared__ uint64_t kbuf[ML ? (kFbThreads / 32) * kMlBuf : kFbThreads];
  __shared__ int s_n, s_base;
  const int tid = threadIdx.x, lane = tid & 31;
  const int b = blockIdx.x / P.blocks_per_img;
  int blk = blockIdx.x - b * P.blocks_per_img;
  int l = 0;
  while (l + 1 < P.nl && blk >= P.lv[l + 1].blk_off) ++l;
  blk -= P.lv[l].blk_off;
  const HeadLevelDev& L = P.lv[l];
  const int ncells = L.ny * L.nx;
  const int cpt = kFbThreads / P.na;                           // cells per tile (thread = (anchor, cell))
  const int cell_begin = blk * P.cells_per_block;
  const int cell_end = min(ncells, cell_begin + P.cells_per_block);
  const int pitch = L.c_stride + 1;
  const TL* lbase = reinterpret_cast<const TL*>(L.logits) + (size_t)b * ncells * L.c_stride;
  const int a = tid / cpt, cl = tid - a * cpt;
  for (int c0 = cell_begin; c0 < cell_end; c0 += cpt) {
    const int nc_tile = min(cpt, cell_end - c0);
    if (tid == 0) s_n = 0;
    // ---- stage nc_tile * c_stride contiguous values
    const TL* src = lbase + (size_t)c0 * L.c_stride;
    const int nvec = nc_tile * L.c_stride / 4;
    for (int i0 = 0; i0 < nvec; i0 += 8 * kFbThreads) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * kFbThreads + tid;
        if (i < nvec) v[u] = Vec4Load<TL>::load(src + 4 * (size_t)i);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * kFbThreads + tid;
        if (i < nvec) {
          const int e = 4 * i;
          // c_stride % 4 == 0: one cell per vector; e / c_stride by multiplication (e * c_stride < 2^32)
          const int cell = L.magic_cs ? (int)__umulhi((unsigned)e, L.magic_cs) : e, off = e - cell * L.c_stride;
          float* d = tile + cell * pitch + off;
          // lanes l, l+8, l+16, l+24 sit 32 words apart: each group of eight lanes stores its four values in an order
          // rotated by lane/8, so that the four store instructions are bank-conflict free
          const int rr = lane >> 3;
          float a0 = v[u].x, a1 = v[u].y, a2 = v[u].z, a3 = v[u].w;
          if (rr & 1) { const float t0 = a0; a0 = a1; a1 = a2; a2 = a3; a3 = t0; }
          if (rr & 2) { const float t0 = a0, t1 = a1; a0 = a2; a1 = a3; a2 = t0; a3 = t1; }
          d[rr & 3] = a0; d[(rr + 1) & 3] = a1; d[(rr + 2) & 3] = a2; d[(rr + 3) & 3] = a3;
        }
      }
    }
    __syncthreads();
    if (ML) {
      uint64_t* wbuf = kbuf + (tid >> 5) * kMlBuf;
      int n_w = 0;                                             // warp-uniform
      auto flush = [&]() {
        __syncwarp();
        int base = 0;
        if (lane == 0 && n_w) base = atomicAdd(P.count + b, n_w);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (int i = lane; i < n_w; i += 32) {
          const int slot = base + i;
          if (slot < P.cap) P.keys[(size_t)b * P.cap + slot] = wbuf[i];
          else if (P.overflow) *P.overflow = 1;
        }
        __syncwarp();
        n_w = 0;
      };
      const bool live = a < P.na && cl < nc_tile;
      const float* r = tile + (live ? cl * pitch + a * P.no : 0);
      const int cell = c0 + cl;
      const int row = L.row_off + a * ncells + cell;
      float obj = 0.f, gate = INFINITY;
      if (live) {
        obj = sigmoid_rn_(r[4]);
        if (obj > P.conf_thres) {                              // general.py:637,653
          const float q = __fdividef(P.conf_thres, obj);       // in (0, 1); only a starting point, the proof is below
          float g = fminf(__logf(__fdividef(q, 1.0f - q)), 30.0f) - 0.01f;
          gate = -INFINITY;
#pragma unroll
          for (int attempt = 0; attempt < 2; ++attempt) {
            const float s_hi = __uint_as_float(__float_as_uint(sigmoid_rn_(g)) + 8u);
            if (__fmul_rn(s_hi, obj) <= P.conf_thres) { gate = g; break; }
            g -= 0.25f;
          }
          if (!(P.conf_thres >= 0.f)) gate = -INFINITY;        // negative / NaN thresholds: no shortcut
        }
      }
      bool any_emit = false;
      for (int c = 0; c < P.nc; c += 4) {
        float v[4];
        bool pass = false;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          v[u] = (live && c + u < P.nc) ? r[5 + c + u] : -INFINITY;
          pass |= !(v[u] < gate);
        }
        if (!__any_sync(0xffffffffu, pass)) continue;
        if (n_w + 128 > kMlBuf) flush();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          bool e = false;
          float conf = 0.f;
          if (!(v[u] < gate)) {
            conf = __fmul_rn(sigmoid_rn_(v[u]), obj);          // general.py:673
            e = conf > P.conf_thres && class_allowed(P, c + u);  // general.py:680, 687-688
          }
          const unsigned m = __ballot_sync(0xffffffffu, e);
          if (e) {
            const uint32_t idx = (uint32_t)row * (uint32_t)P.nc + (uint32_t)(c + u);
            wbuf[n_w + __popc(m & ((1u << lane) - 1u))] = ((uint64_t)(~__float_as_uint(conf)) << 32) | idx;
            any_emit = true;
          }
          n_w += __popc(m);
        }
      }
      flush();
      if (any_emit) {
        const int gy = L.magic_nx ? (int)__umulhi((unsigned)cell, L.magic_nx) : cell;
        const int gx = cell - gy * L.nx;
        const float sx = sigmoid_rn_(r[0]), sy = sigmoid_rn_(r[1]), sw = sigmoid_rn_(r[2]), sh = sigmoid_rn_(r[3]);
        const float bx = __fmul_rn(__fadd_rn(__fsub_rn(__fmul_rn(sx, 2.f), 0.5f), (float)gx), L.stride);
        const float by = __fmul_rn(__fadd_rn(__fsub_rn(__fmul_rn(sy, 2.f), 0.5f), (float)gy), L.stride);
        const float w2 = __fmul_rn(sw, 2.f), h2 = __fmul_rn(sh, 2.f);
        *reinterpret_cast<float4*>(P.boxes + ((size_t)b * P.A + row) * 4) =
            make_float4(bx, by, __fmul_rn(__fmul_rn(w2, w2), L.anchor_wh[2 * a]), __fmul_rn(__fmul_rn(h2, h2), L.anchor_wh[2 * a + 1]));
      }
      __syncthreads();                                         // the tile is re-staged next
      continue;
    }
    bool emit = false;
    unsigned long long key = 0ull;
    if (a < P.na && cl < nc_tile) {
      const int cell = c0 + cl;
      const int gy = L.magic_nx ? (int)__umulhi((unsigned)cell, L.magic_nx) : cell;
      const int gx = cell - gy * L.nx;
      const int row = L.row_off + a * ncells + cell;
      emit = fb_process_row<true>(P, tile + cl * pitch + a * P.no, b, row, gx, gy, L.stride, L.anchor_wh[2 * a],
                                  L.anchor_wh[2 * a + 1], key);
    }
    const unsigned mk = __ballot_sync(0xffffffffu, emit);
    int pos = 0;
    if (lane == 0 && mk) pos = atomicAdd(&s_n, __popc(mk));
    pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(mk & ((1u << lane) - 1u));
    if (emit) kbuf[pos] = key;
    __syncthreads();
    if (tid == 0) s_base = s_n ? atomicAdd(P.count + b, s_n) : 0;
    __syncthreads();
    if (tid < s_n) {
      const int slot = s_base + tid;
      if (slot < P.cap) P.keys[(size_t)b * P.cap + slot] = kbuf[tid];
      else if (P.overflow) *P.overflow = 1;
    }
    __syncthreads();
  }
}

}  // namespace yb

namespace yb {

// ------------------------------------------------------------------------------------------------
// k_sort_nms: one CTA (1024 threads) per image.
// ------------------------------------------------------------------------------------------------
constexpr int kSortThreads = 1024;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kChunk = 256;           // candidates resolved per greedy round
constexpr int kChunkWords = kChunk / 32;

struct NmsParams {
  const float* boxsrc;   // xywh rows: pred (stride no) or fused boxes (stride 4)
  int box_stride;
  const uint16_t* cls_of_row;
  const int* count;
  uint64_t* keys_a;
  uint64_t* keys_b;
  int B, A, nc, cap;
  int multi_label, agnostic;
  int max_nms, max_det;
  float thr_f;           // largest float T with (double)T <= iou_thres: (double)iou > thr  <=>  iou > T
  int thr_nonneg;
  float* out;            // [B][max_det][6]
  int* counts_out;       // [B]
  int band_cap;          // keys per band (<= kBandCap): sized with the shared memory of the launch
  unsigned long long* dbg;
  int has_peer;          // yb_decode_nms_gather: detections also go straight into every peer's receive buffer
  yb_peer_gather peer;
};

__device__ __forceinline__ bool iou_suppresses(float ix1, float iy1, float ix2, float iy2, float iarea,
                                               float jx1, float jy1, float jx2, float jy2, float jarea,
                                               float thr_f, int thr_nonneg) {
  const float w = fmaxf(0.0f, __fsub_rn(fminf(ix2, jx2), fmaxf(ix1, jx1)));
  const float h = fmaxf(0.0f, __fsub_rn(fminf(iy2, jy2), fmaxf(iy1, jy1)));
  const float inter = __fmul_rn(w, h);
  if (thr_nonneg && inter == 0.0f) return false;  // iou is 0 or NaN: never > T when T >= 0
  const float ovr = __fdiv_rn(inter, __fsub_rn(__fadd_rn(iarea, jarea), inter));
  return ovr > thr_f;
}

// Candidates are consumed in bands of descending score: a band is picked by a radix select on the score bits
// (at least kBandMin keys, at most kBandCap so that it sorts inside shared memory), sorted, and fed to the greedy
// loop, which usually reaches max_det inside the first band.  Bands are disjoint key ranges taken in key order and
// every key is distinct, so the sequence of candidates is exactly that of one full sort (general.py:695-699).
#ifdef YB_NMS_TIMING
#define YB_TS(slot) do { if (threadIdx.x == 0 && (slot) < 64) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); P.dbg[blockIdx.x * 64 + (slot)] = t_; } } while (0)
#else
#define YB_TS(slot) do { } while (0)
#endif
constexpr int kBuckets = 128;            // class buckets of the kept set (k_sort_nms)
constexpr float kBucketExtent = 4000.0f; // < kMaxWH with room for the rounding of x + cls * 4096
__device__ __forceinline__ int ordered_int(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ordered_float(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }
constexpr int kBandMin = 1024;
constexpr int kBandCap = 4096;

// Stable LSD radix sort (8-bit digits) of n 64-bit keys by the whole block; src/dst may be shared or global.
// Returns the buffer holding the sorted keys.  whist: [kSortWarps][256] ints of scratch.
// `consider` masks the key bits that take part (digits outside it are skipped like constant digits).
__device__ uint64_t* block_radix_sort(uint64_t* src, uint64_t* dst, int n, int* whist,
                                      unsigned long long consider = ~0ull) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ unsigned long long s_and, s_or;
  __shared__ int digit_tot[256];
  __syncthreads();
  if (tid == 0) { s_and = ~0ull; s_or = 0ull; }
  __syncthreads();
  unsigned long long a = ~0ull, o = 0ull;
  for (int i = tid; i < n; i += (int)blockDim.x) { unsigned long long k = src[i]; a &= k; o |= k; }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    a &= __shfl_xor_sync(0xffffffffu, a, s);
    o |= __shfl_xor_sync(0xffffffffu, o, s);
  }
  if (lane == 0) { atomicAnd(&s_and, a); atomicOr(&s_or, o); }
  __syncthreads();
  const unsigned long long varying = (s_and ^ s_or) & consider;
  // contiguous segment of the array owned by each warp (keeps the scatter stable)
  const int per_warp = ((n + ((int)blockDim.x >> 5) - 1) / ((int)blockDim.x >> 5) + 31) / 32 * 32;
  const int seg0 = min(n, warp * per_warp), seg1 = min(n, seg0 + per_warp);
  for (int pass = 0; pass < 8 && n > 1; ++pass) {
    const int sh = pass * 8;
    if (((varying >> sh) & 0xffull) == 0) continue;  // digit identical for every key
    for (int i = tid; i < ((int)blockDim.x >> 5) * 256; i += (int)blockDim.x) whist[i] = 0;
    __syncthreads();
    int* myh = whist + warp * 256;
    for (int i = seg0 + lane; i - lane < seg1; i += 32) {
      const bool valid = i < seg1;
      const int d = valid ? (int)((src[i] >> sh) & 0xff) : 256 + lane;  // unique dummy digit per idle lane
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      if (valid && lane == __ffs(peers) - 1) myh[d] += __popc(peers);
      __syncwarp();
    }
    __syncthreads();
    // exclusive scan in (digit major, warp minor) order: thread t<256 owns digit t
    if (tid < 256) {
      int s = 0;
      for (int w = 0; w < ((int)blockDim.x >> 5); ++w) { int c = whist[w * 256 + tid]; whist[w * 256 + tid] = s; s += c; }
      digit_tot[tid] = s;
    }
    __syncthreads();
    if (warp == 0) {  // scan the 256 digit totals
      int carry = 0;
      for (int base = 0; base < 256; base += 32) {
        int v = digit_tot[base + lane], x = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, s); if (lane >= s) x += y; }
        digit_tot[base + lane] = carry + x - v;
        carry += __shfl_sync(0xffffffffu, x, 31);
      }
    }
    __syncthreads();
    if (tid < 256) {
      const int base = digit_tot[tid];
      for (int w = 0; w < ((int)blockDim.x >> 5); ++w) whist[w * 256 + tid] += base;
    }
    __syncthreads();
    for (int i = seg0 + lane; i - lane < seg1; i += 32) {
      const bool valid = i < seg1;
      unsigned long long k = valid ? src[i] : 0ull;
      const int d = valid ? (int)((k >> sh) & 0xff) : 256 + lane;
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      int pos = 0;
      if (valid) pos = myh[d] + __popc(peers & ((1u << lane) - 1u));
      __syncwarp();
      if (valid && lane == __ffs(peers) - 1) myh[d] += __popc(peers);
      __syncwarp();
      if (valid) dst[pos] = k;
    }
    __syncthreads();
    uint64_t* t = src; src = dst; dst = t;
  }
  __syncthreads();
  return src;
}

// Radix select on the high key word h = ~score_bits (ascending h = descending score): the smallest T such that
// the keys with lo < h <= T (all keys with h <= T when !have_lo) number at least kBandMin -- refined over up to
// three digit levels (12 + 12 + 8 bits) while that count exceeds kBandCap.  If fewer than kBandMin keys are left,
// T = 0xffffffff takes them all.  Returns the band size through s_out[0] and T through s_out[1].
// hist: 4096 ints of scratch.
__device__ void select_band(const uint64_t* keys, int n, uint32_t lo, bool have_lo, int* hist, uint32_t* s_out,
                            int band_cap) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ int s_wsum[kSortWarps];
  __shared__ int s_found, s_bin, s_cum_in, s_cum_ex, s_total;
  uint32_t prefix = 0;
  int pbits = 0, below = 0;
  for (int level = 0; level < 3; ++level) {
    const int nb = level < 2 ? 12 : 8;
    const int shift = level == 0 ? 20 : level == 1 ? 8 : 0;
    const int nbins = 1 << nb;
    __syncthreads();
    for (int i = tid; i < nbins; i += (int)blockDim.x) hist[i] = 0;
    if (tid == 0) s_found = 0;
    __syncthreads();
    // four keys per thread in flight (the keys come from L2: one dependent load per iteration would serialise)
    for (int i0 = tid; i0 < n; i0 += 4 * (int)blockDim.x) {
      uint32_t hv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * (int)blockDim.x;
        hv[u] = i < n ? (uint32_t)(keys[i] >> 32) : 0u;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * (int)blockDim.x;
        const uint32_t h = hv[u];
        if (i >= n) continue;
        if (have_lo && h <= lo) continue;
        if (pbits && (h >> (32 - pbits)) != prefix) continue;
        atomicAdd(&hist[(h >> shift) & (uint32_t)(nbins - 1)], 1);
      }
    }
    __syncthreads();
    // block scan: thread t owns bins [per*t, per*t + per)
    const int per = nbins / (int)blockDim.x > 0 ? nbins / (int)blockDim.x : 1;
    int loc[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // per <= 4096 / 512
    int sum = 0;
    if (tid * per < nbins) {
      for (int q = 0; q < per; ++q) { loc[q] = hist[tid * per + q]; sum += loc[q]; }
    }
    int incl = sum;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += y; }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int v = lane < ((int)blockDim.x >> 5) ? s_wsum[lane] : 0, x = v;
#pragma unroll
      for (int s = 1; s < 32; s <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, s); if (lane >= s) x += y; }
      s_wsum[lane] = x - v;
      if (lane == 31) s_total = x;
    }
    __syncthreads();
    int c = below + s_wsum[warp] + incl - sum;   // eligible keys ordered before this thread's first bin
    if (tid * per < nbins) {
      for (int q = 0; q < per; ++q) {
        const int cn = c + loc[q];
        if (c < kBandMin && cn >= kBandMin) { s_found = 1; s_bin = tid * per + q; s_cum_in = cn; s_cum_ex = c; }
        c = cn;
      }
    }
    __syncthreads();
    if (!s_found) {   // fewer than kBandMin keys left under this prefix (only possible at level 0): take all
      if (tid == 0) { s_out[0] = (uint32_t)(below + s_total); s_out[1] = 0xffffffffu; }
      __syncthreads();
      return;
    }
    const uint32_t j = (uint32_t)s_bin;
    if (s_cum_in <= band_cap || level == 2) {
      if (tid == 0) {
        s_out[0] = (uint32_t)s_cum_in;
        s_out[1] = (((prefix << nb) | j) << shift) | ((shift ? (1u << shift) : 1u) - 1u);
      }
      __syncthreads();
      return;
    }
    below = s_cum_ex;
    prefix = (prefix << nb) | j;
    pbits += nb;
  }
}

__global__ void __launch_bounds__(kSortThreads, 1) k_sort_nms(const NmsParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31;
  const int parts = (int)blockDim.x / kChunk, pshift = parts == 4 ? 2 : 1;   // threads per candidate in phases A/B
  int n = P.count[b];
  if (n > P.cap) n = P.cap;
  uint64_t* gkeys = P.keys_a + (size_t)b * P.cap;
  uint64_t* gtmp = P.keys_b + (size_t)b * P.cap;
  const int m_total = min(n, P.max_nms);  // general.py:698-699

  // =============================== greedy NMS ================================================
  float4* cbox = reinterpret_cast<float4*>(smem_raw);   // chunk: offset boxes (x1, y1, x2, y2), one 16-byte load per test
  float* kx1 = reinterpret_cast<float*>(cbox + kChunk);
  float* ky1 = kx1 + P.max_det;
  float* kx2 = ky1 + P.max_det;
  float* ky2 = kx2 + P.max_det;
  float* kar = ky2 + P.max_det;
  float* car = kar + P.max_det;
  float* craw = car + kChunk;    // [kChunk][4] un-offset xyxy
  float* cconf = craw + 4 * kChunk;
  float* ccls = cconf + kChunk;
  uint32_t* cmask = reinterpret_cast<uint32_t*>(ccls + kChunk);  // [kChunk][kChunkWords]
  uint32_t* kmeta = cmask + kChunk * kChunkWords;   // [max_det] class << 16 | next kept box of the same class bucket
  unsigned char* calive = reinterpret_cast<unsigned char*>(kmeta + P.max_det);
  __shared__ uint32_t s_alive[kChunkWords], s_new[kChunkWords];
  __shared__ int s_K, s_cnt, s_longrun;
  __shared__ uint32_t s_sel[2];
  // Class buckets: with the 4096-px class offset (general.py:702) two boxes of different classes cannot intersect as
  // long as the un-offset x extent of everything seen so far stays below kBucketExtent, so a candidate only needs the
  // kept boxes of its own class.  Kept boxes are chained per bucket (class & 127); the mode is dropped for good (full
  // scan of the kept set, always correct) if the extent grows beyond the limit, for agnostic NMS, or a negative threshold.
  __shared__ int s_khead[kBuckets], s_kcnt[kBuckets];
  __shared__ int s_bucket_ok, s_xmin, s_xmax;
  const size_t nms_bytes = (size_t)((calive + kChunk) - smem_raw + 15) / 16 * 16;
  uint64_t* band_a = reinterpret_cast<uint64_t*>(smem_raw + nms_bytes);
  const int band_cap = P.band_cap;
  uint64_t* band_b = band_a + band_cap;
  int* scratch = reinterpret_cast<int*>(band_b + band_cap);   // radix histograms [((int)blockDim.x >> 5)][256] / select [4096]
  if (tid == 0) {
    s_K = 0;
    s_bucket_ok = (P.thr_nonneg && !P.agnostic && P.max_det < 0xffff) ? 1 : 0;
    s_xmin = 0x7fffffff;
    s_xmax = (int)0x80000000;
  }
  if (tid < kBuckets) { s_khead[tid] = 0xffff; s_kcnt[tid] = 0; }
  __syncthreads();

  int consumed = 0;
  uint32_t lo = 0;
  bool have_lo = false;
  while (consumed < m_total) {
    if (s_K >= P.max_det) break;
    // ---- next band of candidates, sorted
    const uint64_t* src;
    int band_n;
    uint32_t T = 0xffffffffu;
    if (!have_lo && n <= band_cap) {
      band_n = n;
    } else {
      YB_TS(0);
      select_band(gkeys, n, lo, have_lo, scratch, s_sel, band_cap);
      YB_TS(1);
      band_n = (int)s_sel[0];
      T = s_sel[1];
    }
    if (band_n > band_cap) {
      // more equal scores than a band holds: sort everything once in global memory and walk the rest of it
      // (the `consumed` smallest keys are exactly the bands already processed)
      src = block_radix_sort(gkeys, gtmp, n, scratch) + consumed;
      band_n = n - consumed;
    } else {
      __syncthreads();
      if (tid == 0) s_cnt = 0;
      __syncthreads();
      for (int i0 = 0; i0 < n; i0 += 4 * (int)blockDim.x) {      // four keys per thread in flight
        unsigned long long kv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * (int)blockDim.x + tid;
          kv[u] = i < n ? gkeys[i] : 0ull;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * (int)blockDim.x + tid;
          const uint32_t h = (uint32_t)(kv[u] >> 32);
          const bool sel = i < n && (!have_lo || h > lo) && h <= T;
          const unsigned mk = __ballot_sync(0xffffffffu, sel);
          int pos = 0;
          if (lane == 0 && mk) pos = atomicAdd(&s_cnt, __popc(mk));
          pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(mk & ((1u << lane) - 1u));
          if (sel) band_a[pos] = kv[u];
        }
      }
      __syncthreads();
      YB_TS(2);
      // The band is sorted on the score word only (2-3 digit passes instead of 5); keys with equal scores -- rare, a
      // handful per band -- are then put into index order in place by the thread that owns the start of their run.
      // Runs longer than 32 (synthetic all-equal inputs) fall back to a full 64-bit stable sort of the band.
      {
        uint64_t* sorted = block_radix_sort(band_a, band_b, band_n, scratch, 0xffffffff00000000ull);
        if (tid == 0) s_longrun = 0;
        __syncthreads();
        for (int i = tid; i < band_n; i += (int)blockDim.x) {
          const uint32_t h = (uint32_t)(sorted[i] >> 32);
          if (i + 1 < band_n && (uint32_t)(sorted[i + 1] >> 32) == h && (i == 0 || (uint32_t)(sorted[i - 1] >> 32) != h)) {
            int e = i + 2;
            while (e < band_n && (uint32_t)(sorted[e] >> 32) == h) ++e;
            if (e - i > 32) {
              s_longrun = 1;
            } else {
              for (int a = i + 1; a < e; ++a) {          // insertion sort of the run on the full key (index order)
                const uint64_t k = sorted[a];
                int q = a - 1;
                while (q >= i && sorted[q] > k) { sorted[q + 1] = sorted[q]; --q; }
                sorted[q + 1] = k;
              }
            }
          }
        }
        __syncthreads();
        if (s_longrun) sorted = block_radix_sort(sorted, sorted == band_a ? band_b : band_a, band_n, scratch);
        src = sorted;
      }
      YB_TS(3);
      lo = T;
      have_lo = true;
    }
    const int m = min(band_n, m_total - consumed);

  for (int base = 0; base < m; base += kChunk) {
    YB_TS(4 + base / kChunk);
    const int K = s_K;
    if (K >= P.max_det) break;
    const int cnt = min(kChunk, m - base);
    // ---- load chunk: un-offset xyxy, class offset, offset boxes + area (general.py:676,702-703)
    int ex_lo = 0x7fffffff, ex_hi = (int)0x80000000;
    if (tid < cnt) {
      const unsigned long long key = src[base + tid];
      const uint32_t idx = (uint32_t)key;
      const float conf = __uint_as_float(~(uint32_t)(key >> 32));
      int row, cls;
      if (P.multi_label) { row = (int)(idx / (uint32_t)P.nc); cls = (int)(idx - (uint32_t)row * P.nc); }
      else { row = (int)idx; cls = P.cls_of_row[(size_t)b * P.A + row]; }
      const float* bp = P.boxsrc + ((size_t)b * P.A + row) * P.box_stride;
      const float x = bp[0], y = bp[1], hw = __fmul_rn(bp[2], 0.5f), hh = __fmul_rn(bp[3], 0.5f);  // w/2 exact
      const float x1 = __fsub_rn(x, hw), y1 = __fsub_rn(y, hh), x2 = __fadd_rn(x, hw), y2 = __fadd_rn(y, hh);
      const float c = P.agnostic ? 0.0f : __fmul_rn((float)cls, kMaxWH);
      const float ox1 = __fadd_rn(x1, c), oy1 = __fadd_rn(y1, c), ox2 = __fadd_rn(x2, c), oy2 = __fadd_rn(y2, c);
      cbox[tid] = make_float4(ox1, oy1, ox2, oy2);
      car[tid] = __fmul_rn(__fsub_rn(ox2, ox1), __fsub_rn(oy2, oy1));
      craw[4 * tid + 0] = x1; craw[4 * tid + 1] = y1; craw[4 * tid + 2] = x2; craw[4 * tid + 3] = y2;
      cconf[tid] = conf; ccls[tid] = (float)cls;
      if (x1 != x1 || x2 != x2) s_bucket_ok = 0;
      ex_lo = ordered_int(x1);
      ex_hi = ordered_int(x2);
    }
    if (tid < kChunk) {      // whole warps: one shared-memory atomic per warp instead of one per candidate
      ex_lo = __reduce_min_sync(0xffffffffu, ex_lo);
      ex_hi = __reduce_max_sync(0xffffffffu, ex_hi);
      if (lane == 0) { atomicMin(&s_xmin, ex_lo); atomicMax(&s_xmax, ex_hi); }
    }
    __syncthreads();
    if (tid == 0 && s_bucket_ok && !(__fsub_rn(ordered_float(s_xmax), ordered_float(s_xmin)) <= kBucketExtent)) s_bucket_ok = 0;
    __syncthreads();
    const bool bucket_mode = s_bucket_ok != 0;
    if (base == 0) YB_TS(50);
    // ---- phase A: candidate j vs every kept box (parts = blockDim / kChunk = 4 or 2 threads per candidate)
    {
      const int j = tid >> pshift, part = tid & (parts - 1);
      bool sup = false;
      if (j < cnt) {
        const float4 jb = cbox[j];
        const float jx1 = jb.x, jy1 = jb.y, jx2 = jb.z, jy2 = jb.w, ja = car[j];
        const uint32_t cj = (uint32_t)ccls[j];
        // every thread of the candidate walks the whole chain (dependent loads), so the chain only pays when it is
        // much shorter than this thread's share of the kept array (one dominant class -> array scan)
        if (bucket_mode && s_kcnt[cj & (kBuckets - 1)] * parts * 2 <= K) {
          int k = s_khead[cj & (kBuckets - 1)], t = 0;
          while (k != 0xffff && !sup) {
            const uint32_t meta = kmeta[k];
            if ((meta >> 16) == cj && (t++ & (parts - 1)) == part)
              sup = iou_suppresses(kx1[k], ky1[k], kx2[k], ky2[k], kar[k], jx1, jy1, jx2, jy2, ja, P.thr_f, P.thr_nonneg);
            k = (int)(meta & 0xffffu);
          }
        } else {
          for (int k = part; k < K && !sup; k += parts)
            sup = iou_suppresses(kx1[k], ky1[k], kx2[k], ky2[k], kar[k], jx1, jy1, jx2, jy2, ja, P.thr_f, P.thr_nonneg);
        }
      }
      unsigned s = sup ? 1u : 0u;
      s |= __shfl_xor_sync(0xffffffffu, s, 1);
      if (parts == 4) s |= __shfl_xor_sync(0xffffffffu, s, 2);
      if (part == 0) calive[j] = (j < cnt && !s) ? 1 : 0;
    }
    __syncthreads();
    if (base == 0) YB_TS(51);
    if (tid < kChunkWords) {
      uint32_t w = 0;
      for (int q = 0; q < 32; ++q) w |= (uint32_t)calive[tid * 32 + q] << q;
      s_alive[tid] = w;
      s_new[tid] = 0;
    }
    // ---- phase B: intra-chunk suppression bit matrix, rows of live candidates only.  A warp owns a row i; lane q tests
    // candidate j = w*32 + q of word w (its box comes from consecutive shared-memory words, row i's box is uniform) and the
    // warp ballot IS the matrix word.  Only words that hold some j > i are produced (phase C reads nothing else).
    {
      const int warp = tid >> 5, nwarps = (int)blockDim.x >> 5;
      for (int i = warp; i < cnt; i += nwarps) {
        if (!calive[i]) continue;
        const float4 ib = cbox[i];
        const float ix1 = ib.x, iy1 = ib.y, ix2 = ib.z, iy2 = ib.w, ia = car[i], ci = ccls[i];
        for (int w = i >> 5; w < kChunkWords; ++w) {
          const int j = w * 32 + lane;
          bool sup = false;
          if (j > i && j < cnt && calive[j] && (!bucket_mode || ccls[j] == ci)) {
            const float4 jb = cbox[j];
            sup = iou_suppresses(ix1, iy1, ix2, iy2, ia, jb.x, jb.y, jb.z, jb.w, car[j], P.thr_f, P.thr_nonneg);
          }
          const uint32_t word = __ballot_sync(0xffffffffu, sup);
          if (lane == 0) cmask[i * kChunkWords + w] = word;
        }
      }
    }
    __syncthreads();
    if (base == 0) YB_TS(52);
    // ---- phase C: resolve the chunk in order, warp 0 cooperatively, one 32-candidate word at a time.  Lane q holds row
    // (w*32 + q) of the bit matrix.  If no candidate of the word suppresses another one, all of them are kept at once;
    // otherwise the greedy order is walked with ffs + one shuffle per kept box (no shared-memory round trip on the
    // dependent chain).  The rows of the boxes kept in the word are OR-reduced across the warp into the removed masks
    // of the later words.
    if (tid < 32) {
      uint32_t removed[kChunkWords];
#pragma unroll
      for (int w = 0; w < kChunkWords; ++w) removed[w] = 0;
      int Kc = K;
#pragma unroll
      for (int w = 0; w < kChunkWords; ++w) {
        const uint32_t alive_w = s_alive[w];
        uint32_t rowv[kChunkWords];
        const bool live = (alive_w >> lane) & 1u;            // rows of dead candidates were never written
#pragma unroll
        for (int ww = 0; ww < kChunkWords; ++ww) rowv[ww] = (live && ww >= w) ? cmask[(w * 32 + lane) * kChunkWords + ww] : 0u;
        const uint32_t cand = alive_w & ~removed[w];
        const int room = P.max_det - Kc;
        uint32_t kept = 0;
        if (cand && room > 0) {
          const bool cq = (cand >> lane) & 1u;
          if (!__any_sync(0xffffffffu, cq && (rowv[w] & cand)) && __popc(cand) <= room) {
            kept = cand;
          } else {
            uint32_t rem = cand;
            int left = room;
            while (rem && left > 0) {
              const int q = __ffs(rem) - 1;
              kept |= 1u << q;
              --left;
              const uint32_t sup = __shfl_sync(0xffffffffu, rowv[w], q);
              rem &= ~sup & ~((2u << q) - 1u);
            }
          }
          Kc += __popc(kept);
        }
        if (lane == 0) s_new[w] = kept;
        const bool mine = (kept >> lane) & 1u;
#pragma unroll
        for (int ww = 0; ww < kChunkWords; ++ww)
          if (ww > w) removed[ww] |= __reduce_or_sync(0xffffffffu, mine ? rowv[ww] : 0u);
      }
      if (lane == 0) s_K = Kc;
    }
    __syncthreads();
    if (base == 0) YB_TS(53);
    // ---- phase D: append the newly kept boxes to the kept set and write output rows
    if (tid < cnt && ((s_new[tid >> 5] >> (tid & 31)) & 1u)) {
      int pos = K + __popc(s_new[tid >> 5] & ((1u << (tid & 31)) - 1u));
      for (int w = 0; w < (tid >> 5); ++w) pos += __popc(s_new[w]);
      const float4 nb = cbox[tid];
      kx1[pos] = nb.x; ky1[pos] = nb.y; kx2[pos] = nb.z; ky2[pos] = nb.w; kar[pos] = car[tid];
      if (bucket_mode) {
        const uint32_t cj = (uint32_t)ccls[tid];
        kmeta[pos] = (cj << 16) | (uint32_t)atomicExch(&s_khead[cj & (kBuckets - 1)], pos);
        atomicAdd(&s_kcnt[cj & (kBuckets - 1)], 1);
      }
      float* o = P.out + ((size_t)b * P.max_det + pos) * 6;
      o[0] = craw[4 * tid + 0]; o[1] = craw[4 * tid + 1]; o[2] = craw[4 * tid + 2]; o[3] = craw[4 * tid + 3];
      o[4] = cconf[tid]; o[5] = ccls[tid];
      if (P.has_peer) {   // the same row over NVLink into every peer's block for this rank (P2P stores; self included)
        const size_t roff = (((size_t)P.peer.rank * P.B + b) * P.max_det + pos) * 6;
        for (int q = 0; q < P.peer.n_peers; ++q) {
          float* po = P.peer.out[q] + roff;
          po[0] = o[0]; po[1] = o[1]; po[2] = o[2]; po[3] = o[3]; po[4] = o[4]; po[5] = o[5];
        }
      }
    }
    __syncthreads();
  }
    YB_TS(40);
    consumed += band_n;
  }
  if (tid == 0) P.counts_out[b] = s_K;
  if (P.has_peer) {
    // Completion: every CTA publishes its image, the LAST one to finish raises this rank's arrival flag on every peer
    // (fence.sys + release store), which is what yb_peer_wait polls.  Slot reuse is the caller's protocol (dist.py).
    __syncthreads();
    if (tid == 0) {
      for (int q = 0; q < P.peer.n_peers; ++q) P.peer.counts[q][P.peer.rank * P.B + b] = s_K;
      __threadfence_system();
      const int prev = atomicAdd(P.peer.done_counter, 1);
      if (prev == (int)gridDim.x - 1) {
        *P.peer.done_counter = 0;
        __threadfence_system();
        for (int q = 0; q < P.peer.n_peers; ++q)
          asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(P.peer.flags[q] + P.peer.rank), "r"(P.peer.step) : "memory");
      }
    }
  }
}

// One warp: lane r spins on this rank's arrival flag of rank r until it reaches `step` (acquire, system scope).
// A bounded spin (timeout_ns of globaltimer) reports through *status instead of hanging the device, and poisons the
// late sender's block of the receive buffer (counts = -1) so that stale rows can never be read as detections.
__global__ void k_peer_wait(const uint32_t* flags, int n_peers, uint32_t step, int* status, int* counts, int B,
                            unsigned long long timeout_ns) {
  const int r = threadIdx.x;
  if (r >= n_peers) return;
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + r) : "memory");
    if ((int32_t)(v - step) >= 0) break;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) {
      if (status) atomicOr(status, 1 << 8);
      if (counts) for (int i = 0; i < B; ++i) counts[r * B + i] = -1;
      break;
    }
    __nanosleep(200);
  }
}

// ------------------------------------------------------------------------------------------------
struct NmsWorkspace {
  int* count; uint64_t* keys_a; uint64_t* keys_b; uint16_t* cls_of_row; float* boxes;
  size_t total;
};

static NmsWorkspace carve(void* ws, int B, int A, int cap, bool fused) {
  NmsWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  char* p = (char*)ws;
  w.count = (int*)(p + take((size_t)B * 4));
  w.keys_a = (uint64_t*)(p + take((size_t)B * cap * 8));
  w.keys_b = (uint64_t*)(p + take((size_t)B * cap * 8));
  w.cls_of_row = (uint16_t*)(p + take((size_t)B * A * 2));
  w.boxes = (float*)(p + take(fused ? (size_t)B * A * 16 : 0));
  w.total = off;
  return w;
}

static int resolve_cap(int A, int nc, int multi_label, int cand_cap) {
  long long full = multi_label ? (long long)A * nc : A;
  if (cand_cap > 0 && cand_cap < full) return cand_cap;
  return (int)(full > 0x7fffffffLL ? 0x7fffffff : full);
}

static int nms_common(FilterParams& F, bool fused, int logits_dtype, double iou_thres, int agnostic, int max_nms,
                      int max_det, const int* classes, int n_classes, void* workspace, size_t workspace_bytes,
                      float* out, int* counts, int* overflow, cudaStream_t s, const yb_peer_gather* peer = nullptr) {
  const int B = F.B, A = F.A;
  if (peer) {
    YB_CHECK_ARG(peer->n_peers > 0 && peer->n_peers <= YB_MAX_PEERS && peer->rank >= 0 && peer->rank < peer->n_peers &&
                 peer->done_counter && peer->step > 0, "nms: bad peer-gather table");
    for (int q = 0; q < peer->n_peers; ++q)
      YB_CHECK_ARG(peer->out[q] && peer->counts[q] && peer->flags[q], "nms: peer %d has a NULL buffer", q);
  }
  YB_CHECK_ARG(B > 0 && A > 0 && F.no > 5, "nms: need no >= 6 (nc >= 1)");
  YB_CHECK_ARG(F.no <= 256, "nms: no > 256 unsupported");
  YB_CHECK_ARG(max_det > 0 && max_det <= 4096, "nms: max_det must be in [1,4096]");
  YB_CHECK_ARG(max_nms > 0, "nms: max_nms must be positive");
  YB_CHECK_ARG(out && counts && workspace, "nms: null output/workspace");
  YB_CHECK_ARG((long long)A * F.nc < 0xffffffffLL, "nms: A*nc must fit 32 bits");
  F.use_class_mask = 0;
  memset(F.class_mask, 0, sizeof(F.class_mask));
  if (classes) {  // general.py:687-688 (an empty list filters everything, like tensor([]) any() == False)
    F.use_class_mask = 1;
    for (int i = 0; i < n_classes; ++i) {
      if (classes[i] >= 0 && classes[i] < kMaxClassWords * 32) F.class_mask[classes[i] >> 5] |= 1u << (classes[i] & 31);
    }
  }
  NmsWorkspace W = carve(workspace, B, A, F.cap, fused);
  YB_CHECK_ARG((uintptr_t)workspace % 256 == 0, "nms: workspace must be 256-byte aligned");
  if (W.total > workspace_bytes) {
    set_error("nms: workspace too small (%zu < %zu)", workspace_bytes, W.total);
    return YB_ERR_WORKSPACE;
  }
  F.count = W.count; F.keys = W.keys_a; F.cls_of_row = W.cls_of_row; F.boxes = W.boxes; F.overflow = overflow;
  YB_CUDA(cudaMemsetAsync(W.count, 0, (size_t)B * 4, s));
  if (overflow) YB_CUDA(cudaMemsetAsync(overflow, 0, 4, s));
  {
    // ~8 resident blocks per SM over the whole batch, at least 64 rows per block
    int bpi = (num_sms() * 8 + B - 1) / B;
    if (bpi > (A + 63) / 64) bpi = (A + 63) / 64;
    if (bpi < 1) bpi = 1;
    F.rows_per_block = (A + bpi - 1) / bpi;
    F.blocks_per_img = (A + F.rows_per_block - 1) / F.rows_per_block;
    const int grid = B * F.blocks_per_img;
    const size_t fb_smem = (size_t)kFbThreads * (F.no | 1) * 4;
    bool cells_ok = fused && F.na >= 1 && F.na <= 8;
    int max_cs = 0;
    for (int l = 0; cells_ok && l < F.nl; ++l) {
      const HeadLevelDev& L = F.lv[l];
      const size_t es = L.dtype == YB_F32 ? 4 : 2;
      cells_ok = L.c_stride % 4 == 0 && L.c_stride >= F.na * F.no && (uintptr_t)L.logits % 16 == 0 &&
                 ((size_t)L.c_stride * es) % 16 == 0;
      if (L.c_stride > max_cs) max_cs = L.c_stride;
    }
    const int cpt = cells_ok ? kFbThreads / F.na : 0;
    const size_t cells_smem = cells_ok ? (size_t)cpt * (max_cs + 1) * 4 : 0;
    if (cells_ok && cells_smem <= 100 * 1024) {
      // cells per block: ~8 resident blocks per SM over the batch, whole tiles
      long long total_cells = 0;
      for (int l = 0; l < F.nl; ++l) total_cells += (long long)F.lv[l].ny * F.lv[l].nx;
      long long want = (total_cells * B + (long long)num_sms() * 8 - 1) / ((long long)num_sms() * 8);
      int cpb = (int)((want + cpt - 1) / cpt) * cpt;
      if (cpb < cpt) cpb = cpt;
      F.cells_per_block = cpb;
      int off = 0;
      for (int l = 0; l < F.nl; ++l) {
        F.lv[l].blk_off = off;
        off += (F.lv[l].ny * F.lv[l].nx + cpb - 1) / cpb;
      }
      F.blocks_per_img = off;
      const int grid3 = B * off;
#define YB_CELLS_LAUNCH(T, ML)                                                                                            \
      do {                                                                                                               \
        YB_CUDA(cudaFuncSetAttribute(k_filter_cells<T, ML>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cells_smem)); \
        k_filter_cells<T, ML><<<grid3, kFbThreads, cells_smem, s>>>(F);                                                  \
      } while (0)
      if (logits_dtype == YB_F32) { if (F.multi_label) YB_CELLS_LAUNCH(float, true); else YB_CELLS_LAUNCH(float, false); }
      else { if (F.multi_label) YB_CELLS_LAUNCH(__nv_bfloat16, true); else YB_CELLS_LAUNCH(__nv_bfloat16, false); }
#undef YB_CELLS_LAUNCH
    } else if (!fused && !F.multi_label && F.nc > 1 && (F.no & 1) && ((long long)A * F.no) % 4 == 0 &&
               (uintptr_t)F.pred % 16 == 0 && (size_t)kBkRows * F.no * 4 * 2 <= 200 * 1024 &&
               getenv("YB_NMS_NO_BULK") == nullptr) {
      // decoded predictions, best class: bulk-copy ring (k_filter_bulk)
      const int tile_floats = kBkRows * F.no;
      int stages = (int)((200 * 1024) / ((size_t)tile_floats * 4));
      if (stages > 4) stages = 4;
      const int tiles_img = (A + kBkRows - 1) / kBkRows;
      int bpi3 = (num_sms() * 8 + B - 1) / B;                 // ~8 blocks per SM over the batch
      if (bpi3 > (tiles_img + 1) / 2) bpi3 = (tiles_img + 1) / 2;   // at least two tiles per block
      if (bpi3 < 1) bpi3 = 1;
      F.rows_per_block = ((tiles_img + bpi3 - 1) / bpi3) * kBkRows;
      F.blocks_per_img = (A + F.rows_per_block - 1) / F.rows_per_block;
      const size_t bk_bytes = (size_t)stages * tile_floats * 4;
      YB_CUDA(cudaFuncSetAttribute(k_filter_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bk_bytes));
      k_filter_bulk<<<B * F.blocks_per_img, kBkRows, bk_bytes, s>>>(F, stages, tile_floats);
    } else if (!F.multi_label && fb_smem <= 160 * 1024) {
      // best-class mode: one thread per row over a staged tile
      int bpi2 = (num_sms() * 8 + B - 1) / B;
      if (bpi2 > (A + kFbThreads - 1) / kFbThreads) bpi2 = (A + kFbThreads - 1) / kFbThreads;
      if (bpi2 < 1) bpi2 = 1;
      F.rows_per_block = ((A + bpi2 - 1) / bpi2 + kFbThreads - 1) / kFbThreads * kFbThreads;
      F.blocks_per_img = (A + F.rows_per_block - 1) / F.rows_per_block;
      const int grid2 = B * F.blocks_per_img;
      const int nvb = (F.no + 31) / 32;
#define YB_FB_LAUNCH_T(FU, T, NVV)                                                                                     \
      do {                                                                                                             \
        YB_CUDA(cudaFuncSetAttribute(k_filter_best<FU, T, NVV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_smem)); \
        k_filter_best<FU, T, NVV><<<grid2, kFbThreads, fb_smem, s>>>(F);                                               \
      } while (0)
#define YB_FB_LAUNCH(NVV)                                                     \
      do {                                                                    \
        if (!fused) YB_FB_LAUNCH_T(false, float, NVV);                        \
        else if (logits_dtype == YB_F32) YB_FB_LAUNCH_T(true, float, NVV);    \
        else YB_FB_LAUNCH_T(true, __nv_bfloat16, NVV);                        \
      } while (0)
      if (nvb <= 1) YB_FB_LAUNCH(1);
      else if (nvb == 2) YB_FB_LAUNCH(2);
      else if (nvb == 3) YB_FB_LAUNCH(3);
      else if (nvb == 4) YB_FB_LAUNCH(4);
      else YB_FB_LAUNCH(8);
#undef YB_FB_LAUNCH
#undef YB_FB_LAUNCH_T
    } else {
      const int nv = (F.no + 31) / 32;
#define YB_FILTER_LAUNCH(NVV)                                                              \
      do {                                                                                   \
        if (!fused) k_filter<false, float, NVV><<<grid, 256, 0, s>>>(F);                     \
        else if (logits_dtype == YB_F32) k_filter<true, float, NVV><<<grid, 256, 0, s>>>(F); \
        else k_filter<true, __nv_bfloat16, NVV><<<grid, 256, 0, s>>>(F);                     \
      } while (0)
      if (nv <= 1) YB_FILTER_LAUNCH(1);
      else if (nv == 2) YB_FILTER_LAUNCH(2);
      else if (nv == 3) YB_FILTER_LAUNCH(3);
      else if (nv == 4) YB_FILTER_LAUNCH(4);
      else YB_FILTER_LAUNCH(8);
#undef YB_FILTER_LAUNCH
    }
    YB_LAUNCH_CHECK();
  }
  NmsParams N;
  N.boxsrc = fused ? W.boxes : F.pred;
  N.box_stride = fused ? 4 : F.no;
  N.cls_of_row = W.cls_of_row; N.count = W.count; N.keys_a = W.keys_a; N.keys_b = W.keys_b;
  N.B = B; N.A = A; N.nc = F.nc; N.cap = F.cap; N.multi_label = F.multi_label; N.agnostic = agnostic;
  N.max_nms = max_nms; N.max_det = max_det;
  float T = (float)iou_thres;  // largest float with (double)T <= iou_thres
  if ((double)T > iou_thres) T = nextafterf(T, -INFINITY);
  N.thr_f = T; N.thr_nonneg = (T >= 0.0f) ? 1 : 0;
  N.out = out; N.counts_out = counts;
  N.dbg = reinterpret_cast<unsigned long long*>(W.keys_b);
  N.has_peer = peer ? 1 : 0;
  if (peer) N.peer = *peer; else memset(&N.peer, 0, sizeof(N.peer));
  size_t sm_nms = (size_t)max_det * 6 * 4 + (size_t)kChunk * (5 + 4 + 2) * 4 + (size_t)kChunk * kChunkWords * 4 + kChunk;
  // More images than SMs: 512-thread CTAs with 3072-key bands fit two per SM, so the whole batch is resident at once
  // and one CTA's serial phases hide behind the other's; otherwise one 1024-thread CTA per image.
  int threads = kSortThreads;
  N.band_cap = kBandCap;
  static const char* force = getenv("YB_NMS_SORT_THREADS");
  const size_t smem_half = (sm_nms + 15) / 16 * 16 + 2 * (size_t)3072 * 8 + (size_t)4096 * 4;
  if ((force ? atoi(force) == 512 : B > num_sms()) && smem_half <= 112 * 1024) {
    threads = 512;
    N.band_cap = 3072;
  }
  const size_t hist_bytes = (size_t)(threads / 32) * 256 * 4 > 4096 * 4 ? (size_t)(threads / 32) * 256 * 4 : 4096 * 4;
  size_t smem = (sm_nms + 15) / 16 * 16 + 2 * (size_t)N.band_cap * 8 + hist_bytes;
  YB_CUDA(cudaFuncSetAttribute(k_sort_nms, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_sort_nms<<<B, threads, smem, s>>>(N);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

}  // namespace yb

using namespace yb;

extern "C" size_t yb_nms_workspace_bytes(int B, int A, int no, int multi_label, int cand_cap) {
  if (B <= 0 || A <= 0 || no <= 5) return 0;
  multi_label = multi_label && (no - 5 > 1);
  int cap = resolve_cap(A, no - 5, multi_label, cand_cap);
  return carve(nullptr, B, A, cap, true).total;
}

extern "C" int yb_nms(const float* pred, int B, int A, int no, float conf_thres, double iou_thres,
                      const int* classes, int n_classes, int agnostic, int multi_label, int max_nms, int max_det,
                      int cand_cap, void* workspace, size_t workspace_bytes, float* out, int* counts,
                      int* overflow, void* stream) {
  YB_CHECK_ARG(pred, "yb_nms: null pred");
  FilterParams F;
  memset(&F, 0, sizeof(F));
  F.pred = pred; F.B = B; F.A = A; F.no = no; F.nc = no - 5; F.conf_thres = conf_thres;
  F.multi_label = (multi_label && F.nc > 1) ? 1 : 0;  // general.py:645
  F.cap = resolve_cap(A, F.nc, F.multi_label, cand_cap);
  return nms_common(F, false, YB_F32, iou_thres, agnostic, max_nms, max_det, classes, n_classes, workspace,
                    workspace_bytes, out, counts, overflow, (cudaStream_t)stream);
}

// ---- yb_nms_static: yb_nms with its workspace in library-owned static device memory (no allocation anywhere).
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The array is part of the module
// image: every GPU that loads the library holds kStaticWsBytes until exit.  All calls share it and it keeps the last
// call's contents -- nms_common clears the per-image candidate counters (the only part that must start at zero) on
// every call; calls that overlap on the GPU (two streams, two host threads) race on it.
constexpr size_t kStaticWsBytes = (size_t)16 << 20;      // 16 MiB: e.g. 32 images x 25 200 rows in best-class mode
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];

// In host code `g_static_ws` is only a host-side stand-in for the symbol, NOT a device address: the device address is
// looked up once per device (with that device current) and cached.
static int static_ws_ptr(void** out) {
  constexpr int kMaxDev = 64;
  static void* cache[kMaxDev] = {nullptr};
  int dev = 0;
  YB_CUDA(cudaGetDevice(&dev));
  YB_CHECK_ARG(dev >= 0 && dev < kMaxDev, "yb_nms_static: device ordinal %d out of range", dev);
  if (!cache[dev]) {
    void* p = nullptr;
    YB_CUDA(cudaGetSymbolAddress(&p, g_static_ws));
    YB_CHECK_ARG(p != nullptr && (uintptr_t)p % 256 == 0, "yb_nms_static: bad symbol address");
    cache[dev] = p;
  }
  *out = cache[dev];
  return YB_OK;
}

extern "C" size_t yb_nms_static_capacity(void) { return kStaticWsBytes; }

extern "C" int yb_nms_static(const float* pred, int B, int A, int no, float conf_thres, double iou_thres,
                             const int* classes, int n_classes, int agnostic, int multi_label, int max_nms, int max_det,
                             float* out, int* counts, int* overflow, void* stream) {
  YB_CHECK_ARG(pred, "yb_nms_static: null pred");
  YB_CHECK_ARG(B > 0 && A > 0 && no > 5, "yb_nms_static: need B, A > 0 and no >= 6");
  FilterParams F;
  memset(&F, 0, sizeof(F));
  F.pred = pred; F.B = B; F.A = A; F.no = no; F.nc = no - 5; F.conf_thres = conf_thres;
  F.multi_label = (multi_label && F.nc > 1) ? 1 : 0;  // general.py:645
  F.cap = resolve_cap(A, F.nc, F.multi_label, 0);
  // size check in 64-bit host arithmetic BEFORE anything is launched
  const double need_est = (double)B * 4 + 2.0 * (double)B * (double)F.cap * 8 + (double)B * (double)A * 2 + 4 * 256;
  if (need_est > (double)kStaticWsBytes || carve(nullptr, B, A, F.cap, false).total > kStaticWsBytes) {
    set_error("yb_nms_static: B=%d A=%d no=%d multi_label=%d needs %.0f bytes of scratch, the static array holds %zu "
              "(use yb_nms with a caller workspace)", B, A, no, F.multi_label, need_est, kStaticWsBytes);
    return YB_ERR_WORKSPACE;
  }
  void* ws = nullptr;
  const int rc = static_ws_ptr(&ws);
  if (rc) return rc;
  return nms_common(F, false, YB_F32, iou_thres, agnostic, max_nms, max_det, classes, n_classes, ws, kStaticWsBytes, out,
                    counts, overflow, (cudaStream_t)stream);
}

extern "C" int yb_decode_nms_gather(const yb_head_level* levels, int nl, int na, int no, float conf_thres,
                                    double iou_thres, const int* classes, int n_classes, int agnostic, int multi_label,
                                    int max_nms, int max_det, int cand_cap, void* workspace, size_t workspace_bytes,
                                    float* out, int* counts, int* overflow, const yb_peer_gather* peer, void* stream) {
  YB_CHECK_ARG(levels && nl > 0 && nl <= YB_MAX_LEVELS && na > 0 && na <= YB_MAX_ANCHORS, "yb_decode_nms: bad levels");
  FilterParams F;
  memset(&F, 0, sizeof(F));
  int A = 0;
  const int B = levels[0].logits.N;
  const int dt = levels[0].logits.dtype;
  for (int l = 0; l < nl; ++l) {
    const yb_tensor& t = levels[l].logits;
    YB_CHECK_ARG(view_ok(t) && t.N == B && t.C >= na * no && t.dtype == dt, "yb_decode_nms: level %d invalid", l);
    YB_CHECK_ARG(dt == YB_F32 || dt == YB_BF16, "yb_decode_nms: logits must be fp32/bf16");
    HeadLevelDev& L = F.lv[l];
    L.logits = t.data; L.dtype = t.dtype; L.ny = t.H; L.nx = t.W; L.c_stride = t.c_stride;
    L.stride = levels[l].stride;
    memcpy(L.anchor_wh, levels[l].anchor_wh, sizeof(L.anchor_wh));
    L.row_off = A;
    L.magic_cells = t.H * t.W > 1 ? (unsigned)(0x100000000ULL / (unsigned long long)(t.H * t.W)) + 1u : 0u;
    L.magic_nx = t.W > 1 ? (unsigned)(0x100000000ULL / (unsigned long long)t.W) + 1u : 0u;
    L.magic_cs = t.c_stride > 1 ? (unsigned)(0x100000000ULL / (unsigned long long)t.c_stride) + 1u : 0u;
    YB_CHECK_ARG((long long)na * t.H * t.W * t.H * t.W < 0x100000000LL, "yb_decode_nms: level %d too large", l);
    A += na * t.H * t.W;
  }
  F.nl = nl; F.na = na; F.B = B; F.A = A; F.no = no; F.nc = no - 5; F.conf_thres = conf_thres;
  F.multi_label = (multi_label && F.nc > 1) ? 1 : 0;
  F.cap = resolve_cap(A, F.nc, F.multi_label, cand_cap);
  return nms_common(F, true, dt, iou_thres, agnostic, max_nms, max_det, classes, n_classes, workspace,
                    workspace_bytes, out, counts, overflow, (cudaStream_t)stream, peer);
}

extern "C" int yb_decode_nms(const yb_head_level* levels, int nl, int na, int no, float conf_thres,
                             double iou_thres, const int* classes, int n_classes, int agnostic, int multi_label,
                             int max_nms, int max_det, int cand_cap, void* workspace, size_t workspace_bytes,
                             float* out, int* counts, int* overflow, void* stream) {
  return yb_decode_nms_gather(levels, nl, na, no, conf_thres, iou_thres, classes, n_classes, agnostic, multi_label, max_nms,
                              max_det, cand_cap, workspace, workspace_bytes, out, counts, overflow, nullptr, stream);
}

extern "C" int yb_peer_wait(const uint32_t* flags, int n_peers, uint32_t step, int* status, int32_t* counts, int B,
                            unsigned timeout_ms, void* stream) {
  YB_CHECK_ARG(flags && n_peers > 0 && n_peers <= YB_MAX_PEERS && n_peers <= 32 && (!counts || B > 0),
               "yb_peer_wait: bad arguments");
  const unsigned long long ns = (unsigned long long)(timeout_ms ? timeout_ms : 20000u) * 1000000ull;
  k_peer_wait<<<1, 32, 0, (cudaStream_t)stream>>>(flags, n_peers, step, status, counts, B, ns);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

// ---- peer memory: one cudaMalloc'd region per rank, shared with the other ranks of the box through CUDA IPC
extern "C" int yb_peer_alloc(size_t bytes, void** ptr, unsigned char handle[64]) {
  YB_CHECK_ARG(bytes > 0 && ptr && handle, "yb_peer_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  YB_CUDA(cudaMalloc(&p, bytes));
  YB_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    return YB_ERR_CUDA;
  }
  memcpy(handle, &h, 64);
  *ptr = p;
  return YB_OK;
}

extern "C" int yb_peer_open(const unsigned char handle[64], void** ptr) {
  YB_CHECK_ARG(handle && ptr, "yb_peer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  YB_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return YB_OK;
}

extern "C" int yb_peer_close(void* ptr) {
  if (ptr) YB_CUDA(cudaIpcCloseMemHandle(ptr));
  return YB_OK;
}

extern "C" int yb_peer_free(void* ptr) {
  if (ptr) YB_CUDA(cudaFree(ptr));
  return YB_OK;
}
