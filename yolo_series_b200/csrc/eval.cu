// Evaluation-side matching: box_iou (reference utils/general.py:464-486) and the per-image
// prediction -> target assignment of test.py:179-210, on the device so test.py's batch loop needs no host round trip
// between NMS and the statistics.
//
// k_match: one CTA per image.
//   1. this image's targets are a contiguous range of `targets` (rows [img, cls, x, y, w, h], grouped by image like
//      collate_fn builds them, utils/datasets.py:621-626) found by binary search; a separate check raises *status
//      if the grouping does not hold.  Each target box goes xywh -> xyxy (general.py:275-282) -> native space
//      (scale_coords, general.py:340-361) into shared memory.
//   2. thread per prediction: native-space box (the `predn` of test.py:142-144), best IoU over the targets of its own
//      class (first index on ties = torch.max(1); a NaN IoU poisons the row like torch.max).
//   3. the greedy pass of test.py:199-208 in prediction order (NMS output is sorted by confidence): a prediction
//      whose best IoU exceeds iouv[0] claims its target unless an earlier prediction did; its row of `correct` is
//      best_iou > iouv.  Sequential by definition -> one thread, <= max_det steps, everything in shared memory.
// All arithmetic is single fp32 operations in the reference's order, so `correct` is bit-for-bit the reference's.
#include "common.cuh"

namespace yb {
namespace {

__device__ __forceinline__ float iou_ref(float ax1, float ay1, float ax2, float ay2, float area_a, float bx1, float by1,
                                         float bx2, float by2, float area_b) {
  const float iw = fmaxf(__fsub_rn(fminf(ax2, bx2), fmaxf(ax1, bx1)), 0.0f);
  const float ih = fmaxf(__fsub_rn(fminf(ay2, by2), fmaxf(ay1, by1)), 0.0f);
  const float inter = __fmul_rn(iw, ih);
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_a, area_b), inter));
}

__device__ __forceinline__ float to_native(float v, float pad, float gain, float limit) {
  v = __fdiv_rn(__fsub_rn(v, pad), gain);
  if (v == v) v = fminf(fmaxf(v, 0.0f), limit);
  return v;
}

__global__ void __launch_bounds__(256) k_box_iou(const float* __restrict__ b1, int n, int s1, const float* __restrict__ b2,
                                                 int m, int s2, float* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const float* q = b2 + (size_t)j * s2;
  const float qx1 = q[0], qy1 = q[1], qx2 = q[2], qy2 = q[3];
  const float qa = __fmul_rn(__fsub_rn(qx2, qx1), __fsub_rn(qy2, qy1));
  for (int i = blockIdx.y; i < n; i += gridDim.y) {
    const float* p = b1 + (size_t)i * s1;
    const float px1 = p[0], py1 = p[1], px2 = p[2], py2 = p[3];
    const float pa = __fmul_rn(__fsub_rn(px2, px1), __fsub_rn(py2, py1));
    out[(size_t)i * m + j] = iou_ref(px1, py1, px2, py2, pa, qx1, qy1, qx2, qy2, qa);
  }
}

__global__ void k_check_grouped(const float* __restrict__ targets, int T, int B, int* status) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const float a = targets[(size_t)t * 6];
  bool bad = !(a >= 0.0f && a < (float)B) || a != floorf(a);
  if (t + 1 < T) bad |= targets[(size_t)(t + 1) * 6] < a;
  if (bad) atomicOr(status, 1);
}

__device__ __forceinline__ int lower_bound_img(const float* targets, int T, float key) {
  int lo = 0, hi = T;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (targets[(size_t)mid * 6] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// dynamic smem: float tb[cap][5] (x1,y1,x2,y2,area) | float tcls[cap] | float best[max_det] | int besti[max_det] | u8 taken[cap]
__global__ void __launch_bounds__(256) k_match(const float* __restrict__ dets, const int32_t* __restrict__ counts, int max_det,
                                               const float* __restrict__ targets, int T, const float* __restrict__ table,
                                               const float* __restrict__ iouv, int niou, int cap,
                                               unsigned char* __restrict__ correct, unsigned char* __restrict__ detected,
                                               int* status) {
  extern __shared__ float sm[];
  float* tb = sm;
  float* tcls = tb + (size_t)cap * 5;
  float* best = tcls + cap;
  int* besti = reinterpret_cast<int*>(best + max_det);
  unsigned char* taken = reinterpret_cast<unsigned char*>(besti + max_det);
  const int b = blockIdx.x, tid = threadIdx.x;
  const float gain = table[5 * b], padx = table[5 * b + 1], pady = table[5 * b + 2], w0 = table[5 * b + 3],
              h0 = table[5 * b + 4];
  const int lo = lower_bound_img(targets, T, (float)b), hi = lower_bound_img(targets, T, (float)(b + 1));
  int nt = hi - lo;
  if (nt > cap) {          // cannot happen when cap == T; kept for a caller-limited cap
    if (tid == 0) atomicOr(status, 2);
    nt = cap;
  }
  for (int j = tid; j < nt; j += blockDim.x) {
    const float* t = targets + (size_t)(lo + j) * 6;
    const float hw = __fdiv_rn(t[4], 2.0f), hh = __fdiv_rn(t[5], 2.0f);
    const float x1 = to_native(__fsub_rn(t[2], hw), padx, gain, w0), y1 = to_native(__fsub_rn(t[3], hh), pady, gain, h0);
    const float x2 = to_native(__fadd_rn(t[2], hw), padx, gain, w0), y2 = to_native(__fadd_rn(t[3], hh), pady, gain, h0);
    tb[5 * j] = x1; tb[5 * j + 1] = y1; tb[5 * j + 2] = x2; tb[5 * j + 3] = y2;
    tb[5 * j + 4] = __fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1));
    tcls[j] = t[1];
    taken[j] = 0;
  }
  const int np = min(counts ? counts[b] : max_det, max_det);
  __syncthreads();
  for (int p = tid; p < max_det; p += blockDim.x) {
    unsigned char* c = correct + ((size_t)b * max_det + p) * niou;
    for (int k = 0; k < niou; ++k) c[k] = 0;
    if (p >= np) continue;
    const float* d = dets + ((size_t)b * max_det + p) * 6;
    const float x1 = to_native(d[0], padx, gain, w0), y1 = to_native(d[1], pady, gain, h0);
    const float x2 = to_native(d[2], padx, gain, w0), y2 = to_native(d[3], pady, gain, h0);
    const float area = __fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1)), pc = d[5];
    float bv = -1.0f;
    int bi = -1;
    bool nan = false;
    for (int j = 0; j < nt; ++j) {
      if (tcls[j] != pc) continue;
      const float v = iou_ref(x1, y1, x2, y2, area, tb[5 * j], tb[5 * j + 1], tb[5 * j + 2], tb[5 * j + 3], tb[5 * j + 4]);
      if (v != v) nan = true;
      else if (bi < 0 || v > bv) { bv = v; bi = j; }
    }
    best[p] = nan ? -1.0f : bv;      // NaN > iouv[0] is false: the row never matches
    besti[p] = nan ? -1 : bi;
  }
  __syncthreads();
  if (tid == 0) {
    const float thr0 = iouv[0];
    int found = 0;
    for (int p = 0; p < np && found < nt; ++p) {
      const int j = besti[p];
      if (j < 0 || !(best[p] > thr0) || taken[j]) continue;
      taken[j] = 1;
      ++found;
      unsigned char* c = correct + ((size_t)b * max_det + p) * niou;
      for (int k = 0; k < niou; ++k) c[k] = best[p] > iouv[k] ? 1 : 0;
    }
  }
  if (detected) {
    __syncthreads();
    for (int j = tid; j < nt; j += blockDim.x) detected[lo + j] = taken[j];
  }
}

}  // namespace
}  // namespace yb

extern "C" int yb_box_iou(const float* box1, int n, int stride1, const float* box2, int m, int stride2, float* iou,
                          void* stream) {
  using namespace yb;
  YB_CHECK_ARG(n >= 0 && m >= 0 && stride1 >= 4 && stride2 >= 4, "yb_box_iou: bad sizes");
  if (n == 0 || m == 0) return YB_OK;
  YB_CHECK_ARG(box1 && box2 && iou, "yb_box_iou: NULL pointer");
  const int rows = n < 4 * num_sms() ? n : 4 * num_sms();
  dim3 grid(ceil_div(m, 256), rows), block(256);
  k_box_iou<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(box1, n, stride1, box2, m, stride2, iou);
  YB_LAUNCH_CHECK();
  return YB_OK;
}

extern "C" int yb_match_detections(const float* dets, const int32_t* counts, int B, int max_det, const float* targets,
                                   int T, const float* table, const float* iouv, int niou, uint8_t* correct,
                                   uint8_t* detected, int* status, void* stream) {
  using namespace yb;
  YB_CHECK_ARG(dets && table && iouv && correct && status && B > 0 && max_det > 0 && niou > 0 && T >= 0,
               "yb_match_detections: bad arguments");
  YB_CHECK_ARG(T == 0 || targets, "yb_match_detections: targets is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int cap = T > 0 ? T : 1;
  const size_t smem = (size_t)cap * 6 * sizeof(float) + (size_t)max_det * 8 + (size_t)cap;
  YB_CHECK_ARG(smem <= 200 * 1024, "yb_match_detections: %d targets + max_det %d need %zu B of shared memory (> 200 KB); "
               "split the batch", T, max_det, smem);
  if (smem > 48 * 1024) YB_CUDA(cudaFuncSetAttribute(k_match, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  if (T > 0) {
    k_check_grouped<<<ceil_div(T, 256), 256, 0, st>>>(targets, T, B, status);
    YB_LAUNCH_CHECK();
  }
  k_match<<<B, 256, smem, st>>>(dets, counts, max_det, targets, T, table, iouv, niou, cap, correct, detected, status);
  YB_LAUNCH_CHECK();
  return YB_OK;
}
