/* ORACLE (test infrastructure, NOT product code).
 * C restatement of the greedy suppression used by the reference's NMS:
 *   utils/general.py:704  i = torchvision.ops.nms(boxes, scores, iou_thres)
 * torchvision is a third-party dependency that is not vendored under /root/reference and is not
 * version-pinned by it (no requirements.txt); the image has torchvision 0.26.0.  Its published CPU
 * algorithm (torchvision/csrc/ops/cpu/nms_kernel.cpp, nms_kernel_impl<float>):
 *   areas = (x2-x1)*(y2-y1); order = stable descending argsort(scores);
 *   for i in order: if suppressed[i] continue; keep i;
 *     for later j in order: w = max(0, min(x2)-max(x1)); h likewise; inter = w*h;
 *       ovr = inter / (area_i + area_j - inter);  if (ovr > iou_threshold [double]) suppressed[j] = 1
 * Pinned by tests/test_oracle_nms.py against torchvision.ops.nms itself (available on both boxes).
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (oracle/build.py).  Input boxes must already be in
 * descending-score order (the caller does the stable sort). */
#include <stdint.h>
#include <stdlib.h>

static inline float fmaxf_(float a, float b) { return a < b ? b : a; } /* std::max(a,b) */
static inline float fminf_(float a, float b) { return b < a ? b : a; } /* std::min(a,b) */

int64_t yb_oracle_nms_sorted(const float* boxes /* n x 4, xyxy, sorted */, int64_t n, double iou_threshold,
                             int64_t max_keep, int64_t* keep /* out, capacity n */) {
  unsigned char* sup = (unsigned char*)calloc((size_t)(n > 0 ? n : 1), 1);
  float* area = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
  int64_t nk = 0;
  for (int64_t i = 0; i < n; ++i) area[i] = (boxes[4 * i + 2] - boxes[4 * i]) * (boxes[4 * i + 3] - boxes[4 * i + 1]);
  for (int64_t i = 0; i < n; ++i) {
    if (sup[i]) continue;
    keep[nk++] = i;
    if (max_keep > 0 && nk >= max_keep) break; /* later keeps cannot change earlier ones */
    const float ix1 = boxes[4 * i], iy1 = boxes[4 * i + 1], ix2 = boxes[4 * i + 2], iy2 = boxes[4 * i + 3];
    const float ia = area[i];
    for (int64_t j = i + 1; j < n; ++j) {
      if (sup[j]) continue;
      const float xx1 = fmaxf_(ix1, boxes[4 * j]), yy1 = fmaxf_(iy1, boxes[4 * j + 1]);
      const float xx2 = fminf_(ix2, boxes[4 * j + 2]), yy2 = fminf_(iy2, boxes[4 * j + 3]);
      const float w = fmaxf_(0.0f, xx2 - xx1), h = fmaxf_(0.0f, yy2 - yy1);
      const float inter = w * h;
      const float ovr = inter / (ia + area[j] - inter);
      if ((double)ovr > iou_threshold) sup[j] = 1;
    }
  }
  free(sup);
  free(area);
  return nk;
}
