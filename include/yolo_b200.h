/*
 * yolo_b200.h -- C ABI of libyolo_b200.so: the B200 (sm_100a) implementation of the
 * YOLOv7 inference hot path of qbxlvnf11/yolo-series.
 *
 * The reference has no FFI layer: its boundary is two Python callables,
 *   Model.forward(x, augment, profile)                   models/yolo.py:581-631
 *   non_max_suppression(prediction, conf_thres, ...)     utils/general.py:628-720
 * plus the head decode inside IDetect/IAuxDetect/Detect (models/yolo.py:42-77,
 * 140-176, 364-399).  The Python shim in yolo_series_b200/ keeps those exact
 * signatures and calls the entry points below through ctypes.  Every entry point
 * takes raw device pointers + sizes + a cudaStream_t (passed as void*), never frees
 * caller memory, performs no hidden synchronisation, and returns 0 on success or a
 * negative yb_status; yb_last_error() gives the message (per thread).
 *
 * Layout conventions: activations are NHWC ("channels last") with an explicit
 * channel stride so that a tensor may be a channel slice of a wider concat buffer
 * (replaces Concat.forward = torch.cat(x, 1), models/common.py:61-62).
 */
#ifndef YOLO_B200_H
#define YOLO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YB_ABI_VERSION 4

typedef enum {
  YB_OK = 0,
  YB_ERR_INVALID = -1,   /* bad argument / unsupported shape */
  YB_ERR_CUDA = -2,      /* a CUDA runtime/driver call failed */
  YB_ERR_WORKSPACE = -3, /* caller workspace too small */
  YB_ERR_OVERFLOW = -4,  /* candidate capacity exceeded (see yb_nms) */
  YB_ERR_UNSUPPORTED = -5
} yb_status;

typedef enum { YB_F32 = 0, YB_BF16 = 1, YB_U8 = 2 } yb_dtype;
typedef enum { YB_ACT_NONE = 0, YB_ACT_SILU = 1, YB_ACT_LEAKY = 2 } yb_act;

const char* yb_last_error(void);
int yb_abi_version(void);

/* ------------------------------------------------------------------ tensors */
/* NHWC view: element (n,h,w,c) lives at data + ((n*H + h)*W + w)*c_stride + c   (in elements) */
typedef struct {
  void* data;       /* device pointer, already offset to channel 0 of the view */
  int32_t dtype;    /* yb_dtype */
  int32_t N, H, W, C;
  int32_t c_stride; /* channels of the underlying buffer (>= C) */
} yb_tensor;

/* ------------------------------------------------------------------ conv
 * Replaces Conv.fuseforward = act(conv2d(x, W', b')) (models/common.py:110-111), the fused
 * RepConv (common.py:498-500) and the 1x1 head convs (models/yolo.py:145), with BN / RepVGG /
 * Implicit folding done on the host at load time (utils/torch_utils.py:181-201,
 * common.py:584-643, yolo.py:178-190).
 *   k in {1,3}, stride in {1,2}, pad = k/2, groups = 1.
 *   weight layout: [Cout_pad][k*k][Cin_pad]  ("KRSC"), dtype = x.dtype; bias: fp32 [Cout_pad].
 *   upsample = 2 writes every output pixel to the 2x2 block it maps to under
 *   nn.Upsample(None, 2, 'nearest') (cfg/training/yolov7.yaml:78) -- y then has 2H x 2W.
 *   bf16 x  -> tcgen05/TMEM implicit-GEMM kernel (fp32 accumulate); y may be bf16 or fp32.
 *   fp32 x  -> fp32 CUDA-core reference-mode kernel; y fp32.
 */
typedef struct {
  yb_tensor x, y;
  const void* weight;
  const float* bias;
  int32_t cin_pad, cout_pad; /* padded extents of the weight array */
  int32_t k, stride;
  int32_t act;               /* yb_act */
  float leaky_slope;
  int32_t upsample;          /* 1 or 2 */
  /* optional residual: y = act(conv) + res (Shortcut, models/common.py:85-86); data==NULL -> none */
  yb_tensor res;
  /* tuning overrides, 0 = auto */
  int32_t tile_n, stages;
  /* CTA pairs (clusters of two CTAs, tcgen05 cta_group::2: one M = 256 MMA per pair, each CTA loads half of the
   * weight tile): 0 = cost model decides, 1 = force, -1 = never */
  int32_t pair;
  /* optional second output: MaxPool2d(2, 2) of y (MP, models/common.py:30-36) written by the same kernel from the
   * tile it has just staged -- the full-resolution tensor is not read again.  data == NULL -> none.  bf16, same
   * channel count as y, H/2 x W/2 (floor); bf16 1x1 / generic-kernel convs without upsample only. */
  yb_tensor pool;
} yb_conv_desc;

int yb_conv2d(const yb_conv_desc* d, void* stream);

/* Reusable plan for one bf16 convolution: the TMA descriptors and tile geometry that yb_conv2d derives on
 * every call are built once (what the engine does per op).  Tensors named by the descriptor must stay valid. */
typedef struct yb_conv_plan yb_conv_plan;
int yb_conv_plan_create(const yb_conv_desc* d, yb_conv_plan** out);
int yb_conv_plan_launch(const yb_conv_plan* p, void* stream);
void yb_conv_plan_destroy(yb_conv_plan* p);

/* ------------------------------------------------------------------ pooling / data movement */
/* MP: MaxPool2d(2,2) (models/common.py:30-36) */
int yb_maxpool2x2(const yb_tensor* x, const yb_tensor* y, void* stream);
/* SP: MaxPool2d(k, 1, k/2) with -inf padding (models/common.py:39-45), k odd <= 13 */
int yb_maxpool_s1(const yb_tensor* x, const yb_tensor* y, int k, void* stream);
/* SPPCSPC.m: the k=5,9,13 stride-1 pools of one tensor in a single pass (models/common.py:271,278) */
int yb_spp3(const yb_tensor* x, const yb_tensor* y5, const yb_tensor* y9, const yb_tensor* y13, void* stream);
/* nn.Upsample(None, 2, 'nearest') stand-alone */
int yb_upsample2x(const yb_tensor* x, const yb_tensor* y, void* stream);
/* channel-slice copy (Concat fallback when a producer cannot write in place) */
int yb_copy(const yb_tensor* x, const yb_tensor* y, void* stream);
/* Shortcut: y = a + b (models/common.py:85-86) */
int yb_add(const yb_tensor* a, const yb_tensor* b, const yb_tensor* y, void* stream);

/* Input staging: NCHW image (uint8 0..255 or float 0..1, B x C x H x W) -> NHWC y (bf16 or fp32,
 * y.C >= C, extra channels zero).  uint8 is scaled by 1/255 (detect.py:100-104).
 * reorg != 0 fuses ReOrg (models/common.py:52-53): y is (B, H/2, W/2, >= 4C) with channel
 * block order (even,even),(odd,even),(even,odd),(odd,odd) in (row,col) parity. */
int yb_stage_input(const void* img, int img_dtype, int B, int C, int H, int W, int reorg,
                   const yb_tensor* y, void* stream);

/* Fused staging + first conv for the bf16 path: reads the NCHW uint8/fp32 image directly (optionally
 * through ReOrg), 3x3 conv (stride 1|2) + bias + activation, writes NHWC bf16.  Same result as
 * yb_stage_input followed by yb_conv2d on bf16 operands (weights [Cout][9][cin_pad] bf16). */
int yb_stem_conv(const void* img, int img_dtype, int B, int C, int H, int W, int reorg,
                 const void* weight, const float* bias, int cin_pad, int stride, int act,
                 float leaky_slope, const yb_tensor* y, void* stream);

/* ------------------------------------------------------------------ head decode
 * IDetect.fuseforward inference branch (models/yolo.py:145-162; arithmetic SURVEY.md A.5).
 * Each level: logits NHWC (B, ny, nx, >= na*no), channel index a*no + o.
 * pred: (B, A, no) fp32, A = na * sum(ny*nx), rows level-major then (a, gy, gx).
 * raw[l] (nullable array / nullable entries): (B, na, ny, nx, no) fp32 un-activated logits
 * (the second return value, yolo.py:147-148).
 */
#define YB_MAX_LEVELS 8
#define YB_MAX_ANCHORS 8
typedef struct {
  yb_tensor logits;
  float stride;
  float anchor_wh[2 * YB_MAX_ANCHORS]; /* anchor_grid in pixels */
} yb_head_level;

int yb_decode(const yb_head_level* levels, int nl, int na, int no, float* pred, float* const* raw,
              void* stream);

/* ------------------------------------------------------------------ NMS
 * non_max_suppression (utils/general.py:628-720) + torchvision.ops.nms (general.py:704):
 *   candidate filter obj > conf_thres, conf = obj*cls (nc==1: conf = obj), xywh->xyxy,
 *   best-class or multi-label expansion, optional class whitelist, stable descending sort,
 *   top-max_nms cut, class-offset (4096 px) greedy IoU suppression with the threshold compared
 *   in double, max_det cut.  Results are bit-identical to the reference on identical fp32 input.
 * pred: (B, A, no) fp32.  out: (B, max_det, 6) fp32 rows [x1,y1,x2,y2,conf,cls]; counts: (B) int32.
 * classes: host array (nullable).  cand_cap: per-image candidate capacity the workspace was sized
 *   for (0 -> A, or A*nc when multi_label); if an image produces more, YB_ERR_OVERFLOW is reported
 *   through *overflow (device int, nullable) -- the call itself never synchronises.
 */
size_t yb_nms_workspace_bytes(int B, int A, int no, int multi_label, int cand_cap);
int yb_nms(const float* pred, int B, int A, int no, float conf_thres, double iou_thres,
           const int* classes, int n_classes, int agnostic, int multi_label, int max_nms, int max_det,
           int cand_cap, void* workspace, size_t workspace_bytes, float* out, int* counts,
           int* overflow, void* stream);

/* yb_nms with NO allocation and NO caller workspace.  This entry point has no counterpart in the reference; it was
 * added for callers that cannot call cudaMalloc (CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The scratch yb_nms takes from the caller lives in a file-scope static
 * __device__ array of yb_nms_static_capacity() = 16 MiB (held by every GPU that loads the library until exit), e.g.
 * 32 images x 25 200 rows in best-class mode.  Larger problems (B*(18*A + 4) + ~1 KiB bytes in best-class mode,
 * B*(16*A*nc + 2*A + 4) with multi_label) are rejected with YB_ERR_WORKSPACE before anything is launched.  Same
 * results as yb_nms.  All calls share the array: calls that overlap on the GPU (different streams / threads) race;
 * serialise them on one stream.  The function itself calls no cudaMalloc / cudaFree or allocating routine. */
size_t yb_nms_static_capacity(void);
int yb_nms_static(const float* pred, int B, int A, int no, float conf_thres, double iou_thres,
                  const int* classes, int n_classes, int agnostic, int multi_label, int max_nms, int max_det,
                  float* out, int* counts, int* overflow /* device, nullable */, void* stream);

/* Fused decode -> candidate filter -> NMS: same result as yb_decode followed by yb_nms, without
 * materialising pred in HBM (the north-star's "one fused decode kernel ... threshold compaction"). */
int yb_decode_nms(const yb_head_level* levels, int nl, int na, int no, float conf_thres,
                  double iou_thres, const int* classes, int n_classes, int agnostic, int multi_label,
                  int max_nms, int max_det, int cand_cap, void* workspace, size_t workspace_bytes,
                  float* out, int* counts, int* overflow, void* stream);

/* ------------------------------------------------------------------ multi-GPU gather fused into NMS
 * Image-sharded inference ends with ONE exchange: every rank's fixed-size detection buffers go to every rank.
 * yb_decode_nms_gather does it inside the NMS kernel: each kept row is stored, as it is produced, into the receive
 * buffer of every GPU of the box over NVLink (peer-to-peer stores into CUDA-IPC mapped memory), and the last CTA
 * raises the sender's arrival flag on every peer (fence.sys + st.release.sys).  yb_peer_wait is the receive side:
 * one warp polls this rank's flags until every sender has reached `step`.  No NCCL call, no packing kernels.
 * A sender that does not arrive within `timeout_ms` (0 = 20 s) is reported, never silently ignored: bit 8 of *status
 * is set and the B entries of that sender's block of `counts` (this rank's receive buffer, nullable) are overwritten
 * with -1, so a stale block can not be mistaken for detections.
 *   out[p]    peer p's receive buffer, (n_peers, B, max_det, 6) fp32: this rank writes block `rank`
 *   counts[p] peer p's (n_peers, B) int32;  flags[p]: peer p's (n_peers) uint32 arrival counters
 *   step      strictly increasing per use of a buffer set (> 0); done_counter: local device int, zero-initialised
 * Buffer sets are double-buffered by the caller (yolo_series_b200/dist.py: PeerGather).  yb_peer_alloc / open / close /
 * free manage the cudaMalloc'd, IPC-exported regions (handle = 64-byte cudaIpcMemHandle_t).
 */
#define YB_MAX_PEERS 16
typedef struct {
  int32_t n_peers, rank;
  float* out[YB_MAX_PEERS];
  int32_t* counts[YB_MAX_PEERS];
  uint32_t* flags[YB_MAX_PEERS];
  uint32_t step;
  int32_t* done_counter;
} yb_peer_gather;

int yb_decode_nms_gather(const yb_head_level* levels, int nl, int na, int no, float conf_thres,
                         double iou_thres, const int* classes, int n_classes, int agnostic, int multi_label,
                         int max_nms, int max_det, int cand_cap, void* workspace, size_t workspace_bytes,
                         float* out, int* counts, int* overflow, const yb_peer_gather* peer /* nullable */,
                         void* stream);
int yb_peer_wait(const uint32_t* flags, int n_peers, uint32_t step, int* status /* device, nullable */,
                 int32_t* counts /* device (n_peers, B), nullable */, int B, unsigned timeout_ms, void* stream);
int yb_peer_alloc(size_t bytes, void** ptr, unsigned char handle[64]);
int yb_peer_open(const unsigned char handle[64], void** ptr);
int yb_peer_close(void* ptr);
int yb_peer_free(void* ptr);

/* ------------------------------------------------------------------ engine (Model.forward)
 * A flat launch list produced by the host-side graph lowering of Model.model
 * (models/yolo.py:601-631 forward_once).  Tensors and weights are caller-owned device memory.
 */
typedef enum {
  YB_OP_CONV = 1, YB_OP_MAXPOOL2 = 2, YB_OP_POOL_S1 = 3, YB_OP_SPP3 = 4, YB_OP_UPSAMPLE2 = 5,
  YB_OP_COPY = 6, YB_OP_ADD = 7,
  YB_OP_STEM = 8 /* in0 = raw NCHW image tensor (dtype u8|f32, c_stride ignored); tile_n carries the ReOrg flag */
} yb_op_kind;

typedef struct {
  int32_t kind;           /* yb_op_kind */
  int32_t in0, in1;       /* tensor indices (-1 = unused) */
  int32_t out0, out1, out2;
  int32_t k, stride, act, upsample;
  float leaky_slope;
  int32_t weight;         /* index into weights/biases (-1 = none) */
  int32_t cin_pad, cout_pad;
  int32_t tile_n, stages; /* 0 = auto */
  int32_t pair;           /* yb_conv_desc.pair */
  int32_t pool_out;       /* YB_OP_CONV: tensor index of the fused 2x2 max-pool output (yb_conv_desc.pool), -1 = none */
} yb_op;

typedef struct yb_engine yb_engine;

int yb_engine_create(const yb_tensor* tensors, int n_tensors, const yb_op* ops, int n_ops,
                     const void* const* weights, const float* const* biases, int n_weights,
                     yb_engine** out);
/* launch every op on `stream`; use_graph != 0 replays a CUDA graph captured on first use */
int yb_engine_forward(yb_engine* e, int use_graph, void* stream);
/* eager run with a CUDA event between consecutive ops; ms_per_op[n_ops] (synchronises `stream`) */
int yb_engine_profile(yb_engine* e, float* ms_per_op, void* stream);
int yb_engine_num_launches(const yb_engine* e);
/* name of the CUDA kernel op `i` launches ("k_conv_tc", "k_conv3_tc", "k_maxpool2", ...); NULL if out of range.
 * Used by bench.py to group the per-op times of yb_engine_profile into per-kernel roofline entries. */
const char* yb_engine_op_kernel(const yb_engine* e, int i);
void yb_engine_destroy(yb_engine* e);

/* ------------------------------------------------------------------ pre-processing
 * letterbox (utils/datasets.py:1277-1307) fused with the detect-side conversion (datasets.py:198-199,
 * detect.py:100-104): bilinear resize with the arithmetic of cv2.resize(INTER_LINEAR) on uint8 (11-bit
 * fixed-point weights; an exact 2x reduction is the 2x2 box average), constant border, optional
 * BGR->RGB swap and HWC->CHW.  The geometry (new_w/new_h/top/left, datasets.py:1283-1304) is computed
 * by the caller; images of one call may have different source sizes.
 */
typedef struct {
  const void* src;        /* device, uint8, HWC, 3 channels */
  int64_t src_pitch;      /* bytes between source rows */
  int32_t src_h, src_w;
  int32_t new_w, new_h;   /* resized size before padding (== src size: no resampling) */
  int32_t top, left;      /* where the resized image sits in the destination */
} yb_letterbox_src;

#define YB_LAYOUT_HWC 0   /* dst (B, dst_h, dst_w, 3): what letterbox() itself returns */
#define YB_LAYOUT_CHW 1   /* dst (B, 3, dst_h, dst_w): what the model reads (datasets.py:198) */

int yb_letterbox(const yb_letterbox_src* srcs /* host array */, int B, void* dst, int dst_h, int dst_w,
                 int layout, int swap_rb, const uint8_t border[3] /* in source channel order */,
                 void* stream);

/* ------------------------------------------------------------------ post-NMS rescale
 * scale_coords + clip_coords (utils/general.py:340-361) and, if do_round, the .round() of detect.py:183
 * on detection rows [x1,y1,x2,y2,...] of `row_stride` floats, in place, fp32:
 *   x = clamp((x - pad_x) / gain, 0, w0), y = clamp((y - pad_y) / gain, 0, h0)   (true division).
 * table (device): per image [gain, pad_x, pad_y, w0, h0].  counts (device, nullable): rows of image b that
 * are valid (NMS `counts`); NULL = all max_rows rows.
 */
int yb_scale_coords(float* dets, int B, int max_rows, int row_stride, const int32_t* counts,
                    const float* table, int do_round, void* stream);

/* ------------------------------------------------------------------ TTA / ensemble
 * Model.forward(augment=True) (models/yolo.py:582-597) runs the forward at scales 1 / 0.83 / 0.67 (the 0.83 pass
 * left-right flipped), de-scales / de-flips each pass's boxes and concatenates the passes along dim 1;
 * Ensemble.forward (models/experimental.py:74-81) concatenates the outputs of several models the same way.
 *
 * yb_scale_img: scale_img (utils/torch_utils.py:247-257) fused with the flip of yolo.py:588:
 *   dst (B, C, dst_h, dst_w) fp32 = F.pad(F.interpolate(flip(src), size=(h, w), mode='bilinear',
 *   align_corners=False), [0, dst_w - w, 0, dst_h - h], value=pad_value);  src (B, C, H, W) fp32, or uint8
 *   (scaled by 1/255 first, detect.py:103).  flip: 0 none, 2 up-down, 3 left-right (the reference's dim numbers).
 *   Bit-identical to ATen's CPU bilinear kernel (torch 2.11) on fp32 input.
 * yb_decode_tta: yb_decode of one pass into rows [row_off, row_off + A_pass) of pred (B, rows_total, no) with
 *   xywh /= scale; flip 2: y = img_h - y; flip 3: x = img_w - x   (yolo.py:591-595).  No raw output (the
 *   reference returns None for it).
 */
typedef struct {
  int32_t rows_total; /* rows per image of the destination pred (sum over all passes / ensemble members) */
  int32_t row_off;    /* first destination row of this pass */
  float scale;        /* de-scale divisor si; 1 = none */
  int32_t flip;       /* 0, 2 or 3 */
  float img_h, img_w; /* size of the un-augmented input (yolo.py:583) */
} yb_tta_pass;

int yb_scale_img(const void* src, int src_dtype, int B, int C, int H, int W, int flip, int h, int w, float* dst,
                 int dst_h, int dst_w, float pad_value, void* stream);
int yb_decode_tta(const yb_head_level* levels, int nl, int na, int no, float* pred, const yb_tta_pass* pass,
                  void* stream);

/* ------------------------------------------------------------------ evaluation matching
 * yb_box_iou: box_iou (utils/general.py:464-486): iou (n, m) fp32 of xyxy boxes, rows of stride1 / stride2 floats.
 * yb_match_detections: the per-image statistics of test.py:179-210 for a whole batch in one launch.
 *   dets (B, max_det, 6) + counts (B, nullable): NMS output in inference space (not modified).
 *   targets (T, 6) fp32 rows [image, class, x, y, w, h], pixels in inference space (test.py:109), grouped by image
 *   in ascending order (what collate_fn produces, utils/datasets.py:621-626).
 *   table (B, 5): per image [gain, pad_x, pad_y, w0, h0] of scale_coords (test.py:144,190; general.py:340-361).
 *   iouv (niou): IoU thresholds (test.py:92: linspace(0.5, 0.95, 10)).
 *   correct (B, max_det, niou) uint8 0/1 = the reference's `correct` per image (rows >= counts[b] are 0).
 *   detected (T, nullable) uint8: which targets were claimed.
 *   status (device int, caller-zeroed): bit 0 = targets not grouped by image / index out of range.
 * Bit-identical to the reference (fp32 IoU, first-maximum ties, greedy claim in confidence order).
 */
int yb_box_iou(const float* box1, int n, int stride1, const float* box2, int m, int stride2, float* iou,
               void* stream);
int yb_match_detections(const float* dets, const int32_t* counts, int B, int max_det, const float* targets, int T,
                        const float* table, const float* iouv, int niou, uint8_t* correct, uint8_t* detected,
                        int* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* YOLO_B200_H */
