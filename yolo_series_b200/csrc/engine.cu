// Launch-list executor behind Model.forward (reference models/yolo.py:601-631 forward_once).
// The host-side lowering (yolo_series_b200/plan.py) turns the cfg graph into a flat list of
// primitive ops over NHWC tensor views; this file resolves every bf16 conv to a pre-built TMA /
// tcgen05 plan once, then replays the list on a stream, optionally as a captured CUDA graph so the
// ~100-270 node Python interpreter loop of the reference becomes one graph launch.
#include <vector>

#include "common.cuh"

struct yb_engine {
  std::vector<yb_tensor> tensors;
  std::vector<yb_op> ops;
  std::vector<const void*> weights;
  std::vector<const float*> biases;
  std::vector<yb::ConvTcPlan*> plans;  // per op (nullptr unless bf16 conv)
  std::vector<yb::Conv3Plan*> plans3;  // per op (3x3 s1 slab kernel), takes precedence
  std::vector<yb::ConvS2pPlan*> plans_s2p;  // per op (3x3 s2 pair-slab kernel for 32-channel inputs)
  cudaGraphExec_t graph_exec = nullptr;
  cudaStream_t capture_stream = nullptr;
  int launches = 0;
};

namespace yb {

static int run_ops(yb_engine* e, cudaStream_t s, cudaEvent_t* evs = nullptr) {
  int launches = 0;
  for (size_t i = 0; i < e->ops.size(); ++i) {
    if (evs) cudaEventRecord(evs[i], s);
    const yb_op& o = e->ops[i];
    auto T = [&](int idx) -> const yb_tensor* { return idx >= 0 ? &e->tensors[idx] : nullptr; };
    int rc = YB_OK;
    switch (o.kind) {
      case YB_OP_CONV: {
        if (e->plans_s2p[i]) {
          rc = conv3s2p_launch(e->plans_s2p[i], s);
        } else if (e->plans3[i]) {
          rc = conv3_tc_launch(e->plans3[i], s);
        } else if (e->plans[i]) {
          rc = conv_tc_launch(e->plans[i], s);
        } else {
          yb_conv_desc d;
          memset(&d, 0, sizeof(d));
          d.x = *T(o.in0); d.y = *T(o.out0);
          d.weight = e->weights[o.weight]; d.bias = e->biases[o.weight];
          d.cin_pad = o.cin_pad; d.cout_pad = o.cout_pad; d.k = o.k; d.stride = o.stride; d.act = o.act;
          d.leaky_slope = o.leaky_slope; d.upsample = o.upsample;
          rc = conv_fp32_launch(d, s);
        }
        break;
      }
      case YB_OP_MAXPOOL2: rc = yb_maxpool2x2(T(o.in0), T(o.out0), s); break;
      case YB_OP_POOL_S1: rc = yb_maxpool_s1(T(o.in0), T(o.out0), o.k, s); break;
      case YB_OP_SPP3: rc = yb_spp3(T(o.in0), T(o.out0), T(o.out1), T(o.out2), s); launches += 0; break;
      case YB_OP_UPSAMPLE2: rc = yb_upsample2x(T(o.in0), T(o.out0), s); break;
      case YB_OP_COPY: rc = yb_copy(T(o.in0), T(o.out0), s); break;
      case YB_OP_ADD: rc = yb_add(T(o.in0), T(o.in1), T(o.out0), s); break;
      case YB_OP_STEM: {
        const yb_tensor* im = T(o.in0);
        rc = yb_stem_conv(im->data, im->dtype, im->N, im->C, im->H, im->W, o.tile_n, e->weights[o.weight],
                          e->biases[o.weight], o.cin_pad, o.stride, o.act, o.leaky_slope, T(o.out0), s);
        break;
      }
      default: set_error("engine: unknown op kind %d at %zu", o.kind, i); return YB_ERR_INVALID;
    }
    if (rc) return rc;
    ++launches;
  }
  if (evs) cudaEventRecord(evs[e->ops.size()], s);
  e->launches = launches;
  return YB_OK;
}

}  // namespace yb

using namespace yb;

extern "C" int yb_engine_create(const yb_tensor* tensors, int n_tensors, const yb_op* ops, int n_ops,
                                const void* const* weights, const float* const* biases, int n_weights,
                                yb_engine** out) {
  YB_CHECK_ARG(tensors && ops && out && n_tensors > 0 && n_ops > 0, "yb_engine_create: bad arguments");
  yb_engine* e = new yb_engine();
  e->tensors.assign(tensors, tensors + n_tensors);
  e->ops.assign(ops, ops + n_ops);
  if (n_weights > 0) {
    e->weights.assign(weights, weights + n_weights);
    e->biases.assign(biases, biases + n_weights);
  }
  e->plans.assign(n_ops, nullptr);
  e->plans3.assign(n_ops, nullptr);
  e->plans_s2p.assign(n_ops, nullptr);
  for (int i = 0; i < n_ops; ++i) {
    const yb_op& o = e->ops[i];
    auto ok = [&](int idx) { return idx >= -1 && idx < n_tensors; };
    if (!ok(o.in0) || !ok(o.in1) || !ok(o.out0) || !ok(o.out1) || !ok(o.out2) || o.in0 < 0 || o.out0 < 0) {
      set_error("yb_engine_create: op %d has an invalid tensor index", i);
      yb_engine_destroy(e);
      return YB_ERR_INVALID;
    }
    if (o.kind != YB_OP_CONV && o.kind != YB_OP_STEM) continue;
    if (o.weight < 0 || o.weight >= n_weights) {
      set_error("yb_engine_create: op %d has an invalid weight index", i);
      yb_engine_destroy(e);
      return YB_ERR_INVALID;
    }
    if (o.kind == YB_OP_STEM || e->tensors[o.in0].dtype != YB_BF16) continue;
    yb_conv_desc d;
    memset(&d, 0, sizeof(d));
    d.x = e->tensors[o.in0]; d.y = e->tensors[o.out0];
    d.weight = e->weights[o.weight]; d.bias = e->biases[o.weight];
    d.cin_pad = o.cin_pad; d.cout_pad = o.cout_pad; d.k = o.k; d.stride = o.stride; d.act = o.act;
    d.leaky_slope = o.leaky_slope; d.upsample = o.upsample; d.tile_n = o.tile_n; d.stages = o.stages; d.pair = o.pair;
    if (o.kind == YB_OP_CONV && o.in1 >= 0) {   // residual: y = act(conv(in0)) + in1 (a Shortcut folded into its producer)
      d.res = e->tensors[o.in1];
      d.tile_n = d.tile_n > 0 ? -d.tile_n : (d.tile_n == 0 ? -1 : d.tile_n);
    }
    const bool pooled = o.kind == YB_OP_CONV && o.pool_out >= 0;
    if (pooled) {
      if (o.pool_out >= n_tensors) { set_error("yb_engine_create: op %d has an invalid pool_out index", i); yb_engine_destroy(e); return YB_ERR_INVALID; }
      d.pool = e->tensors[o.pool_out];
      d.tile_n = d.tile_n > 0 ? -d.tile_n : (d.tile_n == 0 ? -1 : d.tile_n);   // the fused pool lives in the generic kernel
    }
    int rc = d.tile_n < 0 ? YB_ERR_UNSUPPORTED : conv3_tc_plan_create(d, &e->plans3[i]);
    if (rc == YB_ERR_UNSUPPORTED && d.tile_n >= 0) rc = conv3s2p_plan_create(d, &e->plans_s2p[i]);
    if (rc == YB_ERR_UNSUPPORTED) {
      if (d.tile_n < 0) d.tile_n = -d.tile_n == 1 ? 0 : -d.tile_n;   // negative tile_n forces the generic kernel
      rc = conv_tc_plan_create(d, &e->plans[i]);
    }
    if (rc) { yb_engine_destroy(e); return rc; }
  }
  *out = e;
  return YB_OK;
}

extern "C" int yb_engine_forward(yb_engine* e, int use_graph, void* stream) {
  YB_CHECK_ARG(e, "yb_engine_forward: null engine");
  cudaStream_t s = (cudaStream_t)stream;
  if (!use_graph) return run_ops(e, s);
  if (!e->graph_exec) {
    // warm every kernel once outside capture (cudaFuncSetAttribute etc. are not capturable)
    int rc = run_ops(e, s);
    if (rc) return rc;
    YB_CUDA(cudaStreamSynchronize(s));
    if (!e->capture_stream) YB_CUDA(cudaStreamCreateWithFlags(&e->capture_stream, cudaStreamNonBlocking));
    cudaGraph_t g = nullptr;
    YB_CUDA(cudaStreamBeginCapture(e->capture_stream, cudaStreamCaptureModeThreadLocal));
    rc = run_ops(e, e->capture_stream);
    cudaError_t ce = cudaStreamEndCapture(e->capture_stream, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    if (ce != cudaSuccess) { set_error("engine: graph capture failed: %s", cudaGetErrorString(ce)); return YB_ERR_CUDA; }
    ce = cudaGraphInstantiate(&e->graph_exec, g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) { set_error("engine: graph instantiate failed: %s", cudaGetErrorString(ce)); return YB_ERR_CUDA; }
  }
  YB_CUDA(cudaGraphLaunch(e->graph_exec, s));
  return YB_OK;
}

extern "C" int yb_engine_profile(yb_engine* e, float* ms_per_op, void* stream) {
  YB_CHECK_ARG(e && ms_per_op, "yb_engine_profile: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const size_t n = e->ops.size();
  std::vector<cudaEvent_t> evs(n + 1);
  for (auto& ev : evs) YB_CUDA(cudaEventCreate(&ev));
  int rc = run_ops(e, s, evs.data());
  if (!rc) {
    cudaError_t ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) { set_error("yb_engine_profile: %s", cudaGetErrorString(ce)); rc = YB_ERR_CUDA; }
  }
  for (size_t i = 0; i < n && !rc; ++i) cudaEventElapsedTime(&ms_per_op[i], evs[i], evs[i + 1]);
  for (auto& ev : evs) cudaEventDestroy(ev);
  return rc;
}

extern "C" int yb_engine_num_launches(const yb_engine* e) { return e ? e->launches : 0; }

extern "C" const char* yb_engine_op_kernel(const yb_engine* e, int i) {
  if (!e || i < 0 || i >= (int)e->ops.size()) return nullptr;
  const yb_op& o = e->ops[i];
  switch (o.kind) {
    case YB_OP_CONV:
      return e->plans_s2p[i] ? "k_conv3s2p_tc" : e->plans3[i] ? conv3_tc_kernel_name(e->plans3[i]) : e->plans[i] ? "k_conv_tc" : "k_conv_fp32";
    case YB_OP_STEM: {
      const yb_tensor& im = e->tensors[o.in0];
      const yb_tensor& y = e->tensors[o.out0];
      const bool tc = !o.tile_n && im.C == 3 && y.C == 32 && im.W % 4 == 0 && (uintptr_t)im.data % 16 == 0;
      const bool tcr = o.tile_n && im.C == 3 && (y.C == 64 || y.C == 80) && o.stride == 1 && im.W % 8 == 0 &&
                       im.H % 2 == 0 && (uintptr_t)im.data % 16 == 0 && getenv("YB_STEM_MMA_SYNC") == nullptr;
      return tc ? "k_stem_tc" : tcr ? "k_stemr_tc" : "k_stem";
    }
    case YB_OP_MAXPOOL2: return "k_maxpool2";
    case YB_OP_POOL_S1: return "k_pool_s1";
    case YB_OP_SPP3: return "k_spp3";
    case YB_OP_UPSAMPLE2: return "k_upsample2";
    case YB_OP_COPY: case YB_OP_ADD: return "k_copy_add";
    default: return "?";
  }
}

extern "C" void yb_engine_destroy(yb_engine* e) {
  if (!e) return;
  for (auto* p : e->plans) conv_tc_plan_destroy(p);
  for (auto* p : e->plans3) conv3_tc_plan_destroy(p);
  for (auto* p : e->plans_s2p) conv3s2p_plan_destroy(p);
  if (e->graph_exec) cudaGraphExecDestroy(e->graph_exec);
  if (e->capture_stream) cudaStreamDestroy(e->capture_stream);
  delete e;
}
