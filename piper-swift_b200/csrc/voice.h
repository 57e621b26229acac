// Voice binding: recognises the Piper-VITS graph from initializer names/shapes plus the
// Conv/ConvTranspose node attributes (the reference interprets the nodes one by one,
// GraphExecutor.swift:227-264; this runtime binds them into a fixed stage pipeline) and
// repacks every weight once into the layout its kernel reads (the reference re-decodes all
// 401 initializers on every call, GraphExecutor.swift:187-189).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "onnx_reader.h"

namespace piper {

struct ConvAttr {
  int stride = 1, dilation = 1, group = 1, pad_begin = -1, pad_end = -1;
};

struct HostTensor {
  std::vector<int64_t> dims;
  std::vector<float> f32;
};

// One Conv / ConvTranspose node of the graph: what ties an initializer to its module when the initializer itself is
// anonymous.  A real Piper export removes weight-norm only from `dec` (HiFi-GAN); the flow's WN convolutions keep
// weight_g / weight_v, which ONNX constant folding turns into initializers called `onnx::Conv_NNNN` - only the node
// name (`/flow/flows.0/enc/in_layers.0/Conv`) and the bias input (`flow.flows.0.enc.in_layers.0.bias`) still carry the
// module path.  The reference never needs names (it interprets node by node, GraphExecutor.swift:227-264).
struct ConvNodeRef {
  std::string node_name, weight_input, bias_input;
  ConvAttr attr;
};

// A graph node as the binder needs it (route A copies it out of the parsed file, route B receives it through
// piper_b200_add_node): names only, plus the Conv / ConvTranspose attributes.
struct GraphNodeRef {
  std::string op_type, name;
  std::vector<std::string> inputs, outputs;
  ConvAttr attr;
};

// fp32 SIMT conv weights: [phases][taps][cin][ldw] (+ bias[ldw]); see kernels.h ConvArgs
struct ConvW {
  float* w = nullptr;
  float* b = nullptr;
  int cin = 0, cout = 0, k = 1, dil = 1, ldw = 0;
  int taps = 1, phases = 1, ct_pad = 0;   // transposed conv: taps = k/stride, phases = stride
  int pad_l = 0;
};
struct LNW {
  float* gamma = nullptr;
  float* beta = nullptr;
  int c = 0;
};
struct DDSLayerW {
  float* w_sep = nullptr;   // [k][C]
  float* b_sep = nullptr;
  LNW n1, n2;
  ConvW c1x1;
  int k = 3, dil = 1;
};
struct EncLayerW {
  ConvW qkv, o, ffn1, ffn2;
  float* rel_k = nullptr;   // [2w+1][kc]
  float* rel_v = nullptr;
  LNW n1, n2;
};
struct ConvFlowW {
  float* pre_w = nullptr;   // [C]
  float* pre_b = nullptr;
  std::vector<DDSLayerW> dds;
  ConvW proj;               // C -> 3*bins-1
  int target = 0;           // which of the two z channels the spline transforms
};
struct CouplingW {
  ConvW pre;                // half -> H (input channels permuted when the layer sits behind a Flip)
  std::vector<ConvW> in_layers;
  std::vector<ConvW> res;   // res part of res_skip (cols [0,H)); empty cout for the last layer
  std::vector<ConvW> skip;  // skip part
  ConvW post;               // H -> half, NEGATED so that x1 - m is an add
  float* cond_w = nullptr;  // [2H*layers][gin] original layout for the speaker GEMV
  float* cond_b = nullptr;
  int in_off = 0, out_off = 0;   // channel offsets of the conditioning / updated half
};
struct ResBlockW {
  std::vector<ConvW> c1, c2;   // ResBlock1: pairs; ResBlock2: c1 only
};
struct StageW {
  ConvW up;                 // transposed conv, polyphase packed
  std::vector<ResBlockW> rbs;
  int ch = 0, stride = 1;
};

struct VoiceHParams {
  int n_vocab = 0, hidden = 0, filter = 0, n_heads = 0, n_layers = 0, ffn_kernel = 0, window = 0;
  int dp_layers = 0, dp_kernel = 0, dp_bins = 0;
  std::vector<int> dp_flow_ids, flow_ids;
  int flow_kernel = 0, flow_wn_layers = 0;
  int resblock = 2, dec_c0 = 0;
  std::vector<std::pair<int, int>> ups;   // (kernel, stride)
  std::vector<int> rb_kernels;
  std::vector<std::vector<int>> rb_dil;
  int gin = 0, n_speakers = 1;
  int hop = 1;
  int64_t n_params = 0;
  double dec_flops_per_frame = 0, flow_flops_per_frame = 0, enc_flops_per_id = 0;
};

struct Voice {
  VoiceHParams hp;
  // host-side staging (route A fills it from the ONNX file, route B from set_weight calls)
  std::map<std::string, HostTensor> tensors;
  std::map<std::string, ConvAttr> conv_attrs;
  std::vector<GraphNodeRef> nodes;       // Conv / ConvTranspose / Gather / Sub / Mul / Exp nodes of the graph
  std::vector<ConvNodeRef> conv_nodes;   // resolved by finalize(): canonical `<module>.weight` / `.bias` aliases
  std::set<std::string> folded_sources;  // initializers that live on under a canonical alias (not counted twice)
  bool finalized = false;

  // device weights
  float* emb = nullptr;
  std::vector<EncLayerW> enc;
  ConvW enc_proj;
  float* emb_g = nullptr;
  ConvW dp_pre, dp_proj;
  std::vector<DDSLayerW> dp_dds;
  float* dp_cond_w = nullptr;
  float* dp_cond_b = nullptr;
  std::vector<ConvFlowW> dp_flows;   // execution order
  int dp_logw_channel = 0;           // which physical z channel holds logw after the flips
  float ea_m = 0.f, ea_logs = 0.f;
  float ea_scale = 0.f;              // exp(-logs) when the export folded it into a constant (ea_has_scale), else unused
  bool ea_has_scale = false;
  std::vector<CouplingW> flows;      // execution order
  ConvW dec_pre;
  float* dec_cond_w = nullptr;
  float* dec_cond_b = nullptr;
  std::vector<StageW> stages;
  float* post_w = nullptr;           // [k][C]
  int post_k = 7;
  float post_b = 0.f;                // dec.conv_post.bias (stock Piper: bias=False -> 0)

  std::vector<void*> allocs;

  void add_onnx(const OnnxModel& m);   // copies initializers + the graph nodes the binder needs
  void add_node(GraphNodeRef g);
  void resolve_graph();                // conv attributes, the `sid`-named embedding table, the folded ElementwiseAffine scale
  void resolve_conv_nodes();           // anonymous conv weights -> `<module>.weight`; weight_g / weight_v pairs folded
  void finalize();                     // throws std::runtime_error on an unrecognised graph
  void release();
  ~Voice() { release(); }
};

}  // namespace piper
