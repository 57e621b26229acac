#include "voice.h"

#include <cmath>
#include <cstring>
#include <stdexcept>

namespace piper {
namespace {

#define CUDA_OK(x)                                                                         \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess)                                                                 \
      throw std::runtime_error(std::string("cuda: ") + cudaGetErrorString(e_) + " at " #x); \
  } while (0)

int round_up(int v, int m) { return (v + m - 1) / m * m; }

struct Binder {
  Voice& v;
  explicit Binder(Voice& voice) : v(voice) {}

  bool has(const std::string& n) const { return v.tensors.count(n) != 0; }
  const HostTensor& get(const std::string& n) const {
    auto it = v.tensors.find(n);
    if (it == v.tensors.end())
      throw std::runtime_error("unsupported: initializer '" + n + "' not found (not a Piper VITS export?)");
    return it->second;
  }
  ConvAttr attr(const std::string& wname) const {
    auto it = v.conv_attrs.find(wname);
    return it == v.conv_attrs.end() ? ConvAttr{} : it->second;
  }
  float* upload(const std::vector<float>& h) {
    float* d = nullptr;
    CUDA_OK(cudaMalloc(&d, std::max<size_t>(h.size(), 1) * sizeof(float)));
    v.allocs.push_back(d);
    if (!h.empty()) CUDA_OK(cudaMemcpy(d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
    return d;
  }
  float* upload_vec(const std::string& n, int expect = -1) {
    const HostTensor& t = get(n);
    if (expect >= 0 && (int)t.f32.size() != expect)
      throw std::runtime_error("unsupported: '" + n + "' has " + std::to_string(t.f32.size()) +
                               " elements, expected " + std::to_string(expect));
    return upload(t.f32);
  }
  LNW ln(const std::string& p, int c) {
    LNW o;
    o.gamma = upload_vec(p + ".gamma", c);
    o.beta = upload_vec(p + ".beta", c);
    o.c = c;
    return o;
  }

  // ONNX Conv weight [cout][cin][k] -> [k][cin][ldw]; optional input-channel permutation
  // (reversal), output-row reversal, column sub-range and sign flip (used to fold the flow's
  // Flip layers and the coupling subtraction into the weights).
  ConvW conv(const std::string& p, int pad_l_override = -1, bool rev_in = false, bool rev_out = false,
             int co_begin = 0, int co_count = -1, float sign = 1.f) {
    const HostTensor& w = get(p + ".weight");
    if (w.dims.size() != 3) throw std::runtime_error("unsupported: '" + p + ".weight' is not rank 3");
    const int cout_all = (int)w.dims[0], cin = (int)w.dims[1], k = (int)w.dims[2];
    const int cout = co_count < 0 ? cout_all - co_begin : co_count;
    const ConvAttr a = attr(p + ".weight");
    if (a.group != 1) throw std::runtime_error("unsupported: grouped conv '" + p + "' on the dense path");
    if (cout_all <= 0 || cin <= 0 || k <= 0 || w.f32.size() != (size_t)cout_all * cin * k || co_begin < 0 || cout <= 0 ||
        co_begin + cout > cout_all)
      throw std::runtime_error("unsupported: '" + p + ".weight' holds " + std::to_string(w.f32.size()) +
                               " values for dims " + std::to_string(cout_all) + "x" + std::to_string(cin) + "x" + std::to_string(k));
    ConvW o;
    o.cin = cin; o.cout = cout; o.k = k; o.taps = k; o.dil = a.dilation;
    o.ldw = round_up(cout, 64);
    o.pad_l = pad_l_override >= 0 ? pad_l_override : (a.pad_begin >= 0 ? a.pad_begin : a.dilation * (k - 1) / 2);
    std::vector<float> pk((size_t)k * cin * o.ldw, 0.f);
    for (int co = 0; co < cout; ++co) {
      const int src_co = co_begin + (rev_out ? cout - 1 - co : co);
      for (int ci = 0; ci < cin; ++ci) {
        const int src_ci = rev_in ? cin - 1 - ci : ci;
        for (int t = 0; t < k; ++t)
          pk[((size_t)t * cin + ci) * o.ldw + co] = sign * w.f32[((size_t)src_co * cin + src_ci) * k + t];
      }
    }
    o.w = upload(pk);
    std::vector<float> b(o.ldw, 0.f);
    if (has(p + ".bias")) {
      const HostTensor& bt = get(p + ".bias");
      if ((int)bt.f32.size() != cout_all) throw std::runtime_error("unsupported: '" + p + ".bias' does not match the weight");
      for (int co = 0; co < cout; ++co)
        b[co] = sign * bt.f32[co_begin + (rev_out ? cout - 1 - co : co)];
    }
    o.b = upload(b);
    return o;
  }

  // fused q|k|v 1x1 convs -> one [1][H][3H] weight
  ConvW qkv(const std::string& p, int H) {
    ConvW o;
    o.cin = H; o.cout = 3 * H; o.k = 1; o.taps = 1; o.ldw = round_up(3 * H, 64);
    std::vector<float> pk((size_t)H * o.ldw, 0.f), b(o.ldw, 0.f);
    const char* names[3] = {".conv_q", ".conv_k", ".conv_v"};
    for (int s = 0; s < 3; ++s) {
      const HostTensor& w = get(p + names[s] + ".weight");
      const HostTensor& bt = get(p + names[s] + ".bias");
      if ((int)w.f32.size() != H * H || (int)bt.f32.size() != H) throw std::runtime_error("unsupported: attention conv shape in " + p);
      for (int co = 0; co < H; ++co) {
        b[s * H + co] = bt.f32[co];
        for (int ci = 0; ci < H; ++ci) pk[(size_t)ci * o.ldw + s * H + co] = w.f32[(size_t)co * H + ci];
      }
    }
    o.w = upload(pk);
    o.b = upload(b);
    return o;
  }

  // ConvTranspose weight [cin][cout][k], stride u, pad p -> polyphase [u][k/u][cin][ldw]:
  // out[q*u+ph] = sum_j in[q + (ph+p)/u - j] * W[:, :, (ph+p)%u + j*u]
  ConvW conv_transpose(const std::string& p, int stride_fallback) {
    const HostTensor& w = get(p + ".weight");
    if (w.dims.size() != 3) throw std::runtime_error("unsupported: '" + p + ".weight' is not rank 3");
    const int cin = (int)w.dims[0], cout = (int)w.dims[1], k = (int)w.dims[2];
    if (cin <= 0 || cout <= 0 || k <= 0 || w.f32.size() != (size_t)cin * cout * k)
      throw std::runtime_error("unsupported: '" + p + ".weight' element count does not match its dims");
    const ConvAttr a = attr(p + ".weight");
    const int u = v.conv_attrs.count(p + ".weight") ? a.stride : stride_fallback;
    if (u <= 0 || k % u != 0)
      throw std::runtime_error("unsupported: ConvTranspose '" + p + "' kernel not a multiple of stride");
    const int pad = a.pad_begin >= 0 ? a.pad_begin : (k - u) / 2;
    ConvW o;
    o.cin = cin; o.cout = cout; o.k = k; o.phases = u; o.taps = k / u; o.ct_pad = pad;
    o.ldw = round_up(cout, 64);
    std::vector<float> pk((size_t)u * o.taps * cin * o.ldw, 0.f);
    for (int ph = 0; ph < u; ++ph) {
      const int kk0 = (ph + pad) % u;
      for (int j = 0; j < o.taps; ++j)
        for (int ci = 0; ci < cin; ++ci)
          for (int co = 0; co < cout; ++co)
            pk[(((size_t)ph * o.taps + j) * cin + ci) * o.ldw + co] = w.f32[((size_t)ci * cout + co) * k + kk0 + j * u];
    }
    o.w = upload(pk);
    std::vector<float> b(o.ldw, 0.f);
    if (has(p + ".bias")) {
      const HostTensor& bt = get(p + ".bias");
      if ((int)bt.f32.size() != cout) throw std::runtime_error("unsupported: '" + p + ".bias' does not match the weight");
      for (int co = 0; co < cout; ++co) b[co] = bt.f32[co];
    }
    o.b = upload(b);
    return o;
  }

  std::vector<DDSLayerW> dds(const std::string& p, int C, int n_layers) {
    std::vector<DDSLayerW> out;
    for (int i = 0; i < n_layers; ++i) {
      DDSLayerW l;
      const std::string sp = p + ".convs_sep." + std::to_string(i);
      const HostTensor& w = get(sp + ".weight");   // [C][1][k]
      if (w.dims.size() != 3 || w.dims[2] <= 0 || w.f32.size() != (size_t)C * (size_t)w.dims[2])
        throw std::runtime_error("unsupported: depthwise weight '" + sp + ".weight' shape");
      l.k = (int)w.dims[2];
      const ConvAttr a = attr(sp + ".weight");
      l.dil = v.conv_attrs.count(sp + ".weight") ? a.dilation : (int)std::lround(std::pow((double)l.k, i));
      std::vector<float> pk((size_t)l.k * C);
      for (int c = 0; c < C; ++c)
        for (int t = 0; t < l.k; ++t) pk[(size_t)t * C + c] = w.f32[(size_t)c * l.k + t];
      l.w_sep = upload(pk);
      l.b_sep = upload_vec(sp + ".bias", C);
      l.n1 = ln(p + ".norms_1." + std::to_string(i), C);
      l.n2 = ln(p + ".norms_2." + std::to_string(i), C);
      l.c1x1 = conv(p + ".convs_1x1." + std::to_string(i));
      out.push_back(l);
    }
    return out;
  }
};

}  // namespace

void Voice::add_onnx(const OnnxModel& m) {
  for (const auto& t : m.initializers) {
    if (t.dtype != 1) continue;   // int64 initializers (sid, shape constants) are not weights
    HostTensor h;
    h.dims = t.dims;
    h.f32 = t.f32;
    tensors[t.name] = std::move(h);
  }
  for (const auto& n : m.nodes) {
    // float values of Constant nodes live under their output name (an export without constant folding keeps scalars there)
    if (n.op_type == "Constant" && !n.outputs.empty()) {
      if (auto x = n.attr("value"); x && x->has_t && x->t.dtype == 1 && !tensors.count(n.outputs[0])) {
        HostTensor h;
        h.dims = x->t.dims;
        h.f32 = x->t.f32;
        tensors[n.outputs[0]] = std::move(h);
        folded_sources.insert(n.outputs[0]);
      }
      continue;
    }
    GraphNodeRef g;
    g.op_type = n.op_type; g.name = n.name; g.inputs = n.inputs; g.outputs = n.outputs;
    if (auto x = n.attr("strides"); x && !x->ints.empty()) g.attr.stride = (int)x->ints[0];
    if (auto x = n.attr("dilations"); x && !x->ints.empty()) g.attr.dilation = (int)x->ints[0];
    if (auto x = n.attr("group")) g.attr.group = (int)x->i;
    if (auto x = n.attr("pads"); x && x->ints.size() >= 2) {
      g.attr.pad_begin = (int)x->ints[0];
      g.attr.pad_end = (int)x->ints[1];
    }
    add_node(std::move(g));
  }
}

// Route A (add_onnx) and route B (piper_b200_add_node) both end here: only the node kinds the binder needs are kept
void Voice::add_node(GraphNodeRef g) {
  static const char* kept[] = {"Conv", "ConvTranspose", "Gather", "Sub", "Mul", "Exp"};
  for (const char* k : kept)
    if (g.op_type == k) {
      nodes.push_back(std::move(g));
      return;
    }
}

// What the graph says about initializers that are not named after their module (real piper_train exports)
void Voice::resolve_graph() {
  for (const GraphNodeRef& n : nodes) {
    if ((n.op_type != "Conv" && n.op_type != "ConvTranspose") || n.inputs.size() < 2) continue;
    conv_attrs[n.inputs[1]] = n.attr;
    ConvNodeRef r;
    r.node_name = n.name;
    r.weight_input = n.inputs[1];
    if (n.inputs.size() >= 3) r.bias_input = n.inputs[2];
    r.attr = n.attr;
    conv_nodes.push_back(std::move(r));
  }
  // The phoneme embedding of a real single-speaker export is NOT called enc_p.emb.weight: piper_train passes
  // input_names = [input, input_lengths, scales, sid] with three real inputs, so the exporter gives the spare name
  // "sid" to the first parameter - the embedding table (that is the `sid` initializer ONNXParsingTests.swift:31
  // finds; the first node is Gather(sid, input), :35).  Bind the table through that Gather.
  if (!tensors.count("enc_p.emb.weight"))
    for (const GraphNodeRef& n : nodes)
      if (n.op_type == "Gather" && n.inputs.size() == 2 && n.inputs[1] == "input") {
        auto it = tensors.find(n.inputs[0]);
        if (it != tensors.end() && it->second.dims.size() == 2) {
          tensors["enc_p.emb.weight"] = it->second;
          folded_sources.insert(n.inputs[0]);
        }
        break;
      }
  // ElementwiseAffine of the duration predictor, reverse: (z - m) * exp(-logs).  Constant folding replaces -logs (or the
  // whole factor) by an anonymous constant and drops `dp.flows.0.logs`: follow Sub(., dp.flows.0.m) to the Mul that scales it.
  if (!tensors.count("dp.flows.0.logs") && !tensors.count("dp.flows.0.inv_scale")) {
    std::string diff;
    for (const GraphNodeRef& n : nodes)
      if (n.op_type == "Sub" && n.inputs.size() == 2 && n.inputs[1] == "dp.flows.0.m" && !n.outputs.empty())
        diff = n.outputs[0];
    for (const GraphNodeRef& n : nodes) {
      if (diff.empty() || n.op_type != "Mul" || n.inputs.size() != 2) continue;
      const std::string* other = n.inputs[0] == diff ? &n.inputs[1] : (n.inputs[1] == diff ? &n.inputs[0] : nullptr);
      if (!other) continue;
      // (a) only the negation was folded: Mul(diff, Exp(onnx::Exp_NNNN)) with the initializer holding -logs
      bool found = false;
      for (const GraphNodeRef& e : nodes)
        if (e.op_type == "Exp" && e.inputs.size() == 1 && !e.outputs.empty() && e.outputs[0] == *other) {
          auto it = tensors.find(e.inputs[0]);
          if (it != tensors.end() && !it->second.f32.empty()) {
            HostTensor lg = it->second;
            for (float& v : lg.f32) v = -v;
            tensors["dp.flows.0.logs"] = std::move(lg);
            folded_sources.insert(e.inputs[0]);
            found = true;
          }
        }
      if (found) break;
      // (b) the whole factor exp(-logs) was folded into an initializer / Constant
      auto it = tensors.find(*other);
      if (it != tensors.end() && (it->second.f32.size() == 2 || it->second.f32.size() == 1)) {   // a mask has T entries
        HostTensor h = it->second;
        tensors["dp.flows.0.inv_scale"] = std::move(h);
        break;
      }
    }
  }
}

namespace {
// "/flow/flows.0/enc/in_layers.0/Conv" (also ".../Conv_1", ".../ConvTranspose") -> "flow.flows.0.enc.in_layers.0"
std::string module_of_node_name(const std::string& name) {
  if (name.empty() || name[0] != '/') return "";
  const size_t cut = name.rfind('/');
  if (cut == 0 || cut == std::string::npos || name.compare(cut + 1, 4, "Conv") != 0) return "";
  std::string s = name.substr(1, cut - 1);
  for (char& c : s)
    if (c == '/') c = '.';
  return s;
}
bool ends_with(const std::string& s, const char* suf) {
  const size_t n = std::strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}
}  // namespace

void Voice::resolve_conv_nodes() {
  // (1) weight-norm pairs kept as parameters (export without constant folding, or route B passing the raw state
  //     dict): w = g * v / ||v||, the norm over every axis but the first (torch.nn.utils.weight_norm, dim = 0)
  std::vector<std::string> pairs;
  for (const auto& kv : tensors)
    if (ends_with(kv.first, ".weight_v")) pairs.push_back(kv.first.substr(0, kv.first.size() - 9));
  for (const std::string& p : pairs) {
    if (tensors.count(p + ".weight") || !tensors.count(p + ".weight_g")) continue;
    const HostTensor& v = tensors[p + ".weight_v"];
    const HostTensor& g = tensors[p + ".weight_g"];
    if (v.dims.empty() || v.dims[0] <= 0 || (int64_t)g.f32.size() != v.dims[0] || v.f32.size() % (size_t)v.dims[0])
      throw std::runtime_error("unsupported: weight_g / weight_v shapes of '" + p + "' do not match");
    const size_t rows = (size_t)v.dims[0], inner = v.f32.size() / rows;
    HostTensor w;
    w.dims = v.dims;
    w.f32.resize(v.f32.size());
    for (size_t r = 0; r < rows; ++r) {
      double ss = 0.0;
      for (size_t i = 0; i < inner; ++i) ss += (double)v.f32[r * inner + i] * (double)v.f32[r * inner + i];
      const double sc = (double)g.f32[r] / std::sqrt(ss);
      for (size_t i = 0; i < inner; ++i) w.f32[r * inner + i] = (float)((double)v.f32[r * inner + i] * sc);
    }
    tensors[p + ".weight"] = std::move(w);
    folded_sources.insert(p + ".weight_v");
    folded_sources.insert(p + ".weight_g");
  }
  // (2) anonymous folded weights: the literal `<module>.weight` initializer is the first choice, else whatever
  //     initializer the module's Conv node reads; the module path comes from the named bias input or the node name
  for (const ConvNodeRef& r : conv_nodes) {
    std::string mod;
    if (ends_with(r.weight_input, ".weight") && tensors.count(r.weight_input)) mod = r.weight_input.substr(0, r.weight_input.size() - 7);
    else if (ends_with(r.bias_input, ".bias")) mod = r.bias_input.substr(0, r.bias_input.size() - 5);
    else mod = module_of_node_name(r.node_name);
    if (mod.empty()) continue;
    conv_attrs[mod + ".weight"] = r.attr;
    if (!tensors.count(mod + ".weight")) {
      auto it = tensors.find(r.weight_input);
      if (it != tensors.end()) {
        tensors[mod + ".weight"] = it->second;
        folded_sources.insert(r.weight_input);
      }
    }
    if (!r.bias_input.empty() && !tensors.count(mod + ".bias")) {
      auto it = tensors.find(r.bias_input);
      if (it != tensors.end()) {
        tensors[mod + ".bias"] = it->second;
        folded_sources.insert(r.bias_input);
      }
    }
  }
}

void Voice::release() {
  for (void* p : allocs) cudaFree(p);
  allocs.clear();
  finalized = false;
}

void Voice::finalize() {
  if (finalized) return;
  resolve_graph();
  resolve_conv_nodes();
  Binder B(*this);
  auto count = [&](const std::string& pre, const std::string& post) {
    int i = 0;
    while (B.has(pre + std::to_string(i) + post)) ++i;
    return i;
  };
  // ---- hyper-parameters from shapes (+ node attributes) ------------------------------------
  const HostTensor& embt = B.get("enc_p.emb.weight");
  if (embt.dims.size() != 2) throw std::runtime_error("unsupported: enc_p.emb.weight is not rank 2");
  hp.n_vocab = (int)embt.dims[0];
  hp.hidden = (int)embt.dims[1];
  const int H = hp.hidden;
  const HostTensor& relk = B.get("enc_p.encoder.attn_layers.0.emb_rel_k");
  if (relk.dims.size() != 3) throw std::runtime_error("unsupported: emb_rel_k is not rank 3");
  hp.window = ((int)relk.dims[1] - 1) / 2;
  const int kc = (int)relk.dims[2];
  hp.n_heads = H / kc;
  hp.n_layers = count("enc_p.encoder.attn_layers.", ".conv_q.weight");
  const HostTensor& f1 = B.get("enc_p.encoder.ffn_layers.0.conv_1.weight");
  hp.filter = (int)f1.dims[0];
  hp.ffn_kernel = (int)f1.dims[2];
  if (H > 256 || H % 4 != 0 || kc > 128) throw std::runtime_error("unsupported: hidden size outside kernel limits");
  hp.dp_layers = count("dp.convs.convs_sep.", ".weight");
  hp.dp_kernel = (int)B.get("dp.convs.convs_sep.0.weight").dims[2];
  for (int i = 0; i < 32; ++i) {
    if (B.has("dp.flows." + std::to_string(i) + ".proj.weight")) hp.dp_flow_ids.push_back(i);
    if (B.has("flow.flows." + std::to_string(i) + ".pre.weight")) hp.flow_ids.push_back(i);
  }
  if (hp.dp_flow_ids.empty() || hp.flow_ids.empty())
    throw std::runtime_error("unsupported: no dp.flows.*/flow.flows.* initializers (not a Piper VITS export)");
  const int nparam = (int)B.get("dp.flows." + std::to_string(hp.dp_flow_ids[0]) + ".proj.weight").dims[0];
  hp.dp_bins = (nparam + 1) / 3;
  if (hp.dp_bins > 16) throw std::runtime_error("unsupported: more than 16 spline bins");
  const std::string fl0 = "flow.flows." + std::to_string(hp.flow_ids[0]);
  hp.flow_kernel = (int)B.get(fl0 + ".enc.in_layers.0.weight").dims[2];
  hp.flow_wn_layers = count(fl0 + ".enc.in_layers.", ".weight");
  hp.dec_c0 = (int)B.get("dec.conv_pre.weight").dims[0];
  hp.resblock = B.has("dec.resblocks.0.convs1.0.weight") ? 1 : 2;
  const int n_ups = count("dec.ups.", ".weight");
  const std::string rbc = hp.resblock == 1 ? ".convs1." : ".convs.";
  const int n_rb = count("dec.resblocks.", rbc + "0.weight");
  if (n_ups == 0 || n_rb % n_ups != 0) throw std::runtime_error("unsupported: HiFi-GAN resblock count");
  const int nk = n_rb / n_ups;
  hp.gin = B.has("emb_g.weight") ? (int)B.get("emb_g.weight").dims[1] : 0;
  hp.n_speakers = hp.gin ? (int)B.get("emb_g.weight").dims[0] : 1;
  hp.n_params = 0;
  for (auto& kv : tensors)   // aliases made by resolve_conv_nodes() count once, under their canonical name
    if (!folded_sources.count(kv.first)) hp.n_params += (int64_t)kv.second.f32.size();

  // ---- text encoder ---------------------------------------------------------------------------
  emb = B.upload(embt.f32);
  const int fpl = (hp.ffn_kernel - 1) / 2;   // FFN "same" padding is an explicit Pad node: (k-1)/2 left, k/2 right
  for (int i = 0; i < hp.n_layers; ++i) {
    const std::string p = "enc_p.encoder.attn_layers." + std::to_string(i);
    EncLayerW l;
    l.qkv = B.qkv(p, H);
    l.o = B.conv(p + ".conv_o");
    l.rel_k = B.upload_vec(p + ".emb_rel_k", (2 * hp.window + 1) * kc);
    l.rel_v = B.upload_vec(p + ".emb_rel_v", (2 * hp.window + 1) * kc);
    l.n1 = B.ln("enc_p.encoder.norm_layers_1." + std::to_string(i), H);
    l.n2 = B.ln("enc_p.encoder.norm_layers_2." + std::to_string(i), H);
    l.ffn1 = B.conv("enc_p.encoder.ffn_layers." + std::to_string(i) + ".conv_1", fpl);
    l.ffn2 = B.conv("enc_p.encoder.ffn_layers." + std::to_string(i) + ".conv_2", fpl);
    enc.push_back(l);
  }
  enc_proj = B.conv("enc_p.proj");
  if (enc_proj.cout != 2 * H) throw std::runtime_error("unsupported: enc_p.proj must emit 2*hidden channels");
  if (hp.gin) emb_g = B.upload_vec("emb_g.weight");

  // ---- stochastic duration predictor (reverse order, one ConvFlow dropped) ------------------------
  dp_pre = B.conv("dp.pre");
  dp_proj = B.conv("dp.proj");
  dp_dds = B.dds("dp.convs", H, hp.dp_layers);
  if (hp.gin) {
    dp_cond_w = B.upload_vec("dp.cond.weight", H * hp.gin);
    dp_cond_b = B.upload_vec("dp.cond.bias", H);
  }
  {
    struct Step { int kind; int id; };   // 0 EA, 1 ConvFlow, 2 Flip
    // forward order [EA 0, ConvFlow 1, Flip 2, ConvFlow 3, Flip 4, ...]; reverse drops ConvFlow 1 ("a useless vflow") but
    // keeps its Flip.  An exported graph no longer holds dp.flows.1's weights (unused -> not exported), the synthetic
    // state dict does: put 1 in the sequence either way so that exactly that one is dropped.
    std::vector<Step> seq{{0, 0}};
    std::vector<int> cf_ids = hp.dp_flow_ids;
    if (cf_ids.empty() || cf_ids[0] != 1) cf_ids.insert(cf_ids.begin(), 1);
    for (int id : cf_ids) {
      seq.push_back({1, id});
      seq.push_back({2, 0});
    }
    std::vector<Step> rev(seq.rbegin(), seq.rend());
    if (rev.size() >= 2) rev.erase(rev.end() - 2);   // "remove a useless vflow"
    int L0 = 0, L1 = 1;   // physical channel holding logical channel 0 / 1
    for (const Step& s : rev) {
      if (s.kind == 2) {
        std::swap(L0, L1);
      } else if (s.kind == 1) {
        const std::string p = "dp.flows." + std::to_string(s.id);
        ConvFlowW f;
        f.pre_w = B.upload_vec(p + ".pre.weight", H);
        f.pre_b = B.upload_vec(p + ".pre.bias", H);
        f.dds = B.dds(p + ".convs", H, hp.dp_layers);
        f.proj = B.conv(p + ".proj");
        f.target = L1;
        dp_flows.push_back(f);
      } else {
        const HostTensor& m = B.get("dp.flows.0.m");
        if (m.f32.empty()) throw std::runtime_error("unsupported: dp.flows.0.m is empty");
        ea_m = m.f32[0];
        if (B.has("dp.flows.0.logs")) {
          const HostTensor& lg = B.get("dp.flows.0.logs");
          if (lg.f32.empty()) throw std::runtime_error("unsupported: dp.flows.0.logs is empty");
          ea_logs = lg.f32[0];
        } else {   // exp(-logs) folded into a constant by the exporter (add_onnx located it)
          const HostTensor& sc = B.get("dp.flows.0.inv_scale");
          ea_scale = sc.f32[0];
          ea_has_scale = true;
        }
        dp_logw_channel = L0;
      }
    }
  }

  // ---- residual coupling flow (reverse); Flip layers folded into the weights ---------------------
  {
    const int half = H / 2;
    int flips = 0;
    for (auto it = hp.flow_ids.rbegin(); it != hp.flow_ids.rend(); ++it) {
      ++flips;   // each coupling layer is preceded by one Flip in reverse order
      const bool primed = flips % 2 == 1;
      const std::string p = "flow.flows." + std::to_string(*it);
      CouplingW c;
      c.in_off = primed ? half : 0;
      c.out_off = primed ? 0 : half;
      c.pre = B.conv(p + ".pre", -1, primed, false);
      for (int l = 0; l < hp.flow_wn_layers; ++l) {
        c.in_layers.push_back(B.conv(p + ".enc.in_layers." + std::to_string(l)));
        const std::string rs = p + ".enc.res_skip_layers." + std::to_string(l);
        if (l < hp.flow_wn_layers - 1) {
          c.res.push_back(B.conv(rs, -1, false, false, 0, H));
          c.skip.push_back(B.conv(rs, -1, false, false, H, H));
        } else {
          c.res.push_back(ConvW{});
          c.skip.push_back(B.conv(rs));
        }
      }
      c.post = B.conv(p + ".post", -1, false, primed, 0, -1, -1.f);
      if (hp.gin) {
        c.cond_w = B.upload_vec(p + ".enc.cond_layer.weight", 2 * H * hp.flow_wn_layers * hp.gin);
        c.cond_b = B.upload_vec(p + ".enc.cond_layer.bias", 2 * H * hp.flow_wn_layers);
      }
      flows.push_back(c);
    }
    // an odd number of Flips leaves z channel-reversed: fold it into conv_pre's input channels
    dec_pre = B.conv("dec.conv_pre", -1, flips % 2 == 1);
  }

  // ---- HiFi-GAN generator ------------------------------------------------------------------------------
  if (hp.gin) {
    dec_cond_w = B.upload_vec("dec.cond.weight", hp.dec_c0 * hp.gin);
    dec_cond_b = B.upload_vec("dec.cond.bias", hp.dec_c0);
  }
  hp.hop = 1;
  double rows = 1.0;   // rows per frame at the current stage
  hp.dec_flops_per_frame = 2.0 * dec_pre.cin * dec_pre.cout * dec_pre.k;
  for (int i = 0; i < n_ups; ++i) {
    StageW s;
    const std::string up = "dec.ups." + std::to_string(i);
    const int kk = (int)B.get(up + ".weight").dims[2];
    s.up = B.conv_transpose(up, kk / 2);
    s.stride = s.up.phases;
    s.ch = s.up.cout;
    hp.ups.push_back({kk, s.stride});
    hp.dec_flops_per_frame += rows * 2.0 * s.up.cin * s.up.cout * s.up.k;
    rows *= s.stride;
    hp.hop *= s.stride;
    for (int j = 0; j < nk; ++j) {
      const std::string rb = "dec.resblocks." + std::to_string(i * nk + j);
      ResBlockW r;
      int m = 0;
      while (B.has(rb + rbc + std::to_string(m) + ".weight")) {
        r.c1.push_back(B.conv(rb + rbc + std::to_string(m)));
        hp.dec_flops_per_frame += rows * 2.0 * r.c1.back().cin * r.c1.back().cout * r.c1.back().k;
        if (hp.resblock == 1) {
          r.c2.push_back(B.conv(rb + ".convs2." + std::to_string(m)));
          hp.dec_flops_per_frame += rows * 2.0 * r.c2.back().cin * r.c2.back().cout * r.c2.back().k;
        }
        ++m;
      }
      if (i == 0) {
        hp.rb_kernels.push_back(r.c1[0].k);
        std::vector<int> d;
        for (auto& c : r.c1) d.push_back(c.dil);
        hp.rb_dil.push_back(d);
      }
      s.rbs.push_back(r);
    }
    stages.push_back(s);
  }
  {
    const HostTensor& w = B.get("dec.conv_post.weight");   // [1][C][k]
    const int C = (int)w.dims[1];
    post_k = (int)w.dims[2];
    if (w.dims.size() != 3 || w.dims[0] != 1 || w.f32.size() != (size_t)w.dims[1] * (size_t)w.dims[2] ||
        C != stages.back().ch || C % 4 != 0) throw std::runtime_error("unsupported: dec.conv_post channel count");
    std::vector<float> pk((size_t)post_k * C);
    for (int c = 0; c < C; ++c)
      for (int t = 0; t < post_k; ++t) pk[(size_t)t * C + c] = w.f32[(size_t)c * post_k + t];
    post_w = B.upload(pk);
    if (B.has("dec.conv_post.bias")) {
      const HostTensor& bt = B.get("dec.conv_post.bias");
      if (bt.f32.size() != 1) throw std::runtime_error("unsupported: dec.conv_post.bias must hold one value");
      post_b = bt.f32[0];
    }
    hp.dec_flops_per_frame += rows * 2.0 * C * post_k;
  }
  // algorithmic FLOP tables
  {
    const int half = H / 2;
    double wn = 0;
    for (int l = 0; l < hp.flow_wn_layers; ++l)
      wn += 2.0 * H * 2 * H * hp.flow_kernel + 2.0 * H * (l < hp.flow_wn_layers - 1 ? 2 * H : H);
    hp.flow_flops_per_frame = hp.flow_ids.size() * (2.0 * half * H + wn + 2.0 * H * half);
    double enc = hp.n_layers * (4 * 2.0 * H * H + 2 * 2.0 * H * hp.filter * hp.ffn_kernel) + 2.0 * H * 2 * H;
    double ddsf = hp.dp_layers * (2.0 * H * hp.dp_kernel + 2.0 * H * H);
    double dp = 2 * 2.0 * H * H + ddsf + dp_flows.size() * (2.0 * H + ddsf + 2.0 * H * nparam);
    hp.enc_flops_per_id = enc + dp;
  }
  finalized = true;
}

}  // namespace piper
