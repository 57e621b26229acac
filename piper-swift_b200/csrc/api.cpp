// C ABI (include/piper_b200.h).  No exception leaves this file.
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>

#include "../../include/piper_b200.h"
#include "onnx_reader.h"
#include "pipeline.h"

using piper::Context;

struct piper_b200_ctx {
  Context c;
};

namespace {
thread_local std::string g_error;

int classify(const std::string& msg) {
  if (msg.rfind("io:", 0) == 0) return PIPER_B200_ERR_IO;
  if (msg.rfind("onnx:", 0) == 0 || msg.rfind("json:", 0) == 0 || msg.rfind("config:", 0) == 0)
    return PIPER_B200_ERR_DECODE;
  if (msg.rfind("unsupported:", 0) == 0) return PIPER_B200_ERR_UNSUPPORTED;
  if (msg.rfind("capacity", 0) == 0) return PIPER_B200_ERR_CAPACITY;
  if (msg.rfind("state:", 0) == 0) return PIPER_B200_ERR_STATE;
  if (msg.rfind("cuda:", 0) == 0) return PIPER_B200_ERR_CUDA;
  return PIPER_B200_ERR_INVALID_ARGUMENT;
}

template <typename F>
int guarded(piper_b200_ctx* ctx, F&& f) {
  try {
    return f();
  } catch (const std::exception& e) {
    if (ctx) ctx->c.last_error = e.what();
    g_error = e.what();
    return classify(e.what());
  } catch (...) {
    if (ctx) ctx->c.last_error = "unknown exception";
    g_error = "unknown exception";
    return PIPER_B200_ERR_CUDA;
  }
}
}  // namespace

extern "C" {

int piper_b200_abi_version(void) { return PIPER_B200_ABI_VERSION; }

const char* piper_b200_last_error(const piper_b200_ctx* ctx) {
  return ctx ? ctx->c.last_error.c_str() : g_error.c_str();
}

int piper_b200_device_count(int* count) {
  if (!count) {
    g_error = "invalid argument: count is null";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  *count = 0;
  const cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) {
    cudaGetLastError();
    *count = 0;
    g_error = std::string("cuda: ") + cudaGetErrorString(e);
    return PIPER_B200_ERR_CUDA;
  }
  return PIPER_B200_OK;
}

int piper_b200_create(int device, piper_b200_ctx** out) {
  if (!out) {
    g_error = "invalid argument: out is null";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  *out = nullptr;
  piper_b200_ctx* ctx = nullptr;
  int rc = guarded(nullptr, [&] {
    ctx = new piper_b200_ctx();
    ctx->c.init(device);
    return PIPER_B200_OK;
  });
  if (rc != PIPER_B200_OK) {
    if (ctx) {
      ctx->c.destroy();
      delete ctx;
    }
    return rc;
  }
  *out = ctx;
  return PIPER_B200_OK;
}

void piper_b200_destroy(piper_b200_ctx* ctx) {
  if (!ctx) return;
  try {
    ctx->c.destroy();
  } catch (...) {
  }
  delete ctx;
}

int piper_b200_load_voice(piper_b200_ctx* ctx, const char* model_path, const char* config_path) {
  if (!ctx || !model_path || !config_path) {
    g_error = "invalid argument: null context or path";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    Context& c = ctx->c;
    if (c.voice.finalized) throw std::runtime_error("state: a voice is already loaded in this context");
    std::shared_ptr<piper::OnnxModel> mp = std::make_shared<piper::OnnxModel>(piper::load_onnx(model_path));
    const piper::OnnxModel& m = *mp;
    // the graph signature the reference's loader test pins (ONNXParsingTests.swift:22-27)
    auto has = [](const std::vector<std::string>& v, const char* n) {
      for (auto& s : v)
        if (s == n) return true;
      return false;
    };
    if (!has(m.inputs, "input") || !has(m.inputs, "input_lengths") || !has(m.inputs, "scales") ||
        !has(m.outputs, "output"))
      throw std::runtime_error("unsupported: graph signature is not input/input_lengths/scales -> output");
    c.config = piper::PiperConfig::load(config_path);
    c.have_config = true;
    c.onnx = mp;
    c.gexec.reset();
    c.generic_only = false;
    try {
      c.voice.add_onnx(m);
      c.finalize_voice();
      if (c.config.num_speakers > 1 && c.voice.hp.gin == 0)
        throw std::runtime_error("unsupported: config says multi-speaker but the graph has no emb_g.weight");
    } catch (const std::exception& e) {
      // Not the Piper-VITS pattern the fused pipeline binds: like the reference, which interprets any graph of its ops
      // node by node (GraphExecutor.swift:591-2663), the context falls back to the generic op executor.  A graph with an
      // op outside that set fails there, at the first synthesize, with PIPER_B200_ERR_UNSUPPORTED.
      const std::string msg = e.what();
      if (msg.rfind("unsupported", 0) != 0) throw;
      c.voice = piper::Voice();
      c.generic_only = true;
      c.generic_reason = msg;
    }
    return PIPER_B200_OK;
  });
}

int piper_b200_set_weight(piper_b200_ctx* ctx, const char* name, int32_t dtype, const int64_t* dims, int32_t rank,
                          const void* data, size_t bytes) {
  if (!ctx || !name || (rank > 0 && !dims) || (!data && bytes)) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    Context& c = ctx->c;
    if (c.voice.finalized) throw std::runtime_error("state: voice already finalized");
    if (rank < 0 || rank > 8) throw std::runtime_error("onnx: tensor rank out of range for '" + std::string(name) + "'");
    int64_t n = 1;
    for (int i = 0; i < rank; ++i) {   // no negative and no overflowing extents (a hostile caller must not wrap the byte check)
      if (dims[i] < 0 || (dims[i] > 0 && n > ((int64_t)1 << 40) / dims[i]))
        throw std::runtime_error("onnx: bad dims for '" + std::string(name) + "'");
      n *= dims[i];
    }
    if (dtype != PIPER_B200_DTYPE_F32) {
      // int64/int32/bool initializers (sid, shape constants) carry no weights on this path
      if (dtype == PIPER_B200_DTYPE_I64 || dtype == PIPER_B200_DTYPE_I32 || dtype == PIPER_B200_DTYPE_BOOL)
        return PIPER_B200_OK;
      throw std::runtime_error("onnx: unsupported tensor data_type " + std::to_string(dtype) + " for '" + name + "'");
    }
    if ((size_t)n * 4 != bytes)
      throw std::runtime_error("onnx: raw_data size mismatch for '" + std::string(name) + "'");
    piper::HostTensor t;
    t.dims.assign(dims, dims + rank);
    t.f32.resize((size_t)n);
    std::memcpy(t.f32.data(), data, bytes);
    c.voice.tensors[name] = std::move(t);
    return PIPER_B200_OK;
  });
}

int piper_b200_set_conv_attr(piper_b200_ctx* ctx, const char* weight_name, int32_t stride, int32_t dilation,
                             int32_t group, int32_t pad_begin, int32_t pad_end) {
  if (!ctx || !weight_name) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    if (ctx->c.voice.finalized) throw std::runtime_error("state: voice already finalized");
    piper::ConvAttr a;
    a.stride = stride; a.dilation = dilation; a.group = group; a.pad_begin = pad_begin; a.pad_end = pad_end;
    ctx->c.voice.conv_attrs[weight_name] = a;
    return PIPER_B200_OK;
  });
}

int piper_b200_add_node(piper_b200_ctx* ctx, const char* op_type, const char* name, const char* const* inputs,
                        int32_t n_inputs, const char* const* outputs, int32_t n_outputs, int32_t stride, int32_t dilation,
                        int32_t group, int32_t pad_begin, int32_t pad_end) {
  if (!ctx || !op_type || n_inputs < 0 || n_outputs < 0 || (n_inputs && !inputs) || (n_outputs && !outputs)) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    if (ctx->c.voice.finalized) throw std::runtime_error("state: voice already finalized");
    piper::GraphNodeRef g;
    g.op_type = op_type;
    g.name = name ? name : "";
    for (int i = 0; i < n_inputs; ++i) g.inputs.push_back(inputs[i] ? inputs[i] : "");
    for (int i = 0; i < n_outputs; ++i) g.outputs.push_back(outputs[i] ? outputs[i] : "");
    g.attr.stride = stride > 0 ? stride : 1;
    g.attr.dilation = dilation > 0 ? dilation : 1;
    g.attr.group = group > 0 ? group : 1;
    g.attr.pad_begin = pad_begin; g.attr.pad_end = pad_end;
    ctx->c.voice.add_node(std::move(g));
    return PIPER_B200_OK;
  });
}

int piper_b200_set_config(piper_b200_ctx* ctx, int32_t sample_rate, int32_t num_speakers, float noise_scale,
                          float length_scale, float noise_w) {
  if (!ctx) {
    g_error = "invalid argument: null context";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  Context& c = ctx->c;
  c.config.sample_rate = sample_rate;
  c.config.num_speakers = num_speakers;
  c.config.noise_scale = noise_scale;
  c.config.length_scale = length_scale;
  c.config.noise_w = noise_w;
  c.have_config = true;
  return PIPER_B200_OK;
}

int piper_b200_finalize(piper_b200_ctx* ctx) {
  if (!ctx) {
    g_error = "invalid argument: null context";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    ctx->c.finalize_voice();
    return PIPER_B200_OK;
  });
}

int piper_b200_get_voice_info(const piper_b200_ctx* ctx, piper_b200_voice_info* out) {
  if (!ctx || !out) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  const Context& c = ctx->c;
  if (c.generic_only) {   // an unrecognised graph: only what the config says is known
    std::memset(out, 0, sizeof(*out));
    out->sample_rate = c.config.sample_rate;
    out->num_speakers = c.config.num_speakers;
    out->noise_scale = c.config.noise_scale;
    out->length_scale = c.config.length_scale;
    out->noise_w = c.config.noise_w;
    return PIPER_B200_OK;
  }
  if (!c.voice.finalized) {
    g_error = "state: voice not finalized";
    return PIPER_B200_ERR_STATE;
  }
  const piper::VoiceHParams& hp = c.voice.hp;
  std::memset(out, 0, sizeof(*out));
  out->sample_rate = c.config.sample_rate;
  out->num_symbols = hp.n_vocab;
  out->num_speakers = hp.n_speakers;
  out->hidden = hp.hidden;
  out->hop = hp.hop;
  out->dec_channels0 = hp.dec_c0;
  out->resblock = hp.resblock;
  out->n_ups = (int)hp.ups.size();
  out->noise_scale = c.config.noise_scale;
  out->length_scale = c.config.length_scale;
  out->noise_w = c.config.noise_w;
  out->n_params = hp.n_params;
  out->dec_flops_per_frame = hp.dec_flops_per_frame;
  out->flow_flops_per_frame = hp.flow_flops_per_frame;
  out->enc_flops_per_id = hp.enc_flops_per_id;
  return PIPER_B200_OK;
}

int piper_b200_set_precision(piper_b200_ctx* ctx, int32_t precision) {
  if (!ctx || precision < 0 || precision > PIPER_B200_PRECISION_F16) {
    g_error = "invalid argument: precision";
    if (ctx) ctx->c.last_error = g_error;
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  ctx->c.precision = precision;
  return PIPER_B200_OK;
}

int piper_b200_set_stream(piper_b200_ctx* ctx, void* cuda_stream) {
  if (!ctx) {
    g_error = "invalid argument: null context";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  ctx->c.stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->c.own_stream;
  return PIPER_B200_OK;
}

int piper_b200_set_output_buffer(piper_b200_ctx* ctx, void* host, size_t capacity_bytes) {
  if (!ctx || (host && capacity_bytes == 0)) {
    g_error = "invalid argument: null context or empty buffer";
    if (ctx) ctx->c.last_error = g_error;
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    Context& c = ctx->c;
    if (host) {   // pageable memory would turn every ranged copy into a synchronous staging copy: refuse it
      if (cudaSetDevice(c.device) != cudaSuccess) throw std::runtime_error("cuda: cudaSetDevice failed");
      cudaPointerAttributes at{};
      const cudaError_t e = cudaPointerGetAttributes(&at, host);
      if (e != cudaSuccess || at.type != cudaMemoryTypeHost) {
        cudaGetLastError();
        throw std::runtime_error("invalid argument: the output buffer is not page-locked (use piper_b200_host_alloc or "
                                 "piper_b200_host_register)");
      }
    }
    c.sstate.active = false;
    c.out_host = host;
    c.out_cap = host ? capacity_bytes : 0;
    return PIPER_B200_OK;
  });
}

int piper_b200_host_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) {
    g_error = "invalid argument: null out or zero bytes";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  *out = nullptr;
  const cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocPortable);
  if (e != cudaSuccess) {
    cudaGetLastError();
    g_error = std::string("cuda: cudaHostAlloc failed: ") + cudaGetErrorString(e);
    return PIPER_B200_ERR_CUDA;
  }
  return PIPER_B200_OK;
}

void piper_b200_host_free(void* p) {
  if (p && cudaFreeHost(p) != cudaSuccess) cudaGetLastError();
}

int piper_b200_host_register(void* p, size_t bytes) {
  if (!p || bytes == 0) {
    g_error = "invalid argument: null pointer or zero bytes";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    cudaGetLastError();
    g_error = std::string("cuda: cudaHostRegister failed: ") + cudaGetErrorString(e);
    return PIPER_B200_ERR_CUDA;
  }
  return PIPER_B200_OK;
}

int piper_b200_host_unregister(void* p) {
  if (!p) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  const cudaError_t e = cudaHostUnregister(p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    g_error = std::string("cuda: cudaHostUnregister failed: ") + cudaGetErrorString(e);
    return PIPER_B200_ERR_CUDA;
  }
  return PIPER_B200_OK;
}

int piper_b200_synthesize(piper_b200_ctx* ctx, const piper_b200_request* req, piper_b200_result* res) {
  if (!ctx || !req || !res) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    if (ctx->c.generic_only || (req->flags & PIPER_B200_FLAG_GENERIC)) return ctx->c.synthesize_generic(*req, *res);
    return ctx->c.synthesize(*req, *res);
  });
}

int piper_b200_stream_next(piper_b200_ctx* ctx, int64_t max_frames, piper_b200_result* res, int32_t* is_final) {
  if (!ctx || !res || !is_final) {
    g_error = "invalid argument: null context, result or is_final";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] { return ctx->c.stream_next(max_frames, *res, is_final); });
}

int piper_b200_get_timings(const piper_b200_ctx* ctx, piper_b200_timings* out) {
  if (!ctx || !out) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  *out = ctx->c.timings;
  return PIPER_B200_OK;
}

int piper_b200_get_stage(piper_b200_ctx* ctx, const char* name, int32_t utt, float* dst, int64_t capacity,
                         int64_t* channels, int64_t* length) {
  if (!ctx || !name || !channels || !length) {
    g_error = "invalid argument: null pointer";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    const piper::StageDump& d = ctx->c.dump;
    if (!d.valid) throw std::runtime_error("state: no stages retained (PIPER_B200_FLAG_KEEP_STAGES)");
    if (utt < 0 || utt >= (int)d.utt_len.size()) throw std::runtime_error("invalid argument: utterance index");
    const std::string n = name;
    const int H = d.H;
    const float* src = nullptr;
    int64_t C = 0, len = 0, ld = 0, row0 = 0, coff = 0;
    if (n == "x") { src = d.x.data(); C = H; ld = H; len = d.utt_len[utt]; row0 = d.utt_start[utt]; }
    else if (n == "m_p") { src = d.stats.data(); C = H; ld = 2 * H; len = d.utt_len[utt]; row0 = d.utt_start[utt]; }
    else if (n == "logs_p") { src = d.stats.data(); C = H; ld = 2 * H; coff = H; len = d.utt_len[utt]; row0 = d.utt_start[utt]; }
    else if (n == "logw") { src = d.logw.data(); C = 1; ld = 1; len = d.utt_len[utt]; row0 = d.utt_start[utt]; }
    else if (n == "z_p") { src = d.zp.data(); C = H; ld = H; len = d.y_len[utt]; row0 = d.fstart[utt]; }
    else if (n == "z") { src = d.z.data(); C = H; ld = H; len = d.y_len[utt]; row0 = d.fstart[utt]; }
    else throw std::runtime_error("invalid argument: unknown stage '" + n + "'");
    *channels = C;
    *length = len;
    if (!dst) return PIPER_B200_OK;
    if (capacity < C * len) throw std::runtime_error("invalid argument: stage buffer too small");
    for (int64_t c = 0; c < C; ++c)
      for (int64_t t = 0; t < len; ++t) dst[c * len + t] = src[(row0 + t) * ld + coff + c];
    return PIPER_B200_OK;
  });
}

int piper_b200_length_regulate(piper_b200_ctx* ctx, const float* logw, const int64_t* lengths, int32_t batch,
                               float length_scale, int32_t* durations, int64_t* frame_lengths,
                               int32_t* frame_index, int64_t frame_capacity) {
  if (!ctx) {
    g_error = "invalid argument: null context";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    return ctx->c.length_regulate(logw, lengths, batch, length_scale, durations, frame_lengths, frame_index,
                                  frame_capacity);
  });
}

int piper_b200_ipa_to_ids(piper_b200_ctx* ctx, const char* ipa_utf8, int64_t* ids, int64_t capacity, int64_t* n_ids) {
  if (!ctx || !ipa_utf8 || !n_ids || capacity < 0 || (capacity > 0 && !ids)) {
    g_error = "invalid argument";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    Context& c = ctx->c;
    if (!c.have_config || c.config.phoneme_id_map.empty())
      throw std::runtime_error("state: no phoneme_id_map (load the voice with piper_b200_load_voice)");
    const std::vector<int64_t> v = c.config.ipa_to_ids(ipa_utf8);
    *n_ids = (int64_t)v.size();
    const int64_t n = (int64_t)v.size() < capacity ? (int64_t)v.size() : capacity;
    for (int64_t i = 0; i < n; ++i) ids[i] = v[(size_t)i];
    if ((int64_t)v.size() > capacity) throw std::runtime_error("capacity: id buffer too small");
    return PIPER_B200_OK;
  });
}

int piper_b200_config_ipa_to_ids(const char* config_path, const char* ipa_utf8, int64_t* ids, int64_t capacity,
                                 int64_t* n_ids) {
  if (!config_path || !ipa_utf8 || !n_ids || capacity < 0 || (capacity > 0 && !ids)) {
    g_error = "invalid argument";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(nullptr, [&] {
    const piper::PiperConfig cfg = piper::PiperConfig::load(config_path);
    const std::vector<int64_t> v = cfg.ipa_to_ids(ipa_utf8);
    *n_ids = (int64_t)v.size();
    const int64_t n = (int64_t)v.size() < capacity ? (int64_t)v.size() : capacity;
    for (int64_t i = 0; i < n; ++i) ids[i] = v[(size_t)i];
    if ((int64_t)v.size() > capacity) throw std::runtime_error("capacity: id buffer too small");
    return PIPER_B200_OK;
  });
}

int piper_b200_random_normal(piper_b200_ctx* ctx, uint64_t seed, int64_t count, float* dst_host) {
  if (!ctx || !dst_host || count < 0) {
    g_error = "invalid argument";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    Context& c = ctx->c;
    if (cudaSetDevice(c.device) != cudaSuccess) throw std::runtime_error("cuda: cudaSetDevice failed");
    c.invalidate_results();   // the arenas are about to be reused: a pending stream / device-resident result dies here
    c.back.reserve((size_t)count * 4 + 1024);
    c.back.reset();
    float* d = c.back.alloc<float>(count);
    piper::Exec ex;
    ex.st = c.stream;
    piper::launch_random_normal(ex, d, count, (uint32_t)(seed & 0xffffffffu));
    cudaError_t e = cudaMemcpyAsync(dst_host, d, (size_t)count * 4, cudaMemcpyDeviceToHost, c.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    if (e != cudaSuccess) throw std::runtime_error(std::string("cuda: ") + cudaGetErrorString(e));
    return PIPER_B200_OK;
  });
}

int piper_b200_selftest_conv(piper_b200_ctx* ctx, int32_t mode, int32_t cin, int32_t cout, int32_t k,
                             int32_t dilation, int32_t stride, int32_t rows, int32_t precision, float* out4) {
  if (!ctx || !out4 || mode < 0 || mode > 5 || cin <= 0 || cout <= 0 || k <= 0 || dilation <= 0 || rows <= 0) {
    g_error = "invalid argument";
    if (ctx) ctx->c.last_error = g_error;
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  ctx->c.invalidate_results();
  const int rc = ctx->c.selftest_conv(mode, cin, cout, k, dilation, stride, rows, precision, out4);
  if (rc != PIPER_B200_OK) g_error = ctx->c.last_error;
  return rc;
}

int piper_b200_conv1d(piper_b200_ctx* ctx, int32_t transposed, const float* x, int64_t rows, const float* w, const float* bias,
                      int32_t cin, int32_t cout, int32_t k, int32_t dilation, int32_t stride, int32_t precision, float* out) {
  if (!ctx || !x || !w || !out || cin <= 0 || cout <= 0 || k <= 0 || dilation <= 0 || stride <= 0 || rows <= 0) {
    g_error = "invalid argument";
    if (ctx) ctx->c.last_error = g_error;
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  ctx->c.invalidate_results();
  const int rc = ctx->c.conv1d_op(transposed, x, rows, w, bias, cin, cout, k, dilation, stride, precision, out);
  if (rc != PIPER_B200_OK) g_error = ctx->c.last_error;
  return rc;
}

int piper_b200_run_graph(piper_b200_ctx* ctx, const int64_t* phoneme_ids, int64_t n_ids, int64_t speaker_id, const float* scales3,
                         uint64_t seed, const char* value_name, void* out, size_t out_capacity_bytes, int32_t* dtype, int64_t* dims8,
                         int32_t* rank) {
  if (!ctx || !phoneme_ids || n_ids <= 0 || !scales3) {
    g_error = "invalid argument: null context, ids or scales";
    if (ctx) ctx->c.last_error = g_error;
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  return guarded(ctx, [&] {
    return ctx->c.run_graph(phoneme_ids, n_ids, speaker_id, scales3, seed, value_name, out, out_capacity_bytes, dtype, dims8, rank);
  });
}

}  // extern "C"
