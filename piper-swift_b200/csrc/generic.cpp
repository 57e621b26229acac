// The generic op executor behind the C ABI (graph_exec.h): piper_b200_synthesize for graphs the fused pipeline does not
// recognise (or on request, PIPER_B200_FLAG_GENERIC) and piper_b200_run_graph, the counterpart of
// GraphExecutor.execute(inputs:maxNodeIndex:overrides:) (GraphExecutor.swift:75-152): run the graph, read any value.
#include <chrono>
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "pipeline.h"

namespace piper {

namespace {
bool has_name(const std::vector<std::string>& v, const char* n) {
  for (auto& s : v)
    if (s == n) return true;
  return false;
}
}  // namespace

static void run_one(Context& c, Exec& ex, const int64_t* ids, int64_t n_ids, int64_t sid, const float* scales3, uint32_t seed) {
  if (!c.onnx) throw std::runtime_error("state: the generic executor needs a model loaded with piper_b200_load_voice");
  if (!c.gexec) c.gexec.reset(new GraphExec(*c.onnx));
  const int64_t len = n_ids;
  std::vector<GraphExec::Feed> feeds;
  feeds.push_back({"input", 7, {1, n_ids}, ids});
  feeds.push_back({"input_lengths", 7, {1}, &len});
  feeds.push_back({"scales", 1, {3}, scales3});
  if (has_name(c.onnx->inputs, "sid")) feeds.push_back({"sid", 7, {1}, &sid});
  c.gexec->run(ex, feeds, seed);
}

int Context::run_graph(const int64_t* ids, int64_t n_ids, int64_t sid, const float* scales3, uint64_t seed, const char* value_name,
                       void* out, size_t cap_bytes, int32_t* dtype, int64_t* dims8, int32_t* rank) {
  if (cudaSetDevice(device) != cudaSuccess) throw std::runtime_error("cuda: cannot select the device");
  invalidate_results();
  Exec ex;
  ex.st = stream;
  run_one(*this, ex, ids, n_ids, sid, scales3, (uint32_t)(seed & 0xffffffffu));
  const std::string name = value_name ? value_name : (onnx->outputs.empty() ? "output" : onnx->outputs[0]);
  const GTensor* t = gexec->find(name);
  if (!t) throw std::runtime_error("invalid argument: the graph has no value named '" + name + "'");
  if (dtype) *dtype = t->dtype;
  if (rank) *rank = (int32_t)t->shape.size();
  if (dims8)
    for (size_t i = 0; i < 8; ++i) dims8[i] = i < t->shape.size() ? t->shape[i] : 0;
  if (out) {
    if (t->bytes() > cap_bytes)
      throw std::runtime_error("capacity: value '" + name + "' needs " + std::to_string(t->bytes()) + " bytes");
    gexec->fetch(ex, *t, out);
  }
  timings = piper_b200_timings{};
  timings.kernel_launches = ex.launches;
  return PIPER_B200_OK;
}

int Context::synthesize_generic(const piper_b200_request& rq, piper_b200_result& rs) {
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  const uint32_t unsupported_flags = PIPER_B200_FLAG_IDS_ON_DEVICE | PIPER_B200_FLAG_KEEP_ON_DEVICE | PIPER_B200_FLAG_KEEP_STAGES |
                                     PIPER_B200_FLAG_STREAM | PIPER_B200_FLAG_PCM16 | PIPER_B200_FLAG_PCM16_ONLY;
  if (rq.flags & unsupported_flags)
    throw std::runtime_error("unsupported: the generic op executor returns float host audio only (no device / stream / PCM16 modes)");
  if (rq.noise_dp || rq.noise_main)
    throw std::runtime_error("unsupported: the generic op executor draws its noise inside the graph (RandomNormalLike), it takes no overrides");
  if (rq.batch <= 0 || !rq.phoneme_ids || !rq.lengths) throw std::runtime_error("invalid argument: empty batch or null ids/lengths");
  if (cudaSetDevice(device) != cudaSuccess) throw std::runtime_error("cuda: cannot select the device");
  invalidate_results();
  const float scales[3] = {std::isnan(rq.noise_scale) ? config.noise_scale : rq.noise_scale,
                           std::isnan(rq.length_scale) ? config.length_scale : rq.length_scale,
                           std::isnan(rq.noise_w) ? config.noise_w : rq.noise_w};
  Exec ex;
  ex.st = stream;
  cudaEventRecord(ev[0], stream);
  generic_audio.clear();
  sample_offsets.assign((size_t)rq.batch + 1, 0);
  frame_lengths.assign((size_t)rq.batch, 0);
  int64_t id_off = 0;
  const int hop = voice.finalized ? voice.hp.hop : 0;
  for (int b = 0; b < rq.batch; ++b) {
    const int64_t T = rq.lengths[b];
    if (T <= 0) throw std::runtime_error("invalid argument: utterance " + std::to_string(b) + " has length " + std::to_string(T));
    run_one(*this, ex, rq.phoneme_ids + id_off, T, rq.speaker_ids ? rq.speaker_ids[b] : 0, scales, (uint32_t)(rq.seed & 0xffffffffu));
    const GTensor* t = gexec->find(onnx->outputs.empty() ? "output" : onnx->outputs[0]);
    if (!t || t->dtype != 1) throw std::runtime_error("unsupported: the graph output is not a float32 tensor");
    const size_t n = (size_t)t->numel(), at = generic_audio.size();
    generic_audio.resize(at + n);
    gexec->fetch(ex, *t, generic_audio.data() + at);
    sample_offsets[(size_t)b + 1] = (int64_t)(at + n);
    frame_lengths[(size_t)b] = hop ? (int64_t)n / hop : 0;
    id_off += T;
  }
  cudaEventRecord(ev[5], stream);
  cudaStreamSynchronize(stream);
  durations.assign((size_t)id_off, 0);   // ceil(w) is an inner value of the graph: read it with piper_b200_run_graph
  const size_t bytes = generic_audio.size() * 4;
  float* dst;
  if (out_host) {
    if (bytes > out_cap) throw std::runtime_error("capacity: the output buffer holds " + std::to_string(out_cap) + " bytes, this result needs " + std::to_string(bytes));
    dst = static_cast<float*>(out_host);
  } else {
    h_audio.reserve(bytes + 16);
    dst = static_cast<float*>(h_audio.p);
  }
  std::memcpy(dst, generic_audio.data(), bytes);
  rs = piper_b200_result{};
  rs.audio = dst;
  rs.sample_offsets = sample_offsets.data();
  rs.frame_lengths = frame_lengths.data();
  rs.durations = durations.data();
  rs.total_samples = (int64_t)generic_audio.size();
  for (int64_t f : frame_lengths) rs.total_frames += f;
  timings = piper_b200_timings{};
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ev[0], ev[5]);
  timings.gpu_ms = ms;
  timings.kernel_launches = ex.launches;
  timings.d2h_bytes = (int64_t)bytes;
  timings.wall_ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
  return PIPER_B200_OK;
}

}  // namespace piper
