// Runtime context and the fixed stage pipeline that replaces GraphExecutor.executeOutput
// (GraphExecutor.swift:156-327): no node interpretation, no per-call initializer decode, no
// value table - ~150 fused launches over packed rows, one host sync (after the length
// regulator publishes the frame counts).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "../../include/piper_b200.h"
#include "conv_mrf.h"
#include "conv_tc.h"
#include "kernels.h"
#include "piper_config.h"
#include <memory>

#include "graph_exec.h"
#include "voice.h"

namespace piper {

// grow-only device arena with bump allocation per call
struct Arena {
  char* base = nullptr;
  size_t cap = 0, off = 0;
  void reset() { off = 0; }
  void reserve(size_t bytes);   // may free + reallocate (contents are dead at that point)
  template <typename T>
  T* alloc(size_t n) {
    const size_t bytes = (n * sizeof(T) + 255) & ~size_t(255);
    if (off + bytes > cap) return nullptr;
    T* p = reinterpret_cast<T*>(base + off);
    off += bytes;
    return p;
  }
  void release();
};

struct PinnedBuf {
  void* p = nullptr;
  size_t cap = 0;
  void reserve(size_t bytes);
  void release();
};

struct StageDump {
  std::vector<float> x, stats, logw, zp, z;   // packed-row layouts
  std::vector<int> utt_start, utt_len;
  std::vector<int64_t> fstart, y_len;
  int H = 0;
  bool valid = false;
};

// 16-bit tensor-core weights of the flow + HiFi-GAN convolutions (built on first use of a
// 16-bit precision mode; back_tc.cpp)
struct TcCoupling {
  TcW pre, post;
  std::vector<TcW> in_layers;
  std::vector<TcW> rs;      // res|skip merged (last layer: skip only)
  // skip-free form: res[l] alone (l < L-1) and ONE final convolution over the concatenated gated activations of all
  // layers, post o (sum_l skip_l) folded into its weights: m = sum_l (W_post W_skip_l) acts_l + (W_post sum_l b_skip_l + b_post)
  std::vector<TcW> res;
  TcW post_fused;           // cin = L * H (w == nullptr: not built)
  float* post_fused_w32 = nullptr;   // fp32 [L*H][half] + bias [half] behind it (source of post_fused, kept for the bias)
};
struct TcResBlock {
  std::vector<TcW> c1, c2;
};
struct TcStage {
  TcW up;
  TcW up_merged;   // the same ConvTranspose as one convolution over the union of the phases' row offsets (w == nullptr: not built)
  std::vector<TcResBlock> rbs;
  // fused MRF stage (conv_mrf.cu): all resblock convolutions of the stage packed as [tap][cout][cin]
  bool mrf = false;
  void* mrf_w = nullptr;
  int mrf_taps = 0;
  int post_tap_base = -1;   // last stage: first of the conv_post taps appended to mrf_w (-1: none)
  int tap_base[MRF_MAX_RB][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
  // ResBlock1 stage as fused units (conv_mrf.h, unit_mode): one launch per (resblock, dilation) pair of convolutions;
  // rb1_base[j * units + u] = first tap of conv1, conv2 follows it
  bool rb1 = false;
  int rb1_units = 0;
  void* rb1_w = nullptr;
  std::vector<int> rb1_base;
};
// text-encoder convolutions as fp32-emulating tensor-core GEMMs (bf16x6 split operands, conv_tc.h)
struct TcEncLayer {
  TcW qkv, o, ffn1, ffn2;
  void* relk16 = nullptr;   // emb_rel_k as f16x3 split planes [16 rows][2][96] (rows beyond 2w+1 zero) for attention_tc.cu
};
struct TcVoice {
  bool built = false;
  int f16 = 0;
  std::vector<TcEncLayer> enc;   // empty when PIPER_B200_FRONT_TC=0 or a shape does not fit
  // duration predictor 1x1 convolutions in the f16x3 form (empty: CUDA-core kernels): dp_pre, dp_dds[*].c1x1, dp_proj,
  // then dp_flows[f].dds[*].c1x1 in execution order
  std::vector<TcW> sdp;
  std::vector<TcCoupling> flows;
  TcW dec_pre;
  std::vector<TcStage> stages;
  std::vector<void*> allocs;
  void build(Exec& ex, const Voice& v, int f16);
  void release();
};
TcW make_tcw(Exec& ex, std::vector<void*>& allocs, const ConvW& a, const ConvW* b, int f16);
TcW make_tcw_split(Exec& ex, std::vector<void*>& allocs, const ConvW& a, int mode);   // fp32 emulation: 1 = bf16x6, 2 = f16x3 (conv_tc.h)

struct Context {
  int device = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;   // audio of finished row ranges goes to the host here while the generator's last stage runs
  cudaEvent_t ev_chunk[8] = {};
  cudaEvent_t ev_join = nullptr;
  Voice voice;
  TcVoice tc;
  PiperConfig config;
  bool have_config = false;
  int precision = PIPER_B200_PRECISION_FP32;
  std::string last_error;

  Arena front, back;
  PinnedBuf h_in, h_meta, h_audio, h_pcm;
  std::vector<int64_t> sample_offsets, frame_lengths;
  std::vector<int32_t> durations;
  piper_b200_timings timings{};
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  StageDump dump;
  Prof prof;
  float* d_audio_dense = nullptr;   // inside `back`
  void* out_host = nullptr;         // caller-owned page-locked output (piper_b200_set_output_buffer), else h_audio / h_pcm
  size_t out_cap = 0;

  int gap_p = 0, gap_f = 0;         // zero rows between utterances (phoneme / frame level)

  // generic ONNX-op executor (graph_exec.h): the parsed model of route A is kept; a graph the fused pipeline
  // does not recognise runs through it (generic_only), a recognised one on request (PIPER_B200_FLAG_GENERIC)
  std::shared_ptr<OnnxModel> onnx;
  std::unique_ptr<GraphExec> gexec;
  bool generic_only = false;
  std::string generic_reason;
  std::vector<float> generic_audio;
  int synthesize_generic(const piper_b200_request& rq, piper_b200_result& rs);
  int run_graph(const int64_t* ids, int64_t n_ids, int64_t sid, const float* scales3, uint64_t seed, const char* value_name, void* out,
                size_t cap_bytes, int32_t* dtype, int64_t* dims8, int32_t* rank);

  // incremental decoding (PIPER_B200_FLAG_STREAM + piper_b200_stream_next): the flow output stays in `back`
  struct StreamState {
    bool active = false, use_tc = false, want_pcm = false;
    int Rf = 0, f0 = 0, halo = 0;       // packed frame rows, first row of the utterance, generator halo (frames)
    int64_t y_len = 0, next = 0;        // frames of the utterance / frames already returned
    float* z = nullptr;
    void* z16 = nullptr;
    const int* frow_utt = nullptr;
    const float* cond_dec = nullptr;
    size_t arena_mark = 0;
    int64_t chunk_offsets[2] = {0, 0};
    int64_t chunk_frames = 0;
  } sstate;
  int stream_next(int64_t max_frames, piper_b200_result& rs, int32_t* is_final);
  // every entry point that resets or regrows `front` / `back` ends a pending streamed utterance and a device-resident
  // result: the next stream_next then reports ERR_STATE instead of decoding over reused memory
  void invalidate_results() { sstate.active = false; d_audio_dense = nullptr; }

  void init(int dev);
  void destroy();
  void finalize_voice();            // voice.finalize() + derived layout constants
  int synthesize(const piper_b200_request& rq, piper_b200_result& rs);
  int selftest_conv(int mode, int cin, int cout, int k, int dil, int stride, int rows, int prec, float* out4);
  int conv1d_op(int transposed, const float* x, int64_t rows, const float* w, const float* bias, int cin, int cout, int k, int dil,
                int stride, int prec, float* out);
  int length_regulate(const float* logw, const int64_t* lengths, int batch, float length_scale,
                      int32_t* durations, int64_t* frame_lengths, int32_t* frame_index, int64_t frame_capacity);
};

// fp32 CUDA-core implementation of flow + HiFi-GAN on packed rows (pipeline_back_f32.cpp)
// The last generator stage can run as several launches over row ranges that begin at utterance starts (the rows before
// such a boundary are zero gap rows, so every valid sample is bit-identical to the single launch); after each range is
// enqueued the caller is told, and can gather and ship that range's audio on another stream while the next one computes.
struct AudioChunks {
  std::vector<int> frame_row;                          // ascending, [0] = 0, back() = rows; each one an utterance's first frame row
  std::function<void(int chunk, float* packed)> done;  // ranges 0..chunk are enqueued on ex.st; packed = [rows * hop]
};
struct BackInputs {
  const AudioChunks* chunks = nullptr;
  const float* zp = nullptr;        // [Rf][H] fp32, gap rows zero
  const int* frow_utt = nullptr;    // [Rf]
  int Rf = 0, batch = 0;
  const float* cond_flow = nullptr; // [n_flows][batch][2H*layers] or null
  const float* cond_dec = nullptr;  // [batch][C0] or null
};
// Both back ends are two steps that can also be called separately (incremental decoding, synthesize.cpp):
//   flow    : updates zp in place to the flow output z (and, for the tensor-core back end, returns the 16-bit
//             operand copy of z the decoder's first convolution reads)
//   decoder : HiFi-GAN over ANY window of packed frame rows [row0, row0 + rows): rows outside the window read
//             as zeros, so only outputs at least `decoder_halo_frames` rows inside the window are exact
struct FlowOut {
  float* z = nullptr;         // [Rf][H] fp32 (== in.zp)
  void* z16 = nullptr;        // [Rf][H] 16-bit (tensor-core back end) or null
};
struct DecoderIn {
  const AudioChunks* chunks = nullptr;
  const float* z = nullptr;   // window start
  const void* z16 = nullptr;
  const int* frow_utt = nullptr;
  int rows = 0, batch = 0;
  const float* cond_dec = nullptr;
};
FlowOut run_flow_f32(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep);
float* run_decoder_f32(Context& ctx, Exec& ex, const DecoderIn& in);
FlowOut run_flow_tc(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep);
float* run_decoder_tc(Context& ctx, Exec& ex, const DecoderIn& in);
int decoder_halo_frames(const Voice& v);   // receptive half-width of the generator, in frames
// returns packed audio [Rf*hop] (device, inside ctx.back); z_out (if non-null) receives the flow output
float* run_back_f32(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep);
size_t back_f32_bytes(const Context& ctx, int Rf, int batch);
// tcgen05 implementation (back_tc.cpp); same contract
float* run_back_tc(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep);
size_t back_tc_bytes(const Context& ctx, int Rf, int batch);

}  // namespace piper
