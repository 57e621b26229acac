// Context::synthesize - the hot path behind PiperMetalRuntime.synthesize
// (PiperMetalRuntime.swift:62-80 -> GraphExecutor.executeOutput, GraphExecutor.swift:156-327).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "pipeline.h"

namespace piper {

#define CUDA_OK(x)                                                                         \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess)                                                                 \
      throw std::runtime_error(std::string("cuda: ") + cudaGetErrorString(e_) + " at " #x); \
  } while (0)

namespace {

ConvArgs conv_args(const ConvW& w, const float* in, int ldin, float* out, int ldout, int R, const int* row_utt) {
  ConvArgs a;
  a.in = in; a.ldin = ldin;
  a.w = w.w; a.ldw = w.ldw; a.bias = w.b;
  a.out = out; a.ldout = ldout;
  a.row_utt = row_utt; a.valid_div = 1;
  a.R_in = R; a.Q = R;
  a.cin = w.cin; a.cout = w.cout; a.taps = w.taps;
  a.tap_step = w.dil; a.in_off = -w.pad_l;
  return a;
}

// a split (fp32-emulating) tensor-core convolution over the packed phoneme rows; `in` = [R][3 * cin] bf16 (bf16x6)
// or [R][2 * cin] fp16 (f16x3, which needs a second accumulator: at most 128 columns per tile)
TcConvArgs split_args(const TcW& w, const void* in, int R, const int* row_utt) {
  TcConvArgs a;
  const int planes = w.split == 2 ? 2 : 3;
  a.in = in; a.ldin = planes * w.cin; a.R_in = R; a.cin = w.cin;
  a.w = &w; a.Q = R;
  a.row_utt = row_utt; a.valid_div = 1;
  a.f16 = w.split == 2 ? 1 : 0;
  a.tap_step = w.dil; a.in_off = -w.pad_l;
  a.n_groups = 1;
  a.nt = 32;
  for (int nt : {256, 192, 128, 96, 64, 32})
    if ((w.split != 2 || nt <= 128) && w.cout % nt == 0) { a.nt = nt; break; }
  // Small calls (one utterance, or the 32 utterances per GPU of the 8-GPU strong-scaling case): with the widest column tile
  // a launch has fewer tiles than SMs and its duration is one CTA's chain (set-up, weight stream through one SM, an
  // epilogue of nt / 32 chunks).  Narrower tiles spread the weights and the epilogue over more SMs; the accumulation
  // order of an output element does not depend on the tile width, so results are bit-identical.
  {
    const int m_tiles = (R + 127) / 128;
    for (int nt : {96, 64, 32})
      if (nt < a.nt && m_tiles * (w.cout / a.nt) * 2 <= 148 && w.cout % nt == 0) a.nt = nt;
  }
  a.n_tiles = w.cout / a.nt;
  a.e[0].bias = w.bias;
  return a;
}

// DDSConv: per layer depthwise conv + LN + GELU, 1x1 conv, LN + GELU, residual
void launch_split_act(Exec& ex, int mode, const float* src, int ldsrc, int64_t R, int C, void* dst, int cpad) {
  if (mode == 2) launch_split2(ex, src, ldsrc, R, C, dst, cpad);
  else launch_split3(ex, src, ldsrc, R, C, dst, cpad);
}

// tcw != null: the 1x1 convolutions run in the f16x3 tensor-core form (tcw[i] = layer i), the depthwise kernel
// writes their split operand (t1s) directly; h_split != null: the last layer also emits the split of h
void run_dds(Exec& ex, const std::vector<DDSLayerW>& layers, float* h, float* t1, float* t2, const int* row_utt,
             int R, int C, const TcW* tcw = nullptr, uint16_t* t1s = nullptr, uint16_t* h_split = nullptr) {
  for (size_t li = 0; li < layers.size(); ++li) {
    const DDSLayerW& l = layers[li];
    if (tcw) {
      launch_dds_sep_ln_gelu(ex, h, l.w_sep, l.b_sep, l.n1.gamma, l.n1.beta, nullptr, row_utt, R, C, l.k, l.dil, t1s);
      TcConvArgs a = split_args(tcw[li], t1s, R, row_utt);
      a.e[0].out_f32 = t2; a.e[0].ld_f32 = C;
      launch_conv_tc(ex, a);
    } else {
      launch_dds_sep_ln_gelu(ex, h, l.w_sep, l.b_sep, l.n1.gamma, l.n1.beta, t1, row_utt, R, C, l.k, l.dil);
      launch_conv_f32(ex, conv_args(l.c1x1, t1, C, t2, C, R, row_utt));
    }
    LNArgs n;
    n.x = t2; n.gamma = l.n2.gamma; n.beta = l.n2.beta; n.out = h; n.res = h; n.row_utt = row_utt;
    n.R = R; n.C = C; n.gelu = 1;
    if (li + 1 == layers.size()) n.out_split2 = h_split;
    launch_layernorm(ex, n);
  }
}

double ms_between(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0.f;
  cudaEventElapsedTime(&ms, a, b);
  return ms;
}

}  // namespace

int Context::synthesize(const piper_b200_request& rq, piper_b200_result& rs) {
  using clk = std::chrono::steady_clock;
  const auto t_wall0 = clk::now();
  if (!voice.finalized) { last_error = "state: voice not finalized"; return PIPER_B200_ERR_STATE; }
  if (rq.batch <= 0 || !rq.phoneme_ids || !rq.lengths) {
    last_error = "invalid argument: empty batch or null ids/lengths";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  const bool use_tc = precision != PIPER_B200_PRECISION_FP32;
  const int want_f16 = precision == PIPER_B200_PRECISION_F16 ? 1 : 0;
  const VoiceHParams& hp = voice.hp;
  const int B = rq.batch, H = hp.hidden;
  const bool ids_dev = rq.flags & PIPER_B200_FLAG_IDS_ON_DEVICE;
  const bool keep_dev = rq.flags & PIPER_B200_FLAG_KEEP_ON_DEVICE;
  const bool keep_stages = rq.flags & PIPER_B200_FLAG_KEEP_STAGES;
  const bool pcm_only = rq.flags & PIPER_B200_FLAG_PCM16_ONLY;
  const bool want_pcm = (rq.flags & PIPER_B200_FLAG_PCM16) || pcm_only;
  const bool stream_begin = rq.flags & PIPER_B200_FLAG_STREAM;
  sstate.active = false;
  d_audio_dense = nullptr;
  if (stream_begin && (rq.batch != 1 || keep_dev || keep_stages || pcm_only)) {
    last_error = "invalid argument: PIPER_B200_FLAG_STREAM needs batch == 1 and float host output";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  if (pcm_only && (keep_dev || keep_stages)) {
    last_error = "invalid argument: PIPER_B200_FLAG_PCM16_ONLY is a host-output mode";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  if (rq.batch > 65535) {   // the batch index is a grid dimension of the per-utterance kernels
    last_error = "capacity: at most 65535 utterances per call (got " + std::to_string(rq.batch) + ")";
    return PIPER_B200_ERR_CAPACITY;
  }
  const float noise_scale = std::isnan(rq.noise_scale) ? config.noise_scale : rq.noise_scale;
  const float length_scale = std::isnan(rq.length_scale) ? config.length_scale : rq.length_scale;
  const float noise_w = std::isnan(rq.noise_w) ? config.noise_w : rq.noise_w;
  const uint32_t seed = (uint32_t)(rq.seed & 0xffffffffu);

  // ---- host-side layout of the phoneme rows ---------------------------------------------------
  std::vector<int> utt_start(B), utt_len(B);
  std::vector<int64_t> id_off(B + 1, 0);
  int64_t rows = 0;
  int max_len = 0;
  for (int b = 0; b < B; ++b) {
    const int64_t T = rq.lengths[b];
    if (T <= 0 || T > (1 << 20)) {
      last_error = "invalid argument: utterance " + std::to_string(b) + " has length " + std::to_string(T);
      return PIPER_B200_ERR_INVALID_ARGUMENT;
    }
    utt_start[b] = (int)rows;
    utt_len[b] = (int)T;
    id_off[b + 1] = id_off[b] + T;
    rows += T + (b + 1 < B ? gap_p : 0);
    max_len = std::max(max_len, (int)T);
  }
  if (rows > (1 << 26)) { last_error = "capacity: too many phoneme rows"; return PIPER_B200_ERR_CAPACITY; }
  const int Rp = (int)rows;
  const int64_t n_ids = id_off[B];
  if (hp.gin && !rq.speaker_ids) {
    last_error = "invalid argument: multi-speaker voice needs speaker_ids";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }

  try {
    CUDA_OK(cudaSetDevice(device));
    Exec ex;
    ex.st = stream;
    prof.reset();
    if (rq.flags & PIPER_B200_FLAG_PROFILE) ex.prof = &prof;
    timings = piper_b200_timings{};
    if (use_tc && (!tc.built || tc.f16 != want_f16)) {
      try {
        tc.build(ex, voice, want_f16);
      } catch (const std::exception& e) {
        tc.release();
        last_error = e.what();
        cudaGetLastError();
        return last_error.rfind("unsupported", 0) == 0 ? PIPER_B200_ERR_UNSUPPORTED : PIPER_B200_ERR_CUDA;
      }
    }
    const int L = hp.flow_wn_layers;
    const int nflow = (int)voice.flows.size();
    const int condw = hp.gin ? (H + nflow * 2 * H * L + hp.dec_c0) : 0;

    // ---- front arena -------------------------------------------------------------------------
    size_t fb = 0;
    fb += (size_t)Rp * 4 * (H * 8 + 3 * H + hp.filter + 2 * H + 32 + 2 + 1) + (size_t)Rp * 4 * 4;
    fb += (size_t)n_ids * 8 * 2 + (size_t)Rp * 8 + (size_t)B * (8 * 6 + 4 * 4) + (size_t)B * condw * 4;
    if (rq.noise_dp) fb += (size_t)n_ids * 2 * 4;
    const bool front_tc = use_tc && !tc.enc.empty();
    if (front_tc) fb += (size_t)Rp * 2 * 3 * (2 * H + hp.filter) + 3 * 256;   // split operands of x, att and the FFN mid
    // attention on the tensor cores (attention_tc.cu): the qkv convolution also writes its output as f16x3 split planes
    static const int att_tc_env = [] { const char* e = getenv("PIPER_B200_ATT_TC"); return e ? atoi(e) : 1; }();
    const bool att_tc = front_tc && att_tc_env && tc.enc[0].qkv.split == 2 && tc.enc[0].relk16 &&
                        attention_tc_supported(H, hp.n_heads, hp.window);
    if (att_tc) fb += (size_t)Rp * 6 * H * 2 + 256;
    const size_t att_scratch_bytes = attention_scratch_bytes((int)Rp, max_len, H, hp.n_heads);
    fb += att_scratch_bytes + 256;
    fb += 64 * 256;
    front.reserve(fb);
    front.reset();
    Arena& F = front;
    int* d_utt_start = F.alloc<int>(B);
    int* d_utt_len = F.alloc<int>(B);
    int64_t* d_id_off = F.alloc<int64_t>(B + 1);
    int64_t* d_sids = F.alloc<int64_t>(B);
    int64_t* d_ids = F.alloc<int64_t>(n_ids);
    int64_t* d_ids_packed = F.alloc<int64_t>(Rp);
    int* p_utt = F.alloc<int>(Rp);
    float* x = F.alloc<float>((size_t)Rp * H);
    float* qkv = F.alloc<float>((size_t)Rp * 3 * H);
    float* att = F.alloc<float>((size_t)Rp * H);
    float* y = F.alloc<float>((size_t)Rp * H);
    float* mid = F.alloc<float>((size_t)Rp * hp.filter);
    float* stats = F.alloc<float>((size_t)Rp * 2 * H);
    uint16_t* xs = front_tc ? F.alloc<uint16_t>((size_t)Rp * 3 * H) : nullptr;
    uint16_t* atts = front_tc ? F.alloc<uint16_t>((size_t)Rp * 3 * H) : nullptr;
    uint16_t* mids = front_tc ? F.alloc<uint16_t>((size_t)Rp * 3 * hp.filter) : nullptr;
    uint16_t* qkvs = att_tc ? F.alloc<uint16_t>((size_t)Rp * 6 * H) : nullptr;
    if (att_tc && !qkvs) throw std::runtime_error("capacity: front arena sizing bug (qkv split planes)");
    if (front_tc && (!xs || !atts || !mids)) throw std::runtime_error("capacity: front arena sizing bug (split operands)");
    float* dph = F.alloc<float>((size_t)Rp * H);
    float* dpc = F.alloc<float>((size_t)Rp * H);
    float* hh = F.alloc<float>((size_t)Rp * H);
    float* t1 = F.alloc<float>((size_t)Rp * H);
    float* t2 = F.alloc<float>((size_t)Rp * H);
    float* params = F.alloc<float>((size_t)Rp * 32);
    float* att_scratch = att_scratch_bytes ? F.alloc<float>(att_scratch_bytes / 4) : nullptr;
    if (att_scratch_bytes && !att_scratch) throw std::runtime_error("capacity: front arena sizing bug (attention scratch)");
    float* zdp = F.alloc<float>((size_t)Rp * 2);
    float* logw = F.alloc<float>(Rp);
    int* dur = F.alloc<int>(Rp);
    int* cum = F.alloc<int>(Rp);
    int64_t* d_ylen = F.alloc<int64_t>(B);
    int64_t* d_fstart = F.alloc<int64_t>(B + 1);
    int64_t* d_soff = F.alloc<int64_t>(B + 1);
    float* d_noise_dp = rq.noise_dp ? F.alloc<float>((size_t)n_ids * 2) : nullptr;
    float* cond = condw ? F.alloc<float>((size_t)B * condw) : nullptr;
    if (!d_soff || (rq.noise_dp && !d_noise_dp) || (condw && !cond))
      throw std::runtime_error("capacity: front arena sizing bug");

    // ---- H2D ------------------------------------------------------------------------------------
    CUDA_OK(cudaEventRecord(ev[0], stream));
    size_t h2d = 0;
    {
      const size_t meta = (size_t)B * 4 * 2 + (size_t)(B + 1) * 8 + (size_t)B * 8;
      h_in.reserve(meta + (ids_dev ? 0 : (size_t)n_ids * 8) + 64);
      char* hp_ = static_cast<char*>(h_in.p);
      std::memcpy(hp_, utt_start.data(), (size_t)B * 4);
      std::memcpy(hp_ + (size_t)B * 4, utt_len.data(), (size_t)B * 4);
      std::memcpy(hp_ + (size_t)B * 8, id_off.data(), (size_t)(B + 1) * 8);
      char* sid_h = hp_ + (size_t)B * 8 + (size_t)(B + 1) * 8;
      if (rq.speaker_ids) std::memcpy(sid_h, rq.speaker_ids, (size_t)B * 8);
      else std::memset(sid_h, 0, (size_t)B * 8);
      CUDA_OK(cudaMemcpyAsync(d_utt_start, hp_, (size_t)B * 4, cudaMemcpyHostToDevice, stream));
      CUDA_OK(cudaMemcpyAsync(d_utt_len, hp_ + (size_t)B * 4, (size_t)B * 4, cudaMemcpyHostToDevice, stream));
      CUDA_OK(cudaMemcpyAsync(d_id_off, hp_ + (size_t)B * 8, (size_t)(B + 1) * 8, cudaMemcpyHostToDevice, stream));
      CUDA_OK(cudaMemcpyAsync(d_sids, sid_h, (size_t)B * 8, cudaMemcpyHostToDevice, stream));
      h2d += meta;
      const int64_t* ids_src = rq.phoneme_ids;
      if (!ids_dev) {
        char* ids_h = sid_h + (size_t)B * 8;
        std::memcpy(ids_h, rq.phoneme_ids, (size_t)n_ids * 8);
        CUDA_OK(cudaMemcpyAsync(d_ids, ids_h, (size_t)n_ids * 8, cudaMemcpyHostToDevice, stream));
        h2d += (size_t)n_ids * 8;
        ids_src = d_ids;
      }
      if (rq.noise_dp) {
        CUDA_OK(cudaMemcpyAsync(d_noise_dp, rq.noise_dp, (size_t)n_ids * 2 * 4, cudaMemcpyHostToDevice, stream));
        h2d += (size_t)n_ids * 8;
      }
      CUDA_OK(cudaMemsetAsync(d_ids_packed, 0xff, (size_t)Rp * 8, stream));   // -1 on gap rows
      CUDA_OK(cudaMemsetAsync(p_utt, 0xff, (size_t)Rp * 4, stream));
      launch_pack_ids(ex, ids_src, d_id_off, d_utt_start, d_utt_len, B, d_ids_packed, p_utt, hp.n_vocab);
    }
    CUDA_OK(cudaEventRecord(ev[1], stream));

    // ---- speaker conditioning vectors ----------------------------------------------------------------
    float* cond_dp = nullptr;
    float* cond_flow = nullptr;
    float* cond_dec = nullptr;
    if (condw) {
      cond_dp = cond;
      cond_flow = cond + (size_t)B * H;
      cond_dec = cond_flow + (size_t)nflow * B * 2 * H * L;
      launch_cond_gemv(ex, voice.emb_g, d_sids, voice.dp_cond_w, voice.dp_cond_b, cond_dp, H, B, hp.gin, H, hp.n_speakers);
      for (int f = 0; f < nflow; ++f)
        launch_cond_gemv(ex, voice.emb_g, d_sids, voice.flows[f].cond_w, voice.flows[f].cond_b,
                         cond_flow + (size_t)f * B * 2 * H * L, 2 * H * L, B, hp.gin, 2 * H * L, hp.n_speakers);
      launch_cond_gemv(ex, voice.emb_g, d_sids, voice.dec_cond_w, voice.dec_cond_b, cond_dec, hp.dec_c0, B, hp.gin,
                       hp.dec_c0, hp.n_speakers);
    }

    // ---- text encoder ------------------------------------------------------------------
    launch_embed(ex, d_ids_packed, voice.emb, x, Rp, H, hp.n_vocab, std::sqrt((float)H));
    CUDA_OK(cudaMemsetAsync(att, 0, (size_t)Rp * H * 4, stream));
    for (size_t li = 0; front_tc && li < voice.enc.size(); ++li) {
      // tensor-core form: every large convolution reads three-term bf16 splits of its fp32 input (fp32 emulation,
      // conv_tc.h) - same graph, same fp32 tensors between the kernels
      const EncLayerW& l = voice.enc[li];
      const TcEncLayer& t = tc.enc[li];
      const int sm = t.qkv.split, planes = sm == 2 ? 2 : 3;
      const bool fused_split = sm == 2;   // f16x3: LayerNorm / attention emit the split operand themselves
      if (!fused_split || li == 0) launch_split_act(ex, sm, x, H, Rp, H, xs, H);
      if (fused_split && li == 0) CUDA_OK(cudaMemsetAsync(atts, 0, (size_t)Rp * 2 * H * 2, stream));   // gap rows
      {
        TcConvArgs a = split_args(t.qkv, xs, Rp, p_utt);
        a.e[0].out_f32 = qkv; a.e[0].ld_f32 = 3 * H;
        if (att_tc) { a.e[0].out_split = qkvs; a.e[0].ld_split = 6 * H; a.e[0].split_c = 3 * H; }
        launch_conv_tc(ex, a);
      }
      bool att_split = false;
      if (att_tc) {   // utterances of at most 128 ids on the tensor cores, longer ones through the CUDA-core kernel
        launch_attention_tc(ex, qkvs, t.relk16, l.rel_v, nullptr, atts, d_utt_start, d_utt_len, B, Rp, H, hp.n_heads, hp.window);
        att_split = true;
        if (max_len > 128)
          launch_attention(ex, qkv, l.rel_k, l.rel_v, att, d_utt_start, d_utt_len, B, max_len, H, hp.n_heads, hp.window, atts,
                           &att_split, att_scratch, 128);
      } else
      launch_attention(ex, qkv, l.rel_k, l.rel_v, att, d_utt_start, d_utt_len, B, max_len, H, hp.n_heads, hp.window,
                       fused_split ? atts : nullptr, &att_split, att_scratch);
      if (!att_split) launch_split_act(ex, sm, att, H, Rp, H, atts, H);
      {
        TcConvArgs a = split_args(t.o, atts, Rp, p_utt);
        a.e[0].out_f32 = y; a.e[0].ld_f32 = H;
        launch_conv_tc(ex, a);
      }
      LNArgs n1;
      n1.x = x; n1.y = y; n1.gamma = l.n1.gamma; n1.beta = l.n1.beta; n1.out = x; n1.row_utt = p_utt; n1.R = Rp; n1.C = H;
      if (fused_split) n1.out_split2 = xs;
      launch_layernorm(ex, n1);
      if (!fused_split) launch_split_act(ex, sm, x, H, Rp, H, xs, H);
      {
        TcConvArgs a = split_args(t.ffn1, xs, Rp, p_utt);
        a.e[0].relu = 1;
        a.e[0].out_split = mids; a.e[0].ld_split = planes * hp.filter; a.e[0].split_c = hp.filter;
        launch_conv_tc(ex, a);
      }
      {
        TcConvArgs a = split_args(t.ffn2, mids, Rp, p_utt);
        a.e[0].out_f32 = y; a.e[0].ld_f32 = H;
        launch_conv_tc(ex, a);
      }
      LNArgs n2 = n1;
      n2.gamma = l.n2.gamma; n2.beta = l.n2.beta;
      launch_layernorm(ex, n2);
    }
    for (size_t li = 0; !front_tc && li < voice.enc.size(); ++li) {
      const EncLayerW& l = voice.enc[li];
      launch_conv_f32(ex, conv_args(l.qkv, x, H, qkv, 3 * H, Rp, p_utt));
      launch_attention(ex, qkv, l.rel_k, l.rel_v, att, d_utt_start, d_utt_len, B, max_len, H, hp.n_heads, hp.window,
                       nullptr, nullptr, att_scratch);
      launch_conv_f32(ex, conv_args(l.o, att, H, y, H, Rp, p_utt));
      LNArgs n1;
      n1.x = x; n1.y = y; n1.gamma = l.n1.gamma; n1.beta = l.n1.beta; n1.out = x; n1.row_utt = p_utt; n1.R = Rp; n1.C = H;
      launch_layernorm(ex, n1);
      ConvArgs f1 = conv_args(l.ffn1, x, H, mid, hp.filter, Rp, p_utt);
      f1.relu = 1;
      launch_conv_f32(ex, f1);
      launch_conv_f32(ex, conv_args(l.ffn2, mid, hp.filter, y, H, Rp, p_utt));
      LNArgs n2 = n1;
      n2.gamma = l.n2.gamma; n2.beta = l.n2.beta;
      launch_layernorm(ex, n2);
    }
    launch_conv_f32(ex, conv_args(voice.enc_proj, x, H, stats, 2 * H, Rp, p_utt));

    // ---- stochastic duration predictor, reverse ---------------------------------------------------
    {
      // f16x3 tensor-core form of the 1x1 convolutions (TcVoice::sdp): xs still holds the split of the encoder output
      // (the last LayerNorm wrote it); atts serves as the depthwise kernel's split output, xs then as the split of dph
      const bool sdp_tc = front_tc && !tc.sdp.empty();
      const TcW* sw = sdp_tc ? tc.sdp.data() : nullptr;
      const size_t nd = voice.dp_dds.size();
      if (sdp_tc) {
        TcConvArgs a = split_args(sw[0], xs, Rp, p_utt);
        a.uttbias = cond_dp; a.ldub = H;
        a.e[0].out_f32 = dph; a.e[0].ld_f32 = H;
        launch_conv_tc(ex, a);
        run_dds(ex, voice.dp_dds, dph, t1, t2, p_utt, Rp, H, sw + 1, atts, xs);
        TcConvArgs b = split_args(sw[1 + nd], xs, Rp, p_utt);
        b.e[0].out_f32 = dpc; b.e[0].ld_f32 = H;
        launch_conv_tc(ex, b);
      } else {
        ConvArgs a = conv_args(voice.dp_pre, x, H, dph, H, Rp, p_utt);
        a.uttbias = cond_dp; a.ldub = H;
        launch_conv_f32(ex, a);
        run_dds(ex, voice.dp_dds, dph, t1, t2, p_utt, Rp, H);
        launch_conv_f32(ex, conv_args(voice.dp_proj, dph, H, dpc, H, Rp, p_utt));
      }
      size_t sdp_next = 2 + nd;
      CUDA_OK(cudaMemsetAsync(zdp, 0, (size_t)Rp * 2 * 4, stream));
      launch_noise_dp(ex, zdp, d_noise_dp, d_utt_start, d_utt_len, d_id_off, B, Rp, seed, noise_w);
      const float inv_sqrt_fc = 1.0f / std::sqrt((float)H);
      for (const ConvFlowW& f : voice.dp_flows) {
        launch_convflow_pre(ex, zdp + (size_t)(1 - f.target) * Rp, f.pre_w, f.pre_b, dpc, hh, p_utt, Rp, H);
        run_dds(ex, f.dds, hh, t1, t2, p_utt, Rp, H, sdp_tc ? sw + sdp_next : nullptr, atts);
        sdp_next += f.dds.size();
        launch_conv_f32(ex, conv_args(f.proj, hh, H, params, 32, Rp, p_utt));
        launch_spline_inverse(ex, zdp + (size_t)f.target * Rp, params, 32, p_utt, Rp, hp.dp_bins, 5.0f, inv_sqrt_fc);
      }
    }

    // ---- length regulator ----------------------------------------------------------------------
    launch_durations(ex, zdp + (size_t)voice.dp_logw_channel * Rp, voice.ea_m, voice.ea_logs, voice.ea_scale, voice.ea_has_scale,
                     length_scale,
                     d_utt_start, d_utt_len, B, logw, dur, cum, d_ylen);
    launch_frame_layout(ex, d_ylen, B, gap_f, d_fstart);
    h_meta.reserve((size_t)(2 * B + 2) * 8 + (size_t)Rp * 4);
    int64_t* h_ylen = static_cast<int64_t*>(h_meta.p);
    int64_t* h_fstart = h_ylen + B;
    int32_t* h_dur = reinterpret_cast<int32_t*>(h_fstart + B + 1);
    CUDA_OK(cudaMemcpyAsync(h_ylen, d_ylen, (size_t)B * 8, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaMemcpyAsync(h_fstart, d_fstart, (size_t)(B + 1) * 8, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaMemcpyAsync(h_dur, dur, (size_t)Rp * 4, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaEventRecord(ev[2], stream));
    CUDA_OK(cudaStreamSynchronize(stream));   // the one data-dependent host sync: frame counts size the back end
    size_t d2h = (size_t)(2 * B + 1) * 8 + (size_t)Rp * 4;

    const int64_t Rf64 = h_fstart[B] - gap_f;
    if (Rf64 <= 0 || Rf64 * hp.hop > (int64_t)2000000000) {
      last_error = "capacity: predicted durations give " + std::to_string(Rf64) + " frame rows";
      return PIPER_B200_ERR_CAPACITY;
    }
    const int Rf = (int)Rf64;
    frame_lengths.assign(h_ylen, h_ylen + B);
    sample_offsets.assign(B + 1, 0);
    int64_t total_frames = 0;
    for (int b = 0; b < B; ++b) {
      if (rq.noise_main && h_ylen[b] > rq.noise_main_stride) {
        last_error = "invalid argument: noise_main_stride " + std::to_string(rq.noise_main_stride) +
                     " < frames of utterance " + std::to_string(b) + " (" + std::to_string(h_ylen[b]) + ")";
        return PIPER_B200_ERR_INVALID_ARGUMENT;
      }
      total_frames += h_ylen[b];
      sample_offsets[b + 1] = sample_offsets[b] + h_ylen[b] * hp.hop;
    }
    const int64_t total_samples = sample_offsets[B];
    // host destination of the audio: the caller's buffer (piper_b200_set_output_buffer) or the library's pinned one
    const size_t out_bytes = (size_t)total_samples * (pcm_only ? 2 : 4);
    if (out_host && !keep_dev && !stream_begin && out_bytes > out_cap) {
      last_error = "capacity: the output buffer holds " + std::to_string(out_cap) + " bytes, this result needs " +
                   std::to_string(out_bytes);
      return PIPER_B200_ERR_CAPACITY;
    }
    durations.resize(n_ids);
    for (int b = 0; b < B; ++b)
      std::memcpy(durations.data() + id_off[b], h_dur + utt_start[b], (size_t)utt_len[b] * 4);

    // ---- back arena ---------------------------------------------------------------------------------------------
    size_t bb = use_tc ? back_tc_bytes(*this, Rf, B) : back_f32_bytes(*this, Rf, B);
    bb += (size_t)Rf * H * 4 * 2 + (size_t)Rf * 4 * 2 + (size_t)total_samples * 6 + 64 * 256;
    if (rq.noise_main) bb += (size_t)B * H * rq.noise_main_stride * 4;
    back.reserve(bb);
    back.reset();
    Arena& K = back;
    int* frow_utt = K.alloc<int>(Rf);
    int* frow_src = K.alloc<int>(Rf);
    float* zp = K.alloc<float>((size_t)Rf * H);
    float* z_keep = keep_stages ? K.alloc<float>((size_t)Rf * H) : nullptr;
    float* dense = K.alloc<float>(total_samples);
    int16_t* pcm = want_pcm ? K.alloc<int16_t>(total_samples) : nullptr;
    float* d_noise_main = nullptr;
    if (rq.noise_main) {
      d_noise_main = K.alloc<float>((size_t)B * H * rq.noise_main_stride);
      if (!d_noise_main) throw std::runtime_error("capacity: back arena sizing bug (noise)");
      CUDA_OK(cudaMemcpyAsync(d_noise_main, rq.noise_main, (size_t)B * H * rq.noise_main_stride * 4,
                              cudaMemcpyHostToDevice, stream));
      h2d += (size_t)B * H * rq.noise_main_stride * 4;
    }
    if (!dense) throw std::runtime_error("capacity: back arena sizing bug");
    CUDA_OK(cudaMemcpyAsync(d_soff, sample_offsets.data(), (size_t)(B + 1) * 8, cudaMemcpyHostToDevice, stream));

    launch_frame_map(ex, d_fstart, d_ylen, cum, d_utt_start, d_utt_len, B, Rf, frow_utt, frow_src);
    launch_expand_prior(ex, stats, frow_utt, frow_src, d_fstart, d_ylen, d_noise_main, rq.noise_main_stride, seed,
                        noise_scale, zp, Rf, H);
    if (keep_stages) {
      dump = StageDump{};
      dump.H = H;
      dump.utt_start = utt_start; dump.utt_len = utt_len;
      dump.fstart.assign(h_fstart, h_fstart + B + 1);
      dump.y_len.assign(h_ylen, h_ylen + B);
      dump.x.resize((size_t)Rp * H); dump.stats.resize((size_t)Rp * 2 * H); dump.logw.resize(Rp);
      dump.zp.resize((size_t)Rf * H); dump.z.resize((size_t)Rf * H);
      CUDA_OK(cudaMemcpyAsync(dump.x.data(), x, dump.x.size() * 4, cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaMemcpyAsync(dump.stats.data(), stats, dump.stats.size() * 4, cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaMemcpyAsync(dump.logw.data(), logw, dump.logw.size() * 4, cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaMemcpyAsync(dump.zp.data(), zp, dump.zp.size() * 4, cudaMemcpyDeviceToHost, stream));
    }
    bool audio_shipped = false;
    BackInputs bi;
    bi.zp = zp; bi.frow_utt = frow_utt; bi.Rf = Rf; bi.batch = B;
    bi.cond_flow = cond_flow; bi.cond_dec = cond_dec;
    CUDA_OK(cudaEventRecord(ev[3], stream));
    if (stream_begin) {   // flow only; the generator runs chunk by chunk in stream_next()
      const FlowOut fo = use_tc ? run_flow_tc(*this, ex, bi, nullptr) : run_flow_f32(*this, ex, bi, nullptr);
      sstate = StreamState{};
      sstate.use_tc = use_tc; sstate.want_pcm = want_pcm;
      sstate.Rf = Rf; sstate.f0 = (int)h_fstart[0]; sstate.halo = decoder_halo_frames(voice);
      sstate.y_len = h_ylen[0]; sstate.next = 0;
      sstate.z = fo.z; sstate.z16 = fo.z16; sstate.frow_utt = frow_utt; sstate.cond_dec = cond_dec;
      sstate.arena_mark = back.off;
    } else {
      // Host-bound audio of a large batch leaves in row ranges: while the generator's last stage works on range c+1,
      // range c is gathered and copied to pinned memory on the copy stream.  Ranges shrink towards the end
      // (weights K, K-1, .. 1), so that the copy left exposed after the last launch is the smallest one.
      static const int chunks_env = [] { const char* e = getenv("PIPER_B200_D2H_CHUNKS"); return e ? atoi(e) : 3; }();
      AudioChunks chunks;
      std::vector<int> chunk_utt;
      if (use_tc && !keep_dev && !keep_stages && (!want_pcm || pcm_only) && chunks_env > 1 && B >= 4 * chunks_env &&
          total_samples >= ((int64_t)1 << 22)) {
        const int Kc = std::min(chunks_env, 8);
        const double wsum = 0.5 * Kc * (Kc + 1);
        chunks.frame_row.push_back(0);
        chunk_utt.push_back(0);
        double acc = 0;
        for (int c = 0, b = 0; c + 1 < Kc; ++c) {
          acc += (Kc - c) / wsum;
          while (b < B && (double)h_fstart[b] < acc * (double)Rf) ++b;
          if (b >= B) break;
          if (b > chunk_utt.back()) { chunk_utt.push_back(b); chunks.frame_row.push_back((int)h_fstart[b]); }
        }
        chunk_utt.push_back(B);
        chunks.frame_row.push_back(Rf);
      }
      if (chunks.frame_row.size() > 2) {
        // before anything is enqueued: a growing pinned buffer synchronises
        char* dst_host = static_cast<char*>(out_host);
        if (!dst_host) {
          PinnedBuf& hb = pcm_only ? h_pcm : h_audio;
          hb.reserve(out_bytes);
          dst_host = static_cast<char*>(hb.p);
        }
        chunks.done = [&, dst_host](int c, float* packed_audio) {
          const int b0 = chunk_utt[(size_t)c], b1 = chunk_utt[(size_t)c + 1];
          const int64_t s0 = sample_offsets[(size_t)b0], s1 = sample_offsets[(size_t)b1];
          CUDA_OK(cudaEventRecord(ev_chunk[c], stream));
          CUDA_OK(cudaStreamWaitEvent(copy_stream, ev_chunk[c], 0));
          Exec exc;
          exc.st = copy_stream;
          launch_gather_audio(exc, packed_audio, d_fstart + b0, d_ylen + b0, d_soff + b0, b1 - b0, hp.hop, dense, pcm,
                              s1 - s0);
          ex.launches += exc.launches;
          if (pcm_only)
            CUDA_OK(cudaMemcpyAsync(dst_host + (size_t)s0 * 2, pcm + s0, (size_t)(s1 - s0) * 2, cudaMemcpyDeviceToHost,
                                    copy_stream));
          else
            CUDA_OK(cudaMemcpyAsync(dst_host + (size_t)s0 * 4, dense + s0, (size_t)(s1 - s0) * 4, cudaMemcpyDeviceToHost,
                                    copy_stream));
        };
        bi.chunks = &chunks;
        run_back_tc(*this, ex, bi, z_keep);
        CUDA_OK(cudaEventRecord(ev_join, copy_stream));
        CUDA_OK(cudaStreamWaitEvent(stream, ev_join, 0));
        audio_shipped = true;
      } else {
        float* packed = use_tc ? run_back_tc(*this, ex, bi, z_keep) : run_back_f32(*this, ex, bi, z_keep);
        launch_gather_audio(ex, packed, d_fstart, d_ylen, d_soff, B, hp.hop, dense, pcm, total_samples);
      }
    }
    CUDA_OK(cudaEventRecord(ev[4], stream));
    if (keep_stages) {
      CUDA_OK(cudaMemcpyAsync(dump.z.data(), z_keep, dump.z.size() * 4, cudaMemcpyDeviceToHost, stream));
      dump.valid = true;
    }

    // ---- D2H ---------------------------------------------------------------------------------------------------------
    rs = piper_b200_result{};
    if (stream_begin) {
      rs.audio = nullptr;
    } else if (keep_dev) {
      rs.audio = dense;
      rs.pcm16 = pcm;
    } else if (pcm_only) {
      void* dst = out_host;
      if (!dst) { h_pcm.reserve(out_bytes); dst = h_pcm.p; }
      if (!audio_shipped) CUDA_OK(cudaMemcpyAsync(dst, pcm, out_bytes, cudaMemcpyDeviceToHost, stream));
      d2h += out_bytes;
      rs.pcm16 = static_cast<const int16_t*>(dst);
    } else {
      void* dst = out_host;
      if (!dst) { h_audio.reserve(out_bytes); dst = h_audio.p; }
      if (!audio_shipped) CUDA_OK(cudaMemcpyAsync(dst, dense, out_bytes, cudaMemcpyDeviceToHost, stream));
      d2h += out_bytes;
      rs.audio = static_cast<const float*>(dst);
      if (want_pcm) {
        h_pcm.reserve((size_t)total_samples * 2);
        CUDA_OK(cudaMemcpyAsync(h_pcm.p, pcm, (size_t)total_samples * 2, cudaMemcpyDeviceToHost, stream));
        d2h += (size_t)total_samples * 2;
        rs.pcm16 = static_cast<const int16_t*>(h_pcm.p);
      }
    }
    CUDA_OK(cudaEventRecord(ev[5], stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    CUDA_OK(cudaGetLastError());
    if (stream_begin) sstate.active = true;
    d_audio_dense = dense;
    rs.sample_offsets = sample_offsets.data();
    rs.frame_lengths = frame_lengths.data();
    rs.durations = durations.data();
    rs.total_samples = total_samples;
    rs.total_frames = total_frames;

    timings.h2d_ms = ms_between(ev[0], ev[1]);
    timings.front_ms = ms_between(ev[1], ev[2]);
    timings.decoder_ms = ms_between(ev[3], ev[4]);   // flow + decoder; split refined by the tc back end
    timings.flow_ms = 0;
    timings.d2h_ms = ms_between(ev[4], ev[5]);
    timings.gpu_ms = ms_between(ev[0], ev[5]);
    timings.kernel_launches = ex.launches;
    timings.h2d_bytes = (int64_t)h2d;
    timings.d2h_bytes = (int64_t)d2h;
    timings.dec_flops = hp.dec_flops_per_frame * (double)total_frames;
    timings.flow_flops = hp.flow_flops_per_frame * (double)total_frames;
    // PIPER_B200_PROF_DUMP=1 (with PIPER_B200_FLAG_PROFILE): one stderr line per profiled launch, in launch order
    static const bool prof_dump = [] { const char* e = getenv("PIPER_B200_PROF_DUMP"); return e && atoi(e) != 0; }();
    for (const Prof::Span& sp : prof.spans) {
      if (prof_dump)
        fprintf(stderr, "prof cls %d ms %.4f gflop %.3f tflops %.1f\n", sp.cls, ms_between(sp.a, sp.b), sp.flops * 1e-9,
                sp.flops * 1e-9 / std::max(ms_between(sp.a, sp.b), 1e-6));
      timings.class_ms[sp.cls] += ms_between(sp.a, sp.b);
      timings.class_flops[sp.cls] += sp.flops;
      timings.class_launches[sp.cls] += 1;
    }
    timings.wall_ms = std::chrono::duration<double, std::milli>(clk::now() - t_wall0).count();
    return PIPER_B200_OK;
  } catch (const std::exception& e) {
    last_error = e.what();
    cudaGetLastError();
    return last_error.rfind("capacity", 0) == 0 ? PIPER_B200_ERR_CAPACITY : PIPER_B200_ERR_CUDA;
  }
}

// Incremental decoding: the generator over frames [next, next + max_frames) of the streamed utterance plus the
// receptive-field halo on both sides; window rows outside the packed frame range do not exist (read as zeros by
// the kernels' bounds handling), rows of the halo are recomputed by the neighbouring chunk.
int Context::stream_next(int64_t max_frames, piper_b200_result& rs, int32_t* is_final) {
  if (!sstate.active) {
    last_error = "state: no streamed utterance (call synthesize with PIPER_B200_FLAG_STREAM first)";
    return PIPER_B200_ERR_STATE;
  }
  if (max_frames <= 0 || !is_final) {
    last_error = "invalid argument: max_frames must be positive";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  using clk = std::chrono::steady_clock;
  const auto t_wall0 = clk::now();
  try {
    CUDA_OK(cudaSetDevice(device));
    Exec ex;
    ex.st = stream;
    const VoiceHParams& hp = voice.hp;
    const int H = hp.hidden;
    const int64_t n0 = sstate.next, n1 = std::min(sstate.y_len, n0 + max_frames);
    const int w0 = (int)std::max<int64_t>(0, sstate.f0 + n0 - sstate.halo);
    const int w1 = (int)std::min<int64_t>(sstate.Rf, sstate.f0 + n1 + sstate.halo);
    back.off = sstate.arena_mark;
    DecoderIn d;
    d.z = sstate.z + (size_t)w0 * H;
    d.z16 = sstate.z16 ? static_cast<const uint16_t*>(sstate.z16) + (size_t)w0 * H : nullptr;
    d.frow_utt = sstate.frow_utt + w0;
    d.rows = w1 - w0; d.batch = 1; d.cond_dec = sstate.cond_dec;
    CUDA_OK(cudaEventRecord(ev[0], stream));
    float* packed = sstate.use_tc ? run_decoder_tc(*this, ex, d) : run_decoder_f32(*this, ex, d);
    const int64_t n_samples = (n1 - n0) * hp.hop;
    const float* src = packed + (size_t)(sstate.f0 + n0 - w0) * hp.hop;
    h_audio.reserve((size_t)n_samples * 4);
    CUDA_OK(cudaMemcpyAsync(h_audio.p, src, (size_t)n_samples * 4, cudaMemcpyDeviceToHost, stream));
    if (sstate.want_pcm) {   // PCM16 of the chunk with the same rule as the one-shot call (WavFileWriter.swift:20-30)
      int16_t* d_pcm = back.alloc<int16_t>((size_t)n_samples);
      if (!d_pcm) throw std::runtime_error("capacity: back arena too small (stream pcm16)");
      launch_pcm16(ex, src, d_pcm, n_samples);
      h_pcm.reserve((size_t)n_samples * 2);
      CUDA_OK(cudaMemcpyAsync(h_pcm.p, d_pcm, (size_t)n_samples * 2, cudaMemcpyDeviceToHost, stream));
    }
    CUDA_OK(cudaEventRecord(ev[5], stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    CUDA_OK(cudaGetLastError());
    rs = piper_b200_result{};
    rs.audio = static_cast<const float*>(h_audio.p);
    if (sstate.want_pcm) rs.pcm16 = static_cast<const int16_t*>(h_pcm.p);
    sstate.chunk_offsets[0] = 0; sstate.chunk_offsets[1] = n_samples;
    sstate.chunk_frames = n1 - n0;
    rs.sample_offsets = sstate.chunk_offsets;
    rs.frame_lengths = &sstate.chunk_frames;
    rs.total_samples = n_samples;
    rs.total_frames = n1 - n0;
    sstate.next = n1;
    *is_final = n1 >= sstate.y_len ? 1 : 0;
    if (*is_final) sstate.active = false;
    timings = piper_b200_timings{};
    timings.gpu_ms = timings.decoder_ms = ms_between(ev[0], ev[5]);
    timings.kernel_launches = ex.launches;
    timings.d2h_bytes = n_samples * (sstate.want_pcm ? 6 : 4);
    timings.wall_ms = std::chrono::duration<double, std::milli>(clk::now() - t_wall0).count();
    return PIPER_B200_OK;
  } catch (const std::exception& e) {
    last_error = e.what();
    cudaGetLastError();
    sstate.active = false;
    return last_error.rfind("capacity", 0) == 0 ? PIPER_B200_ERR_CAPACITY : PIPER_B200_ERR_CUDA;
  }
}

}  // namespace piper
