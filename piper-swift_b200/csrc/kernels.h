// Launch interface of the sm_100a kernels (definitions in kernels_f32.cu / conv_tc.cu).
// Activations are "packed rows": all utterances of a batch lie one after another along
// the time axis, separated by zero gap rows; a buffer is [rows][channels] fp32
// (channels contiguous).  Every producer writes ALL rows [0,R) (zeros on gap rows), so a
// convolution's "same" zero padding and the graph's x*x_mask multiplications both fall out
// of the layout.  Rows outside [0,R) are never read (guarded).
#pragma once
#include <cuda_runtime.h>

#include <utility>
#include <cstdint>
#include <mutex>
#include <vector>

namespace piper {

// One-time set-up PER DEVICE at a launch site: cudaFuncSetAttribute (the raised dynamic shared-memory limit) belongs to
// the device that is current when it is called, so a process driving several GPUs (PiperCUDAPool: one context per
// device, one host thread each) must repeat it on each of them.  `static PerDeviceOnce once; once.run([&] {...});`
struct PerDeviceOnce {
  std::mutex mu;
  uint64_t done = 0;
  template <typename F>
  void run(F&& f) {
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    std::lock_guard<std::mutex> lk(mu);
    if (done & bit) return;
    f();
    done |= bit;
  }
};

// optional per-kernel-class device timing (PIPER_B200_FLAG_PROFILE): CUDA events around every launch of
// the three GEMM-class kernels, summed after the call's final synchronisation
struct Prof {
  enum { CONV_TC = 0, MRF = 1, CONV_F32 = 2, NCLS = 3 };
  struct Span { cudaEvent_t a, b; int cls; double flops; };
  std::vector<Span> spans;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;
  cudaEvent_t get() {
    if (used == pool.size()) {
      cudaEvent_t e = nullptr;
      cudaEventCreate(&e);
      pool.push_back(e);
    }
    return pool[used++];
  }
  void reset() { used = 0; spans.clear(); }
  void release() { for (cudaEvent_t e : pool) cudaEventDestroy(e); pool.clear(); reset(); }
};

struct Exec {
  cudaStream_t st = nullptr;
  int64_t launches = 0;
  Prof* prof = nullptr;
};

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------------------
// A synthesize call is ~145 dependent launches; at small batches (one utterance, or 32 per GPU in the 8-GPU strong-scaling
// case) the step is the sum of launch gaps and kernel prologues, not of work.  Kernels launched through launch_pdl carry
// cudaLaunchAttributeProgrammaticStreamSerialization: their CTAs may be scheduled while the preceding kernel of the stream
// still runs, execute their prologue (barrier init, TMEM allocation, tensor-map prefetch, constant staging), and block in
// PDL_WAIT() until the predecessor has completed and its writes are visible.  PDL_TRIGGER() at the top of a kernel tells
// the scheduler that the NEXT kernel may be staged.  Rules: a kernel launched through launch_pdl MUST execute PDL_WAIT()
// before its first read of anything an earlier kernel wrote (and before its first write of anything an earlier kernel
// reads); plain <<<>>> launches stay fully ordered.  PIPER_B200_PDL=0 turns the attribute off (plain launches).
#if defined(__CUDACC__)
#define PDL_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
#define PDL_TRIGGER() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
#endif
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

struct ProfScope {
  Exec& ex; cudaEvent_t b = nullptr;
  ProfScope(Exec& e, int cls, double flops) : ex(e) {
    if (!ex.prof) return;
    cudaEvent_t a = ex.prof->get();
    b = ex.prof->get();
    cudaEventRecord(a, ex.st);
    ex.prof->spans.push_back({a, b, cls, flops});
  }
  ~ProfScope() { if (b) cudaEventRecord(b, ex.st); }
};

// out[orow][co] = epi( bias[co] + uttbias[utt][co] + sum_{tap,ci} act(in[q+in_off+tap*tap_step][ci]) * w[tap][ci][co] )
//   orow = q*out_mul + out_add;  phases>1 => transposed conv, one polyphase sub-filter per blockIdx.z
struct ConvArgs {
  const float* in = nullptr;  int ldin = 0;
  const float* w = nullptr;   int ldw = 0;       // [phases][taps][cin][ldw]
  const float* bias = nullptr;
  float* out = nullptr;       int ldout = 0;
  const float* res = nullptr; int ldres = 0;     // added after activation (may alias out)
  const float* uttbias = nullptr; int ldub = 0;  // [batch][ldub] per-utterance bias (speaker conditioning)
  const int* row_utt = nullptr;                  // utterance id (>=0) or -1 per base row
  int valid_div = 1;                             // out row r is valid iff row_utt[r / valid_div] >= 0
  int R_in = 0, Q = 0;                           // input rows; number of q positions
  int cin = 0, cout = 0, taps = 1;
  int tap_step = 1, in_off = 0;
  int out_mul = 1, out_add = 0;
  int phases = 1, ct_pad = 0;                    // transposed conv: stride = phases
  int in_act = 0; float in_slope = 0.f;          // leaky-relu on the input operand
  int relu = 0;                                  // relu on (acc + bias)
  int accum = 0;                                 // out += value
  float div = 0.f;                               // if != 0: out = value / div (after accum)
};
void launch_conv_f32(Exec& ex, const ConvArgs& a);

struct LNArgs {
  const float* x = nullptr;   // [R][C]
  const float* y = nullptr;   // optional addend before the norm (x + y)
  const float* gamma = nullptr; const float* beta = nullptr;
  float* out = nullptr;       // [R][C]
  const float* res = nullptr; // optional: out = res + f(LN(.)) (DDSConv residual)
  const int* row_utt = nullptr;
  int R = 0, C = 0;
  int gelu = 0;
  // optional second output: the f16x3 split operand of `out` (conv_tc.h, TcW::split == 2), [R][2 * C] fp16 planes
  // [hi | lo'] - what the next tensor-core convolution reads (saves the separate split pass)
  uint16_t* out_split2 = nullptr;
};
void launch_layernorm(Exec& ex, const LNArgs& a);

void launch_embed(Exec& ex, const int64_t* ids_packed, const float* emb, float* x, int R, int H,
                  int n_vocab, float scale);

// relative-position multi-head attention over qkv [R][3H] (q|k|v), per utterance
// out_split2 != null: the head-dim-96 kernel writes the f16x3 split operand [R][2 * H] instead of fp32 `out`
// (other head dims: fp32 `out`, the caller splits)
void launch_attention(Exec& ex, const float* qkv, const float* rel_k, const float* rel_v, float* out,
                      const int* utt_start, const int* utt_len, int batch, int max_len, int H,
                      int n_heads, int window, uint16_t* out_split2 = nullptr, bool* wrote_split = nullptr,
                      float* scratch = nullptr, int skip_upto = 0);   // skip_upto: utterances of at most this many ids are left alone
// The same attention layer on the tensor cores (attention_tc.cu) for utterances of at most 128 ids and head dimension 96:
// q / k / v come as the f16x3 split planes [rows][2][3H] the qkv convolution writes beside its fp32 output, the
// relative-key table as split planes [16][2][96] (TcEncLayer::relk16); longer utterances are skipped (launch_attention
// with skip_upto = 128 takes them).
bool attention_tc_supported(int H, int n_heads, int window);
void launch_attention_tc(Exec& ex, const uint16_t* qkvs, const void* relk16, const float* rel_v, float* out,
                         uint16_t* out_split2, const int* utt_start, const int* utt_len, int batch, int rows, int H, int n_heads,
                         int window);
// bytes of `scratch` that launch_attention needs for `rows` packed rows when the longest utterance has max_len ids
// (0 for short utterances: keys are split over several CTAs only from 512 ids on)
size_t attention_scratch_bytes(int rows, int max_len, int H, int n_heads);

// depthwise conv (k taps, dilation d) + channel LayerNorm + GELU, DDSConv first half
void launch_dds_sep_ln_gelu(Exec& ex, const float* x, const float* w_sep /*[k][C]*/, const float* b_sep,
                            const float* gamma, const float* beta, float* out, const int* row_utt,
                            int R, int C, int k, int dil, uint16_t* out_split2 = nullptr);   // split: [R][2 * C] instead of `out`

// h[r][c] = w[c]*z[r] + b[c] + cond[r][c]   (ConvFlow.pre on one channel + DDSConv's `x + g`)
void launch_convflow_pre(Exec& ex, const float* z, const float* w, const float* b, const float* cond,
                         float* h, const int* row_utt, int R, int C);

// inverse rational-quadratic spline with linear tails on one channel; params [R][ldp] (3*bins-1 used)
void launch_spline_inverse(Exec& ex, float* z_target, const float* params, int ldp, const int* row_utt,
                           int R, int bins, float tail_bound, float inv_sqrt_filter);

// reference RandomNormalLike (elementwise.metal:132-163): per utterance gid = c*len + t
void launch_noise_dp(Exec& ex, float* z /*[2][R]*/, const float* ext /*ragged [2,T_b] or null*/,
                     const int* utt_start, const int* utt_len, const int64_t* ext_off, int batch, int R,
                     uint32_t seed, float noise_w);

// durations: logw = (z0 - m)*exp(-logs); w = exp(logw)*length_scale; ceil; cumsum; y_len
// (use_scale: exp(-logs) arrives as the constant `ea_scale` the exporter folded it into, ea_logs is unused)
void launch_durations(Exec& ex, const float* z0, float ea_m, float ea_logs, float ea_scale, bool use_scale,
                      float length_scale, const int* utt_start, const int* utt_len, int batch, float* logw_out,
                      int* dur /*[R]*/, int* cum /*[R]*/, int64_t* y_len /*[batch]*/);
// logw given directly (parity entry point)
void launch_durations_from_logw(Exec& ex, const float* logw, float length_scale, const int* utt_start,
                                const int* utt_len, int batch, int* dur, int* cum, int64_t* y_len);
// frame layout: fstart[b] = exclusive scan of (y_len + gap); total rows -> fstart[batch]
void launch_frame_layout(Exec& ex, const int64_t* y_len, int batch, int gap, int64_t* fstart);
// per frame row: utterance id (or -1) and source phoneme row (or -1)
void launch_frame_map(Exec& ex, const int64_t* fstart, const int64_t* y_len, const int* cum,
                      const int* utt_start, const int* utt_len, int batch, int Rf, int* frow_utt,
                      int* frow_src);
// z_p[r][c] = m_p[src][c] + noise * exp(logs_p[src][c]) * noise_scale
void launch_expand_prior(Exec& ex, const float* stats /*[Rp][2H]*/, const int* frow_utt,
                         const int* frow_src, const int64_t* fstart, const int64_t* y_len,
                         const float* ext_noise, int64_t ext_stride, uint32_t seed, float noise_scale,
                         float* zp, int Rf, int H);

// acts = tanh(xin[:, :H] + g) * sigmoid(xin[:, H:] + g)
void launch_wn_gate(Exec& ex, const float* xin /*[R][2H]*/, const float* cond /*[B][ldc] or null*/,
                    int ldc, const int* row_utt, float* acts, int R, int H);

// out[r] = tanh(sum_{tap,c} lrelu(in[r+tap-pad][c], slope) * w[tap][c]) and optional pcm16
void launch_conv_post_tanh(Exec& ex, const float* in, int C, const float* w, int k, float slope,
                           const int* row_utt, int valid_div, float* out, int R, float bias = 0.f);

// speaker conditioning: out[b][co] = bias[co] + sum_ci W[co][ci] * emb_g[sid[b]][ci]
void launch_cond_gemv(Exec& ex, const float* emb_g, const int64_t* sids, const float* W, const float* bias,
                      float* out, int ldo, int batch, int gin, int cout, int n_speakers);

// gather per-utterance audio out of the packed sample rows into a dense concatenation
void launch_gather_audio(Exec& ex, const float* packed, const int64_t* fstart, const int64_t* y_len,
                         const int64_t* sample_off, int batch, int hop, float* dense, int16_t* pcm16,
                         int64_t total_samples);

// int16 PCM of a sample range: Int(clamp(v, -1, 1) * 32767), truncation (WavFileWriter.swift:20-30)
void launch_pcm16(Exec& ex, const float* src, int16_t* dst, int64_t n);

void launch_pack_ids(Exec& ex, const int64_t* ids, const int64_t* id_off, const int* utt_start,
                     const int* utt_len, int batch, int64_t* ids_packed, int* row_utt, int n_vocab);
void launch_random_normal(Exec& ex, float* out, int64_t count, uint32_t seed);
void launch_fill_i32(Exec& ex, int* p, int v, int64_t n);

}  // namespace piper
