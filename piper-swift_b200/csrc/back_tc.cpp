// Residual-coupling flow (reverse) + HiFi-GAN generator on the tcgen05 implicit-GEMM kernel
// (conv_tc.cu): the throughput back end.  Same dataflow as back_f32.cpp -
// the graph's Slice/Split/Conv/Add/Tanh/Sigmoid/Mul/Sub/Concat and LeakyRelu/ConvTranspose/
// Conv/Add/Div nodes (GraphExecutor.swift:1739-1860, 2017-2069) - but every dense convolution
// runs on the tensor cores with 16-bit operands and fp32 accumulation, the WN gate is the
// epilogue of the in_layer convolution, res|skip is one GEMM with two column groups, and
// every producer also emits the 16-bit (leaky-relu'd) operand copy its consumer reads.
// Residual streams, skip sums and the cross-resblock mean stay in fp32.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <stdexcept>

#include "conv_tc.h"
#include "pipeline.h"

namespace piper {

namespace {

#define CUDA_OK(x)                                                                         \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess)                                                                 \
      throw std::runtime_error(std::string("cuda: ") + cudaGetErrorString(e_) + " at " #x); \
  } while (0)

int round_up(int v, int m) { return (v + m - 1) / m * m; }

// columns per CTA: at most 128 so that two CTAs fit per SM
int pick_nt(int cout) {
  for (int nt : {128, 96, 64, 48, 32, 16})
    if (cout % nt == 0) return nt;
  return 16;
}

}  // namespace

TcW make_tcw(Exec& ex, std::vector<void*>& allocs, const ConvW& a, const ConvW* b, int f16) {
  TcW t;
  t.cin = a.cin;
  t.kb = a.cin >= 64 ? 64 : 32;
  t.cin_pad = round_up(a.cin, t.kb);
  t.cout = a.cout + (b ? b->cout : 0);
  t.taps = a.taps; t.phases = a.phases; t.ct_pad = a.ct_pad; t.dil = a.dil; t.pad_l = a.pad_l;
  t.bias = a.b;
  if (!tc_conv_supported(a.cin, a.cout) || (b && (!tc_conv_supported(b->cin, b->cout) || b->cin != a.cin ||
                                                   b->taps != a.taps || b->phases != a.phases)))
    throw std::runtime_error("unsupported: convolution " + std::to_string(a.cin) + "->" + std::to_string(a.cout) +
                             " does not fit the tensor-core tile constraints (use precision fp32)");
  const size_t n = (size_t)t.phases * t.taps * t.cout * t.cin_pad;
  CUDA_OK(cudaMalloc(&t.w, n * 2));
  allocs.push_back(t.w);
  launch_pack_tc_weights(ex, a.w, a.ldw, a.phases * a.taps, a.cin, a.cout, t.w, t.cout, 0, t.cin_pad, f16);
  if (b) launch_pack_tc_weights(ex, b->w, b->ldw, b->phases * b->taps, b->cin, b->cout, t.w, t.cout, a.cout, t.cin_pad, f16);
  return t;
}

// bf16x6 fp32-emulation weights (conv_tc.h: TcW::split); cin must be a whole number of K blocks
TcW make_tcw_split(Exec& ex, std::vector<void*>& allocs, const ConvW& a, int mode) {
  TcW t;
  t.cin = a.cin;
  t.kb = a.cin % 64 == 0 ? 64 : 32;
  t.cin_pad = a.cin;
  t.cout = a.cout;
  t.taps = a.taps; t.phases = a.phases; t.ct_pad = a.ct_pad; t.dil = a.dil; t.pad_l = a.pad_l;
  t.bias = a.b;
  t.split = mode == 2 ? 2 : 1;
  if (a.cin % 32 || a.cout % 32 || a.phases != 1)
    throw std::runtime_error("unsupported: convolution " + std::to_string(a.cin) + "->" + std::to_string(a.cout) +
                             " does not fit the split tensor-core tiles");
  const size_t n = (size_t)t.taps * t.cout * t.cin_pad * (t.split == 2 ? 2 : 6);
  CUDA_OK(cudaMalloc(&t.w, n * 2));
  allocs.push_back(t.w);
  if (t.split == 2) launch_pack_tc_weights_split2(ex, a.w, a.ldw, a.taps, a.cin, a.cout, t.w, t.cout, 0, t.cin_pad);
  else launch_pack_tc_weights_split(ex, a.w, a.ldw, a.taps, a.cin, a.cout, t.w, t.cout, 0, t.cin_pad);
  return t;
}

void TcVoice::build(Exec& ex, const Voice& v, int f16_) {
  release();
  f16 = f16_;
  for (const CouplingW& c : v.flows) {
    TcCoupling t;
    t.pre = make_tcw(ex, allocs, c.pre, nullptr, f16);
    const int L = (int)c.in_layers.size();
    for (int l = 0; l < L; ++l) {
      t.in_layers.push_back(make_tcw(ex, allocs, c.in_layers[l], nullptr, f16));
      if (l < L - 1) t.rs.push_back(make_tcw(ex, allocs, c.res[l], &c.skip[l], f16));
      else t.rs.push_back(make_tcw(ex, allocs, c.skip[l], nullptr, f16));
    }
    t.post = make_tcw(ex, allocs, c.post, nullptr, f16);
    {   // skip-free form (see TcCoupling): fold post into the skip convolutions
      static const int noskip_env = [] { const char* e = getenv("PIPER_B200_FLOW_NOSKIP"); return e ? atoi(e) : 1; }();
      const int H = c.post.cin, half = c.post.cout;
      bool ok = noskip_env && L >= 1 && c.post.taps == 1 && c.post.phases == 1 && (L * H) % 64 == 0;
      for (int l = 0; ok && l < L; ++l) ok = c.skip[l].taps == 1 && c.skip[l].cin == H && c.skip[l].cout == H;
      if (ok) {
        for (int l = 0; l + 1 < L; ++l) t.res.push_back(make_tcw(ex, allocs, c.res[l], nullptr, f16));
        float* wf = nullptr;
        CUDA_OK(cudaMalloc(&wf, ((size_t)L * H * half + half) * 4));
        allocs.push_back(wf);
        float* bf = wf + (size_t)L * H * half;
        if (c.post.b) CUDA_OK(cudaMemcpyAsync(bf, c.post.b, (size_t)half * 4, cudaMemcpyDeviceToDevice, ex.st));
        else CUDA_OK(cudaMemsetAsync(bf, 0, (size_t)half * 4, ex.st));
        for (int l = 0; l < L; ++l)
          launch_fold_1x1(ex, c.skip[l].w, c.skip[l].ldw, c.skip[l].b, c.post.w, c.post.ldw, H, H, half, wf, half, l * H, bf);
        ConvW fw;
        fw.w = wf; fw.b = bf; fw.cin = L * H; fw.cout = half; fw.k = 1; fw.dil = 1; fw.ldw = half;
        t.post_fused = make_tcw(ex, allocs, fw, nullptr, f16);
        t.post_fused_w32 = wf;
      }
    }
    flows.push_back(t);
  }
  dec_pre = make_tcw(ex, allocs, v.dec_pre, nullptr, f16);
  {   // text encoder: the four large convolutions of every layer, fp32 emulated (same weights in f16 and bf16 mode)
    static const int front_env = [] { const char* e = getenv("PIPER_B200_FRONT_TC"); return e ? atoi(e) : 1; }();
    bool ok = front_env != 0;
    for (const EncLayerW& l : v.enc)
      for (const ConvW* c : {&l.qkv, &l.o, &l.ffn1, &l.ffn2})
        ok = ok && c->cin % 32 == 0 && c->cout % 32 == 0 && c->phases == 1;
    enc.clear();
    if (ok)
      for (const EncLayerW& l : v.enc) {
        TcEncLayer t;
        // PIPER_B200_FRONT_SPLIT: 2 = f16x3 (two fp16 planes, three products; default), 1 = bf16x6
        static const int split_env = [] { const char* e = getenv("PIPER_B200_FRONT_SPLIT"); return e ? atoi(e) : 2; }();
        const int mode = split_env == 1 ? 1 : 2;
        t.qkv = make_tcw_split(ex, allocs, l.qkv, mode);
        t.o = make_tcw_split(ex, allocs, l.o, mode);
        t.ffn1 = make_tcw_split(ex, allocs, l.ffn1, mode);
        t.ffn2 = make_tcw_split(ex, allocs, l.ffn2, mode);
        const int kc = v.hp.hidden / std::max(v.hp.n_heads, 1), nrel = 2 * v.hp.window + 1;
        if (mode == 2 && kc == 96 && nrel <= 16) {   // relative-key table for the tensor-core attention
          std::vector<float> rk((size_t)nrel * kc);
          CUDA_OK(cudaMemcpy(rk.data(), l.rel_k, rk.size() * 4, cudaMemcpyDeviceToHost));
          std::vector<uint16_t> h((size_t)16 * 2 * kc, 0);
          for (int m = 0; m < nrel; ++m)
            for (int d = 0; d < kc; ++d) {
              const float x = std::min(std::max(rk[(size_t)m * kc + d], -65504.f), 65504.f);
              const __half hi = __float2half_rn(x);
              const __half lo = __float2half_rn((x - __half2float(hi)) * 2048.f);
              std::memcpy(&h[(size_t)m * 2 * kc + d], &hi, 2);
              std::memcpy(&h[(size_t)m * 2 * kc + kc + d], &lo, 2);
            }
          CUDA_OK(cudaMalloc(&t.relk16, h.size() * 2));
          allocs.push_back(t.relk16);
          CUDA_OK(cudaMemcpy(t.relk16, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
        }
        enc.push_back(t);
      }
    // stochastic duration predictor: its 1x1 convolutions in the f16x3 form (PIPER_B200_SDP_TC=0: CUDA cores)
    static const int sdp_env = [] { const char* e = getenv("PIPER_B200_SDP_TC"); return e ? atoi(e) : 1; }();
    sdp.clear();
    if (ok && sdp_env && !enc.empty() && enc[0].qkv.split == 2) {
      std::vector<const ConvW*> cs;
      cs.push_back(&v.dp_pre);
      for (const DDSLayerW& l : v.dp_dds) cs.push_back(&l.c1x1);
      cs.push_back(&v.dp_proj);
      for (const ConvFlowW& f : v.dp_flows)
        for (const DDSLayerW& l : f.dds) cs.push_back(&l.c1x1);
      bool fits = true;
      for (const ConvW* c : cs) fits = fits && c->cin % 32 == 0 && c->cout % 32 == 0 && c->phases == 1 && c->taps == 1;
      if (fits)
        for (const ConvW* c : cs) sdp.push_back(make_tcw_split(ex, allocs, *c, 2));
    }
  }
  for (const StageW& s : v.stages) {
    TcStage t;
    t.up = make_tcw(ex, allocs, s.up, nullptr, f16);
    {   // merged polyphase form: column block ph * cout .. of union tap u holds W[ph][tap] when the phase's tap reads
        // input row q + first_off + u, zeros otherwise (tap t of phase ph reads row q + (ph + ct_pad) / phases - t)
      const ConvW& c = s.up;
      const int P = c.phases, T = c.taps;
      static const int merge_env = [] { const char* e = getenv("PIPER_B200_UPS_MERGED"); return e ? atoi(e) : 1; }();
      if (merge_env && P > 1 && P <= 16 && c.cout % 32 == 0 && (P * c.cout) % 128 == 0) {
        int omin = 1 << 30, omax = -(1 << 30);
        for (int ph = 0; ph < P; ++ph) {
          const int o = (ph + c.ct_pad) / P;
          omin = std::min(omin, o - (T - 1));
          omax = std::max(omax, o);
        }
        const int U = omax - omin + 1;
        TcW m = t.up;
        m.w = nullptr;
        m.taps = U; m.phases = 1; m.cout = P * c.cout;
        m.phase_cols = c.cout; m.merged_phases = P; m.first_off = omin;
        const size_t n = (size_t)U * m.cout * m.cin_pad;
        CUDA_OK(cudaMalloc(&m.w, n * 2));
        allocs.push_back(m.w);
        CUDA_OK(cudaMemsetAsync(m.w, 0, n * 2, ex.st));
        for (int ph = 0; ph < P; ++ph) {
          m.phase_tapmask[ph] = 0;
          for (int tt = 0; tt < T; ++tt) {
            const int u = (ph + c.ct_pad) / P - tt - omin;
            m.phase_tapmask[ph] |= 1u << u;
            launch_pack_tc_weights(ex, c.w + (size_t)(ph * T + tt) * c.cin * c.ldw, c.ldw, 1, c.cin, c.cout,
                                   static_cast<uint16_t*>(m.w) + (size_t)u * m.cout * m.cin_pad, m.cout, ph * c.cout,
                                   m.cin_pad, f16);
          }
        }
        t.up_merged = m;
      }
    }
    for (const ResBlockW& rb : s.rbs) {
      if (rb.c2.empty() && rb.c1.size() > 2)
        throw std::runtime_error("unsupported: ResBlock2 with more than two dilations");
      TcResBlock r;
      for (const ConvW& c : rb.c1) r.c1.push_back(make_tcw(ex, allocs, c, nullptr, f16));
      for (const ConvW& c : rb.c2) r.c2.push_back(make_tcw(ex, allocs, c, nullptr, f16));
      t.rbs.push_back(r);
    }
    {   // fused MRF stage when every resblock is a two-convolution ResBlock2 with 32, 64 or 128 channels
      const int nk = (int)s.rbs.size();
      int kk[MRF_MAX_RB] = {0, 0, 0, 0}, d0[MRF_MAX_RB] = {0, 0, 0, 0}, d1[MRF_MAX_RB] = {0, 0, 0, 0};
      bool ok = nk >= 1 && nk <= MRF_MAX_RB;
      for (int j = 0; ok && j < nk; ++j) {
        const ResBlockW& rb = s.rbs[j];
        ok = rb.c2.empty() && rb.c1.size() == 2 && rb.c1[0].k == rb.c1[1].k && rb.c1[0].cin == s.ch &&
             rb.c1[0].cout == s.ch && rb.c1[0].pad_l == rb.c1[0].dil * (rb.c1[0].k - 1) / 2 &&
             rb.c1[1].pad_l == rb.c1[1].dil * (rb.c1[1].k - 1) / 2;
        if (ok) { kk[j] = rb.c1[0].k; d0[j] = rb.c1[0].dil; d1[j] = rb.c1[1].dil; }
      }
      if (ok && mrf_tc_supported(s.ch, nk, kk, d0, d1)) {
        int taps = 0;
        for (int j = 0; j < nk; ++j)
          for (int c = 0; c < 2; ++c) { t.tap_base[j][c] = taps; taps += kk[j]; }
        const size_t tap_elems = (size_t)s.ch * s.ch;
        // last stage: conv_post joins the stage's taps (conv_mrf.h, post_tap_base) - tap t is a C x C matrix whose cout
        // row 0 holds post_w[t][:], the other rows are zero
        const bool last = &s == &v.stages.back();
        const int post_taps = (last && mrf_post_supported(s.ch, v.post_k) && v.post_w) ? v.post_k : 0;
        CUDA_OK(cudaMalloc(&t.mrf_w, (size_t)(taps + post_taps) * tap_elems * 2));
        allocs.push_back(t.mrf_w);
        if (post_taps) {
          std::vector<float> pw((size_t)post_taps * s.ch);
          CUDA_OK(cudaMemcpy(pw.data(), v.post_w, pw.size() * 4, cudaMemcpyDeviceToHost));
          std::vector<uint16_t> h((size_t)post_taps * tap_elems, 0);
          for (int tt = 0; tt < post_taps; ++tt)
            for (int c = 0; c < s.ch; ++c) {
              const float x = pw[(size_t)tt * s.ch + c];
              uint16_t b;
              if (f16) { const __half hv = __float2half_rn(x); std::memcpy(&b, &hv, 2); }
              else { const __nv_bfloat16 hv = __float2bfloat16_rn(x); std::memcpy(&b, &hv, 2); }
              h[(size_t)tt * tap_elems + c] = b;
            }
          CUDA_OK(cudaMemcpy(static_cast<uint16_t*>(t.mrf_w) + (size_t)taps * tap_elems, h.data(), h.size() * 2,
                             cudaMemcpyHostToDevice));
          t.post_tap_base = taps;
        }
        for (int j = 0; j < nk; ++j)
          for (int c = 0; c < 2; ++c) {
            const ConvW& cw = s.rbs[j].c1[c];
            launch_pack_tc_weights(ex, cw.w, cw.ldw, cw.taps, cw.cin, cw.cout,
                                   static_cast<uint16_t*>(t.mrf_w) + (size_t)t.tap_base[j][c] * tap_elems, s.ch, 0, s.ch, f16);
          }
        t.mrf_taps = taps + post_taps;
        t.mrf = true;
      }
    }
    {   // ResBlock1 stage (dilated conv -> lrelu -> plain conv, residual; three per resblock) as fused units
      static const int rb1_env = [] { const char* e = getenv("PIPER_B200_RB1_FUSED"); return e ? atoi(e) : 1; }();
      const int nk = (int)s.rbs.size();
      const int U = nk ? (int)s.rbs[0].c1.size() : 0;
      // PIPER_B200_RB1_ONLY_C=<channels>: debugging aid, fuse only the stage of that width
      static const int only_c = [] { const char* e = getenv("PIPER_B200_RB1_ONLY_C"); return e ? atoi(e) : 0; }();
      bool ok = rb1_env && !t.mrf && nk >= 1 && U >= 1 && (s.ch == 32 || s.ch == 64 || s.ch == 128) &&
                (only_c == 0 || only_c == s.ch);
      int taps = 0;
      for (int j = 0; ok && j < nk; ++j) {
        const ResBlockW& rb = s.rbs[j];
        ok = (int)rb.c1.size() == U && (int)rb.c2.size() == U;
        for (int u = 0; ok && u < U; ++u) {
          const ConvW &a = rb.c1[u], &b = rb.c2[u];
          const int k1 = a.k, dd0 = a.dil, dd1 = b.dil;
          ok = a.k == b.k && a.cin == s.ch && a.cout == s.ch && b.cin == s.ch && b.cout == s.ch && a.phases == 1 &&
               b.phases == 1 && a.pad_l == a.dil * (a.k - 1) / 2 && b.pad_l == b.dil * (b.k - 1) / 2 &&
               mrf_tc_supported(s.ch, 1, &k1, &dd0, &dd1);
          taps += 2 * a.k;
        }
      }
      if (ok) {
        const size_t tap_elems = (size_t)s.ch * s.ch;
        CUDA_OK(cudaMalloc(&t.rb1_w, (size_t)taps * tap_elems * 2));
        allocs.push_back(t.rb1_w);
        int base = 0;
        for (int j = 0; j < nk; ++j)
          for (int u = 0; u < U; ++u) {
            t.rb1_base.push_back(base);
            for (const ConvW* cw : {&s.rbs[j].c1[u], &s.rbs[j].c2[u]}) {
              launch_pack_tc_weights(ex, cw->w, cw->ldw, cw->taps, cw->cin, cw->cout,
                                     static_cast<uint16_t*>(t.rb1_w) + (size_t)base * tap_elems, s.ch, 0, s.ch, f16);
              base += cw->k;
            }
          }
        t.rb1 = true;
        t.rb1_units = U;
      }
    }
    stages.push_back(t);
  }
  CUDA_OK(cudaStreamSynchronize(ex.st));
  built = true;
}

void TcVoice::release() {
  for (void* p : allocs) cudaFree(p);
  allocs.clear();
  flows.clear();
  stages.clear();
  enc.clear();
  sdp.clear();
  built = false;
}

namespace {

size_t per_row_max_of(const Voice& v) {
  size_t m = v.hp.dec_c0;
  int U = 1;
  for (auto& s : v.stages) {
    U *= s.stride;
    m = std::max(m, (size_t)U * s.ch);
  }
  return m;
}

// plain "same" convolution over R rows
TcConvArgs tc_args(const TcW& w, const void* in, int ldin, int R, const int* row_utt, int valid_div, int f16) {
  TcConvArgs a;
  a.in = in; a.ldin = ldin; a.R_in = R; a.cin = w.cin;
  a.w = &w;
  a.Q = R;
  a.tap_step = w.dil; a.in_off = -w.pad_l;
  a.out_mul = 1;
  a.n_groups = 1;
  a.nt = pick_nt(w.cout);
  a.n_tiles = w.cout / a.nt;
  a.e[0].bias = w.bias;
  a.row_utt = row_utt; a.valid_div = valid_div;
  a.f16 = f16;
  return a;
}

// Small calls (one utterance): a launch with fewer tiles than SMs lasts as long as one CTA's chain, and two thirds of
// that chain is the epilogue of its nt / 32 column chunks.  Narrower column tiles spread the chunks over more SMs; the
// accumulation order of an output element does not depend on the tile width, so results are bit-identical (same rule as
// split_args in synthesize.cpp).  `cout` = columns per group.
void narrow_small(TcConvArgs& a, int cout) {
  const int m_tiles = (a.Q + 127) / 128;
  for (int nt : {64, 32})
    if (nt < a.nt && m_tiles * (cout / a.nt) * 2 <= 148 && cout % nt == 0) a.nt = nt;
  a.n_tiles = cout / a.nt;
}

}  // namespace

size_t back_tc_bytes(const Context& ctx, int Rf, int batch) {
  const Voice& v = ctx.voice;
  const int H = v.hp.hidden;
  const size_t P = per_row_max_of(v);
  size_t bytes = 0;
  bytes += (size_t)Rf * H * (4 * 2 + 2 * 4);       // h, skip fp32; zb, hb, actsb, skipb 16-bit
  bytes += (size_t)Rf * H * 2 * (size_t)std::max(0, v.hp.flow_wn_layers - 1);   // gated activations of every WN layer
  bytes += (size_t)Rf * P * (4 * 3 + 2 * 5);       // xu, yb, xs fp32; actA, actB, xub, ybb, tbb 16-bit
  bytes += 3 * (128 * 1280 * 2 + 4096);            // pad rows of the planar operand layouts
  bytes += (size_t)H * 2 * 256;                    // pad rows of the flow's group-planar residual copy
  bytes += (size_t)Rf * v.hp.hop * 4;              // packed audio
  bytes += 32 * 256;
  (void)batch;
  return bytes;
}

FlowOut run_flow_tc(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep) {
  Voice& v = ctx.voice;
  TcVoice& tv = ctx.tc;
  Arena& A = ctx.back;
  const int f16 = tv.f16;
  const int H = v.hp.hidden, Rf = in.Rf;
  const int L = v.hp.flow_wn_layers;
  float* z = const_cast<float*>(in.zp);   // fp32, updated in place one half at a time
  float* h = A.alloc<float>((size_t)Rf * H);
  float* skip = A.alloc<float>((size_t)Rf * H);
  uint16_t* zb = A.alloc<uint16_t>((size_t)Rf * H);
  // The 16-bit copy of the WN residual stream is group-planar [H / 8][hb_rows][8] when every in_layer can take its
  // operand as a row window (conv_tc.h, TcConvArgs::in_planar): a k = 5 in_layer then fetches each row once instead
  // of once per tap (-31 % L2 -> SM bytes).  Off by default: the launch time did not change (DESIGN.md section 10 -
  // those launches are bound by the MMA / epilogue interplay inside the SM, not by operand bytes).
  static const int flow_win_env = [] { const char* e = getenv("PIPER_B200_FLOW_WINDOW"); return e ? atoi(e) : 0; }();
  bool hb_planar = flow_win_env != 0 && H % 64 == 0 && !tv.flows.empty();
  for (const TcCoupling& t : tv.flows)
    for (const TcW& w : t.in_layers)
      if (w.taps < 2 || w.kb != 64 || w.split || w.cin != w.cin_pad || w.cin != H || (w.taps - 1) * w.dil > 64 || w.pad_l > 8 ||
          w.phases != 1)
        hb_planar = false;
  const int hb_pad = 8;
  const int64_t hb_rows = hb_planar ? (int64_t)Rf + hb_pad + 248 : Rf;
  uint16_t* hb = A.alloc<uint16_t>((size_t)hb_rows * H);
  const bool noskip = !tv.flows.empty() && tv.flows[0].post_fused.w != nullptr;
  const int LA = noskip ? L : 1;                       // the skip-free form keeps the gated activations of every layer
  uint16_t* actsb = A.alloc<uint16_t>((size_t)Rf * H * LA);
  uint16_t* skipb = A.alloc<uint16_t>((size_t)Rf * H);
  if (!h || !skip || !zb || !hb || !actsb || !skipb) throw std::runtime_error("capacity: back arena too small (flow)");

  // ---- flow ------------------------------------------------------------------------------
  if (hb_planar) {   // zero rows around the signal: taps of the first / last row tile read them
    CUDA_OK(cudaMemset2DAsync(hb, (size_t)hb_rows * 16, 0, (size_t)hb_pad * 16, H / 8, ex.st));
    CUDA_OK(cudaMemset2DAsync(hb + ((size_t)hb_pad + Rf) * 8, (size_t)hb_rows * 16, 0, (size_t)248 * 16, H / 8, ex.st));
  }
  auto hb_out = [&](TcEpi& e) {
    e.out_act = hb; e.ld_act = H;
    if (hb_planar) { e.act_planar = 1; e.act_rows = hb_rows; e.act_pad = hb_pad; }
  };
  launch_to_act16(ex, z, zb, (int64_t)Rf * H, 1.f, f16);
  for (size_t f = 0; f < v.flows.size(); ++f) {
    const CouplingW& c = v.flows[f];
    const TcCoupling& t = tv.flows[f];
    const float* cond = in.cond_flow ? in.cond_flow + f * (size_t)in.batch * 2 * H * L : nullptr;
    {
      TcConvArgs a = tc_args(t.pre, zb + c.in_off, H, Rf, in.frow_utt, 1, f16);
      a.e[0].out_f32 = h; a.e[0].ld_f32 = H;
      hb_out(a.e[0]);
      narrow_small(a, t.pre.cout);
      launch_conv_tc(ex, a);
    }
    for (int l = 0; l < L; ++l) {
      {   // in_layer + (cond) + tanh*sigmoid gate
        TcConvArgs a = tc_args(t.in_layers[l], hb, H, Rf, in.frow_utt, 1, f16);
        if (hb_planar) { a.in_planar = hb; a.in_rows = hb_rows; a.in_pad = hb_pad; }
        a.gate = 1; a.n_groups = 2;
        a.nt = pick_nt(H);
        a.n_tiles = H / a.nt;
        narrow_small(a, H);
        a.e[0].row0 = 0; a.e[0].bias = t.in_layers[l].bias;
        a.e[1].row0 = H; a.e[1].bias = t.in_layers[l].bias + H;
        a.e[0].out_act = actsb + (noskip ? l * H : 0); a.e[0].ld_act = H * LA;
        if (cond) {
          a.uttbias = cond; a.ldub = 2 * H * L;
          a.e[0].ub_off = l * 2 * H; a.e[1].ub_off = l * 2 * H + H;
        }
        launch_conv_tc(ex, a);
      }
      if (noskip) {   // res only: x = (x + res_acts) * mask; the skip path is folded into the final convolution
        if (l < L - 1) {
          TcConvArgs a = tc_args(t.res[l], actsb + l * H, H * LA, Rf, in.frow_utt, 1, f16);
          a.e[0].out_f32 = h; a.e[0].ld_f32 = H; a.e[0].res = h; a.e[0].ld_res = H;
          hb_out(a.e[0]);
          narrow_small(a, t.res[l].cout);
          launch_conv_tc(ex, a);
        }
      } else {   // res | skip
        const bool last = l == L - 1;
        TcConvArgs a = tc_args(t.rs[l], actsb, H, Rf, in.frow_utt, 1, f16);
        a.nt = pick_nt(H); a.n_tiles = H / a.nt;
        if (!last) {
          a.n_groups = 2;
          a.e[0].row0 = 0; a.e[0].bias = c.res[l].b;
          a.e[0].out_f32 = h; a.e[0].ld_f32 = H; a.e[0].res = h; a.e[0].ld_res = H;   // x = (x + res_acts) * mask
          hb_out(a.e[0]);
          a.e[1].row0 = H; a.e[1].bias = c.skip[l].b;
          a.e[1].out_f32 = skip; a.e[1].ld_f32 = H; a.e[1].accum = l > 0;             // output += skip_acts
        } else {
          a.e[0].row0 = 0; a.e[0].bias = c.skip[l].b;
          a.e[0].out_f32 = skip; a.e[0].ld_f32 = H; a.e[0].accum = l > 0;
          a.e[0].out_act = skipb; a.e[0].ld_act = H;
        }
        launch_conv_tc(ex, a);
      }
    }
    {   // post (weights negated): x1 + (-m), masked
      TcConvArgs a = noskip ? tc_args(t.post_fused, actsb, H * LA, Rf, in.frow_utt, 1, f16)
                            : tc_args(t.post, skipb, H, Rf, in.frow_utt, 1, f16);
      a.e[0].out_f32 = z + c.out_off; a.e[0].ld_f32 = H;
      a.e[0].res = z + c.out_off; a.e[0].ld_res = H;
      a.e[0].out_act = zb + c.out_off; a.e[0].ld_act = H;
      narrow_small(a, noskip ? t.post_fused.cout : t.post.cout);
      launch_conv_tc(ex, a);
    }
  }
  if (z_keep) CUDA_OK(cudaMemcpyAsync(z_keep, z, (size_t)Rf * H * 4, cudaMemcpyDeviceToDevice, ex.st));
  FlowOut out;
  out.z = z; out.z16 = zb;
  return out;
}

float* run_back_tc(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep) {
  const FlowOut f = run_flow_tc(ctx, ex, in, z_keep);
  DecoderIn d;
  d.z = f.z; d.z16 = f.z16; d.frow_utt = in.frow_utt; d.rows = in.Rf; d.batch = in.batch; d.cond_dec = in.cond_dec;
  d.chunks = in.chunks;
  return run_decoder_tc(ctx, ex, d);
}

// ---- HiFi-GAN ------------------------------------------------------------------------------
float* run_decoder_tc(Context& ctx, Exec& ex, const DecoderIn& in) {
  Voice& v = ctx.voice;
  TcVoice& tv = ctx.tc;
  Arena& A = ctx.back;
  const int f16 = tv.f16;
  const int H = v.hp.hidden, Rf = in.rows;
  const uint16_t* zb = static_cast<const uint16_t*>(in.z16);
  const size_t P = per_row_max_of(v);
  float* xu = A.alloc<float>((size_t)Rf * P);
  float* yb = A.alloc<float>((size_t)Rf * P);
  float* xs = A.alloc<float>((size_t)Rf * P);
  uint16_t* act[2] = {A.alloc<uint16_t>((size_t)Rf * P), A.alloc<uint16_t>((size_t)Rf * P)};
  uint16_t* xub = A.alloc<uint16_t>((size_t)Rf * P + 128 * 1280);   // + pad rows of the planar layout
  uint16_t* ybb = A.alloc<uint16_t>((size_t)Rf * P + 128 * 1280);   // also the ping / pong planar operands of ResBlock1 units
  uint16_t* tbb = A.alloc<uint16_t>((size_t)Rf * P + 128 * 1280);
  float* audio = A.alloc<float>((size_t)Rf * v.hp.hop);
  if (!xu || !yb || !xs || !act[0] || !act[1] || !xub || !ybb || !tbb || !audio)
    throw std::runtime_error("capacity: back arena too small (decoder)");

  int cur = 0;
  bool post_fused = false, chunked = false;
  {
    TcConvArgs a = tc_args(tv.dec_pre, zb, H, Rf, in.frow_utt, 1, f16);
    a.uttbias = in.cond_dec; a.ldub = v.hp.dec_c0;
    a.e[0].out_act = act[cur]; a.e[0].ld_act = v.hp.dec_c0; a.e[0].act_slope = 0.1f;
    narrow_small(a, tv.dec_pre.cout);
    launch_conv_tc(ex, a);
  }
  int R = Rf, U = 1;
  const int nk = (int)v.stages[0].rbs.size();
  for (size_t si = 0; si < v.stages.size(); ++si) {
    const StageW& s = v.stages[si];
    const TcStage& ts = tv.stages[si];
    const bool last_stage = si + 1 == v.stages.size();
    const int Rn = R * s.stride;
    U *= s.stride;
    const int C = s.ch;
    static const int mrf_env = [] { const char* e = getenv("PIPER_B200_MRF"); return e ? atoi(e) : 1; }();
    static const int c128_env = [] { const char* e = getenv("PIPER_B200_MRF_C128"); return e ? atoi(e) : 1; }();
    const bool use_rb1 = ts.rb1;
    const bool use_mrf = (ts.mrf && mrf_env && (C < 128 || c128_env)) || use_rb1;   // both read the planar lrelu(x) operand
    // fused stages read lrelu(x) as a group-planar tensor [C/8][act_rows][8] with zero pad rows around the signal
    const int act_pad = 128;
    const int64_t act_rows = (int64_t)Rn + act_pad + 1152;
    if (use_mrf) {
      CUDA_OK(cudaMemset2DAsync(xub, (size_t)act_rows * 16, 0, (size_t)act_pad * 16, C / 8, ex.st));
      CUDA_OK(cudaMemset2DAsync(xub + ((size_t)act_pad + Rn) * 8, (size_t)act_rows * 16, 0, (size_t)1152 * 16, C / 8, ex.st));
    }
    {   // LeakyRelu(0.1) [already in the operand copy] + ConvTranspose, one polyphase GEMM per phase
      static const int persist_env = [] { const char* e = getenv("PIPER_B200_TC_PERSIST"); return e ? atoi(e) : 1; }();
      const bool merged = ts.up_merged.w != nullptr && persist_env;
      TcConvArgs a;
      a.in = act[cur]; a.ldin = ts.up.cin; a.R_in = R; a.cin = ts.up.cin;
      a.Q = R; a.out_mul = s.stride;
      a.n_groups = 1;
      if (merged) {   // all phases of a row tile in one pass: N = phases * C columns, union of the row offsets as taps
        a.w = &ts.up_merged;
        a.tap_step = 1; a.in_off = ts.up_merged.first_off;
        const int total = ts.up_merged.cout;
        a.nt = total % 256 == 0 ? 256 : 128;
        a.n_tiles = total / a.nt;
      } else {
        a.w = &ts.up;
        a.tap_step = -1;
        a.nt = pick_nt(C); a.n_tiles = C / a.nt;
      }
      a.e[0].bias = ts.up.bias;
      a.e[0].out_f32 = use_mrf ? nullptr : xu; a.e[0].ld_f32 = C;   // the fused stage takes its residual from the 16-bit copy
      a.e[0].out_act = xub; a.e[0].ld_act = C; a.e[0].act_slope = 0.1f;
      if (use_mrf) { a.e[0].act_planar = 1; a.e[0].act_rows = act_rows; a.e[0].act_pad = act_pad; }
      a.row_utt = in.frow_utt; a.valid_div = U;
      a.f16 = f16;
      launch_conv_tc(ex, a);
    }
    uint16_t* xsb = act[cur ^ 1];
    if (use_rb1) {   // ResBlock1: one fused launch per (resblock, dilation) unit; the unit chain runs through 16-bit planar
                     // operands (ping / pong), the cross-resblock mean accumulates in xs
      uint16_t* pp[2] = {ybb, tbb};
      for (int b = 0; b < 2; ++b) {
        CUDA_OK(cudaMemset2DAsync(pp[b], (size_t)act_rows * 16, 0, (size_t)act_pad * 16, C / 8, ex.st));
        CUDA_OK(cudaMemset2DAsync(pp[b] + ((size_t)act_pad + Rn) * 8, (size_t)act_rows * 16, 0, (size_t)1152 * 16, C / 8, ex.st));
      }
      const int Uu = ts.rb1_units;
      for (int j = 0; j < nk; ++j) {
        const uint16_t* src = xub;
        for (int u = 0; u < Uu; ++u) {
          const ConvW &c1 = s.rbs[j].c1[u], &c2 = s.rbs[j].c2[u];
          MrfArgs a;
          a.R = Rn; a.C = C; a.nk = 1; a.unit_mode = 1;
          a.k[0] = c1.k; a.d0[0] = c1.dil; a.d1[0] = c2.dil;
          a.tap_base[0][0] = 0; a.tap_base[0][1] = c1.k;            // this unit's 2k taps only
          a.b1[0] = c1.b; a.b2[0] = c2.b;
          a.w = static_cast<const uint16_t*>(ts.rb1_w) + (size_t)ts.rb1_base[j * Uu + u] * C * C;
          a.total_taps = c1.k + c2.k;
          a.x_act = src; a.act_rows = act_rows; a.act_pad = act_pad;
          a.row_utt = in.frow_utt; a.valid_div = U;
          a.f16 = f16;
          if (u + 1 < Uu) {
            a.out_planar = pp[u & 1]; a.out_rows = act_rows; a.out_pad = act_pad; a.act_slope = 0.1f;
            src = pp[u & 1];
          } else {
            a.out_f32 = xs; a.accum = j > 0;
            if (j == nk - 1) {
              a.out_div = (float)nk;
              if (!last_stage) { a.out_f32 = j > 0 ? xs : nullptr; a.out_act = xsb; a.act_slope = 0.1f; }
            }
          }
          launch_mrf_tc(ex, a);
        }
      }
    } else if (use_mrf) {   // whole MRF stage in one persistent kernel, intermediates on chip
      MrfArgs a;
      a.R = Rn; a.C = C; a.nk = nk;
      for (int j = 0; j < nk; ++j) {
        a.k[j] = s.rbs[j].c1[0].k; a.d0[j] = s.rbs[j].c1[0].dil; a.d1[j] = s.rbs[j].c1[1].dil;
        a.tap_base[j][0] = ts.tap_base[j][0]; a.tap_base[j][1] = ts.tap_base[j][1];
        a.b1[j] = s.rbs[j].c1[0].b; a.b2[j] = s.rbs[j].c1[1].b;
      }
      a.w = ts.mrf_w; a.total_taps = ts.mrf_taps;
      a.x_act = xub; a.act_rows = act_rows; a.act_pad = act_pad;
      a.row_utt = in.frow_utt; a.valid_div = U;
      static const int post_env = [] { const char* e = getenv("PIPER_B200_MRF_POST"); return e ? atoi(e) : 1; }();
      if (last_stage && post_env && mrf_post_supported(C, v.post_k)) {   // conv_post + tanh as the stage kernel's epilogue
        a.post_w = v.post_w; a.post_k = v.post_k; a.post_slope = 0.01f; a.post_out = audio; a.post_bias = v.post_b;
        a.post_tap_base = ts.post_tap_base;
        post_fused = true;
      } else if (last_stage) a.out_f32 = xs;
      else { a.out_act = xsb; a.act_slope = 0.1f; }
      a.f16 = f16;
      const AudioChunks* ch = in.chunks;
      if (post_fused && ch && ch->frame_row.size() > 2 && ch->frame_row.front() == 0 && ch->frame_row.back() == Rf) {
        // row ranges between utterance starts: the window origin, the validity table and the output move with the range,
        // the planar operand keeps its plane stride; rows before a range start are gap rows (zero either way)
        chunked = true;
        for (size_t c = 0; c + 1 < ch->frame_row.size(); ++c) {
          const int f0 = ch->frame_row[c], f1 = ch->frame_row[c + 1];
          if (f1 <= f0) throw std::runtime_error("decoder: audio chunk boundaries are not ascending");
          MrfArgs b = a;
          b.R = (f1 - f0) * U;
          b.act_pad = act_pad + f0 * U;
          b.row_utt = in.frow_utt + f0;
          b.post_out = audio + (size_t)f0 * U;
          b.no_pdl = 1;
          launch_mrf_tc(ex, b);
          if (ch->done) ch->done((int)c, audio);
        }
      } else
        launch_mrf_tc(ex, a);
    } else
    for (int j = 0; j < nk; ++j) {
      const TcResBlock& rb = ts.rbs[j];
      const int M = (int)rb.c1.size();
      const float* y = xu;
      const uint16_t* y16 = xub;
      for (int m = 0; m < M; ++m) {
        const bool last = m == M - 1;
        const uint16_t* src16 = y16;
        if (!rb.c2.empty()) {   // ResBlock1: dilated conv -> lrelu -> plain conv
          TcConvArgs a1 = tc_args(rb.c1[m], y16, C, Rn, in.frow_utt, U, f16);
          a1.e[0].out_act = tbb; a1.e[0].ld_act = C; a1.e[0].act_slope = 0.1f;
          launch_conv_tc(ex, a1);
          src16 = tbb;
        }
        TcConvArgs a = tc_args(rb.c2.empty() ? rb.c1[m] : rb.c2[m], src16, C, Rn, in.frow_utt, U, f16);
        a.e[0].res = y; a.e[0].ld_res = C;
        if (!last) {
          a.e[0].out_f32 = yb; a.e[0].ld_f32 = C;
          a.e[0].out_act = ybb; a.e[0].ld_act = C; a.e[0].act_slope = 0.1f;
        } else {
          a.e[0].out_f32 = xs; a.e[0].ld_f32 = C;
          a.e[0].accum = j > 0;
          if (j == nk - 1) {
            a.e[0].div = (float)nk;
            if (!last_stage) { a.e[0].out_act = xsb; a.e[0].ld_act = C; a.e[0].act_slope = 0.1f; }
          }
        }
        launch_conv_tc(ex, a);
        y = yb;
        y16 = ybb;
      }
    }
    cur ^= 1;
    R = Rn;
  }
  if (!post_fused) launch_conv_post_tanh(ex, xs, v.stages.back().ch, v.post_w, v.post_k, 0.01f, in.frow_utt, U, audio, R, v.post_b);
  if (in.chunks && !chunked && in.chunks->done)   // this voice's last stage is not the fused form: everything is enqueued now
    for (size_t c = 0; c + 1 < in.chunks->frame_row.size(); ++c) in.chunks->done((int)c, audio);
  return audio;
}

}  // namespace piper
