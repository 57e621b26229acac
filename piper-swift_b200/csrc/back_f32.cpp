// Residual-coupling flow (reverse) + HiFi-GAN generator in fp32 on CUDA cores: the parity-mode
// back end.  Node-for-node it replaces, per coupling layer: Slice(step -1)
// [folded into the weights], Split, Conv 1x1, 4x{Conv k5, Add cond, Split, Tanh, Sigmoid, Mul,
// Conv 1x1, Add, Add}, Conv 1x1, Sub, Mul mask, Concat; and per generator stage: LeakyRelu,
// ConvTranspose, resblocks {LeakyRelu, Conv, [LeakyRelu, Conv], Add}, Add, Div, then
// LeakyRelu, Conv, Tanh (GraphExecutor.swift:1739-1860, 2017-2069).
#include <stdexcept>

#include "pipeline.h"

namespace piper {

namespace {

ConvArgs conv_args(const ConvW& w, const float* in, int ldin, float* out, int ldout, int R,
                   const int* row_utt, int valid_div) {
  ConvArgs a;
  a.in = in; a.ldin = ldin;
  a.w = w.w; a.ldw = w.ldw; a.bias = w.b;
  a.out = out; a.ldout = ldout;
  a.row_utt = row_utt; a.valid_div = valid_div;
  a.R_in = R; a.Q = R;
  a.cin = w.cin; a.cout = w.cout; a.taps = w.taps;
  a.tap_step = w.dil; a.in_off = -w.pad_l;
  return a;
}

}  // namespace

size_t back_f32_bytes(const Context& ctx, int Rf, int batch) {
  const Voice& v = ctx.voice;
  const int H = v.hp.hidden;
  size_t per_row_max = 0;
  int U = 1;
  for (auto& s : v.stages) {
    U *= s.stride;
    per_row_max = std::max(per_row_max, (size_t)U * s.ch);
  }
  per_row_max = std::max(per_row_max, (size_t)v.hp.dec_c0);
  size_t bytes = 0;
  bytes += (size_t)Rf * H * 4 * 4;          // h, acts, skip + slack
  bytes += (size_t)Rf * 2 * H * 4;          // xin
  bytes += (size_t)Rf * per_row_max * 4 * 5;   // decoder ping-pong buffers
  bytes += (size_t)Rf * v.hp.hop * 4;       // packed audio
  bytes += 16 * 256;
  (void)batch;
  return bytes;
}

FlowOut run_flow_f32(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep) {
  Voice& v = ctx.voice;
  Arena& A = ctx.back;
  const int H = v.hp.hidden, Rf = in.Rf;
  const int L = v.hp.flow_wn_layers;
  float* z = const_cast<float*>(in.zp);   // updated in place, half at a time
  float* h = A.alloc<float>((size_t)Rf * H);
  float* xin = A.alloc<float>((size_t)Rf * 2 * H);
  float* acts = A.alloc<float>((size_t)Rf * H);
  float* skip = A.alloc<float>((size_t)Rf * H);
  if (!h || !xin || !acts || !skip) throw std::runtime_error("capacity: back arena too small (flow)");

  // ---- flow ------------------------------------------------------------------------------
  for (size_t f = 0; f < v.flows.size(); ++f) {
    const CouplingW& c = v.flows[f];
    const float* cond = in.cond_flow ? in.cond_flow + f * (size_t)in.batch * 2 * H * L : nullptr;
    launch_conv_f32(ex, conv_args(c.pre, z + c.in_off, H, h, H, Rf, in.frow_utt, 1));
    for (int l = 0; l < L; ++l) {
      launch_conv_f32(ex, conv_args(c.in_layers[l], h, H, xin, 2 * H, Rf, in.frow_utt, 1));
      launch_wn_gate(ex, xin, cond ? cond + (size_t)l * 2 * H : nullptr, 2 * H * L, in.frow_utt, acts, Rf, H);
      if (l < L - 1) {
        ConvArgs r = conv_args(c.res[l], acts, H, h, H, Rf, in.frow_utt, 1);
        r.res = h; r.ldres = H;             // x = (x + res_acts) * mask
        launch_conv_f32(ex, r);
      }
      ConvArgs s = conv_args(c.skip[l], acts, H, skip, H, Rf, in.frow_utt, 1);
      s.accum = l > 0;                      // output = output + skip_acts
      launch_conv_f32(ex, s);
    }
    ConvArgs p = conv_args(c.post, skip, H, z + c.out_off, H, Rf, in.frow_utt, 1);
    p.res = z + c.out_off; p.ldres = H;     // x1 + (-m): post weights are stored negated
    launch_conv_f32(ex, p);
  }
  if (z_keep) {
    if (cudaMemcpyAsync(z_keep, z, (size_t)Rf * H * 4, cudaMemcpyDeviceToDevice, ex.st) != cudaSuccess)
      throw std::runtime_error("cuda: copy of flow output failed");
  }

  FlowOut out;
  out.z = z;
  return out;
}

float* run_back_f32(Context& ctx, Exec& ex, const BackInputs& in, float* z_keep) {
  const FlowOut f = run_flow_f32(ctx, ex, in, z_keep);
  DecoderIn d;
  d.z = f.z; d.frow_utt = in.frow_utt; d.rows = in.Rf; d.batch = in.batch; d.cond_dec = in.cond_dec;
  return run_decoder_f32(ctx, ex, d);
}

// ---- HiFi-GAN ------------------------------------------------------------------------------
float* run_decoder_f32(Context& ctx, Exec& ex, const DecoderIn& in) {
  Voice& v = ctx.voice;
  Arena& A = ctx.back;
  const int H = v.hp.hidden, Rf = in.rows;
  const float* z = in.z;
  size_t per_row_max = v.hp.dec_c0;
  {
    int U = 1;
    for (auto& s : v.stages) {
      U *= s.stride;
      per_row_max = std::max(per_row_max, (size_t)U * s.ch);
    }
  }
  float* buf[5];
  for (auto& b : buf) {
    b = A.alloc<float>((size_t)Rf * per_row_max);
    if (!b) throw std::runtime_error("capacity: back arena too small (decoder)");
  }
  float* audio = A.alloc<float>((size_t)Rf * v.hp.hop);
  if (!audio) throw std::runtime_error("capacity: back arena too small (audio)");

  float* cur = buf[0];
  {
    ConvArgs a = conv_args(v.dec_pre, z, H, cur, v.hp.dec_c0, Rf, in.frow_utt, 1);
    a.uttbias = in.cond_dec; a.ldub = v.hp.dec_c0;
    launch_conv_f32(ex, a);
  }
  int R = Rf, U = 1, cur_i = 0;
  const int nk = (int)v.stages[0].rbs.size();
  for (auto& s : v.stages) {
    float* xu = buf[(cur_i + 1) % 5];
    float* yb = buf[(cur_i + 2) % 5];
    float* tb = buf[(cur_i + 3) % 5];
    float* xs = buf[(cur_i + 4) % 5];
    const int Rn = R * s.stride;
    U *= s.stride;
    {
      ConvArgs a;
      a.in = cur; a.ldin = s.up.cin;
      a.w = s.up.w; a.ldw = s.up.ldw; a.bias = s.up.b;
      a.out = xu; a.ldout = s.ch;
      a.row_utt = in.frow_utt; a.valid_div = U;
      a.R_in = R; a.Q = R;
      a.cin = s.up.cin; a.cout = s.up.cout; a.taps = s.up.taps;
      a.tap_step = -1; a.out_mul = s.stride;
      a.phases = s.up.phases; a.ct_pad = s.up.ct_pad;
      a.in_act = 1; a.in_slope = 0.1f;
      launch_conv_f32(ex, a);
    }
    const int C = s.ch;
    for (int j = 0; j < nk; ++j) {
      const ResBlockW& rb = s.rbs[j];
      const int M = (int)rb.c1.size();
      const float* y = xu;
      for (int m = 0; m < M; ++m) {
        const bool last = m == M - 1;
        float* dst = last ? xs : yb;
        if (!rb.c2.empty()) {
          ConvArgs a1 = conv_args(rb.c1[m], y, C, tb, C, Rn, in.frow_utt, U);
          a1.in_act = 1; a1.in_slope = 0.1f;
          launch_conv_f32(ex, a1);
          ConvArgs a2 = conv_args(rb.c2[m], tb, C, dst, C, Rn, in.frow_utt, U);
          a2.in_act = 1; a2.in_slope = 0.1f;
          a2.res = y; a2.ldres = C;
          if (last) { a2.accum = j > 0; a2.div = (j == nk - 1) ? (float)nk : 0.f; }
          launch_conv_f32(ex, a2);
        } else {
          ConvArgs a1 = conv_args(rb.c1[m], y, C, dst, C, Rn, in.frow_utt, U);
          a1.in_act = 1; a1.in_slope = 0.1f;
          a1.res = y; a1.ldres = C;
          if (last) { a1.accum = j > 0; a1.div = (j == nk - 1) ? (float)nk : 0.f; }
          launch_conv_f32(ex, a1);
        }
        y = dst;
      }
    }
    cur = xs;
    cur_i = (cur_i + 4) % 5;
    R = Rn;
  }
  launch_conv_post_tanh(ex, cur, v.stages.back().ch, v.post_w, v.post_k, 0.01f, in.frow_utt, U, audio, R, v.post_b);
  return audio;
}

}  // namespace piper
