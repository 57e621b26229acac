// Kernel self-test behind piper_b200_selftest_conv: one convolution of a given shape on random
// data through the tcgen05 kernel (conv_tc.cu) and through the fp32 CUDA-core kernel
// (kernels_f32.cu) on the same operands.  The counterpart of the reference's
// PiperMetalMicrobench (Microbench/PiperMetalMicrobench.swift:24-73): it exercises single
// kernels outside the graph.  Operand values are multiples of 1/64 in [-1,1], exact in bf16 and
// fp16, so both kernels see identical inputs and differ only by fp32 summation order.
#include <cuda_fp16.h>

#include <cmath>
#include <cstring>
#include <random>
#include <stdexcept>
#include <vector>

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

struct DevBuf {
  void* p = nullptr;
  explicit DevBuf(size_t bytes) { CUDA_OK(cudaMalloc(&p, bytes ? bytes : 16)); CUDA_OK(cudaMemset(p, 0, bytes ? bytes : 16)); }
  ~DevBuf() { cudaFree(p); }
  template <typename T> T* as() { return static_cast<T*>(p); }
};

float from16(uint16_t b, int f16) {
  if (f16) {
    __half h;
    std::memcpy(&h, &b, 2);
    return __half2float(h);
  }
  uint32_t u = (uint32_t)b << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
}  // namespace

// mode 0: conv (+bias +residual), 1: WN gate (cout = 2*H), 2: res|skip two-group (cout = 2*H, skip accumulates),
// mode 3: transposed conv (stride, k multiple of stride)
// modes 4 / 5: fp32 emulation on the tensor cores (bf16x6 split operands, conv_tc.h) against the fp32 CUDA-core
// kernel, both measured against a float64 host reference on full-mantissa random data.
// out4 = {err(tc), err(simt), tc ms (incl. the activation split), simt ms}; err = rms (4) or max-abs (5).
static void selftest_split(Context& c, int mode, int cin, int cout, int k, int dil, int rows, float* out4) {
  Exec ex;
  ex.st = c.stream;
  std::mt19937 rng(77u + cin * 7 + cout * 13 + k * 31);
  std::normal_distribution<float> nd(0.f, 1.f);
  const int ldw = (cout + 63) / 64 * 64;
  std::vector<float> hw((size_t)k * cin * ldw, 0.f), hb(ldw, 0.f), hx((size_t)rows * cin);
  const float ws = 1.0f / std::sqrt((float)(cin * k));
  for (int t = 0; t < k; ++t)
    for (int ci = 0; ci < cin; ++ci)
      for (int co = 0; co < cout; ++co) hw[((size_t)t * cin + ci) * ldw + co] = nd(rng) * ws;
  for (int co = 0; co < cout; ++co) hb[co] = nd(rng) * 0.1f;
  for (auto& v : hx) v = nd(rng);
  std::vector<int> hutt(rows, 0);
  DevBuf dw(hw.size() * 4), db(hb.size() * 4), dx(hx.size() * 4), dxs(hx.size() * 6), dutt(rows * 4);
  DevBuf dref((size_t)rows * cout * 4), dtc((size_t)rows * cout * 4);
  CUDA_OK(cudaMemcpy(dw.p, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(db.p, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(dx.p, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(dutt.p, hutt.data(), rows * 4, cudaMemcpyHostToDevice));
  ConvW cw;
  cw.w = dw.as<float>(); cw.b = db.as<float>();
  cw.cin = cin; cw.cout = cout; cw.k = k; cw.dil = dil; cw.ldw = ldw;
  cw.taps = k; cw.phases = 1; cw.ct_pad = 0; cw.pad_l = dil * (k - 1) / 2;
  std::vector<void*> allocs;
  const int smode = [] { const char* e = getenv("PIPER_B200_SELFTEST_SPLIT"); return e && atoi(e) == 1 ? 1 : 2; }();
  TcW tw = make_tcw_split(ex, allocs, cw, smode);
  cudaEvent_t e0, e1, e2;
  CUDA_OK(cudaEventCreate(&e0)); CUDA_OK(cudaEventCreate(&e1)); CUDA_OK(cudaEventCreate(&e2));
  ConvArgs r;
  r.in = dx.as<float>(); r.ldin = cin;
  r.w = cw.w; r.ldw = ldw; r.bias = cw.b;
  r.out = dref.as<float>(); r.ldout = cout;
  r.row_utt = dutt.as<int>(); r.valid_div = 1;
  r.R_in = rows; r.Q = rows; r.cin = cin; r.cout = cout; r.taps = k;
  r.tap_step = dil; r.in_off = -cw.pad_l;
  TcConvArgs a;
  a.in = dxs.p; a.ldin = (smode == 2 ? 2 : 3) * cin; a.R_in = rows; a.cin = cin;
  a.w = &tw; a.Q = rows;
  a.row_utt = dutt.as<int>(); a.valid_div = 1;
  a.f16 = smode == 2 ? 1 : 0; a.tap_step = dil; a.in_off = -cw.pad_l;
  a.n_groups = 1;
  a.nt = 32;
  for (int nt : {256, 192, 128, 96, 64, 32}) if ((smode != 2 || nt <= 128) && cout % nt == 0) { a.nt = nt; break; }
  if (const char* e = getenv("PIPER_B200_SELFTEST_NT")) { const int nt = atoi(e); if (nt >= 32 && cout % nt == 0) a.nt = nt; }
  a.n_tiles = cout / a.nt;
  a.e[0].bias = cw.b; a.e[0].out_f32 = dtc.as<float>(); a.e[0].ld_f32 = cout;
  for (int rep = 0; rep < 2; ++rep) {   // second pass is the timed one
    CUDA_OK(cudaEventRecord(e0, c.stream));
    launch_conv_f32(ex, r);
    CUDA_OK(cudaEventRecord(e1, c.stream));
    if (smode == 2) launch_split2(ex, dx.as<float>(), cin, rows, cin, dxs.p, cin);
    else launch_split3(ex, dx.as<float>(), cin, rows, cin, dxs.p, cin);
    launch_conv_tc(ex, a);
    CUDA_OK(cudaEventRecord(e2, c.stream));
  }
  CUDA_OK(cudaStreamSynchronize(c.stream));
  CUDA_OK(cudaGetLastError());
  float ms_ref = 0, ms_tc = 0;
  cudaEventElapsedTime(&ms_ref, e0, e1);
  cudaEventElapsedTime(&ms_tc, e1, e2);
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
  std::vector<float> href((size_t)rows * cout), htc((size_t)rows * cout);
  CUDA_OK(cudaMemcpy(href.data(), dref.p, href.size() * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(htc.data(), dtc.p, htc.size() * 4, cudaMemcpyDeviceToHost));
  for (void* p : allocs) cudaFree(p);
  // float64 reference on a sample of rows (every row near the edges, then a stride)
  double se_tc = 0, se_ref = 0, mx_tc = 0, mx_ref = 0;
  size_t cnt = 0;
  const int step = rows > 512 ? rows / 256 : 1;
  for (int q = 0; q < rows; q += (q < 8 || q >= rows - 8) ? 1 : step) {
    for (int co = 0; co < cout; ++co) {
      double acc = hb[co];
      for (int t = 0; t < k; ++t) {
        const int ir = q - cw.pad_l + t * dil;
        if (ir < 0 || ir >= rows) continue;
        const float* xr = &hx[(size_t)ir * cin];
        const float* wr = &hw[(size_t)t * cin * ldw + co];
        for (int ci = 0; ci < cin; ++ci) acc += (double)xr[ci] * (double)wr[(size_t)ci * ldw];
      }
      const double d1 = std::fabs((double)htc[(size_t)q * cout + co] - acc);
      const double d2 = std::fabs((double)href[(size_t)q * cout + co] - acc);
      se_tc += d1 * d1; se_ref += d2 * d2;
      mx_tc = std::max(mx_tc, std::isnan(d1) ? 1e30 : d1); mx_ref = std::max(mx_ref, d2);
      ++cnt;
    }
  }
  out4[0] = mode == 4 ? (float)std::sqrt(se_tc / cnt) : (float)mx_tc;
  out4[1] = mode == 4 ? (float)std::sqrt(se_ref / cnt) : (float)mx_ref;
  out4[2] = ms_tc; out4[3] = ms_ref;
}

int Context::selftest_conv(int mode, int cin, int cout, int k, int dil, int stride, int rows, int prec, float* out4) {
  try {
    CUDA_OK(cudaSetDevice(device));
    if (mode == 4 || mode == 5) {
      if (cin % 32 || cout % 32 || k < 1 || rows < 1) throw std::runtime_error("invalid argument: split selftest shape");
      selftest_split(*this, mode, cin, cout, k, dil, rows, out4);
      return PIPER_B200_OK;
    }
    if (prec != PIPER_B200_PRECISION_BF16 && prec != PIPER_B200_PRECISION_F16)
      throw std::runtime_error("invalid argument: selftest precision must be bf16 or f16");
    const int f16 = prec == PIPER_B200_PRECISION_F16;
    Exec ex;
    ex.st = stream;
    std::mt19937 rng(1234u + cin * 7 + cout * 13 + k * 31 + mode);
    auto rnd = [&]() { return (float)((int)(rng() % 129) - 64) / 64.0f; };
    const bool ct = mode == 3;
    const int phases = ct ? stride : 1;
    const int taps = ct ? k / stride : k;
    if (ct && (stride < 1 || k % stride)) throw std::runtime_error("invalid argument: k % stride");
    const int ldw = (cout + 63) / 64 * 64;
    // SIMT layout [phases][taps][cin][ldw]
    std::vector<float> hw((size_t)phases * taps * cin * ldw, 0.f), hb(ldw, 0.f);
    for (int pt = 0; pt < phases * taps; ++pt)
      for (int ci = 0; ci < cin; ++ci)
        for (int co = 0; co < cout; ++co) hw[((size_t)pt * cin + ci) * ldw + co] = rnd() * 0.25f;
    for (int co = 0; co < cout; ++co) hb[co] = rnd();
    const int ldin = cin + 32;   // a strided view, like the coupling halves
    std::vector<float> hx((size_t)rows * ldin);
    for (auto& v : hx) v = rnd();
    const int Rout = rows * phases;
    const int H = cout / 2;
    std::vector<float> hres((size_t)Rout * cout);
    for (auto& v : hres) v = rnd();
    std::vector<int> hutt(rows, 0);
    for (int r = 0; r < rows; ++r) hutt[r] = (r % 97 == 50) ? -1 : 0;   // a few masked rows

    DevBuf dw(hw.size() * 4), db(hb.size() * 4), dx(hx.size() * 4), dx16(hx.size() * 2), dres(hres.size() * 4);
    DevBuf dutt(rows * 4), dref((size_t)Rout * cout * 4), dtc((size_t)Rout * cout * 4), dact((size_t)Rout * cout * 2);
    DevBuf dref2((size_t)Rout * cout * 4), dtc2((size_t)Rout * cout * 4);
    CUDA_OK(cudaMemcpy(dtc2.p, hres.data(), hres.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(dw.p, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(db.p, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(dx.p, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(dres.p, hres.data(), hres.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(dutt.p, hutt.data(), rows * 4, cudaMemcpyHostToDevice));
    launch_to_act16(ex, dx.as<float>(), dx16.p, (int64_t)hx.size(), 1.f, f16);

    ConvW cw;
    cw.w = dw.as<float>(); cw.b = db.as<float>();
    cw.cin = cin; cw.cout = cout; cw.k = k; cw.dil = dil; cw.ldw = ldw;
    cw.taps = taps; cw.phases = phases; cw.ct_pad = ct ? (k - stride) / 2 : 0;
    cw.pad_l = dil * (k - 1) / 2;
    std::vector<void*> allocs;
    TcW tw = make_tcw(ex, allocs, cw, nullptr, f16);

    cudaEvent_t e0, e1, e2;
    CUDA_OK(cudaEventCreate(&e0)); CUDA_OK(cudaEventCreate(&e1)); CUDA_OK(cudaEventCreate(&e2));

    // ---- reference: fp32 CUDA-core kernel -----------------------------------------------------
    ConvArgs r;
    r.in = dx.as<float>(); r.ldin = ldin;
    r.w = cw.w; r.ldw = ldw; r.bias = cw.b;
    r.out = dref.as<float>(); r.ldout = cout;
    r.row_utt = dutt.as<int>(); r.valid_div = phases;
    r.R_in = rows; r.Q = rows; r.cin = cin; r.cout = cout; r.taps = taps;
    if (ct) { r.tap_step = -1; r.out_mul = stride; r.phases = phases; r.ct_pad = cw.ct_pad; }
    else { r.tap_step = dil; r.in_off = -cw.pad_l; }
    if (mode == 0 || mode == 3) { r.res = dres.as<float>(); r.ldres = cout; }
    CUDA_OK(cudaMemcpy(dref2.p, hres.data(), hres.size() * 4, cudaMemcpyHostToDevice));   // "skip" accumulator seed
    CUDA_OK(cudaEventRecord(e0, stream));
    if (mode == 2) {
      // res half: out[:, :H] = conv + res ; skip half: accumulates into dref2[:, H:]
      ConvArgs a = r; a.cout = H; a.res = dres.as<float>(); a.ldres = cout;
      launch_conv_f32(ex, a);
      ConvArgs b = r; b.cout = H; b.w = cw.w + H; b.bias = cw.b + H; b.out = dref2.as<float>() + H; b.accum = 1;
      launch_conv_f32(ex, b);
    } else {
      launch_conv_f32(ex, r);
      if (mode == 1) launch_wn_gate(ex, dref.as<float>(), nullptr, 0, dutt.as<int>(), dref2.as<float>(), rows, H);
    }
    CUDA_OK(cudaEventRecord(e1, stream));

    // ---- tensor-core kernel ---------------------------------------------------------------------
    TcConvArgs a;
    a.in = dx16.p; a.ldin = ldin; a.R_in = rows; a.cin = cin;
    a.w = &tw; a.Q = rows;
    a.row_utt = dutt.as<int>(); a.valid_div = phases;
    a.f16 = f16;
    if (ct) { a.tap_step = -1; a.out_mul = stride; }
    else { a.tap_step = dil; a.in_off = -cw.pad_l; }
    auto pick = [](int c) { for (int nt : {128, 96, 64, 48, 32, 16}) if (c % nt == 0) return nt; return 16; };
    if (mode == 1) {
      a.gate = 1; a.n_groups = 2; a.nt = pick(H); a.n_tiles = H / a.nt;
      a.e[0].row0 = 0; a.e[0].bias = cw.b; a.e[1].row0 = H; a.e[1].bias = cw.b + H;
      a.e[0].out_act = dact.p; a.e[0].ld_act = H;
    } else if (mode == 2) {
      a.n_groups = 2; a.nt = pick(H); a.n_tiles = H / a.nt;
      a.e[0].row0 = 0; a.e[0].bias = cw.b;
      a.e[0].out_f32 = dtc.as<float>(); a.e[0].ld_f32 = cout; a.e[0].res = dres.as<float>(); a.e[0].ld_res = cout;
      a.e[1].row0 = H; a.e[1].bias = cw.b + H; a.e[1].col0 = H;
      a.e[1].out_f32 = dtc2.as<float>(); a.e[1].ld_f32 = cout; a.e[1].accum = 1;
    } else {
      a.n_groups = 1; a.nt = pick(cout); a.n_tiles = cout / a.nt;
      a.e[0].bias = cw.b;
      a.e[0].out_f32 = dtc.as<float>(); a.e[0].ld_f32 = cout;
      a.e[0].res = dres.as<float>(); a.e[0].ld_res = cout;
      a.e[0].out_act = dact.p; a.e[0].ld_act = cout;
    }
    // timing probes: PIPER_B200_SELFTEST_NT overrides the column tile, PIPER_B200_SELFTEST_REPEAT launches the (idempotent)
    // kernel n times and reports the mean
    if (const char* e = getenv("PIPER_B200_SELFTEST_NT")) {
      const int nt = atoi(e), width = mode == 0 || mode == 3 ? cout : H;
      if (nt >= 16 && width % nt == 0) { a.nt = nt; a.n_tiles = width / nt; }
    }
    int repeat = 1;
    if (const char* e = getenv("PIPER_B200_SELFTEST_REPEAT")) repeat = std::max(1, atoi(e));
    launch_conv_tc(ex, a);
    if (repeat > 1) {
      CUDA_OK(cudaEventRecord(e1, stream));
      for (int i = 0; i < repeat; ++i) launch_conv_tc(ex, a);
    }
    CUDA_OK(cudaEventRecord(e2, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    CUDA_OK(cudaGetLastError());
    float ms_ref = 0, ms_tc = 0;
    cudaEventElapsedTime(&ms_ref, e0, e1);
    cudaEventElapsedTime(&ms_tc, e1, e2);
    ms_tc /= (float)repeat;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);

    // ---- compare ----------------------------------------------------------------------------------
    double max_diff = 0, max_ref = 0;
    auto cmp = [&](const std::vector<float>& x, const std::vector<float>& y, int c0, int c1, int ld, int R) {
      for (int rr = 0; rr < R; ++rr)
        for (int c = c0; c < c1; ++c) {
          const double d = std::fabs((double)x[(size_t)rr * ld + c] - y[(size_t)rr * ld + c]);
          if (!(d <= max_diff)) max_diff = std::isnan(d) ? 1e30 : d;
          max_ref = std::max(max_ref, (double)std::fabs(y[(size_t)rr * ld + c]));
        }
    };
    std::vector<float> href((size_t)Rout * cout), htc((size_t)Rout * cout);
    if (mode == 1) {
      std::vector<uint16_t> h16((size_t)rows * H);
      std::vector<float> g((size_t)rows * H), gr((size_t)rows * H);
      CUDA_OK(cudaMemcpy(h16.data(), dact.p, h16.size() * 2, cudaMemcpyDeviceToHost));
      CUDA_OK(cudaMemcpy(gr.data(), dref2.p, gr.size() * 4, cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < g.size(); ++i) g[i] = from16(h16[i], f16);
      cmp(g, gr, 0, H, H, rows);
    } else if (mode == 2) {
      CUDA_OK(cudaMemcpy(href.data(), dref.p, href.size() * 4, cudaMemcpyDeviceToHost));
      CUDA_OK(cudaMemcpy(htc.data(), dtc.p, htc.size() * 4, cudaMemcpyDeviceToHost));
      cmp(htc, href, 0, H, cout, Rout);
      CUDA_OK(cudaMemcpy(href.data(), dref2.p, href.size() * 4, cudaMemcpyDeviceToHost));
      CUDA_OK(cudaMemcpy(htc.data(), dtc2.p, htc.size() * 4, cudaMemcpyDeviceToHost));
      cmp(htc, href, H, cout, cout, Rout);
    } else {
      CUDA_OK(cudaMemcpy(href.data(), dref.p, href.size() * 4, cudaMemcpyDeviceToHost));
      CUDA_OK(cudaMemcpy(htc.data(), dtc.p, htc.size() * 4, cudaMemcpyDeviceToHost));
      cmp(htc, href, 0, cout, cout, Rout);
      // the 16-bit operand copy must be the rounded fp32 output
      std::vector<uint16_t> h16((size_t)Rout * cout);
      CUDA_OK(cudaMemcpy(h16.data(), dact.p, h16.size() * 2, cudaMemcpyDeviceToHost));
      const double rel = f16 ? 1.0 / 1024 : 1.0 / 128;
      for (size_t i = 0; i < h16.size(); ++i) {
        const double d = std::fabs((double)from16(h16[i], f16) - htc[i]);
        if (d > rel * std::fabs((double)htc[i]) + 1e-6) max_diff = std::max(max_diff, d);
      }
    }
    for (void* p : allocs) cudaFree(p);
    out4[0] = (float)max_diff; out4[1] = (float)max_ref; out4[2] = ms_tc; out4[3] = ms_ref;
    return PIPER_B200_OK;
  } catch (const std::exception& e) {
    last_error = e.what();
    cudaGetLastError();
    if (last_error.rfind("invalid", 0) == 0) return PIPER_B200_ERR_INVALID_ARGUMENT;
    return last_error.rfind("unsupported", 0) == 0 ? PIPER_B200_ERR_UNSUPPORTED : PIPER_B200_ERR_CUDA;
  }
}

// Op-level known-answer hook behind piper_b200_conv1d: ONE Conv / ConvTranspose node on caller-supplied operands in the
// ONNX layouts (Conv W [cout][cin][k], ConvTranspose W [cin][cout][k]; "same" padding dil*(k-1)/2 resp. (k-stride)/2,
// the only forms a Piper-VITS graph holds), so that a test can set the product kernels against
// torch.nn.functional.conv1d / conv_transpose1d.  fp32 -> conv_f32_*_kernel, f16 / bf16 -> conv_tc_persist_kernel
// (tcgen05) on the 16-bit rounding of x and W with fp32 accumulation.  Reference kernels: conv1d_f32
// (conv1d.metal:28-71), convtranspose1d_f32 (conv1d.metal:97-142).
int Context::conv1d_op(int transposed, const float* x, int64_t rows64, const float* w, const float* bias, int cin, int cout, int k,
                       int dil, int stride, int prec, float* out) {
  try {
    CUDA_OK(cudaSetDevice(device));
    if (rows64 <= 0 || rows64 > (1 << 24)) throw std::runtime_error("invalid argument: rows");
    const int rows = (int)rows64;
    const bool ct = transposed != 0;
    if (ct && (stride < 1 || k % stride || dil != 1)) throw std::runtime_error("invalid argument: ConvTranspose needs k % stride == 0, dilation 1");
    if (!ct && (stride != 1 || (k & 1) == 0)) throw std::runtime_error("invalid argument: Conv needs stride 1 and an odd kernel");
    const bool use_tc = prec != PIPER_B200_PRECISION_FP32;
    if (use_tc && (cin % 32 || cout % 16)) throw std::runtime_error("invalid argument: tensor-core conv needs cin % 32 == 0, cout % 16 == 0");
    const int f16 = prec == PIPER_B200_PRECISION_F16;
    Exec ex;
    ex.st = stream;
    const int phases = ct ? stride : 1, taps = ct ? k / stride : k;
    const int ldw = (cout + 63) / 64 * 64;
    const int pad = ct ? (k - stride) / 2 : 0;
    std::vector<float> pk((size_t)phases * taps * cin * ldw, 0.f), hb(ldw, 0.f);
    if (ct) {   // same repack as Binder::conv_transpose (voice.cpp): phase ph, tap j <- kernel position (ph + pad) % u + j * u
      for (int ph = 0; ph < phases; ++ph) {
        const int kk0 = (ph + pad) % stride;
        for (int j = 0; j < taps; ++j)
          for (int ci = 0; ci < cin; ++ci)
            for (int co = 0; co < cout; ++co)
              pk[(((size_t)ph * taps + j) * cin + ci) * ldw + co] = w[((size_t)ci * cout + co) * k + kk0 + j * stride];
      }
    } else {
      for (int co = 0; co < cout; ++co)
        for (int ci = 0; ci < cin; ++ci)
          for (int t = 0; t < k; ++t) pk[((size_t)t * cin + ci) * ldw + co] = w[((size_t)co * cin + ci) * k + t];
    }
    for (int co = 0; co < cout && bias; ++co) hb[co] = bias[co];
    const int Rout = rows * phases;
    DevBuf dw(pk.size() * 4), db(hb.size() * 4), dx((size_t)rows * cin * 4), dx16((size_t)rows * cin * 2), dutt((size_t)rows * 4);
    DevBuf dout((size_t)Rout * cout * 4);
    CUDA_OK(cudaMemcpy(dw.p, pk.data(), pk.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(db.p, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(dx.p, x, (size_t)rows * cin * 4, cudaMemcpyHostToDevice));
    ConvW cw;
    cw.w = dw.as<float>(); cw.b = db.as<float>();
    cw.cin = cin; cw.cout = cout; cw.k = k; cw.dil = dil; cw.ldw = ldw;
    cw.taps = taps; cw.phases = phases; cw.ct_pad = pad; cw.pad_l = dil * (k - 1) / 2;
    std::vector<void*> allocs;
    if (!use_tc) {
      ConvArgs r;
      r.in = dx.as<float>(); r.ldin = cin;
      r.w = cw.w; r.ldw = ldw; r.bias = cw.b;
      r.out = dout.as<float>(); r.ldout = cout;
      r.row_utt = dutt.as<int>(); r.valid_div = phases;   // all rows belong to utterance 0 (DevBuf zero-fills)
      r.R_in = rows; r.Q = rows; r.cin = cin; r.cout = cout; r.taps = taps;
      if (ct) { r.tap_step = -1; r.out_mul = stride; r.phases = phases; r.ct_pad = pad; }
      else { r.tap_step = dil; r.in_off = -cw.pad_l; }
      launch_conv_f32(ex, r);
    } else {
      launch_to_act16(ex, dx.as<float>(), dx16.p, (int64_t)rows * cin, 1.f, f16);
      TcW tw = make_tcw(ex, allocs, cw, nullptr, f16);
      TcConvArgs a;
      a.in = dx16.p; a.ldin = cin; a.R_in = rows; a.cin = cin;
      a.w = &tw; a.Q = rows;
      a.row_utt = dutt.as<int>(); a.valid_div = phases;
      a.f16 = f16;
      if (ct) { a.tap_step = -1; a.out_mul = stride; }
      else { a.tap_step = dil; a.in_off = -cw.pad_l; }
      a.n_groups = 1;
      a.nt = 16;
      for (int nt : {128, 96, 64, 48, 32, 16}) if (cout % nt == 0) { a.nt = nt; break; }
      a.n_tiles = cout / a.nt;
      a.e[0].bias = cw.b;
      a.e[0].out_f32 = dout.as<float>(); a.e[0].ld_f32 = cout;
      launch_conv_tc(ex, a);
    }
    CUDA_OK(cudaStreamSynchronize(stream));
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaMemcpy(out, dout.p, (size_t)Rout * cout * 4, cudaMemcpyDeviceToHost));
    for (void* p : allocs) cudaFree(p);
    return PIPER_B200_OK;
  } catch (const std::exception& e) {
    last_error = e.what();
    cudaGetLastError();
    if (last_error.rfind("invalid", 0) == 0) return PIPER_B200_ERR_INVALID_ARGUMENT;
    return last_error.rfind("unsupported", 0) == 0 ? PIPER_B200_ERR_UNSUPPORTED : PIPER_B200_ERR_CUDA;
  }
}

}  // namespace piper
