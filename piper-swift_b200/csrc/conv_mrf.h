// Fused multi-receptive-field stage of the HiFi-GAN generator on tcgen05 (sm_100a).
// One persistent kernel computes, for a whole generator stage with ResBlock2 blocks,
//     out = (1/nk) * sum_r  y2_r,   y1_r = x + conv_{k_r,d0_r}(lrelu(x)) + b1_r,
//                                   y2_r = y1_r + conv_{k_r,d1_r}(lrelu(y1_r)) + b2_r
// i.e. the graph's LeakyRelu/Conv/Add chains, the Add across kernels and the Div
// (GraphExecutor.swift:1739-1810, 2047-2069; conv1d.metal:28-71) with every intermediate on
// chip: operands in planar shared memory, y1 and the cross-resblock sum in fp32 in tensor memory.
// The stage input is read ONCE, as the 16-bit lrelu(x) operand; the residual term x is recovered from
// it (x = v >= 0 ? v : 10 v), so x enters with 16-bit rounding like every conv operand does.
// HBM traffic per element: read lrelu(x) (2 bytes), write the stage output once - instead of ~12
// bytes per element per convolution.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.h"

namespace piper {

constexpr int MRF_MAX_RB = 4;

struct MrfArgs {
  int R = 0;                     // rows (time positions) of this stage
  int C = 0;                     // channels (32 or 64)
  int nk = 0;                    // resblocks
  int k[MRF_MAX_RB] = {0, 0, 0, 0};
  int d0[MRF_MAX_RB] = {0, 0, 0, 0}, d1[MRF_MAX_RB] = {0, 0, 0, 0};
  int tap_base[MRF_MAX_RB][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};   // first tap (row block of C rows) of conv (r, j) in `w`
  const float* b1[MRF_MAX_RB] = {nullptr, nullptr, nullptr, nullptr};
  const float* b2[MRF_MAX_RB] = {nullptr, nullptr, nullptr, nullptr};
  const void* w = nullptr;       // 16-bit weights [total_taps][C cout][C cin]
  int total_taps = 0;
  // the stage input as 16-bit lrelu(x, 0.1), group-planar [C/8][act_rows][8]; signal row r at plane row act_pad + r; the rows before
  // act_pad and after act_pad + R must be zero (act_pad >= 128, act_rows >= R + act_pad + 1152)
  const void* x_act = nullptr;
  int64_t act_rows = 0; int act_pad = 0;
  const int* row_utt = nullptr;  int valid_div = 1;
  void* out_act = nullptr;       // optional 16-bit lrelu(out, act_slope) [R][C]
  float act_slope = 1.f;
  float* out_f32 = nullptr;      // optional fp32 out [R][C]
  // optional fused conv_post (last stage, C = 32): audio[r] = tanh(sum_t sum_c post_w[t][c] * lrelu(out[r + t - pad][c],
  // post_slope)) - the final LeakyRelu / Conv(C -> 1, k <= 7, no bias) / Tanh of the generator
  // (GraphExecutor.swift:2047-2069, 1739-1810, 2017) as the stage kernel's output epilogue; the stage output itself
  // is then never written.  post_w = fp32 [post_k][C].
  const float* post_w = nullptr; int post_k = 0; float post_slope = 0.01f; float post_bias = 0.f;
  float* post_out = nullptr;     // fp32 [R]
  // >= 0: `w` also holds post_k taps from this tap index on, each a C x C tap whose cout row 0 is post_w[t][:] in 16 bits
  // and whose other rows are zero: conv_post then runs on the tensor cores as one more convolution of the stage
  // (lrelu(out) as a 16-bit operand, fp32 accumulation) when all taps are resident in shared memory
  int post_tap_base = -1;
  // ResBlock1 unit (nk = 1, k[0] taps for both convolutions, d0 = dilation of conv1, d1 = 1):
  //   out = x + conv2(lrelu(conv1(lrelu x) + b1)) + b2          (GraphExecutor.swift:2047-2069 chains, HiFi-GAN ResBlock1)
  // x enters as the 16-bit lrelu(x) operand (like every fused stage); the next unit reads `out_planar`, the last unit of a
  // resblock adds into / finishes the cross-resblock mean in out_f32 (accum, out_div)
  int unit_mode = 0;
  int accum = 0;                 // out += previous contents of out_f32
  float out_div = 0.f;           // != 0: out /= out_div (after accum), then out_f32 / out_act / out_planar are written
  void* out_planar = nullptr;    // optional 16-bit lrelu(out, act_slope) in the group-planar layout of x_act
  int64_t out_rows = 0; int out_pad = 0;
  int f16 = 0;
  int no_pdl = 0;                // 1: plain stream-ordered launch instead of a programmatic dependent launch (kernels.h)
};

bool mrf_tc_supported(int C, int nk, const int* k, const int* d0, const int* d1);
bool mrf_post_supported(int C, int post_k);
void launch_mrf_tc(Exec& ex, const MrfArgs& a);

}  // namespace piper
