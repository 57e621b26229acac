// Tensor-core (tcgen05 / TMEM / TMA) implicit-GEMM convolution for sm_100a: launch interface.
// Replaces conv1d_f32 / convtranspose1d_f32 (conv1d.metal:28-71, 97-142) and matmul_f32
// (matmul.metal:22-49) for the dense convolutions of the flow and the HiFi-GAN generator
//  when the context runs in a 16-bit operand precision mode.
//
// Operands are 16-bit (bf16 or fp16) "packed rows" [rows][channels] (channels contiguous, so
// the GEMM A operand is K-major in shared memory exactly as TMA delivers it); weights are
// repacked once to [phase][tap][cout][cin_pad] (K-major B operand).  Accumulation is fp32
// in TMEM; the epilogue applies bias / per-utterance bias / residual / cross-resblock
// accumulate / divide / gated tanh*sigmoid in fp32 and writes an fp32 tensor and/or the
// 16-bit (optionally leaky-relu'd) operand copy the next convolution reads.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.h"

namespace piper {

// 16-bit weights of one convolution: [phases*taps][cout][cin_pad]
struct TcW {
  void* w = nullptr;        // bf16 or fp16
  const float* bias = nullptr;   // fp32 [>=cout] (shared with the fp32 ConvW)
  int cin = 0, cin_pad = 0, cout = 0;
  int taps = 1, phases = 1, ct_pad = 0, dil = 1, pad_l = 0;
  int kb = 64;              // K elements per pipeline stage (64 -> 128B swizzle, 32 -> 64B swizzle)
  // polyphase ConvTranspose merged into ONE convolution (persistent kernel only): the weight rows are
  // [union tap][phase * phase_cols + c][cin_pad], zero where a phase does not use a row offset; column n
  // belongs to output row q * stride + n / phase_cols, channel n % phase_cols.  taps = union taps,
  // phases = 1, first_off = input row offset of union tap 0 (tap step +1).
  int phase_cols = 0, merged_phases = 0, first_off = 0;
  uint32_t phase_tapmask[16] = {0};   // per phase: bit u set when union tap u carries weights of that phase  // fp32 emulation ("bf16x6"): both operands are split into three bf16 terms x = x0 + x1 + x2 (24 mantissa bits) and
  // the six products with i + j <= 2 are accumulated in fp32, smallest first.  The weight rows then hold six
  // planes of cin_pad channels [B2 B1 B0 B1 B0 B0]; the activation operand is [rows][3 * cin_pad] = [A0 A1 A2] and
  // the kernel's producer reads planes [A0 A1 A2 A0 A1 A0] against them.
  // split = 2 ("f16x3"): two fp16 planes x = hi + lo' * 2^-11 (hi = fp16(x), lo' = fp16((x - hi) * 2^11): 22-23 mantissa
  // bits, the scaling keeps lo' out of the fp16 subnormals) and THREE products in two accumulators,
  // D0 = A_hi B_hi and D1 = A_hi B_lo' + A_lo' B_hi, combined as D0 + 2^-11 D1 in the epilogue (the dropped lo lo term
  // is below 2^-24 like the dropped bf16 terms).  Weight rows hold [B_hi | B_lo'], activations [A_hi | A_lo'];
  // half the tensor work of bf16x6.  Inputs must stay inside the fp16 range (they are clamped to +-65504).
  int split = 0;
};

struct TcEpi {
  // a column group: `nt * n_tiles` consecutive weight rows starting at row0 (per phase/tap)
  int row0 = 0;                  // first weight row (cout index) of the group
  const float* bias = nullptr;   // group-local: bias[j] belongs to weight row row0 + j (may be null)
  int ub_off = 0;                // offset of the group's column 0 inside a uttbias row
  int col0 = 0;                  // output column where the group's column 0 lands
  float* out_f32 = nullptr; int ld_f32 = 0;
  void* out_act = nullptr;  int ld_act = 0;     // 16-bit copy, leaky-relu(act_slope) applied
  // act_planar: out_act is "group-planar" [C/8][act_rows][8] (8-channel groups as planes, each row 16 bytes;
  // the layout conv_mrf.cu bulk-copies into its no-swizzle operand window); row r lives at plane row act_pad + r
  int act_planar = 0; int64_t act_rows = 0; int act_pad = 0;
  const float* res = nullptr; int ld_res = 0;   // fp32 residual added to the value (same col0)
  int accum = 0;                 // value += out_f32 (previous contents)
  float div = 0.f;               // if != 0: value /= div (after accum)
  float act_slope = 1.f;
  int relu = 0;                  // max(acc + bias, 0) before residual / accumulate
  // three-term bf16 split of the final fp32 value (operand of a `split` convolution): plane t of channel c at
  // out_split[row * ld_split + t * split_c + col0 + c]
  void* out_split = nullptr; int ld_split = 0, split_c = 0;
};

struct TcConvArgs {
  const void* in = nullptr;      // 16-bit activations [R_in][ldin], view of `cin` channels
  int ldin = 0, R_in = 0, cin = 0;
  // optional: the same activations group-planar [cin / 8][in_rows][8], row r at plane row in_pad + r, zero rows around
  // the signal (at least -in_off in front, 128 + the tap span behind).  Plain multi-tap convolutions then load each
  // K block's rows ONCE as a window and realise the taps as descriptor row offsets (persistent kernel); `in` is unused.
  const void* in_planar = nullptr; int64_t in_rows = 0; int in_pad = 0;
  const TcW* w = nullptr;
  int Q = 0;                     // output positions (per phase)
  int tap_step = 1, in_off = 0;  // input row of tap t for position q: q + in_off + t*tap_step
  int out_mul = 1;               // output row = q*out_mul + phase
  int n_groups = 1;              // 1, or 2 = two independent column groups per CTA (res|skip, tanh|sigmoid)
  int nt = 0;                    // columns per group handled by one CTA (UMMA N), multiple of 16, <= 256
  int n_tiles = 1;               // CTAs along the column axis (each covers nt columns of every group)
  int gate = 0;                  // 1: out_act = tanh(g0) * sigmoid(g1) (WN gate), only e[0] outputs used
  TcEpi e[2];
  const float* uttbias = nullptr; int ldub = 0;   // per-utterance bias [batch][ldub], indexed like bias
  const int* row_utt = nullptr;  int valid_div = 1;
  int f16 = 0;                   // 0 = bf16 operands, 1 = fp16 operands
  int window_mode = 1;           // 1: load each K block's activation window once and shift taps by descriptor offsets
};

// returns false (and sets err) when the shape cannot run on this kernel
bool tc_conv_supported(int cin, int cout);
void launch_conv_tc(Exec& ex, const TcConvArgs& a);

// one-time weight repack from the fp32 SIMT layout [phases][taps][cin][ldw] (device) into dst rows
// [ (ph*taps+t)*cout_total + row_off + co ][cin_pad]
void launch_pack_tc_weights(Exec& ex, const float* w_f32, int ldw, int phases_taps, int cin, int cout,
                            void* dst, int cout_total, int row_off, int cin_pad, int f16);
// split-operand helpers (fp32 emulation): weights [phases*taps][cout][6 * cin_pad], activations [R][3 * cpad]
void launch_pack_tc_weights_split(Exec& ex, const float* w_f32, int ldw, int phases_taps, int cin, int cout,
                                  void* dst, int cout_total, int row_off, int cin_pad);
// out[(row_off + ci) * ldo + co] = sum_m A[ci * lda + m] * B[m * ldb + co]   (ci < n, m < k, co < cout)  and, when
// bias_out != null, bias_out[co] (+)= sum_m ba[m] * B[m * ldb + co]: folds a 1x1 convolution (A, ba) into the 1x1
// convolution (B) that follows it.  One-time weight preparation.
void launch_fold_1x1(Exec& ex, const float* A, int lda, const float* ba, const float* B, int ldb, int n, int k, int cout,
                     float* out, int ldo, int row_off, float* bias_out);
void launch_split3(Exec& ex, const float* src, int ldsrc, int64_t R, int C, void* dst, int cpad);
// f16x3 helpers (TcW::split == 2): weights [phases*taps][cout][2 * cin_pad], activations [R][2 * cpad]
void launch_pack_tc_weights_split2(Exec& ex, const float* w_f32, int ldw, int phases_taps, int cin, int cout,
                                   void* dst, int cout_total, int row_off, int cin_pad);
void launch_split2(Exec& ex, const float* src, int ldsrc, int64_t R, int C, void* dst, int cpad);
// fp32 -> 16-bit copy of an activation tensor [R][C] (optional leaky-relu), used at stage boundaries
void launch_to_act16(Exec& ex, const float* src, void* dst, int64_t n, float slope, int f16);

}  // namespace piper
