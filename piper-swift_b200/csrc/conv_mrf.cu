// Fused MRF stage (see conv_mrf.h).  Persistent kernel, one CTA per SM, 160 + 128*MRF_EG threads:
//   warp 0      : TMA producer - the 16-bit lrelu(x) window ((m+1)*128 rows, halo included, group-planar
//                 without swizzle) and the weight taps (ring of C x C tiles; resident when they all fit)
//   warps 1..4 : tcgen05.mma issuers (warp 1 also owns the TMEM allocation).  Warps 1, 2 issue conv1 of every
//                 resblock (tile i into D_i, fresh; even / odd tiles), warps 3, 4 issue conv2 (accumulating into
//                 S_i on top of sum_r y1_r; even / odd tiles): one warp needs ~55+ cycles per small-N MMA plus a
//                 fixed cost per tile, so a single stream left the tensor pipe waiting for the issuer (timeline,
//                 tools/mrf_trace.py); tap shifts are 16-byte-granular descriptor offsets into the planar windows
//   warps 5..   : MRF_EG epilogue warpgroups over (tile, 32-column half) units.  After conv1:
//                 y1 = D + x + b1 (fp32), S += y1 (TMEM read-modify-write), lrelu(y1)*mask -> 16-bit ->
//                 planar smem operand of conv2.  After the last resblock: out = (S + sum b2) / nk -> global.
// A window of W = 128*m rows yields V = W - 2*HALO valid rows (HALO = max_r pad0_r + pad1_r):
// rows closer than HALO to a window edge see truncated receptive fields and are recomputed by
// the neighbouring window.
#include "conv_mrf.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "tc_common.cuh"

namespace piper {

using namespace tc;

namespace {

constexpr int MRF_G = 64;          // upper bound of the guard rows on both sides of the operand windows (p.G >= largest tap shift)
constexpr int MRF_MAX_M = 8;
constexpr int MRF_MAX_RING = 40;   // >= the 30 + 7 resident taps of a 32-channel stage with its conv_post taps
constexpr uint32_t MRF_TRACE_EVENTS = 4096;

struct MrfParams {
  int R, C, nk, m, V, halo, n_windows, G;   // G: guard rows actually used (multiple of 8, <= MRF_G)
  int GX;                                   // guard rows of the lrelu(x) window (conv1 shifts only; <= G)
  int tap_outer;                            // weight ring shorter than a convolution: one issuing warp, taps outermost
  // entry nk of k / d1 / tap_base[.][1] describes the conv_post pseudo pair (post_tc)
  int k[MRF_MAX_RB + 1], d0[MRF_MAX_RB + 1], d1[MRF_MAX_RB + 1];
  int tap_base[MRF_MAX_RB + 1][2];
  const float* b1[MRF_MAX_RB];
  const float* b2[MRF_MAX_RB];
  const uint16_t* x_act; int64_t act_rows; int act_pad;   // group-planar 16-bit lrelu(x): [C/8][act_rows][8]
  const int* row_utt; int valid_div;
  int valid_shift;   // log2(valid_div) when it is a power of two (every stock voice), else -1: a shift instead of a ~20-instruction division per row lookup
  void* out_act; float act_slope;
  float* out_f32;
  const float* post_w; int post_k; float post_slope; float* post_out;   // fused conv_post + tanh (C = 32 only)
  float post_bias;
  // conv_post on the tensor cores (resident weights only): after the last resblock the epilogue writes lrelu(out) as one
  // more 16-bit operand window, the conv2 stream runs the post_k taps (a C x C tap whose row 0 holds the post weights)
  // into S_i, and the output step reads column 0
  int post_tc;
  int f16, ring, resident;
  int mma_split;              // 1: two issuing warps per convolution stream (even / odd tiles)
  // ResBlock1 unit (nk = 1): out = x + conv2(lrelu(conv1(lrelu x) + b1)) + b2 - the residual does not enter the
  // conv2 operand: Y1 = lrelu(D + b1), S starts from x
  int unit_mode;
  int accum;                  // stage output += previous contents of out_f32 (cross-resblock sum kept in HBM)
  float out_div;              // != 0: stage output /= out_div (after accum)
  uint16_t* out_planar; int64_t out_rows; int out_pad;   // optional group-planar 16-bit lrelu(out, act_slope), like x_act
  uint32_t tmem_cols;
  uint32_t* trace;            // debug timeline of CTA 0 (PIPER_B200_MRF_TRACE), else null
};

// UNIT: ResBlock1 unit semantics (MrfParams::unit_mode) and the accumulate / planar stage outputs - a compile-time
// switch so that the ResBlock2 instantiations keep their register allocation
template <int C, int MRF_EG, bool UNIT = false>   // channels, epilogue warpgroups
__global__ void __launch_bounds__(160 + 128 * MRF_EG, 1) mrf_tc_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                                const MrfParams p) {
  PDL_TRIGGER();   // the next kernel may be staged; PDL_WAIT() below, after the set-up that touches constants only
  extern __shared__ uint8_t smem_raw[];
  // 1 KB alignment by pointer arithmetic on the shared array: an integer round-trip hides the address space from the
  // compiler, which then emits generic LD / ST for every shared-memory access of the epilogues
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int m = p.m;
  // operand windows are group-planar without swizzle: [C/8 groups][(m+1)*128 rows][16 bytes]; a tap shift is
  // a 16-byte-granular descriptor offset (row-shifted descriptors into SWIZZLED tiles are ~100x slower
  // unless the shift is a multiple of the 8-row swizzle atom - measured, profiles/r01_tc_align_probe.log)
  constexpr int KS = C / 16;                          // tcgen05 K steps per tap
  constexpr int NP = C / 8;                           // planes
  constexpr int tap_bytes = C * C * 2;
  constexpr int KB = C > 64 ? 64 : C;                 // channels per swizzled weight sub-tile (128-byte rows at most)
  constexpr int KH = C / KB;                          // K sub-tiles per tap
  constexpr int sub_bytes = C * KB * 2;
  const int plane_x = (m * 128 + 2 * p.GX) * 16;      // lrelu(x) window: W + 2*GX rows
  const int plane_bytes = (m * 128 + 2 * p.G) * 16;   // conv2 operand window: W + 2*G rows
  const int win_x = NP * plane_x;
  const int win_bytes = NP * plane_bytes;
  uint8_t* X0 = smem;
  uint8_t* Y1 = smem + win_x;
  uint8_t* WR = Y1 + win_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(WR + (size_t)p.ring * tap_bytes);
  uint64_t* x_full = bars;
  uint64_t* x_empty = bars + 1;
  uint64_t* w_full = bars + 2;
  uint64_t* w_empty = w_full + MRF_MAX_RING;
  uint64_t* d_full = w_empty + MRF_MAX_RING;
  uint64_t* d_empty = d_full + MRF_MAX_M;
  uint64_t* y_ready = d_empty + MRF_MAX_M;
  uint64_t* s_full = y_ready + MRF_MAX_M;
  uint64_t* e_ready = s_full + MRF_MAX_M;    // fused conv_post: tile i's edge partial sums are published
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(e_ready + MRF_MAX_M);
  float* sbias = reinterpret_cast<float*>(tmem_slot + 4);   // [nk][C] b1 then [C] sum_r b2
  // fused conv_post: weights centred in a 7-tap frame [7][32], then the partial sums of the three first / last rows
  // of every 32-row block [2 window parities][m*4][2][3][7] (what the neighbouring warps need of each other)
  float* spost = sbias + (MRF_MAX_RB + 1) * C;
  float* sedge = spost + 7 * 32;

  // warp index and TMEM base go through a shuffle so that the compiler can prove them warp-uniform: the
  // role branches become uniform and the MMA descriptors live in uniform registers (without this every
  // tcgen05.mma pays for several R2UR moves - measured 100 vs 55 cycles per issue, tools/micro/umma_bench.cu)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  // debug timeline (build with -DPIPER_MRF_TRACE, run with PIPER_B200_MRF_TRACE=<path>, read with
  // tools/mrf_trace.py): lane 0 of every warp of CTA 0 appends (tag<<24 | n<<8 | i, clock) to its own region
#ifdef PIPER_MRF_TRACE
  uint32_t trace_n = 0;
  auto TRACE = [&](uint32_t tag, uint32_t n, uint32_t i) {
    if (p.trace && blockIdx.x == 0 && lane == 0 && trace_n < MRF_TRACE_EVENTS) {
      uint32_t* t = p.trace + ((size_t)warp * MRF_TRACE_EVENTS + trace_n) * 2;
      t[0] = (tag << 24) | ((n & 0xFFFFu) << 8) | (i & 0xFFu);
      t[1] = (uint32_t)clock64();
      ++trace_n;
    }
  };
#else
  auto TRACE = [](uint32_t, uint32_t, uint32_t) {};
#endif

  // zero the conv2 operand window once (its guard rows are never written afterwards)
  for (int i = threadIdx.x; i < win_bytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(Y1)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = threadIdx.x; i < (p.nk + 1) * C; i += blockDim.x) {
    const int r = i / C, c = i - r * C;
    float v = 0.f;
    if (r < p.nk) v = p.b1[r][c];
    else for (int q = 0; q < p.nk; ++q) v += p.b2[q][c];
    sbias[i] = v;
  }
  if (p.post_out) {
    const int shift = 3 - (p.post_k - 1) / 2;
    for (int i = threadIdx.x; i < 7 * 32; i += blockDim.x) {
      const int t = i / 32 - shift, c = i & 31;
      spost[i] = (t >= 0 && t < p.post_k && c < C) ? p.post_w[t * C + c] : 0.f;
    }
  }
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_w);
    mbar_init(x_full, 1);
    // the last conv1's commit (of every conv1 issuer) + every epilogue warp (they read the residual from X0)
    mbar_init(x_empty, (p.mma_split ? 2 : 1) + 4 * MRF_EG);
    for (int s = 0; s < p.ring; ++s) {
      mbar_init(&w_full[s], 1);
      // every issuing warp arrives once per slot and lap: the owners of the tap (the stream's one or two warps) through
      // the commit of their MMAs, the warps of the OTHER stream after observing the tap's arrival (skip() below)
      mbar_init(&w_empty[s], p.tap_outer ? 1 : (p.mma_split ? 4 : 2));
    }
    for (int i = 0; i < m; ++i) {
      mbar_init(&d_full[i], 1);
      mbar_init(&d_empty[i], 4 * (C / 32));   // one arrival per epilogue warp of every column half
      mbar_init(&y_ready[i], 4 * (C / 32));
      mbar_init(&s_full[i], 1);
      mbar_init(&e_ready[i], 4);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t d_col = 0, s_col = (uint32_t)(m * C);
  PDL_WAIT();      // the operand window, the validity table and (unit form) the accumulated output are read from here on
#ifdef PIPER_MBAR_WATCHDOG
  if (blockIdx.x == 0 && threadIdx.x == 0)
    printf("mrf_tc_kernel<%d> m %d nk %d ring %d resident %d tap_outer %d split %d unit %d: barriers x_full @%u x_empty +8 w_full +16 "
           "w_empty +%d d_full +%d d_empty +%d y_ready +%d s_full +%d e_ready +%d\n", C, m, p.nk, p.ring, p.resident, p.tap_outer,
           p.mma_split, p.unit_mode, smem_u32(bars), 16 + 8 * MRF_MAX_RING, 16 + 16 * MRF_MAX_RING, 16 + 16 * MRF_MAX_RING + 8 * MRF_MAX_M,
           16 + 16 * MRF_MAX_RING + 16 * MRF_MAX_M, 16 + 16 * MRF_MAX_RING + 24 * MRF_MAX_M, 16 + 16 * MRF_MAX_RING + 32 * MRF_MAX_M);
#endif

  // Convolution pairs are numbered n = it*nk + r over the windows this CTA owns.  The issue order is
  //   conv1(0);  then for n = 0,1,..:  conv2(n) and conv1(n+1)
  // so that the epilogue of pair n+1 never waits for the whole of conv2(n): with resident weights the two
  // are interleaved tile by tile, with the streamed weight ring conv1(n+1) goes first as a block (it only
  // needs D_i back from the epilogue of pair n) and conv2(n) follows.  Producer and MMA warp walk the same
  // sequence, which keeps the single producer thread deadlock-free.
  const int n_it = p.n_windows > (int)blockIdx.x ? (p.n_windows - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const uint32_t n_total = (uint32_t)(n_it * p.nk);

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0 && n_it > 0) {
      uint32_t slot = 0, wph = 0;
      auto load_x = [&](int it) {
        const int w0 = ((int)blockIdx.x + it * (int)gridDim.x) * p.V - p.halo;
        TRACE(20u, (uint32_t)it, 0u);
        mbar_wait(x_empty, ((uint32_t)it & 1u) ^ 1u);
        TRACE(21u, (uint32_t)it, 0u);
        mbar_expect_tx(x_full, (uint32_t)win_x);
        for (int g = 0; g < NP; ++g)
          bulk_load(X0 + (size_t)g * plane_x,
                    p.x_act + ((size_t)g * p.act_rows + (size_t)(p.act_pad + w0 - p.GX)) * 8, (uint32_t)plane_x, x_full);
      };
      auto load_w = [&](int r, int j) {
        for (int t = 0; t < p.k[r]; ++t) {
          TRACE(22u, (uint32_t)(r * 2 + j), (uint32_t)t);
          mbar_wait(&w_empty[slot], wph ^ 1u);
          TRACE(23u, (uint32_t)(r * 2 + j), (uint32_t)t);
          mbar_expect_tx(&w_full[slot], (uint32_t)tap_bytes);
#pragma unroll
          for (int kh = 0; kh < KH; ++kh)
            tma_load_2d(&tmap_w, &w_full[slot], WR + (size_t)slot * tap_bytes + (size_t)kh * sub_bytes, kh * KB,
                        (p.tap_base[r][j] + t) * C);
          if (++slot == (uint32_t)p.ring) { slot = 0; wph ^= 1u; }
        }
      };
      load_x(0);
      if (p.resident) {
        for (int t = 0; t < p.ring; ++t) {          // ring == total taps: slot t holds tap t for the whole kernel
          mbar_expect_tx(&w_full[t], (uint32_t)tap_bytes);
#pragma unroll
          for (int kh = 0; kh < KH; ++kh)
            tma_load_2d(&tmap_w, &w_full[t], WR + (size_t)t * tap_bytes + (size_t)kh * sub_bytes, kh * KB, t * C);
        }
        for (int it = 1; it < n_it; ++it) load_x(it);
      } else {
        load_w(0, 0);
        int it = 0, r = 0;
        for (uint32_t n = 0; n < n_total; ++n) {
          int it1 = it, r1 = r + 1;
          if (r1 == p.nk) { r1 = 0; ++it1; }
          if (n + 1 < n_total) {
            if (r1 == 0) load_x(it1);
            load_w(r1, 0);
          }
          load_w(r, 1);
          it = it1; r = r1;
        }
      }
    }
  } else if (warp <= 4) {
    // ===== MMA issuers: the whole warp runs the (warp-uniform) control flow, one elected lane issues =====
    const uint32_t idesc = make_idesc(C, p.f16);
    // descriptor address units are 16 bytes: one operand row = one unit
    const uint32_t plane16_x = (uint32_t)plane_x >> 4, plane16_y = (uint32_t)plane_bytes >> 4;
    const uint64_t a_hi_x = make_smem_desc_ns(0, (uint32_t)plane_x, 128), a_hi_y = make_smem_desc_ns(0, (uint32_t)plane_bytes, 128);
    const uint64_t b_hi = make_smem_desc(0, KB);
    const uint32_t x0_lo = (smem_u32(X0) >> 4) + (uint32_t)p.GX, y1_lo = (smem_u32(Y1) >> 4) + (uint32_t)p.G;
    const uint32_t wr_lo = smem_u32(WR) >> 4;
    constexpr uint32_t tap16 = tap_bytes >> 4;
    const uint32_t ring = (uint32_t)p.ring;
    // warp 1 issues every conv1, warp 2 every conv2: the two streams only meet through the epilogue's
    // barriers, and one warp alone cannot issue the small N = C MMAs fast enough (tools/micro/umma_bench.cu).
    // In streaming mode both walk the producer's ring order  c1(0); { c1(n+1), c2(n) }  and skip the
    // other stream's slots.
    const int J = (warp - 1) >> 1;          // stream: 0 = conv1, 1 = conv2
    const int half = (warp - 1) & 1;        // tile parity this warp issues (mma_split), or all tiles (warp of half 0)
    const int istep = p.mma_split ? 2 : 1;
    const bool idle = (p.tap_outer && warp != 1) || (!p.mma_split && half == 1);
    uint32_t slot = 0, wph = 0;      // ring position of the next convolution in ring order (streaming mode)
    auto advance = [&](int k) {
      slot += (uint32_t)k;
      while (slot >= ring) { slot -= ring; wph ^= 1u; }
    };
    // A convolution of the OTHER stream.  Its ring slots are not simply stepped over: this warp waits for every tap to
    // land and then arrives on the slot's w_empty like an owner would.  An mbarrier parity wait only tells "phase L
    // done" from "phase L - 1 done"; a warp that steps over a whole lap of a slot (the other stream's convolution
    // covered it) and then waits for lap L + 1 of that slot with the parity of L + 1 = parity of L - 1 passes as soon
    // as it arrives, on weights of the wrong convolution, and its release then corrupts the producer's w_empty count -
    // the round-1 hang of the k = 11 ResBlock1 units (ring 11..14 < 2k; tools/hang_repro.py, mbarrier watchdog:
    // producers stuck on w_empty of the third lap).  Observing every slot in every lap keeps each issuing warp within
    // one phase of every w_full barrier, and a slot cannot be refilled before all issuing warps have passed it, so
    // no barrier can run two phases ahead of a waiter either.
    auto skip = [&](int k) {
      for (int t = 0; t < k; ++t) {
        mbar_wait(&w_full[slot], wph);
        if (lane == 0) mbar_arrive(&w_empty[slot]);
        if (++slot == ring) { slot = 0; wph ^= 1u; }
      }
      __syncwarp();
    };
    // K step ks of a weight tap: 32 bytes further inside the swizzled rows, next sub-tile after KB channels
    auto b_step = [&](int ks) { return (uint32_t)((ks / (KB / 16)) * (sub_bytes >> 4) + 2 * (ks % (KB / 16))); };

    // one convolution over all tiles: per tile k taps x KS tcgen05 K steps, then the commit for the epilogue
    auto conv = [&](uint32_t n, int r, int j, bool first_use, bool fresh = false) {
      const int k = p.k[r];
      const int dil = j ? p.d1[r] : p.d0[r];
      const uint32_t slot0 = p.resident ? (uint32_t)p.tap_base[r][j] : slot;
      const uint32_t wph0 = p.resident ? 0u : wph;
      // streamed ring: a tap's slot goes back to the producer with the MMAs of this warp's LAST tile (not after the whole
      // convolution), so that the next convolution's taps are on their way while the last tile is still being issued
      const int i_first = half * (istep - 1);
      const int i_last = i_first < m ? i_first + ((m - 1 - i_first) / istep) * istep : -1;
      for (int i = i_first; i < m; i += istep) {
        TRACE(1u + (uint32_t)j * 4u, n, (uint32_t)i);
        if (j == 0) {
          mbar_wait(&d_empty[i], (n & 1u) ^ 1u);           // epilogue of the previous pair drained D_i
        } else {                                           // conv2 operand rows of tiles i-1, i, i+1
          if (istep == 2) {
            if (i > 0) mbar_wait(&y_ready[i - 1], n & 1u);
            mbar_wait(&y_ready[i], n & 1u);
          } else if (i == 0) {
            mbar_wait(&y_ready[0], n & 1u);
          }
          if (i + 1 < m) mbar_wait(&y_ready[i + 1], n & 1u);
        }
        tc_fence_after();
        TRACE(2u + (uint32_t)j * 4u, n, (uint32_t)i);
        const uint32_t acc = tmem_base + (j ? s_col : d_col) + (uint32_t)(i * C);
        const uint64_t a_hi = j ? a_hi_y : a_hi_x;
        const uint32_t plane16 = j ? plane16_y : plane16_x;
        uint32_t a_lo = (j ? y1_lo : x0_lo) - (uint32_t)(((k - 1) / 2) * dil) + (uint32_t)(128 * i);
        uint32_t s = slot0, ph = wph0;
        for (int t = 0; t < k; ++t) {
          if (first_use && i < istep) {
            TRACE(24u, n, (uint32_t)(t * 2 + j));
            mbar_wait(&w_full[s], ph);
            tc_fence_after();
            TRACE(25u, n, (uint32_t)(t * 2 + j));
          }
          const uint32_t b_lo = wr_lo + s * tap16;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)   // K step of 16 channels = two planes of A, 32 bytes of B
              umma_f16(acc, a_hi | (uint64_t)(a_lo + (uint32_t)ks * 2u * plane16), b_hi | (uint64_t)(b_lo + b_step(ks)),
                       idesc, ((j == 1 && !fresh) || t > 0 || ks > 0) ? 1u : 0u);
            if (!p.resident && i == i_last) umma_commit(&w_empty[s]);
          }
          a_lo += (uint32_t)dil;
          if (++s == ring) { s = 0; ph ^= 1u; }
        }
        if (elect_one()) umma_commit(j ? &s_full[i] : &d_full[i]);
        TRACE(3u + (uint32_t)j * 4u, n, (uint32_t)i);
      }
      if (!p.resident) {
        if (i_last < 0) {   // a warp without tiles (m = 1, odd warp) still owes its arrival on every slot
          uint32_t s = slot, ph = wph;
          for (int t = 0; t < k; ++t) {
            mbar_wait(&w_full[s], ph);
            if (lane == 0) mbar_arrive(&w_empty[s]);
            if (++s == ring) { s = 0; ph ^= 1u; }
          }
          __syncwarp();
        }
        advance(k);
      }
    };

    // Weight ring shorter than a convolution (C = 128: two 32 KB taps): taps outermost, every tap used by all tiles
    // and handed back at once; warp 1 alone walks the producer's order  c1(0); { c1(n+1), c2(n) }  (a second warp
    // would have to wait several ring generations ahead, which an mbarrier parity cannot express)
    auto conv_to = [&](uint32_t n, int r, int j) {
      const int k = p.k[r];
      const int dil = j ? p.d1[r] : p.d0[r];
      TRACE(1u + (uint32_t)j * 4u, n, 0u);
      for (int i = 0; i < m; ++i) {
        if (j == 0) mbar_wait(&d_empty[i], (n & 1u) ^ 1u);
        else mbar_wait(&y_ready[i], n & 1u);
      }
      tc_fence_after();
      TRACE(2u + (uint32_t)j * 4u, n, 0u);
      const uint64_t a_hi = j ? a_hi_y : a_hi_x;
      const uint32_t plane16 = j ? plane16_y : plane16_x;
      const uint32_t a_base = (j ? y1_lo : x0_lo) - (uint32_t)(((k - 1) / 2) * dil);
      for (int t = 0; t < k; ++t) {
        TRACE(24u, n, (uint32_t)(t * 2 + j));
        mbar_wait(&w_full[slot], wph);
        tc_fence_after();
        TRACE(25u, n, (uint32_t)(t * 2 + j));
        const uint32_t b_lo = wr_lo + slot * tap16;
        if (elect_one()) {   // (warp-uniform tile loops with one election per tile measured 4 % slower here)
          for (int i = 0; i < m; ++i) {
            const uint32_t acc = tmem_base + (j ? s_col : d_col) + (uint32_t)(i * C);
            const uint32_t a_lo = a_base + (uint32_t)(t * dil + 128 * i);
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)
              umma_f16(acc, a_hi | (uint64_t)(a_lo + (uint32_t)ks * 2u * plane16), b_hi | (uint64_t)(b_lo + b_step(ks)),
                       idesc, (j == 1 || t > 0 || ks > 0) ? 1u : 0u);
          }
          umma_commit(&w_empty[slot]);
        }
        __syncwarp();
        if (++slot == ring) { slot = 0; wph ^= 1u; }
      }
      if (elect_one())
        for (int i = 0; i < m; ++i) umma_commit(j ? &s_full[i] : &d_full[i]);
      __syncwarp();
      TRACE(3u + (uint32_t)j * 4u, n, 0u);
    };

    int it = 0, r = 0;
    if (idle) {
      // nothing to issue: a conv stream with a single issuer, or the taps-outermost mode
    } else if (p.tap_outer) {
      if (n_it > 0) {
        mbar_wait(x_full, 0u);
        conv_to(0u, 0, 0);
        if (p.nk == 1 && elect_one()) umma_commit(x_empty);
        for (uint32_t n = 0; n < n_total; ++n) {
          int it1 = it, r1 = r + 1;
          if (r1 == p.nk) { r1 = 0; ++it1; }
          if (n + 1 < n_total) {
            if (r1 == 0) mbar_wait(x_full, (uint32_t)it1 & 1u);
            conv_to(n + 1u, r1, 0);
            if (r1 == p.nk - 1 && elect_one()) umma_commit(x_empty);
          }
          conv_to(n, r, 1);
          it = it1; r = r1;
        }
      }
    } else if (J == 0) {
      if (n_it > 0) {
        mbar_wait(x_full, 0u);
        conv(0u, 0, 0, true);
        if (p.nk == 1 && elect_one()) umma_commit(x_empty);
      }
      for (uint32_t n = 0; n + 1 < n_total; ++n) {
        int it1 = it, r1 = r + 1;
        if (r1 == p.nk) { r1 = 0; ++it1; }
        if (r1 == 0) mbar_wait(x_full, (uint32_t)it1 & 1u);
        conv(n + 1u, r1, 0, !p.resident || it1 == 0);
        // the lrelu(x) window may be refilled once the last conv1 reading it has been issued
        if (r1 == p.nk - 1 && elect_one()) umma_commit(x_empty);
        if (!p.resident) skip(p.k[r]);                       // c2(n) belongs to the other stream
        it = it1; r = r1;
      }
    } else if (p.post_tc) {
      // resident weights: conv2 of every pair, then the conv_post pseudo pair (index nk) of the window, fresh into S_i;
      // the y_ready / s_full phases count nk + 1 steps per window
      uint32_t n2 = 0;
      for (int itw = 0; itw < n_it; ++itw)
        for (int rr = 0; rr <= p.nk; ++rr, ++n2) conv(n2, rr, 1, itw == 0, rr == p.nk);
    } else {
      if (!p.resident && n_it > 0) skip(p.k[0]);             // c1(0)
      for (uint32_t n = 0; n < n_total; ++n) {
        int it1 = it, r1 = r + 1;
        if (r1 == p.nk) { r1 = 0; ++it1; }
        if (!p.resident && n + 1 < n_total) skip(p.k[r1]);   // c1(n+1)
        conv(n, r, 1, !p.resident || it == 0);
        it = it1; r = r1;
      }
    }
  } else {
    // ===== epilogue: MRF_EG groups of four warps.  A unit of work is (tile, 32-column half); group g takes
    // units g, g + MRF_EG, ...  Every thread owns one row and 32 columns: one 128-byte line of the fp32
    // residual, two TMEM loads of D, and two planes of the conv2 operand. =====
    constexpr int HALVES = C / 32;
    const int group = (warp - 5) >> 2;
    const int quarter = warp & 3;                        // TMEM lane quarter this warp may access
    const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
    const float inv_nk = 1.0f / (float)p.nk;
    const int units = m * HALVES;
    uint32_t n = 0;    // pair counter of the conv1 side (d_full / d_empty phases)
    uint32_t n2 = 0;   // step counter of the conv2 side (y_ready / s_full phases): nk per window, nk + 1 with post_tc
    for (int it = 0; it < n_it; ++it) {
      const int w0 = ((int)blockIdx.x + it * (int)gridDim.x) * p.V - p.halo;
      mbar_wait(x_full, (uint32_t)it & 1u);   // the epilogue reads the window itself (residual)
      for (int r = 0; r < p.nk; ++r, ++n, ++n2) {
        const float* b1 = sbias + r * C;
        for (int u = group; u < units; u += MRF_EG) {
          const int i = u / HALVES, cb = (u % HALVES) * 32;
          const int row_in_win = 128 * i + quarter * 32 + lane;
          const int grow = w0 + row_in_win;
          const bool inr = grow >= 0 && grow < p.R;
          // the residual x0 is recovered from the 16-bit lrelu(x0) operand window (x0 = v >= 0 ? v : 10 v): no
          // fp32 copy of the stage input exists.  Window rows and the validity flag do not depend on the MMAs
          const uint32_t row_off = (uint32_t)((p.G + row_in_win) * 16);
          const uint32_t row_off_x = (uint32_t)((p.GX + row_in_win) * 16);
          int ru = 0;
          if (inr && p.row_utt) ru = p.row_utt[p.valid_shift >= 0 ? (grow >> p.valid_shift) : (grow / p.valid_div)];
          TRACE(10u, n, (uint32_t)u);
          if (n2 > 0) {   // conv2 of the previous step must have finished with S_i and with Y1 rows of tiles i-1..i+1
            for (int ii = (i > 0 ? i - 1 : 0); ii <= (i + 1 < m ? i + 1 : m - 1); ++ii)
              mbar_wait(&s_full[ii], (n2 - 1u) & 1u);
          }
          TRACE(11u, n, (uint32_t)u);
          mbar_wait(&d_full[i], n & 1u);
          tc_fence_after();
          TRACE(12u, n, (uint32_t)u);
          asm volatile("" : "+r"(ru));                   // keep the flag's load in flight across the waits
          const bool valid = inr && ru >= 0;
          const uint32_t dcol = tmem_base + lane_sel + d_col + (uint32_t)(i * C + cb);
          const uint32_t scol = tmem_base + lane_sel + s_col + (uint32_t)(i * C + cb);
          // two chunks of 16 columns (= two operand planes) each, processed one after the other to keep the
          // register footprint small enough for MRF_EG = 4 warpgroups
#pragma unroll 1
          for (int hc = 0; hc < 2; ++hc) {
            const int c0 = cb + hc * 16;
            const uint8_t* xrow = X0 + (size_t)(c0 >> 3) * plane_x + row_off_x;
            const uint4 xa = *reinterpret_cast<const uint4*>(xrow);
            const uint4 xb = *reinterpret_cast<const uint4*>(xrow + plane_x);
            uint32_t t0[16], t2[16];
            __syncwarp();
            tmem_ld16_async(dcol + (uint32_t)(hc * 16), t0);
            if (r > 0) tmem_ld16_async(scol + (uint32_t)(hc * 16), t2);
#ifndef PIPER_MRF_EPI_SCALAR
            if (!UNIT) {
              // packed form (FADD2 / FMUL2 on fp32 pairs, the leaky relu of the operand in the 16-bit type): the 16
              // epilogue warps are bound by issue slots, this is ~40 % fewer instructions per chunk than the scalar form
              uint64_t d2[8];
              {
                const uint32_t w[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
                const ulonglong2* b2 = reinterpret_cast<const ulonglong2*>(b1 + c0);
                const uint64_t ten = f2pack(10.f, 10.f);
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                  const ulonglong2 bb = b2[q4];
                  const float2 x01 = unpack2(w[2 * q4], p.f16), x23 = unpack2(w[2 * q4 + 1], p.f16);
                  const float2 t01 = f2unpack(f2mul(f2pack(x01.x, x01.y), ten)), t23 = f2unpack(f2mul(f2pack(x23.x, x23.y), ten));
                  d2[2 * q4] = f2add(bb.x, f2pack(fminf(x01.x, t01.x), fminf(x01.y, t01.y)));
                  d2[2 * q4 + 1] = f2add(bb.y, f2pack(fminf(x23.x, t23.x), fminf(x23.y, t23.y)));
                }
              }
              tmem_wait_ld();
              float dv[16];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                d2[e] = f2add(d2[e], f2pack(__uint_as_float(t0[2 * e]), __uint_as_float(t0[2 * e + 1])));
                const float2 v = f2unpack(d2[e]);
                dv[2 * e] = v.x; dv[2 * e + 1] = v.y;
              }
              if (r > 0) {
                float sv[16];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const float2 v = f2unpack(f2add(d2[e], f2pack(__uint_as_float(t2[2 * e]), __uint_as_float(t2[2 * e + 1]))));
                  sv[2 * e] = v.x; sv[2 * e + 1] = v.y;
                }
                tmem_st16(scol + (uint32_t)(hc * 16), sv);
              } else {
                tmem_st16(scol + (uint32_t)(hc * 16), dv);
              }
              const uint32_t vm = valid ? 0xffffffffu : 0u;
              uint32_t pk[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) pk[e] = lrelu_pack2(dv[2 * e], dv[2 * e + 1], 0.1f, p.f16) & vm;
              uint8_t* y = Y1 + (size_t)(c0 >> 3) * plane_bytes + row_off;   // planes c0/8 and c0/8 + 1
              *reinterpret_cast<uint4*>(y) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              *reinterpret_cast<uint4*>(y + plane_bytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
              continue;
            }
#endif
            float d[16], xr[16];
            {   // x0 + b1 while the TMEM loads are in flight; x0 = min(v, 10 v) undoes lrelu(., 0.1)
              const uint32_t w[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
              const float4* b4 = reinterpret_cast<const float4*>(b1 + c0);
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) {
                const float4 bb = b4[q4];
                const float2 x01 = unpack2(w[2 * q4], p.f16), x23 = unpack2(w[2 * q4 + 1], p.f16);
                if (UNIT) {
                  xr[4 * q4] = fminf(x01.x, x01.x * 10.f);
                  xr[4 * q4 + 1] = fminf(x01.y, x01.y * 10.f);
                  xr[4 * q4 + 2] = fminf(x23.x, x23.x * 10.f);
                  xr[4 * q4 + 3] = fminf(x23.y, x23.y * 10.f);
                  d[4 * q4] = bb.x; d[4 * q4 + 1] = bb.y; d[4 * q4 + 2] = bb.z; d[4 * q4 + 3] = bb.w;
                } else {
                  d[4 * q4] = bb.x + fminf(x01.x, x01.x * 10.f);
                  d[4 * q4 + 1] = bb.y + fminf(x01.y, x01.y * 10.f);
                  d[4 * q4 + 2] = bb.z + fminf(x23.x, x23.x * 10.f);
                  d[4 * q4 + 3] = bb.w + fminf(x23.y, x23.y * 10.f);
                }
              }
            }
            tmem_wait_ld();
#pragma unroll
            for (int e = 0; e < 16; ++e) d[e] += __uint_as_float(t0[e]);
            if (UNIT) {                       // ResBlock1 unit: the stream starts from x, conv2 accumulates onto it
              tmem_st16(scol + (uint32_t)(hc * 16), xr);
            } else if (r > 0) {
              float sv[16];
#pragma unroll
              for (int e = 0; e < 16; ++e) sv[e] = __uint_as_float(t2[e]) + d[e];
              tmem_st16(scol + (uint32_t)(hc * 16), sv);
            } else {
              tmem_st16(scol + (uint32_t)(hc * 16), d);
            }
            uint32_t pk[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)   // lrelu(v, 0.1) = max(v, 0.1 v)
              pk[e] = valid ? pack2(fmaxf(d[2 * e], d[2 * e] * 0.1f), fmaxf(d[2 * e + 1], d[2 * e + 1] * 0.1f), p.f16) : 0u;
            uint8_t* y = Y1 + (size_t)(c0 >> 3) * plane_bytes + row_off;   // planes c0/8 and c0/8 + 1
            *reinterpret_cast<uint4*>(y) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            *reinterpret_cast<uint4*>(y + plane_bytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
          tmem_wait_st();
          fence_proxy_async();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&y_ready[i]);
            mbar_arrive(&d_empty[i]);
          }
          TRACE(13u, n, (uint32_t)u);
        }
        if (r == p.nk - 1) {   // this warp is done with the lrelu(x) window
          __syncwarp();
          if (lane == 0) mbar_arrive(x_empty);
        }
      }
      // ---- stage output: (S + sum_r b2_r) / nk over the window's valid rows ----
      const float* b2s = sbias + p.nk * C;
      if (C == 32 && p.post_out && p.post_tc) {
        // conv_post + tanh with the taps on the tensor cores.  Step 1 (per unit = tile): v = lrelu((S + sum b2) / nk,
        // post_slope) * mask -> 16-bit -> the conv2 operand window (what a conv1 epilogue does with y1); the conv2
        // stream then runs the post taps over it, fresh into S_i.  Step 2: column 0 of S_i is the pre-tanh sample.
        // The accumulation order of a sample (taps, K steps) does not depend on its row's position in the window.
        constexpr int MAXU = (MRF_MAX_M + MRF_EG - 1) / MRF_EG;
        uint32_t vmask = 0u;
#pragma unroll
        for (int ui = 0; ui < MAXU; ++ui) {
          const int i = group + ui * MRF_EG;
          if (i < units) {
            const int row_in_win = 128 * i + quarter * 32 + lane;
            const int grow = w0 + row_in_win;
            const bool inr = grow >= 0 && grow < p.R;
            int ru = 0;
            if (inr && p.row_utt) ru = p.row_utt[p.valid_shift >= 0 ? (grow >> p.valid_shift) : (grow / p.valid_div)];
            TRACE(14u, n2, (uint32_t)i);
            for (int ii = (i > 0 ? i - 1 : 0); ii <= (i + 1 < m ? i + 1 : m - 1); ++ii)
              mbar_wait(&s_full[ii], (n2 - 1u) & 1u);
            tc_fence_after();
            TRACE(15u, n2, (uint32_t)i);
            asm volatile("" : "+r"(ru));
            const bool valid = inr && ru >= 0;
            vmask |= (valid ? 1u : 0u) << ui;
            const uint32_t scol = tmem_base + lane_sel + s_col + (uint32_t)(i * C);
            uint32_t t0[16], t1[16];
            __syncwarp();
            tmem_ld16_async(scol, t0);
            tmem_ld16_async(scol + 16u, t1);
            tmem_wait_ld();
            uint32_t pk[16];
            {
              const uint64_t inv2 = f2pack(inv_nk, inv_nk);
              const uint32_t vm = valid ? 0xffffffffu : 0u;
              const ulonglong2* bq = reinterpret_cast<const ulonglong2*>(b2s);
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) {
                const ulonglong2 ba = bq[q4], bc = bq[4 + q4];
                const float2 a01 = f2unpack(f2mul(f2add(f2pack(__uint_as_float(t0[4 * q4]), __uint_as_float(t0[4 * q4 + 1])), ba.x), inv2));
                const float2 a23 = f2unpack(f2mul(f2add(f2pack(__uint_as_float(t0[4 * q4 + 2]), __uint_as_float(t0[4 * q4 + 3])), ba.y), inv2));
                const float2 c01 = f2unpack(f2mul(f2add(f2pack(__uint_as_float(t1[4 * q4]), __uint_as_float(t1[4 * q4 + 1])), bc.x), inv2));
                const float2 c23 = f2unpack(f2mul(f2add(f2pack(__uint_as_float(t1[4 * q4 + 2]), __uint_as_float(t1[4 * q4 + 3])), bc.y), inv2));
                pk[2 * q4] = lrelu_pack2(a01.x, a01.y, p.post_slope, p.f16) & vm;
                pk[2 * q4 + 1] = lrelu_pack2(a23.x, a23.y, p.post_slope, p.f16) & vm;
                pk[8 + 2 * q4] = lrelu_pack2(c01.x, c01.y, p.post_slope, p.f16) & vm;
                pk[8 + 2 * q4 + 1] = lrelu_pack2(c23.x, c23.y, p.post_slope, p.f16) & vm;
              }
            }
            uint8_t* y = Y1 + (size_t)((p.G + row_in_win) * 16);
#pragma unroll
            for (int g = 0; g < 4; ++g)
              *reinterpret_cast<uint4*>(y + (size_t)g * plane_bytes) = make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
            fence_proxy_async();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&y_ready[i]);
            TRACE(16u, n2, (uint32_t)i);
          }
        }
#pragma unroll
        for (int ui = 0; ui < MAXU; ++ui) {
          const int i = group + ui * MRF_EG;
          if (i < units) {
            const int row_in_win = 128 * i + quarter * 32 + lane;
            const int grow = w0 + row_in_win;
            const bool owned = row_in_win >= p.halo && row_in_win < p.halo + p.V && grow >= 0 && grow < p.R;
            mbar_wait(&s_full[i], n2 & 1u);
            tc_fence_after();
            __syncwarp();
            const float sum = tmem_ld1(tmem_base + lane_sel + s_col + (uint32_t)(i * C));
            if (owned) p.post_out[grow] = ((vmask >> ui) & 1u) ? tanhf(sum + p.post_bias) : 0.f;
            tc_fence_before();
          }
        }
        ++n2;
        continue;
      }
      const uint32_t nl = n2 - 1u;
      if (C == 32 && p.post_out) {
        // Fused conv_post + tanh.  Every thread owns one row with all 32 channels: it forms the seven per-tap channel
        // sums q_t = sum_c w[t][c] * lrelu(out[row][c]); audio[r] = tanh(sum_t q_t[r + t - 3]) then needs the q of the
        // three rows on either side - lanes of the same warp (shuffles) or, at the 32-row block edges, of a neighbouring
        // warp (through `sedge`, published per tile on `e_ready`: only neighbouring tiles couple, and `sedge` is double
        // buffered by window parity so that a fast warp cannot overwrite what a slow neighbour still reads).  The tap
        // sum always runs t = 0..6 whatever the source, so a sample does not depend on where its row falls in a window.
        float* sedge_w = sedge + (it & 1) * (4 * m * 42);
        constexpr int MAXU = (MRF_MAX_M + MRF_EG - 1) / MRF_EG;
        float q[MAXU][7];
        uint32_t vmask = 0u;
#pragma unroll
        for (int ui = 0; ui < MAXU; ++ui) {
          const int i = group + ui * MRF_EG;               // unit = tile (one 32-column half)
          if (i < units) {
            const int row_in_win = 128 * i + quarter * 32 + lane;
            const int grow = w0 + row_in_win;
            const bool inr = grow >= 0 && grow < p.R;
            int ru = 0;
            if (inr && p.row_utt) ru = p.row_utt[p.valid_shift >= 0 ? (grow >> p.valid_shift) : (grow / p.valid_div)];
            TRACE(14u, nl, (uint32_t)i);
            mbar_wait(&s_full[i], nl & 1u);
            tc_fence_after();
            TRACE(15u, nl, (uint32_t)i);
            asm volatile("" : "+r"(ru));
            const bool valid = inr && ru >= 0;
            vmask |= (valid ? 1u : 0u) << ui;
            const uint32_t scol = tmem_base + lane_sel + s_col + (uint32_t)(i * C);
            uint32_t t0[16], t1[16];
            __syncwarp();
            tmem_ld16_async(scol, t0);
            tmem_ld16_async(scol + 16u, t1);
            tmem_wait_ld();
            float x[32];
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              x[e] = valid ? lrelu((__uint_as_float(t0[e]) + b2s[e]) * inv_nk, p.post_slope) : 0.f;
              x[16 + e] = valid ? lrelu((__uint_as_float(t1[e]) + b2s[16 + e]) * inv_nk, p.post_slope) : 0.f;
            }
#pragma unroll
            for (int t = 0; t < 7; ++t) {
              const float4* w4 = reinterpret_cast<const float4*>(spost + t * 32);
              float a = 0.f;
#pragma unroll
              for (int c4 = 0; c4 < 8; ++c4) {
                const float4 w = w4[c4];
                a = fmaf(x[4 * c4], w.x, a);
                a = fmaf(x[4 * c4 + 1], w.y, a);
                a = fmaf(x[4 * c4 + 2], w.z, a);
                a = fmaf(x[4 * c4 + 3], w.w, a);
              }
              q[ui][t] = a;
            }
            const int blk = i * 4 + quarter;
            if (lane < 3 || lane >= 29) {
              float* e = sedge_w + ((blk * 2 + (lane < 3 ? 0 : 1)) * 3 + (lane < 3 ? lane : lane - 29)) * 7;
#pragma unroll
              for (int t = 0; t < 7; ++t) e[t] = q[ui][t];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&e_ready[i]);
          }
        }
#pragma unroll
        for (int ui = 0; ui < MAXU; ++ui) {
          const int i = group + ui * MRF_EG;
          if (i < units) {
            const int row_in_win = 128 * i + quarter * 32 + lane;
            const int grow = w0 + row_in_win;
            const bool owned = row_in_win >= p.halo && row_in_win < p.halo + p.V && grow >= 0 && grow < p.R;
            const int blk = i * 4 + quarter;
            for (int ii = (i > 0 ? i - 1 : 0); ii <= (i + 1 < m ? i + 1 : m - 1); ++ii)
              mbar_wait(&e_ready[ii], (uint32_t)it & 1u);
            float sum = 0.f;
#pragma unroll
            for (int t = 0; t < 7; ++t) {
              const int src = lane + t - 3;
              float v = __shfl_sync(0xffffffffu, q[ui][t], src & 31);
              if (src < 0) v = blk > 0 ? sedge_w[(((blk - 1) * 2 + 1) * 3 + (src + 3)) * 7 + t] : 0.f;
              else if (src > 31) v = blk + 1 < 4 * m ? sedge_w[(((blk + 1) * 2) * 3 + (src - 32)) * 7 + t] : 0.f;
              sum += v;
            }
            if (owned) p.post_out[grow] = ((vmask >> ui) & 1u) ? tanhf(sum + p.post_bias) : 0.f;
            TRACE(16u, nl, (uint32_t)i);
          }
        }
      } else
      for (int u = group; u < units; u += MRF_EG) {
        const int i = u / HALVES, cb = (u % HALVES) * 32;
        const int row_in_win = 128 * i + quarter * 32 + lane;
        const int grow = w0 + row_in_win;
        const bool owned = row_in_win >= p.halo && row_in_win < p.halo + p.V && grow >= 0 && grow < p.R;
        int ru = 0;
        if (owned && p.row_utt) ru = p.row_utt[p.valid_shift >= 0 ? (grow >> p.valid_shift) : (grow / p.valid_div)];
        TRACE(14u, nl, (uint32_t)u);
        mbar_wait(&s_full[i], nl & 1u);
        tc_fence_after();
        TRACE(15u, nl, (uint32_t)u);
        asm volatile("" : "+r"(ru));
        const bool valid = owned && ru >= 0;
        const uint32_t scol = tmem_base + lane_sel + s_col + (uint32_t)(i * C + cb);
        float s0[16], s1[16];
        {
          uint32_t t0[16], t1[16];
          __syncwarp();
          tmem_ld16_async(scol, t0);
          tmem_ld16_async(scol + 16u, t1);
          tmem_wait_ld();
#pragma unroll
          for (int e = 0; e < 16; ++e) { s0[e] = __uint_as_float(t0[e]); s1[e] = __uint_as_float(t1[e]); }
        }
        if (!UNIT && owned && p.out_act && !p.out_f32) {
          // ResBlock2 stage whose only output is the next ConvTranspose's 16-bit operand: packed fp32 pairs, leaky
          // relu in the 16-bit type
          const uint64_t inv2 = f2pack(inv_nk, inv_nk);
          const uint32_t vm = valid ? 0xffffffffu : 0u;
          const ulonglong2* bq = reinterpret_cast<const ulonglong2*>(b2s + cb);
          uint32_t pk[16];
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const ulonglong2 ba = bq[q4], bc = bq[4 + q4];
            const float2 a01 = f2unpack(f2mul(f2add(f2pack(s0[4 * q4], s0[4 * q4 + 1]), ba.x), inv2));
            const float2 a23 = f2unpack(f2mul(f2add(f2pack(s0[4 * q4 + 2], s0[4 * q4 + 3]), ba.y), inv2));
            const float2 c01 = f2unpack(f2mul(f2add(f2pack(s1[4 * q4], s1[4 * q4 + 1]), bc.x), inv2));
            const float2 c23 = f2unpack(f2mul(f2add(f2pack(s1[4 * q4 + 2], s1[4 * q4 + 3]), bc.y), inv2));
            pk[2 * q4] = lrelu_pack2(a01.x, a01.y, p.act_slope, p.f16) & vm;
            pk[2 * q4 + 1] = lrelu_pack2(a23.x, a23.y, p.act_slope, p.f16) & vm;
            pk[8 + 2 * q4] = lrelu_pack2(c01.x, c01.y, p.act_slope, p.f16) & vm;
            pk[8 + 2 * q4 + 1] = lrelu_pack2(c23.x, c23.y, p.act_slope, p.f16) & vm;
          }
          uint4* dst = reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out_act) + (size_t)grow * C + cb);
#pragma unroll
          for (int g = 0; g < 4; ++g) dst[g] = make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
        } else if (owned) {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            s0[e] = valid ? (s0[e] + b2s[cb + e]) * inv_nk : 0.f;
            s1[e] = valid ? (s1[e] + b2s[cb + 16 + e]) * inv_nk : 0.f;
          }
          if (UNIT && p.accum) {   // cross-resblock sum of a ResBlock1 stage lives in out_f32
            const float4* a4 = reinterpret_cast<const float4*>(p.out_f32 + (size_t)grow * C + cb);
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const float4 a = a4[q4], b = a4[4 + q4];
              s0[4 * q4] += a.x; s0[4 * q4 + 1] += a.y; s0[4 * q4 + 2] += a.z; s0[4 * q4 + 3] += a.w;
              s1[4 * q4] += b.x; s1[4 * q4 + 1] += b.y; s1[4 * q4 + 2] += b.z; s1[4 * q4 + 3] += b.w;
            }
          }
          if (UNIT && p.out_div != 0.f) {
#pragma unroll
            for (int e = 0; e < 16; ++e) { s0[e] = s0[e] / p.out_div; s1[e] = s1[e] / p.out_div; }
          }
          if (UNIT && !valid) {
#pragma unroll
            for (int e = 0; e < 16; ++e) { s0[e] = 0.f; s1[e] = 0.f; }
          }
          if (UNIT && p.out_planar) {   // operand of the next fused launch: planes cb/8 .. cb/8 + 3, one 16-byte piece per row each
            uint16_t* pl = p.out_planar + ((size_t)(cb >> 3) * p.out_rows + (size_t)(p.out_pad + grow)) * 8;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const float* v = g < 2 ? s0 + 8 * g : s1 + 8 * (g - 2);
              uint32_t pk[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) pk[e] = pack2(lrelu(v[2 * e], p.act_slope), lrelu(v[2 * e + 1], p.act_slope), p.f16);
              *reinterpret_cast<uint4*>(pl + (size_t)g * p.out_rows * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
          }
          if (p.out_f32) {
            float4* o4 = reinterpret_cast<float4*>(p.out_f32 + (size_t)grow * C + cb);
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              o4[q4] = make_float4(s0[4 * q4], s0[4 * q4 + 1], s0[4 * q4 + 2], s0[4 * q4 + 3]);
              o4[4 + q4] = make_float4(s1[4 * q4], s1[4 * q4 + 1], s1[4 * q4 + 2], s1[4 * q4 + 3]);
            }
          }
          if (p.out_act) {
            uint32_t pk[16];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              pk[e] = pack2(lrelu(s0[2 * e], p.act_slope), lrelu(s0[2 * e + 1], p.act_slope), p.f16);
              pk[8 + e] = pack2(lrelu(s1[2 * e], p.act_slope), lrelu(s1[2 * e + 1], p.act_slope), p.f16);
            }
            uint4* dst = reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out_act) + (size_t)grow * C + cb);
#pragma unroll
            for (int g = 0; g < 4; ++g) dst[g] = make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
          }
        }
        tc_fence_before();
        TRACE(16u, nl, (uint32_t)u);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

}  // namespace

bool mrf_tc_supported(int C, int nk, const int* k, const int* d0, const int* d1) {
  if (C != 32 && C != 64 && C != 128) return false;
  if (nk < 1 || nk > MRF_MAX_RB) return false;
  int taps = 0;
  for (int r = 0; r < nk; ++r) {
    if (k[r] < 1 || (k[r] & 1) == 0) return false;
    const int h = (k[r] - 1) / 2;
    if (h * d0[r] > MRF_G || h * d1[r] > MRF_G) return false;
    if (h * (d0[r] + d1[r]) > 60) return false;          // keeps V = 128*m - 2*halo positive for m >= 2
    taps += 2 * k[r];
  }
  return taps > 0;
}

bool mrf_post_supported(int C, int post_k) { return C == 32 && post_k >= 1 && post_k <= 7 && (post_k & 1) == 1; }

void launch_mrf_tc(Exec& ex, const MrfArgs& a) {
  if (a.R <= 0) return;
  const bool post = a.post_out != nullptr;
  if (post && (!mrf_post_supported(a.C, a.post_k) || !a.post_w))
    throw std::runtime_error("conv_mrf: fused conv_post needs C = 32 and an odd kernel of at most 7 taps");
  if (!mrf_tc_supported(a.C, a.nk, a.k, a.d0, a.d1)) throw std::runtime_error("conv_mrf: unsupported stage shape");
  static int num_sms = 0;
  static PerDeviceOnce once;
  once.run([&] {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || num_sms <= 0)
      throw std::runtime_error("cuda: cannot query the SM count");
    if (cudaFuncSetAttribute(mrf_tc_kernel<32, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<64, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<32, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<64, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<32, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<64, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<32, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<64, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<128, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<128, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<128, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<32, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<64, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<128, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess)
      throw std::runtime_error("cuda: cannot raise dynamic shared memory limit for mrf_tc_kernel");
  });
  MrfParams p{};
  p.R = a.R; p.C = a.C; p.nk = a.nk;
  int halo = 0, kmax = 0;
  for (int r = 0; r < a.nk; ++r) {
    p.k[r] = a.k[r]; p.d0[r] = a.d0[r]; p.d1[r] = a.d1[r];
    p.tap_base[r][0] = a.tap_base[r][0]; p.tap_base[r][1] = a.tap_base[r][1];
    p.b1[r] = a.b1[r]; p.b2[r] = a.b2[r];
    halo = std::max(halo, (a.k[r] - 1) / 2 * (a.d0[r] + a.d1[r]));
    kmax = std::max(kmax, a.k[r]);
  }
  // with the fused conv_post the owned rows also need correct neighbours (post_k - 1) / 2 rows away
  if (post) halo += (a.post_k - 1) / 2;
  p.halo = halo;
  const int row_bytes = a.C * 2, tap_bytes = a.C * a.C * 2;
  if (a.act_pad < MRF_G + 64 || a.act_rows < (int64_t)a.R + a.act_pad + 128 * (MRF_MAX_M + 1))
    throw std::runtime_error("conv_mrf: planar operand needs zero pad rows around the signal");
  // window height: TMEM holds D and S for every tile (2*m*C columns <= 512); smem holds two operand
  // windows of (m+1)*128 rows plus at least kmax weight taps
  int shift = 0;   // conv2 shifts: guard rows of the conv2 operand window
  for (int r = 0; r < a.nk; ++r) shift = std::max(shift, (a.k[r] - 1) / 2 * a.d1[r]);
  static const int g_env = [] { const char* e = getenv("PIPER_B200_MRF_G"); return e ? atoi(e) : 0; }();
  int G = (shift + 7) & ~7;
  if (g_env >= G && g_env <= MRF_G && g_env % 8 == 0) G = g_env;
  p.G = G;
  // the lrelu(x) window only sees the conv1 shifts
  int shift_x = 0;
  for (int r = 0; r < a.nk; ++r) shift_x = std::max(shift_x, (a.k[r] - 1) / 2 * a.d0[r]);
  static const int gx_env = [] { const char* e = getenv("PIPER_B200_MRF_GX"); return e ? atoi(e) : 1; }();
  const int GX = gx_env ? ((shift_x + 7) & ~7) : std::max(G, (shift_x + 7) & ~7);
  p.GX = GX;
  // C = 128: a tap is 32 KB, the ring holds two of them and the taps-outermost issue order is used
  // ResBlock1 units whose 2k taps are not resident also take the taps-outermost order (a single issuing warp): with
  // k = 11 a streamed ring would hold barely more than one convolution
  static const int unit_stream_env0 = [] { const char* e = getenv("PIPER_B200_MRF_UNIT_STREAM"); return e ? atoi(e) : 1; }();
  const int need_ring = (a.C > 64 || (a.unit_mode && !unit_stream_env0)) ? 2 : kmax;
  int m = std::min(MRF_MAX_M, 512 / (2 * a.C));
  auto post_bytes = [&](int mm) { return post ? (7 * 32 + 2 * mm * 4 * 2 * 3 * 7) * 4 + 64 : 0; };
  auto fits = [&](int mm) {
    return (2 * mm * 128 + 2 * G + 2 * GX) * row_bytes + need_ring * tap_bytes + 4096 + post_bytes(mm) <= TC_SMEM_MAX;
  };
  while (m > 2 && !fits(m)) --m;
  // small inputs: prefer more, shorter windows so that every SM has one
  while (m > 2 && (a.R + (128 * m - 2 * halo) - 1) / (128 * m - 2 * halo) < num_sms) --m;
  static const int m_env = [] { const char* e = getenv("PIPER_B200_MRF_M"); return e ? atoi(e) : 0; }();
  if (m_env >= 2 && m_env <= MRF_MAX_M && 2 * m_env * a.C <= 512 && fits(m_env)) m = m_env;
  static const int m64_env = [] { const char* e = getenv("PIPER_B200_MRF_M64"); return e ? atoi(e) : 0; }();
  if (a.C == 64 && m64_env >= 2 && 2 * m64_env * a.C <= 512 && fits(m64_env)) m = m64_env;
  if (!fits(m)) throw std::runtime_error("conv_mrf: stage does not fit in shared memory");
  p.m = m;
  p.V = 128 * m - 2 * halo;
  p.n_windows = (a.R + p.V - 1) / p.V;
  const int win_bytes = (m * 128 + G + GX) * row_bytes;   // mean of the two windows
  int ring = (TC_SMEM_MAX - 4096 - post_bytes(m) - 2 * win_bytes) / tap_bytes;
  if (ring > MRF_MAX_RING) ring = MRF_MAX_RING;
  p.resident = a.total_taps <= ring ? 1 : 0;
  if (p.resident) ring = a.total_taps;
  p.ring = ring;
  // ResBlock1 units take the streamed-ring order with two issuing warps per stream whenever a whole convolution fits in
  // the ring (the form that hung in round 1 until the issuing warps observed every slot, see skip() in the kernel);
  // PIPER_B200_MRF_UNIT_STREAM=0 restores the taps-outermost single-issuer order for them
  static const int unit_stream_env = [] { const char* e = getenv("PIPER_B200_MRF_UNIT_STREAM"); return e ? atoi(e) : 1; }();
  p.tap_outer = (!p.resident && (ring < kmax || (a.unit_mode && !unit_stream_env))) ? 1 : 0;
  static const int split_env = [] { const char* e = getenv("PIPER_B200_MRF_SPLIT"); return e ? atoi(e) : 1; }();
  p.mma_split = (split_env && !p.tap_outer) ? 1 : 0;
  p.x_act = static_cast<const uint16_t*>(a.x_act); p.act_rows = a.act_rows; p.act_pad = a.act_pad;
  p.row_utt = a.row_utt; p.valid_div = a.valid_div < 1 ? 1 : a.valid_div;
  p.valid_shift = -1;
  if ((p.valid_div & (p.valid_div - 1)) == 0) { p.valid_shift = 0; while ((1 << p.valid_shift) < p.valid_div) ++p.valid_shift; }
  p.out_act = a.out_act; p.act_slope = a.act_slope; p.out_f32 = a.out_f32;
  if (a.unit_mode && a.nk != 1) throw std::runtime_error("conv_mrf: a ResBlock1 unit is one convolution pair");
  if (!a.unit_mode && (a.accum || a.out_div != 0.f || a.out_planar))
    throw std::runtime_error("conv_mrf: accum / out_div / out_planar belong to the ResBlock1 unit form");
  if (a.accum && !a.out_f32) throw std::runtime_error("conv_mrf: accum needs out_f32");
  p.unit_mode = a.unit_mode; p.accum = a.accum; p.out_div = a.out_div;
  p.out_planar = static_cast<uint16_t*>(a.out_planar); p.out_rows = a.out_rows; p.out_pad = a.out_pad;
  p.post_w = a.post_w; p.post_k = a.post_k; p.post_slope = a.post_slope; p.post_out = a.post_out; p.post_bias = a.post_bias;
  static const int post_tc_env = [] { const char* e = getenv("PIPER_B200_MRF_POST_TC"); return e ? atoi(e) : 1; }();
  p.post_tc = (post && a.post_tap_base >= 0 && p.resident && !a.unit_mode && a.nk < MRF_MAX_RB + 1 && post_tc_env) ? 1 : 0;
  if (p.post_tc) {
    p.k[a.nk] = a.post_k; p.d0[a.nk] = 1; p.d1[a.nk] = 1;
    p.tap_base[a.nk][0] = a.post_tap_base; p.tap_base[a.nk][1] = a.post_tap_base;
    if ((a.post_k - 1) / 2 > G) throw std::runtime_error("conv_mrf: conv_post taps exceed the guard rows");
  }
  p.f16 = a.f16;
  uint32_t cols = 32;
  while ((int)cols < 2 * m * a.C) cols <<= 1;
  p.tmem_cols = cols;
  const CUtensorMap tw = make_map(a.w, (uint64_t)a.C, (uint64_t)a.total_taps * a.C, (uint64_t)a.C, std::min(a.C, 64), a.C, a.f16);
  const size_t smem = (size_t)2 * win_bytes + (size_t)ring * tap_bytes + 1024 + 3072 + (size_t)post_bytes(m);
  if (smem > (size_t)TC_SMEM_MAX) throw std::runtime_error("conv_mrf: shared memory budget exceeded");
  const int grid = std::min(p.n_windows, num_sms);
  ProfScope ps(ex, Prof::MRF, 2.0 * a.R * a.C * a.C * (double)a.total_taps);
#ifdef PIPER_MRF_TRACE
  static const char* trace_path = getenv("PIPER_B200_MRF_TRACE");
#else
  static const char* trace_path = nullptr;
#endif
  const size_t trace_words = (size_t)(160 + 128 * 5) / 32 * MRF_TRACE_EVENTS * 2;
  if (trace_path) {
    if (cudaMalloc(&p.trace, trace_words * 4) != cudaSuccess) throw std::runtime_error("conv_mrf: trace buffer");
    cudaMemsetAsync(p.trace, 0, trace_words * 4, ex.st);
  }
  static const int eg_env = [] { const char* e = getenv("PIPER_B200_MRF_EG"); return e ? atoi(e) : 0; }();
  // epilogue warpgroups: four, three at C = 64 (re-swept after the packed epilogue of round 2: 1.99 -> 1.93 ms for the
  // C = 64 stage with three, C = 32 is slower with three or five, profiles/r02_mrf_sweep_eg.log)
  const int eg = eg_env >= 2 && eg_env <= 5 ? eg_env : (a.C == 64 ? 3 : 4);
  const int threads = 160 + 128 * eg;
  // plain launch on request (no_pdl): the row-range launches of the last stage must not be staged early - their CTAs would
  // take the SMs the instant the previous range retires, ahead of the copy stream's gather kernel that waits for that range
  auto L = [&](auto kernel, int nthreads) {
    if (a.no_pdl) kernel<<<grid, nthreads, smem, ex.st>>>(tw, p);
    else launch_pdl(kernel, dim3(grid), dim3(nthreads), smem, ex.st, tw, p);
  };
  if (a.unit_mode) {
    if (a.C == 128) L(mrf_tc_kernel<128, 4, true>, 160 + 128 * 4);
    else if (a.C == 64) L(mrf_tc_kernel<64, 4, true>, 160 + 128 * 4);
    else L(mrf_tc_kernel<32, 4, true>, 160 + 128 * 4);
  } else if (a.C == 128) {
    if (eg == 2) L(mrf_tc_kernel<128, 2>, threads);
    else if (eg == 3) L(mrf_tc_kernel<128, 3>, threads);
    else L(mrf_tc_kernel<128, 4>, 160 + 128 * 4);
  } else if (a.C == 32) {
    if (eg == 2) L(mrf_tc_kernel<32, 2>, threads);
    else if (eg == 3) L(mrf_tc_kernel<32, 3>, threads);
    else if (eg == 4) L(mrf_tc_kernel<32, 4>, threads);
    else L(mrf_tc_kernel<32, 5>, threads);
  } else {
    if (eg == 2) L(mrf_tc_kernel<64, 2>, threads);
    else if (eg == 3) L(mrf_tc_kernel<64, 3>, threads);
    else if (eg == 4) L(mrf_tc_kernel<64, 4>, threads);
    else L(mrf_tc_kernel<64, 5>, threads);
  }
  ++ex.launches;
  if (trace_path) {   // debug only: dump "C m warp tag n i clock" lines, one file per stage width
    std::vector<uint32_t> h(trace_words);
    cudaStreamSynchronize(ex.st);
    cudaMemcpy(h.data(), p.trace, trace_words * 4, cudaMemcpyDeviceToHost);
    cudaFree(p.trace);
    const std::string fn = std::string(trace_path) + "." + std::to_string(a.C);
    if (FILE* f = fopen(fn.c_str(), "w")) {
      for (size_t w = 0; w < trace_words / 2 / MRF_TRACE_EVENTS; ++w)
        for (size_t e = 0; e < MRF_TRACE_EVENTS; ++e) {
          const uint32_t a0 = h[(w * MRF_TRACE_EVENTS + e) * 2], c = h[(w * MRF_TRACE_EVENTS + e) * 2 + 1];
          if (!a0) break;
          fprintf(f, "%d %d %zu %u %u %u %u\n", a.C, p.m, w, a0 >> 24, (a0 >> 8) & 0xFFFFu, a0 & 0xFFu, c);
        }
      fclose(f);
    }
  }
}

}  // namespace piper
