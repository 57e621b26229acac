// Micro-benchmark (debug aid, not part of the library): back-to-back tcgen05.mma throughput on one CTA per SM
// for the operand layouts the conv kernels use.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I piper-swift_b200/csrc -I include \
//        tools/micro/umma_bench.cu -o gpurun_out/umma_bench
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include "tc_common.cuh"
using namespace piper::tc;

// a_mode 0: A swizzle-128B K-major (64 channels per row); 1: no-swizzle group-planar (LBO = plane, SBO = 128)
__global__ void __launch_bounds__(128, 1) umma_bench(long long* out, int N, int a_mode, int shift16, int kb_b, int iters,
                                                      int per_tap, int f16) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 160 * 1024 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(&slot, 256); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 0) {
    const uint32_t idesc = make_idesc(N, f16);
    const uint32_t plane_bytes = 1152 * 16;
    const uint32_t a_addr = smem_u32(smem) + 1024, b_addr = smem_u32(smem) + 128 * 1024;
    const uint64_t a_hi = a_mode ? make_smem_desc_ns(0, plane_bytes, 128) : make_smem_desc(0, 64);
    const uint64_t b_hi = make_smem_desc(0, kb_b);
    const uint32_t a_lo0 = a_addr >> 4, b_lo0 = b_addr >> 4;
    const uint32_t kstep_a = a_mode ? 2u * (plane_bytes >> 4) : 2u;
    long long t0 = clock64();
    uint32_t a_lo = a_lo0;
    for (int it = 0; it < iters; ++it) {
      if (elect_one()) {
        for (int ks = 0; ks < per_tap; ++ks)
          umma_f16(tmem, a_hi | (uint64_t)(a_lo + ks * kstep_a), b_hi | (uint64_t)(b_lo0 + 2u * ks), idesc, 1u);
      }
      a_lo += (uint32_t)shift16;
      if ((it & 7) == 7) a_lo = a_lo0;
    }
    long long t1 = clock64();
    if (elect_one()) umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}


// issue-cost probe: PT MMAs per elected block (compile-time unrolled); style 0 = elect_one per block in a
// warp-uniform loop, 1 = lane 0 predicate per block, 2 = the whole loop inside one elected thread,
// 3 = like 0 but loop-invariant descriptors
template <int PT, int STYLE>
__global__ void __launch_bounds__(128, 1) umma_issue(long long* out, int N, int iters) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 160 * 1024 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(&slot, 256); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 0) {
    const uint32_t idesc = make_idesc(N, 1);
    const uint32_t plane_bytes = 1152 * 16;
    const uint32_t a_addr = smem_u32(smem) + 1024, b_addr = smem_u32(smem) + 128 * 1024;
    const uint64_t a_hi = make_smem_desc_ns(0, plane_bytes, 128);
    const uint64_t b_hi = make_smem_desc(0, 64);
    const uint32_t a_lo0 = a_addr >> 4, b_lo0 = b_addr >> 4;
    const uint32_t kstep_a = 2u * (plane_bytes >> 4);
    long long t0 = clock64();
    uint32_t a_lo = a_lo0;
    if (STYLE == 2) {
      if (elect_one()) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
          for (int ks = 0; ks < PT; ++ks)
            umma_f16(tmem, a_hi | (uint64_t)(a_lo + (ks & 3) * kstep_a), b_hi | (uint64_t)(b_lo0 + 2u * (ks & 3)), idesc, 1u);
          a_lo += 1u;
          if ((it & 7) == 7) a_lo = a_lo0;
        }
      }
      __syncwarp();
    } else {
      for (int it = 0; it < iters; ++it) {
        const bool me = STYLE == 1 ? lane == 0 : elect_one();
        if (me) {
#pragma unroll
          for (int ks = 0; ks < PT; ++ks)
            umma_f16(tmem, a_hi | (uint64_t)(a_lo + (ks & 3) * kstep_a), b_hi | (uint64_t)(b_lo0 + 2u * (ks & 3)), idesc, 1u);
        }
        if (STYLE != 3) {
          a_lo += 1u;
          if ((it & 7) == 7) a_lo = a_lo0;
        }
      }
    }
    long long t1 = clock64();
    if (elect_one()) umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int PT, int STYLE>
void run_issue(long long* d, int N) {
  const int iters = 2048;
  cudaFuncSetAttribute(umma_issue<PT, STYLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int rep = 0; rep < 2; ++rep) {
    umma_issue<PT, STYLE><<<148, 128, 200 * 1024>>>(d, N, iters);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return; }
  }
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("issue N=%3d per_block=%2d style=%d : issue %7.1f total %7.1f cycles/mma, %7.1f cycles/block\n", N, PT, STYLE,
         h[0] / (double)(iters * PT), h[1] / (double)(iters * PT), h[1] / (double)iters);
}

// the MRF tile loop, verbatim structure: runtime k taps, ring slots, accumulate flag, commit per tile.
// VAR 0: as in conv_mrf.cu; 1: accumulate flag constant; 2: no per-tile commit; 3: b_lo by increment (no multiply)
template <int KS, int VAR>
__global__ void __launch_bounds__(128, 1) umma_tile_loop(long long* out, int N, int tiles, int k, int dil, int ring_) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar[8];
  __shared__ uint32_t slot_;
  for (int i = threadIdx.x; i < 160 * 1024 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  const int warp = VAR >= 4 ? __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) : (int)(threadIdx.x >> 5);
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bar[i], 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(&slot_, 256); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = VAR >= 4 ? __shfl_sync(0xffffffffu, slot_, 0) : slot_;
  if (warp == 0) {
    const uint32_t idesc = make_idesc(N, 1);
    const uint32_t plane_bytes = 1152 * 16;
    const uint32_t plane16 = plane_bytes >> 4;
    const uint64_t a_hi = make_smem_desc_ns(0, plane_bytes, 128);
    const uint64_t b_hi = make_smem_desc(0, KS * 16);
    const uint32_t x0_lo = ((smem_u32(smem) + 1024) >> 4) + 64, wr_lo = (smem_u32(smem) + 128 * 1024) >> 4;
    const uint32_t tap16 = (uint32_t)(KS * 16 * KS * 16 * 2) >> 4;
    const uint32_t ring = (uint32_t)ring_;
    long long t0 = clock64();
    uint32_t slot0 = 0;
    for (int tl = 0; tl < tiles; ++tl) {
      const int i = tl & 7, j = (tl >> 3) & 1;
      const uint32_t acc = tmem_base + (uint32_t)((i & (KS == 2 ? 7 : 3)) * KS * 16);
      uint32_t a_lo = x0_lo - (uint32_t)(((k - 1) / 2) * dil) + (uint32_t)(128 * i);
      uint32_t s = slot0, ph = 0;
      uint32_t b_inc = wr_lo + s * tap16;
      for (int t = 0; t < k; ++t) {
        const uint32_t b_lo = VAR == 3 ? b_inc : wr_lo + s * tap16;
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < KS; ++ks)
            umma_f16(acc, a_hi | (uint64_t)(a_lo + (uint32_t)ks * 2u * plane16), b_hi | (uint64_t)(b_lo + 2u * ks), idesc,
                     VAR == 1 ? 1u : ((j == 1 || t > 0 || ks > 0) ? 1u : 0u));
        }
        a_lo += (uint32_t)dil;
        b_inc += tap16;
        if (++s == ring) { s = 0; ph ^= 1u; b_inc = wr_lo; }
      }
      if (VAR != 2 && elect_one()) umma_commit(&bar[i]);
      if ((tl & 7) == 7) slot0 = (slot0 + 3) % ring;
      if (ph == 77) out[3] = 1;
    }
    long long t1 = clock64();
    if (elect_one()) umma_commit(&bar[0]);
    // parity of bar[0] depends on the number of commits; just wait for the last phase via a fresh barrier
    __syncwarp();
    long long spin = 0;
    while (clock64() - t1 < 200000) ++spin;
    long long t2 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = spin; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 256);
}

template <int KS, int VAR>
void run_tile(long long* d, int k) {
  const int tiles = 1024;
  cudaFuncSetAttribute(umma_tile_loop<KS, VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int rep = 0; rep < 2; ++rep) {
    umma_tile_loop<KS, VAR><<<148, 128, 200 * 1024>>>(d, KS * 16, tiles, k, 2, KS == 2 ? 30 : 7);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return; }
  }
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("tile loop C=%d k=%d var=%d : issue %7.1f cycles/mma, %7.1f cycles/tap, %7.1f cycles/tile\n", KS * 16, k, VAR,
         h[0] / (double)(tiles * k * KS), h[0] / (double)(tiles * k), h[0] / (double)tiles);
}

int main() {
  { long long* d0; cudaMalloc(&d0, 32);
    if (getenv("TILE_ONLY")) {
      for (int k : {3, 5, 7}) { run_tile<2, 0>(d0, k); run_tile<2, 4>(d0, k); run_tile<4, 0>(d0, k); run_tile<4, 4>(d0, k); }
      return 0;
    }
    for (int N : {32, 64}) {
      run_issue<1, 0>(d0, N); run_issue<2, 0>(d0, N); run_issue<4, 0>(d0, N); run_issue<8, 0>(d0, N); run_issue<16, 0>(d0, N);
      run_issue<1, 1>(d0, N); run_issue<2, 1>(d0, N); run_issue<4, 1>(d0, N); run_issue<16, 1>(d0, N);
      run_issue<1, 2>(d0, N); run_issue<2, 2>(d0, N); run_issue<4, 2>(d0, N); run_issue<16, 2>(d0, N);
      run_issue<2, 3>(d0, N); run_issue<4, 3>(d0, N);
    }
    if (getenv("ISSUE_ONLY")) return 0; }
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(umma_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 4096;
  printf("N a_mode shift16 kb_b per_tap f16 : issue_cycles/mma total_cycles/mma  (math cycles at 4096 FMA/clk)\n");
  int Ns[] = {32, 64, 96, 128, 192, 256};
  for (int f16 = 1; f16 >= 0; --f16)
  for (int N : Ns)
    for (int a_mode = 0; a_mode < 2; ++a_mode)
      for (int shift : {0, 8, 1, 3})
        for (int per_tap : {2, 4}) {
          if (a_mode == 0 && (shift == 1 || shift == 3)) continue;   // swizzled tiles only shift by whole atoms here
          if (!f16 && (N != 32 || per_tap != 2)) continue;
          const int kb_b = per_tap == 2 ? 32 : 64;
          const int sh = a_mode == 0 ? shift * 8 : shift;   // swizzled rows are 128 bytes = 8 units
          for (int rep = 0; rep < 2; ++rep) {
            umma_bench<<<148, 128, 200 * 1024>>>(d, N, a_mode, sh, kb_b, iters, per_tap, f16);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          }
          long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
          const double n = (double)iters * per_tap;
          printf("%3d %d %d %2d %d %d : %7.1f %7.1f  (%.0f)\n", N, a_mode, shift, kb_b, per_tap, f16, h[0] / n, h[1] / n,
                 128.0 * N * 16 / 4096);
        }
  return 0;
}
