// umma_pacing.cu — measures what the prefill kernel's MMA mix can sustain on this GPU.
// Every CTA (1 or 2 per SM, grid = 148 * ctas_per_sm) has ONE thread issuing tcgen05.mma back to back
// over random bf16 operands resident in shared memory / TMEM; optional extra warps run an ex2 loop
// (the softmax's MUFU load).  Reported: SM cycles per MMA (clock64 inside the kernel) and chip
// TFLOP/s over the launch (CUDA events), so the pacing law AND the clock the chip holds under this
// load are both visible.
//   kind 0  SS  M128 N64   K16   (Q K^T, 64-key tile)
//   kind 1  SS  M128 N128  K16   (Q K^T, 128-key tile)
//   kind 2  SS  M128 N256  K16
//   kind 3  TS  M128 N128  K16   (P V, A in TMEM, B MN-major)
//   kind 4  mix 8 x SS N64 + 4 x TS N128      (one 128x64 attention tile)
//   kind 5  mix 8 x SS N128 + 8 x TS N128     (one 128x128 attention tile)
//   kind 6  ex2 only (no MMA): MUFU throughput per SM
//   kind 7  tcgen05.ld only: TMEM read throughput per SM
// Build: make -C probes umma_pacing      Run (B200): probes/umma_pacing
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../minivllm_b200/csrc/tcgen05.cuh"

using namespace mvb200;

constexpr int kThreads = 160;  // warp 0 issues, warps 1..4 optionally run ex2

__global__ void __launch_bounds__(kThreads)
pacing_kernel(int kind, int tiles, int tmem_cols, int ex2_warps, unsigned long long* cycles, float* sink) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t sA = base;                 // 32 KB: [2 halves][128 rows][128 B]
  const uint32_t sB = base + 32768;         // 64 KB: up to [2 halves][256 rows][128 B]
  const uint32_t bar = sB + 65536, slot = bar + 16;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // random bf16 in [-2, 2): realistic toggle rates for the power/clock behaviour
  uint32_t* w = reinterpret_cast<uint32_t*>(gen);
  uint32_t x = 0x9E3779B9u * (blockIdx.x * kThreads + threadIdx.x + 1);
  for (int i = threadIdx.x; i < (32768 + 65536) / 4; i += kThreads) {
    x = x * 1664525u + 1013904223u;
    const uint32_t lo = 0x3F00u | ((x >> 8) & 0x80FFu), hi = 0x3F00u | ((x >> 16) & 0x80FFu);
    w[i] = lo | (hi << 16);
  }
  fence_proxy_async_smem();
  if (warp == 0) {
    tmem_alloc(slot, tmem_cols);
    tmem_relinquish();
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (slot - base));
  if (warp >= 1) {  // seed the TMEM columns the TS variants read as A
    const uint32_t la = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
    uint32_t r[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = w[(threadIdx.x * 32 + i) & 8191];
    for (int c = 0; c < 128; c += 32) tmem_st32(la + c, r);
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (kind == 7) {
    // TMEM read throughput: warps 1..4 (one per SM sub-partition) stream tcgen05.ld.32x32b.x32
    if (warp >= 1) {
      const uint32_t la = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
      uint32_t r[32], acc = 0;
      const long long t0 = clock64();
      for (int t = 0; t < tiles; ++t) {
#pragma unroll
        for (int c = 0; c < 128; c += 32) {
          tmem_ld32(la + c, r);
          tmem_wait_ld();
          acc ^= r[0] ^ r[31];
        }
      }
      if (acc == 0x12345u) sink[0] = 1.f;
      if (threadIdx.x == 32) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
    }
  } else if (threadIdx.x == 0 && kind != 6) {
    const uint32_t colS = 0, colO = tmem_cols >= 512 ? 256 : 128;
    const long long t0 = clock64();
    auto qk = [&](int n, uint32_t dcol) {
      const uint32_t idesc = umma_idesc(128, n, true, false, false);
      const uint32_t bhalf = n * 128;
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        const uint64_t a = umma_smem_desc(sA + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024);
        const uint64_t b = umma_smem_desc(sB + (kk >> 2) * bhalf + (kk & 3) * 32, 16, 1024);
        umma_ss(tmem + dcol, a, b, idesc, kk > 0);
      }
    };
    auto pv = [&](int keys, uint32_t pcol) {
      const uint32_t idesc = umma_idesc(128, 128, true, false, true);
      for (int kk = 0; kk < keys / 16; ++kk) {
        const uint64_t b = umma_smem_desc(sB + kk * 16 * 128, keys * 128, 1024);
        umma_ts(tmem + colO, tmem + pcol + kk * 8, b, idesc, kk > 0);
      }
    };
    for (int t = 0; t < tiles; ++t) {
      switch (kind) {
        case 0: qk(64, colS + (t & 1) * 64); break;
        case 1: qk(128, colS); break;
        case 2: qk(256, colS); break;
        case 3: pv(128, colS); break;
        case 4: qk(64, colS + (t & 1) * 64); pv(64, colS + ((t + 1) & 1) * 64); break;
        default: qk(128, colS + (tmem_cols >= 512 ? (t & 1) * 128 : 0)); pv(128, colS + (tmem_cols >= 512 ? ((t + 1) & 1) * 128 : 0)); break;
      }
    }
    umma_commit(bar);
    mbar_wait(bar, 0);
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  } else if (warp >= 1 && warp <= ex2_warps) {
    // MUFU load: 8 independent chains per thread, `tiles` * 64 ex2 per thread
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = -0.001f * (lane + i + 1);
    const long long t0 = clock64();
    for (int t = 0; t < tiles * 8; ++t) {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = v[i] - 1.0f;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += v[i];
    if (s == 12345.f) sink[0] = s;
    if (kind == 6 && threadIdx.x == 32) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, tmem_cols);
}

int main(int argc, char** argv) {
  const int ms_target = argc > 1 ? atoi(argv[1]) : 20;
  unsigned long long* dcyc;
  float* dsink;
  cudaMalloc(&dcyc, 296 * 8);
  cudaMalloc(&dsink, 4);
  const int smem = 32768 + 65536 + 64 + 1024;
  cudaFuncSetAttribute(pacing_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  struct Cfg { int kind, ctas, ex2; const char* name; double mma_per_tile, flop_per_tile; };
  const double f64 = 2.0 * 128 * 64 * 16, f128 = 2.0 * 128 * 128 * 16, f256 = 2.0 * 128 * 256 * 16;
  const Cfg cfgs[] = {
      {0, 1, 0, "SS N64            1 CTA/SM", 8, 8 * f64},   {0, 2, 0, "SS N64            2 CTA/SM", 8, 8 * f64},
      {1, 1, 0, "SS N128           1 CTA/SM", 8, 8 * f128},  {1, 2, 0, "SS N128           2 CTA/SM", 8, 8 * f128},
      {2, 1, 0, "SS N256           1 CTA/SM", 8, 8 * f256},  {3, 1, 0, "TS N128           1 CTA/SM", 8, 8 * f128},
      {3, 2, 0, "TS N128           2 CTA/SM", 8, 8 * f128},  {4, 1, 0, "mix 128x64 tile   1 CTA/SM", 12, 8 * f64 + 4 * f128},
      {4, 2, 0, "mix 128x64 tile   2 CTA/SM", 12, 8 * f64 + 4 * f128},
      {5, 1, 0, "mix 128x128 tile  1 CTA/SM", 16, 16 * f128}, {5, 2, 0, "mix 128x128 tile  2 CTA/SM", 16, 16 * f128},
      {4, 2, 4, "mix 128x64 + ex2  2 CTA/SM", 12, 8 * f64 + 4 * f128},
      {5, 1, 4, "mix 128x128 + ex2 1 CTA/SM", 16, 16 * f128},
      {7, 1, 4, "LDTM.x32 4 warps  1 CTA/SM", 0, 0},       {7, 2, 4, "LDTM.x32 4 warps  2 CTA/SM", 0, 0},
      {6, 1, 4, "ex2 only 4 warps  1 CTA/SM", 0, 0},         {6, 2, 4, "ex2 only 4 warps  2 CTA/SM", 0, 0},
  };
  printf("%-30s %10s %12s %12s %10s\n", "config", "cyc/MMA", "flop/cyc/SM", "TFLOP/s", "eff MHz");
  for (const Cfg& c : cfgs) {
    const int grid = 148 * c.ctas;
    const int cols = c.ctas == 1 ? 512 : 256;
    int tiles = 2000;
    float ms = 0;
    std::vector<unsigned long long> cyc(grid);
    for (int pass = 0; pass < 3; ++pass) {  // pass 0 calibrates, 1 warms, 2 is reported
      cudaEventRecord(e0);
      pacing_kernel<<<grid, kThreads, smem>>>(c.kind, tiles, cols, c.ex2, dcyc, dsink);
      cudaEventRecord(e1);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: CUDA error %s\n", c.name, cudaGetErrorString(e)); return 2; }
      cudaEventElapsedTime(&ms, e0, e1);
      if (pass == 0) tiles = (int)(tiles * (double)ms_target / ms) + 1;
    }
    cudaMemcpy(cyc.data(), dcyc, grid * 8, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (auto v : cyc) avg += (double)v;
    avg /= grid;
    if (c.kind == 7) {
      printf("%-30s %10.1f cycles per LDTM.x32 (4 KB per warp) => %.0f B/cyc/SM  (%d iters, %.2f ms)\n", c.name,
             avg / (tiles * 4.0), 4.0 * 4096 * c.ctas / (avg / (tiles * 4.0)), tiles, ms);
    } else if (c.kind == 6) {
      const double ex2 = (double)tiles * 8 * 8 * 128 * c.ctas;  // per SM
      printf("%-30s %10s %12.2f %12s %10.0f   (ex2/cyc/SM; %d tiles, %.2f ms)\n", c.name, "-", ex2 / avg, "-",
             avg / (ms * 1e3), tiles, ms);
    } else {
      const double flops = c.flop_per_tile * tiles * grid;
      printf("%-30s %10.1f %12.0f %12.1f %10.0f   (%d tiles, %.2f ms)\n", c.name, avg / (tiles * c.mma_per_tile) * 1.0,
             c.flop_per_tile * tiles * c.ctas / avg, flops / (ms * 1e-3) / 1e12, avg / (ms * 1e3), tiles, ms);
    }
  }
  return 0;
}
