// pipe_rates.cu — reciprocal issue throughput (cycles per warp instruction, per SM sub-partition) of the
// instructions the prefill softmax is made of, with 1, 2 and 4 warps per sub-partition (sm_100a).
// Decides which forms the softmax should use (profiles/r02_pipe_rates.txt).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define OPS 8      // independent chains per thread
#define ITERS 512

template <int kOp>
__device__ __forceinline__ void body(float (&a)[2 * OPS], float s, float m) {
#pragma unroll
  for (int i = 0; i < OPS; ++i) {
    float& x = a[2 * i];
    float& y = a[2 * i + 1];
    if (kOp == 0) {  // FFMA2 reg pair * broadcast scalar + broadcast scalar
      asm volatile("{.reg .b64 t, u, v; mov.b64 t, {%0,%1}; mov.b64 u, {%2,%2}; mov.b64 v, {%3,%3};"
                   " fma.rn.f32x2 t, t, u, v; mov.b64 {%0,%1}, t;}" : "+f"(x), "+f"(y) : "f"(s), "f"(m));
    } else if (kOp == 1) {  // two scalar FFMA (register operands)
      asm volatile("fma.rn.f32 %0, %0, %2, %3; fma.rn.f32 %1, %1, %2, %3;" : "+f"(x), "+f"(y) : "f"(s), "f"(m));
    } else if (kOp == 2) {  // two scalar FFMA, immediate multiplier
      asm volatile("fma.rn.f32 %0, %0, 0f3F800001, %2; fma.rn.f32 %1, %1, 0f3F800001, %2;" : "+f"(x), "+f"(y) : "f"(m));
    } else if (kOp == 3) {  // FADD2
      asm volatile("{.reg .b64 t, u; mov.b64 t, {%0,%1}; mov.b64 u, {%2,%2}; add.rn.f32x2 t, t, u; mov.b64 {%0,%1}, t;}"
                   : "+f"(x), "+f"(y) : "f"(m));
    } else if (kOp == 4) {  // two scalar FADD
      asm volatile("add.rn.f32 %0, %0, %2; add.rn.f32 %1, %1, %2;" : "+f"(x), "+f"(y) : "f"(m));
    } else if (kOp == 5) {  // two MUFU.EX2
      asm volatile("ex2.approx.ftz.f32 %0, %0; ex2.approx.ftz.f32 %1, %1;" : "+f"(x), "+f"(y));
    } else if (kOp == 6) {  // F2FP pack to bf16x2 (+ the move back so that the chain continues)
      uint32_t r;
      asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x), "f"(y));
      x = __uint_as_float(r);
    } else if (kOp == 7) {  // 3-input max
      asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(y), "f"(s));
    } else if (kOp == 8) {  // ex2 on a packed bf16 pair (one MUFU for two values?)
      uint32_t r = __float_as_uint(x);
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r));
      x = __uint_as_float(r);
    } else if (kOp == 9) {  // integer shift-add (the exponent insert of the polynomial exp2)
      uint32_t r = __float_as_uint(x), q = __float_as_uint(y);
      asm volatile("mad.lo.u32 %0, %1, 8388608, %0;" : "+r"(r) : "r"(q));
      x = __uint_as_float(r);
    } else if (kOp == 11) {  // the softmax inner step of one score pair: FFMA2, 2 x MUFU.EX2, FADD2, F2FP
      uint32_t r;
      asm volatile("{.reg .b64 t, u, v, w; .reg .f32 a, b; mov.b64 t, {%0,%1}; mov.b64 u, {%3,%3}; mov.b64 v, {%4,%4};"
                   " fma.rn.f32x2 t, t, u, v; mov.b64 {a,b}, t; ex2.approx.ftz.f32 a, a; ex2.approx.ftz.f32 b, b;"
                   " mov.b64 w, {a,b}; add.rn.f32x2 t, t, w; cvt.rn.bf16x2.f32 %2, b, a; mov.b64 {%0,%1}, t;}"
                   : "+f"(x), "+f"(y), "=r"(r) : "f"(s), "f"(m));
      y += __uint_as_float(r & 0x7f800000u) * 0.f;
    } else if (kOp == 12) {  // same without the exponentials (FMA + ALU pipes only)
      uint32_t r;
      asm volatile("{.reg .b64 t, u, v; mov.b64 t, {%0,%1}; mov.b64 u, {%3,%3}; mov.b64 v, {%4,%4};"
                   " fma.rn.f32x2 t, t, u, v; add.rn.f32x2 t, t, v; cvt.rn.bf16x2.f32 %2, %1, %0; mov.b64 {%0,%1}, t;}"
                   : "+f"(x), "+f"(y), "=r"(r) : "f"(s), "f"(m));
      y += __uint_as_float(r & 0x7f800000u) * 0.f;
    } else if (kOp == 10) {  // FMUL2
      asm volatile("{.reg .b64 t, u; mov.b64 t, {%0,%1}; mov.b64 u, {%2,%2}; mul.rn.f32x2 t, t, u; mov.b64 {%0,%1}, t;}"
                   : "+f"(x), "+f"(y) : "f"(s));
    }
  }
}

template <int kOp>
__global__ void probe(float* out, long long* cycles, float s, float m) {
  float a[2 * OPS];
#pragma unroll
  for (int i = 0; i < 2 * OPS; ++i) a[i] = threadIdx.x * 1e-3f + i;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) body<kOp>(a, s, m);
  const long long t1 = clock64();
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 2 * OPS; ++i) acc += a[i];
  out[threadIdx.x] = acc;
  if ((threadIdx.x & 31) == 0) cycles[threadIdx.x >> 5] = t1 - t0;
}

template <int kOp>
void run(const char* name, int instr_per_chain, float* out, long long* cyc) {
  printf("%-44s", name);
  for (int warps_per_smsp : {1, 2, 4}) {
    const int threads = 128 * warps_per_smsp;
    probe<kOp><<<1, threads>>>(out, cyc, 1.0000001f, 1e-9f);
    cudaDeviceSynchronize();
    long long h[16];
    cudaMemcpy(h, cyc, sizeof(long long) * (threads / 32), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < threads / 32; ++i) mx = h[i] > mx ? h[i] : mx;
    // warp instructions issued per sub-partition
    const double instr = (double)ITERS * OPS * instr_per_chain * warps_per_smsp;
    printf("  w=%d: %6.2f clk/instr", warps_per_smsp, mx / instr);
  }
  printf("\n");
}

int main() {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 4096 * 4);
  cudaMalloc(&cyc, 64 * 8);
  printf("cycles per warp instruction per SM sub-partition (one CTA on one SM)\n");
  run<0>("FFMA2  (pair * bcast + bcast)", 1, out, cyc);
  run<1>("FFMA   (reg, reg, reg) x2", 2, out, cyc);
  run<2>("FFMA   (reg, imm, reg) x2", 2, out, cyc);
  run<3>("FADD2", 1, out, cyc);
  run<4>("FADD x2", 2, out, cyc);
  run<10>("FMUL2", 1, out, cyc);
  run<5>("MUFU.EX2 x2", 2, out, cyc);
  run<8>("MUFU.EX2.bf16x2", 1, out, cyc);
  run<6>("F2FP.BF16.PACK_AB", 1, out, cyc);
  run<7>("FMNMX3", 1, out, cyc);
  run<9>("IMAD (x * 2^23 + y)", 1, out, cyc);
  printf("mixes: cycles per score PAIR (FFMA2 + 2 MUFU + FADD2 + F2FP [+1 FFMA]); alone these cost 3 x 2.1 + 2 x 8\n");
  run<11>("softmax pair step (all pipes)", 1, out, cyc);
  run<12>("softmax pair step without MUFU", 1, out, cyc);
  cudaError_t e = cudaGetLastError();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
