// softmax_math.cuh — packed fp32x2 arithmetic and the two exp2 flavours used by the prefill softmax warps.
#pragma once

#include "common.cuh"

namespace mvb200 {
namespace pf {

// ---- packed fp32x2 arithmetic (Blackwell FFMA2/FADD2: two lanes per issue slot) ----
MVB_DEV uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
MVB_DEV void unpack2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
MVB_DEV uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
MVB_DEV uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
MVB_DEV float ex2_mufu(float x) {  // MUFU.EX2, flush-to-zero: no denormal fix-up code around it
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 2^x for a pair on the FMA pipe (relieves the MUFU unit, which otherwise ties with the tensor pipe):
// round-to-nearest split x = n + f, f in [-0.5, 0.5], degree-3 minimax polynomial for 2^f (max rel.
// error 7.5e-5, far inside bf16's 2^-9), exponent add on the integer pipe.  x is clamped to >= -126.
MVB_DEV void ex2_poly2(uint64_t x, float& y0, float& y1) {
  float x0, x1;
  unpack2(x, x0, x1);
  x = pack2(fmaxf(x0, -126.f), fmaxf(x1, -126.f));
  const uint64_t magic = pack2(12582912.f, 12582912.f);          // 1.5 * 2^23
  const uint64_t t = add2(x, magic);                              // integer part lands in the low mantissa bits
  const uint64_t n = add2(t, pack2(-12582912.f, -12582912.f));
  const uint64_t f = fma2(n, pack2(-1.f, -1.f), x);               // x - n
  uint64_t p = fma2(pack2(0.05517141f, 0.05517141f), f, pack2(0.24261075f, 0.24261075f));
  p = fma2(p, f, pack2(0.69326097f, 0.69326097f));
  p = fma2(p, f, pack2(0.99992812f, 0.99992812f));
  float p0, p1, t0, t1;
  unpack2(p, p0, p1);
  unpack2(t, t0, t1);
  y0 = __int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23));
  y1 = __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23));
}

}  // namespace pf
}  // namespace mvb200
