// Packed fp32 pairs for sm_100a: add/sub/mul/fma.rn.f32x2 (SASS FADD2 / FMUL2 / FFMA2) issue ONE instruction for two
// independent fp32 operations.  The FMA pipe still spends two cycles on it (tools/micro/ffma2.cu: same FLOP rate as scalar
// FFMA), so this buys issue slots, not flops: the per-pixel Gauss-Newton arithmetic is bound by instruction issue
// (profiles/r1_hessian_v4_ncu_full.txt), and running two records per thread in the two halves of every register pair
// halves its instruction count.  .lo = record A, .hi = record B.
#pragma once

namespace m3r {

struct V2 {
  unsigned long long v;
};

__device__ __forceinline__ V2 pk(float a, float b) {
  V2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ V2 bc(float a) { return pk(a, a); }
__device__ __forceinline__ float lo(V2 a) { return __uint_as_float((unsigned)(a.v & 0xffffffffull)); }
__device__ __forceinline__ float hi(V2 a) { return __uint_as_float((unsigned)(a.v >> 32)); }

__device__ __forceinline__ V2 fma2(V2 a, V2 b, V2 c) {
  V2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
// c += a * b in place: the same virtual register is read and written, so loop-carried accumulators need no copies
__device__ __forceinline__ void fma2_acc(V2& c, V2 a, V2 b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c.v) : "l"(a.v), "l"(b.v));
}
__device__ __forceinline__ V2 mul2(V2 a, V2 b) {
  V2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ V2 add2(V2 a, V2 b) {
  V2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ V2 sub2(V2 a, V2 b) {  // FADD2 with a negated operand pair
  V2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}

// single MUFU instructions (no denormal pre-scaling, no Newton step): 1-2 ulp, inputs here are never denormal
__device__ __forceinline__ float rsqrt_approx(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ V2 rsqrt2(V2 a) { return pk(rsqrt_approx(lo(a)), rsqrt_approx(hi(a))); }
__device__ __forceinline__ V2 rcp2(V2 a) { return pk(rcp_approx(lo(a)), rcp_approx(hi(a))); }
__device__ __forceinline__ V2 rcp_abs2(V2 a) { return pk(rcp_approx(fabsf(lo(a))), rcp_approx(fabsf(hi(a)))); }
__device__ __forceinline__ V2 min2(V2 a, float m) { return pk(fminf(lo(a), m), fminf(hi(a), m)); }
__device__ __forceinline__ V2 max2(V2 a, float m) { return pk(fmaxf(lo(a), m), fmaxf(hi(a), m)); }

}  // namespace m3r
