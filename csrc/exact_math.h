// exact_math.h — transcendental functions written only with IEEE-754 fp32/fp64
// add/mul/fma/div, so the CPU oracle and the fp32 "exact" CUDA kernel produce
// the same bits.  libm's expf/tanhf and CUDA's differ by an ulp or two, which
// is enough to flip a sample at a CDF boundary (ref:322-333); sharing one
// definition removes that source of divergence.
//
// Accuracy: b200_expf < 1 ulp over the range that does not overflow;
// b200_tanhf < 4 ulp.  The sigmoid keeps the reference's fp64 add + divide
// (ref:216: `1.0 / (1.0 + expf(-x))`).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDA_ARCH__)
#define B200_HD __host__ __device__ __forceinline__
#define B200_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define B200_MUL(a, b) __fmul_rn((a), (b))
#define B200_ADD(a, b) __fadd_rn((a), (b))
#elif defined(__CUDACC__)
#define B200_HD __host__ __device__ inline
#define B200_FMA(a, b, c) fmaf((a), (b), (c))
#define B200_MUL(a, b) ((a) * (b))
#define B200_ADD(a, b) ((a) + (b))
#else
#define B200_HD inline
#define B200_FMA(a, b, c) fmaf((a), (b), (c))
#define B200_MUL(a, b) ((a) * (b))
#define B200_ADD(a, b) ((a) + (b))
#endif

B200_HD float b200_bits_to_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// e^x.  Cody–Waite reduction x = n*ln2 + r, |r| <= ln2/2, degree-6 polynomial
// in Horner form with fma, exact scaling by 2^n in two steps.
B200_HD float b200_expf(float x) {
  if (x > 88.72283905206835f) return b200_bits_to_float(0x7f800000u);  // +inf
  if (x < -103.97208f) return 0.0f;
  if (x != x) return x;
  const float kLog2e = 1.44269504088896341f;
  const float kLn2Hi = 0.693145751953125f;        // 0x3f317200
  const float kLn2Lo = 1.42860682030941723e-6f;   // ln2 - kLn2Hi
  float t = B200_MUL(x, kLog2e);
  // round to nearest integer without relying on the rounding-mode intrinsic
  float n = floorf(B200_ADD(t, 0.5f));
  float r = B200_FMA(-n, kLn2Hi, x);
  r = B200_FMA(-n, kLn2Lo, r);
  float p = 1.0f / 720.0f;
  p = B200_FMA(p, r, 1.0f / 120.0f);
  p = B200_FMA(p, r, 1.0f / 24.0f);
  p = B200_FMA(p, r, 1.0f / 6.0f);
  p = B200_FMA(p, r, 0.5f);
  p = B200_FMA(p, r, 1.0f);
  p = B200_FMA(p, r, 1.0f);
  int ni = (int)n;
  // 2^ni as two exact factors so that results in the subnormal range and near
  // overflow are still formed by plain multiplies.
  int n1 = ni / 2, n2 = ni - n1;
  float s1 = b200_bits_to_float((uint32_t)(n1 + 127) << 23);
  float s2 = b200_bits_to_float((uint32_t)(n2 + 127) << 23);
  return B200_MUL(B200_MUL(p, s1), s2);
}

// tanh(x): odd polynomial for small |x| (avoids the 1 - 2/(e+1) cancellation),
// exp-based otherwise, saturating to +-1 above 9.02.
B200_HD float b200_tanhf(float x) {
  float ax = fabsf(x);
  if (x != x) return x;
  float y;
  if (ax < 0.35f) {
    float x2 = B200_MUL(ax, ax);
    // tanh(x) = x * (1 + x2*(c1 + x2*(c2 + ...))), Taylor series through x^13
    // (truncation < 2e-8 relative below 0.35)
    float p = 0.00359212803657248f;                  //  21844/6081075
    p = B200_FMA(p, x2, -0.00886323552990220f);      // -1382/155925
    p = B200_FMA(p, x2, 0.0218694885361552f);        //  62/2835
    p = B200_FMA(p, x2, -0.0539682539682540f);       // -17/315
    p = B200_FMA(p, x2, 0.133333333333333f);         //  2/15
    p = B200_FMA(p, x2, -0.333333333333333f);        // -1/3
    p = B200_MUL(p, x2);
    y = B200_FMA(ax, p, ax);
  } else if (ax < 9.02f) {
    float e = b200_expf(B200_MUL(2.0f, ax));
    y = B200_ADD(1.0f, -(2.0f / B200_ADD(e, 1.0f)));
  } else {
    y = 1.0f;
  }
  return x < 0.0f ? -y : y;
}

// ref:216 — double literals promote the add and the divide to fp64.
B200_HD float b200_sigmoidf(float x) {
  return (float)(1.0 / (1.0 + (double)b200_expf(-x)));
}
