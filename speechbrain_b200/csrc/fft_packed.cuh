// Codelets of the v3 fused frame+window+rFFT(400) kernel, written once over a
// value type T:
//   T = float      host reference build (tests/host/frame400_v3_check.cpp, g++)
//   T = sbfe::f2   device build: TWO FRAMES per thread in one 64-bit register
//                  pair, every add/mul/fma is one packed FADD2/FMUL2/FFMA2
//                  (sm_100a f32x2 math): half the issue slots of scalar code.
//
// n_fft = 400 = 20 x 20 Cooley-Tukey.  Stage 1 (role n2): windowed real 20-point
// DFT over n1 of x[20*n1 + n2], times W400^(n2*k1).  Stage 2 (role k1): 20-point
// DFT over n2, split by one radix-2 DIF step into two independent 10-point
// halves (even / odd k2) so that stage 2 has 20 equal work units like stage 1.
// All 20/10/5-point transforms are Good-Thomas prime-factor (no inner twiddles).
//
// Replaces the cuFFT / pocketfft call behind torch.stft in the reference
// (speechbrain/processing/features.py:161-172).
#pragma once

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define SBFE_HD __host__ __device__ __forceinline__
#define SBFE_D __device__ __forceinline__
#else
#define SBFE_HD inline
#endif

namespace sbfe {

// ---- scalar instantiation ---------------------------------------------------
SBFE_HD float fmas(float a, float s, float c) { return a * s + c; }   // a*s + c, s scalar
SBFE_HD float fmat(float a, float b, float c) { return a * b + c; }   // a*b + c
SBFE_HD float muls(float a, float s) { return a * s; }

#if defined(__CUDACC__)
// ---- packed pair of floats (two frames) ---------------------------------------
struct f2 {
  float2 v;
};
SBFE_D f2 mk2(float a, float b) { return f2{make_float2(a, b)}; }
SBFE_D f2 operator+(f2 a, f2 b) { return f2{__fadd2_rn(a.v, b.v)}; }
SBFE_D f2 operator-(f2 a) { return f2{make_float2(-a.v.x, -a.v.y)}; }  // folds into a neg modifier
SBFE_D f2 operator-(f2 a, f2 b) { return f2{__fadd2_rn(a.v, make_float2(-b.v.x, -b.v.y))}; }
SBFE_D f2 muls(f2 a, float s) { return f2{__fmul2_rn(a.v, make_float2(s, s))}; }
SBFE_D f2 fmas(f2 a, float s, f2 c) { return f2{__ffma2_rn(a.v, make_float2(s, s), c.v)}; }
SBFE_D f2 fmat(f2 a, f2 b, f2 c) { return f2{__ffma2_rn(a.v, b.v, c.v)}; }
#endif

template <class T>
struct cx {
  T re, im;
};

// cos/sin(2*pi/5), cos/sin(4*pi/5)
#define SBFE_PC1 0.30901699437494745f
#define SBFE_PC2 (-0.80901699437494745f)
#define SBFE_PS1 0.95105651629515353f
#define SBFE_PS2 0.58778525229247314f

// y * (tr + i*ti), scalar twiddle: 4 ops
template <class T>
SBFE_HD cx<T> cmuls(cx<T> y, float tr, float ti) {
  const T p = muls(y.im, ti);
  const T q = muls(y.im, tr);
  return cx<T>{fmas(y.re, tr, -p), fmas(y.re, ti, q)};
}

// ---- radix-5, complex, in place: 36 ops -------------------------------------
template <class T>
SBFE_HD void dft5(cx<T>& a0, cx<T>& a1, cx<T>& a2, cx<T>& a3, cx<T>& a4) {
  const T t1r = a1.re + a4.re, t1i = a1.im + a4.im;
  const T t2r = a2.re + a3.re, t2i = a2.im + a3.im;
  const T t3r = a1.re - a4.re, t3i = a1.im - a4.im;
  const T t4r = a2.re - a3.re, t4i = a2.im - a3.im;
  const T m1r = fmas(t2r, SBFE_PC2, fmas(t1r, SBFE_PC1, a0.re));
  const T m1i = fmas(t2i, SBFE_PC2, fmas(t1i, SBFE_PC1, a0.im));
  const T m2r = fmas(t2r, SBFE_PC1, fmas(t1r, SBFE_PC2, a0.re));
  const T m2i = fmas(t2i, SBFE_PC1, fmas(t1i, SBFE_PC2, a0.im));
  const T n1r = fmas(t4r, SBFE_PS2, muls(t3r, SBFE_PS1));
  const T n1i = fmas(t4i, SBFE_PS2, muls(t3i, SBFE_PS1));
  const T n2r = fmas(t4r, -SBFE_PS1, muls(t3r, SBFE_PS2));
  const T n2i = fmas(t4i, -SBFE_PS1, muls(t3i, SBFE_PS2));
  a0 = cx<T>{a0.re + t1r + t2r, a0.im + t1i + t2i};
  a1 = cx<T>{m1r + n1i, m1i - n1r};  // m1 - i*n1
  a4 = cx<T>{m1r - n1i, m1i + n1r};
  a2 = cx<T>{m2r + n2i, m2i - n2r};
  a3 = cx<T>{m2r - n2i, m2i + n2r};
}

// ---- radix-5, real input: X0 (real), X1, X2 (X4 = conj X1, X3 = conj X2): 14 ops
template <class T>
SBFE_HD void rdft5(T a0, T a1, T a2, T a3, T a4, T& x0, cx<T>& x1, cx<T>& x2) {
  const T t1 = a1 + a4, t2 = a2 + a3, t3 = a1 - a4, t4 = a2 - a3;
  x0 = a0 + t1 + t2;
  x1 = cx<T>{fmas(t2, SBFE_PC2, fmas(t1, SBFE_PC1, a0)), fmas(t4, -SBFE_PS2, muls(t3, -SBFE_PS1))};
  x2 = cx<T>{fmas(t2, SBFE_PC1, fmas(t1, SBFE_PC2, a0)), fmas(t4, SBFE_PS1, muls(t3, -SBFE_PS2))};
}

// ---- complex 10-point DFT (prime factor 2 x 5), natural order in and out: 92 ops
// n = (5a + 2b) mod 10, k = (5a' + 6b') mod 10.
template <class T>
SBFE_HD void dft10(const cx<T> (&c)[10], cx<T> (&X)[10]) {
  cx<T> u0[5], u1[5];
#define SBFE_R2(b, i0, i1)                                   \
  u0[b] = cx<T>{c[i0].re + c[i1].re, c[i0].im + c[i1].im};   \
  u1[b] = cx<T>{c[i0].re - c[i1].re, c[i0].im - c[i1].im};
  SBFE_R2(0, 0, 5) SBFE_R2(1, 2, 7) SBFE_R2(2, 4, 9) SBFE_R2(3, 6, 1) SBFE_R2(4, 8, 3)
#undef SBFE_R2
  dft5(u0[0], u0[1], u0[2], u0[3], u0[4]);
  dft5(u1[0], u1[1], u1[2], u1[3], u1[4]);
  X[0] = u0[0]; X[6] = u0[1]; X[2] = u0[2]; X[8] = u0[3]; X[4] = u0[4];
  X[5] = u1[0]; X[1] = u1[1]; X[7] = u1[2]; X[3] = u1[3]; X[9] = u1[4];
}

// ---- real-input 20-point DFT (prime factor 4 x 5): X[0..10] ----------------------
// Column c uses x[(4c + 5a) mod 20], a = 0..3.  WINDOWED: the inputs are raw
// samples and w[] the window taps, folded into the first butterflies with FMAs
// (a0 +- a2 = x0*w0 +- x2*w2 costs 3 ops instead of 4).
template <bool WINDOWED, class T>
SBFE_HD void rdft20(const T (&x)[20], const float (&w)[20], cx<T> (&X)[11]) {
  T e0[5], e2[5];
  cx<T> o1[5];
#define SBFE_RR4(c)                                                                      \
  {                                                                                      \
    constexpr int i0 = (4 * (c)) % 20, i1 = (4 * (c) + 5) % 20, i2 = (4 * (c) + 10) % 20, \
                  i3 = (4 * (c) + 15) % 20;                                              \
    T s02, d02, s13, d31;                                                                \
    if (WINDOWED) {                                                                      \
      const T p0 = muls(x[i0], w[i0]), p1 = muls(x[i1], w[i1]);                          \
      s02 = fmas(x[i2], w[i2], p0);                                                      \
      d02 = fmas(x[i2], -w[i2], p0);                                                     \
      s13 = fmas(x[i3], w[i3], p1);                                                      \
      d31 = fmas(x[i3], w[i3], -p1);                                                     \
    } else {                                                                             \
      s02 = x[i0] + x[i2];                                                               \
      d02 = x[i0] - x[i2];                                                               \
      s13 = x[i1] + x[i3];                                                               \
      d31 = x[i3] - x[i1];                                                               \
    }                                                                                    \
    e0[c] = s02 + s13;                                                                   \
    e2[c] = s02 - s13;                                                                   \
    o1[c] = cx<T>{d02, d31};                                                             \
  }
  SBFE_RR4(0) SBFE_RR4(1) SBFE_RR4(2) SBFE_RR4(3) SBFE_RR4(4)
#undef SBFE_RR4
  T r0;
  cx<T> c1, c2;
  // k1' = 0: bins 16*b' mod 20 = 0, 16, 12 -> X[4] = conj X[16], X[8] = conj X[12]
  rdft5(e0[0], e0[1], e0[2], e0[3], e0[4], r0, c1, c2);
  X[0] = cx<T>{r0, r0};  // X[0], X[10] are real: .im is a dummy, callers use .re only
  X[4] = cx<T>{c1.re, -c1.im};
  X[8] = cx<T>{c2.re, -c2.im};
  // k1' = 2: bins (10 + 16*b') mod 20 = 10, 6, 2
  rdft5(e2[0], e2[1], e2[2], e2[3], e2[4], r0, c1, c2);
  X[10] = cx<T>{r0, r0};
  X[6] = c1;
  X[2] = c2;
  // k1' = 1: bins (5 + 16*b') mod 20 = 5, 1, 17, 13, 9
  dft5(o1[0], o1[1], o1[2], o1[3], o1[4]);
  X[5] = o1[0];
  X[1] = o1[1];
  X[3] = cx<T>{o1[2].re, -o1[2].im};
  X[7] = cx<T>{o1[3].re, -o1[3].im};
  X[9] = o1[4];
}

// ---- stage 2, k1 in 1..9, half h: one radix-2 DIF step then a 10-point DFT -------
// c[n] = (a[n] + sigma*a[n+10]) * t[n];  sigma = +1, t = 1 (h = 0: even k2 = 2j)
//                                         sigma = -1, t = W20^n (h = 1: odd k2 = 2j+1)
// sigma / t are per-lane data so that both halves of a warp run the same code.
template <class T>
SBFE_HD void stage2_half(const cx<T> (&a)[20], float sigma, const float (&tr)[10],
                         const float (&ti)[10], cx<T> (&C)[10]) {
  cx<T> c[10];
#pragma unroll
  for (int n = 0; n < 10; ++n) {
    const cx<T> u{fmas(a[n + 10].re, sigma, a[n].re), fmas(a[n + 10].im, sigma, a[n].im)};
    c[n] = cmuls(u, tr[n], ti[n]);
  }
  dft10(c, C);
}

// ---- stage 2, k1 = 10: y real, X[10 + 20*k2] = sum_n y[n] W40^(n(2*k2+1)) ------------
// even k2 = 2j: DFT10 of c[n] = W40^n (y[n] - i*y[n+10]); odd k2 = 19 - 2j by conjugate
// symmetry, so C[j] carries bin 10+40j (j < 5) or its mirror 390-40j (j >= 5).
template <class T>
SBFE_HD void stage2_k10(const T (&y)[20], cx<T> (&C)[10]) {
  constexpr float cr[10] = {1.0f, 0.98768834059513777f, 0.95105651629515357f, 0.89100652418836787f,
                            0.80901699437494742f, 0.70710678118654752f, 0.58778525229247313f,
                            0.45399049973954675f, 0.30901699437494742f, 0.15643446504023087f};
  constexpr float sn[10] = {0.0f, 0.15643446504023087f, 0.30901699437494742f, 0.45399049973954675f,
                            0.58778525229247313f, 0.70710678118654752f, 0.80901699437494742f,
                            0.89100652418836787f, 0.95105651629515357f, 0.98768834059513777f};
  cx<T> c[10];
#pragma unroll
  for (int n = 0; n < 10; ++n) {
    c[n].re = fmas(y[n + 10], -sn[n], muls(y[n], cr[n]));
    c[n].im = fmas(y[n + 10], -cr[n], muls(y[n], -sn[n]));
  }
  dft10(c, C);
}


// ---- radix-4 butterfly, forward (W4 = -i), in place: y_q = sum_m a_m W4^(m q) ---------------------
template <class T>
SBFE_HD void radix4(cx<T>& a0, cx<T>& a1, cx<T>& a2, cx<T>& a3) {
  const T s02r = a0.re + a2.re, s02i = a0.im + a2.im, d02r = a0.re - a2.re, d02i = a0.im - a2.im;
  const T s13r = a1.re + a3.re, s13i = a1.im + a3.im, d13r = a1.re - a3.re, d13i = a1.im - a3.im;
  a0 = cx<T>{s02r + s13r, s02i + s13i};
  a2 = cx<T>{s02r - s13r, s02i - s13i};
  a1 = cx<T>{d02r + d13i, d02i - d13r};   // (a0 - a2) - i (a1 - a3)
  a3 = cx<T>{d02r - d13i, d02i + d13r};   // (a0 - a2) + i (a1 - a3)
}

// ---- complex 16-point DFT (4 x 4 Cooley-Tukey), natural order in and out: X[k] = out[k] --------
// n = j + 4 m, k = q + 4 r:  W16^(nk) = W16^(jq) W4^(jr) W4^(mq).  144 adds + 8 twiddle products.
template <class T>
SBFE_HD void dft16(const cx<T> (&in)[16], cx<T> (&X)[16]) {
  constexpr float c1 = 0.92387953251128674f, s1 = 0.38268343236508977f, h = 0.70710678118654752f;
  cx<T> a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = in[i];
#pragma unroll
  for (int j = 0; j < 4; ++j) radix4(a[j], a[j + 4], a[j + 8], a[j + 12]);   // a[j + 4 q] = b[j][q]
  a[5] = cmuls(a[5], c1, -s1);                       // W16^1
  a[9] = cmuls(a[9], h, -h);                         // W16^2
  a[13] = cmuls(a[13], s1, -c1);                     // W16^3
  a[6] = cmuls(a[6], h, -h);                         // W16^2
  a[10] = cx<T>{a[10].im, -a[10].re};                // W16^4 = -i
  a[14] = cmuls(a[14], -h, -h);                      // W16^6
  a[7] = cmuls(a[7], s1, -c1);                       // W16^3
  a[11] = cmuls(a[11], -h, -h);                      // W16^6
  a[15] = cmuls(a[15], -c1, s1);                     // W16^9
#pragma unroll
  for (int q = 0; q < 4; ++q) radix4(a[4 * q], a[4 * q + 1], a[4 * q + 2], a[4 * q + 3]);
#pragma unroll
  for (int k = 0; k < 16; ++k) X[k] = a[4 * (k & 3) + (k >> 2)];             // a[4 q + r] = X[q + 4 r]
}

// bin (0..200) that output j of a stage-2 unit lands in (mirrored above 200)
SBFE_HD int stage2_bin(int k1, int h, int j) {
  const int k = k1 + 20 * (2 * j + h);
  return k <= 200 ? k : 400 - k;
}

}  // namespace sbfe
