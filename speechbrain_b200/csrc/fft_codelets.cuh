// Register-resident DFT codelets for the fused frame+window+rFFT(400) kernel.
//
// n_fft = 400 = 20 x 20 (Cooley-Tukey with one twiddle stage); each 20-point
// transform is a 4 x 5 Good-Thomas prime-factor transform (no inner twiddles).
// All codelets are straight-line code on scalars so that every value lives in a
// register; they compile for host too (tests/test_codelets_host.py builds a g++
// harness that checks them against a naive DFT without needing a GPU).
//
// Replaces the cuFFT/pocketfft call behind torch.stft in the reference
// (speechbrain/processing/features.py:161-172).
#pragma once

#if defined(__CUDACC__)
#define SBFE_HD __host__ __device__ __forceinline__
#else
#define SBFE_HD inline
#endif

namespace sbfe {

struct cpx {
  float re, im;
};

SBFE_HD cpx cmul(cpx a, cpx b) {
  return cpx{a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re};
}

// cos/sin(2*pi/5), cos/sin(4*pi/5)
#define SBFE_C1 0.30901699437494745f
#define SBFE_C2 (-0.80901699437494745f)
#define SBFE_S1 0.95105651629515353f
#define SBFE_S2 0.58778525229247314f

// ---- radix-4, forward (W4 = -i), in place ---------------------------------
SBFE_HD void dft4(cpx& a0, cpx& a1, cpx& a2, cpx& a3) {
  const float t0r = a0.re + a2.re, t0i = a0.im + a2.im;
  const float t1r = a0.re - a2.re, t1i = a0.im - a2.im;
  const float t2r = a1.re + a3.re, t2i = a1.im + a3.im;
  const float t3r = a1.re - a3.re, t3i = a1.im - a3.im;
  a0 = cpx{t0r + t2r, t0i + t2i};
  a2 = cpx{t0r - t2r, t0i - t2i};
  a1 = cpx{t1r + t3i, t1i - t3r};  // t1 - i*t3
  a3 = cpx{t1r - t3i, t1i + t3r};  // t1 + i*t3
}

// ---- radix-5, forward, in place --------------------------------------------
SBFE_HD void dft5(cpx& a0, cpx& a1, cpx& a2, cpx& a3, cpx& a4) {
  const float t1r = a1.re + a4.re, t1i = a1.im + a4.im;
  const float t2r = a2.re + a3.re, t2i = a2.im + a3.im;
  const float t3r = a1.re - a4.re, t3i = a1.im - a4.im;
  const float t4r = a2.re - a3.re, t4i = a2.im - a3.im;
  const float m1r = a0.re + SBFE_C1 * t1r + SBFE_C2 * t2r;
  const float m1i = a0.im + SBFE_C1 * t1i + SBFE_C2 * t2i;
  const float m2r = a0.re + SBFE_C2 * t1r + SBFE_C1 * t2r;
  const float m2i = a0.im + SBFE_C2 * t1i + SBFE_C1 * t2i;
  const float n1r = SBFE_S1 * t3r + SBFE_S2 * t4r, n1i = SBFE_S1 * t3i + SBFE_S2 * t4i;
  const float n2r = SBFE_S2 * t3r - SBFE_S1 * t4r, n2i = SBFE_S2 * t3i - SBFE_S1 * t4i;
  a0 = cpx{a0.re + t1r + t2r, a0.im + t1i + t2i};
  a1 = cpx{m1r + n1i, m1i - n1r};  // m1 - i*n1
  a4 = cpx{m1r - n1i, m1i + n1r};
  a2 = cpx{m2r + n2i, m2i - n2r};
  a3 = cpx{m2r - n2i, m2i + n2r};
}

// real-input radix-5: returns X0 (real) and X1, X2 (X4 = conj X1, X3 = conj X2)
SBFE_HD void rdft5(float a0, float a1, float a2, float a3, float a4, float& x0, cpx& x1,
                   cpx& x2) {
  const float t1 = a1 + a4, t2 = a2 + a3, t3 = a1 - a4, t4 = a2 - a3;
  x0 = a0 + t1 + t2;
  x1 = cpx{a0 + SBFE_C1 * t1 + SBFE_C2 * t2, -(SBFE_S1 * t3 + SBFE_S2 * t4)};
  x2 = cpx{a0 + SBFE_C2 * t1 + SBFE_C1 * t2, -(SBFE_S2 * t3 - SBFE_S1 * t4)};
}

// ---- complex 20-point DFT (prime-factor 4 x 5) -----------------------------
// in : x[n], natural order.   out: X[k], natural order (in place).
// Input map n = (5*n1 + 4*n2) mod 20, output map k = (5*k1 + 16*k2) mod 20.
SBFE_HD void dft20(cpx (&x)[20]) {
  // 5 radix-4 transforms over n1 (for each n2): elements (4*n2 + 5*n1) mod 20
#define SBFE_R4(n2)                                                                  \
  dft4(x[(4 * (n2)) % 20], x[(4 * (n2) + 5) % 20], x[(4 * (n2) + 10) % 20],         \
       x[(4 * (n2) + 15) % 20])
  SBFE_R4(0); SBFE_R4(1); SBFE_R4(2); SBFE_R4(3); SBFE_R4(4);
#undef SBFE_R4
  // now slot (4*n2 + 5*k1) mod 20 holds the k1-th output of column n2.
  // 4 radix-5 transforms over n2 (for each k1)
#define SBFE_R5(k1)                                                                  \
  dft5(x[(5 * (k1)) % 20], x[(5 * (k1) + 4) % 20], x[(5 * (k1) + 8) % 20],          \
       x[(5 * (k1) + 12) % 20], x[(5 * (k1) + 16) % 20])
  SBFE_R5(0); SBFE_R5(1); SBFE_R5(2); SBFE_R5(3);
#undef SBFE_R5
  // slot (5*k1 + 4*k2) mod 20 holds X[(5*k1 + 16*k2) mod 20]; un-permute.
  // k2 -> slot offset 4*k2, bin offset 16*k2: slot s = 5*k1 + 4*k2 maps to bin
  // 5*k1 + 16*k2 = s + 12*k2 (mod 20).
  cpx y[20];
#define SBFE_MV(k1, k2) y[(5 * (k1) + 16 * (k2)) % 20] = x[(5 * (k1) + 4 * (k2)) % 20]
#define SBFE_MV5(k1) SBFE_MV(k1, 0); SBFE_MV(k1, 1); SBFE_MV(k1, 2); SBFE_MV(k1, 3); SBFE_MV(k1, 4)
  SBFE_MV5(0); SBFE_MV5(1); SBFE_MV5(2); SBFE_MV5(3);
#undef SBFE_MV5
#undef SBFE_MV
#pragma unroll
  for (int i = 0; i < 20; ++i) x[i] = y[i];
}

// ---- real-input 20-point DFT -----------------------------------------------
// in : x[0..19] real.  out: X[0..10] (X[0], X[10] have zero imaginary part).
SBFE_HD void rdft20(const float (&x)[20], cpx (&X)[11]) {
  // radix-4 over n1 on real data, column n2 uses x[(4*n2 + 5*n1) mod 20]:
  //   k1=0: e0 = (a0+a2)+(a1+a3)   k1=2: e2 = (a0+a2)-(a1+a3)   (real)
  //   k1=1: (a0-a2) - i(a1-a3)     k1=3 = conj(k1=1)
  float e0[5], e2[5];
  cpx o1[5];
#define SBFE_RR4(n2)                                                                 \
  {                                                                                  \
    const float a0 = x[(4 * (n2)) % 20], a1 = x[(4 * (n2) + 5) % 20];               \
    const float a2 = x[(4 * (n2) + 10) % 20], a3 = x[(4 * (n2) + 15) % 20];         \
    const float s02 = a0 + a2, s13 = a1 + a3;                                        \
    e0[n2] = s02 + s13;                                                              \
    e2[n2] = s02 - s13;                                                              \
    o1[n2] = cpx{a0 - a2, a3 - a1};                                                  \
  }
  SBFE_RR4(0) SBFE_RR4(1) SBFE_RR4(2) SBFE_RR4(3) SBFE_RR4(4)
#undef SBFE_RR4
  // k1 = 0: bins 16*k2 mod 20 -> k2=0:0, 1:16, 2:12  => X[4]=conj(k2=1), X[8]=conj(k2=2)
  float r0;
  cpx c1, c2;
  rdft5(e0[0], e0[1], e0[2], e0[3], e0[4], r0, c1, c2);
  X[0] = cpx{r0, 0.f};
  X[4] = cpx{c1.re, -c1.im};
  X[8] = cpx{c2.re, -c2.im};
  // k1 = 2: bins (10 + 16*k2) mod 20 -> k2=0:10, 1:6, 2:2
  rdft5(e2[0], e2[1], e2[2], e2[3], e2[4], r0, c1, c2);
  X[10] = cpx{r0, 0.f};
  X[6] = c1;
  X[2] = c2;
  // k1 = 1: bins (5 + 16*k2) mod 20 -> k2=0:5, 1:1, 2:17, 3:13, 4:9
  dft5(o1[0], o1[1], o1[2], o1[3], o1[4]);
  X[5] = o1[0];
  X[1] = o1[1];
  X[3] = cpx{o1[2].re, -o1[2].im};  // conj X[17]
  X[7] = cpx{o1[3].re, -o1[3].im};  // conj X[13]
  X[9] = o1[4];
}

}  // namespace sbfe
