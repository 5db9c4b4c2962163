// Host emulation of the two-stage 20x20 real-FFT dataflow used by the fused
// kernel (stage 1: real 20-pt DFT per residue n2 + twiddle; stage 2: 20-pt DFT
// per k1), checked against a naive double DFT of the windowed frame.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "../../speechbrain_b200/csrc/fft_codelets.cuh"
using namespace sbfe;

int main() {
  const double PI = std::acos(-1.0);
  srand(3);
  float s[400], w[400];
  for (int i = 0; i < 400; ++i) {
    s[i] = (float)((double)rand() / RAND_MAX * 2 - 1);
    w[i] = (float)(0.54 - 0.46 * std::cos(2 * PI * i / 400.0));
  }
  // twiddle table tw[k1][n2] = exp(-2*pi*i*k1*n2/400), k1 = 0..10
  cpx tw[11][20];
  for (int k1 = 0; k1 <= 10; ++k1)
    for (int n2 = 0; n2 < 20; ++n2) {
      double a = -2 * PI * k1 * n2 / 400.0;
      tw[k1][n2] = cpx{(float)std::cos(a), (float)std::sin(a)};
    }
  // exchange buffer: E[n2][0] = Y0 (real), E[n2][1] = Y10 (real, untwiddled),
  // E[n2][2*k1], E[n2][2*k1+1] = twiddled Y[k1], k1 = 1..9
  float E[20][20];
  for (int n2 = 0; n2 < 20; ++n2) {
    float x[20];
    for (int n1 = 0; n1 < 20; ++n1) x[n1] = s[20 * n1 + n2] * w[20 * n1 + n2];
    cpx Y[11];
    rdft20(x, Y);
    E[n2][0] = Y[0].re;
    E[n2][1] = Y[10].re;
    for (int k1 = 1; k1 <= 9; ++k1) {
      cpx t = cmul(Y[k1], tw[k1][n2]);
      E[n2][2 * k1] = t.re;
      E[n2][2 * k1 + 1] = t.im;
    }
  }
  float P[201];
  for (int i = 0; i < 201; ++i) P[i] = -1.f;
  // k1 = 0: real input
  {
    float x[20];
    for (int n2 = 0; n2 < 20; ++n2) x[n2] = E[n2][0];
    cpx X[11];
    rdft20(x, X);
    for (int k2 = 0; k2 <= 10; ++k2) P[20 * k2] = X[k2].re * X[k2].re + X[k2].im * X[k2].im;
  }
  // k1 = 10: real input times twiddle, outputs k2 = 0..9
  {
    cpx z[20];
    for (int n2 = 0; n2 < 20; ++n2) z[n2] = cpx{E[n2][1] * tw[10][n2].re, E[n2][1] * tw[10][n2].im};
    dft20(z);
    for (int k2 = 0; k2 < 10; ++k2) P[10 + 20 * k2] = z[k2].re * z[k2].re + z[k2].im * z[k2].im;
  }
  for (int k1 = 1; k1 <= 9; ++k1) {
    cpx z[20];
    for (int n2 = 0; n2 < 20; ++n2) z[n2] = cpx{E[n2][2 * k1], E[n2][2 * k1 + 1]};
    dft20(z);
    for (int k2 = 0; k2 < 20; ++k2) {
      int k = k1 + 20 * k2;
      int bin = k <= 200 ? k : 400 - k;
      P[bin] = z[k2].re * z[k2].re + z[k2].im * z[k2].im;
    }
  }
  double worst = 0, scale = 0;
  for (int k = 0; k <= 200; ++k) {
    double sr = 0, si = 0;
    for (int n = 0; n < 400; ++n) {
      double a = -2 * PI * k * n / 400.0, v = (double)s[n] * (double)w[n];
      sr += v * std::cos(a);
      si += v * std::sin(a);
    }
    double p = sr * sr + si * si;
    scale = std::fmax(scale, p);
    worst = std::fmax(worst, std::fabs(p - P[k]));
  }
  std::printf("max_abs_err %.3e (max power %.3e)\n", worst, scale);
  return worst < 1e-5 * scale ? 0 : 1;
}
