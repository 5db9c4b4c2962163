// Host emulation of the v3 kernel's arithmetic for one frame (T = float): every
// stage-1 role, the exchange, every stage-2 unit and the bin mapping, against a
// naive double-precision windowed DFT power spectrum.  Built by
// tests/test_library_cpu.py with g++ (no GPU needed).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../speechbrain_b200/csrc/fft_packed.cuh"
using namespace sbfe;

int main() {
  const double PI = std::acos(-1.0);
  srand(11);
  double worst = 0;
  for (int trial = 0; trial < 20; ++trial) {
    float x[400], win[400];
    for (int i = 0; i < 400; ++i) {
      x[i] = (float)((double)rand() / RAND_MAX * 2 - 1);
      win[i] = (float)(0.54 - 0.46 * std::cos(2 * PI * i / 400.0));
    }
    // exchange E[n2][slot]: slot 0 = (y0, y10) reals, slots 1..9 complex (twiddled)
    float Ere[20][10], Eim[20][10];
    for (int n2 = 0; n2 < 20; ++n2) {
      float xs[20], ws[20];
      for (int n1 = 0; n1 < 20; ++n1) { xs[n1] = x[20 * n1 + n2]; ws[n1] = win[20 * n1 + n2]; }
      cx<float> Y[11];
      rdft20<true>(xs, ws, Y);
      Ere[n2][0] = Y[0].re; Eim[n2][0] = Y[10].re;
      for (int k1 = 1; k1 <= 9; ++k1) {
        const double a = -2 * PI * (k1 * n2) / 400.0;
        const cx<float> t = cmuls(Y[k1], (float)std::cos(a), (float)std::sin(a));
        Ere[n2][k1] = t.re; Eim[n2][k1] = t.im;
      }
    }
    std::vector<double> P(201, -1.0);
    // k1 = 0
    {
      float y[20], dummy[20] = {};
      for (int n2 = 0; n2 < 20; ++n2) y[n2] = Ere[n2][0];
      cx<float> X[11];
      rdft20<false>(y, dummy, X);
      for (int k2 = 0; k2 <= 10; ++k2) {
        const double im = (k2 == 0 || k2 == 10) ? 0.0 : X[k2].im;
        P[20 * k2] = (double)X[k2].re * X[k2].re + im * im;
      }
    }
    // k1 = 10
    {
      float y[20];
      for (int n2 = 0; n2 < 20; ++n2) y[n2] = Eim[n2][0];
      cx<float> C[10];
      stage2_k10(y, C);
      for (int j = 0; j < 10; ++j) {
        const int bin = j < 5 ? 10 + 40 * j : 390 - 40 * j;
        if (P[bin] >= 0) { std::printf("bin %d written twice\n", bin); return 2; }
        P[bin] = (double)C[j].re * C[j].re + (double)C[j].im * C[j].im;
      }
    }
    for (int k1 = 1; k1 <= 9; ++k1)
      for (int h = 0; h < 2; ++h) {
        cx<float> a[20], C[10];
        for (int n2 = 0; n2 < 20; ++n2) a[n2] = cx<float>{Ere[n2][k1], Eim[n2][k1]};
        float tr[10], ti[10];
        for (int n = 0; n < 10; ++n) {
          tr[n] = h ? (float)std::cos(2 * PI * n / 20.0) : 1.f;
          ti[n] = h ? (float)-std::sin(2 * PI * n / 20.0) : 0.f;
        }
        stage2_half(a, h ? -1.f : 1.f, tr, ti, C);
        for (int j = 0; j < 10; ++j) {
          const int bin = stage2_bin(k1, h, j);
          if (bin < 0 || bin > 200 || P[bin] >= 0) { std::printf("bad bin %d\n", bin); return 2; }
          P[bin] = (double)C[j].re * C[j].re + (double)C[j].im * C[j].im;
        }
      }
    for (int k = 0; k <= 200; ++k) {
      double sr = 0, si = 0;
      for (int n = 0; n < 400; ++n) {
        const double a = -2 * PI * k * n / 400.0, v = (double)x[n] * win[n];
        sr += v * std::cos(a);
        si += v * std::sin(a);
      }
      const double ref = sr * sr + si * si;
      if (P[k] < 0) { std::printf("bin %d never written\n", k); return 2; }
      worst = std::fmax(worst, std::fabs(P[k] - ref) / (1.0 + ref));
    }
  }
  std::printf("max_rel_err %.3e\n", worst);
  return worst < 2e-5 ? 0 : 1;
}
