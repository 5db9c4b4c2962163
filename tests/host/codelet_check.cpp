// Host-side check of the DFT codelets against a naive double-precision DFT.
// Built and run by tests/test_codelets_host.py (g++, no GPU needed).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "../../speechbrain_b200/csrc/fft_codelets.cuh"
using namespace sbfe;

static double frand() { return (double)rand() / RAND_MAX * 2.0 - 1.0; }

int main() {
  const double PI = std::acos(-1.0);
  double worst = 0;
  srand(7);
  for (int trial = 0; trial < 50; ++trial) {
    // complex 20-point
    cpx x[20];
    double xr[20], xi[20];
    for (int i = 0; i < 20; ++i) { xr[i] = frand(); xi[i] = frand(); x[i] = cpx{(float)xr[i], (float)xi[i]}; xr[i] = x[i].re; xi[i] = x[i].im; }
    dft20(x);
    for (int k = 0; k < 20; ++k) {
      double sr = 0, si = 0;
      for (int n = 0; n < 20; ++n) {
        double a = -2 * PI * k * n / 20.0;
        sr += xr[n] * std::cos(a) - xi[n] * std::sin(a);
        si += xr[n] * std::sin(a) + xi[n] * std::cos(a);
      }
      worst = std::fmax(worst, std::fabs(sr - x[k].re));
      worst = std::fmax(worst, std::fabs(si - x[k].im));
    }
    // real 20-point
    float r[20];
    cpx R[11];
    for (int i = 0; i < 20; ++i) r[i] = (float)frand();
    rdft20(r, R);
    for (int k = 0; k <= 10; ++k) {
      double sr = 0, si = 0;
      for (int n = 0; n < 20; ++n) {
        double a = -2 * PI * k * n / 20.0;
        sr += r[n] * std::cos(a);
        si += r[n] * std::sin(a);
      }
      worst = std::fmax(worst, std::fabs(sr - R[k].re));
      worst = std::fmax(worst, std::fabs(si - R[k].im));
    }
  }
  std::printf("max_abs_err %.3e\n", worst);
  return worst < 2e-5 ? 0 : 1;
}
