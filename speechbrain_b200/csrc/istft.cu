// ISTFT.forward (speechbrain/processing/features.py:283-324) = torch.istft for one-sided spectra:
// per-frame inverse real DFT (1/N), window, overlap-add, division by the overlap-added squared
// window, centre trim, cut / zero-pad to the requested length.  Next row N4 of SURVEY.md 8f.
//   frames               spec [B, T, F, 2] -> windowed frames [B, T, N] (workspace): the Stockham
//                        FFT kernel of generic.cu in inverse mode when n_fft / 2 = 2^a 5^b (400,
//                        512, ...), else istft_frames_kernel: direct inverse DFT from a
//                        shared-memory twiddle table (any even n_fft)
//   istft_ola_kernel     gather formulation of the overlap-add: every output sample sums the
//                        <= ceil(N / hop) frames that cover it and divides by the window
//                        envelope computed the same way (no atomics, deterministic)
#include <cmath>

#include "common.cuh"

namespace sbfe {

bool irfft_frames_supported(int n_fft);   // generic.cu: the Stockham kernel in inverse mode
int launch_irfft_frames(const float*, int64_t, int, const float*, float*, cudaStream_t);

constexpr int kIstftFrames = 4;
constexpr int kIstftThreads = 256;

// smem: cs[N] (cos, -sin)(2 pi j / N) + kIstftFrames spectra of F complex bins
__global__ void __launch_bounds__(kIstftThreads)
istft_frames_kernel(const float2* __restrict__ spec, int64_t n_frames, int n_fft,
                    const float* __restrict__ window, float* __restrict__ frames) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ __align__(16) float smem[];
  float2* s_cs = reinterpret_cast<float2*>(smem);
  float2* s_x = s_cs + n_fft;
  const int bins = n_fft / 2 + 1;
  const int64_t f0 = (int64_t)blockIdx.x * kIstftFrames;
  const int64_t left = n_frames - f0;
  const int nf = left < kIstftFrames ? (int)left : kIstftFrames;
  for (int j = threadIdx.x; j < n_fft; j += kIstftThreads) {
    float sn, cs;
    sincospif(2.0f * (float)j / (float)n_fft, &sn, &cs);
    s_cs[j] = make_float2(cs, -sn);
  }
  for (int i = threadIdx.x; i < nf * bins; i += kIstftThreads) s_x[i] = __ldg(spec + f0 * bins + i);
  __syncthreads();
  const float inv_n = 1.0f / (float)n_fft;
  const int half = n_fft / 2;
  for (int n = threadIdx.x; n < n_fft; n += kIstftThreads) {
    const float w = __ldg(window + n) * inv_n;
    const float sign = (n & 1) ? -1.f : 1.f;
    for (int f = 0; f < nf; ++f) {
      const float2* x = s_x + f * bins;
      // the imaginary parts of bins 0 and N/2 do not exist for a real signal (C2R ignores them)
      float acc = 0.f;
      int idx = n;
      for (int k = 1; k < half; ++k) {
        const float2 cs = s_cs[idx];
        acc = fmaf(x[k].x, cs.x, acc);
        acc = fmaf(x[k].y, cs.y, acc);   // -Im * sin: cs.y = -sin
        idx += n;
        if (idx >= n_fft) idx -= n_fft;
      }
      const float y = fmaf(2.f, acc, x[0].x + sign * x[half].x);
      frames[(f0 + f) * n_fft + n] = y * w;
    }
  }
}

// IDX = int32_t when every index of one utterance fits 31 bits (the usual case: the 64-bit
// divisions otherwise cost more than the memory traffic), int64_t for the rest.
template <typename IDX>
__global__ void __launch_bounds__(256)
istft_ola_kernel(const float* __restrict__ frames, int64_t T64, int n_fft, int hop,
                 const float* __restrict__ window, int64_t start64, int64_t out_len64,
                 float* __restrict__ out) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.y;
  const IDX T = (IDX)T64, start = (IDX)start64, out_len = (IDX)out_len64;
  const IDX full = (IDX)n_fft + (IDX)hop * (T - 1);
  const float* fb = frames + b * T64 * n_fft;
  float* ob = out + b * out_len64;
  for (IDX s = (IDX)(blockIdx.x * blockDim.x + threadIdx.x); s < out_len;
       s += (IDX)(gridDim.x * blockDim.x)) {
    const IDX u = s + start;
    float v = 0.f;
    if (u < full) {
      IDX t_hi = u / hop;
      if (t_hi > T - 1) t_hi = T - 1;
      // first frame that still covers u: ceil((u - n_fft + 1) / hop), 0 when that is negative
      const IDX t_lo = u >= (IDX)n_fft ? (u - (IDX)n_fft + (IDX)hop) / (IDX)hop : 0;
      float acc = 0.f, env = 0.f;
      IDX n = u - t_lo * (IDX)hop;                 // position inside frame t_lo, then n - hop ...
      const float* fr = fb + (int64_t)t_lo * n_fft;
      for (IDX t = t_lo; t <= t_hi; ++t, n -= (IDX)hop, fr += n_fft) {
        const float w = __ldg(window + n);
        acc += fr[n];
        env = fmaf(w, w, env);
      }
      v = acc / env;
    }
    ob[s] = v;
  }
}

int64_t istft_workspace_bytes(int64_t B, int64_t T, int n_fft) {
  return B * T * (int64_t)n_fft * (int64_t)sizeof(float);
}

int launch_istft(const float* spec, int64_t B, int64_t T, int n_fft, int hop, const float* window,
                 int center, int64_t out_len, float* out, void* ws, int64_t ws_bytes,
                 cudaStream_t stream) {
  if (B == 0 || out_len == 0) return 0;
  if (n_fft > 8192) return SBFE_E_TOOBIG;
  if (ws_bytes < istft_workspace_bytes(B, T, n_fft)) return SBFE_E_NOSPACE;
  const int64_t n_frames = B * T;
  const int64_t ctas = (n_frames + kIstftFrames - 1) / kIstftFrames;
  if (ctas > 0x7fffffff || B > 65535) return SBFE_E_TOOBIG;
  const size_t smem = ((size_t)2 * n_fft + (size_t)2 * kIstftFrames * (n_fft / 2 + 1)) * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(istft_frames_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  float* frames = reinterpret_cast<float*>(ws);
  if (n_frames > 0 && irfft_frames_supported(n_fft)) {   // n_fft / 2 = 2^a 5^b: the FFT
    int rc = launch_irfft_frames(spec, n_frames, n_fft, window, frames, stream);
    if (rc) return rc;
  } else if (n_frames > 0) {
    int rc = launch_pdl(istft_frames_kernel, dim3((unsigned)ctas), dim3(kIstftThreads), smem, stream,
                        reinterpret_cast<const float2*>(spec), n_frames, n_fft, window, frames);
    if (rc) return rc;
  }
  int64_t gx = (out_len + 255) / 256;
  if (gx > 148 * 8) gx = 148 * 8;
  const int64_t start = center ? n_fft / 2 : 0;
  const bool small = (int64_t)n_fft + (int64_t)hop * T + out_len + start < 0x7fffffffLL;
  if (small)
    return launch_pdl(istft_ola_kernel<int32_t>, dim3((unsigned)gx, (unsigned)B), dim3(256), 0, stream,
                      (const float*)frames, T, n_fft, hop, window, start, out_len, out);
  return launch_pdl(istft_ola_kernel<int64_t>, dim3((unsigned)gx, (unsigned)B), dim3(256), 0, stream,
                    (const float*)frames, T, n_fft, hop, window, start, out_len, out);
}

}  // namespace sbfe
