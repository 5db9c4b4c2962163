// Fused frame + window + real-FFT(400) + power + sparse mel + dB kernel.
//
// Replaces, in one pass over the waveform and without materialising the complex
// STFT or the power spectrogram in HBM:
//   STFT.forward          speechbrain/processing/features.py:133-188
//   spectral_magnitude    speechbrain/processing/features.py:327-356 (power=1)
//   Filterbank.forward    speechbrain/processing/features.py:490-551
//   _amplitude_to_DB      speechbrain/processing/features.py:701-702 (log part;
//                         the per-utterance top_db clamp needs the max over the
//                         whole utterance, which this kernel only accumulates)
//
// Mapping: one CTA = 32 consecutive frames of one utterance, LANE = FRAME.
// Every warp works on the same 32 frames but on a different slice of the
// transform, so window values, twiddles and mel weights are warp-uniform
// (constant-bank / broadcast loads) and every shared-memory access is
// lane-contiguous, i.e. bank-conflict free:
//   stage 0  waveform span -> smem, pitch (hop | 1) per hop so that the
//            lane stride (= hop samples) is odd
//   stage 1  warp w: residues n2 in {w, w+10}: 20-pt real DFT over n1 of
//            x[20*n1 + n2] * win, times twiddle W400^(n2*k1) -> exchange buffer
//   stage 2  warp w: k1 = w (warp 0 also k1 = 10): 20-pt DFT over n2 ->
//            |X[k1 + 20*k2]|^2 -> power buffer P[bin][lane]
//   stage 3  warp w: filters m = w, w+10, ...: sparse dot with P, dB, max
//   stage 4  [32 x M] tile -> global, coalesced
#include <cmath>
#include <cstring>

#include "common.cuh"
#include "fft_codelets.cuh"

namespace sbfe {

constexpr int kTile = 32;        // frames per CTA (= lanes)
constexpr int kWarps = 10;
constexpr int kThreads = kWarps * 32;
constexpr int kNfft = 400;
constexpr int kBins = 201;
constexpr int kMaxSmem = (31 * 400 + 400 + 40 + 400 * 32) * 4;  // hop = 400 worst case

struct Fbank400Params {
  float window[kNfft];   // zero-padded to n_fft
  cpx tw[11][20];        // W400^(k1*n2), k1 = 0..10
  DbParams db;
  int hop;
  int n_mels;
  unsigned tiles;        // frame tiles per utterance
  int64_t L, wav_stride, T;
  const float* wav;
  float* out;
  sbfe_maxkey_t* utt_max;
  const int32_t* mel_start;
  const int32_t* mel_count;
  const int32_t* mel_offset;
  const float* mel_weight;
};

__host__ __device__ inline int region_a_floats(int hop) {
  const int pad = (hop & 1) ? 0 : 1;
  const int span = (kTile - 1) * hop + kNfft;
  const int wav_floats = span + (span / hop) * pad + 1;
  const int p_floats = kBins * kTile;
  return wav_floats > p_floats ? wav_floats : p_floats;
}

template <int HOP>
__global__ void __launch_bounds__(kThreads, 2)
fbank400_kernel(const __grid_constant__ Fbank400Params prm) {
  extern __shared__ __align__(16) float smem[];
  const int hop = HOP ? HOP : prm.hop;
  const int pad = (hop & 1) ? 0 : 1;
  const int pitch = hop + pad;
  float* s_a = smem;                          // waveform span, later P[bin][lane]
  float* s_e = smem + region_a_floats(hop);   // exchange [200 pairs][32 lanes] float2, later out tile
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t b = blockIdx.x / prm.tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * prm.tiles) * kTile;

  // ---- stage 0: stage the waveform span (zero centre padding) ---------------
  {
    const int span = (kTile - 1) * hop + kNfft;
    const int64_t g0 = t0 * hop - kNfft / 2;
    const float* w = prm.wav + b * prm.wav_stride;
    for (int i = threadIdx.x; i < span; i += kThreads) {
      const int64_t g = g0 + i;
      const float v = (g >= 0 && g < prm.L) ? __ldg(w + g) : 0.f;
      s_a[i + (i / hop) * pad] = v;
    }
  }
  __syncthreads();

  // ---- stage 1: real 20-pt DFT over n1 for residue n2, twiddle --------------
  float2* s_e2 = reinterpret_cast<float2*>(s_e);
#pragma unroll 1
  for (int n2 = warp; n2 < 20; n2 += kWarps) {
    float x[20];
    const float* src = s_a + lane * pitch;
#pragma unroll
    for (int n1 = 0; n1 < 20; ++n1) {
      const int j = 20 * n1 + n2;
      x[n1] = src[j + (j / hop) * pad] * prm.window[j];
    }
    cpx y[11];
    rdft20(x, y);
    float2* dst = s_e2 + (n2 * 10) * kTile + lane;
    dst[0] = make_float2(y[0].re, y[10].re);
#pragma unroll
    for (int k1 = 1; k1 <= 9; ++k1) {
      const cpx t = cmul(y[k1], prm.tw[k1][n2]);
      dst[k1 * kTile] = make_float2(t.re, t.im);
    }
  }
  __syncthreads();

  // ---- stage 2: 20-pt DFT over n2 for each k1, power -> P[bin][lane] --------
  float* s_p = s_a;
  if (warp == 0) {
    float x0[20];
    cpx z[20];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) {
      const float2 v = s_e2[(n2 * 10) * kTile + lane];
      x0[n2] = v.x;
      z[n2] = cpx{v.y * prm.tw[10][n2].re, v.y * prm.tw[10][n2].im};
    }
    cpx y[11];
    rdft20(x0, y);
#pragma unroll
    for (int k2 = 0; k2 <= 10; ++k2)
      s_p[(20 * k2) * kTile + lane] = y[k2].re * y[k2].re + y[k2].im * y[k2].im;
    dft20(z);
#pragma unroll
    for (int k2 = 0; k2 < 10; ++k2)
      s_p[(10 + 20 * k2) * kTile + lane] = z[k2].re * z[k2].re + z[k2].im * z[k2].im;
  } else {
    const int k1 = warp;  // 1..9
    cpx z[20];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) {
      const float2 v = s_e2[(n2 * 10 + k1) * kTile + lane];
      z[n2] = cpx{v.x, v.y};
    }
    dft20(z);
#pragma unroll
    for (int k2 = 0; k2 < 20; ++k2) {
      const int k = k1 + 20 * k2;
      const int bin = (k2 <= 9) ? k : kNfft - k;  // k <= 200 <=> k2 <= 9 for k1 in 1..9
      s_p[bin * kTile + lane] = z[k2].re * z[k2].re + z[k2].im * z[k2].im;
    }
  }
  __syncthreads();

  // ---- stage 3: sparse mel projection + dB, tile [32][M] in smem -------------
  const int M = prm.n_mels;
  const int opitch = M | 1;
  float* s_o = s_e;
  const bool frame_ok = (t0 + lane) < prm.T;
  float vmax = -INFINITY;
#pragma unroll 1
  for (int m = warp; m < M; m += kWarps) {
    const int start = __ldg(prm.mel_start + m), cnt = __ldg(prm.mel_count + m);
    const float* wt = prm.mel_weight + __ldg(prm.mel_offset + m);
    const float* p = s_p + start * kTile + lane;
    float acc = 0.f;
#pragma unroll 4
    for (int i = 0; i < cnt; ++i) acc = fmaf(__ldg(wt + i), p[i * kTile], acc);
    const float v = to_db(acc, prm.db);
    s_o[lane * opitch + m] = v;
    if (frame_ok) vmax = fmaxf(vmax, v);
  }
  vmax = warp_max(vmax);
  if (lane == 0 && vmax > -INFINITY && prm.utt_max != nullptr)
    atomicMax(prm.utt_max + b, max_key_encode(vmax));
  __syncthreads();

  // ---- stage 4: coalesced store of the valid part of the tile ----------------
  const int64_t left = prm.T - t0;
  const int nvalid = left < kTile ? (int)left : kTile;
  float* o = prm.out + (b * prm.T + t0) * M;
  const int total = nvalid * M;
  for (int idx = threadIdx.x; idx < total; idx += kThreads) {
    const int t = idx / M, m = idx - t * M;
    o[idx] = s_o[t * opitch + m];
  }
}

struct TwTable400 {
  cpx tw[11][20];
  TwTable400() {
    const double two_pi = 6.283185307179586476925286766559;
    for (int k1 = 0; k1 <= 10; ++k1)
      for (int n2 = 0; n2 < 20; ++n2) {
        const double a = -two_pi * (double)(k1 * n2) / 400.0;
        tw[k1][n2] = cpx{(float)cos(a), (float)sin(a)};
      }
  }
};

// Host launcher (called from api.cu).  desc has been validated by the caller.
int launch_fbank400(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                    int64_t wav_stride, float* out, sbfe_maxkey_t* utt_max,
                    cudaStream_t stream) {
  Fbank400Params prm;
  static const TwTable400 table;  // C++11 magic static: thread-safe, immutable
  memcpy(prm.tw, table.tw, sizeof(prm.tw));
  for (int i = 0; i < kNfft; ++i) prm.window[i] = d->h_window[i];
  prm.db = DbParams{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                    d->log_mel};
  prm.hop = d->hop;
  prm.n_mels = d->n_mels;
  prm.L = L;
  prm.wav_stride = wav_stride;
  prm.T = 1 + L / d->hop;
  prm.wav = wav;
  prm.out = out;
  prm.utt_max = utt_max;
  prm.mel_start = d->d_mel_start;
  prm.mel_count = d->d_mel_count;
  prm.mel_offset = d->d_mel_offset;
  prm.mel_weight = d->d_mel_weight;

  const int64_t tiles = (prm.T + kTile - 1) / kTile;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  prm.tiles = (unsigned)tiles;
  const size_t smem = (size_t)(region_a_floats(d->hop) + 400 * kTile) * sizeof(float);
  dim3 grid((unsigned)(tiles * B));
  // opt in to > 48 KB dynamic smem once per process (hop <= 400 bounds the size)
  static bool attr_done[64] = {};  // per device; a duplicate set is harmless
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_done[dev]) {
    cudaError_t e = cudaFuncSetAttribute(fbank400_kernel<160>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(fbank400_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               kMaxSmem);
    if (e != cudaSuccess) return (int)e;
    if (dev >= 0 && dev < 64) attr_done[dev] = true;
  }
  if (d->hop == 160) {
    fbank400_kernel<160><<<grid, kThreads, smem, stream>>>(prm);
  } else {
    fbank400_kernel<0><<<grid, kThreads, smem, stream>>>(prm);
  }
  return check_last();
}

}  // namespace sbfe
