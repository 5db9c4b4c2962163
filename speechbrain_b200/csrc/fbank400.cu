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
//   stage 0  waveform span -> smem (128-bit loads, all in flight before the
//            stores), pitch (hop | 1) per hop so that the lane stride is odd;
//            mel tables -> smem
//   stage 1  warp w: residues n2 in {w, w+10}: 20-pt real DFT over n1 of
//            x[20*n1 + n2] * win, times twiddle W400^(n2*k1) -> exchange buffer
//   stage 2  warp w: k1 = w (warp 0 also k1 = 10): 20-pt DFT over n2 ->
//            |X[k1 + 20*k2]|^2 -> power buffer P[bin][lane]
//   stage 3  warp w: filters m = w, w+10, ...: sparse dot with P, dB, max
//   stage 4  [32 x M] tile -> global, 128-bit coalesced stores
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
constexpr int kLdIters = 5;      // 128-bit loads in flight per thread in stage 0

struct Fbank400Params {
  float window[kNfft];   // zero-padded to n_fft
  cpx tw[11][20];        // W400^(k1*n2), k1 = 0..10
  DbParams db;
  int hop;
  int n_mels;
  int mel_nnz;
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
  const int wav_floats = span + (span / hop) * pad + 4;
  const int p_floats = kBins * kTile;
  return ((wav_floats > p_floats ? wav_floats : p_floats) + 3) & ~3;
}
// dynamic smem: region A | exchange E (400 x 32) | mel meta int2[M] | mel weights
__host__ __device__ inline size_t fbank400_smem_bytes(int hop, int n_mels, int nnz) {
  return (size_t)(region_a_floats(hop) + 400 * kTile + 2 * n_mels + nnz) * sizeof(float);
}

template <int HOP>
__global__ void __launch_bounds__(kThreads, 2)
fbank400_kernel(const __grid_constant__ Fbank400Params prm) {
  extern __shared__ __align__(16) float smem[];
  const int hop = HOP ? HOP : prm.hop;
  const int pad = (hop & 1) ? 0 : 1;
  const int pitch = hop + pad;
  const int M = prm.n_mels;
  float* s_a = smem;                          // waveform span, later P[bin][lane]
  float* s_e = smem + region_a_floats(hop);   // exchange [200 pairs][32 lanes] float2, later out tile
  int2* s_meta = reinterpret_cast<int2*>(s_e + 400 * kTile);  // {start | count << 16, offset}
  float* s_w = reinterpret_cast<float*>(s_meta + M);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t b = blockIdx.x / prm.tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * prm.tiles) * kTile;

  // ---- stage 0: stage the waveform span (zero centre padding) ---------------
  {
    const int span = (kTile - 1) * hop + kNfft;
    const int64_t g0 = t0 * hop - kNfft / 2;
    // valid (inside the utterance) part of the span: i in [lo, hi)
    const int lo = g0 < 0 ? (int)(-g0 < span ? -g0 : span) : 0;
    const int64_t rem = prm.L - g0;
    const int hi = rem <= 0 ? 0 : (rem < span ? (int)rem : span);
    const float* w = prm.wav + b * prm.wav_stride + g0;  // only dereferenced inside [lo, hi)
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(w) & 15) == 0) && ((hop & 3) == 0);
    for (int base = 0; base < span; base += kThreads * 4 * kLdIters) {
      float4 v[kLdIters];
#pragma unroll
      for (int k = 0; k < kLdIters; ++k) {
        const int i = base + (k * kThreads + threadIdx.x) * 4;
        if (vec_ok && i >= lo && i + 3 < hi) {
          v[k] = __ldg(reinterpret_cast<const float4*>(w + i));
        } else {
          v[k].x = (i >= lo && i < hi) ? __ldg(w + i) : 0.f;
          v[k].y = (i + 1 >= lo && i + 1 < hi) ? __ldg(w + i + 1) : 0.f;
          v[k].z = (i + 2 >= lo && i + 2 < hi) ? __ldg(w + i + 2) : 0.f;
          v[k].w = (i + 3 >= lo && i + 3 < hi) ? __ldg(w + i + 3) : 0.f;
        }
      }
#pragma unroll
      for (int k = 0; k < kLdIters; ++k) {
        const int i = base + (k * kThreads + threadIdx.x) * 4;
        if (i < span) {  // span is a multiple of 4 whenever hop is; else guard each element
          if ((hop & 3) == 0) {
            float* d = s_a + i + (i / hop) * pad;
            d[0] = v[k].x; d[1] = v[k].y; d[2] = v[k].z; d[3] = v[k].w;
          } else {
            s_a[i + (i / hop) * pad] = v[k].x;
            if (i + 1 < span) s_a[i + 1 + ((i + 1) / hop) * pad] = v[k].y;
            if (i + 2 < span) s_a[i + 2 + ((i + 2) / hop) * pad] = v[k].z;
            if (i + 3 < span) s_a[i + 3 + ((i + 3) / hop) * pad] = v[k].w;
          }
        }
      }
    }
    for (int m = threadIdx.x; m < M; m += kThreads)
      s_meta[m] = make_int2(__ldg(prm.mel_start + m) | (__ldg(prm.mel_count + m) << 16),
                            __ldg(prm.mel_offset + m));
    for (int i = threadIdx.x; i < prm.mel_nnz; i += kThreads) s_w[i] = __ldg(prm.mel_weight + i);
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
  const int opitch = M | 1;
  float* s_o = s_e;
  const bool frame_ok = (t0 + lane) < prm.T;
  float vmax = -INFINITY;
#pragma unroll 1
  for (int m = warp; m < M; m += kWarps) {
    const int2 meta = s_meta[m];
    const int cnt = meta.x >> 16;
    const float* wt = s_w + meta.y;
    const float* p = s_p + (meta.x & 0xffff) * kTile + lane;
    float acc = 0.f;
#pragma unroll 2
    for (int i = 0; i < cnt; ++i) acc = fmaf(wt[i], p[i * kTile], acc);
    const float v = to_db(acc, prm.db);
    s_o[lane * opitch + m] = v;
    vmax = fmaxf(vmax, v);
  }
  if (!frame_ok) vmax = -INFINITY;
  vmax = warp_max(vmax);
  if (lane == 0 && vmax > -INFINITY && prm.utt_max != nullptr)
    atomicMax(prm.utt_max + b, max_key_encode(vmax));
  __syncthreads();

  // ---- stage 4: coalesced store of the valid part of the tile ----------------
  const int64_t left = prm.T - t0;
  const int nvalid = left < kTile ? (int)left : kTile;
  float* o = prm.out + (b * prm.T + t0) * M;
  const int total = nvalid * M;
  if ((M & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
    // thread -> 4 consecutive outputs; (t, m) advanced without divisions
    int e = threadIdx.x * 4;
    int t = e / M, m = e - t * M;
    const int dt = (kThreads * 4) / M, dm = (kThreads * 4) - dt * M;
    for (; e < total; e += kThreads * 4) {
      const float* s = s_o + t * opitch + m;
      *reinterpret_cast<float4*>(o + e) = make_float4(s[0], s[1], s[2], s[3]);
      t += dt;
      m += dm;
      if (m >= M) { m -= M; ++t; }
    }
  } else {
    int e = threadIdx.x;
    int t = e / M, m = e - t * M;
    const int dt = kThreads / M, dm = kThreads - dt * M;
    for (; e < total; e += kThreads) {
      o[e] = s_o[t * opitch + m];
      t += dt;
      m += dm;
      if (m >= M) { m -= M; ++t; }
    }
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

constexpr int kMaxSmem = 200 * 1024;

bool fbank400_supported(const sbfe_fbank_desc* d) {
  return d->n_fft == 400 && d->hop <= 400 && d->n_mels <= 384 && d->mel_nnz >= 0 &&
         fbank400_smem_bytes(d->hop, d->n_mels, d->mel_nnz) <= 100 * 1024;
}

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
  prm.mel_nnz = d->mel_nnz;
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
  const size_t smem = fbank400_smem_bytes(d->hop, d->n_mels, d->mel_nnz);
  dim3 grid((unsigned)(tiles * B));
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
