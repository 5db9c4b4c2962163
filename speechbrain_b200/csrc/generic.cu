// General (any n_fft / hop) kernels: direct-DFT STFT, spectral magnitude and the
// mel projection on a materialised spectrogram.  They back the stand-alone
// STFT / spectral_magnitude / Filterbank modules and Fbank/MFCC with
// n_fft != 400; the default hot path is fbank400.cu.
//   STFT.forward        speechbrain/processing/features.py:133-188
//   spectral_magnitude  speechbrain/processing/features.py:327-356
//   Filterbank.forward  speechbrain/processing/features.py:490-551, 701-702
#include <cmath>

#include "common.cuh"
#include "fft_packed.cuh"

namespace sbfe {

// sample g of a signal of L samples under torch.stft's padding: zeros, or (pad_mode="reflect")
// the mirror image about the first / last sample (processing/features.py:161-172)
__device__ __forceinline__ float padded_sample(const float* __restrict__ w, int64_t g, int64_t L,
                                               int reflect) {
  if (reflect) {
    if (g < 0) g = -g;
    if (g >= L) g = 2 * (L - 1) - g;
  }
  return (g >= 0 && g < L) ? __ldg(w + g) : 0.f;
}

constexpr int kGenFrames = 4;     // frames per CTA
constexpr int kGenThreads = 256;

// smem: twiddle table cs[n_fft] (cos, -sin) + kGenFrames windowed frames.
__global__ void __launch_bounds__(kGenThreads)
stft_generic_kernel(const float* __restrict__ wav, int64_t L, int64_t wav_stride, int64_t T,
                    const float* __restrict__ window, int n_fft, int hop,
                    float* __restrict__ out, int power_only, unsigned tiles, int start, int reflect,
                    int twosided) {
  extern __shared__ __align__(16) float smem[];
  float2* s_cs = reinterpret_cast<float2*>(smem);
  float* s_fr = smem + 2 * n_fft;
  const int bins = n_fft / 2 + 1;
  const int64_t b = blockIdx.x / tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * tiles) * kGenFrames;
  const float* w = wav + b * wav_stride;
  for (int j = threadIdx.x; j < n_fft; j += kGenThreads) {
    float s, c;
    sincospif(2.0f * (float)j / (float)n_fft, &s, &c);
    s_cs[j] = make_float2(c, -s);
  }
  for (int i = threadIdx.x; i < kGenFrames * n_fft; i += kGenThreads) {
    const int f = i / n_fft, n = i - f * n_fft;
    const int64_t g = (t0 + f) * hop + start + n;
    s_fr[i] = padded_sample(w, g, L, reflect) * __ldg(window + n);
  }
  __syncthreads();
  for (int p = threadIdx.x; p < kGenFrames * bins; p += kGenThreads) {
    const int f = p / bins, k = p - f * bins;
    if (t0 + f >= T) continue;
    const float* x = s_fr + f * n_fft;
    float re = 0.f, im = 0.f;
    int idx = 0;
    for (int n = 0; n < n_fft; ++n) {
      const float2 cs = s_cs[idx];
      re = fmaf(x[n], cs.x, re);
      im = fmaf(x[n], cs.y, im);
      idx += k;
      if (idx >= n_fft) idx -= n_fft;
    }
    const int obins = twosided ? n_fft : bins;
    const int64_t o = ((b * T + t0 + f) * obins + k);
    if (power_only) {
      out[o] = re * re + im * im;
      if (twosided && k > 0 && k < n_fft - k) out[o + n_fft - 2 * k] = re * re + im * im;
    } else {
      reinterpret_cast<float2*>(out)[o] = make_float2(re, im);
      if (twosided && k > 0 && k < n_fft - k)   // X[n_fft - k] = conj X[k]
        reinterpret_cast<float2*>(out)[o + n_fft - 2 * k] = make_float2(re, -im);
    }
  }
}

// ---- power-of-two n_fft (256 / 512 / 1024 ...): real FFT through a half-size complex FFT ------
// z[n] = x[2n] + i x[2n+1] (windowed); Stockham autosort FFT of size M = N / 2 in shared memory
// (ping-pong buffers, radix-4 passes and one radix-2 pass when log2 M is odd: natural order in
// and out, no bit reversal), then the real-input untangle
//   X[k] = (Z[k] + conj Z[M-k]) / 2 + W_N^k (Z[k] - conj Z[M-k]) / (2i).
// O(N log N) where stft_generic_kernel is O(N^2); same outputs.
// smem: W_N table cs[N] (cos, -sin) + 2 x kPow2Frames complex buffers of N/2.
constexpr int kPow2Frames = 16;   // frames per CTA: four radix-4 butterflies per thread and pass

__device__ __forceinline__ float2 cmulf(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

// MEL: the Fbank path - the power spectrum stays in shared memory and goes straight through the
// sparse mel projection + dB (Filterbank.forward :490-551, 701-702): nothing but the waveform is
// read from and nothing but the [frames, n_mels] tile is written to HBM.
struct MelArgs {
  int M;
  const int32_t* start;
  const int32_t* count;
  const int32_t* offset;
  const float* weight;
  DbParams db;
  sbfe_maxkey_t* utt_max;
};

// radices of the Stockham passes for M = n_fft / 2 = 2^a 5^b: 4 ... 4 [2] 5 ... 5
struct FftPlan {
  int n;
  unsigned char radix[20];
};

static bool make_fft_plan(int n_fft, FftPlan* plan) {
  int m = n_fft / 2;
  plan->n = 0;
  if (n_fft < 16 || (n_fft & 1)) return false;
  int twos = 0, fives = 0;
  while (m % 2 == 0) { m /= 2; ++twos; }
  while (m % 5 == 0) { m /= 5; ++fives; }
  if (m != 1) return false;
  for (; twos >= 2; twos -= 2) plan->radix[plan->n++] = 4;
  if (twos) plan->radix[plan->n++] = 2;
  for (; fives > 0; --fives) plan->radix[plan->n++] = 5;
  return true;
}

// MODE 0: STFT (complex or power spectrum to HBM); 1: Fbank (power -> mel -> dB, MelArgs);
// 2: INVERSE real FFT for ISTFT - `wav` is the one-sided spectrum [T][N/2 + 1][2] of T frames,
//    `out` the windowed time frames [T][N]: Z[k] = E[k] + i O[k] is rebuilt from X (the inverse of
//    the untangle step), conj(Z) goes through the SAME forward passes (IFFT(Z) = conj(FFT(conj Z)) / M)
//    and x[2n] = Re z[n], x[2n+1] = Im z[n].
template <int MODE>
__global__ void __launch_bounds__(kGenThreads)
stft_pow2_kernel(const float* __restrict__ wav, int64_t L, int64_t wav_stride, int64_t T,
                 const float* __restrict__ window, int n_fft, const FftPlan plan, int hop,
                 float* __restrict__ out, int power_only, unsigned tiles, const MelArgs mel,
                 int start, int reflect, int twosided) {
  constexpr bool MEL = MODE == 1, INV = MODE == 2;
  extern __shared__ __align__(16) float smem[];
  const int M = n_fft >> 1;
  float2* s_cs = reinterpret_cast<float2*>(smem);        // W_N^j, j < N
  float2* s_a = s_cs + n_fft;                            // [kPow2Frames][M]
  float2* s_b = s_a + kPow2Frames * M;                   // [kPow2Frames][M]
  const int bins = M + 1;
  const int64_t b = blockIdx.x / tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * tiles) * kPow2Frames;
  const float* w = wav + b * wav_stride;
  for (int j = threadIdx.x; j < n_fft; j += kGenThreads) {
    float sn, cs;
    sincospif(2.0f * (float)j / (float)n_fft, &sn, &cs);
    s_cs[j] = make_float2(cs, -sn);
  }
  if (INV) {
    // thread = bin k: Z[k] = E + i O with E = (X[k] + conj X[M-k]) / 2, O = conj(W^k) (X[k] - conj X[M-k]) / 2;
    // conj(Z) is stored (the imaginary parts of X[0] and X[M] do not exist for a real signal)
    const float2* spec = reinterpret_cast<const float2*>(wav);
    for (int k = threadIdx.x; k < M; k += kGenThreads) {
      float sn, cs;
      sincospif(2.0f * (float)k / (float)n_fft, &sn, &cs);
      for (int f = 0; f < kPow2Frames; ++f) {
        float2 z = make_float2(0.f, 0.f);
        if (t0 + f < T) {
          const float2* x = spec + (t0 + f) * bins;
          float2 a = __ldg(x + k), c = __ldg(x + (M - k));
          if (k == 0) { a.y = 0.f; c.y = 0.f; }
          const float ex = 0.5f * (a.x + c.x), ey = 0.5f * (a.y - c.y);
          const float dx = 0.5f * (a.x - c.x), dy = 0.5f * (a.y + c.y);
          const float ox = dx * cs - dy * sn, oy = dx * sn + dy * cs;   // (cos + i sin) * D
          z = make_float2(ex - oy, -(ey + ox));                         // conj(E + i O)
        }
        s_a[f * M + k] = z;
      }
    }
  }
  // thread = sample pair n (its two window taps are loaded once), loop over the frames: no
  // division, no per-element window load
  for (int n = threadIdx.x; n < M && !INV; n += kGenThreads) {
    const float w0 = __ldg(window + 2 * n), w1 = __ldg(window + 2 * n + 1);
    int64_t g = t0 * hop + start + 2 * n;
#pragma unroll 4
    for (int f = 0; f < kPow2Frames; ++f, g += hop) {
      const float x0 = padded_sample(w, g, L, reflect) * w0;
      const float x1 = padded_sample(w, g + 1, L, reflect) * w1;
      s_a[f * M + n] = make_float2(x0, x1);
    }
  }
  __syncthreads();
  float2* src = s_a;
  float2* dst = s_b;
  // Stockham passes: radix R with Ns = the size already transformed:
  //   v[r] = in[j + r M/R] * W_(Ns R)^(k r), k = j mod Ns;  out[(j / Ns) Ns R + k + r Ns] = DFT_R(v)[r]
  // The power-of-two passes come first (Ns is then a power of two: shifts), the radix-5 passes last.
  int Ns = 1;
  for (int ps = 0; ps < plan.n; ++ps) {
    const int R = plan.radix[ps];
    const int q = M / R;
    const int tw1 = n_fft / (Ns * R);                    // table step of W_(Ns R)
    // thread -> (frame slot, butterfly): with q < 256 butterflies per frame several frames run
    // side by side (one division per thread and pass)
    const int fpar = q >= kGenThreads ? 1 : kGenThreads / q;
    const int f_first = q >= kGenThreads ? 0 : threadIdx.x / q;
    const int j_first = threadIdx.x - f_first * q;
    const int j_step = q >= kGenThreads ? kGenThreads : q;   // q < 256: one butterfly per frame
    const int f_end = f_first < fpar ? kPow2Frames : 0;
    if (R == 4) {
      const int ls = 31 - __clz(Ns);
      for (int f = f_first; f < f_end; f += fpar)
        for (int j = j_first; j < q; j += j_step) {
          const int k = j & (Ns - 1);
          const int tstep = k * tw1;
          const float2* in = src + f * M + j;
          float2 v0 = in[0], v1 = in[q], v2 = in[2 * q], v3 = in[3 * q];
          v1 = cmulf(v1, s_cs[tstep]);
          v2 = cmulf(v2, s_cs[2 * tstep]);
          v3 = cmulf(v3, s_cs[3 * tstep]);
          const float2 a0 = make_float2(v0.x + v2.x, v0.y + v2.y), a1 = make_float2(v0.x - v2.x, v0.y - v2.y);
          const float2 a2 = make_float2(v1.x + v3.x, v1.y + v3.y);
          const float2 a3 = make_float2(v1.y - v3.y, v3.x - v1.x);   // -i (v1 - v3)
          float2* o = dst + f * M + (((j >> ls) << (ls + 2)) + k);
          if (Ns == 1) {   // first pass: the four outputs are adjacent - two 128-bit stores
            float4* o4 = reinterpret_cast<float4*>(o);
            o4[0] = make_float4(a0.x + a2.x, a0.y + a2.y, a1.x + a3.x, a1.y + a3.y);
            o4[1] = make_float4(a0.x - a2.x, a0.y - a2.y, a1.x - a3.x, a1.y - a3.y);
          } else {
            o[0] = make_float2(a0.x + a2.x, a0.y + a2.y);
            o[Ns] = make_float2(a1.x + a3.x, a1.y + a3.y);
            o[2 * Ns] = make_float2(a0.x - a2.x, a0.y - a2.y);
            o[3 * Ns] = make_float2(a1.x - a3.x, a1.y - a3.y);
          }
        }
    } else if (R == 2) {
      const int ls = 31 - __clz(Ns);
      for (int f = f_first; f < f_end; f += fpar)
        for (int j = j_first; j < q; j += j_step) {
          const int k = j & (Ns - 1);
          const float2* in = src + f * M + j;
          const float2 v0 = in[0], v1 = cmulf(in[q], s_cs[k * tw1]);
          float2* o = dst + f * M + (((j >> ls) << (ls + 1)) + k);
          o[0] = make_float2(v0.x + v1.x, v0.y + v1.y);
          o[Ns] = make_float2(v0.x - v1.x, v0.y - v1.y);
        }
    } else {  // R == 5
      for (int f = f_first; f < f_end; f += fpar)
        for (int j = j_first; j < q; j += j_step) {
          const int jb = j / Ns, k = j - jb * Ns;
          const int tstep = k * tw1;
          const float2* in = src + f * M + j;
          cx<float> v[5];
#pragma unroll
          for (int r = 0; r < 5; ++r) {
            float2 t = in[r * q];
            if (r) t = cmulf(t, s_cs[r * tstep]);
            v[r] = cx<float>{t.x, t.y};
          }
          dft5(v[0], v[1], v[2], v[3], v[4]);
          float2* o = dst + f * M + (jb * Ns * 5 + k);
#pragma unroll
          for (int r = 0; r < 5; ++r) o[r * Ns] = make_float2(v[r].re, v[r].im);
        }
    }
    Ns *= R;
    __syncthreads();
    float2* t = src; src = dst; dst = t;
  }
  const int64_t left = T - t0;
  const int nf = left < kPow2Frames ? (int)left : kPow2Frames;
  if (INV) {   // x[2n] = Re z[n] / M, x[2n+1] = Im z[n] / M with z = conj(result); times the window
    const float inv_m = 1.0f / (float)M;
    for (int n = threadIdx.x; n < M; n += kGenThreads) {
      const float w0 = __ldg(window + 2 * n) * inv_m, w1 = -__ldg(window + 2 * n + 1) * inv_m;
      for (int f = 0; f < nf; ++f) {
        const float2 r = src[f * M + n];
        reinterpret_cast<float2*>(out + (t0 + f) * n_fft)[n] = make_float2(r.x * w0, r.y * w1);
      }
    }
    return;
  }
  // untangle: thread = bin k (its twiddle is loaded once), loop over the frames; the Nyquist
  // bin k = M of frame f goes to thread f of the last pass
  for (int k0 = threadIdx.x; k0 < M + kGenThreads; k0 += kGenThreads) {
    int k = k0, f_lo = 0, f_hi = nf;
    if (k0 >= M) {                       // extra slots: (frame k0 - M, bin M)
      if (k0 - M >= nf) break;
      k = M;
      f_lo = k0 - M;
      f_hi = f_lo + 1;
    }
    const float2 tw = s_cs[k];           // W_N^k (k = M: -1)
    for (int f = f_lo; f < f_hi; ++f) {
      const float2* z = src + f * M;
      const float2 zk = z[k == M ? 0 : k], zm = z[k == 0 ? 0 : M - k];
      const float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);   // even part
      const float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x); // odd: (Zk - conj Zm) / 2i
      const float re = er + orr * tw.x - oi * tw.y, im = ei + orr * tw.y + oi * tw.x;
      if (MEL) {
        reinterpret_cast<float*>(dst)[f * bins + k] = re * re + im * im;  // idle buffer: [frame][bins]
        continue;
      }
      const int obins = twosided ? n_fft : bins;
      const int64_t o = ((b * T + t0 + f) * obins + k);
      if (power_only) {
        out[o] = re * re + im * im;
        if (twosided && k > 0 && k < M) out[o + n_fft - 2 * k] = re * re + im * im;
      } else {
        reinterpret_cast<float2*>(out)[o] = make_float2(re, im);
        if (twosided && k > 0 && k < M)   // X[n_fft - k] = conj X[k]
          reinterpret_cast<float2*>(out)[o + n_fft - 2 * k] = make_float2(re, -im);
      }
    }
  }
  if (MEL) {
    __syncthreads();
    // thread = filter m (its table entry is loaded once) x frame group: lanes walk over m
    const float* s_p = reinterpret_cast<const float*>(dst);
    float vmax = -INFINITY;
    float* o = out + (b * T + t0) * mel.M;
    const int fgroups = kGenThreads / mel.M > 0 ? kGenThreads / mel.M : 1;
    const int fg = threadIdx.x / mel.M;
    for (int m = threadIdx.x - fg * mel.M; m < mel.M && fg < fgroups; m += kGenThreads) {
      const int cnt = __ldg(mel.count + m);
      const float* wt = mel.weight + __ldg(mel.offset + m);
      const float* x0 = s_p + __ldg(mel.start + m);
      for (int f = fg; f < nf; f += fgroups) {
        const float* x = x0 + f * bins;
        float acc = 0.f;
        for (int i = 0; i < cnt; ++i) acc = fmaf(__ldg(wt + i), x[i], acc);
        const float v = to_db(acc, mel.db);
        o[f * mel.M + m] = v;
        vmax = fmaxf(vmax, v);
      }
    }
    vmax = warp_max(vmax);
    if ((threadIdx.x & 31) == 0 && vmax > -INFINITY && mel.utt_max != nullptr)
      atomicMax(mel.utt_max + b, max_key_encode(vmax));
  }
}

static bool pow2_fits(int n_fft) {   // n_fft / 2 = 2^a 5^b and the ping-pong buffers fit
  FftPlan plan;
  return make_fft_plan(n_fft, &plan) &&
         (size_t)(2 + 2 * kPow2Frames) * n_fft * sizeof(float) <= 200 * 1024;
}

// Fbank with a power-of-two n_fft in one launch (any hop, any filter shape: the CSR tables); the
// idle frame buffer [16][N/2] float2 always holds the power tile [16][N/2 + 1] floats
bool fbank_pow2_supported(const sbfe_fbank_desc* d) { return pow2_fits(d->n_fft) && d->n_mels >= 1; }

int launch_fbank_pow2(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                      int64_t wav_stride, float* out, sbfe_maxkey_t* utt_max, cudaStream_t stream) {
  const int64_t T = 1 + L / d->hop;
  const int64_t tiles = (T + kPow2Frames - 1) / kPow2Frames;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  FftPlan plan;
  make_fft_plan(d->n_fft, &plan);
  const size_t smem = (size_t)(2 + 2 * kPow2Frames) * d->n_fft * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(stft_pow2_kernel<1>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  MelArgs mel{d->n_mels, d->d_mel_start, d->d_mel_count, d->d_mel_offset, d->d_mel_weight,
              DbParams{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                       d->log_mel},
              utt_max};
  stft_pow2_kernel<1><<<(unsigned)(tiles * B), kGenThreads, smem, stream>>>(
      wav, L, wav_stride, T, d->d_window, d->n_fft, plan, d->hop, out, 1, (unsigned)tiles, mel,
      -d->n_fft / 2, 0, 0);
  return check_last();
}

// ISTFT step 1 through the FFT: spec [n_frames][n_fft/2+1][2] -> windowed frames [n_frames][n_fft]
bool irfft_frames_supported(int n_fft) { return pow2_fits(n_fft); }

int launch_irfft_frames(const float* spec, int64_t n_frames, int n_fft, const float* window,
                        float* frames, cudaStream_t stream) {
  if (n_frames == 0) return 0;
  const int64_t tiles = (n_frames + kPow2Frames - 1) / kPow2Frames;
  if (tiles > 0x7fffffff) return SBFE_E_TOOBIG;
  FftPlan plan;
  make_fft_plan(n_fft, &plan);
  const size_t smem = (size_t)(2 + 2 * kPow2Frames) * n_fft * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(stft_pow2_kernel<2>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  stft_pow2_kernel<2><<<(unsigned)tiles, kGenThreads, smem, stream>>>(
      spec, 0, 0, n_frames, window, n_fft, plan, 1, frames, 0, (unsigned)tiles, MelArgs{}, 0, 0, 0);
  return check_last();
}

// flags (include/sbfe.h): SBFE_STFT_NO_CENTER 1, SBFE_STFT_REFLECT 2, SBFE_STFT_TWOSIDED 4
int64_t stft_frames(const sbfe_fbank_desc* d, int64_t L, int flags) {
  if (flags & 1) return L < d->n_fft ? 0 : 1 + (L - d->n_fft) / d->hop;
  return 1 + L / d->hop;
}

bool stft400_power_supported(const sbfe_fbank_desc* d);   // fbank400_v3.cu
int launch_stft400_power(const sbfe_fbank_desc*, const float*, int64_t, int64_t, int64_t, float*, int,
                         cudaStream_t);
bool stft512_power_supported(const sbfe_fbank_desc* d);   // fbank512.cu
int launch_stft512_power(const sbfe_fbank_desc*, const float*, int64_t, int64_t, int64_t, float*, int,
                         cudaStream_t);

int launch_stft_generic(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                        int64_t wav_stride, float* out, int power_only, int flags,
                        cudaStream_t stream) {
  const int64_t T = stft_frames(d, L, flags);
  if (T == 0) return 0;
  // the recipes' transforms (n_fft 400 / 512, hop 160, default mode), power only: the fused
  // kernels' FFTs (SBFE_STFT400_OFF: the Stockham kernel below, for the tests' cross-check)
  if (flags == 0 && getenv("SBFE_STFT400_OFF") == nullptr) {
    if (stft400_power_supported(d))   // power spectrum or the complex STFT itself
      return launch_stft400_power(d, wav, B, L, wav_stride, out, power_only ? 0 : 1, stream);
    if (stft512_power_supported(d))
      return launch_stft512_power(d, wav, B, L, wav_stride, out, power_only ? 0 : 1, stream);
  }
  const int start = (flags & 1) ? 0 : -d->n_fft / 2;
  const int reflect = (flags & 2) ? 1 : 0, twosided = (flags & 4) ? 1 : 0;
  const int64_t tiles = (T + kGenFrames - 1) / kGenFrames;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  if (d->n_fft > 4096) return SBFE_E_TOOBIG;
  // power of two (and the ping-pong buffers fit in shared memory): the FFT kernel
  const size_t smem2 = (size_t)(2 + 2 * kPow2Frames) * d->n_fft * sizeof(float);
  if (pow2_fits(d->n_fft)) {
    FftPlan plan;
    make_fft_plan(d->n_fft, &plan);
    const int64_t tiles2 = (T + kPow2Frames - 1) / kPow2Frames;
    cudaError_t e2 = cudaFuncSetAttribute(stft_pow2_kernel<0>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
    if (e2 != cudaSuccess) return (int)e2;
    stft_pow2_kernel<0><<<(unsigned)(tiles2 * B), kGenThreads, smem2, stream>>>(
        wav, L, wav_stride, T, d->d_window, d->n_fft, plan, d->hop, out, power_only, (unsigned)tiles2,
        MelArgs{}, start, reflect, twosided);
    return check_last();
  }
  const size_t smem = (size_t)(2 + kGenFrames) * d->n_fft * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(stft_generic_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  stft_generic_kernel<<<(unsigned)(tiles * B), kGenThreads, smem, stream>>>(
      wav, L, wav_stride, T, d->d_window, d->n_fft, d->hop, out, power_only, (unsigned)tiles, start,
      reflect, twosided);
  return check_last();
}

__global__ void spectral_magnitude_kernel(const float2* __restrict__ in, int64_t n, float power,
                                          int take_log, float eps, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float2 v = in[i];
    float s = v.x * v.x + v.y * v.y;
    if (power < 1.f) s += eps;
    if (power != 1.f) s = powf(s, power);
    out[i] = take_log ? logf(s + eps) : s;
  }
}

int launch_spectral_magnitude(const float* in, int64_t n, float power, int take_log, float eps,
                              float* out, cudaStream_t stream) {
  if (n == 0) return 0;
  int64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  spectral_magnitude_kernel<<<(unsigned)blocks, 256, 0, stream>>>(
      reinterpret_cast<const float2*>(in), n, power, take_log, eps, out);
  return check_last();
}

constexpr int kMelFrames = 8;
constexpr int kMelThreads = 256;

__global__ void __launch_bounds__(kMelThreads)
mel_db_kernel(const float* __restrict__ spec, int64_t T, int bins, int M,
              const int32_t* __restrict__ mel_start, const int32_t* __restrict__ mel_count,
              const int32_t* __restrict__ mel_offset, const float* __restrict__ mel_weight,
              DbParams db, float* __restrict__ out, sbfe_maxkey_t* utt_max, unsigned tiles) {
  extern __shared__ __align__(16) float s_spec[];  // [kMelFrames][bins]
  const int64_t b = blockIdx.x / tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * tiles) * kMelFrames;
  const int64_t left = T - t0;
  const int nf = left < kMelFrames ? (int)left : kMelFrames;
  const float* src = spec + (b * T + t0) * bins;
  for (int i = threadIdx.x; i < nf * bins; i += kMelThreads) s_spec[i] = __ldg(src + i);
  __syncthreads();
  float vmax = -INFINITY;
  float* dst = out + (b * T + t0) * M;
  for (int p = threadIdx.x; p < nf * M; p += kMelThreads) {
    const int f = p / M, m = p - f * M;
    const int start = mel_start[m], cnt = mel_count[m];
    const float* wt = mel_weight + mel_offset[m];
    const float* x = s_spec + f * bins + start;
    float acc = 0.f;
    for (int i = 0; i < cnt; ++i) acc = fmaf(__ldg(wt + i), x[i], acc);
    const float v = to_db(acc, db);
    dst[p] = v;
    vmax = fmaxf(vmax, v);
  }
  vmax = warp_max(vmax);
  if ((threadIdx.x & 31) == 0 && vmax > -INFINITY && utt_max != nullptr)
    atomicMax(utt_max + b, max_key_encode(vmax));
}

static int launch_mel_db_sparse(const sbfe_fbank_desc* d, const float* spec, int64_t B, int64_t T,
                                float* out, sbfe_maxkey_t* utt_max, cudaStream_t stream);

// ---- dense filterbanks on the tensor cores ------------------------------------------------------
// Gaussian filters (speechbrain/processing/features.py:637-658) make Filterbank.forward's matmul
// (:548) a genuinely dense [frames x bins] . [bins x M] contraction (triangular / rectangular
// ones are ~2 % dense and stay on the sparse loops).  mma.sync.m16n8k8 TF32 with fp32 accumulators;
// SPLIT = 3 is the 3xTF32 scheme (x = hi + lo, both TF32: hi*hi + hi*lo + lo*hi, error ~2^-21
// relative: inside the fp32 tolerance), SPLIT = 1 a single TF32 pass (10-bit mantissas: the
// stated looser bound, include/sbfe.h mel_precision).
// Why mma.sync and not tcgen05 / TMEM: K = bins (201) and N = M (<= 128) are tiny and the kernel
// reads 4 * bins and writes 4 * M bytes per frame - it is bound by HBM (or L2), not by the tensor
// pipe (3 x 2 * bins * M flop per frame = 0.1 MFLOP against 1.1 KB: 87 flop/byte, a B200 is at
// > 280); a UMMA tile needs 128 rows in shared memory through descriptors, its accumulator comes
// back through tcgen05.ld for an epilogue that is one lg2 and a store.  Warp-level MMA fed from a
// conflict-free fp32 tile does the same work with none of that and the same memory time.
// Persistent CTAs (one per SM): the [M][bins] weights stay in shared memory, frame tiles are
// double-buffered with cp.async.
constexpr int kMmaRows = 64;       // frames per tile
constexpr int kMmaThreads = 512;   // 16 warps = 4 row groups of 16 x 4 column groups
constexpr int kMmaColGroups = 4;
constexpr int kMmaMaxNT = 4;       // n-tiles of 8 filters per warp: M <= 128
constexpr int kMmaDensity = 5;          // n_fft 400: dense path from 1/5 stored weights (see mel_dense_rule)
constexpr int kMmaDensitySlowPath = 10;
bool fbank400_supported(const sbfe_fbank_desc* d);   // fbank400.cu

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int SPLIT, int NT>      // NT: n-tiles of 8 filters per warp (ceil(M / 32))
__global__ void __launch_bounds__(kMmaThreads, 1)
mel_mma_kernel(const float* __restrict__ spec, int64_t rows_total, int64_t T, int bins, int M,
               int kpad, int npad, const int32_t* __restrict__ mel_start,
               const int32_t* __restrict__ mel_count, const int32_t* __restrict__ mel_offset,
               const float* __restrict__ mel_weight, DbParams db, float* __restrict__ out,
               sbfe_maxkey_t* utt_max, int nbuf) {
  extern __shared__ __align__(16) float s_mma[];
  const int pitch = kpad + 4;                      // = 4 mod 8: fragment loads hit 32 distinct banks
  float* s_w = s_mma;                              // [npad][pitch]: W^T, zero outside each filter
  float* s_a = s_w + (size_t)npad * pitch;         // 2 x [kMmaRows][pitch]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  __shared__ unsigned s_um[kMmaRows];              // max keys of the utterances a tile meets
  if (tid < kMmaRows) s_um[tid] = 0u;
  for (int i = tid; i < npad * pitch; i += kMmaThreads) s_w[i] = 0.f;
  for (int i = tid; i < nbuf * kMmaRows * pitch; i += kMmaThreads) s_a[i] = 0.f;   // k >= bins stays 0
  __syncthreads();
  for (int m = warp; m < M; m += kMmaThreads / 32) {
    const int st = __ldg(mel_start + m), cnt = __ldg(mel_count + m);
    const float* wt = mel_weight + __ldg(mel_offset + m);
    for (int i = lane; i < cnt; i += 32) s_w[m * pitch + st + i] = __ldg(wt + i);
  }
  const int64_t n_tiles = (rows_total + kMmaRows - 1) / kMmaRows;
  auto stage = [&](int64_t tile, float* dst) {     // rows of a tile are one contiguous block of spec
    const int64_t r0 = tile * kMmaRows;
    const int nr = rows_total - r0 < kMmaRows ? (int)(rows_total - r0) : kMmaRows;
    const float* src = spec + r0 * bins;
    const int total = nr * bins;
    int r = tid / bins, k = tid - r * bins;
    const int dr = kMmaThreads / bins, dk = kMmaThreads - dr * bins;
    for (int i = tid; i < total; i += kMmaThreads) {
      const unsigned d = (unsigned)__cvta_generic_to_shared(dst + r * pitch + k);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(src + i) : "memory");
      r += dr;
      k += dk;
      if (k >= bins) { k -= bins; ++r; }
    }
  };
  int64_t tile = blockIdx.x;
  if (tile < n_tiles) stage(tile, s_a);
  asm volatile("cp.async.commit_group;\n" ::: "memory");
  const int rg = warp & 3, ch = warp >> 2;
  const int nt0 = ch * NT;                         // npad = 32 NT: every warp has NT n-tiles
  constexpr int my_nt = NT;
  for (int it = 0; tile < n_tiles; ++it, tile += gridDim.x) {
    // nbuf == 2: the next tile flies while this one is multiplied; 1 (large n_fft: the second
    // buffer does not fit): it is requested once this one is done
    float* a_buf = s_a + (size_t)(nbuf == 2 ? (it & 1) : 0) * kMmaRows * pitch;
    if (nbuf == 2 && tile + gridDim.x < n_tiles)
      stage(tile + gridDim.x, s_a + (size_t)((it & 1) ^ 1) * kMmaRows * pitch);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    if (nbuf == 2) asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    else asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();                               // this tile (and, first time, the weights) is in
    float acc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
    const float* ap = a_buf + (rg * 16 + g) * pitch + t;
    const float* ap8 = ap + 8 * pitch;
    const float* bp = s_w + (nt0 * 8 + g) * pitch + t;
    const int bstep = 8 * pitch;
#pragma unroll 2
    for (int k0 = 0; k0 < kpad; k0 += 8) {
      const float af[4] = {ap[k0], ap8[k0], ap[k0 + 4], ap8[k0 + 4]};
      uint32_t ah[4], al[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ah[i] = to_tf32(af[i]);
        if (SPLIT == 3) al[i] = to_tf32(af[i] - __uint_as_float(ah[i]));
      }
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        if (n < my_nt) {
          const float bf[2] = {bp[n * bstep + k0], bp[n * bstep + k0 + 4]};
          uint32_t bh[2], bl[2];
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            bh[i] = to_tf32(bf[i]);
            if (SPLIT == 3) bl[i] = to_tf32(bf[i] - __uint_as_float(bh[i]));
          }
          if (SPLIT == 3) {                        // the small terms first
            mma_tf32(acc[n], al, bh);
            mma_tf32(acc[n], ah, bl);
          }
          mma_tf32(acc[n], ah, bh);
        }
      }
    }
    // epilogue: dB, store, utterance maxima.  c0 c1: row g, filters 2t, 2t+1; c2 c3: row g + 8
    const int64_t row_a = tile * kMmaRows + rg * 16 + g, row_b = row_a + 8;
    float vmax_a = -INFINITY, vmax_b = -INFINITY;
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      if (n < my_nt) {
        const int m0 = (nt0 + n) * 8 + 2 * t;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int m = m0 + (i & 1);
          const int64_t row = (i & 2) ? row_b : row_a;
          if (m < M && row < rows_total) {
            const float v = to_db(acc[n][i], db);
            out[row * M + m] = v;
            if (i & 2) vmax_b = fmaxf(vmax_b, v);
            else vmax_a = fmaxf(vmax_a, v);
          }
        }
      }
    }
    // utterance maxima: through shared memory first (a tile of 64 rows meets one or two
    // utterances; same-address global atomics from every quad would serialise in L2)
    if (utt_max != nullptr) {
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1) {           // the four lanes of a row
        vmax_a = fmaxf(vmax_a, __shfl_xor_sync(0xffffffffu, vmax_a, o));
        vmax_b = fmaxf(vmax_b, __shfl_xor_sync(0xffffffffu, vmax_b, o));
      }
      const int64_t b_first = (tile * kMmaRows) / T;
      if (t == 0) {
        if (vmax_a > -INFINITY) atomicMax(&s_um[row_a / T - b_first], max_key_encode(vmax_a));
        if (vmax_b > -INFINITY) atomicMax(&s_um[row_b / T - b_first], max_key_encode(vmax_b));
      }
      __syncthreads();
      if (tid < kMmaRows) {
        const unsigned k = s_um[tid];
        if (k) {
          atomicMax(utt_max + b_first + tid, k);
          s_um[tid] = 0u;
        }
      }
    }
    __syncthreads();                               // a_buf is free for the tile after next
    if (nbuf == 1 && tile + gridDim.x < n_tiles) stage(tile + gridDim.x, s_a);
  }
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// dense enough for the tensor-core kernel, and its weights + two frame tiles fit in shared memory
bool mel_umma_ok(const sbfe_fbank_desc* d);          // mel_umma.cu
int launch_mel_umma(const sbfe_fbank_desc*, const float*, int64_t, int64_t, float*, sbfe_maxkey_t*,
                    cudaStream_t);

// does this filterbank count as dense?  SBFE_MEL_MMA = 0: never, 2: always (experiments, tests)
static bool mel_dense_rule(const sbfe_fbank_desc* d) {
  const int bins = d->n_fft / 2 + 1;
  const char* sw = getenv("SBFE_MEL_MMA");
  if (sw != nullptr && sw[0] == '0') return false;
  if (sw != nullptr && sw[0] == '2') return true;
  // where a fused n_fft = 400 kernel takes the configuration, its sparse loops cost about
  // 36 us + 0.02 us per stored weight on the cfg2 batch (triangular-80: 44 us, gaussian-80 at 18 %
  // stored weights: 96 us) against 92 us for power STFT + tcgen05 projection whatever the
  // density: the dense path from 1/5 on; elsewhere (n_fft 512 ...) 10 % is enough
  const int density = fbank400_supported(d) ? kMmaDensity : kMmaDensitySlowPath;
  return (int64_t)density * d->mel_nnz >= (int64_t)bins * d->n_mels;
}

static bool mel_mma_ok(const sbfe_fbank_desc* d, size_t* smem_out, int* nbuf_out = nullptr) {
  const int bins = d->n_fft / 2 + 1;
  const int kpad = (bins + 7) & ~7, npad = (d->n_mels + 31) & ~31;   // 4 column groups x NT x 8
  size_t smem = ((size_t)npad + 2 * kMmaRows) * (kpad + 4) * sizeof(float);
  int nbuf = 2;
  if (smem > 220 * 1024) {
    nbuf = 1;
    smem = ((size_t)npad + kMmaRows) * (kpad + 4) * sizeof(float);
  }
  if (nbuf_out) *nbuf_out = nbuf;
  if (smem_out) *smem_out = smem;
  if (npad > 8 * kMmaColGroups * kMmaMaxNT || smem > 220 * 1024) return false;
  return mel_dense_rule(d);
}

// Fbank with a dense filterbank = STFT power spectrum (workspace) + the tensor-core projection
bool mel_dense_mma(const sbfe_fbank_desc* d) {
  return (mel_dense_rule(d) && mel_umma_ok(d)) || mel_mma_ok(d, nullptr);
}

int launch_mel_db(const sbfe_fbank_desc* d, const float* spec, int64_t B, int64_t T, float* out,
                  sbfe_maxkey_t* utt_max, cudaStream_t stream) {
  if (B == 0 || T == 0) return 0;
  size_t smem_mma = 0;
  int nbuf = 2;
  // dense: tcgen05 + TMEM (mel_umma.cu) where its stages fit, else the warp-level mma.sync kernel
  if (mel_dense_rule(d) && mel_umma_ok(d)) return launch_mel_umma(d, spec, B, T, out, utt_max, stream);
  if (mel_mma_ok(d, &smem_mma, &nbuf)) {
    const int bins = d->n_fft / 2 + 1;
    const int kpad = (bins + 7) & ~7, npad = (d->n_mels + 31) & ~31;
    const int nt = npad / (8 * kMmaColGroups);                       // 1 .. kMmaMaxNT
    using Kern = decltype(&mel_mma_kernel<3, 1>);
    static const Kern table[2][kMmaMaxNT] = {
        {mel_mma_kernel<3, 1>, mel_mma_kernel<3, 2>, mel_mma_kernel<3, 3>, mel_mma_kernel<3, 4>},
        {mel_mma_kernel<1, 1>, mel_mma_kernel<1, 2>, mel_mma_kernel<1, 3>, mel_mma_kernel<1, 4>}};
    Kern k = table[d->mel_precision == 1 ? 1 : 0][nt - 1];
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_mma);
    int dev = 0, sms = 0;
    if (e == cudaSuccess) e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return (int)e;
    const int64_t rows = B * T, n_tiles = (rows + kMmaRows - 1) / kMmaRows;
    const DbParams db{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                      d->log_mel};
    k<<<(unsigned)(n_tiles < sms ? n_tiles : sms), kMmaThreads, smem_mma, stream>>>(
        spec, rows, T, bins, d->n_mels, kpad, npad, d->d_mel_start, d->d_mel_count, d->d_mel_offset,
        d->d_mel_weight, db, out, utt_max, nbuf);
    return check_last();
  }
  return launch_mel_db_sparse(d, spec, B, T, out, utt_max, stream);
}

static int launch_mel_db_sparse(const sbfe_fbank_desc* d, const float* spec, int64_t B, int64_t T, float* out,
                  sbfe_maxkey_t* utt_max, cudaStream_t stream) {
  if (B == 0 || T == 0) return 0;
  const int bins = d->n_fft / 2 + 1;
  const int64_t tiles = (T + kMelFrames - 1) / kMelFrames;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  const size_t smem = (size_t)kMelFrames * bins * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(mel_db_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  const DbParams db{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                    d->log_mel};
  mel_db_kernel<<<(unsigned)(tiles * B), kMelThreads, smem, stream>>>(
      spec, T, bins, d->n_mels, d->d_mel_start, d->d_mel_count, d->d_mel_offset, d->d_mel_weight,
      db, out, utt_max, (unsigned)tiles);
  return check_last();
}

}  // namespace sbfe
