// Time-domain augmentation kernels: the step right before the feature path in the recipes
// (recipes/LibriSpeech/ASR/transformer/train.py:58-63), SURVEY.md 8f row N1, second half.
//   DropChunk   speechbrain/processing/speech_augmentation.py:1013-1173
//   DropFreq    :909-1010 (the notch filters are composed on the host; this is the convolution)
//   DoClip      :1176-1228
//   Resample    :511-820  (SpeedPerturb :435-508 picks one of them)
// The random draws stay the reference's torch calls on the host (same generator streams); these
// kernels do the data movement the reference spreads over per-sentence Python loops, slice
// assignments and one conv1d + conv_transpose1d + pad + add per polyphase branch.
#include "common.cuh"

namespace sbfe {

constexpr int kTdThreads = 256;

// ---- DropChunk: out = x with [start, start + len) of utterance b zeroed (or filled with the
// caller's uniform draws scaled to +- noise_max[b]) for each of its chunks -----------------------
// d_chunk = [n_chunks][4] int32 {utterance, start (in time steps), length, offset into d_noise};
// chunks of one utterance are applied in the order given (later ones overwrite earlier ones,
// like the reference's loop of slice assignments); C = channels per time step.
// grid = n_chunks x ceil(max_len * C / (256 * 4)); out must already hold a copy of x.
__global__ void __launch_bounds__(kTdThreads)
drop_chunks_kernel(float* __restrict__ out, int64_t L, int C, const int32_t* __restrict__ chunk,
                   const float* __restrict__ noise, const float* __restrict__ noise_max) {
  pdl_wait();
  pdl_launch_dependents();
  const int32_t* c = chunk + 4 * blockIdx.x;
  const int64_t b = c[0];
  int64_t s = c[1], n = c[2];
  if (s < 0) { n += s; s = 0; }                      // python slices clip at the ends
  if (s + n > L) n = L - s;
  const int64_t total = n * C;
  float* o = out + (b * L + s) * C;
  const float* u = noise != nullptr ? noise + (int64_t)c[3] * C : nullptr;
  const float nm = noise_max != nullptr ? noise_max[b] : 0.f;
  for (int64_t i = (int64_t)blockIdx.y * kTdThreads + threadIdx.x; i < total;
       i += (int64_t)gridDim.y * kTdThreads)
    o[i] = u != nullptr ? 2.f * nm * u[i] - nm : 0.f;   // 2 * noise_max * rand - noise_max (:1169)
}

// Overlapping chunks of one utterance must land in order: one launch per "round" (the j-th chunk
// of every utterance), which the host arranges; chunks inside a launch never share an utterance.
int launch_drop_chunks(float* out, int64_t L, int C, const int32_t* chunk, int n_chunks,
                       int64_t max_len, const float* noise, const float* noise_max,
                       cudaStream_t stream) {
  if (n_chunks == 0 || max_len <= 0) return 0;
  int64_t by = (max_len * C + kTdThreads * 4 - 1) / (kTdThreads * 4);
  if (by > 65535) by = 65535;
  return launch_pdl(drop_chunks_kernel, dim3((unsigned)n_chunks, (unsigned)by), dim3(kTdThreads), 0,
                    stream, out, L, C, chunk, noise, noise_max);
}

// ---- FIR with zero padding ("same" size): out[b][t] = sum_k h[k] x[b][t + k - pad] ---------------
// = torch.nn.functional.conv1d on the zero-padded signal, which is what convolve1d(padding = pad)
// does (signal_processing.py:157-277: no kernel flip).  A CTA stages kFirTile + K - 1 samples in
// shared memory; 4 outputs per thread share their window loads.
constexpr int kFirTile = 4 * kTdThreads;   // outputs per CTA
constexpr int kFirMaxTaps = 1024;

__global__ void __launch_bounds__(kTdThreads)
fir_same_kernel(const float* __restrict__ x, int64_t L, const float* __restrict__ h, int K, int pad,
                float* __restrict__ out, unsigned tiles) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ float s_fir[];                   // [K] taps | [kFirTile + K - 1] samples
  float* s_h = s_fir;
  float* s_x = s_fir + K;
  const int64_t b = blockIdx.x / tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * tiles) * kFirTile;
  const float* xb = x + b * L;
  for (int k = threadIdx.x; k < K; k += kTdThreads) s_h[k] = __ldg(h + k);
  for (int i = threadIdx.x; i < kFirTile + K - 1; i += kTdThreads) {
    const int64_t g = t0 + i - pad;
    s_x[i] = (g >= 0 && g < L) ? __ldg(xb + g) : 0.f;
  }
  __syncthreads();
  // thread -> outputs t0 + tid + {0, 256, 512, 768}: consecutive lanes read consecutive words
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  const float* w = s_x + threadIdx.x;
  for (int k = 0; k < K; ++k) {
    const float hk = s_h[k];
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[j] = fmaf(hk, w[k + j * kTdThreads], acc[j]);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int64_t t = t0 + threadIdx.x + j * kTdThreads;
    if (t < L) out[b * L + t] = acc[j];
  }
}

int launch_fir_same(const float* x, int64_t B, int64_t L, const float* h, int K, int pad, float* out,
                    cudaStream_t stream) {
  if (B == 0 || L == 0) return 0;
  if (K < 1 || K > kFirMaxTaps || pad < 0 || pad >= K) return SBFE_E_INVAL;
  const int64_t tiles = (L + kFirTile - 1) / kFirTile;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  const size_t smem = (size_t)(2 * K + kFirTile - 1) * sizeof(float);
  return launch_pdl(fir_same_kernel, dim3((unsigned)(tiles * B)), dim3(kTdThreads), smem, stream, x, L,
                    h, K, pad, out, (unsigned)tiles);
}

// ---- DoClip: out = clamp(x, -v, v) ---------------------------------------------------------------
__global__ void __launch_bounds__(kTdThreads)
clip_kernel(const float* __restrict__ x, int64_t n, float v, float* __restrict__ out) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t stride = (int64_t)gridDim.x * kTdThreads;
  int64_t i = (int64_t)blockIdx.x * kTdThreads + threadIdx.x;
  if ((n & 3) == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
    for (; i < n / 4; i += stride) {
      float4 a = __ldg(reinterpret_cast<const float4*>(x) + i);
      a.x = fminf(fmaxf(a.x, -v), v);   // torch.clamp: max with the lower bound first, NaN kept
      a.y = fminf(fmaxf(a.y, -v), v);
      a.z = fminf(fmaxf(a.z, -v), v);
      a.w = fminf(fmaxf(a.w, -v), v);
      reinterpret_cast<float4*>(out)[i] = a;
    }
  } else {
    for (; i < n; i += stride) out[i] = fminf(fmaxf(__ldg(x + i), -v), v);
  }
}

int launch_clip(const float* x, int64_t n, float v, float* out, cudaStream_t stream) {
  if (n == 0) return 0;
  int64_t blocks = (n / 4 + kTdThreads - 1) / kTdThreads + 1;
  if (blocks > 148 * 16) blocks = 148 * 16;
  return launch_pdl(clip_kernel, dim3((unsigned)blocks), dim3(kTdThreads), 0, stream, x, n, v, out);
}

// ---- Resample: polyphase band-limited interpolation (Kaldi's LinearResample) --------------------
// Output sample o = n * P + i (phase i < P) is sum_k w[i][k] x[first[i] + n * S + k], x zero outside
// [0, L): one thread per output sample instead of, per phase, slice + pad + conv1d(stride S) +
// conv_transpose1d(stride P) + pad + add into a zero tensor (speech_augmentation.py:600-665).
__global__ void __launch_bounds__(kTdThreads)
resample_kernel(const float* __restrict__ x, int64_t L, const float* __restrict__ w,
                const int32_t* __restrict__ first, int P, int S, int W, int64_t L_out,
                float* __restrict__ out) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ float s_rs[];                    // [P][W] weights | [P] first indices
  int* s_first = reinterpret_cast<int*>(s_rs + P * W);
  for (int i = threadIdx.x; i < P * W; i += kTdThreads) s_rs[i] = __ldg(w + i);
  for (int i = threadIdx.x; i < P; i += kTdThreads) s_first[i] = __ldg(first + i);
  __syncthreads();
  const int64_t b = blockIdx.y;
  const float* xb = x + b * L;
  for (int64_t o = (int64_t)blockIdx.x * kTdThreads + threadIdx.x; o < L_out;
       o += (int64_t)gridDim.x * kTdThreads) {
    const int64_t n = o / P;
    const int i = (int)(o - n * P);
    const int64_t base = (int64_t)s_first[i] + n * S;
    const float* wi = s_rs + i * W;
    float acc = 0.f;
    if (base >= 0 && base + W <= L) {
      for (int k = 0; k < W; ++k) acc = fmaf(wi[k], __ldg(xb + base + k), acc);
    } else {
      for (int k = 0; k < W; ++k) {
        const int64_t g = base + k;
        if (g >= 0 && g < L) acc = fmaf(wi[k], __ldg(xb + g), acc);
      }
    }
    out[b * L_out + o] = acc;
  }
}

int launch_resample(const float* x, int64_t B, int64_t L, const float* w, const int32_t* first, int P,
                    int S, int W, int64_t L_out, float* out, cudaStream_t stream) {
  if (B == 0 || L_out == 0) return 0;
  if (P < 1 || S < 1 || W < 1 || B > 65535) return SBFE_E_INVAL;
  const size_t smem = (size_t)P * W * sizeof(float) + (size_t)P * sizeof(int);
  if (smem > 48 * 1024) return SBFE_E_TOOBIG;
  int64_t bx = (L_out + kTdThreads - 1) / kTdThreads;
  if (bx > 148 * 8) bx = 148 * 8;
  return launch_pdl(resample_kernel, dim3((unsigned)bx, (unsigned)B), dim3(kTdThreads), smem, stream, x,
                    L, w, first, P, S, W, L_out, out);
}

// ---- DropFreq's composite band-reject filter on the device (speech_augmentation.py:977-1001 +
// notch_filter, signal_processing.py:373-430): the reference builds it with ~40 small ATen ops per
// call; here one CTA does it from the drawn frequencies (kernel arguments: no copy).
// Per frequency f: low-pass = sinc(3 f x) * blackman, normalised to sum 1; high-pass =
// sinc(3 (f + 2 w) x) * blackman, normalised to sum -1, + 1 at the centre; notch = their sum; the
// running filter is cross-correlated with it (conv1d, zero padding = half the length).
constexpr int kNotchMaxFreq = 16;
constexpr int kNotchMaxLen = 255;
struct NotchFreqs { float f[kNotchMaxFreq]; };

__global__ void __launch_bounds__(256)
notch_taps_kernel(const NotchFreqs fr, int n, float width, int len, float* __restrict__ taps) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float s_cur[kNotchMaxLen], s_notch[kNotchMaxLen], s_lo[kNotchMaxLen], s_hi[kNotchMaxLen];
  __shared__ float s_sum[2];
  const int t = threadIdx.x, pad = len / 2;
  if (t < len) s_cur[t] = t == pad ? 1.f : 0.f;
  for (int j = 0; j < n; ++j) {
    const float f = fr.f[j] + width;                 // notch_freq += notch_width
    if (t < len) {
      const float x = (float)(t - pad);
      const float w = 0.42f - 0.5f * cosf(6.283185307179586f * (float)t / (float)len) +
                      0.08f * cosf(12.566370614359172f * (float)t / (float)len);   // periodic blackman
      const float a = 3.f * (f - width) * x, b = 3.f * (f + width) * x;
      s_lo[t] = (t == pad ? 1.f : sinf(a) / a) * w;
      s_hi[t] = (t == pad ? 1.f : sinf(b) / b) * w;
    }
    __syncthreads();
    if (t < 2) {                                      // fixed order: deterministic
      const float* v = t == 0 ? s_lo : s_hi;
      float acc = 0.f;
      for (int i = 0; i < len; ++i) acc += v[i];
      s_sum[t] = acc;
    }
    __syncthreads();
    if (t < len) s_notch[t] = s_lo[t] / s_sum[0] + (s_hi[t] / -s_sum[1] + (t == pad ? 1.f : 0.f));
    __syncthreads();
    float acc = 0.f;
    if (t < len)
      for (int k = 0; k < len; ++k) {
        const int i = t + k - pad;
        if (i >= 0 && i < len) acc = fmaf(s_cur[i], s_notch[k], acc);
      }
    __syncthreads();
    if (t < len) s_cur[t] = acc;
    __syncthreads();
  }
  if (t < len) taps[t] = s_cur[t];
}

int launch_notch_taps(const float* h_freqs, int n, float width, int len, float* taps, cudaStream_t stream) {
  if (n < 0 || n > kNotchMaxFreq || len < 1 || len > kNotchMaxLen || (len & 1) == 0) return SBFE_E_INVAL;
  NotchFreqs fr = {};
  for (int j = 0; j < n; ++j) fr.f[j] = h_freqs[j];
  return launch_pdl(notch_taps_kernel, dim3(1), dim3(256), 0, stream, fr, n, width, len, taps);
}

// ---- AddNoise with white noise (speech_augmentation.py:126-168) --------------------------------
// compute_amplitude (signal_processing.py:51-56) sums |x| over the WHOLE padded row and divides by
// the valid length; partial sums of kAbsChunk samples, added in a fixed order by the consumer
// (deterministic: no float atomics).
constexpr int kAbsChunk = 16384;

__global__ void __launch_bounds__(kTdThreads)
abs_sums_kernel(const float* __restrict__ x, int64_t L, float* __restrict__ partials, int chunks) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.y;
  const int64_t lo = (int64_t)blockIdx.x * kAbsChunk;
  const int64_t hi = lo + kAbsChunk < L ? lo + kAbsChunk : L;
  const float* xb = x + b * L;
  float acc = 0.f;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kTdThreads) acc += fabsf(__ldg(xb + i));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float s_part[kTdThreads / 32];
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kTdThreads / 32; ++w) t += s_part[w];
    partials[b * chunks + blockIdx.x] = t;
  }
}

// out = x * (1 - f) + (f * amp) * noise, three roundings like the reference's eager ops
// (noisy *= 1 - f; noisy += new_noise_amplitude * white_noise), f = 1 / (10^(snr / 20) + 1),
// amp = sum|x| / (lengths * L); absmax[b] (uint bits of max |out|, zeroed by the caller) for
// normalize=True
__global__ void __launch_bounds__(kTdThreads)
add_noise_kernel(const float* __restrict__ x, const float* __restrict__ noise, int64_t L,
                 const float* __restrict__ partials, int chunks, const float* __restrict__ lengths,
                 const float* __restrict__ snr, float* __restrict__ out, unsigned* absmax) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.y;
  __shared__ float s_keep, s_namp;
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int c = 0; c < chunks; ++c) t += partials[b * chunks + c];
    const float amp = t / (__ldg(lengths + b) * (float)L);
    const float f = 1.f / (powf(10.f, __ldg(snr + b) / 20.f) + 1.f);
    s_keep = 1.f - f;
    s_namp = f * amp;
  }
  __syncthreads();
  const float keep = s_keep, namp = s_namp;
  const float* xb = x + b * L;
  const float* nb = noise + b * L;
  float* ob = out + b * L;
  float mx = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * kTdThreads + threadIdx.x; i < L; i += (int64_t)gridDim.x * kTdThreads) {
    const float v = __fadd_rn(__fmul_rn(__ldg(xb + i), keep), __fmul_rn(namp, __ldg(nb + i)));
    ob[i] = v;
    mx = fmaxf(mx, fabsf(v));
  }
  if (absmax != nullptr) {
    const unsigned k = __reduce_max_sync(0xffffffffu, __float_as_uint(mx));   // non-negative floats order as uints
    if ((threadIdx.x & 31) == 0 && k) atomicMax(absmax + b, k);
  }
}

// normalize=True (:164-168): x / clamp(absmax, min = 1)
__global__ void __launch_bounds__(kTdThreads)
peak_normalize_kernel(float* x, int64_t L, const unsigned* __restrict__ absmax) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.y;
  const float d = fmaxf(__uint_as_float(__ldg(absmax + b)), 1.f);
  if (d == 1.f) return;
  float* xb = x + b * L;
  for (int64_t i = (int64_t)blockIdx.x * kTdThreads + threadIdx.x; i < L; i += (int64_t)gridDim.x * kTdThreads)
    xb[i] = xb[i] / d;
}

int64_t abs_chunks(int64_t L) { return (L + kAbsChunk - 1) / kAbsChunk; }

int launch_abs_sums(const float* x, int64_t B, int64_t L, float* partials, cudaStream_t stream) {
  if (B == 0 || L == 0) return 0;
  const int64_t chunks = abs_chunks(L);
  if (B > 65535 || chunks > 0x7fffffff) return SBFE_E_TOOBIG;
  return launch_pdl(abs_sums_kernel, dim3((unsigned)chunks, (unsigned)B), dim3(kTdThreads), 0, stream, x, L,
                    partials, (int)chunks);
}

int launch_add_noise(const float* x, const float* noise, int64_t B, int64_t L, const float* partials,
                     const float* lengths, const float* snr, float* out, unsigned* absmax,
                     cudaStream_t stream) {
  if (B == 0 || L == 0) return 0;
  const int64_t chunks = abs_chunks(L);
  if (B > 65535 || chunks > 0x7fffffff) return SBFE_E_TOOBIG;
  int64_t bx = (L + 4 * kTdThreads - 1) / (4 * kTdThreads);
  if (bx > 148 * 8) bx = 148 * 8;
  return launch_pdl(add_noise_kernel, dim3((unsigned)bx, (unsigned)B), dim3(kTdThreads), 0, stream, x, noise,
                    L, partials, (int)chunks, lengths, snr, out, absmax);
}

int launch_peak_normalize(float* x, int64_t B, int64_t L, const unsigned* absmax, cudaStream_t stream) {
  if (B == 0 || L == 0) return 0;
  if (B > 65535) return SBFE_E_TOOBIG;
  int64_t bx = (L + 4 * kTdThreads - 1) / (4 * kTdThreads);
  if (bx > 148 * 8) bx = 148 * 8;
  return launch_pdl(peak_normalize_kernel, dim3((unsigned)bx, (unsigned)B), dim3(kTdThreads), 0, stream, x, L,
                    absmax);
}


// ---- AddBabble (speech_augmentation.py:861-906): the batch's own signals as babble noise --------
// babble[b] = x[b-1] + x[b-2] + ... + x[b-S] (rows modulo B, added in that order like the
// reference's in-place roll-and-add), scaled to the drawn SNR against x[b].  Two passes: the sums
// of |babble| per 16384 samples (fixed-order partials, like abs_sums_kernel), then the mix, which
// recomputes the babble from the S rows (the batch sits in L2).
__device__ __forceinline__ float babble_at(const float* __restrict__ x, int64_t B, int64_t L, int64_t b,
                                           int S, int64_t i) {
  int64_t r = b - 1 < 0 ? b - 1 + B : b - 1;
  float acc = __ldg(x + r * L + i);
  for (int s = 1; s < S; ++s) {
    r = r - 1 < 0 ? r - 1 + B : r - 1;
    acc = __fadd_rn(acc, __ldg(x + r * L + i));
  }
  return acc;
}

__global__ void __launch_bounds__(kTdThreads)
babble_abs_sums_kernel(const float* __restrict__ x, int64_t B, int64_t L, int S,
                       float* __restrict__ partials, int chunks) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.y;
  const int64_t lo = (int64_t)blockIdx.x * kAbsChunk;
  const int64_t hi = lo + kAbsChunk < L ? lo + kAbsChunk : L;
  float acc = 0.f;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kTdThreads) acc += fabsf(babble_at(x, B, L, b, S, i));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float s_part[kTdThreads / 32];
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kTdThreads / 32; ++w) t += s_part[w];
    partials[b * chunks + blockIdx.x] = t;
  }
}

// out = x * (1 - f) + babble * (f * amp / (bamp + 1e-14)): amp = sum|x[b]| / (len[b] L),
// bamp = sum|babble[b]| / max_{s = 1..S} (len[b - s] L), f = 1 / (10^(snr / 20) + 1)
__global__ void __launch_bounds__(kTdThreads)
add_babble_kernel(const float* __restrict__ x, int64_t B, int64_t L, int S,
                  const float* __restrict__ clean_partials, const float* __restrict__ babble_partials,
                  int chunks, const float* __restrict__ lengths, const float* __restrict__ snr,
                  float* __restrict__ out) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.y;
  __shared__ float s_keep, s_scale;
  if (threadIdx.x == 0) {
    float tc = 0.f, tb = 0.f;
    for (int c = 0; c < chunks; ++c) {
      tc += clean_partials[b * chunks + c];
      tb += babble_partials[b * chunks + c];
    }
    const float amp = tc / (__ldg(lengths + b) * (float)L);
    float blen = 0.f;
    int64_t r = b;
    for (int s = 0; s < S; ++s) {
      r = r - 1 < 0 ? r - 1 + B : r - 1;
      blen = fmaxf(blen, __ldg(lengths + r) * (float)L);
    }
    const float bamp = tb / blen;
    const float f = 1.f / (powf(10.f, __ldg(snr + b) / 20.f) + 1.f);
    s_keep = 1.f - f;
    s_scale = (f * amp) / (bamp + 1e-14f);
  }
  __syncthreads();
  const float keep = s_keep, scale = s_scale;
  const float* xb = x + b * L;
  float* ob = out + b * L;
  for (int64_t i = (int64_t)blockIdx.x * kTdThreads + threadIdx.x; i < L; i += (int64_t)gridDim.x * kTdThreads)
    ob[i] = __fadd_rn(__fmul_rn(__ldg(xb + i), keep), __fmul_rn(babble_at(x, B, L, b, S, i), scale));
}

int launch_add_babble(const float* x, int64_t B, int64_t L, int S, float* clean_partials,
                      float* babble_partials, const float* lengths, const float* snr, float* out,
                      cudaStream_t stream) {
  if (B == 0 || L == 0) return 0;
  const int64_t chunks = abs_chunks(L);
  if (B > 65535 || chunks > 0x7fffffff || S < 1) return SBFE_E_INVAL;
  int rc = launch_abs_sums(x, B, L, clean_partials, stream);
  if (rc) return rc;
  rc = launch_pdl(babble_abs_sums_kernel, dim3((unsigned)chunks, (unsigned)B), dim3(kTdThreads), 0, stream, x,
                  B, L, S, babble_partials, (int)chunks);
  if (rc) return rc;
  int64_t bx = (L + 4 * kTdThreads - 1) / (4 * kTdThreads);
  if (bx > 148 * 8) bx = 148 * 8;
  return launch_pdl(add_babble_kernel, dim3((unsigned)bx, (unsigned)B), dim3(kTdThreads), 0, stream, x, B, L, S,
                    (const float*)clean_partials, (const float*)babble_partials, (int)chunks, lengths, snr,
                    out);
}

}  // namespace sbfe
