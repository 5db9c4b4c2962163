// Post stage: top_db clamp -> DCT -> deltas (x2, concatenated) -> context window,
// fused in one pass so that only the final [B, T, F_out] tensor is written.
//   _amplitude_to_DB clamp  speechbrain/processing/features.py:706-710
//   DCT.forward             speechbrain/processing/features.py:764-786
//   Deltas.forward          speechbrain/processing/features.py:820-852
//   ContextWindow.forward   speechbrain/processing/features.py:897-937
//   composition             speechbrain/lobes/features.py:141-146, 274-280
#include "common.cuh"

namespace sbfe {

// ---- pure clamp (Fbank without deltas/context), float4-vectorised -----------
__global__ void clamp_kernel(const float* __restrict__ x, int64_t per_utt,
                             const sbfe_maxkey_t* __restrict__ utt_max, float top_db,
                             float* __restrict__ out, unsigned chunks) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t b = blockIdx.x / chunks;
  const int64_t c = blockIdx.x - b * chunks;
  const float floor_b = clamp_floor(utt_max, b, top_db);
  const float* src = x + b * per_utt;
  float* dst = out + b * per_utt;
  const int64_t chunk = (per_utt + chunks - 1) / chunks;
  const int64_t lo = c * chunk, hi = min(per_utt, lo + chunk);
  const bool vec = ((reinterpret_cast<uintptr_t>(src + lo) | reinterpret_cast<uintptr_t>(dst + lo)) & 15) == 0;
  if (vec) {
    const int64_t n4 = (hi - lo) / 4;
    const float4* s4 = reinterpret_cast<const float4*>(src + lo);
    float4* d4 = reinterpret_cast<float4*>(dst + lo);
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
      float4 v = s4[i];
      v.x = fmaxf(v.x, floor_b); v.y = fmaxf(v.y, floor_b);
      v.z = fmaxf(v.z, floor_b); v.w = fmaxf(v.w, floor_b);
      d4[i] = v;
    }
    for (int64_t i = lo + n4 * 4 + threadIdx.x; i < hi; i += blockDim.x)
      dst[i] = fmaxf(src[i], floor_b);
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) dst[i] = fmaxf(src[i], floor_b);
  }
}

struct PostParams {
  const float* x;
  float* out;
  const sbfe_maxkey_t* utt_max;
  const float* dct;  // [F_in][F0] or null
  int64_t T;
  float top_db, inv_denom;
  int F_in, F0;      // F0 = features after the optional DCT
  int deltas, n;     // deltas: 0 none, 1 [x, d, dd], 2 d only;  n = (window-1)/2
  int context, left, right;
  int tile;          // frames per CTA
  unsigned tiles;
};

constexpr int kPostThreads = 256;

__global__ void __launch_bounds__(kPostThreads) post_kernel(const PostParams p) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ __align__(16) float smem[];
  const int n2 = p.deltas == 1 ? 2 * p.n : (p.deltas ? p.n : 0);  // base-feature halo per side
  const int n1 = p.deltas == 1 ? p.n : 0;                         // delta halo per side
  const int ctxl = p.context ? p.left : 0, ctxr = p.context ? p.right : 0;
  const int C = ctxl + ctxr + 1;
  const int fr_out = p.tile + ctxl + ctxr;     // frames whose features feed the outputs
  const int fr_d = fr_out + 2 * n1;            // delta frames staged
  const int fr_b = fr_out + 2 * n2;            // base frames staged
  float* s_base = smem;                        // [fr_b][F0]
  float* s_d = s_base + fr_b * p.F0;           // [fr_d][F0]   (deltas only)
  float* s_dd = s_d + (p.deltas ? fr_d * p.F0 : 0);   // [fr_out][F0]  (deltas == 1 only)
  float* s_in = s_dd + (p.deltas == 1 ? fr_out * p.F0 : 0);  // [fr_b][F_in] (DCT only)
  float* s_dct = s_in + (p.dct ? fr_b * p.F_in : 0);    // [F_in][F0]

  const int64_t b = blockIdx.x / p.tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * p.tiles) * p.tile;
  const int64_t T = p.T;
  const float floor_b = clamp_floor(p.utt_max, b, p.top_db);
  const float* xb = p.x + b * T * p.F_in;
  const int64_t base_lo = t0 - ctxl - n2;      // unclamped frame index of s_base row 0
  const int64_t d_lo = t0 - ctxl - n1;
  const int64_t o_lo = t0 - ctxl;

  // stage clamped input rows (replicate-clamped frame index)
  float* s_stage = p.dct ? s_in : s_base;
  for (int i = threadIdx.x; i < fr_b * p.F_in; i += kPostThreads) {
    const int r = i / p.F_in, f = i - r * p.F_in;
    int64_t u = base_lo + r;
    u = u < 0 ? 0 : (u > T - 1 ? T - 1 : u);
    s_stage[i] = fmaxf(__ldg(xb + u * p.F_in + f), floor_b);
  }
  if (p.dct)
    for (int i = threadIdx.x; i < p.F_in * p.F0; i += kPostThreads) s_dct[i] = __ldg(p.dct + i);
  __syncthreads();
  if (p.dct) {
    for (int i = threadIdx.x; i < fr_b * p.F0; i += kPostThreads) {
      const int r = i / p.F0, j = i - r * p.F0;
      const float* row = s_in + r * p.F_in;
      float acc = 0.f;
      for (int m = 0; m < p.F_in; ++m) acc = fmaf(row[m], s_dct[m * p.F0 + j], acc);
      s_base[i] = acc;
    }
    __syncthreads();
  }
  if (p.deltas) {
    // delta at frame v = clamp(u): sum_k k * base[clamp(v + k)] / denom
    for (int i = threadIdx.x; i < fr_d * p.F0; i += kPostThreads) {
      const int r = i / p.F0, f = i - r * p.F0;
      int64_t v = d_lo + r;
      v = v < 0 ? 0 : (v > T - 1 ? T - 1 : v);
      const float* col = s_base + (v - base_lo) * p.F0 + f;
      float acc = 0.f;
      for (int k = 1; k <= p.n; ++k) acc = fmaf((float)k, col[k * p.F0] - col[-k * p.F0], acc);
      s_d[i] = acc * p.inv_denom;
    }
    __syncthreads();
    if (p.deltas == 1)
    for (int i = threadIdx.x; i < fr_out * p.F0; i += kPostThreads) {
      const int r = i / p.F0, f = i - r * p.F0;
      int64_t v = o_lo + r;
      v = v < 0 ? 0 : (v > T - 1 ? T - 1 : v);
      const float* col = s_d + (v - d_lo) * p.F0 + f;
      float acc = 0.f;
      for (int k = 1; k <= p.n; ++k) acc = fmaf((float)k, col[k * p.F0] - col[-k * p.F0], acc);
      s_dd[i] = acc * p.inv_denom;
    }
    __syncthreads();
  }
  // gather: out[t, f*C + c] = feat[t + c - left][f], zero outside [0, T)
  const int nfeat = p.F0 * (p.deltas == 1 ? 3 : 1);
  const int F_out = nfeat * C;
  const int64_t left_frames = T - t0;
  const int nvalid = left_frames < p.tile ? (int)left_frames : p.tile;
  float* ob = p.out + (b * T + t0) * F_out;
  for (int i = threadIdx.x; i < nvalid * F_out; i += kPostThreads) {
    const int t = i / F_out, j = i - t * F_out;
    const int f = j / C, c = j - f * C;
    const int64_t u = t0 + t + c - ctxl;
    float v = 0.f;
    if (u >= 0 && u < T) {
      const int r = t + c;  // row relative to o_lo
      if (p.deltas == 2) v = s_d[(r + n1) * p.F0 + f];
      else if (f < p.F0) v = s_base[(r + n2) * p.F0 + f];
      else if (f < 2 * p.F0) v = s_d[(r + n1) * p.F0 + (f - p.F0)];
      else v = s_dd[r * p.F0 + (f - 2 * p.F0)];
    }
    ob[i] = v;
  }
}

// ---- fast variant: same stages, but the features of a frame are laid out as one row
// [base | delta | delta-delta] and the context gather writes 128-bit vectors whose four
// (feature, context) sources are resolved ONCE per thread instead of two integer divisions per
// output element (the generic kernel above is instruction-bound on exactly that: 660 outputs
// per frame for the MFCC default).  Requires F_out % 4 == 0.
// Template arguments fix the configuration at compile time (0 / -1: taken from PostParams at run
// time): with the MFCC default <23, 20, 2, 5, 5, 1, 1, 32> every loop below unrolls and every
// row / column index is a constant - the preparatory phases (staging, DCT, deltas, feature
// rows) were 73 % of the generic kernel's instructions although they touch 3 % of its bytes.
// Row pitch of the [frames][features] staging rows the context gather reads: the smallest
// P >= nfeat with P % 4 == 2.  A lane's q-th source is (c, f) = ((4v + q) % C, (4v + q) / C) at
// c * P + f; counted over all warps of the gather this pitch needs 1.5 - 2.0 shared-memory
// wavefronts per load where an odd pitch needs 2.5 - 5.6 (MFCC default: 1.83 against 2.67).
__host__ __device__ inline int post_feat_pitch(int nfeat) { return nfeat + ((6 - (nfeat & 3)) & 3); }

template <int FIN, int F0C, int NC, int LC, int RC, int DELTAS, int DCTC, int TILE>
__global__ void post_fast_kernel_t(const PostParams p) {
  const int F_in = FIN > 0 ? FIN : p.F_in;
  const int F0 = F0C > 0 ? F0C : p.F0;
  const int n = NC > 0 ? NC : p.n;
  const int deltas = DELTAS >= 0 ? DELTAS : p.deltas;
  const bool dct = DCTC >= 0 ? DCTC != 0 : p.dct != nullptr;
  const int tile = TILE > 0 ? TILE : p.tile;
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ __align__(16) float smem[];
  const int nthr = blockDim.x;
  const int n2 = deltas == 1 ? 2 * n : (deltas ? n : 0);
  const int n1 = deltas == 1 ? n : 0;
  const int ctxl = LC >= 0 ? LC : (p.context ? p.left : 0), ctxr = RC >= 0 ? RC : (p.context ? p.right : 0);
  const int C = ctxl + ctxr + 1;
  const int fr_out = tile + ctxl + ctxr;
  const int fr_d = fr_out + 2 * n1;
  const int fr_b = fr_out + 2 * n2;
  const int nfeat = F0 * (deltas == 1 ? 3 : 1);
  const int P = post_feat_pitch(nfeat);                  // row pitch of the feature rows
  float* s_dct = smem;                                   // [F_in][F0] first: read as float4
  float* s_feat = s_dct + (dct ? F_in * F0 : 0);         // [fr_out][P]
  float* s_base = s_feat + fr_out * P;                   // [fr_b][F0]
  float* s_d = s_base + fr_b * F0;                     // [fr_d][F0]   (deltas only)
  float* s_in = s_d + (deltas ? fr_d * F0 : 0);      // [fr_b][F_in] (DCT only)

  const int64_t b = blockIdx.x / p.tiles;
  const int64_t t0 = (int64_t)(blockIdx.x - b * p.tiles) * tile;
  const int64_t T = p.T;
  const float floor_b = clamp_floor(p.utt_max, b, p.top_db);
  const float* xb = p.x + b * T * F_in;
  const int64_t base_lo = t0 - ctxl - n2;
  const int64_t d_lo = t0 - ctxl - n1;
  const int64_t o_lo = t0 - ctxl;

  // staging loops walk (row, feature) incrementally from the thread's first element: one integer
  // division per thread and phase instead of one per element, 32-bit row arithmetic
  const int Ti = (int)T, base_lo_i = (int)base_lo, d_lo_i = (int)d_lo, o_lo_i = (int)o_lo;
  float* s_stage = dct ? s_in : s_base;
  // the DCT matrix rides in the same memory round trip as the input rows
  float dv[2] = {0.f, 0.f};
  if (dct) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
      if (threadIdx.x + k * nthr < F_in * F0) dv[k] = __ldg(p.dct + threadIdx.x + k * nthr);
  }
  if (base_lo_i >= 0 && base_lo_i + fr_b <= Ti) {
    // interior tile: the fr_b input rows are one contiguous run of the utterance - a flat copy,
    // all of a thread's loads in flight before its first store
    const float* src = xb + (int64_t)base_lo_i * F_in;
    const int total = fr_b * F_in;
    constexpr int kFlight = 4;
    for (int i0 = threadIdx.x; i0 < total; i0 += kFlight * nthr) {
      float v[kFlight];
#pragma unroll
      for (int k = 0; k < kFlight; ++k)
        if (i0 + k * nthr < total) v[k] = __ldg(src + i0 + k * nthr);
#pragma unroll
      for (int k = 0; k < kFlight; ++k)
        if (i0 + k * nthr < total) s_stage[i0 + k * nthr] = fmaxf(v[k], floor_b);
    }
  } else {
    int r = threadIdx.x / F_in, f = threadIdx.x - r * F_in;
    const int dr = nthr / F_in, df = nthr - dr * F_in;
    while (r < fr_b) {
      int u = base_lo_i + r;
      u = u < 0 ? 0 : (u > Ti - 1 ? Ti - 1 : u);
      s_stage[r * F_in + f] = fmaxf(__ldg(xb + (int64_t)u * F_in + f), floor_b);
      r += dr; f += df;
      if (f >= F_in) { f -= F_in; ++r; }
    }
  }
  if (dct) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
      if (threadIdx.x + k * nthr < F_in * F0) s_dct[threadIdx.x + k * nthr] = dv[k];
    for (int i = threadIdx.x + 2 * nthr; i < F_in * F0; i += nthr) s_dct[i] = __ldg(p.dct + i);
  }
  __syncthreads();
  if (dct) {
    if ((F0 & 3) == 0) {
      // thread = (row, 4 coefficients): one row value + one 128-bit matrix load per 4 FMAs
      const int jg = F0 >> 2;
      for (int i = threadIdx.x; i < fr_b * jg; i += nthr) {
        const int r = i / jg, j4 = (i - r * jg) * 4;
        const float* row = s_in + r * F_in;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int m = 0; m < F_in; ++m) {
          const float xv = row[m];
          const float4 w = *reinterpret_cast<const float4*>(s_dct + m * F0 + j4);
          acc.x = fmaf(xv, w.x, acc.x); acc.y = fmaf(xv, w.y, acc.y);
          acc.z = fmaf(xv, w.z, acc.z); acc.w = fmaf(xv, w.w, acc.w);
        }
        float* dst = s_base + r * F0 + j4;
        dst[0] = acc.x; dst[1] = acc.y; dst[2] = acc.z; dst[3] = acc.w;
      }
    } else {
      const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarps = nthr >> 5;
      for (int r = wid; r < fr_b; r += nwarps) {
        const float* row = s_in + r * F_in;
        for (int j = lane; j < F0; j += 32) {
          float acc = 0.f;
          for (int m = 0; m < F_in; ++m) acc = fmaf(row[m], s_dct[m * F0 + j], acc);
          s_base[r * F0 + j] = acc;
        }
      }
    }
    __syncthreads();
  }
  const int r_first = threadIdx.x / F0, f_first = threadIdx.x - r_first * F0;
  const int dr0 = nthr / F0, df0 = nthr - dr0 * F0;
  if (deltas) {
    int r = r_first, f = f_first;
    while (r < fr_d) {
      int v = d_lo_i + r;
      v = v < 0 ? 0 : (v > Ti - 1 ? Ti - 1 : v);
      const float* col = s_base + (v - base_lo_i) * F0 + f;
      float acc = 0.f;
      for (int k = 1; k <= n; ++k) acc = fmaf((float)k, col[k * F0] - col[-k * F0], acc);
      s_d[r * F0 + f] = acc * p.inv_denom;
      r += dr0; f += df0;
      if (f >= F0) { f -= F0; ++r; }
    }
    __syncthreads();
  }
  // feature rows [base | d | dd] (deltas == 1), [d] (deltas == 2) or [base]
  {
    int r = r_first, f = f_first;
    while (r < fr_out) {
      float* dst = s_feat + r * P + f;
      if (deltas == 2) {
        dst[0] = s_d[(r + n1) * F0 + f];
      } else {
        dst[0] = s_base[(r + n2) * F0 + f];
        if (deltas == 1) {
          dst[F0] = s_d[(r + n1) * F0 + f];
          int v = o_lo_i + r;
          v = v < 0 ? 0 : (v > Ti - 1 ? Ti - 1 : v);
          const float* col = s_d + (v - d_lo_i) * F0 + f;
          float acc = 0.f;
          for (int k = 1; k <= n; ++k) acc = fmaf((float)k, col[k * F0] - col[-k * F0], acc);
          dst[2 * F0] = acc * p.inv_denom;
        }
      }
      r += dr0; f += df0;
      if (f >= F0) { f -= F0; ++r; }
    }
  }
  __syncthreads();
  // gather: out[t, f*C + c] = feat[t + c - left][f], zero outside [0, T)
  const int F_out = nfeat * C, nvec = F_out >> 2;
  const int64_t left_frames = T - t0;
  const int nvalid = left_frames < tile ? (int)left_frames : tile;
  const bool interior = o_lo >= 0 && t0 + nvalid - 1 + ctxr <= T - 1;
  float* ob = p.out + (b * T + t0) * F_out;
  const int groups = nthr / nvec > 0 ? nthr / nvec : 1;   // row groups when nvec is small
  const int g = threadIdx.x / nvec;
  for (int v = threadIdx.x - g * nvec; v < nvec && g < groups; v += nthr) {
    int off[4], cc[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int j = 4 * v + q, f = j / C, c = j - f * C;
      off[q] = c * P + f;
      cc[q] = c;
    }
    // four source pointers and the destination advance by constants: no multiplies in the loop
    const float* s0 = s_feat + g * P + off[0];
    const float* s1 = s_feat + g * P + off[1];
    const float* s2 = s_feat + g * P + off[2];
    const float* s3 = s_feat + g * P + off[3];
    float* o = ob + (int64_t)g * F_out + 4 * v;
    const int sstep = groups * P;
    const int64_t ostep = (int64_t)groups * F_out;
    if (interior) {
#pragma unroll 4
      for (int t = g; t < nvalid; t += groups) {
        *reinterpret_cast<float4*>(o) = make_float4(*s0, *s1, *s2, *s3);
        s0 += sstep; s1 += sstep; s2 += sstep; s3 += sstep;
        o += ostep;
      }
    } else {
      for (int t = g; t < nvalid; t += groups) {
        float4 r = make_float4(*s0, *s1, *s2, *s3);
        const int u0 = o_lo_i + t;
        if (u0 + cc[0] < 0 || u0 + cc[0] >= Ti) r.x = 0.f;
        if (u0 + cc[1] < 0 || u0 + cc[1] >= Ti) r.y = 0.f;
        if (u0 + cc[2] < 0 || u0 + cc[2] >= Ti) r.z = 0.f;
        if (u0 + cc[3] < 0 || u0 + cc[3] >= Ti) r.w = 0.f;
        *reinterpret_cast<float4*>(o) = r;
        s0 += sstep; s1 += sstep; s2 += sstep; s3 += sstep;
        o += ostep;
      }
    }
  }
}

constexpr auto post_fast_kernel = post_fast_kernel_t<0, 0, 0, -1, -1, -1, -1, 0>;

int launch_post(const float* x, int64_t B, int64_t T, int F_in, const sbfe_maxkey_t* utt_max,
                float top_db, const float* dct, int n_dct, int deltas, int delta_window,
                int context, int left, int right, float* out, cudaStream_t stream) {
  if (B == 0 || T == 0) return 0;
  if (!context) left = right = 0;
  if (!dct && !deltas && !(left || right)) {
    const int64_t per_utt = T * F_in;
    int64_t chunks = (per_utt + 8191) / 8192;
    if (chunks * B > 0x7fffffff) return SBFE_E_TOOBIG;
    return launch_pdl(clamp_kernel, dim3((unsigned)(chunks * B)), dim3(256), 0, stream, x, per_utt,
                      utt_max, top_db, out, (unsigned)chunks);
  }
  if (out == x) return SBFE_E_INVAL;
  PostParams p;
  p.x = x; p.out = out; p.utt_max = utt_max; p.dct = dct; p.T = T; p.top_db = top_db;
  p.F_in = F_in; p.F0 = dct ? n_dct : F_in;
  p.deltas = deltas == 2 ? 2 : (deltas ? 1 : 0);
  p.n = (delta_window - 1) / 2;
  if (p.deltas && p.n < 1) return SBFE_E_INVAL;
  p.inv_denom = p.deltas ? 1.0f / ((float)(p.n * (p.n + 1) * (2 * p.n + 1)) / 3.0f) : 0.f;
  p.context = context ? 1 : 0; p.left = left; p.right = right;
  const int n2 = p.deltas == 1 ? 2 * p.n : (p.deltas ? p.n : 0), n1 = p.deltas == 1 ? p.n : 0;
  {  // vectorised gather (post_fast_kernel) when the output rows are whole float4s
    const int nfeat = p.F0 * (p.deltas == 1 ? 3 : 1), C = left + right + 1;
    const int F_out = nfeat * C;
    if ((F_out & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
      // 64-frame tiles halve the halo work; 32-frame tiles when that leaves too few CTAs
      for (int tile = ((T + 63) / 64) * B >= 1184 ? 64 : 32; tile >= 8; tile /= 2) {
        const int fr_out = tile + left + right, fr_d = fr_out + 2 * n1, fr_b = fr_out + 2 * n2;
        size_t fl = (size_t)fr_out * post_feat_pitch(nfeat) + (size_t)fr_b * p.F0;
        if (p.deltas) fl += (size_t)fr_d * p.F0;
        if (dct) fl += (size_t)fr_b * F_in + (size_t)F_in * p.F0;
        const size_t smem = fl * sizeof(float);
        if (smem > 100 * 1024) continue;
        p.tile = tile;
        const int64_t tiles = (T + tile - 1) / tile;
        if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
        p.tiles = (unsigned)tiles;
        if (smem > 48 * 1024) {
          cudaError_t e = cudaFuncSetAttribute(
              post_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
          if (e != cudaSuccess) return (int)e;
        }
        const int nvec = F_out / 4;
        // whole row groups of nvec threads (rounded to warps), up to 512 threads
        const int rowthr = ((nvec + 31) / 32) * 32;
        const int threads = rowthr >= 512 ? (rowthr > 1024 ? 1024 : rowthr) : rowthr * (512 / rowthr);
        // the MFCC default (23 mel -> 20 DCT, window 5, context 5 + 5) has its own instantiation
        const bool mfcc = dct && F_in == 23 && p.F0 == 20 && p.n == 2 && p.deltas == 1 && left == 5 &&
                          right == 5;
        if (mfcc && tile == 32)
          return launch_pdl(post_fast_kernel_t<23, 20, 2, 5, 5, 1, 1, 32>, dim3((unsigned)(tiles * B)),
                            dim3(threads), smem, stream, p);
        if (mfcc && tile == 64)
          return launch_pdl(post_fast_kernel_t<23, 20, 2, 5, 5, 1, 1, 64>, dim3((unsigned)(tiles * B)),
                            dim3(threads), smem, stream, p);
        return launch_pdl(post_fast_kernel, dim3((unsigned)(tiles * B)), dim3(threads), smem, stream,
                          p);
      }
    }
  }
  size_t smem = 0;
  int tile = 32;
  for (;; tile /= 2) {
    const int fr_out = tile + left + right, fr_d = fr_out + 2 * n1, fr_b = fr_out + 2 * n2;
    size_t fl = (size_t)fr_b * p.F0;
    if (p.deltas) fl += (size_t)fr_d * p.F0;
    if (p.deltas == 1) fl += (size_t)fr_out * p.F0;
    if (dct) fl += (size_t)fr_b * F_in + (size_t)F_in * p.F0;
    smem = fl * sizeof(float);
    if (smem <= 200 * 1024 || tile == 1) break;
  }
  if (smem > 200 * 1024) return SBFE_E_TOOBIG;
  p.tile = tile;
  const int64_t tiles = (T + tile - 1) / tile;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  p.tiles = (unsigned)tiles;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  return launch_pdl(post_kernel, dim3((unsigned)(tiles * B)), dim3(kPostThreads), smem, stream, p);
}

}  // namespace sbfe
