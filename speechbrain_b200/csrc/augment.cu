// SpecAugment on features (speechbrain/lobes/augment.py:32-201): time warp (two bicubic /
// bilinear resamplings of the time axis around a random centre, :116-150) followed by
// frequency and time masks (:152-201), as ONE pass over [B, T, F].
//
// The random draws (centre c, warped centre w, mask positions / lengths) are made by the
// caller with the reference's own torch.randint calls, so the module consumes the same
// generator stream; this file only moves data.  HBM-bound: with a warp every element is read
// once (4 source rows per output row, neighbours share them through L1) and written once;
// without a warp only the masked elements are written and nothing is read.
#include "common.cuh"

namespace sbfe {

constexpr int kAugThreads = 256;
constexpr int kAugRows = 48;       // output rows per CTA: the per-CTA plan (masks, row taps) is ~400 instructions, amortised
constexpr int kAugMaxMasks = 16;   // masks per axis kept in shared memory

struct AugParams {
  const float* x;
  float* out;
  int64_t T;
  int F;
  int warp_c, warp_w, warp_mode;   // warp_c < 0: no warp; mode 0 = bicubic, 1 = bilinear
  const int64_t* fpos; const int64_t* flen; int n_f;   // [B, n_f]
  const int64_t* tpos; const int64_t* tlen; int n_t;   // [B, n_t]
  const float* d_fill;             // optional device scalar (mean fill), else `fill`
  float fill;
};

// aten/src/ATen/native/UpSample.h: cubic convolution with A = -0.75
__device__ __forceinline__ float cubic1(float x) {
  const float A = -0.75f;
  return ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f;
}
__device__ __forceinline__ float cubic2(float x) {
  const float A = -0.75f;
  return ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A;
}

struct RowPlan {
  int src[4];      // absolute source rows (clamped to the segment), then scaled by F: element offsets
  float w[4];
  int masked;      // time-masked row
};

// output row t of utterance b -> source rows / weights (align_corners=True resampling of
// [0, c) onto [0, w) and of [c, T) onto [w, T))
__device__ inline void plan_row(const AugParams& p, int64_t t, RowPlan& r) {
  if (p.warp_c < 0) {
    r.src[0] = r.src[1] = r.src[2] = r.src[3] = (int)t;
    r.w[0] = 0.f; r.w[1] = 1.f; r.w[2] = 0.f; r.w[3] = 0.f;
    return;
  }
  const int c = p.warp_c, w = p.warp_w;
  int seg_lo, seg_n, out_n, o;
  if (t < w) { seg_lo = 0; seg_n = c; out_n = w; o = (int)t; }
  else { seg_lo = c; seg_n = (int)(p.T - c); out_n = (int)(p.T - w); o = (int)(t - w); }
  const float scale = out_n > 1 ? (float)(seg_n - 1) / (float)(out_n - 1) : 0.f;
  const float real = scale * (float)o;
  if (p.warp_mode == 0) {
    const float fl = floorf(real);
    const int i = (int)fl;
    const float tt = real - fl;
    r.w[0] = cubic2(tt + 1.f);
    r.w[1] = cubic1(tt);
    r.w[2] = cubic1(1.f - tt);
    r.w[3] = cubic2((1.f - tt) + 1.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int j = i - 1 + k;
      j = j < 0 ? 0 : (j > seg_n - 1 ? seg_n - 1 : j);
      r.src[k] = seg_lo + j;
    }
  } else {  // bilinear (UpSampleBilinear2d, align_corners)
    const int i = (int)real;
    const int i1 = i + (i < seg_n - 1 ? 1 : 0);
    const float l1 = real - (float)i;
    r.src[0] = r.src[3] = seg_lo + i;
    r.src[1] = seg_lo + i;
    r.src[2] = seg_lo + i1;
    r.w[0] = 0.f; r.w[1] = 1.f - l1; r.w[2] = l1; r.w[3] = 0.f;
  }
}

constexpr int kAugMaxCols = 1024;  // column groups whose mask bits are cached in shared memory

template <int VEC>
__global__ void __launch_bounds__(kAugThreads) specaug_kernel(const AugParams p) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ RowPlan s_row[kAugRows];
  __shared__ int s_m[4 * kAugMaxMasks];        // fpos | fend | tpos | tend of this utterance
  __shared__ unsigned char s_cm[kAugMaxCols];  // per column group: VEC mask bits
  const int64_t b = blockIdx.y;
  const int64_t t_lo = (int64_t)blockIdx.x * kAugRows;
  const int rows = p.T - t_lo < kAugRows ? (int)(p.T - t_lo) : kAugRows;
  const int per_row = p.F / VEC;
  if (threadIdx.x < p.n_f) {
    const int64_t a = p.fpos[b * p.n_f + threadIdx.x];
    s_m[threadIdx.x] = (int)a;
    s_m[kAugMaxMasks + threadIdx.x] = (int)(a + p.flen[b * p.n_f + threadIdx.x]);
  }
  if (threadIdx.x >= 32 && threadIdx.x - 32 < p.n_t) {
    const int i = threadIdx.x - 32;
    const int64_t a = p.tpos[b * p.n_t + i];
    s_m[2 * kAugMaxMasks + i] = (int)a;
    s_m[3 * kAugMaxMasks + i] = (int)(a + p.tlen[b * p.n_t + i]);
  }
  __syncthreads();
  if (threadIdx.x < rows) {
    RowPlan r;
    const int64_t t = t_lo + threadIdx.x;
    plan_row(p, t, r);
    int masked = 0;
    for (int i = 0; i < p.n_t; ++i)
      masked |= (s_m[2 * kAugMaxMasks + i] <= t) & (t < s_m[3 * kAugMaxMasks + i]);
    r.masked = masked;
#pragma unroll
    for (int k = 0; k < 4; ++k) r.src[k] *= p.F;   // 32-bit element offsets inside the utterance
    s_row[threadIdx.x] = r;
  }
  auto col_bits = [&](int f) {
    unsigned bits = 0;
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      bool m = false;
      for (int i = 0; i < p.n_f; ++i) m |= (s_m[i] <= f + v) & (f + v < s_m[kAugMaxMasks + i]);
      bits |= (unsigned)m << v;
    }
    return bits;
  };
  const bool cached = per_row <= kAugMaxCols;
  if (cached)
    for (int j = threadIdx.x; j < per_row; j += kAugThreads) s_cm[j] = (unsigned char)col_bits(j * VEC);
  __syncthreads();
  const float fill = p.d_fill ? __ldg(p.d_fill) : p.fill;
  const float* xb = p.x + b * p.T * p.F;
  float* ob = p.out + (b * p.T + t_lo) * p.F;
  const bool warp = p.warp_c >= 0;
  constexpr unsigned kAll = (1u << VEC) - 1u;
  // thread = (row r0 + k * step, column group j): one division per thread, none per element
  const int lanes = per_row < kAugThreads ? per_row : kAugThreads;
  const int step = kAugThreads / lanes;
  const int r0 = threadIdx.x / lanes, j0 = threadIdx.x - r0 * lanes;
  if (r0 >= step) return;
  for (int j = j0; j < per_row; j += lanes) {
    const int f = j * VEC;
    const unsigned cbits = cached ? s_cm[j] : col_bits(f);
#pragma unroll 2
    for (int rr = r0; rr < rows; rr += step) {
      const RowPlan& r = s_row[rr];
      const unsigned bits = r.masked ? kAll : cbits;
      if (!warp && bits == 0u) continue;   // mask-only mode works in place: untouched elements stay
      float y[VEC];
      if (bits == kAll) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) y[v] = fill;
      } else {
        if (VEC == 4) {
          const float4 a0 = *reinterpret_cast<const float4*>(xb + r.src[1] + f);  // may alias out
          if (warp) {
            const float4 am = __ldg(reinterpret_cast<const float4*>(xb + r.src[0] + f));
            const float4 a1 = __ldg(reinterpret_cast<const float4*>(xb + r.src[2] + f));
            const float4 a2 = __ldg(reinterpret_cast<const float4*>(xb + r.src[3] + f));
            y[0] = fmaf(a2.x, r.w[3], fmaf(a1.x, r.w[2], fmaf(a0.x, r.w[1], am.x * r.w[0])));
            y[1] = fmaf(a2.y, r.w[3], fmaf(a1.y, r.w[2], fmaf(a0.y, r.w[1], am.y * r.w[0])));
            y[2] = fmaf(a2.z, r.w[3], fmaf(a1.z, r.w[2], fmaf(a0.z, r.w[1], am.z * r.w[0])));
            y[3] = fmaf(a2.w, r.w[3], fmaf(a1.w, r.w[2], fmaf(a0.w, r.w[1], am.w * r.w[0])));
          } else {
            y[0] = a0.x; y[1] = a0.y; y[2] = a0.z; y[3] = a0.w;
          }
        } else {
#pragma unroll
          for (int v = 0; v < VEC; ++v) {
            const float a0 = xb[r.src[1] + f + v];
            if (warp) {
              const float am = __ldg(xb + r.src[0] + f + v);
              const float a1 = __ldg(xb + r.src[2] + f + v);
              const float a2 = __ldg(xb + r.src[3] + f + v);
              y[v] = fmaf(a2, r.w[3], fmaf(a1, r.w[2], fmaf(a0, r.w[1], am * r.w[0])));
            } else {
              y[v] = a0;
            }
          }
        }
#pragma unroll
        for (int v = 0; v < VEC; ++v) y[v] = ((bits >> v) & 1u) ? fill : y[v];
      }
      float* o = ob + rr * p.F + f;
      if (VEC == 4) {
        *reinterpret_cast<float4*>(o) = make_float4(y[0], y[1], y[2], y[3]);
      } else {
#pragma unroll
        for (int v = 0; v < VEC; ++v) o[v] = y[v];
      }
    }
  }
}

int launch_specaugment(const float* x, int64_t B, int64_t T, int F, int warp_c, int warp_w,
                       int warp_mode, const int64_t* fpos, const int64_t* flen, int n_f,
                       const int64_t* tpos, const int64_t* tlen, int n_t, const float* d_fill,
                       float fill, float* out, cudaStream_t stream) {
  if (B == 0 || T == 0 || F == 0) return 0;
  if (n_f < 0 || n_t < 0 || n_f > kAugMaxMasks || n_t > kAugMaxMasks) return SBFE_E_INVAL;
  if ((n_f > 0 && (!fpos || !flen)) || (n_t > 0 && (!tpos || !tlen))) return SBFE_E_INVAL;
  if (warp_c >= 0) {
    if (out == x) return SBFE_E_INVAL;   // every row is resampled: not an in-place operation
    if (warp_c < 1 || warp_c >= T || warp_w < 1 || warp_w >= T || warp_mode < 0 || warp_mode > 1)
      return SBFE_E_INVAL;
  } else if (out != x) {
    return SBFE_E_INVAL;                 // mask-only mode writes the masked elements of x
  }
  const int64_t tiles = (T + kAugRows - 1) / kAugRows;
  if (tiles > 0x7fffffff || B > 65535 || T * F > 0x7fffffff) return SBFE_E_TOOBIG;
  AugParams p;
  p.x = x; p.out = out; p.T = T; p.F = F;
  p.warp_c = warp_c; p.warp_w = warp_w; p.warp_mode = warp_mode;
  p.fpos = fpos; p.flen = flen; p.n_f = n_f;
  p.tpos = tpos; p.tlen = tlen; p.n_t = n_t;
  p.d_fill = d_fill; p.fill = fill;
  const bool vec = (F % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  const dim3 grid((unsigned)tiles, (unsigned)B);
  return launch_pdl(vec ? specaug_kernel<4> : specaug_kernel<1>, grid, dim3(kAugThreads), 0, stream,
                    p);
}

// ---- mean of a whole tensor (replace_with_zero=False fills with x.mean(), augment.py:196) ----
// two launches, double partial sums, result as a device scalar: no host round trip
constexpr int kMeanThreads = 256;
constexpr int kMeanCtas = 296;

__global__ void __launch_bounds__(kMeanThreads) mean_partial_kernel(const float* __restrict__ x,
                                                                   int64_t n, double* part) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ double s_red[kMeanThreads];
  double s = 0.0;
  const int64_t stride = (int64_t)gridDim.x * kMeanThreads;
  for (int64_t i = (int64_t)blockIdx.x * kMeanThreads + threadIdx.x; i < n; i += stride)
    s += (double)__ldg(x + i);
  s_red[threadIdx.x] = s;
  __syncthreads();
  for (int o = kMeanThreads / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_red[threadIdx.x] += s_red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = s_red[0];
}

__global__ void mean_final_kernel(const double* part, int parts, int64_t n, float* out) {
  pdl_wait();
  pdl_launch_dependents();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < parts; ++i) s += part[i];
    *out = (float)(s / (double)n);
  }
}

int64_t mean_workspace_bytes() { return (int64_t)kMeanCtas * sizeof(double); }

int launch_mean(const float* x, int64_t n, float* d_mean, void* ws, int64_t ws_bytes,
                cudaStream_t stream) {
  if (n <= 0) return SBFE_E_INVAL;
  if (ws_bytes < mean_workspace_bytes()) return SBFE_E_NOSPACE;
  int rc = launch_pdl(mean_partial_kernel, dim3(kMeanCtas), dim3(kMeanThreads), 0, stream, x, n,
                      (double*)ws);
  if (rc) return rc;
  return launch_pdl(mean_final_kernel, dim3(1), dim3(32), 0, stream, (const double*)ws, kMeanCtas,
                    n, d_mean);
}

}  // namespace sbfe
