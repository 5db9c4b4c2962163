// InputNormalization kernels (speechbrain/processing/features.py:1000-1145).
//   stats   : per (utterance, feature) mean and unbiased std over the first
//             n_b = rint(lengths[b] * T) frames                  (:1020-1031, 1120-1145)
//   reduce  : sum over utterances -> the [2F+1] all-reduce payload (:1083-1085)
//   update  : running global statistics                          (:1090-1114)
//   apply   : (x - mean) / std over ALL T frames, padding too     (:1035, 1088, 1116)
// The reference does this with a Python loop over sentences (a host sync per
// sentence); here nothing leaves the device.
#include <atomic>

#include <cuda_bf16.h>

#include "common.cuh"

namespace sbfe {

constexpr int kNormThreads = 256;

// feature stores: float32, or bfloat16 (round to nearest even) for the opt-in reduced-precision
// output (sbfe.h: SBFE_OUT_BF16) - half the bytes for the consumer and for a device -> host copy
template <bool BF16>
__device__ __forceinline__ void store4(void* out, size_t i4, float4 r) {
  if (BF16) {
    const __nv_bfloat162 lo = __floats2bfloat162_rn(r.x, r.y), hi = __floats2bfloat162_rn(r.z, r.w);
    uint2 v;
    v.x = *reinterpret_cast<const unsigned*>(&lo);
    v.y = *reinterpret_cast<const unsigned*>(&hi);
    reinterpret_cast<uint2*>(out)[i4] = v;
  } else {
    reinterpret_cast<float4*>(out)[i4] = r;
  }
}
template <bool BF16>
__device__ __forceinline__ void store1(void* out, size_t i, float r) {
  if (BF16) reinterpret_cast<__nv_bfloat16*>(out)[i] = __float2bfloat16_rn(r);
  else reinterpret_cast<float*>(out)[i] = r;
}
constexpr int kStatChunk = 256;  // frames per CTA in the statistics pass

__device__ __forceinline__ int valid_frames(const float* lengths, int64_t b, int64_t T) {
  // torch.round(lengths[b] * T).int(): float32 product, round-half-to-even (:1023)
  int n = __float2int_rn(lengths[b] * (float)T);
  return n < 0 ? 0 : (n > T ? (int)T : n);
}

// partial sums about the pivot K = x[b, 0, f]:  ws[b][chunk][f] = (sum, sumsq)
__global__ void __launch_bounds__(kNormThreads)
norm_stats_partial_kernel(const float* __restrict__ x, const float* __restrict__ lengths,
                          int64_t T, int F, const sbfe_maxkey_t* __restrict__ utt_max,
                          float top_db, double2* __restrict__ ws, unsigned chunks, int bx) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ double2 s_red[kNormThreads];
  const int64_t b = blockIdx.x / chunks;
  const int c = blockIdx.x - b * chunks;
  const int n = valid_frames(lengths, b, T);
  const int t_lo = c * kStatChunk;
  const int t_hi = min(n, t_lo + kStatChunk);
  const int by = kNormThreads / bx;
  const int tx = threadIdx.x % bx, ty = threadIdx.x / bx;
  const float floor_b = clamp_floor(utt_max, b, top_db);
  const float* xb = x + b * T * F;
  for (int f0 = 0; f0 < F; f0 += bx) {
    const int f = f0 + tx;
    // double accumulators: the pass is memory-bound and B200 runs fp64 adds at
    // half the fp32 rate, so exactness of the statistics is free
    double s1 = 0.0, s2 = 0.0;
    if (f < F && t_lo < t_hi) {
      const float pivot = fmaxf(__ldg(xb + f), floor_b);
      for (int t = t_lo + ty; t < t_hi; t += by) {
        const double d = (double)(fmaxf(__ldg(xb + (int64_t)t * F + f), floor_b) - pivot);
        s1 += d;
        s2 = fma(d, d, s2);
      }
    }
    s_red[threadIdx.x] = make_double2(s1, s2);
    __syncthreads();
    if (ty == 0 && f < F) {
      double a1 = 0.0, a2 = 0.0;
      for (int r = 0; r < by; ++r) {
        a1 += s_red[r * bx + tx].x;
        a2 += s_red[r * bx + tx].y;
      }
      ws[((int64_t)b * chunks + c) * F + f] = make_double2(a1, a2);
    }
    __syncthreads();
  }
}

__global__ void norm_stats_final_kernel(const float* __restrict__ x,
                                        const float* __restrict__ lengths, int64_t B, int64_t T,
                                        int F, int mean_norm, int std_norm, float eps,
                                        const sbfe_maxkey_t* __restrict__ utt_max, float top_db,
                                        const double2* __restrict__ ws, unsigned chunks,
                                        float* __restrict__ mean, float* __restrict__ stdv) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= B * F) return;
  const int64_t b = i / F;
  const int f = (int)(i - b * F);
  const int n = valid_frames(lengths, b, T);
  double s1 = 0.0, s2 = 0.0;
  for (unsigned c = 0; c < chunks; ++c) {
    const double2 v = ws[((int64_t)b * chunks + c) * F + f];
    s1 += v.x;
    s2 += v.y;
  }
  const float floor_b = clamp_floor(utt_max, b, top_db);
  const double pivot = (T > 0) ? (double)fmaxf(x[b * T * F + f], floor_b) : 0.0;
  // n == 0: mean of an empty slice is NaN, n == 1: unbiased std is NaN (torch)
  const double m = (n > 0) ? pivot + s1 / n : (double)NAN;
  double var = (n > 1) ? (s2 - s1 * s1 / n) / (n - 1) : (double)NAN;
  if (var < 0.0) var = 0.0;
  float sd = (float)sqrt(var);
  if (sd == sd) sd = fmaxf(sd, eps);  // torch.max propagates NaN
  mean[i] = mean_norm ? (float)m : 0.f;
  stdv[i] = std_norm ? sd : 1.f;
}

// Partial moments merged with Chan's update (float: <= a few hundred well-conditioned merges).
struct Moments {
  float n, mean, m2;
};
__device__ __forceinline__ void merge(Moments& a, float nb, float mb, float m2b) {
  if (nb <= 0.f) return;
  const float n = a.n + nb, d = mb - a.mean, r = __fdividef(nb, n);
  a.mean = fmaf(d, r, a.mean);
  a.m2 += fmaf(d * d, a.n * r, m2b);
  a.n = n;
}


// ---- statistics from the TILE RECORDS of the fused Fbank kernel -------------------------------
// After sbfe_fbank_forward with tile records (fbank400_v3.cu) record j of an utterance holds, for
// every column over the existing frames of tile j, {mean[F], M2[F], min[F]} of the UN-clamped
// values plus the tile's max key; x itself is top_db-clamped.  A column whose minimum is not below
// the floor (rebuilt from the keys) was not changed by the clamp: its record is exact.  The
// statistics over the first n frames are then the records of the n / tile whole tiles merged
// (Chan) with the n % tile rows of the tile the valid length cuts - 3 F floats per 32 frames
// instead of a pass over T rows; a column the clamp did change (rare inside the valid part) is
// recomputed from x.  Replaces the per-sentence Python loop of processing/features.py:1020-1031
// and _compute_current_stats :1120-1145.

// moments of column `cp` (stride F) over cnt rows, every value through the clamp: the slow path of
// a record that cannot be used (two passes like torch's)
__device__ __noinline__ float2 column_moments_clamped(const float* cp, int F, int cnt, float floor_b) {
  float sm = 0.f;
  for (int r = 0; r < cnt; ++r) sm += fmaxf(__ldcg(cp + (size_t)r * F), floor_b);
  const float mean = sm / (float)cnt;
  float q = 0.f;
  for (int r = 0; r < cnt; ++r) {
    const float d = fmaxf(__ldcg(cp + (size_t)r * F), floor_b) - mean;
    q = fmaf(d, d, q);
  }
  return make_float2(mean, q);
}

// Whole CTA must call.  Thread = (column, tile group): CW = min(F, blockDim) columns per pass,
// G = blockDim / CW groups; a group takes tiles g, g + G, ... and rows g, g + G, ... of the cut
// tile.  `s_red`: 3 * blockDim.x floats.  Writes s_mean[F] / s_std[F] (shared or global) and ends
// with a barrier.  Returns the floor (-inf without a top_db).
__device__ __forceinline__ float tile_stats_block(const float* xb, int n, int T, int F, int tile,
                                                  int tiles, const float* tsb0, float top_db,
                                                  int mean_norm, int std_norm, float eps,
                                                  float* s_mean, float* s_std, float* s_red) {
  const int nthr = blockDim.x;
  const int CW = F < nthr ? F : nthr;            // columns per pass
  const int G = nthr / CW;                       // tile groups
  const int col = threadIdx.x % CW, grp = threadIdx.x / CW;
  const int rec = (int)tile_rec_words(F);
  const int whole = (n >= T) ? tiles : n / tile; // tiles whose existing frames are all valid
  const int r_cut = whole * tile;
  const int ncut = (whole < tiles && r_cut < n) ? n - r_cut : 0;   // rows of the cut tile
  __shared__ unsigned s_key;
  if (threadIdx.x == 0) s_key = 0u;
  __syncthreads();
  if (top_db < INFINITY) {
    unsigned key = 0u;
    for (int i = threadIdx.x; i < tiles; i += nthr) {
      const unsigned k2 = __ldcg(reinterpret_cast<const unsigned*>(tsb0) + (size_t)i * rec + 3 * F);
      key = k2 > key ? k2 : key;
    }
    key = __reduce_max_sync(0xffffffffu, key);
    if ((threadIdx.x & 31) == 0 && key) atomicMax(&s_key, key);
  }
  __syncthreads();
  const float floor_b = s_key ? max_key_decode(s_key) - top_db : -INFINITY;
  constexpr int kSpec = 4;
  for (int f0 = 0; f0 < F; f0 += CW) {
    const int f = f0 + col;
    const bool live = f < F && grp < G;
    Moments acc{0.f, 0.f, 0.f};
    if (live) {
      const float* tsb = tsb0 + f;
      float pm[kSpec], pq[kSpec], pn[kSpec];
      for (int j0 = grp; j0 < whole; j0 += kSpec * G) {   // batches: every load before the first merge
#pragma unroll
        for (int k = 0; k < kSpec; ++k) {
          const int j = j0 + k * G;
          if (j < whole) {
            pm[k] = __ldcg(tsb + (size_t)j * rec);
            pq[k] = __ldcg(tsb + (size_t)j * rec + F);
            pn[k] = __ldcg(tsb + (size_t)j * rec + 2 * F);
          }
        }
#pragma unroll
        for (int k = 0; k < kSpec; ++k) {
          const int j = j0 + k * G;
          if (j < whole) {
            const int cnt = T - j * tile < tile ? T - j * tile : tile;
            float2 mq = make_float2(pm[k], pq[k]);
            if (pn[k] < floor_b)
              mq = column_moments_clamped(xb + (size_t)j * tile * F + f, F, cnt, floor_b);
            merge(acc, (float)cnt, mq.x, mq.y);
          }
        }
      }
      if (grp < ncut) {      // x is clamped already when this kernel may read it: max() is a no-op then
        const float* colp = xb + (size_t)r_cut * F + f;
        const float pivot = fmaxf(__ldcg(colp), floor_b);
        float s1 = 0.f, s2 = 0.f;
        int c = 0;
        for (int r = grp; r < ncut; r += G, ++c) {
          const float d = fmaxf(__ldcg(colp + (size_t)r * F), floor_b) - pivot;
          s1 += d;
          s2 = fmaf(d, d, s2);
        }
        const float mloc = s1 / (float)c;
        merge(acc, (float)c, pivot + mloc, fmaxf(fmaf(-mloc, s1, s2), 0.f));
      }
    }
    if (G > 1) {
      s_red[threadIdx.x] = acc.n;
      s_red[nthr + threadIdx.x] = acc.mean;
      s_red[2 * nthr + threadIdx.x] = acc.m2;
      __syncthreads();
      if (live && grp == 0)
        for (int g = 1; g < G; ++g)
          merge(acc, s_red[g * CW + col], s_red[nthr + g * CW + col], s_red[2 * nthr + g * CW + col]);
    }
    if (live && grp == 0) {
      // n == 0: mean of an empty slice is NaN, n == 1: unbiased std is NaN (torch)
      const float m = (acc.n > 0.f) ? acc.mean : NAN;
      float sd = (acc.n > 1.f) ? sqrtf(fmaxf(acc.m2, 0.f) / (acc.n - 1.f)) : NAN;
      if (sd == sd) sd = fmaxf(sd, eps);  // torch.max propagates NaN
      s_mean[f] = mean_norm ? m : 0.f;
      s_std[f] = std_norm ? sd : 1.f;
    }
    __syncthreads();
  }
  return floor_b;
}

constexpr int kTileNormThreads = 512;

// mean / std [B, F] from the tile records: grid = B
__global__ void __launch_bounds__(kTileNormThreads)
norm_stats_tiles_kernel(const float* __restrict__ x, const float* __restrict__ lengths, int64_t T,
                        int F, int tile, int tiles, const float* __restrict__ tile_stats,
                        float top_db, int mean_norm, int std_norm, float eps,
                        float* __restrict__ mean, float* __restrict__ stdv) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float s_red[3 * kTileNormThreads];
  const int64_t b = blockIdx.x;
  const int n = valid_frames(lengths, b, T);
  tile_stats_block(x + b * T * F, n, (int)T, F, tile, tiles,
                   tile_stats + (size_t)b * tiles * tile_rec_words(F), top_db, mean_norm, std_norm,
                   eps, mean + b * F, stdv + b * F, s_red);
}

// Sentence normalisation in ONE launch (processing/features.py:1020-1035): every CTA rebuilds its
// utterance's statistics from the records (2 F floats per tile, L2-resident: cheaper than a
// second launch and its dependent round trips) and normalises its share of the utterance - half
// of its vectors of x are requested first and fly while the statistics are merged.
//
// BESIDE_CLAMP: the launch directly before this one in the stream may be the clamp kernel of the
// sbfe_fbank_forward call that produced x and the records (or anything else that leaves lengths
// alone and writes x / the records only the way that kernel does: x = max(x, floor) with the same
// floor; moments of records whose minimum is below the floor).  This kernel then does not wait
// for it before it READS: every value of x goes through max(., floor) - the same before and after
// that write -, the floor is rebuilt from the tile keys, and a record whose minimum is below the
// floor is not trusted (its column is recomputed from x).  It waits right before its first
// store, which also keeps completion transitive for whatever follows (common.cuh).  Everything
// before that direct predecessor has completed when this kernel starts, so the records and x as
// the producing kernel left them are visible.
// grid = B * chunks; VEC = 4 when F % 4 == 0 and x / out are 16-byte aligned; F <= blockDim.
constexpr int kSentIters = 8;    // vectors per thread
constexpr int kSentSpec = 4;     // records per thread and batch

template <int VEC, bool BESIDE_CLAMP, bool BF16OUT>
__global__ void __launch_bounds__(kTileNormThreads, 2)
norm_sentence_tiles_kernel(const float* x, const float* __restrict__ lengths, int T, int F,
                           int tile, int tiles, float* tile_stats, float top_db,
                           int mean_norm, int std_norm, float eps, void* out, unsigned chunks,
                           int inplace) {
  SBFE_TL(2, 0);
  if (!BESIDE_CLAMP) {
    pdl_wait();
    pdl_launch_dependents();
    SBFE_TL(2, 1);
  }
  extern __shared__ __align__(16) float s_dyn[];   // mean[F] | std[F] | 3 * blockDim scratch | bitmask
  __shared__ unsigned s_key, s_ticket;
  float* s_mean = s_dyn;
  float* s_std = s_dyn + F;
  float* s_red = s_dyn + 2 * F;
  // IN PLACE with more than one CTA per utterance: tiles whose rows are READ for the statistics
  // (the tile the length cuts, tiles with a column below the floor) - see the end of the kernel
  unsigned* s_held = reinterpret_cast<unsigned*>(s_red + 3 * kTileNormThreads);
#ifdef SBFE_EXP_NOGUARD   // experiment build: shows that tests/test_gpu_r2.py catches the race
  const bool guard = false;
#else
  const bool guard = inplace != 0 && chunks > 1;
#endif
  unsigned* const arrivals =
      reinterpret_cast<unsigned*>(tile_stats + (size_t)(blockIdx.x / chunks) * tiles * tile_rec_words(F)) +
      3 * F + 3;
  bool held_any = false;
  if (guard)
    for (int i = threadIdx.x; i < (tiles + 31) / 32; i += kTileNormThreads) s_held[i] = 0u;
  const unsigned b = blockIdx.x / chunks;
  const int c0 = blockIdx.x - b * chunks;
  const float* xb = x + (size_t)b * T * F;
  float floor_b = -INFINITY;
  // ---- the utterance's statistics from its records (F <= blockDim: one pass over the columns).
  // x and the records may be written beside us (BESIDE_CLAMP): read through L2 (ld.cg).
  {
    if (threadIdx.x == 0) s_key = 0u;
    const int n = valid_frames(lengths, b, T);
    const int G = kTileNormThreads / F;            // tile groups
    const int col = threadIdx.x % F, grp = threadIdx.x / F;
    const int rec = (int)tile_rec_words(F);
    const float* tsb0 = tile_stats + (size_t)b * tiles * rec;
    const float* tsb = tsb0 + col;
    const bool live = grp < G;
    // the first batch does not wait for the valid length (`tiles`, not `whole`): lengths, keys
    // and records share one round trip
    float pm[kSentSpec], pq[kSentSpec], pn[kSentSpec];
    if (live) {
#pragma unroll
      for (int k = 0; k < kSentSpec; ++k) {
        const int j = grp + k * G;
        if (j < tiles) {
          pm[k] = __ldcg(tsb + j * rec);
          pq[k] = __ldcg(tsb + j * rec + F);
          pn[k] = __ldcg(tsb + j * rec + 2 * F);
        }
      }
    }
    unsigned key = 0u;
    if (top_db < INFINITY)
      for (int i = threadIdx.x; i < tiles; i += kTileNormThreads) {
        const unsigned k2 = __ldcg(reinterpret_cast<const unsigned*>(tsb0) + i * rec + 3 * F);
        key = k2 > key ? k2 : key;
      }
    __syncthreads();                               // s_key = 0 is in
    key = __reduce_max_sync(0xffffffffu, key);
    if ((threadIdx.x & 31) == 0 && key) atomicMax(&s_key, key);
    __syncthreads();
    if (s_key) floor_b = max_key_decode(s_key) - top_db;
    SBFE_TL(4, 0);   // keys + first records arrived, floor known
    Moments acc{0.f, 0.f, 0.f};
    if (live) {
      const int whole = (n >= T) ? tiles : n / tile;
      const int r_cut = whole * tile;
      const int ncut = (whole < tiles && r_cut < n) ? n - r_cut : 0;
      const float* colp = xb + (size_t)r_cut * F + col;
      float pivot = 0.f, cv = 0.f;
      if (ncut > 0) {                              // rows of the cut tile: second round trip
        pivot = __ldcg(colp);
        if (grp < ncut) cv = __ldcg(colp + grp * F);
      }
      for (int j0 = grp;;) {
#pragma unroll
        for (int k = 0; k < kSentSpec; ++k) {
          const int j = j0 + k * G;
          if (j < whole) {
            const int cnt = T - j * tile < tile ? T - j * tile : tile;
            // a record below the floor (values the clamp changes, inside the valid part: rare)
            // is not used - its moments may be the clamp kernel's or the producer's
            float2 mq = make_float2(pm[k], pq[k]);
            if (pn[k] < floor_b) {
              mq = column_moments_clamped(xb + (size_t)j * tile * F + col, F, cnt, floor_b);
              if (guard) {
                atomicOr(&s_held[j >> 5], 1u << (j & 31));
                held_any = true;
              }
            }
            merge(acc, (float)cnt, mq.x, mq.y);
          }
        }
        j0 += kSentSpec * G;
        if (j0 >= whole) break;
#pragma unroll
        for (int k = 0; k < kSentSpec; ++k) {
          const int j = j0 + k * G;
          if (j < whole) {
            pm[k] = __ldcg(tsb + j * rec);
            pq[k] = __ldcg(tsb + j * rec + F);
            pn[k] = __ldcg(tsb + j * rec + 2 * F);
          }
        }
      }
      if (guard && ncut > 0 && threadIdx.x == 0) {
        atomicOr(&s_held[whole >> 5], 1u << (whole & 31));
        held_any = true;
      }
      if (grp < ncut) {
        pivot = fmaxf(pivot, floor_b);
        float d = fmaxf(cv, floor_b) - pivot, s1 = d, s2 = d * d;
        int c = 1;
        for (int r = grp + G; r < ncut; r += G, ++c) {
          d = fmaxf(__ldcg(colp + r * F), floor_b) - pivot;
          s1 += d;
          s2 = fmaf(d, d, s2);
        }
        const float mloc = s1 / (float)c;
        merge(acc, (float)c, pivot + mloc, fmaxf(fmaf(-mloc, s1, s2), 0.f));
      }
    }
    SBFE_TL(5, 0);   // per-thread merges (all records, cut rows) done
    s_red[threadIdx.x] = acc.n;
    s_red[kTileNormThreads + threadIdx.x] = acc.mean;
    s_red[2 * kTileNormThreads + threadIdx.x] = acc.m2;
    __syncthreads();   // every thread's reads of x for the statistics are complete
    unsigned ticket = 0u;
    if (guard && threadIdx.x == 0) ticket = atomicAdd(arrivals, 1u);   // in flight during the merge below
    if (threadIdx.x < F) {
      for (int g = 1; g < G; ++g)
        merge(acc, s_red[g * F + col], s_red[kTileNormThreads + g * F + col],
              s_red[2 * kTileNormThreads + g * F + col]);
      // n == 0: mean of an empty slice is NaN, n == 1: unbiased std is NaN (torch)
      const float m = (acc.n > 0.f) ? acc.mean : NAN;
      float sd = (acc.n > 1.f) ? sqrtf(fmaxf(acc.m2, 0.f) / (acc.n - 1.f)) : NAN;
      if (sd == sd) sd = fmaxf(sd, eps);  // torch.max propagates NaN
      s_mean[col] = mean_norm ? m : 0.f;
      s_std[col] = std_norm ? sd : 1.f;
    }
    if (guard && threadIdx.x == 0) {
      if (ticket == chunks - 1) *arrivals = 0u;
      s_ticket = ticket;
    }
    held_any = __syncthreads_or(held_any) != 0;   // block-uniform from here on (false without guard)
  }
  const bool last = held_any && s_ticket == chunks - 1;

  // ---- in place, rows that the statistics READ: the CTAs of an utterance all read them, and
  // one of them owns them in the pass below - it must not overwrite them while a sibling (a later
  // wave of the grid, a slower SM) still computes its statistics.  Every CTA draws a ticket AFTER
  // its own statistics (word 3F + 3 of the utterance's first record: zero from the producer,
  // zero again afterwards); nobody stores the held tiles in the pass below; the CTA that drew the
  // last ticket - every sibling has finished reading - normalises them itself.  (A sibling's
  // reads are complete when it draws: their values went into its statistics; the last CTA's
  // stores depend on the ticket it drew: no fences.)  The ticket is drawn between the two
  // barriers of the column merge: its round trip costs no barrier of its own.
  // ---- (max(x, floor) - mean) / std over the CTA's share of the utterance: all loads first
  const int FV = F / VEC;
  const int nvec = T * FV;                         // < 2^30 (launcher)
  const int i0 = c0 * (kTileNormThreads * kSentIters) + threadIdx.x;
  const size_t ob = (size_t)b * T * F;            // element offset of the utterance in out
  float4 v4[VEC == 4 ? kSentIters : 1];
  float v1[VEC == 4 ? 1 : kSentIters];
#pragma unroll
  for (int k = 0; k < kSentIters; ++k) {
    const int i = i0 + k * kTileNormThreads;
    if (i < nvec) {
      if (VEC == 4) v4[k] = __ldcg(reinterpret_cast<const float4*>(xb + 4 * (size_t)i));
      else v1[k] = __ldcg(xb + i);
    }
  }
  SBFE_TL(3, 0);          // statistics done, vectors requested
  // which of them this thread stores (bit k): inside the utterance and, in place, not in a held
  // tile - worked out while the loads are in flight
  const int row0 = i0 / FV, col0 = i0 - row0 * FV;
  const int drow = kTileNormThreads / FV, dcol = kTileNormThreads - drow * FV;
  unsigned live8 = 0u;
  {
    const int tshift = (tile & (tile - 1)) == 0 ? __ffs(tile) - 1 : -1;
    int r = row0, c = col0;
#pragma unroll
    for (int k = 0; k < kSentIters; ++k) {
      bool mine = i0 + k * kTileNormThreads < nvec;
      if (held_any && mine) {
        const int j = tshift >= 0 ? r >> tshift : r / tile;
        mine = ((s_held[j >> 5] >> (j & 31)) & 1u) == 0u;
      }
      live8 |= (mine ? 1u : 0u) << k;
      r += drow;
      c += dcol;
      if (c >= FV) {
        c -= FV;
        ++r;
      }
    }
  }
  if (BESIDE_CLAMP) {   // nothing has been written so far
    pdl_wait();
    pdl_launch_dependents();
    SBFE_TL(2, 1);
  }
  if (last) {   // the held tiles (a few of the utterance): their loads join the ones in flight
    for (int j = 0; j < tiles; ++j) {
      if (((s_held[j >> 5] >> (j & 31)) & 1u) == 0u) continue;
      const int r_lo = j * tile;
      const int cnt = (T - r_lo < tile ? T - r_lo : tile) * FV;   // vectors of the tile
      const int v0 = r_lo * FV;
      constexpr int kBatch = 2;                                    // loads first, then the stores
      for (int e0 = threadIdx.x; e0 < cnt; e0 += kBatch * kTileNormThreads) {
        float4 h4[VEC == 4 ? kBatch : 1];
        float h1[VEC == 4 ? 1 : kBatch];
#pragma unroll
        for (int k = 0; k < kBatch; ++k) {
          const int e = e0 + k * kTileNormThreads;
          if (e < cnt) {
            if (VEC == 4) h4[k] = __ldcg(reinterpret_cast<const float4*>(xb) + v0 + e);
            else h1[k] = __ldcg(xb + v0 + e);
          }
        }
#pragma unroll
        for (int k = 0; k < kBatch; ++k) {
          const int e = e0 + k * kTileNormThreads;
          if (e < cnt) {
            const int c = e % FV;
            if (VEC == 4) {
              const float4 mm = *reinterpret_cast<const float4*>(s_mean + 4 * c);
              const float4 ss = *reinterpret_cast<const float4*>(s_std + 4 * c);
              float4 r;
              r.x = (fmaxf(h4[k].x, floor_b) - mm.x) / ss.x;
              r.y = (fmaxf(h4[k].y, floor_b) - mm.y) / ss.y;
              r.z = (fmaxf(h4[k].z, floor_b) - mm.z) / ss.z;
              r.w = (fmaxf(h4[k].w, floor_b) - mm.w) / ss.w;
              store4<BF16OUT>(out, ob / 4 + (size_t)(v0 + e), r);
            } else {
              store1<BF16OUT>(out, ob + (size_t)(v0 + e), (fmaxf(h1[k], floor_b) - s_mean[c]) / s_std[c]);
            }
          }
        }
      }
    }
  }
  int col = col0;
#pragma unroll
  for (int k = 0; k < kSentIters; ++k) {
    const int i = i0 + k * kTileNormThreads;
    if ((live8 >> k) & 1u) {
      if (VEC == 4) {
        const float4 mm = *reinterpret_cast<const float4*>(s_mean + 4 * col);
        const float4 ss = *reinterpret_cast<const float4*>(s_std + 4 * col);
        float4 r;
        r.x = (fmaxf(v4[k].x, floor_b) - mm.x) / ss.x;
        r.y = (fmaxf(v4[k].y, floor_b) - mm.y) / ss.y;
        r.z = (fmaxf(v4[k].z, floor_b) - mm.z) / ss.z;
        r.w = (fmaxf(v4[k].w, floor_b) - mm.w) / ss.w;
        store4<BF16OUT>(out, ob / 4 + (size_t)i, r);
      } else {
        store1<BF16OUT>(out, ob + (size_t)i, (fmaxf(v1[k], floor_b) - s_mean[col]) / s_std[col]);
      }
    }
    col += dcol;
    if (col >= FV) col -= FV;
  }
  SBFE_TL(2, 2);
}

// grid = ceil(F / 32) CTAs of 32 columns x 8 utterance groups
__global__ void __launch_bounds__(256)
norm_reduce_kernel(const float* __restrict__ mean, const float* __restrict__ stdv, int64_t B, int F,
                   float* __restrict__ cur) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float s_a[256], s_s[256];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int f = blockIdx.x * 32 + tx;
  float a = 0.f, s = 0.f;
  if (f < F)
    for (int64_t b = ty; b < B; b += 8) {
      a += __ldg(mean + b * F + f);
      s += __ldg(stdv + b * F + f);
    }
  s_a[threadIdx.x] = a;
  s_s[threadIdx.x] = s;
  __syncthreads();
  if (ty == 0 && f < F) {
    for (int r = 1; r < 8; ++r) {
      a += s_a[r * 32 + tx];
      s += s_s[r * 32 + tx];
    }
    cur[f] = a;
    cur[F + f] = s;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) cur[2 * F] = (float)B;
}

__global__ void norm_update_kernel(const float* __restrict__ cur, int F, int first, float weight,
                                   float* __restrict__ gmean, float* __restrict__ gstd) {
  pdl_wait();
  pdl_launch_dependents();
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const float n = cur[2 * F];
  const float cm = cur[f] / n, cs = cur[F + f] / n;
  if (first) {
    gmean[f] = cm;
    gstd[f] = cs;
  } else {
    gmean[f] = (1.f - weight) * gmean[f] + weight * cm;
    gstd[f] = (1.f - weight) * gstd[f] + weight * cs;
  }
}

// out = (max(x, floor_b) - mean) / std.  VEC = 4 when F % 4 == 0 and 16 B aligned.
// An utterance is treated as a flat array of T*F/VEC vectors; a CTA takes kApplyIters
// consecutive blocks of kNormThreads vectors, all loads issued before the first store.
constexpr int kApplyIters = 8;

template <int VEC, bool BF16OUT>
__global__ void __launch_bounds__(kNormThreads)
norm_apply_kernel(const float* x, int64_t T, int F, const float* __restrict__ mean,
                  const float* __restrict__ stdv, int per_utt,
                  const sbfe_maxkey_t* __restrict__ utt_max, float top_db, void* out,
                  unsigned chunks) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ __align__(16) float s_ms[];  // mean[F] | std[F]
  const int64_t b = blockIdx.x / chunks;
  const int c = blockIdx.x - b * chunks;
  const int FV = F / VEC;
  const int64_t nvec = T * FV;
  const int64_t i0 = (int64_t)c * (kNormThreads * kApplyIters) + threadIdx.x;
  const float* xb = x + b * T * F;
  const size_t ob = (size_t)b * T * F;
  int col = (int)(i0 % FV);
  const int dcol = kNormThreads % FV;
  float4 v4[VEC == 4 ? kApplyIters : 1];
  float v1[VEC == 4 ? 1 : kApplyIters];
#pragma unroll
  for (int k = 0; k < kApplyIters; ++k) {
    const int64_t i = i0 + (int64_t)k * kNormThreads;
    if (i < nvec) {
      if (VEC == 4) v4[k] = *reinterpret_cast<const float4*>(xb + 4 * i);
      else v1[k] = xb[i];
    }
  }
  const float* m = mean + (per_utt ? b * F : 0);
  const float* s = stdv + (per_utt ? b * F : 0);
  for (int f = threadIdx.x; f < F; f += kNormThreads) {
    s_ms[f] = __ldcg(m + f);
    s_ms[F + f] = __ldcg(s + f);
  }
  const float floor_b = clamp_floor(utt_max, b, top_db);
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kApplyIters; ++k) {
    const int64_t i = i0 + (int64_t)k * kNormThreads;
    if (i < nvec) {
      if (VEC == 4) {
        const float4 mm = *reinterpret_cast<const float4*>(s_ms + 4 * col);
        const float4 ss = *reinterpret_cast<const float4*>(s_ms + F + 4 * col);
        float4 r;
        r.x = (fmaxf(v4[k].x, floor_b) - mm.x) / ss.x;
        r.y = (fmaxf(v4[k].y, floor_b) - mm.y) / ss.y;
        r.z = (fmaxf(v4[k].z, floor_b) - mm.z) / ss.z;
        r.w = (fmaxf(v4[k].w, floor_b) - mm.w) / ss.w;
        store4<BF16OUT>(out, ob / 4 + (size_t)i, r);
      } else {
        store1<BF16OUT>(out, ob + (size_t)i, (fmaxf(v1[k], floor_b) - s_ms[col]) / s_ms[F + col]);
      }
    }
    col += dcol;
    if (col >= FV) col -= FV;
  }
}

static int pick_bx(int cols) {  // power of two <= 256 covering `cols`
  int bx = 1;
  while (bx < cols && bx < kNormThreads) bx <<= 1;
  return bx;
}

int64_t norm_workspace_bytes(int64_t B, int64_t T, int F) {
  const int64_t chunks = (T + kStatChunk - 1) / kStatChunk;
  return B * (chunks > 0 ? chunks : 1) * F * (int64_t)sizeof(double2);
}

int launch_norm_stats(const float* x, const float* lengths, int64_t B, int64_t T, int F,
                      int mean_norm, int std_norm, float eps, const sbfe_maxkey_t* utt_max,
                      float top_db, float* mean, float* stdv, void* ws, int64_t ws_bytes,
                      cudaStream_t stream) {
  if (B == 0 || F == 0) return 0;
  const int64_t chunks = (T + kStatChunk - 1) / kStatChunk;
  if (ws_bytes < norm_workspace_bytes(B, T, F)) return SBFE_E_NOSPACE;
  if (chunks * B > 0x7fffffff || B * F > 0x7fffffff) return SBFE_E_TOOBIG;
  if (chunks > 0) {
    int rc = launch_pdl(norm_stats_partial_kernel, dim3((unsigned)(chunks * B)), dim3(kNormThreads),
                        0, stream, x, lengths, T, F, utt_max, top_db, (double2*)ws,
                        (unsigned)chunks, pick_bx(F));
    if (rc) return rc;
  }
  return launch_pdl(norm_stats_final_kernel, dim3((unsigned)((B * F + 255) / 256)), dim3(256), 0,
                    stream, x, lengths, B, T, F, mean_norm, std_norm, eps, utt_max, top_db,
                    (const double2*)ws, (unsigned)chunks, mean, stdv);
}

// ---- the no-allocation variant of launch_norm_stats ---------------------------------------------
// Same two kernels; the partial sums live in a file-scope device array instead of a caller-provided
// workspace, for callers that cannot allocate (CUDA graph capture, serving loops).  Every GPU that
// loads the library holds these SBFE_NORM_STATIC_WS_BYTES until the process exits.  The array's
// name is a host-side stand-in, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.  Nothing has to start at zero:
// norm_stats_partial_kernel writes every (utterance, chunk, column) slot that
// norm_stats_final_kernel reads.  All calls share the array, so two calls that overlap on the
// GPU (different streams) race; calls on one stream are ordered.
static __device__ __align__(256) unsigned char g_norm_static_ws[SBFE_NORM_STATIC_WS_BYTES];

int launch_norm_stats_static(const float* x, const float* lengths, int64_t B, int64_t T, int F,
                             int mean_norm, int std_norm, float eps, const sbfe_maxkey_t* utt_max,
                             float top_db, float* mean, float* stdv, cudaStream_t stream) {
  if (B == 0 || F == 0) return 0;
  if (norm_workspace_bytes(B, T, F) > (int64_t)SBFE_NORM_STATIC_WS_BYTES) return SBFE_E_STATIC_LIMIT;
  static std::atomic<void*> addr[64];   // per device; the only transition is null -> the address
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  void* ws = (dev >= 0 && dev < 64) ? addr[dev].load(std::memory_order_acquire) : nullptr;
  if (ws == nullptr) {
    e = cudaGetSymbolAddress(&ws, g_norm_static_ws);
    if (e != cudaSuccess || ws == nullptr) return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
    if (dev >= 0 && dev < 64) addr[dev].store(ws, std::memory_order_release);
  }
  return launch_norm_stats(x, lengths, B, T, F, mean_norm, std_norm, eps, utt_max, top_db, mean,
                           stdv, ws, (int64_t)SBFE_NORM_STATIC_WS_BYTES, stream);
}

int launch_norm_stats_tiles(const float* x, const float* lengths, int64_t B, int64_t T, int F,
                            int mean_norm, int std_norm, float eps, const float* tile_stats,
                            int tile, float top_db, float* mean, float* stdv, cudaStream_t stream) {
  if (B == 0 || F == 0) return 0;
  if (B > 0x7fffffff || T > 0x7fffffff) return SBFE_E_TOOBIG;
  const int tiles = (int)((T + tile - 1) / tile);
  return launch_pdl(norm_stats_tiles_kernel, dim3((unsigned)B), dim3(kTileNormThreads), 0, stream, x,
                    lengths, T, F, tile, tiles, tile_stats, top_db >= 0.f ? top_db : INFINITY, mean_norm,
                    std_norm, eps, mean, stdv);
}

int launch_norm_reduce(const float* mean, const float* stdv, int64_t B, int F, float* cur,
                       cudaStream_t stream) {
  return launch_pdl(norm_reduce_kernel, dim3((F + 31) / 32), dim3(256), 0, stream, mean, stdv, B, F, cur);
}

int launch_norm_update(const float* cur, int F, int first, float weight, float* gmean,
                       float* gstd, cudaStream_t stream) {
  if (F == 0) return 0;
  return launch_pdl(norm_update_kernel, dim3((F + 127) / 128), dim3(128), 0, stream, cur, F, first,
                    weight, gmean, gstd);
}

int launch_norm_apply(const float* x, int64_t B, int64_t T, int F, const float* mean,
                      const float* stdv, int per_utt, const sbfe_maxkey_t* utt_max, float top_db,
                      int out_bf16, void* out, cudaStream_t stream) {
  if (B == 0 || T == 0 || F == 0) return 0;
  const bool vec = (F % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0 &&
                   (B == 1 || (T * F) % 4 == 0);
  const int64_t nvec = T * (vec ? F / 4 : F);
  const int64_t per_cta = (int64_t)kNormThreads * kApplyIters;
  const int64_t chunks = (nvec + per_cta - 1) / per_cta;
  if (chunks * B > 0x7fffffff) return SBFE_E_TOOBIG;
  const size_t smem = (size_t)2 * F * sizeof(float);
  if (smem > 48 * 1024) return SBFE_E_TOOBIG;
  auto k = out_bf16 ? (vec ? norm_apply_kernel<4, true> : norm_apply_kernel<1, true>)
                    : (vec ? norm_apply_kernel<4, false> : norm_apply_kernel<1, false>);
  return launch_pdl(k, dim3((unsigned)(chunks * B)), dim3(kNormThreads), smem, stream, x, T, F, mean,
                    stdv, per_utt, utt_max, top_db, out, (unsigned)chunks);
}

// sentence normalisation from the tile records, one launch
int launch_norm_sentence_tiles(const float* x, const float* lengths, int64_t B, int64_t T, int F,
                               int mean_norm, int std_norm, float eps, float* tile_stats,
                               int tile, float top_db, int beside_clamp, int out_bf16, void* out,
                               cudaStream_t stream) {
  if (B == 0 || T == 0 || F == 0) return 0;
  if (B > 0x7fffffff || T > 0x7fffffff) return SBFE_E_TOOBIG;
  const int tiles = (int)((T + tile - 1) / tile);
  const bool vec = (F % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  const int64_t nvec = T * (vec ? F / 4 : F);
  const int64_t per_cta = (int64_t)kTileNormThreads * kSentIters;
  const int64_t chunks = (nvec + per_cta - 1) / per_cta;
  if (chunks * B > 0x7fffffff) return SBFE_E_TOOBIG;
  if (F > kTileNormThreads || nvec > 0x3fffffff) return SBFE_E_TOOBIG;
  const bool inplace = out == static_cast<const void*>(x) && !out_bf16;
  const size_t smem = ((size_t)2 * F + 3 * kTileNormThreads + (inplace ? (tiles + 31) / 32 : 0)) * sizeof(float);
  if (smem > 48 * 1024) return SBFE_E_TOOBIG;
  // the flag presumes that call's last launch was its clamp kernel: there is none without a top_db
  const bool beside = beside_clamp != 0 && top_db >= 0.f && pdl_enabled();
  using Kern = decltype(&norm_sentence_tiles_kernel<4, true, false>);
  static const Kern table[8] = {
      norm_sentence_tiles_kernel<1, false, false>, norm_sentence_tiles_kernel<4, false, false>,
      norm_sentence_tiles_kernel<1, true, false>,  norm_sentence_tiles_kernel<4, true, false>,
      norm_sentence_tiles_kernel<1, false, true>,  norm_sentence_tiles_kernel<4, false, true>,
      norm_sentence_tiles_kernel<1, true, true>,   norm_sentence_tiles_kernel<4, true, true>};
  Kern k = table[(out_bf16 ? 4 : 0) + (beside ? 2 : 0) + (vec ? 1 : 0)];
  return launch_pdl(k, dim3((unsigned)(chunks * B)), dim3(kTileNormThreads), smem, stream, x, lengths,
                    (int)T, F, tile, tiles, tile_stats, top_db >= 0.f ? top_db : INFINITY, mean_norm,
                    std_norm, eps, out, (unsigned)chunks, inplace ? 1 : 0);
}

// ---- speaker mode (processing/features.py:1037-1081): running per-speaker statistics ---------
// The dictionary lives on the host (python ints -> table rows); this kernel replays the batch in
// order for every feature column: row[i] < 0: unseen speaker outside training -> the sentence's
// own statistics; mode[i] = 2: first sight -> the row becomes the sentence's statistics;
// mode[i] = 1: row = (1 - w) * row + w * current (two roundings and an add, like the reference's
// eager ops: no FMA contraction); mode[i] = 0: read only.  out[i] = the row AFTER its update,
// which is what the reference normalises sentence i with.
__global__ void norm_speaker_update_kernel(const float* __restrict__ mean,
                                           const float* __restrict__ stdv, int B, int F,
                                           const int32_t* __restrict__ row,
                                           const int32_t* __restrict__ mode,
                                           const float* __restrict__ weight, float* tab_mean,
                                           float* tab_std, float* __restrict__ out_mean,
                                           float* __restrict__ out_std) {
  pdl_wait();
  pdl_launch_dependents();
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  for (int i = 0; i < B; ++i) {   // in batch order: a speaker may appear more than once
    const int r = row[i];
    float m = mean[(int64_t)i * F + f], s = stdv[(int64_t)i * F + f];
    if (r >= 0) {
      float* tm = tab_mean + (int64_t)r * F + f;
      float* tsd = tab_std + (int64_t)r * F + f;
      const int md = mode[i];
      if (md == 1) {
        const float w = weight[i], k = weight[B + i];   // w and (1 - w), rounded on the host
        m = __fadd_rn(__fmul_rn(k, *tm), __fmul_rn(w, m));
        s = __fadd_rn(__fmul_rn(k, *tsd), __fmul_rn(w, s));
      } else if (md == 0) {
        m = *tm;
        s = *tsd;
      }
      if (md != 0) {
        *tm = m;
        *tsd = s;
      }
    }
    out_mean[(int64_t)i * F + f] = m;
    out_std[(int64_t)i * F + f] = s;
  }
}

int launch_norm_speaker_update(const float* mean, const float* stdv, int64_t B, int F,
                               const int32_t* row, const int32_t* mode, const float* weight,
                               float* tab_mean, float* tab_std, float* out_mean, float* out_std,
                               cudaStream_t stream) {
  if (B == 0 || F == 0) return 0;
  if (B > 0x7fffffff) return SBFE_E_TOOBIG;
  return launch_pdl(norm_speaker_update_kernel, dim3((F + 127) / 128), dim3(128), 0, stream, mean,
                    stdv, (int)B, F, row, mode, weight, tab_mean, tab_std, out_mean, out_std);
}

#ifdef SBFE_EXP_TIMELINE
SBFE_TL_DEFINE_SYNC(tl_sync_norm)
#endif

}  // namespace sbfe
