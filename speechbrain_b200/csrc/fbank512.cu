// Fused frame + window + real-FFT(512) + power + sparse mel + dB kernel (n_fft = 512, hop = 160:
// what `Fbank(n_fft=512)` builds at 16 kHz; any win_length <= 512).  Same contract and the same
// tile records as fbank400_v3.cu:
//   STFT.forward          speechbrain/processing/features.py:133-188
//   spectral_magnitude    speechbrain/processing/features.py:327-356 (power=1)
//   Filterbank.forward    speechbrain/processing/features.py:490-551
//   _amplitude_to_DB      speechbrain/processing/features.py:701-702 (log part)
//
// The 512-point real DFT of a frame is the 256-point complex DFT of z[n] = x[2n] + i x[2n+1]
// followed by the real-input untangle; 256 = 16 x 16 Cooley-Tukey with BOTH stages the same
// register-resident 16-point codelet (fft_packed.cuh: dft16, two radix-4 levels):
//   stage 1  thread = (frame pair p, residue n2): windowed z[16 n1 + n2] over n1 -> DFT16 ->
//            times W256^(n2 k1) -> exchange[n2][k1][p] (float4 {re0, re1, im0, im1})
//   stage 2  thread = (frame pair p, k1): DFT16 over n2 -> Z[k1 + 16 k2], written back IN PLACE
//            (slot k of the exchange holds Z[k]: the thread reads and writes the same 16 slots)
//   untangle thread = (p, k1), its own Z[k] still in registers, the partner Z[256 - k] from the
//            exchange: X[k] = (Z[k] + conj Z[256-k]) / 2 - (i / 2) W512^k (Z[k] - conj Z[256-k]),
//            |X|^2 -> P[bin][32] over the (dead) waveform buffer
// then the mel / dB / store / tile-record stages of the 400-point kernel.  Like there a thread owns
// the adjacent frames (2p, 2p+1) as packed f32x2 lanes and they share samples: hop = 5 * 32, so
// tap pair n1 of the second frame is tap pair n1 + 5 of the first - 21 64-bit loads serve both.
// 16 units x 16 pairs = 256 threads; 2 CTAs / SM (107.7 KB of shared memory, <= 128 registers).
// One waveform buffer (the power tile overlays it): a CTA's staging is not overlapped with its own
// transform, the SM's other CTA covers it.
#include <atomic>
#include <cmath>

#include "common.cuh"
#include "fft_packed.cuh"

namespace sbfe {

namespace v512 {

constexpr int kTile = 32;
constexpr int kPairs = 16;
constexpr int kUnits = 16;
constexpr int kWarps = kUnits / 2;
constexpr int kThreads = kWarps * 32;
constexpr int kHop = 160;
constexpr int kNfft = 512;
constexpr int kHalf = kNfft / 2;                       // complex FFT size
constexpr int kBins = kHalf + 1;
constexpr float kImSign = 1.f;   // COMPLEX: sign that turns the untangle's imaginary parts into torch.stft's
constexpr int kSpan = (kTile - 1) * kHop + kNfft;      // 5472 samples per tile
constexpr int kChunks = kSpan / 2;                     // 8-byte cp.async chunks
constexpr int kCpIters = (kChunks + kThreads - 1) / kThreads;
constexpr int kPad2 = 2;                               // floats behind every pair of hops
constexpr int kPitch2 = 2 * kHop + kPad2;              // 322 = 2 mod 32: conflict-free 64-bit reads
constexpr int kPRows = 260;                            // power rows incl. 3 zero rows (padded mel reads)
constexpr int kRegionA = kPRows * kTile;               // >= padded span (5506)
constexpr int kRegionE = kHalf * kPairs * 4;           // exchange, float4 entries
constexpr int kTabW = kNfft, kTabTw = 2 * kHalf, kTabUt = 2 * kHalf;

struct Params {
  DbParams db;
  int n_mels, mel_nnz;
  unsigned tiles_per_utt, n_tiles;
  int64_t L, wav_stride, T;
  const float* wav;
  const float* window;     // device, [512], zero-padded / centred to n_fft
  float* out;
  sbfe_maxkey_t* utt_max;
  float* tile_stats;       // optional [n_tiles] records of tile_rec_words(M) (common.cuh)
  const int32_t* mel_start;
  const int32_t* mel_count;
  const int32_t* mel_offset;
  const float* mel_weight;
};

__host__ __device__ inline size_t smem_floats(int n_mels, int nnz) {
  return (size_t)kRegionA + kRegionE + kTabW + kTabTw + kTabUt + ((2 * n_mels + 3) & ~3) +
         ((nnz + 3) & ~3);
}

__device__ __forceinline__ void cp8(float* smem_dst, const float* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp4(float* smem_dst, const float* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes)
               : "memory");
}

// one tile's waveform span into `dst` (sample i at i + 2 * (i / 320)); samples outside [0, L) are
// zero-filled by the copy engine (src-size < cp-size) = the reference's zero centre padding
__device__ __forceinline__ void stage_tile(const Params& prm, unsigned b, unsigned j, float* dst, int tid) {
  const float* row = prm.wav + (int64_t)b * prm.wav_stride;
  const int64_t g0 = (int64_t)j * (kTile * kHop) - kNfft / 2;     // even
  const float* wsrc = row + g0;                                   // dereferenced inside [lo, hi) only
  const int lo = g0 < 0 ? (int)(-g0) : 0;
  const int64_t rem = prm.L - g0;
  const int hi = rem <= 0 ? 0 : (rem < kSpan ? (int)rem : kSpan);
  if ((reinterpret_cast<uintptr_t>(row) & 7) == 0) {
#pragma unroll 1
    for (int k = 0; k < kCpIters; ++k) {
      const int i = 2 * (k * kThreads + tid);
      if (i < kSpan) {
        int n = hi - i;
        n = (i < lo || n < 0) ? 0 : (n > 2 ? 2 : n);
        cp8(dst + i + kPad2 * (i / (2 * kHop)), n ? wsrc + i : prm.window, 4 * n);
      }
    }
  } else {
    for (int i = tid; i < kSpan; i += kThreads) {
      const bool ok = i >= lo && i < hi;
      cp4(dst + i + kPad2 * (i / (2 * kHop)), ok ? wsrc + i : prm.window, ok ? 4 : 0);
    }
  }
}

// POWER: STFT + spectral_magnitude(power=1) of the default mode - the [32 frames][257 bins] tile of
// the power spectrum is the output (M = 257, no filterbank, no records); see fbank400_v3.cu
// COMPLEX: the STFT itself, out [B, T, 257, 2].  The untangle works on PAIRS (k, 256 - k): the
// thread that owns Z[k], k = k1 + 16 k2 with k2 < 8, reads the partner Z[256 - k] (a row with
// k2 >= 8 of unit 16 - k1) and writes X[k] AND X[256 - k] in place - every exchange row is then read
// and written by one half-warp only, so the entries can be rotated inside their row (slot (p + k) &
// 15, one __syncwarp between a row's reads and writes) for the conflict-free store pass of
// fbank400_v3.cu.
template <int MC, bool POWER = false, bool COMPLEX = false>
__global__ void __launch_bounds__(kThreads, 2) fbank512_kernel(const __grid_constant__ Params prm) {
  extern __shared__ __align__(16) float smem[];
  float* s_a = smem;                       // staged waveform, later P[bin][32]
  float* s_e = s_a + kRegionA;             // exchange (Z in place), later the output tile
  float* s_win = s_e + kRegionE;           // [n2][n1] (w[2n], w[2n+1]), n = 16 n1 + n2
  float* s_tw = s_win + kTabW;             // [n2][k1] (re, im) of W256^(n2 k1)
  float* s_ut = s_tw + kTabTw;             // [k] (cos, sin)(2 pi k / 512)
  int2* s_meta = reinterpret_cast<int2*>(s_ut + kTabUt);
  const int M = COMPLEX ? 2 * kBins : (POWER ? kBins : (MC > 0 ? MC : prm.n_mels));   // floats per output row
  float* s_w = reinterpret_cast<float*>(s_meta) + ((2 * M + 3) & ~3);

  __shared__ unsigned s_tile_key;
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int p = lane & 15, h = lane >> 4;
  const int unit = 2 * warp + h;

  // ---- once per CTA: constant tables (may overlap the preceding kernel's tail) ----------------
  for (int i = tid; i < kHalf; i += kThreads) {
    const int n1 = i >> 4, n2 = i & 15;                   // sample pair n = 16 n1 + n2
    reinterpret_cast<float2*>(s_win)[n2 * 16 + n1] =
        make_float2(__ldg(prm.window + 2 * i), __ldg(prm.window + 2 * i + 1));
    // twiddle table entry i = residue * 16 + k1 (residue = i >> 4, k1 = i & 15): W256^(residue k1)
    float sn, cs;
    sincospif(2.0f * (float)(n1 * n2) / (float)kHalf, &sn, &cs);
    reinterpret_cast<float2*>(s_tw)[i] = make_float2(cs, -sn);
    sincospif(2.0f * (float)i / (float)kNfft, &sn, &cs);
    reinterpret_cast<float2*>(s_ut)[i] = make_float2(cs, sn);
  }
  for (int m = tid; !POWER && !COMPLEX && m < M; m += kThreads)
    s_meta[m] = make_int2(__ldg(prm.mel_start + m) | (__ldg(prm.mel_count + m) << 16),
                          __ldg(prm.mel_offset + m));
  for (int i = tid; i < prm.mel_nnz; i += kThreads) s_w[i] = __ldg(prm.mel_weight + i);
  for (int i = tid; i < (kPRows - kBins) * kTile; i += kThreads) s_a[kBins * kTile + i] = 0.f;
  pdl_wait();  // from here on the waveform is read and out / tile records are written

  float db_scale = prm.db.scale, db_amin = prm.db.amin, db_off = prm.db.offset,
        db_floor = prm.db.floor_db;
  asm volatile("" : "+f"(db_scale), "+f"(db_amin), "+f"(db_off), "+f"(db_floor));
  const bool log_mel = prm.db.log_mel != 0;
  float4* s_e4 = reinterpret_cast<float4*>(s_e);
  const int opitch = M | 1;
  float* s_o = s_e;
  const int st_t4 = (tid * 4) / M, st_m4 = tid * 4 - st_t4 * M;
  const int st_dt4 = (kThreads * 4) / M, st_dm4 = kThreads * 4 - st_dt4 * M;
  const int st_t1 = tid / M, st_m1 = tid - st_t1 * M;
  const int st_dt1 = kThreads / M, st_dm1 = kThreads - st_dt1 * M;

  auto mel_plan = [&](int mb, int (&G)[4], int (&W)[4], int (&P)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = mb + i * kUnits + ((i & 1) ? kUnits - 1 - unit : unit);
      const int2 meta = m < M ? s_meta[m] : make_int2(0, 0);
      G[i] = __reduce_max_sync(0xffffffffu, ((meta.x >> 16) + 3) >> 2);
      W[i] = meta.y;
      P[i] = (meta.x & 0xffff) * kPairs;
    }
  };
  int planG[4], planW[4], planP[4];
  __syncthreads();   // tables are complete
  if (!POWER && !COMPLEX) mel_plan(0, planG, planW, planP);

#pragma unroll 1
  for (unsigned tile = blockIdx.x; tile < prm.n_tiles; tile += gridDim.x) {
    const unsigned bu = tile / prm.tiles_per_utt;
    const unsigned ju = tile - bu * prm.tiles_per_utt;
    const int64_t b = bu;
    const int64_t t0 = (int64_t)ju * kTile;
    if (tile + gridDim.x >= prm.n_tiles) pdl_launch_dependents();
    // s_a is free: every thread passed the barrier behind the mel stage of the previous tile
    stage_tile(prm, bu, ju, s_a, tid);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    if (tid == 0) s_tile_key = 0u;
    __syncthreads();   // the span is in; the previous tile's readers of s_o (store, statistics) are done

    // ---- stage 1: unit = residue n2 ----------------------------------------------------------
    {
      const int n2 = unit;
      cx<f2> a[16];
      {
        const float* xa = s_a + p * kPitch2 + 2 * n2;
        float2 xs[21];
#pragma unroll
        for (int m = 0; m < 21; ++m)   // sample 32 m + 2 n2 of the pair's span: + 2 per 320 crossed
          xs[m] = *reinterpret_cast<const float2*>(xa + 32 * m + (m >= 10 ? kPad2 : 0) + (m >= 20 ? kPad2 : 0));
        const float2* w2 = reinterpret_cast<const float2*>(s_win) + n2 * 16;
#pragma unroll
        for (int n1 = 0; n1 < 16; ++n1) {
          const float2 w = w2[n1];
          a[n1] = cx<f2>{muls(mk2(xs[n1].x, xs[n1 + 5].x), w.x), muls(mk2(xs[n1].y, xs[n1 + 5].y), w.y)};
        }
      }
      cx<f2> Y[16];
      dft16(a, Y);
      const float2* tw = reinterpret_cast<const float2*>(s_tw) + n2 * 16;
      float4* dst = s_e4 + (n2 * 16) * kPairs + p;
      dst[0] = make_float4(Y[0].re.v.x, Y[0].re.v.y, Y[0].im.v.x, Y[0].im.v.y);
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) {
        const float2 t2 = tw[k1];
        const cx<f2> t = cmuls(Y[k1], t2.x, t2.y);
        dst[k1 * kPairs] = make_float4(t.re.v.x, t.re.v.y, t.im.v.x, t.im.v.y);
      }
    }
    __syncthreads();

    // ---- stage 2: unit = k1: DFT16 over n2, Z[k1 + 16 k2] back in place ---------------------
    const int k1 = unit;
    cx<f2> Z[16];
    {
      cx<f2> a[16];
      const float4* src = s_e4 + k1 * kPairs + p;
#pragma unroll
      for (int n2 = 0; n2 < 16; ++n2) {
        const float4 q = src[n2 * 16 * kPairs];
        a[n2] = cx<f2>{mk2(q.x, q.y), mk2(q.z, q.w)};
      }
      dft16(a, Z);
      float4* dst = s_e4 + k1 * kPairs + p;
#pragma unroll
      for (int k2 = 0; k2 < 16; ++k2)
        dst[k2 * 16 * kPairs] = make_float4(Z[k2].re.v.x, Z[k2].re.v.y, Z[k2].im.v.x, Z[k2].im.v.y);
    }
    __syncthreads();

    if constexpr (COMPLEX) {
      // ---- untangle in pairs, X in place (see the head of the kernel); X = (S - i W D) / 2 --------
      const float2* ut = reinterpret_cast<const float2*>(s_ut);
      const unsigned half_mask = h ? 0xffff0000u : 0x0000ffffu;   // a unit = one half-warp
      auto put = [&](int k, const f2 xr, const f2 xi) {
        s_e4[k * kPairs + ((p + k) & (kPairs - 1))] =
            make_float4(0.5f * xr.v.x, kImSign * 0.5f * xi.v.x, 0.5f * xr.v.y, kImSign * 0.5f * xi.v.y);
      };
#pragma unroll
      for (int k2 = 0; k2 < 8; ++k2) {
        const int k = k1 + 16 * k2;
        if (k == 0) {          // X[0] = Re Z0 + Im Z0 and X[256] = Re Z0 - Im Z0 share row 0 (both real)
          const f2 x0 = Z[0].re + Z[0].im, xn = Z[0].re - Z[0].im;
          s_e4[p] = make_float4(x0.v.x, xn.v.x, x0.v.y, xn.v.y);   // row 0 is nobody's partner row
          continue;
        }
        const int kp = kHalf - k;
        const float4 q = s_e4[kp * kPairs + p];             // Z[256 - k]; its conjugate is (br, -bi)
        __syncwarp(half_mask);                              // the half-warp has read the row: it is free
        const f2 br = mk2(q.x, q.y), bi = mk2(q.z, q.w);
        const f2 sr = Z[k2].re + br, si = Z[k2].im - bi;
        const f2 dr = Z[k2].re - br, di = Z[k2].im + bi;
        const float2 cs = ut[k];                            // V = W D, W = (c, -s)
        const f2 vr = fmas(dr, cs.x, muls(di, cs.y)), vi = fmas(di, cs.x, muls(dr, -cs.y));
        put(k, sr + vi, si - vr);                           // X[k]       = (S - i V) / 2
        put(kp, sr - vi, mk2(0.f, 0.f) - si - vr);          // X[256 - k] = (conj S - i conj V) / 2
      }
      if (k1 == 0) {           // k = 128 is its own partner
        const f2 sr = Z[8].re + Z[8].re, si = mk2(0.f, 0.f), di = Z[8].im + Z[8].im;
        const float2 cs = ut[128];                          // D = (0, 2 Im Z): V = W D
        const f2 vr = muls(di, cs.y), vi = muls(di, cs.x);
        put(128, sr + vi, si - vr);                         // row 128: read by nobody (Z[8] is in registers)
      }
      __syncthreads();
      // ---- store pass: frame t = row of 257 (re, im) pairs, 8-byte stores, lanes over bins -------
      const int64_t left = prm.T - t0;
      const int nvalid = left < kTile ? (int)left : kTile;
      float2* o2 = reinterpret_cast<float2*>(prm.out) + (b * prm.T + t0) * kBins;
      for (int i = tid; i < nvalid * kBins; i += kThreads) {
        const int t = i / kBins, k = i - t * kBins;
        const int q = t >> 1, odd = t & 1;
        float2 v;
        if (k == 0 || k == kHalf) {
          const float4 e = s_e4[q];
          v = make_float2(k == 0 ? (odd ? e.z : e.x) : (odd ? e.w : e.y), 0.f);
        } else {
          v = reinterpret_cast<const float2*>(s_e4 + k * kPairs + ((q + k) & (kPairs - 1)))[odd];
        }
        o2[i] = v;
      }
      continue;   // the barrier behind the next tile's staging orders these reads before its stage 1
    }
    // ---- untangle + power: bins k1 + 16 k2 (and 256 from Z[0]) --------------------------------
    {
      float2* prow = reinterpret_cast<float2*>(s_a) + p;   // row stride = 16 float2
      const float2* ut = reinterpret_cast<const float2*>(s_ut);
#pragma unroll
      for (int k2 = 0; k2 < 16; ++k2) {
        const int k = k1 + 16 * k2;
        const int kp = (kHalf - k) & (kHalf - 1);
        const float4 q = s_e4[kp * kPairs + p];
        const f2 br = mk2(q.x, q.y), bi = mk2(q.z, q.w);   // Z[256 - k]; its conjugate is (br, -bi)
        const f2 sr = Z[k2].re + br, si = Z[k2].im - bi;
        const f2 dr = Z[k2].re - br, di = Z[k2].im + bi;
        const float2 cs = ut[k];
        // X = S - i W D, W = (c, -s):  Xr = Sr + (c Di - s Dr),  Xi = Si - (c Dr + s Di)
        const f2 xr = fmas(di, cs.x, fmas(dr, -cs.y, sr));
        const f2 xi = fmas(dr, -cs.x, fmas(di, -cs.y, si));
        const f2 pw = muls(fmat(xi, xi, f2{__fmul2_rn(xr.v, xr.v)}), 0.25f);
        prow[k * kPairs] = pw.v;
      }
      if (k1 == 0) {   // Nyquist bin: X[256] = Re Z[0] - Im Z[0]
        const f2 x = Z[0].re - Z[0].im;
        prow[kHalf * kPairs] = __fmul2_rn(x.v, x.v);
      }
    }
    __syncthreads();

    // ---- stage 3: sparse mel projection + dB -> output tile [32][M] in smem ---------------------
    if constexpr (POWER) {
      // P[bin][32] -> tile [frame][bin], pitch 257 (odd): conflict-free reads and writes
      for (int i = tid; i < kBins * kTile; i += kThreads) s_o[(i & (kTile - 1)) * opitch + (i >> 5)] = s_a[i];
    } else {
      float vmax0 = -INFINITY, vmax1 = -INFINITY;
      const float2* pp0 = reinterpret_cast<const float2*>(s_a) + p;
#pragma unroll 1
      for (int mb = 0; mb < M; mb += 4 * kUnits) {
        int G[4];
        const float4* wt[4];
        const float2* pp[4];
        f2 acc[4];
        int W[4], P[4];
        if (mb == 0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) { G[i] = planG[i]; W[i] = planW[i]; P[i] = planP[i]; }
        } else {
          mel_plan(mb, G, W, P);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          wt[i] = reinterpret_cast<const float4*>(s_w + W[i]);
          pp[i] = pp0 + P[i];
          acc[i] = mk2(0.f, 0.f);
        }
#define SBFE_MEL_GROUP(i, g)                                                    \
  {                                                                             \
    const float4 w4 = wt[i][g];                                                 \
    acc[i] = fmas(f2{pp[i][(4 * (g) + 0) * kPairs]}, w4.x, acc[i]);             \
    acc[i] = fmas(f2{pp[i][(4 * (g) + 1) * kPairs]}, w4.y, acc[i]);             \
    acc[i] = fmas(f2{pp[i][(4 * (g) + 2) * kPairs]}, w4.z, acc[i]);             \
    acc[i] = fmas(f2{pp[i][(4 * (g) + 3) * kPairs]}, w4.w, acc[i]);             \
  }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (g < G[i]) SBFE_MEL_GROUP(i, g)
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll 1
          for (int g = 4; g < G[i]; ++g) SBFE_MEL_GROUP(i, g)
        }
#undef SBFE_MEL_GROUP
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int m = mb + i * kUnits + ((i & 1) ? kUnits - 1 - unit : unit);
          float v0 = acc[i].v.x, v1 = acc[i].v.y;
          if (log_mel) {
            const float l0 = fmaf(fast_log2(v0), db_scale, -db_off);
            const float l1 = fmaf(fast_log2(v1), db_scale, -db_off);
            v0 = (v0 > db_amin) ? l0 : db_floor;
            v1 = (v1 > db_amin) ? l1 : db_floor;
          }
          if (m < M) {
            s_o[(2 * p) * opitch + m] = v0;
            s_o[(2 * p + 1) * opitch + m] = v1;
            vmax0 = fmaxf(vmax0, v0);
            vmax1 = fmaxf(vmax1, v1);
          }
        }
      }
      const float vmax = fmaxf((t0 + 2 * p) < prm.T ? vmax0 : -INFINITY,
                               (t0 + 2 * p + 1) < prm.T ? vmax1 : -INFINITY);
      const unsigned key = vmax > -INFINITY ? max_key_encode(vmax) : 0u;
      const unsigned wk = __reduce_max_sync(0xffffffffu, key);
      if (lane == 0 && wk) {
        if (prm.tile_stats != nullptr) atomicMax(&s_tile_key, wk);
        if (prm.utt_max != nullptr) atomicMax(prm.utt_max + b, wk);
      }
    }
    __syncthreads();

    // ---- stage 4: store of the valid part of the tile + tile record -----------------------------
    const int64_t left = prm.T - t0;
    const int nvalid = left < kTile ? (int)left : kTile;
    {
      float* o = prm.out + (b * prm.T + t0) * M;
      const int total = nvalid * M;
      if ((M & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
        int e = tid * 4;
        int t = st_t4, m = st_m4;
        for (; e < total; e += kThreads * 4) {
          const float* s = s_o + t * opitch + m;
          *reinterpret_cast<float4*>(o + e) = make_float4(s[0], s[1], s[2], s[3]);
          t += st_dt4;
          m += st_dm4;
          if (m >= M) { m -= M; ++t; }
        }
      } else {
        int e = tid;
        int t = st_t1, m = st_m1;
        for (; e < total; e += kThreads) {
          o[e] = s_o[t * opitch + m];
          t += st_dt1;
          m += st_dm1;
          if (m >= M) { m -= M; ++t; }
        }
      }
    }
    if (prm.tile_stats != nullptr) {
      float* ts = prm.tile_stats + (size_t)tile * tile_rec_words(M);
      if (tid < kTileKeys) reinterpret_cast<unsigned*>(ts)[3 * M + tid] = tid == 0 ? s_tile_key : 0u;
      // thread = (column, quarter of the rows); rows past the utterance's end count as the
      // column's first value (neutral)
      const int q = tid & 3;
      const float inv = 1.f / (float)nvalid;
      for (int m0 = 0; m0 < M; m0 += kThreads / 4) {
        const int m = m0 + (tid >> 2);
        const bool live = m < M;
        const float* col = s_o + (live ? m : 0);
        const float x0 = col[0];
        const float* cq = col + q * (kTile / 4) * opitch;
        float s1 = 0.f, s2 = 0.f, mn = x0;
#pragma unroll
        for (int r = 0; r < kTile / 4; ++r) {
          const float v = (q * (kTile / 4) + r < nvalid) ? cq[r * opitch] : x0;
          const float d = v - x0;
          s1 += d;
          s2 = fmaf(d, d, s2);
          mn = fminf(mn, v);
        }
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
          s1 += __shfl_xor_sync(0xffffffffu, s1, o);
          s2 += __shfl_xor_sync(0xffffffffu, s2, o);
          mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        }
        if (live && q < 3)
          ts[q * M + m] = q == 0 ? fmaf(s1, inv, x0)
                                 : (q == 1 ? fmaxf(fmaf(-s1 * inv, s1, s2), 0.f) : mn);
      }
    }
  }
}

constexpr int kMaxSmem = 110 * 1024;  // two CTAs per SM

using KernelFn = void (*)(Params);
inline KernelFn pick_kernel(int n_mels) {
  switch (n_mels) {
    case 80: return fbank512_kernel<80>;
    case 40: return fbank512_kernel<40>;
    default: return fbank512_kernel<0>;
  }
}

}  // namespace v512

bool fbank512_supported(const sbfe_fbank_desc* d) {
  return d->n_fft == v512::kNfft && d->hop == v512::kHop && d->mel_pad4 == 2 && d->n_mels >= 1 &&
         d->n_mels <= 384 && d->mel_nnz >= 0 &&
         v512::smem_floats(d->n_mels, d->mel_nnz) * sizeof(float) <= (size_t)v512::kMaxSmem;
}

// the clamp launch of mode (b) is shared with the 400-point kernel (fbank400_v3.cu)
int launch_clamp_tiles(const sbfe_fbank_desc* d, int64_t B, int64_t T, float* out, const float* tile_stats,
                       float top_db, cudaStream_t stream);

static int launch_512(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                      int64_t wav_stride, float* out, sbfe_maxkey_t* utt_max, float* tile_stats,
                      float top_db, int power, cudaStream_t stream) {   // power: 1 = |X|^2, 2 = X
  using namespace v512;
  Params prm;
  prm.db = DbParams{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                    power ? 0 : d->log_mel};
  prm.n_mels = power == 2 ? 2 * kBins : (power ? kBins : d->n_mels);
  prm.mel_nnz = power ? 0 : d->mel_nnz;
  prm.L = L;
  prm.wav_stride = wav_stride;
  prm.T = 1 + L / kHop;
  prm.wav = wav;
  prm.window = d->d_window;
  prm.out = out;
  prm.utt_max = utt_max;
  prm.tile_stats = tile_stats;
  prm.mel_start = d->d_mel_start;
  prm.mel_count = d->d_mel_count;
  prm.mel_offset = d->d_mel_offset;
  prm.mel_weight = d->d_mel_weight;
  const int64_t tiles = (prm.T + kTile - 1) / kTile;
  if (tiles * B > 0x7fffffff || prm.T > 0x7fffffff) return SBFE_E_TOOBIG;
  prm.tiles_per_utt = (unsigned)tiles;
  prm.n_tiles = (unsigned)(tiles * B);
  const size_t smem = smem_floats(power == 2 ? 0 : prm.n_mels, prm.mel_nnz) * sizeof(float);   // COMPLEX: no tables
  static std::atomic<int> per_sm_cache[64];   // 0 -> value, idempotent (see fbank400_v3.cu)
  int dev = 0, sms = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return (int)e;
  const bool cached = dev >= 0 && dev < 64;
  int per_sm = cached ? per_sm_cache[dev].load(std::memory_order_acquire) : 0;
  if (per_sm == 0) {
    per_sm = 2;
    for (int mc : {0, 80, 40}) {
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(pick_kernel(mc), cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
      int n = 0;
      if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, pick_kernel(mc), kThreads, (size_t)kMaxSmem);
      if (e == cudaSuccess && n < per_sm) per_sm = n;
    }
    {
      int n = 0;
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(fbank512_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 kMaxSmem);
      if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fbank512_kernel<0, true>, kThreads,
                                                          (size_t)kMaxSmem);
      if (e == cudaSuccess && n < per_sm) per_sm = n;
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(fbank512_kernel<0, false, true>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
      if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fbank512_kernel<0, false, true>, kThreads,
                                                          (size_t)kMaxSmem);
      if (e == cudaSuccess && n < per_sm) per_sm = n;
    }
    if (e != cudaSuccess) return (int)e;
    if (per_sm < 1) return (int)cudaErrorLaunchOutOfResources;
    if (cached) per_sm_cache[dev].store(per_sm, std::memory_order_release);
  }
  const unsigned cap = (unsigned)(per_sm * sms);
  const unsigned grid = prm.n_tiles < cap ? prm.n_tiles : cap;
  KernelFn kern = power == 2 ? fbank512_kernel<0, false, true>
                             : (power ? fbank512_kernel<0, true> : pick_kernel(d->n_mels));
  int rc = launch_pdl(kern, dim3(grid), dim3(kThreads), (size_t)smem, stream, prm);
  if (rc == 0 && !power && tile_stats != nullptr && top_db >= 0.f)
    rc = launch_clamp_tiles(d, B, prm.T, out, tile_stats, top_db, stream);
  return rc;
}

int launch_fbank512(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                    int64_t wav_stride, float* out, sbfe_maxkey_t* utt_max, float* tile_stats,
                    float top_db, cudaStream_t stream) {
  return launch_512(d, wav, B, L, wav_stride, out, utt_max, tile_stats, top_db, 0, stream);
}

// STFT + power spectrum of the default mode at n_fft 512 / hop 160 (see fbank400_v3.cu)
bool stft512_power_supported(const sbfe_fbank_desc* d) {
  return d->n_fft == v512::kNfft && d->hop == v512::kHop &&
         v512::smem_floats(v512::kBins, 0) * sizeof(float) <= (size_t)v512::kMaxSmem;
}
// complex != 0: out [B, T, 257, 2]; else the power spectrum [B, T, 257]
int launch_stft512_power(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                         int64_t wav_stride, float* out, int complex, cudaStream_t stream) {
  return launch_512(d, wav, B, L, wav_stride, out, nullptr, nullptr, -1.f, complex ? 2 : 1, stream);
}

}  // namespace sbfe
