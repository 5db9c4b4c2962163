// v3 of the fused frame + window + real-FFT(400) + power + sparse mel + dB kernel
// (n_fft = 400, hop = 160: every recipe default).  Same contract as fbank400.cu:
//   STFT.forward          speechbrain/processing/features.py:133-188
//   spectral_magnitude    speechbrain/processing/features.py:327-356 (power=1)
//   Filterbank.forward    speechbrain/processing/features.py:490-551
//   _amplitude_to_DB      speechbrain/processing/features.py:701-702 (log part)
//
// What changed against v2 (profiles/r1_baseline.md: 31.5 M warp-instructions, only
// 31 % of them FP32; ALU pipe > FMA pipe):
//  * PACKED f32x2 MATH: a thread owns TWO frames (2p, 2p+1) of the 32-frame tile and
//    every add/mul/fma is one FADD2/FMUL2/FFMA2 on a 64-bit register pair.
//  * 20 EQUAL WORK UNITS IN BOTH FFT STAGES: thread = (frame pair p = lane & 15,
//    unit s = 2*warp + (lane >> 4)); stage 1 unit = residue n2, stage 2 unit =
//    (k1, even/odd half) after one radix-2 DIF step (fft_packed.cuh).  v2 left 9 of
//    20 role slots idle in stage 2.
//  * ROLE CONSTANTS ARE DATA, NOT CODE: window taps / twiddles of a unit come from
//    shared memory with 128-bit loads (one address per half-warp, broadcast), so the
//    two half-warps of a warp run identical instructions; all shared-memory offsets
//    are compile-time immediates (no index arithmetic in the hot loops).
//  * VECTOR EXCHANGE: stage 1 -> stage 2 goes through float4 {re0, re1, im0, im1}
//    entries; both halves of a warp read the same entry (broadcast).
#include <atomic>
#include <cmath>
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "fft_packed.cuh"

namespace sbfe {

namespace v3 {

#include "fbank400_tables.inc"

constexpr int kTile = 32;            // frames per tile
constexpr int kPairs = 16;           // frame pairs (2p, 2p + 1)
constexpr int kUnits = 20;           // work units per stage
constexpr int kWarps = kUnits / 2;
constexpr int kThreads = kWarps * 32;
constexpr int kHop = 160;
// staged waveform: hops (2q, 2q + 1) back to back, 2 floats of padding behind every PAIR of hops:
// sample i of the span sits at i + 2 * (i / 320).  A thread's frame pair is (2p, 2p + 1) - adjacent
// frames share 12 of their 20 samples per residue (hop = 8 * 20) - and lane p starts at p * 322:
// banks 2p (+1 for the other half-warp's residue), conflict-free scalar reads.
constexpr int kPad2 = 2;
constexpr int kPitch2 = 2 * kHop + kPad2;       // 322 floats per pair of hops
constexpr int kPass = 2 * kPitch2;              // 4 hops = one pass of the staging copy
constexpr int kNfft = 400;
constexpr int kBins = 201;
constexpr int kSpan = (kTile - 1) * kHop + kNfft;                // 5360 samples per tile
constexpr int kChunks = kSpan / 2;                                // 8-byte cp.async chunks
constexpr int kCpIters = (kChunks + kThreads - 1) / kThreads;    // 9
constexpr int kPRows = 204;          // power rows incl. 3 zero rows for padded mel reads
constexpr int kRegionA = kPRows * kTile;                          // >= padded span (5428)
constexpr int kRegionE = 200 * kPairs * 4;                        // exchange, float4 entries
constexpr int kTabW = 400, kTabTw = 400, kTabT2 = 40;
constexpr float kImSign = 1.f;       // COMPLEX: sign that turns stage 2's imaginary parts into torch.stft's

struct Params {
  DbParams db;
  int n_mels, mel_nnz;
  unsigned tiles_per_utt, n_tiles;
  int64_t L, wav_stride, T;
  const float* wav;        // float32 samples, or int16 PCM (PCM16 instantiation)
  float scale;             // PCM16: sample value = scale * int16 (folded into the window table)
  const int64_t* offsets;  // ragged input: utterance b starts at wav + offsets[b] ...
  const int32_t* nsamples; // ... and holds nsamples[b] <= L samples (the rest of [0, L) is zero)
  const float* window;     // device, [400], zero-padded / centred to n_fft
  float* out;
  sbfe_maxkey_t* utt_max;
  float* tile_stats;       // optional [n_tiles] records of tile_rec_words(M) (common.cuh)
  float top_db;            // mode (b): floor = utterance maximum - top_db (inf: no clamp)
  const int32_t* mel_start;
  const int32_t* mel_count;
  const int32_t* mel_offset;
  const float* mel_weight;
};

__host__ __device__ inline size_t smem_floats(int n_mels, int nnz) {
  return (size_t)2 * kRegionA + kRegionE + kTabW + kTabTw + kTabT2 + ((2 * n_mels + 3) & ~3) +
         ((nnz + 3) & ~3);
}

__device__ __forceinline__ void cp_async8(float* smem_dst, const float* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// asynchronous copy of one tile's waveform span into `dst` (layout above); samples
// outside [0, L) are zero-filled by the copy engine (src-size < cp-size).
// Chunk c = k*kThreads + tid covers samples i = 2c, 2c+1: with 80 chunks per hop and
// kThreads = 4*80 the thread's sample / smem offsets are base + k*{640, 648}: immediates.
// Returns true when the staging used the aligned chunk mapping (thread -> chunks k*kThreads + tid),
// which lets a thread look at its own copies before the barrier (all-zero tile detection).
// Everything but `thr_src` (= wav + ibase - n_fft/2, per thread, loop-invariant) is uniform
// over the CTA: the interior tile costs one 64-bit add per thread besides its copies.
template <bool PCM16>
__device__ __forceinline__ bool prefetch_tile(const Params& prm, unsigned b, unsigned j, float* dst,
                                              int tid, int ibase, int dbase) {
  static_assert(kThreads == 4 * (kHop / 2), "chunk mapping assumes 4 hops per pass");
  using S = typename std::conditional<PCM16, int16_t, float>::type;
  constexpr int kCp = PCM16 ? 4 : 8;                         // bytes per chunk (2 samples)
  const int64_t row0 = prm.offsets ? __ldg(prm.offsets + b) : (int64_t)b * prm.wav_stride;
  const int64_t Lb = prm.nsamples ? (int64_t)__ldg(prm.nsamples + b) : prm.L;
  const int64_t t0s = (int64_t)j * (kTile * kHop);            // first sample of frame t0, + n_fft/2
  const S* row = reinterpret_cast<const S*>(prm.wav) + row0;
  const bool al = (reinterpret_cast<uintptr_t>(row) & (kCp - 1)) == 0;
  S* d = reinterpret_cast<S*>(dst) + dbase;
  if (al && j != 0 && t0s + (kSpan - kNfft / 2) <= Lb) {     // interior tile: no guards at all
    const S* src = row + t0s + (ibase - kNfft / 2);
#pragma unroll
    for (int k = 0; k < kCpIters; ++k)
      if (k * kThreads + tid < kChunks) {
        if constexpr (PCM16) cp_async4(d + k * kPass, src + k * (4 * kHop), 4);
        else cp_async8(d + k * kPass, src + k * (4 * kHop), 8);
      }
    return true;
  }
  const int64_t g0 = t0s - kNfft / 2;                         // even
  const S* wsrc = row + g0;                                   // dereferenced inside [lo, hi) only
  const int lo = g0 < 0 ? (int)(-g0) : 0;                     // 0 or 200
  const int64_t rem = Lb - g0;
  const int hi = rem <= 0 ? 0 : (rem < kSpan ? (int)rem : kSpan);
  if (al) {
#pragma unroll 1
    for (int k = 0; k < kCpIters; ++k) {
      if (k * kThreads + tid < kChunks) {
        const int i = ibase + k * (4 * kHop);
        int n = hi - i;
        n = (i < lo || n < 0) ? 0 : (n > 2 ? 2 : n);
        const void* g = n ? (const void*)(wsrc + i) : (const void*)prm.window;
        if constexpr (PCM16) cp_async4(d + k * kPass, g, 2 * n);
        else cp_async8(d + k * kPass, reinterpret_cast<const float*>(g), 4 * n);
      }
    }
  } else if constexpr (PCM16) {
    // odd sample offset: plain 16-bit copies (ordered by the barrier that opens the tile)
    int16_t* d16 = reinterpret_cast<int16_t*>(dst);
    for (int i = tid; i < kSpan; i += kThreads)
      d16[i + (i / (2 * kHop)) * kPad2] = (i >= lo && i < hi) ? wsrc[i] : (int16_t)0;
  } else {
    for (int i = tid; i < kSpan; i += kThreads) {
      const bool ok = i >= lo && i < hi;
      cp_async4(dst + i + (i / (2 * kHop)) * kPad2, ok ? (const void*)(wsrc + i) : (const void*)prm.window,
                ok ? 4 : 0);
    }
  }
  return al;
}


// ---- top_db clamp of the finished utterances (mode (b)) ---------------------------------------
// Reference semantics: x = max(x, max(x[b]) - top_db) over the whole padded utterance
// (processing/features.py:706-710).  The floor of utterance b exists once its last tile has been
// produced, so the clamp is a second, small launch of the same call (clamp_tiles_kernel below):
// it reads the tile records only (keys -> floor, column minima -> "does this tile hold anything
// below it") and refills constant tiles (the zero padding behind an utterance) without reading
// them.  The records keep the moments of the UN-clamped values; a consumer trusts a record's
// column only when its minimum is not below the floor (norm.cu).  Finishing the utterances inside the producing launch instead (every CTA clamping its
// own tiles once a per-utterance counter says the utterance is complete) was built and measured:
// the hand-shake needs >= 3 dependent L2 round trips behind the last tile (maximum, count, poll)
// plus one for the counters' reset, 45.9 us against 43.9 us for the two launches on the cfg2
// batch (profiles/r2_inlaunch_finish.md) - dropped.

// Clamp tile j of utterance b at floor_b in place.  The tile's record is left alone: its moments
// stay those of the un-clamped values, and every consumer checks the record's column minimum
// against the floor before it trusts them (norm.cu) - so a reader that runs beside this kernel
// never meets a half-written record.  Whole CTA calls; no barrier inside.
template <int MC, int NT>
__device__ __forceinline__ void clamp_tile(const Params& prm, unsigned b, unsigned j, float floor_b,
                                           bool constant, int tid) {
  const int M = MC > 0 ? MC : prm.n_mels;
  const int64_t r_lo = (int64_t)j * kTile;
  const int rows = prm.T - r_lo < kTile ? (int)(prm.T - r_lo) : kTile;
  float* p = prm.out + ((int64_t)b * prm.T + r_lo) * M;
  const int total = rows * M;
  const bool vec = (M & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0;
  if (constant) {                   // every value of the tile is the dB floor (all-zero input)
    const float v = fmaxf(prm.db.log_mel ? prm.db.floor_db : 0.f, floor_b);
    if (vec) {
      for (int e = tid * 4; e < total; e += NT * 4)
        *reinterpret_cast<float4*>(p + e) = make_float4(v, v, v, v);
    } else {
      for (int e = tid; e < total; e += NT) p[e] = v;
    }
    return;
  }
  if (vec) {
    for (int e = tid * 4; e < total; e += NT * 4) {
      float4 v = __ldcg(reinterpret_cast<const float4*>(p + e));
      if (fminf(fminf(v.x, v.y), fminf(v.z, v.w)) < floor_b) {
        v.x = fmaxf(v.x, floor_b);
        v.y = fmaxf(v.y, floor_b);
        v.z = fmaxf(v.z, floor_b);
        v.w = fmaxf(v.w, floor_b);
        *reinterpret_cast<float4*>(p + e) = v;
      }
    }
  } else {
    for (int e = tid; e < total; e += NT) {
      const float v = __ldcg(p + e);
      if (v < floor_b) p[e] = floor_b;
    }
  }
}

#ifndef SBFE_EXP_MINB
#define SBFE_EXP_MINB 2
#endif
// MC > 0: n_mels is the compile-time constant MC (every M-dependent index, walk and trip count
// folds away: the recipes' 80 / 40 / 23 filters); MC == 0: any n_mels at run time
// POWER: the kernel is the STFT + spectral_magnitude(power=1) of the default mode (features.py:
// 133-188, 327-356): the tile [32 frames][201 bins] of the power spectrum leaves through the same
// output path, no filterbank (M = 201, no tables, no records)
// COMPLEX: the STFT itself, out [B, T, 201, 2].  Stage 2 writes X[k] = (re, im) of its frame pair IN
// PLACE over the exchange entries its own warp has just read (a warp's 20 rows hold exactly its 20
// bins x 16 pairs), rotated inside the row so that the store pass below reads them conflict-free.
template <bool PCM16, int MC, bool POWER = false, bool COMPLEX = false>
__global__ void __launch_bounds__(kThreads, SBFE_EXP_MINB) fbank400_v3_kernel(const __grid_constant__ Params prm) {
  extern __shared__ __align__(16) float smem[];
  float* s_a = smem;                       // 2 x [staged waveform | P[bin][32]]
  float* s_e = s_a + 2 * kRegionA;         // exchange, later the output tile
  float* s_tw = s_e + kRegionE;            // [n2][20]: window taps of residue n2 (n1 order)
  float* s_tt = s_tw + kTabW;              // [n2][20]: (re, im) of W400^(k1*n2), k1 = 1..9, 2 pad
  float* s_t2 = s_tt + kTabTw;             // [h][20]
  int2* s_meta = reinterpret_cast<int2*>(s_t2 + kTabT2);
  const int M = COMPLEX ? 2 * kBins : (POWER ? kBins : (MC > 0 ? MC : prm.n_mels));   // floats per output row
  float* s_w = reinterpret_cast<float*>(s_meta) + ((2 * M + 3) & ~3);  // 16-byte aligned
  int* s_row = reinterpret_cast<int*>(s_meta);   // COMPLEX: exchange row that holds bin k

  __shared__ unsigned s_tile_key;  // max key of the current tile (tile records)
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int p = lane & 15, h = lane >> 4;
  const int unit = 2 * warp + h;

  // ---- once per CTA: the constant tables (they are not produced by the preceding kernel, so
  // this part may overlap its tail: see pdl_wait in common.cuh), then the first prefetch ------
  const int ibase = kHop * (tid / (kHop / 2)) + 2 * (tid % (kHop / 2));
  const int dbase = ibase + kPad2 * (tid / kHop);   // tid / 160 = hop pair of the pass
  unsigned tile = blockIdx.x;
  // every table word this thread is responsible for is REQUESTED first and stored afterwards:
  // one global-memory round trip for the whole prologue instead of one per table
  {
    static_assert(kTabW <= 2 * kThreads && kTabT2 <= kThreads, "two table words per thread");
    const bool second = tid + kThreads < kTabW;
    const float w0 = __ldg(prm.window + tid), w1 = second ? __ldg(prm.window + tid + kThreads) : 0.f;
    const float c0 = kTw400[tid], c1 = second ? kTw400[tid + kThreads] : 0.f;
    const float t2 = tid < kTabT2 ? kT2[tid] : 0.f;
    int ms = 0, mc = 0, mo = 0;
    if (!POWER && !COMPLEX && tid < M) {
      ms = __ldg(prm.mel_start + tid);
      mc = __ldg(prm.mel_count + tid);
      mo = __ldg(prm.mel_offset + tid);
    }
    const float g0 = tid < prm.mel_nnz ? __ldg(prm.mel_weight + tid) : 0.f;
    const float g1 = tid + kThreads < prm.mel_nnz ? __ldg(prm.mel_weight + tid + kThreads) : 0.f;
    auto put_w = [&](int i, float w, float c) {
      const int n1 = i / 20, n2 = i - n1 * 20;   // window[i] is tap n1 of residue n2
      s_tw[n2 * 20 + n1] = PCM16 ? w * prm.scale : w;
      s_tt[i] = c;
    };
    put_w(tid, w0, c0);
    if (second) put_w(tid + kThreads, w1, c1);
    if (tid < kTabT2) s_t2[tid] = t2;
    if (!POWER && !COMPLEX && tid < M) s_meta[tid] = make_int2(ms | (mc << 16), mo);
    if (tid < prm.mel_nnz) s_w[tid] = g0;
    if (tid + kThreads < prm.mel_nnz) s_w[tid + kThreads] = g1;
  }
  for (int m = tid + kThreads; !POWER && !COMPLEX && m < M; m += kThreads)       // more than 320 filters
    s_meta[m] = make_int2(__ldg(prm.mel_start + m) | (__ldg(prm.mel_count + m) << 16),
                          __ldg(prm.mel_offset + m));
  for (int i = tid + 2 * kThreads; i < prm.mel_nnz; i += kThreads) s_w[i] = __ldg(prm.mel_weight + i);
  if (tid < 2 * (kPRows - kBins) * kTile) {
    const int r = tid / ((kPRows - kBins) * kTile), i = tid - r * (kPRows - kBins) * kTile;
    s_a[r * kRegionA + kBins * kTile + i] = 0.f;
  }
  SBFE_TL(0, 0);
  pdl_wait();  // from here on the waveform is read and out / tile records are written
  SBFE_TL(0, 1);
  // (utterance, first frame) of the tile in flight: one division per tile, carried over
  unsigned nb = tile / prm.tiles_per_utt;
  unsigned nj = tile - nb * prm.tiles_per_utt;               // tile index inside the utterance
  const unsigned step_b = gridDim.x / prm.tiles_per_utt;     // a grid stride in (utterance, tile)
  const unsigned step_j = gridDim.x - step_b * prm.tiles_per_utt;
  int nt0 = (int)nj * kTile;
  bool chunked = false, nxt_chunked = false;   // staging mode of the current / next tile (uniform)
  if (tile < prm.n_tiles) nxt_chunked = prefetch_tile<PCM16>(prm, nb, nj, s_a, tid, ibase, dbase);
  cp_async_commit();
  float db_scale = prm.db.scale, db_amin = prm.db.amin, db_off = prm.db.offset,
        db_floor = prm.db.floor_db;
  // keep them in registers (the compiler otherwise re-reads the constant bank under predicates)
  asm volatile("" : "+f"(db_scale), "+f"(db_amin), "+f"(db_off), "+f"(db_floor));
  const bool log_mel = prm.db.log_mel != 0;
  float4* s_e4 = reinterpret_cast<float4*>(s_e);
  const int opitch = M | 1;
  float* s_o = s_e;
  // stage-4 walk of this thread over the [frames][M] tile (vector / scalar), tile-invariant
  const int st_t4 = (tid * 4) / M, st_m4 = tid * 4 - st_t4 * M;
  const int st_dt4 = (kThreads * 4) / M, st_dm4 = kThreads * 4 - st_dt4 * M;
  const int st_t1 = tid / M, st_m1 = tid - st_t1 * M;
  const int st_so_lin = tid + st_t1;   // M | 1 == M + 1 for the even compile-time sizes
  const int st_dt1 = kThreads / M, st_dm1 = kThreads - st_dt1 * M;

  // stage-3 plan of this thread (tile-invariant): its four filters' group counts, weight and
  // power-row offsets.  Odd slots run backwards over the units: filters widen with m, so unit u
  // gets (narrow, wide, narrow, wide)-ish work and the warps reach the barrier together.  The
  // warp's two filters (2k, 2k+1) store the same number of 4-weight groups (mel_pad4 == 2), so
  // the trip count is warp-uniform (a uniform register): no divergence bookkeeping.
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
  __syncthreads();   // s_meta is complete
  if (!POWER && !COMPLEX) mel_plan(0, planG, planW, planP);

#pragma unroll 1
  for (int it = 0; tile < prm.n_tiles; ++it, tile += gridDim.x) {
    float* s_wave = s_a + (it & 1) * kRegionA;
    const int64_t b = nb;
    const int64_t t0 = nt0;
    chunked = nxt_chunked;
    {  // next tile's waveform flies while this one is transformed
      const unsigned nxt = tile + gridDim.x;
      if (nxt >= prm.n_tiles) pdl_launch_dependents();  // last tile of this CTA: let the next kernel in
      nb += step_b;
      nj += step_j;
      if (nj >= prm.tiles_per_utt) { nj -= prm.tiles_per_utt; ++nb; }
      nt0 = (int)nj * kTile;
#ifndef SBFE_EXP_NOPREFETCH
      if (nxt < prm.n_tiles)
        nxt_chunked = prefetch_tile<PCM16>(prm, nb, nj, s_a + ((it & 1) ^ 1) * kRegionA, tid, ibase, dbase);
#endif
      cp_async_commit();
    }
    cp_async_wait<1>();
    // ---- all-zero span (padding behind the utterance, digital silence): the transform of zeros
    // is zero, every output is exactly the floor (or 0 for linear mel) - write that and move on.
    // Each thread probes the first chunk it staged itself (visible to it after wait_group); the
    // vote rides on the barrier that opens the tile; only a tile whose probes are all zero pays
    // for the full check.  Bit-identical to the full path: 0 * w accumulates to +0, +0 <= amin.
    {
      auto chunk_nonzero = [&](int k) -> int {
        if constexpr (PCM16) {
          return *reinterpret_cast<const int*>(reinterpret_cast<const int16_t*>(s_wave) + dbase +
                                               k * kPass) != 0;
        } else {
          const float2 v = *reinterpret_cast<const float2*>(s_wave + dbase + k * kPass);
          return (v.x != 0.f) | (v.y != 0.f);
        }
      };
      int nz = chunked ? chunk_nonzero(0) : 1;
      // the barrier: this tile's span (all threads' copies) is in smem; s_e / s_o are free
      const int any_nz = __syncthreads_or(nz);
      if (any_nz == 0) {
        nz = 0;
#pragma unroll 1
        for (int k = 1; k < kCpIters; ++k)
          if (k * kThreads + tid < kChunks) nz |= chunk_nonzero(k);
        if (__syncthreads_or(nz) == 0) {
          const int64_t left = prm.T - t0;
          const int nvalid = left < kTile ? (int)left : kTile;
          const float v = log_mel ? db_floor : 0.f;
          float* o = prm.out + (b * prm.T + t0) * M;
          const int total = nvalid * M;
          if ((M & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
            for (int e = tid * 4; e < total; e += kThreads * 4)
              *reinterpret_cast<float4*>(o + e) = make_float4(v, v, v, v);
          } else {
            for (int e = tid; e < total; e += kThreads) o[e] = v;
          }
          const unsigned key = max_key_encode(v);
          if (prm.tile_stats != nullptr) {
            float* ts = prm.tile_stats + (size_t)tile * tile_rec_words(M);
            for (int m = tid; m < M; m += kThreads) {
              ts[m] = v;
              ts[M + m] = 0.f;
              ts[2 * M + m] = v;
            }
            // word 1 of the keys: flags, bit 0 = constant tile (the clamp launch refills it unread)
            if (tid < kTileKeys)
              reinterpret_cast<unsigned*>(ts)[3 * M + tid] = tid == 0 ? key : (tid == 1 ? 1u : 0u);
          }
          if (prm.utt_max != nullptr && tid == 0) atomicMax(prm.utt_max + b, key);
          continue;
        }
      }
    }

    if (tid == 0) s_tile_key = 0u;  // two barriers before stage 3 raises it, read after the third

    // ---- stage 1: unit = residue n2: windowed real DFT20 over n1, twiddle, exchange --------
    {
      const int n2 = unit;
      float wv[20], tw[20];
      {
        const float4* w4 = reinterpret_cast<const float4*>(s_tw + n2 * 20);
        const float4* t4 = reinterpret_cast<const float4*>(s_tt + n2 * 20);
#pragma unroll
        for (int q = 0; q < 5; ++q) {
          const float4 a = w4[q], c = t4[q];
          wv[4 * q] = a.x; wv[4 * q + 1] = a.y; wv[4 * q + 2] = a.z; wv[4 * q + 3] = a.w;
          tw[4 * q] = c.x; tw[4 * q + 1] = c.y; tw[4 * q + 2] = c.z; tw[4 * q + 3] = c.w;
        }
      }
      // frames (2p, 2p + 1): tap n1 of the second is sample 20 (n1 + 8) + n2 of the first's span,
      // so 28 loads serve both (not 40)
      f2 x[20];
      {
        float xs[28];
        if constexpr (PCM16) {   // int16 -> float is exact; the 2^-15 scale sits in the window taps
          const int16_t* xa = reinterpret_cast<const int16_t*>(s_wave) + p * kPitch2 + n2;
#pragma unroll
          for (int m = 0; m < 28; ++m) xs[m] = (float)xa[20 * m + (m >= 16 ? kPad2 : 0)];
        } else {
          const float* xa = s_wave + p * kPitch2 + n2;
#pragma unroll
          for (int m = 0; m < 28; ++m) xs[m] = xa[20 * m + (m >= 16 ? kPad2 : 0)];  // 20 m + n2 >= 320 <=> m >= 16
        }
#pragma unroll
        for (int n1 = 0; n1 < 20; ++n1) x[n1] = mk2(xs[n1], xs[n1 + 8]);
      }
      cx<f2> Y[11];
      rdft20<true>(x, wv, Y);
      float4* dst = s_e4 + (n2 * 10) * kPairs + p;
      dst[0] = make_float4(Y[0].re.v.x, Y[0].re.v.y, Y[10].re.v.x, Y[10].re.v.y);
#pragma unroll
      for (int k1 = 1; k1 <= 9; ++k1) {
        const cx<f2> t = cmuls(Y[k1], tw[2 * (k1 - 1)], tw[2 * (k1 - 1) + 1]);
        dst[k1 * kPairs] = make_float4(t.re.v.x, t.re.v.y, t.im.v.x, t.im.v.y);
      }
    }
    __syncthreads();

#ifdef SBFE_EXP_SKIP234
    if (s_e[tid] == 1.2345e-30f) prm.out[0] = 1.f;
    continue;
#endif
    // ---- stage 2: unit = (k1, half): DFT over n2, power -> P[bin][2p .. 2p+1] ---------------
    float* s_p = s_wave;
    {
      float2* prow = reinterpret_cast<float2*>(s_p) + p;  // row stride = 16 float2
      if (warp != 0) {
        const int k1 = warp;
        float tr[10], ti[10];
        {
          const float4* c4 = reinterpret_cast<const float4*>(s_t2 + h * 20);
          float tmp[20];
#pragma unroll
          for (int q = 0; q < 5; ++q) {
            const float4 a = c4[q];
            tmp[4 * q] = a.x; tmp[4 * q + 1] = a.y; tmp[4 * q + 2] = a.z; tmp[4 * q + 3] = a.w;
          }
#pragma unroll
          for (int n = 0; n < 10; ++n) { tr[n] = tmp[n]; ti[n] = tmp[10 + n]; }
        }
        const float sigma = h ? -1.f : 1.f;
        const float4* src = s_e4 + k1 * kPairs + p;
        cx<f2> a[20];
#pragma unroll
        for (int n2 = 0; n2 < 20; ++n2) {
          const float4 q = src[n2 * 10 * kPairs];
          a[n2] = cx<f2>{mk2(q.x, q.y), mk2(q.z, q.w)};
        }
        cx<f2> C[10];
        stage2_half(a, sigma, tr, ti, C);
        // bin of output j: k = k1 + 20h + 40j, mirrored above 200
        if constexpr (COMPLEX) {
          __syncwarp();   // every lane of the warp has read its exchange entries: the rows are free
#pragma unroll
          for (int j = 0; j < 10; ++j) {
            const int kk = k1 + 20 * h + 40 * j;
            const int bin = j < 5 ? kk : kNfft - kk;          // X[400 - k] = conj X[k]
            const float sg = (j < 5 ? 1.f : -1.f) * kImSign;
            const int row = (10 * h + j) * 10 + k1;
            s_e4[row * kPairs + ((p + row) & (kPairs - 1))] =
                make_float4(C[j].re.v.x, sg * C[j].im.v.x, C[j].re.v.y, sg * C[j].im.v.y);
            if (p == 0) s_row[bin] = row;
          }
        } else {
        float2* pr = prow + (k1 + 20 * h) * kPairs;
#pragma unroll
        for (int j = 0; j < 10; ++j) {
          const f2 pw = fmat(C[j].im, C[j].im, f2{__fmul2_rn(C[j].re.v, C[j].re.v)});
          if (j < 5) pr[40 * j * kPairs] = pw.v;
        }
        float2* pm = prow + (kNfft - k1 - 20 * h) * kPairs;
#pragma unroll
        for (int j = 5; j < 10; ++j) {
          const f2 pw = fmat(C[j].im, C[j].im, f2{__fmul2_rn(C[j].re.v, C[j].re.v)});
          pm[-40 * j * kPairs] = pw.v;
        }
        }
      } else {
        const float4* src = s_e4 + p;
        f2 y[20];
#pragma unroll
        for (int n2 = 0; n2 < 20; ++n2) {
          const float4 q = src[n2 * 10 * kPairs];
          y[n2] = h ? mk2(q.z, q.w) : mk2(q.x, q.y);
        }
        if constexpr (COMPLEX) __syncwarp();   // both halves have read rows 20 n2: they are free
        if (h == 0) {  // k1 = 0: bins 20*k2
          const float none[20] = {};
          cx<f2> X[11];
          rdft20<false>(y, none, X);
          if constexpr (COMPLEX) {
            // bins 0 and 200 are real: they share row 0 as (X0, X200) of the two frames
            s_e4[p] = make_float4(X[0].re.v.x, X[10].re.v.x, X[0].re.v.y, X[10].re.v.y);
#pragma unroll
            for (int k2 = 1; k2 <= 9; ++k2) {
              const int row = k2 * 10;
              s_e4[row * kPairs + ((p + row) & (kPairs - 1))] = make_float4(
                  X[k2].re.v.x, kImSign * X[k2].im.v.x, X[k2].re.v.y, kImSign * X[k2].im.v.y);
              if (p == 0) s_row[20 * k2] = row;
            }
          } else {
          prow[0] = __fmul2_rn(X[0].re.v, X[0].re.v);
          prow[200 * kPairs] = __fmul2_rn(X[10].re.v, X[10].re.v);
#pragma unroll
          for (int k2 = 1; k2 <= 9; ++k2)
            prow[20 * k2 * kPairs] =
                fmat(X[k2].im, X[k2].im, f2{__fmul2_rn(X[k2].re.v, X[k2].re.v)}).v;
          }
        } else {  // k1 = 10: bins 10 + 20*k2
          cx<f2> C[10];
          stage2_k10(y, C);
          if constexpr (COMPLEX) {
#pragma unroll
            for (int j = 0; j < 10; ++j) {
              const int bin = j < 5 ? 10 + 40 * j : 390 - 40 * j;
              const float sg = (j < 5 ? 1.f : -1.f) * kImSign;
              const int row = (10 + j) * 10;
              s_e4[row * kPairs + ((p + row) & (kPairs - 1))] =
                  make_float4(C[j].re.v.x, sg * C[j].im.v.x, C[j].re.v.y, sg * C[j].im.v.y);
              if (p == 0) s_row[bin] = row;
            }
          } else {
#pragma unroll
          for (int j = 0; j < 10; ++j) {
            const int bin = j < 5 ? 10 + 40 * j : 390 - 40 * j;
            prow[bin * kPairs] = fmat(C[j].im, C[j].im, f2{__fmul2_rn(C[j].re.v, C[j].re.v)}).v;
          }
          }
        }
      }
    }
    __syncthreads();

#ifdef SBFE_EXP_SKIP34
    if (s_p[tid] == 1.2345e-30f) prm.out[0] = 1.f;
    continue;
#endif
    if constexpr (COMPLEX) {
      // ---- store pass: frame t = row of 201 (re, im) pairs, 8-byte stores, lanes over bins.
      // Entry of bin k, pair q: row s_row[k], slot (q + row) & 15 - 16 consecutive bins of a warp
      // sit in 16 different slots
      const int64_t left = prm.T - t0;
      const int nvalid = left < kTile ? (int)left : kTile;
      float2* o2 = reinterpret_cast<float2*>(prm.out) + (b * prm.T + t0) * kBins;
      for (int i = tid; i < nvalid * kBins; i += kThreads) {
        const int t = i / kBins, k = i - t * kBins;
        const int q = t >> 1, odd = t & 1;
        float2 v;
        if (k == 0 || k == kNfft / 2) {
          const float4 e = s_e4[q];
          v = make_float2(k == 0 ? (odd ? e.z : e.x) : (odd ? e.w : e.y), 0.f);
        } else {
          const int row = s_row[k];
          v = reinterpret_cast<const float2*>(s_e4 + row * kPairs + ((q + row) & (kPairs - 1)))[odd];
        }
        o2[i] = v;
      }
      continue;   // the barrier at the top of the next iteration orders these reads before its stage 1
    }
    // ---- stage 3: sparse mel projection + dB -> output tile [32][M] in smem -----------------
    if constexpr (POWER) {
      // P[bin][32] -> tile [frame][bin], pitch 201: a warp reads 32 consecutive words of one bin
      // and writes 32 distinct banks (201 = 9 mod 32)
      for (int i = tid; i < kBins * kTile; i += kThreads) s_o[(i & (kTile - 1)) * opitch + (i >> 5)] = s_p[i];
    } else {
      float vmax0 = -INFINITY, vmax1 = -INFINITY;
      const float2* pp0 = reinterpret_cast<const float2*>(s_p) + p;
      // four filters (m, m+20, m+40, m+60) in flight per thread: their loads and FMA chains
      // are independent, which hides the shared-memory latency a one-filter loop exposes
#pragma unroll 1
      for (int mb = 0; mb < M; mb += 4 * kUnits) {   // block-uniform trip count
        int G[4];
        const float4* wt[4];
        const float2* pp[4];
        f2 acc[4];
        int W[4], P[4];
        if (mb == 0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) { G[i] = planG[i]; W[i] = planW[i]; P[i] = planP[i]; }
        } else {
          mel_plan(mb, G, W, P);   // M > 80: the later passes plan on the fly
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
          if (log_mel) {  // branch-free: log2 of everything, select afterwards
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
      {
        const float vmax = fmaxf((t0 + 2 * p) < prm.T ? vmax0 : -INFINITY,
                                 (t0 + 2 * p + 1) < prm.T ? vmax1 : -INFINITY);
        const unsigned key = vmax > -INFINITY ? max_key_encode(vmax) : 0u;
        // one warp reduction, then one atomic per warp: the tile maximum via a shared word
        // (stored with the tile record after the barrier) and / or the utterance maximum
        const unsigned wk = __reduce_max_sync(0xffffffffu, key);
        if (lane == 0 && wk) {
          if (prm.tile_stats != nullptr) atomicMax(&s_tile_key, wk);
          if (prm.utt_max != nullptr) atomicMax(prm.utt_max + b, wk);
        }
      }
    }
    __syncthreads();

#ifdef SBFE_EXP_SKIP4
    if (s_o[tid] == 1.2345e-30f) prm.out[0] = 1.f;
    continue;
#endif
    // ---- stage 4: coalesced store of the valid part of the tile ---------------------------------
    {
      const int64_t left = prm.T - t0;
      const int nvalid = left < kTile ? (int)left : kTile;
      float* o = prm.out + (b * prm.T + t0) * M;
      const int total = nvalid * M;
      if (MC > 0 && kThreads % MC == 0) {
        // a pass of the CTA covers whole rows: element e = tid + k * kThreads sits at smem offset
        // e + e / M = (tid + tid / M) + k * (kThreads + kThreads / M).  One float per lane:
        // consecutive lanes read consecutive banks (the row pitch M | 1 only shifts by one per
        // row) and a warp stores 128 contiguous bytes - the float4 walk below reads its four
        // floats at a lane stride of 4 words, 2.75 wavefronts per load on this tile shape
        constexpr int kRowsPass = MC > 0 ? kThreads / MC : 1;
        constexpr int kPasses = MC > 0 ? (kTile * MC) / kThreads : 1;
        const float* s = s_o + st_so_lin;
#pragma unroll
        for (int k = 0; k < kPasses; ++k)
          if (tid + k * kThreads < total) o[tid + k * kThreads] = s[k * (kThreads + kRowsPass)];
      } else if ((M & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
        int e = tid * 4;
        int t = st_t4, m = st_m4;
        const int dt = st_dt4, dm = st_dm4;
        for (; e < total; e += kThreads * 4) {
          const float* s = s_o + t * opitch + m;
          *reinterpret_cast<float4*>(o + e) = make_float4(s[0], s[1], s[2], s[3]);
          t += dt;
          m += dm;
          if (m >= M) { m -= M; ++t; }
        }
      } else {
        int e = tid;
        int t = st_t1, m = st_m1;
        const int dt = st_dt1, dm = st_dm1;
        for (; e < total; e += kThreads) {
          o[e] = s_o[t * opitch + m];
          t += dt;
          m += dm;
          if (m >= M) { m -= M; ++t; }
        }
      }
    }
    // ---- per-tile column statistics (un-clamped values) for InputNormalization --------------------
    // mean / M2 (sum of squared deviations) / min over the tile's existing frames, so that the
    // normalisation needs no second pass over the features (norm.cu: tile_stats_block, norm_sentence_tiles_kernel)
    if (prm.tile_stats != nullptr) {
      if (tid < kTileKeys)
        reinterpret_cast<unsigned*>(prm.tile_stats)[(size_t)tile * tile_rec_words(M) + 3 * M + tid] =
            tid == 0 ? s_tile_key : 0u;
      const int64_t left = prm.T - t0;
      const int nvalid = left < kTile ? (int)left : kTile;
      float* ts = prm.tile_stats + (size_t)tile * tile_rec_words(M);
      // thread = (column, quarter of the rows): 4 partial sums about the column's first value,
      // merged with two shuffles; lanes (8 columns x 4 quarters) hit 32 distinct banks.
      // FULL (all 32 frames exist: every tile but an utterance's last) carries no row guards.
      const int q = tid & 3;
      auto column_stats = [&](auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
        const float inv = FULL ? 1.f / kTile : 1.f / (float)nvalid;
        for (int m0 = 0; m0 < M; m0 += kThreads / 4) {
          const int m = m0 + (tid >> 2);
          const bool live = m < M;
          const float* col = s_o + (live ? m : 0);
          const float x0 = col[0];
          const float* cq = col + q * (kTile / 4) * opitch;
          float v[kTile / 4];
#pragma unroll
          for (int r = 0; r < kTile / 4; ++r)
            v[r] = (FULL || q * (kTile / 4) + r < nvalid) ? cq[r * opitch] : x0;  // x0: neutral
          float s1 = 0.f, s2 = 0.f, mn = x0;
#pragma unroll
          for (int r = 0; r < kTile / 4; ++r) {
            const float d = v[r] - x0;
            s1 += d;
            s2 = fmaf(d, d, s2);
            mn = fminf(mn, v[r]);
          }
#pragma unroll
          for (int o = 1; o <= 2; o <<= 1) {
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
          }
          if (live && q < 3) {
            const float r = q == 0 ? fmaf(s1, inv, x0)
                                   : (q == 1 ? fmaxf(fmaf(-s1 * inv, s1, s2), 0.f) : mn);
            ts[q * M + m] = r;
          }
        }
      };
      if (nvalid == kTile) column_stats(std::true_type{});
      else column_stats(std::false_type{});
    }
    // the barrier at the top of the next iteration orders these s_o reads before its stage 1
  }
  cp_async_wait<0>();
  SBFE_TL(0, 2);
}

// grid = B x ceil(tiles / 4) CTAs of kClampThreads, four tiles each (the keys are read once per CTA).
// The utterance's keys and the tile's own column minima / flags are requested together (one
// round trip); a tile with nothing below the floor costs nothing else.  The kernel lets its
// successor start right after its own wait: sbfe_norm_sentence_tiles runs beside it (norm.cu).
constexpr int kClampThreads = 128;

constexpr int kClampTilesPerCta = 2;

template <int MC>
__global__ void __launch_bounds__(kClampThreads) clamp_tiles_kernel(const __grid_constant__ Params prm) {
  SBFE_TL(1, 0);
  pdl_wait();
  SBFE_TL(1, 1);
  pdl_launch_dependents();
  const int M = MC > 0 ? MC : prm.n_mels;
  const int tid = threadIdx.x;
  const int tiles = (int)prm.tiles_per_utt;
  const unsigned groups = (prm.tiles_per_utt + kClampTilesPerCta - 1) / kClampTilesPerCta;
  const unsigned b = blockIdx.x / groups, j0 = (blockIdx.x - b * groups) * kClampTilesPerCta;
  const size_t rec = tile_rec_words(M);
  const float* tsb = prm.tile_stats + (size_t)b * tiles * rec;
  __shared__ unsigned s_key;
  if (tid == 0) s_key = 0u;
  unsigned key = 0u;
  for (int i = tid; i < tiles; i += kClampThreads) {
    const unsigned k = __ldg(reinterpret_cast<const unsigned*>(tsb) + (size_t)i * rec + 3 * M);
    key = k > key ? k : key;
  }
  // the CTA's tiles: column minima and flags requested with the keys (one round trip)
  float mn[kClampTilesPerCta];
  unsigned flags[kClampTilesPerCta];
#pragma unroll
  for (int q = 0; q < kClampTilesPerCta; ++q) {
    mn[q] = INFINITY;
    flags[q] = 0u;
    if (j0 + q < prm.tiles_per_utt) {
      const float* ts = tsb + (size_t)(j0 + q) * rec;
      for (int m = tid; m < M; m += kClampThreads) mn[q] = fminf(mn[q], __ldg(ts + 2 * M + m));
      flags[q] = __ldg(reinterpret_cast<const unsigned*>(ts) + 3 * M + 1);
    }
  }
  __syncthreads();
  key = __reduce_max_sync(0xffffffffu, key);
  if ((tid & 31) == 0 && key) atomicMax(&s_key, key);
  __syncthreads();
  if (s_key == 0u) { SBFE_TL(1, 2); return; }
  const float floor_b = max_key_decode(s_key) - prm.top_db;
#pragma unroll
  for (int q = 0; q < kClampTilesPerCta; ++q) {
    if (__syncthreads_or(mn[q] < floor_b) == 0) continue;    // block-uniform (also past the last tile: inf)
    clamp_tile<MC, kClampThreads>(prm, b, j0 + q, floor_b, (flags[q] & 1u) != 0u, tid);
  }
  SBFE_TL(1, 2);
}

constexpr int kMaxSmem = 110 * 1024;  // two CTAs per SM

using KernelFn = void (*)(Params);
// the instantiation for this sample type / filter count (fixed-M builds for the recipes' sizes)
inline KernelFn pick_kernel(int pcm16, int n_mels) {
  switch (n_mels) {
    case 80: return pcm16 ? fbank400_v3_kernel<true, 80> : fbank400_v3_kernel<false, 80>;
    case 40: return pcm16 ? fbank400_v3_kernel<true, 40> : fbank400_v3_kernel<false, 40>;
    case 23: return pcm16 ? fbank400_v3_kernel<true, 23> : fbank400_v3_kernel<false, 23>;
    default: return pcm16 ? fbank400_v3_kernel<true, 0> : fbank400_v3_kernel<false, 0>;
  }
}
inline KernelFn pick_clamp(int n_mels) {
  switch (n_mels) {
    case 80: return clamp_tiles_kernel<80>;
    case 40: return clamp_tiles_kernel<40>;
    case 23: return clamp_tiles_kernel<23>;
    default: return clamp_tiles_kernel<0>;
  }
}

}  // namespace v3

// hop 160, filter tables in the padded layout (every filter's weights start at a multiple
// of 4 and are zero-padded to a multiple of 4), everything fits in half an SM's smem
bool fbank400_v3_supported(const sbfe_fbank_desc* d) {
  return d->n_fft == 400 && d->hop == v3::kHop && d->mel_pad4 == 2 && d->n_mels >= 1 &&
         d->n_mels <= 384 && d->mel_nnz >= 0 &&  // output tile [32][M|1] must fit in region E
        
         v3::smem_floats(d->n_mels, d->mel_nnz) * sizeof(float) <= (size_t)v3::kMaxSmem;
}

static int launch_v3(const sbfe_fbank_desc* d, const void* wav, int pcm16, float scale,
                     const int64_t* offsets, const int32_t* nsamples, int64_t B, int64_t L,
                     int64_t wav_stride, float* out, sbfe_maxkey_t* utt_max, float* tile_stats,
                     float top_db, int power, cudaStream_t stream) {   // power: 1 = |X|^2, 2 = X
  using namespace v3;
  Params prm;
  prm.db = DbParams{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                    power ? 0 : d->log_mel};
  prm.n_mels = power == 2 ? 2 * kBins : (power ? kBins : d->n_mels);
  prm.mel_nnz = power ? 0 : d->mel_nnz;
  prm.L = L;
  prm.wav_stride = wav_stride;
  prm.T = 1 + L / kHop;
  prm.wav = reinterpret_cast<const float*>(wav);
  prm.scale = scale;
  prm.offsets = offsets;
  prm.nsamples = nsamples;
  prm.window = d->d_window;
  prm.out = out;
  prm.utt_max = utt_max;
  prm.tile_stats = tile_stats;
  prm.top_db = top_db >= 0.f ? top_db : INFINITY;
  prm.mel_start = d->d_mel_start;
  prm.mel_count = d->d_mel_count;
  prm.mel_offset = d->d_mel_offset;
  prm.mel_weight = d->d_mel_weight;
  const int64_t tiles = (prm.T + kTile - 1) / kTile;
  if (tiles * B > 0x7fffffff || prm.T > 0x7fffffff) return SBFE_E_TOOBIG;
  prm.tiles_per_utt = (unsigned)tiles;
  prm.n_tiles = (unsigned)(tiles * B);
  const size_t smem = smem_floats(prm.n_mels, prm.mel_nnz) * sizeof(float);
  // once per device and process: the opt-in to > 48 KB of dynamic shared memory and the number
  // of these CTAs an SM holds.  The cache is a lock-free per-device word whose only transition is
  // 0 -> the value after idempotent calls: no call depends on another call's state (sbfe.h).
  static std::atomic<int> per_sm_cache[64];
  int dev = 0, sms = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return (int)e;
  const bool cached = dev >= 0 && dev < 64;
  int per_sm = cached ? per_sm_cache[dev].load(std::memory_order_acquire) : 0;
  if (per_sm == 0) {
    per_sm = 2;
    for (int pcm = 0; pcm < 2; ++pcm)
      for (int mc : {0, 80, 40, 23}) {
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(pick_kernel(pcm, mc), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kMaxSmem);
        int n = 0;
        if (e == cudaSuccess)
          e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, pick_kernel(pcm, mc), kThreads,
                                                            (size_t)kMaxSmem);
        if (e == cudaSuccess && n < per_sm) per_sm = n;
      }
    {
      int n = 0;
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(fbank400_v3_kernel<false, 0, true>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
      if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fbank400_v3_kernel<false, 0, true>, kThreads,
                                                          (size_t)kMaxSmem);
      if (e == cudaSuccess && n < per_sm) per_sm = n;
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(fbank400_v3_kernel<false, 0, false, true>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
      if (e == cudaSuccess)
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fbank400_v3_kernel<false, 0, false, true>,
                                                          kThreads, (size_t)kMaxSmem);
      if (e == cudaSuccess && n < per_sm) per_sm = n;
    }
    if (e != cudaSuccess) return (int)e;
    if (per_sm < 1) return (int)cudaErrorLaunchOutOfResources;
    if (cached) per_sm_cache[dev].store(per_sm, std::memory_order_release);
  }
  // persistent: every CTA strides over the tiles
  const unsigned cap = (unsigned)(per_sm * sms);
  const unsigned grid = prm.n_tiles < cap ? prm.n_tiles : cap;
  KernelFn kern = power == 2 ? fbank400_v3_kernel<false, 0, false, true>
                             : (power ? fbank400_v3_kernel<false, 0, true> : pick_kernel(pcm16, d->n_mels));
  int rc = launch_pdl(kern, dim3(grid), dim3(kThreads), (size_t)smem, stream, prm);
  if (rc == 0 && !power && tile_stats != nullptr && prm.top_db < INFINITY) {
    rc = launch_pdl(pick_clamp(d->n_mels), dim3((unsigned)(((prm.tiles_per_utt + kClampTilesPerCta - 1) / kClampTilesPerCta) * (prm.n_tiles / prm.tiles_per_utt))), dim3(kClampThreads), 0, stream, prm);
  }
  return rc;
}

int launch_fbank400_v3(const sbfe_fbank_desc* d, const void* wav, int pcm16, float scale,
                       const int64_t* offsets, const int32_t* nsamples, int64_t B, int64_t L,
                       int64_t wav_stride, float* out, sbfe_maxkey_t* utt_max, float* tile_stats,
                       float top_db, cudaStream_t stream) {
  return launch_v3(d, wav, pcm16, scale, offsets, nsamples, B, L, wav_stride, out, utt_max, tile_stats,
                   top_db, 0, stream);
}

// STFT + power spectrum of the default mode (centre, zero padding, one-sided) at n_fft 400 / hop
// 160: the fused kernel's transform with the [32 frames][201 bins] tile as its output
bool stft400_power_supported(const sbfe_fbank_desc* d) {
  return d->n_fft == 400 && d->hop == v3::kHop &&
         v3::smem_floats(2 * v3::kBins, 0) * sizeof(float) <= (size_t)v3::kMaxSmem;
}
// complex != 0: out [B, T, 201, 2]; else the power spectrum [B, T, 201]
int launch_stft400_power(const sbfe_fbank_desc* d, const float* wav, int64_t B, int64_t L,
                         int64_t wav_stride, float* out, int complex, cudaStream_t stream) {
  return launch_v3(d, wav, 0, 1.f, nullptr, nullptr, B, L, wav_stride, out, nullptr, nullptr, -1.f,
                   complex ? 2 : 1, stream);
}

// the clamp launch of mode (b) for any kernel that emits the same 32-frame tile records
int launch_clamp_tiles(const sbfe_fbank_desc* d, int64_t B, int64_t T, float* out, const float* tile_stats,
                       float top_db, cudaStream_t stream) {
  using namespace v3;
  Params prm = {};
  prm.db = DbParams{d->multiplier * 0.30102999566398120f, d->amin, d->db_offset, d->floor_db,
                    d->log_mel};
  prm.n_mels = d->n_mels;
  prm.T = T;
  prm.out = out;
  prm.tile_stats = const_cast<float*>(tile_stats);
  prm.top_db = top_db;
  const int64_t tiles = (T + kTile - 1) / kTile;
  if (tiles * B > 0x7fffffff) return SBFE_E_TOOBIG;
  prm.tiles_per_utt = (unsigned)tiles;
  prm.n_tiles = (unsigned)(tiles * B);
  return launch_pdl(pick_clamp(d->n_mels), dim3((unsigned)(((prm.tiles_per_utt + kClampTilesPerCta - 1) / kClampTilesPerCta) * (prm.n_tiles / prm.tiles_per_utt))), dim3(kClampThreads), 0, stream, prm);
}

#ifdef SBFE_EXP_TIMELINE
SBFE_TL_DEFINE_SYNC(tl_sync_v3)
#endif

}  // namespace sbfe
