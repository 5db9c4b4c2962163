/* sbfe.h - C ABI of the B200-native SpeechBrain feature frontend ("sbfe").
 *
 * One shared library (speechbrain_b200/csrc/libsbfe.so) of hand-written sm_100a
 * kernels.  The reference (SpeechBrain, pure Python on torch) has no FFI; the
 * drop-in boundary on its side is the Python classes named below, and each entry
 * point here is what the replacement class binds (ctypes stubs: INTEGRATION.md).
 *
 * Conventions
 *  - every pointer named d_* is a DEVICE pointer owned by the caller; h_* is host.
 *  - all tensors are contiguous float32 unless a stride argument is given.
 *  - `stream` is a cudaStream_t passed as void*; every call only enqueues work:
 *    no allocation, no synchronisation; thread-safe: no call reads state another call
 *    wrote (the only process-wide data are immutable tables and a lock-free once-per-device
 *    mark that an idempotent cudaFuncSetAttribute has been made).
 *  - return value: 0 ok; <0 invalid argument (SBFE_E_*); >0 a cudaError_t.
 *  - frame count T = 1 + L / hop (zero centre padding of n_fft/2 per side).
 */
#ifndef SBFE_H_
#define SBFE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SBFE_ABI_VERSION 14

#if defined(__GNUC__)
#define SBFE_API __attribute__((visibility("default")))
#else
#define SBFE_API
#endif

#define SBFE_E_INVAL (-1)    /* bad size / null pointer / unsupported combination */
#define SBFE_E_TOOBIG (-2)   /* a dimension exceeds what the kernels index */
#define SBFE_E_NOSPACE (-3)  /* workspace too small */
#define SBFE_E_STATIC_LIMIT (-4) /* sbfe_norm_stats_static: needs more than SBFE_NORM_STATIC_WS_BYTES */

/* Utterance-maximum keys: order-preserving uint32 encoding of a float so that
 * atomicMax works for negative dB values.  0 == "no value yet". */
typedef uint32_t sbfe_maxkey_t;

/* STFT + mel filterbank configuration.
 * Mirrors STFT.__init__ (speechbrain/processing/features.py:101-131) and
 * Filterbank.__init__ (:415-488) after the host has turned them into tables. */
typedef struct sbfe_fbank_desc {
  int32_t n_fft;        /* DFT size; bins = n_fft/2 + 1 */
  int32_t hop;          /* hop in samples */
  int32_t n_mels;       /* number of filters M */
  int32_t log_mel;      /* 1: amplitude_to_DB (without the top_db clamp), 0: linear */
  float multiplier;     /* 10 for power spectrogram, else 20           (:451-454) */
  float amin;           /* clamp before the log                        (:701) */
  float db_offset;      /* multiplier * log10(max(amin, ref_value))    (:446,702) */
  float floor_db;       /* value emitted when x <= amin (= multiplier*log10(amin) - db_offset,
                           rounded on the host like torch does; -100 by default) */
  const float* h_window;     /* HOST [n_fft]: window, zero-padded/centred to n_fft */
  const float* d_window;     /* DEVICE copy of the same */
  /* filterbank matrix [bins, M] in compressed-column form: filter m covers bins
   * [mel_start[m], mel_start[m]+mel_count[m]) with weights
   * d_mel_weight[mel_offset[m] ...]  (triangular: <= ~2*band/df entries;
   * gaussian: all bins). */
  const int32_t* d_mel_start;   /* [M] */
  const int32_t* d_mel_count;   /* [M] */
  const int32_t* d_mel_offset;  /* [M] */
  const float* d_mel_weight;    /* [nnz] */
  int32_t mel_nnz;              /* total number of stored weights (padding included) */
  int32_t mel_pad4;             /* 0: plain CSR (always valid, slower kernel).
                                   1: every mel_offset[m] is a multiple of 4 and each filter's
                                   weights are followed by zeros up to a multiple of 4 entries
                                   (mel_count stays the unpadded count).
                                   2: as 1, and filters (2k, 2k+1) both store 4*G weights with
                                   G = max over the pair of ceil(mel_count / 4), zero filled, with
                                   mel_start[m] + 4*G <= 4*ceil(bins / 4) for both.  Needed by the
                                   fused n_fft = 400 / hop = 160 kernel (128-bit weight loads,
                                   warp-uniform trip counts). */
  int32_t mel_precision;        /* dense filterbanks only (gaussian: at least half of the
                                   [bins, M] matrix is stored): they run as a tensor-core GEMM
                                   (mma.sync TF32, fp32 accumulation).  0: 3xTF32 split, inside
                                   the fp32 tolerance of the path (max-abs 1e-4 dB).  1: one TF32
                                   pass - inputs rounded to 10-bit mantissas: |err| <= 2e-2 dB
                                   stated bound (typically 3e-3).  Sparse filterbanks ignore it. */
} sbfe_fbank_desc;

SBFE_API int sbfe_abi_version(void);
SBFE_API const char* sbfe_error_string(int code);

/* Number of frames torch.stft(center=True) yields for L samples (features.py:161). */
SBFE_API int64_t sbfe_num_frames(int64_t n_samples, int32_t hop);

/* Scratch bytes sbfe_fbank_forward needs for shapes [B, L] (0 on the fused
 * n_fft == 400 path; power spectrogram [B,T,bins] otherwise). */
SBFE_API int64_t sbfe_fbank_workspace_bytes(const sbfe_fbank_desc* desc, int64_t B, int64_t L);

/* Fbank hot path = STFT.forward + spectral_magnitude(power=1) + Filterbank.forward
 * (lobes/features.py:138-140; processing/features.py:133-188, 327-356, 490-551, 691-712).
 * wav [B, L] (row stride wav_stride) -> out [B, T, M].
 *
 * Two modes:
 *  (a) d_tile_stats == NULL: out holds the dB values WITHOUT the
 *      per-utterance top_db clamp; d_utt_max[b] (zeroed by the caller first; may be NULL) receives
 *      the key of max_{t,m} out[b] for a later clamp (sbfe_post_forward).  top_db is ignored.
 *  (b) tile records, only where sbfe_fbank_tile_frames(desc) = R > 0 (the fused n_fft 400 and
 *      n_fft 512 kernels, hop 160): d_tile_stats = [B][ceil(T/R)] records of W = sbfe_fbank_tile_stats_floats(desc)
 *      = 3*M + 4 words.  When the call's work completes, out IS top_db-clamped
 *      (x = max(x, max_b - top_db), features.py:706-710, the maximum taken over all T frames of
 *      the padded utterance; top_db < 0: no clamp), and record j of utterance b holds
 *        [0, M)    mean  } of every feature column over the existing frames of tile j, BEFORE
 *        [M, 2M)   M2    } the clamp (M2 = sum of squared deviations),
 *        [2M, 3M)  the column minima of tile j before the clamp: a column whose minimum is not
 *                  below the utterance's floor was not changed by the clamp, its moments hold,
 *        [3M]      max key of tile j, [3M+1] flags (bit 0: the tile's input was all zero),
 *        [3M+2], [3M+3] zero (small counters of the consumers; never above any key).
 *      sbfe_norm_stats_tiles / sbfe_norm_sentence_tiles turn these into InputNormalization
 *      statistics for any valid length: 2 M floats per R frames instead of a pass over out.
 *      The clamp is a second, small kernel launched by this call (it reads the records only
 *      and touches just the tiles that hold a value below the floor).  d_utt_max may be NULL. */
SBFE_API int sbfe_fbank_forward(const sbfe_fbank_desc* desc, const float* d_wav, int64_t B, int64_t L,
                       int64_t wav_stride, float* d_out, sbfe_maxkey_t* d_utt_max,
                       float* d_tile_stats, float top_db, void* d_workspace,
                       int64_t workspace_bytes, void* stream);

/* Ragged ingest (the step before the path: speechbrain/dataio/batch.py:20-190 pads on the CPU
 * and copies the padding over PCIe).  The batch arrives as ONE packed buffer holding only the
 * valid samples: utterance b = d_packed[d_offsets[b] .. + d_nsamples[b]) (offsets multiples of
 * 4 samples for 16-byte aligned rows), d_nsamples[b] <= L.  Result = sbfe_fbank_forward on the
 * zero-padded [B, L] batch, bit for bit; the padding is never materialised (the copy engine of
 * the fused kernel zero-fills).  Only where sbfe_fbank_ragged_supported(desc) != 0, else
 * SBFE_E_INVAL. */
SBFE_API int sbfe_fbank_forward_ragged(const sbfe_fbank_desc* desc, const float* d_packed,
                              const int64_t* d_offsets, const int32_t* d_nsamples, int64_t B,
                              int64_t L, float* d_out, sbfe_maxkey_t* d_utt_max,
                              float* d_tile_stats, float top_db, void* stream);

/* The same with 16-bit PCM samples (what a .wav / .flac file holds; the reference converts them
 * to float32 on the CPU, dataio/dataio.py read_audio via torchaudio.load, and ships 4 bytes per
 * sample over PCIe): sample value = scale * d_packed[i], scale > 0 (1/32768 for torchaudio's
 * normalisation).  int16 -> float32 is exact and the scale is folded into the window taps, so
 * for a power-of-two scale the result equals sbfe_fbank_forward_ragged on the converted
 * float32 samples bit for bit.  Offsets are multiples of 4 samples. */
SBFE_API int sbfe_fbank_forward_ragged_pcm16(const sbfe_fbank_desc* desc, const int16_t* d_packed,
                                    const int64_t* d_offsets, const int32_t* d_nsamples,
                                    int64_t B, int64_t L, float scale, float* d_out,
                                    sbfe_maxkey_t* d_utt_max, float* d_tile_stats, float top_db,
                                    void* stream);

SBFE_API int32_t sbfe_fbank_tile_frames(const sbfe_fbank_desc* desc);
/* 1 where sbfe_fbank_forward_ragged[_pcm16] take this configuration (fused n_fft 400 / hop 160). */
SBFE_API int32_t sbfe_fbank_ragged_supported(const sbfe_fbank_desc* desc);
SBFE_API int32_t sbfe_fbank_tile_stats_floats(const sbfe_fbank_desc* desc);

/* InputNormalization statistics (features.py:1020-1031, 1120-1145; same contract as
 * sbfe_norm_stats) from the tile records of a mode-(b) sbfe_fbank_forward whose output d_x has not
 * been modified since: mean / std [B, F] over the first rint(lengths[b] * T) frames = the records
 * of the whole tiles below that length merged with the rows of the tile it cuts (columns the
 * clamp changed are recomputed from d_x).  top_db: what that sbfe_fbank_forward call was given. */
SBFE_API int sbfe_norm_stats_tiles(const float* d_x, const float* d_lengths, int64_t B, int64_t T,
                          int32_t F, int32_t mean_norm, int32_t std_norm, float eps,
                          const float* d_tile_stats, int32_t tile_frames, float top_db,
                          float* d_mean, float* d_std, void* stream);
/* The same followed by out = (x - mean_b) / std_b over all T frames ("sentence" mode,
 * features.py:1020-1035; in place when d_out == d_x), in ONE launch: every CTA rebuilds the
 * statistics of its utterance from the records while its share of x is in flight.
 * top_db: the value the producing sbfe_fbank_forward call was given (< 0: none).
 * flags: SBFE_BESIDE_CLAMP - the caller states that nothing has been launched into `stream`
 *   since that sbfe_fbank_forward call that writes d_x, the records or d_lengths (launches that
 *   do not touch them are fine).  The kernel then does not wait for the call's clamp launch
 *   before it reads: it applies the clamp to every value it reads itself (same result before
 *   and after), rebuilds the floor from the records' keys, recomputes what a record below the
 *   floor would have told it, and waits only before its first store.  Without the flag it waits
 *   first, like every other kernel of the library.  Results are identical either way.
 *   SBFE_OUT_BF16 - d_out is a bfloat16 [B, T, F] tensor (must not be d_x): the opt-in
 *   reduced-precision output, rounded to nearest even from the float32 result, i.e. relative
 *   error <= 2^-8 of each normalised value (stated bound; the statistics stay float32).
 * In place, the rows the statistics are read from (the tile the length cuts, tiles with values
 * below the floor) are written by the CTA of the utterance that finishes its statistics last: the
 * CTAs count themselves in word 3F + 3 of the utterance's first record (zero from the producer,
 * zero again when the call completes) - the one reason d_tile_stats is not const; two in-place
 * calls must not share records concurrently.  Out of place needs none of this and is the faster
 * call (about 6 us per 32 x 16 s batch). */
#define SBFE_BESIDE_CLAMP 1
#define SBFE_OUT_BF16 2
SBFE_API int sbfe_norm_sentence_tiles(const float* d_x, const float* d_lengths, int64_t B, int64_t T,
                             int32_t F, int32_t mean_norm, int32_t std_norm, float eps,
                             float* d_tile_stats, int32_t tile_frames, float top_db,
                             int32_t flags, void* d_out, void* stream);

/* STFT.forward for [B, L] input (processing/features.py:133-188):
 * out [B, T, bins, 2] (re, im), or, with power_only != 0, [B, T, bins] = re^2+im^2.
 * flags select torch.stft's other modes (STFT.__init__ :101-112):
 *   SBFE_STFT_NO_CENTER  center=False: frame t starts at t * hop, T = 1 + (L - n_fft) / hop
 *                        (0 when L < n_fft); default: frame t is centred on t * hop, T = 1 + L / hop
 *   SBFE_STFT_REFLECT    pad_mode="reflect" instead of zeros (centred only; needs L > n_fft / 2)
 *   SBFE_STFT_TWOSIDED   onesided=False: bins = n_fft (X[n_fft - k] = conj X[k]) instead of
 *                        n_fft / 2 + 1
 * sbfe_stft_frames gives T for the same flags. */
#define SBFE_STFT_NO_CENTER 1
#define SBFE_STFT_REFLECT 2
#define SBFE_STFT_TWOSIDED 4
SBFE_API int64_t sbfe_stft_frames(const sbfe_fbank_desc* desc, int64_t L, int32_t flags);
SBFE_API int sbfe_stft_forward(const sbfe_fbank_desc* desc, const float* d_wav, int64_t B, int64_t L,
                      int64_t wav_stride, float* d_out, int32_t power_only, int32_t flags,
                      void* stream);

/* ---- next row: ISTFT.forward (processing/features.py:283-324) = torch.istft
 * for one-sided spectra: d_spec [B, T, n_fft/2 + 1, 2] -> d_out [B, out_len].  Per frame the
 * inverse real DFT (1/n_fft; the imaginary parts of bins 0 and n_fft/2 are ignored), times
 * d_window [n_fft] (the caller centres a shorter window inside n_fft), overlap-add at t * hop,
 * division by the overlap-added squared window, then samples [start, start + out_len) with
 * start = n_fft/2 if center else 0; samples past n_fft + hop * (T - 1) are zero.  The caller
 * checks that the window envelope is non-zero over that range (torch.istft raises otherwise).
 * Workspace: sbfe_istft_workspace_bytes (the windowed frames). */
SBFE_API int64_t sbfe_istft_workspace_bytes(int64_t B, int64_t T, int32_t n_fft);
SBFE_API int sbfe_istft_forward(const float* d_spec, int64_t B, int64_t T, int32_t n_fft, int32_t hop,
                       const float* d_window, int32_t center, int64_t out_len, float* d_out,
                       void* d_workspace, int64_t workspace_bytes, void* stream);

/* spectral_magnitude (processing/features.py:327-356) on n complex pairs. */
SBFE_API int sbfe_spectral_magnitude(const float* d_stft, int64_t n, float power, int32_t take_log,
                            float eps, float* d_out, void* stream);

/* Filterbank.forward on a spectrogram [B, T, bins] (processing/features.py:490-551,
 * 691-702), again without the top_db clamp; d_utt_max may be NULL. */
SBFE_API int sbfe_filterbank_forward(const sbfe_fbank_desc* desc, const float* d_spec, int64_t B,
                            int64_t T, float* d_out, sbfe_maxkey_t* d_utt_max, void* stream);

/* Post stage, any subset of (in this order):
 *   top_db clamp  x = max(x, max_b - top_db)          (features.py:706-710)
 *   DCT           x @ d_dct[F_in, n_dct]              (features.py:764-786)
 *   Deltas        deltas=1: [x, d(x), d(d(x))]; deltas=2: d(x) only   (features.py:820-852; lobes:141-144)
 *   ContextWindow out[t, f*C + c] = x[t + c - left, f] (features.py:897-937)
 * x [B, T, F_in] -> out [B, T, F_out], F_out = (n_dct or F_in) * (3 if deltas) * (left+right+1 if context).
 * d_utt_max NULL => no clamp; d_dct NULL => no DCT; out may alias x only when
 * no DCT/deltas/context is requested (pure clamp). */
SBFE_API int sbfe_post_forward(const float* d_x, int64_t B, int64_t T, int32_t F_in,
                      const sbfe_maxkey_t* d_utt_max, float top_db, const float* d_dct,
                      int32_t n_dct, int32_t deltas, int32_t delta_window, int32_t context,
                      int32_t left, int32_t right, float* d_out, void* stream);

/* InputNormalization statistics (processing/features.py:1020-1031, 1120-1145).
 * x [B, T, F], d_lengths [B] relative lengths; n_b = rint(lengths[b] * T) frames.
 * Writes mean[B, F] and std[B, F] (unbiased, floored at eps; mean_norm == 0 -> 0,
 * std_norm == 0 -> 1).  With d_utt_max != NULL the top_db clamp is applied to x
 * on the fly (fused frontend).  Workspace: sbfe_norm_workspace_bytes(B, T, F). */
SBFE_API int64_t sbfe_norm_workspace_bytes(int64_t B, int64_t T, int32_t F);
SBFE_API int sbfe_norm_stats(const float* d_x, const float* d_lengths, int64_t B, int64_t T, int32_t F,
                    int32_t mean_norm, int32_t std_norm, float eps,
                    const sbfe_maxkey_t* d_utt_max, float top_db, float* d_mean, float* d_std,
                    void* d_workspace, int64_t workspace_bytes, void* stream);

/* sbfe_norm_stats without a workspace argument and without any allocation.  No counterpart in
 * the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernels, same results as sbfe_norm_stats; the partial sums go to a device array built into
 * the library (SBFE_NORM_STATIC_WS_BYTES, held by every GPU that loads the library until the
 * process exits).  Limit: sbfe_norm_workspace_bytes(B, T, F) = 16 * B * ceil(T / 256) * F must
 * not exceed SBFE_NORM_STATIC_WS_BYTES (e.g. 32 x 16 s x 80 mel = 280 KiB, 128 x 30 s x 80 mel =
 * 1.9 MiB), else SBFE_E_STATIC_LIMIT and nothing is launched.  The call makes no cudaMalloc /
 * cudaFree / allocating library call (first use on a device resolves the array's address with
 * cudaGetSymbolAddress).  All calls share the one array: calls whose kernels may overlap on the
 * GPU (different streams, no ordering between them) race; calls on one stream are safe.  Nothing
 * in the array has to be cleared between calls (every slot read was written by the same call). */
#define SBFE_NORM_STATIC_WS_BYTES (2 * 1024 * 1024)
SBFE_API int sbfe_norm_stats_static(const float* d_x, const float* d_lengths, int64_t B, int64_t T,
                           int32_t F, int32_t mean_norm, int32_t std_norm, float eps,
                           const sbfe_maxkey_t* d_utt_max, float top_db, float* d_mean,
                           float* d_std, void* stream);

/* batch / global statistics (processing/features.py:1083-1114):
 * d_cur[2F+1] = {sum_b mean_b, sum_b std_b, B} (the all-reduce payload).  */
SBFE_API int sbfe_norm_reduce(const float* d_mean, const float* d_std, int64_t B, int32_t F,
                     float* d_cur, void* stream);
/* running update: glob = first ? cur/n : (1-w)*glob + w*cur/n, n = d_cur[2F]. */
SBFE_API int sbfe_norm_update(const float* d_cur, int32_t F, int32_t first, float weight,
                     float* d_glob_mean, float* d_glob_std, void* stream);

/* out = (x - mean) / std (processing/features.py:1035, 1088, 1116).
 * per_utt != 0: mean/std are [B, F]; else [F] shared.  In place when out == x.
 * d_utt_max != NULL applies the top_db clamp to x first.  flags: 0 or SBFE_OUT_BF16 (d_out is
 * bfloat16 [B, T, F], not d_x). */
SBFE_API int sbfe_norm_apply(const float* d_x, int64_t B, int64_t T, int32_t F, const float* d_mean,
                    const float* d_std, int32_t per_utt, const sbfe_maxkey_t* d_utt_max,
                    float top_db, int32_t flags, void* d_out, void* stream);

/* ---- next row: speaker-mode statistics (processing/features.py:1037-1081) --
 * The speaker dictionary stays on the host (ids -> rows of d_table_mean / d_table_std [rows, F]);
 * this call replays the batch in order: for sentence i with current statistics
 * d_mean[i], d_std[i] ([B, F], from sbfe_norm_stats*):
 *   d_row[i] < 0              no table entry is used: out[i] = the sentence's own statistics;
 *   d_mode[i] = 2             first sight of the speaker: table[row] = current;
 *   d_mode[i] = 1             table[row] = (1 - w) * table[row] + w * current, with
 *                             w = d_weight[i] and (1 - w) = d_weight[B + i] (both rounded by the
 *                             caller; evaluated as two products and a sum, like the reference);
 *   d_mode[i] = 0             read only;
 * and out[i] = table[row] after the update - what sbfe_norm_apply(per_utt = 1) then uses. */
SBFE_API int sbfe_norm_speaker_update(const float* d_mean, const float* d_std, int64_t B, int32_t F,
                             const int32_t* d_row, const int32_t* d_mode, const float* d_weight,
                             float* d_table_mean, float* d_table_std, float* d_out_mean,
                             float* d_out_std, void* stream);

/* ---- next row: SpecAugment on features ----------------------------------
 * speechbrain/lobes/augment.py:32-201.  x [B, T, F] -> out [B, T, F]:
 *   time warp (:116-150)  rows [0, c) resampled onto [0, w) and rows [c, T) onto [w, T) along
 *                         time with torch.nn.functional.interpolate(align_corners=True)
 *                         semantics; warp_mode 0 = bicubic (A = -0.75, clamped taps),
 *                         1 = bilinear; warp_c < 0: no warp;
 *   masks (:152-201)      element (t, f) of utterance b becomes the fill value when
 *                         fpos[b,i] <= f < fpos[b,i] + flen[b,i] for some i < n_fmask or
 *                         tpos[b,i] <= t < tpos[b,i] + tlen[b,i] for some i < n_tmask.
 * The draws c, w, *pos, *len are the caller's (the reference's torch.randint calls, int64).
 * Fill = *d_fill when d_fill != NULL (device scalar, e.g. from sbfe_mean), else `fill`.
 * With a warp d_out must differ from d_x; without one d_out must equal d_x and only the
 * masked elements are written (nothing is read).  n_*mask <= 16. */
SBFE_API int sbfe_specaugment_forward(const float* d_x, int64_t B, int64_t T, int32_t F,
                             int32_t warp_c, int32_t warp_w, int32_t warp_mode,
                             const int64_t* d_fpos, const int64_t* d_flen, int32_t n_fmask,
                             const int64_t* d_tpos, const int64_t* d_tlen, int32_t n_tmask,
                             const float* d_fill, float fill, float* d_out, void* stream);

/* ---- next row (second half): time-domain augmentation ------------------------
 * speechbrain/processing/speech_augmentation.py; lobes/augment.py:204-299 (TimeDomainSpecAugment)
 * chains SpeedPerturb -> DropFreq -> DropChunk right before compute_features.  The random draws
 * are the caller's (the reference's torch calls); these are the data movements. */

/* DropChunk (:1013-1173).  d_out [B, L, channels] already holds a copy of the input; chunk i =
 * d_chunks[4 i ..] = {utterance, start, length (time steps), offset into d_noise (time steps)}
 * is zeroed, or, with d_noise / d_noise_max [B] given, filled with
 * 2 * noise_max[b] * d_noise[...] - noise_max[b] (:1163-1170; d_noise: the uniform draws).
 * Chunks of one call must belong to different utterances or not overlap (the caller issues one
 * call per round of overlapping chunks, in the reference's order). */
SBFE_API int sbfe_wav_drop_chunks(float* d_out, int64_t B, int64_t L, int32_t channels,
                         const int32_t* d_chunks, int32_t n_chunks, int64_t max_len,
                         const float* d_noise, const float* d_noise_max, void* stream);
/* DropFreq's filtering (:994-1007) = convolve1d(x, taps, padding = pad) of
 * signal_processing.py:157-277: out[b][t] = sum_k taps[k] x[b][t + k - pad], x zero outside
 * [0, L) (conv1d, no kernel flip); x, out [B, L]; n_taps <= 1024. */
SBFE_API int sbfe_wav_fir(const float* d_x, int64_t B, int64_t L, const float* d_taps, int32_t n_taps,
                 int32_t pad, float* d_out, void* stream);
/* DropFreq's composite filter (:977-1001, notch_filter of signal_processing.py:373-430) built on
 * the device from the drawn frequencies: h_freqs is a HOST array of n_freqs <= 16 normalised
 * frequencies (copied into the launch's arguments), filter_len odd <= 255; d_taps [filter_len] is
 * what sbfe_wav_fir takes (pad = filter_len / 2). */
SBFE_API int sbfe_wav_notch_taps(const float* h_freqs, int32_t n_freqs, float width, int32_t filter_len,
                        float* d_taps, void* stream);
/* DoClip (:1176-1228): out = clamp(x, -value, value) over n floats; in place when d_out == d_x. */
SBFE_API int sbfe_wav_clip(const float* d_x, int64_t n, float value, float* d_out, void* stream);
/* AddNoise with white noise (csv_file = None; :126-168): the caller draws snr_db [B] and the noise
 * [B, L] with the reference's torch.rand / torch.randn_like calls.
 *   sbfe_wav_abs_sums   d_partials [B][sbfe_wav_abs_chunks(L)]: sums of |x| per 16384 samples (the
 *                       numerator of compute_amplitude, signal_processing.py:51-56, which sums the
 *                       whole padded row); the consumer adds them in index order
 *   sbfe_wav_add_noise  out = x * (1 - f) + (f * amp) * noise, f = 1 / (10^(snr_db / 20) + 1),
 *                       amp = sum|x| / (d_lengths[b] * L), rounded like the reference's three eager
 *                       ops; in place when d_out == d_x; d_absmax (NULL, or [B] zeroed by the
 *                       caller) receives the bits of max |out[b]| for normalize = True
 *   sbfe_wav_peak_normalize  x[b] /= max(absmax[b], 1)  (:164-168) */
SBFE_API int64_t sbfe_wav_abs_chunks(int64_t L);
SBFE_API int sbfe_wav_abs_sums(const float* d_x, int64_t B, int64_t L, float* d_partials, void* stream);
SBFE_API int sbfe_wav_add_noise(const float* d_x, const float* d_noise, int64_t B, int64_t L,
                       const float* d_partials, const float* d_lengths, const float* d_snr_db,
                       float* d_out, uint32_t* d_absmax, void* stream);
SBFE_API int sbfe_wav_peak_normalize(float* d_x, int64_t B, int64_t L, const uint32_t* d_absmax,
                            void* stream);
/* AddBabble.forward (:861-906): babble[b] = x[b-1] + ... + x[b-speakers] (rows modulo B, added in
 * that order), out = x * (1 - f) + babble * (f * amp / (bamp + 1e-14)) with amp = sum|x[b]| /
 * (d_lengths[b] * L), bamp = sum|babble[b]| / max_s(d_lengths[b-s] * L), f as in
 * sbfe_wav_add_noise.  d_partials: scratch of 2 * B * sbfe_wav_abs_chunks(L) floats; d_out must not
 * be d_x (every row is read by `speakers` other rows). */
SBFE_API int sbfe_wav_add_babble(const float* d_x, int64_t B, int64_t L, int32_t speakers, float* d_partials,
                        const float* d_lengths, const float* d_snr_db, float* d_out, void* stream);
/* Resample._perform_resample (:600-665), Kaldi's LinearResample: output sample n * phases + i =
 * sum_k d_weights[i][k] x[d_first[i] + n * in_stride + k] (x zero outside [0, L)), k < width;
 * x [B, L] -> out [B, L_out].  Weights and first indices: Resample._indices_and_weights (:740-820). */
SBFE_API int sbfe_wav_resample(const float* d_x, int64_t B, int64_t L, const float* d_weights,
                      const int32_t* d_first, int32_t phases, int32_t in_stride, int32_t width,
                      int64_t L_out, float* d_out, void* stream);

/* mean over n floats as a device scalar (x.mean() of augment.py:196); two small launches,
 * double accumulation, workspace of sbfe_mean_workspace_bytes() bytes. */
SBFE_API int64_t sbfe_mean_workspace_bytes(void);
SBFE_API int sbfe_mean(const float* d_x, int64_t n, float* d_mean, void* d_workspace,
              int64_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SBFE_H_ */
