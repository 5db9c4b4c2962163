// extern "C" entry points of libsbfe.so (declared in include/sbfe.h).
// Validation + dispatch only; kernels live in fbank400.cu / generic.cu /
// post.cu / norm.cu.  Nothing here allocates, synchronises or keeps state.
#include "common.cuh"

namespace sbfe {
int launch_fbank400(const sbfe_fbank_desc*, const float*, int64_t, int64_t, int64_t, float*,
                    sbfe_maxkey_t*, cudaStream_t);
int launch_fbank400_v3(const sbfe_fbank_desc*, const void*, int, float, const int64_t*,
                       const int32_t*, int64_t, int64_t, int64_t, float*, sbfe_maxkey_t*, float*,
                       float, cudaStream_t);
int launch_norm_stats_tiles(const float*, const float*, int64_t, int64_t, int, int, int, float,
                            const float*, int, float, float*, float*, cudaStream_t);
int launch_norm_sentence_tiles(const float*, const float*, int64_t, int64_t, int, int, int, float,
                               float*, int, float, int, int, void*, cudaStream_t);
int64_t istft_workspace_bytes(int64_t, int64_t, int);
int launch_istft(const float*, int64_t, int64_t, int, int, const float*, int, int64_t, float*, void*,
                 int64_t, cudaStream_t);
bool fbank512_supported(const sbfe_fbank_desc*);
int launch_fbank512(const sbfe_fbank_desc*, const float*, int64_t, int64_t, int64_t, float*,
                    sbfe_maxkey_t*, float*, float, cudaStream_t);
bool fbank_pow2_supported(const sbfe_fbank_desc*);
bool mel_dense_mma(const sbfe_fbank_desc*);
int launch_fbank_pow2(const sbfe_fbank_desc*, const float*, int64_t, int64_t, int64_t, float*,
                      sbfe_maxkey_t*, cudaStream_t);
int64_t stft_frames(const sbfe_fbank_desc*, int64_t, int);
int launch_stft_generic(const sbfe_fbank_desc*, const float*, int64_t, int64_t, int64_t, float*,
                        int, int, cudaStream_t);
int launch_spectral_magnitude(const float*, int64_t, float, int, float, float*, cudaStream_t);
int launch_mel_db(const sbfe_fbank_desc*, const float*, int64_t, int64_t, float*, sbfe_maxkey_t*,
                  cudaStream_t);
int launch_post(const float*, int64_t, int64_t, int, const sbfe_maxkey_t*, float, const float*,
                int, int, int, int, int, int, float*, cudaStream_t);
int64_t norm_workspace_bytes(int64_t, int64_t, int);
int launch_norm_stats(const float*, const float*, int64_t, int64_t, int, int, int, float,
                      const sbfe_maxkey_t*, float, float*, float*, void*, int64_t, cudaStream_t);
int launch_norm_stats_static(const float*, const float*, int64_t, int64_t, int, int, int, float,
                             const sbfe_maxkey_t*, float, float*, float*, cudaStream_t);
int launch_norm_reduce(const float*, const float*, int64_t, int, float*, cudaStream_t);
int launch_norm_update(const float*, int, int, float, float*, float*, cudaStream_t);
int launch_norm_apply(const float*, int64_t, int64_t, int, const float*, const float*, int,
                      const sbfe_maxkey_t*, float, int, void*, cudaStream_t);

static bool desc_ok(const sbfe_fbank_desc* d, bool need_mel) {
  if (!d || d->n_fft < 2 || (d->n_fft & 1) || d->hop < 1 || !d->d_window || !d->h_window)
    return false;
  if (need_mel && (d->n_mels < 1 || !d->d_mel_start || !d->d_mel_count || !d->d_mel_offset ||
                   !d->d_mel_weight))
    return false;
  return true;
}
// the fused kernel covers n_fft == 400, hop <= 400 and filterbanks whose tables fit in smem
int launch_specaugment(const float*, int64_t, int64_t, int, int, int, int, const int64_t*,
                       const int64_t*, int, const int64_t*, const int64_t*, int, const float*, float,
                       float*, cudaStream_t);
int launch_drop_chunks(float*, int64_t, int, const int32_t*, int, int64_t, const float*, const float*,
                       cudaStream_t);
int launch_fir_same(const float*, int64_t, int64_t, const float*, int, int, float*, cudaStream_t);
int launch_clip(const float*, int64_t, float, float*, cudaStream_t);
int launch_notch_taps(const float*, int, float, int, float*, cudaStream_t);
int64_t abs_chunks(int64_t);
int launch_abs_sums(const float*, int64_t, int64_t, float*, cudaStream_t);
int launch_add_noise(const float*, const float*, int64_t, int64_t, const float*, const float*, const float*,
                     float*, unsigned*, cudaStream_t);
int launch_peak_normalize(float*, int64_t, int64_t, const unsigned*, cudaStream_t);
int launch_add_babble(const float*, int64_t, int64_t, int, float*, float*, const float*, const float*, float*,
                      cudaStream_t);
int launch_resample(const float*, int64_t, int64_t, const float*, const int32_t*, int, int, int,
                    int64_t, float*, cudaStream_t);
int64_t mean_workspace_bytes();
int launch_mean(const float*, int64_t, float*, void*, int64_t, cudaStream_t);
int launch_norm_speaker_update(const float*, const float*, int64_t, int, const int32_t*,
                               const int32_t*, const float*, float*, float*, float*, float*,
                               cudaStream_t);
bool fbank400_supported(const sbfe_fbank_desc* d);
// packed-f32x2 kernel: additionally hop == 160 and the padded filter layout (mel_pad4)
bool fbank400_v3_supported(const sbfe_fbank_desc* d);
static bool fused_ok(const sbfe_fbank_desc* d) {
  return fbank400_v3_supported(d) || fbank400_supported(d);
}
}  // namespace sbfe

using namespace sbfe;

#ifdef SBFE_EXP_TIMELINE
namespace sbfe {
int tl_sync_v3(unsigned long long*, int);
int tl_sync_norm(unsigned long long*, int);
}
// kernels 0, 1 (producer, clamp) live in fbank400_v3.cu, 2, 3 (normalisation) in norm.cu
extern "C" __attribute__((visibility("default"))) int sbfe_exp_timeline(unsigned long long* h_out, int reset) {
  unsigned long long a[24], b[24];
  int rc = sbfe::tl_sync_v3(a, reset) | sbfe::tl_sync_norm(b, reset);
  for (int i = 0; i < 24; ++i) h_out[i] = i < 6 ? a[i] : b[i];
  return rc;
}
#endif

extern "C" {

int sbfe_abi_version(void) { return SBFE_ABI_VERSION; }

const char* sbfe_error_string(int code) {
  if (code == 0) return "ok";
  if (code == SBFE_E_INVAL) return "sbfe: invalid argument";
  if (code == SBFE_E_TOOBIG) return "sbfe: dimension too large for the kernels";
  if (code == SBFE_E_NOSPACE) return "sbfe: workspace too small";
  if (code == SBFE_E_STATIC_LIMIT)
    return "sbfe: request needs more scratch than the built-in array of the no-allocation entry "
           "point holds (SBFE_NORM_STATIC_WS_BYTES); use the variant that takes a workspace";
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "sbfe: unknown error";
}

int64_t sbfe_num_frames(int64_t n_samples, int32_t hop) {
  return hop > 0 && n_samples >= 0 ? 1 + n_samples / hop : -1;
}

int64_t sbfe_fbank_workspace_bytes(const sbfe_fbank_desc* d, int64_t B, int64_t L) {
  if (!d || d->hop < 1 || B < 0 || L < 0) return SBFE_E_INVAL;
  if (!mel_dense_mma(d) && (fused_ok(d) || fbank512_supported(d) || fbank_pow2_supported(d))) return 0;
  return B * (1 + L / d->hop) * (int64_t)(d->n_fft / 2 + 1) * (int64_t)sizeof(float);
}

int32_t sbfe_fbank_tile_frames(const sbfe_fbank_desc* d) {
  // a dense filterbank goes through the tensor-core projection, not the fused kernels
  return (d && desc_ok(d, true) && (fbank400_v3_supported(d) || fbank512_supported(d)) &&
          !mel_dense_mma(d)) ? 32 : 0;
}

int32_t sbfe_fbank_ragged_supported(const sbfe_fbank_desc* d) {
  return (d && desc_ok(d, true) && fbank400_v3_supported(d) && !mel_dense_mma(d)) ? 1 : 0;
}

int32_t sbfe_fbank_tile_stats_floats(const sbfe_fbank_desc* d) {
  return sbfe_fbank_tile_frames(d) ? (int32_t)tile_rec_words(d->n_mels) : 0;
}

int sbfe_fbank_forward(const sbfe_fbank_desc* d, const float* d_wav, int64_t B, int64_t L,
                       int64_t wav_stride, float* d_out, sbfe_maxkey_t* d_utt_max,
                       float* d_tile_stats, float top_db, void* d_ws,
                       int64_t ws_bytes, void* stream) {
  if (!desc_ok(d, true) || B < 0 || L < 0 || wav_stride < L)
    return SBFE_E_INVAL;
  if (B == 0) return 0;
  if ((!d_wav && L > 0) || !d_out) return SBFE_E_INVAL;
  cudaStream_t s = (cudaStream_t)stream;
  if (fbank400_v3_supported(d) && !mel_dense_mma(d))
    return launch_fbank400_v3(d, d_wav, 0, 1.f, nullptr, nullptr, B, L, wav_stride, d_out, d_utt_max,
                              d_tile_stats, top_db, s);
  if (fbank512_supported(d) && !mel_dense_mma(d))
    return launch_fbank512(d, d_wav, B, L, wav_stride, d_out, d_utt_max, d_tile_stats, top_db, s);
  if (d_tile_stats) return SBFE_E_INVAL;  // only where sbfe_fbank_tile_frames() != 0
  // dense (gaussian) filterbank: power spectrum to the workspace, projection on the tensor cores
  if (!mel_dense_mma(d)) {
    if (fused_ok(d)) return launch_fbank400(d, d_wav, B, L, wav_stride, d_out, d_utt_max, s);
    if (fbank_pow2_supported(d))   // n_fft = 256 / 512 / 1024 ...: FFT + mel + dB in one launch
      return launch_fbank_pow2(d, d_wav, B, L, wav_stride, d_out, d_utt_max, s);
  }
  if (!d_ws || ws_bytes < sbfe_fbank_workspace_bytes(d, B, L)) return SBFE_E_NOSPACE;
  int rc = launch_stft_generic(d, d_wav, B, L, wav_stride, (float*)d_ws, 1, 0, s);
  if (rc) return rc;
  return launch_mel_db(d, (const float*)d_ws, B, 1 + L / d->hop, d_out, d_utt_max, s);
}

int sbfe_fbank_forward_ragged(const sbfe_fbank_desc* d, const float* d_packed,
                              const int64_t* d_offsets, const int32_t* d_nsamples, int64_t B,
                              int64_t L, float* d_out, sbfe_maxkey_t* d_utt_max,
                              float* d_tile_stats, float top_db, void* stream) {
  if (!desc_ok(d, true) || B < 0 || L < 0) return SBFE_E_INVAL;
  if (B == 0) return 0;
  if (!d_packed || !d_offsets || !d_nsamples || !d_out) return SBFE_E_INVAL;
  if (!sbfe_fbank_ragged_supported(d)) return SBFE_E_INVAL;  // pad on the device and use sbfe_fbank_forward
  return launch_fbank400_v3(d, d_packed, 0, 1.f, d_offsets, d_nsamples, B, L, 0, d_out, d_utt_max,
                            d_tile_stats, top_db, (cudaStream_t)stream);
}

int sbfe_fbank_forward_ragged_pcm16(const sbfe_fbank_desc* d, const int16_t* d_packed,
                                    const int64_t* d_offsets, const int32_t* d_nsamples, int64_t B,
                                    int64_t L, float scale, float* d_out, sbfe_maxkey_t* d_utt_max,
                                    float* d_tile_stats, float top_db,
                                    void* stream) {
  if (!desc_ok(d, true) || B < 0 || L < 0 || !(scale > 0.f))
    return SBFE_E_INVAL;
  if (B == 0) return 0;
  if (!d_packed || !d_offsets || !d_nsamples || !d_out) return SBFE_E_INVAL;
  if (!sbfe_fbank_ragged_supported(d)) return SBFE_E_INVAL;  // convert on the device and use sbfe_fbank_forward
  return launch_fbank400_v3(d, d_packed, 1, scale, d_offsets, d_nsamples, B, L, 0, d_out, d_utt_max,
                            d_tile_stats, top_db, (cudaStream_t)stream);
}

int64_t sbfe_istft_workspace_bytes(int64_t B, int64_t T, int32_t n_fft) {
  if (B < 0 || T < 0 || n_fft < 2 || (n_fft & 1)) return SBFE_E_INVAL;
  return istft_workspace_bytes(B, T, n_fft);
}

int sbfe_istft_forward(const float* d_spec, int64_t B, int64_t T, int32_t n_fft, int32_t hop,
                       const float* d_window, int32_t center, int64_t out_len, float* d_out,
                       void* d_ws, int64_t ws_bytes, void* stream) {
  if (B < 0 || T < 1 || n_fft < 2 || (n_fft & 1) || hop < 1 || out_len < 0) return SBFE_E_INVAL;
  if (B == 0 || out_len == 0) return 0;
  if (!d_spec || !d_window || !d_out || !d_ws) return SBFE_E_INVAL;
  return launch_istft(d_spec, B, T, n_fft, hop, d_window, center, out_len, d_out, d_ws, ws_bytes,
                      (cudaStream_t)stream);
}

int64_t sbfe_stft_frames(const sbfe_fbank_desc* d, int64_t L, int32_t flags) {
  if (!desc_ok(d, false) || L < 0) return SBFE_E_INVAL;
  return stft_frames(d, L, flags);
}

int sbfe_stft_forward(const sbfe_fbank_desc* d, const float* d_wav, int64_t B, int64_t L,
                      int64_t wav_stride, float* d_out, int32_t power_only, int32_t flags,
                      void* stream) {
  if (!desc_ok(d, false) || B < 0 || L < 0 || wav_stride < L || (flags & ~7)) return SBFE_E_INVAL;
  // reflection needs the mirror image inside the signal (torch.stft raises otherwise); reflect
  // without centring pads nothing
  if ((flags & SBFE_STFT_REFLECT) && !(flags & SBFE_STFT_NO_CENTER) && L <= d->n_fft / 2)
    return SBFE_E_INVAL;
  if (B == 0) return 0;
  if ((!d_wav && L > 0) || !d_out) return SBFE_E_INVAL;
  return launch_stft_generic(d, d_wav, B, L, wav_stride, d_out, power_only, flags,
                             (cudaStream_t)stream);
}

int sbfe_spectral_magnitude(const float* d_stft, int64_t n, float power, int32_t take_log,
                            float eps, float* d_out, void* stream) {
  if (n < 0 || (n > 0 && (!d_stft || !d_out))) return SBFE_E_INVAL;
  return launch_spectral_magnitude(d_stft, n, power, take_log, eps, d_out, (cudaStream_t)stream);
}

int sbfe_filterbank_forward(const sbfe_fbank_desc* d, const float* d_spec, int64_t B, int64_t T,
                            float* d_out, sbfe_maxkey_t* d_utt_max, void* stream) {
  if (!d || d->n_fft < 0 || (d->n_fft & 1) || d->n_mels < 1 || !d->d_mel_start || !d->d_mel_count ||
      !d->d_mel_offset || !d->d_mel_weight || B < 0 || T < 0)
    return SBFE_E_INVAL;
  if (B == 0 || T == 0) return 0;
  if (!d_spec || !d_out) return SBFE_E_INVAL;
  return launch_mel_db(d, d_spec, B, T, d_out, d_utt_max, (cudaStream_t)stream);
}

int sbfe_post_forward(const float* d_x, int64_t B, int64_t T, int32_t F_in,
                      const sbfe_maxkey_t* d_utt_max, float top_db, const float* d_dct,
                      int32_t n_dct, int32_t deltas, int32_t delta_window, int32_t context,
                      int32_t left, int32_t right, float* d_out, void* stream) {
  if (B < 0 || T < 0 || F_in < 1 || left < 0 || right < 0) return SBFE_E_INVAL;
  if (d_dct && (n_dct < 1 || n_dct > F_in)) return SBFE_E_INVAL;
  if (B == 0 || T == 0) return 0;
  if (!d_x || !d_out) return SBFE_E_INVAL;
  return launch_post(d_x, B, T, F_in, d_utt_max, top_db, d_dct, n_dct, deltas, delta_window,
                     context, left, right, d_out, (cudaStream_t)stream);
}

int sbfe_norm_stats_tiles(const float* d_x, const float* d_lengths, int64_t B, int64_t T, int32_t F,
                          int32_t mean_norm, int32_t std_norm, float eps,
                          const float* d_tile_stats, int32_t tile_frames, float top_db,
                          float* d_mean, float* d_std, void* stream) {
  if (B < 0 || T < 0 || F < 0 || tile_frames < 1) return SBFE_E_INVAL;
  if (B == 0 || F == 0) return 0;
  if (!d_x || !d_lengths || !d_mean || !d_std || !d_tile_stats) return SBFE_E_INVAL;
  return launch_norm_stats_tiles(d_x, d_lengths, B, T, F, mean_norm, std_norm, eps, d_tile_stats,
                                 tile_frames, top_db, d_mean, d_std, (cudaStream_t)stream);
}

int sbfe_norm_sentence_tiles(const float* d_x, const float* d_lengths, int64_t B, int64_t T,
                             int32_t F, int32_t mean_norm, int32_t std_norm, float eps,
                             float* d_tile_stats, int32_t tile_frames, float top_db,
                             int32_t flags, void* d_out, void* stream) {
  if (B < 0 || T < 0 || F < 0 || tile_frames < 1) return SBFE_E_INVAL;
  if (B == 0 || F == 0) return 0;
  if (!d_x || !d_lengths || !d_tile_stats || !d_out) return SBFE_E_INVAL;
  if (flags & ~(SBFE_BESIDE_CLAMP | SBFE_OUT_BF16)) return SBFE_E_INVAL;
  if ((flags & SBFE_OUT_BF16) && d_out == (const void*)d_x) return SBFE_E_INVAL;
  return launch_norm_sentence_tiles(d_x, d_lengths, B, T, F, mean_norm, std_norm, eps, d_tile_stats,
                                    tile_frames, top_db, (flags & SBFE_BESIDE_CLAMP) != 0,
                                    (flags & SBFE_OUT_BF16) != 0, d_out, (cudaStream_t)stream);
}

int64_t sbfe_norm_workspace_bytes(int64_t B, int64_t T, int32_t F) {
  if (B < 0 || T < 0 || F < 0) return SBFE_E_INVAL;
  return norm_workspace_bytes(B, T, F);
}

int sbfe_norm_stats(const float* d_x, const float* d_lengths, int64_t B, int64_t T, int32_t F,
                    int32_t mean_norm, int32_t std_norm, float eps,
                    const sbfe_maxkey_t* d_utt_max, float top_db, float* d_mean, float* d_std,
                    void* d_ws, int64_t ws_bytes, void* stream) {
  if (B < 0 || T < 0 || F < 0) return SBFE_E_INVAL;
  if (B == 0 || F == 0) return 0;
  if (!d_x || !d_lengths || !d_mean || !d_std || !d_ws) return SBFE_E_INVAL;
  return launch_norm_stats(d_x, d_lengths, B, T, F, mean_norm, std_norm, eps, d_utt_max, top_db,
                           d_mean, d_std, d_ws, ws_bytes, (cudaStream_t)stream);
}

int sbfe_norm_stats_static(const float* d_x, const float* d_lengths, int64_t B, int64_t T, int32_t F,
                           int32_t mean_norm, int32_t std_norm, float eps,
                           const sbfe_maxkey_t* d_utt_max, float top_db, float* d_mean,
                           float* d_std, void* stream) {
  if (B < 0 || T < 0 || F < 0) return SBFE_E_INVAL;
  if (B == 0 || F == 0) return 0;
  if (!d_x || !d_lengths || !d_mean || !d_std) return SBFE_E_INVAL;
  return launch_norm_stats_static(d_x, d_lengths, B, T, F, mean_norm, std_norm, eps, d_utt_max,
                                  top_db, d_mean, d_std, (cudaStream_t)stream);
}

int sbfe_norm_reduce(const float* d_mean, const float* d_std, int64_t B, int32_t F, float* d_cur,
                     void* stream) {
  if (B < 0 || F < 1 || !d_mean || !d_std || !d_cur) return SBFE_E_INVAL;
  return launch_norm_reduce(d_mean, d_std, B, F, d_cur, (cudaStream_t)stream);
}

int sbfe_norm_update(const float* d_cur, int32_t F, int32_t first, float weight,
                     float* d_glob_mean, float* d_glob_std, void* stream) {
  if (F < 1 || !d_cur || !d_glob_mean || !d_glob_std) return SBFE_E_INVAL;
  return launch_norm_update(d_cur, F, first, weight, d_glob_mean, d_glob_std,
                            (cudaStream_t)stream);
}

int sbfe_norm_apply(const float* d_x, int64_t B, int64_t T, int32_t F, const float* d_mean,
                    const float* d_std, int32_t per_utt, const sbfe_maxkey_t* d_utt_max,
                    float top_db, int32_t flags, void* d_out, void* stream) {
  if (B < 0 || T < 0 || F < 0 || (flags & ~SBFE_OUT_BF16)) return SBFE_E_INVAL;
  if (B == 0 || T == 0 || F == 0) return 0;
  if (!d_x || !d_out || !d_mean || !d_std) return SBFE_E_INVAL;
  if ((flags & SBFE_OUT_BF16) && d_out == (const void*)d_x) return SBFE_E_INVAL;
  return launch_norm_apply(d_x, B, T, F, d_mean, d_std, per_utt, d_utt_max, top_db,
                           (flags & SBFE_OUT_BF16) != 0, d_out, (cudaStream_t)stream);
}

int sbfe_norm_speaker_update(const float* d_mean, const float* d_std, int64_t B, int32_t F,
                             const int32_t* d_row, const int32_t* d_mode, const float* d_weight,
                             float* d_table_mean, float* d_table_std, float* d_out_mean,
                             float* d_out_std, void* stream) {
  if (B < 0 || F < 0) return SBFE_E_INVAL;
  if (B == 0 || F == 0) return 0;
  if (!d_mean || !d_std || !d_row || !d_mode || !d_weight || !d_out_mean || !d_out_std)
    return SBFE_E_INVAL;
  return launch_norm_speaker_update(d_mean, d_std, B, F, d_row, d_mode, d_weight, d_table_mean,
                                    d_table_std, d_out_mean, d_out_std, (cudaStream_t)stream);
}

int sbfe_specaugment_forward(const float* d_x, int64_t B, int64_t T, int32_t F, int32_t warp_c,
                             int32_t warp_w, int32_t warp_mode, const int64_t* d_fpos,
                             const int64_t* d_flen, int32_t n_fmask, const int64_t* d_tpos,
                             const int64_t* d_tlen, int32_t n_tmask, const float* d_fill,
                             float fill, float* d_out, void* stream) {
  if (B < 0 || T < 0 || F < 0) return SBFE_E_INVAL;
  if (B == 0 || T == 0 || F == 0) return 0;
  if (!d_x || !d_out) return SBFE_E_INVAL;
  return launch_specaugment(d_x, B, T, F, warp_c, warp_w, warp_mode, d_fpos, d_flen, n_fmask,
                            d_tpos, d_tlen, n_tmask, d_fill, fill, d_out, (cudaStream_t)stream);
}

int sbfe_wav_drop_chunks(float* d_out, int64_t B, int64_t L, int32_t channels,
                         const int32_t* d_chunks, int32_t n_chunks, int64_t max_len,
                         const float* d_noise, const float* d_noise_max, void* stream) {
  if (B < 0 || L < 0 || channels < 1 || n_chunks < 0 || (d_noise != nullptr) != (d_noise_max != nullptr))
    return SBFE_E_INVAL;
  if (B == 0 || L == 0 || n_chunks == 0) return 0;
  if (!d_out || !d_chunks) return SBFE_E_INVAL;
  return launch_drop_chunks(d_out, L, channels, d_chunks, n_chunks, max_len, d_noise, d_noise_max,
                            (cudaStream_t)stream);
}

int sbfe_wav_fir(const float* d_x, int64_t B, int64_t L, const float* d_taps, int32_t n_taps,
                 int32_t pad, float* d_out, void* stream) {
  if (B < 0 || L < 0) return SBFE_E_INVAL;
  if (B == 0 || L == 0) return 0;
  if (!d_x || !d_taps || !d_out || d_x == d_out) return SBFE_E_INVAL;
  return launch_fir_same(d_x, B, L, d_taps, n_taps, pad, d_out, (cudaStream_t)stream);
}

int sbfe_wav_clip(const float* d_x, int64_t n, float value, float* d_out, void* stream) {
  if (n < 0 || !(value >= 0.f)) return SBFE_E_INVAL;
  if (n == 0) return 0;
  if (!d_x || !d_out) return SBFE_E_INVAL;
  return launch_clip(d_x, n, value, d_out, (cudaStream_t)stream);
}

int sbfe_wav_notch_taps(const float* h_freqs, int32_t n_freqs, float width, int32_t filter_len,
                        float* d_taps, void* stream) {
  if ((n_freqs > 0 && !h_freqs) || !d_taps) return SBFE_E_INVAL;
  return launch_notch_taps(h_freqs, n_freqs, width, filter_len, d_taps, (cudaStream_t)stream);
}

int64_t sbfe_wav_abs_chunks(int64_t L) { return L < 0 ? SBFE_E_INVAL : abs_chunks(L); }

int sbfe_wav_abs_sums(const float* d_x, int64_t B, int64_t L, float* d_partials, void* stream) {
  if (B < 0 || L < 0) return SBFE_E_INVAL;
  if (B == 0 || L == 0) return 0;
  if (!d_x || !d_partials) return SBFE_E_INVAL;
  return launch_abs_sums(d_x, B, L, d_partials, (cudaStream_t)stream);
}

int sbfe_wav_add_noise(const float* d_x, const float* d_noise, int64_t B, int64_t L,
                       const float* d_partials, const float* d_lengths, const float* d_snr_db,
                       float* d_out, uint32_t* d_absmax, void* stream) {
  if (B < 0 || L < 0) return SBFE_E_INVAL;
  if (B == 0 || L == 0) return 0;
  if (!d_x || !d_noise || !d_partials || !d_lengths || !d_snr_db || !d_out) return SBFE_E_INVAL;
  return launch_add_noise(d_x, d_noise, B, L, d_partials, d_lengths, d_snr_db, d_out, d_absmax,
                          (cudaStream_t)stream);
}

int sbfe_wav_add_babble(const float* d_x, int64_t B, int64_t L, int32_t speakers, float* d_partials,
                        const float* d_lengths, const float* d_snr_db, float* d_out, void* stream) {
  if (B < 0 || L < 0 || speakers < 1) return SBFE_E_INVAL;
  if (B == 0 || L == 0) return 0;
  if (!d_x || !d_partials || !d_lengths || !d_snr_db || !d_out || d_out == d_x) return SBFE_E_INVAL;
  return launch_add_babble(d_x, B, L, speakers, d_partials, d_partials + B * abs_chunks(L), d_lengths,
                           d_snr_db, d_out, (cudaStream_t)stream);
}

int sbfe_wav_peak_normalize(float* d_x, int64_t B, int64_t L, const uint32_t* d_absmax, void* stream) {
  if (B < 0 || L < 0) return SBFE_E_INVAL;
  if (B == 0 || L == 0) return 0;
  if (!d_x || !d_absmax) return SBFE_E_INVAL;
  return launch_peak_normalize(d_x, B, L, d_absmax, (cudaStream_t)stream);
}

int sbfe_wav_resample(const float* d_x, int64_t B, int64_t L, const float* d_weights,
                      const int32_t* d_first, int32_t phases, int32_t in_stride, int32_t width,
                      int64_t L_out, float* d_out, void* stream) {
  if (B < 0 || L < 0 || L_out < 0) return SBFE_E_INVAL;
  if (B == 0 || L_out == 0) return 0;
  if (!d_x || !d_weights || !d_first || !d_out) return SBFE_E_INVAL;
  return launch_resample(d_x, B, L, d_weights, d_first, phases, in_stride, width, L_out, d_out,
                         (cudaStream_t)stream);
}

int64_t sbfe_mean_workspace_bytes(void) { return mean_workspace_bytes(); }

int sbfe_mean(const float* d_x, int64_t n, float* d_mean, void* d_workspace,
              int64_t workspace_bytes, void* stream) {
  if (!d_x || !d_mean || !d_workspace) return SBFE_E_INVAL;
  return launch_mean(d_x, n, d_mean, d_workspace, workspace_bytes, (cudaStream_t)stream);
}

}  // extern "C"
