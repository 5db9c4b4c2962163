// Shared device helpers for the sbfe kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/sbfe.h"

namespace sbfe {

// Order-preserving float -> uint32 key (0 is reserved for "empty").
__device__ __forceinline__ uint32_t max_key_encode(float v) {
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float max_key_decode(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}
// floor for the top_db clamp of utterance b; -inf when clamping is off
__device__ __forceinline__ float clamp_floor(const sbfe_maxkey_t* utt_max, int64_t b,
                                             float top_db) {
  if (utt_max == nullptr) return -INFINITY;
  return max_key_decode(utt_max[b]) - top_db;
}

// Per-tile record of the fused Fbank kernel (include/sbfe.h, sbfe_fbank_tile_stats_floats):
// [3][F] floats = column mean / M2 / min, then kTileKeys words of which the first is the max key
// of the tile (the rest pads the record to 16 bytes); consumers rebuild the utterance maximum
// from the keys: no global atomics, no memset.
constexpr int kTileKeys = 4;
__host__ __device__ inline size_t tile_rec_words(int F) { return (size_t)3 * F + kTileKeys; }

// top_db floor of utterance b from the tile records (whole CTA must call; uses one barrier pair)
__device__ __forceinline__ float block_floor_from_tiles(const float* tile_stats, int64_t b, int tiles,
                                                        int F, float top_db,
                                                        unsigned* key_out = nullptr) {
  __shared__ unsigned s_floor_key;
  if (threadIdx.x == 0) s_floor_key = 0u;
  __syncthreads();
  const size_t rec = tile_rec_words(F);
  const unsigned* base = reinterpret_cast<const unsigned*>(tile_stats) + (size_t)b * tiles * rec + 3 * F;
  unsigned k = 0u;
  for (int i = threadIdx.x; i < tiles * kTileKeys; i += blockDim.x) {
    const unsigned v = __ldg(base + (size_t)(i / kTileKeys) * rec + (i % kTileKeys));
    k = v > k ? v : k;
  }
  k = __reduce_max_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0 && k) atomicMax(&s_floor_key, k);
  __syncthreads();
  const unsigned key = s_floor_key;
  if (key_out != nullptr) *key_out = key;   // the utterance's max key (0: no record held one)
  return key ? max_key_decode(key) - top_db : -INFINITY;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// amplitude_to_DB without the top_db clamp (processing/features.py:701-702).
// log10 through the MUFU lg2 unit: |err| <= ~2^-22 * |log2 x| -> < 3e-5 dB.
struct DbParams {
  float scale;     // multiplier * log10(2)
  float amin;
  float offset;    // multiplier * log10(max(amin, ref))
  float floor_db;  // exact host-rounded value for x <= amin
  int log_mel;
};
// x > amin >= ~1e-38 here, so the flush-to-zero form is exact enough and branch-free
__device__ __forceinline__ float fast_log2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float to_db(float x, const DbParams& p) {
  if (!p.log_mel) return x;
  // !(x > amin) also catches NaN-free zero/negative inputs exactly like clamp()
  return (x > p.amin) ? fmaf(fast_log2(x), p.scale, -p.offset) : p.floor_db;
}

inline int check_last(void) {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : (int)e;
}

// ---- programmatic dependent launch (stream order kept, launch latency overlapped) ----------
// A step is four short kernels back to back; each one's launch + prologue would otherwise sit
// on the critical path.  Every kernel of this library (1) executes pdl_wait() before it reads
// or writes anything another kernel may touch - it returns once the preceding kernel in the
// stream has completed and its writes are visible - and (2) calls pdl_launch_dependents()
// AFTER that wait, so that the NEXT kernel may be scheduled early and park in its own
// pdl_wait().  Two invariants follow: completion is transitive along the stream (every kernel
// waits before it exits), and when a kernel starts, everything before its DIRECT predecessor
// has completed (the predecessor triggered only after its own wait).  The second one lets a
// kernel that reads nothing its direct predecessor writes run beside it and wait at its END
// instead (norm_apply_kernel<., true> behind norm_stats_prefix_kernel, norm.cu).  Kernels of other
// libraries never trigger early, so mixing with them degrades to ordinary stream order.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

#ifdef SBFE_EXP_TIMELINE
// experiment build (profiles/exp_timeline.sh): per kernel id, min start / min wait-return / max end
// of its CTAs in %globaltimer ns; one copy of the stamps per translation unit
static __device__ unsigned long long g_timeline[3 * 8];
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void tl_mark(int kernel, int what) {
  if (threadIdx.x == 0) {
    if (what < 2) atomicMin(&g_timeline[3 * kernel + what], gtimer());
    else atomicMax(&g_timeline[3 * kernel + 2], gtimer());
  }
}
#define SBFE_TL(kernel, what) tl_mark(kernel, what)
// host side of one translation unit's copy: read (and optionally re-arm) the stamps
#define SBFE_TL_DEFINE_SYNC(name)                                                             \
  int name(unsigned long long* h_out, int reset) {                                            \
    cudaDeviceSynchronize();                                                                  \
    cudaError_t e = cudaMemcpyFromSymbol(h_out, g_timeline, sizeof(unsigned long long) * 24); \
    if (reset) {                                                                              \
      unsigned long long init[24];                                                            \
      for (int i = 0; i < 24; ++i) init[i] = (i % 3 == 2) ? 0ull : ~0ull;                     \
      cudaMemcpyToSymbol(g_timeline, init, sizeof(init));                                     \
    }                                                                                         \
    return (int)e;                                                                            \
  }
#else
#define SBFE_TL(kernel, what)
#endif

// SBFE_PDL=0 in the environment turns the launch attribute off (plain stream order).
inline bool pdl_enabled() {
  static const int on = [] {
    const char* v = getenv("SBFE_PDL");
    return (v != nullptr && v[0] == '0') ? 0 : 1;
  }();
  return on != 0;
}

template <typename... KArgs, typename... Args>
inline int launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                      cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return (int)e;
  }
  return check_last();
}

}  // namespace sbfe
