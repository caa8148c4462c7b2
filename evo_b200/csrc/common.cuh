// Shared device/host helpers for the evo_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "../../include/evo_b200.h"

namespace evo {

// ------------------------------------------------------------------------------------------
// host-side error plumbing
// ------------------------------------------------------------------------------------------
void set_error(const char *fmt, ...);
int check_cuda(cudaError_t e, const char *what);
int sm_count();

#define EVO_CUDA(call)                                     \
  do {                                                     \
    int _rc = ::evo::check_cuda((call), #call);            \
    if (_rc) return _rc;                                   \
  } while (0)
void count_launch();
#define EVO_LAUNCH_CHECK(name)        \
  do {                                \
    ::evo::count_launch();            \
    EVO_CUDA(cudaGetLastError());     \
  } while (0)
#define EVO_REQUIRE(cond, msg)                             \
  do {                                                     \
    if (!(cond)) {                                         \
      ::evo::set_error("%s: %s", __func__, msg);          \
      return EVO_ERR_ARG;                                  \
    }                                                      \
  } while (0)

// ------------------------------------------------------------------------------------------
// Philox-4x32-10 counter RNG (Salmon et al. 2011).  Key = 64-bit seed, counter = 4 x 32 bit.
// Counter convention used everywhere: (global datapoint index lo, hi | purpose << 24,
// iteration, draw index) so that streams do not depend on the number of ranks.
// ------------------------------------------------------------------------------------------
struct Philox {
  uint32_t k0, k1;
  __host__ __device__ explicit Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t &hi, uint32_t &lo) {
#ifdef __CUDA_ARCH__
    hi = __umulhi(a, b);
    lo = a * b;
#else
    uint64_t p = (uint64_t)a * b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
#endif
  }
  __host__ __device__ inline uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      uint32_t h0, l0, h1, l1;
      mulhilo(0xD2511F53u, c0, h0, l0);
      mulhilo(0xCD9E8D57u, c2, h1, l1);
      uint32_t n0 = h1 ^ c1 ^ a, n1 = l1, n2 = h0 ^ c3 ^ b, n3 = l0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      a += 0x9E3779B9u;
      b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

// purposes (upper byte of counter word 1)
enum : uint32_t { RNG_INIT = 1, RNG_SELECT = 2, RNG_MUTATE = 3, RNG_CROSS = 4, RNG_FLIP2 = 5 };

struct RngCtx {
  Philox ph;
  uint32_t n_lo, n_hi, iter;
  __device__ RngCtx(uint64_t seed, int64_t n_global, uint32_t iteration)
      : ph(seed), n_lo((uint32_t)n_global), n_hi((uint32_t)((uint64_t)n_global >> 32) & 0xFFFFu), iter(iteration) {}
  // purpose: 8 bit, sub: 8 bit (generation), idx: 32 bit draw index
  __device__ inline uint4 draw(uint32_t purpose, uint32_t sub, uint32_t idx) const {
    return ph(n_lo, n_hi | (purpose << 24) | ((sub & 0xFFu) << 16), iter, idx);
  }
};

__device__ __forceinline__ double u01(uint32_t a, uint32_t b) {
  // 53-bit uniform in (0,1)
  uint64_t v = (((uint64_t)a << 32) | b) >> 11;
  return ((double)v + 0.5) * (1.0 / 9007199254740992.0);
}
__device__ __forceinline__ uint32_t urange(uint32_t r, uint32_t n) { return __umulhi(r, n); }  // [0,n)

// ------------------------------------------------------------------------------------------
// packed states
// ------------------------------------------------------------------------------------------
__host__ __device__ inline int words_for(int H) { return (H + 31) >> 5; }

// cheap 32-bit mixing hash (murmur3 finaliser style); equality is always verified on the words
__device__ __forceinline__ uint32_t hash_state(const uint32_t *w, int Hw) {
  uint32_t h = 0x9E3779B9u;
  for (int i = 0; i < Hw; ++i) {
    uint32_t k = w[i] * 0xCC9E2D51u;
    k = __funnelshift_l(k, k, 15) * 0x1B873593u;
    h ^= k;
    h = __funnelshift_l(h, h, 13) * 5u + 0xE6546B64u;
  }
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ bool same_state(const uint32_t *a, const uint32_t *b, int Hw) {
  for (int i = 0; i < Hw; ++i)
    if (a[i] != b[i]) return false;
  return true;
}
// lexicographic order with latent 0 most significant (what np.unique yields on the reference's
// void view of bool rows, eas.py:252-256): returns true if a < b
__device__ __forceinline__ bool lex_less(const uint32_t *a, const uint32_t *b, int Hw) {
  for (int i = 0; i < Hw; ++i) {
    if (a[i] != b[i]) return __brev(a[i]) < __brev(b[i]);
  }
  return false;
}
__device__ __forceinline__ int popc_state(const uint32_t *w, int Hw) {
  int c = 0;
  for (int i = 0; i < Hw; ++i) c += __popc(w[i]);
  return c;
}

// lpj_reset_check (_models.py:567-596): NaN -> lowest f64, +-inf -> B_max = 0.  flags: 1 nan, 2 small, 4 inf
__device__ __forceinline__ double lpj_sanitise(double v, unsigned &flags) {
  if (isnan(v)) { flags |= 1u; return -DBL_MAX; }
  if (isinf(v)) { flags |= (v < 0 ? 6u : 4u); return 0.0; }
  return v;
}

// ------------------------------------------------------------------------------------------
// group (T threads = one datapoint) primitives.  T is a multiple of 32; T == blockDim.x.
// ------------------------------------------------------------------------------------------
template <int T>
__device__ __forceinline__ void gsync() {
  if (T == 32) __syncwarp(); else __syncthreads();
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// scratch: >= T/32 doubles of shared memory
template <int T>
__device__ __forceinline__ double group_sum(double v, double *scratch) {
  v = warp_sum(v);
  if (T == 32) return v;
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) scratch[w] = v;
  __syncthreads();
  double r = 0;
#pragma unroll
  for (int i = 0; i < T / 32; ++i) r += scratch[i];
  return r;
}
template <int T>
__device__ __forceinline__ double group_max(double v, double *scratch) {
  v = warp_max(v);
  if (T == 32) return v;
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) scratch[w] = v;
  __syncthreads();
  double r = scratch[0];
#pragma unroll
  for (int i = 1; i < T / 32; ++i) r = fmax(r, scratch[i]);
  return r;
}
template <int T>
__device__ __forceinline__ int group_sum_i(int v, int *scratch) {
  v = warp_sum_i(v);
  if (T == 32) return v;
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) scratch[w] = v;
  __syncthreads();
  int r = 0;
#pragma unroll
  for (int i = 0; i < T / 32; ++i) r += scratch[i];
  return r;
}
template <int T>
__device__ __forceinline__ unsigned group_or(unsigned v, int *scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v |= __shfl_xor_sync(0xffffffffu, v, o);
  if (T == 32) return v;
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) scratch[w] = (int)v;
  __syncthreads();
  unsigned r = 0;
#pragma unroll
  for (int i = 0; i < T / 32; ++i) r |= (unsigned)scratch[i];
  return r;
}

// argmax over the group of (key, idx) with ties broken towards the smaller idx; idx < 0 = none.
struct KeyIdx {
  double key;
  int idx;
};
__device__ __forceinline__ KeyIdx better(KeyIdx a, KeyIdx b) {
  if (b.idx < 0) return a;
  if (a.idx < 0) return b;
  if (b.key > a.key || (b.key == a.key && b.idx < a.idx)) return b;
  return a;
}
template <int T>
__device__ __forceinline__ KeyIdx group_argmax(KeyIdx v, double *sk, int *si) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    KeyIdx other;
    other.key = __shfl_xor_sync(0xffffffffu, v.key, o);
    other.idx = __shfl_xor_sync(0xffffffffu, v.idx, o);
    v = better(v, other);
  }
  if (T == 32) return v;
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) { sk[w] = v.key; si[w] = v.idx; }
  __syncthreads();
  KeyIdx r; r.key = sk[0]; r.idx = si[0];
#pragma unroll
  for (int i = 1; i < T / 32; ++i) { KeyIdx o; o.key = sk[i]; o.idx = si[i]; r = better(r, o); }
  return r;
}

// exclusive prefix sum of a 0/1 flag over the group; returns (offset, total)
template <int T>
__device__ __forceinline__ int group_scan_flag(bool flag, int &total, int *scratch) {
  unsigned m = __ballot_sync(0xffffffffu, flag);
  int l = threadIdx.x & 31, w = threadIdx.x >> 5;
  int off = __popc(m & ((1u << l) - 1u));
  if (T == 32) { total = __popc(m); return off; }
  __syncthreads();
  if (l == 0) scratch[w] = __popc(m);
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < T / 32; ++i) { if (i < w) base += scratch[i]; tot += scratch[i]; }
  total = tot;
  return base + off;
}

}  // namespace evo
