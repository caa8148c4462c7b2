// K-state container kernels: bit packing and initial K^n (evo/variational/utils.py:19-152).
#include "common.cuh"

namespace evo {

// one warp packs 32 latents of one row per ballot
__global__ void pack_kernel(const uint8_t *__restrict__ bools, uint32_t *__restrict__ packed, int64_t rows, int H,
                            int Hw) {
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int64_t total = rows * Hw;
  for (int64_t item = warp; item < total; item += nwarps) {
    int64_t r = item / Hw;
    int w = (int)(item - r * Hw);
    int h = (w << 5) + lane;
    bool bit = (h < H) && bools[r * (int64_t)H + h] != 0;
    uint32_t word = __ballot_sync(0xffffffffu, bit);
    if (lane == 0) packed[item] = word;
  }
}

__global__ void unpack_kernel(const uint32_t *__restrict__ packed, uint8_t *__restrict__ bools, int64_t rows, int H,
                              int Hw) {
  int64_t total = rows * (int64_t)H;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / H;
    int h = (int)(i - r * H);
    bools[i] = (packed[r * Hw + (h >> 5)] >> (h & 31)) & 1u;
  }
}

// S distinct Bernoulli(p) states per datapoint, none equal to the permanent all-zero state when
// S_perm > 0; with a background unit the last latent is always on (utils.py:96-138).  One warp per
// datapoint; a state is redrawn until it differs from all earlier ones of that datapoint.
// `use_hash`: a 64-bit hash of every accepted state of the warp's datapoint is kept in shared memory and the words are
// compared only on a hash match (same accept / reject decisions as the plain word compare, S^2 Hw / 2 fewer loads).
__global__ void init_states_kernel(uint32_t *__restrict__ states, int64_t N, int S, int H, int Hw, int S_perm,
                                   int background, double p, uint64_t seed, int64_t n_offset, int use_hash) {
  extern __shared__ unsigned long long hash_sm[];  // [warps per block][S]
  unsigned long long *hk = hash_sm + (size_t)(threadIdx.x >> 5) * S;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int Hm = background ? H - 1 : H;
  const uint32_t thr = (uint32_t)fmin(4294967295.0, p * 4294967296.0);
  for (int64_t n = warp; n < N; n += nwarps) {
    RngCtx rng(seed, n_offset + n, 0u);
    uint32_t *base = states + n * (int64_t)S * Hw;
    uint32_t attempt = 0;
    for (int k = 0; k < S;) {
      // lanes build the words of candidate state k cooperatively
      unsigned long long hv = 0;
      for (int w = lane; w < Hw; w += 32) {
        uint32_t word = 0;
        for (int q = 0; q < 8; ++q) {
          uint4 r = rng.draw(RNG_INIT, 0, (attempt * (uint32_t)Hw + (uint32_t)w) * 8u + (uint32_t)q);
          uint32_t rr[4] = {r.x, r.y, r.z, r.w};
          for (int j = 0; j < 4; ++j) {
            int h = (w << 5) + q * 4 + j;
            if (h < Hm && rr[j] < thr) word |= 1u << (q * 4 + j);
          }
        }
        if (background && ((H - 1) >> 5) == w) word |= 1u << ((H - 1) & 31);
        base[(int64_t)k * Hw + w] = word;
        hv += ((unsigned long long)word + 1ull) * (0x9E3779B97F4A7C15ull + 2ull * (unsigned long long)w);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) hv += __shfl_xor_sync(0xffffffffu, hv, o);
      hv ^= hv >> 29;
      __syncwarp();
      ++attempt;
      // reject duplicates / permanent state
      bool bad = false;
      if (S_perm > 0) {
        int nz = 0;
        for (int w = lane; w < Hw; w += 32) nz |= (base[(int64_t)k * Hw + w] != 0);
        bad = !__any_sync(0xffffffffu, nz);
      }
      for (int j = lane; j < k && !bad; j += 32) {
        if (use_hash && hk[j] != hv) continue;
        bool same = true;
        for (int w = 0; w < Hw; ++w) same &= (base[(int64_t)j * Hw + w] == base[(int64_t)k * Hw + w]);
        if (same) bad = true;
      }
      bad = __any_sync(0xffffffffu, bad);
      // with p tiny and S large the space of likely states is small; raise p after many failures
      if (!bad) {
        if (use_hash && lane == 0) hk[k] = hv;
        ++k;
      }
      __syncwarp();
    }
  }
}

}  // namespace evo

using namespace evo;

extern "C" int evo_states_pack(const uint8_t *bools, uint32_t *packed, int64_t rows, int H, void *stream) {
  EVO_REQUIRE(rows >= 0 && H > 0, "bad sizes");
  if (rows == 0) return EVO_OK;
  int Hw = words_for(H);
  int64_t items = rows * Hw;
  int blocks = (int)((items * 32 + 255) / 256 < (int64_t)sm_count() * 16 ? (items * 32 + 255) / 256 : (int64_t)sm_count() * 16);
  pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(bools, packed, rows, H, Hw);
  EVO_LAUNCH_CHECK("pack");
  return EVO_OK;
}

extern "C" int evo_states_unpack(const uint32_t *packed, uint8_t *bools, int64_t rows, int H, void *stream) {
  EVO_REQUIRE(rows >= 0 && H > 0, "bad sizes");
  if (rows == 0) return EVO_OK;
  int Hw = words_for(H);
  int64_t total = rows * (int64_t)H;
  int blocks = (int)((total + 255) / 256 < (int64_t)sm_count() * 16 ? (total + 255) / 256 : (int64_t)sm_count() * 16);
  unpack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(packed, bools, rows, H, Hw);
  EVO_LAUNCH_CHECK("unpack");
  return EVO_OK;
}

extern "C" int evo_states_init(uint32_t *states, int64_t N, int S, int H, int S_perm, int background, double p,
                               uint64_t seed, int64_t n_offset, void *stream) {
  EVO_REQUIRE(N >= 0 && S > 0 && H > 0, "bad sizes");
  EVO_REQUIRE(p > 0.0 && p < 1.0, "p must be in (0,1)");
  int Hm = background ? H - 1 : H;
  // the number of distinct states must be reachable
  EVO_REQUIRE(Hm >= 31 || (1ll << Hm) - (S_perm ? 1 : 0) >= S, "S exceeds the number of distinct states");
  if (N == 0) return EVO_OK;
  int Hw = words_for(H);
  int64_t warps = N;
  int blocks = (int)((warps * 32 + 127) / 128 < (int64_t)sm_count() * 16 ? (warps * 32 + 127) / 128 : (int64_t)sm_count() * 16);
  const size_t hbytes = (size_t)4 * S * sizeof(unsigned long long);
  const int use_hash = hbytes <= 48 * 1024;
  init_states_kernel<<<blocks, 128, use_hash ? hbytes : 0, (cudaStream_t)stream>>>(states, N, S, H, Hw, S_perm, background, p, seed,
                                                                                  n_offset, use_hash);
  EVO_LAUNCH_CHECK("init_states");
  return EVO_OK;
}
