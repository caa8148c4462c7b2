// Device-side data generation (Model.generate_data evo/models/_models.py:73-99 -> BSC.generate_from_hidden
// bsc.py:27-57, SSSC.generate_from_hidden sssc.py:66-102) and an FMA throughput micro-benchmark (the
// FP32 / FP64 SIMT denominators of the ES3C roofline).
//
// A dataset row is a pure function of (seed, GLOBAL row index): every rank generates its own shard of ONE
// dataset, whatever the number of ranks.  oracle/datagen.py restates the same definition in NumPy (the CPU
// arm of bench.py draws its sample from it).  Draws (Philox-4x32-10, counter = (n_lo, n_hi | purpose << 24, 0, idx)):
//   s_h      purpose GEN_S,     idx = h / 4, component h % 4:  s_h = (r < floor(pi_h 2^32))
//   n_h      purpose GEN_Z,     idx = h:      Box-Muller cosine branch of (u01(x,y), u01(z,w))
//   eps_d    purpose GEN_NOISE, idx = d / 2:  Box-Muller, cosine branch for even d, sine branch for odd d
//   BSC:  y_d = sum_{h in s, ascending} W_dh + sigma eps_d
//   ES3C: z_A = mu_A + L n_A with L = chol(Psi_AA) (or sqrt(diag Psi) n for a diagonal Psi),
//         y_d = sum_{h in A, ascending} W_dh z_h + sigma eps_d
#include "common.cuh"

namespace evo {

enum : uint32_t { RNG_GEN_S = 6, RNG_GEN_Z = 7, RNG_GEN_NOISE = 8 };
constexpr int GEN_WARPS = 4, GEN_CHOL = 32;

__device__ __forceinline__ void box_muller(uint4 r, double &c, double &s) {
  const double u1 = u01(r.x, r.y), u2 = u01(r.z, r.w);
  const double rad = sqrt(-2.0 * log(u1));
  double sn, cs;
  sincospi(2.0 * u2, &sn, &cs);
  c = rad * cs;
  s = rad * sn;
}

struct GenArgs {
  const double *Wt, *pies, *mus, *Psi;
  double pi, sigma;
  int64_t N, n_offset;
  int D, H, Hw, psi_mode;  // psi_mode: 0 = BSC (z = 1), 1 = diagonal Psi, 2 = full Psi (Cholesky of Psi_AA, a <= 32)
  uint64_t seed;
  float *Y, *Z, *Ymean;
  uint32_t *S;
  int32_t *overflow;  // rows whose support exceeded GEN_CHOL under psi_mode 2 (they fall back to the diagonal)
};

__global__ void __launch_bounds__(32 * GEN_WARPS) generate_kernel(GenArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int H = a.H, D = a.D, Hw = a.Hw;
  const size_t per_warp = (size_t)8 * H + (((size_t)4 * Hw + 7) & ~(size_t)7) + (size_t)8 * GEN_CHOL * (GEN_CHOL + 1) + 64;
  unsigned char *mine = smem_raw + (size_t)warp * per_warp;
  double *z = reinterpret_cast<double *>(mine);                       // [H]
  double *L = z + H;                                                  // [GEN_CHOL][GEN_CHOL + 1]
  uint16_t *act = reinterpret_cast<uint16_t *>(L + GEN_CHOL * (GEN_CHOL + 1));  // [32]
  uint32_t *sw = reinterpret_cast<uint32_t *>(act + 32);            // [Hw]
  for (int64_t n = (int64_t)blockIdx.x * GEN_WARPS + warp; n < a.N; n += (int64_t)gridDim.x * GEN_WARPS) {
    RngCtx rng(a.seed, a.n_offset + n, 0u);
    // ---- s ~ Bernoulli(pi_h) ------------------------------------------------------------------------------
    for (int w = lane; w < Hw; w += 32) {
      uint32_t word = 0;
      for (int q = 0; q < 8; ++q) {
        const uint4 r = rng.draw(RNG_GEN_S, 0, (uint32_t)w * 8u + (uint32_t)q);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
        for (int j = 0; j < 4; ++j) {
          const int h = (w << 5) + q * 4 + j;
          if (h < H) {
            const double p = a.pies ? a.pies[h] : a.pi;
            const uint32_t thr = (uint32_t)fmin(4294967295.0, fmax(p, 0.0) * 4294967296.0);
            if (rr[j] < thr) word |= 1u << (q * 4 + j);
          }
        }
      }
      sw[w] = word;
      if (a.S) a.S[n * Hw + w] = word;
    }
    __syncwarp();
    // ---- slab values ---------------------------------------------------------------------------------------
    int na = 0;
    for (int w = 0; w < Hw; ++w) na += __popc(sw[w]);
    for (int h = lane; h < H; h += 32) {
      double v = 0.0;
      if ((sw[h >> 5] >> (h & 31)) & 1u) {
        if (a.psi_mode == 0) v = 1.0;
        else {
          double c, s;
          box_muller(rng.draw(RNG_GEN_Z, 0, (uint32_t)h), c, s);
          v = c;  // standard normal for now
          if (a.psi_mode == 1 || na > GEN_CHOL) v = a.mus[h] + sqrt(fmax(a.Psi[(int64_t)h * H + h], 0.0)) * c;
        }
      }
      z[h] = v;
    }
    __syncwarp();
    if (a.psi_mode == 2 && na > GEN_CHOL && lane == 0 && a.overflow) atomicAdd(a.overflow, 1);
    if (a.psi_mode == 2 && na > 0 && na <= GEN_CHOL) {
      // z_A = mu_A + chol(Psi_AA) n_A  (lower triangle of Psi is read; lane = row)
      if (lane == 0) {
        int i = 0;
        for (int w = 0; w < Hw; ++w) { uint32_t m = sw[w]; while (m) { act[i++] = (uint16_t)((w << 5) + __ffs(m) - 1); m &= m - 1; } }
      }
      __syncwarp();
      const int ld = GEN_CHOL + 1;
      if (lane < na)
        for (int j = 0; j <= lane; ++j) L[lane * ld + j] = a.Psi[(int64_t)act[lane] * H + act[j]];
      __syncwarp();
      for (int k = 0; k < na; ++k) {
        if (lane == k) {
          double d = L[k * ld + k];
          for (int j = 0; j < k; ++j) d -= L[k * ld + j] * L[k * ld + j];
          L[k * ld + k] = sqrt(fmax(d, 0.0));
        }
        __syncwarp();
        if (lane > k && lane < na) {
          double v = L[lane * ld + k];
          for (int j = 0; j < k; ++j) v -= L[lane * ld + j] * L[k * ld + j];
          const double dk = L[k * ld + k];
          L[lane * ld + k] = dk > 0.0 ? v / dk : 0.0;
        }
        __syncwarp();
      }
      double zi = 0.0;
      if (lane < na) {
        zi = a.mus[act[lane]];
        for (int j = 0; j <= lane; ++j) zi += L[lane * ld + j] * z[act[j]];
      }
      __syncwarp();
      if (lane < na) z[act[lane]] = zi;
      __syncwarp();
    }
    if (a.Z)
      for (int h = lane; h < H; h += 32) a.Z[n * H + h] = (float)z[h];
    // ---- y = W z + sigma eps ------------------------------------------------------------------------------------
    for (int d = lane; d < D; d += 32) {
      double acc = 0.0;
      for (int w = 0; w < Hw; ++w) {
        uint32_t m = sw[w];
        while (m) {
          const int h = (w << 5) + __ffs(m) - 1;
          m &= m - 1;
          acc = __dadd_rn(acc, __dmul_rn(__ldg(a.Wt + (int64_t)h * D + d), z[h]));
        }
      }
      double c, s;
      box_muller(rng.draw(RNG_GEN_NOISE, 0, (uint32_t)(d >> 1)), c, s);
      if (a.Ymean) a.Ymean[n * D + d] = (float)acc;
      a.Y[n * D + d] = (float)__dadd_rn(acc, __dmul_rn(a.sigma, (d & 1) ? s : c));
    }
    __syncwarp();
  }
}

// ---- standard_init on the device (_models.py:206-283, sssc.py:104-197) ---------------------------------------------
// column moments of the (masked) data: out[0..D) = sum_n m y, out[D..2D) = sum_n m y^2, out[2D..3D) = sum_n m
__global__ void __launch_bounds__(256) column_moments_kernel(const float *__restrict__ Y, const uint8_t *__restrict__ mask,
                                                             int64_t N, int D, double *__restrict__ out) {
  // thread = column (coalesced rows), CTA = slab of rows
  const int64_t rows_per = (N + gridDim.y - 1) / gridDim.y;
  const int64_t lo = blockIdx.y * rows_per, hi = lo + rows_per < N ? lo + rows_per : N;
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= D) return;
  double s = 0.0, s2 = 0.0, c = 0.0;
  for (int64_t n = lo; n < hi; ++n) {
    const bool ok = !mask || mask[n * D + d];
    if (ok) { const double v = (double)Y[n * D + d]; s += v; s2 += v * v; c += 1.0; }
  }
  atomicAdd(out + d, s);
  atomicAdd(out + D + d, s2);
  atomicAdd(out + 2 * D + d, c);
}
// out[i][j] = (row_mean ? row_mean[i] : 0) + scale * N(0,1)   (rows x cols), Philox keyed by (seed, element index)
__global__ void __launch_bounds__(256) normal_fill_kernel(double *__restrict__ out, int64_t rows, int cols,
                                                          const double *__restrict__ row_mean, double scale, uint64_t seed) {
  const Philox ph(seed);
  const int64_t tot = rows * cols, pairs = (tot + 1) / 2;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < pairs; t += (int64_t)gridDim.x * blockDim.x) {
    double c, s;
    box_muller(ph((uint32_t)t, (uint32_t)(t >> 32), 0x1A17u, 0u), c, s);
    const int64_t i0 = 2 * t, i1 = 2 * t + 1;
    out[i0] = (row_mean ? row_mean[i0 / cols] : 0.0) + scale * c;
    if (i1 < tot) out[i1] = (row_mean ? row_mean[i1 / cols] : 0.0) + scale * s;
  }
}

// ---- FMA throughput: 8 independent accumulator chains per thread, `iters` x 8 FMAs ---------------------------------
template <typename T>
__global__ void __launch_bounds__(256) fma_peak_kernel(T *out, int iters, T x0) {
  T a0 = x0, a1 = x0 + 1, a2 = x0 + 2, a3 = x0 + 3, a4 = x0 + 4, a5 = x0 + 5, a6 = x0 + 6, a7 = x0 + 7;
  const T m = (T)0.9999, c = (T)1e-3;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      a0 = a0 * m + c; a1 = a1 * m + c; a2 = a2 * m + c; a3 = a3 * m + c;
      a4 = a4 * m + c; a5 = a5 * m + c; a6 = a6 * m + c; a7 = a7 * m + c;
    }
  }
  const T r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (r == (T)-1) out[0] = r;  // never true: keeps the chains alive
}

}  // namespace evo

using namespace evo;

extern "C" int evo_generate_data(const double *Wt, const double *pies, double pi, const double *mus, const double *Psi,
                                 int psi_diag, double sigma, int64_t N, int64_t n_offset, int D, int H, uint64_t seed,
                                 float *Y, uint32_t *S_out, float *Z_out, float *Ymean, int32_t *overflow, void *stream) {
  EVO_REQUIRE(N >= 0 && D > 0 && H > 0 && H <= 65535, "bad sizes");
  EVO_REQUIRE(Wt && Y, "Wt / Y null");
  EVO_REQUIRE((Psi == nullptr) == (mus == nullptr), "ES3C needs both mus and Psi; BSC neither");
  EVO_REQUIRE(sigma >= 0.0, "sigma must be non-negative");
  if (N == 0) return EVO_OK;
  GenArgs a;
  a.Wt = Wt; a.pies = pies; a.mus = mus; a.Psi = Psi; a.pi = pi; a.sigma = sigma;
  a.N = N; a.n_offset = n_offset; a.D = D; a.H = H; a.Hw = words_for(H);
  a.psi_mode = Psi ? (psi_diag ? 1 : 2) : 0;
  a.seed = seed; a.Y = Y; a.Z = Z_out; a.Ymean = Ymean; a.S = S_out; a.overflow = overflow;
  const size_t per_warp = (size_t)8 * H + (((size_t)4 * a.Hw + 7) & ~(size_t)7) + (size_t)8 * GEN_CHOL * (GEN_CHOL + 1) + 64;
  const size_t smem = per_warp * GEN_WARPS;
  if (smem > 227 * 1024) {
    set_error("evo_generate_data: H = %d needs %zu bytes of shared memory", H, smem);
    return EVO_ERR_UNSUPPORTED;
  }
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(generate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  EVO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, generate_kernel, 32 * GEN_WARPS, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (int64_t)sm_count() * per_sm;
  const int64_t need = (N + GEN_WARPS - 1) / GEN_WARPS;
  if (g > need) g = need;
  generate_kernel<<<(int)g, 32 * GEN_WARPS, smem, (cudaStream_t)stream>>>(a);
  EVO_LAUNCH_CHECK("generate");
  return EVO_OK;
}

extern "C" int evo_fma_peak(int dtype, int iters, double *tflops_out) {
  EVO_REQUIRE(dtype == 32 || dtype == 64, "dtype must be 32 or 64");
  EVO_REQUIRE(iters > 0 && tflops_out, "bad arguments");
  const int blocks = sm_count() * 8, threads = 256;
  void *buf = nullptr;
  EVO_CUDA(cudaMalloc(&buf, 64));
  cudaEvent_t e0, e1;
  EVO_CUDA(cudaEventCreate(&e0));
  EVO_CUDA(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {  // first repetition warms up
    EVO_CUDA(cudaEventRecord(e0, 0));
    if (dtype == 32) fma_peak_kernel<float><<<blocks, threads>>>((float *)buf, iters, 1.0f);
    else fma_peak_kernel<double><<<blocks, threads>>>((double *)buf, iters, 1.0);
    EVO_CUDA(cudaEventRecord(e1, 0));
    EVO_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    EVO_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  count_launch();
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  const double flop = 2.0 * 64.0 * (double)iters * (double)blocks * threads;
  *tflops_out = flop / (best * 1e-3) / 1e12;
  return EVO_OK;
}

extern "C" int evo_column_moments(const float *Y, const uint8_t *mask, int64_t N, int D, double *out, void *stream) {
  EVO_REQUIRE(Y && out && N >= 0 && D > 0, "bad arguments");
  if (N == 0) return EVO_OK;
  int slabs = (int)((N + 511) / 512);
  const int cap = sm_count() * 8;
  if (slabs > cap) slabs = cap;
  dim3 grid((D + 255) / 256, slabs);
  column_moments_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(Y, mask, N, D, out);
  EVO_LAUNCH_CHECK("column_moments");
  return EVO_OK;
}

extern "C" int evo_normal_fill(double *out, int64_t rows, int cols, const double *row_mean, double scale, uint64_t seed,
                               void *stream) {
  EVO_REQUIRE(out && rows >= 0 && cols > 0, "bad arguments");
  if (rows == 0) return EVO_OK;
  const int64_t pairs = (rows * cols + 1) / 2;
  int64_t b = (pairs + 255) / 256, cap = (int64_t)sm_count() * 8;
  normal_fill_kernel<<<(int)(b < cap ? b : cap), 256, 0, (cudaStream_t)stream>>>(out, rows, cols, row_mean, scale, seed);
  EVO_LAUNCH_CHECK("normal_fill");
  return EVO_OK;
}
