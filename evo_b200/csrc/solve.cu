// Replicated parameter updates of the M-step on the device (no cuSOLVER / cuBLAS, no host fallback):
//   evo_bsc_solve     BSC.M_step parameter solve      evo/models/bsc.py:226-275
//   evo_es3c_update   SSSC.EM_step parameter updates  evo/models/sssc.py:688-770
// Both reduce to  A X = B  with an H x H matrix that is symmetric positive definite (BSC: Wq = sum <s s'>;
// ES3C: sum <s s'> + eps I) or close to it (ES3C: sum <sz sz'>, not symmetric once the reference's element-wise
// Psi update has made Psi non-symmetric).  They are solved by BLOCK GAUSSIAN ELIMINATION on the augmented matrix
// [A | B] with 32 x 32 diagonal blocks: each diagonal block of the running Schur complement is inverted
// explicitly (one warp, Gauss-Jordan with partial pivoting inside the block), which turns the panel scaling,
// the trailing update and the back substitution into plain f64 GEMMs (one strided SIMT kernel).  No pivoting
// ACROSS blocks: stable for (near-)SPD matrices, whose Schur complements stay (near-)SPD.  A pivot below
// 1e-13 x the largest diagonal entry raises `info` (block index + 1): the caller then takes the minimum-norm /
// pseudo-inverse route the reference takes for singular statistics (bsc.py:236-250, sssc.py:695-709).
#include <cooperative_groups.h>
#include <stdlib.h>

#include "common.cuh"

namespace evo {

constexpr int SB = 32;           // elimination block
constexpr int GT = 64, GK = 32;  // GEMM tile (k chunk = the elimination block: one load phase per panel / update tile)

struct GemmArgs {
  const double *A; int64_t ar, ac;  // A(i,k) = A[i*ar + k*ac]
  const double *B; int64_t br, bc;  // B(k,j) = B[k*br + j*bc]
  double *C; int64_t ldc;           // C(i,j) = C[i*ldc + j]
  int M, N, K;
  double alpha, beta;               // C = beta C + alpha A B   (beta == 0: C is not read)
};

// 64 x 64 tile of C per call, 256 threads x (4 x 4), k in steps of 16.  A tile whose C aliases A (N <= 64) or B
// (M <= 64) is safe: every operand read of the tile is done before it is written.
__device__ __forceinline__ void gemm_tile(const GemmArgs &g, int ti, int tj, double (*sa)[GT + 1], double (*sb)[GT + 1]) {
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int i0 = ti * GT, j0 = tj * GT;
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
  for (int k0 = 0; k0 < g.K; k0 += GK) {
    __syncthreads();
    for (int t = threadIdx.x; t < GT * GK; t += 256) {
      int r, k;
      if (g.ac == 1) { k = t % GK; r = t / GK; } else { r = t % GT; k = t / GT; }
      sa[k][r] = (i0 + r < g.M && k0 + k < g.K) ? g.A[(int64_t)(i0 + r) * g.ar + (int64_t)(k0 + k) * g.ac] : 0.0;
    }
    for (int t = threadIdx.x; t < GT * GK; t += 256) {
      int c, k;
      if (g.bc == 1) { c = t % GT; k = t / GT; } else { k = t % GK; c = t / GK; }
      sb[k][c] = (j0 + c < g.N && k0 + k < g.K) ? g.B[(int64_t)(k0 + k) * g.br + (int64_t)(j0 + c) * g.bc] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GK; ++k) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sa[k][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sb[k][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
    }
  }
  __syncthreads();  // in-place products: all operand reads of this tile are done
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = i0 + ty + 16 * i, c = j0 + tx + 16 * j;
      if (r < g.M && c < g.N) {
        double *p = g.C + (int64_t)r * g.ldc + c;
        *p = (g.beta == 0.0 ? 0.0 : g.beta * *p) + g.alpha * acc[i][j];
      }
    }
}
__global__ void __launch_bounds__(256) gemm_f64_kernel(GemmArgs g) {
  __shared__ double sa[GK][GT + 1], sb[GK][GT + 1];
  gemm_tile(g, blockIdx.y, blockIdx.x, sa, sb);
}

static int gemm(const GemmArgs &g, cudaStream_t st) {
  if (g.M <= 0 || g.N <= 0) return EVO_OK;
  dim3 grid((g.N + GT - 1) / GT, (g.M + GT - 1) / GT);
  gemm_f64_kernel<<<grid, 256, 0, st>>>(g);
  EVO_LAUNCH_CHECK("gemm_f64");
  return EVO_OK;
}

// Inverse of the 32 x 32 diagonal block M[r0.., r0..] (leading dimension ld) by Gauss-Jordan with partial
// pivoting, one warp, lane = row, the row of [A | R] in REGISTERS (both loops fully unrolled).  Rows are not swapped:
// the lane chosen as pivot for column k ends up holding row k of the inverse.  inv (32 x 32, row-major) <- block^-1;
// *info = blk + 1 (first failing block) when a pivot is <= tol * scale[0] in magnitude or not finite.
__device__ __forceinline__ void diag_inv_warp(const double *__restrict__ Mx, int64_t ld, int r0, double *inv, double *prow,
                                              double thr, bool &bad) {
  const int lane = threadIdx.x & 31;
  double a[SB], r[SB];
#pragma unroll
  for (int c = 0; c < SB; ++c) {
    a[c] = Mx[(int64_t)(r0 + lane) * ld + r0 + c];
    r[c] = (lane == c) ? 1.0 : 0.0;
  }
  bool used = false;
  int mycol = 0;
  double mypiv = 1.0;
#pragma unroll
  for (int k = 0; k < SB; ++k) {
    double mag = used ? -1.0 : fabs(a[k]);
    int p = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double om = __shfl_xor_sync(0xffffffffu, mag, o);
      const int op = __shfl_xor_sync(0xffffffffu, p, o);
      if (om > mag || (om == mag && op < p)) { mag = om; p = op; }
    }
    if (lane == p) {
#pragma unroll
      for (int c = 0; c < SB; ++c) { prow[c] = a[c]; prow[SB + c] = r[c]; }
      used = true; mycol = k; mypiv = a[k];
    }
    __syncwarp();
    const double piv = prow[k];
    if (!(fabs(piv) > thr) || !isfinite(piv)) bad = true;
    if (lane != p) {
      const double f = a[k] / piv;
#pragma unroll
      for (int c = k + 1; c < SB; ++c) a[c] -= f * prow[c];
#pragma unroll
      for (int c = 0; c < SB; ++c) r[c] -= f * prow[SB + c];
    }
    __syncwarp();
  }
  const double d = 1.0 / mypiv;
#pragma unroll
  for (int c = 0; c < SB; ++c) inv[mycol * SB + c] = r[c] * d;
}
__global__ void __launch_bounds__(32) diag_inv_kernel(const double *__restrict__ Mx, int64_t ld, int r0, double *__restrict__ inv,
                                                      const double *__restrict__ scale, double tol, int blk, int *info) {
  __shared__ double prow[2 * SB];
  bool bad = false;
  diag_inv_warp(Mx, ld, r0, inv, prow, tol * fabs(scale[0]), bad);
  if (bad && threadIdx.x == 0) atomicCAS(info, 0, blk + 1);
}

// The whole elimination as ONE persistent cooperative kernel: grid barriers (~ 2 us) instead of ~ 5 dependent
// launches per block step.  Every CTA inverts the diagonal block itself (redundantly, into shared memory), so the
// panel scaling needs no barrier after the inversion; CTA 0 publishes the inverse for the back substitution.
__global__ void __launch_bounds__(256, 1) block_solve_kernel(double *Mx, int Hp, int Wd, double *invs, const double *scale,
                                                            double tol, int *info) {
  __shared__ double sa[GK][GT + 1], sb[GK][GT + 1];
  __shared__ double prow[2 * SB];
  __shared__ double sinv[SB * SB];
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const int nb = Hp / SB, nrhs = Wd - Hp, G = gridDim.x, me = blockIdx.x;
  const int64_t ld = Wd;
  const double thr = tol * fabs(scale[0]);
  for (int k = 0; k < nb; ++k) {
    const int r0 = k * SB;
    __syncthreads();
    if (k == 0) {  // first block: everybody inverts it (nothing to overlap with)
      if (threadIdx.x < 32) {
        bool bad = false;
        diag_inv_warp(Mx, ld, r0, sinv, prow, thr, bad);
        if (bad && me == 0 && threadIdx.x == 0) atomicCAS(info, 0, k + 1);
      }
      __syncthreads();
      if (me == 0)
        for (int t = threadIdx.x; t < SB * SB; t += 256) invs[t] = sinv[t];
    } else {  // look-ahead: CTA 0 inverted this block during the previous trailing update and published it
      for (int t = threadIdx.x; t < SB * SB; t += 256) sinv[t] = invs[(size_t)k * SB * SB + t];
      __syncthreads();
    }
    const int below = Hp - r0 - SB;
    if (below > 0) {
      // panel: L_ik = M_ik inv (in place); each CTA scales the row tiles it will also use... (any split works)
      GemmArgs p{Mx + (int64_t)(r0 + SB) * ld + r0, ld, 1, sinv, SB, 1, Mx + (int64_t)(r0 + SB) * ld + r0, ld, below, SB, SB, 1.0, 0.0};
      const int pt = (below + GT - 1) / GT;
      for (int t = me; t < pt; t += G) gemm_tile(p, t, 0, sa, sb);
    }
    grid.sync();
    if (below > 0) {
      GemmArgs u{Mx + (int64_t)(r0 + SB) * ld + r0, ld, 1, Mx + (int64_t)r0 * ld + r0 + SB, ld, 1,
                 Mx + (int64_t)(r0 + SB) * ld + r0 + SB, ld, below, Wd - r0 - SB, SB, -1.0, 1.0};
      const int ti = (u.M + GT - 1) / GT, tj = (u.N + GT - 1) / GT;
      if (me == 0) {
        // the tile holding the next diagonal block first, then its inverse -- while the other CTAs finish the update
        gemm_tile(u, 0, 0, sa, sb);
        __syncthreads();
        if (threadIdx.x < 32) {
          bool bad = false;
          diag_inv_warp(Mx, ld, r0 + SB, invs + (size_t)(k + 1) * SB * SB, prow, thr, bad);
          if (bad && threadIdx.x == 0) atomicCAS(info, 0, k + 2);
        }
        __threadfence();
        if (G == 1)
          for (int t = 1; t < ti * tj; ++t) gemm_tile(u, t / tj, t % tj, sa, sb);
      } else {  // the other CTAs share the remaining tiles: CTA 0 is busy with the inversion (the critical path)
        for (int t = me; t < ti * tj; t += G - 1) gemm_tile(u, t / tj, t % tj, sa, sb);
      }
      grid.sync();
    }
  }
  __threadfence();
  grid.sync();  // the last block's inverse (published by CTA 0 without a barrier behind it) is visible to everyone
  for (int k = nb - 1; k >= 0; --k) {
    const int r0 = k * SB;
    double *Rk = Mx + (int64_t)r0 * ld + Hp;
    GemmArgs x{invs + (size_t)k * SB * SB, SB, 1, Rk, ld, 1, Rk, ld, SB, nrhs, SB, 1.0, 0.0};
    const int xt = (nrhs + GT - 1) / GT;
    for (int t = me; t < xt; t += G) gemm_tile(x, 0, t, sa, sb);
    grid.sync();
    if (r0 > 0) {
      GemmArgs u{Mx + r0, ld, 1, Rk, ld, 1, Mx + Hp, ld, r0, nrhs, SB, -1.0, 1.0};
      const int ti = (u.M + GT - 1) / GT, tj = (u.N + GT - 1) / GT;
      for (int t = me; t < ti * tj; t += G) gemm_tile(u, t / tj, t % tj, sa, sb);
      grid.sync();
    }
  }
}

// Block Gaussian elimination of [A | B] in place (Mx: Hp x Wd row-major, Hp a multiple of 32, A = the first Hp
// columns).  On return the right part holds X = A^-1 B.  invs: Hp/32 blocks of 32 x 32.
static int block_solve(double *Mx, int Hp, int Wd, double *invs, const double *scale, int *info, cudaStream_t st) {
  const int nb = Hp / SB;
  const int64_t ld = Wd;
  if (!getenv("EVO_SOLVE_MULTI")) {  // one persistent cooperative kernel when the device supports it
    int dev = 0, coop = 0, per_sm = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess &&
        coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, block_solve_kernel, 256, 0) == cudaSuccess && per_sm >= 1) {
      const int tiles = ((Hp + GT - 1) / GT) * ((Wd + GT - 1) / GT);
      int grid = sm_count() * per_sm;
      if (grid > tiles) grid = tiles;
      if (grid < 1) grid = 1;
      double tol = 1e-13;
      void *args[] = {&Mx, &Hp, &Wd, &invs, (void *)&scale, &tol, &info};
      EVO_CUDA(cudaLaunchCooperativeKernel((const void *)block_solve_kernel, dim3(grid), dim3(256), args, 0, st));
      count_launch();
      return EVO_OK;
    }
    cudaGetLastError();
  }
  for (int k = 0; k < nb; ++k) {
    const int r0 = k * SB;
    double *inv = invs + (size_t)k * SB * SB;
    diag_inv_kernel<<<1, 32, 0, st>>>(Mx, ld, r0, inv, scale, 1e-13, k, info);
    EVO_LAUNCH_CHECK("diag_inv");
    const int below = Hp - r0 - SB;
    if (below <= 0) break;
    // panel: L_ik = M_ik inv  (in place)
    GemmArgs p{Mx + (int64_t)(r0 + SB) * ld + r0, ld, 1, inv, SB, 1, Mx + (int64_t)(r0 + SB) * ld + r0, ld, below, SB, SB, 1.0, 0.0};
    int rc = gemm(p, st);
    if (rc) return rc;
    // trailing update (matrix and right-hand sides at once): M_ij -= L_ik M_kj, j > k
    GemmArgs t{Mx + (int64_t)(r0 + SB) * ld + r0, ld, 1, Mx + (int64_t)r0 * ld + r0 + SB, ld, 1,
               Mx + (int64_t)(r0 + SB) * ld + r0 + SB, ld, below, Wd - r0 - SB, SB, -1.0, 1.0};
    rc = gemm(t, st);
    if (rc) return rc;
  }
  // back substitution, right-looking: X_k = inv_k R_k, then R_i -= M_ik X_k for every block row i above (wide, short
  // GEMMs instead of 32-row dot products with a long k loop)
  const int nrhs = Wd - Hp;
  for (int k = nb - 1; k >= 0; --k) {
    const int r0 = k * SB;
    double *Rk = Mx + (int64_t)r0 * ld + Hp;
    GemmArgs x{invs + (size_t)k * SB * SB, SB, 1, Rk, ld, 1, Rk, ld, SB, nrhs, SB, 1.0, 0.0};
    int rc = gemm(x, st);
    if (rc) return rc;
    if (r0 > 0) {
      GemmArgs u{Mx + r0, ld, 1, Rk, ld, 1, Mx + Hp, ld, r0, nrhs, SB, -1.0, 1.0};
      rc = gemm(u, st);
      if (rc) return rc;
    }
  }
  return EVO_OK;
}

// ---- loaders ----------------------------------------------------------------------------------------------
// Mx[i][j] <- A (H x H): mode 0 = lower triangle mirrored (BSC Wq), 1 = as is, 2 = transposed; + add_diag on the
// diagonal; identity on the padding.  scale[0] <- max |diagonal| (non-negative doubles order like their bit patterns).
__global__ void load_matrix_kernel(const double *__restrict__ A, int H, int Hp, int mode, double add_diag, double *__restrict__ Mx,
                                   int64_t ld, unsigned long long *scale_bits) {
  const int64_t tot = (int64_t)Hp * Hp;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / Hp), j = (int)(t % Hp);
    double v;
    if (i < H && j < H) {
      if (mode == 0) v = (j <= i) ? A[(int64_t)i * H + j] : A[(int64_t)j * H + i];
      else if (mode == 1) v = A[(int64_t)i * H + j];
      else v = A[(int64_t)j * H + i];
      if (i == j) {
        v += add_diag;
        atomicMax(scale_bits, (unsigned long long)__double_as_longlong(fabs(v)));
      }
    } else {
      v = (i == j) ? 1.0 : 0.0;
    }
    Mx[(int64_t)i * ld + j] = v;
  }
}
// right-hand sides: Mx[i][Hp + j] <- B (H x nrhs, leading dimension ldb), or the identity when B == NULL; zero padding
__global__ void load_rhs_kernel(const double *__restrict__ B, int64_t ldb, int H, int Hp, int nrhs, double *__restrict__ Mx, int64_t ld) {
  const int64_t tot = (int64_t)Hp * nrhs;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / nrhs), j = (int)(t % nrhs);
    double v = 0.0;
    if (i < H) v = B ? B[(int64_t)i * ldb + j] : (i == j ? 1.0 : 0.0);
    Mx[(int64_t)i * ld + Hp + j] = v;
  }
}

static int ew_grid(int64_t n) {
  int64_t b = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

// ---- BSC finish: W = X' (bsc.py:251), pies / pi (bsc.py:254-262), sigma (bsc.py:264-275) -----------------------------
// scal: [0] info (block + 1 of the first failed pivot, 0 = ok)  [1] pi  [2] sigma  [3] 1 if every entry of X is finite
__global__ void bsc_finish_kernel(const double *__restrict__ Mx, int64_t ld, int Hp, int D, int H, const double *__restrict__ stats,
                                  double N_total, int learn_W, int background, double n_obs_total, double sigma_old,
                                  double *__restrict__ W, double *__restrict__ pies, const int *__restrict__ info,
                                  double *__restrict__ scal, int *finite_flag) {
  const int64_t tot = (int64_t)D * H;
  bool ok = true;
  if (learn_W) {
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
      const int d = (int)(t / H), h = (int)(t % H);
      const double v = Mx[(int64_t)h * ld + Hp + d];
      ok &= isfinite(v);
      if (info[0] == 0) W[t] = v;  // a failed factorisation leaves W to the caller's fallback
    }
  }
  if (!ok) atomicExch(finite_flag, 0);
  if (blockIdx.x == 0) {
    __shared__ double red[8];
    const double *p_in = stats + (int64_t)H * D + (int64_t)H * H;
    double s = 0.0;
    for (int h = threadIdx.x; h < H; h += blockDim.x) {
      double v = p_in[h] / N_total;
      if (background && h == H - 1) v = 1.0 - 1.1e-5;
      pies[h] = v;
      s += v;
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double tot_p = 0.0;
      for (int i = 0; i < (int)(blockDim.x >> 5); ++i) tot_p += red[i];
      const double sig = p_in[H];
      scal[0] = (double)info[0];
      scal[1] = tot_p / H;
      scal[2] = n_obs_total >= 0.0 ? sqrt((sig + n_obs_total * sigma_old * sigma_old) / N_total / D) : sqrt(sig / N_total / D);
    }
  }
}
__global__ void set_int_kernel(int *p, int v) { *p = v; }
__global__ void flag_to_scal_kernel(const int *flag, double *scal, int slot) { scal[slot] = (double)flag[0]; }

struct SolveWs {
  double *Mx, *invs, *scale, *aux, *aux2;
  int *info, *finite;
  int Hp, Wd;
};
static int64_t ws_doubles(int H, int nrhs, int64_t extra) {
  const int64_t Hp = (H + SB - 1) / SB * SB;
  return Hp * (Hp + nrhs) + Hp * SB + 8 + extra;
}
static SolveWs carve_ws(void *workspace, int H, int nrhs) {
  SolveWs w;
  w.Hp = (H + SB - 1) / SB * SB;
  w.Wd = w.Hp + nrhs;
  double *p = reinterpret_cast<double *>(workspace);
  w.scale = p;                                   // [1] (+ padding)
  w.info = reinterpret_cast<int *>(p + 1);       // [2 ints]
  w.finite = w.info + 1;
  p += 8;
  w.Mx = p; p += (int64_t)w.Hp * w.Wd;
  w.invs = p; p += (int64_t)w.Hp * SB;
  w.aux = p;
  w.aux2 = nullptr;
  return w;
}

// ---- ES3C element-wise updates (sssc.py:712-741) -------------------------------------------------------------------------
__global__ void es3c_vec_kernel(const double *__restrict__ sum_s, const double *__restrict__ sum_sz, int H, double N_total,
                                int learn_pies, int learn_mus, int background, double eps_pies, double eps_mus,
                                double *__restrict__ pies, double *__restrict__ mus) {
  for (int h = blockIdx.x * blockDim.x + threadIdx.x; h < H; h += gridDim.x * blockDim.x) {
    if (learn_pies) {
      double p = sum_s[h] / N_total;
      if (p <= eps_pies) p = eps_pies;
      if (p >= 1.0 - eps_pies) p = 1.0 - eps_pies;
      if (background && h == H - 1) p = 1.0 - 1.1e-5;
      pies[h] = p;
    }
    if (learn_mus) mus[h] = sum_sz[h] / (sum_s[h] + eps_mus);
  }
}
// Psi_new = (mu mu' o sum_ss + sum_szsz - 2 mu[:,None] o sum_s_sz) o inv(sum_ss + eps I): ELEMENT-WISE, as the reference
// computes it (sssc.py:732-739).  inv = right part of the eliminated workspace.
__global__ void es3c_psi_kernel(const double *__restrict__ sum_ss, const double *__restrict__ sum_szsz, const double *__restrict__ sum_s_sz,
                                const double *__restrict__ mus, const double *__restrict__ Mx, int64_t ld, int Hp, int H,
                                double *__restrict__ Psi, int *finite_flag) {
  const int64_t tot = (int64_t)H * H;
  bool ok = true;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / H), j = (int)(t % H);
    const double psi = mus[i] * mus[j] * sum_ss[t] + sum_szsz[t] - 2.0 * mus[i] * sum_s_sz[t];
    const double v = psi * Mx[(int64_t)i * ld + Hp + j];
    ok &= isfinite(v);
    Psi[t] = v;
  }
  if (!ok) atomicExch(finite_flag, 0);
}
// W[d][h] <- X[h][d] (X = right part of the workspace); finite check
__global__ void transpose_out_kernel(const double *__restrict__ Mx, int64_t ld, int Hp, int D, int H, double *__restrict__ W,
                                     const int *__restrict__ info, int *finite_flag) {
  const int64_t tot = (int64_t)D * H;
  bool ok = true;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
    const int d = (int)(t / H), h = (int)(t % H);
    const double v = Mx[(int64_t)h * ld + Hp + d];
    ok &= isfinite(v);
    if (info[0] == 0) W[t] = v;
  }
  if (!ok) atomicExch(finite_flag, 0);
}
// out[0] += sum_ij P[i][j] Q[j][i]   ( = trace(P Q) )
__global__ void trace_prod_kernel(const double *__restrict__ P, const double *__restrict__ Q, int H, double *out) {
  const int64_t tot = (int64_t)H * H;
  double s = 0.0;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(t / H), j = (int)(t % H);
    s += P[t] * Q[(int64_t)j * H + i];
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0 && s != 0.0) atomicAdd(out, s);
}
// scal: [0] info of the W solve  [1] info of the Psi solve  [2] sigma2_new  [3] all outputs finite
__global__ void es3c_scal_kernel(const int *infoW, const int *infoP, const int *finite, const double *tr, int incomplete, double ysq,
                                 const double *trace_stat, double n_obs_total, double sigma2_old, double N_total, int D,
                                 double eps_sigma2, int learn_sigma2, double *scal) {
  scal[0] = (double)infoW[0];
  scal[1] = (double)infoP[0];
  double s2 = sigma2_old;
  if (learn_sigma2) {
    if (incomplete) s2 = ((ysq - trace_stat[0]) + n_obs_total * sigma2_old) / N_total / D + eps_sigma2;
    else s2 = (ysq - tr[0]) / N_total / D + eps_sigma2;
  }
  scal[2] = s2;
  scal[3] = (double)finite[0];
}

}  // namespace evo

using namespace evo;

extern "C" int64_t evo_bsc_solve_workspace_bytes(int D, int H) {
  if (D <= 0 || H <= 0) return 0;
  return 8 * ws_doubles(H, D, 0);
}

extern "C" int evo_bsc_solve(const double *stats, double N_total, int D, int H, int learn_W, int background,
                             double n_obs_total, double sigma_old, double *W, double *pies, double *scal, void *workspace,
                             void *stream) {
  EVO_REQUIRE(stats && pies && scal && workspace, "null buffer");
  EVO_REQUIRE(D > 0 && H > 0 && N_total > 0.0, "bad sizes");
  EVO_REQUIRE(!learn_W || W, "W null");
  EVO_REQUIRE(((uintptr_t)workspace & 15) == 0, "workspace must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SolveWs w = carve_ws(workspace, H, D);
  EVO_CUDA(cudaMemsetAsync(workspace, 0, 64, st));  // scale, info
  set_int_kernel<<<1, 1, 0, st>>>(w.finite, 1);  // a kernel, not a copy from the host stack: capturable in a CUDA graph
  EVO_LAUNCH_CHECK("set_int");
  if (learn_W) {
    const double *Wp = stats, *Wq = stats + (int64_t)H * D;
    load_matrix_kernel<<<ew_grid((int64_t)w.Hp * w.Hp), 256, 0, st>>>(Wq, H, w.Hp, 0, 0.0, w.Mx, w.Wd,
                                                                     reinterpret_cast<unsigned long long *>(w.scale));
    EVO_LAUNCH_CHECK("load_matrix");
    load_rhs_kernel<<<ew_grid((int64_t)w.Hp * D), 256, 0, st>>>(Wp, D, H, w.Hp, D, w.Mx, w.Wd);
    EVO_LAUNCH_CHECK("load_rhs");
    int rc = block_solve(w.Mx, w.Hp, w.Wd, w.invs, w.scale, w.info, st);
    if (rc) return rc;
  }
  bsc_finish_kernel<<<ew_grid((int64_t)D * H), 256, 0, st>>>(w.Mx, w.Wd, w.Hp, D, H, stats, N_total, learn_W, background,
                                                             n_obs_total, sigma_old, W, pies, w.info, scal, w.finite);
  EVO_LAUNCH_CHECK("bsc_finish");
  flag_to_scal_kernel<<<1, 1, 0, st>>>(w.finite, scal, 3);
  EVO_LAUNCH_CHECK("flag_to_scal");
  return EVO_OK;
}

extern "C" int64_t evo_es3c_update_workspace_bytes(int D, int H) {
  if (D <= 0 || H <= 0) return 0;
  const int nrhs = D > H ? D : H;
  return 8 * ws_doubles(H, nrhs, (int64_t)H * H + 16);
}

extern "C" int evo_es3c_update(const evo_es3c_update_args *x, void *stream) {
  EVO_REQUIRE(x && x->stats && x->W && x->pies && x->mus && x->Psi && x->scal && x->workspace, "null buffer");
  EVO_REQUIRE(x->D > 0 && x->H > 0 && x->N_total > 0.0, "bad sizes");
  EVO_REQUIRE(((uintptr_t)x->workspace & 15) == 0, "workspace must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int H = x->H, D = x->D;
  const int64_t H2 = (int64_t)H * H;
  const double *sum_s = x->stats, *sum_sz = sum_s + H, *sum_ss = sum_sz + H, *sum_szsz = sum_ss + H2;
  const double *sum_s_sz = sum_szsz + H2, *sum_outer = sum_s_sz + H2, *WpT = sum_outer + H2, *trace = WpT + (int64_t)H * D;
  const int nrhs = D > H ? D : H;
  SolveWs w = carve_ws(x->workspace, H, nrhs);
  double *WtW = w.aux;             // [H][H]
  double *tr = w.aux + H2;         // [1]
  int *infoP = reinterpret_cast<int *>(tr + 1);  // second info word (Psi solve)
  EVO_CUDA(cudaMemsetAsync(x->workspace, 0, 64, st));
  EVO_CUDA(cudaMemsetAsync(tr, 0, 16, st));
  set_int_kernel<<<1, 1, 0, st>>>(w.finite, 1);  // a kernel, not a copy from the host stack: capturable in a CUDA graph
  EVO_LAUNCH_CHECK("set_int");
  // ---- W = Wp inv(sum_szsz)  <=>  sum_szsz' W' = WpT  (sssc.py:692-694) --------------------------------------------------
  if (x->learn_W) {
    SolveWs ww = w;
    ww.Wd = w.Hp + D;
    load_matrix_kernel<<<ew_grid((int64_t)w.Hp * w.Hp), 256, 0, st>>>(sum_szsz, H, w.Hp, 2, 0.0, w.Mx, ww.Wd,
                                                                     reinterpret_cast<unsigned long long *>(w.scale));
    EVO_LAUNCH_CHECK("load_matrix");
    load_rhs_kernel<<<ew_grid((int64_t)w.Hp * D), 256, 0, st>>>(WpT, D, H, w.Hp, D, w.Mx, ww.Wd);
    EVO_LAUNCH_CHECK("load_rhs");
    int rc = block_solve(w.Mx, w.Hp, ww.Wd, w.invs, w.scale, w.info, st);
    if (rc) return rc;
    transpose_out_kernel<<<ew_grid((int64_t)D * H), 256, 0, st>>>(w.Mx, ww.Wd, w.Hp, D, H, x->W, w.info, w.finite);
    EVO_LAUNCH_CHECK("transpose_out");
  }
  // ---- pies, mus (sssc.py:712-727) -----------------------------------------------------------------------------------------------
  es3c_vec_kernel<<<ew_grid(H), 256, 0, st>>>(sum_s, sum_sz, H, x->N_total, x->learn_pies, x->learn_mus, x->background,
                                             x->eps_pies, x->eps_mus, x->pies, x->mus);
  EVO_LAUNCH_CHECK("es3c_vec");
  // ---- sigma2, complete data: trace(sum <sz><sz>' W'W) with the NEW W (sssc.py:759-768) -- before Psi reuses the workspace ------
  if (x->learn_sigma2 && !x->incomplete) {
    GemmArgs g{x->W, 1, H, x->W, H, 1, WtW, H, H, H, D, 1.0, 0.0};  // W'W: A(i,k) = W[k][i]
    int rc = gemm(g, st);
    if (rc) return rc;
    trace_prod_kernel<<<ew_grid(H2), 256, 0, st>>>(sum_outer, WtW, H, tr);
    EVO_LAUNCH_CHECK("trace_prod");
  }
  // ---- Psi (sssc.py:729-741) ------------------------------------------------------------------------------------------------------
  if (x->learn_Psi) {
    SolveWs wp = w;
    wp.Wd = w.Hp + H;
    EVO_CUDA(cudaMemsetAsync(w.scale, 0, 8, st));
    load_matrix_kernel<<<ew_grid((int64_t)w.Hp * w.Hp), 256, 0, st>>>(sum_ss, H, w.Hp, 1, x->eps_Psi, w.Mx, wp.Wd,
                                                                     reinterpret_cast<unsigned long long *>(w.scale));
    EVO_LAUNCH_CHECK("load_matrix");
    load_rhs_kernel<<<ew_grid((int64_t)w.Hp * H), 256, 0, st>>>(nullptr, 0, H, w.Hp, H, w.Mx, wp.Wd);
    EVO_LAUNCH_CHECK("load_rhs");
    int rc = block_solve(w.Mx, w.Hp, wp.Wd, w.invs, w.scale, infoP, st);
    if (rc) return rc;
    es3c_psi_kernel<<<ew_grid(H2), 256, 0, st>>>(sum_ss, sum_szsz, sum_s_sz, x->mus, w.Mx, wp.Wd, w.Hp, H, x->Psi, w.finite);
    EVO_LAUNCH_CHECK("es3c_psi");
  }
  es3c_scal_kernel<<<1, 1, 0, st>>>(w.info, infoP, w.finite, tr, x->incomplete, x->ysq, trace, x->n_obs_total, x->sigma2_old,
                                    x->N_total, D, x->eps_sigma2, x->learn_sigma2, x->scal);
  EVO_LAUNCH_CHECK("es3c_scal");
  return EVO_OK;
}

// ---- evo_bsc_em_advance: the host's share of an EM iteration (check_params + E_step_precompute + logging) on the device --
// One thread.  rec row: [0..8) acc, [8..12) scal, [12] ljc, [13] pi, [14] sigma, [15] iteration -- all of the iteration
// that just ran; then theta <- the clamped new parameters and their precomputed scalars, iteration + 1.
__global__ void bsc_em_advance_kernel(const double *__restrict__ acc, const double *__restrict__ scal, double *theta, double *rec,
                                      int max_rec, int D, int H, int learn_pi, int learn_sigma, double pi_lo, double pi_hi,
                                      double sigma_lo, double sigma_hi) {
  const int r = (int)theta[7];
  if (rec && r < max_rec) {
    double *row = rec + (int64_t)r * 16;
    for (int i = 0; i < 8; ++i) row[i] = acc[i];
    for (int i = 0; i < 4; ++i) row[8 + i] = scal[i];
    row[12] = theta[6]; row[13] = theta[4]; row[14] = theta[5]; row[15] = theta[3];
  }
  theta[7] = (double)(r + 1);
  double pi = theta[4], sigma = theta[5];
  if (learn_pi) pi = scal[1];
  if (learn_sigma) sigma = scal[2] == scal[2] ? scal[2] : 0.0;  // NaN root of a negative statistic: clamped below
  if (pi < pi_lo) pi = pi_lo;                                   // check_params (_models.py:102-159)
  if (pi >= pi_hi) pi = pi_hi;
  if (sigma < sigma_lo) sigma = sigma_lo;
  if (sigma >= sigma_hi) sigma = sigma_hi;
  theta[4] = pi; theta[5] = sigma;
  theta[0] = -1.0 / 2.0 / sigma / sigma;                         // E_step_precompute (bsc.py:100-125)
  theta[1] = log(pi / (1.0 - pi));
  theta[2] = pi * (double)H;
  theta[6] = (double)H * log(1.0 - pi) - (double)D / 2.0 * log(2.0 * 3.14159265358979323846 * sigma * sigma);
  theta[3] += 1.0;
}

extern "C" int evo_bsc_em_advance(const double *acc, const double *scal, double *theta_dev, double *rec, int max_rec, int D,
                                  int H, int learn_pi, int learn_sigma, double pi_lo, double pi_hi, double sigma_lo,
                                  double sigma_hi, void *stream) {
  EVO_REQUIRE(acc && scal && theta_dev && D > 0 && H > 0, "bad arguments");
  bsc_em_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(acc, scal, theta_dev, rec, max_rec, D, H, learn_pi, learn_sigma, pi_lo,
                                                           pi_hi, sigma_lo, sigma_hi);
  EVO_LAUNCH_CHECK("bsc_em_advance");
  return EVO_OK;
}

// ---- evo_bsc_solve_static: the BSC solve for callers that may not allocate -----------------------------------------------------
// No counterpart in the reference; added (CUDA-graph capture, serving loops and
// harnesses that forbid cudaMalloc).  Same computation as evo_bsc_solve, with the elimination workspace in a file-scope
// device array instead of a caller buffer.  Every GPU that loads the library holds the array until the process exits.
// All calls on a device SHARE it: calls that overlap on the GPU (two streams) race; serialise them.  It keeps the last
// call's contents; the solve clears its 64-byte header (scale, info, finite) itself and overwrites [A | B] and the
// block inverses in full before reading them, so nothing else needs to start at zero.
static __device__ __align__(256) unsigned char g_solve_static[EVO_BSC_SOLVE_STATIC_BYTES];

namespace {
constexpr int kMaxDev = 64;
void *g_static_ptr[kMaxDev];  // device address of g_solve_static, per device (host code must not pass the symbol itself)
bool g_static_set[kMaxDev];
}  // namespace

extern "C" int evo_bsc_solve_static(const double *stats, double N_total, int D, int H, int learn_W, int background,
                                    double n_obs_total, double sigma_old, double *W, double *pies, double *scal, void *stream) {
  EVO_REQUIRE(D > 0 && H > 0, "bad sizes");
  const int64_t need = evo_bsc_solve_workspace_bytes(D, H);
  if (need > (int64_t)EVO_BSC_SOLVE_STATIC_BYTES) {
    set_error("evo_bsc_solve_static: D=%d H=%d needs %lld workspace bytes, the static scratch holds %lld "
              "(EVO_BSC_SOLVE_STATIC_BYTES); use evo_bsc_solve with a caller workspace",
              D, H, (long long)need, (long long)EVO_BSC_SOLVE_STATIC_BYTES);
    return EVO_ERR_ARG;
  }
  int dev = -1;
  EVO_CUDA(cudaGetDevice(&dev));
  EVO_REQUIRE(dev >= 0 && dev < kMaxDev, "device ordinal out of range");
  if (!g_static_set[dev]) {  // once per device, with that device current
    void *p = nullptr;
    EVO_CUDA(cudaGetSymbolAddress(&p, g_solve_static));
    EVO_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    g_static_ptr[dev] = p;
    g_static_set[dev] = true;
  }
  return evo_bsc_solve(stats, N_total, D, H, learn_W, background, n_obs_total, sigma_old, W, pies, scal, g_static_ptr[dev],
                       stream);
}
