// Binary Sparse Coding kernels: Gram tables, stand-alone log-joint, M-step sufficient statistics,
// data estimator.  Reference: evo/models/bsc.py, evo/models/_models.py.
#include "common.cuh"
#include "evaluators.cuh"
#include "stats.cuh"

namespace evo {

// ------------------------------------------------------------------------------------------
// G = W'W  (H,H) f64.  W is (D,H) row-major.  32x32 output tile per CTA, 16x16 threads x 2x2.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gram_kernel(const double *__restrict__ W, double *__restrict__ G, int D, int H) {
  __shared__ double sa[16][33], sb[16][33];
  int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  if (c0 > r0) return;  // lower triangle of tiles, mirrored on store
  double acc[2][2] = {{0, 0}, {0, 0}};
  for (int d0 = 0; d0 < D; d0 += 16) {
    for (int i = threadIdx.x; i < 16 * 32; i += 256) {
      int dd = i >> 5, c = i & 31;
      int d = d0 + dd;
      sa[dd][c] = (d < D && r0 + c < H) ? W[(int64_t)d * H + r0 + c] : 0.0;
      sb[dd][c] = (d < D && c0 + c < H) ? W[(int64_t)d * H + c0 + c] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int dd = 0; dd < 16; ++dd) {
      double a0 = sa[dd][ty], a1 = sa[dd][ty + 16], b0 = sb[dd][tx], b1 = sb[dd][tx + 16];
      acc[0][0] += a0 * b0; acc[0][1] += a0 * b1; acc[1][0] += a1 * b0; acc[1][1] += a1 * b1;
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      int r = r0 + ty + 16 * i, c = c0 + tx + 16 * j;
      if (r < H && c < H) {
        G[(int64_t)r * H + c] = acc[i][j];
        G[(int64_t)c * H + r] = acc[i][j];
      }
    }
}

// ynorm2[n] = |y_n|^2 in f64, one warp per row
__global__ void ynorm_kernel(const float *__restrict__ Y, double *__restrict__ out, int64_t N, int D) {
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t n = warp; n < N; n += nwarps) {
    double s = 0.0;
    for (int d = lane; d < D; d += 32) { double v = (double)Y[n * D + d]; s += v * v; }
    s = warp_sum(s);
    if (lane == 0) out[n] = s;
  }
}

// B = Y W : (N,D) f32 x (D,H) f64 -> (N,H) f64.  64x64 tile per CTA, 256 threads x 4x4, f64 FMA.
__global__ void __launch_bounds__(256) wty_kernel(const float *__restrict__ Y, const double *__restrict__ W,
                                                  double *__restrict__ B, int64_t N, int D, int H) {
  __shared__ double sy[16][65];   // [k][row]
  __shared__ double sw[16][65];   // [k][col]
  int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  int64_t ntr = (N + 63) / 64;
  int ntc = (H + 63) / 64;
  for (int64_t tile = blockIdx.x; tile < ntr * ntc; tile += gridDim.x) {
    int64_t r0 = (tile / ntc) * 64;
    int c0 = (int)(tile % ntc) * 64;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
    for (int d0 = 0; d0 < D; d0 += 16) {
      __syncthreads();
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        int rr = i >> 4, kk = i & 15;  // Y row-major: consecutive threads along d
        int64_t r = r0 + rr;
        int d = d0 + kk;
        sy[kk][rr] = (r < N && d < D) ? (double)Y[r * D + d] : 0.0;
      }
      for (int i = threadIdx.x; i < 16 * 64; i += 256) {
        int kk = i >> 6, cc = i & 63;
        int d = d0 + kk;
        sw[kk][cc] = (d < D && c0 + cc < H) ? W[(int64_t)d * H + c0 + cc] : 0.0;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < 16; ++kk) {
        double a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = sy[kk][ty + 16 * i];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = sw[kk][tx + 16 * j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int64_t r = r0 + ty + 16 * i;
        int c = c0 + tx + 16 * j;
        if (r < N && c < H) B[r * H + c] = acc[i][j];
      }
  }
}

// ------------------------------------------------------------------------------------------
// posterior weights of one datapoint into shared memory: q[k] = exp(lpj_k - max) / sum over all
// S + S_perm columns (bsc.py:176-177,219).  Returns nothing; q has S + S_perm entries.
// ------------------------------------------------------------------------------------------
template <int T>
__device__ void posterior(const double *__restrict__ row, int cols, double *q, double *red_d) {
  double mx = -DBL_MAX;
  for (int i = threadIdx.x; i < cols; i += T) { double v = row[i]; q[i] = v; mx = fmax(mx, v); }
  mx = group_max<T>(mx, red_d);
  double se = 0.0;
  for (int i = threadIdx.x; i < cols; i += T) { double e = exp(q[i] - mx); q[i] = e; se += e; }
  se = group_sum<T>(se, red_d);
  double inv = 1.0 / se;
  for (int i = threadIdx.x; i < cols; i += T) q[i] *= inv;
  __syncthreads();
}

constexpr double Q_SKIP = 1e-18;  // posterior mass below f64 resolution of the sums it would enter

// ------------------------------------------------------------------------------------------
// BSC M-step statistics (bsc.py:193-223).  One CTA per datapoint (persistent loop):
//   es_h = sum_k q_k s_kh                (smem, f64)
//   pies += es ; Wp[h,:] += es_h y_n ; Wq[h,h'] += sum_k q_k s_kh s_kh' (h' <= h) ;
//   sigma += sum_k q_k |y - W s_k|^2 = sum_k q_k (lpj_k - pil_bar |s_k|) / pre1   (same theta)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// One WARP per datapoint (persistent loop; no CTA barrier anywhere).  Lane k owns state k of a batch of
// 32: it lists the state's active latents once (SS_LS entries; longer states take the bit-scan path)
// and issues one global f64 atomic per pair.  <s>_n is gathered latent by latent over the union of the
// states' active latents (lane per latent, fixed summation order k = 0..S-1), so no shared-memory atomics.
constexpr int SS_LS = 32, SS_LD = 34;  // list capacity / row stride in uint16 (17 words: conflict-free rows)
__host__ __device__ inline size_t suffstats_warp_bytes(int H, int Hw, int cols) {
  size_t b = (size_t)8 * H + (size_t)8 * cols;       // es, q
  b += ((size_t)4 * Hw + 7) & ~(size_t)7;             // union words
  b += ((size_t)2 * H + 7) & ~(size_t)7;              // act
  b += (size_t)2 * 32 * SS_LD;                        // per-lane state lists
  b += ((size_t)2 * cols + 7) & ~(size_t)7;           // live states
  return (b + 15) & ~(size_t)15;
}

__global__ void __launch_bounds__(256) bsc_suffstats_kernel(const uint32_t *__restrict__ states,
                                                          const double *__restrict__ lpj,
                                                          const float *__restrict__ Ym, int64_t N, int D, int H, int Hw,
                                                          int S, int S_perm, double pre1, double pil_bar,
                                                          double *__restrict__ stats, float *__restrict__ Es_out,
                                                          WpLists lists, const int32_t *__restrict__ n_list,
                                                          const double *__restrict__ theta_dev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (theta_dev) { pre1 = theta_dev[0]; pil_bar = theta_dev[1]; }  // CUDA-graph replay: scalars from device memory
  const int cols = S + S_perm;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
  unsigned char *mine = smem_raw + (size_t)warp * suffstats_warp_bytes(H, Hw, cols);
  double *es = reinterpret_cast<double *>(mine);  // [H]
  double *q = es + H;                              // [cols]
  uint32_t *uni = reinterpret_cast<uint32_t *>(q + cols);
  uint16_t *act = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(uni) + (((size_t)4 * Hw + 7) & ~(size_t)7));
  uint16_t *slist = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(act) + (((size_t)2 * H + 7) & ~(size_t)7));
  uint16_t *my_list = slist + lane * SS_LD;
  uint16_t *lk = slist + 32 * SS_LD;  // [S] states that carry posterior mass (q > Q_SKIP), ascending
  double *Wp = stats, *Wq = stats + (int64_t)H * D, *pies = Wq + (int64_t)H * H, *sig = pies + H;
  double sig_acc = 0.0;
  for (int h = lane; h < H; h += 32) es[h] = 0.0;  // kept zero between datapoints (active entries reset below)
  __syncwarp();
  // with n_list only the datapoints listed there ([0] = count, [1..]): the ones the fused E-step tail did not cover
  const int64_t n_items = n_list ? (int64_t)n_list[0] : N;
  for (int64_t it = (int64_t)blockIdx.x * wpc + warp; it < n_items; it += (int64_t)gridDim.x * wpc) {
    const int64_t n = n_list ? (int64_t)n_list[1 + it] : it;
    // ---- posterior weights (bsc.py:176-177,219) ------------------------------------------------
    const double *row = lpj + n * cols;
    double mx = -DBL_MAX;
    for (int i = lane; i < cols; i += 32) { double v = row[i]; q[i] = v; mx = fmax(mx, v); }
    mx = warp_max_d(mx);
    double se = 0.0;
    for (int i = lane; i < cols; i += 32) { double e = exp(q[i] - mx); q[i] = e; se += e; }
    se = warp_sum_d(se);
    const double inv = 1.0 / se;
    for (int i = lane; i < cols; i += 32) q[i] *= inv;
    __syncwarp();
    // ---- states that carry posterior mass; the others (q <= 1e-18) are below the f64 resolution of
    //      every sum they would enter and are not even read ------------------------------------------
    const uint32_t *st = states + n * (int64_t)S * Hw;
    int nlive = 0;
    for (int k0 = 0; k0 < S; k0 += 32) {
      const int k = k0 + lane;
      const bool f = k < S && q[S_perm + k] > Q_SKIP;
      const unsigned m = __ballot_sync(0xffffffffu, f);
      if (f) lk[nlive + __popc(m & ((1u << lane) - 1u))] = (uint16_t)k;
      nlive += __popc(m);
    }
    __syncwarp();
    // ---- per live state: sigma, Wq pairs -------------------------------------------------------
    double sg = 0.0;
#pragma unroll 1
    for (int i0 = lane; i0 < nlive; i0 += 32) {
      const int k = lk[i0];
      const double qk = q[S_perm + k];
      const uint32_t *sk = st + (int64_t)k * Hw;
      int a = 0;
      for (int w = 0; w < Hw; ++w) {
        uint32_t m = sk[w];
        while (m) {
          const int bp = __ffs(m) - 1;
          m &= m - 1;
          if (a < SS_LS) my_list[a] = (uint16_t)((w << 5) + bp);
          ++a;
        }
      }
      sg += qk * (row[S_perm + k] - pil_bar * (double)a) / pre1;
      if (a <= SS_LS) {
        for (int i = 0; i < a; ++i) {
          double *wrow = Wq + (int64_t)my_list[i] * H;
          for (int j = 0; j <= i; ++j) atomicAdd(wrow + my_list[j], qk);  // h' <= h
        }
      } else {  // long state: enumerate the pairs from the bits
        for (int w1 = 0; w1 < Hw; ++w1) {
          uint32_t m1 = sk[w1];
          while (m1) {
            const int b1 = __ffs(m1) - 1;
            m1 &= m1 - 1;
            double *wrow = Wq + (int64_t)((w1 << 5) + b1) * H;
            for (int w2 = 0; w2 <= w1; ++w2) {
              uint32_t m2 = sk[w2];
              if (w2 == w1) m2 &= (b1 == 31) ? 0xFFFFFFFFu : ((2u << b1) - 1u);
              while (m2) {
                const int b2 = __ffs(m2) - 1;
                m2 &= m2 - 1;
                atomicAdd(wrow + (w2 << 5) + b2, qk);
              }
            }
          }
        }
      }
    }
    if (S_perm > 0 && lane == 0) sg += q[0] * row[0] / pre1;  // all-zero state: |y|^2
    sig_acc += sg;
    // ---- <s>_n over the union of the live states' active latents --------------------------------
    for (int w = lane; w < Hw; w += 32) {
      uint32_t u = 0u;
      for (int i = 0; i < nlive; ++i) u |= st[(int64_t)lk[i] * Hw + w];
      uni[w] = u;
    }
    __syncwarp();
    int nu = 0;  // union, ascending -> act
    for (int w0 = 0; w0 < Hw; w0 += 32) {
      uint32_t word = (w0 + lane < Hw) ? uni[w0 + lane] : 0u;
      const int c = __popc(word);
      int pre = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
      int pos = nu + pre - c;
      while (word) {
        const int bp = __ffs(word) - 1;
        word &= word - 1;
        act[pos++] = (uint16_t)(((w0 + lane) << 5) + bp);
      }
      nu += __shfl_sync(0xffffffffu, pre, 31);
    }
    __syncwarp();
    for (int i = lane; i < nu; i += 32) {
      const int h = act[i], w = h >> 5, bit = h & 31;
      double e = 0.0;
      for (int j = 0; j < nlive; ++j) {
        const int k = lk[j];
        if ((st[(int64_t)k * Hw + w] >> bit) & 1u) e += q[S_perm + k];
      }
      es[h] = e;  // > 0: the latent is active in a live state
    }
    __syncwarp();
    const int na = nu;
    if (Es_out)
      for (int h = lane; h < H; h += 32) Es_out[n * H + h] = (float)es[h];
    for (int i = lane; i < na; i += 32) atomicAdd(pies + act[i], es[act[i]]);
    if (lists.et_hi) {
      // tensor-core Wp: 2^14 <s> as fp16 hi + lo into the transposed planes (zero elsewhere)
      for (int i = lane; i < na; i += 32) {
        const int h = act[i];
        const double v = es[h] * WP_TC_SCALE;
        const __half hi = __double2half(v);
        lists.et_hi[(int64_t)h * lists.et_ld + n] = hi;
        lists.et_lo[(int64_t)h * lists.et_ld + n] = __double2half(v - (double)__half2float(hi));
      }
    } else if (lists.off) {
      // hand the sparse <s>_n (ascending h) to the Wp pass; overflow entries are accumulated here
      const int nl = na < WP_LCAP ? na : WP_LCAP;
      for (int t = lane; t <= WP_WARPS; t += 32) {
        const int hb = (H + WP_WARPS - 1) / WP_WARPS, bound = t * hb;
        int lo = 0, hi = nl;  // first entry with h >= bound
        while (lo < hi) { int mid = (lo + hi) >> 1; if (act[mid] < bound) lo = mid + 1; else hi = mid; }
        lists.off[n * WP_OFF + t] = (uint8_t)lo;
      }
      for (int i = lane; i < nl; i += 32) {
        lists.h[n * WP_LCAP + i] = act[i];
        lists.e[n * WP_LCAP + i] = es[act[i]];
      }
    }
    for (int i = lists.et_hi ? na : (lists.off ? WP_LCAP : 0); i < na; ++i) {  // not listed: straight to Wp
      const double e = es[act[i]];
      double *wrow = Wp + (int64_t)act[i] * D;
      for (int d = lane; d < D; d += 32) atomicAdd(wrow + d, e * (double)Ym[n * D + d]);
    }
    __syncwarp();
    for (int i = lane; i < na; i += 32) es[act[i]] = 0.0;
    __syncwarp();
  }
  sig_acc = warp_sum_d(sig_acc);
  if (lane == 0 && sig_acc != 0.0) atomicAdd(sig, sig_acc);
}

// ------------------------------------------------------------------------------------------
// Wp[h, :] += sum_n <s_h>_n y_n  (bsc.py:204-206) from the sparse lists of the statistics pass.
// CTA = (slice of WP_CW observables, contiguous chunk of datapoints).  The H x WP_CW slice of Wp is
// accumulated in SHARED memory (f64); warp w owns the latents h = w (mod 16), so no two warps ever
// touch the same cell and no atomics are needed until the single flush at the end.  Lists and the
// y slice of WP_BATCH datapoints arrive through a cp.async double buffer (one HBM latency per batch,
// hidden behind the previous batch).
// ------------------------------------------------------------------------------------------
constexpr int WP_BATCH = 16, WP_THREADS = 32 * WP_WARPS;
struct WpBatch {
  double e[WP_BATCH][WP_LCAP];
  float y[WP_BATCH][WP_CW];
  uint16_t h[WP_BATCH][WP_LCAP];
  uint8_t off[WP_BATCH][WP_OFF];
};
__device__ __forceinline__ void cp_async16(void *dst, const void *src, bool pred) {
  unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  int sz = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async4(void *dst, const void *src, bool pred) {
  unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  int sz = pred ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(sz) : "memory");
}

__global__ void __launch_bounds__(WP_THREADS) bsc_wp_kernel(WpLists lists, const float *__restrict__ Ym, int64_t N, int D,
                                                            int H, int n_slices, double *__restrict__ Wp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WpBatch *buf = reinterpret_cast<WpBatch *>(smem_raw);               // [2]
  double *acc = reinterpret_cast<double *>(smem_raw + 2 * sizeof(WpBatch));  // [H][WP_CW]
  const int slice = blockIdx.x % n_slices, chunk = blockIdx.x / n_slices, chunks = gridDim.x / n_slices;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t lo = N * chunk / chunks, hi = N * (chunk + 1) / chunks;
  const int nb = (int)((hi - lo + WP_BATCH - 1) / WP_BATCH);
  const int col0 = slice * WP_CW;
  for (int i = threadIdx.x; i < H * WP_CW; i += WP_THREADS) acc[i] = 0.0;

  auto issue = [&](int b) {
    WpBatch &B = buf[b & 1];
    const int64_t n0 = lo + (int64_t)b * WP_BATCH;
    for (int i = threadIdx.x; i < WP_BATCH * WP_LCAP / 2; i += WP_THREADS) {  // e: 16 B = 2 doubles
      const bool ok = n0 + i / (WP_LCAP / 2) < hi;
      cp_async16(&B.e[0][0] + 2 * i, ok ? lists.e + n0 * WP_LCAP + 2 * i : lists.e, ok);
    }
    for (int i = threadIdx.x; i < WP_BATCH * WP_LCAP / 8; i += WP_THREADS) {  // h: 16 B = 8 uint16
      const bool ok = n0 + i / (WP_LCAP / 8) < hi;
      cp_async16(&B.h[0][0] + 8 * i, ok ? lists.h + n0 * WP_LCAP + 8 * i : lists.h, ok);
    }
    for (int i = threadIdx.x; i < WP_BATCH * WP_CW; i += WP_THREADS) {        // y slice, 4 B each (any D)
      const int j = i / WP_CW, c = i % WP_CW;
      const bool ok = n0 + j < hi && col0 + c < D;
      cp_async4(&B.y[j][c], ok ? Ym + (n0 + j) * D + col0 + c : Ym, ok);
    }
    for (int i = threadIdx.x; i < WP_BATCH * WP_OFF / 4; i += WP_THREADS) {   // offsets (zero-filled past the end)
      const bool ok = n0 + i / (WP_OFF / 4) < hi;
      cp_async4(&B.off[0][0] + 4 * i, ok ? lists.off + n0 * WP_OFF + 4 * i : lists.off, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  if (nb > 0) issue(0);
  for (int b = 0; b < nb; ++b) {
    if (b + 1 < nb) {
      issue(b + 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const WpBatch &B = buf[b & 1];
#pragma unroll 2
    for (int j = 0; j < WP_BATCH; ++j) {
      const int o0 = B.off[j][warp], o1 = B.off[j][warp + 1];
      if (o0 == o1) continue;
      const double y0 = (double)B.y[j][lane], y1 = (double)B.y[j][lane + 32];
      for (int i = o0; i < o1; ++i) {
        double *row = acc + (size_t)B.h[j][i] * WP_CW;
        const double ee = B.e[j][i];
        row[lane] += ee * y0;
        row[lane + 32] += ee * y1;
      }
    }
    __syncthreads();  // this buffer is refilled by the load issued in the next iteration
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H * WP_CW; i += WP_THREADS) {
    const int h = i / WP_CW, c = col0 + (i % WP_CW);
    const double v = acc[i];
    if (c < D && v != 0.0) atomicAdd(Wp + (int64_t)h * D + c, v);
  }
}

// y_rec[n, ~keep] = W[~keep] <s>_n  (Model.reconstruct _models.py:614-665 + BSC.modelmean bsc.py:279-287)
template <int T>
__global__ void __launch_bounds__(T) bsc_reconstruct_kernel(const uint32_t *__restrict__ states,
                                                            const double *__restrict__ lpj, const double *__restrict__ W,
                                                            const float *__restrict__ Y, const uint8_t *__restrict__ keep,
                                                            const uint8_t *__restrict__ infer, int64_t N, int D, int H,
                                                            int Hw, int S, int S_perm, float *__restrict__ Yrec) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ double red_d[8];
  __shared__ int red_i[8];
  double *es = reinterpret_cast<double *>(smem_raw);  // [H]
  double *q = es + H;                                  // [S + S_perm]
  const int cols = S + S_perm;
  for (int64_t n = blockIdx.x; n < N; n += gridDim.x) {
    __syncthreads();
    int any = 0;
    if (infer) {
      for (int d = threadIdx.x; d < D; d += T) any |= infer[n * D + d];
      any = (int)group_or<T>((unsigned)any, red_i);
    } else {
      any = 1;
    }
    if (!any) {  // no usable entry: the reference leaves the row as the copy of y
      for (int d = threadIdx.x; d < D; d += T) Yrec[n * D + d] = Y[n * D + d];
      continue;
    }
    for (int h = threadIdx.x; h < H; h += T) es[h] = 0.0;
    posterior<T>(lpj + n * cols, cols, q, red_d);
    const uint32_t *st = states + n * (int64_t)S * Hw;
    for (int k = threadIdx.x; k < S; k += T) {
      double qk = q[S_perm + k];
      if (!(qk > Q_SKIP)) continue;
      const uint32_t *sk = st + (int64_t)k * Hw;
      for (int w = 0; w < Hw; ++w) {
        uint32_t m = sk[w];
        while (m) { int b = __ffs(m) - 1; m &= m - 1; atomicAdd(&es[(w << 5) + b], qk); }
      }
    }
    __syncthreads();
    for (int d = threadIdx.x; d < D; d += T) {
      float out = Y[n * D + d];
      if (!keep || !keep[n * D + d]) {
        double acc = 0.0;
        const double *Wr = W + (int64_t)d * H;
        for (int h = 0; h < H; ++h) { double e = es[h]; if (e != 0.0) acc += e * Wr[h]; }
        out = (float)acc;
      }
      Yrec[n * D + d] = out;
    }
  }
}

// sum_n logsumexp_k lpj[n][k]   (_models.py:544-547): one warp per row
__global__ void free_energy_kernel(const double *__restrict__ lpj, int64_t N, int cols, int64_t ld, double *out) {
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  double acc = 0.0;
  for (int64_t n = warp; n < N; n += nwarps) {
    const double *row = lpj + n * ld;
    double mx = -DBL_MAX;
    for (int i = lane; i < cols; i += 32) mx = fmax(mx, row[i]);
    mx = warp_max(mx);
    double se = 0.0;
    for (int i = lane; i < cols; i += 32) se += exp(row[i] - mx);
    se = warp_sum(se);
    acc += mx + log(se);
  }
  if (lane == 0 && acc != 0.0) atomicAdd(out, acc);
}

}  // namespace evo

using namespace evo;

static int grid_for(const void *kernel, int T, size_t smem, int64_t N) {
  int per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, T, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  int64_t g = (int64_t)sm_count() * per_sm;
  if (g > N) g = N;
  if (g < 1) g = 1;
  return (int)g;
}

extern "C" int evo_bsc_prepare(const double *W, const float *Y, int64_t N, int D, int H, double *G, double *B,
                               double *ynorm2, void *stream) {
  EVO_REQUIRE(N >= 0 && D > 0 && H > 0, "bad sizes");
  cudaStream_t st = (cudaStream_t)stream;
  if (G) {
    dim3 grid((H + 31) / 32, (H + 31) / 32);
    gram_kernel<<<grid, 256, 0, st>>>(W, G, D, H);
    EVO_LAUNCH_CHECK("gram");
  }
  if (N == 0) return EVO_OK;
  if (ynorm2) {
    int blocks = (int)((N * 32 + 255) / 256 < (int64_t)sm_count() * 8 ? (N * 32 + 255) / 256 : (int64_t)sm_count() * 8);
    ynorm_kernel<<<blocks, 256, 0, st>>>(Y, ynorm2, N, D);
    EVO_LAUNCH_CHECK("ynorm");
  }
  if (B) {
    int64_t tiles = ((N + 63) / 64) * ((H + 63) / 64);
    int blocks = (int)(tiles < (int64_t)sm_count() * 4 ? tiles : (int64_t)sm_count() * 4);
    wty_kernel<<<blocks, 256, 0, st>>>(Y, W, B, N, D, H);
    EVO_LAUNCH_CHECK("wty");
  }
  return EVO_OK;
}

template <class Eval>
static int launch_lpj(Eval ev, const uint32_t *states, int shared_states, int64_t N, int Sp, int Hw, double *out,
                      int64_t ld, unsigned long long *counters, cudaStream_t st) {
  size_t smem = ev.smem_bytes(Sp <= 32 ? 32 : 128);
  if (Sp <= 32) {
    auto k = lpj_kernel<Eval, 32>;
    if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid_for((const void *)k, 32, smem, N), 32, smem, st>>>(ev, states, shared_states, N, Sp, Hw, out, ld, counters);
  } else {
    auto k = lpj_kernel<Eval, 128>;
    if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid_for((const void *)k, 128, smem, N), 128, smem, st>>>(ev, states, shared_states, N, Sp, Hw, out, ld, counters);
  }
  EVO_LAUNCH_CHECK("lpj");
  return EVO_OK;
}

extern "C" int evo_bsc_lpj(int eval, const uint32_t *states, int shared_states, int64_t N, int Sp, int D, int H,
                           const double *W, const float *Y, const uint8_t *mask, const double *G, const double *B,
                           const double *ynorm2, double pre1, double pil_bar, double *lpj_out, int64_t ld_lpj,
                           unsigned long long *reset_counters, void *stream) {
  EVO_REQUIRE(N >= 0 && Sp > 0 && D > 0 && H > 0 && ld_lpj >= Sp, "bad sizes");
  if (N == 0) return EVO_OK;
  int Hw = words_for(H);
  cudaStream_t st = (cudaStream_t)stream;
  if (eval == EVO_EVAL_BSC_GRAM) {
    EVO_REQUIRE(G && B && ynorm2, "gram evaluator needs G, B, ynorm2");
    EVO_REQUIRE(mask == nullptr, "gram evaluator needs complete data");
    BscGramEval ev{G, B, ynorm2, pre1, pil_bar, H, Hw, nullptr, nullptr, 0.0};
    return launch_lpj(ev, states, shared_states, N, Sp, Hw, lpj_out, ld_lpj, reset_counters, st);
  } else if (eval == EVO_EVAL_BSC_DIRECT) {
    EVO_REQUIRE(W && Y, "direct evaluator needs W, Y");
    BscDirectEval ev{W, Y, mask, pre1, pil_bar, D, H, Hw, nullptr, nullptr, 0.0};
    return launch_lpj(ev, states, shared_states, N, Sp, Hw, lpj_out, ld_lpj, reset_counters, st);
  }
  set_error("evo_bsc_lpj: unknown evaluator %d", eval);
  return EVO_ERR_ARG;
}

template <class Eval>
static int launch_full(Eval ev, int64_t N, int Hm, uint32_t or_mask, double *out, cudaStream_t st) {
  constexpr int T = 128;
  const size_t smem = (sizeof(uint32_t) + sizeof(double)) * FULL_B + ev.smem_bytes(T);
  auto k = full_logsumexp_kernel<Eval, T>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k<<<grid_for((const void *)k, T, smem, N), T, smem, st>>>(ev, N, Hm, or_mask, out);
  EVO_LAUNCH_CHECK("full_logsumexp");
  return EVO_OK;
}

extern "C" int evo_bsc_full_logsumexp(int eval, int64_t N, int D, int H, int background, const double *W, const float *Y,
                                      const uint8_t *mask, const double *G, const double *B, const double *ynorm2,
                                      double pre1, double pil_bar, double *out, void *stream) {
  EVO_REQUIRE(N >= 0 && D > 0 && H > 0 && out, "bad arguments");
  const int Hm = background ? H - 1 : H;
  EVO_REQUIRE(H <= 32 && Hm <= 24, "exact likelihood: at most 24 enumerated latents (2^24 states per datapoint)");
  if (N == 0) return EVO_OK;
  const uint32_t or_mask = background ? (1u << (H - 1)) : 0u;
  cudaStream_t st = (cudaStream_t)stream;
  if (eval == EVO_EVAL_BSC_GRAM) {
    EVO_REQUIRE(G && B && ynorm2 && !mask, "gram evaluator needs G, B, ynorm2 and complete data");
    BscGramEval ev{G, B, ynorm2, pre1, pil_bar, H, 1, nullptr, nullptr, 0.0};
    return launch_full(ev, N, Hm, or_mask, out, st);
  } else if (eval == EVO_EVAL_BSC_DIRECT) {
    EVO_REQUIRE(W && Y, "direct evaluator needs W, Y");
    BscDirectEval ev{W, Y, mask, pre1, pil_bar, D, H, 1, nullptr, nullptr, 0.0};
    return launch_full(ev, N, Hm, or_mask, out, st);
  }
  set_error("evo_bsc_full_logsumexp: unknown evaluator %d", eval);
  return EVO_ERR_ARG;
}

extern "C" int evo_free_energy(const double *lpj, int64_t N, int cols, int64_t ld, double *out, void *stream) {
  EVO_REQUIRE(N >= 0 && cols > 0 && ld >= cols, "bad sizes");
  if (N == 0) return EVO_OK;
  int blocks = (int)((N * 32 + 255) / 256 < (int64_t)sm_count() * 8 ? (N * 32 + 255) / 256 : (int64_t)sm_count() * 8);
  free_energy_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(lpj, N, cols, ld, out);
  EVO_LAUNCH_CHECK("free_energy");
  return EVO_OK;
}

static size_t wp_smem(int H) { return 2 * sizeof(WpBatch) + (size_t)H * WP_CW * sizeof(double); }
static bool wp_lists_applicable(int H) { return wp_smem(H) <= 227 * 1024 && H <= 65535; }

int evo::bsc_suffstats_pass1(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H, int S,
                             int S_perm, double pre1, double pil_bar, double *stats, float *Es_out, WpLists lists,
                             cudaStream_t stream, const int32_t *n_list, const double *theta_dev) {
  const int Hw = words_for(H);
  EVO_REQUIRE(H <= 65535, "H > 65535 not supported by the statistics kernel");
  const size_t per_warp = suffstats_warp_bytes(H, Hw, S + S_perm);
  int wpc = (int)((200 * 1024) / per_warp);
  if (wpc > 8) wpc = 8;
  if (wpc < 1) {
    set_error("evo_bsc_suffstats: H = %d needs %zu bytes of shared memory per datapoint", H, per_warp);
    return EVO_ERR_UNSUPPORTED;
  }
  const size_t smem = per_warp * wpc;
  auto k = bsc_suffstats_kernel;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k<<<grid_for((const void *)k, 32 * wpc, smem, (N + wpc - 1) / wpc), 32 * wpc, smem, stream>>>(
      states, lpj, Ym, N, D, H, Hw, S, S_perm, pre1, pil_bar, stats, Es_out, lists, n_list, theta_dev);
  EVO_LAUNCH_CHECK("bsc_suffstats");
  return EVO_OK;
}

extern "C" int64_t evo_bsc_suffstats_workspace_bytes(int64_t N, int D, int H) {
  (void)D;
  if (N <= 0 || !wp_lists_applicable(H)) return 0;
  return (int64_t)(N * (WP_LCAP * (sizeof(double) + sizeof(uint16_t)) + WP_OFF) + 64);
}

extern "C" int evo_bsc_suffstats(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H,
                                 int S, int S_perm, int allzero, double pre1, double pil_bar, double *stats,
                                 float *Es_out, void *workspace, int64_t workspace_bytes, void *stream) {
  (void)allzero;
  EVO_REQUIRE(N >= 0 && D > 0 && H > 0 && S > 0, "bad sizes");
  if (N == 0) return EVO_OK;
  WpLists lists{};
  if (workspace && wp_lists_applicable(H) && workspace_bytes >= evo_bsc_suffstats_workspace_bytes(N, D, H)) {
    unsigned char *p = reinterpret_cast<unsigned char *>(workspace);
    EVO_REQUIRE(((uintptr_t)p & 15) == 0, "workspace must be 16-byte aligned");
    lists.e = reinterpret_cast<double *>(p);
    lists.h = reinterpret_cast<uint16_t *>(p + (size_t)N * WP_LCAP * sizeof(double));
    lists.off = p + (size_t)N * WP_LCAP * (sizeof(double) + sizeof(uint16_t));
  }
  int rc = bsc_suffstats_pass1(states, lpj, Ym, N, D, H, S, S_perm, pre1, pil_bar, stats, Es_out, lists,
                               (cudaStream_t)stream);
  if (rc) return rc;
  if (lists.off) {
    int n_slices = (D + WP_CW - 1) / WP_CW;
    int chunks = sm_count() / n_slices;
    if (chunks < 1) chunks = 1;
    if ((int64_t)chunks > N) chunks = (int)N;
    size_t wsm = wp_smem(H);
    if (wsm > 48 * 1024)
      EVO_CUDA(cudaFuncSetAttribute(bsc_wp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
    bsc_wp_kernel<<<n_slices * chunks, WP_THREADS, wsm, (cudaStream_t)stream>>>(lists, Ym, N, D, H, n_slices, stats);
    EVO_LAUNCH_CHECK("bsc_wp");
  }
  return EVO_OK;
}

extern "C" int evo_bsc_reconstruct(const uint32_t *states, const double *lpj, const double *W, const float *Y,
                                   const uint8_t *recon_keep, const uint8_t *infer_mask, int64_t N, int D, int H, int S,
                                   int S_perm, float *Yrec, void *stream) {
  EVO_REQUIRE(N >= 0 && D > 0 && H > 0 && S > 0, "bad sizes");
  if (N == 0) return EVO_OK;
  int Hw = words_for(H);
  size_t smem = sizeof(double) * ((size_t)H + S + S_perm);
  constexpr int T = 128;
  auto k = bsc_reconstruct_kernel<T>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k<<<grid_for((const void *)k, T, smem, N), T, smem, (cudaStream_t)stream>>>(states, lpj, W, Y, recon_keep, infer_mask,
                                                                              N, D, H, Hw, S, S_perm, Yrec);
  EVO_LAUNCH_CHECK("bsc_reconstruct");
  return EVO_OK;
}
