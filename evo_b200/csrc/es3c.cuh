// ES3C (spike-and-slab sparse coding, reference class SSSC) -- warp-per-state small dense systems.
//
// For a state with support A (a = |A| <= 32) the reference (evo/models/sssc.py:242-326) forms
//   M = W_A' W_A / s2 + Psi_A^-1,  Lambda = M^-1,  C_det = log|det M| + log|det Psi_A|,
//   lpj = -1/2 (C_det + r' C^-1 r) + sum_{h in A} pil_bar_h,   r = y_obs - W_A mu_A.
// Here the same numbers come from N = I + (1/s2) G_AA Psi_A  (M = N Psi_A^-1, so C_det = log|det N|,
// Lambda = Psi_A N^-1 and r'C^-1 r = |r|^2/s2 - v' Lambda v with v = W_A' r / s2): one LU with partial
// pivoting per state, no inverse of Psi_A, valid for the indefinite Psi the reference's element-wise
// update can produce.  G_AA = W_A' W_A over the observed rows: gathered from the shared
// G = W'W for complete data, accumulated from W' rows under the datapoint's mask otherwise.
#pragma once
#include "common.cuh"
#include "estep.cuh"

namespace evo {

struct Es3cEval {
  // complete data
  const double *G;       // [H][H]
  const double *B;       // [N][H]  Y W
  const double *ynorm2;  // [N]
  // incomplete data
  const double *Wt;      // [H][D]  W'
  const float *Y;        // [N][D]  (0 where masked)
  const uint8_t *mask;   // [N][D]
  // model
  const double *Psi, *mus, *pil_bar;  // [H][H], [H], [H]
  double s2i;
  int D, H, Hw, masked;  // masked: 0 complete data; 1 incomplete, Gram accumulated from W' rows under the mask per state;
                         // 2 incomplete with tables: G_AA = (W'W)_AA - sum_{d missing} W_dA' W_dA, b = (Y W)_A (Y is 0 where
                         // missing) -- |missing| instead of |observed| terms per pair, and no per-state reductions over D
  double *scratch;       // global per-warp workspaces for 16 < a <= 32 (or NULL)
  int scratch_warps;     // capacity of `scratch` in warps
  int max_batch;         // largest cnt eval_batch will see (index lists of the small / large systems); 0 = no lists
  // shared-memory budget of the launching kernel (set by the host; shared memory per datapoint decides how many
  // datapoints an SM holds): the E-step wants the packed 8-lane area and b staged, the statistics pass -- warp and
  // thread paths only -- a smaller warp area and no staging
  int as_cap, as_ld;     // warp path in shared memory: systems up to as_cap x as_cap, leading dimension as_ld
  int warp_bytes;        // per-warp area
  int stage_b;           // stage this datapoint's row of B in shared memory (else read it through L1 / L2)
  // per datapoint (setup)
  double *vec;           // smem: b [H] (complete) or y [D] (incomplete)
  uint8_t *m;            // smem [D] (incomplete)
  unsigned char *wsm;    // smem per-warp area base
  unsigned char *lsm;    // smem: int counters[2] + uint16 small[max_batch] + uint16 big[max_batch]
  double yn2;
  uint16_t *miss;        // smem (masked == 2): missing observables of this datapoint, ascending
  const int *nmiss_ptr;  // smem: their number

  // sum over the missing observables of W_dh W_dh' (the part of (W'W)_hh' this datapoint must not count)
  __device__ __forceinline__ double miss_corr(int h, int h2) const {
    const int nm = *nmiss_ptr;
    const double *wa = Wt + (int64_t)h * D, *wb = Wt + (int64_t)h2 * D;
    double c = 0.0;
    for (int q = 0; q < nm; ++q) { const int d = miss[q]; c += __ldg(wa + d) * __ldg(wb + d); }
    return c;
  }

  static constexpr int AS = 16, LDS = 17;   // shared-memory path
  static constexpr int AG = 32, LDGL = 33;  // global-scratch path
  static constexpr int G8 = 8, LD8 = 9;      // packed path: 4 states per warp, 8 lanes each (a <= 8)
  static constexpr size_t G8_BYTES = 16 + 3 * G8 * LD8 * sizeof(double);
  static constexpr size_t WARP_SMEM = 4 * G8_BYTES > 64 + 3 * AS * LDS * sizeof(double) ? 4 * G8_BYTES : 64 + 3 * AS * LDS * sizeof(double);
  __host__ void budget_estep() { as_cap = AS; as_ld = LDS; warp_bytes = (int)WARP_SMEM; stage_b = 1; }
  __host__ void budget_stats() { as_cap = AS; as_ld = LDS; warp_bytes = (int)WARP_SMEM; stage_b = 0; }
  static constexpr size_t WARP_SCRATCH = 3 * AG * LDGL;  // doubles

  __host__ __device__ size_t head_bytes() const {
    if (masked == 1) return align8((size_t)D * sizeof(double)) + align8((size_t)D);
    return (stage_b ? align8((size_t)H * sizeof(double)) : 0) + (masked == 2 ? align8(48 + (size_t)2 * D) : 0);
  }
  __host__ __device__ size_t list_bytes() const { return max_batch > 0 ? align8(16 + (size_t)6 * max_batch) : 0; }
  static constexpr bool kList = false;
  // resident threads per SM the register allocation should allow: shared memory holds 4 datapoints of 128 threads, so
  // up to 128 registers per thread cost nothing (ptxas otherwise settles on 96 registers and spills in the solvers)
  static constexpr int kThreadsPerSM = 512;
  __host__ __device__ size_t smem_bytes(int T, bool = false) const { return head_bytes() + (size_t)(T / 32) * warp_bytes + list_bytes(); }
  // measured (cfg3/cfg4, 64 states + 64 children): complete data is fastest with 4 warps and more resident
  // CTAs, masked data (a D-long Gram gather per system) with 8 warps per datapoint
  __host__ int threads_hint(int work) const { return work < 64 ? 64 : (masked == 1 ? 256 : 128); }

  template <int T>
  __device__ void setup(int64_t n, unsigned char *smem) {
    vec = reinterpret_cast<double *>(smem);
    if (masked == 1) {
      m = smem + align8((size_t)D * sizeof(double));
      for (int d = threadIdx.x; d < D; d += T) {
        uint8_t ok = mask[n * (int64_t)D + d];
        m[d] = ok;
        vec[d] = ok ? (double)Y[n * (int64_t)D + d] : 0.0;
      }
    } else {
      const double *src = B + n * (int64_t)H;
      if (stage_b) {
        for (int i = threadIdx.x; i < H; i += T) vec[i] = __ldg(src + i);
      } else {
        vec = const_cast<double *>(src);  // read-only from here on
      }
      yn2 = __ldg(ynorm2 + n);
      if (masked == 2) {  // ordered list of this datapoint's missing observables: [8 ints scratch | count | uint16 list]
        int *scr = reinterpret_cast<int *>(smem + (stage_b ? align8((size_t)H * sizeof(double)) : 0));
        miss = reinterpret_cast<uint16_t *>(scr + 12);
        int cnt = 0;
        for (int d0 = 0; d0 < D; d0 += T) {
          const int d = d0 + threadIdx.x;
          const bool f = d < D && !mask[n * (int64_t)D + d];
          int tot;
          const int off = group_scan_flag<T>(f, tot, scr);
          if (f) miss[cnt + off] = (uint16_t)d;
          cnt += tot;
          gsync<T>();
        }
        if (threadIdx.x == 0) scr[8] = cnt;
        nmiss_ptr = scr + 8;
      }
    }
    wsm = smem + head_bytes();
    lsm = wsm + (size_t)(T / 32) * warp_bytes;
  }

  // |y_obs|^2 for incomplete data (needs the loaded y); call after the post-setup barrier
  __device__ double masked_yn2() const {
    double s = 0.0;
    for (int d = (threadIdx.x & 31); d < D; d += 32) s += vec[d] * vec[d];
    return warp_sum(s);
  }

  // ---- active-latent list of one state, warp-cooperative; returns a (may exceed 32) -------------
  __device__ __forceinline__ int active_list(const uint32_t *st, uint16_t *idx) const {
    const int lane = threadIdx.x & 31;
    int base = 0;
    for (int w0 = 0; w0 < Hw; w0 += 32) {
      uint32_t word = (w0 + lane < Hw) ? st[w0 + lane] : 0u;
      int c = __popc(word);
      int pre = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
      int pos = base + pre - c;
      while (word) {
        int bp = __ffs(word) - 1;
        word &= word - 1;
        if (pos < 32) idx[pos] = (uint16_t)(((w0 + lane) << 5) + bp);
        ++pos;
      }
      base += __shfl_sync(0xffffffffu, pre, 31);
    }
    __syncwarp();
    return base;
  }

  // workspaces for a state of size a: three matrices with leading dimension ld
  __device__ __forceinline__ bool workspace(int a, uint16_t *&idx, double *&A1, double *&A2, double *&A3, int &ld) const {
    const int warp = threadIdx.x >> 5;
    unsigned char *base = wsm + (size_t)warp * warp_bytes;
    idx = reinterpret_cast<uint16_t *>(base);
    if (a <= as_cap) {
      A1 = reinterpret_cast<double *>(base + 64);
      A2 = A1 + as_cap * as_ld;
      A3 = A2 + as_cap * as_ld;
      ld = as_ld;
      return true;
    }
    int gw = blockIdx.x * (blockDim.x >> 5) + warp;
    if (a > AG || scratch == nullptr || gw >= scratch_warps) return false;
    A1 = scratch + (size_t)gw * WARP_SCRATCH;
    A2 = A1 + AG * LDGL;
    A3 = A2 + AG * LDGL;
    ld = LDGL;
    return true;
  }

  // States too large for a warp's own area (as_cap < a <= AG) can borrow the areas of ALL warps of the CTA, one state at a
  // time, while the other warps wait at a barrier -- shared memory instead of the global scratch, whose dependent
  // round trips to L2 made 3 % of the states most of the statistics pass at density (the E-step keeps the
  // global scratch: there the serial phase cost more than it saved).  cta_cap: the largest a that fits.
  __device__ __forceinline__ int cta_cap(int nwarps) const {
    const int bytes = nwarps * warp_bytes - 64;
    int a = AG;
    while (a > 0 && 3 * a * (a + 1) * (int)sizeof(double) > bytes) --a;
    return a;
  }
  __device__ __forceinline__ void workspace_cta(int cap, uint16_t *&idx, double *&A1, double *&A2, double *&A3, int &ld) const {
    idx = reinterpret_cast<uint16_t *>(wsm);
    A1 = reinterpret_cast<double *>(wsm + 64);
    ld = cap + 1;
    A2 = A1 + cap * ld;
    A3 = A2 + cap * ld;
  }

  // Gathers G_AA -> A1, Psi_AA -> A2 and returns this lane's b_i (W_A' y)_i; lane i <-> latent idx[i].
  __device__ __forceinline__ double gather(int a, const uint16_t *idx, double *A1, double *A2, int ld, double &ynorm) const {
    const int lane = threadIdx.x & 31;
    double bi = 0.0;
    if (masked == 2) {
      // (W'W)_AA minus the missing observables' share, the warp striding over the missing list for every pair
      const int nm = *nmiss_ptr;
      for (int i = 0; i < a; ++i) {
        const double *wa = Wt + (int64_t)idx[i] * D;
        for (int j = 0; j <= i; ++j) {
          const double *wb = Wt + (int64_t)idx[j] * D;
          double c = 0.0;
          for (int q = lane; q < nm; q += 32) { const int d = miss[q]; c += __ldg(wa + d) * __ldg(wb + d); }
          c = warp_sum(c);
          if (lane == 0) {
            const double v = __ldg(G + (int64_t)idx[i] * H + idx[j]) - c;
            A1[i * ld + j] = v;
            A1[j * ld + i] = v;
          }
        }
      }
      if (lane < a) bi = vec[idx[lane]];
      ynorm = yn2;
    } else if (masked == 0) {
      if (lane < a) {
        int h = idx[lane];
        const double *Gr = G + (int64_t)h * H;
        for (int l = 0; l < a; ++l) A1[lane * ld + l] = __ldg(Gr + idx[l]);
        bi = vec[h];
      }
      ynorm = yn2;
    } else {
      for (int i = 0; i < a; ++i) {
        const double *wi = Wt + (int64_t)idx[i] * D;
        double pb = 0.0;
        for (int d = lane; d < D; d += 32) pb += __ldg(wi + d) * vec[d];
        pb = warp_sum(pb);
        if (lane == i) bi = pb;
        for (int j = 0; j <= i; ++j) {
          const double *wj = Wt + (int64_t)idx[j] * D;
          double p = 0.0;
          for (int d = lane; d < D; d += 32)
            if (m[d]) p += __ldg(wi + d) * __ldg(wj + d);
          p = warp_sum(p);
          if (lane == 0) { A1[i * ld + j] = p; A1[j * ld + i] = p; }
        }
      }
      ynorm = yn2;
    }
    if (lane < a) {
      const double *Pr = Psi + (int64_t)idx[lane] * H;
      for (int l = 0; l < a; ++l) A2[lane * ld + l] = __ldg(Pr + idx[l]);
    }
    __syncwarp();
    return bi;
  }

  // LU with partial pivoting of the a x a matrix A3 (lane = row), solving A3 x = rhs in place.
  // Returns log|det| ; x_i is left in `rhs` of lane i.
  __device__ __forceinline__ double lu_solve(int a, double *A3, int ld, double &rhs) const {
    const int lane = threadIdx.x & 31;
    const bool act = lane < a;
    // log|det| = log|prod piv|: the product is carried as mantissa x 2^esum (one log per state
    // instead of one per pivot); pivots that are 0 / denormal / non-finite take the direct log.
    double prod = 1.0, extra = 0.0;
    int esum = 0;
    for (int k = 0; k < a; ++k) {
      // partial pivoting: largest |A3[i][k]|, i >= k, found by one redux on the f32 magnitude bits with
      // the (inverted) lane in the 5 low bits -- any near-maximal pivot is as good numerically
      unsigned key = 0u;
      if (act && lane >= k) key = (__float_as_uint(fabsf((float)A3[lane * ld + k])) & ~31u) | (unsigned)(31 - lane);
      const int p = 31 - (int)(__reduce_max_sync(0xffffffffu, key) & 31u);
      if (p != k) {
        if (act) { double t = A3[k * ld + lane]; A3[k * ld + lane] = A3[p * ld + lane]; A3[p * ld + lane] = t; }
        double rk = __shfl_sync(0xffffffffu, rhs, k), rp = __shfl_sync(0xffffffffu, rhs, p);
        if (lane == k) rhs = rp;
        if (lane == p) rhs = rk;
      }
      __syncwarp();
      double piv = A3[k * ld + k];
      {
        const int ex = (__double2hiint(piv) >> 20) & 0x7ff;
        if (ex == 0 || ex == 0x7ff) {
          extra += log(fabs(piv));
        } else {
          prod *= __hiloint2double((__double2hiint(piv) & 0x800fffff) | 0x3ff00000, __double2loint(piv));
          esum += ex - 1023;
          const int e2 = (__double2hiint(prod) >> 20) & 0x7ff;  // prod stays in [1, 4): renormalise
          prod = __hiloint2double((__double2hiint(prod) & 0x800fffff) | 0x3ff00000, __double2loint(prod));
          esum += e2 - 1023;
        }
      }
      double rhs_k = __shfl_sync(0xffffffffu, rhs, k);
      if (act && lane > k) {
        double f = A3[lane * ld + k] / piv;
        for (int c = k + 1; c < a; ++c) A3[lane * ld + c] -= f * A3[k * ld + c];
        rhs -= f * rhs_k;
      }
      __syncwarp();
    }
    for (int k = a - 1; k >= 0; --k) {
      double xk = __shfl_sync(0xffffffffu, rhs, k) / A3[k * ld + k];
      if (lane == k) rhs = xk;
      if (act && lane < k) rhs -= A3[lane * ld + k] * xk;
    }
    return log(fabs(prod)) + (double)esum * 0.693147180559945309417232121458 + extra;
  }

  // N = I + s2i A1 A2 -> A3 ; also v_i, |r|^2 and w = Psi_A' v  (v' Lambda v = w . N^-1 v)
  __device__ __forceinline__ void build(int a, const uint16_t *idx, const double *A1, const double *A2, double *A3, int ld,
                                        double bi, double ynorm, double &v, double &rn2, double &w, double &mu) const {
    const int lane = threadIdx.x & 31;
    const bool act = lane < a;
    mu = act ? __ldg(mus + idx[lane]) : 0.0;
    double t = 0.0;
    for (int l = 0; l < a; ++l) {
      double ml = __shfl_sync(0xffffffffu, mu, l);
      if (act) t += A1[lane * ld + l] * ml;
    }
    v = act ? s2i * (bi - t) : 0.0;
    rn2 = ynorm + warp_sum(act ? mu * (t - 2.0 * bi) : 0.0);
    w = 0.0;
    for (int l = 0; l < a; ++l) {
      double vl = __shfl_sync(0xffffffffu, v, l);
      if (act) w += A2[l * ld + lane] * vl;  // (Psi_A' v)_i: Psi need not be symmetric after the reference's update
    }
    if (act) {
      for (int j = 0; j < a; ++j) {
        double acc = 0.0;
        for (int l = 0; l < a; ++l) acc += A1[lane * ld + l] * A2[l * ld + j];
        A3[lane * ld + j] = (lane == j ? 1.0 : 0.0) + s2i * acc;
      }
    }
    __syncwarp();
  }

  // lpj of one state, all 32 lanes of the warp cooperate; result valid in every lane
  __device__ __forceinline__ double warp_lpj(const uint32_t *st) const {
    const int lane = threadIdx.x & 31;
    uint16_t *idx;
    double *A1, *A2, *A3;
    int ld;
    workspace(0, idx, A1, A2, A3, ld);
    int a = active_list(st, idx);
    if (a == 0) return -0.5 * s2i * yn2;
    if (!workspace(a, idx, A1, A2, A3, ld)) return nan("");  // more than 32 active latents: rejected
    double ynorm;
    double bi = gather(a, idx, A1, A2, ld, ynorm);
    double v, rn2, w, mu;
    build(a, idx, A1, A2, A3, ld, bi, ynorm, v, rn2, w, mu);
    double x = v;
    double logdet = lu_solve(a, A3, ld, x);
    const bool act = lane < a;
    double quad = s2i * rn2 - warp_sum(act ? w * x : 0.0);
    double prior = warp_sum(act ? __ldg(pil_bar + idx[lane]) : 0.0);
    __syncwarp();
    return -0.5 * (logdet + quad) + prior;
  }

  // ---- packed path: one state per 8 lanes (a <= 8), four states per warp -------------------------
  // Same arithmetic as warp_lpj (gather -> N = I + s2i G_AA Psi_A -> LU with partial pivoting -> quad
  // form), with every shuffle / redux / sync restricted to the group's 8-lane mask, so the four
  // groups of a warp run their own trip counts.
  // W lanes per state (a <= W), 32 / W states per warp; instantiated for W = 8 (W = 4 was measured: it only paid before the
  // thread path covered 3..4 active latents)
  template <int W>
  __device__ __forceinline__ double gw_sum(unsigned gm, double v) const {
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gm, v, o, W);
    return v;
  }
  template <int W>
  __device__ __forceinline__ double group_lpj(const uint32_t *st) const {
    constexpr int G8 = W, LD8 = W + 1;  // (the names of the 8-lane form)
    constexpr size_t GBYTES = 16 + 3 * W * (W + 1) * sizeof(double);
    const int lane = threadIdx.x & 31, sl = lane & (W - 1), gid = lane / W;
    const unsigned gm = ((1u << W) - 1u) << (W * gid);
    unsigned char *base = wsm + (size_t)(threadIdx.x >> 5) * warp_bytes + (size_t)gid * GBYTES;
    uint16_t *idx = reinterpret_cast<uint16_t *>(base);
    double *A1 = reinterpret_cast<double *>(base + 16), *A2 = A1 + G8 * LD8, *A3 = A2 + G8 * LD8;
    // active latents (the caller guarantees |s| <= 8)
    int a = 0;
    for (int w0 = 0; w0 < Hw; w0 += W) {
      uint32_t word = (w0 + sl < Hw) ? st[w0 + sl] : 0u;
      const int c = __popc(word);
      int pre = c;
#pragma unroll
      for (int o = 1; o < W; o <<= 1) { int t = __shfl_up_sync(gm, pre, o, W); if (sl >= o) pre += t; }
      int pos = a + pre - c;
      while (word) {
        int bp = __ffs(word) - 1;
        word &= word - 1;
        if (pos < G8) idx[pos] = (uint16_t)(((w0 + sl) << 5) + bp);
        ++pos;
      }
      a += __shfl_sync(gm, pre, W - 1, W);
    }
    __syncwarp(gm);
    if (a == 0) return -0.5 * s2i * yn2;
    const bool act = sl < a;
    // gather G_AA -> A1, Psi_AA -> A2, b_i
    double bi = 0.0;
    if (masked != 1) {
      if (act) {
        const int h = idx[sl];
        const double *Gr = G + (int64_t)h * H;
        if (masked == 2) {  // (W'W)_hl minus the missing observables' share, l <= sl: one walk over the missing list
          double c[W];
#pragma unroll
          for (int l = 0; l < W; ++l) c[l] = 0.0;
          const int nm = *nmiss_ptr;
          const double *wh = Wt + (int64_t)h * D;
          for (int q = 0; q < nm; ++q) {
            const int d = miss[q];
            const double x = __ldg(wh + d);
#pragma unroll
            for (int l = 0; l < W; ++l)
              if (l <= sl) c[l] += x * __ldg(Wt + (int64_t)idx[l] * D + d);
          }
#pragma unroll
          for (int l = 0; l < W; ++l) {
            if (l <= sl) {
              const double v = __ldg(Gr + idx[l]) - c[l];
              A1[sl * LD8 + l] = v;
              A1[l * LD8 + sl] = v;
            }
          }
        } else {
          for (int l = 0; l < a; ++l) A1[sl * LD8 + l] = __ldg(Gr + idx[l]);
        }
        bi = vec[h];
      }
    } else {
      for (int i = 0; i < a; ++i) {
        const double *wi = Wt + (int64_t)idx[i] * D;
        double pb = 0.0;
        for (int d = sl; d < D; d += W) pb += __ldg(wi + d) * vec[d];
        pb = gw_sum<W>(gm, pb);
        if (sl == i) bi = pb;
        for (int j = 0; j <= i; ++j) {
          const double *wj = Wt + (int64_t)idx[j] * D;
          double p = 0.0;
          for (int d = sl; d < D; d += W)
            if (m[d]) p += __ldg(wi + d) * __ldg(wj + d);
          p = gw_sum<W>(gm, p);
          if (sl == 0) { A1[i * LD8 + j] = p; A1[j * LD8 + i] = p; }
        }
      }
    }
    if (act) {
      const double *Pr = Psi + (int64_t)idx[sl] * H;
      for (int l = 0; l < a; ++l) A2[sl * LD8 + l] = __ldg(Pr + idx[l]);
    }
    __syncwarp(gm);
    // v, |r|^2, w = Psi_A' v, N -> A3
    const double mu = act ? __ldg(mus + idx[sl]) : 0.0;
    double t = 0.0;
    for (int l = 0; l < a; ++l) {
      double ml = __shfl_sync(gm, mu, l, W);
      if (act) t += A1[sl * LD8 + l] * ml;
    }
    const double v = act ? s2i * (bi - t) : 0.0;
    const double rn2 = yn2 + gw_sum<W>(gm, act ? mu * (t - 2.0 * bi) : 0.0);
    double w = 0.0;
    for (int l = 0; l < a; ++l) {
      double vl = __shfl_sync(gm, v, l, W);
      if (act) w += A2[l * LD8 + sl] * vl;
    }
    if (act) {
      for (int j = 0; j < a; ++j) {
        double acc = 0.0;
        for (int l = 0; l < a; ++l) acc += A1[sl * LD8 + l] * A2[l * LD8 + j];
        A3[sl * LD8 + j] = (sl == j ? 1.0 : 0.0) + s2i * acc;
      }
    }
    __syncwarp(gm);
    // LU with partial pivoting, rhs = v
    double rhs = v, prod = 1.0, extra = 0.0;
    int esum = 0;
    for (int k = 0; k < a; ++k) {
      unsigned key = 0u;
      if (act && sl >= k) key = (__float_as_uint(fabsf((float)A3[sl * LD8 + k])) & ~(unsigned)(W - 1)) | (unsigned)(W - 1 - sl);
      const int p = W - 1 - (int)(__reduce_max_sync(gm, key) & (unsigned)(W - 1));
      if (p != k) {
        if (act) { double x = A3[k * LD8 + sl]; A3[k * LD8 + sl] = A3[p * LD8 + sl]; A3[p * LD8 + sl] = x; }
        double rk = __shfl_sync(gm, rhs, k, W), rp = __shfl_sync(gm, rhs, p, W);
        if (sl == k) rhs = rp;
        if (sl == p) rhs = rk;
      }
      __syncwarp(gm);
      const double piv = A3[k * LD8 + k];
      {
        const int ex = (__double2hiint(piv) >> 20) & 0x7ff;
        if (ex == 0 || ex == 0x7ff) {
          extra += log(fabs(piv));
        } else {
          prod *= __hiloint2double((__double2hiint(piv) & 0x800fffff) | 0x3ff00000, __double2loint(piv));
          esum += ex - 1023;
          const int e2 = (__double2hiint(prod) >> 20) & 0x7ff;
          prod = __hiloint2double((__double2hiint(prod) & 0x800fffff) | 0x3ff00000, __double2loint(prod));
          esum += e2 - 1023;
        }
      }
      const double rhs_k = __shfl_sync(gm, rhs, k, W);
      if (act && sl > k) {
        const double f = A3[sl * LD8 + k] / piv;
        for (int c = k + 1; c < a; ++c) A3[sl * LD8 + c] -= f * A3[k * LD8 + c];
        rhs -= f * rhs_k;
      }
      __syncwarp(gm);
    }
    for (int k = a - 1; k >= 0; --k) {
      const double xk = __shfl_sync(gm, rhs, k, W) / A3[k * LD8 + k];
      if (sl == k) rhs = xk;
      if (act && sl < k) rhs -= A3[sl * LD8 + k] * xk;
    }
    const double logdet = log(fabs(prod)) + (double)esum * 0.693147180559945309417232121458 + extra;
    const double quad = s2i * rn2 - gw_sum<W>(gm, act ? w * rhs : 0.0);
    const double prior = gw_sum<W>(gm, act ? __ldg(pil_bar + idx[sl]) : 0.0);
    __syncwarp(gm);
    return -0.5 * (logdet + quad) + prior;
  }

  // Posterior moments of a state with a <= W active latents by W lanes (32 / W states per warp): the statistics pass's
  // counterpart of group_lpj -- same N and v, then the inverse instead of the factorisation.
  template <int W>
  __device__ __forceinline__ void group_moments(const uint32_t *st, double qk, double *xs, double *xsz, double *sum_ss,
                                                double *sum_szsz) const {
    constexpr int G8 = W, LD8 = W + 1;  // (the names of the 8-lane form)
    constexpr size_t GBYTES = 16 + 3 * W * (W + 1) * sizeof(double);
    const int lane = threadIdx.x & 31, sl = lane & (W - 1), gid = lane / W;
    const unsigned gm = ((1u << W) - 1u) << (W * gid);
    unsigned char *base = wsm + (size_t)(threadIdx.x >> 5) * warp_bytes + (size_t)gid * GBYTES;
    uint16_t *idx = reinterpret_cast<uint16_t *>(base);
    double *A1 = reinterpret_cast<double *>(base + 16), *A2 = A1 + G8 * LD8, *A3 = A2 + G8 * LD8;
    // active latents (the caller guarantees |s| <= 8)
    int a = 0;
    for (int w0 = 0; w0 < Hw; w0 += W) {
      uint32_t word = (w0 + sl < Hw) ? st[w0 + sl] : 0u;
      const int c = __popc(word);
      int pre = c;
#pragma unroll
      for (int o = 1; o < W; o <<= 1) { int t = __shfl_up_sync(gm, pre, o, W); if (sl >= o) pre += t; }
      int pos = a + pre - c;
      while (word) {
        int bp = __ffs(word) - 1;
        word &= word - 1;
        if (pos < G8) idx[pos] = (uint16_t)(((w0 + sl) << 5) + bp);
        ++pos;
      }
      a += __shfl_sync(gm, pre, W - 1, W);
    }
    __syncwarp(gm);
    if (a == 0) return;
    const bool act = sl < a;
    // gather G_AA -> A1, Psi_AA -> A2, b_i
    double bi = 0.0;
    if (masked != 1) {
      if (act) {
        const int h = idx[sl];
        const double *Gr = G + (int64_t)h * H;
        if (masked == 2) {  // (W'W)_hl minus the missing observables' share, l <= sl: one walk over the missing list
          double c[W];
#pragma unroll
          for (int l = 0; l < W; ++l) c[l] = 0.0;
          const int nm = *nmiss_ptr;
          const double *wh = Wt + (int64_t)h * D;
          for (int q = 0; q < nm; ++q) {
            const int d = miss[q];
            const double x = __ldg(wh + d);
#pragma unroll
            for (int l = 0; l < W; ++l)
              if (l <= sl) c[l] += x * __ldg(Wt + (int64_t)idx[l] * D + d);
          }
#pragma unroll
          for (int l = 0; l < W; ++l) {
            if (l <= sl) {
              const double v = __ldg(Gr + idx[l]) - c[l];
              A1[sl * LD8 + l] = v;
              A1[l * LD8 + sl] = v;
            }
          }
        } else {
          for (int l = 0; l < a; ++l) A1[sl * LD8 + l] = __ldg(Gr + idx[l]);
        }
        bi = vec[h];
      }
    } else {
      for (int i = 0; i < a; ++i) {
        const double *wi = Wt + (int64_t)idx[i] * D;
        double pb = 0.0;
        for (int d = sl; d < D; d += W) pb += __ldg(wi + d) * vec[d];
        pb = gw_sum<W>(gm, pb);
        if (sl == i) bi = pb;
        for (int j = 0; j <= i; ++j) {
          const double *wj = Wt + (int64_t)idx[j] * D;
          double p = 0.0;
          for (int d = sl; d < D; d += W)
            if (m[d]) p += __ldg(wi + d) * __ldg(wj + d);
          p = gw_sum<W>(gm, p);
          if (sl == 0) { A1[i * LD8 + j] = p; A1[j * LD8 + i] = p; }
        }
      }
    }
    if (act) {
      const double *Pr = Psi + (int64_t)idx[sl] * H;
      for (int l = 0; l < a; ++l) A2[sl * LD8 + l] = __ldg(Pr + idx[l]);
    }
    __syncwarp(gm);
    // v, |r|^2, w = Psi_A' v, N -> A3
    const double mu = act ? __ldg(mus + idx[sl]) : 0.0;
    double t = 0.0;
    for (int l = 0; l < a; ++l) {
      double ml = __shfl_sync(gm, mu, l, W);
      if (act) t += A1[sl * LD8 + l] * ml;
    }
    const double v = act ? s2i * (bi - t) : 0.0;
    const double rn2 = yn2 + gw_sum<W>(gm, act ? mu * (t - 2.0 * bi) : 0.0);
    double w = 0.0;
    for (int l = 0; l < a; ++l) {
      double vl = __shfl_sync(gm, v, l, W);
      if (act) w += A2[l * LD8 + sl] * vl;
    }
    if (act) {
      for (int j = 0; j < a; ++j) {
        double acc = 0.0;
        for (int l = 0; l < a; ++l) acc += A1[sl * LD8 + l] * A2[l * LD8 + j];
        A3[sl * LD8 + j] = (sl == j ? 1.0 : 0.0) + s2i * acc;
      }
    }
    __syncwarp(gm);
    // Gauss-Jordan with partial pivoting on [N | R] (R = A1: G_AA is no longer needed), x <- N^-1 v -- gj_inverse on W lanes
    double x = v;
    double *R = A1;
    if (act)
      for (int c = 0; c < a; ++c) R[sl * LD8 + c] = (sl == c) ? 1.0 : 0.0;
    __syncwarp(gm);
    for (int k = 0; k < a; ++k) {
      unsigned key = 0u;
      if (act && sl >= k) key = (__float_as_uint(fabsf((float)A3[sl * LD8 + k])) & ~(unsigned)(W - 1)) | (unsigned)(W - 1 - sl);
      const int p = W - 1 - (int)(__reduce_max_sync(gm, key) & (unsigned)(W - 1));
      if (p != k) {
        if (act) {
          double tt = A3[k * LD8 + sl]; A3[k * LD8 + sl] = A3[p * LD8 + sl]; A3[p * LD8 + sl] = tt;
          tt = R[k * LD8 + sl]; R[k * LD8 + sl] = R[p * LD8 + sl]; R[p * LD8 + sl] = tt;
        }
        const double xk = __shfl_sync(gm, x, k, W), xp = __shfl_sync(gm, x, p, W);
        if (sl == k) x = xp;
        if (sl == p) x = xk;
      }
      __syncwarp(gm);
      const double piv = A3[k * LD8 + k];
      const double x_k = __shfl_sync(gm, x, k, W);
      if (act && sl != k) {
        const double f = A3[sl * LD8 + k] / piv;
        for (int c = k + 1; c < a; ++c) A3[sl * LD8 + c] -= f * A3[k * LD8 + c];
        for (int c = 0; c < a; ++c) R[sl * LD8 + c] -= f * R[k * LD8 + c];
        x -= f * x_k;
      }
      __syncwarp(gm);
    }
    if (act) {
      const double inv = 1.0 / A3[sl * LD8 + sl];
      for (int c = 0; c < a; ++c) R[sl * LD8 + c] *= inv;
      x *= inv;
    }
    __syncwarp(gm);
    // kappa = Psi_A x + mu ; Lambda = Psi_A N^-1 ; accumulate with weight qk (as the warp path of es3c_suffstats_kernel)
    double kappa = 0.0;
    for (int l = 0; l < a; ++l) {
      const double xl = __shfl_sync(gm, x, l, W);
      if (act) kappa += A2[sl * LD8 + l] * xl;
    }
    kappa += mu;
    const int h = act ? idx[sl] : 0;
    if (act) {
      atomicAdd(&xs[h], qk);
      atomicAdd(&xsz[h], qk * kappa);
    }
    for (int j = 0; j < a; ++j) {
      const double kj = __shfl_sync(gm, kappa, j, W);
      if (act) {
        double lam = 0.0;
        for (int l = 0; l < a; ++l) lam += A2[sl * LD8 + l] * R[l * LD8 + j];
        const int64_t off = (int64_t)h * H + idx[j];
        atomicAdd(sum_szsz + off, qk * (lam + kappa * kj));
        atomicAdd(sum_ss + off, qk);
      }
    }
    __syncwarp(gm);
  }

  // ---- states with at most SMALL_A active latents: one THREAD per state, everything in registers -------------------
  // Same quantities as build() / lu_solve() -- N = I + s2i G_AA Psi_A, v = s2i (b - G mu), |r|^2, w = Psi_A' v,
  // lpj = -1/2 (log|det N| + s2i |r|^2 - w' N^-1 v) + sum pil -- on a fixed SMALL_A x SMALL_A system padded with the
  // identity (fully unrolled LU with partial pivoting, no divergence between states of different size).  Most states
  // of a sparse model are this small, and a warp (or 8 lanes) per such state is mostly idle.
  static constexpr int SMALL_A = 4;
  static constexpr int SMALL_M = 2;  // the same bound for the moments of the statistics pass (3 costs it two resident CTAs)
  __device__ __forceinline__ double small_lpj(const uint32_t *st) const {
    constexpr int MA = SMALL_A;
    int hs[MA];
    int a = 0;
#pragma unroll
    for (int i = 0; i < MA; ++i) hs[i] = 0;
    for (int w = 0; w < Hw; ++w) {
      uint32_t m = st[w];
      while (m) {
        const int h = (w << 5) + __ffs(m) - 1;
        m &= m - 1;
#pragma unroll
        for (int i = 0; i < MA; ++i) if (i == a) hs[i] = h;
        ++a;
      }
    }
    if (a == 0) return -0.5 * s2i * yn2;
    double g[MA][MA], ps[MA][MA], mu[MA], b[MA];
    double prior = 0.0;
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      const bool oi = i < a;
      mu[i] = oi ? __ldg(mus + hs[i]) : 0.0;
      b[i] = oi ? vec[hs[i]] : 0.0;
      prior += oi ? __ldg(pil_bar + hs[i]) : 0.0;
#pragma unroll
      for (int j = 0; j < MA; ++j) {
        const bool on = oi && j < a;
        ps[i][j] = on ? __ldg(Psi + (int64_t)hs[i] * H + hs[j]) : 0.0;
        if (j <= i) {
          const double v = on ? __ldg(G + (int64_t)hs[i] * H + hs[j]) : 0.0;
          g[i][j] = v;
          g[j][i] = v;
        }
      }
    }
    if (masked == 2) {  // minus the missing observables' share of every pair: ONE walk over the list, a loads per entry
      const int nm = *nmiss_ptr;
      double c[MA][MA];
#pragma unroll
      for (int i = 0; i < MA; ++i)
#pragma unroll
        for (int j = 0; j < MA; ++j) c[i][j] = 0.0;
      for (int q = 0; q < nm; ++q) {
        const int d = miss[q];
        double wv[MA];
#pragma unroll
        for (int i = 0; i < MA; ++i) wv[i] = i < a ? __ldg(Wt + (int64_t)hs[i] * D + d) : 0.0;
#pragma unroll
        for (int i = 0; i < MA; ++i)
#pragma unroll
          for (int j = 0; j <= i; ++j) c[i][j] += wv[i] * wv[j];
      }
#pragma unroll
      for (int i = 0; i < MA; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) { g[i][j] -= c[i][j]; if (j < i) g[j][i] -= c[i][j]; }
    }
    double t[MA], v[MA], w[MA], n[MA][MA];
    double rn2 = yn2;
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      double ti = 0.0;
#pragma unroll
      for (int l = 0; l < MA; ++l) ti += g[i][l] * mu[l];
      t[i] = ti;
      v[i] = s2i * (b[i] - ti);
      rn2 += mu[i] * (ti - 2.0 * b[i]);
    }
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      double wi = 0.0;
#pragma unroll
      for (int l = 0; l < MA; ++l) wi += ps[l][i] * v[l];  // (Psi_A' v)_i: Psi need not be symmetric
      w[i] = wi;
#pragma unroll
      for (int j = 0; j < MA; ++j) {
        double acc = 0.0;
#pragma unroll
        for (int l = 0; l < MA; ++l) acc += g[i][l] * ps[l][j];
        n[i][j] = (i == j ? 1.0 : 0.0) + s2i * acc;
      }
    }
    // LU with partial pivoting, right-hand side v (in place); padded rows are unit rows and never chosen over a
    // larger pivot of the live block except by exact ties, which is harmless
    double logdet = 0.0;
#pragma unroll
    for (int k = 0; k < MA; ++k) {
      int p = k;
      double best = fabs(n[k][k]);
#pragma unroll
      for (int i = k + 1; i < MA; ++i) { const double m2 = fabs(n[i][k]); if (m2 > best) { best = m2; p = i; } }
#pragma unroll
      for (int i = k + 1; i < MA; ++i) {
        if (i == p) {
#pragma unroll
          for (int c = 0; c < MA; ++c) { const double x = n[k][c]; n[k][c] = n[i][c]; n[i][c] = x; }
          const double x = v[k]; v[k] = v[i]; v[i] = x;
        }
      }
      const double piv = n[k][k];
      logdet += log(fabs(piv));
#pragma unroll
      for (int i = k + 1; i < MA; ++i) {
        const double f = n[i][k] / piv;
#pragma unroll
        for (int c = k + 1; c < MA; ++c) n[i][c] -= f * n[k][c];
        v[i] -= f * v[k];
      }
    }
#pragma unroll
    for (int k = MA - 1; k >= 0; --k) {
      double x = v[k];
#pragma unroll
      for (int c = k + 1; c < MA; ++c) x -= n[k][c] * v[c];
      v[k] = x / n[k][k];
    }
    double wx = 0.0;
#pragma unroll
    for (int i = 0; i < MA; ++i) wx += w[i] * v[i];
    return -0.5 * (logdet + s2i * rn2 - wx) + prior;
  }

  // Posterior moments of a state with at most SMALL_A active latents by ONE thread (the statistics pass, sssc.py:553-646):
  // x = N^-1 v, kappa = Psi_A x + mu_A, Lambda = Psi_A N^-1 from the same N, v as small_lpj (LU with partial pivoting,
  // unit padding), accumulated with weight qk into the per-datapoint vectors xs / xsz (shared memory) and the global
  // pair sums sum_ss / sum_szsz.  Returns |s|; nothing is done for states larger than SMALL_M.
  __device__ __forceinline__ int small_moments(const uint32_t *st, double qk, double *xs, double *xsz, double *sum_ss,
                                                double *sum_szsz) const {
    constexpr int MA = SMALL_M;
    int hs[MA];
    int a = 0;
#pragma unroll
    for (int i = 0; i < MA; ++i) hs[i] = 0;
    for (int w = 0; w < Hw; ++w) {
      uint32_t m = st[w];
      while (m) {
        const int h = (w << 5) + __ffs(m) - 1;
        m &= m - 1;
#pragma unroll
        for (int i = 0; i < MA; ++i) if (i == a) hs[i] = h;
        ++a;
      }
    }
    if (a > MA || a == 0) return a;  // larger states: the caller's group / warp paths
    double g[MA][MA], ps[MA][MA], mu[MA], b[MA];
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      const bool oi = i < a;
      mu[i] = oi ? __ldg(mus + hs[i]) : 0.0;
      b[i] = oi ? vec[hs[i]] : 0.0;
#pragma unroll
      for (int j = 0; j < MA; ++j) {
        const bool on = oi && j < a;
        ps[i][j] = on ? __ldg(Psi + (int64_t)hs[i] * H + hs[j]) : 0.0;
        if (j <= i) {
          const double v = on ? __ldg(G + (int64_t)hs[i] * H + hs[j]) : 0.0;
          g[i][j] = v;
          g[j][i] = v;
        }
      }
    }
    if (masked == 2) {  // minus the missing observables' share of every pair: ONE walk over the list, a loads per entry
      const int nm = *nmiss_ptr;
      double c[MA][MA];
#pragma unroll
      for (int i = 0; i < MA; ++i)
#pragma unroll
        for (int j = 0; j < MA; ++j) c[i][j] = 0.0;
      for (int q = 0; q < nm; ++q) {
        const int d = miss[q];
        double wv[MA];
#pragma unroll
        for (int i = 0; i < MA; ++i) wv[i] = i < a ? __ldg(Wt + (int64_t)hs[i] * D + d) : 0.0;
#pragma unroll
        for (int i = 0; i < MA; ++i)
#pragma unroll
          for (int j = 0; j <= i; ++j) c[i][j] += wv[i] * wv[j];
      }
#pragma unroll
      for (int i = 0; i < MA; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) { g[i][j] -= c[i][j]; if (j < i) g[j][i] -= c[i][j]; }
    }
    double v[MA], n[MA][MA], r[MA][MA];
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      double ti = 0.0;
#pragma unroll
      for (int l = 0; l < MA; ++l) ti += g[i][l] * mu[l];
      v[i] = s2i * (b[i] - ti);
#pragma unroll
      for (int j = 0; j < MA; ++j) {
        double acc = 0.0;
#pragma unroll
        for (int l = 0; l < MA; ++l) acc += g[i][l] * ps[l][j];
        n[i][j] = (i == j ? 1.0 : 0.0) + s2i * acc;
        r[i][j] = (i == j ? 1.0 : 0.0);
      }
    }
    // Gauss-Jordan with partial pivoting on [N | I | v]
#pragma unroll
    for (int k = 0; k < MA; ++k) {
      int p = k;
      double best = fabs(n[k][k]);
#pragma unroll
      for (int i = k + 1; i < MA; ++i) { const double m2 = fabs(n[i][k]); if (m2 > best) { best = m2; p = i; } }
#pragma unroll
      for (int i = k + 1; i < MA; ++i) {
        if (i == p) {
#pragma unroll
          for (int c = 0; c < MA; ++c) {
            double x = n[k][c]; n[k][c] = n[i][c]; n[i][c] = x;
            x = r[k][c]; r[k][c] = r[i][c]; r[i][c] = x;
          }
          const double x = v[k]; v[k] = v[i]; v[i] = x;
        }
      }
      const double piv = n[k][k];
#pragma unroll
      for (int i = 0; i < MA; ++i) {
        if (i == k) continue;
        const double f = n[i][k] / piv;
#pragma unroll
        for (int c = 0; c < MA; ++c) { n[i][c] -= f * n[k][c]; r[i][c] -= f * r[k][c]; }
        v[i] -= f * v[k];
      }
    }
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      const double inv = 1.0 / n[i][i];
#pragma unroll
      for (int c = 0; c < MA; ++c) r[i][c] *= inv;
      v[i] *= inv;
    }
    double kappa[MA];
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      double kq = mu[i];
#pragma unroll
      for (int l = 0; l < MA; ++l) kq += ps[i][l] * v[l];
      kappa[i] = kq;
    }
#pragma unroll
    for (int i = 0; i < MA; ++i) {
      if (i >= a) continue;
      atomicAdd(&xs[hs[i]], qk);
      atomicAdd(&xsz[hs[i]], qk * kappa[i]);
#pragma unroll
      for (int j = 0; j < MA; ++j) {
        if (j >= a) continue;
        double lam = 0.0;
#pragma unroll
        for (int l = 0; l < MA; ++l) lam += ps[i][l] * r[l][j];
        const int64_t off = (int64_t)hs[i] * H + hs[j];
        atomicAdd(sum_szsz + off, qk * (lam + kappa[i] * kappa[j]));
        atomicAdd(sum_ss + off, qk);
      }
    }
    return a;
  }

  template <int T, bool LIST = false>
  __device__ void eval_batch(const uint32_t *st, int ld, int cnt, double *out, unsigned &flags) {
    if (masked == 1) yn2 = masked_yn2();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // split the batch by |s|: systems with a <= 8 are packed four per warp (without list space every
    // state goes to the warp-per-state solver)
    const bool listed = max_batch > 0 && cnt <= max_batch;
    int *cn = reinterpret_cast<int *>(lsm);
    uint16_t *small = reinterpret_cast<uint16_t *>(lsm + 16), *big = small + max_batch, *tiny = big + max_batch;
    int ns = 0, nb = cnt, nt = 0;
    if (listed) {
      const int tiny_max = masked == 1 ? -1 : SMALL_A;  // the register path reads the Gram tables
      gsync<T>();
      if (threadIdx.x == 0) { cn[0] = 0; cn[1] = 0; cn[2] = 0; }
      gsync<T>();
#pragma unroll 1
      for (int i = threadIdx.x; i < cnt; i += T) {
        const int a = popc_state(st + (size_t)i * ld, Hw);
        if (a <= tiny_max) tiny[atomicAdd(&cn[2], 1)] = (uint16_t)i;
        else if (a <= G8) small[atomicAdd(&cn[0], 1)] = (uint16_t)i;
        else big[atomicAdd(&cn[1], 1)] = (uint16_t)i;
      }
      gsync<T>();
      ns = cn[0], nb = cn[1], nt = cn[2];
    }
#pragma unroll 1
    for (int j = threadIdx.x; j < nt; j += T) {
      const int i = tiny[j];
      out[i] = lpj_sanitise(small_lpj(st + (size_t)i * ld), flags);
    }
#pragma unroll 1
    for (int q = warp; q * 4 < ns; q += T / 32) {
      const int j = q * 4 + (lane >> 3);
      if (j < ns) {  // whole 8-lane groups take or skip this branch together
        const int i = small[j];
        double v = group_lpj<8>(st + (size_t)i * ld);
        v = lpj_sanitise(v, flags);
        if ((lane & 7) == 0) out[i] = v;
      }
    }
    __syncwarp();
#pragma unroll 1
    for (int q = warp; q < nb; q += T / 32) {
      const int i = listed ? (int)big[q] : q;
      double v = warp_lpj(st + (size_t)i * ld);
      v = lpj_sanitise(v, flags);
      if (lane == 0) out[i] = v;
    }
  }
  __device__ double eval_zero(unsigned &flags) {
    if (masked == 1) yn2 = masked_yn2();
    return lpj_sanitise(-0.5 * s2i * yn2, flags);
  }
};

int es3c_launch_estep(const EstepDev &dv, const evo_estep_args *x, cudaStream_t st);

}  // namespace evo
