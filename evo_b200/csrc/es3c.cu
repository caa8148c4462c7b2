// ES3C kernels: E-step launcher (evaluator in es3c.cuh), posterior moments / sufficient statistics
// and data estimator of SSSC.EM_step (evo/models/sssc.py:553-646).
#include "es3c.cuh"

namespace evo {

// global workspaces for systems with 17..32 active latents: one registration PER DEVICE (a process may
// hold models on several GPUs); looked up by the current device at launch time
constexpr int MAX_DEV = 64;
static double *g_scratch_tab[MAX_DEV];
static int g_scratch_warps_tab[MAX_DEV];
static void current_scratch(double *&scratch, int &warps) {
  int dev = 0;
  scratch = nullptr; warps = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) return;
  scratch = __atomic_load_n(&g_scratch_tab[dev], __ATOMIC_ACQUIRE);
  warps = scratch ? __atomic_load_n(&g_scratch_warps_tab[dev], __ATOMIC_ACQUIRE) : 0;
}

int es3c_launch_estep(const EstepDev &dv, const evo_estep_args *x, cudaStream_t st) {
  EVO_REQUIRE(x->Psi && x->mus && x->pil_bar_h, "es3c evaluator needs Psi, mus, pil_bar_h");
  EVO_REQUIRE(x->sigma2 > 0.0, "sigma2 must be positive");
  Es3cEval ev;
  memset(&ev, 0, sizeof(ev));
  ev.budget_estep();
  ev.Psi = x->Psi; ev.mus = x->mus; ev.pil_bar = x->pil_bar_h;
  ev.s2i = (double)(1.0L / (long double)x->sigma2);
  ev.D = dv.D; ev.H = dv.H; ev.Hw = dv.Hw;
  current_scratch(ev.scratch, ev.scratch_warps);
  if (x->mask) {
    EVO_REQUIRE(x->W && x->Y, "es3c with incomplete data needs Wt (in W) and Y");
    ev.masked = 1; ev.Wt = x->W; ev.Y = x->Y; ev.mask = x->mask;
    if (x->G && x->B && x->ynorm2 && dv.D <= 65535) { ev.masked = 2; ev.G = x->G; ev.B = x->B; ev.ynorm2 = x->ynorm2; }
  } else {
    EVO_REQUIRE(x->G && x->B && x->ynorm2, "es3c with complete data needs G, B, ynorm2");
    ev.masked = 0; ev.G = x->G; ev.B = x->B; ev.ynorm2 = x->ynorm2;
  }
  ev.max_batch = dv.S > dv.Cg ? dv.S : dv.Cg;
  return launch_estep_any(dv, ev, st);
}

// ------------------------------------------------------------------------------------------
// Gauss-Jordan with partial pivoting on [A3 | R] (lane = row): R <- A3^-1 R for an a x a right-hand
// side R that starts as the identity, and x <- A3^-1 x for the vector carried in registers.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void gj_inverse(int a, double *A3, double *R, int ld, double &x) {
  const int lane = threadIdx.x & 31;
  const bool act = lane < a;
  if (act)
    for (int c = 0; c < a; ++c) R[lane * ld + c] = (lane == c) ? 1.0 : 0.0;
  __syncwarp();
  for (int k = 0; k < a; ++k) {
    unsigned key = 0u;
    if (act && lane >= k) key = (__float_as_uint(fabsf((float)A3[lane * ld + k])) & ~31u) | (unsigned)(31 - lane);
    const int p = 31 - (int)(__reduce_max_sync(0xffffffffu, key) & 31u);
    if (p != k) {
      if (act) {
        double t = A3[k * ld + lane]; A3[k * ld + lane] = A3[p * ld + lane]; A3[p * ld + lane] = t;
        t = R[k * ld + lane]; R[k * ld + lane] = R[p * ld + lane]; R[p * ld + lane] = t;
      }
      double xk = __shfl_sync(0xffffffffu, x, k), xp = __shfl_sync(0xffffffffu, x, p);
      if (lane == k) x = xp;
      if (lane == p) x = xk;
    }
    __syncwarp();
    double piv = A3[k * ld + k];
    double x_k = __shfl_sync(0xffffffffu, x, k);
    if (act && lane != k) {  // eliminate column k from every other row
      double f = A3[lane * ld + k] / piv;
      for (int c = k + 1; c < a; ++c) A3[lane * ld + c] -= f * A3[k * ld + c];
      for (int c = 0; c < a; ++c) R[lane * ld + c] -= f * R[k * ld + c];
      x -= f * x_k;
    }
    __syncwarp();
  }
  if (act) {
    double inv = 1.0 / A3[lane * ld + lane];
    for (int c = 0; c < a; ++c) R[lane * ld + c] *= inv;
    x *= inv;
  }
  __syncwarp();
}


// Posterior moments of one state by one warp (lane = row), accumulated with weight qk.  cta > 0: the warp owns the areas of
// all warps of the CTA for this call (Es3cEval::workspace_cta).
__device__ __forceinline__ void warp_state_moments(const Es3cEval &ev, const uint32_t *st, double qk, int cta, int H, double *xs,
                                                   double *xsz, double *sum_ss, double *sum_szsz) {
  const int lane = threadIdx.x & 31;
  {
      uint16_t *idx;
      double *A1, *A2, *A3;
      int ld;
      if (cta > 0) ev.workspace_cta(cta, idx, A1, A2, A3, ld); else ev.workspace(0, idx, A1, A2, A3, ld);
      int na = ev.active_list(st, idx);
      if (na == 0) return;
      if (!(cta > 0 && na <= cta) && !ev.workspace(na, idx, A1, A2, A3, ld)) return;  // rejected state (lowest lpj, zero weight)
      double ynorm;
      double bi = ev.gather(na, idx, A1, A2, ld, ynorm);
      double v, rn2, w, mu;
      ev.build(na, idx, A1, A2, A3, ld, bi, ynorm, v, rn2, w, mu);
      double x = v;
      gj_inverse(na, A3, A1, ld, x);  // A1 <- N^-1, x <- N^-1 v
      const bool on = lane < na;
      // kappa = Psi_A x + mu ; Lambda = Psi_A N^-1
      double kappa = 0.0;
      for (int l = 0; l < na; ++l) {
        double xl = __shfl_sync(0xffffffffu, x, l);
        if (on) kappa += A2[lane * ld + l] * xl;
      }
      kappa += mu;
      int h = on ? idx[lane] : 0;
      if (on) {
        atomicAdd(&xs[h], qk);
        atomicAdd(&xsz[h], qk * kappa);
      }
      for (int j = 0; j < na; ++j) {
        double kj = __shfl_sync(0xffffffffu, kappa, j);
        if (on) {
          double lam = 0.0;
          for (int l = 0; l < na; ++l) lam += A2[lane * ld + l] * A1[l * ld + j];
          int64_t off = (int64_t)h * H + idx[j];
          atomicAdd(sum_szsz + off, qk * (lam + kappa * kj));
          atomicAdd(sum_ss + off, qk);
        }
      }
      __syncwarp();
      }
}

template <int T>
__global__ void __launch_bounds__(T) es3c_suffstats_kernel(evo_es3c_stats_args a, Es3cEval ev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ double red_d[8];
  __shared__ int red_i[8];
  __shared__ int n_act, n_g8, n_big, n_huge;
  const int H = a.H, D = a.D, S = a.S, Hw = ev.Hw, cols = a.S + a.S_perm;
  const int64_t H2 = (int64_t)H * H;
  // smem: xs[H] xsz[H] q[cols] ybuf[D] act[H] | evaluator area
  double *xs = reinterpret_cast<double *>(smem_raw);
  double *xsz = xs + H;
  double *q = xsz + H;
  double *ybuf = q + cols;
  int *act = reinterpret_cast<int *>(ybuf + D);
  uint16_t *g8list = reinterpret_cast<uint16_t *>(act + H + (H & 1));  // live states with 3..8 active latents
  uint16_t *biglist = g8list + ((S + 3) & ~3);                          // ... with up to as_cap (or beyond the CTA area)
  uint16_t *hugelist = biglist + ((S + 3) & ~3);                        // ... in between: the CTA-wide shared workspace
  unsigned char *ev_smem = reinterpret_cast<unsigned char *>(hugelist + ((S + 3) & ~3));
  const int ccap = ev.cta_cap(T / 32);
  double *sum_s = a.stats, *sum_sz = sum_s + H, *sum_ss = sum_sz + H, *sum_szsz = sum_ss + H2;
  double *sum_s_sz = sum_szsz + H2, *sum_outer = sum_s_sz + H2, *WpT = sum_outer + H2, *trace = WpT + (int64_t)H * D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double trace_acc = 0.0;
  for (int64_t n = blockIdx.x; n < a.N; n += gridDim.x) {
    __syncthreads();
    for (int h = threadIdx.x; h < H; h += T) { xs[h] = 0.0; xsz[h] = 0.0; }
    ev.template setup<T>(n, ev_smem);
    // posterior over all S + S_perm columns (the tiny eps of sssc.py:602 is below f64 resolution here)
    {
      const double *row = a.lpj + n * cols;
      double mx = -DBL_MAX;
      for (int i = threadIdx.x; i < cols; i += T) { double v = row[i]; q[i] = v; mx = fmax(mx, v); }
      mx = group_max<T>(mx, red_d);
      double se = 0.0;
      for (int i = threadIdx.x; i < cols; i += T) { double e = exp(q[i] - mx); q[i] = e; se += e; }
      se = group_sum<T>(se, red_d);
      double inv = 1.0 / (se + DBL_MIN);
      for (int i = threadIdx.x; i < cols; i += T) q[i] *= inv;
    }
    __syncthreads();
    if (ev.masked == 1) ev.yn2 = ev.masked_yn2();
    const uint32_t *stn = a.states + n * (int64_t)S * Hw;
    // ---- per-state moments: states with <= SMALL_M active latents (most of a sparse model) by one THREAD each, ... -----
    if (threadIdx.x == 0) { n_g8 = 0; n_big = 0; n_huge = 0; }
    __syncthreads();
    for (int k = threadIdx.x; k < S; k += T) {
      const double qk = q[a.S_perm + k];
      if (!(qk > a.q_skip)) continue;
      int na = ev.masked != 1 ? ev.small_moments(stn + (int64_t)k * Hw, qk, xs, xsz, sum_ss, sum_szsz)
                              : popc_state(stn + (int64_t)k * Hw, Hw);
      if (na == 0 || (ev.masked != 1 && na <= Es3cEval::SMALL_M)) continue;
      if (na <= Es3cEval::G8) g8list[atomicAdd(&n_g8, 1)] = (uint16_t)k;
      else if (na <= ev.as_cap || na > ccap) biglist[atomicAdd(&n_big, 1)] = (uint16_t)k;
      else hugelist[atomicAdd(&n_huge, 1)] = (uint16_t)k;
    }
    __syncthreads();
    // ---- ... up to 8 latents by 8 lanes each (four states per warp) ... ---------------------------------------------------
    for (int g = warp; g * 4 < n_g8; g += T / 32) {
      const int j = g * 4 + (lane >> 3);
      if (j < n_g8) {  // whole 8-lane groups take or skip this branch together
        const int k = g8list[j];
        ev.template group_moments<8>(stn + (int64_t)k * Hw, q[a.S_perm + k], xs, xsz, sum_ss, sum_szsz);
      }
    }
    __syncwarp();
    // ---- ... the others by one warp each -------------------------------------------------------------------------------
    for (int b = warp; b < n_big; b += T / 32) {
      const int k = biglist[b];
      warp_state_moments(ev, stn + (int64_t)k * Hw, q[a.S_perm + k], 0, H, xs, xsz, sum_ss, sum_szsz);
    }
    if (n_huge > 0) {  // as_cap < |s| <= ccap: one state at a time in the CTA's whole workspace, by warp 0
      __syncthreads();
      if (warp == 0)
        for (int b = 0; b < n_huge; ++b) {
          const int k = hugelist[b];
          warp_state_moments(ev, stn + (int64_t)k * Hw, q[a.S_perm + k], ccap, H, xs, xsz, sum_ss, sum_szsz);
        }
    }
    __syncthreads();
    // ---- per-datapoint vectors -> global sums -----------------------------------------------------
    if (threadIdx.x == 0) n_act = 0;
    __syncthreads();
    for (int base = 0; base < H; base += T) {
      int h = base + threadIdx.x;
      bool f = h < H && xs[h] != 0.0;
      int tot;
      int off = group_scan_flag<T>(f, tot, red_i);
      if (f) act[n_act + off] = h;
      __syncthreads();
      if (threadIdx.x == 0) n_act += tot;
      __syncthreads();
    }
    const int na = n_act;
    for (int i = threadIdx.x; i < na; i += T) {
      int h = act[i];
      atomicAdd(sum_s + h, xs[h]);
      atomicAdd(sum_sz + h, xsz[h]);
    }
    for (int i = warp; i < na; i += T / 32) {  // rows over warps, columns over lanes: no integer division per pair
      const int h = act[i];
      const double xh = xs[h], zh = xsz[h];
      for (int j = lane; j < na; j += 32) {
        const int h2 = act[j];
        if (a.learn_psi) atomicAdd(sum_s_sz + (int64_t)h * H + h2, xh * xsz[h2]);
        if (!ev.masked) atomicAdd(sum_outer + (int64_t)h * H + h2, zh * xsz[h2]);
      }
    }
    // estimator, masked trace, W numerator (threads over d, Wt rows are contiguous in d)
    for (int d = threadIdx.x; d < D; d += T) {
      float yraw = a.Y[n * (int64_t)D + d];
      double est = 0.0;
      bool need_est = ev.masked || a.Yrec != nullptr;
      if (need_est)
        for (int i = 0; i < na; ++i) est += __ldg(a.Wt + (int64_t)act[i] * D + d) * xsz[act[i]];
      bool obs = !a.mask || a.mask[n * (int64_t)D + d];
      if (ev.masked && obs) trace_acc += est * est;
      double yuse = (double)yraw;
      if (a.Yrec) {
        bool keep = a.recon_keep && a.recon_keep[n * (int64_t)D + d];
        float yr = keep ? yraw : (float)est;
        a.Yrec[n * (int64_t)D + d] = yr;
        if (ev.masked) yuse = (double)yr;
      }
      ybuf[d] = yuse;
    }
    __syncthreads();
    {  // all (latent, observable) pairs over all threads (D need not divide T); (i, d) advanced without a division
      int i = threadIdx.x / D, d = threadIdx.x - i * D;
      while (i < na) {
        const int h = act[i];
        atomicAdd(WpT + (int64_t)h * D + d, xsz[h] * ybuf[d]);
        d += T;
        while (d >= D) { d -= D; ++i; }
      }
    }
  }
  trace_acc = group_sum<T>(trace_acc, red_d);
  if (threadIdx.x == 0 && trace_acc != 0.0) atomicAdd(trace, trace_acc);
}

}  // namespace evo

using namespace evo;

extern "C" int evo_es3c_set_scratch(double *scratch, int warps) {
  int dev = 0;
  EVO_CUDA(cudaGetDevice(&dev));
  EVO_REQUIRE(dev >= 0 && dev < MAX_DEV, "device index out of range");
  __atomic_store_n(&g_scratch_warps_tab[dev], scratch ? warps : 0, __ATOMIC_RELEASE);
  __atomic_store_n(&g_scratch_tab[dev], scratch, __ATOMIC_RELEASE);
  return EVO_OK;
}

extern "C" int evo_es3c_suffstats(const evo_es3c_stats_args *x, void *stream) {
  EVO_REQUIRE(x != nullptr, "null args");
  EVO_REQUIRE(x->N >= 0 && x->D > 0 && x->H > 0 && x->S > 0, "bad sizes");
  EVO_REQUIRE(x->states && x->lpj && x->Y && x->Wt && x->Psi && x->mus && x->stats, "null table");
  EVO_REQUIRE(x->sigma2 > 0.0, "sigma2 must be positive");
  if (x->N == 0) return EVO_OK;
  Es3cEval ev;
  memset(&ev, 0, sizeof(ev));
  ev.budget_stats();
  ev.Psi = x->Psi; ev.mus = x->mus; ev.pil_bar = x->mus;  // pil_bar is not used by the moments
  ev.s2i = (double)(1.0L / (long double)x->sigma2);
  ev.D = x->D; ev.H = x->H; ev.Hw = words_for(x->H);
  current_scratch(ev.scratch, ev.scratch_warps);
  ev.Wt = x->Wt; ev.Y = x->Y;
  if (x->mask) {
    ev.masked = 1; ev.mask = x->mask;
    if (x->G && x->B && x->ynorm2 && x->D <= 65535) { ev.masked = 2; ev.G = x->G; ev.B = x->B; ev.ynorm2 = x->ynorm2; }
  } else {
    EVO_REQUIRE(x->G && x->B && x->ynorm2, "complete data needs G, B, ynorm2");
    ev.masked = 0; ev.G = x->G; ev.B = x->B; ev.ynorm2 = x->ynorm2;
  }
  constexpr int T = 128;
  int H = x->H, cols = x->S + x->S_perm;
  size_t smem = sizeof(double) * ((size_t)2 * H + cols + x->D) + sizeof(int) * (size_t)(H + (H & 1)) +
                3 * sizeof(uint16_t) * (size_t)((x->S + 3) & ~3) + ev.smem_bytes(T);
  auto k = es3c_suffstats_kernel<T>;
  if (smem > 227 * 1024) {
    set_error("evo_es3c_suffstats: %zu bytes of shared memory exceed 227 KB", smem);
    return EVO_ERR_UNSUPPORTED;
  }
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  EVO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, T, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (int64_t)sm_count() * per_sm;
  if (g > x->N) g = x->N;
  k<<<(int)g, T, smem, (cudaStream_t)stream>>>(*x, ev);
  EVO_LAUNCH_CHECK("es3c_suffstats");
  return EVO_OK;
}

extern "C" int evo_es3c_lpj(const uint32_t *states, int shared_states, int64_t N, int Sp, int D, int H,
                            const double *Wt, const float *Y, const uint8_t *mask, const double *G, const double *B,
                            const double *ynorm2, const double *Psi, const double *mus, const double *pil_bar_h,
                            double sigma2, double *lpj_out, int64_t ld_lpj, unsigned long long *reset_counters,
                            void *stream) {
  EVO_REQUIRE(N >= 0 && Sp > 0 && D > 0 && H > 0 && ld_lpj >= Sp, "bad sizes");
  EVO_REQUIRE(Psi && mus && pil_bar_h && sigma2 > 0.0, "model tables");
  if (N == 0) return EVO_OK;
  Es3cEval ev;
  memset(&ev, 0, sizeof(ev));
  ev.budget_estep();
  ev.Psi = Psi; ev.mus = mus; ev.pil_bar = pil_bar_h;
  ev.s2i = (double)(1.0L / (long double)sigma2);
  ev.D = D; ev.H = H; ev.Hw = words_for(H);
  current_scratch(ev.scratch, ev.scratch_warps);
  if (mask) {
    EVO_REQUIRE(Wt && Y, "incomplete data needs Wt and Y");
    ev.masked = 1; ev.Wt = Wt; ev.Y = Y; ev.mask = mask;
    if (G && B && ynorm2 && D <= 65535) { ev.masked = 2; ev.G = G; ev.B = B; ev.ynorm2 = ynorm2; }
  } else {
    EVO_REQUIRE(G && B && ynorm2, "complete data needs G, B, ynorm2");
    ev.masked = 0; ev.G = G; ev.B = B; ev.ynorm2 = ynorm2;
  }
  constexpr int T = 128;
  ev.max_batch = Sp <= 16384 ? Sp : 0;
  size_t smem = ev.smem_bytes(T);
  auto k = lpj_kernel<Es3cEval, T>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  EVO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, T, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (int64_t)sm_count() * per_sm;
  if (g > N) g = N;
  k<<<(int)g, T, smem, (cudaStream_t)stream>>>(ev, states, shared_states, N, Sp, ev.Hw, lpj_out, ld_lpj, reset_counters);
  EVO_LAUNCH_CHECK("es3c_lpj");
  return EVO_OK;
}

extern "C" int evo_es3c_full_logsumexp(int64_t N, int D, int H, int background, const double *Wt, const float *Y,
                                       const uint8_t *mask, const double *G, const double *B, const double *ynorm2,
                                       const double *Psi, const double *mus, const double *pil_bar_h, double sigma2,
                                       double *out, void *stream) {
  EVO_REQUIRE(N >= 0 && D > 0 && H > 0 && out, "bad arguments");
  EVO_REQUIRE(Psi && mus && pil_bar_h && sigma2 > 0.0, "model tables");
  const int Hm = background ? H - 1 : H;
  EVO_REQUIRE(H <= 32 && Hm <= 24, "exact likelihood: at most 24 enumerated latents (2^24 states per datapoint)");
  if (N == 0) return EVO_OK;
  Es3cEval ev;
  memset(&ev, 0, sizeof(ev));
  ev.budget_estep();
  ev.Psi = Psi; ev.mus = mus; ev.pil_bar = pil_bar_h;
  ev.s2i = (double)(1.0L / (long double)sigma2);
  ev.D = D; ev.H = H; ev.Hw = 1;
  current_scratch(ev.scratch, ev.scratch_warps);
  if (mask) {
    EVO_REQUIRE(Wt && Y, "incomplete data needs Wt and Y");
    ev.masked = 1; ev.Wt = Wt; ev.Y = Y; ev.mask = mask;
    if (G && B && ynorm2 && D <= 65535) { ev.masked = 2; ev.G = G; ev.B = B; ev.ynorm2 = ynorm2; }
  } else {
    EVO_REQUIRE(G && B && ynorm2, "complete data needs G, B, ynorm2");
    ev.masked = 0; ev.G = G; ev.B = B; ev.ynorm2 = ynorm2;
  }
  ev.max_batch = FULL_B;
  constexpr int T = 128;
  const size_t smem = (sizeof(uint32_t) + sizeof(double)) * FULL_B + ev.smem_bytes(T);
  auto k = full_logsumexp_kernel<Es3cEval, T>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  EVO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, T, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (int64_t)sm_count() * per_sm;
  if (g > N) g = N;
  const uint32_t or_mask = background ? (1u << (H - 1)) : 0u;
  k<<<(int)g, T, smem, (cudaStream_t)stream>>>(ev, N, Hm, or_mask, out);
  EVO_LAUNCH_CHECK("es3c_full_logsumexp");
  return EVO_OK;
}
