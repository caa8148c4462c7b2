// Log-pseudo-joint evaluators.  Each evaluator is a small POD passed by value to the kernels with
//   smem_bytes()                      dynamic shared memory it needs per datapoint group
//   setup<T>(n, smem)                 cooperative per-datapoint load (caller syncs afterwards)
//   eval_batch<T>(st, ld, cnt, out, fl)   cooperative evaluation of cnt packed states (row stride ld words) -> out[0..cnt)
//   eval_zero()                       lpj of the all-zero state (permanent state), per thread
#pragma once
#include "common.cuh"

namespace evo {

// ------------------------------------------------------------------------------------------
// BSC, sparse quadratic form (complete data):
//   |y - W s|^2 = |y|^2 - 2 sum_{h in s} b_h + sum_{h,h' in s} G_hh',  b = W'y, G = W'W
// combined in f64 (the three terms cancel catastrophically on well-fitting states).
// Replaces BSC.log_pseudo_joint (evo/models/bsc.py:79-97) for complete data.
// ------------------------------------------------------------------------------------------
struct BscGramEval {
  const double *G;       // [H][H]
  const double *B;       // [N][H]
  const double *ynorm2;  // [N]
  double pre1, pil_bar;
  int H, Hw;
  // per datapoint (set in setup)
  double *b;       // smem [H] (packed-bits forms)
  uint16_t *idx;   // smem [MAXA][T] active-latent lists, one column per thread
  double yn2;
  const double *bg;  // this datapoint's row of B in global memory (the list form reads it through L1/L2)
  const double *theta;  // CUDA-graph replay: pre1 / pil_bar are re-read from theta[0..1] (device memory); NULL = the members
  __device__ __forceinline__ double pre1_now() const { return theta ? theta[0] : pre1; }
  __device__ __forceinline__ double pil_bar_now() const { return theta ? theta[1] : pil_bar; }
  static constexpr int MAXA = 24;
  static constexpr bool kList = true;  // can evaluate index-list rows (estep.cuh LIST representation)
  static constexpr int kThreadsPerSM = 0;  // launch bounds: no minimum number of resident CTAs

  __host__ __device__ size_t smem_bytes(int T, bool list = false) const {
    // list form: no staging at all -- a state touches |s| entries of the row, and the shared memory
    // it would take (8 KB at H = 1024) costs a resident CTA per SM
    return list ? 0 : (size_t)H * sizeof(double) + (size_t)MAXA * T * sizeof(uint16_t);
  }
  __host__ int threads_hint(int work) const { return work <= 32 ? 32 : (work <= 64 ? 64 : 128); }

  template <int T, bool LIST = false>
  __device__ void setup(int64_t n, unsigned char *smem) {
    const double *src = B + n * (int64_t)H;
    bg = src;
    yn2 = __ldg(ynorm2 + n);
    if (LIST) return;
    b = reinterpret_cast<double *>(smem);
    idx = reinterpret_cast<uint16_t *>(smem + (size_t)H * sizeof(double));
    for (int i = threadIdx.x; i < H; i += T) b[i] = __ldg(src + i);
  }

  // pre1 (|y|^2 - 2 sum b + sum G) + pil_bar |s| with the contraction spelled out, so that every
  // evaluation form (bits, scratch lists, list rows) rounds alike
  __device__ __forceinline__ double combine(double sb, double sg, int a) const {
    return __fma_rn(pre1_now(), __dadd_rn(__fma_rn(-2.0, sb, yn2), sg), __dmul_rn(pil_bar_now(), (double)a));
  }

  // generic form: any number of active latents, no scratch
  __device__ __forceinline__ double eval_generic(const uint32_t *st) const {
    double sb = 0.0, sg = 0.0;
    int a = 0;
    for (int w1 = 0; w1 < Hw; ++w1) {
      uint32_t m1 = st[w1];
      while (m1) {
        int b1 = __ffs(m1) - 1;
        m1 &= m1 - 1;
        int h = (w1 << 5) + b1;
        ++a;
        sb += b[h];
        const double *Gr = G + (int64_t)h * H;
        double t = 0.0;
        for (int w2 = 0; w2 <= w1; ++w2) {
          uint32_t m2 = st[w2];
          if (w2 == w1) m2 &= (1u << b1) - 1u;
          while (m2) {
            int b2 = __ffs(m2) - 1;
            m2 &= m2 - 1;
            t += __ldg(Gr + (w2 << 5) + b2);
          }
        }
        sg += __ldg(Gr + h) + 2.0 * t;
      }
    }
    return combine(sb, sg, a);
  }

  // list form: one pass over the words collects the active latents, then a dense a x a triangle
  template <int T>
  __device__ __forceinline__ double eval_one(const uint32_t *st) const {
    uint16_t *my = idx + threadIdx.x;
    int a = 0;
    for (int w = 0; w < Hw; ++w) {
      uint32_t m = st[w];
      while (m) {
        int bp = __ffs(m) - 1;
        m &= m - 1;
        if (a < MAXA) my[a * T] = (uint16_t)((w << 5) + bp);
        ++a;
      }
    }
    if (a > MAXA) return eval_generic(st);
    double sb = 0.0, sg = 0.0;
    for (int i = 0; i < a; ++i) {
      int h = my[i * T];
      sb += b[h];
      const double *Gr = G + (int64_t)h * H;
      double t = 0.0;
#pragma unroll 4
      for (int j = 0; j < i; ++j) t += __ldg(Gr + my[j * T]);
      sg += __ldg(Gr + h) + 2.0 * t;
    }
    return combine(sb, sg, a);
  }

  // index-list row (16 x uint16: ascending active latents, count in e[15]); same summation order as eval_one
  __device__ __forceinline__ double eval_list(const uint32_t *row) const {
    const uint16_t *e = reinterpret_cast<const uint16_t *>(row);
    const int a = e[15];
    double sb = 0.0, sg = 0.0;
    for (int i = 0; i < a; ++i) {
      const int h = e[i];
      sb += __ldg(bg + h);
      const double *Gr = G + (int64_t)h * H;
      double t = 0.0;
#pragma unroll 4
      for (int j = 0; j < i; ++j) t += __ldg(Gr + e[j]);
      sg += __ldg(Gr + h) + 2.0 * t;
    }
    return combine(sb, sg, a);
  }

  template <int T, bool LIST = false>
  __device__ void eval_batch(const uint32_t *st, int ld, int cnt, double *out, unsigned &flags) const {
    for (int i = threadIdx.x; i < cnt; i += T)
      out[i] = lpj_sanitise(LIST ? eval_list(st + (size_t)i * ld) : eval_one<T>(st + (size_t)i * ld), flags);
  }
  __device__ double eval_zero(unsigned &flags) const { return lpj_sanitise(pre1_now() * yn2, flags); }
};

// ------------------------------------------------------------------------------------------
// BSC, explicit masked residual -- the reference's own formula (bsc.py:88-95), f64 throughout.
// Used for incomplete data (the Gram of the observed rows differs per datapoint) and as the
// on-device cross-check of the Gram form.
// ------------------------------------------------------------------------------------------
struct BscDirectEval {
  const double *W;      // [D][H]
  const float *Y;       // [N][D]
  const uint8_t *mask;  // [N][D] or NULL
  double pre1, pil_bar;
  int D, H, Hw;
  double *y;       // smem [D]  (0 where masked)
  uint8_t *m;      // smem [D]
  double yn2;
  const double *theta;  // as in BscGramEval
  __device__ __forceinline__ double pre1_now() const { return theta ? theta[0] : pre1; }
  __device__ __forceinline__ double pil_bar_now() const { return theta ? theta[1] : pil_bar; }

  static constexpr bool kList = false;
  static constexpr int kThreadsPerSM = 0;
  __host__ __device__ size_t smem_bytes(int, bool = false) const { return (size_t)D * sizeof(double) + (((size_t)D + 7) & ~(size_t)7); }
  __host__ int threads_hint(int work) const { return work <= 32 ? 32 : (work <= 64 ? 64 : 128); }

  template <int T>
  __device__ void setup(int64_t n, unsigned char *smem) {
    y = reinterpret_cast<double *>(smem);
    m = smem + (size_t)D * sizeof(double);
    for (int d = threadIdx.x; d < D; d += T) {
      uint8_t ok = mask ? mask[n * (int64_t)D + d] : (uint8_t)1;
      m[d] = ok;
      y[d] = ok ? (double)Y[n * (int64_t)D + d] : 0.0;
    }
  }

  __device__ __forceinline__ double eval_one(const uint32_t *st) const {
    double resid = 0.0;
    for (int d = 0; d < D; ++d) {
      if (!m[d]) continue;
      const double *Wr = W + (int64_t)d * H;
      double mean = 0.0;
      for (int w = 0; w < Hw; ++w) {
        uint32_t mm = st[w];
        while (mm) {
          int bpos = __ffs(mm) - 1;
          mm &= mm - 1;
          mean += __ldg(Wr + (w << 5) + bpos);
        }
      }
      double r = mean - y[d];
      resid += r * r;
    }
    return pre1_now() * resid + pil_bar_now() * (double)popc_state(st, Hw);
  }

  template <int T, bool LIST = false>
  __device__ void eval_batch(const uint32_t *st, int ld, int cnt, double *out, unsigned &flags) const {
    for (int i = threadIdx.x; i < cnt; i += T) out[i] = lpj_sanitise(eval_one(st + (size_t)i * ld), flags);
  }
  __device__ double eval_zero(unsigned &flags) const {
    double s = 0.0;
    for (int d = 0; d < D; ++d) s += y[d] * y[d];
    return lpj_sanitise(pre1_now() * s, flags);
  }
};

// ------------------------------------------------------------------------------------------
// stand-alone lpj (log_pseudo_joint of any evaluator, bsc.py:79-97 / sssc.py:242-326 + lpj_reset_check _models.py:567-596)
// ------------------------------------------------------------------------------------------
template <class Eval, int T>
__global__ void __launch_bounds__(T) lpj_kernel(Eval ev, const uint32_t *__restrict__ states, int shared_states,
                                                int64_t N, int Sp, int Hw, double *__restrict__ out, int64_t ld,
                                                unsigned long long *counters) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int red_i[8];
  unsigned long long c_nan = 0, c_small = 0, c_inf = 0;
  for (int64_t n = blockIdx.x; n < N; n += gridDim.x) {
    __syncthreads();
    ev.template setup<T>(n, smem_raw);
    __syncthreads();
    unsigned fl = 0;
    const uint32_t *st = shared_states ? states : states + n * (int64_t)Sp * Hw;
    ev.template eval_batch<T>(st, Hw, Sp, out + n * ld, fl);
    fl = group_or<T>(fl, red_i);
    if (threadIdx.x == 0 && fl) { if (fl & 1) ++c_nan; else if (fl & 2) ++c_small; else ++c_inf; }
  }
  if (threadIdx.x == 0 && counters) {
    if (c_nan) atomicAdd(counters + 0, c_nan);
    if (c_small) atomicAdd(counters + 1, c_small);
    if (c_inf) atomicAdd(counters + 2, c_inf);
  }
}

// ------------------------------------------------------------------------------------------
// Exact log-likelihood term (free_energy(full=True), _models.py:334-451 over the state table of
// variational/utils.py:58-69): sum_n logsumexp over ALL 2^Hm states, enumerated on the fly from the state index
// (one packed word, Hm <= 24; or_mask = the always-on background bit) instead of materialising the table
// ((2^H, H) bools in the reference, which caps it at H < 12).  One CTA per datapoint, batches of FULL_B states
// through the evaluator's own eval_batch, running (max, sum) per thread.
// ------------------------------------------------------------------------------------------
constexpr int FULL_B = 1024;
template <class Eval, int T>
__global__ void __launch_bounds__(T) full_logsumexp_kernel(Eval ev, int64_t N, int Hm, uint32_t or_mask, double *out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ double red_d[8];
  uint32_t *st = reinterpret_cast<uint32_t *>(smem_raw);
  double *val = reinterpret_cast<double *>(smem_raw + sizeof(uint32_t) * FULL_B);
  unsigned char *ev_smem = smem_raw + (sizeof(uint32_t) + sizeof(double)) * FULL_B;
  const uint32_t total = 1u << Hm;
  double acc = 0.0;
  for (int64_t n = blockIdx.x; n < N; n += gridDim.x) {
    __syncthreads();
    ev.template setup<T>(n, ev_smem);
    __syncthreads();
    double m = -DBL_MAX, ssum = 0.0;
    for (uint32_t base = 0; base < total; base += FULL_B) {
      const int cnt = (int)(total - base < (uint32_t)FULL_B ? total - base : (uint32_t)FULL_B);
      for (int i = threadIdx.x; i < cnt; i += T) st[i] = (base + (uint32_t)i) | or_mask;
      __syncthreads();
      unsigned fl = 0;
      ev.template eval_batch<T>(st, 1, cnt, val, fl);
      __syncthreads();
      for (int i = threadIdx.x; i < cnt; i += T) {
        const double v = val[i];
        if (v > m) { ssum = ssum * exp(m - v) + 1.0; m = v; } else { ssum += exp(v - m); }
      }
      __syncthreads();
    }
    const double gm = group_max<T>(m, red_d);
    const double gs = group_sum<T>(ssum > 0.0 ? ssum * exp(m - gm) : 0.0, red_d);
    if (threadIdx.x == 0) acc += gm + log(gs);
  }
  if (threadIdx.x == 0 && acc != 0.0) atomicAdd(out, acc);
}

}  // namespace evo
