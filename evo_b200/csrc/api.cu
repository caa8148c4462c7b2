// C-ABI glue: error reporting, E-step launcher, stand-alone top-K replacement.
#include <stdarg.h>

#include "common.cuh"
#include "estep.cuh"
#include "estep_flip.cuh"
#include "es3c.cuh"

namespace evo {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int check_cuda(cudaError_t e, const char *what) {
  if (e == cudaSuccess) return EVO_OK;
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return EVO_ERR_CUDA;
}
static unsigned long long g_launches = 0;
void count_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); }
unsigned long long launches() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }
int sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ------------------------------------------------------------------------------------------
// vary_Kn alone (evo/variational/utils.py:231-337).  One CTA per datapoint; new candidates that
// equal a permanent state, a member of K^n or an earlier candidate are dropped; ties between
// equal lpj are broken lexicographically (latent 0 most significant), which is the order the
// reference's np.unique + stable sort produce.
// ------------------------------------------------------------------------------------------
template <int T>
__global__ void __launch_bounds__(T) select_topk_kernel(uint32_t *__restrict__ states, double *__restrict__ lpj,
                                                        int64_t N, int S, int S_perm, int Hw, int Mprime,
                                                        const uint32_t *__restrict__ new_states,
                                                        const double *__restrict__ lpj_new,
                                                        const int32_t *__restrict__ n_new, int U,
                                                        int32_t *__restrict__ out_counts) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ double red_d[8];
  __shared__ int red_i[8];
  unsigned char *dead = smem_raw;       // [U] 1 = duplicate / taken
  unsigned char *taken_old = dead + U;  // [S]
  for (int64_t n = blockIdx.x; n < N; n += gridDim.x) {
    __syncthreads();
    uint32_t *old = states + n * (int64_t)S * Hw;
    double *lrow = lpj + n * (int64_t)(S + S_perm) + S_perm;
    const uint32_t *nw = new_states + n * (int64_t)U * Hw;
    const double *lnew = lpj_new + n * (int64_t)U;
    int un = n_new ? n_new[n] : U;
    int uniq = 0;
    for (int i = threadIdx.x; i < S; i += T) taken_old[i] = 0;
    for (int i = threadIdx.x; i < un; i += T) {
      const uint32_t *si = nw + (int64_t)i * Hw;
      bool dup = (S_perm > 0 && popc_state(si, Hw) == 0);
      for (int k = 0; k < S && !dup; ++k) dup = same_state(old + (int64_t)k * Hw, si, Hw);
      for (int j = 0; j < i && !dup; ++j) dup = same_state(nw + (int64_t)j * Hw, si, Hw);
      dead[i] = dup;
      uniq += !dup;
    }
    uniq = group_sum_i<T>(uniq, red_i);
    __syncthreads();
    int Mp = uniq < Mprime ? uniq : Mprime;
    int n_sub = 0;
    for (int j = 0; j < Mp; ++j) {
      // best remaining new (ties: lexicographically smaller state first)
      KeyIdx bn; bn.key = 0; bn.idx = -1;
      for (int i = threadIdx.x; i < un; i += T) {
        if (dead[i]) continue;
        double v = lnew[i];
        if (bn.idx < 0 || v > bn.key ||
            (v == bn.key && lex_less(nw + (int64_t)i * Hw, nw + (int64_t)bn.idx * Hw, Hw))) { bn.key = v; bn.idx = i; }
      }
      // group reduction with the same tie rule
      for (int o = 16; o > 0; o >>= 1) {
        KeyIdx ot; ot.key = __shfl_xor_sync(0xffffffffu, bn.key, o); ot.idx = __shfl_xor_sync(0xffffffffu, bn.idx, o);
        if (ot.idx >= 0 && (bn.idx < 0 || ot.key > bn.key ||
                            (ot.key == bn.key && ot.idx != bn.idx &&
                             lex_less(nw + (int64_t)ot.idx * Hw, nw + (int64_t)bn.idx * Hw, Hw)))) bn = ot;
      }
      if (T > 32) {
        __shared__ double sk[T / 32];
        __shared__ int si[T / 32];
        __syncthreads();
        if ((threadIdx.x & 31) == 0) { sk[threadIdx.x >> 5] = bn.key; si[threadIdx.x >> 5] = bn.idx; }
        __syncthreads();
        bn.key = sk[0]; bn.idx = si[0];
        for (int w = 1; w < T / 32; ++w) {
          if (si[w] >= 0 && (bn.idx < 0 || sk[w] > bn.key ||
                             (sk[w] == bn.key && lex_less(nw + (int64_t)si[w] * Hw, nw + (int64_t)bn.idx * Hw, Hw)))) {
            bn.key = sk[w]; bn.idx = si[w];
          }
        }
      }
      KeyIdx wo; wo.key = 0; wo.idx = -1;
      for (int i = threadIdx.x; i < S; i += T)
        if (!taken_old[i]) { KeyIdx c; c.key = -lrow[i]; c.idx = i; wo = better(wo, c); }
      wo = group_argmax<T>(wo, red_d, red_i);
      if (bn.idx < 0 || wo.idx < 0 || !(bn.key > -wo.key)) break;
      __syncthreads();
      for (int w = threadIdx.x; w < Hw; w += T) old[(int64_t)wo.idx * Hw + w] = nw[(int64_t)bn.idx * Hw + w];
      if (threadIdx.x == 0) { lrow[wo.idx] = bn.key; taken_old[wo.idx] = 1; dead[bn.idx] = 1; }
      ++n_sub;
      __syncthreads();
    }
    if (threadIdx.x == 0 && out_counts) { out_counts[2 * n] = uniq; out_counts[2 * n + 1] = n_sub; }
  }
}

}  // namespace evo

using namespace evo;

extern "C" const char *evo_last_error(void) { return g_err; }
extern "C" int evo_version(void) { return 100; }
extern "C" int evo_device_sm_count(void) { return sm_count(); }
extern "C" unsigned long long evo_launch_count(void) { return launches(); }

// fills the derived sizes of an E-step launch (children capacity, hash table, pool) from the caller's arguments
static int estep_dims(const evo_estep_args *x, EstepDev &dv);

static int estep_impl(const evo_estep_args *x, const evo_evolve_args *t, void *stream) {
  EVO_REQUIRE(x != nullptr, "null args");
  EVO_REQUIRE(x->N >= 0 && x->H > 0 && x->S > 0 && x->D > 0, "bad sizes");
  EVO_REQUIRE(x->S_perm == 0 || x->S_perm == 1, "S_perm must be 0 or 1 (only the all-zero state is implemented)");
  EVO_REQUIRE(!(x->background && x->S_perm), "background unit and permanent all-zero state are exclusive");
  EVO_REQUIRE(x->n_generations >= 0 && x->n_generations < 256, "bad n_generations");
  EVO_REQUIRE(x->states && x->lpj, "states/lpj null");
  EVO_REQUIRE(x->selection == EVO_SEL_FIT || x->selection == EVO_SEL_RAND || (t && x->selection == EVO_SEL_ALL),
              "bad selection");
  EVO_REQUIRE(x->H <= 65535 && x->S <= 65535, "H and S must not exceed 65535 (latent / state indices are 16 bit)");
  EVO_REQUIRE(x->mutation >= EVO_MUT_RANDFLIP && x->mutation <= EVO_MUT_CROSS_SPARSEFLIP, "bad mutation");
  if (x->N == 0) return EVO_OK;
  EstepDev dv;
  memset(&dv, 0, sizeof(dv));
  dv.N = x->N; dv.n_offset = x->n_offset;
  dv.D = x->D; dv.H = x->H; dv.Hw = words_for(x->H); dv.S = x->S; dv.S_perm = x->S_perm;
  dv.Mprime = x->Mprime > 0 ? x->Mprime : x->S;
  dv.n_parents = x->n_parents; dv.n_children = x->n_children; dv.n_generations = x->n_generations;
  dv.selection = x->selection; dv.mutation = x->mutation; dv.background = x->background;
  dv.p_bf = x->bitflip_prob; dv.sparseness = x->sparseness;
  dv.seed = x->seed; dv.iteration = x->iteration;
  dv.states = x->states; dv.lpj = x->lpj;
  dv.inj_children = x->inj_children; dv.inj_count = x->inj_count; dv.inj_cmax = x->inj_cmax;
  dv.acc = x->acc;
  {
    int rc = estep_dims(x, dv);
    if (rc) return rc;
  }
  dv.theta_dev = x->theta_dev;
  EVO_REQUIRE(!x->theta_dev || x->eval == EVO_EVAL_BSC_GRAM || x->eval == EVO_EVAL_BSC_DIRECT,
              "theta_dev is for the BSC evaluators");
  if (x->fs_rest) {
    EVO_REQUIRE(x->fs_stats && x->fs_wp_workspace, "fused statistics need fs_stats, fs_wp_workspace and fs_rest");
    EVO_REQUIRE(x->N < 2147483647LL, "fused statistics: N must fit int32");
    WpLists pl{};
    wp_tc_planes(x->fs_wp_workspace, x->N, x->D, x->H, &pl);
    dv.fs_Wq = x->fs_stats + (int64_t)x->H * x->D;
    dv.fs_pies = dv.fs_Wq + (int64_t)x->H * x->H;
    dv.fs_sig = dv.fs_pies + x->H;
    dv.fs_hi = pl.et_hi; dv.fs_lo = pl.et_lo; dv.fs_ld = pl.et_ld;
    dv.fs_rest = x->fs_rest;
  }
  if (t) {
    dv.generic_only = 1;
    dv.evolve_only = t->evolve_only; dv.lpj_given = t->lpj_given; dv.skip_eval = t->skip_eval; dv.gen0 = t->gen0;
    dv.pool0 = t->pool0; dv.pool0_n = t->pool0_n;
    dv.tr_parents = t->tr_parents; dv.tr_npar = t->tr_npar; dv.tr_children = t->tr_children; dv.tr_count = t->tr_count;
    dv.tr_pos = t->tr_pos;
    dv.ev_states = t->ev_states; dv.ev_lpj = t->ev_lpj; dv.ev_count = t->ev_count; dv.ev_pool = t->ev_pool;
    dv.ev_pool_n = t->ev_pool_n;
    EVO_REQUIRE(!t->pool0 || t->pool0_n, "pool0 needs pool0_n");
    EVO_REQUIRE(!t->tr_parents || t->tr_npar, "tr_parents needs tr_npar");
    EVO_REQUIRE(!t->tr_children || t->tr_count, "tr_children needs tr_count");
    EVO_REQUIRE(!t->evolve_only || (t->ev_states && t->ev_lpj && t->ev_count), "evolve_only needs ev_states, ev_lpj, ev_count");
    EVO_REQUIRE(!t->ev_pool || t->ev_pool_n, "ev_pool needs ev_pool_n");
    EVO_REQUIRE(t->gen0 >= 0 && t->gen0 + x->n_generations < 256, "bad gen0");
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (x->eval == EVO_EVAL_BSC_GRAM) {
    EVO_REQUIRE(x->G && x->B && x->ynorm2, "gram evaluator needs G, B, ynorm2");
    EVO_REQUIRE(x->mask == nullptr, "gram evaluator needs complete data");
    BscGramEval ev{x->G, x->B, x->ynorm2, x->pre1, x->pil_bar, dv.H, dv.Hw, nullptr, nullptr, 0.0};
    ev.theta = x->theta_dev;
    return launch_estep_any(dv, ev, st);
  } else if (x->eval == EVO_EVAL_BSC_DIRECT) {
    EVO_REQUIRE(x->W && x->Y, "direct evaluator needs W, Y");
    BscDirectEval ev{x->W, x->Y, x->mask, x->pre1, x->pil_bar, dv.D, dv.H, dv.Hw, nullptr, nullptr, 0.0};
    ev.theta = x->theta_dev;
    return launch_estep_any(dv, ev, st);
  } else if (x->eval == EVO_EVAL_ES3C) {
    return es3c_launch_estep(dv, x, st);
  }
  set_error("evo_estep: unknown evaluator %d", x->eval);
  return EVO_ERR_ARG;
}

static int estep_dims(const evo_estep_args *x, EstepDev &dv) {
  int P = x->n_parents < x->S ? x->n_parents : x->S;
  int Cg;
  if (x->inj_children) {
    EVO_REQUIRE(x->inj_count && x->inj_cmax > 0, "injected children need counts and a capacity");
    Cg = x->inj_cmax;
  } else {
    EVO_REQUIRE(x->n_generations == 0 || x->n_parents >= 1, "n_parents must be >= 1");
    if (x->mutation == EVO_MUT_RANDFLIP || x->mutation == EVO_MUT_SPARSEFLIP) {
      EVO_REQUIRE(x->n_generations == 0 || x->n_children >= 1, "n_children must be >= 1");
      Cg = P * x->n_children;
      EVO_REQUIRE(Cg <= 1024, "at most 1024 children per generation");
    } else {
      Cg = P * (P - 1);
    }
    if (x->mutation == EVO_MUT_SPARSEFLIP || x->mutation == EVO_MUT_CROSS_SPARSEFLIP)
      EVO_REQUIRE(x->bitflip_prob > 0.0, "Please specify the bitflip probability");  // eas.py:68
  }
  if (Cg < 1) Cg = 1;
  dv.Cg = Cg;
  dv.Ctot = Cg * (x->n_generations > 0 ? x->n_generations : 1);
  int need = 2 * (dv.S + dv.Ctot);
  int HT = 64;
  while (HT < need) HT <<= 1;
  dv.HT = HT;
  dv.PK = dv.S > Cg ? dv.S : Cg;
  auto pow2mask = [](int n) { uint32_t m = 1; while ((int)m < n) m <<= 1; return m - 1u; };
  dv.Hs = dv.Hw | 1;
  dv.Hr = dv.Hw;
  dv.ovf_list = x->overflow_list;
  dv.n_list = nullptr;
  dv.mask_c = pow2mask(dv.Ctot); dv.mask_s = pow2mask(dv.S); dv.mask_p = pow2mask(dv.PK);
  return EVO_OK;
}

extern "C" int evo_estep(const evo_estep_args *x, void *stream) { return estep_impl(x, nullptr, stream); }

extern "C" int evo_evolve(const evo_evolve_args *t, void *stream) {
  EVO_REQUIRE(t != nullptr, "null args");
  return estep_impl(&t->e, t, stream);
}

extern "C" int evo_evolve_dims(const evo_estep_args *x, int32_t *Cg, int32_t *Ctot, int32_t *PK) {
  EVO_REQUIRE(x != nullptr && x->S > 0 && x->H > 0, "bad sizes");
  EstepDev dv;
  memset(&dv, 0, sizeof(dv));
  dv.S = x->S; dv.H = x->H; dv.Hw = words_for(x->H);
  int rc = estep_dims(x, dv);
  if (rc) return rc;
  if (Cg) *Cg = dv.Cg;
  if (Ctot) *Ctot = dv.Ctot;
  if (PK) *PK = dv.PK;
  return EVO_OK;
}

extern "C" int evo_select_topk(uint32_t *states, double *lpj, int64_t N, int S, int S_perm, int H, int Mprime,
                               const uint32_t *new_states, const double *lpj_new, const int32_t *n_new, int U,
                               int32_t *out_counts, void *stream) {
  EVO_REQUIRE(N >= 0 && S > 0 && H > 0 && U >= 0, "bad sizes");
  if (N == 0 || U == 0) {
    if (out_counts && N > 0) EVO_CUDA(cudaMemsetAsync(out_counts, 0, sizeof(int32_t) * 2 * N, (cudaStream_t)stream));
    return EVO_OK;
  }
  int Hw = words_for(H);
  size_t smem = (size_t)U + S + 16;
  constexpr int T = 64;
  auto k = select_topk_kernel<T>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t g = (int64_t)sm_count() * 8;
  if (g > N) g = N;
  k<<<(int)g, T, smem, (cudaStream_t)stream>>>(states, lpj, N, S, S_perm, Hw, Mprime > 0 ? Mprime : S, new_states,
                                               lpj_new, n_new, U, out_counts);
  EVO_LAUNCH_CHECK("select_topk");
  return EVO_OK;
}
