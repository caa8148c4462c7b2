/* evo_b200 -- C-ABI of the B200-native E-step / M-step hot path of tvlearn/evo.
 *
 * The reference (pure Python, /root/reference) has no FFI; the entry points below are what its
 * model classes would bind for this path.  Each one names the reference function it replaces
 * (file:line relative to the reference root).  INTEGRATION.md shows the ctypes binding.
 *
 * Conventions
 *   - every function returns 0 on success, a negative code on error; the message is available
 *     from evo_last_error() (thread-local).
 *   - all pointers are DEVICE pointers unless the name says host; the caller owns every buffer.
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no hidden syncs.
 *   - K-state sets are bit-packed: uint32 [N][S][Hw], Hw = ceil(H/32), latent h lives in word
 *     h/32 bit h%32, pad bits are zero.  The reference stores them as (N,S,H) bool
 *     (evo/variational/utils.py:94-150).
 *   - lpj is float64 [N][S + S_perm], permanent states first (utils.py:94, _models.py:507-512).
 *   - Y is float32 [N][D]; masks are uint8 [N][D] (1 = usable), NULL = complete data.
 *   - W is float64 [D][H] row-major, exactly the reference's model_params["W"].
 */
#ifndef EVO_B200_H
#define EVO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EVO_OK 0
#define EVO_ERR_ARG (-1)
#define EVO_ERR_CUDA (-2)
#define EVO_ERR_UNSUPPORTED (-3)

/* parent selection (evo/variational/eas.py:138-150) */
#define EVO_SEL_FIT 0
#define EVO_SEL_RAND 1
/* mutation operators (eas.py:10-135) */
#define EVO_MUT_RANDFLIP 0
#define EVO_MUT_SPARSEFLIP 1
#define EVO_MUT_CROSS 2
#define EVO_MUT_CROSS_RANDFLIP 3
#define EVO_MUT_CROSS_SPARSEFLIP 4

/* log-joint evaluators */
#define EVO_EVAL_BSC_GRAM 0   /* sparse quadratic form s'Gs - 2 s'b (complete data)          */
#define EVO_EVAL_BSC_DIRECT 1 /* explicit masked residual (missing data / reference formula) */
#define EVO_EVAL_ES3C 2       /* per-state Woodbury / Cholesky                                */

const char *evo_last_error(void);
int evo_version(void);
/* number of SMs of the current device (grid sizing) */
int evo_device_sm_count(void);
/* kernels launched by this library since load (bench.py reports the count of the timed region) */
unsigned long long evo_launch_count(void);

/* ---- K-state container (evo/variational/utils.py:19-152) ------------------------------- */
/* (rows,H) bool bytes -> (rows,Hw) packed words, and back */
int evo_states_pack(const uint8_t *bools, uint32_t *packed, int64_t rows, int H, void *stream);
int evo_states_unpack(const uint32_t *packed, uint8_t *bools, int64_t rows, int H, void *stream);
/* Initial K^n: S distinct Bernoulli(p) states per datapoint, never equal to a permanent state
 * (utils.py:100-138).  Philox stream keyed by (seed, n_offset + n). */
int evo_states_init(uint32_t *states, int64_t N, int S, int H, int S_perm, int background, double p,
                    uint64_t seed, int64_t n_offset, void *stream);

/* ---- BSC log-pseudo-joint (evo/models/bsc.py:60-125, _models.py:567-596) ------------------ */
/* State-independent tables for the Gram form: G = W'W (H,H) f64, B = Y W (N,H) f64,
 * ynorm2[n] = |y_n|^2 f64.  Pass ynorm2 == NULL to skip it (Y is constant over EM iterations). */
int evo_bsc_prepare(const double *W, const float *Y, int64_t N, int D, int H, double *G, double *B,
                    double *ynorm2, void *stream);
/* The same table B = Y W on the tensor cores (tcgen05, TMA-fed): Y and W are scaled by powers of two
 * and split into fp16 hi + lo planes, three passes (hi*hi, hi*lo, lo*hi) accumulate in fp32 TMEM, the
 * epilogue writes f64.  |B_tc - B_f64| <= 2e-6 max|B| (measured <= 1e-6); lpj from it agrees with the
 * f64 reference to <= 1e-5 relative.  Workspace (256-byte aligned) holds the planes:
 *   evo_bsc_tables_tc_set_data  once per dataset (Y planes), evo_bsc_tables_tc once per parameter set. */
int64_t evo_bsc_tables_tc_workspace_bytes(int64_t N, int D, int H);
int evo_bsc_tables_tc_set_data(const float *Y, int64_t N, int D, int H, void *workspace, void *stream);
int evo_bsc_tables_tc(const double *W, int64_t N, int D, int H, void *workspace, double *B, void *stream);
/* lpj[n][k] = pre1 * |y_n - W s_nk|^2_obs + pil_bar |s_nk|, sanitised as lpj_reset_check.
 * eval = EVO_EVAL_BSC_GRAM uses (G,B,ynorm2); EVO_EVAL_BSC_DIRECT uses (W,Y,mask).
 * states: [N][Sp][Hw] or, with shared_states != 0, one [Sp][Hw] table used for every datapoint
 * (the reference's free_energy(full=True) state table, _models.py:385-431).
 * lpj_out has row stride ld_lpj doubles; reset_counters = 3 uint64 (nan, small, inf) or NULL. */
int evo_bsc_lpj(int eval, const uint32_t *states, int shared_states, int64_t N, int Sp, int D, int H,
                const double *W, const float *Y, const uint8_t *mask, const double *G, const double *B,
                const double *ynorm2, double pre1, double pil_bar, double *lpj_out, int64_t ld_lpj,
                unsigned long long *reset_counters, void *stream);

/* Dense tensor-core form of the same log-pseudo-joint (bsc.py:88-95: `np.dot(states, W.T) - y`, squared
 * and summed over the usable entries) as ONE fused tcgen05 GEMM  C[(n,k)][d] = sum_h s_nkh W[d][h]
 * whose epilogue forms pre1 * sum_d m_nd (C - y_nd)^2 + pil_bar |s| straight from TMEM; the N*S x D
 * residual is never written.  W is scaled by a power of two and split into two fp16 planes (hi + lo,
 * 22 significant bits; the binary operand is exact), both accumulated in fp32: lpj agrees with the f64
 * reference to <= 1e-5 relative (measured <= 2e-6).
 *   evo_bsc_lpj_gemm_workspace_bytes: size of the caller-owned workspace (256-byte aligned) for (D,H).
 *   evo_bsc_lpj_gemm_prepare: once per parameter set -- fills the workspace from W (two kernels).
 *   evo_bsc_lpj_gemm: same meaning of states / shared_states / mask / lpj_out / ld_lpj / reset_counters
 *   as evo_bsc_lpj (counters are incremented per sanitised state). */
int64_t evo_bsc_lpj_gemm_workspace_bytes(int D, int H);
int evo_bsc_lpj_gemm_prepare(const double *W, int D, int H, void *workspace, void *stream);
int evo_bsc_lpj_gemm(const uint32_t *states, int shared_states, int64_t N, int Sp, int D, int H,
                     const void *workspace, const float *Y, const uint8_t *mask, double pre1, double pil_bar,
                     double *lpj_out, int64_t ld_lpj, unsigned long long *reset_counters, void *stream);

/* ---- fused E-step (Model.E_step _models.py:454-565 = lpj of K^n + evolve_states eas.py:153-313
 *      + vary_Kn utils.py:231-337 + free-energy tail _models.py:544-547) ---------------------- */
typedef struct evo_estep_args {
  /* sizes */
  int64_t N;         /* datapoints on this rank */
  int64_t n_offset;  /* global index of row 0 (RNG keying: results do not depend on #ranks) */
  int32_t D, H, S, S_perm, Mprime;
  /* evolutionary hyper-parameters (container keys, utils.py:210-226) */
  int32_t n_parents, n_children, n_generations, selection, mutation, background;
  double bitflip_prob; /* p_bf */
  double sparseness;   /* theta["piH"] */
  /* RNG */
  uint64_t seed;
  uint32_t iteration;
  int32_t eval; /* EVO_EVAL_* */
  /* K sets, in place */
  uint32_t *states; /* [N][S][Hw] */
  double *lpj;      /* [N][S+S_perm] */
  /* injected candidates (parity tests): children[n][g][c] packed, count[n][g]; NULL = RNG mode */
  const uint32_t *inj_children; /* [N][G][inj_cmax][Hw] */
  const int32_t *inj_count;     /* [N][G] */
  int32_t inj_cmax;
  int32_t _pad0;
  /* model tables */
  const double *W;       /* [D][H]           (direct, es3c) */
  const float *Y;        /* [N][D]           (direct, es3c) */
  const uint8_t *mask;   /* [N][D] or NULL   (direct, es3c) */
  const double *G;       /* [H][H]           (gram) */
  const double *B;       /* [N][H]           (gram) */
  const double *ynorm2;  /* [N]              (gram) */
  double pre1, pil_bar;  /* BSC scalars (bsc.py:111-113) */
  /* ES3C tables (sssc.py:329-366): pil_bar_h (H), mus (H), Psi (H,H), sigma2 */
  const double *pil_bar_h, *mus, *Psi;
  double sigma2;
  /* outputs: acc[0] += sum_n logsumexp_k lpj_nk, acc[1] += #new unique, acc[2] += #substituted,
   * acc[3..5] += lpj_reset_check counters (nan, small, inf), acc[6] += #log-joints evaluated */
  double *acc;
  /* optional workspace of N + 1 int32 (or NULL).  With it, the BSC Gram E-step in RNG mode with randflip
   * keeps states in shared memory as lists of at most 15 active latents (32 bytes per state instead of
   * 4*Hw); datapoints that do not fit are collected here and redone by the packed-bits kernel in the same
   * call.  Results do not depend on which kernel handled a datapoint. */
  int32_t *overflow_list;
  /* FUSED BSC STATISTICS (optional; all three NULL = off).  With them, the randflip fast path of the Gram evaluator
   * adds each datapoint's share of Wq, pies and sigma (bsc.py:193-223) to `fs_stats` in the tail of the E-step --
   * K^n and its log-joints are still in shared memory there -- and writes <s>_n into the fp16 planes of the
   * tensor-core Wp workspace `fs_wp_workspace` (cleared beforehand by evo_bsc_wp_tc_begin).  Datapoints it does not
   * cover (packed-bits overflow pass, unions of more than 384 latents, more substitutions than scratch rows; every
   * datapoint if another kernel variant runs) are listed in `fs_rest` ([0] = count, N + 1 int32) for
   * evo_bsc_suffstats_tc_finish, which adds their share and runs the Wp GEMM. */
  double *fs_stats;
  void *fs_wp_workspace;
  int32_t *fs_rest;
  /* CUDA-GRAPH REPLAY (optional, BSC evaluators).  NULL: the scalars above are used.  Otherwise the kernels read
   * pre1, pil_bar, sparseness and iteration from theta_dev[0..3] (device memory, written by evo_bsc_em_advance), so a
   * captured launch stays valid while the parameters change from iteration to iteration. */
  const double *theta_dev;
} evo_estep_args;
int evo_estep(const evo_estep_args *args, void *stream);

/* ---- evolve_states (evo/variational/eas.py:153-313) as an entry point of its own ----------------------
 * The generation loop of evo_estep -- parent selection (eas.py:138-150), mutation (eas.py:10-135),
 * de-duplication against (permanent, K^n, earlier new) (eas.py:252-258), lpj of the new states (eas.py:267),
 * next pool = new + re-discovered (eas.py:271-295) -- with what evo_estep keeps on-chip made visible:
 *   - operator trace: the selected parents, the children exactly as mutated (before de-duplication) and the
 *     positions / cut points that produced them, per datapoint and generation;
 *   - evolve_only: the K sets, lpj and acc are left untouched and the new unique states of all generations
 *     come back with their lpj (what the reference's evolve_states returns, eas.py:313), together with the
 *     parent pool the next generation would select from.
 * `e` has the meaning it has for evo_estep (RNG mode, or injected children); selection may also be
 * EVO_SEL_ALL.  Always runs the generic kernel (every operator), never the randflip fast paths, and draws
 * the same Philox streams: with the same seed / iteration / n_offset the children are those evo_estep makes.
 * Buffer extents: evo_evolve_dims (Cg = children capacity per generation, Ctot = Cg * generations,
 * PK = pool capacity = max(S, Cg)).  Caller zeroes tr_count / tr_npar / ev_count (generations that do not
 * run leave them untouched). */
#define EVO_SEL_ALL 2 /* every pool member in pool order (first n_parents): operator calls on explicit parents */
typedef struct evo_evolve_args {
  evo_estep_args e;
  int32_t evolve_only; /* 1: do not touch e.states / e.lpj / e.acc; fill ev_* */
  int32_t lpj_given;   /* 1: e.lpj holds this_lpj of K^n (eas.py:196); K^n is not re-evaluated */
  int32_t skip_eval;   /* 1: new states come back un-evaluated (lpj 0): the caller applies its own eval_lpj */
  int32_t gen0;        /* index of the first generation of this call (Philox stream keying) */
  const int32_t *pool0;   /* [N][S] first parent pool as indices into K^n, pool0_n[n] entries; NULL = all of K^n */
  const int32_t *pool0_n; /* [N] */
  uint32_t *tr_parents;  /* [N][G][n_parents][Hw] in selection order, or NULL */
  int32_t *tr_npar;      /* [N][G] */
  uint32_t *tr_children; /* [N][G][Cg][Hw], or NULL */
  int32_t *tr_count;     /* [N][G] */
  int32_t *tr_pos;       /* [N][G][Cg][2]: randflip (flipped position, -1); cross* (cut point, extra flipped
                          * position or -1); sparseflip (-1, -1); or NULL */
  uint32_t *ev_states;   /* [N][Ctot][Hw] new unique states, generation by generation, first occurrence order */
  double *ev_lpj;        /* [N][Ctot] */
  int32_t *ev_count;     /* [N][1 + G]: total, then per generation */
  int32_t *ev_pool;      /* [N][PK] next parent pool: index < S = member of K^n, else S + index into ev_states */
  int32_t *ev_pool_n;    /* [N] */
} evo_evolve_args;
int evo_evolve_dims(const evo_estep_args *e, int32_t *Cg, int32_t *Ctot, int32_t *PK);
int evo_evolve(const evo_evolve_args *args, void *stream);

/* vary_Kn alone (utils.py:231-337) for one batch of datapoints: new candidates [N][U][Hw] with
 * lpj_new [N][U], n_new[n] valid rows each; updates states/lpj in place; out_counts[n] =
 * (#new unique, #substituted). */
int evo_select_topk(uint32_t *states, double *lpj, int64_t N, int S, int S_perm, int H, int Mprime,
                    const uint32_t *new_states, const double *lpj_new, const int32_t *n_new, int U,
                    int32_t *out_counts, void *stream);

/* sum_n logsumexp_k lpj[n][0..cols) accumulated into *out (f64) (_models.py:544-547) */
int evo_free_energy(const double *lpj, int64_t N, int cols, int64_t ld, double *out, void *stream);

/* ---- BSC M-step statistics (bsc.py:176-223) and estimator (_models.py:614-665) ------------ */
/* stats layout (f64): [Wp H*D | Wq H*H | pies H | sigma_sum 1]; accumulated (+=), caller zeroes.
 * Ym = y or y_reconstructed (f32 [N][D]).  Only the lower triangle (h' <= h) of Wq is written.
 * workspace (optional, 16-byte aligned, evo_bsc_suffstats_workspace_bytes(N, D, H) bytes; 0 = not
 * applicable for this H): with it, Wp is accumulated by a second kernel in shared-memory slices from
 * per-datapoint sparse <s> lists instead of one global atomic per (datapoint, latent, observable). */
int64_t evo_bsc_suffstats_workspace_bytes(int64_t N, int D, int H);
int evo_bsc_suffstats(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D,
                      int H, int S, int S_perm, int allzero, double pre1, double pil_bar,
                      double *stats, float *Es_out, void *workspace, int64_t workspace_bytes, void *stream);
/* The same statistics with Wp = sum_n <s>_n y_n' (bsc.py:204-206) on the tensor cores: the statistics pass
 * writes 2^14 <s> as fp16 hi + lo into transposed planes, Y' = 2^ey y lives there as hi + lo planes too
 * (evo_bsc_wp_tc_set_data, once per dataset / per reconstructed Ym), and a split-K tcgen05 GEMM (three
 * passes hi*hi, hi*lo, lo*hi, fp32 accumulation over <= 4096 datapoints per work item, f64 atomics across
 * work items) adds into stats.  Wq, pies and sigma are f64 as in evo_bsc_suffstats.
 * |Wp_tc - Wp_f64| <= 1e-5 max|Wp| (measured <= 2e-6).  Workspace: 256-byte aligned. */
int64_t evo_bsc_wp_tc_workspace_bytes(int64_t N, int D, int H);
int evo_bsc_wp_tc_set_data(const float *Ym, int64_t N, int D, int H, void *workspace, void *stream);
int evo_bsc_suffstats_tc(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H,
                         int S, int S_perm, double pre1, double pil_bar, double *stats, float *Es_out,
                         void *workspace, void *stream);
/* The two halves of evo_bsc_suffstats_tc around a fused E-step (evo_estep_args.fs_*): `begin` clears the <s> planes of
 * the workspace BEFORE evo_estep; `finish` adds the statistics of the datapoints listed in `rest` ([0] = count; NULL =
 * all N) and runs the split-K Wp GEMM.  theta_dev != NULL: pre1 / pil_bar are read from theta_dev[0..1] on the device
 * (CUDA-graph replay, see evo_bsc_em_advance) and the two arguments are ignored. */
int evo_bsc_wp_tc_begin(int64_t N, int D, int H, void *workspace, void *stream);
int evo_bsc_suffstats_tc_finish(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H,
                                int S, int S_perm, double pre1, double pil_bar, double *stats, void *workspace,
                                const int32_t *rest, const double *theta_dev, void *stream);
/* y_rec[n][d] = sum_h <s_h>_n W[d][h] where recon_mask[n][d] == 0 (x false); other entries copy y;
 * rows whose infer_mask is all zero are left untouched (_models.py:646-648). */
int evo_bsc_reconstruct(const uint32_t *states, const double *lpj, const double *W, const float *Y,
                        const uint8_t *recon_keep, const uint8_t *infer_mask, int64_t N, int D, int H,
                        int S, int S_perm, float *Yrec, void *stream);

/* ---- replicated parameter updates on the device (no library solver, no host fallback) ----------------------------
 * Both solve A X = B by block Gaussian elimination of [A | B] with explicitly inverted 32 x 32 diagonal blocks
 * (csrc/solve.cu).  `info` (scal[0]; for ES3C also scal[1]) is 0 when every pivot exceeded 1e-13 x the largest
 * diagonal entry, else 1 + the index of the first failing block: the statistics are (numerically) singular and the
 * caller takes the reference's minimum-norm route (bsc.py:233-250 lstsq / pinv; sssc.py:695-709).
 *
 * evo_bsc_solve -- BSC.M_step after the all-reduce (evo/models/bsc.py:226-275).  stats = the reduced buffer of
 * evo_bsc_suffstats [Wp H*D | Wq H*H (lower triangle) | pies H | sigma 1].  Writes W [D][H] = (Wq^-1 Wp)' (only if
 * learn_W and info == 0), pies [H] = sum / N_total (background: last = 1 - 1.1e-5), scal[4] = {info, pi = mean pies,
 * sigma = sqrt(sig / (N D)) or sqrt((sig + n_obs_total sigma_old^2) / (N D)) for incomplete data (n_obs_total >= 0),
 * 1 if W is finite}.  The caller keeps the parameters it does not learn. */
int64_t evo_bsc_solve_workspace_bytes(int D, int H);
int evo_bsc_solve(const double *stats, double N_total, int D, int H, int learn_W, int background, double n_obs_total,
                  double sigma_old, double *W, double *pies, double *scal, void *workspace, void *stream);
/* evo_bsc_em_advance -- the host's share of BSC.step between two iterations, on the device, so that a whole EM
 * iteration can be captured in ONE CUDA graph and replayed without a host round trip: logs the iteration that just ran
 * into rec[r][16] = {acc[8] (reduced E-step scalars), scal[4] (evo_bsc_solve), ljc, pi, sigma, iteration}, r = theta_dev[7]
 * (rows beyond max_rec are dropped; rec may be NULL), then applies check_params (_models.py:102-159: lo <= v, v >= hi -> hi)
 * to the new pi / sigma (scal[1], scal[2]; kept when not learned) and E_step_precompute (bsc.py:100-125):
 *   theta_dev[8] = {pre1, pil_bar, piH, iteration + 1, pi, sigma, ljc, r + 1}
 * which evo_estep (evo_estep_args.theta_dev) and evo_bsc_suffstats_tc_finish read.  Complete data only. */
int evo_bsc_em_advance(const double *acc, const double *scal, double *theta_dev, double *rec, int max_rec, int D, int H,
                       int learn_pi, int learn_sigma, double pi_lo, double pi_hi, double sigma_lo, double sigma_hi,
                       void *stream);
/* evo_bsc_solve_static -- NO counterpart in the reference: added for callers
 * that cannot allocate device memory (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments, outputs and error behaviour as evo_bsc_solve without the workspace: the elimination scratch is a
 * file-scope device array of EVO_BSC_SOLVE_STATIC_BYTES that every GPU loading the library holds until exit.  Sizes
 * with evo_bsc_solve_workspace_bytes(D, H) > EVO_BSC_SOLVE_STATIC_BYTES (e.g. beyond D = H = 512) are rejected with
 * EVO_ERR_ARG.  The call performs no cudaMalloc / cudaFree.  All calls on a device share the array: calls that overlap
 * on the GPU race -- keep them on one stream or order them with events. */
#define EVO_BSC_SOLVE_STATIC_BYTES (6 * 1024 * 1024)
int evo_bsc_solve_static(const double *stats, double N_total, int D, int H, int learn_W, int background,
                         double n_obs_total, double sigma_old, double *W, double *pies, double *scal, void *stream);
/* evo_es3c_update -- SSSC.EM_step parameter updates (evo/models/sssc.py:688-770) from the reduced buffer of
 * evo_es3c_suffstats.  W / pies / mus / Psi hold the current parameters and are overwritten where learned:
 *   W = Wp inv(sum_szsz) (:692-694), pies = clip(sum_s / N, eps, 1 - eps) (:712-721), mus = sum_sz / (sum_s + eps_mus)
 *   (:724-727), Psi = (mu mu' o sum_ss + sum_szsz - 2 mu[:,None] o sum_s_sz) o inv(sum_ss + eps_Psi I) -- the reference's
 *   ELEMENT-WISE product (:729-741) --, sigma2 (:744-770): complete data (ysq - trace(sum_outer W'W)) / (N D) + eps with
 *   the NEW W, incomplete data ((ysq - trace statistic) + n_obs_total sigma2_old) / (N D) + eps.
 * scal[4] = {info of the W solve, info of the Psi solve, sigma2_new, 1 if all outputs are finite}. */
typedef struct evo_es3c_update_args {
  int32_t D, H;
  int32_t learn_W, learn_pies, learn_mus, learn_Psi, learn_sigma2, background, incomplete, _pad;
  double N_total, n_obs_total, ysq, sigma2_old;
  double eps_pies, eps_mus, eps_Psi, eps_sigma2;
  const double *stats;
  double *W, *pies, *mus, *Psi; /* [D][H], [H], [H], [H][H] */
  double *scal;                 /* [4] */
  void *workspace;              /* evo_es3c_update_workspace_bytes(D, H), 16-byte aligned */
} evo_es3c_update_args;
int64_t evo_es3c_update_workspace_bytes(int D, int H);
int evo_es3c_update(const evo_es3c_update_args *args, void *stream);

/* ---- ES3C (class SSSC) ------------------------------------------------------------------------
 * The E-step is evo_estep with eval = EVO_EVAL_ES3C: complete data uses (G, B, ynorm2) from
 * evo_bsc_prepare, incomplete data uses Wt = W' [H][D] passed in `W`, plus pil_bar_h, mus, Psi, sigma2
 * (sssc.py:242-366).  States with more than 32 active latents are rejected (lowest lpj).
 * `es3c_scratch` (f64, es3c_scratch_warps * 3168 doubles) serves states with 17..32 active latents. */
int evo_es3c_set_scratch(double *scratch, int warps);
/* SSSC.log_pseudo_joint (sssc.py:242-326) for states [N][Sp][Hw] (or one shared [Sp][Hw] table) */
int evo_es3c_lpj(const uint32_t *states, int shared_states, int64_t N, int Sp, int D, int H, const double *Wt,
                 const float *Y, const uint8_t *mask, const double *G, const double *B, const double *ynorm2,
                 const double *Psi, const double *mus, const double *pil_bar_h, double sigma2, double *lpj_out,
                 int64_t ld_lpj, unsigned long long *reset_counters, void *stream);

/* Sufficient statistics + data estimator of SSSC.EM_step (sssc.py:553-646) over the post-selection
 * K sets.  stats (f64, accumulated, caller zeroes), H2 = H*H:
 *   [sum_s H | sum_sz H | sum_ss H2 | sum_szsz H2 | sum_s_sz H2 | sum_szsz_outer H2 | WpT H*D | trace 1]
 * sum_s_sz only if learn_psi; sum_szsz_outer for complete data, trace = sum_n |W_obs <sz>_n|^2 for
 * incomplete data (the reference's D x D accumulator is only ever traced, sssc.py:752).
 * Yrec (may be NULL): y_rec[n][d] = recon_keep[n][d] ? y : (W <sz>_n)_d  (sssc.py:614-627); with
 * incomplete data WpT accumulates y_rec (sssc.py:630-634). */
typedef struct evo_es3c_stats_args {
  int64_t N;
  int32_t D, H, S, S_perm, learn_psi, _pad;
  const uint32_t *states; /* [N][S][Hw] */
  const double *lpj;      /* [N][S+S_perm] */
  const float *Y;         /* [N][D] (0 where masked) */
  const uint8_t *mask;    /* [N][D] or NULL */
  const uint8_t *recon_keep; /* [N][D] or NULL (= reconstruct every entry) */
  const double *G, *B, *ynorm2; /* complete data */
  const double *Wt;       /* [H][D] */
  const double *Psi, *mus;
  double sigma2;
  double *stats;
  float *Yrec;            /* [N][D] or NULL */
  double q_skip;          /* states with posterior weight q <= q_skip are not expanded (their slab moments cost a
                           * factorisation each).  The mus update divides by sum_n <s_h> + 2.2e-16 (sssc.py:726), so the
                           * weight dropped per latent must stay far below that: 1e-30 bounds the relative change of
                           * every mu_h by 64 N 1e-30 / 2.2e-16 (< 1e-7 up to N = 3e5); 0 expands every state */
} evo_es3c_stats_args;
int evo_es3c_suffstats(const evo_es3c_stats_args *args, void *stream);

/* ---- exact log-likelihood (Model.free_energy(full=True), evo/models/_models.py:334-451; state table
 * evo/variational/utils.py:58-69) -- *out += sum_n logsumexp over ALL 2^Hm states of the model's log-pseudo-joint, the
 * states enumerated on the fly from their index (no table: the reference materialises (2^H, H) bools and stops at
 * H < 12).  Hm = H (or H - 1 with a background unit, whose bit is forced on) <= 24, H <= 32.  Arguments as for
 * evo_bsc_lpj / evo_es3c_lpj. */
int evo_bsc_full_logsumexp(int eval, int64_t N, int D, int H, int background, const double *W, const float *Y,
                           const uint8_t *mask, const double *G, const double *B, const double *ynorm2, double pre1,
                           double pil_bar, double *out, void *stream);
int evo_es3c_full_logsumexp(int64_t N, int D, int H, int background, const double *Wt, const float *Y, const uint8_t *mask,
                            const double *G, const double *B, const double *ynorm2, const double *Psi, const double *mus,
                            const double *pil_bar_h, double sigma2, double *out, void *stream);

/* ---- communicator: the one collective of an EM iteration (replaces evo/utils/parallel.py:45-67 and the ~15
 * Allreduce / Bcast calls per iteration of bsc.py:230-274, sssc.py:671-690, _models.py:139,156) -----------------------
 * NCCL resolved at run time (the copy already loaded by the process, else libnccl.so.2, else evo_comm_load(path)).
 * evo_comm_unique_id fills 128 bytes on one rank; the caller ships them to every rank (any host channel); each rank
 * calls evo_comm_init on its own device (collective).  evo_comm_allreduce: in-place SUM of `count` float64 on `stream`. */
int evo_comm_load(const char *libnccl_path);
int evo_comm_unique_id(void *id128);
int evo_comm_init(int rank, int world, const void *id128, void **comm_out);
int evo_comm_allreduce(void *comm, double *buf, int64_t count, void *stream);
int evo_comm_destroy(void *comm);

/* ---- device-side setup ----------------------------------------------------------------
 * Model.generate_data (evo/models/_models.py:73-99) -> BSC.generate_from_hidden (bsc.py:27-57) /
 * SSSC.generate_from_hidden (sssc.py:66-102): s_h ~ Bernoulli(pies[h]) (pies NULL: the scalar `pi`),
 * BSC (mus == Psi == NULL): y = W s + sigma N(0, I);  ES3C: z_A ~ N(mus_A, Psi_AA), y = W_A z_A + sigma N(0, I)
 * with sigma = sqrt(sigma2); psi_diag != 0 uses the diagonal of Psi only, otherwise the Cholesky factor of
 * Psi_AA (supports of more than 32 latents fall back to the diagonal and are counted in *overflow).
 * Row n is a pure function of (seed, n_offset + n): Philox-4x32-10 counter streams documented in
 * csrc/datagen.cu and restated in oracle/datagen.py -- shards of one dataset do not depend on the rank count.
 * Wt = W' [H][D] f64.  Outputs: Y f32 [N][D]; optional S_out packed [N][Hw], Z_out f32 [N][H], Ymean f32 [N][D]. */
int evo_generate_data(const double *Wt, const double *pies, double pi, const double *mus, const double *Psi,
                      int psi_diag, double sigma, int64_t N, int64_t n_offset, int D, int H, uint64_t seed,
                      float *Y, uint32_t *S_out, float *Z_out, float *Ymean, int32_t *overflow, void *stream);
/* standard_init on the device (evo/models/_models.py:206-283, sssc.py:104-197): column moments of the (masked) data,
 * out[3 D] += {sum m y, sum m y^2, sum m} (accumulated: the caller zeroes and all-reduces), and the initial dictionary
 * out[i][j] = row_mean[i] + scale N(0,1) ([rows][cols] f64; row_mean may be NULL) from a Philox stream keyed by
 * (seed, element) -- every rank fills the same values from the same seed, no broadcast. */
int evo_column_moments(const float *Y, const uint8_t *mask, int64_t N, int D, double *out, void *stream);
int evo_normal_fill(double *out, int64_t rows, int cols, const double *row_mean, double scale, uint64_t seed, void *stream);
/* FMA throughput of the SIMT pipes in TFLOP/s (dtype 32 or 64; 8 independent chains per thread, best of 3,
 * synchronous, default stream): the measured denominator of the ES3C roofline (bench.py). */
int evo_fma_peak(int dtype, int iters, double *tflops_out);

/* ---- overlapping-patch pipeline of the image drivers (tvutil.prepost OverlappingPatches /
 *      MultiDimOverlappingPatches .get / .set_and_merge with mean_merger / median_merger; call sites
 *      examples/image-denoising/main.py:85-88,169-177, examples/image-inpainting/main.py:86-88,174-182) ----
 * image f32 [C][Hi][Wi]; patches f32 [N][D], D = C*ph*pw (channel, patch row, patch column), N = ny*nx
 * top-left positions (multiples of shift that fit), row-major.  merge method: 0 = nan-mean, 1 = nan-median
 * over the estimates of all patches covering a pixel (<= 256 per pixel for the median); uncovered = NaN. */
int64_t evo_patches_count(int Hi, int Wi, int ph, int pw, int shift);
int evo_patches_extract(const float *image, int C, int Hi, int Wi, int ph, int pw, int shift, float *patches,
                        void *stream);
int evo_patches_merge(const float *patches, int C, int Hi, int Wi, int ph, int pw, int shift, int method,
                      float *image, void *stream);
/* *out += sum_i (a_i - b_i)^2 in f64 (PSNR = 10 log10(R^2 n / sum), examples/image-denoising/utils.py:8-18) */
int evo_sum_sqdiff(const float *a, const float *b, int64_t n, double *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* EVO_B200_H */
