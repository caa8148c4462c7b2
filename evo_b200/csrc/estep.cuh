// Fused truncated-variational E-step for one datapoint per thread group (T threads = one CTA):
//   lpj of K^n under the current parameters       (_models.py:507-513)
//   evolutionary generations: select -> mutate -> packed-bit hash de-duplication -> lpj of the
//   genuinely new states                            (evo/variational/eas.py:153-313)
//   top-K replacement (j-th best new <-> j-th worst old slot)   (evo/variational/utils.py:231-337)
//   free-energy tail sum_n logsumexp_k lpj_nk        (_models.py:544-547)
// The K^n block, the children of all generations, the hash table and the lpj values live in
// shared memory; HBM sees one read of K^n (+ the evaluator's per-datapoint table row), one write
// of the lpj row and writes of the substituted states only.
#pragma once
#include <stdlib.h>

#include "common.cuh"
#include "evaluators.cuh"
#include "stats.cuh"

namespace evo {

struct EstepDev {
  int64_t N, n_offset;
  int D, H, Hw, S, S_perm, Mprime;
  int Hs;    // row stride (words) of a state in SHARED memory, odd => one state per thread walks its
             // words without bank conflicts (global rows stay Hw)
  int Hr;    // words of one state row in shared memory: Hw (packed bits) or LIST_WORDS (index lists)
  int32_t *ovf_list;      // LIST kernel: [0] = count, [1..] datapoints it could not represent (or NULL)
  const int32_t *n_list;  // process only these datapoints ([0] = count, [1..]); NULL = all N
  int n_parents, n_children, n_generations, selection, mutation, background;
  double p_bf, sparseness;
  uint64_t seed;
  uint32_t iteration;
  uint32_t *states;
  double *lpj;
  const uint32_t *inj_children;
  const int32_t *inj_count;
  int inj_cmax;
  double *acc;
  // ---- generic kernel only: operator trace / stand-alone evolve (evo_evolve, include/evo_b200.h) ----
  int generic_only;  // any of the fields below is in use: the FAST / LIST variants are not launched
  int evolve_only, lpj_given, skip_eval, gen0;
  const int32_t *pool0, *pool0_n;
  uint32_t *tr_parents, *tr_children;
  int32_t *tr_npar, *tr_count, *tr_pos;
  uint32_t *ev_states;
  double *ev_lpj;
  int32_t *ev_count, *ev_pool, *ev_pool_n;
  // ---- fused BSC statistics in the tail of the randflip fast path (estep_flip.cuh); fs_rest == NULL: off --------------
  double *fs_Wq, *fs_pies, *fs_sig;
  __half *fs_hi, *fs_lo;  // transposed fp16 planes of 2^14 <s>: [h][n], row stride fs_ld
  int64_t fs_ld;
  int32_t *fs_rest;       // [0] = count, [1..]: datapoints the tail did not cover
  // ---- CUDA-graph replay: per-iteration scalars from device memory (evo_bsc_em_advance); NULL = the fields above ----
  const double *theta_dev;  // [0] pre1 [1] pil_bar [2] piH [3] iteration
  // derived
  int Cg;    // children capacity per generation
  int Ctot;  // Cg * n_generations
  int HT;    // hash table slots (power of two)
  int PK;    // capacity of pool / key arrays = max(S, Cg)
  uint32_t mask_c, mask_s, mask_p;  // index-code masks (pow2 - 1) for Ctot, S, PK (order_key)
};

// CUDA-graph replay: the per-iteration scalars are re-read from device memory at their points of use (theta_dev NULL: the
// kernel arguments; the BSC evaluators do the same for pre1 / pil_bar).  The argument structs are never written.
__device__ __forceinline__ uint32_t th_iteration(const EstepDev &a) { return a.theta_dev ? (uint32_t)a.theta_dev[3] : a.iteration; }
__device__ __forceinline__ double th_sparseness(const EstepDev &a) { return a.theta_dev ? a.theta_dev[2] : a.sparseness; }

constexpr int TAG_EMPTY = -1;
constexpr int TAG_CHILD = 1 << 24;

struct EstepSmem {
  uint32_t *st_uniq;   // [(S + Ctot)][Hs]
  uint32_t *st_child;  // [Cg][Hs]
  double *lpj_uniq;    // [S + Ctot]
  double *keys;        // [PK] order keys of the parent-selection pool
  double *vkeys;       // [Ctot + S] order keys of vary_Kn (aliases st_child / table)
  double *red_d;       // [8]
  int *table;          // [HT]
  int *pool;           // [PK]
  int *slot;           // [Cg]   hash slot / flip positions
  int *parents;        // [n_parents]
  int *red_i;          // [8]
  int *sh;             // [8] shared scalars
  unsigned char *redisc;  // [S + Ctot]
  int *rnk;               // [Ctot] rank of each eligible new state among the new ones (by lpj, descending)
  int *old_by_rank;       // [S] slot of the r-th worst old state
  int *eidx;              // [Ctot] indices of the new states that beat the worst old one (compacted)
  unsigned char *status;  // [Cg]
  unsigned char *eval_smem;
};

__host__ __device__ inline size_t align8(size_t x) { return (x + 7) & ~(size_t)7; }
__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

__host__ __device__ inline size_t estep_carve(const EstepDev &a, unsigned char *base, EstepSmem *s) {
  size_t off = 0;
  size_t U = (size_t)a.S + a.Ctot;
#define CARVE(field, type, count)                                  \
  do {                                                             \
    if (s) s->field = reinterpret_cast<type *>(base + off);       \
    off = align16(off + sizeof(type) * (size_t)(count));          \
  } while (0)
  CARVE(lpj_uniq, double, U);
  CARVE(keys, double, a.PK + 2);  // order keys of the selection pool
  CARVE(red_d, double, 8);
  CARVE(st_uniq, uint32_t, U * a.Hs);
  // Two phases share one region: the generations need the children and the hash table, the top-K
  // replacement afterwards (vary_Kn) needs order keys of the new and the old states and their ranks --
  // nothing of the first set is read once the generation loop has ended.
  {
    const size_t base_off = off;
    CARVE(st_child, uint32_t, (size_t)a.Cg * a.Hs);
    CARVE(table, int, a.HT);
    const size_t end_gen = off;
    off = base_off;
    CARVE(vkeys, double, (size_t)((a.Ctot + 1) & ~1) + a.S + 2);  // new [Ctot] then old [S]
    CARVE(rnk, int, a.Ctot > 0 ? a.Ctot : 1);
    CARVE(old_by_rank, int, a.S);
    CARVE(eidx, int, a.Ctot > 0 ? a.Ctot : 1);
    if (off < end_gen) off = end_gen;
  }
  CARVE(pool, int, a.PK);
  CARVE(slot, int, a.Cg > 0 ? a.Cg : 1);
  CARVE(parents, int, a.n_parents > 0 ? a.n_parents : 1);
  CARVE(red_i, int, 8);
  CARVE(sh, int, 8);
  CARVE(redisc, unsigned char, U);
  CARVE(status, unsigned char, a.Cg > 0 ? a.Cg : 1);
#undef CARVE
  if (s) s->eval_smem = base + off;
  return off;
}

// ---- sparse state rows (LIST representation) --------------------------------------------------
// A state with at most LIST_CAP active latents is kept in shared memory as 16 uint16: the active
// indices ascending in e[0..14] (0xFFFF padding) and the count in e[15] -- 32 bytes instead of
// 4*Hw (132 at H = 1024).  Hashing, comparing and copying treat the row as LIST_WORDS opaque words;
// only load, bit flip, evaluation and write-back know the layout.  A datapoint whose K^n or children
// do not fit is handed, untouched, to the packed-bits kernel through `ovf_list`.
constexpr int LIST_WORDS = 8, LIST_CAP = 15;

// returns false on overflow (the row is then unspecified)
__device__ __forceinline__ bool list_toggle(uint32_t *row, int pos) {
  uint16_t *e = reinterpret_cast<uint16_t *>(row);
  const int cnt = e[15];
  int i = 0;
  while (i < cnt && (int)e[i] < pos) ++i;
  if (i < cnt && (int)e[i] == pos) {
    for (int j = i; j < cnt - 1; ++j) e[j] = e[j + 1];
    e[cnt - 1] = 0xFFFFu;
    e[15] = (uint16_t)(cnt - 1);
    return true;
  }
  if (cnt >= LIST_CAP) return false;
  for (int j = cnt; j > i; --j) e[j] = e[j - 1];
  e[i] = (uint16_t)pos;
  e[15] = (uint16_t)(cnt + 1);
  return true;
}
// packed bits (global, Hg words) -> list row, one WARP per state; returns false on overflow
__device__ __forceinline__ bool list_from_bits(const uint32_t *src, int Hg, uint32_t *row) {
  uint16_t *e = reinterpret_cast<uint16_t *>(row);
  const int lane = threadIdx.x & 31;
  if (lane < 16) e[lane] = 0xFFFFu;
  __syncwarp();
  int base = 0;
  for (int w0 = 0; w0 < Hg; w0 += 32) {
    uint32_t word = (w0 + lane < Hg) ? src[w0 + lane] : 0u;
    const int c = __popc(word);
    int pre = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
    int pos = base + pre - c;
    while (word) {
      int bp = __ffs(word) - 1;
      word &= word - 1;
      if (pos < LIST_CAP) e[pos] = (uint16_t)(((w0 + lane) << 5) + bp);
      ++pos;
    }
    base += __shfl_sync(0xffffffffu, pre, 31);
  }
  __syncwarp();
  if (lane == 0) e[15] = (uint16_t)(base <= LIST_CAP ? base : LIST_CAP);
  __syncwarp();
  return base <= LIST_CAP;
}
// the same conversion by ONE thread (short rows: fewer instructions than a warp scan per state)
__device__ __forceinline__ bool list_from_bits_thread(const uint32_t *src, int Hg, uint32_t *row) {
  uint16_t *e = reinterpret_cast<uint16_t *>(row);
#pragma unroll
  for (int w = 0; w < LIST_WORDS; ++w) row[w] = 0xFFFFFFFFu;
  int cnt = 0;
  for (int w0 = 0; w0 < Hg; w0 += 4) {
    uint32_t x[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) x[q] = (w0 + q < Hg) ? __ldg(src + w0 + q) : 0u;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t m = x[q];
      while (m) {
        const int bp = __ffs(m) - 1;
        m &= m - 1;
        if (cnt < LIST_CAP) e[cnt] = (uint16_t)(((w0 + q) << 5) + bp);
        ++cnt;
      }
    }
  }
  e[15] = (uint16_t)(cnt <= LIST_CAP ? cnt : LIST_CAP);
  return cnt <= LIST_CAP;
}
// list row -> packed bits (Hg words), one thread
__device__ __forceinline__ void list_to_bits(const uint32_t *row, int Hg, uint32_t *dst) {
  const uint16_t *e = reinterpret_cast<const uint16_t *>(row);
  const int cnt = e[15];
  int p = 0;
  for (int w = 0; w < Hg; ++w) {
    uint32_t x = 0u;
    while (p < cnt && (e[p] >> 5) == w) { x |= 1u << (e[p] & 31); ++p; }
    dst[w] = x;
  }
}

// ---- hash table over packed states --------------------------------------------------------
// (LIST rows have a compile-time width, so hashing and comparing unroll fully)
template <bool LIST>
__device__ __forceinline__ const uint32_t *tag_state(const EstepDev &a, const EstepSmem &s, int tag) {
  const int hs = LIST ? LIST_WORDS + 1 : a.Hs;
  return tag >= TAG_CHILD ? s.st_child + (size_t)(tag - TAG_CHILD) * hs : s.st_uniq + (size_t)tag * hs;
}
// read-only probe: index of the uniq entry equal to st, or -1
template <bool LIST>
__device__ __forceinline__ int table_find(const EstepDev &a, const EstepSmem &s, const uint32_t *st) {
  const int hw = LIST ? LIST_WORDS : a.Hr;
  uint32_t mask = (uint32_t)a.HT - 1u;
  uint32_t p = (uint32_t)hash_state(st, hw) & mask;
  for (int it = 0; it < a.HT; ++it) {
    int tag = s.table[p];
    if (tag == TAG_EMPTY) return -1;
    if (same_state(tag_state<LIST>(a, s, tag), st, hw)) return tag;
    p = (p + 1) & mask;
  }
  return -1;
}
// concurrent insert; equal states converge on one slot holding the smallest tag.  Returns the slot.
template <bool LIST>
__device__ __forceinline__ int table_insert(const EstepDev &a, const EstepSmem &s, const uint32_t *st, int tag) {
  const int hw = LIST ? LIST_WORDS : a.Hr;
  uint32_t mask = (uint32_t)a.HT - 1u;
  uint32_t p = (uint32_t)hash_state(st, hw) & mask;
  for (int it = 0; it < a.HT; ++it) {
    int cur = s.table[p];
    if (cur == TAG_EMPTY) {
      cur = atomicCAS(&s.table[p], TAG_EMPTY, tag);
      if (cur == TAG_EMPTY) return (int)p;
    }
    if (same_state(tag_state<LIST>(a, s, cur), st, hw)) {
      atomicMin(&s.table[p], tag);
      return (int)p;
    }
    p = (p + 1) & mask;
  }
  return -1;  // table full: cannot happen, HT >= 2 * (S + Ctot)
}

// ---- rank counting on order keys ------------------------------------------------------------
// One f64 compare per pair: the low mantissa bits of a key carry the element's index, so keys are
// pairwise distinct and values that agree in all higher bits (|rel. difference| < 2^-44 for
// U <= 256) are ordered by index -- the smaller index compares LARGER when `hi` is true (ranking by
// value descending, index ascending), SMALLER otherwise (value ascending, index ascending).
__device__ __forceinline__ double order_key(double v, uint32_t idx, uint32_t mask, bool hi) {
  uint64_t b = (uint64_t)__double_as_longlong(v);
  const bool neg = (b >> 63) != 0;
  const uint32_t code = (hi != neg) ? (mask - idx) : idx;
  return __longlong_as_double((long long)((b & ~(uint64_t)mask) | code));
}
__device__ __forceinline__ int count_greater(const double *keys, int n, double k) {
  int r = 0, j = 0;
#pragma unroll 2
  for (; j + 4 <= n; j += 4) {
    const double2 x = *reinterpret_cast<const double2 *>(keys + j);
    const double2 y = *reinterpret_cast<const double2 *>(keys + j + 2);
    r += (int)(x.x > k) + (int)(x.y > k) + (int)(y.x > k) + (int)(y.y > k);
  }
#pragma unroll 1
  for (; j < n; ++j) r += (int)(keys[j] > k);
  return r;
}
__device__ __forceinline__ int count_less(const double *keys, int n, double k) {
  int r = 0, j = 0;
#pragma unroll 2
  for (; j + 4 <= n; j += 4) {
    const double2 x = *reinterpret_cast<const double2 *>(keys + j);
    const double2 y = *reinterpret_cast<const double2 *>(keys + j + 2);
    r += (int)(x.x < k) + (int)(x.y < k) + (int)(y.x < k) + (int)(y.y < k);
  }
#pragma unroll 1
  for (; j < n; ++j) r += (int)(keys[j] < k);
  return r;
}

// ---- mutation helpers -----------------------------------------------------------------------
__device__ __forceinline__ void flip_bit(uint32_t *st, int pos) { st[pos >> 5] ^= 1u << (pos & 31); }

// randflip positions (eas.py:27-42): C distinct uniformly drawn positions per parent; child c = p * C + j flips slot[c].
// Every child draws its own position (Philox draw index c * 64 + round); in rounds, a child that collides with a LOWER
// sibling's current position redraws, until no collision is left.  Relabelling the positions maps trajectories of this
// process onto equally likely ones, so the final tuple is uniform over the ordered C-tuples of distinct positions --
// the distribution of the reference's sequential draw -- while all children of all parents draw in parallel.
// Shared by every kernel variant (same streams => same children).  Ends with a group barrier.
template <int T>
__device__ __forceinline__ void draw_flip_positions(const RngCtx &rng, uint32_t gr, int take, int C, int Hm, int *slot, int *red_i) {
  const int tid = threadIdx.x, nc = take * C;
  _Pragma("unroll 1") for (int c = tid; c < nc; c += T)
    slot[c] = (int)urange(rng.draw(RNG_MUTATE, gr, (uint32_t)c * 64u).x, (uint32_t)Hm);
  gsync<T>();
  if (C > Hm) return;  // cannot be distinct (degenerate input): keep the first draw
  for (uint32_t round = 1; round < 64u; ++round) {
    unsigned need = 0;
    int it = 0;
    _Pragma("unroll 1") for (int c = tid; c < nc; c += T, ++it) {
      const int j = c % C, b0 = c - j, pos = slot[c];
      bool dup = false;
      for (int q = 0; q < j; ++q) dup |= (slot[b0 + q] == pos);
      if (dup) need |= 1u << (it & 31);
    }
    if (!group_or<T>(need, red_i)) break;  // (barriers inside: every read of this round is done)
    it = 0;
    _Pragma("unroll 1") for (int c = tid; c < nc; c += T, ++it)
      if ((need >> (it & 31)) & 1u) slot[c] = (int)urange(rng.draw(RNG_MUTATE, gr, (uint32_t)c * 64u + round).x, (uint32_t)Hm);
    gsync<T>();
  }
}

// sparseflip (eas.py:46-100): independent flips, p1 on on-bits, p0 on off-bits, over the first Hm
// latents.  Sampled by thinning a Bernoulli(pmax) process visited with geometric jumps -- same
// distribution, O(#candidates) draws instead of O(H).
__device__ inline void sparseflip_state(uint32_t *st, int Hm, double sparseness, double p_bf, const RngCtx &rng,
                                        uint32_t purpose, uint32_t g, uint32_t c) {
  int a = 0;
  for (int i = 0; i < Hm; i += 32) {
    uint32_t w = st[i >> 5];
    if (Hm - i < 32) w &= (1u << (Hm - i)) - 1u;
    a += __popc(w);
  }
  double H = (double)Hm, s = (double)a, tiny = 1e-100;
  double alpha = (H - s) * (H * p_bf - (sparseness - s)) / ((sparseness - s + H * p_bf) * s + tiny);
  double p0 = H * p_bf / (H + (alpha - 1.0) * s + tiny);
  double p1 = alpha * p0;
  p0 = fmin(fmax(p0, 0.0), 1.0);
  p1 = fmin(fmax(p1, 0.0), 1.0);
  if (!(p0 == p0)) p0 = 0.0;
  if (!(p1 == p1)) p1 = 0.0;
  double pmax = fmax(p0, p1);
  if (pmax <= 0.0) return;
  double inv_log = (pmax < 1.0) ? 1.0 / log1p(-pmax) : 0.0;
  uint32_t draw_idx = c * 4096u;
  long pos = -1;
  for (;;) {
    uint4 r = rng.draw(purpose, g, draw_idx++);
    double u = u01(r.x, r.y);
    long gap = (pmax < 1.0) ? (long)floor(log(u) * inv_log) : 0;
    pos += 1 + gap;
    if (pos >= Hm) break;
    bool on = (st[pos >> 5] >> (pos & 31)) & 1u;
    double pb = on ? p1 : p0;
    double v = u01(r.z, r.w);
    if (v * pmax < pb) flip_bit(st, (int)pos);
  }
}

// ------------------------------------------------------------------------------------------
// FAST = the RNG hot path only (randflip mutation, no injected candidates, no background unit): the
// other operators are compiled out, which shrinks the kernel's instruction footprint.
// LIST = states live in shared memory as index lists (see above); requires FAST and an evaluator with eval_batch<T, true>.
template <class Eval, int T, bool FAST = false, bool LIST = false>
__global__ void __launch_bounds__(T, Eval::kThreadsPerSM / T) estep_kernel(EstepDev a, Eval ev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  EstepSmem s;
  estep_carve(a, smem_raw, &s);
  const int tid = threadIdx.x;
  const int Hw = LIST ? LIST_WORDS : a.Hr, Hg = a.Hw, Hs = LIST ? LIST_WORDS + 1 : a.Hs, S = a.S;  // Hw: words of a shared-memory row, Hg: of a global row
  const bool bgu = FAST ? false : (a.background != 0);
  const int Hm = bgu ? a.H - 1 : a.H;
  double acc_F = 0.0, acc_U = 0.0, acc_sub = 0.0, acc_eval = 0.0;
  unsigned long long cnt_nan = 0, cnt_small = 0, cnt_inf = 0;

  const int64_t n_total = a.n_list ? (int64_t)a.n_list[0] : a.N;
  for (int64_t it = blockIdx.x; it < n_total; it += gridDim.x) {
    const int64_t n = a.n_list ? (int64_t)a.n_list[1 + it] : it;
    RngCtx rng(a.seed, a.n_offset + n, th_iteration(a));
    gsync<T>();  // previous datapoint fully done with shared memory
    // ---- load K^n, evaluator tables ---------------------------------------------------------
    {
      const uint32_t *src = a.states + n * (int64_t)S * Hg;
      if (LIST) {
        if (tid == 0) s.sh[0] = 0;
        gsync<T>();
        if (Hg <= 64) {
          _Pragma("unroll 1") for (int i = tid; i < S; i += T)
            if (!list_from_bits_thread(src + (size_t)i * Hg, Hg, s.st_uniq + (size_t)i * Hs)) s.sh[0] = 1;
        } else {
          for (int i = tid >> 5; i < S; i += T / 32)
            if (!list_from_bits(src + (size_t)i * Hg, Hg, s.st_uniq + (size_t)i * Hs) && (tid & 31) == 0) s.sh[0] = 1;
        }
      } else {
        {  // global rows of Hw words -> shared rows of Hs words; (row, word) advanced without a division per element
          int r = tid / Hw, c = tid - r * Hw;
          const int dr = T / Hw, dc = T - dr * Hw;
          _Pragma("unroll 1") for (int i = tid; i < S * Hw; i += T) {
            s.st_uniq[r * Hs + c] = src[i];
            r += dr; c += dc;
            if (c >= Hw) { c -= Hw; ++r; }
          }
        }
      }
      _Pragma("unroll 1") for (int i = tid; i < a.HT; i += T) s.table[i] = TAG_EMPTY;
      _Pragma("unroll 1") for (int i = tid; i < S + a.Ctot; i += T) s.redisc[i] = 0;
      if constexpr (LIST) ev.template setup<T, true>(n, s.eval_smem);
      else ev.template setup<T>(n, s.eval_smem);
    }
    gsync<T>();
    if (LIST && s.sh[0]) {  // a state of K^n has more than LIST_CAP active latents: not for this kernel
      if (tid == 0) a.ovf_list[1 + atomicAdd(a.ovf_list, 1)] = (int32_t)n;
      continue;
    }
    // ---- lpj of the permanent state ---------------------------------------------------------------
    double lpj_perm = 0.0;
    if (a.S_perm > 0) {
      unsigned f0 = 0;
      lpj_perm = ev.eval_zero(f0);
      if (tid == 0 && f0) { if (f0 & 1) ++cnt_nan; else if (f0 & 2) ++cnt_small; else ++cnt_inf; }
    }

    int U = 0;        // accepted new unique states so far
    int pool_n = S;   // generation 0 selects from K^n

    // pass g = -1 evaluates K^n, passes 0.. are the generations: ONE call site of the evaluator, so its
    // code is instantiated once (instruction-cache footprint)
    for (int g = -1; g < a.n_generations; ++g) {
      int nc = 0, u_g = 0;
      const uint32_t gr = FAST ? (uint32_t)g : (uint32_t)(g + a.gen0);  // generation index of the Philox streams
      if (g >= 0) {
      if (!FAST && a.inj_children) {
        // ---- injected candidates (parity runs) -------------------------------------------------
        nc = a.inj_count[n * (int64_t)a.n_generations + g];
        if (nc > a.Cg) nc = a.Cg;
        const uint32_t *src = a.inj_children + ((n * (int64_t)a.n_generations + g) * a.inj_cmax) * (int64_t)Hg;
        _Pragma("unroll 1") for (int i = tid; i < nc * Hw; i += T) s.st_child[(i / Hw) * Hs + (i % Hw)] = src[i];
      } else {
        // ---- parent selection (eas.py:138-150) -------------------------------------------------
        int take = pool_n < a.n_parents ? pool_n : a.n_parents;
        {
          double mn = DBL_MAX;
          if (a.selection == EVO_SEL_FIT) {
            _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) mn = fmin(mn, s.lpj_uniq[s.pool[i]]);
            mn = -group_max<T>(-mn, s.red_d);
          }
          double shift = 2.0 * fmin(mn, 0.0);
          if (!FAST && a.selection == EVO_SEL_ALL) {  // explicit parents: the pool in its own order
            _Pragma("unroll 1") for (int i = tid; i < take; i += T) s.parents[i] = s.pool[i];
          } else {
          _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) {
            uint4 r = rng.draw(RNG_SELECT, gr, (uint32_t)i);
            double u = u01(r.x, r.y);
            double key;
            if (a.selection == EVO_SEL_FIT) {
              double w = s.lpj_uniq[s.pool[i]] - shift;
              // Efraimidis-Spirakis weighted sampling without replacement: keep the largest log(u)/w
              key = (w > 0.0) ? (double)(__logf((float)u) / (float)w) : -DBL_MAX;
            } else {
              key = u;
            }
            s.keys[i] = order_key(key, (uint32_t)i, a.mask_p, true);
          }
          gsync<T>();
          // the `take` largest keys, by rank counting (strict total order: key desc, index asc)
          _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) {
            int r = count_greater(s.keys, pool_n, s.keys[i]);
            if (r < take) s.parents[r] = s.pool[i];
          }
          }
          gsync<T>();
          if (!FAST && a.tr_parents) {
            uint32_t *dst = a.tr_parents + ((n * (int64_t)a.n_generations + g) * a.n_parents) * (int64_t)Hg;
            _Pragma("unroll 1") for (int i = tid; i < take * Hw; i += T)
              dst[i] = s.st_uniq[(size_t)s.parents[i / Hw] * Hs + (i % Hw)];
            if (tid == 0) a.tr_npar[n * (int64_t)a.n_generations + g] = take;
          }
        }
        // ---- mutation (eas.py:10-135) -----------------------------------------------------------
        const int mut = FAST ? (int)EVO_MUT_RANDFLIP : a.mutation;
        if (FAST || mut == EVO_MUT_RANDFLIP || mut == EVO_MUT_SPARSEFLIP) {
          const int C = a.n_children;
          nc = take * C;
          if (FAST || mut == EVO_MUT_RANDFLIP) draw_flip_positions<T>(rng, gr, take, C, Hm, s.slot, s.red_i);
          _Pragma("unroll 1") for (int c = tid; c < nc; c += T) {
            int p = c / C;
            const uint32_t *src = s.st_uniq + (size_t)s.parents[p] * Hs;
            uint32_t *dst = s.st_child + (size_t)c * Hs;
            for (int w = 0; w < Hw; ++w) dst[w] = src[w];
            if (bgu) dst[(a.H - 1) >> 5] &= ~(1u << ((a.H - 1) & 31));
            if (LIST) { if (!list_toggle(dst, s.slot[c])) s.sh[0] = 1; }
            else if (FAST || mut == EVO_MUT_RANDFLIP) flip_bit(dst, s.slot[c]);
            else sparseflip_state(dst, Hm, th_sparseness(a), a.p_bf, rng, RNG_MUTATE, gr, (uint32_t)c);
            if (bgu) dst[(a.H - 1) >> 5] |= 1u << ((a.H - 1) & 31);
            if (!FAST && a.tr_pos) {
              int32_t *tp = a.tr_pos + ((n * (int64_t)a.n_generations + g) * a.Cg + c) * 2;
              tp[0] = (mut == EVO_MUT_RANDFLIP) ? s.slot[c] : -1; tp[1] = -1;
            }
          }
        } else {
          // one-point crossover of every unordered parent pair, two children per pair
          int npairs = take * (take - 1) / 2;
          nc = 2 * npairs;
          _Pragma("unroll 1") for (int q = tid; q < npairs; q += T) {
            int i = 0, rem = q;
            while (rem >= take - 1 - i) { rem -= take - 1 - i; ++i; }
            int j = i + 1 + rem;
            uint4 r = rng.draw(RNG_CROSS, gr, (uint32_t)q);
            int cp = 1 + (Hm > 1 ? (int)urange(r.x, (uint32_t)(Hm - 1)) : 0);  // cp in {1..Hm-1}
            const uint32_t *pa = s.st_uniq + (size_t)s.parents[i] * Hs;
            const uint32_t *pb = s.st_uniq + (size_t)s.parents[j] * Hs;
            uint32_t *c0 = s.st_child + (size_t)(2 * q) * Hs, *c1 = c0 + Hs;
            for (int w = 0; w < Hw; ++w) {
              int lo = w << 5;
              uint32_t head = cp >= lo + 32 ? 0xFFFFFFFFu : (cp <= lo ? 0u : ((1u << (cp - lo)) - 1u));
              c0[w] = (pa[w] & head) | (pb[w] & ~head);
              c1[w] = (pb[w] & head) | (pa[w] & ~head);
            }
            for (int k = 0; k < 2; ++k) {
              uint32_t *dst = k ? c1 : c0;
              uint32_t cidx = (uint32_t)(2 * q + k);
              if (bgu) dst[(a.H - 1) >> 5] &= ~(1u << ((a.H - 1) & 31));
              int fp = -1;
              if (mut == EVO_MUT_CROSS_RANDFLIP) {
                uint4 r2 = rng.draw(RNG_FLIP2, gr, cidx);
                fp = (int)urange(r2.x, (uint32_t)Hm);
                flip_bit(dst, fp);
              } else if (mut == EVO_MUT_CROSS_SPARSEFLIP) {
                sparseflip_state(dst, Hm, th_sparseness(a), a.p_bf, rng, RNG_FLIP2, gr, cidx);
              }
              if (a.tr_pos) {
                int32_t *tp = a.tr_pos + ((n * (int64_t)a.n_generations + g) * a.Cg + cidx) * 2;
                tp[0] = cp; tp[1] = fp;
              }
              if (bgu) dst[(a.H - 1) >> 5] |= 1u << ((a.H - 1) & 31);
            }
          }
        }
      }
      gsync<T>();
      if (LIST && s.sh[0]) break;  // a child does not fit the list form: the packed-bits kernel redoes this datapoint
      if (!FAST && a.tr_children) {
        uint32_t *dst = a.tr_children + ((n * (int64_t)a.n_generations + g) * a.Cg) * (int64_t)Hg;
        _Pragma("unroll 1") for (int i = tid; i < nc * Hw; i += T) dst[i] = s.st_child[(i / Hw) * Hs + (i % Hw)];
        if (tid == 0) a.tr_count[n * (int64_t)a.n_generations + g] = nc;
      }

      // ---- de-duplicate against (permanent, K^n, earlier new) and among the children ----------
      _Pragma("unroll 1") for (int c = tid; c < nc; c += T) {
        const uint32_t *st = s.st_child + (size_t)c * Hs;
        unsigned char stt = 1;  // candidate
        if (a.S_perm > 0 && (LIST ? reinterpret_cast<const uint16_t *>(st)[15] == 0 : popc_state(st, Hw) == 0)) stt = 0;
        else {
          int hit = table_find<LIST>(a, s, st);
          if (hit >= 0) { s.redisc[hit] = 1; stt = 0; }
        }
        s.status[c] = stt;
      }
      gsync<T>();
      _Pragma("unroll 1") for (int c = tid; c < nc; c += T)
        if (s.status[c]) s.slot[c] = table_insert<LIST>(a, s, s.st_child + (size_t)c * Hs, TAG_CHILD + c);
      gsync<T>();
      // first occurrence of each new state wins; compact winners (child order) behind st_uniq
      for (int base = 0; base < nc; base += T) {
        int c = base + tid;
        bool win = false;
        int sl = -1;
        if (c < nc && s.status[c]) { sl = s.slot[c]; win = (sl >= 0 && s.table[sl] == TAG_CHILD + c); }
        int tot;
        int off = group_scan_flag<T>(win, tot, s.red_i);
        if (win) {
          int idx = S + U + u_g + off;
          const uint32_t *src = s.st_child + (size_t)c * Hs;
          uint32_t *dst = s.st_uniq + (size_t)idx * Hs;
          for (int w = 0; w < Hw; ++w) dst[w] = src[w];
          s.status[c] = 2;
          s.table[sl] = idx;  // from now on a uniq entry (losers of this state compare against a child tag and fail)
        }
        u_g += tot;
        gsync<T>();
      }

      }  // g >= 0

      // ---- lpj of K^n (g = -1) or of the genuinely new states -----------------------------------
      {
        const int first = g < 0 ? 0 : S + U, cnt = g < 0 ? S : u_g;
        unsigned fl = 0;
        if (!FAST && g < 0 && a.lpj_given) {  // this_lpj supplied by the caller (eas.py:196)
          const double *row = a.lpj + n * (int64_t)(S + a.S_perm) + a.S_perm;
          _Pragma("unroll 1") for (int i = tid; i < S; i += T) s.lpj_uniq[i] = row[i];
        } else if (!FAST && g >= 0 && a.skip_eval) {  // the caller evaluates the new states itself
          _Pragma("unroll 1") for (int i = tid; i < cnt; i += T) s.lpj_uniq[first + i] = 0.0;
        } else if (cnt > 0) {
          ev.template eval_batch<T, LIST>(s.st_uniq + (size_t)first * Hs, Hs, cnt, s.lpj_uniq + first, fl);
          fl = group_or<T>(fl, s.red_i);
          if (tid == 0 && fl) { if (fl & 1) ++cnt_nan; else if (fl & 2) ++cnt_small; else ++cnt_inf; }
        }
      }
      gsync<T>();
      if (g < 0) {
        _Pragma("unroll 1") for (int i = tid; i < S; i += T) {
          table_insert<LIST>(a, s, s.st_uniq + (size_t)i * Hs, i);
          s.pool[i] = i;
        }
        gsync<T>();
        if (!FAST && a.pool0) {  // caller-supplied first pool (generation-by-generation evolve_states)
          pool_n = a.pool0_n[n];
          if (pool_n > S) pool_n = S;
          _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) s.pool[i] = a.pool0[n * (int64_t)S + i];
          gsync<T>();
        }
        continue;
      }
      if (!FAST && a.ev_count && tid == 0) a.ev_count[n * (int64_t)(1 + a.n_generations) + 1 + g] = u_g;
      // ---- next pool: this round's new states, then the re-discovered old ones -----------------
      int pn = 0;
      _Pragma("unroll 1") for (int i = tid; i < u_g && i < a.PK; i += T) s.pool[i] = S + U + i;
      pn = u_g < a.PK ? u_g : a.PK;
      for (int base = 0; base < S + U; base += T) {
        int i = base + tid;
        bool f = (i < S + U) && s.redisc[i];
        int tot;
        int off = group_scan_flag<T>(f, tot, s.red_i);
        if (f) {
          if (pn + off < a.PK) s.pool[pn + off] = i;
          s.redisc[i] = 0;
        }
        pn += tot;
        if (pn > a.PK) pn = a.PK;
        gsync<T>();
      }
      U += u_g;
      pool_n = pn;
      if (pool_n == 0) break;  // "No new and unique states" (eas.py:306-308)
    }
    if (LIST && s.sh[0]) {  // nothing of this datapoint has been written yet
      if (tid == 0) a.ovf_list[1 + atomicAdd(a.ovf_list, 1)] = (int32_t)n;
      continue;
    }

    if (!FAST && a.evolve_only) {  // evolve_states alone (eas.py:313): hand the new states back, leave K^n as it is
      gsync<T>();
      uint32_t *dst = a.ev_states + n * (int64_t)a.Ctot * Hg;
      _Pragma("unroll 1") for (int i = tid; i < U * Hw; i += T) dst[i] = s.st_uniq[(size_t)(S + i / Hw) * Hs + (i % Hw)];
      _Pragma("unroll 1") for (int i = tid; i < U; i += T) a.ev_lpj[n * (int64_t)a.Ctot + i] = s.lpj_uniq[S + i];
      if (a.ev_pool) {
        _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) a.ev_pool[n * (int64_t)a.PK + i] = s.pool[i];
        if (tid == 0) a.ev_pool_n[n] = pool_n;
      }
      if (tid == 0) a.ev_count[n * (int64_t)(1 + a.n_generations)] = U;
      continue;
    }
    // ---- vary_Kn: j-th best new replaces j-th worst old while it is better ----------------------
    // Ranks by counting (strict total orders: new by lpj desc / index asc, old by lpj asc / slot
    // asc).  New state of rank r is accepted iff r < M' and it beats the old state of rank r; the
    // test is monotone in r, so this equals the reference's pairwise walk (utils.py:295-319).
    int n_sub = 0;
    {
      int Mp = U < a.Mprime ? U : a.Mprime;
      if (Mp > 0) {
        double mn = DBL_MAX;
        _Pragma("unroll 1") for (int i = tid; i < S; i += T) mn = fmin(mn, s.lpj_uniq[i]);
        mn = -group_max<T>(-mn, s.red_d);
        // Only candidates above the worst old state can ever be accepted.  They are compacted (in index
        // order) and ranked among themselves: every other candidate is smaller than all of them, so the
        // ranks are those of the full list at E^2 instead of U^2 compares -- late in training E is a few.
        const double *ln = s.lpj_uniq + S;
        double *kn = s.vkeys, *ko = s.vkeys + ((a.Ctot + 1) & ~1);
        int E = 0;
        for (int base = 0; base < U; base += T) {
          const int i = base + tid;
          const bool f = i < U && ln[i] > mn;
          int tot;
          const int off = group_scan_flag<T>(f, tot, s.red_i);
          if (f) {
            kn[E + off] = order_key(ln[i], (uint32_t)i, a.mask_c, true);
            s.eidx[E + off] = i;
          }
          E += tot;
        }
        int acc_local = 0;
        if (E > 0) {  // uniform over the group
          _Pragma("unroll 1") for (int k = tid; k < S; k += T) ko[k] = order_key(s.lpj_uniq[k], (uint32_t)k, a.mask_s, false);
          gsync<T>();
          _Pragma("unroll 1") for (int e = tid; e < E; e += T) s.rnk[e] = count_greater(kn, E, kn[e]);
          _Pragma("unroll 1") for (int k = tid; k < S; k += T) {
            int r = count_less(ko, S, ko[k]);
            if (r < Mp) s.old_by_rank[r] = k;
          }
          gsync<T>();
          _Pragma("unroll 1") for (int e = tid; e < E; e += T) {
            int r = s.rnk[e];
            if (r >= Mp) continue;
            const int i = s.eidx[e];
            int k = s.old_by_rank[r];
            double v = ln[i];
            if (!(v > s.lpj_uniq[k])) continue;
            const uint32_t *src = s.st_uniq + (size_t)(S + i) * Hs;
            uint32_t *dst = s.st_uniq + (size_t)k * Hs;
            uint32_t *gdst = a.states + (n * (int64_t)S + k) * Hg;
            if (LIST) {
              for (int w = 0; w < Hw; ++w) dst[w] = src[w];
              list_to_bits(src, Hg, gdst);
            } else {
              for (int w = 0; w < Hw; ++w) { uint32_t x = src[w]; dst[w] = x; gdst[w] = x; }
            }
            s.rnk[e] = -1 - k;  // remembered; lpj written after the barrier (other threads still compare against it)
            ++acc_local;
          }
          gsync<T>();
          _Pragma("unroll 1") for (int e = tid; e < E; e += T) { int r = s.rnk[e]; if (r < 0) s.lpj_uniq[-1 - r] = ln[s.eidx[e]]; }
        }
        n_sub = group_sum_i<T>(acc_local, s.red_i);
        gsync<T>();
      }
    }
    // ---- write lpj row, free-energy contribution ------------------------------------------------
    {
      double *row = a.lpj + n * (int64_t)(S + a.S_perm);
      double mx = -DBL_MAX;
      _Pragma("unroll 1") for (int i = tid; i < S; i += T) { double v = s.lpj_uniq[i]; row[a.S_perm + i] = v; mx = fmax(mx, v); }
      if (a.S_perm > 0) { if (tid == 0) row[0] = lpj_perm; mx = fmax(mx, lpj_perm); }
      mx = group_max<T>(mx, s.red_d);
      double se = 0.0;
      _Pragma("unroll 1") for (int i = tid; i < S; i += T) se += exp(s.lpj_uniq[i] - mx);
      if (a.S_perm > 0 && tid == 0) se += exp(lpj_perm - mx);
      se = group_sum<T>(se, s.red_d);
      if (tid == 0) {
        acc_F += mx + log(se);
        acc_U += (double)U;
        acc_sub += (double)n_sub;
        acc_eval += (double)(S + U);
      }
    }
  }
  if (tid == 0 && a.acc) {
    atomicAdd(a.acc + 0, acc_F);
    atomicAdd(a.acc + 1, acc_U);
    atomicAdd(a.acc + 2, acc_sub);
    if (cnt_nan) atomicAdd(a.acc + 3, (double)cnt_nan);
    if (cnt_small) atomicAdd(a.acc + 4, (double)cnt_small);
    if (cnt_inf) atomicAdd(a.acc + 5, (double)cnt_inf);
    atomicAdd(a.acc + 6, acc_eval);
  }
}

// ---- launch: threads per datapoint group -------------------------------------------------------
// T in {32, 64, 128, 256}: the evaluator's hint for the work per datapoint (max(S, children per
// generation)); EVO_ESTEP_T overrides it (tuning runs).  Persistent grid = SMs x resident CTAs.
// Kernel variants: generic (every operator, injected candidates), FAST (randflip RNG path), and for
// evaluators that read index lists FAST+LIST followed by a FAST packed-bits launch over the datapoints
// the list form could not hold (device-side list, normally empty).
template <class Eval, int TT, bool FAST, bool LIST>
static int launch_estep_variant(const EstepDev &dv, Eval ev, int64_t items, cudaStream_t st) {
  size_t smem = estep_carve(dv, nullptr, nullptr) + ev.smem_bytes(TT, LIST);
  if (smem > 227 * 1024) {
    set_error("evo_estep: %zu bytes of shared memory per datapoint exceed 227 KB", smem);
    return EVO_ERR_UNSUPPORTED;
  }
  auto k = estep_kernel<Eval, TT, FAST, LIST>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  EVO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, TT, smem));
  if (per_sm < 1) {
    set_error("evo_estep: %zu bytes of shared memory per datapoint exceed the SM", smem);
    return EVO_ERR_UNSUPPORTED;
  }
  int64_t g = (int64_t)sm_count() * per_sm;
  if (g > items) g = items;
  if (g < 1) g = 1;
  k<<<(int)g, TT, smem, st>>>(dv, ev);
  EVO_LAUNCH_CHECK("estep");
  return EVO_OK;
}

// fused statistics not covered by the kernel that runs: every datapoint goes to the separate statistics pass
static __global__ void fs_rest_all_kernel(int32_t *rest, int64_t N) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i <= N; i += (int64_t)gridDim.x * blockDim.x)
    rest[i] = i == 0 ? (int32_t)N : (int32_t)(i - 1);
}
static int fs_rest_all(const EstepDev &dv, cudaStream_t st) {
  if (!dv.fs_rest) return EVO_OK;
  int64_t b = (dv.N + 256) / 256, cap = (int64_t)sm_count() * 8;
  fs_rest_all_kernel<<<(int)(b < cap ? b : cap), 256, 0, st>>>(dv.fs_rest, dv.N);
  EVO_LAUNCH_CHECK("fs_rest_all");
  return EVO_OK;
}

// randflip fast path over index lists with implicit children (estep_flip.cuh)
struct BscGramEval;
template <int TT>
static int launch_estep_flip(const EstepDev &dv, BscGramEval ev, cudaStream_t st);

template <class Eval, int TT>
static int launch_estep_T(const EstepDev &dv, Eval ev, bool fast, cudaStream_t st) {
  if constexpr (Eval::kList) {
    const int PB = dv.n_parents * (dv.n_generations > 1 ? dv.n_generations - 1 : 0);
    if (fast && dv.ovf_list && dv.H < 65535 && dv.S + PB < 65535 && dv.N < 2147483647LL && dv.n_children <= dv.H &&
        !getenv("EVO_ESTEP_BITS")) {
      EVO_CUDA(cudaMemsetAsync(dv.ovf_list, 0, sizeof(int32_t), st));
      if (dv.fs_rest) EVO_CUDA(cudaMemsetAsync(dv.fs_rest, 0, sizeof(int32_t), st));
      int rc = launch_estep_flip<TT>(dv, ev, st);
      if (rc) return rc;
      EstepDev db = dv;
      db.n_list = dv.ovf_list;
      db.ovf_list = nullptr;
      // full persistent grid: the number of overflowed datapoints is only known on the device
      return launch_estep_variant<Eval, TT, true, false>(db, ev, dv.N, st);
    }
  }
  EstepDev db = dv;
  db.ovf_list = nullptr;
  {
    int rc = fs_rest_all(dv, st);
    if (rc) return rc;
  }
  if (fast) return launch_estep_variant<Eval, TT, true, false>(db, ev, dv.N, st);
  return launch_estep_variant<Eval, TT, false, false>(db, ev, dv.N, st);
}

template <class Eval>
static int launch_estep_any(const EstepDev &dv, Eval ev, cudaStream_t st) {
  int work = dv.S > dv.Cg ? dv.S : dv.Cg;
  int T = ev.threads_hint(work);
  const bool fast = !dv.inj_children && dv.mutation == EVO_MUT_RANDFLIP && !dv.background && !dv.generic_only &&
                    dv.selection != EVO_SEL_ALL && !getenv("EVO_ESTEP_GENERIC");
  if (const char *e = getenv("EVO_ESTEP_T")) { int t = atoi(e); if (t > 0) T = t; }
  if (T <= 32) return launch_estep_T<Eval, 32>(dv, ev, fast, st);
  if (T <= 64) return launch_estep_T<Eval, 64>(dv, ev, fast, st);
  if (T <= 128) return launch_estep_T<Eval, 128>(dv, ev, fast, st);
  return launch_estep_T<Eval, 256>(dv, ev, fast, st);
}

}  // namespace evo
