// Fused E-step, randflip fast path (BSC Gram evaluator, RNG mode, no background unit): the same algorithm and the
// same Philox streams as estep_kernel (estep.cuh) -- lpj of K^n, generations of select -> mutate -> de-duplicate ->
// evaluate, top-K replacement, free-energy tail (_models.py:454-565, eas.py:153-313, utils.py:231-337) -- with a
// representation built for single-bit children:
//   * explicit states (K^n, the parents selected in generations >= 1) are index LISTS: 31 ascending uint16 latents and
//     a count, 64 bytes whatever H is (132 bytes per packed state at H = 1024);
//   * a child is IMPLICIT: (explicit parent row, flipped position) -- 4 bytes; it is materialised only if it becomes a
//     parent or is accepted into K^n;
//   * every state carries its sums sb = sum_{h in s} b_h and sg = sum_{h,h' in s} G_hh' (lpj = pre1 (|y|^2 - 2 sb + sg)
//     + pil_bar |s|), so a child costs O(|s|) loads: sb' = sb +- b_p, sg' = sg + G_pp +- 2 sum_{h in parent} G_ph
//     instead of the O(|s|^2) walk of a fresh evaluation (K^n itself is evaluated in full: theta changed);
//   * de-duplication hashes SETS: h(s) = sum_{h in s} mix(h) mod 2^32, so h(child) = h(parent) +- mix(p); a table hit
//     on the hash is always verified exactly (merge walk over the two parent lists);
//   * shared memory per datapoint decides how many datapoints an SM holds (the kernel is latency-bound): arrays that
//     are dead after the generations (sb, sg, hash) are reused by vary_Kn and by the fused statistics, and sb / sg are
//     not kept for the children of the last generation.
// Datapoints whose K^n or children need more than 31 active latents are handed, untouched, to the packed-bits kernel
// through `ovf_list` (same protocol as before).  Results do not depend on which kernel handled a datapoint: K sets
// bit-identical, lpj equal to the full evaluation within rounding (<= 1e-12 relative; tests/test_evolve_gpu.py chains
// this kernel to the generic kernel, to injected children and to the NumPy oracle).
#pragma once
#include "estep.cuh"

namespace evo {

constexpr int FL_CAP = 31, FL_WORDS = 16, FL_STRIDE = 17;  // list row: 32 x uint16 (+ one pad word: conflict-free rows)

struct FlipSmem {
  double *lpj;               // [S + Ctot]
  double *sb, *sg;           // [S + Ctot - Cg]: the children of the LAST generation never become parents
  double *keys;              // [PK + 2] selection keys; later vkeys of vary_Kn
  double *vkeys;             // [Ctot + S + 2] (aliases the hash table)
  double *red_d;             // [8]
  uint32_t *hash;            // [S + Ctot] low half of the 64-bit set hash (additive mod 2^32; hits are verified exactly)
  double *es_u;              // fused statistics: <s> over the union (in the dead table region, behind old_by_rank)
  int es_cap;
  uint32_t *rows;            // [S + PB][FL_STRIDE]
  int *table;                // [HT]
  int *pool;                 // [PK]
  int *slot;                 // [Cg] flip positions, then hash slots
  int *parents;              // [n_parents] uniq index of the selected parents
  int *par_row;              // [n_parents] their explicit row
  int *red_i, *sh;           // [8], [8]
  int *rnk, *old_by_rank, *eidx;  // vary_Kn (alias the table region)
  uint16_t *nw_par, *nw_pos;      // [Ctot] implicit new states: explicit parent row, flipped position
  unsigned char *cnt;        // [S + Ctot] |s|
  unsigned char *redisc;     // [S + Ctot]
  unsigned char *status;     // [Cg]
};

__host__ __device__ inline size_t flip_carve(const EstepDev &a, unsigned char *base, FlipSmem *s) {
  size_t off = 0;
  const size_t U = (size_t)a.S + a.Ctot;
  const int PB = a.n_parents * (a.n_generations > 1 ? a.n_generations - 1 : 0);
#define CARVE(field, type, count)                                  \
  do {                                                             \
    if (s) s->field = reinterpret_cast<type *>(base + off);       \
    off = align16(off + sizeof(type) * (size_t)(count));          \
  } while (0)
  CARVE(lpj, double, U);
  // [sb | sg | hash] are dead once the generations are over: vary_Kn's scratch (vkeys, rnk, eidx) aliases them when it
  // fits, and the fused statistics carve their scratch rows from the same span
  const size_t dead0 = off;
  CARVE(sb, double, U - a.Cg);
  CARVE(sg, double, U - a.Cg);
  CARVE(hash, uint32_t, U);
  const size_t dead1 = off;
  CARVE(keys, double, a.PK + 2);
  CARVE(red_d, double, 8);
  CARVE(rows, uint32_t, ((size_t)a.S + PB) * FL_STRIDE);
  const size_t n_vk = (size_t)((a.Ctot + 1) & ~1) + a.S + 2, n_ct = a.Ctot > 0 ? a.Ctot : 1;
  const bool alias = align16(sizeof(double) * n_vk) + 2 * align16(sizeof(int) * n_ct) <= dead1 - dead0;
  if (alias) {
    const size_t keep = off;
    off = dead0;
    CARVE(vkeys, double, n_vk);
    CARVE(rnk, int, n_ct);
    CARVE(eidx, int, n_ct);
    off = keep;
  }
  {
    const size_t base_off = off;
    CARVE(table, int, a.HT);
    const size_t end_gen = off;
    off = base_off;
    if (!alias) {
      CARVE(vkeys, double, n_vk);
      CARVE(rnk, int, n_ct);
      CARVE(eidx, int, n_ct);
    }
    CARVE(old_by_rank, int, a.S);
    if (s) {
      s->es_u = reinterpret_cast<double *>(base + off);
      s->es_cap = end_gen > off ? (int)((end_gen - off) / sizeof(double)) : 0;
    }
    if (off < end_gen) off = end_gen;
  }
  CARVE(pool, int, a.PK);
  CARVE(slot, int, a.Cg > 0 ? a.Cg : 1);
  CARVE(parents, int, a.n_parents > 0 ? a.n_parents : 1);
  CARVE(par_row, int, a.n_parents > 0 ? a.n_parents : 1);
  CARVE(red_i, int, 8);
  CARVE(sh, int, 8);
  CARVE(nw_par, uint16_t, a.Ctot > 0 ? a.Ctot : 1);
  CARVE(nw_pos, uint16_t, a.Ctot > 0 ? a.Ctot : 1);
  CARVE(cnt, unsigned char, U);
  CARVE(redisc, unsigned char, U);
  CARVE(status, unsigned char, a.Cg > 0 ? a.Cg : 1);
#undef CARVE
  return off;
}

// ---- 31-entry list rows --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mix_latent(uint32_t h) {  // splitmix64 finaliser of the latent index
  uint64_t z = ((uint64_t)h + 1ull) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// packed bits -> list row by ONE thread; returns the count (> FL_CAP = does not fit; the row is then unspecified)
__device__ __forceinline__ int flist_from_bits_thread(const uint32_t *src, int Hg, uint32_t *row) {
  uint16_t *e = reinterpret_cast<uint16_t *>(row);
#pragma unroll
  for (int w = 0; w < FL_WORDS; ++w) row[w] = 0xFFFFFFFFu;
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
        if (cnt < FL_CAP) e[cnt] = (uint16_t)(((w0 + q) << 5) + bp);
        ++cnt;
      }
    }
  }
  e[31] = (uint16_t)(cnt <= FL_CAP ? cnt : FL_CAP);
  return cnt;
}
// the same by one WARP (long rows); every lane returns the count
__device__ __forceinline__ int flist_from_bits_warp(const uint32_t *src, int Hg, uint32_t *row) {
  uint16_t *e = reinterpret_cast<uint16_t *>(row);
  const int lane = threadIdx.x & 31;
  e[lane] = 0xFFFFu;
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
      const int bp = __ffs(word) - 1;
      word &= word - 1;
      if (pos < FL_CAP) e[pos] = (uint16_t)(((w0 + lane) << 5) + bp);
      ++pos;
    }
    base += __shfl_sync(0xffffffffu, pre, 31);
  }
  __syncwarp();
  if (lane == 0) e[31] = (uint16_t)(base <= FL_CAP ? base : FL_CAP);
  __syncwarp();
  return base;
}
// dst <- src with position p toggled (sorted insert / remove); the caller guarantees the result fits
__device__ __forceinline__ void flist_toggle_copy(const uint32_t *src, int p, uint32_t *dst) {
  const uint16_t *s = reinterpret_cast<const uint16_t *>(src);
  uint16_t *d = reinterpret_cast<uint16_t *>(dst);
  const int cnt = s[31];
  int i = 0, j = 0;
  bool present = false;
  for (; i < cnt; ++i) {
    const int h = s[i];
    if (h == p) { present = true; continue; }
    if (!present && h > p && j == i) d[j++] = (uint16_t)p;  // insert before the first larger entry
    d[j++] = (uint16_t)h;
  }
  if (!present && j == cnt) d[j++] = (uint16_t)p;  // larger than every entry
  for (int k = j; k < 31; ++k) d[k] = 0xFFFFu;
  d[31] = (uint16_t)j;
}
// list (row with position p toggled; p < 0: the row itself) -> packed bits (Hg words), one thread
__device__ __forceinline__ void flist_to_bits_toggled(const uint32_t *row, int p, int Hg, uint32_t *dst) {
  const uint16_t *e = reinterpret_cast<const uint16_t *>(row);
  const int cnt = e[31];
  int q = 0;
  for (int w = 0; w < Hg; ++w) {
    uint32_t x = 0u;
    while (q < cnt && (e[q] >> 5) == w) { x |= 1u << (e[q] & 31); ++q; }
    if (p >= 0 && (p >> 5) == w) x ^= 1u << (p & 31);
    dst[w] = x;
  }
}
// is  rowA (+) rowB == {p} (+) {q}  as sets?  (p, q < 0: absent)   rowA != rowB as rows of distinct states
__device__ __forceinline__ bool symdiff_is(const uint32_t *rowA, const uint32_t *rowB, int p, int q) {
  const uint16_t *a = reinterpret_cast<const uint16_t *>(rowA), *b = reinterpret_cast<const uint16_t *>(rowB);
  if (p == q) p = q = -1;  // the same flip on both sides cancels: equal iff the rows are equal
  const int na = a[31], nb = b[31];
  int i = 0, j = 0, nd = 0;
  bool gotp = (p < 0), gotq = (q < 0);
  while (i < na || j < nb) {
    const int x = i < na ? a[i] : 0x10000, y = j < nb ? b[j] : 0x10000;
    if (x == y) { ++i; ++j; continue; }
    const int d = x < y ? x : y;
    if (x < y) ++i; else ++j;
    if (d == p && !gotp) gotp = true;
    else if (d == q && !gotq) gotq = true;
    else return false;
    if (++nd > 2) return false;
  }
  return gotp && gotq;
}

// state idx of the unique set: explicit row and pending flip (-1 = none)
struct FlipRef { int row, pos; };
__device__ __forceinline__ FlipRef flip_ref(const FlipSmem &s, int S, int idx) {
  FlipRef r;
  if (idx < S) { r.row = idx; r.pos = -1; } else { r.row = s.nw_par[idx - S]; r.pos = s.nw_pos[idx - S]; }
  return r;
}
__device__ __forceinline__ bool flip_equal(const FlipSmem &s, int S, int i, int j) {
  if (s.hash[i] != s.hash[j] || s.cnt[i] != s.cnt[j]) return false;
  const FlipRef a = flip_ref(s, S, i), b = flip_ref(s, S, j);
  if (a.row == b.row) return a.pos == b.pos;
  return symdiff_is(s.rows + (size_t)a.row * FL_STRIDE, s.rows + (size_t)b.row * FL_STRIDE, a.pos, b.pos);
}
// concurrent insert of uniq index idx; equal states converge on one slot holding the smallest index.  Returns the slot.
__device__ __forceinline__ int flip_insert(const EstepDev &a, const FlipSmem &s, int idx) {
  const uint32_t mask = (uint32_t)a.HT - 1u;
  uint32_t p = s.hash[idx] & mask;
  for (int it = 0; it < a.HT; ++it) {
    int cur = s.table[p];
    if (cur == TAG_EMPTY) {
      cur = atomicCAS(&s.table[p], TAG_EMPTY, idx);
      if (cur == TAG_EMPTY) return (int)p;
    }
    if (cur == idx || flip_equal(s, a.S, cur, idx)) {
      atomicMin(&s.table[p], idx);
      return (int)p;
    }
    p = (p + 1) & mask;
  }
  return -1;
}

// ---- fused BSC statistics (bsc.py:176-223) in the E-step tail ---------------------------------------------------------
// K^n (lists) and its final log-joints are still in shared memory: posterior weights, sigma, <s>_n over the union of
// the live states' latents, the Wq pair sums and the fp16 hi / lo planes of <s>_n for the tensor-core Wp GEMM are
// formed here instead of by a second pass that re-reads lpj and the K sets from HBM.  Slots replaced in this E-step are
// implicit (parent row, position): they are materialised into scratch rows carved from the dead sb / sg arrays first.
// Returns this datapoint's sigma term (valid on thread 0); a datapoint that does not fit (scratch rows, union
// capacity) is appended to fs_rest untouched.
constexpr double FS_Q_SKIP = 1e-18;  // = Q_SKIP of the statistics pass (bsc.cu)
template <int T>
__device__ __forceinline__ double flip_fused_stats(const EstepDev &a, const FlipSmem &s, const BscGramEval &ev, int64_t n,
                                                   double mx, double se, double lpj_perm) {
  const int tid = threadIdx.x, S = a.S, Hw = a.Hw;
  uint32_t *scr = reinterpret_cast<uint32_t *>(s.sb);  // sb, sg, hash (and vary_Kn's scratch over them): dead by now
  const int R = (int)((reinterpret_cast<unsigned char *>(s.keys) - reinterpret_cast<unsigned char *>(s.sb)) / (FL_STRIDE * 4));
  uint32_t *ubm = reinterpret_cast<uint32_t *>(s.keys);                        // union bitmap [Hw]
  const int keys_bytes = (a.PK + 2) * 8, bm_bytes = (4 * Hw + 7) & ~7;
  uint16_t *act = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(s.keys) + bm_bytes);
  double *es_u = s.es_u;                                                       // the dead hash-table region
  int capU = keys_bytes > bm_bytes ? (keys_bytes - bm_bytes) / 2 : 0;
  if (capU > s.es_cap) capU = s.es_cap;
  int *frow = s.old_by_rank;  // scratch row of a replaced slot, -1 = the slot's own row
  bool ok = capU >= 32;
  // ---- scratch rows for the replaced slots ----------------------------------------------------------------------------
  int nrep = 0;
  for (int base = 0; base < S; base += T) {
    const int k = base + tid;
    const bool f = k < S && s.pool[k] > 0;
    int tot;
    const int off = group_scan_flag<T>(f, tot, s.red_i);
    if (k < S) frow[k] = f ? nrep + off : -1;
    nrep += tot;
    gsync<T>();
  }
  if (nrep > R) ok = false;
  if (ok) {
    _Pragma("unroll 1") for (int k = tid; k < S; k += T) {
      if (frow[k] >= 0) {
        const int i = s.pool[k] - 1;
        flist_toggle_copy(s.rows + (size_t)s.nw_par[i] * FL_STRIDE, s.nw_pos[i], scr + (size_t)frow[k] * FL_STRIDE);
      }
    }
    _Pragma("unroll 1") for (int w = tid; w < Hw; w += T) ubm[w] = 0u;
  }
  gsync<T>();
  // ---- posterior weights, union bitmap, sigma ----------------------------------------------------------------------------
  const double inv = 1.0 / se;
  double sg = 0.0;
  if (ok) {
    _Pragma("unroll 1") for (int k = tid; k < S; k += T) {
      const double qk = exp(s.lpj[k] - mx) * inv;
      if (!(qk > FS_Q_SKIP)) continue;
      const uint16_t *e = reinterpret_cast<const uint16_t *>(frow[k] >= 0 ? scr + (size_t)frow[k] * FL_STRIDE : s.rows + (size_t)k * FL_STRIDE);
      const int na = e[31];
      for (int x = 0; x < na; ++x) atomicOr(&ubm[e[x] >> 5], 1u << (e[x] & 31));
      sg += qk * (s.lpj[k] - ev.pil_bar_now() * (double)na) / ev.pre1_now();
    }
    if (a.S_perm > 0 && tid == 0) sg += exp(lpj_perm - mx) * inv * lpj_perm / ev.pre1_now();  // all-zero state: |y|^2
  }
  gsync<T>();
  // ---- ascending list of the union (warp 0) -----------------------------------------------------------------------------------
  if (ok && tid < 32) {
    int nu = 0;
    for (int w0 = 0; w0 < Hw; w0 += 32) {
      uint32_t word = (w0 + tid < Hw) ? ubm[w0 + tid] : 0u;
      const int c = __popc(word);
      int pre = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, pre, o); if (tid >= o) pre += t; }
      int pos = nu + pre - c;
      while (word) {
        const int bp = __ffs(word) - 1;
        word &= word - 1;
        if (pos < capU) act[pos] = (uint16_t)(((w0 + tid) << 5) + bp);
        ++pos;
      }
      nu += __shfl_sync(0xffffffffu, pre, 31);
    }
    if (tid == 0) s.sh[1] = nu;
  }
  gsync<T>();
  const int nu = ok ? s.sh[1] : 0;
  if (nu > capU) ok = false;
  if (!ok) {
    if (tid == 0) a.fs_rest[1 + atomicAdd(a.fs_rest, 1)] = (int32_t)n;
    return 0.0;
  }
  _Pragma("unroll 1") for (int i = tid; i < nu; i += T) es_u[i] = 0.0;
  gsync<T>();
  // ---- <s> over the union and the Wq pair sums ------------------------------------------------------------------------------------
  _Pragma("unroll 1") for (int k = tid; k < S; k += T) {
    const double qk = exp(s.lpj[k] - mx) * inv;
    if (!(qk > FS_Q_SKIP)) continue;
    const uint16_t *e = reinterpret_cast<const uint16_t *>(frow[k] >= 0 ? scr + (size_t)frow[k] * FL_STRIDE : s.rows + (size_t)k * FL_STRIDE);
    const int na = e[31];
    for (int x = 0; x < na; ++x) {
      const int h = e[x];
      int lo = 0, hi = nu;  // position of h in the union list
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (act[mid] < h) lo = mid + 1; else hi = mid; }
      atomicAdd(&es_u[lo], qk);
      double *wrow = a.fs_Wq + (int64_t)h * a.H;
      for (int y = 0; y <= x; ++y) atomicAdd(wrow + e[y], qk);  // h' <= h
    }
  }
  gsync<T>();
  _Pragma("unroll 1") for (int i = tid; i < nu; i += T) {
    const int h = act[i];
    const double v = es_u[i];
    atomicAdd(a.fs_pies + h, v);
    const double sv = v * WP_TC_SCALE;
    const __half hi = __double2half(sv);
    a.fs_hi[(int64_t)h * a.fs_ld + n] = hi;
    a.fs_lo[(int64_t)h * a.fs_ld + n] = __double2half(sv - (double)__half2float(hi));
  }
  return group_sum<T>(sg, s.red_d);
}

template <int T>
__global__ void __launch_bounds__(T) estep_flip_kernel(EstepDev a, BscGramEval ev) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FlipSmem s;
  flip_carve(a, smem_raw, &s);
  const int tid = threadIdx.x;
  const int Hg = a.Hw, S = a.S, Hm = a.H, C = a.n_children;
  double acc_F = 0.0, acc_U = 0.0, acc_sub = 0.0, acc_eval = 0.0, acc_sig = 0.0;
  unsigned long long cnt_nan = 0, cnt_small = 0, cnt_inf = 0;

  for (int64_t n = blockIdx.x; n < a.N; n += gridDim.x) {
    RngCtx rng(a.seed, a.n_offset + n, th_iteration(a));
    const double *bg = ev.B + n * (int64_t)ev.H;
    const double yn2 = __ldg(ev.ynorm2 + n);
    gsync<T>();  // previous datapoint fully done with shared memory
    // ---- load K^n as lists; evaluate it in full (sb, sg, lpj, hash) --------------------------------------------
    if (tid == 0) s.sh[0] = 0;
    _Pragma("unroll 1") for (int i = tid; i < a.HT; i += T) s.table[i] = TAG_EMPTY;
    _Pragma("unroll 1") for (int i = tid; i < S + a.Ctot; i += T) s.redisc[i] = 0;
    gsync<T>();
    {
      const uint32_t *src = a.states + n * (int64_t)S * Hg;
      if (Hg <= 64) {
        _Pragma("unroll 1") for (int i = tid; i < S; i += T)
          if (flist_from_bits_thread(src + (size_t)i * Hg, Hg, s.rows + (size_t)i * FL_STRIDE) > FL_CAP) s.sh[0] = 1;
      } else {
        for (int i = tid >> 5; i < S; i += T / 32)
          if (flist_from_bits_warp(src + (size_t)i * Hg, Hg, s.rows + (size_t)i * FL_STRIDE) > FL_CAP && (tid & 31) == 0) s.sh[0] = 1;
      }
    }
    gsync<T>();
    if (s.sh[0]) {  // a state of K^n has more than FL_CAP active latents: not for this kernel
      if (tid == 0) {
        a.ovf_list[1 + atomicAdd(a.ovf_list, 1)] = (int32_t)n;
        if (a.fs_rest) a.fs_rest[1 + atomicAdd(a.fs_rest, 1)] = (int32_t)n;
      }
      continue;
    }
    unsigned fl0 = 0;
    _Pragma("unroll 1") for (int i = tid; i < S; i += T) {
      const uint16_t *e = reinterpret_cast<const uint16_t *>(s.rows + (size_t)i * FL_STRIDE);
      const int na = e[31];
      double sb = 0.0, sg = 0.0;
      uint32_t hs = 0;
      for (int x = 0; x < na; ++x) {
        const int h = e[x];
        hs += (uint32_t)mix_latent((uint32_t)h);
        sb += __ldg(bg + h);
        const double *Gr = ev.G + (int64_t)h * ev.H;
        double t = 0.0;
#pragma unroll 4
        for (int y = 0; y < x; ++y) t += __ldg(Gr + e[y]);
        sg += __ldg(Gr + h) + 2.0 * t;
      }
      s.sb[i] = sb; s.sg[i] = sg; s.hash[i] = hs; s.cnt[i] = (unsigned char)na;
      s.lpj[i] = lpj_sanitise(__fma_rn(ev.pre1_now(), __dadd_rn(__fma_rn(-2.0, sb, yn2), sg), __dmul_rn(ev.pil_bar_now(), (double)na)), fl0);
      s.pool[i] = i;
    }
    fl0 = group_or<T>(fl0, s.red_i);
    if (tid == 0 && fl0) { if (fl0 & 1) ++cnt_nan; else if (fl0 & 2) ++cnt_small; else ++cnt_inf; }
    double lpj_perm = 0.0;
    if (a.S_perm > 0) {
      unsigned f0 = 0;
      lpj_perm = lpj_sanitise(ev.pre1_now() * yn2, f0);
      if (tid == 0 && f0) { if (f0 & 1) ++cnt_nan; else if (f0 & 2) ++cnt_small; else ++cnt_inf; }
    }
    gsync<T>();
    _Pragma("unroll 1") for (int i = tid; i < S; i += T) flip_insert(a, s, i);
    gsync<T>();

    int U = 0, pool_n = S;
    for (int g = 0; g < a.n_generations; ++g) {
      // ---- parent selection (eas.py:138-150): identical keys and ranking to estep_kernel ------------------------
      const int take = pool_n < a.n_parents ? pool_n : a.n_parents;
      {
        double mn = DBL_MAX;
        if (a.selection == EVO_SEL_FIT) {
          _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) mn = fmin(mn, s.lpj[s.pool[i]]);
          mn = -group_max<T>(-mn, s.red_d);
        }
        const double shift = 2.0 * fmin(mn, 0.0);
        _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) {
          const uint4 r = rng.draw(RNG_SELECT, g, (uint32_t)i);
          const double u = u01(r.x, r.y);
          double key;
          if (a.selection == EVO_SEL_FIT) {
            const double w = s.lpj[s.pool[i]] - shift;
            key = (w > 0.0) ? (double)(__logf((float)u) / (float)w) : -DBL_MAX;
          } else {
            key = u;
          }
          s.keys[i] = order_key(key, (uint32_t)i, a.mask_p, true);
        }
        gsync<T>();
        _Pragma("unroll 1") for (int i = tid; i < pool_n; i += T) {
          const int r = count_greater(s.keys, pool_n, s.keys[i]);
          if (r < take) s.parents[r] = s.pool[i];
        }
        gsync<T>();
      }
      // ---- explicit rows of the parents (new states are materialised into this generation's buffer rows) and the
      //      C distinct flip positions of every parent (eas.py:27-42) --------------------------------------------------
      _Pragma("unroll 1") for (int p = tid; p < take; p += T) {
        const int u = s.parents[p];
        int row = u;
        if (u >= S) {
          row = S + (g - 1) * a.n_parents + p;
          flist_toggle_copy(s.rows + (size_t)s.nw_par[u - S] * FL_STRIDE, s.nw_pos[u - S], s.rows + (size_t)row * FL_STRIDE);
        }
        s.par_row[p] = row;
      }
      draw_flip_positions<T>(rng, (uint32_t)g, take, C, Hm, s.slot, s.red_i);
      const int nc = take * C;
      const bool keep_sums = g + 1 < a.n_generations;  // only states that can still become parents need sb / sg
      // ---- children: implicit (parent row, position), sums and lpj from the parent's in O(|s|) ----------------------
      unsigned fl = 0;
      _Pragma("unroll 1") for (int c = tid; c < nc; c += T) {
        const int p = c / C, pu = s.parents[p], prow = s.par_row[p], pos = s.slot[c];
        const uint16_t *e = reinterpret_cast<const uint16_t *>(s.rows + (size_t)prow * FL_STRIDE);
        const int na = e[31];
        const double *Gr = ev.G + (int64_t)pos * ev.H;
        double t = 0.0;
        bool present = false;
#pragma unroll 4
        for (int x = 0; x < na; ++x) {
          const int h = e[x];
          present |= (h == pos);
          t += __ldg(Gr + h);
        }
        const int idx = S + U + c;
        unsigned char stt = 1;
        if (!present && na >= FL_CAP) { s.sh[0] = 1; stt = 0; }
        const int nn = present ? na - 1 : na + 1;
        const double gpp = __ldg(Gr + pos), bp = __ldg(bg + pos);
        const double sb = present ? s.sb[pu] - bp : s.sb[pu] + bp;
        const double sg = present ? (s.sg[pu] + gpp) - 2.0 * t : (s.sg[pu] + gpp) + 2.0 * t;
        if (keep_sums) { s.sb[idx] = sb; s.sg[idx] = sg; }
        s.cnt[idx] = (unsigned char)nn;
        s.hash[idx] = present ? s.hash[pu] - (uint32_t)mix_latent((uint32_t)pos) : s.hash[pu] + (uint32_t)mix_latent((uint32_t)pos);
        s.nw_par[idx - S] = (uint16_t)prow; s.nw_pos[idx - S] = (uint16_t)pos;
        unsigned f1 = 0;
        s.lpj[idx] = lpj_sanitise(__fma_rn(ev.pre1_now(), __dadd_rn(__fma_rn(-2.0, sb, yn2), sg), __dmul_rn(ev.pil_bar_now(), (double)nn)), f1);
        if (a.S_perm > 0 && nn == 0) stt = 0;  // the permanent all-zero state
        s.status[c] = stt ? (unsigned char)(1u | (f1 << 4)) : 0;  // reset flags ride along: only new unique states count
      }
      gsync<T>();
      if (s.sh[0]) break;  // a child does not fit the list form: the packed-bits kernel redoes this datapoint
      // ---- de-duplicate against (K^n, earlier new) and among the children: first occurrence wins ---------------------
      _Pragma("unroll 1") for (int c = tid; c < nc; c += T)
        if (s.status[c]) s.slot[c] = flip_insert(a, s, S + U + c);
      gsync<T>();
      int u_g = 0;
      for (int base = 0; base < nc; base += T) {
        const int c = base + tid, idx = S + U + c;
        bool win = false;
        int sl = -1, owner = -1;
        if (c < nc && s.status[c]) { sl = s.slot[c]; owner = sl >= 0 ? s.table[sl] : -1; win = (owner == idx); }
        if (c < nc && s.status[c] && owner >= 0 && owner < S + U) s.redisc[owner] = 1;  // reproduced an existing state
        // winners move down to S + U + u_g + off (<= their provisional place): read, barrier, write
        double v_l = 0, v_b = 0, v_g = 0;
        uint32_t v_h = 0;
        unsigned v_pp = 0;
        unsigned char v_c = 0;
        if (win) { fl |= (unsigned)(s.status[c] >> 4);
                   v_l = s.lpj[idx]; v_h = s.hash[idx]; v_c = s.cnt[idx];
                   if (keep_sums) { v_b = s.sb[idx]; v_g = s.sg[idx]; }
                   v_pp = ((unsigned)s.nw_par[idx - S] << 16) | s.nw_pos[idx - S]; }
        int tot;
        const int off = group_scan_flag<T>(win, tot, s.red_i);
        gsync<T>();
        if (win) {
          const int dst = S + U + u_g + off;
          s.lpj[dst] = v_l; s.hash[dst] = v_h; s.cnt[dst] = v_c;
          if (keep_sums) { s.sb[dst] = v_b; s.sg[dst] = v_g; }
          s.nw_par[dst - S] = (uint16_t)(v_pp >> 16); s.nw_pos[dst - S] = (uint16_t)(v_pp & 0xFFFFu);
          s.table[sl] = dst;  // the entry follows its state
        }
        u_g += tot;
        gsync<T>();
      }
      {  // lpj_reset_check counts one event per evaluated batch (the new unique states of this generation)
        fl = group_or<T>(fl, s.red_i);
        if (tid == 0 && fl && u_g > 0) { if (fl & 1) ++cnt_nan; else if (fl & 2) ++cnt_small; else ++cnt_inf; }
      }
      // ---- next pool: this round's new states, then the re-discovered old ones --------------------------------------
      int pn = 0;
      _Pragma("unroll 1") for (int i = tid; i < u_g && i < a.PK; i += T) s.pool[i] = S + U + i;
      pn = u_g < a.PK ? u_g : a.PK;
      for (int base = 0; base < S + U; base += T) {
        const int i = base + tid;
        const bool f = (i < S + U) && s.redisc[i];
        int tot;
        const int off = group_scan_flag<T>(f, tot, s.red_i);
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
    if (s.sh[0]) {  // nothing of this datapoint has been written yet
      if (tid == 0) {
        a.ovf_list[1 + atomicAdd(a.ovf_list, 1)] = (int32_t)n;
        if (a.fs_rest) a.fs_rest[1 + atomicAdd(a.fs_rest, 1)] = (int32_t)n;
      }
      continue;
    }
    if (a.fs_rest) {  // slots replaced below are recorded here (the parent pool is no longer needed)
      _Pragma("unroll 1") for (int k = tid; k < S; k += T) s.pool[k] = 0;
    }

    // ---- vary_Kn (utils.py:295-319): identical ranking to estep_kernel ---------------------------------------------------
    int n_sub = 0;
    {
      const int Mp = U < a.Mprime ? U : a.Mprime;
      if (Mp > 0) {
        double mn = DBL_MAX;
        _Pragma("unroll 1") for (int i = tid; i < S; i += T) mn = fmin(mn, s.lpj[i]);
        mn = -group_max<T>(-mn, s.red_d);
        const double *ln = s.lpj + S;
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
        if (E > 0) {
          _Pragma("unroll 1") for (int k = tid; k < S; k += T) ko[k] = order_key(s.lpj[k], (uint32_t)k, a.mask_s, false);
          gsync<T>();
          _Pragma("unroll 1") for (int e = tid; e < E; e += T) s.rnk[e] = count_greater(kn, E, kn[e]);
          _Pragma("unroll 1") for (int k = tid; k < S; k += T) {
            const int r = count_less(ko, S, ko[k]);
            if (r < Mp) s.old_by_rank[r] = k;
          }
          gsync<T>();
          _Pragma("unroll 1") for (int e = tid; e < E; e += T) {
            const int r = s.rnk[e];
            if (r >= Mp) continue;
            const int i = s.eidx[e], k = s.old_by_rank[r];
            if (!(ln[i] > s.lpj[k])) continue;
            flist_to_bits_toggled(s.rows + (size_t)s.nw_par[i] * FL_STRIDE, s.nw_pos[i], Hg, a.states + (n * (int64_t)S + k) * Hg);
            s.rnk[e] = -1 - k;  // lpj written after the barrier (other threads still compare against the old value)
            if (a.fs_rest) s.pool[k] = i + 1;
            ++acc_local;
          }
          gsync<T>();
          _Pragma("unroll 1") for (int e = tid; e < E; e += T) { const int r = s.rnk[e]; if (r < 0) s.lpj[-1 - r] = ln[s.eidx[e]]; }
        }
        n_sub = group_sum_i<T>(acc_local, s.red_i);
        gsync<T>();
      }
    }
    // ---- write lpj row, free-energy contribution --------------------------------------------------------------------------
    {
      double *row = a.lpj + n * (int64_t)(S + a.S_perm);
      double mx = -DBL_MAX;
      _Pragma("unroll 1") for (int i = tid; i < S; i += T) { const double v = s.lpj[i]; row[a.S_perm + i] = v; mx = fmax(mx, v); }
      if (a.S_perm > 0) { if (tid == 0) row[0] = lpj_perm; mx = fmax(mx, lpj_perm); }
      mx = group_max<T>(mx, s.red_d);
      double se = 0.0;
      _Pragma("unroll 1") for (int i = tid; i < S; i += T) se += exp(s.lpj[i] - mx);
      if (a.S_perm > 0 && tid == 0) se += exp(lpj_perm - mx);
      se = group_sum<T>(se, s.red_d);
      if (tid == 0) {
        acc_F += mx + log(se);
        acc_U += (double)U;
        acc_sub += (double)n_sub;
        acc_eval += (double)(S + U);
      }
      if (a.fs_rest) acc_sig += flip_fused_stats<T>(a, s, ev, n, mx, se, lpj_perm);
    }
  }
  if (tid == 0 && a.fs_rest && acc_sig != 0.0) atomicAdd(a.fs_sig, acc_sig);
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

template <int TT>
static int launch_estep_flip(const EstepDev &dv, BscGramEval ev, cudaStream_t st) {
  const size_t smem = flip_carve(dv, nullptr, nullptr);
  if (smem > 227 * 1024) return EVO_ERR_UNSUPPORTED;
  auto k = estep_flip_kernel<TT>;
  if (smem > 48 * 1024) EVO_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  EVO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, TT, smem));
  if (per_sm < 1) return EVO_ERR_UNSUPPORTED;
  int64_t g = (int64_t)sm_count() * per_sm;
  if (g > dv.N) g = dv.N;
  if (g < 1) g = 1;
  k<<<(int)g, TT, smem, st>>>(dv, ev);
  EVO_LAUNCH_CHECK("estep_flip");
  return EVO_OK;
}

}  // namespace evo
