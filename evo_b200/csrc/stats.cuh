// Shared between bsc.cu (statistics pass) and dense.cu (tensor-core Wp): the hand-over of the sparse
// <s>_n from the per-datapoint pass to whichever kernel accumulates Wp = sum_n <s>_n y_n' (bsc.py:204-206).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace evo {

constexpr int WP_LCAP = 64;   // listed entries per datapoint
constexpr int WP_CW = 64;     // observables per Wp slice
constexpr int WP_WARPS = 32;  // warps of the Wp pass; warp w owns the latents [w*hb, (w+1)*hb), hb = ceil(H/32)
constexpr int WP_OFF = 40;
constexpr double WP_TC_SCALE = 16384.0;  // <s> in [0, 1] -> fp16 hi + lo planes of 2^14 <s>

struct WpLists {
  // (a) f64 path: per-datapoint lists for bsc_wp_kernel
  uint8_t *off;   // [N][WP_WARPS + 1 (+pad to 40)]: the (ascending-h) list of datapoint n restricted to warp w
                  // is [off[w], off[w+1]); off[WP_WARPS] = number of listed entries
  uint16_t *h;    // [N][WP_LCAP]
  double *e;      // [N][WP_LCAP]
  // (b) tensor-core path: transposed fp16 planes E'[h][n] (zeroed by the caller), row stride et_ld
  __half *et_hi, *et_lo;
  int64_t et_ld;
};

// the statistics pass (bsc.cu); lists select where <s>_n goes (all null: straight to Wp by atomics)
int bsc_suffstats_pass1(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H, int S,
                        int S_perm, double pre1, double pil_bar, double *stats, float *Es_out, WpLists lists,
                        cudaStream_t st, const int32_t *n_list = nullptr, const double *theta_dev = nullptr);

// the <s> planes of a tensor-core Wp workspace (dense.cu geometry): et_hi / et_lo / et_ld of `out`
int wp_tc_planes(void *workspace, int64_t N, int D, int H, WpLists *out);

}  // namespace evo
