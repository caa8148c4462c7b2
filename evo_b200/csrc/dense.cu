// Dense BSC log-pseudo-joint on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), sm_100a.
//
//   lpj[n][k] = pre1 * sum_{d in obs} ((s_nk W')_d - y_nd)^2 + pil_bar |s_nk|      (evo/models/bsc.py:79-97)
//
// evaluated as ONE logical GEMM  C[(n,k)][d] = sum_h S[(n,k)][h] W[d][h]  (M = N*S states, N = D,
// K = H) whose accumulator never leaves the SM: the epilogue reads it from TMEM, subtracts y_n,
// squares, reduces over d and writes one f64 per state.  The N*K x D residual of the reference
// (`np.dot(states, W.T) - y`, bsc.py:91-93) is never materialised.
//
// Precision: the binary operand S is exact in fp16.  W (f64) is scaled by a power of two and split
// into two fp16 planes hi + lo (22 significant bits); both planes accumulate into the same fp32
// TMEM accumulator, so the logical GEMM is executed as two tensor-core passes.  Measured against the
// f64 oracle: <= 2e-6 relative on lpj (tests/test_dense_gpu.py states 1e-5).
//
// Roles inside one persistent CTA (320 threads, one CTA per SM):
//   warp 0      TMA producer: W planes, [box_rows x 64] fp16 tiles, SWIZZLE_128B, 5-stage ring
//   warp 1      MMA issuer (one lane): tcgen05.mma.kind::f16, M=128, N<=256, K=16; owns TMEM (512 cols)
//   warps 2-5   epilogue: tcgen05.ld 32x32b -> (acc*2^-e - y)^2 -> f64 row sums -> lpj
//   warps 6-9   A generator: packed state bits -> fp16 {0,1} tiles in the UMMA K-major 128B-swizzle layout
// Pipelines: A ring (3 stages, generic-proxy writes + fence.proxy.async), B ring (5 stages, TMA
// complete_tx), TMEM accumulator double buffer (2 x 256 columns) between MMA and epilogue.
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "stats.cuh"

namespace evo {
namespace dense {

constexpr int BM = 128;           // states per tile (UMMA M)
constexpr int BK = 64;            // latents per k-block: 64 fp16 = 128 B = one swizzle row
constexpr int BN = 256;           // max observables per accumulator stage (UMMA N)
constexpr int A_STAGES = 3;
constexpr int B_STAGES = 5;
constexpr int A_BYTES = BM * BK * 2;   // 16 KB
constexpr int B_BYTES = BN * BK * 2;   // 32 KB
constexpr int N_BARS = 2 * A_STAGES + 2 * B_STAGES + 4;
constexpr int THREADS = 320;
constexpr int Y_STAGE_FLOATS = 1024;  // per epilogue warp: y of its datapoint (D <= 1024)
constexpr size_t SMEM_BYTES = 1024 + (size_t)A_STAGES * A_BYTES + (size_t)B_STAGES * B_BYTES + 4 * Y_STAGE_FLOATS * 4 +
                              8 * N_BARS + 16;

struct Args {
  const uint32_t *states;
  int shared_states;
  int64_t rows;  // N * Sp
  int Sp, D, Dp, H, Hw, KB, n_tiles, box_rows, plane_rows;
  const float *Y;
  const uint8_t *mask;
  const float *descale;  // device scalar 2^-e written by the split kernel
  double pre1, pil_bar;
  double *out;
  int64_t ld;
  unsigned long long *counters;
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B: 8-row x 128 B atoms, SBO = 1024 B between
// 8-row groups, LBO unused (1), descriptor version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D = f32, A = B = f16, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t umma_idesc_f16(int n) {
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

// two state bits -> two fp16 values {0.0, 1.0} packed in one b32
__device__ __forceinline__ uint32_t bits2_to_half2(uint32_t b) {
  return ((b & 1u) * 0x3C00u) | ((b & 2u) * 0x1E000000u);
}

__global__ void __launch_bounds__(THREADS, 1) dense_lpj_kernel(const __grid_constant__ CUtensorMap tmap, Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t *sm = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
  uint8_t *sA = sm;
  uint8_t *sB = sm + (size_t)A_STAGES * A_BYTES;
  uint8_t *sY = sB + (size_t)B_STAGES * B_BYTES;
  uint64_t *bars = reinterpret_cast<uint64_t *>(sY + 4 * Y_STAGE_FLOATS * 4);
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + N_BARS);
  const uint32_t bar0 = smem_u32(bars);
  auto A_FULL = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto A_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(A_STAGES + s); };
  auto B_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * A_STAGES + s); };
  auto B_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * A_STAGES + B_STAGES + s); };
  auto T_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * A_STAGES + 2 * B_STAGES + s); };
  auto T_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * A_STAGES + 2 * B_STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < A_STAGES; ++s) { mbar_init(A_FULL(s), 128); mbar_init(A_EMPTY(s), 1); }
    for (int s = 0; s < B_STAGES; ++s) { mbar_init(B_FULL(s), 1); mbar_init(B_EMPTY(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(T_FULL(s), 1); mbar_init(T_EMPTY(s), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int64_t tiles = (a.rows + BM - 1) / BM;

  if (warp == 0) {
    // ===== TMA producer: W planes ================================================================
    if (lane == 0) {
      int bs = 0;
      uint32_t bph = 0;
      const uint32_t tx = (uint32_t)a.box_rows * 128u;
      for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x)
        for (int j = 0; j < a.n_tiles; ++j)
          for (int kb = 0; kb < a.KB; ++kb)
            for (int p = 0; p < 2; ++p) {
              mbar_wait(B_EMPTY(bs), bph ^ 1u);
              mbar_expect_tx(B_FULL(bs), tx);
              tma_load_2d(smem_u32(sB + (size_t)bs * B_BYTES), &tmap, B_FULL(bs), kb * BK,
                          p * a.plane_rows + j * a.box_rows);
              if (++bs == B_STAGES) { bs = 0; bph ^= 1u; }
            }
    }
  } else if (warp == 1) {
    // ===== MMA issuer ============================================================================
    if (lane == 0) {
      int as = 0, bs = 0, ts = 0;
      uint32_t aph = 0, bph = 0, tph = 0;
      for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x)
        for (int j = 0; j < a.n_tiles; ++j) {
          const int nj = min(BN, a.Dp - j * BN);
          const uint32_t idesc = umma_idesc_f16(nj);
          mbar_wait(T_EMPTY(ts), tph ^ 1u);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)ts * BN;
          for (int kb = 0; kb < a.KB; ++kb) {
            mbar_wait(A_FULL(as), aph);
            const uint64_t adesc = umma_desc_sw128(smem_u32(sA + (size_t)as * A_BYTES));
            for (int p = 0; p < 2; ++p) {
              mbar_wait(B_FULL(bs), bph);
              tc_fence_after();
              const uint64_t bdesc = umma_desc_sw128(smem_u32(sB + (size_t)bs * B_BYTES));
#pragma unroll
              for (int k4 = 0; k4 < BK / 16; ++k4)  // +32 B per K=16 step inside the 128 B swizzle row
                umma_f16(tmem_d, adesc + (uint64_t)(2 * k4), bdesc + (uint64_t)(2 * k4), idesc, (uint32_t)(kb | p | k4));
              umma_commit(B_EMPTY(bs));
              if (++bs == B_STAGES) { bs = 0; bph ^= 1u; }
            }
            umma_commit(A_EMPTY(as));
            if (++as == A_STAGES) { as = 0; aph ^= 1u; }
          }
          umma_commit(T_FULL(ts));
          if (++ts == 2) { ts = 0; tph ^= 1u; }
        }
    }
  } else if (warp < 6) {
    // ===== epilogue: TMEM -> squared residual -> lpj =============================================
    // y of the warp's datapoint is staged in shared memory one tile ahead (coalesced loads issued
    // while the current tile is being reduced), so the reduction never waits on HBM.
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const float descale = __ldg(a.descale);
    float *ys = reinterpret_cast<float *>(sY) + (size_t)q * Y_STAGE_FLOATS;
    const bool can_stage = (a.Sp % 32 == 0) && (a.D % 4 == 0) && a.D <= Y_STAGE_FLOATS && !a.mask;
    const int nv = a.D >> 2;  // float4 per row
    float4 pre[Y_STAGE_FLOATS / 128];
    auto fetch_y = [&](int64_t tile) {
      const int64_t row0 = tile * BM + q * 32;
      if (!can_stage || row0 >= a.rows) return;
      const float4 *src = reinterpret_cast<const float4 *>(a.Y + (row0 / a.Sp) * (int64_t)a.D);
#pragma unroll
      for (int i = 0; i < Y_STAGE_FLOATS / 128; ++i)
        if (lane + 32 * i < nv) pre[i] = __ldg(src + lane + 32 * i);
    };
    auto stash_y = [&]() {
      if (!can_stage) return;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < Y_STAGE_FLOATS / 128; ++i)
        if (lane + 32 * i < nv) reinterpret_cast<float4 *>(ys)[lane + 32 * i] = pre[i];
      __syncwarp();
    };
    fetch_y(blockIdx.x);
    int ts = 0;
    uint32_t tph = 0;
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
      const int64_t gr = t * BM + q * 32 + lane;
      const bool valid = gr < a.rows;
      const int64_t n = valid ? gr / a.Sp : 0;
      const int k = valid ? (int)(gr - n * a.Sp) : 0;
      const float *yrow = a.Y + n * (int64_t)a.D;
      const uint8_t *mrow = a.mask ? a.mask + n * (int64_t)a.D : nullptr;
      const bool vec = ((a.D & 3) == 0) && !mrow;
      stash_y();
      fetch_y(t + gridDim.x);
      double acc = 0.0;
      for (int j = 0; j < a.n_tiles; ++j) {
        const int nj = min(BN, a.Dp - j * BN);
        mbar_wait(T_FULL(ts), tph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)ts * BN;
        uint32_t v[2][16];
        tmem_ld16(taddr, v[0]);
        tmem_ld_wait();
#pragma unroll 1
        for (int c = 0; c < nj; c += 32) {
#pragma unroll
          for (int hb = 0; hb < 2; ++hb) {
            const int cc = c + 16 * hb;
            if (cc >= nj) break;
            if (cc + 16 < nj) tmem_ld16(taddr + (uint32_t)(cc + 16), v[hb ^ 1]);  // next chunk in flight
            const int d0 = j * BN + cc;
            float part = 0.f;
            if (vec && d0 + 16 <= a.D) {
#pragma unroll
              for (int i = 0; i < 16; i += 4) {
                const float4 y4 = can_stage ? *reinterpret_cast<const float4 *>(ys + d0 + i)
                                            : __ldg(reinterpret_cast<const float4 *>(yrow + d0 + i));
                float r0 = fmaf(__uint_as_float(v[hb][i + 0]), descale, -y4.x);
                float r1 = fmaf(__uint_as_float(v[hb][i + 1]), descale, -y4.y);
                float r2 = fmaf(__uint_as_float(v[hb][i + 2]), descale, -y4.z);
                float r3 = fmaf(__uint_as_float(v[hb][i + 3]), descale, -y4.w);
                part = fmaf(r0, r0, part);
                part = fmaf(r1, r1, part);
                part = fmaf(r2, r2, part);
                part = fmaf(r3, r3, part);
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int d = d0 + i;
                if (d < a.D && (!mrow || mrow[d])) {
                  float r = fmaf(__uint_as_float(v[hb][i]), descale, -__ldg(yrow + d));
                  part = fmaf(r, r, part);
                }
              }
            }
            acc += (double)part;
            tmem_ld_wait();
          }
        }
        tc_fence_before();
        mbar_arrive(T_EMPTY(ts));
        if (++ts == 2) { ts = 0; tph ^= 1u; }
      }
      if (valid) {
        const uint32_t *st = a.states + (a.shared_states ? (int64_t)k : gr) * a.Hw;
        int pc = 0;
        for (int w = 0; w < a.Hw; ++w) pc += __popc(__ldg(st + w));
        unsigned fl = 0;
        const double v = lpj_sanitise(a.pre1 * acc + a.pil_bar * (double)pc, fl);
        a.out[n * a.ld + k] = v;
        if (fl && a.counters) atomicAdd(a.counters + ((fl & 1u) ? 0 : ((fl & 2u) ? 1 : 2)), 1ull);
      }
    }
  } else {
    // ===== A generator: packed bits -> fp16 tile, K-major, 128 B swizzle =========================
    const int r = (warp - 6) * 32 + lane;
    uint8_t *rowbase = sA + (size_t)(r >> 3) * 1024 + (size_t)(r & 7) * 128;
    const uint32_t sw = (uint32_t)(r & 7);
    int as = 0;
    uint32_t aph = 0;
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
      const int64_t gr = t * BM + r;
      const bool valid = gr < a.rows;
      const uint32_t *st = a.states + (valid ? (a.shared_states ? gr % a.Sp : gr) : 0) * a.Hw;
      for (int j = 0; j < a.n_tiles; ++j)
        for (int kb = 0; kb < a.KB; ++kb) {
          uint32_t w0 = 0, w1 = 0;
          if (valid) {
            if (2 * kb < a.Hw) w0 = __ldg(st + 2 * kb);
            if (2 * kb + 1 < a.Hw) w1 = __ldg(st + 2 * kb + 1);
          }
          mbar_wait(A_EMPTY(as), aph ^ 1u);
          uint8_t *dst = rowbase + (size_t)as * A_BYTES;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const uint32_t byte = ((c < 4 ? w0 : w1) >> (8 * (c & 3))) & 0xFFu;
            uint4 o;
            o.x = bits2_to_half2(byte);
            o.y = bits2_to_half2(byte >> 2);
            o.z = bits2_to_half2(byte >> 4);
            o.w = bits2_to_half2(byte >> 6);
            *reinterpret_cast<uint4 *>(dst + (((uint32_t)c ^ sw) << 4)) = o;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          mbar_arrive(A_FULL(as));
          if (++as == A_STAGES) { as = 0; aph ^= 1u; }
        }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ---- W -> two scaled fp16 planes ----------------------------------------------------------------
__global__ void wmax_kernel(const double *__restrict__ W, int64_t n, unsigned int *out) {
  float m = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = fabsf((float)W[i]);
    if (v == v && v < INFINITY) m = fmaxf(m, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, __float_as_uint(m));  // non-negative floats order like their bits
}

// planes: [2][plane_rows][Hp] fp16, zero padded.  hdr[0] = max|W| bits (input), hdr[1] = 2^-e (f32, output)
__global__ void wsplit_kernel(const double *__restrict__ W, int D, int H, int plane_rows, int Hp, __half *planes,
                              unsigned int *hdr) {
  const float mx = __uint_as_float(hdr[0]);
  int e = 0;
  if (mx > 0.f) e = 13 - ilogbf(mx);  // max |W| 2^e in [2^13, 2^14): fp16 range with headroom
  e = max(-100, min(100, e));
  const double scale = ldexp(1.0, e);
  const int64_t total = (int64_t)plane_rows * Hp;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int d = (int)(i / Hp), h = (int)(i - (int64_t)d * Hp);
    double v = (d < D && h < H) ? W[(int64_t)d * H + h] * scale : 0.0;
    if (!(v == v)) v = 0.0;
    v = fmin(fmax(v, -65504.0), 65504.0);
    const __half hi = __double2half(v);
    const __half lo = __double2half(v - (double)__half2float(hi));
    planes[i] = hi;
    planes[total + i] = lo;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<float *>(hdr)[1] = (float)ldexp(1.0, -e);
}

struct Geometry {
  int Dp, Hp, KB, n_tiles, box_rows, plane_rows;
  size_t plane_bytes, total_bytes;
};
static Geometry geometry(int D, int H) {
  Geometry g;
  g.Dp = (D + 15) & ~15;
  g.Hp = (H + BK - 1) / BK * BK;
  g.KB = g.Hp / BK;
  g.box_rows = g.Dp < BN ? g.Dp : BN;
  g.n_tiles = (g.Dp + BN - 1) / BN;
  g.plane_rows = g.n_tiles * g.box_rows;
  g.plane_bytes = (size_t)g.plane_rows * g.Hp * sizeof(__half);
  g.total_bytes = 256 + 2 * g.plane_bytes;  // 256-byte header (max bits, descale), planes 256-B aligned
  return g;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}


// =================================================================================================
// B = Y W on the tensor cores (table of the Gram form, evo_bsc_prepare's f64 SIMT GEMM at 2 N D H
// flop is the largest dense contraction of a BSC iteration at D = H = 1024).
//   C[n][h] = sum_d Y[n][d] W[d][h]:  M = N datapoints (tiles of 128), N = H (tiles of <= 256), K = D.
// Both operands are scaled by a power of two and split into fp16 hi + lo planes; three tensor passes
// (hi*hi, hi*lo, lo*hi) accumulate in the same fp32 TMEM tile -- the dropped lo*lo term is 2^-22
// relative.  Y planes are built once per dataset, W' planes once per parameter set.  Epilogue:
// TMEM -> f64 -> B[n][h].  Measured against the f64 kernel: <= 1e-6 of max|B| (tests state 2e-6).
// =================================================================================================
constexpr int T_STAGES = 2;
constexpr int T_STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;  // {Y hi, Y lo, W' hi, W' lo} of one k-block: 96 KB
constexpr int T_THREADS = 192;
constexpr size_t T_SMEM_BYTES = 1024 + (size_t)T_STAGES * T_STAGE_BYTES + 8 * (2 * T_STAGES + 4) + 16;

struct TArgs {
  int64_t N;             // valid output rows (A-operand rows)
  int H, KB, n_tiles, box_rows;  // valid output columns, k-blocks in all, column tiles, B-operand rows per tile
  int64_t a_plane_rows;  // rows of one A plane (rounded up to 128)
  int b_plane_rows;      // rows of one B plane (n_tiles * box_rows)
  int Hp;                // H rounded up to 16
  int k_chunks, kb_chunk;  // split-K: work item = (row tile, column tile, chunk of kb_chunk k-blocks)
  int accumulate;        // 0: out = C (k_chunks == 1); 1: out += C by f64 atomics
  const float *descale;  // device scalar 2^-(ea + eb)
  double *B;             // output, row stride H
  int chunk_major;       // split-K item order: 1 = all output tiles of one k-chunk are neighbours, so the CTAs running at the
                         // same time share the chunk's operand slabs in L2 and every operand byte comes from HBM once
};
// Split-K item order of the Wp GEMM, measured at cfg5 (N = 1M): tile-major 8.8 ms; tile-major with a private f64 tile per CTA
// instead of one atomic per element per item 8.9 ms (the atomics were never the limit); chunk-major 7.5 ms (the operand slabs
// of a k-chunk, 32 MB, stay in L2 while all 32 output tiles use them); chunks of 32 / 128 / 256 k-blocks 7.6 / 7.9 / 7.9 ms.
__device__ __forceinline__ void item_decode(const TArgs &a, int64_t it, int64_t tiles, int64_t &tj, int &kc) {
  if (a.chunk_major) { kc = (int)(it / tiles); tj = it - (int64_t)kc * tiles; }
  else { tj = it / a.k_chunks; kc = (int)(it - tj * a.k_chunks); }
}

__global__ void __launch_bounds__(T_THREADS, 1) table_gemm_kernel(const __grid_constant__ CUtensorMap tmapA,
                                                                  const __grid_constant__ CUtensorMap tmapB, TArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t *sm = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
  uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (size_t)T_STAGES * T_STAGE_BYTES);
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * T_STAGES + 4);
  const uint32_t bar0 = smem_u32(bars);
  auto FULL = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(T_STAGES + s); };
  auto T_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * T_STAGES + s); };
  auto T_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * T_STAGES + 2 + s); };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < T_STAGES; ++s) { mbar_init(FULL(s), 1); mbar_init(EMPTY(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(T_FULL(s), 1); mbar_init(T_EMPTY(s), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int64_t tiles_all = ((a.N + BM - 1) / BM) * a.n_tiles;
  const int64_t items = tiles_all * a.k_chunks;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      int st = 0;
      uint32_t ph = 0;
      const uint32_t tx = 2u * A_BYTES + 2u * (uint32_t)a.box_rows * 128u;
      for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
        int64_t tj;
        int kc;
        item_decode(a, it, tiles_all, tj, kc);
        const int64_t t = tj / a.n_tiles;
        const int j = (int)(tj % a.n_tiles), kb0 = kc * a.kb_chunk, kb1 = min(a.KB, kb0 + a.kb_chunk);
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(EMPTY(st), ph ^ 1u);
            mbar_expect_tx(FULL(st), tx);
            const uint32_t base = smem_u32(sm + (size_t)st * T_STAGE_BYTES);
            tma_load_2d(base, &tmapA, FULL(st), kb * BK, (int)(t * BM));
            tma_load_2d(base + A_BYTES, &tmapA, FULL(st), kb * BK, (int)(a.a_plane_rows + t * BM));
            tma_load_2d(base + 2 * A_BYTES, &tmapB, FULL(st), kb * BK, j * a.box_rows);
            tma_load_2d(base + 2 * A_BYTES + B_BYTES, &tmapB, FULL(st), kb * BK, a.b_plane_rows + j * a.box_rows);
            if (++st == T_STAGES) { st = 0; ph ^= 1u; }
          }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer =====
      int st = 0, ts = 0;
      uint32_t ph = 0, tph = 0;
      for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
          int64_t tj;
          int kc;
          item_decode(a, it, tiles_all, tj, kc);
          const int j = (int)(tj % a.n_tiles);
          const int kb0 = kc * a.kb_chunk, kb1 = min(a.KB, kb0 + a.kb_chunk);
          const int nj = min(BN, a.Hp - j * BN);
          const uint32_t idesc = umma_idesc_f16(nj);
          mbar_wait(T_EMPTY(ts), tph ^ 1u);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)ts * BN;
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(FULL(st), ph);
            tc_fence_after();
            const uint32_t base = smem_u32(sm + (size_t)st * T_STAGE_BYTES);
            const uint64_t ahi = umma_desc_sw128(base), alo = umma_desc_sw128(base + A_BYTES);
            const uint64_t bhi = umma_desc_sw128(base + 2 * A_BYTES), blo = umma_desc_sw128(base + 2 * A_BYTES + B_BYTES);
#pragma unroll
            for (int k4 = 0; k4 < BK / 16; ++k4) umma_f16(tmem_d, ahi + 2 * k4, bhi + 2 * k4, idesc, (uint32_t)((kb - kb0) | k4));
#pragma unroll
            for (int k4 = 0; k4 < BK / 16; ++k4) umma_f16(tmem_d, ahi + 2 * k4, blo + 2 * k4, idesc, 1u);
#pragma unroll
            for (int k4 = 0; k4 < BK / 16; ++k4) umma_f16(tmem_d, alo + 2 * k4, bhi + 2 * k4, idesc, 1u);
            umma_commit(EMPTY(st));
            if (++st == T_STAGES) { st = 0; ph ^= 1u; }
          }
          umma_commit(T_FULL(ts));
          if (++ts == 2) { ts = 0; tph ^= 1u; }
        }
    }
  } else {
    // ===== epilogue: TMEM -> f64 -> B[n][h] =====
    const int q = warp & 3;
    const float descale = __ldg(a.descale);
    int ts = 0;
    uint32_t tph = 0;
    const bool vec2 = (a.H & 1) == 0;
    for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
      int64_t tj;
      int kc;
      item_decode(a, it, tiles_all, tj, kc);
      const int64_t t = tj / a.n_tiles;
      const int j = (int)(tj % a.n_tiles);
      const int64_t n = t * BM + q * 32 + lane;
      const bool valid = n < a.N;
      double *orow = a.B + (valid ? n : 0) * (int64_t)a.H;
      {
        const int nj = min(BN, a.Hp - j * BN);
        mbar_wait(T_FULL(ts), tph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)ts * BN;
        for (int c = 0; c < nj; c += 16) {
          uint32_t v[16];
          tmem_ld16(taddr + (uint32_t)c, v);
          tmem_ld_wait();
          const int h0 = j * BN + c;
          if (valid) {
            if (a.accumulate) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float f = __uint_as_float(v[i]);
                if (h0 + i < a.H && f != 0.f) atomicAdd(orow + h0 + i, (double)f * (double)descale);
              }
            } else if (vec2 && h0 + 16 <= a.H) {
#pragma unroll
              for (int i = 0; i < 16; i += 2)
                *reinterpret_cast<double2 *>(orow + h0 + i) =
                    make_double2((double)(__uint_as_float(v[i]) * descale), (double)(__uint_as_float(v[i + 1]) * descale));
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (h0 + i < a.H) orow[h0 + i] = (double)(__uint_as_float(v[i]) * descale);
            }
          }
        }
        tc_fence_before();
        mbar_arrive(T_EMPTY(ts));
        if (++ts == 2) { ts = 0; tph ^= 1u; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// max |x| of an f32 array (non-negative floats order like their bit patterns)
__global__ void fmax_kernel(const float *__restrict__ X, int64_t n, unsigned int *out) {
  float m = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = fabsf(X[i]);
    if (v == v && v < INFINITY) m = fmaxf(m, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, __float_as_uint(m));
}

__device__ __forceinline__ int scale_exp(float mx) {
  int e = 0;
  if (mx > 0.f) e = 13 - ilogbf(mx);  // max |x| 2^e in [2^13, 2^14)
  return max(-100, min(100, e));
}
__device__ __forceinline__ void split_half(double v, __half &hi, __half &lo) {
  if (!(v == v)) v = 0.0;
  v = fmin(fmax(v, -65504.0), 65504.0);
  hi = __double2half(v);
  lo = __double2half(v - (double)__half2float(hi));
}

// Y (f32 [N][D]) -> planes [2][Np][Dp] fp16, zero padded.  hdr[0] = max|Y| bits; hdr[1] <- 2^-ey (f32)
__global__ void ysplit_kernel(const float *__restrict__ Y, int64_t N, int D, int64_t Np, int Dp, __half *planes,
                              unsigned int *hdr) {
  const int e = scale_exp(__uint_as_float(hdr[0]));
  const double scale = ldexp(1.0, e);
  const int64_t total = Np * Dp;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / Dp;
    const int d = (int)(i - n * Dp);
    __half hi, lo;
    split_half((n < N && d < D) ? (double)Y[n * D + d] * scale : 0.0, hi, lo);
    planes[i] = hi;
    planes[total + i] = lo;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<float *>(hdr)[1] = (float)ldexp(1.0, -e);
}

// W (f64 [D][H]) -> transposed planes [2][Hr][Dp] fp16.  hdr[2] = max|W| bits; hdr[3] <- 2^-ew; hdr[4] <- 2^-(ey+ew)
__global__ void wtsplit_kernel(const double *__restrict__ W, int D, int H, int Hr, int Dp, __half *planes, unsigned int *hdr) {
  const int e = scale_exp(__uint_as_float(hdr[2]));
  const double scale = ldexp(1.0, e);
  const int64_t total = (int64_t)Hr * Dp;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int h = (int)(i / Dp), d = (int)(i - (int64_t)h * Dp);
    __half hi, lo;
    split_half((h < H && d < D) ? W[(int64_t)d * H + h] * scale : 0.0, hi, lo);
    planes[i] = hi;
    planes[total + i] = lo;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float *f = reinterpret_cast<float *>(hdr);
    f[3] = (float)ldexp(1.0, -e);
    f[4] = f[1] * f[3];
  }
}

struct TGeometry {
  int Dp, KB, Hp, n_tiles, box_rows, b_plane_rows;
  int64_t Np;
  size_t y_bytes, w_bytes, total_bytes;
};
static TGeometry tgeometry(int64_t N, int D, int H) {
  TGeometry g;
  g.Dp = (D + BK - 1) / BK * BK;
  g.KB = g.Dp / BK;
  g.Hp = (H + 15) & ~15;
  g.box_rows = g.Hp < BN ? g.Hp : BN;
  g.n_tiles = (g.Hp + BN - 1) / BN;
  g.b_plane_rows = g.n_tiles * g.box_rows;
  g.Np = (N + BM - 1) / BM * BM;
  g.y_bytes = ((size_t)2 * g.Np * g.Dp * sizeof(__half) + 255) & ~(size_t)255;
  g.w_bytes = ((size_t)2 * g.b_plane_rows * g.Dp * sizeof(__half) + 255) & ~(size_t)255;
  g.total_bytes = 256 + g.y_bytes + g.w_bytes;
  return g;
}
static int encode_2d(CUtensorMap *tm, void *base, uint64_t cols, uint64_t rows, uint32_t box_rows) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return EVO_ERR_UNSUPPORTED;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * sizeof(__half)};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult cr = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d)", (int)cr);
    return EVO_ERR_CUDA;
  }
  return EVO_OK;
}

// ---- Wp = <S>' Y on the tensor cores (bsc.py:204-206) ----------------------------------------------
// C[h][d] = sum_n E'[h][n] Y'[d][n]: M = H (tiles of 128), N = D (tiles of <= 256), K = the datapoints,
// split into chunks of kb_chunk k-blocks; every work item adds its fp32 tile into the f64 statistics
// by atomics.  E' = 2^14 <s> (written by the statistics pass), Y' = 2^ey y: fp16 hi + lo planes, K-major.
// Y (f32 [N][D]) -> transposed planes [2][Dr][Np].  hdr[0] = max|Y| bits; hdr[1] <- 2^-ey; hdr[4] <- 2^-(ey+14)
__global__ void __launch_bounds__(256) ytsplit_kernel(const float *__restrict__ Y, int64_t N, int D, int64_t Np, int Dr,
                                                      __half *planes, unsigned int *hdr) {
  // 32 (d) x 32 (n) tiles through shared memory: reads coalesced along d, writes along n
  __shared__ float tile[32][33];
  const int e = scale_exp(__uint_as_float(hdr[0]));
  const float scale = (float)ldexp(1.0, e);  // power of two: exact in f32
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t n_tiles = Np / 32, total = (int64_t)Dr * Np;
  const int d_tiles = (Dr + 31) / 32;
  for (int64_t t = blockIdx.x; t < n_tiles * d_tiles; t += gridDim.x) {
    const int64_t n0 = (t / d_tiles) * 32;
    const int d0 = (int)(t % d_tiles) * 32;
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
      const int64_t n = n0 + r;
      const int d = d0 + tx;
      tile[r][tx] = (n < N && d < D) ? Y[n * D + d] * scale : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
      const int d = d0 + r;
      if (d < Dr) {
        __half hi, lo;
        split_half((double)tile[tx][r], hi, lo);
        const int64_t i = (int64_t)d * Np + n0 + tx;
        planes[i] = hi;
        planes[total + i] = lo;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float *f = reinterpret_cast<float *>(hdr);
    f[1] = (float)ldexp(1.0, -e);
    f[4] = (float)ldexp(1.0, -e - 14);
  }
}

struct WGeometry {
  int64_t Np;  // datapoints rounded up to a k-block
  int KB, Hr, Dp, n_tiles, box_rows, Dr;
  size_t yt_bytes, et_bytes, total_bytes;
};
static WGeometry wgeometry(int64_t N, int D, int H) {
  WGeometry g;
  g.Np = (N + BK - 1) / BK * BK;
  g.KB = (int)(g.Np / BK);
  g.Hr = (H + BM - 1) / BM * BM;
  g.Dp = (D + 15) & ~15;
  g.box_rows = g.Dp < BN ? g.Dp : BN;
  g.n_tiles = (g.Dp + BN - 1) / BN;
  g.Dr = g.n_tiles * g.box_rows;
  g.yt_bytes = ((size_t)2 * g.Dr * g.Np * sizeof(__half) + 255) & ~(size_t)255;
  g.et_bytes = ((size_t)2 * g.Hr * g.Np * sizeof(__half) + 255) & ~(size_t)255;
  g.total_bytes = 256 + g.yt_bytes + g.et_bytes;
  return g;
}

}  // namespace dense
}  // namespace evo

using namespace evo;

extern "C" int64_t evo_bsc_lpj_gemm_workspace_bytes(int D, int H) {
  if (D <= 0 || H <= 0) return 0;
  return (int64_t)dense::geometry(D, H).total_bytes;
}

extern "C" int evo_bsc_lpj_gemm_prepare(const double *W, int D, int H, void *workspace, void *stream) {
  EVO_REQUIRE(W && workspace && D > 0 && H > 0, "bad arguments");
  EVO_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  dense::Geometry g = dense::geometry(D, H);
  unsigned int *hdr = reinterpret_cast<unsigned int *>(workspace);
  __half *planes = reinterpret_cast<__half *>(reinterpret_cast<uint8_t *>(workspace) + 256);
  EVO_CUDA(cudaMemsetAsync(hdr, 0, 256, st));
  int64_t n = (int64_t)D * H;
  int blocks = (int)((n + 255) / 256 < (int64_t)sm_count() * 4 ? (n + 255) / 256 : (int64_t)sm_count() * 4);
  dense::wmax_kernel<<<blocks, 256, 0, st>>>(W, n, hdr);
  EVO_LAUNCH_CHECK("wmax");
  int64_t total = (int64_t)g.plane_rows * g.Hp;
  blocks = (int)((total + 255) / 256 < (int64_t)sm_count() * 4 ? (total + 255) / 256 : (int64_t)sm_count() * 4);
  dense::wsplit_kernel<<<blocks, 256, 0, st>>>(W, D, H, g.plane_rows, g.Hp, planes, hdr);
  EVO_LAUNCH_CHECK("wsplit");
  return EVO_OK;
}

extern "C" int evo_bsc_lpj_gemm(const uint32_t *states, int shared_states, int64_t N, int Sp, int D, int H,
                                const void *workspace, const float *Y, const uint8_t *mask, double pre1, double pil_bar,
                                double *lpj_out, int64_t ld_lpj, unsigned long long *reset_counters, void *stream) {
  EVO_REQUIRE(N >= 0 && Sp > 0 && D > 0 && H > 0 && ld_lpj >= Sp, "bad sizes");
  EVO_REQUIRE(states && workspace && Y && lpj_out, "null pointer");
  if (N == 0) return EVO_OK;
  dense::Geometry g = dense::geometry(D, H);
  dense::EncodeTiledFn enc = dense::encode_fn();
  if (!enc) {
    set_error("evo_bsc_lpj_gemm: cuTensorMapEncodeTiled is not available from the driver");
    return EVO_ERR_UNSUPPORTED;
  }
  uint8_t *planes = const_cast<uint8_t *>(reinterpret_cast<const uint8_t *>(workspace)) + 256;
  CUtensorMap tmap;
  cuuint64_t gdim[2] = {(cuuint64_t)g.Hp, (cuuint64_t)(2 * g.plane_rows)};
  cuuint64_t gstride[1] = {(cuuint64_t)g.Hp * sizeof(__half)};
  cuuint32_t box[2] = {(cuuint32_t)dense::BK, (cuuint32_t)g.box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, planes, gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) {
    set_error("evo_bsc_lpj_gemm: cuTensorMapEncodeTiled failed (%d)", (int)cr);
    return EVO_ERR_CUDA;
  }
  dense::Args a;
  a.states = states;
  a.shared_states = shared_states;
  a.rows = N * Sp;
  a.Sp = Sp; a.D = D; a.Dp = g.Dp; a.H = H; a.Hw = words_for(H); a.KB = g.KB;
  a.n_tiles = g.n_tiles; a.box_rows = g.box_rows; a.plane_rows = g.plane_rows;
  a.Y = Y; a.mask = mask;
  a.descale = reinterpret_cast<const float *>(workspace) + 1;
  a.pre1 = pre1; a.pil_bar = pil_bar;
  a.out = lpj_out; a.ld = ld_lpj; a.counters = reset_counters;
  // per device and cheap: set on every launch (a process may drive several GPUs)
  EVO_CUDA(cudaFuncSetAttribute(dense::dense_lpj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)dense::SMEM_BYTES));
  int64_t tiles = (a.rows + dense::BM - 1) / dense::BM;
  int grid = (int)(tiles < sm_count() ? tiles : sm_count());
  dense::dense_lpj_kernel<<<grid, dense::THREADS, dense::SMEM_BYTES, (cudaStream_t)stream>>>(tmap, a);
  EVO_LAUNCH_CHECK("dense_lpj");
  return EVO_OK;
}

static int grid_1d(int64_t n) {
  int64_t b = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

extern "C" int64_t evo_bsc_tables_tc_workspace_bytes(int64_t N, int D, int H) {
  if (N <= 0 || D <= 0 || H <= 0) return 0;
  return (int64_t)dense::tgeometry(N, D, H).total_bytes;
}

extern "C" int evo_bsc_tables_tc_set_data(const float *Y, int64_t N, int D, int H, void *workspace, void *stream) {
  EVO_REQUIRE(Y && workspace && N > 0 && D > 0 && H > 0, "bad arguments");
  EVO_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  dense::TGeometry g = dense::tgeometry(N, D, H);
  unsigned int *hdr = reinterpret_cast<unsigned int *>(workspace);
  __half *yp = reinterpret_cast<__half *>(reinterpret_cast<uint8_t *>(workspace) + 256);
  EVO_CUDA(cudaMemsetAsync(hdr, 0, 8, st));
  dense::fmax_kernel<<<grid_1d(N * D), 256, 0, st>>>(Y, N * (int64_t)D, hdr);
  EVO_LAUNCH_CHECK("fmax");
  dense::ysplit_kernel<<<grid_1d(g.Np * g.Dp), 256, 0, st>>>(Y, N, D, g.Np, g.Dp, yp, hdr);
  EVO_LAUNCH_CHECK("ysplit");
  return EVO_OK;
}

extern "C" int evo_bsc_tables_tc(const double *W, int64_t N, int D, int H, void *workspace, double *B, void *stream) {
  EVO_REQUIRE(W && workspace && B && N > 0 && D > 0 && H > 0, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  dense::TGeometry g = dense::tgeometry(N, D, H);
  uint8_t *ws = reinterpret_cast<uint8_t *>(workspace);
  unsigned int *hdr = reinterpret_cast<unsigned int *>(ws);
  __half *yp = reinterpret_cast<__half *>(ws + 256);
  __half *wp = reinterpret_cast<__half *>(ws + 256 + g.y_bytes);
  EVO_CUDA(cudaMemsetAsync(hdr + 2, 0, 4, st));
  dense::wmax_kernel<<<grid_1d((int64_t)D * H), 256, 0, st>>>(W, (int64_t)D * H, hdr + 2);
  EVO_LAUNCH_CHECK("wmax");
  dense::wtsplit_kernel<<<grid_1d((int64_t)g.b_plane_rows * g.Dp), 256, 0, st>>>(W, D, H, g.b_plane_rows, g.Dp, wp, hdr);
  EVO_LAUNCH_CHECK("wtsplit");
  CUtensorMap ta, tb;
  int rc = dense::encode_2d(&ta, yp, (uint64_t)g.Dp, (uint64_t)(2 * g.Np), (uint32_t)dense::BM);
  if (rc) return rc;
  rc = dense::encode_2d(&tb, wp, (uint64_t)g.Dp, (uint64_t)(2 * g.b_plane_rows), (uint32_t)g.box_rows);
  if (rc) return rc;
  dense::TArgs a;
  a.N = N; a.H = H; a.KB = g.KB; a.n_tiles = g.n_tiles; a.box_rows = g.box_rows;
  a.a_plane_rows = g.Np; a.b_plane_rows = g.b_plane_rows; a.Hp = g.Hp;
  a.k_chunks = 1; a.kb_chunk = g.KB; a.accumulate = 0;
  a.descale = reinterpret_cast<const float *>(ws) + 4;
  a.B = B;
  a.chunk_major = 0;
  EVO_CUDA(cudaFuncSetAttribute(dense::table_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)dense::T_SMEM_BYTES));
  int64_t tiles = (N + dense::BM - 1) / dense::BM * g.n_tiles;
  int grid = (int)(tiles < sm_count() ? tiles : sm_count());
  dense::table_gemm_kernel<<<grid, dense::T_THREADS, dense::T_SMEM_BYTES, st>>>(ta, tb, a);
  EVO_LAUNCH_CHECK("table_gemm");
  return EVO_OK;
}

extern "C" int64_t evo_bsc_wp_tc_workspace_bytes(int64_t N, int D, int H) {
  if (N <= 0 || D <= 0 || H <= 0) return 0;
  return (int64_t)dense::wgeometry(N, D, H).total_bytes;
}

extern "C" int evo_bsc_wp_tc_set_data(const float *Y, int64_t N, int D, int H, void *workspace, void *stream) {
  EVO_REQUIRE(Y && workspace && N > 0 && D > 0 && H > 0, "bad arguments");
  EVO_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  dense::WGeometry g = dense::wgeometry(N, D, H);
  unsigned int *hdr = reinterpret_cast<unsigned int *>(workspace);
  __half *yt = reinterpret_cast<__half *>(reinterpret_cast<uint8_t *>(workspace) + 256);
  EVO_CUDA(cudaMemsetAsync(hdr, 0, 8, st));
  dense::fmax_kernel<<<grid_1d(N * D), 256, 0, st>>>(Y, N * (int64_t)D, hdr);
  EVO_LAUNCH_CHECK("fmax");
  dense::ytsplit_kernel<<<grid_1d((int64_t)g.Dr * g.Np / 4), 256, 0, st>>>(Y, N, D, g.Np, g.Dr, yt, hdr);
  EVO_LAUNCH_CHECK("ytsplit");
  return EVO_OK;
}

int evo::wp_tc_planes(void *workspace, int64_t N, int D, int H, WpLists *out) {
  dense::WGeometry g = dense::wgeometry(N, D, H);
  uint8_t *ws = reinterpret_cast<uint8_t *>(workspace);
  __half *et = reinterpret_cast<__half *>(ws + 256 + g.yt_bytes);
  out->et_hi = et;
  out->et_lo = et + (size_t)g.Hr * g.Np;
  out->et_ld = g.Np;
  return EVO_OK;
}

extern "C" int evo_bsc_wp_tc_begin(int64_t N, int D, int H, void *workspace, void *stream) {
  EVO_REQUIRE(workspace && N > 0 && D > 0 && H > 0, "bad arguments");
  EVO_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  dense::WGeometry g = dense::wgeometry(N, D, H);
  uint8_t *ws = reinterpret_cast<uint8_t *>(workspace);
  EVO_CUDA(cudaMemsetAsync(ws + 256 + g.yt_bytes, 0, (size_t)2 * g.Hr * g.Np * sizeof(__half), (cudaStream_t)stream));
  return EVO_OK;
}

static int wp_tc_finish(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H, int S,
                        int S_perm, double pre1, double pil_bar, double *stats, float *Es_out, void *workspace,
                        const int32_t *rest, bool all, const double *theta_dev, cudaStream_t st);

extern "C" int evo_bsc_suffstats_tc(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H,
                                    int S, int S_perm, double pre1, double pil_bar, double *stats, float *Es_out,
                                    void *workspace, void *stream) {
  EVO_REQUIRE(states && lpj && stats && workspace && N > 0 && D > 0 && H > 0 && S > 0, "bad arguments");
  int rc = evo_bsc_wp_tc_begin(N, D, H, workspace, stream);
  if (rc) return rc;
  return wp_tc_finish(states, lpj, Ym, N, D, H, S, S_perm, pre1, pil_bar, stats, Es_out, workspace, nullptr, true, nullptr,
                      (cudaStream_t)stream);
}

extern "C" int evo_bsc_suffstats_tc_finish(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D,
                                           int H, int S, int S_perm, double pre1, double pil_bar, double *stats,
                                           void *workspace, const int32_t *rest, const double *theta_dev, void *stream) {
  EVO_REQUIRE(states && lpj && stats && workspace && N > 0 && D > 0 && H > 0 && S > 0, "bad arguments");
  EVO_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  return wp_tc_finish(states, lpj, Ym, N, D, H, S, S_perm, pre1, pil_bar, stats, nullptr, workspace, rest, rest == nullptr,
                      theta_dev, (cudaStream_t)stream);
}

static int wp_tc_finish(const uint32_t *states, const double *lpj, const float *Ym, int64_t N, int D, int H, int S,
                        int S_perm, double pre1, double pil_bar, double *stats, float *Es_out, void *workspace,
                        const int32_t *rest, bool all, const double *theta_dev, cudaStream_t st) {
  dense::WGeometry g = dense::wgeometry(N, D, H);
  uint8_t *ws = reinterpret_cast<uint8_t *>(workspace);
  __half *yt = reinterpret_cast<__half *>(ws + 256);
  __half *et = reinterpret_cast<__half *>(ws + 256 + g.yt_bytes);
  WpLists lists{};
  lists.et_hi = et;
  lists.et_lo = et + (size_t)g.Hr * g.Np;
  lists.et_ld = g.Np;
  int rc = bsc_suffstats_pass1(states, lpj, Ym, N, D, H, S, S_perm, pre1, pil_bar, stats, Es_out, lists, st, all ? nullptr : rest, theta_dev);
  if (rc) return rc;
  CUtensorMap ta, tb;
  rc = dense::encode_2d(&ta, et, (uint64_t)g.Np, (uint64_t)(2 * g.Hr), (uint32_t)dense::BM);
  if (rc) return rc;
  rc = dense::encode_2d(&tb, yt, (uint64_t)g.Np, (uint64_t)(2 * g.Dr), (uint32_t)g.box_rows);
  if (rc) return rc;
  const int base_items = (g.Hr / dense::BM) * g.n_tiles;
  int per = sm_count() / base_items;  // k-chunks that fill the machine once
  if (per < 1) per = 1;
  int kb_chunk = (g.KB + per - 1) / per;
  if (kb_chunk < 8) kb_chunk = 8;     // >= 512 datapoints per item
  if (kb_chunk > 64) kb_chunk = 64;   // <= 4096 datapoints in one fp32 accumulation
  dense::TArgs a;
  a.N = H; a.H = D; a.KB = g.KB; a.n_tiles = g.n_tiles; a.box_rows = g.box_rows;
  a.a_plane_rows = g.Hr; a.b_plane_rows = g.Dr; a.Hp = g.Dp;
  a.kb_chunk = kb_chunk; a.k_chunks = (g.KB + kb_chunk - 1) / kb_chunk; a.accumulate = 1;
  a.descale = reinterpret_cast<const float *>(ws) + 4;
  a.B = stats;  // Wp [H][D] leads the statistics block
  a.chunk_major = 1;
  EVO_CUDA(cudaFuncSetAttribute(dense::table_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)dense::T_SMEM_BYTES));
  const int64_t items = (int64_t)base_items * a.k_chunks;
  int grid = (int)(items < sm_count() ? items : sm_count());
  dense::table_gemm_kernel<<<grid, dense::T_THREADS, dense::T_SMEM_BYTES, st>>>(ta, tb, a);
  EVO_LAUNCH_CHECK("wp_gemm");
  return EVO_OK;
}
