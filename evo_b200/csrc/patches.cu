// Overlapping-patch pipeline either side of the ES3C / BSC path in the denoising and inpainting
// drivers (reference call sites examples/image-denoising/main.py:85-88,169-177,
// examples/image-inpainting/main.py:86-88,174-182).  The implementation lives in tvutil.prepost
// (un-vendored dependency), so these kernels follow the call sites' documented semantics:
//
//   extract  image (C, Hi, Wi) -> patches [N][D], D = C*ph*pw (channel slowest, then patch row, then patch
//            column), N = ny*nx top-left positions (0, shift, 2 shift, ... while the patch fits), row-major;
//   merge    every pixel collects the estimates of all patches covering it (position order) and reduces
//            them by mean (f64 accumulation) or median (mean of the two middle values for an even count);
//            a pixel no patch covers becomes NaN;
//   psnr     sum of squared differences (f64) -- the host forms 10 log10(R^2 / mse) (utils.py:8-18).
//
// All three are HBM-bound byte movers: extract writes N*D*4 bytes once (coalesced along the patch row),
// merge reads them once (each pixel's estimates lie D+-1 floats apart in consecutive patches: the reads
// of neighbouring pixels of a warp fall into the same rows).
#include "common.cuh"

namespace evo {

struct PatchGeom {
  int C, Hi, Wi, ph, pw, shift, ny, nx, D;
};

// one warp per patch: the D image offsets of a patch's elements are tabulated once per CTA in shared
// memory, so the inner loop is one table read, one image read (L2-resident) and one coalesced store
__global__ void __launch_bounds__(256) patches_extract_kernel(const float *__restrict__ img, PatchGeom g,
                                                              float *__restrict__ out) {
  extern __shared__ int off[];  // [D]
  for (int d = threadIdx.x; d < g.D; d += blockDim.x) {
    const int c = d / (g.ph * g.pw), r = (d / g.pw) % g.ph, q = d % g.pw;
    off[d] = (c * g.Hi + r) * g.Wi + q;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t N = (int64_t)g.ny * g.nx, warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; p < N; p += warps) {
    const int ky = (int)(p / g.nx), kx = (int)(p - (int64_t)ky * g.nx);
    const float *src = img + (int64_t)ky * g.shift * g.Wi + kx * g.shift;
    float *dst = out + p * g.D;
    for (int d = lane; d < g.D; d += 32) dst[d] = __ldg(src + off[d]);
  }
}

// range of patch indices k (position = k*shift) along one axis that cover coordinate x: k*shift <= x < k*shift + len
__device__ __forceinline__ void cover_range(int x, int len, int shift, int n, int &k0, int &k1) {
  k0 = x - len + 1 <= 0 ? 0 : (x - len + shift) / shift;  // ceil((x - len + 1) / shift)
  k1 = x / shift;
  if (k1 > n - 1) k1 = n - 1;
}

// one thread per pixel
__global__ void __launch_bounds__(256) patches_merge_mean_kernel(const float *__restrict__ pat, PatchGeom g,
                                                                 float *__restrict__ img) {
  const int64_t total = (int64_t)g.C * g.Hi * g.Wi;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % g.Wi), y = (int)((i / g.Wi) % g.Hi), c = (int)(i / ((int64_t)g.Wi * g.Hi));
    int ky0, ky1, kx0, kx1;
    cover_range(y, g.ph, g.shift, g.ny, ky0, ky1);
    cover_range(x, g.pw, g.shift, g.nx, kx0, kx1);
    double s = 0.0;
    int cnt = 0;
    for (int ky = ky0; ky <= ky1; ++ky)
      for (int kx = kx0; kx <= kx1; ++kx) {
        const int r = y - ky * g.shift, q = x - kx * g.shift;
        s += (double)__ldg(pat + ((int64_t)ky * g.nx + kx) * g.D + ((int64_t)c * g.ph + r) * g.pw + q);
        ++cnt;
      }
    img[i] = cnt ? (float)(s / (double)cnt) : __int_as_float(0x7fc00000);
  }
}

// one warp per pixel: the estimates are spread over the lanes (<= PM_MAX per lane), the median is found by
// radix selection on order-preserving keys (one warp reduction per bit) -- no sorting, no local arrays
constexpr int PM_MAX = 8;  // estimates per lane: up to 256 covering patches per pixel (16x16 patches at shift 1)
__global__ void __launch_bounds__(256) patches_merge_median_kernel(const float *__restrict__ pat, PatchGeom g,
                                                                   float *__restrict__ img) {
  const int lane = threadIdx.x & 31;
  const int64_t total = (int64_t)g.C * g.Hi * g.Wi;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < total; i += warps) {
    const int x = (int)(i % g.Wi), y = (int)((i / g.Wi) % g.Hi), c = (int)(i / ((int64_t)g.Wi * g.Hi));
    int ky0, ky1, kx0, kx1;
    cover_range(y, g.ph, g.shift, g.ny, ky0, ky1);
    cover_range(x, g.pw, g.shift, g.nx, kx0, kx1);
    const int wx = kx1 - kx0 + 1, wy = ky1 - ky0 + 1;
    const int cnt = (wx > 0 && wy > 0) ? wx * wy : 0;
    if (cnt == 0) {
      if (lane == 0) img[i] = __int_as_float(0x7fc00000);
      continue;
    }
    float v[PM_MAX];
    const int per = (cnt + 31) / 32;  // <= PM_MAX (checked on the host)
#pragma unroll
    for (int t = 0; t < PM_MAX; ++t) {
      v[t] = 0.f;
      const int e = t * 32 + lane;
      if (t < per && e < cnt) {
        const int ky = ky0 + e / wx, kx = kx0 + e % wx;
        const int r = y - ky * g.shift, q = x - kx * g.shift;
        v[t] = __ldg(pat + ((int64_t)ky * g.nx + kx) * g.D + ((int64_t)c * g.ph + r) * g.pw + q);
      }
    }
    // order-preserving keys; lanes past the count hold the largest key and never rank
    uint32_t key[PM_MAX];
#pragma unroll
    for (int t = 0; t < PM_MAX; ++t) {
      const uint32_t b = __float_as_uint(v[t]);
      key[t] = (t * 32 + lane < cnt) ? ((b & 0x80000000u) ? ~b : (b | 0x80000000u)) : 0xFFFFFFFFu;
    }
    // lower median = element of rank (cnt-1)/2 by MSB-first radix selection (32 passes, one warp reduction each)
    uint32_t prefix = 0u;
    int k = (cnt - 1) >> 1;
    for (int bit = 31; bit >= 0; --bit) {
      int c0 = 0;  // keys that match the prefix above `bit` and have this bit clear
#pragma unroll
      for (int t = 0; t < PM_MAX; ++t)
        if (t < per) c0 += (((key[t] ^ prefix) >> bit) == 0u) && (t * 32 + lane < cnt);
      c0 = __reduce_add_sync(0xffffffffu, c0);
      if (k >= c0) { k -= c0; prefix |= 1u << bit; }
    }
    // upper median (even count): the same value if enough copies of it reach rank cnt/2, else the next larger key
    int le = 0;
    uint32_t nxt = 0xFFFFFFFFu;
#pragma unroll
    for (int t = 0; t < PM_MAX; ++t)
      if (t < per && t * 32 + lane < cnt) {
        le += key[t] <= prefix;
        if (key[t] > prefix) nxt = min(nxt, key[t]);
      }
    le = __reduce_add_sync(0xffffffffu, le);
    nxt = __reduce_min_sync(0xffffffffu, nxt);
    const uint32_t hi_key = ((cnt >> 1) < le) ? prefix : nxt;
    auto unkey = [](uint32_t x) { return __uint_as_float((x & 0x80000000u) ? (x & 0x7FFFFFFFu) : ~x); };
    const float m_lo = unkey(prefix), m_hi = unkey(hi_key);
    if (lane == 0) img[i] = (m_lo + m_hi) * 0.5f;
  }
}

__global__ void __launch_bounds__(256) sqdiff_kernel(const float *__restrict__ a, const float *__restrict__ b, int64_t n,
                                                     double *out) {
  double s = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double d = (double)a[i] - (double)b[i];
    s += d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0 && s != 0.0) atomicAdd(out, s);
}

static int patch_geom(int C, int Hi, int Wi, int ph, int pw, int shift, PatchGeom &g) {
  if (C <= 0 || Hi <= 0 || Wi <= 0 || ph <= 0 || pw <= 0 || shift <= 0 || ph > Hi || pw > Wi) {
    set_error("patches: bad geometry C=%d image %dx%d patch %dx%d shift %d", C, Hi, Wi, ph, pw, shift);
    return EVO_ERR_ARG;
  }
  g.C = C; g.Hi = Hi; g.Wi = Wi; g.ph = ph; g.pw = pw; g.shift = shift;
  g.ny = (Hi - ph) / shift + 1;
  g.nx = (Wi - pw) / shift + 1;
  g.D = C * ph * pw;
  return EVO_OK;
}
static int grid_for_items(int64_t n) {
  int64_t b = (n + 255) / 256, cap = (int64_t)sm_count() * 8;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

}  // namespace evo

using namespace evo;

extern "C" int64_t evo_patches_count(int Hi, int Wi, int ph, int pw, int shift) {
  if (Hi <= 0 || Wi <= 0 || ph <= 0 || pw <= 0 || shift <= 0 || ph > Hi || pw > Wi) return 0;
  return (int64_t)((Hi - ph) / shift + 1) * ((Wi - pw) / shift + 1);
}

extern "C" int evo_patches_extract(const float *image, int C, int Hi, int Wi, int ph, int pw, int shift, float *patches,
                                   void *stream) {
  EVO_REQUIRE(image && patches, "null pointer");
  PatchGeom g;
  int rc = patch_geom(C, Hi, Wi, ph, pw, shift, g);
  if (rc) return rc;
  const int64_t total = (int64_t)g.ny * g.nx * g.D;
  if ((size_t)g.D * sizeof(int) > 48 * 1024) {
    set_error("patches: D = %d elements per patch not supported (offset table > 48 KB)", g.D);
    return EVO_ERR_UNSUPPORTED;
  }
  patches_extract_kernel<<<grid_for_items(total / 4 + 1), 256, (size_t)g.D * sizeof(int), (cudaStream_t)stream>>>(image, g, patches);
  EVO_LAUNCH_CHECK("patches_extract");
  return EVO_OK;
}

extern "C" int evo_patches_merge(const float *patches, int C, int Hi, int Wi, int ph, int pw, int shift, int method,
                                 float *image, void *stream) {
  EVO_REQUIRE(image && patches, "null pointer");
  EVO_REQUIRE(method == 0 || method == 1, "method: 0 = mean, 1 = median");
  PatchGeom g;
  int rc = patch_geom(C, Hi, Wi, ph, pw, shift, g);
  if (rc) return rc;
  const int64_t pixels = (int64_t)C * Hi * Wi;
  if (method == 0) {
    patches_merge_mean_kernel<<<grid_for_items(pixels), 256, 0, (cudaStream_t)stream>>>(patches, g, image);
    EVO_LAUNCH_CHECK("patches_merge_mean");
  } else {
    const int cover = ((ph + shift - 1) / shift) * ((pw + shift - 1) / shift);
    if (cover > 32 * PM_MAX) {
      set_error("patches: median over %d estimates per pixel (> %d) not supported", cover, 32 * PM_MAX);
      return EVO_ERR_UNSUPPORTED;
    }
    patches_merge_median_kernel<<<grid_for_items(pixels * 32), 256, 0, (cudaStream_t)stream>>>(patches, g, image);
    EVO_LAUNCH_CHECK("patches_merge_median");
  }
  return EVO_OK;
}

extern "C" int evo_sum_sqdiff(const float *a, const float *b, int64_t n, double *out, void *stream) {
  EVO_REQUIRE(a && b && out && n >= 0, "bad arguments");
  if (n == 0) return EVO_OK;
  sqdiff_kernel<<<grid_for_items(n), 256, 0, (cudaStream_t)stream>>>(a, b, n, out);
  EVO_LAUNCH_CHECK("sum_sqdiff");
  return EVO_OK;
}
