// GroupNorm(+SiLU) and LayerNorm over NHWC rows: HBM-bound, 16-byte vectorised, coalesced along channels.
// See include/mdx.h (mdx_groupnorm, mdx_layernorm) for the reference call sites.
//
// GroupNorm = two passes over the activation (the minimum without a persistent cooperative kernel):
//   gn_stats_kernel : each thread owns one 16-byte channel vector and walks pixel rows; per-channel fp32
//                     partial sums -> per-group smem bins -> fp64 atomics into stats[n][group][{sum,sumsq}]
//   gn_apply_kernel : y = x * a_c + b_c (a,b folded from mean/rstd/gamma/beta), optional SiLU, store.
// Both take two sources so the UNet skip concat (openaimodel.py:566-571) is normalised without a copy.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

// 8 consecutive channels per thread, whatever the storage type (fp32: two 16-byte accesses, bf16: one)
template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 8;
  __device__ static void load(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  __device__ static void store(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct Vec<bf16> {
  static constexpr int N = 8;
  __device__ static void load(const bf16* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    float2 f;
    f = unpack_bf16x2(u.x); v[0] = f.x; v[1] = f.y;
    f = unpack_bf16x2(u.y); v[2] = f.x; v[3] = f.y;
    f = unpack_bf16x2(u.z); v[4] = f.x; v[5] = f.y;
    f = unpack_bf16x2(u.w); v[6] = f.x; v[7] = f.y;
  }
  __device__ static void store(bf16* p, const float (&v)[8]) {
    uint4 u;
    u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
    u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};

struct GnP {
  const void* x1; long long ld1; int C1;
  const void* x2; long long ld2; int C2;
  int n, HW, groups, cpg;
  int rows_per_block;
  float eps;
  const float* gamma; const float* beta;
  int silu;
  void* out; long long ldo;
  void* raw; long long ldraw;  // optional un-normalised copy of concat(x1, x2) in the out dtype
  double* stats;
};

// blockDim = (CVr, RP): x = channel-vector index, y = row lane
template <typename T>
__global__ void gn_stats_kernel(const GnP p) {
  constexpr int V = Vec<T>::N;
  __shared__ double bins[64][2];  // fp64 so that the (unordered) atomics cannot change the rounded result
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  for (int i = tid; i < p.groups * 2; i += blockDim.x * blockDim.y) (&bins[0][0])[i] = 0.0;
  __syncthreads();
  const int C = p.C1 + p.C2;
  const int c0 = threadIdx.x * V;
  const int img = blockIdx.y;
  if (c0 < C) {
    const T* src;
    long long ld;
    int cc;
    if (c0 < p.C1) { src = reinterpret_cast<const T*>(p.x1); ld = p.ld1; cc = c0; }
    else { src = reinterpret_cast<const T*>(p.x2); ld = p.ld2; cc = c0 - p.C1; }
    const int r0 = blockIdx.x * p.rows_per_block;
    const int r1 = min(p.HW, r0 + p.rows_per_block);
    float s[V], ss[V];
#pragma unroll
    for (int i = 0; i < V; ++i) { s[i] = 0.f; ss[i] = 0.f; }
    const T* base = src + (long long)img * p.HW * ld + cc;
    for (int r = r0 + threadIdx.y; r < r1; r += blockDim.y) {
      float v[V];
      Vec<T>::load(base + (long long)r * ld, v);
#pragma unroll
      for (int i = 0; i < V; ++i) { s[i] += v[i]; ss[i] = fmaf(v[i], v[i], ss[i]); }
    }
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const int g = (c0 + i) / p.cpg;
      atomicAdd(&bins[g][0], (double)s[i]);
      atomicAdd(&bins[g][1], (double)ss[i]);
    }
  }
  __syncthreads();
  for (int i = tid; i < p.groups * 2; i += blockDim.x * blockDim.y)
    atomicAdd(&p.stats[(long long)img * p.groups * 2 + i], (&bins[0][0])[i]);
}

template <typename T, typename TO>
__global__ void gn_apply_kernel(const GnP p) {
  constexpr int V = Vec<T>::N;
  const int C = p.C1 + p.C2;
  const int c0 = threadIdx.x * V;
  const int img = blockIdx.y;
  if (c0 >= C) return;
  const T* src;
  long long ld;
  int cc;
  if (c0 < p.C1) { src = reinterpret_cast<const T*>(p.x1); ld = p.ld1; cc = c0; }
  else { src = reinterpret_cast<const T*>(p.x2); ld = p.ld2; cc = c0 - p.C1; }
  float a[V], b[V];
  const double cnt = (double)p.HW * p.cpg;
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = c0 + i;
    const int g = c / p.cpg;
    const double sum = p.stats[((long long)img * p.groups + g) * 2];
    const double sq = p.stats[((long long)img * p.groups + g) * 2 + 1];
    const double mean = sum / cnt;
    double var = sq / cnt - mean * mean;
    if (var < 0.0) var = 0.0;
    const float rstd = (float)(1.0 / sqrt(var + (double)p.eps));
    a[i] = rstd * p.gamma[c];
    b[i] = p.beta[c] - (float)mean * a[i];
  }
  const int r0 = blockIdx.x * p.rows_per_block;
  const int r1 = min(p.HW, r0 + p.rows_per_block);
  const T* base = src + (long long)img * p.HW * ld + cc;
  TO* obase = reinterpret_cast<TO*>(p.out) + (long long)img * p.HW * p.ldo + c0;
  TO* rbase = p.raw ? reinterpret_cast<TO*>(p.raw) + (long long)img * p.HW * p.ldraw + c0 : nullptr;
  for (int r = r0 + threadIdx.y; r < r1; r += blockDim.y) {
    float v[V];
    Vec<T>::load(base + (long long)r * ld, v);
    if (rbase) Vec<TO>::store(rbase + (long long)r * p.ldraw, v);
#pragma unroll
    for (int i = 0; i < V; ++i) {
      float y = fmaf(v[i], a[i], b[i]);
      v[i] = p.silu ? silu_f(y) : y;
    }
    Vec<TO>::store(obase + (long long)r * p.ldo, v);
  }
}

// ---------------- LayerNorm: one warp per row, values held in registers ----------------
template <typename T, typename TO, int MAXV>
__global__ void __launch_bounds__(256) layernorm_kernel(const T* __restrict__ x, long long ldx, int rows, int C,
                                                        float eps, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, TO* __restrict__ out,
                                                        long long ldo) {
  constexpr int V = Vec<T>::N;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const int nvec = C / V;
  const T* xr = x + (long long)warp * ldx;
  float v[MAXV][V];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int vi = lane + j * 32;
    if (vi < nvec) {
      Vec<T>::load(xr + vi * V, v[j]);
#pragma unroll
      for (int i = 0; i < V; ++i) s += v[j][i];
    }
  }
  s = warp_sum(s);
  const float mean = s / (float)C;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int vi = lane + j * 32;
    if (vi < nvec) {
#pragma unroll
      for (int i = 0; i < V; ++i) { float d = v[j][i] - mean; q = fmaf(d, d, q); }
    }
  }
  q = warp_sum(q);
  const float rstd = rsqrtf(q / (float)C + eps);
  TO* orow = out + (long long)warp * ldo;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int vi = lane + j * 32;
    if (vi < nvec) {
      float y[V];
#pragma unroll
      for (int i = 0; i < V; ++i) {
        const int c = vi * V + i;
        y[i] = (v[j][i] - mean) * rstd * gamma[c] + beta[c];
      }
      Vec<TO>::store(orow + vi * V, y);
    }
  }
}

template <typename T, typename TO>
int launch_gn(GnP p, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const int C = p.C1 + p.C2;
  const int CV = C / V;
  const int CVr = ((CV + 31) / 32) * 32;
  MDX_CHECK_ARG(CVr <= 1024, "mdx_groupnorm: too many channels (%d)", C);
  int RP = 512 / CVr;
  if (RP < 1) RP = 1;
  if (RP > 16) RP = 16;
  // enough blocks to fill the machine, but at least 4*RP rows each
  int chunks = (p.HW + 4 * RP - 1) / (4 * RP);
  int want = (4 * 148 + p.n - 1) / p.n;
  if (chunks > want) chunks = want;
  if (chunks < 1) chunks = 1;
  p.rows_per_block = (p.HW + chunks - 1) / chunks;
  chunks = (p.HW + p.rows_per_block - 1) / p.rows_per_block;
  dim3 grid(chunks, p.n), block(CVr, RP);
  MDX_CHECK_CUDA(cudaMemsetAsync(p.stats, 0, sizeof(double) * 2 * p.groups * p.n, st));
  gn_stats_kernel<T><<<grid, block, 0, st>>>(p);
  MDX_CHECK_LAUNCH();
  gn_apply_kernel<T, TO><<<grid, block, 0, st>>>(p);
  MDX_CHECK_LAUNCH();
  g_mdx_launches.fetch_add(2, std::memory_order_relaxed);
  return MDX_OK;
}

}  // namespace

extern "C" int mdx_groupnorm(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2,
                             int32_t n, int32_t HW, int32_t groups, float eps, const float* gamma,
                             const float* beta, int32_t silu, void* out, int64_t ldo, void* raw_copy, int64_t ldraw,
                             int32_t in_dtype, int32_t out_dtype, double* stats_scratch, void* stream) {
  MDX_CHECK_ARG(x1 && out && gamma && beta && stats_scratch, "mdx_groupnorm: NULL operand");
  MDX_CHECK_ARG(n > 0 && HW > 0 && C1 > 0 && C2 >= 0, "mdx_groupnorm: empty problem");
  MDX_CHECK_ARG(C2 == 0 || x2 != nullptr, "mdx_groupnorm: x2 is NULL but C2 > 0");
  const int C = C1 + C2;
  MDX_CHECK_ARG(groups > 0 && groups <= 64 && C % groups == 0, "mdx_groupnorm: C=%d not divisible by groups=%d", C,
                groups);
  MDX_CHECK_ARG((in_dtype == MDX_BF16 || in_dtype == MDX_F32) && (out_dtype == MDX_BF16 || out_dtype == MDX_F32),
                "mdx_groupnorm: bad dtype");
  MDX_CHECK_ARG(!(in_dtype == MDX_BF16 && out_dtype == MDX_F32), "mdx_groupnorm: bf16 -> fp32 is not supported");
  MDX_CHECK_ARG(C1 % 8 == 0 && C2 % 8 == 0 && ld1 % 8 == 0 && ld2 % 8 == 0 && ldo % 8 == 0 && ldraw % 8 == 0,
                "mdx_groupnorm: channels and row strides must be multiples of 8");
  MDX_CHECK_ARG(((uintptr_t)x1 & 15) == 0 && ((uintptr_t)x2 & 15) == 0 && ((uintptr_t)out & 15) == 0 &&
                    ((uintptr_t)raw_copy & 15) == 0,
                "mdx_groupnorm: operands must be 16-byte aligned");
  GnP p;
  p.x1 = x1; p.ld1 = ld1; p.C1 = C1; p.x2 = x2; p.ld2 = ld2; p.C2 = C2;
  p.n = n; p.HW = HW; p.groups = groups; p.cpg = C / groups; p.rows_per_block = 0;
  p.eps = eps; p.gamma = gamma; p.beta = beta; p.silu = silu; p.out = out; p.ldo = ldo;
  p.raw = raw_copy; p.ldraw = ldraw; p.stats = stats_scratch;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (in_dtype == MDX_BF16) return launch_gn<bf16, bf16>(p, st);
  if (out_dtype == MDX_BF16) return launch_gn<float, bf16>(p, st);
  return launch_gn<float, float>(p, st);
}

extern "C" int mdx_layernorm(const void* x, int64_t ldx, int32_t rows, int32_t C, float eps, const float* gamma,
                             const float* beta, void* out, int64_t ldo, int32_t in_dtype, int32_t out_dtype,
                             void* stream) {
  MDX_CHECK_ARG(x && out && gamma && beta, "mdx_layernorm: NULL operand");
  MDX_CHECK_ARG(rows > 0 && C > 0, "mdx_layernorm: empty problem");
  MDX_CHECK_ARG((in_dtype == MDX_BF16 || in_dtype == MDX_F32) && (out_dtype == MDX_BF16 || out_dtype == MDX_F32),
                "mdx_layernorm: bad dtype");
  MDX_CHECK_ARG(!(in_dtype == MDX_BF16 && out_dtype == MDX_F32), "mdx_layernorm: bf16 -> fp32 is not supported");
  MDX_CHECK_ARG(C % 8 == 0 && ldx % 8 == 0 && ldo % 8 == 0, "mdx_layernorm: C and strides must be multiples of 8");
  MDX_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0, "mdx_layernorm: 16-byte alignment");
  MDX_CHECK_ARG(C / 8 <= 8 * 32, "mdx_layernorm: C=%d too large (max 2048)", C);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int wpb = 8;
  dim3 grid((rows + wpb - 1) / wpb), block(wpb * 32);
  if (in_dtype == MDX_BF16)
    layernorm_kernel<bf16, bf16, 8><<<grid, block, 0, st>>>((const bf16*)x, ldx, rows, C, eps, gamma, beta, (bf16*)out, ldo);
  else if (out_dtype == MDX_BF16)
    layernorm_kernel<float, bf16, 8><<<grid, block, 0, st>>>((const float*)x, ldx, rows, C, eps, gamma, beta, (bf16*)out, ldo);
  else
    layernorm_kernel<float, float, 8><<<grid, block, 0, st>>>((const float*)x, ldx, rows, C, eps, gamma, beta, (float*)out, ldo);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

// ---------------------------------------------------------------------------------------------------------
// mdx_groupnorm_noalloc: GroupNorm for callers that cannot hand us a scratch buffer.
//
// mdx_groupnorm needs n*groups*2 doubles of device scratch for the per-group sums, and expects the caller to
// own them.  Some callers have no way to obtain device memory at the point of the call: a custom-op body that
// a framework invokes with nothing but its input/output addresses and a stream (the MindSpore `Custom` aot
// form INTEGRATION.md describes gets no workspace), a host thread that must not touch the allocator because
// the framework's memory pool is frozen, or code that runs inside a region where cudaMalloc is forbidden.
// For them the statistics live in the fixed array below, which is part of the library image: every GPU that
// loads libmdx_b200.so carries MDX_GN_NOALLOC_MAX_PAIRS * 16 bytes (256 KiB) of it until the process exits.
//
// The array's name is only a host-side handle in host code, so its device address is looked up with
// cudaGetSymbolAddress, once per device, with that device current, the first time the entry point is called
// there.  That first call does the lookup (it may also trigger the lazy load of this module), so make it before
// starting any stream capture.  Nothing here allocates: no cudaMalloc / cudaFree, no library routine that does.
//
// All calls on a device share the one array and it keeps whatever the previous call left in it.  Each call
// clears the part it uses (on the stream) before accumulating, so back-to-back calls on one stream are safe;
// two calls that can overlap on the GPU (different streams, no ordering between them) race on the array and
// must not be issued.
// ---------------------------------------------------------------------------------------------------------
#define MDX_GN_NOALLOC_MAX_PAIRS 16384  /* n * groups must not exceed this */
static __device__ __align__(256) double g_gn_noalloc_stats[MDX_GN_NOALLOC_MAX_PAIRS * 2];

#include <mutex>

namespace {
constexpr int kMaxDevices = 64;
double* g_gn_noalloc_ptr[kMaxDevices] = {nullptr};
std::mutex g_gn_noalloc_mu;

// Device address of g_gn_noalloc_stats on the CURRENT device (resolved once per device).
int gn_noalloc_scratch(double** out) {
  int dev = -1;
  MDX_CHECK_CUDA(cudaGetDevice(&dev));
  MDX_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "mdx_groupnorm_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_gn_noalloc_mu);
  if (g_gn_noalloc_ptr[dev] == nullptr) {
    double* p = nullptr;
    MDX_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_noalloc_stats));
    MDX_CHECK_ARG(p != nullptr, "mdx_groupnorm_noalloc: cudaGetSymbolAddress returned NULL");
    g_gn_noalloc_ptr[dev] = p;
  }
  *out = g_gn_noalloc_ptr[dev];
  return MDX_OK;
}
}  // namespace

extern "C" int mdx_groupnorm_noalloc(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2,
                                     int32_t C2, int32_t n, int32_t HW, int32_t groups, float eps,
                                     const float* gamma, const float* beta, int32_t silu, void* out, int64_t ldo,
                                     void* raw_copy, int64_t ldraw, int32_t in_dtype, int32_t out_dtype,
                                     void* stream) {
  MDX_CHECK_ARG(n > 0 && groups > 0, "mdx_groupnorm_noalloc: empty problem");
  MDX_CHECK_ARG((long long)n * groups <= MDX_GN_NOALLOC_MAX_PAIRS,
                "mdx_groupnorm_noalloc: n*groups = %lld exceeds the built-in scratch (%d); use mdx_groupnorm with "
                "caller-provided scratch",
                (long long)n * groups, MDX_GN_NOALLOC_MAX_PAIRS);
  double* scratch = nullptr;
  int rc = gn_noalloc_scratch(&scratch);
  if (rc != MDX_OK) return rc;
  // mdx_groupnorm zeroes the n*groups*2 doubles it is about to use (cudaMemsetAsync on `stream`) before the
  // statistics kernel accumulates into them, so the previous call's contents never leak into this one.
  return mdx_groupnorm(x1, ld1, C1, x2, ld2, C2, n, HW, groups, eps, gamma, beta, silu, out, ldo, raw_copy, ldraw,
                       in_dtype, out_dtype, scratch, stream);
}
