// GroupNorm(+SiLU) and LayerNorm over NHWC rows: HBM-bound, 16-byte vectorised, coalesced along channels.
// See include/mdx.h (mdx_groupnorm, mdx_layernorm) for the reference call sites.
//
// GroupNorm = two passes over the activation (the minimum without a persistent cooperative kernel):
//   gn_stats_kernel : each thread owns 8 consecutive channels and walks pixel rows (4 loads in flight); per-channel
//                     fp32 partial sums -> per-group fp64 smem bins -> fp64 atomics into stats[n][group][{sum,sumsq}]
//   gn_apply_kernel : y = x * a_c + b_c (a,b folded from mean/rstd/gamma/beta), optional SiLU, store.
// Both take two sources so the UNet skip concat (openaimodel.py:566-571) is normalised without a copy.
// Thread mapping: a 1-D block of CV*RP threads, CV = C/8 channel vectors x RP row lanes — no padding lanes, so a
// 320-channel row (40 vectors) keeps every thread busy.
//
// LayerNorm: one warp handles R rows at a time, rows held in registers (NV float4 per lane per row), gamma/beta
// loaded once per warp iteration; two-pass mean / variance in fp32.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>
#include <stdlib.h>

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
  int CV, RP;             // channel vectors per row, row lanes per block
  int rows_per_block;
  float eps;
  const float* gamma; const float* beta;
  int silu;
  void* out; long long ldo;
  void* raw; long long ldraw;  // optional un-normalised copy of concat(x1, x2) in the out dtype
  double* stats;
  const float* chan1; const float* chan2;  // optional per-image [slots][2][C] partial {sums | sums of squares}
  int slots1, slots2;                      // written by the producers (mdx_gemm_args.chan_stats)
  const float* ss; long long ldss;  // optional FiLM rows [n][scale(C) | shift(C)]: y = gn(x) * (1 + scale) + shift
};

template <typename T>
__device__ __forceinline__ const T* gn_src(const GnP& p, int c0, int img, long long* ld) {
  if (c0 < p.C1) {
    *ld = p.ld1;
    return reinterpret_cast<const T*>(p.x1) + (long long)img * p.HW * p.ld1 + c0;
  }
  *ld = p.ld2;
  return reinterpret_cast<const T*>(p.x2) + (long long)img * p.HW * p.ld2 + (c0 - p.C1);
}

template <typename T>
__global__ void __launch_bounds__(512) gn_stats_kernel(const GnP p) {
  constexpr int V = 8;
  // per-thread partials for the (at most two) groups its 8 channels touch: {sum_lo, sq_lo, sum_hi, sq_hi}.
  // Requires cpg >= 4 so that 8 channels span <= 3 groups... (cpg >= 8 -> <= 2 groups; smaller cpg handled below)
  __shared__ float4 part[512];
  __shared__ double bins[64][2];
  const int cv = threadIdx.x % p.CV, rl = threadIdx.x / p.CV;
  const int c0 = cv * V;
  const int img = blockIdx.y;
  long long ld;
  const T* base = gn_src<T>(p, c0, img, &ld);
  const int r0 = blockIdx.x * p.rows_per_block;
  const int r1 = min(p.HW, r0 + p.rows_per_block);
  const int RP = p.RP;
  float s[V], ss[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { s[i] = 0.f; ss[i] = 0.f; }
  int r = r0 + rl;
  for (; r + 3 * RP < r1; r += 4 * RP) {
    float v[4][V];
#pragma unroll
    for (int u = 0; u < 4; ++u) Vec<T>::load(base + (long long)(r + u * RP) * ld, v[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int i = 0; i < V; ++i) { s[i] += v[u][i]; ss[i] = fmaf(v[u][i], v[u][i], ss[i]); }
  }
  for (; r < r1; r += RP) {
    float v[V];
    Vec<T>::load(base + (long long)r * ld, v);
#pragma unroll
    for (int i = 0; i < V; ++i) { s[i] += v[i]; ss[i] = fmaf(v[i], v[i], ss[i]); }
  }
  if (p.cpg >= 8) {
    // deterministic block reduction, no atomics: 8 channels touch at most two groups
    const int g_lo = c0 / p.cpg;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < V; ++i) {
      if ((c0 + i) / p.cpg == g_lo) { acc.x += s[i]; acc.y += ss[i]; }
      else { acc.z += s[i]; acc.w += ss[i]; }
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int g = threadIdx.x; g < p.groups; g += blockDim.x) {
      const int cv_first = (g * p.cpg) / V, cv_last = (g * p.cpg + p.cpg - 1) / V;
      double ds = 0.0, dq = 0.0;
      for (int c = cv_first; c <= cv_last; ++c) {
        const int glo = (c * V) / p.cpg;
        for (int l = 0; l < RP; ++l) {
          const float4 q = part[l * p.CV + c];
          if (glo == g) { ds += q.x; dq += q.y; }
          else if (glo + 1 == g) { ds += q.z; dq += q.w; }
        }
      }
      atomicAdd(&p.stats[((long long)img * p.groups + g) * 2], ds);
      atomicAdd(&p.stats[((long long)img * p.groups + g) * 2 + 1], dq);
    }
    return;
  }
  // narrow groups (cpg < 8, e.g. 192 channels): fp64 smem bins with atomics
  for (int i = threadIdx.x; i < p.groups * 2; i += blockDim.x) (&bins[0][0])[i] = 0.0;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int g = (c0 + i) / p.cpg;
    atomicAdd(&bins[g][0], (double)s[i]);
    atomicAdd(&bins[g][1], (double)ss[i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < p.groups * 2; i += blockDim.x)
    atomicAdd(&p.stats[(long long)img * p.groups * 2 + i], (&bins[0][0])[i]);
}

#ifndef MDX_GN_APPLY_BLOCKS
#define MDX_GN_APPLY_BLOCKS 2   // resident blocks per SM the apply kernel is compiled for (register cap 64 at 2)
#endif
template <typename T, typename TO, bool SILU>
__global__ void __launch_bounds__(512, MDX_GN_APPLY_BLOCKS) gn_apply_kernel(const GnP p) {
  constexpr int V = 8;
  const int cv = threadIdx.x % p.CV, rl = threadIdx.x / p.CV;
  const int c0 = cv * V;
  const int img = blockIdx.y;
  long long ld;
  const T* base = gn_src<T>(p, c0, img, &ld);
  float a[V], b[V];
  const double cnt = (double)p.HW * p.cpg;
  // mean / rstd once per distinct group of this thread's channels (fp64 only for E[x^2] - mean^2)
  int g_prev = -1;
  float mean_f = 0.f, rstd_f = 0.f;
  const double inv_cnt = 1.0 / cnt;
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = c0 + i;
    const int g = c / p.cpg;
    if (g != g_prev) {
      g_prev = g;
      const double mean = p.stats[((long long)img * p.groups + g) * 2] * inv_cnt;
      double var = p.stats[((long long)img * p.groups + g) * 2 + 1] * inv_cnt - mean * mean;
      if (var < 0.0) var = 0.0;
      mean_f = (float)mean;
      rstd_f = rsqrtf((float)var + p.eps);
    }
    a[i] = rstd_f * __ldg(p.gamma + c);
    b[i] = __ldg(p.beta + c) - mean_f * a[i];
    if (p.ss) {
      const float* row = p.ss + (long long)img * p.ldss;
      const float sc = 1.f + __ldg(row + c);
      a[i] *= sc;
      b[i] = fmaf(b[i], sc, __ldg(row + p.C1 + p.C2 + c));
    }
  }
  const int r0 = blockIdx.x * p.rows_per_block;
  const int r1 = min(p.HW, r0 + p.rows_per_block);
  const int RP = p.RP;
  TO* obase = reinterpret_cast<TO*>(p.out) + (long long)img * p.HW * p.ldo + c0;
  TO* rbase = p.raw ? reinterpret_cast<TO*>(p.raw) + (long long)img * p.HW * p.ldraw + c0 : nullptr;
  auto finish = [&](int row, float (&v)[V]) {
    if (rbase) Vec<TO>::store(rbase + (long long)row * p.ldraw, v);
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float y = fmaf(v[i], a[i], b[i]);
      if constexpr (!SILU) v[i] = y;
      else if constexpr (sizeof(TO) == 2) v[i] = silu_fast(y);   // bf16 operand of the next GEMM
      else v[i] = silu_f(y);                                     // fp32 check mode
    }
    Vec<TO>::store(obase + (long long)row * p.ldo, v);
  };
  // rows in flight per thread: 4 x 16 B for bf16 input, 2 x 32 B for fp32 input (the same bytes; the fp32 rows cost
  // twice the registers, and the kernel has to stay within 64 registers to keep two blocks resident per SM)
  constexpr int U = sizeof(T) == 2 ? 4 : 2;
  int r = r0 + rl;
  for (; r + (U - 1) * RP < r1; r += U * RP) {
    float v[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u) Vec<T>::load(base + (long long)(r + u * RP) * ld, v[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) finish(r + u * RP, v[u]);
  }
  for (; r < r1; r += RP) {
    float v[V];
    Vec<T>::load(base + (long long)r * ld, v);
    finish(r, v);
  }
}

// Partial sums written by the producing GEMMs' epilogues (mdx_gemm_args.chan_stats: per image [slots][2][C], one slot
// per 32 output rows) -> per-group sums.  One warp per (image, group), fixed summation order (deterministic); a group
// of concat(x1, x2) may take channels from both tensors.
__global__ void __launch_bounds__(128) gn_fold_kernel(const GnP p) {
  __shared__ double red[2][4];
  const int i = blockIdx.x;                       // (image, group)
  const int img = i / p.groups, g = i % p.groups;
  const int c_lo = g * p.cpg;
  // items = (slot, channel of the group); a thread walks them with stride 128, four loads in flight
  float fs[4] = {0.f, 0.f, 0.f, 0.f}, fq[4] = {0.f, 0.f, 0.f, 0.f};
  auto item = [&](int it, float& s, float& q) {
    const int c = c_lo + it % p.cpg;
    const bool first = c < p.C1;
    const int C = first ? p.C1 : p.C2, cc = first ? c : c - p.C1, slots = first ? p.slots1 : p.slots2;
    const int k = it / p.cpg;
    if (k < slots) {
      const float* src = (first ? p.chan1 : p.chan2) + ((long long)img * slots + k) * 2 * C + cc;
      s += src[0];
      q += src[C];
    }
  };
  const int slots_max = p.slots1 > p.slots2 ? p.slots1 : p.slots2;
  const int total = slots_max * p.cpg;
  int it = threadIdx.x;
  for (; it + 3 * 128 < total; it += 4 * 128) {
#pragma unroll
    for (int u = 0; u < 4; ++u) item(it + u * 128, fs[u], fq[u]);
  }
  for (; it < total; it += 128) item(it, fs[0], fq[0]);
  double s = ((double)fs[0] + fs[1]) + ((double)fs[2] + fs[3]);
  double q = ((double)fq[0] + fq[1]) + ((double)fq[2] + fq[3]);
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, m);
    q += __shfl_xor_sync(0xffffffffu, q, m);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    p.stats[(long long)i * 2] = (red[0][0] + red[0][1]) + (red[0][2] + red[0][3]);
    p.stats[(long long)i * 2 + 1] = (red[1][0] + red[1][1]) + (red[1][2] + red[1][3]);
  }
}

// ---------------- LayerNorm: one warp per R rows, rows held in registers ----------------
template <typename T> __device__ __forceinline__ float4 ln_ld4(const T* p);
template <> __device__ __forceinline__ float4 ln_ld4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <> __device__ __forceinline__ float4 ln_ld4<bf16>(const bf16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
  return make_float4(a.x, a.y, b.x, b.y);
}
template <typename T> __device__ __forceinline__ void ln_st4(T* p, float4 v);
template <> __device__ __forceinline__ void ln_st4<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void ln_st4<bf16>(bf16* p, float4 v) {
  uint2 u;
  u.x = pack_bf16x2(v.x, v.y);
  u.y = pack_bf16x2(v.z, v.w);
  *reinterpret_cast<uint2*>(p) = u;
}

template <typename T, typename TO, int NV, int R>
__global__ void __launch_bounds__(256) layernorm_kernel(const T* __restrict__ x, long long ldx, int rows, int C,
                                                        float eps, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, TO* __restrict__ out,
                                                        long long ldo) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long row0 = (long long)warp * R;
  if (row0 >= rows) return;
  const int nvec = C >> 2;
  float4 v[R][NV];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool rok = row0 + r < rows;
    const T* xr = x + (row0 + r) * ldx;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = lane + j * 32;
      v[r][j] = (rok && vi < nvec) ? ln_ld4<T>(xr + vi * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  float mean[R], rstd[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) s += (v[r][j].x + v[r][j].y) + (v[r][j].z + v[r][j].w);
    mean[r] = warp_sum(s) / (float)C;
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      if (lane + j * 32 < nvec) {
        const float dx = v[r][j].x - mean[r], dy = v[r][j].y - mean[r], dz = v[r][j].z - mean[r], dw = v[r][j].w - mean[r];
        q += (dx * dx + dy * dy) + (dz * dz + dw * dw);
      }
    }
    rstd[r] = rsqrtf(warp_sum(q) / (float)C + eps);
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vi = lane + j * 32;
    if (vi < nvec) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + vi * 4));
      const float4 b = __ldg(reinterpret_cast<const float4*>(beta + vi * 4));
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (row0 + r < rows) {
          float4 y;
          y.x = (v[r][j].x - mean[r]) * rstd[r] * g.x + b.x;
          y.y = (v[r][j].y - mean[r]) * rstd[r] * g.y + b.y;
          y.z = (v[r][j].z - mean[r]) * rstd[r] * g.z + b.z;
          y.w = (v[r][j].w - mean[r]) * rstd[r] * g.w + b.w;
          ln_st4<TO>(out + (row0 + r) * ldo + vi * 4, y);
        }
      }
    }
  }
}

template <typename T, typename TO>
int launch_gn(GnP p, cudaStream_t st) {
  const int C = p.C1 + p.C2;
  p.CV = C / 8;
  MDX_CHECK_ARG(p.CV <= 512, "mdx_groupnorm: too many channels (%d)", C);
  static const int blk = getenv("MDX_GN_BLOCK") ? atoi(getenv("MDX_GN_BLOCK")) : 256;   // threads per block: four small blocks per SM hide each other's prologue better than two of 512 (profiles/r02_gn_apply_ab_block.txt)
  p.RP = (p.CV <= blk ? blk : 512) / p.CV;
  if (p.RP < 1) p.RP = 1;
  if (p.RP > 16) p.RP = 16;
  // a few waves of blocks (MDX_GN_APPLY_BLOCKS resident per SM): enough that the last, partly filled wave is cheap,
  // few enough that a block's prologue (statistics -> per-channel scale / shift) is amortised over many rows; each
  // block walks at least 8*RP rows (4-deep unrolled loads twice)
  int chunks = (p.HW + 8 * p.RP - 1) / (8 * p.RP);
  static const int waves = getenv("MDX_GN_WAVES") ? atoi(getenv("MDX_GN_WAVES")) : 4;   // A/B switch
  int want = (MDX_GN_APPLY_BLOCKS * (512 / (p.CV * p.RP) > 1 ? 512 / (p.CV * p.RP) : 1) * waves * mdx_num_sms() + p.n - 1) / p.n;
  if (chunks > want) chunks = want;
  if (chunks < 1) chunks = 1;
  p.rows_per_block = (p.HW + chunks - 1) / chunks;
  chunks = (p.HW + p.rows_per_block - 1) / p.rows_per_block;
  dim3 grid(chunks, p.n), block(p.CV * p.RP);
  if (p.chan1) {
    gn_fold_kernel<<<p.n * p.groups, 128, 0, st>>>(p);
  } else {
    MDX_CHECK_CUDA(cudaMemsetAsync(p.stats, 0, sizeof(double) * 2 * p.groups * p.n, st));
    gn_stats_kernel<T><<<grid, block, 0, st>>>(p);
  }
  MDX_CHECK_LAUNCH();
  if (p.silu) gn_apply_kernel<T, TO, true><<<grid, block, 0, st>>>(p);
  else gn_apply_kernel<T, TO, false><<<grid, block, 0, st>>>(p);
  MDX_CHECK_LAUNCH();
  g_mdx_launches.fetch_add(2, std::memory_order_relaxed);
  return MDX_OK;
}

template <typename T, typename TO, int NV, int R>
void launch_ln_one(const void* x, long long ldx, int rows, int C, float eps, const float* gamma, const float* beta,
                   void* out, long long ldo, cudaStream_t st) {
  const int wpb = 8;
  const long long warps = ((long long)rows + R - 1) / R;
  dim3 grid((unsigned)((warps + wpb - 1) / wpb)), block(wpb * 32);
  layernorm_kernel<T, TO, NV, R><<<grid, block, 0, st>>>((const T*)x, ldx, rows, C, eps, gamma, beta, (TO*)out, ldo);
}

template <typename T, typename TO>
int launch_ln(const void* x, long long ldx, int rows, int C, float eps, const float* gamma, const float* beta,
              void* out, long long ldo, cudaStream_t st) {
  const int need = (C / 4 + 31) / 32;  // float4 vectors per lane
  if (need <= 1) launch_ln_one<T, TO, 1, 4>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else if (need <= 2) launch_ln_one<T, TO, 2, 4>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else if (need <= 3) launch_ln_one<T, TO, 3, 4>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else if (need <= 5) launch_ln_one<T, TO, 5, 2>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else if (need <= 8) launch_ln_one<T, TO, 8, 1>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else if (need <= 10) launch_ln_one<T, TO, 10, 1>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else if (need <= 16) launch_ln_one<T, TO, 16, 1>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  else {
    mdx_set_error("mdx_layernorm: C=%d too large (max 2048)", C);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

}  // namespace


static int gn_run(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2, int32_t n,
                  int32_t HW, int32_t groups, float eps, const float* gamma, const float* beta, int32_t silu,
                  void* out, int64_t ldo, void* raw_copy, int64_t ldraw, int32_t in_dtype, int32_t out_dtype,
                  const float* scale_shift, int64_t ld_ss, double* stats_scratch, void* stream,
                  const float* chan1 = nullptr, int slots1 = 0, const float* chan2 = nullptr, int slots2 = 0) {
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
  p.ss = scale_shift; p.ldss = ld_ss;
  p.chan1 = chan1; p.chan2 = chan2; p.slots1 = slots1; p.slots2 = slots2;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (in_dtype == MDX_BF16) return launch_gn<bf16, bf16>(p, st);
  if (out_dtype == MDX_BF16) return launch_gn<float, bf16>(p, st);
  return launch_gn<float, float>(p, st);
}

extern "C" int mdx_groupnorm(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2,
                             int32_t n, int32_t HW, int32_t groups, float eps, const float* gamma,
                             const float* beta, int32_t silu, void* out, int64_t ldo, void* raw_copy, int64_t ldraw,
                             int32_t in_dtype, int32_t out_dtype, double* stats_scratch, void* stream) {
  return gn_run(x1, ld1, C1, x2, ld2, C2, n, HW, groups, eps, gamma, beta, silu, out, ldo, raw_copy, ldraw, in_dtype,
                out_dtype, nullptr, 0, stats_scratch, stream);
}

extern "C" int mdx_groupnorm_film(const void* x, int64_t ldx, int32_t C, int32_t n, int32_t HW, int32_t groups,
                                  float eps, const float* gamma, const float* beta, const float* scale_shift,
                                  int64_t ld_ss, int32_t silu, void* out, int64_t ldo, int32_t in_dtype,
                                  int32_t out_dtype, double* stats_scratch, void* stream) {
  MDX_CHECK_ARG(scale_shift && ld_ss >= 2 * (int64_t)C, "mdx_groupnorm_film: scale_shift rows must hold 2*C floats");
  return gn_run(x, ldx, C, nullptr, 0, 0, n, HW, groups, eps, gamma, beta, silu, out, ldo, nullptr, 0, in_dtype,
                out_dtype, scale_shift, ld_ss, stats_scratch, stream);
}

extern "C" int mdx_groupnorm_pre(const void* x1, int64_t ld1, int32_t C1, const float* chan1, int32_t slots1,
                                 const void* x2, int64_t ld2, int32_t C2, const float* chan2, int32_t slots2, int32_t n, int32_t HW, int32_t groups,
                                 float eps, const float* gamma, const float* beta, const float* scale_shift,
                                 int64_t ld_ss, int32_t silu, void* out, int64_t ldo, void* raw_copy, int64_t ldraw,
                                 int32_t in_dtype, int32_t out_dtype, double* stats_scratch, void* stream) {
  MDX_CHECK_ARG(chan1 != nullptr && slots1 > 0 && (C2 == 0 || (chan2 != nullptr && slots2 > 0)),
                "mdx_groupnorm_pre: every source needs its channel statistics");
  MDX_CHECK_ARG(scale_shift == nullptr || ld_ss >= 2 * (int64_t)(C1 + C2), "mdx_groupnorm_pre: scale_shift rows too short");
  return gn_run(x1, ld1, C1, x2, ld2, C2, n, HW, groups, eps, gamma, beta, silu, out, ldo, raw_copy, ldraw, in_dtype,
                out_dtype, scale_shift, ld_ss, stats_scratch, stream, chan1, slots1, chan2, slots2);
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
  MDX_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)gamma & 15) == 0 &&
                    ((uintptr_t)beta & 15) == 0,
                "mdx_layernorm: 16-byte alignment");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (in_dtype == MDX_BF16) return launch_ln<bf16, bf16>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  if (out_dtype == MDX_BF16) return launch_ln<float, bf16>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
  return launch_ln<float, float>(x, ldx, rows, C, eps, gamma, beta, out, ldo, st);
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
