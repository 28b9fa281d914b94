// Row softmax and token embedding: the two small operators the once-per-prompt networks (VAE decoder, text tower)
// need beyond the UNet's kernel set.  See include/mdx.h (mdx_softmax_rows, mdx_embedding).
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

// One CTA per row; the row (up to 16 K columns per pass) is read twice from L2/HBM: max+sum online, then normalise.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) softmax_rows_kernel(const TI* __restrict__ x, long long ldx, int cols,
                                                           float scale, TO* __restrict__ out, long long ldo) {
  __shared__ float red[2][8];
  const long long row = blockIdx.x;
  const TI* xr = x + row * ldx;
  TO* orow = out + row * ldo;
  const float sl2 = scale * 1.4426950408889634f;
  float m = -INFINITY, l = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    const float v = to_f32<TI>(xr[c]) * sl2;
    const float mn = fmaxf(m, v);
    l = l * exp2f(m - mn) + exp2f(v - mn);
    m = mn;
  }
  // block-wide (m, l) merge
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float wm = warp_max(m);
  float wl = warp_sum(m == -INFINITY ? 0.f : l * exp2f(m - wm));  // lanes / warps that saw no column contribute 0
  if (lane == 0) { red[0][warp] = wm; red[1][warp] = wl; }
  __syncthreads();
  float bm = -INFINITY;
#pragma unroll
  for (int i = 0; i < 8; ++i) bm = fmaxf(bm, red[0][i]);
  float bl = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) bl += red[0][i] == -INFINITY ? 0.f : red[1][i] * exp2f(red[0][i] - bm);
  const float inv = 1.f / bl;
  for (int c = threadIdx.x; c < cols; c += blockDim.x)
    orow[c] = from_f32<TO>(exp2f(to_f32<TI>(xr[c]) * sl2 - bm) * inv);
}

template <typename TO>
__global__ void embedding_kernel(const float* __restrict__ table, int vocab, int width, const float* __restrict__ pos,
                                 int L, const int32_t* __restrict__ ids, int n, TO* __restrict__ out) {
  const long long total = (long long)n * width;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tok = (int)(i / width), c = (int)(i % width);
    int id = ids[tok];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    out[i] = from_f32<TO>(table[(long long)id * width + c] + pos[(long long)(tok % L) * width + c]);
  }
}

}  // namespace

extern "C" int mdx_softmax_rows(const void* x, int64_t ldx, int64_t rows, int32_t cols, float scale, void* out,
                                int64_t ldo, int32_t in_dtype, int32_t out_dtype, void* stream) {
  MDX_CHECK_ARG(x && out && rows > 0 && cols > 0, "mdx_softmax_rows: bad arguments");
  MDX_CHECK_ARG(rows < (1ll << 31), "mdx_softmax_rows: too many rows");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((unsigned)rows), block(256);
  if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    softmax_rows_kernel<float, float><<<grid, block, 0, st>>>((const float*)x, ldx, cols, scale, (float*)out, ldo);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    softmax_rows_kernel<float, bf16><<<grid, block, 0, st>>>((const float*)x, ldx, cols, scale, (bf16*)out, ldo);
  else if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    softmax_rows_kernel<bf16, bf16><<<grid, block, 0, st>>>((const bf16*)x, ldx, cols, scale, (bf16*)out, ldo);
  else {
    mdx_set_error("mdx_softmax_rows: unsupported dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_embedding(const float* table, int32_t vocab, int32_t width, const float* pos, int32_t L,
                             const int32_t* ids, int32_t n, void* out, int32_t dtype, void* stream) {
  MDX_CHECK_ARG(table && pos && ids && out && vocab > 0 && width > 0 && L > 0 && n > 0, "mdx_embedding: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * width;
  unsigned grid = (unsigned)((total + 255) / 256 > 148 * 16 ? 148 * 16 : (total + 255) / 256);
  if (dtype == MDX_F32)
    embedding_kernel<float><<<grid, 256, 0, st>>>(table, vocab, width, pos, L, ids, n, (float*)out);
  else if (dtype == MDX_BF16)
    embedding_kernel<bf16><<<grid, 256, 0, st>>>(table, vocab, width, pos, L, ids, n, (bf16*)out);
  else {
    mdx_set_error("mdx_embedding: bad dtype %d", dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}
