// Layout movers and small elementwise kernels (all HBM- or latency-bound). See include/mdx.h.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

inline unsigned blocks_for(long long n, int per_block) {
  long long b = (n + per_block - 1) / per_block;
  if (b > 148ll * 32) b = 148ll * 32;  // grid-stride beyond this
  if (b < 1) b = 1;
  return (unsigned)b;
}

// util.py:111-131: freqs = exp(-ln(max_period) * k / half); emb = [cos(t f) | sin(t f)]
__global__ void timestep_embedding_kernel(const float* __restrict__ t, int n, int dim, float max_period,
                                          float* __restrict__ out) {
  const int half = dim / 2;
  const long long total = (long long)n * dim;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / dim), j = (int)(i % dim);
    float v = 0.f;
    if (j < 2 * half) {
      const int k = j < half ? j : j - half;
      const float freq = expf(-logf(max_period) * (float)k / (float)half);
      const float arg = t[b] * freq;
      v = j < half ? cosf(arg) : sinf(arg);
    }
    out[i] = v;
  }
}

template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, int n, int C, int H, int W, int repeat, int Cpad,
                                    T* __restrict__ out) {
  const long long total = (long long)n * repeat * H * W * Cpad;
  const long long HW = (long long)H * W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % Cpad);
    const long long pix = i / Cpad;
    const long long hw = pix % HW;
    const int b = (int)(pix / HW);
    float v = 0.f;
    if (c < C) v = x[((long long)(b % n) * C + c) * HW + hw];
    out[i] = from_f32<T>(v);
  }
}

template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ x, int n, int C, int H, int W, long long ldx,
                                    float* __restrict__ out) {
  const long long HW = (long long)H * W;
  const long long total = (long long)n * C * HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long hw = i % HW;
    const int c = (int)((i / HW) % C);
    const int b = (int)(i / (HW * C));
    out[i] = to_f32<T>(x[((long long)b * HW + hw) * ldx + c]);
  }
}

// 4 channels per thread; TI -> TO conversion on the fly (fp32 trunk -> bf16 conv operand)
template <typename T> __device__ __forceinline__ void ld4(const T* p, float (&v)[4]);
template <> __device__ __forceinline__ void ld4<float>(const float* p, float (&v)[4]) {
  float4 a = *reinterpret_cast<const float4*>(p);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
template <> __device__ __forceinline__ void ld4<bf16>(const bf16* p, float (&v)[4]) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  float2 f = unpack_bf16x2(u.x); v[0] = f.x; v[1] = f.y;
  f = unpack_bf16x2(u.y); v[2] = f.x; v[3] = f.y;
}
template <typename T> __device__ __forceinline__ void st4(T* p, const float (&v)[4]);
template <> __device__ __forceinline__ void st4<float>(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
template <> __device__ __forceinline__ void st4<bf16>(bf16* p, const float (&v)[4]) {
  uint2 u; u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = u;
}

template <typename TI, typename TO>
__global__ void upsample2x_kernel(const TI* __restrict__ x, int n, int H, int W, int C, TO* __restrict__ out) {
  const int CV = C / 4;
  const long long total = (long long)n * 2 * H * 2 * W * CV;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    long long pix = i / CV;
    const int ox = (int)(pix % (2 * W));
    pix /= 2 * W;
    const int oy = (int)(pix % (2 * H));
    const int b = (int)(pix / (2 * H));
    float v[4];
    ld4<TI>(x + ((((long long)b * H + (oy >> 1)) * W + (ox >> 1)) * CV + cv) * 4, v);
    st4<TO>(out + i * 4, v);
  }
}

template <typename T>
__global__ void concat_kernel(const T* __restrict__ x1, long long ld1, int C1, const T* __restrict__ x2,
                              long long ld2, int C2, long long rows, T* __restrict__ out, long long ldo) {
  constexpr int V = 16 / sizeof(T);
  const int CV = (C1 + C2) / V, CV1 = C1 / V;
  const long long total = rows * CV;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    const long long r = i / CV;
    uint4 v = cv < CV1 ? *reinterpret_cast<const uint4*>(x1 + r * ld1 + cv * V)
                       : *reinterpret_cast<const uint4*>(x2 + r * ld2 + (cv - CV1) * V);
    *reinterpret_cast<uint4*>(out + r * ldo + cv * V) = v;
  }
}

template <typename TI, typename TO>
__global__ void im2col3x3_kernel(const TI* __restrict__ x, int n, int H, int W, int C, int stride, int OH, int OW,
                                 TO* __restrict__ out) {
  const int CV = C / 4;
  const long long total = (long long)n * OH * OW * 9 * CV;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    long long r = i / CV;
    const int tap = (int)(r % 9);
    r /= 9;
    const int ox = (int)(r % OW);
    r /= OW;
    const int oy = (int)(r % OH);
    const int b = (int)(r / OH);
    const int iy = oy * stride + tap / 3 - 1, ix = ox * stride + tap % 3 - 1;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (iy >= 0 && iy < H && ix >= 0 && ix < W) ld4<TI>(x + (((long long)b * H + iy) * W + ix) * C + cv * 4, v);
    st4<TO>(out + i * 4, v);
  }
}

template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ x, TO* __restrict__ out, long long count) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x)
    out[i] = from_f32<TO>(to_f32<TI>(x[i]));
}

template <typename T>
__global__ void geglu_kernel(const T* __restrict__ x, long long ldx, long long rows, int I, T* __restrict__ out,
                             long long ldo) {
  const long long total = rows * I;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % I);
    const long long r = i / I;
    const float a = to_f32<T>(x[r * ldx + j]);
    const float g = to_f32<T>(x[r * ldx + I + j]);
    out[r * ldo + j] = from_f32<T>(a * gelu_tanh_f(g));
  }
}

}  // namespace

#define LAUNCHED(nk)                                              \
  do {                                                            \
    g_mdx_launches.fetch_add((nk), std::memory_order_relaxed);    \
    MDX_CHECK_LAUNCH();                                           \
    return MDX_OK;                                                \
  } while (0)

extern "C" int mdx_timestep_embedding(const float* t, int32_t n, int32_t dim, float max_period, float* out,
                                      void* stream) {
  MDX_CHECK_ARG(t && out && n > 0 && dim > 1, "mdx_timestep_embedding: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  timestep_embedding_kernel<<<blocks_for((long long)n * dim, 256), 256, 0, st>>>(t, n, dim, max_period, out);
  LAUNCHED(1);
}

extern "C" int mdx_nchw_to_nhwc_pad(const float* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t repeat,
                                    int32_t Cpad, void* out, int32_t dtype, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && C > 0 && H > 0 && W > 0 && repeat > 0 && Cpad >= C, "mdx_nchw_to_nhwc: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * repeat * H * W * Cpad;
  if (dtype == MDX_BF16)
    nchw_to_nhwc_kernel<bf16><<<blocks_for(total, 256), 256, 0, st>>>(x, n, C, H, W, repeat, Cpad, (bf16*)out);
  else if (dtype == MDX_F32)
    nchw_to_nhwc_kernel<float><<<blocks_for(total, 256), 256, 0, st>>>(x, n, C, H, W, repeat, Cpad, (float*)out);
  else {
    mdx_set_error("mdx_nchw_to_nhwc: bad dtype %d", dtype);
    return MDX_ERR_INVALID;
  }
  LAUNCHED(1);
}
extern "C" int mdx_nchw_to_nhwc(const float* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t repeat, void* out,
                                int32_t dtype, void* stream) {
  return mdx_nchw_to_nhwc_pad(x, n, C, H, W, repeat, C, out, dtype, stream);
}

extern "C" int mdx_nhwc_to_nchw_ld(const void* x, int64_t ldx, int32_t n, int32_t C, int32_t H, int32_t W,
                                   int32_t dtype, float* out, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && C > 0 && H > 0 && W > 0 && ldx >= C, "mdx_nhwc_to_nchw: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * C * H * W;
  if (dtype == MDX_BF16)
    nhwc_to_nchw_kernel<bf16><<<blocks_for(total, 256), 256, 0, st>>>((const bf16*)x, n, C, H, W, ldx, out);
  else if (dtype == MDX_F32)
    nhwc_to_nchw_kernel<float><<<blocks_for(total, 256), 256, 0, st>>>((const float*)x, n, C, H, W, ldx, out);
  else {
    mdx_set_error("mdx_nhwc_to_nchw: bad dtype %d", dtype);
    return MDX_ERR_INVALID;
  }
  LAUNCHED(1);
}
extern "C" int mdx_nhwc_to_nchw(const void* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t dtype, float* out,
                                void* stream) {
  return mdx_nhwc_to_nchw_ld(x, C, n, C, H, W, dtype, out, stream);
}

extern "C" int mdx_upsample_nearest2x(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, void* out,
                                      int32_t in_dtype, int32_t out_dtype, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && H > 0 && W > 0 && C > 0, "mdx_upsample_nearest2x: bad arguments");
  MDX_CHECK_ARG(C % 4 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0, "mdx_upsample_nearest2x: alignment");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * 4 * H * W * (C / 4);
  const unsigned g = blocks_for(total, 256);
  if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    upsample2x_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)x, n, H, W, C, (bf16*)out);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    upsample2x_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, (bf16*)out);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    upsample2x_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, (float*)out);
  else {
    mdx_set_error("mdx_upsample_nearest2x: unsupported dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  LAUNCHED(1);
}

extern "C" int mdx_concat_channels(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2,
                                   int64_t rows, void* out, int64_t ldo, int32_t dtype, void* stream) {
  MDX_CHECK_ARG(x1 && x2 && out && rows > 0 && C1 > 0 && C2 > 0, "mdx_concat_channels: bad arguments");
  MDX_CHECK_ARG(dtype == MDX_BF16 || dtype == MDX_F32, "mdx_concat_channels: bad dtype");
  const int V = dtype == MDX_BF16 ? 8 : 4;
  MDX_CHECK_ARG(C1 % V == 0 && C2 % V == 0 && ld1 % V == 0 && ld2 % V == 0 && ldo % V == 0, "mdx_concat_channels: alignment");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = rows * ((C1 + C2) / V);
  if (dtype == MDX_BF16)
    concat_kernel<bf16><<<blocks_for(total, 256), 256, 0, st>>>((const bf16*)x1, ld1, C1, (const bf16*)x2, ld2, C2, rows, (bf16*)out, ldo);
  else
    concat_kernel<float><<<blocks_for(total, 256), 256, 0, st>>>((const float*)x1, ld1, C1, (const float*)x2, ld2, C2, rows, (float*)out, ldo);
  LAUNCHED(1);
}

extern "C" int mdx_im2col3x3(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, int32_t stride, void* out,
                             int32_t in_dtype, int32_t out_dtype, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && H > 0 && W > 0 && C > 0 && stride > 0, "mdx_im2col3x3: bad arguments");
  MDX_CHECK_ARG(C % 4 == 0, "mdx_im2col3x3: C must be a multiple of 4");
  const int OH = (H + 2 - 3) / stride + 1, OW = (W + 2 - 3) / stride + 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * OH * OW * 9 * (C / 4);
  const unsigned g = blocks_for(total, 256);
  if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    im2col3x3_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)x, n, H, W, C, stride, OH, OW, (bf16*)out);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    im2col3x3_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, stride, OH, OW, (bf16*)out);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    im2col3x3_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, stride, OH, OW, (float*)out);
  else {
    mdx_set_error("mdx_im2col3x3: unsupported dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  LAUNCHED(1);
}

extern "C" int mdx_cast(const void* x, int32_t in_dtype, void* out, int32_t out_dtype, int64_t count, void* stream) {
  MDX_CHECK_ARG(x && out && count > 0, "mdx_cast: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const unsigned g = blocks_for(count, 256);
  if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    cast_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)x, (bf16*)out, count);
  else if (in_dtype == MDX_BF16 && out_dtype == MDX_F32)
    cast_kernel<bf16, float><<<g, 256, 0, st>>>((const bf16*)x, (float*)out, count);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    cast_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, (float*)out, count);
  else if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    cast_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)x, (bf16*)out, count);
  else {
    mdx_set_error("mdx_cast: bad dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  LAUNCHED(1);
}

extern "C" int mdx_geglu(const void* x, int64_t ldx, int32_t rows, int32_t I, void* out, int64_t ldo, int32_t dtype,
                         void* stream) {
  MDX_CHECK_ARG(x && out && rows > 0 && I > 0, "mdx_geglu: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)rows * I;
  if (dtype == MDX_BF16)
    geglu_kernel<bf16><<<blocks_for(total, 256), 256, 0, st>>>((const bf16*)x, ldx, rows, I, (bf16*)out, ldo);
  else if (dtype == MDX_F32)
    geglu_kernel<float><<<blocks_for(total, 256), 256, 0, st>>>((const float*)x, ldx, rows, I, (float*)out, ldo);
  else {
    mdx_set_error("mdx_geglu: bad dtype %d", dtype);
    return MDX_ERR_INVALID;
  }
  LAUNCHED(1);
}
