// Small operators that only the Taichu-GLIDE family needs.  All HBM- or latency-bound.
// See include/mdx.h for the reference call sites of each entry point.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

inline unsigned grid_for(long long n, int per_block = 256) {
  long long b = (n + per_block - 1) / per_block;
  if (b > 148ll * 16) b = 148ll * 16;
  return (unsigned)(b < 1 ? 1 : b);
}

// nn.AvgPool2d(kernel_size=2, stride=2) on NHWC; 4 channels per thread
template <typename TI, typename TO>
__global__ void avgpool2x2_kernel(const TI* __restrict__ x, int n, int H, int W, int C, TO* __restrict__ out) {
  const int OH = H / 2, OW = W / 2, C4 = C / 4;
  const long long total = (long long)n * OH * OW * C4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C4) * 4;
    long long r = i / C4;
    const int ow = (int)(r % OW); r /= OW;
    const int oh = (int)(r % OH);
    const int b = (int)(r / OH);
    const TI* p = x + (((long long)b * H + 2 * oh) * W + 2 * ow) * C + c;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[k] += to_f32<TI>(p[((long long)dy * W + dx) * C + k]);
    TO* o = out + (((long long)b * OH + oh) * OW + ow) * C + c;
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = from_f32<TO>(acc[k] * 0.25f);
  }
}

// dst[b][r][0:width] = src[b][r][0:width]; 16-byte chunks; src_batch_stride 0 broadcasts one block of rows
__global__ void copy_rows_kernel(const uint4* __restrict__ src, long long ld_src, long long bs_src,
                                 uint4* __restrict__ dst, long long ld_dst, long long bs_dst, int nb, int rows,
                                 int wv) {
  const long long total = (long long)nb * rows * wv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % wv);
    const long long rr = i / wv;
    const int r = (int)(rr % rows);
    const int b = (int)(rr / rows);
    dst[b * bs_dst + r * ld_dst + v] = src[b * bs_src + r * ld_src + v];
  }
}

// text2im_model.py:79-84: xf_in = where(mask, table[id] + pos[l], padding[l])
template <typename TO>
__global__ void embedding_masked_kernel(const float* __restrict__ table, int vocab, int width,
                                        const float* __restrict__ pos, const float* __restrict__ pad, int L,
                                        const int32_t* __restrict__ ids, const int32_t* __restrict__ mask, int n,
                                        TO* __restrict__ out) {
  const long long total = (long long)n * width;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tok = (int)(i / width), c = (int)(i % width);
    const long long pl = (long long)(tok % L) * width + c;
    float v;
    if (mask == nullptr || mask[tok] != 0) {
      int id = ids[tok];
      id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
      v = table[(long long)id * width + c] + pos[pl];
    } else {
      v = pad[pl];
    }
    out[i] = from_f32<TO>(v);
  }
}

// ops.ResizeBilinear(size, align_corners=False) with half_pixel_centers=False (its default): src = dst * in/out,
// neighbours clamped to the last row/column.  quantize: gaussian_diffusion.py superResPreprocess applied to the
// source pixels first: round((v + 1) * 127.5) / 127.5 - 1.
__global__ void resize_bilinear_kernel(const float* __restrict__ x, int nc, int H, int W, int OH, int OW, int quantize,
                                       float* __restrict__ out) {
  const long long total = (long long)nc * OH * OW;
  const float sy = (float)H / (float)OH, sx = (float)W / (float)OW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int ox = (int)(i % OW);
    const int oy = (int)((i / OW) % OH);
    const long long pl = i / ((long long)OW * OH);
    const float fy = oy * sy, fx = ox * sx;
    const int y0 = (int)floorf(fy), x0 = (int)floorf(fx);
    const int y1 = min(y0 + 1, H - 1), x1 = min(x0 + 1, W - 1);
    const float wy = fy - y0, wx = fx - x0;
    const float* p = x + pl * H * W;
    float v00 = p[y0 * W + x0], v01 = p[y0 * W + x1], v10 = p[y1 * W + x0], v11 = p[y1 * W + x1];
    if (quantize) {
      v00 = rintf((v00 + 1.f) * 127.5f) / 127.5f - 1.f;
      v01 = rintf((v01 + 1.f) * 127.5f) / 127.5f - 1.f;
      v10 = rintf((v10 + 1.f) * 127.5f) / 127.5f - 1.f;
      v11 = rintf((v11 + 1.f) * 127.5f) / 127.5f - 1.f;
    }
    const float top = v00 + (v01 - v00) * wx, bot = v10 + (v11 - v10) * wx;
    out[i] = top + (bot - top) * wy;
  }
}

// NHWC (padded) image of concat(x1, x2) along channels; x2 may be NULL.  x1 [n] is read `repeat` times; x2 holds n2
// images (n: read `repeat` times as well; n*repeat: one per output image, e.g. [uncond | cond] concat conditioning).
template <typename T>
__global__ void nchw2_to_nhwc_kernel(const float* __restrict__ x1, int C1, const float* __restrict__ x2, int C2, int n,
                                     int n2, int H, int W, int repeat, int Cpad, T* __restrict__ out) {
  const long long HW = (long long)H * W;
  const long long total = (long long)n * repeat * HW * Cpad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % Cpad);
    const long long pix = i / Cpad;
    const long long hw = pix % HW;
    const int bb = (int)(pix / HW);
    const int b = bb % n;
    float v = 0.f;
    if (c < C1) v = x1[((long long)b * C1 + c) * HW + hw];
    else if (c < C1 + C2) v = x2[((long long)(bb % n2) * C2 + (c - C1)) * HW + hw];
    out[i] = from_f32<T>(v);
  }
}

// nn.PReLU(C) on NHWC rows (srgan.py:45,63,80,88): y = x > 0 ? x : slope[c] * x, optional "+ residual" first
// (the `conv1 + conv2` sum of Generator.construct feeds no PReLU, so residual is added AFTER the activation).
template <typename TI, typename TO>
__global__ void prelu_kernel(const TI* __restrict__ x, long long ldx, long long rows, int C, const float* __restrict__ slope,
                             const TI* __restrict__ residual, long long ldr, TO* __restrict__ out, long long ldo) {
  const long long total = rows * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long r = i / C;
    float v = to_f32<TI>(x[r * ldx + c]);
    if (slope) v = v > 0.f ? v : slope[c] * v;
    if (residual) v += to_f32<TI>(residual[r * ldr + c]);
    out[r * ldo + c] = from_f32<TO>(v);
  }
}

// ops.DepthToSpace(2) (srgan.py:62) on NHWC, fused with the PReLU that follows it:
// out[n, 2h+i, 2w+j, c] = prelu(in[n, h, w, (2i+j)*C + c])   (depth index = block offset major, "DCR")
template <typename TI, typename TO>
__global__ void depth_to_space2_kernel(const TI* __restrict__ x, int n, int H, int W, int C, const float* __restrict__ slope,
                                       TO* __restrict__ out) {
  const long long total = (long long)n * 4 * H * W * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long r = i / C;
    const int ow = (int)(r % (2 * W)); r /= 2 * W;
    const int oh = (int)(r % (2 * H));
    const int b = (int)(r / (2 * H));
    const int blk = (oh & 1) * 2 + (ow & 1);
    float v = to_f32<TI>(x[((((long long)b * H + (oh >> 1)) * W + (ow >> 1)) * 4 + blk) * C + c]);
    if (slope) v = v > 0.f ? v : slope[c] * v;
    out[i] = from_f32<TO>(v);
  }
}

struct GlideP {
  const void* model_out; long long ld;
  int cfg; float scale; int mode;
  const float* x; const float* noise;
  float recip, recipm1, c1, c2, min_log, max_log, sqrt_ab_prev, sigma, dir, nonzero;
  float* x_prev; float* pred_xstart;
  int P, C, HW;
};

// gaussian_diffusion.py PMeanVariance.construct (:158-181) + Guider (guider.py:60-70) + PSample (:49-63) or
// DDimSample (:82-103), one pass.  model_out rows: [cond P | uncond P] images, channels [eps C | var C].
template <typename T>
__global__ void glide_step_kernel(const GlideP p) {
  const long long total = (long long)p.P * p.C * p.HW;
  const T* m = reinterpret_cast<const T*>(p.model_out);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int hw = (int)(i % p.HW);
    const int c = (int)((i / p.HW) % p.C);
    const int b = (int)(i / ((long long)p.HW * p.C));
    const long long ic = ((long long)b * p.HW + hw) * p.ld + c;
    float eps = to_f32<T>(m[ic]);
    if (p.cfg) {
      const float eu = to_f32<T>(m[ic + (long long)p.P * p.HW * p.ld]);
      eps = eu + p.scale * (eps - eu);
    }
    const float xv = p.x[i];
    float x0 = p.recip * xv - p.recipm1 * eps;
    x0 = fminf(fmaxf(x0, -1.f), 1.f);
    float xp;
    if (p.mode == 0) {
      const float frac = (to_f32<T>(m[ic + p.C]) + 1.f) * 0.5f;
      const float logvar = frac * p.max_log + (1.f - frac) * p.min_log;
      xp = p.c1 * x0 + p.c2 * xv;
      if (p.nonzero != 0.f) xp += expf(0.5f * logvar) * p.noise[i];
    } else {
      const float e2 = (p.recip * xv - x0) / p.recipm1;
      xp = p.dir * e2 + x0 * p.sqrt_ab_prev;
      if (p.nonzero != 0.f && p.sigma != 0.f) xp += p.sigma * p.noise[i];
    }
    if (p.pred_xstart) p.pred_xstart[i] = x0;
    p.x_prev[i] = xp;
  }
}

}  // namespace

extern "C" int mdx_avgpool2x2(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, void* out, int32_t in_dtype,
                              int32_t out_dtype, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && H > 0 && W > 0 && C > 0, "mdx_avgpool2x2: bad arguments");
  MDX_CHECK_ARG(H % 2 == 0 && W % 2 == 0 && C % 4 == 0, "mdx_avgpool2x2: H, W must be even and C a multiple of 4");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * (H / 2) * (W / 2) * (C / 4);
  const unsigned g = grid_for(total);
  if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    avgpool2x2_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, (float*)out);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    avgpool2x2_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, (bf16*)out);
  else if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    avgpool2x2_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)x, n, H, W, C, (bf16*)out);
  else {
    mdx_set_error("mdx_avgpool2x2: unsupported dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_copy_rows(const void* src, int64_t ld_src_bytes, int64_t batch_stride_src_bytes, void* dst,
                             int64_t ld_dst_bytes, int64_t batch_stride_dst_bytes, int32_t n_batch, int32_t rows,
                             int64_t width_bytes, void* stream) {
  MDX_CHECK_ARG(src && dst && n_batch > 0 && rows > 0 && width_bytes > 0, "mdx_copy_rows: bad arguments");
  MDX_CHECK_ARG(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0 && width_bytes % 16 == 0 &&
                    ld_src_bytes % 16 == 0 && ld_dst_bytes % 16 == 0 && batch_stride_src_bytes % 16 == 0 &&
                    batch_stride_dst_bytes % 16 == 0,
                "mdx_copy_rows: pointers, widths and strides must be multiples of 16 bytes");
  MDX_CHECK_ARG(width_bytes / 16 < (1ll << 31), "mdx_copy_rows: row too wide");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int wv = (int)(width_bytes / 16);
  const long long total = (long long)n_batch * rows * wv;
  copy_rows_kernel<<<grid_for(total), 256, 0, st>>>((const uint4*)src, ld_src_bytes / 16, batch_stride_src_bytes / 16,
                                                    (uint4*)dst, ld_dst_bytes / 16, batch_stride_dst_bytes / 16,
                                                    n_batch, rows, wv);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_embedding_masked(const float* table, int32_t vocab, int32_t width, const float* pos,
                                    const float* pad, int32_t L, const int32_t* ids, const int32_t* mask, int32_t n,
                                    void* out, int32_t dtype, void* stream) {
  MDX_CHECK_ARG(table && pos && ids && out && vocab > 0 && width > 0 && L > 0 && n > 0,
                "mdx_embedding_masked: bad arguments");
  MDX_CHECK_ARG(mask == nullptr || pad != nullptr, "mdx_embedding_masked: a mask needs the padding embedding");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const unsigned g = grid_for((long long)n * width);
  if (dtype == MDX_F32)
    embedding_masked_kernel<float><<<g, 256, 0, st>>>(table, vocab, width, pos, pad, L, ids, mask, n, (float*)out);
  else if (dtype == MDX_BF16)
    embedding_masked_kernel<bf16><<<g, 256, 0, st>>>(table, vocab, width, pos, pad, L, ids, mask, n, (bf16*)out);
  else {
    mdx_set_error("mdx_embedding_masked: bad dtype %d", dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_resize_bilinear(const float* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t OH, int32_t OW,
                                   int32_t quantize_u8, float* out, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && C > 0 && H > 0 && W > 0 && OH > 0 && OW > 0, "mdx_resize_bilinear: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * C * OH * OW;
  resize_bilinear_kernel<<<grid_for(total), 256, 0, st>>>(x, n * C, H, W, OH, OW, quantize_u8, out);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_nchw2_to_nhwc_pad(const float* x1, int32_t C1, const float* x2, int32_t C2, int32_t n, int32_t n2,
                                     int32_t H, int32_t W, int32_t repeat, int32_t Cpad, void* out, int32_t dtype,
                                     void* stream) {
  MDX_CHECK_ARG(C2 == 0 || n2 == n || n2 == n * repeat, "mdx_nchw2_to_nhwc_pad: x2 must hold n or n*repeat images");
  MDX_CHECK_ARG(x1 && out && C1 > 0 && C2 >= 0 && (C2 == 0 || x2) && n > 0 && H > 0 && W > 0 && repeat > 0,
                "mdx_nchw2_to_nhwc_pad: bad arguments");
  MDX_CHECK_ARG(Cpad >= C1 + C2, "mdx_nchw2_to_nhwc_pad: Cpad=%d < C1+C2=%d", Cpad, C1 + C2);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const unsigned g = grid_for((long long)n * repeat * H * W * Cpad);
  if (dtype == MDX_F32)
    nchw2_to_nhwc_kernel<float><<<g, 256, 0, st>>>(x1, C1, x2, C2, n, n2 > 0 ? n2 : n, H, W, repeat, Cpad, (float*)out);
  else if (dtype == MDX_BF16)
    nchw2_to_nhwc_kernel<bf16><<<g, 256, 0, st>>>(x1, C1, x2, C2, n, n2 > 0 ? n2 : n, H, W, repeat, Cpad, (bf16*)out);
  else {
    mdx_set_error("mdx_nchw2_to_nhwc_pad: bad dtype %d", dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_sampler_glide_step(const void* model_out, int64_t ld_out, int32_t model_dtype, int32_t cfg,
                                      float scale, int32_t mode, const float* x, const float* coef_host,
                                      const float* noise, float* x_prev, float* pred_xstart, int32_t P, int32_t C,
                                      int32_t h, int32_t w, void* stream) {
  MDX_CHECK_ARG(model_out && x && coef_host && x_prev, "mdx_sampler_glide_step: NULL operand");
  MDX_CHECK_ARG(P > 0 && C > 0 && h > 0 && w > 0 && ld_out >= 2 * (int64_t)C,
                "mdx_sampler_glide_step: empty problem / ld_out < 2*C");
  MDX_CHECK_ARG(model_dtype == MDX_F32 || model_dtype == MDX_BF16, "mdx_sampler_glide_step: bad dtype");
  MDX_CHECK_ARG(mode == 0 || mode == 1, "mdx_sampler_glide_step: mode must be 0 (p_sample) or 1 (ddim)");
  GlideP p;
  p.model_out = model_out; p.ld = ld_out; p.cfg = cfg; p.scale = scale; p.mode = mode;
  p.x = x; p.noise = noise;
  p.recip = coef_host[0]; p.recipm1 = coef_host[1]; p.c1 = coef_host[2]; p.c2 = coef_host[3];
  p.min_log = coef_host[4]; p.max_log = coef_host[5];
  const float ab = coef_host[6], ab_prev = coef_host[7], eta = coef_host[8];
  p.nonzero = coef_host[9];
  p.sigma = 0.f; p.dir = 0.f; p.sqrt_ab_prev = sqrtf(ab_prev);
  if (mode == 1) {
    MDX_CHECK_ARG(ab > 0.f && ab < 1.f && ab_prev > 0.f && p.recipm1 != 0.f, "mdx_sampler_glide_step: bad alphas");
    // gaussian_diffusion.py:85-91 as written: sigma = eta * sqrt((1-ab_prev)/(1-ab)) * sqrt((1-ab)/ab_prev)
    p.sigma = eta * sqrtf((1.f - ab_prev) / (1.f - ab)) * sqrtf((1.f - ab) / ab_prev);
    const float d2 = 1.f - ab_prev - p.sigma * p.sigma;
    p.dir = sqrtf(d2 > 0.f ? d2 : 0.f);
  }
  const bool needs_noise = p.nonzero != 0.f && (mode == 0 || p.sigma != 0.f);
  MDX_CHECK_ARG(!needs_noise || noise, "mdx_sampler_glide_step: this step draws noise but noise is NULL");
  p.x_prev = x_prev; p.pred_xstart = pred_xstart; p.P = P; p.C = C; p.HW = h * w;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)P * C * h * w;
  if (model_dtype == MDX_BF16) glide_step_kernel<bf16><<<grid_for(total), 256, 0, st>>>(p);
  else glide_step_kernel<float><<<grid_for(total), 256, 0, st>>>(p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_prelu(const void* x, int64_t ldx, int64_t rows, int32_t C, const float* slope, const void* residual,
                         int64_t ldr, void* out, int64_t ldo, int32_t in_dtype, int32_t out_dtype, void* stream) {
  MDX_CHECK_ARG(x && out && rows > 0 && C > 0 && (slope || residual), "mdx_prelu: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const unsigned g = grid_for(rows * C);
  if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    prelu_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, ldx, rows, C, slope, (const float*)residual, ldr, (float*)out, ldo);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    prelu_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)x, ldx, rows, C, slope, (const float*)residual, ldr, (bf16*)out, ldo);
  else if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    prelu_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)x, ldx, rows, C, slope, (const bf16*)residual, ldr, (bf16*)out, ldo);
  else {
    mdx_set_error("mdx_prelu: unsupported dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_depth_to_space2(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, const float* slope,
                                   void* out, int32_t in_dtype, int32_t out_dtype, void* stream) {
  MDX_CHECK_ARG(x && out && n > 0 && H > 0 && W > 0 && C > 0, "mdx_depth_to_space2: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const unsigned g = grid_for((long long)n * 4 * H * W * C);
  if (in_dtype == MDX_F32 && out_dtype == MDX_F32)
    depth_to_space2_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, slope, (float*)out);
  else if (in_dtype == MDX_F32 && out_dtype == MDX_BF16)
    depth_to_space2_kernel<float, bf16><<<g, 256, 0, st>>>((const float*)x, n, H, W, C, slope, (bf16*)out);
  else if (in_dtype == MDX_BF16 && out_dtype == MDX_BF16)
    depth_to_space2_kernel<bf16, bf16><<<g, 256, 0, st>>>((const bf16*)x, n, H, W, C, slope, (bf16*)out);
  else {
    mdx_set_error("mdx_depth_to_space2: unsupported dtypes %d -> %d", in_dtype, out_dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}
