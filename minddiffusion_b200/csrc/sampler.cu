// Sampler arithmetic: classifier-free-guidance combine + latent update, ONE kernel per step.
// See include/mdx.h (mdx_sampler_plms_step, mdx_sampler_dpmpp2m_step and their *_ex forms) for formulas and
// reference lines.  The UNet output arrives NHWC in the model dtype; sampler state (x, eps history, x0 history)
// is fp32 NCHW, the layout the reference's sampler API returns.
//
// Work split: one thread per (image, pixel); it walks the C channels.  The NHWC model output of a pixel is one
// contiguous run (C <= 16 values), every NCHW plane access is coalesced across the warp.
//
// Per-step scalars come either from the launch parameters (host form) or from a row of a device table selected by a
// device step counter (table form): the table form makes a step's launch sequence identical from step to step, so a
// sampler loop can capture it once in a CUDA graph and replay it.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>
#include <stddef.h>

extern std::atomic<long long> g_mdx_launches;

namespace {

constexpr int NC = MDX_STEP_COEFS;

struct StepP {
  const void* model_out; long long ld;
  int cfg; float scale; int model_type;
  const float* x; const float* x_model;
  float* h1; float* h2; float* h3;
  const float* noise;
  float* x_out; float* pred_x0; float* e_t_out;
  int shift_hist;
  int B, C, HW;
  float hc[NC];              // host form: this step's row
  const float* table;        // table form: [rows][NC]
  const int* step;           // table form: row index
};

__device__ __forceinline__ void load_row(const StepP& p, float (&c)[NC]) {
  if (p.table) {
    const float* r = p.table + (long long)(*p.step) * NC;
#pragma unroll
    for (int j = 0; j < NC; ++j) c[j] = __ldg(r + j);
  } else {
#pragma unroll
    for (int j = 0; j < NC; ++j) c[j] = p.hc[j];
  }
}

template <typename T>
__device__ __forceinline__ float combined(const T* m, long long iu, long long half, int cfg, float scale) {
  const float eu = to_f32<T>(m[iu]);
  if (!cfg) return eu;
  const float ec = to_f32<T>(m[iu + half]);
  return eu + scale * (ec - eu);
}

// row: {w0, w1, w2, w3, sqrt_at, sqrt_aprev, dir_coef, sigma, sqrt_1m_at, v_a, v_s, use_noise}
template <typename T>
__global__ void plms_step_kernel(const StepP p) {
  float c[NC];
  load_row(p, c);
  const long long npix = (long long)p.B * p.HW;
  const long long half = (long long)p.B * p.HW * p.ld;
  const T* m = reinterpret_cast<const T*>(p.model_out);
  const float* xm = p.x_model ? p.x_model : p.x;
  const bool noisy = p.noise != nullptr && c[11] != 0.f;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < npix; q += (long long)gridDim.x * blockDim.x) {
    const long long b = q / p.HW;
    const int hw = (int)(q - b * p.HW);
    const long long mrow = q * p.ld;
    for (int ch = 0; ch < p.C; ++ch) {
      const long long i = (b * p.C + ch) * p.HW + hw;
      float e_t = combined<T>(m, mrow + ch, half, p.cfg, p.scale);
      if (p.model_type == 1) e_t = c[9] * e_t + c[10] * xm[i];   // v -> eps at the evaluated (x, t)
      float e = c[0] * e_t;
      float o1 = 0.f, o2 = 0.f;
      if (p.h1) { o1 = p.h1[i]; e += c[1] * o1; }
      if (p.h2) { o2 = p.h2[i]; e += c[2] * o2; }
      if (p.h3) e += c[3] * p.h3[i];
      const float xv = p.x[i];
      const float x0 = (xv - c[8] * e) / c[4];
      float xp = c[5] * x0 + c[6] * e;
      if (noisy) xp += c[7] * p.noise[i];
      if (p.shift_hist) {
        p.h3[i] = o2;
        p.h2[i] = o1;
        p.h1[i] = e_t;
      } else if (p.e_t_out) {
        p.e_t_out[i] = e_t;
      }
      if (p.pred_x0) p.pred_x0[i] = x0;
      p.x_out[i] = xp;
    }
  }
}

// row: {alpha_s, sigma_s, c_x, c_0, c_1, use_prev}
template <typename T>
__global__ void dpm_step_kernel(const StepP p) {
  float c[NC];
  load_row(p, c);
  const long long npix = (long long)p.B * p.HW;
  const long long half = (long long)p.B * p.HW * p.ld;
  const T* m = reinterpret_cast<const T*>(p.model_out);
  const bool prev = p.h1 != nullptr && c[5] != 0.f;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < npix; q += (long long)gridDim.x * blockDim.x) {
    const long long b = q / p.HW;
    const int hw = (int)(q - b * p.HW);
    const long long mrow = q * p.ld;
    for (int ch = 0; ch < p.C; ++ch) {
      const long long i = (b * p.C + ch) * p.HW + hw;
      const float out = combined<T>(m, mrow + ch, half, p.cfg, p.scale);
      const float xv = p.x[i];
      const float noise = p.model_type == 1 ? c[0] * out + c[1] * xv : out;
      const float x0 = (xv - c[1] * noise) / c[0];
      float xn = c[2] * xv + c[3] * x0;
      if (prev) xn += c[4] * (x0 - p.h1[i]);
      p.pred_x0[i] = x0;       // x0_cur; may be the buffer h1 points at (read above, same thread)
      p.x_out[i] = xn;
    }
  }
}

// img = (c0 * x0 + c1 * noise) * mask + (1 - mask) * img          (plms.py:153-156 with q_sample, ddpm.py:180-182)
__global__ void mask_blend_kernel(const float* x0, const float* noise, const float* mask, int mask_c, float c0, float c1,
                                  float* img, int B, int C, int HW) {
  const long long total = (long long)B * C * HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int hw = (int)(i % HW);
    const int ch = (int)((i / HW) % C);
    const long long b = i / ((long long)HW * C);
    const float mk = mask[(b * mask_c + (mask_c == 1 ? 0 : ch)) * HW + hw];
    const float orig = c0 * x0[i] + c1 * noise[i];
    img[i] = orig * mk + (1.f - mk) * img[i];
  }
}

__global__ void step_advance_kernel(int* step, int delta) { *step += delta; }

// out[j][:] = table[*idx + offset][:] for j < n_out
__global__ void gather_row_kernel(const float* table, long long ld, const int* idx, int offset, float* out, int n_out,
                                  int width) {
  const float* r = table + (long long)(*idx + offset) * ld;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_out * width; i += gridDim.x * blockDim.x)
    out[i] = r[i % width];
}

inline unsigned grid_for(long long n) {
  long long b = (n + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  return (unsigned)(b < 1 ? 1 : b);
}

int fill(StepP& p, const mdx_step_args* a, const char* who) {
  MDX_CHECK_ARG(a != nullptr, "%s: args is NULL", who);
  MDX_CHECK_ARG(a->model_out && a->x && a->x_out, "%s: NULL operand", who);
  MDX_CHECK_ARG(a->B > 0 && a->C > 0 && a->h > 0 && a->w > 0 && a->ld_out >= a->C, "%s: empty problem / ld_out < C", who);
  MDX_CHECK_ARG(a->model_dtype == MDX_F32 || a->model_dtype == MDX_BF16, "%s: bad dtype", who);
  MDX_CHECK_ARG(a->model_type == 0 || a->model_type == 1, "%s: model_type must be 0 (eps) or 1 (v)", who);
  MDX_CHECK_ARG((a->coef_host != nullptr) != (a->coef_table != nullptr && a->step_dev != nullptr),
                "%s: pass either coef_host or (coef_table, step_dev)", who);
  p.model_out = a->model_out; p.ld = a->ld_out; p.cfg = a->cfg; p.scale = a->scale; p.model_type = a->model_type;
  p.x = a->x; p.x_model = a->x_model;
  p.h1 = a->hist[0]; p.h2 = a->hist[1]; p.h3 = a->hist[2];
  p.noise = a->noise; p.x_out = a->x_out; p.pred_x0 = a->pred_x0; p.e_t_out = a->e_t_out;
  p.shift_hist = a->shift_hist;
  p.B = a->B; p.C = a->C; p.HW = a->h * a->w;
  p.table = nullptr; p.step = nullptr;
  for (int j = 0; j < NC; ++j) p.hc[j] = 0.f;
  if (a->coef_host) {
    for (int j = 0; j < NC; ++j) p.hc[j] = a->coef_host[j];
  } else {
    p.table = a->coef_table;
    p.step = a->step_dev;
  }
  return MDX_OK;
}

}  // namespace

extern "C" int mdx_sampler_plms_coef_row(const float* w4, const float* coef4, const float* v2, int32_t use_noise,
                                         float* row) {
  MDX_CHECK_ARG(w4 && coef4 && row, "mdx_sampler_plms_coef_row: NULL argument");
  const float a_t = coef4[0], a_prev = coef4[1], sigma = coef4[2], s1m = coef4[3];
  MDX_CHECK_ARG(a_t > 0.f && a_prev > 0.f, "mdx_sampler_plms_coef_row: alphas must be positive");
  for (int j = 0; j < NC; ++j) row[j] = 0.f;
  row[0] = w4[0]; row[1] = w4[1]; row[2] = w4[2]; row[3] = w4[3];
  row[4] = sqrtf(a_t);
  row[5] = sqrtf(a_prev);
  const float dir2 = 1.f - a_prev - sigma * sigma;
  row[6] = sqrtf(dir2 > 0.f ? dir2 : 0.f);
  row[7] = sigma;
  row[8] = s1m;
  row[9] = v2 ? v2[0] : 1.f;
  row[10] = v2 ? v2[1] : 0.f;
  row[11] = (use_noise && sigma != 0.f) ? 1.f : 0.f;
  return MDX_OK;
}

extern "C" int mdx_sampler_plms_step_ex(const mdx_step_args* a, void* stream) {
  StepP p;
  int rc = fill(p, a, "mdx_sampler_plms_step_ex");
  if (rc) return rc;
  if (a->shift_hist)
    MDX_CHECK_ARG(p.h1 && p.h2 && p.h3, "mdx_sampler_plms_step_ex: shift_hist needs all three history buffers");
  if (a->coef_host) {
    MDX_CHECK_ARG(a->coef_host[4] > 0.f, "mdx_sampler_plms_step_ex: sqrt(a_t) must be positive");
    MDX_CHECK_ARG(a->coef_host[11] == 0.f || a->noise, "mdx_sampler_plms_step_ex: this step draws noise but noise is NULL");
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long npix = (long long)a->B * a->h * a->w;
  if (a->model_dtype == MDX_BF16) plms_step_kernel<bf16><<<grid_for(npix), 256, 0, st>>>(p);
  else plms_step_kernel<float><<<grid_for(npix), 256, 0, st>>>(p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_sampler_dpmpp2m_step_ex(const mdx_step_args* a, void* stream) {
  StepP p;
  int rc = fill(p, a, "mdx_sampler_dpmpp2m_step_ex");
  if (rc) return rc;
  MDX_CHECK_ARG(a->pred_x0 != nullptr, "mdx_sampler_dpmpp2m_step_ex: pred_x0 (x0_cur) is required");
  if (a->coef_host) MDX_CHECK_ARG(a->coef_host[0] != 0.f, "mdx_sampler_dpmpp2m_step_ex: alpha_s must be non-zero");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long npix = (long long)a->B * a->h * a->w;
  if (a->model_dtype == MDX_BF16) dpm_step_kernel<bf16><<<grid_for(npix), 256, 0, st>>>(p);
  else dpm_step_kernel<float><<<grid_for(npix), 256, 0, st>>>(p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_sampler_plms_step(const void* eps_model, int64_t ld_eps, int32_t model_dtype, int32_t cfg, float scale,
                                     const float* x, const float* old1, const float* old2, const float* old3,
                                     const float* w_host, const float* coef_host, const float* noise, float* x_prev,
                                     float* pred_x0, float* e_t_out, int32_t B, int32_t C, int32_t h, int32_t w,
                                     void* stream) {
  MDX_CHECK_ARG(eps_model && x && w_host && coef_host && x_prev, "mdx_sampler_plms_step: NULL operand");
  float row[NC];
  int rc = mdx_sampler_plms_coef_row(w_host, coef_host, nullptr, noise != nullptr, row);
  if (rc) return rc;
  mdx_step_args a;
  memset(&a, 0, sizeof(a));
  a.model_out = eps_model; a.ld_out = ld_eps; a.model_dtype = model_dtype; a.cfg = cfg; a.scale = scale;
  a.x = x;
  a.hist[0] = const_cast<float*>(old1); a.hist[1] = const_cast<float*>(old2); a.hist[2] = const_cast<float*>(old3);
  a.noise = noise; a.x_out = x_prev; a.pred_x0 = pred_x0; a.e_t_out = e_t_out;
  a.B = B; a.C = C; a.h = h; a.w = w;
  a.coef_host = row;
  return mdx_sampler_plms_step_ex(&a, stream);
}

extern "C" int mdx_sampler_dpmpp2m_step(const void* eps_model, int64_t ld_eps, int32_t model_dtype, int32_t cfg, float scale,
                                        int32_t model_type, const float* x, const float* x0_prev,
                                        const float* coef_host, float* x_next, float* x0_cur, int32_t B, int32_t C,
                                        int32_t h, int32_t w, void* stream) {
  MDX_CHECK_ARG(eps_model && x && coef_host && x_next && x0_cur, "mdx_sampler_dpmpp2m_step: NULL operand");
  float row[NC];
  for (int j = 0; j < NC; ++j) row[j] = j < 5 ? coef_host[j] : 0.f;
  row[5] = x0_prev ? 1.f : 0.f;
  mdx_step_args a;
  memset(&a, 0, sizeof(a));
  a.model_out = eps_model; a.ld_out = ld_eps; a.model_dtype = model_dtype; a.cfg = cfg; a.scale = scale;
  a.model_type = model_type;
  a.x = x; a.hist[0] = const_cast<float*>(x0_prev);
  a.x_out = x_next; a.pred_x0 = x0_cur;
  a.B = B; a.C = C; a.h = h; a.w = w;
  a.coef_host = row;
  return mdx_sampler_dpmpp2m_step_ex(&a, stream);
}

extern "C" int mdx_sampler_mask_blend(const float* x0, const float* noise, const float* mask, int32_t mask_channels,
                                      float sqrt_ac, float sqrt_1m_ac, float* img, int32_t B, int32_t C, int32_t h,
                                      int32_t w, void* stream) {
  MDX_CHECK_ARG(x0 && noise && mask && img, "mdx_sampler_mask_blend: NULL operand");
  MDX_CHECK_ARG(B > 0 && C > 0 && h > 0 && w > 0, "mdx_sampler_mask_blend: empty problem");
  MDX_CHECK_ARG(mask_channels == 1 || mask_channels == C, "mdx_sampler_mask_blend: mask must have 1 or C channels");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  mask_blend_kernel<<<grid_for((long long)B * C * h * w), 256, 0, st>>>(x0, noise, mask, mask_channels, sqrt_ac,
                                                                       sqrt_1m_ac, img, B, C, h * w);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_step_advance(int32_t* step_dev, int32_t delta, void* stream) {
  MDX_CHECK_ARG(step_dev != nullptr, "mdx_step_advance: NULL counter");
  step_advance_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(step_dev, delta);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_gather_row(const float* table, int64_t ld, const int32_t* idx_dev, int32_t offset, float* out,
                              int32_t n_out, int32_t width, void* stream) {
  MDX_CHECK_ARG(table && idx_dev && out && n_out > 0 && width > 0 && ld >= width, "mdx_gather_row: bad arguments");
  const int total = n_out * width;
  gather_row_kernel<<<(total + 255) / 256 > 64 ? 64 : (total + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      table, ld, idx_dev, offset, out, n_out, width);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int32_t mdx_sizeof_gemm_args(void) { return (int32_t)sizeof(mdx_gemm_args); }
extern "C" int32_t mdx_sizeof_conv_args(void) { return (int32_t)sizeof(mdx_conv_args); }
extern "C" int32_t mdx_sizeof_step_args(void) { return (int32_t)sizeof(mdx_step_args); }
extern "C" int32_t mdx_offsetof_gemm_args_chan_stats(void) { return (int32_t)offsetof(mdx_gemm_args, chan_stats); }
