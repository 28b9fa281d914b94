// Sampler arithmetic: classifier-free-guidance combine + latent update, ONE kernel per step.
// See include/mdx.h (mdx_sampler_plms_step, mdx_sampler_dpmpp2m_step) for formulas and reference lines.
// The UNet output arrives NHWC in the model dtype; sampler state (x, eps history, x0 history) is fp32 NCHW,
// the layout the reference's sampler API returns.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

struct PlmsP {
  const void* eps_model; long long ld_eps;
  int cfg; float scale;
  const float* x; const float* old1; const float* old2; const float* old3;
  float w0, w1, w2, w3;
  float sqrt_at, sqrt_aprev, dir_coef, sigma, sqrt_1m_at;
  const float* noise;
  float* x_prev; float* pred_x0; float* e_t_out;
  int B, C, HW;
};

template <typename T>
__device__ __forceinline__ float model_eps(const void* m, int cfg, float scale, int b, int c, int hw, int B,
                                           long long ld, int HW) {
  const T* e = reinterpret_cast<const T*>(m);
  const long long iu = ((long long)b * HW + hw) * ld + c;
  if (!cfg) return to_f32<T>(e[iu]);
  const float eu = to_f32<T>(e[iu]);
  const float ec = to_f32<T>(e[iu + (long long)B * HW * ld]);
  return eu + scale * (ec - eu);
}

template <typename T>
__global__ void plms_step_kernel(const PlmsP p) {
  const long long total = (long long)p.B * p.C * p.HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int hw = (int)(i % p.HW);
    const int c = (int)((i / p.HW) % p.C);
    const int b = (int)(i / ((long long)p.HW * p.C));
    const float e_t = model_eps<T>(p.eps_model, p.cfg, p.scale, b, c, hw, p.B, p.ld_eps, p.HW);
    float e = p.w0 * e_t;
    if (p.old1) e += p.w1 * p.old1[i];
    if (p.old2) e += p.w2 * p.old2[i];
    if (p.old3) e += p.w3 * p.old3[i];
    const float xv = p.x[i];
    const float x0 = (xv - p.sqrt_1m_at * e) / p.sqrt_at;
    float xp = p.sqrt_aprev * x0 + p.dir_coef * e;
    if (p.noise) xp += p.sigma * p.noise[i];
    if (p.e_t_out) p.e_t_out[i] = e_t;
    if (p.pred_x0) p.pred_x0[i] = x0;
    p.x_prev[i] = xp;
  }
}

struct DpmP {
  const void* eps_model; long long ld_eps;
  int cfg; float scale; int model_type;
  const float* x; const float* x0_prev;
  float alpha_s, sigma_s, cx, c0, c1;
  float* x_next; float* x0_cur;
  int B, C, HW;
};

template <typename T>
__global__ void dpm_step_kernel(const DpmP p) {
  const long long total = (long long)p.B * p.C * p.HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int hw = (int)(i % p.HW);
    const int c = (int)((i / p.HW) % p.C);
    const int b = (int)(i / ((long long)p.HW * p.C));
    const float out = model_eps<T>(p.eps_model, p.cfg, p.scale, b, c, hw, p.B, p.ld_eps, p.HW);
    const float xv = p.x[i];
    const float noise = p.model_type == 1 ? p.alpha_s * out + p.sigma_s * xv : out;
    const float x0 = (xv - p.sigma_s * noise) / p.alpha_s;
    float xn = p.cx * xv + p.c0 * x0;
    if (p.x0_prev) xn += p.c1 * (x0 - p.x0_prev[i]);
    p.x0_cur[i] = x0;
    p.x_next[i] = xn;
  }
}

inline unsigned grid_for(long long n) {
  long long b = (n + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  return (unsigned)(b < 1 ? 1 : b);
}

}  // namespace

extern "C" int mdx_sampler_plms_step(const void* eps_model, int64_t ld_eps, int32_t model_dtype, int32_t cfg, float scale,
                                     const float* x, const float* old1, const float* old2, const float* old3,
                                     const float* w_host, const float* coef_host, const float* noise, float* x_prev,
                                     float* pred_x0, float* e_t_out, int32_t B, int32_t C, int32_t h, int32_t w,
                                     void* stream) {
  MDX_CHECK_ARG(eps_model && x && w_host && coef_host && x_prev, "mdx_sampler_plms_step: NULL operand");
  MDX_CHECK_ARG(B > 0 && C > 0 && h > 0 && w > 0 && ld_eps >= C, "mdx_sampler_plms_step: empty problem / ld_eps < C");
  MDX_CHECK_ARG(model_dtype == MDX_F32 || model_dtype == MDX_BF16, "mdx_sampler_plms_step: bad dtype");
  PlmsP p;
  p.eps_model = eps_model; p.ld_eps = ld_eps; p.cfg = cfg; p.scale = scale;
  p.x = x; p.old1 = old1; p.old2 = old2; p.old3 = old3;
  p.w0 = w_host[0]; p.w1 = w_host[1]; p.w2 = w_host[2]; p.w3 = w_host[3];
  const float a_t = coef_host[0], a_prev = coef_host[1], sigma = coef_host[2], s1m = coef_host[3];
  MDX_CHECK_ARG(a_t > 0.f && a_prev > 0.f, "mdx_sampler_plms_step: alphas must be positive");
  p.sqrt_at = sqrtf(a_t);
  p.sqrt_aprev = sqrtf(a_prev);
  const float dir2 = 1.f - a_prev - sigma * sigma;
  p.dir_coef = sqrtf(dir2 > 0.f ? dir2 : 0.f);
  p.sigma = sigma; p.sqrt_1m_at = s1m;
  p.noise = (sigma != 0.f) ? noise : nullptr;
  p.x_prev = x_prev; p.pred_x0 = pred_x0; p.e_t_out = e_t_out;
  p.B = B; p.C = C; p.HW = h * w;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)B * C * h * w;
  if (model_dtype == MDX_BF16) plms_step_kernel<bf16><<<grid_for(total), 256, 0, st>>>(p);
  else plms_step_kernel<float><<<grid_for(total), 256, 0, st>>>(p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

extern "C" int mdx_sampler_dpmpp2m_step(const void* eps_model, int64_t ld_eps, int32_t model_dtype, int32_t cfg, float scale,
                                        int32_t model_type, const float* x, const float* x0_prev,
                                        const float* coef_host, float* x_next, float* x0_cur, int32_t B, int32_t C,
                                        int32_t h, int32_t w, void* stream) {
  MDX_CHECK_ARG(eps_model && x && coef_host && x_next && x0_cur, "mdx_sampler_dpmpp2m_step: NULL operand");
  MDX_CHECK_ARG(B > 0 && C > 0 && h > 0 && w > 0 && ld_eps >= C, "mdx_sampler_dpmpp2m_step: empty problem / ld_eps < C");
  MDX_CHECK_ARG(model_dtype == MDX_F32 || model_dtype == MDX_BF16, "mdx_sampler_dpmpp2m_step: bad dtype");
  MDX_CHECK_ARG(model_type == 0 || model_type == 1, "mdx_sampler_dpmpp2m_step: model_type must be 0 (noise) or 1 (v)");
  DpmP p;
  p.eps_model = eps_model; p.ld_eps = ld_eps; p.cfg = cfg; p.scale = scale; p.model_type = model_type;
  p.x = x; p.x0_prev = x0_prev;
  p.alpha_s = coef_host[0]; p.sigma_s = coef_host[1]; p.cx = coef_host[2]; p.c0 = coef_host[3]; p.c1 = coef_host[4];
  MDX_CHECK_ARG(p.alpha_s != 0.f, "mdx_sampler_dpmpp2m_step: alpha_s must be non-zero");
  p.x_next = x_next; p.x0_cur = x0_cur; p.B = B; p.C = C; p.HW = h * w;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)B * C * h * w;
  if (model_dtype == MDX_BF16) dpm_step_kernel<bf16><<<grid_for(total), 256, 0, st>>>(p);
  else dpm_step_kernel<float><<<grid_for(total), 256, 0, st>>>(p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}
