// CUDA-core implicit-GEMM convolution / dense layer.  Used for the fp32 check mode (every conv / dense of the
// path) and, in bf16 mode, for the tiny timestep-embedding MLPs.  See include/mdx.h (mdx_conv_direct).
//
// 64x64 output tile per CTA, K step 16, 256 threads, 4x4 register micro-tile, im2col gather done on the fly
// (zero padding, stride, optional nearest-2x upsampled view of the input, optional SiLU on load).
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

struct ConvP {
  const void* x; long long ldx;
  int n, H, W, Cin;       // stored input
  int Hl, Wl;             // logical input (after optional upsample)
  const void* w;
  int Cout, kh, kw, stride, pad, upsample, pre_silu;
  int OH, OW;
  long long M;            // n*OH*OW
  int K;                  // kh*kw*Cin
  void* out; long long ldo;
  const float* bias; const float* rowbias; long long ld_rowbias;
  const void* residual; long long ldr;
  int act;
};

constexpr int TM = 64, TN = 64, TK = 16;

template <typename TI, typename TO>
__global__ void __launch_bounds__(256) conv_direct_kernel(const ConvP p) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const TI* __restrict__ x = reinterpret_cast<const TI*>(p.x);
  const TI* __restrict__ w = reinterpret_cast<const TI*>(p.w);

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long m0 = (long long)blockIdx.x * TM;
  const int n0 = blockIdx.y * TN;

  // this thread's gather row (fixed across the K loop)
  const int lrow = tid >> 2;        // 0..63
  const int lk = (tid & 3) * 4;     // 0,4,8,12
  const long long grow = m0 + lrow;
  const bool row_ok = grow < p.M;
  int img = 0, oy = 0, ox = 0;
  if (row_ok) {
    img = (int)(grow / ((long long)p.OH * p.OW));
    int rem = (int)(grow % ((long long)p.OH * p.OW));
    oy = rem / p.OW;
    ox = rem % p.OW;
  }
  const int bcol = n0 + lrow;
  const bool col_ok = bcol < p.Cout;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < p.K; k0 += TK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int kk = k0 + lk + i;
      float av = 0.f, bv = 0.f;
      if (kk < p.K) {
        if (row_ok) {
          const int tap = kk / p.Cin;
          const int c = kk - tap * p.Cin;
          const int r = tap / p.kw, s = tap - r * p.kw;
          int iy = oy * p.stride + r - p.pad;
          int ix = ox * p.stride + s - p.pad;
          if (iy >= 0 && iy < p.Hl && ix >= 0 && ix < p.Wl) {
            if (p.upsample) { iy >>= 1; ix >>= 1; }
            av = to_f32<TI>(x[(((long long)img * p.H + iy) * p.W + ix) * p.ldx + c]);
            if (p.pre_silu) av = silu_f(av);
          }
        }
        if (col_ok) bv = to_f32<TI>(w[(long long)bcol * p.K + kk]);
      }
      As[lk + i][lrow] = av;
      Bs[lk + i][lrow] = bv;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  TO* __restrict__ out = reinterpret_cast<TO*>(p.out);
  const TO* __restrict__ res = reinterpret_cast<const TO*>(p.residual);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long row = m0 + ty * 4 + i;
    if (row >= p.M) continue;
    const int rimg = (int)(row / ((long long)p.OH * p.OW));
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int col = n0 + tx * 4 + j;
      if (col >= p.Cout) continue;
      float v = acc[i][j];
      if (p.bias) v += p.bias[col];
      if (p.rowbias) v += p.rowbias[(long long)rimg * p.ld_rowbias + col];
      if (p.act) v = apply_act_precise(v, p.act);
      if (res) v += to_f32<TO>(res[row * p.ldr + col]);
      out[row * p.ldo + col] = from_f32<TO>(v);
    }
  }
}

}  // namespace

extern "C" int mdx_conv_direct(const mdx_conv_args* a, void* stream) {
  MDX_CHECK_ARG(a != nullptr, "mdx_conv_direct: args is NULL");
  MDX_CHECK_ARG(a->x && a->w && a->out, "mdx_conv_direct: NULL operand");
  MDX_CHECK_ARG(a->n > 0 && a->H > 0 && a->W > 0 && a->Cin > 0 && a->Cout > 0, "mdx_conv_direct: empty problem");
  MDX_CHECK_ARG(a->kh > 0 && a->kw > 0 && a->stride > 0 && a->pad >= 0, "mdx_conv_direct: bad kernel geometry");
  ConvP p;
  p.x = a->x; p.ldx = a->ldx; p.n = a->n; p.H = a->H; p.W = a->W; p.Cin = a->Cin;
  p.Hl = a->upsample ? 2 * a->H : a->H;
  p.Wl = a->upsample ? 2 * a->W : a->W;
  p.w = a->w; p.Cout = a->Cout; p.kh = a->kh; p.kw = a->kw; p.stride = a->stride; p.pad = a->pad;
  p.upsample = a->upsample; p.pre_silu = a->pre_silu;
  p.OH = (p.Hl + 2 * a->pad - a->kh) / a->stride + 1;
  p.OW = (p.Wl + 2 * a->pad - a->kw) / a->stride + 1;
  MDX_CHECK_ARG(p.OH > 0 && p.OW > 0, "mdx_conv_direct: empty output");
  p.M = (long long)a->n * p.OH * p.OW;
  p.K = a->kh * a->kw * a->Cin;
  p.out = a->out; p.ldo = a->ldo; p.bias = a->bias; p.rowbias = a->rowbias; p.ld_rowbias = a->ld_rowbias;
  p.residual = a->residual; p.ldr = a->ldr; p.act = a->act;
  long long gx = (p.M + TM - 1) / TM;
  MDX_CHECK_ARG(gx < (1ll << 31), "mdx_conv_direct: too many rows");
  dim3 grid((unsigned)gx, (unsigned)((a->Cout + TN - 1) / TN));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (a->in_dtype == MDX_F32 && a->out_dtype == MDX_F32)
    conv_direct_kernel<float, float><<<grid, 256, 0, st>>>(p);
  else if (a->in_dtype == MDX_BF16 && a->out_dtype == MDX_BF16)
    conv_direct_kernel<bf16, bf16><<<grid, 256, 0, st>>>(p);
  else if (a->in_dtype == MDX_BF16 && a->out_dtype == MDX_F32)
    conv_direct_kernel<bf16, float><<<grid, 256, 0, st>>>(p);
  else if (a->in_dtype == MDX_F32 && a->out_dtype == MDX_BF16)
    conv_direct_kernel<float, bf16><<<grid, 256, 0, st>>>(p);
  else {
    mdx_set_error("mdx_conv_direct: unsupported dtype combination in=%d out=%d", a->in_dtype, a->out_dtype);
    return MDX_ERR_INVALID;
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}
