// Multi-head attention o = softmax(q k^T * scale) v, never materialising the [Nq, Nk] score matrix.
// See include/mdx.h (mdx_attention); reference: CrossAttention.construct, ldm/modules/attention.py:117-166.
//
//  bf16: flash-style kernel. CTA = 128 query rows (8 warps x 16 rows) of one (batch, head); K/V streamed in
//        64-key tiles through a 2-stage cp.async ring; S = Q K^T and O += P V on mma.sync.m16n8k16 (bf16 in,
//        fp32 accumulate); online softmax in fp32 registers with exp2 and a pre-folded log2(e) scale.
//  fp32: check-mode kernel, one warp per query row, CUDA cores only.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>
#include <stdlib.h>

extern std::atomic<long long> g_mdx_launches;

namespace {

struct AttnP {
  const void* q; long long ldq;
  const void* k; long long ldk;
  const void* v; long long ldv;
  void* o; long long ldo;
  int batch, heads, Nq, Nk, d;
  float scale;
  int causal;
};

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(smem)));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(smem)));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int FA_BM = 128;  // queries per CTA
constexpr int FA_BN = 64;   // keys per tile

// HD = head dim padded to a multiple of 16.  smem rows are padded by 16 B so ldmatrix is conflict-free.
template <int HD>
__global__ void __launch_bounds__(256, 1) flash_attn_bf16_kernel(const AttnP p) {
  constexpr int LDS = HD + 8;  // elements per smem row
  extern __shared__ __align__(16) uint8_t fa_smem[];
  bf16* sQ = reinterpret_cast<bf16*>(fa_smem);
  bf16* sK = sQ + FA_BM * LDS;            // [2][FA_BN][LDS]
  bf16* sV = sK + 2 * FA_BN * LDS;        // [2][FA_BN][LDS]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int q0 = blockIdx.x * FA_BM;
  const int head = blockIdx.y, b = blockIdx.z;
  const int d = p.d;
  const bf16* gq = reinterpret_cast<const bf16*>(p.q) + (long long)b * p.Nq * p.ldq + head * d;
  const bf16* gk = reinterpret_cast<const bf16*>(p.k) + (long long)b * p.Nk * p.ldk + head * d;
  const bf16* gv = reinterpret_cast<const bf16*>(p.v) + (long long)b * p.Nk * p.ldv + head * d;
  bf16* go = reinterpret_cast<bf16*>(p.o) + (long long)b * p.Nq * p.ldo + head * d;

  constexpr int CH = HD / 8;  // 16-byte chunks per row
  const int dch = d / 8;      // valid chunks
  // ---- Q tile ----
  for (int i = tid; i < FA_BM * CH; i += 256) {
    const int r = i / CH, c = i % CH;
    const bool ok = (q0 + r < p.Nq) && (c < dch);
    cp_async16(sQ + r * LDS + c * 8, ok ? gq + (long long)(q0 + r) * p.ldq + c * 8 : gq, ok);
  }
  auto load_kv = [&](int tile, int stage) {
    const int k0 = tile * FA_BN;
    bf16* dk = sK + stage * FA_BN * LDS;
    bf16* dv = sV + stage * FA_BN * LDS;
    for (int i = tid; i < FA_BN * CH; i += 256) {
      const int r = i / CH, c = i % CH;
      const bool ok = (k0 + r < p.Nk) && (c < dch);
      cp_async16(dk + r * LDS + c * 8, ok ? gk + (long long)(k0 + r) * p.ldk + c * 8 : gk, ok);
      cp_async16(dv + r * LDS + c * 8, ok ? gv + (long long)(k0 + r) * p.ldv + c * 8 : gv, ok);
    }
  };
  const int ntiles = (p.Nk + FA_BN - 1) / FA_BN;
  load_kv(0, 0);
  cp_async_commit();

  float o_acc[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) { o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f; }
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  const float sl2 = p.scale * 1.4426950408889634f;  // scale * log2(e)

  uint32_t qf[HD / 16][4];
  bool q_loaded = false;

  for (int t = 0; t < ntiles; ++t) {
    if (t + 1 < ntiles) load_kv(t + 1, (t + 1) & 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (!q_loaded) {
      q_loaded = true;
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk)
        ldmatrix_x4(qf[kk], sQ + (warp * 16 + (lane & 15)) * LDS + kk * 16 + (lane >> 4) * 8);
    }
    const bf16* tK = sK + (t & 1) * FA_BN * LDS;
    const bf16* tV = sV + (t & 1) * FA_BN * LDS;

    // ---- S = Q K^T (16 x 64 per warp) ----
    float s[FA_BN / 8][4];
#pragma unroll
    for (int j = 0; j < FA_BN / 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
#pragma unroll
      for (int jp = 0; jp < FA_BN / 16; ++jp) {
        uint32_t kf[4];
        ldmatrix_x4(kf, tK + (jp * 16 + (lane & 7) + ((lane >> 4) & 1) * 8) * LDS + kk * 16 + ((lane >> 3) & 1) * 8);
        mma_bf16_16816(s[2 * jp], qf[kk], kf[0], kf[1]);
        mma_bf16_16816(s[2 * jp + 1], qf[kk], kf[2], kf[3]);
      }
    }
    // ---- mask keys beyond Nk, online softmax ----
    const int kbase = t * FA_BN;
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int j = 0; j < FA_BN / 8; ++j) {
      const int key = kbase + j * 8 + 2 * t4;
      if (key >= p.Nk) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
      if (key + 1 >= p.Nk) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
      if (p.causal) {  // key j visible to query i only if j <= i
        const int qr0 = q0 + warp * 16 + g, qr1 = qr0 + 8;
        if (key > qr0) s[j][0] = -INFINITY;
        if (key + 1 > qr0) s[j][1] = -INFINITY;
        if (key > qr1) s[j][2] = -INFINITY;
        if (key + 1 > qr1) s[j][3] = -INFINITY;
      }
      mx[0] = fmaxf(mx[0], fmaxf(s[j][0], s[j][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[j][2], s[j][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    float corr[2], msc[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const float m_new = fmaxf(m_run[r], mx[r]);
      corr[r] = exp2f((m_run[r] - m_new) * sl2);  // m_run = -inf on first tile -> 0
      m_run[r] = m_new;
      msc[r] = m_new * sl2;
      l_run[r] *= corr[r];
    }
    float rs[2] = {0.f, 0.f};
    uint32_t pf[FA_BN / 16][4];
#pragma unroll
    for (int j = 0; j < FA_BN / 8; ++j) {
      const float p0 = exp2f(fmaf(s[j][0], sl2, -msc[0]));
      const float p1 = exp2f(fmaf(s[j][1], sl2, -msc[0]));
      const float p2 = exp2f(fmaf(s[j][2], sl2, -msc[1]));
      const float p3 = exp2f(fmaf(s[j][3], sl2, -msc[1]));
      rs[0] += p0 + p1;
      rs[1] += p2 + p3;
      pf[j >> 1][(j & 1) * 2 + 0] = pack_bf16x2(p0, p1);
      pf[j >> 1][(j & 1) * 2 + 1] = pack_bf16x2(p2, p3);
    }
    l_run[0] += rs[0];
    l_run[1] += rs[1];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
      o_acc[i][0] *= corr[0]; o_acc[i][1] *= corr[0];
      o_acc[i][2] *= corr[1]; o_acc[i][3] *= corr[1];
    }
    // ---- O += P V ----
#pragma unroll
    for (int kk = 0; kk < FA_BN / 16; ++kk) {
#pragma unroll
      for (int dp = 0; dp < HD / 16; ++dp) {
        uint32_t vf[4];
        ldmatrix_x4_trans(vf, tV + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + dp * 16 + ((lane >> 4) & 1) * 8);
        mma_bf16_16816(o_acc[2 * dp], pf[kk], vf[0], vf[1]);
        mma_bf16_16816(o_acc[2 * dp + 1], pf[kk], vf[2], vf[3]);
      }
    }
    __syncthreads();  // everyone done with this stage before it is refilled
  }
  cp_async_wait<0>();

  // ---- finalise: row sums across the quad, normalise, store ----
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const float inv0 = 1.f / l_run[0], inv1 = 1.f / l_run[1];
  const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) {
    const int col = i * 8 + 2 * t4;
    if (col < d) {
      if (row0 < p.Nq)
        *reinterpret_cast<uint32_t*>(go + (long long)row0 * p.ldo + col) = pack_bf16x2(o_acc[i][0] * inv0, o_acc[i][1] * inv0);
      if (row1 < p.Nq)
        *reinterpret_cast<uint32_t*>(go + (long long)row1 * p.ldo + col) = pack_bf16x2(o_acc[i][2] * inv1, o_acc[i][3] * inv1);
    }
  }
}

// ---------------- fp32 check-mode attention: one warp per query row ----------------
__global__ void __launch_bounds__(128) attn_f32_kernel(const AttnP p) {
  __shared__ float sq[4][160];
  __shared__ float sp[4][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qi = blockIdx.x * 4 + warp;
  const int head = blockIdx.y, b = blockIdx.z;
  if (qi >= p.Nq) return;
  const int d = p.d;
  const float* gq = reinterpret_cast<const float*>(p.q) + ((long long)b * p.Nq + qi) * p.ldq + head * d;
  const float* gk = reinterpret_cast<const float*>(p.k) + (long long)b * p.Nk * p.ldk + head * d;
  const float* gv = reinterpret_cast<const float*>(p.v) + (long long)b * p.Nk * p.ldv + head * d;
  float* go = reinterpret_cast<float*>(p.o) + ((long long)b * p.Nq + qi) * p.ldo + head * d;
  for (int i = lane; i < d; i += 32) sq[warp][i] = gq[i] * p.scale;
  __syncwarp();
  float m = -INFINITY, l = 0.f;
  float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};  // d <= 160
  for (int k0 = 0; k0 < p.Nk; k0 += 32) {
    const int key = k0 + lane;
    float s = -INFINITY;
    if (key < p.Nk && !(p.causal && key > qi)) {
      const float* kr = gk + (long long)key * p.ldk;
      float dot = 0.f;
      for (int i = 0; i < d; ++i) dot = fmaf(sq[warp][i], kr[i], dot);
      s = dot;
    }
    const float m_new = fmaxf(m, warp_max(s));
    const float corr = __expf(m - m_new);
    const float pe = (key < p.Nk && !(p.causal && key > qi)) ? __expf(s - m_new) : 0.f;
    l = l * corr + warp_sum(pe);
    m = m_new;
    sp[warp][lane] = pe;
    __syncwarp();
    const int kn = min(32, p.Nk - k0);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int di = lane + 32 * j;
      if (di < d) {
        float a = acc[j] * corr;
        for (int kk = 0; kk < kn; ++kk) a = fmaf(sp[warp][kk], gv[(long long)(k0 + kk) * p.ldv + di], a);
        acc[j] = a;
      }
    }
    __syncwarp();
  }
  const float inv = 1.f / l;
#pragma unroll
  for (int j = 0; j < 5; ++j) {
    const int di = lane + 32 * j;
    if (di < d) go[di] = acc[j] * inv;
  }
}

template <int HD>
int launch_fa(const AttnP& p, cudaStream_t st) {
  constexpr int LDS = HD + 8;
  constexpr int SMEM = (FA_BM + 4 * FA_BN) * LDS * 2;
  static MdxPerDeviceOnce attr;
  if (attr.need()) {
    MDX_CHECK_CUDA(cudaFuncSetAttribute(flash_attn_bf16_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr.mark();
  }
  dim3 grid((p.Nq + FA_BM - 1) / FA_BM, p.heads, p.batch);
  flash_attn_bf16_kernel<HD><<<grid, 256, SMEM, st>>>(p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

}  // namespace

int mdx_attention_tc_d64(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                         void* o, long long ldo, int batch, int heads, int Nq, int Nk, int d, float scale,
                         cudaStream_t st);

extern "C" int mdx_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                             void* o, int64_t ldo, int32_t batch, int32_t heads, int32_t Nq, int32_t Nk, int32_t d,
                             float scale, int32_t dtype, int32_t causal, void* stream) {
  MDX_CHECK_ARG(q && k && v && o, "mdx_attention: NULL operand");
  MDX_CHECK_ARG(batch > 0 && heads > 0 && Nq > 0 && Nk > 0, "mdx_attention: empty problem");
  MDX_CHECK_ARG(d > 0 && d <= 160 && d % 8 == 0, "mdx_attention: head dim %d unsupported (need d %% 8 == 0, d <= 160)", d);
  MDX_CHECK_ARG(heads <= 65535 && batch <= 65535, "mdx_attention: too many heads/batches");
  AttnP p{q, ldq, k, ldk, v, ldv, o, ldo, batch, heads, Nq, Nk, d, scale, causal};
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == MDX_F32) {
    dim3 grid((Nq + 3) / 4, heads, batch);
    attn_f32_kernel<<<grid, 128, 0, st>>>(p);
    g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
    MDX_CHECK_LAUNCH();
    return MDX_OK;
  }
  MDX_CHECK_ARG(dtype == MDX_BF16, "mdx_attention: bad dtype");
  MDX_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 2 == 0, "mdx_attention: row strides must be multiples of 8");
  MDX_CHECK_ARG(((uintptr_t)q & 15) == 0 && ((uintptr_t)k & 15) == 0 && ((uintptr_t)v & 15) == 0 && ((uintptr_t)o & 3) == 0,
                "mdx_attention: operands must be 16-byte aligned");
  // head dim <= 64 (every attention layer of the SD2 UNet is 64; wukong-huahua's 64x64 level is 40, zero-padded to
  // 64 by TMA): tcgen05/TMEM kernels; MDX_ATTN_LEGACY=1 keeps mma.sync
  const bool legacy = getenv("MDX_ATTN_LEGACY") != nullptr;
  if (d <= 64 && !legacy && !causal && ldo % 8 == 0 && ((uintptr_t)o & 15) == 0)
    return mdx_attention_tc_d64(q, ldq, k, ldk, v, ldv, o, ldo, batch, heads, Nq, Nk, d, scale, st);
  if (d <= 48) return launch_fa<48>(p, st);
  if (d <= 64) return launch_fa<64>(p, st);
  if (d <= 80) return launch_fa<80>(p, st);
  if (d <= 96) return launch_fa<96>(p, st);
  if (d <= 128) return launch_fa<128>(p, st);
  return launch_fa<160>(p, st);
}
