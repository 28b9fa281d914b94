// Tensor-core implicit GEMM for sm_100a: TMA (4-D NHWC boxes, zero-filled halos) -> 128B-swizzled smem ring
// -> tcgen05.mma (cta_group::1, M=128, N=BN, bf16 x bf16 -> fp32 in TMEM, double-buffered accumulator)
// -> tcgen05.ld epilogue (bias / per-image emb bias / GEGLU / residual) -> bf16 or fp32 NHWC rows.
//
// One kernel serves every conv3x3 (stride 1), conv1x1 and nn.Dense of the denoising path; see include/mdx.h
// (mdx_gemm_tc) for the reference call sites it replaces.  A convolution is never lowered to an im2col
// buffer: for tap (r,s) the producer asks TMA for the activation box shifted by (r-1, s-1); out-of-image
// coordinates are filled with zeros by the TMA unit, which is exactly pad_mode='pad', padding=1.
//
// Warp roles (256 threads, 1 CTA / SM, persistent over output tiles):
//   warp 0  : TMA producer (one elected thread)
//   warp 1  : MMA issuer  (one elected thread)
//   warp 2  : TMEM allocate / free
//   warps 4-7: epilogue, thread i of warp w owns accumulator row 32*(w-4)+i (TMEM lane)
#include "common.cuh"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB

struct TcParams {
  int nseg;
  int seg_kb[3];    // 64-wide channel blocks per tap
  int seg_taps[3];  // 1 or 9
  int total_kb;
  int n_img, H, W;
  int tn, th, tw;
  int tiles_x, tiles_y;
  int N, n_tiles_n, num_tiles;
  void* out;
  long long ldc;
  int out_f32;
  const float* bias;
  const float* rowbias;
  long long ld_rowbias;
  const void* residual;
  long long ldr;
  int res_f32;
  int geglu;
  int act;
};

template <int BN> struct TcCfg {
  static constexpr int B_STAGE_BYTES = BN * BK * 2;
  static constexpr int STAGES = (BN >= 256) ? 4 : (BN >= 160 ? 5 : (BN >= 128 ? 5 : 7));
  static constexpr int TMEM_COLS = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
  static constexpr int EPI_WARPS = (BN >= 256) ? 4 : 8;      // 4 TMEM lane quadrants x {1,2} column parities
  static constexpr int THREADS = 128 + 32 * EPI_WARPS;
  static constexpr int STG_LD = 36;                           // floats per staging row (32 + 4 pad, 16-B aligned)
  static constexpr int STAGING_OFF = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + 256;
  static constexpr int SMEM_BYTES = STAGING_OFF + EPI_WARPS * 32 * STG_LD * 4 + 1024 /*align*/;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
};

template <int BN>
__global__ void __launch_bounds__(TcCfg<BN>::THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB,
               const TcParams p) {
  using Cfg = TcCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * Cfg::B_STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA0);
    tma_prefetch_desc(&tmB);
    if (p.nseg > 1) tma_prefetch_desc(&tmA1);
    if (p.nseg > 2) tma_prefetch_desc(&tmA2);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 32 * Cfg::EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tiles_per_img_grid = p.tiles_x * p.tiles_y;

  if (warp == 0) {
    if (elect_one()) {
      // ===================== TMA producer =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int n_tile = tile % p.n_tiles_n;
        const int m_tile = tile / p.n_tiles_n;
        const int x0 = (m_tile % p.tiles_x) * p.tw;
        const int y0 = ((m_tile / p.tiles_x) % p.tiles_y) * p.th;
        const int n0 = (m_tile / tiles_per_img_grid) * p.tn;
        int kcoord = 0;
        for (int seg = 0; seg < p.nseg; ++seg) {
          const CUtensorMap* tm = seg == 0 ? &tmA0 : (seg == 1 ? &tmA1 : &tmA2);
          const int taps = p.seg_taps[seg];
          const int nkb = p.seg_kb[seg];
          for (int tap = 0; tap < taps; ++tap) {
            const int dy = taps == 9 ? tap / 3 - 1 : 0;
            const int dx = taps == 9 ? tap % 3 - 1 : 0;
            for (int cb = 0; cb < nkb; ++cb) {
              mbar_wait(&empty[stage], phase ^ 1);
              mbar_arrive_expect_tx(&full[stage], A_STAGE_BYTES + Cfg::B_STAGE_BYTES);
              tma_load_4d(sA + stage * A_STAGE_BYTES, tm, &full[stage], cb * BK, x0 + dx, y0 + dy, n0);
              tma_load_2d(sB + stage * Cfg::B_STAGE_BYTES, &tmB, &full[stage], kcoord, n_tile * BN);
              kcoord += BK;
              if (++stage == STAGES) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // ===================== MMA issuer =====================
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t iter = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++iter) {
        const uint32_t as = iter & 1;
        const uint32_t aph = (iter >> 1) & 1;
        mbar_wait(&tempty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < p.total_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * A_STAGE_BYTES);
          const uint32_t b_base = smem_u32(sB + stage * Cfg::B_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            umma_bf16(d_tmem, umma_desc_sw128(a_base + k * 32), umma_desc_sw128(b_base + k * 32), idesc,
                      (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty[stage]);  // smem slot reusable once these MMAs retire
          if (kb == p.total_kb - 1) umma_commit(&tfull[as]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    // Phase 1 (row-owner layout, what tcgen05.ld gives): thread `lane` of the warp holds accumulator row
    //   32*q + lane; GEGLU's value*gelu(gate) is formed here.  The chunk (<= 32 output columns) goes to a
    //   per-warp smem staging tile.
    // Phase 2 (coalesced layout): 8 lanes cover 32 consecutive columns of one row, 4 rows per pass, so global
    //   reads of the residual and writes of the output are contiguous 128-byte (fp32) / 64-byte (bf16) segments.
    constexpr int NEW = Cfg::EPI_WARPS;
    const int ew = warp - 4;
    const int qw = ew & 3;                 // TMEM lane quadrant this warp may read
    const int cpar = ew >> 2;              // with 8 epilogue warps: even / odd chunks
    const int row = qw * 32 + lane;
    const int xl = row % p.tw;
    const int yl = (row / p.tw) % p.th;
    const int nl = row / (p.tw * p.th);
    float* stg = reinterpret_cast<float*>(smem + Cfg::STAGING_OFF) + ew * (32 * Cfg::STG_LD);
    const int sub_row = lane >> 3;         // phase 2: row within a pass of 4
    const int c4 = (lane & 7) * 4;         // phase 2: first of this lane's 4 columns inside the chunk
    constexpr int OUTW = BN;               // output columns per tile (GEGLU: BN/2)
    uint32_t iter = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++iter) {
      const uint32_t as = iter & 1;
      const uint32_t aph = (iter >> 1) & 1;
      const int n_tile = tile % p.n_tiles_n;
      const int m_tile = tile / p.n_tiles_n;
      const int x = (m_tile % p.tiles_x) * p.tw + xl;
      const int y = ((m_tile / p.tiles_x) % p.tiles_y) * p.th + yl;
      const int n = (m_tile / tiles_per_img_grid) * p.tn + nl;
      const bool valid = (x < p.W) && (y < p.H) && (n < p.n_img);
      const long long pix_own = valid ? ((long long)n * p.H + y) * p.W + x : -1;

      mbar_wait(&tfull[as], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16) + as * BN;

      // pixel index of the row each lane serves in each phase-2 pass (constant over the tile's chunks)
      long long pixs[8];
#pragma unroll
      for (int pass = 0; pass < 8; ++pass) pixs[pass] = __shfl_sync(0xffffffffu, pix_own, pass * 4 + sub_row);

      const int outw = p.geglu ? OUTW / 2 : OUTW;            // output columns this tile produces
      const int n_out = p.geglu ? p.N / 2 : p.N;             // output columns of the whole GEMM
      const int ocol_tile = n_tile * outw;
      if (p.geglu) {
        // GEGLU: per sub-chunk of 16 output columns stage [value16 | gate16], then in the coalesced layout
        // (4 lanes x 4 columns per row, 8 rows per pass) form (value + bv) * gelu(gate + bg) and store bf16/fp32.
        constexpr int HALF = BN / 2;
        const int sr8 = lane >> 2, c4g = (lane & 3) * 4;
        long long pix8[4];
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) pix8[pass] = __shfl_sync(0xffffffffu, pix_own, pass * 8 + sr8);
#pragma unroll 1
        for (int sc = cpar; sc < HALF / 16; sc += (NEW / 4)) {
          const int ocol0 = ocol_tile + sc * 16;
          if (ocol0 >= n_out) break;
          uint32_t rv[16], rg[16];
          tmem_ld16(taddr + sc * 16, rv);
          tmem_ld16(taddr + HALF + sc * 16, rg);
          float4 bv = make_float4(0.f, 0.f, 0.f, 0.f), bg = bv;
          if (p.bias) {  // bias is packed like the weight rows: [value half | gate half] per tile
            bv = __ldg(reinterpret_cast<const float4*>(p.bias + n_tile * BN + sc * 16 + c4g));
            bg = __ldg(reinterpret_cast<const float4*>(p.bias + n_tile * BN + HALF + sc * 16 + c4g));
          }
          tmem_ld_wait();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            *reinterpret_cast<uint4*>(stg + lane * Cfg::STG_LD + 4 * j) = make_uint4(rv[4 * j], rv[4 * j + 1], rv[4 * j + 2], rv[4 * j + 3]);
            *reinterpret_cast<uint4*>(stg + lane * Cfg::STG_LD + 16 + 4 * j) = make_uint4(rg[4 * j], rg[4 * j + 1], rg[4 * j + 2], rg[4 * j + 3]);
          }
          __syncwarp();
#pragma unroll
          for (int pass = 0; pass < 4; ++pass) {
            const int r = pass * 8 + sr8;
            const long long pix = pix8[pass];
            if (pix < 0) continue;
            const float4 a = *reinterpret_cast<const float4*>(stg + r * Cfg::STG_LD + c4g);
            const float4 g = *reinterpret_cast<const float4*>(stg + r * Cfg::STG_LD + 16 + c4g);
            float4 o;
            o.x = (a.x + bv.x) * gelu_tanh_fast(g.x + bg.x);
            o.y = (a.y + bv.y) * gelu_tanh_fast(g.y + bg.y);
            o.z = (a.z + bv.z) * gelu_tanh_fast(g.z + bg.z);
            o.w = (a.w + bv.w) * gelu_tanh_fast(g.w + bg.w);
            if (p.out_f32) {
              *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + pix * p.ldc + ocol0 + c4g) = o;
            } else {
              uint2 u;
              u.x = pack_bf16x2(o.x, o.y);
              u.y = pack_bf16x2(o.z, o.w);
              *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(p.out) + pix * p.ldc + ocol0 + c4g) = u;
            }
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty[as]);
        continue;
      }
      const int nchunks = (outw + 31) / 32;
#pragma unroll 1
      for (int c = cpar; c < nchunks; c += (NEW / 4)) {
        const int cw = min(32, outw - c * 32);               // 32 or 16
        const int ocol0 = ocol_tile + c * 32;
        if (ocol0 >= n_out) break;
        // ---- prefetch (coalesced layout): everything phase 2 adds that does not depend on the accumulator.
        // Issued before the TMEM load so that the 8 x 512 B reads per warp are in flight together. ----
        const bool col_ok = c4 < cw && (ocol0 + c4) < n_out;
        float4 radd[8];
#pragma unroll
        for (int pass = 0; pass < 8; ++pass) {
          float4 acc4 = make_float4(0.f, 0.f, 0.f, 0.f);
          const long long pix = pixs[pass];
          if (col_ok && pix >= 0) {
            if (p.residual) {
              if (p.res_f32) {
                acc4 = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.residual) + pix * p.ldr + ocol0 + c4));
              } else {
                const uint2 u = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(p.residual) + pix * p.ldr + ocol0 + c4));
                const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
                acc4 = make_float4(f0.x, f0.y, f1.x, f1.y);
              }
            }
            if (p.rowbias) {
              const int img = (int)(pix / ((long long)p.H * p.W));
              const float4 rb = __ldg(reinterpret_cast<const float4*>(p.rowbias + (long long)img * p.ld_rowbias + ocol0 + c4));
              acc4.x += rb.x; acc4.y += rb.y; acc4.z += rb.z; acc4.w += rb.w;
            }
          }
          radd[pass] = acc4;
        }
        // ---- phase 1 ----
        float v[32];
        if (!p.geglu) {
          uint32_t r0[16], r1[16];
          tmem_ld16(taddr + c * 32, r0);
          if (cw > 16) tmem_ld16(taddr + c * 32 + 16, r1);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) { v[i] = __uint_as_float(r0[i]); v[16 + i] = cw > 16 ? __uint_as_float(r1[i]) : 0.f; }
        } else {
          constexpr int HALF = BN / 2;
#pragma unroll
          for (int hsel = 0; hsel < 2; ++hsel) {
            if (hsel * 16 < cw) {
              uint32_t rv[16], rg[16];
              tmem_ld16(taddr + c * 32 + hsel * 16, rv);
              tmem_ld16(taddr + HALF + c * 32 + hsel * 16, rg);
              tmem_ld_wait();
              const int bcol_v = n_tile * BN + c * 32 + hsel * 16;   // bias is packed like the weight rows
              const int bcol_g = bcol_v + HALF;
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                float a = __uint_as_float(rv[i]);
                float g = __uint_as_float(rg[i]);
                if (p.bias) { a += __ldg(p.bias + bcol_v + i); g += __ldg(p.bias + bcol_g + i); }
                v[hsel * 16 + i] = a * gelu_tanh_fast(g);
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) v[hsel * 16 + i] = 0.f;
            }
          }
        }
        __syncwarp();  // previous chunk's phase-2 reads of the staging tile are done
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(stg + lane * Cfg::STG_LD + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        // ---- phase 2 ----
        float4 bz = make_float4(0.f, 0.f, 0.f, 0.f);
        if (col_ok && p.bias && !p.geglu) bz = __ldg(reinterpret_cast<const float4*>(p.bias + ocol0 + c4));
#pragma unroll
        for (int pass = 0; pass < 8; ++pass) {
          const int r = pass * 4 + sub_row;
          const long long pix = pixs[pass];
          if (!col_ok || pix < 0) continue;
          float4 o = *reinterpret_cast<const float4*>(stg + r * Cfg::STG_LD + c4);
          if (p.act) {  // activation sits between the bias and the residual (radd carries rowbias + residual)
            o.x = apply_act(o.x + bz.x, p.act) + radd[pass].x; o.y = apply_act(o.y + bz.y, p.act) + radd[pass].y;
            o.z = apply_act(o.z + bz.z, p.act) + radd[pass].z; o.w = apply_act(o.w + bz.w, p.act) + radd[pass].w;
          } else {
            o.x += bz.x + radd[pass].x; o.y += bz.y + radd[pass].y; o.z += bz.z + radd[pass].z; o.w += bz.w + radd[pass].w;
          }
          if (p.out_f32) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + pix * p.ldc + ocol0 + c4) = o;
          } else {
            uint2 u;
            u.x = pack_bf16x2(o.x, o.y);
            u.y = pack_bf16x2(o.z, o.w);
            *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(p.out) + pix * p.ldc + ocol0 + c4) = u;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int make_map_a(CUtensorMap* m, const void* ptr, long long ld, int C, int W, int H, int n, int tw, int th, int tn) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    mdx_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return MDX_ERR_CUDA;
  }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)W * ld * 2, (cuuint64_t)H * W * ld * 2};
  cuuint32_t box[4] = {(cuuint32_t)BK, (cuuint32_t)tw, (cuuint32_t)th, (cuuint32_t)tn};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdx_set_error("cuTensorMapEncodeTiled(A) failed with CUresult %d (ptr=%p ld=%lld C=%d W=%d H=%d n=%d box=%d,%d,%d)",
                  (int)r, ptr, ld, C, W, H, n, tw, th, tn);
    return MDX_ERR_CUDA;
  }
  return MDX_OK;
}

int make_map_b(CUtensorMap* m, const void* ptr, long long K, int N, int bn) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    mdx_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return MDX_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)N};
  cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)bn};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdx_set_error("cuTensorMapEncodeTiled(B) failed with CUresult %d (K=%lld N=%d bn=%d)", (int)r, K, N, bn);
    return MDX_ERR_CUDA;
  }
  return MDX_OK;
}

// Pick the 128-pixel tile (tw x th x tn) that wastes the fewest accumulator rows; prefer wide tiles.
void pick_tile(int W, int H, int n, int* tw_, int* th_, int* tn_) {
  long long best = -1;
  int btw = 128, bth = 1, btn = 1;
  for (int tw = 128; tw >= 1; tw >>= 1) {
    for (int th = 128 / tw; th >= 1; th >>= 1) {
      int tn = 128 / (tw * th);
      long long cover = (long long)((W + tw - 1) / tw) * tw * ((H + th - 1) / th) * th * ((n + tn - 1) / tn) * tn;
      if (best < 0 || cover < best) {
        best = cover;
        btw = tw; bth = th; btn = tn;
      }
    }
  }
  *tw_ = btw; *th_ = bth; *tn_ = btn;
}

int num_sms() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
  }
  return sms;
}

template <int BN>
int launch_tc(const CUtensorMap* maps, const CUtensorMap& mb, const TcParams& p, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    MDX_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        TcCfg<BN>::SMEM_BYTES));
    attr_set = true;
  }
  int grid = p.num_tiles < num_sms() ? p.num_tiles : num_sms();
  gemm_tc_kernel<BN><<<grid, TcCfg<BN>::THREADS, TcCfg<BN>::SMEM_BYTES, st>>>(maps[0], maps[1], maps[2], mb, p);
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

}  // namespace

#include "../../include/mdx.h"

extern "C" int mdx_gemm_tc_pick_block_n(int32_t N, int32_t geglu) {
  if (geglu) {
    int I = N / 2;
    if (I % 80 == 0) return 160;
    if (I % 64 == 0) return 128;
    if (I % 32 == 0) return 64;
    return 0;
  }
  if (N % 160 == 0) return 160;
  if (N % 128 == 0) return 128;
  if (N % 64 == 0) return 64;
  // fewest padded columns, ties to the larger tile
  int best = 64, best_waste = 1 << 30;
  const int cands[4] = {256, 160, 128, 64};
  for (int i = 0; i < 4; ++i) {
    int bn = cands[i];
    int waste = ((N + bn - 1) / bn) * bn - N;
    if (waste < best_waste) {
      best_waste = waste;
      best = bn;
    }
  }
  return best;
}

extern "C" int mdx_gemm_tc(const mdx_gemm_args* a, void* stream) {
  MDX_CHECK_ARG(a != nullptr, "mdx_gemm_tc: args is NULL");
  MDX_CHECK_ARG(a->nseg >= 1 && a->nseg <= 3, "mdx_gemm_tc: nseg must be 1..3 (got %d)", a->nseg);
  MDX_CHECK_ARG(a->n_img > 0 && a->H > 0 && a->W > 0 && a->N > 0, "mdx_gemm_tc: empty problem");
  MDX_CHECK_ARG(a->N % 16 == 0, "mdx_gemm_tc: N must be a multiple of 16 (got %d)", a->N);
  MDX_CHECK_ARG(a->b && a->out, "mdx_gemm_tc: NULL operand");
  MDX_CHECK_ARG(a->out_dtype == MDX_BF16 || a->out_dtype == MDX_F32, "mdx_gemm_tc: bad out_dtype");
  MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->b) & 15) == 0, "mdx_gemm_tc: B must be 16-byte aligned");
  MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->out) & 15) == 0 && a->ldc % 8 == 0,
                "mdx_gemm_tc: out must be 16-byte aligned with ldc %% 8 == 0");
  if (a->residual)
    MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->residual) & 15) == 0 && a->ldr % 8 == 0 &&
                      (a->residual_dtype == MDX_BF16 || a->residual_dtype == MDX_F32),
                  "mdx_gemm_tc: residual must be 16-byte aligned with ldr %% 8 == 0, dtype bf16 or fp32");
  if (a->bias) MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->bias) & 15) == 0, "mdx_gemm_tc: bias alignment");
  if (a->rowbias)
    MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->rowbias) & 15) == 0 && a->ld_rowbias % 4 == 0,
                  "mdx_gemm_tc: rowbias alignment");
  if (a->geglu) MDX_CHECK_ARG(!a->rowbias && !a->residual && !a->act && a->N % 32 == 0, "mdx_gemm_tc: geglu excludes rowbias/residual/act");
  MDX_CHECK_ARG(a->act >= 0 && a->act <= 2, "mdx_gemm_tc: act must be 0, 1 (gelu) or 2 (quick gelu)");
  if (a->act) MDX_CHECK_ARG(!a->rowbias, "mdx_gemm_tc: act excludes rowbias");

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.nseg = a->nseg;
  long long ktot = 0;
  for (int s = 0; s < a->nseg; ++s) {
    MDX_CHECK_ARG(a->a[s] != nullptr, "mdx_gemm_tc: A segment %d is NULL", s);
    MDX_CHECK_ARG(a->seg_c[s] > 0 && a->seg_c[s] % 64 == 0, "mdx_gemm_tc: seg_c[%d]=%d must be a multiple of 64", s,
                  a->seg_c[s]);
    MDX_CHECK_ARG(a->seg_taps[s] == 1 || a->seg_taps[s] == 9, "mdx_gemm_tc: seg_taps[%d] must be 1 or 9", s);
    MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->a[s]) & 15) == 0 && a->lda[s] % 8 == 0 && a->lda[s] >= a->seg_c[s],
                  "mdx_gemm_tc: A segment %d must be 16-byte aligned with lda %% 8 == 0", s);
    p.seg_kb[s] = a->seg_c[s] / 64;
    p.seg_taps[s] = a->seg_taps[s];
    p.total_kb += p.seg_kb[s] * p.seg_taps[s];
    ktot += (long long)a->seg_taps[s] * a->seg_c[s];
  }
  int bn = a->block_n ? a->block_n : mdx_gemm_tc_pick_block_n(a->N, a->geglu);
  MDX_CHECK_ARG(bn == 64 || bn == 128 || bn == 160 || bn == 256, "mdx_gemm_tc: unsupported block_n %d", bn);
  if (a->geglu) MDX_CHECK_ARG((a->N / 2) % (bn / 2) == 0, "mdx_gemm_tc: geglu needs N/2 %% (block_n/2) == 0");

  p.n_img = a->n_img; p.H = a->H; p.W = a->W;
  pick_tile(a->W, a->H, a->n_img, &p.tw, &p.th, &p.tn);
  p.tiles_x = (a->W + p.tw - 1) / p.tw;
  p.tiles_y = (a->H + p.th - 1) / p.th;
  int tiles_n = (a->n_img + p.tn - 1) / p.tn;
  p.N = a->N;
  p.n_tiles_n = (a->N + bn - 1) / bn;
  long long nt = (long long)p.tiles_x * p.tiles_y * tiles_n * p.n_tiles_n;
  MDX_CHECK_ARG(nt < (1ll << 31), "mdx_gemm_tc: too many tiles");
  p.num_tiles = (int)nt;
  p.out = a->out; p.ldc = a->ldc; p.out_f32 = a->out_dtype == MDX_F32;
  p.bias = a->bias; p.rowbias = a->rowbias; p.ld_rowbias = a->ld_rowbias;
  p.residual = a->residual; p.ldr = a->ldr; p.res_f32 = a->residual_dtype == MDX_F32;
  p.geglu = a->geglu;
  p.act = a->act;

  CUtensorMap maps[3];
  for (int s = 0; s < 3; ++s) {
    int src = s < a->nseg ? s : 0;
    int rc = make_map_a(&maps[s], a->a[src], a->lda[src], a->seg_c[src], a->W, a->H, a->n_img, p.tw, p.th, p.tn);
    if (rc) return rc;
  }
  CUtensorMap mb;
  int rc = make_map_b(&mb, a->b, ktot, a->N, bn);
  if (rc) return rc;

  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (bn) {
    case 64: return launch_tc<64>(maps, mb, p, st);
    case 128: return launch_tc<128>(maps, mb, p, st);
    case 160: return launch_tc<160>(maps, mb, p, st);
    default: return launch_tc<256>(maps, mb, p, st);
  }
}
