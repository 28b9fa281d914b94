// Tensor-core implicit GEMM for sm_100a: TMA (4-D NHWC boxes, zero-filled halos) -> 128B-swizzled smem ring
// -> tcgen05.mma (cta_group::1, M=128, N=BN, bf16 x bf16 -> fp32 in TMEM, double-buffered accumulator)
// -> tcgen05.ld epilogue (bias / per-image emb bias / GEGLU / residual) -> bf16 or fp32 NHWC rows.
//
// One kernel serves every conv3x3 (stride 1), conv1x1 and nn.Dense of the denoising path; see include/mdx.h
// (mdx_gemm_tc) for the reference call sites it replaces.  A convolution is never lowered to an im2col
// buffer: for tap (r,s) the producer asks TMA for the activation box shifted by (r-1, s-1); out-of-image
// coordinates are filled with zeros by the TMA unit, which is exactly pad_mode='pad', padding=1.
//
// Warp roles (512 threads, 1 CTA / SM, persistent over output tiles):
//   warp 0   : TMA producer (one elected thread)
//   warp 1   : MMA issuer  (one elected thread)
//   warp 2   : TMEM allocate / free
//   warps 4-15: epilogue; warp w reads TMEM lane quadrant w % 4 (accumulator rows 32*(w%4) .. +31) and takes
//              every third column chunk — three warps per scheduler keep the transpose / store chain busy
#include "common.cuh"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB

// Division by a run-time constant as multiply-high + shift (dividends < 2^31): the tile -> (image, y, x) arithmetic
// runs once per tile in every epilogue warp, and a hardware-less integer divide is ~30 instructions.
struct FastDiv {
  uint32_t d, mul, shr;
};
inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  uint32_t s = 0;
  while ((1ull << s) < d) ++s;                                   // ceil(log2 d)
  f.mul = (uint32_t)(((1ull << 32) * ((1ull << s) - d)) / d + 1);
  f.shr = s;
  return f;
}
__device__ __forceinline__ uint32_t fdiv(uint32_t x, const FastDiv& f) { return (__umulhi(x, f.mul) + x) >> f.shr; }

struct TcParams {
  int nseg;
  int seg_kb[3];    // 64-wide channel blocks per tap
  int seg_taps[3];  // 1, 9 (3x3) or 4 (2x2 window of one phase of a 2x-upsampled 3x3 conv)
  int up_py, up_px; // phase of the upsampled output this launch produces (taps == 4)
  int up;           // 1: output / residual rows are pixels (2y+py, 2x+px) of the [n, 2H, 2W] image
  int stride;       // 1, or 2: H, W below are the OUTPUT image of a stride-2 3x3 conv (the A map strides the input)
  long long hw_out; // pixels per image in the output addressing
  int total_kb;
  int n_img, H, W;
  int tn, th, tw;
  int tiles_x, tiles_y;
  FastDiv fd_tx, fd_ty, fd_tpi, fd_ntn;   // tiles_x, tiles_y, tiles_x * tiles_y, n_tiles_n
  int N, n_tiles_n, num_tiles;
  int m_tiles;
  float* chan_stats;   // optional [n_img][slots][2][N]: per-(image, 32-row slot, channel) sum / sum of squares of the
                       // stored output; slot = (phase *) tile-in-image * 4 + lane quadrant, every slot written once
  int stat_slots;      // slots per image
  int debug;        // MDX_GEMM_DEBUG bit mask (profiling only): 1 = no global stores, 2 = epilogue does nothing
  int b_resident;   // K fits the B ring: each CTA keeps ONE n-tile's weight slice in smem and streams only A tiles
  // Cluster launch (0 = none, else the cluster size 2 or 4).  The persistent schedule gives consecutive blocks
  // consecutive tiles, n fastest, so the CTAs of a cluster hold mc_gn neighbouring n-tiles of mc / mc_gn m-tiles at the
  // same time.  CTAs on the same m-tile each fetch 1 / mc_gn of every A block and multicast it to that group; CTAs on the
  // same n-tile (2 x 2 clusters) do the same with the B block.  The tensor maps then hold the sub-boxes.
  int mc, mc_gn;
  int mc_cy, mc_cx;             // the A tile is cut into (tn / mc_sn) x mc_cy x mc_cx sub-boxes of mc_sn x mc_sy x mc_sx
  int mc_sn, mc_sy, mc_sx;      // images x rows x pixels, numbered in row order of the tile
  int mc_grid_cap;  // host only: CTAs the device keeps resident in clusters of this size (the schedule is persistent)
  int halo;         // CTA-pair form, one 3x3 segment: per (dx, 64-channel block) ONE A box of th + 2 image rows serves the
  int halo_bytes;   // three dy taps as row-offset views (halo_row_bytes apart) of the same smem — a third to a half of
  int halo_row_bytes;   // the A traffic; the A ring (3 stages of 32 KiB) and the B ring (6 stages) run on their own barriers
  int cg2;          // 1: clusters of two as a CTA PAIR (tcgen05 cta_group::2): the pair holds two consecutive m-tiles of one
                    // n-tile, the even CTA issues M = 256 MMAs over both SMs, each CTA loads its own A block and half of
                    // the B block (the B map holds BLOCK_N / 2 rows).  Opt-in (MDX_GEMM_CG2=1), excludes mc.
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
  static constexpr int STAGES = (BN >= 256) ? 3 : (BN >= 192 ? 4 : (BN >= 128 ? 5 : 7));
  // CTA-pair form: a CTA stages only half of each B block, so the same ring space holds more stages
  static constexpr int STAGES_CG2 = (STAGES * (A_STAGE_BYTES + B_STAGE_BYTES)) / (A_STAGE_BYTES + B_STAGE_BYTES / 2);
  static constexpr int TMEM_COLS = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
  static constexpr int EPI_WARPS = 12;                        // 4 TMEM lane quadrants x 3 column shares
  static constexpr int THREADS = 128 + 32 * EPI_WARPS;        // 512: one role warpgroup + three epilogue warpgroups
  static constexpr int STG_BYTES = 3072;                      // per warp: 32 rows x 80 B staging + 512 B bias slice
  static constexpr int STAGING_OFF = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + 256;
  static constexpr int SMEM_BYTES = STAGING_OFF + EPI_WARPS * STG_BYTES + 1024 /*align*/;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
};

// ------------------------------------------------------------------------------------------------------
// Epilogue of one 128 x BN accumulator tile for one warp (TMEM lane quadrant in `taddr`, column share cpar of
// cstep).  tcgen05.ld hands every thread consecutive columns of ITS accumulator row; global memory wants
// consecutive columns of the SAME row across lanes.  The transpose goes through a per-warp staging tile of
// 32 rows x 64 payload bytes (80-byte pitch: conflict-free for 16-byte accesses in both directions):
//   plain bf16 (no residual / per-image bias): 32-column chunks; bias + activation + bf16 packing happen in the
//     row-owner layout, the coalesced phase is LDS.128 -> STG.128 only;
//   general (fp32 out and/or residual / per-image bias): 16-column fp32 chunks; the coalesced phase adds bias,
//     activation and the residual (+ per-image bias), whose loads for chunk c+1 are issued before chunk c is
//     processed.
// Either way 4 lanes cover 64 contiguous bytes of one row and a warp instruction covers 8 rows.
// ------------------------------------------------------------------------------------------------------
template <bool RES_F32>
__device__ __forceinline__ float4 epi_load_add(const TcParams& p, long long pix, int col) {
  float4 acc4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.residual) {
    if (RES_F32) {
      acc4 = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.residual) + pix * p.ldr + col));
    } else {
      const uint2 u = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(p.residual) + pix * p.ldr + col));
      const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
      acc4 = make_float4(f0.x, f0.y, f1.x, f1.y);
    }
  }
  if (p.rowbias) {
    const int img = (int)(pix / p.hw_out);
    const float4 rb = __ldg(reinterpret_cast<const float4*>(p.rowbias + (long long)img * p.ld_rowbias + col));
    acc4.x += rb.x; acc4.y += rb.y; acc4.z += rb.z; acc4.w += rb.w;
  }
  return acc4;
}

template <int BN, bool OUT_F32, bool HAS_ADD>
__device__ __forceinline__ void epi_tile(const TcParams& p, uint32_t taddr, uint32_t stg, int lane, int cpar,
                                         int cstep, int n_tile, long long pix_own, long long stat_slot,
                                         uint64_t* tfull_bar, uint32_t tfull_parity) {
  const int n_out = p.N;
  const int ocol_tile = n_tile * BN;
  const int sr8 = lane >> 2, q4 = lane & 3;
  long long pix8[4];
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) pix8[pass] = __shfl_sync(0xffffffffu, pix_own, pass * 8 + sr8);
  if constexpr (!OUT_F32 && !HAS_ADD) {
    constexpr int NCH = BN / 32;
    mbar_wait(tfull_bar, tfull_parity);
    tc_fence_after();
#pragma unroll 1
    for (int c = cpar; c < NCH; c += cstep) {
      const int ocol0 = ocol_tile + c * 32;
      if (ocol0 >= n_out) break;
      uint32_t r[32];
      tmem_ld32(taddr + c * 32, r);
      float4 b[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        b[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p.bias && ocol0 + 4 * j < n_out) b[j] = __ldg(reinterpret_cast<const float4*>(p.bias + ocol0) + j);
      }
      tmem_ld_wait();
      uint32_t pk[16];
      if (p.act) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          pk[2 * j] = pack_bf16x2(apply_act(__uint_as_float(r[4 * j]) + b[j].x, p.act),
                                  apply_act(__uint_as_float(r[4 * j + 1]) + b[j].y, p.act));
          pk[2 * j + 1] = pack_bf16x2(apply_act(__uint_as_float(r[4 * j + 2]) + b[j].z, p.act),
                                      apply_act(__uint_as_float(r[4 * j + 3]) + b[j].w, p.act));
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          pk[2 * j] = pack_bf16x2(__uint_as_float(r[4 * j]) + b[j].x, __uint_as_float(r[4 * j + 1]) + b[j].y);
          pk[2 * j + 1] = pack_bf16x2(__uint_as_float(r[4 * j + 2]) + b[j].z, __uint_as_float(r[4 * j + 3]) + b[j].w);
        }
      }
      __syncwarp();  // the previous chunk's coalesced reads of the staging tile are done
#pragma unroll
      for (int j = 0; j < 4; ++j) sts128(stg + lane * 80 + 16 * j, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
      __syncwarp();
      const bool cok = ocol0 + q4 * 8 < n_out;
      bf16* obase = reinterpret_cast<bf16*>(p.out) + ocol0 + q4 * 8;
#pragma unroll
      for (int pass = 0; pass < 4; ++pass) {
        const uint4 u = lds128(stg + (pass * 8 + sr8) * 80 + q4 * 16);
        const long long pix = pix8[pass];
        if (cok && pix >= 0 && !(p.debug & 1)) *reinterpret_cast<uint4*>(obase + pix * p.ldc) = u;
      }
    }
  } else {
    constexpr int NCH = BN / 16;
    // residual (+ per-image bias) of this warp's first TWO chunks, requested before the accumulator is even ready:
    // nothing here depends on the MMAs, and two chunks in flight per warp keep enough bytes outstanding for HBM
    float4 radd[4], rnext[4];
    auto load_add = [&](float4 (&dst)[4], int c) {
      const int col = ocol_tile + c * 16 + q4 * 4;
      const bool ok = c < NCH && col < n_out;
#pragma unroll
      for (int pass = 0; pass < 4; ++pass) {
        dst[pass] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok && pix8[pass] >= 0)
          dst[pass] = p.res_f32 ? epi_load_add<true>(p, pix8[pass], col) : epi_load_add<false>(p, pix8[pass], col);
      }
    };
    if constexpr (HAS_ADD) {
      load_add(radd, cpar);
      load_add(rnext, cpar + cstep);
    }
    mbar_wait(tfull_bar, tfull_parity);
    tc_fence_after();
#pragma unroll 1
    for (int c = cpar; c < NCH; c += cstep) {
      const int ocol0 = ocol_tile + c * 16;
      if (ocol0 >= n_out) break;
      const int col = ocol0 + q4 * 4;
      const bool col_ok = col < n_out;
      uint32_t r[16];
      tmem_ld16(taddr + c * 16, r);
      float4 rnext2[4];
      if constexpr (HAS_ADD) load_add(rnext2, c + 2 * cstep);   // two chunks ahead
      float4 bz = make_float4(0.f, 0.f, 0.f, 0.f);
      if (col_ok && p.bias) bz = __ldg(reinterpret_cast<const float4*>(p.bias + col));
      tmem_ld_wait();
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 4; ++j) sts128(stg + lane * 80 + 16 * j, r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
      __syncwarp();
      float cs[4] = {0.f, 0.f, 0.f, 0.f}, cq[4] = {0.f, 0.f, 0.f, 0.f};   // column sums over this lane's rows
#pragma unroll
      for (int pass = 0; pass < 4; ++pass) {
        const uint4 u = lds128(stg + (pass * 8 + sr8) * 80 + q4 * 16);
        float4 o = make_float4(__uint_as_float(u.x) + bz.x, __uint_as_float(u.y) + bz.y, __uint_as_float(u.z) + bz.z,
                               __uint_as_float(u.w) + bz.w);
        if (p.act) {  // activation sits between the bias and the residual
          o.x = apply_act(o.x, p.act); o.y = apply_act(o.y, p.act); o.z = apply_act(o.z, p.act); o.w = apply_act(o.w, p.act);
        }
        if constexpr (HAS_ADD) { o.x += radd[pass].x; o.y += radd[pass].y; o.z += radd[pass].z; o.w += radd[pass].w; }
        const long long pix = pix8[pass];
        if (!col_ok || pix < 0 || (p.debug & 1)) continue;
        if constexpr (OUT_F32) {
          *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + pix * p.ldc + col) = o;
        } else {
          uint2 w;
          w.x = pack_bf16x2(o.x, o.y);
          w.y = pack_bf16x2(o.z, o.w);
          *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(p.out) + pix * p.ldc + col) = w;
          if (p.chan_stats) {   // the statistics are those of the values a reader of `out` sees
            const float2 lo = unpack_bf16x2(w.x), hi = unpack_bf16x2(w.y);
            o = make_float4(lo.x, lo.y, hi.x, hi.y);
          }
        }
        if (p.chan_stats) {
          cs[0] += o.x; cs[1] += o.y; cs[2] += o.z; cs[3] += o.w;
          cq[0] = fmaf(o.x, o.x, cq[0]); cq[1] = fmaf(o.y, o.y, cq[1]);
          cq[2] = fmaf(o.z, o.z, cq[2]); cq[3] = fmaf(o.w, o.w, cq[3]);
        }
      }
      if (p.chan_stats && stat_slot >= 0) {
        // GroupNorm statistics for the consumer of this tensor (mdx_groupnorm_pre): reduce the 8 lanes that hold
        // the same 4 columns (lane = 4*row_lane + q4); lanes 0..3 store the warp's partial sums into its own slot —
        // no atomics, every slot is written exactly once, the consumer adds the slots in a fixed order
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
          for (int m = 4; m < 32; m <<= 1) {
            cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], m);
            cq[k] += __shfl_xor_sync(0xffffffffu, cq[k], m);
          }
        }
        if (lane < 4 && col_ok) {
          float* dst = p.chan_stats + (long long)stat_slot * 2 * p.N + col;
          *reinterpret_cast<float4*>(dst) = make_float4(cs[0], cs[1], cs[2], cs[3]);
          *reinterpret_cast<float4*>(dst + p.N) = make_float4(cq[0], cq[1], cq[2], cq[3]);
        }
      }
      if constexpr (HAS_ADD) {
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) { radd[pass] = rnext[pass]; rnext[pass] = rnext2[pass]; }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------
// Lean epilogue for the layers whose tiles are epilogue-bound (small K: the transformer's projections, the 1x1
// convs): fp32 or bf16 output, optional residual (fp32 trunk or bf16), optional per-image bias row (emb add),
// optional GroupNorm statistics; no activation.  Same data flow as epi_tile's general path (16-column chunks
// through the per-warp staging tile), but
//   * everything that does not change from chunk to chunk is computed once per tile: the four row pointers of a
//     lane for `out` and `residual`, the row-validity mask, the bias pointers;
//   * the variants are template parameters instead of run-time branches inside the chunk loop;
//   * one residual buffer, refilled for the next chunk as soon as the current chunk has consumed it (no buffer is
//     ever copied: a register move of a pending load's destination waits for the load); the first chunk's residual
//     is requested before the accumulator wait.
// The general path executes ~600 instructions per chunk and warp — with three epilogue warps per scheduler the
// instruction stream itself, not HBM, set the tile time of the K = 320 layers; this one executes ~100.
// ------------------------------------------------------------------------------------------------------
template <int RES>
__device__ __forceinline__ void res_load(uint4 (&q)[4], const char* base, const uint32_t (&ro)[4], int eoff, uint32_t ok) {
  constexpr int RSZ = RES == 1 ? 4 : 2;
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) {
    if ((ok >> pass) & 1u) {
      const char* src = base + (size_t)(ro[pass] + (uint32_t)eoff) * RSZ;
      if constexpr (RES == 1) {
        q[pass] = __ldg(reinterpret_cast<const uint4*>(src));
      } else {
        const uint2 t = __ldg(reinterpret_cast<const uint2*>(src));
        q[pass].x = t.x;
        q[pass].y = t.y;
      }
    }
  }
}

#ifndef MDX_EPI_NBUF
#define MDX_EPI_NBUF 2      // residual chunks in flight per warp (named buffers; A/B: 1, 2, 3)
#endif

// Per-tile state of one epilogue warp on the lean path.
struct EpiFastCtx {
  uint32_t oo[4], ro[4];     // element offsets of this lane's four rows in out / residual (first column of chunk 0)
  uint32_t ok;               // bit `pass` set: that row exists
  char* obase;
  const char* rbase;
  const float* bp;           // bias + first column, or NULL
  const float* rbp;          // per-image bias row + first column, or NULL
  float* stats;              // chan_stats slot + first column, or NULL
  int N;
  uint32_t taddr, stg;
  int lane, sr8, q4;
};

// One 16-column chunk `c`: accumulator -> staging tile -> coalesced rows (+ bias, + residual `cur`) -> global.
template <bool OUT_F32, int RES, bool STATS>
__device__ __forceinline__ void epi_fast_chunk(const EpiFastCtx& x, int c, const uint4 (&cur)[4]) {
  constexpr int OSZ = OUT_F32 ? 4 : 2;
  uint32_t r[16];
  tmem_ld16(x.taddr + c * 16, r);
  float4 bz = make_float4(0.f, 0.f, 0.f, 0.f);
  if (x.bp) bz = __ldg(reinterpret_cast<const float4*>(x.bp + c * 16));
  if (x.rbp) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(x.rbp + c * 16));
    bz.x += t.x; bz.y += t.y; bz.z += t.z; bz.w += t.w;
  }
  tmem_ld_wait();
  __syncwarp();   // the previous chunk's coalesced reads of the staging tile are done
#pragma unroll
  for (int j = 0; j < 4; ++j) sts128(x.stg + x.lane * 80 + 16 * j, r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
  __syncwarp();
  float cs[4] = {0.f, 0.f, 0.f, 0.f}, cq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) {
    const uint4 u = lds128(x.stg + (pass * 8 + x.sr8) * 80 + x.q4 * 16);
    float4 o = make_float4(__uint_as_float(u.x) + bz.x, __uint_as_float(u.y) + bz.y, __uint_as_float(u.z) + bz.z,
                           __uint_as_float(u.w) + bz.w);
    if constexpr (RES == 1) {
      o.x += __uint_as_float(cur[pass].x); o.y += __uint_as_float(cur[pass].y);
      o.z += __uint_as_float(cur[pass].z); o.w += __uint_as_float(cur[pass].w);
    } else if constexpr (RES == 2) {
      const float2 f0 = unpack_bf16x2(cur[pass].x), f1 = unpack_bf16x2(cur[pass].y);
      o.x += f0.x; o.y += f0.y; o.z += f1.x; o.w += f1.y;
    }
    if ((x.ok >> pass) & 1u) {
      char* dst = x.obase + (size_t)(x.oo[pass] + (uint32_t)(c * 16)) * OSZ;
      if constexpr (OUT_F32) {
        *reinterpret_cast<float4*>(dst) = o;
      } else {
        uint2 w;
        w.x = pack_bf16x2(o.x, o.y);
        w.y = pack_bf16x2(o.z, o.w);
        *reinterpret_cast<uint2*>(dst) = w;
        if constexpr (STATS) {   // the statistics are those of the values a reader of `out` sees
          const float2 lo = unpack_bf16x2(w.x), hi = unpack_bf16x2(w.y);
          o = make_float4(lo.x, lo.y, hi.x, hi.y);
        }
      }
      if constexpr (STATS) {
        cs[0] += o.x; cs[1] += o.y; cs[2] += o.z; cs[3] += o.w;
        cq[0] = fmaf(o.x, o.x, cq[0]); cq[1] = fmaf(o.y, o.y, cq[1]);
        cq[2] = fmaf(o.z, o.z, cq[2]); cq[3] = fmaf(o.w, o.w, cq[3]);
      }
    }
  }
  if constexpr (STATS) {
    if (x.stats) {   // as in epi_tile: 8 lanes share 4 columns; lanes 0..3 store the warp's partial sums for its slot
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int m = 4; m < 32; m <<= 1) {
          cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], m);
          cq[k] += __shfl_xor_sync(0xffffffffu, cq[k], m);
        }
      }
      if (x.lane < 4) {
        float* dst = x.stats + c * 16;
        *reinterpret_cast<float4*>(dst) = make_float4(cs[0], cs[1], cs[2], cs[3]);
        *reinterpret_cast<float4*>(dst + x.N) = make_float4(cq[0], cq[1], cq[2], cq[3]);
      }
    }
  }
}

template <int BN, bool OUT_F32, int RES, bool STATS>
__device__ __forceinline__ void epi_fast(const TcParams& p, uint32_t taddr, uint32_t stg, int lane, int cpar, int n_tile,
                                         long long pix_own, long long stat_slot, const float* rb_row,
                                         uint64_t* tfull_bar, uint32_t tfull_parity) {
  constexpr int CSTEP = TcCfg<BN>::EPI_WARPS / 4;
  EpiFastCtx x;
  x.lane = lane; x.sr8 = lane >> 2; x.q4 = lane & 3;
  x.taddr = taddr; x.stg = stg; x.N = p.N;
  const int colbase = n_tile * BN + x.q4 * 4;               // this lane's first column of chunk 0
  int nch = p.N - n_tile * BN;                               // 16-column chunks of this tile inside N (N % 16 == 0)
  nch = (nch > BN ? BN : nch) >> 4;
  x.ok = 0;
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) {
    const long long pix = __shfl_sync(0xffffffffu, pix_own, pass * 8 + x.sr8);
    if (pix >= 0) x.ok |= 1u << pass;
    const long long pp = pix < 0 ? 0 : pix;
    x.oo[pass] = (uint32_t)(pp * p.ldc + colbase);           // (the host admits this path only while they fit 32 bits)
    x.ro[pass] = RES ? (uint32_t)(pp * p.ldr + colbase) : 0u;
  }
  x.obase = reinterpret_cast<char*>(p.out);
  x.rbase = reinterpret_cast<const char*>(p.residual);
  x.bp = p.bias ? p.bias + colbase : nullptr;
  x.rbp = rb_row ? rb_row + colbase : nullptr;
  x.stats = (STATS && stat_slot >= 0) ? p.chan_stats + (long long)stat_slot * 2 * p.N + colbase : nullptr;
  int c = cpar;
  // Residual buffers: MDX_EPI_NBUF named register sets, filled NBUF - 1 chunks ahead (with one buffer: refilled right
  // after it was consumed).  The chunk loop is unrolled NBUF times, so a buffer is never copied — a register move of
  // a pending load's destination would wait for the load.  The first chunks' loads go out before the accumulator
  // wait: nothing in them depends on the MMAs.
  uint4 ra[4];
#if MDX_EPI_NBUF >= 2
  uint4 rb[4];
#endif
#if MDX_EPI_NBUF >= 3
  uint4 rc[4];
#endif
  auto load = [&](uint4 (&q)[4], int cc) __attribute__((always_inline)) {
    if constexpr (RES != 0) {
      if (cc < nch) res_load<RES>(q, x.rbase, x.ro, cc * 16, x.ok);
    }
  };
  load(ra, c);
#if MDX_EPI_NBUF >= 3
  load(rb, c + CSTEP);
#endif
  mbar_wait(tfull_bar, tfull_parity);
  tc_fence_after();
#pragma unroll 1
  for (;;) {
#if MDX_EPI_NBUF == 1
    if (c >= nch) break;
    epi_fast_chunk<OUT_F32, RES, STATS>(x, c, ra);
    load(ra, c + CSTEP);
    c += CSTEP;
#elif MDX_EPI_NBUF == 2
    if (c >= nch) break;
    load(rb, c + CSTEP);
    epi_fast_chunk<OUT_F32, RES, STATS>(x, c, ra);
    c += CSTEP;
    if (c >= nch) break;
    load(ra, c + CSTEP);
    epi_fast_chunk<OUT_F32, RES, STATS>(x, c, rb);
    c += CSTEP;
#else
    if (c >= nch) break;
    load(rc, c + 2 * CSTEP);
    epi_fast_chunk<OUT_F32, RES, STATS>(x, c, ra);
    c += CSTEP;
    if (c >= nch) break;
    load(ra, c + 2 * CSTEP);
    epi_fast_chunk<OUT_F32, RES, STATS>(x, c, rb);
    c += CSTEP;
    if (c >= nch) break;
    load(rb, c + 2 * CSTEP);
    epi_fast_chunk<OUT_F32, RES, STATS>(x, c, rc);
    c += CSTEP;
#endif
  }
}

// Epilogue kinds: each is its own kernel instantiation, so the register allocation and code of one epilogue do not
// weigh on another.  EPI_FAST + (OUT_F32 * 6 + RES * 2 + STATS) selects an epi_fast variant.
constexpr int EPI_GENERIC = 0;   // plain bf16 store, activations, profiling switches, ragged per-image bias (epi_tile)
constexpr int EPI_GEGLU = 1;
constexpr int EPI_FAST = 2;

// CG2: the CTA-pair form (tcgen05 cta_group::2).  A template parameter, not a runtime switch: a kernel that contains
// cta_group::2 instructions can only be launched as a cluster ("cluster misconfiguration" otherwise).
template <int BN, int EPI, bool CG2 = false>
__global__ void __launch_bounds__(TcCfg<BN>::THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB,
               const TcParams p) {
  using Cfg = TcCfg<BN>;
  constexpr int STAGES = CG2 ? Cfg::STAGES_CG2 : Cfg::STAGES;
  constexpr int B_BYTES = CG2 ? Cfg::B_STAGE_BYTES / 2 : Cfg::B_STAGE_BYTES;   // B bytes a CTA stages per k-block
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* bfull = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bfull + 1);
  // halo mode (CTA-pair instantiations only; their ring leaves room before the epilogue staging): its own rings
  constexpr int HALO_SA = 3, HALO_SB = 6, HALO_A_STAGE = 32768;
  uint64_t* ha_full = bfull + 2;           // HALO_SA
  uint64_t* ha_empty = ha_full + HALO_SA;  // HALO_SA
  uint64_t* hb_full = ha_empty + HALO_SA;  // HALO_SB
  uint64_t* hb_empty = hb_full + HALO_SB;  // HALO_SB
  uint8_t* sBh = smem + HALO_SA * HALO_A_STAGE;
  static_assert(!CG2 || BN != 160 ||
                    (HALO_SA * HALO_A_STAGE + HALO_SB * (Cfg::B_STAGE_BYTES / 2) <= Cfg::STAGES_CG2 * (A_STAGE_BYTES + Cfg::B_STAGE_BYTES / 2) &&
                     Cfg::STAGES_CG2 * (A_STAGE_BYTES + Cfg::B_STAGE_BYTES / 2) + (2 * Cfg::STAGES_CG2 + 7 + 2 * (HALO_SA + HALO_SB)) * 8 <= Cfg::STAGING_OFF),
                "halo rings");

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
      // multicast: a stage is rewritten by every CTA this one shares A or B with, so all of them must have drained it
      mbar_init(&empty[i], p.mc ? p.mc_gn + ((p.mc / p.mc_gn == 2 && !p.b_resident) ? 1 : 0) : 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], CG2 ? 2 * Cfg::EPI_WARPS : 32 * Cfg::EPI_WARPS);   // pair: one arrival per epilogue warp of both CTAs
    }
    mbar_init(bfull, 1);
    if constexpr (CG2) {
      if (p.halo) {
        for (int i = 0; i < HALO_SA; ++i) { mbar_init(&ha_full[i], 1); mbar_init(&ha_empty[i], 1); }
        for (int i = 0; i < HALO_SB; ++i) { mbar_init(&hb_full[i], 1); mbar_init(&hb_empty[i], 1); }
      }
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (CG2) { tmem_alloc_cg2(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish_cg2(); }
    else { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (CG2 || p.mc) cluster_sync_all();    // the peer's barriers exist before anything is multicast at them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tiles_per_img_grid = p.tiles_x * p.tiles_y;
  // Tile schedule.  Streaming mode: tiles round-robin over CTAs, n fastest (CTAs that run together share the A tile
  // in L2).  Resident mode (small K): CTA b owns n-tile b % n_tiles_n for its whole life and walks m-tiles, so the
  // weight slice is fetched once per CTA instead of once per tile.
  const int cta_per_n = p.b_resident ? (int)gridDim.x / p.n_tiles_n : 1;
  const int pair_rank = CG2 ? (int)cluster_ctarank() : 0;
  auto tile_at = [&](int it, int& m_tile, int& n_tile) -> bool {
    if constexpr (CG2) {                     // pair c walks (two m-tiles, n-tile) items, n fastest
      const long long item = (long long)(blockIdx.x >> 1) + (long long)it * (gridDim.x >> 1);
      if (item >= (p.num_tiles >> 1)) return false;
      const int pm = (int)fdiv((uint32_t)item, p.fd_ntn);
      n_tile = (int)item - pm * p.n_tiles_n;
      m_tile = 2 * pm + pair_rank;
      return true;
    }
    if (p.b_resident) {
      n_tile = (int)blockIdx.x % p.n_tiles_n;
      m_tile = (int)blockIdx.x / p.n_tiles_n + it * cta_per_n;
      return m_tile < p.m_tiles;
    }
    const long long tile = (long long)blockIdx.x + (long long)it * gridDim.x;
    if (tile >= p.num_tiles) return false;
    m_tile = (int)fdiv((uint32_t)tile, p.fd_ntn);
    n_tile = (int)tile - m_tile * p.n_tiles_n;
    return true;
  };

  // Register rebalancing (512 threads start at 128 registers each): the role warpgroup keeps 40, the three epilogue
  // warpgroups take 152 (128 x 40 + 384 x 152 = 63488 <= 65536).  Each setmaxnreg sits at the top of its own branch.
  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
  if (warp == 0) {
    if (elect_one()) {
      // ===================== TMA producer =====================
      int stage = 0;
      uint32_t phase = 0;
      int m_tile, n_tile;
      // cluster rank -> (rn, rm): which n-tile / m-tile of the cluster; rn is also which slice of A this CTA fetches,
      // rm which slice of B
      const int rank = p.mc ? (int)cluster_ctarank() : 0;
      const int gn = p.mc ? p.mc_gn : 1;
      const int rn = rank % gn, rm = rank / gn;
      const int gm = (p.mc && p.mc / p.mc_gn == 2) ? 2 : 1;     // B pieces per block (resident slices are loaded whole, unshared)
      const uint16_t a_mask = (uint16_t)(((1u << gn) - 1u) << (rm * gn));                  // same m-tile
      const uint16_t b_mask = (uint16_t)((gm == 2 ? (1u | (1u << gn)) : 1u) << rn);          // same n-tile
      const int a_slice = rn * (A_STAGE_BYTES / gn);
      const int b_rows = BN / gm, b_slice = rm * (B_BYTES / gm);
      int a_dn = 0, a_dy = 0, a_dx = 0;
      if (p.mc) {
        a_dn = (rn / (p.mc_cy * p.mc_cx)) * p.mc_sn;
        a_dy = ((rn / p.mc_cx) % p.mc_cy) * p.mc_sy;
        a_dx = (rn % p.mc_cx) * p.mc_sx;
      }
      if (p.b_resident && tile_at(0, m_tile, n_tile)) {
        mbar_arrive_expect_tx(bfull, p.total_kb * B_BYTES);
        for (int kb = 0; kb < p.total_kb; ++kb)
          for (int h = 0; h < gm; ++h)        // (the B map holds a 1 / gm box; the resident slice is not shared)
            tma_load_2d(sB + kb * B_BYTES + h * (B_BYTES / gm), &tmB, bfull, kb * BK, n_tile * BN + h * b_rows);
      }
      // pair: the even CTA's barrier collects both CTAs' bytes (each loads A and half of B)
      const uint32_t stage_tx = CG2 ? 2 * (A_STAGE_BYTES + B_BYTES)
                                      : (p.b_resident ? A_STAGE_BYTES : A_STAGE_BYTES + B_BYTES);
      bool halo_done = false;
      if constexpr (CG2) {
        if (p.halo) {
          halo_done = true;
          int sa = 0, sb = 0;
          uint32_t pha = 0, phb = 0;
          const int nkb = p.seg_kb[0];
          for (int it = 0; tile_at(it, m_tile, n_tile); ++it) {
            const int m_row = (int)fdiv((uint32_t)m_tile, p.fd_tx);
            const int x0 = (m_tile - m_row * p.tiles_x) * p.tw;
            const int y0 = (m_row - (int)fdiv((uint32_t)m_row, p.fd_ty) * p.tiles_y) * p.th;
            const int n0 = (int)fdiv((uint32_t)m_tile, p.fd_tpi);
            for (int dxi = 0; dxi < 3; ++dxi) {
              for (int cb = 0; cb < nkb; ++cb) {
                mbar_wait(&ha_empty[sa], pha ^ 1);
                if (pair_rank == 0) mbar_arrive_expect_tx(&ha_full[sa], 2u * (uint32_t)p.halo_bytes);
                tma_load_4d_cg2(smem + sa * HALO_A_STAGE, &tmA0, &ha_full[sa], cb * BK, x0 + dxi - 1, y0 - 1, n0);
                for (int dyi = 0; dyi < 3; ++dyi) {
                  mbar_wait(&hb_empty[sb], phb ^ 1);
                  if (pair_rank == 0) mbar_arrive_expect_tx(&hb_full[sb], 2u * B_BYTES);
                  tma_load_2d_cg2(sBh + sb * B_BYTES, &tmB, &hb_full[sb], ((dyi * 3 + dxi) * nkb + cb) * BK,
                                  n_tile * BN + pair_rank * (BN / 2));
                  if (++sb == HALO_SB) { sb = 0; phb ^= 1; }
                }
                if (++sa == HALO_SA) { sa = 0; pha ^= 1; }
              }
            }
          }
        }
      }
      for (int it = 0; !halo_done && tile_at(it, m_tile, n_tile); ++it) {
        const int m_row = (int)fdiv((uint32_t)m_tile, p.fd_tx);
        const int x0 = (m_tile - m_row * p.tiles_x) * p.tw + a_dx;
        const int y0 = (m_row - (int)fdiv((uint32_t)m_row, p.fd_ty) * p.tiles_y) * p.th + a_dy;
        const int n0 = (int)fdiv((uint32_t)m_tile, p.fd_tpi) * p.tn + a_dn;
        int kcoord = 0;
        for (int seg = 0; seg < p.nseg; ++seg) {
          const CUtensorMap* tm = seg == 0 ? &tmA0 : (seg == 1 ? &tmA1 : &tmA2);
          const int taps = p.seg_taps[seg];
          const int nkb = p.seg_kb[seg];
          for (int tap = 0; tap < taps; ++tap) {
            const int dy = taps == 9 ? tap / 3 - 1 : (taps == 4 ? p.up_py - 1 + (tap >> 1) : 0);
            const int dx = taps == 9 ? tap % 3 - 1 : (taps == 4 ? p.up_px - 1 + (tap & 1) : 0);
            for (int cb = 0; cb < nkb; ++cb) {
              mbar_wait(&empty[stage], phase ^ 1);
              if constexpr (CG2) {
                if (pair_rank == 0) mbar_arrive_expect_tx(&full[stage], stage_tx);
                tma_load_4d_cg2(sA + stage * A_STAGE_BYTES, tm, &full[stage], cb * BK, x0 * p.stride + dx, y0 * p.stride + dy, n0);
                tma_load_2d_cg2(sB + stage * B_BYTES, &tmB, &full[stage], kcoord, n_tile * BN + pair_rank * (BN / 2));
                kcoord += BK;
                if (++stage == STAGES) {
                  stage = 0;
                  phase ^= 1;
                }
                continue;
              }
              mbar_arrive_expect_tx(&full[stage], stage_tx);
              if (p.mc)
                tma_load_4d_mc(sA + stage * A_STAGE_BYTES + a_slice, tm, &full[stage], cb * BK, x0 * p.stride + dx,
                               y0 * p.stride + dy, n0, a_mask);
              else
                tma_load_4d(sA + stage * A_STAGE_BYTES, tm, &full[stage], cb * BK, x0 * p.stride + dx, y0 * p.stride + dy, n0);
              if (!p.b_resident) {
                if (gm == 2)                                   // (streaming only: shared with the CTA on the same n-tile)
                  tma_load_2d_mc(sB + stage * B_BYTES + b_slice, &tmB, &full[stage], kcoord, n_tile * BN + rm * b_rows, b_mask);
                else
                  tma_load_2d(sB + stage * B_BYTES, &tmB, &full[stage], kcoord, n_tile * BN);
              }
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
    if (elect_one() && pair_rank == 0) {
      // ===================== MMA issuer (in pair mode: the even CTA, for both SMs) =====================
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      constexpr uint32_t idesc2 = umma_idesc_bf16(2 * BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t iter = 0;
      int m_tile, n_tile;
      uint16_t mc_mask = 1;
      if (p.mc) {
        const int rank = (int)cluster_ctarank(), gn = p.mc_gn, gm = p.mc / p.mc_gn;
        const int rn = rank % gn, rm = rank / gn;
        mc_mask = (uint16_t)((((1u << gn) - 1u) << (rm * gn)) | (((gm == 2 && !p.b_resident) ? (1u | (1u << gn)) : 1u) << rn));
      }
      bool halo_done = false;
      if constexpr (CG2) {
        if (p.halo) {
          halo_done = true;
          int sa = 0, sb = 0;
          uint32_t pha = 0, phb = 0;
          const int nkb = p.seg_kb[0];
          for (int it = 0; tile_at(it, m_tile, n_tile); ++it, ++iter) {
            const uint32_t as = iter & 1;
            const uint32_t aph = (iter >> 1) & 1;
            mbar_wait(&tempty[as], aph ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + as * BN;
            uint32_t acc = 0;
            for (int dxi = 0; dxi < 3; ++dxi) {
              for (int cb = 0; cb < nkb; ++cb) {
                mbar_wait(&ha_full[sa], pha);
                const uint32_t a_base = smem_u32(smem + sa * HALO_A_STAGE);
                for (int dyi = 0; dyi < 3; ++dyi) {
                  mbar_wait(&hb_full[sb], phb);
                  tc_fence_after();
                  const uint32_t a_view = a_base + (uint32_t)(dyi * p.halo_row_bytes);   // image rows y0 - 1 + dyi ...
                  const uint32_t b_base = smem_u32(sBh + sb * B_BYTES);
#pragma unroll
                  for (int k = 0; k < BK / 16; ++k) {
                    umma_bf16_cg2(d_tmem, umma_desc_sw128(a_view + k * 32), umma_desc_sw128(b_base + k * 32), idesc2, acc);
                    acc = 1;
                  }
                  umma_commit_cg2_mc(&hb_empty[sb], (uint16_t)3);
                  if (++sb == HALO_SB) { sb = 0; phb ^= 1; }
                }
                umma_commit_cg2_mc(&ha_empty[sa], (uint16_t)3);
                if (++sa == HALO_SA) { sa = 0; pha ^= 1; }
              }
            }
            umma_commit_cg2_mc(&tfull[as], (uint16_t)3);
          }
        }
      }
      for (int it = 0; !halo_done && tile_at(it, m_tile, n_tile); ++it, ++iter) {
        const uint32_t as = iter & 1;
        const uint32_t aph = (iter >> 1) & 1;
        if (p.b_resident && it == 0) mbar_wait(bfull, 0);
        mbar_wait(&tempty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < p.total_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + stage * A_STAGE_BYTES);
          const uint32_t b_base = smem_u32(sB + (p.b_resident ? kb : stage) * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if constexpr (CG2)
              umma_bf16_cg2(d_tmem, umma_desc_sw128(a_base + k * 32), umma_desc_sw128(b_base + k * 32), idesc2,
                            (kb | k) != 0 ? 1u : 0u);
            else
              umma_bf16(d_tmem, umma_desc_sw128(a_base + k * 32), umma_desc_sw128(b_base + k * 32), idesc,
                        (kb | k) != 0 ? 1u : 0u);
          }
          if constexpr (CG2) umma_commit_cg2_mc(&empty[stage], (uint16_t)3);   // both CTAs' producers
          else if (p.mc) umma_commit_mc(&empty[stage], mc_mask);   // ... in this CTA and in every CTA that also writes it
          else umma_commit(&empty[stage]);  // smem slot reusable once these MMAs retire
          if (kb == p.total_kb - 1) {
            if constexpr (CG2) umma_commit_cg2_mc(&tfull[as], (uint16_t)3);      // both CTAs' epilogues
            else umma_commit(&tfull[as]);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 152;");
    // ===================== epilogue (see epi_tile) =====================
    constexpr int NEW = Cfg::EPI_WARPS;
    const int ew = warp - 4;
    const int qw = ew & 3;                 // TMEM lane quadrant this warp may read
    const int cpar = ew >> 2;              // column share 0..2 of this warp inside its quadrant
    const int row = qw * 32 + lane;
    const int xl = row % p.tw;
    const int yl = (row / p.tw) % p.th;
    const int nl = row / (p.tw * p.th);
    const uint32_t stg = smem_u32(smem + Cfg::STAGING_OFF) + ew * Cfg::STG_BYTES;
    // chan_stats rides on the general path (its coalesced phase sees fp32 values of whole rows)
    const bool has_add = p.residual != nullptr || p.rowbias != nullptr || p.chan_stats != nullptr;
    uint32_t iter = 0;
    int m_tile, n_tile;
    // GEGLU: the bias slice of this warp's unit stays in registers while the n-tile does not change (weight-resident
    // schedule: the whole kernel) — a global load at the top of every unit sat on the critical path of the epilogue
    [[maybe_unused]] float g_blv = 0.f, g_blg = 0.f;
    [[maybe_unused]] int g_bias_key = -1;
    for (int it = 0; tile_at(it, m_tile, n_tile); ++it, ++iter) {
      const uint32_t as = iter & 1;
      const uint32_t aph = (iter >> 1) & 1;
      const int m_row = (int)fdiv((uint32_t)m_tile, p.fd_tx);             // m_tile / tiles_x
      const int m_img = (int)fdiv((uint32_t)m_tile, p.fd_tpi);            // m_tile / (tiles_x * tiles_y)
      const int x = (m_tile - m_row * p.tiles_x) * p.tw + xl;
      const int y = (m_row - (int)fdiv((uint32_t)m_row, p.fd_ty) * p.tiles_y) * p.th + yl;
      const int n = m_img * p.tn + nl;
      const bool valid = (x < p.W) && (y < p.H) && (n < p.n_img);
      const long long pix_own = !valid ? -1
                                : (p.up ? ((long long)n * 2 * p.H + 2 * y + p.up_py) * (2 * p.W) + 2 * x + p.up_px
                                        : ((long long)n * p.H + y) * p.W + x);

      const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16) + as * BN;
      // statistics slot of this warp's 32 rows: (image, [phase,] tile inside the image, lane quadrant); tn == 1
      long long stat_slot = -1;
      if (p.chan_stats) {
        const int img = m_img;
        const int in_img = (p.up ? (p.up_py * 2 + p.up_px) * tiles_per_img_grid : 0) + (m_tile - m_img * tiles_per_img_grid);
        stat_slot = (long long)img * p.stat_slots + in_img * 4 + qw;
      }

      if constexpr (EPI >= EPI_FAST) {
        constexpr int V = EPI - EPI_FAST;
#ifdef MDX_EPI_L2PF
        if ((V % 6) / 2 != 0 && cpar == 0) {   // next tile's residual rows on their way from HBM to L2
          int m2, n2;
          if (tile_at(it + 1, m2, n2)) {
            const int x2 = (m2 % p.tiles_x) * p.tw + xl;
            const int y2 = ((m2 / p.tiles_x) % p.tiles_y) * p.th + yl;
            const int i2 = (m2 / tiles_per_img_grid) * p.tn + nl;
            if (x2 < p.W && y2 < p.H && i2 < p.n_img) {
              const long long pix2 = p.up ? ((long long)i2 * 2 * p.H + 2 * y2 + p.up_py) * (2 * p.W) + 2 * x2 + p.up_px
                                          : ((long long)i2 * p.H + y2) * p.W + x2;
              const int col2 = n2 * BN;
              const int ncols = (p.N - col2) < BN ? (p.N - col2) : BN;
              const int esz = p.res_f32 ? 4 : 2;
              const char* src = reinterpret_cast<const char*>(p.residual) + (pix2 * p.ldr + col2) * esz;
              asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"((uint32_t)(ncols * esz)) : "memory");
            }
          }
        }
#endif
        // (tn == 1 whenever a per-image bias takes this path: the tile's 128 rows belong to one image)
        const float* rb_row = p.rowbias ? p.rowbias + (long long)m_img * p.ld_rowbias : nullptr;
        epi_fast<BN, (V / 6) != 0, (V % 6) / 2, (V % 2) != 0>(p, taddr, stg, lane, cpar, n_tile, pix_own, stat_slot, rb_row,
                                                             &tfull[as], aph);
      } else if constexpr (EPI == EPI_GEGLU) {
        mbar_wait(&tfull[as], aph);
        tc_fence_after();
        // GEGLU: the tile holds [BN/2 value columns | the matching BN/2 gate columns].  In the row-owner layout a
        // thread forms (value + bv) * gelu(gate + bg) for up to 32 output columns of its row, packs bf16, and the
        // unit goes through the 80-byte-row staging tile to 16-byte coalesced stores (as the plain bf16 path).
        constexpr int HALF = BN / 2;
        constexpr int NU = (HALF + 31) / 32;           // units of 32 output columns (the last one may hold 16)
        const int n_out = p.N / 2;
        const int ocol_tile = n_tile * HALF;
        const int sr8 = lane >> 2, q4 = lane & 3;
        long long pix8[4];
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) pix8[pass] = __shfl_sync(0xffffffffu, pix_own, pass * 8 + sr8);
#pragma unroll 1
        for (int u = cpar; u < NU; u += (NEW / 4)) {
          const int ucol = u * 32;
          const int uw = (HALF - ucol) < 32 ? (HALF - ucol) : 32;       // 32 or 16
          const int ocol0 = ocol_tile + ucol;
          if (ocol0 >= n_out) break;
          const float* bvp = p.bias + n_tile * BN + ucol;   // bias is packed like the weight rows
          const int bias_key = n_tile * NU + u;
          if (bias_key != g_bias_key) {                     // lane <-> column; broadcast through smem below
            g_bias_key = bias_key;
            g_blv = 0.f;
            g_blg = 0.f;
            if (p.bias && lane < uw) {
              g_blv = __ldg(bvp + lane);
              g_blg = __ldg(bvp + HALF + lane);
            }
          }
          const float blv = g_blv, blg = g_blg;
          uint32_t rv[32], rg[32];
          if (uw == 32) {
            tmem_ld32(taddr + ucol, rv);
            tmem_ld32(taddr + HALF + ucol, rg);
          } else {
            tmem_ld16(taddr + ucol, *reinterpret_cast<uint32_t(*)[16]>(&rv[0]));
            tmem_ld16(taddr + HALF + ucol, *reinterpret_cast<uint32_t(*)[16]>(&rg[0]));
          }
          __syncwarp();  // the previous unit's reads of the staging tile and the bias slice are done
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(stg + 2560 + lane * 4), "r"(__float_as_uint(blv)) : "memory");
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(stg + 2688 + lane * 4), "r"(__float_as_uint(blg)) : "memory");
          __syncwarp();
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (4 * j < uw) {
              const uint4 ubv = lds128(stg + 2560 + 16 * j), ubg = lds128(stg + 2688 + 16 * j);
              const float4 bv = make_float4(__uint_as_float(ubv.x), __uint_as_float(ubv.y), __uint_as_float(ubv.z), __uint_as_float(ubv.w));
              const float4 bg = make_float4(__uint_as_float(ubg.x), __uint_as_float(ubg.y), __uint_as_float(ubg.z), __uint_as_float(ubg.w));
              float o0, o1, o2, o3;
              geglu_pair(o0, o1, __uint_as_float(rv[4 * j]), __uint_as_float(rv[4 * j + 1]), __uint_as_float(rg[4 * j]),
                         __uint_as_float(rg[4 * j + 1]), bv.x, bv.y, bg.x, bg.y);
              geglu_pair(o2, o3, __uint_as_float(rv[4 * j + 2]), __uint_as_float(rv[4 * j + 3]),
                         __uint_as_float(rg[4 * j + 2]), __uint_as_float(rg[4 * j + 3]), bv.z, bv.w, bg.z, bg.w);
              pk[2 * j] = pack_bf16x2(o0, o1);
              pk[2 * j + 1] = pack_bf16x2(o2, o3);
            } else {
              pk[2 * j] = 0u;
              pk[2 * j + 1] = 0u;
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) sts128(stg + lane * 80 + 16 * j, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          __syncwarp();
          const bool cok = q4 * 8 < uw && ocol0 + q4 * 8 < n_out;
          bf16* obase = reinterpret_cast<bf16*>(p.out) + ocol0 + q4 * 8;
#pragma unroll
          for (int pass = 0; pass < 4; ++pass) {
            const uint4 w = lds128(stg + (pass * 8 + sr8) * 80 + q4 * 16);
            const long long pix = pix8[pass];
            if (cok && pix >= 0 && !(p.debug & 1)) *reinterpret_cast<uint4*>(obase + pix * p.ldc) = w;
          }
        }
      } else if (p.debug & 2) {
        mbar_wait(&tfull[as], aph);  // profiling only: drop the tile
      } else if (p.out_f32) {
        if (has_add) epi_tile<BN, true, true>(p, taddr, stg, lane, cpar, NEW / 4, n_tile, pix_own, stat_slot, &tfull[as], aph);
        else epi_tile<BN, true, false>(p, taddr, stg, lane, cpar, NEW / 4, n_tile, pix_own, stat_slot, &tfull[as], aph);
      } else {
        if (has_add) epi_tile<BN, false, true>(p, taddr, stg, lane, cpar, NEW / 4, n_tile, pix_own, stat_slot, &tfull[as], aph);
        else epi_tile<BN, false, false>(p, taddr, stg, lane, cpar, NEW / 4, n_tile, pix_own, stat_slot, &tfull[as], aph);
      }
      tc_fence_before();
      if constexpr (CG2) {                      // the even CTA's MMA thread waits for both CTAs' epilogues
        __syncwarp();
        if (lane == 0) {
          if (pair_rank == 0) mbar_arrive(&tempty[as]);
          else mbar_arrive_remote(&tempty[as], 0);
        }
      } else {
        mbar_arrive(&tempty[as]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CG2 || p.mc) cluster_sync_all();    // no CTA leaves while its peer may still signal or write into it
  if (warp == 2) {
    tc_fence_after();
    if constexpr (CG2) tmem_dealloc_cg2(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
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

// W, H: the INPUT image; stride 2 makes the box step over every other pixel (TMA element strides), so that tap (r, s)
// of a stride-2 convolution is one box load starting at (2 x0 + s - 1, 2 y0 + r - 1) — no im2col buffer
int make_map_a(CUtensorMap* m, const void* ptr, long long ld, int C, int W, int H, int n, int tw, int th, int tn,
               int stride = 1) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    mdx_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return MDX_ERR_CUDA;
  }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)W * ld * 2, (cuuint64_t)H * W * ld * 2};
  cuuint32_t box[4] = {(cuuint32_t)BK, (cuuint32_t)(tw * stride), (cuuint32_t)(th * stride), (cuuint32_t)tn};
  cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
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

int num_sms() { return mdx_num_sms(); }

template <int BN, int EPI, bool CG2>
int launch_tc_kernel(const CUtensorMap* maps, const CUtensorMap& mb, const TcParams& q, int grid, cudaStream_t st) {
  static MdxPerDeviceOnce attr;
  if (attr.need()) {
    MDX_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, CG2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        TcCfg<BN>::SMEM_BYTES));
    attr.mark();
  }
  if (CG2 || q.mc) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(TcCfg<BN>::THREADS);
    cfg.dynamicSmemBytes = TcCfg<BN>::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CG2 ? 2u : (unsigned)q.mc; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    MDX_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, EPI, CG2>, maps[0], maps[1], maps[2], mb, q));
  } else {
    gemm_tc_kernel<BN, EPI, CG2><<<grid, TcCfg<BN>::THREADS, TcCfg<BN>::SMEM_BYTES, st>>>(maps[0], maps[1], maps[2], mb, q);
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}

template <int BN, int EPI>
int launch_tc_epi(const CUtensorMap* maps, const CUtensorMap& mb, const TcParams& q, int grid, cudaStream_t st) {
  if (q.cg2) return launch_tc_kernel<BN, EPI, true>(maps, mb, q, grid, st);
  return launch_tc_kernel<BN, EPI, false>(maps, mb, q, grid, st);
}

// How many CTAs of this block shape the device keeps resident when they are launched as clusters of `cs` (a GPC whose
// free SMs are not a multiple of the cluster size leaves some without partners); 0 = cluster launch not possible.
// Cached per device and size.
template <int BN>
int max_clustered_ctas(int cs) {
  static int cached[64][2];
  static MdxPerDeviceOnce once;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || (cs != 2 && cs != 4)) return 0;
  if (once.need()) {
    for (int i = 0; i < 2; ++i) {
      const int c = i == 0 ? 2 : 4;
      int n = 0;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)(c * num_sms()));
      cfg.blockDim = dim3(TcCfg<BN>::THREADS);
      cfg.dynamicSmemBytes = TcCfg<BN>::SMEM_BYTES;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = (unsigned)c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      if (cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI_GENERIC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               TcCfg<BN>::SMEM_BYTES) != cudaSuccess ||
          cudaOccupancyMaxActiveClusters(&n, gemm_tc_kernel<BN, EPI_GENERIC>, &cfg) != cudaSuccess) {
        cudaGetLastError();
        n = 0;
      }
      cached[dev][i] = c * n;
      if (getenv("MDX_GEMM_MC_VERBOSE"))
        fprintf(stderr, "mdx_gemm_tc: BN=%d: %d CTAs resident in clusters of %d (%d SMs)\n", BN, c * n, c, num_sms());
    }
    once.mark();
  }
  return cached[dev][cs == 2 ? 0 : 1];
}

template <int BN>
int launch_tc(const CUtensorMap* maps, const CUtensorMap& mb, const TcParams& p, cudaStream_t st) {
  int grid = p.num_tiles < num_sms() ? p.num_tiles : num_sms();
  if (p.mc && grid > p.mc_grid_cap) grid = p.mc_grid_cap;
  if (p.mc) grid -= grid % p.mc;            // (num_tiles is a multiple of the cluster size, see mdx_gemm_tc)
  if (p.cg2) { if (grid > p.mc_grid_cap) grid = p.mc_grid_cap; grid -= grid & 1; }
  TcParams q = p;
  if (const char* dbg = getenv("MDX_GEMM_DEBUG")) q.debug = atoi(dbg);
  const bool no_resident = getenv("MDX_GEMM_NO_RESIDENT") != nullptr;  // A/B switch for tests and profiling
  // weights stay in smem when the whole K fits the B ring and every CTA would otherwise re-fetch them >= 3 times
  if (!no_resident && !p.cg2 && p.nseg == 1 && p.seg_taps[0] == 1 && p.total_kb <= TcCfg<BN>::STAGES &&
      p.n_tiles_n <= (p.mc ? p.mc_grid_cap : num_sms()) && (long long)p.m_tiles * p.n_tiles_n >= 3ll * num_sms()) {
    q.b_resident = 1;
    const int avail = p.mc && p.mc_grid_cap < num_sms() ? p.mc_grid_cap : num_sms();
    int unit = p.n_tiles_n;                  // whole rows of n-tiles, and whole clusters
    while (p.mc && unit % p.mc) unit += p.n_tiles_n;
    grid = (avail / unit) * unit;
    if (grid == 0) { q.b_resident = 0; grid = p.num_tiles < avail ? p.num_tiles : avail; grid -= p.mc ? grid % p.mc : 0; }
  }
  if (q.cg2 && (grid <= 0 || (grid & 1))) {
    mdx_set_error("mdx_gemm_tc: internal: grid %d does not divide into CTA pairs", grid);
    return MDX_ERR_INVALID;
  }
  if (q.mc && (grid <= 0 || grid % q.mc)) {
    mdx_set_error("mdx_gemm_tc: internal: grid %d does not divide into clusters of %d", grid, q.mc);
    return MDX_ERR_INVALID;
  }
  if (q.geglu) return launch_tc_epi<BN, EPI_GEGLU>(maps, mb, q, grid, st);
  // epi_fast: no activation, no profiling switch, a per-image bias only when a tile never straddles two images, row
  // offsets that fit 32 bits; the plain bf16 store (no add, no statistics) stays on epi_tile's 32-column path
  const bool has_add = q.residual != nullptr || q.rowbias != nullptr || q.chan_stats != nullptr;
  const long long rows_out = (long long)q.n_img * q.hw_out;
  const bool fast_ok = q.act == 0 && q.debug == 0 && (q.rowbias == nullptr || q.tn == 1) && (q.out_f32 || has_add) &&
                       rows_out * q.ldc < (1ll << 32) && (q.residual == nullptr || rows_out * q.ldr < (1ll << 32)) &&
                       getenv("MDX_GEMM_NO_FAST_EPI") == nullptr;
  if (!fast_ok) return launch_tc_epi<BN, EPI_GENERIC>(maps, mb, q, grid, st);
  const int v = (q.out_f32 ? 6 : 0) + (q.residual ? (q.res_f32 ? 1 : 2) : 0) * 2 + (q.chan_stats ? 1 : 0);
  switch (v) {
#define MDX_TC_CASE(V) case V: return launch_tc_epi<BN, EPI_FAST + V>(maps, mb, q, grid, st);
    MDX_TC_CASE(1) MDX_TC_CASE(2) MDX_TC_CASE(3) MDX_TC_CASE(4) MDX_TC_CASE(5) MDX_TC_CASE(6) MDX_TC_CASE(7)
    MDX_TC_CASE(8) MDX_TC_CASE(9) MDX_TC_CASE(10) MDX_TC_CASE(11)
#undef MDX_TC_CASE
    default: return launch_tc_epi<BN, EPI_GENERIC>(maps, mb, q, grid, st);
  }
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
  if (N % 192 == 0) return 192;   // GLIDE's channel counts (192, 384, 576, 768): one to four full tiles
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

extern "C" int mdx_gemm_tc_stat_slots(int32_t n_img, int32_t H, int32_t W, int32_t up) {
  if (n_img <= 0 || H <= 0 || W <= 0) return 0;
  int tw, th, tn;
  pick_tile(W, H, n_img, &tw, &th, &tn);
  if (tn != 1) return 0;
  return ((W + tw - 1) / tw) * ((H + th - 1) / th) * 4 * (up ? 4 : 1);
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
  if (a->geglu)
    MDX_CHECK_ARG(!a->rowbias && !a->residual && !a->act && a->N % 32 == 0 && a->out_dtype == MDX_BF16,
                  "mdx_gemm_tc: geglu excludes rowbias/residual/act and writes bf16");
  MDX_CHECK_ARG(a->act >= 0 && a->act <= 3, "mdx_gemm_tc: act must be 0, 1 (gelu), 2 (quick gelu) or 3 (tanh)");
  if (a->act) MDX_CHECK_ARG(!a->rowbias, "mdx_gemm_tc: act excludes rowbias");

  TcParams p;
  memset(&p, 0, sizeof(p));
  p.nseg = a->nseg;
  long long ktot = 0;
  for (int s = 0; s < a->nseg; ++s) {
    MDX_CHECK_ARG(a->a[s] != nullptr, "mdx_gemm_tc: A segment %d is NULL", s);
    MDX_CHECK_ARG(a->seg_c[s] > 0 && a->seg_c[s] % 64 == 0, "mdx_gemm_tc: seg_c[%d]=%d must be a multiple of 64", s,
                  a->seg_c[s]);
    MDX_CHECK_ARG(a->seg_taps[s] == 1 || a->seg_taps[s] == 9 || (a->seg_taps[s] == 4 && a->up_phase),
                  "mdx_gemm_tc: seg_taps[%d] must be 1 or 9 (or 4 with up_phase)", s);
    MDX_CHECK_ARG((reinterpret_cast<uintptr_t>(a->a[s]) & 15) == 0 && a->lda[s] % 8 == 0 && a->lda[s] >= a->seg_c[s],
                  "mdx_gemm_tc: A segment %d must be 16-byte aligned with lda %% 8 == 0", s);
    p.seg_kb[s] = a->seg_c[s] / 64;
    p.seg_taps[s] = a->seg_taps[s];
    p.total_kb += p.seg_kb[s] * p.seg_taps[s];
    ktot += (long long)a->seg_taps[s] * a->seg_c[s];
  }
  int bn = a->block_n ? a->block_n : mdx_gemm_tc_pick_block_n(a->N, a->geglu);
  MDX_CHECK_ARG(bn == 64 || bn == 128 || bn == 160 || bn == 192 || bn == 256, "mdx_gemm_tc: unsupported block_n %d", bn);
  if (a->geglu) MDX_CHECK_ARG((a->N / 2) % (bn / 2) == 0, "mdx_gemm_tc: geglu needs N/2 %% (block_n/2) == 0");

  const int stride = a->stride == 2 ? 2 : 1;
  MDX_CHECK_ARG(a->stride == 0 || a->stride == 1 || a->stride == 2, "mdx_gemm_tc: stride must be 1 or 2");
  if (stride == 2)
    MDX_CHECK_ARG(a->nseg == 1 && a->seg_taps[0] == 9 && !a->up_phase && !a->geglu,
                  "mdx_gemm_tc: stride 2 needs one 3x3 segment (no up_phase, no geglu)");
  // the image the tiles, the output rows and the statistics slots live in: the conv's output
  const int OH = stride == 2 ? (a->H - 1) / 2 + 1 : a->H, OW = stride == 2 ? (a->W - 1) / 2 + 1 : a->W;
  p.stride = stride;
  p.n_img = a->n_img; p.H = OH; p.W = OW;
  pick_tile(OW, OH, a->n_img, &p.tw, &p.th, &p.tn);
  p.tiles_x = (OW + p.tw - 1) / p.tw;
  p.tiles_y = (OH + p.th - 1) / p.th;
  p.fd_tx = make_fastdiv((uint32_t)p.tiles_x);
  p.fd_ty = make_fastdiv((uint32_t)p.tiles_y);
  p.fd_tpi = make_fastdiv((uint32_t)(p.tiles_x * p.tiles_y));
  int tiles_n = (a->n_img + p.tn - 1) / p.tn;
  p.N = a->N;
  p.n_tiles_n = (a->N + bn - 1) / bn;
  p.fd_ntn = make_fastdiv((uint32_t)p.n_tiles_n);
  long long nt = (long long)p.tiles_x * p.tiles_y * tiles_n * p.n_tiles_n;
  MDX_CHECK_ARG(nt < (1ll << 31), "mdx_gemm_tc: too many tiles");
  p.num_tiles = (int)nt;
  p.m_tiles = p.tiles_x * p.tiles_y * tiles_n;
  if (a->chan_stats) {
    MDX_CHECK_ARG(p.tn == 1, "mdx_gemm_tc: chan_stats needs images of at least 128 pixels (got %d x %d)", OH, OW);
    p.stat_slots = p.tiles_x * p.tiles_y * 4 * (a->up_phase ? 4 : 1);
  }
  p.out = a->out; p.ldc = a->ldc; p.out_f32 = a->out_dtype == MDX_F32;
  p.bias = a->bias; p.rowbias = a->rowbias; p.ld_rowbias = a->ld_rowbias;
  p.residual = a->residual; p.ldr = a->ldr; p.res_f32 = a->residual_dtype == MDX_F32;
  p.geglu = a->geglu;
  p.act = a->act;
  p.chan_stats = a->chan_stats;
  if (a->chan_stats)
    MDX_CHECK_ARG(!a->geglu && (reinterpret_cast<uintptr_t>(a->chan_stats) & 15) == 0,
                  "mdx_gemm_tc: chan_stats excludes geglu and must be 16-byte aligned");
  // (the slot count is checked below, once the tile shape is known)
  p.hw_out = (long long)OH * OW;
  if (a->up_phase) {
    MDX_CHECK_ARG(a->up_phase >= 1 && a->up_phase <= 4 && a->nseg == 1 && a->seg_taps[0] == 4 && !a->geglu,
                  "mdx_gemm_tc: up_phase needs 1..4, one segment with 4 taps, no geglu");
    p.up = 1;
    p.up_py = (a->up_phase - 1) >> 1;
    p.up_px = (a->up_phase - 1) & 1;
    p.hw_out *= 4;
  }

  // Cluster launch: the persistent schedule puts consecutive blocks on consecutive tiles, n fastest, so the CTAs of a
  // cluster of 2 or 4 work on neighbouring n-tiles of the same m-tile (and, when there are only two n-tiles, on two
  // m-tiles) at the same time.  Each A block is then fetched once per m-tile and each B block once per n-tile of the
  // cluster — every CTA loads its share of the box and multicasts it — which takes 22 % (pairs) to 50 % (2 x 2) off
  // the L2 -> SM traffic the large convolutions are bound by (ncu: L2 76 % busy next to a tensor pipe 83 % busy).
  // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
  // hung the GPU: every operand must be an ordinary allocation of the current device, anything else takes the
  // plain launch.
  int btw = p.tw, bth = p.th, btn = p.tn, b_box = bn;
  {
    // Default: convolutions only (any segment with taps) — alternating in-process A/B in the power-capped steady state
    // (tools/gemm_mc_ab.py, profiles/r02_gemm_cluster_ab.txt): 3x3 convs -3 .. -9 %, 1x1 / dense layers 0 .. +2.5 %.
    // MDX_GEMM_MC = 0 / 2 / 4 forces the largest cluster tried for every eligible launch (tests and A/B runs).
    const char* mc_env = getenv("MDX_GEMM_MC");
    bool has_taps = false;
    for (int s = 0; s < a->nseg; ++s) has_taps = has_taps || a->seg_taps[s] > 1;
    const int mc_max = mc_env ? atoi(mc_env) : (has_taps ? 4 : 0);
    int cs = 0, gn = 1;
    if (mc_max >= 4 && p.n_tiles_n % 4 == 0 && p.num_tiles % 4 == 0) { cs = 4; gn = 4; }
    else if (mc_max >= 4 && p.n_tiles_n == 2 && p.num_tiles % 4 == 0) { cs = 4; gn = 2; }
    else if (mc_max >= 2 && p.n_tiles_n % 2 == 0) { cs = 2; gn = 2; }
    int cur = -1;
    bool local_ok = cudaGetDevice(&cur) == cudaSuccess;
    auto local = [&](const void* ptr) {
      cudaPointerAttributes pa;
      if (cudaPointerGetAttributes(&pa, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
      return pa.type == cudaMemoryTypeDevice && pa.device == cur;
    };
    for (int s = 0; local_ok && s < a->nseg; ++s) local_ok = local(a->a[s]);
    if (local_ok) local_ok = local(a->b);
    auto cap_of = [&](int c) {
      switch (bn) {
        case 64: return max_clustered_ctas<64>(c);
        case 128: return max_clustered_ctas<128>(c);
        case 160: return max_clustered_ctas<160>(c);
        case 192: return max_clustered_ctas<192>(c);
        default: return max_clustered_ctas<256>(c);
      }
    };
    // CTA pairs (tcgen05 cta_group::2) take precedence for convolutions with an even number of m-tiles: two consecutive
    // m-tiles of one n-tile per cluster, one M = 256 MMA over both SMs, the B block halved between the two CTAs.
    // Alternating in-process A/B (profiles/r02_gemm_cluster_ab.txt): 3x3 convs -11 .. -17 % against the plain launch
    // (multicast pairs -2 .. -9 %), 1x1 / dense / GEGLU layers +6 .. +37 % — those never take it.  MDX_GEMM_CG2=0 turns
    // it off, =1 forces it for every eligible launch; MDX_GEMM_MC=0 turns every cluster form off.
    const char* cg2_env = getenv("MDX_GEMM_CG2");
    const bool want_cg2 = cg2_env ? atoi(cg2_env) == 1 : (has_taps && !a->geglu && !(mc_env && atoi(mc_env) == 0));
    bool ok = local_ok && cs != 0;
    int cap = 0;
    if (local_ok && want_cg2 && p.m_tiles % 2 == 0) {
      const int cap2 = cap_of(2);
      if (cap2 >= 2 && cap2 * 16 >= num_sms() * 15) {
        p.cg2 = 1;
        p.mc_grid_cap = cap2;
        b_box = bn / 2;
        ok = false;                           // not both
      }
    }
    for (; ok && cs >= 2; cs >>= 1) {         // a few SMs without partners are a fair price; far fewer is not
      cap = cap_of(cs);
      if (cap >= cs && cap * 16 >= num_sms() * 15) break;
      gn = 2;                                 // fall back from 4 to pairs along n
    }
    if (ok && cs >= 2) {
      p.mc = cs;
      p.mc_gn = gn;
      p.mc_grid_cap = cap;
      // A sub-box: halve the outermost dimension the tile spans until 1 / gn is left — its pieces are then consecutive
      // row ranges of the tile, numbered in row order
      for (int g = gn; g > 1; g >>= 1) {
        if (btn > 1) btn >>= 1;
        else if (bth > 1) bth >>= 1;
        else btw >>= 1;
      }
      p.mc_sn = btn; p.mc_sy = bth; p.mc_sx = btw;
      p.mc_cy = p.th / bth; p.mc_cx = p.tw / btw;
      if (cs / gn == 2) b_box = bn / 2;       // 2 x 2: the CTAs on the same n-tile fetch half of the B block each
    }
  }
  // Halo boxes (CTA pairs, one plain 3x3 segment, tiles inside one image): the dy taps of a dx share one A box of
  // th + 2 image rows.  Changes the order in which the taps are accumulated (dx outermost), so results differ from the
  // other launch forms in the last fp32 bits.  MDX_GEMM_HALO=0 turns it off.
  if (p.cg2 && a->nseg == 1 && a->seg_taps[0] == 9 && stride == 1 && !a->up_phase && p.tn == 1 && bn == 160 &&
      p.tw >= 8 && (p.th + 2) * p.tw * 128 <= 32768) {
    const char* h = getenv("MDX_GEMM_HALO");
    if (!(h && atoi(h) == 0)) {
      p.halo = 1;
      p.halo_row_bytes = p.tw * 128;
      p.halo_bytes = (p.th + 2) * p.tw * 128;
      bth = p.th + 2;
    }
  }
  CUtensorMap maps[3];
  for (int s = 0; s < 3; ++s) {
    int src = s < a->nseg ? s : 0;
    int rc = make_map_a(&maps[s], a->a[src], a->lda[src], a->seg_c[src], a->W, a->H, a->n_img, btw, bth, btn, stride);
    if (rc) return rc;
  }
  CUtensorMap mb;
  int rc = make_map_b(&mb, a->b, ktot, a->N, b_box);
  if (rc) return rc;

  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (bn) {
    case 64: return launch_tc<64>(maps, mb, p, st);
    case 128: return launch_tc<128>(maps, mb, p, st);
    case 160: return launch_tc<160>(maps, mb, p, st);
    case 192: return launch_tc<192>(maps, mb, p, st);
    default: return launch_tc<256>(maps, mb, p, st);
  }
}
