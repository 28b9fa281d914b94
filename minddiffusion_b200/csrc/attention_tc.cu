// Flash attention on the 5th-generation tensor cores (tcgen05 + TMEM) for head dim 64 — the shape of every
// attention layer of the stablediffusionv2 UNet (num_head_channels: 64, configs/v2-inference.yaml:31).
// Reference semantics: CrossAttention.construct, ldm/modules/attention.py:134-152:  softmax(q k^T * d^-1/2) v.
//
// One CTA = 128 queries of one (batch, head); keys/values stream through in tiles of 128.
//   warp 0 : TMA producer   Q once, then the K tiles (4-D boxes cut straight out of the fused [B, N, 3C] / [B, L, 2C]
//                           projection buffers, 128-byte swizzle), 3-stage mbarrier ring
//   warp 3 : TMA producer   the V tiles, its own ring — a full V ring never holds back a K load
//   warp 1 : MMA issuer     S_j = Q K_j^T   (M128 x N128 x K64, both operands K-major)        -> TMEM S[j%3]
//   warp 2 : MMA issuer     O_j = P_j V_j   (M128 x N64 x K128, A = P from TMEM, B = V MN-major) -> TMEM O[j&1]
//                           (and TMEM allocate / free: 512 columns, 3 x 128 for S, 2 x 64 for O).  Two issuing threads
//                           because one thread's scalar work between two MMA batches left the tensor pipe idle; the
//                           ordering they lose (tcgen05 operations of different threads are unordered) is restored by
//                           s_empty: the score MMA of tile j + 3 waits for P V of tile j to have read its stage.
//   warps 4-11: softmax, two warpgroups ping-ponging over key tiles (group g owns tiles g, g+2, ...).  Thread =
//                           query row (TMEM lane), so row max / row sum need no shuffles.  One pass over S: the 128
//                           scores of the row into registers, max, exp2 + sum (3 of every 16 pairs on the FMA pipe,
//                           the rest on the MUFU), P written back as bf16 pairs over the S columns (tcgen05.st).  The
//                           running output O[g] stays in TMEM and accumulates over the group's tiles; it is rescaled in
//                           place (tcgen05.ld / st) only when a row's max moves by more than 2^8; the two groups'
//                           (max, sum, O) are merged at the end.
// Keys beyond Nk are masked to -inf (K/V rows there are zero-filled by TMA); query rows beyond Nq are not stored.
#include "common.cuh"
#include <atomic>

extern std::atomic<long long> g_mdx_launches;

namespace {

constexpr int AT_BM = 128;   // queries per CTA
constexpr int AT_BN = 128;   // keys per tile
constexpr int AT_D = 64;
constexpr int AT_KS = 3;     // K ring depth
constexpr int AT_VS = 3;     // V ring depth
constexpr int AT_SS = 3;     // S accumulators in TMEM: Q K^T runs two tiles ahead of the softmax
constexpr int TILE_BYTES = 128 * 64 * 2;  // 16 KiB: one [128 x 64] bf16 operand block

struct AttnTcP {
  int Nq, Nk;
  float sl2;  // scale * log2(e)
  bf16* o;
  long long ldo;
  int d;  // head dim <= 64; the operand tiles are zero-padded to 64 columns by TMA
};

// smem map (offsets from the 1024-aligned base)
constexpr int OFF_Q = 0;
constexpr int OFF_K = OFF_Q + TILE_BYTES;
constexpr int OFF_V = OFF_K + AT_KS * TILE_BYTES;
constexpr int OFF_XCH = OFF_V + AT_VS * TILE_BYTES;        // 32 KiB: group 1 parks its O rows here for the final merge
constexpr int OFF_BAR = OFF_XCH + 2 * TILE_BYTES;
constexpr int OFF_XML = OFF_BAR + 512;
constexpr int AT_SMEM = OFF_XML + 1024 + 1024;

__device__ __forceinline__ void tma_load_3d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ float max3f(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
// packed fp32 pairs (sm_100 FFMA2 / FADD2): one issue slot for two lanes of the softmax's scale-subtract and row sum
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b, float c) {
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\tmov.b64 rc, {%5, %5};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
__device__ __forceinline__ void fadd2(float& d0, float& d1, float a0, float a1) {
  asm("{\n\t.reg .b64 ra, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rd, {%0, %1};\n\tadd.rn.f32x2 rd, rd, ra;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "+f"(d0), "+f"(d1)
      : "f"(a0), "f"(a1));
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
      "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void fmul2(float& d0, float& d1, float b) {
  asm("{\n\t.reg .b64 rb, rd;\n\tmov.b64 rb, {%2, %2};\n\tmov.b64 rd, {%0, %1};\n\tmul.rn.f32x2 rd, rd, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "+f"(d0), "+f"(d1)
      : "f"(b));
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 2^x for a pair of lanes on the FMA pipe instead of the MUFU.  At d = 64 the 128 exponentials per score row take
// twice the MUFU cycles the tile's two MMAs take on the tensor pipe, so a share of them is moved: floor(x) by a
// round-down magic add, 2^frac by a degree-3 fit on [0, 1) (relative error 8.8e-5, well under P's bf16 rounding;
// exact at frac = 0, so the row maximum still maps to 1), floor(x) added into the exponent field as an integer.
#ifndef MDX_ATTN_POLY_NUM
#define MDX_ATTN_POLY_NUM 3                                  // pairs out of every MDX_ATTN_POLY_DEN that take this path
#endif
#ifndef MDX_ATTN_POLY_DEN
#define MDX_ATTN_POLY_DEN 16
#endif
__device__ __forceinline__ void ex2_poly2(float& p0, float& p1, float x0, float x1) {
  x0 = fmaxf(x0, -126.f);                                    // masked keys arrive as -inf; 2^-126 is zero to bf16 P
  x1 = fmaxf(x1, -126.f);
  uint32_t t0, t1, q0, q1;
  asm("{\n\t.reg .b64 x, k, t, f, p, c;\n\t"
      "mov.b64 x, {%4, %5};\n\t"
      "mov.b64 k, {%6, %6};\n\t"
      "add.rm.f32x2 t, x, k;\n\t"                            // low mantissa bits of t = floor(x)
      "sub.rn.f32x2 f, t, k;\n\t"
      "sub.rn.f32x2 f, x, f;\n\t"                            // frac in [0, 1)
      "mov.b64 p, {%7, %7};\n\t"
      "mov.b64 c, {%8, %8};\n\t"
      "fma.rn.f32x2 p, p, f, c;\n\t"
      "mov.b64 c, {%9, %9};\n\t"
      "fma.rn.f32x2 p, p, f, c;\n\t"
      "mov.b64 c, {%10, %10};\n\t"
      "fma.rn.f32x2 p, p, f, c;\n\t"
      "mov.b64 {%0, %1}, t;\n\t"
      "mov.b64 {%2, %3}, p;\n\t}"
      : "=r"(t0), "=r"(t1), "=r"(q0), "=r"(q1)
      : "f"(x0), "f"(x1), "f"(12582912.f), "f"(0.077119089663028717041015625f), "f"(0.227564394474029541015625f),
        "f"(0.695146143436431884765625f), "f"(1.f));
  p0 = __uint_as_float(q0 + (t0 << 23));
  p1 = __uint_as_float(q1 + (t1 << 23));
}
#ifdef MDX_ATTN_TRACE
// debug build only: CTA 0 of the persistent kernel stamps (event, tile, clock) per role into a device array that
// tools/attn_trace.py reads back through mdx_attn_trace_dump
constexpr int AT_TR_N = 4096;                                // records per role
__device__ uint32_t at_trace[4 * AT_TR_N * 2];
#define AT_TR(role, ev, tile)                                                              \
  do {                                                                                     \
    if (blockIdx.x == 0 && tr_n < AT_TR_N) {                                               \
      at_trace[((role) * AT_TR_N + tr_n) * 2] = ((uint32_t)(ev) << 24) | (uint32_t)(tile); \
      at_trace[((role) * AT_TR_N + tr_n) * 2 + 1] = (uint32_t)clock();                     \
      ++tr_n;                                                                              \
    }                                                                                      \
  } while (0)
#else
#define AT_TR(role, ev, tile) do { } while (0)
#endif
// exp2 of the pair with index k of a score row: the MUFU for most pairs, the FMA-pipe form for the chosen share
__device__ __forceinline__ void ex2_pair(int k, float& p0, float& p1, float x0, float x1) {
  if ((k % MDX_ATTN_POLY_DEN) < MDX_ATTN_POLY_NUM) {         // neighbouring pairs: spreading them evenly measured 1.5 % slower
    ex2_poly2(p0, p1, x0, x1);
  } else {
    p0 = ex2_approx(x0);
    p1 = ex2_approx(x1);
  }
}
// MN-major operand (V as B of the P V product): rows of 64 elements (128 B) along N, 8-row swizzle atoms along K,
// 1024 B between atoms (SBO).  Same bit layout as umma_desc_sw128; LBO is unused because N spans one atom.
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr) { return umma_desc_sw128(smem_addr); }
// instruction descriptor: bf16 x bf16 -> fp32, A K-major, B MN-major (bit 16)
__host__ __device__ constexpr uint32_t umma_idesc_bf16_bmn(int M, int N) { return umma_idesc_bf16(M, N) | (1u << 16); }

__global__ void __launch_bounds__(384, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const AttnTcP p) {
  extern __shared__ uint8_t at_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars;                 // 1
  uint64_t* k_full = bars + 1;             // AT_KS
  uint64_t* k_empty = k_full + AT_KS;      // AT_KS
  uint64_t* v_full = k_empty + AT_KS;      // AT_VS
  uint64_t* v_empty = v_full + AT_VS;      // AT_VS
  uint64_t* s_full = v_empty + AT_VS;      // AT_SS
  uint64_t* p_full = s_full + AT_SS;       // AT_SS
  uint64_t* o_full = p_full + AT_SS;       // 2
  uint64_t* o_empty = o_full + 2;          // 2
  uint64_t* s_empty = o_empty + 2;         // AT_SS: P V of the stage has completed, the next score MMA may overwrite it
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + AT_SS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * AT_BM;
  const int head = blockIdx.y, b = blockIdx.z;
  const int T = (p.Nk + AT_BN - 1) / AT_BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < AT_KS; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < AT_VS; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < AT_SS; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 128); }
    for (int i = 0; i < 2; ++i) mbar_init(&o_full[i], 1);
    for (int i = 0; i < AT_SS; ++i) mbar_init(&s_empty[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t TM_S = tmem_base;                    // S[sb] at columns sb*128, sb = tile % 3
  const uint32_t TM_O = tmem_base + AT_SS * AT_BN;    // O[ob] at columns 384 + ob*64

  // Register rebalancing: the producer / MMA / allocator warpgroup needs few registers, the softmax warpgroups hold a
  // whole score row (128) plus the running output row (64) each.  384 threads x 168 = 128 x 40 + 256 x 232.
  // (setmaxnreg is issued at the top of each warpgroup's own branch so that ptxas can size the two regions.)


  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
  // Roles as in attn_tc_persistent_kernel (see the notes there): K and V each have their own producer thread, the
  // score MMAs and the P V MMAs each their own issuing thread with s_empty carrying the one ordering the single
  // issuer used to get from program order; both issue loops are unrolled by the ring depth so that stage and parity
  // are immediates.
  if (warp == 0) {
    if (elect_one()) {
      // ===================== TMA producer: Q, then the K tiles =====================
      mbar_arrive_expect_tx(q_full, TILE_BYTES);
      tma_load_4d(smem + OFF_Q, &tmQ, q_full, 0, head, q0, b);
      uint32_t ks = 0, kph = 1;
      for (int j = 0; j < T; ++j) {
        mbar_wait(&k_empty[ks], kph);
        mbar_arrive_expect_tx(&k_full[ks], TILE_BYTES);
        tma_load_4d(smem + OFF_K + ks * TILE_BYTES, &tmK, &k_full[ks], 0, head, j * AT_BN, b);
        if (++ks == AT_KS) { ks = 0; kph ^= 1; }
      }
    }
  } else if (warp == 3) {
    if (elect_one()) {
      // ===================== TMA producer: the V tiles =====================
      uint32_t vs = 0, vph = 1;
      for (int j = 0; j < T; ++j) {
        mbar_wait(&v_empty[vs], vph);
        mbar_arrive_expect_tx(&v_full[vs], TILE_BYTES);
        tma_load_4d(smem + OFF_V + vs * TILE_BYTES, &tmV, &v_full[vs], 0, head, j * AT_BN, b);
        if (++vs == AT_VS) { vs = 0; vph ^= 1; }
      }
    }
  } else {
    static_assert(AT_KS == 3 && AT_VS == 3 && AT_SS == 3, "the issue loops are unrolled by the common ring depth");
    if (elect_one()) {
      constexpr uint64_t DESC_HI = (uint64_t)((1024u >> 4) | (1u << 14) | (2u << 29)) << 32;   // SBO, version, SW128
      const uint32_t dlo = ((smem_u32(smem) >> 4) & 0x3FFFu) | (1u << 16);  // + (byte offset >> 4): no carry, smem < 256 KiB
      auto desc_at = [&](uint32_t lo, uint32_t byte_off) -> uint64_t { return DESC_HI | (uint64_t)(lo + (byte_off >> 4)); };
      if (warp == 1) {
        // ===================== MMA issuer: scores =====================
        constexpr uint32_t idesc_s = umma_idesc_bf16(AT_BM, AT_BN);
        mbar_wait(q_full, 0);
        int j = 0;
        for (uint32_t r = 0; j < T; ++r) {
          const uint32_t par = r & 1;
#pragma unroll
          for (int st = 0; st < 3; ++st) {
            if (j < T) {
              uint32_t lo = dlo;
              asm volatile("" : "+r"(lo));                   // per tile: keeps ptxas from hoisting (and spilling) 24 descriptors
              mbar_wait(&k_full[st], par);
              if (r >= 1) mbar_wait(&s_empty[st], par ^ 1);   // P V of tile j - 3 has read P out of this stage
              tc_fence_after();
#pragma unroll
              for (int k = 0; k < AT_D / 16; ++k)
                umma_bf16(TM_S + st * AT_BN, desc_at(lo, OFF_Q + k * 32), desc_at(lo, OFF_K + st * TILE_BYTES + k * 32), idesc_s,
                          k != 0 ? 1u : 0u);
              umma_commit(&s_full[st]);
              umma_commit(&k_empty[st]);
              ++j;
            }
          }
        }
      } else {
        // ===================== MMA issuer: P V =====================
        // A = P from TMEM (bf16 pairs over the first 64 columns of the tile's S stage; 16 keys = 8 columns), B = V
        // MN-major.  O[i & 1] accumulates over all of a group's tiles; the softmax rescales it in place when its row
        // max moves.
        constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(AT_BM, AT_D);
        int i = 0;
        for (uint32_t r = 0; i < T; ++r) {
          const uint32_t par = r & 1;
#pragma unroll
          for (int st = 0; st < 3; ++st) {
            if (i < T) {
              const uint32_t ob = i & 1;
              uint32_t lo = dlo;
              asm volatile("" : "+r"(lo));
              mbar_wait(&v_full[st], par);
              mbar_wait(&p_full[st], par);
              tc_fence_after();
#pragma unroll
              for (int kk = 0; kk < AT_BN / 16; ++kk) {
                umma_bf16_ts(TM_O + ob * AT_D, TM_S + st * AT_BN + kk * 8, desc_at(lo, OFF_V + st * TILE_BYTES + kk * 16 * 128),
                             idesc_o, (i >= 2 || kk != 0) ? 1u : 0u);
              }
              umma_commit(&o_full[ob]);
              umma_commit(&v_empty[st]);
              umma_commit(&s_empty[st]);
              ++i;
            }
          }
        }
      }
    }
  }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    // ===================== softmax + output =====================
    // Two warpgroups ping-pong over the key tiles: group g owns tiles j = g, g+2, ... and with them S[g], P[g],
    // O[g].  Each keeps its own running (max, sum, O) — two independent online-softmax streams — merged once at
    // the end.  Two warps per scheduler hide the TMEM-load / MUFU latencies that a single warp cannot.
    const int g = (warp - 4) >> 2;
    const int qw = warp & 3;
    const int row = qw * 32 + lane;                         // query row in the tile == TMEM lane
    const uint32_t lane_off = (uint32_t)(qw * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f;
    const float c = p.sl2;
    const uint32_t to = TM_O + lane_off + g * AT_D;

    int it = 0;  // this group's iteration count (phase bit of its barriers)
    for (int j = g; j < T; j += 2, ++it) {
      const int sb = j % AT_SS;
      mbar_wait(&s_full[sb], (j / AT_SS) & 1);
      tc_fence_after();
      const uint32_t ts = TM_S + lane_off + sb * AT_BN;
      const int key0 = j * AT_BN;
      // ---- one pass over S: all 128 scores of the row into registers (4 TMEM loads in flight, one wait) ----
      uint32_t sr[AT_BN];
#pragma unroll
      for (int cc = 0; cc < AT_BN / 32; ++cc) tmem_ld32(ts + cc * 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[cc * 32]));
      tmem_ld_wait();
      if (key0 + AT_BN > p.Nk) {  // ragged last tile: keys beyond Nk count as -inf
#pragma unroll
        for (int i = 0; i < AT_BN; ++i)
          if (key0 + i >= p.Nk) sr[i] = 0xff800000u;
      }
      // row max: 3-input max (FMNMX3), four independent chains
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < AT_BN; i += 8) {
        mx0 = max3f(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
        mx1 = max3f(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
        mx2 = max3f(mx2, __uint_as_float(sr[i + 4]), __uint_as_float(sr[i + 5]));
        mx3 = max3f(mx3, __uint_as_float(sr[i + 6]), __uint_as_float(sr[i + 7]));
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      // ---- lazy rescale: the running max only moves when the tile's max exceeds it by more than 2^8 in the exp2
      //      domain (P then stays <= 256, exact in fp32 sums and fine in bf16); m = -inf on the first tile ----
      float alpha = 1.f;
      const bool moved = (mx - m_run) * c > 8.f;
      if (moved) {
        alpha = ex2_approx((m_run - mx) * c);              // first tile: exp2(-inf) = 0
        l_run *= alpha;
        m_run = mx;
      }
      if (it >= 1) {
        // P V of this group's previous tile has landed in O[g] (the wait also keeps the phase bookkeeping in step);
        // O[g] accumulates in TMEM across tiles and is touched only when some row of the warp moved its max
        mbar_wait(&o_full[g], (it - 1) & 1);
        if (__any_sync(0xffffffffu, moved)) {
          tc_fence_after();
#pragma unroll
          for (int cc = 0; cc < AT_D / 32; ++cc) {
            uint32_t r[32];
            tmem_ld32(to + cc * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float a0 = __uint_as_float(r[i]), a1 = __uint_as_float(r[i + 1]);
              fmul2(a0, a1, alpha);
              r[i] = __float_as_uint(a0);
              r[i + 1] = __float_as_uint(a1);
            }
            tmem_st32(to + cc * 32, r);
          }
          tmem_st_wait();
        }
      }
      // ---- p = exp2(s*c - m*c), row sum, bf16 P into the swizzled smem operand ----
      float rs8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const float nmsc = -m_run * c;
      uint32_t pk[AT_BN / 2];
#pragma unroll
      for (int ch = 0; ch < AT_BN / 8; ++ch) {                // 16 chunks of 8 keys
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const int i = ch * 8 + h * 2;
          float x0, x1;
          ffma2(x0, x1, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]), c, nmsc);
          float p0, p1;
          ex2_pair(i >> 1, p0, p1, x0, x1);
          fadd2(rs8[2 * h], rs8[2 * h + 1], p0, p1);
          pk[i >> 1] = pack_bf16x2(p0, p1);
        }
      }
      // P (bf16 pairs: key 2c in the low half of column c) over the first 64 columns of this tile's S buffer
      tmem_st32(ts, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
      tmem_st32(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
      tmem_st_wait();
      l_run += ((rs8[0] + rs8[1]) + (rs8[2] + rs8[3])) + ((rs8[4] + rs8[5]) + (rs8[6] + rs8[7]));
      tc_fence_before();
      mbar_arrive(&p_full[sb]);
    }
    // ---- this group's output row: O[g] after its last P V ----
    float o_acc[AT_D];
#pragma unroll
    for (int i = 0; i < AT_D; ++i) o_acc[i] = 0.f;
    if (it >= 1) {
      mbar_wait(&o_full[g], (it - 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int cc = 0; cc < AT_D / 32; ++cc) {
        uint32_t r[32];
        tmem_ld32(to + cc * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[cc * 32 + i] = __uint_as_float(r[i]);
      }
      tc_fence_before();
    }
    // ---- merge the two streams: group 1 parks (m, l, O) in its own (now idle) P buffer, group 0 combines ----
    float* xch = reinterpret_cast<float*>(smem + OFF_XCH);                  // 32 KiB = 128 rows x 64 fp32
    float* xml = reinterpret_cast<float*>(smem + OFF_XML);                  // 128 x {m, l}
    if (g == 1) {
#pragma unroll
      for (int i = 0; i < AT_D; ++i) xch[i * AT_BM + row] = o_acc[i];     // column-major: conflict-free
      xml[row * 2] = m_run;
      xml[row * 2 + 1] = l_run;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (g == 0) {
      const float m1 = xml[row * 2], l1 = xml[row * 2 + 1];
      const float m = fmaxf(m_run, m1);
      const float w0 = ex2_approx((m_run - m) * c);
      const float w1 = (m1 == -INFINITY) ? 0.f : ex2_approx((m1 - m) * c);
      const float inv = 1.f / (l_run * w0 + l1 * w1);
      if (q0 + row < p.Nq) {
        bf16* orow = p.o + ((long long)b * p.Nq + q0 + row) * p.ldo + head * p.d;
#pragma unroll
        for (int v8 = 0; v8 < AT_D / 8; ++v8) {
        if (v8 * 8 >= p.d) break;
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = (o_acc[v8 * 8 + i] * w0 + xch[(v8 * 8 + i) * AT_BM + row] * w1) * inv;
          uint4 u;
          u.x = pack_bf16x2(f[0], f[1]);
          u.y = pack_bf16x2(f[2], f[3]);
          u.z = pack_bf16x2(f[4], f[5]);
          u.w = pack_bf16x2(f[6], f[7]);
          *reinterpret_cast<uint4*>(orow + v8 * 8) = u;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Multi-item form of attn_tc_kernel: a CTA walks `ipc` consecutive (query tile, head, batch) work items
// w = blockIdx.x * ipc + 0 .. ipc-1 (a handful; the grid stays several waves long so that the hardware's dynamic block
// scheduling still evens out SM-to-SM differences — a fully static one-CTA-per-SM split measured slower on long
// sequences).  With one 128-query tile per launch slot, a CTA's fill (TMEM allocation, barrier
// set-up, the first Q / K loads' DRAM latency, the first score MMA) and drain (merge of the two softmax streams, the
// output store) were exposed for every tile — about a tenth of a 32-key-tile item and a third of an 8-key-tile one.
// Here the pipeline simply runs on: the K / V rings, the three score accumulators and all barriers keep one running
// tile count across items, Q is double-buffered, so the producer is already fetching the next item's Q / K / V and
// the MMA thread already issuing its first score tiles while the softmax warps finish and merge the current one.
// Extra hand-offs: q_empty (the item's last score MMA has retired: its Q buffer may be refilled), o_empty[g] (group g
// has read its finished output accumulator: the next item's first P V may overwrite it), and named barrier 2
// (group 0 is done reading group 1's parked rows of the previous item).
// ------------------------------------------------------------------------------------------------------------
constexpr int ATP_OFF_Q = 0;                                   // two Q buffers
constexpr int ATP_OFF_K = ATP_OFF_Q + 2 * TILE_BYTES;
constexpr int ATP_OFF_V = ATP_OFF_K + AT_KS * TILE_BYTES;
constexpr int ATP_OFF_XCH = ATP_OFF_V + AT_VS * TILE_BYTES;
constexpr int ATP_OFF_BAR = ATP_OFF_XCH + 2 * TILE_BYTES;
constexpr int ATP_OFF_XML = ATP_OFF_BAR + 512;
constexpr int ATP_SMEM = ATP_OFF_XML + 1024 + 1024;

struct AttnTcPP {
  AttnTcP a;
  int nqt, heads, total;   // query tiles per (batch, head), heads, work items
  int ipc;                 // work items per CTA (the last CTAs may get fewer)
  int w_mul, w_stride;     // item wi of CTA c is w = c * w_mul + wi * w_stride: (ipc, 1) = consecutive items per CTA,
                           // (1, gridDim.x) = one CTA per SM striding over all items
};

__global__ void __launch_bounds__(384, 1)
attn_tc_persistent_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                          const __grid_constant__ CUtensorMap tmV, const AttnTcPP pp) {
  const AttnTcP& p = pp.a;
  extern __shared__ uint8_t at_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ATP_OFF_BAR);
  uint64_t* q_full = bars;                 // 2
  uint64_t* q_empty = q_full + 2;          // 2
  uint64_t* k_full = q_empty + 2;          // AT_KS
  uint64_t* k_empty = k_full + AT_KS;      // AT_KS
  uint64_t* v_full = k_empty + AT_KS;      // AT_VS
  uint64_t* v_empty = v_full + AT_VS;      // AT_VS
  uint64_t* s_full = v_empty + AT_VS;      // AT_SS
  uint64_t* p_full = s_full + AT_SS;       // AT_SS
  uint64_t* o_full = p_full + AT_SS;       // 2
  uint64_t* o_empty = o_full + 2;          // 2
  uint64_t* s_empty = o_empty + 2;         // AT_SS: P V of the stage has completed, the next score MMA may overwrite it
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + AT_SS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = (p.Nk + AT_BN - 1) / AT_BN;
  const int w_first = (int)blockIdx.x * pp.w_mul;
  int n_items = (pp.total - w_first + pp.w_stride - 1) / pp.w_stride;                  // work items of this CTA (>= 1)
  n_items = n_items < pp.ipc ? n_items : pp.ipc;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < AT_KS; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < AT_VS; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < AT_SS; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 128); }
    for (int i = 0; i < 2; ++i) { mbar_init(&o_full[i], 1); mbar_init(&o_empty[i], 128); }
    for (int i = 0; i < AT_SS; ++i) mbar_init(&s_empty[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t TM_S = tmem_base;
  const uint32_t TM_O = tmem_base + AT_SS * AT_BN;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    // The single-thread roles below sit on the kernel's critical path (the trace in profiles/ shows the tensor pipe
    // idle whenever the issuing thread is between two MMA batches), so each keeps running ring positions and phase
    // bits instead of dividing the flat tile count, and builds an operand descriptor by one add onto a precomputed
    // word.  (Running the loops warp-converged with lane 0 issuing was tried: the compiler then wraps every
    // tcgen05.mma in an elect / broadcast loop, which is slower than this.)
    if (warp == 0) {
      if (elect_one()) {
        // ===================== TMA producer for Q and K: runs ahead across work items =====================
        uint32_t ks = 0, kph = 1;                           // ring position, parity to wait for on k_empty
        for (int wi = 0; wi < n_items; ++wi) {
          const int w = w_first + wi * pp.w_stride;
          const int qt = w % pp.nqt, head = (w / pp.nqt) % pp.heads, b = w / (pp.nqt * pp.heads);
          const int qs = wi & 1;
          mbar_wait(&q_empty[qs], ((wi >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&q_full[qs], TILE_BYTES);
          tma_load_4d(smem + ATP_OFF_Q + qs * TILE_BYTES, &tmQ, &q_full[qs], 0, head, qt * AT_BM, b);
          for (int j = 0; j < T; ++j) {
            mbar_wait(&k_empty[ks], kph);
            mbar_arrive_expect_tx(&k_full[ks], TILE_BYTES);
            tma_load_4d(smem + ATP_OFF_K + ks * TILE_BYTES, &tmK, &k_full[ks], 0, head, j * AT_BN, b);
            if (++ks == AT_KS) { ks = 0; kph ^= 1; }
          }
        }
      }
    } else if (warp == 3) {
      if (elect_one()) {
        // ===================== TMA producer for V: its own thread, so a full V ring never holds back a K load ========
        uint32_t vs = 0, vph = 1;
        for (int wi = 0; wi < n_items; ++wi) {
          const int w = w_first + wi * pp.w_stride;
          const int head = (w / pp.nqt) % pp.heads, b = w / (pp.nqt * pp.heads);
          for (int j = 0; j < T; ++j) {
            mbar_wait(&v_empty[vs], vph);
            mbar_arrive_expect_tx(&v_full[vs], TILE_BYTES);
            tma_load_4d(smem + ATP_OFF_V + vs * TILE_BYTES, &tmV, &v_full[vs], 0, head, j * AT_BN, b);
            if (++vs == AT_VS) { vs = 0; vph ^= 1; }
          }
        }
      }
    } else {
      // ===================== MMA issuers: warp 1 the score products, warp 2 the P V products =====================
      // Two threads, because one thread's address and barrier arithmetic between two batches is what the tensor pipe
      // waited on; tcgen05 operations of different threads are not ordered, so the one hazard the single issuer
      // resolved by program order — the score MMA of tile j + 3 overwriting the S / P stage P V of tile j reads — is
      // carried by s_empty.  Both loops are unrolled by the ring depth: tile 3r + st sits in stage st of every ring
      // with parity r & 1, so barrier addresses and descriptors are immediates on one base.
      static_assert(AT_KS == 3 && AT_VS == 3 && AT_SS == 3, "the issue loops are unrolled by the common ring depth");
      if (elect_one()) {
#ifdef MDX_ATTN_TRACE
        int tr_n = 0;
#endif
        constexpr uint64_t DESC_HI = (uint64_t)((1024u >> 4) | (1u << 14) | (2u << 29)) << 32;   // SBO, version, SW128
        const uint32_t dlo = ((smem_u32(smem) >> 4) & 0x3FFFu) | (1u << 16);  // + (byte offset >> 4): no carry, smem < 256 KiB
        auto desc_at = [&](uint32_t lo, uint32_t byte_off) -> uint64_t { return DESC_HI | (uint64_t)(lo + (byte_off >> 4)); };
        const uint32_t total_tiles = (uint32_t)n_items * (uint32_t)T;
        if (warp == 1) {
          constexpr uint32_t idesc_s = umma_idesc_bf16(AT_BM, AT_BN);
          uint32_t wi = 0, j = 0, jj = 0;                    // (item, tile inside it), flat tile count
          for (uint32_t r = 0; jj < total_tiles; ++r) {
            const uint32_t par = r & 1;
#pragma unroll
            for (int st = 0; st < 3; ++st) {
              if (jj < total_tiles) {
                const uint32_t qs = wi & 1;
                uint32_t lo = dlo;
                asm volatile("" : "+r"(lo));                 // per tile: keeps ptxas from hoisting (and spilling) 24 descriptors
                if (j == 0) mbar_wait(&q_full[qs], (wi >> 1) & 1);
                mbar_wait(&k_full[st], par);
                if (r >= 1) mbar_wait(&s_empty[st], par ^ 1);
                AT_TR(0, 1, jj);
                tc_fence_after();
                const uint32_t q_off = ATP_OFF_Q + qs * TILE_BYTES;
#pragma unroll
                for (int k = 0; k < AT_D / 16; ++k)
                  umma_bf16(TM_S + st * AT_BN, desc_at(lo, q_off + k * 32), desc_at(lo, ATP_OFF_K + st * TILE_BYTES + k * 32),
                            idesc_s, k != 0 ? 1u : 0u);
                umma_commit(&s_full[st]);
                umma_commit(&k_empty[st]);
                if (j == (uint32_t)T - 1) umma_commit(&q_empty[qs]);   // the item's score MMAs are all issued
                AT_TR(0, 2, jj);
                if (++j == (uint32_t)T) { j = 0; ++wi; }
                ++jj;
              }
            }
          }
        } else {
          constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(AT_BM, AT_D);
          uint32_t pwi = 0, pj = 0, i = 0;
          for (uint32_t r = 0; i < total_tiles; ++r) {
            const uint32_t par = r & 1;
#pragma unroll
            for (int st = 0; st < 3; ++st) {
              if (i < total_tiles) {
                const uint32_t ob = pj & 1;
                uint32_t lo = dlo;
                asm volatile("" : "+r"(lo));
                if (pj < 2 && pwi >= 1) {       // first P V of this group in a new item overwrites O[ob]:
                  mbar_wait(&o_empty[ob], (pwi - 1) & 1);    // the group has read the previous item's result out of it
                }
                mbar_wait(&v_full[st], par);
                mbar_wait(&p_full[st], par);
                AT_TR(3, 3, i);
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < AT_BN / 16; ++kk) {
                  umma_bf16_ts(TM_O + ob * AT_D, TM_S + st * AT_BN + kk * 8,
                               desc_at(lo, ATP_OFF_V + st * TILE_BYTES + kk * 16 * 128), idesc_o, (pj >= 2 || kk != 0) ? 1u : 0u);
                }
                umma_commit(&o_full[ob]);
                umma_commit(&v_empty[st]);
                umma_commit(&s_empty[st]);
                AT_TR(3, 5, i);
                if (++pj == (uint32_t)T) { pj = 0; ++pwi; }
                ++i;
              }
            }
          }
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    // ===================== softmax + output (as attn_tc_kernel, phases from running counts) =====================
    const int g = (warp - 4) >> 2;
    const int qw = warp & 3;
    const int row = qw * 32 + lane;
    const uint32_t lane_off = (uint32_t)(qw * 32) << 16;
    const float c = p.sl2;
    const uint32_t to = TM_O + lane_off + g * AT_D;
    float* xch = reinterpret_cast<float*>(smem + ATP_OFF_XCH);
    float* xml = reinterpret_cast<float*>(smem + ATP_OFF_XML);
    uint32_t pvc = 0;                                        // P V completions of this group waited for so far
#ifdef MDX_ATTN_TRACE
    int tr_n = (qw == 0 && lane == 0) ? 0 : AT_TR_N;
#endif
    for (int wi = 0; wi < n_items; ++wi) {
      const int w = w_first + wi * pp.w_stride;
      const int qt = w % pp.nqt, head = (w / pp.nqt) % pp.heads, b = w / (pp.nqt * pp.heads);
      const int q0 = qt * AT_BM;
      float m_run = -INFINITY, l_run = 0.f;
      int it = 0;
      for (int j = g; j < T; j += 2, ++it) {
        const uint32_t jj = (uint32_t)wi * (uint32_t)T + (uint32_t)j;
        const int sb = jj % AT_SS;
        AT_TR(1 + g, 0, jj);
        mbar_wait(&s_full[sb], (jj / AT_SS) & 1);
        AT_TR(1 + g, 1, jj);
        tc_fence_after();
        const uint32_t ts = TM_S + lane_off + sb * AT_BN;
        const int key0 = j * AT_BN;
        uint32_t sr[AT_BN];
#pragma unroll
        for (int cc = 0; cc < AT_BN / 32; ++cc) tmem_ld32(ts + cc * 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[cc * 32]));
        tmem_ld_wait();
        AT_TR(1 + g, 2, jj);
        if (key0 + AT_BN > p.Nk) {
#pragma unroll
          for (int i = 0; i < AT_BN; ++i)
            if (key0 + i >= p.Nk) sr[i] = 0xff800000u;
        }
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < AT_BN; i += 8) {
          mx0 = max3f(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
          mx1 = max3f(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
          mx2 = max3f(mx2, __uint_as_float(sr[i + 4]), __uint_as_float(sr[i + 5]));
          mx3 = max3f(mx3, __uint_as_float(sr[i + 6]), __uint_as_float(sr[i + 7]));
        }
        const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        float alpha = 1.f;
        const bool moved = (mx - m_run) * c > 8.f;
        if (moved) {
          alpha = ex2_approx((m_run - mx) * c);
          l_run *= alpha;
          m_run = mx;
        }
        AT_TR(1 + g, 3, jj);
        if (it >= 1) {
          mbar_wait(&o_full[g], pvc & 1);
          ++pvc;
          if (__any_sync(0xffffffffu, moved)) {
            AT_TR(1 + g, 7, jj);
            tc_fence_after();
#pragma unroll
            for (int cc = 0; cc < AT_D / 32; ++cc) {
              uint32_t r[32];
              tmem_ld32(to + cc * 32, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                float a0 = __uint_as_float(r[i]), a1 = __uint_as_float(r[i + 1]);
                fmul2(a0, a1, alpha);
                r[i] = __float_as_uint(a0);
                r[i + 1] = __float_as_uint(a1);
              }
              tmem_st32(to + cc * 32, r);
            }
            tmem_st_wait();
          }
        }
        AT_TR(1 + g, 4, jj);
        float rs8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float nmsc = -m_run * c;
        uint32_t pk[AT_BN / 2];
#pragma unroll
        for (int ch = 0; ch < AT_BN / 8; ++ch) {
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const int i = ch * 8 + h * 2;
            float x0, x1;
            ffma2(x0, x1, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]), c, nmsc);
            float p0, p1;
            ex2_pair(i >> 1, p0, p1, x0, x1);
            fadd2(rs8[2 * h], rs8[2 * h + 1], p0, p1);
            pk[i >> 1] = pack_bf16x2(p0, p1);
          }
        }
        AT_TR(1 + g, 5, jj);
        tmem_st32(ts, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
        tmem_st32(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
        tmem_st_wait();
        l_run += ((rs8[0] + rs8[1]) + (rs8[2] + rs8[3])) + ((rs8[4] + rs8[5]) + (rs8[6] + rs8[7]));
        tc_fence_before();
        mbar_arrive(&p_full[sb]);
        AT_TR(1 + g, 6, jj);
      }
      // ---- this group's output row of the item, then hand O[g] back to the MMA thread ----
      float o_acc[AT_D];
#pragma unroll
      for (int i = 0; i < AT_D; ++i) o_acc[i] = 0.f;
      if (it >= 1) {
        mbar_wait(&o_full[g], pvc & 1);
        ++pvc;
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < AT_D / 32; ++cc) {
          uint32_t r[32];
          tmem_ld32(to + cc * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o_acc[cc * 32 + i] = __uint_as_float(r[i]);
        }
        tc_fence_before();
      }
      mbar_arrive(&o_empty[g]);
      // ---- merge: group 1 parks (m, l, O), group 0 combines and stores ----
      if (g == 1) {
        if (wi >= 1) asm volatile("bar.sync 2, 256;" ::: "memory");   // group 0 has read the previous item's rows
#pragma unroll
        for (int i = 0; i < AT_D; ++i) xch[i * AT_BM + row] = o_acc[i];
        xml[row * 2] = m_run;
        xml[row * 2 + 1] = l_run;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (g == 0) {
        const float m1 = xml[row * 2], l1 = xml[row * 2 + 1];
        const float m = fmaxf(m_run, m1);
        const float w0 = ex2_approx((m_run - m) * c);
        const float w1 = (m1 == -INFINITY) ? 0.f : ex2_approx((m1 - m) * c);
        const float inv = 1.f / (l_run * w0 + l1 * w1);
        float f[AT_D];
#pragma unroll
        for (int i = 0; i < AT_D; ++i) f[i] = (o_acc[i] * w0 + xch[i * AT_BM + row] * w1) * inv;
        if (wi + 1 < n_items) asm volatile("bar.arrive 2, 256;" ::: "memory");   // the parked rows may be overwritten
        if (q0 + row < p.Nq) {
          bf16* orow = p.o + ((long long)b * p.Nq + q0 + row) * p.ldo + head * p.d;
#pragma unroll
          for (int v8 = 0; v8 < AT_D / 8; ++v8) {
            if (v8 * 8 >= p.d) break;
            uint4 u;
            u.x = pack_bf16x2(f[v8 * 8], f[v8 * 8 + 1]);
            u.y = pack_bf16x2(f[v8 * 8 + 2], f[v8 * 8 + 3]);
            u.z = pack_bf16x2(f[v8 * 8 + 4], f[v8 * 8 + 5]);
            u.w = pack_bf16x2(f[v8 * 8 + 6], f[v8 * 8 + 7]);
            *reinterpret_cast<uint4*>(orow + v8 * 8) = u;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Single-tile variant: Nk <= 128 (the cross-attention over 77 text tokens, attention.py:118-152).  There is no
// key loop to pipeline, so the CTA is small (4 role warps + ONE softmax warpgroup, 48 KiB smem, 128 TMEM columns:
// S, then P over its first half and O over its second) and two CTAs share an SM to overlap each other's
// TMA -> MMA -> softmax -> MMA -> store latency chain.
// ------------------------------------------------------------------------------------------------------------
constexpr int AT1_SMEM = 3 * TILE_BYTES + 256 + 1024;

__global__ void __launch_bounds__(256, 2)
attn_tc_single_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const AttnTcP p) {
  extern __shared__ uint8_t at_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 3 * TILE_BYTES);
  uint64_t* qk_full = bars;       // Q and K landed
  uint64_t* v_full = bars + 1;
  uint64_t* s_full = bars + 2;
  uint64_t* p_full = bars + 3;
  uint64_t* o_full = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * AT_BM;
  const int head = blockIdx.y, b = blockIdx.z;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(qk_full, 1);
    mbar_init(v_full, 1);
    mbar_init(s_full, 1);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int ksteps = (p.Nk + 15) / 16;          // P V only over the 16-key blocks that hold real keys

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 0) {
      if (elect_one()) {
        mbar_arrive_expect_tx(qk_full, 2 * TILE_BYTES);
        tma_load_4d(smem, &tmQ, qk_full, 0, head, q0, b);
        tma_load_4d(smem + TILE_BYTES, &tmK, qk_full, 0, head, 0, b);
        mbar_arrive_expect_tx(v_full, TILE_BYTES);
        tma_load_4d(smem + 2 * TILE_BYTES, &tmV, v_full, 0, head, 0, b);
      }
    } else if (warp == 1) {
      if (elect_one()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(AT_BM, AT_BN);
        constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(AT_BM, AT_D);
        const uint32_t q_base = smem_u32(smem), k_base = smem_u32(smem + TILE_BYTES);
        const uint32_t v_base = smem_u32(smem + 2 * TILE_BYTES);
        mbar_wait(qk_full, 0);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < AT_D / 16; ++k)
          umma_bf16(tmem_base, umma_desc_sw128(q_base + k * 32), umma_desc_sw128(k_base + k * 32), idesc_s, k != 0 ? 1u : 0u);
        umma_commit(s_full);
        mbar_wait(p_full, 0);
        mbar_wait(v_full, 0);
        tc_fence_after();
        for (int kk = 0; kk < ksteps; ++kk)
          umma_bf16_ts(tmem_base + 64, tmem_base + kk * 8, umma_desc_sw128_mn(v_base + kk * 16 * 128), idesc_o,
                       kk != 0 ? 1u : 0u);
        umma_commit(o_full);
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    const int qw = warp & 3;
    const int row = qw * 32 + lane;
    const uint32_t ts = tmem_base + ((uint32_t)(qw * 32) << 16);
    const float c = p.sl2;
    mbar_wait(s_full, 0);
    tc_fence_after();
    uint32_t sr[AT_BN];
#pragma unroll
    for (int cc = 0; cc < AT_BN / 32; ++cc)
      if (cc * 32 < p.Nk) tmem_ld32(ts + cc * 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[cc * 32]));
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < AT_BN; ++i)
      if (i >= p.Nk) sr[i] = 0xff800000u;            // keys beyond Nk (and the column groups not loaded) count as -inf
    float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
    for (int i = 0; i < AT_BN; i += 8) {
      mx0 = max3f(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
      mx1 = max3f(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
      mx2 = max3f(mx2, __uint_as_float(sr[i + 4]), __uint_as_float(sr[i + 5]));
      mx3 = max3f(mx3, __uint_as_float(sr[i + 6]), __uint_as_float(sr[i + 7]));
    }
    const float nmsc = -fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * c;
    float rs8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    uint32_t pk[AT_BN / 2];
#pragma unroll
    for (int i = 0; i < AT_BN; i += 2) {
      float x0, x1;
      ffma2(x0, x1, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]), c, nmsc);
      float p0, p1;
      ex2_pair(i >> 1, p0, p1, x0, x1);
      fadd2(rs8[i & 6], rs8[(i & 6) + 1], p0, p1);
      pk[i >> 1] = pack_bf16x2(p0, p1);
    }
    const float l = ((rs8[0] + rs8[1]) + (rs8[2] + rs8[3])) + ((rs8[4] + rs8[5]) + (rs8[6] + rs8[7]));
    tmem_st32(ts, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
    tmem_st32(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
    tmem_st_wait();
    tc_fence_before();
    mbar_arrive(p_full);
    const float inv = 1.f / l;
    mbar_wait(o_full, 0);
    tc_fence_after();
    uint32_t r[AT_D];
    tmem_ld32(ts + 64, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
    tmem_ld32(ts + 96, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
    tmem_ld_wait();
    tc_fence_before();
    if (q0 + row < p.Nq) {
      bf16* orow = p.o + ((long long)b * p.Nq + q0 + row) * p.ldo + head * p.d;
#pragma unroll
      for (int v8 = 0; v8 < AT_D / 8; ++v8) {
        if (v8 * 8 >= p.d) break;
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(r[v8 * 8]) * inv, __uint_as_float(r[v8 * 8 + 1]) * inv);
        u.y = pack_bf16x2(__uint_as_float(r[v8 * 8 + 2]) * inv, __uint_as_float(r[v8 * 8 + 3]) * inv);
        u.z = pack_bf16x2(__uint_as_float(r[v8 * 8 + 4]) * inv, __uint_as_float(r[v8 * 8 + 5]) * inv);
        u.w = pack_bf16x2(__uint_as_float(r[v8 * 8 + 6]) * inv, __uint_as_float(r[v8 * 8 + 7]) * inv);
        *reinterpret_cast<uint4*>(orow + v8 * 8) = u;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Multi-item form of the single-tile case (Nk <= 128: the cross-attention over 77 text tokens).  One item = one
// 128-query tile against one key tile: a score MMA, one softmax, one P V — a pure latency chain when every item is its
// own CTA.  Here a CTA (one per SM) walks its items and the two softmax warpgroups take ALTERNATE items (group g owns
// items g, g+2, ... with O[g] and no merge), so one group's exponentials run while the other waits for its P V, the
// MMA thread issues score MMAs two items ahead, and the producer keeps three items' K / V and two items' Q in flight.
// Only the key columns that exist are loaded, exponentiated and multiplied (77 keys: 3 of 4 score column groups, 10
// of 16 exp chunks, 5 of 8 P V steps).
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(384, 1)
attn_tc_single_persistent_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                                 const __grid_constant__ CUtensorMap tmV, const AttnTcPP pp) {
  const AttnTcP& p = pp.a;
  extern __shared__ uint8_t at_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ATP_OFF_BAR);
  uint64_t* q_full = bars;                 // 2
  uint64_t* q_empty = q_full + 2;          // 2
  uint64_t* k_full = q_empty + 2;          // AT_KS
  uint64_t* k_empty = k_full + AT_KS;      // AT_KS
  uint64_t* v_full = k_empty + AT_KS;      // AT_VS
  uint64_t* v_empty = v_full + AT_VS;      // AT_VS
  uint64_t* s_full = v_empty + AT_VS;      // AT_SS
  uint64_t* p_full = s_full + AT_SS;       // AT_SS
  uint64_t* o_full = p_full + AT_SS;       // 2
  uint64_t* o_empty = o_full + 2;          // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = (int)gridDim.x;
  const int n_items = (pp.total - (int)blockIdx.x + G - 1) / G;     // items w = blockIdx.x + wi * G of this CTA (>= 1)
  const int ksteps = (p.Nk + 15) / 16;                              // P V only over the 16-key blocks that hold keys

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < AT_KS; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < AT_VS; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < AT_SS; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 128); }
    for (int i = 0; i < 2; ++i) { mbar_init(&o_full[i], 1); mbar_init(&o_empty[i], 128); }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t TM_S = tmem_base;
  const uint32_t TM_O = tmem_base + AT_SS * AT_BN;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 0) {
      if (elect_one()) {
        for (int wi = 0; wi < n_items; ++wi) {
          const int w = (int)blockIdx.x + wi * G;
          const int qt = w % pp.nqt, head = (w / pp.nqt) % pp.heads, b = w / (pp.nqt * pp.heads);
          const int qs = wi & 1, ks = wi % AT_KS, vs = wi % AT_VS;
          mbar_wait(&q_empty[qs], ((wi >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&q_full[qs], TILE_BYTES);
          tma_load_4d(smem + ATP_OFF_Q + qs * TILE_BYTES, &tmQ, &q_full[qs], 0, head, qt * AT_BM, b);
          mbar_wait(&k_empty[ks], ((wi / AT_KS) & 1) ^ 1);
          mbar_arrive_expect_tx(&k_full[ks], TILE_BYTES);
          tma_load_4d(smem + ATP_OFF_K + ks * TILE_BYTES, &tmK, &k_full[ks], 0, head, 0, b);
          mbar_wait(&v_empty[vs], ((wi / AT_VS) & 1) ^ 1);
          mbar_arrive_expect_tx(&v_full[vs], TILE_BYTES);
          tma_load_4d(smem + ATP_OFF_V + vs * TILE_BYTES, &tmV, &v_full[vs], 0, head, 0, b);
        }
      }
    } else if (warp == 1) {
      if (elect_one()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(AT_BM, AT_BN);
        constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(AT_BM, AT_D);
        auto issue_pv = [&](int i) {
          const int pb = i % AT_SS, vs = i % AT_VS, ob = i & 1;
          if (i >= 2) mbar_wait(&o_empty[ob], ((i >> 1) - 1) & 1);   // the group has read item i-2's result out of O[ob]
          mbar_wait(&p_full[pb], (i / AT_SS) & 1);
          mbar_wait(&v_full[vs], (i / AT_VS) & 1);
          tc_fence_after();
          const uint32_t p_tmem = TM_S + pb * AT_BN;
          const uint32_t v_base = smem_u32(smem + ATP_OFF_V + vs * TILE_BYTES);
          for (int kk = 0; kk < ksteps; ++kk)
            umma_bf16_ts(TM_O + ob * AT_D, p_tmem + kk * 8, umma_desc_sw128_mn(v_base + kk * 16 * 128), idesc_o,
                         kk != 0 ? 1u : 0u);
          umma_commit(&o_full[ob]);
          umma_commit(&v_empty[vs]);
        };
        for (int wi = 0; wi < n_items; ++wi) {
          const int ks = wi % AT_KS, sb = wi % AT_SS, qs = wi & 1;
          mbar_wait(&q_full[qs], (wi >> 1) & 1);
          mbar_wait(&k_full[ks], (wi / AT_KS) & 1);
          tc_fence_after();
          const uint32_t q_base = smem_u32(smem + ATP_OFF_Q + qs * TILE_BYTES);
          const uint32_t k_base = smem_u32(smem + ATP_OFF_K + ks * TILE_BYTES);
#pragma unroll
          for (int k = 0; k < AT_D / 16; ++k)
            umma_bf16(TM_S + sb * AT_BN, umma_desc_sw128(q_base + k * 32), umma_desc_sw128(k_base + k * 32), idesc_s,
                      k != 0 ? 1u : 0u);
          umma_commit(&s_full[sb]);
          umma_commit(&k_empty[ks]);
          umma_commit(&q_empty[qs]);
          if (wi >= 2) issue_pv(wi - 2);
        }
        if (n_items >= 2) issue_pv(n_items - 2);
        issue_pv(n_items - 1);
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int g = (warp - 4) >> 2;
    const int qw = warp & 3;
    const int row = qw * 32 + lane;
    const uint32_t lane_off = (uint32_t)(qw * 32) << 16;
    const float c = p.sl2;
    const uint32_t to = TM_O + lane_off + g * AT_D;
    for (int wi = g; wi < n_items; wi += 2) {
      const int w = (int)blockIdx.x + wi * G;
      const int qt = w % pp.nqt, head = (w / pp.nqt) % pp.heads, b = w / (pp.nqt * pp.heads);
      const int q0 = qt * AT_BM;
      const int sb = wi % AT_SS;
      mbar_wait(&s_full[sb], (wi / AT_SS) & 1);
      tc_fence_after();
      const uint32_t ts = TM_S + lane_off + sb * AT_BN;
      uint32_t sr[AT_BN];
#pragma unroll
      for (int cc = 0; cc < AT_BN / 32; ++cc)
        if (cc * 32 < p.Nk) tmem_ld32(ts + cc * 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[cc * 32]));
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < AT_BN; ++i)
        if (i >= p.Nk) sr[i] = 0xff800000u;            // keys beyond Nk (and the column groups not loaded): -inf
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < AT_BN; i += 8) {
        mx0 = max3f(mx0, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]));
        mx1 = max3f(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
        mx2 = max3f(mx2, __uint_as_float(sr[i + 4]), __uint_as_float(sr[i + 5]));
        mx3 = max3f(mx3, __uint_as_float(sr[i + 6]), __uint_as_float(sr[i + 7]));
      }
      const float nmsc = -fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * c;
      float rs8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      uint32_t pk[AT_BN / 2];
#pragma unroll
      for (int ch = 0; ch < AT_BN / 8; ++ch) {
        if (ch * 8 < p.Nk) {                            // warp-uniform: chunks with no real key are skipped
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const int i = ch * 8 + h * 2;
            float x0, x1;
            ffma2(x0, x1, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]), c, nmsc);
            float p0, p1;
            ex2_pair(i >> 1, p0, p1, x0, x1);
            fadd2(rs8[2 * h], rs8[2 * h + 1], p0, p1);
            pk[i >> 1] = pack_bf16x2(p0, p1);
          }
        } else {
#pragma unroll
          for (int h = 0; h < 4; ++h) pk[(ch * 8 + h * 2) >> 1] = 0u;
        }
      }
      const float l = ((rs8[0] + rs8[1]) + (rs8[2] + rs8[3])) + ((rs8[4] + rs8[5]) + (rs8[6] + rs8[7]));
      tmem_st32(ts, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
      tmem_st32(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&p_full[sb]);
      const float inv = 1.f / l;
      mbar_wait(&o_full[g], (wi >> 1) & 1);
      tc_fence_after();
      uint32_t r[AT_D];
      tmem_ld32(to, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
      tmem_ld32(to + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&o_empty[g]);
      if (q0 + row < p.Nq) {
        bf16* orow = p.o + ((long long)b * p.Nq + q0 + row) * p.ldo + head * p.d;
#pragma unroll
        for (int v8 = 0; v8 < AT_D / 8; ++v8) {
          if (v8 * 8 >= p.d) break;
          uint4 u;
          u.x = pack_bf16x2(__uint_as_float(r[v8 * 8]) * inv, __uint_as_float(r[v8 * 8 + 1]) * inv);
          u.y = pack_bf16x2(__uint_as_float(r[v8 * 8 + 2]) * inv, __uint_as_float(r[v8 * 8 + 3]) * inv);
          u.z = pack_bf16x2(__uint_as_float(r[v8 * 8 + 4]) * inv, __uint_as_float(r[v8 * 8 + 5]) * inv);
          u.w = pack_bf16x2(__uint_as_float(r[v8 * 8 + 6]) * inv, __uint_as_float(r[v8 * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + v8 * 8) = u;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn at_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// [batch, rows, heads, d] bf16 view (row stride ld, heads packed d apart): box = 64 channels x 1 head x 128 rows x 1
// batch.  For d < 64 the box runs past the head's d channels; TMA zero-fills those columns, which adds nothing to
// q k^T and leaves the padded output columns (never stored) at zero.
int make_map_qkv(CUtensorMap* m, const void* ptr, long long ld, int d, int heads, int rows, int batch) {
  EncodeTiledFn enc = at_encode_fn();
  if (!enc) {
    mdx_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return MDX_ERR_CUDA;
  }
  cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)heads, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[3] = {(cuuint64_t)d * 2, (cuuint64_t)ld * 2, (cuuint64_t)rows * ld * 2};
  cuuint32_t box[4] = {64, 1, 128, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mdx_set_error("cuTensorMapEncodeTiled(attention) failed with CUresult %d", (int)r);
    return MDX_ERR_CUDA;
  }
  return MDX_OK;
}

}  // namespace

// Internal entry (called by mdx_attention for bf16, d <= 64, d % 8 == 0, not causal).
#ifdef MDX_ATTN_TRACE
extern "C" int mdx_attn_trace_dump(uint32_t* host, int n_words) {
  return (int)cudaMemcpyFromSymbol(host, at_trace, sizeof(uint32_t) * n_words);
}
#endif
int mdx_attention_tc_d64(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                         void* o, long long ldo, int batch, int heads, int Nq, int Nk, int d, float scale,
                         cudaStream_t st) {
  CUtensorMap mq, mk, mv;
  int rc;
  if ((rc = make_map_qkv(&mq, q, ldq, d, heads, Nq, batch))) return rc;
  if ((rc = make_map_qkv(&mk, k, ldk, d, heads, Nk, batch))) return rc;
  if ((rc = make_map_qkv(&mv, v, ldv, d, heads, Nk, batch))) return rc;
  static MdxPerDeviceOnce attr;
  if (attr.need()) {
    MDX_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
    MDX_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_single_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT1_SMEM));
    MDX_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATP_SMEM));
    MDX_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_single_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        ATP_SMEM));
    attr.mark();
  }
  AttnTcP p;
  p.Nq = Nq; p.Nk = Nk; p.sl2 = scale * 1.4426950408889634f;
  p.o = reinterpret_cast<bf16*>(o); p.ldo = ldo; p.d = d;
  dim3 grid((Nq + AT_BM - 1) / AT_BM, heads, batch);
  // Launch form.  Measured in one process (tools/attn_ab.py, CFG batch 128): one CTA per SM striding over all query
  // tiles beats one tile per CTA by 3 % at 4096 keys, 13 % at 1024 and 26 % at 256 (a CTA's fill and drain are
  // overlapped with the neighbouring tiles), but loses 4-9 % on long sequences (9216 keys) and when an SM gets fewer
  // than ~30 tiles; a few consecutive tiles per CTA sits in between.  MDX_ATTN_NO_PERSIST / MDX_ATTN_IPC force a form.
  const bool no_persist = getenv("MDX_ATTN_NO_PERSIST") != nullptr;
  const char* ipc_s = getenv("MDX_ATTN_IPC");                 // items per CTA, 0 = one CTA per SM
  const int ipc_env = ipc_s ? atoi(ipc_s) : -1;
  const int sms = mdx_num_sms();
  const long long total = (long long)grid.x * heads * batch;
  const int T = (Nk + AT_BN - 1) / AT_BN;
  const bool persist = ipc_env >= 0 || (!no_persist && T <= 32 && total >= 32ll * sms);
  if (Nk <= AT_BN && !getenv("MDX_ATTN_NO_SINGLE")) {
    // single key tile (cross-attention): with enough items per SM the multi-item kernel pipelines them
    if (!no_persist && total >= 8ll * sms) {
      AttnTcPP pp;
      pp.a = p;
      pp.nqt = (int)grid.x; pp.heads = heads;
      MDX_CHECK_ARG(total < (1ll << 31), "mdx_attention: too many query tiles");
      pp.total = (int)total; pp.ipc = 0; pp.w_mul = 1; pp.w_stride = sms;
      attn_tc_single_persistent_kernel<<<sms, 384, ATP_SMEM, st>>>(mq, mk, mv, pp);
    } else {
      attn_tc_single_kernel<<<grid, 256, AT1_SMEM, st>>>(mq, mk, mv, p);
    }
  } else if (!persist) {
    attn_tc_kernel<<<grid, 384, AT_SMEM, st>>>(mq, mk, mv, p);
  } else {
    AttnTcPP pp;
    pp.a = p;
    pp.nqt = (int)grid.x; pp.heads = heads;
    MDX_CHECK_ARG(total < (1ll << 31), "mdx_attention: too many query tiles");
    pp.total = (int)total;
    if (ipc_env <= 0) {
      const unsigned g = (unsigned)(total < sms ? total : sms);
      pp.ipc = (int)((total + g - 1) / g); pp.w_mul = 1; pp.w_stride = (int)g;
      attn_tc_persistent_kernel<<<g, 384, ATP_SMEM, st>>>(mq, mk, mv, pp);
    } else {
      const long long ipc = ipc_env > 8 ? 8 : ipc_env;
      pp.ipc = (int)ipc; pp.w_mul = (int)ipc; pp.w_stride = 1;
      attn_tc_persistent_kernel<<<(unsigned)((total + ipc - 1) / ipc), 384, ATP_SMEM, st>>>(mq, mk, mv, pp);
    }
  }
  g_mdx_launches.fetch_add(1, std::memory_order_relaxed);
  MDX_CHECK_LAUNCH();
  return MDX_OK;
}
