// Implicit-GEMM convolution on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM, operands staged in
// shared memory by TMA) for sm_100a.  bf16 channels-last activations, fp32 accumulate.
//
// Replaces the cuDNN / cuBLAS calls behind nn.Conv2d 3x3 / 1x1 and NIN (models/layers.py:96-118, 537-546) in
// NCSNpp.forward / DDPM.forward, with the elementwise work around them fused into the epilogue:
//     out = ((conv(x) + bias[n] + temb[b, n]) + residual) * out_scale
// (Dense_0 time-embedding bias layerspp.py:263, skip add + 1/sqrt(2) layerspp.py:271-274), and the skip-connection
// concat (ncsnpp.py:318) folded into the K loop as a second TMA source.
//
// GEMM view: M = B*H*W output pixels, N = Cout, K = taps * (C1 + C2).  stride 1, "same" padding.
//   A tile (128 pixels x 64 channels of ONE filter tap) = one 4-D TMA box {64 c, TW w, TH h, TB b} over the NHWC
//     activation, started at (c0, w0 + s - pad, h0 + r - pad, b0): the zero padding of the convolution is TMA's
//     out-of-bounds fill, so there is no im2col buffer and no bounds logic in the kernel.  128B swizzle makes the
//     box land in the canonical K-major SW128 layout that the UMMA shared-memory descriptor expects.
//   B tile (BN output channels x 64 k) = a 2-D TMA box over the packed weights [Cout][taps*(C1+C2)].
//   D tile (128 x BN fp32) lives in TMEM; two accumulator buffers so the epilogue of tile i overlaps the MMAs of
//     tile i+1.
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected thread),
// warps 2..9 = epilogue (tcgen05.ld -> bias/temb/residual/scale -> bf16 -> global).  Persistent: grid = min(tiles,
// 148), static round-robin tile schedule.
#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "tc_common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int TC_BM = 128;      // pixels per tile (UMMA M)
constexpr int TC_BK = 64;       // channels per stage (64 bf16 = 128 B = one swizzle row)
constexpr int TC_THREADS = 320;

struct ConvTcParams {
  int B, H, W;                  // activation extent (stride 1: output extent too)
  int C1, C2, Cout;             // Cout = channels of the output tensor (multiple of BN)
  int C3, C4;                   // shortcut sources (0 = none): extra K blocks at the CENTRE tap -- the 1x1 `Conv_2(x)` of a
                                // residual block folded into Conv_1's K loop; weights [Cout][taps*(C1+C2) + C3 + C4]
  int taps_w, pad, stride;      // KW (1 or 3) -- KH == KW; pad = KW / 2 (0 for the stride-2 VALID convs)
  int TW, TH, TB;               // tile box: TW * TH * TB == 128
  int tiles_w, tiles_h;         // M tiles per image row / per image column (TB == 1), else 1
  int num_m, num_n;             // tile counts
  const float* bias;            // [Cout] or null
  const float* temb;            // [B][temb_stride] or null
  int temb_stride;
  const bf16* res;              // [B, H, W, Cout] or null
  float out_scale;
  bf16* out;                    // [B, H, W, Cout]                        (EPI 0)
  float* out_nchw;              // [B, cout_true, H, W] fp32, * sample_mul[b]  (EPI 1: network head, ncsnpp.py:371-379)
  int narrow;                   // EPI 1: 0 = NCHW head above; 8 = [B, H, W, 8] bf16 (+ res of the same layout), the
                                // 3-channel image pyramids of progressive='output_skip' (ncsnpp.py:343-365)
  const float* sample_mul;
  int cout_true;
  // optional by-product for the consumer GroupNorm: per-(image, slot, channel) sum and sum of squares of the bf16
  // OUTPUT values, [B][slots][Cout][2] fp32; slots = 2 * tiles per image when a tile lies inside one image, else 1
  float* stats;
  int stat_slots;
  // GNF variants: the A operand is GroupNorm + SiLU of the raw source(s), applied in shared memory between the TMA load
  // and the MMA.  [B][C1 + C2][2] fp32 = (scale / 2, shift / 2) per (image, channel) from b200_gn_coef.
  const float* gn_coef;
  // batched GEMM form (b200_gemm_tc_splitk): the B operand differs per "image": 3-D weight map {K, N, image}.
  // w_batched == 2 (b200_wgrad_tc): both operands come straight from channels-last tensors and are consumed MN-major
  // (rows = the 64 pixels of the K block, 128-byte rows = 64 channels): A = 64-channel boxes of dY [P][Cout] (mapA1, 2-D),
  // B = 4-D boxes {64 c, w, h, b} of the conv's source(s) (mapX1 / mapX2) shifted by the tap, zero outside the image
  int w_batched;
  int wg_H, wg_W, wg_ks, wg_C1, wg_C2;
  long long wg_Kc;
};

// DS (1x1 convolutions: a handful of K blocks per tile, so the kernel is bound by its epilogue, not by the operand
// pipeline): one operand stage fewer and TWO output staging buffers, so that a pass never waits for the previous
// pass's TMA store to drain the buffer and the residual tile of the next pass is in flight during this one.
template <int BN, bool HALO = false, bool DS = false>
struct TcSmem {
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;            // 16 KB
  static constexpr int B_BYTES = BN * TC_BK * 2;               // 32 KB (BN = 256)
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = ((BN == 256) ? 4 : (BN == 128 ? 6 : 8)) - (DS ? 1 : 0);
  static constexpr int TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;
  // output staging for the TMA-store epilogue: BN/64 boxes of [128 pixels x 64 channels] bf16 (SW128), EPI 0 only
  static constexpr int OUT_OFF = STAGES * STAGE_BYTES;
  static constexpr int OUT_BUF = TC_BM * 128 * 2;              // 128 columns per pass (2 passes for BN = 256)
  static constexpr int OUT_BYTES = (BN >= 128) ? OUT_BUF * (DS ? 2 : 1) : 0;
  static constexpr int BAR_OFF = OUT_OFF + OUT_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024 + 1024;   // barriers, bias+temb table, alignment slack           // barriers + tmem slot, + slack for 1024 B alignment
};

// Halo variant (3x3, BN = 128, tile = 8 wide x 16 high pixels of one image).  With one TMA box per tap an N = 128 tile
// re-fetches its activations 9 times and is bound by SM ingress (~38 B/clk of unique activation bytes measured,
// profiles/r01_conv128_ncu.txt) instead of the tensor pipe.  Here each 64-channel K block loads THREE boxes
// {64 c, 8 w, 18 h} (one per horizontal tap, rows h0-1 .. h0+16) and the vertical taps are descriptor offsets of
// r * 8 pixels * 128 B = r * 1024 B into the same box -- still 1024-byte aligned, so the SW128 phase is unchanged.
// Activation ingress drops from 9 x 16 KB to 3 x 18 KB per K block (x 3.4 instead of x 9).  Weights ride a separate
// ring (one 16 KB box per tap) filled by their own producer warp.
constexpr int HALO_TW = 8, HALO_TH = 16;
template <int BN, bool DS>
struct TcSmem<BN, true, DS> {
  static constexpr int BOX_BYTES = (HALO_TH + 2) * HALO_TW * 128;      // 18 KB
  static constexpr int A_BYTES = 3 * BOX_BYTES;                        // 54 KB per K block
  static constexpr int B_BYTES = BN * TC_BK * 2;
  static constexpr int STAGES = 2;                                     // activation stages
  static constexpr int WGROUP = 1;                                     // taps per weight-ring stage
  static constexpr int WSTAGES = 5;                                    // weight ring (5 x 16 KB: what fits beside 2 x 54 KB)
  static constexpr int W_OFF = STAGES * A_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int OUT_OFF = W_OFF + WSTAGES * B_BYTES;
  static constexpr int OUT_BUF = TC_BM * 128 * 2;
  static constexpr int OUT_BYTES = OUT_BUF;
  static constexpr int BAR_OFF = OUT_OFF + OUT_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024 + 1024;   // barriers, bias+temb table, alignment slack
  static constexpr int STAGE_BYTES = 0;
};

// Pair variant (cta_group::2, Cout tile 256): a 128 x 256 tile streams 48 KB per K block per SM, two thirds of it
// weights that every CTA fetches again for every tile, and L2 cannot feed that to 148 SMs (the MMA issuer waited 33-36 %
// of its time for operand stages, profiles/r01f_conv_tc_roles.txt).  Two SMs of a TPC now run one 256 x 256 tile: each
// loads its own 128 pixels and HALF of the weight tile (32 KB per K block per SM), and six stages fit.
struct TcSmemPair {
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;            // 16 KB
  static constexpr int B_BYTES = 128 * TC_BK * 2;              // 16 KB: this CTA's half of the 256 weight rows
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 6;
  static constexpr int TMEM_COLS = 512;
  static constexpr int OUT_OFF = STAGES * STAGE_BYTES;
  static constexpr int OUT_BUF = TC_BM * 128 * 2;
  static constexpr int OUT_BYTES = OUT_BUF;
  static constexpr int BAR_OFF = OUT_OFF + OUT_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024 + 1024;   // barriers, bias+temb table, alignment slack
};
// Pair + halo (Cout tile 128): a single SM's 128 x 128 x 16 MMA reads 4 KB of A and 4 KB of B per 64 clk, i.e. all of
// the 128 B/clk shared memory delivers; as half of a 256 x 128 pair MMA it reads 4 + 2 KB.  Each CTA loads 64 of the
// 128 weight rows per tap (8 KB boxes); a ring stage holds one filter ROW (3 taps, 24 KB), so the single MMA-issuing
// thread -- whose per-tap barrier traffic is what paces N = 128 tiles -- waits and commits 6 instead of 18 times per tile.
// BN = 256 (halo + pair): the 256 x 256 pair tile with halo reuse of its activations -- per K block and SM 54 KB of
// activations serve all nine taps (instead of 9 x 16 KB) next to 9 x 16 KB of weights: 1.18 MB -> 0.81 MB of L2 reads
// per tile, for the shapes whose MMA issuer still waited 23 % of its time for operand stages.  Per-tap weight ring
// (four 128-clk MMAs per tap hide the barrier round trip), 5 x 16 KB.
template <bool DS, int BN = 128>
struct TcSmemHaloPairT {
  static constexpr int BOX_BYTES = (HALO_TH + 2) * HALO_TW * 128;
  static constexpr int A_BYTES = 3 * BOX_BYTES;
  static constexpr int B_BYTES = (BN / 2) * TC_BK * 2;         // 8 KB (BN = 128) / 16 KB (BN = 256)
  static constexpr int STAGES = 2;
  static constexpr int WGROUP = BN == 128 ? 3 : 1;             // taps per weight-ring stage: one filter row (BN = 128)
  static constexpr int WSTAGES = BN == 128 ? (DS ? 2 : 3) : 5; // DS: a ring stage makes room for the second staging buffer
  static constexpr int W_OFF = STAGES * A_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int OUT_OFF = W_OFF + WSTAGES * WGROUP * B_BYTES;
  static constexpr int OUT_BUF = TC_BM * 128 * 2;
  static constexpr int OUT_BYTES = OUT_BUF * (DS ? 2 : 1);
  static constexpr int BAR_OFF = OUT_OFF + OUT_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024 + 1024;   // barriers, bias+temb table, alignment slack
  static constexpr int STAGE_BYTES = 0;
};
template <int BN, bool HALO, bool DS, bool PAIR> struct TcSmemSel { using type = TcSmem<BN, HALO, DS>; };
template <int BN, bool DS> struct TcSmemSel<BN, false, DS, true> { using type = TcSmemPair; };
template <int BN, bool DS> struct TcSmemSel<BN, true, DS, true> { using type = TcSmemHaloPairT<DS, BN>; };
using TcSmemHaloPair = TcSmemHaloPairT<false>;

// GNF (halo + pair variants only): GroupNorm + SiLU of the INPUT fused into the operand pipeline.  The conv consumes
// the producer's raw output; four extra warps rewrite every landed activation stage in place,
//     h = fma(x, scale/2, shift/2);  y = fma(h, tanh.approx(h), h)          (= silu(x * scale + shift), bit for bit what
// gn_apply_kernel<bf16> writes), and only then is the stage handed to the MMA issuer.  The standalone GroupNorm pass
// (one read + one write of every activation, ~23 % of the C3 forward) disappears.  The three column-shifted boxes of a
// stage overlap: each of the 10 x 18 unique pixels is transformed ONCE by its owner thread, which reads the raw value
// from the one box copy nobody else reads and stores the result into every box that holds that pixel (<= 3 stores).
// Zero padding stays zero: out-of-image pixels (TMA zero fill) are skipped.
constexpr int TC_XF_WARPS = 4;
template <int BN, int EPI, bool HALO = false, bool DS = false, bool PAIR = false, bool GNF = false>
__global__ void __launch_bounds__(TC_THREADS + (HALO ? 32 : 0) + (GNF ? 32 * TC_XF_WARPS : 0), 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap mapA1, const __grid_constant__ CUtensorMap mapA2,
               const __grid_constant__ CUtensorMap mapW, const __grid_constant__ CUtensorMap mapOut,
               const __grid_constant__ CUtensorMap mapRes, const __grid_constant__ CUtensorMap mapX1,
               const __grid_constant__ CUtensorMap mapX2, const ConvTcParams p) {
  using S = typename TcSmemSel<BN, HALO, DS, PAIR>::type;
  static_assert(!PAIR || (EPI == 0 && ((BN == 256 && !DS) || (BN == 128 && HALO))), "pair variants");
  static_assert(!GNF || (HALO && PAIR), "GroupNorm-fused operand pipeline: halo + pair variants only");
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;   // 0 = leader (issues the MMAs)
  constexpr int NWS = HALO ? (PAIR ? TcSmemHaloPairT<DS, BN>::WSTAGES : TcSmem<BN, true>::WSTAGES) : 0;   // weight ring (halo)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar0 = base + S::BAR_OFF;
  // barrier layout (8 B each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base slot
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (S::STAGES + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * S::STAGES + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * S::STAGES + 2 + a); };
  auto wfull_bar = [&](int s) { return bar0 + 8u * (2 * S::STAGES + 4 + s); };
  auto wempty_bar = [&](int s) { return bar0 + 8u * (2 * S::STAGES + 4 + NWS + s); };
  const uint32_t res_bar = bar0 + 8u * (2 * S::STAGES + 4 + 2 * NWS);      // residual tile landed in the staging buffer
  const uint32_t tmem_slot = res_bar + 8u;
  auto xf_bar = [&](int s) { return res_bar + 16u + 8u * s; };             // GNF: stage transformed (leader's copy counts)
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  // the warp index through a shuffle: provably warp-uniform, so the role branches below are convergent for ptxas
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int Ctot = p.C1 + p.C2;
  const int kb_per_tap = Ctot / TC_BK;
  const int taps = p.taps_w * p.taps_w;
  const int kb_extra = (p.C3 + p.C4) / TC_BK;            // shortcut K blocks (centre tap only)
  const int num_kb = taps * kb_per_tap + (HALO ? 0 : kb_extra);
  const int num_tiles = p.num_m * p.num_n;
  // work items: tiles, round robin over CTAs; PAIR: pairs of M tiles (2 pm, 2 pm + 1), round robin over CTA pairs
  const int tile_first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tile_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int tile_end = PAIR ? ((p.num_m + 1) / 2) * p.num_n : num_tiles;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S::STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), PAIR ? 16 : 8); }   // PAIR: both CTAs' epilogue warps
    for (int w = 0; w < NWS; ++w) { mbar_init(wfull_bar(w), 1); mbar_init(wempty_bar(w), 1); }
    mbar_init(res_bar, 1);
    if constexpr (GNF) for (int s = 0; s < S::STAGES; ++s) mbar_init(xf_bar(s), 2 * TC_XF_WARPS);   // both CTAs' warps
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA1) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW) : "memory");
  }
  if (warp == 1) {
    if constexpr (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(S::TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(S::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();   // PAIR: the peer's barriers are initialised too
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  b200_pdl_wait();   // everything above overlapped the previous kernel's tail; its outputs are read from here on
  b200_pdl_trigger();

  auto tile_coords = [&](int tile, int& mt, int& nt) {
    nt = tile % p.num_n; mt = tile / p.num_n;
    if constexpr (PAIR) mt = 2 * mt + (int)cta_rank;      // an odd tail tile lies beyond the batch: TMA zero-fills / clips
  };
  auto m_origin = [&](int mt, int& b0, int& h0, int& w0) {
    if (p.TB > 1) { b0 = mt * p.TB; h0 = 0; w0 = 0; return; }
    const int per_img = p.tiles_w * p.tiles_h;
    b0 = mt / per_img;
    const int r = mt - b0 * per_img;
    h0 = (r / p.tiles_w) * p.TH;
    w0 = (r % p.tiles_w) * p.TW;
  };

  if (warp == 0) {
    // ================= TMA producer =================
    {  // the whole warp runs the loop and waits; one elected lane issues (keeps descriptors in uniform registers)
      const bool elected = tc_elect_one();
      int stage = 0; uint32_t phase = 0;
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        int mt, nt, b0, h0, w0;
        tile_coords(tile, mt, nt);
        m_origin(mt, b0, h0, w0);
        if constexpr (HALO) {
          // activations only: three column-shifted boxes with one halo row above and below per 64-channel block
          for (int kb = 0; kb < kb_per_tap; ++kb) {
            mbar_wait(empty_bar(stage), phase ^ 1);
            if (elected) {
            const uint32_t sa = base + stage * S::A_BYTES;
            const int c = kb * TC_BK;
            // (GNF: each CTA's boxes complete on its OWN barrier -- its own transform warps consume them first)
            if constexpr (PAIR && !GNF) {
              const uint32_t lbar = mapa_cta(full_bar(stage), 0);
              if (cta_rank == 0) mbar_expect_tx(full_bar(stage), 2 * S::A_BYTES);
#pragma unroll
              for (int sx = 0; sx < 3; ++sx) {
                if (c < p.C1) tma_load_4d_2sm(sa + sx * S::BOX_BYTES, &mapA1, lbar, c, w0 + sx - 1, h0 - 1, b0);
                else tma_load_4d_2sm(sa + sx * S::BOX_BYTES, &mapA2, lbar, c - p.C1, w0 + sx - 1, h0 - 1, b0);
              }
            } else {
            mbar_expect_tx(full_bar(stage), S::A_BYTES);
#pragma unroll
            for (int sx = 0; sx < 3; ++sx) {
              if (c < p.C1) tma_load_4d(sa + sx * S::BOX_BYTES, &mapA1, full_bar(stage), c, w0 + sx - 1, h0 - 1, b0);
              else tma_load_4d(sa + sx * S::BOX_BYTES, &mapA2, full_bar(stage), c - p.C1, w0 + sx - 1, h0 - 1, b0);
            }
            }
            }
            if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
          }
          // shortcut K blocks: ONE box (the centre column shift; its row 1 is the centre tap) of the shortcut source
          for (int kx = 0; kx < kb_extra; ++kx) {
            mbar_wait(empty_bar(stage), phase ^ 1);
            if (elected) {
              const uint32_t sa = base + stage * S::A_BYTES + S::BOX_BYTES;
              const int c = kx * TC_BK;
              if constexpr (PAIR && !GNF) {
                const uint32_t lbar = mapa_cta(full_bar(stage), 0);
                if (cta_rank == 0) mbar_expect_tx(full_bar(stage), 2 * S::BOX_BYTES);
                if (c < p.C3) tma_load_4d_2sm(sa, &mapX1, lbar, c, w0, h0 - 1, b0);
                else tma_load_4d_2sm(sa, &mapX2, lbar, c - p.C3, w0, h0 - 1, b0);
              } else {
                mbar_expect_tx(full_bar(stage), S::BOX_BYTES);
                if (c < p.C3) tma_load_4d(sa, &mapX1, full_bar(stage), c, w0, h0 - 1, b0);
                else tma_load_4d(sa, &mapX2, full_bar(stage), c - p.C3, w0, h0 - 1, b0);
              }
            }
            if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
          }
        } else {
        for (int tap = 0; tap < taps; ++tap) {
          const int r = tap / p.taps_w, s = tap - r * p.taps_w;
          for (int kb = 0; kb < kb_per_tap; ++kb) {
            mbar_wait(empty_bar(stage), phase ^ 1);
            if (elected) {
            const uint32_t sa = base + stage * S::STAGE_BYTES, sb = sa + S::A_BYTES;
            const int c = kb * TC_BK;
            if constexpr (PAIR) {
              // both CTAs' boxes complete on the LEADER's barrier, which expects the bytes of the whole pair
              const uint32_t lbar = mapa_cta(full_bar(stage), 0);
              if (cta_rank == 0) mbar_expect_tx(full_bar(stage), 2 * S::STAGE_BYTES);
              if (c < p.C1) tma_load_4d_2sm(sa, &mapA1, lbar, c, w0 * p.stride + s - p.pad, h0 * p.stride + r - p.pad, b0);
              else tma_load_4d_2sm(sa, &mapA2, lbar, c - p.C1, w0 * p.stride + s - p.pad, h0 * p.stride + r - p.pad, b0);
              tma_load_2d_2sm(sb, &mapW, lbar, tap * Ctot + c, nt * BN + (int)cta_rank * 128);
            } else {
            mbar_expect_tx(full_bar(stage), S::STAGE_BYTES);
            if (p.w_batched == 2) {
              // weight gradient: K block = 64 pixels starting at p0; 64-channel boxes, 8 KB each, MN-major
              const long long p0 = (long long)b0 * p.wg_Kc + c;
              const int hw = p.wg_H * p.wg_W;
              const int ib = (int)(p0 / hw), rem = (int)(p0 - (long long)ib * hw);
              const int ih = rem / p.wg_W, iw = rem - ih * p.wg_W;
              const int cin = p.wg_C1 + p.wg_C2, wpad = p.wg_ks / 2;
#pragma unroll
              for (int g = 0; g < TC_BM / 64; ++g) tma_load_2d(sa + g * 8192, &mapA1, full_bar(stage), w0 + g * 64, (int)p0);
#pragma unroll
              for (int g = 0; g < BN / 64; ++g) {
                const int n0 = nt * BN + g * 64;
                const int wtap = n0 / cin, cc = n0 - wtap * cin;
                const int ry = wtap / p.wg_ks, sx = wtap - ry * p.wg_ks;
                if (cc < p.wg_C1) tma_load_4d(sb + g * 8192, &mapX1, full_bar(stage), cc, iw + sx - wpad, ih + ry - wpad, ib);
                else tma_load_4d(sb + g * 8192, &mapX2, full_bar(stage), cc - p.wg_C1, iw + sx - wpad, ih + ry - wpad, ib);
              }
            } else {
            if (c < p.C1) tma_load_4d(sa, &mapA1, full_bar(stage), c, w0 * p.stride + s - p.pad, h0 * p.stride + r - p.pad, b0);
            else tma_load_4d(sa, &mapA2, full_bar(stage), c - p.C1, w0 * p.stride + s - p.pad, h0 * p.stride + r - p.pad, b0);
            if (p.w_batched) tma_load_3d(sb, &mapW, full_bar(stage), tap * Ctot + c, nt * BN, b0);
            else tma_load_2d(sb, &mapW, full_bar(stage), tap * Ctot + c, nt * BN);
            }
            }
            }
            if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
          }
        }
        for (int kx = 0; kx < kb_extra; ++kx) {          // shortcut K blocks: centre tap of the shortcut source
          mbar_wait(empty_bar(stage), phase ^ 1);
          if (elected) {
            const uint32_t sa = base + stage * S::STAGE_BYTES, sb = sa + S::A_BYTES;
            const int c = kx * TC_BK;
            if constexpr (PAIR) {
              const uint32_t lbar = mapa_cta(full_bar(stage), 0);
              if (cta_rank == 0) mbar_expect_tx(full_bar(stage), 2 * S::STAGE_BYTES);
              if (c < p.C3) tma_load_4d_2sm(sa, &mapX1, lbar, c, w0 * p.stride, h0 * p.stride, b0);
              else tma_load_4d_2sm(sa, &mapX2, lbar, c - p.C3, w0 * p.stride, h0 * p.stride, b0);
              tma_load_2d_2sm(sb, &mapW, lbar, taps * Ctot + c, nt * BN + (int)cta_rank * 128);
            } else {
              mbar_expect_tx(full_bar(stage), S::STAGE_BYTES);
              if (c < p.C3) tma_load_4d(sa, &mapX1, full_bar(stage), c, w0 * p.stride, h0 * p.stride, b0);
              else tma_load_4d(sa, &mapX2, full_bar(stage), c - p.C3, w0 * p.stride, h0 * p.stride, b0);
              tma_load_2d(sb, &mapW, full_bar(stage), taps * Ctot + c, nt * BN);
            }
          }
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
        }
      }
    }
  } else if (HALO && warp == 10) {
    // ================= weight producer (halo variant): one [BN x 64] box per (K block, tap) =================
    if constexpr (HALO) {
      {
        const bool elected = tc_elect_one();
        int ws = 0; uint32_t wphase = 0;
        for (int tile = tile_first; tile < tile_end; tile += tile_step) {
          int mt, nt;
          tile_coords(tile, mt, nt);
          for (int kb = 0; kb < kb_per_tap; ++kb)
            for (int tap = 0; tap < 9; tap += S::WGROUP) {
              mbar_wait(wempty_bar(ws), wphase ^ 1);
              if (elected) {
              const uint32_t wdst = base + S::W_OFF + ws * (S::WGROUP * S::B_BYTES);
              if constexpr (PAIR) {
                const uint32_t lbar = mapa_cta(wfull_bar(ws), 0);
                if (cta_rank == 0) mbar_expect_tx(wfull_bar(ws), 2 * S::WGROUP * S::B_BYTES);
#pragma unroll
                for (int g = 0; g < S::WGROUP; ++g)
                  tma_load_2d_2sm(wdst + g * S::B_BYTES, &mapW, lbar, (tap + g) * Ctot + kb * TC_BK,
                                  nt * BN + (int)cta_rank * (BN / 2));
              } else {
              mbar_expect_tx(wfull_bar(ws), S::WGROUP * S::B_BYTES);
#pragma unroll
              for (int g = 0; g < S::WGROUP; ++g)
                tma_load_2d(wdst + g * S::B_BYTES, &mapW, wfull_bar(ws), (tap + g) * Ctot + kb * TC_BK, nt * BN);
              }
              }
              if (++ws == NWS) { ws = 0; wphase ^= 1; }
            }
          for (int kx = 0; kx < kb_extra; ++kx) {        // shortcut K blocks: one weight tile per ring stage
            mbar_wait(wempty_bar(ws), wphase ^ 1);
            if (elected) {
              const uint32_t wdst = base + S::W_OFF + ws * (S::WGROUP * S::B_BYTES);
              if constexpr (PAIR) {
                const uint32_t lbar = mapa_cta(wfull_bar(ws), 0);
                if (cta_rank == 0) mbar_expect_tx(wfull_bar(ws), 2 * S::B_BYTES);
                tma_load_2d_2sm(wdst, &mapW, lbar, 9 * Ctot + kx * TC_BK, nt * BN + (int)cta_rank * (BN / 2));
              } else {
                mbar_expect_tx(wfull_bar(ws), S::B_BYTES);
                tma_load_2d(wdst, &mapW, wfull_bar(ws), 9 * Ctot + kx * TC_BK, nt * BN);
              }
            }
            if (++ws == NWS) { ws = 0; wphase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    // The WHOLE warp runs the loop (waits included) and one elected lane issues: inside an `if (lane == 0)` region
    // ptxas cannot keep the descriptors in uniform registers and wraps every UTCHMMA in an ELECT / 6 x R2UR.BROADCAST /
    // BRA.U.ANY "waterfall" -- ~86 clk per issued MMA against the 64 clk an N = 128 MMA takes to execute.
    if (cta_rank == 0) {
      const bool elected = tc_elect_one();
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      // instruction descriptor: D = f32, A = B = bf16, both K-major, N = BN, M = 128 (PAIR: 256 over the two CTAs)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)((PAIR ? 2 * TC_BM : TC_BM) >> 4) << 24);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      int ws = 0; uint32_t wphase = 0;          // weight ring (halo variant)
      (void)ws; (void)wphase;
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_u + (uint32_t)(acc * BN);
        if constexpr (HALO) {
          for (int kb = 0; kb < kb_per_tap; ++kb) {
            if constexpr (GNF) mbar_wait_cluster(xf_bar(stage), phase); else mbar_wait(full_bar(stage), phase);
            tc_fence_after();
            const uint32_t sa = base + stage * S::A_BYTES;
            const uint64_t da0 = make_sw128_desc(sa);
#pragma unroll
            for (int tap0 = 0; tap0 < 9; tap0 += S::WGROUP) {
              mbar_wait(wfull_bar(ws), wphase);
              tc_fence_after();
              const uint64_t db0 = make_sw128_desc(base + S::W_OFF + ws * (S::WGROUP * S::B_BYTES));
              if (elected) {
#pragma unroll
              for (int g = 0; g < S::WGROUP; ++g) {
                const int tap = tap0 + g, r = tap / 3, sx = tap - r * 3;
                // descriptor start addresses are in 16-byte units: constant offsets per tap once unrolled
                const uint64_t da = da0 + (uint64_t)((sx * S::BOX_BYTES + r * (HALO_TW * 128)) >> 4);
                const uint64_t db = db0 + (uint64_t)((g * S::B_BYTES) >> 4);
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k) {
                  if constexpr (PAIR) tc_mma_bf16_2sm(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | tap | k) != 0 ? 1u : 0u);
                  else tc_mma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | tap | k) != 0 ? 1u : 0u);
                }
              }
              if constexpr (PAIR) tc_commit_2sm(wempty_bar(ws)); else tc_commit(wempty_bar(ws));
              }
              if (++ws == NWS) { ws = 0; wphase ^= 1; }
            }
            if (elected) { if constexpr (PAIR) tc_commit_2sm(empty_bar(stage)); else tc_commit(empty_bar(stage)); }
            if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
          }
          for (int kx = 0; kx < kb_extra; ++kx) {        // shortcut K blocks: centre tap = box 1, row offset 1
            if constexpr (GNF) mbar_wait_cluster(xf_bar(stage), phase); else mbar_wait(full_bar(stage), phase);
            tc_fence_after();
            mbar_wait(wfull_bar(ws), wphase);
            tc_fence_after();
            const uint64_t da = make_sw128_desc(base + stage * S::A_BYTES + S::BOX_BYTES + HALO_TW * 128);
            const uint64_t db = make_sw128_desc(base + S::W_OFF + ws * (S::WGROUP * S::B_BYTES));
            if (elected) {
#pragma unroll
              for (int k = 0; k < TC_BK / 16; ++k) {
                if constexpr (PAIR) tc_mma_bf16_2sm(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, 1u);
                else tc_mma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, 1u);
              }
              if constexpr (PAIR) { tc_commit_2sm(wempty_bar(ws)); tc_commit_2sm(empty_bar(stage)); }
              else { tc_commit(wempty_bar(ws)); tc_commit(empty_bar(stage)); }
            }
            if (++ws == NWS) { ws = 0; wphase ^= 1; }
            if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
          }
        } else {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = base + stage * S::STAGE_BYTES, sb = sa + S::A_BYTES;
          const uint64_t da = make_sw128_desc(sa), db = make_sw128_desc(sb);
          if (elected) {
          if (!PAIR && p.w_batched == 2) {
            // both operands MN-major (weight gradient): a K step of 16 pixels = 16 rows of 128 B = 2 KB; the 64-channel
            // groups along M / N are 8 KB apart (LBO), the 8-pixel atoms along K 1 KB (SBO)
            const uint32_t idesc_mn = idesc | (1u << 15) | (1u << 16);
#pragma unroll
            for (int k = 0; k < TC_BK / 16; ++k)
              tc_mma_bf16(d_tmem, make_sw128_desc_ex(sa + k * 2048, 8192, 1024), make_sw128_desc_ex(sb + k * 2048, 8192, 1024),
                          idesc_mn, (kb | k) != 0 ? 1u : 0u);
          } else {
#pragma unroll
          for (int k = 0; k < TC_BK / 16; ++k) {
            // advance 16 elements (32 B) along K inside the 128 B swizzle row: +2 in the 16-byte-unit address field
            if constexpr (PAIR) tc_mma_bf16_2sm(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
            else tc_mma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
          }
          }
          if constexpr (PAIR) tc_commit_2sm(empty_bar(stage)); else tc_commit(empty_bar(stage));
          }
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
        }
        if (elected) { if constexpr (PAIR) tc_commit_2sm(tfull_bar(acc)); else tc_commit(tfull_bar(acc)); }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp < 10) {
    // ================= epilogue warps (2..9): two per TMEM lane quadrant, each takes half of the columns ==========
    const int q = warp & 3;                        // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;
    const int m_local = q * 32 + lane;
    int acc = 0; uint32_t acc_phase = 0;
    // Residual (skip connection, EPI 0): a thread-per-pixel-row read of `res` touches 32 different 128-byte lines per
    // warp instruction and made the epilogue of conv_1 / NIN_3 / the 1x1 shortcuts the bottleneck of the kernel
    // (profiles/r01f_conv_tc_roles.txt).  Instead the residual tile is fetched by TMA INTO the output staging buffer
    // (same box, same SW128 layout) one pass ahead; each thread then reads the 16-byte units it is about to overwrite.
    uint32_t res_phase = 0;
    auto res_fetch = [&](int tile_, int pass_, int buf_) {           // thread 64 only
      int mt_, nt_, b0_, h0_, w0_;
      tile_coords(tile_, mt_, nt_);
      m_origin(mt_, b0_, h0_, w0_);
      const uint32_t dst = base + S::OUT_OFF + buf_ * S::OUT_BUF;
      mbar_expect_tx(res_bar, 2 * TC_BM * 128);
      tma_load_4d(dst, &mapRes, res_bar, nt_ * BN + pass_ * 128, w0_, h0_, b0_);
      tma_load_4d(dst + TC_BM * 128, &mapRes, res_bar, nt_ * BN + pass_ * 128 + 64, w0_, h0_, b0_);
    };
    if constexpr (EPI == 0) {
      if (p.res && threadIdx.x == 64 && tile_first < tile_end) res_fetch(tile_first, 0, 0);
    }
    int pcount = 0;                                      // passes done by this CTA (DS: staging buffer = pcount & 1)
    // bias[n] + temb[b, n] of the tile's BN columns (tiles inside one image, TB == 1): the 16 broadcast global loads
    // per 32-column chunk sat on the critical path of every epilogue-bound tile (long-scoreboard stalls on the first
    // FADDs, profiles/r01g_conv_tc_ncu_full.txt).  One thread per column now fetches the NEXT tile's value while this
    // tile is drained and drops it into a 1 KB shared table after the last read of the current one.
    float* btab = reinterpret_cast<float*>(smem_raw + (bar0 - smem_u32(smem_raw)) + 256);
    const bool use_tab = EPI == 0 && p.TB == 1 && (p.bias || p.temb);
    const int et_col = (int)threadIdx.x - 64;            // column of the table this thread maintains (< BN)
    auto tab_value = [&](int tile_) -> float {
      int mt_, nt_, b0_, h0_, w0_;
      tile_coords(tile_, mt_, nt_);
      m_origin(mt_, b0_, h0_, w0_);
      float v = 0.0f;
      if (et_col < BN && b0_ < p.B) {
        if (p.bias) v = __ldg(p.bias + nt_ * BN + et_col);
        if (p.temb) v += __ldg(p.temb + (size_t)b0_ * p.temb_stride + nt_ * BN + et_col);
      }
      return v;
    };
    if (use_tab && tile_first < tile_end && et_col < BN) btab[et_col] = tab_value(tile_first);
    for (int tile = tile_first; tile < tile_end; tile += tile_step) {
      int mt, nt, b0, h0, w0;
      tile_coords(tile, mt, nt);
      m_origin(mt, b0, h0, w0);
      const bool tab_more = use_tab && tile + tile_step < tile_end;
      float tab_next = 0.0f;
      if (tab_more) tab_next = tab_value(tile + tile_step);   // in flight while this tile is drained
      const int lw = m_local % p.TW, lh = (m_local / p.TW) % p.TH, lb = m_local / (p.TW * p.TH);
      const int b = b0 + lb, h = h0 + lh, w = w0 + lw;
      const bool valid = b < p.B;
      const float* tb = (p.temb && valid) ? p.temb + (size_t)b * p.temb_stride + nt * BN : nullptr;
      const float* bs = p.bias ? p.bias + nt * BN : nullptr;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      if (use_tab) asm volatile("bar.sync 1, 256;" ::: "memory");   // the table holds this tile's values
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if constexpr (EPI == 1) {
        uint32_t r[16];
        tc_ld16(t_addr, r);
        tc_wait_ld();
        if (valid && half == 0 && p.narrow == 8) {
          const size_t px = (((size_t)b * p.H + h) * p.W + w) * 8;
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[j]) + (bs ? __ldg(bs + j) : 0.0f);
          if (p.res) {
            const uint4 rr = __ldg(reinterpret_cast<const uint4*>(p.res + px));
            const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&rr);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 f = __bfloat1622float2(hh[i]);
              v[2 * i] += f.x; v[2 * i + 1] += f.y;
            }
          }
          uint4 t;
          __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
          for (int i = 0; i < 4; ++i) ho[i] = __floats2bfloat162_rn(v[2 * i] * p.out_scale, v[2 * i + 1] * p.out_scale);
          *reinterpret_cast<uint4*>(p.out + px) = t;
        } else if (valid && half == 0) {
          const float mul = (p.sample_mul ? __ldg(p.sample_mul + b) : 1.0f) * p.out_scale;
          float* op = p.out_nchw + (((size_t)b * p.cout_true) * p.H + h) * p.W + w;
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j < p.cout_true)
              op[(size_t)j * p.H * p.W] = (__uint_as_float(r[j]) + (bs ? __ldg(bs + j) : 0.0f)) * mul;
        }
      } else {
      constexpr int CH = 64;                           // columns per warp per pass (two warps share a lane quadrant)
#pragma unroll 1
      for (int pass = 0; pass < BN / 128; ++pass) {
      const int cbeg = pass * 128 + half * CH;
      const int sofs = S::OUT_OFF + (DS ? (pcount & 1) * S::OUT_BUF : 0);   // this pass's staging buffer
      if (p.res) {
        mbar_wait(res_bar, res_phase);                 // staging buffer holds this pass's residual tile
        res_phase ^= 1;
      }
#pragma unroll 1
      for (int c0 = cbeg; c0 < cbeg + CH; c0 += 32) {
        uint32_t r[32];
        tc_ld32(t_addr + c0, r);
        // this thread's row of the [128 x 64] SW128 box of the 64-channel group: unit u of row r sits at u ^ (r & 7)
        uint8_t* srow = smem_raw + (base - smem_u32(smem_raw)) + sofs + ((c0 >> 6) & 1) * (TC_BM * 128) + m_local * 128;
        uint4 rv[4];
        if (p.res) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            rv[j] = *reinterpret_cast<const uint4*>(srow + (((((c0 & 63) >> 3) + j) ^ (m_local & 7)) << 4));
        }
        tc_wait_ld();
        if (!p.res && c0 == cbeg) {                      // (uniform) staging buffer free: the previous store has read it
          if (threadIdx.x == 64) {
            if constexpr (DS) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the store before last
            else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          }
          asm volatile("bar.sync 1, 256;" ::: "memory");
        }
        if (valid) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          if (use_tab) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 t = *reinterpret_cast<const float4*>(btab + c0 + j);
              v[j] += t.x; v[j + 1] += t.y; v[j + 2] += t.z; v[j + 3] += t.w;
            }
          } else {
          if (bs) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 t = __ldg(reinterpret_cast<const float4*>(bs + c0 + j));
              v[j] += t.x; v[j + 1] += t.y; v[j + 2] += t.z; v[j + 3] += t.w;
            }
          }
          if (tb) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 t = __ldg(reinterpret_cast<const float4*>(tb + c0 + j));
              v[j] += t.x; v[j + 1] += t.y; v[j + 2] += t.z; v[j + 3] += t.w;
            }
          }
          }
          if (p.res) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&rv[j]);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float2 f = __bfloat1622float2(hh[i]);
                v[j * 8 + 2 * i] += f.x; v[j * 8 + 2 * i + 1] += f.y;
              }
            }
          }
          // stage into the box (over the residual units just read)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 t;
            __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              hh[i] = __floats2bfloat162_rn(v[j * 8 + 2 * i] * p.out_scale, v[j * 8 + 2 * i + 1] * p.out_scale);
            const int unit = ((c0 & 63) >> 3) + j;
            *reinterpret_cast<uint4*>(srow + ((unit ^ (m_local & 7)) << 4)) = t;
          }
        }
      }
      if (pass == BN / 128 - 1) {                      // last TMEM read of this tile done: hand the accumulator back
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (PAIR) mbar_arrive_cluster(mapa_cta(tempty_bar(acc), 0));   // the leader's barrier counts both CTAs
          else mbar_arrive(tempty_bar(acc));
        }
      }
      fence_proxy_async_smem();                        // generic-proxy smem writes -> visible to the TMA store
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (tab_more && pass == BN / 128 - 1 && et_col < BN) btab[et_col] = tab_next;
      if (threadIdx.x == 64) {
        tma_store_4d(base + sofs, &mapOut, nt * BN + pass * 128, w0, h0, b0);
        tma_store_4d(base + sofs + TC_BM * 128, &mapOut, nt * BN + pass * 128 + 64, w0, h0, b0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if constexpr (DS) {
          // the OTHER buffer is idle once the previous pass's store has read it (every thread is past that pass's
          // column sums: it went through the barrier above): fetch the next pass's residual tile into it now
          const bool last_pass = pass == BN / 128 - 1;
          if (p.res && (!last_pass || tile + tile_step < tile_end)) {
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            if (last_pass) res_fetch(tile + tile_step, 0, (pcount + 1) & 1); else res_fetch(tile, pass + 1, (pcount + 1) & 1);
          }
        }
      }
      if (p.stats) {
        // column sums over the staged tile: thread = (column, 64-row half); fixed order -> bit-reproducible
        const int et = threadIdx.x - 64, col = et & 127, rh = et >> 7;
        const uint8_t* sbox = smem_raw + (base - smem_u32(smem_raw)) + sofs + (col >> 6) * (TC_BM * 128);
        const int unit = (col & 63) >> 3, sub = (col & 7) * 2;
        const int rows_img = p.TW * p.TH;                  // rows of one image inside the tile (128 when TB == 1)
        const int seg = rows_img < 64 ? rows_img : 64;
        const int ch = nt * BN + pass * 128 + col;
        // row r of the box keeps this column at r * 128 + ((unit ^ (r & 7)) << 4) + sub; segments start at multiples of
        // 8 rows, so the eight swizzled offsets of an 8-row group are fixed per thread.  Even / odd rows accumulate
        // separately (two dependency chains instead of one); the order is fixed, so results stay reproducible.
        int off8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) off8[i] = i * 128 + ((unit ^ i) << 4) + sub;
        for (int r0 = rh * 64; r0 < rh * 64 + 64; r0 += seg) {
          float s1 = 0.f, s2 = 0.f, t1 = 0.f, t2 = 0.f;
          if (seg < 8) {                                   // images smaller than 8 pixels: plain loop
            for (int r = r0; r < r0 + seg; ++r) {
              const float f = __bfloat162float(*reinterpret_cast<const bf16*>(sbox + r * 128 + ((unit ^ (r & 7)) << 4) + sub));
              s1 += f; s2 = fmaf(f, f, s2);
            }
          } else
          for (int r = r0; r < r0 + seg; r += 8) {
            const uint8_t* rb = sbox + r * 128;
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = __bfloat162float(*reinterpret_cast<const bf16*>(rb + off8[i]));
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
              s1 += f[i]; s2 = fmaf(f[i], f[i], s2);
              t1 += f[i + 1]; t2 = fmaf(f[i + 1], f[i + 1], t2);
            }
          }
          s1 += t1; s2 += t2;
          int bi, slot;
          if (p.TB > 1) { bi = b0 + r0 / rows_img; slot = 0; }
          else { bi = b0; slot = ((h0 / p.TH) * p.tiles_w + w0 / p.TW) * 2 + rh; }
          if (bi < p.B) {
            float* sp = p.stats + (((size_t)bi * p.stat_slots + slot) * p.Cout + ch) * 2;
            sp[0] = s1; sp[1] = s2;
          }
        }
      }
      if (!DS && p.res) {
        // next pass's residual tile: once the store above has read the staging buffer and the column sums are done
        const bool last_pass = pass == BN / 128 - 1;
        const bool more = !last_pass || tile + tile_step < tile_end;
        if (p.stats) asm volatile("bar.sync 1, 256;" ::: "memory");
        if (threadIdx.x == 64 && more) {
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          if (last_pass) res_fetch(tile + tile_step, 0, 0); else res_fetch(tile, pass + 1, 0);
        }
      }
      ++pcount;
      }
      }
      if constexpr (EPI == 1) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(acc));
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (GNF && warp >= 11) {
    // ================= GroupNorm + SiLU transform warps (11..14) =================
    if constexpr (GNF) {
      const int tt = (int)threadIdx.x - 11 * 32;                   // 0 .. 127
      const int u = tt & 7;                                        // 16-byte unit of a pixel row = 8 channels
      uint8_t* const sgen = smem_raw + (base - smem_u32(smem_raw));
      int stage = 0; uint32_t phase = 0;
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        int mt, nt, b0, h0, w0;
        tile_coords(tile, mt, nt);
        m_origin(mt, b0, h0, w0);
        const bool live = b0 < p.B;                                // (odd tail tile of a pair: nothing but zero fill)
        for (int kb = 0; kb < kb_per_tap; ++kb) {
          float4 cf[4];                                            // (scale/2, shift/2) of this thread's 8 channels
          if (live) {
            const float4* cp = reinterpret_cast<const float4*>(p.gn_coef + ((size_t)b0 * Ctot + kb * TC_BK + u * 8) * 2);
#pragma unroll
            for (int i = 0; i < 4; ++i) cf[i] = __ldg(cp + i);
          }
          mbar_wait(full_bar(stage), phase);                       // this CTA's three boxes have landed
          if (live) {
            uint8_t* const sa = sgen + stage * S::A_BYTES;
#pragma unroll 4
            for (int q = tt >> 3; q < 10 * (HALO_TH + 2); q += 4 * TC_XF_WARPS) {
              const int r = q / 10, j = q - r * 10;                // unique pixel: row r of the boxes, column w0 - 1 + j
              if ((unsigned)(h0 - 1 + r) < (unsigned)p.H && (unsigned)(w0 - 1 + j) < (unsigned)p.W) {
                const int sbox = j == 0 ? 0 : (j == 9 ? 2 : 1), scol = j - sbox;
                uint8_t* const row = sa + r * (HALO_TW * 128);
                const uint4 v = *reinterpret_cast<const uint4*>(row + sbox * S::BOX_BYTES + scol * 128 + ((u ^ scol) << 4));
                const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
                uint32_t o4[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float h0f = fmaf(__uint_as_float(w4[i] << 16), cf[i].x, cf[i].y);
                  const float h1f = fmaf(__uint_as_float(w4[i] & 0xffff0000u), cf[i].z, cf[i].w);
                  float t0, t1;
                  asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(h0f));
                  asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(h1f));
                  const __nv_bfloat162 pk = __floats2bfloat162_rn(fmaf(h0f, t0, h0f), fmaf(h1f, t1, h1f));
                  o4[i] = *reinterpret_cast<const uint32_t*>(&pk);
                }
                const uint4 o = make_uint4(o4[0], o4[1], o4[2], o4[3]);
                if (j <= 7) *reinterpret_cast<uint4*>(row + j * 128 + ((u ^ j) << 4)) = o;
                if (j >= 1 && j <= 8) *reinterpret_cast<uint4*>(row + S::BOX_BYTES + (j - 1) * 128 + ((u ^ (j - 1)) << 4)) = o;
                if (j >= 2) *reinterpret_cast<uint4*>(row + 2 * S::BOX_BYTES + (j - 2) * 128 + ((u ^ (j - 2)) << 4)) = o;
              }
            }
          }
          fence_proxy_async_smem();                                // generic-proxy writes -> visible to the MMA's reads
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_cta(xf_bar(stage), 0));
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
        for (int kx = 0; kx < kb_extra; ++kx) {          // shortcut K blocks carry the RAW block input: pass them on
          mbar_wait(full_bar(stage), phase);
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_cta(xf_bar(stage), 0));
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  }
  if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if constexpr (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(S::TMEM_COLS));
    else
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(S::TMEM_COLS));
  }
}

}  // namespace b200

// ---- host side ---------------------------------------------------------------------------------------------------
using namespace b200;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct MapKey {
  uint64_t v[14];
  bool operator==(const MapKey& o) const {
    for (int i = 0; i < 14; ++i) if (v[i] != o.v[i]) return false;
    return true;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < 14; ++i) { h ^= k.v[i]; h *= 1099511628211ull; }
    return (size_t)h;
  }
};
std::mutex g_map_mu;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;

// 4-D map over a channels-last activation [B, H, W, C] (bf16); box {64, TW, TH, TB}.
int act_map(CUtensorMap* out, const void* ptr, int B, int H, int W, int C, int TW, int TH, int TB, int estride = 1) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  // with an element stride s the box is given in traversed elements: TW*s columns load TW pixels
  const uint32_t box[4] = {(uint32_t)TC_BK, (uint32_t)(TW * estride), (uint32_t)(TH * estride), (uint32_t)TB};
  const uint32_t es[4] = {1, (uint32_t)estride, (uint32_t)estride, 1};
  return tc_tensor_map(out, ptr, 4, dims, strides, box, 0, es);
}

// 2-D map over packed weights [Cout][Ktot] (bf16, K contiguous); box {64, BN}.
int weight_map(CUtensorMap* out, const void* ptr, int Cout, int Ktot, int BN) {
  const uint64_t dims[2] = {(uint64_t)Ktot, (uint64_t)Cout};
  const uint64_t strides[1] = {(uint64_t)Ktot * 2};
  const uint32_t box[2] = {(uint32_t)TC_BK, (uint32_t)BN};
  return tc_tensor_map(out, ptr, 2, dims, strides, box, 1);
}

bool tile_box(int B, int H, int W, int& TW, int& TH, int& TB, int& tiles_w, int& tiles_h) {
  if (W >= TC_BM) {
    if (W % TC_BM) return false;
    TW = TC_BM; TH = 1; TB = 1; tiles_w = W / TC_BM; tiles_h = H;
    return true;
  }
  if (W <= 0 || TC_BM % W) return false;
  const int rows = TC_BM / W;
  TW = W; tiles_w = 1;
  if (rows <= H) {
    if (H % rows) return false;
    TH = rows; TB = 1; tiles_h = H / rows;
    return true;
  }
  if (rows % H) return false;
  TH = H; TB = rows / H; tiles_h = 1;
  return TB <= 256;
}

struct XMaps { CUtensorMap x1, x2; };     // shortcut sources (unused copies of the A map when p.C3 + p.C4 == 0)

template <int BN, int EPI, bool HALO = false, bool DS = false>
int launch(const CUtensorMap& a1, const CUtensorMap& a2, const CUtensorMap& w, const CUtensorMap& om,
           const ConvTcParams& p, cudaStream_t st, const XMaps* xm = nullptr) {
  // residual tiles are fetched with the geometry of the output boxes (same tensor shape, other base pointer)
  CUtensorMap rm = om;
  if (EPI == 0 && p.res) {
    const int rc = act_map(&rm, p.res, p.B, p.H, p.W, p.Cout, p.TW, p.TH, p.TB);
    if (rc) return rc;
  }
  static int attr_state[B200_MAX_DEVICES];
  if (int rc = b200_set_smem_once(attr_state, conv_tc_kernel<BN, EPI, HALO, DS>, TcSmem<BN, HALO, DS>::TOTAL)) return rc;
  const int tiles = p.num_m * p.num_n;
  const int grid = tiles < B200_NUM_SMS ? tiles : B200_NUM_SMS;
  cudaError_t le = b200_launch_pdl(conv_tc_kernel<BN, EPI, HALO, DS>, dim3(grid), dim3(TC_THREADS + (HALO ? 32 : 0)),
                                   (size_t)TcSmem<BN, HALO, DS>::TOTAL, st, a1, a2, w, om, rm, xm ? xm->x1 : a1,
                                   xm ? xm->x2 : a1, p);
  if (le != cudaSuccess) return -(1000 + (int)le);
  B200_RETURN_LAST();
}

template <int BN, bool HALO, bool DS = false, bool GNF = false>
int launch_pair(const CUtensorMap& a1, const CUtensorMap& a2, const CUtensorMap& w, const CUtensorMap& om,
                const ConvTcParams& p, cudaStream_t st, const XMaps* xm = nullptr) {
  using SP = typename TcSmemSel<BN, HALO, DS, true>::type;
  static int attr_state[B200_MAX_DEVICES];
  auto kern = conv_tc_kernel<BN, 0, HALO, DS, true, GNF>;
  if (int rc = b200_set_smem_once(attr_state, kern, SP::TOTAL)) return rc;
  CUtensorMap rm = om;
  if (p.res) {
    const int rc = act_map(&rm, p.res, p.B, p.H, p.W, p.Cout, p.TW, p.TH, p.TB);
    if (rc) return rc;
  }
  const int pairs = ((p.num_m + 1) / 2) * p.num_n;
  const int clusters = pairs < B200_NUM_SMS / 2 ? pairs : B200_NUM_SMS / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(TC_THREADS + (HALO ? 32 : 0) + (GNF ? 32 * TC_XF_WARPS : 0));
  cfg.dynamicSmemBytes = SP::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;       // see b200_launch_pdl (common.cuh)
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = b200_pdl_enabled() ? 2 : 1;
  cudaError_t le = cudaLaunchKernelEx(&cfg, kern, a1, a2, w, om, rm, xm ? xm->x1 : a1, xm ? xm->x2 : a1, p);
  if (le != cudaSuccess) return -(1000 + (int)le);
  B200_RETURN_LAST();
}

// CTA-pair variant for 3x3 convolutions with 256-wide tiles (B200_CONV_PAIR=0 disables it: A/B measurements)
bool pair_enabled() {
  static const bool on = [] { const char* e = getenv("B200_CONV_PAIR"); return !(e && e[0] == '0'); }();
  return on;
}

bool halo_pair_enabled() {
  static const bool on = [] { const char* e = getenv("B200_CONV_HALO_PAIR"); return !(e && e[0] == '0'); }();
  return on;
}

// halo variant on / off (B200_CONV_HALO=0 disables it: A/B measurements)
bool halo_enabled() {
  static const bool on = [] { const char* e = getenv("B200_CONV_HALO"); return !(e && e[0] == '0'); }();
  return on;
}

// double-buffered output staging for 1x1 convolutions on / off (B200_CONV_DS=0 disables it: A/B measurements)
bool ds_enabled() {
  static const bool on = [] { const char* e = getenv("B200_CONV_DS"); return !(e && e[0] == '0'); }();
  return on;
}

}  // namespace

namespace b200 {
int tc_tensor_map(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides,
                  const uint32_t* box, int l2_256, const uint32_t* estr, int f32) {
  MapKey key;
  for (int i = 0; i < 14; ++i) key.v[i] = 0;
  key.v[0] = (uint64_t)ptr; key.v[1] = (uint64_t)rank | ((uint64_t)l2_256 << 8) | ((uint64_t)(f32 ? 1 : 0) << 12);
  if (estr) for (int i = 0; i < rank; ++i) key.v[1] |= (uint64_t)estr[i] << (16 + 8 * i);
  for (int i = 0; i < rank; ++i) { key.v[2 + i] = dims[i]; key.v[6 + i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) key.v[10 + i] = strides[i];
  {
    std::lock_guard<std::mutex> g(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) { *out = it->second; return B200_OK; }
  }
  EncodeTiledFn enc = get_encode();
  if (!enc) return B200_ERR_DRIVER;
  cuuint64_t d[4], st[3];
  cuuint32_t bx[4], es[4] = {1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; if (estr) es[i] = estr[i]; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides[i];
  CUresult r = enc(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   l2_256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return B200_ERR_DRIVER;
  std::lock_guard<std::mutex> g(g_map_mu);
  g_maps[key] = *out;
  return B200_OK;
}
}  // namespace b200

// slots per image of the `stats` by-product for an output of extent H x W (0: shape not supported)
extern "C" int b200_conv2d_tc_stat_slots(int B, int H, int W) {
  int TW, TH, TB, tw, th;
  if (B <= 0 || !tile_box(B, H, W, TW, TH, TB, tw, th)) return 0;
  return TB > 1 ? 1 : 2 * tw * th;
}

extern "C" int b200_conv2d_tc_supported(int C1, int C2, int B, int H, int W, int Cout, int ksize, int stride,
                                        int out_mode) {
  int TW, TH, TB, tw, th;
  if (stride != 1 || (ksize != 1 && ksize != 3)) return 0;
  if (C1 <= 0 || C1 % TC_BK || C2 < 0 || C2 % TC_BK || Cout <= 0) return 0;
  if (out_mode == 0 ? (Cout % 128 != 0) : out_mode == 1 ? (Cout > 16) : (out_mode != 2 || Cout != 8)) return 0;
  if (B <= 0 || !tile_box(B, H, W, TW, TH, TB, tw, th)) return 0;
  return 1;
}

// Which tile variant a stride-1 convolution of this shape runs as (shared by the launcher and by
// b200_conv2d_tc_gn_supported): BN, and whether it is one of the halo + pair variants.
static void conv_tc_variant(int B, int H, int W, int Cout, int ksize, int out_mode, int& BN, bool& halo_pair) {
  int TW, TH, TB, tw, th;
  tile_box(B, H, W, TW, TH, TB, tw, th);
  const int num_m = TB > 1 ? (B + TB - 1) / TB : B * tw * th;
  BN = out_mode != 0 ? 16 : ((Cout % 256 == 0) ? 256 : 128);
  static const int pair_min = [] { const char* e = getenv("B200_CONV_PAIR_MIN"); return e ? atoi(e) : 48; }();
  const bool pair_small = out_mode == 0 && ksize == 3 && pair_enabled() && ((num_m + 1) / 2) * (Cout / 256) >= pair_min;
  if (BN == 256 && num_m * (Cout / 256) < B200_NUM_SMS && !pair_small) BN = 128;
  static const bool hp256 = [] { const char* e = getenv("B200_CONV_HALO_PAIR256"); return !(e && e[0] == '0'); }();
  const bool halo256 = BN == 256 && hp256 && halo_pair_enabled() && pair_enabled();
  halo_pair = out_mode == 0 && (BN == 128 || halo256) && ksize == 3 && TB == 1 && H % HALO_TH == 0 && W % HALO_TW == 0 &&
              halo_enabled() && halo_pair_enabled();
}

// 1: a 3x3 convolution of this shape can take its input through the GroupNorm-fused operand pipeline (b200_conv2d_tc_gn)
extern "C" int b200_conv2d_tc_gn_supported(int C1, int C2, int B, int H, int W, int Cout, int ksize) {
  if (!b200_conv2d_tc_supported(C1, C2, B, H, W, Cout, ksize, 1, 0)) return 0;
  static const bool on = [] { const char* e = getenv("B200_CONV_GN_FUSE"); return !(e && e[0] == '0'); }();
  int BN; bool hp;
  conv_tc_variant(B, H, W, Cout, ksize, 0, BN, hp);
  return (on && hp) ? 1 : 0;
}

// 1: ... and the fused form is the faster one for this shape (what the engine's planner asks).  Measured at B = 1024 and
// B = 128 (profiles/r02_gn_fusion_shapes.txt): 256-wide pair tiles gain 4-9 % over conv + standalone GroupNorm; 128-wide
// pair tiles LOSE 3-16 %: an N = 128 pair MMA already reads 96 of the 128 B/clk the SM's shared memory delivers (4 KB of
// A + 2 KB of B per 64 clk) and the in-place transform adds ~40 B/clk (one 16-byte load and up to three stores per 8
// channels of each of the 180 unique pixels of a stage) -- those convolutions become shared-memory-bandwidth bound.
extern "C" int b200_conv2d_tc_gn_preferred(int C1, int C2, int B, int H, int W, int Cout, int ksize) {
  if (!b200_conv2d_tc_gn_supported(C1, C2, B, H, W, Cout, ksize)) return 0;
  int BN; bool hp;
  conv_tc_variant(B, H, W, Cout, ksize, 0, BN, hp);
  return BN == 256 ? 1 : 0;
}

static int conv2d_tc_impl(const void* x1, int C1, const void* x2, int C2, const void* w, const float* bias,
                          const float* temb, int temb_stride, const void* res, float out_scale, void* out,
                          int out_mode, const float* sample_mul, float* stats, const float* gn_coef, int B, int H, int W,
                          int Cout, int ksize, cudaStream_t st, const void* xs1 = nullptr, int Cs1 = 0,
                          const void* xs2 = nullptr, int Cs2 = 0);

extern "C" int b200_conv2d_tc(const void* x1, int C1, const void* x2, int C2, const void* w, const float* bias,
                              const float* temb, int temb_stride, const void* res, float out_scale, void* out,
                              int out_mode, const float* sample_mul, float* stats, int B, int H, int W, int Cout,
                              int ksize, cudaStream_t st) {
  return conv2d_tc_impl(x1, C1, x2, C2, w, bias, temb, temb_stride, res, out_scale, out, out_mode, sample_mul, stats,
                        nullptr, B, H, W, Cout, ksize, st);
}

// conv(silu(groupnorm(cat(x1, x2)))) with the normalisation applied inside the operand pipeline: x1 / x2 are the RAW
// producer outputs and gn_coef = [B][C1 + C2][2] (scale / 2, shift / 2) from b200_gn_coef.  Replaces
// `h = act(GroupNorm(x)); h = Conv(h)` of the residual blocks (models/layerspp.py:242-246, 264-266) without writing h.
extern "C" int b200_conv2d_tc_gn(const void* x1, int C1, const void* x2, int C2, const float* gn_coef, const void* w,
                                 const float* bias, const float* temb, int temb_stride, const void* res, float out_scale,
                                 void* out, float* stats, int B, int H, int W, int Cout, int ksize, cudaStream_t st) {
  if (!gn_coef) return B200_ERR_ARG;
  if (!b200_conv2d_tc_gn_supported(C1, C2, B, H, W, Cout, ksize)) return B200_ERR_UNSUPPORTED;
  return conv2d_tc_impl(x1, C1, x2, C2, w, bias, temb, temb_stride, res, out_scale, out, 0, nullptr, stats, gn_coef, B, H,
                        W, Cout, ksize, st);
}

// The second convolution of a residual block with its 1x1 shortcut folded in (models/layerspp.py:266-274):
//     out = (conv3x3(h; W1) + conv1x1(cat(xs1, xs2); W2) + bias) * out_scale,      bias = b1 + b2,
// one GEMM over K = 9 (C1 + C2) + Cs1 + Cs2: the shortcut's K blocks are loaded from the block input at the centre tap
// and meet the stacked weight [W1 | W2] ([Cout][9 (C1 + C2) + Cs1 + Cs2], K contiguous).  `Conv_2`'s own launch, its
// output and the residual read of `Conv_1` disappear.  gn_coef != NULL: h = silu(groupnorm(raw x1 / x2)) as in
// b200_conv2d_tc_gn (the shortcut sources are never normalised).  3x3, stride 1, Cs1 % 64 == Cs2 % 64 == 0.
extern "C" int b200_conv2d_tc_sc(const void* x1, int C1, const void* x2, int C2, const float* gn_coef, const void* xs1,
                                 int Cs1, const void* xs2, int Cs2, const void* w, const float* bias, const float* temb,
                                 int temb_stride, float out_scale, void* out, float* stats, int B, int H, int W, int Cout,
                                 cudaStream_t st) {
  if (!xs1 || Cs1 <= 0) return B200_ERR_ARG;
  if (gn_coef && !b200_conv2d_tc_gn_supported(C1, C2, B, H, W, Cout, 3)) return B200_ERR_UNSUPPORTED;
  return conv2d_tc_impl(x1, C1, x2, C2, w, bias, temb, temb_stride, nullptr, out_scale, out, 0, nullptr, stats, gn_coef,
                        B, H, W, Cout, 3, st, xs1, Cs1, xs2, Cs2);
}

static int conv2d_tc_impl(const void* x1, int C1, const void* x2, int C2, const void* w, const float* bias,
                          const float* temb, int temb_stride, const void* res, float out_scale, void* out,
                          int out_mode, const float* sample_mul, float* stats, const float* gn_coef, int B, int H, int W,
                          int Cout, int ksize, cudaStream_t st, const void* xs1, int Cs1, const void* xs2, int Cs2) {
  if (!x1 || !w || !out || (C2 > 0 && !x2)) return B200_ERR_ARG;
  if (!b200_conv2d_tc_supported(C1, C2, B, H, W, Cout, ksize, 1, out_mode)) return B200_ERR_UNSUPPORTED;
  if ((out_mode == 1 && (temb || res)) || (out_mode == 2 && temb)) return B200_ERR_UNSUPPORTED;
  ConvTcParams p;
  p.B = B; p.H = H; p.W = W; p.C1 = C1; p.C2 = C2; p.Cout = Cout; p.taps_w = ksize; p.pad = ksize / 2; p.stride = 1;
  tile_box(B, H, W, p.TW, p.TH, p.TB, p.tiles_w, p.tiles_h);
  p.num_m = p.TB > 1 ? (B + p.TB - 1) / p.TB : B * p.tiles_w * p.tiles_h;
  int BN = out_mode != 0 ? 16 : ((Cout % 256 == 0) ? 256 : 128);
  // small levels: more, narrower tiles -- unless the CTA-pair variant still fills most of the 74 SM pairs in ONE wave of
  // 256 x 256 tiles (4x4 maps at B = 1024: 64 pairs; two waves of 128-wide tiles took twice as long)
  static const int pair_min = [] { const char* e = getenv("B200_CONV_PAIR_MIN"); return e ? atoi(e) : 48; }();
  const bool pair_small = out_mode == 0 && ksize == 3 && pair_enabled() &&
                          ((p.num_m + 1) / 2) * (Cout / 256) >= pair_min;
  if (BN == 256 && p.num_m * (Cout / 256) < B200_NUM_SMS && !pair_small) BN = 128;
  p.num_n = out_mode != 0 ? 1 : Cout / BN;
  p.narrow = out_mode == 2 ? 8 : 0;
  p.bias = bias; p.temb = temb; p.temb_stride = temb_stride; p.res = (const bf16*)res; p.out_scale = out_scale;
  p.out = (bf16*)out; p.out_nchw = (float*)out; p.sample_mul = sample_mul; p.cout_true = Cout;
  p.stats = out_mode == 0 ? stats : nullptr; p.stat_slots = p.TB > 1 ? 1 : 2 * p.tiles_w * p.tiles_h;
  p.gn_coef = gn_coef; p.w_batched = 0;
  p.C3 = Cs1; p.C4 = Cs2;
  if (Cs1 < 0 || Cs2 < 0 || Cs1 % TC_BK || Cs2 % TC_BK || (Cs1 > 0 && !xs1) || (Cs2 > 0 && !xs2) || (Cs2 > 0 && Cs1 == 0))
    return B200_ERR_ARG;
  if ((Cs1 + Cs2) > 0 && (out_mode != 0 || ksize != 3)) return B200_ERR_UNSUPPORTED;
  const int Ktot = ksize * ksize * (C1 + C2) + Cs1 + Cs2;
  XMaps xm;
  const XMaps* xmp = (Cs1 + Cs2) > 0 ? &xm : nullptr;
  CUtensorMap a1, a2, wm;
  int rc = act_map(&a1, x1, B, H, W, C1, p.TW, p.TH, p.TB);
  if (rc) return rc;
  if (C2 > 0) { rc = act_map(&a2, x2, B, H, W, C2, p.TW, p.TH, p.TB); if (rc) return rc; }
  else a2 = a1;
  rc = weight_map(&wm, w, Cout, Ktot, BN);          // rows beyond Cout are zero-filled by TMA (head: 3 of 16)
  if (rc) return rc;
  if (out_mode != 0) {
    // network head / image-pyramid outputs (16-wide tiles): purely a data-movement launch, and with one box per tap every
    // activation tile lands nine times in shared memory (250 us at B = 1024 for 268 MB of input).  Same halo loader as the
    // wide variants when the geometry allows: three column-shifted boxes per K block serve all nine taps.
    static const bool head_halo = [] { const char* e = getenv("B200_CONV_HEAD_HALO"); return !(e && e[0] == '0'); }();
    if (head_halo && ksize == 3 && p.TB == 1 && H % HALO_TH == 0 && W % HALO_TW == 0 && halo_enabled() && !xmp) {
      p.TW = HALO_TW; p.TH = HALO_TH; p.tiles_w = W / HALO_TW; p.tiles_h = H / HALO_TH;
      p.num_m = B * p.tiles_w * p.tiles_h;
      rc = act_map(&a1, x1, B, H, W, C1, HALO_TW, HALO_TH + 2, 1);
      if (rc) return rc;
      if (C2 > 0) { rc = act_map(&a2, x2, B, H, W, C2, HALO_TW, HALO_TH + 2, 1); if (rc) return rc; }
      else a2 = a1;
      return launch<16, 1, true>(a1, a2, wm, a1, p, st);
    }
    return launch<16, 1>(a1, a2, wm, a1, p, st);
  }
  CUtensorMap om;
  static const bool hp256 = [] { const char* e = getenv("B200_CONV_HALO_PAIR256"); return !(e && e[0] == '0'); }();
  const bool halo256 = BN == 256 && hp256 && halo_pair_enabled() && pair_enabled();
  if ((BN == 128 || halo256) && ksize == 3 && p.TB == 1 && H % HALO_TH == 0 && W % HALO_TW == 0 && halo_enabled()) {
    // N = 128 tiles are bound by activation ingress, not by the tensor pipe: 8 x 16 pixel tiles with halo reuse
    p.TW = HALO_TW; p.TH = HALO_TH; p.tiles_w = W / HALO_TW; p.tiles_h = H / HALO_TH;
    p.num_m = B * p.tiles_w * p.tiles_h;
    p.stat_slots = 2 * p.tiles_w * p.tiles_h;
    rc = act_map(&a1, x1, B, H, W, C1, HALO_TW, HALO_TH + 2, 1);
    if (rc) return rc;
    if (C2 > 0) { rc = act_map(&a2, x2, B, H, W, C2, HALO_TW, HALO_TH + 2, 1); if (rc) return rc; }
    else a2 = a1;
    rc = act_map(&om, out, B, H, W, Cout, HALO_TW, HALO_TH, 1);
    if (rc) return rc;
    if (xmp) {
      rc = act_map(&xm.x1, xs1, B, H, W, Cs1, HALO_TW, HALO_TH + 2, 1);
      if (rc) return rc;
      if (Cs2 > 0) { rc = act_map(&xm.x2, xs2, B, H, W, Cs2, HALO_TW, HALO_TH + 2, 1); if (rc) return rc; }
      else xm.x2 = xm.x1;
    }
    if (BN == 256) {
      CUtensorMap wh;                                     // each CTA of the pair loads 128 of the 256 weight rows
      rc = weight_map(&wh, w, Cout, Ktot, 128);
      if (rc) return rc;
      if (gn_coef) return launch_pair<256, true, false, true>(a1, a2, wh, om, p, st, xmp);
      return launch_pair<256, true>(a1, a2, wh, om, p, st, xmp);
    }
    if (halo_pair_enabled()) {
      CUtensorMap wh;                                     // each CTA of the pair loads 64 of the 128 weight rows
      rc = weight_map(&wh, w, Cout, Ktot, 64);
      if (rc) return rc;
      // B200_CONV_HALO_PAIR_DS=1: second staging buffer instead of the third weight-ring stage for convs with a
      // residual (measured: 128->128 0.304 -> 0.299 ms, 256->128 0.505 -> 0.529 ms; off)
      static const bool hp_ds = [] { const char* e = getenv("B200_CONV_HALO_PAIR_DS"); return e && e[0] == '1'; }();
      if (gn_coef) return launch_pair<128, true, false, true>(a1, a2, wh, om, p, st, xmp);
      if (res && hp_ds) return launch_pair<128, true, true>(a1, a2, wh, om, p, st, xmp);
      return launch_pair<128, true>(a1, a2, wh, om, p, st, xmp);
    }
    if (gn_coef) return B200_ERR_UNSUPPORTED;
    return launch<128, 0, true>(a1, a2, wm, om, p, st, xmp);
  }
  if (gn_coef) return B200_ERR_UNSUPPORTED;
  rc = act_map(&om, out, B, H, W, Cout, p.TW, p.TH, p.TB);
  if (rc) return rc;
  if (xmp) {
    rc = act_map(&xm.x1, xs1, B, H, W, Cs1, p.TW, p.TH, p.TB);
    if (rc) return rc;
    if (Cs2 > 0) { rc = act_map(&xm.x2, xs2, B, H, W, Cs2, p.TW, p.TH, p.TB); if (rc) return rc; }
    else xm.x2 = xm.x1;
  }
  if (ksize == 3 && BN == 256 && pair_enabled()) {
    CUtensorMap wh;                                       // each CTA of the pair loads 128 of the 256 weight rows
    rc = weight_map(&wh, w, Cout, Ktot, 128);
    if (rc) return rc;
    return launch_pair<256, false>(a1, a2, wh, om, p, st, xmp);
  }
  if (ksize == 1 && ds_enabled())     // epilogue-bound: double-buffered staging
    return BN == 256 ? launch<256, 0, false, true>(a1, a2, wm, om, p, st) : launch<128, 0, false, true>(a1, a2, wm, om, p, st);
  return BN == 256 ? launch<256, 0>(a1, a2, wm, om, p, st, xmp) : launch<128, 0>(a1, a2, wm, om, p, st, xmp);
}

// 3x3 stride-2 VALID convolution (zero fill beyond the input extent) on the same kernel: the TMA box walks the input
// with element strides {1, 2, 2, 1}.  Covers conv_downsample_2d's F.conv2d(stride=2) on the FIR-padded (H+1)x(W+1)
// tensor (models/up_or_down_sampling.py:174-178) and DDPM's pad(0,1,0,1) + conv3x3 stride 2 (models/layers.py:596-608).
extern "C" int b200_conv2d_tc_s2_supported(int C, int B, int Ho, int Wo, int Cout) {
  int TW, TH, TB, tw, th;
  if (C <= 0 || C % TC_BK || Cout <= 0 || Cout % 128 || B <= 0) return 0;
  if (!tile_box(B, Ho, Wo, TW, TH, TB, tw, th)) return 0;
  return (TW * 2 <= 256 && TH * 2 <= 256) ? 1 : 0;
}

extern "C" int b200_conv2d_tc_s2(const void* x, int C, const void* w, const float* bias, const void* res,
                                 float out_scale, void* out, float* stats, int B, int Hin, int Win, int Ho, int Wo,
                                 int Cout, cudaStream_t st) {
  if (!x || !w || !out) return B200_ERR_ARG;
  if (!b200_conv2d_tc_s2_supported(C, B, Ho, Wo, Cout)) return B200_ERR_UNSUPPORTED;
  ConvTcParams p;
  p.B = B; p.H = Ho; p.W = Wo; p.C1 = C; p.C2 = 0; p.Cout = Cout; p.taps_w = 3; p.pad = 0; p.stride = 2;
  tile_box(B, Ho, Wo, p.TW, p.TH, p.TB, p.tiles_w, p.tiles_h);
  p.num_m = p.TB > 1 ? (B + p.TB - 1) / p.TB : B * p.tiles_w * p.tiles_h;
  int BN = (Cout % 256 == 0) ? 256 : 128;
  if (BN == 256 && p.num_m * (Cout / 256) < B200_NUM_SMS) BN = 128;
  p.num_n = Cout / BN;
  p.bias = bias; p.temb = nullptr; p.temb_stride = 0; p.res = (const bf16*)res; p.out_scale = out_scale;
  p.out = (bf16*)out; p.out_nchw = nullptr; p.sample_mul = nullptr; p.cout_true = Cout; p.narrow = 0;
  p.stats = stats; p.stat_slots = p.TB > 1 ? 1 : 2 * p.tiles_w * p.tiles_h; p.gn_coef = nullptr; p.C3 = 0; p.C4 = 0; p.w_batched = 0;
  CUtensorMap a1, wm, om;
  int rc = act_map(&a1, x, B, Hin, Win, C, p.TW, p.TH, p.TB, 2);
  if (rc) return rc;
  rc = weight_map(&wm, w, Cout, 9 * C, BN);
  if (rc) return rc;
  rc = act_map(&om, out, B, Ho, Wo, Cout, p.TW, p.TH, p.TB);
  if (rc) return rc;
  return BN == 256 ? launch<256, 0>(a1, a1, wm, om, p, st) : launch<128, 0>(a1, a1, wm, om, p, st);
}

// Split-K batched GEMM on the 1x1 path of the same kernel: out[s][m][n] = sum over the s-th K chunk of a[m][k] bm[n][k]
// (bf16 in / out, fp32 accumulate; a and bm row-major with row strides lda / ldb elements; K / S a multiple of 64).
// The weight gradient dW[co][(tap, ci)] = sum_p dY^T[co][p] XT[(tap, ci)][p] is a GEMM over K = B H W pixels with only
// (Cout / 128) x (9 Cin / BN) output tiles: run as ONE K loop per tile it kept 9-18 of the 148 SMs busy for 2048 K
// blocks (0.45 ms per layer whatever its size, 18.7 of the 59 ms of a C3 training step).  Here every K chunk is an
// "image" of the 1x1 convolution with its own B operand (3-D weight map), so tiles x S CTAs share the loop; the S
// partial products are summed by b200_sum_partials.
extern "C" int b200_gemm_tc_splitk(const void* a, long long lda, const void* bm, long long ldb, void* out, int M, int N,
                                   long long K, int S, cudaStream_t st) {
  if (!a || !bm || !out || S <= 0 || S > 65535 || lda < K || ldb < K) return B200_ERR_ARG;
  if (M <= 0 || M % TC_BM || N <= 0 || N % 128 || K <= 0 || K % S || (K / S) % TC_BK || lda % 8 || ldb % 8)
    return B200_ERR_UNSUPPORTED;
  const long long Kc = K / S;
  ConvTcParams p;
  p.B = S; p.H = 1; p.W = M; p.C1 = (int)Kc; p.C2 = 0; p.Cout = N; p.taps_w = 1; p.pad = 0; p.stride = 1;
  p.TW = TC_BM; p.TH = 1; p.TB = 1; p.tiles_w = M / TC_BM; p.tiles_h = 1;
  p.num_m = S * p.tiles_w;
  const int BN = (N % 256 == 0) ? 256 : 128;
  p.num_n = N / BN;
  p.narrow = 0;
  p.bias = nullptr; p.temb = nullptr; p.temb_stride = 0; p.res = nullptr; p.out_scale = 1.0f;
  p.out = (bf16*)out; p.out_nchw = nullptr; p.sample_mul = nullptr; p.cout_true = N;
  p.stats = nullptr; p.stat_slots = 2 * p.tiles_w; p.gn_coef = nullptr; p.C3 = 0; p.C4 = 0; p.w_batched = 1;
  CUtensorMap am, wm, om;
  {
    const uint64_t dims[4] = {(uint64_t)Kc, (uint64_t)M, 1, (uint64_t)S};
    const uint64_t strides[3] = {(uint64_t)lda * 2, (uint64_t)M * lda * 2, (uint64_t)Kc * 2};
    const uint32_t box[4] = {(uint32_t)TC_BK, (uint32_t)TC_BM, 1u, 1u};
    if (int rc = tc_tensor_map(&am, a, 4, dims, strides, box, 0)) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)Kc, (uint64_t)N, (uint64_t)S};
    const uint64_t strides[2] = {(uint64_t)ldb * 2, (uint64_t)Kc * 2};
    const uint32_t box[3] = {(uint32_t)TC_BK, (uint32_t)BN, 1u};
    if (int rc = tc_tensor_map(&wm, bm, 3, dims, strides, box, 1)) return rc;
  }
  if (int rc = act_map(&om, out, S, 1, M, N, p.TW, p.TH, p.TB)) return rc;
  return BN == 256 ? launch<256, 0, false, true>(am, am, wm, om, p, st) : launch<128, 0, false, true>(am, am, wm, om, p, st);
}

// Weight gradient of a stride-1 "same" convolution with no transposed copy of anything:
//   out[s][co][(tap, ci)] = sum over the s-th chunk of pixels p of gy[p][co] * x[p + tap][ci],
// gy = dY [P][M = Cout] and x1 / x2 = the conv's source(s) [P][C], all channels-last as the engine holds them (P = B H W).
// Both MMA operands are consumed MN-major: a K block is 64 pixels = 64 rows of 128 bytes (64 channels); for the B operand
// those rows are a 4-D TMA box {64 c, w, h, b} of x shifted by the tap with zero fill outside the image -- the forward
// conv's im2col trick with the roles of pixels and channels exchanged.  Replaces b200_transpose2d (dY^T) + b200_im2col_t
// (which wrote 9x the activation: 9.8 of the 47 ms of a C3 training step) + b200_gemm_tc_splitk.
static bool wgrad_geometry(int M, int C1, int C2, int B, int H, int W, int ksize, int S, int& BN, int& TWk, int& THk, int& TBk) {
  if (S <= 0 || S > 65535 || B <= 0 || H <= 0 || W <= 0 || C1 <= 0 || C2 < 0) return false;
  const long long P = (long long)B * H * W;
  const int N = ksize * ksize * (C1 + C2);
  if ((ksize != 1 && ksize != 3) || M <= 0 || M % TC_BM || P % S || (P / S) % TC_BK || P > 0x7fffffffLL) return false;
  if (C1 % 64 || C2 % 64) return false;
  BN = N % 256 == 0 ? 256 : 128;
  if (N % BN) return false;
  // a K block = 64 consecutive pixels = a {TWk, THk, TBk} box of whole rows / whole images
  TWk = W < TC_BK ? W : TC_BK;
  if (TC_BK % TWk || W % TWk) return false;
  const int rows = TC_BK / TWk;
  THk = rows < H ? rows : H;
  if (rows % THk || H % THk) return false;
  TBk = rows / THk;
  if (TWk * THk * TBk != TC_BK || (TBk > 1 && (THk != H || TWk != W)) || B % TBk) return false;
  return true;
}

extern "C" int b200_wgrad_tc_supported(int M, int C1, int C2, int B, int H, int W, int ksize, int S) {
  int BN, TWk, THk, TBk;
  return wgrad_geometry(M, C1, C2, B, H, W, ksize, S, BN, TWk, THk, TBk) ? 1 : 0;
}

extern "C" int b200_wgrad_tc(const void* gy, const void* x1, int C1, const void* x2, int C2, void* out, int M, int B,
                             int H, int W, int ksize, int S, cudaStream_t st) {
  if (!gy || !x1 || !out || (C2 > 0 && !x2)) return B200_ERR_ARG;
  int BN, TWk, THk, TBk;
  if (!wgrad_geometry(M, C1, C2, B, H, W, ksize, S, BN, TWk, THk, TBk)) return B200_ERR_UNSUPPORTED;
  const long long P = (long long)B * H * W;
  const int N = ksize * ksize * (C1 + C2);
  const long long Kc = P / S;
  ConvTcParams p;
  p.B = S; p.H = 1; p.W = M; p.C1 = (int)Kc; p.C2 = 0; p.Cout = N; p.taps_w = 1; p.pad = 0; p.stride = 1;
  p.TW = TC_BM; p.TH = 1; p.TB = 1; p.tiles_w = M / TC_BM; p.tiles_h = 1;
  p.num_m = S * p.tiles_w; p.num_n = N / BN; p.narrow = 0;
  p.bias = nullptr; p.temb = nullptr; p.temb_stride = 0; p.res = nullptr; p.out_scale = 1.0f;
  p.out = (bf16*)out; p.out_nchw = nullptr; p.sample_mul = nullptr; p.cout_true = N;
  p.stats = nullptr; p.stat_slots = 2 * p.tiles_w; p.gn_coef = nullptr; p.C3 = 0; p.C4 = 0;
  p.w_batched = 2; p.wg_H = H; p.wg_W = W; p.wg_ks = ksize; p.wg_C1 = C1; p.wg_C2 = C2; p.wg_Kc = Kc;
  CUtensorMap am, om;
  XMaps xm;
  {
    const uint64_t dims[2] = {(uint64_t)M, (uint64_t)P};               // dY [P][Cout]: boxes of 64 channels x 64 pixels
    const uint64_t strides[1] = {(uint64_t)M * 2};
    const uint32_t box[2] = {64u, (uint32_t)TC_BK};
    if (int rc = tc_tensor_map(&am, gy, 2, dims, strides, box, 0)) return rc;
  }
  auto xmap = [&](CUtensorMap* m, const void* ptr, int C) {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
    const uint32_t box[4] = {64u, (uint32_t)TWk, (uint32_t)THk, (uint32_t)TBk};
    return tc_tensor_map(m, ptr, 4, dims, strides, box, 0);
  };
  if (int rc = xmap(&xm.x1, x1, C1)) return rc;
  if (C2 > 0) { if (int rc = xmap(&xm.x2, x2, C2)) return rc; } else xm.x2 = xm.x1;
  if (int rc = act_map(&om, out, S, 1, M, N, p.TW, p.TH, p.TB)) return rc;
  return BN == 256 ? launch<256, 0, false, true>(am, am, xm.x1, om, p, st, &xm)
                   : launch<128, 0, false, true>(am, am, xm.x1, om, p, st, &xm);
}
