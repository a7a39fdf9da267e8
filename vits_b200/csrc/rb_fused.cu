// Fused ResBlock2 iteration (modules.py:250-260) as ONE persistent tcgen05 kernel:
//
//     y_in = leaky_relu(x)  (16-bit, time-major)          [single HBM read]
//       -> c1: dilated conv C->C (implicit GEMM, taps = row-shifted views of one smem slab)
//       -> + cond bias, tanh(a)*sigmoid(b)  -> G tile (C/2 ch) written to SHARED memory
//       -> c2: conv C/2->C over the smem G tile (taps = row-shifted views again)
//       -> + bias + x (recovered from the same smem slab) -> leaky_relu -> y_out   [single HBM write]
//
// The residual stream lives in HBM only as y = fp16(leaky_relu(x, 0.1)); leaky_relu is
// invertible (x = y >= 0 ? y : 10*y), so the one tensor serves both as the next conv's
// operand and as the residual.  Per iteration and element the kernel moves 2 B in + 2 B
// out instead of the 14 B of the layer-at-a-time plan.
//
// Output rows per tile: T_out = 128 - (k-1): c1 computes the 128 G rows
// [t0-p2, t0-p2+128) that c2's taps need for outputs [t0, t0+T_out), p2 = (k-1)/2.
//
// The kernel is specialised at compile time on (C, k) — 4 stages x 3 kernel sizes of the
// reference config — so that the two MMA-issuing warps run straight-line code: every UMMA
// descriptor is `base + compile-time offset`, no per-tap constant reloads or integer
// division (the first version spent ~500 cycles of uniform-datapath instructions per tap
// around 128 cycles of tensor work; profiles/r01_c_ncu_rb_fused_full_summary.txt).
//
// Roles (12 or 20 warps; the second epilogue group exists when there are >= 2 accumulator sets):
//   warp 0      activation-slab TMA producer
//   warp 1      c1 MMA issuer (and c2 in streamed-weight mode)
//   warp 2      c2 MMA issuer (resident-weight mode: small N needs two issuing threads)
//   warp 3      weight TMA producer
//   warps 4-7   epilogue 1: TMEM -> bias + cond + gate -> G tile in smem (swizzled like TMA)
//   warps 8-11  epilogue 2: TMEM -> bias + residual + leaky_relu -> staged rows -> TMA store
//   warps 12-15 / 16-19  second gate / store group: the two groups take alternate tiles (C <= 64) or split the columns (C >= 128)
// Every CTA (pair) works through one contiguous run of tiles, so the batch row - its conditioning vector, length and
// tile range, all cached in registers - changes a handful of times per launch.
// Weights are resident in smem when they fit (C <= 64) and streamed through one ring in
// the issuing warp's fixed order otherwise.  G tiles and both TMEM accumulators are
// NBUF-deep (4 for C <= 64, where a tile is ~500 tensor cycles and the mbarrier / commit
// round trips between the four stages dominate; 2 for C = 128; 1 for C = 256):
// c1(n+1) | gate(n) | c2(n) | store(n-1) overlap on consecutive tiles.
// For C >= 64 the store epilogue re-reads the residual rows from global memory (L2 hits:
// TMA fetched them microseconds earlier), so a slab is released as soon as c1's MMAs
// retire and 2-3 slabs are enough to keep loads ahead of the tensor pipe.
#include "common.h"
#include "kernels.h"
#include "ptx.h"
#include "tc_common.cuh"

#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace vitsb {

// Timing-experiment hooks (scripts/gpu_ablate.sh, VITSB_RB_TRACE): compiled in only with -DVITSB_RB_DEBUG,
// because even uniform run-time branches in the epilogue loops cost ~5 % of the kernel.
#ifdef VITSB_RB_DEBUG
#define RB_DBG(mask) ((rt.dbg & (mask)) != 0)
#define RB_TRACE_ON 1
#else
#define RB_DBG(mask) (false)
#define RB_TRACE_ON 0
#endif

// Biases / conditioning vectors: shared memory from C = 128 up, global / L1 below.  With contiguous tile runs per CTA the
// batch row - hence the L1 lines - stays the same for tens of tiles, and for C <= 64 every LDS competes with the N <= 64
// MMAs' operand reads for the shared-memory pipe (ncu, per launch: C = 64 k = 3 / 7 / 11 59.6 / 77.0 / 107.4 -> 57.1 / 70.2 /
// 100.9 us when its biases moved from shared memory to global; C = 128 loses 3-4 % there).
#ifndef VITSB_BIAS_SMEM_MINC
#define VITSB_BIAS_SMEM_MINC 128
#endif
constexpr int kBiasSmemMinC = VITSB_BIAS_SMEM_MINC;      // channel count from which the epilogues read biases from shared memory
#ifndef VITSB_RB_EARLY_MAXC
#define VITSB_RB_EARLY_MAXC 256
#endif
constexpr int kEarlyMaxC = VITSB_RB_EARLY_MAXC;           // epilogues hand an accumulator back as soon as its last chunk is in registers (0: at the end of the tile)
// Requesting a chunk's bias / conditioning values one chunk ahead (the first chunk's before the accumulator wait) was
// measured and is off: the values stay live across the wait and the kernels got 1-10 % slower (C = 32 / k = 3: 49.3 -> 55 us).
#ifdef VITSB_RB_BIAS_AHEAD
constexpr bool kBiasAhead = true;
#else
constexpr bool kBiasAhead = false;
#endif
constexpr int kRbMaxB = kMaxBatchRows;
constexpr uint32_t kRbSmemBudget = 220 * 1024;
constexpr int kRbMaxSlab = 16, kRbBars = 97;

template <int C, int KT>
struct RbT {
  static constexpr int C2 = C / 2;
  static constexpr int BK1 = C >= 64 ? 64 : C, nCB1 = C / BK1, rb1 = BK1 * 2, NK1 = BK1 / 16;
  static constexpr int BK2 = C2 >= 64 ? 64 : C2, nCB2 = C2 / BK2, rb2 = BK2 * 2, NK2 = BK2 / 16;
  static constexpr int P2 = (KT - 1) / 2, T_OUT = kTileM - (KT - 1);
  static constexpr int RG = (kTileM + (KT - 1) + 7) & ~7;
  static constexpr uint32_t G_BYTES = ((uint32_t)RG * rb2 + 1023u) & ~1023u;     // per channel block
  // CTA-pair mode (cluster of 2, tcgen05 cta_group::2) for the streamed-weight sizes: one UMMA covers the
  // 128-row tiles of BOTH CTAs (M = 256) and each CTA holds half of the weight rows, so the L2 -> smem
  // weight stream per SM halves (at M = 128 it needs ~64 B/clk/SM, above what L2 can feed 148 SMs)
  static constexpr bool PAIR = C >= 128;
  static constexpr int NCL = PAIR ? 2 : 1;
  static constexpr uint32_t W1_TX = (uint32_t)(C / NCL) * rb1, W2_TX = (uint32_t)(C / NCL) * rb2;  // bytes per (cb, tap) weight box, per CTA
  static constexpr uint32_t W_SLOT = W1_TX > W2_TX ? W1_TX : W2_TX;               // ring slot (streamed mode)
  static constexpr int NBOX1 = KT * nCB1, NBOX2 = KT * nCB2;
  static constexpr uint32_t W_ALL = NBOX1 * W1_TX + NBOX2 * W2_TX;
#ifndef VITSB_RB_NBUF32
#define VITSB_RB_NBUF32 8
#endif
  static constexpr int NBUF = C <= 32 ? VITSB_RB_NBUF32 : (C <= 64 ? 4 : (C <= 128 ? 2 : 1));   // G tiles / TMEM accumulator sets
  static constexpr uint32_t G_SET = nCB2 * G_BYTES;
  static constexpr bool STATIONARY = C <= 64;                                     // W_ALL <= 135 KB: resident weights, two issuers
#ifndef VITSB_RB_RES_GLOBAL_MINC
#define VITSB_RB_RES_GLOBAL_MINC 64
#endif
  static constexpr bool RES_GLOBAL = C >= VITSB_RB_RES_GLOBAL_MINC;                                     // residual rows re-read from global (L2)
  static constexpr uint32_t TMEM_COLS = 2 * NBUF * C <= 256 ? 256 : 512;
  // epilogue warp groups: with >= 2 accumulator sets, two groups of (4 gate + 4 store) warps take alternate
  // tiles - the epilogues are latency / issue bound (one warp per scheduler reaches IPC ~0.3), not MUFU bound
  // C >= 128 (COOP): the pipeline is bound by the LATENCY of one tile's epilogues (few accumulator sets), so both
  // groups work on EVERY tile and split its columns; C <= 64 (MMA-bound, 4 sets): groups take alternate tiles.
  static constexpr bool COOP = C >= 128;
#ifndef VITSB_RB_NGRP_SMALL
#define VITSB_RB_NGRP_SMALL 2
#endif
  // C <= 64: the epilogue groups take alternate tiles; C >= 128 splits columns between two groups.  A third group for
  // C <= 64 (-DVITSB_RB_NGRP_SMALL=3: 28 warps, 72 registers, no spills) was measured in round 2: s3 -2 %, s2 +-0 -
  // the narrow stages are not short of epilogue warps (shared-memory bandwidth of the N <= 64 MMAs)
  static constexpr int NGRP = C <= 64 ? VITSB_RB_NGRP_SMALL : 2;
  static constexpr int TSTEP = COOP ? 1 : NGRP;                                   // tile stride of one epilogue group
  static constexpr int NARR = 4 * NCL * (COOP ? NGRP : 1);                        // epilogue-warp arrivals per barrier phase
  static constexpr int THREADS = 128 + NGRP * 256;
  // mid-iteration output rows are staged in shared memory (one swizzled [128][CB_O] block per store group)
  // and written by TMA stores: a thread owns a ROW, so direct global stores touch 32 different lines per
  // instruction.  C = 64 / k = 11 has no smem left next to its 135 KB of resident weights and stores directly.
  static constexpr bool STAGE_OUT = !(C == 64 && KT == 11);
  static constexpr int CB_O = C >= 64 ? 64 : C, RB_O = CB_O * 2, NBLK_O = C / CB_O;
  static constexpr uint32_t STG_BYTES = (uint32_t)kTileM * RB_O;                   // per store group
  static constexpr uint32_t STG_ALL = STAGE_OUT ? NGRP * STG_BYTES : 0u;
  static_assert(2 * NBUF * C <= 512, "TMEM columns");
};

// run-time geometry (depends on the dilation and the smem left after the fixed parts)
struct RbRt {
  int dil, p1;               // c1 dilation, (k-1)*dil/2
  int R1;                    // slab rows (multiple of 8)
  uint32_t slab_bytes;       // per channel block (multiple of 1024)
  int nSlab;                 // activation slab buffers (1..8)
  int nW;                    // weight ring slots (streamed mode, 2..8)
  uint32_t idesc;
  int trace_t0;              // first tile recorded by the timeline trace
  int dbg;                   // VITSB_RB_DBG ablation mask (timing experiments only; results are wrong when set)
};

template <typename OpT, int C, int KT>
__global__ void __launch_bounds__((RbT<C, KT>::THREADS), 1)
rb_fused_kernel(const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmW1,
                const __grid_constant__ CUtensorMap tmW2, const RbArgs a, const RbRt rt) {
  using G = RbT<C, KT>;
  constexpr bool kEarly = C <= kEarlyMaxC;
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_tile_prefix[kRbMaxB + 1];
  __shared__ uint32_t s_tmem;
  // Bias vectors staged in shared memory: with ~220 KB of the SM's 256 KB carved out as smem the L1 is a few KB
  // and the residual stream flushes it, so per-chunk bias loads from global paid an L2 round trip each.
  constexpr int kSB = C >= kBiasSmemMinC ? C : 4;                   // small C reads its biases from global / L1 (measured faster there)
  __shared__ __align__(16) float s_bias2[kSB];            // conv2 bias
  __shared__ __align__(16) float s_bias1[G::NGRP][kSB];         // per epilogue group: conv1 bias + conditioning of the current batch row

  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t slab_set = G::nCB1 * rt.slab_bytes;
  const uint32_t sSlab = smem_base;                                            // [nSlab][nCB1][slab_bytes]
  const uint32_t sG = sSlab + rt.nSlab * slab_set;                             // [NBUF][nCB2][G_BYTES]
  const uint32_t sW = sG + G::NBUF * G::G_SET;                                 // all boxes (c1 then c2), or the ring
  const uint32_t sW2 = sW + G::NBOX1 * G::W1_TX;                               // resident mode: c2 boxes
  constexpr bool kStage = G::STAGE_OUT;                                        // TMA-stored output rows
  const uint32_t sStage = sW + (G::STATIONARY ? G::W_ALL : rt.nW * G::W_SLOT); // [NGRP][128][CB_O] output staging
  const uint32_t bars = sStage + (kStage ? G::STG_ALL : 0u);
  // mbarriers: up to kRbMaxSlab slabs, 8 accumulator / G-tile sets, 8 weight-ring slots (kRbBars in total)
  auto slab_full = [&](int i) { return bars + 8u * i; };
  auto slab_empty = [&](int i) { return bars + 8u * (16 + i); };
  auto acc1_full = [&](int i) { return bars + 8u * (32 + i); };
  auto acc1_empty = [&](int i) { return bars + 8u * (40 + i); };
  auto acc2_full = [&](int i) { return bars + 8u * (48 + i); };
  auto acc2_empty = [&](int i) { return bars + 8u * (56 + i); };
  auto g_full = [&](int i) { return bars + 8u * (64 + i); };
  auto g_empty = [&](int i) { return bars + 8u * (72 + i); };
  const uint32_t w_all = bars + 8u * 80;
  auto w_full = [&](int i) { return bars + 8u * (81 + i); };
  auto w_empty = [&](int i) { return bars + 8u * (89 + i); };

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform role id
  const int rank = G::PAIR ? (int)ptx::cluster_ctarank() : 0;
  const int unit_id = blockIdx.x / G::NCL, n_units = gridDim.x / G::NCL;     // a unit = one CTA, or one CTA pair

  // ---- tile table: tiles of batch row b cover output rows [0, min(L, valid_b + halo)) ----
  if (warp == 2) build_tile_prefix(s_tile_prefix, a.B, a.len, a.len_mul, a.L, G::T_OUT, lane);
  if (C >= kBiasSmemMinC)
    for (int i = threadIdx.x; i < C; i += blockDim.x) s_bias2[i] = a.bias2[i];
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmY); ptx::prefetch_tmap(&tmW1); ptx::prefetch_tmap(&tmW2);
    if (kStage) ptx::prefetch_tmap(&tmO);
    // slab release: c1's MMA commit, plus the 4 store-epilogue warps when they read the residual from the slab
    for (int i = 0; i < kRbMaxSlab; ++i) { ptx::mbar_init(slab_full(i), 1); ptx::mbar_init(slab_empty(i), G::RES_GLOBAL ? 1 : 5); }
    ptx::mbar_init(w_all, 1);
    for (int i = 0; i < 8; ++i) {
      // pair mode: the leader's copies collect the arrivals of both CTAs' epilogue warps
      ptx::mbar_init(acc1_full(i), 1);  ptx::mbar_init(acc1_empty(i), G::NARR);
      ptx::mbar_init(acc2_full(i), 1);  ptx::mbar_init(acc2_empty(i), G::NARR);
      ptx::mbar_init(g_full(i), G::NARR);        ptx::mbar_init(g_empty(i), 1);
    }
    for (int i = 0; i < 8; ++i) { ptx::mbar_init(w_full(i), 1); ptx::mbar_init(w_empty(i), 1); }
    ptx::mbar_fence_init();
  }
  if (warp == 1) { if (G::PAIR) ptx::tmem_alloc_2sm(ptx::smem_u32(&s_tmem), G::TMEM_COLS); else ptx::tmem_alloc(ptx::smem_u32(&s_tmem), G::TMEM_COLS); }
  ptx::tc_fence_before();
  if (G::PAIR) ptx::cluster_sync_all(); else __syncthreads();      // barrier inits visible to the peer before any remote signal
  ptx::tc_fence_after();
  ptx::pdl_launch_dependents();      // the next kernel may begin its own prologue on SMs this grid has left
  ptx::pdl_wait();                   // everything above overlapped the previous kernel's tail; its outputs are needed from here on
  const uint32_t tmem_base = s_tmem;
  // timeline trace of CTA 0 (VITSB_RB_TRACE): trace[(role * 16 + tile) * 2 + kind] = clock64()
  auto trace = [&](int role, int tile, int kind) {      // rt.trace_t0: first traced tile of the CTA's sequence (VITSB_RB_TRACE "C,k,count,first")
    tile -= rt.trace_t0;
    if (RB_TRACE_ON && a.trace && blockIdx.x == 0 && tile >= 0 && tile < 16 && lane == 0) a.trace[(role * 16 + tile) * 2 + kind] = clock64();
  };
  const int total_tiles = s_tile_prefix[a.B];
  const int total_units = (total_tiles + G::NCL - 1) / G::NCL;      // pair mode: tiles 2u and 2u+1 (the last may be missing)
  // Tile schedule: every CTA (or pair) takes a CONTIGUOUS run of units.  With a round-robin schedule consecutive tiles of a
  // CTA are n_units apart while a batch row has 64-640 tiles at the C2 sizes, so the batch row changed on every second
  // tile or so - and with it the conditioning vector of the gate epilogue (two group barriers around an exposed global
  // load: the 1.5-1.9 k cycle set-up gap per tile in the debug timeline, 36 % of the tiles) and the cached length.
#ifndef VITSB_RB_ROUND_ROBIN
  const int units_q = total_units / n_units, units_r = total_units - units_q * n_units;
  const int unit_base = unit_id * units_q + (unit_id < units_r ? unit_id : units_r), unit_step = 1;
  const int my_tiles = RB_DBG(256) ? 0 : units_q + (unit_id < units_r ? 1 : 0);
#else
  const int unit_base = unit_id, unit_step = n_units;
  const int my_tiles = RB_DBG(256) ? 0 : (total_units > unit_id ? (total_units - 1 - unit_id) / n_units + 1 : 0);
#endif
  constexpr uint32_t acc1_col = 0, acc2_col = G::NBUF * C;       // TMEM columns: [acc1 x NBUF][acc2 x NBUF]

  // static schedule; b only grows along a CTA's tile sequence
  // epilogue group of this warp (warps 4-11: group 0, warps 12-19: group 1) and its tile stride
  const int grp = warp >= 4 ? (warp - 4) >> 3 : 0;
  const bool is_epi = warp >= 4;
  const int tstep = is_epi ? G::TSTEP : 1;
  const int n_first = is_epi && !G::COOP ? grp : 0;          // first tile (of this CTA's sequence) the warp works on
  int cur_unit = unit_base + n_first * unit_step, cur_b = 0, cur_tile = 0;
  // Tile range [cur_lo, cur_hi) of batch row cur_b, kept in registers: under the epilogues' and the tensor core's
  // shared-memory traffic a dependent LDS takes ~500 cycles (debug timeline, phases 8-9), and the table walk plus the
  // t0 lookup were two of them in series at the top of every tile of every role.  Now the table is read only when a
  // CTA's tile sequence crosses into another batch row.
  int cur_lo = s_tile_prefix[0], cur_hi = s_tile_prefix[1];
  bool cur_dead = false;                          // pair mode, odd tile count: rank 1 of the last pair runs the protocol on zeros
  auto tile_seek = [&]() {
    const int t = G::PAIR ? cur_unit * 2 + rank : cur_unit;
    cur_dead = t >= total_tiles;
    cur_tile = cur_dead ? total_tiles - 1 : t;
    while (cur_hi <= cur_tile) { ++cur_b; cur_lo = cur_hi; cur_hi = s_tile_prefix[cur_b + 1]; }
  };
  auto tile_t0 = [&]() { return cur_dead ? a.L + 4096 : (cur_tile - cur_lo) * G::T_OUT; };   // dead: every row out of range
  auto tile_next = [&]() { cur_unit += tstep * unit_step; };

  if (warp == 0) {
    // =================== activation-slab TMA producer ===================
    const bool leader = ptx::elect_one();
    const int halo = G::P2 + rt.p1, nSlab = rt.nSlab;
    const uint32_t slab_tx = G::nCB1 * (uint32_t)rt.R1 * G::rb1, slab_bytes = rt.slab_bytes;
    Cursor sl{0, 0};
    for (int n = 0; n < my_tiles; ++n, tile_next()) {
      tile_seek();
      const int t0 = tile_t0();
      ptx::mbar_wait(slab_empty(sl.idx), sl.phase ^ 1u);
      trace(0, n, 0);
      if (leader && RB_DBG(64)) {
        if (rank == 0) ptx::mbar_arrive(slab_full(sl.idx));
      } else if (leader) {
        if (rank == 0) ptx::mbar_expect_tx(slab_full(sl.idx), slab_tx * G::NCL);      // pair: both CTAs' slabs land on the leader's barrier
#pragma unroll
        for (int cb = 0; cb < G::nCB1; ++cb) {
          if (G::PAIR) ptx::tma_load_3d_2sm(sSlab + sl.idx * slab_set + cb * slab_bytes, &tmY, slab_full(sl.idx), cb * G::BK1, t0 - halo, cur_b);
          else ptx::tma_load_3d(sSlab + sl.idx * slab_set + cb * slab_bytes, &tmY, slab_full(sl.idx), cb * G::BK1, t0 - halo, cur_b);
        }
      }
      sl.next(nSlab);
    }
    __syncwarp();
  } else if (warp == 3) {
    // =================== weight TMA producer ===================
    const bool leader = ptx::elect_one();
    if (G::STATIONARY) {
      // every (conv, channel block, tap) box stays resident: one barrier, loaded once per CTA
      if (leader && my_tiles > 0) {
        ptx::mbar_expect_tx(w_all, G::W_ALL);
        for (int cb = 0; cb < G::nCB1; ++cb)
          for (int j = 0; j < KT; ++j)
            ptx::tma_load_2d(sW + (cb * KT + j) * G::W1_TX, &tmW1, w_all, j * C + cb * G::BK1, 0);
        for (int cb = 0; cb < G::nCB2; ++cb)
          for (int j = 0; j < KT; ++j)
            ptx::tma_load_2d(sW2 + (cb * KT + j) * G::W2_TX, &tmW2, w_all, j * G::C2 + cb * G::BK2, 0);
      }
    } else {
      // one ring, filled in the issuing warp's consumption order: c1(0), [c1(n), c2(n-1)]..., c2(last)
      const int nW = rt.nW;
      Cursor wr{0, 0};
      auto load_w1 = [&]() {
        for (int cb = 0; cb < G::nCB1; ++cb)
          for (int j = 0; j < KT; ++j) {
            ptx::mbar_wait(w_empty(wr.idx), wr.phase ^ 1u);
            if (leader) {
              if RB_DBG(128) { if (rank == 0) ptx::mbar_arrive(w_full(wr.idx)); }
              else if (rank == 0) ptx::mbar_expect_tx(w_full(wr.idx), G::W1_TX * G::NCL);
              if RB_DBG(128) {}
              else if (G::PAIR) ptx::tma_load_2d_2sm(sW + wr.idx * G::W_SLOT, &tmW1, w_full(wr.idx), j * C + cb * G::BK1, rank * (C / 2));
              else ptx::tma_load_2d(sW + wr.idx * G::W_SLOT, &tmW1, w_full(wr.idx), j * C + cb * G::BK1, 0);
            }
            wr.next(nW);
          }
      };
      auto load_w2 = [&]() {
        for (int cb = 0; cb < G::nCB2; ++cb)
          for (int j = 0; j < KT; ++j) {
            ptx::mbar_wait(w_empty(wr.idx), wr.phase ^ 1u);
            if (leader) {
              if RB_DBG(128) { if (rank == 0) ptx::mbar_arrive(w_full(wr.idx)); }
              else if (rank == 0) ptx::mbar_expect_tx(w_full(wr.idx), G::W2_TX * G::NCL);
              if RB_DBG(128) {}
              else if (G::PAIR) ptx::tma_load_2d_2sm(sW + wr.idx * G::W_SLOT, &tmW2, w_full(wr.idx), j * G::C2 + cb * G::BK2, rank * (C / 2));
              else ptx::tma_load_2d(sW + wr.idx * G::W_SLOT, &tmW2, w_full(wr.idx), j * G::C2 + cb * G::BK2, 0);
            }
            wr.next(nW);
          }
      };
      for (int n = 0; n <= my_tiles; ++n) {
        if (n < my_tiles) load_w1();
        if (n >= 1) load_w2();
      }
    }
    __syncwarp();
  } else if (warp == 1 || warp == 2) {
    // =================== MMA issuers (warp-uniform loops, elected lane issues) ===================
    // resident weights: warp 1 issues c1, warp 2 issues c2, independently;
    // streamed weights: warp 1 issues both in ring order c1(n), c2(n-1); warp 2 idles.
    if (my_tiles > 0 && (warp == 1 || G::STATIONARY) && rank == 0) {
      const bool leader = ptx::elect_one();
      const int nSlab = rt.nSlab, nW = rt.nW;
      const uint32_t slab_bytes = rt.slab_bytes, idesc = rt.idesc;
      const uint32_t tap16 = ((uint32_t)rt.dil * G::rb1) >> 4;                  // c1 tap stride in 16-byte units
      const uint64_t dhi1 = ptx::umma_smem_desc(0, G::rb1, 0), dhi2 = ptx::umma_smem_desc(0, G::rb2, 0);
      Cursor sl{0, 0}, ab{0, 0}, gb{0, 0}, wr{0, 0};
      int tr_c1 = 0, tr_c2 = 0;
      if (G::STATIONARY) { ptx::mbar_wait(w_all, 0); ptx::tc_fence_after(); }
      // pair mode: one instruction drives both CTAs' tensor cores; commits arrive on both CTAs' barrier copies
      auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t id, uint32_t acc) {
        if RB_DBG(8) return;
        if (G::PAIR) ptx::umma_f16_2sm(d, ad, bd, id, acc); else ptx::umma_f16(d, ad, bd, id, acc);
      };
      auto commit = [&](uint32_t bar) { if (G::PAIR) ptx::umma_commit_2sm(bar); else ptx::umma_commit(bar); };
      auto do_c1 = [&]() {
        ptx::mbar_wait(slab_full(sl.idx), sl.phase);
        trace(1, tr_c1, 0);
        ptx::mbar_wait(acc1_empty(ab.idx), ab.phase ^ 1u);
        ptx::tc_fence_after();
        trace(6, tr_c1, 0);
        const uint32_t d_tmem = tmem_base + acc1_col + ab.idx * C;
        const uint32_t a16 = ((sSlab + sl.idx * slab_set) & 0x3FFFFu) >> 4;
#pragma unroll
        for (int cb = 0; cb < G::nCB1; ++cb) {
#pragma unroll
          for (int j = 0; j < KT; ++j) {
            uint32_t b16;
            if (G::STATIONARY) {
              b16 = ((sW + (cb * KT + j) * G::W1_TX) & 0x3FFFFu) >> 4;
            } else {
              ptx::mbar_wait(w_full(wr.idx), wr.phase);
              ptx::tc_fence_after();
              b16 = ((sW + wr.idx * G::W_SLOT) & 0x3FFFFu) >> 4;
            }
            // descriptors differ only in their 14-bit start-address field (16-byte units)
            const uint64_t ad0 = dhi1 | (uint64_t)(a16 + cb * (slab_bytes >> 4) + j * tap16), bd0 = dhi1 | (uint64_t)b16;
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < G::NK1; ++kk)
                mma(d_tmem, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (cb | j | kk) != 0 ? 1u : 0u);
              if (!G::STATIONARY) commit(w_empty(wr.idx));
            }
            if (!G::STATIONARY) wr.next(nW);
          }
        }
        if (leader) {
          commit(acc1_full(ab.idx));
          commit(slab_empty(sl.idx));                // c1 no longer reads the slab
        }
        trace(1, tr_c1, 1); ++tr_c1;
        sl.next(nSlab); ab.next(G::NBUF);
      };
      auto do_c2 = [&]() {
        ptx::mbar_wait(g_full(gb.idx), gb.phase);
        trace(2, tr_c2, 0);
        ptx::mbar_wait(acc2_empty(gb.idx), gb.phase ^ 1u);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc2_col + gb.idx * C;
        const uint32_t a16 = ((sG + gb.idx * G::G_SET) & 0x3FFFFu) >> 4;
#pragma unroll
        for (int cb = 0; cb < G::nCB2; ++cb) {
#pragma unroll
          for (int j = 0; j < KT; ++j) {
            uint32_t b16;
            if (G::STATIONARY) {
              b16 = ((sW2 + (cb * KT + j) * G::W2_TX) & 0x3FFFFu) >> 4;
            } else {
              ptx::mbar_wait(w_full(wr.idx), wr.phase);
              ptx::tc_fence_after();
              b16 = ((sW + wr.idx * G::W_SLOT) & 0x3FFFFu) >> 4;
            }
            const uint64_t ad0 = dhi2 | (uint64_t)(a16 + cb * (G::G_BYTES >> 4) + j * (G::rb2 >> 4)), bd0 = dhi2 | (uint64_t)b16;
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < G::NK2; ++kk)
                mma(d_tmem, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (cb | j | kk) != 0 ? 1u : 0u);
              if (!G::STATIONARY) commit(w_empty(wr.idx));
            }
            if (!G::STATIONARY) wr.next(nW);
          }
        }
        if (leader) {
          commit(acc2_full(gb.idx));
          commit(g_empty(gb.idx));
        }
        trace(2, tr_c2, 1); ++tr_c2;
        gb.next(G::NBUF);
      };
      if (G::STATIONARY) {
        if (warp == 1) { for (int n = 0; n < my_tiles; ++n) do_c1(); }
        else           { for (int n = 0; n < my_tiles; ++n) do_c2(); }
      } else {
        for (int n = 0; n <= my_tiles; ++n) {
          if (n < my_tiles) do_c1();
          if (n >= 1) do_c2();
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4 && ((warp - 4) & 4) == 0) {
    // =================== epilogue 1: gate -> G tile in shared memory ===================
    const int q = warp & 3, r = q * 32 + lane;
    constexpr int NCH = G::C2 / 16;              // 16-gate chunks per row
    // bias source: shared memory for C >= 128 (the L1 left next to ~220 KB of smem is tiny and the residual stream
    // flushes it); global / L1 for small C, where the smem bandwidth belongs to the MMA operands (N <= 64 is smem-bound)
    constexpr bool kBiasSmem = C >= kBiasSmemMinC;
    const uint32_t sb1 = ptx::smem_u32(&s_bias1[grp][0]);
    int bias_b = -1;                             // batch row whose cond/bias vector is staged in s_bias1[grp] (or held in pa / pb)
    // small C: the whole cond/bias row lives in registers (32 for C = 32), loaded when the batch row changes
    constexpr bool kBiasRegs = !kBiasSmem && G::C2 <= 32;
    float pa[kBiasRegs ? G::C2 : 1], pb[kBiasRegs ? G::C2 : 1];
    int len_b = -1, valid = a.L;                 // valid rows of batch row len_b (one global load per batch row, not per tile)
    constexpr int NCH_G = G::COOP ? NCH / G::NGRP : NCH;      // chunks of a row this group handles
    const int ch0 = G::COOP ? grp * NCH_G : 0;
    static_assert(!G::COOP || NCH % G::NGRP == 0, "column split");
    for (int n = n_first; n < my_tiles; n += G::TSTEP, tile_next()) {
      const Cursor gb{n % G::NBUF, (uint32_t)(n / G::NBUF) & 1u};
      // debug timeline: phases 8-11 = the hand-over from the traced tile to this group's next one (top of loop, set-up done,
      // accumulator full, G buffer free)
      auto gtrace = [&](int phase) {
        if (RB_TRACE_ON && (warp & 7) == 4 && n == rt.trace_t0 + 4 + n_first + G::TSTEP && a.trace && blockIdx.x == 0 && lane == 0)
          a.trace[(5 * 16 + phase) * 2 + grp] = clock64();
      };
      gtrace(8);
      tile_seek();
      const int b = cur_b, t0 = tile_t0();
      if (a.len && b != len_b) { valid = a.len[b] * a.len_mul; len_b = b; }
      const int t = t0 - G::P2 + r;
      const bool live = t >= 0 && t < valid;     // G outside the sequence is the conv's zero padding
      const float* gbias = a.bias1 + (long long)b * a.bias1_bstride;
      if (kBiasSmem && b != bias_b) {            // uniform over the group's 4 warps (same tile sequence)
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");      // everyone is done with the old vector
        const float* bias = a.bias1 + (long long)b * a.bias1_bstride;
        for (int i = r; i < C; i += 128) s_bias1[grp][i] = bias[i];
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        bias_b = b;
      }
      if (kBiasRegs && b != bias_b) {            // the row stays in registers until the batch row changes
        bias_b = b;
#pragma unroll
        for (int i = 0; i < G::C2; i += 4) {
          const float4 x = *reinterpret_cast<const float4*>(gbias + i), y = *reinterpret_cast<const float4*>(gbias + G::C2 + i);
          pa[kBiasRegs ? i : 0] = x.x; pa[kBiasRegs ? i + 1 : 0] = x.y; pa[kBiasRegs ? i + 2 : 0] = x.z; pa[kBiasRegs ? i + 3 : 0] = x.w;
          pb[kBiasRegs ? i : 0] = y.x; pb[kBiasRegs ? i + 1 : 0] = y.y; pb[kBiasRegs ? i + 2 : 0] = y.z; pb[kBiasRegs ? i + 3 : 0] = y.w;
        }
      }
      const int s = gb.idx;
      // bias / conditioning values of a 16-gate chunk: requested BEFORE the accumulator wait for the first chunk and right
      // after the previous chunk's math for the others - under this kernel's shared-memory load a dependent LDS takes
      // ~500 cycles (debug timeline), which used to sit between the accumulator becoming full and the first TMEM load.
      float ba[16], bb[16];
      auto load_bias = [&](int c0) {
        if (!RB_DBG(2048)) {
          if (kBiasSmem) {
            lds16(sb1 + 4 * c0, ba);
            lds16(sb1 + 4 * (G::C2 + c0), bb);
          } else if (kBiasRegs) {
#pragma unroll
            for (int i = 0; i < 16; ++i) { ba[i] = pa[kBiasRegs ? c0 + i : 0]; bb[i] = pb[kBiasRegs ? c0 + i : 0]; }
          } else {
            load16(gbias + c0, ba);
            load16(gbias + G::C2 + c0, bb);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) ba[i] = bb[i] = 0.f;
        }
      };
      if (kBiasAhead) load_bias(16 * ch0);
      gtrace(9);
      ptx::mbar_wait(acc1_full(s), gb.phase);
      gtrace(10);
      ptx::mbar_wait(g_empty(s), gb.phase ^ 1u);
      gtrace(11);
      ptx::tc_fence_after();
      if ((warp & 3) == 0) trace(3, n, 0);
      // per-phase timestamps of warps 4 and 12 (quarter 0 of each gate group) on tile 4 (role 5: slot = phase, kind = group)
      auto ptrace = [&](int phase) {
        if (RB_TRACE_ON && (warp & 7) == 4 && n == rt.trace_t0 + 4 + n_first && a.trace && blockIdx.x == 0 && lane == 0 && phase < 16)
          a.trace[(5 * 16 + phase) * 2 + grp] = clock64();
      };
      ptrace(0);
      const uint32_t trow = tmem_base + acc1_col + s * C + (static_cast<uint32_t>(q * 32) << 16);
      const uint32_t grow = sG + s * G::G_SET + r * G::rb2;
#pragma unroll
      for (int cc = 0; cc < NCH_G; ++cc) {
        const int c0 = 16 * (ch0 + cc);
        if (!kBiasAhead) load_bias(c0);
        uint32_t ra[16], rb[16];
        if (!RB_DBG(1024)) {
          ptx::tmem_ld16(trow + c0, ra);
          ptx::tmem_ld16(trow + G::C2 + c0, rb);
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) ra[i] = rb[i] = 0u;
        }
        if (kEarly && cc == NCH_G - 1) {         // c1's accumulator is in registers: hand it back before the last chunk's math
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) { if (G::PAIR) ptx::mbar_arrive_leader(acc1_empty(s)); else ptx::mbar_arrive(acc1_empty(s)); }
        }
        ptrace(1 + 3 * cc);                    // bias + TMEM loads done
        uint32_t p[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if RB_DBG(4) { p[i] = ra[i] ^ rb[i]; continue; }     // ablation: no gate math
          const float v0 = gate_fast(__uint_as_float(ra[2 * i]) + ba[2 * i], __uint_as_float(rb[2 * i]) + bb[2 * i]);
          const float v1 = gate_fast(__uint_as_float(ra[2 * i + 1]) + ba[2 * i + 1], __uint_as_float(rb[2 * i + 1]) + bb[2 * i + 1]);
          p[i] = live ? Pack16<OpT>::pack2(v0, v1) : 0u;
        }
        if (kBiasAhead && cc + 1 < NCH_G) load_bias(c0 + 16);
        // channel c0..c0+15 of row r -> block cb2, two 16-byte chunks, swizzled like a TMA write would
        const int cb2 = c0 / G::BK2, cin = c0 - cb2 * G::BK2;
        const uint32_t rowaddr = grow + cb2 * G::G_BYTES + cin * 2;
        ptrace(2 + 3 * cc);                    // gate math done
        if (!RB_DBG(4096)) {
          st_shared_v4(swz_addr<G::rb2>(rowaddr), p[0], p[1], p[2], p[3]);
          st_shared_v4(swz_addr<G::rb2>(rowaddr + 16), p[4], p[5], p[6], p[7]);
        }
        ptrace(3 + 3 * cc);                    // G stores issued
      }
      ptrace(13);
      fence_proxy_async_smem();                  // generic-proxy smem writes -> visible to this SM's tensor core (async proxy)
      ptrace(15);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (G::PAIR) ptx::mbar_arrive_leader(g_full(s)); else ptx::mbar_arrive(g_full(s));
        if (!kEarly) { if (G::PAIR) ptx::mbar_arrive_leader(acc1_empty(s)); else ptx::mbar_arrive(acc1_empty(s)); }
      }
      ptrace(14);
      if ((warp & 3) == 0) trace(3, n, 1);
    }
  } else if (warp >= 8) {
    // =================== epilogue 2: bias + residual + leaky_relu -> global ===================
    // y_out = op(leaky_relu(c2 + bias + x)).  The MRF average over the branches' final y streams (models.py:311-314) is
    // a separate coalesced pass (mrf_sum_kernel / conv_post's tile load): summing the other branches here cost one
    // exposed DRAM round trip per 16-column chunk of this row-per-thread epilogue.
    const int q = warp & 3, r = q * 32 + lane;
    const int nSlab = rt.nSlab;
    const uint32_t slab_bytes = rt.slab_bytes;
    const int srow = r + G::P2 + rt.p1;          // slab row holding y_in[t]
    constexpr int NCH = C / 16;                  // 16-column chunks per row
    const uint32_t sb2 = ptx::smem_u32(&s_bias2[0]);
    constexpr int NCH_S = G::COOP ? NCH / G::NGRP : NCH;      // chunks of a row this group handles (COOP: column split)
    const int ch0 = G::COOP ? grp * NCH_S : 0;
#ifndef VITSB_RB_PF64
#define VITSB_RB_PF64 2                          // C = 64 keeps two register sets (ping-pong, below): 2 chunks each, -1..-4 % against 4
#endif
#ifndef VITSB_RB_PF128
#define VITSB_RB_PF128 2                         // C = 128: two chunks in flight (k = 11: -2.8 % against 4, others neutral)
#endif
#ifndef VITSB_RB_PF256
#define VITSB_RB_PF256 4
#endif
    constexpr int PF = NCH_S < 4 ? NCH_S : (C == 64 ? VITSB_RB_PF64 : (C == 128 ? VITSB_RB_PF128 : (C == 256 ? VITSB_RB_PF256 : 4)));    // residual chunks kept in flight (2 x uint4 each; 96-register cap with 20 warps)
    Cursor sl{0, 0};
    int len_b = -1, valid = a.L;                 // valid rows of batch row len_b (one global load per batch row, not per tile)
    if (!G::COOP) for (int i = 0; i < grp; ++i) sl.next(nSlab);
    // output staging (kStage): row r of this group's block, written in place of a TMA-load image; thread 0 of the
    // group issues the stores and is the one that waits for them (bulk async-groups are per thread)
    const uint32_t stg = sStage + grp * G::STG_BYTES, stg_row = stg + r * G::RB_O;
    const bool issuer = r == 0;
    constexpr int CH_BLK = G::CB_O / 16;         // 16-column chunks per staged block
    static_assert(!G::COOP || (NCH_S % CH_BLK == 0 && NCH_S % PF == 0), "column split must be block aligned");
    // Residual rows from global memory (C >= 64) are software-pipelined ACROSS tiles: while the last PF chunks of a tile
    // are processed, the first PF chunks of this group's NEXT tile are already being loaded into the registers just
    // freed.  Loading them at the top of the next tile exposed one L2 round trip (~1.2 k cycles) per tile on the group
    // that sets the tile period of the k = 3 shapes (4.1 k cycles of work + that gap = the 5.3 k cycles measured).
#ifdef VITSB_RB_NO_LOOKAHEAD
    constexpr bool kLook = false;                 // A/B switch: load each tile's first chunks at the top of that tile
#else
    constexpr bool kLook = G::RES_GLOBAL;
#endif
    // Two register sets for the residual rows: the tile in hand reads `yq`, the look-ahead loads of this group's next tile
    // land in `yn`, and the tile loop is unrolled by two with the sets swapped.  With one loop-carried set ptxas rotated the
    // ~32 registers at the back edge, and the first of those moves waited for the loads just issued (9 % of the store
    // warps' samples, ahead of the accumulator release).
    uint4 yqA[PF][2], yqB[PF][2];
    // row pointer of this thread in tile nn of the CTA's sequence (nullptr: no such tile / row outside the tensor)
    auto res_row = [&](int nn) -> const uint4* {
      if (nn >= my_tiles) return nullptr;
      const int unit = unit_base + nn * unit_step;
      const int tt = G::PAIR ? unit * 2 + rank : unit;
      if (tt >= total_tiles) return nullptr;
      int bb = cur_b, lo = cur_lo;
      if (tt >= cur_hi) {                           // the next tile is in a later batch row: walk the table (rare)
        while (s_tile_prefix[bb + 1] <= tt) ++bb;
        lo = s_tile_prefix[bb];
      }
      const int tn = (tt - lo) * G::T_OUT + r;
      if (r >= G::T_OUT || tn >= a.L) return nullptr;
      return reinterpret_cast<const uint4*>(reinterpret_cast<const OpT*>(a.y_res) + ((long long)bb * a.L + tn) * C);
    };
    auto load_res = [&](const uint4* row, int c, uint4 (&dst)[2]) {
      dst[0] = dst[1] = make_uint4(0u, 0u, 0u, 0u);
      if (row && !RB_DBG(2)) { dst[0] = ldg_nc_v4_pinned(row + 2 * c); dst[1] = ldg_nc_v4_pinned(row + 2 * c + 1); }
    };
    if (kLook) {
      const uint4* row0 = res_row(n_first);
#pragma unroll
      for (int c = 0; c < PF; ++c) load_res(row0, ch0 + c, yqA[c]);
    }
    auto store_tile = [&](const int n, uint4 (&yq)[PF][2], uint4 (&yn)[PF][2]) {
      const Cursor ab{n % G::NBUF, (uint32_t)(n / G::NBUF) & 1u};
      tile_seek();
      const int b = cur_b, t0 = tile_t0();
      if (a.len && b != len_b) { valid = a.len[b] * a.len_mul; len_b = b; }
      const int t = t0 + r;
      const bool in_range = r < G::T_OUT && t < a.L;
      const bool live = t < valid;
      const int s = ab.idx, ss = sl.idx;
      const long long orow = ((long long)b * a.L + t) * C;
      const uint4* yres = in_range ? reinterpret_cast<const uint4*>(reinterpret_cast<const OpT*>(a.y_res) + orow) : nullptr;
      const uint4* ynext = kLook ? res_row(n + G::TSTEP) : nullptr;       // this group's next tile
      const uint32_t yrow = sSlab + ss * slab_set + srow * G::rb1;
      auto load_y = [&](int c, uint4 (&dst)[2]) {
        if (G::RES_GLOBAL) {
          load_res(yres, c, dst);
        } else {
          const int c0 = 16 * c, cb = c0 / G::BK1, cin = c0 - cb * G::BK1;
          const uint32_t yaddr = yrow + cb * slab_bytes + cin * 2;
          dst[0] = ld_shared_v4(swz_addr<G::rb1>(yaddr)); dst[1] = ld_shared_v4(swz_addr<G::rb1>(yaddr + 16));
        }
      };
      if (!kLook) {
        if (!G::RES_GLOBAL) ptx::mbar_wait(slab_full(ss), sl.phase);      // acquire the TMA-written slab for the residual reads
#pragma unroll
        for (int c = 0; c < PF; ++c) load_y(ch0 + c, yq[c]);
      }
      // conv2 bias of a chunk: requested before the accumulator wait (first chunk) / after the previous chunk's math
      float bv[16];
      auto load_bias2 = [&](int c0) { if (C >= kBiasSmemMinC) lds16(sb2 + 4 * c0, bv); else load16(a.bias2 + c0, bv); };
      constexpr bool kAhead2 = kBiasAhead && C < 256;          // C = 256: 4 residual chunks in flight leave no registers for it (spills)
      if (kAhead2) load_bias2(16 * ch0);
      ptx::mbar_wait(acc2_full(s), ab.phase);
      ptx::tc_fence_after();
      if ((warp & 3) == 0) trace(4, n, 0);
      const uint32_t trow = tmem_base + acc2_col + s * C + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll
      for (int cc = 0; cc < NCH_S; ++cc) {
        const int c = ch0 + cc, c0 = 16 * c;
        if (!kAhead2) load_bias2(c0);
        uint32_t ra[16];
        if (!RB_DBG(32)) { ptx::tmem_ld16(trow + c0, ra); ptx::tmem_ld_wait(); }
        else {
#pragma unroll
          for (int i = 0; i < 16; ++i) ra[i] = 0u;
        }
        if (kEarly && cc == NCH_S - 1) {         // the accumulator is in registers: hand it back before the math / staging / store tail
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) { if (G::PAIR) ptx::mbar_arrive_leader(acc2_empty(s)); else ptx::mbar_arrive(acc2_empty(s)); }
        }
        const uint4 y0 = yq[cc % PF][0], y1 = yq[cc % PF][1];
        if (cc + PF < NCH_S) load_y(c + PF, yq[cc % PF]);
        else if (kLook) load_res(ynext, ch0 + cc + PF - NCH_S, yn[cc % PF]);      // first chunks of the next tile
        const uint32_t yw[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
        float v[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          // inverse of leaky_relu(., 0.1): x = y >= 0 ? y : 10 y = min(y, 10 y)
#ifdef VITSB_NO_MIX16
          const float2 y = Pack16<OpT>::unpack2(yw[i]);
          v[2 * i] = (__uint_as_float(ra[2 * i]) + bv[2 * i]) + fminf(y.x, 10.f * y.x);
          v[2 * i + 1] = (__uint_as_float(ra[2 * i + 1]) + bv[2 * i + 1]) + fminf(y.y, 10.f * y.y);
#else
          v[2 * i] = __uint_as_float(ra[2 * i]) + bv[2 * i];
          v[2 * i + 1] = __uint_as_float(ra[2 * i + 1]) + bv[2 * i + 1];
          add_inv_lrelu2<OpT>(yw[i], v[2 * i], v[2 * i + 1]);      // mixed-precision add / fma straight from the packed halves
#endif
        }
        if (kAhead2 && cc + 1 < NCH_S) load_bias2(c0 + 16);
        if (!G::RES_GLOBAL && cc == NCH_S - 1) {   // every residual value read from the slab has been consumed: release it before
          __syncwarp();                            // the staging / store tail (measured neutral: more or earlier slabs do not move C = 32)
          if (lane == 0) ptx::mbar_arrive(slab_empty(ss));
        }
        if (kStage) {
          if (cc % CH_BLK == 0) {                // block buffer reuse: the previous store must have read it
            if (issuer) ptx::bulk_wait_read0();
            asm volatile("bar.sync %0, 128;" ::"r"(4 + grp) : "memory");
          }
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            // leaky_relu(v, 0.1) = max(v, 0.1 v); rows beyond len[b] are stored as zeros
            const uint32_t w = Pack16<OpT>::pack2(fmaxf(v[2 * i], 0.1f * v[2 * i]), fmaxf(v[2 * i + 1], 0.1f * v[2 * i + 1]));
            pk[i] = live ? w : 0u;
          }
          const uint32_t saddr = stg_row + (cc % CH_BLK) * 32;
          st_shared_v4(swz_addr<G::RB_O>(saddr), pk[0], pk[1], pk[2], pk[3]);
          st_shared_v4(swz_addr<G::RB_O>(saddr + 16), pk[4], pk[5], pk[6], pk[7]);
          if (cc % CH_BLK == CH_BLK - 1) {       // block complete: rows [t0, t0 + T_OUT) x 64 channels, clipped at L by the map
            fence_proxy_async_smem();
            asm volatile("bar.sync %0, 128;" ::"r"(4 + grp) : "memory");
            if (issuer && t0 < a.L && !RB_DBG(1)) {
              ptx::tma_store_3d(&tmO, stg, (c / CH_BLK) * G::CB_O, t0, b);
              ptx::bulk_commit();
            }
          }
        } else {
          if (in_range && !RB_DBG(1)) {
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              // leaky_relu(v, 0.1) = max(v, 0.1 v); rows beyond len[b] are stored as zeros
              const uint32_t w = Pack16<OpT>::pack2(fmaxf(v[2 * i], 0.1f * v[2 * i]), fmaxf(v[2 * i + 1], 0.1f * v[2 * i + 1]));
              pk[i] = live ? w : 0u;
            }
            const long long orow_st = RB_DBG(8192) ? (long long)r * C : orow;   // ablation: every tile stores to the same rows (no HBM write stream)
            uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<OpT*>(a.y_out) + orow_st + c0);
            d[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            d[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
        }
      }
      if (!kEarly) {
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (G::PAIR) ptx::mbar_arrive_leader(acc2_empty(s)); else ptx::mbar_arrive(acc2_empty(s)); }
      }
      if ((warp & 3) == 0) trace(4, n, 1);
#pragma unroll
      for (int i = 0; i < G::TSTEP; ++i) sl.next(nSlab);
    };
#ifdef VITSB_RB_PING_ALL
    constexpr bool kPing = kLook;
#else
    constexpr bool kPing = kLook && C == 64;      // measured per shape (ncu): C = 64 -3..-6 %, C = 128 / 256 +3..+5 % (column-split groups)
#endif
    if (kPing) {
      for (int n = n_first; n < my_tiles;) {
        store_tile(n, yqA, yqB); n += G::TSTEP; tile_next();
        if (n >= my_tiles) break;
        store_tile(n, yqB, yqA); n += G::TSTEP; tile_next();
      }
    } else {
      for (int n = n_first; n < my_tiles; n += G::TSTEP, tile_next()) store_tile(n, yqA, yqA);
    }
    if (kStage && issuer) ptx::bulk_wait0();     // staged rows are in global memory before the CTA retires
  }

  ptx::tc_fence_before();
  if (G::PAIR) ptx::cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if (G::PAIR) ptx::tmem_dealloc_2sm(tmem_base, G::TMEM_COLS); else ptx::tmem_dealloc(tmem_base, G::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
bool rb_fused_supported(int C, int k, int dil, int B) {
  if (!(C == 32 || C == 64 || C == 128 || C == 256)) return false;
  if (!(k == 3 || k == 7 || k == 11)) return false;
  if (dil < 1 || B > kRbMaxB) return false;
  if (((kTileM + (k - 1) * dil + 7) & ~7) > 256) return false;   // TMA box rows
  return true;
}

namespace {

template <typename OpT, int C, int KT>
cudaError_t launch_t(const RbLaunch& L, int op, const LaunchOpts& lo, cudaStream_t stream) {
  using G = RbT<C, KT>;
  RbRt rt{};
  rt.dil = L.dil; rt.p1 = (KT - 1) * L.dil / 2;
  rt.R1 = (kTileM + (KT - 1) * L.dil + 7) & ~7;
  rt.slab_bytes = ((uint32_t)rt.R1 * G::rb1 + 1023u) & ~1023u;
  const uint32_t slab_set = G::nCB1 * rt.slab_bytes;
  const uint32_t fixed = 1024 + 8 * kRbBars + 64 + G::NBUF * G::G_SET + G::STG_ALL;
  if (G::STATIONARY) {
    if (fixed + G::W_ALL + 2 * slab_set > kRbSmemBudget) return cudaErrorNotSupported;
    rt.nSlab = (int)((kRbSmemBudget - fixed - G::W_ALL) / slab_set);
    // slab lifetime: load + c1 only (RES_GLOBAL), or the whole 4-stage pipeline (C = 32 reads the residual from the slab:
    // ~6.7 k cycles from TMA issue to the store epilogue's release, so the tile period is that latency / nSlab - measured
    // 1.1 k cycles per tile with 6 slabs against ~0.4 k of tensor work; use all 8 barrier slots)
    // C = 64 (slab released after c1): measured per k with the rb_nslab option - 3 slabs for k = 3 (4: +2 %, 6: +12 %), 2 for k = 7 (-3 %)
    const int want64 = KT == 3 ? 3 : (KT == 7 ? 2 : 4);
    const int want = lo.rb_nslab > 0 ? (lo.rb_nslab > kRbMaxSlab ? kRbMaxSlab : lo.rb_nslab) : (G::RES_GLOBAL ? want64 : G::NBUF + 4);
    if (rt.nSlab > want) rt.nSlab = want;
  } else {
    // streamed weights: slabs first (2 when the accumulators are double-buffered), the rest to the ring
    rt.nSlab = G::NBUF >= 2 ? 2 : 1;
    if (fixed + rt.nSlab * slab_set + 2 * G::W_SLOT > kRbSmemBudget) rt.nSlab = 1;
    if (fixed + rt.nSlab * slab_set + 2 * G::W_SLOT > kRbSmemBudget) return cudaErrorNotSupported;
    rt.nW = (int)((kRbSmemBudget - fixed - rt.nSlab * slab_set) / G::W_SLOT);
    if (rt.nW > 8) rt.nW = 8;
  }
  rt.idesc = ptx::umma_idesc_f16(op == OP_BF16, kTileM * G::NCL, C);
#ifdef VITSB_RB_DEBUG
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("VITSB_RB_DBG"); dbg = e ? atoi(e) : 0; } rt.dbg = dbg; }
#endif

  TmaEncodeFn enc = tma_encode_fn();
  if (!enc) return cudaErrorNotSupported;
  const CUtensorMapDataType dt = op == OP_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  auto swz = [](int rb) { return rb == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (rb == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B); };
  CUtensorMap tmY, tmO, tmW1, tmW2;
  cuuint32_t estr[3] = {1, 1, 1};
  bool ok = true;
  {
    cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)L.a.L, (cuuint64_t)L.B};
    cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)L.a.L * C * 2};
    cuuint32_t box[3] = {(cuuint32_t)G::BK1, (cuuint32_t)rt.R1, 1};
    ok &= enc(&tmY, dt, 3, const_cast<void*>(L.y_in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz(G::rb1),
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  }
  {
    // output rows of a tile: [T_OUT][CB_O] boxes of y_out (only dereferenced by the staged mid-iteration kernels)
    cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)L.a.L, (cuuint64_t)L.B};
    cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)L.a.L * C * 2};
    cuuint32_t box[3] = {(cuuint32_t)G::CB_O, (cuuint32_t)G::T_OUT, 1};
    void* base = L.a.y_out ? L.a.y_out : const_cast<void*>(L.y_in);
    ok &= enc(&tmO, dt, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz(G::RB_O),
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)KT * C, (cuuint64_t)C};
    cuuint64_t strides[1] = {(cuuint64_t)KT * C * 2};
    cuuint32_t box[2] = {(cuuint32_t)G::BK1, (cuuint32_t)(C / G::NCL)};
    ok &= enc(&tmW1, dt, 2, const_cast<void*>(L.w1), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz(G::rb1),
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)KT * G::C2, (cuuint64_t)C};
    cuuint64_t strides[1] = {(cuuint64_t)KT * G::C2 * 2};
    cuuint32_t box[2] = {(cuuint32_t)G::BK2, (cuuint32_t)(C / G::NCL)};
    ok &= enc(&tmW2, dt, 2, const_cast<void*>(L.w2), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz(G::rb2),
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  }
  if (!ok) { fprintf(stderr, "rb_fused: tensor map encode failed (C=%d k=%d)\n", C, KT); return cudaErrorInvalidValue; }
  const size_t wbytes = G::STATIONARY ? (size_t)G::W_ALL : (size_t)rt.nW * G::W_SLOT;
  const size_t smem = 1024 + (size_t)rt.nSlab * slab_set + (size_t)G::NBUF * G::G_SET + wbytes + G::STG_ALL + 8 * kRbBars + 64;
  const int num_sms = lo.num_sms > 0 ? lo.num_sms : 148;
  static PerDeviceOnce smem_optin;     // one per template instantiation; the opt-in itself is per device
  const cudaError_t attr_err = smem_optin.run([] {
    // opt-in limit is 227 KB per CTA INCLUDING the kernel's static shared memory (tile table)
    cudaFuncAttributes fa{};
    cudaError_t e = cudaFuncGetAttributes(&fa, rb_fused_kernel<OpT, C, KT>);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(rb_fused_kernel<OpT, C, KT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               232448 - (int)fa.sharedSizeBytes);
    return e;
  });
  if (attr_err != cudaSuccess) { fprintf(stderr, "rb_fused: cudaFuncSetAttribute failed: %s\n", cudaGetErrorString(attr_err)); return attr_err; }
  // upper bound on tiles -> grid (persistent: at most one CTA per SM)
  const long long max_tiles = (long long)L.B * ((L.a.L + G::T_OUT - 1) / G::T_OUT);
  const long long max_units = (max_tiles + G::NCL - 1) / G::NCL;
  const int sm_units = num_sms / G::NCL;
  int units = (int)(max_units < sm_units ? max_units : sm_units);
  if (units < 1) units = 1;
  const int grid = units * G::NCL;
  RbArgs args = L.a;
  args.y_res = L.y_in;
  static long long* d_trace = nullptr;
  static int trace_C = -1, trace_K = 0, trace_left = 0, trace_first = 0;
#ifdef VITSB_RB_DEBUG
  if (trace_C < 0) {
    const char* e = getenv("VITSB_RB_TRACE");      // "C,k,count": dump the CTA-0 timeline of the first `count` launches of that shape
    trace_C = 0;
    if (e && sscanf(e, "%d,%d,%d,%d", &trace_C, &trace_K, &trace_left, &trace_first) >= 3) cudaMalloc(&d_trace, 7 * 16 * 2 * sizeof(long long));
  }
#endif
  const bool tracing = d_trace && trace_C == C && trace_K == KT && trace_left > 0;
  if (tracing) { cudaMemsetAsync(d_trace, 0, 7 * 16 * 2 * sizeof(long long), stream); args.trace = d_trace; rt.trace_t0 = trace_first; }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(G::THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = G::NCL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = lo.pdl ? 2 : 1;
  cudaError_t le = cudaLaunchKernelEx(&cfg, rb_fused_kernel<OpT, C, KT>, tmY, tmO, tmW1, tmW2, args, rt);
  if (le == cudaSuccess) le = cudaGetLastError();
  if (tracing && le == cudaSuccess) {
    --trace_left;
    long long h[7 * 16 * 2];
    cudaStreamSynchronize(stream);
    cudaMemcpy(h, d_trace, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = 0;
    for (long long v : h) if (v && (!t0 || v < t0)) t0 = v;
    const char* names[7] = {"slab.issue", "c1(slab_ok,done)", "c2(g_ok,done)", "gate(start,end)", "store(start,end)", "gate0 chunk phases", "c1.acc_ok"};
    fprintf(stderr, "rb_fused trace C=%d k=%d dil=%d nSlab=%d nW=%d (cycles since first event, CTA 0)\n", C, KT, L.dil, rt.nSlab, rt.nW);
    for (int r = 0; r < 7; ++r) {
      fprintf(stderr, "  %-18s", names[r]);
      for (int t = 0; t < (r == 5 ? 16 : 8); ++t) {
        const long long x = h[(r * 16 + t) * 2], y = h[(r * 16 + t) * 2 + 1];
        if (!x && !y) { if (r == 5) continue; break; }
        fprintf(stderr, " [%lld,%lld]", x ? x - t0 : -1, y ? y - t0 : -1);
      }
      fprintf(stderr, "\n");
    }
  }
  if (le != cudaSuccess)
    fprintf(stderr, "rb_fused launch failed (%s): C=%d k=%d dil=%d grid=%d smem=%zu nSlab=%d nW=%d\n", cudaGetErrorString(le),
            C, KT, L.dil, grid, smem, rt.nSlab, rt.nW);
  return le;
}

template <typename OpT>
cudaError_t dispatch(const RbLaunch& L, int op, const LaunchOpts& lo, cudaStream_t s) {
  if (!L.a.y_out) return cudaErrorInvalidValue;
#define VITSB_RB_CASE(CC, KK) if (L.C == CC && L.k == KK) return launch_t<OpT, CC, KK>(L, op, lo, s);
  VITSB_RB_CASE(32, 3) VITSB_RB_CASE(32, 7) VITSB_RB_CASE(32, 11)
  VITSB_RB_CASE(64, 3) VITSB_RB_CASE(64, 7) VITSB_RB_CASE(64, 11)
  VITSB_RB_CASE(128, 3) VITSB_RB_CASE(128, 7) VITSB_RB_CASE(128, 11)
  VITSB_RB_CASE(256, 3) VITSB_RB_CASE(256, 7) VITSB_RB_CASE(256, 11)
#undef VITSB_RB_CASE
  return cudaErrorNotSupported;
}

}  // namespace

cudaError_t launch_rb_fused(const RbLaunch& L, int op, const LaunchOpts& lo, cudaStream_t stream) {
  if (op == OP_F32 || !rb_fused_supported(L.C, L.k, L.dil, L.B)) return cudaErrorNotSupported;
  return op == OP_BF16 ? dispatch<__nv_bfloat16>(L, op, lo, stream) : dispatch<__half>(L, op, lo, stream);
}

}  // namespace vitsb
