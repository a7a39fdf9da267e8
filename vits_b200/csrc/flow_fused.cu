// One reverse ResidualCouplingLayer (modules.py:362-375, mean-only) with its WN (modules.py:157-182) as ONE
// persistent tcgen05 kernel, one CTA per 128-row time tile:
//
//     h  = pre(x0)                                               TMEM columns [0, 256)      fp32, resident
//     for l < n_layers:
//        a    = in_layers[l](h) + cond[b, l]      (k = 5, gated)   TMEM accumulator, 64 gate channels per pass
//        acts = tanh(a[:H]) * sigmoid(a[H:])                      [hi | lo] 16-bit tile in SHARED memory
//        h   += res_skip[l][:H](acts)             (l < last)      accumulated IN PLACE in TMEM by the MMA
//        m   += (post o res_skip[l][H:])(acts)                    TMEM columns [256, 352)
//     x1 -= m + b_m                                              fp32 flow state + [hi | lo] operand copy, in place
//
// The skip sum of the WN is only ever consumed by the 1x1 `post` conv (modules.py:366), which is linear, so the host
// composes post o skip_l into one 96 x 256 matrix per layer (engine_weights.cu) and the 256-channel skip accumulator
// never exists: the coupling mean m is accumulated directly.  That is what makes the state fit TMEM
// (h 256 + m 96 + gate accumulator 128 columns of 512), so that h, acts and m never touch HBM: a coupling reads the
// x0 half of z once and read-modify-writes the x1 half once.
//
// A tile computes rows [t0 - halo, t0 - halo + 128) of every layer and keeps the inner T_OUT = 128 - 2 * halo rows,
// halo = 2 * n_layers (each k = 5 layer consumes 2 rows per side): rows whose receptive field left the tile are
// garbage but row-independent, and rows outside [0, len[b]) are stored as zeros in the h operand, which is exactly the
// conv zero padding / x_mask of the reference.
//
// Roles (12 warps): warps 0 and 3 TMA producers (x0 rows; every weight box through one ring, in the MMA warp's fixed
// consumption order, alternate boxes each), warp 1 tcgen05.mma issuer (pair mode: the leader CTA's, for both CTAs), warp 2
// L2 prefetch of the coupling's weights, warps 4-11 epilogues (TMEM -> registers -> shared / global).
// Taps of the k = 5 conv are row-shifted UMMA descriptors over the h slab in shared memory (132 rows), as in rb_fused.
// Precision follows DESIGN.md section 3: in_layers single term on hi(h); pre / res / (post o skip) 3-term split
// (ah + al) * Wh + ah * Wl; exact-form gate (ex2 / rcp), because the split operand needs more than tanh.approx gives.
#include "common.h"
#include "kernels.h"
#include "ptx.h"
#include "tc_common.cuh"

#include <cstdio>
#include <cstdlib>
#include <type_traits>

namespace vitsb {

namespace ff {
constexpr int H = 256, IC = 192, HALF = 96, KT = 5, XW = 128;       // hidden, inter, inter / 2, flow kernel, x0 window columns
constexpr int NPASS = H / 64;                                      // in_layer passes: 64 gate channels = 128 packed columns each
constexpr int SLAB_ROWS = kTileM + KT - 1;                         // 132
constexpr uint32_t HBLK = ((uint32_t)SLAB_ROWS * 128u + 1023u) & ~1023u;   // bytes per 64-channel block of the h slab
constexpr uint32_t TILE_B = kTileM * 128u;                         // one [128 x 64] 16-bit operand tile
constexpr uint32_t SLOT_B = 2 * TILE_B;                            // acts slot: hi tile + lo tile
constexpr int NST1 = 5;                                            // ring stages per CTA (16 KB each; pair mode: 10 of 8 KB)
constexpr int NST_MAX = 2 * NST1;
constexpr uint32_t RING_B = NST1 * TILE_B;
constexpr uint32_t TM_H = 0, TM_M = 256, TM_ACC = 384, TM_COLS = 512;
constexpr int THREADS = 384;
constexpr int RS_ROWS = H + HALF;                                  // rows per layer of the res / (post o skip) matrix
constexpr int MAX_LAYERS = 4;

// barrier slots (8 bytes each)
enum { B_FULL = 0, B_EMPTY = B_FULL + NST_MAX, B_PRE = B_EMPTY + NST_MAX, B_AFULL = B_PRE + 2, B_AEMPTY = B_AFULL + 2,
       B_ACCF = B_AEMPTY + 2, B_ACCE, B_HREADY, B_HFULL, B_MREADY = B_HFULL + NPASS, B_MEMPTY, B_COUNT };

__host__ __device__ inline uint32_t smem_bytes(int nl) {
  return 1024u + NPASS * HBLK + 2 * SLOT_B + RING_B + (uint32_t)nl * 2 * H * 4u + 8u * B_COUNT + 64u;
}
}  // namespace ff

// PAIR: clusters of two CTAs, tcgen05 cta_group::2.  One MMA covers the 128-row tiles of both CTAs (M = 256) and each CTA
// holds HALF of every weight box, so the L2 -> shared-memory weight stream per SM halves.  That stream was the bound: a
// CTA needs 1.3 MB of in_layer weights per layer against ~20 k cycles of MMA, 64 B/clk, and the ring sustained ~33 B/clk
// (the issuing warp spent half of its time waiting for stages to fill: warp-state samples, profiles/).  The leader CTA's
// warp 1 issues for both; commits are multicast to both CTAs' barrier copies; the peer's producer / epilogue warps signal
// the leader's barriers remotely.  A pair whose second tile does not exist runs the protocol on a dead tile (no stores).
template <typename OpT, bool PAIR>
__global__ void __launch_bounds__(ff::THREADS, 1)
flow_fused_kernel(const __grid_constant__ CUtensorMap tmZ, const __grid_constant__ CUtensorMap tmPre,
                  const __grid_constant__ CUtensorMap tmIn, const __grid_constant__ CUtensorMap tmRS,
                  const __grid_constant__ CUtensorMap tmRSm, const FlowFusedArgs a) {
  using namespace ff;
  constexpr int NCL = PAIR ? 2 : 1;
  constexpr int NST = NST1 * NCL;                          // ring stages
  constexpr uint32_t STAGE_B = TILE_B / NCL;               // bytes of a stage in THIS CTA (its half of the box)
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_prefix[kMaxBatchRows + 1];
  __shared__ uint32_t s_tmem;

  const int nl = a.n_layers, halo = 2 * nl, t_out = kTileM - 2 * halo;
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sH = smem_base;                          // [NPASS][HBLK]   h slab, slab row s = tile row s - 2
  const uint32_t sA = sH + NPASS * HBLK;                  // [2][hi | lo]     acts chunk slots (and the x0 operand tiles of `pre`)
  const uint32_t sR = sA + 2 * SLOT_B;                    // [NST][STAGE_B]  weight ring
  const uint32_t sC = sR + RING_B;                        // fp32 [nl][2H]   conditioning + in_layer bias of the current batch row
  const uint32_t bars = sC + (uint32_t)nl * 2 * H * 4u;
  auto bar = [&](int i) { return bars + 8u * i; };
  float* s_cond = reinterpret_cast<float*>(smem_raw + (sC - ptx::smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int rank = PAIR ? (int)ptx::cluster_ctarank() : 0;
  const int unit_id = blockIdx.x / NCL, n_units = gridDim.x / NCL;      // a unit = one CTA, or one CTA pair

  // the four slab rows outside the tile are only ever read into rows that are discarded; keep them finite
  for (int i = threadIdx.x; i < NPASS * 4 * 8; i += blockDim.x) {
    const int blk = i >> 5, r4 = (i >> 3) & 3, ch = i & 7;
    const int srow = r4 < 2 ? r4 : SLAB_ROWS - 4 + r4;
    st_shared_v4(sH + blk * HBLK + srow * 128 + ch * 16, 0u, 0u, 0u, 0u);
  }
  fence_proxy_async_smem();
  if (warp == 2) build_tile_prefix(s_prefix, a.B, a.len, 1, a.T, t_out, lane, 0);
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmZ); ptx::prefetch_tmap(&tmPre); ptx::prefetch_tmap(&tmIn); ptx::prefetch_tmap(&tmRS); ptx::prefetch_tmap(&tmRSm);
    for (int s = 0; s < NST; ++s) { ptx::mbar_init(bar(B_FULL + s), 1); ptx::mbar_init(bar(B_EMPTY + s), 1); }
    // pair mode: the leader's copies collect the arrivals of both CTAs' epilogue warps
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(bar(B_PRE + s), 1); ptx::mbar_init(bar(B_AFULL + s), 8 * NCL); ptx::mbar_init(bar(B_AEMPTY + s), 1); }
    ptx::mbar_init(bar(B_ACCF), 1);   ptx::mbar_init(bar(B_ACCE), 8 * NCL);
    ptx::mbar_init(bar(B_HREADY), 1);
    for (int kb = 0; kb < NPASS; ++kb) ptx::mbar_init(bar(B_HFULL + kb), 4 * NCL);      // one per 64-channel block of the h slab
    ptx::mbar_init(bar(B_MREADY), 1); ptx::mbar_init(bar(B_MEMPTY), 8 * NCL);
    ptx::mbar_fence_init();
  }
  if (warp == 1) { if (PAIR) ptx::tmem_alloc_2sm(ptx::smem_u32(&s_tmem), TM_COLS); else ptx::tmem_alloc(ptx::smem_u32(&s_tmem), TM_COLS); }
  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync_all(); else __syncthreads();      // barrier inits visible to the peer before any remote signal
  ptx::tc_fence_after();
  const uint32_t tmem_base = s_tmem;
  // Programmatic dependent launch: everything above (and the L2 weight prefetch below, which reads constants only) may
  // overlap the previous kernel's tail; every other role waits here for that kernel's writes (z state, cond vectors).
  ptx::pdl_launch_dependents();
  if (warp != 2 && warp != 3) ptx::pdl_wait();      // warp 2 (L2 prefetch) and warp 3 (weight boxes) read constants only
  // clock64 timeline of CTA 0's first tile (debug build only): trace[slot] = cycles
#ifdef VITSB_RB_DEBUG
  auto trace = [&](int slot) { if (a.trace && blockIdx.x == 0 && lane == 0 && slot < 256) a.trace[slot] = clock64(); };
#else
  auto trace = [&](int) {};
#endif

  const int total_tiles = s_prefix[a.B];
  const int total_units = (total_tiles + NCL - 1) / NCL;  // pair mode: tiles 2u and 2u + 1 (the last may be missing)
  const int my_tiles = total_units > unit_id ? (total_units - 1 - unit_id) / n_units + 1 : 0;
  int cur_b = 0;
  // returns false for the dead tile of an odd pair: same coordinates as the last real tile, nothing stored
  auto tile_coords = [&](int n, int& b, int& t0) -> bool {        // b only grows along a CTA's tile sequence
    int tile = (unit_id + n * n_units) * NCL + rank;
    const bool real = tile < total_tiles;
    if (!real) tile = total_tiles - 1;
    while (s_prefix[cur_b + 1] <= tile) ++cur_b;
    b = cur_b;
    t0 = (tile - s_prefix[cur_b]) * t_out;
    return real;
  };
  const int fills_per_tile = 1 + 2 * nl;                  // fills of one acts slot per tile: x0 tiles of `pre` + two gate passes per layer

  if (warp == 2) {
    // =================== L2 prefetch of this coupling's weights ===================
    // Every CTA streams the same 6.8 MB of weights, in lockstep, through a ring that holds ~2 k cycles of tensor work:
    // when a box misses L2 every CTA waits out the HBM latency together (first timeline: 305-350 cycles per stage
    // against 256 of MMA).  Here the grid pulls the whole set into L2 once, each CTA a 1 / gridDim slice, roughly in
    // the order of first use; the ring loads that follow are L2 hits.
    constexpr uint32_t CH = 8192;
    const char* wbase = reinterpret_cast<const char*>(a.w);
    const size_t seg_off[3] = {a.fw.w_pre * 2, a.fw.w_in * 2, a.fw.w_rs * 2};
    const size_t seg_len[3] = {(size_t)H * 2 * XW * 2, (size_t)nl * 2 * H * KT * H * 2, (size_t)nl * RS_ROWS * 2 * H * 2};
    const int lane64 = lane;
    const size_t nch[3] = {(seg_len[0] + CH - 1) / CH, (seg_len[1] + CH - 1) / CH, (seg_len[2] + CH - 1) / CH};
    // chunk g of the concatenated segments goes to CTA g % gridDim, lane (g / gridDim) % 64
    for (size_t g = blockIdx.x + (size_t)gridDim.x * lane64; g < nch[0] + nch[1] + nch[2]; g += (size_t)gridDim.x * 32) {
      const int sg = g < nch[0] ? 0 : (g < nch[0] + nch[1] ? 1 : 2);
      const size_t c = g - (sg == 0 ? 0 : (sg == 1 ? nch[0] : nch[0] + nch[1]));
      const size_t off = c * CH;
      const uint32_t bytes = (uint32_t)(seg_len[sg] - off < CH ? seg_len[sg] - off : CH);
      ptx::prefetch_l2_bulk(wbase + seg_off[sg] + off, bytes);
    }
  } else if (warp == 0 || warp == 3) {
    // =================== TMA producers: x0 operand tiles + every weight box, in the MMA warp's order ===================
    // Two warps walk the same box sequence and each issues every other box (warp 0 also the x0 tiles): one warp needed
    // ~330 cycles per box (barrier poll, expect_tx, coordinate set-up, all serial uniform-datapath code) against 256 cycles
    // of MMA per ring stage, and the issuing warp spent half of its time waiting for stages to fill (warp-state samples).
    const bool leader = ptx::elect_one();
    const int my_par = warp == 0 ? 0 : 1;
    int seq = 0;
    Cursor rg{0, 0};
    auto put = [&](const CUtensorMap* m, int col, int row, uint32_t bytes) {
      if ((seq & 1) == my_par) {
        ptx::mbar_wait(bar(B_EMPTY + rg.idx), rg.phase ^ 1u);
        if (leader) {
          // pair: this CTA loads its half of the box rows; both halves complete on the leader's barrier
          // (`bytes` = this CTA's share: box rows / NCL, 128 bytes each)
          if (rank == 0) ptx::mbar_expect_tx(bar(B_FULL + rg.idx), bytes * NCL);
          if (PAIR) ptx::tma_load_2d_2sm(sR + rg.idx * STAGE_B, m, bar(B_FULL + rg.idx), col, row + rank * (int)(bytes / 128u));
          else ptx::tma_load_2d(sR + rg.idx * STAGE_B, m, bar(B_FULL + rg.idx), col, row);
        }
      }
      ++seq;
      rg.next(NST);
    };
    for (int n = 0; n < my_tiles; ++n) {
      int b, t0;
      tile_coords(n, b, t0);
      // x0 rows [t0 - halo, +128) as [hi | lo] 64-column tiles into the two acts slots (free once the previous tile's last
      // res / m MMAs have retired); rows outside [0, T) are zero-filled by TMA
      for (int s = 0; s < 2 && warp == 0; ++s) {
        const int k = n * fills_per_tile;                 // fill index of this slot; needs completion k - 1 of its empty barrier
        ptx::mbar_wait(bar(B_AEMPTY + s), (uint32_t)(k & 1) ^ 1u);
        if (leader) {
          if (rank == 0) ptx::mbar_expect_tx(bar(B_PRE + s), SLOT_B * NCL);
          if (PAIR) {
            ptx::tma_load_3d_2sm(sA + s * SLOT_B, &tmZ, bar(B_PRE + s), a.fw.xwin + 64 * s, t0 - halo, b);
            ptx::tma_load_3d_2sm(sA + s * SLOT_B + TILE_B, &tmZ, bar(B_PRE + s), IC + a.fw.xwin + 64 * s, t0 - halo, b);
          } else {
            ptx::tma_load_3d(sA + s * SLOT_B, &tmZ, bar(B_PRE + s), a.fw.xwin + 64 * s, t0 - halo, b);
            ptx::tma_load_3d(sA + s * SLOT_B + TILE_B, &tmZ, bar(B_PRE + s), IC + a.fw.xwin + 64 * s, t0 - halo, b);
          }
        }
      }
      // pre: [Wh | Wl] over the x0 window, two 128-row halves of the H output channels
      for (int part = 0; part < 2; ++part)
        for (int kb = 0; kb < 2; ++kb)
          for (int nh = 0; nh < 2; ++nh) put(&tmPre, part * XW + kb * 64, nh * 128, STAGE_B);
      for (int l = 0; l < nl; ++l) {
        const bool last = l == nl - 1;
        auto put_in = [&](int p) {
          for (int kbi = 0; kbi < NPASS; ++kbi) {           // block order 0, 2, 1, 3: the order the h epilogue completes them in
            const int kb = ((kbi & 1) << 1) | (kbi >> 1);
            for (int j = 0; j < KT; ++j) put(&tmIn, j * H + kb * 64, l * 2 * H + p * 128, STAGE_B);
          }
        };
        auto put_rs = [&](int p) {
          for (int part = 0; part < 2; ++part) {          // Wh then Wl
            const int col = part * H + 64 * p;
            if (!last) { put(&tmRS, col, l * RS_ROWS, STAGE_B); put(&tmRS, col, l * RS_ROWS + 128, STAGE_B); }
            put(&tmRSm, col, l * RS_ROWS + H, HALF * 128u / NCL);
          }
        };
        put_in(0);
        for (int p = 1; p < NPASS; ++p) { put_in(p); put_rs(p - 1); }
        put_rs(NPASS - 1);
      }
    }
    __syncwarp();
  } else if (warp == 1 && rank == 0) {
    // =================== MMA issuer (pair mode: the leader CTA's, for both CTAs) ===================
    const bool leader = ptx::elect_one();
    const uint64_t dhi = ptx::umma_smem_desc(0, 128, 0);
    const bool bf = sizeof(OpT) == 2 && !std::is_same<OpT, __half>::value;
    const uint32_t id128 = ptx::umma_idesc_f16(bf, kTileM * NCL, 128), id96 = ptx::umma_idesc_f16(bf, kTileM * NCL, HALF);
    auto d16 = [&](uint32_t addr) { return dhi | (uint64_t)((addr & 0x3FFFFu) >> 4); };
    Cursor rg{0, 0};
    // one ring stage against one [128 x 64] A tile: 4 K = 16 steps
    auto group = [&](uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool fresh) {
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          if (PAIR) ptx::umma_f16_2sm(d_tmem, adesc + 2 * kk, bdesc + 2 * kk, idesc, (fresh && kk == 0) ? 0u : 1u);
          else ptx::umma_f16(d_tmem, adesc + 2 * kk, bdesc + 2 * kk, idesc, (fresh && kk == 0) ? 0u : 1u);
        }
      }
    };
    auto stage_wait = [&]() -> uint64_t {
      ptx::mbar_wait(bar(B_FULL + rg.idx), rg.phase);
      ptx::tc_fence_after();
      return d16(sR + rg.idx * STAGE_B);
    };
    auto commit = [&](int b_) { if (leader) { if (PAIR) ptx::umma_commit_2sm(bar(b_)); else ptx::umma_commit(bar(b_)); } };
    auto stage_done = [&]() {
      commit(B_EMPTY + rg.idx);
      rg.next(NST);
    };
    for (int n = 0; n < my_tiles; ++n) {
      // ---- pre: h = x0 * Wpre (3-term split), TMEM h columns written fresh ----
      if (n == 0) trace(0);
      ptx::mbar_wait(bar(B_PRE + 0), n & 1);
      ptx::mbar_wait(bar(B_PRE + 1), n & 1);
      ptx::tc_fence_after();
      if (n == 0) trace(1);
      for (int part = 0; part < 2; ++part)
        for (int kb = 0; kb < 2; ++kb)
          for (int nh = 0; nh < 2; ++nh) {
            const uint64_t bd = stage_wait();
            const uint32_t d = tmem_base + TM_H + nh * 128;
            group(d, d16(sA + kb * SLOT_B), bd, id128, part == 0 && kb == 0);          // hi * (Wh | Wl)
            if (part == 0) group(d, d16(sA + kb * SLOT_B + TILE_B), bd, id128, false);  // lo * Wh
            stage_done();
          }
      commit(B_AEMPTY + 0); commit(B_AEMPTY + 1);
      commit(B_HREADY);
      if (n == 0) trace(2);
      for (int l = 0; l < nl; ++l) {
        const bool last = l == nl - 1;
        const int hh = n * nl + l;
        auto do_in = [&](int p) {
          const int ai = (n * nl + l) * NPASS + p;        // accumulator use index: needs completion ai - 1 of acc_empty
          ptx::mbar_wait(bar(B_ACCE), (uint32_t)(ai & 1) ^ 1u);
          ptx::tc_fence_after();
          if (n == 0) trace(16 + l * 16 + p * 2);           // in(p) may start
          const uint32_t d = tmem_base + TM_ACC;
          for (int kbi = 0; kbi < NPASS; ++kbi) {
            const int kb = ((kbi & 1) << 1) | (kbi >> 1);
            if (p == 0) {                                   // first pass of the layer: the h blocks arrive one by one
              ptx::mbar_wait(bar(B_HFULL + kb), hh & 1);
              ptx::tc_fence_after();
              if (n == 0 && kbi == 0) trace(3 + l);         // first h block of layer l ready
            }
            for (int j = 0; j < KT; ++j) {
              const uint64_t bd = stage_wait();
              group(d, d16(sH + kb * HBLK + j * 128), bd, id128, j == 0 && kbi == 0);    // tap j = slab shifted by j rows
              stage_done();
            }
          }
          commit(B_ACCF);
          if (n == 0) trace(16 + l * 16 + p * 2 + 1);       // in(p) issued
        };
        auto do_rs = [&](int p) {
          const int s = p & 1;
          const int fi = (n * nl + l) * 2 + (p >> 1);     // gate fill index of this slot
          ptx::mbar_wait(bar(B_AFULL + s), fi & 1);
          if (l == 0 && p == 0) ptx::mbar_wait(bar(B_MEMPTY), (uint32_t)(n & 1) ^ 1u);   // previous tile's m has been read
          ptx::tc_fence_after();
          if (n == 0) trace(16 + l * 16 + 8 + p * 2);       // rs(p) may start
          const uint64_t ahi = d16(sA + s * SLOT_B), alo = d16(sA + s * SLOT_B + TILE_B);
          for (int part = 0; part < 2; ++part) {
            if (!last) {
              for (int nh = 0; nh < 2; ++nh) {
                const uint64_t bd = stage_wait();
                const uint32_t d = tmem_base + TM_H + nh * 128;
                group(d, ahi, bd, id128, false);                                          // h += ...: accumulates in place
                if (part == 0) group(d, alo, bd, id128, false);
                stage_done();
              }
            }
            const uint64_t bd = stage_wait();
            const uint32_t d = tmem_base + TM_M;
            group(d, ahi, bd, id96, l == 0 && p == 0 && part == 0);
            if (part == 0) group(d, alo, bd, id96, false);
            stage_done();
          }
          commit(B_AEMPTY + s);
          if (n == 0) trace(16 + l * 16 + 8 + p * 2 + 1);   // rs(p) issued
        };
        do_in(0);
        for (int p = 1; p < NPASS; ++p) { do_in(p); do_rs(p - 1); }
        do_rs(NPASS - 1);
        commit(last ? B_MREADY : B_HREADY);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // =================== epilogues: 8 warps, warp = (TMEM lane quarter q, column half hf) ===================
    const int q = warp & 3, hf = (warp - 4) >> 2, r = q * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const float* bcum = a.bias + a.fw.b_cum;
    const float* bm = a.bias + a.fw.b_m;
    int staged_b = -1;
    auto arrive = [&](int b_) { if (PAIR) ptx::mbar_arrive_leader(bar(b_)); else ptx::mbar_arrive(bar(b_)); };
    for (int n = 0; n < my_tiles; ++n) {
      int b, t0;
      const bool real = tile_coords(n, b, t0);
      const int valid = a.len ? a.len[b] : a.T;
      const int t = t0 - halo + r;
      const bool live = real && t >= 0 && t < valid;
      if (b != staged_b) {                                // uniform over the 8 warps (same tile sequence)
        asm volatile("bar.sync 1, 256;" ::: "memory");    // everyone is done with the previous row's vector
        const float* cv = a.cond + (long long)b * a.cond_bstride + a.fw.cond_off;
        for (int i = threadIdx.x - 128; i < nl * 2 * H; i += 256) s_cond[i] = cv[i];
        asm volatile("bar.sync 1, 256;" ::: "memory");
        staged_b = b;
      }
      for (int l = 0; l < nl; ++l) {
        // ---- h operand of layer l: TMEM h + cumulative bias -> 16-bit slab rows (zero outside the sequence) ----
        ptx::mbar_wait(bar(B_HREADY), (n * nl + l) & 1);
        ptx::tc_fence_after();
        if (n == 0 && warp == 4) trace(96 + l * 16);         // h epilogue starts
#pragma unroll 1
        for (int half_i = 0; half_i < 2; ++half_i) {           // one 64-channel block of the slab per iteration
          const int kb = hf * 2 + half_i;
          uint32_t ra[4][16];
#pragma unroll
          for (int c = 0; c < 4; ++c) ptx::tmem_ld16(tmem_base + TM_H + lane_base + kb * 64 + c * 16, ra[c]);
          float bv[4][16];
#pragma unroll
          for (int c = 0; c < 4; ++c) load16(bcum + l * H + kb * 64 + c * 16, bv[c]);
          ptx::tmem_ld_wait();
          const uint32_t rowaddr = sH + kb * HBLK + (r + 2) * 128;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint32_t w = Pack16<OpT>::pack2(__uint_as_float(ra[c][2 * i]) + bv[c][2 * i], __uint_as_float(ra[c][2 * i + 1]) + bv[c][2 * i + 1]);
              pk[i] = live ? w : 0u;
            }
            st_shared_v4(swz_addr<128>(rowaddr + c * 32), pk[0], pk[1], pk[2], pk[3]);
            st_shared_v4(swz_addr<128>(rowaddr + c * 32 + 16), pk[4], pk[5], pk[6], pk[7]);
          }
          fence_proxy_async_smem();
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) arrive(B_HFULL + kb);
        }
        if (n == 0 && warp == 4) trace(96 + l * 16 + 1);     // h epilogue done
        // ---- gate passes: acts[64 p .. 64 p + 63] = tanh(a + cond_a) * sigmoid(b + cond_b) as [hi | lo] tiles ----
        for (int p = 0; p < NPASS; ++p) {
          const int s = p & 1;
          const int ai = (n * nl + l) * NPASS + p;
          const int k = n * fills_per_tile + 1 + 2 * l + (p >> 1);
          ptx::mbar_wait(bar(B_ACCF), ai & 1);
          ptx::tc_fence_after();
          if (n == 0 && warp == 4) trace(96 + l * 16 + 2 + p * 2);     // gate(p) starts
          const float* cb = s_cond + l * 2 * H + p * 128;
          // drain this warp's 32 gate channels (tanh + sigmoid halves) into registers and hand the accumulator back
          // at once: the MUFU-bound gate math below then runs under the next pass's MMAs
          uint32_t ra[2][16], rb[2][16];
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            ptx::tmem_ld16(tmem_base + TM_ACC + lane_base + hf * 32 + c * 16, ra[c]);
            ptx::tmem_ld16(tmem_base + TM_ACC + lane_base + 64 + hf * 32 + c * 16, rb[c]);
          }
          ptx::tmem_ld_wait();
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) arrive(B_ACCE);
          ptx::mbar_wait(bar(B_AEMPTY + s), (uint32_t)(k & 1) ^ 1u);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const int g0 = hf * 32 + c * 16;            // gate channel inside the pass
            float ba[16], bb[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) { ba[i] = cb[g0 + i]; bb[i] = cb[64 + g0 + i]; }
            uint32_t ph[8], pl[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float v0 = gate_exact3(__uint_as_float(ra[c][2 * i]) + ba[2 * i], __uint_as_float(rb[c][2 * i]) + bb[2 * i]);
              float v1 = gate_exact3(__uint_as_float(ra[c][2 * i + 1]) + ba[2 * i + 1], __uint_as_float(rb[c][2 * i + 1]) + bb[2 * i + 1]);
              if (!live) { v0 = 0.f; v1 = 0.f; }
              ph[i] = Pack16<OpT>::pack2(v0, v1);
              const float2 hi = Pack16<OpT>::unpack2(ph[i]);
              pl[i] = Pack16<OpT>::pack2(v0 - hi.x, v1 - hi.y);
            }
            const uint32_t addr = sA + s * SLOT_B + r * 128 + g0 * 2;
            st_shared_v4(swz_addr<128>(addr), ph[0], ph[1], ph[2], ph[3]);
            st_shared_v4(swz_addr<128>(addr + 16), ph[4], ph[5], ph[6], ph[7]);
            st_shared_v4(swz_addr<128>(addr + TILE_B), pl[0], pl[1], pl[2], pl[3]);
            st_shared_v4(swz_addr<128>(addr + TILE_B + 16), pl[4], pl[5], pl[6], pl[7]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) arrive(B_AFULL + s);
          if (n == 0 && warp == 4) trace(96 + l * 16 + 2 + p * 2 + 1); // gate(p) done
        }
      }
      // ---- x1 -= m + b_m on the tile's inner rows (modules.py:370-372; logs == 0), fp32 state + [hi | lo] operand copy ----
      ptx::mbar_wait(bar(B_MREADY), n & 1);
      ptx::tc_fence_after();
      if (n == 0 && warp == 4) trace(12);                    // m ready
      const bool mine = live && r >= halo && r < kTileM - halo;
      const long long row = (long long)b * a.T + t;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int ch = hf * 48 + c * 16;
        uint32_t ra[16];
        ptx::tmem_ld16(tmem_base + TM_M + lane_base + ch, ra);
        ptx::tmem_ld_wait();
        if (mine) {
          float bv[16], x[16];
          load16(bm + ch, bv);
          float* zp = a.zf + row * IC + a.fw.coff + ch;
          load16(zp, x);
#pragma unroll
          for (int i = 0; i < 16; ++i) x[i] = x[i] - (__uint_as_float(ra[i]) + bv[i]);
          store16(zp, x);
          store_op16<OpT>(reinterpret_cast<OpT*>(a.zop) + row * (2 * IC) + a.fw.coff + ch, x, 1.f, IC);
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) arrive(B_MEMPTY);
      if (n == 0 && warp == 4) trace(13);                    // tile done
    }
  }

  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if (PAIR) ptx::tmem_dealloc_2sm(tmem_base, TM_COLS); else ptx::tmem_dealloc(tmem_base, TM_COLS);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
bool flow_fused_supported(int inter, int hidden, int kernel, int n_layers, const int* dilation_rates, int n_flows, int op) {
  if (op != OP_F16 && op != OP_BF16) return false;
  if (inter != ff::IC || hidden != ff::H || kernel != ff::KT) return false;
  if (n_layers < 1 || n_layers > ff::MAX_LAYERS) return false;
  for (int i = 0; i < n_flows; ++i)
    if (dilation_rates[i] != 1) return false;
  return ff::smem_bytes(n_layers) + sizeof(int) * (kMaxBatchRows + 2) <= 232448;
}

int flow_fused_tile_rows(int n_layers) { return kTileM - 4 * n_layers; }

cudaError_t launch_flow_fused(const FlowFusedArgs& a, int op, const LaunchOpts& lo, cudaStream_t stream) {
  using namespace ff;
  if (op != OP_F16 && op != OP_BF16) return cudaErrorNotSupported;
  if (a.n_layers < 1 || a.n_layers > MAX_LAYERS || a.B < 1 || a.B > kMaxBatchRows) return cudaErrorNotSupported;
  TmaEncodeFn enc = tma_encode_fn();
  if (!enc) return cudaErrorNotSupported;
  const CUtensorMapDataType dt = op == OP_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const int nl = a.n_layers;
#ifdef VITSB_FF_NO_PAIR
  const bool pair = false;
#else
  const bool pair = lo.pair;
#endif
  const int ncl = pair ? 2 : 1;
  CUtensorMap tmZ, tmPre, tmIn, tmRS, tmRSm;
  cuuint32_t estr[3] = {1, 1, 1};
  bool ok = true;
  {
    cuuint64_t dims[3] = {(cuuint64_t)2 * IC, (cuuint64_t)a.T, (cuuint64_t)a.B};
    cuuint64_t strides[2] = {(cuuint64_t)2 * IC * 2, (cuuint64_t)a.T * 2 * IC * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)kTileM, 1};
    ok &= enc(&tmZ, dt, 3, a.zop, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  }
  auto wmap = [&](CUtensorMap* m, size_t off, int K, int rows, int box_rows) {
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    void* p = const_cast<char*>(reinterpret_cast<const char*>(a.w) + off * 2);
    ok &= enc(m, dt, 2, p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  };
  // pair mode: a CTA loads its half of the rows of every weight box
  wmap(&tmPre, a.fw.w_pre, 2 * XW, H, 128 / ncl);
  wmap(&tmIn, a.fw.w_in, KT * H, nl * 2 * H, 128 / ncl);
  wmap(&tmRS, a.fw.w_rs, 2 * H, nl * RS_ROWS, 128 / ncl);
  wmap(&tmRSm, a.fw.w_rs, 2 * H, nl * RS_ROWS, HALF / ncl);
  if (!ok) { fprintf(stderr, "flow_fused: tensor map encode failed\n"); return cudaErrorInvalidValue; }
  const size_t smem = smem_bytes(nl);
  static PerDeviceOnce smem_optin;
  const cudaError_t ae = smem_optin.run([] {
    cudaError_t e = cudaSuccess;
    cudaFuncAttributes fa{};
    auto optin = [&](auto kern) {
      if (e != cudaSuccess) return;
      e = cudaFuncGetAttributes(&fa, kern);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - (int)fa.sharedSizeBytes);
    };
    optin(flow_fused_kernel<__half, false>); optin(flow_fused_kernel<__half, true>);
    optin(flow_fused_kernel<__nv_bfloat16, false>); optin(flow_fused_kernel<__nv_bfloat16, true>);
    return e;
  });
  if (ae != cudaSuccess) { fprintf(stderr, "flow_fused: cudaFuncSetAttribute failed: %s\n", cudaGetErrorString(ae)); return ae; }
  const int t_out = flow_fused_tile_rows(nl);
  const long long max_tiles = a.m_tiles > 0 ? a.m_tiles : (long long)a.B * ((a.T + t_out - 1) / t_out);
  const int num_sms = lo.num_sms > 0 ? lo.num_sms : 148;
  const long long max_units = (max_tiles + ncl - 1) / ncl;
  const int sm_units = num_sms / ncl;
  int units = (int)(max_units < sm_units ? max_units : sm_units);
  if (units < 1) units = 1;
  const int grid = units * ncl;
  FlowFusedArgs args = a;
#ifdef VITSB_RB_DEBUG
  static long long* d_trace = nullptr;
  static int trace_left = -1;
  if (trace_left < 0) { const char* e = getenv("VITSB_FF_TRACE"); trace_left = e ? atoi(e) : 0; if (trace_left > 0) cudaMalloc(&d_trace, 256 * sizeof(long long)); }
  const bool tracing = d_trace && trace_left > 0;
  if (tracing) { cudaMemsetAsync(d_trace, 0, 256 * sizeof(long long), stream); args.trace = d_trace; }
#endif
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = ncl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = lo.pdl ? 2 : 1;
  cudaError_t le;
  if (pair) le = op == OP_BF16 ? cudaLaunchKernelEx(&cfg, flow_fused_kernel<__nv_bfloat16, true>, tmZ, tmPre, tmIn, tmRS, tmRSm, args)
                               : cudaLaunchKernelEx(&cfg, flow_fused_kernel<__half, true>, tmZ, tmPre, tmIn, tmRS, tmRSm, args);
  else le = op == OP_BF16 ? cudaLaunchKernelEx(&cfg, flow_fused_kernel<__nv_bfloat16, false>, tmZ, tmPre, tmIn, tmRS, tmRSm, args)
                          : cudaLaunchKernelEx(&cfg, flow_fused_kernel<__half, false>, tmZ, tmPre, tmIn, tmRS, tmRSm, args);
  if (le == cudaSuccess) le = cudaGetLastError();
#ifdef VITSB_RB_DEBUG
  if (tracing && le == cudaSuccess) {
    --trace_left;
    long long h[256];
    cudaStreamSynchronize(stream);
    cudaMemcpy(h, d_trace, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = h[0];
    auto rel = [&](int i) { return h[i] ? h[i] - t0 : -1; };
    fprintf(stderr, "flow_fused trace (cycles, CTA 0 tile 0; grid=%d): pre_wait %lld pre_issued %lld m_ready %lld tile_done %lld\n", grid, rel(1), rel(2), rel(12), rel(13));
    for (int l = 0; l < nl; ++l) {
      fprintf(stderr, "  L%d mma: h_full %lld |", l, rel(3 + l));
      for (int p = 0; p < 4; ++p) fprintf(stderr, " in%d[%lld,%lld]", p, rel(16 + l * 16 + p * 2), rel(16 + l * 16 + p * 2 + 1));
      for (int p = 0; p < 4; ++p) fprintf(stderr, " rs%d[%lld,%lld]", p, rel(16 + l * 16 + 8 + p * 2), rel(16 + l * 16 + 8 + p * 2 + 1));
      fprintf(stderr, "\n  L%d epi: h[%lld,%lld]", l, rel(96 + l * 16), rel(96 + l * 16 + 1));
      for (int p = 0; p < 4; ++p) fprintf(stderr, " g%d[%lld,%lld]", p, rel(96 + l * 16 + 2 + p * 2), rel(96 + l * 16 + 2 + p * 2 + 1));
      fprintf(stderr, "\n");
    }
  }
#endif
  if (le != cudaSuccess) fprintf(stderr, "flow_fused launch failed (%s): grid=%d smem=%zu nl=%d\n", cudaGetErrorString(le), grid, smem, nl);
  return le;
}

}  // namespace vitsb
