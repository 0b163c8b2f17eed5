// tcgen05 / TMEM / bulk-TMA implicit-GEMM encoder (AT_PREC_BF16, AT_PREC_FP16).
//
// Replaces the cuDNN calls behind the reference backbone (model_utils.py:107-138,
// backbone_resnet.py:14-158).  Design ("shift-GEMM on a padded planar layout"):
//
//  * Activations live in HBM as 16-bit NC/8HW8 planes: plane cg holds channels [8cg, 8cg+8) of
//    every pixel as one 16-byte unit; pixels of the whole crop batch are laid out in ONE padded
//    linear order  g = img*Hp*Wp + y*Wp + x  with Hp = H+1, Wp = W+1: one zero column and one
//    zero row per image are shared between neighbours (x = W is the left pad of the next row,
//    y = H the top pad of the next image), plus zero guards in front of / behind the batch.
//  * With that layout a 3x3 tap (r,s) of a stride-1 convolution is a pure SHIFT of the pixel
//    index by (r-1)*Wp + (s-1).  A CTA loads the input pixels [g0 - Wp - 1, g0 + 128 + Wp + 1)
//    of each channel plane ONCE with 1-D bulk TMA (cp.async.bulk, contiguous in HBM and in
//    shared memory) and issues all 9 taps as tcgen05.mma instructions whose A descriptors differ
//    only in their start address (no-swizzle K-major canonical layout: LBO = plane stride,
//    SBO = 128 B).  No im2col buffer, no re-load per tap, no tensor maps.
//  * Stride-2 convolutions read a 4-phase (space-to-depth) copy of their input that the
//    producing layer's epilogue writes directly, so they are shift-GEMMs too.
//  * GEMM view: M = 128 padded-linear pixels per tile (TMEM lanes), N = Cout (padded to 16),
//    K = taps * Cin in chunks of <= 64 channels.  Accumulators live in TMEM (fp32), 2-4 stages.
//    Warp roles: warp 0 weight producer (bulk TMA ring, or resident weights when they fit),
//    warp 1 single-thread MMA issuer, warp 2 TMEM allocator + activation producer, warps 4-11
//    two epilogue groups (tcgen05.ld -> +bias +residual, ReLU -> 16-bit pack -> coalesced 16 B
//    stores, pads written as zeros so the layout invariant is kept without memsets).
//  * Persistent grid of one CTA per SM (148), tiles round-robin so neighbouring tiles (which
//    share halo pixels) are in flight together and the halo re-reads hit L2.
#include <algorithm>
#include <mutex>
#include <stdlib.h>
#include <string.h>
#include <utility>
#include <vector>

#include "conv.cuh"
#include "tc_ptx.cuh"

namespace abt {
namespace tc {

constexpr int kTileM = 128;
constexpr int kThreads = 384;
// Register cap.  The register file is split per SM sub-partition (4 x 16384): 3 warps x 160 regs
// leave 1024 registers per sub-partition -- room for one 32-register warp of another kernel (the
// source-box pull of the next step, crop_resample.cu) to stay co-resident.  At 168 (what
// __launch_bounds__(384, 1) allows) a resident side warp blocks the launch of the next conv CTA.
constexpr int kMaxRegs = 160;
constexpr int kMaxTaps = 9;
constexpr int kSmemBudget = 220 * 1024;
constexpr int kS2dC = 64;                                  // channels of the x-space-to-depth view of layer 1
// S2D weight blocks, compact: the centre taps (di = 0) as [8][64][8] (8 KB), the side taps only with the four K groups
// they use ([4][64][8], 4 KB: K groups 4-7 for di = -1, 0-3 for di = +1) - in 4 KB units tap t starts at kS2dUnit[t]
constexpr int64_t kS2dUnitBytes = 4096;
constexpr int kS2dUnits = 12;
constexpr int64_t kS2dMainBytes = kS2dUnits * kS2dUnitBytes;
constexpr int64_t kS2dIdBytes = kS2dC * kS2dC * 2;


// q = n / d for the bounded n of this kernel: q = umulhi(n, magic) >> shift (host: make_divider)
struct Divider {
  uint32_t magic, shift, d;
};

struct ConvTcParams {
  const uint4* a_src[4];
  long long a_plane_px;
  const uint4* w;
  const float* bias;
  const uint4* res;
  long long res_plane_px;
  uint4* out;
  long long out_plane_px;
  uint4* out_ph[4];
  long long out_ph_plane_px;
  long long out_ph_stride;   // units between consecutive phase images (S2D epilogue)
  float* out_f32;
  // folded 1x1 stride-2 shortcut (first block of layers 2-4): sc_nk extra K chunks read phase image
  // (0,0) of the block input at shift 0 and accumulate W_sc x into the same accumulator
  const uint4* sc_src;
  long long sc_plane_px;
  const float* bias2;     // second bias added to `bias` (the shortcut's BN shift)
  int sc_nk, sc_planes, sc_ksteps;
  uint32_t sc_a_bytes, sc_w_bytes;
  const float* res_f32;   // MODE 4: optional fp32 NHWC residual (channel stride f32_ld)
  float out_scale;        // MODE 4: accumulator scale (undoes the 2^8 pre-scale of the split weights)
  int split_planes;       // MODE 4: channel groups per part of the [hi | lo | hi] split output
  int f32_ld;             // MODE 4: channel stride of out_f32 / res_f32
  int out_mode, relu, fmt;
  int n_src, n_taps, n_kchunks, kc;
  int n, cout;
  int tap_src[kMaxTaps], tap_shift[kMaxTaps];
  int min_shift, npix;
  int H, W, Hp, Wp, Hp2, Wp2;
  long long n_img, P;
  int mt;
  long long n_tiles;
  int ra, ws, as, w_resident;
  uint32_t smem_a_off, smem_w_off, smem_bias_off, smem_bar_off;
  uint32_t a_slot_bytes, a_plane_bytes, w_block_bytes, tmem_cols, acc_cols;
  Divider div_hw, div_wp;  // g / (Hp*Wp), rem / Wp
  int slab_shift;          // log2(32-column slabs per accumulator) (0 when not a power of two: MT == 1 then)
  long long* prof;  // optional [grid][16] cycle counters (debug)
  int dbg;          // debug knobs (AT_TC_DBG): 1 no epilogue, 2 no A copies
  int32_t* status;  // optional device status word: bit 1 (value 2) |= an fp16 activation of this layer saturated / is NaN
  // Work items of conv_tc_kernel: tiles [0, t_full) as a whole, then every remaining tile (the partial last wave of the
  // persistent grid) as n_split = 2^split_shift items of n / n_split output channels each, so that the last wave costs
  // 1 / n_split of a tile time instead of a whole one (layer 3: 9.13 -> 10 waves without it).  n_split = 1: plain tiles.
  int t_full, n_split, split_shift, n_items;
  // cta_group::2 (conv_tc2_kernel): each CTA of a pair holds n / 2 output channels of every weight block
  int bias_wrap;          // > 0: bias index = channel % bias_wrap (S2D: both pixel parities share the layer's 32 biases)
  uint32_t w_src_bytes;   // bytes of a whole weight block in global memory (= block stride)
  uint32_t w_kg_bytes;    // bytes of one 8-channel K group of a whole block in global memory (n * 16)
};

// Ring cursor without divisions: slot index + phase parity.
struct Ring {
  int slot = 0;
  uint32_t phase = 0;
  __device__ __forceinline__ void advance(int n) {
    if (++slot == n) { slot = 0; phase ^= 1u; }
  }
};

__device__ __forceinline__ uint32_t fast_div(uint32_t n, const Divider& dv) { return __umulhi(n, dv.magic) >> dv.shift; }

// one 32- or 16-column slab of an accumulator row -> +bias (+residual) -> ReLU -> 16-bit pack -> store
struct EpiCoord {
  bool valid;
  long long g;      // padded-linear pixel index (planar modes)
  long long o1;     // mode 1: index inside a phase image; mode 2: float index of channel 0
  int ph;
};

// MODE: 0 planar out, 1 planar out + residual, 2 phase-split out + residual, 3 fp32 NHWC out,
// 5 phase-split out without residual (the residual was folded into the accumulator as an identity tap),
// 4 error-compensated ("fp16 x 3") conv of the tail: the input planes are [hi | lo | hi] halves of
// fp32 activations, the weights [hi | hi | lo] halves pre-scaled by 2^8; the epilogue rescales, adds
// bias and an optional fp32 NHWC residual, applies ReLU, and writes fp32 NHWC and/or the split planes
// of the next conv.
// WRES: the layer's weights stay resident in shared memory (else they stream through a ring).
// S2D: layer 1 as a 64 -> 64 conv over the x-space-to-depth view of its 48 x 48 x 32 maps (rows of 24 pixel PAIRS,
// channels [even pixel 0..31 | odd pixel 0..31] = planes 0-3 | 4-7): one accumulator row holds two output pixels
// (TMEM columns [0,32) | [32,64)), so the A operand - what an N = 32 MMA spends its time reading - is read once
// for both.  Tap (dy, di) of that view carries W[dy][dx = 2 di + pi - po]: the centre column is dense (K = 64,
// N = 64), di = -1 only connects odd inputs to even outputs (K steps 2-3, N = columns 0-31) and di = +1 even inputs
// to odd outputs (K steps 0-1, columns 32-63); the other halves are structural zeros and are not issued.
template <int NTAPS, int KSTEPS, int MT, int WRES, int MODE, int S2D = 0>
__global__ void __maxnreg__(kMaxRegs) conv_tc_kernel(const __grid_constant__ ConvTcParams p) {
  static_assert(!S2D || (NTAPS == 9 && KSTEPS == 4 && WRES == 1 && (MODE == 0 || MODE == 5)), "S2D: 3x3, 64 -> 64, resident weights");
  extern __shared__ __align__(128) uint8_t smem[];
  using namespace ptx;
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar0 = sbase + p.smem_bar_off;
  const uint32_t a_full = bar0, a_empty = a_full + 8 * p.ra;
  const uint32_t w_full = a_empty + 8 * p.ra, w_empty = w_full + 8 * p.ws;
  const uint32_t acc_full = w_empty + 8 * p.ws, acc_empty = acc_full + 8 * p.as;
  const uint32_t tmem_slot = acc_empty + 8 * p.as;
  float* bias_s = reinterpret_cast<float*>(smem + p.smem_bias_off);
  if (p.prof != nullptr && threadIdx.x == 0) p.prof[blockIdx.x * 16 + 5] = (long long)globaltimer_ns();
  // programmatic dependent launch: the next conv of the stream may start its prologue (barriers,
  // TMEM, bias, resident weights: nothing that depends on this grid) as soon as SMs free up
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.ra; ++i) { mbar_init(a_full + 8 * i, 1); mbar_init(a_empty + 8 * i, 1); }
    for (int i = 0; i < p.ws; ++i) { mbar_init(w_full + 8 * i, 1); mbar_init(w_empty + 8 * i, 1); }
    for (int i = 0; i < p.as; ++i) { mbar_init(acc_full + 8 * i, 1); mbar_init(acc_empty + 8 * i, 4); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < p.n; i += kThreads)
    bias_s[i] = i < p.cout ? p.bias[p.bias_wrap > 0 ? i % p.bias_wrap : i] + (p.bias2 != nullptr ? p.bias2[i] : 0.f) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + (tmem_slot - sbase));

  const int first_tile = blockIdx.x;   // first work item of this CTA; items are strided by the grid size
  const int n_tiles = p.n_items;
  const int n_my = first_tile < n_tiles ? (n_tiles - first_tile + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  // work item -> (tile, channel slice sp); sp < 0: the whole tile
  auto item_tile = [&](int w, int& sp) -> int {
    if (w < p.t_full) { sp = -1; return w; }
    const int r = w - p.t_full;
    sp = r & (p.n_split - 1);
    return p.t_full + (r >> p.split_shift);
  };
  const int n_part = p.n >> p.split_shift;   // output channels of a slice
  const int nk = p.n_kchunks;
  // NTAPS == 8 is the stem GEMM: 8 MMAs (4 pooling phases x 2 rows), one 8-channel input plane,
  // the two K halves of an MMA are two neighbouring pixels (LBO = 16 B), 4 accumulators.
  constexpr bool kStem = NTAPS == 8;
  constexpr int kPlanesPerChunk = kStem ? 1 : KSTEPS * 2;
  constexpr int kTilePx = kTileM * MT;

  if (warp == 0) {
    if (lane == 0) {  // ---------------- weight producer
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.w);
      const int nmain = nk * NTAPS;
      const int nblocks = nmain + p.sc_nk;  // the shortcut blocks follow the main ones at the same stride
      if (S2D) {  // compact blocks, loaded as 4 KB units (12 + 2 for the identity)
        if (n_my > 0)
          for (int u = 0; u < kS2dUnits + 2 * p.sc_nk; ++u) {
            mbar_expect_tx(w_full + 8 * u, (uint32_t)kS2dUnitBytes);
            bulk_g2s(sbase + p.smem_w_off + u * (uint32_t)kS2dUnitBytes, wsrc + (size_t)u * kS2dUnitBytes, (uint32_t)kS2dUnitBytes,
                     w_full + 8 * u);
          }
      } else if (WRES) {
        if (n_my > 0)
          for (int b = 0; b < nblocks; ++b) {
            const uint32_t bytes = b < nmain ? p.w_block_bytes : p.sc_w_bytes;
            mbar_expect_tx(w_full + 8 * b, bytes);
            bulk_g2s(sbase + p.smem_w_off + b * p.w_block_bytes, wsrc + (size_t)b * p.w_block_bytes, bytes, w_full + 8 * b);
          }
      } else {
        Ring r;
        for (int it = 0; it < n_my; ++it)
          for (int b = 0; b < nblocks; ++b) {
            const uint32_t bytes = b < nmain ? p.w_block_bytes : p.sc_w_bytes;
            mbar_wait(w_empty + 8 * r.slot, r.phase ^ 1u);
            mbar_expect_tx(w_full + 8 * r.slot, bytes);
            bulk_g2s(sbase + p.smem_w_off + r.slot * p.w_block_bytes, wsrc + (size_t)b * p.w_block_bytes, bytes,
                     w_full + 8 * r.slot);
            r.advance(p.ws);
          }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: the whole warp runs the control flow converged (warp-uniform
    // values stay in uniform registers), one elected lane issues tcgen05.mma / commit.  Taps,
    // K steps and sub-tiles are compile-time so the issue block is straight-line code: two
    // uniform adds per descriptor, no loop control between the MMAs.
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t idesc = umma_idesc(kTileM, p.n, p.fmt);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);                       // SBO = 128 B, version 1
    const uint32_t a_lo0 = ((kStem ? 1u : (p.a_plane_bytes >> 4)) << 16) + ((sbase + p.smem_a_off) >> 4);
    const uint32_t b_lo0 = ((uint32_t)p.n << 16) + ((sbase + p.smem_w_off) >> 4);   // LBO = n*16 B
    const uint32_t a_slot16 = p.a_slot_bytes >> 4, w_blk16 = p.w_block_bytes >> 4;
    const uint32_t a_kstep16 = (2 * p.a_plane_bytes) >> 4, b_kstep16 = (uint32_t)(2 * p.n);
    uint32_t tap_off[NTAPS];
#pragma unroll
    for (int t = 0; t < NTAPS; ++t)
      tap_off[t] = ((uint32_t)p.tap_src[t] * kPlanesPerChunk * p.a_plane_bytes >> 4) + (uint32_t)(p.tap_shift[t] - p.min_shift);
    Ring ra, rw, rc;
    long long t_acc = 0, t_a = 0, t_w = 0, t_start = clock64(), t0;
    const bool prof = p.prof != nullptr && (p.dbg & 16);
    if (p.prof != nullptr && lane == 0) p.prof[blockIdx.x * 16 + 7] = (long long)globaltimer_ns();
    const uint32_t idesc_full = idesc, idesc_part = umma_idesc(kTileM, n_part, p.fmt);
    const uint32_t idesc_half = umma_idesc(kTileM, 32, p.fmt);   // S2D side taps
    (void)idesc_half;
    for (int it = 0; it < n_my; ++it) {
      int sp;
      item_tile(first_tile + it * (int)gridDim.x, sp);
      const uint32_t idesc = sp < 0 ? idesc_full : idesc_part;
      const uint32_t b_sp = sp < 0 ? 0u : (uint32_t)(sp * n_part);   // rows of B skipped (16-byte units)
      if (prof) t0 = clock64();
      mbar_wait(acc_empty + 8 * rc.slot, rc.phase ^ 1u);
      if (prof) t_acc += clock64() - t0;
      tc_fence_after();
      const uint32_t d_base = tm + rc.slot * p.acc_cols;
      for (int kci = 0; kci < nk; ++kci) {
        if (prof) t0 = clock64();
        mbar_wait(a_full + 8 * ra.slot, ra.phase);
        if (prof) t_a += clock64() - t0;
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + ra.slot * a_slot16;
        const uint32_t acc0 = kci != 0;
        if (WRES) {
          if (it == 0) {
            for (int t = 0; t < (S2D ? kS2dUnits : NTAPS); ++t) mbar_wait(w_full + 8 * (kci * NTAPS + t), 0);
            tc_fence_after();
          }
          if (elect_one()) {
            const uint32_t b_lo = b_lo0 + kci * NTAPS * w_blk16 + b_sp;
#pragma unroll
            for (int ti = 0; ti < NTAPS; ++ti) {
              // S2D issues the dense centre tap of a row first, so that the first MMA of a tile initialises all 64 columns
              const int t = S2D ? (ti % 3 == 0 ? ti + 1 : (ti % 3 == 1 ? ti - 1 : ti)) : ti;
              const int di = S2D ? t % 3 - 1 : 0;
#pragma unroll
              for (int j = 0; j < KSTEPS; ++j) {
                if (S2D && ((di < 0 && j < 2) || (di > 0 && j >= 2))) continue;   // structural zeros
                const uint32_t nsh = (S2D && di > 0) ? 32u : 0u;                     // output columns / B rows of this tap
                // B block of (tap, K step): dense blocks at a uniform stride, or the compact S2D units
                const uint32_t boff = S2D ? (uint32_t)((t / 3 * 4 + (di < 0 ? 0 : (di > 0 ? 3 : 1))) * (kS2dUnitBytes / 16)) +
                                                (uint32_t)(j - (di < 0 ? 2 : 0)) * b_kstep16
                                          : (uint32_t)(t * KSTEPS + j) * b_kstep16;
                const uint64_t bd = ((uint64_t)desc_hi << 32) | (b_lo + boff + nsh);
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                  const uint64_t ad = ((uint64_t)desc_hi << 32) | (a_lo + tap_off[t] + j * a_kstep16 + m * kTileM);
                  if (kStem)
                    umma_f16(d_base + (t >> 1) * p.n, ad, bd, idesc, (uint32_t)(t & 1));
                  else if (S2D)
                    umma_f16(d_base + m * p.n + nsh, ad, bd, di != 0 ? idesc_half : idesc, (ti | j) != 0 ? 1u : acc0);
                  else
                    umma_f16(d_base + m * p.n, ad, bd, idesc, (t | j) != 0 ? 1u : acc0);
                }
              }
            }
            umma_commit(a_empty + 8 * ra.slot);
          }
          __syncwarp();
        } else {
#pragma unroll
          for (int t = 0; t < NTAPS; ++t) {
            if (prof) t0 = clock64();
            mbar_wait(w_full + 8 * rw.slot, rw.phase);
            if (prof) t_w += clock64() - t0;
            tc_fence_after();
            if (elect_one()) {
              const uint32_t b_lo = b_lo0 + rw.slot * w_blk16 + b_sp;
#pragma unroll
              for (int j = 0; j < KSTEPS; ++j) {
                const uint64_t bd = ((uint64_t)desc_hi << 32) | (b_lo + j * b_kstep16);
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                  const uint64_t ad = ((uint64_t)desc_hi << 32) | (a_lo + tap_off[t] + j * a_kstep16 + m * kTileM);
                  umma_f16(d_base + m * p.n, ad, bd, idesc, (t | j) != 0 ? 1u : acc0);
                }
              }
              umma_commit(w_empty + 8 * rw.slot);
              if (t == NTAPS - 1) umma_commit(a_empty + 8 * ra.slot);
            }
            __syncwarp();
            rw.advance(p.ws);
          }
        }
        ra.advance(p.ra);
      }
      // folded shortcut: one tap at shift 0 over the chunks of the block input's phase image (0,0)
      for (int sc = 0; sc < p.sc_nk; ++sc) {
        mbar_wait(a_full + 8 * ra.slot, ra.phase);
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + ra.slot * a_slot16 + (uint32_t)(-p.min_shift);
        uint32_t b_lo;
        if (S2D) {
          if (it == 0) {
            mbar_wait(w_full + 8 * kS2dUnits, 0);
            mbar_wait(w_full + 8 * (kS2dUnits + 1), 0);
            tc_fence_after();
          }
          b_lo = b_lo0 + (uint32_t)(kS2dMainBytes / 16);
        } else if (WRES) {
          if (it == 0) {
            mbar_wait(w_full + 8 * (nk * NTAPS + sc), 0);
            tc_fence_after();
          }
          b_lo = b_lo0 + (nk * NTAPS + sc) * w_blk16 + b_sp;
        } else {
          mbar_wait(w_full + 8 * rw.slot, rw.phase);
          tc_fence_after();
          b_lo = b_lo0 + rw.slot * w_blk16 + b_sp;
        }
        if (elect_one()) {
          for (int j = 0; j < p.sc_ksteps; ++j) {
            const uint64_t bd = ((uint64_t)desc_hi << 32) | (b_lo + j * b_kstep16);
#pragma unroll
            for (int m = 0; m < MT; ++m) {
              const uint64_t ad = ((uint64_t)desc_hi << 32) | (a_lo + j * a_kstep16 + m * kTileM);
              umma_f16(d_base + m * p.n, ad, bd, idesc, 1u);
            }
          }
          if (!WRES) umma_commit(w_empty + 8 * rw.slot);
          umma_commit(a_empty + 8 * ra.slot);
        }
        __syncwarp();
        if (!WRES) rw.advance(p.ws);
        ra.advance(p.ra);
      }
      if (elect_one()) umma_commit(acc_full + 8 * rc.slot);
      __syncwarp();
      rc.advance(p.as);
    }
    if (p.prof != nullptr && lane == 0) {
      long long* o = p.prof + blockIdx.x * 16;
      o[0] = clock64() - t_start; o[1] = t_acc; o[2] = t_a; o[3] = t_w; o[4] = n_my;
    }
  } else if (warp == 2) {
    if (lane == 0) {  // ---------------- activation producer
      Ring ra;
      pdl_wait();  // the previous layer's output must be complete before the first load
      for (int it = 0; it < n_my; ++it) {
        int sp_unused;
        const long long g0 = (long long)item_tile(first_tile + it * (int)gridDim.x, sp_unused) * kTilePx + p.min_shift;
        for (int kci = 0; kci < nk; ++kci) {
          if (p.dbg & 8) mbar_wait_sleep(a_empty + 8 * ra.slot, ra.phase ^ 1u, 100);
          else mbar_wait(a_empty + 8 * ra.slot, ra.phase ^ 1u);
          if ((p.dbg & 2) && (it > 0 || kci >= p.ra)) { mbar_arrive(a_full + 8 * ra.slot); ra.advance(p.ra); continue; }
          mbar_expect_tx(a_full + 8 * ra.slot, p.a_slot_bytes);
          uint32_t dst = sbase + p.smem_a_off + ra.slot * p.a_slot_bytes;
          for (int s = 0; s < p.n_src; ++s) {
            const uint4* src = p.a_src[s] + ((long long)(kci * kPlanesPerChunk) * p.a_plane_px + g0);
            for (int pl = 0; pl < kPlanesPerChunk; ++pl, dst += p.a_plane_bytes, src += p.a_plane_px)
              bulk_g2s(dst, src, p.a_plane_bytes, a_full + 8 * ra.slot);
          }
          ra.advance(p.ra);
        }
        for (int sc = 0; sc < p.sc_nk; ++sc) {
          mbar_wait(a_empty + 8 * ra.slot, ra.phase ^ 1u);
          mbar_expect_tx(a_full + 8 * ra.slot, p.sc_a_bytes);
          uint32_t dst = sbase + p.smem_a_off + ra.slot * p.a_slot_bytes;
          const uint4* src = p.sc_src + ((long long)(sc * p.sc_planes) * p.sc_plane_px + g0);
          for (int pl = 0; pl < p.sc_planes; ++pl, dst += p.a_plane_bytes, src += p.sc_plane_px)
            bulk_g2s(dst, src, p.a_plane_bytes, a_full + 8 * ra.slot);
          ra.advance(p.ra);
        }
      }
    }
  } else if (warp >= 4) {  // ---------------- epilogue: two groups of 4 warps, alternating tiles
    const int grp = (warp - 4) >> 2;
    const int qd = warp & 3;  // TMEM lane quarter this warp may access
    Ring rc;
    if (grp == 1) rc.advance(p.as);
    long long e_wait = 0, e_start = clock64();
    const bool prof = p.prof != nullptr && (p.dbg & 16);
    const int ncg = (p.cout + 7) >> 3;
    constexpr bool has_res = MODE == 1 || MODE == 2;
    constexpr int out_mode = (MODE == 2 || MODE == 5) ? 1 : (MODE == 3 ? 2 : (MODE == 4 ? 3 : 0));
    const int n_slabs_full = (p.n + 31) >> 5;  // 32-column slabs (the last one may hold 16)
    // slabs of one work item in flat order f = m * n_slabs + sl (per item: see below)
    // Residual ring: 4 slots of one slab (4 channel groups x 16 B per lane).  Slot f & 3 holds flat
    // slab f; after slab f is consumed the slot is refilled with slab f + 4 of the same tile or,
    // past the end, with the matching slab of this group's NEXT tile, so every global load is
    // issued at least 4 slabs (usually one whole tile) before its use.  Needs S < 4 or S % 4 == 0,
    // which holds for every residual layer (n_slabs is 1, 2, 4 or 8).  Pads and the tail behind
    // the last image are readable (zeros / finite stale data) and masked by `valid` below.
    uint4 rr[4][4];
    uint32_t ovf = 0u;  // OR of overflow_probe over every fp16 word this thread stores (valid pixels only)
    pdl_wait();  // before the first residual load / output store
    auto res_load = [&](uint4 (&dst)[4], uint32_t tile_gbase, int f) {
      const int m = f >> p.slab_shift, sl = f - (m << p.slab_shift);
      const uint32_t g = tile_gbase + m * kTileM;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (sl * 4 + i < ncg) dst[i] = __ldg(p.res + (long long)(sl * 4 + i) * p.res_plane_px + g);
    };
    if (has_res && grp < n_my) {
      const uint32_t gb = (uint32_t)(first_tile + grp * (int)gridDim.x) * kTilePx + qd * 32 + lane;
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (k < MT * n_slabs_full) res_load(rr[k], gb, k);   // (residual layers are never channel-sliced)
    }
    for (int it = grp; it < n_my; it += 2) {
      int sp;
      const int tile = item_tile(first_tile + it * (int)gridDim.x, sp);
      // a channel slice: n_part channels starting at c_off, accumulated in TMEM columns [0, n_part) of each sub-tile
      const int c_off = sp < 0 ? 0 : sp * n_part;
      const int n_slabs = sp < 0 ? n_slabs_full : n_slabs_full >> p.split_shift;
      const int slab_shift = sp < 0 ? p.slab_shift : p.slab_shift - p.split_shift;
      const int S = MT * n_slabs;
      const uint32_t gbase = (uint32_t)tile * kTilePx + qd * 32 + lane;
      const bool has_next = it + 2 < n_my;
      const uint32_t gbase_next = gbase + 2u * gridDim.x * kTilePx;
      const long long e0 = prof ? clock64() : 0;
      mbar_wait(acc_full + 8 * rc.slot, rc.phase);
      if (prof) e_wait += clock64() - e0;
      tc_fence_after();
      if (!(p.dbg & 1)) {
        bool valid = false;
        uint32_t g = 0, taddr = 0;
        long long o1 = 0;
        uint4* out_ph = nullptr;
#pragma unroll 1
        for (int f0 = 0; f0 < S; f0 += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int f = f0 + k;
            if (f >= S) break;
            const int m = MT == 1 ? 0 : (MODE == 4 ? (f >= n_slabs ? 1 : 0) : f >> slab_shift);
            const int sl = MT == 1 ? f : (MODE == 4 ? f - m * n_slabs : f - (m << slab_shift));
            if (sl == 0) {  // new sub-tile: decode the pixel of this lane
              g = gbase + m * kTileM;
              const uint32_t img = fast_div(g, p.div_hw);
              const uint32_t rem = g - img * p.div_hw.d;
              const uint32_t y = fast_div(rem, p.div_wp), x = rem - y * p.div_wp.d;
              valid = (long long)g < p.P && y < (uint32_t)p.H && x < (uint32_t)p.W;
              taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + rc.slot * p.acc_cols + m * p.n;
              if (out_mode == 1 && S2D) {   // x counts pixel pairs: the column parity is the channel half (below)
                out_ph = p.out_ph[(y & 1) * 2];
                o1 = (long long)img * (p.Hp2 * p.Wp2) + (y >> 1) * p.Wp2 + x;
              } else if (out_mode == 1) {
                out_ph = p.out_ph[(y & 1) * 2 + (x & 1)];
                o1 = (long long)img * (p.Hp2 * p.Wp2) + (y >> 1) * p.Wp2 + (x >> 1);
              } else if (out_mode == 2) {
                o1 = ((long long)(img * p.H + y) * p.W + x) * p.cout;
              } else if (out_mode == 3) {
                o1 = ((long long)(img * p.H + y) * p.W + x) * p.f32_ld;
              }
            }
            const int c0 = sl << 5;
            const bool wide = c_off + c0 + 32 <= p.n;
            uint32_t v[32];
            if (wide) tmem_ld_x32(taddr + c0, v);
            else tmem_ld_x16(taddr + c0, *reinterpret_cast<uint32_t(*)[16]>(v));
            if (kStem) {  // 2x2 max-pool = elementwise max of the four phase accumulators
#pragma unroll 1
              for (int q = 1; q < 4; ++q) {
                uint32_t u[32];
                tmem_ld_wait();
                tmem_ld_x32(taddr + q * p.n + c0, u);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(fmaxf(__uint_as_float(v[i]), __uint_as_float(u[i])));
              }
            }
            tmem_ld_wait();
            uint4 pk[4];
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              const int cg = ((c_off + c0) >> 3) + h;
              float o[8];
              const float4 b0 = *reinterpret_cast<const float4*>(bias_s + (cg < ncg ? cg : 0) * 8);
              const float4 b1 = *reinterpret_cast<const float4*>(bias_s + (cg < ncg ? cg : 0) * 8 + 4);
              if (MODE == 4) {
                const float sc = p.out_scale;
                o[0] = fmaf(__uint_as_float(v[h * 8 + 0]), sc, b0.x); o[1] = fmaf(__uint_as_float(v[h * 8 + 1]), sc, b0.y);
                o[2] = fmaf(__uint_as_float(v[h * 8 + 2]), sc, b0.z); o[3] = fmaf(__uint_as_float(v[h * 8 + 3]), sc, b0.w);
                o[4] = fmaf(__uint_as_float(v[h * 8 + 4]), sc, b1.x); o[5] = fmaf(__uint_as_float(v[h * 8 + 5]), sc, b1.y);
                o[6] = fmaf(__uint_as_float(v[h * 8 + 6]), sc, b1.z); o[7] = fmaf(__uint_as_float(v[h * 8 + 7]), sc, b1.w);
                const bool live = valid && cg < ncg && (wide || h < 2);
                const bool full = cg * 8 + 8 <= p.cout;  // the last group of a 76-channel tensor holds 4 channels
                if (p.res_f32 != nullptr && live) {
                  const float4 r0 = __ldg(reinterpret_cast<const float4*>(p.res_f32 + o1 + cg * 8));
                  o[0] += r0.x; o[1] += r0.y; o[2] += r0.z; o[3] += r0.w;
                  if (full) {
                    const float4 r1 = __ldg(reinterpret_cast<const float4*>(p.res_f32 + o1 + cg * 8 + 4));
                    o[4] += r1.x; o[5] += r1.y; o[6] += r1.z; o[7] += r1.w;
                  }
                }
                if (p.relu) {
#pragma unroll
                  for (int i = 0; i < 8; ++i) o[i] = fmaxf(o[i], 0.f);
                }
                if (p.out_f32 != nullptr && live) {
                  float4* op = reinterpret_cast<float4*>(p.out_f32 + o1 + cg * 8);
                  op[0] = make_float4(o[0], o[1], o[2], o[3]);
                  if (full) op[1] = make_float4(o[4], o[5], o[6], o[7]);
                }
                if (p.out != nullptr && cg < p.split_planes && (wide || h < 2)) {
                  uint4 hi = make_uint4(0u, 0u, 0u, 0u), lo = hi;
                  if (valid) {  // x = hi + lo with 2 x 11 significant bits; pads stay zero
                    float a, b;
                    hi.x = pack2(o[0], o[1], 0); unpack2(hi.x, 0, a, b); lo.x = pack2(o[0] - a, o[1] - b, 0);
                    hi.y = pack2(o[2], o[3], 0); unpack2(hi.y, 0, a, b); lo.y = pack2(o[2] - a, o[3] - b, 0);
                    hi.z = pack2(o[4], o[5], 0); unpack2(hi.z, 0, a, b); lo.z = pack2(o[4] - a, o[5] - b, 0);
                    hi.w = pack2(o[6], o[7], 0); unpack2(hi.w, 0, a, b); lo.w = pack2(o[6] - a, o[7] - b, 0);
                    ovf |= overflow_probe(hi.x) | overflow_probe(hi.y) | overflow_probe(hi.z) | overflow_probe(hi.w);
                  }
                  uint4* op = p.out + (long long)cg * p.out_plane_px + g;
                  const long long part = (long long)p.split_planes * p.out_plane_px;
                  op[0] = hi; op[part] = lo; op[2 * part] = hi;
                }
                continue;
              }
              o[0] = __uint_as_float(v[h * 8 + 0]) + b0.x; o[1] = __uint_as_float(v[h * 8 + 1]) + b0.y;
              o[2] = __uint_as_float(v[h * 8 + 2]) + b0.z; o[3] = __uint_as_float(v[h * 8 + 3]) + b0.w;
              o[4] = __uint_as_float(v[h * 8 + 4]) + b1.x; o[5] = __uint_as_float(v[h * 8 + 5]) + b1.y;
              o[6] = __uint_as_float(v[h * 8 + 6]) + b1.z; o[7] = __uint_as_float(v[h * 8 + 7]) + b1.w;
              if (has_res) {
                float a, b;
                unpack2(rr[k][h].x, p.fmt, a, b); o[0] += a; o[1] += b;
                unpack2(rr[k][h].y, p.fmt, a, b); o[2] += a; o[3] += b;
                unpack2(rr[k][h].z, p.fmt, a, b); o[4] += a; o[5] += b;
                unpack2(rr[k][h].w, p.fmt, a, b); o[6] += a; o[7] += b;
              }
              if (p.relu) {
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaxf(o[i], 0.f);
              }
              if (out_mode == 2) {
                if (valid && cg < ncg && (wide || h < 2)) {
                  float4* op = reinterpret_cast<float4*>(p.out_f32 + o1 + cg * 8);
                  op[0] = make_float4(o[0], o[1], o[2], o[3]);
                  op[1] = make_float4(o[4], o[5], o[6], o[7]);
                }
              } else {
                pk[h] = make_uint4(0u, 0u, 0u, 0u);
                if (valid) {
                  pk[h].x = pack2(o[0], o[1], p.fmt);
                  pk[h].y = pack2(o[2], o[3], p.fmt);
                  pk[h].z = pack2(o[4], o[5], p.fmt);
                  pk[h].w = pack2(o[6], o[7], p.fmt);
                  if (p.fmt == 0)
                    ovf |= overflow_probe(pk[h].x) | overflow_probe(pk[h].y) | overflow_probe(pk[h].z) | overflow_probe(pk[h].w);
                }
              }
            }
            // refill this ring slot before the stores go out
            if (has_res) {
              if (f + 4 < S) res_load(rr[k], gbase, f + 4);
              else if (has_next) res_load(rr[k], gbase_next, S >= 4 ? f + 4 - S : f);
            }
            if (out_mode != 2 && out_mode != 3) {
#pragma unroll
              for (int h = 0; h < 4; ++h) {
                const int cg = ((c_off + c0) >> 3) + h;
                if (cg < ncg && (wide || h < 2)) {
                  if (out_mode == 0)
                    p.out[(long long)cg * p.out_plane_px + g] = pk[h];
                  else if (valid && S2D)   // planes 0-3 = even columns -> phase (y&1, 0), planes 4-7 -> phase (y&1, 1)
                    (out_ph + (cg >> 2) * p.out_ph_stride)[(long long)(cg & 3) * p.out_ph_plane_px + o1] = pk[h];
                  else if (valid)
                    out_ph[(long long)cg * p.out_ph_plane_px + o1] = pk[h];
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + 8 * rc.slot);
      rc.advance(p.as);
      rc.advance(p.as);
    }
    if (p.status != nullptr && (ovf & kOverflowProbeMask) != 0u) atomicOr(p.status, 2);
    if (p.prof != nullptr && lane == 0 && qd == 0) {
      long long* o = p.prof + blockIdx.x * 16 + 8 + grp * 2;
      o[0] = clock64() - e_start; o[1] = e_wait;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    __syncwarp();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
  if (p.prof != nullptr && threadIdx.x == 0) p.prof[blockIdx.x * 16 + 6] = (long long)globaltimer_ns();
}

#include "conv_tc2.cuh"

// ------------------------------------------------------------------------------------------
// geometry
struct Level {
  int H, W, Hp, Wp, C;
  long long n_img, P, guard, plane_px;
  long long planes() const { return C / 8; }
  long long buf_units() const { return planes() * plane_px; }  // 16-byte units
};

Level make_level(int H, int W, int C, long long n_img) {
  Level L;
  L.H = H; L.W = W; L.Hp = H + 1; L.Wp = W + 1; L.C = C; L.n_img = n_img;
  L.P = n_img * L.Hp * L.Wp;
  L.guard = round_up(L.Wp + 2, 8) + 8;
  L.plane_px = L.guard + round_up(L.P, 1024) + L.guard;
  return L;
}

// zero the front guard, everything behind the last image, and (optionally) the pad cells
__global__ void zero_halo_kernel(uint4* base, long long n_planes, Level L, int zero_pads) {
  pdl_prologue();
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  const long long tail = L.plane_px - L.guard - L.P;  // cells behind the images (incl. back guard)
  const long long pads_per_img = zero_pads ? (L.Hp + L.Wp - 1) : 0;
  const long long per_plane = L.guard + tail + pads_per_img * L.n_img;
  const long long total = per_plane * n_planes;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long pl = i / per_plane;
    long long k = i - pl * per_plane;
    uint4* plane = base + pl * L.plane_px;
    if (k < L.guard) {
      plane[k] = z;
    } else if ((k -= L.guard) < tail) {
      plane[L.guard + L.P + k] = z;
    } else {
      k -= tail;
      const long long img = k / pads_per_img;
      const int c = (int)(k - img * pads_per_img);
      // pad column x = W for y in [0,H), then pad row y = H for x in [0,Wp)
      const long long cell = c < L.H ? (long long)c * L.Wp + L.W : (long long)L.H * L.Wp + (c - L.H);
      plane[L.guard + img * L.Hp * L.Wp + cell] = z;
    }
  }
}

int zero_halo(uint4* base, long long n_planes, const Level& L, bool pads, cudaStream_t s) {
  const long long per_plane = L.guard + (L.plane_px - L.guard - L.P) + (pads ? (L.Hp + L.Wp - 1) * L.n_img : 0);
  const long long total = per_plane * n_planes;
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 8);
  launch_pdl(zero_halo_kernel, dim3(blocks), dim3(256), 0, s, base, n_planes, L, pads ? 1 : 0);
  return check_launch("zero_halo_kernel");
}

// ------------------------------------------------------------------------------------------
// packed 16-bit weights: per layer, blocks [kchunk][tap], each block [kc/8][n_pad][8]
struct TcLayer {
  int64_t offset;  // bytes into the tc blob
  int n_pad, kc, nk;
  int64_t s2d_offset;  // layer 1 only: the same conv as 64 -> 64 blocks over the x-space-to-depth view (-1: none)
};
bool is_s2d_layer_host(int i) { return i >= L_LAYER1 && i < L_LAYER1 + 4; }
TcLayer g_tc[L_COUNT];
int64_t g_tc_bytes = 0;
std::once_flag g_tc_once;

// conv2 of the first block of layers 2-4: the 1x1 stride-2 shortcut (the next table row) can be folded in
bool is_fold_host(int i) { return i == L_LAYER2 + 1 || i == L_LAYER3 + 1 || i == L_LAYER4 + 1; }

// conv2 of the other (identity-shortcut) blocks: the residual x can be accumulated as a 10th tap
// with identity weights (extra K chunks at shift 0 over the block input)
bool is_identity_host(int i) {
  return i == L_LAYER1 + 1 || i == L_LAYER1 + 3 || i == L_LAYER2 + 4 || i == L_LAYER2 + 6 || i == L_LAYER3 + 4 ||
         i == L_LAYER3 + 6 || i == L_LAYER3 + 8 || i == L_LAYER3 + 10 || i == L_LAYER4 + 4;
}

void build_tc_table() {
  const ConvLayer* T = layer_table();
  int64_t off = 0;
  for (int i = 0; i < L_COUNT; ++i) {
    g_tc[i].offset = -1;
    if (i > L_PROJ) continue;
    if (i == L_STEM) {  // stem GEMM: 8 blocks (4 pooling phases x 2 rows) of [2 pixels][32][8]
      g_tc[i].n_pad = 32; g_tc[i].kc = 16; g_tc[i].nk = 1;
      g_tc[i].offset = off;
      off += 8 * 1024;          // GEMM stem blocks (tcgen05 variant)
      off += 16 * 32 * 8;       // mma.sync B fragments (stem_mma_kernel)
      continue;
    }
    g_tc[i].n_pad = (int)round_up(T[i].cout, 16);
    g_tc[i].kc = T[i].cin < 64 ? T[i].cin : 64;
    g_tc[i].nk = T[i].cin / g_tc[i].kc;
    g_tc[i].offset = off;
    off += (int64_t)T[i].ksize * T[i].ksize * T[i].cin * g_tc[i].n_pad * 2;
    if (is_fold_host(i)) off += (int64_t)T[i + 1].cin * g_tc[i].n_pad * 2;  // copy of the shortcut's blocks (layer i + 1)
    if (is_identity_host(i)) off += (int64_t)T[i].cout * g_tc[i].n_pad * 2;    // identity blocks
    off = round_up(off, 128);
    g_tc[i].s2d_offset = -1;
    if (is_s2d_layer_host(i)) {
      g_tc[i].s2d_offset = off;
      off += kS2dMainBytes + (is_identity_host(i) ? kS2dIdBytes : 0);
      off = round_up(off, 128);
    }
  }
  g_tc_bytes = off + 128;
}
const TcLayer* tc_table() {
  std::call_once(g_tc_once, build_tc_table);
  return g_tc;
}

__global__ void prepack_identity_kernel(uint16_t* __restrict__ out, int c, int n_pad, int kc, int fmt) {
  const long long total = (long long)c * n_pad;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int e = (int)(i & 7);
  long long r = i >> 3;
  const int nn = (int)(r % n_pad); r /= n_pad;
  const int kk = (int)(r % (kc / 8));
  const int kci = (int)(r / (kc / 8));
  out[i] = ptx::pack1(kci * kc + kk * 8 + e == nn ? 1.0f : 0.0f, fmt);
}

// Layer-1 weights [9][32][coutP] -> nine [8][64][8] blocks of the x-space-to-depth view: input channel
// ci' = pi * 32 + ci (pi = parity of the input column), output nn = po * 32 + co, tap (dy, di) carries
// W[dy][dx] with dx = 2 di + pi - po when that is a tap of the 3x3 kernel and zero otherwise.
__global__ void prepack_s2d_kernel(const float* __restrict__ w, uint16_t* __restrict__ out, int coutP, int fmt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int)(kS2dMainBytes / 2)) return;
  const int unit = i / 2048;                       // 2048 elements per 4 KB unit
  const int row3 = unit / 4, u3 = unit % 4;        // a kernel row (dy) is 4 units: di = -1 | di = 0 (two) | di = +1
  const int t = row3 * 3 + (u3 == 0 ? 0 : (u3 == 3 ? 2 : 1));
  const int loc = i - (row3 * 4 + (u3 == 0 ? 0 : (u3 == 3 ? 3 : 1))) * 2048;   // element inside the tap's block
  const int e = loc & 7, nn = (loc >> 3) % kS2dC;
  const int kk = loc / (kS2dC * 8) + (t % 3 == 0 ? 4 : 0);   // di = -1 stores K groups 4-7
  const int cip = kk * 8 + e, pi = cip >> 5, ci = cip & 31, po = nn >> 5, co = nn & 31;
  const int dy = t / 3, dx = 2 * (t % 3 - 1) + pi - po;
  const float v = (dx >= -1 && dx <= 1) ? w[((long long)(dy * 3 + dx + 1) * 32 + ci) * coutP + co] : 0.f;
  out[i] = ptx::pack1(v, fmt);
}

__global__ void prepack_kernel(const float* __restrict__ w /*[taps][cin][coutP]*/, uint16_t* __restrict__ out, int taps,
                               int cin, int cout, int coutP, int n_pad, int kc, int fmt) {
  const long long total = (long long)taps * cin * n_pad;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  // i -> (kci, t, kk, nn, e)
  const int e = (int)(i & 7);
  long long r = i >> 3;
  const int nn = (int)(r % n_pad); r /= n_pad;
  const int kk = (int)(r % (kc / 8)); r /= (kc / 8);
  const int t = (int)(r % taps);
  const int kci = (int)(r / taps);
  const int ci = kci * kc + kk * 8 + e;
  const float v = nn < cout ? w[((long long)t * cin + ci) * coutP + nn] : 0.f;
  out[i] = ptx::pack1(v, fmt);
}

// ------------------------------------------------------------------------------------------
// Stem as a tensor-core GEMM (conv3x3(1->32)+BN+ReLU+maxpool2, model_utils.py:119-124).
// Input plane S: cell g of the level-1 geometry holds the 2x2 block of crop pixels under pooled
// pixel g as channels (pa*2+pb) of one 16-byte unit (4 values + 4 zeros).  The conv output at
// pooling phase (oy,ox) is then a 2x2-shift GEMM over S; one MMA covers two horizontally
// neighbouring cells (its two K halves, LBO = 16 B).  The epilogue takes the max of the four
// phase accumulators (= the max-pool), adds the bias and applies ReLU.
__global__ void prepack_stem_kernel(const float* __restrict__ w /*[9][32]*/, uint16_t* __restrict__ out, int fmt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 8 * 512) return;
  const int e = i & 7, nn = (i >> 3) & 31, h = (i >> 8) & 1, t = i >> 9;
  const int op = t >> 1, q = t & 1, oy = op >> 1, ox = op & 1;
  const int dy = (oy == 0 ? -1 : 0) + q, dx = (ox == 0 ? -1 : 0) + h;
  const int pa = e >> 1, pb = e & 1;
  const int r = 2 * dy + pa - oy + 1, c = 2 * dx + pb - ox + 1;
  float v = 0.f;
  if (e < 4 && r >= 0 && r <= 2 && c >= 0 && c <= 2) v = w[(r * 3 + c) * 32 + nn];
  out[i] = ptx::pack1(v, fmt);
}

template <typename TIn>
__global__ void __launch_bounds__(256) stem_pack_kernel(const TIn* __restrict__ crops, uint4* __restrict__ out /*px0*/, Level L,
                                                        int fmt) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= L.P) return;
  const int HWp = L.Hp * L.Wp;
  const int img = (int)(g / HWp);
  const int rem = (int)(g - (long long)img * HWp);
  const int py = rem / L.Wp, px = rem - py * L.Wp;
  uint4 pk = make_uint4(0u, 0u, 0u, 0u);
  if (py < L.H && px < L.W) {
    const int W = 2 * L.W;
    const TIn* ip = crops + (size_t)img * (2 * L.H) * W + (2 * py) * W + 2 * px;
    float v00, v01, v10, v11;
    if (sizeof(TIn) == 1) {
      const uchar2 a = __ldg(reinterpret_cast<const uchar2*>(ip)), b = __ldg(reinterpret_cast<const uchar2*>(ip + W));
      v00 = __fdiv_rn((float)a.x, 255.0f); v01 = __fdiv_rn((float)a.y, 255.0f);
      v10 = __fdiv_rn((float)b.x, 255.0f); v11 = __fdiv_rn((float)b.y, 255.0f);
    } else {
      const float2 a = __ldg(reinterpret_cast<const float2*>(ip)), b = __ldg(reinterpret_cast<const float2*>(ip + W));
      v00 = a.x; v01 = a.y; v10 = b.x; v11 = b.y;
    }
    pk.x = ptx::pack2(v00, v01, fmt);
    pk.y = ptx::pack2(v10, v11, fmt);
  }
  out[g] = pk;
}

// stem: conv3x3(1->32)+BN+ReLU+maxpool2 (model_utils.py:119-124), fp32 math, writes the level-1
// planar layout (pads as zeros).  64 consecutive padded cells x 4 channel groups per block
// iteration; the 72 weights of a thread's channel group stay in registers across iterations.
constexpr int kStemIters = 8;
template <typename TIn>
__global__ void __launch_bounds__(256) stem_tc_kernel(const TIn* __restrict__ crops, const float* __restrict__ w,
                                                      const float* __restrict__ bias, uint4* __restrict__ out /*plane0,px0*/,
                                                      Level L, int fmt) {
  const int cg = threadIdx.x >> 6;
  float wr[9][8], br[8];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int ch = 0; ch < 8; ++ch) wr[t][ch] = __ldg(w + t * 32 + cg * 8 + ch);
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) br[ch] = __ldg(bias + cg * 8 + ch);
  const int HWp = L.Hp * L.Wp;
  const int H = 2 * L.H, W = 2 * L.W;
  for (int k = 0; k < kStemIters; ++k) {
    const long long g = ((long long)blockIdx.x * kStemIters + k) * 64 + (threadIdx.x & 63);
    if (g >= L.P) return;
    const int img = (int)(g / HWp);
    const int rem = (int)(g - (long long)img * HWp);
    const int py = rem / L.Wp, px = rem - py * L.Wp;
    uint4 pk = make_uint4(0u, 0u, 0u, 0u);
    if (py < L.H && px < L.W) {
      const TIn* ip = crops + (size_t)img * H * W;
      float patch[4][4];
#pragma unroll
      for (int dy = 0; dy < 4; ++dy)
#pragma unroll
        for (int dx = 0; dx < 4; ++dx) {
          const int y = 2 * py - 1 + dy, x = 2 * px - 1 + dx;
          float v = 0.f;
          if ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) {
            if (sizeof(TIn) == 1)
              v = __fdiv_rn((float)__ldg(ip + y * W + x), 255.0f);
            else
              v = (float)__ldg(ip + y * W + x);
          }
          patch[dy][dx] = v;
        }
      float res[8];
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        float best = 0.f;  // ReLU then max == max(0, max of the four)
#pragma unroll
        for (int oy = 0; oy < 2; ++oy)
#pragma unroll
          for (int ox = 0; ox < 2; ++ox) {
            float acc = br[ch];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int q = 0; q < 3; ++q) acc = fmaf(patch[oy + r][ox + q], wr[r * 3 + q][ch], acc);
            best = fmaxf(best, acc);
          }
        res[ch] = best;
      }
      pk.x = ptx::pack2(res[0], res[1], fmt);
      pk.y = ptx::pack2(res[2], res[3], fmt);
      pk.z = ptx::pack2(res[4], res[5], fmt);
      pk.w = ptx::pack2(res[6], res[7], fmt);
    }
    out[(long long)cg * L.plane_px + g] = pk;
  }
}

#include "stem_mma.cuh"

// ------------------------------------------------------------------------------------------
// layout converters (debug entry point + tests)
// fp32 NHWC [n,H,W,C] -> planar 16-bit.  phase_split: H,W are the FULL-resolution sizes and the
// four (H/2 x W/2) phase images are written to base + ph * phase_units.
__global__ void pack_planar_kernel(const float* __restrict__ in, uint4* __restrict__ base /*plane0,px0*/, Level L,
                                   int phase_split, long long phase_units, int fullH, int fullW, int fmt) {
  const long long total = L.n_img * fullH * fullW * (L.C / 8);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int cg = (int)(i % (L.C / 8));
  long long r = i / (L.C / 8);
  const int x = (int)(r % fullW); r /= fullW;
  const int y = (int)(r % fullH);
  const long long img = r / fullH;
  const float* ip = in + ((img * fullH + y) * fullW + x) * L.C + cg * 8;
  uint4 pk;
  pk.x = ptx::pack2(ip[0], ip[1], fmt);
  pk.y = ptx::pack2(ip[2], ip[3], fmt);
  pk.z = ptx::pack2(ip[4], ip[5], fmt);
  pk.w = ptx::pack2(ip[6], ip[7], fmt);
  if (phase_split) {
    const int ph = (y & 1) * 2 + (x & 1);
    base[ph * phase_units + (long long)cg * L.plane_px + img * L.Hp * L.Wp + (y >> 1) * L.Wp + (x >> 1)] = pk;
  } else {
    base[(long long)cg * L.plane_px + img * L.Hp * L.Wp + y * L.Wp + x] = pk;
  }
}

__global__ void unpack_planar_kernel(const uint4* __restrict__ base, float* __restrict__ out, Level L, int cout, int fmt) {
  const long long total = L.n_img * L.H * L.W * (L.C / 8);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int cg = (int)(i % (L.C / 8));
  long long r = i / (L.C / 8);
  const int x = (int)(r % L.W); r /= L.W;
  const int y = (int)(r % L.H);
  const long long img = r / L.H;
  const uint4 pk = base[(long long)cg * L.plane_px + img * L.Hp * L.Wp + y * L.Wp + x];
  float o[8];
  ptx::unpack2(pk.x, fmt, o[0], o[1]);
  ptx::unpack2(pk.y, fmt, o[2], o[3]);
  ptx::unpack2(pk.z, fmt, o[4], o[5]);
  ptx::unpack2(pk.w, fmt, o[6], o[7]);
  float* op = out + ((img * L.H + y) * L.W + x) * cout + cg * 8;
  for (int k = 0; k < 8; ++k)
    if (cg * 8 + k < cout) op[k] = o[k];
}

// ------------------------------------------------------------------------------------------
// launcher
struct ConvIo {
  const uint4* a_base;       // plane 0 / pixel 0 of source 0 (phase 0)
  long long a_plane_px;
  long long a_phase_units;   // distance between phase images (stride-2 inputs)
  const uint4* res;          // plane 0 / pixel 0, or null
  long long res_plane_px;
  uint4* out;                // mode 0 / mode 1 (phase 0) plane 0 / pixel 0
  long long out_plane_px;
  long long out_phase_units;
  float* out_f32;
  int out_mode;              // 0 planar, 1 phase-split, 2 fp32 NHWC, 3 tail "fp16 x 3" (kernel MODE 4)
  bool relu;
  const float* res_f32;      // out_mode 3
  float out_scale;
  int split_planes, f32_ld;
  int sc_layer;              // > 0: fold this 1x1 stride-2 shortcut layer into the accumulator ...
  const uint4* sc_src;       // ... reading phase image (0,0) of the block input (plane 0 / pixel 0)
  long long sc_plane_px;
  bool s2d;                  // layer 1 over its x-space-to-depth view (the Level passed is the 48 x 24 x 64 one)
};

// A convolution that is not a row of the backbone layer table (the pose-head convs of the tail).
struct TcConvDesc {
  const char* name;
  int ksize, stride, cout, n_pad, kc, nk;
  const void* w;       // packed blocks [kchunk][tap][kc/8][n_pad][8]
  const float* bias;
};

int g_mt_override = 0;

// q = n / d as umulhi(n, magic) >> shift; exact for n < 2^23 and the divisors of this kernel
Divider make_divider(uint32_t d) {
  Divider dv;
  const int s = d >= 256 ? 40 : 32;
  const unsigned long long m = ((1ull << s) + d - 1) / d;
  dv.magic = (uint32_t)m;
  dv.shift = (uint32_t)(s - 32);
  dv.d = d;
  return dv;
}
long long* g_prof = nullptr;
int g_prof_seq = -1;  // >= 0: launch k writes its counters at g_prof + k * 148 * 16
thread_local int32_t* g_status = nullptr;  // status word of the encoder call in progress on this thread

// "The encoder is under way" event: recorded on the encoder's stream right after layer1.0.conv1 has been launched, and
// `release_stream` is made to wait for it.  Side-stream work released at that point (the source-box pull of the NEXT step)
// meets an SM array that is full of conv CTAs and spreads over their leftover registers, instead of piling up on
// the SMs that happen to be free at a kernel boundary and pushing a CTA of the 148-CTA persistent conv grids into a second wave.
// `which`: one event per device and purpose (0 = release point, 1 = front stream -> encoder stream)
int order_streams(cudaStream_t from, cudaStream_t to, int which) {
  static std::mutex mu;
  static cudaEvent_t ev[2][64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  cudaEvent_t e;
  {
    std::lock_guard<std::mutex> lock(mu);
    cudaEvent_t& slot = ev[which & 1][dev & 63];
    if (slot == nullptr && cudaEventCreateWithFlags(&slot, cudaEventDisableTiming) != cudaSuccess) {
      set_error("at_encoder_forward: cudaEventCreate: %s", cudaGetErrorString(cudaGetLastError()));
      return AT_ERR_CUDA;
    }
    e = slot;
  }
  if (cudaEventRecord(e, from) != cudaSuccess || cudaStreamWaitEvent(to, e, 0) != cudaSuccess) {
    set_error("at_encoder_forward: stream ordering event: %s", cudaGetErrorString(cudaGetLastError()));
    return AT_ERR_CUDA;
  }
  return AT_OK;
}
int release_after_launch(cudaStream_t s, cudaStream_t release_stream) { return order_streams(s, release_stream, 0); }


// conv_tc2_kernel (CTA pairs): ring / residency configuration and launch.  `p` arrives with the layer, tap, source
// and shortcut fields filled by launch_conv_tc.
constexpr int kSmemBudget2 = 226 * 1024;

int launch_conv_tc2(ConvTcParams p, const char* name, const Level& OL, int out_mode, int mt, int max_shift, cudaStream_t stream) {
  const int taps = 9;
  const int nblocks = p.n_kchunks * taps + p.sc_nk;
  p.w_src_bytes = (uint32_t)(p.kc * p.n * 2);
  p.w_kg_bytes = (uint32_t)(p.n * 16);
  p.w_block_bytes = p.w_src_bytes / 2;   // per CTA: half of the output channels
  p.sc_w_bytes /= 2;
  const int64_t misc = 3072;             // bias + barriers + alignment
  auto a_slot = [&](int m) { return (int64_t)p.n_src * (p.kc / 8) * round_up(kTileM * m + max_shift - p.min_shift, 8) * 16; };
  // weights resident when at least 2 activation slots still fit (smaller tiles first give up their second sub-tile)
  // experiment knobs: AT_TC2_RES=0 never keeps weights resident, AT_TC2_WS=<n> ring depth of streamed weights,
  // AT_TC2_MINRA=<n> activation slots a resident configuration must leave (default 2)
  static const int knob_res = getenv("AT_TC2_RES") ? atoi(getenv("AT_TC2_RES")) : 1;
  static const int knob_ws = getenv("AT_TC2_WS") ? atoi(getenv("AT_TC2_WS")) : 4;
  static const int knob_minra = getenv("AT_TC2_MINRA") ? atoi(getenv("AT_TC2_MINRA")) : 2;
  int res_mt = 0;
  for (int m = mt; m >= 1 && knob_res; m >>= 1)
    if ((int64_t)nblocks * p.w_block_bytes + knob_minra * a_slot(m) + misc <= kSmemBudget2) { res_mt = m; break; }
  int64_t ra;
  if (res_mt > 0) {
    mt = res_mt; p.w_resident = 1; p.ws = nblocks;
    ra = (kSmemBudget2 - misc - (int64_t)nblocks * p.w_block_bytes) / a_slot(mt);
  } else {
    p.w_resident = 0; p.ws = knob_ws;
    for (;; mt >>= 1) {
      ra = (kSmemBudget2 - misc - (int64_t)p.ws * p.w_block_bytes) / a_slot(mt);
      if (ra >= 3 || mt == 1) break;
    }
    while (ra < 2 && p.ws > 2) { --p.ws; ra = (kSmemBudget2 - misc - (int64_t)p.ws * p.w_block_bytes) / a_slot(mt); }
  }
  if (ra < 2) {
    set_error("conv_tc2: layer %s does not fit shared memory", name);
    return AT_ERR_UNSUPPORTED;
  }
  if (ra > 8) ra = 8;
  p.mt = mt; p.ra = (int)ra;
  p.npix = (int)round_up(kTileM * mt + max_shift - p.min_shift, 8);
  p.a_plane_bytes = (uint32_t)p.npix * 16u;
  p.a_slot_bytes = (uint32_t)a_slot(mt);
  p.sc_a_bytes = (uint32_t)p.sc_planes * p.a_plane_bytes;
  const int64_t w_bytes = (int64_t)p.ws * p.w_block_bytes;
  p.acc_cols = (uint32_t)round_up(mt * p.n, 32);
  p.as = (int)(512 / p.acc_cols);
  if (p.as > 4) p.as = 4;
  uint32_t cols = 32;
  while (cols < p.as * p.acc_cols) cols <<= 1;
  p.tmem_cols = cols;
  p.smem_a_off = 0;
  p.smem_w_off = (uint32_t)round_up((int64_t)p.ra * p.a_slot_bytes, 128);
  p.smem_bias_off = (uint32_t)round_up(p.smem_w_off + w_bytes, 128);
  p.smem_bar_off = (uint32_t)round_up(p.smem_bias_off + p.n * 4, 128);
  const uint32_t smem_bytes = p.smem_bar_off + 8 * (3 * p.ra + 3 * p.ws + 2 * p.as) + 16;
  p.n_tiles = (OL.P + kTileM * mt - 1) / (kTileM * mt);
  const long long n_pair_tiles = (p.n_tiles + 1) / 2;
  const int grid = 2 * (int)std::min<long long>(n_pair_tiles, kNumSMs / 2);
  if (grid == 0) return AT_OK;
  p.div_hw = make_divider((uint32_t)(OL.Hp * OL.Wp));
  p.div_wp = make_divider((uint32_t)OL.Wp);
  const int ksteps = p.kc >> 4;
  const int mode = out_mode == 1 ? 5 : 0;
  using Kernel = void (*)(const ConvTcParams);
  Kernel fn = nullptr;
#define AT_TC2_CASE(K, M, R, O) \
  if (ksteps == K && mt == M && p.w_resident == R && mode == O) fn = conv_tc2_kernel<K, M, R, O>;
#define AT_TC2_CASES(K, M) AT_TC2_CASE(K, M, 0, 0) AT_TC2_CASE(K, M, 0, 5) AT_TC2_CASE(K, M, 1, 0) AT_TC2_CASE(K, M, 1, 5)
  AT_TC2_CASES(2, 1) AT_TC2_CASES(2, 2) AT_TC2_CASES(4, 1) AT_TC2_CASES(4, 2)
#undef AT_TC2_CASES
#undef AT_TC2_CASE
  if (fn == nullptr) {
    set_error("conv_tc2: no kernel for kc=%d mt=%d resident=%d mode=%d", p.kc, mt, p.w_resident, mode);
    return AT_ERR_UNSUPPORTED;
  }
  {
    static std::mutex mu;
    static std::vector<std::pair<int, Kernel>> configured;
    int dev = 0;
    cudaGetDevice(&dev);
    const std::pair<int, Kernel> key(dev, fn);
    std::lock_guard<std::mutex> lock(mu);
    if (std::find(configured.begin(), configured.end(), key) == configured.end()) {
      cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget2 + 1024);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      if (e != cudaSuccess) {
        set_error("cudaFuncSetAttribute(conv_tc2_kernel): %s", cudaGetErrorString(e));
        return AT_ERR_CUDA;
      }
      configured.push_back(key);
    }
  }
  static const bool use_pdl = getenv("AT_TC_PDL") == nullptr || atoi(getenv("AT_TC_PDL")) != 0;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = use_pdl ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, fn, p);
  if (e != cudaSuccess) {
    count_launch();
    set_error("conv_tc2_kernel launch (%s, smem %u): %s", name, smem_bytes, cudaGetErrorString(e));
    return AT_ERR_CUDA;
  }
  return check_launch("conv_tc2_kernel");
}

// `out_level`: geometry of the OUTPUT (= geometry of the input for stride 1, = geometry of each
// input phase image for stride 2).  `next`: geometry of the phase images written in mode 1.
int launch_conv_tc(int layer_id, const float* blob, const void* tc_blob, const Level& OL, const Level* next,
                   const ConvIo& io, int fmt, cudaStream_t stream, const TcConvDesc* desc = nullptr) {
  ConvLayer L;
  TcLayer T;
  if (desc != nullptr) {
    L.name = desc->name; L.cin = desc->kc * desc->nk; L.cout = desc->cout; L.ksize = desc->ksize; L.stride = desc->stride;
    L.weight_offset = 0; L.bias_offset = 0; L.bf16_offset = -1;
    T.offset = 0; T.n_pad = desc->n_pad; T.kc = desc->kc; T.nk = desc->nk;
  } else {
    L = layer_table()[layer_id];
    T = tc_table()[layer_id];
  }
  if (io.s2d) {
    if (desc != nullptr || T.s2d_offset < 0 || L.ksize != 3 || L.stride != 1 || L.cin != 32 || L.cout != 32 || io.res != nullptr ||
        io.sc_layer > 0 || (io.out_mode != 0 && io.out_mode != 1)) {
      set_error("conv_tc: layer %s has no space-to-depth form", L.name);
      return AT_ERR_UNSUPPORTED;
    }
    L.cin = L.cout = kS2dC;
    T.offset = T.s2d_offset; T.n_pad = kS2dC; T.kc = kS2dC; T.nk = 1;
  }
  ConvTcParams p;
  memset(&p, 0, sizeof(p));
  p.bias_wrap = io.s2d ? 32 : 0;
  const bool is_stem = desc == nullptr && layer_id == L_STEM;
  const int taps = is_stem ? 8 : L.ksize * L.ksize;
  p.n_taps = taps;
  p.kc = T.kc; p.n_kchunks = T.nk; p.n = T.n_pad; p.cout = L.cout;
  p.H = OL.H; p.W = OL.W; p.Hp = OL.Hp; p.Wp = OL.Wp; p.n_img = OL.n_img; p.P = OL.P;
  int max_shift;
  if (is_stem) {
    p.n_src = 1;
    for (int t = 0; t < 8; ++t) {
      const int op = t >> 1, q = t & 1, oy = op >> 1, ox = op & 1;
      p.tap_src[t] = 0;
      p.tap_shift[t] = ((oy == 0 ? -1 : 0) + q) * OL.Wp + (ox == 0 ? -1 : 0);
    }
    p.min_shift = -(OL.Wp + 1);
    max_shift = OL.Wp + 1;
  } else if (L.stride == 1) {
    p.n_src = 1;
    for (int t = 0; t < taps; ++t) {
      p.tap_src[t] = 0;
      p.tap_shift[t] = taps == 9 ? (t / 3 - 1) * OL.Wp + (t % 3 - 1) : 0;
    }
    p.min_shift = taps == 9 ? -(OL.Wp + 1) : 0;
    max_shift = taps == 9 ? OL.Wp + 1 : 0;
  } else {
    // tap (r,s) reads input (2y+r-1, 2x+s-1): phase (r!=1, s!=1), shifted by -1 row/col when r/s == 0
    p.n_src = taps == 9 ? 4 : 1;
    for (int t = 0; t < taps; ++t) {
      const int r = taps == 9 ? t / 3 : 1, s = taps == 9 ? t % 3 : 1;
      p.tap_src[t] = (r != 1) * 2 + (s != 1);
      p.tap_shift[t] = (r == 0 ? -OL.Wp : 0) + (s == 0 ? -1 : 0);
    }
    p.min_shift = taps == 9 ? -(OL.Wp + 1) : 0;
    max_shift = 0;
  }
  for (int s = 0; s < p.n_src; ++s) p.a_src[s] = io.a_base + s * io.a_phase_units;
  p.a_plane_px = io.a_plane_px;
  p.w = desc != nullptr ? reinterpret_cast<const uint4*>(desc->w)
                        : reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(tc_blob) + T.offset);
  p.bias = desc != nullptr ? desc->bias : blob + L.bias_offset;
  p.res_f32 = io.res_f32; p.out_scale = io.out_scale; p.split_planes = io.split_planes; p.f32_ld = io.f32_ld;
  p.res = io.res; p.res_plane_px = io.res_plane_px;
  p.out = io.out; p.out_plane_px = io.out_plane_px;
  p.out_mode = io.out_mode; p.relu = io.relu ? 1 : 0; p.fmt = fmt;
  p.out_f32 = io.out_f32;
  p.prof = g_prof;
  p.status = g_status;
  { const char* e = getenv("AT_TC_DBG"); p.dbg = e ? atoi(e) : 0; }
  if (g_prof != nullptr && g_prof_seq >= 0) p.prof = g_prof + (long long)(g_prof_seq++) * kNumSMs * 16;
  if (io.out_mode == 1) {
    for (int s = 0; s < 4; ++s) p.out_ph[s] = io.out + s * io.out_phase_units;
    p.out_ph_stride = io.out_phase_units;
    p.out_ph_plane_px = io.out_plane_px;
    p.Hp2 = next->Hp; p.Wp2 = next->Wp;
  }
  // tile / ring configuration
  // sub-tiles (128 pixels each) per tile: measured best per shape on B200 (scripts/tc_prof.py)
  int mt = p.n == 64 ? (io.res != nullptr ? 1 : 2) : (p.n <= 128 ? 2 : 1);
  if (io.out_mode == 3) mt = 1;  // pose head: 2 tiles per CTA overlap epilogue and MMAs (measured: 0.31 vs 0.34 ms tail)
  {
    // experiment knobs: AT_TC_MT<N> = sub-tiles per tile for layers with padded Cout == N
    char name[32];
    snprintf(name, sizeof(name), "AT_TC_MT%d", p.n);
    const char* e = getenv(name);
    if (e) mt = atoi(e);
  }
  if (g_mt_override > 0) mt = g_mt_override;
  if (p.n_src > 1 || is_stem) mt = 1;  // the 4-phase stride-2 inputs already fill the A ring
  if (mt > 2) mt = 2;
  if (taps == 1) mt = 1;
  while (mt > 1 && round_up(mt * p.n, 32) > 256) mt >>= 1;  // keep >= 2 accumulator stages
  if (io.sc_layer == -1) {  // identity shortcut: x accumulated through identity weight blocks
    if (desc != nullptr || !is_identity_host(layer_id) || L.cin != L.cout) {  // (S2D: the 64 x 64 identity follows the nine blocks)
      set_error("conv_tc: layer %s has no identity blocks", L.name);
      return AT_ERR_UNSUPPORTED;
    }
    p.sc_nk = L.cout / p.kc; p.sc_planes = p.kc / 8; p.sc_ksteps = p.kc / 16;
    p.sc_src = io.sc_src; p.sc_plane_px = io.sc_plane_px;
    p.sc_w_bytes = (uint32_t)(p.kc * p.n * 2);
  }
  if (io.sc_layer > 0) {
    const ConvLayer& S = layer_table()[io.sc_layer];
    const int kc_sc = S.cin < 64 ? S.cin : 64;
    if (desc != nullptr || S.ksize != 1 || S.cout != L.cout || kc_sc > p.kc || (S.cin / kc_sc > 1 && kc_sc != p.kc)) {
      set_error("conv_tc: cannot fold shortcut layer %s into %s", S.name, L.name);
      return AT_ERR_UNSUPPORTED;
    }
    p.sc_nk = S.cin / kc_sc; p.sc_planes = kc_sc / 8; p.sc_ksteps = kc_sc / 16;
    p.sc_src = io.sc_src; p.sc_plane_px = io.sc_plane_px;
    p.bias2 = blob + S.bias_offset;
    p.sc_w_bytes = (uint32_t)(kc_sc * p.n * 2);
  }
  {
    // CTA pairs (cta_group::2) for the 3x3 convs with >= 64 output channels: opt-in through AT_TC_CG2=<mask> (A/B)
    // bit mask by padded Cout: 1 -> 64 (layer 2), 2 -> 128 (layer 3), 4 -> 256 (layer 4)
    // Default 0: measured on B200 (profiles/r2_cg2_experiments.log) the pair kernel is correct (all conv / encoder / parity
    // tests pass with mask 7) but not faster than conv_tc_kernel in any ring configuration (2.49 vs 2.44 ms per 2048 crops
    // at best, layer 3 with resident weights): the relay / multicast-commit turnaround costs more than the halved B reads save.
    static const int cg2_mask = getenv("AT_TC_CG2") == nullptr ? 0 : atoi(getenv("AT_TC_CG2"));
    const bool use_cg2 = (cg2_mask & (p.n == 64 ? 1 : p.n == 128 ? 2 : 4)) != 0;
    if (use_cg2 && desc == nullptr && !is_stem && !io.s2d && taps == 9 && p.n >= 64 && p.n % 32 == 0 && io.res == nullptr &&
        (io.out_mode == 0 || io.out_mode == 1) && p.cout == p.n)
      return launch_conv_tc2(p, L.name, OL, io.out_mode, mt, max_shift, stream);
  }
  // resident weight blocks (S2D: 4 KB units, 12 for the nine compact tap blocks + 2 for the identity)
  const int nblocks = io.s2d ? kS2dUnits + 2 * p.sc_nk : p.n_kchunks * taps + p.sc_nk;
  const int64_t misc = 4096;
  int64_t w_bytes = 0, ra = 0;
  for (;; mt >>= 1) {
    p.mt = mt;
    p.npix = (int)round_up(kTileM * mt + max_shift - p.min_shift, 8);
    p.a_plane_bytes = (uint32_t)p.npix * 16u;
    p.a_slot_bytes = (uint32_t)(is_stem ? 1 : p.n_src * (p.kc / 8)) * p.a_plane_bytes;
    p.w_block_bytes = io.s2d ? (uint32_t)kS2dUnitBytes : (uint32_t)(p.kc * p.n * 2);
    p.sc_a_bytes = (uint32_t)p.sc_planes * p.a_plane_bytes;
    p.w_resident = (int64_t)nblocks * p.w_block_bytes <= 100 * 1024;
    static const int ws_stream = getenv("AT_TC_WS") != nullptr ? std::max(2, atoi(getenv("AT_TC_WS"))) : 4;   // A/B knob: weight ring depth
    p.ws = p.w_resident ? nblocks : ws_stream;
    while (!p.w_resident && p.ws > 2 && (int64_t)p.ws * p.w_block_bytes > 128 * 1024) --p.ws;
    w_bytes = (int64_t)p.ws * p.w_block_bytes;
    ra = (kSmemBudget - w_bytes - misc) / p.a_slot_bytes;
    while (ra < 2 && !p.w_resident && p.ws > 2) {
      --p.ws;
      w_bytes = (int64_t)p.ws * p.w_block_bytes;
      ra = (kSmemBudget - w_bytes - misc) / p.a_slot_bytes;
    }
    if (ra >= 2 || mt == 1) break;
  }
  p.acc_cols = (uint32_t)round_up((is_stem ? 4 : mt) * p.n, 32);
  p.as = (int)(512 / p.acc_cols);
  if (p.as > 4) p.as = 4;
  uint32_t cols = 32;
  while (cols < p.as * p.acc_cols) cols <<= 1;
  p.tmem_cols = cols;
  if (ra > 8) ra = 8;
  {
    static const int ra_max = getenv("AT_TC_RA_MAX") != nullptr ? atoi(getenv("AT_TC_RA_MAX")) : 8;   // A/B knob: ring depth cap
    if (ra > ra_max && ra_max >= 2) ra = ra_max;
  }
  if (ra < 2) {
    set_error("conv_tc: layer %s does not fit shared memory (a_slot %u B, weights %lld B)", L.name, p.a_slot_bytes,
              (long long)w_bytes);
    return AT_ERR_UNSUPPORTED;
  }
  p.ra = (int)ra;
  p.smem_a_off = 0;
  p.smem_w_off = (uint32_t)round_up((int64_t)p.ra * p.a_slot_bytes, 128);
  p.smem_bias_off = (uint32_t)round_up(p.smem_w_off + w_bytes, 128);
  p.smem_bar_off = (uint32_t)round_up(p.smem_bias_off + p.n * 4, 128);
  const uint32_t smem_bytes = p.smem_bar_off + 8 * (2 * p.ra + 2 * p.ws + 2 * p.as) + 16;
  p.n_tiles = (OL.P + kTileM * mt - 1) / (kTileM * mt);
  // the partial last wave of the persistent grid, sliced by output channels (AT_TC_NSPLIT=0 switches it off for A/B)
  p.t_full = (int)p.n_tiles; p.n_split = 1; p.split_shift = 0;
  {
    static const bool nsplit = getenv("AT_TC_NSPLIT") == nullptr || atoi(getenv("AT_TC_NSPLIT")) != 0;
    const int rest = (int)(p.n_tiles % kNumSMs);
    if (nsplit && rest > 0 && !is_stem && !io.s2d && io.res == nullptr && io.out_mode != 3 && io.out_mode != 2 && p.cout == p.n && p.n % 64 == 0) {
      int sh = 0;
      while ((p.n >> (sh + 1)) % 32 == 0 && (rest << (sh + 1)) <= kNumSMs) ++sh;   // slices of >= 32 channels, one wave
      if (sh > 0) { p.t_full = (int)p.n_tiles - rest; p.split_shift = sh; p.n_split = 1 << sh; }
    }
  }
  p.n_items = p.t_full + ((int)p.n_tiles - p.t_full) * p.n_split;
  const int grid = (int)std::min<long long>(p.n_items, kNumSMs);
  if (grid == 0) return AT_OK;
  const int ksteps = p.kc >> 4;
  p.div_hw = make_divider((uint32_t)(OL.Hp * OL.Wp));
  p.div_wp = make_divider((uint32_t)OL.Wp);
  {
    const int n_slabs = (p.n + 31) / 32;
    p.slab_shift = 0;
    while ((1 << p.slab_shift) < n_slabs) ++p.slab_shift;
    if ((1 << p.slab_shift) != n_slabs && io.out_mode != 3 && (mt > 1 || io.res != nullptr)) {
      set_error("conv_tc: layer %s: %d slabs per accumulator need mt == 1 and no residual", L.name, n_slabs);
      return AT_ERR_UNSUPPORTED;
    }
  }
  const int mode = io.out_mode == 3 ? 4 : io.out_mode == 2 ? 3 : (io.res != nullptr ? (io.out_mode == 1 ? 2 : 1) : (io.out_mode == 1 ? 5 : 0));
  using Kernel = void (*)(const ConvTcParams);
  Kernel fn = nullptr;
#define AT_TC_CASE(T, K, M, R, O) \
  if (!io.s2d && taps == T && ksteps == K && mt == M && p.w_resident == R && mode == O) fn = conv_tc_kernel<T, K, M, R, O>;
  if (io.s2d && p.w_resident == 1 && mode == 0 && mt == 2) fn = conv_tc_kernel<9, 4, 2, 1, 0, 1>;
  if (io.s2d && p.w_resident == 1 && mode == 5 && mt == 2) fn = conv_tc_kernel<9, 4, 2, 1, 5, 1>;
  if (io.s2d && p.w_resident == 1 && mode == 0 && mt == 1) fn = conv_tc_kernel<9, 4, 1, 1, 0, 1>;
  if (io.s2d && p.w_resident == 1 && mode == 5 && mt == 1) fn = conv_tc_kernel<9, 4, 1, 1, 5, 1>;
#define AT_TC_CASE3(T, K, M, R) \
  AT_TC_CASE(T, K, M, R, 0) AT_TC_CASE(T, K, M, R, 1) AT_TC_CASE(T, K, M, R, 2) AT_TC_CASE(T, K, M, R, 5)
  AT_TC_CASE(8, 1, 1, 1, 0)
  AT_TC_CASE3(9, 2, 1, 1) AT_TC_CASE3(9, 2, 2, 1)
  AT_TC_CASE3(9, 4, 1, 1) AT_TC_CASE3(9, 4, 2, 1)
  AT_TC_CASE3(9, 4, 1, 0) AT_TC_CASE3(9, 4, 2, 0)
  AT_TC_CASE(1, 2, 1, 1, 0) AT_TC_CASE(1, 4, 1, 1, 0) AT_TC_CASE(1, 4, 1, 1, 3)
  AT_TC_CASE(9, 5, 1, 0, 4) AT_TC_CASE(9, 5, 2, 0, 4)
#undef AT_TC_CASE3
#undef AT_TC_CASE
  if (fn == nullptr) {
    set_error("conv_tc: no kernel for taps=%d kc=%d mt=%d resident=%d mode=%d", taps, p.kc, mt, p.w_resident, mode);
    return AT_ERR_UNSUPPORTED;
  }
  {
    static std::mutex mu;
    static std::vector<std::pair<int, Kernel>> configured;  // function attributes are per device
    int dev = 0;
    cudaGetDevice(&dev);
    const std::pair<int, Kernel> key(dev, fn);
    std::lock_guard<std::mutex> lock(mu);
    if (std::find(configured.begin(), configured.end(), key) == configured.end()) {
      cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget + 4096);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      if (e != cudaSuccess) {
        set_error("cudaFuncSetAttribute(conv_tc_kernel): %s", cudaGetErrorString(e));
        return AT_ERR_CUDA;
      }
      configured.push_back(key);
    }
  }
  {
    // Programmatic dependent launch.  Backbone layers: off by default since round 2 - with one 220 KB CTA per SM the next
    // layer's CTAs cannot become resident before their predecessors exit anyway, and their resident-weight loads then
    // compete with the stragglers: 2.36 -> 2.33 ms per 2048 crops without it (ABAB on two boxes, AT_TC_PDL=1 for A/B).
    // The pose-head convs (out_mode 3) sit in the tail's chain of short kernels and follow AT_TAIL_PDL (default on).
    static const bool pdl_backbone = getenv("AT_TC_PDL") != nullptr && atoi(getenv("AT_TC_PDL")) != 0;
    static const bool pdl_tail = getenv("AT_TAIL_PDL") == nullptr || atoi(getenv("AT_TAIL_PDL")) != 0;
    const bool use_pdl = io.out_mode == 3 ? pdl_tail : pdl_backbone;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = use_pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, p);
    if (e != cudaSuccess) {
      count_launch();
      set_error("conv_tc_kernel launch: %s", cudaGetErrorString(e));
      return AT_ERR_CUDA;
    }
  }
  return check_launch("conv_tc_kernel");
}

// ------------------------------------------------------------------------------------------
// encoder orchestration
constexpr int kChunkCrops = 2048;

struct Plan {
  Level lv[4];   // activation geometry per resolution level (48, 24, 12, 6)
  int ch[4];     // channels of the level (lv[0].C is 64 in the space-to-depth view of level 0)
  long long buf_units[4];   // one activation buffer per level (16-byte units)
  long long ps_units[3];    // one phase image set member (phase stride) for the outputs of levels 0..2
  long long off_lv[4];      // workspace offsets (units): 4 buffers per level (X, T, Y, D)
  long long off_ps[3];
  long long total_units;
};

// AT_TC_S2D=0: layer 1 as plain 32 -> 32 convs (the round-1 form) instead of the x-space-to-depth form (conv_tc_kernel S2D)
bool s2d_enabled() {
  static const bool on = (getenv("AT_TC_S2D") == nullptr || atoi(getenv("AT_TC_S2D")) != 0) &&
                         (getenv("AT_STEM_MMA") == nullptr || atoi(getenv("AT_STEM_MMA")) != 0);
  return on;
}

Plan make_plan(long long nc) {
  Plan P;
  const int HW[4] = {48, 24, 12, 6}, C[4] = {32, 64, 128, 256};
  long long off = 0;
  for (int l = 0; l < 4; ++l) {
    P.lv[l] = l == 0 && s2d_enabled() ? make_level(48, 24, kS2dC, nc) : make_level(HW[l], HW[l], C[l], nc);
    P.ch[l] = C[l];
    P.buf_units[l] = P.lv[l].buf_units();
    P.off_lv[l] = off;
    off += 4 * P.buf_units[l];
  }
  for (int l = 0; l < 3; ++l) {
    // phase images of level l's output have the geometry of level l+1 but level l's channel count
    Level ph = make_level(HW[l + 1], HW[l + 1], C[l], nc);
    P.ps_units[l] = ph.buf_units();
    P.off_ps[l] = off;
    off += 4 * P.ps_units[l];
  }
  P.total_units = off;
  return P;
}

}  // namespace tc

int encoder_chunk_crops() { return tc::kChunkCrops; }

int64_t encoder_workspace_bytes_bf16(int c) {
  const long long nc = c < tc::kChunkCrops ? (c > 0 ? c : 1) : tc::kChunkCrops;
  return tc::make_plan(nc).total_units * 16 + 256;
}

// Zero what the kernels rely on but never write: the guard in front of every plane (the top /
// left padding of the first image) and the pad cells of the phase-split buffers.
int encoder_prepare_workspace(void* workspace, long long nc, cudaStream_t s) {
  using namespace tc;
  uint4* ws = reinterpret_cast<uint4*>(round_up((int64_t)(uintptr_t)workspace, 128));
  const Plan P = make_plan(nc);
  int rc;
  for (int l = 0; l < 4; ++l)
    if ((rc = zero_halo(ws + P.off_lv[l], 4 * P.lv[l].planes(), P.lv[l], false, s))) return rc;
  for (int l = 0; l < 3; ++l) {
    Level ph = make_level(P.lv[l + 1].H, P.lv[l + 1].W, P.ch[l], nc);
    if ((rc = zero_halo(ws + P.off_ps[l], 4 * ph.planes(), ph, true, s))) return rc;
  }
  return AT_OK;
}

// prepared_c > 0: the caller guarantees that the workspace was prepared for prepared_c crops per
// chunk (encoder_prepare_workspace) and only written by this function since.
int encoder_forward_bf16(const float* blob, const void* tc_blob, const void* crops, int crop_dtype, int c, float* features,
                         void* workspace, int64_t workspace_bytes, int fmt, int prepared_c, int32_t* status,
                         cudaStream_t front_stream, cudaStream_t release_stream, cudaStream_t s) {
  using namespace tc;
  struct StatusScope {  // every conv launch of this call reports fp16 saturation into `status`
    explicit StatusScope(int32_t* p) { g_status = p; }
    ~StatusScope() { g_status = nullptr; }
  } status_scope(status);
  const ConvLayer* T = layer_table();
  (void)T;
  uint4* ws = reinterpret_cast<uint4*>(round_up((int64_t)(uintptr_t)workspace, 128));
  for (int c0 = 0; c0 < c; c0 += kChunkCrops) {
    const long long nc = std::min(c - c0, kChunkCrops);
    const Plan P = make_plan(nc);
    int rc;
    bool prepared_now = false;
    if (nc != prepared_c) {
      if ((rc = encoder_prepare_workspace(workspace, nc, s))) return rc;
      prepared_c = (int)nc;
      prepared_now = true;
    }
    // Front stream: the caller produced the crops there and the stem of the first chunk may run there too, next to whatever
    // `s` still has queued in front of this call (BatchedHandTracker: the previous step's tail); `s` joins before layer 1.
    // Not when this call had to prepare the workspace on `s` (the stem writes it), and not for later chunks (their stem
    // overwrites planes the previous chunk's convs on `s` still read): then `s` simply waits for the front stream first.
    const bool stem_on_front = front_stream != nullptr && front_stream != s && c0 == 0 && !prepared_now;
    if (front_stream != nullptr && front_stream != s && c0 == 0 && !stem_on_front && (rc = order_streams(front_stream, s, 1))) return rc;
    Level phl[3];
    for (int l = 0; l < 3; ++l) phl[l] = make_level(P.lv[l + 1].H, P.lv[l + 1].W, P.ch[l], nc);
    // One span of the pass: the stem (l_begin == 0), levels [l_begin, l_end) and the projection (l_end == 4) over
    // crops [cs, cs + n) of the chunk.  A span over part of the chunk addresses the same planes at a crop offset
    // (the planes keep the chunk's stride, tiles are counted from the span's first pixel).
    auto run_span = [&](int l_begin, int l_end, long long cs, long long n, bool release) -> int {
    auto part = [&](const Level& L) { Level R = L; R.n_img = n; R.P = n * L.Hp * L.Wp; return R; };
    auto buf = [&](int l, int i) { return ws + P.off_lv[l] + i * P.buf_units[l] + P.lv[l].guard + cs * P.lv[l].Hp * P.lv[l].Wp; };
    auto ps_buf = [&](int l) { return ws + P.off_ps[l] + phl[l].guard + cs * phl[l].Hp * phl[l].Wp; };
    Level phs[3];
    for (int l = 0; l < 3; ++l) phs[l] = part(phl[l]);
    const long long cg0 = c0 + cs;  // first crop of the span in the caller's arrays
    const bool s2d = s2d_enabled();
    // stem.  Default: mma.sync kernel straight from the crops (registers, no TMEM read-back, no pack kernel).
    // AT_STEM_MMA=0: pack the crops into the 2x2-block plane S (first plane of the unused D buffer of level 0),
    // then the tcgen05 stem GEMM writes X of level 0 (A/B).
    static const bool stem_mma = getenv("AT_STEM_MMA") == nullptr || atoi(getenv("AT_STEM_MMA")) != 0;
    cudaStream_t ss = s;   // stream of the stem launch
    if (l_begin == 0 && stem_on_front && cs == 0) {
      if (stem_mma) ss = front_stream;
      else if ((rc = order_streams(front_stream, s, 1))) return rc;
    }
    if (l_begin > 0) {
    } else if (stem_mma) {
      const Level L0 = part(P.lv[0]);
      const uint2* wf = reinterpret_cast<const uint2*>(reinterpret_cast<const uint8_t*>(tc_blob) + tc_table()[L_STEM].offset + 8 * 1024);
      const float* bias = blob + layer_table()[L_STEM].bias_offset;
      const int grid = (int)std::min<long long>(n, 2 * kNumSMs);
#define AT_STEM_LAUNCH(TIN, F)                                                                                        \
  do {                                                                                                              \
    AT_CARVEOUT_ONCE((stem_mma_kernel<TIN, F>));                                                                    \
    launch_pdl(stem_mma_kernel<TIN, F>, dim3(grid), dim3(kStemThreads), 0, ss, (const TIN*)crops + (size_t)cg0 * 96 * 96, (int)n, \
               wf, bias, buf(0, 0), L0, g_status, s2d ? 1 : 0);                                                      \
  } while (0)
      if (crop_dtype == AT_U8) { if (fmt) AT_STEM_LAUNCH(uint8_t, 1); else AT_STEM_LAUNCH(uint8_t, 0); }
      else { if (fmt) AT_STEM_LAUNCH(float, 1); else AT_STEM_LAUNCH(float, 0); }
#undef AT_STEM_LAUNCH
      if ((rc = check_launch("stem_mma_kernel"))) return rc;
      if (ss != s && (rc = order_streams(ss, s, 1))) return rc;
    } else {
      const Level L0 = part(P.lv[0]);
      const unsigned blocks = (unsigned)((L0.P + 255) / 256);
      if (crop_dtype == AT_U8) {
        AT_CARVEOUT_ONCE(stem_pack_kernel<uint8_t>);
        stem_pack_kernel<uint8_t><<<blocks, 256, 0, s>>>((const uint8_t*)crops + (size_t)cg0 * 96 * 96, buf(0, 3), L0, fmt);
      } else {
        AT_CARVEOUT_ONCE(stem_pack_kernel<float>);
        stem_pack_kernel<float><<<blocks, 256, 0, s>>>((const float*)crops + (size_t)cg0 * 96 * 96, buf(0, 3), L0, fmt);
      }
      if ((rc = check_launch("stem_pack_kernel"))) return rc;
      ConvIo io;
      memset(&io, 0, sizeof(io));
      io.a_base = buf(0, 3); io.a_plane_px = L0.plane_px;
      io.out = buf(0, 0); io.out_plane_px = L0.plane_px; io.out_mode = 0; io.relu = true;
      if ((rc = launch_conv_tc(L_STEM, blob, tc_blob, L0, nullptr, io, fmt, s))) return rc;
    }
    const int nblocks[4] = {2, 3, 5, 2};
    // fold the 1x1 stride-2 shortcut of the first block of layers 2-4 into that block's conv2
    // (3 launches, their output writes and the residual reads go away); AT_TC_FOLD=0 keeps them apart
    static const bool fold = getenv("AT_TC_FOLD") == nullptr || atoi(getenv("AT_TC_FOLD")) != 0;
    // identity residuals of the other blocks: accumulated as a 10th tap with identity weights instead of being
    // added in the epilogue (level bit mask; measured per level on B200: -0.04 / -0.02 / -0.05 / -0.005 ms)
    static const int fold_id = getenv("AT_TC_FOLD_ID") != nullptr ? atoi(getenv("AT_TC_FOLD_ID")) : 15;
    int id = L_LAYER1;
    for (int l = 0; l < l_begin; ++l) id += 2 * nblocks[l] + (l > 0 ? 1 : 0);
    int cur = 0;  // index (0 = X, 2 = Y) of the buffer holding the current activation of the level
    for (int l = l_begin; l < l_end; ++l) {
      const Level LV = part(P.lv[l]);
      for (int b = 0; b < nblocks[l]; ++b) {
        const bool first = l > 0 && b == 0;
        const bool last = b == nblocks[l] - 1 && l < 3;
        ConvIo io;
        memset(&io, 0, sizeof(io));
        io.s2d = s2d && l == 0;
        uint4* Tm = buf(l, 1);
        const uint4* resid;
        if (first) {
          // conv1 (3x3 stride 2) and the 1x1 stride-2 shortcut both read the phase-split output of level l-1
          const uint4* src = ps_buf(l - 1);
          io.a_base = src; io.a_plane_px = phl[l - 1].plane_px; io.a_phase_units = P.ps_units[l - 1];
          io.out = Tm; io.out_plane_px = LV.plane_px; io.out_mode = 0; io.relu = true;
          if ((rc = launch_conv_tc(id, blob, tc_blob, LV, nullptr, io, fmt, s))) return rc;
          if (fold) {
            resid = nullptr;  // the shortcut conv is accumulated inside conv2 (below)
          } else {
            io.out = buf(l, 3); io.relu = false;
            if ((rc = launch_conv_tc(id + 2, blob, tc_blob, LV, nullptr, io, fmt, s))) return rc;
            resid = buf(l, 3);
          }
          cur = 2;  // so that the block output goes to X
        } else {
          io.a_base = buf(l, cur); io.a_plane_px = LV.plane_px;
          io.out = Tm; io.out_plane_px = LV.plane_px; io.out_mode = 0; io.relu = true;
          if ((rc = launch_conv_tc(id, blob, tc_blob, LV, nullptr, io, fmt, s))) return rc;
          if (release && release_stream != nullptr && l == 0 && b == 0 &&   // last chunk: layer1.0.conv1 is launched
              (rc = release_after_launch(s, release_stream)))
            return rc;
          resid = buf(l, cur);
        }
        memset(&io, 0, sizeof(io));
        io.s2d = s2d && l == 0;
        io.a_base = Tm; io.a_plane_px = LV.plane_px;
        io.res = resid; io.res_plane_px = LV.plane_px; io.relu = true;
        if (!first && (fold_id >> l) & 1) {  // identity residual as a 10th tap
          io.res = nullptr;
          io.sc_layer = -1; io.sc_src = resid; io.sc_plane_px = LV.plane_px;
        }
        if (first && fold) {
          io.sc_layer = id + 2;
          io.sc_src = ps_buf(l - 1);  // phase (0,0) of the previous level's output
          io.sc_plane_px = phl[l - 1].plane_px;
        }
        if (last) {
          io.out = ps_buf(l); io.out_plane_px = phl[l].plane_px; io.out_phase_units = P.ps_units[l];
          io.out_mode = 1;
          if ((rc = launch_conv_tc(id + 1, blob, tc_blob, LV, &phs[l], io, fmt, s))) return rc;
        } else {
          cur = 2 - cur;
          io.out = buf(l, cur); io.out_plane_px = LV.plane_px; io.out_mode = 0;
          if ((rc = launch_conv_tc(id + 1, blob, tc_blob, LV, nullptr, io, fmt, s))) return rc;
        }
        id += first ? 3 : 2;
      }
      cur = 0;
    }
    // proj 1x1 256->72, fp32 NHWC features.  The last block of level 3 wrote X or Y: recompute which.
    if (l_end == 4) {
      // level 3 has 2 blocks: block 0 (first) writes X (cur: 2 -> 0), block 1 writes Y
      ConvIo io;
      memset(&io, 0, sizeof(io));
      io.a_base = buf(3, 2); io.a_plane_px = P.lv[3].plane_px;
      io.out_mode = 2; io.out_f32 = features + (size_t)cg0 * 36 * 72; io.relu = false;
      if ((rc = launch_conv_tc(L_PROJ, blob, tc_blob, part(P.lv[3]), nullptr, io, fmt, s))) return rc;
    }
    return AT_OK;
    };  // run_span
    // AT_TC_SUB=<crops>: levels [0, AT_TC_SUB_LEVELS) run span by span so that a span's activations (layer 1: 154 KB per
    // crop and tensor) stay in the 126 MB L2 between the convs of a block; the deeper levels run over the whole chunk.
    static const int sub = getenv("AT_TC_SUB") != nullptr ? atoi(getenv("AT_TC_SUB")) : 0;
    static const int sub_levels = getenv("AT_TC_SUB_LEVELS") != nullptr ? std::min(4, std::max(1, atoi(getenv("AT_TC_SUB_LEVELS")))) : 1;
    const bool last_chunk = c0 + kChunkCrops >= c;
    if (sub > 0 && sub < nc) {
      for (long long cs = 0; cs < nc; cs += sub)
        if ((rc = run_span(0, sub_levels, cs, std::min<long long>(sub, nc - cs), last_chunk && cs + sub >= nc))) return rc;
      if (sub_levels < 4 && (rc = run_span(sub_levels, 4, 0, nc, false))) return rc;
    } else if ((rc = run_span(0, 4, 0, nc, last_chunk))) {
      return rc;
    }
  }
  (void)workspace_bytes;
  return AT_OK;
}

// ------------------------------------------------------------------------------------------
// Pose-head 3x3 convolutions (model_utils.py:195-208: two BasicBlocks at 76 / 72 channels on 6x6
// maps) as error-compensated fp16 x 3 shift-GEMMs on tcgen05.  x = x_hi + x_lo and w = w_hi + w_lo
// in fp16 (2 x 11 significant bits); the conv runs over 3 x 80 virtual input channels
//   [x_hi | x_lo | x_hi] * [w_hi | w_hi | w_lo]  =  x_hi w_hi + x_lo w_hi + x_hi w_lo
// accumulated in fp32 in TMEM: the same arithmetic contract as conv_nhwc_f16x3 (conv_f32.cu), at the
// UMMA rate instead of the mma.sync rate.  Weights are pre-scaled by 2^8 (their lo halves would fall
// into the fp16 denormals otherwise); the epilogue rescales.
namespace tc {

constexpr int kHeadCp = 80;             // channels per part (76 / 72 padded to a multiple of 16)
constexpr int kHeadParts = 3;
constexpr float kHeadWScale = 256.0f;

__global__ void head_prepack_kernel(const float* __restrict__ w /*[9][cin][coutP]*/, uint16_t* __restrict__ out, int cin,
                                    int cout, int coutP) {
  const int total = kHeadParts * 9 * kHeadCp * kHeadCp;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  // i -> (part = k-chunk, tap, kk, nn, e): the block order prepack_kernel uses
  const int e = i & 7;
  int r = i >> 3;
  const int nn = r % kHeadCp; r /= kHeadCp;
  const int kk = r % (kHeadCp / 8); r /= (kHeadCp / 8);
  const int t = r % 9;
  const int part = r / 9;
  const int ci = kk * 8 + e;
  float v = 0.f;
  if (ci < cin && nn < cout) v = w[((long long)t * cin + ci) * coutP + nn] * kHeadWScale;
  const __half hi = __float2half_rn(v);
  const __half lo = __float2half_rn(v - __half2float(hi));
  const __half o = part == 2 ? lo : hi;
  out[i] = *reinterpret_cast<const uint16_t*>(&o);
}

// fp32 NHWC [n,6,6,hc] -> split planes [hi | lo | hi] x 10 channel groups on the padded linear pixel
// order; pad cells are written as zeros.
__global__ void head_split_pack_kernel(const float* __restrict__ in, uint4* __restrict__ base /*plane0,px0*/, Level L,
                                       int hc) {
  pdl_prologue();
  // 32-bit index arithmetic (the launcher checks the range): the 64-bit divisions were most of this kernel
  const uint32_t total = (uint32_t)L.P * (kHeadCp / 8);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const uint32_t g = i / (kHeadCp / 8);
  const int cg = (int)(i - g * (kHeadCp / 8));
  const uint32_t HWp = (uint32_t)(L.Hp * L.Wp);
  const uint32_t img = g / HWp;
  const int rem = (int)(g - img * HWp);
  const int y = rem / L.Wp, x = rem - y * L.Wp;
  uint4 hi = make_uint4(0u, 0u, 0u, 0u), lo = hi;
  if (y < L.H && x < L.W) {
    const float* ip = in + ((long long)(img * L.H + y) * L.W + x) * hc + cg * 8;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = cg * 8 + k < hc ? ip[k] : 0.f;
    float a, b;
    hi.x = ptx::pack2(v[0], v[1], 0); ptx::unpack2(hi.x, 0, a, b); lo.x = ptx::pack2(v[0] - a, v[1] - b, 0);
    hi.y = ptx::pack2(v[2], v[3], 0); ptx::unpack2(hi.y, 0, a, b); lo.y = ptx::pack2(v[2] - a, v[3] - b, 0);
    hi.z = ptx::pack2(v[4], v[5], 0); ptx::unpack2(hi.z, 0, a, b); lo.z = ptx::pack2(v[4] - a, v[5] - b, 0);
    hi.w = ptx::pack2(v[6], v[7], 0); ptx::unpack2(hi.w, 0, a, b); lo.w = ptx::pack2(v[6] - a, v[7] - b, 0);
  }
  uint4* op = base + (long long)cg * L.plane_px + g;
  const long long part = (long long)(kHeadCp / 8) * L.plane_px;
  op[0] = hi; op[part] = lo; op[2 * part] = hi;
}

Level head_level(long long n) { return make_level(6, 6, kHeadParts * kHeadCp, n); }

}  // namespace tc

int64_t head_tc_weight_bytes() { return (int64_t)tc::kHeadParts * 9 * tc::kHeadCp * tc::kHeadCp * 2; }

int64_t head_tc_workspace_bytes(int n) { return 2 * tc::head_level(n > 0 ? n : 1).buf_units() * 16 + 256; }

int head_tc_prepack(const ConvLayer& L, const float* blob, void* dst, cudaStream_t s) {
  AT_REQUIRE(L.ksize == 3 && L.cin <= tc::kHeadCp && L.cout <= tc::kHeadCp, "head_tc_prepack: not a pose-head 3x3 conv");
  const int total = tc::kHeadParts * 9 * tc::kHeadCp * tc::kHeadCp;
  tc::head_prepack_kernel<<<(total + 255) / 256, 256, 0, s>>>(blob + L.weight_offset, reinterpret_cast<uint16_t*>(dst), L.cin,
                                                              L.cout, (int)round_up(L.cout, 4));
  return check_launch("head_prepack_kernel");
}

// The two BasicBlocks of the pose head: x [n,6,6,hc] fp32 NHWC (also the first residual) -> y (same
// shape; may alias x).  h2: fp32 scratch of the same shape.  wq[i]: head_tc_prepack output of conv i.
int head_tc_forward(const ConvLayer* L4, const float* blob, const void* const* wq, float* x, float* h2, float* y, int hc,
                    int n, void* workspace, cudaStream_t s) {
  using namespace tc;
  const Level HL = head_level(n);
  uint4* ws = reinterpret_cast<uint4*>(round_up((int64_t)(uintptr_t)workspace, 128));
  uint4* A = ws;
  uint4* B = ws + HL.buf_units();
  int rc;
  // the guards in front of / behind the planes are read as halo and never written by the kernels
  if ((rc = zero_halo(A, 2 * HL.planes(), HL, false, s))) return rc;
  {
    const long long total = HL.P * (kHeadCp / 8);
    AT_REQUIRE(total < (1LL << 31), "head_tc_forward: too many samples for one call");
    launch_pdl(head_split_pack_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, s, x, A + HL.guard, HL, hc);
    if ((rc = check_launch("head_split_pack_kernel"))) return rc;
  }
  auto conv = [&](int i, const uint4* in, const float* res, float* out_f32, uint4* out_split) -> int {
    TcConvDesc d;
    d.name = L4[i].name; d.ksize = 3; d.stride = 1; d.cout = L4[i].cout; d.n_pad = kHeadCp; d.kc = kHeadCp; d.nk = kHeadParts;
    d.w = wq[i]; d.bias = blob + L4[i].bias_offset;
    ConvIo io;
    memset(&io, 0, sizeof(io));
    io.a_base = in + HL.guard; io.a_plane_px = HL.plane_px;
    io.out = out_split != nullptr ? out_split + HL.guard : nullptr; io.out_plane_px = HL.plane_px;
    io.out_f32 = out_f32; io.res_f32 = res; io.out_mode = 3; io.relu = true;
    io.out_scale = 1.0f / kHeadWScale; io.split_planes = kHeadCp / 8; io.f32_ld = hc;
    return launch_conv_tc(-1, blob, nullptr, HL, nullptr, io, 0, s, &d);
  };
  if ((rc = conv(0, A, nullptr, nullptr, B))) return rc;   // block 0 conv1 + ReLU
  if ((rc = conv(1, B, x, h2, A))) return rc;              // block 0 conv2 + x, ReLU
  if ((rc = conv(2, A, nullptr, nullptr, B))) return rc;   // block 1 conv1 + ReLU
  return conv(3, B, h2, y, nullptr);                       // block 1 conv2 + h2, ReLU
}

}  // namespace abt

extern "C" int64_t at_model_bf16_blob_bytes(void) {
  abt::tc::tc_table();
  return abt::tc::g_tc_bytes;
}

extern "C" int at_model_prepack_tc(const float* blob, void* tc_blob, int precision, void* stream) {
  using namespace abt;
  AT_REQUIRE(blob && tc_blob, "at_model_prepack_tc: NULL pointer");
  AT_REQUIRE(precision == AT_PREC_BF16 || precision == AT_PREC_FP16, "at_model_prepack_tc: precision must be bf16 or fp16");
  const ConvLayer* T = layer_table();
  const tc::TcLayer* TT = tc::tc_table();
  tc::prepack_stem_kernel<<<16, 256, 0, (cudaStream_t)stream>>>(
      blob + T[L_STEM].weight_offset, reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(tc_blob) + TT[L_STEM].offset),
      precision == AT_PREC_BF16 ? 1 : 0);
  {
    int rc = check_launch("prepack_stem_kernel");
    if (rc) return rc;
  }
  tc::prepack_stem_frag_kernel<<<2, 256, 0, (cudaStream_t)stream>>>(
      blob + T[L_STEM].weight_offset, reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(tc_blob) + TT[L_STEM].offset + 8 * 1024),
      precision == AT_PREC_BF16 ? 1 : 0);
  {
    int rc = check_launch("prepack_stem_frag_kernel");
    if (rc) return rc;
  }
  for (int i = L_LAYER1; i <= L_PROJ; ++i) {
    const int taps = T[i].ksize * T[i].ksize;
    const long long total = (long long)taps * T[i].cin * TT[i].n_pad;
    tc::prepack_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        blob + T[i].weight_offset, reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(tc_blob) + TT[i].offset), taps,
        T[i].cin, T[i].cout, (int)round_up(T[i].cout, 4), TT[i].n_pad, TT[i].kc, precision == AT_PREC_BF16 ? 1 : 0);
    int rc = check_launch("prepack_kernel");
    if (rc) return rc;
    if (tc::is_identity_host(i)) {
      const int64_t main_bytes = (int64_t)taps * T[i].cin * TT[i].n_pad * 2;
      const long long tot = (long long)T[i].cout * TT[i].n_pad;
      tc::prepack_identity_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
          reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(tc_blob) + TT[i].offset + main_bytes), T[i].cout, TT[i].n_pad,
          TT[i].kc, precision == AT_PREC_BF16 ? 1 : 0);
      if ((rc = check_launch("prepack_identity_kernel"))) return rc;
    }
    if (TT[i].s2d_offset >= 0) {  // layer 1: the x-space-to-depth form (+ its 64 x 64 identity block)
      uint16_t* dst = reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(tc_blob) + TT[i].s2d_offset);
      tc::prepack_s2d_kernel<<<(unsigned)((tc::kS2dMainBytes / 2 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
          blob + T[i].weight_offset, dst, (int)round_up(T[i].cout, 4), precision == AT_PREC_BF16 ? 1 : 0);
      if ((rc = check_launch("prepack_s2d_kernel"))) return rc;
      if (tc::is_identity_host(i)) {
        tc::prepack_identity_kernel<<<(tc::kS2dC * tc::kS2dC + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
            dst + tc::kS2dMainBytes / 2, tc::kS2dC, tc::kS2dC, tc::kS2dC, precision == AT_PREC_BF16 ? 1 : 0);
        if ((rc = check_launch("prepack_identity_kernel"))) return rc;
      }
    }
    if (tc::is_fold_host(i)) {  // second copy of the block's 1x1 shortcut weights, behind conv2's own blocks
      const ConvLayer& S = T[i + 1];
      const int64_t main_bytes = (int64_t)taps * T[i].cin * TT[i].n_pad * 2;
      const long long tot = (long long)S.cin * TT[i].n_pad;
      tc::prepack_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
          blob + S.weight_offset, reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(tc_blob) + TT[i].offset + main_bytes), 1,
          S.cin, S.cout, (int)round_up(S.cout, 4), TT[i].n_pad, S.cin < 64 ? S.cin : 64, precision == AT_PREC_BF16 ? 1 : 0);
      if ((rc = check_launch("prepack_kernel"))) return rc;
    }
  }
  return AT_OK;
}

extern "C" int at_model_prepack_bf16(const float* blob, void* bf16_blob, void* stream) {
  return at_model_prepack_tc(blob, bf16_blob, AT_PREC_BF16, stream);
}

// Debug / test entry: one backbone conv layer through the tcgen05 path on fp32 NHWC tensors.
// in [n,H,W,cin] -> out [n,Ho,Wo,cout]; res (optional) [n,Ho,Wo,cout].  workspace: at_debug_conv_tc_workspace_bytes.
/* Debug: when non-NULL every conv_tc launch writes [148][16] cycle counters there
 * (MMA warp: total, wait acc_empty, wait a_full, wait w_full, tiles; epilogue groups at +8: total, wait). */
extern "C" void at_debug_set_prof(long long* dev_ptr) { abt::tc::g_prof = dev_ptr; abt::tc::g_prof_seq = -1; }
extern "C" void at_debug_set_prof_seq(long long* dev_ptr) { abt::tc::g_prof = dev_ptr; abt::tc::g_prof_seq = 0; }

extern "C" int64_t at_debug_conv_tc_workspace_bytes(int n_img, int H, int W) {
  using namespace abt;
  tc::Level a = tc::make_level(H, W, 256, n_img);
  return a.buf_units() * 16 * 16 + 4096;
}

extern "C" int at_debug_conv_tc(const float* blob, const void* tc_blob, int layer_id, const float* in, int n_img, int H,
                                int W, const float* res, int relu, int precision, int mt, float* out, void* workspace,
                                int64_t workspace_bytes, void* stream_) {
  using namespace abt;
  AT_REQUIRE(layer_id >= L_LAYER1 && layer_id <= L_PROJ, "at_debug_conv_tc: layer_id must be a backbone conv");
  AT_REQUIRE(precision == AT_PREC_BF16 || precision == AT_PREC_FP16, "at_debug_conv_tc: bad precision");
  AT_REQUIRE(workspace_bytes >= at_debug_conv_tc_workspace_bytes(n_img, H, W), "at_debug_conv_tc: workspace too small");
  cudaStream_t s = (cudaStream_t)stream_;
  const int fmt = precision == AT_PREC_BF16 ? 1 : 0;
  const ConvLayer& L = layer_table()[layer_id];
  const int Ho = H / L.stride, Wo = W / L.stride;
  tc::Level OL = tc::make_level(Ho, Wo, (int)round_up(L.cout, 8), n_img);   // output geometry
  tc::Level IL = tc::make_level(Ho, Wo, L.cin, n_img);                      // input (phase) geometry
  uint4* ws = reinterpret_cast<uint4*>(round_up((int64_t)(uintptr_t)workspace, 128));
  const long long quarter = at_debug_conv_tc_workspace_bytes(n_img, H, W) / 16 / 4;
  uint4* a_buf = ws;                 // up to 4 phase images
  uint4* r_buf = ws + 2 * quarter;
  uint4* o_buf = ws + 3 * quarter;
  int rc;
  const int n_ph = L.stride == 2 ? 4 : 1;
  for (int ph = 0; ph < n_ph; ++ph)
    if ((rc = tc::zero_halo(a_buf + ph * IL.buf_units(), IL.planes(), IL, true, s))) return rc;
  {
    const long long total = (long long)n_img * H * W * (L.cin / 8);
    tc::pack_planar_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(in, a_buf + IL.guard, IL, L.stride == 2,
                                                                            IL.buf_units(), H, W, fmt);
    if ((rc = check_launch("pack_planar_kernel"))) return rc;
  }
  tc::ConvIo io;
  memset(&io, 0, sizeof(io));
  if (res != nullptr) {
    if ((rc = tc::zero_halo(r_buf, OL.planes(), OL, true, s))) return rc;
    const long long total = (long long)n_img * Ho * Wo * (OL.C / 8);
    AT_REQUIRE(L.cout % 8 == 0, "at_debug_conv_tc: residual needs cout % 8 == 0");
    tc::pack_planar_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(res, r_buf + OL.guard, OL, 0, 0, Ho, Wo, fmt);
    if ((rc = check_launch("pack_planar_kernel"))) return rc;
    io.res = r_buf + OL.guard; io.res_plane_px = OL.plane_px;
  }
  io.a_base = a_buf + IL.guard; io.a_plane_px = IL.plane_px; io.a_phase_units = IL.buf_units();
  io.relu = relu != 0;
  tc::g_mt_override = mt;
  if (layer_id == L_PROJ) {
    io.out_mode = 2; io.out_f32 = out;
    rc = tc::launch_conv_tc(layer_id, blob, tc_blob, OL, nullptr, io, fmt, s);
    tc::g_mt_override = 0;
    return rc;
  }
  if ((rc = tc::zero_halo(o_buf, OL.planes(), OL, false, s))) return rc;
  io.out = o_buf + OL.guard; io.out_plane_px = OL.plane_px; io.out_mode = 0;
  rc = tc::launch_conv_tc(layer_id, blob, tc_blob, OL, nullptr, io, fmt, s);
  tc::g_mt_override = 0;
  if (rc) return rc;
  const long long total = (long long)n_img * Ho * Wo * (OL.C / 8);
  tc::unpack_planar_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(o_buf + OL.guard, out, OL, L.cout, fmt);
  return check_launch("unpack_planar_kernel");
}
