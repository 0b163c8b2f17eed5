// The shift-GEMM of conv_tc_kernel on CTA PAIRS (tcgen05 cta_group::2) for the 3x3 convolutions with 64 or
// more output channels (layers 2-4).  Included by conv_tc.cu inside namespace abt::tc.
//
// conv_tc_kernel is bound by shared-memory bandwidth there, not by the tensor pipe: per MMA (128 pixels x N x
// 16) a CTA reads the A tile (4 KB) and the whole B tile (32 N bytes), and layers 3-4, whose weights do not
// fit next to the activation ring, also stream all of them through shared memory for every tile (measured
// 88-97 cycles per N = 128 MMA against 64 of math).  A pair issues ONE MMA of M = 256: each
// CTA contributes its own 128 pixels (its own tile of the padded pixel order) and only HALF of the output
// channels of each weight block, so B reads and weight traffic per CTA halve and layer 3's weights become
// resident.  Roles per CTA as in conv_tc_kernel; differences:
//   * only the leader (cluster rank 0) issues MMAs / commits (multicast to the same barrier of both CTAs);
//   * the peer's warp 1 relays "my A slot / my weight block has landed" to the leader's pa_full / pw_full
//     barriers (remote mbarrier arrive); the leader waits for its own and the peer's before it issues;
//   * the epilogue warps of both CTAs arrive on the LEADER's acc_empty barrier (count 8);
//   * weight blocks stay in the standard prepacked layout: a CTA copies its half of each 8-channel K group.
// 3x3 taps only (+ the folded shortcut / identity chunks), no epilogue residual, output modes 0 (planar) and
// 5 (phase split).  Barrier layout behind smem_bar_off: a_full[ra] a_empty[ra] w_full[ws] w_empty[ws]
// acc_full[as] acc_empty[as] pa_full[ra] pw_full[ws] tmem_slot.
#pragma once

template <int KSTEPS, int MT, int WRES, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(kMaxRegs) conv_tc2_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  using namespace ptx;
  constexpr int NTAPS = 9;
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const uint32_t bar0 = sbase + p.smem_bar_off;
  const uint32_t a_full = bar0, a_empty = a_full + 8 * p.ra;
  const uint32_t w_full = a_empty + 8 * p.ra, w_empty = w_full + 8 * p.ws;
  const uint32_t acc_full = w_empty + 8 * p.ws, acc_empty = acc_full + 8 * p.as;
  const uint32_t pa_full = acc_empty + 8 * p.as, pw_full = pa_full + 8 * p.ra;
  const uint32_t tmem_slot = pw_full + 8 * p.ws;
  float* bias_s = reinterpret_cast<float*>(smem + p.smem_bias_off);
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.ra; ++i) { mbar_init(a_full + 8 * i, 1); mbar_init(a_empty + 8 * i, 1); mbar_init(pa_full + 8 * i, 1); }
    for (int i = 0; i < p.ws; ++i) { mbar_init(w_full + 8 * i, 1); mbar_init(w_empty + 8 * i, 1); mbar_init(pw_full + 8 * i, 1); }
    for (int i = 0; i < p.as; ++i) { mbar_init(acc_full + 8 * i, 1); mbar_init(acc_empty + 8 * i, 8); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc_cg2(tmem_slot, p.tmem_cols);
    tmem_relinquish_cg2();
  }
  for (int i = threadIdx.x; i < p.n; i += kThreads)
    bias_s[i] = i < p.cout ? p.bias[i] + (p.bias2 != nullptr ? p.bias2[i] : 0.f) : 0.f;
  tc_fence_before();
  cluster_sync_all();   // barriers of BOTH CTAs are initialised before anybody arrives on them remotely
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + (tmem_slot - sbase));

  const int pair = blockIdx.x >> 1, n_pairs = (int)gridDim.x >> 1;
  const int n_tiles = (int)p.n_tiles;
  const int n_pair_tiles = (n_tiles + 1) >> 1;
  const int n_my = pair < n_pair_tiles ? (n_pair_tiles - pair + n_pairs - 1) / n_pairs : 0;   // the same in both CTAs
  const int nk = p.n_kchunks;
  constexpr int kPlanesPerChunk = KSTEPS * 2;
  constexpr int kTilePx = kTileM * MT;
  const int nmain = nk * NTAPS;
  const int nblocks = nmain + p.sc_nk;
  auto my_tile = [&](int it) { return 2 * (pair + it * n_pairs) + (int)rank; };

  if (warp == 0) {
    if (lane == 0) {  // ---------------- weight producer: my half of the output channels of every block
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.w);
      const uint32_t half_kg = p.w_kg_bytes >> 1;
      auto load_block = [&](int b, uint32_t dst, uint32_t bar) {
        const uint32_t bytes = b < nmain ? p.w_block_bytes : p.sc_w_bytes;   // per CTA
        const int kgs = (int)(bytes / half_kg);
        mbar_expect_tx(bar, bytes);
        const uint8_t* src = wsrc + (size_t)b * p.w_src_bytes + rank * half_kg;
        for (int kg = 0; kg < kgs; ++kg) bulk_g2s(dst + kg * half_kg, src + (size_t)kg * p.w_kg_bytes, half_kg, bar);
      };
      if (WRES) {
        if (n_my > 0)
          for (int b = 0; b < nblocks; ++b) load_block(b, sbase + p.smem_w_off + b * p.w_block_bytes, w_full + 8 * b);
      } else {
        Ring r;
        for (int it = 0; it < n_my; ++it)
          for (int b = 0; b < nblocks; ++b) {
            mbar_wait(w_empty + 8 * r.slot, r.phase ^ 1u);
            load_block(b, sbase + p.smem_w_off + r.slot * p.w_block_bytes, w_full + 8 * r.slot);
            r.advance(p.ws);
          }
      }
    }
  }
  else if (warp == 1 && rank == 0) {
    // ---------------- MMA issuer (leader): M = 256 = my 128 pixels + the peer's, N = all output channels
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t idesc = umma_idesc(2 * kTileM, p.n, p.fmt);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);
    const uint32_t nl = (uint32_t)p.n >> 1;                                       // B rows held by each CTA
    const uint32_t a_lo0 = ((p.a_plane_bytes >> 4) << 16) + ((sbase + p.smem_a_off) >> 4);
    const uint32_t b_lo0 = (nl << 16) + ((sbase + p.smem_w_off) >> 4);            // LBO = (n/2)*16 B
    const uint32_t a_slot16 = p.a_slot_bytes >> 4, w_blk16 = p.w_block_bytes >> 4;
    const uint32_t a_kstep16 = (2 * p.a_plane_bytes) >> 4, b_kstep16 = 2 * nl;
    uint32_t tap_off[NTAPS];
#pragma unroll
    for (int t = 0; t < NTAPS; ++t)
      tap_off[t] = ((uint32_t)p.tap_src[t] * kPlanesPerChunk * p.a_plane_bytes >> 4) + (uint32_t)(p.tap_shift[t] - p.min_shift);
    Ring ra, rw, rc;
    for (int it = 0; it < n_my; ++it) {
      mbar_wait_cluster(acc_empty + 8 * rc.slot, rc.phase ^ 1u);
      tc_fence_after();
      const uint32_t d_base = tm + rc.slot * p.acc_cols;
      for (int kci = 0; kci < nk; ++kci) {
        mbar_wait(a_full + 8 * ra.slot, ra.phase);
        mbar_wait_cluster(pa_full + 8 * ra.slot, ra.phase);
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + ra.slot * a_slot16;
        const uint32_t acc0 = kci != 0;
        if (WRES) {
          if (it == 0) {
            for (int t = 0; t < NTAPS; ++t) {
              mbar_wait(w_full + 8 * (kci * NTAPS + t), 0);
              mbar_wait_cluster(pw_full + 8 * (kci * NTAPS + t), 0);
            }
            tc_fence_after();
          }
          if (elect_one()) {
            const uint32_t b_lo = b_lo0 + kci * NTAPS * w_blk16;
#pragma unroll
            for (int t = 0; t < NTAPS; ++t) {
#pragma unroll
              for (int j = 0; j < KSTEPS; ++j) {
                const uint64_t bd = ((uint64_t)desc_hi << 32) | (b_lo + t * w_blk16 + j * b_kstep16);
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                  const uint64_t ad = ((uint64_t)desc_hi << 32) | (a_lo + tap_off[t] + j * a_kstep16 + m * kTileM);
                  umma_f16_cg2(d_base + m * p.n, ad, bd, idesc, (t | j) != 0 ? 1u : acc0);
                }
              }
            }
            umma_commit_cg2(a_empty + 8 * ra.slot);
          }
          __syncwarp();
        } else {
#pragma unroll
          for (int t = 0; t < NTAPS; ++t) {
            mbar_wait(w_full + 8 * rw.slot, rw.phase);
            mbar_wait_cluster(pw_full + 8 * rw.slot, rw.phase);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t b_lo = b_lo0 + rw.slot * w_blk16;
#pragma unroll
              for (int j = 0; j < KSTEPS; ++j) {
                const uint64_t bd = ((uint64_t)desc_hi << 32) | (b_lo + j * b_kstep16);
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                  const uint64_t ad = ((uint64_t)desc_hi << 32) | (a_lo + tap_off[t] + j * a_kstep16 + m * kTileM);
                  umma_f16_cg2(d_base + m * p.n, ad, bd, idesc, (t | j) != 0 ? 1u : acc0);
                }
              }
              umma_commit_cg2(w_empty + 8 * rw.slot);
              if (t == NTAPS - 1) umma_commit_cg2(a_empty + 8 * ra.slot);
            }
            __syncwarp();
            rw.advance(p.ws);
          }
        }
        ra.advance(p.ra);
      }
      // folded shortcut / identity residual: one tap at shift 0 over extra K chunks
      for (int sc = 0; sc < p.sc_nk; ++sc) {
        mbar_wait(a_full + 8 * ra.slot, ra.phase);
        mbar_wait_cluster(pa_full + 8 * ra.slot, ra.phase);
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + ra.slot * a_slot16 + (uint32_t)(-p.min_shift);
        uint32_t b_lo;
        if (WRES) {
          if (it == 0) {
            mbar_wait(w_full + 8 * (nmain + sc), 0);
            mbar_wait_cluster(pw_full + 8 * (nmain + sc), 0);
            tc_fence_after();
          }
          b_lo = b_lo0 + (nmain + sc) * w_blk16;
        } else {
          mbar_wait(w_full + 8 * rw.slot, rw.phase);
          mbar_wait_cluster(pw_full + 8 * rw.slot, rw.phase);
          tc_fence_after();
          b_lo = b_lo0 + rw.slot * w_blk16;
        }
        if (elect_one()) {
          for (int j = 0; j < p.sc_ksteps; ++j) {
            const uint64_t bd = ((uint64_t)desc_hi << 32) | (b_lo + j * b_kstep16);
#pragma unroll
            for (int m = 0; m < MT; ++m) {
              const uint64_t ad = ((uint64_t)desc_hi << 32) | (a_lo + j * a_kstep16 + m * kTileM);
              umma_f16_cg2(d_base + m * p.n, ad, bd, idesc, 1u);
            }
          }
          if (!WRES) umma_commit_cg2(w_empty + 8 * rw.slot);
          umma_commit_cg2(a_empty + 8 * ra.slot);
        }
        __syncwarp();
        if (!WRES) rw.advance(p.ws);
        ra.advance(p.ra);
      }
      if (elect_one()) umma_commit_cg2(acc_full + 8 * rc.slot);
      __syncwarp();
      rc.advance(p.as);
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ---------------- relay (peer): tell the leader which of MY operands have landed, in its order
      const uint32_t l_pa = mapa(pa_full, 0), l_pw = mapa(pw_full, 0);
      Ring ra, rw;
      for (int it = 0; it < n_my; ++it) {
        for (int kci = 0; kci < nk; ++kci) {
          mbar_wait(a_full + 8 * ra.slot, ra.phase);
          mbar_arrive_cluster(l_pa + 8 * ra.slot);
          if (WRES) {
            if (it == 0)
              for (int t = 0; t < NTAPS; ++t) {
                mbar_wait(w_full + 8 * (kci * NTAPS + t), 0);
                mbar_arrive_cluster(l_pw + 8 * (kci * NTAPS + t));
              }
          } else {
            for (int t = 0; t < NTAPS; ++t) {
              mbar_wait(w_full + 8 * rw.slot, rw.phase);
              mbar_arrive_cluster(l_pw + 8 * rw.slot);
              rw.advance(p.ws);
            }
          }
          ra.advance(p.ra);
        }
        for (int sc = 0; sc < p.sc_nk; ++sc) {
          mbar_wait(a_full + 8 * ra.slot, ra.phase);
          mbar_arrive_cluster(l_pa + 8 * ra.slot);
          if (WRES) {
            if (it == 0) { mbar_wait(w_full + 8 * (nmain + sc), 0); mbar_arrive_cluster(l_pw + 8 * (nmain + sc)); }
          } else {
            mbar_wait(w_full + 8 * rw.slot, rw.phase);
            mbar_arrive_cluster(l_pw + 8 * rw.slot);
            rw.advance(p.ws);
          }
          ra.advance(p.ra);
        }
      }
    }
  }
  else if (warp == 2) {
    if (lane == 0) {  // ---------------- activation producer: MY tile (a pair's odd tile may lie behind the last one)
      Ring ra;
      pdl_wait();
      for (int it = 0; it < n_my; ++it) {
        const int tile = my_tile(it) < n_tiles ? my_tile(it) : 0;   // dummy tile: any readable pixels, its results are dropped
        const long long g0 = (long long)tile * kTilePx + p.min_shift;
        for (int kci = 0; kci < nk; ++kci) {
          mbar_wait(a_empty + 8 * ra.slot, ra.phase ^ 1u);
          mbar_expect_tx(a_full + 8 * ra.slot, p.a_slot_bytes);
          uint32_t dst = sbase + p.smem_a_off + ra.slot * p.a_slot_bytes;
          for (int sidx = 0; sidx < p.n_src; ++sidx) {
            const uint4* src = p.a_src[sidx] + ((long long)(kci * kPlanesPerChunk) * p.a_plane_px + g0);
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
    const int qd = warp & 3;
    Ring rc;
    if (grp == 1) rc.advance(p.as);
    const int ncg = (p.cout + 7) >> 3;
    constexpr int out_mode = MODE == 5 ? 1 : 0;
    const int n_slabs = p.n >> 5;   // n is 64, 128 or 256
    const uint32_t acc_empty_leader = rank == 0 ? acc_empty : mapa(acc_empty, 0);
    uint32_t ovf = 0u;
    pdl_wait();
    for (int it = grp; it < n_my; it += 2) {
      const int tile = my_tile(it);
      const uint32_t gbase = (uint32_t)tile * kTilePx + qd * 32 + lane;
      mbar_wait(acc_full + 8 * rc.slot, rc.phase);
      tc_fence_after();
      if (tile < n_tiles) {
#pragma unroll 1
        for (int m = 0; m < MT; ++m) {
          const uint32_t g = gbase + m * kTileM;
          const uint32_t img = fast_div(g, p.div_hw);
          const uint32_t rem = g - img * p.div_hw.d;
          const uint32_t y = fast_div(rem, p.div_wp), x = rem - y * p.div_wp.d;
          const bool valid = (long long)g < p.P && y < (uint32_t)p.H && x < (uint32_t)p.W;
          const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + rc.slot * p.acc_cols + m * p.n;
          uint4* out_ph = nullptr;
          long long o1 = 0;
          if (out_mode == 1) {
            out_ph = p.out_ph[(y & 1) * 2 + (x & 1)];
            o1 = (long long)img * (p.Hp2 * p.Wp2) + (y >> 1) * p.Wp2 + (x >> 1);
          }
#pragma unroll 1
          for (int sl = 0; sl < n_slabs; ++sl) {
            const int c0 = sl << 5;
            uint32_t v[32];
            tmem_ld_x32(taddr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              const int cg = (c0 >> 3) + h;
              const float4 b0 = *reinterpret_cast<const float4*>(bias_s + cg * 8);
              const float4 b1 = *reinterpret_cast<const float4*>(bias_s + cg * 8 + 4);
              float o[8];
              o[0] = __uint_as_float(v[h * 8 + 0]) + b0.x; o[1] = __uint_as_float(v[h * 8 + 1]) + b0.y;
              o[2] = __uint_as_float(v[h * 8 + 2]) + b0.z; o[3] = __uint_as_float(v[h * 8 + 3]) + b0.w;
              o[4] = __uint_as_float(v[h * 8 + 4]) + b1.x; o[5] = __uint_as_float(v[h * 8 + 5]) + b1.y;
              o[6] = __uint_as_float(v[h * 8 + 6]) + b1.z; o[7] = __uint_as_float(v[h * 8 + 7]) + b1.w;
              if (p.relu) {
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaxf(o[i], 0.f);
              }
              uint4 pk = make_uint4(0u, 0u, 0u, 0u);
              if (valid) {
                pk.x = pack2(o[0], o[1], p.fmt); pk.y = pack2(o[2], o[3], p.fmt);
                pk.z = pack2(o[4], o[5], p.fmt); pk.w = pack2(o[6], o[7], p.fmt);
                if (p.fmt == 0) ovf |= overflow_probe(pk.x) | overflow_probe(pk.y) | overflow_probe(pk.z) | overflow_probe(pk.w);
              }
              if (cg < ncg) {
                if (out_mode == 0) p.out[(long long)cg * p.out_plane_px + g] = pk;
                else if (valid) out_ph[(long long)cg * p.out_ph_plane_px + o1] = pk;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) mbar_arrive(acc_empty_leader + 8 * rc.slot);
        else mbar_arrive_cluster(acc_empty_leader + 8 * rc.slot);
      }
      rc.advance(p.as);
      rc.advance(p.as);
    }
    if (p.status != nullptr && (ovf & kOverflowProbeMask) != 0u) atomicOr(p.status, 2);
  }
  tc_fence_before();
  cluster_sync_all();   // both CTAs are done with both TMEMs (and with each other's barriers)
  if (warp == 2) tmem_dealloc_cg2(tmem_base, p.tmem_cols);
}
