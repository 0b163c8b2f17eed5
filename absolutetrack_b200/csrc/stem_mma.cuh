// Stem on the legacy tensor-core path (mma.sync), accumulators in REGISTERS.
// Included by conv_tc.cu inside namespace abt::tc (needs Level, ptx::pack1/pack2/overflow_probe).
//
// conv3x3(1->32)+BN+ReLU+maxpool2 (model_utils.py:119-124) has 0.5 % of the encoder's FLOPs but every
// pre-pool value has to be compared: on tcgen05 that is four fp32 accumulators per output read back from
// TMEM (512 B per pooled pixel at 64 B/clk/SM: 150 us per 2048 crops before anything else), plus a separate
// pack kernel.  Here one m16n8k16 group covers 16 pooled pixels of a row: K = the 4x4 input patch around the
// 2x2 block of a pooled pixel (exactly 16 values), N = 4 pooling phases x 32 channels.  The four phase
// accumulators of a channel land in the same thread, so the pool is three FMNMX in registers; bias, ReLU,
// 16-bit pack and the planar store follow (s2d: into the x-space-to-depth view of level 0 - rows of 24 pixel
// pairs, planes 0-3 = even columns, 4-7 = odd columns; L is then the 48 x 24 x 64 level).  The crop is staged once per CTA as 16-bit values (the /255
// through a 256-entry table, bit-identical to the fp32 division the other paths use) in words
// (col 2m-1, col 2m), so every A fragment register is one aligned LDS.32.  Writes the level-0 planar layout
// including its zero pad cells.
#pragma once

constexpr int kStemPW = 50;        // words per staged row (49 used)
constexpr int kStemRows = 98;      // input rows -1 .. 96
constexpr int kStemThreads = 256;

// B fragments of mma.m16n8k16 (col-major K x N): [n-block b = phase q * 4 + channel block][lane] -> 2 registers
__global__ void prepack_stem_frag_kernel(const float* __restrict__ w /*[9][32]*/, uint2* __restrict__ out /*[16][32]*/, int fmt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 16 * 32) return;
  const int b = i >> 5, lane = i & 31;
  const int col = b * 8 + (lane >> 2), q = col >> 5, ch = col & 31, oy = q >> 1, ox = q & 1;
  auto wv = [&](int k) -> float {   // weight of patch position k = dy*4+dx for output phase (oy,ox)
    const int dy = k >> 2, dx = k & 3, r = dy - oy, t = dx - ox;
    return (r >= 0 && r <= 2 && t >= 0 && t <= 2) ? w[(r * 3 + t) * 32 + ch] : 0.f;
  };
  const int k0 = 2 * (lane & 3);
  out[i] = make_uint2(ptx::pack2(wv(k0), wv(k0 + 1), fmt), ptx::pack2(wv(k0 + 8), wv(k0 + 9), fmt));
}

template <int FMT>
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], const uint2& b) {
  if (FMT)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

template <typename TIn, int FMT>
__global__ void __launch_bounds__(kStemThreads, 2)
stem_mma_kernel(const TIn* __restrict__ crops, int n_crops, const uint2* __restrict__ wfrag, const float* __restrict__ bias,
                uint4* __restrict__ out /*plane 0, pixel 0*/, Level L, int32_t* __restrict__ status, int s2d) {
  __shared__ __align__(16) uint32_t tile[kStemRows * kStemPW];
  __shared__ uint16_t lut[256];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int r = lane >> 2, j = lane & 3;
  ptx::pdl_launch_dependents();   // layer1.0.conv1 may set up while this grid drains (it waits before its first load)
  uint2 bf[16];
#pragma unroll
  for (int b = 0; b < 16; ++b) bf[b] = __ldg(wfrag + b * 32 + lane);
  float bs[4][2];
#pragma unroll
  for (int cb = 0; cb < 4; ++cb) { bs[cb][0] = __ldg(bias + cb * 8 + 2 * j); bs[cb][1] = __ldg(bias + cb * 8 + 2 * j + 1); }
  if (sizeof(TIn) == 1) lut[tid] = ptx::pack1(__fdiv_rn((float)tid, 255.0f), FMT);
  // the border rows (input rows -1 and 96) stay zero; the border columns are rewritten as zeros with every crop
  for (int i = tid; i < 2 * kStemPW; i += kStemThreads) tile[i < kStemPW ? i : (kStemRows - 1) * kStemPW + (i - kStemPW)] = 0u;
  const int HWp = L.Hp * L.Wp;
  uint32_t ovf = 0u;
  ptx::pdl_wait();   // the crops come from the previous kernel of the stream (weights, bias and the table above do not)
  for (int crop = blockIdx.x; crop < n_crops; crop += gridDim.x) {
    __syncthreads();  // the previous crop's fragments have been read (and the table / border rows are written)
    // ---- stage: thread (row y, half h) converts 48 pixels (+1) into words (col 2m-1, col 2m)
    if (tid < 192) {
      const int y = tid >> 1, h = tid & 1;
      const TIn* ip = crops + (size_t)crop * 96 * 96 + y * 96 + 48 * h;
      uint32_t* tw = tile + (y + 1) * kStemPW + 25 * h;   // half 0: words 0..24, half 1: words 25..48
      uint16_t v[49];
      if (sizeof(TIn) == 1) {
        const uint4* ip4 = reinterpret_cast<const uint4*>(ip);
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const uint4 u = __ldg(ip4 + q);
          const uint32_t wds[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int e = 0; e < 16; ++e) v[16 * q + e] = lut[(wds[e >> 2] >> (8 * (e & 3))) & 0xFFu];
        }
        v[48] = h == 0 ? lut[(uint8_t)__ldg(reinterpret_cast<const uint8_t*>(ip) + 48)] : (uint16_t)0;
      } else {
        const float4* ip4 = reinterpret_cast<const float4*>(ip);
#pragma unroll
        for (int q = 0; q < 12; ++q) {
          const float4 f = __ldg(ip4 + q);
          v[4 * q] = ptx::pack1(f.x, FMT); v[4 * q + 1] = ptx::pack1(f.y, FMT);
          v[4 * q + 2] = ptx::pack1(f.z, FMT); v[4 * q + 3] = ptx::pack1(f.w, FMT);
        }
        v[48] = h == 0 ? ptx::pack1((float)__ldg(ip + 48), FMT) : (uint16_t)0;
      }
      if (h == 0) {
        tw[0] = (uint32_t)v[0] << 16;                                        // (col -1 = 0, col 0)
#pragma unroll
        for (int m = 1; m <= 24; ++m) tw[m] = (uint32_t)v[2 * m - 1] | ((uint32_t)v[2 * m] << 16);
      } else {
#pragma unroll
        for (int m = 0; m < 24; ++m) tw[m] = (uint32_t)v[2 * m + 1] | ((uint32_t)v[2 * m + 2] << 16);   // word 25+m = (col 49+2m, col 50+2m); col 96 = 0
      }
    }
    __syncthreads();
    // ---- compute: warp w takes pooled rows w, w+8, ..., three groups of 16 pooled pixels per row
    const long long img_base = (long long)crop * HWp;
    for (int py = warp; py < 48; py += 8) {
#pragma unroll 1
      for (int g = 0; g < 3; ++g) {
        const int px0 = 16 * g;
        const uint32_t* ta = tile + (2 * py + (j >> 1)) * kStemPW + px0 + r + (j & 1);
        const uint32_t a[4] = {ta[0], ta[8], ta[2 * kStemPW], ta[2 * kStemPW + 8]};
        float best[4][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
#pragma unroll
          for (int cb = 0; cb < 4; ++cb) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            mma_16816<FMT>(acc, a, bf[q * 4 + cb]);
#pragma unroll
            for (int e = 0; e < 4; ++e) best[cb][e] = q == 0 ? acc[e] : fmaxf(best[cb][e], acc[e]);
          }
        }
        // pooled pixels px0 + r and px0 + r + 8 of row py
        const long long g0 = s2d ? img_base + (long long)py * L.Wp + (px0 >> 1) + (r >> 1) : img_base + (long long)py * L.Wp + px0 + r;
        const int pl0 = s2d ? (r & 1) * 4 : 0;
        const int second = s2d ? 4 * 4 : 8 * 4;   // words to the second pixel (4 pairs / 8 pixels further)
#pragma unroll
        for (int cb = 0; cb < 4; ++cb) {
          const uint32_t lo = ptx::pack2(fmaxf(best[cb][0] + bs[cb][0], 0.f), fmaxf(best[cb][1] + bs[cb][1], 0.f), FMT);
          const uint32_t hi = ptx::pack2(fmaxf(best[cb][2] + bs[cb][0], 0.f), fmaxf(best[cb][3] + bs[cb][1], 0.f), FMT);
          if (FMT == 0) ovf |= ptx::overflow_probe(lo) | ptx::overflow_probe(hi);
          uint32_t* op = reinterpret_cast<uint32_t*>(out + (long long)(pl0 + cb) * L.plane_px + g0) + j;
          op[0] = lo;
          op[second] = hi;   // pooled pixel + 8
        }
      }
      if (lane < (s2d ? 32 : 16))  // the pad column x = W of this row (all 4 / 8 planes)
        reinterpret_cast<uint32_t*>(out + (long long)(lane >> 2) * L.plane_px + img_base + (long long)py * L.Wp + L.W)[lane & 3] = 0u;
    }
    const int npl = s2d ? 8 : 4;
    for (int i = tid; i < L.Wp * npl * 4; i += kStemThreads)  // the pad row y = H
      reinterpret_cast<uint32_t*>(out + (long long)((i >> 2) % npl) * L.plane_px + img_base + (long long)L.H * L.Wp + i / (npl * 4))[i & 3] = 0u;
  }
  if (status != nullptr && (ovf & ptx::kOverflowProbeMask) != 0u) atomicOr(status, 2);
}
