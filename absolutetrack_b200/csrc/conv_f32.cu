// float32 NHWC implicit-GEMM convolution on CUDA cores (the reference-exact encoder/tail path,
// AT_PREC_FP32) + the fused stem (conv3x3(1->32)+BN+ReLU+maxpool2).
//
// Replaces the cuDNN/oneDNN calls behind backbone_resnet.py:56-72 (BasicBlock),
// model_utils.py:119-135 (stem, proj), :141-163 (fusion), :195-208 (head), temporal.py:31-37.
// GEMM view: M = N*Ho*Wo output pixels, N = Cout, K = kh*kw*Cin.  BatchNorm is folded into the
// weights on the host; the epilogue fuses bias + residual + ReLU and can write into a channel
// slice of a wider NHWC buffer (concat for free).
#include <stdlib.h>

#include <cuda_fp16.h>

#include "conv.cuh"

namespace abt {

namespace {

constexpr int BK = 16;
constexpr int kConvThreads = 256;

struct ConvArgs {
  const float* in;
  const float* w;     // [taps][Cin][CoutP]
  const float* bias;  // [CoutP]
  const float* res;   // optional residual, NHWC with res_ld channels per pixel
  float* out;
  int N, H, W, Cin, Ho, Wo, Cout, CoutP;
  int ksize, stride, pad;
  int in_ld, in_coff, out_ld, out_coff, res_ld;
  int relu;
  int M;  // N*Ho*Wo
};

template <int BM, int BN>
__global__ void __launch_bounds__(kConvThreads) conv_nhwc_f32_kernel(ConvArgs a) {
  constexpr int TN_THREADS = BN / 4;
  constexpr int A_F4 = BM * BK / 4 / kConvThreads;  // float4 loads of A per thread (1 or 2)
  constexpr int LDA = BM + 4;
  static_assert(BM * BN / 16 == kConvThreads, "tile/threads mismatch");
  __shared__ __align__(16) float As[2][BK][LDA];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int tx = tid % TN_THREADS, ty = tid / TN_THREADS;

  // A-load bookkeeping: this thread always loads the same pixel row(s), 4 consecutive k
  int a_row[A_F4], a_kq[A_F4], a_img[A_F4], a_oy[A_F4], a_ox[A_F4];
  bool a_ok[A_F4];
#pragma unroll
  for (int i = 0; i < A_F4; ++i) {
    const int l = tid + i * kConvThreads;
    a_row[i] = l / 4;
    a_kq[i] = l % 4;
    const int m = m0 + a_row[i];
    a_ok[i] = m < a.M;
    const int mm = a_ok[i] ? m : 0;
    a_img[i] = mm / (a.Ho * a.Wo);
    const int rem = mm - a_img[i] * (a.Ho * a.Wo);
    a_oy[i] = rem / a.Wo;
    a_ox[i] = rem - a_oy[i] * a.Wo;
  }
  const int b_row = tid / TN_THREADS, b_c4 = tid % TN_THREADS;
  const bool b_thread = b_row < BK;
  const bool vec_in = ((a.Cin | a.in_ld | a.in_coff) & 3) == 0;

  const int kchunks = (a.Cin + BK - 1) / BK;
  const int taps = a.ksize * a.ksize;
  const int steps = taps * kchunks;

  float4 ra[A_F4];
  float4 rb;

  auto load_step = [&](int step) {
    const int tap = step / kchunks;
    const int k0 = (step - tap * kchunks) * BK;
    const int r = tap / a.ksize, s = tap - r * a.ksize;
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      const int iy = a_oy[i] * a.stride + r - a.pad;
      const int ix = a_ox[i] * a.stride + s - a.pad;
      const int k = k0 + a_kq[i] * 4;
      if (a_ok[i] && (unsigned)iy < (unsigned)a.H && (unsigned)ix < (unsigned)a.W && k < a.Cin) {
        const float* p = a.in + ((int64_t)(a_img[i] * a.H + iy) * a.W + ix) * a.in_ld + a.in_coff + k;
        if (vec_in) {
          v = __ldg(reinterpret_cast<const float4*>(p));
        } else {
          v.x = __ldg(p);
          if (k + 1 < a.Cin) v.y = __ldg(p + 1);
          if (k + 2 < a.Cin) v.z = __ldg(p + 2);
          if (k + 3 < a.Cin) v.w = __ldg(p + 3);
        }
      }
      ra[i] = v;
    }
    rb = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b_thread) {
      const int k = k0 + b_row;
      const int n = n0 + b_c4 * 4;
      if (k < a.Cin && n < a.CoutP)
        rb = __ldg(reinterpret_cast<const float4*>(a.w + ((int64_t)tap * a.Cin + k) * a.CoutP + n));
    }
  };
  auto store_step = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      As[buf][a_kq[i] * 4 + 0][a_row[i]] = ra[i].x;
      As[buf][a_kq[i] * 4 + 1][a_row[i]] = ra[i].y;
      As[buf][a_kq[i] * 4 + 2][a_row[i]] = ra[i].z;
      As[buf][a_kq[i] * 4 + 3][a_row[i]] = ra[i].w;
    }
    if (b_thread) *reinterpret_cast<float4*>(&Bs[buf][b_row][b_c4 * 4]) = rb;
  };

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  load_step(0);
  store_step(0);
  __syncthreads();
  for (int step = 0; step < steps; ++step) {
    const int buf = step & 1;
    if (step + 1 < steps) load_step(step + 1);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 av = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float am[4] = {av.x, av.y, av.z, av.w};
      const float bn[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(am[i], bn[j], acc[i][j]);
    }
    if (step + 1 < steps) store_step(buf ^ 1);
    __syncthreads();
  }

  // epilogue: bias + residual + relu
  const int n = n0 + tx * 4;
  if (n >= a.Cout) return;
  float bias[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) bias[j] = (n + j < a.CoutP) ? __ldg(a.bias + n + j) : 0.f;
  const bool vec_out = ((a.out_ld | a.out_coff) & 3) == 0 && (n + 3 < a.Cout);
  const bool vec_res = a.res != nullptr && (a.res_ld & 3) == 0 && (n + 3 < a.Cout);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= a.M) continue;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = acc[i][j] + bias[j];
    if (a.res != nullptr) {
      const float* rp = a.res + (int64_t)m * a.res_ld + n;
      if (vec_res) {
        const float4 rv = __ldg(reinterpret_cast<const float4*>(rp));
        v[0] += rv.x; v[1] += rv.y; v[2] += rv.z; v[3] += rv.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n + j < a.Cout) v[j] += __ldg(rp + j);
      }
    }
    if (a.relu) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    float* op = a.out + (int64_t)m * a.out_ld + a.out_coff + n;
    if (vec_out) {
      *reinterpret_cast<float4*>(op) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (n + j < a.Cout) op[j] = v[j];
    }
  }
}


// ------------------------------------------------------------------------------------------
// Error-compensated TF32 tensor-core variant ("3xTF32") for the small 6x6 tail convolutions.
// Every fp32 operand x is split into hi = tf32(x), lo = tf32(x - hi); the product is accumulated
// as a_lo*b_hi + a_hi*b_lo + a_hi*b_hi in fp32, which keeps ~2^-21 relative accuracy per product
// (fp32-level: the tail cannot afford a 10-bit mantissa) at 1/3 of the TF32 rate.
// mma.sync.m16n8k8 (register fragments): the tail is 1 % of the network's FLOPs and its tiles are
// tiny; the tcgen05/TMEM machinery of the encoder buys nothing here.
// CTA tile BM=128 pixels x BN (64 or 96) channels, 8 warps as 4(M) x 2(N), BK = 16.
// hi = x with the low 13 mantissa bits cleared (a valid tf32), lo = x - hi (exact in fp32; the
// tensor core ignores its low 13 bits).  Two full-rate ALU ops per value instead of two cvt.rna
// on the conversion pipe; the dropped terms (a_lo*b_lo, the truncation of lo) are ~2^-20 relative.
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xFFFFE000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int BN>
__global__ void __launch_bounds__(kConvThreads) conv_nhwc_tf32x3_kernel(ConvArgs a) {
  constexpr int BM = 128;
  constexpr int LDA = BM + 8, LDB = BN + 8;      // +8: conflict-free fragment reads (bank = 8*t + g)
  constexpr int WN = BN / 2;                     // warp tile 32 x WN
  constexpr int NT = WN / 8;                     // n8 tiles per warp (4 or 6)
  constexpr int A_F4 = BM * BK / 4 / kConvThreads;  // 2
  constexpr int B_F4 = (BK * BN / 4 + kConvThreads - 1) / kConvThreads;  // 1 or 2
  __shared__ __align__(16) float As[2][BK][LDA];
  __shared__ __align__(16) float Bs[2][BK][LDB];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp & 3, wn = warp >> 2;
  const int g = lane >> 2, t = lane & 3;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;

  int a_row[A_F4], a_kq[A_F4], a_img[A_F4], a_oy[A_F4], a_ox[A_F4];
  bool a_ok[A_F4];
#pragma unroll
  for (int i = 0; i < A_F4; ++i) {
    const int l = tid + i * kConvThreads;
    a_row[i] = l / 4;
    a_kq[i] = l % 4;
    const int m = m0 + a_row[i];
    a_ok[i] = m < a.M;
    const int mm = a_ok[i] ? m : 0;
    a_img[i] = mm / (a.Ho * a.Wo);
    const int rem = mm - a_img[i] * (a.Ho * a.Wo);
    a_oy[i] = rem / a.Wo;
    a_ox[i] = rem - a_oy[i] * a.Wo;
  }
  const bool vec_in = ((a.Cin | a.in_ld | a.in_coff) & 3) == 0;
  const int kchunks = (a.Cin + BK - 1) / BK;
  const int steps = a.ksize * a.ksize * kchunks;

  float4 ra[A_F4], rb[B_F4];
  auto load_step = [&](int step) {
    const int tap = step / kchunks;
    const int k0 = (step - tap * kchunks) * BK;
    const int r = tap / a.ksize, s = tap - r * a.ksize;
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      const int iy = a_oy[i] * a.stride + r - a.pad;
      const int ix = a_ox[i] * a.stride + s - a.pad;
      const int k = k0 + a_kq[i] * 4;
      if (a_ok[i] && (unsigned)iy < (unsigned)a.H && (unsigned)ix < (unsigned)a.W && k < a.Cin) {
        const float* p = a.in + ((int64_t)(a_img[i] * a.H + iy) * a.W + ix) * a.in_ld + a.in_coff + k;
        if (vec_in) {
          v = __ldg(reinterpret_cast<const float4*>(p));
        } else {
          v.x = __ldg(p);
          if (k + 1 < a.Cin) v.y = __ldg(p + 1);
          if (k + 2 < a.Cin) v.z = __ldg(p + 2);
          if (k + 3 < a.Cin) v.w = __ldg(p + 3);
        }
      }
      ra[i] = v;
    }
#pragma unroll
    for (int i = 0; i < B_F4; ++i) {
      const int l = tid + i * kConvThreads;
      const int row = l / (BN / 4), c4 = l - row * (BN / 4);
      rb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < BK) {
        const int k = k0 + row, n = n0 + c4 * 4;
        if (k < a.Cin && n < a.CoutP)
          rb[i] = __ldg(reinterpret_cast<const float4*>(a.w + ((int64_t)tap * a.Cin + k) * a.CoutP + n));
      }
    }
  };
  auto store_step = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      As[buf][a_kq[i] * 4 + 0][a_row[i]] = ra[i].x;
      As[buf][a_kq[i] * 4 + 1][a_row[i]] = ra[i].y;
      As[buf][a_kq[i] * 4 + 2][a_row[i]] = ra[i].z;
      As[buf][a_kq[i] * 4 + 3][a_row[i]] = ra[i].w;
    }
#pragma unroll
    for (int i = 0; i < B_F4; ++i) {
      const int l = tid + i * kConvThreads;
      const int row = l / (BN / 4), c4 = l - row * (BN / 4);
      if (row < BK) *reinterpret_cast<float4*>(&Bs[buf][row][c4 * 4]) = rb[i];
    }
  };

  float acc[2][NT][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;

  load_step(0);
  store_step(0);
  __syncthreads();
  for (int step = 0; step < steps; ++step) {
    const int buf = step & 1;
    if (step + 1 < steps) load_step(step + 1);
#pragma unroll
    for (int kk = 0; kk < BK; kk += 8) {
      uint32_t ah[2][4], al[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int mr = wm * 32 + i * 16 + g;
        const float v[4] = {As[buf][kk + t][mr], As[buf][kk + t][mr + 8], As[buf][kk + t + 4][mr], As[buf][kk + t + 4][mr + 8]};
#pragma unroll
        for (int e = 0; e < 4; ++e) split_tf32(v[e], ah[i][e], al[i][e]);
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const int nc = wn * WN + j * 8 + g;
        const float v[2] = {Bs[buf][kk + t][nc], Bs[buf][kk + t + 4][nc]};
        uint32_t bh[2], bl[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) split_tf32(v[e], bh[e], bl[e]);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          mma_tf32(acc[i][j], al[i], bh);
          mma_tf32(acc[i][j], ah[i], bl);
          mma_tf32(acc[i][j], ah[i], bh);
        }
      }
    }
    if (step + 1 < steps) store_step(buf ^ 1);
    __syncthreads();
  }

  // epilogue: bias + residual + relu; fragment (i,j): rows g / g+8, columns 2t, 2t+1
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int n = n0 + wn * WN + j * 8 + 2 * t;
    if (n >= a.Cout) continue;
    const bool two = n + 1 < a.Cout;
    const float b0 = __ldg(a.bias + n), b1 = two ? __ldg(a.bias + n + 1) : 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int m = m0 + wm * 32 + i * 16 + g + h * 8;
        if (m >= a.M) continue;
        float v0 = acc[i][j][h * 2] + b0, v1 = acc[i][j][h * 2 + 1] + b1;
        if (a.res != nullptr) {
          const float* rp = a.res + (int64_t)m * a.res_ld + n;
          v0 += __ldg(rp);
          if (two) v1 += __ldg(rp + 1);
        }
        if (a.relu) {
          v0 = fmaxf(v0, 0.f);
          v1 = fmaxf(v1, 0.f);
        }
        float* op = a.out + (int64_t)m * a.out_ld + a.out_coff + n;
        if (two && (((a.out_ld | a.out_coff) & 1) == 0)) {
          *reinterpret_cast<float2*>(op) = make_float2(v0, v1);
        } else {
          op[0] = v0;
          if (two) op[1] = v1;
        }
      }
  }
}

// ---- error-compensated fp16 ("fp16 x 3") implicit GEMM ------------------------------------------
// x = hi + lo with hi = fp16(x), lo = fp16(x - hi): 22 significant bits, the same as the TF32 split
// above (both formats carry a 10-bit mantissa), but mma.m16n8k16 covers twice the K of m16n8k8 per
// instruction: half the tensor instructions for the same product hi*hi + lo*hi + hi*lo.
// Operands are split ONCE -- activations when they are staged into shared memory, weights at model
// load (conv_f16x3_split_weights: [part][tap][n][k] halves, k contiguous, scaled by 2^kWScaleLog2 so
// that the lo part of small folded-BN weights stays out of the fp16 denormals) -- and the fragments
// come from ldmatrix.  Tiles as above: 128 x BN x 16, 8 warps as 4 x 2.
constexpr int kWScaleLog2 = 8;
constexpr int LDK = BK + 8;  // halves per smem row: 48 B, conflict-free for ldmatrix (16-byte chunk index 3r mod 8)

__device__ __forceinline__ float clamp_h(float x) { return fminf(fmaxf(x, -65504.f), 65504.f); }
__device__ __forceinline__ void split_h2(float x, float y, __half2& hi, __half2& lo) {
  x = clamp_h(x); y = clamp_h(y);
  hi = __floats2half2_rn(x, y);
  const float2 hf = __half22float2(hi);
  lo = __floats2half2_rn(x - hf.x, y - hf.y);
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_ptr);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma_f16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void split_weights_kernel(const float* __restrict__ w, __half* __restrict__ dst, int taps, int Cin, int Cout,
                                     int CoutP, int Np, int Kp) {
  const int64_t total = (int64_t)taps * Np * Kp;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int k = (int)(i % Kp);
  const int n = (int)((i / Kp) % Np);
  const int tap = (int)(i / ((int64_t)Kp * Np));
  float v = 0.f;
  if (k < Cin && n < Cout) v = w[((int64_t)tap * Cin + k) * CoutP + n] * (float)(1 << kWScaleLog2);
  v = clamp_h(v);
  const __half hi = __float2half_rn(v);
  dst[i] = hi;
  dst[total + i] = __float2half_rn(v - __half2float(hi));
}

struct ConvArgsH {
  ConvArgs a;
  const __half* wsplit;  // [2][taps][Np][Kp]
  int Np, Kp;
};

template <int BN>
__global__ void __launch_bounds__(kConvThreads) conv_nhwc_f16x3_kernel(ConvArgsH ah_) {
  // programmatic dependent launch: the index set-up and the first weight tile (constants) run while the previous kernel
  // of the stream drains; its output - this conv's input - is first read behind griddepcontrol.wait below
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const ConvArgs& a = ah_.a;
  constexpr int BM = 128;
  constexpr int WN = BN / 2;   // warp tile 32 x WN
  constexpr int NT = WN / 8;   // n8 tiles per warp (4 or 6)
  constexpr int A_F4 = BM * BK / 4 / kConvThreads;                 // 2
  constexpr int B_V = (2 * BN * 2 + kConvThreads - 1) / kConvThreads;  // 16-byte loads of B per thread (1 or 2)
  __shared__ __align__(16) __half As[2][2][BM][LDK];  // [buffer][hi/lo][row][k]
  __shared__ __align__(16) __half Bs[2][2][BN][LDK];  // [buffer][hi/lo][n][k]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp & 3, wn = warp >> 2;
  const int g = lane >> 2, t = lane & 3;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int taps = a.ksize * a.ksize;

  int a_row[A_F4], a_kq[A_F4], a_img[A_F4], a_oy[A_F4], a_ox[A_F4];
  bool a_ok[A_F4];
#pragma unroll
  for (int i = 0; i < A_F4; ++i) {
    const int l = tid + i * kConvThreads;
    a_row[i] = l / 4;
    a_kq[i] = l % 4;
    const int m = m0 + a_row[i];
    a_ok[i] = m < a.M;
    const int mm = a_ok[i] ? m : 0;
    a_img[i] = mm / (a.Ho * a.Wo);
    const int rem = mm - a_img[i] * (a.Ho * a.Wo);
    a_oy[i] = rem / a.Wo;
    a_ox[i] = rem - a_oy[i] * a.Wo;
  }
  const bool vec_in = ((a.Cin | a.in_ld | a.in_coff) & 3) == 0;
  const int kchunks = ah_.Kp / BK;
  const int steps = taps * kchunks;

  float4 ra[A_F4];
  uint4 rb[B_V];
  auto load_b = [&](int step) {
    const int tap = step / kchunks;
    const int k0 = (step - tap * kchunks) * BK;
#pragma unroll
    for (int i = 0; i < B_V; ++i) {
      const int l = tid + i * kConvThreads;
      rb[i] = make_uint4(0u, 0u, 0u, 0u);
      if (l < 2 * BN * 2) {
        const int part = l / (BN * 2), rem = l - part * (BN * 2);
        const int n = rem >> 1, h = rem & 1;
        rb[i] = __ldg(reinterpret_cast<const uint4*>(
            ah_.wsplit + (((int64_t)part * taps + tap) * ah_.Np + n0 + n) * ah_.Kp + k0 + h * 8));
      }
    }
  };
  auto load_a = [&](int step) {
    const int tap = step / kchunks;
    const int k0 = (step - tap * kchunks) * BK;
    const int r = tap / a.ksize, s = tap - r * a.ksize;
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      const int iy = a_oy[i] * a.stride + r - a.pad;
      const int ix = a_ox[i] * a.stride + s - a.pad;
      const int k = k0 + a_kq[i] * 4;
      if (a_ok[i] && (unsigned)iy < (unsigned)a.H && (unsigned)ix < (unsigned)a.W && k < a.Cin) {
        const float* p = a.in + ((int64_t)(a_img[i] * a.H + iy) * a.W + ix) * a.in_ld + a.in_coff + k;
        if (vec_in) {
          v = __ldg(reinterpret_cast<const float4*>(p));
        } else {
          v.x = __ldg(p);
          if (k + 1 < a.Cin) v.y = __ldg(p + 1);
          if (k + 2 < a.Cin) v.z = __ldg(p + 2);
          if (k + 3 < a.Cin) v.w = __ldg(p + 3);
        }
      }
      ra[i] = v;
    }
  };
  auto load_step = [&](int step) { load_a(step); load_b(step); };
  auto store_step = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      __half2 h0, l0, h1, l1;
      split_h2(ra[i].x, ra[i].y, h0, l0);
      split_h2(ra[i].z, ra[i].w, h1, l1);
      uint2 hv, lv;
      hv.x = *reinterpret_cast<uint32_t*>(&h0); hv.y = *reinterpret_cast<uint32_t*>(&h1);
      lv.x = *reinterpret_cast<uint32_t*>(&l0); lv.y = *reinterpret_cast<uint32_t*>(&l1);
      *reinterpret_cast<uint2*>(&As[buf][0][a_row[i]][a_kq[i] * 4]) = hv;
      *reinterpret_cast<uint2*>(&As[buf][1][a_row[i]][a_kq[i] * 4]) = lv;
    }
#pragma unroll
    for (int i = 0; i < B_V; ++i) {
      const int l = tid + i * kConvThreads;
      if (l < 2 * BN * 2) {
        const int part = l / (BN * 2), rem = l - part * (BN * 2);
        *reinterpret_cast<uint4*>(&Bs[buf][part][rem >> 1][(rem & 1) * 8]) = rb[i];
      }
    }
  };

  float acc[2][NT][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;

  // ldmatrix lane addressing: A x4 = (rows 0-7,k0) (rows 8-15,k0) (rows 0-7,k8) (rows 8-15,k8);
  // B x4 = (n 0-7,k0) (n 0-7,k8) (n 8-15,k0) (n 8-15,k8)
  const int a_lrow = (lane & 7) + ((lane >> 3) & 1) * 8, a_lk = (lane >> 4) * 8;
  const int b_lrow = (lane & 7) + (lane >> 4) * 8, b_lk = ((lane >> 3) & 1) * 8;

  load_b(0);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  load_a(0);
  store_step(0);
  __syncthreads();
  for (int step = 0; step < steps; ++step) {
    const int buf = step & 1;
    if (step + 1 < steps) load_step(step + 1);
    uint32_t fa[2][2][4];  // [m tile][hi/lo][4]
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int p = 0; p < 2; ++p) ldmatrix_x4(fa[i][p], &As[buf][p][wm * 32 + i * 16 + a_lrow][a_lk]);
#pragma unroll
    for (int jj = 0; jj < NT / 2; ++jj) {
      uint32_t fb[2][4];  // [hi/lo][b0,b1 of tile 2jj, b0,b1 of tile 2jj+1]
#pragma unroll
      for (int p = 0; p < 2; ++p) ldmatrix_x4(fb[p], &Bs[buf][p][wn * WN + jj * 16 + b_lrow][b_lk]);
#pragma unroll
      for (int jh = 0; jh < 2; ++jh)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          float(&d)[4] = acc[i][jj * 2 + jh];
          mma_f16(d, fa[i][1], fb[0][jh * 2], fb[0][jh * 2 + 1]);  // lo * hi
          mma_f16(d, fa[i][0], fb[1][jh * 2], fb[1][jh * 2 + 1]);  // hi * lo
          mma_f16(d, fa[i][0], fb[0][jh * 2], fb[0][jh * 2 + 1]);  // hi * hi
        }
    }
    if (step + 1 < steps) store_step(buf ^ 1);
    __syncthreads();
  }

  // epilogue: undo the weight scale, bias + residual + relu; fragment (i,j): rows g / g+8, columns 2t, 2t+1
  constexpr float kUnscale = 1.0f / (float)(1 << kWScaleLog2);
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int n = n0 + wn * WN + j * 8 + 2 * t;
    if (n >= a.Cout) continue;
    const bool two = n + 1 < a.Cout;
    const float b0 = __ldg(a.bias + n), b1 = two ? __ldg(a.bias + n + 1) : 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int m = m0 + wm * 32 + i * 16 + g + h * 8;
        if (m >= a.M) continue;
        float v0 = fmaf(acc[i][j][h * 2], kUnscale, b0), v1 = fmaf(acc[i][j][h * 2 + 1], kUnscale, b1);
        if (a.res != nullptr) {
          const float* rp = a.res + (int64_t)m * a.res_ld + n;
          v0 += __ldg(rp);
          if (two) v1 += __ldg(rp + 1);
        }
        if (a.relu) {
          v0 = fmaxf(v0, 0.f);
          v1 = fmaxf(v1, 0.f);
        }
        float* op = a.out + (int64_t)m * a.out_ld + a.out_coff + n;
        if (two && (((a.out_ld | a.out_coff) & 1) == 0)) {
          *reinterpret_cast<float2*>(op) = make_float2(v0, v1);
        } else {
          op[0] = v0;
          if (two) op[1] = v1;
        }
      }
  }
}

// Stem: conv3x3(1->32, pad 1, bias) + BN (folded) + ReLU + maxpool 2x2   (model_utils.py:119-124)
// in: crops [c,96,96] (u8 grey levels -> /255, or f32); out NHWC [c,48,48,32] f32.
// One thread = one pooled pixel x 8 output channels (4 lanes cover the 32 channels).
template <typename TIn>
__global__ void __launch_bounds__(256) stem_kernel(const TIn* __restrict__ crops, const float* __restrict__ w,
                                                   const float* __restrict__ bias, float* __restrict__ out, int c,
                                                   int H, int W) {
  __shared__ float sw[9 * 32];
  __shared__ float sb[32];
  for (int i = threadIdx.x; i < 9 * 32; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < 32) sb[threadIdx.x] = bias[threadIdx.x];
  __syncthreads();
  const int Hp = H / 2, Wp = W / 2;
  const int64_t total = (int64_t)c * Hp * Wp * 4;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int cg = (int)(gid & 3);  // channel group of 8
  const int64_t pix = gid >> 2;
  const int px = (int)(pix % Wp);
  const int py = (int)((pix / Wp) % Hp);
  const int64_t img = pix / ((int64_t)Wp * Hp);
  const TIn* ip = crops + img * H * W;
  float patch[4][4];
#pragma unroll
  for (int dy = 0; dy < 4; ++dy)
#pragma unroll
    for (int dx = 0; dx < 4; ++dx) {
      const int y = 2 * py - 1 + dy, x = 2 * px - 1 + dx;
      float v = 0.f;
      if ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) {
        if (sizeof(TIn) == 1)
          v = __fdiv_rn((float)ip[y * W + x], 255.0f);
        else
          v = (float)ip[y * W + x];
      }
      patch[dy][dx] = v;
    }
  float res[8];
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) {
    const int co = cg * 8 + ch;
    float best = 0.f;  // ReLU then max == max(0, max of the four)
#pragma unroll
    for (int oy = 0; oy < 2; ++oy)
#pragma unroll
      for (int ox = 0; ox < 2; ++ox) {
        float s = sb[co];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int q = 0; q < 3; ++q) s = fmaf(patch[oy + r][ox + q], sw[(r * 3 + q) * 32 + co], s);
        best = fmaxf(best, s);
      }
    res[ch] = best;
  }
  float4* op = reinterpret_cast<float4*>(out + pix * 32 + cg * 8);
  op[0] = make_float4(res[0], res[1], res[2], res[3]);
  op[1] = make_float4(res[4], res[5], res[6], res[7]);
}

}  // namespace

int conv_nhwc_f32(const ConvLayer& L, const float* blob, const float* in, int in_ld, int in_coff, const float* res,
                  int res_ld, float* out, int out_ld, int out_coff, int N, int H, int W, bool relu,
                  cudaStream_t stream) {
  ConvArgs a;
  a.in = in;
  a.w = blob + L.weight_offset;
  a.bias = blob + L.bias_offset;
  a.res = res;
  a.out = out;
  a.N = N; a.H = H; a.W = W; a.Cin = L.cin;
  a.ksize = L.ksize; a.stride = L.stride; a.pad = L.ksize / 2;
  a.Ho = (H + 2 * a.pad - L.ksize) / L.stride + 1;
  a.Wo = (W + 2 * a.pad - L.ksize) / L.stride + 1;
  a.Cout = L.cout;
  a.CoutP = (int)round_up(L.cout, 4);
  a.in_ld = in_ld; a.in_coff = in_coff; a.out_ld = out_ld; a.out_coff = out_coff; a.res_ld = res_ld;
  a.relu = relu ? 1 : 0;
  const int64_t M = (int64_t)N * a.Ho * a.Wo;
  if (M == 0) return AT_OK;
  if (M > 0x7fffffffLL) {
    set_error("conv_nhwc_f32: M too large");
    return AT_ERR_INVALID_ARGUMENT;
  }
  a.M = (int)M;
  if (L.cout <= 32) {
    dim3 grid((unsigned)((M + 127) / 128), (unsigned)((L.cout + 31) / 32));
    conv_nhwc_f32_kernel<128, 32><<<grid, kConvThreads, 0, stream>>>(a);
  } else {
    dim3 grid((unsigned)((M + 63) / 64), (unsigned)((L.cout + 63) / 64));
    conv_nhwc_f32_kernel<64, 64><<<grid, kConvThreads, 0, stream>>>(a);
  }
  return check_launch("conv_nhwc_f32_kernel");
}

// Same contract as conv_nhwc_f32, on the tensor cores with error-compensated TF32 (fp32-level
// accuracy).  Used by the tail (csrc/tail.cu); AT_TAIL_SIMT=1 routes back to the CUDA-core kernel.
int conv_nhwc_tf32x3(const ConvLayer& L, const float* blob, const float* in, int in_ld, int in_coff, const float* res,
                     int res_ld, float* out, int out_ld, int out_coff, int N, int H, int W, bool relu,
                     cudaStream_t stream) {
  static const bool simt = getenv("AT_TAIL_SIMT") != nullptr && atoi(getenv("AT_TAIL_SIMT")) != 0;
  if (simt) return conv_nhwc_f32(L, blob, in, in_ld, in_coff, res, res_ld, out, out_ld, out_coff, N, H, W, relu, stream);
  ConvArgs a;
  a.in = in;
  a.w = blob + L.weight_offset;
  a.bias = blob + L.bias_offset;
  a.res = res;
  a.out = out;
  a.N = N; a.H = H; a.W = W; a.Cin = L.cin;
  a.ksize = L.ksize; a.stride = L.stride; a.pad = L.ksize / 2;
  a.Ho = (H + 2 * a.pad - L.ksize) / L.stride + 1;
  a.Wo = (W + 2 * a.pad - L.ksize) / L.stride + 1;
  a.Cout = L.cout;
  a.CoutP = (int)round_up(L.cout, 4);
  a.in_ld = in_ld; a.in_coff = in_coff; a.out_ld = out_ld; a.out_coff = out_coff; a.res_ld = res_ld;
  a.relu = relu ? 1 : 0;
  const int64_t M = (int64_t)N * a.Ho * a.Wo;
  if (M == 0) return AT_OK;
  if (M > 0x7fffffffLL) {
    set_error("conv_nhwc_tf32x3: M too large");
    return AT_ERR_INVALID_ARGUMENT;
  }
  a.M = (int)M;
  // 96-wide tiles for Cout in (64, 96]; otherwise 64-wide tiles
  if (L.cout > 64 && L.cout <= 96) {
    dim3 grid((unsigned)((M + 127) / 128), 1);
    conv_nhwc_tf32x3_kernel<96><<<grid, kConvThreads, 0, stream>>>(a);
  } else {
    dim3 grid((unsigned)((M + 127) / 128), (unsigned)((L.cout + 63) / 64));
    conv_nhwc_tf32x3_kernel<64><<<grid, kConvThreads, 0, stream>>>(a);
  }
  return check_launch("conv_nhwc_tf32x3_kernel");
}

namespace {
inline int f16x3_bn(const ConvLayer& L) { return (L.cout > 64 && L.cout <= 96) ? 96 : 64; }
inline int f16x3_np(const ConvLayer& L) { return (int)round_up(L.cout, f16x3_bn(L)); }
inline int f16x3_kp(const ConvLayer& L) { return (int)round_up(L.cin, BK); }
}  // namespace

int64_t conv_f16x3_weight_halves(const ConvLayer& L) {
  return 2ll * L.ksize * L.ksize * f16x3_np(L) * f16x3_kp(L);
}

int conv_f16x3_split_weights(const ConvLayer& L, const float* blob, void* dst, cudaStream_t stream) {
  const int taps = L.ksize * L.ksize, Np = f16x3_np(L), Kp = f16x3_kp(L);
  const int64_t total = (int64_t)taps * Np * Kp;
  split_weights_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
      blob + L.weight_offset, (__half*)dst, taps, L.cin, L.cout, (int)round_up(L.cout, 4), Np, Kp);
  return check_launch("split_weights_kernel");
}

int conv_nhwc_f16x3(const ConvLayer& L, const float* blob, const void* wsplit, const float* in, int in_ld, int in_coff,
                    const float* res, int res_ld, float* out, int out_ld, int out_coff, int N, int H, int W, bool relu,
                    cudaStream_t stream) {
  ConvArgsH h;
  ConvArgs& a = h.a;
  a.in = in;
  a.w = blob + L.weight_offset;
  a.bias = blob + L.bias_offset;
  a.res = res;
  a.out = out;
  a.N = N; a.H = H; a.W = W; a.Cin = L.cin;
  a.ksize = L.ksize; a.stride = L.stride; a.pad = L.ksize / 2;
  a.Ho = (H + 2 * a.pad - L.ksize) / L.stride + 1;
  a.Wo = (W + 2 * a.pad - L.ksize) / L.stride + 1;
  a.Cout = L.cout;
  a.CoutP = (int)round_up(L.cout, 4);
  a.in_ld = in_ld; a.in_coff = in_coff; a.out_ld = out_ld; a.out_coff = out_coff; a.res_ld = res_ld;
  a.relu = relu ? 1 : 0;
  h.wsplit = (const __half*)wsplit;
  h.Np = f16x3_np(L);
  h.Kp = f16x3_kp(L);
  const int64_t M = (int64_t)N * a.Ho * a.Wo;
  if (M == 0) return AT_OK;
  if (M > 0x7fffffffLL) {
    set_error("conv_nhwc_f16x3: M too large");
    return AT_ERR_INVALID_ARGUMENT;
  }
  a.M = (int)M;
  if (f16x3_bn(L) == 96) {
    dim3 grid((unsigned)((M + 127) / 128), (unsigned)(h.Np / 96));
    launch_pdl(conv_nhwc_f16x3_kernel<96>, dim3(grid), dim3(kConvThreads), 0, stream, h);
  } else {
    dim3 grid((unsigned)((M + 127) / 128), (unsigned)(h.Np / 64));
    launch_pdl(conv_nhwc_f16x3_kernel<64>, dim3(grid), dim3(kConvThreads), 0, stream, h);
  }
  return check_launch("conv_nhwc_f16x3_kernel");
}

int stem_f32(const ConvLayer& L, const float* blob, const void* crops, int crop_dtype, float* out, int c,
             cudaStream_t stream) {
  const int H = 96, W = 96;
  const int64_t total = (int64_t)c * (H / 2) * (W / 2) * 4;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (crop_dtype == AT_U8)
    stem_kernel<uint8_t><<<blocks, 256, 0, stream>>>((const uint8_t*)crops, blob + L.weight_offset,
                                                      blob + L.bias_offset, out, c, H, W);
  else
    stem_kernel<float><<<blocks, 256, 0, stream>>>((const float*)crops, blob + L.weight_offset, blob + L.bias_offset,
                                                    out, c, H, W);
  return check_launch("stem_kernel");
}

}  // namespace abt
