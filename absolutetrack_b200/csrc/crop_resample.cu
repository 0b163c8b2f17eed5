// Perspective-crop resampler: fused  dst-pixel -> fisheye source map  +  OpenCV fixed-point
// bilinear gather  +  /255 and dtype convert.   Replaces tracker._warp_image
// (lib/tracker/tracker.py:61-89) and the /255 of tracker.py:332; and, second kernel, the
// torch_data homography resampler (lib/batched_dataset/data_transform.py:78-144).
//
// Roofline: HBM-bound by design (source bounding box bytes + 96*96*sizeof(out) per crop).
// One CTA handles a quarter of one crop (24 rows), threads walk dst pixels in raster order so
// output stores are fully coalesced and the gathered source taps of a warp are neighbours
// (they hit the same few 128B lines through L1).
#include <stdlib.h>

#include "at_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kSplit = 4;  // CTAs per crop

template <typename T>
struct MapParams {
  T M[9], d[3], fdx, fdy, cdx, cdy, fsx, fsy, csx, csy, k1, k2, k3, k4, p1, p2, k5, k6;
};

template <typename T>
__device__ __forceinline__ T t_sqrt(T x);
template <>
__device__ __forceinline__ float t_sqrt<float>(float x) { return sqrtf(x); }
template <>
__device__ __forceinline__ double t_sqrt<double>(double x) { return sqrt(x); }
template <typename T>
__device__ __forceinline__ T t_atan2(T y, T x);
template <>
__device__ __forceinline__ float t_atan2<float>(float y, float x) { return atan2f(y, x); }
template <>
__device__ __forceinline__ double t_atan2<double>(double y, double x) { return atan2(y, x); }

// Map one dst pixel to float32 source window coordinates, in the reference's operation order:
// camera.py:320-329,69-75 (pinhole unproject + normalise, eps 5.43e-20), :296-306 (rotate),
// :80-85 (arctan projection, eps 2^-128), :122-143 (Fisheye62), :308-312, tracker.py:78-82.
template <typename T>
__device__ __forceinline__ void map_pixel(const MapParams<T>& P, int x, int y, float& wx, float& wy) {
  const T qx = (T(x) - P.cdx) / P.fdx;
  const T qy = (T(y) - P.cdy) / P.fdy;
  T nrm = t_sqrt(qx * qx + qy * qy + T(1));
  nrm = nrm > T(5.43e-20) ? nrm : T(5.43e-20);
  const T vx = qx / nrm, vy = qy / nrm, vz = T(1) / nrm;
  const T ex = P.M[0] * vx + P.M[1] * vy + P.M[2] * vz + P.d[0];
  const T ey = P.M[3] * vx + P.M[4] * vy + P.M[5] * vz + P.d[1];
  const T ez = P.M[6] * vx + P.M[7] * vy + P.M[8] * vz + P.d[2];
  if (ez < T(0)) {
    wx = -1.0f;
    wy = -1.0f;
    return;
  }
  const T r = t_sqrt(ex * ex + ey * ey);
  const T eps = T(2.938735877055719e-39);  // 2^-128
  const T s = t_atan2(r, ez) / (r > eps ? r : eps);
  const T px = ex * s, py = ey * s;
  const T pi2 = T(9.869604401089358);
  T r2 = px * px + py * py;
  r2 = r2 > pi2 ? pi2 : r2;
  const T r4 = r2 * r2, r6 = r2 * r4, r8 = r4 * r4, r10 = r4 * r6, r12 = r6 * r6;
  const T radial = T(1) + P.k1 * r2 + P.k2 * r4 + P.k3 * r6 + P.k4 * r8 + P.k5 * r10 + P.k6 * r12;
  const T ux = px * radial, uy = py * radial;
  const T x2 = ux * ux, y2 = uy * uy, xy = ux * uy, rr = x2 + y2;
  const T dx = ux + (T(2) * P.p2 * xy + P.p1 * (rr + T(2) * x2));
  const T dy = uy + (T(2) * P.p1 * xy + P.p2 * (rr + T(2) * y2));
  wx = (float)(dx * P.fsx + P.csx);
  wy = (float)(dy * P.fsy + P.csy);
}

// cv2.remap(INTER_LINEAR, BORDER_CONSTANT=0) on uint8: 1/32-px coordinates, 15-bit weights.
__device__ __forceinline__ int remap_fixed_u8(const uint8_t* __restrict__ img, int64_t row_stride, int W, int H,
                                              float wx, float wy) {
  const int sx = __float2int_rn(wx * 32.0f);
  const int sy = __float2int_rn(wy * 32.0f);
  const int ix = min(max(sx >> 5, -32768), 32767);
  const int iy = min(max(sy >> 5, -32768), 32767);
  const int fx = sx & 31, fy = sy & 31;
  const bool x0 = (unsigned)ix < (unsigned)W, x1 = (unsigned)(ix + 1) < (unsigned)W;
  const bool y0 = (unsigned)iy < (unsigned)H, y1 = (unsigned)(iy + 1) < (unsigned)H;
  const uint8_t* r0 = img + (int64_t)iy * row_stride + ix;
  const uint8_t* r1 = r0 + row_stride;
  const int t00 = (y0 && x0) ? __ldg(r0) : 0;
  const int t01 = (y0 && x1) ? __ldg(r0 + 1) : 0;
  const int t10 = (y1 && x0) ? __ldg(r1) : 0;
  const int t11 = (y1 && x1) ? __ldg(r1 + 1) : 0;
  const int acc = ((32 - fy) * (32 - fx) * t00 + (32 - fy) * fx * t01 + fy * (32 - fx) * t10 + fy * fx * t11) * 32;
  return (acc + 16384) >> 15;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
crop_resample_kernel(const uint8_t* __restrict__ src, int64_t frame_stride, int64_t view_stride, int64_t row_stride,
                     const AtCropParams* __restrict__ params, int dst_w, int dst_h, void* __restrict__ out,
                     int out_dtype) {
  __shared__ MapParams<T> P;
  __shared__ int s_meta[4];
  const int crop = blockIdx.x;
  const AtCropParams& g = params[crop];
  if (threadIdx.x < 9) P.M[threadIdx.x] = (T)g.M[threadIdx.x];
  if (threadIdx.x < 3) P.d[threadIdx.x] = (T)g.d[threadIdx.x];
  if (threadIdx.x == 32) {
    P.fdx = (T)g.f_dst[0]; P.fdy = (T)g.f_dst[1]; P.cdx = (T)g.c_dst[0]; P.cdy = (T)g.c_dst[1];
    P.fsx = (T)g.f_src[0]; P.fsy = (T)g.f_src[1]; P.csx = (T)g.c_src[0]; P.csy = (T)g.c_src[1];
    P.k1 = (T)g.k[0]; P.k2 = (T)g.k[1]; P.k3 = (T)g.k[2]; P.k4 = (T)g.k[3];
    P.p1 = (T)g.k[4]; P.p2 = (T)g.k[5]; P.k5 = (T)g.k[6]; P.k6 = (T)g.k[7];
    s_meta[0] = g.frame; s_meta[1] = g.view; s_meta[2] = g.src_w; s_meta[3] = g.src_h;
  }
  __syncthreads();
  const int W = s_meta[2], H = s_meta[3];
  const uint8_t* img = src + (int64_t)s_meta[0] * frame_stride + (int64_t)s_meta[1] * view_stride;
  const int npix = dst_w * dst_h;
  const int per = (npix + kSplit - 1) / kSplit;
  const int p_begin = blockIdx.y * per;
  const int p_end = min(npix, p_begin + per);
  const int64_t obase = (int64_t)crop * npix;
  for (int p = p_begin + threadIdx.x; p < p_end; p += kThreads) {
    const int y = p / dst_w, x = p - y * dst_w;
    float wx, wy;
    map_pixel<T>(P, x, y, wx, wy);
    const int v = remap_fixed_u8(img, row_stride, W, H, wx, wy);
    if (out_dtype == AT_U8) {
      ((uint8_t*)out)[obase + p] = (uint8_t)v;
    } else if (out_dtype == AT_F32) {
      ((float*)out)[obase + p] = __fdiv_rn((float)v, 255.0f);
    } else {
      ((__nv_bfloat16*)out)[obase + p] = __float2bfloat16_rn(__fdiv_rn((float)v, 255.0f));
    }
  }
}


// ------------------------------------------------------------------------------------------
// Fast fp32 map (AT_MAP_FP32): same formulas, but reciprocal / rsqrt / polynomial-atan instead of
// IEEE divisions, sqrtf and atan2f, constants folded per crop, 4 consecutive dst pixels per thread
// (independent chains hide the gather latency; one 4-pixel store).  Map error vs the fp64 reference
// stays ~1e-4 px (tests: <= 1 grey level on smooth images).
struct FastParams {
  float M[9], d[3];
  float ifdx, ifdy, qx0, qy0;  // q = x*ifd + q0  ==  (x - c_dst) / f_dst
  float fsx, fsy, csx, csy, k1, k2, k3, k4, p1, p2, k5, k6;
};

// atan(a) for a in [0,1]: Abramowitz-Stegun 4.4.49 (|err| <= 2e-8)
__device__ __forceinline__ float atan01(float a) {
  const float z = a * a;
  float p = 0.0028662257f;
  p = fmaf(p, z, -0.0161657367f);
  p = fmaf(p, z, 0.0429096138f);
  p = fmaf(p, z, -0.0752896400f);
  p = fmaf(p, z, 0.1065626393f);
  p = fmaf(p, z, -0.1420889944f);
  p = fmaf(p, z, 0.1999355085f);
  p = fmaf(p, z, -0.3333314528f);
  return fmaf(p * z, a, a);
}

__device__ __forceinline__ void map_pixel_fast(const FastParams& P, float qx, float qy, float& wx, float& wy) {
  const float inv = rsqrtf(fmaf(qx, qx, fmaf(qy, qy, 1.0f)));
  const float vx = qx * inv, vy = qy * inv, vz = inv;
  const float ex = fmaf(P.M[0], vx, fmaf(P.M[1], vy, fmaf(P.M[2], vz, P.d[0])));
  const float ey = fmaf(P.M[3], vx, fmaf(P.M[4], vy, fmaf(P.M[5], vz, P.d[1])));
  const float ez = fmaf(P.M[6], vx, fmaf(P.M[7], vy, fmaf(P.M[8], vz, P.d[2])));
  const float r2e = fmaf(ex, ex, ey * ey);
  const float rinv = rsqrtf(fmaxf(r2e, 1e-30f));
  const float r = r2e * rinv;
  // theta = atan2(r, ez) with r, ez >= 0
  const float ezp = fmaxf(ez, 0.0f);  // ez < 0 is overridden below (branch-free: 4 chains interleave)
  const float mn = fminf(r, ezp), mx = fmaxf(r, ezp);
  float th = atan01(__fdividef(mn, fmaxf(mx, 1e-30f)));
  th = r > ezp ? 1.57079632679f - th : th;
  const float s = th * rinv;  // camera.py:80-85: atan2(r, z) / max(r, eps)
  const float px = ex * s, py = ey * s;
  float r2 = fmaf(px, px, py * py);
  r2 = fminf(r2, 9.869604401089358f);
  float radial = P.k6;
  radial = fmaf(radial, r2, P.k5);
  radial = fmaf(radial, r2, P.k4);
  radial = fmaf(radial, r2, P.k3);
  radial = fmaf(radial, r2, P.k2);
  radial = fmaf(radial, r2, P.k1);
  radial = fmaf(radial, r2, 1.0f);
  const float ux = px * radial, uy = py * radial;
  const float x2 = ux * ux, y2 = uy * uy, xy = ux * uy, rr = x2 + y2;
  const float dx = ux + fmaf(2.0f * P.p2, xy, P.p1 * fmaf(2.0f, x2, rr));
  const float dy = uy + fmaf(2.0f * P.p1, xy, P.p2 * fmaf(2.0f, y2, rr));
  wx = ez < 0.0f ? -1.0f : fmaf(dx, P.fsx, P.csx);  // tracker.py:78-80
  wy = ez < 0.0f ? -1.0f : fmaf(dy, P.fsy, P.csy);
}

constexpr int kFastPx = 4;  // dst pixels per thread

__global__ void __launch_bounds__(kThreads)
crop_resample_fast_kernel(const uint8_t* __restrict__ src, int64_t frame_stride, int64_t view_stride, int64_t row_stride,
                          const AtCropParams* __restrict__ params, int dst_w, int dst_h, void* __restrict__ out,
                          int out_dtype) {
  __shared__ FastParams P;
  __shared__ int s_meta[4];
  const int crop = blockIdx.x;
  const AtCropParams& g = params[crop];
  if (threadIdx.x < 9) P.M[threadIdx.x] = (float)g.M[threadIdx.x];
  if (threadIdx.x < 3) P.d[threadIdx.x] = (float)g.d[threadIdx.x];
  if (threadIdx.x == 32) {
    P.ifdx = 1.0f / (float)g.f_dst[0]; P.ifdy = 1.0f / (float)g.f_dst[1];
    P.qx0 = -(float)g.c_dst[0] * P.ifdx; P.qy0 = -(float)g.c_dst[1] * P.ifdy;
    P.fsx = (float)g.f_src[0]; P.fsy = (float)g.f_src[1]; P.csx = (float)g.c_src[0]; P.csy = (float)g.c_src[1];
    P.k1 = (float)g.k[0]; P.k2 = (float)g.k[1]; P.k3 = (float)g.k[2]; P.k4 = (float)g.k[3];
    P.p1 = (float)g.k[4]; P.p2 = (float)g.k[5]; P.k5 = (float)g.k[6]; P.k6 = (float)g.k[7];
    s_meta[0] = g.frame; s_meta[1] = g.view; s_meta[2] = g.src_w; s_meta[3] = g.src_h;
  }
  __syncthreads();
  const int W = s_meta[2], H = s_meta[3];
  const uint8_t* img = src + (int64_t)s_meta[0] * frame_stride + (int64_t)s_meta[1] * view_stride;
  const int groups_per_row = dst_w / kFastPx;  // host guarantees dst_w % 4 == 0 on this path
  const int ngroups = groups_per_row * dst_h;
  const int per = (ngroups + kSplit - 1) / kSplit;
  const int g_begin = blockIdx.y * per, g_end = min(ngroups, g_begin + per);
  const int64_t obase = (int64_t)crop * dst_w * dst_h;
  for (int gi = g_begin + threadIdx.x; gi < g_end; gi += kThreads) {
    const int y = gi / groups_per_row, x0 = (gi - y * groups_per_row) * kFastPx;
    const float qy = fmaf((float)y, P.ifdy, P.qy0);
    int v[kFastPx];
#pragma unroll
    for (int i = 0; i < kFastPx; ++i) {
      float wx, wy;
      map_pixel_fast(P, fmaf((float)(x0 + i), P.ifdx, P.qx0), qy, wx, wy);
      v[i] = remap_fixed_u8(img, row_stride, W, H, wx, wy);
    }
    const int64_t o = obase + (int64_t)y * dst_w + x0;
    if (out_dtype == AT_U8) {
      *reinterpret_cast<uchar4*>((uint8_t*)out + o) = make_uchar4((uint8_t)v[0], (uint8_t)v[1], (uint8_t)v[2], (uint8_t)v[3]);
    } else if (out_dtype == AT_F32) {
      *reinterpret_cast<float4*>((float*)out + o) = make_float4(__fdiv_rn((float)v[0], 255.0f), __fdiv_rn((float)v[1], 255.0f),
                                                                 __fdiv_rn((float)v[2], 255.0f), __fdiv_rn((float)v[3], 255.0f));
    } else {
      __nv_bfloat16* ob = (__nv_bfloat16*)out + o;
#pragma unroll
      for (int i = 0; i < kFastPx; ++i) ob[i] = __float2bfloat16_rn(__fdiv_rn((float)v[i], 255.0f));
    }
  }
}


// ------------------------------------------------------------------------------------------
// Shared-memory-staged resampler (default for AT_MAP_FP32): one CTA per crop.
//  1. the fast map is evaluated on the crop's BOUNDARY pixels; their min/max (the map is
//     continuous, so the image of the boundary bounds the image of the crop) give the source
//     bounding box, clipped to the image and widened by the bilinear footprint;
//  2. the box is loaded row by row with coalesced 4-byte loads into shared memory - every source
//     byte crosses the memory system ONCE.  `src` may be device memory or PINNED HOST memory
//     (zero-copy ingest: only the ~30 KB box of each crop crosses PCIe instead of the 305 KB view);
//  3. the fixed-point bilinear gather reads its 4 taps from shared memory.
// Crops whose box does not fit (very wide crops), that look behind the camera, or whose source
// layout is not 4-byte aligned take the direct global-memory gather instead.
constexpr int kStageBytes = 72 * 1024;

__device__ __forceinline__ int remap_fixed_u8_smem(const uint8_t* __restrict__ tile, int pitch, int x0a, int y0, int bw,
                                                   int bh, int W, int H, float wx, float wy) {
  const int sx = __float2int_rn(wx * 32.0f);
  const int sy = __float2int_rn(wy * 32.0f);
  const int ix = min(max(sx >> 5, -32768), 32767);
  const int iy = min(max(sy >> 5, -32768), 32767);
  const int fx = sx & 31, fy = sy & 31;
  const bool x0 = (unsigned)ix < (unsigned)W, x1 = (unsigned)(ix + 1) < (unsigned)W;
  const bool y0v = (unsigned)iy < (unsigned)H, y1v = (unsigned)(iy + 1) < (unsigned)H;
  // box-relative tap coordinates, each clamped into the tile: taps outside the image are masked by
  // the border tests above, and a (never observed) map excursion cannot read outside the tile
  const int cx0 = min(max(ix - x0a, 0), bw - 1), cx1 = min(max(ix + 1 - x0a, 0), bw - 1);
  const int cy0 = min(max(iy - y0, 0), bh - 1) * pitch, cy1 = min(max(iy + 1 - y0, 0), bh - 1) * pitch;
  const int t00 = (y0v && x0) ? tile[cy0 + cx0] : 0;
  const int t01 = (y0v && x1) ? tile[cy0 + cx1] : 0;
  const int t10 = (y1v && x0) ? tile[cy1 + cx0] : 0;
  const int t11 = (y1v && x1) ? tile[cy1 + cx1] : 0;
  const int acc = ((32 - fy) * (32 - fx) * t00 + (32 - fy) * fx * t01 + fy * (32 - fx) * t10 + fy * fx * t11) * 32;
  return (acc + 16384) >> 15;
}

__global__ void __launch_bounds__(kThreads)
crop_resample_staged_kernel(const uint8_t* __restrict__ src, int64_t frame_stride, int64_t view_stride, int64_t row_stride,
                            const AtCropParams* __restrict__ params, int dst_w, int dst_h, void* __restrict__ out,
                            int out_dtype, unsigned long long* __restrict__ staged_bytes) {
  extern __shared__ __align__(16) uint8_t tile[];
  __shared__ FastParams P;
  __shared__ int s_meta[4];
  __shared__ float s_red[kThreads / 32][4];
  __shared__ int s_behind;
  __shared__ int s_box[6];  // x0a, y0, box width (bytes, multiple of 4), box height, direct flag
  const int crop = blockIdx.x;
  const AtCropParams& g = params[crop];
  if (threadIdx.x < 9) P.M[threadIdx.x] = (float)g.M[threadIdx.x];
  if (threadIdx.x < 3) P.d[threadIdx.x] = (float)g.d[threadIdx.x];
  if (threadIdx.x == 32) {
    P.ifdx = 1.0f / (float)g.f_dst[0]; P.ifdy = 1.0f / (float)g.f_dst[1];
    P.qx0 = -(float)g.c_dst[0] * P.ifdx; P.qy0 = -(float)g.c_dst[1] * P.ifdy;
    P.fsx = (float)g.f_src[0]; P.fsy = (float)g.f_src[1]; P.csx = (float)g.c_src[0]; P.csy = (float)g.c_src[1];
    P.k1 = (float)g.k[0]; P.k2 = (float)g.k[1]; P.k3 = (float)g.k[2]; P.k4 = (float)g.k[3];
    P.p1 = (float)g.k[4]; P.p2 = (float)g.k[5]; P.k5 = (float)g.k[6]; P.k6 = (float)g.k[7];
    s_meta[0] = g.frame; s_meta[1] = g.view; s_meta[2] = g.src_w; s_meta[3] = g.src_h;
    s_behind = 0;
  }
  __syncthreads();
  const int W = s_meta[2], H = s_meta[3];
  const uint8_t* img = src + (int64_t)s_meta[0] * frame_stride + (int64_t)s_meta[1] * view_stride;

  // ---- 1. source bounding box from the boundary pixels
  float mnx = 3.0e38f, mny = 3.0e38f, mxx = -3.0e38f, mxy = -3.0e38f;
  const int nb = 2 * (dst_w + dst_h);
  for (int i = threadIdx.x; i < nb; i += kThreads) {
    int x, y;
    if (i < dst_w) { x = i; y = 0; }
    else if (i < 2 * dst_w) { x = i - dst_w; y = dst_h - 1; }
    else if (i < 2 * dst_w + dst_h) { x = 0; y = i - 2 * dst_w; }
    else { x = dst_w - 1; y = i - 2 * dst_w - dst_h; }
    float wx, wy;
    map_pixel_fast(P, fmaf((float)x, P.ifdx, P.qx0), fmaf((float)y, P.ifdy, P.qy0), wx, wy);
    if (wx == -1.0f && wy == -1.0f) s_behind = 1;  // ez < 0 somewhere on the boundary (or exactly (-1,-1))
    mnx = fminf(mnx, wx); mxx = fmaxf(mxx, wx);
    mny = fminf(mny, wy); mxy = fmaxf(mxy, wy);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
  }
  if ((threadIdx.x & 31) == 0) {
    float* r = s_red[threadIdx.x >> 5];
    r[0] = mnx; r[1] = mxx; r[2] = mny; r[3] = mxy;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kThreads / 32; ++w) {
      mnx = fminf(mnx, s_red[w][0]); mxx = fmaxf(mxx, s_red[w][1]);
      mny = fminf(mny, s_red[w][2]); mxy = fmaxf(mxy, s_red[w][3]);
    }
    // taps: ix = round(32 wx) >> 5 in [floor(wx - 1/64), floor(wx + 1/64)], and ix + 1; 1 px of slack
    const float lim = 1.0e6f;
    mnx = fminf(fmaxf(mnx, -lim), lim); mxx = fminf(fmaxf(mxx, -lim), lim);
    mny = fminf(fmaxf(mny, -lim), lim); mxy = fminf(fmaxf(mxy, -lim), lim);
    int x0 = max(0, (int)floorf(mnx) - 2), x1 = min(W - 1, (int)floorf(mxx) + 3);
    int y0 = max(0, (int)floorf(mny) - 2), y1 = min(H - 1, (int)floorf(mxy) + 3);
    int direct = s_behind;
    int x0a = x0 & ~3, bw = 0, bh = 0;
    if (x1 < x0 || y1 < y0) {
      x0a = 0; y0 = 0; bw = 4; bh = 2;   // box outside the image: every tap is the constant border
      x0 = 0;
    } else {
      bw = ((x1 + 1 - x0a) + 3) & ~3;
      bh = y1 - y0 + 1;
      if (bh < 2) { if (y0 > 0) --y0; bh = 2; }
      if (bw < 4) bw = 4;
    }
    const bool aligned = ((W | (int)(view_stride & 3) | (int)(row_stride & 3) | (int)(frame_stride & 3)) & 3) == 0 &&
                         (((uintptr_t)src) & 3) == 0;
    // 16-byte loads when every row of the box has the same 16-byte phase and the widened box stays
    // inside the (multi-view) source row: 4x fewer, 4x larger requests (matters over PCIe)
    int vec16 = 0;
    if (aligned && (row_stride & 15) == 0 && (frame_stride & 15) == 0 && x1 >= x0 && y1 >= y0) {
      const long long view_off = (long long)s_meta[1] * view_stride;       // offset of the view inside its row
      const int phase = (int)(((uintptr_t)src + (uintptr_t)view_off + (uintptr_t)x0a) & 15);
      const int nx0 = x0a - phase;
      const int nbw = ((x0a + bw - nx0) + 15) & ~15;
      if (view_off + nx0 >= 0 && view_off + nx0 + nbw <= row_stride && (long long)nbw * bh <= kStageBytes) {
        x0a = nx0; bw = nbw; vec16 = 1;
      }
    }
    if (!aligned || (long long)bw * bh > kStageBytes || (!vec16 && x0a + bw > ((W + 3) & ~3)) || y0 + bh > H) direct = 1;
    s_box[0] = x0a; s_box[1] = y0; s_box[2] = bw; s_box[3] = bh; s_box[4] = direct; s_box[5] = vec16;
    if (staged_bytes != nullptr) atomicAdd(staged_bytes, direct ? (unsigned long long)(4ll * dst_w * dst_h) : (unsigned long long)((long long)bw * bh));
  }
  __syncthreads();
  const int x0a = s_box[0], y0 = s_box[1], bw = s_box[2], bh = s_box[3];
  const bool direct = s_box[4] != 0;

  // ---- 2. stage the box (coalesced 32-bit loads; rows are 4-byte aligned by construction)
  if (!direct && s_box[5]) {
    const int wpr = bw >> 4;
    const int total = wpr * bh;
    uint4* t128 = reinterpret_cast<uint4*>(tile);
    for (int i = threadIdx.x; i < total; i += kThreads) {
      const int r = i / wpr, c = i - r * wpr;
      t128[i] = __ldg(reinterpret_cast<const uint4*>(img + (int64_t)(y0 + r) * row_stride + x0a) + c);
    }
  } else if (!direct) {
    const int wpr = bw >> 2;
    const int total = wpr * bh;
    uint32_t* t32 = reinterpret_cast<uint32_t*>(tile);
    for (int i = threadIdx.x; i < total; i += kThreads) {
      const int r = i / wpr, c = i - r * wpr;
      const int yy = y0 + r;
      uint32_t v = 0u;
      if (yy < H) v = __ldg(reinterpret_cast<const uint32_t*>(img + (int64_t)yy * row_stride + x0a) + c);
      t32[i] = v;
    }
  }
  __syncthreads();

  // ---- 3. map + gather, 4 consecutive dst pixels per thread
  const int groups_per_row = dst_w / kFastPx;
  const int ngroups = groups_per_row * dst_h;
  const int64_t obase = (int64_t)crop * dst_w * dst_h;
  for (int gi = threadIdx.x; gi < ngroups; gi += kThreads) {
    const int y = gi / groups_per_row, xg = (gi - y * groups_per_row) * kFastPx;
    const float qy = fmaf((float)y, P.ifdy, P.qy0);
    int v[kFastPx];
#pragma unroll
    for (int i = 0; i < kFastPx; ++i) {
      float wx, wy;
      map_pixel_fast(P, fmaf((float)(xg + i), P.ifdx, P.qx0), qy, wx, wy);
      v[i] = direct ? remap_fixed_u8(img, row_stride, W, H, wx, wy)
                    : remap_fixed_u8_smem(tile, bw, x0a, y0, bw, bh, W, H, wx, wy);
    }
    const int64_t o = obase + (int64_t)y * dst_w + xg;
    if (out_dtype == AT_U8) {
      *reinterpret_cast<uchar4*>((uint8_t*)out + o) = make_uchar4((uint8_t)v[0], (uint8_t)v[1], (uint8_t)v[2], (uint8_t)v[3]);
    } else if (out_dtype == AT_F32) {
      *reinterpret_cast<float4*>((float*)out + o) = make_float4(__fdiv_rn((float)v[0], 255.0f), __fdiv_rn((float)v[1], 255.0f),
                                                                 __fdiv_rn((float)v[2], 255.0f), __fdiv_rn((float)v[3], 255.0f));
    } else {
      __nv_bfloat16* ob = (__nv_bfloat16*)out + o;
#pragma unroll
      for (int i = 0; i < kFastPx; ++i) ob[i] = __float2bfloat16_rn(__fdiv_rn((float)v[i], 255.0f));
    }
  }
}

// ------------------------------------------------------------------------------------------
// Lattice-map resampler (default for AT_MAP_FP32 at the model's crop size): the staged kernel with
// the per-pixel map replaced by a bicubic interpolation of the map itself.
//  1. the fast fp32 map is evaluated only on a 4-pixel lattice ((W/4+3) x (H/4+3) points, one ring
//     outside the crop) and stored pre-multiplied by 32 (the fixed-point scale of cv2.remap);
//  2. the source box is bounded by the lattice points covering the crop and staged through shared
//     memory exactly like the staged kernel (device or PINNED HOST source);
//  3. a thread walks 12-pixel row segments: 6 lattice columns are interpolated down to its row
//     (4-point Lagrange in y), then each pixel is a 4-point Lagrange blend in x - weights are
//     compile-time constants - followed by the fixed-point bilinear gather from shared memory.
// The pinhole->Fisheye62 map is smooth on the scale of the focal length: measured on the
// recording's geometry the interpolation error at a 4-pixel lattice is 1.6e-4 px, the same class as
// the fp32 map's own 1.2e-4 px.  ~20 instead of ~70 instructions per pixel for the
// map.  Crops that look behind the camera keep the exact per-pixel map; boxes that do not fit the
// staging buffer gather from global memory.
constexpr int kLatStep = 4;
constexpr int kLatMaxDim = 96 / kLatStep + 3;   // 27
constexpr int kLatPitch = kLatMaxDim + 1;        // 28
constexpr int kLatSeg = 12;                      // pixels per thread unit (3 lattice cells)
constexpr int kLatStageBytes = 46 * 1024;        // 4 CTAs per SM (with 8.4 KB of static shared memory each)
constexpr float kLatMinFocal = 80.0f;            // narrower crops only (see the kernel)
// Exactness.  cv2.remap only sees rint(32 * map): the interpolated fp32 map u~ (in 1/32 px) gives the
// reference's result unless the exact value lies on the other side of a rounding boundary k + 1/2, which
// requires |u~ - (k + 1/2)| < |u~ - u|.  |u~ - u| <= 0.0078 was measured over every pixel of the recording's
// geometry (numpy emulation of this arithmetic against the float64 map); every pixel inside kLatBand of a
// boundary (about 6 % of them) is queued and re-evaluated with the reference's own float64 operation
// sequence (map_pixel<double>) after the fast pass, so the kernel's output is bit-identical to
// tracker._warp_image.  kLatMaxFlagged bounds the queue; on overflow the whole crop is re-evaluated.
constexpr float kLatBand = 0.016f;
constexpr int kLatMaxFlagged = 1024;
constexpr float kLatClamp = 1048576.0f;          // |32 * map| stored on the lattice (keeps the magic-number rounding valid)
constexpr float kRoundMagic = 12582912.0f;       // 1.5 * 2^23: (x + magic) - magic == rint(x) for |x| < 2^22

// Gather from the staged tile with pre-scaled coordinates (32 * w): rint -> 1/32-px fixed point.
// The tile is stored with one pitch + 16 bytes of padding in front and one pitch behind, so the
// address of tap (ix, iy) may be formed without per-coordinate clamps: taps outside the image get
// zero weight (cv2's constant-0 border) and whatever byte their address holds is irrelevant; one
// clamp of the linear address keeps every access inside the buffer.  koff = (1 - y0) * pitch - x0a + 16.
// round-half-even of a pre-scaled coordinate without F2I, plus the distance to the nearest integer
__device__ __forceinline__ int round_fixed(float v, float& dist) {
  const float r = v + kRoundMagic;
  const int i = __float_as_int(r) - 0x4B400000;
  dist = v - (r - kRoundMagic);
  return i;
}

// FLAG = true: *near is set when a coordinate lies within kLatBand of a rounding boundary (see kLatBand)
template <bool FLAG>
__device__ __forceinline__ int remap_fixed_u8_smem_s(const uint8_t* __restrict__ tile, int pitch, int koff, int amax, int W,
                                                     int H, float sxf, float syf, bool* near = nullptr) {
  int sx, sy;
  if (FLAG) {
    float dx, dy;
    sx = round_fixed(sxf, dx);
    sy = round_fixed(syf, dy);
    *near = fmaxf(fabsf(dx), fabsf(dy)) > 0.5f - kLatBand;
  } else {
    sx = __float2int_rn(sxf);
    sy = __float2int_rn(syf);
  }
  const int ix = sx >> 5, iy = sy >> 5;
  const int fx = sx & 31, fy = sy & 31;
  const int wx1 = (unsigned)(ix + 1) < (unsigned)W ? fx : 0, wx0 = (unsigned)ix < (unsigned)W ? 32 - fx : 0;
  const int wy1 = (unsigned)(iy + 1) < (unsigned)H ? fy : 0, wy0 = (unsigned)iy < (unsigned)H ? 32 - fy : 0;
  const int a = min(max(iy * pitch + ix + koff, 0), amax);
  const uint8_t* t0 = tile + a;
  const uint8_t* t1 = t0 + pitch;
  const int acc = wy0 * (wx0 * t0[0] + wx1 * t0[1]) + wy1 * (wx0 * t1[0] + wx1 * t1[1]);
  return (acc + 512) >> 10;   // == ((acc * 32) + 16384) >> 15
}

template <bool FLAG>
__device__ __forceinline__ int remap_fixed_u8_s(const uint8_t* __restrict__ img, int64_t row_stride, int W, int H, float sxf,
                                                float syf, bool* near = nullptr) {
  int sx, sy;
  if (FLAG) {
    float dx, dy;
    sx = round_fixed(sxf, dx);
    sy = round_fixed(syf, dy);
    *near = fmaxf(fabsf(dx), fabsf(dy)) > 0.5f - kLatBand;
  } else {
    sx = __float2int_rn(sxf);
    sy = __float2int_rn(syf);
  }
  const int ix = min(max(sx >> 5, -32768), 32767);
  const int iy = min(max(sy >> 5, -32768), 32767);
  const int fx = sx & 31, fy = sy & 31;
  const bool x0 = (unsigned)ix < (unsigned)W, x1 = (unsigned)(ix + 1) < (unsigned)W;
  const bool y0 = (unsigned)iy < (unsigned)H, y1 = (unsigned)(iy + 1) < (unsigned)H;
  const uint8_t* r0 = img + (int64_t)iy * row_stride + ix;
  const uint8_t* r1 = r0 + row_stride;
  const int t00 = (y0 && x0) ? __ldg(r0) : 0;
  const int t01 = (y0 && x1) ? __ldg(r0 + 1) : 0;
  const int t10 = (y1 && x0) ? __ldg(r1) : 0;
  const int t11 = (y1 && x1) ? __ldg(r1 + 1) : 0;
  const int acc = (32 - fy) * ((32 - fx) * t00 + fx * t01) + fy * ((32 - fx) * t10 + fx * t11);
  return (acc + 512) >> 10;
}

// 4-point Lagrange weights for nodes (-1, 0, 1, 2) at t = k/4
__device__ __forceinline__ float2 lag4(int k, const float2& a, const float2& b, const float2& c, const float2& d) {
  float w0, w1, w2, w3;
  if (k == 0) return b;
  if (k == 1) { w0 = -0.0546875f; w1 = 0.8203125f; w2 = 0.2734375f; w3 = -0.0390625f; }
  else if (k == 2) { w0 = -0.0625f; w1 = 0.5625f; w2 = 0.5625f; w3 = -0.0625f; }
  else { w0 = -0.0390625f; w1 = 0.2734375f; w2 = 0.8203125f; w3 = -0.0546875f; }
  return make_float2(fmaf(w0, a.x, fmaf(w1, b.x, fmaf(w2, c.x, w3 * d.x))), fmaf(w0, a.y, fmaf(w1, b.y, fmaf(w2, c.y, w3 * d.y))));
}

// one output element in the crop's dtype
__device__ __forceinline__ void store_level(void* __restrict__ out, int out_dtype, int64_t idx, int v) {
  if (out_dtype == AT_U8) ((uint8_t*)out)[idx] = (uint8_t)v;
  else if (out_dtype == AT_F32) ((float*)out)[idx] = __fdiv_rn((float)v, 255.0f);
  else ((__nv_bfloat16*)out)[idx] = __float2bfloat16_rn(__fdiv_rn((float)v, 255.0f));
}

__global__ void __launch_bounds__(kThreads, 4)
crop_resample_lattice_kernel(const uint8_t* __restrict__ src, int64_t frame_stride, int64_t view_stride, int64_t row_stride,
                             const AtCropParams* __restrict__ params, int dst_w, int dst_h, void* __restrict__ out,
                             int out_dtype, unsigned long long* __restrict__ staged_bytes, int exact) {
  extern __shared__ __align__(16) uint8_t tile[];
  __shared__ FastParams P;
  __shared__ MapParams<double> PD;   // the reference's float64 map, for the boundary pixels
  __shared__ int s_meta[4];
  __shared__ float s_red[kThreads / 32][4];
  __shared__ int s_behind;
  __shared__ int s_nflag;
  __shared__ int s_box[6];
  __shared__ float2 s_lat[kLatMaxDim * kLatPitch];
  __shared__ unsigned short s_list[kLatMaxFlagged];
  abt::pdl_prologue();   // (programmatic dependent launch: resident while the previous kernel of the stream drains)
  const int crop = blockIdx.x;
  const AtCropParams& g = params[crop];
  if (threadIdx.x < 9) { P.M[threadIdx.x] = (float)g.M[threadIdx.x]; PD.M[threadIdx.x] = g.M[threadIdx.x]; }
  if (threadIdx.x < 3) { P.d[threadIdx.x] = (float)g.d[threadIdx.x]; PD.d[threadIdx.x] = g.d[threadIdx.x]; }
  if (threadIdx.x == 32) {
    P.ifdx = 1.0f / (float)g.f_dst[0]; P.ifdy = 1.0f / (float)g.f_dst[1];
    P.qx0 = -(float)g.c_dst[0] * P.ifdx; P.qy0 = -(float)g.c_dst[1] * P.ifdy;
    P.fsx = (float)g.f_src[0]; P.fsy = (float)g.f_src[1]; P.csx = (float)g.c_src[0]; P.csy = (float)g.c_src[1];
    P.k1 = (float)g.k[0]; P.k2 = (float)g.k[1]; P.k3 = (float)g.k[2]; P.k4 = (float)g.k[3];
    P.p1 = (float)g.k[4]; P.p2 = (float)g.k[5]; P.k5 = (float)g.k[6]; P.k6 = (float)g.k[7];
    s_meta[0] = g.frame; s_meta[1] = g.view; s_meta[2] = g.src_w; s_meta[3] = g.src_h;
    // the interpolation error grows like (lattice step / focal length)^4: hand crops have f >= 90 px
    // (1.6e-4 px); anything much wider keeps the exact per-pixel map, like crops that look behind the camera
    s_behind = fminf((float)g.f_dst[0], (float)g.f_dst[1]) < kLatMinFocal ? 1 : 0;
    s_nflag = 0;
  }
  if (threadIdx.x == 64) {
    PD.fdx = g.f_dst[0]; PD.fdy = g.f_dst[1]; PD.cdx = g.c_dst[0]; PD.cdy = g.c_dst[1];
    PD.fsx = g.f_src[0]; PD.fsy = g.f_src[1]; PD.csx = g.c_src[0]; PD.csy = g.c_src[1];
    PD.k1 = g.k[0]; PD.k2 = g.k[1]; PD.k3 = g.k[2]; PD.k4 = g.k[3];
    PD.p1 = g.k[4]; PD.p2 = g.k[5]; PD.k5 = g.k[6]; PD.k6 = g.k[7];
  }
  __syncthreads();
  const int W = s_meta[2], H = s_meta[3];
  const uint8_t* img = src + (int64_t)s_meta[0] * frame_stride + (int64_t)s_meta[1] * view_stride;

  // ---- 1. the map on the lattice (coordinates -4, 0, 4, ..., W + 4) and the source box of the crop
  const int nlx = dst_w / kLatStep + 3, nly = dst_h / kLatStep + 3;
  float mnx = 3.0e38f, mny = 3.0e38f, mxx = -3.0e38f, mxy = -3.0e38f;
  for (int i = threadIdx.x; i < nlx * nly; i += kThreads) {
    const int ly = i / nlx, lx = i - ly * nlx;
    float wx, wy;
    map_pixel_fast(P, fmaf((float)(kLatStep * (lx - 1)), P.ifdx, P.qx0), fmaf((float)(kLatStep * (ly - 1)), P.ifdy, P.qy0), wx, wy);
    if (wx == -1.0f && wy == -1.0f) s_behind = 1;  // ez < 0 at a lattice point: the map is not smooth there
    s_lat[ly * kLatPitch + lx] = make_float2(fminf(fmaxf(wx * 32.0f, -kLatClamp), kLatClamp),
                                             fminf(fmaxf(wy * 32.0f, -kLatClamp), kLatClamp));
    if (lx >= 1 && lx <= nlx - 2 && ly >= 1 && ly <= nly - 2) {  // the points spanning [0, W] x [0, H]
      mnx = fminf(mnx, wx); mxx = fmaxf(mxx, wx);
      mny = fminf(mny, wy); mxy = fmaxf(mxy, wy);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
  }
  if ((threadIdx.x & 31) == 0) {
    float* r = s_red[threadIdx.x >> 5];
    r[0] = mnx; r[1] = mxx; r[2] = mny; r[3] = mxy;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kThreads / 32; ++w) {
      mnx = fminf(mnx, s_red[w][0]); mxx = fmaxf(mxx, s_red[w][1]);
      mny = fminf(mny, s_red[w][2]); mxy = fmaxf(mxy, s_red[w][3]);
    }
    const float lim = 1.0e6f;
    mnx = fminf(fmaxf(mnx, -lim), lim); mxx = fminf(fmaxf(mxx, -lim), lim);
    mny = fminf(fmaxf(mny, -lim), lim); mxy = fminf(fmaxf(mxy, -lim), lim);
    int x0 = max(0, (int)floorf(mnx) - 2), x1 = min(W - 1, (int)floorf(mxx) + 3);
    int y0 = max(0, (int)floorf(mny) - 2), y1 = min(H - 1, (int)floorf(mxy) + 3);
    int direct = 0;
    int x0a = x0 & ~3, bw = 0, bh = 0;
    if (x1 < x0 || y1 < y0) {
      x0a = 0; y0 = 0; bw = 4; bh = 2;   // box outside the image: every tap is the constant border
      x0 = 0;
    } else {
      bw = ((x1 + 1 - x0a) + 3) & ~3;
      bh = y1 - y0 + 1;
      if (bh < 2) { if (y0 > 0) --y0; bh = 2; }
      if (bw < 4) bw = 4;
    }
    const bool aligned = ((W | (int)(view_stride & 3) | (int)(row_stride & 3) | (int)(frame_stride & 3)) & 3) == 0 &&
                         (((uintptr_t)src) & 3) == 0;
    int vec16 = 0;
    if (aligned && (row_stride & 15) == 0 && (frame_stride & 15) == 0 && x1 >= x0 && y1 >= y0) {
      const long long view_off = (long long)s_meta[1] * view_stride;
      const int phase = (int)(((uintptr_t)src + (uintptr_t)view_off + (uintptr_t)x0a) & 15);
      const int nx0 = x0a - phase;
      const int nbw = ((x0a + bw - nx0) + 15) & ~15;
      if (view_off + nx0 >= 0 && view_off + nx0 + nbw <= row_stride && (long long)nbw * (bh + 2) + 32 <= kLatStageBytes) {
        x0a = nx0; bw = nbw; vec16 = 1;
      }
    }
    if (!aligned || (long long)bw * (bh + 2) + 32 > kLatStageBytes || (!vec16 && x0a + bw > ((W + 3) & ~3)) || y0 + bh > H) direct = 1;
    s_box[0] = x0a; s_box[1] = y0; s_box[2] = bw; s_box[3] = bh; s_box[4] = direct; s_box[5] = vec16;
    if (staged_bytes != nullptr)
      atomicAdd(staged_bytes, (direct || s_behind) ? (unsigned long long)(4ll * dst_w * dst_h) : (unsigned long long)((long long)bw * bh));
  }
  __syncthreads();
  const int x0a = s_box[0], y0 = s_box[1], bw = s_box[2], bh = s_box[3];
  const bool behind = s_behind != 0;
  const bool direct = s_box[4] != 0 || behind;

  // ---- 2. stage the box (data starts one pitch + 16 bytes into the buffer: see remap_fixed_u8_smem_s)
  uint8_t* tdata = tile + bw + 16;
  if (!direct && s_box[5]) {
    const int wpr = bw >> 4;
    const int total = wpr * bh;
    uint4* t128 = reinterpret_cast<uint4*>(tdata);
    for (int i = threadIdx.x; i < total; i += kThreads) {
      const int r = i / wpr, c = i - r * wpr;
      t128[i] = __ldg(reinterpret_cast<const uint4*>(img + (int64_t)(y0 + r) * row_stride + x0a) + c);
    }
  } else if (!direct) {
    const int wpr = bw >> 2;
    const int total = wpr * bh;
    uint32_t* t32 = reinterpret_cast<uint32_t*>(tdata);
    for (int i = threadIdx.x; i < total; i += kThreads) {
      const int r = i / wpr, c = i - r * wpr;
      const int yy = y0 + r;
      uint32_t v = 0u;
      if (yy < H) v = __ldg(reinterpret_cast<const uint32_t*>(img + (int64_t)yy * row_stride + x0a) + c);
      t32[i] = v;
    }
  }
  __syncthreads();

  const int npix = dst_w * dst_h;
  const int64_t obase = (int64_t)crop * npix;
  const int koff = (1 - y0) * bw - x0a + 16;       // tap address = iy * pitch + ix + koff (padded tile)
  const int amax = (bh + 1) * bw + 16 - 2;          // last address whose 2x2 footprint is inside the buffer
  // one pixel through the reference's own float64 operation sequence (bit-identical map), then the same gather
  auto exact_pixel = [&](int p) {
    const int y = p / dst_w, x = p - y * dst_w;
    float wx, wy;
    map_pixel<double>(PD, x, y, wx, wy);
    const int v = direct ? remap_fixed_u8(img, row_stride, W, H, wx, wy)
                         : remap_fixed_u8_smem_s<false>(tile, bw, koff, amax, W, H, wx * 32.0f, wy * 32.0f);
    store_level(out, out_dtype, obase + p, v);
  };
  if (behind && exact) {  // no smooth map to interpolate: every pixel through the float64 map (rare crops)
    for (int p = threadIdx.x; p < npix; p += kThreads) exact_pixel(p);
    return;
  }

  // ---- 3. interpolate the map and gather: units of 12 pixels in a row
  const int segs = dst_w / kLatSeg;
  const int nunits = segs * dst_h;
  const int lane = threadIdx.x & 31;
  for (int u0 = 0; u0 < nunits; u0 += kThreads) {
    const int u = u0 + threadIdx.x;
    const int y = segs == 8 ? u >> 3 : u / segs, seg = u - y * segs;   // 96-pixel rows: no division
    const int xs = seg * kLatSeg;
    unsigned near_mask = 0u;
    if (u < nunits) {
      int v[kLatSeg];
      if (behind) {  // uniform per CTA (exact == 0 only)
        const float qy = fmaf((float)y, P.ifdy, P.qy0);
#pragma unroll
        for (int i = 0; i < kLatSeg; ++i) {
          float wx, wy;
          map_pixel_fast(P, fmaf((float)(xs + i), P.ifdx, P.qx0), qy, wx, wy);
          v[i] = remap_fixed_u8_s<false>(img, row_stride, W, H, wx * 32.0f, wy * 32.0f);
        }
      } else {
        const int cy = y >> 2, ty = y & 3;
        const float2* L = s_lat + cy * kLatPitch + 3 * seg;
        float2 col[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) col[j] = lag4(ty, L[j], L[kLatPitch + j], L[2 * kLatPitch + j], L[3 * kLatPitch + j]);
        if (!direct) {
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 m = lag4(t, col[c], col[c + 1], col[c + 2], col[c + 3]);
              bool near;
              v[c * 4 + t] = remap_fixed_u8_smem_s<true>(tile, bw, koff, amax, W, H, m.x, m.y, &near);
              near_mask |= near ? 1u << (c * 4 + t) : 0u;
            }
        } else {
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 m = lag4(t, col[c], col[c + 1], col[c + 2], col[c + 3]);
              bool near;
              v[c * 4 + t] = remap_fixed_u8_s<true>(img, row_stride, W, H, m.x, m.y, &near);
              near_mask |= near ? 1u << (c * 4 + t) : 0u;
            }
        }
      }
      const int64_t o = obase + (int64_t)y * dst_w + xs;
      if (out_dtype == AT_U8) {
        uchar4* op = reinterpret_cast<uchar4*>((uint8_t*)out + o);
#pragma unroll
        for (int q = 0; q < 3; ++q)
          op[q] = make_uchar4((uint8_t)v[4 * q], (uint8_t)v[4 * q + 1], (uint8_t)v[4 * q + 2], (uint8_t)v[4 * q + 3]);
      } else if (out_dtype == AT_F32) {
        float4* op = reinterpret_cast<float4*>((float*)out + o);
#pragma unroll
        for (int q = 0; q < 3; ++q)
          op[q] = make_float4(__fdiv_rn((float)v[4 * q], 255.0f), __fdiv_rn((float)v[4 * q + 1], 255.0f),
                              __fdiv_rn((float)v[4 * q + 2], 255.0f), __fdiv_rn((float)v[4 * q + 3], 255.0f));
      } else {
        __nv_bfloat16* ob = (__nv_bfloat16*)out + o;
#pragma unroll
        for (int i = 0; i < kLatSeg; ++i) ob[i] = __float2bfloat16_rn(__fdiv_rn((float)v[i], 255.0f));
      }
    }
    if (exact) {  // queue the pixels next to a rounding boundary (warp-aggregated append)
      const int cnt = __popc(near_mask);
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      if (total) {
        int base = 0;
        if (lane == 31) base = atomicAdd(&s_nflag, total);
        base = __shfl_sync(0xffffffffu, base, 31);
        int pos = base + incl - cnt;
        while (near_mask) {
          const int b = __ffs(near_mask) - 1;
          near_mask &= near_mask - 1;
          if (pos < kLatMaxFlagged) s_list[pos] = (unsigned short)(y * dst_w + xs + b);
          ++pos;
        }
      }
    }
  }
  if (!exact) return;

  // ---- 4. the queued pixels again, through the float64 map (their stores follow the fast pass's: barrier)
  __syncthreads();
  const int nf = s_nflag;
  if (nf > kLatMaxFlagged) {
    for (int p = threadIdx.x; p < npix; p += kThreads) exact_pixel(p);
  } else {
    for (int i = threadIdx.x; i < nf; i += kThreads) exact_pixel((int)s_list[i]);
  }
}

// ---- source-box pull (host -> device ingest without copying whole frames) -------------------------
// For every crop: bound the source region its map touches (boundary pixels of the crop, as in the
// staged kernel), then copy just those rows/columns from `src` (pinned host memory, read over PCIe)
// into `dst`, a device buffer with the SAME layout.  The ordinary resampler then runs on `dst`: bytes
// outside the boxes are never read.  Spans are widened to 16-byte address alignment -- harmless,
// because source and destination share one layout.
//
// The kernel is built to hide behind the previous step's encoder: a 384-thread x 168-register
// conv_tc CTA with ~220 KB of shared memory leaves 1024 registers, ~6 KB of shared memory and warp
// slots free on its SM.  So: ONE warp per CTA, 32 registers, and the bytes in flight live in a small
// shared-memory ring filled by cp.async.ca (LDGSTS; the .cg form reads PCIe at half the rate)
// instead of in registers.  Each lane reads back only the 16-byte slots it filled itself, so
// cp.async.wait_group is the only synchronisation in the copy loop.  The launch also asks for the
// maximum shared-memory carve-out, like the conv kernels: see at_common.cuh.
constexpr int kPullStages = 2;        // ring halves

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// word i of the box -> byte offset of its 16-byte word, or 0xFFFFFFFF when the word is past the row / buffer
__device__ __forceinline__ uint32_t pull_word_offset(int i, int total, int wpr, uint32_t base, uint32_t row_stride,
                                                     uint32_t bw, uint32_t tb) {
  if (i >= total) return 0xFFFFFFFFu;
  const int r = i / wpr, c = i - r * wpr;
  const uint32_t row0 = base + (uint32_t)r * row_stride;
  const uint32_t addr = (row0 & ~15u) + ((uint32_t)c << 4);
  return (addr < row0 + bw && addr + 16u <= tb) ? addr : 0xFFFFFFFFu;
}

// Bulk-TMA variant of the pull (AT_PULL_TMA=1, A/B): the same boxes, moved as chunks of <= 256 bytes of a row by the
// copy engine of the SM - cp.async.bulk from the pinned frames into a 6-slot shared-memory ring (mbarrier completion),
// cp.async.bulk back out to the device buffer - issued by ONE thread; no LSU traffic, no registers holding data.
// Up to kTmaAhead loads in flight per CTA.  The other lanes only help with the box bounds.
constexpr int kTmaSlots = 6, kTmaAhead = 4, kTmaChunk = 256;

__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__global__ void __launch_bounds__(32)
crop_pull_boxes_tma_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int64_t total_bytes,
                           int64_t frame_stride, int64_t view_stride, int64_t row_stride,
                           const AtCropParams* __restrict__ params, int n_crops, int dst_w, int dst_h,
                           unsigned long long* __restrict__ pulled_bytes) {
  __shared__ __align__(128) uint8_t ring[kTmaSlots * kTmaChunk];
  __shared__ __align__(8) uint64_t bars[kTmaSlots];
  __shared__ uint32_t c_addr[kTmaSlots], c_bytes[kTmaSlots];
  __shared__ FastParams P;
  using namespace abt::ptx;
  const int lane = threadIdx.x;
  const uint32_t ring_s = smem_u32(ring), bar_s = smem_u32(bars);
  if (lane == 0) {
    for (int i = 0; i < kTmaSlots; ++i) mbar_init(bar_s + 8 * i, 1);
    fence_barrier_init();
  }
  __syncwarp();
  uint32_t produced = 0, consumed = 0;   // chunk counters of lane 0 (slot = counter % kTmaSlots, parity = counter / kTmaSlots & 1)
  unsigned long long my_bytes = 0ull;
  for (int crop = blockIdx.x; crop < n_crops; crop += gridDim.x) {
    const AtCropParams& g = params[crop];
    __syncwarp();
    if (lane < 9) P.M[lane] = (float)g.M[lane];
    if (lane < 3) P.d[lane] = (float)g.d[lane];
    if (lane == 31) {
      P.ifdx = 1.0f / (float)g.f_dst[0]; P.ifdy = 1.0f / (float)g.f_dst[1];
      P.qx0 = -(float)g.c_dst[0] * P.ifdx; P.qy0 = -(float)g.c_dst[1] * P.ifdy;
      P.fsx = (float)g.f_src[0]; P.fsy = (float)g.f_src[1]; P.csx = (float)g.c_src[0]; P.csy = (float)g.c_src[1];
      P.k1 = (float)g.k[0]; P.k2 = (float)g.k[1]; P.k3 = (float)g.k[2]; P.k4 = (float)g.k[3];
      P.p1 = (float)g.k[4]; P.p2 = (float)g.k[5]; P.k5 = (float)g.k[6]; P.k6 = (float)g.k[7];
    }
    __syncwarp();
    const int W = g.src_w, H = g.src_h;
    float mnx = 3.0e38f, mny = 3.0e38f, mxx = -3.0e38f, mxy = -3.0e38f;
    bool behind = false;
    const int nb = 2 * (dst_w + dst_h);
    for (int i = lane; i < nb; i += 32) {
      int x, y;
      if (i < dst_w) { x = i; y = 0; }
      else if (i < 2 * dst_w) { x = i - dst_w; y = dst_h - 1; }
      else if (i < 2 * dst_w + dst_h) { x = 0; y = i - 2 * dst_w; }
      else { x = dst_w - 1; y = i - 2 * dst_w - dst_h; }
      float wx, wy;
      map_pixel_fast(P, fmaf((float)x, P.ifdx, P.qx0), fmaf((float)y, P.ifdy, P.qy0), wx, wy);
      behind |= (wx == -1.0f && wy == -1.0f);
      mnx = fminf(mnx, wx); mxx = fmaxf(mxx, wx);
      mny = fminf(mny, wy); mxy = fmaxf(mxy, wy);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
      mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    }
    behind = __any_sync(0xffffffffu, behind);
    const float lim = 1.0e6f;
    mnx = fminf(fmaxf(mnx, -lim), lim); mxx = fminf(fmaxf(mxx, -lim), lim);
    mny = fminf(fmaxf(mny, -lim), lim); mxy = fminf(fmaxf(mxy, -lim), lim);
    int x0 = max(0, (int)floorf(mnx) - 2), x1 = min(W - 1, (int)floorf(mxx) + 3);   // the same box as crop_pull_boxes_kernel
    int y0 = max(0, (int)floorf(mny) - 2), y1 = min(H - 1, (int)floorf(mxy) + 3);
    if (behind) { x0 = 0; y0 = 0; x1 = W - 1; y1 = H - 1; }
    const int bw = x1 - x0 + 1, bh = y1 - y0 + 1;
    if (bw <= 0 || bh <= 0 || lane != 0) continue;
    const uint32_t base = (uint32_t)((int64_t)g.frame * frame_stride + (int64_t)g.view * view_stride +
                                     (int64_t)y0 * row_stride + x0);
    const uint32_t tb = (uint32_t)total_bytes & ~15u, rs = (uint32_t)row_stride;
    // chunk generator: rows top to bottom, each row's 16-byte-aligned span [s, e) in pieces of kTmaChunk bytes
    int r = 0;
    uint32_t s = base & ~15u, e = min((base + (uint32_t)bw + 15u) & ~15u, tb);
    auto drain_one = [&]() {
      const uint32_t slot = consumed % kTmaSlots;
      mbar_wait(bar_s + 8 * slot, (consumed / kTmaSlots) & 1u);
      bulk_s2g(dst + c_addr[slot], ring_s + slot * kTmaChunk, c_bytes[slot]);
      bulk_commit();
      ++consumed;
    };
    while (r < bh) {
      if (s >= e) {   // next row
        if (++r >= bh) break;
        const uint32_t a = base + (uint32_t)r * rs;
        s = a & ~15u; e = min((a + (uint32_t)bw + 15u) & ~15u, tb);
        continue;
      }
      const uint32_t n = min(e - s, (uint32_t)kTmaChunk);
      if (produced - consumed >= (uint32_t)kTmaAhead) drain_one();
      // the slot about to be refilled was stored kTmaSlots - kTmaAhead >= 2 commits ago: all but the newest group have read it
      bulk_wait_read<1>();
      const uint32_t slot = produced % kTmaSlots;
      c_addr[slot] = s; c_bytes[slot] = n;
      mbar_expect_tx(bar_s + 8 * slot, n);
      bulk_g2s(ring_s + slot * kTmaChunk, src + s, n, bar_s + 8 * slot);
      ++produced;
      my_bytes += n;
      s += n;
    }
    while (consumed < produced) drain_one();
  }
  if (lane == 0) {
    bulk_wait_all();
    if (pulled_bytes != nullptr && my_bytes) atomicAdd(pulled_bytes, my_bytes);
  }
}

template <int kPullPerLane>  // 16-byte slots per lane per stage: ring = 2 x kPullPerLane x 512 B
__global__ void __maxnreg__(32)
crop_pull_boxes_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int64_t total_bytes,
                       int64_t frame_stride, int64_t view_stride, int64_t row_stride,
                       const AtCropParams* __restrict__ params, int n_crops, int dst_w, int dst_h,
                       unsigned long long* __restrict__ pulled_bytes) {
  __shared__ __align__(16) uint8_t ring[kPullStages * kPullPerLane * 32 * 16];
  __shared__ FastParams P;
  const int lane = threadIdx.x;
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring) + (uint32_t)lane * 16u;
  uint32_t my_words = 0u;
  for (int crop = blockIdx.x; crop < n_crops; crop += gridDim.x) {
    const AtCropParams& g = params[crop];
    __syncwarp();
    if (lane < 9) P.M[lane] = (float)g.M[lane];
    if (lane < 3) P.d[lane] = (float)g.d[lane];
    if (lane == 31) {
      P.ifdx = 1.0f / (float)g.f_dst[0]; P.ifdy = 1.0f / (float)g.f_dst[1];
      P.qx0 = -(float)g.c_dst[0] * P.ifdx; P.qy0 = -(float)g.c_dst[1] * P.ifdy;
      P.fsx = (float)g.f_src[0]; P.fsy = (float)g.f_src[1]; P.csx = (float)g.c_src[0]; P.csy = (float)g.c_src[1];
      P.k1 = (float)g.k[0]; P.k2 = (float)g.k[1]; P.k3 = (float)g.k[2]; P.k4 = (float)g.k[3];
      P.p1 = (float)g.k[4]; P.p2 = (float)g.k[5]; P.k5 = (float)g.k[6]; P.k6 = (float)g.k[7];
    }
    __syncwarp();
    const int W = g.src_w, H = g.src_h;
    float mnx = 3.0e38f, mny = 3.0e38f, mxx = -3.0e38f, mxy = -3.0e38f;
    bool behind = false;
    const int nb = 2 * (dst_w + dst_h);
    for (int i = lane; i < nb; i += 32) {
      int x, y;
      if (i < dst_w) { x = i; y = 0; }
      else if (i < 2 * dst_w) { x = i - dst_w; y = dst_h - 1; }
      else if (i < 2 * dst_w + dst_h) { x = 0; y = i - 2 * dst_w; }
      else { x = dst_w - 1; y = i - 2 * dst_w - dst_h; }
      float wx, wy;
      map_pixel_fast(P, fmaf((float)x, P.ifdx, P.qx0), fmaf((float)y, P.ifdy, P.qy0), wx, wy);
      behind |= (wx == -1.0f && wy == -1.0f);
      mnx = fminf(mnx, wx); mxx = fmaxf(mxx, wx);
      mny = fminf(mny, wy); mxy = fmaxf(mxy, wy);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
      mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    }
    behind = __any_sync(0xffffffffu, behind);
    const float lim = 1.0e6f;
    mnx = fminf(fmaxf(mnx, -lim), lim); mxx = fminf(fmaxf(mxx, -lim), lim);
    mny = fminf(fmaxf(mny, -lim), lim); mxy = fminf(fmaxf(mxy, -lim), lim);
    // taps: ix = round(32 wx) >> 5 and ix + 1 (see the staged kernel); 2 px of slack
    int x0 = max(0, (int)floorf(mnx) - 2), x1 = min(W - 1, (int)floorf(mxx) + 3);
    int y0 = max(0, (int)floorf(mny) - 2), y1 = min(H - 1, (int)floorf(mxy) + 3);
    if (behind) { x0 = 0; y0 = 0; x1 = W - 1; y1 = H - 1; }  // part of the crop looks behind the camera: whole view
    const int bw = x1 - x0 + 1, bh = y1 - y0 + 1;
    if (bw <= 0 || bh <= 0) continue;                         // nothing inside the image (warp-uniform)
    const uint32_t base = (uint32_t)((int64_t)g.frame * frame_stride + (int64_t)g.view * view_stride +
                                     (int64_t)y0 * row_stride + x0);
    const int wpr = (bw + 15 + 15) >> 4;  // 16-byte words per row, upper bound over row phases
    const int total = wpr * bh;
    const uint32_t tb = (uint32_t)total_bytes, rs = (uint32_t)row_stride;
    constexpr int kStageWords = 32 * kPullPerLane;
    const int nst = (total + kStageWords - 1) / kStageWords;
    // prologue: fill kPullStages - 1 stages
    for (int st = 0; st < nst + kPullStages - 1; ++st) {
      if (st < nst) {
        const uint32_t sbase = ring_s + (uint32_t)(st % kPullStages) * (kStageWords * 16);
#pragma unroll
        for (int u = 0; u < kPullPerLane; ++u) {
          const uint32_t a = pull_word_offset(st * kStageWords + u * 32 + lane, total, wpr, base, rs, (uint32_t)bw, tb);
          if (a != 0xFFFFFFFFu) cp_async16(sbase + u * 512, src + a);
        }
      }
      cp_async_commit();
      const int dr = st - (kPullStages - 1);  // stage to drain
      if (dr >= 0) {
        cp_async_wait<kPullStages - 1>();
        const uint8_t* sb = ring + lane * 16 + (dr % kPullStages) * (kStageWords * 16);
#pragma unroll
        for (int u = 0; u < kPullPerLane; ++u) {
          const uint32_t a = pull_word_offset(dr * kStageWords + u * 32 + lane, total, wpr, base, rs, (uint32_t)bw, tb);
          if (a != 0xFFFFFFFFu) {
            *reinterpret_cast<uint4*>(dst + a) = *reinterpret_cast<const uint4*>(sb + u * 512);
            ++my_words;
          }
        }
      }
    }
    cp_async_wait<0>();
  }
  if (pulled_bytes != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_words += __shfl_xor_sync(0xffffffffu, my_words, o);
    if (lane == 0 && my_words) atomicAdd(pulled_bytes, 16ull * my_words);
  }
}

// data_transform.py:78-144, float64 map (numpy promotes float32 x int32 -> float64 at :99).
__global__ void __launch_bounds__(kThreads)
resample_homography_kernel(const float* __restrict__ images, int H, int W, const float* __restrict__ xfs, int oh,
                           int ow, float* __restrict__ out, float out_divisor) {
  const int n = blockIdx.x;
  __shared__ double m[12];
  if (threadIdx.x < 12) m[threadIdx.x] = (double)xfs[(int64_t)n * 16 + threadIdx.x];
  __syncthreads();
  const float* img = images + (int64_t)n * H * W;
  const int npix = oh * ow;
  const int per = (npix + kSplit - 1) / kSplit;
  const int p_begin = blockIdx.y * per, p_end = min(npix, p_begin + per);
  for (int p = p_begin + threadIdx.x; p < p_end; p += kThreads) {
    const int v = p / ow, u = p - v * ow;
    const double du = u, dv = v;
    // tensordot order (:99): sum over (u, v, 1) then + t (:101)
    const double gx = (m[0] * du + m[1] * dv + m[2]) + m[3];
    const double gy = (m[4] * du + m[5] * dv + m[6]) + m[7];
    const double gz = (m[8] * du + m[9] * dv + m[10]) + m[11];
    const double x = gx / gz, y = gy / gz;
    float res = 0.0f;
    if (x >= 0.0 && x < (double)(W - 1) && y >= 0.0 && y < (double)(H - 1)) {
      const int x0 = (int)x, y0 = (int)y, x1 = x0 + 1, y1 = y0 + 1;
      const double f00 = img[(int64_t)y0 * W + x0], f01 = img[(int64_t)y1 * W + x0];
      const double f10 = img[(int64_t)y0 * W + x1], f11 = img[(int64_t)y1 * W + x1];
      const double val = (((f00 * (x1 - x)) * (y1 - y) + (f10 * (x - x0)) * (y1 - y)) + (f01 * (x1 - x)) * (y - y0)) +
                         (f11 * (x - x0)) * (y - y0);
      res = (float)(val / (double)((x1 - x0) * (y1 - y0)));
    }
    out[(int64_t)n * npix + p] = __fdiv_rn(res, out_divisor);
  }
}

}  // namespace

extern "C" int at_crop_resample_fisheye62_u8(const uint8_t* src, int64_t frame_stride, int64_t view_stride,
                                             int64_t row_stride, const AtCropParams* params, int n_crops,
                                             int dst_w, int dst_h, void* out, int out_dtype, int map_mode,
                                             void* stream) {
  return at_crop_resample_fisheye62_u8_ex(src, frame_stride, view_stride, row_stride, params, n_crops, dst_w, dst_h, out,
                                          out_dtype, map_mode, nullptr, stream);
}

extern "C" int at_crop_resample_fisheye62_u8_ex(const uint8_t* src, int64_t frame_stride, int64_t view_stride,
                                                int64_t row_stride, const AtCropParams* params, int n_crops,
                                                int dst_w, int dst_h, void* out, int out_dtype, int map_mode,
                                                unsigned long long* staged_bytes, void* stream) {
  AT_REQUIRE(n_crops >= 0, "at_crop_resample_fisheye62_u8: n_crops < 0");
  if (n_crops == 0) return AT_OK;
  AT_REQUIRE(src && params && out, "at_crop_resample_fisheye62_u8: NULL pointer");
  AT_REQUIRE(dst_w > 0 && dst_h > 0 && dst_w <= 4096 && dst_h <= 4096, "at_crop_resample_fisheye62_u8: bad crop size");
  AT_REQUIRE(out_dtype == AT_U8 || out_dtype == AT_F32 || out_dtype == AT_BF16,
             "at_crop_resample_fisheye62_u8: bad out_dtype %d", out_dtype);
  AT_REQUIRE(map_mode == AT_MAP_FP32 || map_mode == AT_MAP_FP64, "at_crop_resample_fisheye62_u8: bad map_mode %d",
             map_mode);
  // pinned host source (zero-copy ingest) -> the staged kernel: every source byte crosses PCIe once
  bool src_on_host = false;
  {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, src) == cudaSuccess) src_on_host = attr.type == cudaMemoryTypeHost;
    else cudaGetLastError();
  }
  AT_REQUIRE(!src_on_host || (map_mode == AT_MAP_FP32 && dst_w % 4 == 0),
             "at_crop_resample_fisheye62_u8: a host (pinned) source needs AT_MAP_FP32 and dst_w % 4 == 0");
  dim3 grid((unsigned)n_crops, kSplit);
  if (map_mode == AT_MAP_FP64) {
    crop_resample_kernel<double><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        src, frame_stride, view_stride, row_stride, params, dst_w, dst_h, out, out_dtype);
  } else if (dst_w % kLatSeg == 0 && dst_h % kLatStep == 0 && dst_w <= 96 && dst_h <= 96 &&
             (getenv("AT_CROP_LATTICE") == nullptr || atoi(getenv("AT_CROP_LATTICE")) != 0)) {
    // the model's crop size: bicubic lattice map + staged gather, for device and pinned-host sources alike
    AT_ONCE_PER_DEVICE(
        cudaFuncSetAttribute(crop_resample_lattice_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kLatStageBytes);
        cudaFuncSetAttribute(crop_resample_lattice_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    static const int exact = getenv("AT_CROP_EXACT") == nullptr || atoi(getenv("AT_CROP_EXACT")) != 0;  // 0: A/B only
    abt::launch_pdl(crop_resample_lattice_kernel, dim3((unsigned)n_crops), dim3(kThreads), kLatStageBytes, (cudaStream_t)stream,
                    src, frame_stride, view_stride, row_stride, params, dst_w, dst_h, out, out_dtype, staged_bytes, exact);
  } else if (dst_w % kFastPx == 0 && (src_on_host || staged_bytes != nullptr || getenv("AT_CROP_STAGED") != nullptr)) {
    AT_ONCE_PER_DEVICE(
        cudaFuncSetAttribute(crop_resample_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStageBytes);
        cudaFuncSetAttribute(crop_resample_staged_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    crop_resample_staged_kernel<<<(unsigned)n_crops, kThreads, kStageBytes, (cudaStream_t)stream>>>(
        src, frame_stride, view_stride, row_stride, params, dst_w, dst_h, out, out_dtype, staged_bytes);
  } else if (dst_w % kFastPx == 0) {
    crop_resample_fast_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(src, frame_stride, view_stride, row_stride,
                                                                           params, dst_w, dst_h, out, out_dtype);
  } else {
    crop_resample_kernel<float><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        src, frame_stride, view_stride, row_stride, params, dst_w, dst_h, out, out_dtype);
  }
  return abt::check_launch("crop_resample_kernel");
}

extern "C" int at_resample_homography_f32(const float* images, int n, int H, int W, const float* xfs, int oh, int ow,
                                          float* out, float out_divisor, void* stream) {
  AT_REQUIRE(n >= 0, "at_resample_homography_f32: n < 0");
  if (n == 0) return AT_OK;
  AT_REQUIRE(images && xfs && out, "at_resample_homography_f32: NULL pointer");
  AT_REQUIRE(H > 1 && W > 1 && oh > 0 && ow > 0, "at_resample_homography_f32: bad image size");
  AT_REQUIRE(out_divisor != 0.0f, "at_resample_homography_f32: out_divisor is 0");
  dim3 grid((unsigned)n, kSplit);
  resample_homography_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(images, H, W, xfs, oh, ow, out, out_divisor);
  return abt::check_launch("resample_homography_kernel");
}

extern "C" int at_crop_pull_source_boxes_u8(const uint8_t* src, uint8_t* dst, int64_t n_frames, int64_t frame_stride,
                                            int64_t view_stride, int64_t row_stride, const AtCropParams* params,
                                            int n_crops, int dst_w, int dst_h, int max_ctas,
                                            unsigned long long* pulled_bytes, void* stream) {
  AT_REQUIRE(n_crops >= 0 && n_frames >= 0, "at_crop_pull_source_boxes_u8: negative count");
  if (n_crops == 0 || n_frames == 0) return AT_OK;
  AT_REQUIRE(src && dst && params, "at_crop_pull_source_boxes_u8: NULL pointer");
  AT_REQUIRE(dst_w > 0 && dst_h > 0 && dst_w <= 4096 && dst_h <= 4096, "at_crop_pull_source_boxes_u8: bad crop size");
  AT_REQUIRE((((uintptr_t)src | (uintptr_t)dst) & 15) == 0 && (frame_stride & 15) == 0 && frame_stride > 0,
             "at_crop_pull_source_boxes_u8: source, destination and frame stride must be 16-byte aligned");
  AT_REQUIRE(n_frames * frame_stride < (1ll << 32) - 16 && row_stride >= 0 && view_stride >= 0,
             "at_crop_pull_source_boxes_u8: at most 4 GiB of frames per call, non-negative strides");
  {
    cudaPointerAttributes attr;
    const bool ok = cudaPointerGetAttributes(&attr, src) == cudaSuccess &&
                    (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged);
    if (!ok) cudaGetLastError();
    AT_REQUIRE(ok, "at_crop_pull_source_boxes_u8: the source must be pinned host memory (or device memory)");
  }
  int grid = max_ctas > 0 ? max_ctas : abt::kNumSMs;
  if (grid > n_crops) grid = n_crops;
  // Bytes in flight = grid x depth x 1 KB.  Measured on B200 (scripts/pull_time.py, bench.py): PCIe
  // saturates (~36-39 GB/s on these short rows) from ~150 KB in flight; every SM-issued PCIe read
  // also slows the co-running encoder a little (scripts/overlap_probe.cu), and more requests in
  // flight only make that worse -> one warp per SM, 1 KB each.  AT_PULL_DEPTH (1, 2, 4, 8) to tune.
  static const int depth = getenv("AT_PULL_DEPTH") ? atoi(getenv("AT_PULL_DEPTH")) : 1;
  static const bool pull_tma = getenv("AT_PULL_TMA") != nullptr && atoi(getenv("AT_PULL_TMA")) != 0;
  if (pull_tma) {
    AT_CARVEOUT_ONCE(crop_pull_boxes_tma_kernel);
    crop_pull_boxes_tma_kernel<<<grid, 32, 0, (cudaStream_t)stream>>>(src, dst, n_frames * frame_stride, frame_stride, view_stride,
                                                                      row_stride, params, n_crops, dst_w, dst_h, pulled_bytes);
    return abt::check_launch("crop_pull_boxes_tma_kernel");
  }
#define AT_PULL_LAUNCH(D)                                                                                         \
  do {                                                                                                            \
    AT_CARVEOUT_ONCE(crop_pull_boxes_kernel<D>);                                                                  \
    crop_pull_boxes_kernel<D><<<grid, 32, 0, (cudaStream_t)stream>>>(src, dst, n_frames * frame_stride,           \
                                                                     frame_stride, view_stride, row_stride, params, \
                                                                     n_crops, dst_w, dst_h, pulled_bytes);        \
  } while (0)
  if (depth == 1) AT_PULL_LAUNCH(1);
  else if (depth == 2) AT_PULL_LAUNCH(2);
  else if (depth == 8) AT_PULL_LAUNCH(8);
  else AT_PULL_LAUNCH(4);
#undef AT_PULL_LAUNCH
  return abt::check_launch("crop_pull_boxes_kernel");
}
