// Hand forward kinematics + linear-blend landmark skinning (float32), one thread per hand.
//
// Replaces lib/common/hand_skinning.py:17-209 (skin_landmarks and helpers) and the right-hand
// mirror of lib/tracker/perspective_crop.py:40-51.  Roofline: HBM-bound, 404 B per hand with
// a shared hand model (22+16 floats in, 63 floats out), 1,688 B with per-sample models.
//
// Layout: the 17 skinning frames (3x4 each) of every hand in the block live in shared memory,
// element-major ([204][kBlock]) so that lanes never conflict; landmarks are staged through the
// same shared memory for fully coalesced 128-bit-friendly stores.
#include <algorithm>

#include "at_common.cuh"
#include "fk_math.cuh"

namespace {

using abt::Xf;
using abt::compose;
using abt::joint_local;

constexpr int kBlock = 64;         // hands (= threads) per CTA
constexpr int kFrameFloats = 12;   // 3x4 rigid frame
constexpr int kNumFrames = 17;     // root + wrist + 3 per finger (hand.py:21)
constexpr int kSmemFloats = kNumFrames * kFrameFloats * kBlock;  // 13056 floats = 51 KB

__device__ __forceinline__ void store_frame(float* __restrict__ s, int frame, const Xf& x) {
  float* p = s + frame * kFrameFloats * kBlock + threadIdx.x;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[(i * 4 + 0) * kBlock] = x.r[i * 3 + 0];
    p[(i * 4 + 1) * kBlock] = x.r[i * 3 + 1];
    p[(i * 4 + 2) * kBlock] = x.r[i * 3 + 2];
    p[(i * 4 + 3) * kBlock] = x.t[i];
  }
}

__global__ void __launch_bounds__(kBlock) fk_skin_kernel(const float* __restrict__ joint_angles,
                                                         const float* __restrict__ wrist,
                                                         const int64_t* __restrict__ hand_idx, AtHandModel m,
                                                         int64_t n, float* __restrict__ landmarks, float tscale,
                                                         float* __restrict__ wrist_out) {
  extern __shared__ float smem[];
  const int64_t base = (int64_t)blockIdx.x * kBlock;
  const int64_t h = base + threadIdx.x;
  const bool active = h < n;
  const int64_t hm = m.per_sample ? h : 0;

  float out[63];
  if (active) {
    const float* ja = joint_angles + h * 22;
    const float* wp = wrist + h * 16;
    const float* axes = m.joint_rotation_axes + hm * 66;
    const float* rest = m.joint_rest_positions + hm * 66;
    Xf W;
    const float sgn = (hand_idx != nullptr && hand_idx[h] == 1) ? -1.0f : 1.0f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      W.r[i * 3 + 0] = sgn * wp[i * 4 + 0];
      W.r[i * 3 + 1] = wp[i * 4 + 1];
      W.r[i * 3 + 2] = wp[i * 4 + 2];
      W.t[i] = wp[i * 4 + 3] * tscale;
    }
    if (wrist_out != nullptr) {  // the input transform with its translation in the output unit
      float* wo = wrist_out + h * 16;
#pragma unroll
      for (int i = 0; i < 16; ++i) wo[i] = (i < 12 && (i & 3) == 3) ? wp[i] * tscale : wp[i];
    }
    store_frame(smem, 0, W);
    store_frame(smem, 1, W);
#pragma unroll 1
    for (int f = 0; f < 5; ++f) {
      Xf cur = W;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int j = 4 * f + i;
        Xf T, nxt;
        joint_local(axes + 3 * j, rest + 3 * j, ja[j], T);
        compose(cur, T, nxt);
        cur = nxt;
        if (i >= 1) store_frame(smem, 2 + 3 * f + (i - 1), cur);
      }
    }
    // skinning: sum_k F[idx_k] * (w_k * [p,1])  (zero weights dropped, hand_skinning.py:92-94)
    const int K = m.n_weights;
    const float* lm = m.landmark_rest_positions + hm * 63;
    const float* bw = m.landmark_rest_bone_weights + hm * 21 * K;
    const int32_t* bi = m.landmark_rest_bone_indices + hm * 21 * K;
#pragma unroll
    for (int l = 0; l < 21; ++l) {
      const float px = lm[3 * l], py = lm[3 * l + 1], pz = lm[3 * l + 2];
      float ax = 0.f, ay = 0.f, az = 0.f;
      for (int k = 0; k < K; ++k) {
        const float w = bw[l * K + k];
        if (w != 0.0f) {
          int b = bi[l * K + k];
          b = min(max(b, 0), kNumFrames - 1);
          const float* F = smem + b * kFrameFloats * kBlock + threadIdx.x;
          const float qx = px * w, qy = py * w, qz = pz * w;
          ax += F[0 * kBlock] * qx + F[1 * kBlock] * qy + F[2 * kBlock] * qz + F[3 * kBlock] * w;
          ay += F[4 * kBlock] * qx + F[5 * kBlock] * qy + F[6 * kBlock] * qz + F[7 * kBlock] * w;
          az += F[8 * kBlock] * qx + F[9 * kBlock] * qy + F[10 * kBlock] * qz + F[11 * kBlock] * w;
        }
      }
      out[3 * l] = ax;
      out[3 * l + 1] = ay;
      out[3 * l + 2] = az;
    }
  }
  __syncthreads();  // all frame reads of the block are done; reuse smem for output staging
  if (active) {
    float* so = smem + threadIdx.x * 63;  // stride 63 (odd): conflict-free
#pragma unroll
    for (int i = 0; i < 63; ++i) so[i] = out[i];
  }
  __syncthreads();
  const int64_t n_here = min((int64_t)kBlock, n - base);
  const int total = (int)n_here * 63;
  float* dst = landmarks + base * 63;
  for (int i = threadIdx.x; i < total; i += kBlock) dst[i] = smem[i];
}


// ------------------------------------------------------------------------------------------
// Per-sample hand models (config 3: every sample brings its own skeleton, 1,284 B of model per hand next to 404 B
// of pose in / landmarks out): one WARP per hand.  With one thread per hand (fk_skin_kernel) each thread walks its
// own 1.3 KB of model with 4-byte loads 1.3 KB apart from its neighbour's - 0.11 of HBM.  Here lane j < 20 loads
// joint j (axis, rest position, angle: the warp's loads are contiguous) and builds its local transform (MUFU rsqrt /
// sin / cos as in the shared-model path), so the 20 Rodrigues formulas of a hand run in parallel; the chain products
// T_0 ... T_i of every finger are a two-round prefix scan over its four lanes (shuffles), then one product with the
// wrist; lane l < 21 skins landmark l, fetching the frames of its bones from the lanes that hold them, and stores its
// three coordinates (252 contiguous bytes per warp).  No shared memory.  (Products are associated differently from
// hand_skinning.py:35-53's left-to-right loop: ~1e-4 mm at most, tests/test_gpu_kernels.py bounds it at 1e-3 mm.)
constexpr int kWarpFkThreads = 256;

__device__ __forceinline__ void shfl_xf(const Xf& v, int src, Xf& o) {
#pragma unroll
  for (int e = 0; e < 9; ++e) o.r[e] = __shfl_sync(0xffffffffu, v.r[e], src);
#pragma unroll
  for (int e = 0; e < 3; ++e) o.t[e] = __shfl_sync(0xffffffffu, v.t[e], src);
}

template <int KMAX>   // upper bound of n_weights the loops are unrolled for (4: the usual 3-weight models; 8: the maximum)
__global__ void __launch_bounds__(kWarpFkThreads, 4) fk_skin_warp_kernel(const float* __restrict__ joint_angles,
                                                                         const float* __restrict__ wrist,
                                                                         const int64_t* __restrict__ hand_idx, AtHandModel m,
                                                                         int64_t n, float* __restrict__ landmarks, float tscale,
                                                                         float* __restrict__ wrist_out) {
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = (int64_t)gridDim.x * (kWarpFkThreads / 32);
  const int K = m.n_weights;
  const int pos = lane & 3;  // position of joint `lane` in its finger chain
  for (int64_t h = (int64_t)blockIdx.x * (kWarpFkThreads / 32) + (threadIdx.x >> 5); h < n; h += nwarps) {
    const int64_t hm = m.per_sample ? h : 0;
    const float* wp = wrist + h * 16;
    Xf W;
    const float sgn = (hand_idx != nullptr && hand_idx[h] == 1) ? -1.0f : 1.0f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      W.r[i * 3 + 0] = sgn * wp[i * 4 + 0];
      W.r[i * 3 + 1] = wp[i * 4 + 1];
      W.r[i * 3 + 2] = wp[i * 4 + 2];
      W.t[i] = wp[i * 4 + 3] * tscale;
    }
    if (wrist_out != nullptr && lane < 16)
      wrist_out[h * 16 + lane] = (lane < 12 && (lane & 3) == 3) ? wp[lane] * tscale : wp[lane];
    // everything this lane reads from global memory is requested up front (one exposed latency per hand)
    const int l = lane < 21 ? lane : 0;
    const float* lm = m.landmark_rest_positions + hm * 63 + 3 * l;
    const float px = lm[0], py = lm[1], pz = lm[2];
    const float* bw = m.landmark_rest_bone_weights + (hm * 21 + l) * K;
    const int32_t* bi = m.landmark_rest_bone_indices + (hm * 21 + l) * K;
    float wk[KMAX];
    int bk[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      wk[k] = k < K ? bw[k] : 0.0f;
      bk[k] = k < K ? bi[k] : 0;
    }
    const int j = lane < 20 ? lane : 0;
    Xf T;
    {
      const float* axis = m.joint_rotation_axes + hm * 66 + 3 * j;
      const float* rest = m.joint_rest_positions + hm * 66 + 3 * j;
      const float angle = joint_angles[h * 22 + j];
      const float wx = axis[0] * angle, wy = axis[1] * angle, wz = axis[2] * angle;
      const float th2 = fmaxf(wx * wx + wy * wy + wz * wz, 1e-4f);   // so3_exp_map's clamp (hand_skinning.py:46)
      const float inv = rsqrtf(th2);
      float sn, cs;
      __sincosf(th2 * inv, &sn, &cs);
      const float f1 = inv * sn, f2 = inv * inv * (1.0f - cs);
      T.r[0] = 1.0f - f2 * (wy * wy + wz * wz); T.r[1] = f2 * (wx * wy) - f1 * wz;        T.r[2] = f2 * (wx * wz) + f1 * wy;
      T.r[3] = f2 * (wx * wy) + f1 * wz;        T.r[4] = 1.0f - f2 * (wx * wx + wz * wz); T.r[5] = f2 * (wy * wz) - f1 * wx;
      T.r[6] = f2 * (wx * wz) - f1 * wy;        T.r[7] = f2 * (wy * wz) + f1 * wx;        T.r[8] = 1.0f - f2 * (wx * wx + wy * wy);
      const float px = rest[0], py = rest[1], pz = rest[2];
      T.t[0] = px - (T.r[0] * px + T.r[1] * py + T.r[2] * pz);
      T.t[1] = py - (T.r[3] * px + T.r[4] * py + T.r[5] * pz);
      T.t[2] = pz - (T.r[6] * px + T.r[7] * py + T.r[8] * pz);
    }
    // chains: inclusive prefix products over the four lanes of a finger, then the wrist in front
    Xf cur, up;
    shfl_xf(T, (lane + 31) & 31, up);
    if (pos >= 1) compose(up, T, cur); else cur = T;
    shfl_xf(cur, (lane + 30) & 31, up);
    if (pos >= 2) { Xf t2; compose(up, cur, t2); cur = t2; }
    { Xf t3; compose(W, cur, t3); cur = t3; }
    // skinning: sum_k F[idx_k] * (w_k * [p,1])  (zero weights dropped, hand_skinning.py:92-94)
    float ax = 0.f, ay = 0.f, az = 0.f;
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      if (k >= K) break;
      const float w = wk[k];
      const int b = min(max(bk[k], 0), kNumFrames - 1);
      // frames 0, 1 = the wrist; frame 2 + 3 f + (i - 1), i = 1..3, is held by lane 4 f + i
      const int fq = ((b - 2) * 11) >> 5;   // (b - 2) / 3 for b - 2 in [0, 14]
      const int src = b < 2 ? lane : 4 * fq + (b - 2 - 3 * fq) + 1;
      Xf F;
      shfl_xf(cur, src, F);
      if (b < 2) F = W;
      if (w != 0.0f) {
        const float qx = px * w, qy = py * w, qz = pz * w;
        ax += F.r[0] * qx + F.r[1] * qy + F.r[2] * qz + F.t[0] * w;
        ay += F.r[3] * qx + F.r[4] * qy + F.r[5] * qz + F.t[1] * w;
        az += F.r[6] * qx + F.r[7] * qy + F.r[8] * qz + F.t[2] * w;
      }
    }
    if (lane < 21) {
      float* o = landmarks + h * 63 + 3 * lane;
      o[0] = ax; o[1] = ay; o[2] = az;
    }
  }
}

// ------------------------------------------------------------------------------------------
// Shared-hand-model fast path (eval scripts, frame-stream tracker): the model is uniform across the
// CTA, so per joint the constant parts of the Rodrigues formula are computed once per CTA:
//   R = I + s1*K + s2*K^2,  t = p - R p = -(s1*K p + s2*K^2 p),  K = skew(axis),
//   s1 = angle*sin(th)/th,  s2 = angle^2*(1-cos(th))/th^2,  th = sqrt(max(|axis|^2 angle^2, 1e-4))
// (identical to so3_exp_map(axis*angle) of hand_skinning.py:46).  The 17 frames never leave
// registers: landmarks are visited frame by frame through a CSR table built once per CTA
// (bone index -> landmarks) and accumulated in shared memory, which doubles as the staging buffer
// for 16-byte input loads and output stores.  252 B of shared memory and 80 registers per hand -> 768 hands per
// SM, run as 8 CTAs of 96 (small CTAs: while one copies its tile in or out, seven compute; measured
// 0.438 / 0.456 / 0.472 / 0.477 of HBM at 3 x 256 / 4 x 192 / 6 x 128 / 8 x 96).
constexpr int kFast = 96;   // hands per CTA (>= 96: threads 0..19 and the last 64 build the tables)
constexpr int kFastCtas = 8; // per SM: 8 x 24.2 KB of shared memory, 8 x 96 x 80 registers
// shared buffer: first the flat [hands][22] joint-angle stage, then the flat [hands][63] landmark
// accumulator = the exact global layout, so both copies are straight 16-byte moves; the per-thread
// rows have odd (63) / 8-byte-friendly (22) strides -> conflict-free LDS.64 / STS.32
constexpr int kOutFloats = kFast * 63;

struct JointConst {  // 24 floats, 16-byte aligned: K (3), K^2 (6 sym), K p (3), K^2 p (3), |axis|^2
  float4 v[6];
};

struct FastTables {
  JointConst jc[20];
  float4 ent_q[64];        // (w p, w) of entry: out (+)= R (w p) + w t
  int ent_off[64];         // 3 * landmark; bit 31 set: first contribution to that landmark (store)
  int2 fs[kNumFrames];     // CSR row of frame b: entries [x, y)
  int lm_has[21];
  int n_missing;           // landmarks without any non-zero weight (normally 0: the zero-fill loop is skipped)
};

// every landmark attached to frame b: out[l] (+)= w * (F [p,1])   (hand_skinning._lbs :56-67)
__device__ __forceinline__ void emit_frame(const FastTables& tb, float* __restrict__ row, int b, const Xf& F) {
  const int2 span = tb.fs[b];  // one LDS.64
  for (int e = span.x; e < span.y; ++e) {
    const float4 q = tb.ent_q[e];
    const int off = tb.ent_off[e];
    const float ox = fmaf(F.r[0], q.x, fmaf(F.r[1], q.y, fmaf(F.r[2], q.z, F.t[0] * q.w)));
    const float oy = fmaf(F.r[3], q.x, fmaf(F.r[4], q.y, fmaf(F.r[5], q.z, F.t[1] * q.w)));
    const float oz = fmaf(F.r[6], q.x, fmaf(F.r[7], q.y, fmaf(F.r[8], q.z, F.t[2] * q.w)));
    float* o = row + (off & 0xff);
    if (off < 0) {
      o[0] = ox; o[1] = oy; o[2] = oz;
    } else {
      o[0] += ox; o[1] += oy; o[2] += oz;
    }
  }
}

__global__ void __launch_bounds__(kFast, kFastCtas) fk_skin_shared_kernel(const float* __restrict__ joint_angles,
                                                                  const float* __restrict__ wrist,
                                                                  const int64_t* __restrict__ hand_idx, AtHandModel m,
                                                                  int64_t n, float* __restrict__ landmarks, float tscale,
                                                                  float* __restrict__ wrist_out) {
  abt::pdl_prologue();
  extern __shared__ __align__(16) float smem[];
  __shared__ FastTables tb;
  __shared__ int s_b[64];
  __shared__ int s_valid[64];
  const int tid = threadIdx.x;
  const int K = m.n_weights;
  const int npairs = 21 * K;  // <= 63 for K <= 3 (host checks)

  // ---- per-CTA tables
  if (tid < 20) {
    const float ax = m.joint_rotation_axes[3 * tid], ay = m.joint_rotation_axes[3 * tid + 1], az = m.joint_rotation_axes[3 * tid + 2];
    const float px = m.joint_rest_positions[3 * tid], py = m.joint_rest_positions[3 * tid + 1], pz = m.joint_rest_positions[3 * tid + 2];
    const float a2 = ax * ax + ay * ay + az * az;
    // K^2 = a a^T - |a|^2 I (symmetric): xx, yy, zz, xy, xz, yz
    const float kxx = -(ay * ay + az * az), kyy = -(ax * ax + az * az), kzz = -(ax * ax + ay * ay);
    const float kxy = ax * ay, kxz = ax * az, kyz = ay * az;
    const float kpx = ay * pz - az * py, kpy = az * px - ax * pz, kpz = ax * py - ay * px;  // K p = a x p
    const float k2x = kxx * px + kxy * py + kxz * pz, k2y = kxy * px + kyy * py + kyz * pz, k2z = kxz * px + kyz * py + kzz * pz;
    JointConst& J = tb.jc[tid];
    J.v[0] = make_float4(ax, ay, az, a2);
    J.v[1] = make_float4(kxx, kyy, kzz, 0.f);
    J.v[2] = make_float4(kxy, kxz, kyz, 0.f);
    J.v[3] = make_float4(kpx, kpy, kpz, 0.f);
    J.v[4] = make_float4(k2x, k2y, k2z, 0.f);
    J.v[5] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (tid >= kFast - 64) {  // the last 64 threads build the CSR table (threads 0..19: joint constants)
    const int e = tid - (kFast - 64);
    int valid = 0, b = 0;
    if (e < npairs) {
      valid = m.landmark_rest_bone_weights[e] != 0.0f;   // zero weights dropped (hand_skinning.py:92-94)
      b = min(max(m.landmark_rest_bone_indices[e], 0), kNumFrames - 1);
    }
    s_b[e] = b;
    s_valid[e] = valid;
  }
  __syncthreads();
  if (tid >= kFast - 64) {  // the last 64 threads build the CSR table (threads 0..19: joint constants)
    const int e = tid - (kFast - 64);
    if (e < npairs && s_valid[e]) {
      const int b = s_b[e], l = e / K;
      int pos = 0, first = 1;
      for (int j = 0; j < npairs; ++j) {
        if (!s_valid[j]) continue;
        const bool before = s_b[j] < b || (s_b[j] == b && j < e);
        pos += before;
        if (before && j / K == l) first = 0;
      }
      const float w = m.landmark_rest_bone_weights[e];
      const float* lp = m.landmark_rest_positions + 3 * l;
      tb.ent_q[pos] = make_float4(lp[0] * w, lp[1] * w, lp[2] * w, w);
      tb.ent_off[pos] = (3 * l) | (first ? (int)0x80000000 : 0);
    }
    if (e < kNumFrames) {
      int lo = 0, hi = 0;
      for (int j = 0; j < npairs; ++j) {
        lo += s_valid[j] && s_b[j] < e;
        hi += s_valid[j] && s_b[j] <= e;
      }
      tb.fs[e] = make_int2(lo, hi);
    }
    if (e < 21) {
      int any = 0;
      for (int k = 0; k < K; ++k) any |= s_valid[e * K + k];
      tb.lm_has[e] = any;
    }
    if (e == 21) {
      int miss = 0;
      for (int l = 0; l < 21; ++l) {
        int any = 0;
        for (int k = 0; k < K; ++k) any |= s_valid[l * K + k];
        miss += !any;
      }
      tb.n_missing = miss;
    }
  }
  // tile bases are multiples of 16 B (96 hands x 88 / 64 / 252 B), so alignment depends on the pointers only
  const bool vec_ok = (((uintptr_t)joint_angles | (uintptr_t)wrist | (uintptr_t)landmarks) & 15) == 0;
  const bool wo_vec = ((uintptr_t)wrist_out & 15) == 0;
  // ---- persistent loop over tiles of kFast hands (the tables above are built once per CTA)
  const int64_t n_tiles = (n + kFast - 1) / kFast;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
  const int64_t base = tile * kFast;
  const int n_here = (int)min((int64_t)kFast, n - base);
  // the wrist rows (64 B per hand, used once) come straight from global memory as three float4,
  // issued first so that their latency overlaps the joint-angle staging below
  const bool active = tid < n_here;
  Xf W;
  if (active) {
    const float sgn = (hand_idx != nullptr && hand_idx[base + tid] == 1) ? -1.0f : 1.0f;
    const float* gw = wrist + (base + tid) * 16;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      float4 r;
      if (vec_ok) r = __ldg(reinterpret_cast<const float4*>(gw) + i);
      else r = make_float4(__ldg(gw + 4 * i), __ldg(gw + 4 * i + 1), __ldg(gw + 4 * i + 2), __ldg(gw + 4 * i + 3));
      r.w *= tscale;
      W.r[i * 3 + 0] = sgn * r.x; W.r[i * 3 + 1] = r.y; W.r[i * 3 + 2] = r.z; W.t[i] = r.w;
      if (wrist_out != nullptr) {
        float* wo = wrist_out + (base + tid) * 16 + 4 * i;
        if (wo_vec) *reinterpret_cast<float4*>(wo) = r;
        else { wo[0] = r.x; wo[1] = r.y; wo[2] = r.z; wo[3] = r.w; }
      }
    }
    if (wrist_out != nullptr) {
      float* wo = wrist_out + (base + tid) * 16 + 12;
      wo[0] = __ldg(gw + 12); wo[1] = __ldg(gw + 13); wo[2] = __ldg(gw + 14); wo[3] = __ldg(gw + 15);
    }
  }
  // coalesced input staging: the tile's joint angles [n_here,22] as one flat 16-byte copy
  {
    const float* gja = joint_angles + base * 22;
    const int tot = n_here * 22;
    int done = 0;
    if (vec_ok) {
      const int nv = tot >> 2;
      const float4* g4 = reinterpret_cast<const float4*>(gja);
      float4* s4 = reinterpret_cast<float4*>(smem);
      for (int i = tid; i < nv; i += kFast) s4[i] = __ldg(g4 + i);
      done = nv << 2;
    }
    for (int i = done + tid; i < tot; i += kFast) smem[i] = __ldg(gja + i);
  }
  __syncthreads();
  float ang[20];
  if (active) {
    const float2* a2 = reinterpret_cast<const float2*>(smem + tid * 22);
#pragma unroll
    for (int j = 0; j < 10; ++j) {
      const float2 v = a2[j];
      ang[2 * j] = v.x; ang[2 * j + 1] = v.y;
    }
  }
  __syncthreads();  // inputs are in registers: the buffer becomes the flat [hands][63] landmark accumulator
  float* row = smem + tid * 63;
  if (active) {
    if (tb.n_missing)
      for (int l = 0; l < 21; ++l)
        if (!tb.lm_has[l]) { row[3 * l] = 0.f; row[3 * l + 1] = 0.f; row[3 * l + 2] = 0.f; }
    emit_frame(tb, row, 0, W);
    emit_frame(tb, row, 1, W);
#pragma unroll 1  // (fully unrolled: 0.414 instead of 0.439 of HBM at 4M hands - instruction cache)
    for (int f = 0; f < 5; ++f) {
      Xf cur = W;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const JointConst& J = tb.jc[4 * f + i];
        const float4 c0 = J.v[0], c1 = J.v[1], c2 = J.v[2], c3 = J.v[3], c4 = J.v[4];
        const float a = ang[i];  // the angles of the current finger are rotated into ang[0..3]
        // th = sqrt(max(|w|^2, 1e-4)) in [0.01, ~5]: MUFU rsqrt / sin / cos are accurate to ~4e-7
        // absolute there, far inside the 1e-3 mm landmark tolerance (tests/test_gpu_kernels.py)
        const float th2 = fmaxf(c0.w * a * a, 1e-4f);
        const float inv = rsqrtf(th2);
        const float th = th2 * inv;
        float sn, cs;
        __sincosf(th, &sn, &cs);
        const float s1 = a * (inv * sn);
        const float s2 = (a * a) * (inv * inv * (1.0f - cs));
        Xf T, nxt;
        T.r[0] = 1.0f + s2 * c1.x;        T.r[1] = s2 * c2.x - s1 * c0.z;  T.r[2] = s2 * c2.y + s1 * c0.y;
        T.r[3] = s2 * c2.x + s1 * c0.z;   T.r[4] = 1.0f + s2 * c1.y;       T.r[5] = s2 * c2.z - s1 * c0.x;
        T.r[6] = s2 * c2.y - s1 * c0.y;   T.r[7] = s2 * c2.z + s1 * c0.x;  T.r[8] = 1.0f + s2 * c1.z;
        T.t[0] = -(s1 * c3.x + s2 * c4.x);
        T.t[1] = -(s1 * c3.y + s2 * c4.y);
        T.t[2] = -(s1 * c3.z + s2 * c4.z);
        compose(cur, T, nxt);
        cur = nxt;
        if (i >= 1) emit_frame(tb, row, 2 + 3 * f + (i - 1), cur);
      }
#pragma unroll
      for (int q = 0; q < 16; ++q) ang[q] = ang[q + 4];
    }
  }
  __syncthreads();
  // the accumulator already has the global layout: one flat 16-byte copy
  {
    float* dst = landmarks + base * 63;
    const int tot = n_here * 63;
    int done = 0;
    if (vec_ok) {
      const int nv = tot >> 2;
      float4* d4 = reinterpret_cast<float4*>(dst);
      const float4* s4 = reinterpret_cast<const float4*>(smem);
      for (int i = tid; i < nv; i += kFast) d4[i] = s4[i];
      done = nv << 2;
    }
    for (int i = done + tid; i < tot; i += kFast) dst[i] = smem[i];
  }
  __syncthreads();  // the buffer is reused as the input stage of the next tile
  }
}

// ------------------------------------------------------------------------------------------
// Evaluation sums (run_eval_known_skeleton.py:92-93 / load_eval.py:26-44 bookkeeping): one warp per
// hand, lanes = landmarks / joints; per-CTA partial sums, one fp64 atomic per CTA and quantity.
constexpr int kEvalWarps = 8;
__global__ void __launch_bounds__(kEvalWarps * 32)
eval_accumulate_kernel(const float* __restrict__ lm, const float* __restrict__ gt, const float* __restrict__ ja,
                       const float* __restrict__ gja, const uint8_t* __restrict__ valid, int64_t n, double* __restrict__ sums) {
  abt::pdl_prologue();
  __shared__ float s_part[kEvalWarps][3];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float e_sum = 0.f, a_sum = 0.f, cnt = 0.f;
  for (int64_t h = (int64_t)blockIdx.x * kEvalWarps + warp; h < n; h += (int64_t)gridDim.x * kEvalWarps) {
    if (valid != nullptr && valid[h] == 0) continue;
    float d = 0.f, a = 0.f;
    if (lane < 21) {
      const float* p = lm + h * 63 + lane * 3;
      const float* q = gt + h * 63 + lane * 3;
      const float dx = p[0] - q[0], dy = p[1] - q[1], dz = p[2] - q[2];
      d = sqrtf(dx * dx + dy * dy + dz * dz);
    }
    if (ja != nullptr && lane < 20) a = fabsf(ja[h * 22 + lane] - gja[h * 22 + lane]);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      d += __shfl_xor_sync(0xffffffffu, d, o);
      a += __shfl_xor_sync(0xffffffffu, a, o);
    }
    e_sum += d * (1.0f / 21.0f);
    a_sum += a * (57.29577951308232f / 20.0f);
    cnt += 1.0f;
  }
  if (lane == 0) { s_part[warp][0] = e_sum; s_part[warp][1] = a_sum; s_part[warp][2] = cnt; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < kEvalWarps; ++w) t += (double)s_part[w][threadIdx.x];
    if (t != 0.0) atomicAdd(sums + threadIdx.x, t);
  }
}


// ------------------------------------------------------------------------------------------
// Evaluation table of one recording (load_eval._compute_metrics :26-44 and the PCK counts of
// metric_utils._PCK_curve :34-44) as sums that add up across recordings and ranks.  One warp per
// hand-frame, lanes = landmarks.  float64 throughout like the reference's numpy arrays, and every
// sum is taken in numpy's order (norm: (x^2 + y^2) + z^2; mean over 21: the 8-lane pairwise block
// of numpy's add.reduce, then the 5 remaining terms), so a hand-frame's error is bit-identical to
// the reference's and the integer PCK counts are exact.
constexpr int kEvalMaxThresholds = 128;

__device__ __forceinline__ double norm3_np(double x, double y, double z) {
  return sqrt(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)));
}

// numpy's pairwise add.reduce of 21 contiguous doubles held by lanes 0..20; result valid in lane 0
__device__ __forceinline__ double sum21_np(double a, int lane) {
  const unsigned full = 0xffffffffu;
  double r = __dadd_rn(a, __shfl_down_sync(full, a, 8));       // r[j] = a[j] + a[j+8], j < 8
  r = __dadd_rn(r, __shfl_down_sync(full, r, 1));              // r0+r1, r2+r3, ...
  r = __dadd_rn(r, __shfl_down_sync(full, r, 2));              // (r0+r1)+(r2+r3), (r4+r5)+(r6+r7)
  r = __dadd_rn(r, __shfl_down_sync(full, r, 4));
#pragma unroll
  for (int i = 16; i < 21; ++i) r = __dadd_rn(r, __shfl_sync(full, a, i));
  (void)lane;
  return r;
}

__global__ void __launch_bounds__(256)
eval_recording_kernel(const float* __restrict__ tracked, const float* __restrict__ gt, const uint8_t* __restrict__ valid,
                      int64_t n_hands, int64_t T, const double* __restrict__ thresholds, int n_thr,
                      double* __restrict__ sums, unsigned long long* __restrict__ counts) {
  __shared__ double s_sum[3];
  __shared__ unsigned long long s_cnt[3 + kEvalMaxThresholds];
  for (int i = threadIdx.x; i < 3 + kEvalMaxThresholds; i += blockDim.x) s_cnt[i] = 0ull;
  if (threadIdx.x < 3) s_sum[threadIdx.x] = 0.0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t n = n_hands * T;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  double thr[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) thr[k] = (lane + 32 * k < n_thr) ? thresholds[lane + 32 * k] : -1.0;
  double se = 0.0, sa = 0.0, sg = 0.0;
  unsigned nv = 0, na = 0, c[4] = {0, 0, 0, 0};
  for (int64_t idx = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); idx < n; idx += warps) {
    if (!valid[idx]) continue;  // warp-uniform
    const int64_t t = idx % T;
    const bool tri = t + 2 < T && valid[idx + 1] && valid[idx + 2];  // frames t, t+1, t+2 of the same hand
    double d = 0.0, a = 0.0, g = 0.0;
    if (lane < 21) {
      const float* p = tracked + (idx * 21 + lane) * 3;
      const float* q = gt + (idx * 21 + lane) * 3;
      d = norm3_np((double)q[0] - (double)p[0], (double)q[1] - (double)p[1], (double)q[2] - (double)p[2]);
      if (tri) {
        double v[3], w[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {  // pts[t] + pts[t+2] - 2 pts[t+1]  (load_eval.py:30)
          v[k] = __dadd_rn(__dadd_rn((double)p[k], (double)p[k + 126]), -2.0 * (double)p[k + 63]);
          w[k] = __dadd_rn(__dadd_rn((double)q[k], (double)q[k + 126]), -2.0 * (double)q[k + 63]);
        }
        a = norm3_np(v[0], v[1], v[2]);
        g = norm3_np(w[0], w[1], w[2]);
      }
    }
    const double e = __shfl_sync(0xffffffffu, sum21_np(d, lane), 0) / 21.0;
    se += e;
    nv += 1;
    if (tri) {
      sa += sum21_np(a, lane) / 21.0;
      sg += sum21_np(g, lane) / 21.0;
      na += 1;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) c[k] += (e <= thr[k]) ? 1u : 0u;  // metric_utils.py:40 (thr = -1: never)
  }
  if (lane == 0) {
    atomicAdd(&s_sum[0], se); atomicAdd(&s_sum[1], sa); atomicAdd(&s_sum[2], sg);
    atomicAdd(&s_cnt[0], (unsigned long long)nv); atomicAdd(&s_cnt[2], (unsigned long long)na);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (c[k]) atomicAdd(&s_cnt[3 + lane + 32 * k], (unsigned long long)c[k]);
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) s_cnt[1] = (unsigned long long)n;
  __syncthreads();
  if (threadIdx.x < 3 && s_sum[threadIdx.x] != 0.0) atomicAdd(sums + threadIdx.x, s_sum[threadIdx.x]);
  for (int i = threadIdx.x; i < 3 + n_thr; i += blockDim.x)
    if (s_cnt[i]) atomicAdd(counts + i, s_cnt[i]);
}

}  // namespace

extern "C" int at_eval_recording_f32(const float* tracked_keypoints, const float* gt_keypoints, const uint8_t* valid_tracking,
                                     int64_t n_hands, int64_t n_frames, const double* thresholds, int n_thresholds,
                                     double* sums, int64_t* counts, void* stream) {
  AT_REQUIRE(n_hands >= 0 && n_frames >= 0, "at_eval_recording_f32: negative size");
  AT_REQUIRE(n_thresholds >= 0 && n_thresholds <= kEvalMaxThresholds, "at_eval_recording_f32: at most 128 thresholds");
  if (n_hands * n_frames == 0) return AT_OK;
  AT_REQUIRE(tracked_keypoints && gt_keypoints && valid_tracking && sums && counts, "at_eval_recording_f32: NULL pointer");
  AT_REQUIRE(n_thresholds == 0 || thresholds != nullptr, "at_eval_recording_f32: thresholds is NULL");
  const int64_t n = n_hands * n_frames;
  const int64_t blocks = std::min<int64_t>((n + 7) / 8, 4 * abt::kNumSMs);
  eval_recording_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      tracked_keypoints, gt_keypoints, valid_tracking, n_hands, n_frames, thresholds, n_thresholds, sums,
      reinterpret_cast<unsigned long long*>(counts));
  return abt::check_launch("eval_recording_kernel");
}

extern "C" int at_eval_accumulate_f32(const float* landmarks, const float* gt_landmarks, const float* joint_angles,
                                      const float* gt_joint_angles, const uint8_t* valid, int64_t n, double* sums,
                                      void* stream) {
  AT_REQUIRE(n >= 0, "at_eval_accumulate_f32: n < 0");
  if (n == 0) return AT_OK;
  AT_REQUIRE(landmarks && gt_landmarks && sums, "at_eval_accumulate_f32: NULL pointer");
  AT_REQUIRE((joint_angles == nullptr) == (gt_joint_angles == nullptr), "at_eval_accumulate_f32: give both joint-angle arrays or none");
  const int64_t blocks = std::min<int64_t>((n + kEvalWarps - 1) / kEvalWarps, 4 * abt::kNumSMs);
  abt::launch_pdl(eval_accumulate_kernel, dim3((unsigned)blocks), dim3(kEvalWarps * 32), 0, (cudaStream_t)stream, landmarks, gt_landmarks, joint_angles,
                                                                                        gt_joint_angles, valid, n, sums);
  return abt::check_launch("eval_accumulate_kernel");
}

extern "C" int at_fk_skin_scaled_f32(const float* joint_angles, const float* wrist_xforms, const int64_t* hand_idx,
                                     const AtHandModel* model, int64_t n, float translation_scale, float* wrist_out,
                                     float* landmarks, void* stream);

extern "C" int at_fk_skin_f32(const float* joint_angles, const float* wrist_xforms, const int64_t* hand_idx,
                              const AtHandModel* model, int64_t n, float* landmarks, void* stream) {
  return at_fk_skin_scaled_f32(joint_angles, wrist_xforms, hand_idx, model, n, 1.0f, nullptr, landmarks, stream);
}

extern "C" int at_fk_skin_scaled_f32(const float* joint_angles, const float* wrist_xforms, const int64_t* hand_idx,
                                     const AtHandModel* model, int64_t n, float translation_scale, float* wrist_out,
                                     float* landmarks, void* stream) {
  AT_REQUIRE(model != nullptr, "at_fk_skin_f32: model is NULL");
  AT_REQUIRE(n >= 0, "at_fk_skin_f32: n < 0");
  if (n == 0) return AT_OK;
  AT_REQUIRE(joint_angles && wrist_xforms && landmarks, "at_fk_skin_f32: NULL data pointer");
  AT_REQUIRE(model->joint_rotation_axes && model->joint_rest_positions && model->landmark_rest_positions &&
                 model->landmark_rest_bone_weights && model->landmark_rest_bone_indices,
             "at_fk_skin_f32: NULL hand-model pointer");
  AT_REQUIRE(model->n_weights >= 1 && model->n_weights <= 8, "at_fk_skin_f32: n_weights out of range");
  const int smem_bytes = kSmemFloats * (int)sizeof(float);
  AT_ONCE_PER_DEVICE(cudaFuncSetAttribute(fk_skin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  if (!model->per_sample && model->n_weights <= 3) {
    constexpr int fast_smem = kOutFloats * (int)sizeof(float);
    AT_ONCE_PER_DEVICE(
        cudaFuncSetAttribute(fk_skin_shared_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, fast_smem);
        cudaFuncSetAttribute(fk_skin_shared_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    const int64_t fb = std::min<int64_t>((n + kFast - 1) / kFast, kFastCtas * abt::kNumSMs);  // persistent
    abt::launch_pdl(fk_skin_shared_kernel, dim3((unsigned)fb), dim3(kFast), fast_smem, (cudaStream_t)stream, 
        joint_angles, wrist_xforms, hand_idx, *model, n, landmarks, translation_scale, wrist_out);
    return abt::check_launch("fk_skin_shared_kernel");
  }
  // per-sample models: one warp per hand (AT_FK_WARP=0: the one-thread-per-hand kernel, for A/B)
  static const bool warp_fk = getenv("AT_FK_WARP") == nullptr || atoi(getenv("AT_FK_WARP")) != 0;
  if (model->per_sample && warp_fk) {
    const int64_t wb = std::min<int64_t>((n + kWarpFkThreads / 32 - 1) / (kWarpFkThreads / 32), 24LL * abt::kNumSMs);
    if (model->n_weights <= 4)
      fk_skin_warp_kernel<4><<<(unsigned)wb, kWarpFkThreads, 0, (cudaStream_t)stream>>>(joint_angles, wrist_xforms, hand_idx, *model, n,
                                                                                      landmarks, translation_scale, wrist_out);
    else
      fk_skin_warp_kernel<8><<<(unsigned)wb, kWarpFkThreads, 0, (cudaStream_t)stream>>>(joint_angles, wrist_xforms, hand_idx, *model, n,
                                                                                      landmarks, translation_scale, wrist_out);
    return abt::check_launch("fk_skin_warp_kernel");
  }
  const int64_t blocks = (n + kBlock - 1) / kBlock;
  AT_REQUIRE(blocks <= 0x7fffffffLL, "at_fk_skin_f32: n too large");
  fk_skin_kernel<<<(unsigned)blocks, kBlock, smem_bytes, (cudaStream_t)stream>>>(
      joint_angles, wrist_xforms, hand_idx, *model, n, landmarks, translation_scale, wrist_out);
  return abt::check_launch("fk_skin_kernel");
}
