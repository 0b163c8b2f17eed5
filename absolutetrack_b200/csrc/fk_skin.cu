// Hand forward kinematics + linear-blend landmark skinning (float32), one thread per hand.
//
// Replaces lib/common/hand_skinning.py:17-209 (skin_landmarks and helpers) and the right-hand
// mirror of lib/tracker/perspective_crop.py:40-51.  Roofline: HBM-bound, 404 B per hand with
// a shared hand model (22+16 floats in, 63 floats out), 1,688 B with per-sample models.
//
// Layout: the 17 skinning frames (3x4 each) of every hand in the block live in shared memory,
// element-major ([204][kBlock]) so that lanes never conflict; landmarks are staged through the
// same shared memory for fully coalesced 128-bit-friendly stores.
#include "at_common.cuh"

namespace {

constexpr int kBlock = 64;         // hands (= threads) per CTA
constexpr int kFrameFloats = 12;   // 3x4 rigid frame
constexpr int kNumFrames = 17;     // root + wrist + 3 per finger (hand.py:21)
constexpr int kSmemFloats = kNumFrames * kFrameFloats * kBlock;  // 13056 floats = 51 KB

struct Xf {  // rigid 3x4: rotation r (row-major) + translation t
  float r[9];
  float t[3];
};

// so3_exp_map of pytorch3d (hand_skinning.py:46) applied to axis*angle, then
// T = [R, p - R p] (hand_skinning.py:47-51).
__device__ __forceinline__ void joint_local(const float* __restrict__ axis, const float* __restrict__ rest,
                                            float angle, Xf& T) {
  const float wx = axis[0] * angle, wy = axis[1] * angle, wz = axis[2] * angle;
  const float n2 = wx * wx + wy * wy + wz * wz;
  const float th = sqrtf(fmaxf(n2, 1e-4f));
  const float inv = 1.0f / th;
  float s, c;
  sincosf(th, &s, &c);
  const float f1 = inv * s;
  const float f2 = inv * inv * (1.0f - c);
  // K = skew(w); K^2 = w w^T - |w|^2 I
  T.r[0] = 1.0f + f2 * (-(wy * wy + wz * wz));
  T.r[1] = f1 * (-wz) + f2 * (wx * wy);
  T.r[2] = f1 * (wy) + f2 * (wx * wz);
  T.r[3] = f1 * (wz) + f2 * (wx * wy);
  T.r[4] = 1.0f + f2 * (-(wx * wx + wz * wz));
  T.r[5] = f1 * (-wx) + f2 * (wy * wz);
  T.r[6] = f1 * (-wy) + f2 * (wx * wz);
  T.r[7] = f1 * (wx) + f2 * (wy * wz);
  T.r[8] = 1.0f + f2 * (-(wx * wx + wy * wy));
  const float px = rest[0], py = rest[1], pz = rest[2];
  T.t[0] = px - (T.r[0] * px + T.r[1] * py + T.r[2] * pz);
  T.t[1] = py - (T.r[3] * px + T.r[4] * py + T.r[5] * pz);
  T.t[2] = pz - (T.r[6] * px + T.r[7] * py + T.r[8] * pz);
}

__device__ __forceinline__ void compose(const Xf& a, const Xf& b, Xf& o) {  // o = a * b
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j)
      o.r[i * 3 + j] = a.r[i * 3 + 0] * b.r[0 * 3 + j] + a.r[i * 3 + 1] * b.r[1 * 3 + j] + a.r[i * 3 + 2] * b.r[2 * 3 + j];
    o.t[i] = a.r[i * 3 + 0] * b.t[0] + a.r[i * 3 + 1] * b.t[1] + a.r[i * 3 + 2] * b.t[2] + a.t[i];
  }
}

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
                                                         int64_t n, float* __restrict__ landmarks) {
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
      W.t[i] = wp[i * 4 + 3];
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

}  // namespace

extern "C" int at_fk_skin_f32(const float* joint_angles, const float* wrist_xforms, const int64_t* hand_idx,
                              const AtHandModel* model, int64_t n, float* landmarks, void* stream) {
  AT_REQUIRE(model != nullptr, "at_fk_skin_f32: model is NULL");
  AT_REQUIRE(n >= 0, "at_fk_skin_f32: n < 0");
  if (n == 0) return AT_OK;
  AT_REQUIRE(joint_angles && wrist_xforms && landmarks, "at_fk_skin_f32: NULL data pointer");
  AT_REQUIRE(model->joint_rotation_axes && model->joint_rest_positions && model->landmark_rest_positions &&
                 model->landmark_rest_bone_weights && model->landmark_rest_bone_indices,
             "at_fk_skin_f32: NULL hand-model pointer");
  AT_REQUIRE(model->n_weights >= 1 && model->n_weights <= 8, "at_fk_skin_f32: n_weights out of range");
  static bool attr_set = false;
  const int smem_bytes = kSmemFloats * (int)sizeof(float);
  if (!attr_set) {
    cudaFuncSetAttribute(fk_skin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    attr_set = true;
  }
  const int64_t blocks = (n + kBlock - 1) / kBlock;
  AT_REQUIRE(blocks <= 0x7fffffffLL, "at_fk_skin_f32: n too large");
  fk_skin_kernel<<<(unsigned)blocks, kBlock, smem_bytes, (cudaStream_t)stream>>>(
      joint_angles, wrist_xforms, hand_idx, *model, n, landmarks);
  return at::check_launch("fk_skin_kernel");
}
