// Rigid-frame helpers shared by the FK/skinning kernels (fk_skin.cu) and the crop-camera
// generator (crop_cameras.cu): float32, the reference's operation order
// (lib/common/hand_skinning.py:35-53 with pytorch3d so3_exp_map, :17-32).
#pragma once
#include "at_common.cuh"

namespace abt {

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

}  // namespace abt
