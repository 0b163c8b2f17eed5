// Crop-camera generation + model-input assembly on the device.
//
// Replaces HandTracker.gen_crop_cameras (lib/tracker/tracker.py:222-260) and the per-crop half of
// HandTracker._make_inputs (tracker.py:315-368); see include/absolutetrack_b200.h section (5) for
// the reference functions folded in.  This is the 14 ms/frame of host Python (4 FK calls, 4
// fisheye projections, look-at, intrinsics per hand) that dominates run_eval_* once the network is
// on the GPU.
//
// Three launches, all tiny and latency-bound (a pose is ~25 kFLOP, 1.7 KB of output):
//   1. crop_cameras_pose_kernel   one WARP per input pose: float32 FK of the 1-3 pose sets (lane =
//      (set, finger) chain), skinning of the 21-63 crop points (lane = landmark), float64 Fisheye62
//      projection of the 21 landmarks into every camera (lane = (camera, landmark)), camera
//      selection, look-at + intrinsics per selected camera (lanes = crop points for the bounds).
//      Results go to per-pose slots in the workspace.
//   2. crop_cameras_scan_kernel   one CTA: exclusive scan of the per-pose view counts -> output
//      offsets and the (n, c) counts the host needs to size the next launches.
//   3. crop_cameras_emit_kernel   one warp per pose: copy the slots to their compacted position
//      (the reference's order), sample_range / hand_idx / memory_idx / use_memory, history update.
#include <mutex>

#include "at_common.cuh"
#include "fk_math.cuh"

namespace {

using abt::Xf;
using abt::compose;
using abt::joint_local;

constexpr int kWarpsPerCta = 4;
constexpr int kMaxCams = 8;
constexpr int kMaxViews = 4;
constexpr int kNumFrames = 17;

struct __align__(16) CropSlot {  // one selected view of one pose
  AtCropParams params;           // 256
  double c2w[16];                // 128
  float extrinsics[16];          // 64
  float intrinsics[9];           // 36
  int32_t cam;                   // 4
  int32_t pad[2];                // 8  -> 496 bytes
};
static_assert(sizeof(CropSlot) == 496, "CropSlot layout");

struct PoseHeader {  // per pose: views selected (0 = dropped), error flag, output offsets
  int32_t nsel, err, crop_off, sample_off;
};

__host__ __device__ inline int64_t header_bytes(int n_poses) { return abt::round_up((int64_t)n_poses * sizeof(PoseHeader), 256); }

// camera.py:296-300 world_to_eye (R^T (p - t)), :80-85 arctan projection, :122-143 Fisheye62
// distortion, :308-312 window coordinates - float64, the reference's operation order.
__device__ __forceinline__ bool landmark_visible(const double* __restrict__ c2w, const AtRigCamera& cam, float px,
                                                 float py, float pz) {
  const double vx = (double)px - c2w[3], vy = (double)py - c2w[7], vz = (double)pz - c2w[11];
  const double ex = vx * c2w[0] + vy * c2w[4] + vz * c2w[8];
  const double ey = vx * c2w[1] + vy * c2w[5] + vz * c2w[9];
  const double ez = vx * c2w[2] + vy * c2w[6] + vz * c2w[10];
  const double r = sqrt(ex * ex + ey * ey);
  const double eps = 2.938735877055719e-39;  // 2^-128
  const double s = atan2(r, ez) / (r > eps ? r : eps);
  const double qx = ex * s, qy = ey * s;
  const double pi2 = 9.869604401089358;
  double r2 = qx * qx + qy * qy;
  r2 = r2 > pi2 ? pi2 : r2;
  const double r4 = r2 * r2, r6 = r2 * r4, r8 = r4 * r4, r10 = r4 * r6, r12 = r6 * r6;
  const double radial =
      1.0 + cam.k[0] * r2 + cam.k[1] * r4 + cam.k[2] * r6 + cam.k[3] * r8 + cam.k[6] * r10 + cam.k[7] * r12;
  const double ux = qx * radial, uy = qy * radial;
  const double x2 = ux * ux, y2 = uy * uy, xy = ux * uy, rr = x2 + y2;
  const double p1 = cam.k[4], p2 = cam.k[5];
  const double dx = ux + (2.0 * p2 * xy + p1 * (rr + 2.0 * x2));
  const double dy = uy + (2.0 * p1 * xy + p2 * (rr + 2.0 * y2));
  const double wx = dx * cam.f[0] + cam.c[0];
  const double wy = dy * cam.f[1] + cam.c[1];
  return wx >= 0.0 && wx <= (double)(cam.width - 1) && wy >= 0.0 && wy <= (double)(cam.height - 1) && ez > 0.0;
}

__device__ __forceinline__ void mat3_mul(const double* a, const double* b, double* o) {  // o = a b (row-major 3x3)
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) o[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j];
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
crop_cameras_pose_kernel(AtCropCameraInputs in, AtHandModel hm, AtCropCameraOpts opt, PoseHeader* __restrict__ hdr,
                         CropSlot* __restrict__ slots) {
  __shared__ float s_frames[kWarpsPerCta][3][kNumFrames][12];
  __shared__ float s_pts[kWarpsPerCta][63][3];
  __shared__ int s_count[kWarpsPerCta][kMaxCams];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x * kWarpsPerCta + warp;
  if (p >= in.n_poses) return;  // whole warps leave together; only __syncwarp below
  const int frame = in.pose_frame[p];
  const int hand = in.pose_hand[p];
  const bool valid = in.pose_valid == nullptr || in.pose_valid[p] != 0;
  const int nsets = opt.num_crop_points / 21;
  float(*frames)[kNumFrames][12] = s_frames[warp];
  float(*pts)[3] = s_pts[warp];

  // ---- 1. FK chains: lane = set*5 + finger (hand_skinning.py:17-32,100-127)
  if (lane < kMaxCams) s_count[warp][lane] = 0;
  if (lane < nsets * 5) {
    const int set = lane / 5, f = lane - set * 5;
    const float* wp = in.wrist_xforms + (int64_t)p * 16;
    const float sgn = hand == 1 ? -1.0f : 1.0f;  // perspective_crop.py:46-48: xf[:, 0] *= -1
    Xf W;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      W.r[i * 3 + 0] = sgn * wp[i * 4 + 0];
      W.r[i * 3 + 1] = wp[i * 4 + 1];
      W.r[i * 3 + 2] = wp[i * 4 + 2];
      W.t[i] = wp[i * 4 + 3];
    }
    if (f == 0) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) frames[set][0][i * 4 + j] = frames[set][1][i * 4 + j] = W.r[i * 3 + j];
        frames[set][0][i * 4 + 3] = frames[set][1][i * 4 + 3] = W.t[i];
      }
    }
    Xf cur = W;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int j = 4 * f + i;
      // set 0: the given pose; set 1: neutral pose; set 2: open hand (perspective_crop.py:113-131)
      const float a = set == 0 ? in.joint_angles[(int64_t)p * 22 + j] : (set == 1 ? in.neutral_joint_angles[j] : 0.0f);
      Xf T, nxt;
      joint_local(hm.joint_rotation_axes + 3 * j, hm.joint_rest_positions + 3 * j, a, T);
      compose(cur, T, nxt);
      cur = nxt;
      if (i >= 1) {
        float* F = frames[set][2 + 3 * f + (i - 1)];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          F[r * 4 + 0] = cur.r[r * 3 + 0]; F[r * 4 + 1] = cur.r[r * 3 + 1]; F[r * 4 + 2] = cur.r[r * 3 + 2];
          F[r * 4 + 3] = cur.t[r];
        }
      }
    }
  }
  __syncwarp();

  // ---- 2. skinning of the crop points: lane -> (set, landmark) (hand_skinning.py:56-97)
  const int K = hm.n_weights;
  for (int idx = lane; idx < nsets * 21; idx += 32) {
    const int set = idx / 21, l = idx - set * 21;
    const float px = hm.landmark_rest_positions[3 * l], py = hm.landmark_rest_positions[3 * l + 1],
                pz = hm.landmark_rest_positions[3 * l + 2];
    float ax = 0.f, ay = 0.f, az = 0.f;
    for (int k = 0; k < K; ++k) {
      const float w = hm.landmark_rest_bone_weights[l * K + k];
      if (w != 0.0f) {
        int b = hm.landmark_rest_bone_indices[l * K + k];
        b = min(max(b, 0), kNumFrames - 1);
        const float* F = frames[set][b];
        const float qx = px * w, qy = py * w, qz = pz * w;
        ax += F[0] * qx + F[1] * qy + F[2] * qz + F[3] * w;
        ay += F[4] * qx + F[5] * qy + F[6] * qz + F[7] * w;
        az += F[8] * qx + F[9] * qy + F[10] * qz + F[11] * w;
      }
    }
    pts[idx][0] = ax; pts[idx][1] = ay; pts[idx][2] = az;
  }
  __syncwarp();

  // ---- 3. visibility of the 21 landmarks of the given pose in every camera (perspective_crop.py:54-86)
  const double* c2w_frame = in.camera_to_world + (int64_t)frame * in.n_cams * 16;
  for (int idx = lane; idx < in.n_cams * 21; idx += 32) {
    const int cam = idx / 21, l = idx - cam * 21;
    if (landmark_visible(c2w_frame + cam * 16, in.cameras[cam], pts[l][0], pts[l][1], pts[l][2]))
      atomicAdd(&s_count[warp][cam], 1);
  }
  __syncwarp();

  // ---- 4. selection (every lane computes the same list)
  int sel[kMaxViews];
  int nsel = 0;
  {
    int keep[kMaxCams], nkeep = 0;
    for (int cam = 0; cam < in.n_cams; ++cam)
      if (s_count[warp][cam] >= opt.min_required_vis_landmarks) keep[nkeep++] = cam;
    if (!opt.sort_camera_index) {  // most visible first, stable (list.sort(reverse=True, key=count))
      for (int i = 1; i < nkeep; ++i) {
        const int c = keep[i];
        int j = i - 1;
        while (j >= 0 && s_count[warp][keep[j]] < s_count[warp][c]) { keep[j + 1] = keep[j]; --j; }
        keep[j + 1] = c;
      }
    }
    nsel = min(nkeep, opt.max_view_num);
    for (int i = 0; i < kMaxViews; ++i) sel[i] = i < nsel ? keep[i] : 0;
  }
  if (!valid) nsel = 0;  // tracker.py:236-237

  // ---- 5. bounding-box centre of the crop points, float32 like the numpy arrays (crop.py:57)
  const int npts = nsets * 21;
  float mn[3], mx[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float lo = INFINITY, hi = -INFINITY;
    for (int i = lane; i < npts; i += 32) { lo = fminf(lo, pts[i][a]); hi = fmaxf(hi, pts[i][a]); }
    mn[a] = warp_min(lo); mx[a] = warp_max(hi);
  }
  const double ctr[3] = {(double)((mn[0] + mx[0]) / 2.0f), (double)((mn[1] + mx[1]) / 2.0f), (double)((mn[2] + mx[2]) / 2.0f)};

  int err = 0;
  for (int s = 0; s < nsel; ++s) {
    const int cam = sel[s];
    const double* cw = c2w_frame + cam * 16;
    double R[9], t[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      R[i * 3] = cw[i * 4]; R[i * 3 + 1] = cw[i * 4 + 1]; R[i * 3 + 2] = cw[i * 4 + 2]; t[i] = cw[i * 4 + 3];
    }
    // affine.py:59-60: centre in the original eye frame, unit z direction
    const double vx = ctr[0] - t[0], vy = ctr[1] - t[1], vz = ctr[2] - t[2];
    double lx = R[0] * vx + R[3] * vy + R[6] * vz, ly = R[1] * vx + R[4] * vy + R[7] * vz, lz = R[2] * vx + R[5] * vy + R[8] * vz;
    const double ln = sqrt(lx * lx + ly * ly + lz * lz);
    lx /= ln; ly /= ln; lz /= ln;
    // affine.py:34-44 from_two_vectors((0,0,1), z_dir)
    const double bn = fmax(5.43e-20, sqrt(lx * lx + ly * ly + lz * lz));
    const double bx = lx / bn, by = ly / bn, bz = lz / bn;
    const double ax_[3] = {-by, bx, 0.0};  // cross((0,0,1), b)
    const double s2 = ax_[0] * ax_[0] + ax_[1] * ax_[1];
    const double g = (1.0 - bz) / fmax(s2, 1e-15);
    // K = skew(v), v = (v0, v1, 0):  K = [[0,0,v1],[0,0,-v0],[-v1,v0,0]];  K^2 = [[-v1^2, v0 v1, 0],[v0 v1, -v0^2, 0],[0,0,-(v0^2+v1^2)]]
    const double v0 = ax_[0], v1 = ax_[1];
    const double dR[9] = {1.0 - g * v1 * v1, g * v0 * v1,       v1,
                          g * v0 * v1,       1.0 - g * v0 * v0, -v0,
                          -v1,               v0,                1.0 - g * s2};
    double Rn[9], Rz[9], Rd[9];
    mat3_mul(R, dR, Rd);
    double sn, cs;
    sincos(in.cameras[cam].angle_deg * 0.017453292519943295, &sn, &cs);
    Rz[0] = cs; Rz[1] = -sn; Rz[2] = 0; Rz[3] = sn; Rz[4] = cs; Rz[5] = 0; Rz[6] = 0; Rz[7] = 0; Rz[8] = 1;
    mat3_mul(Rd, Rz, Rn);  // new eye->world rotation (affine.py:66-74)
    // new world->eye = [Rn^T, -Rn^T t]; right hands: first row negated (crop.py:63-66)
    const double mir = (opt.mirror_right_hand && hand == 1) ? -1.0 : 1.0;
    double Wr[9], Wt[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double m = i == 0 ? mir : 1.0;
      Wr[i * 3] = m * Rn[i]; Wr[i * 3 + 1] = m * Rn[3 + i]; Wr[i * 3 + 2] = m * Rn[6 + i];
      Wt[i] = -(Wr[i * 3] * t[0] + Wr[i * 3 + 1] * t[1] + Wr[i * 3 + 2] * t[2]);
    }
    // crop.py:15-28: the largest focal that keeps every crop point inside the centred image
    double big = 0.0;
    int bad = 0;
    for (int i = lane; i < npts; i += 32) {
      const double x = pts[i][0], y = pts[i][1], z = pts[i][2];
      const double ex = Wr[0] * x + Wr[1] * y + Wr[2] * z + Wt[0];
      const double ey = Wr[3] * x + Wr[4] * y + Wr[5] * z + Wt[1];
      const double ez = Wr[6] * x + Wr[7] * y + Wr[8] * z + Wt[2];
      big = fmax(big, fmax(fabs(ex / ez), fabs(ey / ez)));
      bad |= ez < 0.0001;
    }
    big = warp_max(big);
    bad = __any_sync(0xffffffffu, bad);
    const double cx = ((double)opt.crop_w - 1.0) / 2.0, cy = ((double)opt.crop_h - 1.0) / 2.0;
    const double fx0 = cx / big, fy0 = cy / big;
    if (bad || fx0 < opt.min_focal || fy0 < opt.min_focal) err = 1;
    const double fx = opt.focal_multiplier * fx0, fy = opt.focal_multiplier * fy0;

    if (lane == 0) {
      CropSlot& o = slots[(int64_t)p * opt.max_view_num + s];
      const AtRigCamera& rc = in.cameras[cam];
      o.cam = cam;
      // camera_to_world of the crop camera = inv(new world->eye): rotation Wr^T, same position
      double Rdst[9];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Rdst[i * 3 + j] = Wr[j * 3 + i];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        o.c2w[i * 4] = Rdst[i * 3]; o.c2w[i * 4 + 1] = Rdst[i * 3 + 1]; o.c2w[i * 4 + 2] = Rdst[i * 3 + 2]; o.c2w[i * 4 + 3] = t[i];
        o.extrinsics[i * 4] = (float)Wr[i * 3]; o.extrinsics[i * 4 + 1] = (float)Wr[i * 3 + 1];
        o.extrinsics[i * 4 + 2] = (float)Wr[i * 3 + 2]; o.extrinsics[i * 4 + 3] = (float)(Wt[i] * 0.001);
      }
      o.c2w[12] = 0; o.c2w[13] = 0; o.c2w[14] = 0; o.c2w[15] = 1;
      o.extrinsics[12] = 0; o.extrinsics[13] = 0; o.extrinsics[14] = 0; o.extrinsics[15] = 1;
      o.intrinsics[0] = (float)fx; o.intrinsics[1] = 0; o.intrinsics[2] = (float)cx;
      o.intrinsics[3] = 0; o.intrinsics[4] = (float)fy; o.intrinsics[5] = (float)cy;
      o.intrinsics[6] = 0; o.intrinsics[7] = 0; o.intrinsics[8] = 1;
      AtCropParams& q = o.params;
      q.frame = frame; q.view = cam; q.src_w = rc.width; q.src_h = rc.height;
      // camera.crop_kernel_params: M = R_src^T R_dst, d = R_src^T (t_dst - t_src) = 0 (same position)
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) q.M[i * 3 + j] = R[i] * Rdst[j] + R[3 + i] * Rdst[3 + j] + R[6 + i] * Rdst[6 + j];
      q.d[0] = q.d[1] = q.d[2] = 0.0;
      q.f_dst[0] = fx; q.f_dst[1] = fy; q.c_dst[0] = cx; q.c_dst[1] = cy;
      q.f_src[0] = rc.f[0]; q.f_src[1] = rc.f[1]; q.c_src[0] = rc.c[0]; q.c_src[1] = rc.c[1];
#pragma unroll
      for (int i = 0; i < 8; ++i) q.k[i] = rc.k[i];
      q.reserved[0] = q.reserved[1] = 0.0;
    }
  }
  if (lane == 0) {
    // gen_crop_cameras drops hands with fewer than min_num_crops views (tracker.py:254-260)
    PoseHeader h;
    h.nsel = (nsel >= max(opt.min_num_crops, 1)) ? nsel : 0;
    h.err = err;
    h.crop_off = 0;
    h.sample_off = 0;
    hdr[p] = h;
  }
}

// Exclusive scan of (views, emitted) over the poses, one CTA of 1024 threads, chunked.
__global__ void __launch_bounds__(1024) crop_cameras_scan_kernel(PoseHeader* __restrict__ hdr, int n_poses,
                                                                int32_t* __restrict__ counts) {
  __shared__ int s_c[32], s_n[32], s_e[32];
  __shared__ int base_c, base_n, base_e;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { base_c = 0; base_n = 0; base_e = 0; }
  __syncthreads();
  for (int start = 0; start < n_poses; start += 1024) {
    const int p = start + tid;
    const int c = p < n_poses ? hdr[p].nsel : 0;
    const int n = c > 0 ? 1 : 0;
    const int e = p < n_poses ? hdr[p].err : 0;
    int ic = c, in_ = n, ie = e;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int tc = __shfl_up_sync(0xffffffffu, ic, o), tn = __shfl_up_sync(0xffffffffu, in_, o),
                te = __shfl_up_sync(0xffffffffu, ie, o);
      if (lane >= o) { ic += tc; in_ += tn; ie += te; }
    }
    if (lane == 31) { s_c[warp] = ic; s_n[warp] = in_; s_e[warp] = ie; }
    __syncthreads();
    if (warp == 0) {
      int wc = s_c[lane], wn = s_n[lane], we = s_e[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int tc = __shfl_up_sync(0xffffffffu, wc, o), tn = __shfl_up_sync(0xffffffffu, wn, o),
                  te = __shfl_up_sync(0xffffffffu, we, o);
        if (lane >= o) { wc += tc; wn += tn; we += te; }
      }
      s_c[lane] = wc; s_n[lane] = wn; s_e[lane] = we;
    }
    __syncthreads();
    const int pre_c = base_c + (warp ? s_c[warp - 1] : 0) + ic - c;
    const int pre_n = base_n + (warp ? s_n[warp - 1] : 0) + in_ - n;
    if (p < n_poses) { hdr[p].crop_off = pre_c; hdr[p].sample_off = pre_n; }
    __syncthreads();
    if (tid == 0) { base_c += s_c[31]; base_n += s_n[31]; base_e += s_e[31]; }
    __syncthreads();
  }
  if (tid == 0) { counts[0] = base_n; counts[1] = base_c; counts[2] = base_e; counts[3] = 0; }
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
crop_cameras_emit_kernel(AtCropCameraInputs in, AtCropCameraOpts opt, AtCropCameraOutputs out,
                         const PoseHeader* __restrict__ hdr, const CropSlot* __restrict__ slots) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x * kWarpsPerCta + warp;
  if (p >= in.n_poses) return;
  const PoseHeader h = hdr[p];
  const int slot = in.pose_frame[p] * 2 + in.pose_hand[p];
  if (lane == 0 && out.tracking_history != nullptr) {
    const uint8_t before = out.tracking_history[slot];
    if (h.nsel > 0) out.use_memory[h.sample_off] = before;
    out.tracking_history[slot] = h.nsel > 0;  // tracker.py:409-410
  } else if (lane == 0 && h.nsel > 0) {
    out.use_memory[h.sample_off] = 0;
  }
  if (h.nsel == 0) return;
  if (lane == 0) {
    out.sample_range[2 * h.sample_off] = h.crop_off;
    out.sample_range[2 * h.sample_off + 1] = h.crop_off + h.nsel;
    out.hand_idx[h.sample_off] = in.pose_hand[p];
    out.memory_idx[h.sample_off] = slot;
    if (out.pose_index != nullptr) out.pose_index[h.sample_off] = p;
  }
  for (int s = 0; s < h.nsel; ++s) {
    const CropSlot& src = slots[(int64_t)p * opt.max_view_num + s];
    const int c = h.crop_off + s;
    // 256-byte parameter record: 16 lanes x 16 B
    if (lane < 16) reinterpret_cast<uint4*>(out.crop_params + c)[lane] = reinterpret_cast<const uint4*>(&src.params)[lane];
    if (lane < 16) out.extrinsics[c * 16 + lane] = src.extrinsics[lane];
    if (lane >= 16 && lane < 25) out.intrinsics[c * 9 + (lane - 16)] = src.intrinsics[lane - 16];
    if (out.crop_camera_to_world != nullptr && lane < 16) out.crop_camera_to_world[(int64_t)c * 16 + lane] = src.c2w[lane];
    if (out.crop_camera_index != nullptr && lane == 31) out.crop_camera_index[c] = src.cam;
  }
}

// ------------------------------------------------------------------------------------------
// Config-3 preprocessing: data_transform._gen_crop_matrices (:145-209) + _compute_resample_matrix
// (:57-75) for a batch of pinhole views, one warp per (frame, view).  The views are rigid
// world->eye transforms (the reference inverts them with numpy; here R^T), the crop camera keeps
// the view's position, so the resample matrix has no translation:
//   resample = K_orig [R_orig R_new^T | 0] K_new^-1   (new image -> original image).
__global__ void __launch_bounds__(kWarpsPerCta * 32)
crop_matrices_kernel(const float* __restrict__ orig_extrinsics, const float* __restrict__ orig_intrinsics,
                     const float* __restrict__ crop_points, const uint8_t* __restrict__ mirror, int64_t n_frames,
                     int n_views, int n_pts, int crop_w, int crop_h, float focal_multiplier, float min_focal,
                     float* __restrict__ extrinsics_xf, float* __restrict__ new_intrinsics,
                     float* __restrict__ resample_xf, int* __restrict__ n_errors) {
  const int lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  if (item >= n_frames * n_views) return;
  const int64_t frame = item / n_views;
  const float* E = orig_extrinsics + item * 16;
  const float* pts = crop_points + frame * n_pts * 3;
  double Rw[9], tw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    Rw[i * 3] = E[i * 4]; Rw[i * 3 + 1] = E[i * 4 + 1]; Rw[i * 3 + 2] = E[i * 4 + 2]; tw[i] = E[i * 4 + 3];
  }
  // crop.py:57: centre of the points' bounding box
  float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
  for (int i = lane; i < n_pts; i += 32)
#pragma unroll
    for (int k = 0; k < 3; ++k) { mn[k] = fminf(mn[k], pts[i * 3 + k]); mx[k] = fmaxf(mx[k], pts[i * 3 + k]); }
  double ctr[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) ctr[k] = ((double)warp_min(mn[k]) + (double)warp_max(mx[k])) / 2.0;
  // affine.py:59-60: the centre in the view's eye frame, unit z direction
  double lx = Rw[0] * ctr[0] + Rw[1] * ctr[1] + Rw[2] * ctr[2] + tw[0];
  double ly = Rw[3] * ctr[0] + Rw[4] * ctr[1] + Rw[5] * ctr[2] + tw[1];
  double lz = Rw[6] * ctr[0] + Rw[7] * ctr[1] + Rw[8] * ctr[2] + tw[2];
  const double ln = sqrt(lx * lx + ly * ly + lz * lz);
  lx /= ln; ly /= ln; lz /= ln;
  // affine.py:34-44 from_two_vectors((0,0,1), z_dir); the camera angle is 0 (data_transform.py:199)
  const double bn = fmax(5.43e-20, sqrt(lx * lx + ly * ly + lz * lz));
  const double bx = lx / bn, by = ly / bn, bz = lz / bn;
  const double v0 = -by, v1 = bx;
  const double s2 = v0 * v0 + v1 * v1;
  const double g = (1.0 - bz) / fmax(s2, 1e-15);
  const double dR[9] = {1.0 - g * v1 * v1, g * v0 * v1,       v1,
                        g * v0 * v1,       1.0 - g * v0 * v0, -v0,
                        -v1,               v0,                1.0 - g * s2};
  // new world->eye rotation = dR^T Rw (inverse of R_e2w dR), first row negated when mirrored (crop.py:63-66)
  const double mir = (mirror != nullptr && mirror[frame]) ? -1.0 : 1.0;
  double Wr[9], Wt[3], pos[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) pos[k] = -(Rw[k] * tw[0] + Rw[3 + k] * tw[1] + Rw[6 + k] * tw[2]);  // eye position in the world
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double m = i == 0 ? mir : 1.0;
#pragma unroll
    for (int j = 0; j < 3; ++j) Wr[i * 3 + j] = m * (dR[i] * Rw[j] + dR[3 + i] * Rw[3 + j] + dR[6 + i] * Rw[6 + j]);
    Wt[i] = -(Wr[i * 3] * pos[0] + Wr[i * 3 + 1] * pos[1] + Wr[i * 3 + 2] * pos[2]);
  }
  // crop.py:15-28
  double big = 0.0;
  int bad = 0;
  for (int i = lane; i < n_pts; i += 32) {
    const double x = pts[i * 3], y = pts[i * 3 + 1], z = pts[i * 3 + 2];
    const double ex = Wr[0] * x + Wr[1] * y + Wr[2] * z + Wt[0];
    const double ey = Wr[3] * x + Wr[4] * y + Wr[5] * z + Wt[1];
    const double ez = Wr[6] * x + Wr[7] * y + Wr[8] * z + Wt[2];
    big = fmax(big, fmax(fabs(ex / ez), fabs(ey / ez)));
    bad |= ez < 0.0001;
  }
  big = warp_max(big);
  bad = __any_sync(0xffffffffu, bad);
  const double cx = ((double)crop_w - 1.0) / 2.0, cy = ((double)crop_h - 1.0) / 2.0;
  const double fx0 = cx / big, fy0 = cy / big;
  if (lane == 0 && (bad || fx0 < min_focal || fy0 < min_focal)) atomicAdd(n_errors, 1);
  const double fx = focal_multiplier * fx0, fy = focal_multiplier * fy0;
  if (lane == 0) {
    float* X = extrinsics_xf + item * 16;
    float* K = new_intrinsics + item * 9;
    float* S = resample_xf + item * 16;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      X[i * 4] = (float)Wr[i * 3]; X[i * 4 + 1] = (float)Wr[i * 3 + 1]; X[i * 4 + 2] = (float)Wr[i * 3 + 2]; X[i * 4 + 3] = (float)Wt[i];
    }
    X[12] = 0.f; X[13] = 0.f; X[14] = 0.f; X[15] = 1.f;
    K[0] = (float)fx; K[1] = 0.f; K[2] = (float)cx; K[3] = 0.f; K[4] = (float)fy; K[5] = (float)cy; K[6] = 0.f; K[7] = 0.f; K[8] = 1.f;
    // A = R_orig R_new^T; K_new^-1 = [[1/fx, 0, -cx/fx], [0, 1/fy, -cy/fy], [0, 0, 1]]
    const float* Ko = orig_intrinsics + item * 9;
    const double ofx = Ko[0], ofy = Ko[4], ocx = Ko[2], ocy = Ko[5];  // uv_to_window_matrix of f, c (:181-195)
    double A[9], B[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[i * 3 + j] = Rw[i * 3] * Wr[j * 3] + Rw[i * 3 + 1] * Wr[j * 3 + 1] + Rw[i * 3 + 2] * Wr[j * 3 + 2];
#pragma unroll
    for (int i = 0; i < 3; ++i) {  // B = A K_new^-1
      B[i * 3] = A[i * 3] / fx;
      B[i * 3 + 1] = A[i * 3 + 1] / fy;
      B[i * 3 + 2] = A[i * 3 + 2] - A[i * 3] * (cx / fx) - A[i * 3 + 1] * (cy / fy);
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {  // K_orig B
      S[j] = (float)(ofx * B[j] + ocx * B[6 + j]);
      S[4 + j] = (float)(ofy * B[3 + j] + ocy * B[6 + j]);
      S[8 + j] = (float)B[6 + j];
    }
    S[3] = 0.f; S[7] = 0.f; S[11] = 0.f; S[12] = 0.f; S[13] = 0.f; S[14] = 0.f; S[15] = 1.f;
  }
}

}  // namespace

extern "C" int64_t at_crop_cameras_workspace_bytes(int n_poses, int max_view_num) {
  if (n_poses < 0 || max_view_num < 1 || max_view_num > kMaxViews) return AT_ERR_INVALID_ARGUMENT;
  return header_bytes(n_poses) + (int64_t)n_poses * max_view_num * (int64_t)sizeof(CropSlot);
}

extern "C" int at_gen_crop_cameras(const AtCropCameraInputs* in, const AtHandModel* hand_model,
                                   const AtCropCameraOpts* opts, const AtCropCameraOutputs* out, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
  AT_REQUIRE(in && hand_model && opts && out, "at_gen_crop_cameras: NULL argument struct");
  AT_REQUIRE(in->n_poses >= 0 && in->n_frames >= 0, "at_gen_crop_cameras: negative size");
  AT_REQUIRE(in->n_cams >= 1 && in->n_cams <= kMaxCams, "at_gen_crop_cameras: n_cams must be 1..8");
  AT_REQUIRE(opts->num_crop_points == 21 || opts->num_crop_points == 42 || opts->num_crop_points == 63,
             "at_gen_crop_cameras: num_crop_points must be 21, 42 or 63 (perspective_crop.py:111)");
  AT_REQUIRE(opts->max_view_num >= 1 && opts->max_view_num <= kMaxViews, "at_gen_crop_cameras: max_view_num must be 1..4");
  AT_REQUIRE(opts->crop_w > 1 && opts->crop_h > 1, "at_gen_crop_cameras: crop size");
  AT_REQUIRE(out->counts != nullptr, "at_gen_crop_cameras: counts is NULL");
  AT_REQUIRE(!hand_model->per_sample, "at_gen_crop_cameras: one shared hand model per call");
  AT_REQUIRE(hand_model->n_weights >= 1 && hand_model->n_weights <= 8, "at_gen_crop_cameras: n_weights out of range");
  cudaStream_t st = (cudaStream_t)stream;
  if (in->n_poses == 0) {
    cudaMemsetAsync(out->counts, 0, 4 * sizeof(int32_t), st);
    return AT_OK;
  }
  AT_REQUIRE(in->joint_angles && in->wrist_xforms && in->pose_frame && in->pose_hand && in->camera_to_world && in->cameras,
             "at_gen_crop_cameras: NULL input pointer");
  AT_REQUIRE(opts->num_crop_points == 21 || in->neutral_joint_angles, "at_gen_crop_cameras: neutral_joint_angles is NULL");
  AT_REQUIRE(hand_model->joint_rotation_axes && hand_model->joint_rest_positions && hand_model->landmark_rest_positions &&
                 hand_model->landmark_rest_bone_weights && hand_model->landmark_rest_bone_indices,
             "at_gen_crop_cameras: NULL hand-model pointer");
  AT_REQUIRE(out->crop_params && out->intrinsics && out->extrinsics && out->sample_range && out->hand_idx &&
                 out->memory_idx && out->use_memory,
             "at_gen_crop_cameras: NULL output pointer");
  const int64_t need = at_crop_cameras_workspace_bytes(in->n_poses, opts->max_view_num);
  if (workspace == nullptr || workspace_bytes < need) {
    abt::set_error("at_gen_crop_cameras: workspace of %lld bytes needed, %lld given", (long long)need, (long long)workspace_bytes);
    return AT_ERR_WORKSPACE_TOO_SMALL;
  }
  PoseHeader* hdr = reinterpret_cast<PoseHeader*>(workspace);
  CropSlot* slots = reinterpret_cast<CropSlot*>(reinterpret_cast<char*>(workspace) + header_bytes(in->n_poses));
  const unsigned blocks = (unsigned)((in->n_poses + kWarpsPerCta - 1) / kWarpsPerCta);
  // these launches run on the ingest stream next to the previous step's encoder: same carve-out
  // as the conv kernels (at_common.cuh)
  AT_CARVEOUT_ONCE(crop_cameras_pose_kernel);
  AT_CARVEOUT_ONCE(crop_cameras_scan_kernel);
  AT_CARVEOUT_ONCE(crop_cameras_emit_kernel);
  crop_cameras_pose_kernel<<<blocks, kWarpsPerCta * 32, 0, st>>>(*in, *hand_model, *opts, hdr, slots);
  int rc = abt::check_launch("crop_cameras_pose_kernel");
  if (rc) return rc;
  crop_cameras_scan_kernel<<<1, 1024, 0, st>>>(hdr, in->n_poses, out->counts);
  rc = abt::check_launch("crop_cameras_scan_kernel");
  if (rc) return rc;
  crop_cameras_emit_kernel<<<blocks, kWarpsPerCta * 32, 0, st>>>(*in, *opts, *out, hdr, slots);
  return abt::check_launch("crop_cameras_emit_kernel");
}

extern "C" int at_gen_crop_matrices_f32(const float* orig_extrinsics, const float* orig_intrinsics, const float* crop_points,
                                        const uint8_t* mirror_image, int64_t n_frames, int n_views, int n_pts, int crop_w,
                                        int crop_h, float focal_multiplier, float* extrinsics_xf, float* new_intrinsics,
                                        float* resample_xf, int* n_errors, void* stream) {
  AT_REQUIRE(n_frames >= 0 && n_views >= 1 && n_pts >= 1, "at_gen_crop_matrices_f32: sizes");
  AT_REQUIRE(crop_w > 1 && crop_h > 1, "at_gen_crop_matrices_f32: crop size");
  if (n_frames == 0) return AT_OK;
  AT_REQUIRE(orig_extrinsics && orig_intrinsics && crop_points && extrinsics_xf && new_intrinsics && resample_xf && n_errors,
             "at_gen_crop_matrices_f32: NULL pointer");
  const int64_t items = n_frames * n_views;
  AT_REQUIRE(items <= 0x7fffffffLL, "at_gen_crop_matrices_f32: too many views");
  const unsigned blocks = (unsigned)((items + kWarpsPerCta - 1) / kWarpsPerCta);
  crop_matrices_kernel<<<blocks, kWarpsPerCta * 32, 0, (cudaStream_t)stream>>>(
      orig_extrinsics, orig_intrinsics, crop_points, mirror_image, n_frames, n_views, n_pts, crop_w, crop_h,
      focal_multiplier, 5.0f, extrinsics_xf, new_intrinsics, resample_xf, n_errors);
  return abt::check_launch("crop_matrices_kernel");
}

// ------------------------------------------------------------------------------------------
// at_gen_crop_cameras_static: the same three launches with the between-pass scratch (pose headers + per-view
// slots) in a file-scope device array instead of a caller-provided workspace.  No counterpart in the
// reference; added for callers that cannot allocate device memory
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the
// library holds the array (about 4 MB) until the process exits.
namespace {

constexpr int kStaticMaxPoses = AT_CROP_CAMERAS_STATIC_MAX_POSES;
constexpr int64_t kStaticScratchBytes =
    (kStaticMaxPoses * (int64_t)sizeof(PoseHeader) + 255) / 256 * 256 + (int64_t)kStaticMaxPoses * kMaxViews * (int64_t)sizeof(CropSlot);

static __device__ __align__(256) unsigned char g_crop_cameras_scratch[kStaticScratchBytes];

// In host code `g_crop_cameras_scratch` is only a stand-in for the symbol, not a device address: the address
// is looked up once per device, with that device current, and cached.
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
void* g_scratch_ptr[kMaxDevices] = {};

int static_scratch_address(void** ptr) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    abt::set_error("at_gen_crop_cameras_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return AT_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (g_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_crop_cameras_scratch);
    if (e != cudaSuccess || p == nullptr) {
      abt::set_error("at_gen_crop_cameras_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return AT_ERR_CUDA;
    }
    g_scratch_ptr[dev] = p;
  }
  *ptr = g_scratch_ptr[dev];
  return AT_OK;
}

}  // namespace

extern "C" int at_gen_crop_cameras_static(const AtCropCameraInputs* in, const AtHandModel* hand_model,
                                          const AtCropCameraOpts* opts, const AtCropCameraOutputs* out, void* stream) {
  AT_REQUIRE(in && hand_model && opts && out, "at_gen_crop_cameras_static: NULL argument struct");
  AT_REQUIRE(opts->max_view_num >= 1 && opts->max_view_num <= kMaxViews, "at_gen_crop_cameras_static: max_view_num must be 1..4");
  if (in->n_poses > kStaticMaxPoses) {
    abt::set_error("at_gen_crop_cameras_static: %d poses exceed the static scratch of AT_CROP_CAMERAS_STATIC_MAX_POSES = %d "
                   "(use at_gen_crop_cameras with a caller-provided workspace)", in->n_poses, kStaticMaxPoses);
    return AT_ERR_WORKSPACE_TOO_SMALL;
  }
  void* scratch = nullptr;
  const int rc = static_scratch_address(&scratch);
  if (rc) return rc;
  // Nothing in the scratch has to start at zero: pass 1 writes the header of every pose and the slots of
  // every selected view before pass 2 / 3 read them, so the previous call's contents are never observed.
  return at_gen_crop_cameras(in, hand_model, opts, out, scratch, kStaticScratchBytes, stream);
}
