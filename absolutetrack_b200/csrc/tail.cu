// Per-hand "tail" of the network at 6x6 resolution: FTL into the canonical space, 2-view fusion,
// SimpleConvRNN step, skeleton features, pose head, decoders (Procrustes/SVD) and wrist->world.
//
// Replaces feature_extractor.py:61-141, model_utils.py:57-104,166-192 (FTL), temporal.py:51-139,
// skeleton_encoder.py:43-53, regressor.py:76-186, model_utils.py:17-54 (procrustes_align),
// umetrack_model.py:77-97,123-242.  Features are NHWC: [sample, 36 pixels, channels].
// The small 4x4 / 3x3 matrix algebra runs in double on one thread per hand (the reference uses
// fp32 torch.inverse / svd; the differences are ~1e-7 relative); feature math is fp32.
#include <stdlib.h>

#include "conv.cuh"

namespace abt {
namespace {

constexpr int kPix = 36;
constexpr int kFeat = 72;
constexpr int kMem = 18;

// ---------------------------------------------------------------- small dense algebra (double)
__device__ void mat4_mul(const double* a, const double* b, double* o) {
  double r[16];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += a[i * 4 + k] * b[k * 4 + j];
      r[i * 4 + j] = s;
    }
  for (int i = 0; i < 16; ++i) o[i] = r[i];
}

// general 4x4 inverse, Gauss-Jordan with partial pivoting (torch.inverse equivalent)
__device__ bool mat4_inv(const double* a, double* o) {
  double m[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      m[i][j] = a[i * 4 + j];
      m[i][4 + j] = (i == j) ? 1.0 : 0.0;
    }
  bool ok = true;
  for (int c = 0; c < 4; ++c) {
    int piv = c;
    for (int r = c + 1; r < 4; ++r)
      if (fabs(m[r][c]) > fabs(m[piv][c])) piv = r;
    if (m[piv][c] == 0.0) {
      ok = false;
      continue;
    }
    if (piv != c)
      for (int j = 0; j < 8; ++j) {
        const double t = m[c][j];
        m[c][j] = m[piv][j];
        m[piv][j] = t;
      }
    const double inv = 1.0 / m[c][c];
    for (int j = 0; j < 8; ++j) m[c][j] *= inv;
    for (int r = 0; r < 4; ++r)
      if (r != c) {
        const double f = m[r][c];
        if (f != 0.0)
          for (int j = 0; j < 8; ++j) m[r][j] -= f * m[c][j];
      }
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) o[i * 4 + j] = m[i][4 + j];
  return ok;
}

__device__ double det3(const double* m) {
  return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}

// Kabsch rotation R = V diag(1,1,det(V U^T)) U^T for H = U S V^T   (model_utils.py:41-50),
// via one-sided Jacobi (Hestenes) on H: H V = U S.
__device__ void kabsch_rotation(const double* H, double* R) {
  double A[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  for (int i = 0; i < 9; ++i) A[i] = H[i];
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int i = 0; i < 3; ++i) {
          alpha += A[i * 3 + p] * A[i * 3 + p];
          beta += A[i * 3 + q] * A[i * 3 + q];
          gamma += A[i * 3 + p] * A[i * 3 + q];
        }
        if (gamma == 0.0) continue;
        off = fmax(off, fabs(gamma) / sqrt(fmax(alpha * beta, 1e-300)));
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int i = 0; i < 3; ++i) {
          const double ap = A[i * 3 + p], aq = A[i * 3 + q];
          A[i * 3 + p] = c * ap - s * aq;
          A[i * 3 + q] = s * ap + c * aq;
          const double vp = V[i * 3 + p], vq = V[i * 3 + q];
          V[i * 3 + p] = c * vp - s * vq;
          V[i * 3 + q] = s * vp + c * vq;
        }
      }
    if (off < 1e-15) break;
  }
  // singular values = column norms; order columns by decreasing norm
  double sv[3];
  int ord[3] = {0, 1, 2};
  for (int j = 0; j < 3; ++j) sv[j] = sqrt(A[j] * A[j] + A[3 + j] * A[3 + j] + A[6 + j] * A[6 + j]);
  for (int i = 0; i < 2; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (sv[ord[j]] > sv[ord[i]]) {
        const int t = ord[i];
        ord[i] = ord[j];
        ord[j] = t;
      }
  double U[9], Vs[9];
  for (int k = 0; k < 3; ++k) {
    const int j = ord[k];
    for (int i = 0; i < 3; ++i) Vs[i * 3 + k] = V[i * 3 + j];
    const double inv = sv[j] > 0 ? 1.0 / sv[j] : 0.0;
    for (int i = 0; i < 3; ++i) U[i * 3 + k] = A[i * 3 + j] * inv;
  }
  const double s0 = sv[ord[0]];
  if (!(s0 > 0)) {  // H == 0: rotation undefined, return identity
    for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0 : 0.0;
    return;
  }
  if (sv[ord[1]] <= 1e-14 * s0) {  // rank 1: complete u2 with any unit vector orthogonal to u1
    const double ux = U[0], uy = U[3], uz = U[6];
    double ax = fabs(ux) < 0.9 ? 1.0 : 0.0, ay = fabs(ux) < 0.9 ? 0.0 : 1.0, az = 0.0;
    const double d = ax * ux + ay * uy + az * uz;
    ax -= d * ux; ay -= d * uy; az -= d * uz;
    const double nn = sqrt(ax * ax + ay * ay + az * az);
    U[1] = ax / nn; U[4] = ay / nn; U[7] = az / nn;
  }
  if (sv[ord[2]] <= 1e-14 * s0) {  // rank <= 2: u3 = u1 x u2 (sign fixed by the det correction)
    U[2] = U[3] * U[7] - U[6] * U[4];
    U[5] = U[6] * U[1] - U[0] * U[7];
    U[8] = U[0] * U[4] - U[3] * U[1];
  }
  const double d = (det3(Vs) * det3(U)) >= 0 ? 1.0 : -1.0;  // det(V U^T)
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      R[i * 3 + j] = Vs[i * 3 + 0] * U[j * 3 + 0] + Vs[i * 3 + 1] * U[j * 3 + 1] + d * Vs[i * 3 + 2] * U[j * 3 + 2];
}

__device__ __forceinline__ void load4x4(const float* p, double* o) {
  for (int i = 0; i < 16; ++i) o[i] = (double)p[i];
}

// ---------------------------------------------------------------- kernels

// SkeletonEncoder (skeleton_encoder.py:43-53): cat(axes, rest)[22,6] -> Linear(132->144) -> BN -> ReLU
__global__ void skeleton_kernel(const float* __restrict__ axes, const float* __restrict__ rest,
                                const float* __restrict__ w, const float* __restrict__ b,
                                float* __restrict__ out /*[n_skel,144]*/) {
  pdl_prologue();
  __shared__ float in[132];
  const int s = blockIdx.x;
  for (int k = threadIdx.x; k < 132; k += blockDim.x) {
    const int j = k / 6, e = k % 6;
    in[k] = e < 3 ? axes[(s * 22 + j) * 3 + e] : rest[(s * 22 + j) * 3 + (e - 3)];
  }
  __syncthreads();
  const int o = threadIdx.x;
  if (o < 144) {
    float acc = b[o];
    for (int k = 0; k < 132; ++k) acc = fmaf(in[k], w[k * 144 + o], acc);
    out[s * 144 + o] = fmaxf(acc, 0.f);
  }
}

struct PrepArgs {
  AtTailInputs in;
  AtTemporalState st;
  float* fuse_in;   // [n,36,144]
  float* temp_in;   // [n,36,90]
  int32_t* status;
};

// Per hand sample: view transforms (feature_extractor.py:61-87, model_utils.py:166-192), FTL of the
// per-view features into the canonical space (feature_extractor.py:123-128) laid out as the fusion
// input, single-view shortcut (feature_extractor.py:89-94), and the memory warp / reset of
// temporal.py:51-91 written into channels [0,18) of the temporal input.
__global__ void __launch_bounds__(128, 7) tail_prepare_kernel(PrepArgs a) {   // 1,024 CTAs (one per sample) fit ONE wave of 148 x 7
  pdl_prologue();
  __shared__ float X[2][12];  // per view: 3x3 R (row-major) + t
  __shared__ float Xm[12];
  __shared__ int s_info[4];   // r0, nv, mem slot, use_memory
  const int i = blockIdx.x;
  // The float64 set-up (three 4x4 inverses, six products) is three independent chains - the transform of view 0, of
  // view 1 and of the memory features; lane 0 of warps 0, 1 and 2 take one each (one thread doing all three was most of
  // this kernel's time).  Every chain keeps its operation order, so the matrices are bit-identical to the serial version.
  const int64_t r0 = a.in.sample_range[2 * i], r1 = a.in.sample_range[2 * i + 1];
  const int64_t m = a.in.memory_idx[i];
  int nv = (int)(r1 - r0);
  const bool bad = r0 < 0 || r1 > a.in.c || nv < 1 || nv > 2 || m < 0 || m >= a.st.capacity;
  if (bad) nv = 0;
  const int use_memory = bad ? 0 : (int)a.in.use_memory[i];
  const int role = (threadIdx.x & 31) == 0 ? (int)(threadIdx.x >> 5) : -1;   // 0, 1: views; 2: memory
  if (role == 0) {
    if (bad) atomicOr(a.status, 1);  // AT_STATUS_BAD_SAMPLE_RANGE (the encoder ORs its own bit into the same word)
    s_info[0] = (int)r0; s_info[1] = nv; s_info[2] = (int)m; s_info[3] = use_memory;
  }
  if (!bad && role >= 0 && role <= 2) {
    double E0[16], S0[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    load4x4(a.in.extrinsics + r0 * 16, E0);
    S0[10] = (double)(a.in.intrinsics[r0 * 9] / 200.0f);  // compute_singlev_xfs (model_utils.py:166-192)
    if (role < 2 && nv == 2) {
      const int v = role;
      double S0i[16];
      mat4_inv(S0, S0i);
      double Ev[16], Evi[16], Sv[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, T[16];
      load4x4(a.in.extrinsics + (r0 + v) * 16, Ev);
      Sv[10] = (double)(a.in.intrinsics[(r0 + v) * 9] / 200.0f);
      mat4_inv(Ev, Evi);
      mat4_mul(Evi, Sv, T);  // xf_to_world = inv(E_v) @ S_v
      mat4_mul(E0, T, T);    // xf_0 @ ...
      mat4_mul(S0i, T, T);   // s_0 @ ...
      for (int r = 0; r < 3; ++r) {
        for (int q = 0; q < 3; ++q) X[v][r * 3 + q] = (float)T[r * 4 + q];
        X[v][9 + r] = (float)T[r * 4 + 3];
      }
    } else if (role == 0) {  // single view: FTL by S_v only (feature_extractor.py:89-94)
      for (int r = 0; r < 3; ++r) {
        for (int q = 0; q < 3; ++q) X[0][r * 3 + q] = (float)S0[r * 4 + q];
        X[0][9 + r] = 0.f;
      }
    }
    if (role == 2) {
      if (use_memory) {  // prev_cam0 -> cur_cam0 (temporal.py:69-73)
        double P[16], Pi[16], T[16];
        load4x4(a.st.prev_extrinsics + m * 16, P);
        mat4_inv(P, Pi);
        mat4_mul(E0, Pi, T);
        for (int r = 0; r < 3; ++r) {
          for (int q = 0; q < 3; ++q) Xm[r * 3 + q] = (float)T[r * 4 + q];
          Xm[9 + r] = (float)T[r * 4 + 3];
        }
      }
      for (int k = 0; k < 16; ++k) a.st.prev_extrinsics[m * 16 + k] = a.in.extrinsics[r0 * 16 + k];  // temporal.py:89
    }
  }
  __syncthreads();
  const int use_mem = s_info[3];
  if (nv == 0) return;
  // image features: 24 xyz triplets (channels k, k+24, k+48) per pixel per view
  for (int idx = threadIdx.x; idx < nv * kPix * 24; idx += blockDim.x) {
    const int v = idx / (kPix * 24);
    const int rem = idx - v * kPix * 24;
    const int p = rem / 24, k = rem - p * 24;
    const float* f = a.in.features + ((int64_t)(r0 + v) * kPix + p) * kFeat;
    const float x = f[k], y = f[k + 24], z = f[k + 48];
    const float* R = X[v];
    const float ox = R[0] * x + R[1] * y + R[2] * z + R[9];
    const float oy = R[3] * x + R[4] * y + R[5] * z + R[10];
    const float oz = R[6] * x + R[7] * y + R[8] * z + R[11];
    if (nv == 2) {
      float* o = a.fuse_in + ((int64_t)i * kPix + p) * 144 + v * kFeat;
      o[k] = ox; o[k + 24] = oy; o[k + 48] = oz;
    } else {
      float* o = a.temp_in + ((int64_t)i * kPix + p) * 90 + kMem;
      o[k] = ox; o[k + 24] = oy; o[k + 48] = oz;
    }
  }
  // memory features: 6 triplets (channels k, k+6, k+12) per pixel
  for (int idx = threadIdx.x; idx < kPix * 6; idx += blockDim.x) {
    const int p = idx / 6, k = idx - p * 6;
    float ox = 0.f, oy = 0.f, oz = 0.f;
    if (use_mem) {
      const float* f = a.st.mem_features + ((int64_t)m * kPix + p) * kMem;
      const float x = f[k], y = f[k + 6], z = f[k + 12];
      ox = Xm[0] * x + Xm[1] * y + Xm[2] * z + Xm[9];
      oy = Xm[3] * x + Xm[4] * y + Xm[5] * z + Xm[10];
      oz = Xm[6] * x + Xm[7] * y + Xm[8] * z + Xm[11];
    }
    float* o = a.temp_in + ((int64_t)i * kPix + p) * 90;
    o[k] = ox; o[k + 6] = oy; o[k + 12] = oz;
  }
}

// After the fusion convs: FTL canonical -> cam0 by S_0 = diag(1,1,fx0/200,1)
// (feature_extractor.py:137-139), written into channels [18,90) of the temporal input.
// Single-view samples were already written by tail_prepare_kernel.
__global__ void post_fusion_kernel(const float* __restrict__ fused /*[n,36,72]*/, const float* __restrict__ intrinsics,
                                   const int64_t* __restrict__ sample_range, int n, int c,
                                   float* __restrict__ temp_in /*[n,36,90]*/) {
  pdl_prologue();
  // one CTA per sample (the per-element version spent its time in 64-bit index divisions: 18 us for 10 MB)
  const int i = blockIdx.x;
  if (i >= n) return;
  const int64_t r0 = sample_range[2 * i], r1 = sample_range[2 * i + 1];
  if (r1 - r0 != 2 || r0 < 0 || r1 > c) return;
  const float sz = intrinsics[r0 * 9] / 200.0f;
  const float* src = fused + (int64_t)i * kPix * kFeat;
  float* dst = temp_in + (int64_t)i * kPix * 90 + kMem;
  for (int idx = threadIdx.x; idx < kPix * kFeat; idx += blockDim.x) {
    const int p = idx / kFeat, ch = idx - p * kFeat;
    const float v = src[idx];
    dst[p * 90 + ch] = ch >= 48 ? sz * v : v;
  }
}

// After the temporal convs (temporal.py:84-91,127-137): out[:, :18] -> memory[memory_idx],
// out[:, 18:] -> head input channels [0,72); skeleton features -> channels [72,76).
__global__ void split_kernel(const float* __restrict__ t_out /*[n,36,90]*/, const int64_t* __restrict__ memory_idx,
                             int64_t capacity, float* __restrict__ mem /*[cap,36,18]*/,
                             const float* __restrict__ skel /*[n_skel,144] or null*/, int n_skel, int n, int head_c,
                             float* __restrict__ head_in /*[n,36,head_c]*/) {
  pdl_prologue();
  // one CTA per sample, 32-bit index arithmetic (see post_fusion_kernel)
  const int i = blockIdx.x;
  if (i >= n) return;
  const int64_t m = memory_idx[i];
  const bool m_ok = m >= 0 && m < capacity;
  const float* src = t_out + (int64_t)i * kPix * 90;
  float* hin = head_in + (int64_t)i * kPix * head_c;
  float* mrow = mem + (m_ok ? m : 0) * kPix * kMem;
  for (int idx = threadIdx.x; idx < kPix * 90; idx += blockDim.x) {
    const int p = idx / 90, ch = idx - p * 90;
    const float v = src[idx];
    if (ch < kMem) {
      if (m_ok) mrow[p * kMem + ch] = v;
    } else {
      hin[p * head_c + (ch - kMem)] = v;
    }
  }
  const int extra = head_c - kFeat;  // skeleton channels 0..3, feature index sc*36 + p (View((-1,4,6,6)))
  const float* sk = skel != nullptr ? skel + (int64_t)(n_skel == 1 ? 0 : i) * 144 : nullptr;
  for (int idx = threadIdx.x; idx < kPix * extra; idx += blockDim.x) {
    const int p = idx / extra, sc = idx - p * extra;
    hin[p * head_c + kFeat + sc] = sk[sc * kPix + p];
  }
}

struct DecodeArgs {
  const float* head_out;  // [n,36,ld]
  int ld, n_out;          // 62 (known skeleton) or 63
  const float* extrinsics;
  const int64_t* sample_range;
  const int64_t* hand_idx;
  int n, c, predict_scale;
  AtTailOutputs out;
};

// AdaptiveAvgPool2d + decoders (regressor.py:76-121,163-186) + wrist->world (umetrack_model.py:77-90)
__global__ void __launch_bounds__(64) decode_kernel(DecodeArgs a) {
  pdl_prologue();
  __shared__ float f[64];
  __shared__ double s_E0i[16];   // inv(E_cam0): computed by warp 0 while warp 1 runs the Procrustes fit
  __shared__ int s_have_E0;
  const int i = blockIdx.x;
  const int t = threadIdx.x;
  if (t < a.n_out) {
    float s = 0.f;
    for (int p = 0; p < kPix; ++p) s += a.head_out[((int64_t)i * kPix + p) * a.ld + t];
    f[t] = s / 36.0f;
  }
  __syncthreads();
  if (t == 0) {
    const int64_t r0 = a.sample_range[2 * i];
    s_have_E0 = r0 >= 0 && r0 < a.c;
    if (s_have_E0) {
      double E0[16];
      load4x4(a.extrinsics + r0 * 16, E0);
      mat4_inv(E0, s_E0i);
    }
  }
  if (t < 22) a.out.joint_angles[i * 22 + t] = t < 20 ? f[t] : 0.0f;  // decode_joint_angles
  const int sig0 = 41 + (a.predict_scale ? 1 : 0);
  if (t < 21) {  // decode_landmark_unc_sigmas: clamp(softplus(x), 1e-5); softplus threshold 20
    const float x = f[sig0 + t];
    const float sp = x > 20.0f ? x : log1pf(expf(x));
    a.out.sigmas[i * 21 + t] = fmaxf(sp, 1e-5f);
  }
  double Wc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1};
  if (t == 32) {
    if (a.predict_scale && a.out.skel_scales) a.out.skel_scales[i] = expf(f[41]);  // decode_skel_scales
    // decode_wrist_xfs_svd: Procrustes from the fixed 7-point template (regressor.py:19-47)
    const double q = 0.1 / sqrt(2.0);
    const double A[7][3] = {{0, 0, 0}, {0.1, 0, 0}, {0, 0.1, 0}, {0, 0, 0.1}, {-q, -q, 0}, {-q, 0, -q}, {0, -q, -q}};
    double am[3] = {0, 0, 0}, bm[3] = {0, 0, 0}, B[7][3];
    for (int k = 0; k < 7; ++k)
      for (int e = 0; e < 3; ++e) {
        // the template is float32 in the reference (torch .float())
        B[k][e] = (double)f[20 + 3 * k + e];
        am[e] += (double)(float)A[k][e];
        bm[e] += B[k][e];
      }
    for (int e = 0; e < 3; ++e) {
      am[e] /= 7.0;
      bm[e] /= 7.0;
    }
    double H[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < 7; ++k)
      for (int r = 0; r < 3; ++r)
        for (int cc = 0; cc < 3; ++cc) H[r * 3 + cc] += ((double)(float)A[k][r] - am[r]) * (B[k][cc] - bm[cc]);
    double R[9];
    kabsch_rotation(H, R);
    for (int r = 0; r < 3; ++r) {
      for (int cc = 0; cc < 3; ++cc) Wc[r * 4 + cc] = R[r * 3 + cc];
      Wc[r * 4 + 3] = bm[r] - (R[r * 3] * am[0] + R[r * 3 + 1] * am[1] + R[r * 3 + 2] * am[2]);
    }
  }
  __syncthreads();   // s_E0i
  if (t == 32) {
    // wrist in world = inv(E_cam0) @ wrist_in_cam0; right hand: negate column 0
    double Ww[16];
    if (s_have_E0) {
      mat4_mul(s_E0i, Wc, Ww);
    } else {
      for (int k = 0; k < 16; ++k) Ww[k] = Wc[k];
    }
    const double sgn = a.hand_idx[i] == 1 ? -1.0 : 1.0;
    for (int r = 0; r < 4; ++r)
      for (int cc = 0; cc < 4; ++cc) a.out.wrist_xfs[i * 16 + r * 4 + cc] = (float)(cc == 0 ? sgn * Ww[r * 4 + cc] : Ww[r * 4 + cc]);
  }
}

// workspace carve-up (floats per sample)
constexpr int64_t kWsFuseIn = kPix * 144, kWsF1 = kPix * 108, kWsF2 = kPix * 72, kWsF3 = kPix * 72;
constexpr int64_t kWsT = kPix * 90;   // x4: temp_in, t1, t2, t3
constexpr int64_t kWsH = kPix * 76;   // x3: head_in, h1, h2
constexpr int64_t kWsHo = kPix * 64;
constexpr int64_t kWsSkel = 144;
constexpr int64_t kWsPerSample = kWsFuseIn + kWsF1 + kWsF2 + kWsF3 + 4 * kWsT + 3 * kWsH + kWsHo + kWsSkel;

}  // namespace
}  // namespace abt

extern "C" int64_t at_tail_workspace_bytes(int n) {
  if (n < 0) return AT_ERR_INVALID_ARGUMENT;
  // fp32 NHWC activations of the tail + the split fp16 planes of the tcgen05 pose head
  return (int64_t)sizeof(float) * abt::kWsPerSample * (int64_t)n + 256 + abt::head_tc_workspace_bytes(n);
}

namespace abt {
namespace {
// byte offset of layer `id`'s pre-split fp16 weights inside the tail operand blob (layers L_FUSION.. in order)
int64_t f16_blob_offset(int id) {
  const ConvLayer* T = layer_table();
  int64_t off = 0;
  for (int i = L_FUSION; i < id; ++i)
    if (i != L_SKEL) off += round_up(conv_f16x3_weight_halves(T[i]) * 2, 256);
  return off;
}
}  // namespace
}  // namespace abt

namespace abt {
namespace {
// the tcgen05 operands of the 8 pose-head 3x3 convs (4 per head) follow the mma.sync operands
int64_t head_tc_blob_offset(int id) {
  const int k = id >= L_HEAD_U ? 4 + (id - L_HEAD_U) : id - L_HEAD_K;
  return round_up(f16_blob_offset(L_COUNT), 256) + (int64_t)k * round_up(head_tc_weight_bytes(), 256);
}
}  // namespace
}  // namespace abt

extern "C" int64_t at_tail_f16_blob_bytes(void) { return abt::head_tc_blob_offset(abt::L_HEAD_U + 4); }

extern "C" int at_tail_prepack_f16(const float* blob, void* f16_blob, void* stream) {
  using namespace abt;
  AT_REQUIRE(blob && f16_blob, "at_tail_prepack_f16: NULL argument");
  AT_REQUIRE((((uintptr_t)f16_blob) & 15) == 0, "at_tail_prepack_f16: the operand blob must be 16-byte aligned");
  const ConvLayer* T = layer_table();
  for (int i = L_FUSION; i < L_COUNT; ++i) {
    if (i == L_SKEL) continue;
    int rc = conv_f16x3_split_weights(T[i], blob, (uint8_t*)f16_blob + f16_blob_offset(i), (cudaStream_t)stream);
    if (rc) return rc;
  }
  for (int hb : {(int)L_HEAD_K, (int)L_HEAD_U})
    for (int i = 0; i < 4; ++i) {
      int rc = head_tc_prepack(T[hb + i], blob, (uint8_t*)f16_blob + head_tc_blob_offset(hb + i), (cudaStream_t)stream);
      if (rc) return rc;
    }
  return AT_OK;
}

extern "C" int at_tail_forward(const float* blob, const AtTailInputs* in, AtTemporalState* state,
                               const AtTailOutputs* out, void* workspace, int64_t workspace_bytes, void* stream_) {
  return at_tail_forward_f16(blob, nullptr, in, state, out, workspace, workspace_bytes, stream_);
}

extern "C" int at_tail_forward_f16(const float* blob, const void* f16_blob, const AtTailInputs* in,
                                   AtTemporalState* state, const AtTailOutputs* out, void* workspace,
                                   int64_t workspace_bytes, void* stream_) {
  using namespace abt;
  AT_REQUIRE(blob && in && state && out, "at_tail_forward: NULL argument");
  AT_REQUIRE((((uintptr_t)f16_blob) & 15) == 0, "at_tail_forward: the fp16 operand blob must be 16-byte aligned");
  const int n = in->n, c = in->c;
  AT_REQUIRE(n >= 0 && c >= 0, "at_tail_forward: negative size");
  if (n == 0) return AT_OK;
  AT_REQUIRE(in->features && in->intrinsics && in->extrinsics && in->sample_range && in->memory_idx &&
                 in->use_memory && in->hand_idx,
             "at_tail_forward: NULL input pointer");
  AT_REQUIRE(out->joint_angles && out->wrist_xfs && out->sigmas && out->status, "at_tail_forward: NULL output pointer");
  AT_REQUIRE(state->mem_features && state->prev_extrinsics && state->capacity > 0,
             "at_tail_forward: temporal state not allocated");
  const bool known = in->head_mode == AT_HEAD_KNOWN_SKELETON;
  AT_REQUIRE(known || in->head_mode == AT_HEAD_PREDICT_SCALE, "at_tail_forward: bad head_mode %d", in->head_mode);
  if (known) {
    AT_REQUIRE(in->skel_axes && in->skel_rest, "at_tail_forward: skeleton data required for the known-skeleton head");
    AT_REQUIRE(in->n_skel == 1 || in->n_skel == n, "at_tail_forward: n_skel must be 1 or n");
  } else {
    AT_REQUIRE(out->skel_scales, "at_tail_forward: skel_scales output required for the scale head");
  }
  if (workspace_bytes < at_tail_workspace_bytes(n) || workspace == nullptr) {
    set_error("at_tail_forward: workspace too small (%lld < %lld)", (long long)workspace_bytes,
              (long long)at_tail_workspace_bytes(n));
    return AT_ERR_WORKSPACE_TOO_SMALL;
  }
  cudaStream_t s = (cudaStream_t)stream_;
  const ConvLayer* T = layer_table();
  float* w = (float*)workspace;
  float* fuse_in = w;  w += kWsFuseIn * n;
  float* f1 = w;       w += kWsF1 * n;
  float* f2 = w;       w += kWsF2 * n;
  float* f3 = w;       w += kWsF3 * n;
  float* temp_in = w;  w += kWsT * n;
  float* t1 = w;       w += kWsT * n;
  float* t2 = w;       w += kWsT * n;
  float* t3 = w;       w += kWsT * n;
  float* head_in = w;  w += kWsH * n;
  float* h1 = w;       w += kWsH * n;
  float* h2 = w;       w += kWsH * n;
  float* ho = w;       w += kWsHo * n;
  float* skel = w;
  void* head_ws = (uint8_t*)workspace + (int64_t)sizeof(float) * kWsPerSample * (int64_t)n + 256;

  // every conv of the tail: error-compensated fp16 x 3 tensor-core GEMM when the pre-split operand
  // blob is given, else the TF32 x 3 kernel on the fp32 blob (both fp32-grade; see conv_f32.cu)
  static const bool force_tf32 = getenv("AT_TAIL_TF32") != nullptr && atoi(getenv("AT_TAIL_TF32")) != 0;
  const bool use_f16 = f16_blob != nullptr && !force_tf32;
  auto conv = [&](int id, const float* x, int x_ld, const float* res, int res_ld, float* y, int y_ld, bool relu) -> int {
    if (use_f16)
      return conv_nhwc_f16x3(T[id], blob, (const uint8_t*)f16_blob + f16_blob_offset(id), x, x_ld, 0, res, res_ld, y, y_ld,
                             0, n, 6, 6, relu, s);
    return conv_nhwc_tf32x3(T[id], blob, x, x_ld, 0, res, res_ld, y, y_ld, 0, n, 6, 6, relu, s);
  };

  PrepArgs pa;
  pa.in = *in; pa.st = *state; pa.fuse_in = fuse_in; pa.temp_in = temp_in; pa.status = out->status;
  launch_pdl(tail_prepare_kernel, dim3(n), dim3(128), 0, s, pa);
  int rc = check_launch("tail_prepare_kernel");
  if (rc) return rc;
  // 2-view fusion (model_utils.py:141-163).  Rows of single-view samples are computed on
  // whatever the workspace holds and ignored afterwards.
  if ((rc = conv(L_FUSION + 0, fuse_in, 144, nullptr, 0, f1, 108, true))) return rc;
  if ((rc = conv(L_FUSION + 1, f1, 108, nullptr, 0, f2, 72, true))) return rc;
  if ((rc = conv(L_FUSION + 2, f2, 72, nullptr, 0, f3, 72, false))) return rc;
  {
    launch_pdl(post_fusion_kernel, dim3((unsigned)n), dim3(256), 0, s, f3, in->intrinsics, in->sample_range, n, c, temp_in);
    if ((rc = check_launch("post_fusion_kernel"))) return rc;
  }
  // SimpleConvRNN step (temporal.py:76-91): 3x conv1x1(90->90), ReLU after the first two
  if ((rc = conv(L_TEMPORAL + 0, temp_in, 90, nullptr, 0, t1, 90, true))) return rc;
  if ((rc = conv(L_TEMPORAL + 1, t1, 90, nullptr, 0, t2, 90, true))) return rc;
  if ((rc = conv(L_TEMPORAL + 2, t2, 90, nullptr, 0, t3, 90, false))) return rc;
  const int hc = known ? 76 : 72;
  if (known) {
    const ConvLayer& L = T[L_SKEL];
    launch_pdl(skeleton_kernel, dim3(in->n_skel), dim3(160), 0, s, in->skel_axes, in->skel_rest, blob + L.weight_offset,
                                               blob + L.bias_offset, skel);
    if ((rc = check_launch("skeleton_kernel"))) return rc;
  }
  {
    launch_pdl(split_kernel, dim3((unsigned)n), dim3(256), 0, s, t3, in->memory_idx, state->capacity, state->mem_features,
                                                               known ? skel : nullptr, in->n_skel, n, hc, head_in);
    if ((rc = check_launch("split_kernel"))) return rc;
  }
  // pose head (model_utils.py:195-208): 2 BasicBlocks + conv1x1 + global average pool
  const int hb = known ? L_HEAD_K : L_HEAD_U;
  // the four 3x3 convs: tcgen05 shift-GEMMs on split fp16 planes (conv_tc.cu) unless AT_TAIL_HEAD_TC=0
  const char* head_env = getenv("AT_TAIL_HEAD_TC");  // read per call: tests toggle it
  const bool head_tc = head_env == nullptr || atoi(head_env) != 0;
  if (use_f16 && head_tc) {
    const void* wq[4];
    for (int i = 0; i < 4; ++i) wq[i] = (const uint8_t*)f16_blob + head_tc_blob_offset(hb + i);
    if ((rc = head_tc_forward(T + hb, blob, wq, head_in, h2, head_in, hc, n, head_ws, s))) return rc;
  } else {
    if ((rc = conv(hb + 0, head_in, hc, nullptr, 0, h1, hc, true))) return rc;
    if ((rc = conv(hb + 1, h1, hc, head_in, hc, h2, hc, true))) return rc;
    if ((rc = conv(hb + 2, h2, hc, nullptr, 0, h1, hc, true))) return rc;
    if ((rc = conv(hb + 3, h1, hc, h2, hc, head_in, hc, true))) return rc;
  }
  if ((rc = conv(hb + 4, head_in, hc, nullptr, 0, ho, 64, false))) return rc;
  DecodeArgs da;
  da.head_out = ho; da.ld = 64; da.n_out = known ? 62 : 63;
  da.extrinsics = in->extrinsics; da.sample_range = in->sample_range; da.hand_idx = in->hand_idx;
  da.n = n; da.c = c; da.predict_scale = known ? 0 : 1; da.out = *out;
  launch_pdl(decode_kernel, dim3(n), dim3(64), 0, s, da);
  return check_launch("decode_kernel");
}
