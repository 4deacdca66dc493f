// Fused rotation-parameterisation + forward kinematics, forward and backward: one warp per frame,
// one lane per joint, the kinematic chain walked level by level with warp shuffles.
//
// Restates (does not copy) third_party/smplx/smplx/lbs.py:
//   batch_rodrigues :299-333, transform_mat + batch_rigid_transform :336-405, the pose feature :226-231,
//   vertices2joints :255-272 folded with blend_shapes :275-296 into J0 + Jdirs.beta (model.cu),
// and pytorch3d's rotation_6d_to_matrix(matrix_to_rotation_6d(X)) / axis_angle_to_matrix([0,0,z]) as the
// closures call them (src/video_mocap/optimization.py:192-202,335-340,672-679).
#include "common.cuh"

#define FULL 0xffffffffu

struct LaneFK {
  float raw[9];   // the parameter as given (3 used for axis-angle)
  float M[9];     // after the optional yaw pre-rotation (root only), before orthonormalisation
  float R[9];     // effective local rotation
  float Jr[3];    // rest joint
  float rel[3];   // rest joint relative to the parent
  float pRG[9];   // parent's global rotation (identity for the root)
  float RG[9];    // global rotation
  float tG[3];    // global translation = posed joint
  float cz, sz;   // cos/sin of the yaw (root only)
};

__device__ __forceinline__ void mat3_mul(const float* a, const float* b, float* c) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) c[i * 3 + k] = a[i * 3] * b[k] + a[i * 3 + 1] * b[3 + k] + a[i * 3 + 2] * b[6 + k];
}

// lbs.py:299-333 (note the +1e-8 inside the norm, and axis = r / angle, not (r+1e-8)/angle)
__device__ __forceinline__ void rodrigues_fwd(const float* r, float* R) {
  float qx = r[0] + 1e-8f, qy = r[1] + 1e-8f, qz = r[2] + 1e-8f;
  float ang = sqrtf(qx * qx + qy * qy + qz * qz);
  float nx = r[0] / ang, ny = r[1] / ang, nz = r[2] / ang;
  float s = sinf(ang), c = cosf(ang);
  float K[9] = {0.f, -nz, ny, nz, 0.f, -nx, -ny, nx, 0.f};
  float KK[9];
  mat3_mul(K, K, KK);
  float oc = 1.f - c;
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = ((i % 4 == 0) ? 1.f : 0.f) + s * K[i] + oc * KK[i];
}

__device__ __forceinline__ void rodrigues_bwd(const float* r, const float* gR, float* gr) {
  float qx = r[0] + 1e-8f, qy = r[1] + 1e-8f, qz = r[2] + 1e-8f;
  float ang = sqrtf(qx * qx + qy * qy + qz * qz);
  float inv = 1.f / ang;
  float nx = r[0] * inv, ny = r[1] * inv, nz = r[2] * inv;
  float s = sinf(ang), c = cosf(ang), oc = 1.f - c;
  float K[9] = {0.f, -nz, ny, nz, 0.f, -nx, -ny, nx, 0.f};
  float KK[9];
  mat3_mul(K, K, KK);
  float gth = 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) gth += gR[i] * (c * K[i] + s * KK[i]);
  // gK = s*gR + oc*(gR K^T + K^T gR)
  float Kt[9] = {K[0], K[3], K[6], K[1], K[4], K[7], K[2], K[5], K[8]};
  float t1[9], t2[9], gK[9];
  mat3_mul(gR, Kt, t1);
  mat3_mul(Kt, gR, t2);
#pragma unroll
  for (int i = 0; i < 9; ++i) gK[i] = s * gR[i] + oc * (t1[i] + t2[i]);
  float gn[3] = {gK[7] - gK[5], gK[2] - gK[6], gK[3] - gK[1]};
  float dotnr = gn[0] * r[0] + gn[1] * r[1] + gn[2] * r[2];
  gth -= dotnr * inv * inv;
  float q[3] = {qx, qy, qz};
#pragma unroll
  for (int k = 0; k < 3; ++k) gr[k] = gn[k] * inv + gth * q[k] * inv;
}

// pytorch3d rotation_6d_to_matrix(matrix_to_rotation_6d(X)): rows 0 and 1, F.normalize eps 1e-12
__device__ __forceinline__ void sixd_fwd(const float* X, float* R) {
  float n1 = fmaxf(sqrtf(X[0] * X[0] + X[1] * X[1] + X[2] * X[2]), 1e-12f);
  float b1[3] = {X[0] / n1, X[1] / n1, X[2] / n1};
  float d = b1[0] * X[3] + b1[1] * X[4] + b1[2] * X[5];
  float u[3] = {X[3] - d * b1[0], X[4] - d * b1[1], X[5] - d * b1[2]};
  float n2 = fmaxf(sqrtf(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]), 1e-12f);
  float b2[3] = {u[0] / n2, u[1] / n2, u[2] / n2};
  R[0] = b1[0]; R[1] = b1[1]; R[2] = b1[2];
  R[3] = b2[0]; R[4] = b2[1]; R[5] = b2[2];
  R[6] = b1[1] * b2[2] - b1[2] * b2[1];
  R[7] = b1[2] * b2[0] - b1[0] * b2[2];
  R[8] = b1[0] * b2[1] - b1[1] * b2[0];
}

__device__ __forceinline__ void cross3(const float* a, const float* b, float* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

__device__ __forceinline__ void sixd_bwd(const float* X, const float* gR, float* gX) {
  float n1r = sqrtf(X[0] * X[0] + X[1] * X[1] + X[2] * X[2]);
  float n1 = fmaxf(n1r, 1e-12f);
  float b1[3] = {X[0] / n1, X[1] / n1, X[2] / n1};
  const float* a2 = X + 3;
  float d = b1[0] * a2[0] + b1[1] * a2[1] + b1[2] * a2[2];
  float u[3] = {a2[0] - d * b1[0], a2[1] - d * b1[1], a2[2] - d * b1[2]};
  float n2r = sqrtf(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
  float n2 = fmaxf(n2r, 1e-12f);
  float b2[3] = {u[0] / n2, u[1] / n2, u[2] / n2};
  float gb1[3] = {gR[0], gR[1], gR[2]}, gb2[3] = {gR[3], gR[4], gR[5]};
  const float* gb3 = gR + 6;
  float t[3];
  cross3(b2, gb3, t);            // b3 = b1 x b2  ->  g_b1 += b2 x g_b3
  gb1[0] += t[0]; gb1[1] += t[1]; gb1[2] += t[2];
  cross3(gb3, b1, t);            //               ->  g_b2 += g_b3 x b1
  gb2[0] += t[0]; gb2[1] += t[1]; gb2[2] += t[2];
  float gu[3];
  if (n2r > 1e-12f) {
    float p = gb2[0] * b2[0] + gb2[1] * b2[1] + gb2[2] * b2[2];
#pragma unroll
    for (int k = 0; k < 3; ++k) gu[k] = (gb2[k] - p * b2[k]) / n2;
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) gu[k] = gb2[k] / n2;
  }
  float gd = -(gu[0] * b1[0] + gu[1] * b1[1] + gu[2] * b1[2]);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    gX[3 + k] = gu[k] + gd * b1[k];
    gb1[k] += -d * gu[k] + gd * a2[k];
  }
  if (n1r > 1e-12f) {
    float p = gb1[0] * b1[0] + gb1[1] * b1[1] + gb1[2] * b1[2];
#pragma unroll
    for (int k = 0; k < 3; ++k) gX[k] = (gb1[k] - p * b1[k]) / n1;
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) gX[k] = gb1[k] / n1;
  }
  gX[6] = gX[7] = gX[8] = 0.f;
}

// everything a lane needs about its joint; called by all 32 lanes of the warp (lanes >= J idle but
// take part in the shuffles)
__device__ __forceinline__ void fk_lane_forward(const uuo_tree& tree, const float* __restrict__ J0,
                                                const float* __restrict__ Jdirs, int NB,
                                                const float* __restrict__ betas_f, const float* __restrict__ go_f,
                                                const float* __restrict__ bp_f, int pose_mode, const float* yaw_f,
                                                int lane, LaneFK& s) {
  const int J = tree.J;
  const bool act = lane < J;
  const int jl = act ? lane : 0;
  const int ne = (pose_mode == UUO_POSE_AXISANGLE) ? 3 : 9;
  const float* src = (jl == 0) ? go_f : bp_f + (size_t)(jl - 1) * ne;
#pragma unroll
  for (int i = 0; i < 9; ++i) s.raw[i] = (i < ne) ? src[i] : 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) s.M[i] = s.raw[i];
  s.cz = 1.f; s.sz = 0.f;
  if (yaw_f != nullptr && jl == 0 && pose_mode != UUO_POSE_AXISANGLE) {
    float z = *yaw_f;
    s.cz = cosf(z); s.sz = sinf(z);
    // Rz(z) @ raw
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      s.M[k] = s.cz * s.raw[k] - s.sz * s.raw[3 + k];
      s.M[3 + k] = s.sz * s.raw[k] + s.cz * s.raw[3 + k];
      s.M[6 + k] = s.raw[6 + k];
    }
  }
  if (pose_mode == UUO_POSE_AXISANGLE) rodrigues_fwd(s.raw, s.R);
  else if (pose_mode == UUO_POSE_RAW6D) sixd_fwd(s.M, s.R);
  else {
#pragma unroll
    for (int i = 0; i < 9; ++i) s.R[i] = s.M[i];
  }
  // rest joint (lbs.py:209-213 folded)
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float a = J0[jl * 3 + k];
    for (int l = 0; l < NB; ++l) a += Jdirs[(jl * 3 + k) * NB + l] * betas_f[l];
    s.Jr[k] = a;
  }
  const int par = act ? tree.parents[jl] : -1;
  const int psrc = par >= 0 ? par : lane;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float pj = __shfl_sync(FULL, s.Jr[k], psrc);
    s.rel[k] = par >= 0 ? s.Jr[k] - pj : s.Jr[k];          // lbs.py:378-379
  }
  // chain (lbs.py:387-393) level by level
#pragma unroll
  for (int i = 0; i < 9; ++i) { s.RG[i] = s.R[i]; s.pRG[i] = (i % 4 == 0) ? 1.f : 0.f; }
#pragma unroll
  for (int k = 0; k < 3; ++k) s.tG[k] = s.rel[k];
  const int mydepth = act ? tree.depth[jl] : -1;
  for (int d = 1; d <= tree.max_depth; ++d) {
    float pR[9], pt[3];
#pragma unroll
    for (int i = 0; i < 9; ++i) pR[i] = __shfl_sync(FULL, s.RG[i], psrc);
#pragma unroll
    for (int k = 0; k < 3; ++k) pt[k] = __shfl_sync(FULL, s.tG[k], psrc);
    if (mydepth == d) {
      mat3_mul(pR, s.R, s.RG);
#pragma unroll
      for (int k = 0; k < 3; ++k)
        s.tG[k] = pR[k * 3] * s.rel[0] + pR[k * 3 + 1] * s.rel[1] + pR[k * 3 + 2] * s.rel[2] + pt[k];
#pragma unroll
      for (int i = 0; i < 9; ++i) s.pRG[i] = pR[i];
    }
  }
}

__global__ void __launch_bounds__(128) fk_forward_kernel(
    uuo_tree tree, const float* __restrict__ J0, const float* __restrict__ Jdirs, int NB, int P, int KP,
    const float* __restrict__ betas, int betas_rows, const float* __restrict__ go, const float* __restrict__ bp,
    int pose_mode, const float* __restrict__ yaw, const float* __restrict__ transl, int F, int n_joints_out,
    float* __restrict__ coef, float* __restrict__ A, float* __restrict__ joints) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (f >= F) return;
  const int J = tree.J;
  const int ne = (pose_mode == UUO_POSE_AXISANGLE) ? 3 : 9;
  // betas_rows in {1, G, F} with G | F: row of frame f (G stacked hypothesis groups of F / G consecutive frames each)
  const float* betas_f = betas + (size_t)(betas_rows == 1 ? 0 : f / (F / betas_rows)) * NB;
  LaneFK s;
  fk_lane_forward(tree, J0, Jdirs, NB, betas_f, go + (size_t)f * ne, bp + (size_t)f * (J - 1) * ne, pose_mode,
                  yaw ? yaw + f : nullptr, lane, s);
  float* cf = coef + (size_t)f * KP;
  if (lane < J) {
    if (lane >= 1) {
#pragma unroll
      for (int i = 0; i < 9; ++i) cf[(lane - 1) * 9 + i] = s.R[i] - ((i % 4 == 0) ? 1.f : 0.f);   // lbs.py:227
    }
    float* a = A + ((size_t)f * J + lane) * 12;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      float rj = s.RG[r * 3] * s.Jr[0] + s.RG[r * 3 + 1] * s.Jr[1] + s.RG[r * 3 + 2] * s.Jr[2];
      reinterpret_cast<float4*>(a)[r] = make_float4(s.RG[r * 3], s.RG[r * 3 + 1], s.RG[r * 3 + 2], s.tG[r] - rj);
    }
    if (joints) {
      float* jo = joints + ((size_t)f * n_joints_out + lane) * 3;
#pragma unroll
      for (int k = 0; k < 3; ++k) jo[k] = s.tG[k] + (transl ? transl[(size_t)f * 3 + k] : 0.f);
    }
  }
  for (int k = P + lane; k < KP; k += 32) cf[k] = (k < P + NB) ? betas_f[k - P] : 0.f;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

__global__ void __launch_bounds__(128) fk_backward_kernel(
    uuo_tree tree, const float* __restrict__ J0, const float* __restrict__ Jdirs, int NB, int P, int KP,
    const float* __restrict__ betas, int betas_rows, const float* __restrict__ go, const float* __restrict__ bp,
    int pose_mode, const float* __restrict__ yaw, int F, int n_joints_out, const float* __restrict__ g_joints,
    const float* __restrict__ gA_parts, int n_ga, const float* __restrict__ gcoef_parts, int n_gc,
    float* __restrict__ gbetas_f, float* __restrict__ g_go, float* __restrict__ g_bp, float* __restrict__ g_yaw,
    float* __restrict__ g_transl) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (f >= F) return;
  const int J = tree.J;
  const bool act = lane < J;
  const int jl = act ? lane : 0;
  const int ne = (pose_mode == UUO_POSE_AXISANGLE) ? 3 : 9;
  // betas_rows in {1, G, F} with G | F: row of frame f (G stacked hypothesis groups of F / G consecutive frames each)
  const float* betas_f = betas + (size_t)(betas_rows == 1 ? 0 : f / (F / betas_rows)) * NB;
  LaneFK s;
  fk_lane_forward(tree, J0, Jdirs, NB, betas_f, go + (size_t)f * ne, bp + (size_t)f * (J - 1) * ne, pose_mode,
                  yaw ? yaw + f : nullptr, lane, s);

  // upstream: dA (3x4 per joint) summed over the deterministic partials, posed-joint grads
  float gAm[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) gAm[i] = 0.f;
  float gsum[3] = {0.f, 0.f, 0.f};
  for (int p = 0; p < n_ga; ++p) {
    const float* src = gA_parts + ((size_t)p * F + f) * UUO_GA_STRIDE;
    if (act) {
      const float4* s4 = reinterpret_cast<const float4*>(src + jl * 12);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        float4 v = s4[r];
        gAm[r * 4] += v.x; gAm[r * 4 + 1] += v.y; gAm[r * 4 + 2] += v.z; gAm[r * 4 + 3] += v.w;
      }
    }
    if (lane < 3) gsum[0] += src[288 + lane];
  }
  float gtG[3], gRG[9], gJr[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    float gj = (g_joints && act) ? g_joints[((size_t)f * n_joints_out + jl) * 3 + r] : 0.f;
    float gAt = act ? gAm[r * 4 + 3] : 0.f;
    gtG[r] = gAt + gj;
#pragma unroll
    for (int c = 0; c < 3; ++c) gRG[r * 3 + c] = act ? gAm[r * 4 + c] - gAt * s.Jr[c] : 0.f;   // A_t = tG - RG Jr
  }
#pragma unroll
  for (int c = 0; c < 3; ++c)
    gJr[c] = act ? -(s.RG[c] * gAm[3] + s.RG[3 + c] * gAm[7] + s.RG[6 + c] * gAm[11]) : 0.f;

  // g_transl = sum_v g_vertices incl. extra-joint rows (partials, lanes 0..2) + sum over the J posed joints of g_joints
  if (g_transl) {
    float tj[3] = {0.f, 0.f, 0.f};
    if (g_joints) {
      for (int q = lane; q < J; q += 32) {   // extra joints are vertices: already inside the vertex partials
        const float* gq = g_joints + ((size_t)f * n_joints_out + q) * 3;
        tj[0] += gq[0]; tj[1] += gq[1]; tj[2] += gq[2];
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) tj[k] = warp_sum(tj[k]);
    float mine = gsum[0];   // lane k (<3) holds component k
    if (lane < 3) g_transl[(size_t)f * 3 + lane] = mine + (lane == 0 ? tj[0] : (lane == 1 ? tj[1] : tj[2]));
  }

  // reverse sweep over the tree: deepest level first
  const int mydepth = act ? tree.depth[jl] : -1;
  float gR[9], grel[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) gR[i] = 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) grel[k] = 0.f;
  int ch[UUO_MAX_CHILD];
#pragma unroll
  for (int c = 0; c < UUO_MAX_CHILD; ++c) ch[c] = act ? tree.children[jl][c] : -1;
  for (int d = tree.max_depth; d >= 1; --d) {
    // contribution of a depth-d joint to its parent: gRG_p += gRG R^T + gtG (x) rel ; gtG_p += gtG
    float cR[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        cR[r * 3 + c] = gRG[r * 3] * s.R[c * 3] + gRG[r * 3 + 1] * s.R[c * 3 + 1] + gRG[r * 3 + 2] * s.R[c * 3 + 2] +
                        gtG[r] * s.rel[c];
    if (mydepth == d) {
      // own local grads: gR = pRG^T gRG ; grel = pRG^T gtG
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          gR[r * 3 + c] = s.pRG[r] * gRG[c] + s.pRG[3 + r] * gRG[3 + c] + s.pRG[6 + r] * gRG[6 + c];
        grel[r] = s.pRG[r] * gtG[0] + s.pRG[3 + r] * gtG[1] + s.pRG[6 + r] * gtG[2];
      }
    }
#pragma unroll
    for (int c = 0; c < UUO_MAX_CHILD; ++c) {
      if (!((tree.slot_mask[d] >> c) & 1)) continue;      // no joint has a depth-d child in this slot (uniform)
      const int cj = ch[c];
      const bool valid = cj >= 0 && tree.depth[cj >= 0 ? cj : 0] == d;
      const int srcl = valid ? cj : lane;
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        float v = __shfl_sync(FULL, cR[i], srcl);
        if (valid) gRG[i] += v;
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        float v = __shfl_sync(FULL, gtG[k], srcl);
        if (valid) gtG[k] += v;
      }
    }
  }
  if (mydepth == 0) {
#pragma unroll
    for (int i = 0; i < 9; ++i) gR[i] = gRG[i];
#pragma unroll
    for (int k = 0; k < 3; ++k) grel[k] = gtG[k];
  }
  // rel = Jr - Jr[parent]
#pragma unroll
  for (int k = 0; k < 3; ++k) gJr[k] += grel[k];
#pragma unroll
  for (int c = 0; c < UUO_MAX_CHILD; ++c) {
    const int cj = ch[c];
    const int srcl = cj >= 0 ? cj : lane;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float v = __shfl_sync(FULL, grel[k], srcl);
      if (cj >= 0) gJr[k] -= v;
    }
  }

  // coefficient grads (posedirs/shapedirs transpose GEMM partials)
  float gpf[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) gpf[i] = 0.f;
  float gb_lane = 0.f;   // lane l < NB: d/d beta_l from the shape blend
  for (int p = 0; p < n_gc; ++p) {
    const float* src = gcoef_parts + ((size_t)p * F + f) * KP;
    if (act && jl >= 1) {
#pragma unroll
      for (int i = 0; i < 9; ++i) gpf[i] += src[(jl - 1) * 9 + i];
    }
    if (lane < NB) gb_lane += src[P + lane];
  }
  if (act && jl >= 1) {
#pragma unroll
    for (int i = 0; i < 9; ++i) gR[i] += gpf[i];
  }

  // betas: Jr = J0 + Jdirs . beta
  if (gbetas_f) {
    for (int l = 0; l < NB; ++l) {
      float v = 0.f;
      if (act) v = gJr[0] * Jdirs[(jl * 3 + 0) * NB + l] + gJr[1] * Jdirs[(jl * 3 + 1) * NB + l] +
                   gJr[2] * Jdirs[(jl * 3 + 2) * NB + l];
      v = warp_sum(v);
      if (lane == l) gb_lane += v;
    }
    if (lane < NB) gbetas_f[(size_t)f * 16 + lane] = gb_lane;
  }

  // parameterisation backward
  if (act) {
    float gM[9];
    if (pose_mode == UUO_POSE_AXISANGLE) {
      float gr[3];
      rodrigues_bwd(s.raw, gR, gr);
      float* dst = (jl == 0) ? (g_go ? g_go + (size_t)f * 3 : nullptr)
                             : (g_bp ? g_bp + ((size_t)f * (J - 1) + (jl - 1)) * 3 : nullptr);
      if (dst) { dst[0] = gr[0]; dst[1] = gr[1]; dst[2] = gr[2]; }
    } else {
      if (pose_mode == UUO_POSE_RAW6D) sixd_bwd(s.M, gR, gM);
      else {
#pragma unroll
        for (int i = 0; i < 9; ++i) gM[i] = gR[i];
      }
      float graw[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) graw[i] = gM[i];
      if (yaw != nullptr && jl == 0) {
        // M = Rz raw  ->  g_raw = Rz^T gM ; g_z = sum gM .* (dRz/dz raw)
        float gz = 0.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          graw[k] = s.cz * gM[k] + s.sz * gM[3 + k];
          graw[3 + k] = -s.sz * gM[k] + s.cz * gM[3 + k];
          gz += gM[k] * (-s.sz * s.raw[k] - s.cz * s.raw[3 + k]) + gM[3 + k] * (s.cz * s.raw[k] - s.sz * s.raw[3 + k]);
        }
        if (g_yaw) g_yaw[f] = gz;
      }
      float* dst = (jl == 0) ? (g_go ? g_go + (size_t)f * 9 : nullptr)
                             : (g_bp ? g_bp + ((size_t)f * (J - 1) + (jl - 1)) * 9 : nullptr);
      if (dst) {
#pragma unroll
        for (int i = 0; i < 9; ++i) dst[i] = graw[i];
      }
    }
  }
}

// deterministic reduction of per-frame beta grads to the shared row (body_models.py:371-376 broadcast).
// ONE CTA of 1024 threads: thread t sums frames t, t + 1024, ... for all 16 slots of a frame (one 64 B row per load,
// rows of consecutive threads are consecutive), then a fixed-order tree over the threads.  (The r1g version ran one
// 256-thread CTA per beta with 4 B loads 64 B apart: 14 us at F = 5000, 3 % of a fitting iteration.)
// one CTA per betas row: g_betas[g] = sum over the row's F frames (frames g F .. g F + F - 1)
__global__ void __launch_bounds__(1024) reduce_gbetas_kernel(const float* __restrict__ gbetas_f, int F, int NB,
                                                             float* __restrict__ g_betas) {
  gbetas_f += (size_t)blockIdx.x * F * 16;
  g_betas += (size_t)blockIdx.x * NB;
  __shared__ float sh[32][17];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = 0.f;
  for (int f = tid; f < F; f += 1024) {
    const float4* src = reinterpret_cast<const float4*>(gbetas_f + (size_t)f * 16);
    const float4 x0 = __ldg(src), x1 = __ldg(src + 1), x2 = __ldg(src + 2), x3 = __ldg(src + 3);
    a[0] += x0.x; a[1] += x0.y; a[2] += x0.z; a[3] += x0.w; a[4] += x1.x; a[5] += x1.y; a[6] += x1.z; a[7] += x1.w;
    a[8] += x2.x; a[9] += x2.y; a[10] += x2.z; a[11] += x2.w; a[12] += x3.x; a[13] += x3.y; a[14] += x3.z; a[15] += x3.w;
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
  }
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < 16; ++i) sh[w][i] = a[i];
  }
  __syncthreads();
  if (tid < NB) {
    float t = 0.f;
    for (int q = 0; q < 32; ++q) t += sh[q][tid];
    g_betas[tid] = t;
  }
}

__global__ void copy_gbetas_kernel(const float* __restrict__ gbetas_f, int F, int NB, float* __restrict__ g_betas) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < F * NB) g_betas[i] = gbetas_f[(size_t)(i / NB) * 16 + (i % NB)];
}

int launch_fk_forward(const uuo_model* m, const float* betas, int betas_rows, const float* go, const float* bp,
                      int pose_mode, const float* yaw, const float* transl, int F, float* joints,
                      const uuo_ws& ws, cudaStream_t st) {
  uuo_tree t = make_tree(m);
  dim3 grid((F + 3) / 4), block(128);
  { ProfScope ps_("fk_forward", st);
  fk_forward_kernel<<<grid, block, 0, st>>>(t, m->J0, m->Jdirs, m->NB, m->P, m->KP, betas, betas_rows, go, bp,
                                            pose_mode, yaw, transl, F, m->J + m->NE, ws.coef, ws.A, joints);
  } UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_fk_backward(const uuo_model* m, const float* betas, int betas_rows, const float* go, const float* bp,
                       int pose_mode, const float* yaw, int F, const float* g_joints, const uuo_ws& ws,
                       int n_ga_parts, int n_gc_parts, const float*, float* g_betas, float* g_go, float* g_bp,
                       float* g_yaw, float* g_transl, cudaStream_t st) {
  uuo_tree t = make_tree(m);
  dim3 grid((F + 3) / 4), block(128);
  { ProfScope ps_("fk_backward", st);
  fk_backward_kernel<<<grid, block, 0, st>>>(t, m->J0, m->Jdirs, m->NB, m->P, m->KP, betas, betas_rows, go, bp,
                                             pose_mode, yaw, F, m->J + m->NE, g_joints, ws.gA, n_ga_parts,
                                             ws.gcoef, n_gc_parts, g_betas ? ws.gbetas : nullptr, g_go, g_bp,
                                             g_yaw, g_transl);
  } UUO_LAUNCH_CHECK();
  if (g_betas) {
    if (betas_rows < F) reduce_gbetas_kernel<<<betas_rows, 1024, 0, st>>>(ws.gbetas, F / betas_rows, m->NB, g_betas);
    else copy_gbetas_kernel<<<(F * m->NB + 255) / 256, 256, 0, st>>>(ws.gbetas, F, m->NB, g_betas);
    UUO_LAUNCH_CHECK();
  }
  return UUO_OK;
}
