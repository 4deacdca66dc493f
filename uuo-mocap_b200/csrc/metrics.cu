// Evaluation metrics of the step right after the fitting path:
//   /root/reference/src/video_mocap/evaluation/metrics.py:49-135  MPJPE / MPJVE / V2V (+ joint subsets)
//   :138-190 compute_similarity_transform (batched orthogonal Procrustes: 3x3 SVD per frame) for the PA- variants
//   :26-46   compute_marker_to_surface_distance (igl.signed_distance there; here the unsigned point-to-triangle
//            distance, which is what mean(abs(signed distance)) measures)
// Sums are block partials added in a fixed order by the last kernel (deterministic, no float atomics).
#include "common.cuh"

// ---- mean over (frame, selected point) of || (a - b) ||, optionally on frame-to-frame velocities ---------------
// partial[b] = sum over this block's items; items = (F - velocity) * n_sel
__global__ void __launch_bounds__(256) point_error_kernel(const float* __restrict__ a, const float* __restrict__ b, int F,
                                                          int P, const int32_t* __restrict__ sel, int n_sel, int velocity,
                                                          float freq, float* __restrict__ partial) {
  __shared__ float red[256];
  const size_t n_items = (size_t)(F - velocity) * n_sel;
  float acc = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_items; i += (size_t)gridDim.x * blockDim.x) {
    const int f = (int)(i / n_sel), s = (int)(i - (size_t)f * n_sel);
    const int p = sel ? sel[s] : s;
    const float* pa = a + ((size_t)f * P + p) * 3;
    const float* pb = b + ((size_t)f * P + p) * 3;
    float e[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (velocity) {   // (x[t+1] - x[t]) * freq on both, then the difference (metrics.py:70-73)
        const float va = (pa[(size_t)P * 3 + k] - pa[k]) * freq, vb = (pb[(size_t)P * 3 + k] - pb[k]) * freq;
        e[k] = va - vb;
      } else {
        e[k] = pa[k] - pb[k];
      }
    }
    acc += sqrtf(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = red[0];
}

__global__ void finish_mean_kernel(const float* __restrict__ partial, int n, double count, float* __restrict__ out) {
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += (double)partial[i];
  out[0] = (float)(s / count);
}

// ---- batched similarity transform (Procrustes) ---------------------------------------------------------------
// one-sided Jacobi SVD of a 3x3 matrix in double: K = U diag(s) V^T
__device__ void svd3(const double K[9], double U[9], double S[3], double V[9]) {
  double A[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) { A[i] = K[i]; V[i] = (i % 4 == 0) ? 1.0 : 0.0; }
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int i = 0; i < 3; ++i) {
          alpha += A[i * 3 + p] * A[i * 3 + p];
          beta += A[i * 3 + q] * A[i * 3 + q];
          gamma += A[i * 3 + p] * A[i * 3 + q];
        }
        off = fmax(off, fabs(gamma) / sqrt(fmax(alpha * beta, 1e-300)));
        if (fabs(gamma) < 1e-300) continue;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
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
  for (int j = 0; j < 3; ++j) {
    S[j] = sqrt(A[j] * A[j] + A[3 + j] * A[3 + j] + A[6 + j] * A[6 + j]);
    for (int i = 0; i < 3; ++i) U[i * 3 + j] = S[j] > 1e-300 ? A[i * 3 + j] / S[j] : (i == j ? 1.0 : 0.0);
  }
  // descending singular values like torch.svd (the construction below only needs the LAST one to be the smallest)
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2 - a; ++b)
      if (S[b] < S[b + 1]) {
        double t = S[b]; S[b] = S[b + 1]; S[b + 1] = t;
        for (int i = 0; i < 3; ++i) {
          t = U[i * 3 + b]; U[i * 3 + b] = U[i * 3 + b + 1]; U[i * 3 + b + 1] = t;
          t = V[i * 3 + b]; V[i * 3 + b] = V[i * 3 + b + 1]; V[i * 3 + b + 1] = t;
        }
      }
}

__device__ __forceinline__ double det3(const double M[9]) {
  return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// S1_hat[f] = scale * R * S1[f] + t with (scale, R, t) the similarity transform closest to S2[f]  (metrics.py:138-190)
__global__ void __launch_bounds__(128) similarity_transform_kernel(const float* __restrict__ S1, const float* __restrict__ S2,
                                                                   int F, int N, float* __restrict__ out) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const float* a = S1 + (size_t)f * N * 3;
  const float* b = S2 + (size_t)f * N * 3;
  double mu1[3] = {0, 0, 0}, mu2[3] = {0, 0, 0};
  for (int i = 0; i < N; ++i)
    for (int k = 0; k < 3; ++k) { mu1[k] += a[i * 3 + k]; mu2[k] += b[i * 3 + k]; }
  for (int k = 0; k < 3; ++k) { mu1[k] /= N; mu2[k] /= N; }
  double var1 = 0, K[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};       // K = X1 X2^T  (3x3)
  for (int i = 0; i < N; ++i) {
    double x1[3], x2[3];
    for (int k = 0; k < 3; ++k) { x1[k] = a[i * 3 + k] - mu1[k]; x2[k] = b[i * 3 + k] - mu2[k]; var1 += x1[k] * x1[k]; }
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) K[r * 3 + c] += x1[r] * x2[c];
  }
  double U[9], S[3], V[9];
  svd3(K, U, S, V);
  // Z fixes det(R) = +1: R = V Z U^T, Z = diag(1, 1, sign(det(U V^T)))
  double UVt[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) UVt[r * 3 + c] = U[r * 3] * V[c * 3] + U[r * 3 + 1] * V[c * 3 + 1] + U[r * 3 + 2] * V[c * 3 + 2];
  const double d = det3(UVt);
  const double z = d > 0 ? 1.0 : (d < 0 ? -1.0 : 0.0);
  double R[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) R[r * 3 + c] = V[r * 3] * U[c * 3] + V[r * 3 + 1] * U[c * 3 + 1] + z * V[r * 3 + 2] * U[c * 3 + 2];
  double trace = 0;                                          // trace(R K)
  for (int r = 0; r < 3; ++r)
    for (int k = 0; k < 3; ++k) trace += R[r * 3 + k] * K[k * 3 + r];
  const double scale = trace / var1;
  double t[3];
  for (int r = 0; r < 3; ++r) t[r] = mu2[r] - scale * (R[r * 3] * mu1[0] + R[r * 3 + 1] * mu1[1] + R[r * 3 + 2] * mu1[2]);
  float* o = out + (size_t)f * N * 3;
  for (int i = 0; i < N; ++i)
    for (int r = 0; r < 3; ++r)
      o[i * 3 + r] = (float)(scale * (R[r * 3] * a[i * 3] + R[r * 3 + 1] * a[i * 3 + 1] + R[r * 3 + 2] * a[i * 3 + 2]) + t[r]);
}

// ---- marker-to-surface distance: unsigned distance from every marker to the closest triangle of the frame's mesh ----
// One CTA per frame: the frame's vertices staged in shared memory, faces streamed from L2 (shared by all frames),
// a thread owns faces tid, tid+blockDim, ... and 8 markers at a time; per-marker minima merged with atomicMin on the
// bit pattern of the (non-negative) squared distance.
#define PM_MG 8
__device__ __forceinline__ float point_tri_dist2(const float p[3], const float a[3], const float ab[3], const float ac[3]) {
  // closest point on a triangle (Ericson, Real-Time Collision Detection, 5.1.5), returned as the squared distance
  float ap[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
  const float d1 = ab[0] * ap[0] + ab[1] * ap[1] + ab[2] * ap[2], d2 = ac[0] * ap[0] + ac[1] * ap[1] + ac[2] * ap[2];
  float v = 0.f, w = 0.f;                                    // closest = a + v ab + w ac
  const float abab = ab[0] * ab[0] + ab[1] * ab[1] + ab[2] * ab[2], acac = ac[0] * ac[0] + ac[1] * ac[1] + ac[2] * ac[2];
  const float abac = ab[0] * ac[0] + ab[1] * ac[1] + ab[2] * ac[2];
  const float d3 = d1 - abab, d4 = d2 - abac;                // ab.(p-b), ac.(p-b)
  const float d5 = d1 - abac, d6 = d2 - acac;                // ab.(p-c), ac.(p-c)
  if (d1 <= 0.f && d2 <= 0.f) { v = 0.f; w = 0.f; }
  else if (d3 >= 0.f && d4 <= d3) { v = 1.f; w = 0.f; }
  else {
    const float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { v = d1 / (d1 - d3); w = 0.f; }
    else if (d6 >= 0.f && d5 <= d6) { v = 0.f; w = 1.f; }
    else {
      const float vb = d5 * d2 - d1 * d6;
      if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { v = 0.f; w = d2 / (d2 - d6); }
      else {
        const float va = d3 * d6 - d5 * d4;
        if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
          w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
          v = 1.f - w;
        } else {
          const float denom = 1.f / (va + vb + vc);
          v = vb * denom;
          w = vc * denom;
        }
      }
    }
  }
  const float dx = ap[0] - v * ab[0] - w * ac[0], dy = ap[1] - v * ab[1] - w * ac[1], dz = ap[2] - v * ab[2] - w * ac[2];
  return dx * dx + dy * dy + dz * dz;
}

__global__ void __launch_bounds__(256) point_mesh_kernel(const float* __restrict__ points, const float* __restrict__ vertices,
                                                         const int32_t* __restrict__ faces, int F, int M, int V, int NF,
                                                         float* __restrict__ dist) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* vs = reinterpret_cast<float*>(smem_raw);                       // [3V]
  unsigned* best = reinterpret_cast<unsigned*>(vs + ((3 * V + 3) & ~3)); // [M] bits of the squared distance
  float* pm = reinterpret_cast<float*>(best + M);                        // [3M]
  const int f = blockIdx.x, tid = threadIdx.x;
  for (int i = tid; i < 3 * V; i += blockDim.x) vs[i] = vertices[(size_t)f * 3 * V + i];
  for (int i = tid; i < M; i += blockDim.x) best[i] = 0x7f800000u;
  for (int i = tid; i < 3 * M; i += blockDim.x) pm[i] = points[(size_t)f * 3 * M + i];
  __syncthreads();
  for (int m0 = 0; m0 < M; m0 += PM_MG) {
    float b[PM_MG];
#pragma unroll
    for (int i = 0; i < PM_MG; ++i) b[i] = __int_as_float(0x7f800000);
    for (int t = tid; t < NF; t += blockDim.x) {
      const int i0 = __ldg(faces + 3 * t), i1 = __ldg(faces + 3 * t + 1), i2 = __ldg(faces + 3 * t + 2);
      const float a[3] = {vs[3 * i0], vs[3 * i0 + 1], vs[3 * i0 + 2]};
      const float ab[3] = {vs[3 * i1] - a[0], vs[3 * i1 + 1] - a[1], vs[3 * i1 + 2] - a[2]};
      const float ac[3] = {vs[3 * i2] - a[0], vs[3 * i2 + 1] - a[1], vs[3 * i2 + 2] - a[2]};
#pragma unroll
      for (int i = 0; i < PM_MG; ++i) {
        const int m = min(m0 + i, M - 1);
        const float p[3] = {pm[3 * m], pm[3 * m + 1], pm[3 * m + 2]};
        b[i] = fminf(b[i], point_tri_dist2(p, a, ab, ac));
      }
    }
#pragma unroll
    for (int i = 0; i < PM_MG; ++i)
      if (m0 + i < M) atomicMin(best + m0 + i, __float_as_uint(fmaxf(b[i], 0.f)));
  }
  __syncthreads();
  for (int i = tid; i < M; i += blockDim.x) dist[(size_t)f * M + i] = sqrtf(__uint_as_float(best[i]));
}

// ---- entry points ----
extern "C" int uuo_point_error_mean(const float* a, const float* b, int F, int P, const int32_t* sel, int n_sel,
                                    int velocity, float freq, float* scratch, float* out, void* stream) {
  UUO_CHECK_ARG(a && b && scratch && out, "NULL pointer");
  UUO_CHECK_ARG(F > (velocity ? 1 : 0) && P > 0, "bad sizes");
  if (!sel) n_sel = P;
  UUO_CHECK_ARG(n_sel > 0, "empty selection");
  const size_t items = (size_t)(F - (velocity ? 1 : 0)) * n_sel;
  int blocks = (int)((items + 255) / 256);
  if (blocks > 1024) blocks = 1024;                      // scratch holds 1024 partial sums
  point_error_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a, b, F, P, sel, n_sel, velocity ? 1 : 0, freq, scratch);
  finish_mean_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(scratch, blocks, (double)items, out);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_similarity_transform(const float* S1, const float* S2, int F, int N, float* out, void* stream) {
  UUO_CHECK_ARG(S1 && S2 && out && F > 0 && N > 0, "bad arguments");
  similarity_transform_kernel<<<(F + 127) / 128, 128, 0, (cudaStream_t)stream>>>(S1, S2, F, N, out);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_point_mesh_distance(const float* points, const float* vertices, const int32_t* faces, int F, int M, int V,
                                       int NF, float* dist, void* stream) {
  UUO_CHECK_ARG(points && vertices && faces && dist, "NULL pointer");
  UUO_CHECK_ARG(F >= 0 && M >= 0 && V > 0 && NF > 0, "bad sizes");
  if (F == 0 || M == 0) return UUO_OK;
  const size_t smem = (size_t)((3 * V + 3) & ~3) * 4 + (size_t)M * 4 + (size_t)3 * M * 4;
  UUO_CHECK_ARG(smem <= 227 * 1024, "V / M too large for the shared-memory frame");
  UUO_DYN_SMEM(point_mesh_kernel, smem);
  ProfScope ps_("point_mesh_distance", (cudaStream_t)stream);
  point_mesh_kernel<<<F, 256, smem, (cudaStream_t)stream>>>(points, vertices, faces, F, M, V, NF, dist);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}
