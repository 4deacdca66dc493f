// Linear blend skinning, forward and backward (lbs.py:237-250), dense J = 24 weights.
//   T[f,v] = sum_j W[v,j] A[f,j]          (3x4)
//   out    = T[:, :3] v_posed + T[:, 3] + transl
// Forward / d(v_posed) kernels: a thread owns VPT vertices (their 24 weights live in registers) and
// walks the frames of the tile; the frame's 24 relative transforms are broadcast from shared memory
// as float4.  d(A) kernel: a thread owns 4 joints x 12 entries of one frame and walks the vertices of
// its split; g and v_posed tiles are staged [vertex][frame] in shared memory.
#include "common.cuh"

#include <stdlib.h>
// UUO_SKIN_SIMT=1 selects the exact-fp32 SIMT kernels (debug cross-check of the tensor-core path, tools/gemm_check.py)
// (only in a build with -DUUO_DEBUG_SIMT; the release library has one skinning path, the SIMT kernels are compiled out)
#ifdef UUO_DEBUG_SIMT
static bool skin_simt() {
  const char* e = getenv("UUO_SKIN_SIMT");
  return e && e[0] == '1';
}
#else
static constexpr bool skin_simt() { return false; }
#endif

// number of d(A) partial slabs launch_skin_backward writes to ws.gA
int uuo_ga_parts(const uuo_model* m, bool with_g_joints) {
  if (skin_simt()) return UUO_NSPLIT_GA;
  return 2 * uuo_bwdA_splits(m) + ((with_g_joints && m->NE > 0) ? 1 : 0);
}

#define SK_THREADS 128
#define SK_VPT 2
#define SK_FB 16

#ifdef UUO_DEBUG_SIMT
__global__ void __launch_bounds__(SK_THREADS) skin_forward_kernel(
    const float* __restrict__ W, const float* __restrict__ A, const float* __restrict__ vposed, int NP,
    const float* __restrict__ transl, int F, int V, float* __restrict__ vertices) {
  __shared__ __align__(16) float4 As[SK_FB][UUO_SKIN_J * 3];
  const int tid = threadIdx.x;
  const int f0 = blockIdx.y * SK_FB;
  const int nf = min(SK_FB, F - f0);
  {
    const float4* src = reinterpret_cast<const float4*>(A + (size_t)f0 * UUO_SKIN_J * 12);
    float4* dst = &As[0][0];
    for (int i = tid; i < nf * UUO_SKIN_J * 3; i += SK_THREADS) dst[i] = src[i];
  }
  int vid[SK_VPT];
  float w[SK_VPT][UUO_SKIN_J];
#pragma unroll
  for (int q = 0; q < SK_VPT; ++q) {
    vid[q] = blockIdx.x * (SK_THREADS * SK_VPT) + q * SK_THREADS + tid;
    const int vv = min(vid[q], V - 1);
#pragma unroll
    for (int j4 = 0; j4 < UUO_SKIN_J / 4; ++j4) {
      float4 t = *reinterpret_cast<const float4*>(W + (size_t)vv * UUO_SKIN_J + j4 * 4);
      w[q][j4 * 4] = t.x; w[q][j4 * 4 + 1] = t.y; w[q][j4 * 4 + 2] = t.z; w[q][j4 * 4 + 3] = t.w;
    }
  }
  __syncthreads();
  for (int fi = 0; fi < nf; ++fi) {
    const int f = f0 + fi;
    float vp[SK_VPT][3];
#pragma unroll
    for (int q = 0; q < SK_VPT; ++q) {
      const int vv = min(vid[q], V - 1);
      const float* p = vposed + (size_t)f * NP + vv * 3;
      vp[q][0] = p[0]; vp[q][1] = p[1]; vp[q][2] = p[2];
    }
    float T[SK_VPT][12];
#pragma unroll
    for (int q = 0; q < SK_VPT; ++q)
#pragma unroll
      for (int i = 0; i < 12; ++i) T[q][i] = 0.f;
#pragma unroll
    for (int j = 0; j < UUO_SKIN_J; ++j) {
      float4 r0 = As[fi][j * 3], r1 = As[fi][j * 3 + 1], r2 = As[fi][j * 3 + 2];
#pragma unroll
      for (int q = 0; q < SK_VPT; ++q) {
        const float wj = w[q][j];
        T[q][0] = fmaf(wj, r0.x, T[q][0]); T[q][1] = fmaf(wj, r0.y, T[q][1]);
        T[q][2] = fmaf(wj, r0.z, T[q][2]); T[q][3] = fmaf(wj, r0.w, T[q][3]);
        T[q][4] = fmaf(wj, r1.x, T[q][4]); T[q][5] = fmaf(wj, r1.y, T[q][5]);
        T[q][6] = fmaf(wj, r1.z, T[q][6]); T[q][7] = fmaf(wj, r1.w, T[q][7]);
        T[q][8] = fmaf(wj, r2.x, T[q][8]); T[q][9] = fmaf(wj, r2.y, T[q][9]);
        T[q][10] = fmaf(wj, r2.z, T[q][10]); T[q][11] = fmaf(wj, r2.w, T[q][11]);
      }
    }
    float tr[3] = {0.f, 0.f, 0.f};
    if (transl) { tr[0] = transl[(size_t)f * 3]; tr[1] = transl[(size_t)f * 3 + 1]; tr[2] = transl[(size_t)f * 3 + 2]; }
#pragma unroll
    for (int q = 0; q < SK_VPT; ++q) {
      if (vid[q] < V) {
        float* o = vertices + ((size_t)f * V + vid[q]) * 3;
#pragma unroll
        for (int r = 0; r < 3; ++r)
          o[r] = (T[q][r * 4] * vp[q][0] + T[q][r * 4 + 1] * vp[q][1] + T[q][r * 4 + 2] * vp[q][2] + T[q][r * 4 + 3]) + tr[r];
      }
    }
  }
}
#endif  // UUO_DEBUG_SIMT

// joints[f, J+e] = vertices[f, extra_ids[e]]   (vertex_joint_selector.py:73-80; transl already inside)
__global__ void extra_joints_kernel(const float* __restrict__ vertices, const int32_t* __restrict__ extra_ids,
                                    int F, int V, int J, int NE, float* __restrict__ joints) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= F * NE) return;
  int f = i / NE, e = i % NE;
  const float* src = vertices + ((size_t)f * V + extra_ids[e]) * 3;
  float* dst = joints + ((size_t)f * (J + NE) + J + e) * 3;
  dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
}

__device__ __forceinline__ void load_g(const float* __restrict__ g_vertices, const float* __restrict__ g_joints,
                                       const int32_t* __restrict__ extra_ids, const uint8_t* __restrict__ vflag,
                                       int f, int v, int V, int J, int NE, float* g) {
  if (g_vertices) {
    const float* p = g_vertices + ((size_t)f * V + v) * 3;
    g[0] = p[0]; g[1] = p[1]; g[2] = p[2];
  } else {
    g[0] = g[1] = g[2] = 0.f;
  }
  if (g_joints && vflag[v]) {
    for (int e = 0; e < NE; ++e)
      if (extra_ids[e] == v) {
        const float* q = g_joints + ((size_t)f * (J + NE) + J + e) * 3;
        g[0] += q[0]; g[1] += q[1]; g[2] += q[2];
      }
  }
}

// g_vposed[f,v] = T[:3,:3]^T g[f,v]
#ifdef UUO_DEBUG_SIMT
__global__ void __launch_bounds__(SK_THREADS) skin_backward_vposed_kernel(
    const float* __restrict__ W, const float* __restrict__ A, const float* __restrict__ g_vertices,
    const float* __restrict__ g_joints, const int32_t* __restrict__ extra_ids, const uint8_t* __restrict__ vflag,
    int F, int V, int NE, int NP, float* __restrict__ gvposed) {
  __shared__ __align__(16) float4 As[SK_FB][UUO_SKIN_J * 3];
  const int tid = threadIdx.x;
  const int f0 = blockIdx.y * SK_FB;
  const int nf = min(SK_FB, F - f0);
  {
    const float4* src = reinterpret_cast<const float4*>(A + (size_t)f0 * UUO_SKIN_J * 12);
    float4* dst = &As[0][0];
    for (int i = tid; i < nf * UUO_SKIN_J * 3; i += SK_THREADS) dst[i] = src[i];
  }
  int vid[SK_VPT];
  float w[SK_VPT][UUO_SKIN_J];
#pragma unroll
  for (int q = 0; q < SK_VPT; ++q) {
    vid[q] = blockIdx.x * (SK_THREADS * SK_VPT) + q * SK_THREADS + tid;
    const int vv = min(vid[q], V - 1);
#pragma unroll
    for (int j4 = 0; j4 < UUO_SKIN_J / 4; ++j4) {
      float4 t = *reinterpret_cast<const float4*>(W + (size_t)vv * UUO_SKIN_J + j4 * 4);
      w[q][j4 * 4] = t.x; w[q][j4 * 4 + 1] = t.y; w[q][j4 * 4 + 2] = t.z; w[q][j4 * 4 + 3] = t.w;
    }
  }
  __syncthreads();
  for (int fi = 0; fi < nf; ++fi) {
    const int f = f0 + fi;
    float g[SK_VPT][3];
#pragma unroll
    for (int q = 0; q < SK_VPT; ++q) {
      if (vid[q] < V) load_g(g_vertices, g_joints, extra_ids, vflag, f, vid[q], V, UUO_SKIN_J, NE, g[q]);
      else g[q][0] = g[q][1] = g[q][2] = 0.f;
    }
    float T[SK_VPT][9];
#pragma unroll
    for (int q = 0; q < SK_VPT; ++q)
#pragma unroll
      for (int i = 0; i < 9; ++i) T[q][i] = 0.f;
#pragma unroll
    for (int j = 0; j < UUO_SKIN_J; ++j) {
      float4 r0 = As[fi][j * 3], r1 = As[fi][j * 3 + 1], r2 = As[fi][j * 3 + 2];
#pragma unroll
      for (int q = 0; q < SK_VPT; ++q) {
        const float wj = w[q][j];
        T[q][0] = fmaf(wj, r0.x, T[q][0]); T[q][1] = fmaf(wj, r0.y, T[q][1]); T[q][2] = fmaf(wj, r0.z, T[q][2]);
        T[q][3] = fmaf(wj, r1.x, T[q][3]); T[q][4] = fmaf(wj, r1.y, T[q][4]); T[q][5] = fmaf(wj, r1.z, T[q][5]);
        T[q][6] = fmaf(wj, r2.x, T[q][6]); T[q][7] = fmaf(wj, r2.y, T[q][7]); T[q][8] = fmaf(wj, r2.z, T[q][8]);
      }
    }
#pragma unroll
    for (int q = 0; q < SK_VPT; ++q) {
      // columns [3V, NP) of the row are written as zeros by the threads just past V (the GEMM reads them)
      float* o = gvposed + (size_t)f * NP + vid[q] * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (vid[q] * 3 + c < NP) o[c] = T[q][c] * g[q][0] + T[q][3 + c] * g[q][1] + T[q][6 + c] * g[q][2];
    }
  }
}

// dA[f,j,a,b] = sum_v W[v,j] g[f,v,a] * [v_posed[f,v]; 1][b]  -> deterministic partial per vertex split
#define GA_FB 32      // frames per block (one lane per frame)
#define GA_VC 32      // vertices per staged chunk
#define GA_THREADS 192  // 6 joint groups of 4 joints x 32 frames

__global__ void __launch_bounds__(GA_THREADS) skin_backward_A_kernel(
    const float* __restrict__ W, const float* __restrict__ vposed, const float* __restrict__ g_vertices,
    const float* __restrict__ g_joints, const int32_t* __restrict__ extra_ids, int F, int V, int NE, int NP,
    int v_per_split, float* __restrict__ gA_parts) {
  __shared__ __align__(16) float4 gs[GA_VC][GA_FB + 1];
  __shared__ __align__(16) float4 vs[GA_VC][GA_FB + 1];
  __shared__ __align__(16) float Ws[GA_VC][UUO_SKIN_J];
  const int tid = threadIdx.x;
  const int lane = tid & 31, jg = tid >> 5;   // jg = joint group (0..5), lane = frame within the tile
  const int f0 = blockIdx.x * GA_FB;
  const int split = blockIdx.y;
  const int v_begin = split * v_per_split;
  const int v_end = min(V, v_begin + v_per_split);
  const int J = UUO_SKIN_J;

  float acc[4][12];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[j][i] = 0.f;
  float gsum[3] = {0.f, 0.f, 0.f};

  for (int vc = v_begin; vc < v_end; vc += GA_VC) {
    const int nv = min(GA_VC, v_end - vc);
    __syncthreads();
    // stage: 32 frames x nv vertices x 3 floats of g and v_posed; weights of the chunk
    for (int i = tid; i < GA_FB * GA_VC * 3; i += GA_THREADS) {
      const int fi = i / (GA_VC * 3), r = i % (GA_VC * 3);
      const int vl = r / 3, k = r % 3;
      const int f = f0 + fi;
      float gv = 0.f, pv = 0.f;
      if (f < F && vl < nv) {
        if (g_vertices) gv = g_vertices[((size_t)f * V + vc) * 3 + r];
        pv = vposed[(size_t)f * NP + (size_t)vc * 3 + r];
      }
      reinterpret_cast<float*>(&gs[vl][fi])[k] = gv;
      reinterpret_cast<float*>(&vs[vl][fi])[k] = pv;
    }
    for (int i = tid; i < GA_VC * J; i += GA_THREADS) {
      const int vl = i / J;
      Ws[vl][i % J] = (vl < nv) ? W[(size_t)(vc + vl) * J + (i % J)] : 0.f;
    }
    __syncthreads();
    if (g_joints && tid < GA_FB && f0 + tid < F) {
      for (int e = 0; e < NE; ++e) {
        const int ev = extra_ids[e];
        if (ev >= vc && ev < vc + nv) {
          const float* q = g_joints + ((size_t)(f0 + tid) * (J + NE) + J + e) * 3;
          float4 t = gs[ev - vc][tid];
          t.x += q[0]; t.y += q[1]; t.z += q[2];
          gs[ev - vc][tid] = t;
        }
      }
    }
    __syncthreads();
    for (int vl = 0; vl < nv; ++vl) {
      const float4 g = gs[vl][lane];
      const float4 p = vs[vl][lane];
      const float4 w4 = *reinterpret_cast<const float4*>(&Ws[vl][jg * 4]);
      const float wj[4] = {w4.x, w4.y, w4.z, w4.w};
      if (jg == 0) { gsum[0] += g.x; gsum[1] += g.y; gsum[2] += g.z; }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a0 = wj[j] * g.x, a1 = wj[j] * g.y, a2 = wj[j] * g.z;
        acc[j][0] = fmaf(a0, p.x, acc[j][0]); acc[j][1] = fmaf(a0, p.y, acc[j][1]);
        acc[j][2] = fmaf(a0, p.z, acc[j][2]); acc[j][3] += a0;
        acc[j][4] = fmaf(a1, p.x, acc[j][4]); acc[j][5] = fmaf(a1, p.y, acc[j][5]);
        acc[j][6] = fmaf(a1, p.z, acc[j][6]); acc[j][7] += a1;
        acc[j][8] = fmaf(a2, p.x, acc[j][8]); acc[j][9] = fmaf(a2, p.y, acc[j][9]);
        acc[j][10] = fmaf(a2, p.z, acc[j][10]); acc[j][11] += a2;
      }
    }
  }
  const int f = f0 + lane;
  if (f < F) {
    float* dst = gA_parts + ((size_t)split * F + f) * UUO_GA_STRIDE;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float4* d4 = reinterpret_cast<float4*>(dst + (jg * 4 + j) * 12);
      d4[0] = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
      d4[1] = make_float4(acc[j][4], acc[j][5], acc[j][6], acc[j][7]);
      d4[2] = make_float4(acc[j][8], acc[j][9], acc[j][10], acc[j][11]);
    }
    if (jg == 0) *reinterpret_cast<float4*>(dst + 288) = make_float4(gsum[0], gsum[1], gsum[2], 0.f);
  }
}
#endif  // UUO_DEBUG_SIMT

// ---- extra-joint rows of g_joints (vertex_joint_selector.py:73-80): joints[J+e] = vertices[extra_ids[e]], so their
// upstream gradient lands on NU <= NE distinct vertices.  The tensor-core kernels see g_vertices only; these two tiny
// kernels add the remaining rank-NU terms: g_vposed[f, v_u] += R(f,v_u)^T gj_u and one more dA slab.
__global__ void extras_vposed_kernel(const float* __restrict__ W, const float* __restrict__ A,
                                     const float* __restrict__ g_joints, const int32_t* __restrict__ xu_vert,
                                     const int32_t* __restrict__ xu_start, const int32_t* __restrict__ xu_list, int F,
                                     int NU, int J, int NE, int NP, float* __restrict__ gvposed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= F * NU) return;
  const int f = i / NU, u = i % NU, v = xu_vert[u];
  float gj[3] = {0.f, 0.f, 0.f};
  for (int q = xu_start[u]; q < xu_start[u + 1]; ++q) {
    const float* p = g_joints + ((size_t)f * (J + NE) + J + xu_list[q]) * 3;
    gj[0] += p[0]; gj[1] += p[1]; gj[2] += p[2];
  }
  float R[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int j = 0; j < J; ++j) {
    const float w = W[(size_t)v * J + j];
    const float* a = A + ((size_t)f * J + j) * 12;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) R[r * 3 + c] = fmaf(w, a[r * 4 + c], R[r * 3 + c]);
  }
  float* o = gvposed + (size_t)f * NP + (size_t)v * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) o[c] += R[c] * gj[0] + R[3 + c] * gj[1] + R[6 + c] * gj[2];
}

__global__ void extras_dA_kernel(const float* __restrict__ W, const float* __restrict__ vposed,
                                 const float* __restrict__ g_joints, const int32_t* __restrict__ xu_vert,
                                 const int32_t* __restrict__ xu_start, const int32_t* __restrict__ xu_list, int F, int NU,
                                 int J, int NE, int NP, float* __restrict__ slab) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= F * (J + 1)) return;
  const int f = i / (J + 1), j = i % (J + 1);
  float acc[12] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int u = 0; u < NU; ++u) {
    const int v = xu_vert[u];
    float gj[3] = {0.f, 0.f, 0.f};
    for (int q = xu_start[u]; q < xu_start[u + 1]; ++q) {
      const float* p = g_joints + ((size_t)f * (J + NE) + J + xu_list[q]) * 3;
      gj[0] += p[0]; gj[1] += p[1]; gj[2] += p[2];
    }
    const float w = j < J ? W[(size_t)v * J + j] : 1.f;
    const float* vp = vposed + (size_t)f * NP + (size_t)v * 3;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float a = w * gj[r];
      acc[r * 4] = fmaf(a, vp[0], acc[r * 4]); acc[r * 4 + 1] = fmaf(a, vp[1], acc[r * 4 + 1]);
      acc[r * 4 + 2] = fmaf(a, vp[2], acc[r * 4 + 2]); acc[r * 4 + 3] += a;
    }
  }
  float* dst = slab + (size_t)f * UUO_GA_STRIDE;
  if (j < J) {
#pragma unroll
    for (int e = 0; e < 12; ++e) dst[j * 12 + e] = acc[e];
  } else {
    dst[288] = acc[3]; dst[289] = acc[7]; dst[290] = acc[11]; dst[291] = 0.f;
  }
}

int launch_skin_forward(const uuo_model* m, const float* transl, int F, float* vertices, float* joints,
                        const uuo_ws& ws, cudaStream_t st) {
  if (!skin_simt()) {
    int rc;
    { ProfScope ps_("pack_transforms", st); if ((rc = launch_pack_transforms(m, F, ws, st))) return rc; }
    { ProfScope ps_("skin_forward", st); if ((rc = launch_skin_tc(m, 0, F, ws.vposed, nullptr, transl, vertices, ws, st))) return rc; }
  }
#ifdef UUO_DEBUG_SIMT
  else {
  const int vpb = SK_THREADS * SK_VPT;
  dim3 grid((m->V + vpb - 1) / vpb, (F + SK_FB - 1) / SK_FB);
  { ProfScope ps_("skin_forward", st);
  skin_forward_kernel<<<grid, SK_THREADS, 0, st>>>(m->W, ws.A, ws.vposed, m->NP, transl, F, m->V, vertices);
  } UUO_LAUNCH_CHECK();
  }
#endif
  if (joints && m->NE > 0) {
    int n = F * m->NE;
    extra_joints_kernel<<<(n + 255) / 256, 256, 0, st>>>(vertices, m->extra_ids, F, m->V, m->J, m->NE, joints);
    UUO_LAUNCH_CHECK();
  }
  return UUO_OK;
}

int launch_skin_backward(const uuo_model* m, int F, const float* g_vertices, const float* g_joints,
                         const uuo_ws& ws, cudaStream_t st) {
  if (!skin_simt()) {
    int rc;
    ProfScope ps_("skin_backward_vposed", st);
    if ((rc = launch_skin_tc(m, 1, F, g_vertices, g_joints, nullptr, ws.gvposed, ws, st))) return rc;
    if (g_joints && m->NE > 0) {
      const int n = F * m->NU;
      extras_vposed_kernel<<<(n + 127) / 128, 128, 0, st>>>(m->W, ws.A, g_joints, m->xu_vert, m->xu_start, m->xu_list, F,
                                                            m->NU, m->J, m->NE, m->NP, ws.gvposed);
      UUO_LAUNCH_CHECK();
    }
  }
#ifdef UUO_DEBUG_SIMT
  else { ProfScope ps_("skin_backward_vposed", st);
  const int vpb = SK_THREADS * SK_VPT;
  // cover every column of the padded row so the pad is written as zeros
  dim3 grid1((m->NP / 3 + 1 + vpb - 1) / vpb, (F + SK_FB - 1) / SK_FB);
  skin_backward_vposed_kernel<<<grid1, SK_THREADS, 0, st>>>(m->W, ws.A, g_vertices, g_joints, m->extra_ids, m->vflag,
                                                            F, m->V, m->NE, m->NP, ws.gvposed);
  }
#endif
  UUO_LAUNCH_CHECK();
  if (!skin_simt()) {
    ProfScope ps_("skin_backward_A", st);
    int rc = launch_skin_bwdA_tc(m, F, g_vertices, g_joints, ws, st);
    if (rc) return rc;
    if (g_joints && m->NE > 0) {
      const int n = F * (m->J + 1);
      float* slab = ws.gA + (size_t)2 * uuo_bwdA_splits(m) * F * UUO_GA_STRIDE;
      extras_dA_kernel<<<(n + 127) / 128, 128, 0, st>>>(m->W, ws.vposed, g_joints, m->xu_vert, m->xu_start, m->xu_list, F,
                                                        m->NU, m->J, m->NE, m->NP, slab);
      UUO_LAUNCH_CHECK();
    }
    return UUO_OK;
  }
#ifdef UUO_DEBUG_SIMT
  const int vps = ((m->V + UUO_NSPLIT_GA - 1) / UUO_NSPLIT_GA + GA_VC - 1) / GA_VC * GA_VC;
  dim3 grid2((F + GA_FB - 1) / GA_FB, UUO_NSPLIT_GA);
  { ProfScope ps_("skin_backward_A", st);
  skin_backward_A_kernel<<<grid2, GA_THREADS, 0, st>>>(m->W, ws.vposed, g_vertices, g_joints, m->extra_ids, F, m->V,
                                                       m->NE, m->NP, vps, ws.gA);
  } UUO_LAUNCH_CHECK();
#endif
  return UUO_OK;
}
