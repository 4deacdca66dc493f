// Row-sparse LBS: forward of R selected vertices per frame and backward of a gradient that touches
// only R vertex rows per frame.  This is what the fitting closures need:
//   marker stage  : virtual markers = vertices[:, idx]   (optimization.py:345-351, a gather)
//   chamfer stage : d loss / d vertices is non-zero only at the argmin rows (pytorch3d knn backward)
// Nothing of size [F,V] is touched.  One warp per frame, frames independent, rows accumulated in order
// (deterministic).  The per-row arithmetic (blend-shape dot products of length K = P + NB, the 24-joint
// transform blend, the 3x4 apply) runs ONE ROW PER LANE -- no cross-lane reductions; the frame's coefficients
// and transforms are staged in shared memory and read as broadcasts.  Only the two reductions OVER rows of the
// backward (d coef, d A) run row by row with lanes across the coefficient / joint axis.
#include <algorithm>

#include "common.cuh"

#define FULL 0xffffffffu
#define ROWS_WARPS 4
#define ROWS_MAXKP 256      // KP <= 256
#define ROWS_CHUNK 64       // rows staged per pass of the backward

// rows[] layout: rows_shared == 0: [F][R] (per frame); == 1: [R] (all frames); == G > 1: [G][R], frame f belongs to
// group f / (F / G) (stacked hypotheses: every block of F / G frames has its own correspondences)
__device__ __forceinline__ size_t rows_base(int rows_shared, int F, int f, int R) {
  if (rows_shared == 0) return (size_t)f * R;
  if (rows_shared == 1) return 0;
  return (size_t)(f / (F / rows_shared)) * R;
}

struct __align__(16) RowRec {
  float vp[3];
  int v;
  float gvp[3];
  float pad0;
  float g[3];
  float pad1;
};

// v_posed of row v for this lane: vt + coef . Bt[3v+c]   (lbs.py:226-233 for one vertex)
__device__ __forceinline__ void lane_vposed(const float* __restrict__ vt, const float* __restrict__ Bt, int KP,
                                            const float4* __restrict__ s_coef4, int v, float* vp) {
  const float4* b0 = reinterpret_cast<const float4*>(Bt + (size_t)(3 * v) * KP);
  const float4* b1 = b0 + (KP >> 2);
  const float4* b2 = b1 + (KP >> 2);
  float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll 4
  for (int k = 0; k < (KP >> 2); ++k) {
    const float4 c = s_coef4[k];
    const float4 x = __ldg(b0 + k), y = __ldg(b1 + k), z = __ldg(b2 + k);
    a0 = fmaf(c.x, x.x, a0); a0 = fmaf(c.y, x.y, a0); a0 = fmaf(c.z, x.z, a0); a0 = fmaf(c.w, x.w, a0);
    a1 = fmaf(c.x, y.x, a1); a1 = fmaf(c.y, y.y, a1); a1 = fmaf(c.z, y.z, a1); a1 = fmaf(c.w, y.w, a1);
    a2 = fmaf(c.x, z.x, a2); a2 = fmaf(c.y, z.y, a2); a2 = fmaf(c.z, z.z, a2); a2 = fmaf(c.w, z.w, a2);
  }
  vp[0] = a0 + vt[3 * v];
  vp[1] = a1 + vt[3 * v + 1];
  vp[2] = a2 + vt[3 * v + 2];
}

// blended transform of row v for this lane: T[ab] = sum_j W[v][j] A[j][ab]   (lbs.py:237-243), NAB = 12 or 9 (rotation only)
template <bool WITH_T>
__device__ __forceinline__ void lane_blend(const float* __restrict__ W, int J, const float4* __restrict__ s_A4, int v,
                                           float* T /*[12]*/) {
#pragma unroll
  for (int i = 0; i < 12; ++i) T[i] = 0.f;
  const float4* w4 = reinterpret_cast<const float4*>(W + (size_t)v * J);   // J == 24: rows are 96 B, 16 B aligned
  for (int j4 = 0; j4 < (J >> 2); ++j4) {
    const float4 w = __ldg(w4 + j4);
    const float ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 r0 = s_A4[(j4 * 4 + i) * 3], r1 = s_A4[(j4 * 4 + i) * 3 + 1], r2 = s_A4[(j4 * 4 + i) * 3 + 2];
      T[0] = fmaf(ws[i], r0.x, T[0]); T[1] = fmaf(ws[i], r0.y, T[1]); T[2] = fmaf(ws[i], r0.z, T[2]);
      T[4] = fmaf(ws[i], r1.x, T[4]); T[5] = fmaf(ws[i], r1.y, T[5]); T[6] = fmaf(ws[i], r1.z, T[6]);
      T[8] = fmaf(ws[i], r2.x, T[8]); T[9] = fmaf(ws[i], r2.y, T[9]); T[10] = fmaf(ws[i], r2.z, T[10]);
      if (WITH_T) {
        T[3] = fmaf(ws[i], r0.w, T[3]); T[7] = fmaf(ws[i], r1.w, T[7]); T[11] = fmaf(ws[i], r2.w, T[11]);
      }
    }
  }
}

// stage the frame's coefficients [KP] and relative transforms [J][12] into this warp's shared memory
__device__ __forceinline__ void stage_frame(const float* __restrict__ coef, const float* __restrict__ A, int KP, int J,
                                            int f, int lane, float* s_coef, float* s_A) {
  const float4* c4 = reinterpret_cast<const float4*>(coef + (size_t)f * KP);
  for (int i = lane; i < (KP >> 2); i += 32) reinterpret_cast<float4*>(s_coef)[i] = __ldg(c4 + i);
  const float4* a4 = reinterpret_cast<const float4*>(A + (size_t)f * J * 12);
  for (int i = lane; i < J * 3; i += 32) reinterpret_cast<float4*>(s_A)[i] = __ldg(a4 + i);
  __syncwarp();
}

__global__ void __launch_bounds__(ROWS_WARPS * 32) rows_forward_kernel(
    const float* __restrict__ vt, const float* __restrict__ Bt, const float* __restrict__ W, int KP, int J,
    const float* __restrict__ coef, const float* __restrict__ A, const float* __restrict__ transl, int F,
    const int32_t* __restrict__ rows, int rows_shared, int R, float* __restrict__ out_rows,
    float* __restrict__ rowbuf /* [F][ROWS_CHUNK][4] v_posed of the rows, or NULL */) {
  __shared__ __align__(16) float s_coef_all[ROWS_WARPS][ROWS_MAXKP];
  __shared__ __align__(16) float s_A_all[ROWS_WARPS][UUO_SKIN_J * 12];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int f = blockIdx.x * ROWS_WARPS + w;
  if (f >= F) return;
  stage_frame(coef, A, KP, J, f, lane, s_coef_all[w], s_A_all[w]);
  const float4* s_coef4 = reinterpret_cast<const float4*>(s_coef_all[w]);
  const float4* s_A4 = reinterpret_cast<const float4*>(s_A_all[w]);
  float tr[3] = {0.f, 0.f, 0.f};
  if (transl) { tr[0] = transl[(size_t)f * 3]; tr[1] = transl[(size_t)f * 3 + 1]; tr[2] = transl[(size_t)f * 3 + 2]; }
  for (int r0 = 0; r0 < R; r0 += 32) {
    const int r = r0 + lane;
    if (r >= R) break;
    const int v = rows[rows_base(rows_shared, F, f, R) + r];
    float vp[3], T[12];
    lane_vposed(vt, Bt, KP, s_coef4, v, vp);
    if (rowbuf) reinterpret_cast<float4*>(rowbuf)[(size_t)f * ROWS_CHUNK + r] = make_float4(vp[0], vp[1], vp[2], 0.f);
    lane_blend<true>(W, J, s_A4, v, T);
    float* o = out_rows + ((size_t)f * R + r) * 3;
#pragma unroll
    for (int a = 0; a < 3; ++a)
      o[a] = fmaf(T[a * 4], vp[0], fmaf(T[a * 4 + 1], vp[1], fmaf(T[a * 4 + 2], vp[2], T[a * 4 + 3]))) + tr[a];
  }
}

// rowbuf[f][r] = v_posed rows of a dense forward (ws.vposed), picked by position
__global__ void cache_vposed_rows_kernel(const float* __restrict__ vposed, int NP, int F, const int32_t* __restrict__ pos,
                                         int R, float* __restrict__ rowbuf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= F * R) return;
  const int f = i / R, r = i - f * R;
  const float* p = vposed + (size_t)f * NP + 3 * pos[i];
  reinterpret_cast<float4*>(rowbuf)[(size_t)f * ROWS_CHUNK + r] = make_float4(p[0], p[1], p[2], 0.f);
}

template <bool CACHED>
__global__ void __launch_bounds__(ROWS_WARPS * 32, 9) rows_backward_kernel(
    const float* __restrict__ vt, const float* __restrict__ Bt, const float* __restrict__ W, int KP, int J,
    const float* __restrict__ coef, const float* __restrict__ A, int F, const int32_t* __restrict__ rows,
    int rows_shared, int R, const float* __restrict__ g_rows, const float* __restrict__ rowbuf,
    float* __restrict__ gA_part, float* __restrict__ gcoef_part) {
  __shared__ __align__(16) float s_coef_all[ROWS_WARPS][ROWS_MAXKP];
  __shared__ __align__(16) float s_A_all[ROWS_WARPS][UUO_SKIN_J * 12];
  __shared__ RowRec s_rec_all[ROWS_WARPS][ROWS_CHUNK];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int f = blockIdx.x * ROWS_WARPS + w;
  if (f >= F) return;
  stage_frame(coef, A, KP, J, f, lane, s_coef_all[w], s_A_all[w]);
  const float4* s_coef4 = reinterpret_cast<const float4*>(s_coef_all[w]);
  const float4* s_A4 = reinterpret_cast<const float4*>(s_A_all[w]);
  RowRec* rec = s_rec_all[w];
  // d coef: lane owns k = 4 lane .. 4 lane + 3 and 128 + 4 lane .. (KP <= 256): two 16 B loads per Bt row
  float4 gc0 = make_float4(0.f, 0.f, 0.f, 0.f), gc1 = gc0;
  const bool has1 = 128 + 4 * lane < KP;
  const int kp4 = KP >> 2;
  float ga[12];                // d A[j = lane][ab]
#pragma unroll
  for (int i = 0; i < 12; ++i) ga[i] = 0.f;
  float gs = 0.f;              // lane a < 3: sum of g over rows (d transl)

  for (int c0 = 0; c0 < R; c0 += ROWS_CHUNK) {
    const int nc = min(ROWS_CHUNK, R - c0);
    // ---- A. one row per lane: v_posed, rotation part of the blended transform, d v_posed = T^T g
    for (int r0 = 0; r0 < nc; r0 += 32) {
      const int r = r0 + lane;
      if (r < nc) {
        const int v = rows[rows_base(rows_shared, F, f, R) + c0 + r];
        const float* gp = g_rows + ((size_t)f * R + c0 + r) * 3;
        const float g0 = gp[0], g1 = gp[1], g2 = gp[2];
        RowRec o;
        o.v = v; o.g[0] = g0; o.g[1] = g1; o.g[2] = g2; o.pad0 = 0.f; o.pad1 = 0.f;
        if (g0 == 0.f && g1 == 0.f && g2 == 0.f) {        // masked / dropped markers contribute nothing
          o.v = -1;
          o.vp[0] = o.vp[1] = o.vp[2] = o.gvp[0] = o.gvp[1] = o.gvp[2] = 0.f;
        } else {
          float T[12];
          if (CACHED) {        // R <= ROWS_CHUNK: written by the forward of these very rows
            const float4 c = reinterpret_cast<const float4*>(rowbuf)[(size_t)f * ROWS_CHUNK + r];
            o.vp[0] = c.x; o.vp[1] = c.y; o.vp[2] = c.z;
          } else {
            lane_vposed(vt, Bt, KP, s_coef4, v, o.vp);
          }
          lane_blend<false>(W, J, s_A4, v, T);
#pragma unroll
          for (int c = 0; c < 3; ++c) o.gvp[c] = T[c] * g0 + T[4 + c] * g1 + T[8 + c] * g2;
        }
        rec[r] = o;
      }
    }
    __syncwarp();
    // ---- B. reductions over rows: lanes across coefficients (d coef) and joints (d A)
    for (int r = 0; r < nc; ++r) {
      const float4 q0 = reinterpret_cast<const float4*>(rec + r)[0];   // vp, v
      const int v = __float_as_int(q0.w);
      if (v < 0) continue;                                              // warp-uniform
      const float4 q1 = reinterpret_cast<const float4*>(rec + r)[1];   // gvp
      const float4 q2 = reinterpret_cast<const float4*>(rec + r)[2];   // g
      const float4* b4 = reinterpret_cast<const float4*>(Bt + (size_t)(3 * v) * KP) + lane;
      {
        const float4 x = __ldg(b4), y = __ldg(b4 + kp4), z = __ldg(b4 + 2 * kp4);
        gc0.x = fmaf(q1.z, z.x, fmaf(q1.y, y.x, fmaf(q1.x, x.x, gc0.x)));
        gc0.y = fmaf(q1.z, z.y, fmaf(q1.y, y.y, fmaf(q1.x, x.y, gc0.y)));
        gc0.z = fmaf(q1.z, z.z, fmaf(q1.y, y.z, fmaf(q1.x, x.z, gc0.z)));
        gc0.w = fmaf(q1.z, z.w, fmaf(q1.y, y.w, fmaf(q1.x, x.w, gc0.w)));
      }
      if (has1) {
        const float4 x = __ldg(b4 + 32), y = __ldg(b4 + kp4 + 32), z = __ldg(b4 + 2 * kp4 + 32);
        gc1.x = fmaf(q1.z, z.x, fmaf(q1.y, y.x, fmaf(q1.x, x.x, gc1.x)));
        gc1.y = fmaf(q1.z, z.y, fmaf(q1.y, y.y, fmaf(q1.x, x.y, gc1.y)));
        gc1.z = fmaf(q1.z, z.z, fmaf(q1.y, y.z, fmaf(q1.x, x.z, gc1.z)));
        gc1.w = fmaf(q1.z, z.w, fmaf(q1.y, y.w, fmaf(q1.x, x.w, gc1.w)));
      }
      if (lane < 3) gs += lane == 0 ? q2.x : (lane == 1 ? q2.y : q2.z);
      // dA[j][a][b] += W[v][j] g[a] vh[b],  vh = (v_posed, 1)
      const float wj = lane < J ? __ldg(W + (size_t)v * J + lane) : 0.f;
      const float gg[3] = {q2.x, q2.y, q2.z};
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const float wg = wj * gg[a];
        ga[a * 4] = fmaf(wg, q0.x, ga[a * 4]);
        ga[a * 4 + 1] = fmaf(wg, q0.y, ga[a * 4 + 1]);
        ga[a * 4 + 2] = fmaf(wg, q0.z, ga[a * 4 + 2]);
        ga[a * 4 + 3] += wg;
      }
    }
    __syncwarp();
  }
  float* dstA = gA_part + (size_t)f * UUO_GA_STRIDE;
  if (lane < J) {
    float4* d4 = reinterpret_cast<float4*>(dstA + lane * 12);
    d4[0] = make_float4(ga[0], ga[1], ga[2], ga[3]);
    d4[1] = make_float4(ga[4], ga[5], ga[6], ga[7]);
    d4[2] = make_float4(ga[8], ga[9], ga[10], ga[11]);
  }
  if (lane < 4) dstA[J * 12 + lane] = (lane < 3) ? gs : 0.f;
  float4* dstC = reinterpret_cast<float4*>(gcoef_part + (size_t)f * KP);
  if (4 * lane < KP) dstC[lane] = gc0;
  if (has1) dstC[32 + lane] = gc1;
}

static int rows_check(const uuo_model* m) {
  if (m->KP > ROWS_MAXKP || (m->KP & 31)) { uuo_set_error("rows path: KP must be a multiple of 32, <= 256"); return UUO_ERR_ARG; }
  if (m->J != UUO_SKIN_J) { uuo_set_error("rows path: J must be 24"); return UUO_ERR_ARG; }
  return UUO_OK;
}

int launch_rows_forward(const uuo_model* m, const float* transl, int F, const int32_t* rows, int rows_shared,
                        int R, float* out_rows, const uuo_ws& ws, cudaStream_t st) {
  int rc = rows_check(m);
  if (rc) return rc;
  float* rowbuf = R <= ROWS_CHUNK ? ws.rowbuf : nullptr;
  const int G = rows_shared > 1 ? rows_shared : 1;           // rows_shared = G > 1: G contiguous frame groups, rows[G][R]
  if (F % G) { uuo_set_error("rows path: F must be a multiple of the number of row groups"); return UUO_ERR_ARG; }
  ProfScope ps_("rows_forward", st);
  // rows shared by all frames (of a group): the blend-shape part is a GEMM on the tensor cores (rows_tc.cu)
  if (rows_tc_usable(m, F, rows_shared, R)) return launch_rows_forward_tc(m, transl, F, rows, rows_shared, R, out_rows, ws, st);
  rows_forward_kernel<<<(F + ROWS_WARPS - 1) / ROWS_WARPS, ROWS_WARPS * 32, 0, st>>>(
      m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, transl, F, rows, rows_shared, R, out_rows, rowbuf);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_cache_vposed_rows(const uuo_model* m, int F, const int32_t* pos, int R, const uuo_ws& ws, cudaStream_t st) {
  if (R > ROWS_CHUNK) { uuo_set_error("cache_vposed_rows: R must be <= 64"); return UUO_ERR_ARG; }
  cache_vposed_rows_kernel<<<(F * R + 255) / 256, 256, 0, st>>>(ws.vposed, m->NP, F, pos, R, ws.rowbuf);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_rows_backward(const uuo_model* m, int F, const int32_t* rows, int rows_shared, int R,
                         const float* g_rows, int vposed_cached, const uuo_ws& ws, cudaStream_t st) {
  int rc = rows_check(m);
  if (rc) return rc;
  if (vposed_cached && R > ROWS_CHUNK) { uuo_set_error("rows backward: cached v_posed needs R <= 64"); return UUO_ERR_ARG; }
  ProfScope ps_("rows_backward", st);
  if (vposed_cached && rows_tc_usable(m, F, rows_shared, R)) return launch_rows_backward_tc(m, F, rows, rows_shared, R, g_rows, ws, st);
  if (vposed_cached)
    rows_backward_kernel<true><<<(F + ROWS_WARPS - 1) / ROWS_WARPS, ROWS_WARPS * 32, 0, st>>>(
        m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, F, rows, rows_shared, R, g_rows, ws.rowbuf, ws.gA, ws.gcoef);
  else
    rows_backward_kernel<false><<<(F + ROWS_WARPS - 1) / ROWS_WARPS, ROWS_WARPS * 32, 0, st>>>(
        m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, F, rows, rows_shared, R, g_rows, ws.rowbuf, ws.gA, ws.gcoef);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}
