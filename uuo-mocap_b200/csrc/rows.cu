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

// ---- forward for rows SHARED by all frames (marker stage: the same M vertices in every frame) ----
// Persistent CTAs keep the 3R selected rows of Bt in shared memory (pitch KP+1: conflict-free when lanes
// walk different rows at the same k), so the blend-shape part is a small GEMM out of shared memory:
// lanes across the 3R outputs, the frame's coefficients broadcast.  Then one row per lane for the blend.
#define ROWSS_WARPS 16
#define ROWSS_NI 6            // 3R <= 192
#define ROWSS_NF 2            // frames per warp pass: every Bt value read from shared memory feeds two frames

struct rowss_layout {
  int pitch, n3;
  size_t off_vt, off_w, off_wj, off_wc, off_coef, off_A, off_vp, total;
};
__host__ __device__ inline rowss_layout rowss_make_layout(int KP, int R, int J) {
  rowss_layout L;
  L.pitch = KP + 1;
  L.n3 = 3 * R;
  L.off_vt = (size_t)L.n3 * L.pitch * 4;
  L.off_w = L.off_vt + (size_t)((L.n3 + 3) & ~3) * 4;
  L.off_wj = L.off_w + (size_t)R * J * 4;                       // non-zero weights of a row | their joints | how many
  L.off_wc = (L.off_wj + (size_t)R * J + 3) & ~(size_t)3;
  L.off_coef = (L.off_wc + (size_t)R * 4 + 15) & ~(size_t)15;
  L.off_A = L.off_coef + (size_t)ROWSS_WARPS * ROWSS_NF * ROWS_MAXKP * 4;
  L.off_vp = L.off_A + (size_t)ROWSS_WARPS * ROWSS_NF * UUO_SKIN_J * 12 * 4;
  L.total = L.off_vp + (size_t)ROWSS_WARPS * ROWSS_NF * 32 * ROWSS_NI * 4;
  return L;
}

__global__ void __launch_bounds__(ROWSS_WARPS * 32, 1) rows_forward_shared_kernel(
    const float* __restrict__ vt, const float* __restrict__ Bt, const float* __restrict__ W, int KP, int J,
    const float* __restrict__ coef, const float* __restrict__ A, const float* __restrict__ transl, int F,
    const int32_t* __restrict__ rows, int R, float* __restrict__ out_rows, float* __restrict__ rowbuf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const rowss_layout L = rowss_make_layout(KP, R, J);
  float* s_B = reinterpret_cast<float*>(smem_raw);                     // [3R][KP+1]
  float* s_vt = reinterpret_cast<float*>(smem_raw + L.off_vt);         // [3R]
  float* s_W = reinterpret_cast<float*>(smem_raw + L.off_w);           // [R][J] the row's non-zero weights (ascending joint:
  uint8_t* s_Wj = smem_raw + L.off_wj;                                 // [R][J]  the dense order; a zero weight adds exactly
  int* s_Wc = reinterpret_cast<int*>(smem_raw + L.off_wc);             // [R]     nothing, so skipping it changes no bit)
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  float* s_coef = reinterpret_cast<float*>(smem_raw + L.off_coef) + w * ROWSS_NF * ROWS_MAXKP;     // [NF][KP]
  float* s_A = reinterpret_cast<float*>(smem_raw + L.off_A) + w * ROWSS_NF * UUO_SKIN_J * 12;      // [NF][J][12]
  float* s_vp = reinterpret_cast<float*>(smem_raw + L.off_vp) + w * ROWSS_NF * 32 * ROWSS_NI;      // [NF][3R]
  const int n3 = L.n3, pitch = L.pitch;
  if (gridDim.y > 1) {       // grouped rows: group g = blockIdx.y owns frames [g F, (g+1) F) and rows[g][R]
    const size_t fb = (size_t)blockIdx.y * F;
    rows += (size_t)blockIdx.y * R;
    coef += fb * KP;
    A += fb * J * 12;
    if (transl) transl += fb * 3;
    out_rows += fb * R * 3;
    if (rowbuf) rowbuf += fb * ROWS_CHUNK * 4;
  }
  for (int i = tid; i < n3 * KP; i += ROWSS_WARPS * 32) {
    const int n = i / KP, k = i - n * KP;
    s_B[n * pitch + k] = __ldg(Bt + (size_t)(3 * rows[n / 3] + n % 3) * KP + k);
  }
  for (int i = tid; i < n3; i += ROWSS_WARPS * 32) s_vt[i] = vt[3 * rows[i / 3] + i % 3];
  if (tid < R) {
    const float* wr = W + (size_t)rows[tid] * J;
    int c = 0;
    for (int j = 0; j < J; ++j) {
      const float x = wr[j];
      if (x != 0.f) { s_W[tid * J + c] = x; s_Wj[tid * J + c] = (uint8_t)j; ++c; }
    }
    s_Wc[tid] = c;
  }
  __syncthreads();
  for (int f0 = (blockIdx.x * ROWSS_WARPS + w) * ROWSS_NF; f0 < F; f0 += gridDim.x * ROWSS_WARPS * ROWSS_NF) {
#pragma unroll
    for (int u = 0; u < ROWSS_NF; ++u)       // an odd tail repeats the last frame (computed twice, stored once)
      stage_frame(coef, A, KP, J, min(f0 + u, F - 1), lane, s_coef + u * ROWS_MAXKP, s_A + u * UUO_SKIN_J * 12);
    // v_posed of the 3R outputs: lane owns n = lane + 32 i
    float acc[ROWSS_NF][ROWSS_NI];
    const float* bp[ROWSS_NI];
#pragma unroll
    for (int i = 0; i < ROWSS_NI; ++i) {
#pragma unroll
      for (int u = 0; u < ROWSS_NF; ++u) acc[u][i] = 0.f;
      bp[i] = s_B + min(lane + 32 * i, n3 - 1) * pitch;
    }
    for (int k4 = 0; k4 < (KP >> 2); ++k4) {
      float4 c[ROWSS_NF];
#pragma unroll
      for (int u = 0; u < ROWSS_NF; ++u) c[u] = reinterpret_cast<const float4*>(s_coef + u * ROWS_MAXKP)[k4];
#pragma unroll
      for (int i = 0; i < ROWSS_NI; ++i) {
        if (32 * i < n3) {
          const float* b = bp[i] + 4 * k4;
          const float b0 = b[0], b1 = b[1], b2 = b[2], b3 = b[3];
#pragma unroll
          for (int u = 0; u < ROWSS_NF; ++u) {
            acc[u][i] = fmaf(c[u].x, b0, acc[u][i]);
            acc[u][i] = fmaf(c[u].y, b1, acc[u][i]);
            acc[u][i] = fmaf(c[u].z, b2, acc[u][i]);
            acc[u][i] = fmaf(c[u].w, b3, acc[u][i]);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < ROWSS_NI; ++i) {
      const int n = lane + 32 * i;
      if (n < n3) {
#pragma unroll
        for (int u = 0; u < ROWSS_NF; ++u) s_vp[u * 32 * ROWSS_NI + n] = acc[u][i] + s_vt[n];
      }
    }
    __syncwarp();
    for (int u = 0; u < ROWSS_NF; ++u) {
      const int f = f0 + u;
      if (f >= F) break;
      const float4* s_A4 = reinterpret_cast<const float4*>(s_A + u * UUO_SKIN_J * 12);
      const float* vpu = s_vp + u * 32 * ROWSS_NI;
      float tr[3] = {0.f, 0.f, 0.f};
      if (transl) { tr[0] = transl[(size_t)f * 3]; tr[1] = transl[(size_t)f * 3 + 1]; tr[2] = transl[(size_t)f * 3 + 2]; }
      for (int r = lane; r < R; r += 32) {
        const float vp0 = vpu[3 * r], vp1 = vpu[3 * r + 1], vp2 = vpu[3 * r + 2];
        if (rowbuf) reinterpret_cast<float4*>(rowbuf)[(size_t)f * ROWS_CHUNK + r] = make_float4(vp0, vp1, vp2, 0.f);
        float T[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) T[i] = 0.f;
        const int cnt = s_Wc[r];
        for (int i = 0; i < cnt; ++i) {
          const float wj = s_W[r * J + i];
          const int j = s_Wj[r * J + i];
          const float4 r0 = s_A4[j * 3], r1 = s_A4[j * 3 + 1], r2 = s_A4[j * 3 + 2];
          T[0] = fmaf(wj, r0.x, T[0]); T[1] = fmaf(wj, r0.y, T[1]); T[2] = fmaf(wj, r0.z, T[2]); T[3] = fmaf(wj, r0.w, T[3]);
          T[4] = fmaf(wj, r1.x, T[4]); T[5] = fmaf(wj, r1.y, T[5]); T[6] = fmaf(wj, r1.z, T[6]); T[7] = fmaf(wj, r1.w, T[7]);
          T[8] = fmaf(wj, r2.x, T[8]); T[9] = fmaf(wj, r2.y, T[9]); T[10] = fmaf(wj, r2.z, T[10]); T[11] = fmaf(wj, r2.w, T[11]);
        }
        float* o = out_rows + ((size_t)f * R + r) * 3;
#pragma unroll
        for (int a = 0; a < 3; ++a)
          o[a] = fmaf(T[a * 4], vp0, fmaf(T[a * 4 + 1], vp1, fmaf(T[a * 4 + 2], vp2, T[a * 4 + 3]))) + tr[a];
      }
    }
    __syncwarp();
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

// ---- backward for rows SHARED by all frames with their v_posed cached (marker stage) ----
// Same arithmetic, in the same order, as rows_backward_kernel<true> (bit-equal results), but the 3R selected rows of Bt
// stay in shared memory for the whole launch (persistent CTAs, like rows_forward_shared_kernel) instead of being read
// from L2 by every frame (121 KB per frame), and the skinning weights are walked through their non-zeros only: per row
// (ELL) for the blended rotation, per joint (CSC, rows ascending = the dense summation order; an exact zero weight adds
// exactly nothing) for d A.  A warp pass takes ROWSB_NF frames so every Bt value read feeds that many frames.
#define ROWSB_WARPS 16
#define ROWSB_NF 2
#define ROWSB_NK (ROWS_MAXKP / 32)

struct rowsb_layout {
  int pitch, n3, n3p;
  size_t off_ellw, off_ellj, off_ellc, off_cscw, off_cscr, off_cscc, off_warp, warp_floats, total;
};
__host__ __device__ inline rowsb_layout rowsb_make_layout(int KP, int R, int J) {
  rowsb_layout L;
  L.pitch = KP + 1;
  L.n3 = 3 * R;
  L.n3p = (L.n3 + 3) & ~3;
  size_t off = (size_t)L.n3 * L.pitch * 4;
  auto take = [&](size_t bytes) { const size_t o = off; off = (off + bytes + 15) & ~(size_t)15; return o; };
  L.off_ellw = take((size_t)R * J * 4);
  L.off_ellj = take((size_t)R * J);
  L.off_ellc = take((size_t)R * 4);
  L.off_cscw = take((size_t)J * R * 4);
  L.off_cscr = take((size_t)J * R);
  L.off_cscc = take((size_t)J * 4);
  L.off_warp = off;
  L.warp_floats = (size_t)ROWSB_NF * (UUO_SKIN_J * 12 + 3 * L.n3p);        // per frame: A | gvp | g | vp
  L.total = off + (size_t)ROWSB_WARPS * L.warp_floats * 4;
  return L;
}

__global__ void __launch_bounds__(ROWSB_WARPS * 32, 1) rows_backward_shared_kernel(
    const float* __restrict__ Bt, const float* __restrict__ W, int KP, int J, const float* __restrict__ A, int F,
    const int32_t* __restrict__ rows, int R, const float* __restrict__ g_rows, const float* __restrict__ rowbuf,
    float* __restrict__ gA_part, float* __restrict__ gcoef_part) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const rowsb_layout L = rowsb_make_layout(KP, R, J);
  float* s_B = reinterpret_cast<float*>(smem_raw);                               // [3R][KP+1]
  float* s_ellw = reinterpret_cast<float*>(smem_raw + L.off_ellw);               // [R][J] non-zero weights of a row
  uint8_t* s_ellj = smem_raw + L.off_ellj;                                       // [R][J] their joints (ascending)
  int* s_ellc = reinterpret_cast<int*>(smem_raw + L.off_ellc);                   // [R] how many
  float* s_cscw = reinterpret_cast<float*>(smem_raw + L.off_cscw);               // [J][R] non-zero weights of a joint
  uint8_t* s_cscr = smem_raw + L.off_cscr;                                       // [J][R] their rows (ascending)
  int* s_cscc = reinterpret_cast<int*>(smem_raw + L.off_cscc);                   // [J] how many
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int n3 = L.n3, n3p = L.n3p, pitch = L.pitch;
  float* s_warp = reinterpret_cast<float*>(smem_raw + L.off_warp) + (size_t)w * L.warp_floats;
  const int fr_floats = UUO_SKIN_J * 12 + 3 * n3p;
  if (gridDim.y > 1) {       // grouped rows: group g = blockIdx.y owns frames [g F, (g+1) F) and rows[g][R]
    const size_t fb = (size_t)blockIdx.y * F;
    rows += (size_t)blockIdx.y * R;
    A += fb * J * 12;
    g_rows += fb * R * 3;
    rowbuf += fb * ROWS_CHUNK * 4;
    gA_part += fb * UUO_GA_STRIDE;
    gcoef_part += fb * KP;
  }
  for (int i = tid; i < n3 * KP; i += ROWSB_WARPS * 32) {
    const int n = i / KP, k = i - n * KP;
    s_B[n * pitch + k] = __ldg(Bt + (size_t)(3 * rows[n / 3] + n % 3) * KP + k);
  }
  if (tid < R) {
    const float* wr = W + (size_t)rows[tid] * J;
    int c = 0;
    for (int j = 0; j < J; ++j) {
      const float x = wr[j];
      if (x != 0.f) { s_ellw[tid * J + c] = x; s_ellj[tid * J + c] = (uint8_t)j; ++c; }
    }
    s_ellc[tid] = c;
  } else if (tid >= 64 && tid < 64 + J) {
    const int j = tid - 64;
    int c = 0;
    for (int r = 0; r < R; ++r) {
      const float x = W[(size_t)rows[r] * J + j];
      if (x != 0.f) { s_cscw[j * R + c] = x; s_cscr[j * R + c] = (uint8_t)r; ++c; }
    }
    s_cscc[j] = c;
  }
  __syncthreads();

  for (int f0 = (blockIdx.x * ROWSB_WARPS + w) * ROWSB_NF; f0 < F; f0 += gridDim.x * ROWSB_WARPS * ROWSB_NF) {
    // ---- A. one row per lane: rotation part of the blended transform, d v_posed = T^T g
#pragma unroll
    for (int u = 0; u < ROWSB_NF; ++u) {        // an odd tail repeats the last frame (computed twice, stored once)
      const int f = min(f0 + u, F - 1);
      float* s_A = s_warp + u * fr_floats;
      float* s_gvp = s_A + UUO_SKIN_J * 12;
      float* s_g = s_gvp + n3p;
      float* s_vp = s_g + n3p;
      const float4* a4 = reinterpret_cast<const float4*>(A + (size_t)f * J * 12);
      for (int i = lane; i < J * 3; i += 32) reinterpret_cast<float4*>(s_A)[i] = __ldg(a4 + i);
      __syncwarp();
      const float4* s_A4 = reinterpret_cast<const float4*>(s_A);
      for (int r = lane; r < R; r += 32) {
        const float* gp = g_rows + ((size_t)f * R + r) * 3;
        const float g0 = gp[0], g1 = gp[1], g2 = gp[2];
        const float4 c = reinterpret_cast<const float4*>(rowbuf)[(size_t)f * ROWS_CHUNK + r];
        float T[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) T[i] = 0.f;
        const int cnt = s_ellc[r];
        for (int i = 0; i < cnt; ++i) {
          const float wj = s_ellw[r * J + i];
          const int j = s_ellj[r * J + i];
          const float4 r0 = s_A4[j * 3], r1 = s_A4[j * 3 + 1], r2 = s_A4[j * 3 + 2];
          T[0] = fmaf(wj, r0.x, T[0]); T[1] = fmaf(wj, r0.y, T[1]); T[2] = fmaf(wj, r0.z, T[2]);
          T[3] = fmaf(wj, r1.x, T[3]); T[4] = fmaf(wj, r1.y, T[4]); T[5] = fmaf(wj, r1.z, T[5]);
          T[6] = fmaf(wj, r2.x, T[6]); T[7] = fmaf(wj, r2.y, T[7]); T[8] = fmaf(wj, r2.z, T[8]);
        }
        const bool off = g0 == 0.f && g1 == 0.f && g2 == 0.f;      // masked / dropped markers contribute nothing
        s_g[3 * r] = g0; s_g[3 * r + 1] = g1; s_g[3 * r + 2] = g2;
        s_vp[3 * r] = off ? 0.f : c.x; s_vp[3 * r + 1] = off ? 0.f : c.y; s_vp[3 * r + 2] = off ? 0.f : c.z;
#pragma unroll
        for (int q = 0; q < 3; ++q) s_gvp[3 * r + q] = off ? 0.f : T[q] * g0 + T[3 + q] * g1 + T[6 + q] * g2;
      }
    }
    __syncwarp();
    // ---- B1. d coef[k] = sum_n gvp[n] Bt[n][k]: lane owns k = lane + 32 i
    {
      float acc[ROWSB_NF][ROWSB_NK];
#pragma unroll
      for (int u = 0; u < ROWSB_NF; ++u)
#pragma unroll
        for (int i = 0; i < ROWSB_NK; ++i) acc[u][i] = 0.f;
      const float* gv0 = s_warp + UUO_SKIN_J * 12;
#pragma unroll 4
      for (int n = 0; n < n3; ++n) {
        float b[ROWSB_NK];
        const float* bn = s_B + n * pitch + lane;
#pragma unroll
        for (int i = 0; i < ROWSB_NK; ++i) b[i] = (32 * i < KP) ? bn[32 * i] : 0.f;      // KP is a multiple of 32
#pragma unroll
        for (int u = 0; u < ROWSB_NF; ++u) {
          const float gv = gv0[u * fr_floats + n];
#pragma unroll
          for (int i = 0; i < ROWSB_NK; ++i)
            if (32 * i < KP) acc[u][i] = fmaf(gv, b[i], acc[u][i]);
        }
      }
#pragma unroll
      for (int u = 0; u < ROWSB_NF; ++u) {
        if (f0 + u >= F) break;
        float* dstC = gcoef_part + (size_t)(f0 + u) * KP + lane;
#pragma unroll
        for (int i = 0; i < ROWSB_NK; ++i)
          if (32 * i < KP) dstC[32 * i] = acc[u][i];
      }
    }
    // ---- B2. dA[j = lane][a][b] = sum_r W[r][j] g[r][a] (v_posed[r], 1)[b] over the rows where W[r][j] != 0
    {
      float ga[ROWSB_NF][12], gs[ROWSB_NF];
#pragma unroll
      for (int u = 0; u < ROWSB_NF; ++u) {
        gs[u] = 0.f;
#pragma unroll
        for (int i = 0; i < 12; ++i) ga[u][i] = 0.f;
      }
      const int cnt = lane < J ? s_cscc[lane] : 0;
      for (int i = 0; i < cnt; ++i) {
        const float wj = s_cscw[lane * R + i];
        const int r = s_cscr[lane * R + i];
#pragma unroll
        for (int u = 0; u < ROWSB_NF; ++u) {
          const float* sg = s_warp + u * fr_floats + UUO_SKIN_J * 12 + n3p + 3 * r;
          const float* sv = sg + n3p;
          const float v0 = sv[0], v1 = sv[1], v2 = sv[2];
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            const float wg = wj * sg[a];
            ga[u][a * 4] = fmaf(wg, v0, ga[u][a * 4]);
            ga[u][a * 4 + 1] = fmaf(wg, v1, ga[u][a * 4 + 1]);
            ga[u][a * 4 + 2] = fmaf(wg, v2, ga[u][a * 4 + 2]);
            ga[u][a * 4 + 3] += wg;
          }
        }
      }
      if (lane < 3) {
        for (int r = 0; r < R; ++r)
#pragma unroll
          for (int u = 0; u < ROWSB_NF; ++u) gs[u] += s_warp[u * fr_floats + UUO_SKIN_J * 12 + n3p + 3 * r + lane];
      }
#pragma unroll
      for (int u = 0; u < ROWSB_NF; ++u) {
        if (f0 + u >= F) break;
        float* dstA = gA_part + (size_t)(f0 + u) * UUO_GA_STRIDE;
        if (lane < J) {
          float4* d4 = reinterpret_cast<float4*>(dstA + lane * 12);
          d4[0] = make_float4(ga[u][0], ga[u][1], ga[u][2], ga[u][3]);
          d4[1] = make_float4(ga[u][4], ga[u][5], ga[u][6], ga[u][7]);
          d4[2] = make_float4(ga[u][8], ga[u][9], ga[u][10], ga[u][11]);
        }
        if (lane < 4) dstA[J * 12 + lane] = (lane < 3) ? gs[u] : 0.f;
      }
    }
    __syncwarp();
  }
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
  const rowss_layout L = rowss_make_layout(m->KP, R, m->J);
  ProfScope ps_("rows_forward", st);
  const int G = rows_shared > 1 ? rows_shared : 1;           // rows_shared = G > 1: G contiguous frame groups, rows[G][R]
  if (F % G) { uuo_set_error("rows path: F must be a multiple of the number of row groups"); return UUO_ERR_ARG; }
  if (rows_shared && 3 * R <= 32 * ROWSS_NI && L.total <= 227 * 1024 && F / G >= 64) {
    UUO_DYN_SMEM(rows_forward_shared_kernel, L.total);
    const int n_sm = uuo_sm_count();
    const int Fg = F / G;
    const int need = (Fg + ROWSS_WARPS * ROWSS_NF - 1) / (ROWSS_WARPS * ROWSS_NF);
    const int per_group = std::max(1, std::min(need, n_sm / G));      // one CTA per SM at most: a single wave
    rows_forward_shared_kernel<<<dim3(per_group, G), ROWSS_WARPS * 32, L.total, st>>>(
        m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, transl, Fg, rows, R, out_rows, rowbuf);
  } else {
    rows_forward_kernel<<<(F + ROWS_WARPS - 1) / ROWS_WARPS, ROWS_WARPS * 32, 0, st>>>(
        m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, transl, F, rows, rows_shared, R, out_rows, rowbuf);
  }
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
  const int G = rows_shared > 1 ? rows_shared : 1;
  const rowsb_layout LB = rowsb_make_layout(m->KP, R, m->J);
  if (vposed_cached && rows_shared && F % G == 0 && F / G >= 64 && LB.total <= 227 * 1024 && m->J <= 255) {
    // rows shared by all frames (of a group): the selected Bt rows stay in shared memory
    UUO_DYN_SMEM(rows_backward_shared_kernel, LB.total);
    const int n_sm = uuo_sm_count();
    const int Fg = F / G;
    const int need = (Fg + ROWSB_WARPS * ROWSB_NF - 1) / (ROWSB_WARPS * ROWSB_NF);
    const int per_group = std::max(1, std::min(need, n_sm / G));      // one CTA per SM at most: a single wave
    ProfScope ps_("rows_backward", st);
    rows_backward_shared_kernel<<<dim3(per_group, G), ROWSB_WARPS * 32, LB.total, st>>>(
        m->Bt, m->W, m->KP, m->J, ws.A, Fg, rows, R, g_rows, ws.rowbuf, ws.gA, ws.gcoef);
    UUO_LAUNCH_CHECK();
    return UUO_OK;
  }
  { ProfScope ps_("rows_backward", st);
  if (vposed_cached)
    rows_backward_kernel<true><<<(F + ROWS_WARPS - 1) / ROWS_WARPS, ROWS_WARPS * 32, 0, st>>>(
        m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, F, rows, rows_shared, R, g_rows, ws.rowbuf, ws.gA, ws.gcoef);
  else
    rows_backward_kernel<false><<<(F + ROWS_WARPS - 1) / ROWS_WARPS, ROWS_WARPS * 32, 0, st>>>(
        m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, F, rows, rows_shared, R, g_rows, ws.rowbuf, ws.gA, ws.gcoef);
  } UUO_LAUNCH_CHECK();
  return UUO_OK;
}
