// Row-sparse LBS: forward of R selected vertices per frame and backward of a gradient that touches
// only R vertex rows per frame.  This is what the fitting closures need:
//   marker stage  : virtual markers = vertices[:, idx]   (optimization.py:345-351, a gather)
//   chamfer stage : d loss / d vertices is non-zero only at the argmin rows (pytorch3d knn backward)
// One warp per frame; lanes span the K = P + NB coefficients for the blend-shape dot products and the
// 288 entries of dA for the skinning backward.  Frames are independent, rows are accumulated in order
// (deterministic).  Nothing of size [F,V] is touched.
#include "common.cuh"

#define FULL 0xffffffffu
#define ROWS_MAXK 8   // KP <= 256

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// v_posed of one vertex: vt + coef . Bt[3v+c]   (all lanes get the 3 results)
__device__ __forceinline__ void row_vposed(const float* __restrict__ vt, const float* __restrict__ Bt, int KP,
                                           const float* cf /*[ROWS_MAXK] per lane*/, int v, int lane, float* vp) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float* b = Bt + (size_t)(3 * v + c) * KP;
    float a = 0.f;
#pragma unroll
    for (int q = 0; q < ROWS_MAXK; ++q) {
      int k = lane + 32 * q;
      if (k < KP) a = fmaf(cf[q], b[k], a);
    }
    vp[c] = wsum(a) + vt[3 * v + c];
  }
}

__global__ void __launch_bounds__(128) rows_forward_kernel(
    const float* __restrict__ vt, const float* __restrict__ Bt, const float* __restrict__ W, int KP, int J,
    const float* __restrict__ coef, const float* __restrict__ A, const float* __restrict__ transl, int F,
    const int32_t* __restrict__ rows, int rows_shared, int R, float* __restrict__ out_rows) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (f >= F) return;
  float cf[ROWS_MAXK];
#pragma unroll
  for (int q = 0; q < ROWS_MAXK; ++q) {
    int k = lane + 32 * q;
    cf[q] = (k < KP) ? coef[(size_t)f * KP + k] : 0.f;
  }
  const float* Af = A + (size_t)f * J * 12;
  for (int r = 0; r < R; ++r) {
    const int v = rows_shared ? rows[r] : rows[(size_t)f * R + r];
    float vp[3];
    row_vposed(vt, Bt, KP, cf, v, lane, vp);
    // lane i < 12 owns T entry i
    float t = 0.f;
    if (lane < 12) {
      for (int j = 0; j < J; ++j) t = fmaf(W[(size_t)v * J + j], Af[j * 12 + lane], t);
    }
    // out[a] = T[a][0..2].vp + T[a][3]
    float contrib = 0.f;
    if (lane < 12) {
      int b = lane & 3;
      contrib = (b < 3) ? t * vp[b] : t;
    }
    // sum the 4 lanes of each row a
    contrib += __shfl_xor_sync(FULL, contrib, 1);
    contrib += __shfl_xor_sync(FULL, contrib, 2);
    if (lane < 12 && (lane & 3) == 0) {
      int a = lane >> 2;
      out_rows[((size_t)f * R + r) * 3 + a] = contrib + (transl ? transl[(size_t)f * 3 + a] : 0.f);
    }
  }
}

__global__ void __launch_bounds__(128) rows_backward_kernel(
    const float* __restrict__ vt, const float* __restrict__ Bt, const float* __restrict__ W, int KP, int J,
    const float* __restrict__ coef, const float* __restrict__ A, int F, const int32_t* __restrict__ rows,
    int rows_shared, int R, const float* __restrict__ g_rows, float* __restrict__ gA_part,
    float* __restrict__ gcoef_part) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (f >= F) return;
  float cf[ROWS_MAXK], gc[ROWS_MAXK];
#pragma unroll
  for (int q = 0; q < ROWS_MAXK; ++q) {
    int k = lane + 32 * q;
    cf[q] = (k < KP) ? coef[(size_t)f * KP + k] : 0.f;
    gc[q] = 0.f;
  }
  float ga[9];   // dA entries e = lane + 32*q, e < 288
#pragma unroll
  for (int q = 0; q < 9; ++q) ga[q] = 0.f;
  float gs = 0.f;  // lane a < 3: sum of g over rows
  const float* Af = A + (size_t)f * J * 12;
  for (int r = 0; r < R; ++r) {
    const int v = rows_shared ? rows[r] : rows[(size_t)f * R + r];
    const float* gp = g_rows + ((size_t)f * R + r) * 3;
    const float g0 = gp[0], g1 = gp[1], g2 = gp[2];
    if (g0 == 0.f && g1 == 0.f && g2 == 0.f) continue;   // masked / dropped markers (warp-uniform)
    float vp[3];
    row_vposed(vt, Bt, KP, cf, v, lane, vp);
    if (lane < 3) gs += (lane == 0 ? g0 : (lane == 1 ? g1 : g2));
    // T[:3,:3] entries on lanes 0..8 (row a = lane/3, col b = lane%3)
    float t = 0.f;
    if (lane < 9) {
      int a = lane / 3, b = lane % 3;
      for (int j = 0; j < J; ++j) t = fmaf(W[(size_t)v * J + j], Af[j * 12 + a * 4 + b], t);
    }
    // g_vposed[c] = sum_a T[a][c] g[a]
    float gvp[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float t0 = __shfl_sync(FULL, t, c), t1 = __shfl_sync(FULL, t, 3 + c), t2 = __shfl_sync(FULL, t, 6 + c);
      gvp[c] = t0 * g0 + t1 * g1 + t2 * g2;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* b = Bt + (size_t)(3 * v + c) * KP;
#pragma unroll
      for (int q = 0; q < ROWS_MAXK; ++q) {
        int k = lane + 32 * q;
        if (k < KP) gc[q] = fmaf(gvp[c], b[k], gc[q]);
      }
    }
    // dA[j][a][b] += W[v][j] g[a] vh[b]
#pragma unroll
    for (int q = 0; q < 9; ++q) {
      int e = lane + 32 * q;
      int j = e / 12, ab = e % 12, a = ab >> 2, b = ab & 3;
      float ga_ = (a == 0 ? g0 : (a == 1 ? g1 : g2));
      float vh = (b == 0 ? vp[0] : (b == 1 ? vp[1] : (b == 2 ? vp[2] : 1.f)));
      ga[q] = fmaf(W[(size_t)v * J + j] * ga_, vh, ga[q]);
    }
  }
  float* dstA = gA_part + (size_t)f * UUO_GA_STRIDE;
#pragma unroll
  for (int q = 0; q < 9; ++q) dstA[lane + 32 * q] = ga[q];
  if (lane < 4) dstA[288 + lane] = (lane < 3) ? gs : 0.f;
  float* dstC = gcoef_part + (size_t)f * KP;
#pragma unroll
  for (int q = 0; q < ROWS_MAXK; ++q) {
    int k = lane + 32 * q;
    if (k < KP) dstC[k] = gc[q];
  }
}

int launch_rows_forward(const uuo_model* m, const float* transl, int F, const int32_t* rows, int rows_shared,
                        int R, float* out_rows, const uuo_ws& ws, cudaStream_t st) {
  if (m->KP > 32 * ROWS_MAXK) { uuo_set_error("rows path: KP too large"); return UUO_ERR_ARG; }
  { ProfScope ps_("rows_forward", st);
  rows_forward_kernel<<<(F + 3) / 4, 128, 0, st>>>(m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, transl, F, rows,
                                                   rows_shared, R, out_rows);
  } UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_rows_backward(const uuo_model* m, int F, const int32_t* rows, int rows_shared, int R,
                         const float* g_rows, const uuo_ws& ws, cudaStream_t st) {
  if (m->KP > 32 * ROWS_MAXK) { uuo_set_error("rows path: KP too large"); return UUO_ERR_ARG; }
  { ProfScope ps_("rows_backward", st);
  rows_backward_kernel<<<(F + 3) / 4, 128, 0, st>>>(m->vt, m->Bt, m->W, m->KP, m->J, ws.coef, ws.A, F, rows,
                                                    rows_shared, R, g_rows, ws.gA, ws.gcoef);
  } UUO_LAUNCH_CHECK();
  return UUO_OK;
}
