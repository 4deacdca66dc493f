// Rows SHARED by all frames (marker stage: the same M vertices in every frame, optimization.py:335-351) with the
// blend-shape part on the tensor cores.  The selected rows are fixed, so
//   v_posed_sel[F][3R] = coef[F][KP] . Bt_sel[3R][KP]^T + vt_sel          (forward)
//   d coef[F][KP]      = d v_posed_sel[F][3R] . Bt_sel[3R][KP]            (backward)
// are two small dense GEMMs: they run through umma_gemm_kernel<3> (3xTF32, fp32 accuracy) with operand images of the
// selected rows that sel_build_kernel writes in the layout that kernel expects (a few hundred KB, rebuilt by every
// call: the library cannot know when the caller changes `rows`).  What is left per frame - the blended transform of
// each row through its non-zero weights, the 3x4 apply, d v_posed = T^T g, d A - is one warp per frame with little
// shared memory, so many CTAs fit an SM (the shared-memory kernels of rows.cu hold 121 KB of Bt rows: one CTA per SM,
// latency bound).  Used when rows_shared == 1; row groups (stacked hypotheses) stay on the kernels of rows.cu.
#include <algorithm>

#include "common.cuh"
#include "umma.cuh"

namespace {

constexpr int RT_WARPS = 8;
constexpr int RT_CHUNK = 64;                    // rowbuf is [F][64][4] (rows.cu: ROWS_CHUNK)
constexpr int MAXR = 64, MAXJ = UUO_SKIN_J;
// ws.sel_img carve-up (floats)
constexpr int OFF_IMGF = 0;                     // [KP/32][2][NTf*32]   B image of the forward  (rows n = 3r + c, K = k)
constexpr int OFF_IMGB = 8 * 2 * 192 * 32;      // [NTf/32][2][KP*32]   B image of the backward (rows k, K = n)
constexpr int OFF_BIAS = OFF_IMGB + 6 * 2 * 256 * 32;   // [192] vt of the selected rows
constexpr int OFF_ELLW = OFF_BIAS + 192;        // [R][J] non-zero weights of a row (ascending joint)
constexpr int OFF_ELLJ = OFF_ELLW + MAXR * MAXJ;
constexpr int OFF_CSCW = OFF_ELLJ + MAXR * MAXJ;   // [J][R] non-zero weights of a joint (ascending row)
constexpr int OFF_CSCR = OFF_CSCW + MAXR * MAXJ;
constexpr int OFF_CNT = OFF_CSCR + MAXR * MAXJ;    // [64] per row, then [24] per joint
static_assert(OFF_CNT + 128 <= UUO_SEL_IMG_FLOATS, "sel_img too small");

__device__ __forceinline__ int sw_off(int row, int kk) { return row * 32 + ((((kk >> 2) ^ (row & 7))) << 2) + (kk & 3); }

__global__ void __launch_bounds__(256) sel_build_kernel(const float* __restrict__ vt, const float* __restrict__ Bt,
                                                        const float* __restrict__ W, int KP, int J,
                                                        const int32_t* __restrict__ rows, int R, int NTf,
                                                        float* __restrict__ img) {
  const int n3 = 3 * R, kbf = KP / 32, kbb = NTf / 32;
  rows += (size_t)blockIdx.y * R;                              // group blockIdx.y: its own rows and its own image
  img += (size_t)blockIdx.y * UUO_SEL_IMG_FLOATS;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  for (int i = tid; i < kbf * NTf * 32; i += nthr) {
    const int kb = i / (NTf * 32), rem = i - kb * NTf * 32, n = rem >> 5, kk = rem & 31;
    const float x = n < n3 ? __ldg(Bt + (size_t)(3 * rows[n / 3] + n % 3) * KP + kb * 32 + kk) : 0.f;
    const float hi = umma::to_tf32(x);
    float* t = img + OFF_IMGF + (size_t)kb * 2 * NTf * 32;
    t[sw_off(n, kk)] = hi;
    t[NTf * 32 + sw_off(n, kk)] = umma::to_tf32(x - hi);
  }
  for (int i = tid; i < kbb * KP * 32; i += nthr) {
    const int kb = i / (KP * 32), rem = i - kb * KP * 32, k = rem >> 5, nn = rem & 31, n = kb * 32 + nn;
    const float x = n < n3 ? __ldg(Bt + (size_t)(3 * rows[n / 3] + n % 3) * KP + k) : 0.f;
    const float hi = umma::to_tf32(x);
    float* t = img + OFF_IMGB + (size_t)kb * 2 * KP * 32;
    t[sw_off(k, nn)] = hi;
    t[KP * 32 + sw_off(k, nn)] = umma::to_tf32(x - hi);
  }
  for (int n = tid; n < 192; n += nthr) img[OFF_BIAS + n] = n < n3 ? vt[3 * rows[n / 3] + n % 3] : 0.f;
  if (blockIdx.x == 0) {
    int* cnt = reinterpret_cast<int*>(img + OFF_CNT);
    int* ellj = reinterpret_cast<int*>(img + OFF_ELLJ);
    int* cscr = reinterpret_cast<int*>(img + OFF_CSCR);
    const int t = threadIdx.x;
    if (t < R) {
      const float* wr = W + (size_t)rows[t] * J;
      int c = 0;
      for (int j = 0; j < J; ++j) {
        const float x = wr[j];
        if (x != 0.f) { img[OFF_ELLW + t * J + c] = x; ellj[t * J + c] = j; ++c; }
      }
      cnt[t] = c;
    } else if (t >= 64 && t < 64 + J) {
      const int j = t - 64;
      int c = 0;
      for (int r = 0; r < R; ++r) {
        const float x = W[(size_t)rows[r] * J + j];
        if (x != 0.f) { img[OFF_CSCW + j * R + c] = x; cscr[j * R + c] = r; ++c; }
      }
      cnt[64 + j] = c;
    }
  }
}

// out_rows[f][r] = T(f, r) (v_posed_sel[f][r], 1) + transl[f]; also leaves v_posed_sel in rowbuf for the backward
__global__ void __launch_bounds__(RT_WARPS * 32) rows_tc_fwd_kernel(const float* __restrict__ A, int J,
                                                                    const float* __restrict__ transl, int F, int R,
                                                                    const float* __restrict__ sel_vp,
                                                                    const float* __restrict__ img,
                                                                    float* __restrict__ out_rows,
                                                                    float* __restrict__ rowbuf) {
  __shared__ __align__(16) float s_A_all[RT_WARPS][UUO_SKIN_J * 12];
  __shared__ float s_ellw[MAXR * MAXJ];
  __shared__ int s_ellj[MAXR * MAXJ];
  __shared__ int s_ellc[MAXR];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (gridDim.y > 1) {       // row groups: group g = blockIdx.y owns frames [g F, (g+1) F) and image g
    const size_t fb = (size_t)blockIdx.y * F;
    img += (size_t)blockIdx.y * UUO_SEL_IMG_FLOATS;
    A += fb * J * 12;
    if (transl) transl += fb * 3;
    sel_vp += fb * UUO_SEL_LD;
    out_rows += fb * R * 3;
    if (rowbuf) rowbuf += fb * RT_CHUNK * 4;
  }
  for (int i = tid; i < R * J; i += RT_WARPS * 32) {
    s_ellw[i] = img[OFF_ELLW + i];
    s_ellj[i] = reinterpret_cast<const int*>(img + OFF_ELLJ)[i];
  }
  if (tid < R) s_ellc[tid] = reinterpret_cast<const int*>(img + OFF_CNT)[tid];
  __syncthreads();
  float* s_A = s_A_all[w];
  const float4* s_A4 = reinterpret_cast<const float4*>(s_A);
  for (int f = blockIdx.x * RT_WARPS + w; f < F; f += gridDim.x * RT_WARPS) {
    const float4* a4 = reinterpret_cast<const float4*>(A + (size_t)f * J * 12);
    for (int i = lane; i < J * 3; i += 32) reinterpret_cast<float4*>(s_A)[i] = __ldg(a4 + i);
    __syncwarp();
    float tr[3] = {0.f, 0.f, 0.f};
    if (transl) { tr[0] = transl[(size_t)f * 3]; tr[1] = transl[(size_t)f * 3 + 1]; tr[2] = transl[(size_t)f * 3 + 2]; }
    for (int r = lane; r < R; r += 32) {
      const float* vp = sel_vp + (size_t)f * UUO_SEL_LD + 3 * r;
      const float vp0 = vp[0], vp1 = vp[1], vp2 = vp[2];
      if (rowbuf) reinterpret_cast<float4*>(rowbuf)[(size_t)f * RT_CHUNK + r] = make_float4(vp0, vp1, vp2, 0.f);
      float T[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) T[i] = 0.f;
      const int cnt = s_ellc[r];
      for (int i = 0; i < cnt; ++i) {
        const float wj = s_ellw[r * J + i];
        const int j = s_ellj[r * J + i];
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
    __syncwarp();
  }
}

// d v_posed_sel (-> sel_gvp, the A operand of the backward GEMM, padding columns = 0) and d A of one frame per warp;
// v_posed of the rows comes from rowbuf (the forward, or uuo_lbs_cache_vposed_rows)
__global__ void __launch_bounds__(RT_WARPS * 32) rows_tc_bwd_kernel(const float* __restrict__ A, int J, int F, int R,
                                                                    int kcols, const float* __restrict__ g_rows,
                                                                    const float* __restrict__ rowbuf,
                                                                    const float* __restrict__ img,
                                                                    float* __restrict__ sel_gvp,
                                                                    float* __restrict__ gA_part) {
  __shared__ __align__(16) float s_A_all[RT_WARPS][UUO_SKIN_J * 12];
  __shared__ float s_g_all[RT_WARPS][3 * MAXR], s_vp_all[RT_WARPS][3 * MAXR];
  __shared__ float s_ellw[MAXR * MAXJ], s_cscw[MAXR * MAXJ];
  __shared__ uint8_t s_ellj[MAXR * MAXJ], s_cscr[MAXR * MAXJ];
  __shared__ int s_cnt[128];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (gridDim.y > 1) {
    const size_t fb = (size_t)blockIdx.y * F;
    img += (size_t)blockIdx.y * UUO_SEL_IMG_FLOATS;
    A += fb * J * 12;
    g_rows += fb * R * 3;
    rowbuf += fb * RT_CHUNK * 4;
    sel_gvp += fb * UUO_SEL_LD;
    gA_part += fb * UUO_GA_STRIDE;
  }
  for (int i = tid; i < R * J; i += RT_WARPS * 32) {
    s_ellw[i] = img[OFF_ELLW + i];
    s_ellj[i] = (uint8_t) reinterpret_cast<const int*>(img + OFF_ELLJ)[i];
    s_cscw[i] = img[OFF_CSCW + i];
    s_cscr[i] = (uint8_t) reinterpret_cast<const int*>(img + OFF_CSCR)[i];
  }
  if (tid < 128) s_cnt[tid] = reinterpret_cast<const int*>(img + OFF_CNT)[tid];
  __syncthreads();
  float* s_A = s_A_all[w];
  float* s_g = s_g_all[w];
  float* s_vp = s_vp_all[w];
  const float4* s_A4 = reinterpret_cast<const float4*>(s_A);
  const int n3 = 3 * R;
  for (int f = blockIdx.x * RT_WARPS + w; f < F; f += gridDim.x * RT_WARPS) {
    const float4* a4 = reinterpret_cast<const float4*>(A + (size_t)f * J * 12);
    for (int i = lane; i < J * 3; i += 32) reinterpret_cast<float4*>(s_A)[i] = __ldg(a4 + i);
    __syncwarp();
    float* gv = sel_gvp + (size_t)f * UUO_SEL_LD;
    for (int r = lane; r < R; r += 32) {
      const float* gp = g_rows + ((size_t)f * R + r) * 3;
      const float g0 = gp[0], g1 = gp[1], g2 = gp[2];
      const float4 c = reinterpret_cast<const float4*>(rowbuf)[(size_t)f * RT_CHUNK + r];
      float T[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) T[i] = 0.f;
      const int cnt = s_cnt[r];
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
      for (int q = 0; q < 3; ++q) gv[3 * r + q] = off ? 0.f : T[q] * g0 + T[3 + q] * g1 + T[6 + q] * g2;
    }
    for (int n = n3 + lane; n < kcols; n += 32) gv[n] = 0.f;   // K padding of the GEMM operand: 0 x 0, never garbage
    __syncwarp();
    // dA[j = lane][a][b] = sum_r W[r][j] g[r][a] (v_posed[r], 1)[b] over the rows where W[r][j] != 0, rows ascending
    float ga[12], gs = 0.f;
#pragma unroll
    for (int i = 0; i < 12; ++i) ga[i] = 0.f;
    const int cnt = lane < J ? s_cnt[64 + lane] : 0;
    for (int i = 0; i < cnt; ++i) {
      const float wj = s_cscw[lane * R + i];
      const int r = s_cscr[lane * R + i];
      const float v0 = s_vp[3 * r], v1 = s_vp[3 * r + 1], v2 = s_vp[3 * r + 2];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const float wg = wj * s_g[3 * r + a];
        ga[a * 4] = fmaf(wg, v0, ga[a * 4]);
        ga[a * 4 + 1] = fmaf(wg, v1, ga[a * 4 + 1]);
        ga[a * 4 + 2] = fmaf(wg, v2, ga[a * 4 + 2]);
        ga[a * 4 + 3] += wg;
      }
    }
    if (lane < 3)
      for (int r = 0; r < R; ++r) gs += s_g[3 * r + lane];
    float* dstA = gA_part + (size_t)f * UUO_GA_STRIDE;
    if (lane < J) {
      float4* d4 = reinterpret_cast<float4*>(dstA + lane * 12);
      d4[0] = make_float4(ga[0], ga[1], ga[2], ga[3]);
      d4[1] = make_float4(ga[4], ga[5], ga[6], ga[7]);
      d4[2] = make_float4(ga[8], ga[9], ga[10], ga[11]);
    }
    if (lane < 4) dstA[J * 12 + lane] = (lane < 3) ? gs : 0.f;
    __syncwarp();
  }
}

int sel_build(const uuo_model* m, const int32_t* rows, int R, int NTf, int G, const uuo_ws& ws, cudaStream_t st) {
  sel_build_kernel<<<dim3(G > 1 ? 8 : 64, G), 256, 0, st>>>(m->vt, m->Bt, m->W, m->KP, m->J, rows, R, NTf, ws.sel_img);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

dim3 frame_grid(int Fg, int G) {
  const int need = (Fg + RT_WARPS - 1) / RT_WARPS;
  return dim3(std::max(1, std::min(need, 8 * uuo_sm_count() / G)), G);
}

}  // namespace

// true when the tensor-core path takes this call (launch_rows_forward / launch_rows_backward ask first):
// rows_shared = 1 (all frames) or G row groups of at least 64 frames each, as many as the workspace has images for
bool rows_tc_usable(const uuo_model* m, int F, int rows_shared, int R) {
  if (rows_shared < 1 || R < 1 || R > MAXR || m->J != UUO_SKIN_J || (m->KP & 31) || m->KP > 256) return false;
  return F % rows_shared == 0 && F / rows_shared >= 64 && rows_shared <= uuo_sel_images(F);
}

int launch_rows_forward_tc(const uuo_model* m, const float* transl, int F, const int32_t* rows, int G, int R,
                           float* out_rows, const uuo_ws& ws, cudaStream_t st) {
  const int NTf = (3 * R + 31) / 32 * 32;
  int rc = sel_build(m, rows, R, NTf, G, ws, st);
  if (rc) return rc;
  if ((rc = launch_umma_one_tile(ws.coef, m->KP, F, ws.sel_img + OFF_IMGF, NTf, m->KP / 32, ws.sel_img + OFF_BIAS,
                                 ws.sel_vp, UUO_SEL_LD, G, UUO_SEL_IMG_FLOATS, UUO_SEL_IMG_FLOATS, st))) return rc;
  rows_tc_fwd_kernel<<<frame_grid(F / G, G), RT_WARPS * 32, 0, st>>>(ws.A, m->J, transl, F / G, R, ws.sel_vp, ws.sel_img,
                                                                    out_rows, ws.rowbuf);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_rows_backward_tc(const uuo_model* m, int F, const int32_t* rows, int G, int R, const float* g_rows,
                            const uuo_ws& ws, cudaStream_t st) {
  const int NTf = (3 * R + 31) / 32 * 32;
  int rc = sel_build(m, rows, R, NTf, G, ws, st);     // the forward of this iteration may have been a dense one
  if (rc) return rc;
  rows_tc_bwd_kernel<<<frame_grid(F / G, G), RT_WARPS * 32, 0, st>>>(ws.A, m->J, F / G, R, NTf, g_rows, ws.rowbuf, ws.sel_img,
                                                                    ws.sel_gvp, ws.gA);
  UUO_LAUNCH_CHECK();
  return launch_umma_one_tile(ws.sel_gvp, UUO_SEL_LD, F, ws.sel_img + OFF_IMGB, m->KP, NTf / 32, nullptr, ws.gcoef, m->KP,
                              G, UUO_SEL_IMG_FLOATS, 0, st);
}
