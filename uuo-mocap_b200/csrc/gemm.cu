// The two dense contractions of lbs(): exact-fp32 SIMT path.
//   forward  v_posed[F,3V] = v_template + [pose_feature | betas][F,KP] @ Bmat[KP,NP]
//            (lbs.py:209 blend_shapes + lbs.py:229-233 posedirs matmul, one GEMM of K = P + NB)
//   backward g_coef[F,KP]  = g_vposed[F,NP] @ Bt[NP,KP]   (split-K, deterministic partials)
// Register-tiled 16x16 threads, TM x TN outputs per thread, BK = 8, register-prefetched double buffer.
#include <stdlib.h>

#include "common.cuh"

#ifdef UUO_DEBUG_SIMT   // exact-fp32 SIMT cross-check of the tensor-core GEMMs: debug builds only (tools/gemm_check.py)
template <int TM, int TN>
__global__ void __launch_bounds__(256) sgemm_kernel(const float* __restrict__ Amat, int lda, int M,
                                                    const float* __restrict__ Bm, int ldb, int Kchunk,
                                                    float* __restrict__ C, int ldc, int N,
                                                    const float* __restrict__ bias, size_t c_split_stride) {
  constexpr int BM = 16 * TM, BN = 16 * TN, BK = 8;
  constexpr int A_F4 = BM * BK / 4;   // float4 loads of the A tile
  constexpr int B_F4 = BK * BN / 4;
  constexpr int A_PER = (A_F4 + 255) / 256, B_PER = (B_F4 + 255) / 256;
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int k_begin = blockIdx.z * Kchunk;
  C += (size_t)blockIdx.z * c_split_stride;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  float4 ra[A_PER], rb[B_PER];
  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < A_PER; ++i) {
      int q = tid + i * 256;
      if (q < A_F4) {
        int row = q / (BK / 4), c4 = q % (BK / 4);
        int gr = min(m0 + row, M - 1);
        ra[i] = *reinterpret_cast<const float4*>(Amat + (size_t)gr * lda + k0 + c4 * 4);
      }
    }
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
      int q = tid + i * 256;
      if (q < B_F4) {
        int row = q / (BN / 4), c4 = q % (BN / 4);
        rb[i] = *reinterpret_cast<const float4*>(Bm + (size_t)(k0 + row) * ldb + n0 + c4 * 4);
      }
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_PER; ++i) {
      int q = tid + i * 256;
      if (q < A_F4) {
        int row = q / (BK / 4), c4 = q % (BK / 4);
        As[buf][c4 * 4 + 0][row] = ra[i].x;
        As[buf][c4 * 4 + 1][row] = ra[i].y;
        As[buf][c4 * 4 + 2][row] = ra[i].z;
        As[buf][c4 * 4 + 3][row] = ra[i].w;
      }
    }
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
      int q = tid + i * 256;
      if (q < B_F4) {
        int row = q / (BN / 4), c4 = q % (BN / 4);
        *reinterpret_cast<float4*>(&Bs[buf][row][c4 * 4]) = rb[i];
      }
    }
  };

  const int nk = Kchunk / BK;
  gload(k_begin);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload(k_begin + (kt + 1) * BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      if constexpr (TM % 4 == 0) {
#pragma unroll
        for (int i = 0; i < TM; i += 4) {
          float4 v = *reinterpret_cast<const float4*>(&As[buf][kk][ty * TM + i]);
          a[i] = v.x; a[i + 1] = v.y; a[i + 2] = v.z; a[i + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < TM; ++i) a[i] = As[buf][kk][ty * TM + i];
      }
      if constexpr (TN % 4 == 0) {
#pragma unroll
        for (int j = 0; j < TN; j += 4) {
          float4 v = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4 + 16 * j]);
          b[j] = v.x; b[j + 1] = v.y; b[j + 2] = v.z; b[j + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < TN; ++j) b[j] = Bs[buf][kk][tx + 16 * j];
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int gr = m0 + ty * TM + i;
    if (gr >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int gc = n0 + ((TN % 4 == 0) ? (tx * 4 + (j & 3) + 64 * (j >> 2)) : (tx + 16 * j));
      if (gc < N) C[(size_t)gr * ldc + gc] = acc[i][j] + (bias ? bias[gc] : 0.f);
    }
  }
}

#endif  // UUO_DEBUG_SIMT

// The release library has ONE GEMM path (tcgen05).  A build with -DUUO_DEBUG_SIMT keeps the SIMT kernels above and
// lets UUO_GEMM_SIMT=1 select them per call (debug cross-check of the tensor-core path).
#ifndef UUO_DEBUG_SIMT
int launch_vposed_gemm(const uuo_model* m, int F, const uuo_ws& ws, bool betas_shared, cudaStream_t st) {
  return launch_vposed_gemm_umma(m, F, ws, betas_shared, st);
}
int launch_gcoef_gemm(const uuo_model* m, int F, const uuo_ws& ws, int* n_parts, cudaStream_t st) {
  *n_parts = uuo_gcoef_splits(m, F);
  return launch_gcoef_gemm_umma(m, F, ws, st);
}
#else
static bool use_simt() {
  const char* e = getenv("UUO_GEMM_SIMT");   // read per call so one process can compare both paths
  return e && e[0] == '1';
}

int launch_vposed_gemm(const uuo_model* m, int F, const uuo_ws& ws, bool betas_shared, cudaStream_t st) {
  if (!use_simt()) return launch_vposed_gemm_umma(m, F, ws, betas_shared, st);
  // M = F, K = KP, N = NP (padded columns are zeros in Bmat and vt)
  dim3 grid(m->NP / 128, (F + 127) / 128, 1);
  { ProfScope ps_("vposed_gemm", st);
  sgemm_kernel<8, 8><<<grid, 256, 0, st>>>(ws.coef, m->KP, F, m->Bmat, m->NP, m->KP, ws.vposed, m->NP, m->NP,
                                           m->vt, 0);
  } UUO_LAUNCH_CHECK();
  return UUO_OK;
}

// returns the number of partial slabs written to ws.gcoef through *n_parts
int launch_gcoef_gemm(const uuo_model* m, int F, const uuo_ws& ws, int* n_parts, cudaStream_t st) {
  if (!use_simt()) { *n_parts = uuo_gcoef_splits(m, F); return launch_gcoef_gemm_umma(m, F, ws, st); }
  // M = F, K = NP split in 8 chunks, N = KP
  const int nsplit = 8;
  *n_parts = nsplit;
  const int kchunk = m->NP / nsplit;   // NP is a multiple of 256 -> chunk multiple of 32
  dim3 grid((m->KP + 111) / 112, (F + 127) / 128, nsplit);
  { ProfScope ps_("gcoef_gemm", st);
  sgemm_kernel<8, 7><<<grid, 256, 0, st>>>(ws.gvposed, m->NP, F, m->Bt, m->KP, kchunk, ws.gcoef, m->KP, m->KP,
                                           nullptr, (size_t)F * m->KP);
  } UUO_LAUNCH_CHECK();
  return UUO_OK;
}
#endif  // UUO_DEBUG_SIMT
