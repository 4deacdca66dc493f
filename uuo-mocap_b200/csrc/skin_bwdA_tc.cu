// d(A) of linear blend skinning on the tensor cores:
//   dA[f,j,(r,c)] = sum_v W[v,j] g[f,v,r] [v_posed[f,v]; 1][c]        (adjoint of lbs.py:237-250)
//   dtransl-like row  gsum[f,r] = sum_v g[f,v,r]                      (a 25th all-ones "joint" column)
// is the GEMM  D[(f,e), j] = Q[(f,e), v] . W'[j, v]^T  contracted over the vertices:
//   M = 128 rows = 10 frames x 12 entries (120 used), N = 32 (24 joints, ones, pad), K = vertices (blocks of 32).
// Q = g (x) [v_posed;1] never exists in HBM: 12 producer warps (4 groups of 3, group = pipeline stage) read g and
// v_posed once, form the 12 products per (frame, vertex), split them hi/lo for 3xTF32 and store the SWIZZLE_128B
// K-major image straight into shared memory.  W'^T is a model constant image fetched by bulk copy.  The vertex range
// is split into UUO_NSPLIT_GA deterministic partial slabs (summed by fk_backward).
#include <algorithm>

#include "common.cuh"
#include "umma.cuh"

namespace {

constexpr int FT = 10;                   // frames per tile -> 120 of the 128 MMA rows
constexpr int BK = 32;                   // vertices per K block
constexpr int NJ = 32;                   // MMA N
constexpr int Q_BYTES = 128 * BK * 4;    // 16 KB per half
constexpr int WT_BYTES = NJ * BK * 4;    // 4 KB per half
constexpr int STAGE_BYTES = 2 * Q_BYTES + 2 * WT_BYTES;   // 40 KB
constexpr int STAGES = 4;
constexpr int EPI_WARPS = 4, GROUP_WARPS = 3;
constexpr int WARP_MMA = EPI_WARPS, WARP_LOAD = EPI_WARPS + 1, WARP_PROD0 = EPI_WARPS + 2;
constexpr int NTHREADS = (EPI_WARPS + 2 + STAGES * GROUP_WARPS) * 32;   // 576
constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + 256;

struct BwdAArgs {
  const float* WTimg;      // [kb_total][2][32*32]
  const float* vposed;     // [F][NP]
  const float* g_vertices; // [F][V][3] or null
  const float* g_joints;   // [F][J+NE][3] or null
  const int32_t* extra_ids;
  const uint8_t* vflag;    // [NP/3]
  float* gA;               // [splits][F][UUO_GA_STRIDE]
  int F, V, NP, NE, J, n_ftiles, splits, kb_per_split;
};

__global__ void __launch_bounds__(NTHREADS, 1) skin_bwdA_tc_kernel(const BwdAArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = umma::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + STAGES * STAGE_BYTES);
  // bars: [0,S) full_q, [S,2S) full_w, [2S,3S) empty, [3S,3S+2) tmem_full, [3S+2,3S+4) tmem_empty
  const uint32_t bar0 = umma::smem_u32(bars);
  auto FULL_Q = [&](int s) { return bar0 + 8u * s; };
  auto FULL_W = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto EMPTY = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto TFULL = [&](int q) { return bar0 + 8u * (3 * STAGES + q); };
  auto TEMPTY = [&](int q) { return bar0 + 8u * (3 * STAGES + 2 + q); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = a.n_ftiles * a.splits;
  const int my_items = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // items b, b+G, ...
  const uint32_t total_kb = (uint32_t)my_items * a.kb_per_split;

  // rows 120..127 of every Q tile are never written by the producers: clear them once
  for (int i = threadIdx.x; i < STAGES * 2 * 8 * 32; i += NTHREADS) {
    const int s = i / (2 * 8 * 32), r = i % (2 * 8 * 32);
    reinterpret_cast<float*>(sm + s * STAGE_BYTES + (r / 256) * Q_BYTES + 120 * 128)[r % 256] = 0.f;
  }
  umma::fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      umma::mbar_init(FULL_Q(s), GROUP_WARPS);
      umma::mbar_init(FULL_W(s), 1);
      umma::mbar_init(EMPTY(s), 1);
    }
    for (int q = 0; q < 2; ++q) { umma::mbar_init(TFULL(q), 1); umma::mbar_init(TEMPTY(q), EPI_WARPS); }
    umma::fence_barrier_init();
  }
  if (warp == WARP_MMA) umma::tmem_alloc(umma::smem_u32(tmem_slot), 64);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= WARP_PROD0) {
    // ===== Q producers: group = stage; a thread owns (frame fl, 4 consecutive vertices) of its group's K blocks =====
    const int pw = warp - WARP_PROD0, grp = pw / GROUP_WARPS, wg = pw % GROUP_WARPS;
    const int fl = wg * 4 + (lane >> 3), c4 = lane & 7;
    const bool fl_ok = fl < FT;
    float g[12], p[12], gn[12], pn[12];
    auto load = [&](uint32_t it, float (&gg)[12], float (&pp)[12]) {
#pragma unroll
      for (int i = 0; i < 12; ++i) { gg[i] = 0.f; pp[i] = 0.f; }
      if (it >= total_kb || !fl_ok) return;
      const int w = blockIdx.x + (int)(it / a.kb_per_split) * gridDim.x;
      const int ft = w / a.splits, split = w % a.splits;
      const int kb = split * a.kb_per_split + (int)(it % a.kb_per_split);
      const int f = ft * FT + fl, v0 = kb * BK + c4 * 4;
      if (f >= a.F) return;
      const float4* vp4 = reinterpret_cast<const float4*>(a.vposed + (size_t)f * a.NP + (size_t)v0 * 3);
      const float4 p0 = __ldg(vp4), p1 = __ldg(vp4 + 1), p2 = __ldg(vp4 + 2);
      pp[0] = p0.x; pp[1] = p0.y; pp[2] = p0.z; pp[3] = p0.w; pp[4] = p1.x; pp[5] = p1.y;
      pp[6] = p1.z; pp[7] = p1.w; pp[8] = p2.x; pp[9] = p2.y; pp[10] = p2.z; pp[11] = p2.w;
      if (a.g_vertices) {
        const float* gp = a.g_vertices + ((size_t)f * a.V + v0) * 3;
        const int nvalid = min(12, (a.V - v0) * 3);
#pragma unroll
        for (int i = 0; i < 12; ++i)
          if (i < nvalid) gg[i] = __ldg(gp + i);
      }
      if (a.g_joints) {
        const uint32_t fl4 = *reinterpret_cast<const uint32_t*>(a.vflag + v0);
        if (fl4) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if ((fl4 >> (8 * i)) & 0xff)
              for (int e = 0; e < a.NE; ++e)
                if (a.extra_ids[e] == v0 + i) {
                  const float* q = a.g_joints + ((size_t)f * (a.J + a.NE) + a.J + e) * 3;
                  gg[i * 3] += q[0]; gg[i * 3 + 1] += q[1]; gg[i * 3 + 2] += q[2];
                }
        }
      }
    };
    load(grp, g, p);
    for (uint32_t it = grp; it < total_kb; it += STAGES) {
      load(it + STAGES, gn, pn);
      umma::mbar_wait(EMPTY(grp), ((it / STAGES) & 1) ^ 1);
      if (fl_ok) {
        uint8_t* q_hi = sm + grp * STAGE_BYTES;
        uint8_t* q_lo = q_hi + Q_BYTES;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int row = fl * 12 + r * 4 + c;
            float4 q;   // the 4 vertices of this thread = 4 consecutive K elements
            q.x = c < 3 ? g[r] * p[c] : g[r];
            q.y = c < 3 ? g[3 + r] * p[3 + c] : g[3 + r];
            q.z = c < 3 ? g[6 + r] * p[6 + c] : g[6 + r];
            q.w = c < 3 ? g[9 + r] * p[9 + c] : g[9 + r];
            float4 h, l;
            umma::split_tf32(q.x, h.x, l.x);
            umma::split_tf32(q.y, h.y, l.y);
            umma::split_tf32(q.z, h.z, l.z);
            umma::split_tf32(q.w, h.w, l.w);
            const int off = row * 128 + ((c4 ^ (row & 7)) << 4);
            *reinterpret_cast<float4*>(q_hi + off) = h;
            *reinterpret_cast<float4*>(q_lo + off) = l;
          }
      }
      umma::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(FULL_Q(grp));
#pragma unroll
      for (int i = 0; i < 12; ++i) { g[i] = gn[i]; p[i] = pn[i]; }
    }
  } else if (warp == WARP_LOAD) {
    if (lane == 0) {
      for (uint32_t it = 0; it < total_kb; ++it) {
        const int s = it % STAGES;
        const int w = blockIdx.x + (int)(it / a.kb_per_split) * gridDim.x;
        const int kb = (w % a.splits) * a.kb_per_split + (int)(it % a.kb_per_split);
        umma::mbar_wait(EMPTY(s), ((it / STAGES) & 1) ^ 1);
        umma::mbar_arrive_expect_tx(FULL_W(s), 2 * WT_BYTES);
        umma::bulk_g2s(base + s * STAGE_BYTES + 2 * Q_BYTES, a.WTimg + (size_t)kb * 2 * NJ * BK, 2 * WT_BYTES, FULL_W(s));
      }
    }
  } else if (warp == WARP_MMA) {
    if (lane == 0) {
      const uint32_t idesc = umma::idesc_tf32_m128(NJ);
      uint32_t it = 0;
      for (int item = 0; item < my_items; ++item) {
        const uint32_t acc = item & 1;
        umma::mbar_wait(TEMPTY(acc), ((item >> 1) & 1) ^ 1);
        umma::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * NJ;
        for (int kl = 0; kl < a.kb_per_split; ++kl, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          umma::mbar_wait(FULL_Q(s), ph);
          umma::mbar_wait(FULL_W(s), ph);
          umma::tc_fence_after();
          const uint32_t q_hi = base + s * STAGE_BYTES, q_lo = q_hi + Q_BYTES;
          const uint32_t w_hi = q_hi + 2 * Q_BYTES, w_lo = w_hi + WT_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            const uint64_t dqh = umma::smem_desc_sw128(q_hi + k * 32), dql = umma::smem_desc_sw128(q_lo + k * 32);
            const uint64_t dwh = umma::smem_desc_sw128(w_hi + k * 32), dwl = umma::smem_desc_sw128(w_lo + k * 32);
            umma::mma_tf32(d_tmem, dqh, dwh, idesc, (kl > 0 || k > 0) ? 1u : 0u);
            umma::mma_tf32(d_tmem, dql, dwh, idesc, 1u);
            umma::mma_tf32(d_tmem, dqh, dwl, idesc, 1u);
          }
          umma::mma_commit(EMPTY(s));
        }
        umma::mma_commit(TFULL(acc));
      }
    }
  } else {
    // ===== epilogue: lane = row (fl, e); 32 columns = joints (+ ones column) =====
    const int row = warp * 32 + lane, fl = row / 12, e = row % 12;
    for (int item = 0; item < my_items; ++item) {
      const int w = blockIdx.x + item * gridDim.x;
      const int ft = w / a.splits, split = w % a.splits;
      const uint32_t acc = item & 1;
      umma::mbar_wait(TFULL(acc), (item >> 1) & 1);
      umma::tc_fence_after();
      uint32_t r[32];
      umma::tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + acc * NJ, r);
      umma::tmem_ld_wait();
      umma::tc_fence_before();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(TEMPTY(acc));
      const int f = ft * FT + fl;
      if (row < FT * 12 && f < a.F) {
        float* dst = a.gA + ((size_t)split * a.F + f) * UUO_GA_STRIDE;
#pragma unroll
        for (int j = 0; j < UUO_SKIN_J; ++j) dst[j * 12 + e] = __uint_as_float(r[j]);
        if ((e & 3) == 3) dst[288 + (e >> 2)] = __uint_as_float(r[UUO_SKIN_J]);
        if (e == 0) dst[291] = 0.f;
      }
    }
  }

  umma::tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    __syncwarp();
    umma::tmem_dealloc(tmem_base, 64);
  }
}

}  // namespace

int uuo_bwdA_splits(const uuo_model* m) {
  const int kbT = m->NP / 3 / BK;
  for (int s = UUO_NSPLIT_GA; s > 1; --s)
    if (kbT % s == 0) return s;
  return 1;
}

int launch_skin_bwdA_tc(const uuo_model* m, int F, const float* g_vertices, const float* g_joints, const uuo_ws& ws,
                        cudaStream_t st) {
  static bool attr_set = false;
  static int num_sms = 148;
  if (!attr_set) {
    UUO_CUDA(cudaFuncSetAttribute(skin_bwdA_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    int dev = 0;
    UUO_CUDA(cudaGetDevice(&dev));
    UUO_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    attr_set = true;
  }
  BwdAArgs a;
  a.WTimg = m->WTimg; a.vposed = ws.vposed; a.g_vertices = g_vertices; a.g_joints = g_joints;
  a.extra_ids = m->extra_ids; a.vflag = m->vflag; a.gA = ws.gA;
  a.F = F; a.V = m->V; a.NP = m->NP; a.NE = m->NE; a.J = m->J;
  a.n_ftiles = (F + FT - 1) / FT;
  a.splits = uuo_bwdA_splits(m);
  a.kb_per_split = m->NP / 3 / BK / a.splits;
  const int items = a.n_ftiles * a.splits;
  skin_bwdA_tc_kernel<<<std::min(items, num_sms), NTHREADS, SMEM_BYTES, st>>>(a);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}
