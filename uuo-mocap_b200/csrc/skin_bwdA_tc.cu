// d(A) of linear blend skinning on the tensor cores:
//   dA[f,j,(r,c)] = sum_v W[v,j] g[f,v,r] [v_posed[f,v]; 1][c]        (adjoint of lbs.py:237-250)
//   dtransl-like row  gsum[f,r] = sum_v g[f,v,r]                      (a 25th all-ones "joint" row)
// with g = g_vertices (the extra-joint rows of g_joints are added by extras_backward, skin.cu, as one more slab)
// as the GEMM  D[j, (f,e)] = W'[j, v] . Q[(f,e), v]^T  contracted over the vertices (K, blocks of 32):
//   A operand (M = 64): rows 0..31 = tf32 hi of [W | 1 | 0]^T, rows 32..63 = its lo part (one constant 8 KB image per
//   K block, bulk-copied);  B operand (N = 240 = 20 frames x 12 entries): Q = g (x) [v_posed;1], which never exists in
//   HBM: 15 producer warps (3 groups of 5, group = pipeline stage) read g and v_posed once, form the 12 products per
//   (frame, vertex), split them hi/lo and store the SWIZZLE_128B K-major image straight into shared memory.
// Two MMAs per k-step (B = Q_hi, B = Q_lo) give all four hi/lo cross terms: accumulator rows 0..31 hold W_hi.Q and
// rows 32..63 hold W_lo.Q; they are written as two partial slabs.  The vertex range is split into uuo_bwdA_splits()
// chunks, so fk_backward sums 2 x splits deterministic slabs.
// Measured on B200: an SS-mode MMA costs ~max(128 rows of A, M.N/256) cycles, so the frame axis (large) sits on N.
#include <algorithm>

#include "common.cuh"
#include "umma.cuh"

namespace {

constexpr int FT = 20;                   // frames per tile
constexpr int NQ = FT * 12;              // 240 MMA columns
constexpr int BK = 32;                   // vertices per K block
constexpr int Q_BYTES = NQ * BK * 4;     // 30 KB per half
constexpr int WW_BYTES = 64 * BK * 4;    // 8 KB: [hi rows 0..31 | lo rows 32..63]
constexpr int STAGE_BYTES = 2 * Q_BYTES + WW_BYTES + 1024;   // 69 KB (keeps every tile 1024 B aligned)
constexpr int STAGES = 3;
constexpr int EPI_WARPS = 4, GROUP_WARPS = 5;
constexpr int WARP_MMA = EPI_WARPS, WARP_LOAD = EPI_WARPS + 1, WARP_PROD0 = EPI_WARPS + 2;
constexpr int NTHREADS = (EPI_WARPS + 2 + STAGES * GROUP_WARPS) * 32;   // 672
constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + 256;

struct BwdAArgs {
  const float* WTimg;      // [kb_total][2][32*32]
  const float* vposed;     // [F][NP]
  const float* g_vertices; // [F][V][3] or null
  const float* g_joints;   // [F][J+NE][3] or null
  const int32_t* extra_ids;
  const uint8_t* vflag;    // [NP/3]
  float* gA;               // [2*splits][F][UUO_GA_STRIDE]
  int F, V, NP, NE, J, n_ftiles, splits, kb_per_split;
};

// kind::tf32, fp32 accumulate, K-major operands, M = 64
__device__ __forceinline__ uint32_t idesc_tf32_m64(uint32_t n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((n >> 3) << 17) | ((64u >> 4) << 24);
}

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

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      umma::mbar_init(FULL_Q(s), GROUP_WARPS);
      umma::mbar_init(FULL_W(s), 1);
      umma::mbar_init(EMPTY(s), 1);
    }
    for (int q = 0; q < 2; ++q) { umma::mbar_init(TFULL(q), 1); umma::mbar_init(TEMPTY(q), EPI_WARPS); }
    umma::fence_barrier_init();
  }
  if (warp == WARP_MMA) umma::tmem_alloc(umma::smem_u32(tmem_slot), 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= WARP_PROD0) {
    // ===== Q producers: group = stage; a thread owns (frame fl, 4 consecutive vertices) of its group's K blocks =====
    const int pw = warp - WARP_PROD0, grp = pw / GROUP_WARPS, wg = pw % GROUP_WARPS;
    const int fl = wg * 4 + (lane >> 3), c4 = lane & 7;
    const bool g_vec2 = (a.V & 1) == 0 && (reinterpret_cast<uintptr_t>(a.g_vertices) & 7) == 0;   // 8 B aligned rows -> float2 loads
    // position of a K block in this CTA's sequence: (item index, block within the split), advanced STAGES at a time
    struct Pos { int item, kl, f, v0; bool live; };
    auto resolve = [&](Pos& q) {
      q.live = q.item < my_items;
      if (!q.live) return;
      const int w = blockIdx.x + q.item * gridDim.x;
      q.f = (w / a.splits) * FT + fl;
      q.v0 = ((w % a.splits) * a.kb_per_split + q.kl) * BK + c4 * 4;
    };
    auto advance = [&](Pos& q) {
      q.kl += STAGES;
      bool moved = false;
      while (q.kl >= a.kb_per_split) { q.kl -= a.kb_per_split; ++q.item; moved = true; }
      if (moved) resolve(q); else q.v0 += STAGES * BK;
    };
    auto load = [&](const Pos& q, float (&gg)[12], float (&pp)[12], uint32_t& ff) {
      ff = 0u;
#pragma unroll
      for (int i = 0; i < 12; ++i) { gg[i] = 0.f; pp[i] = 0.f; }
      if (!q.live || q.f >= a.F) return;
      const float4* vp4 = reinterpret_cast<const float4*>(a.vposed + (size_t)q.f * a.NP + (size_t)q.v0 * 3);
      const float4 p0 = __ldg(vp4), p1 = __ldg(vp4 + 1), p2 = __ldg(vp4 + 2);
      pp[0] = p0.x; pp[1] = p0.y; pp[2] = p0.z; pp[3] = p0.w; pp[4] = p1.x; pp[5] = p1.y;
      pp[6] = p1.z; pp[7] = p1.w; pp[8] = p2.x; pp[9] = p2.y; pp[10] = p2.z; pp[11] = p2.w;
      if (a.g_vertices) {
        const float* gp = a.g_vertices + ((size_t)q.f * a.V + q.v0) * 3;
        const int nvalid = min(12, (a.V - q.v0) * 3);
        if (g_vec2) {
#pragma unroll
          for (int i = 0; i < 6; ++i)
            if (2 * i < nvalid) {
              const float2 t = __ldg(reinterpret_cast<const float2*>(gp) + i);
              gg[2 * i] = t.x; gg[2 * i + 1] = t.y;
            }
        } else {
#pragma unroll
          for (int i = 0; i < 12; ++i)
            if (i < nvalid) gg[i] = __ldg(gp + i);
        }
      }
    };
    // One K block of this group: prefetch the group's next block into (gn, pn), then form Q from (g, p).
    // The loop below alternates two register sets (no copies), so the loads of the next block stay in flight
    // across the whole compute phase instead of being waited on at a register move.
    Pos cur{0, grp, 0, 0, false}, nxt;
    uint32_t n = 0;
    auto body = [&](float (&g)[12], float (&p)[12], float (&gn)[12], float (&pn)[12]) {
      uint32_t unused;
      nxt = cur;
      advance(nxt);
      load(nxt, gn, pn, unused);
      umma::mbar_wait(EMPTY(grp), (n & 1) ^ 1);
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
      umma::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(FULL_Q(grp));
      cur = nxt;
      ++n;
    };
    float gA[12], pA[12], gB[12], pB[12];
    while (cur.kl >= a.kb_per_split) { cur.kl -= a.kb_per_split; ++cur.item; }
    resolve(cur);
    { uint32_t unused; load(cur, gA, pA, unused); }
    while (cur.live) {
      body(gA, pA, gB, pB);
      if (!cur.live) break;
      body(gB, pB, gA, pA);
    }
  } else if (warp == WARP_LOAD) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = 0; item < my_items; ++item) {
        const int w = blockIdx.x + item * gridDim.x;
        const int kb0 = (w % a.splits) * a.kb_per_split;
        for (int kl = 0; kl < a.kb_per_split; ++kl, ++it) {
          const int s = it % STAGES;
          umma::mbar_wait(EMPTY(s), ((it / STAGES) & 1) ^ 1);
          umma::mbar_arrive_expect_tx(FULL_W(s), WW_BYTES);
          umma::bulk_g2s(base + s * STAGE_BYTES + 2 * Q_BYTES, a.WTimg + (size_t)(kb0 + kl) * 2 * 32 * BK, WW_BYTES, FULL_W(s));
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // the warp stays converged; one elected lane issues
    const uint32_t idesc = idesc_tf32_m64(NQ);
    uint32_t it = 0;
    for (int item = 0; item < my_items; ++item) {
      const uint32_t acc = item & 1;
      umma::mbar_wait(TEMPTY(acc), ((item >> 1) & 1) ^ 1);
      const uint32_t d_tmem = tmem_base + acc * 256;
      for (int kl = 0; kl < a.kb_per_split; ++kl, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        umma::mbar_wait(FULL_Q(s), ph);
        umma::mbar_wait(FULL_W(s), ph);
        umma::tc_fence_after();
        if (umma::elect_one()) {
          const uint32_t q_hi = base + s * STAGE_BYTES, q_lo = q_hi + Q_BYTES, ww = q_hi + 2 * Q_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            const uint64_t dw = umma::smem_desc_sw128(ww + k * 32);
            umma::mma_tf32(d_tmem, dw, umma::smem_desc_sw128(q_hi + k * 32), idesc, (kl > 0 || k > 0) ? 1u : 0u);
            umma::mma_tf32(d_tmem, dw, umma::smem_desc_sw128(q_lo + k * 32), idesc, 1u);
          }
          umma::mma_commit(EMPTY(s));
          if (kl + 1 == a.kb_per_split) umma::mma_commit(TFULL(acc));
        }
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue.  M = 64 accumulator rows live in TMEM lanes 32*(row/16) + row%16:
    //   warp 0 lanes 0..15 = W_hi rows j 0..15, warp 1 = j 16..31, warp 2 / 3 = the W_lo rows of the same joints =====
    const int j = (warp & 1) * 16 + lane, part = warp >> 1;
    const bool j_ok = lane < 16 && j <= UUO_SKIN_J;
    for (int item = 0; item < my_items; ++item) {
      const int w = blockIdx.x + item * gridDim.x;
      const int ft = w / a.splits, split = w % a.splits;
      const uint32_t acc = item & 1;
      umma::mbar_wait(TFULL(acc), (item >> 1) & 1);
      umma::tc_fence_after();
      float* slab = a.gA + (size_t)(split * 2 + part) * a.F * UUO_GA_STRIDE;
#pragma unroll 1
      for (int st = 0; st < FT / 4; ++st) {
        uint32_t r[48];
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + acc * 256 + st * 48;
        umma::tmem_ld16(taddr, r);
        umma::tmem_ld16(taddr + 16, r + 16);
        umma::tmem_ld16(taddr + 32, r + 32);
        umma::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int f = ft * FT + st * 4 + k;
          if (!j_ok || f >= a.F) continue;
          float* dst = slab + (size_t)f * UUO_GA_STRIDE;
          if (j < UUO_SKIN_J) {
            float4* d4 = reinterpret_cast<float4*>(dst + j * 12);
#pragma unroll
            for (int q = 0; q < 3; ++q)
              d4[q] = make_float4(__uint_as_float(r[k * 12 + q * 4]), __uint_as_float(r[k * 12 + q * 4 + 1]),
                                  __uint_as_float(r[k * 12 + q * 4 + 2]), __uint_as_float(r[k * 12 + q * 4 + 3]));
          } else {   // the ones row: sum_v g[f,v,r] sits in entries (r,3)
            *reinterpret_cast<float4*>(dst + 288) = make_float4(__uint_as_float(r[k * 12 + 3]), __uint_as_float(r[k * 12 + 7]),
                                                                __uint_as_float(r[k * 12 + 11]), 0.f);
          }
        }
      }
      umma::tc_fence_before();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(TEMPTY(acc));
    }
  }

  umma::tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    __syncwarp();
    umma::tmem_dealloc(tmem_base, 512);
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
  UUO_DYN_SMEM(skin_bwdA_tc_kernel, SMEM_BYTES);
  const int num_sms = uuo_sm_count();
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
