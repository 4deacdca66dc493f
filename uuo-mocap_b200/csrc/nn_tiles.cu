// Marker -> vertex correspondence, pruned (same results as nn_argmin_kernel in nn.cu, bit for bit).
//
// The reference's knn_points(K=1) (src/video_mocap/losses/chamfer_distance.py:5-21 -> pytorch3d) scans all
// V vertices for every marker: 45 x 6890 x ~12 fp32 instructions per frame, 8x above what the 82.7 KB of
// vertices cost to stream from HBM.  Here the vertices are visited through a model-constant permutation
// `order` (uuo_model_vertex_order: leaves of a balanced kd-tree over the template) so that every 32
// consecutive entries -- a tile, one per lane -- stay spatially close under any pose, and every 32
// consecutive tiles -- a round -- form a larger coherent region.  One persistent CTA per SM, one frame at a time:
//   1. the frame's vertices are copied once into shared memory (cp.async, 16 B body + 4 B head/tail); the copy
//      of frame f+1 is issued as soon as step 2 of frame f has drained the staging buffer, so it overlaps step 3,
//   2. one warp per tile permutes the tile into a float4 (x, y, z, id) array in visiting order and builds the
//      tile's axis-aligned box (redux.sync.min/max.f32); one warp per round builds the round's box,
//   3. one warp per marker computes the distance to the boxes WITH THE SAME fp32 OPERATION SEQUENCE AS THE
//      EXACT DISTANCE (sub, mul, add, all round-to-nearest, no FMA).  Rounding is monotone, so this value
//      never exceeds the exact distance of any vertex inside the box; a round / tile is evaluated only if
//      its bound is <= the best exact distance so far (ties included), starting from the closest box.
// The result is the lexicographic minimum of (distance, vertex id), i.e. the FIRST minimum of the
// reference's scan.  HBM traffic: the frame's vertices once (82 680 B) + markers + 8 B per marker out.
#include "common.cuh"
#include "umma.cuh"

__device__ __forceinline__ float nt_sqdist(float mx, float my, float mz, float vx, float vy, float vz) {
  const float dx = __fsub_rn(mx, vx), dy = __fsub_rn(my, vy), dz = __fsub_rn(mz, vz);
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
__device__ __forceinline__ float redux_min_f32(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float redux_max_f32(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ void cp_async4(void* smem, const void* g) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(g));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(g));
}

// smem carve-up: sorted float4 [ntp*32] | stage[3V + 8] floats | box[6][ntp] floats | rbox[6][32] floats | ord u16 [ntp*32]
struct nt_layout {
  int nr, ntp;
  size_t off_stage, off_box, off_rbox, off_ord, total;
};
__host__ __device__ inline nt_layout nt_make_layout(int V) {
  nt_layout L;
  const int ntiles = (V + 31) / 32;
  L.nr = (ntiles + 31) / 32;
  L.ntp = L.nr * 32;
  L.off_stage = (size_t)L.ntp * 32 * 16;
  L.off_box = L.off_stage + (size_t)((3 * V + 8 + 3) & ~3) * 4;
  L.off_rbox = L.off_box + (size_t)6 * L.ntp * 4;
  L.off_ord = L.off_rbox + 6 * 32 * 4;
  L.total = L.off_ord + (size_t)L.ntp * 32 * 2;
  return L;
}

// distance from the marker to box t of a [6][n] box table, with the exact distance's operation sequence
__device__ __forceinline__ float nt_boxdist(const float* __restrict__ b, int n, int t, float mx, float my, float mz) {
  // per axis one of fl(lo - m), fl(m - hi), 0: the rounded difference the exact distance takes for a vertex on that
  // face, and |fl(m - v)| of any vertex inside the box is >= it (rounding is monotone)
  const float dx = fmaxf(fmaxf(__fsub_rn(b[t], mx), __fsub_rn(mx, b[3 * n + t])), 0.f);
  const float dy = fmaxf(fmaxf(__fsub_rn(b[n + t], my), __fsub_rn(my, b[4 * n + t])), 0.f);
  const float dz = fmaxf(fmaxf(__fsub_rn(b[2 * n + t], mz), __fsub_rn(mz, b[5 * n + t])), 0.f);
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// lane-local lexicographic minimum of (distance, vertex id) over this lane's vertex of tile t
__device__ __forceinline__ void nt_eval(const float4* __restrict__ sorted, int t, int lane, float mx, float my, float mz,
                                        float& d_l, int& i_l) {
  const float4 p = sorted[t * 32 + lane];
  const float d = nt_sqdist(mx, my, mz, p.x, p.y, p.z);
  const int id = __float_as_int(p.w);
  if (d < d_l || (d == d_l && id < i_l)) { d_l = d; i_l = id; }
}

// warp-wide lexicographic minimum; result uniform
__device__ __forceinline__ void nt_reduce(float d_l, int i_l, float& best, int& besti) {
  best = redux_min_f32(d_l);
  besti = __reduce_min_sync(0xffffffffu, d_l == best ? i_l : 0x7fffffff);
}

__global__ void __launch_bounds__(1024, 1) nn_tiles_kernel(const float* __restrict__ markers,
                                                           const float* __restrict__ vertices,
                                                           const int32_t* __restrict__ order, int F, int M, int V,
                                                           float* __restrict__ dist2, int32_t* __restrict__ idx) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x, nwarp = nthr >> 5;
  const nt_layout L = nt_make_layout(V);
  const int nr = L.nr, ntp = L.ntp, nflt = 3 * V;
  float4* sorted = reinterpret_cast<float4*>(smem_raw);                     // [ntp*32] (x, y, z, id) in visiting order
  float* stage = reinterpret_cast<float*>(smem_raw + L.off_stage);          // the frame as it lies in HBM
  float* box = reinterpret_cast<float*>(smem_raw + L.off_box);              // [6][ntp]: lo xyz, hi xyz per tile
  float* rbox = reinterpret_cast<float*>(smem_raw + L.off_rbox);            // [6][32]: per round of 32 tiles
  uint16_t* ord = reinterpret_cast<uint16_t*>(smem_raw + L.off_ord);        // [ntp*32]; the tail repeats order[V-1]
  for (int i = tid; i < ntp * 32; i += nthr) ord[i] = (uint16_t)__ldg(order + min(i, V - 1));
  const float INF = __int_as_float(0x7f800000);

  // stage frame f.  Normal case: 1-D bulk copies (cp.async.bulk -> mbarrier) of the frame's 16 B-aligned hull -- deep
  // in flight without occupying L1 request slots (cp.async tops out near 6 B/cycle/SM here).  The hull of the first /
  // last frame may stick out of the tensor when its ends are not 16 B aligned: those frames use cp.async (16 B body +
  // 4 B head / tail).  Either way the smem offset keeps the source's alignment modulo 16 B.
  __shared__ __align__(8) uint64_t load_bar;
  const uint32_t bar = umma::smem_u32(&load_bar);
  if (tid == 0) { umma::mbar_init(bar, 1); umma::fence_barrier_init(); }
  __syncthreads();
  const uintptr_t vbase = reinterpret_cast<uintptr_t>(vertices), vend = vbase + (size_t)F * nflt * 4;
  auto hull_ok = [&](int f) {
    const uintptr_t a = vbase + (size_t)f * nflt * 4;
    return (a & ~(uintptr_t)15) >= vbase && ((a + (size_t)nflt * 4 + 15) & ~(uintptr_t)15) <= vend;
  };
  auto issue_load = [&](int f) {
    const float* src = vertices + (size_t)f * nflt;
    if (hull_ok(f)) {
      if (tid == 0) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(src), lo = a & ~(uintptr_t)15;
        const uint32_t bytes = (uint32_t)(((a + (size_t)nflt * 4 + 15) & ~(uintptr_t)15) - lo);
        umma::fence_proxy_async_smem();            // generic-proxy reads of the staging buffer are done (barrier above)
        umma::mbar_arrive_expect_tx(bar, bytes);
        const uint32_t chunk = ((bytes / 8) + 15) & ~15u;
        for (uint32_t off = 0; off < bytes; off += chunk)
          umma::bulk_g2s(umma::smem_u32(stage) + off, reinterpret_cast<const void*>(lo + off), min(chunk, bytes - off), bar);
      }
      return;
    }
    const int mis = (int)((reinterpret_cast<uintptr_t>(src) & 15) >> 2);
    float* raw = stage + mis;
    const int head = min(nflt, (4 - mis) & 3);
    const int body = (nflt - head) >> 2;
    const int tail0 = head + 4 * body;
    if (tid < head) cp_async4(raw + tid, src + tid);
    for (int i = tid; i < body; i += nthr) cp_async16(raw + head + 4 * i, src + head + 4 * i);
    if (tid < nflt - tail0) cp_async4(raw + tail0 + tid, src + tail0 + tid);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  uint32_t parity = 0;
  if ((int)blockIdx.x < F) issue_load(blockIdx.x);

  for (int f = blockIdx.x; f < F; f += gridDim.x) {
    if (hull_ok(f)) { umma::mbar_wait(bar, parity); parity ^= 1; }
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                   // frame f staged; every warp is done with the previous frame
    const float* raw = stage + (int)((reinterpret_cast<uintptr_t>(vertices + (size_t)f * nflt) & 15) >> 2);
    // ---- 2. permute into visiting order + tile boxes: one warp per tile, one lane per vertex (consecutive lanes
    // write consecutive float4: no bank conflicts on the store; the gather from the staged frame is the random part)
    for (int t = warp; t < ntp; t += nwarp) {
      const int e = t * 32 + lane;
      const int id = ord[e];
      const float* p = raw + 3 * id;
      const float x = p[0], y = p[1], z = p[2];
      sorted[e] = make_float4(x, y, z, __int_as_float(id));
      const float lx = redux_min_f32(x), ly = redux_min_f32(y), lz = redux_min_f32(z);
      const float hx = redux_max_f32(x), hy = redux_max_f32(y), hz = redux_max_f32(z);
      if (lane == 0) {
        box[t] = lx; box[ntp + t] = ly; box[2 * ntp + t] = lz;
        box[3 * ntp + t] = hx; box[4 * ntp + t] = hy; box[5 * ntp + t] = hz;
      }
    }
    __syncthreads();
    // the staging buffer is free again: the next frame streams in while the markers are matched
    if (f + (int)gridDim.x < F) issue_load(f + gridDim.x);
    // ---- 2b. round boxes (union of 32 consecutive tile boxes)
    for (int r = warp; r < nr; r += nwarp) {
      const int t = r * 32 + lane;
      const float lx = redux_min_f32(box[t]), ly = redux_min_f32(box[ntp + t]), lz = redux_min_f32(box[2 * ntp + t]);
      const float hx = redux_max_f32(box[3 * ntp + t]), hy = redux_max_f32(box[4 * ntp + t]),
                  hz = redux_max_f32(box[5 * ntp + t]);
      if (lane == 0) {
        rbox[r] = lx; rbox[32 + r] = ly; rbox[64 + r] = lz;
        rbox[96 + r] = hx; rbox[128 + r] = hy; rbox[160 + r] = hz;
      }
    }
    __syncthreads();
    // ---- 3. one warp per marker
    for (int m = warp; m < M; m += nwarp) {
      const float* mp = markers + ((size_t)f * M + m) * 3;
      const float mx = __ldg(mp), my = __ldg(mp + 1), mz = __ldg(mp + 2);
      const float rlb = lane < nr ? nt_boxdist(rbox, 32, lane, mx, my, mz) : INF;
      // closest round, closest tile in it: an exact distance that bounds everything else
      const float rmin = redux_min_f32(rlb);
      const unsigned bm = __ballot_sync(0xffffffffu, lane < nr && rlb == rmin);
      const int r0 = bm ? __ffs(bm) - 1 : 0;
      float lb = nt_boxdist(box, ntp, r0 * 32 + lane, mx, my, mz);
      const float tmin = redux_min_f32(lb);
      const unsigned bt = __ballot_sync(0xffffffffu, lb == tmin);
      const int j0 = bt ? __ffs(bt) - 1 : 0;
      float d_l = INF;
      int i_l = 0x7fffffff;
      nt_eval(sorted, r0 * 32 + j0, lane, mx, my, mz, d_l, i_l);
      float best;
      int besti;
      nt_reduce(d_l, i_l, best, besti);
      // every other tile whose bound does not exceed that distance (NaN bounds are evaluated, never pruned);
      // lanes keep their own running minimum, one warp reduction per round
      unsigned pend = __ballot_sync(0xffffffffu, lane < nr && !(rlb > best));
      int r = r0;
      unsigned b = __ballot_sync(0xffffffffu, !(lb > best)) & ~(1u << j0);
      pend &= ~(1u << r0);
      while (true) {
        const bool any = b != 0;
        while (b) {
          const int j = __ffs(b) - 1;
          b &= b - 1;
          nt_eval(sorted, r * 32 + j, lane, mx, my, mz, d_l, i_l);
        }
        if (any) nt_reduce(d_l, i_l, best, besti);
        // next pending round that still can hold something closer
        bool found = false;
        while (pend) {
          r = __ffs(pend) - 1;
          pend &= pend - 1;
          if (__shfl_sync(0xffffffffu, rlb, r) > best) continue;
          found = true;
          break;
        }
        if (!found) break;
        lb = nt_boxdist(box, ntp, r * 32 + lane, mx, my, mz);
        b = __ballot_sync(0xffffffffu, !(lb > best));
      }
      if (lane == 0) {
        // all-NaN distances leave (inf, none): the exhaustive scan of nn.cu reports (inf, 0) there
        dist2[(size_t)f * M + m] = best;
        idx[(size_t)f * M + m] = besti == 0x7fffffff ? 0 : besti;
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

extern "C" int uuo_nn_argmin_ordered(const float* markers, const float* vertices, const int32_t* order, int F, int M,
                                     int V, float* dist2, int32_t* idx, void* stream) {
  UUO_CHECK_ARG(F >= 0 && M >= 0 && V > 0, "bad sizes");
  if (F == 0 || M == 0) return UUO_OK;                     // empty inputs: nothing to read, pointers may be NULL
  UUO_CHECK_ARG(markers && vertices && dist2 && idx, "NULL pointer");
  const nt_layout L = nt_make_layout(V);
  const size_t smem = L.total;
  if (!order || V > 65535 || L.nr > 32 || smem > 227 * 1024)      // no order / does not fit: the exhaustive kernel
    return uuo_nn_argmin(markers, vertices, F, M, V, dist2, idx, stream);
  if (F == 0 || M == 0) return UUO_OK;
  UUO_DYN_SMEM(nn_tiles_kernel, smem);
  const int n_sm = uuo_sm_count();
  {
    // one persistent CTA per SM; warps = markers / passes so that every pass over the markers is (almost) full
    const int passes = (M + 31) / 32;
    int warps = (M + passes - 1) / passes;
    if (warps < 8) warps = 8;
    const int grid = F < n_sm ? F : n_sm;
    ProfScope ps_("nn_argmin_tiles", (cudaStream_t)stream);
    nn_tiles_kernel<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(markers, vertices, order, F, M, V, dist2, idx);
  }
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

// =====================================================================================================
// Presorted variant: the vertices already lie in visiting order (the fitting loop evaluates lbs() on a twin
// model whose vertices are renumbered by uuo_model_vertex_order), so a tile is 32 CONSECUTIVE vertices and
// nothing is gathered: the staged frame is read with stride-3 words (conflict-free) straight from the
// staging buffer.  ids[p] = original vertex id of position p (ties break on the ORIGINAL numbering).
//   - one persistent CTA per SM, 32 warps; frames double-buffered with 1-D bulk copies (cp.async.bulk ->
//     mbarrier), the next frame's markers / weights / hints prefetched with cp.async, so the HBM stream of
//     frame f+1 overlaps all of frame f;
//   - step 2: tile boxes, one lane per half tile; round boxes = union of 32 tile boxes (CREDUX);
//   - step 3: markers are handed out to warps one by one.  A warp evaluates a FIRST tile (the hinted one, else the
//     tile with the smallest bound) -> an exact distance best0 that bounds everything else; then every other tile
//     whose box bound does not exceed best0 is evaluated by the same warp, one vertex per lane, each lane keeping
//     its own lexicographic minimum of (distance, original id); ONE warp reduction per marker at the end gives
//     the reference's first minimum.  (r1g kept a shared (marker, tile) pair list drained by all warps: with a
//     hint only ~4 tiles per marker survive and the list's atomics, fences and waiting cost more than they saved.)
// =====================================================================================================
// per-phase cycle counters of CTA 0 (tools/nn_phases.py): compile with -DNP_TIMING
#ifdef NP_TIMING
__device__ unsigned long long np_phase_cycles[8];
#define NP_TICK(i)                                                     \
  do {                                                                 \
    if (blockIdx.x == 0 && threadIdx.x == 0) {                         \
      const long long now_ = clock64();                                \
      np_phase_cycles[i] += (unsigned long long)(now_ - tick_);        \
      tick_ = now_;                                                    \
    }                                                                  \
  } while (0)
extern "C" int uuo_debug_nn_phases(unsigned long long* out, int reset) {
  if (out) cudaMemcpyFromSymbol(out, np_phase_cycles, sizeof(np_phase_cycles));
  if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(np_phase_cycles, z, sizeof(z)); }
  return 0;
}
#else
#define NP_TICK(i) do { } while (0)
#endif

#define NP_THREADS 1024
#define NP_GROUP 512         // threads per half-CTA
#define NP_MAXM 256          // markers per pass over a frame

struct np_layout {
  int nr, ntp, stage_floats;
  size_t off_box, off_rbox, off_mk, off_mw, off_mh, off_ids, total;
};
__host__ __device__ inline np_layout np_make_layout(int V) {
  np_layout L;
  const int ntiles = (V + 31) / 32;
  L.nr = (ntiles + 31) / 32;
  L.ntp = L.nr * 32;
  L.stage_floats = (3 * V + 8 + 3) & ~3;
  L.off_box = (size_t)2 * L.stage_floats * 4;
  L.off_rbox = L.off_box + (size_t)2 * 6 * L.ntp * 4;  // boxes, round boxes, markers: one set per half-CTA
  L.off_mk = L.off_rbox + 2 * 6 * 32 * 4;
  L.off_mw = L.off_mk + 2 * NP_MAXM * 3 * 4;
  L.off_mh = L.off_mw + 2 * NP_MAXM * 4;
  L.off_ids = L.off_mh + 2 * NP_MAXM * 4;
  L.total = L.off_ids + (size_t)L.ntp * 32 * 2;
  return L;
}

__global__ void __launch_bounds__(NP_THREADS, 1) nn_presorted_kernel(const float* __restrict__ markers,
                                                                     const float* __restrict__ vertices,
                                                                     const int32_t* __restrict__ ids,
                                                                     const float* __restrict__ wts,
                                                                     const int32_t* hint, int F, int M,
                                                                     int V, float* __restrict__ dist2,
                                                                     int32_t* __restrict__ idx,
                                                                     int32_t* pos_out) {   // may alias hint
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // two half-CTAs of 16 warps work on ALTERNATE frames (group g owns staging buffer g, box set g, marker buffers g and
  // named barrier 1 + g): one group's boxes, barrier tails and load wait overlap the other group's markers
  const int tid = threadIdx.x, lane = tid & 31, g = tid >> 9, gt = tid & (NP_GROUP - 1), warp = gt >> 5, nthr = NP_GROUP,
            nwarp = NP_GROUP / 32;
  auto gbar = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "n"(NP_GROUP) : "memory"); };
  const np_layout L = np_make_layout(V);
  const int nr = L.nr, ntp = L.ntp, nflt = 3 * V;
  float* stage0 = reinterpret_cast<float*>(smem_raw);
  float* box = reinterpret_cast<float*>(smem_raw + L.off_box) + (size_t)g * 6 * L.ntp;
  float* rbox = reinterpret_cast<float*>(smem_raw + L.off_rbox) + g * 6 * 32;
  float* mk0 = reinterpret_cast<float*>(smem_raw + L.off_mk);
  float* mw0 = reinterpret_cast<float*>(smem_raw + L.off_mw);      // marker weights: a marker of weight 0 is not searched
  int* mh0 = reinterpret_cast<int*>(smem_raw + L.off_mh);          // hinted row per marker
  unsigned short* ids_s = reinterpret_cast<unsigned short*>(smem_raw + L.off_ids);   // original id of every row
  for (int i = tid; i < ntp * 32; i += NP_THREADS) {
    const int p = min(i, V - 1);
    ids_s[i] = (unsigned short)(ids ? __ldg(ids + p) : p);
  }
  __shared__ __align__(8) uint64_t load_bar[2];
  __shared__ unsigned marker_next[2];  // markers handed out to the group's warps so far
  const float INF = __int_as_float(0x7f800000);
  if (tid == 0) {
    umma::mbar_init(umma::smem_u32(&load_bar[0]), 1);
    umma::mbar_init(umma::smem_u32(&load_bar[1]), 1);
    umma::fence_barrier_init();
  }
  __syncthreads();
  const uintptr_t vbase = reinterpret_cast<uintptr_t>(vertices), vend = vbase + (size_t)F * nflt * 4;
  // can frame f's 16 B hull be copied in bulk (does it stay inside the tensor)?  With frames of >= 16 B only the first
  // and the last frame can fail, so the per-frame test is two compares (it ran on every thread twice per frame)
  const bool big = (size_t)nflt * 4 >= 16;
  const bool first_ok = (vbase & 15) == 0, last_ok = (vend & 15) == 0;
  auto hull_ok = [&](int f) {
    if (big) return (f > 0 || first_ok) && (f < F - 1 || last_ok);
    const uintptr_t a = vbase + (size_t)f * nflt * 4;
    return (a & ~(uintptr_t)15) >= vbase && ((a + (size_t)nflt * 4 + 15) & ~(uintptr_t)15) <= vend;
  };
  // stage frame f into buffer b (see nn_tiles_kernel: bulk copy of the 16 B hull, cp.async for out-of-tensor hulls)
  // and, with cp.async, the first NP_MAXM markers of that frame with their weights and hints
  auto issue_load = [&](int f, int b) {
    {
      const int mc = min(NP_MAXM, M);
      for (int i = gt; i < mc * 3; i += nthr) cp_async4(mk0 + b * NP_MAXM * 3 + i, markers + (size_t)f * M * 3 + i);
      if (wts)
        for (int i = gt; i < mc; i += nthr) cp_async4(mw0 + b * NP_MAXM + i, wts + (size_t)f * M + i);
      if (hint)
        for (int i = gt; i < mc; i += nthr) cp_async4(mh0 + b * NP_MAXM + i, hint + (size_t)f * M + i);
    }
    const float* src = vertices + (size_t)f * nflt;
    float* stage = stage0 + (size_t)b * L.stage_floats;
    if (hull_ok(f)) {
      if (gt == NP_GROUP - 32) {
        const uint32_t bar = umma::smem_u32(&load_bar[b]);
        const uintptr_t a = reinterpret_cast<uintptr_t>(src), lo = a & ~(uintptr_t)15;
        const uint32_t bytes = (uint32_t)(((a + (size_t)nflt * 4 + 15) & ~(uintptr_t)15) - lo);
        umma::fence_proxy_async_smem();
        umma::mbar_arrive_expect_tx(bar, bytes);
        const uint32_t chunk = ((bytes / 8) + 15) & ~15u;
        for (uint32_t off = 0; off < bytes; off += chunk)
          umma::bulk_g2s(umma::smem_u32(stage) + off, reinterpret_cast<const void*>(lo + off), min(chunk, bytes - off), bar);
      }
    } else {
      const int mis = (int)((reinterpret_cast<uintptr_t>(src) & 15) >> 2);
      float* raw = stage + mis;
      const int head = min(nflt, (4 - mis) & 3);
      const int body = (nflt - head) >> 2;
      const int tail0 = head + 4 * body;
      if (gt < head) cp_async4(raw + gt, src + gt);
      for (int i = gt; i < body; i += nthr) cp_async16(raw + head + 4 * i, src + head + 4 * i);
      if (gt < nflt - tail0) cp_async4(raw + tail0 + gt, src + tail0 + gt);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  uint32_t parity = 0;
  const int fstep = 2 * (int)gridDim.x, f0 = (int)blockIdx.x + g * (int)gridDim.x, buf = g;
  if (f0 < F) issue_load(f0, buf);

#ifdef NP_TIMING
  long long tick_ = clock64();
#endif
  for (int f = f0; f < F; f += fstep) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ONE warp of the half polls the copy's mbarrier, the others sleep in the named barrier behind it (16 spinning
    // warps took ~14 % of the kernel's issue slots away from the other half's markers)
    if (hull_ok(f)) {
      if (warp == 0) umma::mbar_wait(umma::smem_u32(&load_bar[buf]), parity);
      parity ^= 1;
    }
    gbar();
    NP_TICK(0);   // wait for the frame
    if (gt == 0) marker_next[g] = 0;
    const float* raw = stage0 + (size_t)buf * L.stage_floats +
                       (int)((reinterpret_cast<uintptr_t>(vertices + (size_t)f * nflt) & 15) >> 2);
    float* mk = mk0 + buf * NP_MAXM * 3;
    float* mw = mw0 + buf * NP_MAXM;
    int* mh = mh0 + buf * NP_MAXM;
    // ---- 2. tile boxes: one lane per HALF tile (16 vertices walked serially), 16 tiles per warp.  Lane l starts its
    // walk at vertex (l & 15): at every step the 16 lanes of a half read 16 different residues of 3j mod 32 and the
    // other half sits 48 words = 16 banks further, so the stride-96-word tiles never collide on a bank.
    for (int tb = warp * 16; tb < ntp; tb += nwarp * 16) {
      const int t = tb + (lane & 15), rot = lane & 15;
      const int p0 = t * 32 + (lane >> 4) * 16;
      float lx = INF, ly = INF, lz = INF, hx = -INF, hy = -INF, hz = -INF;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float* p = raw + 3 * min(p0 + ((i + rot) & 15), V - 1);     // the tail repeats the last vertex
        const float x = p[0], y = p[1], z = p[2];
        lx = fminf(lx, x); hx = fmaxf(hx, x);
        ly = fminf(ly, y); hy = fmaxf(hy, y);
        lz = fminf(lz, z); hz = fmaxf(hz, z);
      }
      lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, 16)); hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, 16));
      ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, 16)); hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, 16));
      lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, 16)); hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, 16));
      if (lane < 16) {
        box[t] = lx; box[ntp + t] = ly; box[2 * ntp + t] = lz;
        box[3 * ntp + t] = hx; box[4 * ntp + t] = hy; box[5 * ntp + t] = hz;
      }
    }
    gbar();
    NP_TICK(1);   // tile boxes
    for (int r = warp; r < nr; r += nwarp) {
      const int t = r * 32 + lane;
      const float lx = redux_min_f32(box[t]), ly = redux_min_f32(box[ntp + t]), lz = redux_min_f32(box[2 * ntp + t]);
      const float hx = redux_max_f32(box[3 * ntp + t]), hy = redux_max_f32(box[4 * ntp + t]),
                  hz = redux_max_f32(box[5 * ntp + t]);
      if (lane == 0) {
        rbox[r] = lx; rbox[32 + r] = ly; rbox[64 + r] = lz;
        rbox[96 + r] = hx; rbox[128 + r] = hy; rbox[160 + r] = hz;
      }
    }
    for (int m0 = 0; m0 < M; m0 += NP_MAXM) {
      const int mc = min(NP_MAXM, M - m0);
      if (m0 > 0) {                                        // later chunks of a large marker set: fetched in place
        for (int i = gt; i < mc * 3; i += nthr) mk[i] = __ldg(markers + ((size_t)f * M + m0) * 3 + i);
        for (int i = gt; i < mc; i += nthr) {
          if (wts) mw[i] = __ldg(wts + (size_t)f * M + m0 + i);
          if (hint) mh[i] = hint[(size_t)f * M + m0 + i];   // a plain load: pos_out may be the same buffer
        }
        if (gt == 0) marker_next[g] = 0;
      }
      gbar();                                              // round boxes, markers, counter
      NP_TICK(2);
      // ---- 3. markers, handed out one by one (their costs differ a lot)
      for (;;) {
        int m = 0;
        if (lane == 0) m = (int)atomicAdd(&marker_next[g], 1u);
        m = __shfl_sync(0xffffffffu, m, 0);
        if (m >= mc) break;
        const size_t o = (size_t)f * M + m0 + m;
        if (wts && mw[m] == 0.0f) {            // weight 0: contributes nothing to the loss or its gradient
          if (lane == 0) {
            dist2[o] = 0.0f;
            idx[o] = ids_s[0];
            if (pos_out) pos_out[o] = 0;
          }
          continue;
        }
        const float mx = mk[3 * m], my = mk[3 * m + 1], mz = mk[3 * m + 2];
        const float rlb = lane < nr ? nt_boxdist(rbox, 32, lane, mx, my, mz) : INF;
        // the first tile: the hinted one (where the previous iteration of the fit found this marker's vertex: it
        // nearly always holds the nearest vertex again, so best0 is tight and few other tiles survive the bound),
        // else the tile with the smallest bound in the round with the smallest bound.  Which tile goes first never
        // changes the result, only how many tiles are left.
        int first = -1;
        if (hint) {
          const int h = mh[m];
          if (h >= 0 && h < V) first = h >> 5;
        }
        if (first < 0) {
          const float rmin = redux_min_f32(rlb);
          const unsigned bm = __ballot_sync(0xffffffffu, lane < nr && rlb == rmin);
          const int r0 = bm ? __ffs(bm) - 1 : 0;
          const float lb0 = nt_boxdist(box, ntp, r0 * 32 + lane, mx, my, mz);
          const float tmin = redux_min_f32(lb0);
          const unsigned bt = __ballot_sync(0xffffffffu, lb0 == tmin);
          first = r0 * 32 + (bt ? __ffs(bt) - 1 : 0);
        }
        // lane-local lexicographic minimum of (distance, original id << 16 | row); inf / NaN never win (the
        // exhaustive scan's strict '<' against an initial +inf never accepts them either)
        float bd = INF;
        unsigned bk = 0xffffffffu;
        {
          const int pos = min(first * 32 + lane, V - 1);
          const float* p = raw + 3 * pos;
          const float d = nt_sqdist(mx, my, mz, p[0], p[1], p[2]);
          if (d < INF) { bd = d; bk = ((unsigned)ids_s[pos] << 16) | (unsigned)pos; }
        }
        const float best0 = redux_min_f32(bd);
        unsigned pend = __ballot_sync(0xffffffffu, lane < nr && !(rlb > best0));     // NaN bounds are never pruned
        while (pend) {
          const int r = __ffs(pend) - 1;
          pend &= pend - 1;
          const float lb = nt_boxdist(box, ntp, r * 32 + lane, mx, my, mz);
          unsigned cb = __ballot_sync(0xffffffffu, !(lb > best0) && r * 32 + lane != first);
#ifdef NP_TIMING
          if (blockIdx.x == 0 && lane == 0) atomicAdd(&np_phase_cycles[6], (unsigned long long)__popc(cb));
#endif
          while (cb) {
            const int j = __ffs(cb) - 1;
            cb &= cb - 1;
            const int pos = min((r * 32 + j) * 32 + lane, V - 1);
            const float* p = raw + 3 * pos;
            const float d = nt_sqdist(mx, my, mz, p[0], p[1], p[2]);
            const unsigned k = ((unsigned)ids_s[pos] << 16) | (unsigned)pos;
            if (d < bd || (d == bd && k < bk && d < INF)) { bd = d; bk = k; }
          }
        }
        const float dm = redux_min_f32(bd);
        const unsigned kk = __reduce_min_sync(0xffffffffu, bd == dm ? bk : 0xffffffffu);
        if (lane == 0) {
          const bool none = kk == 0xffffffffu;             // no finite distance: (inf, 0) like the exhaustive scan
          dist2[o] = none ? INF : dm;
          idx[o] = none ? 0 : (int)(kk >> 16);
          if (pos_out) pos_out[o] = none ? 0 : (int)(kk & 0xffffu);
        }
      }
      NP_TICK(3);   // markers (this warp's)
      gbar();                                              // every warp is done with the markers / this frame's buffers
      NP_TICK(4);   // waiting for the other warps
    }
    // the group's buffers are free: stream its next frame in (the other group computes meanwhile)
    if (f + fstep < F) issue_load(f + fstep, buf);
    NP_TICK(5);
#ifdef NP_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 0) np_phase_cycles[7] += 1;
#endif
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

static int nn_presorted_launch(const float* markers, const float* vertices, const int32_t* ids, const float* weights,
                               const int32_t* hint, int F, int M, int V, float* dist2, int32_t* idx, int32_t* pos,
                               void* stream) {
  UUO_CHECK_ARG(F >= 0 && M >= 0 && V > 0, "bad sizes");
  const np_layout L = np_make_layout(V);
  UUO_CHECK_ARG(V < 65535 && L.nr <= 32 && L.total <= 227 * 1024 && M < 65536,
                "V too large for the presorted correspondence kernel (two staged frames must fit in shared memory)");
  if (F == 0 || M == 0) return UUO_OK;                     // empty inputs: nothing to read, pointers may be NULL
  UUO_CHECK_ARG(markers && vertices && dist2 && idx, "NULL pointer");
  UUO_DYN_SMEM(nn_presorted_kernel, L.total);
  const int n_sm = uuo_sm_count();
  {
    ProfScope ps_("nn_argmin_presorted", (cudaStream_t)stream);
    nn_presorted_kernel<<<F < n_sm ? F : n_sm, NP_THREADS, L.total, (cudaStream_t)stream>>>(markers, vertices, ids, weights,
                                                                                          hint, F, M, V, dist2, idx, pos);
  }
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_nn_argmin_presorted(const float* markers, const float* vertices, const int32_t* ids, int F, int M,
                                       int V, float* dist2, int32_t* idx, int32_t* pos, void* stream) {
  return nn_presorted_launch(markers, vertices, ids, nullptr, nullptr, F, M, V, dist2, idx, pos, stream);
}

extern "C" int uuo_nn_argmin_presorted_fit(const float* markers, const float* vertices, const int32_t* ids,
                                           const float* weights, const int32_t* hint, int F, int M, int V, float* dist2,
                                           int32_t* idx, int32_t* pos, void* stream) {
  return nn_presorted_launch(markers, vertices, ids, weights, hint, F, M, V, dist2, idx, pos, stream);
}
