// Linear blend skinning on the tensor cores (lbs.py:237-250), dense J = 24 weights.
//   T[f,v,e] = sum_j W[v,j] A[f,j,e]   (e = 12 entries of the 3x4 relative transform)
// is the GEMM  D[v, (f,e)] = W[v, j] . Aimg[(f,e), j]^T  with M = 128 vertices (TMEM lane = vertex), N = 16 frames x 12
// entries = 192 columns, K = 24 (three kind::tf32 k-steps), 3xTF32 split for fp32 accuracy.  The epilogue thread of
// lane v reads its 12 blended entries per frame straight from TMEM and applies them:
//   mode 0 (forward)      vertices[f,v] = T[:, :3] v_posed[f,v] + T[:, 3] + transl[f]
//   mode 1 (d v_posed)    g_vposed[f,v] = T[:, :3]^T g_vertices[f,v]   (extra-joint rows of g_joints: extras_backward, skin.cu)
// Operands: W is a model constant (pre-split, pre-swizzled image, resident in smem for the whole CTA); the transforms
// are packed once per call into the same image format by pack_transforms_kernel and fetched with one 48 KB bulk copy
// per frame tile.  Accumulators are double buffered in TMEM so MMAs of tile i+1 overlap the epilogue of tile i.
#include <algorithm>

#include "common.cuh"
#include "umma.cuh"

namespace {

constexpr int VT = 128;                 // vertices per tile (MMA M)
constexpr int NF = 16;                  // frames per tile
constexpr int NCOL = NF * 12;           // 192 accumulator columns (MMA N)
constexpr int W_BYTES = VT * 128;       // 16 KB per half (hi / lo)
constexpr int T_BYTES = NCOL * 128;     // 24 KB per half
constexpr int T_STAGES = 2;              // transform images (bulk copy -> MMA)
constexpr int X_STAGES = 3;              // per-vertex inputs of a tile (cp.async -> epilogue)
constexpr int XROW = 400;                 // smem slot of one input row: 384 B + up to 12 B alignment shift
constexpr int XW_BYTES = 8 * XROW + 128;  // per warp and stage: 8 frame rows (+ 8 transl rows)
constexpr int X_BYTES = 8 * XW_BYTES;     // all epilogue warps
constexpr int EPI_WARPS = 8;
constexpr int WARP_MMA = EPI_WARPS, WARP_LOAD = EPI_WARPS + 1;
constexpr int NTHREADS = (EPI_WARPS + 2) * 32;   // 320
constexpr int SMEM_BYTES = 1024 + 2 * W_BYTES + T_STAGES * 2 * T_BYTES + X_STAGES * X_BYTES + 256;

// Aimg[ft][half][row = fl*12 + e][32] (SWIZZLE_128B), element k = joint
__global__ void pack_transforms_kernel(const float* __restrict__ A, int F, int n_ftiles, float* __restrict__ Aimg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;   // one 16 B chunk: (ft, row, c)
  if (i >= n_ftiles * NCOL * 8) return;
  const int c = i & 7, row = (i >> 3) % NCOL, ft = (i >> 3) / NCOL;
  const int fl = row / 12, e = row % 12, f = ft * NF + fl;
  float4 h = make_float4(0.f, 0.f, 0.f, 0.f), l = h;
  if (f < F && c < UUO_SKIN_J / 4) {
    const float* src = A + ((size_t)f * UUO_SKIN_J + c * 4) * 12 + e;
    umma::split_tf32(src[0], h.x, l.x);
    umma::split_tf32(src[12], h.y, l.y);
    umma::split_tf32(src[24], h.z, l.z);
    umma::split_tf32(src[36], h.w, l.w);
  }
  float* hi = Aimg + (size_t)ft * 2 * NCOL * 32;
  const int off = row * 32 + ((c ^ (row & 7)) << 2);
  *reinterpret_cast<float4*>(hi + off) = h;
  *reinterpret_cast<float4*>(hi + NCOL * 32 + off) = l;
}

struct SkinArgs {
  const float* Wimg;      // [vtiles][2][128*32]
  const float* Aimg;      // [ftiles][2][192*32]
  const float* in;        // mode 0: v_posed [F][NP];  mode 1: g_vertices [F][V][3] (may be null)
  const float* g_joints;  // mode 1: [F][J+NE][3] or null
  const int32_t* extra_ids;
  const uint8_t* vflag;
  const float* transl;    // mode 0: [F][3] or null
  float* out;             // mode 0: vertices [F][V][3];  mode 1: g_vposed [F][NP]
  int F, V, NP, NE, J, n_ftiles, ft_per_chunk;
};

template <int MODE>
__global__ void __launch_bounds__(NTHREADS, 1) skin_tc_kernel(const SkinArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = umma::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - raw);
  const uint32_t w_smem = base, t_smem = base + 2 * W_BYTES, x_smem = t_smem + T_STAGES * 2 * T_BYTES;
  uint8_t* x_ptr = sm + 2 * W_BYTES + T_STAGES * 2 * T_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(x_ptr + X_STAGES * X_BYTES);
  const uint32_t bar0 = umma::smem_u32(bars);
  // transform ring (bulk copy -> MMA), input ring (cp.async -> epilogue), accumulator ring (MMA -> epilogue)
  auto TF_FULL = [&](int s) { return bar0 + 8u * s; };
  auto TF_EMPTY = [&](int s) { return bar0 + 8u * (T_STAGES + s); };
  auto X_FULL = [&](int s) { return bar0 + 8u * (2 * T_STAGES + s); };
  auto X_EMPTY = [&](int s) { return bar0 + 8u * (2 * T_STAGES + X_STAGES + s); };
  auto TFULL = [&](int q) { return bar0 + 8u * (2 * T_STAGES + 2 * X_STAGES + q); };
  auto TEMPTY = [&](int q) { return bar0 + 8u * (2 * T_STAGES + 2 * X_STAGES + 2 + q); };
  const uint32_t WBAR = bar0 + 8u * (2 * T_STAGES + 2 * X_STAGES + 4);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * T_STAGES + 2 * X_STAGES + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int vt = blockIdx.x;
  const int ft0 = blockIdx.y * a.ft_per_chunk, ft1 = min(ft0 + a.ft_per_chunk, a.n_ftiles);

  if (threadIdx.x == 0) {
    for (int s = 0; s < T_STAGES; ++s) { umma::mbar_init(TF_FULL(s), 1); umma::mbar_init(TF_EMPTY(s), 1); }
    for (int q = 0; q < 2; ++q) { umma::mbar_init(TFULL(q), 1); umma::mbar_init(TEMPTY(q), EPI_WARPS); }
    umma::mbar_init(WBAR, 1);
    umma::fence_barrier_init();
  }
  if (warp == WARP_MMA) umma::tmem_alloc(umma::smem_u32(tmem_slot), 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == WARP_LOAD) {
    // ===== loader: bulk copies of W (once) and of the transform image of every frame tile =====
    if (lane == 0) {
      umma::mbar_arrive_expect_tx(WBAR, 2 * W_BYTES);
      umma::bulk_g2s(w_smem, a.Wimg + (size_t)vt * 2 * VT * 32, 2 * W_BYTES, WBAR);
    }
    if (lane == 0) {
      uint32_t it = 0;
      for (int ft = ft0; ft < ft1; ++ft, ++it) {
        const int ts = it % T_STAGES;
        umma::mbar_wait(TF_EMPTY(ts), ((it / T_STAGES) & 1) ^ 1);
        umma::mbar_arrive_expect_tx(TF_FULL(ts), 2 * T_BYTES);
        umma::bulk_g2s(t_smem + ts * 2 * T_BYTES, a.Aimg + (size_t)ft * 2 * NCOL * 32, 2 * T_BYTES, TF_FULL(ts));
      }
    }
  } else if (warp == WARP_MMA) {
    // the warp stays converged; one elected lane issues
    const uint32_t idesc = umma::idesc_tf32_m128(NCOL);
    umma::mbar_wait(WBAR, 0);
    uint32_t it = 0;
    for (int ft = ft0; ft < ft1; ++ft, ++it) {
      const int s = it % T_STAGES;
      const uint32_t acc = it & 1;
      umma::mbar_wait(TEMPTY(acc), ((it >> 1) & 1) ^ 1);
      umma::mbar_wait(TF_FULL(s), (it / T_STAGES) & 1);
      umma::tc_fence_after();
      if (umma::elect_one()) {
        const uint32_t d_tmem = tmem_base + acc * 256;
        const uint32_t t_hi = t_smem + s * 2 * T_BYTES, t_lo = t_hi + T_BYTES;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const uint64_t dwh = umma::smem_desc_sw128(w_smem + k * 32), dwl = umma::smem_desc_sw128(w_smem + W_BYTES + k * 32);
          const uint64_t dth = umma::smem_desc_sw128(t_hi + k * 32), dtl = umma::smem_desc_sw128(t_lo + k * 32);
          umma::mma_tf32(d_tmem, dwh, dth, idesc, k > 0 ? 1u : 0u);
          umma::mma_tf32(d_tmem, dwl, dth, idesc, 1u);
          umma::mma_tf32(d_tmem, dwh, dtl, idesc, 1u);
        }
        umma::mma_commit(TF_EMPTY(s));
        umma::mma_commit(TFULL(acc));
      }
      __syncwarp();
    }
  } else {
    // ===== epilogue: warp quadrant = TMEM lanes, warp group = which 8 of the tile's 16 frames =====
    const int gq = warp >> 2, quad = warp & 3;
    const int vl = quad * 32 + lane, v = vt * VT + vl;
    const bool v_ok = v < a.V;
    // Per-warp input ring: the warp copies exactly the rows it will consume (8 frames x 32 vertices x 12 B = 384 B
    // contiguous per frame, + the 8 transl rows) with cp.async, X_STAGES - 1 tiles ahead; consumption is warp-local
    // (cp.async.wait_group + __syncwarp), so no register prefetch sits on the load scoreboard.
    uint8_t* xw = x_ptr + (size_t)warp * X_STAGES * XW_BYTES;
    const uint32_t xw_s = x_smem + warp * X_STAGES * XW_BYTES;
    // A row (frame f, the warp's 32 vertices) is 384 contiguous bytes whose address is only 4/8 B aligned for the
    // [F][V][3] tensors: copy its 16 B-aligned hull (<= 25 chunks) and remember the shift; slots are 400 B apart.
    const long pitch = MODE == 0 ? (long)a.NP : (long)a.V * 3;
    const int v_first = vt * VT + quad * 32;
    const int row_bytes = MODE == 0 ? 384 : max(0, min(384, (a.V - v_first) * 12));   // bytes of the row that exist
    const uintptr_t in_base = reinterpret_cast<uintptr_t>(a.in) + (size_t)v_first * 12;
    const uintptr_t in_end = reinterpret_cast<uintptr_t>(a.in) + (size_t)a.F * pitch * 4;
    const size_t pitch4 = (size_t)pitch * 4;
    // MODE 0 reads v_posed: rows of NP floats (a multiple of 64 B) from a 16 B aligned workspace - no shift, no hull
    auto row_shift = [&](int f) { return MODE == 0 ? 0u : (uint32_t)((in_base + (size_t)f * pitch * 4) & 15); };
    auto issue = [&](int ft, int stage) {
      if (ft < ft1) {
        const uint32_t dst0 = xw_s + stage * XW_BYTES;
        const int f0 = ft * NF + gq * 8;
        const uintptr_t row0 = in_base + (size_t)f0 * pitch4;
        if (MODE == 0) {                 // 8 frames x 24 aligned 16 B chunks = 6 per lane
#pragma unroll
          for (int i = 0; i < 6; ++i) {
            const int q = lane + 32 * i, fi = q / 24, ch = q - fi * 24;
            const bool ok = f0 + fi < a.F;
            umma::cp_async16(dst0 + fi * XROW + ch * 16,
                             ok ? reinterpret_cast<const void*>(row0 + (size_t)fi * pitch4 + ch * 16) : (const void*)a.Wimg,
                             ok ? 16u : 0u);
          }
        } else
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const int q = lane + 32 * i;
          if (q < 8 * 25) {
            const int fi = q / 25, ch = q - fi * 25;
            const uintptr_t row = row0 + (size_t)fi * pitch4;
            const uint32_t shift = (uint32_t)(row & 15);
            const uintptr_t src = row - shift + ch * 16;
            const bool ok = a.in != nullptr && f0 + fi < a.F && ch * 16 < (int)shift + row_bytes;
            const uint32_t n = ok ? (uint32_t)min((uintptr_t)16, in_end - src) : 0u;
            umma::cp_async16(dst0 + fi * XROW + ch * 16, ok ? reinterpret_cast<const void*>(src) : (const void*)a.Wimg, n);
          }
        }
        if (MODE == 0 && lane < 24) {   // transl of the warp's 8 frames, one float per lane (caller memory: 4 B aligned)
          const long e0 = ((long)ft * NF + gq * 8) * 3 + lane;
          const bool ok = a.transl != nullptr && e0 < (long)a.F * 3;
          umma::cp_async4(dst0 + 8 * XROW + lane * 4, ok ? (const void*)(a.transl + e0) : (const void*)a.Wimg, ok ? 4u : 0u);
        }
      }
      umma::cp_async_commit();
    };
    uint32_t it = 0;
#pragma unroll
    for (int d = 0; d < X_STAGES - 1; ++d) issue(ft0 + d, d);
    for (int ft = ft0; ft < ft1; ++ft, ++it) {
      const int fbase = ft * NF + gq * 8;
      issue(ft + X_STAGES - 1, (it + X_STAGES - 1) % X_STAGES);
      const uint32_t acc = it & 1;
      const float* xin = reinterpret_cast<const float*>(xw + (it % X_STAGES) * XW_BYTES);
      umma::cp_async_wait<X_STAGES - 1>();
      __syncwarp();
      umma::mbar_wait(TFULL(acc), (it >> 1) & 1);
      umma::tc_fence_after();
      // output row of the tile's first frame for this lane's vertex; frames are out_pitch floats apart
      const size_t out_pitch = MODE == 0 ? (size_t)a.V * 3 : (size_t)a.NP;
      float* const o_tile = a.out + (size_t)fbase * out_pitch + (size_t)v * 3;
#pragma unroll
      for (int st = 0; st < 2; ++st) {
        uint32_t r[48];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * 256 + (gq * 8 + st * 4) * 12;
        umma::tmem_ld16(taddr, r);
        umma::tmem_ld16(taddr + 16, r + 16);
        umma::tmem_ld16(taddr + 32, r + 32);
        float x[4][3];
#pragma unroll
        for (int k = 0; k < 4; ++k) {   // lane stride 12 B: conflict-free
          const int i = st * 4 + k;
          const float* xr = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(xin) + i * XROW + row_shift(fbase + i)) + lane * 3;
          x[k][0] = v_ok ? xr[0] : 0.f; x[k][1] = v_ok ? xr[1] : 0.f; x[k][2] = v_ok ? xr[2] : 0.f;
        }
        umma::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int i = st * 4 + k, f = fbase + i;
          float T[12];
#pragma unroll
          for (int e = 0; e < 12; ++e) T[e] = __uint_as_float(r[k * 12 + e]);
          if (f < a.F) {
            if (MODE == 0) {
              if (v_ok) {
                const float* tr = xin + 8 * (XROW / 4) + i * 3;   // broadcast read
                float* o = o_tile + (size_t)i * out_pitch;
#pragma unroll
                for (int rr = 0; rr < 3; ++rr)
                  __stcs(o + rr, (T[rr * 4] * x[k][0] + T[rr * 4 + 1] * x[k][1] + T[rr * 4 + 2] * x[k][2] + T[rr * 4 + 3]) + tr[rr]);   // streamed
              }
            } else {
              // vertex slots past V have zero weights -> zeros, which also clears the row padding the GEMM reads
              float* o = o_tile + (size_t)i * out_pitch;
              const float g0 = x[k][0], g1 = x[k][1], g2 = x[k][2];
#pragma unroll
              for (int c = 0; c < 3; ++c) o[c] = T[c] * g0 + T[4 + c] * g1 + T[8 + c] * g2;
            }
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

int launch_pack_transforms(const uuo_model* m, int F, const uuo_ws& ws, cudaStream_t st) {
  const int n_ft = (F + NF - 1) / NF;
  const int n = n_ft * NCOL * 8;
  pack_transforms_kernel<<<(n + 255) / 256, 256, 0, st>>>(ws.A, F, n_ft, ws.Aimg);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_skin_tc(const uuo_model* m, int mode, int F, const float* in, const float* g_joints, const float* transl,
                   float* out, const uuo_ws& ws, cudaStream_t st) {
  UUO_DYN_SMEM(skin_tc_kernel<0>, SMEM_BYTES);
  UUO_DYN_SMEM(skin_tc_kernel<1>, SMEM_BYTES);
  const int num_sms = uuo_sm_count();
  SkinArgs a;
  a.Wimg = m->Wimg; a.Aimg = ws.Aimg; a.in = in; a.g_joints = g_joints; a.extra_ids = m->extra_ids; a.vflag = m->vflag;
  a.transl = transl; a.out = out;
  a.F = F; a.V = m->V; a.NP = m->NP; a.NE = m->NE; a.J = m->J;
  a.n_ftiles = (F + NF - 1) / NF;
  const int vtiles = m->NP / 3 / VT;
  // frame chunks: about three waves of CTAs, each walking a contiguous run of frame tiles with W resident
  int chunks = std::max(1, std::min(a.n_ftiles, (3 * num_sms) / vtiles));
  a.ft_per_chunk = (a.n_ftiles + chunks - 1) / chunks;
  chunks = (a.n_ftiles + a.ft_per_chunk - 1) / a.ft_per_chunk;
  dim3 grid(vtiles, chunks);
  if (mode == 0) skin_tc_kernel<0><<<grid, NTHREADS, SMEM_BYTES, st>>>(a);
  else skin_tc_kernel<1><<<grid, NTHREADS, SMEM_BYTES, st>>>(a);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}
