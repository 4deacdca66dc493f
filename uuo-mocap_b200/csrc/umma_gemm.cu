// The two dense contractions of lbs() on the 5th-gen tensor cores (tcgen05.mma kind::tf32, TMEM accumulators),
// 3xTF32 split so the result stays at fp32 accuracy (vertices 1e-5 / gradients 1e-4 contract):
//     x = hi + lo,  hi = tf32(x), lo = tf32(x - hi);   A.B ~= Ahi.Bhi + Alo.Bhi + Ahi.Blo
//   forward  v_posed[F,NP] = v_template + coef[F,KP] . Bmat[KP,NP]        (lbs.py:209 + lbs.py:229-233)
//   backward g_coef[F,KP]  = g_vposed[F,NP] . Bmat[KP,NP]^T  (split-K, deterministic partial slabs)
// Both are C[M,N] = A[M,K] . B[N,K]^T with K-major operands.
//   * B is a model constant: uuo_model_create stores it pre-split (hi|lo) and pre-swizzled, tile by tile, exactly as
//     the shared-memory image the MMA descriptor expects, so one 1-D bulk copy (UBLKCP) per stage fetches it.
//   * A changes every call: 8 producer warps read the fp32 rows (coalesced 128 B), split them and store the
//     SWIZZLE_128B image of the hi and lo tiles.
//   * one thread issues the MMAs; accumulators are double buffered in TMEM (2 x 256 columns) so the epilogue of
//     tile i (tcgen05.ld -> smem transpose -> coalesced stores, + v_template) overlaps the MMAs of tile i+1.
// Persistent: one CTA per SM walks work items (n_tile, k_split, m_tile) round robin.
#include <algorithm>

#include "common.cuh"
#include "umma.cuh"

namespace {

constexpr int BM = 128;             // rows of A per tile (frames)
constexpr int BK = 32;              // fp32 elements per K block = one 128 B swizzle row
constexpr int NT_MAX = 256;         // max columns per tile
constexpr int A_TILE_BYTES = BM * BK * 4;            // 16 KB (hi) + 16 KB (lo)
constexpr int B_TILE_BYTES_MAX = NT_MAX * BK * 4;    // 32 KB (hi) + 32 KB (lo)
// 3 terms: a stage holds A hi|lo + B hi|lo = 96 KB -> 2 stages.  2 terms: B's lo half is never fetched, a stage is
// 64 KB -> 3 stages, which is what hides the fill latency of a stage (split + store of A, L2 round trip of B) behind
// the 8 MMAs of the other two
template <int TERMS> struct Cfg {
  static constexpr int STAGES = TERMS == 3 ? 2 : 3;
  static constexpr int STAGE_BYTES = 2 * A_TILE_BYTES + (TERMS == 3 ? 2 : 1) * B_TILE_BYTES_MAX;
};
constexpr int EPI_WARPS = 4, PROD_WARPS = 8;
constexpr int WARP_MMA = EPI_WARPS, WARP_BLOAD = EPI_WARPS + 1, WARP_PROD0 = EPI_WARPS + 2;
constexpr int NTHREADS = (EPI_WARPS + 2 + PROD_WARPS) * 32;   // 448
constexpr int STAGING_FLOATS = 32 * 32;              // per epilogue warp, XOR-swizzled transpose tile
template <int TERMS> constexpr int smem_bytes() {
  return 1024 /*align slack*/ + Cfg<TERMS>::STAGES * Cfg<TERMS>::STAGE_BYTES + EPI_WARPS * STAGING_FLOATS * 4 + 256;
}

struct GemmArgs {
  const float* A; long lda; int M;   // fp32 row-major, K contiguous
  const float* Bimg;                 // [n_tiles][kb_total][2][NT*32] swizzled images
  float* C; long ldc; long split_stride;
  const float* bias;                 // [n_tiles*NT] or null
  int NT, n_tiles, m_tiles, kb_total, k_splits, kb_per_split, n_valid;
  int m_group;                       // m tiles per group of the item order (see decode_item)
  int terms;                         // 3: Ahi.Bhi + Alo.Bhi + Ahi.Blo;  2: B's lo half is neither fetched nor multiplied
  int k_zero_from;                   // columns k >= this of A are read as 0 (-1: none)
  // row GROUPS (rows_tc.cu, stacked hypotheses): group_rows > 0 = consecutive blocks of that many rows of A / C, each
  // with its own B image and bias; an m tile never straddles two groups (mt_per_group tiles per group, the last ragged)
  int group_rows, mt_per_group;
  long b_group_stride, bias_group_stride;
};

// rows of an m tile: first global row, and how many rows from `local` on are real
__device__ __forceinline__ void tile_rows(const GemmArgs& g, int mt, int local, long& row, int& nreal, int& grp) {
  if (g.group_rows > 0) {
    grp = mt / g.mt_per_group;
    const int rl = (mt - grp * g.mt_per_group) * BM + local;
    row = (long)grp * g.group_rows + rl;
    nreal = g.group_rows - rl;
  } else {
    grp = 0;
    row = (long)mt * BM + local;
    nreal = g.M - (int)row;
  }
}

// Work item w -> (m tile, n tile, k split).  m tiles are walked in GROUPS of g.m_group: inside a group the CTAs sweep
// all its m tiles for one n tile, then the next n tile, ... so the A rows of the group (m_group x 128 rows) are re-read
// from L2 for every n tile instead of from HBM (ncu at F = 100 000, one group: 5.6 GB of DRAM reads for 0.11 GB of
// operands, because a full sweep over 782 m tiles pushes ~190 MB through the 126 MB L2 between two uses of a row).
__device__ __forceinline__ void decode_item(const GemmArgs& g, int w, int& mt, int& nt, int& ks) {
  const int per_group = g.m_group * g.n_tiles * g.k_splits;
  const int grp = w / per_group, m0 = grp * g.m_group;
  const int mg = min(g.m_group, g.m_tiles - m0);          // the last group may be short
  int r = w - grp * per_group;
  mt = m0 + r % mg;
  r /= mg;
  ks = r % g.k_splits;
  nt = r / g.k_splits;
}

template <int TERMS>
__global__ void __launch_bounds__(NTHREADS, 1) umma_gemm_kernel(const GemmArgs g) {
  constexpr int STAGES = Cfg<TERMS>::STAGES, STAGE_BYTES = Cfg<TERMS>::STAGE_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = umma::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - raw);
  float* staging = reinterpret_cast<float*>(sm + STAGES * STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + STAGES * STAGE_BYTES + EPI_WARPS * STAGING_FLOATS * 4);
  // bars: [0,S) full_a, [S,2S) full_b, [2S,3S) empty, [3S,3S+2) tmem_full, [3S+2,3S+4) tmem_empty
  const uint32_t bar0 = umma::smem_u32(bars);
  auto FULL_A = [&](int s) { return bar0 + 8u * s; };
  auto FULL_B = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto EMPTY = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto TFULL = [&](int a) { return bar0 + 8u * (3 * STAGES + a); };
  auto TEMPTY = [&](int a) { return bar0 + 8u * (3 * STAGES + 2 + a); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = g.m_tiles * g.n_tiles * g.k_splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      umma::mbar_init(FULL_A(s), PROD_WARPS);
      umma::mbar_init(FULL_B(s), 1);
      umma::mbar_init(EMPTY(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      umma::mbar_init(TFULL(a), 1);
      umma::mbar_init(TEMPTY(a), EPI_WARPS);
    }
    umma::fence_barrier_init();
  }
  if (warp == WARP_MMA) umma::tmem_alloc(umma::smem_u32(tmem_slot), 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t b_tile_bytes = (uint32_t)g.NT * BK * 4;

  if (warp >= WARP_PROD0) {
    // ===== A producers: fp32 rows -> (hi, lo) SWIZZLE_128B images; global loads run PF K-blocks ahead =====
    constexpr int PF = 3;
    const int tp = threadIdx.x - WARP_PROD0 * 32;
    int w = blockIdx.x, mt = 0, nt = 0, ks = 0, kb = 0, kb_end = 0;
    bool live = w < n_items;
    if (live) { decode_item(g, w, mt, nt, ks); kb = ks * g.kb_per_split; kb_end = min(kb + g.kb_per_split, g.kb_total); }
    auto gload = [&](float4 (&v)[4], int& kb_of) {   // loads the block the iterator points at, then advances it
      kb_of = kb;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int q = tp + 256 * i, row = q >> 3, c = q & 7;
        long gr;
        int nreal, grp;
        tile_rows(g, mt, row, gr, nreal, grp);
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (nreal > 0) v[i] = __ldg(reinterpret_cast<const float4*>(g.A + gr * g.lda + (long)kb * BK + c * 4));
      }
      if (++kb >= kb_end) {
        w += gridDim.x;
        live = w < n_items;
        if (live) { decode_item(g, w, mt, nt, ks); kb = ks * g.kb_per_split; kb_end = min(kb + g.kb_per_split, g.kb_total); }
      }
    };
    uint32_t loaded = 0, it = 0;
    // split + store one block.  Nothing touches the loaded registers before this point: the k_zero_from masking used
    // to sit in gload, right behind the __ldg, so every load was waited for at once and the prefetch distance was lost
    // (ncu source page: 24 % of the kernel's stall samples on that line)
    auto consume = [&](float4 (&v)[4], int kb_of) {
      const int s = it % STAGES;
      umma::mbar_wait(EMPTY(s), ((it / STAGES) & 1) ^ 1);
      uint8_t* a_hi = sm + s * STAGE_BYTES;
      uint8_t* a_lo = a_hi + A_TILE_BYTES;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int q = tp + 256 * i, row = q >> 3, c = q & 7;
        const int off = row * 128 + ((c ^ (row & 7)) << 4);
        float4 x = v[i];
        if (g.k_zero_from >= 0) {
          const int k0 = kb_of * BK + c * 4;
          if (k0 + 3 >= g.k_zero_from) {
            if (k0 >= g.k_zero_from) x.x = 0.f;
            if (k0 + 1 >= g.k_zero_from) x.y = 0.f;
            if (k0 + 2 >= g.k_zero_from) x.z = 0.f;
            x.w = 0.f;
          }
        }
        float4 h, l;
        umma::split_tf32(x.x, h.x, l.x);
        umma::split_tf32(x.y, h.y, l.y);
        umma::split_tf32(x.z, h.z, l.z);
        umma::split_tf32(x.w, h.w, l.w);
        *reinterpret_cast<float4*>(a_hi + off) = h;
        *reinterpret_cast<float4*>(a_lo + off) = l;
      }
      umma::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(FULL_A(s));
      ++it;
    };
    // a ring of PF + 1 register buffers with STATIC roles (the loop is unrolled over the ring): slot C is consumed while
    // slot L receives the block PF ahead; no register moves, so a load is first waited for PF blocks after its issue
    float4 b0[4], b1[4], b2[4], b3[4];
    int k0_ = 0, k1_ = 0, k2_ = 0, k3_ = 0;
    static_assert(PF == 3, "the ring below is written for PF = 3");
    if (live) { gload(b0, k0_); ++loaded; }
    if (live) { gload(b1, k1_); ++loaded; }
    if (live) { gload(b2, k2_); ++loaded; }
#define UUO_GEMM_STEP(C, KC, L, KL)                        \
    {                                                      \
      if (it >= loaded) break;                             \
      if (live) { gload(L, KL); ++loaded; }                \
      consume(C, KC);                                      \
    }
    while (true) {
      UUO_GEMM_STEP(b0, k0_, b3, k3_)
      UUO_GEMM_STEP(b1, k1_, b0, k0_)
      UUO_GEMM_STEP(b2, k2_, b1, k1_)
      UUO_GEMM_STEP(b3, k3_, b2, k2_)
    }
#undef UUO_GEMM_STEP
  } else if (warp == WARP_BLOAD) {
    // ===== B loader: one bulk copy per stage brings the pre-swizzled (hi|lo) tile =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
        int mt, nt, ks;
        decode_item(g, w, mt, nt, ks);
        const int kb0 = ks * g.kb_per_split, kb1 = min(kb0 + g.kb_per_split, g.kb_total);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % STAGES;
          umma::mbar_wait(EMPTY(s), ((it / STAGES) & 1) ^ 1);
          const uint32_t nbytes = (TERMS == 3 ? 2u : 1u) * b_tile_bytes;       // hi | lo are contiguous: hi alone = first half
          umma::mbar_arrive_expect_tx(FULL_B(s), nbytes);
          const float* src = g.Bimg + ((size_t)nt * g.kb_total + kb) * (size_t)(2 * g.NT * BK) +
                             (g.group_rows > 0 ? (size_t)(mt / g.mt_per_group) * g.b_group_stride : 0);
          umma::bulk_g2s(base + s * STAGE_BYTES + 2 * A_TILE_BYTES, src, nbytes, FULL_B(s));
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // ===== MMA issuer: the warp stays converged, one elected lane issues (uniform-register operands) =====
    const uint32_t idesc = umma::idesc_tf32_m128((uint32_t)g.NT);
    uint32_t it = 0, item = 0;
    for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++item) {
      int mt, nt, ks;
      decode_item(g, w, mt, nt, ks);
      const int kb0 = ks * g.kb_per_split, kb1 = min(kb0 + g.kb_per_split, g.kb_total);
      const uint32_t acc = item & 1;
      umma::mbar_wait(TEMPTY(acc), ((item >> 1) & 1) ^ 1);
      const uint32_t d_tmem = tmem_base + acc * NT_MAX;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        umma::mbar_wait(FULL_A(s), ph);
        umma::mbar_wait(FULL_B(s), ph);
        umma::tc_fence_after();
        if (umma::elect_one()) {
          const uint32_t a_hi = base + s * STAGE_BYTES, a_lo = a_hi + A_TILE_BYTES;
          const uint32_t b_hi = a_hi + 2 * A_TILE_BYTES, b_lo = b_hi + b_tile_bytes;
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            const uint64_t dah = umma::smem_desc_sw128(a_hi + k * 32), dal = umma::smem_desc_sw128(a_lo + k * 32);
            const uint64_t dbh = umma::smem_desc_sw128(b_hi + k * 32), dbl = umma::smem_desc_sw128(b_lo + k * 32);
            umma::mma_tf32(d_tmem, dah, dbh, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            umma::mma_tf32(d_tmem, dal, dbh, idesc, 1u);
            if (TERMS == 3) umma::mma_tf32(d_tmem, dah, dbl, idesc, 1u);
          }
          umma::mma_commit(EMPTY(s));                     // smem slot is free once these MMAs have read it
          if (kb + 1 == kb1) umma::mma_commit(TFULL(acc));   // accumulator complete
        }
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue warps 0..3: TMEM lanes 32*warp .. +31 =====
    float* stg = staging + warp * STAGING_FLOATS;
    uint32_t item = 0;
    for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++item) {
      int mt, nt, ks;
      decode_item(g, w, mt, nt, ks);
      const uint32_t acc = item & 1;
      umma::mbar_wait(TFULL(acc), (item >> 1) & 1);
      umma::tc_fence_after();
      float* Cw = g.C + (size_t)ks * g.split_stride;
      long row0;
      int nreal, grp;
      tile_rows(g, mt, warp * 32, row0, nreal, grp);
      const int nrows = min(32, nreal);        // may be <= 0 for a ragged last tile
      const int sub = lane >> 3, c4 = lane & 7;
      for (int c = 0; c < g.NT / 32; ++c) {
        uint32_t r[32];
        umma::tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + acc * NT_MAX + c * 32, r);
        umma::tmem_ld_wait();
        // lane = row: 8 x 16 B chunks, chunk position XOR (row & 7) -> conflict-free both ways
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) =
              make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                          __uint_as_float(r[4 * j + 3]));
        __syncwarp();
        const int col = nt * g.NT + c * 32 + c4 * 4;
        if (col < g.n_valid) {
          float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
          if (g.bias) bv = __ldg(reinterpret_cast<const float4*>(g.bias + (size_t)grp * g.bias_group_stride + col));
          float* dst = Cw + (size_t)row0 * g.ldc + col;
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) {
            const int rl = rr * 4 + sub;     // 4 rows per instruction, 8 lanes x 16 B = one 128 B line per row
            float4 v = *reinterpret_cast<const float4*>(stg + rl * 32 + ((c4 ^ (rl & 7)) << 2));
            v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
            if (rl < nrows) __stcs(reinterpret_cast<float4*>(dst + (size_t)rl * g.ldc), v);   // streamed: keep A / B in L2
          }
        }
        __syncwarp();
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

static inline float host_tf32(float x) {   // round to nearest (ties away) like cvt.rna.tf32.f32
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7f800000u) != 0x7f800000u) u += 0x1000u;
  u &= 0xffffe000u;
  float r;
  memcpy(&r, &u, 4);
  return r;
}

void uuo_build_b_image(std::vector<float>& out, const float* src, int rows, int K, int NT) {
  const int n_tiles = (rows + NT - 1) / NT, kbT = K / BK;
  out.assign((size_t)n_tiles * kbT * 2 * NT * BK, 0.f);
  for (int t = 0; t < n_tiles; ++t)
    for (int kb = 0; kb < kbT; ++kb) {
      float* hi = out.data() + ((size_t)t * kbT + kb) * 2 * NT * BK;
      float* lo = hi + (size_t)NT * BK;
      for (int r = 0; r < NT; ++r) {
        const int gr = t * NT + r;
        if (gr >= rows) continue;
        for (int k = 0; k < BK; ++k) {
          const float x = src[(size_t)gr * K + kb * BK + k];
          const float h = host_tf32(x);
          const int off = r * 32 + (((k >> 2) ^ (r & 7)) << 2) + (k & 3);
          hi[off] = h;
          lo[off] = host_tf32(x - h);
        }
      }
    }
}

static int launch_umma(GemmArgs g, cudaStream_t st) {
  const int items = g.m_tiles * g.n_tiles * g.k_splits;
  const int grid = std::min(items, uuo_sm_count());
  g.m_group = std::min(g.m_tiles, 2 * grid);               // <= 296 m tiles = 34 MB of fp32 rows at K = 224 (1x and 4x: slower)
  if (g.terms == 2) {
    UUO_DYN_SMEM(umma_gemm_kernel<2>, smem_bytes<2>());
    umma_gemm_kernel<2><<<grid, NTHREADS, smem_bytes<2>(), st>>>(g);
  } else {
    UUO_DYN_SMEM(umma_gemm_kernel<3>, smem_bytes<3>());
    umma_gemm_kernel<3><<<grid, NTHREADS, smem_bytes<3>(), st>>>(g);
  }
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

// number of deterministic K-split slabs the backward GEMM writes for F frames (<= UUO_NSPLIT_GC)
int uuo_gcoef_splits(const uuo_model* m, int F) {
  const int m_tiles = (F + BM - 1) / BM, kbT = m->NP / BK;
  int want = std::max(1, std::min(UUO_NSPLIT_GC, (2 * uuo_sm_count_of(m->device) + m_tiles - 1) / m_tiles));
  const int per = (kbT + want - 1) / want;
  return (kbT + per - 1) / per;
}

// v_shaped[c] = v_template[c] + sum_l betas[l] shapedirs[c][l]  (lbs.py:209, 275-296) for ONE betas row, exact fp32
__global__ void __launch_bounds__(256) vshaped_kernel(const float* __restrict__ vt, const float* __restrict__ Bmat, int NP,
                                                      int P, int NB, const float* __restrict__ betas,
                                                      float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  float a = 0.f;
  for (int l = 0; l < NB; ++l) a = fmaf(__ldg(betas + l), Bmat[(size_t)(P + l) * NP + c], a);
  out[c] = vt[c] + a;          // the offset first, then the template: the order of lbs.py:209 (v_template + blend_shapes)
}

int launch_vshaped(const uuo_model* m, const float* betas, const uuo_ws& ws, cudaStream_t st) {
  vshaped_kernel<<<(m->NP + 255) / 256, 256, 0, st>>>(m->vt, m->Bmat, m->NP, m->P, m->NB, betas, ws.vshaped);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

int launch_vposed_gemm_umma(const uuo_model* m, int F, const uuo_ws& ws, bool betas_shared, cudaStream_t st) {
  GemmArgs g;
  g.A = ws.coef; g.lda = m->KP; g.M = F;
  g.Bimg = m->Bimg_fwd;
  g.C = ws.vposed; g.ldc = m->NP; g.split_stride = 0;
  g.bias = m->vt; g.terms = 3; g.k_zero_from = -1;
  g.group_rows = 0; g.mt_per_group = 1; g.b_group_stride = 0; g.bias_group_stride = 0;
  if (betas_shared) {
    // betas are shared by all frames (the fitting loop, optimization.py:196-202): the shape blend is frame independent,
    // so it is evaluated once in exact fp32 and enters as the bias; the tensor cores see only the pose rows, where the
    // constant operand's rounded hi half is enough (posedirs offsets are ~1e-2 of a vertex coordinate: the dropped
    // A.B_lo term is < 1e-6 of it), which also halves the B bytes every tile pulls through L2
    g.bias = ws.vshaped; g.terms = 2; g.k_zero_from = m->P;
  }
  g.NT = 256; g.n_tiles = m->NP / 256; g.m_tiles = (F + BM - 1) / BM;
  g.kb_total = m->KP / BK; g.k_splits = 1; g.kb_per_split = g.kb_total; g.n_valid = m->NP;
  ProfScope ps_("vposed_gemm", st);
  return launch_umma(g, st);
}

// C[M][ldc] = A[M][lda] . B^T (+ bias) for ONE n tile of NT columns (a multiple of 32, <= 256) whose 3xTF32 image
// [kb_total][2][NT*32] the caller built (rows_tc.cu: the selected rows of the marker stage); K = 32 kb_total.
// groups > 1: M = groups x (M / groups) rows, group g uses the image at Bimg + g b_group_stride (bias likewise)
int launch_umma_one_tile(const float* A, long lda, int M, const float* Bimg, int NT, int kb_total, const float* bias,
                         float* C, long ldc, int groups, long b_group_stride, long bias_group_stride, cudaStream_t st) {
  GemmArgs g;
  g.A = A; g.lda = lda; g.M = M;
  g.Bimg = Bimg;
  g.C = C; g.ldc = ldc; g.split_stride = 0;
  g.bias = bias; g.terms = 3; g.k_zero_from = -1;
  g.NT = NT; g.n_tiles = 1; g.m_tiles = (M + BM - 1) / BM;
  g.group_rows = 0; g.mt_per_group = 1; g.b_group_stride = 0; g.bias_group_stride = 0;
  if (groups > 1) {
    g.group_rows = M / groups;
    g.mt_per_group = (g.group_rows + BM - 1) / BM;
    g.m_tiles = groups * g.mt_per_group;
    g.b_group_stride = b_group_stride; g.bias_group_stride = bias_group_stride;
  }
  g.kb_total = kb_total; g.k_splits = 1; g.kb_per_split = kb_total; g.n_valid = NT;
  return launch_umma(g, st);
}

int launch_gcoef_gemm_umma(const uuo_model* m, int F, const uuo_ws& ws, cudaStream_t st) {
  GemmArgs g;
  g.A = ws.gvposed; g.lda = m->NP; g.M = F;
  g.Bimg = m->Bimg_bwd;
  g.C = ws.gcoef; g.ldc = m->KP; g.split_stride = (long)F * m->KP;
  g.bias = nullptr; g.terms = 3; g.k_zero_from = -1;
  g.group_rows = 0; g.mt_per_group = 1; g.b_group_stride = 0; g.bias_group_stride = 0;
  g.NT = m->KP; g.n_tiles = 1; g.m_tiles = (F + BM - 1) / BM;
  g.kb_total = m->NP / BK; g.k_splits = uuo_gcoef_splits(m, F);
  g.kb_per_split = (g.kb_total + g.k_splits - 1) / g.k_splits; g.n_valid = m->KP;
  ProfScope ps_("gcoef_gemm", st);
  return launch_umma(g, st);
}
