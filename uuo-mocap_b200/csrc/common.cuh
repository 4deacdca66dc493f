// Shared declarations of libuuomocap (sm_100a only).  See include/uuo_mocap.h for the C-ABI
// and DESIGN.md for the data layout in HBM.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../include/uuo_mocap.h"

#define UUO_MAX_J 32       // one lane per joint in the FK kernels
#define UUO_MAX_CHILD 4    // children per joint handled by the reverse tree sweep
#define UUO_SKIN_J 24      // the dense skinning kernels are compiled for the SMPL joint count
#define UUO_NSPLIT_GA 4    // vertex splits of the skinning-backward dA reduction (deterministic partials)
#define UUO_NSPLIT_GC 16   // K splits of the posedirs-transpose GEMM (deterministic partials)
#define UUO_SEL_LD 192     // 3 R <= 192 selected row coordinates per frame (R <= 64)
#define UUO_SEL_IMG_FLOATS (8 * 2 * 192 * 32 + 6 * 2 * 256 * 32 + 192 + 3 * 64 * 24 + 3 * 24 * 64 + 256)
#define UUO_GA_STRIDE 292  // 288 dA entries + 3 (sum_v g_v) + 1 pad

struct uuo_model {
  int V, J, NB, P, NE, device;
  int KC;      // P + NB coefficients per frame: [pose_feature(P), betas(NB)]
  int KP;      // KC rounded up to 16 (row pitch of coef / Bt)
  int NP;      // 3V rounded up to 128 (row pitch of v_posed / Bmat)
  int max_nnz; // max non-zeros in a skinning-weight row
  int max_depth;
  // device constants
  float* vt;      // [NP]      v_template, zero padded
  float* Bmat;    // [KP][NP]  rows 0..P-1 = posedirs (lbs.py:229-231), rows P..P+NB-1 = shapedirs^T (lbs.py:295)
  float* Bt;      // [NP][KP]  transpose of Bmat (row = one vertex coordinate)
  float* J0;      // [J][3]    J_regressor @ v_template           (lbs.py:213 folded, beta independent part)
  float* Jdirs;   // [J][3][NB] J_regressor @ shapedirs
  float* W;       // [V][J]    skinning weights
  int32_t* ell_j; // [V][4]    joint ids of the 4 largest weights (valid when max_nnz <= 4)
  float* ell_w;   // [V][4]
  int32_t* extra_ids;  // [NE]
  uint8_t* vflag;      // [V] 1 when some extra joint is this vertex
  int NU;              // distinct vertices among the extra joints
  int32_t* xu_vert;    // [NU]   the vertex
  int32_t* xu_start;   // [NU+1] CSR into xu_list
  int32_t* xu_list;    // [NE]   extra-joint indices e aliasing that vertex
  int32_t* vorder;     // [V]    spatially coherent vertex permutation (nn_tiles.cu)
  float* Bimg_fwd;     // [NP/256][KP/32][2][256*32] tf32 hi|lo SWIZZLE_128B tiles of Bt   (umma_gemm.cu)
  float* Wimg;         // [NP/384][2][128*32] tf32 hi|lo SWIZZLE_128B tiles of W (row = vertex, K = joint, padded to 32)
  float* WTimg;        // [NP/96][2][32*32] tf32 hi|lo tiles of [W | 1]^T (row = joint, K = vertex)  (skin_bwdA_tc.cu)
  float* Bimg_bwd;     // [1][NP/32][2][KP*32]       tf32 hi|lo SWIZZLE_128B tiles of Bmat
  // small tables passed by value to the FK kernels
  int8_t parents[UUO_MAX_J];
  int8_t depth[UUO_MAX_J];
  int8_t children[UUO_MAX_J][UUO_MAX_CHILD];
};

// tree tables by value (kernel parameter)
struct uuo_tree {
  int8_t parents[UUO_MAX_J];
  int8_t depth[UUO_MAX_J];
  int8_t children[UUO_MAX_J][UUO_MAX_CHILD];
  int J, max_depth;
  uint8_t slot_mask[UUO_MAX_J + 1];   // per depth d: bit c set when some joint's child slot c holds a joint of depth d
};

static inline uuo_tree make_tree(const uuo_model* m) {
  uuo_tree t;
  memcpy(t.parents, m->parents, sizeof(t.parents));
  memcpy(t.depth, m->depth, sizeof(t.depth));
  memcpy(t.children, m->children, sizeof(t.children));
  t.J = m->J;
  t.max_depth = m->max_depth;
  memset(t.slot_mask, 0, sizeof(t.slot_mask));
  for (int j = 0; j < m->J; ++j)
    for (int c = 0; c < UUO_MAX_CHILD; ++c) {
      const int cj = m->children[j][c];
      if (cj >= 0) t.slot_mask[m->depth[cj]] |= (uint8_t)(1u << c);
    }
  return t;
}

// workspace carve-up (all offsets in floats, every region 16-byte aligned)
struct uuo_ws {
  int F, FP;
  float* coef;    // [FP][KP]
  float* A;       // [F][J][12]   rows of [R | t]   (lbs.py:403-405 relative transforms, 3x4 part)
  float* vposed;  // [F][NP]
  float* gvposed; // [F][NP]
  float* gA;      // [<= 2*NSPLIT_GA + 1][F][GA_STRIDE]  (+1: the extra-joint slab)
  float* gcoef;   // [NSPLIT_GC][F][KP]
  float* gbetas;  // [F][16]
  float* rowbuf;  // [F][R<=64][4]  scratch of the row-sparse paths
  float* Aimg;    // [ceil(F/16)][2][192*32] tf32 hi|lo swizzled image of the relative transforms (skin_tc.cu)
  float* vshaped; // [NP]  v_template + shapedirs . betas when ONE betas row is shared by all frames (exact fp32)
  // rows shared by all frames on the tensor cores (rows_tc.cu)
  float* sel_vp;   // [F][UUO_SEL_LD]  v_posed of the selected rows (3r + c), columns >= 3R are padding
  float* sel_gvp;  // [F][UUO_SEL_LD]  d v_posed of the selected rows, padding columns written as 0
  float* sel_img;  // [uuo_sel_images(F)][UUO_SEL_IMG_FLOATS] per row group: both operand images + bias + sparse weight tables of the current row set
  size_t total_floats;
};

// operand images of selected row sets the workspace holds: one per row group of >= 64 frames, at most 64
static inline int uuo_sel_images(int F) { return F < 128 ? 1 : (F / 64 > 64 ? 64 : F / 64); }

static inline size_t align4(size_t x) { return (x + 3) & ~(size_t)3; }

static inline uuo_ws carve_ws(const uuo_model* m, int F, void* base) {
  uuo_ws w;
  w.F = F;
  w.FP = (F + 127) / 128 * 128;
  size_t off = 0;
  float* b = (float*)base;
  auto take = [&](size_t n) { float* p = b ? b + off : nullptr; off += align4(n); return p; };
  w.coef = take((size_t)w.FP * m->KP);
  w.A = take((size_t)F * m->J * 12);
  w.vposed = take((size_t)F * m->NP);
  w.gvposed = take((size_t)F * m->NP);
  w.gA = take((size_t)(2 * UUO_NSPLIT_GA + 1) * F * UUO_GA_STRIDE);   // tensor-core path: hi and lo slab per vertex split
  w.gcoef = take((size_t)UUO_NSPLIT_GC * F * m->KP);
  w.gbetas = take((size_t)F * 16);
  w.rowbuf = take((size_t)F * 64 * 4);
  w.Aimg = take((size_t)((F + 15) / 16) * 2 * 192 * 32);
  w.vshaped = take((size_t)m->NP);
  w.sel_vp = take((size_t)F * UUO_SEL_LD);
  w.sel_gvp = take((size_t)F * UUO_SEL_LD);
  w.sel_img = take((size_t)uuo_sel_images(F) * UUO_SEL_IMG_FLOATS);
  w.total_floats = off;
  return w;
}

void uuo_set_error(const char* fmt, ...);

#define UUO_CHECK_ARG(cond, msg)                 \
  do {                                           \
    if (!(cond)) {                               \
      uuo_set_error("%s: %s", __func__, msg);    \
      return UUO_ERR_ARG;                        \
    }                                            \
  } while (0)

#define UUO_CUDA(call)                                                                   \
  do {                                                                                   \
    cudaError_t e__ = (call);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      uuo_set_error("%s: %s failed: %s", __func__, #call, cudaGetErrorString(e__));      \
      return UUO_ERR_CUDA;                                                               \
    }                                                                                    \
  } while (0)

#define UUO_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    cudaError_t e__ = cudaGetLastError();                                                \
    if (e__ != cudaSuccess) {                                                            \
      uuo_set_error("%s: kernel launch failed: %s", __func__, cudaGetErrorString(e__));  \
      return UUO_ERR_CUDA;                                                               \
    }                                                                                    \
  } while (0)

// per-device launch configuration (devcfg.cu)
int uuo_sm_count();                                          // SMs of the current device
int uuo_sm_count_of(int device);
int uuo_ensure_dyn_smem(const void* func, size_t bytes);     // opt-in to > 48 KB dynamic smem, once per (device, kernel)
#define UUO_DYN_SMEM(kernel, bytes)                                        \
  do {                                                                     \
    int rc__ = uuo_ensure_dyn_smem((const void*)(kernel), (size_t)(bytes)); \
    if (rc__) return rc__;                                                 \
  } while (0)

// per-kernel event timing (prof.cu); a scope records an event before and after the launches inside it
bool uuo_prof_enabled();
void uuo_prof_mark(const char* name, cudaStream_t st);
struct ProfScope {
  const char* name; cudaStream_t st;
  ProfScope(const char* n, cudaStream_t s) : name(n), st(s) { if (uuo_prof_enabled()) uuo_prof_mark(name, st); }
  ~ProfScope() { if (uuo_prof_enabled()) uuo_prof_mark(name, st); }
};

// ---- launchers implemented in the .cu files (host functions, return uuo_status) ----
int launch_fk_forward(const uuo_model* m, const float* betas, int betas_rows, const float* go, const float* bp,
                      int pose_mode, const float* yaw, const float* transl, int F, float* joints,
                      const uuo_ws& ws, cudaStream_t st);
int launch_fk_backward(const uuo_model* m, const float* betas, int betas_rows, const float* go, const float* bp,
                       int pose_mode, const float* yaw, int F, const float* g_joints, const uuo_ws& ws,
                       int n_ga_parts, int n_gc_parts, const float* g_rows_sum /*unused*/,
                       float* g_betas, float* g_go, float* g_bp, float* g_yaw, float* g_transl, cudaStream_t st);
// host: rows x K fp32 (K contiguous, pitch ld = K) -> [rows/NT][K/32][2][NT*32] swizzled tf32 hi|lo tiles
void uuo_build_b_image(std::vector<float>& out, const float* src, int rows, int K, int NT);
// betas_shared != NULL: ONE betas row for all frames -> v_shaped is evaluated once in exact fp32 (launch_vshaped) and
// only the pose rows go through the tensor cores
// (launch_vshaped must have filled ws.vshaped when betas_shared is set)
int launch_vposed_gemm(const uuo_model* m, int F, const uuo_ws& ws, bool betas_shared, cudaStream_t st);
int launch_vshaped(const uuo_model* m, const float* betas, const uuo_ws& ws, cudaStream_t st);
int launch_gcoef_gemm(const uuo_model* m, int F, const uuo_ws& ws, int* n_parts, cudaStream_t st);
int launch_vposed_gemm_umma(const uuo_model* m, int F, const uuo_ws& ws, bool betas_shared, cudaStream_t st);
int launch_gcoef_gemm_umma(const uuo_model* m, int F, const uuo_ws& ws, cudaStream_t st);
int uuo_gcoef_splits(const uuo_model* m, int F);
int launch_pack_transforms(const uuo_model* m, int F, const uuo_ws& ws, cudaStream_t st);
int launch_skin_tc(const uuo_model* m, int mode, int F, const float* in, const float* g_joints, const float* transl,
                   float* out, const uuo_ws& ws, cudaStream_t st);
int launch_umma_one_tile(const float* A, long lda, int M, const float* Bimg, int NT, int kb_total, const float* bias,
                         float* C, long ldc, int groups, long b_group_stride, long bias_group_stride, cudaStream_t st);
bool rows_tc_usable(const uuo_model* m, int F, int rows_shared, int R);
int launch_rows_forward_tc(const uuo_model* m, const float* transl, int F, const int32_t* rows, int G, int R,
                           float* out_rows, const uuo_ws& ws, cudaStream_t st);
int launch_rows_backward_tc(const uuo_model* m, int F, const int32_t* rows, int G, int R, const float* g_rows,
                            const uuo_ws& ws, cudaStream_t st);
int launch_skin_bwdA_tc(const uuo_model* m, int F, const float* g_vertices, const float* g_joints, const uuo_ws& ws,
                        cudaStream_t st);
int uuo_bwdA_splits(const uuo_model* m);
int uuo_ga_parts(const uuo_model* m, bool with_g_joints);
int launch_skin_forward(const uuo_model* m, const float* transl, int F, float* vertices, float* joints,
                        const uuo_ws& ws, cudaStream_t st);
int launch_skin_backward(const uuo_model* m, int F, const float* g_vertices, const float* g_joints,
                         const uuo_ws& ws, cudaStream_t st);
int launch_rows_forward(const uuo_model* m, const float* transl, int F, const int32_t* rows, int rows_shared,
                        int R, float* out_rows, const uuo_ws& ws, cudaStream_t st);
int launch_rows_backward(const uuo_model* m, int F, const int32_t* rows, int rows_shared, int R,
                         const float* g_rows, int vposed_cached, const uuo_ws& ws, cudaStream_t st);
int launch_cache_vposed_rows(const uuo_model* m, int F, const int32_t* pos, int R, const uuo_ws& ws, cudaStream_t st);
