// extern "C" entry points of the SMPL layer (include/uuo_mocap.h): argument checks + kernel sequencing.
#include "common.cuh"

static int check_common(const uuo_model* m, const float* betas, int betas_rows, const float* go, const float* bp,
                        int pose_mode, int F, void* workspace) {
  UUO_CHECK_ARG(m != nullptr, "model is NULL");
  UUO_CHECK_ARG(F > 0, "F must be positive");
  UUO_CHECK_ARG(betas && go && bp, "betas / global_orient / body_pose is NULL");
  UUO_CHECK_ARG(betas_rows >= 1 && betas_rows <= F && F % betas_rows == 0, "betas_rows must be 1, F or a divisor of F");
  UUO_CHECK_ARG(pose_mode >= 0 && pose_mode <= 2, "bad pose_mode");
  UUO_CHECK_ARG(workspace != nullptr, "workspace is NULL");
  UUO_CHECK_ARG(((uintptr_t)workspace & 15) == 0, "workspace must be 16-byte aligned");
  return UUO_OK;
}

extern "C" int uuo_lbs_forward(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                               const float* body_pose, int pose_mode, const float* yaw, const float* transl, int F,
                               float* vertices, float* joints, void* workspace, void* stream) {
  int rc = check_common(m, betas, betas_rows, global_orient, body_pose, pose_mode, F, workspace);
  if (rc) return rc;
  UUO_CHECK_ARG(vertices != nullptr, "vertices is NULL");
  cudaStream_t st = (cudaStream_t)stream;
  uuo_ws ws = carve_ws(m, F, workspace);
  if ((rc = launch_fk_forward(m, betas, betas_rows, global_orient, body_pose, pose_mode, yaw, transl, F, joints, ws, st))) return rc;
  const bool shared = betas_rows == 1;
  if (shared && (rc = launch_vshaped(m, betas, ws, st))) return rc;
  // (r2: running GEMM -> skinning in 1024-frame chunks so that v_posed is read back from L2 was measured and lost:
  // vposed_gemm 0.187 -> 0.289 ms, skin_forward 0.176 -> 0.225 ms at F = 5000 -- the persistent kernels' tails
  // (648 items over 148 SMs per chunk) cost more than the HBM read saves)
  if ((rc = launch_vposed_gemm(m, F, ws, shared, st))) return rc;
  if ((rc = launch_skin_forward(m, transl, F, vertices, joints, ws, st))) return rc;
  return UUO_OK;
}

extern "C" int uuo_lbs_backward(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                                const float* body_pose, int pose_mode, const float* yaw, int F,
                                const float* g_vertices, const float* g_joints, void* workspace, float* g_betas,
                                float* g_global_orient, float* g_body_pose, float* g_yaw, float* g_transl,
                                void* stream) {
  int rc = check_common(m, betas, betas_rows, global_orient, body_pose, pose_mode, F, workspace);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  uuo_ws ws = carve_ws(m, F, workspace);
  if ((rc = launch_skin_backward(m, F, g_vertices, g_joints, ws, st))) return rc;
  int n_gc = 0;
  if ((rc = launch_gcoef_gemm(m, F, ws, &n_gc, st))) return rc;
  return launch_fk_backward(m, betas, betas_rows, global_orient, body_pose, pose_mode, yaw, F, g_joints, ws,
                            uuo_ga_parts(m, g_joints != nullptr), n_gc, nullptr, g_betas, g_global_orient, g_body_pose, g_yaw,
                            g_transl, st);
}

extern "C" int uuo_lbs_forward_rows(const uuo_model* m, const float* betas, int betas_rows,
                                    const float* global_orient, const float* body_pose, int pose_mode,
                                    const float* yaw, const float* transl, int F, const int32_t* rows,
                                    int rows_shared, int R, float* out_rows, float* joints, void* workspace,
                                    void* stream) {
  int rc = check_common(m, betas, betas_rows, global_orient, body_pose, pose_mode, F, workspace);
  if (rc) return rc;
  UUO_CHECK_ARG(rows && out_rows && R > 0, "rows / out_rows is NULL or R <= 0");
  UUO_CHECK_ARG(joints == nullptr || m->NE == 0, "joints output of the rows path would need the extra-joint vertices: pass NULL");
  cudaStream_t st = (cudaStream_t)stream;
  uuo_ws ws = carve_ws(m, F, workspace);
  if ((rc = launch_fk_forward(m, betas, betas_rows, global_orient, body_pose, pose_mode, yaw, transl, F, joints, ws, st))) return rc;
  return launch_rows_forward(m, transl, F, rows, rows_shared, R, out_rows, ws, st);
}

static int backward_rows_impl(int vposed_cached, const uuo_model* m, const float* betas, int betas_rows,
                                     const float* global_orient, const float* body_pose, int pose_mode,
                                     const float* yaw, int F, const int32_t* rows, int rows_shared, int R,
                                     const float* g_rows, const float* g_joints, void* workspace, float* g_betas,
                                     float* g_global_orient, float* g_body_pose, float* g_yaw, float* g_transl,
                                     void* stream) {
  int rc = check_common(m, betas, betas_rows, global_orient, body_pose, pose_mode, F, workspace);
  if (rc) return rc;
  UUO_CHECK_ARG(rows && g_rows && R > 0, "rows / g_rows is NULL or R <= 0");
  UUO_CHECK_ARG(g_joints == nullptr || m->NE == 0, "g_joints on the rows path would touch the extra-joint vertices: pass NULL");
  cudaStream_t st = (cudaStream_t)stream;
  uuo_ws ws = carve_ws(m, F, workspace);
  if ((rc = launch_rows_backward(m, F, rows, rows_shared, R, g_rows, vposed_cached, ws, st))) return rc;
  return launch_fk_backward(m, betas, betas_rows, global_orient, body_pose, pose_mode, yaw, F, g_joints, ws, 1, 1,
                            nullptr, g_betas, g_global_orient, g_body_pose, g_yaw, g_transl, st);
}

#define UUO_ROWS_ARGS m, betas, betas_rows, global_orient, body_pose, pose_mode, yaw, F, rows, rows_shared, R, g_rows, g_joints, \
                      workspace, g_betas, g_global_orient, g_body_pose, g_yaw, g_transl, stream

extern "C" int uuo_lbs_backward_rows(const uuo_model* m, const float* betas, int betas_rows,
                                     const float* global_orient, const float* body_pose, int pose_mode,
                                     const float* yaw, int F, const int32_t* rows, int rows_shared, int R,
                                     const float* g_rows, const float* g_joints, void* workspace, float* g_betas,
                                     float* g_global_orient, float* g_body_pose, float* g_yaw, float* g_transl,
                                     void* stream) {
  return backward_rows_impl(0, UUO_ROWS_ARGS);
}

extern "C" int uuo_lbs_backward_rows_cached(const uuo_model* m, const float* betas, int betas_rows,
                                            const float* global_orient, const float* body_pose, int pose_mode,
                                            const float* yaw, int F, const int32_t* rows, int rows_shared, int R,
                                            const float* g_rows, const float* g_joints, void* workspace, float* g_betas,
                                            float* g_global_orient, float* g_body_pose, float* g_yaw, float* g_transl,
                                            void* stream) {
  UUO_CHECK_ARG(R <= 64, "the cached v_posed rows hold at most 64 rows per frame");
  return backward_rows_impl(1, UUO_ROWS_ARGS);
}

extern "C" int uuo_lbs_cache_vposed_rows(const uuo_model* m, int F, const int32_t* pos, int R, void* workspace,
                                         void* stream) {
  UUO_CHECK_ARG(m && pos && workspace && F > 0 && R > 0 && R <= 64, "bad arguments (R must be in [1,64])");
  uuo_ws ws = carve_ws(m, F, workspace);
  return launch_cache_vposed_rows(m, F, pos, R, ws, (cudaStream_t)stream);
}
