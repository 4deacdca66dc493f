/* uuo_mocap.h -- C-ABI of libuuomocap.so: the B200-native (sm_100a) SMPL marker-fitting hot path.
 *
 * Every entry point replaces a reference call site (paths relative to /root/reference); the
 * reference has no FFI of its own (it is pure Python/PyTorch), so these are the functions a
 * ctypes binding inside src/video_mocap/utils/smpl.py and src/video_mocap/optimization.py
 * would bind.
 *
 * Conventions
 *   - every function returns 0 on success, a negative uuo_status on failure; uuo_last_error()
 *     gives a thread-local message.  Nothing throws across the boundary.
 *   - all tensor pointers are CALLER-OWNED DEVICE memory on the model's device, contiguous
 *     row-major fp32 (indices int32) unless the parameter says "host".  The library owns only
 *     the uuo_model constants.  Workspaces are caller-allocated (uuo_*_workspace_bytes).
 *   - functions enqueue on `stream` (a cudaStream_t passed as void*) and return without
 *     synchronising; a model is immutable after creation and may be shared across streams.
 *   - there is no CPU fallback: without a CUDA device every compute call fails with
 *     UUO_ERR_CUDA.
 */
#ifndef UUO_MOCAP_H
#define UUO_MOCAP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct uuo_model uuo_model;

enum uuo_status {
  UUO_OK = 0,
  UUO_ERR_ARG = -1,     /* bad argument (null pointer, bad size, unsupported shape) */
  UUO_ERR_CUDA = -2,    /* a CUDA runtime call or kernel launch failed */
  UUO_ERR_ALLOC = -3    /* device or host allocation failed */
};

/* How the pose arrays are to be read.
 *   UUO_POSE_ROTMAT : [F,J,3,3] matrices used as they are        (smplx pose2rot=False,
 *                     third_party/smplx/smplx/lbs.py:226-231)
 *   UUO_POSE_AXISANGLE : [F,J,3] axis-angle through batch_rodrigues (pose2rot=True, lbs.py:218-225,299-333)
 *   UUO_POSE_RAW6D  : [F,J,3,3] free 3x3 parameters re-orthonormalised by pytorch3d's
 *                     rotation_6d_to_matrix(matrix_to_rotation_6d(X)) (rows 0,1 Gram-Schmidt) as
 *                     the fitting closures do before every SMPL call
 *                     (src/video_mocap/optimization.py:196-202,335-340), fused into the kernel. */
enum uuo_pose_mode { UUO_POSE_ROTMAT = 0, UUO_POSE_AXISANGLE = 1, UUO_POSE_RAW6D = 2 };

const char* uuo_last_error(void);
int uuo_version(void);

/* ---- model: replaces smplx.SMPL.__init__ buffer registration (body_models.py:144-269) ----
 * HOST pointers: v_template[V,3], shapedirs[V,3,NB], posedirs[P,3V] (P = 9*(J-1), the layout
 * smplx registers), J_regressor[J,V], lbs_weights[V,J], parents[J] (parents[0] = -1),
 * extra_joint_ids[NE] (vertex_joint_selector.py:36-68).  Copies everything to `device`. */
int uuo_model_create(const float* v_template, const float* shapedirs, const float* posedirs,
                     const float* J_regressor, const float* lbs_weights, const int32_t* parents,
                     const int32_t* extra_joint_ids, int V, int J, int NB, int P, int NE, int device,
                     uuo_model** out);
void uuo_model_destroy(uuo_model* m);
/* what: 0=V 1=J 2=NB 3=P 4=NE 5=device 6=max nnz per skinning-weight row */
int uuo_model_info(const uuo_model* m, int what);

/* ---- SMPL layer: replaces smplx.SMPL.forward -> lbs() (body_models.py:315-399, lbs.py:156-252)
 * and its autograd backward.  The workspace written by forward is what backward reads. */
size_t uuo_lbs_workspace_bytes(const uuo_model* m, int F);

/* betas[betas_rows,NB], betas_rows = 1 (one row shared by all frames, body_models.py:371-376), F (one per frame), or a
 * divisor G of F (G stacked hypothesis groups of F/G consecutive frames, each with its own row: the batched MHT loop of
 * multimodal.py:462-606); gradients g_betas[betas_rows,NB] are summed over the frames of a row;
 * global_orient[F,1,...], body_pose[F,J-1,...] per pose_mode; yaw[F] or NULL: when given the root
 * matrix is Rz(yaw[f]) @ global_orient[f] (optimization.py:192,672-679) before pose_mode is applied;
 * transl[F,3] or NULL.  Outputs vertices[F,V,3], joints[F,J+NE,3]. */
int uuo_lbs_forward(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                    const float* body_pose, int pose_mode, const float* yaw, const float* transl, int F,
                    float* vertices, float* joints, void* workspace, void* stream);

/* g_vertices[F,V,3] / g_joints[F,J+NE,3] may be NULL (treated as zero).  Outputs (any may be NULL):
 * g_betas[betas_rows,NB], g_global_orient / g_body_pose (shape of the inputs), g_yaw[F], g_transl[F,3].
 * Inputs of the forward call must be passed again unchanged. */
/* lbs() forward that allocates NOTHING and takes no workspace (no counterpart in the reference: added
 * for callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, results and errors as uuo_lbs_forward; the workspace is a file-scope
 * device array of UUO_STATIC_WORKSPACE_BYTES (8 MiB, held by every GPU that loads the library until the process
 * exits), which is enough for uuo_lbs_workspace_bytes(m, F) of about 35 frames of an SMPL-sized model; a larger F is
 * rejected with UUO_ERR_ARG and a message that states both sizes.  The model's device must be current.  Every call
 * shares that array (it keeps the previous call's contents; the forward writes each region before reading it, so
 * nothing needs clearing): calls that overlap on one GPU race, serialise them on one stream.  A uuo_lbs_backward on
 * the same frames needs a caller-allocated workspace and uuo_lbs_forward. */
#define UUO_STATIC_WORKSPACE_BYTES (8u << 20)
int uuo_lbs_forward_static(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                           const float* body_pose, int pose_mode, const float* yaw, const float* transl, int F,
                           float* vertices, float* joints, void* stream);
int uuo_lbs_backward(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                     const float* body_pose, int pose_mode, const float* yaw, int F,
                     const float* g_vertices, const float* g_joints, void* workspace,
                     float* g_betas, float* g_global_orient, float* g_body_pose, float* g_yaw,
                     float* g_transl, void* stream);

/* Backward for a gradient that touches only R vertex rows per frame (what the chamfer and marker
 * losses produce): g_rows[F,R,3].  Same outputs as uuo_lbs_backward; never touches a dense [F,V,3] gradient.
 * rows_shared selects the layout of rows (here and in the other *_rows entry points):
 *   0: rows[F,R], one set per frame;  1: rows[R], the same set for every frame;
 *   G > 1: rows[G,R], frame f uses set f / (F/G) (G hypotheses stacked along the frame axis, F % G == 0). */
int uuo_lbs_backward_rows(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                          const float* body_pose, int pose_mode, const float* yaw, int F,
                          const int32_t* rows, int rows_shared, int R, const float* g_rows,
                          const float* g_joints, void* workspace,
                          float* g_betas, float* g_global_orient, float* g_body_pose, float* g_yaw,
                          float* g_transl, void* stream);

/* The same backward when the workspace already holds v_posed of these rows (R <= 64), which saves re-evaluating
 * the 3R blend-shape dot products: either uuo_lbs_forward_rows ran on this workspace with the SAME rows, or
 * uuo_lbs_cache_vposed_rows picked them out of the v_posed a dense uuo_lbs_forward left in the workspace
 * (pos[F,R] = row of that forward's vertex array, e.g. the `pos` output of uuo_nn_argmin_presorted). */
int uuo_lbs_backward_rows_cached(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                                 const float* body_pose, int pose_mode, const float* yaw, int F,
                                 const int32_t* rows, int rows_shared, int R, const float* g_rows,
                                 const float* g_joints, void* workspace,
                                 float* g_betas, float* g_global_orient, float* g_body_pose, float* g_yaw,
                                 float* g_transl, void* stream);
int uuo_lbs_cache_vposed_rows(const uuo_model* m, int F, const int32_t* pos, int R, void* workspace, void* stream);

/* Forward of R selected vertices only (marker stage, optimization.py:345-351 is a gather):
 * out_rows[F,R,3]; joints[F,J+NE,3] may be NULL. */
int uuo_lbs_forward_rows(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                         const float* body_pose, int pose_mode, const float* yaw, const float* transl, int F,
                         const int32_t* rows, int rows_shared, int R, float* out_rows, float* joints,
                         void* workspace, void* stream);

/* ---- correspondence: replaces pytorch3d knn_points(K=1) inside chamfer_distance as called by
 * src/video_mocap/losses/chamfer_distance.py:5-21.  dist2[F,M] = min_v ((dx*dx+dy*dy)+dz*dz) with every
 * operation rounded to fp32 (no FMA), idx[F,M] = the FIRST vertex attaining it. */
int uuo_nn_argmin(const float* markers, const float* vertices, int F, int M, int V, float* dist2,
                  int32_t* idx, void* stream);

/* Same outputs as uuo_nn_argmin, bit for bit, without scanning every vertex: `order[V]` (device) is a
 * permutation of the vertex ids in which consecutive entries are spatially close (uuo_model_vertex_order
 * returns the model's: dominant joint, then Morton code of the template).  Tiles of 32 consecutive entries
 * are bounded per frame by their axis-aligned box; a tile is evaluated only when the distance to its box,
 * computed with the exact distance's own fp32 operation sequence (monotone rounding => a true lower
 * bound), does not exceed the best distance so far.  order == NULL falls back to uuo_nn_argmin. */
int uuo_nn_argmin_ordered(const float* markers, const float* vertices, const int32_t* order, int F, int M,
                          int V, float* dist2, int32_t* idx, void* stream);
/* device pointer owned by the model, [V] int32 */
const int32_t* uuo_model_vertex_order(const uuo_model* m);
/* the same permutation copied to HOST memory out[V] */
int uuo_model_vertex_order_host(const uuo_model* m, int32_t* out);

/* Same outputs again for vertices that ALREADY lie in visiting order: the fitting loop evaluates lbs() on a twin
 * model whose vertices are renumbered by uuo_model_vertex_order, so a tile is 32 consecutive vertices and nothing
 * is gathered.  ids[V] (device, or NULL = identity) maps a position to the ORIGINAL vertex id: idx[F,M] reports
 * original ids and ties break on them, exactly like uuo_nn_argmin on the un-permuted vertices; pos[F,M] (may be
 * NULL) receives the winning position, i.e. the row of `vertices` that holds the matched vertex. */
int uuo_nn_argmin_presorted(const float* markers, const float* vertices, const int32_t* ids, int F, int M, int V,
                            float* dist2, int32_t* idx, int32_t* pos, void* stream);
/* The same search as the fitting loop runs it, iteration after iteration (weighted_chamfer_distance,
 * losses/chamfer_distance.py:5-21, inside closure_stage_chamfer, optimization.py:232-239).  Both extras may be NULL.
 *   weights[F,M]: the per-marker chamfer weights (the marker mask).  A marker whose weight is exactly 0 multiplies
 *     its distance by 0 in the loss and in every gradient, so it is not searched: dist2 = 0, pos = 0, idx = ids[0].
 *   hint[F,M]: any guess of pos[F,M], normally the previous iteration's output (hint == pos is allowed: the same
 *     buffer in and out).  The tile of the hinted row is evaluated first, which tightens the pruning bound; values
 *     outside [0,V) are ignored.  The outputs never depend on the hint.
 * Every marker of non-zero weight gets exactly the outputs of uuo_nn_argmin_presorted. */
int uuo_nn_argmin_presorted_fit(const float* markers, const float* vertices, const int32_t* ids, const float* weights,
                                const int32_t* hint, int F, int M, int V, float* dist2, int32_t* idx, int32_t* pos,
                                void* stream);

/* get_marker_mask (optimization.py:703-715): mask[F,M] = (|x|+|y|+|z| != 0) as float 0/1. */
int uuo_marker_mask(const float* markers, int n_markers_total, float* mask, void* stream);

/* weighted single-directional chamfer value and its gradient rows
 * (chamfer_distance.py:5-21 + pytorch3d chamfer_distance weights / batch_reduction="mean"):
 * loss_sum[0] += scale * sum_{f,m} w*d2 ; g_rows[F,M,3] = scale * w * (-2) * (marker - vertex[idx]),
 * where the caller passes scale = weight / sum(w) (sum over ALL shards).
 */
int uuo_chamfer_rows(const float* markers, const float* vertices, const float* mask, const int32_t* idx,
                     int F, int M, int V, float scale, float* loss_sum, float* g_rows, void* stream);

/* MarkerLoss (src/video_mocap/losses/losses.py:43-51) on gathered rows virt[F,M,3]:
 * loss_sum[0] += scale * sum (||m - v|| - dist)^2 * w ; g_virt[F,M,3] its gradient * scale. */
int uuo_marker_loss_rows(const float* markers, const float* virt, const float* mask, int n, float dist,
                         float scale, float* loss_sum, float* g_virt, void* stream);

/* compute_nearest_points(use_mean) (optimization.py:464-486,595-603): acc[M,V] += sum over the
 * frames with frame_mask==1 of ||vertex - marker||_2 (frames added in order), then
 * uuo_mean_argmin divides by count and takes the first argmin per marker. */
int uuo_mean_dist_accumulate(const float* markers, const float* vertices, const float* frame_mask, int F,
                             int M, int V, float* acc, void* stream);
int uuo_mean_argmin(const float* acc, int M, int V, float count, int32_t* idx, void* stream);
/* The same accumulation for callers that only want the per-marker argmin (the reference keeps nothing else of the
 * [M,V] matrix: optimization.py:596-603).  bound[M] must be >= the smallest COMPLETE sum of each marker (over all the
 * frames that will be added to acc, on every shard).  A sum that has passed bound[m] cannot be the minimum any more
 * (the terms are >= 0 and fp32 addition is monotone), so the kernel stops evaluating it: on return an entry is either
 * bit-equal to what uuo_mean_dist_accumulate leaves there, or a partial sum > bound[m].  The argmin is unchanged.
 * uuo_mean_dist_candidate_sums gives the material for the bound: sums[m] (fp64, device) = the sum of the distances
 * marker m <-> vertex cand[m] over the frames with frame_mask == 1; shards add their sums, and
 * bound = sum * (1 + frames * 2^-23) covers the rounding of the in-order fp32 accumulation. */
int uuo_mean_dist_accumulate_bounded(const float* markers, const float* vertices, const float* frame_mask, int F,
                                     int M, int V, const float* bound, float* acc, void* stream);
int uuo_mean_dist_candidate_sums(const float* markers, const float* vertices, const float* frame_mask, int F, int M,
                                 int V, const int32_t* cand, double* sums, void* stream);

/* ---- prior / regulariser terms of the closures (optimization.py:243-256,363-374) ----
 * g[i] += scale * 2 * (x[i]-y[i]) ; loss_sum[0] += scale * sum (x-y)^2 (caller folds 1/numel in scale) */
int uuo_sqdiff_accumulate(const float* x, const float* y, size_t n, float scale, float* loss_sum, float* g,
                          void* stream);
/* "temporal" term exactly as written (optimization.py:368-374): e_t = x[t+2] - 2 x[t+1] - x[t], x is
 * [F,D] (one shard of the sequence).  prev2[2,D] = the two rows before row 0 and next2[2,D] = the two
 * rows after row F-1 as received from the neighbouring shards (NULL at the ends of the sequence); with
 * them every shard computes the full gradient of its own rows, so no gradient is sent back.  The shard
 * adds to loss_sum the terms e_t whose t is one of its rows. */
int uuo_temporal_accumulate(const float* x, const float* prev2, const float* next2, int F, int D, float scale,
                            float* loss_sum, float* g, void* stream);
/* mm[F,3] = mean over ALL M markers of the frame (optimization.py:248; dropped markers included). */
/* Part fits (src/video_mocap/markers/markers_utils.py:441-444, 463: chamfer against vertices[:, vertex_indices] of one
 * candidate sub-tree): G candidates are stacked along the frame axis, frames_per_group frames each; keep[G][V] (uint8, row
 * numbering of `vertices`) marks a candidate's vertex subset.  Rows that are not kept are overwritten with far_value, so
 * the nearest-vertex kernels, run over the whole array, return the nearest KEPT vertex. */
int uuo_mask_vertices(float* vertices, const uint8_t* keep, int F, int V, int frames_per_group, float far_value, void* stream);
int uuo_marker_mean(const float* markers, int F, int M, float* mm, void* stream);
/* "trans_vel" (optimization.py:246-250): mse(trans[1:]-trans[:-1], mm[1:]-mm[:-1]).  prev[6] / next[6] =
 * [trans(3), mm(3)] of the frame before row 0 / after row F-1 (NULL at the ends of the sequence). */
int uuo_transvel_accumulate(const float* trans, const float* mm, const float* prev, const float* next, int F,
                            float scale, float* loss_sum, float* g_trans, void* stream);

/* ---- frame-range shards over peer-mapped device memory (NVLink P2P; csrc/peer.cu) ----
 * Replaces nothing in the reference (it is single-device): this is the exchange step the
 * sharded closures need -- the sum over shards of [d/d betas | loss] (betas are shared by all frames,
 * src/video_mocap/optimization.py:165-174,255-258) and the 1-/2-frame halo of the temporal terms (:246-250,368-374) --
 * done by kernels that write into the neighbour's mailbox, so an iteration holds no host-issued collective and can be
 * captured in a CUDA graph on every rank.  One uuo_comm per rank; all ranks issue the same sequence of calls.
 *   uuo_comm_create          allocates this rank's mailbox on `device` (library-owned memory)
 *   uuo_comm_export          host buffer of uuo_comm_handle_bytes() bytes <- the mailbox's CUDA IPC handle
 *   uuo_comm_connect         handles_host[world][handle_bytes] of all ranks (other processes; own entry ignored)
 *   uuo_comm_connect_local   peers[world] created in THIS process (tests, or one process driving several GPUs)
 *   uuo_comm_all_reduce_sum  x[n] (n <= 1024) <- sum over ranks, added in rank order on every rank
 *   uuo_comm_halo_exchange   x[F,D]: prev[r*ld + col + d] <- the left neighbour's last n rows, next[...] <- the right
 *                            neighbour's first n rows (n*D <= 1024); prev / next untouched at the ends of the sequence
 *   uuo_comm_group_*         the same two exchanges for ALL ranks of a world that lives on one device, as ONE cooperative
 *                            launch (block r = rank r): separate launches that wait on one another are not guaranteed to be
 *                            co-resident on one GPU.  comms[world] linked with uuo_comm_connect_local; arrays indexed by rank */
typedef struct uuo_comm uuo_comm;
int uuo_comm_create(int rank, int world, int device, uuo_comm** out);
int uuo_comm_set_timeout(uuo_comm* c, double seconds);   /* a wait on a peer longer than this traps (default 60 s) */
int uuo_comm_handle_bytes(void);
int uuo_comm_export(uuo_comm* c, void* handle_host);
int uuo_comm_connect(uuo_comm* c, const void* handles_host);
int uuo_comm_connect_local(uuo_comm* c, uuo_comm* const* peers);
void uuo_comm_destroy(uuo_comm* c);
int uuo_comm_all_reduce_sum(uuo_comm* c, float* x, int n, void* stream);
int uuo_comm_halo_exchange(uuo_comm* c, const float* x, int F, int D, int n, float* prev, float* next, int ld, int col,
                           void* stream);
int uuo_comm_group_all_reduce_sum(uuo_comm* const* comms, float* const* xs, int n, int world, void* stream);
int uuo_comm_group_halo_exchange(uuo_comm* const* comms, const float* const* xs, const int* Fs, int D, int n,
                                 float* const* prevs, float* const* nexts, int ld, int col, int world, void* stream);

/* ---- fused Adam over a flat parameter block (north-star optimiser; torch.optim.Adam semantics) */
int uuo_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                  float beta1, float beta2, float eps, int step, void* stream);

/* The same update with the step number in DEVICE memory: uses *step_counter + 1 as `step`, then increments the counter
 * on the stream.  Nothing on the host changes from one iteration to the next, so closure + update can be captured
 * once in a CUDA graph and replayed (fit.run_adam). */
int uuo_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                      float beta1, float beta2, float eps, int32_t* step_counter, void* stream);

/* ---- evaluation metrics, the step right after the path (src/video_mocap/evaluation/metrics.py) ----
 * out[0] = mean over frames and selected points of ||a - b||_2 (compute_MPJPE / compute_V2V, :49-62,126-135; sel[n_sel]
 * = point ids or NULL for all P); velocity != 0: on (x[t+1]-x[t])*freq instead (compute_MPJVE, :65-87).
 * scratch: 1024 floats of device memory. */
int uuo_point_error_mean(const float* a, const float* b, int F, int P, const int32_t* sel, int n_sel, int velocity,
                         float freq, float* scratch, float* out, void* stream);
/* compute_similarity_transform (:138-190): out[F,N,3] = S1 after the similarity transform (scale, rotation with
 * det +1, translation) that brings it closest to S2, one 3x3 SVD per frame. */
int uuo_similarity_transform(const float* S1, const float* S2, int F, int N, float* out, void* stream);
/* compute_marker_to_surface_distance (:26-46, igl.signed_distance there): dist[F,M] = unsigned distance from every
 * point to the closest triangle of the frame's mesh; faces[NF,3] int32 shared by all frames. */
int uuo_point_mesh_distance(const float* points, const float* vertices, const int32_t* faces, int F, int M, int V,
                            int NF, float* dist, void* stream);

/* ---- measurement aid (bench.py): per-kernel CUDA-event timing on the launching stream.
 * uuo_profile_enable(1) clears and starts recording, (0) stops.  uuo_profile_read(slot, ...) returns the
 * number of slots; for a valid slot it waits for its events and reports the summed duration and count. */
int uuo_profile_enable(int on);
int uuo_profile_read(int slot, char* name, int name_cap, float* total_ms, int* count);

#ifdef __cplusplus
}
#endif
#endif /* UUO_MOCAP_H */
