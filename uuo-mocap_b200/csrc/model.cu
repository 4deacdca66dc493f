// Model constants: host-side folding and upload.  Replaces the buffer registration of
// smplx.SMPL.__init__ (third_party/smplx/smplx/body_models.py:144-269).
#include <stdarg.h>

#include <algorithm>
#include <new>
#include <vector>

#include "common.cuh"

static thread_local char g_err[512] = "";

void uuo_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* uuo_last_error(void) { return g_err; }
extern "C" int uuo_version(void) { return 100; }

template <typename T>
static int upload(T** dst, const std::vector<T>& src) {
  UUO_CUDA(cudaMalloc((void**)dst, src.size() * sizeof(T)));
  UUO_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return UUO_OK;
}

extern "C" int uuo_model_create(const float* v_template, const float* shapedirs, const float* posedirs,
                                const float* J_regressor, const float* lbs_weights, const int32_t* parents,
                                const int32_t* extra_joint_ids, int V, int J, int NB, int P, int NE,
                                int device, uuo_model** out) {
  UUO_CHECK_ARG(out != nullptr, "out is NULL");
  *out = nullptr;
  UUO_CHECK_ARG(v_template && shapedirs && posedirs && J_regressor && lbs_weights && parents, "NULL array");
  UUO_CHECK_ARG(V > 0 && V < (1 << 24), "bad V");
  UUO_CHECK_ARG(J == UUO_SKIN_J, "this build supports J == 24 (SMPL) only");
  UUO_CHECK_ARG(P == 9 * (J - 1), "P must be 9*(J-1)");
  UUO_CHECK_ARG(NB >= 1 && NB <= 16, "NB must be in [1,16]");
  UUO_CHECK_ARG(NE >= 0 && (NE == 0 || extra_joint_ids), "bad extra joints");
  UUO_CHECK_ARG(parents[0] < 0, "parents[0] must be -1");
  for (int j = 1; j < J; ++j) UUO_CHECK_ARG(parents[j] >= 0 && parents[j] < j, "parents must precede children");
  for (int e = 0; e < NE; ++e) UUO_CHECK_ARG(extra_joint_ids[e] >= 0 && extra_joint_ids[e] < V, "extra joint id out of range");

  int ndev = 0;
  UUO_CUDA(cudaGetDeviceCount(&ndev));
  UUO_CHECK_ARG(device >= 0 && device < ndev, "no such CUDA device");
  UUO_CUDA(cudaSetDevice(device));

  uuo_model* m = new (std::nothrow) uuo_model();
  if (!m) { uuo_set_error("uuo_model_create: host allocation failed"); return UUO_ERR_ALLOC; }
  memset(m, 0, sizeof(*m));
  m->V = V; m->J = J; m->NB = NB; m->P = P; m->NE = NE; m->device = device;
  m->KC = P + NB;
  m->KP = (m->KC + 31) / 32 * 32;      // K blocks of the tensor-core GEMMs are 32 fp32 (one 128 B swizzle row)
  m->NP = (3 * V + 255) / 256 * 256;   // forward GEMM column tiles are 256 wide
  const int KP = m->KP, NP = m->NP, N = 3 * V;

  // kinematic tree tables
  memset(m->children, -1, sizeof(m->children));
  int nchild[UUO_MAX_J] = {0};
  m->max_depth = 0;
  for (int j = 0; j < UUO_MAX_J; ++j) { m->parents[j] = -1; m->depth[j] = 0; }
  for (int j = 1; j < J; ++j) {
    int p = parents[j];
    m->parents[j] = (int8_t)p;
    m->depth[j] = (int8_t)(m->depth[p] + 1);
    m->max_depth = std::max(m->max_depth, (int)m->depth[j]);
    if (nchild[p] >= UUO_MAX_CHILD) {
      delete m;
      uuo_set_error("uuo_model_create: joint %d has more than %d children", p, UUO_MAX_CHILD);
      return UUO_ERR_ARG;
    }
    m->children[p][nchild[p]++] = (int8_t)j;
  }

  std::vector<float> vt(NP, 0.f), Bmat((size_t)KP * NP, 0.f), Bt((size_t)NP * KP, 0.f);
  for (int n = 0; n < N; ++n) vt[n] = v_template[n];
  for (int p = 0; p < P; ++p)
    for (int n = 0; n < N; ++n) Bmat[(size_t)p * NP + n] = posedirs[(size_t)p * N + n];
  for (int n = 0; n < N; ++n)
    for (int l = 0; l < NB; ++l) Bmat[(size_t)(P + l) * NP + n] = shapedirs[(size_t)n * NB + l];
  for (int k = 0; k < KP; ++k)
    for (int n = 0; n < NP; ++n) Bt[(size_t)n * KP + k] = Bmat[(size_t)k * NP + n];

  // fold the joint regressor (lbs.py:209-213): J = Jreg @ (v_template + shapedirs . beta)
  std::vector<float> J0((size_t)J * 3), Jdirs((size_t)J * 3 * NB);
  for (int j = 0; j < J; ++j)
    for (int k = 0; k < 3; ++k) {
      double a = 0.0;
      for (int v = 0; v < V; ++v) a += (double)J_regressor[(size_t)j * V + v] * v_template[v * 3 + k];
      J0[j * 3 + k] = (float)a;
      for (int l = 0; l < NB; ++l) {
        double s = 0.0;
        for (int v = 0; v < V; ++v)
          s += (double)J_regressor[(size_t)j * V + v] * shapedirs[((size_t)v * 3 + k) * NB + l];
        Jdirs[(j * 3 + k) * NB + l] = (float)s;
      }
    }

  std::vector<float> W((size_t)V * J);
  std::vector<int32_t> ell_j((size_t)V * 4, 0);
  std::vector<float> ell_w((size_t)V * 4, 0.f);
  m->max_nnz = 0;
  for (int v = 0; v < V; ++v) {
    int nnz = 0;
    for (int j = 0; j < J; ++j) {
      float w = lbs_weights[(size_t)v * J + j];
      W[(size_t)v * J + j] = w;
      if (w != 0.f) {
        if (nnz < 4) { ell_j[v * 4 + nnz] = j; ell_w[v * 4 + nnz] = w; }
        ++nnz;
      }
    }
    m->max_nnz = std::max(m->max_nnz, nnz);
  }
  // tensor-core operand images (umma_gemm.cu): forward B rows = vertex coordinates (K = KP),
  // backward B rows = coefficients (K = NP); both pre-split into tf32 hi | lo and pre-swizzled.
  std::vector<float> img_fwd, img_bwd;
  uuo_build_b_image(img_fwd, Bt.data(), NP, KP, 256);
  uuo_build_b_image(img_bwd, Bmat.data(), KP, NP, KP);
  UUO_CHECK_ARG(KP <= 256 && KP % 16 == 0, "KP must be <= 256");
  // skinning weights as an MMA operand: rows = the NP/3 vertex slots (zeros past V), K = 32 (J = 24 + zero pad)
  std::vector<float> Wpad((size_t)(NP / 3) * 32, 0.f), img_w;
  for (int v = 0; v < V; ++v)
    for (int j = 0; j < J; ++j) Wpad[(size_t)v * 32 + j] = lbs_weights[(size_t)v * J + j];
  uuo_build_b_image(img_w, Wpad.data(), NP / 3, 32, 128);
  // [W | 1]^T for the d(A) contraction over vertices: 32 rows (24 joints, a ones row that yields sum_v g, zero pad)
  std::vector<float> WT((size_t)32 * (NP / 3), 0.f), img_wt;
  for (int v = 0; v < V; ++v) {
    for (int j = 0; j < J; ++j) WT[(size_t)j * (NP / 3) + v] = lbs_weights[(size_t)v * J + j];
    WT[(size_t)J * (NP / 3) + v] = 1.f;
  }
  uuo_build_b_image(img_wt, WT.data(), 32, NP / 3, 32);

  // spatially coherent visiting order for the pruned correspondence kernel (nn_tiles.cu): leaves of a balanced
  // kd-tree over the template -- split the longest axis so that the left part holds a multiple of 32 vertices;
  // consecutive leaves (tiles of 32) share ancestors, so 32 consecutive tiles are a coherent region as well
  std::vector<int32_t> vorder(V);
  {
    for (int v = 0; v < V; ++v) vorder[v] = v;
    std::vector<std::pair<int, int>> stack(1, std::make_pair(0, V));
    while (!stack.empty()) {
      const int b0 = stack.back().first, e0 = stack.back().second;
      stack.pop_back();
      const int n = e0 - b0;
      if (n <= 32) continue;
      float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
      for (int i = b0; i < e0; ++i)
        for (int k = 0; k < 3; ++k) {
          const float x = v_template[vorder[i] * 3 + k];
          lo[k] = std::min(lo[k], x); hi[k] = std::max(hi[k], x);
        }
      int ax = 0;
      for (int k = 1; k < 3; ++k)
        if (hi[k] - lo[k] > hi[ax] - lo[ax]) ax = k;
      const int mid = b0 + 32 * (((n + 31) / 32) / 2);
      std::stable_sort(vorder.begin() + b0, vorder.begin() + e0, [&](int32_t a, int32_t c) {
        return v_template[a * 3 + ax] < v_template[c * 3 + ax];
      });
      stack.push_back(std::make_pair(b0, mid));
      stack.push_back(std::make_pair(mid, e0));
    }
  }

  std::vector<uint8_t> vflag(NP / 3, 0);   // padded: read 4 flags at a time
  std::vector<int32_t> ex(std::max(NE, 1), 0);
  for (int e = 0; e < NE; ++e) { ex[e] = extra_joint_ids[e]; vflag[ex[e]] = 1; }

  // distinct vertices among the extra joints (CSR), for the extra-joint backward kernels
  std::vector<int32_t> xu_vert, xu_start(1, 0), xu_list;
  for (int e = 0; e < NE; ++e) {
    bool seen = false;
    for (int q = 0; q < e; ++q) seen |= ex[q] == ex[e];
    if (seen) continue;
    xu_vert.push_back(ex[e]);
    for (int q = e; q < NE; ++q)
      if (ex[q] == ex[e]) xu_list.push_back(q);
    xu_start.push_back((int32_t)xu_list.size());
  }
  m->NU = (int)xu_vert.size();
  if (xu_vert.empty()) { xu_vert.push_back(0); xu_list.push_back(0); }

  int rc = UUO_OK;
  if ((rc = upload(&m->xu_vert, xu_vert)) || (rc = upload(&m->xu_start, xu_start)) || (rc = upload(&m->xu_list, xu_list))) {
    uuo_model_destroy(m);
    return rc;
  }
  if ((rc = upload(&m->vt, vt)) || (rc = upload(&m->Bmat, Bmat)) || (rc = upload(&m->Bt, Bt)) ||
      (rc = upload(&m->J0, J0)) || (rc = upload(&m->Jdirs, Jdirs)) || (rc = upload(&m->W, W)) ||
      (rc = upload(&m->ell_j, ell_j)) || (rc = upload(&m->ell_w, ell_w)) ||
      (rc = upload(&m->extra_ids, ex)) || (rc = upload(&m->vflag, vflag)) ||
      (rc = upload(&m->Bimg_fwd, img_fwd)) || (rc = upload(&m->Bimg_bwd, img_bwd)) ||
      (rc = upload(&m->Wimg, img_w)) || (rc = upload(&m->WTimg, img_wt)) || (rc = upload(&m->vorder, vorder))) {
    uuo_model_destroy(m);
    return rc;
  }
  *out = m;
  return UUO_OK;
}

extern "C" void uuo_model_destroy(uuo_model* m) {
  if (!m) return;
  cudaFree(m->vt); cudaFree(m->Bmat); cudaFree(m->Bt); cudaFree(m->J0); cudaFree(m->Jdirs);
  cudaFree(m->W); cudaFree(m->ell_j); cudaFree(m->ell_w); cudaFree(m->extra_ids); cudaFree(m->vflag);
  cudaFree(m->Bimg_fwd); cudaFree(m->Bimg_bwd); cudaFree(m->Wimg); cudaFree(m->WTimg); cudaFree(m->xu_vert); cudaFree(m->xu_start); cudaFree(m->xu_list);
  cudaFree(m->vorder);
  delete m;
}

extern "C" const int32_t* uuo_model_vertex_order(const uuo_model* m) { return m ? m->vorder : nullptr; }

extern "C" int uuo_model_vertex_order_host(const uuo_model* m, int32_t* out) {
  UUO_CHECK_ARG(m && out, "NULL pointer");
  UUO_CUDA(cudaMemcpy(out, m->vorder, (size_t)m->V * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return UUO_OK;
}

extern "C" int uuo_model_info(const uuo_model* m, int what) {
  if (!m) return UUO_ERR_ARG;
  switch (what) {
    case 0: return m->V;
    case 1: return m->J;
    case 2: return m->NB;
    case 3: return m->P;
    case 4: return m->NE;
    case 5: return m->device;
    case 6: return m->max_nnz;
    case 7: return m->KP;
    case 8: return m->NP;
    default: return UUO_ERR_ARG;
  }
}

extern "C" size_t uuo_lbs_workspace_bytes(const uuo_model* m, int F) {
  if (!m || F <= 0) return 0;
  uuo_ws w = carve_ws(m, F, nullptr);
  return w.total_floats * sizeof(float);
}
