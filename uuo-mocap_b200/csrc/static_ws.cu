// uuo_lbs_forward_static: lbs() forward for callers that cannot allocate device memory (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The workspace the forward kernels need (uuo_lbs_workspace_bytes) lives in a file-scope device
// array of UUO_STATIC_WORKSPACE_BYTES that every GPU loading this library holds until the process exits.
//
// In host code `g_static_ws` is only a stand-in for the symbol, NOT a device address: the address comes from
// cudaGetSymbolAddress, asked once per device with that device current, and that pointer is what the kernels get.
// The array is shared by every call and keeps the previous call's contents.  The forward kernels write each workspace
// region before they read it (the ordinary entry point runs on uninitialised torch.empty memory), so nothing has to be
// cleared; two calls that overlap on one GPU (different streams, or concurrent graph launches) race on it.
#include <mutex>

#include "common.cuh"

namespace {

__device__ __align__(256) unsigned char g_static_ws[UUO_STATIC_WORKSPACE_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
void* g_addr[MAX_DEVICES];      // device -> address of g_static_ws on that device (fixed table: no host allocation either)

// address of the array on the CURRENT device
int static_ws_address(int dev, void** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_addr[dev]) {
    void* p = nullptr;
    UUO_CUDA(cudaGetSymbolAddress(&p, g_static_ws));
    if (!p) { uuo_set_error("uuo_lbs_forward_static: cudaGetSymbolAddress returned NULL"); return UUO_ERR_CUDA; }
    g_addr[dev] = p;
  }
  *out = g_addr[dev];
  return UUO_OK;
}

}  // namespace

extern "C" int uuo_lbs_forward_static(const uuo_model* m, const float* betas, int betas_rows, const float* global_orient,
                                      const float* body_pose, int pose_mode, const float* yaw, const float* transl,
                                      int F, float* vertices, float* joints, void* stream) {
  UUO_CHECK_ARG(m != nullptr, "model is NULL");
  UUO_CHECK_ARG(F > 0, "F must be positive");
  const size_t need = uuo_lbs_workspace_bytes(m, F);
  if (need > (size_t)UUO_STATIC_WORKSPACE_BYTES) {
    uuo_set_error("uuo_lbs_forward_static: F = %d frames need %zu bytes of workspace, the static workspace holds %zu "
                  "(use uuo_lbs_forward with a caller-allocated workspace)", F, need, (size_t)UUO_STATIC_WORKSPACE_BYTES);
    return UUO_ERR_ARG;
  }
  int dev = -1;
  UUO_CUDA(cudaGetDevice(&dev));
  UUO_CHECK_ARG(dev >= 0 && dev < MAX_DEVICES, "device index out of range");
  UUO_CHECK_ARG(dev == m->device, "the model's device must be the current device");
  void* ws = nullptr;
  int rc = static_ws_address(dev, &ws);
  if (rc) return rc;
  return uuo_lbs_forward(m, betas, betas_rows, global_orient, body_pose, pose_mode, yaw, transl, F, vertices, joints,
                         ws, stream);
}
