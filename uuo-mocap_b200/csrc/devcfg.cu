// Per-device launch configuration.  cudaFuncSetAttribute and the SM count belong to a DEVICE, and one process may
// hold models on several GPUs (uuo_model_create takes a device index), so both are cached per device under a mutex
// instead of in process-wide statics.
#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

static std::mutex g_mu;
static std::map<int, int> g_sms;                                        // device -> SM count
static std::map<std::pair<int, const void*>, size_t> g_dyn_smem;        // (device, kernel) -> opted-in bytes

int uuo_sm_count_of(int device) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_sms.find(device);
  if (it != g_sms.end()) return it->second;
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return 1;                     // no usable device: the launch that follows reports the real error
  }
  g_sms[device] = n;
  return n;
}

int uuo_sm_count() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 1; }
  return uuo_sm_count_of(dev);
}

int uuo_ensure_dyn_smem(const void* func, size_t bytes) {
  int dev = 0;
  UUO_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_mu);
  size_t& have = g_dyn_smem[std::make_pair(dev, func)];
  if (bytes > have) {
    UUO_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    have = bytes;
  }
  return UUO_OK;
}
