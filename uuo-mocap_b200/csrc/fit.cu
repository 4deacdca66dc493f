// Loss terms of the fitting closures (src/video_mocap/optimization.py:187-275, 329-394) on row-sparse
// data, and the fused Adam step.  Everything here is O(F*M) or O(F*D): tiny next to the LBS kernels.
#include "common.cuh"

__device__ __forceinline__ float block_sum_256(float v) {
  __shared__ float sh[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x < 8) r = sh[threadIdx.x];
  if (threadIdx.x < 32) {
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  }
  return r;   // valid in thread 0
}

// optimization.py:703-715
__global__ void marker_mask_kernel(const float* __restrict__ markers, int n, float* __restrict__ mask) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = markers + (size_t)i * 3;
  mask[i] = ((fabsf(p[0]) + fabsf(p[1])) + fabsf(p[2])) != 0.f ? 1.f : 0.f;
}

__global__ void __launch_bounds__(256) chamfer_rows_kernel(const float* __restrict__ markers,
                                                           const float* __restrict__ vertices,
                                                           const float* __restrict__ mask,
                                                           const int32_t* __restrict__ idx, int F, int M, int V,
                                                           float scale, float* __restrict__ loss_sum,
                                                           float* __restrict__ g_rows) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  float l = 0.f;
  if (i < F * M) {
    const int f = i / M;
    const float w = mask ? mask[i] : 1.f;
    const float* mp = markers + (size_t)i * 3;
    const float* vp = vertices + ((size_t)f * V + idx[i]) * 3;
    const float dx = mp[0] - vp[0], dy = mp[1] - vp[1], dz = mp[2] - vp[2];
    l = scale * w * ((dx * dx + dy * dy) + dz * dz);
    if (g_rows) {
      const float c = -2.f * scale * w;       // d/dv of |m - v|^2 (pytorch3d knn_points_backward, p2 side)
      g_rows[(size_t)i * 3] = c * dx; g_rows[(size_t)i * 3 + 1] = c * dy; g_rows[(size_t)i * 3 + 2] = c * dz;
    }
  }
  l = block_sum_256(l);
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, l);
}

// losses/losses.py:43-51
__global__ void __launch_bounds__(256) marker_loss_kernel(const float* __restrict__ markers,
                                                          const float* __restrict__ virt,
                                                          const float* __restrict__ mask, int n, float dist,
                                                          float scale, float* __restrict__ loss_sum,
                                                          float* __restrict__ g_virt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  float l = 0.f;
  if (i < n) {
    const float w = mask ? mask[i] : 1.f;
    const float* mp = markers + (size_t)i * 3;
    const float* vp = virt + (size_t)i * 3;
    const float dx = mp[0] - vp[0], dy = mp[1] - vp[1], dz = mp[2] - vp[2];
    const float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float e = nrm - dist;
    l = scale * w * e * e;
    if (g_virt) {
      // d/dv (|m-v| - d)^2 = 2 e * (-(m-v)/|m-v|); torch.norm backward gives 0 at |.| == 0
      const float c = nrm > 0.f ? -2.f * scale * w * e / nrm : 0.f;
      g_virt[(size_t)i * 3] = c * dx; g_virt[(size_t)i * 3 + 1] = c * dy; g_virt[(size_t)i * 3 + 2] = c * dz;
    }
  }
  l = block_sum_256(l);
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, l);
}

__global__ void __launch_bounds__(256) sqdiff_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                     size_t n, float scale, float* __restrict__ loss_sum,
                                                     float* __restrict__ g) {
  float l = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float d = x[i] - y[i];
    l += d * d;
    if (g) g[i] += 2.f * scale * d;
  }
  l = block_sum_256(l * scale);
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, l);
}

// optimization.py:368-374, sign as written: e_t = x[t+2] - 2 x[t+1] - x[t].  Rows outside [0,F) come
// from the halo rows of the neighbouring shards (prev = rows -2,-1 ; next = rows F, F+1) when given.
__device__ __forceinline__ bool trow(const float* x, const float* prev, const float* next, int F, int D, int t,
                                     int d, float& out) {
  if (t >= 0 && t < F) { out = x[(size_t)t * D + d]; return true; }
  if (t < 0 && t >= -2 && prev) { out = prev[(size_t)(t + 2) * D + d]; return true; }
  if (t >= F && t < F + 2 && next) { out = next[(size_t)(t - F) * D + d]; return true; }
  out = 0.f;
  return false;
}

__global__ void __launch_bounds__(256) temporal_kernel(const float* __restrict__ x, const float* __restrict__ prev,
                                                       const float* __restrict__ next, int F, int D, float scale,
                                                       float* __restrict__ loss_sum, float* __restrict__ g) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float l = 0.f;
  if (i < (size_t)F * D) {
    const int t = (int)(i / D), d = (int)(i % D);
    float xm2, xm1, x0, xp1, xp2;
    const bool vm2 = trow(x, prev, next, F, D, t - 2, d, xm2), vm1 = trow(x, prev, next, F, D, t - 1, d, xm1);
    trow(x, prev, next, F, D, t, d, x0);
    const bool vp1 = trow(x, prev, next, F, D, t + 1, d, xp1), vp2 = trow(x, prev, next, F, D, t + 2, d, xp2);
    float gi = 0.f;
    if (vp1 && vp2) { const float e = xp2 - 2.f * xp1 - x0; l = e * e; gi -= e; }          // e_t, own term
    if (vm1 && vp1) { const float e = xp1 - 2.f * x0 - xm1; gi -= 2.f * e; }               // e_{t-1}
    if (vm2 && vm1) { const float e = x0 - 2.f * xm1 - xm2; gi += e; }                      // e_{t-2}
    if (g) g[i] += 2.f * scale * gi;
  }
  l = block_sum_256(l * scale);
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, l);
}

// mean over ALL markers of a frame (optimization.py:248, torch.mean(markers, dim=1))
__global__ void marker_mean_kernel(const float* __restrict__ markers, int F, int M, float* __restrict__ mm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= F * 3) return;
  int f = i / 3, k = i % 3;
  float a = 0.f;
  for (int m = 0; m < M; ++m) a += markers[((size_t)f * M + m) * 3 + k];
  mm[i] = a / (float)M;
}

// optimization.py:246-250: r_t = (trans[t+1]-trans[t]) - (mm[t+1]-mm[t]); halo = [trans(3), mm(3)] of the
// neighbouring frame.  The shard owns the terms r_t with t in [0,F) whose t+1 exists.
__global__ void __launch_bounds__(256) transvel_kernel(const float* __restrict__ trans, const float* __restrict__ mm,
                                                       const float* __restrict__ prev, const float* __restrict__ next,
                                                       int F, float scale, float* __restrict__ loss_sum,
                                                       float* __restrict__ g_trans) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  float l = 0.f;
  if (i < F * 3) {
    const int t = i / 3, k = i % 3;
    const float x0 = trans[i] - mm[i];
    float gi = 0.f;
    if (t + 1 < F || next) {
      const float x1 = (t + 1 < F) ? trans[i + 3] - mm[i + 3] : next[k] - next[3 + k];
      const float r = x1 - x0;
      l = r * r;
      gi -= r;
    }
    if (t > 0 || prev) {
      const float xm = (t > 0) ? trans[i - 3] - mm[i - 3] : prev[k] - prev[3 + k];
      gi += x0 - xm;
    }
    if (g_trans) g_trans[i] += 2.f * scale * gi;
  }
  l = block_sum_256(l * scale);
  if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, l);
}

// torch.optim.Adam (single-tensor path, no amsgrad / weight decay / maximize)
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, size_t n, float beta1,
                                                   float beta2, float eps, float step_size, float bc2_sqrt) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = m[i] + (gi - m[i]) * (1.f - beta1);              // lerp_(grad, 1-beta1)
    const float vi = v[i] * beta2 + (1.f - beta2) * gi * gi;          // mul_(beta2).addcmul_(g,g,1-beta2)
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// Same update with the step number held in DEVICE memory (read here, incremented by adam_count_kernel right after), so one
// closure + Adam iteration can be captured once in a CUDA graph and replayed: nothing on the host changes per step.
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, size_t n, float lr,
                                                       float beta1, float beta2, float eps,
                                                       const int32_t* __restrict__ step_counter) {
  __shared__ float s_step_size, s_bc2_sqrt;
  if (threadIdx.x == 0) {
    const double step = (double)(*step_counter + 1);
    s_step_size = (float)((double)lr / (1.0 - pow((double)beta1, step)));
    s_bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, step));
  }
  __syncthreads();
  const float step_size = s_step_size, bc2_sqrt = s_bc2_sqrt;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = m[i] + (gi - m[i]) * (1.f - beta1);
    const float vi = v[i] * beta2 + (1.f - beta2) * gi * gi;
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}
__global__ void adam_count_kernel(int32_t* step_counter) { *step_counter += 1; }

static inline int nblocks(size_t n) {
  const size_t b = (n + 255) / 256, cap = (size_t)uuo_sm_count() * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

// rows of `vertices` that the frame's group does not keep are moved far away, so a nearest-vertex search over the
// whole array behaves like a search over the group's vertex subset (markers_utils.py:441-444, 463: vertices[:, vertex_indices])
__global__ void __launch_bounds__(256) mask_vertices_kernel(float* __restrict__ vertices, const uint8_t* __restrict__ keep,
                                                            size_t n, int V, int frames_per_group, float far_value) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t f = i / V;
  const int row = (int)(i - f * V);
  if (!keep[(f / frames_per_group) * V + row]) {
    float* p = vertices + i * 3;
    p[0] = far_value; p[1] = far_value; p[2] = far_value;
  }
}

extern "C" int uuo_mask_vertices(float* vertices, const uint8_t* keep, int F, int V, int frames_per_group, float far_value,
                                 void* stream) {
  UUO_CHECK_ARG(vertices && keep && V > 0 && frames_per_group > 0, "bad arguments");
  UUO_CHECK_ARG(F % frames_per_group == 0, "F must be a multiple of frames_per_group");
  if (F == 0) return UUO_OK;
  const size_t n = (size_t)F * V;
  ProfScope ps_("mask_vertices", (cudaStream_t)stream);
  mask_vertices_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(vertices, keep, n, V, frames_per_group,
                                                                                     far_value);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_marker_mask(const float* markers, int n, float* mask, void* stream) {
  UUO_CHECK_ARG(markers && mask && n >= 0, "bad arguments");
  if (n == 0) return UUO_OK;
  marker_mask_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(markers, n, mask);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_chamfer_rows(const float* markers, const float* vertices, const float* mask, const int32_t* idx,
                                int F, int M, int V, float scale, float* loss_sum, float* g_rows, void* stream) {
  UUO_CHECK_ARG(markers && vertices && idx, "NULL pointer");
  if (F * M == 0) return UUO_OK;
  ProfScope ps_("chamfer_rows", (cudaStream_t)stream);
  chamfer_rows_kernel<<<(F * M + 255) / 256, 256, 0, (cudaStream_t)stream>>>(markers, vertices, mask, idx, F, M, V,
                                                                             scale, loss_sum, g_rows);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_marker_loss_rows(const float* markers, const float* virt, const float* mask, int n, float dist,
                                    float scale, float* loss_sum, float* g_virt, void* stream) {
  UUO_CHECK_ARG(markers && virt, "NULL pointer");
  if (n == 0) return UUO_OK;
  ProfScope ps_("marker_loss_rows", (cudaStream_t)stream);
  marker_loss_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(markers, virt, mask, n, dist, scale, loss_sum,
                                                                        g_virt);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_sqdiff_accumulate(const float* x, const float* y, size_t n, float scale, float* loss_sum,
                                     float* g, void* stream) {
  UUO_CHECK_ARG(x && y, "NULL pointer");
  if (n == 0) return UUO_OK;
  ProfScope ps_("sqdiff", (cudaStream_t)stream);
  sqdiff_kernel<<<nblocks(n), 256, 0, (cudaStream_t)stream>>>(x, y, n, scale, loss_sum, g);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_temporal_accumulate(const float* x, const float* prev2, const float* next2, int F, int D,
                                       float scale, float* loss_sum, float* g, void* stream) {
  UUO_CHECK_ARG(x && F >= 0 && D > 0, "bad arguments");
  if (F == 0) return UUO_OK;
  size_t n = (size_t)F * D;
  ProfScope ps_("temporal", (cudaStream_t)stream);
  temporal_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, prev2, next2, F, D, scale,
                                                                                 loss_sum, g);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_marker_mean(const float* markers, int F, int M, float* mm, void* stream) {
  UUO_CHECK_ARG(markers && mm && M > 0, "bad arguments");
  if (F == 0) return UUO_OK;
  marker_mean_kernel<<<(F * 3 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(markers, F, M, mm);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_transvel_accumulate(const float* trans, const float* mm, const float* prev, const float* next,
                                       int F, float scale, float* loss_sum, float* g_trans, void* stream) {
  UUO_CHECK_ARG(trans && mm, "NULL pointer");
  if (F == 0) return UUO_OK;
  ProfScope ps_("transvel", (cudaStream_t)stream);
  transvel_kernel<<<(F * 3 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(trans, mm, prev, next, F, scale, loss_sum,
                                                                         g_trans);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                             float beta1, float beta2, float eps, int step, void* stream) {
  UUO_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && step >= 1, "bad arguments");
  if (n == 0) return UUO_OK;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  ProfScope ps_("adam_step", (cudaStream_t)stream);
  adam_kernel<<<nblocks(n), 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, beta1, beta2, eps,
                                                            (float)((double)lr / bc1), (float)sqrt(bc2));
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                                 float beta1, float beta2, float eps, int32_t* step_counter, void* stream) {
  UUO_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && step_counter, "NULL pointer");
  if (n == 0) return UUO_OK;
  ProfScope ps_("adam_step", (cudaStream_t)stream);
  adam_dev_kernel<<<nblocks(n), 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps,
                                                                step_counter);
  adam_count_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_counter);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}
