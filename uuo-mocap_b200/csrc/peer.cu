// Frame-range shards talking over peer-mapped device memory (NVLink / NVSwitch P2P), so that one fitting
// iteration -- closure + Adam -- has no host-issued collective in it and can be captured in a CUDA graph on
// every rank (per iteration the shards exchange ONE small sum (d/d betas + loss,
// optimization.py:255-258 with betas shared by all frames) and, with the temporal terms on, a 1- or 2-frame
// halo with the two neighbours (optimization.py:246-250, 368-374)).
//
// Every rank owns one `Mailbox` in its own HBM; the other ranks map it (cudaIpcOpenMemHandle between
// processes, the plain pointer inside one process).  A sender WRITES its payload into the receiver's mailbox,
// fences (system scope), then writes the receiver's flag = epoch; the receiver spins on its OWN memory.
// Slots are double-buffered by epoch parity: a rank can only be one epoch ahead of a peer it exchanges with,
// because passing epoch e needs that peer's flag for e, which the peer writes after it finished reading e-1.
// The epoch counters live in device memory, so a replayed graph advances them like eager launches do.
// All ranks must issue the same sequence of comm calls.  A wait that lasts longer than the communicator's time limit
// (60 s unless uuo_comm_set_timeout says otherwise) traps instead of hanging the GPU.
#include "common.cuh"

#define UUO_COMM_MAXW 16
#define UUO_COMM_MAXRED 1024
#define UUO_COMM_MAXHALO 1024
#define UUO_COMM_TIMEOUT_NS 60000000000ull   // default; uuo_comm_set_timeout changes it per communicator

struct Mailbox {
  uint32_t ar_flag[2][UUO_COMM_MAXW];
  uint32_t halo_flag[2][2];                       // [parity][0: written by the left neighbour, 1: by the right one]
  uint32_t ar_epoch, halo_epoch, pad_[2];         // owner only
  float ar_data[2][UUO_COMM_MAXW][UUO_COMM_MAXRED];
  float halo_data[2][2][UUO_COMM_MAXHALO];
};

struct PeerTable { Mailbox* p[UUO_COMM_MAXW]; };

struct uuo_comm {
  int rank, world, device;
  Mailbox* mine;
  PeerTable peers;
  bool opened[UUO_COMM_MAXW];   // mapped through IPC (to be closed)
  bool connected;
  unsigned long long timeout_ns;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t e, unsigned long long timeout_ns) {
  const unsigned long long t0 = globaltimer_ns();
  while ((int32_t)(ld_acquire_sys(p) - e) < 0) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > timeout_ns) {
      printf("uuo_comm: a peer did not arrive at epoch %u within the time limit\n", e);
      __trap();
    }
  }
}

// x[n] <- sum over ranks of x[n], summed in rank order on every rank (identical bits everywhere)
__device__ __forceinline__ void all_reduce_body(const PeerTable& peers, Mailbox* mine, int rank, int world,
                                                float* __restrict__ x, int n, unsigned long long timeout_ns) {
  const uint32_t e = mine->ar_epoch + 1;
  const int par = e & 1;
  for (int p = 0; p < world; ++p) {
    float* dst = peers.p[p]->ar_data[par][rank];
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = x[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < world) {
    st_release_sys(&peers.p[threadIdx.x]->ar_flag[par][rank], e);
    wait_flag(&mine->ar_flag[par][threadIdx.x], e, timeout_ns);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float s = 0.f;
    for (int q = 0; q < world; ++q) s += __ldcg(&mine->ar_data[par][q][i]);
    x[i] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) mine->ar_epoch = e;
}

__global__ void __launch_bounds__(256) comm_all_reduce_kernel(PeerTable peers, Mailbox* mine, int rank, int world,
                                                              float* __restrict__ x, int n,
                                                              unsigned long long timeout_ns) {
  all_reduce_body(peers, mine, rank, world, x, n, timeout_ns);
}

// prev[r*ld + col + d] <- row (F_left - n + r) of the left neighbour, next[...] <- row r of the right neighbour
__device__ __forceinline__ void halo_body(const PeerTable& peers, Mailbox* mine, int rank, int world,
                                          const float* __restrict__ x, int F, int D, int n, float* __restrict__ prev,
                                          float* __restrict__ next, int ld, int col, unsigned long long timeout_ns) {
  const uint32_t e = mine->halo_epoch + 1;
  const int par = e & 1;
  const int cnt = n * D;
  const bool has_l = rank > 0, has_r = rank + 1 < world;
  if (has_l) {   // my first rows are the left neighbour's "next"
    float* dst = peers.p[rank - 1]->halo_data[par][1];
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) dst[i] = x[i];
  }
  if (has_r) {   // my last rows are the right neighbour's "prev"
    float* dst = peers.p[rank + 1]->halo_data[par][0];
    const float* src = x + (size_t)(F - n) * D;
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) dst[i] = src[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0 && has_l) {
    st_release_sys(&peers.p[rank - 1]->halo_flag[par][1], e);
    wait_flag(&mine->halo_flag[par][0], e, timeout_ns);
  }
  if (threadIdx.x == 32 && has_r) {
    st_release_sys(&peers.p[rank + 1]->halo_flag[par][0], e);
    wait_flag(&mine->halo_flag[par][1], e, timeout_ns);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int r = i / D, d = i % D;
    if (has_l && prev) prev[(size_t)r * ld + col + d] = __ldcg(&mine->halo_data[par][0][i]);
    if (has_r && next) next[(size_t)r * ld + col + d] = __ldcg(&mine->halo_data[par][1][i]);
  }
  __syncthreads();
  if (threadIdx.x == 0) mine->halo_epoch = e;
}

__global__ void __launch_bounds__(256) comm_halo_kernel(PeerTable peers, Mailbox* mine, int rank, int world,
                                                        const float* __restrict__ x, int F, int D, int n,
                                                        float* __restrict__ prev, float* __restrict__ next, int ld,
                                                        int col, unsigned long long timeout_ns) {
  halo_body(peers, mine, rank, world, x, F, D, n, prev, next, ld, col, timeout_ns);
}

// ---- all ranks of a world that lives on ONE device, as ONE cooperative launch: block r plays rank r -------------
// Separate launches that wait on one another are not guaranteed to run at the same time on one GPU, so a process that
// drives several shards of one device (the single-GPU parity tests) runs the exchange of all its ranks in one grid
// whose blocks are co-resident by construction.  Same mailboxes, epochs, flags and code as the per-rank kernels.
struct GroupArgs {
  PeerTable peers;                       // the same table on every rank (all mailboxes are local)
  float* x[UUO_COMM_MAXW];
  const float* hx[UUO_COMM_MAXW];
  float* prev[UUO_COMM_MAXW];
  float* next[UUO_COMM_MAXW];
  int F[UUO_COMM_MAXW];
};

__global__ void __launch_bounds__(256) comm_group_all_reduce_kernel(GroupArgs a, int world, int n,
                                                                    unsigned long long timeout_ns) {
  const int r = blockIdx.x;
  all_reduce_body(a.peers, a.peers.p[r], r, world, a.x[r], n, timeout_ns);
}

__global__ void __launch_bounds__(256) comm_group_halo_kernel(GroupArgs a, int world, int D, int n, int ld, int col,
                                                              unsigned long long timeout_ns) {
  const int r = blockIdx.x;
  halo_body(a.peers, a.peers.p[r], r, world, a.hx[r], a.F[r], D, n, a.prev[r], a.next[r], ld, col, timeout_ns);
}

extern "C" int uuo_comm_create(int rank, int world, int device, uuo_comm** out) {
  UUO_CHECK_ARG(out != nullptr, "out is NULL");
  UUO_CHECK_ARG(world >= 1 && world <= UUO_COMM_MAXW && rank >= 0 && rank < world, "bad rank / world (<= 16 ranks)");
  int prev_dev = 0;
  UUO_CUDA(cudaGetDevice(&prev_dev));
  UUO_CUDA(cudaSetDevice(device));
  uuo_comm* c = new uuo_comm();
  c->rank = rank; c->world = world; c->device = device; c->connected = false;
  c->timeout_ns = UUO_COMM_TIMEOUT_NS;
  // load both kernels NOW: with lazy module loading the first launch of a kernel may have to wait for running kernels
  // to finish, and a running exchange kernel waits for its peers
  cudaFuncAttributes fa;
  UUO_CUDA(cudaFuncGetAttributes(&fa, comm_all_reduce_kernel));
  UUO_CUDA(cudaFuncGetAttributes(&fa, comm_halo_kernel));
  for (int i = 0; i < UUO_COMM_MAXW; ++i) { c->peers.p[i] = nullptr; c->opened[i] = false; }
  if (cudaMalloc(&c->mine, sizeof(Mailbox)) != cudaSuccess) {
    delete c;
    cudaSetDevice(prev_dev);
    uuo_set_error("uuo_comm_create: cudaMalloc of the mailbox failed");
    return UUO_ERR_ALLOC;
  }
  UUO_CUDA(cudaMemset(c->mine, 0, sizeof(Mailbox)));
  UUO_CUDA(cudaDeviceSynchronize());
  c->peers.p[rank] = c->mine;
  if (world == 1) c->connected = true;
  UUO_CUDA(cudaSetDevice(prev_dev));
  *out = c;
  return UUO_OK;
}

extern "C" int uuo_comm_set_timeout(uuo_comm* c, double seconds) {
  UUO_CHECK_ARG(c && seconds > 0.0 && seconds < 86400.0, "bad arguments");
  c->timeout_ns = (unsigned long long)(seconds * 1e9);
  return UUO_OK;
}

extern "C" int uuo_comm_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int uuo_comm_export(uuo_comm* c, void* handle_host) {
  UUO_CHECK_ARG(c && handle_host, "NULL pointer");
  cudaIpcMemHandle_t h;
  UUO_CUDA(cudaIpcGetMemHandle(&h, c->mine));
  memcpy(handle_host, &h, sizeof(h));
  return UUO_OK;
}

extern "C" int uuo_comm_connect(uuo_comm* c, const void* handles_host) {
  UUO_CHECK_ARG(c && handles_host, "NULL pointer");
  int prev_dev = 0;
  UUO_CUDA(cudaGetDevice(&prev_dev));
  UUO_CUDA(cudaSetDevice(c->device));
  const cudaIpcMemHandle_t* hs = (const cudaIpcMemHandle_t*)handles_host;
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank || c->peers.p[p]) continue;
    void* ptr = nullptr;
    UUO_CUDA(cudaIpcOpenMemHandle(&ptr, hs[p], cudaIpcMemLazyEnablePeerAccess));
    c->peers.p[p] = (Mailbox*)ptr;
    c->opened[p] = true;
  }
  c->connected = true;
  UUO_CUDA(cudaSetDevice(prev_dev));
  return UUO_OK;
}

extern "C" int uuo_comm_connect_local(uuo_comm* c, uuo_comm* const* peers) {
  UUO_CHECK_ARG(c && peers, "NULL pointer");
  for (int p = 0; p < c->world; ++p) {
    UUO_CHECK_ARG(peers[p] && peers[p]->rank == p && peers[p]->world == c->world, "peers[p] must be rank p of this world");
    if (peers[p]->device != c->device) {
      int can = 0;
      UUO_CUDA(cudaDeviceCanAccessPeer(&can, c->device, peers[p]->device));
      UUO_CHECK_ARG(can, "the peer's device is not P2P-accessible");
      int prev_dev = 0;
      UUO_CUDA(cudaGetDevice(&prev_dev));
      UUO_CUDA(cudaSetDevice(c->device));
      cudaError_t e = cudaDeviceEnablePeerAccess(peers[p]->device, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
      cudaSetDevice(prev_dev);
      UUO_CUDA(e);
    }
    c->peers.p[p] = peers[p]->mine;
  }
  c->connected = true;
  return UUO_OK;
}

extern "C" void uuo_comm_destroy(uuo_comm* c) {
  if (!c) return;
  int prev_dev = 0;
  cudaGetDevice(&prev_dev);
  cudaSetDevice(c->device);
  for (int p = 0; p < c->world; ++p)
    if (c->opened[p] && c->peers.p[p]) cudaIpcCloseMemHandle(c->peers.p[p]);
  cudaFree(c->mine);
  cudaSetDevice(prev_dev);
  delete c;
}

extern "C" int uuo_comm_all_reduce_sum(uuo_comm* c, float* x, int n, void* stream) {
  UUO_CHECK_ARG(c && x, "NULL pointer");
  UUO_CHECK_ARG(c->connected, "uuo_comm_connect[_local] has not run");
  UUO_CHECK_ARG(n >= 0 && n <= UUO_COMM_MAXRED, "n must be <= 1024 floats");
  if (c->world == 1 || n == 0) return UUO_OK;
  ProfScope ps_("comm_all_reduce", (cudaStream_t)stream);
  comm_all_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(c->peers, c->mine, c->rank, c->world, x, n,
                                                                      c->timeout_ns);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

extern "C" int uuo_comm_halo_exchange(uuo_comm* c, const float* x, int F, int D, int n, float* prev, float* next,
                                      int ld, int col, void* stream) {
  UUO_CHECK_ARG(c && x, "NULL pointer");
  UUO_CHECK_ARG(c->connected, "uuo_comm_connect[_local] has not run");
  UUO_CHECK_ARG(n >= 1 && D >= 1 && (size_t)n * D <= UUO_COMM_MAXHALO, "halo must be <= 1024 floats");
  UUO_CHECK_ARG(F >= n, "a shard must hold at least n frames");
  UUO_CHECK_ARG(ld >= col + D && col >= 0, "bad ld / col");
  UUO_CHECK_ARG((c->rank == 0 || prev) && (c->rank + 1 == c->world || next), "prev / next buffer is NULL");
  if (c->world == 1) return UUO_OK;
  ProfScope ps_("comm_halo", (cudaStream_t)stream);
  comm_halo_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(c->peers, c->mine, c->rank, c->world, x, F, D, n, prev, next,
                                                        ld, col, c->timeout_ns);
  UUO_LAUNCH_CHECK();
  return UUO_OK;
}

static int group_check(uuo_comm* const* comms, int world) {
  UUO_CHECK_ARG(comms && world >= 1 && world <= UUO_COMM_MAXW, "bad arguments");
  for (int r = 0; r < world; ++r) {
    UUO_CHECK_ARG(comms[r] && comms[r]->rank == r && comms[r]->world == world && comms[r]->connected,
                  "comms[r] must be the connected rank r of this world");
    UUO_CHECK_ARG(comms[r]->device == comms[0]->device, "a group launch needs every rank on one device");
  }
  return UUO_OK;
}

extern "C" int uuo_comm_group_all_reduce_sum(uuo_comm* const* comms, float* const* xs, int n, int world, void* stream) {
  int rc = group_check(comms, world);
  if (rc) return rc;
  UUO_CHECK_ARG(xs && n >= 0 && n <= UUO_COMM_MAXRED, "n must be <= 1024 floats");
  if (world == 1 || n == 0) return UUO_OK;
  GroupArgs a;
  memset(&a, 0, sizeof(a));
  a.peers = comms[0]->peers;
  for (int r = 0; r < world; ++r) { UUO_CHECK_ARG(xs[r], "xs[r] is NULL"); a.x[r] = xs[r]; }
  unsigned long long to = comms[0]->timeout_ns;
  void* args[] = {&a, &world, &n, &to};
  UUO_CUDA(cudaLaunchCooperativeKernel((const void*)comm_group_all_reduce_kernel, dim3(world), dim3(256), args, 0,
                                       (cudaStream_t)stream));
  return UUO_OK;
}

extern "C" int uuo_comm_group_halo_exchange(uuo_comm* const* comms, const float* const* xs, const int* Fs, int D, int n,
                                            float* const* prevs, float* const* nexts, int ld, int col, int world,
                                            void* stream) {
  int rc = group_check(comms, world);
  if (rc) return rc;
  UUO_CHECK_ARG(xs && Fs && prevs && nexts, "NULL pointer");
  UUO_CHECK_ARG(n >= 1 && D >= 1 && (size_t)n * D <= UUO_COMM_MAXHALO, "halo must be <= 1024 floats");
  UUO_CHECK_ARG(ld >= col + D && col >= 0, "bad ld / col");
  if (world == 1) return UUO_OK;
  GroupArgs a;
  memset(&a, 0, sizeof(a));
  a.peers = comms[0]->peers;
  for (int r = 0; r < world; ++r) {
    UUO_CHECK_ARG(xs[r] && Fs[r] >= n, "every shard must hold at least n frames");
    UUO_CHECK_ARG((r == 0 || prevs[r]) && (r + 1 == world || nexts[r]), "prev / next buffer is NULL");
    a.hx[r] = xs[r]; a.F[r] = Fs[r]; a.prev[r] = prevs[r]; a.next[r] = nexts[r];
  }
  unsigned long long to = comms[0]->timeout_ns;
  void* args[] = {&a, &world, &D, &n, &ld, &col, &to};
  UUO_CUDA(cudaLaunchCooperativeKernel((const void*)comm_group_halo_kernel, dim3(world), dim3(256), args, 0,
                                       (cudaStream_t)stream));
  return UUO_OK;
}
