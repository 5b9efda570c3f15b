// Host-buffer entry points and the NCCL gradient all-reduce of libdrrt_b200.so.
//
// The `_host` functions are what a binding without device memory management (the
// reference's TracerS-style CPU callers, or a non-torch host program) calls: they
// own a per-thread stream and a grow-only set of device buffers, copy in, run the
// same kernels as the device entry points, copy out and synchronise.
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/drrt_b200.h"
#include "drrt_launch.h"

extern "C" void drrt_set_error_(const char *msg);

namespace {

void set_err(const std::string &m) { drrt_set_error_(m.c_str()); }

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct HostCtx {
  cudaStream_t stream = nullptr;
  DevBuf rif, ray[4], out[2], grad, stats;
  bool cache_grid = false;
  const float *cached_ptr = nullptr;
  int cw = 0, ch = 0, cd = 0;
  ~HostCtx() {}
};

thread_local HostCtx g_ctx;

int hfail(int code, const char *what, cudaError_t e) {
  set_err(std::string(what) + ": " + cudaGetErrorString(e));
  return code;
}

#define H_TRY(expr)                                          \
  do {                                                       \
    cudaError_t e_ = (expr);                                 \
    if (e_ != cudaSuccess) return hfail(DRRT_ERR_CUDA, #expr, e_); \
  } while (0)

int upload_grid(HostCtx &c, const float *rif, int W, int H, int D) {
  size_t bytes = sizeof(float) * (size_t) W * H * D;
  bool hit = c.cache_grid && c.cached_ptr == rif && c.cw == W && c.ch == H && c.cd == D && c.rif.cap >= bytes;
  if (hit) return DRRT_OK;
  H_TRY(c.rif.reserve(bytes));
  H_TRY(cudaMemcpyAsync(c.rif.p, rif, bytes, cudaMemcpyHostToDevice, c.stream));
  c.cached_ptr = rif; c.cw = W; c.ch = H; c.cd = D;
  return DRRT_OK;
}

int ensure_stream(HostCtx &c) {
  if (!c.stream) H_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  return DRRT_OK;
}

}  // namespace

extern "C" {


void drrt_host_cache_grid(int enable) { g_ctx.cache_grid = enable != 0; }
void drrt_host_invalidate_grid(void) { g_ctx.cached_ptr = nullptr; }

void drrt_host_release(void) {
  HostCtx &c = g_ctx;
  if (c.stream) cudaStreamSynchronize(c.stream);
  c.rif.release();
  for (auto &b : c.ray) b.release();
  for (auto &b : c.out) b.release();
  c.grad.release();
  c.stats.release();
  if (c.stream) cudaStreamDestroy(c.stream);
  c.stream = nullptr;
  c.cached_ptr = nullptr;
}

int drrt_trace_host(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                    int64_t N, int layout, float h, float ds, float *xt, float *vt,
                    int64_t *stats, unsigned flags) {
  HostCtx &c = g_ctx;
  if (W <= 0 || H <= 0 || D <= 0 || N < 0 || !rif) {
    set_err("drrt_trace_host: invalid argument");
    return DRRT_ERR_INVALID_ARGUMENT;
  }
  int rc = ensure_stream(c);
  if (rc) return rc;
  if ((rc = upload_grid(c, rif, W, H, D))) return rc;
  size_t rb = sizeof(float) * 3 * (size_t) N;
  if (N > 0) {
    H_TRY(c.ray[0].reserve(rb)); H_TRY(c.ray[1].reserve(rb));
    H_TRY(c.out[0].reserve(rb)); H_TRY(c.out[1].reserve(rb));
    H_TRY(cudaMemcpyAsync(c.ray[0].p, pos, rb, cudaMemcpyHostToDevice, c.stream));
    H_TRY(cudaMemcpyAsync(c.ray[1].p, vel, rb, cudaMemcpyHostToDevice, c.stream));
  }
  H_TRY(c.stats.reserve(2 * sizeof(int64_t)));
  H_TRY(cudaMemsetAsync(c.stats.p, 0, 2 * sizeof(int64_t), c.stream));
  rc = drrt_trace((const float *) c.rif.p, W, H, D, (const float *) c.ray[0].p, (const float *) c.ray[1].p,
                  N, layout, h, ds, (float *) c.out[0].p, (float *) c.out[1].p, (int64_t *) c.stats.p, flags,
                  (drrt_stream_t) c.stream);
  if (rc) return rc;
  if (N > 0) {
    H_TRY(cudaMemcpyAsync(xt, c.out[0].p, rb, cudaMemcpyDeviceToHost, c.stream));
    H_TRY(cudaMemcpyAsync(vt, c.out[1].p, rb, cudaMemcpyDeviceToHost, c.stream));
  }
  int64_t st[2] = {0, 0};
  H_TRY(cudaMemcpyAsync(st, c.stats.p, sizeof st, cudaMemcpyDeviceToHost, c.stream));
  H_TRY(cudaStreamSynchronize(c.stream));
  if (stats) { stats[0] = st[0]; stats[1] = st[1]; }
  return DRRT_OK;
}

int drrt_backtrace_host(const float *rif, int W, int H, int D, const float *xt, const float *vt,
                        const float *dx, const float *dv, int64_t N, int layout, float h,
                        float ds, float *grad, int64_t *stats, unsigned flags) {
  HostCtx &c = g_ctx;
  if (W <= 0 || H <= 0 || D <= 0 || N < 0 || !rif || !grad) {
    set_err("drrt_backtrace_host: invalid argument");
    return DRRT_ERR_INVALID_ARGUMENT;
  }
  int rc = ensure_stream(c);
  if (rc) return rc;
  if ((rc = upload_grid(c, rif, W, H, D))) return rc;
  const int64_t G = (int64_t) W * H * D;
  size_t rb = sizeof(float) * 3 * (size_t) N, gb = sizeof(float) * (size_t) G;
  const float *src[4] = {xt, vt, dx, dv};
  if (N > 0)
    for (int k = 0; k < 4; ++k) {
      H_TRY(c.ray[k].reserve(rb));
      H_TRY(cudaMemcpyAsync(c.ray[k].p, src[k], rb, cudaMemcpyHostToDevice, c.stream));
    }
  size_t sb = drrt_backtrace_scratch_bytes(G, flags);
  H_TRY(c.grad.reserve(gb + sb + 256));
  void *scratch = sb ? (void *) ((char *) c.grad.p + ((gb + 255) / 256) * 256) : nullptr;
  if (flags & DRRT_ACCUMULATE) H_TRY(cudaMemcpyAsync(c.grad.p, grad, gb, cudaMemcpyHostToDevice, c.stream));
  H_TRY(c.stats.reserve(2 * sizeof(int64_t)));
  H_TRY(cudaMemsetAsync(c.stats.p, 0, 2 * sizeof(int64_t), c.stream));
  rc = drrt_backtrace((const float *) c.rif.p, W, H, D, (const float *) c.ray[0].p, (const float *) c.ray[1].p,
                      (const float *) c.ray[2].p, (const float *) c.ray[3].p, N, layout, h, ds,
                      (float *) c.grad.p, scratch, sb, (int64_t *) c.stats.p, flags, (drrt_stream_t) c.stream);
  if (rc) return rc;
  H_TRY(cudaMemcpyAsync(grad, c.grad.p, gb, cudaMemcpyDeviceToHost, c.stream));
  int64_t st[2] = {0, 0};
  H_TRY(cudaMemcpyAsync(st, c.stats.p, sizeof st, cudaMemcpyDeviceToHost, c.stream));
  H_TRY(cudaStreamSynchronize(c.stream));
  if (stats) { stats[0] = st[0]; stats[1] = st[1]; }
  return DRRT_OK;
}

// ---------------------------------------------------------------------------
// NCCL (resolved at run time so the library loads on hosts without NCCL)
// ---------------------------------------------------------------------------
typedef struct { char internal[DRRT_UNIQUE_ID_BYTES]; } nccl_uid_t;
typedef int (*fn_uid_t)(nccl_uid_t *);
typedef int (*fn_init_t)(void **, int, nccl_uid_t, int);
typedef int (*fn_destroy_t)(void *);
typedef int (*fn_allreduce_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*fn_reducescatter_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*fn_allgather_t)(const void *, void *, size_t, int, void *, cudaStream_t);
typedef const char *(*fn_errstr_t)(int);

static struct {
  std::once_flag once;
  void *h = nullptr;
  fn_uid_t uid = nullptr;
  fn_init_t init = nullptr;
  fn_destroy_t destroy = nullptr;
  fn_allreduce_t allreduce = nullptr;
  fn_reducescatter_t reducescatter = nullptr;
  fn_allgather_t allgather = nullptr;
  fn_errstr_t errstr = nullptr;
} g_nccl;

static bool nccl_load() {
  std::call_once(g_nccl.once, [] {
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
      g_nccl.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (g_nccl.h) break;
    }
    if (!g_nccl.h) return;
    g_nccl.uid = (fn_uid_t) dlsym(g_nccl.h, "ncclGetUniqueId");
    g_nccl.init = (fn_init_t) dlsym(g_nccl.h, "ncclCommInitRank");
    g_nccl.destroy = (fn_destroy_t) dlsym(g_nccl.h, "ncclCommDestroy");
    g_nccl.allreduce = (fn_allreduce_t) dlsym(g_nccl.h, "ncclAllReduce");
    g_nccl.reducescatter = (fn_reducescatter_t) dlsym(g_nccl.h, "ncclReduceScatter");
    g_nccl.allgather = (fn_allgather_t) dlsym(g_nccl.h, "ncclAllGather");
    g_nccl.errstr = (fn_errstr_t) dlsym(g_nccl.h, "ncclGetErrorString");
  });
  return g_nccl.h && g_nccl.uid && g_nccl.init && g_nccl.destroy && g_nccl.allreduce;
}

static int nccl_fail(const char *what, int rc) {
  set_err(std::string(what) + ": " + (g_nccl.errstr ? g_nccl.errstr(rc) : "nccl error"));
  return DRRT_ERR_NCCL;
}

int drrt_comm_unique_id(char id[DRRT_UNIQUE_ID_BYTES]) {
  if (!nccl_load()) { set_err("libnccl.so.2 not found"); return DRRT_ERR_NCCL; }
  nccl_uid_t u;
  int rc = g_nccl.uid(&u);
  if (rc) return nccl_fail("ncclGetUniqueId", rc);
  memcpy(id, u.internal, DRRT_UNIQUE_ID_BYTES);
  return DRRT_OK;
}

int drrt_comm_init_rank(void **comm, int nranks, const char id[DRRT_UNIQUE_ID_BYTES], int rank) {
  if (!nccl_load()) { set_err("libnccl.so.2 not found"); return DRRT_ERR_NCCL; }
  nccl_uid_t u;
  memcpy(u.internal, id, DRRT_UNIQUE_ID_BYTES);
  int rc = g_nccl.init(comm, nranks, u, rank);
  if (rc) return nccl_fail("ncclCommInitRank", rc);
  return DRRT_OK;
}

int drrt_comm_destroy(void *comm) {
  if (!nccl_load()) { set_err("libnccl.so.2 not found"); return DRRT_ERR_NCCL; }
  int rc = g_nccl.destroy(comm);
  if (rc) return nccl_fail("ncclCommDestroy", rc);
  return DRRT_OK;
}

// One sum all-reduce of the private gradient grids:
// ncclFloat32 = 7, ncclSum = 0.
int drrt_allreduce_grad(float *grad, size_t count, void *comm, drrt_stream_t stream) {
  if (!nccl_load()) { set_err("libnccl.so.2 not found"); return DRRT_ERR_NCCL; }
  int rc = g_nccl.allreduce(grad, grad, count, 7, 0, comm, (cudaStream_t) stream);
  if (rc) return nccl_fail("ncclAllReduce", rc);
  return DRRT_OK;
}

// ---------------------------------------------------------------------------
// Peer mappings for drrt_adam_step_peers: CUDA IPC handles of the allocations that hold each rank's grid and
// gradient.  The caller's buffers usually sit inside a larger allocation (a caching allocator's segment), so the
// handle is taken for the allocation's base (cuMemGetAddressRange) and travels with the offset; an allocation is
// opened once per process however many buffers live in it.
// ---------------------------------------------------------------------------
typedef int (*fn_addr_range_t)(unsigned long long *, size_t *, unsigned long long);
static fn_addr_range_t cu_addr_range() {
  static fn_addr_range_t f = [] {
    void *h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    return h ? (fn_addr_range_t) dlsym(h, "cuMemGetAddressRange_v2") : (fn_addr_range_t) nullptr;
  }();
  return f;
}

static std::mutex g_ipc_mu;
static std::vector<std::pair<cudaIpcMemHandle_t, void *>> g_ipc_open;

int drrt_ipc_export(const void *ptr, char handle[DRRT_IPC_HANDLE_BYTES], uint64_t *offset) {
  static_assert(sizeof(cudaIpcMemHandle_t) == DRRT_IPC_HANDLE_BYTES, "IPC handle size");
  if (!ptr || !handle || !offset) { set_err("ipc_export: NULL argument"); return DRRT_ERR_INVALID_ARGUMENT; }
  fn_addr_range_t range = cu_addr_range();
  if (!range) { set_err("libcuda.so.1: cuMemGetAddressRange not found"); return DRRT_ERR_CUDA; }
  unsigned long long base = 0;
  size_t size = 0;
  if (range(&base, &size, (unsigned long long) (uintptr_t) ptr) != 0) {
    set_err("ipc_export: not a device allocation"); return DRRT_ERR_CUDA;
  }
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, (void *) (uintptr_t) base);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_err(std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e) +
            " (the buffer must come from cudaMalloc, not from a virtual-memory / async pool)");
    return DRRT_ERR_CUDA;
  }
  memcpy(handle, &h, sizeof h);
  *offset = (uint64_t) ((uintptr_t) ptr - (uintptr_t) base);
  return DRRT_OK;
}

int drrt_ipc_open(const char handle[DRRT_IPC_HANDLE_BYTES], uint64_t offset, void **ptr) {
  if (!handle || !ptr) { set_err("ipc_open: NULL argument"); return DRRT_ERR_INVALID_ARGUMENT; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  std::lock_guard<std::mutex> lk(g_ipc_mu);
  void *base = nullptr;
  for (auto &o : g_ipc_open)
    if (memcmp(&o.first, &h, sizeof h) == 0) base = o.second;
  if (!base) {
    cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      set_err(std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
      return DRRT_ERR_CUDA;
    }
    g_ipc_open.emplace_back(h, base);
  }
  *ptr = (char *) base + offset;
  return DRRT_OK;
}

int drrt_ipc_close_all(void) {
  std::lock_guard<std::mutex> lk(g_ipc_mu);
  for (auto &o : g_ipc_open) cudaIpcCloseMemHandle(o.second);
  g_ipc_open.clear();
  cudaGetLastError();
  return DRRT_OK;
}

// Sharded optimiser step: in-place reduce-scatter of the private gradient grids
// (recvbuff = sendbuff + rank * recvcount is NCCL's in-place form) and in-place all-gather of the updated slabs.
int drrt_reduce_scatter_grad(float *grad, size_t slab_count, int rank, void *comm, drrt_stream_t stream) {
  if (!nccl_load() || !g_nccl.reducescatter) { set_err("libnccl.so.2 not found"); return DRRT_ERR_NCCL; }
  if (!grad || rank < 0) { set_err("reduce_scatter: bad argument"); return DRRT_ERR_INVALID_ARGUMENT; }
  int rc = g_nccl.reducescatter(grad, grad + (size_t) rank * slab_count, slab_count, 7, 0, comm, (cudaStream_t) stream);
  if (rc) return nccl_fail("ncclReduceScatter", rc);
  return DRRT_OK;
}

int drrt_allgather_grid(float *grid, size_t slab_count, int rank, void *comm, drrt_stream_t stream) {
  if (!nccl_load() || !g_nccl.allgather) { set_err("libnccl.so.2 not found"); return DRRT_ERR_NCCL; }
  if (!grid || rank < 0) { set_err("allgather: bad argument"); return DRRT_ERR_INVALID_ARGUMENT; }
  int rc = g_nccl.allgather(grid + (size_t) rank * slab_count, grid, slab_count, 7, comm, (cudaStream_t) stream);
  if (rc) return nccl_fail("ncclAllGather", rc);
  return DRRT_OK;
}

}  // extern "C"
