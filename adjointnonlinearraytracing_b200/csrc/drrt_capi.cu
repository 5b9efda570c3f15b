// C ABI of libdrrt_b200.so (see include/drrt_b200.h for the contract and the
// reference interface each entry point replaces).
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <string>

#include "../../include/drrt_b200.h"
#include <math.h>

#include "drrt_launch.h"

using namespace drrt;

namespace {

thread_local std::string g_err;
std::atomic<int64_t> g_launches{0};

int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CU_TRY(expr)                                                              \
  do {                                                                            \
    cudaError_t e_ = (expr);                                                      \
    if (e_ != cudaSuccess) return fail(DRRT_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

int check_launch(const char *what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(DRRT_ERR_CUDA, "%s launch failed: %s", what, cudaGetErrorString(e));
  return DRRT_OK;
}

// Loop caps are evaluated in float and truncated, reference src/tracer.cpp:51,
// :262, :417 (`int max_steps = 4 * h * hmax(res) / delta_s`).
int cap_grid(float mult, float h, int W, int H, int D, float ds) {
  int m = W > H ? W : H;
  if (D > m) m = D;
  float f = mult * h * (float) m / ds;
  if (!(f > 0.f)) return 0;
  if (f > 2.0e9f) return 2000000000;
  return (int) f;
}

int cap_cable(float length, float ds) {  // src/tracer.cpp:332, :544
  float f = 4 * length / ds;
  if (!(f > 0.f)) return 0;
  if (f > 2.0e9f) return 2000000000;
  return (int) f;
}

int make_grid(GridP &g, const float *rif, int W, int H, int D, float h) {
  if (W <= 0 || H <= 0 || D <= 0 || (int64_t) W * H * D > (int64_t) INT_MAX)
    return fail(DRRT_ERR_RESOLUTION, "Resolution doesn't match data");
  // volume::eval_grad, reference src/volume.cpp:117-124
  if (!(W == 1 && H == 1 && D == 1) && (W < 2 || H < 2))
    return fail(DRRT_ERR_RESOLUTION, "volume: invalid resolution!");
  if (!rif) return fail(DRRT_ERR_INVALID_ARGUMENT, "rif is NULL");
  if (!(h > 0.f)) return fail(DRRT_ERR_INVALID_ARGUMENT, "h must be positive");
  g.data = rif;
  g.W = W; g.H = H; g.D = D;
  g.h = h;
  g.rcp_h = 1.0f / h;
  g.hix = (float) (W - 1) * h;
  g.hiy = (float) (H - 1) * h;
  g.hiz = (float) (D - 1) * h;
  return DRRT_OK;
}

int make_cable(CableP &c, const float *rif, int res, float radius, float length) {
  if (res < 2) return fail(DRRT_ERR_RESOLUTION, "cable profile needs at least 2 samples");
  if (!rif) return fail(DRRT_ERR_INVALID_ARGUMENT, "rif is NULL");
  c.data = rif;
  c.res = res;
  c.radius = radius;
  c.length = length;
  c.h = radius / (float) (res - 1);
  c.r2 = radius * radius;
  return DRRT_OK;
}

int check_rays(int64_t N, int layout, float ds) {
  if (N < 0) return fail(DRRT_ERR_INVALID_ARGUMENT, "negative ray count");
  if (layout != DRRT_SOA && layout != DRRT_AOS)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "layout must be DRRT_SOA (1) or DRRT_AOS (3)");
  if (!(ds > 0.f)) return fail(DRRT_ERR_INVALID_ARGUMENT, "ds must be positive");
  return DRRT_OK;
}

constexpr size_t kFixedTail = 16;

// 0 = off, -1 = measured on the device, 2^k = forced (DRRT_INTERLEAVE(n)).
// Default: off with the face-carry flush (measured: profiles/r05_facecarry.log -- once the
// shared face stays in registers the strided ray assignment no longer pays), measured on
// the device without it (DRRT_NO_FACE_CARRY = the r03/r04 configuration).
int interleave_of(unsigned flags) {
  if (flags & DRRT_NO_INTERLEAVE) return 0;
  unsigned n = (flags >> 12) & 0x3fu;
  if (n) return (int) n;
  if (flags & DRRT_AUTO_INTERLEAVE) return -1;
  return (flags & DRRT_NO_FACE_CARRY) ? -1 : 0;
}

// 0 fast polynomial, 1 reference-literal, 2 exact-trajectory (default)
int arith_of(unsigned flags) {
  if (flags & DRRT_REFERENCE_ARITH) return 1;
  if (flags & DRRT_FAST_ARITH) return 0;
  return 2;
}  // max_bits word, padded

// Two rays per thread in drrt_trace (trace_pair_kernel): forced by DRRT_PAIR_RAYS / DRRT_NO_PAIR_RAYS.
int pair_of(unsigned flags, int64_t N) {
  if (flags & DRRT_NO_PAIR_RAYS) return 0;
  if (flags & DRRT_PAIR_RAYS) return 1;
  (void) N;
  return 0;
}

// Gather from a shared-memory copy of the grid (trace_kernel / backtrace_kernel <..., SM>): grids of up to 50 000 voxels
// (200 KB; 36^3) and batches of at most 512 rays per SM.  OPT-IN (DRRT_SMEM_GRID): measured slower than the default on
// the 32^3 / 256^2 configuration (adjoint 0.167 vs 0.153 ms on the lens field, 0.126 vs 0.110 ms on the bump field,
// forward 0.084 vs 0.073 ms; profiles/r12_smem_grid_rejected.log) -- 148 CTAs x 128 KB of grid copy and one 448-thread CTA
// per SM at 90 registers cost more than the L2 round trips they remove.
int smem_grid_of(unsigned flags, int64_t G, int64_t N) {
  if (!(flags & DRRT_SMEM_GRID)) return 0;
  return G * 4 <= 200 * 1024 && N <= 148 * 512 ? 1 : 0;
}

int trace_common(int kind, const float *rif, const float *sdf, int W, int H, int D,
                 const float *pos, const float *vel, const float *aux0, const float *aux1,
                 int64_t N, int layout, float h, float ds, float *xt, float *vt,
                 uint8_t *not_escaped, float *dist2, int64_t *stats, unsigned flags,
                 cudaStream_t s) {
  TraceArgs a;
  memset(&a, 0, sizeof a);
  int rc = make_grid(a.grid, rif, W, H, D, h);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, ds))) return rc;
  if (N == 0) return DRRT_OK;
  if (!pos || !vel || !xt || !vt) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL ray buffer");
  a.sdf = sdf; a.pos = pos; a.vel = vel; a.aux0 = aux0; a.aux1 = aux1;
  a.N = N; a.layout = layout; a.ds = ds;
  a.max_steps = cap_grid(kind == 3 ? 2.0f : 4.0f, h, W, H, D, ds);
  a.xt = xt; a.vt = vt; a.not_escaped = not_escaped; a.dist2 = dist2; a.stats = stats;
  void *tail = nullptr;
  if (kind >= 1) {
    // SoA scratch for the batch-coupled tails (trace_plane, trace_target, trace_sdf): x, v (3N floats each), iters (N), max (1)
    size_t bytes = sizeof(float) * 6 * (size_t) N + sizeof(int) * ((size_t) N + 4);
    CU_TRY(cudaMallocAsync(&tail, bytes, s));
    a.tail_x = (float *) tail;
    a.tail_v = a.tail_x + 3 * N;
    a.tail_iters = (int *) (a.tail_v + 3 * N);
    a.tail_max_iters = a.tail_iters + N;
    CU_TRY(cudaMemsetAsync(a.tail_max_iters, 0, sizeof(int), s));
  }
  a.pair = pair_of(flags, N);
  a.smem_grid = smem_grid_of(flags, (int64_t) W * H * D, N);
  a.one = 1.0f;
  g_launches += launch_trace(kind, arith_of(flags), a, s);
  rc = check_launch("trace");
  if (tail) cudaFreeAsync(tail, s);
  return rc;
}

// Private gradient copies.  Every float add to ONE address is serialised chip-wide at ~0.7 G adds/s
// however the adds are packed into requests (profiles/r09_l2_probe_hot.log), so a batch that
// focuses (a lens) or crosses (tomography views) spends its time queueing on a few voxels.  CTA b
// therefore adds into copy b % R and one pass sums the copies: R x the throughput on the hot voxels at
// no cost inside the ray loop.  R is bounded by the copies' footprint (2^23 floats = 32 MiB, L2-
// resident next to the grid) and by the batch (no more than ~8 rays' worth of voxels to sum).
// Measured (profiles/r09_replicas.log): Luneburg 32^3 adjoint 332 -> 152 us, 64^3 lens 1234 -> 646 us,
// fuel-injection views 4.82 -> 4.18 ms, non-focusing 64^3 491 -> 397 us; nothing at 129^3 / 256^3
// (those are issue-bound, profiles/r09_replicas_big.log), which the footprint bound leaves alone.
constexpr int kMaxReplicas = 16;
constexpr int64_t kReplicaFloats = (int64_t) 1 << 23;

int replicas_for_grid(int64_t G, unsigned flags) {
  if ((flags & (DRRT_DETERMINISTIC | DRRT_NO_REPLICAS)) || G < 4096) return 1;
  const int forced = (int) ((flags >> 20) & 0xfu);  // DRRT_REPLICAS(n): the caller's choice, any grid size
  if (forced > 1) return forced;
  const int64_t r = kReplicaFloats / G;
  return (int) (r < 1 ? 1 : r > kMaxReplicas ? kMaxReplicas : r);
}

int backtrace_common(bool sdf_kind, const float *rif, const float *sdf, int W, int H, int D,
                     const float *xt, const float *vt, const float *dx, const float *dv, int64_t N,
                     int layout, float h, float ds, float *grad, void *scratch, size_t scratch_bytes, int64_t *stats,
                     unsigned flags, cudaStream_t s, const float *pos0 = nullptr, float *dpos = nullptr,
                     float *dvel = nullptr) {
  AdjArgs a;
  memset(&a, 0, sizeof a);
  a.pos0 = pos0; a.dpos = dpos; a.dvel = dvel;
  int rc = make_grid(a.grid, rif, W, H, D, h);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, ds))) return rc;
  if (!grad) return fail(DRRT_ERR_INVALID_ARGUMENT, "grad is NULL");
  const int64_t G = (int64_t) W * H * D;
  const bool det = (flags & DRRT_DETERMINISTIC) != 0;
  const bool accumulate = (flags & DRRT_ACCUMULATE) != 0;
  if (det && !scratch) return fail(DRRT_ERR_INVALID_ARGUMENT, "DRRT_DETERMINISTIC needs a scratch buffer");
  // The scratch buffer is sized by the caller with drrt_backtrace_scratch_bytes(G, flags) for THIS flag word; a
  // buffer sized under other flags (e.g. for the int64 grid, then reused for private float copies) is caught here.
  if (det && scratch_bytes < sizeof(unsigned long long) * (size_t) G + kFixedTail)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "scratch buffer smaller than drrt_backtrace_scratch_bytes(grid_elems, flags)");
  if (!accumulate) CU_TRY(cudaMemsetAsync(grad, 0, sizeof(float) * (size_t) G, s));  // src/tracer.cpp:401
  if (N == 0) return DRRT_OK;
  if (!xt || !vt || !dx || !dv) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL ray buffer");
  a.sdf = sdf; a.xt = xt; a.vt = vt; a.dx = dx; a.dv = dv;
  a.N = N; a.layout = layout; a.ds = ds; a.stats = stats;
  a.max_steps = cap_grid(2.0f, h, W, H, D, ds);
  a.agg_policy = (int) ((flags >> 8) & 0xfu);  // experimental: DRRT_AGG_RUN(n), 0 = adaptive
  if (a.agg_policy == 15) a.agg_policy = 32;   // DRRT_AGG_RUN_WARP
  a.agg_t8 = 6; a.agg_twarp = 12;  // adaptive policy thresholds (profiles/r09_agg_policies.log)
  // The face carry halves the atomics of a cell change and pays while cell changes are frequent; from
  // ~6 steps per cell on its branchy flush costs more than the atomics it saves (Luneburg yaml shape:
  // ds = h/4 7.2 vs 8.9 ms with / without, h/8 12.6 vs 11.1, h/32 39.3 vs 37.4;
  // profiles/r09_facecarry_vs_step.log), so it is chosen from h/ds unless a flag decides.
  a.smem_grid = smem_grid_of(flags, G, N);
  a.face_carry = (flags & DRRT_NO_FACE_CARRY) ? 0 : (flags & DRRT_FACE_CARRY) ? 1 : (h < 6.0f * ds ? 1 : 0);
  unsigned long long *fixed = nullptr;
  unsigned *max_bits = nullptr;
  if (det) {
    fixed = (unsigned long long *) scratch;
    max_bits = (unsigned *) ((char *) scratch + sizeof(unsigned long long) * (size_t) G);
    CU_TRY(cudaMemsetAsync(scratch, 0, sizeof(unsigned long long) * (size_t) G + kFixedTail, s));
    g_launches += launch_fixed_max(dx, dv, 3 * N, max_bits, s);
  }
  int R = scratch ? replicas_for_grid(G, flags) : 1;
  if (R > 1 && scratch_bytes < sizeof(float) * (size_t) G * (size_t) (R - 1)) {
    if ((flags >> 20) & 0xfu)  // DRRT_REPLICAS(n) asked for by the caller: an undersized buffer is an error
      return fail(DRRT_ERR_INVALID_ARGUMENT, "scratch buffer smaller than drrt_backtrace_scratch_bytes(grid_elems, flags)");
    R = 1 + (int) (scratch_bytes / (sizeof(float) * (size_t) G));  // the library's own choice: as many copies as fit
  }
  if (R > 1 && !((flags >> 20) & 0xfu)) {
    const int64_t by_rays = 8 * N / G;
    if (by_rays < R) R = by_rays < 1 ? 1 : (int) by_rays;
  }
  if (R > 1) {
    a.replicas = R; a.replica_base = (float *) scratch; a.replica_stride = G;
    CU_TRY(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t) G * (size_t) (R - 1), s));
  }
  g_launches += launch_backtrace(sdf_kind, arith_of(flags), (flags & DRRT_WARP_AGGREGATION) != 0, interleave_of(flags), a, grad, fixed, max_bits, s);
  if (R > 1) g_launches += launch_reduce_replicas(grad, (const float *) scratch, G, R - 1, s);
  if (det) g_launches += launch_fixed_to_float(fixed, max_bits, grad, G, accumulate, s);
  return check_launch("backtrace");
}

}  // namespace

extern "C" {

int drrt_abi_version(void) { return DRRT_ABI_VERSION; }
const char *drrt_last_error(void) { return g_err.c_str(); }
// internal: lets drrt_host.cu report through the same per-thread slot
void drrt_set_error_(const char *msg) { g_err = msg ? msg : ""; }
int64_t drrt_launch_count(void) { return g_launches.load(); }

int drrt_trace(const float *rif, int W, int H, int D, const float *pos, const float *vel,
               int64_t N, int layout, float h, float ds, float *xt, float *vt, int64_t *stats,
               unsigned flags, drrt_stream_t stream) {
  return trace_common(0, rif, nullptr, W, H, D, pos, vel, nullptr, nullptr, N, layout, h, ds, xt, vt,
                      nullptr, nullptr, stats, flags, (cudaStream_t) stream);
}

int drrt_trace_plane(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                     const float *pln_o, const float *pln_d, int64_t N, int layout, float h,
                     float ds, float *xt, float *vt, uint8_t *not_escaped, int64_t *stats,
                     unsigned flags, drrt_stream_t stream) {
  if (N > 0 && (!pln_o || !pln_d || !not_escaped)) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL plane buffer");
  return trace_common(1, rif, nullptr, W, H, D, pos, vel, pln_o, pln_d, N, layout, h, ds, xt, vt,
                      not_escaped, nullptr, stats, flags, (cudaStream_t) stream);
}

int drrt_trace_target(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                      const float *target, int64_t N, int layout, float h, float ds, float *xt,
                      float *vt, float *dist2, int64_t *stats, unsigned flags,
                      drrt_stream_t stream) {
  if (N > 0 && (!target || !dist2)) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL target buffer");
  return trace_common(2, rif, nullptr, W, H, D, pos, vel, target, nullptr, N, layout, h, ds, xt, vt,
                      nullptr, dist2, stats, flags, (cudaStream_t) stream);
}

int drrt_trace_sdf(const float *rif, const float *sdf, int W, int H, int D, const float *pos,
                   const float *vel, int64_t N, int layout, float h, float ds, float *xt,
                   float *vt, int64_t *stats, unsigned flags, drrt_stream_t stream) {
  if (!sdf) return fail(DRRT_ERR_INVALID_ARGUMENT, "sdf is NULL");
  return trace_common(3, rif, sdf, W, H, D, pos, vel, nullptr, nullptr, N, layout, h, ds, xt, vt,
                      nullptr, nullptr, stats, flags, (cudaStream_t) stream);
}

int drrt_trace_cable(const float *rif, int res, float radius, float length, const float *pos,
                     const float *vel, const float *target, int64_t N, int layout, float ds,
                     float *xt, float *vt, float *dist2, int64_t *stats, unsigned flags,
                     drrt_stream_t stream) {
  CableArgs a;
  memset(&a, 0, sizeof a);
  int rc = make_cable(a.cable, rif, res, radius, length);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, ds))) return rc;
  if (N == 0) return DRRT_OK;
  if (!pos || !vel || !target || !xt || !vt || !dist2) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL ray buffer");
  a.pos = pos; a.vel = vel; a.target = target; a.N = N; a.layout = layout; a.ds = ds;
  a.max_steps = cap_cable(length, ds);
  a.xt = xt; a.vt = vt; a.dist2 = dist2; a.stats = stats;
  g_launches += launch_trace_cable((flags & DRRT_REFERENCE_ARITH) != 0, a, (cudaStream_t) stream);
  return check_launch("trace_cable");
}

size_t drrt_backtrace_scratch_bytes(int64_t grid_elems, unsigned flags) {
  if (grid_elems <= 0) return 0;
  if (flags & DRRT_DETERMINISTIC) return sizeof(unsigned long long) * (size_t) grid_elems + kFixedTail;
  const int R = replicas_for_grid(grid_elems, flags);
  return R > 1 ? sizeof(float) * (size_t) grid_elems * (size_t) (R - 1) : 0;
}

int drrt_backtrace(const float *rif, int W, int H, int D, const float *xt, const float *vt,
                   const float *dx, const float *dv, int64_t N, int layout, float h, float ds,
                   float *grad, void *scratch, size_t scratch_bytes, int64_t *stats, unsigned flags,
                   drrt_stream_t stream) {
  return backtrace_common(false, rif, nullptr, W, H, D, xt, vt, dx, dv, N, layout, h, ds, grad,
                          scratch, scratch_bytes, stats, flags, (cudaStream_t) stream);
}

int drrt_backtrace_sdf(const float *rif, const float *sdf, int W, int H, int D, const float *xt,
                       const float *vt, const float *dx, const float *dv, int64_t N, int layout,
                       float h, float ds, float *grad, void *scratch, size_t scratch_bytes, int64_t *stats,
                       unsigned flags, drrt_stream_t stream) {
  if (!sdf) return fail(DRRT_ERR_INVALID_ARGUMENT, "sdf is NULL");
  return backtrace_common(true, rif, sdf, W, H, D, xt, vt, dx, dv, N, layout, h, ds, grad, scratch,
                          scratch_bytes, stats, flags, (cudaStream_t) stream);
}

int drrt_backtrace_rays(const float *rif, const float *sdf, int W, int H, int D, const float *xt,
                        const float *vt, const float *dx, const float *dv, const float *pos0, int64_t N,
                        int layout, float h, float ds, float *grad, float *dpos, float *dvel,
                        void *scratch, size_t scratch_bytes, int64_t *stats, unsigned flags, drrt_stream_t stream) {
  if (N > 0 && (!dpos || !dvel)) return fail(DRRT_ERR_INVALID_ARGUMENT, "dpos / dvel is NULL");
  return backtrace_common(sdf != nullptr, rif, sdf, W, H, D, xt, vt, dx, dv, N, layout, h, ds, grad,
                          scratch, scratch_bytes, stats, flags, (cudaStream_t) stream, pos0, dpos, dvel);
}

int drrt_backtrace_cable(const float *rif, int res, float radius, float length, const float *xt,
                         const float *vt, const float *dx, const float *dv, int64_t N, int layout,
                         float ds, float *grad, void *scratch, size_t scratch_bytes, int64_t *stats, unsigned flags,
                         drrt_stream_t stream) {
  cudaStream_t s = (cudaStream_t) stream;
  CableArgs a;
  memset(&a, 0, sizeof a);
  int rc = make_cable(a.cable, rif, res, radius, length);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, ds))) return rc;
  if (!grad) return fail(DRRT_ERR_INVALID_ARGUMENT, "grad is NULL");
  const bool det = (flags & DRRT_DETERMINISTIC) != 0;
  const bool accumulate = (flags & DRRT_ACCUMULATE) != 0;
  if (det && !scratch) return fail(DRRT_ERR_INVALID_ARGUMENT, "DRRT_DETERMINISTIC needs a scratch buffer");
  if (det && scratch_bytes < sizeof(unsigned long long) * (size_t) res + kFixedTail)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "scratch buffer smaller than drrt_backtrace_scratch_bytes(res, flags)");
  if (!accumulate) CU_TRY(cudaMemsetAsync(grad, 0, sizeof(float) * (size_t) res, s));
  if (N == 0) return DRRT_OK;
  if (!xt || !vt || !dx || !dv) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL ray buffer");
  a.pos = xt; a.vel = vt; a.dx = dx; a.dv = dv; a.N = N; a.layout = layout; a.ds = ds;
  a.max_steps = cap_cable(length, ds);
  a.stats = stats;
  unsigned long long *fixed = nullptr;
  unsigned *max_bits = nullptr;
  if (det) {
    fixed = (unsigned long long *) scratch;
    max_bits = (unsigned *) ((char *) scratch + sizeof(unsigned long long) * (size_t) res);
    CU_TRY(cudaMemsetAsync(scratch, 0, sizeof(unsigned long long) * (size_t) res + kFixedTail, s));
    g_launches += launch_fixed_max(dx, dv, 3 * N, max_bits, s);
  }
  g_launches += launch_backtrace_cable((flags & DRRT_REFERENCE_ARITH) != 0, a, grad, fixed, max_bits, s);
  if (det) g_launches += launch_fixed_to_float(fixed, max_bits, grad, res, accumulate, s);
  return check_launch("backtrace_cable");
}

namespace {
int sensor_common(bool backward, const float *x, const float *v, const float *e, float e0, int64_t N, int layout,
                  const float *plane_p, const float *plane_n, const float *t1, const float *t2, int res,
                  float span, unsigned flags, float *image, const float *dimage, float *dx, float *dv,
                  float *de, cudaStream_t s) {
  const bool far = (flags & DRRT_SENSOR_FAR) != 0;
  if (N < 0) return fail(DRRT_ERR_INVALID_ARGUMENT, "negative ray count");
  if (layout != DRRT_SOA && layout != DRRT_AOS) return fail(DRRT_ERR_INVALID_ARGUMENT, "layout must be DRRT_SOA or DRRT_AOS");
  if (res <= 0 || !(span > 0.f)) return fail(DRRT_ERR_INVALID_ARGUMENT, "sensor: res and span must be positive");
  if (!plane_n || !t1 || !t2 || (!far && !plane_p)) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL plane vector");
  SensorArgs a;
  memset(&a, 0, sizeof a);
  for (int k = 0; k < 3; ++k) {
    a.p[k] = plane_p ? plane_p[k] : 0.f; a.n[k] = plane_n[k]; a.t1[k] = t1[k]; a.t2[k] = t2[k];
  }
  a.x = x; a.v = v; a.e = e; a.e0 = e0; a.N = N; a.layout = layout; a.res = res;
  a.offset = 0.5f * span;
  a.rcp_h = (float) res / span;
  if (!backward) {
    if (!image) return fail(DRRT_ERR_INVALID_ARGUMENT, "image is NULL");
    if (!(flags & DRRT_ACCUMULATE)) CU_TRY(cudaMemsetAsync(image, 0, sizeof(float) * (size_t) res * res, s));
  } else if (!dimage) {
    return fail(DRRT_ERR_INVALID_ARGUMENT, "dimage is NULL");
  }
  if (N == 0) return DRRT_OK;
  if (!v || (!far && !x)) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL ray buffer");
  a.image = image; a.dimage = dimage; a.dx = far ? nullptr : dx; a.dv = dv; a.de = de;
  if (backward) {
    if (far && dx) CU_TRY(cudaMemsetAsync(dx, 0, sizeof(float) * 3 * (size_t) N, s));  // far field ignores x
    g_launches += launch_sensor_splat_bwd(far, a, s);
  } else {
    g_launches += launch_sensor_splat(far, a, s);
  }
  return check_launch("sensor_splat");
}
}  // namespace

namespace {
int gather_common(bool backward, const float *x, const float *v, int64_t N, int layout, const float *plane_p,
                  const float *plane_n, const float *t1, const float *t2, const float *tex, int res, float span,
                  unsigned flags, float *f, const float *df, float *dx, float *dv, cudaStream_t s) {
  const bool far = (flags & DRRT_SENSOR_FAR) != 0;
  if (N < 0) return fail(DRRT_ERR_INVALID_ARGUMENT, "negative ray count");
  if (layout != DRRT_SOA && layout != DRRT_AOS) return fail(DRRT_ERR_INVALID_ARGUMENT, "layout must be DRRT_SOA or DRRT_AOS");
  if (res <= 0 || !(span > 0.f)) return fail(DRRT_ERR_INVALID_ARGUMENT, "sensor: res and span must be positive");
  if (!plane_n || !t1 || !t2 || (!far && !plane_p) || !tex) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL plane vector or texture");
  if (N == 0) return DRRT_OK;
  if (!v || (!far && !x) || (backward ? !df : !f)) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL ray buffer");
  SensorArgs a;
  memset(&a, 0, sizeof a);
  for (int k = 0; k < 3; ++k) {
    a.p[k] = plane_p ? plane_p[k] : 0.f; a.n[k] = plane_n[k]; a.t1[k] = t1[k]; a.t2[k] = t2[k];
  }
  a.x = x; a.v = v; a.N = N; a.layout = layout; a.res = res;
  a.offset = 0.5f * span;
  a.rcp_h = (float) res / span;
  a.dimage = tex;
  if (backward) {
    a.e = df; a.dx = far ? nullptr : dx; a.dv = dv;
    if (far && dx) CU_TRY(cudaMemsetAsync(dx, 0, sizeof(float) * 3 * (size_t) N, s));
    g_launches += launch_sensor_gather_bwd(far, a, s);
  } else {
    a.de = f;
    g_launches += launch_sensor_gather(far, a, s);
  }
  return check_launch("sensor_gather");
}
}  // namespace

int drrt_sensor_gather(const float *x, const float *v, int64_t N, int layout, const float *plane_p,
                       const float *plane_n, const float *t1, const float *t2, const float *tex, int res,
                       float span, float *f, unsigned flags, drrt_stream_t stream) {
  return gather_common(false, x, v, N, layout, plane_p, plane_n, t1, t2, tex, res, span, flags, f, nullptr, nullptr,
                       nullptr, (cudaStream_t) stream);
}

int drrt_sensor_gather_backward(const float *x, const float *v, int64_t N, int layout, const float *plane_p,
                                const float *plane_n, const float *t1, const float *t2, const float *tex, int res,
                                float span, const float *df, float *dx, float *dv, unsigned flags,
                                drrt_stream_t stream) {
  return gather_common(true, x, v, N, layout, plane_p, plane_n, t1, t2, tex, res, span, flags, nullptr, df, dx, dv,
                       (cudaStream_t) stream);
}

int drrt_sensor_splat(const float *x, const float *v, const float *e, float e0, int64_t N, int layout,
                      const float *plane_p, const float *plane_n, const float *t1, const float *t2, int res,
                      float span, float *image, unsigned flags, drrt_stream_t stream) {
  return sensor_common(false, x, v, e, e0, N, layout, plane_p, plane_n, t1, t2, res, span, flags, image, nullptr,
                       nullptr, nullptr, nullptr, (cudaStream_t) stream);
}

int drrt_sensor_splat_backward(const float *x, const float *v, const float *e, float e0, int64_t N, int layout,
                               const float *plane_p, const float *plane_n, const float *t1, const float *t2,
                               int res, float span, const float *dimage, float *dx, float *dv, float *de,
                               unsigned flags, drrt_stream_t stream) {
  return sensor_common(true, x, v, e, e0, N, layout, plane_p, plane_n, t1, t2, res, span, flags, nullptr, dimage,
                       dx, dv, de, (cudaStream_t) stream);
}

int drrt_source_rand(int kind, const float *offset, uint64_t seed, const float *view_rot, int nviews, int P0,
                     int P1, int spp, float width, float sensor_dist, float cone_angle_deg, int independent,
                     const float *post_rot, float post_span, float *x, float *v, float *planes, uint8_t *keep,
                     float *tpv, float *target, drrt_stream_t stream) {
  if (kind < DRRT_SOURCE_PLANE || kind > DRRT_SOURCE_POINT_LATTICE)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "ray source: unknown kind");
  if (nviews < 0 || P0 <= 0 || P1 <= 0 || spp <= 0) return fail(DRRT_ERR_INVALID_ARGUMENT, "ray source: bad view / pixel / sample count");
  if (nviews == 0) return DRRT_OK;
  if (!view_rot || !x || !v || !planes) return fail(DRRT_ERR_INVALID_ARGUMENT, "ray source: NULL buffer");
  SourceArgs a;
  memset(&a, 0, sizeof a);
  a.kind = kind;
  a.cone_cos = (float) cos(0.5 * (double) cone_angle_deg * 3.14159265358979323846 / 180.0);
  a.cone_half = (float) (0.5 * (double) cone_angle_deg * 3.14159265358979323846 / 180.0);
  a.offset = offset; a.seed = seed; a.P0 = P0; a.P1 = P1; a.spp = spp;
  a.rays_per_view = (int64_t) spp * P0 * P1;
  a.width = width; a.sensor_dist = sensor_dist; a.independent = independent;
  if (post_rot) {
    a.has_post = 1; a.post_span = post_span;
    memcpy(a.post_rot, post_rot, sizeof(float) * 9);
  }
  for (int first = 0; first < nviews; first += kSourceMaxViews) {
    const int nv = nviews - first < kSourceMaxViews ? nviews - first : kSourceMaxViews;
    a.first_view = first; a.nviews = nv;
    memcpy(a.view_rot, view_rot + 9 * (size_t) first, sizeof(float) * 9 * (size_t) nv);
    const size_t r0 = (size_t) first * (size_t) a.rays_per_view;
    a.x = x + 3 * r0; a.v = v + 3 * r0; a.planes = planes + 9 * r0;
    a.keep = keep ? keep + r0 : nullptr;
    a.tpv = tpv ? tpv + r0 : nullptr;
    a.target = target ? target + 3 * r0 : nullptr;
    g_launches += launch_plane_source(a, (cudaStream_t) stream);
  }
  return check_launch("ray_source");
}

int drrt_plane_source_rand(const float *offset, uint64_t seed, const float *view_rot, int nviews, int P0, int P1,
                           int spp, float width, float sensor_dist, int independent, const float *post_rot,
                           float post_span, float *x, float *v, float *planes, uint8_t *keep,
                           drrt_stream_t stream) {
  return drrt_source_rand(DRRT_SOURCE_PLANE, offset, seed, view_rot, nviews, P0, P1, spp, width, sensor_dist, 0.f,
                          independent, post_rot, post_span, x, v, planes, keep, nullptr, nullptr, stream);
}

int drrt_adam_step_slab(float *n, const float *grad, float *exp_avg, float *exp_avg_sq, int S0, int S1, int S2,
                        int i0_begin, int i0_count, int step, float lr, float beta1, float beta2, float eps,
                        float clamp_min, unsigned flags, drrt_stream_t stream) {
  if (S0 <= 0 || S1 <= 0 || S2 <= 0) return fail(DRRT_ERR_RESOLUTION, "Resolution doesn't match data");
  if (i0_begin < 0 || i0_count < 0 || (int64_t) i0_begin + i0_count > S0)
    return fail(DRRT_ERR_RESOLUTION, "slab outside the grid");
  if (i0_count == 0) return DRRT_OK;
  if (!n || !grad || !exp_avg || !exp_avg_sq) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL grid buffer");
  if (step < 1) return fail(DRRT_ERR_INVALID_ARGUMENT, "step counts from 1 (torch.optim.Adam)");
  int r = launch_adam_step(n, grad, exp_avg, exp_avg_sq, S0, S1, S2, i0_begin, i0_count, step, lr, beta1, beta2, eps,
                           clamp_min, !(flags & DRRT_ADAM_KEEP_BOUNDARY), !(flags & DRRT_ADAM_NO_CLAMP),
                           (cudaStream_t) stream);
  if (r < 0) return fail(DRRT_ERR_RESOLUTION, "grid too large for one launch");
  g_launches += r;
  return check_launch("adam_step");
}

int drrt_adam_step_peers(float *const *n_replicas, const float *const *grad_replicas, int nranks, int rank,
                         float *exp_avg_slab, float *exp_avg_sq_slab, int S0, int S1, int S2, int i0_begin,
                         int i0_count, int step, float lr, float beta1, float beta2, float eps, float clamp_min,
                         unsigned flags, drrt_stream_t stream) {
  if (S0 <= 0 || S1 <= 0 || S2 <= 0) return fail(DRRT_ERR_RESOLUTION, "Resolution doesn't match data");
  if (i0_begin < 0 || i0_count < 0 || (int64_t) i0_begin + i0_count > S0)
    return fail(DRRT_ERR_RESOLUTION, "slab outside the grid");
  if (i0_count == 0) return DRRT_OK;
  if (!n_replicas || !grad_replicas || !exp_avg_slab || !exp_avg_sq_slab)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL grid buffer");
  if (step < 1) return fail(DRRT_ERR_INVALID_ARGUMENT, "step counts from 1 (torch.optim.Adam)");
  int r = launch_adam_step_peers(n_replicas, grad_replicas, nranks, rank, exp_avg_slab, exp_avg_sq_slab, S0, S1, S2,
                                 i0_begin, i0_count, step, lr, beta1, beta2, eps, clamp_min,
                                 !(flags & DRRT_ADAM_KEEP_BOUNDARY), !(flags & DRRT_ADAM_NO_CLAMP), (cudaStream_t) stream);
  if (r == -2) return fail(DRRT_ERR_INVALID_ARGUMENT, "adam_step_peers: 1..16 ranks, non-NULL replica pointers");
  if (r < 0) return fail(DRRT_ERR_RESOLUTION, "grid too large for one launch");
  g_launches += r;
  return check_launch("adam_step_peers");
}

int drrt_adam_step_multimem(float *n_multicast, const float *grad_multicast, const float *n_local, float *exp_avg_slab,
                            float *exp_avg_sq_slab, int S0, int S1, int S2, int i0_begin, int i0_count, int step, float lr,
                            float beta1, float beta2, float eps, float clamp_min, unsigned flags, drrt_stream_t stream) {
  if (S0 <= 0 || S1 <= 0 || S2 <= 0) return fail(DRRT_ERR_RESOLUTION, "Resolution doesn't match data");
  if (i0_begin < 0 || i0_count < 0 || (int64_t) i0_begin + i0_count > S0)
    return fail(DRRT_ERR_RESOLUTION, "slab outside the grid");
  if (i0_count == 0) return DRRT_OK;
  if (!n_multicast || !grad_multicast || !n_local || !exp_avg_slab || !exp_avg_sq_slab)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL grid buffer");
  if (step < 1) return fail(DRRT_ERR_INVALID_ARGUMENT, "step counts from 1 (torch.optim.Adam)");
  int r = launch_adam_step_multimem(n_multicast, grad_multicast, n_local, exp_avg_slab, exp_avg_sq_slab, S0, S1, S2, i0_begin,
                                    i0_count, step, lr, beta1, beta2, eps, clamp_min, !(flags & DRRT_ADAM_KEEP_BOUNDARY),
                                    !(flags & DRRT_ADAM_NO_CLAMP), (cudaStream_t) stream);
  if (r == -2) return fail(DRRT_ERR_INVALID_ARGUMENT, "adam_step_multimem: rows of a multiple of 4 floats, 16-byte aligned buffers");
  if (r < 0) return fail(DRRT_ERR_RESOLUTION, "grid too large for one launch");
  g_launches += r;
  return check_launch("adam_step_multimem");
}

int drrt_adam_step(float *n, const float *grad, float *exp_avg, float *exp_avg_sq, int S0, int S1, int S2,
                   int step, float lr, float beta1, float beta2, float eps, float clamp_min,
                   unsigned flags, drrt_stream_t stream) {
  return drrt_adam_step_slab(n, grad, exp_avg, exp_avg_sq, S0, S1, S2, 0, S0, step, lr, beta1, beta2, eps, clamp_min,
                             flags, stream);
}

int drrt_upres_volume(const float *src, int s0, int s1, int s2, float *dst, int d0, int d1, int d2,
                      drrt_stream_t stream) {
  if (s0 <= 0 || s1 <= 0 || s2 <= 0 || d0 <= 0 || d1 <= 0 || d2 <= 0)
    return fail(DRRT_ERR_RESOLUTION, "Resolution doesn't match data");
  if (!src || !dst) return fail(DRRT_ERR_INVALID_ARGUMENT, "NULL grid buffer");
  g_launches += launch_upres_volume(src, s0, s1, s2, dst, d0, d1, d2, (cudaStream_t) stream);
  return check_launch("upres_volume");
}

int drrt_volume_eval(const float *rif, int W, int H, int D, const float *pos, int64_t N,
                     int layout, float h, float *n, float *grad, float *hess, unsigned flags,
                     drrt_stream_t stream) {
  GridP g;
  int rc = make_grid(g, rif, W, H, D, h);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, 1.0f))) return rc;
  g_launches += launch_volume_eval(g, pos, N, layout, n, grad, hess, !(flags & DRRT_FAST_ARITH),
                                   (cudaStream_t) stream);
  return check_launch("volume_eval");
}

int drrt_volume_splat(float *data, int W, int H, int D, const float *pos, const float *val,
                      const float *grad, int64_t N, int layout, float h, drrt_stream_t stream) {
  GridP g;
  int rc = make_grid(g, data, W, H, D, h);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, 1.0f))) return rc;
  g_launches += launch_volume_splat(g, data, pos, val, grad, N, layout, (cudaStream_t) stream);
  return check_launch("volume_splat");
}

int drrt_cylinder_eval(const float *rif, int res, float radius, float length, const float *pos,
                       int64_t N, int layout, float *n, float *grad, float *hess,
                       drrt_stream_t stream) {
  CableP c;
  int rc = make_cable(c, rif, res, radius, length);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, 1.0f))) return rc;
  g_launches += launch_cylinder_eval(c, pos, N, layout, n, grad, hess, (cudaStream_t) stream);
  return check_launch("cylinder_eval");
}

int drrt_cylinder_splat(float *data, int res, float radius, float length, const float *pos,
                        const float *val, const float *grad, int64_t N, int layout,
                        drrt_stream_t stream) {
  CableP c;
  int rc = make_cable(c, data, res, radius, length);
  if (rc) return rc;
  if ((rc = check_rays(N, layout, 1.0f))) return rc;
  g_launches += launch_cylinder_splat(c, data, pos, val, grad, N, layout, (cudaStream_t) stream);
  return check_launch("cylinder_splat");
}

// In-switch sum of the gradient grid (drrt_misc_kernels.cu: allreduce_multimem_kernel).
int drrt_allreduce_multimem(float *multicast_grad, size_t count, int rank, int nranks, drrt_stream_t stream) {
  if (!multicast_grad || nranks <= 0 || rank < 0 || rank >= nranks)
    return fail(DRRT_ERR_INVALID_ARGUMENT, "allreduce_multimem: bad argument");
  if (count % 4 || ((uintptr_t) multicast_grad & 15))
    return fail(DRRT_ERR_INVALID_ARGUMENT, "allreduce_multimem: count must be a multiple of 4 floats and the buffer 16-byte aligned");
  const int64_t n4 = (int64_t) (count / 4);
  const int64_t per = (n4 + nranks - 1) / nranks;
  const int64_t first = per * rank;
  const int64_t mine = first >= n4 ? 0 : (first + per > n4 ? n4 - first : per);
  g_launches += launch_allreduce_multimem(multicast_grad, first, mine, (cudaStream_t) stream);
  return check_launch("allreduce_multimem");
}

// ---------------------------------------------------------------------------
// drrt_backtrace_noalloc: drrt_backtrace for callers that cannot allocate (code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch that drrt_backtrace takes from its caller (the int64 grid of DRRT_DETERMINISTIC, or the private gradient
// copies of small grids) lives in a file-scope device array of a fixed, modest size: every GPU that loads the library
// holds it (17.2 MB) until the process exits.  `g_noalloc_scratch` is a HOST-side stand-in in host code, not a device
// address: its device address is resolved once per device with cudaGetSymbolAddress (with that device current) and that
// pointer is what the kernels get.  Every call shares the array and finds the previous call's contents in it: the part
// a call uses is cleared (cudaMemsetAsync on `stream`) before its kernels run, and calls that overlap on the GPU
// (different streams, no ordering between them) race.
// ---------------------------------------------------------------------------
namespace {
constexpr size_t kNoAllocScratchBytes = sizeof(unsigned long long) * (size_t) DRRT_NOALLOC_MAX_VOXELS + 256;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];
constexpr int kNoAllocMaxDevices = 64;
std::atomic<void *> g_noalloc_ptr[kNoAllocMaxDevices];

// Device address of the scratch array on the CURRENT device (resolved once per device).
int noalloc_scratch(void **out) {
  int dev = -1;
  CU_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(DRRT_ERR_INVALID_ARGUMENT, "drrt_backtrace_noalloc: device index out of range");
  void *p = g_noalloc_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return fail(DRRT_ERR_CUDA, "drrt_backtrace_noalloc: cudaGetSymbolAddress(scratch): %s", cudaGetErrorString(e));
    g_noalloc_ptr[dev].store(p, std::memory_order_release);  // the same address whichever thread resolves it
  }
  *out = p;
  return DRRT_OK;
}
}  // namespace

int drrt_backtrace_noalloc(const float *rif, int W, int H, int D, const float *xt, const float *vt, const float *dx,
                           const float *dv, int64_t N, int layout, float h, float ds, float *grad, int64_t *stats,
                           unsigned flags, drrt_stream_t stream) {
  if (W <= 0 || H <= 0 || D <= 0) return fail(DRRT_ERR_RESOLUTION, "volume: invalid resolution!");
  const int64_t G = (int64_t) W * H * D;
  if (G > (int64_t) DRRT_NOALLOC_MAX_VOXELS)
    return fail(DRRT_ERR_INVALID_ARGUMENT,
                "drrt_backtrace_noalloc: a grid of %lld voxels exceeds the built-in scratch (DRRT_NOALLOC_MAX_VOXELS = %d); "
                "use drrt_backtrace with a caller-owned scratch buffer", (long long) G, (int) DRRT_NOALLOC_MAX_VOXELS);
  void *scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  // the ray interleave's device-side estimate takes a stream-ordered allocation: never on this path
  flags = (flags & ~(DRRT_AUTO_INTERLEAVE | (0x3fu << 12))) | DRRT_NO_INTERLEAVE;
  return backtrace_common(false, rif, nullptr, W, H, D, xt, vt, dx, dv, N, layout, h, ds, grad, scratch,
                          kNoAllocScratchBytes, stats, flags, (cudaStream_t) stream);
}

}  // extern "C"
