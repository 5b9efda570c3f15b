/*
 * drrt_b200.h -- C ABI of libdrrt_b200.so: the eikonal ray tracer and its adjoint
 * as hand-written sm_100a CUDA kernels.
 *
 * Each entry point replaces one method of the reference's pybind11 classes
 * drrt.TracerC / TracerD / TracerS (reference src/drrt.cpp:28-58; C++ signatures
 * include/tracer.h:15-89).  Argument order follows the reference method; enoki
 * array arguments become plain pointers + sizes.
 *
 * Conventions
 *   - float32 everywhere; the grid is flat, idx = (z*H + y)*W + x with
 *     (W,H,D) = the `res` triple the reference passes (src/volume.cpp:134).
 *   - Ray vectors: `layout` = DRRT_AOS for (N,3) row-major (what torch hands to
 *     core/tracer.py) or DRRT_SOA for [3][N] (enoki's internal layout).
 *   - Functions without the `_host` suffix take DEVICE pointers and enqueue work
 *     on `stream` (a cudaStream_t; NULL = legacy default stream) without
 *     synchronising.  The caller owns every buffer; the only memory the device entry points
 *     obtain themselves is stream-ordered and short-lived (cudaMallocAsync / cudaFreeAsync on
 *     `stream`, no synchronisation): 28 N + 16 bytes of per-ray tail state in drrt_trace_plane /
 *     _target / _sdf (the batch-coupled iterations of the reference, src/tracer.cpp:144-152,
 *     :216-231, :279-303) and one int for the opt-in ray interleave of drrt_backtrace.
 *   - `_host` functions take HOST pointers (pinned or pageable), copy in, run,
 *     copy out and synchronise before returning.
 *   - Return value: DRRT_OK or a negative error code; drrt_last_error() gives the
 *     message of the calling thread's last failure.  The messages for size /
 *     resolution mismatches are the reference's (src/volume.cpp:28,37,115,124).
 *   - `stats` (nullable, device pointer for device entry points, host pointer for
 *     `_host` ones) receives two int64: [0] += live ray-steps integrated,
 *     [1] += rays that never exited (the reference only prints
 *     "failed to exit all rays", src/tracer.cpp:89-96).  Device stats are
 *     accumulated, not overwritten: zero them yourself.
 */
#ifndef DRRT_B200_H
#define DRRT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DRRT_ABI_VERSION 2

typedef void *drrt_stream_t; /* cudaStream_t */

enum {
  DRRT_OK = 0,
  DRRT_ERR_INVALID_ARGUMENT = -1,
  DRRT_ERR_RESOLUTION = -2, /* "Resolution doesn't match data" / "volume: invalid resolution!" */
  DRRT_ERR_CUDA = -3,
  DRRT_ERR_NCCL = -4
};

enum { DRRT_SOA = 1, DRRT_AOS = 3 };

/* flags (bitwise or) */
enum {
  /* Arithmetic.  Default (neither bit set): "exact trajectory" -- everything that
   * decides where a ray goes (cell lookup, n, grad n, the v and x updates, every
   * branch) is evaluated in the reference's literal operation order with no FMA
   * contraction, so ray endpoints, directions and the cells visited are
   * bit-identical to a float32 CPU evaluation of the reference formulas; the
   * terms that only feed the linear gradient accumulation (Hessian, splat
   * weights, lambda/mu updates) use factored FMA forms, which changes the
   * gradient at the 1e-6 relative level (same class as atomic reordering).
   * DRRT_REFERENCE_ARITH: literal operation order everywhere, 16 atomics per
   * step like the reference (slow; for audits).
   * DRRT_FAST_ARITH: the trilinear cell as a factored polynomial everywhere
   * (~1.5x faster forward); float32 rounding then differs from the reference and a
   * few rays per 10^4 take the other branch at a cell face. */
  DRRT_REFERENCE_ARITH = 1u << 0,
  DRRT_FAST_ARITH = 1u << 3,
  /* Adjoint only, A/B switches for profiling (defaults: face-carry flush when ds > h/6, private
   * gradient copies on small grids, no interleave, no warp aggregation).
   * DRRT_NO_INTERLEAVE: ray i on thread i instead of the strided assignment that puts
   * the lanes of a warp into different cells.
   * DRRT_WARP_AGGREGATION: pre-sum lanes that flush the same cell with a segmented
   * shuffle reduction before the atomics (the alternative to interleaving).
   * bits 8-11  DRRT_AGG_RUN(n), n in {2,4,8} or DRRT_AGG_RUN_WARP: fixed maximum run
   *            length for the aggregation; 0 = adaptive (whole warp when all flushing
   *            lanes sit in one cell -- focused beams).
   * bits 12-17 DRRT_INTERLEAVE(n), n in {1,2,4,8,16,32}: fixed interleave; 0 = measured. */
  DRRT_WARP_AGGREGATION = 1u << 4,
  DRRT_NO_INTERLEAVE = 1u << 5,
  /* DRRT_NO_FACE_CARRY: flush all eight corner accumulators on every cell change instead of
   * keeping the four of the face shared with the next cell in registers (A/B switch; this
   * also restores the device-measured interleave as the default, i.e. the r03/r04 kernel).
   * DRRT_AUTO_INTERLEAVE: measure the interleave on the device even with the face carry. */
  DRRT_NO_FACE_CARRY = 1u << 7,
  /* DRRT_FACE_CARRY: use the face carry whatever the step size (the default uses it when a ray takes
   * fewer than 6 steps per cell, ds > h/6, where it pays). */
  DRRT_FACE_CARRY = 1u << 24,
  /* DRRT_NO_REPLICAS: one gradient grid for all CTAs even where the library would use private
   * copies (grids up to 2^23 / 2 voxels; see drrt_backtrace_scratch_bytes). */
  DRRT_AUTO_INTERLEAVE = 1u << 18,
  DRRT_NO_REPLICAS = 1u << 19,
  /* bits 20-23  DRRT_REPLICAS(n), n in 2..15: force n gradient copies whatever the grid size
   *             (a beam focused inside a LARGE grid; the scratch buffer grows to (n-1) grids). */
  /* drrt_trace only (literal arithmetic): DRRT_PAIR_RAYS traces two rays per thread so that every
   * float32 operation of the reference formulas issues as packed FP32 (FMUL2 / FFMA2, one instruction
   * per two rays; results bit-identical to one ray per thread); DRRT_NO_PAIR_RAYS forces one ray per
   * thread.  Neither bit: the library's choice ("Ray pairs"). */
  /* Opt-in: drrt_trace / drrt_backtrace on grids of up to 50 000 voxels with at most 512 rays per SM gather from a
   * per-CTA shared-memory copy of the grid (results identical; measured SLOWER than the default on the 32^3 / 256^2
   * configuration, profiles/r12_smem_grid_rejected.log, hence not the default). */
  DRRT_SMEM_GRID = 1u << 27,
  DRRT_PAIR_RAYS = 1u << 25,
  DRRT_NO_PAIR_RAYS = 1u << 26,
  /* Adjoint only: accumulate the gradient in 64-bit fixed point so the result is
   * bit-stable from run to run regardless of atomic ordering (and of the ray
   * order / interleave). */
  DRRT_DETERMINISTIC = 1u << 1,
  /* Adjoint only: do NOT zero-fill `grad` first (the reference always returns a
   * fresh grid, src/tracer.cpp:401-403; set this to accumulate several ray
   * batches into one grid). */
  DRRT_ACCUMULATE = 1u << 2
};

#define DRRT_AGG_RUN(n) ((unsigned) (n) << 8)
#define DRRT_AGG_RUN_WARP (15u << 8)
#define DRRT_REPLICAS(n) ((unsigned) (n) << 20)
#define DRRT_INTERLEAVE(n) ((unsigned) (n) << 12)

int drrt_abi_version(void);
const char *drrt_last_error(void);
/* Number of kernels launched by this library in this process so far. */
int64_t drrt_launch_count(void);

/* ---- forward traces -------------------------------------------------------
 * drrt_trace        <- TracerC.trace        (src/drrt.cpp:51, src/tracer.cpp:35-100)
 * drrt_trace_plane  <- TracerC.trace_pln    (src/drrt.cpp:52, src/tracer.cpp:102-172)
 * drrt_trace_target <- TracerC.trace_target (src/drrt.cpp:54, src/tracer.cpp:174-242)
 * drrt_trace_sdf    <- TracerC.trace_sdf    (src/drrt.cpp:53, src/tracer.cpp:244-310)
 * drrt_trace_cable  <- TracerC.trace_cable  (src/drrt.cpp:55, src/tracer.cpp:312-382)
 */
int drrt_trace(const float *rif, int W, int H, int D, const float *pos, const float *vel,
               int64_t N, int layout, float h, float ds, float *xt, float *vt, int64_t *stats,
               unsigned flags, drrt_stream_t stream);

int drrt_trace_plane(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                     const float *pln_o, const float *pln_d, int64_t N, int layout, float h,
                     float ds, float *xt, float *vt, uint8_t *not_escaped, int64_t *stats,
                     unsigned flags, drrt_stream_t stream);

int drrt_trace_target(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                      const float *target, int64_t N, int layout, float h, float ds, float *xt,
                      float *vt, float *dist2, int64_t *stats, unsigned flags,
                      drrt_stream_t stream);

int drrt_trace_sdf(const float *rif, const float *sdf, int W, int H, int D, const float *pos,
                   const float *vel, int64_t N, int layout, float h, float ds, float *xt,
                   float *vt, int64_t *stats, unsigned flags, drrt_stream_t stream);

int drrt_trace_cable(const float *rif, int res, float radius, float length, const float *pos,
                     const float *vel, const float *target, int64_t N, int layout, float ds,
                     float *xt, float *vt, float *dist2, int64_t *stats, unsigned flags,
                     drrt_stream_t stream);

/* ---- adjoints -------------------------------------------------------------
 * drrt_backtrace       <- TracerC.backtrace       (src/drrt.cpp:56, src/tracer.cpp:384-440)
 * drrt_backtrace_sdf   <- TracerC.backtrace_sdf   (src/drrt.cpp:57, src/tracer.cpp:443-509)
 * drrt_backtrace_cable <- TracerC.backtrace_cable (src/drrt.cpp:58, src/tracer.cpp:511-567)
 * `grad` has W*H*D (resp. `res`) floats.  `scratch`: a device buffer of
 * drrt_backtrace_scratch_bytes(grid_elems, flags) bytes, or NULL when that is 0.
 * DRRT_DETERMINISTIC needs it (the int64 grid).  Without that flag it holds the private
 * gradient copies of small grids (CTA b adds into copy b % R, one pass sums them: R x
 * the atomic throughput on the voxels where a focused or crossing batch piles up);
 * passing NULL there is allowed and simply means one copy.
 */
size_t drrt_backtrace_scratch_bytes(int64_t grid_elems, unsigned flags);

int drrt_backtrace(const float *rif, int W, int H, int D, const float *xt, const float *vt,
                   const float *dx, const float *dv, int64_t N, int layout, float h, float ds,
                   float *grad, void *scratch, size_t scratch_bytes, int64_t *stats, unsigned flags,
                   drrt_stream_t stream);

int drrt_backtrace_sdf(const float *rif, const float *sdf, int W, int H, int D, const float *xt,
                       const float *vt, const float *dx, const float *dv, int64_t N, int layout,
                       float h, float ds, float *grad, void *scratch, size_t scratch_bytes, int64_t *stats,
                       unsigned flags, drrt_stream_t stream);

int drrt_backtrace_cable(const float *rif, int res, float radius, float length, const float *xt,
                         const float *vt, const float *dx, const float *dv, int64_t N, int layout,
                         float ds, float *grad, void *scratch, size_t scratch_bytes, int64_t *stats, unsigned flags,
                         drrt_stream_t stream);

/* drrt_backtrace_noalloc: NO counterpart in the reference -- added for callers
 * that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * drrt_backtrace without the scratch argument: the scratch (the int64 grid of DRRT_DETERMINISTIC, else the private
 * gradient copies of small grids, as many as fit) is a static device array inside the library, so the call makes no
 * cudaMalloc / cudaFree / cudaMallocAsync and uses no allocating library routine.  Limits and rules:
 *   - grids of at most DRRT_NOALLOC_MAX_VOXELS voxels (129^3); larger grids are refused with
 *     DRRT_ERR_INVALID_ARGUMENT (use drrt_backtrace with your own scratch);
 *   - every GPU that loads the library holds the array (8 * DRRT_NOALLOC_MAX_VOXELS + 256 bytes = 17.2 MB) until exit;
 *   - all calls share the array, which keeps the last call's contents: each call clears the part it uses on `stream`
 *     before its kernels; calls that OVERLAP on the GPU (different streams with no ordering between them) race;
 *   - the ray interleave is off on this path (DRRT_INTERLEAVE / DRRT_AUTO_INTERLEAVE are ignored: its estimate
 *     allocates); results equal drrt_backtrace's with the same flags | DRRT_NO_INTERLEAVE (bit-identical in
 *     deterministic mode, atomic-order noise otherwise);
 *   - the first call on a device resolves the array's device address (cudaGetSymbolAddress; with lazy module loading the
 *     driver loads the module then): make one call per device before starting a graph capture.
 * Inputs and outputs stay the caller's, as for drrt_backtrace. */
#define DRRT_NOALLOC_MAX_VOXELS 2146689
int drrt_backtrace_noalloc(const float *rif, int W, int H, int D, const float *xt, const float *vt, const float *dx,
                           const float *dv, int64_t N, int layout, float h, float ds, float *grad, int64_t *stats,
                           unsigned flags, drrt_stream_t stream);

/* drrt_backtrace_rays <- the ray-state half of ADTracerC.backward / ADSDFTracerC.backward
 * (reference core/tracer.py:53-60, :247-291: enoki.gradient(ctx.x), enoki.gradient(ctx.v)):
 * the same adjoint pass as drrt_backtrace (sdf == NULL) / drrt_backtrace_sdf, which in
 * addition writes dL/dpos and dL/dvel of the LAUNCH state (N x 3 each, `layout`) from the
 * final co-state: dpos = lambda, dvel = mu + (k - 1) ds lambda, k = number of straight
 * steps the forward ray took outside the volume before its first in-bounds sample
 * (recovered from `pos0`, the launch positions; pos0 == NULL means k = 0, i.e. rays
 * launched on their first in-bounds sample).  Exact for rays launched outside or on the
 * boundary of the volume -- the only kind the reference's sources produce. */
int drrt_backtrace_rays(const float *rif, const float *sdf, int W, int H, int D, const float *xt,
                        const float *vt, const float *dx, const float *dv, const float *pos0, int64_t N,
                        int layout, float h, float ds, float *grad, float *dpos, float *dvel,
                        void *scratch, size_t scratch_bytes, int64_t *stats, unsigned flags, drrt_stream_t stream);

/* ---- volume primitives (parity probes; reference src/test.cpp:61-101) -------
 * drrt_volume_eval   <- volume::eval_grad + eval_hess (src/volume.cpp:101-181, :40-99)
 *                       hess = (N,3): d2/dxdy, d2/dxdz, d2/dydz
 * drrt_volume_splat  <- volume::splat (src/volume.cpp:182-244), accumulates into `data`
 * drrt_cylinder_eval <- cylinder_volume::eval_grad + eval_hess
 *                       (src/cylinder_volume.cpp:26-111), hess = (N,9) row-major
 * drrt_cylinder_splat<- cylinder_volume::splat (src/cylinder_volume.cpp:113-148)
 */
int drrt_volume_eval(const float *rif, int W, int H, int D, const float *pos, int64_t N,
                     int layout, float h, float *n, float *grad, float *hess, unsigned flags,
                     drrt_stream_t stream);
int drrt_volume_splat(float *data, int W, int H, int D, const float *pos, const float *val,
                      const float *grad, int64_t N, int layout, float h, drrt_stream_t stream);
int drrt_cylinder_eval(const float *rif, int res, float radius, float length, const float *pos,
                       int64_t N, int layout, float *n, float *grad, float *hess,
                       drrt_stream_t stream);
int drrt_cylinder_splat(float *data, int res, float radius, float length, const float *pos,
                        const float *val, const float *grad, int64_t N, int layout,
                        drrt_stream_t stream);

/* ---- optimiser step and multires up-res (the callers' side of the adjoint) -----
 * drrt_adam_step    <- one iteration of the loop body of multires_opt, reference
 *                      core/optimizer.py:61-68, fused into one pass over the grid:
 *                      n.grad[boundary] = 0 (mask = everything but [1:-1,1:-1,1:-1]),
 *                      torch.optim.Adam step (no weight decay / amsgrad; `step` counts
 *                      from 1), n.clamp_(min = clamp_min).  (S0,S1,S2) = n.shape.
 * drrt_upres_volume <- upres_scene / grid.upres_volume (core/optimizer.py:7-10,
 *                      core/grid.py:318-330): trilinear resample on linspace(0,1)
 *                      lattices, evaluated in float64 like the reference. */
enum { DRRT_ADAM_KEEP_BOUNDARY = 1u << 0, DRRT_ADAM_NO_CLAMP = 1u << 1 };
int drrt_adam_step(float *n, const float *grad, float *exp_avg, float *exp_avg_sq, int S0, int S1, int S2,
                   int step, float lr, float beta1, float beta2, float eps, float clamp_min,
                   unsigned flags, drrt_stream_t stream);
/* The same step on planes [i0_begin, i0_begin + i0_count) of the (S0,S1,S2) grid only: all four pointers address
 * the FIRST element of that slab (each rank's share of the sharded optimiser step, see drrt_reduce_scatter_grad);
 * the zeroed boundary is that of the whole grid. */
int drrt_adam_step_slab(float *n, const float *grad, float *exp_avg, float *exp_avg_sq, int S0, int S1, int S2,
                        int i0_begin, int i0_count, int step, float lr, float beta1, float beta2, float eps,
                        float clamp_min, unsigned flags, drrt_stream_t stream);
/* The sharded step as ONE kernel over peer memory (NVLink P2P) instead of reduce-scatter + slab step + all-gather:
 * `n_replicas` / `grad_replicas` are HOST arrays of `nranks` device pointers to every rank's full grid and full
 * private gradient grid (peer mappings of the other ranks' buffers, e.g. CUDA IPC).  For each voxel of this rank's
 * slab the kernel sums the nranks gradients in rank order, applies the step with this rank's slab of the Adam state
 * and writes the new value into all nranks replicas.  The caller must order it after every rank's adjoint and before
 * every rank's next read of the grid (two stream-ordered barriers; dist.PeerShardedAdam). */
int drrt_adam_step_peers(float *const *n_replicas, const float *const *grad_replicas, int nranks, int rank,
                         float *exp_avg_slab, float *exp_avg_sq_slab, int S0, int S1, int S2, int i0_begin,
                         int i0_count, int step, float lr, float beta1, float beta2, float eps, float clamp_min,
                         unsigned flags, drrt_stream_t stream);
/* The same sharded step with the NVSwitch doing the gradient sum and the broadcast (NVLink SHARP): `n_multicast` /
 * `grad_multicast` are the MULTICAST addresses of the IOR grid and of the gradient grid (one virtual address bound to
 * every rank's copy; the caller creates the mappings, e.g. torch.distributed._symmetric_memory), `n_local` this rank's
 * own copy of the grid.  One multimem.ld_reduce and one multimem.st per float4 of the slab; S2 % 4 == 0.  Ordering
 * as for drrt_adam_step_peers (dist.SwitchShardedAdam). */
int drrt_adam_step_multimem(float *n_multicast, const float *grad_multicast, const float *n_local, float *exp_avg_slab,
                            float *exp_avg_sq_slab, int S0, int S1, int S2, int i0_begin, int i0_count, int step, float lr,
                            float beta1, float beta2, float eps, float clamp_min, unsigned flags, drrt_stream_t stream);
int drrt_upres_volume(const float *src, int s0, int s1, int s2, float *dst, int d0, int d1, int d2,
                      drrt_stream_t stream);

/* ---- sensor epilogue (the step after every trace in the reference's drivers) -------
 * drrt_sensor_splat <- sensor.generate_sensor (reference core/sensor.py:5-28: rays to the
 *                      sensor plane :195-202, |v.n| foreshortening, projection on the
 *                      tangents, Grid.Splat(average=False) core/grid.py:133-151 with the 4x4
 *                      radial-tent stencil core/grid.py:38-66, :80-85), and with
 *                      DRRT_SENSOR_FAR <- sensor.generate_inf_sensor (core/sensor.py:31-54;
 *                      `span` is then 2 sin(angle_span / 2), x and plane_p are ignored).
 *                      `image` is res x res, row index along t1 (= n x t2), column along t2;
 *                      zero-filled first unless DRRT_ACCUMULATE.  `e` = per-ray radiance
 *                      (device, N floats) or NULL for the constant e0.  The four plane
 *                      vectors are HOST pointers to 3 floats.
 * drrt_sensor_splat_backward <- what torch autograd derives for the same call: given
 *                      dL/dimage writes dL/dx, dL/dv (N x 3, `layout`) and dL/de (N); each
 *                      output is nullable. */
enum { DRRT_SENSOR_FAR = 1u << 6 };
int drrt_sensor_splat(const float *x, const float *v, const float *e, float e0, int64_t N, int layout,
                      const float *plane_p, const float *plane_n, const float *t1, const float *t2, int res,
                      float span, float *image, unsigned flags, drrt_stream_t stream);
int drrt_sensor_splat_backward(const float *x, const float *v, const float *e, float e0, int64_t N, int layout,
                               const float *plane_p, const float *plane_n, const float *t1, const float *t2,
                               int res, float span, const float *dimage, float *dx, float *dv, float *de,
                               unsigned flags, drrt_stream_t stream);

/* drrt_sensor_gather <- sensor.get_sdf_vals_near (core/sensor.py:116-134) and, with
 *     DRRT_SENSOR_FAR, sensor.get_sdf_vals_far (:137-155; `span` = 2 sin(angle_span/2), the RAW
 *     direction is projected, as in the reference): Grid.Get (core/grid.py:106-121) of the
 *     res x res texture `tex` (device) at each ray's plane coordinates, over the same 4x4
 *     radial-tent stencil with indices clamped to the texture.  f: N floats.
 * drrt_sensor_gather_backward <- its torch-autograd backward w.r.t. the rays: given dL/df
 *     (N floats) writes dL/dx, dL/dv (each nullable). */
int drrt_sensor_gather(const float *x, const float *v, int64_t N, int layout, const float *plane_p,
                       const float *plane_n, const float *t1, const float *t2, const float *tex, int res,
                       float span, float *f, unsigned flags, drrt_stream_t stream);
int drrt_sensor_gather_backward(const float *x, const float *v, int64_t N, int layout, const float *plane_p,
                                const float *plane_n, const float *t1, const float *t2, const float *tex, int res,
                                float span, const float *df, float *dx, float *dv, unsigned flags,
                                drrt_stream_t stream);

/* ---- on-device ray generation (the step before every trace in the reference's drivers) ----
 * drrt_plane_source_rand <- source.plane_source3_rand + rotate_pts_to_source (reference
 *     core/source.py:54-69, :275-293) for `nviews` views in one call, which is what
 *     rand_rays_in_sphere (:352-357) and rand_rays_cube (:398-412) concatenate, optionally
 *     followed by random_rotate_ic (:555-563).
 *     view_rot: HOST, nviews x 9 row-major rotation matrices (rotate_ray3, :303-312).
 *     offset:   DEVICE uniforms in [0,1) laid out [nviews][2*spp][P0][P1] (the reference's
 *               torch.rand(2*spp, P0, P1) per view), or NULL to draw them in-kernel from
 *               Philox4x32-10 keyed by (seed, view, sample).
 *     post_rot: HOST 3x3 applied about (post_span/2)^3 to rays and planes, or NULL.
 *     Outputs (device): x, v (n,3); planes (n,3,3) = sensor origin, normal, tangent per ray;
 *     n = nviews*spp*P0*P1, ray index = ((view*spp + s)*P0 + i)*P1 + j.  keep (nullable, n
 *     bytes): 1 where the un-rotated launch point lies inside the circle of diameter `width`
 *     (the reference's circle=True filter, :277-280; the caller compacts). */
int drrt_plane_source_rand(const float *offset, uint64_t seed, const float *view_rot, int nviews, int P0, int P1,
                           int spp, float width, float sensor_dist, int independent, const float *post_rot,
                           float post_span, float *x, float *v, float *planes, uint8_t *keep,
                           drrt_stream_t stream);
/* drrt_source_rand: the same call for the other random sources; kind =
 *   DRRT_SOURCE_PLANE  plane_source3_rand (as above),
 *   DRRT_SOURCE_POINT  <- source.point_source3_rand (core/source.py:72-104, what
 *                      rand_ptrays_in_sphere :360-365 concatenates): rays from the point
 *                      (0,-w/2,0) through a jittered pixel grid; keep = its circle filter,
 *   DRRT_SOURCE_CONE   <- source.cone_source3_rand + hatbox_sample (:186-202, :531-545, what
 *                      rand_ptcone_in_sphere :388-396 concatenates): directions uniform in the
 *                      cone of full angle cone_angle_deg about the view axis; the two uniform
 *                      planes of `offset` are the reference's z and theta draws.
 *   DRRT_SOURCE_AREA_CONE <- source.area_source3_cone (:153-183, what rand_cone_in_sphere :379-386
 *                      concatenates): a cone of directions from every jittered pixel of the
 *                      launch plane; DRRT_SOURCE_AREA_BIAS <- source.area_source3_rand_bias
 *                      (:107-150, rand_area_in_sphere :368-376): each jittered pixel aims at a
 *                      random point of the sensor.  Both use FOUR uniform planes per view in
 *                      `offset` ([nviews][4][spp*P0*P1]: the pixel jitter, then the reference's
 *                      two later draws) and write tpv (N floats: sensor_dist / v.y) and, for
 *                      AREA_BIAS, target (N x 3: the reference's `xt`); both nullable. */
enum { DRRT_SOURCE_PLANE = 0, DRRT_SOURCE_POINT = 1, DRRT_SOURCE_CONE = 2, DRRT_SOURCE_AREA_CONE = 3,
       DRRT_SOURCE_AREA_BIAS = 4,
       /* <- source.point_source3 (core/source.py:29-51, what rand_rays_cube(src_type='point') uses): a point source with
        * a REGULAR lattice of directions theta_i x phi_j in [-cone_angle/2, cone_angle/2]; no jitter (offset / seed are
        * ignored).  Pass P0 = pixels[0] * floor(sqrt(spp)), P1 = pixels[1] * floor(sqrt(spp)), spp = 1. */
       DRRT_SOURCE_POINT_LATTICE = 5 };
int drrt_source_rand(int kind, const float *offset, uint64_t seed, const float *view_rot, int nviews, int P0,
                     int P1, int spp, float width, float sensor_dist, float cone_angle_deg, int independent,
                     const float *post_rot, float post_span, float *x, float *v, float *planes, uint8_t *keep,
                     float *tpv, float *target, drrt_stream_t stream);

/* ---- host-buffer entry points (what a non-torch binding of the reference API
 * calls; device memory is managed inside the library and cached per thread) --- */
int drrt_trace_host(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                    int64_t N, int layout, float h, float ds, float *xt, float *vt,
                    int64_t *stats, unsigned flags);
int drrt_backtrace_host(const float *rif, int W, int H, int D, const float *xt, const float *vt,
                        const float *dx, const float *dv, int64_t N, int layout, float h,
                        float ds, float *grad, int64_t *stats, unsigned flags);
/* Keep `rif` resident on the device between _host calls of this thread: when
 * enabled, the grid is uploaded only if its host pointer or dimensions differ
 * from the previous call, or after drrt_host_invalidate_grid().  Default: off
 * (upload on every call). */
void drrt_host_cache_grid(int enable);
void drrt_host_invalidate_grid(void);
/* Free this thread's cached device buffers and stream. */
void drrt_host_release(void);

/* ---- multi-GPU: one private gradient grid per rank, summed over NVLink ------
 * The communicator is an ncclComm_t created by the caller or by
 * drrt_comm_init_rank() from an id all ranks share. */
#define DRRT_UNIQUE_ID_BYTES 128
int drrt_comm_unique_id(char id[DRRT_UNIQUE_ID_BYTES]);
int drrt_comm_init_rank(void **comm, int nranks, const char id[DRRT_UNIQUE_ID_BYTES], int rank);
int drrt_comm_destroy(void *comm);
int drrt_allreduce_grad(float *grad, size_t count, void *comm, drrt_stream_t stream);
/* The same sum INSIDE the NVSwitch, as this library's own kernel (no NCCL): `multicast_grad` is the multicast
 * address of the gradient grid -- one virtual address bound to every rank's copy (NVLink SHARP; the caller creates the
 * mapping, e.g. with torch.distributed._symmetric_memory) -- and rank r sums slab r with multimem.ld_reduce and
 * broadcasts it with multimem.st.  count % 4 == 0.  The caller must order the call after every rank's adjoint and
 * before any rank reads the result (a barrier on `stream` on each side); single node. */
int drrt_allreduce_multimem(float *multicast_grad, size_t count, int rank, int nranks, drrt_stream_t stream);
/* Sharded optimiser step for grids whose Adam pass is worth splitting (the 512^3
 * configuration): the same bytes on the wire as the all-reduce, 1/nranks of the optimiser work and state.
 *   drrt_reduce_scatter_grad: in place; afterwards grad[rank * slab_count, (rank + 1) * slab_count) holds the
 *                             sum over ranks of that slab (the rest of `grad` is scratch).  count = nranks * slab_count.
 *   drrt_adam_step_slab       on that slab (above), then
 *   drrt_allgather_grid:      in place; every rank contributes grid[rank * slab_count, ...) and receives the rest. */
int drrt_reduce_scatter_grad(float *grad, size_t slab_count, int rank, void *comm, drrt_stream_t stream);
/* Peer mappings for drrt_adam_step_peers (CUDA IPC).  drrt_ipc_export: handle of the device allocation that holds
 * `ptr` and ptr's offset inside it (the buffer must come from cudaMalloc).  drrt_ipc_open, in another process on the
 * same node: the address `ptr` has there, usable by kernels of the calling thread's current device (peer access is
 * enabled by the driver); an allocation is opened once per process.  drrt_ipc_close_all unmaps everything. */
#define DRRT_IPC_HANDLE_BYTES 64
int drrt_ipc_export(const void *ptr, char handle[DRRT_IPC_HANDLE_BYTES], uint64_t *offset);
int drrt_ipc_open(const char handle[DRRT_IPC_HANDLE_BYTES], uint64_t offset, void **ptr);
int drrt_ipc_close_all(void);
int drrt_allgather_grid(float *grid, size_t slab_count, int rank, void *comm, drrt_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DRRT_B200_H */
