// Host-visible argument blocks and launchers shared by the kernel translation
// units and the C ABI (drrt_capi.cu).
#pragma once
#include <cuda_runtime.h>
#include <limits.h>
#include <stdint.h>

#include "../../include/drrt_b200.h"
#include "drrt_device.cuh"

#define DRRT_LAYOUT_SOA 1
#define DRRT_LAYOUT_AOS 3

namespace drrt {

#ifndef DRRT_TRACE_THREADS
#define DRRT_TRACE_THREADS 128
#endif
#ifndef DRRT_ADJ_THREADS
#define DRRT_ADJ_THREADS 128
#endif
constexpr int kTraceThreads = DRRT_TRACE_THREADS;
constexpr int kAdjThreads = DRRT_ADJ_THREADS;
#ifndef DRRT_ADJ_MIN_BLOCKS
#define DRRT_ADJ_MIN_BLOCKS 7
#endif
#ifndef DRRT_TRACE_MIN_BLOCKS
#define DRRT_TRACE_MIN_BLOCKS 6  // measured (profiles/r06_tune_occupancy.log): 74 registers / 6 CTAs beat 56 / 9
#endif
#ifndef DRRT_TRACE_PAIR_MIN_BLOCKS
#define DRRT_TRACE_PAIR_MIN_BLOCKS 4
#endif
constexpr int kTracePairMinBlocks = DRRT_TRACE_PAIR_MIN_BLOCKS;
constexpr int kAdjMinBlocks = DRRT_ADJ_MIN_BLOCKS;
constexpr int kTraceMinBlocks = DRRT_TRACE_MIN_BLOCKS;

struct TraceArgs {
  GridP grid;
  const float *sdf;
  const float *pos, *vel;
  const float *aux0, *aux1;  // plane origin / normal, or target
  int64_t N;
  int layout;
  float ds;
  int max_steps;
  float *xt, *vt;
  uint8_t *not_escaped;
  float *dist2;
  int64_t *stats;
  // trace_target tail state (SoA scratch)
  float *tail_x, *tail_v;
  int *tail_iters, *tail_max_iters;
  int pair;   // trace only: two rays per thread (trace_pair_kernel)
  int smem_grid;  // small grid + small batch: gather from a shared-memory copy of the grid (one CTA per SM)
  float one;  // 1.0f, opaque to the compiler (PairK)
};

struct AdjArgs {
  GridP grid;
  const float *sdf;
  const float *xt, *vt, *dx, *dv;
  int64_t N;
  int layout;
  float ds;
  int max_steps;
  int agg_policy;  // warp aggregation: 0 adaptive, 1 off, 2/4/8/32 fixed maximum run length
  int agg_t8, agg_twarp;  // adaptive: lanes per cell from which runs of 8 / the whole warp are summed
  int smem_grid;   // small grid + small batch: gather from a shared-memory copy of the grid (one CTA per SM)
  int face_carry;  // keep the accumulators of the face shared with the next cell in registers
  int replicas;    // > 1: CTA b adds into gradient copy b % replicas (copy 0 = grad, the others in `replica_base`)
  float *replica_base;
  int64_t replica_stride;  // floats per copy (= W*H*D)
  const int *interleave_log2;  // device: log2 of the ray interleave, or NULL (identity mapping)
  int64_t *stats;
  // ray gradients: dL/dpos, dL/dvel of the launch state; pos0 nullable
  const float *pos0;
  float *dpos, *dvel;
};

struct CableArgs {
  CableP cable;
  const float *pos, *vel, *target;  // adjoint: xt, vt, dx(=target slot unused)
  const float *dx, *dv;
  int64_t N;
  int layout;
  float ds;
  int max_steps;
  float *xt, *vt, *dist2;
  int64_t *stats;
};

// sensor epilogue, drrt_sensor_kernels.cu
struct SensorArgs {
  const float *x, *v;   // ray state at the volume exit (x unused in far-field mode)
  const float *e;       // per-ray radiance, or NULL -> e0
  float e0;
  int64_t N;
  int layout;
  float p[3], n[3], t1[3], t2[3];  // plane origin, normal, tangents (t1 = n x t2)
  int res;
  float offset;         // span / 2 (near) or sin(angle_span / 2) (far)
  float rcp_h;          // res / span resp. res / (2 offset)
  float *image;         // forward: res x res, accumulated into
  const float *dimage;  // backward: dL/dimage
  float *dx, *dv, *de;  // backward outputs (each nullable)
};
int launch_sensor_splat(bool far, const SensorArgs &a, cudaStream_t s);
int launch_sensor_splat_bwd(bool far, const SensorArgs &a, cudaStream_t s);
// Grid.Get on the plane: texture in a.dimage, per-ray output in a.de (forward); upstream dL/df in a.e (backward)
int launch_sensor_gather(bool far, const SensorArgs &a, cudaStream_t s);
int launch_sensor_gather_bwd(bool far, const SensorArgs &a, cudaStream_t s);

// on-device ray generation, drrt_source_kernels.cu
constexpr int kSourceMaxViews = 32;  // kinds: DRRT_SOURCE_* of include/drrt_b200.h  // views per launch (their rotations travel as kernel parameters)
struct SourceArgs {
  const float *offset;   // device uniforms [total views][2*spp][P0][P1], or NULL -> Philox(seed)
  uint64_t seed;
  int first_view, nviews;  // this launch covers views [first_view, first_view + nviews)
  int P0, P1, spp;
  int64_t rays_per_view;   // spp * P0 * P1
  float width, sensor_dist;
  int kind;                // DRRT_SOURCE_*
  float cone_cos;          // cone source: cos(cone_angle / 2)
  float cone_half;         // point lattice: cone_angle / 2 in radians
  int independent;
  int has_post;
  float post_rot[9], post_span;
  float view_rot[9 * kSourceMaxViews];
  float *x, *v, *planes;   // outputs of this launch: (n,3), (n,3), (n,3,3)
  uint8_t *keep;           // circle mask of this launch, or NULL
  float *tpv, *target;     // area sources: per-ray sensor_dist / v.y and target point (n,3); nullable
};
int launch_plane_source(const SourceArgs &a, cudaStream_t s);

// kind: 0 plain, 1 plane, 2 target, 3 sdf.  Returns the number of kernels launched.
// arith: 0 fast polynomial, 1 reference-literal, 2 exact-trajectory (default)
int launch_trace(int kind, int arith, const TraceArgs &a, cudaStream_t s);
// interleave: 0 off, -1 measured on the device, 2^k forced
int launch_backtrace(bool sdf, int arith, bool warp_agg, int interleave, const AdjArgs &a, float *grad,
                     unsigned long long *fixed, const unsigned *fixed_max_bits, cudaStream_t s);
int launch_trace_cable(bool ref_arith, const CableArgs &a, cudaStream_t s);
int launch_backtrace_cable(bool ref_arith, const CableArgs &a, float *grad,
                           unsigned long long *fixed, const unsigned *fixed_max_bits,
                           cudaStream_t s);

// deterministic mode helpers
int launch_fixed_max(const float *dx, const float *dv, int64_t n3, unsigned *max_bits,
                     cudaStream_t s);
// in-switch sum of slab [first4, first4 + count4) (float4 units) of the multicast-mapped gradient grid
int launch_allreduce_multimem(float *mc, int64_t first4, int64_t count4, cudaStream_t s);
int launch_reduce_replicas(float *grad, const float *replicas, int64_t G, int extra, cudaStream_t s);
int launch_fixed_to_float(const unsigned long long *fixed, const unsigned *max_bits, float *grad,
                          int64_t G, bool accumulate, cudaStream_t s);

// optimiser step / up-res
// the buffers hold planes [i0_begin, i0_begin + i0_count) of the (S0, S1, S2) grid
int launch_adam_step(float *n, const float *grad, float *m, float *v, int S0, int S1, int S2, int i0_begin,
                     int i0_count, int step,
                     float lr, float beta1, float beta2, float eps, float clamp_min, bool zero_boundary,
                     bool clamp, cudaStream_t s);
// sharded step over peer memory: host arrays of `nranks` device pointers to every rank's full grid / gradient
constexpr int kMaxPeers = 16;
int launch_adam_step_peers(float *const *n_replicas, const float *const *grad_replicas, int nranks, int self,
                           float *m, float *v, int S0, int S1, int S2, int i0_begin, int i0_count, int step,
                           float lr, float beta1, float beta2, float eps, float clamp_min, bool zero_boundary,
                           bool clamp, cudaStream_t s);
// the sharded step with the switch doing the gradient sum and the broadcast of the new slab (multicast addresses)
int launch_adam_step_multimem(float *n_mc, const float *g_mc, const float *n_local, float *m, float *v, int S0, int S1,
                              int S2, int i0_begin, int i0_count, int step, float lr, float beta1, float beta2, float eps,
                              float clamp_min, bool zero_boundary, bool clamp, cudaStream_t s);
int launch_upres_volume(const float *src, int s0, int s1, int s2, float *dst, int d0, int d1, int d2,
                        cudaStream_t s);

// volume probes
int launch_volume_eval(const GridP &g, const float *pos, int64_t N, int layout, float *n,
                       float *grad, float *hess, bool ref_arith, cudaStream_t s);
int launch_volume_splat(const GridP &g, float *data, const float *pos, const float *val,
                        const float *grad, int64_t N, int layout, cudaStream_t s);
int launch_cylinder_eval(const CableP &c, const float *pos, int64_t N, int layout, float *n,
                         float *grad, float *hess, cudaStream_t s);
int launch_cylinder_splat(const CableP &c, float *data, const float *pos, const float *val,
                          const float *grad, int64_t N, int layout, cudaStream_t s);

}  // namespace drrt
