// Small helper kernels: deterministic-mode fixed-point plumbing and the
// single-evaluation probes of the volume primitives used by the parity tests.
#include <cstdlib>

#include "drrt_device.cuh"
#include "drrt_launch.h"

namespace drrt {

// max(|dx|, |dv|) over the batch as a float bit pattern (non-negative floats
// order like unsigned integers), consumed by AccumFixed::prepare().
__global__ void fixed_max_kernel(const float *dx, const float *dv, int64_t n3, unsigned *max_bits) {
  unsigned m = 0;
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n3;
       i += (int64_t) gridDim.x * blockDim.x) {
    m = max(m, __float_as_uint(fabsf(dx[i])));
    m = max(m, __float_as_uint(fabsf(dv[i])));
  }
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicMax(max_bits, m);
}

__global__ void fixed_to_float_kernel(const unsigned long long *fixed, const unsigned *max_bits,
                                      float *grad, int64_t G, bool accumulate) {
  const float inv = 1.0f / AccumFixed::scale_from_bits(*max_bits);
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < G;
       i += (int64_t) gridDim.x * blockDim.x) {
    // int64 -> double -> scaled float: a single rounding to float at the end
    float v = (float) ((double) (long long) fixed[i] * (double) inv);
    grad[i] = accumulate ? grad[i] + v : v;
  }
}

int launch_fixed_max(const float *dx, const float *dv, int64_t n3, unsigned *max_bits,
                     cudaStream_t s) {
  if (n3 <= 0) return 0;
  int nb = (int) ((n3 + 1023) / 1024);
  if (nb > 148 * 8) nb = 148 * 8;
  fixed_max_kernel<<<nb, 256, 0, s>>>(dx, dv, n3, max_bits);
  return 1;
}

// grad[i] += sum of the `extra` private copies (the adjoint's per-CTA gradient copies)
__global__ void __launch_bounds__(256) reduce_replicas_kernel(float *__restrict__ grad, const float *__restrict__ rep,
                                                              int64_t G, int extra) {
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t) gridDim.x * blockDim.x) {
    float sum = grad[i];
    for (int r = 0; r < extra; ++r) sum += rep[(int64_t) r * G + i];
    grad[i] = sum;
  }
}

int launch_reduce_replicas(float *grad, const float *replicas, int64_t G, int extra, cudaStream_t s) {
  if (G <= 0 || extra <= 0) return 0;
  int nb = (int) ((G + 255) / 256);
  if (nb > 148 * 8) nb = 148 * 8;
  reduce_replicas_kernel<<<nb, 256, 0, s>>>(grad, replicas, G, extra);
  return 1;
}

// ---------------------------------------------------------------------------
// Gradient sum INSIDE the NVSwitch (the collective of the multi-GPU path as our own kernel).
// `mc` is the MULTICAST address of the gradient grid: one virtual address bound to every rank's copy of the buffer
// (NVLink SHARP; the mapping comes from the caller, e.g. torch's symmetric memory).  Rank r owns slab r of the grid: for
// each float4 of its slab ONE multimem.ld_reduce makes the switch fetch the P copies and return their sum, and ONE
// multimem.st makes it write the sum back into all P copies.  Per GPU the links carry 1/P of the grid in (sums) and 1/P
// out (stores) -- against (P-1)/P each way for a ring or the P2P kernel of drrt_adam_step_peers -- and nothing is staged.
// The caller orders it after every rank's adjoint and before every rank's next read (two barriers on the stream).
// ---------------------------------------------------------------------------
template <int U>
__global__ void __launch_bounds__(512) allreduce_multimem_kernel(float *mc, int64_t first4, int64_t count4) {
  // U independent ld_reduce in flight per thread before the first store: the round trip through the switch is ~2-3 us
  const int64_t stride = (int64_t) gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i0 < count4; i0 += stride * U) {
    float a[U], b[U], c[U], d[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < count4)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(a[u]), "=f"(b[u]), "=f"(c[u]), "=f"(d[u]) : "l"(mc + 4 * (first4 + i)) : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < count4)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     :: "l"(mc + 4 * (first4 + i)), "f"(a[u]), "f"(b[u]), "f"(c[u]), "f"(d[u]) : "memory");
    }
  }
}

int launch_allreduce_multimem(float *mc, int64_t first4, int64_t count4, cudaStream_t s) {
  if (count4 <= 0) return 0;
  // dev knobs (tools/gpu_switch_allreduce.py): DRRT_MM_UNROLL in {1,2,4,8}, DRRT_MM_CTAS per SM
  static const int unroll = [] { const char *e = getenv("DRRT_MM_UNROLL"); return e ? atoi(e) : 4; }();
  static const int ctas = [] { const char *e = getenv("DRRT_MM_CTAS"); return e ? atoi(e) : 4; }();
  int64_t nb = (count4 + 512 * unroll - 1) / (512 * unroll);
  if (nb > 148 * ctas) nb = 148 * ctas;
  if (nb < 1) nb = 1;
  switch (unroll) {
    case 1: allreduce_multimem_kernel<1><<<(unsigned) nb, 512, 0, s>>>(mc, first4, count4); break;
    case 2: allreduce_multimem_kernel<2><<<(unsigned) nb, 512, 0, s>>>(mc, first4, count4); break;
    case 8: allreduce_multimem_kernel<8><<<(unsigned) nb, 512, 0, s>>>(mc, first4, count4); break;
    default: allreduce_multimem_kernel<4><<<(unsigned) nb, 512, 0, s>>>(mc, first4, count4); break;
  }
  return 1;
}

int launch_fixed_to_float(const unsigned long long *fixed, const unsigned *max_bits, float *grad,
                          int64_t G, bool accumulate, cudaStream_t s) {
  if (G <= 0) return 0;
  int nb = (int) ((G + 1023) / 1024);
  if (nb > 148 * 8) nb = 148 * 8;
  fixed_to_float_kernel<<<nb, 256, 0, s>>>(fixed, max_bits, grad, G, accumulate);
  return 1;
}

// ---------------------------------------------------------------------------
// Probes
// ---------------------------------------------------------------------------
template <bool REF>
__global__ void volume_eval_kernel(GridP g, const float *pos, int64_t N, int layout, float *n,
                                   float *grad, float *hess) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  f3 p = ld3(pos, r, N, layout);
  CellIdx c = cell_of(g, p);
  int idx[8];
  float v[8];
  corner_indices(g, c.cx(), c.cy(), c.cz(), idx);
  load_corners(g, idx, v);
  float nn; f3 nx, hm;
  if (REF) {
    eval_grad_ref(g, v, c.wx, c.wy, c.wz, nn, nx);
    hm = eval_hess_ref(g, v, c.wx, c.wy, c.wz);
  } else {
    CellPoly cp = make_poly(v);
    f3 gg, hh;
    eval_poly(cp, c.wx, c.wy, c.wz, nn, gg, hh);
    float r2 = g.rcp_h * g.rcp_h;
    nx.x = gg.x * g.rcp_h; nx.y = gg.y * g.rcp_h; nx.z = gg.z * g.rcp_h;
    hm.x = hh.x * r2; hm.y = hh.y * r2; hm.z = hh.z * r2;
  }
  n[r] = nn;
  st3(grad, r, N, layout, nx);
  st3(hess, r, N, layout, hm);
}

// volume::splat, src/volume.cpp:182-244 (fused form; the order of the float
// atomics is not the CPU's serial order).
__global__ void volume_splat_kernel(GridP g, float *data, const float *pos, const float *val,
                                    const float *grad, int64_t N, int layout) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  f3 p = ld3(pos, r, N, layout), gr = ld3(grad, r, N, layout);
  CellIdx c = cell_of(g, p);
  int idx[8];
  corner_indices(g, c.cx(), c.cy(), c.cz(), idx);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  splat_accumulate(val[r], gr, c.wx, c.wy, c.wz, acc);
#pragma unroll
  for (int k = 0; k < 8; ++k) atomicAdd(data + idx[k], acc[k]);
}

int launch_volume_eval(const GridP &g, const float *pos, int64_t N, int layout, float *n,
                       float *grad, float *hess, bool ref_arith, cudaStream_t s) {
  if (N <= 0) return 0;
  unsigned nb = (unsigned) ((N + 127) / 128);
  if (ref_arith) volume_eval_kernel<true><<<nb, 128, 0, s>>>(g, pos, N, layout, n, grad, hess);
  else volume_eval_kernel<false><<<nb, 128, 0, s>>>(g, pos, N, layout, n, grad, hess);
  return 1;
}

int launch_volume_splat(const GridP &g, float *data, const float *pos, const float *val,
                        const float *grad, int64_t N, int layout, cudaStream_t s) {
  if (N <= 0) return 0;
  volume_splat_kernel<<<(unsigned) ((N + 127) / 128), 128, 0, s>>>(g, data, pos, val, grad, N, layout);
  return 1;
}

}  // namespace drrt
