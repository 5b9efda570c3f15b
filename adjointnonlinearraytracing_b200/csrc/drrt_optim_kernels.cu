// the optimiser step that follows every adjoint call in the
// reference's coarse-to-fine loop (reference core/optimizer.py:55-68) and the
// trilinear up-res between levels (core/optimizer.py:7-10 -> core/grid.py:318-330).
//
// adam_step_kernel fuses, in one pass over the grid,
//     n.grad[boundary] = 0            core/optimizer.py:61-62 (mask = all but [1:-1,1:-1,1:-1])
//     torch.optim.Adam step           core/optimizer.py:64   (no weight decay, no amsgrad)
//     n.clamp_(min = 1)               core/optimizer.py:67
// which the reference runs as ~10 separate elementwise kernels.  HBM-bound:
// 4 reads + 3 writes of 4 B per voxel = 28 B per voxel.
#include "drrt_device.cuh"
#include "drrt_launch.h"

namespace drrt {

struct AdamP {
  float lr_over_bc1;    // lr / (1 - beta1^t)
  float rsqrt_bc2;      // 1 / sqrt(1 - beta2^t)
  float beta1, beta2, eps, clamp_min;
  int zero_boundary, clamp;
};

__device__ __forceinline__ void adam_one(float &n, float g, float &m, float &v, const AdamP &p) {
  // torch._single_tensor_adam: exp_avg.lerp_(grad, 1 - beta1);
  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value = 1 - beta2);
  // denom = exp_avg_sq.sqrt() / sqrt(bias_correction2) + eps; param.addcdiv_(exp_avg, denom, -step_size)
  m = m + (1.0f - p.beta1) * (g - m);
  v = v * p.beta2 + (1.0f - p.beta2) * g * g;
  const float denom = sqrtf(v) * p.rsqrt_bc2 + p.eps;
  n = n - p.lr_over_bc1 * (m / denom);
  if (p.clamp) n = fmaxf(n, p.clamp_min);
}

// One row (fixed i0, i1) per `bpr` consecutive blocks, VEC consecutive i2 per thread.
template <int VEC>
__global__ void __launch_bounds__(128) adam_step_kernel(float *__restrict__ n, const float *__restrict__ grad,
                                                        float *__restrict__ m, float *__restrict__ v, int S0,
                                                        int S1, int S2, int i0_begin, int bpr, AdamP p) {
  // the buffers hold planes [i0_begin, i0_begin + count) of an (S0, S1, S2) grid: the whole grid, or one
  // rank's slab of the sharded optimiser step (boundary planes are those of the WHOLE grid)
  const int row = blockIdx.x / bpr;  // (i0 - i0_begin) * S1 + i1
  const int i0 = row / S1 + i0_begin, i1 = row - (row / S1) * S1;
  const int i2 = ((blockIdx.x - row * bpr) * blockDim.x + threadIdx.x) * VEC;
  if (i2 >= S2) return;
  const bool row_boundary = p.zero_boundary && (i0 == 0 || i0 == S0 - 1 || i1 == 0 || i1 == S1 - 1);
  const size_t base = (size_t) row * S2 + i2;
  if (VEC == 4) {
    float4 nn = *reinterpret_cast<float4 *>(n + base), gg = *reinterpret_cast<const float4 *>(grad + base);
    float4 mm = *reinterpret_cast<float4 *>(m + base), vv = *reinterpret_cast<float4 *>(v + base);
    if (row_boundary) gg = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.zero_boundary && i2 == 0) gg.x = 0.f;
    if (p.zero_boundary && i2 + 3 == S2 - 1) gg.w = 0.f;
    adam_one(nn.x, gg.x, mm.x, vv.x, p); adam_one(nn.y, gg.y, mm.y, vv.y, p);
    adam_one(nn.z, gg.z, mm.z, vv.z, p); adam_one(nn.w, gg.w, mm.w, vv.w, p);
    *reinterpret_cast<float4 *>(n + base) = nn;
    *reinterpret_cast<float4 *>(m + base) = mm;
    *reinterpret_cast<float4 *>(v + base) = vv;
  } else {
    float nn = n[base], gg = grad[base], mm = m[base], vv = v[base];
    if (row_boundary || (p.zero_boundary && (i2 == 0 || i2 == S2 - 1))) gg = 0.f;
    adam_one(nn, gg, mm, vv, p);
    n[base] = nn; m[base] = mm; v[base] = vv;
  }
}

int launch_adam_step(float *n, const float *grad, float *m, float *v, int S0, int S1, int S2, int i0_begin,
                     int i0_count, int step,
                     float lr, float beta1, float beta2, float eps, float clamp_min, bool zero_boundary,
                     bool clamp, cudaStream_t s) {
  if ((int64_t) i0_count * S1 * S2 <= 0) return 0;
  AdamP p;
  const double bc1 = 1.0 - pow((double) beta1, (double) step), bc2 = 1.0 - pow((double) beta2, (double) step);
  p.lr_over_bc1 = (float) ((double) lr / bc1);
  p.rsqrt_bc2 = (float) (1.0 / sqrt(bc2));
  p.beta1 = beta1; p.beta2 = beta2; p.eps = eps; p.clamp_min = clamp_min;
  p.zero_boundary = zero_boundary; p.clamp = clamp;
  const bool vec = (S2 % 4 == 0) && ((((uintptr_t) n | (uintptr_t) grad | (uintptr_t) m | (uintptr_t) v) & 15) == 0);
  const int bpr = (S2 / (vec ? 4 : 1) + 127) / 128;
  const int64_t nblocks = (int64_t) i0_count * S1 * bpr;
  if (nblocks > 0x7fffffffLL) return -1;
  if (vec) adam_step_kernel<4><<<(unsigned) nblocks, 128, 0, s>>>(n, grad, m, v, S0, S1, S2, i0_begin, bpr, p);
  else adam_step_kernel<1><<<(unsigned) nblocks, 128, 0, s>>>(n, grad, m, v, S0, S1, S2, i0_begin, bpr, p);
  return 1;
}

// ---------------------------------------------------------------------------
// Sharded optimiser step over peer memory: ONE kernel in place of
// reduce-scatter + slab Adam + all-gather.  Every rank owns the planes [i0_begin, i0_begin + count)
// of the grid; for each voxel of its slab it loads the P private gradients straight from the
// peers' grids over NVLink (P2P loads, summed in rank order, so the result does not depend on
// which rank owns the slab), applies boundary-zero + Adam + clamp with its slab of the Adam
// state, and stores the new value into all P replicas of the grid (P2P stores).  The caller
// brackets the launch with two stream-ordered barriers (dist.PeerShardedAdam).
// ---------------------------------------------------------------------------
struct PeerPtrs {
  float *n[kMaxPeers];
  const float *g[kMaxPeers];
  int P;
};

template <int VEC>
__global__ void __launch_bounds__(128) adam_step_peers_kernel(PeerPtrs pp, float *__restrict__ m, float *__restrict__ v,
                                                              int S0, int S1, int S2, int i0_begin, int bpr, AdamP p) {
  const int row = blockIdx.x / bpr;  // (i0 - i0_begin) * S1 + i1
  const int i0 = row / S1 + i0_begin, i1 = row - (row / S1) * S1;
  const int i2 = ((blockIdx.x - row * bpr) * blockDim.x + threadIdx.x) * VEC;
  if (i2 >= S2) return;
  const bool row_boundary = p.zero_boundary && (i0 == 0 || i0 == S0 - 1 || i1 == 0 || i1 == S1 - 1);
  const size_t local = (size_t) row * S2 + i2;                    // within the slab (Adam state)
  const size_t base = ((size_t) i0 * S1 + i1) * S2 + i2;          // within the whole grid (replicas)
  if (VEC == 4) {
    // all peer loads are issued before the first add (NVLink latency is ~2 us: keep them in flight together)
    float4 t[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
      if (r < pp.P) t[r] = __ldcs(reinterpret_cast<const float4 *>(pp.g[r] + base));
    float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
      if (r < pp.P) { gg.x += t[r].x; gg.y += t[r].y; gg.z += t[r].z; gg.w += t[r].w; }
    float4 nn = *reinterpret_cast<const float4 *>(pp.n[0] + base);  // replicas are identical: any copy will do
    float4 mm = *reinterpret_cast<float4 *>(m + local), vv = *reinterpret_cast<float4 *>(v + local);
    if (row_boundary) gg = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.zero_boundary && i2 == 0) gg.x = 0.f;
    if (p.zero_boundary && i2 + 3 == S2 - 1) gg.w = 0.f;
    adam_one(nn.x, gg.x, mm.x, vv.x, p); adam_one(nn.y, gg.y, mm.y, vv.y, p);
    adam_one(nn.z, gg.z, mm.z, vv.z, p); adam_one(nn.w, gg.w, mm.w, vv.w, p);
    for (int r = 0; r < pp.P; ++r) *reinterpret_cast<float4 *>(pp.n[r] + base) = nn;
    *reinterpret_cast<float4 *>(m + local) = mm;
    *reinterpret_cast<float4 *>(v + local) = vv;
  } else {
    float gg = 0.f;
    for (int r = 0; r < pp.P; ++r) gg += pp.g[r][base];
    float nn = pp.n[0][base], mm = m[local], vv = v[local];
    if (row_boundary || (p.zero_boundary && (i2 == 0 || i2 == S2 - 1))) gg = 0.f;
    adam_one(nn, gg, mm, vv, p);
    for (int r = 0; r < pp.P; ++r) pp.n[r][base] = nn;
    m[local] = mm; v[local] = vv;
  }
}

int launch_adam_step_peers(float *const *n_replicas, const float *const *grad_replicas, int nranks, int self,
                           float *m, float *v, int S0, int S1, int S2, int i0_begin, int i0_count, int step,
                           float lr, float beta1, float beta2, float eps, float clamp_min, bool zero_boundary,
                           bool clamp, cudaStream_t s) {
  if ((int64_t) i0_count * S1 * S2 <= 0) return 0;
  if (nranks < 1 || nranks > kMaxPeers || self < 0 || self >= nranks) return -2;
  PeerPtrs pp;
  pp.P = nranks;
  uintptr_t bits = (uintptr_t) m | (uintptr_t) v;
  for (int r = 0; r < nranks; ++r) {
    pp.n[r] = n_replicas[r]; pp.g[r] = grad_replicas[r];
    if (!pp.n[r] || !pp.g[r]) return -2;
    bits |= (uintptr_t) pp.n[r] | (uintptr_t) pp.g[r];
  }
  // read this rank's own copy of n (local HBM) rather than a peer's
  float *t = pp.n[0]; pp.n[0] = pp.n[self]; pp.n[self] = t;
  AdamP p;
  const double bc1 = 1.0 - pow((double) beta1, (double) step), bc2 = 1.0 - pow((double) beta2, (double) step);
  p.lr_over_bc1 = (float) ((double) lr / bc1);
  p.rsqrt_bc2 = (float) (1.0 / sqrt(bc2));
  p.beta1 = beta1; p.beta2 = beta2; p.eps = eps; p.clamp_min = clamp_min;
  p.zero_boundary = zero_boundary; p.clamp = clamp;
  const bool vec = (S2 % 4 == 0) && ((bits & 15) == 0);
  const int bpr = (S2 / (vec ? 4 : 1) + 127) / 128;
  const int64_t nblocks = (int64_t) i0_count * S1 * bpr;
  if (nblocks > 0x7fffffffLL) return -1;
  if (vec) adam_step_peers_kernel<4><<<(unsigned) nblocks, 128, 0, s>>>(pp, m, v, S0, S1, S2, i0_begin, bpr, p);
  else adam_step_peers_kernel<1><<<(unsigned) nblocks, 128, 0, s>>>(pp, m, v, S0, S1, S2, i0_begin, bpr, p);
  return 1;
}

// ---------------------------------------------------------------------------
// The same sharded step with the SWITCH doing the sum and the broadcast (NVLink SHARP): `g_mc` and `n_mc` are the
// multicast addresses of the gradient grid and of the IOR grid (one virtual address bound to every rank's copy).  For each
// float4 of its slab a rank issues ONE multimem.ld_reduce (the switch fetches the P private gradients and returns their
// sum), applies boundary-zero + Adam + clamp with the local slab of the Adam state and its local copy of n, and issues ONE
// multimem.st that lands the new value in all P replicas.  Per GPU the links carry the grid once out (its copy of the
// gradient, pulled by the switch) and once in (the updated slabs) instead of 2 x (P-1)/P x in each direction for the
// P2P kernel above or a ring.  Rows of S2 % 4 == 0 floats only (the caller falls back to the P2P kernel otherwise).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) adam_step_multimem_kernel(float *n_mc, const float *g_mc, const float *__restrict__ n_local,
                                                                 float *__restrict__ m, float *__restrict__ v, int S0, int S1,
                                                                 int S2, int i0_begin, int bpr, AdamP p) {
  const int row = blockIdx.x / bpr;  // (i0 - i0_begin) * S1 + i1
  const int i0 = row / S1 + i0_begin, i1 = row - (row / S1) * S1;
  const int i2 = ((blockIdx.x - row * bpr) * blockDim.x + threadIdx.x) * 4;
  if (i2 >= S2) return;
  const bool row_boundary = p.zero_boundary && (i0 == 0 || i0 == S0 - 1 || i1 == 0 || i1 == S1 - 1);
  const size_t local = (size_t) row * S2 + i2;
  const size_t base = ((size_t) i0 * S1 + i1) * S2 + i2;
  float4 gg;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(gg.x), "=f"(gg.y), "=f"(gg.z), "=f"(gg.w) : "l"(g_mc + base) : "memory");
  float4 nn = *reinterpret_cast<const float4 *>(n_local + base);
  float4 mm = *reinterpret_cast<float4 *>(m + local), vv = *reinterpret_cast<float4 *>(v + local);
  if (row_boundary) gg = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.zero_boundary && i2 == 0) gg.x = 0.f;
  if (p.zero_boundary && i2 + 3 == S2 - 1) gg.w = 0.f;
  adam_one(nn.x, gg.x, mm.x, vv.x, p); adam_one(nn.y, gg.y, mm.y, vv.y, p);
  adam_one(nn.z, gg.z, mm.z, vv.z, p); adam_one(nn.w, gg.w, mm.w, vv.w, p);
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               :: "l"(n_mc + base), "f"(nn.x), "f"(nn.y), "f"(nn.z), "f"(nn.w) : "memory");
  *reinterpret_cast<float4 *>(m + local) = mm;
  *reinterpret_cast<float4 *>(v + local) = vv;
}

int launch_adam_step_multimem(float *n_mc, const float *g_mc, const float *n_local, float *m, float *v, int S0, int S1,
                              int S2, int i0_begin, int i0_count, int step, float lr, float beta1, float beta2, float eps,
                              float clamp_min, bool zero_boundary, bool clamp, cudaStream_t s) {
  if ((int64_t) i0_count * S1 * S2 <= 0) return 0;
  if (S2 % 4 || (((uintptr_t) n_mc | (uintptr_t) g_mc | (uintptr_t) n_local | (uintptr_t) m | (uintptr_t) v) & 15)) return -2;
  AdamP p;
  const double bc1 = 1.0 - pow((double) beta1, (double) step), bc2 = 1.0 - pow((double) beta2, (double) step);
  p.lr_over_bc1 = (float) ((double) lr / bc1);
  p.rsqrt_bc2 = (float) (1.0 / sqrt(bc2));
  p.beta1 = beta1; p.beta2 = beta2; p.eps = eps; p.clamp_min = clamp_min;
  p.zero_boundary = zero_boundary; p.clamp = clamp;
  const int bpr = (S2 / 4 + 127) / 128;
  const int64_t nblocks = (int64_t) i0_count * S1 * bpr;
  if (nblocks > 0x7fffffffLL) return -1;
  adam_step_multimem_kernel<<<(unsigned) nblocks, 128, 0, s>>>(n_mc, g_mc, n_local, m, v, S0, S1, S2, i0_begin, bpr, p);
  return 1;
}

// ---------------------------------------------------------------------------
// upres_volume (core/grid.py:318-330 -> Grid.GetLinear :227-273), evaluated like the
// reference in float64 (core/optimizer.py:10 converts to double first):
//   x_k = linspace(0, 1, d_k)[i_k];  norm = x / h, h = 1 / max(s_0 - 1, 1);
//   x0 = floor(norm); w0 = clip(norm - x0, 0, 1); indices clipped to [0, s_0 - 1]
// (the reference clips every axis with res[0]: cubic volumes).
// ---------------------------------------------------------------------------
__device__ __forceinline__ double linspace01(int i, int steps) {
  // at::linspace: step = (end - start) / (steps - 1); first half start + i*step, second half end - (steps-1-i)*step
  if (steps <= 1) return 0.0;
  const double step = 1.0 / (double) (steps - 1);
  return i < steps / 2 ? (double) i * step : 1.0 - (double) (steps - 1 - i) * step;
}

template <typename T>
__global__ void __launch_bounds__(128) upres_volume_kernel(const T *__restrict__ src, int s0, int s1, int s2,
                                                           T *__restrict__ dst, int d0, int d1, int d2) {
  const int64_t G = (int64_t) d0 * d1 * d2;
  const int nvox = s0 - 1 > 1 ? s0 - 1 : 1;
  const double h = 1.0 / (double) nvox;
  for (int64_t o = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; o < G; o += (int64_t) gridDim.x * blockDim.x) {
    const int i2 = (int) (o % d2), i1 = (int) ((o / d2) % d1), i0 = (int) (o / ((int64_t) d2 * d1));
    const double nx[3] = {linspace01(i0, d0) / h, linspace01(i1, d1) / h, linspace01(i2, d2) / h};
    int lo[3], hi[3];
    double w[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double f = floor(nx[k]);
      double ww = nx[k] - f;
      w[k] = ww < 0.0 ? 0.0 : (ww > 1.0 ? 1.0 : ww);
      const long long x0 = (long long) f;
      lo[k] = (int) (x0 < 0 ? 0 : (x0 > s0 - 1 ? s0 - 1 : x0));
      hi[k] = (int) (x0 + 1 < 0 ? 0 : (x0 + 1 > s0 - 1 ? s0 - 1 : x0 + 1));
    }
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {  // corner order of torch.meshgrid(arange(2) x 3): first axis slowest
      const int b0 = (c >> 2) & 1, b1 = (c >> 1) & 1, b2 = c & 1;
      const int j0 = b0 ? hi[0] : lo[0], j1 = b1 ? hi[1] : lo[1], j2 = b2 ? hi[2] : lo[2];
      const double ww = (b0 ? w[0] : 1.0 - w[0]) * (b1 ? w[1] : 1.0 - w[1]) * (b2 ? w[2] : 1.0 - w[2]);
      acc += (double) src[((int64_t) j0 * s1 + j1) * s2 + j2] * ww;
    }
    dst[o] = (T) acc;
  }
}

int launch_upres_volume(const float *src, int s0, int s1, int s2, float *dst, int d0, int d1, int d2,
                        cudaStream_t s) {
  const int64_t G = (int64_t) d0 * d1 * d2;
  if (G <= 0) return 0;
  int nb = (int) ((G + 127) / 128);
  if (nb > 148 * 16) nb = 148 * 16;
  upres_volume_kernel<float><<<nb, 128, 0, s>>>(src, s0, s1, s2, dst, d0, d1, d2);
  return 1;
}

}  // namespace drrt
