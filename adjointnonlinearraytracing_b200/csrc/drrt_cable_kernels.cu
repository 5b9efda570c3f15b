// Fibre ("cable") variant: radially symmetric 1-D IOR profile n(r) around the y
// axis (reference src/cylinder_volume.cpp), forward trace_cable
// (src/tracer.cpp:312-382) and adjoint backtrace_cable (:511-567).
//
// Arithmetic follows the reference's operation order with strictly rounded
// operations; the profile has <= a few hundred samples, so the cost of a step is
// the ALU work, and every ray of the batch splats into the same few addresses.
// The adjoint therefore accumulates into a per-CTA shared-memory copy of the
// profile and adds it to global memory once per CTA.
#include "drrt_device.cuh"
#include "drrt_launch.h"

namespace drrt {

constexpr float kCylEps = 1e-6f;  // FLOAT_EPSILON, src/cylinder_volume.cpp:15
constexpr int kCableSmemMax = 8192;

struct CylCell {
  int idx0, idx1;
  float w0, w1, r;
  f3 xs;
};

// src/cylinder_volume.cpp:37-48 (w0 is measured from the CLAMPED idx0)
__device__ __forceinline__ CylCell cyl_cell(const CableP &c, f3 p) {
  CylCell q;
  q.xs.x = sub_(p.x, c.radius); q.xs.y = 0.f; q.xs.z = sub_(p.z, c.radius);
  float d = fma_(q.xs.z, q.xs.z, fma_(q.xs.y, q.xs.y, mul_(q.xs.x, q.xs.x)));
  q.r = __fsqrt_rn(d);
  float rm = __fdiv_rn(q.r, c.h);
  int fl = (int) floorf(rm);
  q.idx0 = max(min(fl, c.res - 1), 0);
  q.idx1 = max(min(q.idx0 + 1, c.res - 1), 0);
  q.w0 = sub_(rm, (float) q.idx0);
  q.w1 = sub_(1.0f, q.w0);
  return q;
}

// normalize(xs) = xs * rsqrt(squared_norm(xs))
__device__ __forceinline__ f3 cyl_xhat(const CylCell &q) {
  float d = fma_(q.xs.z, q.xs.z, fma_(q.xs.y, q.xs.y, mul_(q.xs.x, q.xs.x)));
  float rs = __frcp_rn(__fsqrt_rn(d));
  f3 h = {mul_(q.xs.x, rs), mul_(q.xs.y, rs), mul_(q.xs.z, rs)};
  return h;
}

// cylinder_volume::eval_grad (:26-59) and eval_hess (:61-111); gathers unmasked.
// Hm = (H00, H02, H20, H22); every other entry of the reference's matrix is 0.
__device__ __forceinline__ void cyl_eval(const CableP &c, const CylCell &q, float &n, f3 &nx, float Hm[4]) {
  float val0 = __ldg(c.data + q.idx0), val1 = __ldg(c.data + q.idx1);
  n = add_(mul_(val0, q.w1), mul_(val1, q.w0));
  float rx = __fdiv_rn(sub_(val1, val0), c.h);
  f3 xh = cyl_xhat(q);
  bool small = q.r < kCylEps;
  nx.x = small ? 0.f : mul_(rx, xh.x);
  nx.y = small ? 0.f : mul_(rx, xh.y);
  nx.z = small ? 0.f : mul_(rx, xh.z);
  if (Hm) {
    if (small) { xh.x = xh.y = xh.z = 0.f; }
    float s = __fdiv_rn(rx, q.r);
    Hm[0] = small ? 0.f : mul_(sub_(1.0f, mul_(xh.x, xh.x)), s);
    Hm[1] = small ? 0.f : mul_(-mul_(xh.x, xh.z), s);
    Hm[2] = small ? 0.f : mul_(-mul_(xh.z, xh.x), s);
    Hm[3] = small ? 0.f : mul_(sub_(1.0f, mul_(xh.z, xh.z)), s);
  }
}

// cylinder_volume::inbounds, :150-156
__device__ __forceinline__ bool cyl_inbounds(const CableP &c, f3 p) {
  float px = sub_(p.x, c.radius), pz = sub_(p.z, c.radius);
  float r = add_(mul_(px, px), mul_(pz, pz));
  return (r < c.r2) & (p.y < c.length) & (p.y >= 0.f);
}

// cylinder_volume::escaped, :158-170
__device__ __forceinline__ bool cyl_escaped(const CableP &c, f3 p, f3 v) {
  float px = sub_(p.x, c.radius), pz = sub_(p.z, c.radius);
  bool esc_length = ((p.y < 0.f) & (v.y < 0.f)) | ((p.y > c.length) & (v.y > 0.f));
  bool out_radius = add_(mul_(px, px), mul_(pz, pz)) >= c.r2;
  bool esc_radius = add_(mul_(px, v.x), mul_(pz, v.z)) > 0.f;
  return (out_radius & esc_radius) | esc_length;
}

__device__ __forceinline__ float cdist2(f3 x, f3 t) {
  float dx = sub_(x.x, t.x), dy = sub_(x.y, t.y), dz = sub_(x.z, t.z);
  return fma_(dz, dz, fma_(dy, dy, mul_(dx, dx)));
}

// Tracer::trace_cable, src/tracer.cpp:312-382.  Updates are masked by `active`
// (:353-354), so an escaped ray is frozen and per-ray exit is exact.
__global__ void __launch_bounds__(kTraceThreads) trace_cable_kernel(CableArgs a) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  int64_t steps = 0;
  int not_exited = 0;
  if (r < a.N) {
    const CableP c = a.cable;
    const float ds = a.ds;
    f3 x = ld3(a.pos, r, a.N, a.layout), v = ld3(a.vel, r, a.N, a.layout);
    f3 t = ld3(a.target, r, a.N, a.layout);
    f3 xt = x, vt = v;
    float best = cdist2(x, t);    // :340
    bool inside = cyl_inbounds(c, x);  // :344
    bool esc = false;
    for (int i = 0; i < a.max_steps; ++i) {
      CylCell q = cyl_cell(c, x);
      float n; f3 nx;
      cyl_eval(c, q, n, nx, nullptr);  // :351
      float dsn = mul_(ds, n);
      v.x = fma_(dsn, nx.x, v.x); v.y = fma_(dsn, nx.y, v.y); v.z = fma_(dsn, nx.z, v.z);
      x.x = fma_(ds, v.x, x.x); x.y = fma_(ds, v.y, x.y); x.z = fma_(ds, v.z, x.z);
      ++steps;
      float cur = cdist2(x, t);  // :356
      if (cur < best) { xt = x; vt = v; best = cur; }
      bool cur_inside = cyl_inbounds(c, x);
      esc = (inside & !cur_inside) | cyl_escaped(c, x, v);  // :360-362
      inside = cur_inside;
      if (esc) break;
    }
    if (!esc) not_exited = 1;
    st3(a.xt, r, a.N, a.layout, xt);
    st3(a.vt, r, a.N, a.layout, vt);
    a.dist2[r] = best;
  }
  if (a.stats) {
    steps = warp_sum_i64(steps);
    int ne = __reduce_add_sync(0xffffffffu, not_exited);
    if ((threadIdx.x & 31) == 0) {
      if (steps) atomicAdd((unsigned long long *) a.stats, (unsigned long long) steps);
      if (ne) atomicAdd((unsigned long long *) a.stats + 1, (unsigned long long) ne);
    }
  }
}

// Tracer::backtrace_cable, src/tracer.cpp:511-567, with cylinder_volume::splat
// (src/cylinder_volume.cpp:113-148) accumulated in shared memory per CTA.
//
// REF = false (default): the four splat terms of a step are fused into two (one per sample) and
// kept in registers while the ray stays between the same two radial samples -- with
// ds = R / res / 2 along the fibre axis a ray changes its radial bin rarely, and every ray of a CTA
// hits the same few shared-memory words, so the atomics were the bottleneck (32-way same-address
// conflicts).  REF = true: the reference's four adds per step, in its order.
template <class Accum, bool REF>
__global__ void __launch_bounds__(kAdjThreads) backtrace_cable_kernel(CableArgs a, Accum out_in) {
  typedef typename Accum::T T;
  extern __shared__ unsigned long long smem_raw[];
  T *sm = reinterpret_cast<T *>(smem_raw);
  Accum out = out_in;
  out.prepare();
  const CableP c = a.cable;
  const bool use_sm = c.res <= kCableSmemMax;
  if (use_sm) {
    for (int k = threadIdx.x; k < c.res; k += blockDim.x) sm[k] = (T) 0;
    __syncthreads();
  }
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  int64_t steps = 0;
  if (r < a.N) {
    const float ds = a.ds;
    f3 x = ld3(a.pos, r, a.N, a.layout), v = ld3(a.vel, r, a.N, a.layout);
    f3 la = ld3(a.dx, r, a.N, a.layout), mu = ld3(a.dv, r, a.N, a.layout);
    mu.x = add_(mu.x, mul_(ds, la.x)); mu.y = add_(mu.y, mul_(ds, la.y)); mu.z = add_(mu.z, mul_(ds, la.z));  // :537
    f3 nv = {-v.x, -v.y, -v.z};
    bool active = !cyl_escaped(c, x, nv);  // :540-541
    int cur0 = -1, cur1 = -1;   // fused mode: the radial samples the register accumulators belong to
    float acc0 = 0.f, acc1 = 0.f;
    auto flush = [&]() {
      if (cur0 < 0) return;
      if (use_sm) { atomicAdd(sm + cur0, out.q(acc0)); atomicAdd(sm + cur1, out.q(acc1)); }
      else { out.add(cur0, acc0); out.add(cur1, acc1); }
    };
    for (int i = 0; active && i < a.max_steps; ++i) {
      x.x = fma_(-ds, v.x, x.x); x.y = fma_(-ds, v.y, x.y); x.z = fma_(-ds, v.z, x.z);  // :547
      CylCell q = cyl_cell(c, x);
      float n, Hm[4]; f3 nx;
      cyl_eval(c, q, n, nx, Hm);  // :548-549
      float mdsn = mul_(-ds, n);
      v.x = fma_(mdsn, nx.x, v.x); v.y = fma_(mdsn, nx.y, v.y); v.z = fma_(mdsn, nx.z, v.z);  // :550
      ++steps;
      nv.x = -v.x; nv.y = -v.y; nv.z = -v.z;
      active = !cyl_escaped(c, x, nv);  // :552
      if (!active) break;
      float dn = fma_(mu.z, nx.z, fma_(mu.y, nx.y, mul_(mu.x, nx.x)));  // :557
      f3 dnx = {mul_(mul_(n, mu.x), ds), mul_(mul_(n, mu.y), ds), mul_(mul_(n, mu.z), ds)};
      // splat: src/cylinder_volume.cpp:139-147
      float val = mul_(dn, ds);
      f3 xh = cyl_xhat(q);
      float gv = fma_(dnx.z, xh.z, fma_(dnx.y, xh.y, mul_(dnx.x, xh.x)));
      if (q.r < kCylEps) gv = 0.f;
      float t0 = mul_(val, q.w1), t1 = mul_(val, q.w0);
      float t2 = __fdiv_rn(-gv, c.h), t3 = __fdiv_rn(gv, c.h);
      if (!REF) {
        if (q.idx0 != cur0 || q.idx1 != cur1) {
          flush();
          cur0 = q.idx0; cur1 = q.idx1; acc0 = acc1 = 0.f;
        }
        acc0 += t0 + t2; acc1 += t1 + t3;
      } else if (use_sm) {
        atomicAdd(sm + q.idx0, out.q(t0)); atomicAdd(sm + q.idx1, out.q(t1));
        atomicAdd(sm + q.idx0, out.q(t2)); atomicAdd(sm + q.idx1, out.q(t3));
      } else {
        out.add(q.idx0, t0); out.add(q.idx1, t1); out.add(q.idx0, t2); out.add(q.idx1, t3);
      }
      // :561-562; (n*H)*mu with the reference's column-sum order (matrix.h:170-178)
      float h0 = fma_(mul_(n, Hm[1]), mu.z, fma_(mul_(n, 0.f), mu.y, mul_(mul_(n, Hm[0]), mu.x)));
      float h2 = fma_(mul_(n, Hm[3]), mu.z, fma_(mul_(n, 0.f), mu.y, mul_(mul_(n, Hm[2]), mu.x)));
      float h1 = 0.f;
      la.x = add_(la.x, mul_(ds, add_(mul_(dn, nx.x), h0)));
      la.y = add_(la.y, mul_(ds, add_(mul_(dn, nx.y), h1)));
      la.z = add_(la.z, mul_(ds, add_(mul_(dn, nx.z), h2)));
      mu.x = add_(mu.x, mul_(ds, la.x)); mu.y = add_(mu.y, mul_(ds, la.y)); mu.z = add_(mu.z, mul_(ds, la.z));
    }
    if (!REF) flush();
  }
  if (use_sm) {
    __syncthreads();
    for (int k = threadIdx.x; k < c.res; k += blockDim.x) {
      T s = sm[k];
      if (s != (T) 0) atomicAdd(out.dst + k, s);
    }
  }
  if (a.stats) {
    steps = warp_sum_i64(steps);
    if ((threadIdx.x & 31) == 0 && steps) atomicAdd((unsigned long long *) a.stats, (unsigned long long) steps);
  }
}

static inline unsigned blocks_for(int64_t n, int threads) { return (unsigned) ((n + threads - 1) / threads); }

int launch_trace_cable(bool, const CableArgs &a, cudaStream_t s) {
  if (a.N <= 0) return 0;
  trace_cable_kernel<<<blocks_for(a.N, kTraceThreads), kTraceThreads, 0, s>>>(a);
  return 1;
}

int launch_backtrace_cable(bool ref_arith, const CableArgs &a, float *grad, unsigned long long *fixed,
                           const unsigned *fixed_max_bits, cudaStream_t s) {
  if (a.N <= 0) return 0;
  unsigned nb = blocks_for(a.N, kAdjThreads);
  int n_sm = a.cable.res <= kCableSmemMax ? a.cable.res : 0;
  if (fixed) {
    AccumFixed acc{fixed, fixed_max_bits, 1.0f};
    // the int64 copy of a profile with more than 6144 samples exceeds the 48 KiB default of dynamic shared memory
    static const bool optin = [] {
      cudaFuncSetAttribute(backtrace_cable_kernel<AccumFixed, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int) (kCableSmemMax * sizeof(unsigned long long)));
      cudaFuncSetAttribute(backtrace_cable_kernel<AccumFixed, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int) (kCableSmemMax * sizeof(unsigned long long)));
      return true;
    }();
    (void) optin;
    if (ref_arith) backtrace_cable_kernel<AccumFixed, true><<<nb, kAdjThreads, n_sm * sizeof(unsigned long long), s>>>(a, acc);
    else backtrace_cable_kernel<AccumFixed, false><<<nb, kAdjThreads, n_sm * sizeof(unsigned long long), s>>>(a, acc);
  } else {
    AccumFloat acc{grad};
    if (ref_arith) backtrace_cable_kernel<AccumFloat, true><<<nb, kAdjThreads, n_sm * sizeof(float), s>>>(a, acc);
    else backtrace_cable_kernel<AccumFloat, false><<<nb, kAdjThreads, n_sm * sizeof(float), s>>>(a, acc);
  }
  return 1;
}

// Probes of the cylinder primitives (reference src/test.cpp:82-101).
__global__ void cylinder_eval_kernel(CableP c, const float *pos, int64_t N, int layout, float *n,
                                     float *grad, float *hess) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  f3 p = ld3(pos, r, N, layout);
  CylCell q = cyl_cell(c, p);
  float nn, Hm[4]; f3 nx;
  cyl_eval(c, q, nn, nx, Hm);
  n[r] = nn;
  st3(grad, r, N, layout, nx);
  float *H = hess + 9 * r;
  H[0] = Hm[0]; H[1] = 0.f; H[2] = Hm[1];
  H[3] = 0.f;   H[4] = 0.f; H[5] = 0.f;
  H[6] = Hm[2]; H[7] = 0.f; H[8] = Hm[3];
}

__global__ void cylinder_splat_kernel(CableP c, float *data, const float *pos, const float *val,
                                      const float *grad, int64_t N, int layout) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  f3 p = ld3(pos, r, N, layout), gr = ld3(grad, r, N, layout);
  CylCell q = cyl_cell(c, p);
  f3 xh = cyl_xhat(q);
  float gv = fma_(gr.z, xh.z, fma_(gr.y, xh.y, mul_(gr.x, xh.x)));
  if (q.r < kCylEps) gv = 0.f;
  atomicAdd(data + q.idx0, mul_(val[r], q.w1));
  atomicAdd(data + q.idx1, mul_(val[r], q.w0));
  atomicAdd(data + q.idx0, __fdiv_rn(-gv, c.h));
  atomicAdd(data + q.idx1, __fdiv_rn(gv, c.h));
}

int launch_cylinder_eval(const CableP &c, const float *pos, int64_t N, int layout, float *n,
                         float *grad, float *hess, cudaStream_t s) {
  if (N <= 0) return 0;
  cylinder_eval_kernel<<<(unsigned) ((N + 127) / 128), 128, 0, s>>>(c, pos, N, layout, n, grad, hess);
  return 1;
}

int launch_cylinder_splat(const CableP &c, float *data, const float *pos, const float *val,
                          const float *grad, int64_t N, int layout, cudaStream_t s) {
  if (N <= 0) return 0;
  cylinder_splat_kernel<<<(unsigned) ((N + 127) / 128), 128, 0, s>>>(c, data, pos, val, grad, N, layout);
  return 1;
}

}  // namespace drrt
