// the sensor epilogue that follows every trace in the
// reference's drivers (reference core/image_opt.py:97-118, core/luneburg_opt.py:100-122):
//
//   near field  sensor.generate_sensor       core/sensor.py:5-28
//                 = trace_rays_to_plane       core/sensor.py:195-202
//                 + |v.n| foreshortening      core/sensor.py:17-18
//                 + projection on (t1, t2)    core/sensor.py:21-22, get_tan_vecs :218-229
//                 + Grid.Splat(average=False) core/grid.py:133-151 over the 4x4 radial-tent
//                   stencil of Grid.index_values / rbf_tent (core/grid.py:38-66, :80-85)
//   far field   sensor.generate_inf_sensor   core/sensor.py:31-54 (normalised direction
//                 projected on (t1, t2), no foreshortening)
//
// The reference runs this as ~40 torch elementwise / index kernels over (N,16) temporaries
// and differentiates it with torch autograd.  Here: one forward kernel (one ray per thread,
// <= 16 float REDs into the res x res image, which is L2-resident) and one backward kernel
// that gathers dL/dimage on the same stencil and writes dL/dx, dL/dv (and dL/de) directly.
#include "drrt_device.cuh"
#include "drrt_launch.h"

namespace drrt {

namespace {

constexpr float kSqrt2 = 1.41421356237309504880f;

struct Stencil {
  float na, nb;   // norm_x = xn / h - 0.5 (core/grid.py:39)
  int ia, ib;     // floor(norm_x) (core/grid.py:41)
};

// Everything a ray needs before the stencil: plane-space coordinates and, for the
// backward pass, the pieces of the chain rule.
struct Proj {
  float ua, ub;   // xn (near) or vn (far), core/sensor.py:21-22 / :47-48
  float fe;       // e * |v.n| (near) or e (far)
  float fs_sign;  // sign(v.n) (near)
  float t, ndv;   // near: ray parameter of the plane hit and n.v
  float vnorm;    // far: |v|
};

// NORM: far field only -- project the normalised direction (generate_inf_sensor) or the raw
// one (get_sdf_vals_far, core/sensor.py:150-151)
template <bool FAR, bool NORM = true>
__device__ __forceinline__ Proj project(const SensorArgs &a, f3 x, f3 v, float e) {
  Proj p;
  const f3 n = {a.n[0], a.n[1], a.n[2]}, t1 = {a.t1[0], a.t1[1], a.t1[2]}, t2 = {a.t2[0], a.t2[1], a.t2[2]};
  if (!FAR) {
    const f3 o = {a.p[0], a.p[1], a.p[2]};
    // trace_rays_to_plane: t = n.(p - x) / n.v; xp = x + t v
    p.ndv = n.x * v.x + n.y * v.y + n.z * v.z;
    p.t = (n.x * (o.x - x.x) + n.y * (o.y - x.y) + n.z * (o.z - x.z)) / p.ndv;
    const f3 d = {fmaf(p.t, v.x, x.x) - o.x, fmaf(p.t, v.y, x.y) - o.y, fmaf(p.t, v.z, x.z) - o.z};
    p.ua = d.x * t1.x + d.y * t1.y + d.z * t1.z + a.offset;
    p.ub = d.x * t2.x + d.y * t2.y + d.z * t2.z + a.offset;
    p.fe = e * fabsf(p.ndv);
    p.fs_sign = p.ndv > 0.f ? 1.f : (p.ndv < 0.f ? -1.f : 0.f);
    p.vnorm = 1.f;
  } else {
    p.vnorm = NORM ? sqrtf(v.x * v.x + v.y * v.y + v.z * v.z) : 1.f;
    const float r = 1.f / p.vnorm;
    p.ua = (v.x * t1.x + v.y * t1.y + v.z * t1.z) * r + a.offset;
    p.ub = (v.x * t2.x + v.y * t2.y + v.z * t2.z) * r + a.offset;
    p.fe = e;
    p.fs_sign = 0.f; p.t = 0.f; p.ndv = 1.f;
  }
  return p;
}

__device__ __forceinline__ Stencil stencil_of(const SensorArgs &a, const Proj &p) {
  Stencil s;
  s.na = p.ua * a.rcp_h - 0.5f;
  s.nb = p.ub * a.rcp_h - 0.5f;
  s.ia = (int) floorf(s.na);
  s.ib = (int) floorf(s.nb);
  return s;
}

// radial tent weights of the 4x4 taps (core/grid.py:80-85): w = max(sqrt2 - r, 0)
__device__ __forceinline__ float tap_weights(const Stencil &s, float w[16]) {
  float W = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float da = s.na - (float) (s.ia - 1 + i);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float db = s.nb - (float) (s.ib - 1 + j);
      const float r = sqrtf(da * da + db * db);
      w[4 * i + j] = fmaxf(kSqrt2 - r, 0.f);
      W += w[4 * i + j];
    }
  }
  return W;
}

__device__ __forceinline__ bool finite_ray(const Proj &p) {
  // rays parallel to the plane (n.v = 0) give inf/nan coordinates; torch's .long() of those
  // is out of range for every tap, so they contribute nothing
  return isfinite(p.ua) && isfinite(p.ub) && fabsf(p.ua) < 1e9f && fabsf(p.ub) < 1e9f;
}

template <bool FAR>
__global__ void __launch_bounds__(256) sensor_splat_kernel(SensorArgs a) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.N) return;
  const f3 x = FAR ? f3{0, 0, 0} : ld3(a.x, r, a.N, a.layout);
  const f3 v = ld3(a.v, r, a.N, a.layout);
  const float e = a.e ? a.e[r] : a.e0;
  const Proj p = project<FAR>(a, x, v, e);
  if (!finite_ray(p)) return;
  const Stencil s = stencil_of(a, p);
  float w[16];
  const float W = tap_weights(s, w);
  const float scale = p.fe / W;  // average=False: we = w / w.sum (core/grid.py:145-146)
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ia = s.ia - 1 + i;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ib = s.ib - 1 + j;
      const float val = w[4 * i + j] * scale;
      // per-tap bounds mask (core/grid.py:139-141); zero-weight taps add nothing
      if (val != 0.f && ia >= 0 && ia < a.res && ib >= 0 && ib < a.res)
        atomicAdd(a.image + (size_t) ia * a.res + ib, val);
    }
  }
}

// Backward of the above for L = sum(dimage * image):
//   c_k = fe w_k / W;  A = sum_inb G_k w_k;  dL/dnorm = fe (A' W - A W') / W^2 with
//   w_k' = -d_k / r_k on the support;  dL/dfe = A / W.
//
// GATHER: backward of sensor_gather_kernel instead (Grid.Get, core/grid.py:106-121): the tap
// values are the texture itself at CLAMPED indices (no bounds mask, core/grid.py:51), the
// per-ray factor is the upstream gradient dL/df (passed in `e`), and there is no
// foreshortening term.
template <bool FAR, bool GATHER>
__global__ void __launch_bounds__(256) sensor_splat_bwd_kernel(SensorArgs a) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.N) return;
  const f3 x = FAR ? f3{0, 0, 0} : ld3(a.x, r, a.N, a.layout);
  const f3 v = ld3(a.v, r, a.N, a.layout);
  const float e = a.e ? a.e[r] : a.e0;
  Proj p = project<FAR, !GATHER>(a, x, v, e);
  if (GATHER) { p.fe = e; p.fs_sign = 0.f; }
  f3 gx = {0, 0, 0}, gv = {0, 0, 0};
  float ge = 0.f;
  if (finite_ray(p)) {
    const Stencil s = stencil_of(a, p);
    float W = 0.f, A = 0.f, Aa = 0.f, Ab = 0.f, Wa = 0.f, Wb = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int ia = s.ia - 1 + i;
      const float da = s.na - (float) ia;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int ib = s.ib - 1 + j;
        const float db = s.nb - (float) ib;
        const float rr = sqrtf(da * da + db * db);
        const float w = fmaxf(kSqrt2 - rr, 0.f);
        W += w;
        if (w > 0.f) {
          const float ir = rr > 0.f ? 1.f / rr : 0.f;  // d|d|/dd at 0: torch's subgradient 0
          const float wa = -da * ir, wb = -db * ir;
          Wa += wa; Wb += wb;
          if (GATHER) {
            const int ca = min(max(ia, 0), a.res - 1), cb = min(max(ib, 0), a.res - 1);
            const float G = a.dimage[(size_t) ca * a.res + cb];
            A = fmaf(G, w, A); Aa = fmaf(G, wa, Aa); Ab = fmaf(G, wb, Ab);
          } else if (ia >= 0 && ia < a.res && ib >= 0 && ib < a.res) {
            const float G = a.dimage[(size_t) ia * a.res + ib];
            A = fmaf(G, w, A); Aa = fmaf(G, wa, Aa); Ab = fmaf(G, wb, Ab);
          }
        }
      }
    }
    const float iW = 1.f / W;
    const float AoW = A * iW;
    // d/dnorm, then norm = u / h - 0.5
    const float k = p.fe * iW * a.rcp_h;
    const float ga = k * (Aa - AoW * Wa), gb = k * (Ab - AoW * Wb);
    const f3 q = {a.t1[0] * ga + a.t2[0] * gb, a.t1[1] * ga + a.t2[1] * gb, a.t1[2] * ga + a.t2[2] * gb};
    if (!FAR) {
      // xp = x + t v, t = n.(p - x) / n.v:  dL/dx = q - n (q.v)/(n.v), dL/dv = t dL/dx
      const float c = (q.x * v.x + q.y * v.y + q.z * v.z) / p.ndv;
      gx.x = q.x - a.n[0] * c; gx.y = q.y - a.n[1] * c; gx.z = q.z - a.n[2] * c;
      // fe = e |v.n|
      const float f = p.fs_sign * e * AoW;
      gv.x = fmaf(p.t, gx.x, a.n[0] * f); gv.y = fmaf(p.t, gx.y, a.n[1] * f); gv.z = fmaf(p.t, gx.z, a.n[2] * f);
      ge = fabsf(p.ndv) * AoW;
    } else if (GATHER) {
      gv = q;  // vn = v . T, no normalisation (core/sensor.py:150)
    } else {
      // vn = (v / |v|) . T:  dL/dv = (q - vhat (vhat.q)) / |v|
      const float iv = 1.f / p.vnorm;
      const f3 vh = {v.x * iv, v.y * iv, v.z * iv};
      const float c = vh.x * q.x + vh.y * q.y + vh.z * q.z;
      gv.x = (q.x - vh.x * c) * iv; gv.y = (q.y - vh.y * c) * iv; gv.z = (q.z - vh.z * c) * iv;
      ge = AoW;
    }
  }
  if (a.dx) st3(a.dx, r, a.N, a.layout, gx);
  if (a.dv) st3(a.dv, r, a.N, a.layout, gv);
  if (a.de) a.de[r] = ge;
}

// Grid.Get on the sensor plane: get_sdf_vals_near / get_sdf_vals_far (core/sensor.py:116-155):
// f = sum_k tex[clamp(idx_k)] w_k / sum_k w_k over the same 4x4 radial-tent stencil.
template <bool FAR>
__global__ void __launch_bounds__(256) sensor_gather_kernel(SensorArgs a) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.N) return;
  const f3 x = FAR ? f3{0, 0, 0} : ld3(a.x, r, a.N, a.layout);
  const f3 v = ld3(a.v, r, a.N, a.layout);
  const Proj p = project<FAR, false>(a, x, v, 1.f);
  float f = 0.f;
  if (finite_ray(p)) {
    const Stencil s = stencil_of(a, p);
    float w[16];
    const float W = tap_weights(s, w);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int ca = min(max(s.ia - 1 + i, 0), a.res - 1);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int cb = min(max(s.ib - 1 + j, 0), a.res - 1);
        if (w[4 * i + j] != 0.f) f = fmaf(a.dimage[(size_t) ca * a.res + cb], w[4 * i + j], f);
      }
    }
    f /= W;
  }
  a.de[r] = f;
}

}  // namespace

int launch_sensor_gather(bool far, const SensorArgs &a, cudaStream_t s) {
  if (a.N <= 0) return 0;
  const unsigned nb = (unsigned) ((a.N + 255) / 256);
  if (far) sensor_gather_kernel<true><<<nb, 256, 0, s>>>(a);
  else sensor_gather_kernel<false><<<nb, 256, 0, s>>>(a);
  return 1;
}

int launch_sensor_gather_bwd(bool far, const SensorArgs &a, cudaStream_t s) {
  if (a.N <= 0) return 0;
  const unsigned nb = (unsigned) ((a.N + 255) / 256);
  if (far) sensor_splat_bwd_kernel<true, true><<<nb, 256, 0, s>>>(a);
  else sensor_splat_bwd_kernel<false, true><<<nb, 256, 0, s>>>(a);
  return 1;
}

int launch_sensor_splat(bool far, const SensorArgs &a, cudaStream_t s) {
  if (a.N <= 0) return 0;
  const unsigned nb = (unsigned) ((a.N + 255) / 256);
  if (far) sensor_splat_kernel<true><<<nb, 256, 0, s>>>(a);
  else sensor_splat_kernel<false><<<nb, 256, 0, s>>>(a);
  return 1;
}

int launch_sensor_splat_bwd(bool far, const SensorArgs &a, cudaStream_t s) {
  if (a.N <= 0) return 0;
  const unsigned nb = (unsigned) ((a.N + 255) / 256);
  if (far) sensor_splat_bwd_kernel<true, false><<<nb, 256, 0, s>>>(a);
  else sensor_splat_bwd_kernel<false, false><<<nb, 256, 0, s>>>(a);
  return 1;
}

}  // namespace drrt
