// TEST INFRASTRUCTURE ONLY -- not product code.
//
// C-ABI wrapper around the *unmodified* reference's CPU TAPE-AUTODIFF tracer:
// drrt::Tracer<true,false> (the reference instantiates it at src/tracer.cpp:569) on
// enoki's DiffArray<DynamicArray<Packet<float>>> (instantiated by the reference's
// vendored enoki at ext/enoki/src/autodiff/autodiff.cpp:1229-1230).  It is what the
// reference's ADTracerC / ADTracerS return as d rif, d pos, d vel
// (core/tracer.py:45-66: set_gradient(outx), set_gradient(outv), FloatD.backward(),
// gradient(rif), gradient(x), gradient(v)) and what the paper checks its adjoint against.
//
// Like ref_shim.cpp, the reference's translation units are NOT copied into the
// repository: oracle/build_ref.sh streams them into the git-ignored oracle/_ref/gen/ with
// only their explicit-instantiation lines removed, and compiles enoki's autodiff.cpp from
// where it lies under /root/reference.  All buffers are HOST float32, rays AoS (N,3).
#include <cstdint>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#include "gen/volume.cpp"
#include "gen/cylinder_volume.cpp"
#include "gen/tracer.cpp"

// enoki's CUDA back end is referenced only by dead code on this path.
namespace enoki {
void cuda_dec_ref_ext(uint32_t) { std::abort(); }
void cuda_inc_ref_ext(uint32_t) { std::abort(); }
size_t cuda_var_size(uint32_t) { std::abort(); }
void cuda_fetch_element(void *, uint32_t, size_t, size_t) { std::abort(); }
uint32_t cuda_trace_append(EnokiType, const char *) { std::abort(); }
uint32_t cuda_trace_append(EnokiType, const char *, uint32_t) { std::abort(); }
uint32_t cuda_trace_append(EnokiType, const char *, uint32_t, uint32_t) { std::abort(); }
uint32_t cuda_trace_append(EnokiType, const char *, uint32_t, uint32_t, uint32_t) { std::abort(); }
}  // namespace enoki

namespace drrt {
template struct volume<true, false>;
template struct cylinder_volume<true, false>;
template class Tracer<true, false>;
}  // namespace drrt

using FloatP_ = drrt::Float<false, false>;   // DynamicArray<Packet<float>>
using FloatA_ = drrt::Float<true, false>;    // DiffArray<DynamicArray<Packet<float>>>
using Vec3A_ = drrt::Vector3f<true, false>;
using TracerA_ = drrt::Tracer<true, false>;

static thread_local std::string g_err;

static FloatA_ load1(const float *p, int64_t n) { return FloatA_(FloatP_::copy(p, (size_t) n)); }

static FloatP_ comp(const float *p, int64_t n, int k) {
  FloatP_ a = enoki::empty<FloatP_>((size_t) n);
  for (int64_t i = 0; i < n; ++i) a.coeff(i) = p[3 * i + k];
  return a;
}

static Vec3A_ load3(const float *p, int64_t n) {
  return Vec3A_(FloatA_(comp(p, n, 0)), FloatA_(comp(p, n, 1)), FloatA_(comp(p, n, 2)));
}

static void store1p(const FloatP_ &v, float *p, int64_t n) {
  const size_t s = enoki::slices(v);
  for (int64_t i = 0; i < n; ++i) p[i] = s == 0 ? 0.f : v.coeff(s == 1 ? 0 : i);
}

static void store3p(const FloatP_ &x, const FloatP_ &y, const FloatP_ &z, float *p, int64_t n) {
  const size_t sx = enoki::slices(x), sy = enoki::slices(y), sz = enoki::slices(z);
  for (int64_t i = 0; i < n; ++i) {
    p[3 * i + 0] = sx == 0 ? 0.f : x.coeff(sx == 1 ? 0 : i);
    p[3 * i + 1] = sy == 0 ? 0.f : y.coeff(sy == 1 ? 0 : i);
    p[3 * i + 2] = sz == 0 ? 0.f : z.coeff(sz == 1 ? 0 : i);
  }
}

extern "C" {

const char *refad_last_error() { return g_err.c_str(); }

// ADTracerC forward + backward (core/tracer.py:17-66) on the CPU tape:
//   (xt, vt) = Tracer<true,false>().trace(rif, res, pos, vel, h, ds)
//   set_gradient(xt, seed_x); set_gradient(vt, seed_v); backward()
//   -> d_rif[G], d_pos[N,3], d_vel[N,3]
// `sdf` non-NULL selects trace_sdf (ADSDFTracerC, core/tracer.py:247-291).
int refad_trace_backward(const float *rif, const float *sdf, int W, int H, int D, const float *pos,
                         const float *vel, int64_t N, float h, float ds, const float *seed_x,
                         const float *seed_v, float *xt, float *vt, float *d_rif, float *d_pos,
                         float *d_vel) {
  try {
    const int64_t G = (int64_t) W * H * D;
    FloatA_ r = load1(rif, G);
    Vec3A_ p = load3(pos, N), v = load3(vel, N);
    enoki::set_requires_gradient(r, true);
    enoki::set_requires_gradient(p, true);
    enoki::set_requires_gradient(v, true);
    std::pair<Vec3A_, Vec3A_> out;
    if (sdf) {
      FloatA_ s = load1(sdf, G);
      out = TracerA_().trace_sdf(r, s, drrt::ScalarVector3i(W, H, D), p, v, h, ds);
    } else {
      out = TracerA_().trace(r, drrt::ScalarVector3i(W, H, D), p, v, h, ds);
    }
    store3p(enoki::detach(out.first.x()), enoki::detach(out.first.y()), enoki::detach(out.first.z()), xt, N);
    store3p(enoki::detach(out.second.x()), enoki::detach(out.second.y()), enoki::detach(out.second.z()), vt, N);
    for (int k = 0; k < 3; ++k) {
      // an output component that never touched the tape (e.g. vt of a ray that flew
      // straight) has no node; seeding it is a no-op in the reference's Python as well
      if (enoki::requires_gradient(out.first[k])) enoki::set_gradient(out.first[k], comp(seed_x, N, k));
      if (enoki::requires_gradient(out.second[k])) enoki::set_gradient(out.second[k], comp(seed_v, N, k));
    }
    enoki::backward<FloatA_>();
    store1p(enoki::gradient(r), d_rif, G);
    store3p(enoki::gradient(p.x()), enoki::gradient(p.y()), enoki::gradient(p.z()), d_pos, N);
    store3p(enoki::gradient(v.x()), enoki::gradient(v.y()), enoki::gradient(v.z()), d_vel, N);
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
  return 0;
}

}  // extern "C"
