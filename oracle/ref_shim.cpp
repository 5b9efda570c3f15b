// TEST INFRASTRUCTURE ONLY -- not product code.
//
// C-ABI wrapper around the *unmodified* reference implementation's CPU
// instantiation (drrt::Tracer<false,false>, volume<false,false>,
// cylinder_volume<false,false>; the reference instantiates them itself at
// src/tracer.cpp:570, src/volume.cpp:281, src/cylinder_volume.cpp:180 and binds
// them as `TracerS`, src/drrt.cpp:38-45).
//
// The three reference translation units are NOT copied into this repository:
// oracle/build_ref.sh streams them from /root/reference through a filter that
// only drops their explicit-instantiation lines (the <*,true> CUDA
// instantiations need enoki's CUDA runtime) into the git-ignored oracle/_ref/gen/
// and this file #includes the result.  All buffers are HOST float32, ray
// vectors are AoS (N,3) like the torch tensors in core/tracer.py.
#include <cstdint>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#include "gen/volume.cpp"
#include "gen/cylinder_volume.cpp"
#include "gen/tracer.cpp"

// enoki's CUDA back end is referenced only by dead code on this path
// (src/tracer.cpp:16-33 and the 1x1x1 branch src/volume.cpp:117-121).
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
template struct volume<false, false>;
template struct cylinder_volume<false, false>;
template class Tracer<false, false>;
}  // namespace drrt

using FloatS_ = drrt::Float<false, false>;
using Vec3S_ = drrt::Vector3f<false, false>;
using TracerS_ = drrt::Tracer<false, false>;

static thread_local std::string g_err;

static FloatS_ load1(const float *p, int64_t n) { return FloatS_::copy(p, (size_t) n); }

static Vec3S_ load3(const float *p, int64_t n) {
  FloatS_ a = enoki::empty<FloatS_>((size_t) n), b = enoki::empty<FloatS_>((size_t) n),
          c = enoki::empty<FloatS_>((size_t) n);
  for (int64_t i = 0; i < n; ++i) {
    a.coeff(i) = p[3 * i + 0];
    b.coeff(i) = p[3 * i + 1];
    c.coeff(i) = p[3 * i + 2];
  }
  return Vec3S_(a, b, c);
}

static void store3(const Vec3S_ &v, float *p, int64_t n) {
  for (int64_t i = 0; i < n; ++i) {
    p[3 * i + 0] = v.x().coeff(i);
    p[3 * i + 1] = v.y().coeff(i);
    p[3 * i + 2] = v.z().coeff(i);
  }
}

static void store1(const FloatS_ &v, float *p, int64_t n) {
  if ((int64_t) enoki::slices(v) == 1 && n > 1) {
    for (int64_t i = 0; i < n; ++i) p[i] = v.coeff(0);
  } else {
    for (int64_t i = 0; i < n; ++i) p[i] = v.coeff(i);
  }
}

#define REF_TRY try {
#define REF_CATCH                                  \
  }                                                \
  catch (const std::exception &e) {                \
    g_err = e.what();                              \
    return -1;                                     \
  }                                                \
  return 0;

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

// Which ISA this build of the reference was compiled for ("v3" = AVX2+FMA,
// "v4" = AVX-512); enoki's rcp() differs between the two.
const char *ref_isa() {
#if defined(__AVX512F__)
  return "avx512";
#elif defined(__AVX2__)
  return "avx2";
#else
  return "sse";
#endif
}

// The library's own rcp(h) as volume::eval_grad computes it (src/volume.cpp:128 on the size-1 array h_,
// include/volume.h:49): enoki's packet rcp is rcp14 + one Newton step on AVX-512 and rcp + Newton on AVX2,
// neither always equal to the correctly rounded 1/h.  Tests use it to tell which build shares the GPU's
// 1/h for a given spacing (lattice-exact samples select their cell by the last ulp of p * rcp(h)).
float ref_rcp(float h) {
  FloatS_ hh(h);
  FloatS_ r = enoki::rcp(hh);
  return r.coeff(0);
}

int ref_trace(const float *rif, int W, int H, int D, const float *pos, const float *vel, int64_t N,
              float h, float ds, float *xt, float *vt) {
  REF_TRY
  FloatS_ r = load1(rif, (int64_t) W * H * D);
  Vec3S_ p = load3(pos, N), v = load3(vel, N);
  auto out = TracerS_().trace(r, drrt::ScalarVector3i(W, H, D), p, v, h, ds);
  store3(out.first, xt, N);
  store3(out.second, vt, N);
  REF_CATCH
}

int ref_trace_plane(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                    const float *pln_o, const float *pln_d, int64_t N, float h, float ds, float *xt,
                    float *vt, uint8_t *not_escaped) {
  REF_TRY
  FloatS_ r = load1(rif, (int64_t) W * H * D);
  Vec3S_ p = load3(pos, N), v = load3(vel, N), po = load3(pln_o, N), pd = load3(pln_d, N);
  auto out = TracerS_().trace_plane(r, drrt::ScalarVector3i(W, H, D), p, v, po, pd, h, ds);
  store3(std::get<0>(out), xt, N);
  store3(std::get<1>(out), vt, N);
  auto &m = std::get<2>(out);
  for (int64_t i = 0; i < N; ++i) not_escaped[i] = m.coeff(enoki::slices(m) == 1 ? 0 : i) ? 1 : 0;
  REF_CATCH
}

int ref_trace_target(const float *rif, int W, int H, int D, const float *pos, const float *vel,
                     const float *target, int64_t N, float h, float ds, float *xt, float *vt,
                     float *dist2) {
  REF_TRY
  FloatS_ r = load1(rif, (int64_t) W * H * D);
  Vec3S_ p = load3(pos, N), v = load3(vel, N), t = load3(target, N);
  auto out = TracerS_().trace_target(r, drrt::ScalarVector3i(W, H, D), p, v, t, h, ds);
  store3(std::get<0>(out), xt, N);
  store3(std::get<1>(out), vt, N);
  store1(std::get<2>(out), dist2, N);
  REF_CATCH
}

int ref_trace_sdf(const float *rif, const float *sdf, int W, int H, int D, const float *pos,
                  const float *vel, int64_t N, float h, float ds, float *xt, float *vt) {
  REF_TRY
  FloatS_ r = load1(rif, (int64_t) W * H * D), s = load1(sdf, (int64_t) W * H * D);
  Vec3S_ p = load3(pos, N), v = load3(vel, N);
  auto out = TracerS_().trace_sdf(r, s, drrt::ScalarVector3i(W, H, D), p, v, h, ds);
  store3(out.first, xt, N);
  store3(out.second, vt, N);
  REF_CATCH
}

int ref_trace_cable(const float *rif, int res, float radius, float length, const float *pos,
                    const float *vel, const float *target, int64_t N, float ds, float *xt,
                    float *vt, float *dist2) {
  REF_TRY
  FloatS_ r = load1(rif, res);
  Vec3S_ p = load3(pos, N), v = load3(vel, N), t = load3(target, N);
  auto out = TracerS_().trace_cable(r, radius, length, p, v, t, ds);
  store3(std::get<0>(out), xt, N);
  store3(std::get<1>(out), vt, N);
  store1(std::get<2>(out), dist2, N);
  REF_CATCH
}

int ref_backtrace(const float *rif, int W, int H, int D, const float *xt, const float *vt,
                  const float *dx, const float *dv, int64_t N, float h, float ds, float *grad) {
  REF_TRY
  int64_t G = (int64_t) W * H * D;
  FloatS_ r = load1(rif, G);
  Vec3S_ a = load3(xt, N), b = load3(vt, N), c = load3(dx, N), d = load3(dv, N);
  FloatS_ g = TracerS_().backtrace(r, drrt::ScalarVector3i(W, H, D), a, b, c, d, h, ds);
  store1(g, grad, G);
  REF_CATCH
}

int ref_backtrace_sdf(const float *rif, const float *sdf, int W, int H, int D, const float *xt,
                      const float *vt, const float *dx, const float *dv, int64_t N, float h,
                      float ds, float *grad) {
  REF_TRY
  int64_t G = (int64_t) W * H * D;
  FloatS_ r = load1(rif, G), s = load1(sdf, G);
  Vec3S_ a = load3(xt, N), b = load3(vt, N), c = load3(dx, N), d = load3(dv, N);
  FloatS_ g = TracerS_().backtrace_sdf(r, s, drrt::ScalarVector3i(W, H, D), a, b, c, d, h, ds);
  store1(g, grad, G);
  REF_CATCH
}

int ref_backtrace_cable(const float *rif, int res, float radius, float length, const float *xt,
                        const float *vt, const float *dx, const float *dv, int64_t N, float ds,
                        float *grad) {
  REF_TRY
  FloatS_ r = load1(rif, res);
  Vec3S_ a = load3(xt, N), b = load3(vt, N), c = load3(dx, N), d = load3(dv, N);
  FloatS_ g = TracerS_().backtrace_cable(r, radius, length, a, b, c, d, ds);
  store1(g, grad, res);
  REF_CATCH
}

// volume::{eval_grad, eval_hess, splat} probes (fixtures of src/test.cpp:61-80).
int ref_volume_eval(const float *rif, int W, int H, int D, const float *pos, int64_t N, float h,
                    float *n, float *grad /*N,3*/, float *hess /*N,3: xy,xz,yz*/) {
  REF_TRY
  FloatS_ r = load1(rif, (int64_t) W * H * D);
  Vec3S_ p = load3(pos, N);
  drrt::volume<false, false> vol(drrt::ScalarVector3i(W, H, D), r, h);
  enoki::mask_t<FloatS_> m = enoki::full<enoki::mask_t<FloatS_>>(true, (size_t) N);
  auto out = vol.eval_grad(p, m);
  auto Hs = vol.eval_hess(p, m);
  store1(out.first, n, N);
  store3(out.second, grad, N);
  for (int64_t i = 0; i < N; ++i) {
    hess[3 * i + 0] = Hs(0, 1).coeff(i);
    hess[3 * i + 1] = Hs(0, 2).coeff(i);
    hess[3 * i + 2] = Hs(1, 2).coeff(i);
  }
  REF_CATCH
}

int ref_volume_splat(float *data, int W, int H, int D, const float *pos, const float *val,
                     const float *grad, int64_t N, float h) {
  REF_TRY
  int64_t G = (int64_t) W * H * D;
  FloatS_ r = load1(data, G);
  Vec3S_ p = load3(pos, N), g = load3(grad, N);
  FloatS_ v = load1(val, N);
  drrt::volume<false, false> vol(drrt::ScalarVector3i(W, H, D), r, h);
  enoki::mask_t<FloatS_> m = enoki::full<enoki::mask_t<FloatS_>>(true, (size_t) N);
  vol.splat(p, v, g, m);
  store1(vol.get_data(), data, G);
  REF_CATCH
}

// cylinder_volume probes (fixture of src/test.cpp:82-101).
int ref_cylinder_eval(const float *rif, int res, float radius, float length, const float *pos,
                      int64_t N, float *n, float *grad /*N,3*/, float *hess /*N,9 row-major*/) {
  REF_TRY
  FloatS_ r = load1(rif, res);
  Vec3S_ p = load3(pos, N);
  drrt::cylinder_volume<false, false> vol(r, radius, length);
  enoki::mask_t<FloatS_> m = enoki::full<enoki::mask_t<FloatS_>>(true, (size_t) N);
  auto out = vol.eval_grad(p, m);
  auto Hs = vol.eval_hess(p, m);
  store1(out.first, n, N);
  store3(out.second, grad, N);
  for (int64_t i = 0; i < N; ++i)
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        auto &e = Hs(a, b);
        hess[9 * i + 3 * a + b] = e.coeff(enoki::slices(e) == 1 ? 0 : i);
      }
  REF_CATCH
}

int ref_cylinder_splat(float *data, int res, float radius, float length, const float *pos,
                       const float *val, const float *grad, int64_t N) {
  REF_TRY
  FloatS_ r = load1(data, res);
  Vec3S_ p = load3(pos, N), g = load3(grad, N);
  FloatS_ v = load1(val, N);
  drrt::cylinder_volume<false, false> vol(r, radius, length);
  enoki::mask_t<FloatS_> m = enoki::full<enoki::mask_t<FloatS_>>(true, (size_t) N);
  vol.splat(p, v, g, m);
  store1(vol.get_data(), data, res);
  REF_CATCH
}

}  // extern "C"
