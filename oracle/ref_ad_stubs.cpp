// TEST INFRASTRUCTURE ONLY.  Link-time stubs for the three members of enoki's CUDA tape
// (Tape<CUDAArray<float>>::get, ::dec_ref_ext and ~DiffArray<CUDAArray<float>>) that the
// reference's *dead* 1x1x1-grid branch (src/volume.cpp:117-121, `Vector3fD(0,0,0)`) drags
// into the CPU autodiff build of oracle/ref_ad_shim.cpp.  enoki defines them only when it is
// built with its CUDA back end (ext/enoki/src/autodiff/autodiff.cpp, `#if defined(ENOKI_CUDA)`),
// which needs a GPU driver at load time, and its header declares them `extern template`, so
// they are supplied here under their Itanium-mangled names.  They abort: nothing on the CPU
// tape path reaches them.
#include <cstdlib>

extern "C" {
void *_ZN5enoki4TapeINS_9CUDAArrayIfEEE3getEv() { std::abort(); }
void _ZN5enoki4TapeINS_9CUDAArrayIfEEE11dec_ref_extEj(void *, unsigned) { std::abort(); }
void _ZN5enoki9DiffArrayINS_9CUDAArrayIfEEED1Ev(void *) { std::abort(); }
}
