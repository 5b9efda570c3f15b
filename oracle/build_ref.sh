#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the reference's own CPU tracer
# (Tracer<false,false>) from the sources where they lie under /root/reference
# into the git-ignored oracle/_ref/.  Nothing from the reference is committed:
# the three .cpp files are streamed through a filter that drops only their
# explicit template instantiations into oracle/_ref/gen/.
# Two ISA builds are produced because the .so travels to a GPU box whose host
# CPU may differ from this container's:
#   libdrrt_ref_v3.so  -march=x86-64-v3 (AVX2+FMA)
#   libdrrt_ref_v4.so  -march=x86-64-v4 (AVX-512)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${DRRT_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
  echo "build_ref: $REF not present; keeping prebuilt files in $OUT" >&2
  exit 0
fi
mkdir -p "$OUT/gen"
for f in volume cylinder_volume tracer; do
  grep -v -E '^template (struct|class) (volume|cylinder_volume|Tracer)<' "$REF/src/$f.cpp" > "$OUT/gen/$f.cpp"
done
CXXFLAGS="-std=c++17 -O3 -DNDEBUG -fPIC -shared -w -I$OUT -I$REF/include -I$REF/ext/enoki/include"
for v in v3 v4; do
  if [ ! -f "$OUT/libdrrt_ref_$v.so" ] || [ "$HERE/ref_shim.cpp" -nt "$OUT/libdrrt_ref_$v.so" ]; then
    g++ $CXXFLAGS -march=x86-64-$v "$HERE/ref_shim.cpp" -o "$OUT/libdrrt_ref_$v.so" &
  fi
done
# The reference's CPU TAPE-AUTODIFF tracer, Tracer<true,false> (src/tracer.cpp:569), with enoki's
# autodiff back end compiled from where it lies (ext/enoki/src/autodiff/autodiff.cpp; the CPU
# instantiations are at :1229-1230).  One portable ISA build: it pins gradients, it is never timed.
if [ ! -f "$OUT/libdrrt_refad_v3.so" ] || [ "$HERE/ref_ad_shim.cpp" -nt "$OUT/libdrrt_refad_v3.so" ] \
   || [ "$HERE/ref_ad_stubs.cpp" -nt "$OUT/libdrrt_refad_v3.so" ]; then
  g++ -std=c++17 -O2 -DNDEBUG -fPIC -shared -w -I"$OUT" -I"$REF/include" -I"$REF/ext/enoki/include" \
      -march=x86-64-v3 -DENOKI_AUTODIFF_BUILD=1 "$HERE/ref_ad_shim.cpp" "$HERE/ref_ad_stubs.cpp" \
      "$REF/ext/enoki/src/autodiff/autodiff.cpp" -o "$OUT/libdrrt_refad_v3.so" &
fi
wait
ls -la "$OUT"/*.so
