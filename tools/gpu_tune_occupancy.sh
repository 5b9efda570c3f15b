# Dev script (GPU box): forward / adjoint times of occupancy-tuning builds of the grid kernels
# (lib/var_<tag>/libdrrt_b200.so, built with -DDRRT_ADJ_MIN_BLOCKS / -DDRRT_TRACE_MIN_BLOCKS).
for tag in base ${TAGS:-}; do
  if [ $tag = base ]; then unset DRRT_B200_LIB; else export DRRT_B200_LIB=$PWD/adjointnonlinearraytracing_b200/lib/var_$tag/libdrrt_b200.so; fi
  [ -n "${DRRT_B200_LIB:-}" ] && [ ! -f "$DRRT_B200_LIB" ] && continue
  python - <<PY
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from adjointnonlinearraytracing_b200 import ops, scenes as sc
from tools.gpu_configs import timeit, multiview_rays
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
for name, ng, (pos, vel) in (("caustic", 256, sc.rays_collimated(2048)), ("cone", 256, sc.rays_cone(2048, spread=0.05)),
                             ("fuel", 65, multiview_rays(32, 64, 16))):
    rif = T(sc.field_bump(ng)); h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = T(pos), T(vel); one = torch.ones_like(pos)
    xt, vt = ops.trace(rif, pos, vel, h, ds)
    tf = timeit(lambda: ops.trace(rif, pos, vel, h, ds)); tb = timeit(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds))
    print(f"$tag {name:8s} fwd {tf:7.3f} ms  adj {tb:7.3f} ms", flush=True)
PY
done
