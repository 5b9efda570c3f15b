// Forward eikonal trace and adjoint backtrace through a regular IOR grid.
//
// One ray per thread, the whole integration loop in registers, per-ray exit:
// the reference instead launches one JIT kernel + one host-synchronising
// all()/none() reduction per step (reference src/tracer.cpp:82, :426) and
// streams every ray's state through global memory each step.
//
// Per-ray early exit is exact for trace / trace_plane / trace_sdf / backtrace*:
// once a ray is `escaped` (forward) or not `active` (adjoint) none of its
// outputs can change again (gathers are masked to 0 outside, so it flies
// straight away from a convex box).  trace_target is the exception -- see
// trace_target_tail_kernel.
#include "drrt_device.cuh"
#include "drrt_launch.h"

namespace drrt {

enum { K_PLAIN = 0, K_PLANE = 1, K_TARGET = 2, K_SDF = 3 };

// Arithmetic flavours (see drrt_device.cuh and DESIGN.md section 4):
//   A_FAST    factored per-cell polynomial everywhere
//   A_REF     the reference's literal operation order everywhere
//   A_EXACT   (default) the ray trajectory -- cell lookup, n, grad n, the v and x
//             updates, every branch -- in the reference's literal operation order,
//             so endpoints and the set of visited cells are bit-identical to A_REF;
//             the quantities that only feed the (linear) gradient accumulation --
//             Hessian, splat weights, lambda/mu -- in factored FMA form.
enum { A_FAST = 0, A_REF = 1, A_EXACT = 2 };

// Register cache of the cell a ray is in: the floor coordinates it was filled for
// (floats; NaN = empty, so the first comparison reports a change) and the cell's
// data in the form the arithmetic flavour wants.
struct CellTag {
  float fx, fy, fz;
  __device__ __forceinline__ void reset() { fx = __int_as_float(0x7fc00000); fy = fz = 0.f; }
  __device__ __forceinline__ bool differs(const CellIdx &c) const {
    return (c.fx != fx) | (c.fy != fy) | (c.fz != fz);
  }
  __device__ __forceinline__ void set(const CellIdx &c) { fx = c.fx; fy = c.fy; fz = c.fz; }
};

// SM: the grid is a shared-memory copy (small grids)
template <int AR, bool SM = false>
struct CellCache;

template <bool SM>
struct CellCache<A_REF, SM> : CellTag {
  f32x2 V[4];  // x-adjacent corner pairs
  __device__ __forceinline__ CellAddr refill(const GridP &g, const CellIdx &c) {
    set(c);
    return gather_cell<SM>(g, c.cx(), c.cy(), c.cz(), V);
  }
  __device__ __forceinline__ void eval(const GridP &g, const CellIdx &c, float &n, f3 &nx, f3 *hm) const {
    eval_grad_ref_x2(g, V, c.wx, c.wy, c.wz, n, nx);
    if (hm) {
      const float v[8] = {lo2(V[0]), hi2(V[0]), lo2(V[1]), hi2(V[1]), lo2(V[2]), hi2(V[2]), lo2(V[3]), hi2(V[3])};
      *hm = eval_hess_ref(g, v, c.wx, c.wy, c.wz);
    }
  }
};

template <bool SM>
struct CellCache<A_FAST, SM> : CellTag {
  CellPoly p;
  __device__ __forceinline__ CellAddr refill(const GridP &g, const CellIdx &c) {
    set(c);
    f32x2 V[4];
    CellAddr a = gather_cell<SM>(g, c.cx(), c.cy(), c.cz(), V);
    const float v[8] = {lo2(V[0]), hi2(V[0]), lo2(V[1]), hi2(V[1]), lo2(V[2]), hi2(V[2]), lo2(V[3]), hi2(V[3])};
    p = make_poly(v);
    return a;
  }
  // nx and hm are returned in world units
  __device__ __forceinline__ void eval(const GridP &g, const CellIdx &c, float &n, f3 &nx, f3 *hm) const {
    f3 gg, hh;
    eval_poly(p, c.wx, c.wy, c.wz, n, gg, hh);
    nx.x = gg.x * g.rcp_h; nx.y = gg.y * g.rcp_h; nx.z = gg.z * g.rcp_h;
    if (hm) {
      float r2 = g.rcp_h * g.rcp_h;
      hm->x = hh.x * r2; hm->y = hh.y * r2; hm->z = hh.z * r2;
    }
  }
};

template <bool SM>
struct CellCache<A_EXACT, SM> : CellTag {
  f32x2 V[4];  // x-adjacent corner pairs
  // mixed second differences in the reference's association order
  // (src/volume.cpp:79-90; they cancel to ~1e-4 of the voxel values, so the order
  // of the subtractions matters at the 1e-3 relative level), pre-scaled by 1/h^2:
  // d?0 at w = 0 and the slope d?1 - d?0.
  float hxy0, hxys, hxz0, hxzs, hyz0, hyzs;
  __device__ __forceinline__ CellAddr refill(const GridP &g, const CellIdx &c) {
    set(c);
    CellAddr a = gather_cell<SM>(g, c.cx(), c.cy(), c.cz(), V);
    const float r2 = mul_(g.rcp_h, g.rcp_h);
    const float v000 = lo2(V[0]), v100 = hi2(V[0]), v010 = lo2(V[1]), v110 = hi2(V[1]);
    const float v001 = lo2(V[2]), v101 = hi2(V[2]), v011 = lo2(V[3]), v111 = hi2(V[3]);
    float a0 = add_(sub_(sub_(v110, v010), v100), v000), a1 = add_(sub_(sub_(v111, v011), v101), v001);
    float b0 = add_(sub_(sub_(v101, v001), v100), v000), b1 = add_(sub_(sub_(v111, v011), v110), v010);
    float c0 = add_(sub_(sub_(v011, v001), v010), v000), c1 = add_(sub_(sub_(v111, v101), v110), v100);
    hxy0 = a0 * r2; hxys = (a1 - a0) * r2;
    hxz0 = b0 * r2; hxzs = (b1 - b0) * r2;
    hyz0 = c0 * r2; hyzs = (c1 - c0) * r2;
    return a;
  }
  __device__ __forceinline__ void eval(const GridP &g, const CellIdx &c, float &n, f3 &nx, f3 *hm) const {
    eval_grad_ref_x2(g, V, c.wx, c.wy, c.wz, n, nx);
    if (hm) {
      hm->x = fmaf(hxys, c.wz, hxy0);
      hm->y = fmaf(hxzs, c.wy, hxz0);
      hm->z = fmaf(hyzs, c.wx, hyz0);
    }
  }
};

// Trilinear value only (sdf lookups), src/volume.cpp:162-163 under the caller's mask.
template <int AR, class CC = CellCache<AR>>
__device__ __forceinline__ float eval_value(const GridP &g, CC &cc, f3 x) {
  CellIdx c = cell_of(g, x);
  if (cc.differs(c)) cc.refill(g, c);
  float n; f3 nx;
  cc.eval(g, c, n, nx, nullptr);
  return n;
}

__device__ __forceinline__ float dist2_ref(f3 x, f3 t) {
  float dx = sub_(x.x, t.x), dy = sub_(x.y, t.y), dz = sub_(x.z, t.z);
  return fma_(dz, dz, fma_(dy, dy, mul_(dx, dx)));  // squared_norm: mul, fmadd, fmadd
}

// ---------------------------------------------------------------------------
// Forward: Tracer::trace (src/tracer.cpp:35-100), trace_plane (:102-172),
// trace_target (:174-242, per-ray part), trace_sdf (:244-310).
// ---------------------------------------------------------------------------
// Small grids (SM): a batch of <= ~10^5 rays leaves 3-4 warps per scheduler, too few to hide the L2 round trip of the
// gather that follows every cell change -- it is 28 % of the forward kernel's stall samples at 32^3 / 256^2
// (profiles/r12_luneburg32_full.txt).  When the whole grid fits a CTA's shared memory (<= 50 000 voxels) the kernel
// runs as one CTA of up to 512 rays per SM, copies the grid in once and gathers from shared memory.
constexpr int kSmemGridThreads = 512;
extern __shared__ float s_grid[];

template <bool SM>
__device__ __forceinline__ GridP grid_in_smem(GridP g) {
  if (SM) {
    const int G = g.W * g.H * g.D;
    for (int i = threadIdx.x; i < G; i += blockDim.x) s_grid[i] = __ldg(g.data + i);
    __syncthreads();
    g.data = s_grid;
  }
  return g;
}

template <int KIND, int AR, bool SM = false>
__global__ void __launch_bounds__(SM ? kSmemGridThreads : kTraceThreads, SM ? 1 : KIND == K_PLAIN ? kTraceMinBlocks : 8) trace_kernel(TraceArgs a) {
  constexpr bool REF = AR != A_FAST;
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = r < a.N;
  int steps = 0;  // per ray < max_steps < 2^31
  int not_exited = 0;
  const GridP g = grid_in_smem<SM>(a.grid);
  if (valid) {
    const float ds = a.ds;
    f3 x = ld3(a.pos, r, a.N, a.layout), v = ld3(a.vel, r, a.N, a.layout);
    f3 xt = x, vt = v;
    f3 aux0 = {0, 0, 0}, aux1 = {0, 0, 0};
    float best = 0.f;
    if (KIND == K_PLANE) { aux0 = ld3(a.aux0, r, a.N, a.layout); aux1 = ld3(a.aux1, r, a.N, a.layout); }
    if (KIND == K_TARGET) { aux0 = ld3(a.aux0, r, a.N, a.layout); best = dist2_ref(x, aux0); }
    bool inside = inbounds(g, x);  // :61
    bool esc = false, crossed = false;
    CellCache<AR, SM> cc;
    CellCache<AR> sc;  // sdf grid (K_SDF only)
    cc.reset(); sc.reset();
    GridP sg = g;
    sg.data = a.sdf;
    int i = 0;
    if (a.max_steps > 0) do {  // single exit test at the bottom: !esc && i < max_steps
      if (inside) {  // masked gather: outside lanes see n = 0, grad n = 0 and fly straight
        CellIdx c = cell_of(g, x);
        if (cc.differs(c)) cc.refill(g, c);
        float n; f3 nx;
        cc.eval(g, c, n, nx, nullptr);
        float dsn = REF ? mul_(ds, n) : ds * n;
        v.x = fmaf(dsn, nx.x, v.x); v.y = fmaf(dsn, nx.y, v.y); v.z = fmaf(dsn, nx.z, v.z);  // :70
      }
      x.x = fmaf(ds, v.x, x.x); x.y = fmaf(ds, v.y, x.y); x.z = fmaf(ds, v.z, x.z);  // :71
      const bool inb = inbounds(g, x);
      bool cur_inside = inb;  // :73
      if (KIND == K_PLANE) {  // :144-145
        float dd = mul_(sub_(x.x, aux0.x), aux1.x);
        dd = fma_(sub_(x.y, aux0.y), aux1.y, dd);
        dd = fma_(sub_(x.z, aux0.z), aux1.z, dd);
        cur_inside = inb & !(dd > 0.f);
      }
      if (KIND == K_SDF) {  // :287-288, evaluated under the previous `inside`
        cur_inside = inside ? (eval_value<AR>(sg, sc, x) < 0.f) : false;
      }
      if (KIND == K_TARGET) {  // :216-227
        float cur = dist2_ref(x, aux0);
        if (cur < best) { xt = x; vt = v; best = cur; }
      }
      // :74-76.  Nothing can end the ray while it stays inside (cross needs !cur_inside,
      // escaped() implies !inbounds), so the exit logic sits behind one rarely taken branch.
      ++i;
      if (cur_inside & inb) {
        inside = true;  // :86 (fast path: still inside, nothing can end the ray)
      } else {
        crossed = inside & !cur_inside;
        // :79-80 capture: a crossing always ends the ray's loop, so (x, v) at loop exit
        // ARE the captured exit state; the copy is done once after the loop.
        esc = crossed | (!inb && escaped(g, x, v));
        inside = cur_inside;  // :86
      }
    } while (!esc && i < a.max_steps);
    steps = i;
    if (KIND != K_TARGET && crossed) { xt = x; vt = v; }
    if (!esc) {
      not_exited = 1;
      if (KIND == K_PLAIN || KIND == K_PLANE) xt = x;  // :95, :167 (vt keeps the launch value)
    }
    st3(a.xt, r, a.N, a.layout, xt);
    st3(a.vt, r, a.N, a.layout, vt);
    if (KIND == K_PLANE) a.not_escaped[r] = esc ? 0 : 1;  // :171
    if (KIND == K_TARGET) {
      a.dist2[r] = best;
      // state for the batch-coupled tail
      st3(a.tail_x, r, a.N, DRRT_LAYOUT_SOA, x);
      st3(a.tail_v, r, a.N, DRRT_LAYOUT_SOA, v);
      a.tail_iters[r] = i;
      atomicMax(a.tail_max_iters, i);
    }
    if (KIND == K_PLANE) {
      // A ray stopped by its plane INSIDE the grid flies on straight (masked gathers) while the batch's loop
      // runs; if it moves against the plane normal it comes back to the near side, is "inside" again, bends
      // again and its next crossing overwrites the captured state (src/tracer.cpp:144-152 has no
      // `!escaped` on `cross`): state for trace_plane_tail_kernel.  With v.d > 0 the plane distance only grows.
      const float vd = fmaf(v.z, aux1.z, fmaf(v.y, aux1.y, v.x * aux1.x));
      const float tiny = 1e-5f * sqrtf((v.x * v.x + v.y * v.y + v.z * v.z) * (aux1.x * aux1.x + aux1.y * aux1.y + aux1.z * aux1.z));
      const bool more = esc && inbounds(g, x) && !(vd > tiny);
      a.tail_iters[r] = more ? i : 0x7fffffff;
      if (more) {
        st3(a.tail_x, r, a.N, DRRT_LAYOUT_SOA, x);
        st3(a.tail_v, r, a.N, DRRT_LAYOUT_SOA, v);
      }
    }
    if (KIND == K_SDF) {
      // A ray that left the GRID while still inside the sdf (escaped, not crossed, inside) keeps being
      // integrated by the reference -- clamped samples beyond the box -- for as long as the batch's loop
      // runs, and its crossing is captured then (src/tracer.cpp:279-303): state for trace_sdf_tail_kernel.
      const bool more = esc && !crossed && inside;
      a.tail_iters[r] = more ? i : 0x7fffffff;
      if (more) {
        st3(a.tail_x, r, a.N, DRRT_LAYOUT_SOA, x);
        st3(a.tail_v, r, a.N, DRRT_LAYOUT_SOA, v);
      }
    }
  }
  if (KIND == K_SDF || KIND == K_PLANE) {  // the batch's loop runs until its last ray has escaped (or to the cap)
    const int wmax = __reduce_max_sync(0xffffffffu, valid ? steps : 0);
    if ((threadIdx.x & 31) == 0 && wmax) atomicMax(a.tail_max_iters, wmax);
  }
  if (a.stats) {
    const int64_t wsteps = warp_sum_i64((int64_t) steps);
    int ne = __reduce_add_sync(0xffffffffu, not_exited);
    if ((threadIdx.x & 31) == 0) {
      if (wsteps) atomicAdd((unsigned long long *) a.stats, (unsigned long long) wsteps);
      if (ne) atomicAdd((unsigned long long *) a.stats + 1, (unsigned long long) ne);
    }
  }
}

// ---------------------------------------------------------------------------
// Forward trace, two rays per thread (Tracer::trace only, literal arithmetic).
// trace_kernel is bound by instruction issue, and two thirds of its instructions are
// the float32 operations of the reference formulas; with ray A in the low and ray B in
// the high half of every f32x2 each of them -- multiplies and adds alike -- issues once
// for two rays (FMUL2 / FFMA2, see PairK).  Everything that is per ray stays per ray: the
// cell cache tags and refills, the in-bounds tests and the exit, so the results are
// bit-identical to trace_kernel<K_PLAIN, A_REF>.  Lane l of warp w takes rays 64 w + l and
// 64 w + 32 + l (same coalescing as one ray per lane; neighbours in the batch, so the two
// rays of a thread enter and leave the volume within a few steps of each other).
// ---------------------------------------------------------------------------
__device__ __forceinline__ f32x2 with_lo(f32x2 v, float lo) { return pk2(lo, hi2(v)); }
__device__ __forceinline__ f32x2 with_hi(f32x2 v, float hi) { return pk2(lo2(v), hi); }
__device__ __forceinline__ f3 lo3(const f32x2 p[3]) { f3 r = {lo2(p[0]), lo2(p[1]), lo2(p[2])}; return r; }
__device__ __forceinline__ f3 hi3(const f32x2 p[3]) { f3 r = {hi2(p[0]), hi2(p[1]), hi2(p[2])}; return r; }

// A ray of the pair has left the loop: write its outputs (src/tracer.cpp:79-80, :95).
__device__ __forceinline__ void pair_store_exit(const TraceArgs &a, int64_t r, bool crossed, f3 x, f3 v) {
  if (crossed) {
    st3(a.xt, r, a.N, a.layout, x);
    st3(a.vt, r, a.N, a.layout, v);
  } else {  // escaped without ever having been inside: the launch state is returned
    st3(a.xt, r, a.N, a.layout, ld3(a.pos, r, a.N, a.layout));
    st3(a.vt, r, a.N, a.layout, ld3(a.vel, r, a.N, a.layout));
  }
}

__global__ void __launch_bounds__(kTraceThreads, kTracePairMinBlocks) trace_pair_kernel(TraceArgs a) {
  const unsigned lane = threadIdx.x & 31;
  const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t rA = gw * 64 + lane, rB = rA + 32;
  const GridP g = a.grid;
  PairK K;
  K.one = pk2(a.one, a.one);
  K.mone = pk2(-a.one, -a.one);
  const f32x2 DS = pk2(a.ds, a.ds), RH = pk2(g.rcp_h, g.rcp_h);
  bool liveA = rA < a.N, liveB = rB < a.N;
  f32x2 X[3] = {0ull, 0ull, 0ull}, V[3] = {0ull, 0ull, 0ull};
  if (liveA) {
    const f3 x = ld3(a.pos, rA, a.N, a.layout), v = ld3(a.vel, rA, a.N, a.layout);
    X[0] = with_lo(X[0], x.x); X[1] = with_lo(X[1], x.y); X[2] = with_lo(X[2], x.z);
    V[0] = with_lo(V[0], v.x); V[1] = with_lo(V[1], v.y); V[2] = with_lo(V[2], v.z);
  }
  if (liveB) {
    const f3 x = ld3(a.pos, rB, a.N, a.layout), v = ld3(a.vel, rB, a.N, a.layout);
    X[0] = with_hi(X[0], x.x); X[1] = with_hi(X[1], x.y); X[2] = with_hi(X[2], x.z);
    V[0] = with_hi(V[0], v.x); V[1] = with_hi(V[1], v.y); V[2] = with_hi(V[2], v.z);
  }
  bool insA = liveA && inbounds(g, lo3(X)), insB = liveB && inbounds(g, hi3(X));  // :61
  int stepsA = 0, stepsB = 0;
  CellTag tagA, tagB;
  tagA.reset(); tagB.reset();
  f32x2 Vp[8] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull};
  int i = 0;
  while ((liveA | liveB) && i < a.max_steps) {
    if (insA | insB) {  // masked gather: outside rays see n = 0, grad n = 0 and fly straight
      const f32x2 px = mul2(X[0], RH), py = mul2(X[1], RH), pz = mul2(X[2], RH);
      CellIdx cA, cB;
      cA.fx = floorf(lo2(px)); cA.fy = floorf(lo2(py)); cA.fz = floorf(lo2(pz));
      cB.fx = floorf(hi2(px)); cB.fy = floorf(hi2(py)); cB.fz = floorf(hi2(pz));
      const f32x2 wx = K.sub(px, pk2(cA.fx, cB.fx)), wy = K.sub(py, pk2(cA.fy, cB.fy)), wz = K.sub(pz, pk2(cA.fz, cB.fz));
      if (insA && tagA.differs(cA)) {
        tagA.set(cA);
        f32x2 T[4];
        gather_cell(g, cA.cx(), cA.cy(), cA.cz(), T);
#pragma unroll
        for (int j = 0; j < 4; ++j) { Vp[2 * j] = with_lo(Vp[2 * j], lo2(T[j])); Vp[2 * j + 1] = with_lo(Vp[2 * j + 1], hi2(T[j])); }
      }
      if (insB && tagB.differs(cB)) {
        tagB.set(cB);
        f32x2 T[4];
        gather_cell(g, cB.cx(), cB.cy(), cB.cz(), T);
#pragma unroll
        for (int j = 0; j < 4; ++j) { Vp[2 * j] = with_hi(Vp[2 * j], lo2(T[j])); Vp[2 * j + 1] = with_hi(Vp[2 * j + 1], hi2(T[j])); }
      }
      f32x2 n, nx[3];
      eval_grad_ref_pair(K, RH, Vp, wx, wy, wz, n, nx);
      const f32x2 dsn = mul2(DS, n);
      const f32x2 n0 = fma2(dsn, nx[0], V[0]), n1 = fma2(dsn, nx[1], V[1]), n2 = fma2(dsn, nx[2], V[2]);  // :70
      if (insA & insB) {
        V[0] = n0; V[1] = n1; V[2] = n2;
      } else if (insA) {
        V[0] = with_lo(V[0], lo2(n0)); V[1] = with_lo(V[1], lo2(n1)); V[2] = with_lo(V[2], lo2(n2));
      } else {
        V[0] = with_hi(V[0], hi2(n0)); V[1] = with_hi(V[1], hi2(n1)); V[2] = with_hi(V[2], hi2(n2));
      }
    }
    X[0] = fma2(DS, V[0], X[0]); X[1] = fma2(DS, V[1], X[1]); X[2] = fma2(DS, V[2], X[2]);  // :71
    ++i;
    const bool inbA = inbounds(g, lo3(X)), inbB = inbounds(g, hi3(X));
    if (inbA & inbB & liveA & liveB) {
      insA = insB = true;  // :86 (fast path: nothing can end a ray while it stays inside)
    } else {
      if (liveA) {
        if (!inbA) {
          const bool crossed = insA;  // :74
          if (crossed || escaped(g, lo3(X), lo3(V))) {  // :75-80
            pair_store_exit(a, rA, crossed, lo3(X), lo3(V));
            liveA = false; stepsA = i;
          }
        }
        insA = liveA & inbA;
      }
      if (liveB) {
        if (!inbB) {
          const bool crossed = insB;
          if (crossed || escaped(g, hi3(X), hi3(V))) {
            pair_store_exit(a, rB, crossed, hi3(X), hi3(V));
            liveB = false; stepsB = i;
          }
        }
        insB = liveB & inbB;
      }
    }
  }
  int not_exited = 0;
  if (liveA) {  // :89-96: current x, launch v
    st3(a.xt, rA, a.N, a.layout, lo3(X));
    st3(a.vt, rA, a.N, a.layout, ld3(a.vel, rA, a.N, a.layout));
    stepsA = i; ++not_exited;
  }
  if (liveB) {
    st3(a.xt, rB, a.N, a.layout, hi3(X));
    st3(a.vt, rB, a.N, a.layout, ld3(a.vel, rB, a.N, a.layout));
    stepsB = i; ++not_exited;
  }
  if (a.stats) {
    const int64_t wsteps = warp_sum_i64((int64_t) stepsA + (int64_t) stepsB);
    int ne = __reduce_add_sync(0xffffffffu, not_exited);
    if (lane == 0) {
      if (wsteps) atomicAdd((unsigned long long *) a.stats, (unsigned long long) wsteps);
      if (ne) atomicAdd((unsigned long long *) a.stats + 1, (unsigned long long) ne);
    }
  }
}

// trace_target couples rays: the reference keeps testing every ray for a new
// closest approach until the whole batch has escaped (src/tracer.cpp:216-231),
// so an already-escaped ray's result depends on how long the slowest ray keeps
// the loop alive.  After escape a ray is outside the box, its gathers are masked
// to zero and it moves on the straight line x += ds*v; this kernel replays
// exactly those remaining iterations (up to the batch-wide count found by
// trace_kernel<K_TARGET>) and keeps updating the closest sample.
__global__ void __launch_bounds__(kTraceThreads) trace_target_tail_kernel(TraceArgs a) {
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.N) return;
  const int total = *a.tail_max_iters;
  int i = a.tail_iters[r];
  if (i >= total) return;
  const float ds = a.ds;
  f3 x = ld3(a.tail_x, r, a.N, DRRT_LAYOUT_SOA), v = ld3(a.tail_v, r, a.N, DRRT_LAYOUT_SOA);
  f3 t = ld3(a.aux0, r, a.N, a.layout);
  float best = a.dist2[r];
  bool upd = false;
  f3 bx = x;
  for (; i < total; ++i) {
    x.x = fmaf(ds, v.x, x.x); x.y = fmaf(ds, v.y, x.y); x.z = fmaf(ds, v.z, x.z);
    float cur = dist2_ref(x, t);
    if (cur < best) { best = cur; bx = x; upd = true; }
  }
  if (upd) {
    st3(a.xt, r, a.N, a.layout, bx);
    st3(a.vt, r, a.N, a.layout, v);
    a.dist2[r] = best;
  }
}

// trace_sdf couples rays the same way in one corner case: a ray that leaves the grid while the sdf
// still says "inside" (an object reaching the boundary of the volume) has escaped but not crossed; the
// reference's loop goes on until every ray of the batch has escaped, and while it does this ray keeps
// stepping with `inside` set -- refractive index and sdf sampled at clamped positions beyond the box --
// and the first sdf >= 0 sample is captured as its exit state (src/tracer.cpp:279-303).  This kernel
// replays those iterations up to the batch-wide count found by trace_kernel<K_SDF>.
template <int AR>
__global__ void __launch_bounds__(kTraceThreads) trace_sdf_tail_kernel(TraceArgs a) {
  constexpr bool REF = AR != A_FAST;
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.N) return;
  const int total = *a.tail_max_iters;
  int i = a.tail_iters[r];
  if (i >= total) return;
  const GridP g = a.grid;
  GridP sg = g;
  sg.data = a.sdf;
  const float ds = a.ds;
  f3 x = ld3(a.tail_x, r, a.N, DRRT_LAYOUT_SOA), v = ld3(a.tail_v, r, a.N, DRRT_LAYOUT_SOA);
  CellCache<AR> cc, sc;
  cc.reset(); sc.reset();
  for (; i < total; ++i) {
    CellIdx c = cell_of(g, x);
    if (cc.differs(c)) cc.refill(g, c);
    float n; f3 nx;
    cc.eval(g, c, n, nx, nullptr);
    float dsn = REF ? mul_(ds, n) : ds * n;
    v.x = fmaf(dsn, nx.x, v.x); v.y = fmaf(dsn, nx.y, v.y); v.z = fmaf(dsn, nx.z, v.z);
    x.x = fmaf(ds, v.x, x.x); x.y = fmaf(ds, v.y, x.y); x.z = fmaf(ds, v.z, x.z);
    if (!(eval_value<AR>(sg, sc, x) < 0.f)) {  // crossed: captured, and never inside again
      st3(a.xt, r, a.N, a.layout, x);
      st3(a.vt, r, a.N, a.layout, v);
      return;
    }
  }
}

// trace_plane, same coupling (see trace_kernel<K_PLANE>): continue a ray that its plane stopped inside the
// grid, for as long as the batch's loop runs -- straight while beyond the plane, bending again once it is
// back on the near side -- and capture every later crossing like the reference does.
template <int AR>
__global__ void __launch_bounds__(kTraceThreads) trace_plane_tail_kernel(TraceArgs a) {
  constexpr bool REF = AR != A_FAST;
  const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.N) return;
  const int total = *a.tail_max_iters;
  int i = a.tail_iters[r];
  if (i >= total) return;
  const GridP g = a.grid;
  const float ds = a.ds;
  f3 x = ld3(a.tail_x, r, a.N, DRRT_LAYOUT_SOA), v = ld3(a.tail_v, r, a.N, DRRT_LAYOUT_SOA);
  const f3 o = ld3(a.aux0, r, a.N, a.layout), d = ld3(a.aux1, r, a.N, a.layout);
  CellCache<AR> cc;
  cc.reset();
  bool inside = false;  // it has just crossed
  for (; i < total; ++i) {
    if (inside) {
      CellIdx c = cell_of(g, x);
      if (cc.differs(c)) cc.refill(g, c);
      float n; f3 nx;
      cc.eval(g, c, n, nx, nullptr);
      float dsn = REF ? mul_(ds, n) : ds * n;
      v.x = fmaf(dsn, nx.x, v.x); v.y = fmaf(dsn, nx.y, v.y); v.z = fmaf(dsn, nx.z, v.z);
    }
    x.x = fmaf(ds, v.x, x.x); x.y = fmaf(ds, v.y, x.y); x.z = fmaf(ds, v.z, x.z);
    const bool inb = inbounds(g, x);
    float dd = mul_(sub_(x.x, o.x), d.x);
    dd = fma_(sub_(x.y, o.y), d.y, dd);
    dd = fma_(sub_(x.z, o.z), d.z, dd);
    const bool cur_inside = inb & !(dd > 0.f);
    if (inside & !cur_inside) {
      st3(a.xt, r, a.N, a.layout, x);
      st3(a.vt, r, a.N, a.layout, v);
    }
    inside = cur_inside;
    if (!inb) return;  // out of the (convex) grid on a straight line: never inside again
  }
}

// ---------------------------------------------------------------------------
// Adjoint: Tracer::backtrace (src/tracer.cpp:384-440), backtrace_sdf (:443-509).
// The ray is re-integrated backwards from its stored exit state; no tape.
//
// Gradient splat.
//   A_REF: the reference's 16 terms, one atomic each (src/volume.cpp:217-243).
//   A_EXACT / A_FAST: the two terms per corner are fused (8 values) and those 8
//   values are ACCUMULATED IN REGISTERS while the ray stays in one cell; they are
//   flushed when it moves on (ds = h/2: ~2 steps per cell).  With AGG the flush is
//   warp-aggregated: lanes that flush the same cell in the same step (neighbouring
//   sensor pixels share cells -- 8 rays per cell in the 2048^2 / 256^3 workload)
//   are summed with a segmented shuffle reduction and one lane per run issues the
//   8 RED.ADD.F32, so same-address conflicts inside a warp (which the LSU
//   serialises into one L2 sector operation per lane, profiles/r01_adj_full.txt:
//   9.7 sectors per RED request) disappear.
// ---------------------------------------------------------------------------
constexpr unsigned kFull = 0xffffffffu;
constexpr int kAggMin = 4;   // fewer flushing lanes than this: plain atomics

template <class Accum>
__device__ __forceinline__ void flush_direct(const Accum &out, const CellAddr &ca, const float acc[8]) {
#pragma unroll
  for (int k = 0; k < 8; ++k) out.add(ca.at(k), acc[k]);
}

// Face carry.  When a ray steps into a face-adjacent cell, four of the eight corner
// accumulators belong to the face the two cells share: they stay in registers (moved to
// the new cell's slots) and only the four corners left behind are flushed -- half the
// atomics per cell visit along the dominant direction of travel.  Adjacency is read off
// the addresses themselves (same strides, base shifted by exactly one stride), so clamped
// cells at and beyond the grid boundary simply take the full flush.
// Returns true when accumulators were carried (P still holds unflushed data).
template <class Accum>
__device__ __forceinline__ bool flush_carry(const Accum &out, const CellAddr &ca, const CellAddr &na, f32x2 P[4]) {
  const bool same = (na.dx == ca.dx) & (na.dy == ca.dy) & (na.dz == ca.dz);
  const int d = na.base - ca.base;
  if (same && ca.dz != 0 && d == -ca.dz) {        // z-: old z0 face becomes the new z1 face
    out.add(ca.at(4), lo2(P[2])); out.add(ca.at(5), hi2(P[2])); out.add(ca.at(6), lo2(P[3])); out.add(ca.at(7), hi2(P[3]));
    P[2] = P[0]; P[3] = P[1]; P[0] = P[1] = 0ull;
  } else if (same && ca.dz != 0 && d == ca.dz) {  // z+
    out.add(ca.at(0), lo2(P[0])); out.add(ca.at(1), hi2(P[0])); out.add(ca.at(2), lo2(P[1])); out.add(ca.at(3), hi2(P[1]));
    P[0] = P[2]; P[1] = P[3]; P[2] = P[3] = 0ull;
  } else if (same && ca.dy != 0 && d == -ca.dy) {  // y-
    out.add(ca.at(2), lo2(P[1])); out.add(ca.at(3), hi2(P[1])); out.add(ca.at(6), lo2(P[3])); out.add(ca.at(7), hi2(P[3]));
    P[1] = P[0]; P[3] = P[2]; P[0] = P[2] = 0ull;
  } else if (same && ca.dy != 0 && d == ca.dy) {   // y+
    out.add(ca.at(0), lo2(P[0])); out.add(ca.at(1), hi2(P[0])); out.add(ca.at(4), lo2(P[2])); out.add(ca.at(5), hi2(P[2]));
    P[0] = P[1]; P[2] = P[3]; P[1] = P[3] = 0ull;
  } else if (same && ca.dx != 0 && d == -ca.dx) {  // x-: x0 corners become x1 corners
#pragma unroll
    for (int j = 0; j < 4; ++j) { out.add(ca.at(2 * j + 1), hi2(P[j])); P[j] = pk2(0.f, lo2(P[j])); }
  } else if (same && ca.dx != 0 && d == ca.dx) {   // x+
#pragma unroll
    for (int j = 0; j < 4; ++j) { out.add(ca.at(2 * j), lo2(P[j])); P[j] = pk2(hi2(P[j]), 0.f); }
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) { out.add(ca.at(2 * j), lo2(P[j])); out.add(ca.at(2 * j + 1), hi2(P[j])); }
    P[0] = P[1] = P[2] = P[3] = 0ull;
    return false;
  }
  return true;
}

__device__ __forceinline__ void unpack8(const f32x2 P[4], float acc[8]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) { acc[2 * j] = lo2(P[j]); acc[2 * j + 1] = hi2(P[j]); }
}

// Segmented shuffle reduction over runs of consecutive lanes that sit in the same cell (lanes
// that do not flush this step carry zeros, so they do not cut a run), runs cut at multiples of RUN
// lanes: log2(RUN) rounds of 8 SHFL + 8 FADD.  Returns true on the lane that now holds the sum of
// a run with at least one flushing lane (`fm` = ballot of the flushing lanes).
template <int RUN>
__device__ __forceinline__ bool reduce_runs(float acc[8], bool first, unsigned lane, unsigned fm) {
  const bool head = first || (lane & (RUN - 1)) == 0;
  const unsigned stops = __ballot_sync(kFull, head);
  const unsigned above = stops & ~((2u << lane) - 1u);
  const int end = above ? __ffs(above) - 1 : 32;
#pragma unroll
  for (int off = 1; off < RUN; off <<= 1) {
    const bool take = (int) (lane + off) < end;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = __shfl_down_sync(kFull, acc[k], off);
      if (take) acc[k] += t;
    }
  }
  const unsigned span = (end == 32 ? kFull : (1u << end) - 1u) & ~((1u << lane) - 1u);
  return head && (fm & span) != 0;
}

// Per-warp aggregation state: the run length in use and a flush-event counter.
struct AggState {
  int run, events;
};

// Warp-uniform: every lane of the warp calls this with its own `fl`; `fm` = ballot of `fl` (non-zero).
// `policy`: 2, 4, 8, 32 = fixed maximum run length; 0 = adaptive.
// Returns true when the flushing lanes' accumulators were summed and written here, false when the
// caller should flush each lane on its own (the plain path, with the face carry).
//
// Measured (tools/gpu_tune_agg.py, profiles/r03_tune_agg.log): a same-address
// conflict inside one RED costs less than a shuffle round until ~4 lanes collide,
// so the best run length is about (lanes per cell) / 4: none at 4 rays per cell,
// 2 at 8, 4 at 16, 8 at 32.  The adaptive policy re-measures the average number of
// consecutive lanes per cell every 8th flush of the warp and picks accordingly.
// The exception is a warp whose flushing lanes ALL sit in one cell (a focused beam): every lane-op on
// one address is serialised chip-wide at ~0.7 G/s however it is packed into requests
// (profiles/r09_l2_probe_hot.log), so there the whole warp is summed first (run 32).
template <class Accum>
__device__ __forceinline__ bool flush_aggregated(const Accum &out, const CellAddr &ca, const f32x2 P[4],
                                                 bool fl, unsigned fm, unsigned lane, int policy, int2 t, AggState &st) {
  const int nfl = __popc(fm);
  const bool probe = policy == 0 && (st.events++ & 7) == 0;
  if (policy != 0) st.run = policy;
  if ((st.run == 1 && !probe) || nfl < kAggMin) return false;
  // two lanes address the same 8 voxels iff corner 0 and corner 7 coincide
  const int k0 = ca.base;
  const int k7 = ca.base + ca.dx + ca.dy + ca.dz;
  const int p0 = __shfl_up_sync(kFull, k0, 1), p7 = __shfl_up_sync(kFull, k7, 1);
  const bool first = lane == 0 || p0 != k0 || p7 != k7;  // first lane of a same-cell run
  if (probe) {
    const int nruns = __popc(__ballot_sync(kFull, first));
    st.run = (nfl >= t.y * nruns) ? 32 : (nfl >= t.x * nruns) ? 8 : 1;
    if (st.run == 1) return false;
  }
  float acc[8];
  unpack8(P, acc);
  if (!fl) {
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  }
  bool head;
  if (st.run == 32) head = reduce_runs<32>(acc, first, lane, fm);
  else if (st.run == 8) head = reduce_runs<8>(acc, first, lane, fm);
  else if (st.run == 4) head = reduce_runs<4>(acc, first, lane, fm);
  else head = reduce_runs<2>(acc, first, lane, fm);
  if (head) flush_direct(out, ca, acc);
  return true;
}

// ---------------------------------------------------------------------------
// Ray interleaving.  Neighbouring sensor pixels share grid cells (8 rays per cell
// along x at 2048^2 rays / 256^3 voxels), so with ray i on lane i the lanes of a
// warp splat into the same few addresses and every RED is replayed once per
// colliding lane.  Instead of repairing that with shuffles, the adjoint assigns
// rays to lanes with a stride: with interleave IL = 2^sh, IL consecutive warps
// share a tile of 32*IL consecutive rays and lane l of warp w takes ray
// tile*32*IL + l*IL + w.  The lanes of one warp then sit in 32 DIFFERENT,
// x-adjacent cells: gathers and REDs of a warp cover consecutive addresses (full
// sectors, no same-address collisions), at no instruction cost.  IL is measured
// on the device from the batch itself (average number of consecutive rays per
// cell in a sample), so no host synchronisation is needed.
// ---------------------------------------------------------------------------
constexpr int kInterleaveMax = 32;

__global__ void __launch_bounds__(256) estimate_interleave_kernel(GridP g, const float *pos, int64_t N,
                                                                  int layout, int forced_log2, int *out) {
  if (forced_log2 >= 0) {
    if (threadIdx.x == 0) *out = forced_log2;
    return;
  }
  __shared__ int s_breaks;
  if (threadIdx.x == 0) s_breaks = 0;
  __syncthreads();
  const int S = (int) (N < 8192 ? N : 8192);
  const int64_t start = (N - S) / 2;
  int breaks = 0;
  for (int i = threadIdx.x + 1; i < S; i += blockDim.x) {
    CellIdx a = cell_of(g, ld3(pos, start + i - 1, N, layout)), b = cell_of(g, ld3(pos, start + i, N, layout));
    breaks += (a.fx != b.fx) | (a.fy != b.fy) | (a.fz != b.fz);
  }
  breaks = __reduce_add_sync(kFull, breaks);
  if ((threadIdx.x & 31) == 0) atomicAdd(&s_breaks, breaks);
  __syncthreads();
  if (threadIdx.x == 0) {
    const float m = (float) S / (float) (s_breaks + 1);  // average consecutive rays per cell
    int sh = 0;
    while ((2 << sh) <= kInterleaveMax && (float) (2 << sh) <= 1.1f * m) ++sh;
    *out = sh;
  }
}

// RG ("ray gradients"): also return dL/dpos and dL/dvel of the LAUNCH state.
// With a_k = dL/dx_k, b_k = dL/dv_k the recursion below holds lambda = a_k and
// mu = b_{k-1} after sample x_k has been processed; before x_k the forward ray flew
// straight (gathers masked to zero outside the box, src/tracer.cpp:66-71), where
// a is constant and b_{j-1} = b_j + ds a_j.  So dL/dpos = lambda and
// dL/dvel = mu + (k - 1) ds lambda, with k the number of straight steps between the
// launch position and the last processed sample (k = 0 when `pos0` is not given, which
// is exact for rays launched on the first in-bounds sample).
template <bool SDF, int AR, class Accum, bool AGG, bool RG = false, bool SM = false>
__global__ void __launch_bounds__(SM ? kSmemGridThreads : kAdjThreads, SM ? 1 : (SDF || AR == A_REF || RG) ? 5 : AGG ? kAdjMinBlocks - 1 : kAdjMinBlocks) backtrace_kernel(AdjArgs a, Accum out_in) {
  constexpr bool REF = AR == A_REF;     // literal 16-atomic splat and lambda/mu order
  constexpr bool TRAJ_REF = AR != A_FAST;
  Accum out = out_in;
  out.prepare();
  if (a.replicas > 1) out.select((int) (blockIdx.x % (unsigned) a.replicas), a.replica_base, a.replica_stride);
  const unsigned lane = threadIdx.x & 31;
  int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (a.interleave_log2) {  // see "Ray interleaving" above
    const int sh = *a.interleave_log2;
    const int64_t gw = r >> 5;
    r = ((((gw >> sh) << 5) + lane) << sh) | (gw & ((1 << sh) - 1));
  }
  const bool valid = r < a.N;
  int steps = 0;  // per ray < max_steps < 2^31
  const GridP g = grid_in_smem<SM>(a.grid);
  const float ds = a.ds;
  f3 x = {0, 0, 0}, v = {0, 0, 0}, la = {0, 0, 0}, mu = {0, 0, 0};
  bool active = false;
  CellCache<AR, SM> cc;
  CellCache<AR> sc;
  cc.reset(); sc.reset();
  GridP sg = g;
  sg.data = a.sdf;
  bool outside = false;
  f3 xl = {0, 0, 0}, vl = {0, 0, 0};  // RG: (x_k, v_k) of the last sample that updated lambda/mu
  // RG: a ray launched INSIDE the region the adjoint integrates through (trace_sdf's rays start inside the
  // object, src/tracer.cpp:271-277) is re-integrated past its launch point; the backward samples retrace the
  // forward ones, so the sample within half a step of pos0 is x_0 and (lambda, mu) after it are dL/dx_0, dL/dv_0.
  f3 p0 = {0, 0, 0}, lcap = {0, 0, 0}, mcap = {0, 0, 0};
  bool got = false;
  if (valid) {
    x = ld3(a.xt, r, a.N, a.layout); v = ld3(a.vt, r, a.N, a.layout);
    xl = x; vl = v;
    if (RG && a.pos0) p0 = ld3(a.pos0, r, a.N, a.layout);
    la = ld3(a.dx, r, a.N, a.layout); mu = ld3(a.dv, r, a.N, a.layout);
    mu.x = add_(mu.x, mul_(ds, la.x)); mu.y = add_(mu.y, mul_(ds, la.y)); mu.z = add_(mu.z, mul_(ds, la.z));  // :410
    f3 nv = {-v.x, -v.y, -v.z};
    active = !escaped(g, x, nv);  // :413-414
    if (SDF) {  // :476-477
      float dist = active ? eval_value<AR>(sg, sc, x) : 0.f;
      outside = dist >= 0.f;
    }
  }
  CellAddr ca = {0, 0, 0, 0};
  f32x2 P[4] = {0ull, 0ull, 0ull, 0ull};  // accumulator pairs (acc[2j], acc[2j+1])
  AggState agg = {1, 0};
  bool have = false;
  for (int i = 0; i < a.max_steps; ++i) {
    if (AGG) {
      if (!__any_sync(kFull, active)) break;
    } else if (!active) {
      break;
    }
    CellIdx c = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bool chg = false;
    if (active) {
      x.x = fmaf(-ds, v.x, x.x); x.y = fmaf(-ds, v.y, x.y); x.z = fmaf(-ds, v.z, x.z);  // :420
      c = cell_of(g, x);
      chg = cc.differs(c);
    }
    // the new cell's gather is issued BEFORE the old cell's accumulators are flushed, so the
    // L2 latency of the face the ray moves towards overlaps the address arithmetic and REDs
    CellAddr nca = ca;
    if (chg) nca = cc.refill(g, c);
    if (!REF) {
      const bool fl = have && (chg || !active);
      bool summed = false;  // AGG: the warp summed and wrote this flush
      if (AGG) {
        const unsigned fm = __ballot_sync(kFull, fl);
        if (fm) summed = flush_aggregated(out, ca, P, fl, fm, lane, a.agg_policy, make_int2(a.agg_t8, a.agg_twarp), agg);
      }
      if (fl) {
        if (summed) {
          P[0] = P[1] = P[2] = P[3] = 0ull;
          have = false;
        } else if (a.face_carry && active) {  // active && fl implies chg: the ray moves on to cell nca
          have = flush_carry(out, ca, nca, P);
        } else {
          float acc[8];
          unpack8(P, acc);
          flush_direct(out, ca, acc);
          P[0] = P[1] = P[2] = P[3] = 0ull;
          have = false;
        }
      }
    }
    ca = nca;
    if (active) {
      float n; f3 nx, hm;
      cc.eval(g, c, n, nx, &hm);  // :421-422 (same 8 voxels serve grad and Hessian)
      float mdsn = TRAJ_REF ? mul_(-ds, n) : -ds * n;
      v.x = fmaf(mdsn, nx.x, v.x); v.y = fmaf(mdsn, nx.y, v.y); v.z = fmaf(mdsn, nx.z, v.z);  // :423
      ++steps;
      f3 nv = {-v.x, -v.y, -v.z};
      active = inbounds(g, x) || !escaped(g, x, nv);  // :425
      if (SDF) {  // :488-497
        float dist = eval_value<AR>(sg, sc, x);
        bool now_out = dist >= 0.f;
        active = active & !(!outside & now_out);
        outside = now_out;
      }
      if (active) {  // :426 per ray
        if (REF) {
          float dn = fma_(mu.z, nx.z, fma_(mu.y, nx.y, mul_(mu.x, nx.x)));  // :430
          f3 dnx = {mul_(mul_(n, mu.x), ds), mul_(mul_(n, mu.y), ds), mul_(mul_(n, mu.z), ds)};  // :431-432
          float t[16];
          splat_terms_ref(mul_(dn, ds), dnx, c.wx, c.wy, c.wz, t);
#pragma unroll
          for (int k = 0; k < 16; ++k) out.add(ca.at(k & 7), t[k]);
          // :434  la += ds * (dn * nx + (n * Hess) * mu), matrix.h:170-178 column sums
          float nhxy = mul_(n, hm.x), nhxz = mul_(n, hm.y), nhyz = mul_(n, hm.z);
          float h0 = fma_(nhxz, mu.z, mul_(nhxy, mu.y));
          float h1 = fma_(nhyz, mu.z, mul_(nhxy, mu.x));
          float h2 = fma_(nhyz, mu.y, mul_(nhxz, mu.x));
          la.x = add_(la.x, mul_(ds, add_(mul_(dn, nx.x), h0)));
          la.y = add_(la.y, mul_(ds, add_(mul_(dn, nx.y), h1)));
          la.z = add_(la.z, mul_(ds, add_(mul_(dn, nx.z), h2)));
          mu.x = add_(mu.x, mul_(ds, la.x)); mu.y = add_(mu.y, mul_(ds, la.y)); mu.z = add_(mu.z, mul_(ds, la.z));  // :435
        } else {
          float dn = fmaf(mu.z, nx.z, fmaf(mu.y, nx.y, mu.x * nx.x));
          float nds = n * ds;
          f3 dnx = {nds * mu.x, nds * mu.y, nds * mu.z};
          splat_accumulate_x2(dn * ds, dnx, c.wx, c.wy, c.wz, P);
          have = true;
          float nhxy = n * hm.x, nhxz = n * hm.y, nhyz = n * hm.z;
          float h0 = fmaf(nhxz, mu.z, nhxy * mu.y);
          float h1 = fmaf(nhyz, mu.z, nhxy * mu.x);
          float h2 = fmaf(nhyz, mu.y, nhxz * mu.x);
          la.x = fmaf(ds, fmaf(dn, nx.x, h0), la.x);
          la.y = fmaf(ds, fmaf(dn, nx.y, h1), la.y);
          la.z = fmaf(ds, fmaf(dn, nx.z, h2), la.z);
          mu.x = fmaf(ds, la.x, mu.x); mu.y = fmaf(ds, la.y, mu.y); mu.z = fmaf(ds, la.z, mu.z);
        }
        if (RG) {
          xl = x; vl = v;
          if (a.pos0 && !got) {
            const float ex = x.x - p0.x, ey = x.y - p0.y, ez = x.z - p0.z;
            if (ex * ex + ey * ey + ez * ez <= 0.25f * ds * ds * (v.x * v.x + v.y * v.y + v.z * v.z)) {
              got = true; lcap = la; mcap = mu;
            }
          }
        }
      }
    }
  }
  if (!REF) {
    bool summed = false;
    if (AGG) {
      const unsigned fm = __ballot_sync(kFull, have);
      if (fm) summed = flush_aggregated(out, ca, P, have, fm, lane, a.agg_policy, make_int2(a.agg_t8, a.agg_twarp), agg);
    }
    if (have && !summed) {
      float acc[8];
      unpack8(P, acc);
      flush_direct(out, ca, acc);
    }
  }
  if (RG && valid) {
    float k = 0.f;
    if (a.pos0) {
      const f3 x0 = p0;
      const float num = (xl.x - x0.x) * vl.x + (xl.y - x0.y) * vl.y + (xl.z - x0.z) * vl.z;
      const float den = ds * (vl.x * vl.x + vl.y * vl.y + vl.z * vl.z);
      k = den > 0.f ? rintf(num / den) : 0.f;
    }
    float t = (k - 1.f) * ds;
    if (got) { la = lcap; mu = mcap; t = -ds; }
    st3(a.dpos, r, a.N, a.layout, la);
    f3 dv0 = {fmaf(t, la.x, mu.x), fmaf(t, la.y, mu.y), fmaf(t, la.z, mu.z)};
    st3(a.dvel, r, a.N, a.layout, dv0);
  }
  if (a.stats) {
    const int64_t wsteps = warp_sum_i64((int64_t) steps);
    if (lane == 0 && wsteps) atomicAdd((unsigned long long *) a.stats, (unsigned long long) wsteps);
  }
}

// ---------------------------------------------------------------------------
// Launchers
// ---------------------------------------------------------------------------
static inline unsigned blocks_for(int64_t n, int threads) { return (unsigned) ((n + threads - 1) / threads); }

// shared-memory grid variants: rays spread over the 148 SMs, one CTA each
constexpr int kSmemGridBytes = 200 * 1024;
static inline int smem_grid_threads(int64_t n) {
  int64_t t = ((n + 147) / 148 + 31) / 32 * 32;
  return (int) (t < 64 ? 64 : t > kSmemGridThreads ? kSmemGridThreads : t);
}

template <int KIND>
static void launch_trace_kind(int arith, const TraceArgs &a, cudaStream_t s) {
  const int nt = kTraceThreads;  // 32 / 64 / 128 threads per CTA time the same, also below one wave (profiles/r09_cta_size.log)
  unsigned nb = blocks_for(a.N, nt);
  // the forward pass is all trajectory: A_EXACT and A_REF are the same kernel
  if (KIND == K_PLAIN && arith != A_FAST && a.pair) {
    trace_pair_kernel<<<blocks_for(((a.N + 63) / 64) * 32, nt), nt, 0, s>>>(a);
    return;
  }
  if (KIND == K_PLAIN && arith != A_FAST && a.smem_grid) {  // small grid, small batch: one CTA per SM, grid in shared memory
    const int t = smem_grid_threads(a.N);
    const size_t bytes = sizeof(float) * (size_t) a.grid.W * a.grid.H * a.grid.D;
    static const bool optin = [] {
      cudaFuncSetAttribute(trace_kernel<K_PLAIN, A_REF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemGridBytes);
      return true;
    }();
    (void) optin;
    trace_kernel<K_PLAIN, A_REF, true><<<blocks_for(a.N, t), t, bytes, s>>>(a);
    return;
  }
  if (arith == A_FAST) trace_kernel<KIND, A_FAST><<<nb, nt, 0, s>>>(a);
  else trace_kernel<KIND, A_REF><<<nb, nt, 0, s>>>(a);
}

int launch_trace(int kind, int arith, const TraceArgs &a, cudaStream_t s) {
  if (a.N <= 0) return 0;
  switch (kind) {
    case K_PLAIN: launch_trace_kind<K_PLAIN>(arith, a, s); return 1;
    case K_PLANE:
      launch_trace_kind<K_PLANE>(arith, a, s);
      if (arith == A_FAST) trace_plane_tail_kernel<A_FAST><<<blocks_for(a.N, kTraceThreads), kTraceThreads, 0, s>>>(a);
      else trace_plane_tail_kernel<A_REF><<<blocks_for(a.N, kTraceThreads), kTraceThreads, 0, s>>>(a);
      return 2;
    case K_SDF:
      launch_trace_kind<K_SDF>(arith, a, s);
      if (arith == A_FAST) trace_sdf_tail_kernel<A_FAST><<<blocks_for(a.N, kTraceThreads), kTraceThreads, 0, s>>>(a);
      else trace_sdf_tail_kernel<A_REF><<<blocks_for(a.N, kTraceThreads), kTraceThreads, 0, s>>>(a);
      return 2;
    case K_TARGET:
      launch_trace_kind<K_TARGET>(arith, a, s);
      trace_target_tail_kernel<<<blocks_for(a.N, kTraceThreads), kTraceThreads, 0, s>>>(a);
      return 2;
  }
  return 0;
}

template <bool SDF, class Accum, bool AGG>
static void launch_backtrace_acc(int arith, const AdjArgs &a, Accum acc, cudaStream_t s) {
  // with interleaving the thread count is rounded up to whole tiles of the largest interleave
  const int64_t tile = 32 * kInterleaveMax;
  const int64_t nthreads = a.interleave_log2 ? ((a.N + tile - 1) / tile) * tile : a.N;
  const int nt = kAdjThreads;
  unsigned nb = blocks_for(nthreads, nt);
  if (a.dpos) {  // ray gradients: no warp aggregation variant
    if (arith == A_REF) backtrace_kernel<SDF, A_REF, Accum, false, true><<<nb, nt, 0, s>>>(a, acc);
    else if (arith == A_FAST) backtrace_kernel<SDF, A_FAST, Accum, false, true><<<nb, nt, 0, s>>>(a, acc);
    else backtrace_kernel<SDF, A_EXACT, Accum, false, true><<<nb, nt, 0, s>>>(a, acc);
    return;
  }
  if (!SDF && !AGG && arith == A_EXACT && a.smem_grid && !a.interleave_log2) {  // see trace_kernel<..., SM>
    const int t = smem_grid_threads(a.N);
    const size_t bytes = sizeof(float) * (size_t) a.grid.W * a.grid.H * a.grid.D;
    static const bool optin = [] {
      cudaFuncSetAttribute(backtrace_kernel<false, A_EXACT, Accum, false, false, true>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemGridBytes);
      return true;
    }();
    (void) optin;
    backtrace_kernel<false, A_EXACT, Accum, false, false, true><<<blocks_for(a.N, t), t, bytes, s>>>(a, acc);
    return;
  }
  if (arith == A_REF) backtrace_kernel<SDF, A_REF, Accum, false><<<nb, nt, 0, s>>>(a, acc);
  else if (arith == A_FAST) backtrace_kernel<SDF, A_FAST, Accum, AGG><<<nb, nt, 0, s>>>(a, acc);
  else backtrace_kernel<SDF, A_EXACT, Accum, AGG><<<nb, nt, 0, s>>>(a, acc);
}

int launch_backtrace(bool sdf, int arith, bool warp_agg, int interleave, const AdjArgs &a_in, float *grad,
                     unsigned long long *fixed, const unsigned *fixed_max_bits, cudaStream_t s) {
  if (a_in.N <= 0) return 0;
  AdjArgs a = a_in;
  int launched = 1;
  a.interleave_log2 = nullptr;
  // interleave: 0 = off, -1 = measured on the device, 2^k = forced
  int *slot = nullptr;  // per-call, stream-ordered: nothing is shared between calls or streams
  if (interleave != 0 && a.N >= 4096 && cudaMallocAsync((void **) &slot, sizeof(int), s) == cudaSuccess) {
    int forced = -1;
    if (interleave > 0) { forced = 0; while ((2 << forced) <= interleave && (2 << forced) <= kInterleaveMax) ++forced; }
    estimate_interleave_kernel<<<1, 256, 0, s>>>(a.grid, a.xt, a.N, a.layout, forced, slot);
    a.interleave_log2 = slot;
    ++launched;
  }
  if (fixed) {
    // integer accumulation must not depend on how lanes are grouped: no float pre-summation
    AccumFixed acc{fixed, fixed_max_bits, 1.0f};
    if (sdf) launch_backtrace_acc<true, AccumFixed, false>(arith, a, acc, s);
    else launch_backtrace_acc<false, AccumFixed, false>(arith, a, acc, s);
  } else {
    AccumFloat acc{grad};
    if (sdf) {
      if (warp_agg) launch_backtrace_acc<true, AccumFloat, true>(arith, a, acc, s);
      else launch_backtrace_acc<true, AccumFloat, false>(arith, a, acc, s);
    } else {
      if (warp_agg) launch_backtrace_acc<false, AccumFloat, true>(arith, a, acc, s);
      else launch_backtrace_acc<false, AccumFloat, false>(arith, a, acc, s);
    }
  }
  if (slot) cudaFreeAsync(slot, s);
  return launched;
}

}  // namespace drrt
