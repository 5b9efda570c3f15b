// Device-side building blocks of the eikonal tracer (sm_100a).
//
// Semantics follow the reference's volume / cylinder_volume classes
// (reference src/volume.cpp, src/cylinder_volume.cpp); file:line citations sit
// on each function.  Two arithmetic flavours, selected at compile time:
//
//   REF = true   the reference's exact operation order, every multiply/add a
//                separately rounded IEEE operation (__fmul_rn/__fadd_rn are
//                never contracted by nvcc) and FMA only where the reference
//                writes fmadd.  Bit-identical to a float32 CPU evaluation.
//   REF = false  the trilinear cell as a polynomial in the fractional
//                coordinates, a0 + ax*wx + ... + axyz*wx*wy*wz, whose 8
//                coefficients are cached in registers while the ray stays in
//                the cell: 11 FMAs give n and grad n, 1 more the Hessian.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace drrt {

struct GridP {
  const float *data;
  int W, H, D;
  float h, rcp_h;
  float hix, hiy, hiz;  // (res_k - 1) * h, src/volume.cpp:252-254
};

struct CableP {
  const float *data;
  int res;
  float radius, length;
  float h;   // radius / (res - 1), src/cylinder_volume.cpp:42
  float r2;  // radius * radius
};

struct f3 {
  float x, y, z;
};

__device__ __forceinline__ f3 ld3(const float *p, int64_t i, int64_t n, int layout) {
  f3 r;
  if (layout == 3) {
    r.x = p[3 * i]; r.y = p[3 * i + 1]; r.z = p[3 * i + 2];
  } else {
    r.x = p[i]; r.y = p[n + i]; r.z = p[2 * n + i];
  }
  return r;
}

__device__ __forceinline__ void st3(float *p, int64_t i, int64_t n, int layout, f3 v) {
  if (layout == 3) {
    p[3 * i] = v.x; p[3 * i + 1] = v.y; p[3 * i + 2] = v.z;
  } else {
    p[i] = v.x; p[n + i] = v.y; p[2 * n + i] = v.z;
  }
}

// strictly rounded helpers (no contraction)
__device__ __forceinline__ float mul_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// volume::inbounds, src/volume.cpp:246-256
__device__ __forceinline__ bool inbounds(const GridP &g, f3 p) {
  return (p.x >= 0.f) & (p.y >= 0.f) & (p.z >= 0.f) & (p.x < g.hix) & (p.y < g.hiy) & (p.z < g.hiz);
}

// volume::escaped, src/volume.cpp:258-271
__device__ __forceinline__ bool escaped(const GridP &g, f3 p, f3 v) {
  bool ex = ((p.x < 0.f) & (v.x < 0.f)) | ((p.x >= g.hix) & (v.x > 0.f));
  bool ey = ((p.y < 0.f) & (v.y < 0.f)) | ((p.y >= g.hiy) & (v.y > 0.f));
  bool ez = ((p.z < 0.f) & (v.z < 0.f)) | ((p.z >= g.hiz) & (v.z > 0.f));
  return ex | ey | ez;
}

// Cell lookup, src/volume.cpp:128-141: fractional weights from the UNCLAMPED
// floor, both corner indices clamped to [0, res-1].
struct CellIdx {
  float fx, fy, fz;   // unclamped floor(p / h), kept as floats: the per-step "same cell?" test
                      // compares these, the integer conversion happens only on a refill
  float wx, wy, wz;   // w0
  __device__ __forceinline__ int cx() const { return (int) fx; }
  __device__ __forceinline__ int cy() const { return (int) fy; }
  __device__ __forceinline__ int cz() const { return (int) fz; }
};

__device__ __forceinline__ CellIdx cell_of(const GridP &g, f3 p) {
  CellIdx c;
  float px = mul_(p.x, g.rcp_h), py = mul_(p.y, g.rcp_h), pz = mul_(p.z, g.rcp_h);
  c.fx = floorf(px); c.fy = floorf(py); c.fz = floorf(pz);
  c.wx = sub_(px, c.fx); c.wy = sub_(py, c.fy); c.wz = sub_(pz, c.fz);
  return c;
}

// The 8 flat corner indices in the reference's order 000 100 010 110 001 101 011
// 111 (first digit = x), src/volume.cpp:134-141.
__device__ __forceinline__ void corner_indices(const GridP &g, int cx, int cy, int cz, int idx[8]) {
  int x0 = max(min(cx, g.W - 1), 0), x1 = max(min(cx + 1, g.W - 1), 0);
  int y0 = max(min(cy, g.H - 1), 0), y1 = max(min(cy + 1, g.H - 1), 0);
  int z0 = max(min(cz, g.D - 1), 0), z1 = max(min(cz + 1, g.D - 1), 0);
  int r00 = (z0 * g.H + y0) * g.W, r10 = (z0 * g.H + y1) * g.W;
  int r01 = (z1 * g.H + y0) * g.W, r11 = (z1 * g.H + y1) * g.W;
  idx[0] = r00 + x0; idx[1] = r00 + x1; idx[2] = r10 + x0; idx[3] = r10 + x1;
  idx[4] = r01 + x0; idx[5] = r01 + x1; idx[6] = r11 + x0; idx[7] = r11 + x1;
}

// Compact form of the same 8 addresses: base = idx[0] and the three (clamped)
// strides, idx[k] = base + (k&1 ? dx : 0) + (k&2 ? dy : 0) + (k&4 ? dz : 0).
struct CellAddr {
  int base, dx, dy, dz;
  __device__ __forceinline__ int at(int k) const {
    return base + ((k & 1) ? dx : 0) + ((k & 2) ? dy : 0) + ((k & 4) ? dz : 0);
  }
};

__device__ __forceinline__ CellAddr corner_addr(const GridP &g, int cx, int cy, int cz) {
  int x0 = max(min(cx, g.W - 1), 0), x1 = max(min(cx + 1, g.W - 1), 0);
  int y0 = max(min(cy, g.H - 1), 0), y1 = max(min(cy + 1, g.H - 1), 0);
  int z0 = max(min(cz, g.D - 1), 0), z1 = max(min(cz + 1, g.D - 1), 0);
  CellAddr a;
  a.base = (z0 * g.H + y0) * g.W + x0;
  a.dx = x1 - x0;
  a.dy = (y1 - y0) * g.W;
  a.dz = (z1 - z0) * g.H * g.W;
  return a;
}

__device__ __forceinline__ void load_corners(const GridP &g, const CellAddr &a, float v[8]) {
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = __ldg(g.data + a.at(k));
}

__device__ __forceinline__ void load_corners(const GridP &g, const int idx[8], float v[8]) {
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = __ldg(g.data + idx[k]);
}

// ---------------------------------------------------------------------------
// Reference-order arithmetic (REF = true)
// ---------------------------------------------------------------------------

// volume::eval_grad, src/volume.cpp:152-178, from already-gathered corners.
__device__ __forceinline__ void eval_grad_ref(const GridP &g, const float v[8], float w0x, float w0y,
                                              float w0z, float &n, f3 &nx) {
  float w1x = sub_(1.0f, w0x), w1y = sub_(1.0f, w0y), w1z = sub_(1.0f, w0z);
  float v000 = v[0], v100 = v[1], v010 = v[2], v110 = v[3];
  float v001 = v[4], v101 = v[5], v011 = v[6], v111 = v[7];
  float w000 = mul_(mul_(w1x, w1y), w1z), w100 = mul_(mul_(w0x, w1y), w1z);
  float w010 = mul_(mul_(w1x, w0y), w1z), w110 = mul_(mul_(w0x, w0y), w1z);
  float w001 = mul_(mul_(w1x, w1y), w0z), w101 = mul_(mul_(w0x, w1y), w0z);
  float w011 = mul_(mul_(w1x, w0y), w0z), w111 = mul_(mul_(w0x, w0y), w0z);
  float s = mul_(w000, v000);
  s = add_(s, mul_(w100, v100)); s = add_(s, mul_(w010, v010)); s = add_(s, mul_(w110, v110));
  s = add_(s, mul_(w001, v001)); s = add_(s, mul_(w101, v101)); s = add_(s, mul_(w011, v011));
  s = add_(s, mul_(w111, v111));
  n = s;
#define DRRT_Q(va, wa, wb) mul_(mul_(va, wa), wb)
#define DRRT_S4(a, b, c, d) add_(add_(add_(a, b), c), d)
  float gx = sub_(DRRT_S4(DRRT_Q(v100, w1y, w1z), DRRT_Q(v101, w1y, w0z), DRRT_Q(v110, w0y, w1z), DRRT_Q(v111, w0y, w0z)),
                  DRRT_S4(DRRT_Q(v000, w1y, w1z), DRRT_Q(v001, w1y, w0z), DRRT_Q(v010, w0y, w1z), DRRT_Q(v011, w0y, w0z)));
  float gy = sub_(DRRT_S4(DRRT_Q(v010, w1x, w1z), DRRT_Q(v011, w1x, w0z), DRRT_Q(v110, w0x, w1z), DRRT_Q(v111, w0x, w0z)),
                  DRRT_S4(DRRT_Q(v000, w1x, w1z), DRRT_Q(v001, w1x, w0z), DRRT_Q(v100, w0x, w1z), DRRT_Q(v101, w0x, w0z)));
  float gz = sub_(DRRT_S4(DRRT_Q(v001, w1x, w1y), DRRT_Q(v011, w1x, w0y), DRRT_Q(v101, w0x, w1y), DRRT_Q(v111, w0x, w0y)),
                  DRRT_S4(DRRT_Q(v000, w1x, w1y), DRRT_Q(v010, w1x, w0y), DRRT_Q(v100, w0x, w1y), DRRT_Q(v110, w0x, w0y)));
#undef DRRT_Q
#undef DRRT_S4
  nx.x = mul_(gx, g.rcp_h); nx.y = mul_(gy, g.rcp_h); nx.z = mul_(gz, g.rcp_h);
}

// ---------------------------------------------------------------------------
// Packed FP32 (Blackwell FMUL2 / FFMA2: two independent IEEE-rounded float
// operations per issue slot).  The kernels are issue-bound on the literal
// reference arithmetic (profiles/r02_fwd_full.txt: 86 % issue-active, 64 of ~150
// instructions per step are FMULs), so the multiplies are paired up.
// ptxas contracts mul.rn.f32x2 feeding add/sub.rn.f32x2 into FFMA2 even with
// explicit .rn and --fmad=false (checked in SASS), which would change the
// rounding; packed products are therefore only ever consumed by further packed
// multiplies or by SCALAR add.rn.f32, which ptxas never contracts.
// ---------------------------------------------------------------------------
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float lo2(f32x2 v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  (void) hi;
  return lo;
}
__device__ __forceinline__ float hi2(f32x2 v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  (void) lo;
  return hi;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2s(f32x2 a, float b) { return mul2(a, pk2(b, b)); }
// contraction allowed / intended (non-literal arithmetic only)
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// Corner values as x-adjacent pairs: V[j] = (v[2j], v[2j+1]) = (x0, x1) at
// (y,z) = (0,0), (1,0), (0,1), (1,1).
// SM: the grid pointer addresses a shared-memory copy of the whole grid (small grids, see trace_kernel): plain loads,
// which the compiler turns into LDS; otherwise the non-coherent global path.
template <bool SM>
__device__ __forceinline__ float ldv(const float *p) { return SM ? *p : __ldg(p); }

template <bool SM = false>
__device__ __forceinline__ void load_corner_pairs(const GridP &g, const CellAddr &a, f32x2 V[4]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) V[j] = pk2(ldv<SM>(g.data + a.at(2 * j)), ldv<SM>(g.data + a.at(2 * j + 1)));
}

// Address + gather of one cell.  Cells strictly inside the grid (all but the rays'
// first and last steps) take the fast path: unit strides, four row pointers, the
// x+1 neighbour as an immediate offset.
template <bool SM = false>
__device__ __forceinline__ CellAddr gather_cell(const GridP &g, int cx, int cy, int cz, f32x2 V[4]) {
  CellAddr a;
  const bool interior = ((unsigned) cx < (unsigned) (g.W - 1)) & ((unsigned) cy < (unsigned) (g.H - 1)) &
                        ((unsigned) cz < (unsigned) (g.D - 1));
  if (interior) {
    a.base = (cz * g.H + cy) * g.W + cx;
    a.dx = 1; a.dy = g.W; a.dz = g.H * g.W;
    const float *p00 = g.data + a.base, *p10 = p00 + a.dy, *p01 = p00 + a.dz, *p11 = p01 + a.dy;
    V[0] = pk2(ldv<SM>(p00), ldv<SM>(p00 + 1));
    V[1] = pk2(ldv<SM>(p10), ldv<SM>(p10 + 1));
    V[2] = pk2(ldv<SM>(p01), ldv<SM>(p01 + 1));
    V[3] = pk2(ldv<SM>(p11), ldv<SM>(p11 + 1));
  } else {
    a = corner_addr(g, cx, cy, cz);
    load_corner_pairs<SM>(g, a, V);
  }
  return a;
}

// volume::eval_grad, src/volume.cpp:152-178 -- same operations, same order and
// same roundings as eval_grad_ref above, with the 64 multiplies issued as 30
// FMUL2 (+3 FMUL); every add is a scalar FADD in the reference's order.
__device__ __forceinline__ void eval_grad_ref_x2(const GridP &g, const f32x2 V[4], float w0x, float w0y,
                                                 float w0z, float &n, f3 &nx) {
  const float w1x = sub_(1.0f, w0x), w1y = sub_(1.0f, w0y), w1z = sub_(1.0f, w0z);
  const f32x2 A = pk2(w1x, w0x);
  // n = sum_k w_k v_k, w_ijk = (wx_i * wy_j) * wz_k
  const f32x2 Py0 = mul2s(A, w1y), Py1 = mul2s(A, w0y);
  const f32x2 T00 = mul2(mul2s(Py0, w1z), V[0]), T10 = mul2(mul2s(Py1, w1z), V[1]);
  const f32x2 T01 = mul2(mul2s(Py0, w0z), V[2]), T11 = mul2(mul2s(Py1, w0z), V[3]);
  float s = lo2(T00);
  s = add_(s, hi2(T00)); s = add_(s, lo2(T10)); s = add_(s, hi2(T10));
  s = add_(s, lo2(T01)); s = add_(s, hi2(T01)); s = add_(s, lo2(T11)); s = add_(s, hi2(T11));
  n = s;
  // d/dx: (v * wy) * wz, hi = x1 (plus), lo = x0 (minus)
  const f32x2 X00 = mul2s(mul2s(V[0], w1y), w1z), X01 = mul2s(mul2s(V[2], w1y), w0z);
  const f32x2 X10 = mul2s(mul2s(V[1], w0y), w1z), X11 = mul2s(mul2s(V[3], w0y), w0z);
  const float gx = sub_(add_(add_(add_(hi2(X00), hi2(X01)), hi2(X10)), hi2(X11)),
                        add_(add_(add_(lo2(X00), lo2(X01)), lo2(X10)), lo2(X11)));
  // d/dy and d/dz share the first products v * wx
  const f32x2 F00 = mul2(V[0], A), F10 = mul2(V[1], A), F01 = mul2(V[2], A), F11 = mul2(V[3], A);
  const f32x2 G10 = mul2s(F10, w1z), G11 = mul2s(F11, w0z), G00 = mul2s(F00, w1z), G01 = mul2s(F01, w0z);
  const float gy = sub_(add_(add_(add_(lo2(G10), lo2(G11)), hi2(G10)), hi2(G11)),
                        add_(add_(add_(lo2(G00), lo2(G01)), hi2(G00)), hi2(G01)));
  const f32x2 H01 = mul2s(F01, w1y), H11 = mul2s(F11, w0y), H00 = mul2s(F00, w1y), H10 = mul2s(F10, w0y);
  const float gz = sub_(add_(add_(add_(lo2(H01), lo2(H11)), hi2(H01)), hi2(H11)),
                        add_(add_(add_(lo2(H00), lo2(H10)), hi2(H00)), hi2(H10)));
  nx.x = mul_(gx, g.rcp_h); nx.y = mul_(gy, g.rcp_h); nx.z = mul_(gz, g.rcp_h);
}

// ---------------------------------------------------------------------------
// Two rays per thread ("ray pairs"): ray A in the low half and ray B in the high
// half of every f32x2, so that ALL of the literal arithmetic -- multiplies and
// adds -- issues as packed FP32 (one FMUL2 / FFMA2 per two rays).  Adds are
// written as fma(a, 1, b), which rounds exactly like add.rn (a * 1 is exact), with
// the 1 passed in from the host as an opaque kernel argument so that ptxas can
// neither fold it nor contract a producing multiply into the add.
// ---------------------------------------------------------------------------
struct PairK {
  f32x2 one, mone;  // (1, 1), (-1, -1)
  __device__ __forceinline__ f32x2 add(f32x2 a, f32x2 b) const { return fma2(a, one, b); }    // a + b
  __device__ __forceinline__ f32x2 sub(f32x2 a, f32x2 b) const { return fma2(b, mone, a); }   // a - b
};

// volume::eval_grad, src/volume.cpp:152-178, for two rays at once: the same
// operations in the same order with the same roundings as eval_grad_ref.
// Vp[k] = (corner k of ray A's cell, corner k of ray B's cell), reference corner order.
__device__ __forceinline__ void eval_grad_ref_pair(const PairK &K, f32x2 rcp_h, const f32x2 Vp[8], f32x2 w0x,
                                                   f32x2 w0y, f32x2 w0z, f32x2 &n, f32x2 nx[3]) {
  const f32x2 w1x = K.sub(K.one, w0x), w1y = K.sub(K.one, w0y), w1z = K.sub(K.one, w0z);
  {
    const f32x2 p00 = mul2(w1x, w1y), p10 = mul2(w0x, w1y), p01 = mul2(w1x, w0y), p11 = mul2(w0x, w0y);
    f32x2 s = mul2(mul2(p00, w1z), Vp[0]);
    s = K.add(s, mul2(mul2(p10, w1z), Vp[1]));
    s = K.add(s, mul2(mul2(p01, w1z), Vp[2]));
    s = K.add(s, mul2(mul2(p11, w1z), Vp[3]));
    s = K.add(s, mul2(mul2(p00, w0z), Vp[4]));
    s = K.add(s, mul2(mul2(p10, w0z), Vp[5]));
    s = K.add(s, mul2(mul2(p01, w0z), Vp[6]));
    s = K.add(s, mul2(mul2(p11, w0z), Vp[7]));
    n = s;
  }
#define DRRT_Q2(va, wa, wb) mul2(mul2(va, wa), wb)
#define DRRT_S42(a, b, c, d) K.add(K.add(K.add(a, b), c), d)
  const f32x2 gx = K.sub(DRRT_S42(DRRT_Q2(Vp[1], w1y, w1z), DRRT_Q2(Vp[5], w1y, w0z), DRRT_Q2(Vp[3], w0y, w1z), DRRT_Q2(Vp[7], w0y, w0z)),
                         DRRT_S42(DRRT_Q2(Vp[0], w1y, w1z), DRRT_Q2(Vp[4], w1y, w0z), DRRT_Q2(Vp[2], w0y, w1z), DRRT_Q2(Vp[6], w0y, w0z)));
  // d/dy and d/dz share the first products v * wx
  f32x2 F[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) F[k] = mul2(Vp[k], (k & 1) ? w0x : w1x);
  const f32x2 gy = K.sub(DRRT_S42(mul2(F[2], w1z), mul2(F[6], w0z), mul2(F[3], w1z), mul2(F[7], w0z)),
                         DRRT_S42(mul2(F[0], w1z), mul2(F[4], w0z), mul2(F[1], w1z), mul2(F[5], w0z)));
  const f32x2 gz = K.sub(DRRT_S42(mul2(F[4], w1y), mul2(F[6], w0y), mul2(F[5], w1y), mul2(F[7], w0y)),
                         DRRT_S42(mul2(F[0], w1y), mul2(F[2], w0y), mul2(F[1], w1y), mul2(F[3], w0y)));
#undef DRRT_Q2
#undef DRRT_S42
  nx[0] = mul2(gx, rcp_h); nx[1] = mul2(gy, rcp_h); nx[2] = mul2(gz, rcp_h);
}

// enoki lerp(a,b,t) = fmadd(b, t, fnmadd(a, t, a))
__device__ __forceinline__ float lerp_ref(float a, float b, float t) { return fma_(b, t, fma_(-a, t, a)); }

// volume::eval_hess, src/volume.cpp:79-98: (dxdy, dxdz, dydz), / h / h as two
// multiplications by rcp(h).
__device__ __forceinline__ f3 eval_hess_ref(const GridP &g, const float v[8], float w0x, float w0y,
                                            float w0z) {
  float v000 = v[0], v100 = v[1], v010 = v[2], v110 = v[3];
  float v001 = v[4], v101 = v[5], v011 = v[6], v111 = v[7];
  float dxdy = lerp_ref(add_(sub_(sub_(v110, v010), v100), v000), add_(sub_(sub_(v111, v011), v101), v001), w0z);
  float dxdz = lerp_ref(add_(sub_(sub_(v101, v001), v100), v000), add_(sub_(sub_(v111, v011), v110), v010), w0y);
  float dydz = lerp_ref(add_(sub_(sub_(v011, v001), v010), v000), add_(sub_(sub_(v111, v101), v110), v100), w0x);
  f3 hm;
  hm.x = mul_(mul_(dxdy, g.rcp_h), g.rcp_h);
  hm.y = mul_(mul_(dxdz, g.rcp_h), g.rcp_h);
  hm.z = mul_(mul_(dydz, g.rcp_h), g.rcp_h);
  return hm;
}

// The 16 terms of volume::splat, src/volume.cpp:217-243 (no 1/h on the gradient
// weights -- the reference's quirk, kept).
__device__ __forceinline__ void splat_terms_ref(float val, f3 gr, float w0x, float w0y, float w0z,
                                                float t[16]) {
  float w1x = sub_(1.0f, w0x), w1y = sub_(1.0f, w0y), w1z = sub_(1.0f, w0z);
  t[0] = mul_(mul_(mul_(val, w1x), w1y), w1z);
  t[1] = mul_(mul_(mul_(val, w0x), w1y), w1z);
  t[2] = mul_(mul_(mul_(val, w1x), w0y), w1z);
  t[3] = mul_(mul_(mul_(val, w0x), w0y), w1z);
  t[4] = mul_(mul_(mul_(val, w1x), w1y), w0z);
  t[5] = mul_(mul_(mul_(val, w0x), w1y), w0z);
  t[6] = mul_(mul_(mul_(val, w1x), w0y), w0z);
  t[7] = mul_(mul_(mul_(val, w0x), w0y), w0z);
  float gx = gr.x, gy = gr.y, gz = gr.z, mx = -gr.x;
#define DRRT_P(a, b, c) mul_(mul_(a, b), c)
  t[8]  = sub_(sub_(DRRT_P(mx, w1y, w1z), DRRT_P(gy, w1x, w1z)), DRRT_P(gz, w1x, w1y));
  t[9]  = sub_(sub_(DRRT_P(gx, w1y, w1z), DRRT_P(gy, w0x, w1z)), DRRT_P(gz, w0x, w1y));
  t[10] = sub_(add_(DRRT_P(mx, w0y, w1z), DRRT_P(gy, w1x, w1z)), DRRT_P(gz, w1x, w0y));
  t[11] = sub_(add_(DRRT_P(gx, w0y, w1z), DRRT_P(gy, w0x, w1z)), DRRT_P(gz, w0x, w0y));
  t[12] = add_(sub_(DRRT_P(mx, w1y, w0z), DRRT_P(gy, w1x, w0z)), DRRT_P(gz, w1x, w1y));
  t[13] = add_(sub_(DRRT_P(gx, w1y, w0z), DRRT_P(gy, w0x, w0z)), DRRT_P(gz, w0x, w1y));
  t[14] = add_(add_(DRRT_P(mx, w0y, w0z), DRRT_P(gy, w1x, w0z)), DRRT_P(gz, w1x, w0y));
  t[15] = add_(add_(DRRT_P(gx, w0y, w0z), DRRT_P(gy, w0x, w0z)), DRRT_P(gz, w0x, w0y));
#undef DRRT_P
}

// ---------------------------------------------------------------------------
// Polynomial cell (REF = false)
// ---------------------------------------------------------------------------
struct CellPoly {
  float a0, ax, ay, az, axy, axz, ayz, axyz;
};

__device__ __forceinline__ CellPoly make_poly(const float v[8]) {
  CellPoly c;
  c.a0 = v[0];
  c.ax = v[1] - v[0];
  c.ay = v[2] - v[0];
  c.az = v[4] - v[0];
  float bx = v[3] - v[2];  // x-difference on the y1 edge
  float cxx = v[5] - v[4]; // x-difference on the z1 edge
  float dxx = v[7] - v[6]; // x-difference on the y1 z1 edge
  c.axy = bx - c.ax;
  c.axz = cxx - c.ax;
  c.ayz = (v[6] - v[4]) - c.ay;
  c.axyz = (dxx - cxx) - c.axy;
  return c;
}

// n, grad n (in cell units: multiply by rcp_h for world units) and the three
// mixed second partials (cell units: multiply by rcp_h^2).
__device__ __forceinline__ void eval_poly(const CellPoly &c, float wx, float wy, float wz, float &n,
                                          f3 &g, f3 &hm) {
  float pxy = fmaf(c.axyz, wz, c.axy);
  float pxz = fmaf(c.axyz, wy, c.axz);
  float pyz = fmaf(c.axyz, wx, c.ayz);
  float tx = fmaf(c.axz, wz, c.ax);
  float ty = fmaf(c.ayz, wz, c.ay);
  float tz = fmaf(c.ayz, wy, c.az);
  g.x = fmaf(pxy, wy, tx);
  g.y = fmaf(pxy, wx, ty);
  g.z = fmaf(pxz, wx, tz);
  n = fmaf(wx, g.x, fmaf(wy, ty, fmaf(c.az, wz, c.a0)));
  hm.x = pxy; hm.y = pxz; hm.z = pyz;
}

// Fused splat: contribution of (val, gr) to the 8 corners, i.e. term[k] +
// term[8+k] of splat_terms_ref, factored as
//   c_ijk = A_i * wy_j * wz_k + wx_i * (gy*s_j*wz_k + gz*s_k*wy_j),
//   A_i = val*wx_i + s_i*gx,  s_0 = -1, s_1 = +1,  w_0 = 1 - w, w_1 = w.
__device__ __forceinline__ void splat_accumulate(float val, f3 gr, float wx, float wy, float wz,
                                                 float acc[8]) {
  float ux = 1.0f - wx, uy = 1.0f - wy, uz = 1.0f - wz;
  float A0 = fmaf(val, ux, -gr.x), A1 = fmaf(val, wx, gr.x);
  float B00 = uy * uz, B10 = wy * uz, B01 = uy * wz, B11 = wy * wz;
  float gyuz = gr.y * uz, gywz = gr.y * wz, gzuy = gr.z * uy, gzwy = gr.z * wy;
  float C00 = -gyuz - gzuy, C10 = gyuz - gzwy, C01 = gzuy - gywz, C11 = gywz + gzwy;
  acc[0] = fmaf(ux, C00, fmaf(A0, B00, acc[0]));
  acc[1] = fmaf(wx, C00, fmaf(A1, B00, acc[1]));
  acc[2] = fmaf(ux, C10, fmaf(A0, B10, acc[2]));
  acc[3] = fmaf(wx, C10, fmaf(A1, B10, acc[3]));
  acc[4] = fmaf(ux, C01, fmaf(A0, B01, acc[4]));
  acc[5] = fmaf(wx, C01, fmaf(A1, B01, acc[5]));
  acc[6] = fmaf(ux, C11, fmaf(A0, B11, acc[6]));
  acc[7] = fmaf(wx, C11, fmaf(A1, B11, acc[7]));
}

// The same fused splat on x-adjacent accumulator pairs P[j] = (acc[2j], acc[2j+1])
// with FFMA2 (contraction is intended here: this is not the literal arithmetic).
__device__ __forceinline__ void splat_accumulate_x2(float val, f3 gr, float wx, float wy, float wz,
                                                    f32x2 P[4]) {
  const float ux = 1.0f - wx, uy = 1.0f - wy, uz = 1.0f - wz;
  const f32x2 X = pk2(ux, wx);
  const f32x2 A = fma2(pk2(val, val), X, pk2(-gr.x, gr.x));
  const float B00 = uy * uz, B10 = wy * uz, B01 = uy * wz, B11 = wy * wz;
  const float gyuz = gr.y * uz, gywz = gr.y * wz, gzuy = gr.z * uy, gzwy = gr.z * wy;
  const float C00 = -gyuz - gzuy, C10 = gyuz - gzwy, C01 = gzuy - gywz, C11 = gywz + gzwy;
  P[0] = fma2(X, pk2(C00, C00), fma2(A, pk2(B00, B00), P[0]));
  P[1] = fma2(X, pk2(C10, C10), fma2(A, pk2(B10, B10), P[1]));
  P[2] = fma2(X, pk2(C01, C01), fma2(A, pk2(B01, B01), P[2]));
  P[3] = fma2(X, pk2(C11, C11), fma2(A, pk2(B11, B11), P[3]));
}

// ---------------------------------------------------------------------------
// Gradient accumulation back ends
// ---------------------------------------------------------------------------
// float atomics (red.global.add.f32) or 64-bit fixed point for the
// deterministic mode: integer addition is associative, so the sum is
// independent of the order in which threads arrive.
struct AccumFloat {
  typedef float T;
  float *dst;
  __device__ __forceinline__ void prepare() {}
  // gradient copy `r` of this CTA (0 = dst itself, r >= 1 = copy r - 1 behind `base`)
  __device__ __forceinline__ void select(int r, float *base, int64_t stride) { if (r) dst = base + (int64_t) (r - 1) * stride; }
  __device__ __forceinline__ float q(float v) const { return v; }
  __device__ __forceinline__ void add(int idx, float v) const { atomicAdd(dst + idx, v); }
};

struct AccumFixed {
  typedef unsigned long long T;
  unsigned long long *dst;
  const unsigned *max_bits;  // bit pattern of max(|dx|,|dv|) over the batch (device)
  float scale;               // set by prepare(): a power of two
  // scale = 2^(34 - e) with max = m * 2^e, m in [0.5, 1): every thread derives the
  // same value, contributions keep ~34 bits below the largest seed and ~29 bits of
  // headroom above it before the int64 sum could overflow.
  static __device__ __forceinline__ float scale_from_bits(unsigned bits) {
    float m = __uint_as_float(bits);
    if (!(m > 0.f) || !(m < 3.0e38f)) return 1.0f;
    int e;
    frexpf(m, &e);
    return exp2f((float) (34 - e));
  }
  __device__ __forceinline__ void prepare() { scale = scale_from_bits(*max_bits); }
  __device__ __forceinline__ void select(int, float *, int64_t) {}  // integer sums need no copies
  __device__ __forceinline__ unsigned long long q(float v) const {
    return (unsigned long long) __float2ll_rn(v * scale);
  }
  __device__ __forceinline__ void add(int idx, float v) const { atomicAdd(dst + idx, q(v)); }
};

__device__ __forceinline__ int64_t warp_sum_i64(int64_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace drrt
