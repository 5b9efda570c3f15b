/*
 * TEST INFRASTRUCTURE ONLY -- not product code.
 *
 * Plain-C CPU restatement of the reference's eikonal tracer hot path
 * (ArjunTeh/AdjointNonlinearRayTracing): forward trace variants, adjoint
 * backtrace variants, and the two volume types they evaluate.  Every function
 * cites the reference file:line it follows.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg may load this; the product (the CUDA library
 * behind include/drrt_b200.h) never does.
 *
 * Parity pinning: the reference ships no tests or golden vectors for this path
 * (SURVEY.md section 4), so this restatement is pinned against the reference's own
 * CPU instantiation Tracer<false,false> compiled from /root/reference by
 * oracle/build_ref.sh (oracle/_ref/libdrrt_ref_v{3,4}.so) -- see
 * tests/test_oracle_vs_ref.py and the committed fixtures under tests/golden/.
 *
 * Layout: all buffers are host memory; ray vectors are AoS (N,3) exactly like the
 * torch tensors core/tracer.py hands to enoki; the grid is flat with
 * idx = (z*H + y)*W + x (src/volume.cpp:134).
 *
 * Arithmetic: `real` is float (default) or double (-DORACLE_DOUBLE, used only to
 * arbitrate float32 disagreements).  fmadd() in the reference is a hardware FMA
 * (`fma_` here); everything else is a separate rounded operation, so build this
 * file with -ffp-contract=off.  enoki's rcp()/rsqrt() are approximate+Newton on
 * the CPU and rcp.approx on the GPU; here they are exact 1/x and 1/sqrt(x).
 * The structure is kept literal: outer loop over integration steps, inner loop
 * over rays, batch-wide early exit -- so the batch-dependent quirks of
 * trace_target are reproduced by construction.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef ORACLE_DOUBLE
typedef double real;
#define fma_ fma
#define floor_ floor
#define sqrt_ sqrt
#define SYM(name) orc64_##name
#else
typedef float real;
#define fma_ fmaf
#define floor_ floorf
#define sqrt_ sqrtf
#define SYM(name) orc_##name
#endif

typedef struct {
  int W, H, D;
  real h;
  real rcp_h;  /* rcp(h_), src/volume.cpp:128 */
  real hi[3];  /* (res_k - 1) * h_, src/volume.cpp:252-254 */
  real *data;  /* mutable for splat targets */
} vol_t;

static vol_t vol_make(real *data, int W, int H, int D, real h) {
  vol_t g;
  g.W = W; g.H = H; g.D = D; g.h = h;
  g.rcp_h = (real) 1 / h;
  g.hi[0] = (real) (W - 1) * h;
  g.hi[1] = (real) (H - 1) * h;
  g.hi[2] = (real) (D - 1) * h;
  g.data = data;
  return g;
}

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* Cell lookup shared by eval_grad / eval_hess / splat: src/volume.cpp:128-141,
 * :55-68, :200-214.  floor2int rounds toward -inf; both corners are clamped. */
typedef struct { int idx[8]; real w0[3], w1[3]; } cell_t;

static inline cell_t vol_cell(const vol_t *g, const real p[3]) {
  cell_t c;
  int p0[3], p1[3];
  const int res[3] = {g->W, g->H, g->D};
  for (int k = 0; k < 3; ++k) {
    real pm = p[k] * g->rcp_h;
    real fl = floor_(pm);
    int ip = (int) fl;
    c.w0[k] = pm - (real) ip;
    c.w1[k] = (real) 1 - c.w0[k];
    p0[k] = imax(imin(ip, res[k] - 1), 0);
    p1[k] = imax(imin(ip + 1, res[k] - 1), 0);
  }
  /* order: 000 100 010 110 001 101 011 111 (first digit = x) */
  for (int k = 0; k < 8; ++k) {
    int x = (k & 1) ? p1[0] : p0[0];
    int y = (k & 2) ? p1[1] : p0[1];
    int z = (k & 4) ? p1[2] : p0[2];
    c.idx[k] = (z * g->H + y) * g->W + x;
  }
  return c;
}

/* volume::eval_grad, src/volume.cpp:101-181.  Masked-off lanes gather 0
 * (ext/enoki/include/enoki/cuda.h:855-857; CPU masked gather likewise). */
static inline void vol_eval_grad(const vol_t *g, const real p[3], int mask, real *n, real nx[3]) {
  cell_t c = vol_cell(g, p);
  real v[8];
  for (int k = 0; k < 8; ++k) v[k] = mask ? g->data[c.idx[k]] : (real) 0;
  const real *w0 = c.w0, *w1 = c.w1;
  real v000 = v[0], v100 = v[1], v010 = v[2], v110 = v[3];
  real v001 = v[4], v101 = v[5], v011 = v[6], v111 = v[7];
  real w000 = w1[0] * w1[1] * w1[2], w100 = w0[0] * w1[1] * w1[2];
  real w010 = w1[0] * w0[1] * w1[2], w110 = w0[0] * w0[1] * w1[2];
  real w001 = w1[0] * w1[1] * w0[2], w101 = w0[0] * w1[1] * w0[2];
  real w011 = w1[0] * w0[1] * w0[2], w111 = w0[0] * w0[1] * w0[2];
  *n = w000 * v000 + w100 * v100 + w010 * v010 + w110 * v110 +
       w001 * v001 + w101 * v101 + w011 * v011 + w111 * v111;
  real gx = (v100 * w1[1] * w1[2] + v101 * w1[1] * w0[2] + v110 * w0[1] * w1[2] + v111 * w0[1] * w0[2]) -
            (v000 * w1[1] * w1[2] + v001 * w1[1] * w0[2] + v010 * w0[1] * w1[2] + v011 * w0[1] * w0[2]);
  real gy = (v010 * w1[0] * w1[2] + v011 * w1[0] * w0[2] + v110 * w0[0] * w1[2] + v111 * w0[0] * w0[2]) -
            (v000 * w1[0] * w1[2] + v001 * w1[0] * w0[2] + v100 * w0[0] * w1[2] + v101 * w0[0] * w0[2]);
  real gz = (v001 * w1[0] * w1[1] + v011 * w1[0] * w0[1] + v101 * w0[0] * w1[1] + v111 * w0[0] * w0[1]) -
            (v000 * w1[0] * w1[1] + v010 * w1[0] * w0[1] + v100 * w0[0] * w1[1] + v110 * w0[0] * w0[1]);
  nx[0] = gx * g->rcp_h;
  nx[1] = gy * g->rcp_h;
  nx[2] = gz * g->rcp_h;
}

/* enoki lerp(a,b,t) = fmadd(b, t, fnmadd(a, t, a)), array_math.h:1351-1354 */
static inline real lerp_(real a, real b, real t) { return fma_(b, t, fma_(-a, t, a)); }

/* volume::eval_hess, src/volume.cpp:40-99: only mixed partials are non-zero.
 * Returns (dxdy, dxdz, dydz); H / h_ / h_ is two multiplications by rcp(h)
 * (ext/enoki/include/enoki/matrix.h:197-203). */
static inline void vol_eval_hess(const vol_t *g, const real p[3], int mask, real hx[3]) {
  cell_t c = vol_cell(g, p);
  real v[8];
  for (int k = 0; k < 8; ++k) v[k] = mask ? g->data[c.idx[k]] : (real) 0;
  real v000 = v[0], v100 = v[1], v010 = v[2], v110 = v[3];
  real v001 = v[4], v101 = v[5], v011 = v[6], v111 = v[7];
  real dxdy = lerp_(v110 - v010 - v100 + v000, v111 - v011 - v101 + v001, c.w0[2]);
  real dxdz = lerp_(v101 - v001 - v100 + v000, v111 - v011 - v110 + v010, c.w0[1]);
  real dydz = lerp_(v011 - v001 - v010 + v000, v111 - v101 - v110 + v100, c.w0[0]);
  hx[0] = dxdy * g->rcp_h * g->rcp_h;
  hx[1] = dxdz * g->rcp_h * g->rcp_h;
  hx[2] = dydz * g->rcp_h * g->rcp_h;
}

/* One ray's 16 contributions of volume::splat, src/volume.cpp:182-244, in the
 * reference's scatter_add order: term t in 0..7 is `val * w_t` to corner t
 * (:217-224), term 8..15 the gradient weights (:227-243).  NOTE the reference's
 * gradient term carries no 1/h -- reproduced. */
static inline void vol_splat_terms(const vol_t *g, const real p[3], real val, const real gr[3],
                                   int idx[8], real term[16]) {
  cell_t c = vol_cell(g, p);
  const real *w0 = c.w0, *w1 = c.w1;
  for (int k = 0; k < 8; ++k) idx[k] = c.idx[k];
  term[0] = val * w1[0] * w1[1] * w1[2];
  term[1] = val * w0[0] * w1[1] * w1[2];
  term[2] = val * w1[0] * w0[1] * w1[2];
  term[3] = val * w0[0] * w0[1] * w1[2];
  term[4] = val * w1[0] * w1[1] * w0[2];
  term[5] = val * w0[0] * w1[1] * w0[2];
  term[6] = val * w1[0] * w0[1] * w0[2];
  term[7] = val * w0[0] * w0[1] * w0[2];
  real gx = gr[0], gy = gr[1], gz = gr[2];
  term[8]  = -gx * w1[1] * w1[2] - gy * w1[0] * w1[2] - gz * w1[0] * w1[1];
  term[9]  =  gx * w1[1] * w1[2] - gy * w0[0] * w1[2] - gz * w0[0] * w1[1];
  term[10] = -gx * w0[1] * w1[2] + gy * w1[0] * w1[2] - gz * w1[0] * w0[1];
  term[11] =  gx * w0[1] * w1[2] + gy * w0[0] * w1[2] - gz * w0[0] * w0[1];
  term[12] = -gx * w1[1] * w0[2] - gy * w1[0] * w0[2] + gz * w1[0] * w1[1];
  term[13] =  gx * w1[1] * w0[2] - gy * w0[0] * w0[2] + gz * w0[0] * w1[1];
  term[14] = -gx * w0[1] * w0[2] + gy * w1[0] * w0[2] + gz * w1[0] * w0[1];
  term[15] =  gx * w0[1] * w0[2] + gy * w0[0] * w0[2] + gz * w0[0] * w0[1];
}

/* volume::inbounds, src/volume.cpp:246-256 */
static inline int vol_inbounds(const vol_t *g, const real p[3]) {
  return p[0] >= 0 && p[1] >= 0 && p[2] >= 0 && p[0] < g->hi[0] && p[1] < g->hi[1] && p[2] < g->hi[2];
}

/* volume::escaped, src/volume.cpp:258-271 */
static inline int vol_escaped(const vol_t *g, const real p[3], const real v[3]) {
  int e = 0;
  for (int k = 0; k < 3; ++k)
    e |= (p[k] < 0 && v[k] < 0) || (p[k] >= g->hi[k] && v[k] > 0);
  return e;
}

/* ------------------------------------------------------------------------ */
/* cylinder_volume, src/cylinder_volume.cpp                                  */
/* ------------------------------------------------------------------------ */

#define CYL_EPS ((real) 1e-6f) /* FLOAT_EPSILON, src/cylinder_volume.cpp:15 */

typedef struct {
  int res;
  real radius, length;
  real h; /* radius_ / (res-1), src/cylinder_volume.cpp:42 */
  real *data;
} cyl_t;

static cyl_t cyl_make(real *data, int res, real radius, real length) {
  cyl_t c;
  c.res = res; c.radius = radius; c.length = length;
  c.h = radius / (real) (res - 1);
  c.data = data;
  return c;
}

typedef struct { int idx0, idx1; real w0, w1, r; real xs[3]; } cyl_cell_t;

/* radial lookup shared by eval_grad/eval_hess/splat: src/cylinder_volume.cpp:37-48.
 * NOTE w0 is measured from the *clamped* idx0 (":48"), unlike the grid volume. */
static inline cyl_cell_t cyl_cell(const cyl_t *c, const real p[3]) {
  cyl_cell_t q;
  q.xs[0] = p[0] - c->radius;
  q.xs[1] = 0;
  q.xs[2] = p[2] - c->radius;
  /* norm = sqrt(dot): x*x, fmadd(y,y,.), fmadd(z,z,.)  (array_static.h:948-957) */
  real d = q.xs[0] * q.xs[0];
  d = fma_(q.xs[1], q.xs[1], d);
  d = fma_(q.xs[2], q.xs[2], d);
  q.r = sqrt_(d);
  real rm = q.r / c->h;
  int fl = (int) floor_(rm);
  q.idx0 = imax(imin(fl, c->res - 1), 0);
  q.idx1 = imax(imin(q.idx0 + 1, c->res - 1), 0);
  q.w0 = rm - (real) q.idx0;
  q.w1 = (real) 1 - q.w0;
  return q;
}

/* normalize(xs) = xs * rsqrt(squared_norm(xs)), array_router.h:665-667 */
static inline void cyl_xhat(const cyl_cell_t *q, real xhat[3]) {
  real d = q->xs[0] * q->xs[0];
  d = fma_(q->xs[1], q->xs[1], d);
  d = fma_(q->xs[2], q->xs[2], d);
  real rs = (real) 1 / sqrt_(d);
  xhat[0] = q->xs[0] * rs;
  xhat[1] = q->xs[1] * rs;
  xhat[2] = q->xs[2] * rs;
}

/* cylinder_volume::eval_grad, src/cylinder_volume.cpp:26-59 (gathers unmasked) */
static inline void cyl_eval_grad(const cyl_t *c, const real p[3], real *n, real nx[3]) {
  cyl_cell_t q = cyl_cell(c, p);
  real val0 = c->data[q.idx0], val1 = c->data[q.idx1];
  *n = val0 * q.w1 + val1 * q.w0;
  real rx = (val1 - val0) / c->h;
  real xhat[3];
  cyl_xhat(&q, xhat);
  int small = q.r < CYL_EPS;
  for (int k = 0; k < 3; ++k) nx[k] = small ? (real) 0 : rx * xhat[k];
}

/* cylinder_volume::eval_hess, src/cylinder_volume.cpp:61-111; row-major 3x3 */
static inline void cyl_eval_hess(const cyl_t *c, const real p[3], real Hm[9]) {
  cyl_cell_t q = cyl_cell(c, p);
  real val0 = c->data[q.idx0], val1 = c->data[q.idx1];
  real rx = (val1 - val0) / c->h;
  real xhat[3];
  cyl_xhat(&q, xhat);
  int small = q.r < CYL_EPS;
  if (small) xhat[0] = xhat[1] = xhat[2] = 0;
  real s = rx / q.r;
  for (int k = 0; k < 9; ++k) Hm[k] = 0;
  Hm[0] = ((real) 1 - xhat[0] * xhat[0]) * s;
  Hm[2] = (-(xhat[0] * xhat[2])) * s;
  Hm[6] = (-(xhat[2] * xhat[0])) * s;
  Hm[8] = ((real) 1 - xhat[2] * xhat[2]) * s;
  if (small) for (int k = 0; k < 9; ++k) Hm[k] = 0;
}

/* cylinder_volume::splat, src/cylinder_volume.cpp:113-148: four scatter_adds */
static inline void cyl_splat_terms(const cyl_t *c, const real p[3], real val, const real gr[3],
                                   int idx[2], real term[4]) {
  cyl_cell_t q = cyl_cell(c, p);
  real xhat[3];
  cyl_xhat(&q, xhat);
  idx[0] = q.idx0; idx[1] = q.idx1;
  term[0] = val * q.w1;
  term[1] = val * q.w0;
  real gv = gr[0] * xhat[0];
  gv = fma_(gr[1], xhat[1], gv);
  gv = fma_(gr[2], xhat[2], gv);
  if (q.r < CYL_EPS) gv = 0;
  term[2] = -gv / c->h;
  term[3] = gv / c->h;
}

/* cylinder_volume::inbounds, src/cylinder_volume.cpp:150-156 */
static inline int cyl_inbounds(const cyl_t *c, const real p[3]) {
  real px = p[0] - c->radius, pz = p[2] - c->radius;
  real r = px * px + pz * pz;
  return (r < c->radius * c->radius) && (p[1] < c->length) && (p[1] >= 0);
}

/* cylinder_volume::escaped, src/cylinder_volume.cpp:158-170 */
static inline int cyl_escaped(const cyl_t *c, const real p[3], const real v[3]) {
  real px = p[0] - c->radius, pz = p[2] - c->radius;
  int esc_length = (p[1] < 0 && v[1] < 0) || (p[1] > c->length && v[1] > 0);
  int out_radius = (px * px + pz * pz) >= (c->radius * c->radius);
  int esc_radius = (px * v[0] + pz * v[2]) > 0;
  return (out_radius && esc_radius) || esc_length;
}

/* ------------------------------------------------------------------------ */
/* Tracer, src/tracer.cpp                                                    */
/* ------------------------------------------------------------------------ */

/* loop caps are evaluated in float and truncated (src/tracer.cpp:51, :262, :417) */
static int cap_grid(float mult, float h, int W, int H, int D, float ds) {
  int m = W > H ? W : H;
  if (D > m) m = D;
  return (int) (mult * h * (float) m / ds);
}

typedef struct {
  real *x, *v;          /* (N,3) working state */
  unsigned char *inside, *escaped, *active;
} fwd_state_t;

static fwd_state_t fwd_alloc(int64_t N, const real *pos, const real *vel) {
  fwd_state_t s;
  s.x = (real *) malloc(sizeof(real) * 3 * (size_t) (N ? N : 1));
  s.v = (real *) malloc(sizeof(real) * 3 * (size_t) (N ? N : 1));
  s.inside = (unsigned char *) calloc((size_t) (N ? N : 1), 1);
  s.escaped = (unsigned char *) calloc((size_t) (N ? N : 1), 1);
  s.active = (unsigned char *) malloc((size_t) (N ? N : 1));
  memcpy(s.x, pos, sizeof(real) * 3 * (size_t) N);
  memcpy(s.v, vel, sizeof(real) * 3 * (size_t) N);
  memset(s.active, 1, (size_t) N);
  return s;
}

static void fwd_free(fwd_state_t *s) {
  free(s->x); free(s->v); free(s->inside); free(s->escaped); free(s->active);
}

enum { K_PLAIN = 0, K_PLANE = 1, K_TARGET = 2 };

/* Shared body of Tracer::trace (src/tracer.cpp:35-100), trace_plane (:102-172)
 * and trace_target (:174-242); they differ only where `kind` is tested. */
static int64_t trace_grid(int kind, const real *rif, int W, int H, int D, const real *pos,
                          const real *vel, const real *aux0 /*pln_o | target*/,
                          const real *aux1 /*pln_d*/, int64_t N, real h, real ds_, real *xt,
                          real *vt, unsigned char *not_escaped, real *dist2) {
  vol_t grid = vol_make((real *) rif, W, H, D, h);
  int max_steps = cap_grid(4.0f, (float) h, W, H, D, (float) ds_);   /* :51 */
  fwd_state_t s = fwd_alloc(N, pos, vel);
  memcpy(xt, pos, sizeof(real) * 3 * (size_t) N);                     /* :56-57 */
  memcpy(vt, vel, sizeof(real) * 3 * (size_t) N);
  const real ds = ds_;
  int64_t live_steps = 0;
  for (int64_t r = 0; r < N; ++r) {
    s.inside[r] = (unsigned char) vol_inbounds(&grid, s.x + 3 * r);   /* :61 */
    if (kind == K_TARGET) {                                           /* :200 */
      real d[3] = {s.x[3 * r] - aux0[3 * r], s.x[3 * r + 1] - aux0[3 * r + 1], s.x[3 * r + 2] - aux0[3 * r + 2]};
      dist2[r] = fma_(d[2], d[2], fma_(d[1], d[1], d[0] * d[0]));
    }
  }
  for (int i = 0; i < max_steps; ++i) {                               /* :66 */
    int all_escaped = 1;
    for (int64_t r = 0; r < N; ++r) {
      real *x = s.x + 3 * r, *v = s.v + 3 * r;
      if (!s.escaped[r]) ++live_steps;
      real n, nx[3];
      vol_eval_grad(&grid, x, s.inside[r], &n, nx);                   /* :68 */
      real dsn = ds * n;
      for (int k = 0; k < 3; ++k) v[k] = fma_(dsn, nx[k], v[k]);      /* :70 */
      for (int k = 0; k < 3; ++k) x[k] = fma_(ds, v[k], x[k]);        /* :71 */
      int cur_inside = vol_inbounds(&grid, x);                        /* :73 */
      if (kind == K_PLANE) {                                          /* :144-145 */
        const real *o = aux0 + 3 * r, *d = aux1 + 3 * r;
        real dd = (x[0] - o[0]) * d[0];
        dd = fma_(x[1] - o[1], d[1], dd);
        dd = fma_(x[2] - o[2], d[2], dd);
        cur_inside = cur_inside && !(dd > 0);
      }
      int cross = s.inside[r] && !cur_inside;                         /* :74 */
      s.escaped[r] |= (unsigned char) cross;                          /* :75 */
      s.escaped[r] |= (unsigned char) vol_escaped(&grid, x, v);       /* :76 */
      s.active[r] &= (unsigned char) !s.escaped[r];                   /* :77 */
      if (kind == K_TARGET) {                                         /* :216-227 */
        const real *t = aux0 + 3 * r;
        real d[3] = {x[0] - t[0], x[1] - t[1], x[2] - t[2]};
        real cur = fma_(d[2], d[2], fma_(d[1], d[1], d[0] * d[0]));
        if (cur < dist2[r]) {
          for (int k = 0; k < 3; ++k) { xt[3 * r + k] = x[k]; vt[3 * r + k] = v[k]; }
          dist2[r] = cur;
        }
      } else if (cross) {                                             /* :79-80 */
        for (int k = 0; k < 3; ++k) { xt[3 * r + k] = x[k]; vt[3 * r + k] = v[k]; }
      }
      all_escaped &= s.escaped[r];
      s.inside[r] = (unsigned char) cur_inside;                       /* :86 (harmless before break) */
    }
    if (all_escaped) break;                                           /* :82 */
  }
  if (kind != K_TARGET) {                                             /* :89-96; :238 is commented out */
    int any_active = 0;
    for (int64_t r = 0; r < N; ++r) any_active |= s.active[r];
    if (any_active)
      for (int64_t r = 0; r < N; ++r)
        if (!s.escaped[r])
          for (int k = 0; k < 3; ++k) xt[3 * r + k] = s.x[3 * r + k];
  }
  if (kind == K_PLANE)                                                /* :171 */
    for (int64_t r = 0; r < N; ++r) not_escaped[r] = (unsigned char) !s.escaped[r];
  fwd_free(&s);
  return live_steps;
}

int64_t SYM(trace)(const real *rif, int W, int H, int D, const real *pos, const real *vel,
                   int64_t N, real h, real ds, real *xt, real *vt) {
  return trace_grid(K_PLAIN, rif, W, H, D, pos, vel, 0, 0, N, h, ds, xt, vt, 0, 0);
}

int64_t SYM(trace_plane)(const real *rif, int W, int H, int D, const real *pos, const real *vel,
                         const real *pln_o, const real *pln_d, int64_t N, real h, real ds,
                         real *xt, real *vt, unsigned char *not_escaped) {
  return trace_grid(K_PLANE, rif, W, H, D, pos, vel, pln_o, pln_d, N, h, ds, xt, vt, not_escaped, 0);
}

int64_t SYM(trace_target)(const real *rif, int W, int H, int D, const real *pos, const real *vel,
                          const real *target, int64_t N, real h, real ds, real *xt, real *vt,
                          real *dist2) {
  return trace_grid(K_TARGET, rif, W, H, D, pos, vel, target, 0, N, h, ds, xt, vt, 0, dist2);
}

/* Tracer::trace_sdf, src/tracer.cpp:244-310.  `inside` starts as the grid's
 * inbounds (:272) and thereafter is `sdf(x) < 0` evaluated under the previous
 * `inside` mask (:287-288), so a ray that is once not-inside stays so. */
int64_t SYM(trace_sdf)(const real *rif, const real *sdf, int W, int H, int D, const real *pos,
                       const real *vel, int64_t N, real h, real ds_, real *xt, real *vt) {
  vol_t grid = vol_make((real *) rif, W, H, D, h);
  vol_t sdfv = vol_make((real *) sdf, W, H, D, h);
  int max_steps = cap_grid(2.0f, (float) h, W, H, D, (float) ds_);   /* :262 */
  fwd_state_t s = fwd_alloc(N, pos, vel);
  memcpy(xt, pos, sizeof(real) * 3 * (size_t) N);
  memcpy(vt, vel, sizeof(real) * 3 * (size_t) N);
  const real ds = ds_;
  int64_t live_steps = 0;
  for (int64_t r = 0; r < N; ++r) s.inside[r] = (unsigned char) vol_inbounds(&grid, s.x + 3 * r);
  /* :276-277 only seed `active`, which no output depends on */
  for (int i = 0; i < max_steps; ++i) {
    int all_escaped = 1;
    for (int64_t r = 0; r < N; ++r) {
      real *x = s.x + 3 * r, *v = s.v + 3 * r;
      if (!s.escaped[r]) ++live_steps;
      real n, nx[3], dist, distx[3];
      vol_eval_grad(&grid, x, s.inside[r], &n, nx);                   /* :282 */
      real dsn = ds * n;
      for (int k = 0; k < 3; ++k) v[k] = fma_(dsn, nx[k], v[k]);
      for (int k = 0; k < 3; ++k) x[k] = fma_(ds, v[k], x[k]);
      vol_eval_grad(&sdfv, x, s.inside[r], &dist, distx);             /* :287 */
      int cur_inside = dist < 0;                                      /* :288 */
      int cross = s.inside[r] && !cur_inside;
      s.escaped[r] |= (unsigned char) cross;
      s.escaped[r] |= (unsigned char) vol_escaped(&grid, x, v);
      if (cross)
        for (int k = 0; k < 3; ++k) { xt[3 * r + k] = x[k]; vt[3 * r + k] = v[k]; }
      all_escaped &= s.escaped[r];
      s.inside[r] = (unsigned char) cur_inside;
    }
    if (all_escaped) break;
  }
  fwd_free(&s);
  return live_steps;
}

/* Tracer::trace_cable, src/tracer.cpp:312-382: updates masked by `active`
 * (:353-354), closest approach to `target` recorded like trace_target. */
int64_t SYM(trace_cable)(const real *rif, int res, real radius, real length, const real *pos,
                         const real *vel, const real *target, int64_t N, real ds, real *xt,
                         real *vt, real *dist2) {
  cyl_t cable = cyl_make((real *) rif, res, radius, length);
  int max_steps = (int) (4 * (float) length / (float) ds);           /* :332 */
  fwd_state_t s = fwd_alloc(N, pos, vel);
  memcpy(xt, pos, sizeof(real) * 3 * (size_t) N);
  memcpy(vt, vel, sizeof(real) * 3 * (size_t) N);
  int64_t live_steps = 0;
  for (int64_t r = 0; r < N; ++r) {
    const real *x = s.x + 3 * r, *t = target + 3 * r;
    s.inside[r] = (unsigned char) cyl_inbounds(&cable, x);            /* :344 */
    real d[3] = {x[0] - t[0], x[1] - t[1], x[2] - t[2]};
    dist2[r] = fma_(d[2], d[2], fma_(d[1], d[1], d[0] * d[0]));       /* :340 */
  }
  for (int i = 0; i < max_steps; ++i) {
    int all_escaped = 1;
    for (int64_t r = 0; r < N; ++r) {
      real *x = s.x + 3 * r, *v = s.v + 3 * r;
      const real *t = target + 3 * r;
      if (s.active[r]) ++live_steps;
      real n, nx[3];
      cyl_eval_grad(&cable, x, &n, nx);                               /* :351 */
      if (s.active[r]) {                                              /* :353-354 */
        real dsn = ds * n;
        for (int k = 0; k < 3; ++k) v[k] = fma_(dsn, nx[k], v[k]);
        for (int k = 0; k < 3; ++k) x[k] = fma_(ds, v[k], x[k]);
      }
      real d[3] = {x[0] - t[0], x[1] - t[1], x[2] - t[2]};
      real cur = fma_(d[2], d[2], fma_(d[1], d[1], d[0] * d[0]));     /* :356 */
      int closer = cur < dist2[r];
      int cur_inside = cyl_inbounds(&cable, x);
      int cross = s.inside[r] && !cur_inside;
      s.escaped[r] |= (unsigned char) cross;
      s.escaped[r] |= (unsigned char) cyl_escaped(&cable, x, v);
      s.active[r] &= (unsigned char) !s.escaped[r];
      if (closer) {                                                   /* :365-367 */
        for (int k = 0; k < 3; ++k) { xt[3 * r + k] = x[k]; vt[3 * r + k] = v[k]; }
        dist2[r] = cur;
      }
      all_escaped &= s.escaped[r];
      s.inside[r] = (unsigned char) cur_inside;
    }
    if (all_escaped) break;
  }
  fwd_free(&s);
  return live_steps;
}

/* ------------------------------------------------------------------------ */
/* Adjoint: Tracer::backtrace*, src/tracer.cpp:384-567                        */
/* ------------------------------------------------------------------------ */

typedef struct {
  real *x, *v, *la, *mu;
  unsigned char *active, *outside;
} adj_state_t;

static adj_state_t adj_alloc(int64_t N, const real *xt, const real *vt, const real *dx,
                             const real *dv, real ds) {
  adj_state_t s;
  size_t n3 = 3 * (size_t) (N ? N : 1);
  s.x = (real *) malloc(sizeof(real) * n3);
  s.v = (real *) malloc(sizeof(real) * n3);
  s.la = (real *) malloc(sizeof(real) * n3);
  s.mu = (real *) malloc(sizeof(real) * n3);
  s.active = (unsigned char *) calloc((size_t) (N ? N : 1), 1);
  s.outside = (unsigned char *) calloc((size_t) (N ? N : 1), 1);
  memcpy(s.x, xt, sizeof(real) * 3 * (size_t) N);
  memcpy(s.v, vt, sizeof(real) * 3 * (size_t) N);
  for (int64_t i = 0; i < 3 * N; ++i) {
    s.la[i] = dx[i];                       /* :409 */
    real t = ds * dx[i];
    s.mu[i] = dv[i] + t;                   /* :410 */
  }
  return s;
}

static void adj_free(adj_state_t *s) {
  free(s->x); free(s->v); free(s->la); free(s->mu); free(s->active); free(s->outside);
}

/* la/mu update shared by all three adjoints (:434-435, :503-504, :561-562).
 * Hm is a full row-major 3x3; (n*H)*mu follows matrix.h:170-178 (column sums
 * with fmadd). */
static inline void adj_update(real *la, real *mu, real ds, real dn, real n, const real nx[3],
                              const real Hm[9]) {
  real Hmu[3];
  for (int i = 0; i < 3; ++i) {
    real acc = (n * Hm[3 * i + 0]) * mu[0];
    acc = fma_(n * Hm[3 * i + 1], mu[1], acc);
    acc = fma_(n * Hm[3 * i + 2], mu[2], acc);
    Hmu[i] = acc;
  }
  for (int k = 0; k < 3; ++k) {
    real t = dn * nx[k] + Hmu[k];
    la[k] = la[k] + ds * t;
  }
  for (int k = 0; k < 3; ++k) mu[k] = mu[k] + ds * la[k];
}

/* Shared body of Tracer::backtrace (src/tracer.cpp:384-440) and backtrace_sdf
 * (:443-509; sdf != NULL).  The gradient grid is zero-filled (:401-403) and
 * returned in `grad`.  Within a step the 16 scatter_adds of volume::splat are
 * applied term-major, ray-minor -- the CPU reference's exact summation order
 * (ext/enoki/include/enoki/dynamic.h:518-535). */
static int64_t backtrace_grid(const real *rif, const real *sdf, int W, int H, int D,
                              const real *xt, const real *vt, const real *dx, const real *dv,
                              int64_t N, real h, real ds, real *grad, const real *pos0, real *dpos,
                              real *dvel) {
  int64_t G = (int64_t) W * H * D;
  vol_t grid = vol_make((real *) rif, W, H, D, h);
  vol_t gradv = vol_make(grad, W, H, D, h);
  vol_t sdfv = vol_make((real *) sdf, W, H, D, h);
  memset(grad, 0, sizeof(real) * (size_t) G);
  adj_state_t s = adj_alloc(N, xt, vt, dx, dv, ds);
  int *sidx = (int *) malloc(sizeof(int) * 8 * (size_t) (N ? N : 1));
  real *sterm = (real *) malloc(sizeof(real) * 16 * (size_t) (N ? N : 1));
  /* ray gradients (not in the reference's backtrace; see backtrace_rays below):
   * (x_k, v_k) of the last sample that updated lambda / mu */
  real *xl = 0, *vl = 0, *cap = 0;   /* cap: (la, mu) captured at the launch sample, [6] per ray, + flag */
  unsigned char *got = 0;
  if (dpos && pos0) {
    cap = (real *) malloc(sizeof(real) * 6 * (size_t) (N ? N : 1));
    got = (unsigned char *) calloc((size_t) (N ? N : 1), 1);
  }
  if (dpos) {
    xl = (real *) malloc(sizeof(real) * 3 * (size_t) (N ? N : 1));
    vl = (real *) malloc(sizeof(real) * 3 * (size_t) (N ? N : 1));
    memcpy(xl, xt, sizeof(real) * 3 * (size_t) N);
    memcpy(vl, vt, sizeof(real) * 3 * (size_t) N);
  }
  int64_t live_steps = 0;
  for (int64_t r = 0; r < N; ++r) {
    real nv[3] = {-s.v[3 * r], -s.v[3 * r + 1], -s.v[3 * r + 2]};
    s.active[r] = (unsigned char) !vol_escaped(&grid, s.x + 3 * r, nv);   /* :413-414 */
    if (sdf) {                                                            /* :476-477 */
      real dist, distx[3];
      vol_eval_grad(&sdfv, s.x + 3 * r, s.active[r], &dist, distx);
      s.outside[r] = (unsigned char) (dist >= 0);
    }
  }
  int max_steps = cap_grid(2.0f, (float) h, W, H, D, (float) ds);         /* :417 */
  for (int i = 0; i < max_steps; ++i) {
    int any_active = 0;
    for (int64_t r = 0; r < N; ++r) {
      if (!s.active[r]) continue;   /* inactive lanes never influence an output again */
      ++live_steps;
      real *x = s.x + 3 * r, *v = s.v + 3 * r;
      for (int k = 0; k < 3; ++k) x[k] = fma_(-ds, v[k], x[k]);           /* :420 */
      real n, nx[3], hx[3];
      vol_eval_grad(&grid, x, 1, &n, nx);                                 /* :421 */
      vol_eval_hess(&grid, x, 1, hx);                                     /* :422 */
      real mdsn = -ds * n;
      for (int k = 0; k < 3; ++k) v[k] = fma_(mdsn, nx[k], v[k]);         /* :423 */
      real nv[3] = {-v[0], -v[1], -v[2]};
      int act = !vol_escaped(&grid, x, nv);                               /* :425 */
      if (sdf) {                                                          /* :488-497 */
        real dist, distx[3];
        vol_eval_grad(&sdfv, x, 1, &dist, distx);
        int cross = !s.outside[r] && (dist >= 0);
        act = act && !cross;
        s.outside[r] = (unsigned char) (dist >= 0);
      }
      s.active[r] = (unsigned char) act;
      any_active |= act;
    }
    if (!any_active) break;                                               /* :426-428 */
    for (int64_t r = 0; r < N; ++r) {
      if (!s.active[r]) continue;
      real *x = s.x + 3 * r, *mu = s.mu + 3 * r, *la = s.la + 3 * r;
      real n, nx[3], hx[3];
      vol_eval_grad(&grid, x, 1, &n, nx);
      vol_eval_hess(&grid, x, 1, hx);
      real dn = fma_(mu[2], nx[2], fma_(mu[1], nx[1], mu[0] * nx[0]));    /* :430 */
      real dnx[3] = {n * mu[0] * ds, n * mu[1] * ds, n * mu[2] * ds};     /* :431-432 */
      vol_splat_terms(&gradv, x, dn * ds, dnx, sidx + 8 * r, sterm + 16 * r);
      real Hm[9] = {0, hx[0], hx[1], hx[0], 0, hx[2], hx[1], hx[2], 0};
      adj_update(la, mu, ds, dn, n, nx, Hm);                              /* :434-435 */
      if (dpos) for (int k = 0; k < 3; ++k) { xl[3 * r + k] = x[k]; vl[3 * r + k] = s.v[3 * r + k]; }
      if (got && !got[r]) {
        /* A ray launched INSIDE the region the adjoint integrates through (trace_sdf's rays start inside
         * the object, src/tracer.cpp:271-277) is re-integrated past its launch point; the backward samples
         * retrace the forward ones, so the sample that coincides with pos0 (within half a step) is x_0 and
         * (la, mu) after it are dL/dx_0, dL/dv_0. */
        const real *p0 = pos0 + 3 * r, *vv = s.v + 3 * r;
        real d2 = (x[0] - p0[0]) * (x[0] - p0[0]) + (x[1] - p0[1]) * (x[1] - p0[1]) + (x[2] - p0[2]) * (x[2] - p0[2]);
        real v2 = vv[0] * vv[0] + vv[1] * vv[1] + vv[2] * vv[2];
        if (d2 <= (real) 0.25 * ds * ds * v2) {
          got[r] = 1;
          for (int k = 0; k < 3; ++k) { cap[6 * r + k] = la[k]; cap[6 * r + 3 + k] = mu[k]; }
        }
      }
    }
    for (int t = 0; t < 16; ++t)                                          /* src/volume.cpp:217-243 */
      for (int64_t r = 0; r < N; ++r)
        if (s.active[r]) grad[sidx[8 * r + (t & 7)]] += sterm[16 * r + t];
  }
  if (dpos) {
    for (int64_t r = 0; r < N; ++r) {
      real k = 0;
      const real *a = xl + 3 * r, *b = vl + 3 * r;
      if (pos0) {
        real num = (a[0] - pos0[3 * r]) * b[0] + (a[1] - pos0[3 * r + 1]) * b[1] + (a[2] - pos0[3 * r + 2]) * b[2];
        real den = ds * (b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
        k = den > 0 ? (real) rint((double) (num / den)) : 0;
      }
      real t = (k - 1) * ds;
      const real *la = s.la + 3 * r, *mu = s.mu + 3 * r;
      if (got && got[r]) { la = cap + 6 * r; mu = cap + 6 * r + 3; t = -ds; }
      for (int c = 0; c < 3; ++c) {
        dpos[3 * r + c] = la[c];
        dvel[3 * r + c] = mu[c] + t * la[c];
      }
    }
    free(xl); free(vl); free(cap); free(got);
  }
  free(sidx); free(sterm);
  adj_free(&s);
  return live_steps;
}

/* NOT a reference function: the adjoint's final co-state mapped to dL/dpos, dL/dvel of
 * the launch state -- what ADTracerC.backward returns through enoki's tape (reference
 * core/tracer.py:53-60).  After sample x_k has been processed la = dL/dx_k and
 * mu = dL/dv_{k-1}; before x_k the forward ray flew straight (masked gathers,
 * src/tracer.cpp:66-71), so dL/dpos = la and dL/dvel = mu + (k-1) ds la with k the
 * number of straight steps from pos0 to x_k (k = 0 without pos0).  Pinned against a
 * float64 tape differentiation of the forward loop in tests (oracle/ad_oracle.py). */
int64_t SYM(backtrace_rays)(const real *rif, const real *sdf, int W, int H, int D, const real *xt,
                            const real *vt, const real *dx, const real *dv, const real *pos0,
                            int64_t N, real h, real ds, real *grad, real *dpos, real *dvel) {
  return backtrace_grid(rif, sdf, W, H, D, xt, vt, dx, dv, N, h, ds, grad, pos0, dpos, dvel);
}

int64_t SYM(backtrace)(const real *rif, int W, int H, int D, const real *xt, const real *vt,
                       const real *dx, const real *dv, int64_t N, real h, real ds, real *grad) {
  return backtrace_grid(rif, 0, W, H, D, xt, vt, dx, dv, N, h, ds, grad, 0, 0, 0);
}

int64_t SYM(backtrace_sdf)(const real *rif, const real *sdf, int W, int H, int D, const real *xt,
                           const real *vt, const real *dx, const real *dv, int64_t N, real h,
                           real ds, real *grad) {
  return backtrace_grid(rif, sdf, W, H, D, xt, vt, dx, dv, N, h, ds, grad, 0, 0, 0);
}

/* Tracer::backtrace_cable, src/tracer.cpp:511-567 */
int64_t SYM(backtrace_cable)(const real *rif, int res, real radius, real length, const real *xt,
                             const real *vt, const real *dx, const real *dv, int64_t N, real ds,
                             real *grad) {
  cyl_t cable = cyl_make((real *) rif, res, radius, length);
  cyl_t gradc = cyl_make(grad, res, radius, length);
  memset(grad, 0, sizeof(real) * (size_t) res);
  adj_state_t s = adj_alloc(N, xt, vt, dx, dv, ds);
  int *sidx = (int *) malloc(sizeof(int) * 2 * (size_t) (N ? N : 1));
  real *sterm = (real *) malloc(sizeof(real) * 4 * (size_t) (N ? N : 1));
  int64_t live_steps = 0;
  for (int64_t r = 0; r < N; ++r) {
    real nv[3] = {-s.v[3 * r], -s.v[3 * r + 1], -s.v[3 * r + 2]};
    s.active[r] = (unsigned char) !cyl_escaped(&cable, s.x + 3 * r, nv);  /* :540-541 */
  }
  int max_steps = (int) (4 * (float) length / (float) ds);               /* :544 */
  for (int i = 0; i < max_steps; ++i) {
    int any_active = 0;
    for (int64_t r = 0; r < N; ++r) {
      if (!s.active[r]) continue;
      ++live_steps;
      real *x = s.x + 3 * r, *v = s.v + 3 * r;
      for (int k = 0; k < 3; ++k) x[k] = fma_(-ds, v[k], x[k]);           /* :547 */
      real n, nx[3];
      cyl_eval_grad(&cable, x, &n, nx);                                   /* :548 */
      real mdsn = -ds * n;
      for (int k = 0; k < 3; ++k) v[k] = fma_(mdsn, nx[k], v[k]);         /* :550 */
      real nv[3] = {-v[0], -v[1], -v[2]};
      s.active[r] = (unsigned char) !cyl_escaped(&cable, x, nv);          /* :552 */
      any_active |= s.active[r];
    }
    if (!any_active) break;
    for (int64_t r = 0; r < N; ++r) {
      if (!s.active[r]) continue;
      real *x = s.x + 3 * r, *mu = s.mu + 3 * r, *la = s.la + 3 * r;
      real n, nx[3], Hm[9];
      cyl_eval_grad(&cable, x, &n, nx);
      cyl_eval_hess(&cable, x, Hm);                                       /* :549 */
      real dn = fma_(mu[2], nx[2], fma_(mu[1], nx[1], mu[0] * nx[0]));    /* :557 */
      real dnx[3] = {n * mu[0] * ds, n * mu[1] * ds, n * mu[2] * ds};
      cyl_splat_terms(&gradc, x, dn * ds, dnx, sidx + 2 * r, sterm + 4 * r);  /* :559 */
      adj_update(la, mu, ds, dn, n, nx, Hm);                              /* :561-562 */
    }
    for (int t = 0; t < 4; ++t)                                           /* src/cylinder_volume.cpp:139-147 */
      for (int64_t r = 0; r < N; ++r)
        if (s.active[r]) grad[sidx[2 * r + (t & 1)]] += sterm[4 * r + t];
  }
  free(sidx); free(sterm);
  adj_free(&s);
  return live_steps;
}

/* ------------------------------------------------------------------------ */
/* Probes of the volume primitives (fixtures of src/test.cpp:61-101)          */
/* ------------------------------------------------------------------------ */

void SYM(volume_eval)(const real *rif, int W, int H, int D, const real *pos, int64_t N, real h,
                      real *n, real *grad, real *hess) {
  vol_t g = vol_make((real *) rif, W, H, D, h);
  for (int64_t r = 0; r < N; ++r) {
    vol_eval_grad(&g, pos + 3 * r, 1, n + r, grad + 3 * r);
    vol_eval_hess(&g, pos + 3 * r, 1, hess + 3 * r);
  }
}

void SYM(volume_splat)(real *data, int W, int H, int D, const real *pos, const real *val,
                       const real *grad, int64_t N, real h) {
  vol_t g = vol_make(data, W, H, D, h);
  int *sidx = (int *) malloc(sizeof(int) * 8 * (size_t) (N ? N : 1));
  real *sterm = (real *) malloc(sizeof(real) * 16 * (size_t) (N ? N : 1));
  for (int64_t r = 0; r < N; ++r)
    vol_splat_terms(&g, pos + 3 * r, val[r], grad + 3 * r, sidx + 8 * r, sterm + 16 * r);
  for (int t = 0; t < 16; ++t)
    for (int64_t r = 0; r < N; ++r) data[sidx[8 * r + (t & 7)]] += sterm[16 * r + t];
  free(sidx); free(sterm);
}

void SYM(cylinder_eval)(const real *rif, int res, real radius, real length, const real *pos,
                        int64_t N, real *n, real *grad, real *hess) {
  cyl_t c = cyl_make((real *) rif, res, radius, length);
  for (int64_t r = 0; r < N; ++r) {
    cyl_eval_grad(&c, pos + 3 * r, n + r, grad + 3 * r);
    cyl_eval_hess(&c, pos + 3 * r, hess + 9 * r);
  }
}

void SYM(cylinder_splat)(real *data, int res, real radius, real length, const real *pos,
                         const real *val, const real *grad, int64_t N) {
  cyl_t c = cyl_make(data, res, radius, length);
  for (int t = 0; t < 4; ++t)
    for (int64_t r = 0; r < N; ++r) {
      int idx[2]; real term[4];
      cyl_splat_terms(&c, pos + 3 * r, val[r], grad + 3 * r, idx, term);
      data[idx[t & 1]] += term[t];
    }
}
