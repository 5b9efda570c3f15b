// on-device ray generation -- the step before every trace in the
// reference's drivers, which build the batch with CPU torch ops and copy it to the GPU each
// iteration (reference core/luneburg_opt.py:53-59, core/image_opt.py:44-52).
//
//   plane_source3_rand     reference core/source.py:54-69  (jittered pixel grid on a plane)
//   point_source3_rand     reference core/source.py:72-104 (point source through a jittered pixel grid)
//   point_source3          reference core/source.py:29-51  (point source, regular lattice of directions)
//   cone_source3_rand      reference core/source.py:186-202 + hatbox_sample :531-545 (uniform cone)
//   area_source3_cone      reference core/source.py:153-183 (cone per jittered pixel; rand_cone_in_sphere)
//   area_source3_rand_bias reference core/source.py:107-150 (pixel -> random sensor point; rand_area_in_sphere)
//   rotate_pts_to_source   reference core/source.py:275-293 (view rotation, launch plane,
//                                                          sensor plane triple)
//   rand_rays_in_sphere    reference core/source.py:352-357 (one plane source per view)
//   rand_rays_cube         reference core/source.py:398-412 (4 + 2 axis views)
//   random_rotate_ic       reference core/source.py:555-563 (one rigid rotation of the batch)
//
// One thread per ray writes x, v and the (3,3) plane triple: 60 B per ray, HBM-write bound.
// The jitter is either read from a caller-supplied uniform array (the layout of the
// reference's torch.rand(2*spp, P0, P1) per view -- used by the parity tests) or drawn
// in-kernel from Philox4x32-10 keyed by (seed, view, sample index).
#include "drrt_device.cuh"
#include "drrt_launch.h"

namespace drrt {

namespace {

__device__ __forceinline__ void philox_round(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3, uint32_t k0,
                                             uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
  const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
  c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
}

// Philox4x32-10 (Salmon et al., SC'11): 4 x 32 random bits for one 128-bit counter.
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    philox_round(c0, c1, c2, c3, k0, k1);
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// uniform in [0, 1) with 24 random bits, like torch.rand for float32
__device__ __forceinline__ float u01(uint32_t bits) { return (float) (bits >> 8) * (1.0f / 16777216.0f); }

__device__ __forceinline__ f3 rot(const float *R, f3 a) {  // a @ R^T = R a, R row-major
  return f3{R[0] * a.x + R[1] * a.y + R[2] * a.z, R[3] * a.x + R[4] * a.y + R[5] * a.z,
            R[6] * a.x + R[7] * a.y + R[8] * a.z};
}

// The 15 output floats of a ray (x, v, the 3 x 3 plane triple; AoS like the reference's tensors) are staged in shared
// memory per CTA and written back with consecutive threads on consecutive floats: 15 coalesced stores of 4 sectors per
// warp instead of 15 strided ones of 12 (x, v) and 36 (planes) sectors.
constexpr int kSourceThreads = 256;

__device__ __forceinline__ void source_ray(const SourceArgs &a, int64_t r, float *sx, float *sv, float *sp);

__global__ void __launch_bounds__(kSourceThreads) plane_source_kernel(SourceArgs a) {
  __shared__ float sx[3 * kSourceThreads], sv[3 * kSourceThreads], sp[9 * kSourceThreads];
  const int64_t n = a.rays_per_view * a.nviews;
  const int64_t r0 = (int64_t) blockIdx.x * kSourceThreads;
  const int64_t r = r0 + threadIdx.x;
  if (r < n) source_ray(a, r, sx + 3 * threadIdx.x, sv + 3 * threadIdx.x, sp + 9 * threadIdx.x);
  __syncthreads();
  const int cnt = (int) (n - r0 < kSourceThreads ? n - r0 : kSourceThreads);
  float *gx = a.x + 3 * r0, *gv = a.v + 3 * r0, *gp = a.planes + 9 * r0;
  for (int i = threadIdx.x; i < 3 * cnt; i += kSourceThreads) { gx[i] = sx[i]; gv[i] = sv[i]; }
  for (int i = threadIdx.x; i < 9 * cnt; i += kSourceThreads) gp[i] = sp[i];
}

__device__ __forceinline__ void source_ray(const SourceArgs &a, int64_t r, float *ox, float *ov, float *op) {
  const int view = (int) (r / a.rays_per_view);
  const int64_t q = r - (int64_t) view * a.rays_per_view;  // (s * P0 + i) * P1 + j
  const int j = (int) (q % a.P1);
  const int i = (int) ((q / a.P1) % a.P0);
  float ua, ub, uc = 0.f, ud = 0.f;
  const bool area = a.kind == DRRT_SOURCE_AREA_CONE || a.kind == DRRT_SOURCE_AREA_BIAS;
  if (a.kind == DRRT_SOURCE_POINT_LATTICE) {
    ua = ub = 0.f;  // no jitter: a regular lattice
  } else if (a.offset) {  // torch.rand(2*spp, P0, P1) of this view: [:spp] jitters axis 0, [spp:] axis 1;
                   // the area sources draw two more planes per view (rand(n) twice / rand(2, n))
    const float *o = a.offset + (int64_t) (a.first_view + view) * (area ? 4 : 2) * a.rays_per_view;
    ua = o[q]; ub = o[a.rays_per_view + q];
    if (area) { uc = o[2 * a.rays_per_view + q]; ud = o[3 * a.rays_per_view + q]; }
  } else {
    const uint4 b = philox4x32_10((uint32_t) q, (uint32_t) (q >> 32), (uint32_t) (a.first_view + view), 0u,
                                  (uint32_t) a.seed, (uint32_t) (a.seed >> 32));
    ua = u01(b.x); ub = u01(b.y); uc = u01(b.z); ud = u01(b.w);
  }
  const float w = a.width;
  const float *R = a.view_rot + 9 * view;
  const float hw = 0.5f * w;
  const f3 pn = rot(R, f3{0.f, 1.f, 0.f});  // plane normal = the view's optical axis
  f3 t = rot(R, f3{0.f, 0.f, 1.f});
  f3 x, v;
  float pd = a.sensor_dist + hw;            // plane_x = pd * plane_v + w/2 (:290, :198-199)
  bool inside = true;
  if (a.kind == DRRT_SOURCE_PLANE) {
    f3 p;
    p.y = 0.f;
    if (a.independent) {  // core/source.py:60-63
      p.x = ua * w - 0.5f * w;
      p.z = ub * w - 0.5f * w;
    } else {              // core/source.py:56-57, :65-68
      p.x = w * ((float) i / (float) a.P0 - 0.5f) + (ua * w) / (float) a.P0;
      p.z = w * ((float) j / (float) a.P1 - 0.5f) + (ub * w) / (float) a.P1;
    }
    inside = sqrtf(p.x * p.x + p.z * p.z) < hw;  // circle: core/source.py:277-280
    x = rot(R, p);                          // :284
    v = pn;                                 // :285
    x.x += hw; x.y += hw; x.z += hw;
    x.x -= w * v.x * 0.5f; x.y -= w * v.y * 0.5f; x.z -= w * v.z * 0.5f;   // :287
  } else if (area) {
    // area_source3_cone (core/source.py:153-183) / area_source3_rand_bias (:107-150): jittered pixel
    // centres, jitter scaled by width / pixels[0] on BOTH axes (:109, :155)
    const float js = w / (float) a.P0;
    const float pa = w * (((float) i + 0.5f) / (float) a.P0 - 0.5f) + (ua - 0.5f) * js;
    const float pb = w * (((float) j + 0.5f) / (float) a.P1 - 0.5f) + (ub - 0.5f) * js;
    f3 pos, d, tgt = {0.f, 0.f, 0.f};
    float tp;
    if (a.kind == DRRT_SOURCE_AREA_CONE) {
      pos = f3{pa, -hw, pb};
      inside = sqrtf(pa * pa + hw * hw + pb * pb) < hw;   // 3-D norm incl. y = -w/2, as written (:163-165)
      const float z = uc * (1.f - a.cone_cos) + a.cone_cos;  // hatbox_sample about (0,1,0), :531-545
      float sn, cs;
      sincosf(6.283185307179586f * ud, &sn, &cs);
      const float sc = sqrtf(fmaxf(1.f - z * z, 0.f));
      d = f3{-cs * sc, z, -sn * sc};
      tp = a.sensor_dist / d.y;                             // :170
    } else {
      inside = sqrtf(pa * pa + pb * pb) < hw;               // :118-120
      tgt = f3{-pa, a.sensor_dist + hw, -pb};               // pt = -pos + (sd + w/2) e_y, :122-124
      const f3 p0 = {pa, -(a.sensor_dist + hw), pb};        // :123
      const f3 to = {(uc - 0.5f) * w - p0.x, hw - p0.y, (ud - 0.5f) * w - p0.z};   // :126-133
      const float inv = 1.f / sqrtf(to.x * to.x + to.y * to.y + to.z * to.z);
      d = f3{to.x * inv, to.y * inv, to.z * inv};
      tp = a.sensor_dist / d.y;                             // :136
      pos = f3{fmaf(tp, d.x, p0.x), fmaf(tp, d.y, p0.y), fmaf(tp, d.z, p0.z)};     // :137
      tgt = rot(R, tgt); tgt.x += hw; tgt.y += hw; tgt.z += hw;                    // :139
    }
    x = rot(R, pos);
    x.x += hw; x.y += hw; x.z += hw;
    v = rot(R, d);
    if (a.tpv) a.tpv[r] = tp;
    if (a.target) {
      if (a.has_post) { const float c = 0.5f * a.post_span; tgt = rot(a.post_rot, f3{tgt.x - c, tgt.y - c, tgt.z - c}); tgt.x += c; tgt.y += c; tgt.z += c; }
      a.target[3 * r] = tgt.x; a.target[3 * r + 1] = tgt.y; a.target[3 * r + 2] = tgt.z;
    }
  } else {
    f3 d;
    if (a.kind == DRRT_SOURCE_POINT_LATTICE) {  // point_source3, core/source.py:29-51
      // theta = linspace(-a, a, P0)[i], phi = linspace(-a, a, P1)[j] (meshgrid 'ij'), the caller passes
      // P0 = pixels[0] * floor(sqrt(spp)), P1 = pixels[1] * floor(sqrt(spp)), spp = 1 (:30-33)
      const float ang = a.cone_half;
      const float th = a.P0 > 1 ? fmaf(2.f * ang, (float) i / (float) (a.P0 - 1), -ang) : -ang;
      const float ph = a.P1 > 1 ? fmaf(2.f * ang, (float) j / (float) (a.P1 - 1), -ang) : -ang;
      float st, ct, sp, cp;
      sincosf(th, &st, &ct); sincosf(ph, &sp, &cp);
      d = f3{ct * sp, ct * cp, st};                                  // :34-36
      const float inv = 1.f / sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);  // :40
      d = f3{d.x * inv, d.y * inv, d.z * inv};
    } else if (a.kind == DRRT_SOURCE_POINT) {      // point_source3_rand, core/source.py:72-104
      // pixel centres jittered by (u - 0.5) -- not scaled by the pixel size in the reference (:73-79)
      const float pa = w * (((float) i + 0.5f) / (float) a.P0 - 0.5f) + (ua - 0.5f);
      const float pb = w * (((float) j + 0.5f) / (float) a.P1 - 0.5f) + (ub - 0.5f);
      inside = sqrtf(pa * pa + pb * pb) < hw;                       // :81-83
      const float inv = 1.f / sqrtf(pa * pa + w * w + pb * pb);     // :85-89
      d = f3{pa * inv, w * inv, pb * inv};
      pd = a.sensor_dist * w * 0.5f;                                // :101
    } else {                                // cone_source3_rand + hatbox_sample, :186-202, :531-545
      const float z = ua * (1.f - a.cone_cos) + a.cone_cos;
      float sn, cs;
      sincosf(6.283185307179586f * ub, &sn, &cs);
      const float sc = sqrtf(fmaxf(1.f - z * z, 0.f));
      // t1 = (0,0,1) x (0,1,0) = (-1,0,0), t2 = t1 x (0,1,0) = (0,0,-1)
      d = f3{-cs * sc, z, -sn * sc};
    }
    x = rot(R, f3{0.f, -hw, 0.f});
    x.x += hw; x.y += hw; x.z += hw;
    v = rot(R, d);
  }
  if (a.keep) a.keep[r] = inside ? 1 : 0;
  f3 px = {pd * pn.x + hw, pd * pn.y + hw, pd * pn.z + hw};
  f3 pnr = pn;
  if (a.has_post) {  // random_rotate_ic, core/source.py:555-563: rotate about the volume centre span / 2
    const float c = 0.5f * a.post_span;
    f3 d = {x.x - c, x.y - c, x.z - c};
    x = rot(a.post_rot, d); x.x += c; x.y += c; x.z += c;
    d = f3{px.x - c, px.y - c, px.z - c};
    px = rot(a.post_rot, d); px.x += c; px.y += c; px.z += c;
    v = rot(a.post_rot, v);
    t = rot(a.post_rot, t);
    pnr = rot(a.post_rot, pn);
  }
  ox[0] = x.x; ox[1] = x.y; ox[2] = x.z;
  ov[0] = v.x; ov[1] = v.y; ov[2] = v.z;
  op[0] = px.x; op[1] = px.y; op[2] = px.z;   // planes[:, 0] sensor origin
  op[3] = pnr.x; op[4] = pnr.y; op[5] = pnr.z;  // planes[:, 1] sensor normal
  op[6] = t.x; op[7] = t.y; op[8] = t.z;      // planes[:, 2] sensor tangent
}

}  // namespace

int launch_plane_source(const SourceArgs &a, cudaStream_t s) {
  const int64_t n = a.rays_per_view * a.nviews;
  if (n <= 0) return 0;
  plane_source_kernel<<<(unsigned) ((n + kSourceThreads - 1) / kSourceThreads), kSourceThreads, 0, s>>>(a);
  return 1;
}

}  // namespace drrt
