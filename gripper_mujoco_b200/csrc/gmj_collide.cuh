// gmj_collide.cuh -- narrow phase for the gripper scene: support maps of box / cylinder /
// ellipsoid (+ sphere sweep for filleted objects), Minkowski Portal Refinement penetration query
// (the algorithm MuJoCo 2.1.5 runs through libccd for convex pairs, tolerance 1e-6, 50 iterations)
// and closed-form ground-plane contacts. One lane tests one gripper geom against the object, so a
// warp covers all 3N segment plates + palm in one pass and the 3 hook plates in a second.
#pragma once
#include "gmj_device.cuh"

template <typename T>
struct Shape { int type; V3<T> size; T round; V3<T> pos; M3<T> R; };

#ifndef GMJ_CINL
#define GMJ_CINL inline
#endif
template <typename T> __device__ __forceinline__ T shape_bound(const Shape<T>& s) {
  T r;
  if (s.type == GMJ_GEOM_BOX) r = sqrtT<T>(dot(s.size, s.size));
  else if (s.type == GMJ_GEOM_CYLINDER) r = sqrtT<T>(s.size.x * s.size.x + s.size.z * s.size.z);
  else r = maxT<T>(s.size.x, maxT<T>(s.size.y, s.size.z));
  return r + s.round;
}

template <typename T> __device__ inline V3<T> support(const Shape<T>& s, V3<T> d) {
  V3<T> dl = mulT(s.R, d), p;
  if (s.type == GMJ_GEOM_BOX) {
    p = mk<T>(dl.x >= 0 ? s.size.x : -s.size.x, dl.y >= 0 ? s.size.y : -s.size.y, dl.z >= 0 ? s.size.z : -s.size.z);
  } else if (s.type == GMJ_GEOM_CYLINDER) {
    T l = sqrtT<T>(dl.x * dl.x + dl.y * dl.y);
    if (l > T(1e-12)) { const T rl = rcpT<T>(l); p.x = s.size.x * dl.x * rl; p.y = s.size.x * dl.y * rl; } else { p.x = 0; p.y = 0; }
    p.z = dl.z >= 0 ? s.size.z : -s.size.z;
  } else {
    T a = s.size.x, b = s.size.y, c = s.size.z;
    T l = sqrtT<T>(a * a * dl.x * dl.x + b * b * dl.y * dl.y + c * c * dl.z * dl.z);
    if (l < T(1e-30)) l = T(1e-30);
    const T rl = rcpT<T>(l);
    p = mk<T>(a * a * dl.x * rl, b * b * dl.y * rl, c * c * dl.z * rl);
  }
  if (s.round > 0) p = p + dl * s.round;
  return mul(s.R, p) + s.pos;
}

template <typename T> struct MV { V3<T> v, a, b; };
template <typename T> __device__ __forceinline__ MV<T> msupport(const Shape<T>& A, const Shape<T>& B, V3<T> dir) {
  MV<T> o;
  o.a = support(A, dir);
  o.b = support(B, dir * T(-1));
  o.v = o.a - o.b;
  return o;
}

// closest point of triangle (a,b,c) to the origin: barycentric weights (Ericson, RTCD 5.1.5)
template <typename T> __device__ inline void tri_closest_origin(V3<T> a, V3<T> b, V3<T> c, T* w) {
  V3<T> ab = b - a, ac = c - a, ap = a * T(-1), bp = b * T(-1), cp = c * T(-1);
  T d1 = dot(ab, ap), d2 = dot(ac, ap);
  if (d1 <= 0 && d2 <= 0) { w[0] = 1; w[1] = 0; w[2] = 0; return; }
  T d3 = dot(ab, bp), d4 = dot(ac, bp);
  if (d3 >= 0 && d4 <= d3) { w[0] = 0; w[1] = 1; w[2] = 0; return; }
  T vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) { T v = d1 / (d1 - d3); w[0] = 1 - v; w[1] = v; w[2] = 0; return; }
  T d5 = dot(ab, cp), d6 = dot(ac, cp);
  if (d6 >= 0 && d5 <= d6) { w[0] = 0; w[1] = 0; w[2] = 1; return; }
  T vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) { T v = d2 / (d2 - d6); w[0] = 1 - v; w[1] = 0; w[2] = v; return; }
  T va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) { T v = (d4 - d3) / ((d4 - d3) + (d5 - d6)); w[0] = 0; w[1] = 1 - v; w[2] = v; return; }
  T den = rcpT<T>(va + vb + vc);
  w[1] = vb * den; w[2] = vc * den; w[0] = 1 - w[1] - w[2];
}

#define GMJ_MPR_TOL 1e-6
#define GMJ_MPR_MAXIT 50

// 1 on penetration; normal points from A to B, depth > 0, pos = contact point
template <typename T>
__device__ GMJ_CINL int mpr_penetration(const Shape<T>& A, const Shape<T>& B, T* depth, V3<T>* normal, V3<T>* pos) {
  MV<T> v0, v1, v2, v3, v4;
  V3<T> n;
  v0.v = A.pos - B.pos; v0.a = A.pos; v0.b = B.pos;
  if (dot(v0.v, v0.v) < T(1e-20)) v0.v = mk<T>(T(1e-5), 0, 0);
  n = normalized(v0.v * T(-1));
  v1 = msupport(A, B, n);
  if (dot(v1.v, n) <= 0) return 0;
  n = cross(v1.v, v0.v);
  if (dot(n, n) < T(1e-24)) {
    n = normalized(v1.v - v0.v);
    *depth = dot(v1.v, n); *normal = n; *pos = (v1.a + v1.b) * T(0.5);
    return 1;
  }
  n = normalized(n);
  v2 = msupport(A, B, n);
  if (dot(v2.v, n) <= 0) return 0;
  n = cross(v1.v - v0.v, v2.v - v0.v);
  if (dot(n, v0.v) > 0) { MV<T> t = v1; v1 = v2; v2 = t; n = n * T(-1); }
  for (int it = 0;; it++) {
    if (it > GMJ_MPR_MAXIT) return 0;
    if (dot(n, n) < T(1e-30)) return 0;
    n = normalized(n);
    v3 = msupport(A, B, n);
    if (dot(v3.v, n) <= 0) return 0;
    if (dot(cross(v1.v, v3.v), v0.v) < 0) { v2 = v3; n = cross(v1.v - v0.v, v3.v - v0.v); continue; }
    if (dot(cross(v3.v, v2.v), v0.v) < 0) { v1 = v3; n = cross(v3.v - v0.v, v2.v - v0.v); continue; }
    break;
  }
  int inside = 0;
  for (int it = 0; it < GMJ_MPR_MAXIT; it++) {
    n = cross(v2.v - v1.v, v3.v - v1.v);
    if (dot(n, n) < T(1e-30)) break;
    n = normalized(n);
    if (dot(n, v1.v) >= 0) inside = 1;
    v4 = msupport(A, B, n);
    T d4 = dot(v4.v, n);
    if (!inside && d4 < 0) return 0;
    T reach = minT<T>(d4 - dot(v1.v, n), minT<T>(d4 - dot(v2.v, n), d4 - dot(v3.v, n)));
    if (reach <= T(GMJ_MPR_TOL)) break;
    V3<T> t1 = cross(v4.v, v0.v);
    if (dot(v1.v, t1) > 0) { if (dot(v2.v, t1) > 0) v1 = v4; else v3 = v4; }
    else { if (dot(v3.v, t1) > 0) v2 = v4; else v1 = v4; }
  }
  if (!inside) return 0;
  T w[3];
  tri_closest_origin(v1.v, v2.v, v3.v, w);
  V3<T> p = v1.v * w[0] + v2.v * w[1] + v3.v * w[2];
  *pos = ((v1.a + v1.b) * w[0] + (v2.a + v2.b) * w[1] + (v3.a + v3.b) * w[2]) * T(0.5);
  T d = sqrtT<T>(dot(p, p));
  if (d > T(1e-12)) *normal = p * rcpT<T>(d);
  else {
    V3<T> nn = normalized(cross(v2.v - v1.v, v3.v - v1.v));
    if (dot(nn, v0.v) > 0) nn = nn * T(-1);
    *normal = nn;
  }
  *depth = d;
  return 1;
}

// mju_makeFrame: rows normal, t1, t2
template <typename T> __device__ inline void make_frame(V3<T> n, T* f) {
  V3<T> y = (absT<T>(n.y) < T(0.5)) ? mk<T>(0, 1, 0) : mk<T>(0, 0, 1);
  T d = dot(n, y);
  y = normalized(y - n * d);
  V3<T> z = cross(n, y);
  f[0] = n.x; f[1] = n.y; f[2] = n.z; f[3] = y.x; f[4] = y.y; f[5] = y.z; f[6] = z.x; f[7] = z.y; f[8] = z.z;
}

// candidate plane points of one shape against the ground plane z = gz (normal +z). Returns count;
// pts[i] = point on the swept surface, d[i] = signed distance.
template <typename T>
__device__ GMJ_CINL int plane_candidates(const Shape<T>& s, T gz, V3<T>* pts, T* d) {
  int n = 0;
  if (s.pos.z - gz - shape_bound(s) > 0) return 0;
  if (s.type == GMJ_GEOM_ELLIPSOID) {
    V3<T> p = support(s, mk<T>(0, 0, -1));
    pts[0] = p; d[0] = p.z - gz; n = 1;
  } else if (s.type == GMJ_GEOM_BOX) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      V3<T> c = mk<T>((i & 1 ? 1 : -1) * s.size.x, (i & 2 ? 1 : -1) * s.size.y, (i & 4 ? 1 : -1) * s.size.z);
      V3<T> p = mul(s.R, c) + s.pos;
      p.z -= s.round;
      pts[n] = p; d[n] = p.z - gz; n++;
    }
  } else {
    V3<T> axis = mk<T>(s.R.m[2], s.R.m[5], s.R.m[8]);
    T prj = -axis.z;
    if (prj < 0) { axis = axis * T(-1); prj = -prj; }
    V3<T> vec = mk<T>(-axis.x * prj, -axis.y * prj, T(-1) - axis.z * prj);
    T len = sqrtT<T>(dot(vec, vec));
    if (len < T(1e-12)) { vec = mk<T>(s.R.m[0], s.R.m[3], s.R.m[6]); len = 1; }
    vec = vec * (s.size.x * rcpT<T>(len));
    V3<T> ax = axis * s.size.z;
    V3<T> p = s.pos + ax + vec; p.z -= s.round; pts[n] = p; d[n] = p.z - gz; n++;
    p = s.pos - ax + vec; p.z -= s.round; pts[n] = p; d[n] = p.z - gz; n++;
    V3<T> v1 = normalized(cross(vec, axis)) * (s.size.x * T(0.8660254037844386));
    p = s.pos + ax - vec * T(0.5) - v1; p.z -= s.round; pts[n] = p; d[n] = p.z - gz; n++;
    p = s.pos + ax - vec * T(0.5) + v1; p.z -= s.round; pts[n] = p; d[n] = p.z - gz; n++;
  }
  return n;
}

// Candidate `i` of plane_candidates() computed alone (one lane per candidate): returns the candidate count n (uniform
// over lanes); p / d are set when i < n, d = 1 (not penetrating) otherwise. Same arithmetic as plane_candidates().
template <typename T>
__device__ GMJ_CINL int plane_candidate_lane(const Shape<T>& s, T gz, int i, V3<T>& p, T& d) {
  d = T(1); p = mk<T>(0, 0, 0);
  if (s.pos.z - gz - shape_bound(s) > 0) return 0;
  int n;
  if (s.type == GMJ_GEOM_ELLIPSOID) {
    n = 1;
    if (i == 0) { p = support(s, mk<T>(0, 0, -1)); d = p.z - gz; }
  } else if (s.type == GMJ_GEOM_BOX) {
    n = 8;
    if (i < 8) {
      V3<T> c = mk<T>((i & 1 ? 1 : -1) * s.size.x, (i & 2 ? 1 : -1) * s.size.y, (i & 4 ? 1 : -1) * s.size.z);
      p = mul(s.R, c) + s.pos;
      p.z -= s.round;
      d = p.z - gz;
    }
  } else {
    n = 4;
    V3<T> axis = mk<T>(s.R.m[2], s.R.m[5], s.R.m[8]);
    T prj = -axis.z;
    if (prj < 0) { axis = axis * T(-1); prj = -prj; }
    V3<T> vec = mk<T>(-axis.x * prj, -axis.y * prj, T(-1) - axis.z * prj);
    T len = sqrtT<T>(dot(vec, vec));
    if (len < T(1e-12)) { vec = mk<T>(s.R.m[0], s.R.m[3], s.R.m[6]); len = 1; }
    vec = vec * (s.size.x * rcpT<T>(len));
    V3<T> ax = axis * s.size.z;
    if (i < 4) {
      if (i == 0) p = s.pos + ax + vec;
      else if (i == 1) p = s.pos - ax + vec;
      else {
        V3<T> v1 = normalized(cross(vec, axis)) * (s.size.x * T(0.8660254037844386));
        p = (i == 2) ? s.pos + ax - vec * T(0.5) - v1 : s.pos + ax - vec * T(0.5) + v1;
      }
      p.z -= s.round;
      d = p.z - gz;
    }
  }
  return n;
}
