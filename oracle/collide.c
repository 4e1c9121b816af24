/* collide.c -- oracle narrow phase for the scene class of the gripper task.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * The reference delegates collision to MuJoCo 2.1.5 (mj_collision inside mj_step1,
 * src/myfunctions.cpp:1893): plane-vs-convex for the ground and libccd MPR
 * (ccdMPRPenetration, tolerance 1e-6, 50 iterations) for every mesh pair, one contact per
 * convex pair. The graspable meshes are restated as analytic primitives (box / cylinder /
 * ellipsoid + sphere sweep for the fillet), so this file is: support maps, an MPR penetration
 * query over them (Minkowski Portal Refinement, G. Snethen, "XenoCollide", Game Programming
 * Gems 7 -- the algorithm libccd implements) and closed-form plane contacts.
 * PARITY UNPINNED against MuJoCo (library absent).
 */
#include <math.h>
#include <string.h>
#include "oracle.h"
#include "vec.h"

typedef struct { int type; double size[3]; double round; double pos[3]; double R[9]; } Shape;

static double shape_bound(const Shape* s) {
  double r;
  if (s->type == GMJ_GEOM_BOX) r = sqrt(s->size[0] * s->size[0] + s->size[1] * s->size[1] + s->size[2] * s->size[2]);
  else if (s->type == GMJ_GEOM_CYLINDER) r = sqrt(s->size[0] * s->size[0] + s->size[2] * s->size[2]);
  else r = fmax(s->size[0], fmax(s->size[1], s->size[2]));
  return r + s->round;
}

/* support point of a shape in world direction d (unit) */
static void support(const Shape* s, const double d[3], double out[3]) {
  double dl[3], p[3];
  mulTv(dl, s->R, d);
  if (s->type == GMJ_GEOM_BOX) {
    for (int k = 0; k < 3; k++) p[k] = dl[k] >= 0 ? s->size[k] : -s->size[k];
  } else if (s->type == GMJ_GEOM_CYLINDER) {
    double l = sqrt(dl[0] * dl[0] + dl[1] * dl[1]);
    if (l > 1e-12) { p[0] = s->size[0] * dl[0] / l; p[1] = s->size[0] * dl[1] / l; }
    else { p[0] = 0; p[1] = 0; }
    p[2] = dl[2] >= 0 ? s->size[2] : -s->size[2];
  } else {
    double a = s->size[0], b = s->size[1], c = s->size[2];
    double l = sqrt(a * a * dl[0] * dl[0] + b * b * dl[1] * dl[1] + c * c * dl[2] * dl[2]);
    if (l < 1e-300) l = 1e-300;
    p[0] = a * a * dl[0] / l; p[1] = b * b * dl[1] / l; p[2] = c * c * dl[2] / l;
  }
  if (s->round > 0) for (int k = 0; k < 3; k++) p[k] += s->round * dl[k];
  mulv(out, s->R, p);
  add3(out, out, s->pos);
}

typedef struct { double v[3], a[3], b[3]; } MV;  /* Minkowski vertex v = a - b */

static void msupport(const Shape* A, const Shape* B, const double dir[3], MV* o) {
  double nd[3] = { -dir[0], -dir[1], -dir[2] };
  support(A, dir, o->a);
  support(B, nd, o->b);
  sub3(o->v, o->a, o->b);
}

/* closest point of triangle (a,b,c) to the origin, barycentric weights out (Ericson, RTCD 5.1.5) */
static void tri_closest_origin(const double a[3], const double b[3], const double c[3], double w[3]) {
  double ab[3], ac[3], ap[3], bp[3], cp[3];
  sub3(ab, b, a); sub3(ac, c, a);
  scl3(ap, a, -1); scl3(bp, b, -1); scl3(cp, c, -1);
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
  if (d1 <= 0 && d2 <= 0) { w[0] = 1; w[1] = 0; w[2] = 0; return; }
  double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
  if (d3 >= 0 && d4 <= d3) { w[0] = 0; w[1] = 1; w[2] = 0; return; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) { double v = d1 / (d1 - d3); w[0] = 1 - v; w[1] = v; w[2] = 0; return; }
  double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  if (d6 >= 0 && d5 <= d6) { w[0] = 0; w[1] = 0; w[2] = 1; return; }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) { double v = d2 / (d2 - d6); w[0] = 1 - v; w[1] = 0; w[2] = v; return; }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    double v = (d4 - d3) / ((d4 - d3) + (d5 - d6)); w[0] = 0; w[1] = 1 - v; w[2] = v; return; }
  double den = 1.0 / (va + vb + vc);
  w[1] = vb * den; w[2] = vc * den; w[0] = 1 - w[1] - w[2];
}

#define MPR_TOL 1e-6
#define MPR_MAXIT 50

/* returns 1 on penetration: normal points from A to B, depth > 0, pos = contact point */
static int mpr_penetration(const Shape* A, const Shape* B, double* depth, double normal[3], double pos[3]) {
  MV v0, v1, v2, v3, v4;
  double n[3], t1[3], t2[3];
  sub3(v0.v, A->pos, B->pos);
  copy3(v0.a, A->pos); copy3(v0.b, B->pos);
  if (dot3(v0.v, v0.v) < 1e-20) { v0.v[0] = 1e-5; v0.v[1] = 0; v0.v[2] = 0; }
  scl3(n, v0.v, -1); normalize3(n);
  msupport(A, B, n, &v1);
  if (dot3(v1.v, n) <= 0) return 0;
  cross3(n, v1.v, v0.v);
  if (dot3(n, n) < 1e-24) {
    /* origin lies on the segment v0 -> v1 */
    sub3(n, v1.v, v0.v); normalize3(n);
    *depth = dot3(v1.v, n);
    copy3(normal, n);
    for (int k = 0; k < 3; k++) pos[k] = 0.5 * (v1.a[k] + v1.b[k]);
    return 1;
  }
  normalize3(n);
  msupport(A, B, n, &v2);
  if (dot3(v2.v, n) <= 0) return 0;
  sub3(t1, v1.v, v0.v); sub3(t2, v2.v, v0.v); cross3(n, t1, t2);
  if (dot3(n, v0.v) > 0) { MV t = v1; v1 = v2; v2 = t; scl3(n, n, -1); }
  /* phase 1: find a portal the origin ray passes through */
  for (int it = 0;; it++) {
    if (it > MPR_MAXIT) return 0;
    if (dot3(n, n) < 1e-30) return 0;
    normalize3(n);
    msupport(A, B, n, &v3);
    if (dot3(v3.v, n) <= 0) return 0;
    cross3(t1, v1.v, v3.v);
    if (dot3(t1, v0.v) < 0) { v2 = v3; sub3(t1, v1.v, v0.v); sub3(t2, v3.v, v0.v); cross3(n, t1, t2); continue; }
    cross3(t1, v3.v, v2.v);
    if (dot3(t1, v0.v) < 0) { v1 = v3; sub3(t1, v3.v, v0.v); sub3(t2, v2.v, v0.v); cross3(n, t1, t2); continue; }
    break;
  }
  /* phase 2: refine the portal until it reaches the surface */
  int inside = 0;
  for (int it = 0; it < MPR_MAXIT; it++) {
    sub3(t1, v2.v, v1.v); sub3(t2, v3.v, v1.v); cross3(n, t1, t2);
    if (dot3(n, n) < 1e-30) break;
    normalize3(n);
    if (dot3(n, v1.v) >= 0) inside = 1;
    msupport(A, B, n, &v4);
    double d4 = dot3(v4.v, n);
    if (!inside && d4 < 0) return 0;
    double r1 = d4 - dot3(v1.v, n), r2 = d4 - dot3(v2.v, n), r3 = d4 - dot3(v3.v, n);
    double reach = fmin(r1, fmin(r2, r3));
    if (reach <= MPR_TOL) break;
    cross3(t1, v4.v, v0.v);
    if (dot3(v1.v, t1) > 0) { if (dot3(v2.v, t1) > 0) v1 = v4; else v3 = v4; }
    else { if (dot3(v3.v, t1) > 0) v2 = v4; else v1 = v4; }
  }
  if (!inside) return 0;
  double w[3], p[3];
  tri_closest_origin(v1.v, v2.v, v3.v, w);
  for (int k = 0; k < 3; k++) {
    p[k] = w[0] * v1.v[k] + w[1] * v2.v[k] + w[2] * v3.v[k];
    pos[k] = 0.5 * (w[0] * (v1.a[k] + v1.b[k]) + w[1] * (v2.a[k] + v2.b[k]) + w[2] * (v3.a[k] + v3.b[k]));
  }
  double d = sqrt(dot3(p, p));
  if (d > 1e-12) { scl3(normal, p, 1.0 / d); }
  else { sub3(t1, v2.v, v1.v); sub3(t2, v3.v, v1.v); cross3(normal, t1, t2); normalize3(normal);
         if (dot3(normal, v0.v) > 0) scl3(normal, normal, -1); }
  *depth = d;
  return 1;
}

/* mju_makeFrame: tangents for a given normal */
static void make_frame(double f[9]) {
  double* n = f; double* y = f + 3; double* z = f + 6;
  if (fabs(n[1]) < 0.5) { y[0] = 0; y[1] = 1; y[2] = 0; } else { y[0] = 0; y[1] = 0; y[2] = 1; }
  double d = dot3(n, y);
  for (int k = 0; k < 3; k++) y[k] -= d * n[k];
  normalize3(y);
  cross3(z, n, y);
}

static void add_contact(OrEnv* e, const double pos[3], const double n[3], double dist, double mu,
                        int b1, int c1, int b2, int c2) {
  if (e->ncon >= GMJ_MAXCON) return;
  OrContact* c = &e->con[e->ncon++];
  copy3(c->pos, pos); copy3(c->frame, n);
  make_frame(c->frame);
  c->dist = dist; c->mu = mu; c->body1 = b1; c->body2 = b2; c->cls1 = c1; c->cls2 = c2;
  c->force[0] = c->force[1] = c->force[2] = 0;
}

/* keep at most 4 deepest of the candidate plane points */
typedef struct { double p[3]; double d; } PlanePt;
static void emit_plane_pts(OrEnv* e, PlanePt* pts, int n, double mu, int body, int cls) {
  static const double up[3] = { 0, 0, 1 };
  for (int k = 0; k < 4 && k < n; k++) {
    int best = -1;
    for (int i = 0; i < n; i++) if (pts[i].d < 0 && (best < 0 || pts[i].d < pts[best].d)) best = i;
    if (best < 0) break;
    /* contact position at the mid-surface like MuJoCo plane contacts (pos - 0.5*dist*normal) */
    double pos[3] = { pts[best].p[0], pts[best].p[1], pts[best].p[2] - 0.5 * pts[best].d };
    add_contact(e, pos, up, pts[best].d, mu, -1, GMJ_CLS_GROUND, body, cls);
    pts[best].d = 1;
  }
}

/* ground plane z = 0 against a shape (geom1 = ground, normal +z) */
static void plane_shape(OrEnv* e, const Shape* s, double mu, int body, int cls) {
  PlanePt pts[8]; int n = 0;
  double rz[3] = { s->R[6], s->R[7], s->R[8] };       /* world z expressed in shape frame */
  if (s->pos[2] - shape_bound(s) > 0) return;
  if (s->type == GMJ_GEOM_ELLIPSOID) {
    double dn[3] = { 0, 0, -1 }, p[3];
    support(s, dn, p);
    pts[0].p[0] = p[0]; pts[0].p[1] = p[1]; pts[0].p[2] = p[2]; pts[0].d = p[2]; n = 1;
  } else if (s->type == GMJ_GEOM_BOX) {
    for (int i = 0; i < 8; i++) {
      double c[3] = { (i & 1 ? 1 : -1) * s->size[0], (i & 2 ? 1 : -1) * s->size[1], (i & 4 ? 1 : -1) * s->size[2] };
      double p[3]; mulv(p, s->R, c); add3(p, p, s->pos);
      pts[n].p[0] = p[0]; pts[n].p[1] = p[1]; pts[n].p[2] = p[2] - s->round; pts[n].d = p[2] - s->round; n++;
    }
  } else {
    /* cylinder, scheme of MuJoCo's plane-cylinder routine: deepest rim point on the near cap,
     * the same generator on the far cap, and two more near-cap rim points at +-120 deg */
    double axis[3] = { s->R[2], s->R[5], s->R[8] };    /* shape z in world */
    double prj = -axis[2];                             /* dot(axis, -normal) */
    if (prj < 0) { scl3(axis, axis, -1); prj = -prj; } /* axis now points towards the plane */
    double vec[3] = { -axis[0] * prj, -axis[1] * prj, -1.0 - axis[2] * prj }; /* -n minus its axis part */
    double len = sqrt(dot3(vec, vec));
    if (len < 1e-12) { vec[0] = s->R[0]; vec[1] = s->R[3]; vec[2] = s->R[6]; len = 1; }
    scl3(vec, vec, s->size[0] / len);
    double ax[3]; scl3(ax, axis, s->size[2]);
    double p[3];
    for (int k = 0; k < 3; k++) p[k] = s->pos[k] + ax[k] + vec[k];
    pts[n].p[0] = p[0]; pts[n].p[1] = p[1]; pts[n].p[2] = p[2] - s->round; pts[n].d = p[2] - s->round; n++;
    for (int k = 0; k < 3; k++) p[k] = s->pos[k] - ax[k] + vec[k];
    pts[n].p[0] = p[0]; pts[n].p[1] = p[1]; pts[n].p[2] = p[2] - s->round; pts[n].d = p[2] - s->round; n++;
    double v1[3]; cross3(v1, vec, axis); normalize3(v1);
    scl3(v1, v1, s->size[0] * sqrt(3.0) * 0.5);
    for (int sgn = -1; sgn <= 1; sgn += 2) {
      for (int k = 0; k < 3; k++) p[k] = s->pos[k] + ax[k] - 0.5 * vec[k] + sgn * v1[k];
      pts[n].p[0] = p[0]; pts[n].p[1] = p[1]; pts[n].p[2] = p[2] - s->round; pts[n].d = p[2] - s->round; n++;
    }
  }
  (void)rz;
  emit_plane_pts(e, pts, n, mu, body, cls);
}

static void object_shape(const OrEnv* e, Shape* s) {
  double Rs[9];
  quat2mat(Rs, e->obj.quat);
  s->type = e->obj.type; s->round = e->obj.round;
  copy3(s->size, e->obj.size);
  copy3(s->pos, e->opos);
  mulm(s->R, e->omat, Rs);
}

static void geom_shape(const OrEnv* e, int g, Shape* s) {
  const GmjModel* m = &e->m;
  int b = m->g_body[g];
  double Rg[9], p[3];
  quat2mat(Rg, m->g_quat[g]);
  s->type = m->g_type[g]; s->round = 0;
  copy3(s->size, m->g_size[g]);
  mulv(p, e->xmat[b], m->g_pos[g]); add3(s->pos, p, e->xpos[b]);
  mulm(s->R, e->xmat[b], Rg);
}

/* contact ordering follows MuJoCo's body order ground < gripper < object:
 * geom1 is the lower one, the frame normal points geom1 -> geom2. */
void or_collide(OrEnv* e) {
  const GmjModel* m = &e->m;
  e->ncon = 0;
  Shape so, sg;
  if (e->has_object) {
    object_shape(e, &so);
    plane_shape(e, &so, fmax(m->ground_friction, e->obj.friction), e->B, GMJ_CLS_OBJECT);
  }
  for (int g = 0; g < m->ngeom; g++) {
    geom_shape(e, g, &sg);
    if (m->g_type[g] == GMJ_GEOM_BOX)
      plane_shape(e, &sg, fmax(m->ground_friction, m->g_friction[g]), m->g_body[g], m->g_class[g]);
  }
  if (e->self_collision) {
    /* gripper <-> gripper (off by default, see oracle.h): every geom pair on different bodies that are not parent and
     * child, geom1 < geom2 like MuJoCo's pair list, which puts them before the gripper <-> object pairs */
    Shape sh;
    for (int g = 0; g < m->ngeom; g++) {
      geom_shape(e, g, &sg);
      double rg = shape_bound(&sg);
      for (int h = g + 1; h < m->ngeom; h++) {
        int bg = m->g_body[g], bh = m->g_body[h];
        if (bg == bh || m->parent[bg] == bh || m->parent[bh] == bg) continue;
        geom_shape(e, h, &sh);
        double d[3]; sub3(d, sh.pos, sg.pos);
        double rr = rg + shape_bound(&sh);
        if (dot3(d, d) > rr * rr) continue;
        double depth, n[3], pos[3];
        if (mpr_penetration(&sg, &sh, &depth, n, pos))
          add_contact(e, pos, n, -depth, fmax(m->g_friction[g], m->g_friction[h]), bg, m->g_class[g], bh, m->g_class[h]);
      }
    }
  }
  if (!e->has_object) return;
  double ro = shape_bound(&so);
  for (int g = 0; g < m->ngeom; g++) {
    geom_shape(e, g, &sg);
    double d[3]; sub3(d, sg.pos, so.pos);
    double rr = ro + shape_bound(&sg);
    if (dot3(d, d) > rr * rr) continue;
    double depth, n[3], pos[3];
    if (mpr_penetration(&sg, &so, &depth, n, pos))
      add_contact(e, pos, n, -depth, fmax(m->g_friction[g], e->obj.friction),
                  m->g_body[g], m->g_class[g], e->B, GMJ_CLS_OBJECT);
  }
}
