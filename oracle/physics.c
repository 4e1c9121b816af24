/* physics.c -- oracle forward dynamics for the gripper task scene.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). PARITY UNPINNED (MuJoCo absent).
 *
 * Restates what the reference gets from MuJoCo 2.1.5 per substep
 * (luke::step = mj_step1 ; control ; mj_step2, src/myfunctions.cpp:1873-1898) following the
 * pipeline in MuJoCo's "Computation" chapter with the model's (default) options
 * (description/mujoco/xacro/mujoco.xacro:6-12: Euler, pyramidal cone, condim 3,
 * solref (0.02,1), solimp (0.9,0.95,0.001,0.5,2)):
 *   kinematics -> composite inertia M -> bias (RNE) -> passive -> actuation (PD control,
 *   src/myfunctions.cpp:1965-2057) -> qacc_smooth -> collision -> constraint rows (welds,
 *   joint limits, pyramidal contacts) -> soft-constraint QP (dual, projected Gauss-Seidel to
 *   convergence; MuJoCo's Newton solver converges to the same unique optimum) -> Euler with
 *   implicit joint damping.
 * Spatial quantities are expressed in world coordinates about the world origin.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "oracle.h"
#include "vec.h"

void or_collide(OrEnv* e);

/* --------------------------------------------------------------------- kinematics */
void or_kinematics(OrEnv* e) {
  const GmjModel* m = &e->m;
  for (int i = 0; i < e->B; i++) {
    double Rb[9], R0[9], P0[3], t[3];
    quat2mat(Rb, m->quat[i]);
    int p = m->parent[i];
    if (p < 0) { memcpy(R0, Rb, sizeof(Rb)); copy3(P0, m->pos[i]); }
    else { mulm(R0, e->xmat[p], Rb); mulv(t, e->xmat[p], m->pos[i]); add3(P0, e->xpos[p], t); }
    mulv(e->xaxis[i], R0, m->axis[i]);
    copy3(e->xanchor[i], P0);
    if (m->jtype[i] == 0) {
      memcpy(e->xmat[i], R0, sizeof(R0));
      copy3(e->xpos[i], P0); axpy3(e->xpos[i], e->qpos[i], e->xaxis[i]);
    } else {
      double Rj[9];
      axisangle2mat(Rj, m->axis[i], e->qpos[i]);
      mulm(e->xmat[i], R0, Rj);
      copy3(e->xpos[i], P0);
    }
  }
  const double* q = e->qpos + e->B;
  copy3(e->opos, q);
  double qq[4] = { q[3], q[4], q[5], q[6] };
  double l = sqrt(qq[0] * qq[0] + qq[1] * qq[1] + qq[2] * qq[2] + qq[3] * qq[3]);
  for (int k = 0; k < 4; k++) qq[k] /= l;
  quat2mat(e->omat, qq);
}

/* spatial helpers: s = (w[3]; v[3]) motion, f = (n[3]; f[3]) force */
static void crossm(double* o, const double* v, const double* s) {     /* v x_m s */
  double a[3], b[3], c[3];
  cross3(a, v, s); cross3(b, v, s + 3); cross3(c, v + 3, s);
  o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; o[3] = b[0] + c[0]; o[4] = b[1] + c[1]; o[5] = b[2] + c[2];
}
static void crossf(double* o, const double* v, const double* f) {     /* v x* f */
  double a[3], b[3], c[3];
  cross3(a, v, f); cross3(b, v + 3, f + 3); cross3(c, v, f + 3);
  o[0] = a[0] + b[0]; o[1] = a[1] + b[1]; o[2] = a[2] + b[2]; o[3] = c[0]; o[4] = c[1]; o[5] = c[2];
}
typedef struct { double m, h[3], I[9]; } SpI;                         /* about the world origin */
static void spi_mul(double* o, const SpI* I, const double* s) {
  double a[3], b[3];
  mulv(a, I->I, s); cross3(b, I->h, s + 3);
  o[0] = a[0] + b[0]; o[1] = a[1] + b[1]; o[2] = a[2] + b[2];
  cross3(b, s, I->h);
  o[3] = I->m * s[3] + b[0]; o[4] = I->m * s[4] + b[1]; o[5] = I->m * s[5] + b[2];
}

/* mass matrix (gripper block dense, nv x nv with the object block diagonal) and bias forces */
static void dynamics(OrEnv* e, double* M, double* bias) {
  const GmjModel* m = &e->m;
  int B = e->B, nv = e->nv;
  static SpI I[GMJ_MAXB], Ic[GMJ_MAXB];
  static double S[GMJ_MAXB][6], v[GMJ_MAXB][6], a[GMJ_MAXB][6], f[GMJ_MAXB][6];
  memset(M, 0, sizeof(double) * nv * nv);
  for (int i = 0; i < B; i++) {
    double c[3], Iw[9], Ib[9], t[9];
    mulv(c, e->xmat[i], m->ipos[i]); add3(c, c, e->xpos[i]);
    const double* in = m->inertia[i];
    Ib[0] = in[0]; Ib[4] = in[1]; Ib[8] = in[2]; Ib[1] = Ib[3] = in[3]; Ib[2] = Ib[6] = in[4]; Ib[5] = Ib[7] = in[5];
    mulm(t, e->xmat[i], Ib);
    for (int r = 0; r < 3; r++) for (int cc = 0; cc < 3; cc++)
      Iw[3 * r + cc] = t[3 * r] * e->xmat[i][3 * cc] + t[3 * r + 1] * e->xmat[i][3 * cc + 1] + t[3 * r + 2] * e->xmat[i][3 * cc + 2];
    double mm = m->mass[i], cc2 = dot3(c, c);
    I[i].m = mm; scl3(I[i].h, c, mm);
    for (int r = 0; r < 3; r++) for (int cc = 0; cc < 3; cc++)
      I[i].I[3 * r + cc] = Iw[3 * r + cc] + mm * ((r == cc ? cc2 : 0.0) - c[r] * c[cc]);
    if (m->jtype[i] == 0) { zero3(S[i]); copy3(S[i] + 3, e->xaxis[i]); }
    else { copy3(S[i], e->xaxis[i]); cross3(S[i] + 3, e->xanchor[i], e->xaxis[i]); }
  }
  /* forward pass: velocities and bias accelerations (gravity as base acceleration +g) */
  for (int i = 0; i < B; i++) {
    int p = m->parent[i];
    double vj[6], sd[6];
    for (int k = 0; k < 6; k++) vj[k] = S[i][k] * e->qvel[i];
    if (p < 0) { for (int k = 0; k < 6; k++) { v[i][k] = vj[k]; a[i][k] = 0; } a[i][5] = m->gravity; }
    else { for (int k = 0; k < 6; k++) { v[i][k] = v[p][k] + vj[k]; a[i][k] = a[p][k]; } }
    crossm(sd, v[i], S[i]);
    for (int k = 0; k < 6; k++) a[i][k] += sd[k] * e->qvel[i];
    double Ia[6], Iv[6], vIv[6];
    spi_mul(Ia, &I[i], a[i]); spi_mul(Iv, &I[i], v[i]); crossf(vIv, v[i], Iv);
    for (int k = 0; k < 6; k++) f[i][k] = Ia[k] + vIv[k];
    Ic[i] = I[i];
  }
  for (int i = B - 1; i >= 0; i--) {
    bias[i] = 0;
    for (int k = 0; k < 6; k++) bias[i] += S[i][k] * f[i][k];
    int p = m->parent[i];
    if (p >= 0) {
      for (int k = 0; k < 6; k++) f[p][k] += f[i][k];
      Ic[p].m += Ic[i].m; add3(Ic[p].h, Ic[p].h, Ic[i].h);
      for (int k = 0; k < 9; k++) Ic[p].I[k] += Ic[i].I[k];
    }
  }
  for (int i = 0; i < B; i++) {
    double F[6];
    spi_mul(F, &Ic[i], S[i]);
    for (int j = i; j >= 0; j = m->parent[j]) {
      double d = 0;
      for (int k = 0; k < 6; k++) d += S[j][k] * F[k];
      M[i * nv + j] = d; M[j * nv + i] = d;
    }
  }
  /* object: free joint, COM at the body origin, translational dofs in the world frame and
   * rotational dofs in the body frame => constant diagonal mass matrix */
  if (e->has_object) {
    const double* w = e->qvel + B + 3;
    double Iw[3] = { e->obj.inertia[0] * w[0], e->obj.inertia[1] * w[1], e->obj.inertia[2] * w[2] }, g[3];
    cross3(g, w, Iw);
    for (int k = 0; k < 3; k++) {
      M[(B + k) * nv + B + k] = e->obj.mass;
      M[(B + 3 + k) * nv + B + 3 + k] = e->obj.inertia[k];
      bias[B + k] = 0; bias[B + 3 + k] = g[k];
    }
    bias[B + 2] = e->obj.mass * m->gravity;
  } else {
    for (int k = 0; k < 6; k++) { M[(B + k) * nv + B + k] = 1.0; bias[B + k] = 0; }
  }
}

/* dense Cholesky M = L L^T (lower, in place) and solve */
static int chol(double* A, int n) {
  for (int j = 0; j < n; j++) {
    double d = A[j * n + j];
    for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
    if (!(d > 0)) return 0;
    d = sqrt(d); A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      double s = A[i * n + j];
      for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = s / d;
    }
  }
  return 1;
}
static void chol_solve(const double* L, int n, double* x) {
  for (int i = 0; i < n; i++) { double s = x[i]; for (int k = 0; k < i; k++) s -= L[i * n + k] * x[k]; x[i] = s / L[i * n + i]; }
  for (int i = n - 1; i >= 0; i--) { double s = x[i]; for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k]; x[i] = s / L[i * n + i]; }
}

/* ---- timed-baseline path (OrEnv::fast_solver): the same factorisations in a FILL-FREE ordering. With the dofs ordered
 * fingers, palm, base, object the mass matrix is block-arrow (finger blocks couple only through the base row), so row
 * i of its Cholesky factor starts at first[i] -- a skyline factorisation does the tree-sparse L D L^T work MuJoCo does,
 * instead of dense nv^3 / 3. Same arithmetic on the non-zero entries; the checker path keeps the dense routines. */
static int chol_sky(double* A, int n, const int* first) {
  for (int j = 0; j < n; j++) {
    double d = A[j * n + j];
    for (int k = first[j]; k < j; k++) d -= A[j * n + k] * A[j * n + k];
    if (!(d > 0)) return 0;
    d = sqrt(d); A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      if (first[i] > j) { A[i * n + j] = 0; continue; }
      double s = A[i * n + j];
      for (int k = first[i] > first[j] ? first[i] : first[j]; k < j; k++) s -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = s / d;
    }
  }
  return 1;
}
static void chol_solve_sky(const double* L, int n, const int* first, const int* perm, double* v) {
  double x[OR_MAXV];
  for (int i = 0; i < n; i++) x[i] = v[perm[i]];
  int i0 = 0; while (i0 < n && x[i0] == 0.0) i0++;               /* leading zeros of the right-hand side stay zero */
  for (int i = i0; i < n; i++) { double s = x[i]; for (int k = first[i] > i0 ? first[i] : i0; k < i; k++) s -= L[i * n + k] * x[k]; x[i] = s / L[i * n + i]; }
  for (int i = n - 1; i >= 0; i--) { x[i] /= L[i * n + i]; const double xi = x[i]; for (int k = first[i]; k < i; k++) x[k] -= L[i * n + k] * xi; }
  for (int i = 0; i < n; i++) v[perm[i]] = x[i];
}
static void sky_setup(const OrEnv* e, const double* M, double* Mp, int* perm, int* first) {
  int nv = e->nv, B = e->B, n = 0;
  for (int i = 1; i < B; i++) perm[n++] = i;                    /* fingers (prismatic, revolute, segments) x 3, palm */
  perm[n++] = 0;                                                 /* base last among the gripper dofs: no fill-in */
  for (int i = B; i < nv; i++) perm[n++] = i;                   /* object */
  for (int i = 0; i < nv; i++) {
    first[i] = i;
    for (int j = 0; j < nv; j++) Mp[i * nv + j] = M[perm[i] * nv + perm[j]];
    for (int j = 0; j < i; j++) if (Mp[i * nv + j] != 0.0) { first[i] = j; break; }
  }
}

/* translational + rotational Jacobian of body b at world point p: rows x,y,z (3 x nv each) */
static void jac_point(const OrEnv* e, int b, const double p[3], double* Jp, double* Jr) {
  const GmjModel* m = &e->m;
  int nv = e->nv, B = e->B;
  memset(Jp, 0, sizeof(double) * 3 * nv); memset(Jr, 0, sizeof(double) * 3 * nv);
  if (b < 0) return;
  if (b == B) {
    for (int k = 0; k < 3; k++) {
      Jp[k * nv + B + k] = 1.0;
      double ax[3] = { e->omat[k], e->omat[3 + k], e->omat[6 + k] }, r[3], c[3];
      sub3(r, p, e->opos); cross3(c, ax, r);
      for (int d = 0; d < 3; d++) { Jp[d * nv + B + 3 + k] = c[d]; Jr[d * nv + B + 3 + k] = ax[d]; }
    }
    return;
  }
  for (int j = b; j >= 0; j = m->parent[j]) {
    if (m->jtype[j] == 0) { for (int d = 0; d < 3; d++) Jp[d * nv + j] = e->xaxis[j][d]; }
    else {
      double r[3], c[3];
      sub3(r, p, e->xanchor[j]); cross3(c, e->xaxis[j], r);
      for (int d = 0; d < 3; d++) { Jp[d * nv + j] = c[d]; Jr[d * nv + j] = e->xaxis[j][d]; }
    }
  }
}

/* solimp impedance d(r) (MuJoCo getimpedance; power-law sigmoid) */
static double impedance(const double* solimp, double pos) {
  double d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  double x = fabs(pos) / width;
  if (x >= 1) return dw;
  if (x <= 0) return d0;
  double y;
  /* a*x^p with a = 1/mid^(p-1); p = 2 in every reference model: squared explicitly, as the device does */
  double a = (x <= mid) ? x / mid : (1 - x) / (1 - mid);
  double ap = (power == 2.0) ? a * a : pow(a, power);
  if (x <= mid) y = ap * mid;
  else y = 1 - ap * (1 - mid);
  return d0 + y * (dw - d0);
}

enum { ROW_EQ = 0, ROW_LIMIT = 1, ROW_CONTACT = 2 };


/* exact solution of (A + R) f + b = w, eq rows w = 0, ineq rows f >= 0, w >= 0, f w = 0, starting from the
 * active set of the given f: block principal pivoting with Murty fallback (finite for SPD matrices) */
static void lcp_polish(int n, const double* A, int lda, const double* R, const double* b, const int* type, double* f) {
  static double K[OR_MAXROW * OR_MAXROW], x[OR_MAXROW], w[OR_MAXROW];
  int act[OR_MAXROW], idx[OR_MAXROW];
  double bscale = 0;
  for (int r = 0; r < n; r++) { act[r] = (type[r] == ROW_EQ) || f[r] > 0; bscale = fmax(bscale, fabs(b[r])); }
  int best = 1 << 30, patience = 3;
  for (int round = 0; round < 200; round++) {
    int m = 0;
    for (int r = 0; r < n; r++) if (act[r]) idx[m++] = r;
    for (int a = 0; a < m; a++) {
      for (int c = 0; c < m; c++) K[a * m + c] = A[idx[a] * lda + idx[c]];
      K[a * m + a] += R[idx[a]];
      x[a] = -b[idx[a]];
    }
    if (m > 0) { if (!chol(K, m)) return; chol_solve(K, m, x); }
    double fscale = 0;
    for (int a = 0; a < m; a++) fscale = fmax(fscale, fabs(x[a]));
    for (int r = 0; r < n; r++) {
      double s = b[r];
      for (int a = 0; a < m; a++) s += A[r * lda + idx[a]] * x[a];
      w[r] = s;
    }
    const double tolf = 1e-12 * (1 + fscale), tolw = 1e-12 * (1 + bscale);
    int ninf = 0, top = -1, bad[OR_MAXROW];
    for (int r = 0, a = 0; r < n; r++) {
      bad[r] = 0;
      if (act[r]) { if (type[r] != ROW_EQ && x[a] < -tolf) bad[r] = 1; a++; }
      else if (w[r] < -tolw) bad[r] = 1;
      if (bad[r]) { ninf++; top = r; }
    }
    if (ninf == 0) {
      for (int r = 0, a = 0; r < n; r++) f[r] = act[r] ? x[a++] : 0.0;
      for (int r = 0; r < n; r++) if (type[r] != ROW_EQ && f[r] < 0) f[r] = 0;
      return;
    }
    int block = 1;
    if (ninf < best) { best = ninf; patience = 3; } else if (patience > 0) patience--; else block = 0;
    if (block) { for (int r = 0; r < n; r++) if (bad[r]) act[r] = !act[r]; }
    else act[top] = !act[top];
  }
}

void or_physics_step(OrEnv* e) {
  const GmjModel* m = &e->m;
  const double dt = e->s.general[GS_mujoco_timestep];
  int B = e->B, nv = e->nv;
  static double M[OR_MAXV * OR_MAXV], L[OR_MAXV * OR_MAXV], bias[OR_MAXV], qfrc[OR_MAXV], qacc_s[OR_MAXV];
  static double J[OR_MAXROW][OR_MAXV], Bm[OR_MAXROW][OR_MAXV], A[OR_MAXROW][OR_MAXROW];
  static double rpos[OR_MAXROW], Rr[OR_MAXROW], aref[OR_MAXROW], bvec[OR_MAXROW], fr[OR_MAXROW];
  static int rtype[OR_MAXROW];
  static double Jp1[3 * OR_MAXV], Jr1[3 * OR_MAXV], Jp2[3 * OR_MAXV], Jr2[3 * OR_MAXV];

  /* after a BADQACC-class failure the env is frozen until reset (is_done reports it) */
  if (e->unstable) { e->time += dt; return; }
  if (e->script) {                 /* scripted observables: the clock advances as in mj_step, nothing is simulated */
    e->time += dt;
    if (e->script_cursor < e->script_rows) {
      memcpy(e->script_cur, e->script + (size_t)e->script_cursor * OR_SCRIPT_ROWLEN, sizeof(e->script_cur));
      e->script_cursor++;
    }
    e->wrist_z = e->script_cur[39];
    if (e->has_object) for (int k = 0; k < 3; k++) e->qpos[B + k] = e->script_cur[40 + k];
    return;
  }
  /* ---- mj_step1: position + velocity stage */
  or_kinematics(e);
  dynamics(e, M, bias);
  or_collide(e);

  /* ---- luke::control (src/myfunctions.cpp:1912-2057): ctrl = -(kp (q - q*) + kd qdot) */
  double ctrl[OR_MAXV];
  memset(ctrl, 0, sizeof(ctrl));
  ctrl[0] = -((e->qpos[0] - e->base_z) * m->base_kp + e->qvel[0] * m->base_kd);
  for (int f = 0; f < 3; f++) {
    int ip = 1 + f * (m->N + 2), ir = ip + 1;
    ctrl[ip] = -((e->qpos[ip] - e->tgt_next.x) * m->kp[0] + e->qvel[ip] * m->kd[0]);
    ctrl[ir] = -((e->qpos[ir] - e->tgt_next.th) * m->kp[1] + e->qvel[ir] * m->kd[1]);
  }
  ctrl[B - 1] = -((e->qpos[B - 1] - e->tgt_next.z) * m->kp[2] + e->qvel[B - 1] * m->kd[2]);

  /* ---- mj_step2: passive + actuation, qacc_smooth */
  for (int i = 0; i < nv; i++) qfrc[i] = -bias[i];
  for (int i = 0; i < B; i++) qfrc[i] += ctrl[i] - m->stiffness[i] * e->qpos[i] - m->damping[i] * e->qvel[i];
  /* calibration tip load: xfrc_applied on the last segment of each finger, applied at the body
   * COM, force = tip_mat * (0,0,-F) (src/myfunctions.cpp:1642-1727) */
  if (e->tip_force != 0.0) {
    for (int f = 0; f < 3; f++) {
      int b = m->tip_body[f];
      double c[3], fv[3], loc[3] = { 0, 0, -e->tip_force };
      mulv(c, e->xmat[b], m->ipos[b]); add3(c, c, e->xpos[b]);
      mulv(fv, m->tip_mat[f], loc);
      jac_point(e, b, c, Jp1, Jr1);
      for (int j = 0; j < nv; j++) qfrc[j] += Jp1[j] * fv[0] + Jp1[nv + j] * fv[1] + Jp1[2 * nv + j] * fv[2];
    }
  }
  static int perm[OR_MAXV], first[OR_MAXV];
  const int sky = e->fast_solver;
  if (sky) { sky_setup(e, M, L, perm, first); if (!chol_sky(L, nv, first)) { e->unstable = 1; return; } }
  else {
    memcpy(L, M, sizeof(double) * nv * nv);
    if (!chol(L, nv)) { e->unstable = 1; return; }
  }
  memcpy(qacc_s, qfrc, sizeof(double) * nv);
  if (sky) chol_solve_sky(L, nv, first, perm, qacc_s); else chol_solve(L, nv, qacc_s);

  /* ---- constraint rows */
  int nr = 0;
  /* welds (equality): 3 position + 3 orientation rows; all-zero rows are dropped (they carry
   * no force in MuJoCo either) */
  for (int w = 0; w < 7; w++) {
    const OrWeld* wd = &e->weld[w];
    if (!wd->on) continue;
    int b1 = wd->body1, b2 = wd->body2;
    double a1[3], r[3], Rt[9], Re[9], rr[3];
    mulv(a1, e->xmat[b1], wd->relpos); add3(a1, a1, e->xpos[b1]);
    sub3(r, e->xpos[b2], a1);
    mulm(Rt, e->xmat[b1], wd->relmat);                 /* target orientation of body2 */
    for (int x = 0; x < 3; x++) for (int y = 0; y < 3; y++)
      Re[3 * x + y] = e->xmat[b2][3 * x] * Rt[3 * y] + e->xmat[b2][3 * x + 1] * Rt[3 * y + 1] + e->xmat[b2][3 * x + 2] * Rt[3 * y + 2];
    rr[0] = 0.5 * (Re[7] - Re[5]); rr[1] = 0.5 * (Re[2] - Re[6]); rr[2] = 0.5 * (Re[3] - Re[1]);
    jac_point(e, b1, a1, Jp1, Jr1);
    jac_point(e, b2, e->xpos[b2], Jp2, Jr2);
    for (int d = 0; d < 6; d++) {
      double nz = 0;
      for (int j = 0; j < nv; j++) {
        double v = (d < 3) ? (Jp2[d * nv + j] - Jp1[d * nv + j]) : (Jr2[(d - 3) * nv + j] - Jr1[(d - 3) * nv + j]);
        J[nr][j] = v; nz += v * v;
      }
      if (nz < 1e-10) continue;   /* |J| < 1e-5: structurally zero row */
      rpos[nr] = (d < 3) ? r[d] : rr[d - 3];
      rtype[nr] = ROW_EQ;
      nr++;
    }
  }
  /* joint limits (margin 0) */
  for (int i = 0; i < B; i++) {
    double lo = e->qpos[i] - m->range_lo[i], hi = m->range_hi[i] - e->qpos[i];
    if (lo < 0 && nr < GMJ_MAXROW) { memset(J[nr], 0, sizeof(double) * nv); J[nr][i] = 1; rpos[nr] = lo; rtype[nr++] = ROW_LIMIT; }
    if (hi < 0 && nr < GMJ_MAXROW) { memset(J[nr], 0, sizeof(double) * nv); J[nr][i] = -1; rpos[nr] = hi; rtype[nr++] = ROW_LIMIT; }
  }
  /* contacts: pyramidal cone, condim 3 -> 4 rows (n + mu t1, n - mu t1, n + mu t2, n - mu t2) */
  for (int c = 0; c < e->ncon; c++) {
    OrContact* cn = &e->con[c];
    if (nr + 4 > GMJ_MAXROW) { e->ncon = c; break; }   /* row cap shared with the device path */
    cn->row = nr;
    jac_point(e, cn->body1, cn->pos, Jp1, Jr1);
    jac_point(e, cn->body2, cn->pos, Jp2, Jr2);
    double Jc[3][OR_MAXV];
    for (int a = 0; a < 3; a++) for (int j = 0; j < nv; j++) {
      double s = 0;
      for (int d = 0; d < 3; d++) s += cn->frame[3 * a + d] * (Jp2[d * nv + j] - Jp1[d * nv + j]);
      Jc[a][j] = s;
    }
    for (int k = 0; k < 4; k++) {
      int t = 1 + (k >> 1); double sg = (k & 1) ? -cn->mu : cn->mu;
      for (int j = 0; j < nv; j++) J[nr][j] = Jc[0][j] + sg * Jc[t][j];
      rpos[nr] = cn->dist; rtype[nr] = ROW_CONTACT; nr++;
    }
  }
  e->nrow = nr;

  /* ---- A = J M^-1 J^T, regulariser R, reference acceleration */
  for (int r = 0; r < nr; r++) {
    memcpy(Bm[r], J[r], sizeof(double) * nv);
    if (sky) chol_solve_sky(L, nv, first, perm, Bm[r]); else chol_solve(L, nv, Bm[r]);
  }
  for (int r = 0; r < nr; r++) for (int c = r; c < nr; c++) {
    double s = 0;
    for (int j = 0; j < nv; j++) s += J[r][j] * Bm[c][j];
    A[r][c] = s; A[c][r] = s;
  }
  {
    double tc = fmax(m->solref[0], 2 * dt), dr = m->solref[1], dmax = m->solimp[1];
    double kk = 1.0 / (dmax * dmax * tc * tc * dr * dr), bb = 2.0 / (dmax * tc);
    for (int r = 0; r < nr; r++) {
      double imp = impedance(m->solimp, rpos[r]);
      double jv = 0, ja = 0;
      for (int j = 0; j < nv; j++) { jv += J[r][j] * e->qvel[j]; ja += J[r][j] * qacc_s[j]; }
      aref[r] = -bb * jv - kk * imp * rpos[r];
      Rr[r] = fmax(1e-15, (1 - imp) / imp * A[r][r]);
      bvec[r] = ja - aref[r];
    }
    /* pyramidal contacts (mj_makeImpedance): every edge of a contact shares
     * R = 2 mu^2 (1-d)/d diagApprox, diagApprox = (1 + mu^2) * A_nn, where A_nn is the diagonal of
     * the pure normal direction n = (e0 + e1)/2 (MuJoCo approximates it by body inverse weights) */
    for (int c = 0; c < e->ncon; c++) {
      int r0 = e->con[c].row; double mu = e->con[c].mu;
      double imp = impedance(m->solimp, rpos[r0]);
      double Ann = 0.25 * (A[r0][r0] + A[r0 + 1][r0 + 1] + 2 * A[r0][r0 + 1]);
      double Rpy = 2 * mu * mu * (1 - imp) / imp * (1 + mu * mu) * Ann;
      for (int k = 0; k < 4; k++) Rr[r0 + k] = fmax(1e-15, Rpy);
    }
  }
  /* ---- dual QP by projected Gauss-Seidel, warm-started, to convergence */
  if (nr == e->nrow_prev) memcpy(fr, e->efc_force_prev, sizeof(double) * nr); else memset(fr, 0, sizeof(double) * nr);
  int it = 0;
  for (; it < 4000 && nr > 0 && !e->fast_solver; it++) {
    double change = 0, scale = 0;
    for (int r = 0; r < nr; r++) {
      double res = bvec[r] + Rr[r] * fr[r];
      for (int c = 0; c < nr; c++) res += A[r][c] * fr[c];
      double fn = fr[r] - res / (A[r][r] + Rr[r]);
      if (rtype[r] != ROW_EQ && fn < 0) fn = 0;
      change = fmax(change, fabs(fn - fr[r])); scale = fmax(scale, fabs(fn));
      fr[r] = fn;
    }
    if (change <= 1e-13 * (1 + scale)) { it++; break; }
  }
  e->niter = it;
  /* ---- polish: the dual has a unique solution (A + R is SPD). Gauss-Seidel stalls on stiff squeezes
   * (4000 sweeps are not always enough), so the active set it found is refined by principal pivoting
   * with dense Cholesky solves until the complementarity conditions hold to 1e-12. */
  if (nr > 0) lcp_polish(nr, &A[0][0], OR_MAXROW, Rr, bvec, rtype, fr);
  memcpy(e->efc_force_prev, fr, sizeof(double) * nr); e->nrow_prev = nr;
  for (int j = 0; j < nv; j++) { e->qacc[j] = qacc_s[j]; }
  for (int r = 0; r < nr; r++) {
    for (int j = 0; j < nv; j++) { e->qacc[j] += Bm[r][j] * fr[r]; qfrc[j] += J[r][j] * fr[r]; }
  }
  /* contact forces in the contact frame (mj_contactForce for pyramidal condim 3) */
  for (int c = 0; c < e->ncon; c++) {
    OrContact* cn = &e->con[c]; const double* f4 = fr + cn->row;
    cn->force[0] = f4[0] + f4[1] + f4[2] + f4[3];
    cn->force[1] = cn->mu * (f4[0] - f4[1]);
    cn->force[2] = cn->mu * (f4[2] - f4[3]);
  }
  /* wrist force sensor z (quirk Q3): S_base . cfrc_int = generalised force supplied at the base
   * joint = actuator + passive (+ limit rows on that joint) */
  e->wrist_z = ctrl[0] - m->damping[0] * e->qvel[0] - m->stiffness[0] * e->qpos[0];
  for (int r = 0; r < nr; r++) if (rtype[r] == ROW_LIMIT) e->wrist_z += J[r][0] * fr[r];

  /* BADQACC (mj_checkAcc): |qacc| > 1e10 or nan */
  for (int j = 0; j < nv; j++) if (!(fabs(e->qacc[j]) < 1e10)) { e->unstable = 1; }

  /* ---- Euler, joint damping implicit: (M + dt D) a = qfrc_smooth + qfrc_constraint */
  for (int i = 0; i < B; i++) M[i * nv + i] += dt * m->damping[i];
  if (sky) {
    sky_setup(e, M, L, perm, first);
    if (!chol_sky(L, nv, first)) { e->unstable = 1; return; }
    chol_solve_sky(L, nv, first, perm, qfrc);
  } else {
    if (!chol(M, nv)) { e->unstable = 1; return; }
    chol_solve(M, nv, qfrc);
  }
  for (int j = 0; j < nv; j++) e->qvel[j] += dt * qfrc[j];
  for (int i = 0; i < B; i++) e->qpos[i] += dt * e->qvel[i];
  if (e->has_object) {
    double* q = e->qpos + B; const double* v = e->qvel + B;
    for (int k = 0; k < 3; k++) q[k] += dt * v[k];
    double w[3] = { v[3], v[4], v[5] };
    double ang = sqrt(dot3(w, w)) * dt;
    if (ang > 0) {
      double ax[3]; copy3(ax, w); normalize3(ax);
      double s = sin(ang / 2), dq[4] = { cos(ang / 2), ax[0] * s, ax[1] * s, ax[2] * s };
      double a0 = q[3], a1 = q[4], a2 = q[5], a3 = q[6];
      q[3] = a0 * dq[0] - a1 * dq[1] - a2 * dq[2] - a3 * dq[3];
      q[4] = a0 * dq[1] + a1 * dq[0] + a2 * dq[3] - a3 * dq[2];
      q[5] = a0 * dq[2] - a1 * dq[3] + a2 * dq[0] + a3 * dq[1];
      q[6] = a0 * dq[3] + a1 * dq[2] - a2 * dq[1] + a3 * dq[0];
    }
    double l = sqrt(q[3] * q[3] + q[4] * q[4] + q[5] * q[5] + q[6] * q[6]);
    for (int k = 3; k < 7; k++) q[k] /= l;
  }
  e->time += dt;
  e->st_substeps++; e->st_contacts += e->ncon; e->st_rows += nr; e->st_iters += it;
}

/* mj_forward subset used by reset/spawn: kinematics + collision, no integration */
void or_forward(OrEnv* e) { or_kinematics(e); or_collide(e); }
