/* vec.h -- small 3-vector / 3x3 helpers for the oracle (row-major matrices). TEST INFRASTRUCTURE ONLY. */
#ifndef ORACLE_VEC_H_
#define ORACLE_VEC_H_
#include <math.h>
static inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void cross3(double* o, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z; }
static inline void add3(double* o, const double* a, const double* b) { o[0] = a[0] + b[0]; o[1] = a[1] + b[1]; o[2] = a[2] + b[2]; }
static inline void sub3(double* o, const double* a, const double* b) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
static inline void scl3(double* o, const double* a, double s) { o[0] = a[0] * s; o[1] = a[1] * s; o[2] = a[2] * s; }
static inline void copy3(double* o, const double* a) { o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; }
static inline void zero3(double* o) { o[0] = o[1] = o[2] = 0; }
static inline void axpy3(double* o, double s, const double* a) { o[0] += s * a[0]; o[1] += s * a[1]; o[2] += s * a[2]; }
static inline double normalize3(double* a) { double l = sqrt(dot3(a, a)); if (l > 0) { a[0] /= l; a[1] /= l; a[2] /= l; } return l; }
/* o = M v */
static inline void mulv(double* o, const double* M, const double* v) {
  double x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2], y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2],
         z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z; }
/* o = M^T v */
static inline void mulTv(double* o, const double* M, const double* v) {
  double x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2], y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2],
         z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z; }
/* o = A B */
static inline void mulm(double* o, const double* A, const double* B) {
  double t[9];
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
    t[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
  for (int k = 0; k < 9; k++) o[k] = t[k]; }
/* o = A^T B */
static inline void mulTm(double* o, const double* A, const double* B) {
  double t[9];
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
    t[3 * r + c] = A[r] * B[c] + A[3 + r] * B[3 + c] + A[6 + r] * B[6 + c];
  for (int k = 0; k < 9; k++) o[k] = t[k]; }
static inline void quat2mat(double* R, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = 1 - 2 * (x * x + y * y); }
/* rotation about unit axis a by angle t */
static inline void axisangle2mat(double* R, const double* a, double t) {
  double c = cos(t), s = sin(t), v = 1 - c;
  R[0] = c + a[0] * a[0] * v; R[1] = a[0] * a[1] * v - a[2] * s; R[2] = a[0] * a[2] * v + a[1] * s;
  R[3] = a[1] * a[0] * v + a[2] * s; R[4] = c + a[1] * a[1] * v; R[5] = a[1] * a[2] * v - a[0] * s;
  R[6] = a[2] * a[0] * v - a[1] * s; R[7] = a[2] * a[1] * v + a[0] * s; R[8] = c + a[2] * a[2] * v; }
#endif
