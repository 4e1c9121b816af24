// gmj_physics.cuh -- one physics substep of one environment, executed by one warp.
//
// Replaces luke::step = mj_step1 ; control ; mj_step2 (src/myfunctions.cpp:1873-1898) for the
// gripper scene class. Pipeline (same semantics as MuJoCo's default options):
//   kinematics (level loop over the tree, parent pose by warp shuffle)
//   spatial inertias / velocities about a fixed reference point; prefix (velocity, bias
//   acceleration) and suffix (composite inertia, force) scans inside each finger chain
//   tree-sparse mass matrix: one dense (N+2)x(N+2) block per finger + base arrow
//   LDL^T with the base ordered last (no fill-in), rows of L held in registers
//   collision: lane-per-geom MPR against the object + closed-form plane contacts
//   constraint rows (welds, limits, pyramidal contacts) stored WHITENED, W = D^-1/2 L^-1 J^T,
//   so A = W W^T and no second copy (M^-1 J^T) is kept
//   dual projected Gauss-Seidel, one lane per row, warm-started
//   Euler with implicit joint damping through a second factorisation (M + h D)
#pragma once
#include "gmj_collide.cuh"
#include "gmj_tail.cuh"

template <typename T, int N, int QC = 40>
struct WarpShared {
  static constexpr int B = 3 * N + 8;
  static constexpr int NW = B + 6;          // whitened row: B gripper columns (lane order) + 6 object columns
  static constexpr int WS = NW | 1;         // odd stride: conflict-free when lanes walk rows
  EnvState<T, QC> e;
  WarpAux<T> ax;
  T W[GMJ_RMAX + 1][WS];            // whitened rows; row nrow holds the whitened smooth force
  // dual solve (wide kernels, -DGMJ_DUAL): packed lower triangle of W W^T, overwritten by its Cholesky factor;
  // primal solve (lane-per-dof kernels): only the PrimalScratch of gmj_primal.cuh lives here
#ifdef GMJ_DUAL
  static constexpr int ASZ = GMJ_RMAX * (GMJ_RMAX + 1) / 2;
#elif defined(GMJ_SELF_COLLISION)
  // + the dense (B + 6)^2 normal matrix of the fallback solve for rows that couple two leaf blocks (gmj_primal.cuh), with
  // its solution vector behind it; it lies over the primal scratch, which that solve does not use
  static constexpr int ASZ = QC == 40 ? (NW * (NW + 1) / 2 + NW + 3 > 3 * (N + 2) * (N + 3) + 8 * B + 36 + NW + 3 + 8
                                          ? NW * (NW + 1) / 2 + NW + 3 : 3 * (N + 2) * (N + 3) + 8 * B + 36 + NW + 3 + 8)
                                      : GMJ_RMAX * (GMJ_RMAX + 1) / 2;
#else
  static constexpr int ASZ = QC == 40 ? 3 * (N + 2) * (N + 3) + 8 * B + 36 + NW + 3 + 8 : GMJ_RMAX * (GMJ_RMAX + 1) / 2;
#endif
  T A[ASZ];
  T u[WS + 1];                      // W^T f scratch of the LCP solver
  T L2[3][N + 2][N + 3];            // second factor, for column access in the back substitution
  T qfk[QC];                        // qpos of the last kinematics pass (weld anchors)
  T objv[32];                       // uniform over the lanes, kept out of registers: object smooth force, 1/sqrt(M), M (6 each), position (3), rotation (9)
  T rpos[GMJ_RMAX], rjv[GMJ_RMAX], rmu[GMJ_RMAX];
  int rtype[GMJ_RMAX], rfirst[GMJ_RMAX], rlane[GMJ_RMAX];
};
// packed lower-triangular index; consecutive rows start at triangular numbers, which are all distinct
// mod 32 over any 32 consecutive rows, so "lane i reads L[i][k]" is bank-conflict free
#define TRI(i, j) (((i) * ((i) + 1) >> 1) + (j))

enum { ROW_EQ = 0, ROW_LIMIT = 1, ROW_CONTACT = 2 };

template <typename T> struct Sp { V3<T> w, v; };           // spatial motion (w; v) or force (n; f)
template <typename T> __device__ __forceinline__ Sp<T> operator+(Sp<T> a, Sp<T> b) { Sp<T> r; r.w = a.w + b.w; r.v = a.v + b.v; return r; }
template <typename T> __device__ __forceinline__ Sp<T> operator*(Sp<T> a, T s) { Sp<T> r; r.w = a.w * s; r.v = a.v * s; return r; }
template <typename T> __device__ __forceinline__ T spdot(Sp<T> a, Sp<T> b) { return dot(a.w, b.w) + dot(a.v, b.v); }
template <typename T> __device__ __forceinline__ Sp<T> spshfl(Sp<T> a, int src) { Sp<T> r; r.w = shfl3(a.w, src); r.v = shfl3(a.v, src); return r; }
template <typename T> __device__ __forceinline__ Sp<T> spup(Sp<T> a, int d) { Sp<T> r; r.w = shflup3(a.w, d); r.v = shflup3(a.v, d); return r; }
template <typename T> __device__ __forceinline__ Sp<T> spdn(Sp<T> a, int d) { Sp<T> r; r.w = shfldn3(a.w, d); r.v = shfldn3(a.v, d); return r; }
template <typename T> __device__ __forceinline__ Sp<T> crossm(Sp<T> v, Sp<T> s) { Sp<T> r; r.w = cross(v.w, s.w); r.v = cross(v.w, s.v) + cross(v.v, s.w); return r; }
template <typename T> __device__ __forceinline__ Sp<T> crossf(Sp<T> v, Sp<T> f) { Sp<T> r; r.w = cross(v.w, f.w) + cross(v.v, f.v); r.v = cross(v.w, f.v); return r; }
// spatial inertia about the reference point: mass, first moment h = m c, symmetric I (xx yy zz xy xz yz)
template <typename T> struct SpI { T m; V3<T> h; T I[6]; };
template <typename T> __device__ __forceinline__ V3<T> symmul(const T* I, V3<T> v) {
  return mk<T>(I[0] * v.x + I[3] * v.y + I[4] * v.z, I[3] * v.x + I[1] * v.y + I[5] * v.z, I[4] * v.x + I[5] * v.y + I[2] * v.z);
}
template <typename T> __device__ __forceinline__ Sp<T> spimul(const SpI<T>& I, Sp<T> s) {
  Sp<T> r; r.w = symmul(I.I, s.w) + cross(I.h, s.v); r.v = s.v * I.m + cross(s.w, I.h); return r;
}
template <typename T> __device__ __forceinline__ SpI<T> spiadd(SpI<T> a, const SpI<T>& b) {
  a.m += b.m; a.h = a.h + b.h;
#pragma unroll
  for (int k = 0; k < 6; k++) a.I[k] += b.I[k];
  return a;
}
template <typename T> __device__ __forceinline__ SpI<T> spishfl_dn(const SpI<T>& a, int d) {
  SpI<T> r; r.m = __shfl_down_sync(GMJ_FULL, a.m, d); r.h = shfldn3(a.h, d);
#pragma unroll
  for (int k = 0; k < 6; k++) r.I[k] = __shfl_down_sync(GMJ_FULL, a.I[k], d);
  return r;
}
template <typename T> __device__ __forceinline__ SpI<T> spishfl(const SpI<T>& a, int src) {
  SpI<T> r; r.m = __shfl_sync(GMJ_FULL, a.m, src); r.h = shfl3(a.h, src);
#pragma unroll
  for (int k = 0; k < 6; k++) r.I[k] = __shfl_sync(GMJ_FULL, a.I[k], src);
  return r;
}

// lane topology (compile-time arithmetic, no loads)
template <int N> struct Topo {
  static constexpr int NL = N + 2, B = 3 * N + 8, PALM = 3 * N + 7;
  int lane, fb, local, depth, parent; bool act, chain, hinge;
  __device__ __forceinline__ Topo(int l) {
    lane = l; act = l < B;
    if (l == 0) { fb = 0; local = 0; depth = 0; parent = 0; chain = false; hinge = false; }
    else if (l < PALM) { int f = (l - 1) / NL; local = (l - 1) - f * NL; fb = 1 + f * NL; depth = local + 1; parent = local == 0 ? 0 : l - 1; chain = true; hinge = local > 0; }
    else if (l == PALM) { fb = PALM; local = 0; depth = 1; parent = 0; chain = false; hinge = false; }
    else { fb = l; local = 0; depth = 99; parent = 0; chain = false; hinge = false; }
  }
};

// LDL^T of the block-arrow mass matrix. In: Mrow[k] = M[lane][fb+k] (k <= local), Mb = M[lane][base],
// lane 0: Mrow[0] = M[0][0]. Out: Mrow[k<local] = L entries, D = pivot of this lane, lb = L[base][lane].
template <typename T, int N>
__device__ __forceinline__ void ldl_factor(const Topo<N>& tp, T* Mrow, T Mb, T& D, T& lb) {
  constexpr int NL = N + 2;
#pragma unroll
  for (int k = 0; k < NL - 1; k++) {
    T Dk = __shfl_sync(GMJ_FULL, Mrow[k], tp.fb + k);
    T Mbk = __shfl_sync(GMJ_FULL, Mb, tp.fb + k);
    // lanes outside the finger chains read a meaningless (often zero) pivot; their quotient is never used, but an
    // IEEE division by 0 / denormal sends the WHOLE warp through ptxas' slow-path subroutine (ncu r1: 12 such calls
    // per substep from this line alone), so they divide by one instead
    T inv = rcpT<T>(tp.chain ? Dk : T(1));
    bool below = tp.chain && tp.local > k;
    T mk_pre = Mrow[k];
    T l = below ? mk_pre * inv : T(0);
#pragma unroll
    for (int j = k + 1; j < NL; j++) {
      T mjk = __shfl_sync(GMJ_FULL, mk_pre, tp.fb + j);
      if (below && tp.local >= j) Mrow[j] -= l * mjk;
    }
    if (below) { Mb -= (Mbk * inv) * mk_pre; Mrow[k] = l; }
  }
  // pivots
  T d = T(1);
#pragma unroll
  for (int k = 0; k < NL; k++) if (tp.local == k) d = Mrow[k];
  const T dsafe = (tp.lane != 0 && tp.act) ? d : T(1);       // same reason: no special-case operands on idle lanes
  const T dinv = rcpT<T>(dsafe);
  T contrib = (tp.lane != 0 && tp.act) ? Mb * Mb * dinv : T(0);
  T s = warp_sum(contrib);
  if (tp.lane == 0) { D = Mrow[0] - s; lb = 0; }
  else { D = d; lb = tp.act ? Mb * dinv : T(0); }
}

// x <- L^-1 x (lane-distributed)
template <typename T, int N>
__device__ __forceinline__ T ldl_fwd(const Topo<N>& tp, const T* Lrow, T lb, T x) {
  constexpr int NL = N + 2;
#pragma unroll
  for (int k = 0; k < NL - 1; k++) {
    T xk = __shfl_sync(GMJ_FULL, x, tp.fb + k);
    if (tp.chain && tp.local > k) x -= Lrow[k] * xk;
  }
  T s = warp_sum((tp.lane != 0 && tp.act) ? lb * x : T(0));
  if (tp.lane == 0) x -= s;
  return x;
}
// K right-hand sides at once: the chain recursion is latency bound (one shuffle + FMA per level), so
// independent columns ride along for free
template <typename T, int N, int K>
__device__ __forceinline__ void ldl_fwdK(const Topo<N>& tp, const T* Lrow, T lb, T (&x)[K]) {
  constexpr int NL = N + 2;
#pragma unroll
  for (int k = 0; k < NL - 1; k++) {
#pragma unroll
    for (int m = 0; m < K; m++) {
      T xk = __shfl_sync(GMJ_FULL, x[m], tp.fb + k);
      if (tp.chain && tp.local > k) x[m] -= Lrow[k] * xk;
    }
  }
  T sv[K];
#pragma unroll
  for (int m = 0; m < K; m++) sv[m] = (tp.lane != 0 && tp.act) ? lb * x[m] : T(0);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int m = 0; m < K; m++) sv[m] += __shfl_xor_sync(GMJ_FULL, sv[m], o);
  }
  if (tp.lane == 0) {
#pragma unroll
    for (int m = 0; m < K; m++) x[m] -= sv[m];
  }
}
// y = L x
template <typename T, int N>
__device__ __forceinline__ T ldl_mulL(const Topo<N>& tp, const T* Lrow, T lb, T x) {
  constexpr int NL = N + 2;
  T acc = x;
#pragma unroll
  for (int k = 0; k < NL - 1; k++) {
    T xk = __shfl_sync(GMJ_FULL, x, tp.fb + k);
    if (tp.chain && tp.local > k) acc += Lrow[k] * xk;
  }
  T s = warp_sum((tp.lane != 0 && tp.act) ? lb * x : T(0));
  if (tp.lane == 0) acc += s;
  return acc;
}

template <typename T> __device__ __forceinline__ T impedance(const T* solimp, T pos) {
  T d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  T x = absT<T>(pos) * rcpT<T>(width);
  if (x >= 1) return dw;
  if (x <= 0) return d0;
  T y;
  // power = 2 in every model of the reference (solimp "0.9 0.95 0.001 0.5 2"): no powf expansion on the hot path
  T a = (x <= mid) ? x * rcpT<T>(mid) : (1 - x) * rcpT<T>(1 - mid);
  T ap = (power == T(2)) ? a * a : powT<T>(a, power);
  if (x <= mid) y = ap * mid;
  else y = 1 - ap * (1 - mid);
  return d0 + y * (dw - d0);
}

// ---------------------------------------------------------------------------------------------
// store an already whitened row (lane value xw) + meta. Jo == nullptr: the row has no object part.
// (Spreading the twelve lane-0 stores over lanes 0-11 was measured 15 % SLOWER: seven divergent paths per row.)
template <typename T, int N>
__device__ __forceinline__ void put_row(WarpShared<T, N>& ws, const Topo<N>& tp, const T* rsqo, int r, T xw,
                                        const T* Jo, int type, T pos, T jv, T mu, int first, int jl) {
  if (tp.act) ws.W[r][tp.lane] = xw;
  if (tp.lane == 0) {
    if (Jo) {
#pragma unroll
      for (int k = 0; k < 6; k++) ws.W[r][3 * N + 8 + k] = Jo[k] * rsqo[k];
    }                                    // (no object part: the columns were cleared at the start of the row phase)
    ws.rtype[r] = type; ws.rpos[r] = pos; ws.rjv[r] = jv; ws.rmu[r] = mu; ws.rfirst[r] = first; ws.rlane[r] = jl;
  }
}

// point Jacobian entry of this lane's dof for a point p on gripper body `bl` (lane index) projected
// on direction d: d . (a x (p - anchor)) for hinges, d . a for slides, 0 if the dof is not an ancestor
template <typename T, int N>
__device__ __forceinline__ T jac_entry(const Topo<N>& tp, int bl, V3<T> p, V3<T> d, V3<T> aw, V3<T> anchor) {
  if (bl < 0 || !tp.act) return T(0);
  bool anc;
  if (tp.lane == 0) anc = true;
  else if (tp.lane == Topo<N>::PALM) anc = (bl == Topo<N>::PALM);
  else anc = (bl >= tp.fb && bl < tp.fb + Topo<N>::NL && bl >= tp.lane);
  if (!anc) return T(0);
  return tp.hinge ? dot(d, cross(aw, p - anchor)) : dot(d, aw);
}


// ------------------------------------------------------------------ A = W W^T, packed lower triangle
// Entry-parallel: the nr (nr + 1) / 2 packed entries are dealt 32 at a time, lane e computes
// A[i][j] = W_i . W_j for its entry (W_i is a broadcast or near-broadcast read, W_j walks rows with an odd
// stride). bcol (optional, per owned row lane / lane + 32) = W_r . W[nr], the extra column of the dual.
template <typename T, int N, int NT = 32, int QC>
__device__ void form_A(WarpShared<T, N, QC>& ws, int nr, int lane, T* bcol) {   // lane in [0, NT): NT threads share the entries
  constexpr int NW = WarpShared<T, N, QC>::NW;
  const int ne = nr * (nr + 1) / 2;
#pragma unroll 1
  for (int e0 = 0; e0 < ne; e0 += NT) {
    const int e = e0 + lane;
    const bool ok = e < ne;
    const int ee = ok ? e : 0;
    int i = (int)((sqrtf(8.0f * (float)ee + 1.0f) - 1.0f) * 0.5f);
    if ((i + 1) * (i + 2) / 2 <= ee) i++;
    if (i * (i + 1) / 2 > ee) i--;
    const int j = ee - i * (i + 1) / 2;
    const T* wi = ws.W[i];
    const T* wj = ws.W[j];
    T a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
    for (int k = 0; k + 3 < NW; k += 4) {
      a0 += wi[k] * wj[k]; a1 += wi[k + 1] * wj[k + 1]; a2 += wi[k + 2] * wj[k + 2]; a3 += wi[k + 3] * wj[k + 3];
    }
#pragma unroll
    for (int k = NW & ~3; k < NW; k++) a0 += wi[k] * wj[k];
    if (ok) ws.A[ee] = (a0 + a2) + (a1 + a3);
  }
  if (bcol) {
#pragma unroll 1
    for (int slot = 0; slot < 2; slot++) {
      const int r = lane + 32 * slot;
      bcol[slot] = 0;
      if (slot == 1 && nr <= 32) break;
      const T* wr = ws.W[r < nr ? r : 0];
      const T* wn = ws.W[nr];
      T a0 = 0, a1 = 0;
#pragma unroll
      for (int k = 0; k + 1 < NW; k += 2) { a0 += wr[k] * wn[k]; a1 += wr[k + 1] * wn[k + 1]; }
      if (NW & 1) a0 += wr[NW - 1] * wn[NW - 1];
      bcol[slot] = r < nr ? a0 + a1 : T(0);
    }
  }
}

// ------------------------------------------------------------------ dense LCP solver (warp)
// (A + R) f + b = w; equality rows: w = 0; inequality rows: f >= 0, w >= 0, f w = 0. Block principal
// pivoting over the active set; each round factors the active principal submatrix in place (packed
// lower triangle of ws.A, left-looking so column j still holds the original entries when it is
// reached) and gets w on the inactive rows from W (w = W (W^T f) + b). A second round has to rebuild
// A from W first; with the warm-started active set that happens in about 1 % of the substeps.
// Lane owns rows lane and lane + 32; b, R, type, f are per owned row. Returns the number of rounds.
// TWO = false: nr <= 32, one row slot per lane -- the usual case (17 rows on average), compiled without the
// second-slot arithmetic and the slot selects
template <typename T, int N, int QC, bool TWO>
__device__ int lcp_solve_impl(WarpShared<T, N, QC>& ws, int nr, int lane, const T* b, const T* R, const int* type, T* f, bool warm) {
  constexpr int NW = WarpShared<T, N, QC>::NW;
  T* A = ws.A;
#ifndef GMJ_LCP_GROW
#define GMJ_LCP_GROW 16
#define GMJ_LCP_GROW_FROM 2
#endif
#ifndef GMJ_LCP_EPS_WIDE
#define GMJ_LCP_EPS_WIDE 2e-6
#endif
  const T eps = sizeof(T) == 8 ? T(1e-12) : (NW > 40 ? T(GMJ_LCP_EPS_WIDE) : T(2e-6));
  const int maxit = 40;
  const int r0 = lane, r1 = lane + 32;
  const bool two = TWO && nr > 32;
  bool act0, act1;
  act0 = r0 < nr && (type[0] == ROW_EQ || (warm ? f[0] > 0 : b[0] < 0));
  act1 = (TWO && r1 < nr) && (type[1] == ROW_EQ || (warm ? f[1] > 0 : b[1] < 0));
  T bscale = warp_max(maxT<T>(r0 < nr ? absT<T>(b[0]) : T(0), (TWO && r1 < nr) ? absT<T>(b[1]) : T(0)));
  int best = 1 << 30, patience = 3, it = 0;
  T x0 = 0, x1 = 0;
  for (;; ) {
    if (it > 0) { __syncwarp(); form_A<T, N>(ws, nr, lane, nullptr); __syncwarp(); }
    it++;
    const unsigned m0 = __ballot_sync(GMJ_FULL, act0), m1 = __ballot_sync(GMJ_FULL, act1);
    // ---- left-looking Cholesky of K_FF (identity on the inactive rows), column j. The rows below the
    // diagonal are dealt to the lanes afresh for every column (row j + lane), so a second pass is only
    // needed while more than 32 rows remain -- not for every column of a system with more than 32 rows.
    T ld0 = 1, ld1 = 1;                       // 1 / L[r][r] of the owned rows
    const int t0 = TRI(r0, 0), t1 = TRI(r1, 0);
    for (int j = 0; j < nr; j++) {
      const bool jact = (!TWO || j < 32) ? ((m0 >> j) & 1) : ((m1 >> (j - 32)) & 1);
      const int tj = TRI(j, 0);
      const T Rj = __shfl_sync(GMJ_FULL, (!TWO || j < 32) ? R[0] : R[1], j & 31);
      T inv = 1;
#pragma unroll 1
      for (int base = j; base < nr; base += 32) {
        if (!TWO && base != j) break;
        const int i = base + lane;
        const bool iact = i < nr && (((!TWO || i < 32) ? (m0 >> i) : (m1 >> (i - 32))) & 1);
        const int ti = TRI(i, 0);
        T sv = 0;
        if (jact && iact) {
          T a = A[ti + j], c = 0, a2 = 0, c2 = 0;       // four chains: the loads, not the FMAs, bound this loop
          int k = 0;
          for (; k + 3 < j; k += 4) {
            a -= A[ti + k] * A[tj + k]; c -= A[ti + k + 1] * A[tj + k + 1];
            a2 -= A[ti + k + 2] * A[tj + k + 2]; c2 -= A[ti + k + 3] * A[tj + k + 3];
          }
          for (; k + 1 < j; k += 2) { a -= A[ti + k] * A[tj + k]; c -= A[ti + k + 1] * A[tj + k + 1]; }
          if (k < j) a -= A[ti + k] * A[tj + k];
          sv = (a + a2) + (c + c2) + (i == j ? Rj : T(0));
        }
        if (base == j) {
          const T dj = __shfl_sync(GMJ_FULL, sv, 0);
          inv = jact ? rsqrtT<T>(maxT<T>(dj, T(1e-30))) : T(1);
        }
        if (i > j && i < nr) A[ti + j] = sv * inv;
      }
      if (r0 == j) ld0 = inv;
      if (TWO && r1 == j) ld1 = inv;
      __syncwarp();
    }
    // ---- forward substitution L y = -b_F
    T y0 = act0 ? -b[0] : T(0), y1 = act1 ? -b[1] : T(0);
    for (int j = 0; j < nr; j++) {
      T yj = __shfl_sync(GMJ_FULL, (!TWO || j < 32) ? y0 * ld0 : y1 * ld1, j & 31);
      if (r0 == j) y0 = yj;
      if (TWO && r1 == j) y1 = yj;
      if (r0 > j && r0 < nr) y0 -= A[t0 + j] * yj;
      if (two && r1 > j && (TWO && r1 < nr)) y1 -= A[t1 + j] * yj;
    }
    // ---- back substitution L^T x = y
    for (int j = nr - 1; j >= 0; j--) {
      T xj = __shfl_sync(GMJ_FULL, (!TWO || j < 32) ? y0 * ld0 : y1 * ld1, j & 31);
      if (r0 == j) y0 = xj;
      if (TWO && r1 == j) y1 = xj;
      const int tj = TRI(j, 0);
      if (r0 < j) y0 -= A[tj + r0] * xj;
      if (two && r1 < j) y1 -= A[tj + r1] * xj;
    }
    x0 = act0 ? y0 : T(0); x1 = act1 ? y1 : T(0);
    // ---- w = W (W^T x) + b on the inactive rows (skipped when every row is active: nothing to check there)
    const bool need_w = __any_sync(GMJ_FULL, (r0 < nr && !act0) || (TWO && r1 < nr && !act1));
    if (need_w) {
      constexpr int NS = (NW + 31) / 32;       // column slots per lane: lane, lane + 32, ... (NW = 86 at N = 24)
      T u[NS];
#pragma unroll
      for (int s = 0; s < NS; s++) u[s] = 0;
      for (int c = 0; c < nr; c++) {
        T xc = __shfl_sync(GMJ_FULL, (!TWO || c < 32) ? x0 : x1, c & 31);
#pragma unroll
        for (int s = 0; s < NS; s++) if (lane + 32 * s < NW) u[s] += ws.W[c][lane + 32 * s] * xc;
      }
#pragma unroll
      for (int s = 0; s < NS; s++) if (lane + 32 * s < NW) ws.u[lane + 32 * s] = u[s];
      __syncwarp();
    }
    T w0 = r0 < nr ? b[0] : T(0), w1 = (TWO && r1 < nr) ? b[1] : T(0);
    if (r0 < nr && !act0) { T a = 0; for (int k = 0; k < NW; k++) a += ws.W[r0][k] * ws.u[k]; w0 += a; }
    if (two && (TWO && r1 < nr) && !act1) { T a = 0; for (int k = 0; k < NW; k++) a += ws.W[r1][k] * ws.u[k]; w1 += a; }
    T fscale = warp_max(maxT<T>(absT<T>(x0), absT<T>(x1)));
    // Safety net against pivoting on rounding noise: the tolerance widens x16 per round from the third round on
    // (capped at 1e-2). Systems that converge in one or two rounds -- 99 % of the substeps at every N -- never see
    // it; without it a noisy system spends up to maxit rounds per substep and holds its lockstep block.
    T epsr = eps;
    for (int k = GMJ_LCP_GROW_FROM; k < it && epsr < T(1e-2); k++) epsr *= GMJ_LCP_GROW;
    const T tolf = epsr * (1 + fscale), tolw = epsr * (1 + bscale);
    bool bad0 = r0 < nr && type[0] != ROW_EQ && (act0 ? x0 < -tolf : w0 < -tolw);
    bool bad1 = (TWO && r1 < nr) && type[1] != ROW_EQ && (act1 ? x1 < -tolf : w1 < -tolw);
    const unsigned b0 = __ballot_sync(GMJ_FULL, bad0), b1 = __ballot_sync(GMJ_FULL, bad1);
    const int ninf = __popc(b0) + __popc(b1);
    if (ninf == 0 || it >= maxit) { if (it > GMJ_LCP_GROW_FROM) it |= 1 << 16; if (ninf != 0) it |= 1 << 17; break; }
    bool block = true;
    if (ninf < best) { best = ninf; patience = 3; }
    else if (patience > 0) patience--;
    else block = false;
    if (block) { if (bad0) act0 = !act0; if (bad1) act1 = !act1; }
    else {                                     // Murty: flip only the infeasible row of largest index
      int top = b1 ? 32 + (31 - __clz(b1)) : (31 - __clz(b0));
      if (r0 == top) act0 = !act0;
      if (TWO && r1 == top) act1 = !act1;
    }
  }
  f[0] = (type[0] != ROW_EQ && x0 < 0) ? T(0) : x0;
  f[1] = (type[1] != ROW_EQ && x1 < 0) ? T(0) : x1;
  return it;
}

template <typename T, int N, int QC>
__device__ __forceinline__ int lcp_solve(WarpShared<T, N, QC>& ws, int nr, int lane, const T* b, const T* R, const int* type, T* f, bool warm) {
  return nr > 32 ? lcp_solve_impl<T, N, QC, true>(ws, nr, lane, b, R, type, f, warm)
                 : lcp_solve_impl<T, N, QC, false>(ws, nr, lane, b, R, type, f, warm);
}

#include "gmj_primal.cuh"

// ------------------------------------------------------------------ block lockstep
// The substep is ~10 k straight-line SASS instructions per warp against a 32 KB L1.5 I-cache, so it
// runs at the speed of instruction fetch. Phase barriers keep the warps of a block on the same code
// so that one fetch stream feeds all of them. Barrier 1 is used (0 is __syncthreads in the caller);
// warps that skip a substep (finished, frozen) call phase_idle() to pass the same barriers.
#ifndef GMJ_PHASE_MASK
#define GMJ_PHASE_MASK 0x3ff
#endif
#define GMJ_NPHASE 9
template <int NT> __device__ __forceinline__ void phase_barrier(int id) {
  if ((GMJ_PHASE_MASK >> id) & 1) asm volatile("bar.sync 1, %0;" :: "n"(NT) : "memory");
}
template <int NT> __device__ __forceinline__ void phase_idle() {
#pragma unroll
  for (int id = 0; id < GMJ_NPHASE; id++) phase_barrier<NT>(id);
}

// Development build -DGMJ_PROF: per-env cycle counters per phase (own work only; prof[10] = waits at the phase
// barriers), read back by gmj_prof_host (profiles/r2_env_cost.py). prof[k] = cycles of the code between phase barrier k
// and k + 1; prof[9] = per-substep tail; prof[11] = the whole substep loop.
#ifdef GMJ_PROF
#define GMJ_PROF_DECL long long prof_t0 = 0
#define GMJ_PROF_END(k) do { if (lane == 0) e.prof[k] += (unsigned)(clock64() - prof_t0); } while (0)
#define GMJ_PROF_BEGIN do { if (lane == 0) prof_t0 = clock64(); } while (0)
#define GMJ_PHASE(k) do { GMJ_PROF_END((k) - 1); GMJ_PROF_BEGIN; phase_barrier<NT>(k); GMJ_PROF_END(10); GMJ_PROF_BEGIN; } while (0)
#else
#define GMJ_PROF_DECL
#define GMJ_PROF_END(k)
#define GMJ_PROF_BEGIN
#define GMJ_PHASE(k) phase_barrier<NT>(k)
#endif

#ifdef GMJ_SELF_COLLISION
// Gripper <-> gripper contacts (M9; the reference's single collision group, src/objecthandler.cpp:135-212): plates and hooks
// of DIFFERENT fingers, and every plate / hook against the palm cylinder. Same-finger pairs that MuJoCo's parent-child
// filter lets through are two or more segments apart and cannot meet within the joint limits: not tested.
// Every finger tests against the next one (0 -> 1, 1 -> 2, 2 -> 0), so each pair of fingers is walked once; geom1 is the
// geom of the lower finger like MuJoCo's pair list. Returns the new contact count.
template <typename T, int N>
__device__ __noinline__ int self_collide(WarpAux<T>& ax, const Shape<T>& sg, const Shape<T>& sh, T bg, T bh, int gtype, int htype,
                                         int gcls, T gfric, int lane, int ncon) {
  constexpr int NL = N + 2, PALM = 3 * N + 7;
  const Topo<N> tp(lane);
  auto fetch = [&](const Shape<T>& s, int src) {
    Shape<T> o; o.type = __shfl_sync(GMJ_FULL, s.type, src); o.round = 0;
    o.size = shfl3(s.size, src); o.pos = shfl3(s.pos, src);
#pragma unroll
    for (int k = 0; k < 9; k++) o.R.m[k] = __shfl_sync(GMJ_FULL, s.R.m[k], src);
    return o;
  };
  // one candidate pair per lane: MPR if the bounding spheres overlap, then the hits are emitted in lane order
  auto pair = [&](bool on, const Shape<T>& mine, T bmine, const Shape<T>& other, int olane, int ocls, T ofr, bool mine_first) {
    int hit = 0; T dep = 0; V3<T> nn = mk<T>(0, 0, 1), pp = mk<T>(0, 0, 0);
    if (on && other.type >= 0) {
      const V3<T> d = other.pos - mine.pos; const T rr = bmine + shape_bound(other);
      if (dot(d, d) <= rr * rr) {                                  // coordinates relative to my geom (FP32 resolution)
        Shape<T> a = mine, b = other; a.pos = mk<T>(0, 0, 0); b.pos = d;
        hit = mine_first ? mpr_penetration<T>(a, b, &dep, &nn, &pp) : mpr_penetration<T>(b, a, &dep, &nn, &pp);
      }
    }
    unsigned hm = __ballot_sync(GMJ_FULL, hit);
    while (hm) {
      const int src = __ffs(hm) - 1; hm &= hm - 1;
      if (ncon < GMJ_MAXCON) {
        if (lane == src) {
          DevContact<T>& cn = ax.con[ncon];
          const V3<T> pos = pp + mine.pos;
          cn.pos[0] = pos.x; cn.pos[1] = pos.y; cn.pos[2] = pos.z;
          make_frame<T>(nn, cn.frame);
          cn.dist = -dep; cn.mu = maxT<T>(gfric, ofr);
          cn.body1 = mine_first ? lane : olane; cn.cls1 = mine_first ? gcls : ocls;
          cn.body2 = mine_first ? olane : lane; cn.cls2 = mine_first ? ocls : gcls;
          cn.force[0] = cn.force[1] = cn.force[2] = 0; cn.row = -1;
        }
        ncon++;
      }
    }
  };
  const int fi = tp.chain ? (lane - 1) / NL : 0, fnext = (fi + 1) % 3;
  const bool first = fi < fnext;                                   // my finger holds geom1 of the pair
  const bool hasg = tp.chain && gtype >= 0, hash = tp.chain && htype >= 0;
#pragma unroll 1
  for (int j = 2; j < NL; j++) {                                   // partner: segment j - 1 of the next finger (+ its hook)
    const int pl = tp.chain ? 1 + fnext * NL + j : lane;
    {                                                              // bounding spheres first: centres and radii only
      const V3<T> pg = shfl3(sg.pos, pl), ph = shfl3(sh.pos, pl);
      const T obg = __shfl_sync(GMJ_FULL, bg, pl), obh = __shfl_sync(GMJ_FULL, bh, pl);
      const int otg = __shfl_sync(GMJ_FULL, gtype, pl), oth = __shfl_sync(GMJ_FULL, htype, pl);
      auto near = [&](bool on, V3<T> a, T ra, int ot, V3<T> b, T rb) { const V3<T> d = b - a; const T rr = ra + rb; return on && ot >= 0 && dot(d, d) <= rr * rr; };
      const bool cand = near(hasg, sg.pos, bg, otg, pg, obg) || near(hasg, sg.pos, bg, oth, ph, obh) ||
                        near(hash, sh.pos, bh, otg, pg, obg) || near(hash, sh.pos, bh, oth, ph, obh);
      if (!__any_sync(GMJ_FULL, cand)) continue;
    }
    const Shape<T> og = fetch(sg, pl), oh = fetch(sh, pl);
    const int ocls = __shfl_sync(GMJ_FULL, gcls, pl); const T ofr = __shfl_sync(GMJ_FULL, gfric, pl);
    pair(hasg, sg, bg, og, pl, ocls, ofr, first); pair(hasg, sg, bg, oh, pl, ocls, ofr, first);
    pair(hash, sh, bh, og, pl, ocls, ofr, first); pair(hash, sh, bh, oh, pl, ocls, ofr, first);
  }
  {                                                                // every finger geom against the palm (geom2: the last geom)
    const Shape<T> op = fetch(sg, PALM);
    const int ocls = __shfl_sync(GMJ_FULL, gcls, PALM); const T ofr = __shfl_sync(GMJ_FULL, gfric, PALM);
    pair(hasg, sg, bg, op, PALM, ocls, ofr, true);
    pair(hash, sh, bh, op, PALM, ocls, ofr, true);
  }
  return ncon;
}
#endif

template <typename T, int N, int NT>
__device__ void physics_substep(WarpShared<T, N>& ws, const DevModel<T>& M, const DevSettings& S, int lane, double dtd, T tip_force) {
  const T dt = (T)dtd;
  constexpr int NL = N + 2, B = 3 * N + 8, PALM = 3 * N + 7;
  const Topo<N> tp(lane);
  EnvState<T>& e = ws.e;
  // after a BADQACC-class failure the env is frozen until the caller resets it (is_done reports it)
  if (e.unstable) { __syncwarp(); if (lane == 0) e.time += dtd; __syncwarp(); phase_idle<NT>(); return; }
  const T gz = -M.zref;                                   // ground plane height in shifted coordinates

  // ------------------------------------------------------------------ kinematics
  GMJ_PROF_DECL;
  phase_barrier<NT>(0); GMJ_PROF_BEGIN;
  T q = tp.act ? e.qpos[lane] : T(0), qd = tp.act ? e.qvel[lane] : T(0);
  ws.qfk[lane] = q;
  M3<T> Rl; V3<T> pl = mk<T>(M.f[MF_pos0][lane], M.f[MF_pos1][lane], M.f[MF_pos2][lane]);
  V3<T> axl = mk<T>(M.f[MF_ax0][lane], M.f[MF_ax1][lane], M.f[MF_ax2][lane]);
#pragma unroll
  for (int k = 0; k < 9; k++) Rl.m[k] = M.f[MF_R0 + k][lane];
  M3<T> R; V3<T> p = mk<T>(0, 0, 0), aw = mk<T>(0, 0, 1), anchor = mk<T>(0, 0, 0);
#pragma unroll
  for (int k = 0; k < 9; k++) R.m[k] = (k % 4 == 0) ? T(1) : T(0);
#ifdef GMJ_GENERIC_FK
#pragma unroll
  for (int d = 0; d <= NL; d++) {
    M3<T> Rp = shflM(R, tp.parent);
    V3<T> pp = shfl3(p, tp.parent);
    if (tp.depth == d) {
      M3<T> R0; V3<T> P0;
      if (d == 0) { R0 = Rl; P0 = pl; } else { R0 = mul(Rp, Rl); P0 = pp + mul(Rp, pl); }
      aw = mul(R0, axl); anchor = P0;
      if (!tp.hinge) { R = R0; p = P0 + aw * q; } else { R = mul(R0, rot_axis(axl, q)); p = P0; }
    }
  }
#else
  // Base, slides, finger revolute and FIRST segment by the depth walk (4 levels). The segments after the first hang on
  // parallel hinges (axis y of the previous segment, offset (L, 0, 0), no fixed rotation: checked on the host,
  // build_model), so the rest of a finger is a PLANAR chain: frame = Rb Roty(sum of angles), anchors = running sum of
  // L (cos, 0, -sin) in the frame Rb of the first segment's joint. Two short scans replace N - 1 levels of the walk
  // (each 12 shuffles + two 3x3 products that only the lanes of that depth needed).
  M3<T> Rb; V3<T> Pb = mk<T>(0, 0, 0);
#pragma unroll
  for (int k = 0; k < 9; k++) Rb.m[k] = 0;
#pragma unroll
  for (int d = 0; d <= 3; d++) {
    M3<T> Rp = shflM(R, tp.parent);
    V3<T> pp = shfl3(p, tp.parent);
    if (tp.depth == d) {
      M3<T> R0; V3<T> P0;
      if (d == 0) { R0 = Rl; P0 = pl; } else { R0 = mul(Rp, Rl); P0 = pp + mul(Rp, pl); }
      aw = mul(R0, axl); anchor = P0;
      if (!tp.hinge) { R = R0; p = P0 + aw * q; } else { R = mul(R0, rot_axis(axl, q)); p = P0; }
      if (d == 3) { Rb = R0; Pb = P0; }
    }
  }
  {
    const bool seg = tp.chain && tp.local >= 2, tail = tp.chain && tp.local >= 3;
    const int s = tp.local - 2;                               // segment index inside the finger (0 = first segment)
    T th = seg ? q : T(0);                                    // inclusive prefix of the segment angles
#pragma unroll
    for (int off = 1; off < N; off <<= 1) { const T t = __shfl_up_sync(GMJ_FULL, th, off); if (seg && s >= off) th += t; }
    const int src = tp.chain ? tp.fb + 2 : lane;
    Rb = shflM(Rb, src); Pb = shfl3(Pb, src);
    T sx, cx; sincosT<T>(th - q, &sx, &cx);                   // frame before this segment's own hinge
    T ux = tail ? pl.x * cx : T(0), uz = tail ? -pl.x * sx : T(0);
#pragma unroll
    for (int off = 1; off < N; off <<= 1) {
      const T tx = __shfl_up_sync(GMJ_FULL, ux, off), tz = __shfl_up_sync(GMJ_FULL, uz, off);
      if (seg && s >= off) { ux += tx; uz += tz; }
    }
    if (tail) {
      T si, ci; sincosT<T>(th, &si, &ci);
      aw = mk<T>(Rb.m[1], Rb.m[4], Rb.m[7]);                  // Rb ey: the same world axis for every hinge of the finger
      anchor = Pb + mk<T>(Rb.m[0] * ux + Rb.m[2] * uz, Rb.m[3] * ux + Rb.m[5] * uz, Rb.m[6] * ux + Rb.m[8] * uz);
#pragma unroll
      for (int r = 0; r < 3; r++) {
        R.m[3 * r] = Rb.m[3 * r] * ci - Rb.m[3 * r + 2] * si;
        R.m[3 * r + 1] = Rb.m[3 * r + 1];
        R.m[3 * r + 2] = Rb.m[3 * r] * si + Rb.m[3 * r + 2] * ci;
      }
      p = anchor;
    }
  }
#endif
  // local force frames for the tally (xmat of finger_i and palm)
  if (tp.chain && tp.local == 1) { int f = (lane - 1) / NL;
#pragma unroll
    for (int k = 0; k < 9; k++) ws.ax.fingerR[f][k] = R.m[k]; }
  if (lane == PALM) {
#pragma unroll
    for (int k = 0; k < 9; k++) ws.ax.palmR[k] = R.m[k]; }
  // object (replicated in every lane)
  // object pose: one lane computes, everybody reads it back from shared memory where it is needed (objv)
  T* const pov = ws.objv + 18; T* const Rov = ws.objv + 21;
  if (lane == 0) {
    pov[0] = e.qpos[32]; pov[1] = e.qpos[33]; pov[2] = e.qpos[34] - M.zref;
    T a = e.qpos[35], b = e.qpos[36], c = e.qpos[37], d = e.qpos[38];
    T l = T(1) / sqrtT<T>(a * a + b * b + c * c + d * d);
    M3<T> Ro = quat2mat<T>(a * l, b * l, c * l, d * l);
#pragma unroll
    for (int k = 0; k < 9; k++) Rov[k] = Ro.m[k];
  }
  const bool has_obj = e.has_object != 0;

  // ------------------------------------------------------------------ spatial quantities
  GMJ_PHASE(1);
  T mass = tp.act ? M.f[MF_mass][lane] : T(0);
  V3<T> c = p + mul(R, mk<T>(M.f[MF_ip0][lane], M.f[MF_ip1][lane], M.f[MF_ip2][lane]));
  SpI<T> I;
  {
    T ib[6] = { M.f[MF_ixx][lane], M.f[MF_iyy][lane], M.f[MF_izz][lane], M.f[MF_ixy][lane], M.f[MF_ixz][lane], M.f[MF_iyz][lane] };
    M3<T> Ib; Ib.m[0] = ib[0]; Ib.m[4] = ib[1]; Ib.m[8] = ib[2]; Ib.m[1] = Ib.m[3] = ib[3]; Ib.m[2] = Ib.m[6] = ib[4]; Ib.m[5] = Ib.m[7] = ib[5];
    M3<T> t = mul(R, Ib);
    // Iw = t R^T (symmetric)
    T Iw[6];
    Iw[0] = t.m[0] * R.m[0] + t.m[1] * R.m[1] + t.m[2] * R.m[2];
    Iw[1] = t.m[3] * R.m[3] + t.m[4] * R.m[4] + t.m[5] * R.m[5];
    Iw[2] = t.m[6] * R.m[6] + t.m[7] * R.m[7] + t.m[8] * R.m[8];
    Iw[3] = t.m[0] * R.m[3] + t.m[1] * R.m[4] + t.m[2] * R.m[5];
    Iw[4] = t.m[0] * R.m[6] + t.m[1] * R.m[7] + t.m[2] * R.m[8];
    Iw[5] = t.m[3] * R.m[6] + t.m[4] * R.m[7] + t.m[5] * R.m[8];
    T cc = dot(c, c);
    I.m = mass; I.h = c * mass;
    I.I[0] = Iw[0] + mass * (cc - c.x * c.x); I.I[1] = Iw[1] + mass * (cc - c.y * c.y); I.I[2] = Iw[2] + mass * (cc - c.z * c.z);
    I.I[3] = Iw[3] - mass * c.x * c.y; I.I[4] = Iw[4] - mass * c.x * c.z; I.I[5] = Iw[5] - mass * c.y * c.z;
    if (!tp.act) {
#pragma unroll
      for (int k = 0; k < 6; k++) I.I[k] = 0;
    }
  }
  Sp<T> Sj;
  if (!tp.act) { Sj.w = mk<T>(0, 0, 0); Sj.v = mk<T>(0, 0, 0); }
  else if (!tp.hinge) { Sj.w = mk<T>(0, 0, 0); Sj.v = aw; }
  else { Sj.w = aw; Sj.v = cross(anchor, aw); }
  // velocities: inclusive prefix inside each chain, then add the base
  Sp<T> v = Sj * qd;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) { Sp<T> t = spup(v, off); if (tp.chain && tp.local >= off) v = v + t; }
  Sp<T> vbase = spshfl(v, 0);
  if (lane != 0 && tp.act) v = v + vbase;
  // bias accelerations
  Sp<T> a = crossm(v, Sj) * qd;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) { Sp<T> t = spup(a, off); if (tp.chain && tp.local >= off) a = a + t; }
  Sp<T> abase = spshfl(a, 0);
  if (lane != 0 && tp.act) a = a + abase;
  a.v.z += M.gravity;
  // forces and composite inertias: inclusive suffix inside each chain, then gather into the base
  Sp<T> f = spimul(I, a) + crossf(v, spimul(I, v));
  SpI<T> Ic = I;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) {
    Sp<T> tf = spdn(f, off); SpI<T> ti = spishfl_dn(Ic, off);
    if (tp.chain && tp.local + off < NL) { f = f + tf; Ic = spiadd(Ic, ti); }
  }
  {
    Sp<T> fs = f; SpI<T> is = Ic;
#pragma unroll
    for (int r = 0; r < 4; r++) {
      int src = (r < 3) ? (1 + r * NL) : PALM;
      Sp<T> tf = spshfl(f, src); SpI<T> ti = spishfl(Ic, src);
      if (lane == 0) { fs = fs + tf; is = spiadd(is, ti); }
    }
    f = fs; Ic = is;
  }
  T bias = spdot(Sj, f);
  // mass matrix rows: M[lane][fb+k] and coupling to the base
  Sp<T> F = spimul(Ic, Sj);
  T Mrow[NL];
#pragma unroll
  for (int k = 0; k < NL; k++) {
    Sp<T> Sk = spshfl(Sj, tp.fb + k);
    Mrow[k] = (k <= tp.local) ? spdot(Sk, F) : T(0);
  }
  Sp<T> Sb = spshfl(Sj, 0);
  T Mb = (lane != 0 && tp.act) ? spdot(Sb, F) : T(0);

  // ------------------------------------------------------------------ control + smooth forces
  GMJ_PHASE(2);
  // luke::control_gripper / control_base (src/myfunctions.cpp:1965-2057)
  T damp = tp.act ? M.f[MF_damp][lane] : T(0), stiff = tp.act ? M.f[MF_stiff][lane] : T(0);
  T ctrl = 0;
  {
    int tsel = (int)M.f[MF_tsel][lane];
    T tgt = 0;
    if (tsel == 1) tgt = (T)e.base_z; else if (tsel == 2) tgt = (T)e.tnext.x; else if (tsel == 3) tgt = (T)e.tnext.th; else if (tsel == 4) tgt = (T)e.tnext.z;
    if (tsel != 0) ctrl = -((q - tgt) * M.f[MF_kp][lane] + qd * M.f[MF_kd][lane]);
  }
  T tau = -bias + ctrl - stiff * q - damp * qd;
  T wrist = ctrl - damp * qd - stiff * q;                 // lane 0: force sensor z (quirk Q3)
  if (tip_force != T(0)) {
    // calibration tip load on the last segment of each finger (src/myfunctions.cpp:1642-1727)
    int fidx = tp.chain ? (lane - 1) / NL : 0;
    int tl = M.tip_lane[fidx];
    V3<T> ct = shfl3(c, tl);
    V3<T> fv = mk<T>(-M.tip_mat[fidx][2] * tip_force, -M.tip_mat[fidx][5] * tip_force, -M.tip_mat[fidx][8] * tip_force);
    Sp<T> fs; fs.w = cross(ct, fv); fs.v = fv;
    if (tp.chain) tau += spdot(Sj, fs);
    // base: all three fingers
    T tb = 0;
#pragma unroll
    for (int ff = 0; ff < 3; ff++) {
      V3<T> c2 = shfl3(c, M.tip_lane[ff]);
      V3<T> f2 = mk<T>(-M.tip_mat[ff][2] * tip_force, -M.tip_mat[ff][5] * tip_force, -M.tip_mat[ff][8] * tip_force);
      Sp<T> s2; s2.w = cross(c2, f2); s2.v = f2;
      tb += spdot(Sj, s2);
    }
    if (lane == 0) tau += tb;
  }
  // object smooth force (free joint: world linear dofs, body-frame angular dofs)
  // 18 values every lane would otherwise hold for the whole substep (ncu r1: spills were 20 % of the L1 traffic)
  T* const tauo = ws.objv; T* const rsqo = ws.objv + 6; T* const Mo = ws.objv + 12;
  if (lane == 0) {
    const T om = e.obj[9];
    const V3<T> oI = mk<T>(e.obj[10], e.obj[11], e.obj[12]);
    const V3<T> wo = mk<T>(e.qvel[35], e.qvel[36], e.qvel[37]);
    V3<T> Iw = mk<T>(oI.x * wo.x, oI.y * wo.y, oI.z * wo.z);
    V3<T> gyr = cross(wo, Iw);
    tauo[0] = 0; tauo[1] = 0; tauo[2] = has_obj ? -om * M.gravity : T(0);
    tauo[3] = has_obj ? -gyr.x : T(0); tauo[4] = has_obj ? -gyr.y : T(0); tauo[5] = has_obj ? -gyr.z : T(0);
    Mo[0] = Mo[1] = Mo[2] = has_obj ? om : T(1);
    Mo[3] = has_obj ? oI.x : T(1); Mo[4] = has_obj ? oI.y : T(1); Mo[5] = has_obj ? oI.z : T(1);
#pragma unroll
    for (int k = 0; k < 6; k++) rsqo[k] = T(1) / sqrtT<T>(Mo[k]);
  }
  __syncwarp();

  // ------------------------------------------------------------------ factorisations
  GMJ_PHASE(3);
  T M2row[NL];
#pragma unroll
  for (int k = 0; k < NL; k++) M2row[k] = Mrow[k] + ((k == tp.local && tp.act) ? dt * damp : T(0));
  T D, lb, D2, lb2;
  ldl_factor<T, N>(tp, Mrow, Mb, D, lb);
  ldl_factor<T, N>(tp, M2row, Mb, D2, lb2);
  const T rsq = tp.act ? rsqrtT<T>(tp.act ? D : T(1)) : T(0);
  int bad = !(D > 0) && tp.act;
  // second factor to shared memory for column access
  if (tp.chain) { int fi = (lane - 1) / NL;
#pragma unroll
    for (int k = 0; k < NL; k++) ws.L2[fi][tp.local][k] = (k < tp.local) ? M2row[k] : T(0); }

  // whitened smooth force w_s = D^-1/2 L^-1 tau
  T wsm = ldl_fwd<T, N>(tp, Mrow, lb, tau) * rsq;
  __syncwarp();

  // ------------------------------------------------------------------ collision
  GMJ_PHASE(4);
  int ncon = 0;
  Shape<T> so;
  so.type = (int)e.obj[0]; so.size = mk<T>(e.obj[1], e.obj[2], e.obj[3]); so.round = e.obj[4];
  {
    M3<T> Ro;
#pragma unroll
    for (int k = 0; k < 9; k++) Ro.m[k] = Rov[k];
    so.pos = mk<T>(pov[0], pov[1], pov[2]); so.R = mul(Ro, quat2mat<T>(e.obj[5], e.obj[6], e.obj[7], e.obj[8]));
  }
  const T ofric = e.obj[13];
  auto add_contact = [&](V3<T> pos, V3<T> n, T dist, T mu, int b1, int c1, int b2, int c2, int slot) {
    DevContact<T>& cn = ws.ax.con[slot];
    cn.pos[0] = pos.x; cn.pos[1] = pos.y; cn.pos[2] = pos.z;
    make_frame<T>(n, cn.frame);
    cn.dist = dist; cn.mu = mu; cn.body1 = b1; cn.body2 = b2; cn.cls1 = c1; cn.cls2 = c2;
    cn.force[0] = cn.force[1] = cn.force[2] = 0; cn.row = -1;
  };
  // (1) object vs ground: lane i forms candidate i (8 box corners / 4 cylinder points / 1 ellipsoid point); the up to four
  // deepest penetrating ones become contacts, deepest first, ties by candidate index. (The serial version kept 8 points
  // per lane in local memory: ncu r2, 2 % of the stall samples.)
  if (has_obj) {
    V3<T> pc; T dc;
    plane_candidate_lane<T>(so, gz, lane, pc, dc);
    int rank = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) { const T dj = __shfl_sync(GMJ_FULL, dc, j); rank += (dj < 0 && (dj < dc || (dj == dc && j < lane))) ? 1 : 0; }
    const bool pen = lane < 8 && dc < 0;
    int np = __popc(__ballot_sync(GMJ_FULL, pen));
    if (np > 4) np = 4;
    if (np > GMJ_MAXCON - ncon) np = GMJ_MAXCON - ncon;
    if (pen && rank < np) {
      V3<T> pp = pc; pp.z -= T(0.5) * dc;
      add_contact(pp, mk<T>(0, 0, 1), dc, maxT<T>(M.ground_friction, ofric), -1, GMJ_CLS_GROUND, B, GMJ_CLS_OBJECT, ncon + rank);
    }
    ncon += np;
    __syncwarp();
  }
  // per-lane geoms: primary (segment plate / palm cylinder) and hook plate on the tip lanes
  Shape<T> sg, sh;
  const int gtype = tp.act ? (int)M.f[MF_gtype][lane] : -1, htype = tp.act ? (int)M.f[MF_htype][lane] : -1;
  const int gcls = (int)M.f[MF_gcls][lane];
  const T gfric = M.f[MF_gfric][lane];
  {
    M3<T> Rg, Rh;
#pragma unroll
    for (int k = 0; k < 9; k++) { Rg.m[k] = M.f[MF_gR0 + k][lane]; Rh.m[k] = M.f[MF_hR0 + k][lane]; }
    sg.type = gtype; sg.round = 0; sg.size = mk<T>(M.f[MF_gs0][lane], M.f[MF_gs1][lane], M.f[MF_gs2][lane]);
    sg.pos = p + mul(R, mk<T>(M.f[MF_gp0][lane], M.f[MF_gp1][lane], M.f[MF_gp2][lane])); sg.R = mul(R, Rg);
    sh.type = htype; sh.round = 0; sh.size = mk<T>(M.f[MF_hs0][lane], M.f[MF_hs1][lane], M.f[MF_hs2][lane]);
    sh.pos = p + mul(R, mk<T>(M.f[MF_hp0][lane], M.f[MF_hp1][lane], M.f[MF_hp2][lane])); sh.R = mul(R, Rh);
  }
  // (2) gripper plates vs ground (rare): lanes with a vertex below the plane emit in lane order,
  // primary geom first then the hook (== geom order of the model)
  {
    V3<T> pts[8]; T dd[8];
    bool lowg = false, lowh = false;
    if (gtype == GMJ_GEOM_BOX) { int n = plane_candidates<T>(sg, gz, pts, dd); for (int i = 0; i < n; i++) lowg |= dd[i] < 0; }
    if (htype == GMJ_GEOM_BOX) { int n = plane_candidates<T>(sh, gz, pts, dd); for (int i = 0; i < n; i++) lowh |= dd[i] < 0; }
    unsigned mg = __ballot_sync(GMJ_FULL, lowg), mh = __ballot_sync(GMJ_FULL, lowh);
    unsigned any = mg | mh;
    while (any) {
      int src = __ffs(any) - 1; any &= any - 1;
      for (int which = 0; which < 2; which++) {
        if (!(((which == 0) ? mg : mh) >> src & 1)) continue;
        int cnt = 0;
        if (lane == src) {
          const Shape<T>& s = which == 0 ? sg : sh;
          int n = plane_candidates<T>(s, gz, pts, dd);
          T mu = maxT<T>(M.ground_friction, gfric);
          for (int k = 0; k < 4 && k < n; k++) {
            int best = -1;
            for (int i = 0; i < n; i++) if (dd[i] < 0 && (best < 0 || dd[i] < dd[best])) best = i;
            if (best < 0) break;
            if (ncon + cnt < GMJ_MAXCON) { V3<T> pp = pts[best]; pp.z -= T(0.5) * dd[best]; add_contact(pp, mk<T>(0, 0, 1), dd[best], mu, -1, GMJ_CLS_GROUND, lane, gcls, ncon + cnt); cnt++; }
            dd[best] = 1;
          }
        }
        ncon += __shfl_sync(GMJ_FULL, cnt, src);
      }
    }
  }
#ifdef GMJ_SELF_COLLISION
  // (2b) gripper <-> gripper (M9): behind an early-out, in its own function so that the substep's registers are not
  // shaped by it. All geom centres lie in their finger's radial plane (120 deg apart), so with every centre further than
  // rho0 from the gripper axis (world z through the origin: the base is a z slide) two centres of different fingers are
  // at least sqrt(3) rho0 apart, and every plate is clear of the palm cylinder (r = 22.5 mm).
  {
    const T bg = gtype >= 0 ? shape_bound(sg) : T(0), bh = htype >= 0 ? shape_bound(sh) : T(0);
    T rho2 = T(1e30);
    if (tp.chain && gtype >= 0) rho2 = sg.pos.x * sg.pos.x + sg.pos.y * sg.pos.y;
    if (tp.chain && htype >= 0) rho2 = minT<T>(rho2, sh.pos.x * sh.pos.x + sh.pos.y * sh.pos.y);
    const T bmax = warp_max(tp.chain ? maxT<T>(bg, bh) : T(0));
    const T rmin2 = -warp_max(-rho2);
    const T rho0 = T(2) * bmax / T(1.7) + T(0.0225);
    if (rmin2 < rho0 * rho0) {
      Shape<T> cg = sg, ch = sh;
      ncon = self_collide<T, N>(ws.ax, cg, ch, bg, bh, gtype, htype, gcls, gfric, lane, ncon);
    }
  }
#endif
  // (3) object vs gripper geoms: one MPR query per lane (and a second pass for the hooks)
  if (has_obj) {
    T ro = shape_bound(so);
    int hitg = 0, hith = 0; T depg = 0, deph = 0; V3<T> ng, nh, pg, ph;
    // queries run in coordinates relative to the object centre (keeps FP32 resolution)
    Shape<T> so0 = so; so0.pos = mk<T>(0, 0, 0);
    bool rung = false, runh = false;
    if (gtype >= 0) {
      V3<T> d = sg.pos - so.pos; T rr = ro + shape_bound(sg);
      if (dot(d, d) <= rr * rr) { Shape<T> s0 = sg; s0.pos = d; hitg = mpr_penetration<T>(s0, so0, &depg, &ng, &pg); rung = true; }
    }
    if (htype >= 0) {
      V3<T> d = sh.pos - so.pos; T rr = ro + shape_bound(sh);
      if (dot(d, d) <= rr * rr) { Shape<T> s0 = sh; s0.pos = d; hith = mpr_penetration<T>(s0, so0, &deph, &nh, &ph); runh = true; }
    }
    { const int passes = (__any_sync(GMJ_FULL, rung) ? 1 : 0) + (__any_sync(GMJ_FULL, runh) ? 1 : 0); if (lane == 0) e.mpr_prev = passes; }
    unsigned mg = __ballot_sync(GMJ_FULL, hitg), mh = __ballot_sync(GMJ_FULL, hith);
    unsigned any = mg | mh;
    while (any) {
      int src = __ffs(any) - 1; any &= any - 1;
      if ((mg >> src) & 1) {
        if (ncon < GMJ_MAXCON) { if (lane == src) add_contact(pg + so.pos, ng, -depg, maxT<T>(gfric, ofric), lane, gcls, B, GMJ_CLS_OBJECT, ncon); ncon++; }
      }
      if ((mh >> src) & 1) {
        if (ncon < GMJ_MAXCON) { if (lane == src) add_contact(ph + so.pos, nh, -deph, maxT<T>(gfric, ofric), lane, gcls, B, GMJ_CLS_OBJECT, ncon); ncon++; }
      }
    }
  }
  else if (lane == 0) e.mpr_prev = 0;
  __syncwarp();

  // ------------------------------------------------------------------ constraint rows
  GMJ_PHASE(5);
  int nr = 0;
  // object columns of every row cleared by the whole warp at once: most rows (welds, limits, gripper-ground contacts)
  // have no object part, and their six zero stores from lane 0 were 2-3 % of the substep's stall samples (ncu r2)
#pragma unroll
  for (int j = 0; j < (6 * GMJ_RMAX + 31) / 32; j++) {
    const int idx = lane + 32 * j;
    if (idx < 6 * GMJ_RMAX) ws.W[idx / 6][3 * N + 8 + idx % 6] = T(0);
  }
  __syncwarp();
  // Welds and joint limits act on ONE joint each: every such row is a multiple of L^-1 e_j, so the unit
  // solves are done four joints at a time and the rows are scaled copies.
  // welds: joint `j` locked at its anchor; rows = world components of the joint axis
  // (weld residual of MuJoCo restricted to the single joint between the two bodies)
  {
    const unsigned won = e.weld_on & 0x7f;
#pragma unroll 1
    for (int g = 0; g < 2; g++) {
      unsigned grp = g == 0 ? (won & 0x0f) : (won >> 4);       // ids 0-3, then 4-6
      if (!grp) continue;
      T x[4];
#pragma unroll
      for (int m = 0; m < 4; m++) {
        const int id = 4 * g + m;
        const int j = (id < 3) ? (1 + id * NL) : ((id < 6) ? (2 + (id - 3) * NL) : PALM);
        x[m] = (((grp >> m) & 1) && lane == j) ? T(1) : T(0);
      }
      ldl_fwdK<T, N, 4>(tp, Mrow, lb, x);
#pragma unroll
      for (int m = 0; m < 4; m++) {
        const int id = 4 * g + m;
        if (id > 6 || !((grp >> m) & 1)) continue;
        const int j = (id < 3) ? (1 + id * NL) : ((id < 6) ? (2 + (id - 3) * NL) : PALM);
        V3<T> aj = shfl3(aw, j);
        T qj = __shfl_sync(GMJ_FULL, q, j), qdj = __shfl_sync(GMJ_FULL, qd, j);
        const bool rev = (id >= 3 && id < 6);
        const T sgn = rev ? T(-1) : T(1);
        T s1, c1; sincosT<T>(qj - e.weld_anchor[id], &s1, &c1);
        const T rho = rev ? s1 : (qj - e.weld_anchor[id]);
        const T comp[3] = { aj.x, aj.y, aj.z };
        const T xw = x[m] * rsq;
#pragma unroll
        for (int d = 0; d < 3; d++) {
          if (comp[d] * comp[d] < T(1e-10)) continue;
          if (nr >= GMJ_RMAX) break;
          put_row<T, N>(ws, tp, rsqo, nr, sgn * comp[d] * xw, nullptr, ROW_EQ, sgn * rho * comp[d], sgn * comp[d] * qdj, T(0), nr, j);
          nr++;
        }
      }
    }
  }
  // joint limits (margin 0)
  {
    T lo = q - M.f[MF_lo][lane], hi = M.f[MF_hi][lane] - q;
    unsigned mlo = __ballot_sync(GMJ_FULL, tp.act && lo < 0), mhi = __ballot_sync(GMJ_FULL, tp.act && hi < 0);
    unsigned any = mlo | mhi;
    while (any) {
      int jj[4]; T x[4];
#pragma unroll
      for (int m = 0; m < 4; m++) {
        jj[m] = any ? __ffs(any) - 1 : -1;
        if (any) any &= any - 1;
        x[m] = (lane == jj[m]) ? T(1) : T(0);
      }
      ldl_fwdK<T, N, 4>(tp, Mrow, lb, x);
#pragma unroll
      for (int m = 0; m < 4; m++) {
        const int j = jj[m];
        if (j < 0) continue;
        T loj = __shfl_sync(GMJ_FULL, lo, j), hij = __shfl_sync(GMJ_FULL, hi, j), qdj = __shfl_sync(GMJ_FULL, qd, j);
        const T xw = x[m] * rsq;
        if (((mlo >> j) & 1) && nr < GMJ_RMAX) { put_row<T, N>(ws, tp, rsqo, nr, xw, nullptr, ROW_LIMIT, loj, qdj, T(1), nr, j); nr++; }
        if (((mhi >> j) & 1) && nr < GMJ_RMAX) { put_row<T, N>(ws, tp, rsqo, nr, -xw, nullptr, ROW_LIMIT, hij, -qdj, T(-1), nr, j); nr++; }
      }
    }
  }
  // contacts: pyramidal, condim 3 -> 4 rows each, all four linear in the three frame rows (n, t1, t2),
  // which are whitened together
#pragma unroll 1
  for (int cidx = 0; cidx < ncon; cidx++) {
    if (nr + 4 > GMJ_RMAX) { ncon = cidx; break; }
    DevContact<T>& cn = ws.ax.con[cidx];
    V3<T> cp = mk<T>(cn.pos[0], cn.pos[1], cn.pos[2]);
    T mu = cn.mu;
    int b1 = cn.body1, b2 = cn.body2;
    T Jc[3], jvc[3];
#pragma unroll
    for (int ax3 = 0; ax3 < 3; ax3++) {
      V3<T> d = mk<T>(cn.frame[3 * ax3], cn.frame[3 * ax3 + 1], cn.frame[3 * ax3 + 2]);
      T j2 = (b2 < B) ? jac_entry<T, N>(tp, b2, cp, d, aw, anchor) : T(0);
      T j1 = jac_entry<T, N>(tp, b1, cp, d, aw, anchor);
      Jc[ax3] = j2 - j1;
      jvc[ax3] = Jc[ax3] * qd;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int ax3 = 0; ax3 < 3; ax3++) jvc[ax3] += __shfl_xor_sync(GMJ_FULL, jvc[ax3], o);
    }
    // object part (body2 == B; linear dofs world, angular dofs body frame), one entry per lane: lane 6 a + c holds the
    // object Jacobian of frame axis a, coordinate c. (Every lane used to form all 18 entries, the four rows' 24 combinations
    // and lane 0 stored them: ncu r2, put_row's object stores alone were 2.8 % of the stall samples.)
    const bool objc = b2 == B;
    T jo = 0;
    const int ja = lane / 6, jc = lane - 6 * ja;
    if (objc && lane < 18) {
      const V3<T> d = mk<T>(cn.frame[3 * ja], cn.frame[3 * ja + 1], cn.frame[3 * ja + 2]);
      if (jc < 3) jo = jc == 0 ? d.x : (jc == 1 ? d.y : d.z);
      else {
        const int k = jc - 3;
        const V3<T> r = cp - mk<T>(pov[0], pov[1], pov[2]);
        const V3<T> axk = mk<T>(Rov[k], Rov[3 + k], Rov[6 + k]);
        jo = dot(d, cross(axk, r));
      }
    }
    if (objc) {                                           // J_obj . v_obj per frame axis: sum over the six lanes of a group
      const T pv = lane < 18 ? jo * e.qvel[32 + jc] : T(0);
      const T t1 = pv + __shfl_down_sync(GMJ_FULL, pv, 1);
      const T t2 = t1 + __shfl_down_sync(GMJ_FULL, t1, 2);
#pragma unroll
      for (int ax3 = 0; ax3 < 3; ax3++) jvc[ax3] += __shfl_sync(GMJ_FULL, t2, 6 * ax3) + __shfl_sync(GMJ_FULL, t1, 6 * ax3 + 4);
    }
    const bool gripper_part = (b1 >= 0 && b1 < B) || (b2 >= 0 && b2 < B);
    T xw[3] = { Jc[0], Jc[1], Jc[2] };
    if (gripper_part) ldl_fwdK<T, N, 3>(tp, Mrow, lb, xw);      // object-ground contacts have no gripper columns
#pragma unroll
    for (int ax3 = 0; ax3 < 3; ax3++) xw[ax3] *= rsq;
    if (lane == 0) cn.row = nr;
    if (objc) {                                           // lane 6 k + c: object column c of pyramid row k = n +- mu t
      const int kr = ja, tr = 1 + (kr >> 1);
      const T sgl = (kr & 1) ? -mu : mu;
      const T jn = __shfl_sync(GMJ_FULL, jo, jc), jt = __shfl_sync(GMJ_FULL, jo, lane < 24 ? 6 * tr + jc : lane);
      if (lane < 24) ws.W[nr + kr][B + jc] = (jn + sgl * jt) * rsqo[jc];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int t = 1 + (k >> 1); const T sg2 = (k & 1) ? -mu : mu;
      put_row<T, N>(ws, tp, rsqo, nr + k, xw[0] + sg2 * xw[t], nullptr, ROW_CONTACT, cn.dist, jvc[0] + sg2 * jvc[t], mu, nr,
#ifdef GMJ_SELF_COLLISION
                    (b1 >= 0 && b1 < B && b2 >= 0 && b2 < B) ? -3 :            // -3: two gripper bodies, any column (dense solve)
#endif
                    gripper_part ? -(10 + ((b1 >= 0 && b1 < B) ? b1 : b2)) : -2);        // -(10 + gripper body lane) / -2: no gripper part
    }
    nr += 4;
  }
  // whitened smooth force as row nr
  if (tp.act) ws.W[nr][lane] = wsm;
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < 6; k++) ws.W[nr][B + k] = tauo[k] * rsqo[k];
    ws.ax.ncon = ncon; ws.ax.nrow = nr;
  }
  __syncwarp();

  // ------------------------------------------------------------------ A = W W^T (+ column W . w_s)
  GMJ_PHASE(6);
  // lane owns rows r0 = lane and r1 = lane + 32
  const int r0 = lane, r1 = lane + 32;
  int iters = 0;
  T g = 0, go = 0;
  const T tc = maxT<T>(M.solref_tc, 2 * dt), dr = M.solref_dr, dmax = M.solimp[1];
  const T kk = rcpT<T>(dmax * dmax * tc * tc * dr * dr), bb = T(2) * rcpT<T>(dmax * tc);
#ifndef GMJ_DUAL
  // ------------------------------------------------------------------ primal active-set solve (gmj_primal.cuh)
  {
    T pf[2] = { 0, 0 };
    GMJ_PHASE(7);
    if (nr > 0) iters = primal_solve<T, N>(ws, tp, M.solimp, kk, bb, nr, lane, e.nrow_prev == nr, e.fprev, pf, g, go);
    __syncwarp();
    if (r0 < nr) e.fprev[r0] = pf[0];
    if (r1 < nr) e.fprev[r1] = pf[1];
    if (lane == 0) { e.nrow_prev = nr; ws.ax.niter = iters; }
    __syncwarp();
    GMJ_PHASE(8);
    // g = whitened constraint force = x - x_s straight from the primal solution (sum_r W[r] f_r of the dual path)
  }
#else
  T bcol[2];
  form_A<T, N>(ws, nr, lane, bcol);
  __syncwarp();
  // regulariser, reference acceleration, PGS state per owned row
  T pf[2], py[2], pb[2], pR[2], pinv[2]; int ptype[2];
  {
#pragma unroll
    for (int slot = 0; slot < 2; slot++) {
      int r = slot ? r1 : r0;
      pf[slot] = 0; py[slot] = 0; pb[slot] = 0; pR[slot] = 1; pinv[slot] = 0; ptype[slot] = ROW_EQ;
      if (r < nr) {
        T pos = ws.rpos[r];
        T imp = impedance<T>(M.solimp, pos);
        T aref = -bb * ws.rjv[r] - kk * imp * pos;
        int type = ws.rtype[r];
        T Arr = ws.A[TRI(r, r)], Rv;
        if (type == ROW_CONTACT) {
          int rf = ws.rfirst[r]; T mu = ws.rmu[r];
          T Ann = T(0.25) * (ws.A[TRI(rf, rf)] + ws.A[TRI(rf + 1, rf + 1)] + 2 * ws.A[TRI(rf + 1, rf)]);
          Rv = 2 * mu * mu * (1 - imp) * rcpT<T>(imp) * (1 + mu * mu) * Ann;
        } else Rv = (1 - imp) * rcpT<T>(imp) * Arr;
        Rv = maxT<T>(Rv, T(1e-15));
        pR[slot] = Rv; pinv[slot] = rcpT<T>(Arr + Rv); pb[slot] = bcol[slot] - aref; ptype[slot] = type;
        pf[slot] = (e.nrow_prev == nr) ? e.fprev[r] : T(0);
      }
    }
  }
  __syncwarp();
  // ------------------------------------------------------------------ dual LCP, direct
  GMJ_PHASE(7);
  // (A + R) f + b = w, equality rows w = 0, inequality rows f >= 0, w >= 0, f w = 0. A + R is SPD so
  // the solution is unique: the oracle reaches it with PGS converged to 1e-13, here it is found by
  // block principal pivoting over the active set (warm-started from the previous substep) with an
  // in-place warp Cholesky per pivot -- a few thousand cycles instead of hundreds of GS sweeps.
  if (nr > 0) iters = lcp_solve<T, N>(ws, nr, lane, pb, pR, ptype, pf, e.nrow_prev == nr);
  if (r0 < nr) e.fprev[r0] = pf[0];
  if (r1 < nr) e.fprev[r1] = pf[1];
  if (lane == 0) { e.nrow_prev = nr; ws.ax.niter = iters; }
  __syncwarp();

  // ------------------------------------------------------------------ constraint force, integration
  GMJ_PHASE(8);
  for (int r = 0; r < nr; r++) {
    T fr = e.fprev[r];
    if (tp.act) g += ws.W[r][lane] * fr;
    if (lane < 6) go += ws.W[r][B + lane] * fr;
  }
#endif
  // tau_c = L D^1/2 g
  T tauc = ldl_mulL<T, N>(tp, Mrow, lb, tp.act ? g * sqrtT<T>(D) : T(0));
  // limit-row forces acting on the base joint enter the wrist force sensor
  {   // lane = row (at most one of the base joint's two limit rows can be active, so the lane sum has one term)
    T wl = 0;
    if (lane < nr && ws.rtype[lane] == ROW_LIMIT && ws.rlane[lane] == 0) wl = ws.rmu[lane] * e.fprev[lane];
    if (lane + 32 < nr && ws.rtype[lane + 32] == ROW_LIMIT && ws.rlane[lane + 32] == 0) wl += ws.rmu[lane + 32] * e.fprev[lane + 32];
    wl = warp_sum(wl);
    if (lane == 0) { wrist += wl; ws.ax.wrist_z = wrist; }
  }
  // contact forces in the contact frame (mj_contactForce, pyramidal condim 3)
  if (lane < ncon) {
    DevContact<T>& cn = ws.ax.con[lane];
    int rr = cn.row;
    T f0 = e.fprev[rr], f1 = e.fprev[rr + 1], f2 = e.fprev[rr + 2], f3 = e.fprev[rr + 3];
    cn.force[0] = f0 + f1 + f2 + f3; cn.force[1] = cn.mu * (f0 - f1); cn.force[2] = cn.mu * (f2 - f3);
  }
  // Euler with implicit damping: (M + h D) a = tau + tau_c
  T x = ldl_fwd<T, N>(tp, M2row, lb2, tau + tauc);
  x = tp.act ? x * rcpT<T>(tp.act ? D2 : T(1)) : T(0);
  {
    T xb = __shfl_sync(GMJ_FULL, x, 0);
    if (lane != 0 && tp.act) x -= lb2 * xb;
    int fi = tp.chain ? (lane - 1) / NL : 0;
#pragma unroll
    for (int k = NL - 1; k >= 1; k--) {
      T xk = __shfl_sync(GMJ_FULL, x, tp.fb + k);
      if (tp.chain && tp.local < k) x -= ws.L2[fi][k][tp.local] * xk;
    }
  }
  bad |= !(absT<T>(x) < T(1e10));
  bad |= !(D2 > 0) && tp.act;
  if (tp.act) { qd += dt * x; q += dt * qd; e.qpos[lane] = q; e.qvel[lane] = qd; }
  // object
  if (has_obj) {
    T ao[6];
#pragma unroll
    for (int k = 0; k < 6; k++) { T gk = __shfl_sync(GMJ_FULL, go, k); ao[k] = (tauo[k] + gk * rcpT<T>(rsqo[k])) * rcpT<T>(Mo[k]); bad |= !(absT<T>(ao[k]) < T(1e10)); }
    if (lane == 0) {
      V3<T> vo = mk<T>(e.qvel[32], e.qvel[33], e.qvel[34]), wo = mk<T>(e.qvel[35], e.qvel[36], e.qvel[37]);
      vo = vo + mk<T>(ao[0], ao[1], ao[2]) * dt; wo = wo + mk<T>(ao[3], ao[4], ao[5]) * dt;
      e.qvel[32] = vo.x; e.qvel[33] = vo.y; e.qvel[34] = vo.z; e.qvel[35] = wo.x; e.qvel[36] = wo.y; e.qvel[37] = wo.z;
      e.qpos[32] += dt * vo.x; e.qpos[33] += dt * vo.y; e.qpos[34] += dt * vo.z;
      e.obj_xy[0] += dtd * (double)vo.x; e.obj_xy[1] += dtd * (double)vo.y;
      T ang = sqrtT<T>(dot(wo, wo)) * dt;
      T a0 = e.qpos[35], a1 = e.qpos[36], a2 = e.qpos[37], a3 = e.qpos[38];
      if (ang > 0) {
        V3<T> axn = normalized(wo);
        T s, cs; sincosT<T>(ang / 2, &s, &cs);
        T d0 = cs, d1 = axn.x * s, d2 = axn.y * s, d3 = axn.z * s;
        T n0 = a0 * d0 - a1 * d1 - a2 * d2 - a3 * d3, n1 = a0 * d1 + a1 * d0 + a2 * d3 - a3 * d2;
        T n2 = a0 * d2 - a1 * d3 + a2 * d0 + a3 * d1, n3 = a0 * d3 + a1 * d2 - a2 * d1 + a3 * d0;
        a0 = n0; a1 = n1; a2 = n2; a3 = n3;
      }
      T l = T(1) / sqrtT<T>(a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3);
      e.qpos[35] = a0 * l; e.qpos[36] = a1 * l; e.qpos[37] = a2 * l; e.qpos[38] = a3 * l;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 6; k++) (void)__shfl_sync(GMJ_FULL, go, k);
  }
  unsigned anybad = __ballot_sync(GMJ_FULL, bad);
  GMJ_PROF_END(8);
  if (lane == 0) {
    if (anybad) e.unstable = 1;
    e.time += dtd;
    e.st_substeps++; e.st_contacts += ncon; e.st_rows += nr; e.st_iters += iters & 0xffff; e.st_relaxed += (iters >> 16) & 1; e.st_capped += (iters >> 17) & 1;
  }
  __syncwarp();
}
