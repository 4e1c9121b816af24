// gmj_primal_wide.cuh -- the primal active-set constraint solve (gmj_primal.cuh) for the wide kernel: 96 threads per
// environment, warp w = finger w (lane l < N + 2 = dof l of the chain), palm on lane 30 of warp 0, base replicated on
// lane 31 of every warp.
//
// Same problem and same pivoting rule as the lane-per-dof kernel: H x = g with H = I + sum_active w_r w_r^T / R_r,
// block-arrow with three (N + 2)^2 chain blocks + the palm as leaves and a 7 x 7 root (base + object 6). Here the three
// chain blocks are factored by the three warps AT THE SAME TIME (the dual solve of round 1 ran on warp 0 alone while
// the other two waited: ncu r2, 34 % of the step in lcp_solve + 8 % in W W^T, stall_barrier 22 %):
//   * thread = leaf dof: its row of the chain block in registers (right-looking elimination, pivot column published per
//     step through the warp's colb buffer as in ldl_factor_w), the 7 root couplings and the right-hand side ride along;
//   * thread = constraint row (tid < nr <= 48): regulariser, activity, residual, force;
//   * Schur complement of the 7 x 7 root summed over all leaves by 28 threads, solved redundantly by every thread.
// Scratch: the factor's chain blocks (packed lower triangles) live in the packed-A storage the dual solve used; the
// couplings, the root system and the solution in the row builders' per-slot joint arrays (sj .. jhi), dead by now.
#pragma once
#include <cstddef>
#include "gmj_primal.cuh"

template <typename T, int N> struct PrimalWideA {              // over ws.A
  static constexpr int NL = N + 2, TRI_ = NL * (NL + 1) / 2;
  T Lp[3][TRI_];
};
template <typename T, int N> struct PrimalWideB {              // over ws.sj .. ws.jhi
  static constexpr int B = 3 * N + 8, NW = B + 6;
  T hc[GMJ_WNT][8];         // per leaf: coupling row scaled by D^-1/2 (7) and y D^-1/2
  T root[36];               // Schur complement: 28 packed entries + 7 right-hand sides
  T x[NW + 2];              // solution (whitened acceleration), W column order
  T fmax[3], bmax[3];       // per-warp max |f| (per round) and max |aref|
  unsigned mask[2][2];      // [0]: active rows 0-31 / 32-47, [1]: infeasible rows
  int touch[GMJ_RMAX];      // per row: bit w = the row has entries in finger w's columns, bit 3 = palm column
};

template <typename T, int N>
__device__ __forceinline__ PrimalWideB<T, N>& primal_wide_b(WideShared<T, N>& ws) {
  static_assert(sizeof(PrimalWideA<T, N>) <= sizeof(ws.A), "chain blocks of the primal factor must fit the packed-A storage");
  // sj, jaw, jq, jqd, jlo, jhi are consecutive arrays of T in WideShared (no padding between them)
  static_assert(sizeof(ws.sj) + sizeof(ws.jaw) + sizeof(ws.jq) + sizeof(ws.jqd) + sizeof(ws.jlo) + sizeof(ws.jhi) >= sizeof(PrimalWideB<T, N>),
                "primal scratch must fit the row builders' joint arrays");
  return *reinterpret_cast<PrimalWideB<T, N>*>(&ws.sj[0][0]);
}

// Per row (thread = row): reference acceleration, regulariser from the row norms, first active set, which leaves the
// row touches. Leaves aref in ws.rpos and 1 / R in ws.rjv (as primal_prepare does).
template <typename T, int N>
__device__ __forceinline__ void primal_prepare_w(WideShared<T, N>& ws, PrimalWideB<T, N>& sb, const T* solimp, T kk, T bb, int nr, int tid,
                                                 bool warm, const T* fprev, bool& act, int& type, T& Rr, T& aref) {
  constexpr int NL = N + 2, B = 3 * N + 8, NW = B + 6;
  act = false; type = ROW_EQ; Rr = 1; aref = 0;
  if (tid < nr) {
    const int r = tid;
    const T pos = ws.rpos[r];
    const T imp = impedance<T>(solimp, pos);
    const T ar = -bb * ws.rjv[r] - kk * imp * pos;
    const int ty = ws.rtype[r];
    const T* wr = ws.W[r];
    const T* xs = ws.W[nr];
    T n0 = 0, n1 = 0, j0 = 0, j1 = 0;
    int touch = 0;
#pragma unroll 1
    for (int g = 0; g < 3; g++) {                       // the three finger column groups, then palm / base / object
      bool nz = false;
      for (int c = g * NL; c < (g + 1) * NL; c++) nz |= wr[c] != T(0);
      touch |= nz ? (1 << g) : 0;
    }
    touch |= (wr[3 * NL] != T(0)) ? 8 : 0;
    if (ty == ROW_CONTACT) {
      const int rf = ws.rfirst[r];
      const T* wa = ws.W[rf]; const T* wb = ws.W[rf + 1];
      for (int c = 0; c + 1 < NW; c += 2) {
        const T u0 = wa[c] + wb[c], u1 = wa[c + 1] + wb[c + 1];
        n0 += u0 * u0; n1 += u1 * u1;
        if (!warm) { j0 += wr[c] * xs[c]; j1 += wr[c + 1] * xs[c + 1]; }
      }
      if (NW & 1) { const T u0 = wa[NW - 1] + wb[NW - 1]; n0 += u0 * u0; if (!warm) j0 += wr[NW - 1] * xs[NW - 1]; }
      const T mu = ws.rmu[r];
      const T Ann = T(0.25) * (n0 + n1);
      Rr = 2 * mu * mu * (1 - imp) * rcpT<T>(imp) * (1 + mu * mu) * Ann;
    } else {
      for (int c = 0; c + 1 < NW; c += 2) {
        n0 += wr[c] * wr[c]; n1 += wr[c + 1] * wr[c + 1];
        if (!warm) { j0 += wr[c] * xs[c]; j1 += wr[c + 1] * xs[c + 1]; }
      }
      if (NW & 1) { n0 += wr[NW - 1] * wr[NW - 1]; if (!warm) j0 += wr[NW - 1] * xs[NW - 1]; }
      Rr = (1 - imp) * rcpT<T>(imp) * (n0 + n1);
    }
    Rr = maxT<T>(Rr, T(1e-15));
    aref = ar; type = ty;
    act = ty == ROW_EQ || (warm ? fprev[r] > 0 : (j0 + j1) - ar < 0);
    sb.touch[r] = touch;
  }
  env_sync<T, N>(ws);                                   // every row has read rpos / rjv / the contact rows it borrows
  if (tid < nr) { ws.rpos[tid] = aref; ws.rjv[tid] = rcpT<T>(Rr); }
}

// 7 x 7 Cholesky + solve of the root system (28 packed entries + 7 right-hand sides), compile-time indices: registers
template <typename T>
__device__ __forceinline__ void root7_solve(const T* root, T (&xr)[7]) {
  T Sm[28];
#pragma unroll
  for (int i = 0; i < 28; i++) Sm[i] = root[i];
#pragma unroll
  for (int m = 0; m < 7; m++) xr[m] = root[28 + m];
  static_for<0, 7>([&](auto jc) {
    constexpr int j = decltype(jc)::value, jj = j * (j + 1) / 2;
    T d = Sm[jj + j];
    static_for<0, j>([&](auto kc) { constexpr int k = decltype(kc)::value; d -= Sm[jj + k] * Sm[jj + k]; });
    const T inv = rsqrtT<T>(maxT<T>(d, T(1e-30)));
    Sm[jj + j] = inv;
    static_for<j + 1, 7>([&](auto ic) {
      constexpr int i = decltype(ic)::value, ii = i * (i + 1) / 2;
      T v = Sm[ii + j];
      static_for<0, j>([&](auto kc) { constexpr int k = decltype(kc)::value; v -= Sm[ii + k] * Sm[jj + k]; });
      Sm[ii + j] = v * inv;
    });
  });
  static_for<0, 7>([&](auto ic) {
    constexpr int i = decltype(ic)::value, ii = i * (i + 1) / 2;
    static_for<0, i>([&](auto kc) { constexpr int k = decltype(kc)::value; xr[i] -= Sm[ii + k] * xr[k]; });
    xr[i] *= Sm[ii + i];
  });
  static_for<0, 7>([&](auto rc) {
    constexpr int i = 6 - decltype(rc)::value, ii = i * (i + 1) / 2;
    static_for<i + 1, 7>([&](auto kc) { constexpr int k = decltype(kc)::value; xr[i] -= Sm[k * (k + 1) / 2 + i] * xr[k]; });
    xr[i] *= Sm[ii + i];
  });
}

// The whole solve for one environment (all 96 threads call it). Out: force of row `tid` in frow, x - x_s of this thread's
// gripper coordinate in dx (base replicas included) and of the object coordinates in dxo (tid 0-5). Returns the number
// of pivoting rounds (+ flags in bits 16 / 17, see primal_solve).
template <typename T, int N>
__device__ int primal_solve_w(WideShared<T, N>& ws, const TopoW<N>& tp, const T* solimp, T kk, T bb, int nr, int tid, bool warm,
                              const T* fprev, T& frow, T& dx, T& dxo) {
  constexpr int NL = N + 2, B = 3 * N + 8, NW = B + 6, BASEC = 3 * NL + 1;
  PrimalWideA<T, N>& sa = *reinterpret_cast<PrimalWideA<T, N>*>(ws.A);
  PrimalWideB<T, N>& sb = primal_wide_b<T, N>(ws);
  const int w = tp.w, lane = tp.lane;
  const bool leaf = tp.chain || tp.palm;
  const T eps = sizeof(T) == 8 ? T(1e-12) : T(2e-6);
  const int maxit = 40;
  bool act; int type; T Rr, aref;
  primal_prepare_w<T, N>(ws, sb, solimp, kk, bb, nr, tid, warm, fprev, act, type, Rr, aref);
  {
    const T am = warp_max(tid < nr ? absT<T>(aref) : T(0));
    const unsigned m = __ballot_sync(GMJ_FULL, act);
    if (lane == 0) { sb.bmax[w] = am; if (w < 2) sb.mask[0][w] = m; }
  }
  env_sync<T, N>(ws);
  const T bscale = maxT<T>(sb.bmax[0], maxT<T>(sb.bmax[1], sb.bmax[2]));
  unsigned m0 = sb.mask[0][0], m1 = sb.mask[0][1];
  const T* xs = ws.W[nr];
  // root entry owned by this thread (tid < 28): e = m (m + 1) / 2 + n, m >= n; root coordinate 0 = base, 1-6 = object
  int em = 0, en = 0;
  { int e = tid < 28 ? tid : 0; while ((em + 1) * (em + 2) / 2 <= e) em++; en = e - em * (em + 1) / 2; }
  const int cm = em == 0 ? BASEC : B + em - 1, cn = en == 0 ? BASEC : B + en - 1;
  int best = 1 << 30, patience = 3, it = 0;
  T xr[7], x = 0, f = 0;
  for (;;) {
    it++;
    // ---- H and g: one rank-one term per active row; a warp only walks the rows that touch its finger (or the palm)
    T Hrow[NL], Hc[7], Hd = T(1), gl = leaf ? xs[tp.col] : T(0);
#pragma unroll
    for (int k = 0; k < NL; k++) Hrow[k] = (tp.chain && k == tp.local) ? T(1) : T(0);
#pragma unroll
    for (int m = 0; m < 7; m++) Hc[m] = 0;
    T Se = (tid < 28 && em == en) ? T(1) : T(0);
    T gr = (tid < 28 && em == en) ? xs[cm] : T(0);
#pragma unroll 1
    for (int r = 0; r < nr; r++) {
      if (!(((r < 32 ? m0 >> r : m1 >> (r - 32))) & 1)) continue;
      const T* wr = ws.W[r];
      const T s = ws.rjv[r], sar = s * ws.rpos[r];        // 1 / R_r and aref_r / R_r (primal_prepare_w)
      const int tch = sb.touch[r];
      const bool mine = ((tch >> w) & 1) != 0, palmrow = w == 0 && (tch & 8);
      if (mine || palmrow) {                              // warp-uniform
        const T wc = leaf ? wr[tp.col] : T(0);
        const T swc = s * wc;
        if (mine) {
          const T* wf = wr + w * NL;
#pragma unroll
          for (int k = 0; k < NL; k++) if (tp.chain && k <= tp.local) Hrow[k] += swc * wf[k];
        }
        if (tp.palm) Hd += swc * wc;
        Hc[0] += swc * wr[BASEC];
#pragma unroll
        for (int m = 1; m < 7; m++) Hc[m] += swc * wr[B + m - 1];
        gl += sar * wc;
      }
      if (tid < 28) { const T a = wr[cm], b = wr[cn]; Se += s * a * b; if (em == en) gr += sar * a; }
    }
    // ---- leaves: right-looking LDL^T of this warp's chain block; pivot column k goes through the warp's colb buffer,
    // the couplings and the right-hand side of the pivot lane by shuffle
#pragma unroll
    for (int k = 0; k < NL - 1; k++) {
      T* col = ws.colb[w][k & 1];
      col[lane] = Hrow[k];
      __syncwarp();
      const T inv = rcpT<T>(tp.chain ? col[k] : T(1));
      const bool below = tp.chain && tp.local > k;
      const T l = below ? Hrow[k] * inv : T(0);
#pragma unroll
      for (int j4 = (k + 1) / 4; j4 < (NL + 3) / 4; j4++) {
        const Quad<T> c = reinterpret_cast<const Quad<T>*>(col)[j4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int j = 4 * j4 + u;
          if (j > k && j < NL) Hrow[j] -= l * c.v[u];
        }
      }
#pragma unroll
      for (int m = 0; m < 7; m++) { const T cp = __shfl_sync(GMJ_FULL, Hc[m], k); Hc[m] -= l * cp; }
      { const T gp = __shfl_sync(GMJ_FULL, gl, k); gl -= l * gp; }
      if (below) Hrow[k] = l;
    }
    T D = tp.palm ? Hd : T(1);
#pragma unroll
    for (int k = 0; k < NL; k++) if (tp.chain && tp.local == k) D = Hrow[k];
    const T dinv = rcpT<T>(D), dis = sqrtT<T>(dinv);
    if (tp.chain) {
      T* Lrow = &sa.Lp[w][tp.local * (tp.local + 1) / 2];
#pragma unroll
      for (int k = 0; k < NL - 1; k++) if (k < tp.local) Lrow[k] = Hrow[k];
    }
#pragma unroll
    for (int m = 0; m < 7; m++) sb.hc[tid][m] = leaf ? Hc[m] * dis : T(0);
    sb.hc[tid][7] = leaf ? gl * dis : T(0);
    env_sync<T, N>(ws);
    // ---- Schur complement on the root: S -= sum_leaves u u^T, g_root -= sum_leaves u y'
    if (tid < 28) {
      T a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll 2
      for (int c = 0; c < GMJ_WNT; c += 2) {
        a0 += sb.hc[c][em] * sb.hc[c][en]; a1 += sb.hc[c + 1][em] * sb.hc[c + 1][en];
        b0 += sb.hc[c][em] * sb.hc[c][7]; b1 += sb.hc[c + 1][em] * sb.hc[c + 1][7];
      }
      sb.root[tid] = Se - (a0 + a1);
      if (em == en) sb.root[28 + em] = gr - (b0 + b1);
    }
    env_sync<T, N>(ws);
    root7_solve<T>(sb.root, xr);
    // ---- leaves back: x_c = (y_c - Hc_c . x_root) / D_c - sum_{below} L[i][c] x_i
    x = gl;
#pragma unroll
    for (int m = 0; m < 7; m++) x -= Hc[m] * xr[m];
    x = leaf ? x * dinv : T(0);
#pragma unroll 2
    for (int k = NL - 1; k >= 1; k--) {
      const T xk = __shfl_sync(GMJ_FULL, x, k);
      if (tp.chain && tp.local < k) x -= sa.Lp[w][k * (k + 1) / 2 + tp.local] * xk;
    }
    if (tp.base) x = xr[0];
    if (tp.own) sb.x[tp.col] = x;
    if (tid < 6) sb.x[B + tid] = xr[1 + tid];
    env_sync<T, N>(ws);
    // ---- residual of row tid: jar = w_r . x - aref_r; force of an active row -jar / R, slack of an inactive row jar
    T jr = 0;
    if (tid < nr) {
      const T* wr = ws.W[tid];
      T a0 = 0, a1 = 0;
      for (int c = 0; c + 1 < NW; c += 2) { a0 += wr[c] * sb.x[c]; a1 += wr[c + 1] * sb.x[c + 1]; }
      if (NW & 1) a0 += wr[NW - 1] * sb.x[NW - 1];
      jr = (a0 + a1) - aref;
    }
    f = act ? -jr * rcpT<T>(Rr) : T(0);
    { const T fm = warp_max(absT<T>(f)); if (lane == 0) sb.fmax[w] = fm; }
    env_sync<T, N>(ws);
    const T fscale = maxT<T>(sb.fmax[0], maxT<T>(sb.fmax[1], sb.fmax[2]));
    T epsr = eps;
    for (int k = GMJ_LCP_GROW_FROM; k < it && epsr < T(1e-2); k++) epsr *= GMJ_LCP_GROW;
    const T tolf = epsr * (1 + fscale), tolw = epsr * (1 + bscale);
    const bool bad = tid < nr && type != ROW_EQ && (act ? f < -tolf : jr < -tolw);
    { const unsigned bm = __ballot_sync(GMJ_FULL, bad); if (lane == 0 && w < 2) sb.mask[1][w] = bm; }
    env_sync<T, N>(ws);
    const unsigned b0 = sb.mask[1][0], b1 = sb.mask[1][1];
    const int ninf = __popc(b0) + __popc(b1);
    if (ninf == 0 || it >= maxit) { if (it > GMJ_LCP_GROW_FROM) it |= 1 << 16; if (ninf != 0) it |= 1 << 17; break; }
    bool block = true;
    if (ninf < best) { best = ninf; patience = 3; }
    else if (patience > 0) patience--;
    else block = false;
    if (block) { if (bad) act = !act; }
    else {                                     // Murty: flip only the infeasible row of largest index
      const int top = b1 ? 32 + (31 - __clz(b1)) : (31 - __clz(b0));
      if (tid == top) act = !act;
    }
    { const unsigned m = __ballot_sync(GMJ_FULL, act); if (lane == 0 && w < 2) sb.mask[0][w] = m; }
    env_sync<T, N>(ws);
    m0 = sb.mask[0][0]; m1 = sb.mask[0][1];
  }
  frow = (type != ROW_EQ && f < 0) ? T(0) : f;
  dx = tp.act ? x - xs[tp.col] : T(0);
  dxo = tid < 6 ? xr[1 + tid] - xs[B + tid] : T(0);
  return it;
}
