// gmj_primal.cuh -- constraint solve of the lane-per-dof kernel in PRIMAL (acceleration) space.
//
// The soft-constraint problem of one substep is, in whitened coordinates x = D^1/2 L^T qacc (M = L D L^T),
//     minimise  1/2 |x - x_s|^2  +  sum_{r active}  (w_r . x - aref_r)^2 / (2 R_r)
// with w_r the whitened constraint rows (ws.W), x_s the whitened smooth force (row nr of ws.W), R_r MuJoCo's
// regulariser; row r is active when it is an equality or its force f_r = -(w_r . x - aref_r) / R_r is positive.
// Round 1 solved the DUAL of this, (A + R) f + b = w with A = W W^T, by block principal pivoting with a dense warp
// Cholesky of the active rows: the r2 per-env cycle counters put W W^T + the factorisation at 34 % of a substep on
// average (17 rows) and at 100 k of 180 k cycles for the envs that hold a lockstep block back (46 rows) -- dense work
// that grows with the square / cube of the row count.
// The primal normal matrix  H = I + sum_active w_r w_r^T / R_r  has the sparsity of the mass matrix instead: a whitened
// row touches one finger chain (or the palm), the base and the object, never two fingers. So H is block-arrow --
// three dense (N + 2)^2 finger blocks + the palm, each coupled only to the 7 root coordinates (base + object 6) -- and
// its factorisation is the tree LDL^T the kernel already runs for M, whatever the number of rows: leaves in registers
// (lane = row, pivot rows by shuffle) with the 7 coupling columns and the right-hand side riding along, then a 7 x 7
// Schur complement solved redundantly by every lane. Rows only enter through the rank-one accumulation of H, which is
// linear in the row count. The pivoting rule (flip every violated row, Murty's least-index rule after three
// non-improving rounds), tolerances and warm start are those of the dual solver; the solution is the same unique one.
#pragma once
#include <type_traits>

// scratch laid over the packed-A storage of WarpShared (unused in primal mode)
// compile-time loop: f(std::integral_constant<int, I>) for I in [I0, I1)
template <int I0, int I1, class F> __device__ __forceinline__ void static_for(F&& f) {
  if constexpr (I0 < I1) { f(std::integral_constant<int, I0>{}); static_for<I0 + 1, I1>(f); }
}

template <typename T, int N> struct PrimalScratch {
  static constexpr int NL = N + 2, B = 3 * N + 8;
  T Lc[3][NL][NL + 1];      // finger blocks of the factor, for the column walk of the back substitution
  T hc[B][8];               // per leaf: coupling row scaled by D^-1/2 (7) and y D^-1/2
  T root[36];               // Schur complement: 28 packed entries + 7 right-hand sides
  T x[WarpShared<T, N>::NW + 3];   // solution (whitened acceleration)
  unsigned tmask[4][2];     // rows (lane, lane + 32) whose gripper part lies in finger 0-2 / on the palm
};

// H x = g for the current active set. act0/act1: this lane's rows lane, lane + 32 active. Returns x in sc.x and this
// lane's entry in xl (gripper coordinate `lane`), xr[7] the root solution (base, object 0-5), replicated.
template <typename T, int N>
__device__ __forceinline__ void primal_factor_solve(WarpShared<T, N>& ws, PrimalScratch<T, N>& sc, const Topo<N>& tp, int nr, int lane,
                                                     unsigned m0, unsigned m1, T& xl, T (&xr)[7]) {
  constexpr int NL = N + 2, B = 3 * N + 8;
  const bool leaf = tp.act && lane != 0;
  // root entry owned by this lane: e = m (m + 1) / 2 + n, m >= n
  int em = 0, en = 0;
  { int e = lane < 28 ? lane : 0; while ((em + 1) * (em + 2) / 2 <= e) em++; en = e - em * (em + 1) / 2; }
  const int cm = em == 0 ? 0 : B + em - 1, cn = en == 0 ? 0 : B + en - 1;      // W columns of the two root coordinates
  const T* xs = ws.W[nr];
  T Hrow[NL], Hc[7], gl = tp.act ? xs[lane] : T(0);
#pragma unroll
  for (int k = 0; k < NL; k++) Hrow[k] = (k == tp.local && leaf) ? T(1) : T(0);
#pragma unroll
  for (int m = 0; m < 7; m++) Hc[m] = 0;
  T Se = (lane < 28 && em == en) ? T(1) : T(0);
  unsigned gact = 0;                                     // leaf blocks (fingers 0-2, palm 3) with at least one active row
  T gr = (lane < 28 && em == en) ? xs[cm] : T(0);
  // ---- H and g: one rank-one term per active row. A row's gripper part lies in ONE leaf block (a finger or the palm),
  // so every finger walks its own list of active rows: the trip count is the longest list, not the row count.
  {
    const int myg = tp.chain ? (lane - 1) / NL : (lane == Topo<N>::PALM ? 3 : -1);
    unsigned a0 = myg >= 0 ? (sc.tmask[myg][0] & m0) : 0u, a1 = myg >= 0 ? (sc.tmask[myg][1] & m1) : 0u;
    int trips = 0;
#pragma unroll
    for (int gq = 0; gq < 4; gq++) {
      const int c = __popc(sc.tmask[gq][0] & m0) + __popc(sc.tmask[gq][1] & m1);
      trips = c > trips ? c : trips; gact |= (c > 0) << gq;
    }
#pragma unroll 1
    for (int it = 0; it < trips; it++) {
      if (a0 | a1) {
        int r;
        if (a0) { r = __ffs(a0) - 1; a0 &= a0 - 1; } else { r = 32 + __ffs(a1) - 1; a1 &= a1 - 1; }
        const T* wr = ws.W[r];
        const T s = ws.rjv[r], sa = s * ws.rpos[r];      // 1 / R_r and aref_r / R_r (primal_prepare)
        const T wc = wr[lane];
        const T swc = s * wc;
#pragma unroll
        for (int k = 0; k < NL; k++) if (k <= tp.local) Hrow[k] += swc * wr[tp.fb + k];
        Hc[0] += swc * wr[0];
#pragma unroll
        for (int m = 1; m < 7; m++) Hc[m] += swc * wr[B + m - 1];
        gl += sa * wc;
      }
    }
  }
  // root block: every active row
  if (lane < 28) {
    unsigned a0 = m0, a1 = m1;
#pragma unroll 1
    while (a0 | a1) {
      int r;
      if (a0) { r = __ffs(a0) - 1; a0 &= a0 - 1; } else { r = 32 + __ffs(a1) - 1; a1 &= a1 - 1; }
      const T* wr = ws.W[r];
      const T s = ws.rjv[r], sa = s * ws.rpos[r];
      const T a = wr[cm], b = wr[cn];
      Se += s * a * b; if (em == en) gr += sa * a;
    }
  }
  // ---- leaves: LDL^T inside each finger chain (pivot row by shuffle), couplings and right-hand side ride along
#pragma unroll
  for (int k = 0; k < NL - 1; k++) {
    const int p = tp.fb + k;
    const T Dk = __shfl_sync(GMJ_FULL, Hrow[k], p);
    const T inv = rcpT<T>(tp.chain ? Dk : T(1));
    const bool below = tp.chain && tp.local > k;
    const T hk = Hrow[k];
    const T l = below ? hk * inv : T(0);
#pragma unroll
    for (int j = k + 1; j < NL; j++) {
      const T hjk = __shfl_sync(GMJ_FULL, hk, tp.fb + j);
      if (below && tp.local >= j) Hrow[j] -= l * hjk;
    }
#pragma unroll
    for (int m = 0; m < 7; m++) { const T cp = __shfl_sync(GMJ_FULL, Hc[m], p); if (below) Hc[m] -= l * cp; }
    { const T gp = __shfl_sync(GMJ_FULL, gl, p); if (below) gl -= l * gp; }
    if (below) Hrow[k] = l;
  }
  T D = T(1);
#pragma unroll
  for (int k = 0; k < NL; k++) if (tp.local == k) D = Hrow[k];
  if (!leaf) D = T(1);
  const T dinv = rcpT<T>(D), dis = sqrtT<T>(dinv);
  // factor columns + scaled couplings to shared memory
  if (tp.chain) { const int fi = (lane - 1) / NL;
#pragma unroll
    for (int k = 0; k < NL; k++) sc.Lc[fi][tp.local][k] = (k < tp.local) ? Hrow[k] : T(0); }
  if (tp.act) {
#pragma unroll
    for (int m = 0; m < 7; m++) sc.hc[lane][m] = leaf ? Hc[m] * dis : T(0);
    sc.hc[lane][7] = leaf ? gl * dis : T(0);
  }
  __syncwarp();
  // ---- Schur complement on the root: S -= sum_leaves u u^T, g_root -= sum_leaves u y'
  if (lane < 28) {
    // only the leaf blocks with an active row have couplings (the others hold zeros); odd lanes sum into a0 / b0,
    // even lanes into a1 / b1 whatever is skipped, so the sums do not depend on the skipping
    T a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
    for (int gq = 0; gq < 4; gq++) {
      if (!((gact >> gq) & 1)) continue;
#pragma unroll
      for (int k = 0; k < (gq < 3 ? NL : 1); k++) {
        const int c = 1 + gq * NL + k;
        const T pa = sc.hc[c][em] * sc.hc[c][en], pb = sc.hc[c][em] * sc.hc[c][7];
        if (c & 1) { a0 += pa; b0 += pb; } else { a1 += pa; b1 += pb; }
      }
    }
    Se -= a0 + a1;
    sc.root[lane] = Se;
    if (em == en) sc.root[28 + em] = gr - (b0 + b1);
  }
  __syncwarp();
  // ---- 7 x 7 Cholesky + solve, every lane redundantly (uniform straight-line code on broadcast loads). The indices are
  // compile-time constants (static_for), so the 28 factor entries stay in registers: with `#pragma unroll` loops the
  // array went to local memory (ncu r2: 4.9 % of all stall samples at N = 8 on its LDL / STL).
  {
    T Sm[28];
#pragma unroll
    for (int i = 0; i < 28; i++) Sm[i] = sc.root[i];
#pragma unroll
    for (int m = 0; m < 7; m++) xr[m] = sc.root[28 + m];
    static_for<0, 7>([&](auto jc) {
      constexpr int j = decltype(jc)::value, jj = j * (j + 1) / 2;
      T d = Sm[jj + j];
      static_for<0, j>([&](auto kc) { constexpr int k = decltype(kc)::value; d -= Sm[jj + k] * Sm[jj + k]; });
      const T inv = rsqrtT<T>(maxT<T>(d, T(1e-30)));
      Sm[jj + j] = inv;                                    // 1 / L_jj
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
  // ---- leaves back: x_c = (y_c - Hc_c . x_root) / D_c - sum_{below} L[i][c] x_i
  T x = gl;
#pragma unroll
  for (int m = 0; m < 7; m++) x -= Hc[m] * xr[m];
  x = leaf ? x * dinv : T(0);
  {
    const int fi = tp.chain ? (lane - 1) / NL : 0;
#pragma unroll
    for (int k = NL - 1; k >= 1; k--) {
      const T xk = __shfl_sync(GMJ_FULL, x, tp.fb + k);
      if (tp.chain && tp.local < k) x -= sc.Lc[fi][k][tp.local] * xk;
    }
  }
  if (lane == 0) x = xr[0];
  xl = x;
  if (tp.act) sc.x[lane] = x;
  if (lane < 6) sc.x[B + lane] = xr[1 + lane];
  __syncwarp();
}

#ifdef GMJ_SELF_COLLISION
// Fallback for row sets with a contact between two gripper bodies (finger <-> finger, hook <-> hook, segment <-> palm):
// such a row couples two leaf blocks, H is no longer block-arrow. H = I + sum_active w w^T / R is then formed densely
// ((B + 6)^2, packed lower triangle over the primal scratch), factored by a left-looking warp Cholesky (lane = row, rows
// lane and lane + 32) and solved for the one right-hand side. Rare, so runtime loops, no unrolling.
template <typename T, int N>
__device__ __noinline__ void primal_dense_solve(WarpShared<T, N>& ws, int nr, int lane, unsigned m0, unsigned m1, T* xv) {
  constexpr int NW = WarpShared<T, N>::NW;
  T* H = ws.A;
  const T* xs = ws.W[nr];
  constexpr int ne = NW * (NW + 1) / 2;
#pragma unroll 1
  for (int e0 = 0; e0 < ne; e0 += 32) {                                  // entries dealt 32 at a time, like form_A
    const int e = e0 + lane; const bool ok = e < ne; const int ee = ok ? e : 0;
    int i = (int)((sqrtf(8.0f * (float)ee + 1.0f) - 1.0f) * 0.5f);
    if ((i + 1) * (i + 2) / 2 <= ee) i++;
    if (i * (i + 1) / 2 > ee) i--;
    const int j = ee - i * (i + 1) / 2;
    T a = i == j ? T(1) : T(0);
    unsigned a0 = m0, a1 = m1;
#pragma unroll 1
    while (a0 | a1) {
      int r;
      if (a0) { r = __ffs(a0) - 1; a0 &= a0 - 1; } else { r = 32 + __ffs(a1) - 1; a1 &= a1 - 1; }
      a += ws.rjv[r] * ws.W[r][i] * ws.W[r][j];
    }
    if (ok) H[ee] = a;
  }
  T g[2];                                                                // right-hand side, rows lane and lane + 32
#pragma unroll
  for (int slot = 0; slot < 2; slot++) {
    const int c = lane + 32 * slot;
    T v = c < NW ? xs[c] : T(0);
    unsigned a0 = m0, a1 = m1;
#pragma unroll 1
    while ((a0 | a1) && c < NW) {
      int r;
      if (a0) { r = __ffs(a0) - 1; a0 &= a0 - 1; } else { r = 32 + __ffs(a1) - 1; a1 &= a1 - 1; }
      v += ws.rjv[r] * ws.rpos[r] * ws.W[r][c];
    }
    g[slot] = v;
  }
  __syncwarp();
#pragma unroll 1
  for (int j = 0; j < NW; j++) {                                         // column j of L, and the forward substitution with it
    T v[2];
#pragma unroll
    for (int slot = 0; slot < 2; slot++) {
      const int i = lane + 32 * slot;
      v[slot] = 0;
      if (i >= j && i < NW) {
        T a = H[TRI(i, j)];
        for (int k = 0; k < j; k++) a -= H[TRI(i, k)] * H[TRI(j, k)];
        v[slot] = a;
      }
    }
    const T d = __shfl_sync(GMJ_FULL, j < 32 ? v[0] : v[1], j & 31);
    const T inv = rsqrtT<T>(maxT<T>(d, T(1e-30)));
    const T gj = __shfl_sync(GMJ_FULL, j < 32 ? g[0] : g[1], j & 31) * inv;      // y_j
    __syncwarp();
#pragma unroll
    for (int slot = 0; slot < 2; slot++) {
      const int i = lane + 32 * slot;
      if (i == j) { H[TRI(j, j)] = inv; g[slot] = gj; }                   // 1 / L_jj on the diagonal
      else if (i > j && i < NW) { const T l = v[slot] * inv; H[TRI(i, j)] = l; g[slot] -= l * gj; }
    }
    __syncwarp();
  }
#pragma unroll 1
  for (int j = NW - 1; j >= 0; j--) {                                    // L^T x = y
    const T xj = __shfl_sync(GMJ_FULL, j < 32 ? g[0] : g[1], j & 31) * H[TRI(j, j)];
#pragma unroll
    for (int slot = 0; slot < 2; slot++) {
      const int i = lane + 32 * slot;
      if (i == j) g[slot] = xj;
      else if (i < j) g[slot] -= H[TRI(j, i)] * xj;
    }
  }
  if (lane < NW) xv[lane] = g[0];
  if (lane + 32 < NW) xv[lane + 32] = g[1];
  __syncwarp();
}
#endif

// Sparse row products: a whitened row is zero outside the base column, ONE leaf block (finger g: columns 1 + g NL ..,
// palm: column 3 N + 7; g = -1: none) and the six object columns. u.v or |u|^2 over exactly those columns.
template <typename T, int N, class F>
__device__ __forceinline__ T row_reduce(int g, F&& term) {
  constexpr int NL = N + 2, B = 3 * N + 8, PALM = 3 * N + 7;
  T a0 = term(0), a1 = 0;
#pragma unroll
  for (int k = 0; k < 6; k += 2) { a0 += term(B + k); a1 += term(B + k + 1); }
  if (g >= 0 && g < 3) {
    const int st = 1 + g * NL;
#pragma unroll
    for (int k = 0; k + 1 < NL; k += 2) { a0 += term(st + k); a1 += term(st + k + 1); }
    if (NL & 1) a0 += term(st + NL - 1);
  } else if (g == 3) a1 += term(PALM);
#ifdef GMJ_SELF_COLLISION
  else if (g == 4) {                         // a row between two gripper bodies: any gripper column
#pragma unroll 1
    for (int c = 1; c < B; c++) a1 += term(c);
  }
#endif
  return a0 + a1;
}

// Per row (lane = row, rows lane and lane + 32): reference acceleration, regulariser from the row norms
// (A_rr = |w_r|^2; contacts use the norm of the whitened NORMAL row, the mean of their first two pyramid rows),
// unconstrained residual for the cold start. Leaves aref in ws.rpos and 1 / R in ws.rjv.
template <typename T, int N>
__device__ __forceinline__ void primal_prepare(WarpShared<T, N>& ws, const T* solimp, T kk, T bb, int nr, int lane, bool warm,
                                               const T* fprev, bool (&act)[2], int (&type)[2], T (&Rr)[2], T (&aref)[2], int (&grp)[2]) {
  constexpr int NL = N + 2, PALM = 3 * N + 7;
#pragma unroll
  for (int slot = 0; slot < 2; slot++) {
    const int r = lane + 32 * slot;
    act[slot] = false; type[slot] = ROW_EQ; Rr[slot] = 1; aref[slot] = 0; grp[slot] = -1;
    if (slot == 1 && nr <= 32) break;
    if (r < nr) {
      const T pos = ws.rpos[r];
      const T imp = impedance<T>(solimp, pos);
      const T ar = -bb * ws.rjv[r] - kk * imp * pos;
      const int ty = ws.rtype[r];
      {                                                   // leaf block the row's gripper part lies in (welds / limits: their
        const int rl = ws.rlane[r];                       // joint lane; contacts: the gripper body's lane, put_row)
        const int bl = ty == ROW_CONTACT ? (rl <= -10 ? -(rl + 10) : 0) : rl;
        if (bl > 0) grp[slot] = bl == PALM ? 3 : (bl - 1) / NL;
#ifdef GMJ_SELF_COLLISION
        if (ty == ROW_CONTACT && rl == -3) grp[slot] = 4;
#endif
      }
      const T* wr = ws.W[r];
      const T* xs = ws.W[nr];
      const int g = grp[slot];
      const T jx = warm ? T(0) : row_reduce<T, N>(g, [&](int c) { return wr[c] * xs[c]; });
      if (ty == ROW_CONTACT) {
        const int rf = ws.rfirst[r];
        const T* wa = ws.W[rf]; const T* wb = ws.W[rf + 1];
        const T mu = ws.rmu[r];
        const T Ann = T(0.25) * row_reduce<T, N>(g, [&](int c) { const T u = wa[c] + wb[c]; return u * u; });
        Rr[slot] = 2 * mu * mu * (1 - imp) * rcpT<T>(imp) * (1 + mu * mu) * Ann;
      } else {
        Rr[slot] = (1 - imp) * rcpT<T>(imp) * row_reduce<T, N>(g, [&](int c) { return wr[c] * wr[c]; });
      }
      Rr[slot] = maxT<T>(Rr[slot], T(1e-15));
      aref[slot] = ar; type[slot] = ty;
      act[slot] = ty == ROW_EQ || (warm ? fprev[r] > 0 : jx - ar < 0);
    }
  }
  __syncwarp();
#pragma unroll
  for (int slot = 0; slot < 2; slot++) {
    const int r = lane + 32 * slot;
    if (r < nr) { ws.rpos[r] = aref[slot]; ws.rjv[r] = rcpT<T>(Rr[slot]); }
  }
  __syncwarp();
}

// The whole solve. Out: force of this lane's rows in f[2], x - x_s of this lane's gripper coordinate in dx and of the
// object coordinates in dxo (lanes 0-5). Returns the number of pivoting rounds.
template <typename T, int N>
__device__ int primal_solve(WarpShared<T, N>& ws, const Topo<N>& tp, const T* solimp, T kk, T bb, int nr, int lane, bool warm,
                            const T* fprev, T (&f)[2], T& dx, T& dxo) {
  constexpr int B = 3 * N + 8;
  static_assert(sizeof(PrimalScratch<T, N>) <= sizeof(ws.A), "primal scratch must fit the packed-A storage");
  PrimalScratch<T, N>& sc = *reinterpret_cast<PrimalScratch<T, N>*>(ws.A);
  const T eps = sizeof(T) == 8 ? T(1e-12) : T(2e-6);
  const int maxit = 40;
  bool act[2]; int type[2]; T Rr[2], aref[2];
  int grp[2];
  primal_prepare<T, N>(ws, solimp, kk, bb, nr, lane, warm, fprev, act, type, Rr, aref, grp);
#pragma unroll
  for (int gq = 0; gq < 4; gq++) {
    const unsigned t0 = __ballot_sync(GMJ_FULL, grp[0] == gq), t1 = __ballot_sync(GMJ_FULL, grp[1] == gq);
    if (lane == 0) { sc.tmask[gq][0] = t0; sc.tmask[gq][1] = t1; }
  }
  __syncwarp();
  const int r0 = lane, r1 = lane + 32;
  const bool two = nr > 32;
  const T* xs = ws.W[nr];
  // scale of the right-hand side for the slack tolerance (the dual solver's |b| = |w_r . x_s - aref_r|): aref dominates
  const T bscale = warp_max(maxT<T>(r0 < nr ? absT<T>(aref[0]) : T(0), (two && r1 < nr) ? absT<T>(aref[1]) : T(0)));
  int best = 1 << 30, patience = 3, it = 0;
  T xl = 0, xr[7];
  const T* xv = sc.x;                        // solution vector the residuals read
#ifdef GMJ_SELF_COLLISION
  // a row between two gripper bodies anywhere in the set: every round of this solve goes through the dense fallback (its
  // matrix lies over the primal scratch; the solution vector sits behind it)
  const bool dense = __any_sync(GMJ_FULL, grp[0] == 4 || grp[1] == 4);
  T* const xdense = ws.A + WarpShared<T, N>::NW * (WarpShared<T, N>::NW + 1) / 2;
  if (dense) xv = xdense;
#endif
  T f0 = 0, f1 = 0;
  for (;; ) {
    it++;
    const unsigned m0 = __ballot_sync(GMJ_FULL, act[0]), m1 = __ballot_sync(GMJ_FULL, act[1]);
#ifdef GMJ_SELF_COLLISION
    if (dense) {
      primal_dense_solve<T, N>(ws, nr, lane, m0, m1, xdense);
      xl = tp.act ? xv[lane] : T(0);
      xr[0] = xv[0];
#pragma unroll
      for (int k = 0; k < 6; k++) xr[1 + k] = xv[B + k];
    } else
#endif
    primal_factor_solve<T, N>(ws, sc, tp, nr, lane, m0, m1, xl, xr);
    // ---- residuals jar_r = w_r . x - aref_r; force of an active row -jar / R, slack of an inactive row jar
    T j0 = 0, j1 = 0;
    if (r0 < nr) { const T* wr = ws.W[r0]; j0 = row_reduce<T, N>(grp[0], [&](int c) { return wr[c] * xv[c]; }) - aref[0]; }
    if (two && r1 < nr) { const T* wr = ws.W[r1]; j1 = row_reduce<T, N>(grp[1], [&](int c) { return wr[c] * xv[c]; }) - aref[1]; }
    f0 = act[0] ? -j0 * rcpT<T>(Rr[0]) : T(0); f1 = act[1] ? -j1 * rcpT<T>(Rr[1]) : T(0);
    const T fscale = warp_max(maxT<T>(absT<T>(f0), absT<T>(f1)));
    T epsr = eps;
    for (int k = GMJ_LCP_GROW_FROM; k < it && epsr < T(1e-2); k++) epsr *= GMJ_LCP_GROW;
    const T tolf = epsr * (1 + fscale), tolw = epsr * (1 + bscale);
    const bool bad0 = r0 < nr && type[0] != ROW_EQ && (act[0] ? f0 < -tolf : j0 < -tolw);
    const bool bad1 = two && r1 < nr && type[1] != ROW_EQ && (act[1] ? f1 < -tolf : j1 < -tolw);
    const unsigned b0 = __ballot_sync(GMJ_FULL, bad0), b1 = __ballot_sync(GMJ_FULL, bad1);
    const int ninf = __popc(b0) + __popc(b1);
    if (ninf == 0 || it >= maxit) { if (it > GMJ_LCP_GROW_FROM) it |= 1 << 16; if (ninf != 0) it |= 1 << 17; break; }
    bool block = true;
    if (ninf < best) { best = ninf; patience = 3; }
    else if (patience > 0) patience--;
    else block = false;
    if (block) { if (bad0) act[0] = !act[0]; if (bad1) act[1] = !act[1]; }
    else {                                     // Murty: flip only the infeasible row of largest index
      const int top = b1 ? 32 + (31 - __clz(b1)) : (31 - __clz(b0));
      if (r0 == top) act[0] = !act[0];
      if (r1 == top) act[1] = !act[1];
    }
    __syncwarp();
  }
  f[0] = (type[0] != ROW_EQ && f0 < 0) ? T(0) : f0;
  f[1] = (type[1] != ROW_EQ && f1 < 0) ? T(0) : f1;
  dx = tp.act ? xl - xs[lane] : T(0);
  dxo = lane < 6 ? xr[1 + lane] - xs[B + lane] : T(0);
  return it;
}
