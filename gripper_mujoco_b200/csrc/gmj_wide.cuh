// gmj_wide.cuh -- the physics substep for long fingers (B = 3N + 8 > 32 dofs, N <= 28; configs[4] runs N = 24).
//
// One environment per 96-thread block, one warp per finger: lane l < N+2 of warp f owns dof l of finger f
// (prismatic, revolute, N segment hinges), lane 30 of warp 0 owns the palm slide and lane 31 of EVERY warp
// holds a replica of the base slide, so all tree recursions stay warp-local (shuffles along the chain,
// base by broadcast from lane 31) exactly as in the lane-per-dof kernel (gmj_physics.cuh). The only
// cross-warp traffic is the base row of the block-arrow mass matrix: composite inertia / force gather, the
// base pivot of the factorisations and the base entry of every triangular solve, each ONE shared-memory
// exchange + block barrier (xsum). The (N+2)x(N+2) finger blocks of both factors live in shared memory
// (rows of 26 do not fit the register file next to everything else); the dual LCP (<= 48 rows) is solved
// by warp 0 with the same warp Cholesky as the narrow kernel. Same pipeline and semantics as
// physics_substep; EnvState keeps qpos/qvel in the reference dof order (object at [88, 95)).
#pragma once
#include "gmj_physics.cuh"

#define GMJ_WQC 96                   // qpos/qvel capacity of the wide state (B <= 88, object at 88)
#define GMJ_WNT 96                   // threads per environment
#define GMJ_XCH 20                   // values per group in one cross-warp exchange

template <int N> struct TopoW {
  static constexpr int NL = N + 2, B = 3 * N + 8, PALM_L = 30, BASE_L = 31;
  static constexpr int PALM_SLOT = PALM_L, BASE_SLOT = BASE_L, OBJ = 96;
  int w, lane, slot, local, parent, dof, col;
  bool chain, palm, base, act, own, hinge;
  __device__ __forceinline__ TopoW(int tid) {
    w = tid >> 5; lane = tid & 31; slot = tid;
    chain = lane < NL; palm = (w == 0 && lane == PALM_L); base = lane == BASE_L;
    act = chain || palm || base; own = chain || palm || (base && w == 0);
    hinge = chain && lane > 0;
    local = chain ? lane : 0;
    parent = (chain && lane > 0) ? lane - 1 : BASE_L;
    dof = base ? 0 : (chain ? 1 + w * NL + lane : (palm ? B - 1 : 0));
    col = chain ? w * NL + lane : (palm ? 3 * NL : 3 * NL + 1);      // column of this dof in a whitened row
  }
};

template <typename T, int N>
struct WideShared : WarpShared<T, N, GMJ_WQC> {
  T L1[3][N + 2][N + 3];            // first factor (M), finger blocks; WarpShared::L2 holds the second (M + h D)
  T sj[GMJ_WNT][6];                 // joint motion subspaces (mass-matrix rows read them by broadcast)
  T jaw[GMJ_WNT][3], jq[GMJ_WNT], jqd[GMJ_WNT], jlo[GMJ_WNT], jhi[GMJ_WNT];   // per-slot joint data for the row builders
  T xch[2][4][GMJ_XCH];             // cross-warp exchange (double buffered): groups = finger 0-2, palm
  int cnt[2][4];                    // contacts per group: gripper-ground, gripper-object
  unsigned limmask[3];              // joint-limit violations per warp
  int ebad, extra;
  int grp;                          // which env group of the block this is (named barrier 1 + grp)
  int mprw[3];                      // MPR passes per warp in the last collision phase
  alignas(32) T colb[3][2][32];     // per-warp column broadcast of the factorisation (double buffered)
};

// Environments per block. FP32: as many as shared memory holds (4 at N = 24), run in LOCKSTEP like the lane-per-dof
// kernel: the substep is ~12 k straight-line SASS instructions per warp, independent 96-thread blocks each streamed
// their own copy (ncu r1: stall_no_inst 41 %); groups that share a block share the fetch stream. The three warps of
// one env meet on named barrier 1 + grp (env_sync), the whole block on barrier 0 at the phase boundaries (wide_phase).
template <typename T, int N> struct WideCfg {
  static constexpr int FIT = (int)((227 * 1024 - 64) / sizeof(WideShared<T, N>));
  static constexpr int EPB = sizeof(T) == 4 ? (FIT >= 4 ? 4 : (FIT >= 1 ? FIT : 1)) : 1;
};
#define GMJ_WNPHASE 6                // block-wide barriers inside one wide substep
template <typename T, int N> __device__ __forceinline__ void env_sync(const WideShared<T, N>& ws) {
  if (WideCfg<T, N>::EPB == 1) __syncthreads();
  else asm volatile("bar.sync %0, %1;" :: "r"(ws.grp + 1), "n"(GMJ_WNT) : "memory");
}
template <typename T, int N> __device__ __forceinline__ void wide_phase() {
  if (WideCfg<T, N>::EPB > 1) __syncthreads();
}
template <typename T, int N> __device__ __forceinline__ void wide_phase_idle() {
#pragma unroll
  for (int k = 0; k < GMJ_WNPHASE; k++) wide_phase<T, N>();
}

// sum K per-thread partials over the whole environment (3 warps); every thread gets the same totals
template <typename T, int N, int K>
__device__ __forceinline__ void xsum(WideShared<T, N>& ws, int& xn, int w, int lane, T (&v)[K]) {
  static_assert(K <= GMJ_XCH, "exchange too wide");
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int m = 0; m < K; m++) v[m] += __shfl_xor_sync(GMJ_FULL, v[m], o);
  }
  T (*buf)[GMJ_XCH] = ws.xch[xn & 1];
  if (lane == 0) {
#pragma unroll
    for (int m = 0; m < K; m++) buf[w][m] = v[m];
  }
  env_sync<T, N>(ws);
#pragma unroll
  for (int m = 0; m < K; m++) v[m] = buf[0][m] + buf[1][m] + buf[2][m];
  xn++;
}

// x <- L^-1 x for K right-hand sides, lane-distributed; P extra per-thread partials ride along in the same
// exchange and come back summed over the environment. Lw = this warp's finger block of the factor.
template <typename T, int N, int K, int P>
__device__ __forceinline__ void fwd_w(WideShared<T, N>& ws, int& xn, const TopoW<N>& tp, const T* Lw, T lb, T (&x)[K], T* pay) {
  constexpr int NL = N + 2, NLP = N + 3;
  bool nz = false;
#pragma unroll
  for (int m = 0; m < K; m++) nz |= (tp.chain && x[m] != T(0));
  if (__any_sync(GMJ_FULL, nz)) {
#pragma unroll 2
    for (int k = 0; k < NL - 1; k++) {
      const T lik = (tp.chain && tp.local > k) ? Lw[tp.local * NLP + k] : T(0);
#pragma unroll
      for (int m = 0; m < K; m++) { T xk = __shfl_sync(GMJ_FULL, x[m], k); x[m] -= lik * xk; }
    }
  }
  T sv[K + P];
#pragma unroll
  for (int m = 0; m < K; m++) sv[m] = (tp.chain || tp.palm) ? lb * x[m] : T(0);
#pragma unroll
  for (int m = 0; m < P; m++) sv[K + m] = pay[m];
  xsum<T, N, K + P>(ws, xn, tp.w, tp.lane, sv);
  if (tp.base) {
#pragma unroll
    for (int m = 0; m < K; m++) x[m] -= sv[m];
  }
#pragma unroll
  for (int m = 0; m < P; m++) pay[m] = sv[K + m];
}

// LDL^T of one finger block + the base arrow. The row of each lane stays in registers (fully unrolled right-looking
// elimination); column k is published once per step through a per-warp shared buffer and read back by
// broadcast, four entries per load.
// In: row[k] = M[lane][k] (k <= lane), Mb = M[lane][base], dself = own diagonal of the palm / base lanes.
// Out (shared): Lw[i][k<i] = L, Lw[i][i] = pivot; D = pivot of this lane, lb = L[base][lane].
template <typename T> struct alignas(4 * sizeof(T)) Quad { T v[4]; };
template <typename T, int N>
__device__ __forceinline__ void ldl_factor_w(WideShared<T, N>& ws, int& xn, const TopoW<N>& tp, const T (&rin)[N + 2], T* Lw, T Mb, T dself, T& D, T& lb) {
  constexpr int NL = N + 2, NLP = N + 3;
  T row[NL];
#pragma unroll
  for (int k = 0; k < NL; k++) row[k] = rin[k];
#pragma unroll
  for (int k = 0; k < NL - 1; k++) {
    T* col = ws.colb[tp.w][k & 1];
    col[tp.lane] = row[k];
    __syncwarp();
    const T inv = rcpT<T>(col[k]);
    const T Mbk = __shfl_sync(GMJ_FULL, Mb, k);
    const bool below = tp.chain && tp.local > k;
    const T mk_pre = below ? row[k] : T(0);
    const T l = mk_pre * inv;                       // 0 on the lanes that are not below the pivot: their rows stay
#pragma unroll
    for (int j4 = (k + 1) / 4; j4 < (NL + 3) / 4; j4++) {
      const Quad<T> c = reinterpret_cast<const Quad<T>*>(col)[j4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int j = 4 * j4 + u;
        if (j > k && j < NL) row[j] -= l * c.v[u];
      }
    }
    Mb -= (Mbk * inv) * mk_pre;
    if (below) row[k] = l;
  }
  T d = dself;
#pragma unroll
  for (int k = 0; k < NL; k++) {
    if (tp.chain && k <= tp.local) Lw[tp.local * NLP + k] = row[k];
    if (tp.chain && k == tp.local) d = row[k];
  }
  const T dinv = rcpT<T>((tp.chain || tp.palm) ? d : T(1));
  T s[1] = { (tp.chain || tp.palm) ? Mb * Mb * dinv : T(0) };
  xsum<T, N, 1>(ws, xn, tp.w, tp.lane, s);
  if (tp.base) { D = dself - s[0]; lb = 0; }
  else { D = d; lb = (tp.chain || tp.palm) ? Mb * dinv : T(0); }
}

// point Jacobian entry of this thread's dof for a point p on the gripper body in slot `bs`, along d
template <typename T, int N>
__device__ __forceinline__ T jac_entry_w(const TopoW<N>& tp, int bs, V3<T> p, V3<T> d, V3<T> aw, V3<T> anchor) {
  if (bs < 0 || bs >= TopoW<N>::OBJ || !tp.act) return T(0);
  bool anc;
  if (tp.base) anc = true;
  else if (tp.palm) anc = (bs == TopoW<N>::PALM_SLOT);
  else anc = ((bs >> 5) == tp.w) && ((bs & 31) < TopoW<N>::NL) && ((bs & 31) >= tp.lane);
  if (!anc) return T(0);
  return tp.hinge ? dot(d, cross(aw, p - anchor)) : dot(d, aw);
}

template <typename T, int N>
__device__ __forceinline__ void put_row_w(WideShared<T, N>& ws, const TopoW<N>& tp, const T* rsqo, int r, T xw,
                                          const T* Jo, int type, T pos, T jv, T mu, int first, int jdof) {
  if (tp.own) ws.W[r][tp.col] = xw;
  if (tp.slot == 0) {
#pragma unroll
    for (int k = 0; k < 6; k++) ws.W[r][3 * N + 8 + k] = Jo[k] * rsqo[k];
    ws.rtype[r] = type; ws.rpos[r] = pos; ws.rjv[r] = jv; ws.rmu[r] = mu; ws.rfirst[r] = first; ws.rlane[r] = jdof;
  }
}

// does this thread hold (a replica of) the dof in slot js? The base slide is replicated in lane 31 of every warp.
template <int N> __device__ __forceinline__ bool holds_slot(const TopoW<N>& tp, int js) {
  return js == TopoW<N>::BASE_SLOT ? tp.base : tp.slot == js;
}
// slot of weld id's joint (0-2 prismatic, 3-5 revolute, 6 palm)
__device__ __forceinline__ int weld_slot_w(int id) { return id < 3 ? 32 * id : (id < 6 ? 32 * (id - 3) + 1 : 30); }

#include "gmj_primal_wide.cuh"

template <typename T, int N>
__device__ void physics_substep_wide(WideShared<T, N>& ws, const DevModel<T, GMJ_WNT>& M, const DevSettings& S, int tid, double dtd, T tip_force) {
  using TW = TopoW<N>;
  constexpr int NL = N + 2, NLP = N + 3, B = 3 * N + 8, OQ = GMJ_WQC - 8;
  const T dt = (T)dtd;
  const TW tp(tid);
  const int lane = tp.lane, w = tp.w;
  EnvState<T, GMJ_WQC>& e = ws.e;
  int xn = 0;
  if (e.unstable) { wide_phase_idle<T, N>(); env_sync<T, N>(ws); if (tid == 0) e.time += dtd; return; }
  const T gz = -M.zref;
  T* L1w = &ws.L1[w][0][0];
  T* L2w = &ws.L2[w][0][0];

  // ------------------------------------------------------------------ kinematics (parallel prefix of rigid transforms)
  T q = tp.own || tp.base ? e.qpos[tp.dof] : T(0), qd = tp.own || tp.base ? e.qvel[tp.dof] : T(0);
  if (tp.own) ws.qfk[tp.dof] = q;
  if (tid == 0) ws.ebad = 0;
  M3<T> Rl; V3<T> pl = mk<T>(M.f[MF_pos0][tid], M.f[MF_pos1][tid], M.f[MF_pos2][tid]);
  const V3<T> axl = mk<T>(M.f[MF_ax0][tid], M.f[MF_ax1][tid], M.f[MF_ax2][tid]);
#pragma unroll
  for (int k = 0; k < 9; k++) Rl.m[k] = M.f[MF_R0 + k][tid];
  // local transform of the body relative to its parent, joint motion included
  M3<T> R = tp.hinge ? mul(Rl, rot_axis(axl, q)) : Rl;
  V3<T> p = tp.hinge ? pl : pl + mul(Rl, axl) * q;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) {
    M3<T> Ru; V3<T> pu;
#pragma unroll
    for (int k = 0; k < 9; k++) Ru.m[k] = __shfl_up_sync(GMJ_FULL, R.m[k], off);
    pu = shflup3(p, off);
    if (tp.chain && tp.local >= off) { p = pu + mul(Ru, p); R = mul(Ru, R); }
  }
  {
    M3<T> Rb = shflM(R, TW::BASE_L); V3<T> pb = shfl3(p, TW::BASE_L);
    if (tp.chain || tp.palm) { p = pb + mul(Rb, p); R = mul(Rb, R); }
  }
  V3<T> aw, anchor;
  {
    // joint frame: parent's world pose composed with the fixed offset
    M3<T> Rp; V3<T> pp;
#pragma unroll
    for (int k = 0; k < 9; k++) Rp.m[k] = __shfl_sync(GMJ_FULL, R.m[k], tp.parent);
    pp = shfl3(p, tp.parent);
    if (tp.base) { aw = mul(Rl, axl); anchor = pl; }
    else { M3<T> R0 = mul(Rp, Rl); aw = mul(R0, axl); anchor = pp + mul(Rp, pl); }
    if (!tp.act) { aw = mk<T>(0, 0, 1); anchor = mk<T>(0, 0, 0); }
  }
  if (tp.chain && tp.local == 1) {
#pragma unroll
    for (int k = 0; k < 9; k++) ws.ax.fingerR[w][k] = R.m[k]; }
  if (tp.palm) {
#pragma unroll
    for (int k = 0; k < 9; k++) ws.ax.palmR[k] = R.m[k]; }
  ws.jaw[tid][0] = aw.x; ws.jaw[tid][1] = aw.y; ws.jaw[tid][2] = aw.z; ws.jq[tid] = q; ws.jqd[tid] = qd;
  const T lim_lo = q - M.f[MF_lo][tid], lim_hi = M.f[MF_hi][tid] - q;
  ws.jlo[tid] = lim_lo; ws.jhi[tid] = lim_hi;
  {
    unsigned ml = __ballot_sync(GMJ_FULL, tp.own && (lim_lo < 0 || lim_hi < 0));
    if (lane == 0) ws.limmask[w] = ml;
  }
  // object (replicated in every thread)
  V3<T> po = mk<T>(e.qpos[OQ], e.qpos[OQ + 1], e.qpos[OQ + 2] - M.zref);
  M3<T> Ro;
  { T a = e.qpos[OQ + 3], b = e.qpos[OQ + 4], c = e.qpos[OQ + 5], d = e.qpos[OQ + 6];
    T l = T(1) / sqrtT<T>(a * a + b * b + c * c + d * d);
    Ro = quat2mat<T>(a * l, b * l, c * l, d * l); }
  V3<T> vo = mk<T>(e.qvel[OQ], e.qvel[OQ + 1], e.qvel[OQ + 2]), wo = mk<T>(e.qvel[OQ + 3], e.qvel[OQ + 4], e.qvel[OQ + 5]);
  const bool has_obj = e.has_object != 0;
  const T om = e.obj[9];
  const V3<T> oI = mk<T>(e.obj[10], e.obj[11], e.obj[12]);

  wide_phase<T, N>();
  // ------------------------------------------------------------------ spatial quantities
  T mass = tp.act ? M.f[MF_mass][tid] : T(0);
  V3<T> c = p + mul(R, mk<T>(M.f[MF_ip0][tid], M.f[MF_ip1][tid], M.f[MF_ip2][tid]));
  SpI<T> I;
  {
    T ib[6] = { M.f[MF_ixx][tid], M.f[MF_iyy][tid], M.f[MF_izz][tid], M.f[MF_ixy][tid], M.f[MF_ixz][tid], M.f[MF_iyz][tid] };
    M3<T> Ib; Ib.m[0] = ib[0]; Ib.m[4] = ib[1]; Ib.m[8] = ib[2]; Ib.m[1] = Ib.m[3] = ib[3]; Ib.m[2] = Ib.m[6] = ib[4]; Ib.m[5] = Ib.m[7] = ib[5];
    M3<T> t = mul(R, Ib);
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
  ws.sj[tid][0] = Sj.w.x; ws.sj[tid][1] = Sj.w.y; ws.sj[tid][2] = Sj.w.z; ws.sj[tid][3] = Sj.v.x; ws.sj[tid][4] = Sj.v.y; ws.sj[tid][5] = Sj.v.z;
  // velocities and bias accelerations: inclusive prefix along the chain, then add the base
  Sp<T> v = Sj * qd;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) { Sp<T> t = spup(v, off); if (tp.chain && tp.local >= off) v = v + t; }
  Sp<T> vbase = spshfl(v, TW::BASE_L);
  if (tp.chain || tp.palm) v = v + vbase;
  Sp<T> a = crossm(v, Sj) * qd;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) { Sp<T> t = spup(a, off); if (tp.chain && tp.local >= off) a = a + t; }
  Sp<T> abase = spshfl(a, TW::BASE_L);
  if (tp.chain || tp.palm) a = a + abase;
  a.v.z += M.gravity;
  // forces and composite inertias: inclusive suffix along the chain, then gather into the base (exchange)
  Sp<T> f = spimul(I, a) + crossf(v, spimul(I, v));
  SpI<T> Ic = I;
#pragma unroll
  for (int off = 1; off < NL; off <<= 1) {
    Sp<T> tf = spdn(f, off); SpI<T> ti = spishfl_dn(Ic, off);
    if (tp.chain && tp.local + off < NL) { f = f + tf; Ic = spiadd(Ic, ti); }
  }
  // calibration tip load on the last segment of each finger (src/myfunctions.cpp:1642-1727)
  T tau_tip = 0, tb_tip = 0;
  if (tip_force != T(0)) {
    V3<T> ct = shfl3(c, NL - 1);
    V3<T> fv = mk<T>(-M.tip_mat[w][2] * tip_force, -M.tip_mat[w][5] * tip_force, -M.tip_mat[w][8] * tip_force);
    Sp<T> fs; fs.w = cross(ct, fv); fs.v = fv;
    if (tp.chain) tau_tip = spdot(Sj, fs);
    if (tp.base) tb_tip = spdot(Sj, fs);
  }
  {
    T (*buf)[GMJ_XCH] = ws.xch[xn & 1];
    const int grp = tp.palm ? 3 : w;
    if ((tp.chain && tp.local == 0) || tp.palm) {
      T* o = buf[grp];
      o[0] = f.w.x; o[1] = f.w.y; o[2] = f.w.z; o[3] = f.v.x; o[4] = f.v.y; o[5] = f.v.z;
      o[6] = Ic.m; o[7] = Ic.h.x; o[8] = Ic.h.y; o[9] = Ic.h.z;
#pragma unroll
      for (int k = 0; k < 6; k++) o[10 + k] = Ic.I[k];
    }
    if (tp.base) buf[w][16] = tb_tip;
    env_sync<T, N>(ws);
    if (tp.base) {
#pragma unroll 1
      for (int g = 0; g < 4; g++) {
        const T* o = buf[g];
        f.w = f.w + mk<T>(o[0], o[1], o[2]); f.v = f.v + mk<T>(o[3], o[4], o[5]);
        Ic.m += o[6]; Ic.h = Ic.h + mk<T>(o[7], o[8], o[9]);
#pragma unroll
        for (int k = 0; k < 6; k++) Ic.I[k] += o[10 + k];
      }
      tau_tip = buf[0][16] + buf[1][16] + buf[2][16];
    }
    xn++;
  }
  T bias = spdot(Sj, f);
  // mass matrix rows into the factor storage: M[lane][k] (k <= local), coupling to the base in Mb
  Sp<T> F = spimul(Ic, Sj);
  const T damp = tp.act ? M.f[MF_damp][tid] : T(0), stiff = tp.act ? M.f[MF_stiff][tid] : T(0);
  T Mrow[NL];
#pragma unroll
  for (int k = 0; k < NL; k++) {
    const T* sk = ws.sj[32 * w + k];
    const T mv = sk[0] * F.w.x + sk[1] * F.w.y + sk[2] * F.w.z + sk[3] * F.v.x + sk[4] * F.v.y + sk[5] * F.v.z;
    Mrow[k] = (tp.chain && k <= tp.local) ? mv : T(0);
  }
  Sp<T> Sb = spshfl(Sj, TW::BASE_L);
  const T Mb = (tp.chain || tp.palm) ? spdot(Sb, F) : T(0);
  const T Mself = spdot(Sj, F);             // own diagonal (used by the palm and base lanes)

  // ------------------------------------------------------------------ control + smooth forces
  T ctrl = 0;
  {
    int tsel = tp.act ? (int)M.f[MF_tsel][tid] : 0;
    T tgt = 0;
    if (tsel == 1) tgt = (T)e.base_z; else if (tsel == 2) tgt = (T)e.tnext.x; else if (tsel == 3) tgt = (T)e.tnext.th; else if (tsel == 4) tgt = (T)e.tnext.z;
    if (tsel != 0) ctrl = -((q - tgt) * M.f[MF_kp][tid] + qd * M.f[MF_kd][tid]);
  }
  T tau = -bias + ctrl - stiff * q - damp * qd + tau_tip;
  T wrist = ctrl - damp * qd - stiff * q;                 // base: force sensor z (quirk Q3)
  T tauo[6], rsqo[6], Mo[6];
  {
    V3<T> Iw = mk<T>(oI.x * wo.x, oI.y * wo.y, oI.z * wo.z);
    V3<T> gyr = cross(wo, Iw);
    tauo[0] = 0; tauo[1] = 0; tauo[2] = has_obj ? -om * M.gravity : T(0);
    tauo[3] = has_obj ? -gyr.x : T(0); tauo[4] = has_obj ? -gyr.y : T(0); tauo[5] = has_obj ? -gyr.z : T(0);
    Mo[0] = Mo[1] = Mo[2] = has_obj ? om : T(1);
    Mo[3] = has_obj ? oI.x : T(1); Mo[4] = has_obj ? oI.y : T(1); Mo[5] = has_obj ? oI.z : T(1);
#pragma unroll
    for (int k = 0; k < 6; k++) rsqo[k] = T(1) / sqrtT<T>(Mo[k]);
  }

  wide_phase<T, N>();
  // ------------------------------------------------------------------ factorisations
  T D, lb, D2, lb2;
  ldl_factor_w<T, N>(ws, xn, tp, Mrow, L1w, Mb, Mself, D, lb);
#pragma unroll
  for (int k = 0; k < NL; k++) if (tp.chain && k == tp.local) Mrow[k] += dt * damp;
  ldl_factor_w<T, N>(ws, xn, tp, Mrow, L2w, Mb, Mself + dt * damp, D2, lb2);
  const T rsq = tp.act ? rsqrtT<T>(tp.act ? D : T(1)) : T(0);
  int bad = !(D > 0) && tp.act;
  // whitened smooth force w_s = D^-1/2 L^-1 tau
  T wsm;
  { T x1[1] = { tp.act ? tau : T(0) }; fwd_w<T, N, 1, 0>(ws, xn, tp, L1w, lb, x1, nullptr); wsm = x1[0] * rsq; }

  wide_phase<T, N>();
  // ------------------------------------------------------------------ collision
  Shape<T> so;
  so.type = (int)e.obj[0]; so.size = mk<T>(e.obj[1], e.obj[2], e.obj[3]); so.round = e.obj[4];
  so.pos = po; so.R = mul(Ro, quat2mat<T>(e.obj[5], e.obj[6], e.obj[7], e.obj[8]));
  const T ofric = e.obj[13];
  auto add_contact = [&](V3<T> pos, V3<T> n, T dist, T mu, int b1, int c1, int b2, int c2, int slot) {
    DevContact<T>& cn = ws.ax.con[slot];
    cn.pos[0] = pos.x; cn.pos[1] = pos.y; cn.pos[2] = pos.z;
    make_frame<T>(n, cn.frame);
    cn.dist = dist; cn.mu = mu; cn.body1 = b1; cn.body2 = b2; cn.cls1 = c1; cn.cls2 = c2;
    cn.force[0] = cn.force[1] = cn.force[2] = 0; cn.row = -1;
  };
  // (1) object vs ground: lane i of every warp forms candidate i (each warp needs the count), warp 0 writes the up to four
  // deepest penetrating ones, deepest first, ties by candidate index
  int n1 = 0;
  if (has_obj) {
    V3<T> pc; T dc;
    plane_candidate_lane<T>(so, gz, lane, pc, dc);
    int rank = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) { const T dj = __shfl_sync(GMJ_FULL, dc, j); rank += (dj < 0 && (dj < dc || (dj == dc && j < lane))) ? 1 : 0; }
    const bool pen = lane < 8 && dc < 0;
    n1 = __popc(__ballot_sync(GMJ_FULL, pen));
    if (n1 > 4) n1 = 4;
    if (n1 > GMJ_MAXCON) n1 = GMJ_MAXCON;
    if (w == 0 && pen && rank < n1) {
      V3<T> pp = pc; pp.z -= T(0.5) * dc;
      add_contact(pp, mk<T>(0, 0, 1), dc, maxT<T>(M.ground_friction, ofric), -1, GMJ_CLS_GROUND, TW::OBJ, GMJ_CLS_OBJECT, rank);
    }
  }
  // per-thread geoms: primary (segment plate / palm cylinder) and the hook plate on the tip lanes
  Shape<T> sg, sh;
  const bool geo = tp.chain || tp.palm;
  const int gtype = geo ? (int)M.f[MF_gtype][tid] : -1, htype = geo ? (int)M.f[MF_htype][tid] : -1;
  const int gcls = (int)M.f[MF_gcls][tid];
  const T gfric = M.f[MF_gfric][tid];
  {
    M3<T> Rg, Rh;
#pragma unroll
    for (int k = 0; k < 9; k++) { Rg.m[k] = M.f[MF_gR0 + k][tid]; Rh.m[k] = M.f[MF_hR0 + k][tid]; }
    sg.type = gtype; sg.round = 0; sg.size = mk<T>(M.f[MF_gs0][tid], M.f[MF_gs1][tid], M.f[MF_gs2][tid]);
    sg.pos = p + mul(R, mk<T>(M.f[MF_gp0][tid], M.f[MF_gp1][tid], M.f[MF_gp2][tid])); sg.R = mul(R, Rg);
    sh.type = htype; sh.round = 0; sh.size = mk<T>(M.f[MF_hs0][tid], M.f[MF_hs1][tid], M.f[MF_hs2][tid]);
    sh.pos = p + mul(R, mk<T>(M.f[MF_hp0][tid], M.f[MF_hp1][tid], M.f[MF_hp2][tid])); sh.R = mul(R, Rh);
  }
  // candidates of this thread: (2) its plates against the ground (up to 4 + 4), (3) its geoms against the object (up to 2)
  int n2g = 0, n2h = 0, hitg = 0, hith = 0; T depg = 0, deph = 0; V3<T> ng, nh, pg, ph;
  {
    V3<T> pts[8]; T dd[8];
    if (gtype == GMJ_GEOM_BOX) { int n = plane_candidates<T>(sg, gz, pts, dd); for (int i = 0; i < n; i++) n2g += dd[i] < 0; }
    if (htype == GMJ_GEOM_BOX) { int n = plane_candidates<T>(sh, gz, pts, dd); for (int i = 0; i < n; i++) n2h += dd[i] < 0; }
    if (n2g > 4) n2g = 4; if (n2h > 4) n2h = 4;
  }
  bool rung = false, runh = false;
  if (has_obj) {
    T ro = shape_bound(so);
    Shape<T> so0 = so; so0.pos = mk<T>(0, 0, 0);
    if (gtype >= 0) {
      V3<T> d = sg.pos - so.pos; T rr = ro + shape_bound(sg);
      if (dot(d, d) <= rr * rr) { Shape<T> s0 = sg; s0.pos = d; hitg = mpr_penetration<T>(s0, so0, &depg, &ng, &pg); rung = true; }
    }
    if (htype >= 0) {
      V3<T> d = sh.pos - so.pos; T rr = ro + shape_bound(sh);
      if (dot(d, d) <= rr * rr) { Shape<T> s0 = sh; s0.pos = d; hith = mpr_penetration<T>(s0, so0, &deph, &nh, &ph); runh = true; }
    }
  }
  // MPR passes this warp ran (load key of the lockstep sort, see sort_key_of): the env's figure is the max over its warps
  { const int passes = (__any_sync(GMJ_FULL, rung) ? 1 : 0) + (__any_sync(GMJ_FULL, runh) ? 1 : 0); if (lane == 31) ws.mprw[w] = passes; }
  // slots in body order: chain of finger 0, 1, 2, then the palm; category (2) before category (3)
  int ncon;
  {
    const int c2 = n2g + n2h, c3 = hitg + hith;
    int s2 = tp.chain ? c2 : 0, s3 = tp.chain ? c3 : 0;       // inclusive scans over the chain lanes
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      int t2 = __shfl_up_sync(GMJ_FULL, s2, off), t3 = __shfl_up_sync(GMJ_FULL, s3, off);
      if (lane >= off) { s2 += t2; s3 += t3; }
    }
    if (lane == 31) { ws.cnt[0][w] = s2; ws.cnt[1][w] = s3; }
    if (tp.palm) { ws.cnt[0][3] = c2; ws.cnt[1][3] = c3; }
    env_sync<T, N>(ws);
    if (tid == 0) { int m = ws.mprw[0]; if (ws.mprw[1] > m) m = ws.mprw[1]; if (ws.mprw[2] > m) m = ws.mprw[2]; e.mpr_prev = m; }
    const int grp = tp.palm ? 3 : w;
    int off2 = n1, off3 = n1 + ws.cnt[0][0] + ws.cnt[0][1] + ws.cnt[0][2] + ws.cnt[0][3];
    for (int g = 0; g < grp; g++) { off2 += ws.cnt[0][g]; off3 += ws.cnt[1][g]; }
    if (tp.chain) { off2 += s2 - c2; off3 += s3 - c3; }
    int all = n1;
#pragma unroll
    for (int g = 0; g < 4; g++) all += ws.cnt[0][g] + ws.cnt[1][g];
    ncon = all < GMJ_MAXCON ? all : GMJ_MAXCON;
    if (geo && c2 > 0) {
      V3<T> pts[8]; T dd[8];
      int slot = off2;
      for (int which = 0; which < 2; which++) {
        const int want = which == 0 ? n2g : n2h;
        if (!want) continue;
        const Shape<T>& s = which == 0 ? sg : sh;
        int n = plane_candidates<T>(s, gz, pts, dd);
        T mu = maxT<T>(M.ground_friction, gfric);
        for (int k = 0; k < 4 && k < n; k++) {
          int best = -1;
          for (int i = 0; i < n; i++) if (dd[i] < 0 && (best < 0 || dd[i] < dd[best])) best = i;
          if (best < 0) break;
          if (slot < GMJ_MAXCON) { V3<T> pp = pts[best]; pp.z -= T(0.5) * dd[best]; add_contact(pp, mk<T>(0, 0, 1), dd[best], mu, -1, GMJ_CLS_GROUND, tid, gcls, slot); }
          slot++;
          dd[best] = 1;
        }
      }
    }
    if (geo && c3 > 0) {
      int slot = off3;
      if (hitg) { if (slot < GMJ_MAXCON) add_contact(pg + so.pos, ng, -depg, maxT<T>(gfric, ofric), tid, gcls, TW::OBJ, GMJ_CLS_OBJECT, slot); slot++; }
      if (hith) { if (slot < GMJ_MAXCON) add_contact(ph + so.pos, nh, -deph, maxT<T>(gfric, ofric), tid, gcls, TW::OBJ, GMJ_CLS_OBJECT, slot); slot++; }
    }
    env_sync<T, N>(ws);
  }

  wide_phase<T, N>();
  // ------------------------------------------------------------------ constraint rows
  int nr = 0;
  const T zero6[6] = { 0, 0, 0, 0, 0, 0 };
  // welds: joint locked at its anchor; rows = world components of the joint axis. One unit solve per weld,
  // all seven in one pass (each is non-zero in one warp only).
  {
    const unsigned won = e.weld_on & 0x7f;
    if (won) {
      T x[7];
#pragma unroll
      for (int m = 0; m < 7; m++) x[m] = (((won >> m) & 1) && holds_slot<N>(tp, weld_slot_w(m))) ? T(1) : T(0);
      fwd_w<T, N, 7, 0>(ws, xn, tp, L1w, lb, x, nullptr);
#pragma unroll
      for (int m = 0; m < 7; m++) {
        if (!((won >> m) & 1)) continue;
        const int js = weld_slot_w(m);
        const T qj = ws.jq[js], qdj = ws.jqd[js];
        const bool rev = (m >= 3 && m < 6);
        const T sgn = rev ? T(-1) : T(1);
        T s1, c1; sincosT<T>(qj - e.weld_anchor[m], &s1, &c1);
        const T rho = rev ? s1 : (qj - e.weld_anchor[m]);
        const T comp[3] = { ws.jaw[js][0], ws.jaw[js][1], ws.jaw[js][2] };
        const T xw = x[m] * rsq;
        const int jdof = m < 3 ? 1 + m * NL : (m < 6 ? 2 + (m - 3) * NL : B - 1);
#pragma unroll
        for (int d = 0; d < 3; d++) {
          if (comp[d] * comp[d] < T(1e-10)) continue;
          if (nr >= GMJ_RMAX) break;
          put_row_w<T, N>(ws, tp, rsqo, nr, sgn * comp[d] * xw, zero6, ROW_EQ, sgn * rho * comp[d], sgn * comp[d] * qdj, T(0), nr, jdof);
          nr++;
        }
      }
    }
  }
  // joint limits (margin 0), in dof order: base, finger 0, 1, 2, palm
  {
    unsigned m0 = ws.limmask[0], m1 = ws.limmask[1], m2 = ws.limmask[2];
    const unsigned chainmask = NL >= 32 ? 0xffffffffu : ((1u << NL) - 1u);
    // ordered list encoded as up to four passes over masks: (base) (w0 chain) (w1 chain) (w2 chain) (palm)
    int pending[4]; int np = 0;
    auto flush = [&]() {
      T x[4];
#pragma unroll
      for (int m = 0; m < 4; m++) x[m] = (m < np && holds_slot<N>(tp, pending[m])) ? T(1) : T(0);
      fwd_w<T, N, 4, 0>(ws, xn, tp, L1w, lb, x, nullptr);
#pragma unroll
      for (int m = 0; m < 4; m++) {
        if (m >= np) continue;
        const int js = pending[m];
        const TW tj(js);
        const T loj = ws.jlo[js], hij = ws.jhi[js], qdj = ws.jqd[js];
        const T xw = x[m] * rsq;
        if (loj < 0 && nr < GMJ_RMAX) { put_row_w<T, N>(ws, tp, rsqo, nr, xw, zero6, ROW_LIMIT, loj, qdj, T(1), nr, tj.dof); nr++; }
        if (hij < 0 && nr < GMJ_RMAX) { put_row_w<T, N>(ws, tp, rsqo, nr, -xw, zero6, ROW_LIMIT, hij, -qdj, T(-1), nr, tj.dof); nr++; }
      }
      np = 0;
    };
    if ((m0 >> TW::BASE_L) & 1) { pending[np++] = TW::BASE_SLOT; }
    for (int ww = 0; ww < 3; ww++) {
      unsigned mm = (ww == 0 ? m0 : (ww == 1 ? m1 : m2)) & chainmask;
      while (mm) {
        int l = __ffs(mm) - 1; mm &= mm - 1;
        pending[np++] = 32 * ww + l;
        if (np == 4) flush();
      }
    }
    if ((m0 >> TW::PALM_L) & 1) { pending[np++] = TW::PALM_SLOT; if (np == 4) flush(); }
    if (np > 0) flush();
  }
  // contacts: pyramidal, condim 3 -> 4 rows each, linear in the three frame rows (n, t1, t2), whitened together
#pragma unroll 1
  for (int cidx = 0; cidx < ncon; cidx++) {
    if (nr + 4 > GMJ_RMAX) { ncon = cidx; break; }
    DevContact<T>& cn = ws.ax.con[cidx];
    V3<T> cp = mk<T>(cn.pos[0], cn.pos[1], cn.pos[2]);
    T mu = cn.mu;
    int b1 = cn.body1, b2 = cn.body2;
    T Jc[3], Joc[3][6], jvc[3];
#pragma unroll
    for (int ax3 = 0; ax3 < 3; ax3++) {
      V3<T> d = mk<T>(cn.frame[3 * ax3], cn.frame[3 * ax3 + 1], cn.frame[3 * ax3 + 2]);
      T j2 = jac_entry_w<T, N>(tp, b2, cp, d, aw, anchor);
      T j1 = jac_entry_w<T, N>(tp, b1, cp, d, aw, anchor);
      Jc[ax3] = j2 - j1;
#pragma unroll
      for (int k = 0; k < 6; k++) Joc[ax3][k] = 0;
      if (b2 == TW::OBJ) {
        Joc[ax3][0] = d.x; Joc[ax3][1] = d.y; Joc[ax3][2] = d.z;
        V3<T> r = cp - po;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          V3<T> axk = mk<T>(Ro.m[k], Ro.m[3 + k], Ro.m[6 + k]);
          Joc[ax3][3 + k] = dot(d, cross(axk, r));
        }
      }
      jvc[ax3] = tp.own ? Jc[ax3] * qd : T(0);
    }
    const bool gripper_part = (b1 >= 0 && b1 < TW::OBJ) || (b2 >= 0 && b2 < TW::OBJ);
    T xw[3] = { Jc[0], Jc[1], Jc[2] };
    if (gripper_part) fwd_w<T, N, 3, 3>(ws, xn, tp, L1w, lb, xw, jvc);      // object-ground contacts have no gripper columns
    else { jvc[0] = jvc[1] = jvc[2] = 0; }
#pragma unroll
    for (int ax3 = 0; ax3 < 3; ax3++) {
      xw[ax3] *= rsq;
      jvc[ax3] += (b2 == TW::OBJ) ? (Joc[ax3][0] * vo.x + Joc[ax3][1] * vo.y + Joc[ax3][2] * vo.z + Joc[ax3][3] * wo.x + Joc[ax3][4] * wo.y + Joc[ax3][5] * wo.z) : T(0);
    }
    if (tid == 0) cn.row = nr;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int t = 1 + (k >> 1); const T sg2 = (k & 1) ? -mu : mu;
      T Jo[6];
#pragma unroll
      for (int i = 0; i < 6; i++) Jo[i] = Joc[0][i] + sg2 * Joc[t][i];
      put_row_w<T, N>(ws, tp, rsqo, nr + k, xw[0] + sg2 * xw[t], Jo, ROW_CONTACT, cn.dist, jvc[0] + sg2 * jvc[t], mu, nr, -1);
    }
    nr += 4;
  }
  // whitened smooth force as row nr
  if (tp.own) ws.W[nr][tp.col] = wsm;
  if (tid == 0) {
#pragma unroll
    for (int k = 0; k < 6; k++) ws.W[nr][B + k] = tauo[k] * rsqo[k];
    ws.ax.ncon = ncon; ws.ax.nrow = nr;
  }
  env_sync<T, N>(ws);

  wide_phase<T, N>();
#ifndef GMJ_DUAL_WIDE
  // ------------------------------------------------------------------ primal active-set solve (gmj_primal_wide.cuh)
  int iters = 0;
  T g = 0, go = 0;
  {
    const T tc = maxT<T>(M.solref_tc, 2 * dt), dr = M.solref_dr, dmax = M.solimp[1];
    const T kk = rcpT<T>(dmax * dmax * tc * tc * dr * dr), bb = T(2) * rcpT<T>(dmax * tc);
    T pf = 0;
    if (nr > 0) iters = primal_solve_w<T, N>(ws, tp, M.solimp, kk, bb, nr, tid, e.nrow_prev == nr, e.fprev, pf, g, go);
    env_sync<T, N>(ws);
    if (tid < nr) e.fprev[tid] = pf;
    if (tid == 0) { e.nrow_prev = nr; ws.ax.niter = iters; }
  }
  env_sync<T, N>(ws);

  wide_phase<T, N>();
  // ------------------------------------------------------------------ constraint force, integration
  // g = whitened constraint force = x - x_s straight from the primal solution
#else
  // ------------------------------------------------------------------ A = W W^T, all 96 threads share the entries
  form_A<T, N, GMJ_WNT>(ws, nr, tid, nullptr);
  env_sync<T, N>(ws);
  // ------------------------------------------------------------------ regulariser + dual LCP: warp 0
  int iters = 0;
  if (w == 0) {
    constexpr int NW = B + 6;
    const int r0 = lane, r1 = lane + 32;
    T bcol[2];
#pragma unroll 1
    for (int slot = 0; slot < 2; slot++) {
      const int r = lane + 32 * slot;
      bcol[slot] = 0;
      if (r >= nr) continue;
      const T* wr = ws.W[r];
      const T* wn = ws.W[nr];
      T a0 = 0, a1 = 0;
      for (int k = 0; k + 1 < NW; k += 2) { a0 += wr[k] * wn[k]; a1 += wr[k + 1] * wn[k + 1]; }
      if (NW & 1) a0 += wr[NW - 1] * wn[NW - 1];
      bcol[slot] = a0 + a1;
    }
    T pf[2], pb[2], pR[2]; int ptype[2];
    T tc = maxT<T>(M.solref_tc, 2 * dt), dr = M.solref_dr, dmax = M.solimp[1];
    T kk = rcpT<T>(dmax * dmax * tc * tc * dr * dr), bb = T(2) * rcpT<T>(dmax * tc);
#pragma unroll
    for (int slot = 0; slot < 2; slot++) {
      int r = slot ? r1 : r0;
      pf[slot] = 0; pb[slot] = 0; pR[slot] = 1; ptype[slot] = ROW_EQ;
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
        pR[slot] = Rv; pb[slot] = bcol[slot] - aref; ptype[slot] = type;
        pf[slot] = (e.nrow_prev == nr) ? e.fprev[r] : T(0);
      }
    }
    __syncwarp();
    if (nr > 0) iters = lcp_solve<T, N>(ws, nr, lane, pb, pR, ptype, pf, e.nrow_prev == nr);
    __syncwarp();
    if (r0 < nr) e.fprev[r0] = pf[0];
    if (r1 < nr) e.fprev[r1] = pf[1];
    if (lane == 0) { e.nrow_prev = nr; ws.ax.niter = iters; }
  }
  env_sync<T, N>(ws);

  wide_phase<T, N>();
  // ------------------------------------------------------------------ constraint force, integration
  T g = 0, go = 0;
  for (int r = 0; r < nr; r++) {
    T fr = e.fprev[r];
    if (tp.act) g += ws.W[r][tp.col] * fr;
    if (tid < 6) go += ws.W[r][B + tid] * fr;
  }
#endif
  // tau_c = L D^1/2 g
  T tauc;
  {
    const T x = tp.act ? g * sqrtT<T>(D) : T(0);
    T acc = x;
#pragma unroll 2
    for (int k = 0; k < NL - 1; k++) {
      T xk = __shfl_sync(GMJ_FULL, x, k);
      if (tp.chain && tp.local > k) acc += L1w[tp.local * NLP + k] * xk;
    }
    T s[1] = { (tp.chain || tp.palm) ? lb * x : T(0) };
    xsum<T, N, 1>(ws, xn, w, lane, s);
    if (tp.base) acc += s[0];
    tauc = acc;
  }
  if (tp.base && w == 0) {
    for (int r = 0; r < nr; r++) if (ws.rtype[r] == ROW_LIMIT && ws.rlane[r] == 0) wrist += ws.rmu[r] * e.fprev[r];
    ws.ax.wrist_z = wrist;
  }
  // contact forces in the contact frame (mj_contactForce, pyramidal condim 3)
  if (tid < ncon) {
    DevContact<T>& cn = ws.ax.con[tid];
    int rr = cn.row;
    T f0 = e.fprev[rr], f1 = e.fprev[rr + 1], f2 = e.fprev[rr + 2], f3 = e.fprev[rr + 3];
    cn.force[0] = f0 + f1 + f2 + f3; cn.force[1] = cn.mu * (f0 - f1); cn.force[2] = cn.mu * (f2 - f3);
  }
  // Euler with implicit damping: (M + h D) a = tau + tau_c
  T x;
  { T x1[1] = { tp.act ? tau + tauc : T(0) }; fwd_w<T, N, 1, 0>(ws, xn, tp, L2w, lb2, x1, nullptr); x = x1[0]; }
  x = tp.act ? x * rcpT<T>(tp.act ? D2 : T(1)) : T(0);
  {
    T xb = __shfl_sync(GMJ_FULL, x, TW::BASE_L);
    if (tp.chain || tp.palm) x -= lb2 * xb;
#pragma unroll 2
    for (int k = NL - 1; k >= 1; k--) {
      T xk = __shfl_sync(GMJ_FULL, x, k);
      if (tp.chain && tp.local < k) x -= L2w[k * NLP + tp.local] * xk;
    }
  }
  bad |= !(absT<T>(x) < T(1e10));
  bad |= !(D2 > 0) && tp.act;
  if (tp.own) { qd += dt * x; q += dt * qd; e.qpos[tp.dof] = q; e.qvel[tp.dof] = qd; }
  if (w == 0) {
    if (has_obj) {
      T ao[6];
#pragma unroll
      for (int k = 0; k < 6; k++) { T gk = __shfl_sync(GMJ_FULL, go, k); ao[k] = (tauo[k] + gk * rcpT<T>(rsqo[k])) * rcpT<T>(Mo[k]); bad |= !(absT<T>(ao[k]) < T(1e10)); }
      if (lane == 0) {
        vo = vo + mk<T>(ao[0], ao[1], ao[2]) * dt; wo = wo + mk<T>(ao[3], ao[4], ao[5]) * dt;
        e.qvel[OQ] = vo.x; e.qvel[OQ + 1] = vo.y; e.qvel[OQ + 2] = vo.z; e.qvel[OQ + 3] = wo.x; e.qvel[OQ + 4] = wo.y; e.qvel[OQ + 5] = wo.z;
        e.qpos[OQ] += dt * vo.x; e.qpos[OQ + 1] += dt * vo.y; e.qpos[OQ + 2] += dt * vo.z;
        e.obj_xy[0] += dtd * (double)vo.x; e.obj_xy[1] += dtd * (double)vo.y;
        T ang = sqrtT<T>(dot(wo, wo)) * dt;
        T a0 = e.qpos[OQ + 3], a1 = e.qpos[OQ + 4], a2 = e.qpos[OQ + 5], a3 = e.qpos[OQ + 6];
        if (ang > 0) {
          V3<T> axn = normalized(wo);
          T s, cs; sincosT<T>(ang / 2, &s, &cs);
          T d0 = cs, d1 = axn.x * s, d2 = axn.y * s, d3 = axn.z * s;
          T n0 = a0 * d0 - a1 * d1 - a2 * d2 - a3 * d3, n1q = a0 * d1 + a1 * d0 + a2 * d3 - a3 * d2;
          T n2 = a0 * d2 - a1 * d3 + a2 * d0 + a3 * d1, n3 = a0 * d3 + a1 * d2 - a2 * d1 + a3 * d0;
          a0 = n0; a1 = n1q; a2 = n2; a3 = n3;
        }
        T l = T(1) / sqrtT<T>(a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3);
        e.qpos[OQ + 3] = a0 * l; e.qpos[OQ + 4] = a1 * l; e.qpos[OQ + 5] = a2 * l; e.qpos[OQ + 6] = a3 * l;
      }
    }
  }
  unsigned anybad = __ballot_sync(GMJ_FULL, bad);
  if (anybad && lane == 0) ws.ebad = 1;
  env_sync<T, N>(ws);
  if (tid == 0) {
    if (ws.ebad) e.unstable = 1;
    e.time += dtd;
    e.st_substeps++; e.st_contacts += ncon; e.st_rows += nr; e.st_iters += iters & 0xffff; e.st_relaxed += (iters >> 16) & 1; e.st_capped += (iters >> 17) & 1;
  }
}
