// Wall-normal stage, segment-parallel version: the same arithmetic as kernels_pencil.cu (which stays
// as the fall-back for NY > 768 and as the cross-check in the tests), reorganised so that NOTHING
// makes a round trip through HBM scratch and a chain is worked on by many threads at once.
//
// A thread owns S consecutive rows (a y-segment) of one real chain (mode, re|im) in REGISTERS.  A
// block is CB chains x nseg = ceil(NY/S) segments; the CB chains of a segment are neighbouring lanes,
// so every global access is still a coalesced row segment of the plane-major layout.
//
// A tridiagonal sweep d'_j = b_j + a_j d'_{j-1} is affine in the value entering the segment:
//   1. each thread sweeps its segment with 0 entering and keeps the prefix products P_r = a_0..a_r,
//   2. publishes (P_last, y_last) in shared memory; after one barrier every thread composes the maps
//      of the segments before it (<= nseg fused multiply-adds) to get its true entering value,
//   3. fixes up its rows: d'_r = y_r + P_r * entering.
// Back substitution is the same from the other end.  For the fixed d/dy matrix the products P_r come
// from a table; for the kk-dependent Helmholtz / implicit matrices the pivots are precomputed per
// mode by pivot_kernel (once for Helmholtz; per substep for the implicit matrix, whose D = nu*beta*dt
// changes) and the products are formed on the fly.
//
// Traffic per substep (P = one compact field): pre 3P + 5P, post 7P + 6P + 1P of pivots, pivots
// 0.5P written: 22.5P against 55P of the scratch version; parallelism x nseg (25 at NY = 193).
#include <vector>
#include <string.h>
#include "pencil_common.cuh"

namespace chan {

struct SegTabs { const double *pf, *pb; const double *pivH, *pivI; };

struct Ctx {
  int tx, seg, nseg, cb, ny, j0, nrow;     // nrow = valid rows of this segment
  double *smH, *smF, *smB, *smW;            // halos, forward maps, backward maps, wall values
};

template <int S> __device__ __forceinline__ bool row_ok(const Ctx& c, int r) { return r < c.nrow; }

// neighbours' edge rows: hm[q][0] = f(j0-1), hm[q][1] = f(j0-2), hp[q][0] = f(j0+S), hp[q][1] = f(j0+S+1)
template <int S, int NR>
__device__ __forceinline__ void halos(const Ctx& c, const double (&f)[NR][S], double (&hm)[NR][2], double (&hp)[NR][2]) {
  const int n = c.nseg * c.cb, me = c.seg * c.cb + c.tx;
#pragma unroll
  for (int q = 0; q < NR; q++) {
    double* b = c.smH + (size_t)q * 4 * n;
    b[0 * n + me] = f[q][0]; b[1 * n + me] = f[q][1]; b[2 * n + me] = f[q][S - 1]; b[3 * n + me] = f[q][S - 2];
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < NR; q++) {
    const double* b = c.smH + (size_t)q * 4 * n;
    bool lo = c.seg > 0, hi = c.seg + 1 < c.nseg;
    hm[q][0] = lo ? b[2 * n + me - c.cb] : 0.0; hm[q][1] = lo ? b[3 * n + me - c.cb] : 0.0;
    hp[q][0] = hi ? b[0 * n + me + c.cb] : 0.0; hp[q][1] = hi ? b[1 * n + me + c.cb] : 0.0;
  }
}

// entering values of a forward (upward) recurrence: compose the maps of segments 0..seg-1
template <int NR>
__device__ __forceinline__ void compose_fwd(const Ctx& c, double A, const double (&Y)[NR], double (&xin)[NR]) {
  const int n = c.nseg * c.cb, me = c.seg * c.cb + c.tx;
  c.smF[me] = A;
#pragma unroll
  for (int q = 0; q < NR; q++) c.smF[(size_t)(1 + q) * n + me] = Y[q];
  __syncthreads();
#pragma unroll
  for (int q = 0; q < NR; q++) xin[q] = 0.0;
  for (int s = 0; s < c.seg; s++) {
    double a = c.smF[s * c.cb + c.tx];
#pragma unroll
    for (int q = 0; q < NR; q++) xin[q] = c.smF[(size_t)(1 + q) * n + s * c.cb + c.tx] + a * xin[q];
  }
}
template <int NR>
__device__ __forceinline__ void compose_bwd(const Ctx& c, double A, const double (&Y)[NR], double (&xin)[NR]) {
  const int n = c.nseg * c.cb, me = c.seg * c.cb + c.tx;
  c.smB[me] = A;
#pragma unroll
  for (int q = 0; q < NR; q++) c.smB[(size_t)(1 + q) * n + me] = Y[q];
  __syncthreads();
#pragma unroll
  for (int q = 0; q < NR; q++) xin[q] = 0.0;
  for (int s = c.nseg - 1; s > c.seg; s--) {
    double a = c.smB[s * c.cb + c.tx];
#pragma unroll
    for (int q = 0; q < NR; q++) xin[q] = c.smB[(size_t)(1 + q) * n + s * c.cb + c.tx] + a * xin[q];
  }
}

// x = d/dy f for NR chains at once, in place (f -> derivative), fp64.  Same folded-LU tables as ElimD1.
template <int S, int NR>
__device__ void d1_seg(const Ctx& c, const D1Tab& t, const SegTabs& st, double (&f)[NR][S]) {
  double hm[NR][2], hp[NR][2];
  halos<S, NR>(c, f, hm, hp);
  const int ny = c.ny;
  double Y[NR], xin[NR];
  // forward: y_r = rhs_r - eL_j y_{r-1}
#pragma unroll
  for (int q = 0; q < NR; q++) {
    double fm2 = hm[q][1], fm1 = hm[q][0], yprev = 0.0;
#pragma unroll
    for (int r = 0; r < S; r++) {
      int j = c.j0 + r;
      double f0 = f[q][r];
      double fp1 = r + 1 < S ? f[q][r + 1] : hp[q][0];
      double fp2 = r + 2 < S ? f[q][r + 2] : (r + 1 < S ? hp[q][0] : hp[q][1]);
      double y = 0.0;
      if (j < ny) {
        double eP = __ldg(t.eP + j), eC = __ldg(t.eC + j), eN = __ldg(t.eN + j);
        if (j == 0) y = eC * f0 + eN * fp1 + eP * fp2;
        else if (j == ny - 1) y = eC * f0 + eP * fm1 + eN * fm2 - __ldg(t.eL + j) * yprev;
        else y = eP * fm1 + eC * f0 + eN * fp1 - __ldg(t.eL + j) * yprev;
      }
      fm2 = fm1; fm1 = f0; f[q][r] = y; yprev = y;
    }
  }
  // the last VALID row carries the map out of the segment
#pragma unroll
  for (int q = 0; q < NR; q++) {
    double y = f[q][0];
#pragma unroll
    for (int r = 1; r < S; r++) if (r < c.nrow) y = f[q][r];
    Y[q] = y;
  }
  const int jl = c.j0 + c.nrow - 1;
  compose_fwd<NR>(c, __ldg(st.pf + jl), Y, xin);
#pragma unroll
  for (int q = 0; q < NR; q++) {
#pragma unroll
    for (int r = 0; r < S; r++) if (r < c.nrow) f[q][r] += __ldg(st.pf + c.j0 + r) * xin[q];
  }
  // backward: x_r = d'_r - cp_j x_{r+1}
#pragma unroll
  for (int q = 0; q < NR; q++) {
    double xn = 0.0;
#pragma unroll
    for (int r = S - 1; r >= 0; r--) {
      if (r < c.nrow) { xn = f[q][r] - __ldg(t.cp + c.j0 + r) * xn; f[q][r] = xn; }
    }
    Y[q] = f[q][0];
  }
  compose_bwd<NR>(c, __ldg(st.pb + c.j0), Y, xin);
#pragma unroll
  for (int q = 0; q < NR; q++) {
#pragma unroll
    for (int r = 0; r < S; r++) if (r < c.nrow) f[q][r] += __ldg(st.pb + c.j0 + r) * xin[q];
  }
}

// Helmholtz (implicit = false) or implicit solve with precomputed pivots piv[j*M + m], in place.
template <int S, int NR>
__device__ void var_seg(const Ctx& c, const D2Tab& t, const double* __restrict__ piv, long M, long m, bool implicit,
                        double kk, double D, double (&f)[NR][S]) {
  double hm[NR][2], hp[NR][2];
  halos<S, NR>(c, f, hm, hp);
  const int ny = c.ny;
  double a_[S], cp_[S];            // a_j = -lo_j * piv_j ; cp_j = up_j * piv_j
  double pvv[S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    a_[r] = 0.0; cp_[r] = 0.0; pvv[r] = 1.0;
    if (j > 0 && j < ny - 1) {
      double rP = __ldg(t.rP + j), rN = __ldg(t.rN + j), wP = __ldg(t.wP + j), wN = __ldg(t.wN + j);
      double lo, up;
      if (implicit) { lo = rP - D * wP + D * kk * rP; up = rN - D * wN + D * kk * rN; }
      else          { lo = wP - kk * rP;              up = wN - kk * rN; }
      if (j == 1) lo = 0.0;
      if (j == ny - 2) up = 0.0;
      double pv = piv[(long)j * M + m];
      pvv[r] = pv; a_[r] = -lo * pv; cp_[r] = up * pv;
    }
  }
  double Y[NR], xin[NR], P[S];
  {
    double p = 1.0;
#pragma unroll
    for (int r = 0; r < S; r++) { p *= a_[r]; P[r] = p; }
  }
#pragma unroll
  for (int q = 0; q < NR; q++) {
    double fm1 = hm[q][0], yprev = 0.0;
#pragma unroll
    for (int r = 0; r < S; r++) {
      int j = c.j0 + r;
      double f0 = f[q][r];
      double fp1 = r + 1 < S ? f[q][r + 1] : hp[q][0];
      double y = 0.0;
      if (j > 0 && j < ny - 1) {
        double rhs = __ldg(t.rN + j) * fp1 + f0 + __ldg(t.rP + j) * fm1;
        y = rhs * pvv[r] + a_[r] * yprev;
      }
      fm1 = f0; f[q][r] = y; yprev = y;
    }
    double y = f[q][0];
#pragma unroll
    for (int r = 1; r < S; r++) if (r < c.nrow) y = f[q][r];
    Y[q] = y;
  }
  {
    double A = P[0];
#pragma unroll
    for (int r = 1; r < S; r++) if (r < c.nrow) A = P[r];
    compose_fwd<NR>(c, A, Y, xin);
  }
#pragma unroll
  for (int q = 0; q < NR; q++) {
#pragma unroll
    for (int r = 0; r < S; r++) if (r < c.nrow) f[q][r] += P[r] * xin[q];
  }
  // backward
  {
    double p = 1.0;
#pragma unroll
    for (int r = S - 1; r >= 0; r--) { if (r < c.nrow) p *= -cp_[r]; P[r] = p; }
  }
#pragma unroll
  for (int q = 0; q < NR; q++) {
    double xn = 0.0;
#pragma unroll
    for (int r = S - 1; r >= 0; r--) if (r < c.nrow) { xn = f[q][r] - cp_[r] * xn; f[q][r] = xn; }
    Y[q] = f[q][0];
  }
  compose_bwd<NR>(c, P[0], Y, xin);
#pragma unroll
  for (int q = 0; q < NR; q++) {
#pragma unroll
    for (int r = 0; r < S; r++) if (r < c.nrow) f[q][r] += P[r] * xin[q];
  }
}

template <typename T, int S> __device__ __forceinline__ void round_vec(double (&f)[S]) {
#pragma unroll
  for (int r = 0; r < S; r++) f[r] = rnd<T>(f[r]);
}

template <typename T>
__device__ __forceinline__ bool setup_seg(const PArgs<T>& a, int cb, Mode& md, Wave<T>& wv, Ctx& c, double* sm, int S) {
  c.tx = threadIdx.x; c.seg = threadIdx.y; c.nseg = blockDim.y; c.cb = cb; c.ny = a.ny;
  c.j0 = c.seg * S; c.nrow = min(S, a.ny - c.j0);
  const int n = c.nseg * cb;
  c.smH = sm; c.smF = c.smH + 8 * n; c.smB = c.smF + 3 * n; c.smW = c.smB + 3 * n;
  long e = (long)blockIdx.x * cb + threadIdx.x;
  bool active = e < 2 * a.M;
  if (!active) e = 2 * a.M - 2 + (e & 1);           // idle lanes shadow the last pair (they never store)
  md.e = e; md.m = e >> 1; md.c = (int)(e & 1);
  int ikx = a.ikx0 + (int)(md.m / a.Kz);
  md.kzi = (int)(md.m % a.Kz);
  md.kxi = ikx <= a.kxmax ? ikx : ikx - a.Kx;
  md.zero_mode = (md.kxi == 0 && md.kzi == 0);
  wv.k1 = a.k1fac * (T)md.kxi; wv.k3 = a.k3fac * (T)md.kzi;
  wv.kk = wv.k1 * wv.k1 + wv.k3 * wv.k3;
  wv.rkk = md.zero_mode ? (T)1 : (T)1 / wv.kk;
  double k1d = (double)a.k1fac * (double)md.kxi, k3d = (double)a.k3fac * (double)md.kzi;
  wv.kkd = k1d * k1d + k3d * k3d;
  wv.sc = md.c ? (T)1 : (T)-1; wv.sp = -wv.sc;
  return active;
}

// =================================================================================
template <typename T, int S, int CB, int MINB, int LB>
__global__ void __launch_bounds__(LB, MINB) pencil_pre_seg_kernel(PArgs<T> a, SegTabs st, PeerOut po) {
  typedef typename Vec2<T>::type T2;
  extern __shared__ double sm_seg[];
  Mode md; Wave<T> wv; Ctx c;
  const bool active = setup_seg<T>(a, CB, md, wv, c, sm_seg, S);
  const long stp = 2 * a.M; const int cc = md.c;
  const bool with_mean = md.zero_mode && a.use_mean;
  double f[2][S];
  T uu[S], ww[S], upp[S], wpp[S], vp[S], vo[S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    uu[r] = ww[r] = upp[r] = wpp[r] = vp[r] = vo[r] = 0;
    if (j < c.ny) {
      long cm = (long)j * a.M + md.m;
      T2 dvz = a.dv[cm], gz = a.g[cm];
      uw_comp<T, T2>(wv, md, md.zero_mode, dvz, gz, cc, uu[r], ww[r]);
      uw_comp<T, T2>(wv, md, md.zero_mode, dvz, gz, cc ^ 1, upp[r], wpp[r]);
      if (with_mean) { uu[r] = cc ? (T)0 : a.mean_u[j]; upp[r] = cc ? a.mean_u[j] : (T)0; }   // writeUmean (nonLinear.c:19-21)
      { T2 vz = a.v[cm]; vp[r] = other<T2, T>(vz, cc); vo[r] = own<T2, T>(vz, cc); }
    }
    f[0][r] = (double)uu[r]; f[1][r] = (double)ww[r];
  }
  d1_seg<S, 2>(c, a.yt.d1[0], st, f);
  if (!active) return;
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    if (j < c.ny) {
      long re = (long)j * stp + md.e;
      T dudy = (T)f[0][r], dwdy = (T)f[1][r], s = wv.sc;
      T o1 = dwdy - s * wv.k3 * vp[r];                  // omega_x = dw/dy - i k3 v
      T o2 = s * (wv.k3 * upp[r] - wv.k1 * wpp[r]);     // omega_y = i (k3 u - k1 w)
      T o3 = -dudy + s * wv.k1 * vp[r];                 // omega_z = -du/dy + i k1 v
      if (po.nranks > 1) {
        // multi-GPU: row j goes straight into the receive buffer of the rank that owns plane j (NVLink
        // store for the other ranks) — this IS the pencil -> plane transpose of channel_cuda_mpi.c:95-128
        int p = 0;
#pragma unroll
        for (int i = 1; i < 8; i++) p += (i < po.nranks && j >= po.part_j[i]) ? 1 : 0;
        long idx = 2 * (po.src_off[p] + (long)(j - po.part_j[p]) * a.M) + md.e;
        ((T*)po.rcv[p][0])[idx] = uu[r]; ((T*)po.rcv[p][1])[idx] = vo[r]; ((T*)po.rcv[p][2])[idx] = ww[r];
        ((T*)po.rcv[p][3])[idx] = o1; ((T*)po.rcv[p][4])[idx] = o2; ((T*)po.rcv[p][5])[idx] = o3;
      } else {
        ((T*)a.u)[re] = uu[r]; ((T*)a.w)[re] = ww[r];
        ((T*)a.ox)[re] = o1; ((T*)a.oy)[re] = o2; ((T*)a.oz)[re] = o3;
      }
      if (md.zero_mode && cc == 0 && a.sc) {                      // readUtau (convolution.c:14-16)
        if (j == 0) a.sc->dudy_wall[0] = (double)dudy;
        if (j == c.ny - 1) a.sc->dudy_wall[1] = (double)dudy;
      }
    }
  }
}

// =================================================================================
template <typename T, int S, int CB, int MINB, int LB>
__global__ void __launch_bounds__(LB, MINB) pencil_post_seg_kernel(PArgs<T> a, SegTabs st) {
  typedef typename Vec2<T>::type T2;
  extern __shared__ double sm_seg[];
  Mode md; Wave<T> wv; Ctx c;
  const bool active = setup_seg<T>(a, CB, md, wv, c, sm_seg, S);
  const long stp = 2 * a.M; const int cc = md.c;
  const T dt = (T)a.sc->dt, s = wv.sc;
  const T* phiR = (const T*)a.phi; const T* gR = (const T*)a.g;
  const T2* H1 = a.ox; const T2* H3 = a.oz; const T* H2R = (const T*)a.oy;
  // ---- hv = -i (k1 H1 + k3 H3) -> d/dy -> R_phi_new = hv' - kk H2 ; R_g_new = i (k3 H1 - k1 H3)
  double rpn[1][S]; T rgn[S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r; rpn[0][r] = 0.0; rgn[r] = 0;
    if (j < c.ny) {
      long cm = (long)j * a.M + md.m;
      T h1p = other<T2, T>(H1[cm], cc), h3p = other<T2, T>(H3[cm], cc);
      rpn[0][r] = (double)(-s * (wv.k1 * h1p + wv.k3 * h3p));
      rgn[r] = s * (wv.k3 * h1p - wv.k1 * h3p);
    }
  }
  d1_seg<S, 1>(c, a.yt.d1[0], st, rpn);
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    if (j < c.ny) rpn[0][r] = (double)((T)rpn[0][r] - wv.kk * H2R[(long)j * stp + md.e]);      // calcHvvkernel (:180-181)
  }
  // ---- d2/dy2 of phi and g = d/dy twice with the cast in between (Q4)
  double f[2][S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r; long re = (long)(j < c.ny ? j : c.ny - 1) * stp + md.e;
    f[0][r] = (double)phiR[re]; f[1][r] = (double)gR[re];
  }
  d1_seg<S, 2>(c, a.yt.d1[0], st, f);
  round_vec<T, S>(f[0]); round_vec<T, S>(f[1]);
  d1_seg<S, 2>(c, a.yt.d1[0], st, f);
  // ---- explicit part of the RK substep (rk_step_1/2, RK3_kernels.cu:76-80,138-142)
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    if (j < c.ny) {
      long re = (long)j * stp + md.e;
      T pyy = (T)f[0][r], gyy = (T)f[1][r];
      T p = phiR[re], g = gR[re];
      T rp_old = ((const T*)a.Rphi)[re], rg_old = ((const T*)a.Rg)[re];
      T rp_new = (T)rpn[0][r];
      T pw = p + dt * (a.alpha * a.nuT * (pyy - wv.kk * p) + a.dseda * rp_old);
      T gw = g + dt * (a.alpha * a.nuT * (gyy - wv.kk * g) + a.dseda * rg_old);
      pw = pw + dt * a.gammad * rp_new;
      gw = gw + dt * a.gammad * rgn[r];
      if (active) { ((T*)a.Rphi)[re] = rp_new; ((T*)a.Rg)[re] = rgn[r]; }
      f[0][r] = (double)pw; f[1][r] = (double)gw;
    } else { f[0][r] = 0.0; f[1][r] = 0.0; }
  }
  // ---- implicit solves (implicitSolver_double), shared pivots
  const double bdt = (double)(dt * a.betha);
  const double Dimp = a.nuD * bdt;
  var_seg<S, 2>(c, a.yt.d2[0], st.pivI, a.M, md.m, true, wv.kkd, Dimp, f);
  double pv[3][S];       // phi_new, v, dv
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    if (j < c.ny && active) ((T*)a.g)[(long)j * stp + md.e] = (T)f[1][r];
    pv[0][r] = rnd<T>(f[0][r]);
  }
  // ---- v = Helmholtz^-1 phi_new ; dv = d/dy v   (bilaplaSolver_double, bilplacSolver_double.cu:330-343)
  {
    double h[1][S];
#pragma unroll
    for (int r = 0; r < S; r++) h[0][r] = pv[0][r];
    var_seg<S, 1>(c, a.yt.d2[0], st.pivH, a.M, md.m, false, wv.kkd, 0.0, h);
#pragma unroll
    for (int r = 0; r < S; r++) { pv[1][r] = rnd<T>(h[0][r]); h[0][r] = pv[1][r]; }
    d1_seg<S, 1>(c, a.yt.d1[0], st, h);
#pragma unroll
    for (int r = 0; r < S; r++) pv[2][r] = rnd<T>(h[0][r]);
  }
  // ---- wall correction (setBoundaryCond): dv at the two walls -> C1, C2 -> phi, v, dv
  const bool correct = (double)dt > 1e-8;
  double c1 = 0.0, c2 = 0.0;
  BcCoef k;
  if (correct) {
    if (c.seg == 0) c.smW[c.tx] = pv[2][0];
    if (c.seg == c.nseg - 1) {
      double v = pv[2][0];
#pragma unroll
      for (int r = 1; r < S; r++) if (r < c.nrow) v = pv[2][r];
      c.smW[CB + c.tx] = v;
    }
    __syncthreads();
    const double dvm = c.smW[c.tx], dvp = c.smW[CB + c.tx];
    k = bc_coef(md.zero_mode ? 1.0 : wv.kkd, a.reD, (double)dt, (double)a.betha);
    BcFun p = bc_functions(k, 1.0, cc), q = bc_functions(k, -1.0, cc);
    double det = q.dv1 * p.dv2 - p.dv1 * q.dv2;
    c1 = p.dv2 * dvm - q.dv2 * dvp; c1 = -c1 / det;
    c2 = -p.dv1 * dvm + q.dv1 * dvp; c2 = -c2 / det;
  }
#pragma unroll
  for (int r = 0; r < S; r++) {
    int j = c.j0 + r;
    const int jc = j < c.ny ? j : c.ny - 1;
    double pp = pv[0][r], vv = pv[1][r], dd = pv[2][r];
    if (correct) {                                   // uniform: the shuffles inside stay converged
      BcFun fn = bc_functions(k, __ldg(a.yt.y + jc), cc);
      pp = pp + c1 * fn.ddv1 + c2 * fn.ddv2;
      dd = dd + c1 * fn.dv1 + c2 * fn.dv2;
      vv = vv + c1 * fn.v1 + c2 * fn.v2;
      if (md.zero_mode) { pp = 0.0; vv = 0.0; dd = 0.0; }
    }
    if (j < c.ny && active) {
      long re = (long)j * stp + md.e;
      ((T*)a.phi)[re] = (T)pp; ((T*)a.v)[re] = (T)vv; ((T*)a.dv)[re] = (T)dd;
    }
  }
}

// pivots of (A - kk M) [implicit = 0] or (M - D(A - kk M)) [implicit = 1], one thread per mode, rows in order
template <typename T>
__global__ void __launch_bounds__(128) pivot_kernel(PArgs<T> a, double* __restrict__ piv, int implicit) {
  long m = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= a.M) return;
  int ikx = a.ikx0 + (int)(m / a.Kz), kz = (int)(m % a.Kz);
  int kx = ikx <= a.kxmax ? ikx : ikx - a.Kx;
  double k1d = (double)a.k1fac * (double)kx, k3d = (double)a.k3fac * (double)kz, kk = k1d * k1d + k3d * k3d;
  double D = 0.0;
  if (implicit) { const T dt = (T)a.sc->dt; D = a.nuD * (double)(dt * a.betha); }
  const D2Tab& t = a.yt.d2[0];
  const int ny = a.ny;
  double cprev = 0.0;
  piv[m] = 1.0;
  for (int j = 1; j < ny - 1; j++) {
    double rP = __ldg(t.rP + j), rN = __ldg(t.rN + j), wP = __ldg(t.wP + j), wC = __ldg(t.wC + j), wN = __ldg(t.wN + j);
    double lo, di, up;
    if (implicit) { lo = rP - D * wP + D * kk * rP; di = 1.0 - D * wC + D * kk * 1.0; up = rN - D * wN + D * kk * rN; }
    else          { lo = wP - kk * rP;              di = wC - kk;                     up = wN - kk * rN; }
    if (j == 1) lo = 0.0;
    if (j == ny - 2) up = 0.0;
    double pv = 1.0 / (di - lo * cprev);
    cprev = up * pv;
    piv[(long)j * a.M + m] = pv;
  }
  piv[(long)(ny - 1) * a.M + m] = 1.0;
}

// ---------------------------------------------------------------------------------
// rows per segment: 8 up to NY = 512 (up to 64 segments of 8 chains per block above NY = 256), 12 up to 768.  The
// 16- and 24-row builds spill (ptxas: 1.4 KB per thread) and stay reachable through CHAN_TUNE_SEG_SHAPE only.
static int seg_len(int ny) { return ny <= 512 ? 8 : (ny <= 768 ? 12 : 0); }
int seg_supported(chan_handle* h) { return seg_len(h->d.ny) != 0 && h->d.ny >= 16; }

int seg_setup(chan_handle* h) {
  const int ny = h->d.ny;
  int S = seg_len(ny);
  if (h->tune.seg_shape > 0) S = h->tune.seg_shape / 100;          // experiment knob: S*100 + CB
  h->seg_S = S;
  // prefix products of the folded d/dy LU (direction 0) inside each segment
  std::vector<double> tab(21 * (size_t)ny), pf(ny), pb(ny);
  CHAN_CUDA(h, cudaMemcpy(tab.data(), h->ytab_store, sizeof(double) * tab.size(), cudaMemcpyDeviceToHost));
  const double* eL = &tab[(1 + 3) * (size_t)ny]; const double* cp = &tab[(1 + 4) * (size_t)ny];
  for (int s0 = 0; s0 < ny; s0 += S) {
    int s1 = s0 + S < ny ? s0 + S : ny;
    double p = 1.0; for (int j = s0; j < s1; j++) { p *= -eL[j]; pf[j] = p; }
    p = 1.0; for (int j = s1 - 1; j >= s0; j--) { p *= -cp[j]; pb[j] = p; }
  }
  CHAN_CUDA(h, cudaMalloc(&h->seg_pf, sizeof(double) * ny)); CHAN_CUDA(h, cudaMalloc(&h->seg_pb, sizeof(double) * ny));
  CHAN_CUDA(h, cudaMemcpy(h->seg_pf, pf.data(), sizeof(double) * ny, cudaMemcpyHostToDevice));
  CHAN_CUDA(h, cudaMemcpy(h->seg_pb, pb.data(), sizeof(double) * ny, cudaMemcpyHostToDevice));
  const size_t pbytes = (size_t)ny * h->d.M * sizeof(double);
  CHAN_CUDA(h, cudaMalloc(&h->pivH, pbytes)); CHAN_CUDA(h, cudaMalloc(&h->pivI, pbytes));
  h->pivH_ready = 0;
  return CHAN_OK;
}

// LB: launch bound = CB x the largest segment count the shape is used for
template <typename T, int S, int CB, int MINB = 1, int LB = CB * 32>
static int launch_seg(chan_handle* h, bool post, int n) {
  PArgs<T> a; fill_pencil_args<T>(h, a);
  static const float kAlpha[3] = { 29.0f / 96.0f, -3.0f / 40.0f, 1.0f / 6.0f }, kDseda[3] = { 0.0f, -17.0f / 60.0f, -5.0f / 12.0f };
  static const float kGamma[3] = { 8.0f / 15.0f, 5.0f / 12.0f, 3.0f / 4.0f }, kBetha[3] = { 37.0f / 160.0f, 5.0f / 24.0f, 1.0f / 6.0f };
  a.n = n; a.alpha = (T)kAlpha[n]; a.dseda = (T)kDseda[n]; a.gammad = (T)kGamma[n]; a.betha = (T)kBetha[n];
  SegTabs st; st.pf = h->seg_pf; st.pb = h->seg_pb; st.pivH = h->pivH; st.pivI = h->pivI;
  const int nseg = (h->d.ny + S - 1) / S;
  if (CB * nseg > LB) return set_err(h, CHAN_ERR_ARG, "segment shape %dx%d does not fit NY=%d", S, CB, h->d.ny);
  const long nchain = 2 * h->d.M;
  dim3 block(CB, nseg), grid((unsigned)((nchain + CB - 1) / CB));
  size_t sm = sizeof(double) * ((size_t)14 * nseg * CB + 2 * CB);
  if (post) {
    unsigned pg = (unsigned)((h->d.M + 127) / 128);
    if (!h->pivH_ready) { pivot_kernel<T><<<pg, 128, 0, h->stream>>>(a, h->pivH, 0); h->launches++; h->pivH_ready = 1; }
    pivot_kernel<T><<<pg, 128, 0, h->stream>>>(a, h->pivI, 1); h->launches++;
    CHAN_CUDA(h, cudaFuncSetAttribute(pencil_post_seg_kernel<T, S, CB, MINB, LB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    pencil_post_seg_kernel<T, S, CB, MINB, LB><<<grid, block, sm, h->stream>>>(a, st);
  } else {
    CHAN_CUDA(h, cudaFuncSetAttribute(pencil_pre_seg_kernel<T, S, CB, MINB, LB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    PeerOut po; fill_peer_out(h, po);
    pencil_pre_seg_kernel<T, S, CB, MINB, LB><<<grid, block, sm, h->stream>>>(a, st, po);
  }
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
template <typename T>
static int launch_seg_any(chan_handle* h, bool post, int n) {
  if (h->tune.seg_shape > 0) {
    switch (h->tune.seg_shape) {
      case 408: return launch_seg<T, 4, 8, 2>(h, post, n);
      case 409: return launch_seg<T, 4, 8, 3>(h, post, n);
      case 416: return launch_seg<T, 4, 16, 1>(h, post, n);
      case 808: return launch_seg<T, 8, 8, 2>(h, post, n);
      case 809: return launch_seg<T, 8, 8, 3>(h, post, n);
      case 804: return launch_seg<T, 8, 4, 4>(h, post, n);
      case 1616: return launch_seg<T, 16, 16>(h, post, n);
      case 2408: return launch_seg<T, 24, 8>(h, post, n);
      default: break;
    }
  }
  if (h->seg_S == 8) return h->d.ny <= 256 ? launch_seg<T, 8, 16>(h, post, n) : launch_seg<T, 8, 8, 1, 512>(h, post, n);
  return launch_seg<T, 12, 8, 1, 512>(h, post, n);
}
int launch_pencil_pre_seg(chan_handle* h) { return h->prec == 4 ? launch_seg_any<float>(h, false, 0) : launch_seg_any<double>(h, false, 0); }
int launch_pencil_post_seg(chan_handle* h, int n) { return h->prec == 4 ? launch_seg_any<float>(h, true, n) : launch_seg_any<double>(h, true, n); }

}  // namespace chan
