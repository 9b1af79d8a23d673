// Wall-normal stage, on-chip version: one WARP (or W warps) per (kx,kz) mode with the 32 lanes running ALONG y.
//
// Why.  The thread-per-chain kernels (kernels_pencil.cu) keep the forward-sweep intermediates of every Thomas solve in
// fp64 scratch in HBM: pencil_post moves 42 P of DRAM traffic for 13 P of algorithmic bytes and is bound by exactly
// that traffic (ncu, profiles/r01_round_end_ncu_kernels.txt); it also needs one thread per chain, which starves a GPU
// that holds few chains (many ranks).  Here a lane owns S consecutive rows of both components of its mode in
// REGISTERS, so nothing but the inputs, the outputs and one pivot table per kk-dependent matrix touches HBM
// (13 P + 1.5 P), and a mode is worked on by 32..96 threads.
//
// How a tridiagonal sweep runs across lanes.  Elimination and back substitution are first-order recurrences
// x_j = b_j + a_j x_{j-1}, i.e. affine maps of the value entering a lane's segment:
//   1. each lane sweeps its S rows with 0 entering and keeps the running products P_r = a_0..a_r (for the fixed d/dy
//      matrix they come from a table),
//   2. the per-lane maps (P_last, x_last) are combined by a Kogge-Stone scan over the warp with shuffles (and over
//      the W warps of a mode through a few shared-memory words and a named barrier),
//   3. each lane adds P_r * entering value to its rows.
// The matrices are real and identical for the re and im parts of a mode and for every right-hand side that shares
// them, so the coefficients, products and the scan of the `a` part are computed once per matrix for up to 4
// right-hand sides.  The Helmholtz / implicit pivots 1/(d_j - l_j c_{j-1}) are a non-linear recurrence: they come
// from a table [mode][y] written by pivot_lanes_kernel (Helmholtz: once per handle; implicit: once per substep, D =
// nu*beta*dt changes), 0.5 P of reads each.
//
// Global memory is plane-major (chan_internal.h): a block of MB modes reads / writes row segments of 2*MB reals
// through a TRANSPOSING staging tile in shared memory (column = one real chain, y contiguous, so that lanes along y
// read with stride S: S odd and a column pitch chosen per type keep both sides free of bank conflicts).
//
// Same arithmetic per row as the thread-per-chain kernels (same tables, same rounding points to the state type);
// results differ from theirs by the reassociation of the recurrences only (<= 1e-13 relative, tests/test_gpu_parity.py).
#include <vector>
#include <string.h>
#include "pencil_common.cuh"

namespace chan {

struct LaneTabs {                       // entry of row j = L*S + r lives at [r*NL + L] (NL = 32*W lanes along y): a warp reads
                                        // one contiguous piece per r; rows past NY hold zeros
  const double *eP, *eC, *eN, *eL, *cp; // folded LU of the compact d/dy matrix, upward sweep (D1Tab of dir 0)
  const double *rP, *rN, *wP, *wC, *wN; // second-derivative compact weights (D2Tab of dir 0)
  const double *y;
  const double *pivH, *pivI;            // [M][S*NL], same order inside a mode
  int NYL, NL, NYP, W, MB;              // padded rows S*NL, lanes along y, column pitch of the staging tile, warps per mode, modes per block
};

// ---- warp / mode-level plumbing -------------------------------------------------------------------------------
struct LCtx {
  int lane, ww, mi, W, L, j0, ny, NL;   // lane, warp within the mode, mode within the block, lanes-along-y index, first row
  double* xch;                          // this mode's exchange words in shared memory: [2 parities][W][XW]
  int par;                              // parity of the next exchange (see scan_maps)
};
enum { XW = 18 };                       // exchange words per warp and parity

__device__ __forceinline__ void mode_bar(const LCtx& c) {
  // the W warps of a mode meet on their own named barrier (ids 1..MB); one warp: nothing to wait for
  if (c.W > 1) asm volatile("bar.sync %0, %1;" ::"r"(1 + c.mi), "r"(c.W * 32) : "memory");
  else __syncwarp();
}
__device__ __forceinline__ double shfl_up_d(double v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
__device__ __forceinline__ double shfl_dn_d(double v, int d) { return __shfl_down_sync(0xffffffffu, v, d); }

// Value entering each lane's segment of the recurrence x = Y + A * x_in, lanes ascending (UP) or descending.
// In: this lane's map (A, Y[q]).  Out: xin[q].
// Kogge-Stone over the warp, but only as deep as the matrix needs: after k doubling steps a lane has composed the maps
// of the 2^k lanes before it, and what it leaves out is scaled by a product of 2^k factors A.  With amax = the largest
// |A| of the warp, k is the first depth with amax^(2^k) < 2^-54 (half an ulp of the result): 3 steps for the d/dy matrix
// (|A| ~ 0.27^S), 1-3 for the implicit and the high-wavenumber Helmholtz matrices, all 5 for smooth Helmholtz modes.
// Several warps per mode exchange their totals through shared memory; the words alternate between two parities so
// that one barrier per scan is enough (a warp cannot be two exchanges ahead of another one of its mode).
template <int NR, bool UP>
__device__ __forceinline__ void scan_maps(LCtx& c, double A, const double (&Yin)[NR], double (&xin)[NR]) {
  static_assert(1 + NR <= XW, "exchange words");
  double Y[NR];
#pragma unroll
  for (int q = 0; q < NR; q++) Y[q] = Yin[q];
  int depth = 5;
  {
    const unsigned mh = __reduce_max_sync(0xffffffffu, (unsigned)__double2hiint(fabs(A)));
    const int lb = 1022 - (int)((mh >> 20) & 0x7ffu);      // amax < 2^-lb
    if (lb > 0) { depth = 0; while ((lb << depth) < 54 && depth < 5) depth++; }
  }
#pragma unroll
  for (int st = 0; st < 5; st++) {
    if (st < depth) {
      const int off = 1 << st;
      const bool on = UP ? (c.lane >= off) : (c.lane + off < 32);
      const double Ap = UP ? shfl_up_d(A, off) : shfl_dn_d(A, off);
      const double Aon = on ? A : 0.0;
#pragma unroll
      for (int q = 0; q < NR; q++) {
        const double Yp = UP ? shfl_up_d(Y[q], off) : shfl_dn_d(Y[q], off);
        Y[q] = fma(Aon, Yp, Y[q]);
      }
      if (on) A *= Ap;
    }
  }
  // exclusive: the composed map of the lanes before this one
  const bool first = UP ? (c.lane == 0) : (c.lane == 31);
  double Ae = UP ? shfl_up_d(A, 1) : shfl_dn_d(A, 1);
  if (first) Ae = 1.0;
#pragma unroll
  for (int q = 0; q < NR; q++) { double Ye = UP ? shfl_up_d(Y[q], 1) : shfl_dn_d(Y[q], 1); xin[q] = first ? 0.0 : Ye; }
  if (c.W > 1) {
    double* slot = c.xch + (size_t)c.par * c.W * XW;
    const bool last = UP ? (c.lane == 31) : (c.lane == 0);
    if (last) {
      slot[c.ww * XW] = A;
#pragma unroll
      for (int q = 0; q < NR; q++) slot[c.ww * XW + 1 + q] = Y[q];
    }
    mode_bar(c);
    double carry[NR];
#pragma unroll
    for (int q = 0; q < NR; q++) carry[q] = 0.0;
    if (UP) { for (int w = 0; w < c.ww; w++) {
#pragma unroll
        for (int q = 0; q < NR; q++) carry[q] = fma(slot[w * XW], carry[q], slot[w * XW + 1 + q]); } }
    else { for (int w = c.W - 1; w > c.ww; w--) {
#pragma unroll
        for (int q = 0; q < NR; q++) carry[q] = fma(slot[w * XW], carry[q], slot[w * XW + 1 + q]); } }
#pragma unroll
    for (int q = 0; q < NR; q++) xin[q] = fma(Ae, carry[q], xin[q]);
    c.par ^= 1;
  }
}

// rows just outside the lane's segment: hm1 = f(j0-1), hm2 = f(j0-2), hp1 = f(j0+S)  (0 outside the grid)
template <int S, int NR>
__device__ __forceinline__ void halos(LCtx& c, const double (&f)[NR][S], double (&hm1)[NR], double (&hm2)[NR], double (&hp1)[NR]) {
  static_assert(3 * NR <= XW, "exchange words");
#pragma unroll
  for (int q = 0; q < NR; q++) {
    hm1[q] = shfl_up_d(f[q][S - 1], 1); hm2[q] = shfl_up_d(f[q][S - 2], 1); hp1[q] = shfl_dn_d(f[q][0], 1);
  }
  if (c.W > 1) {
    double* slot = c.xch + (size_t)c.par * c.W * XW + c.ww * XW;
    if (c.lane == 31) {
#pragma unroll
      for (int q = 0; q < NR; q++) { slot[q] = f[q][S - 1]; slot[NR + q] = f[q][S - 2]; }
    }
    if (c.lane == 0) {
#pragma unroll
      for (int q = 0; q < NR; q++) slot[2 * NR + q] = f[q][0];
    }
    mode_bar(c);
    if (c.lane == 0 && c.ww > 0) {
#pragma unroll
      for (int q = 0; q < NR; q++) { hm1[q] = slot[-XW + q]; hm2[q] = slot[-XW + NR + q]; }
    }
    if (c.lane == 31 && c.ww + 1 < c.W) {
#pragma unroll
      for (int q = 0; q < NR; q++) hp1[q] = slot[XW + 2 * NR + q];
    }
    c.par ^= 1;
  }
  if (c.L == 0) {
#pragma unroll
    for (int q = 0; q < NR; q++) { hm1[q] = 0.0; hm2[q] = 0.0; }
  }
  if (c.L == c.NL - 1) {
#pragma unroll
    for (int q = 0; q < NR; q++) hp1[q] = 0.0;
  }
}

// f <- d/dy f for NR chains (deriv_Y_HO_double, derivatives_nu_double.cu:390-415), fp64, in place
template <int S, int NR>
__device__ __forceinline__ void d1_lanes(LCtx& c, const LaneTabs& t, double (&f)[NR][S]) {
  double hm1[NR], hm2[NR], hp1[NR];
  halos<S, NR>(c, f, hm1, hm2, hp1);
  const int ny = c.ny, NL = c.NL;
  double yprev[NR], xin[NR], P[S];
#pragma unroll
  for (int q = 0; q < NR; q++) yprev[q] = 0.0;
  // elimination: y_j = eP f_(j-1) + eC f_j + eN f_(j+1) - eL y_(j-1); the one-sided wall rows (:85-113) take
  // f_(j+2) in place of f_(j-1) at j = 0 and f_(j-2) in place of f_(j+1) at j = NY-1
  double p = 1.0;
#pragma unroll
  for (int r = 0; r < S; r++) {
    const int j = c.j0 + r, ti = r * NL + c.L;
    const double eP = __ldg(t.eP + ti), eC = __ldg(t.eC + ti), eN = __ldg(t.eN + ti), eL = __ldg(t.eL + ti);
    p *= -eL; P[r] = p;
#pragma unroll
    for (int q = 0; q < NR; q++) {
      const double f0 = f[q][r];
      const double fp1 = r + 1 < S ? f[q][r + 1] : hp1[q];
      const double fprev = (j == 0) ? f[q][r + 2 < S ? r + 2 : S - 1] : hm1[q];
      const double fnext = (j == ny - 1) ? hm2[q] : fp1;
      const double y = fma(eN, fnext, fma(eC, f0, eP * fprev)) - eL * yprev[q];
      hm2[q] = hm1[q]; hm1[q] = f0;
      f[q][r] = y; yprev[q] = y;
    }
  }
  scan_maps<NR, true>(c, p, yprev, xin);
#pragma unroll
  for (int r = 0; r < S; r++) {
#pragma unroll
    for (int q = 0; q < NR; q++) f[q][r] = fma(P[r], xin[q], f[q][r]);
  }
  // back substitution: x_j = y_j - cp_j x_(j+1)
#pragma unroll
  for (int q = 0; q < NR; q++) yprev[q] = 0.0;
  p = 1.0;
#pragma unroll
  for (int r = S - 1; r >= 0; r--) {
    const double cp = __ldg(t.cp + r * NL + c.L);
    p *= -cp; P[r] = p;
#pragma unroll
    for (int q = 0; q < NR; q++) { yprev[q] = fma(-cp, yprev[q], f[q][r]); f[q][r] = yprev[q]; }
  }
  scan_maps<NR, false>(c, p, yprev, xin);
#pragma unroll
  for (int r = 0; r < S; r++) {
#pragma unroll
    for (int q = 0; q < NR; q++) f[q][r] = fma(P[r], xin[q], f[q][r]);
  }
}

// f <- solution of the Helmholtz (A - kk M) x = M f or implicit (M - D(A - kk M)) x = M f system for NR chains of one
// mode (hemholzt_nu_double.cu:119-185, implicitStep_nu_double.cu:119-185); piv = this mode's piece of the pivot table
template <int S, int NR>
__device__ __forceinline__ void var_lanes(LCtx& c, const LaneTabs& t, const double* __restrict__ piv, bool implicit, double kk, double D,
                                          double (&f)[NR][S]) {
  double hm1[NR], hm2[NR], hp1[NR];
  halos<S, NR>(c, f, hm1, hm2, hp1);
  const int ny = c.ny, NL = c.NL;
  double cpv[S], P[S], yprev[NR], xin[NR];
#pragma unroll
  for (int q = 0; q < NR; q++) yprev[q] = 0.0;
  double p = 1.0;
#pragma unroll
  for (int r = 0; r < S; r++) {
    const int j = c.j0 + r, ti = r * NL + c.L;
    const double rP = __ldg(t.rP + ti), rN = __ldg(t.rN + ti), wP = __ldg(t.wP + ti), wN = __ldg(t.wN + ti);
    const double pv = __ldg(piv + ti);                    // 0 on the wall rows and on the padding: x = 0 there
    double lo, up;
    if (implicit) { lo = rP - D * wP + D * kk * rP; up = rN - D * wN + D * kk * rN; }
    else          { lo = wP - kk * rP;              up = wN - kk * rN; }
    if (j == 1) lo = 0.0;
    if (j == ny - 2) up = 0.0;
    const double a = -lo * pv;
    cpv[r] = up * pv;
    p *= a; P[r] = p;
#pragma unroll
    for (int q = 0; q < NR; q++) {
      const double f0 = f[q][r];
      const double fp1 = r + 1 < S ? f[q][r + 1] : hp1[q];
      const double rhs = fma(rP, hm1[q], fma(rN, fp1, f0));
      const double y = fma(a, yprev[q], rhs * pv);
      hm1[q] = f0; f[q][r] = y; yprev[q] = y;
    }
  }
  scan_maps<NR, true>(c, p, yprev, xin);
#pragma unroll
  for (int r = 0; r < S; r++) {
#pragma unroll
    for (int q = 0; q < NR; q++) f[q][r] = fma(P[r], xin[q], f[q][r]);
  }
  p = 1.0;
#pragma unroll
  for (int q = 0; q < NR; q++) yprev[q] = 0.0;
#pragma unroll
  for (int r = S - 1; r >= 0; r--) {
    p *= -cpv[r]; P[r] = p;
#pragma unroll
    for (int q = 0; q < NR; q++) { yprev[q] = fma(-cpv[r], yprev[q], f[q][r]); f[q][r] = yprev[q]; }
  }
  scan_maps<NR, false>(c, p, yprev, xin);
#pragma unroll
  for (int r = 0; r < S; r++) {
#pragma unroll
    for (int q = 0; q < NR; q++) f[q][r] = fma(P[r], xin[q], f[q][r]);
  }
}

// the six functions of setBoundaryCond (bilplacSolver_double.cu:174-217) at one y; same expressions as bc_functions
// (pencil_common.cuh), all four exponentials evaluated here (re and im of a mode share the result)
// exp / expf of arguments far below their underflow thresholds (-745.2 / -103.97) return exactly 0: skip the call there.
// l1 = sqrt(kk + Re/(dt beta)) is in the thousands, so away from the walls a whole warp takes the short way.
__device__ __forceinline__ double exp_or_zero(double x) { return x < -760.0 ? 0.0 : exp(x); }
__device__ __forceinline__ double expf_or_zero(double x) { return x < -110.0 ? 0.0 : (double)expf((float)x); }
__device__ __forceinline__ BcFun bc_functions_all(const BcCoef& k, double y) {
  const double E1m = expf_or_zero(-k.l1 * (1.0 - y)), E1p = exp_or_zero(-k.l1 * (1.0 + y));
  const double E2m = expf_or_zero(-k.l2 * (1.0 - y)), E2p = exp_or_zero(-k.l2 * (1.0 + y));
  const double a1 = (E1m + E1p) * k.ip1, b1 = (E1m - E1p) * k.im1, a1d = (E1m - E1p) * k.ip1, b1d = (E1m + E1p) * k.im1;
  const double a2 = (E2m + E1p) * k.ip2, b2 = (E2m - E2p) * k.im2, a2d = (E2m - E1p) * k.ip2, b2d = (E2m + E2p) * k.im2;
  const double hD = 0.5 * k.D;
  BcFun r;
  r.v1 = hD * (a1 - b1) + hD * (-a2 + b2);
  r.v2 = hD * (a1 + b1) + hD * (-a2 - b2);
  r.dv1 = hD * k.l1 * (a1d - b1d) + hD * k.l2 * (-a2d + b2d);
  r.dv2 = hD * k.l1 * (a1d + b1d) + hD * k.l2 * (-a2d - b2d);
  r.ddv1 = 0.5 * (a1 - b1);
  r.ddv2 = 0.5 * (a1 + b1);
  return r;
}

// ---- staging: global rows (2*MB reals of the block's modes per y) <-> shared columns [chain][NYP] -----------------
// Column position of row j: j itself when S is odd, j + j/S (one pad element per lane segment) when S is even, so that
// lanes along y (segments S or S+1 elements apart) always fall into distinct banks.  Columns hold NY rows; lanes past
// the last row never touch them.  The loads are asynchronous copies (cp.async, one element each — the transposition is
// in the destination address): every thread issues all of its copies back to back and the block waits once.
template <int S> __device__ __forceinline__ int cpos(int j) { return (S & 1) ? j : j + j / S; }
template <int S> __device__ __forceinline__ int cseg(int L) { return L * ((S & 1) ? S : S + 1); }

__device__ __forceinline__ void cp_async_el(double* dst, const double* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_el(float* dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// a thread keeps one column (the block size is a multiple of the column count) and walks down the rows
template <typename T, int S, int NA>
__device__ __forceinline__ void stage_in(T* const (&tile)[NA], const T* const (&src)[NA], long st, long e0, long elim, int lgcol, int ny, int NYP) {
  const int cc = threadIdx.x & ((1 << lgcol) - 1), dj = blockDim.x >> lgcol;
  int j = threadIdx.x >> lgcol;
  const int cb = cc * NYP;
  if (e0 + cc < elim) {
    long go = (long)j * st + e0 + cc; const long dgo = (long)dj * st;
    for (; j < ny; j += dj, go += dgo) {
      const int so = cb + cpos<S>(j);
#pragma unroll
      for (int a = 0; a < NA; a++) cp_async_el(tile[a] + so, src[a] + go);
    }
  } else {
    for (; j < ny; j += dj) {
      const int so = cb + cpos<S>(j);
#pragma unroll
      for (int a = 0; a < NA; a++) tile[a][so] = (T)0;
    }
  }
}
template <typename T, int S, int NA>
__device__ __forceinline__ void stage_out(T* const (&dst)[NA], const T* const (&tile)[NA], long st, long e0, long elim, int lgcol, int ny, int NYP) {
  const int cc = threadIdx.x & ((1 << lgcol) - 1), dj = blockDim.x >> lgcol;
  int j = threadIdx.x >> lgcol;
  const int cb = cc * NYP;
  if (e0 + cc >= elim) return;
  long go = (long)j * st + e0 + cc; const long dgo = (long)dj * st;
  for (; j < ny; j += dj, go += dgo) {
    const int so = cb + cpos<S>(j);
    T v[NA];
#pragma unroll
    for (int a = 0; a < NA; a++) v[a] = tile[a][so];
#pragma unroll
    for (int a = 0; a < NA; a++) dst[a][go] = v[a];
  }
}

extern __shared__ __align__(16) unsigned char lanes_smem[];

// =====================================================================================================================
// pencil_post: H1..3, phi, g, R_phi, R_g -> phi, g, v, dv, R_phi, R_g, in the reference's own order (RK3.c:131-166):
// RKstep_1 (needs phi'', g'' and the old R), calcHvg, RKstep_2, implicit solves, Helmholtz, d/dy, wall correction.
// =====================================================================================================================
template <typename T, int S>
__global__ void __launch_bounds__(512, 1) pencil_post_lanes_kernel(PArgs<T> a, LaneTabs t) {
  const int W = t.W, MB = t.MB, NYP = t.NYP, NYL = t.NYL, ny = a.ny;
  const int ncol = 2 * MB, lgcol = 31 - __clz(ncol);
  T* const tl = (T*)lanes_smem;                                  // 7 tiles of ncol columns
  const size_t tsz = (size_t)ncol * NYP;
  T* const tPhi = tl; T* const tG = tl + tsz; T* const tRp = tl + 2 * tsz; T* const tRg = tl + 3 * tsz;
  T* const tH1 = tl + 4 * tsz; T* const tH2 = tl + 5 * tsz; T* const tH3 = tl + 6 * tsz;
  double* const xch_all = (double*)(tl + 7 * tsz + (((7 * tsz) & 1) ? 1 : 0));   // 8-byte aligned for T = float as well
  const long st = 2 * a.M, e0 = (long)blockIdx.x * ncol, elim = 2 * a.M;
  {
    T* const tA[7] = { tPhi, tG, tH1, tH3, tRp, tRg, tH2 };
    const T* const sA[7] = { (const T*)a.phi, (const T*)a.g, (const T*)a.ox, (const T*)a.oz, (const T*)a.Rphi, (const T*)a.Rg, (const T*)a.oy };
    stage_in<T, S, 7>(tA, sA, st, e0, elim, lgcol, ny, NYP);
    cp_async_commit();
  }
  LCtx c;
  const int warp = threadIdx.x >> 5;
  c.lane = threadIdx.x & 31; c.W = W; c.mi = warp / W; c.ww = warp - c.mi * W; c.L = c.ww * 32 + c.lane; c.j0 = c.L * S; c.ny = ny; c.NL = t.NL;
  c.xch = xch_all + (size_t)c.mi * 2 * W * XW; c.par = 0;
  long m = (long)blockIdx.x * MB + c.mi;
  if (m >= a.M) m = a.M - 1;                                      // idle warps shadow the last mode on zero columns; stage_out skips their chains
  // mode constants (setup_chain of pencil_common.cuh, per mode instead of per chain)
  const int ikx = a.ikx0 + (int)(m / a.Kz), kzi = (int)(m % a.Kz);
  const int kxi = ikx <= a.kxmax ? ikx : ikx - a.Kx;
  const bool zero_mode = (kxi == 0 && kzi == 0);
  const T k1 = a.k1fac * (T)kxi, k3 = a.k3fac * (T)kzi, kk = k1 * k1 + k3 * k3;
  const double k1d = (double)a.k1fac * (double)kxi, k3d = (double)a.k3fac * (double)kzi, kkd = k1d * k1d + k3d * k3d;
  const T dt = (T)a.sc->dt;
  const int col0 = (2 * c.mi) * NYP + cseg<S>(c.L), col1 = col0 + NYP;    // this lane's rows of the re / im columns
  bool rowok[S];
#pragma unroll
  for (int r = 0; r < S; r++) rowok[r] = c.j0 + r < ny;
  cp_async_wait<0>();
  __syncthreads();
  // ---- the three first derivatives that need nothing but the inputs, in one sweep: phi, g (RKstep_1, RK3_kernels.cu:155-177)
  // and hv = -i(k1 H1 + k3 H3) (calcHvg, nonLinear_kernels.cu:112-133)
  double f[4][S], hv[2][S];                                       // phi.re phi.im g.re g.im
  {
    double f6[6][S];
#pragma unroll
    for (int r = 0; r < S; r++) {
      f6[0][r] = rowok[r] ? (double)tPhi[col0 + r] : 0.0; f6[1][r] = rowok[r] ? (double)tPhi[col1 + r] : 0.0;
      f6[2][r] = rowok[r] ? (double)tG[col0 + r] : 0.0;   f6[3][r] = rowok[r] ? (double)tG[col1 + r] : 0.0;
#pragma unroll
      for (int cc = 0; cc < 2; cc++) {
        const T s = cc ? (T)1 : (T)-1;                            // own component of i*z: -im for re, +re for im
        const int oo = (cc ? col0 : col1) + r;                    // the OTHER component
        const T h1p = rowok[r] ? tH1[oo] : (T)0, h3p = rowok[r] ? tH3[oo] : (T)0;
        f6[4 + cc][r] = (double)(T)(-s * (k1 * h1p + k3 * h3p));
      }
    }
    d1_lanes<S, 6>(c, t, f6);
#pragma unroll
    for (int r = 0; r < S; r++) {
#pragma unroll
      for (int q = 0; q < 4; q++) f[q][r] = rnd<T>(f6[q][r]);     // second derivative = d/dy twice with the cast in between (Q4)
      hv[0][r] = f6[4][r]; hv[1][r] = f6[5][r];
    }
  }
  d1_lanes<S, 4>(c, t, f);
#pragma unroll
  for (int r = 0; r < S; r++) {
#pragma unroll
    for (int cc = 0; cc < 2; cc++) {
      const int o = (cc ? col1 : col0) + r, oo = (cc ? col0 : col1) + r;
      const T z = (T)0, s = cc ? (T)1 : (T)-1;
      const T p = rowok[r] ? tPhi[o] : z, g = rowok[r] ? tG[o] : z, rp = rowok[r] ? tRp[o] : z, rg = rowok[r] ? tRg[o] : z;
      const T h1p = rowok[r] ? tH1[oo] : z, h3p = rowok[r] ? tH3[oo] : z, h2 = rowok[r] ? tH2[o] : z;
      const T pyy = (T)f[cc][r], gyy = (T)f[2 + cc][r];
      T pw = p + dt * (a.alpha * a.nuT * (pyy - kk * p) + a.dseda * rp);        // rk_step_1 (RK3_kernels.cu:76-80)
      T gw = g + dt * (a.alpha * a.nuT * (gyy - kk * g) + a.dseda * rg);
      const T rp_new = (T)hv[cc][r] - kk * h2;                                  // calcHvvkernel (nonLinear_kernels.cu:180-181)
      const T rg_new = s * (k3 * h1p - k1 * h3p);                               // h_g = i(k3 H1 - k1 H3) (:132-133)
      pw = pw + dt * a.gammad * rp_new;                                         // rk_step_2 (RK3_kernels.cu:138-142)
      gw = gw + dt * a.gammad * rg_new;
      if (rowok[r]) { tRp[o] = rp_new; tRg[o] = rg_new; }
      f[cc][r] = (double)pw; f[2 + cc][r] = (double)gw;
    }
  }
  // ---- implicitSolver_double on g_w, and on phi_w inside bilaplaSolver_double: one matrix, four right-hand sides
  const double bdt = (double)(dt * a.betha);                      // dt*betha in the state type (implicitStep_nu_double.cu:235)
  const double Dimp = a.nuD * bdt;
  var_lanes<S, 4>(c, t, t.pivI + (size_t)m * NYL, true, kkd, Dimp, f);
  double pv[2][S], vv[2][S], dd[2][S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    if (rowok[r]) { tG[col0 + r] = (T)f[2][r]; tG[col1 + r] = (T)f[3][r]; }
    pv[0][r] = rnd<T>(f[0][r]); pv[1][r] = rnd<T>(f[1][r]);
    vv[0][r] = pv[0][r]; vv[1][r] = pv[1][r];
  }
  // ---- v = Helmholtz^-1 phi, dv = d/dy v (bilplacSolver_double.cu:330-343)
  var_lanes<S, 2>(c, t, t.pivH + (size_t)m * NYL, false, kkd, 0.0, vv);
#pragma unroll
  for (int r = 0; r < S; r++) { vv[0][r] = rnd<T>(vv[0][r]); vv[1][r] = rnd<T>(vv[1][r]); dd[0][r] = vv[0][r]; dd[1][r] = vv[1][r]; }
  d1_lanes<S, 2>(c, t, dd);
#pragma unroll
  for (int r = 0; r < S; r++) { dd[0][r] = rnd<T>(dd[0][r]); dd[1][r] = rnd<T>(dd[1][r]); }
  // ---- wall correction (setBoundaryCond, bilplacSolver_double.cu:6-296); skipped when dt <= 1e-8 (:314, Q6)
  if ((double)dt > 1e-8) {
    // dv at the two walls, every lane of the mode needs them
    double* slot = c.xch + (size_t)c.par * W * XW;
    const int jl = ny - 1, Ll = jl / S, rl = jl - Ll * S;
    if (c.L == 0) { slot[0] = dd[0][0]; slot[1] = dd[1][0]; }
    if (c.L == Ll) {
#pragma unroll
      for (int r = 0; r < S; r++) if (r == rl) { slot[2] = dd[0][r]; slot[3] = dd[1][r]; }
    }
    mode_bar(c);
    const double dvm[2] = { slot[0], slot[1] }, dvp[2] = { slot[2], slot[3] };
    c.par ^= 1;
    if (c.j0 < ny) {                                              // lanes past the last row have nothing to correct
    const BcCoef k = bc_coef(zero_mode ? 1.0 : kkd, a.reD, (double)dt, (double)a.betha);
    const BcFun p = bc_functions_all(k, 1.0), q = bc_functions_all(k, -1.0);
    const double det = q.dv1 * p.dv2 - p.dv1 * q.dv2;
    double c1[2], c2[2];
#pragma unroll
    for (int cc = 0; cc < 2; cc++) {
      c1[cc] = p.dv2 * dvm[cc] - q.dv2 * dvp[cc]; c1[cc] = -c1[cc] / det;
      c2[cc] = -p.dv1 * dvm[cc] + q.dv1 * dvp[cc]; c2[cc] = -c2[cc] / det;
    }
#pragma unroll
    for (int r = 0; r < S; r++) {
      const BcFun fn = bc_functions_all(k, __ldg(t.y + r * t.NL + c.L));
#pragma unroll
      for (int cc = 0; cc < 2; cc++) {
        pv[cc][r] = pv[cc][r] + c1[cc] * fn.ddv1 + c2[cc] * fn.ddv2;
        dd[cc][r] = dd[cc][r] + c1[cc] * fn.dv1 + c2[cc] * fn.dv2;
        vv[cc][r] = vv[cc][r] + c1[cc] * fn.v1 + c2[cc] * fn.v2;
        if (zero_mode) { pv[cc][r] = 0.0; vv[cc][r] = 0.0; dd[cc][r] = 0.0; }        // (:252-262)
      }
    }
    }
  }
#pragma unroll
  for (int r = 0; r < S; r++) {
    if (rowok[r]) {
      tPhi[col0 + r] = (T)pv[0][r]; tPhi[col1 + r] = (T)pv[1][r];
      tH1[col0 + r] = (T)vv[0][r]; tH1[col1 + r] = (T)vv[1][r];       // v and dv leave through the tiles of H1 and H3
      tH3[col0 + r] = (T)dd[0][r]; tH3[col1 + r] = (T)dd[1][r];
    }
  }
  __syncthreads();
  {
    T* const dsts[6] = { (T*)a.phi, (T*)a.g, (T*)a.v, (T*)a.dv, (T*)a.Rphi, (T*)a.Rg };
    const T* const tiles[6] = { tPhi, tG, tH1, tH3, tRp, tRg };
    stage_out<T, S, 6>(dsts, tiles, st, e0, elim, lgcol, ny, NYP);
  }
}

// =====================================================================================================================
// pencil_pre: (dv, g, v, <u>) -> u, w, omega_x, omega_y, omega_z: calcUW (nonLinear_kernels.cu:8-92) + writeUmean
// (meanUevol.c:117-125) + the two deriv_Y of convolution.c:7-11 + readUtau (:14-16) + calcOmega
// (convolution_kernels.cu:7-66).  With several GPUs the outputs are stored locally and a bulk peer copy delivers them
// (launch_peer_scatter); with CHAN_TUNE_PRE_DIRECT=1 the six plane-stage inputs of row j leave the staging tiles straight into
// the receive buffers of the rank that owns plane j instead (NVLink stores of 64-byte row pieces: measured slower).
// =====================================================================================================================
template <typename T, int S>
__global__ void __launch_bounds__(512, 1) pencil_pre_lanes_kernel(PArgs<T> a, LaneTabs t, PeerOut po) {
  typedef typename Vec2<T>::type T2;
  const int W = t.W, MB = t.MB, NYP = t.NYP, ny = a.ny;
  const int ncol = 2 * MB, lgcol = 31 - __clz(ncol);
  T* const tl = (T*)lanes_smem;                                  // 6 tiles: dv -> u, g -> w, v, omega_x, omega_y, omega_z
  const size_t tsz = (size_t)ncol * NYP;
  T* const tU = tl; T* const tW = tl + tsz; T* const tV = tl + 2 * tsz; T* const tOx = tl + 3 * tsz; T* const tOy = tl + 4 * tsz; T* const tOz = tl + 5 * tsz;
  double* const xch_all = (double*)(tl + 7 * tsz + (((7 * tsz) & 1) ? 1 : 0));   // same carve-up as pencil_post
  const long st = 2 * a.M, e0 = (long)blockIdx.x * ncol, elim = 2 * a.M;
  {
    T* const tA[3] = { tU, tW, tV };
    const T* const sA[3] = { (const T*)a.dv, (const T*)a.g, (const T*)a.v };
    stage_in<T, S, 3>(tA, sA, st, e0, elim, lgcol, ny, NYP);
    cp_async_commit();
  }
  LCtx c;
  const int warp = threadIdx.x >> 5;
  c.lane = threadIdx.x & 31; c.W = W; c.mi = warp / W; c.ww = warp - c.mi * W; c.L = c.ww * 32 + c.lane; c.j0 = c.L * S; c.ny = ny; c.NL = t.NL;
  c.xch = xch_all + (size_t)c.mi * 2 * W * XW; c.par = 0;
  long m = (long)blockIdx.x * MB + c.mi;
  const bool live = m < a.M;
  if (!live) m = a.M - 1;
  const int ikx = a.ikx0 + (int)(m / a.Kz), kzi = (int)(m % a.Kz);
  const int kxi = ikx <= a.kxmax ? ikx : ikx - a.Kx;
  const bool zero_mode = (kxi == 0 && kzi == 0);
  const T k1 = a.k1fac * (T)kxi, k3 = a.k3fac * (T)kzi, kk = k1 * k1 + k3 * k3;
  const T rkk = zero_mode ? (T)1 : (T)1 / kk;
  const bool with_mean = zero_mode && a.use_mean;
  const int col0 = (2 * c.mi) * NYP + cseg<S>(c.L), col1 = col0 + NYP;
  bool rowok[S];
#pragma unroll
  for (int r = 0; r < S; r++) rowok[r] = c.j0 + r < ny;
  cp_async_wait<0>();
  __syncthreads();
  // u = i (k1 dv - k3 g)/kk, w = i (k3 dv + k1 g)/kk (uw_comp of pencil_common.cuh, both components of the mode)
  T u[2][S], w[2][S];
  double f[4][S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    const T z = (T)0;
    const T dre = rowok[r] ? tU[col0 + r] : z, dim = rowok[r] ? tU[col1 + r] : z, gre = rowok[r] ? tW[col0 + r] : z, gim = rowok[r] ? tW[col1 + r] : z;
#pragma unroll
    for (int cc = 0; cc < 2; cc++) {
      const T s = cc ? (T)1 : (T)-1, dp = cc ? dre : dim, gp = cc ? gre : gim;      // the other component
      if (sizeof(T) == 8) { u[cc][r] = s * (k1 * dp - k3 * gp) * rkk; w[cc][r] = s * (k3 * dp + k1 * gp) * rkk; }
      else { const T kd = zero_mode ? (T)1 : kk; u[cc][r] = s * (k1 * dp - k3 * gp) / kd; w[cc][r] = s * (k3 * dp + k1 * gp) / kd; }
      if (zero_mode) { u[cc][r] = 0; w[cc][r] = 0; }
    }
    if (with_mean) { u[0][r] = rowok[r] ? a.mean_u[c.j0 + r] : z; u[1][r] = 0; }    // writeUmean (nonLinear.c:19-21)
    f[0][r] = (double)u[0][r]; f[1][r] = (double)u[1][r]; f[2][r] = (double)w[0][r]; f[3][r] = (double)w[1][r];
  }
  d1_lanes<S, 4>(c, t, f);
#pragma unroll
  for (int r = 0; r < S; r++) {
    const T z = (T)0;
    const T vre = rowok[r] ? tV[col0 + r] : z, vim = rowok[r] ? tV[col1 + r] : z;
#pragma unroll
    for (int cc = 0; cc < 2; cc++) {
      const T s = cc ? (T)1 : (T)-1;
      const T dudy = (T)f[cc][r], dwdy = (T)f[2 + cc][r];
      const T vp = cc ? vre : vim, up = u[cc ^ 1][r], wp = w[cc ^ 1][r];
      if (rowok[r]) {
        const int o = (cc ? col1 : col0) + r;
        tOx[o] = dwdy - s * k3 * vp;                    // omega_x = dw/dy - i k3 v   (calcOmegakernel, convolution_kernels.cu:46-53)
        tOy[o] = s * (k3 * up - k1 * wp);               // omega_y = i (k3 u - k1 w)
        tOz[o] = -dudy + s * k1 * vp;                   // omega_z = -du/dy + i k1 v
        tU[o] = u[cc][r]; tW[o] = w[cc][r];
      }
      if (zero_mode && cc == 0 && a.sc && live) {        // readUtau (convolution.c:14-16)
        if (c.j0 + r == 0) a.sc->dudy_wall[0] = (double)dudy;
        if (c.j0 + r == ny - 1) a.sc->dudy_wall[1] = (double)dudy;
      }
    }
  }
  __syncthreads();
  // a thread keeps one column and walks down the rows (stage_out), choosing the destination per row
  const int cc = threadIdx.x & (ncol - 1), dj = blockDim.x >> lgcol;
  if (e0 + cc >= elim) return;
  const int cb = cc * NYP;
  const bool omega_only = a.op == OP_OMEGA;
  for (int j = threadIdx.x >> lgcol; j < ny; j += dj) {
    const int so = cb + cpos<S>(j);
    const T vu = tU[so], vw = tW[so], vv = tV[so], v1 = tOx[so], v2 = tOy[so], v3 = tOz[so];
    if (po.nranks > 1) {
      int p = 0;
#pragma unroll
      for (int i = 1; i < 8; i++) p += (i < po.nranks && j >= po.part_j[i]) ? 1 : 0;
      const long idx = 2 * (po.src_off[p] + (long)(j - po.part_j[p]) * a.M) + e0 + cc;
      ((T*)po.rcv[p][0])[idx] = vu; ((T*)po.rcv[p][1])[idx] = vv; ((T*)po.rcv[p][2])[idx] = vw;
      ((T*)po.rcv[p][3])[idx] = v1; ((T*)po.rcv[p][4])[idx] = v2; ((T*)po.rcv[p][5])[idx] = v3;
    } else {
      const long go = (long)j * st + e0 + cc;
      if (!omega_only) { ((T*)a.u)[go] = vu; ((T*)a.w)[go] = vw; }
      ((T*)a.ox)[go] = v1; ((T*)a.oy)[go] = v2; ((T*)a.oz)[go] = v3;
    }
  }
}

// Pivots of (A - kk M) [implicit = 0] or (M - D(A - kk M)) [implicit = 1] for the upward elimination, one thread per
// mode, rows in order: piv[m*NYL + j] = 1/(d_j - l_j c_(j-1)), 0 on the wall rows (identity rows with zero right-hand
// side) and on the padding.
template <typename T>
__global__ void __launch_bounds__(128) pivot_lanes_kernel(PArgs<T> a, LaneTabs t, double* __restrict__ piv, int implicit) {
  const long m = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= a.M) return;
  const int ikx = a.ikx0 + (int)(m / a.Kz), kz = (int)(m % a.Kz);
  const int kx = ikx <= a.kxmax ? ikx : ikx - a.Kx;
  const double k1d = (double)a.k1fac * (double)kx, k3d = (double)a.k3fac * (double)kz, kk = k1d * k1d + k3d * k3d;
  double D = 0.0;
  if (implicit) { const T dt = (T)a.sc->dt; D = a.nuD * (double)(dt * a.betha); }
  const int ny = a.ny;
  double* row = piv + (size_t)m * t.NYL;
  const int S = t.NYL / t.NL;
  double cprev = 0.0;
  row[0] = 0.0;                                                   // wall rows and padding: zero pivots
  for (int j = ny - 1; j < t.NYL; j++) row[(j % S) * t.NL + j / S] = 0.0;
  for (int j = 1; j < ny - 1; j++) {
    const int ti = (j % S) * t.NL + j / S;
    const double rP = __ldg(t.rP + ti), rN = __ldg(t.rN + ti), wP = __ldg(t.wP + ti), wC = __ldg(t.wC + ti), wN = __ldg(t.wN + ti);
    double lo, di, up;
    if (implicit) { lo = rP - D * wP + D * kk * rP; di = 1.0 - D * wC + D * kk * 1.0; up = rN - D * wN + D * kk * rN; }
    else          { lo = wP - kk * rP;              di = wC - kk;                     up = wN - kk * rN; }
    if (j == 1) lo = 0.0;
    if (j == ny - 2) up = 0.0;
    const double pv = 1.0 / (di - lo * cprev);
    cprev = up * pv;
    row[(j % S) * t.NL + j / S] = pv;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Rows per lane: 4.  Small segments mean several warps per mode: what bounds the number of modes in flight on an SM is the
// shared memory of their staging tiles, so the parallelism has to come from the lanes of each mode (7 rows per lane, one
// warp per mode at NY = 193, ran at 8 warps per SM and spilled: measured no faster, profiles/r02_lanes_notes.txt).
static int lanes_S(const chan_handle*) { return 4; }
static int lanes_W(const chan_handle* h) { const int S = lanes_S(h); return (h->d.ny + 32 * S - 1) / (32 * S); }

int lanes_supported(chan_handle* h) { return h->d.ny >= 8 && lanes_W(h) <= 6; }

static size_t lanes_smem_bytes(const chan_handle* h, int ntiles, int MB, int NYP) {
  const size_t tsz = (size_t)2 * MB * NYP * h->prec;
  return (ntiles * tsz + 8) / 8 * 8 + (size_t)MB * 2 * h->ln.W * XW * sizeof(double);
}

int lanes_setup(chan_handle* h) {
  const int ny = h->d.ny, S = lanes_S(h), W = lanes_W(h), NYL = 32 * W * S;
  h->ln.S = S; h->ln.W = W; h->ln.NYL = NYL;
  const int collen = (S & 1) ? ny : ny + (ny - 1) / S + 1;           // cpos(ny-1) + 1
  // Modes per block (a power of two): the shape that keeps most warps on an SM (shared memory of the 7 staging tiles,
  // 128 registers per thread), among equals the one with more blocks per SM (a block that waits for its tiles leaves the
  // SM to the others), then the wider one (longer row segments in global memory).  CHAN_TUNE_LANES_MB overrides.
  int best = -1, bestMB = 0, bestNYP = 0;
  for (int MB = 8; MB >= 1; MB >>= 1) {
    if (MB * W * 32 > 512) continue;
    // column pitch: the rows of one warp-wide staging access (32 / (2 MB) consecutive y of every chain) must fall into
    // distinct banks: pitch = that row count modulo the bank count in elements (16 for 8-byte, 32 for 4-byte elements)
    const int banks = h->prec == 8 ? 16 : 32;
    int per = (h->prec == 8 ? 16 : 32) / (2 * MB); if (per < 1) per = 1;
    int NYP = collen;
    while (NYP % banks != per % banks) NYP++;
    const size_t sm = lanes_smem_bytes(h, 7, MB, NYP);
    if (sm > 227 * 1024) continue;
    int blocks = (int)((227 * 1024) / (sm + 1024)); const int rb = 512 / (MB * W * 32); if (blocks > rb) blocks = rb;
    if (blocks < 1) continue;
    const int score = blocks * MB * W * 1000 + (blocks > 1 ? 100 : 0) + MB;
    if (h->tune.lanes_mb == MB) { bestMB = MB; bestNYP = NYP; break; }
    if (score > best) { best = score; bestMB = MB; bestNYP = NYP; }
  }
  if (!bestMB) return set_err(h, CHAN_ERR_ARG, "NY=%d too large for the on-chip wall-normal stage", ny);
  h->ln.MB = bestMB; h->ln.NYP = bestNYP;
  // y tables in lane order: row j = L*S + r at [r*NL + L], zeros past NY
  const int NL = 32 * W;
  std::vector<double> tab(21 * (size_t)ny);
  CHAN_CUDA(h, cudaMemcpy(tab.data(), h->ytab_store, sizeof(double) * tab.size(), cudaMemcpyDeviceToHost));
  std::vector<double> p(11 * (size_t)NYL, 0.0);
  // order in ytab_store: y | dir0: eP eC eN eL cp rP rN wP wC wN | dir1 ...
  const int map[11] = { 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 0 };              // eP eC eN eL cp rP rN wP wC wN y
  for (int k = 0; k < 11; k++)
    for (int j = 0; j < ny; j++) p[(size_t)k * NYL + (size_t)(j % S) * NL + j / S] = tab[(size_t)map[k] * ny + j];
  CHAN_CUDA(h, cudaMalloc(&h->ln.tabs, sizeof(double) * p.size()));
  CHAN_CUDA(h, cudaMemcpy(h->ln.tabs, p.data(), sizeof(double) * p.size(), cudaMemcpyHostToDevice));
  const size_t pbytes = (size_t)h->d.M * NYL * sizeof(double);
  CHAN_CUDA(h, cudaMalloc(&h->ln.pivH, pbytes)); CHAN_CUDA(h, cudaMalloc(&h->ln.pivI, pbytes));
  h->ln.pivH_ready = 0;
  return CHAN_OK;
}

static void fill_lane_tabs(const chan_handle* h, LaneTabs& t) {
  const double* b = h->ln.tabs; const size_t n = h->ln.NYL;
  t.eP = b; t.eC = b + n; t.eN = b + 2 * n; t.eL = b + 3 * n; t.cp = b + 4 * n;
  t.rP = b + 5 * n; t.rN = b + 6 * n; t.wP = b + 7 * n; t.wC = b + 8 * n; t.wN = b + 9 * n; t.y = b + 10 * n;
  t.pivH = h->ln.pivH; t.pivI = h->ln.pivI; t.NYL = h->ln.NYL; t.NL = 32 * h->ln.W; t.NYP = h->ln.NYP; t.W = h->ln.W; t.MB = h->ln.MB;
}

static const float kAlphaL[3] = { 29.0f / 96.0f, -3.0f / 40.0f, 1.0f / 6.0f };     // RK3_kernels.cu:24
static const float kDsedaL[3] = { 0.0f, -17.0f / 60.0f, -5.0f / 12.0f };           // RK3_kernels.cu:25
static const float kGammaL[3] = { 8.0f / 15.0f, 5.0f / 12.0f, 3.0f / 4.0f };       // RK3_kernels.cu:113
static const float kBethaL[3] = { 37.0f / 160.0f, 5.0f / 24.0f, 1.0f / 6.0f };     // RK3.c:14

// Pivot tables of substep n on h->stream: they depend on dt (fixed for the whole step) and on n only, so chan_rk_substep
// issues this on the side stream at the START of the substep and the serial recurrence (0.1 ms) runs beside pre and the
// plane stage instead of in front of post.
template <typename T>
static int pivots_lanes_t(chan_handle* h, int n) {
  PArgs<T> a; fill_pencil_args<T>(h, a);
  a.n = n; a.betha = (T)kBethaL[n];
  LaneTabs t; fill_lane_tabs(h, t);
  const unsigned pg = (unsigned)((h->d.M + 127) / 128);
  if (!h->ln.pivH_ready) { pivot_lanes_kernel<T><<<pg, 128, 0, h->stream>>>(a, t, h->ln.pivH, 0); h->launches++; h->ln.pivH_ready = 1; }
  pivot_lanes_kernel<T><<<pg, 128, 0, h->stream>>>(a, t, h->ln.pivI, 1); h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
int launch_lanes_pivots(chan_handle* h, int n) { return h->prec == 4 ? pivots_lanes_t<float>(h, n) : pivots_lanes_t<double>(h, n); }

template <typename T, int S>
static int post_lanes_t(chan_handle* h, int n) {
  PArgs<T> a; fill_pencil_args<T>(h, a);
  a.n = n; a.alpha = (T)kAlphaL[n]; a.dseda = (T)kDsedaL[n]; a.gammad = (T)kGammaL[n]; a.betha = (T)kBethaL[n];
  LaneTabs t; fill_lane_tabs(h, t);
  const size_t sm = lanes_smem_bytes(h, 7, t.MB, t.NYP);
  CHAN_CUDA(h, cudaFuncSetAttribute(pencil_post_lanes_kernel<T, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  const unsigned grid = (unsigned)((h->d.M + t.MB - 1) / t.MB);
  pencil_post_lanes_kernel<T, S><<<grid, t.MB * t.W * 32, sm, h->stream>>>(a, t);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}

template <typename T, int S>
static int pre_lanes_t(chan_handle* h) {
  PArgs<T> a; fill_pencil_args<T>(h, a);
  LaneTabs t; fill_lane_tabs(h, t);
  PeerOut po; fill_peer_out(h, po);
  const size_t sm = lanes_smem_bytes(h, 7, t.MB, t.NYP);
  CHAN_CUDA(h, cudaFuncSetAttribute(pencil_pre_lanes_kernel<T, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  const unsigned grid = (unsigned)((h->d.M + t.MB - 1) / t.MB);
  pencil_pre_lanes_kernel<T, S><<<grid, t.MB * t.W * 32, sm, h->stream>>>(a, t, po);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
int launch_pencil_pre_lanes(chan_handle* h) { return h->prec == 4 ? pre_lanes_t<float, 4>(h) : pre_lanes_t<double, 4>(h); }

int launch_pencil_post_lanes(chan_handle* h, int n) {
  return h->prec == 4 ? post_lanes_t<float, 4>(h, n) : post_lanes_t<double, 4>(h, n);
}

}  // namespace chan
