// Wall-normal (y) stages of the RK3 substep: every compact-FD derivative and
// every tridiagonal Helmholtz / implicit / "bi-Laplacian" solve, fused with the
// pointwise spectral algebra around them.
//
// One thread owns one REAL chain: (mode m, component c = re|im) for all NY rows.
// The matrices are real, so the complex solves of the reference (cuSPARSE
// ZgtsvStridedBatch on complex right-hand sides, derivatives_nu_double.cu:406)
// split into two independent real Thomas sweeps.  Rows are plane-major
// (chan_internal.h), so at every sweep step the 32 lanes of a warp touch 32
// consecutive reals: 128/256-byte fully coalesced transactions, no shared
// memory, no transposes, any NY.
//
// Sweeps alternate direction: the back-substitution of solve k runs in the same
// direction as the elimination of solve k+1, so both are fused in ONE pass over
// y (the elimination lags one row behind).  Forward-sweep intermediates go to
// the fp64 scratch arrays sA/sB/sC (+ pivots sP) and are read back by the very
// thread that wrote them.  The compact first-derivative matrix is the same for
// every mode: its LU is folded into the y tables once (kernels_tables.cu); the
// Helmholtz / implicit pivots depend on kk and are formed on the fly.
//
// Reference semantics mirrored here (file:line under /root/reference/src):
//   pencil_pre   = calcUW (nonLinear_kernels.cu:8-92) + writeUmean
//                  (meanUevol.c:117-125) + the two deriv_Y of convolution.c:7-11
//                  + readUtau (:14-16) + calcOmega (convolution_kernels.cu:7-66)
//   pencil_post  = RKstep_1 (RK3_kernels.cu:155-177) + calcHvg
//                  (nonLinear_kernels.cu:212-232) + RKstep_2 (:179-194) +
//                  implicitSolver_double + bilaplaSolver_double
//                  (implicitStep_nu_double.cu:227, bilplacSolver_double.cu:320)
//                  + the copies/dealias of RK3.c:161-166 (dealias is structural
//                  here: truncated modes are not stored)
// The state type T rounds exactly where the reference's cast_kernel does (after
// every solve) and where its pointwise kernels compute in float.
#include <string.h>
#include "pencil_common.cuh"

namespace chan {

// The sweeps are latency-bound (one dependent row per thread in flight): pull the rows a few steps
// ahead into L2, one request per 128-byte line (lanes 0 and 16 of the warp's 256-byte row segment).
#define CHAN_PF_ROWS 6
__device__ __forceinline__ void pf_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
struct Prefetch {
  bool lead; int ny;
  __device__ __forceinline__ Prefetch(int ny_) : lead((threadIdx.x & 15) == 0), ny(ny_) {}
  // row j + d*CHAN_PF_ROWS of an array with `stride` elements per row; p points at this thread's element of row j
  template <typename E> __device__ __forceinline__ void operator()(const E* p, long stride, int j, int d) const {
    int jn = j + d * CHAN_PF_ROWS;
    if (lead && jn >= 0 && jn < ny) pf_l2(p + (long)d * CHAN_PF_ROWS * stride);
  }
};


// ---- fixed-matrix compact d/dy: elimination with the folded LU ---------------
struct ElimD1 {
  double f1, f2, dprev;
  __device__ __forceinline__ void step(const D1Tab& t, int dir, int jj, int ny, double f,
                                       double* s, long stride, long e) {
    if (jj == 0) { f2 = f; return; }
    if (jj == 1) { f1 = f; return; }
    if (jj == 2) {
      double d0 = __ldg(t.eC) * f2 + __ldg(t.eN) * f1 + __ldg(t.eP) * f;
      s[(long)phys_row(dir, 0, ny) * stride + e] = d0; dprev = d0;
    }
    int r = jj - 1;
    double d = __ldg(t.eP + r) * f2 + __ldg(t.eC + r) * f1 + __ldg(t.eN + r) * f - __ldg(t.eL + r) * dprev;
    s[(long)phys_row(dir, r, ny) * stride + e] = d; dprev = d;
    if (jj == ny - 1) {
      double dl = __ldg(t.eC + jj) * f + __ldg(t.eP + jj) * f1 + __ldg(t.eN + jj) * f2 - __ldg(t.eL + jj) * dprev;
      s[(long)phys_row(dir, jj, ny) * stride + e] = dl;
    }
    f2 = f1; f1 = f;
  }
};

// ---- kk-dependent Helmholtz / implicit matrices: pivots on the fly, NR right-hand sides
template <int NR>
struct ElimVar {
  double f1[NR], f2[NR], dprev[NR], cprev;
  // implicit: (M - D(A - kk M)) x = M f ; helmholtz: (A - kk M) x = M f   (hemholzt_nu_double.cu:156-163,
  // implicitStep_nu_double.cu:156-163); wall rows are identity with zero rhs (:167-185)
  __device__ __forceinline__ void step(const D2Tab& t, int dir, int jj, int ny, const double (&f)[NR],
                                       bool implicit, double kk, double D,
                                       double* const (&s)[NR], long stride, long e,
                                       double* sp, long M, long m) {
    if (jj == 0) {
#pragma unroll
      for (int q = 0; q < NR; q++) f2[q] = f[q];
      return;
    }
    if (jj == 1) {
      long r0 = phys_row(dir, 0, ny);
#pragma unroll
      for (int q = 0; q < NR; q++) { f1[q] = f[q]; s[q][r0 * stride + e] = 0.0; dprev[q] = 0.0; }
      sp[r0 * M + m] = 0.0; cprev = 0.0;
      return;
    }
    int r = jj - 1;
    double rP = __ldg(t.rP + r), rN = __ldg(t.rN + r);
    double wP = __ldg(t.wP + r), wC = __ldg(t.wC + r), wN = __ldg(t.wN + r);
    double lo, di, up;
    if (implicit) { lo = rP - D * wP + D * kk * rP; di = 1.0 - D * wC + D * kk * 1.0; up = rN - D * wN + D * kk * rN; }
    else          { lo = wP - kk * rP;              di = wC - kk;                     up = wN - kk * rN; }
    if (r == 1) lo = 0.0;
    if (r == ny - 2) up = 0.0;
    double piv = 1.0 / (di - lo * cprev);
    double cp = up * piv;
    long rr = phys_row(dir, r, ny);
#pragma unroll
    for (int q = 0; q < NR; q++) {
      double rhs = rN * f[q] + f1[q] + rP * f2[q];
      double d = (rhs - lo * dprev[q]) * piv;
      s[q][rr * stride + e] = d; dprev[q] = d;
      f2[q] = f1[q]; f1[q] = f[q];
    }
    sp[rr * M + m] = cp; cprev = cp;
    if (jj == ny - 1) {
      long rl = phys_row(dir, jj, ny);
#pragma unroll
      for (int q = 0; q < NR; q++) s[q][rl * stride + e] = 0.0;
      sp[rl * M + m] = 0.0;
    }
  }
};

// ---- pointer-walking variants used by pencil_post: address arithmetic was 48 % of its instructions
// (profiles/r01_final_*).  `cur` points at this chain's scratch element of the CURRENT sweep row, `ss` is
// the pointer step of one sweep position (+row stride going up, -row stride going down).
struct ElimD1P {
  double f1, f2, dprev;
  __device__ __forceinline__ void step(const D1Tab& t, int jj, int ny, double f, double* cur, long ss) {
    if (jj == 0) { f2 = f; return; }
    if (jj == 1) { f1 = f; return; }
    if (jj == 2) {
      double d0 = __ldg(t.eC) * f2 + __ldg(t.eN) * f1 + __ldg(t.eP) * f;
      cur[-2 * ss] = d0; dprev = d0;
    }
    const int r = jj - 1;
    double d = __ldg(t.eP + r) * f2 + __ldg(t.eC + r) * f1 + __ldg(t.eN + r) * f - __ldg(t.eL + r) * dprev;
    cur[-ss] = d; dprev = d;
    if (jj == ny - 1) {
      double dl = __ldg(t.eC + jj) * f + __ldg(t.eP + jj) * f1 + __ldg(t.eN + jj) * f2 - __ldg(t.eL + jj) * dprev;
      cur[0] = dl;
    }
    f2 = f1; f1 = f;
  }
};
template <int NR>
struct ElimVarP {
  double f1[NR], f2[NR], dprev[NR], cprev;
  // cur[q]: scratch element of the current sweep row; pcur: pivot element of the current sweep row (step ps)
  __device__ __forceinline__ void step(const D2Tab& t, int jj, int ny, const double (&f)[NR], bool implicit, double kk, double D,
                                       double* const (&cur)[NR], long ss, double* pcur, long ps) {
    if (jj == 0) {
#pragma unroll
      for (int q = 0; q < NR; q++) f2[q] = f[q];
      return;
    }
    if (jj == 1) {
#pragma unroll
      for (int q = 0; q < NR; q++) { f1[q] = f[q]; cur[q][-ss] = 0.0; dprev[q] = 0.0; }
      pcur[-ps] = 0.0; cprev = 0.0;
      return;
    }
    const int r = jj - 1;
    double rP = __ldg(t.rP + r), rN = __ldg(t.rN + r);
    double wP = __ldg(t.wP + r), wC = __ldg(t.wC + r), wN = __ldg(t.wN + r);
    double lo, di, up;
    if (implicit) { lo = rP - D * wP + D * kk * rP; di = 1.0 - D * wC + D * kk * 1.0; up = rN - D * wN + D * kk * rN; }
    else          { lo = wP - kk * rP;              di = wC - kk;                     up = wN - kk * rN; }
    if (r == 1) lo = 0.0;
    if (r == ny - 2) up = 0.0;
    double piv = 1.0 / (di - lo * cprev);
    double cp = up * piv;
#pragma unroll
    for (int q = 0; q < NR; q++) {
      double rhs = rN * f[q] + f1[q] + rP * f2[q];
      double d = (rhs - lo * dprev[q]) * piv;
      cur[q][-ss] = d; dprev[q] = d;
      f2[q] = f1[q]; f1[q] = f[q];
    }
    pcur[-ps] = cp; cprev = cp;
    if (jj == ny - 1) {
#pragma unroll
      for (int q = 0; q < NR; q++) cur[q][0] = 0.0;
      pcur[0] = 0.0;
    }
  }
};

// ---- unfused building blocks (operator-level API, calcVdV) -------------------
// out = d/dy in  (deriv_Y_HO_double, derivatives_nu_double.cu:390-415); in/out are real views [NY][2M]
// Sweep driver of pencil_post.  A thread's rows are strictly sequential, and the hardware issues in
// order, so with "load row j+1, then work on row j" only ONE row of loads is in flight per thread and a
// pass costs one memory latency per row (the ablation in profiles/r01_post_passes.txt: the passes with
// few bytes per row ran at 2-4 TB/s).  Here the loads of U rows are issued back to back, then the U
// dependent row updates follow.  ld(jj) returns the loaded row jj of the sweep, pr(jj, row) works on it.
template <int U, class Load, class Proc>
__device__ __forceinline__ void sweep_rows(int ny, Load ld, Proc pr, int part = 0, int nparts = 1) {
  // part / nparts: batches dealt round-robin to cooperating warps (only for passes whose rows are independent)
  for (int j0 = part * U; j0 < ny; j0 += nparts * U) {
    decltype(ld(0)) r[U];
#pragma unroll
    for (int u = 0; u < U; u++) r[u] = ld(min(j0 + u, ny - 1));        // clamped: a harmless reload
    __syncwarp();          // both lanes of a mode hold their rows before either overwrites shared pivots
#pragma unroll
    for (int u = 0; u < U; u++) if (j0 + u < ny) pr(j0 + u, r[u]);
  }
}

template <typename T>
__device__ void d1_solve(const PArgs<T>& a, const Mode& md, const T* __restrict__ in, T* __restrict__ out, double* s) {
  const long st = 2 * a.M; const int ny = a.ny;
  __syncwarp();
  ElimD1 el;
  for (int jj = 0; jj < ny; jj++) el.step(a.yt.d1[0], 0, jj, ny, (double)in[(long)jj * st + md.e], s, st, md.e);
  double x = 0.0;
  for (int j = ny - 1; j >= 0; j--) {
    x = s[(long)j * st + md.e] - __ldg(a.yt.d1[0].cp + j) * x;
    out[(long)j * st + md.e] = (T)x;
  }
}
template <typename T>
__device__ void var_solve(const PArgs<T>& a, const Mode& md, const Wave<T>& wv, bool implicit, double D,
                          const T* __restrict__ in, T* __restrict__ out, double* s, double* sp) {
  const long st = 2 * a.M; const int ny = a.ny;
  __syncwarp();   // the pivot row is shared by the two components of a mode
  ElimVar<1> el;
  double* const ss[1] = { s };
  for (int jj = 0; jj < ny; jj++) {
    const double f[1] = { (double)in[(long)jj * st + md.e] };
    el.step(a.yt.d2[0], 0, jj, ny, f, implicit, wv.kkd, D, ss, st, md.e, sp, a.M, md.m);
  }
  double x = 0.0;
  for (int j = ny - 1; j >= 0; j--) {
    x = s[(long)j * st + md.e] - sp[(long)j * a.M + md.m] * x;
    out[(long)j * st + md.e] = (T)x;
  }
}

// phi, v, dv: real views; dvm, dvp: own component of dv at j=0, NY-1 (state-type values)
// Double-buffered variant: the loads of batch k+1 are issued before batch k is worked on, so a thread always has
// U..2U rows of loads in flight (for passes whose rows are cheap in registers).  Legal wherever sweep_rows is:
// a pass only ever writes rows it has already loaded.
template <int U, class Load, class Proc>
__device__ __forceinline__ void sweep_rows_db(int ny, Load ld, Proc pr) {
  decltype(ld(0)) nx[U];
#pragma unroll
  for (int u = 0; u < U; u++) nx[u] = ld(min(u, ny - 1));
  for (int j0 = 0; j0 < ny; j0 += U) {
    decltype(ld(0)) r[U];
#pragma unroll
    for (int u = 0; u < U; u++) r[u] = nx[u];
    if (j0 + U < ny) {
#pragma unroll
      for (int u = 0; u < U; u++) nx[u] = ld(min(j0 + U + u, ny - 1));
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < U; u++) if (j0 + u < ny) pr(j0 + u, r[u]);
  }
}

template <typename T>
__device__ void bc_correct(const PArgs<T>& a, const Mode& md, const Wave<T>& wv, T betha, T dt,
                           T* phi, T* v, T* dv, double dvm, double dvp, int part = 0, int nparts = 1) {
  const long st = 2 * a.M; const int ny = a.ny;
  if (!((double)dt > 1e-8)) return;                       // bilplacSolver_double.cu:314 (Q6)
  // mode (0,0) is zeroed (:252-262); it runs the same instructions on a dummy kk so that the
  // shuffles inside bc_functions stay warp-uniform
  const BcCoef k = bc_coef(md.zero_mode ? 1.0 : wv.kkd, a.reD, (double)dt, (double)betha);
  BcFun p = bc_functions(k, 1.0, md.c), q = bc_functions(k, -1.0, md.c);
  double det = q.dv1 * p.dv2 - p.dv1 * q.dv2;
  double c1 = p.dv2 * dvm - q.dv2 * dvp; c1 = -c1 / det;
  double c2 = -p.dv1 * dvm + q.dv1 * dvp; c2 = -c2 / det;
  T* const pp = phi + md.e; T* const pv = v + md.e; T* const pd = dv + md.e;
  struct Row { T p, v, d; };
  auto ld = [&](int j) { Row r; r.p = pp[(long)j * st]; r.v = pv[(long)j * st]; r.d = pd[(long)j * st]; return r; };
  auto pr = [&](int j, const Row& r) {
      BcFun f = bc_functions(k, __ldg(a.yt.y + j), md.c);
      double ppv = (double)r.p, vv = (double)r.v, dd = (double)r.d;
      ppv = ppv + c1 * f.ddv1 + c2 * f.ddv2;
      dd = dd + c1 * f.dv1 + c2 * f.dv2;
      vv = vv + c1 * f.v1 + c2 * f.v2;
      if (md.zero_mode) { ppv = 0.0; vv = 0.0; dd = 0.0; }
      pp[(long)j * st] = (T)ppv; pv[(long)j * st] = (T)vv; pd[(long)j * st] = (T)dd;
    };
  if (nparts == 1) sweep_rows_db<4>(ny, ld, pr);
  else sweep_rows<8>(ny, ld, pr, part, nparts);
}

// =================================================================================
// pencil_pre: (dv, g, v, <u>) -> u, w, omega_x, omega_y, omega_z
// =================================================================================
template <typename T>
__global__ void __launch_bounds__(128, 4) pencil_pre_kernel(PArgs<T> a, PeerOut po) {
  typedef typename Vec2<T>::type T2;
  Mode md; Wave<T> wv;
  if (!setup(a, md, wv)) return;
  const long st = 2 * a.M, M = a.M; const int ny = a.ny; const int c = md.c; const long e = md.e;
  const int top = ny - 1;
  const bool with_mean = md.zero_mode && a.use_mean;
  const bool omega = a.op == OP_OMEGA;                 // operator-level calcOmega: u, w are given
  const int sst = 32;                                  // warp-tiled scratch, see pencil_post
  double* const sA0 = a.sA + (e >> 5) * 32 * ny + (e & 31);
  double* const sB0 = a.sB + (e >> 5) * 32 * ny + (e & 31);
  const T2* const f0 = (omega ? a.u : a.dv) + md.m; const T2* const f1 = (omega ? a.w : a.g) + md.m;
  // pass 1 (up): u, w on the fly -> eliminate d/dy u, d/dy w
  {
    ElimD1P eu, ew;
    struct Row { T2 a0, a1; };
    sweep_rows<4>(ny,
      [&](int j) { Row r; r.a0 = f0[(long)j * M]; r.a1 = f1[(long)j * M]; return r; },
      [&](int j, const Row& r) {
        T u, w;
        if (omega) { u = own<T2, T>(r.a0, c); w = own<T2, T>(r.a1, c); }
        else {
          uw_comp<T, T2>(wv, md, md.zero_mode, r.a0, r.a1, c, u, w);
          if (with_mean) u = c ? (T)0 : a.mean_u[j];                 // writeUmean (nonLinear.c:19-21)
        }
        eu.step(a.yt.d1[0], j, ny, (double)u, sA0 + j * sst, sst);
        ew.step(a.yt.d1[0], j, ny, (double)w, sB0 + j * sst, sst);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 7);
  // pass 2 (down): back-substitute, build omega (calcOmegakernel, convolution_kernels.cu:46-53)
  {
    double xu = 0.0, xw = 0.0;
    const T2* const fv = a.v + md.m;
    struct Row { double dA, dB; T2 v, a0, a1; };       // a0,a1 = (dv,g) or (u,w) of the row
    sweep_rows<4>(ny,
      [&](int jj) {
        const int j = top - jj; Row r;
        r.dA = sA0[j * sst]; r.dB = sB0[j * sst]; r.v = fv[(long)j * M]; r.a0 = f0[(long)j * M]; r.a1 = f1[(long)j * M];
        return r;
      },
      [&](int jj, const Row& r) {
        const int j = top - jj;
        double cp = __ldg(a.yt.d1[0].cp + j);
        xu = r.dA - cp * xu; xw = r.dB - cp * xw;
        T dudy = (T)xu, dwdy = (T)xw;
        T u, w, up, wp;
        if (omega) {
          u = own<T2, T>(r.a0, c); up = other<T2, T>(r.a0, c); w = own<T2, T>(r.a1, c); wp = other<T2, T>(r.a1, c);
        } else {
          uw_comp<T, T2>(wv, md, md.zero_mode, r.a0, r.a1, c, u, w);
          uw_comp<T, T2>(wv, md, md.zero_mode, r.a0, r.a1, c ^ 1, up, wp);
          if (with_mean) { u = c ? (T)0 : a.mean_u[j]; up = c ? a.mean_u[j] : (T)0; }
        }
        T vp = other<T2, T>(r.v, c);
        T s = wv.sc;
        T o1 = dwdy - s * wv.k3 * vp;                    // omega_x = dw/dy - i k3 v
        T o2 = s * (wv.k3 * up - wv.k1 * wp);            // omega_y = i (k3 u - k1 w)
        T o3 = -dudy + s * wv.k1 * vp;                   // omega_z = -du/dy + i k1 v
        if (po.nranks > 1) {
          // multi-GPU: row j goes straight into the receive buffer of the rank that owns plane j (an NVLink
          // store for the other ranks) - the pencil -> plane transpose of channel_cuda_mpi.c:95-128
          int p = 0;
#pragma unroll
          for (int i = 1; i < 8; i++) p += (i < po.nranks && j >= po.part_j[i]) ? 1 : 0;
          long idx = 2 * (po.src_off[p] + (long)(j - po.part_j[p]) * M) + e;
          ((T*)po.rcv[p][0])[idx] = u; ((T*)po.rcv[p][1])[idx] = own<T2, T>(r.v, c); ((T*)po.rcv[p][2])[idx] = w;
          ((T*)po.rcv[p][3])[idx] = o1; ((T*)po.rcv[p][4])[idx] = o2; ((T*)po.rcv[p][5])[idx] = o3;
        } else {
          const long re = (long)j * st + e;
          if (!omega) { ((T*)a.u)[re] = u; ((T*)a.w)[re] = w; }
          ((T*)a.ox)[re] = o1; ((T*)a.oy)[re] = o2; ((T*)a.oz)[re] = o3;
        }
        if (md.zero_mode && c == 0 && a.sc) {            // readUtau (convolution.c:14-16)
          if (j == 0) a.sc->dudy_wall[0] = (double)dudy;
          if (j == ny - 1) a.sc->dudy_wall[1] = (double)dudy;
        }
      });
  }
}

// Delivery of the six plane-stage inputs (u, v, w, omega) to the plane owners over NVLink: the pencil -> plane
// transpose of channel_cuda_mpi.c:95-128.  In the plane-major layout the rows a peer needs are ONE contiguous slab of
// each field (rows [j0_p, j0_p + nyl_p) x all local modes), so the delivery is 6 x S flat copies.  Two things decide its
// speed (measured on 8 B200s, profiles/r02_n8_notes.txt): the packets must be large — stores issued from inside the y
// kernels are 64..128-byte row pieces and reach 240-330 GB/s per GPU — and no two ranks may send to the same owner at
// the same time: block z of rank r copies to rank (r + z) mod S, so the ranks walk the destinations in step, one sender
// per receiver at any moment.
struct ScatterArgs { const void* src[6]; void* dst[8][6]; long src_off[8], dst_off[8], count[8]; int S, rank; };
template <typename V>
__global__ void __launch_bounds__(256) peer_scatter_kernel(ScatterArgs a) {
  const int p = (a.rank + blockIdx.z) % a.S, q = blockIdx.y;
  const V* __restrict__ s = (const V*)a.src[q] + a.src_off[p];
  V* __restrict__ d = (V*)a.dst[p][q] + a.dst_off[p];
  const long n = a.count[p], stride = (long)gridDim.x * blockDim.x * 4;
  for (long i0 = (long)blockIdx.x * blockDim.x * 4 + threadIdx.x; i0 < n; i0 += stride) {
    V v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) { const long i = i0 + (long)k * blockDim.x; if (i < n) v[k] = s[i]; }
#pragma unroll
    for (int k = 0; k < 4; k++) { const long i = i0 + (long)k * blockDim.x; if (i < n) d[i] = v[k]; }
  }
}
static int run_scatter(chan_handle* h, ScatterArgs& a, int nf) {
  const int S = h->cfg.size;
  a.S = S; a.rank = h->cfg.rank;
  if (h->tune.xchg_ce) {
    // CHAN_TUNE_XCHG_CE=1: the same flat copies as cudaMemcpyAsync on four copy streams (copy engines instead of SMs, so
    // that they can run beside the transforms without taking issue slots); same rotation of the destinations
    const size_t cb = 2 * (size_t)h->prec;
    cudaStream_t base = h->stream;
    CHAN_CUDA(h, cudaEventRecord(h->ev_cf, base));
    for (int i = 0; i < 4; i++) CHAN_CUDA(h, cudaStreamWaitEvent(h->xc[i], h->ev_cf, 0));
    for (int z = 0; z < S; z++) {
      const int p = (a.rank + z) % S;
      if (!a.count[p]) continue;
      for (int q = 0; q < nf; q++)
        CHAN_CUDA(h, cudaMemcpyAsync((char*)a.dst[p][q] + a.dst_off[p] * cb, (const char*)a.src[q] + a.src_off[p] * cb, a.count[p] * cb, cudaMemcpyDeviceToDevice, h->xc[z & 3]));
    }
    for (int i = 0; i < 4; i++) { CHAN_CUDA(h, cudaEventRecord(h->ev_cj[i], h->xc[i])); CHAN_CUDA(h, cudaStreamWaitEvent(base, h->ev_cj[i], 0)); }
    return CHAN_OK;
  }
  // copy units of 16 bytes: one fp64 complex, or two fp32 complex when every offset and count is even
  const long per16 = h->prec == 8 ? 1 : 2;
  bool even = true;
  for (int p = 0; p < S; p++) if ((a.src_off[p] | a.dst_off[p] | a.count[p]) % per16) even = false;
  dim3 grid(16, nf, S);
  if (h->prec == 8) peer_scatter_kernel<double2><<<grid, 256, 0, h->stream>>>(a);
  else if (even) { for (int p = 0; p < S; p++) { a.src_off[p] /= 2; a.dst_off[p] /= 2; a.count[p] /= 2; } peer_scatter_kernel<float4><<<grid, 256, 0, h->stream>>>(a); }
  else peer_scatter_kernel<float2><<<grid, 256, 0, h->stream>>>(a);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
int launch_peer_scatter(chan_handle* h, int batch) {
  ScatterArgs a; memset(&a, 0, sizeof(a));
  const void* f[6] = { h->u, h->v, h->w, h->ox, h->oy, h->oz };
  const long M = h->d.M;
  for (int q = 0; q < 6; q++) a.src[q] = f[q];
  for (int p = 0; p < h->cfg.size; p++) {
    const long nyl = h->part_j[p + 1] - h->part_j[p];
    long r0 = 0, nr = nyl;                                   // rows of the destination's plane batch `batch`
    if (batch >= 0) { r0 = (long)batch * h->pb_rank[p]; nr = nyl - r0 < h->pb_rank[p] ? nyl - r0 : h->pb_rank[p]; if (nr < 0) nr = 0; if (r0 > nyl) r0 = nyl; }
    long L[6]; chan_peer_put_layout(h->d.nx, h->d.ny, h->d.nz, h->cfg.size, h->cfg.rank, p, L);     // the layout the CPU tests check
    a.src_off[p] = L[0] + r0 * M; a.dst_off[p] = L[1] + r0 * M; a.count[p] = nr * M;
    for (int q = 0; q < 6; q++) a.dst[p][q] = h->peer_rcv[p][q];
  }
  return run_scatter(h, a, 6);
}
// The way back (transposeXYZ2YZX, channel_cuda_mpi.c:64-93): xfft_fwd leaves the three products in the plane-side send
// buffers, one block per destination rank (SegMap), and each block is one contiguous slab [j0, j0+nyl) x (all modes of the
// destination) of the destination's pencil-side field.
// rows [js, js+pb) of this rank's planes (one plane batch), on h->stream
int launch_peer_scatter_back(chan_handle* h, int js, int pb) {
  ScatterArgs a; memset(&a, 0, sizeof(a));
  for (int q = 0; q < 3; q++) a.src[q] = h->snd[q];
  for (int p = 0; p < h->cfg.size; p++) {
    const long Mp = (long)h->seg.width[p] * h->d.Kz;
    long L[6]; chan_peer_put_layout(h->d.nx, h->d.ny, h->d.nz, h->cfg.size, h->cfg.rank, p, L);
    a.src_off[p] = L[3] + (long)js * Mp; a.dst_off[p] = L[4] + (long)js * Mp; a.count[p] = (long)pb * Mp;
    for (int q = 0; q < 3; q++) a.dst[p][q] = h->peer_H[p][q];
  }
  return run_scatter(h, a, 3);
}

// =================================================================================
// pencil_post: H1..3, phi, g, R_phi, R_g -> phi, g, v, dv, R_phi, R_g   (7 passes)
// =================================================================================
// DEEP = false: 128 registers, four blocks per SM (a GPU full of chains, bandwidth-bound).
// DEEP = true: few chains per SM (many ranks, small grids): occupancy is irrelevant and a pass costs one
// memory latency per batch, so the batches are as deep as 255 registers allow.
#ifndef POST_DB
#define POST_DB 1
#endif
#if POST_DB
#define SWEEP_DB(U) sweep_rows_db<(U) / 2>
#else
#define SWEEP_DB(U) sweep_rows<U>
#endif
template <typename T, bool DEEP>
__global__ void __launch_bounds__(128, DEEP ? 1 : 4) pencil_post_kernel(PArgs<T> a) {
  constexpr int PU1 = DEEP ? 8 : 6, PU2 = DEEP ? 12 : 8, PU3 = DEEP ? 6 : 3, PU4 = DEEP ? 12 : 8, PU5 = DEEP ? 16 : 8, PU6 = DEEP ? 16 : 8;
  typedef typename Vec2<T>::type T2;
  Mode md; Wave<T> wv;
  if (!setup(a, md, wv)) return;
  const long st = 2 * a.M, M = a.M; const int ny = a.ny; const int c = md.c; const long e = md.e;
  const T dt = (T)a.sc->dt;
  const T s = wv.sc;
  const int top = ny - 1;
  // The scratch belongs to this kernel alone, so it is tiled per warp: [32 chains][row][lane] (pivots:
  // [16 modes][row][lane pair]).  A warp then streams through ONE contiguous piece of memory per scratch
  // array instead of touching a 256-byte fragment of every 2M-wide row.
  const int sst = 32, pst = 16;
  double* const sA0 = a.sA + (e >> 5) * 32 * ny + (e & 31);
  double* const sB0 = a.sB + (e >> 5) * 32 * ny + (e & 31);
  double* const sC0 = a.sC + (e >> 5) * 32 * ny + (e & 31);
  double* const sP0 = a.sP + (md.m >> 4) * 16 * ny + (md.m & 15);
  const T2* const H1 = a.ox + md.m; const T2* const H3 = a.oz + md.m; const T* const H2 = (const T*)a.oy + e;
  T* const phi = (T*)a.phi + e; T* const g = (T*)a.g + e;
  // P1 (up): eliminate d/dy of phi, g and hv = -i (k1 H1 + k3 H3)  (calcHvgkernel, nonLinear_kernels.cu:129-130)
  {
    ElimD1P ep, eg, eh;
    struct Row { T2 h1, h3; T p, g; };
    SWEEP_DB(PU1)(ny,
      [&](int j) { Row r; r.h1 = H1[(long)j * M]; r.h3 = H3[(long)j * M]; r.p = phi[(long)j * st]; r.g = g[(long)j * st]; return r; },
      [&](int j, const Row& r) {
        T h1p = other<T2, T>(r.h1, c), h3p = other<T2, T>(r.h3, c);
        T hv = -s * (wv.k1 * h1p + wv.k3 * h3p);
        ep.step(a.yt.d1[0], j, ny, (double)r.p, sA0 + j * sst, sst);
        eg.step(a.yt.d1[0], j, ny, (double)r.g, sB0 + j * sst, sst);
        eh.step(a.yt.d1[0], j, ny, (double)hv, sC0 + j * sst, sst);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 1);
  // P2 (down): phi', g', hv' ; eliminate d/dy of phi', g' (second derivative = d/dy twice, Q4)
  {
    ElimD1P ep, eg;
    double xp = 0.0, xg = 0.0, xh = 0.0;
    struct Row { double dA, dB, dC; };
    SWEEP_DB(PU2)(ny,
      [&](int jj) { int j = top - jj; Row r; r.dA = sA0[j * sst]; r.dB = sB0[j * sst]; r.dC = sC0[j * sst]; return r; },
      [&](int jj, const Row& r) {
        const int j = top - jj;
        double cp = __ldg(a.yt.d1[0].cp + j);
        xp = r.dA - cp * xp; xg = r.dB - cp * xg; xh = r.dC - cp * xh;
        sC0[j * sst] = rnd<T>(xh);
        ep.step(a.yt.d1[1], jj, ny, rnd<T>(xp), sA0 + j * sst, -sst);
        eg.step(a.yt.d1[1], jj, ny, rnd<T>(xg), sB0 + j * sst, -sst);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 2);
  // P3 (up): phi'', g'' -> explicit RK update (rk_step_1/2, RK3_kernels.cu:76-80,138-142) -> eliminate implicit solves
  const double bdt = (double)(dt * a.betha);                         // dt*betha in the state type (:235)
  const double Dimp = a.nuD * bdt;
  {
    ElimVarP<2> ev;
    double xp = 0.0, xg = 0.0;
    T* const Rp = (T*)a.Rphi + e; T* const Rg = (T*)a.Rg + e;
    struct Row { double dA, dB, dC; T2 h1, h3; T h2, p, g, rp, rg; };
    sweep_rows<PU3>(ny,
      [&](int j) {
        Row r; r.dA = sA0[j * sst]; r.dB = sB0[j * sst]; r.dC = sC0[j * sst]; r.h1 = H1[(long)j * M]; r.h3 = H3[(long)j * M];
        r.h2 = H2[(long)j * st]; r.p = phi[(long)j * st]; r.g = g[(long)j * st]; r.rp = Rp[(long)j * st]; r.rg = Rg[(long)j * st];
        return r;
      },
      [&](int j, const Row& r) {
        double cp = __ldg(a.yt.d1[1].cp + (top - j));
        xp = r.dA - cp * xp; xg = r.dB - cp * xg;
        T pyy = (T)xp, gyy = (T)xg;
        T h1p = other<T2, T>(r.h1, c), h3p = other<T2, T>(r.h3, c);
        T rp_new = (T)r.dC - wv.kk * r.h2;                              // calcHvvkernel (:180-181)
        T rg_new = s * (wv.k3 * h1p - wv.k1 * h3p);                     // h_g = i (k3 H1 - k1 H3) (:132-133)
        T pw = r.p + dt * (a.alpha * a.nuT * (pyy - wv.kk * r.p) + a.dseda * r.rp);
        T gw = r.g + dt * (a.alpha * a.nuT * (gyy - wv.kk * r.g) + a.dseda * r.rg);
        pw = pw + dt * a.gammad * rp_new;
        gw = gw + dt * a.gammad * rg_new;
        Rp[(long)j * st] = rp_new; Rg[(long)j * st] = rg_new;
        const double f[2] = { (double)pw, (double)gw };
        double* const cur[2] = { sA0 + j * sst, sB0 + j * sst };
        ev.step(a.yt.d2[0], j, ny, f, true, wv.kkd, Dimp, cur, sst, sP0 + j * pst, pst);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 3);
  // P4 (down): phi_new, g_new ; eliminate Helmholtz (D2 - kk) v = phi_new.  The Helmholtz elimination
  // lags one row ABOVE the row being read, so the batch of rows below still holds the implicit-solve data.
  {
    ElimVarP<1> eh;
    double xp = 0.0, xg = 0.0;
    struct Row { double dA, dB, cp; };
    SWEEP_DB(PU4)(ny,
      [&](int jj) { int j = top - jj; Row r; r.dA = sA0[j * sst]; r.dB = sB0[j * sst]; r.cp = sP0[j * pst]; return r; },
      [&](int jj, const Row& r) {
        const int j = top - jj;
        xp = r.dA - r.cp * xp; xg = r.dB - r.cp * xg;
        g[(long)j * st] = (T)xg;
        phi[(long)j * st] = (T)xp;
        const double f[1] = { rnd<T>(xp) };
        double* const cur[1] = { sA0 + j * sst };
        eh.step(a.yt.d2[1], jj, ny, f, false, wv.kkd, 0.0, cur, -sst, sP0 + j * pst, -pst);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 4);
  // P5 (up): v ; eliminate d/dy v
  {
    ElimD1P ev;
    double xv = 0.0;
    T* const pv = (T*)a.v + e;
    struct Row { double dA, cp; };
    SWEEP_DB(PU5)(ny,
      [&](int j) { Row r; r.dA = sA0[j * sst]; r.cp = sP0[j * pst]; return r; },
      [&](int j, const Row& r) {
        xv = r.dA - r.cp * xv;
        pv[(long)j * st] = (T)xv;
        ev.step(a.yt.d1[0], j, ny, rnd<T>(xv), sA0 + j * sst, sst);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 5);
  // P6 (down): dv
  double dvm = 0.0, dvp = 0.0;
  {
    double x = 0.0;
    T* const pdv = (T*)a.dv + e;
    sweep_rows<PU6>(ny,
      [&](int jj) { return sA0[(top - jj) * sst]; },
      [&](int jj, const double& dA) {
        const int j = top - jj;
        x = dA - __ldg(a.yt.d1[0].cp + j) * x;
        pdv[(long)j * st] = (T)x;
        if (j == top) dvp = rnd<T>(x);
        if (j == 0) dvm = rnd<T>(x);
      });
  }
  CHAN_ABLATE_RETURN(a.dbg_stop == 6);
  // P7: wall correction (skipped when dt <= 1e-8, including the zeroing of mode (0,0))
  bc_correct<T>(a, md, wv, a.betha, dt, (T*)a.phi, (T*)a.v, (T*)a.dv, dvm, dvp);
}

// =================================================================================
// pencil_post for ranks that hold few chains (many GPUs, small grids).  One warp per chain runs out of
// parallelism there: the kernel is then bound by the ~135k dependent instructions a chain executes, not by
// memory.  Here THREE warps share 32 chains, one role each; the passes already hand their data over through
// the (L2-resident) scratch, so the roles only meet at a block barrier between passes:
//   pass   warp 0 (phi)                      warp 1 (g)                    warp 2 (hv)
//   P1     eliminate phi'                    eliminate g'                  eliminate hv'
//   P2     phi' -> eliminate phi''           g' -> eliminate g''           hv'
//   P3     phi'' , RK update, implicit elim. g'', RK update, implicit el.  -
//   P4     phi_new, Helmholtz elimination    g_new                         -
//   P5-6   v, dv                             -                             -
//   P7     wall correction, rows dealt round-robin to the three warps
// Same arithmetic in the same order per chain as pencil_post_kernel (bit-identical results).  The Helmholtz
// pivots go to sC (free after P3) because warp 1 still reads the implicit pivots in sP during P4.
// =================================================================================
template <typename T>
__global__ void __launch_bounds__(96, 2) pencil_post_roles_kernel(PArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  __shared__ double wall[2][32];
  const int role = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Mode md; Wave<T> wv;
  long e = (long)blockIdx.x * 32 + lane;
  // lanes past the last chain leave (2M is even, so re/im pairs stay complete for the shuffles of the wall
  // correction); every warp of a block keeps at least one pair, so the block barriers below stay valid
  if (e >= 2 * a.M) return;
  const bool active = true;
  setup_chain(a, e, md, wv);
  const long st = 2 * a.M, M = a.M; const int ny = a.ny; const int c = md.c;
  const T dt = (T)a.sc->dt;
  const T s = wv.sc;
  const int top = ny - 1;
  const int sst = 32, pst = 16;
  const long tile = (long)blockIdx.x * 32 * ny + lane;     // this block's warp tile of the scratch
  double* const sA0 = a.sA + tile; double* const sB0 = a.sB + tile; double* const sC0 = a.sC + tile;
  double* const sP0 = a.sP + (long)blockIdx.x * 16 * ny + (lane >> 1);
  const T2* const H1 = a.ox + md.m; const T2* const H3 = a.oz + md.m; const T* const H2 = (const T*)a.oy + e;
  T* const phi = (T*)a.phi + e; T* const g = (T*)a.g + e;
  // P1
  if (active) {
    ElimD1P el;
    if (role < 2) {
      const T* const f = role == 0 ? phi : g; double* const sX = role == 0 ? sA0 : sB0;
      sweep_rows<16>(ny, [&](int j) { return f[(long)j * st]; },
                     [&](int j, const T& v) { el.step(a.yt.d1[0], j, ny, (double)v, sX + j * sst, sst); });
    } else {
      struct Row { T2 h1, h3; };
      sweep_rows<8>(ny, [&](int j) { Row r; r.h1 = H1[(long)j * M]; r.h3 = H3[(long)j * M]; return r; },
                    [&](int j, const Row& r) {
                      T h1p = other<T2, T>(r.h1, c), h3p = other<T2, T>(r.h3, c);
                      T hv = -s * (wv.k1 * h1p + wv.k3 * h3p);
                      el.step(a.yt.d1[0], j, ny, (double)hv, sC0 + j * sst, sst);
                    });
    }
  }
  __syncthreads();
  // P2
  if (active) {
    double* const sX = role == 0 ? sA0 : (role == 1 ? sB0 : sC0);
    ElimD1P el; double x = 0.0;
    sweep_rows<16>(ny, [&](int jj) { return sX[(top - jj) * sst]; },
                   [&](int jj, const double& d) {
                     const int j = top - jj;
                     x = d - __ldg(a.yt.d1[0].cp + j) * x;
                     if (role == 2) sX[j * sst] = rnd<T>(x);
                     else el.step(a.yt.d1[1], jj, ny, rnd<T>(x), sX + j * sst, -sst);
                   });
  }
  __syncthreads();
  // P3
  const double bdt = (double)(dt * a.betha);
  const double Dimp = a.nuD * bdt;
  if (active && role < 2) {
    ElimVarP<1> ev; double x = 0.0;
    T* const R = (role == 0 ? (T*)a.Rphi : (T*)a.Rg) + e;
    T* const f = role == 0 ? phi : g; double* const sX = role == 0 ? sA0 : sB0;
    struct Row { double d, dC; T2 h1, h3; T h2, f, r; };
    sweep_rows<6>(ny,
      [&](int j) {
        Row r; r.d = sX[j * sst]; r.f = f[(long)j * st]; r.r = R[(long)j * st];
        if (role == 0) { r.dC = sC0[j * sst]; r.h2 = H2[(long)j * st]; r.h1 = mk0<T2>(); r.h3 = mk0<T2>(); }
        else { r.h1 = H1[(long)j * M]; r.h3 = H3[(long)j * M]; r.dC = 0.0; r.h2 = 0; }
        return r;
      },
      [&](int j, const Row& r) {
        x = r.d - __ldg(a.yt.d1[1].cp + (top - j)) * x;
        T fyy = (T)x, rnew;
        if (role == 0) rnew = (T)r.dC - wv.kk * r.h2;                                   // calcHvvkernel (:180-181)
        else { T h1p = other<T2, T>(r.h1, c), h3p = other<T2, T>(r.h3, c); rnew = s * (wv.k3 * h1p - wv.k1 * h3p); }
        T fw = r.f + dt * (a.alpha * a.nuT * (fyy - wv.kk * r.f) + a.dseda * r.r);
        fw = fw + dt * a.gammad * rnew;
        R[(long)j * st] = rnew;
        const double ff[1] = { (double)fw };
        double* const cur[1] = { sX + j * sst };
        ev.step(a.yt.d2[0], j, ny, ff, true, wv.kkd, Dimp, cur, sst, sP0 + j * pst, pst);   // both roles write the same pivots
      });
  }
  __syncthreads();
  // P4
  if (active && role < 2) {
    ElimVarP<1> eh; double x = 0.0;
    T* const f = role == 0 ? phi : g; double* const sX = role == 0 ? sA0 : sB0;
    struct Row { double d, cp; };
    sweep_rows<16>(ny, [&](int jj) { int j = top - jj; Row r; r.d = sX[j * sst]; r.cp = sP0[j * pst]; return r; },
                   [&](int jj, const Row& r) {
                     const int j = top - jj;
                     x = r.d - r.cp * x;
                     f[(long)j * st] = (T)x;
                     if (role == 0) {
                       const double ff[1] = { rnd<T>(x) };
                       double* const cur[1] = { sX + j * sst };
                       eh.step(a.yt.d2[1], jj, ny, ff, false, wv.kkd, 0.0, cur, -sst, sC0 + j * sst, -sst);
                     }
                   });
  }
  __syncthreads();
  // P5, P6
  if (active && role == 0) {
    {
      ElimD1P ev; double xv = 0.0;
      T* const pv = (T*)a.v + e;
      struct Row { double dA, cp; };
      sweep_rows<16>(ny, [&](int j) { Row r; r.dA = sA0[j * sst]; r.cp = sC0[j * sst]; return r; },
                     [&](int j, const Row& r) {
                       xv = r.dA - r.cp * xv;
                       pv[(long)j * st] = (T)xv;
                       ev.step(a.yt.d1[0], j, ny, rnd<T>(xv), sA0 + j * sst, sst);
                     });
    }
    double x = 0.0, dvm = 0.0, dvp = 0.0;
    T* const pdv = (T*)a.dv + e;
    sweep_rows<16>(ny, [&](int jj) { return sA0[(top - jj) * sst]; },
                   [&](int jj, const double& dA) {
                     const int j = top - jj;
                     x = dA - __ldg(a.yt.d1[0].cp + j) * x;
                     pdv[(long)j * st] = (T)x;
                     if (j == top) dvp = rnd<T>(x);
                     if (j == 0) dvm = rnd<T>(x);
                   });
    wall[0][lane] = dvm; wall[1][lane] = dvp;
  }
  __syncthreads();
  // P7
  if (active) bc_correct<T>(a, md, wv, a.betha, dt, (T*)a.phi, (T*)a.v, (T*)a.dv, wall[0][lane], wall[1][lane], role, 3);
}

// =================================================================================
// generic operator kernel (operator-level C ABI and the RKstep prologue)
// =================================================================================
template <typename T>
__global__ void __launch_bounds__(128) op_kernel(PArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  Mode md; Wave<T> wv;
  if (!setup(a, md, wv)) return;
  const long st = 2 * a.M; const int ny = a.ny; const int c = md.c;
  T* f0 = (T*)a.f0; T* f1 = (T*)a.f1; T* f2 = (T*)a.f2;
  switch (a.op) {
    case OP_DERIV_Y: d1_solve<T>(a, md, f0, f0, a.sA); break;
    case OP_DERIV_YY: d1_solve<T>(a, md, f0, f0, a.sA); d1_solve<T>(a, md, f0, f0, a.sA); break;
    case OP_HELM: var_solve<T>(a, md, wv, false, 0.0, f0, f0, a.sA, a.sP); break;
    case OP_IMPLICIT: {
      double D = a.nuD * (double)(a.op_dt * a.op_betha);
      var_solve<T>(a, md, wv, true, D, f0, f0, a.sA, a.sP);
    } break;
    case OP_VDV:           // calcVdV (RK3.c:52-62): f1 = Helmholtz(f0), f2 = d/dy f1
      var_solve<T>(a, md, wv, false, 0.0, f0, f1, a.sA, a.sP);
      d1_solve<T>(a, md, f1, f2, a.sA);
      break;
    case OP_BILAP: {       // bilaplaSolver_double (bilplacSolver_double.cu:320-348)
      double D = a.nuD * (double)(a.op_dt * a.op_betha);
      var_solve<T>(a, md, wv, true, D, f0, f0, a.sA, a.sP);
      var_solve<T>(a, md, wv, false, 0.0, f0, f1, a.sA, a.sP);
      d1_solve<T>(a, md, f1, f2, a.sA);
      double dvm = (double)f2[md.e], dvp = (double)f2[(long)(ny - 1) * st + md.e];
      bc_correct<T>(a, md, wv, a.op_betha, a.op_dt, f0, f1, f2, dvm, dvp);
    } break;
    case OP_UW:            // calcUW: (f0 = dv, f1 = g) -> (u, w)
      for (int j = 0; j < ny; j++) {
        long cm = (long)j * a.M + md.m; T u, w;
        uw_comp<T, T2>(wv, md, md.zero_mode, a.f0[cm], a.f1[cm], c, u, w);
        ((T*)a.u)[(long)j * st + md.e] = u; ((T*)a.w)[(long)j * st + md.e] = w;
      }
      break;
    case OP_HVG: {         // calcHvg (nonLinear_kernels.cu:212-232): f0 = H1 -> h_v, f2 = H3 -> h_g, f1 = H2
      T s = wv.sc;
      T* tmp = (T*)a.u;    // work field
      for (int j = 0; j < ny; j++) {
        long cm = (long)j * a.M + md.m, re = (long)j * st + md.e;
        T h1p = other<T2, T>(a.f0[cm], c), h3p = other<T2, T>(a.f2[cm], c);
        tmp[re] = -s * (wv.k1 * h1p + wv.k3 * h3p);
        ((T*)a.w)[re] = s * (wv.k3 * h1p - wv.k1 * h3p);
      }
      __syncwarp();   // partner components are read above before anything is overwritten below
      d1_solve<T>(a, md, tmp, tmp, a.sA);
      for (int j = 0; j < ny; j++) {
        long re = (long)j * st + md.e;
        f0[re] = tmp[re] - wv.kk * f1[re];
        f2[re] = ((T*)a.w)[re];
      }
    } break;
  }
}

// ---------------------------------------------------------------------------------
template <typename T>
void fill_pencil_args(chan_handle* h, PArgs<T>& a) {
  typedef typename Vec2<T>::type T2;
  a.phi = (T2*)h->phi; a.g = (T2*)h->g; a.v = (T2*)h->v; a.dv = (T2*)h->dv; a.Rphi = (T2*)h->Rphi; a.Rg = (T2*)h->Rg;
  a.u = (T2*)h->u; a.w = (T2*)h->w; a.ox = (T2*)h->ox; a.oy = (T2*)h->oy; a.oz = (T2*)h->oz;
  a.sA = h->sA; a.sB = h->sB; a.sC = h->sC; a.sP = h->sP;
  a.mean_u = (const T*)h->mean_u; a.sc = h->sc; a.yt = h->yt;
  a.ny = h->d.ny; a.Kz = h->d.Kz; a.ikx0 = h->d.ikx0; a.kxmax = h->d.kxmax; a.Kx = h->d.Kx; a.M = h->d.M;
  // channel.h:53-57: PI2 = 2.0f*3.14159265f, LX = 1.0f*PI2, LZ = 0.5f*PI2 (float arithmetic, widened for double)
  const float PI2 = 2.0f * 3.14159265f, LX = 1.0f * PI2, LZ = 0.5f * PI2;
  a.k1fac = (T)PI2 / (T)LX; a.k3fac = (T)PI2 / (T)LZ;
  const float re = (float)h->cfg.reynolds;                 // const float REYNOLDS (channel.h:61)
  a.nuT = (T)1.0f / (T)re;
  a.nuD = 1.0 / (double)re;
  a.reD = (double)re;
  a.n = 0; a.alpha = a.dseda = a.gammad = a.betha = 0;
  a.op = 0; a.f0 = a.f1 = a.f2 = 0; a.op_betha = 0; a.op_dt = 0; a.use_mean = 1; a.pf_on = h->tune.post_pf; a.dbg_stop = h->tune.post_stop;
}

static const float kAlpha[3] = { 29.0f / 96.0f, -3.0f / 40.0f, 1.0f / 6.0f };     // RK3_kernels.cu:24
static const float kDseda[3] = { 0.0f, -17.0f / 60.0f, -5.0f / 12.0f };           // RK3_kernels.cu:25
static const float kGamma[3] = { 8.0f / 15.0f, 5.0f / 12.0f, 3.0f / 4.0f };       // RK3_kernels.cu:113
static const float kBetha[3] = { 37.0f / 160.0f, 5.0f / 24.0f, 1.0f / 6.0f };     // RK3.c:14

static int chain_block_threads(const chan_handle* h, long nchain) {
  if (h->tune.post_threads > 0) return h->tune.post_threads;
  (void)nchain;
  return 128;      // one-warp blocks deal the chains more evenly but change nothing measurable (both kernels are bound by HBM as a whole)
}
template <typename T>
static int pre_t(chan_handle* h) {
  PArgs<T> a; fill_pencil_args<T>(h, a);
  long nchain = 2 * h->d.M;
  PeerOut po; fill_peer_out(h, po);
  // The warps of these kernels are independent (no shared memory, no block barrier): the block size only sets the granularity
  // at which chains are dealt to the SMs (CHAN_TUNE_POST_THREADS; 32/64/128 measured equal at four grids, round 2).
  const int bt = chain_block_threads(h, nchain);
  pencil_pre_kernel<T><<<(unsigned)((nchain + bt - 1) / bt), bt, 0, h->stream>>>(a, po);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
template <typename T>
static int post_t(chan_handle* h, int n) {
  PArgs<T> a; fill_pencil_args<T>(h, a);
  a.n = n; a.alpha = (T)kAlpha[n]; a.dseda = (T)kDseda[n]; a.gammad = (T)kGamma[n]; a.betha = (T)kBetha[n];
  long nchain = 2 * h->d.M;
  const int bt = h->tune.post_deep ? 32 : chain_block_threads(h, nchain);   // few chains: one warp per block, spread over all SMs
  const unsigned grid = (unsigned)((nchain + bt - 1) / bt);
  if (h->tune.post_deep == 2) pencil_post_roles_kernel<T><<<(unsigned)((nchain + 31) / 32), 96, 0, h->stream>>>(a);
  else if (h->tune.post_deep) pencil_post_kernel<T, true><<<grid, bt, 0, h->stream>>>(a);
  else pencil_post_kernel<T, false><<<grid, bt, 0, h->stream>>>(a);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
template <typename T>
static int op_t(chan_handle* h, int op, void* f0, void* f1, void* f2, double betha, double dt) {
  typedef typename Vec2<T>::type T2;
  PArgs<T> a; fill_pencil_args<T>(h, a);
  a.op = op; a.f0 = (T2*)f0; a.f1 = (T2*)f1; a.f2 = (T2*)f2; a.op_betha = (T)betha; a.op_dt = (T)dt;
  long nchain = 2 * h->d.M;
  unsigned grid = (unsigned)((nchain + 127) / 128);
  if (op == OP_OMEGA) { a.use_mean = 0; a.sc = 0; PeerOut po; memset(&po, 0, sizeof(po)); po.nranks = 1; pencil_pre_kernel<T><<<grid, 128, 0, h->stream>>>(a, po); }
  else op_kernel<T><<<grid, 128, 0, h->stream>>>(a);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}

template void fill_pencil_args<float>(chan_handle*, PArgs<float>&);
template void fill_pencil_args<double>(chan_handle*, PArgs<double>&);

int launch_pencil_pre(chan_handle* h) { return h->prec == 4 ? pre_t<float>(h) : pre_t<double>(h); }
int launch_pencil_post(chan_handle* h, int n) { return h->prec == 4 ? post_t<float>(h, n) : post_t<double>(h, n); }
int launch_op(chan_handle* h, int op, void* f0, void* f1, void* f2, double betha, double dt) {
  return h->prec == 4 ? op_t<float>(h, op, f0, f1, f2, betha, dt) : op_t<double>(h, op, f0, f1, f2, betha, dt);
}

}  // namespace chan
