// Device code shared by the two implementations of the wall-normal stage: kernels_pencil.cu (one
// thread per real chain, sequential sweeps through fp64 scratch) and kernels_pencil_seg.cu (one
// thread per (chain, y-segment), everything in registers).
#pragma once
#include <string.h>
#include "chan_internal.h"

namespace chan {

template <typename T>
struct PArgs {
  typedef typename Vec2<T>::type T2;
  T2 *phi, *g, *v, *dv, *Rphi, *Rg, *u, *w, *ox, *oy, *oz;
  double *sA, *sB, *sC, *sP;
  const T* mean_u;
  Scalars* sc;
  YTabs yt;
  int ny, Kz, ikx0, kxmax, Kx;
  long M;
  T k1fac, k3fac;          // PI2/LX, PI2/LZ in the state type (RK3_kernels.cu:53-54)
  T nuT;                   // 1.0f/REYNOLDS in the state type (RK3_kernels.cu:58)
  double nuD;              // 1.0/REYNOLDS in double (implicitStep_nu_double.cu:133)
  double reD;              // (double)REYNOLDS (bilplacSolver_double.cu:58)
  int n;                   // RK substep
  T alpha, dseda, gammad, betha;
  // generic operator launches
  int op; T2 *f0, *f1, *f2; T op_betha, op_dt; int use_mean; int pf_on, dbg_stop;
};

struct Mode {
  long m; int c; long e;   // mode, component, real chain index
  int kxi, kzi; bool zero_mode;
};

template <typename T>
struct Wave { T k1, k3, kk, rkk; double kkd; T sc; T sp; };   // rkk = 1/kk (1 for the mean mode)   // sc = sign of own component of i*z: -1 (re), +1 (im)

template <typename T>
__device__ __forceinline__ void setup_chain(const PArgs<T>& a, long e, Mode& md, Wave<T>& wv);
template <typename T>
__device__ __forceinline__ bool setup(const PArgs<T>& a, Mode& md, Wave<T>& wv) {
  long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 2 * a.M) return false;
  setup_chain(a, e, md, wv);
  return true;
}
template <typename T>
__device__ __forceinline__ void setup_chain(const PArgs<T>& a, long e, Mode& md, Wave<T>& wv) {
  md.e = e; md.m = e >> 1; md.c = (int)(e & 1);
  int ikx = a.ikx0 + (int)(md.m / a.Kz);
  md.kzi = (int)(md.m % a.Kz);
  md.kxi = ikx <= a.kxmax ? ikx : ikx - a.Kx;
  md.zero_mode = (md.kxi == 0 && md.kzi == 0);
  wv.k1 = a.k1fac * (T)md.kxi;
  wv.k3 = a.k3fac * (T)md.kzi;
  wv.kk = wv.k1 * wv.k1 + wv.k3 * wv.k3;
  wv.rkk = md.zero_mode ? (T)1 : (T)1 / wv.kk;
  double k1d = (double)a.k1fac * (double)md.kxi, k3d = (double)a.k3fac * (double)md.kzi;
  wv.kkd = k1d * k1d + k3d * k3d;
  wv.sc = md.c ? (T)1 : (T)-1;
  wv.sp = -wv.sc;
}

__device__ __forceinline__ int phys_row(int dir, int jj, int ny) { return dir ? ny - 1 - jj : jj; }

template <typename T2, typename T>
__device__ __forceinline__ T own(const T2& z, int c) { return c ? z.y : z.x; }
template <typename T2, typename T>
__device__ __forceinline__ T other(const T2& z, int c) { return c ? z.x : z.y; }

template <typename T> __device__ __forceinline__ double rnd(double x) { return (double)(T)x; }
template <typename T2> __device__ __forceinline__ T2 mk0() { T2 z; z.x = 0; z.y = 0; return z; }

// ---- wall correction of the bi-Laplacian (setBoundaryCond, bilplacSolver_double.cu:6-296)
struct BcFun { double v1, v2, dv1, dv2, ddv1, ddv2; };
// The six functions of bilplacSolver_double.cu:174-217.  Half of the exponentials are fp32 `expf`,
// and the l2 lines take exp(-l1*(1+y)) in their first bracket — mirrored: it defines
// the reference's v.  Row-independent factors (1 +- exp(-2l), their reciprocals) are hoisted into
// BcCoef, and the two threads of a mode (re|im see the same functions) each evaluate two of the four
// exponentials and trade them by shuffle: 4x fewer transcendentals than the formula as written.
struct BcCoef { double l1, l2, D, ip1, im1, ip2, im2; };
__device__ __forceinline__ BcCoef bc_coef(double kkd, double reD, double dt, double betha) {
  BcCoef k;
  k.l1 = sqrt(kkd + reD / (dt * betha));
  k.l2 = sqrt(kkd);
  k.D = 1.0 / (k.l1 * k.l1 - k.l2 * k.l2);
  double e1 = exp(-k.l1 * 2.0), e2 = exp(-k.l2 * 2.0);
  k.ip1 = 1.0 / (1.0 + e1); k.im1 = 1.0 / (1.0 - e1); k.ip2 = 1.0 / (1.0 + e2); k.im2 = 1.0 / (1.0 - e2);
  return k;
}
// c = component of the calling thread; its partner (lane ^ 1) must call with the same y
__device__ __forceinline__ BcFun bc_functions(const BcCoef& k, double y, int c) {
  const double l = c ? k.l2 : k.l1;
  double em = (double)expf((float)(-l * (1.0 - y)));
  double ep = exp(-l * (1.0 + y));
  double om = __shfl_xor_sync(0xffffffffu, em, 1), op = __shfl_xor_sync(0xffffffffu, ep, 1);
  const double E1m = c ? om : em, E1p = c ? op : ep, E2m = c ? em : om, E2p = c ? ep : op;
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

// ---- pointwise pieces ----------------------------------------------------------
// own component of u, w from complex dv, g (calcUWkernel, nonLinear_kernels.cu:53-82)
template <typename T, typename T2>
__device__ __forceinline__ void uw_comp(const Wave<T>& wv, const Mode& md, bool zero_mode, const T2& dvz, const T2& gz, int c,
                                        T& u, T& w) {
  T s = c ? (T)1 : (T)-1;
  T dp = other<T2, T>(dvz, c), gp = other<T2, T>(gz, c);
  // u = i (k1 dv - k3 g)/kk ; w = i (k3 dv + k1 g)/kk, written in the reference's operation order.  In
  // double the six divisions per chain row were the bulk of the `pre` kernels' instructions: multiply by
  // the per-thread reciprocal instead (last-bit difference); the cheap float division stays exact.
  if (sizeof(T) == 8) {
    u = s * (wv.k1 * dp - wv.k3 * gp) * wv.rkk;
    w = s * (wv.k3 * dp + wv.k1 * gp) * wv.rkk;
  } else {
    T kk = zero_mode ? (T)1 : wv.kk;
    u = s * (wv.k1 * dp - wv.k3 * gp) / kk;
    w = s * (wv.k3 * dp + wv.k1 * gp) / kk;
  }
  if (zero_mode) { u = 0; w = 0; }
}


template <typename T>
void fill_pencil_args(chan_handle* h, PArgs<T>& a);

// where the `pre` kernels deliver their rows: locally (a bulk scatter follows when there are several GPUs, the default), or
// straight into the plane owners' receive buffers from inside the kernel (CHAN_TUNE_PRE_DIRECT=1: small packets, slower)
static inline void fill_peer_out(chan_handle* h, PeerOut& po) {
  memset(&po, 0, sizeof(po));
  po.nranks = (h->peer_ready && h->tune.pre_direct) ? h->cfg.size : 1;
  if (po.nranks == 1) return;
  for (int p = 0; p <= h->cfg.size; p++) po.part_j[p] = h->part_j[p];
  for (int p = 0; p < h->cfg.size; p++) {
    po.src_off[p] = (long)(h->part_j[p + 1] - h->part_j[p]) * h->d.ikx0 * h->d.Kz;
    for (int f = 0; f < 6; f++) po.rcv[p][f] = h->peer_rcv[p][f];
  }
}

}  // namespace chan
