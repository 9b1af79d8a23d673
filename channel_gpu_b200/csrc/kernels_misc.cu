// Boundary layout conversion, the mean-flow ODE, the time-step estimate and the
// statistics sums.
#include "chan_internal.h"

namespace chan {

// -------------------------------------------------------------------------------
// reference layout [i][k][j] (j fastest, RK3_kernels.cu:16-18, all NX*NZ modes)
//   <-> compact plane-major [j][ikx*Kz+kz] (kz fastest, active modes only).
// A tiled transpose through shared memory, coalesced on both sides.
// -------------------------------------------------------------------------------
// slab = 1: the dense side is the rank's own slab [Kxl][NZ][NY] in active-kx order (the library's boundary staging
// buffer); slab = 0: the caller's full reference-layout array [NX][NZ][NY].
template <typename T2, bool PACK>
__global__ void relayout_kernel(const T2* __restrict__ src, T2* __restrict__ dst, int nx, int ny, int nz,
                                int Kx, int Kz, int kxmax, int ikx0, int Kxl, int slab) {
  __shared__ T2 tile[32][33];
  const int ikxl = blockIdx.z;                 // local active kx
  const int ikx = ikx0 + ikxl;
  const int kx = ikx <= kxmax ? ikx : ikx - Kx;
  const int i = slab ? ikxl : (kx >= 0 ? kx : kx + nx);
  const int j0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const long M = (long)Kxl * Kz;
  if (PACK) {
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      int k = k0 + r, j = j0 + threadIdx.x;
      if (k < Kz && j < ny) tile[r][threadIdx.x] = src[((size_t)i * nz + k) * ny + j];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      int j = j0 + r, k = k0 + threadIdx.x;
      if (k < Kz && j < ny) dst[(size_t)j * M + (size_t)ikxl * Kz + k] = tile[threadIdx.x][r];
    }
  } else {
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      int j = j0 + r, k = k0 + threadIdx.x;
      if (k < Kz && j < ny) tile[threadIdx.x][r] = src[(size_t)j * M + (size_t)ikxl * Kz + k];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      int k = k0 + r, j = j0 + threadIdx.x;
      if (k < Kz && j < ny) dst[((size_t)i * nz + k) * ny + j] = tile[r][threadIdx.x];
    }
  }
}

template <typename T>
static int relayout_t(chan_handle* h, const void* src, void* dst, bool pack, int slab) {
  typedef typename Vec2<T>::type T2;
  const Dims& d = h->d;
  dim3 grid((d.ny + 31) / 32, (d.Kz + 31) / 32, d.Kxl), block(32, 8);
  if (pack) relayout_kernel<T2, true><<<grid, block, 0, h->stream>>>((const T2*)src, (T2*)dst, d.nx, d.ny, d.nz, d.Kx, d.Kz, d.kxmax, d.ikx0, d.Kxl, slab);
  else {
    // truncated modes come back as exact zeros (dealias.cu:37): clear the dense side first
    CHAN_CUDA(h, cudaMemsetAsync(dst, 0, (size_t)(slab ? d.Kxl : d.nx) * d.nz * d.ny * sizeof(T2), h->stream));
    relayout_kernel<T2, false><<<grid, block, 0, h->stream>>>((const T2*)src, (T2*)dst, d.nx, d.ny, d.nz, d.Kx, d.Kz, d.kxmax, d.ikx0, d.Kxl, slab);
  }
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
int launch_pack(chan_handle* h, const void* s, void* c, int slab) { return h->prec == 4 ? relayout_t<float>(h, s, c, true, slab) : relayout_t<double>(h, s, c, true, slab); }
int launch_unpack(chan_handle* h, const void* c, void* s, int slab) { return h->prec == 4 ? relayout_t<float>(h, c, s, false, slab) : relayout_t<double>(h, c, s, false, slab); }

// -------------------------------------------------------------------------------
// Scalar stage, once per substep after the plane stage: the (kx,kz)=(0,0) mean-flow
// ODE of meanUevol.c (the reference runs it on the host in fp32 with three blocking
// copies per substep) and calcDt (RK3.c:64-109).  State-type
// arithmetic throughout, mirroring the host code statement by statement.
// -------------------------------------------------------------------------------
template <typename T>
struct SArgs {
  typedef typename Vec2<T>::type T2;
  T *u, *N, *uw, *tmp;              // T[NY]; tmp = 6*NY scratch
  const T *ma, *mb;                 // mean_a (NY+2), mean_b (NY)
  const T2* H1;                     // compact field holding H1 (mode (0,0) is element 0 of each row)
  long M; int ny, nx, nz, n;
  int owns_zero;                    // this rank holds kx = 0
  Scalars* sc;
  T alpha, betha, gammad, dseda, reynolds, qvel, cfl;
};

// meanUevol.c:86-114; x = rhs in, solution out; a=lower, b=diag, c=upper (c is destroyed)
template <typename T>
__device__ void tri_inplace(T* x, int n, const T* a, const T* b, T* c) {
  // the recurrences run in registers; the array loads do not depend on them and are hoisted by the unroll
  T cp = c[0] / b[0], xp = x[0] / b[0];
  c[0] = cp; x[0] = xp;
#pragma unroll 4
  for (int i = 1; i < n; i++) {
    T m = (T)(1.0 / (double)(b[i] - a[i] * cp));
    cp = c[i] * m;
    xp = (x[i] - a[i] * xp) * m;
    c[i] = cp; x[i] = xp;
  }
#pragma unroll 4
  for (int i = n - 1; i-- > 0;) { xp = x[i] - c[i] * xp; x[i] = xp; }
}

template <typename T>
__device__ __forceinline__ void w2_coef(T a, T b, T& A, T& B, T& C, T& al, T& be) {   // meanUevol.c:275-283, 355-363
  A = (T)-12.0f * b / (a * a * a - b * b * b - (T)4.0f * a * a * b + (T)4.0f * b * b * a);
  B = (T)12.0f * a / (a * a * a - b * b * b - (T)4.0f * a * a * b + (T)4.0f * b * b * a);
  C = -A - B;
  al = -(-b * b * b - a * b * b + a * a * b) / (a * a * a - (T)4.0f * a * a * b + (T)4.0f * a * b * b - b * b * b);
  be = -(a * a * a + b * a * a - b * b * a) / (a * a * a - (T)4.0f * a * a * b + (T)4.0f * a * b * b - b * b * b);
}

template <typename T>
__global__ void __launch_bounds__(128) scalar_kernel(SArgs<T> a) {
  const int ny = a.ny, tid = threadIdx.x, nt = blockDim.x;
  Scalars* sc = a.sc;
  // work arrays, the mean profile and its increment live in shared memory: the two tridiagonal solves are
  // serial on one thread, and from global memory every dependent access cost an L2 round trip (87 us per call)
  extern __shared__ __align__(16) unsigned char scalar_smem[];
  T* lo = (T*)scalar_smem; T* cd = lo + ny; T* up = cd + ny; T* aux = up + ny; T* nnew = aux + ny; T* su = nnew + ny; T* suw = su + ny;
  const T nx = (T)a.nx, nz = (T)a.nz;
  const T N2 = (T)(a.nx * (2 * a.nz - 2));
  const T DELTA_Y = (T)(2.0f / (float)(ny - 1));
  const T PI2 = (T)(2.0f * 3.14159265f), LX = (T)1.0f * PI2, LZ = (T)0.5f * PI2, LY = (T)2.0f;

  // ---- calcDt after substep 0 (RK3.c:64-109); maxima over valid points only (Q7)
  if (a.n == 0 && tid == 0) {
    double um[3];
    for (int c = 0; c < 3; c++) um[c] = __longlong_as_double((long long)sc->umax_bits[c]);
    sc->umax[0] = um[0]; sc->umax[1] = um[1]; sc->umax[2] = um[2];
    const T pi = (T)acosf(-1.0f);
    T c = (nx / ((T)2.0f * pi * LX)) * fabs((T)um[0] / N2) + fabs((T)um[1] / N2) / DELTA_Y + (nx / ((T)2.0f * pi * LZ)) * fabs((T)um[2] / N2);
    T dtc = a.cfl / c;
    T third = (T)1.0f / (T)3.0f;
    T dtv = a.cfl * a.reynolds / ((T)1.0f / (DELTA_Y * DELTA_Y) + (third * nx / LX) * (third * nx / LX) + (third * nz / LZ) * (third * nz / LZ));
    T dt = fmin(dtc, dtv);
    sc->dt_c = (double)dtc; sc->dt_v = (double)dtv;
    sc->dt_next = sc->fixed_dt > 0.0 ? (double)(T)sc->fixed_dt : (double)dt;
  }
  if (!a.owns_zero) return;
  const T dt = (T)sc->dt;
  for (int j = tid; j < ny; j += nt) su[j] = a.u[j];
  __syncthreads();

  // ---- meanURKstep_1 (meanUevol.c:439-464): uw = (alpha+betha) nu d2u/dy2 + dseda N_old
  for (int j = tid; j < ny; j += nt) {
    T A, B, C, al, be; w2_coef<T>(a.ma[j], a.mb[j], A, B, C, al, be);
    up[j] = al; cd[j] = (T)1.0f; lo[j] = be;
    if (j >= 1 && j < ny - 1) aux[j] = A * su[j + 1] + C * su[j] + B * su[j - 1];
    nnew[j] = a.H1[(size_t)j * a.M].x;                      // readNmean (meanUevol.c:127-133)
  }
  __syncthreads();
  if (tid == 0) {
    lo[0] = (T)0.0f; up[ny - 1] = (T)0.0f;
    {
      T a0 = a.ma[0], b0 = a.ma[ny];
      up[0] = (a0 + b0) / ((T)2.0f * a0 - b0); cd[0] = (T)1.0f;
      T A = (T)6.0f / ((a0 - b0) * ((T)2.0f * a0 - b0));
      T B = (T)-6.0f * a0 / ((a0 * b0 - b0 * b0) * ((T)2.0f * a0 - b0));
      T E = -A - B;
      aux[0] = E * su[0] + A * su[1] + B * su[2];
    }
    {
      T aN = a.mb[ny - 1], bN = a.ma[ny + 1];
      lo[ny - 1] = (aN + bN) / ((T)2.0f * aN - bN); cd[ny - 1] = (T)1.0f;
      T A = (T)6.0f / ((aN - bN) * ((T)2.0f * aN - bN));
      T B = (T)-6.0f * aN / ((aN * bN - bN * bN) * ((T)2.0f * aN - bN));
      T E = -A - B;
      aux[ny - 1] = E * su[ny - 1] + A * su[ny - 2] + B * su[ny - 3];
    }
    tri_inplace<T>(aux, ny, lo, cd, up);
  }
  __syncthreads();
  const T nu = (T)1.0f / a.reynolds;
  for (int j = tid; j < ny; j += nt) {
    T uw = (a.alpha + a.betha) * nu * aux[j] + a.dseda * a.N[j];
    a.N[j] = nnew[j];
    // ---- meanURKstep_2 (meanUevol.c:466-487)
    suw[j] = dt * (uw + a.gammad * nnew[j]);
  }
  __syncthreads();
  // implicitStepMeanU (meanUevol.c:347-437)
  for (int j = tid; j < ny; j += nt) {
    T A, B, C, al, be; w2_coef<T>(a.ma[j], a.mb[j], A, B, C, al, be);
    lo[j] = be - nu * dt * a.betha * B;
    cd[j] = (T)1.0f - nu * dt * a.betha * C;
    up[j] = al - nu * dt * a.betha * A;
    if (j == 0 || j == ny - 1) { lo[j] = 0; cd[j] = (T)1.0f; up[j] = 0; }
    if (j == 1) lo[j] = 0;
    if (j == ny - 2) up[j] = 0;
    aux[j] = (j == 0 || j == ny - 1) ? (T)0.0f : (al * suw[j + 1] + suw[j] + be * suw[j - 1]);
  }
  __syncthreads();
  if (tid == 0) {
    tri_inplace<T>(aux, ny, lo, cd, up);
    for (int j = 0; j < ny; j++) { suw[j] = aux[j]; su[j] = su[j] + aux[j]; }
    // forcing (meanUevol.c:201-221): constant flux; the integral starts at j = 10 (Q12)
    T Qt = (T)0.0f;
    for (int j = 10; j < ny - 1; j++) Qt += (T)0.5f * (su[j + 1] + su[j]) * a.ma[j];
    for (int j = 1; j < ny - 1; j++) su[j] = su[j] + (a.qvel * N2 - Qt) / LY;
    // readUtau (meanUevol.c:178-199)
    for (int c = 0; c < 2; c++) {
      T d = (T)sc->dudy_wall[c] / N2;
      sc->utau[c] = (double)(T)sqrt((T)1.0f / a.reynolds * fabs(d));
    }
  }
  __syncthreads();
  for (int j = tid; j < ny; j += nt) { a.u[j] = su[j]; a.uw[j] = suw[j]; }
}

static const float kAlpha[3] = { 29.0f / 96.0f, -3.0f / 40.0f, 1.0f / 6.0f };
static const float kDseda[3] = { 0.0f, -17.0f / 60.0f, -5.0f / 12.0f };
static const float kGamma[3] = { 8.0f / 15.0f, 5.0f / 12.0f, 3.0f / 4.0f };
static const float kBetha[3] = { 37.0f / 160.0f, 5.0f / 24.0f, 1.0f / 6.0f };

template <typename T>
static int scalar_t(chan_handle* h, int n) {
  typedef typename Vec2<T>::type T2;
  SArgs<T> a;
  a.u = (T*)h->mean_u; a.N = (T*)h->mean_N; a.uw = (T*)h->mean_uw; a.tmp = (T*)h->mean_tmp;
  a.ma = (const T*)h->mean_a; a.mb = (const T*)h->mean_b;
  a.H1 = (const T2*)h->ox; a.M = h->d.M; a.ny = h->d.ny; a.nx = h->d.nx; a.nz = h->d.nz; a.n = n;
  a.owns_zero = (h->d.ikx0 == 0);
  a.sc = h->sc;
  a.alpha = (T)kAlpha[n]; a.betha = (T)kBetha[n]; a.gammad = (T)kGamma[n]; a.dseda = (T)kDseda[n];
  a.reynolds = (T)(float)h->cfg.reynolds; a.qvel = (T)(float)h->cfg.qvelocity; a.cfl = (T)(float)h->cfg.cfl;
  const size_t sm = (size_t)7 * h->d.ny * sizeof(T);
  if (sm > 48 * 1024) CHAN_CUDA(h, cudaFuncSetAttribute(scalar_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  scalar_kernel<T><<<1, 128, sm, h->stream>>>(a);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
int launch_scalar_stage(chan_handle* h, int n) { return h->prec == 4 ? scalar_t<float>(h, n) : scalar_t<double>(h, n); }

// RK3.c:179-180: dt lags one step; maxima restart for the next step's calcDt
__global__ void end_step_kernel(Scalars* sc) {
  sc->time += sc->dt;
  sc->dt = sc->dt_next;
  sc->counter += 1;
  sc->umax_bits[0] = sc->umax_bits[1] = sc->umax_bits[2] = 0ull;
}
int launch_end_step(chan_handle* h) {
  end_step_kernel<<<1, 1, 0, h->stream>>>(h->sc);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}

// -------------------------------------------------------------------------------
// calcSt (statistics.cu:161-243): per-plane sums of -2 v u, 2 u^2, 2 v^2, 2 w^2 over the
// half spectrum (kz = 0 weighted 1/2), re and im parts added as the reference's
// DC-bin trick does (fft.c:239-269).  u, w are rebuilt from (dv, g) on the fly.
// -------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) stats_kernel(const typename Vec2<T>::type* dv, const typename Vec2<T>::type* g,
                                                    const typename Vec2<T>::type* v, double* out, long M, int ny, int Kz,
                                                    int ikx0, int kxmax, int Kx, T k1fac, T k3fac, T n2) {
  typedef typename Vec2<T>::type T2;
  const int j = blockIdx.x;
  double acc[4] = { 0, 0, 0, 0 };
  for (long m = threadIdx.x; m < M; m += blockDim.x) {
    int ikx = ikx0 + (int)(m / Kz), kz = (int)(m % Kz);
    int kx = ikx <= kxmax ? ikx : ikx - Kx;
    T k1 = k1fac * (T)kx, k3 = k3fac * (T)kz, kk = k1 * k1 + k3 * k3;
    T2 d = dv[(size_t)j * M + m], gg = g[(size_t)j * M + m], vv = v[(size_t)j * M + m];
    T2 u, w;
    if (kx == 0 && kz == 0) { u.x = u.y = w.x = w.y = 0; }
    else {
      u.x = -(k1 * d.y - k3 * gg.y) / kk; u.y = (k1 * d.x - k3 * gg.x) / kk;
      w.x = -(k3 * d.y + k1 * gg.y) / kk; w.y = (k3 * d.x + k1 * gg.x) / kk;
    }
    u.x /= n2; u.y /= n2; w.x /= n2; w.y /= n2; vv.x /= n2; vv.y /= n2;
    T wt = kz == 0 ? (T)0.5f : (T)1.0f;
    acc[0] += (double)((T)-2.0f * vv.x * u.x * wt) + (double)((T)-2.0f * vv.y * u.y * wt);
    acc[1] += (double)((T)2.0f * u.x * u.x * wt) + (double)((T)2.0f * u.y * u.y * wt);
    acc[2] += (double)((T)2.0f * vv.x * vv.x * wt) + (double)((T)2.0f * vv.y * vv.y * wt);
    acc[3] += (double)((T)2.0f * w.x * w.x * wt) + (double)((T)2.0f * w.y * w.y * wt);
  }
  __shared__ double red[4][8];
  for (int c = 0; c < 4; c++) {
    double x = acc[c];
    for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) == 0) red[c][threadIdx.x >> 5] = x;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double x = 0; for (int wi = 0; wi < (blockDim.x >> 5); wi++) x += red[threadIdx.x][wi];
    out[(size_t)threadIdx.x * ny + j] = x;
  }
}

template <typename T>
static int stats_t(chan_handle* h, double* out) {
  typedef typename Vec2<T>::type T2;
  const float PI2 = 2.0f * 3.14159265f, LX = 1.0f * PI2, LZ = 0.5f * PI2;
  const Dims& d = h->d;
  stats_kernel<T><<<d.ny, 256, 0, h->stream>>>((const T2*)h->dv, (const T2*)h->g, (const T2*)h->v, out, d.M, d.ny, d.Kz,
                                               d.ikx0, d.kxmax, d.Kx, (T)PI2 / (T)LX, (T)PI2 / (T)LZ, (T)(d.nx * d.nzp));
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
int launch_stats(chan_handle* h, double* out) { return h->prec == 4 ? stats_t<float>(h, out) : stats_t<double>(h, out); }

}  // namespace chan
