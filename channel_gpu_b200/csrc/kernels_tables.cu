// One-time tables: wall-normal mesh coefficients, folded LU of the compact
// first-derivative matrix, FFT twiddles and radix plans.
//
// The mesh expressions are evaluated ON THE DEVICE with CUDA's tanh, in the same
// mixed precision as the reference's kernels (product j*DELTA_Y in the state
// type, tanh in double: channel.h:21,50-51; derivatives_nu_double.cu:60-65,
// 85-113, 227-274; hemholzt_nu_double.cu:136-145), so coefficients agree with a
// reference build to the last bit instead of to 1 ulp of a host libm.  The
// reference recomputes them (4 tanh per thread) in every kernel launch; here
// they are built once.
#include <vector>
#include <math.h>
#include <stdio.h>
#include <stdarg.h>
#include "chan_internal.h"
#include "fft_core.cuh"

namespace chan {

int set_err(chan_handle* h, int code, const char* fmt, ...) {
  if (h) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(h->err, sizeof(h->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

template <typename T>
__device__ __forceinline__ double fmesh_dev(int j, T dy) {   // Fmesh(j*DELTA_Y-1.0)
  T p = (T)j * dy;
  double x = (double)p - 1.0;
  return tanh(2.0 * x) / tanh(2.0);
}
template <typename T>
__device__ __forceinline__ double fmesh_host_style(int j, T dy) {   // Fmesh(j*DELTA_Y-1.0f), meanUevol.c:207
  T p = (T)j * dy - (T)1.0f;
  double x = (double)p;
  return tanh(2.0 * x) / tanh(2.0);
}

// raw[c*ny + j], c = 0..11: y A1 B1 C1 al1 be1 A2 B2 C2 al2 be2 (unused);  wall[8]; mean a,b in T
template <typename T>
__global__ void mesh_kernel(int ny, T dy, double* raw, double* wall, T* mean_a, T* mean_b) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= ny) return;
  double yj = fmesh_dev<T>(j, dy);
  double a = fmesh_dev<T>(j + 1, dy) - yj;
  double b = fmesh_dev<T>(j - 1, dy) - yj;
  raw[0 * ny + j] = yj;
  {
    double A = 2.0 * (2.0 * a * b * b - b * b * b) / ((a * a - a * b) * (a - b) * (a - b));
    double B = 2.0 * (2.0 * b * a * a - a * a * a) / ((b * b - a * b) * (a - b) * (a - b));
    raw[1 * ny + j] = A; raw[2 * ny + j] = B; raw[3 * ny + j] = -A - B;
    raw[4 * ny + j] = b * b / ((a - b) * (a - b));
    raw[5 * ny + j] = a * a / ((a - b) * (a - b));
  }
  {
    double A = -12.0 * b / (a * a * a - b * b * b - 4.0 * a * a * b + 4.0 * b * b * a);
    double B = 12.0 * a / (a * a * a - b * b * b - 4.0 * a * a * b + 4.0 * b * b * a);
    raw[6 * ny + j] = A; raw[7 * ny + j] = B; raw[8 * ny + j] = -A - B;
    raw[9 * ny + j] = -(-b * b * b - a * b * b + a * a * b) / (a * a * a - 4.0 * a * a * b + 4.0 * a * b * b - b * b * b);
    raw[10 * ny + j] = -(a * a * a + b * a * a - b * b * a) / (a * a * a - 4.0 * a * a * b + 4.0 * a * b * b - b * b * b);
  }
  mean_a[j] = (T)(fmesh_host_style<T>(j + 1, dy) - fmesh_host_style<T>(j, dy));
  mean_b[j] = (T)(fmesh_host_style<T>(j - 1, dy) - fmesh_host_style<T>(j, dy));
  if (j == 0) {     // two-point-away differences of the wall closures (meanUevol.c:262,275,307,323)
    mean_a[ny] = (T)(fmesh_host_style<T>(2, dy) - fmesh_host_style<T>(0, dy));
    mean_a[ny + 1] = (T)(fmesh_host_style<T>(ny - 3, dy) - fmesh_host_style<T>(ny - 1, dy));
  }
  if (j == 0) {
    double a0 = fmesh_dev<T>(1, dy) - fmesh_dev<T>(0, dy);
    double b0 = fmesh_dev<T>(2, dy) - fmesh_dev<T>(1, dy);
    wall[0] = -(3.0 * a0 + 2.0 * b0) / (a0 * a0 + a0 * b0);
    wall[1] = ((a0 + b0) * (2.0 * b0 - a0)) / (a0 * b0 * b0);
    wall[2] = a0 * a0 / (b0 * b0 * a0 + b0 * b0 * b0);
    wall[3] = (a0 + b0) / b0;
  }
  if (j == ny - 1) {
    double aN = fmesh_dev<T>(j - 1, dy) - fmesh_dev<T>(j, dy);
    double bN = fmesh_dev<T>(j - 2, dy) - fmesh_dev<T>(j - 1, dy);
    wall[4] = -(3.0 * aN + 2.0 * bN) / (aN * aN + aN * bN);
    wall[5] = ((aN + bN) * (2.0 * bN - aN)) / (aN * bN * bN);
    wall[6] = aN * aN / (bN * bN * aN + bN * bN * bN);
    wall[7] = (aN + bN) / bN;
  }
}

static int perm_pos(const FftPlan& p, int n, int s) {   // where DIF leaves output n / DIT expects input n
  if (s == 0) return 0;
  int r = p.radix[s - 1];
  return (n % r) * p.span[s - 1] + perm_pos(p, n / r, s - 1);
}
int perm_pos_of(const FftPlan& p, int n) { return perm_pos(p, n, p.nstage); }
static int upload_perms(chan_handle* h, FftPlan& p, int** dinv, int** dpos) {
  std::vector<int> inv(p.n), pos(p.n);
  for (int n = 0; n < p.n; n++) { pos[n] = perm_pos(p, n, p.nstage); inv[pos[n]] = n; }
  CHAN_CUDA(h, cudaMalloc(dinv, sizeof(int) * p.n));
  CHAN_CUDA(h, cudaMalloc(dpos, sizeof(int) * p.n));
  CHAN_CUDA(h, cudaMemcpy(*dinv, inv.data(), sizeof(int) * p.n, cudaMemcpyHostToDevice));
  CHAN_CUDA(h, cudaMemcpy(*dpos, pos.data(), sizeof(int) * p.n, cudaMemcpyHostToDevice));
  p.inv = *dinv; p.pos = *dpos;
  return CHAN_OK;
}

// Radix schedule: fewest stages (>= 2) out of {16,14,12,10,8,7,6,5,4,3,2}, then the largest smallest radix.
// (5, 7, 10 and 14 only ever divide lengths with such a factor: the schedules of the 2^a 3^b lengths are what they were.)
// small_last: the last stage is fused with the pointwise product in z and must be 4, 3 or 2
// (6 lines x r_L complex values live in registers).
static bool plan_search(int n, int depth, int maxdepth, int* cur, int* best, int* bestmin, bool small_last, int max_radix) {
  static const int cand[11] = { 16, 14, 12, 10, 8, 7, 6, 5, 4, 3, 2 };
  if (n == 1) {
    if (depth != maxdepth) return false;
    int mn = 99; bool has_small = false;
    for (int i = 0; i < depth; i++) { mn = cur[i] < mn ? cur[i] : mn; if (cur[i] <= 4) has_small = true; }
    if (small_last && !has_small) return false;
    if (mn > *bestmin) { *bestmin = mn; for (int i = 0; i < depth; i++) best[i] = cur[i]; }
    return true;
  }
  if (depth >= maxdepth) return false;
  bool found = false;
  for (int c = 0; c < 11; c++) {
    int r = cand[c];
    if (n % r || r > max_radix) continue;
    if (depth > 0 && r > cur[depth - 1]) continue;      // non-increasing: each multiset once
    cur[depth] = r;
    if (plan_search(n / r, depth + 1, maxdepth, cur, best, bestmin, small_last, max_radix)) found = true;
  }
  return found;
}
void make_plan(FftPlan* p, int n, bool small_last, int max_radix) {
  int cur[16], best[16], bestmin = 0, ns = 0;
  for (int d = 2; d <= 12 && !ns; d++) { bestmin = 0; if (plan_search(n, 0, d, cur, best, &bestmin, small_last, max_radix)) ns = d; }
  p->n = ns ? n : -1; p->nstage = ns;
  if (!ns) return;
  if (small_last) {     // move the largest radix <= 4 to the end
    int li = -1; for (int i = 0; i < ns; i++) if (best[i] <= 4 && (li < 0 || best[i] > best[li])) li = i;
    int v = best[li]; for (int i = li; i + 1 < ns; i++) best[i] = best[i + 1]; best[ns - 1] = v;
  }
  int span = 1;
  for (int i = 0; i < ns; i++) { p->radix[i] = best[i]; p->span[i] = span; p->span_magic[i] = fft::make_fastdiv(span).M; span *= best[i]; }
  p->r_first = ns ? p->radix[0] : 0; p->r_last = ns ? p->radix[ns - 1] : 0;
  p->rfirst_magic = ns ? fft::make_fastdiv(p->radix[0]).M : 0;
  for (int i = 0; i < ns; i++) p->tws[i] = n / (p->span[i] * p->radix[i]);
}

template <typename T>
static int upload_twiddles(chan_handle* h, int n, void** dev) {
  typedef typename Vec2<T>::type T2;
  std::vector<T2> w(n);
  for (int k = 0; k < n; k++) {
    // exact symmetries via octant reduction keep |w| = 1 to the last bit where it matters
    long double ang = 2.0L * 3.141592653589793238462643383279502884L * (long double)k / (long double)n;
    w[k].x = (T)cosl(ang); w[k].y = (T)(-sinl(ang));
  }
  CHAN_CUDA(h, cudaMalloc(dev, sizeof(T2) * n));
  CHAN_CUDA(h, cudaMemcpy(*dev, w.data(), sizeof(T2) * n, cudaMemcpyHostToDevice));
  return CHAN_OK;
}

template <typename T>
static int build_tables_t(chan_handle* h) {
  const int ny = h->d.ny;
  // DELTA_Y = 2.0f/(NY-1) is a FLOAT quotient, widened for the double state (channel.h:51)
  const T dy = (T)(2.0f / (float)(ny - 1));
  double *raw, *wall; T *ma, *mb;
  CHAN_CUDA(h, cudaMalloc(&raw, sizeof(double) * 12 * ny));
  CHAN_CUDA(h, cudaMalloc(&wall, sizeof(double) * 8));
  CHAN_CUDA(h, cudaMalloc(&ma, sizeof(T) * (ny + 2)));
  CHAN_CUDA(h, cudaMalloc(&mb, sizeof(T) * ny));
  mesh_kernel<T><<<(ny + 127) / 128, 128, 0, h->stream>>>(ny, dy, raw, wall, ma, mb);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  std::vector<double> r(12 * ny), wl(8);
  CHAN_CUDA(h, cudaMemcpyAsync(r.data(), raw, sizeof(double) * 12 * ny, cudaMemcpyDeviceToHost, h->stream));
  CHAN_CUDA(h, cudaMemcpyAsync(wl.data(), wall, sizeof(double) * 8, cudaMemcpyDeviceToHost, h->stream));
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  cudaFree(raw); cudaFree(wall);
  h->mean_a = ma; h->mean_b = mb;
  const double *Y = &r[0], *A1 = &r[ny], *B1 = &r[2 * ny], *C1 = &r[3 * ny], *al1 = &r[4 * ny], *be1 = &r[5 * ny];
  const double *A2 = &r[6 * ny], *B2 = &r[7 * ny], *C2 = &r[8 * ny], *al2 = &r[9 * ny], *be2 = &r[10 * ny];

  // 21 arrays: y | dir0: eP eC eN eL cp rP rN wP wC wN | dir1: same
  std::vector<double> t(21 * (size_t)ny);
  for (int j = 0; j < ny; j++) t[j] = Y[j];
  for (int dir = 0; dir < 2; dir++) {
    double* eP = &t[(1 + dir * 10 + 0) * (size_t)ny]; double* eC = eP + ny; double* eN = eC + ny;
    double* eL = eN + ny; double* cp = eL + ny;
    double* rP = cp + ny; double* rN = rP + ny; double* wP = rN + ny; double* wC = wP + ny; double* wN = wC + ny;
    const double* wf = dir ? &wl[4] : &wl[0];    // wall met first in this sweep
    const double* wb = dir ? &wl[0] : &wl[4];    // wall met last
    double cprev = 0.0;
    for (int jj = 0; jj < ny; jj++) {
      int j = dir ? ny - 1 - jj : jj;
      double lo, up, cP, cC, cN;
      if (jj == 0)            { lo = 0.0;  up = wf[3]; cC = wf[0]; cN = wf[1]; cP = wf[2]; }
      else if (jj == ny - 1)  { lo = wb[3]; up = 0.0;  cC = wb[0]; cP = wb[1]; cN = wb[2]; }
      else {
        lo = dir ? al1[j] : be1[j]; up = dir ? be1[j] : al1[j];
        cP = dir ? A1[j] : B1[j];   cN = dir ? B1[j] : A1[j]; cC = C1[j];
      }
      double m = 1.0 / (1.0 - lo * cprev);
      eP[jj] = cP * m; eC[jj] = cC * m; eN[jj] = cN * m; eL[jj] = lo * m; cp[jj] = up * m;
      cprev = cp[jj];
      rP[jj] = dir ? al2[j] : be2[j]; rN[jj] = dir ? be2[j] : al2[j];
      wP[jj] = dir ? A2[j] : B2[j];   wN[jj] = dir ? B2[j] : A2[j]; wC[jj] = C2[j];
    }
  }
  CHAN_CUDA(h, cudaMalloc(&h->ytab_store, sizeof(double) * t.size()));
  CHAN_CUDA(h, cudaMemcpy(h->ytab_store, t.data(), sizeof(double) * t.size(), cudaMemcpyHostToDevice));
  const double* base = h->ytab_store;
  h->yt.y = base;
  for (int dir = 0; dir < 2; dir++) {
    const double* p = base + (1 + dir * 10) * (size_t)ny;
    h->yt.d1[dir].eP = p; h->yt.d1[dir].eC = p + ny; h->yt.d1[dir].eN = p + 2 * ny;
    h->yt.d1[dir].eL = p + 3 * ny; h->yt.d1[dir].cp = p + 4 * ny;
    h->yt.d2[dir].rP = p + 5 * ny; h->yt.d2[dir].rN = p + 6 * ny;
    h->yt.d2[dir].wP = p + 7 * ny; h->yt.d2[dir].wC = p + 8 * ny; h->yt.d2[dir].wN = p + 9 * ny;
  }
  make_plan(&h->px, h->d.nx, false, h->tune.max_radix);
  make_plan(&h->pz, h->d.nzp, true, h->tune.max_radix);
  if (h->px.n < 0 || h->pz.n < 0)
    return set_err(h, CHAN_ERR_ARG, "NX=%d and 2NZ-2=%d must factor into 2s, 3s, 5s and 7s", h->d.nx, h->d.nzp);
  int rc = upload_twiddles<T>(h, h->d.nx, &h->tw_x); if (rc) return rc;
  rc = upload_twiddles<T>(h, h->d.nzp, &h->tw_z); if (rc) return rc;
  h->px.tw = h->tw_x; h->pz.tw = h->tw_z;
  rc = upload_perms(h, h->px, &h->invx, &h->posx); if (rc) return rc;
  rc = upload_perms(h, h->pz, &h->invz, &h->posz); if (rc) return rc;
  return CHAN_OK;
}

int build_tables(chan_handle* h) {
  return h->prec == 4 ? build_tables_t<float>(h) : build_tables_t<double>(h);
}

}  // namespace chan
