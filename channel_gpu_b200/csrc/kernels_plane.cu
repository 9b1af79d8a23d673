// Wall-parallel (x,z) stages of the RK3 substep: the 6 inverse and 3 forward
// 2-D FFTs of convolution.c:24-41, the 3/2-rule zero-pad / truncate
// (dealias.cu:9-45 seen from the other side: only active modes are stored, the
// pad happens in the load and the truncation in the store), the physical-space
// product u x omega (rotorkernel, convolution_kernels.cu:69-148), the velocity
// maxima of calcUmax (fft.c:143-169) and imposeSymetry (imposeSymetry.c:5-18).
//
// All transforms are hand-written shared-memory mixed-radix (4,2,3) FFTs.  cuFFT
// is not linked.  Design points:
//  * PERMUTATION-FREE.  Inverse transforms run decimation-in-frequency (natural
//    order in, digit-reversed order out), forward transforms decimation-in-time
//    with the same radix schedule (digit-reversed in, natural out).  Physical
//    space is only ever used pointwise (products, maxima), so x and z stay
//    digit-reversed there and no reorder pass or index table exists on the hot
//    path.
//  * BATCH-FASTEST shared-memory layout s[pos][line]: the lanes of a warp work
//    on the same butterfly of neighbouring lines, so every stage, whatever its
//    stride, is bank-conflict free.
//  * TWO-FOR-ONE real transforms in z: the same field at two neighbouring x is
//    packed as one complex line (a + i b), halving the z-FFT work; Hermitian
//    unpacking happens in the store.
//  * The physical fields never touch HBM: C2R x6 -> u x omega -> R2C x3 happen
//    inside one kernel per (plane, x-pair).  The x-transformed intermediates
//    Q[f][plane][x][kz] are produced and consumed per PLANE BATCH sized to stay
//    resident in the 126 MB L2.
#include "chan_internal.h"

namespace chan {

template <typename T2> __device__ __forceinline__ T2 mk(decltype(T2().x) a, decltype(T2().x) b) { T2 r; r.x = a; r.y = b; return r; }
template <typename T2> __device__ __forceinline__ T2 cadd(T2 a, T2 b) { return mk<T2>(a.x + b.x, a.y + b.y); }
template <typename T2> __device__ __forceinline__ T2 csub(T2 a, T2 b) { return mk<T2>(a.x - b.x, a.y - b.y); }
// a * w (SIGN<0) or a * conj(w) (SIGN>0)
template <typename T2, int SIGN> __device__ __forceinline__ T2 ctw(T2 a, T2 w) {
  if (SIGN < 0) return mk<T2>(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x);
  return mk<T2>(a.x * w.x + a.y * w.y, a.y * w.x - a.x * w.y);
}
// multiply by SIGN * i
template <typename T2, int SIGN> __device__ __forceinline__ T2 muli(T2 a) {
  return SIGN > 0 ? mk<T2>(-a.y, a.x) : mk<T2>(a.y, -a.x);
}

template <typename T2, int SIGN> __device__ __forceinline__ void bf2(T2* a) {
  T2 t = a[0]; a[0] = cadd(t, a[1]); a[1] = csub(t, a[1]);
}
template <typename T2, int SIGN> __device__ __forceinline__ void bf4(T2* a) {
  T2 t0 = cadd(a[0], a[2]), t1 = csub(a[0], a[2]), t2 = cadd(a[1], a[3]), t3 = muli<T2, SIGN>(csub(a[1], a[3]));
  a[0] = cadd(t0, t2); a[2] = csub(t0, t2); a[1] = cadd(t1, t3); a[3] = csub(t1, t3);
}
template <typename T2, int SIGN> __device__ __forceinline__ void bf3(T2* a) {
  typedef decltype(T2().x) T;
  const T h = (T)0.86602540378443864676372317075294;
  T2 t1 = cadd(a[1], a[2]);
  T2 t2 = mk<T2>(a[0].x - (T)0.5 * t1.x, a[0].y - (T)0.5 * t1.y);
  T2 d = csub(a[1], a[2]);
  T2 t3 = muli<T2, SIGN>(mk<T2>(h * d.x, h * d.y));
  a[0] = cadd(a[0], t1); a[1] = cadd(t2, t3); a[2] = csub(t2, t3);
}

// One radix-R stage over LINES interleaved lines (element (pos, line) at s[pos*STRIDE + line]).
template <typename T2, int R, int LINES, int STRIDE, bool DIF, int SIGN>
__device__ __forceinline__ void fft_stage(T2* s, int n, int m, const T2* __restrict__ tw, int tid, int nthr) {
  const int nb = n / R;
  const int tws = n / (m * R);
  for (int b = tid; b < nb * LINES; b += nthr) {
    int line = b % LINES, bf = b / LINES;
    int g = bf / m, k = bf - g * m;
    T2* p = s + (size_t)(g * m * R + k) * STRIDE + line;
    T2 a[R];
#pragma unroll
    for (int q = 0; q < R; q++) a[q] = p[(size_t)q * m * STRIDE];
    if (!DIF) {
#pragma unroll
      for (int q = 1; q < R; q++) a[q] = ctw<T2, SIGN>(a[q], __ldg(tw + q * k * tws));
    }
    if (R == 2) bf2<T2, SIGN>(a); else if (R == 3) bf3<T2, SIGN>(a); else bf4<T2, SIGN>(a);
    if (DIF) {
#pragma unroll
      for (int q = 1; q < R; q++) a[q] = ctw<T2, SIGN>(a[q], __ldg(tw + q * k * tws));
    }
#pragma unroll
    for (int q = 0; q < R; q++) p[(size_t)q * m * STRIDE] = a[q];
  }
  __syncthreads();
}

template <typename T2, int LINES, int STRIDE, bool DIF, int SIGN>
__device__ __forceinline__ void fft_smem(T2* s, const FftPlan& pl, int tid, int nthr) {
  const T2* tw = (const T2*)pl.tw;
  for (int i = 0; i < pl.nstage; i++) {
    int t = DIF ? pl.nstage - 1 - i : i;
    int r = pl.radix[t], m = pl.span[t];
    if (r == 4) fft_stage<T2, 4, LINES, STRIDE, DIF, SIGN>(s, pl.n, m, tw, tid, nthr);
    else if (r == 2) fft_stage<T2, 2, LINES, STRIDE, DIF, SIGN>(s, pl.n, m, tw, tid, nthr);
    else fft_stage<T2, 3, LINES, STRIDE, DIF, SIGN>(s, pl.n, m, tw, tid, nthr);
  }
}

// -------------------------------------------------------------------------------
template <typename T>
struct XArgs {
  typedef typename Vec2<T>::type T2;
  const T2* src[6]; T2* dst[6];   // compact plane-major fields
  T2* Q;                          // [f][pb][NX][Kz]
  FftPlan px;
  int nx, Kx, Kz, kxmax, j_start, pb; long M;
  T scale;                        // forward only: divide by this when != 0 (normalizekernel, dealias.cu:72-97)
};

extern __shared__ __align__(16) unsigned char smem_raw[];

// x inverse: compact spectral rows -> zero-padded NX line -> DIF inverse -> Q (x digit-reversed)
template <typename T, int TK>
__global__ void __launch_bounds__(256) xfft_inv_kernel(XArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  T2* s = (T2*)smem_raw;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int kz0 = blockIdx.x * TK, jb = blockIdx.y, f = blockIdx.z;
  const int j = a.j_start + jb;
  const T2* src = a.src[f] + (size_t)j * a.M;
  for (int idx = tid; idx < a.nx * TK; idx += nthr) {
    int pos = idx / TK, t = idx - pos * TK, kz = kz0 + t;
    int ikx = pos <= a.kxmax ? pos : (pos >= a.nx - a.kxmax ? pos - a.nx + a.Kx : -1);
    T2 v = mk<T2>(0, 0);
    if (ikx >= 0 && kz < a.Kz) v = src[(size_t)ikx * a.Kz + kz];
    s[idx] = v;
  }
  __syncthreads();
  fft_smem<T2, TK, TK, true, +1>(s, a.px, tid, nthr);
  T2* q = a.Q + ((size_t)(f * a.pb + jb) * a.nx) * a.Kz;
  for (int idx = tid; idx < a.nx * TK; idx += nthr) {
    int pos = idx / TK, t = idx - pos * TK, kz = kz0 + t;
    if (kz < a.Kz) q[(size_t)pos * a.Kz + kz] = s[idx];
  }
}

// x forward: Q (x digit-reversed) -> DIT forward -> truncated compact spectral rows
template <typename T, int TK>
__global__ void __launch_bounds__(256) xfft_fwd_kernel(XArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  T2* s = (T2*)smem_raw;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int kz0 = blockIdx.x * TK, jb = blockIdx.y, f = blockIdx.z;
  const int j = a.j_start + jb;
  const T2* q = a.Q + ((size_t)(f * a.pb + jb) * a.nx) * a.Kz;
  for (int idx = tid; idx < a.nx * TK; idx += nthr) {
    int pos = idx / TK, t = idx - pos * TK, kz = kz0 + t;
    s[idx] = kz < a.Kz ? q[(size_t)pos * a.Kz + kz] : mk<T2>(0, 0);
  }
  __syncthreads();
  fft_smem<T2, TK, TK, false, -1>(s, a.px, tid, nthr);
  T2* dst = a.dst[f] + (size_t)j * a.M;
  for (int idx = tid; idx < a.Kx * TK; idx += nthr) {
    int ikx = idx / TK, t = idx - ikx * TK, kz = kz0 + t;
    int pos = ikx <= a.kxmax ? ikx : ikx - a.Kx + a.nx;
    if (kz < a.Kz) {
      T2 v = s[pos * TK + t];
      if (a.scale != (T)0) { v.x /= a.scale; v.y /= a.scale; }
      dst[(size_t)ikx * a.Kz + kz] = v;
    }
  }
}

// -------------------------------------------------------------------------------
enum { ZMODE_ROTOR = 0, ZMODE_IDENTITY = 1, ZMODE_EXPORT = 2 };

template <typename T>
struct ZArgs {
  typedef typename Vec2<T>::type T2;
  T2* Q; FftPlan pz;
  int nx, nzp, Kz, pb, j_start;
  T n2;                            // (T)N2: the reference divides by the int N2 (convolution_kernels.cu:107-121)
  unsigned long long* umax_bits;   // 3 slots
  // export only
  T* phys; const int* invx; const int* invz; int nz;
};

__device__ __forceinline__ void atomic_max_pos(unsigned long long* slot, double v) {
  atomicMax(slot, (unsigned long long)__double_as_longlong(v));   // v >= 0: bit order == value order
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) zstage_kernel(ZArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  constexpr int NIN = MODE == ZMODE_ROTOR ? 6 : (MODE == ZMODE_IDENTITY ? 2 : 1);
  constexpr int NOUT = MODE == ZMODE_ROTOR ? 3 : (MODE == ZMODE_IDENTITY ? 2 : 0);
  constexpr int ST = NIN;
  T2* s = (T2*)smem_raw;
  __shared__ double red[3][8];
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int xp = blockIdx.x, jb = blockIdx.y;
  const int n = a.nzp;
  const size_t fstride = (size_t)a.pb * a.nx * a.Kz;
  T2* q0 = a.Q + ((size_t)jb * a.nx + 2 * xp) * a.Kz;     // field 0, x = 2xp ; x+1 is Kz further
  for (int idx = tid; idx < n * ST; idx += nthr) s[idx] = mk<T2>(0, 0);
  __syncthreads();
  // two-for-one packing: Z = A(x0) + i A(x1), Hermitian-extended; imaginary part of the kz=0 bin is
  // dropped, as a C2R transform does
  for (int idx = tid; idx < a.Kz * NIN; idx += nthr) {
    int f = idx / a.Kz, k = idx - f * a.Kz;
    T2 A = q0[f * fstride + k], B = q0[f * fstride + a.Kz + k];
    if (k == 0) s[f] = mk<T2>(A.x, B.x);
    else {
      s[(size_t)k * ST + f] = mk<T2>(A.x - B.y, A.y + B.x);
      s[(size_t)(n - k) * ST + f] = mk<T2>(A.x + B.y, B.x - A.y);
    }
  }
  __syncthreads();
  fft_smem<T2, NIN, ST, true, +1>(s, a.pz, tid, nthr);
  if (MODE == ZMODE_ROTOR) {
    T mu = 0, mv = 0, mw = 0;
    for (int i = tid; i < n; i += nthr) {
      T2* p = s + (size_t)i * ST;
      T2 u = p[0], v = p[1], w = p[2], ox = p[3], oy = p[4], oz = p[5];
      mu = fmax(mu, fmax(fabs(u.x), fabs(u.y))); mv = fmax(mv, fmax(fabs(v.x), fabs(v.y))); mw = fmax(mw, fmax(fabs(w.x), fabs(w.y)));
      u.x /= a.n2; v.x /= a.n2; w.x /= a.n2; u.y /= a.n2; v.y /= a.n2; w.y /= a.n2;
      ox.x /= a.n2; oy.x /= a.n2; oz.x /= a.n2; ox.y /= a.n2; oy.y /= a.n2; oz.y /= a.n2;
      p[0] = mk<T2>(v.x * oz.x - w.x * oy.x, v.y * oz.y - w.y * oy.y);
      p[1] = mk<T2>(w.x * ox.x - u.x * oz.x, w.y * ox.y - u.y * oz.y);
      p[2] = mk<T2>(u.x * oy.x - v.x * ox.x, u.y * oy.y - v.y * ox.y);
    }
    // block max -> one atomic per block and component
    double r3[3] = { (double)mu, (double)mv, (double)mw };
#pragma unroll
    for (int c = 0; c < 3; c++) {
      double v = r3[c];
      for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
      if ((tid & 31) == 0) red[c][tid >> 5] = v;
    }
    __syncthreads();
    if (tid < 3) {
      double v = 0; for (int wi = 0; wi < (nthr >> 5); wi++) v = fmax(v, red[tid][wi]);
      atomic_max_pos(a.umax_bits + tid, v);
    }
  }
  if (MODE == ZMODE_EXPORT) {
    const int j = a.j_start + jb;
    const int x0 = a.invx[2 * xp], x1 = a.invx[2 * xp + 1];
    T* o0 = a.phys + ((size_t)j * a.nx + x0) * (2 * a.nz);
    T* o1 = a.phys + ((size_t)j * a.nx + x1) * (2 * a.nz);
    for (int i = tid; i < n; i += nthr) { int z = a.invz[i]; T2 v = s[i]; o0[z] = v.x; o1[z] = v.y; }
    if (tid < 2) { o0[n + tid] = 0; o1[n + tid] = 0; }
    return;
  }
  __syncthreads();
  fft_smem<T2, NOUT, ST, false, -1>(s, a.pz, tid, nthr);
  // Hermitian unpacking of the pair and truncation to kz < Kz
  for (int idx = tid; idx < a.Kz * NOUT; idx += nthr) {
    int f = idx / a.Kz, k = idx - f * a.Kz;
    T2 Yk = s[(size_t)k * ST + f], Yn = s[(size_t)(k ? n - k : 0) * ST + f];
    const T h = (T)0.5;
    q0[f * fstride + k] = mk<T2>(h * (Yk.x + Yn.x), h * (Yk.y - Yn.y));
    q0[f * fstride + a.Kz + k] = mk<T2>(h * (Yk.y + Yn.y), h * (Yn.x - Yk.x));
  }
}

// -------------------------------------------------------------------------------
static int pick_tk(chan_handle* h, size_t elem) {
  for (int tk = 8; tk >= 2; tk >>= 1)
    if ((size_t)h->d.nx * tk * elem <= 200 * 1024) return tk;
  return 0;
}

template <typename T, int TK>
static int launch_x(chan_handle* h, bool inverse, XArgs<T>& a, int nf, int pb) {
  typedef typename Vec2<T>::type T2;
  size_t sm = (size_t)h->d.nx * TK * sizeof(T2);
  dim3 grid((h->d.Kz + TK - 1) / TK, pb, nf);
  if (inverse) {
    CHAN_CUDA(h, cudaFuncSetAttribute(xfft_inv_kernel<T, TK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    xfft_inv_kernel<T, TK><<<grid, 256, sm, h->stream>>>(a);
  } else {
    CHAN_CUDA(h, cudaFuncSetAttribute(xfft_fwd_kernel<T, TK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    xfft_fwd_kernel<T, TK><<<grid, 256, sm, h->stream>>>(a);
  }
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
template <typename T>
static int launch_x_any(chan_handle* h, bool inverse, XArgs<T>& a, int nf, int pb) {
  typedef typename Vec2<T>::type T2;
  int tk = pick_tk(h, sizeof(T2));
  if (tk == 8) return launch_x<T, 8>(h, inverse, a, nf, pb);
  if (tk == 4) return launch_x<T, 4>(h, inverse, a, nf, pb);
  if (tk == 2) return launch_x<T, 2>(h, inverse, a, nf, pb);
  return set_err(h, CHAN_ERR_ARG, "NX=%d too large for the shared-memory x transform", h->d.nx);
}

template <typename T, int MODE>
static int launch_z(chan_handle* h, ZArgs<T>& a, int nin, int pb) {
  typedef typename Vec2<T>::type T2;
  size_t sm = (size_t)h->d.nzp * nin * sizeof(T2);
  if (sm > 220 * 1024) return set_err(h, CHAN_ERR_ARG, "2NZ-2=%d too large for the shared-memory z transform", h->d.nzp);
  CHAN_CUDA(h, cudaFuncSetAttribute(zstage_kernel<T, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  dim3 grid(h->d.nx / 2, pb);
  zstage_kernel<T, MODE><<<grid, 256, sm, h->stream>>>(a);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}

template <typename T>
static void fill_x(chan_handle* h, XArgs<T>& a) {
  typedef typename Vec2<T>::type T2;
  a.Q = (T2*)h->Q; a.px = h->px; a.nx = h->d.nx; a.Kx = h->d.Kx; a.Kz = h->d.Kz; a.kxmax = h->d.kxmax;
  a.M = (long)h->d.Kx * h->d.Kz; a.scale = (T)0;
}
template <typename T>
static void fill_z(chan_handle* h, ZArgs<T>& a) {
  typedef typename Vec2<T>::type T2;
  a.Q = (T2*)h->Q; a.pz = h->pz; a.nx = h->d.nx; a.nzp = h->d.nzp; a.Kz = h->d.Kz; a.nz = h->d.nz;
  a.n2 = (T)(h->d.nx * h->d.nzp);
  a.umax_bits = h->sc->umax_bits;
  a.phys = 0; a.invx = (const int*)h->invx; a.invz = (const int*)h->invz;
}

template <typename T>
static int nonlinear_t(chan_handle* h) {
  typedef typename Vec2<T>::type T2;
  XArgs<T> xi, xf; fill_x(h, xi); fill_x(h, xf);
  const void* in[6] = { h->u, h->v, h->w, h->ox, h->oy, h->oz };
  void* out[3] = { h->ox, h->oy, h->oz };
  for (int f = 0; f < 6; f++) { xi.src[f] = (const T2*)in[f]; xi.dst[f] = 0; xf.src[f] = 0; xf.dst[f] = f < 3 ? (T2*)out[f] : 0; }
  ZArgs<T> z; fill_z(h, z);
  for (int js = 0; js < h->d.nyl; js += h->plane_batch) {
    int pb = h->d.nyl - js < h->plane_batch ? h->d.nyl - js : h->plane_batch;
    xi.j_start = xf.j_start = z.j_start = h->d.j0 + js; xi.pb = xf.pb = z.pb = pb;
    int rc = launch_x_any<T>(h, true, xi, 6, pb); if (rc) return rc;
    rc = launch_z<T, ZMODE_ROTOR>(h, z, 6, pb); if (rc) return rc;
    rc = launch_x_any<T>(h, false, xf, 3, pb); if (rc) return rc;
  }
  return CHAN_OK;
}

template <typename T>
static int symmetrize_t(chan_handle* h, void* fa, void* fb) {
  typedef typename Vec2<T>::type T2;
  XArgs<T> xi, xf; fill_x(h, xi); fill_x(h, xf);
  for (int f = 0; f < 6; f++) { xi.src[f] = 0; xi.dst[f] = 0; xf.src[f] = 0; xf.dst[f] = 0; }
  xi.src[0] = (const T2*)fa; xi.src[1] = (const T2*)fb; xf.dst[0] = (T2*)fa; xf.dst[1] = (T2*)fb;
  xf.scale = (T)(h->d.nx * h->d.nzp);
  ZArgs<T> z; fill_z(h, z);
  for (int js = 0; js < h->d.nyl; js += h->plane_batch) {
    int pb = h->d.nyl - js < h->plane_batch ? h->d.nyl - js : h->plane_batch;
    xi.j_start = xf.j_start = z.j_start = h->d.j0 + js; xi.pb = xf.pb = z.pb = pb;
    int rc = launch_x_any<T>(h, true, xi, 2, pb); if (rc) return rc;
    rc = launch_z<T, ZMODE_IDENTITY>(h, z, 2, pb); if (rc) return rc;
    rc = launch_x_any<T>(h, false, xf, 2, pb); if (rc) return rc;
  }
  return CHAN_OK;
}

template <typename T>
static int export_t(chan_handle* h, const void* f, void* phys) {
  typedef typename Vec2<T>::type T2;
  XArgs<T> xi; fill_x(h, xi);
  for (int k = 0; k < 6; k++) { xi.src[k] = 0; xi.dst[k] = 0; }
  xi.src[0] = (const T2*)f;
  ZArgs<T> z; fill_z(h, z); z.phys = (T*)phys;
  for (int js = 0; js < h->d.nyl; js += h->plane_batch) {
    int pb = h->d.nyl - js < h->plane_batch ? h->d.nyl - js : h->plane_batch;
    xi.j_start = z.j_start = h->d.j0 + js; xi.pb = z.pb = pb;
    int rc = launch_x_any<T>(h, true, xi, 1, pb); if (rc) return rc;
    rc = launch_z<T, ZMODE_EXPORT>(h, z, 1, pb); if (rc) return rc;
  }
  return CHAN_OK;
}

int launch_nonlinear_planes(chan_handle* h) { return h->prec == 4 ? nonlinear_t<float>(h) : nonlinear_t<double>(h); }
int launch_symmetrize(chan_handle* h, void* a, void* b) { return h->prec == 4 ? symmetrize_t<float>(h, a, b) : symmetrize_t<double>(h, a, b); }
int launch_to_physical(chan_handle* h, const void* f, void* phys) { return h->prec == 4 ? export_t<float>(h, f, phys) : export_t<double>(h, f, phys); }

}  // namespace chan
