// chan_deriv_y_pencils_static: compact d/dy of a batch of y-pencils with NO memory
// allocation of any kind inside the call.
//
// Who needs this.  Every other entry point works out of buffers that chan_create
// obtained with cudaMalloc.  Some hosts cannot do that at the moment they need a
// derivative: a solver that has handed the whole device to its own arena (a
// memory-sized Re_tau=2000 run leaves no room, and a second cudaMalloc would
// fail or fragment the arena), a post-processing pass that only borrows the GPU
// between two steps of someone else's job, or a time loop that must not take
// the implicit device-wide synchronisation cudaMalloc/cudaFree bring.  For them
// the Thomas forward-sweep scratch of the d/dy kernel lives in the fixed array
// below, which is part of the library image: each GPU that loads the library
// reserves it once, until the process exits.
//
// Limits and rules
//  * at most CHAN_STATIC_SCRATCH_DOUBLES doubles of scratch: NY * 2 * n_pencils
//    must not exceed it (2 Mi doubles = 16 MiB, e.g. 5433 pencils at NY = 193);
//    larger requests are refused with CHAN_ERR_ARG and a message, nothing runs;
//  * the first call on each device looks up the array's device address (see
//    static_scratch_for_device below).  Make that first call outside any stream
//    capture;
//  * every call on a device uses the same array.  Nothing in it has to start at
//    zero (each scratch entry is written by the forward sweep before the
//    backward sweep reads it), so no clearing pass is issued; but two calls whose
//    kernels overlap on the GPU (different streams, different handles, different
//    host threads) race on it.  Serialise them.
//
//  * u_dev must be complete before the call: the kernel runs on the handle's own
//    non-blocking stream, which is not ordered after the caller's streams.  The
//    call returns after the kernel has finished (it synchronises that stream).
//  * one thread walks one pencil, so its loads are 2*NY elements apart between
//    neighbouring threads (uncoalesced): fine for the small batches this entry
//    point is limited to, not a substitute for the fused path.
//
// The arithmetic is that of chan_deriv_y (same y tables, same sweep order and
// operations per row): fp64 results are identical to it on the same pencils
// (tests/test_zz_no_alloc.py asserts equality).
#include "chan_internal.h"

namespace chan {

#define CHAN_STATIC_SCRATCH_DOUBLES (2u << 20)

static __device__ __align__(256) double g_static_scratch[CHAN_STATIC_SCRATCH_DOUBLES];

// In host code `g_static_scratch` is only a stand-in symbol, not a device pointer: handing it to a
// kernel would fault (or, where the GPU can reach host memory, quietly use the wrong bytes).  The
// real address is asked for once per device, with that device current, and cached.
static const int kMaxDevices = 64;
static double* g_scratch_addr[kMaxDevices];

// The address is looked up for the HANDLE's device (the one its stream and tables live on), whatever
// device the calling thread has current: that device is made current for the lookup and for the
// launch that follows, and the caller's device is restored before returning.
static int static_scratch_for_device(chan_handle* h, double** out) {
  const int dev = h->device;
  if (dev < 0 || dev >= kMaxDevices) return set_err(h, CHAN_ERR_ARG, "device ordinal %d outside the supported range", dev);
  if (!g_scratch_addr[dev]) {
    double* p = 0;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess || !p)
      return set_err(h, CHAN_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", dev, cudaGetErrorString(e));
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return CHAN_OK;
}

// One thread per real chain (pencil p, component c).  u: complex [n_pencils][NY], y fastest (a run of
// consecutive (i,k) pencils of the reference layout), in place.  s: scratch [NY][2*n_pencils].
template <typename T>
__global__ void __launch_bounds__(128) deriv_y_pencils_kernel(T* u, double* s, D1Tab t, int ny, long n_chain) {
  long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_chain) return;
  T* f = u + (e >> 1) * (long)ny * 2 + (e & 1);         // element j at f[2*j]
  // forward sweep with the folded LU (kernels_tables.cu): rows lag two behind the loads
  double f0 = (double)f[0], f1 = (double)f[2], f2 = (double)f[4];
  double d = __ldg(t.eC) * f0 + __ldg(t.eN) * f1 + __ldg(t.eP) * f2;
  s[e] = d;
  for (int j = 1; j < ny - 1; j++) {
    double fp = (double)f[2 * (j - 1)], fc = (double)f[2 * j], fn = (double)f[2 * (j + 1)];
    d = __ldg(t.eP + j) * fp + __ldg(t.eC + j) * fc + __ldg(t.eN + j) * fn - __ldg(t.eL + j) * d;
    s[(long)j * n_chain + e] = d;
  }
  {
    int j = ny - 1;
    double fc = (double)f[2 * j], fp = (double)f[2 * (j - 1)], fn = (double)f[2 * (j - 2)];
    d = __ldg(t.eC + j) * fc + __ldg(t.eP + j) * fp + __ldg(t.eN + j) * fn - __ldg(t.eL + j) * d;
    s[(long)j * n_chain + e] = d;
  }
  // backward sweep, result cast to the state type like cast_kernel (derivatives_nu_double.cu:4-30)
  double x = 0.0;
  for (int j = ny - 1; j >= 0; j--) {
    x = s[(long)j * n_chain + e] - __ldg(t.cp + j) * x;
    f[2 * j] = (T)x;
  }
}

}  // namespace chan

using namespace chan;

extern "C" int chan_deriv_y_pencils_static(chan_handle* h, void* u_dev, long n_pencils) {
  if (!h || !u_dev || n_pencils <= 0) return CHAN_ERR_ARG;
  const int ny = h->d.ny;
  const long n_chain = 2 * n_pencils;
  if ((double)n_chain * (double)ny > (double)CHAN_STATIC_SCRATCH_DOUBLES)
    return set_err(h, CHAN_ERR_ARG,
                   "chan_deriv_y_pencils_static: %ld pencils of NY=%d need %.0f scratch doubles, the static array holds %u "
                   "(at most %ld pencils per call at this NY)",
                   n_pencils, ny, (double)n_chain * ny, (unsigned)CHAN_STATIC_SCRATCH_DOUBLES,
                   (long)(CHAN_STATIC_SCRATCH_DOUBLES / (2u * (unsigned)ny)));
  int caller_dev = -1;
  CHAN_CUDA(h, cudaGetDevice(&caller_dev));
  if (caller_dev != h->device) CHAN_CUDA(h, cudaSetDevice(h->device));
  double* scratch = 0;
  int rc = static_scratch_for_device(h, &scratch);
  if (!rc) {
    unsigned grid = (unsigned)((n_chain + 127) / 128);
    if (h->prec == 4) deriv_y_pencils_kernel<float><<<grid, 128, 0, h->stream>>>((float*)u_dev, scratch, h->yt.d1[0], ny, n_chain);
    else deriv_y_pencils_kernel<double><<<grid, 128, 0, h->stream>>>((double*)u_dev, scratch, h->yt.d1[0], ny, n_chain);
    h->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) rc = set_err(h, CHAN_ERR_CUDA, "chan_deriv_y_pencils_static: %s", cudaGetErrorString(e));
  }
  if (caller_dev != h->device) cudaSetDevice(caller_dev);
  return rc;
}
