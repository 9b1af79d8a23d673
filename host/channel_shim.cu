/* The reference's operator seam implemented on libchannel_b200.
 *
 * Nasrollah/CHANNEL_GPU has no plugin layer: RK3.c, nonLinear.c and RK3_kernels.cu call the free functions that
 * src/channel.h declares (channel.h:175-265).  This file DEFINES those functions — same names, same C++ signatures, same
 * in-place semantics on the caller's device arrays in the reference layout — by forwarding to the C ABI of
 * include/channel_b200.h, so that the reference's own driver links and runs unchanged on top of this library:
 *
 *     reference, kept:      RK3.c  RK3_kernels.cu  nonLinear.c  meanUevol.c  statistics.cu  fft.c  channel_cuda_mpi.c
 *                           setUp.c  check.cu
 *     reference, replaced:  derivatives_nu_double.cu  hemholzt_nu_double.cu  implicitStep_nu_double.cu
 *                           bilplacSolver_double.cu  nonLinear_kernels.cu  convolution.c  convolution_kernels.cu
 *                           dealias.cu  imposeSymetry.c            (all by this file)
 *
 * It is compiled only by `SHIM=1 oracle/build_ref.sh NX NY NZ [32|64]`, against the reference's channel.h where it lies
 * (and through the same float -> double retyping as the reference text for the 64-bit build), into
 * oracle/_ref/refshim_<grid>_f<prec>; tests/test_gpu_shim.py runs that binary against the pure reference build's dumps.
 * A maintainer of the reference who wants the drop-in adds this one file to the Makefile in place of the nine above
 * and links -lchannel_b200.
 */
#include "channel.h"
#include "channel_b200.h"
#include <stdio.h>
#include <stdlib.h>

static chan_handle* g_h = 0;

static chan_handle* shim_handle(domain_t domain) {
  if (g_h) return g_h;
  chan_config c;
  chan_config_default(&c);
  c.nx = NX; c.ny = NY; c.nz = NZ;                          /* compile-time grid of the reference (channel.h:42-44) */
  c.precision = (int)sizeof(float);                         /* 4 as shipped, 8 in the retyped build */
  c.reynolds = REYNOLDS; c.qvelocity = QVELOCITY;
  c.rank = domain.rank; c.size = domain.size;
  if (domain.size != 1) { fprintf(stderr, "channel_shim: one rank per process only (the multi-GPU path takes an NCCL communicator, INTEGRATION.md section 4)\n"); exit(1); }
  if (chan_create(&c, &g_h)) { fprintf(stderr, "channel_shim: chan_create: %s\n", g_h ? chan_last_error(g_h) : "out of memory"); exit(1); }
  return g_h;
}
/* the reference's helpers print and exit(1) (check.cu:3-79): so does the seam */
#define SHIM_CK(call) do { if (call) { fprintf(stderr, "channel_shim: %s: %s\n", #call, chan_last_error(g_h)); exit(1); } } while (0)

/* set-up hooks of the replaced translation units (setUp.c:14-16) */
void setDerivativesDouble(domain_t domain) { shim_handle(domain); }
void setHemholztDouble(domain_t domain) { shim_handle(domain); }
void setImplicitDouble(domain_t domain) { shim_handle(domain); }

void deriv_Y_HO_double(float2* u, domain_t domain) { SHIM_CK(chan_deriv_y(shim_handle(domain), u)); }                 /* channel.h:239 */
void deriv_YY_HO_double(float2* u, domain_t domain) { SHIM_CK(chan_deriv_yy(shim_handle(domain), u)); }               /* channel.h:240 */
void hemholztSolver_double(float2* u, domain_t domain) { SHIM_CK(chan_helmholtz(shim_handle(domain), u)); }           /* channel.h:212 */
void implicitSolver_double(float2* u, float betha, float dt, domain_t domain) {                                       /* channel.h:230 */
  SHIM_CK(chan_implicit(shim_handle(domain), u, (double)betha, (double)dt));
}
void bilaplaSolver_double(float2* ddv, float2* v, float2* dv, float betha, float dt, domain_t domain) {               /* channel.h:265 */
  SHIM_CK(chan_bilaplacian(shim_handle(domain), ddv, v, dv, (double)betha, (double)dt));
}
void calcUW(float2* ux, float2* uz, float2* f, float2* g, domain_t domain) { SHIM_CK(chan_calc_uw(shim_handle(domain), ux, uz, f, g)); }       /* channel.h:221 */
void calcHvg(float2* nl_x, float2* nl_y, float2* nl_z, domain_t domain) { SHIM_CK(chan_calc_hvg(shim_handle(domain), nl_x, nl_y, nl_z)); }    /* channel.h:222 */

void convolution(float2* ux, float2* uy, float2* uz, float2* wx, float2* wy, float2* wz, domain_t domain) {           /* channel.h:179 */
  chan_handle* h = shim_handle(domain);
  /* convolution.c:7-16: the wall shear of the mean mode goes to the host-side mean-flow code before the curl */
  cudaCheck(cudaMemcpy(wz, ux, SIZE, cudaMemcpyDeviceToDevice), domain, "MemInfo1");
  SHIM_CK(chan_deriv_y(h, wz));
  if (RANK == 0) readUtau(wz, domain);
  SHIM_CK(chan_convolution(h, ux, uy, uz, wx, wy, wz));
}

void dealias(float2* u, domain_t domain) { SHIM_CK(chan_dealias(shim_handle(domain), u)); }                            /* channel.h:243 */
void set2zero(float2* u, domain_t domain) { cudaCheck(cudaMemset(u, 0, SIZE), domain, "MemInfo1"); }                   /* channel.h:244 */
static __global__ void shim_normalize_kernel(float2* u, size_t n, float n2) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { u[i].x /= n2; u[i].y /= n2; }                 /* normalizekernel, dealias.cu:72-97 */
}
void normalize(float2* u, domain_t domain) {                                                                           /* channel.h:245 */
  const size_t n = (size_t)NXSIZE * NY * NZ;
  shim_normalize_kernel<<<(unsigned)((n + 255) / 256), 256>>>(u, n, (float)(NX * (2 * NZ - 2)));
  kernelCheck(cudaGetLastError(), domain, "normalize");
}
void imposeSymetry(float2* u, float2* v, domain_t domain) { SHIM_CK(chan_impose_symmetry(shim_handle(domain), u, v)); } /* channel.h:175 */
