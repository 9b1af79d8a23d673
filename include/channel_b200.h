/* libchannel_b200 — C ABI of the B200-native KMM channel-flow RK3 time-stepper.
 *
 * Drop-in boundary for the hot path of Nasrollah/CHANNEL_GPU (one RK3 step of
 * the (phi = lap v, g = omega_y) formulation).  The reference has no FFI layer:
 * its seam is the set of free functions of src/channel.h that RK3.c,
 * nonLinear.c, convolution.c and imposeSymetry.c call.  Each entry point below
 * names the reference interface it replaces (file:line under
 * /root/reference/src).  INTEGRATION.md shows the binding a maintainer adds.
 *
 * Conventions
 *  - plain C, opaque handle, int return codes (0 = CHAN_OK); the reference's
 *    print-and-exit(1) error helpers (check.cu:3-79) become codes plus
 *    chan_last_error().
 *  - "reference layout": complex array [NX][NZ][NY], index h = i*NY*NZ + k*NY + j
 *    (RK3_kernels.cu:16-18), element type float2 (precision 4, the reference as
 *    shipped) or double2 (precision 8), spectral coefficients un-normalised
 *    (x N2 = NX*(2NZ-2), convolution_kernels.cu:93).  NZ counts COMPLEX z modes;
 *    the physical z size is 2NZ-2 (fft.c:17).
 *  - *_dev arguments are DEVICE pointers in the reference layout, updated in
 *    place exactly like the reference functions they replace; inside, the
 *    library keeps only the active modes of the 2/3 rule (dealias.cu:37) in its
 *    own plane-major layout, so truncated modes come back as exact zeros.
 *  - grid sizes are run-time values (the reference bakes -DNX -DNY -DNZ in at
 *    compile time, Makefile:8-11, channel.h:42-44); NY is not limited to 1024
 *    (main.c:79) and NX, NY need no divisibility (channel.h:72-73).
 *  - one handle drives one GPU; no globals; calls on one handle are not
 *    re-entrant (neither is the reference: static handles per translation unit).
 */
#ifndef CHANNEL_B200_H
#define CHANNEL_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct chan_handle chan_handle;

enum {
  CHAN_OK = 0,
  CHAN_ERR_ARG = 1,      /* bad argument / unsupported size                       */
  CHAN_ERR_CUDA = 2,     /* a CUDA runtime call or kernel failed                  */
  CHAN_ERR_ALLOC = 3,    /* out of device or host memory                          */
  CHAN_ERR_STATE = 4     /* call order (e.g. step before fields were set)         */
};

/* Replaces domain_t (channel.h:23-30) and the compile-time constants of
 * channel.h:50-62, RK3.c:68.  chan_config_default() fills the reference's values. */
typedef struct chan_config {
  int nx, ny, nz;        /* run.conf application.NX/NY/NZ (config.c:18-23)        */
  int precision;         /* 4 = float state (as shipped), 8 = double state        */
  double reynolds;       /* REYNOLDS   channel.h:61  (3250)                       */
  double qvelocity;      /* QVELOCITY  channel.h:62  (1.8)                        */
  double cfl;            /* CFL        RK3.c:68      (0.5)                        */
  int rank, size;        /* domain_t.rank/size: slab of a multi-GPU run           */
  int device;            /* CUDA device ordinal, -1 = current (main.c:87)         */
  int plane_batch;       /* y-planes per FFT-stage launch group, 0 = auto         */
  void* nccl_comm;       /* ncclComm_t of the `size` ranks (one per GPU of the box);
                            NULL when size == 1.  Replaces MPI_COMM_WORLD (main.c:25-29) */
} chan_config;

typedef struct chan_scalars {
  double dt, dt_next, dt_c, dt_v;     /* RK3.c:64-109, 114, 179                   */
  double umax[3];                     /* max |u|,|v|,|w| (un-normalised) fft.c:143 */
  double utau[2];                     /* meanUevol.c:178-199                      */
  int counter;                        /* steps taken (RK3.c:180)                  */
  double time;                        /* sum of the dt of the steps taken (RK3.c:126 ignores its `time`) */
} chan_scalars;

/* field selectors for chan_get_field / chan_set_field (globals of RK3.c:4-12) */
enum { CHAN_F_DDV = 0, CHAN_F_G, CHAN_F_V, CHAN_F_DV, CHAN_F_R_DDV, CHAN_F_R_G,
       CHAN_F_U, CHAN_F_W, CHAN_F_WX, CHAN_F_WY, CHAN_F_WZ, CHAN_F_COUNT };

int  chan_config_default(chan_config* cfg);
/* setUp (setUp.c:3-29) + setRK3 (RK3.c:17-50) + setRKmean (meanUevol.c:43-84).
 * Sizes: NX and 2NZ-2 must factor into 2s, 3s, 5s and 7s (the hand-written FFTs have radices 2, 3, 4, 5, 7 and their
 * composites 6, 8, 10, 12, 14, 16; the reference, through cuFFT, takes any size), NX even, NX >= 8, NY >= 8, NZ >= 5; NX*8 and
 * (2NZ-2)*7 complex elements of the state type must fit the 200 KB / 220 KB shared-memory tiles of the x / z
 * transforms (NX <= 12800, 2NZ-2 <= 2011 in fp64).
 * Ownership on failure: whenever *out comes back non-NULL the caller owns it, also when the return code is
 * not CHAN_OK - read chan_last_error(*out), then chan_destroy(*out).  *out is NULL only when the handle
 * itself could not be allocated (CHAN_ERR_ALLOC) or an argument was NULL. */
int  chan_create(const chan_config* cfg, chan_handle** out);
void chan_destroy(chan_handle* h);
const char* chan_last_error(const chan_handle* h);
long chan_launch_count(const chan_handle* h);   /* kernels launched so far */
/* Per-stage device timing with CUDA events on the library's own stream (no
 * counterpart in the reference, which has no timers at all).  Stages:
 * 0 pencil_pre, 1 plane stage (x/z FFTs + u x omega), 2 scalar stage, 3 pencil_post,
 * 4 one whole RK3 step as issued by chan_rk_step (encloses the others);
 * 5 xfft_inv, 6 zstage, 7 xfft_fwd: the three kernels inside stage 1;
 * 8, 9 (several GPUs): the bulk peer copies pencil -> plane (inside stage 0) and plane -> pencil (inside stage 1).
 * chan_get_stage_ms synchronises, returns accumulated milliseconds and launch
 * groups per stage (arrays of CHAN_STAGE_COUNT) since the last call, and resets them.  At most 32768 stage
 * instances are recorded between two reads (a Re_tau=1000 step is 76 of them); when more were issued the call still returns
 * the recorded ones but with CHAN_ERR_STATE, so that a caller never mistakes a partial sum for the whole. */
enum { CHAN_STAGE_PRE = 0, CHAN_STAGE_PLANES, CHAN_STAGE_SCALAR, CHAN_STAGE_POST, CHAN_STAGE_STEP,
       CHAN_STAGE_XINV, CHAN_STAGE_Z, CHAN_STAGE_XFWD, CHAN_STAGE_XCHG_OUT, CHAN_STAGE_XCHG_BACK, CHAN_STAGE_COUNT };
int  chan_set_timing(chan_handle* h, int on);
int  chan_get_stage_ms(chan_handle* h, double* ms_out, long* count_out);

/* ---- state in/out, HOST pointers, reference layout ------------------------- */
/* readData / genRandData + cudaMemcpy H2D (io.c:3-34,74-98; main.c:107-114) */
int chan_set_fields(chan_handle* h, const void* ddv_host, const void* g_host);
/* cudaMemcpy D2H + writeData (io.c:37-72) */
int chan_get_fields(chan_handle* h, void* ddv_host, void* g_host);
/* One field.  With several ranks (chan_config.size > 1) the array is still the FULL [NX][NZ][NY] array of the run,
 * as in the reference where every rank holds rows i of its slab (main.c:47): a rank reads / writes only the rows
 * of its own active-kx slab (rank 0 also zeroes the rows of the truncated kx on output); other rows are not touched. */
int chan_get_field(chan_handle* h, int which, void* host);
int chan_set_field(chan_handle* h, int which, const void* host);
/* The same for a host array that holds just this rank's slab: [Kxl][NZ][NY], rows in active-kx order
 * ikx0 .. ikx0+Kxl-1 with (ikx0, Kxl) = chan_partition(chan_active_kx(NX), size, rank) and kx = ikx <= NX/3 ? ikx :
 * ikx - chan_active_kx(NX).  What a memory-sized multi-GPU run uses: no rank ever holds the dense array
 * (hit_mpi.c:282-328 fans the file out slab by slab for the same reason). */
int chan_get_field_slab(chan_handle* h, int which, void* host_slab);
int chan_set_field_slab(chan_handle* h, int which, const void* host_slab);
/* DEVICE pointers to full reference-layout arrays (the ddv, g of main.c:102-103): the coarse RKstep seam without a
 * bounce through the host.  chan_get_field(s)_dev clears the whole array first and is single-GPU only. */
int chan_set_fields_dev(chan_handle* h, const void* ddv_dev, const void* g_dev);
int chan_get_fields_dev(chan_handle* h, void* ddv_dev, void* g_dev);
int chan_set_field_dev(chan_handle* h, int which, const void* dev);
int chan_get_field_dev(chan_handle* h, int which, void* dev);
/* readU / writeU payload (meanUevol.c:153-176): NY complex, .x = <u>(y)*N2 */
int chan_set_umean(chan_handle* h, const void* umean_host);
int chan_get_umean(chan_handle* h, void* umean_host);
/* N_host of meanUevol.c:127-133: NY reals (double) */
int chan_get_nmean(chan_handle* h, double* out_ny);
int chan_get_scalars(chan_handle* h, chan_scalars* out);
/* >0: use this dt for every following step instead of calcDt (test/bench hook) */
int chan_set_fixed_dt(chan_handle* h, double dt);
/* dt of the step about to run (the reference starts with 0, RK3.c:114) */
int chan_set_dt(chan_handle* h, double dt);

/* ---- checkpoint / restart -------------------------------------------------------
 * No counterpart in the reference: it restarts from the field files written once at the end of a run (main.c:139,
 * io.c:37-72) and loses R_phi, R_g, dt and the step count (RK3.c:114), so a restarted run is not the continued
 * run.  Here one rank's complete state (phi, g, v, dv, R_phi, R_g, mean profile, N_old, every step scalar, the
 * symmetry cadence counter) goes into one host buffer of chan_checkpoint_bytes() bytes; loading it into a handle
 * created with the same grid / precision / rank / size continues bit for bit (tests/test_gpu_checkpoint.py).
 * The payload is versioned and private to the library (compact layout); host/checkpoint.c writes it to a file. */
size_t chan_checkpoint_bytes(const chan_handle* h);
int chan_checkpoint_save(chan_handle* h, void* host_buf);
int chan_checkpoint_load(chan_handle* h, const void* host_buf);

/* ---- the hot path ---------------------------------------------------------- */
/* RKstep prologue (RK3.c:120-123): dealias, imposeSymetry, calcVdV */
int chan_rk_init(chan_handle* h);
/* body of RKstep's while loop (RK3.c:127-191), nsteps times, asynchronous */
int chan_rk_step(chan_handle* h, int nsteps);
/* one pass of the n_step loop (RK3.c:129-167), n = 0,1,2 */
int chan_rk_substep(chan_handle* h, int n);
/* RK3.c:174-179 (symmetry cadence, dt lag, counter) */
int chan_rk_end_step(chan_handle* h);
int chan_sync(chan_handle* h);
/* calcSt + sumElementsComplex (statistics.cu:161-243, fft.c:239-269): the four
 * per-plane sums (Reynolds stress, u, v, w) before the sqrt of the printout */
int chan_statistics(chan_handle* h, double* out_4ny_host);

/* ---- operator level, DEVICE pointers, reference layout, in place ------------ */
int chan_dealias(chan_handle* h, void* u_dev);                                  /* dealias            channel.h:243 */
int chan_deriv_y(chan_handle* h, void* u_dev);                                  /* deriv_Y_HO_double  channel.h:239 */
int chan_deriv_yy(chan_handle* h, void* u_dev);                                 /* deriv_YY_HO_double channel.h:240 */
int chan_helmholtz(chan_handle* h, void* u_dev);                                /* hemholztSolver_double channel.h:212 */
int chan_implicit(chan_handle* h, void* u_dev, double betha, double dt);        /* implicitSolver_double channel.h:230 */
int chan_bilaplacian(chan_handle* h, void* ddv_dev, void* v_dev, void* dv_dev,
                     double betha, double dt);                                  /* bilaplaSolver_double  channel.h:265 */
int chan_calc_uw(chan_handle* h, void* ux_dev, void* uz_dev,
                 const void* f_dev, const void* g_dev);                         /* calcUW             channel.h:221 */
int chan_convolution(chan_handle* h, void* ux_dev, void* uy_dev, void* uz_dev,
                     void* wx_dev, void* wy_dev, void* wz_dev);                 /* convolution        channel.h:179 */
int chan_calc_hvg(chan_handle* h, void* nlx_dev, void* nly_dev, void* nlz_dev); /* calcHvg            channel.h:222 */
int chan_impose_symmetry(chan_handle* h, void* u_dev, void* v_dev);             /* imposeSymetry      channel.h:175 */
/* fftBackwardTranspose (channel.h:116): spectral [NX][NZ][NY] -> real planes
 * [NY][NX][2NZ] (2NZ-2 points + 2 pad floats per row, fft.c:17-21), out of place */
int chan_fft_backward_transpose(chan_handle* h, const void* u_dev, void* phys_dev);

/* ---- allocation-free variant ------------------------------------------------------
 * No counterpart in the reference.  Compact d/dy (the arithmetic of chan_deriv_y) of
 * n_pencils consecutive y-pencils, complex [n_pencils][NY] with y fastest — any run of
 * (i,k) pencils of a reference-layout array — in place, on the handle's stream, for
 * hosts that cannot allocate device memory when they call (all of it already belongs
 * to their own arena, or they may not take cudaMalloc's device-wide synchronisation).
 * The forward-sweep scratch is a fixed 16 MiB array inside the library image, held by
 * every GPU that loads the library until exit: NY*2*n_pencils must not exceed
 * 2097152 doubles, otherwise CHAN_ERR_ARG is returned and nothing runs.  The call makes
 * no cudaMalloc/cudaFree and no allocating library call.  The first call on a device
 * resolves the array's address: issue it before any stream capture.  All calls share
 * the array, so calls that may overlap on the GPU must be serialised by the caller.
 * The kernel runs on the handle's device and private stream whatever device the
 * calling thread has current (it is switched and restored); u_dev must be complete
 * before the call (that stream is not ordered after the caller's), and the call
 * returns once the kernel has finished.  One thread walks one pencil, so the loads
 * are uncoalesced: meant for small batches, not as a substitute for the fused path. */
int chan_deriv_y_pencils_static(chan_handle* h, void* u_dev, long n_pencils);

/* ---- multi-GPU peer-store exchange (no counterpart in the reference, which stages every transpose
 * through host memory and MPI_Alltoall, channel_cuda_mpi.c:75-85) ------------------------------
 * One process per GPU: chan_ipc_export writes CHAN_IPC_BYTES bytes describing this rank's exchange
 * buffers; the host passes them to every other rank (any transport: the python launcher uses
 * torch.distributed), which calls chan_ipc_import(h, rank_of_owner, bytes).  Once every peer has been
 * imported the two exchanges of a substep are bulk peer copies over NVLink straight into the owners'
 * buffers (one contiguous slab per field and destination, destinations rotated by rank), and NCCL only
 * provides two stream-ordered barriers per substep. */
#define CHAN_IPC_BYTES (9 * 64)
int chan_ipc_export(chan_handle* h, void* bytes_out);
int chan_ipc_import(chan_handle* h, int peer_rank, const void* bytes_in);

/* ---- slab decomposition (main.c:47, channel.h:71-79), pure host arithmetic --- */
/* splits n items over `size` ranks as evenly as possible: [*start, *start+*count) */
int chan_partition(int n, int size, int rank, int* start, int* count);
int chan_active_kx(int nx);     /* 2*floor(nx/3)+1 */
/* What rank `rank` of `size` exchanges with `peer` for ONE field in the pencil -> plane direction
 * (the transposeYZX2XYZ of channel_cuda_mpi.c:95-128), in complex elements:
 *   out[0], out[1]  offset / count inside the rank's pencil-side field  F[NY][Kx_local][Kz]  (sent)
 *   out[2], out[3]  offset / count inside its plane-side buffer (per-source blocks)        (received)
 *   out[4], out[5]  first active-kx index and count of the peer's slab
 *   out[6], out[7]  first y plane and count of the peer's slab
 * The plane -> pencil direction swaps the roles of [0,1] and [2,3].  Pure host arithmetic. */
int chan_exchange_layout(int nx, int ny, int nz, int size, int rank, int peer, long* out8);
int chan_active_kz(int nz);     /* floor((2nz-2)/3)+1 */
/* The same exchange as ONE-SIDED copies (what the bulk NVLink peer copies do): rank `rank` writes into the buffers of
 * `dest`.  In complex elements:
 *   out[0], out[1], out[2]  pencil -> plane: source offset in the rank's pencil-side field, destination offset in dest's
 *                           plane-side buffer (the block of source `rank`), count
 *   out[3], out[4], out[5]  plane -> pencil: source offset in the rank's plane-side buffer (the block of dest's kx slab),
 *                           destination offset in dest's pencil-side field (the rows of the rank's planes), count
 * Rows are contiguous on both sides, so a sub-range of planes is the same triple shifted by rows * (modes of the slab). */
int chan_peer_put_layout(int nx, int ny, int nz, int size, int rank, int dest, long* out6);

#ifdef __cplusplus
}
#endif
#endif
