// Internal declarations of libchannel_b200 (not installed; the public C ABI is
// include/channel_b200.h).
//
// DATA LAYOUT.  Every spectral field is stored COMPACT and
// PLANE-MAJOR:  F[j][m],  m = ikx*Kz + kz,  complex T2, with
//   ikx in [0,Kx)  the 2*floor(NX/3)+1 active kx of the 2/3 rule (dealias.cu:37):
//                  kx = ikx <= kxmax ? ikx : ikx - Kx,
//   kz  in [0,Kz)  Kz = floor((2NZ-2)/3)+1,
//   j   in [0,NY)  the wall-normal index (slowest).
// One y-plane of all active modes is contiguous, so the x/z FFT kernels read
// planes directly (no y-pencil <-> xz-plane transpose on one GPU), and the
// y-solves run one thread per REAL chain (mode, re|im) with the 32 lanes of a
// warp on 32 consecutive reals of a plane row: fully coalesced.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/channel_b200.h"

namespace chan {

template <typename T> struct Vec2;
template <> struct Vec2<float>  { typedef float2  type; };
template <> struct Vec2<double> { typedef double2 type; };

// ---- y tables -------------------------------------------------------------
// Direction-generic coefficient tables, indexed by SWEEP position jj.  A sweep
// with dir=0 walks j = jj (up), dir=1 walks j = NY-1-jj (down).
struct D1Tab {            // compact first derivative, fixed matrix, LU folded in
  const double *eP, *eC, *eN;   // rhs weights (prev, centre, next in sweep) * pivot
  const double *eL;             // lower * pivot   (multiplies d'[jj-1])
  const double *cp;             // upper * pivot   (back substitution)
};
struct D2Tab {            // second-derivative compact weights (Helmholtz / implicit)
  const double *rP, *rN;        // rhs weights of prev / next in sweep (alpha2 / beta2)
  const double *wP, *wC, *wN;   // operator weights of prev / centre / next (B2 | A2, C2)
};
struct YTabs {
  D1Tab d1[2];
  D2Tab d2[2];
  const double* y;              // Fmesh(j*DELTA_Y-1.0), physical j
};

// ---- FFT plan -------------------------------------------------------------
struct FftPlan {
  int n;
  int nstage;
  int radix[16];
  int span[16];                 // sub-transform size entering stage t (1, r0, r0*r1, ...)
  unsigned span_magic[16];      // FastDiv magic of span[t] (fft_core.cuh)
  int r_first, r_last;          // radix[0], radix[nstage-1]
  unsigned rfirst_magic;        // FastDiv magic of r_first (row skew of the z transforms; 0 in PlanT.fsk.d turns it off)
  int tws[16];                  // twiddle stride of stage t: n / (span[t] * radix[t])
  const void* tw;               // T2[n]: exp(-2 pi i k / n)
  const int* inv;               // digit-reversed position -> natural index
  const int* pos;               // natural index -> digit-reversed position
};

// Plane-side addressing of one compact field whose rows come from several kx slabs (multi-GPU: one
// segment per source rank, as received; single GPU: one segment = the field itself).  Element
// (row jl, ikx, kz) of segment s = start[s] <= ikx < start[s+1] lives at
//   base[s] + (jl*width[s] + ikx - start[s])*Kz + kz      (in complex elements)
// The segments are the even partition of chan_partition (the first pr slabs have pq+1 rows, the others pq) and the bases are
// linear in (start, width): base[s] = bms*start[s] + bmw*width[s] (plane-side blocks in rank order: bms = nyl*Kz, bmw = 0;
// peers' pencil-side fields: bms = 0, bmw = j0*Kz).  With closed = 1 the FFT load / store stages compute segment, start and
// width of an active kx from that (two multiply-high divisions) instead of a select chain over up to 8 segments.
struct SegMap { int nseg; int start[9]; int width[8]; long base[8]; int closed, pq, pr; unsigned mq, mq1; long bms, bmw; };

// Peer-store exchange (multi-GPU): where the wall-normal `pre` kernel delivers row j of plane-stage
// input field f — straight into the receive buffer of the rank that owns plane j, over NVLink.
struct PeerOut { void* rcv[8][6]; int part_j[9]; int nranks; long src_off[8]; };   // src_off[p] = nyl_p*ikx0*Kz (complex elements)

// ---- handle ---------------------------------------------------------------
struct Dims {
  int nx, ny, nz, nzp;          // NZ = complex z count, nzp = 2NZ-2 physical z points
  int kxmax, Kx, Kz;            // active-mode box
  int ikx0, Kxl;                // this rank's active-kx slab (pencil stages)
  int j0, nyl;                  // this rank's y-plane slab (plane stages)
  long M;                       // Kxl*Kz modes per plane row held by the pencil stages
};

struct Scalars {                // device-resident step scalars (doubles hold T-rounded values)
  double dt;                    // dt of the current step (one-step lag, RK3.c:114,179)
  double dt_next;               // calcDt of substep 0 (RK3.c:143-145)
  double dt_c, dt_v;
  double fixed_dt;              // >0: injected instead of calcDt
  unsigned long long umax_bits[3];   // max |u|,|v|,|w| over physical points, as ordered bits
  double umax[3];               // the same maxima as sampled at substep 0 (what calcDt used)
  double dudy_wall[2];          // d<u>/dy at j=0, NY-1 (convolution.c:14-16)
  double utau[2];
  int    counter;
  double time;                  // sum of the dt of the steps taken (the reference ignores its `time` argument, RK3.c:126)
  unsigned long long umax_scratch[3];   // sink of the maxima of operator-level launches (chan_convolution): they must not feed calcDt
};

}  // namespace chan

struct chan_handle {
  int prec;                     // 4 | 8 : sizeof(state real)
  chan::Dims d;
  chan_config cfg;
  int device;
  cudaStream_t stream;
  cudaStream_t aux;             // side stream: the one-block mean-flow/dt kernel runs beside pencil_post
  cudaStream_t xs;              // side stream of the plane -> pencil peer copies: batch b travels while batch b+1 is transformed
  cudaStream_t xo;              // side stream of the pencil -> plane peer copies: the rows of plane batch b+1 travel while batch b is transformed
  cudaEvent_t ev_batch[64], ev_in[64], ev_xs, ev_pre;
  cudaStream_t xc[4]; cudaEvent_t ev_cf, ev_cj[4];      // copy-engine variant of the peer copies (CHAN_TUNE_XCHG_CE)
  int pb_rank[8], nbatch;       // plane batch of every rank (same formula everywhere) and the largest batch count of any rank
  cudaEvent_t ev_fork, ev_join, ev_sub, ev_piv;
  // state + work fields, compact plane-major, complex T
  void *phi, *g, *v, *dv, *Rphi, *Rg;
  void *u, *w, *ox, *oy, *oz;
  // y-solve scratch: 3 real chains [NY][2M] + pivots [NY][M], double
  double *sA, *sB, *sC, *sP;
  // plane-stage staging Q[f][jb][p][kz], complex T
  void* Q; int plane_batch;
  // tables
  double* ytab_store; chan::YTabs yt;
  // segment-parallel y stage (kernels_pencil_seg.cu): segment length, prefix-product tables of the
  // folded d/dy LU for that length, per-mode pivots of the Helmholtz / implicit matrices [NY][M]
  int seg_S; double *seg_pf, *seg_pb; double *pivH, *pivI; int pivH_ready;
  // on-chip y stage (kernels_pencil_lanes.cu): rows per lane, warps per mode, padded rows, staging column pitch, modes per
  // block, padded y tables, pivot tables [M][NYL] of the Helmholtz / implicit matrices
  struct { int S, W, NYL, NYP, MB; double* tabs; double *pivH, *pivI; int pivH_ready; } ln;
  void *tw_x, *tw_z; chan::FftPlan px, pz;
  int *invx, *invz, *posx, *posz;   // digit-reversal tables of the two plans
  void *mean_a, *mean_b;        // host-mesh differences in T (meanUevol.c:230-231)
  void *mean_u, *mean_N, *mean_uw, *mean_tmp;   // T[NY] (+ scratch)
  chan::Scalars* sc;            // device
  // multi-GPU exchange (comm.cu): NCCL communicator owned by the caller, receive buffers of the 6
  // fields entering the plane stage, send buffers of the 3 leaving it, slab boundaries of all ranks
  void* nccl_comm; void* rcv[6]; void* snd[3];
  // peer-copy exchange: CUDA-IPC-mapped buffers of the other ranks (rcv[6] and the pencil fields H1..3)
  int peer_ready; void* peer_rcv[8][6]; void* peer_H[8][3]; int* barrier_word;
  int part_kx[9], part_j[9];      // part_kx[r] = first active-kx index of rank r (part_kx[size] = Kx), same for planes
  chan::SegMap seg;               // plane-side segmentation (identical for every field)
  // boundary staging (reference layout <-> compact)
  void* io_buf; size_t io_bytes;   // the rank's slab [Kxl][NZ][NY] of the reference layout
  double* stat_buf;             // 4*NY sums of chan_statistics
  struct { int z_threads, z_xp, z_pad, x_threads, x_tk, max_radix, z_stop, x_skip, pencil_seg, seg_shape, post_pf, post_threads, post_stop, post_deep, z_fwdpack, z_inpack, pre_direct, fwd_direct, xchg_pipe, xchg_ce, lanes, lanes_s, lanes_mb; } tune;   // launch shapes (CHAN_TUNE_* env overrides)
  long launches;                // kernels launched so far (bench's gpu_launches)
  long host_counter;            // steps issued (symmetry cadence, RK3.c:174)
  // optional per-stage CUDA-event timing on h->stream (bench.py's roofline figures)
  enum { EV_MAX = 32768 };
  int timing; int n_ev, n_ev_dropped; cudaEvent_t ev0[EV_MAX], ev1[EV_MAX]; int ev_stage[EV_MAX];
  char err[256];
};

// Timing-ablation early returns (scripts/ablate_post.sh, CHAN_TUNE_*_STOP / CHAN_TUNE_X_SKIP) exist only in builds made
// with -DCHAN_ABLATE; the shipped library has none, and chan_create refuses those variables so that a stray setting
// cannot silently turn a run into a no-op.
#ifdef CHAN_ABLATE
#define CHAN_ABLATE_RETURN(cond) do { if (cond) return; } while (0)
#else
#define CHAN_ABLATE_RETURN(cond) do { } while (0)
#endif

namespace chan {
enum { OP_DERIV_Y = 1, OP_DERIV_YY, OP_HELM, OP_IMPLICIT, OP_BILAP, OP_VDV, OP_UW, OP_HVG, OP_OMEGA };
int set_err(chan_handle* h, int code, const char* fmt, ...);
int timing_begin(chan_handle* h, int stage);     // api.cu: CUDA-event pair around a stage (no-op unless chan_set_timing)
void timing_end(chan_handle* h, int id);
#define CHAN_CUDA(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return chan::set_err(h, CHAN_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

// kernels_tables.cu
int build_tables(chan_handle* h);
void make_plan(FftPlan* p, int n, bool small_last, int max_radix = 16);
int perm_pos_of(const FftPlan& p, int n);
// kernels_pencil.cu
int launch_pencil_pre(chan_handle* h);
int launch_pencil_post(chan_handle* h, int n);
int launch_peer_scatter_back(chan_handle* h, int js, int pb);     // the three products -> the pencil owners
int launch_peer_scatter(chan_handle* h, int batch);   // batch b of every destination; -1: everything          // the six plane-stage inputs -> the plane owners (bulk peer copies)
int launch_op(chan_handle* h, int op, void* f0, void* f1, void* f2, double betha, double dt);
// kernels_pencil_seg.cu
int seg_supported(chan_handle* h);
int seg_setup(chan_handle* h);
int launch_pencil_pre_seg(chan_handle* h);
int launch_pencil_post_seg(chan_handle* h, int n);
// kernels_pencil_lanes.cu
int lanes_supported(chan_handle* h);
int lanes_setup(chan_handle* h);
int launch_pencil_post_lanes(chan_handle* h, int n);
int launch_pencil_pre_lanes(chan_handle* h);
int launch_lanes_pivots(chan_handle* h, int n);     // pivot tables of the implicit matrix of substep n (and once the Helmholtz ones), on h->stream
// kernels_plane.cu
int launch_nonlinear_planes(chan_handle* h, bool op_level = false);   // u,v,w,omega -> H (in ox,oy,oz); op_level: inputs sit in the local pencil fields, step scalars untouched
int launch_symmetrize(chan_handle* h, void* a, void* b);     // imposeSymetry on two fields
int launch_to_physical(chan_handle* h, const void* f, void* phys_out);   // debug/parity: one field -> real [NY][NX][nzp]
// kernels_misc.cu
// slab = 0: the dense side is a full reference-layout array; 1: the rank's slab [Kxl][NZ][NY] in active-kx order
int launch_pack(chan_handle* h, const void* ref_layout_dev, void* compact, int slab = 0);      // [i][k][j] -> [j][m]
int launch_unpack(chan_handle* h, const void* compact, void* ref_layout_dev, int slab = 0);    // [j][m] -> [i][k][j] (zeros on truncated modes)
int launch_scalar_stage(chan_handle* h, int n);              // mean flow, dt, bookkeeping
int launch_end_step(chan_handle* h);
// comm.cu
int exchange_to_planes(chan_handle* h, int nf, void* const* pencil_fields);      // -> h->rcv[0..nf)
int exchange_to_pencils(chan_handle* h, int nf, void* const* pencil_fields);     // h->snd[0..nf) ->
int allreduce_umax(chan_handle* h);
int stream_barrier(chan_handle* h);                 // all ranks' work issued so far has completed
int allreduce_sum_host(chan_handle* h, double* dev_buf, int n);
int launch_stats(chan_handle* h, double* out4ny_dev);
}  // namespace chan
