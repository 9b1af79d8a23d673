// Wall-parallel (x,z) stages of the RK3 substep: the 6 inverse and 3 forward
// 2-D FFTs of convolution.c:24-41, the 3/2-rule zero-pad / truncate
// (dealias.cu:9-45 seen from the other side: only active modes are stored, the
// pad happens in the load and the truncation in the store), the physical-space
// product u x omega (rotorkernel, convolution_kernels.cu:69-148), the velocity
// maxima of calcUmax (fft.c:143-169) and imposeSymetry (imposeSymetry.c:5-18).
//
// All transforms are hand-written shared-memory mixed-radix FFTs (radices up to
// 16 in registers, fft_core.cuh); cuFFT is not linked.  Design points (details
// in plane_core.cuh):
//  * first / last stages fused with the global loads and stores; in z the last
//    inverse stage, u x omega and the first forward stage are one register pass,
//    so a 384-point line makes 3 shared-memory round trips instead of 10;
//  * inverse = decimation in time, forward = decimation in frequency with the
//    same radix schedule: physical space is in natural order and no reorder pass
//    exists;
//  * batch-fastest shared-memory layout s[pos][line]: conflict-free at every
//    stride;
//  * two-for-one real transforms in z: the same field at two neighbouring x is
//    one complex line (a + i b);
//  * the physical fields never touch HBM; the x-transformed staging
//    Q[f][plane][x][kz] is produced and consumed per plane batch (as many planes
//    as fit in 1.5 GiB: the kernels are not bandwidth-bound, large launches win
//    over L2 residency).
#include "plane_core.cuh"
// 512 threads x 128 registers = one full register file: the bound leaves the register allocation where (256, 2) had it and
// lets long transforms (2NZ-2 = 1536: one block per SM for shared memory) run with more than 8 warps
#ifndef ZMAXT
#define ZMAXT 512
#endif
#ifndef ZMINB
#define ZMINB 1
#endif

namespace chan {
using namespace fft;

template <typename T>
struct XArgs {
  typedef typename Vec2<T>::type T2;
  const T2* src[6]; SegDst<typename Vec2<T>::type> dst[3];   // plane-side field buffers (see SegMap)
  SegMap seg, seg_out; int j_glob0;        // global index of local plane 0 (export only)
  T2* Q;                          // [f][pb][NX][Kz]
  FftPlan px;
  int nx, Kx, Kz, kxmax, j_start, pb; long M;
  int dbg_skip;                   // timing ablation only (CHAN_TUNE_X_SKIP bit 0: inverse, bit 1: forward returns at once)
  T scale;                        // forward only: divide by this when != 0 (normalizekernel, dealias.cu:72-97)
};

extern __shared__ __align__(16) unsigned char smem_raw[];

__device__ __forceinline__ FastDiv plan_span(const FftPlan& p, int t) { FastDiv f; f.d = (unsigned)p.span[t]; f.M = p.span_magic[t]; return f; }


// x inverse: compact spectral rows -> (zero pad) -> natural-order physical x line, still spectral in z
template <typename T, int TK, bool W5>
__global__ void __launch_bounds__(256, 2) xfft_inv_kernel(XArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  T2* s = (T2*)smem_raw;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int kz0 = blockIdx.x * TK, jb = blockIdx.y, f = blockIdx.z;
  CHAN_ABLATE_RETURN(a.dbg_skip & 1);
  const PlanT<T2> pl = plan_of<T2>(a.px);
  const T2* src = a.src[f];
  T2* q = a.Q + ((size_t)(f * a.pb + jb) * a.nx) * a.Kz;
  FFT_DISPATCH_RADIX(W5, pl.r_first, (x_inv_load_stage0<T2, RR>(s, src, a.seg, a.j_start + jb, pl, TK, kz0, a.Kz, a.Kx, a.kxmax, tid, nthr)));
  __syncthreads();
  for (int t = 1; t + 1 < pl.nstage; t++) { plan_stage_x<T2, +1, true, TK, W5>(s, pl, a.px.radix[t], a.px.span[t], tid, nthr); __syncthreads(); }
  FFT_DISPATCH_RADIX(W5, pl.r_last, (x_inv_last_store<T2, RR>(s, q, pl, TK, kz0, a.Kz, tid, nthr)));
}

// x forward: natural-order x line -> truncated compact spectral rows
template <typename T, int TK, bool W5>
__global__ void __launch_bounds__(256, 2) xfft_fwd_kernel(XArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  T2* s = (T2*)smem_raw;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int kz0 = blockIdx.x * TK, jb = blockIdx.y, f = blockIdx.z;
  CHAN_ABLATE_RETURN(a.dbg_skip & 2);
  const PlanT<T2> pl = plan_of<T2>(a.px);
  const T2* q = a.Q + ((size_t)(f * a.pb + jb) * a.nx) * a.Kz;
  FFT_DISPATCH_RADIX(W5, pl.r_last, (x_fwd_load_first<T2, RR>(s, q, pl, TK, kz0, a.Kz, tid, nthr)));
  __syncthreads();
  for (int t = pl.nstage - 2; t >= 1; t--) { plan_stage_x<T2, -1, false, TK, W5>(s, pl, a.px.radix[t], a.px.span[t], tid, nthr); __syncthreads(); }
  FFT_DISPATCH_RADIX(W5, pl.r_first, (x_fwd_stage0_store<T2, RR>(s, a.dst[f], a.seg_out, a.j_start + jb, pl, TK, kz0, a.Kz, a.Kx, a.kxmax, a.scale, tid, nthr)));
}

// -------------------------------------------------------------------------------
template <typename T>
struct ZArgs {
  typedef typename Vec2<T>::type T2;
  T2* Q; FftPlan pz;
  int dbg_stop;                    // >0: timing ablation only (CHAN_TUNE_Z_STOP): return after phase N
  int nx, nzp, Kz, pb, j_start, XP, SS;   // SS: shared-memory elements per position (>= XP*NIN)
  int SSo, SKo;                    // forward-side row pitch and skew multiplier (plane_core.cuh)
  int sha;                         // inverse-side extra skew row >> sha (-1: none, rows padded to an odd pitch instead)
  FastDiv f_lin, f_lout, f_ng, f_mid, f_kz;   // lines in / out, n/r_first, n/r_last, Kz (index arithmetic without divisions)
  T n2;                            // (T)N2: the reference divides by the int N2 (convolution_kernels.cu:107-121)
  unsigned long long* umax_bits;   // 3 slots
  T* phys; int nz, j_glob0;        // export only: real [NY][NX][2NZ]
};

__device__ __forceinline__ void atomic_max_pos(unsigned long long* slot, double v) {
  atomicMax(slot, (unsigned long long)__double_as_longlong(v));   // v >= 0: bit order == value order
}

template <typename T, int MODE, bool W5>
__global__ void __launch_bounds__(ZMAXT, ZMINB) zstage_kernel(ZArgs<T> a) {
  typedef typename Vec2<T>::type T2;
  constexpr int NIN = MODE == ZM_ROTOR ? 6 : (MODE == ZM_IDENTITY ? 2 : 1);
  constexpr int NOUT = MODE == ZM_ROTOR ? 3 : (MODE == ZM_IDENTITY ? 2 : 0);
  T2* s = (T2*)smem_raw;
  __shared__ unsigned long long red[3];
  if (threadIdx.x < 3) red[threadIdx.x] = 0ull;          // barriers follow before its use
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int xp0 = blockIdx.x * a.XP, jb = blockIdx.y;
  CHAN_ABLATE_RETURN(a.dbg_stop == 9);
  const int nxp_valid = min(a.XP, a.nx / 2 - xp0);
  const int ST = a.XP * NIN;
  const PlanT<T2> pl = plan_of<T2>(a.pz);
  const size_t fstride = (size_t)a.pb * a.nx * a.Kz;
  T2* q0 = a.Q + ((size_t)jb * a.nx + 2 * xp0) * a.Kz;     // field 0, x = 2*xp0
  FFT_DISPATCH_RADIX(W5, pl.r_first, (z_inv_load_stage0<T2, RR, NIN>(s, q0, fstride, pl, a.f_ng, a.XP, a.SS, a.sha, nxp_valid, a.Kz, tid, nthr)));
  __syncthreads();
  CHAN_ABLATE_RETURN(a.dbg_stop == 1);
  for (int t = 1; t + 1 < pl.nstage; t++) { plan_stage<T2, +1, true, W5>(s, pl, a.pz.radix[t], plan_span(a.pz, t), a.pz.tws[t], a.SS, a.f_lin, ST, ST, tid, nthr, 1, a.sha); __syncthreads(); }
  CHAN_ABLATE_RETURN(a.dbg_stop == 2);
  T um[3] = { 0, 0, 0 };
  T* phys = 0; size_t pxs = 0;
  if (MODE == ZM_EXPORT) { pxs = (size_t)2 * a.nz; phys = a.phys + ((size_t)(a.j_glob0 + a.j_start + jb) * a.nx + 2 * xp0) * pxs; }
  switch (pl.r_last) {
    case 2: z_mid_fused<T2, 2, MODE>(s, s, pl, a.f_mid, a.XP, a.SS, a.sha, a.SSo, a.SKo, nxp_valid, a.n2, um, phys, pxs, tid, nthr); break;
    case 3: z_mid_fused<T2, 3, MODE>(s, s, pl, a.f_mid, a.XP, a.SS, a.sha, a.SSo, a.SKo, nxp_valid, a.n2, um, phys, pxs, tid, nthr); break;
    default: z_mid_fused<T2, 4, MODE>(s, s, pl, a.f_mid, a.XP, a.SS, a.sha, a.SSo, a.SKo, nxp_valid, a.n2, um, phys, pxs, tid, nthr); break;
  }
  if (MODE == ZM_EXPORT) return;
  CHAN_ABLATE_RETURN(a.dbg_stop == 3);
  if (MODE == ZM_ROTOR) {      // block max -> one atomic per block and component
    // values are >= 0, so the bit pattern orders like the value: warp maximum of the high words, then of
    // the low words among the lanes that hold it (two redux instead of ten 64-bit shuffle steps)
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const double v = (double)um[c];
      const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
      const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
      const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
      if ((tid & 31) == 0) atomicMax(&red[c], ((unsigned long long)mh << 32) | ml);
    }
  }
  __syncthreads();
  if (MODE == ZM_ROTOR && tid < 3) atomicMax(a.umax_bits + tid, red[tid]);
  for (int t = pl.nstage - 2; t >= 0; t--) { plan_stage<T2, -1, false, W5>(s, pl, a.pz.radix[t], plan_span(a.pz, t), a.pz.tws[t], a.SSo, a.f_lout, NOUT, NOUT, tid, nthr, a.SKo); __syncthreads(); }
  CHAN_ABLATE_RETURN(a.dbg_stop == 4);
  z_unpack_store<T2, NIN, NOUT>(s, q0, fstride, pl, a.f_kz, a.XP, a.SSo, a.SKo, nxp_valid, tid, nthr);
}

// -------------------------------------------------------------------------------
// kz columns per x-transform block: the widest tile that still lets three blocks share an SM (sweeps
// at NX = 384 and 1024, profiles/r01_sweeps.txt); whatever fits when NX is so long that no tile does
static int pick_tk(chan_handle* h, size_t elem) {
  const int top = h->tune.x_tk > 0 ? h->tune.x_tk : 8;
  if (h->tune.x_tk <= 0)
    for (int tk = top; tk >= 2; tk >>= 1)
      if ((size_t)h->d.nx * tk * elem <= 72 * 1024) return tk;
  for (int tk = top; tk >= 2; tk >>= 1)
    if ((size_t)h->d.nx * tk * elem <= 200 * 1024) return tk;
  return 0;
}
// 128 threads (four blocks per SM) up to 3072-element tiles, 256 above (measured at NX = 384 and 1024)
static int pick_x_threads(chan_handle* h, int tk) {
  if (h->tune.x_threads > 0) return h->tune.x_threads;
  return h->d.nx * tk >= 4096 ? 256 : 128;
}

template <typename T, int TK, bool W5>
static int launch_x(chan_handle* h, bool inverse, XArgs<T>& a, int nf, int pb) {
  typedef typename Vec2<T>::type T2;
  size_t sm = (size_t)h->d.nx * TK * sizeof(T2);
  dim3 grid((h->d.Kz + TK - 1) / TK, pb, nf);
  if (inverse) {
    CHAN_CUDA(h, cudaFuncSetAttribute(xfft_inv_kernel<T, TK, W5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    xfft_inv_kernel<T, TK, W5><<<grid, pick_x_threads(h, TK), sm, h->stream>>>(a);
  } else {
    CHAN_CUDA(h, cudaFuncSetAttribute(xfft_fwd_kernel<T, TK, W5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    xfft_fwd_kernel<T, TK, W5><<<grid, pick_x_threads(h, TK), sm, h->stream>>>(a);
  }
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}
template <typename T, bool W5>
static int launch_x_tk(chan_handle* h, bool inverse, XArgs<T>& a, int nf, int pb) {
  typedef typename Vec2<T>::type T2;
  int tk = pick_tk(h, sizeof(T2));
  if (tk == 8) return launch_x<T, 8, W5>(h, inverse, a, nf, pb);
  if (tk == 4) return launch_x<T, 4, W5>(h, inverse, a, nf, pb);
  if (tk == 2) return launch_x<T, 2, W5>(h, inverse, a, nf, pb);
  return set_err(h, CHAN_ERR_ARG, "NX=%d too large for the shared-memory x transform", h->d.nx);
}
// the kernel instances with radix-5/7/10/14 butterflies serve the lengths with a factor 5 or 7 only (FFT_DISPATCH_RADIX)
template <typename T>
static int launch_x_any(chan_handle* h, bool inverse, XArgs<T>& a, int nf, int pb) {
  return (h->d.nx % 5 == 0 || h->d.nx % 7 == 0) ? launch_x_tk<T, true>(h, inverse, a, nf, pb) : launch_x_tk<T, false>(h, inverse, a, nf, pb);
}

template <typename T, int MODE, bool W5>
static int launch_z_w(chan_handle* h, ZArgs<T>& a, int nin, int pb) {
  typedef typename Vec2<T>::type T2;
  // x-pairs per block: enough positions for the register-fused middle stage to occupy the block,
  // bounded by ~96 KB of shared memory so that two blocks fit on an SM
  const int n = h->d.nzp, m = n / h->pz.radix[h->pz.nstage - 1];
  int xp = h->tune.z_xp > 0 ? h->tune.z_xp : (96 + m - 1) / m;
  while (xp > 1 && (size_t)n * (xp * nin + 1) * sizeof(T2) > 96 * 1024) xp--;
  if (xp > h->d.nx / 2) xp = h->d.nx / 2;
  // one thread per item of the register-fused middle stage (m positions x xp pairs), within 128..256
  int threads = h->tune.z_threads;
  // (up to 384: at 2NZ-2 = 1536 only one block fits an SM and 256 threads left it with 8 warps: 14.3 -> 11.1 ms per substep
  // on a 256x633x769 grid with 384, 11.5 with 512)
  if (threads <= 0) { threads = (m * xp + 31) / 32 * 32; threads = threads < 128 ? 128 : (threads > 384 ? 384 : threads); }
  // Row layouts (plane_core.cuh).  Packed rows need the fused middle stage to be a single round (its barrier then
  // separates ALL reads of the inverse side from the writes of the forward side, whatever the two layouts are).
  const int nout = MODE == ZM_ROTOR ? 3 : (MODE == ZM_IDENTITY ? 2 : 1), lout = xp * nout;
  const bool single = m * xp <= threads && MODE == ZM_ROTOR;
  int ss = xp * nin, sha = -1;
  if (!(ss & 1)) {
    if (h->tune.z_inpack && single && m % 16 == 0) sha = (ss % 8 == 0) ? 0 : ((ss % 8 == 4) ? 1 : 2);   // the shifts divide every span
    else if (h->tune.z_pad) ss++;            // an odd pitch keeps the lanes of the middle stage (consecutive rows) apart
  }
  const bool pack = h->tune.z_fwdpack && (lout & 1) && single;
  a.XP = xp; a.SS = ss; a.sha = sha;
  a.SSo = pack ? lout : (ss | 1); a.SKo = pack ? lout : 1;
  const size_t nsk = n / h->pz.radix[0];
  const size_t ext_in = (size_t)n * ss + nsk + (sha >= 0 ? (n >> sha) : 0), ext_out = (size_t)n * a.SSo + a.SKo * nsk + lout;
  size_t sm = (ext_in > ext_out ? ext_in : ext_out) * sizeof(T2);
  if (sm > 220 * 1024) return set_err(h, CHAN_ERR_ARG, "2NZ-2=%d too large for the shared-memory z transform", n);
  a.f_lin = make_fastdiv(xp * nin); a.f_lout = make_fastdiv(xp * (MODE == ZM_ROTOR ? 3 : (MODE == ZM_IDENTITY ? 2 : 1)));
  a.f_ng = make_fastdiv(n / h->pz.r_first); a.f_mid = make_fastdiv(m); a.f_kz = make_fastdiv(h->d.Kz);
  CHAN_CUDA(h, cudaFuncSetAttribute(zstage_kernel<T, MODE, W5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  dim3 grid((h->d.nx / 2 + xp - 1) / xp, pb);
  zstage_kernel<T, MODE, W5><<<grid, threads, sm, h->stream>>>(a);
  h->launches++;
  CHAN_CUDA(h, cudaGetLastError());
  return CHAN_OK;
}

template <typename T, int MODE>
static int launch_z(chan_handle* h, ZArgs<T>& a, int nin, int pb) {
  return (h->d.nzp % 5 == 0 || h->d.nzp % 7 == 0) ? launch_z_w<T, MODE, true>(h, a, nin, pb) : launch_z_w<T, MODE, false>(h, a, nin, pb);
}

template <typename T>
static void fill_x(chan_handle* h, XArgs<T>& a) {
  typedef typename Vec2<T>::type T2;
  a.Q = (T2*)h->Q; a.px = h->px; a.nx = h->d.nx; a.Kx = h->d.Kx; a.Kz = h->d.Kz; a.kxmax = h->d.kxmax;
  a.M = (long)h->d.Kx * h->d.Kz; a.scale = (T)0; a.dbg_skip = h->tune.x_skip; a.seg = h->seg; a.seg_out = h->seg; a.j_glob0 = h->d.j0;
  for (int f = 0; f < 3; f++) for (int g = 0; g < 8; g++) a.dst[f].p[g] = 0;
}
template <typename T>
static void fill_z(chan_handle* h, ZArgs<T>& a) {
  typedef typename Vec2<T>::type T2;
  a.Q = (T2*)h->Q; a.pz = h->pz; a.nx = h->d.nx; a.nzp = h->d.nzp; a.Kz = h->d.Kz; a.nz = h->d.nz;
  a.n2 = (T)(h->d.nx * h->d.nzp);
  a.umax_bits = h->sc->umax_bits;
  a.phys = 0; a.XP = 1; a.SS = 0; a.dbg_stop = h->tune.z_stop; a.j_glob0 = h->d.j0;
}

// forward stores go straight into the owners' pencil-side fields: segment g = rank g, rows are global
template <typename T>
static void peer_out(chan_handle* h, XArgs<T>& xf) {
  typedef typename Vec2<T>::type T2;
  xf.seg_out.bms = 0; xf.seg_out.bmw = (long)h->d.j0 * h->d.Kz;      // bases of the peers' pencil-side fields: j0 * width * Kz
  for (int g = 0; g < h->cfg.size; g++) {
    xf.seg_out.base[g] = (long)h->d.j0 * xf.seg_out.width[g] * h->d.Kz;
    for (int f = 0; f < 3; f++) xf.dst[f].p[g] = (T2*)h->peer_H[g][f];
  }
}

template <typename T>
static int nonlinear_t(chan_handle* h, bool op_level) {
  typedef typename Vec2<T>::type T2;
  XArgs<T> xi, xf; fill_x(h, xi); fill_x(h, xf);
  void* pin[6] = { h->u, h->v, h->w, h->ox, h->oy, h->oz };
  void* pout[3] = { h->ox, h->oy, h->oz };
  const bool multi = h->cfg.size > 1, peer = multi && h->peer_ready;
  // operator-level calls (chan_convolution) left their inputs in the LOCAL pencil fields: nothing has been
  // delivered to the plane owners, so the input side goes through the NCCL exchange even when the peer copies are on
  const bool piped = peer && !op_level && !h->tune.pre_direct && !h->tune.fwd_direct && h->tune.xchg_pipe;   // the inputs arrive batch by batch (chan_rk_substep)
  if (peer && !op_level && !piped) { int rc = stream_barrier(h); if (rc) return rc; }    // every rank's `pre` has delivered
  else if (multi) { int rc = exchange_to_planes(h, 6, pin); if (rc) return rc; }
  for (int f = 0; f < 6; f++) xi.src[f] = (const T2*)(multi ? h->rcv[f] : pin[f]);
  for (int f = 0; f < 3; f++) for (int g = 0; g < 8; g++) xf.dst[f].p[g] = (T2*)(multi ? h->snd[f] : pout[f]);
  const bool direct = peer && h->tune.fwd_direct;          // stores of xfft_fwd go straight into the pencil owners' fields (64-byte pieces)
  if (direct) peer_out<T>(h, xf);
  ZArgs<T> z; fill_z(h, z);
  if (op_level) z.umax_bits = h->sc->umax_scratch;         // the reference's convolution() does not touch calcDt's maxima
  for (int js = 0; js < h->d.nyl; js += h->plane_batch) {
    int pb = h->d.nyl - js < h->plane_batch ? h->d.nyl - js : h->plane_batch;
    xi.j_start = xf.j_start = z.j_start = js; xi.pb = xf.pb = z.pb = pb;
    if (piped) CHAN_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_in[js / h->plane_batch], 0));   // batch arrived from every rank
    int t = timing_begin(h, CHAN_STAGE_XINV);
    int rc = launch_x_any<T>(h, true, xi, 6, pb); if (rc) return rc;
    timing_end(h, t); t = timing_begin(h, CHAN_STAGE_Z);
    rc = launch_z<T, ZM_ROTOR>(h, z, 6, pb); if (rc) return rc;
    timing_end(h, t); t = timing_begin(h, CHAN_STAGE_XFWD);
    rc = launch_x_any<T>(h, false, xf, 3, pb); if (rc) return rc;
    timing_end(h, t);
    if (peer && !direct) {
      // plane -> pencil transpose as bulk peer copies, batch by batch on a side stream: the products of this batch cross
      // NVLink while the next batch is being transformed
      const int ib = js / h->plane_batch;
      cudaStream_t main_stream = h->stream;
      if (ib < 64) {
        CHAN_CUDA(h, cudaEventRecord(h->ev_batch[ib], main_stream)); CHAN_CUDA(h, cudaStreamWaitEvent(h->xs, h->ev_batch[ib], 0));
        // the products of these planes land in the peers' omega fields, over the very rows the peers sent for THIS batch:
        // their copies are complete (the barrier of batch ib preceded the transforms), later batches touch other rows
        h->stream = h->xs;
      }
      t = timing_begin(h, CHAN_STAGE_XCHG_BACK);
      rc = launch_peer_scatter_back(h, js, pb);
      timing_end(h, t);
      h->stream = main_stream;
      if (rc) return rc;
    }
  }
  if (piped) CHAN_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_in[h->nbatch - 1], 0));   // every outgoing copy of this rank is done: the next `pre` may overwrite its sources
  if (peer && !direct) { CHAN_CUDA(h, cudaEventRecord(h->ev_xs, h->xs)); CHAN_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_xs, 0)); }
  if (multi) { int rc = allreduce_umax(h); if (rc) return rc; }               // also the barrier after the peer copies
  if (multi && !peer) { int rc = exchange_to_pencils(h, 3, pout); if (rc) return rc; }
  return CHAN_OK;
}

template <typename T>
static int symmetrize_t(chan_handle* h, void* fa, void* fb) {
  typedef typename Vec2<T>::type T2;
  XArgs<T> xi, xf; fill_x(h, xi); fill_x(h, xf);
  for (int f = 0; f < 6; f++) { xi.src[f] = 0; xf.src[f] = 0; }
  void* pf[2] = { fa, fb };
  const bool multi = h->cfg.size > 1;
  if (multi) { int rc = exchange_to_planes(h, 2, pf); if (rc) return rc; }
  xi.src[0] = (const T2*)(multi ? h->rcv[0] : fa); xi.src[1] = (const T2*)(multi ? h->rcv[1] : fb);
  for (int g = 0; g < 8; g++) { xf.dst[0].p[g] = (T2*)(multi ? h->snd[0] : fa); xf.dst[1].p[g] = (T2*)(multi ? h->snd[1] : fb); }
  xf.scale = (T)(h->d.nx * h->d.nzp);
  ZArgs<T> z; fill_z(h, z);
  for (int js = 0; js < h->d.nyl; js += h->plane_batch) {
    int pb = h->d.nyl - js < h->plane_batch ? h->d.nyl - js : h->plane_batch;
    xi.j_start = xf.j_start = z.j_start = js; xi.pb = xf.pb = z.pb = pb;
    int rc = launch_x_any<T>(h, true, xi, 2, pb); if (rc) return rc;
    rc = launch_z<T, ZM_IDENTITY>(h, z, 2, pb); if (rc) return rc;
    rc = launch_x_any<T>(h, false, xf, 2, pb); if (rc) return rc;
  }
  if (multi) { int rc = exchange_to_pencils(h, 2, pf); if (rc) return rc; }
  return CHAN_OK;
}

template <typename T>
static int export_t(chan_handle* h, const void* f, void* phys) {
  typedef typename Vec2<T>::type T2;
  if (h->cfg.size > 1) return set_err(h, CHAN_ERR_STATE, "chan_fft_backward_transpose is a single-GPU validation entry point");
  XArgs<T> xi; fill_x(h, xi);
  for (int k = 0; k < 6; k++) xi.src[k] = 0;
  xi.src[0] = (const T2*)f;
  ZArgs<T> z; fill_z(h, z); z.phys = (T*)phys;
  for (int js = 0; js < h->d.nyl; js += h->plane_batch) {
    int pb = h->d.nyl - js < h->plane_batch ? h->d.nyl - js : h->plane_batch;
    xi.j_start = z.j_start = js; xi.pb = z.pb = pb;
    int rc = launch_x_any<T>(h, true, xi, 1, pb); if (rc) return rc;
    rc = launch_z<T, ZM_EXPORT>(h, z, 1, pb); if (rc) return rc;
  }
  return CHAN_OK;
}

int launch_nonlinear_planes(chan_handle* h, bool op_level) { return h->prec == 4 ? nonlinear_t<float>(h, op_level) : nonlinear_t<double>(h, op_level); }
int launch_symmetrize(chan_handle* h, void* a, void* b) { return h->prec == 4 ? symmetrize_t<float>(h, a, b) : symmetrize_t<double>(h, a, b); }
int launch_to_physical(chan_handle* h, const void* f, void* phys) { return h->prec == 4 ? export_t<float>(h, f, phys) : export_t<double>(h, f, phys); }

}  // namespace chan
