// C ABI of libchannel_b200 (include/channel_b200.h): handle life cycle, state
// in/out in the reference layout, the RK3 driver entry points and the
// operator-level drop-ins.  Host code only; kernels live in kernels_*.cu.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include "chan_internal.h"

using namespace chan;

static size_t csize(const chan_handle* h) { return 2 * (size_t)h->prec; }                      // bytes per complex
static size_t field_bytes(const chan_handle* h) { return (size_t)h->d.ny * h->d.M * csize(h); } // compact field

extern "C" int chan_partition(int n, int size, int rank, int* start, int* count) {
  if (size <= 0 || rank < 0 || rank >= size || n < 0) return CHAN_ERR_ARG;
  int q = n / size, r = n % size;
  int s = rank * q + (rank < r ? rank : r);
  int c = q + (rank < r ? 1 : 0);
  if (start) *start = s;
  if (count) *count = c;
  return CHAN_OK;
}
extern "C" int chan_active_kx(int nx) { return 2 * (nx / 3) + 1; }
extern "C" int chan_active_kz(int nz) { return (2 * nz - 2) / 3 + 1; }

extern "C" int chan_exchange_layout(int nx, int ny, int nz, int size, int rank, int peer, long* out) {
  if (!out || size < 1 || rank < 0 || rank >= size || peer < 0 || peer >= size) return CHAN_ERR_ARG;
  const int Kx = chan_active_kx(nx), Kz = chan_active_kz(nz);
  int kx0, kxl, j0, nyl, pkx0, pkxl, pj0, pnyl;
  chan_partition(Kx, size, rank, &kx0, &kxl); chan_partition(ny, size, rank, &j0, &nyl);
  chan_partition(Kx, size, peer, &pkx0, &pkxl); chan_partition(ny, size, peer, &pj0, &pnyl);
  const long M = (long)kxl * Kz;
  out[0] = (long)pj0 * M; out[1] = (long)pnyl * M;                 // my rows of the peer's planes: contiguous
  long off = 0;                                                    // blocks of lower-ranked sources come first
  for (int r = 0; r < peer; r++) { int s, c; chan_partition(Kx, size, r, &s, &c); off += (long)nyl * c * Kz; }
  out[2] = off; out[3] = (long)nyl * pkxl * Kz;
  out[4] = pkx0; out[5] = pkxl; out[6] = pj0; out[7] = pnyl;
  return CHAN_OK;
}

// the same exchange seen from the sender: bulk peer copies write straight into the destination's buffers
extern "C" int chan_peer_put_layout(int nx, int ny, int nz, int size, int rank, int dest, long* out) {
  long L[8], Lp[8];
  if (!out) return CHAN_ERR_ARG;
  int rc = chan_exchange_layout(nx, ny, nz, size, rank, dest, L); if (rc) return rc;
  if ((rc = chan_exchange_layout(nx, ny, nz, size, dest, rank, Lp))) return rc;
  out[0] = L[0]; out[1] = Lp[2]; out[2] = L[1];      // pencil -> plane: my rows of dest's planes -> dest's block for source `rank`
  out[3] = L[2]; out[4] = Lp[0]; out[5] = L[3];      // plane -> pencil: my block for dest's kx slab -> rows of my planes in dest's field
  return CHAN_OK;
}

extern "C" int chan_config_default(chan_config* c) {
  if (!c) return CHAN_ERR_ARG;
  memset(c, 0, sizeof(*c));
  c->nx = 128; c->ny = 128; c->nz = 65;            // run.conf:3-5
  c->precision = 8;
  c->reynolds = 3250.0; c->qvelocity = 1.8; c->cfl = 0.5;
  c->rank = 0; c->size = 1; c->device = -1; c->plane_batch = 0;
  return CHAN_OK;
}

static int alloc_dev(chan_handle* h, void** p, size_t bytes) {
  if (cudaMalloc(p, bytes) != cudaSuccess) { cudaGetLastError(); return set_err(h, CHAN_ERR_ALLOC, "cudaMalloc of %zu bytes failed", bytes); }
  CHAN_CUDA(h, cudaMemsetAsync(*p, 0, bytes, h->stream));
  return CHAN_OK;
}

extern "C" int chan_create(const chan_config* cfg, chan_handle** out) {
  if (!cfg || !out) return CHAN_ERR_ARG;
  *out = 0;
  chan_handle* h = new (std::nothrow) chan_handle;
  if (!h) return CHAN_ERR_ALLOC;
  memset(h, 0, sizeof(*h));
  *out = h;                                         // returned even on failure so the caller can read the message
  h->cfg = *cfg;
  if (cfg->precision != 4 && cfg->precision != 8) return set_err(h, CHAN_ERR_ARG, "precision must be 4 or 8, got %d", cfg->precision);
  if (cfg->nx < 8 || cfg->ny < 8 || cfg->nz < 5 || (cfg->nx & 1)) return set_err(h, CHAN_ERR_ARG, "grid %dx%dx%d: need even NX>=8, NY>=8, NZ>=5", cfg->nx, cfg->ny, cfg->nz);
  if (cfg->size < 1 || cfg->size > 8 || cfg->rank < 0 || cfg->rank >= cfg->size) return set_err(h, CHAN_ERR_ARG, "rank %d of %d: 1..8 ranks (one NVSwitch box) are supported", cfg->rank, cfg->size);
  if (cfg->size > 1 && !cfg->nccl_comm) return set_err(h, CHAN_ERR_ARG, "size > 1 needs chan_config.nccl_comm");
  h->nccl_comm = cfg->nccl_comm;
  h->prec = cfg->precision;
  {   // launch shapes and kernel choices; CHAN_TUNE_* environment variables override (scripts/sweep_*.sh, tests)
    struct { const char* name; int* v; int def; bool ablate; } knobs[] = {
      { "CHAN_TUNE_Z_THREADS", &h->tune.z_threads, 0, false },
      { "CHAN_TUNE_Z_XP", &h->tune.z_xp, 0, false },
      { "CHAN_TUNE_Z_PAD", &h->tune.z_pad, 1, false },
      { "CHAN_TUNE_X_THREADS", &h->tune.x_threads, 0, false },
      { "CHAN_TUNE_X_TK", &h->tune.x_tk, 0, false },
      { "CHAN_TUNE_MAX_RADIX", &h->tune.max_radix, 16, false },
      { "CHAN_TUNE_PENCIL_SEG", &h->tune.pencil_seg, -1, false },
      { "CHAN_TUNE_SEG_SHAPE", &h->tune.seg_shape, 0, false },
      { "CHAN_TUNE_POST_PF", &h->tune.post_pf, 0, false },
      { "CHAN_TUNE_POST_THREADS", &h->tune.post_threads, 0, false },
      { "CHAN_TUNE_POST_DEEP", &h->tune.post_deep, -1, false },
      { "CHAN_TUNE_Z_FWDPACK", &h->tune.z_fwdpack, 1, false },
      { "CHAN_TUNE_Z_INPACK", &h->tune.z_inpack, 1, false },
      { "CHAN_TUNE_PRE_DIRECT", &h->tune.pre_direct, 0, false },
      { "CHAN_TUNE_FWD_DIRECT", &h->tune.fwd_direct, 0, false },
      { "CHAN_TUNE_XCHG_PIPE", &h->tune.xchg_pipe, 0, false },
      { "CHAN_TUNE_XCHG_CE", &h->tune.xchg_ce, 0, false },
      { "CHAN_TUNE_LANES", &h->tune.lanes, -1, false },
      { "CHAN_TUNE_LANES_S", &h->tune.lanes_s, 4, false },
      { "CHAN_TUNE_LANES_MB", &h->tune.lanes_mb, 0, false },
      // timing ablation: kernels return early (results are garbage) - only in -DCHAN_ABLATE builds
      { "CHAN_TUNE_Z_STOP", &h->tune.z_stop, 0, true },
      { "CHAN_TUNE_X_SKIP", &h->tune.x_skip, 0, true },
      { "CHAN_TUNE_POST_STOP", &h->tune.post_stop, 0, true } };
    for (auto& k : knobs) {
      const char* e = getenv(k.name);
      *k.v = e ? atoi(e) : k.def;
#ifndef CHAN_ABLATE
      if (k.ablate && e && atoi(e) != 0)
        return set_err(h, CHAN_ERR_ARG, "%s is set but this library was built without -DCHAN_ABLATE: it would make kernels return early", k.name);
#endif
    }
  }
  Dims& d = h->d;
  d.nx = cfg->nx; d.ny = cfg->ny; d.nz = cfg->nz; d.nzp = 2 * cfg->nz - 2;
  d.kxmax = d.nx / 3; d.Kx = 2 * d.kxmax + 1; d.Kz = d.nzp / 3 + 1;          // dealias.cu:37 (integer divisions)
  if (d.Kz > d.nz - 1 || d.kxmax >= d.nx / 2) return set_err(h, CHAN_ERR_ARG, "grid too small for the 2/3 rule");
  // slabs: active kx for the pencil stages, y planes for the plane stage (main.c:47, channel.h:71-79),
  // split as evenly as possible — no divisibility requirement
  for (int r = 0; r <= cfg->size; r++) {
    int s0 = d.Kx, c0 = 0, s1 = d.ny, c1 = 0;
    if (r < cfg->size) { chan_partition(d.Kx, cfg->size, r, &s0, &c0); chan_partition(d.ny, cfg->size, r, &s1, &c1); }
    h->part_kx[r] = s0; h->part_j[r] = s1;
  }
  d.ikx0 = h->part_kx[cfg->rank]; d.Kxl = h->part_kx[cfg->rank + 1] - d.ikx0;
  d.j0 = h->part_j[cfg->rank]; d.nyl = h->part_j[cfg->rank + 1] - d.j0;
  d.M = (long)d.Kxl * d.Kz;
  if (d.Kxl < 1 || d.nyl < 1) return set_err(h, CHAN_ERR_ARG, "grid too small for %d ranks", cfg->size);
  h->seg.nseg = cfg->size;
  { long off = 0;
    for (int r = 0; r < cfg->size; r++) {
      h->seg.start[r] = h->part_kx[r]; h->seg.width[r] = h->part_kx[r + 1] - h->part_kx[r];
      h->seg.base[r] = off; off += (long)d.nyl * h->seg.width[r] * d.Kz;
    }
    h->seg.start[cfg->size] = d.Kx; }
  h->seg.closed = 0;
  if (cfg->size >= 4) {      // from four segments on the closed form beats the select chain (with two the chain is four instructions)
    h->seg.closed = 1; h->seg.pq = d.Kx / cfg->size; h->seg.pr = d.Kx % cfg->size;
    h->seg.mq = (unsigned)(0x100000000ull / (unsigned)h->seg.pq) + 1u; h->seg.mq1 = (unsigned)(0x100000000ull / (unsigned)(h->seg.pq + 1)) + 1u;
    if (h->seg.pq == 1) h->seg.mq = 0;                                  // FastDiv convention: M = 0 encodes d = 1
    h->seg.bms = (long)d.nyl * d.Kz; h->seg.bmw = 0;
  }
  if (cfg->device >= 0) CHAN_CUDA(h, cudaSetDevice(cfg->device));
  CHAN_CUDA(h, cudaGetDevice(&h->device));
  CHAN_CUDA(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  {   // the side stream carries one-block kernels that must not queue behind a GPU-filling kernel of the main stream
    int lo = 0, hi = 0;
    CHAN_CUDA(h, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CHAN_CUDA(h, cudaStreamCreateWithPriority(&h->aux, cudaStreamNonBlocking, hi));
  }
  {   // the peer-copy streams outrank the main stream: their few blocks must not queue behind the thousands of an FFT launch
    int lo = 0, hi = 0;
    CHAN_CUDA(h, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CHAN_CUDA(h, cudaStreamCreateWithPriority(&h->xs, cudaStreamNonBlocking, hi));
    CHAN_CUDA(h, cudaStreamCreateWithPriority(&h->xo, cudaStreamNonBlocking, hi));
  }
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_xs, cudaEventDisableTiming));
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_cf, cudaEventDisableTiming));
  for (int i = 0; i < 4; i++) { CHAN_CUDA(h, cudaStreamCreateWithFlags(&h->xc[i], cudaStreamNonBlocking)); CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_cj[i], cudaEventDisableTiming)); }
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_pre, cudaEventDisableTiming));
  for (int i = 0; i < 64; i++) { CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_batch[i], cudaEventDisableTiming)); CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming)); }
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_sub, cudaEventDisableTiming));
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_piv, cudaEventDisableTiming));
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
  CHAN_CUDA(h, cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
  int rc;
  const size_t fb = field_bytes(h);
  void** fields[11] = { &h->phi, &h->g, &h->v, &h->dv, &h->Rphi, &h->Rg, &h->u, &h->w, &h->ox, &h->oy, &h->oz };
  for (int i = 0; i < 11; i++) if ((rc = alloc_dev(h, fields[i], fb))) return rc;
  const size_t chain = (size_t)d.ny * ((2 * d.M + 31) / 32 * 32) * sizeof(double);   // whole warp tiles (pencil_post scratch layout)
  if ((rc = alloc_dev(h, (void**)&h->sA, chain))) return rc;
  if ((rc = alloc_dev(h, (void**)&h->sB, chain))) return rc;
  if ((rc = alloc_dev(h, (void**)&h->sC, chain))) return rc;
  if ((rc = alloc_dev(h, (void**)&h->sP, chain / 2))) return rc;
  // plane batch: the plane kernels are latency/issue-bound, not bandwidth-bound (profiles/r01_*): larger
  // launches win over L2 residency of the staging, so take as many planes per launch group as fit in
  // 1.5 GiB of staging (all of them up to Re_tau=550; 10 at Re_tau=1000 in fp64)
  const size_t plane_q = (size_t)6 * d.nx * d.Kz * csize(h);
  h->nbatch = 1;
  for (int r = 0; r < cfg->size; r++) {      // the same arithmetic on every rank: senders address the batches of their peers
    const int nyl_r = h->part_j[r + 1] - h->part_j[r];
    int pbr = cfg->plane_batch > 0 ? cfg->plane_batch : (int)((size_t)(1536u << 20) / plane_q);
    if (pbr < 1) pbr = 1;
    if (pbr > nyl_r) pbr = nyl_r;
    // several GPUs: at least four batches (of at least four planes), so that the peer copies of one batch overlap the
    // transforms of the next (kernels_plane.cu)
    if (cfg->size > 1 && cfg->plane_batch <= 0 && h->tune.xchg_pipe) { int q = (nyl_r + 3) / 4; if (q < 4) q = 4; if (pbr > q) pbr = q; }
    h->pb_rank[r] = pbr;
    const int nb = (nyl_r + pbr - 1) / pbr; if (nb > h->nbatch) h->nbatch = nb;
  }
  const int pb = h->pb_rank[cfg->rank];
  h->plane_batch = pb;
  if ((rc = alloc_dev(h, &h->Q, plane_q * pb))) return rc;
  if (cfg->size > 1) {
    const size_t pb_bytes = (size_t)d.nyl * d.Kx * d.Kz * csize(h);
    for (int f = 0; f < 6; f++) if ((rc = alloc_dev(h, &h->rcv[f], pb_bytes))) return rc;
    for (int f = 0; f < 3; f++) if ((rc = alloc_dev(h, &h->snd[f], pb_bytes))) return rc;
    if ((rc = alloc_dev(h, (void**)&h->barrier_word, 256))) return rc;
  }
  const size_t rb = (size_t)d.ny * h->prec;
  if ((rc = alloc_dev(h, &h->mean_u, rb))) return rc;
  if ((rc = alloc_dev(h, &h->mean_N, rb))) return rc;
  if ((rc = alloc_dev(h, &h->mean_uw, rb))) return rc;
  if ((rc = alloc_dev(h, &h->mean_tmp, 6 * rb))) return rc;
  if ((rc = alloc_dev(h, (void**)&h->sc, sizeof(Scalars)))) return rc;
  h->io_bytes = (size_t)d.Kxl * d.nz * d.ny * csize(h);              // the rank's slab of the reference layout
  if ((rc = alloc_dev(h, &h->io_buf, h->io_bytes))) return rc;
  if ((rc = alloc_dev(h, (void**)&h->stat_buf, sizeof(double) * 4 * d.ny))) return rc;
  if ((rc = build_tables(h))) return rc;
  // wall-normal stage: bit 0 = segment-parallel pre kernel, bit 1 = segment-parallel post kernel.
  // Measured on B200 (profiles/): the register version of `pre` always wins; `post` is instruction-
  // and latency-bound in both forms and the register version only pays when there are too few chains
  // to fill the GPU with one thread per chain (small grids, many GPUs).
  // y stage: thread-per-chain kernels (batched row loads, HBM-bound) when there are enough chains to fill the
  // SMs, the segment-parallel register kernels when a rank holds few chains (many ranks / small grids)
  // `pre`: the register kernel below 148*256 chains.  With several GPUs the kernels store locally and a bulk scatter
  // delivers the rows (kernels_pencil.cu: launch_peer_scatter); with CHAN_TUNE_PRE_DIRECT=1 they store straight into the
  // plane owners' buffers, and then the register kernel is forced: the thread-per-chain kernel walks the rows in order, so
  // all ranks would store into the SAME plane owner at the same time (Re_tau=1000 on 8 GPUs: 3.1 ms per substep).
  // `post`: the chain kernels win at every size measured.
  // Round-2 sweep (profiles/sweep_r02.json): between 148*128 and 148*256 chains the thread-per-chain `pre` beats the register
  // kernel at every NY (16641 modes: 0.111 vs 0.151 ms at NY=97, 0.73 vs 1.08 ms at NY=633), and below that the on-chip lanes
  // kernels take over: the register kernels are picked only for in-kernel peer stores.
  if (h->tune.pencil_seg < 0) h->tune.pencil_seg = (cfg->size > 1 && h->tune.pre_direct) ? 1 : 0;
  if (h->tune.post_deep < 0) h->tune.post_deep = (2 * d.M <= 148L * 64) ? 2 : ((2 * d.M < 148L * 128) ? 1 : 0);   // 2: three warps per 32 chains
  if (h->tune.pencil_seg && seg_supported(h)) { if ((rc = seg_setup(h))) return rc; } else h->tune.pencil_seg = 0;
  // on-chip y stage (warps per mode, lanes along y; kernels_pencil_lanes.cu): bit 0 = post, bit 1 = pre.  It needs a
  // third of the DRAM traffic and has 32..160 threads per mode, but executes more instructions per mode than the
  // thread-per-chain kernels: measured on B200 it wins where a rank holds few chains (many GPUs / small grids: Re_tau=550
  // on 8 GPUs: pre 0.080 -> 0.053 ms, post 0.379 -> 0.218 ms per substep) and loses on a GPU full of chains
  // (Re_tau=550 on one GPU: post 0.80 -> 1.04 ms), profiles/r02_lanes_notes.txt.  CHAN_TUNE_LANES overrides.
  if (h->tune.lanes < 0) h->tune.lanes = (2 * d.M < 148L * 192) ? 3 : 0;     // 16.8 k chains at NY=193: step 3.54 -> 2.63 ms; 43.7 k at NY=385: slower
  if (h->tune.lanes && lanes_supported(h)) { if ((rc = lanes_setup(h))) return rc; } else h->tune.lanes = 0;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}

extern "C" void chan_destroy(chan_handle* h) {
  if (!h) return;
  if (h->stream) cudaStreamSynchronize(h->stream);
  for (int r = 0; r < 8; r++) if (r != h->cfg.rank) {
    for (int f = 0; f < 6; f++) if (h->peer_rcv[r][f]) cudaIpcCloseMemHandle(h->peer_rcv[r][f]);
    for (int f = 0; f < 3; f++) if (h->peer_H[r][f]) cudaIpcCloseMemHandle(h->peer_H[r][f]);
  }
  void* ptrs[] = { h->phi, h->g, h->v, h->dv, h->Rphi, h->Rg, h->u, h->w, h->ox, h->oy, h->oz, h->sA, h->sB, h->sC, h->sP,
                   h->Q, h->ytab_store, h->seg_pf, h->seg_pb, h->pivH, h->pivI, h->ln.tabs, h->ln.pivH, h->ln.pivI, h->tw_x, h->tw_z, h->invx, h->invz, h->posx, h->posz, h->mean_a, h->mean_b, h->mean_u, h->mean_N,
                   h->mean_uw, h->mean_tmp, h->sc, h->io_buf, h->stat_buf, h->rcv[0], h->rcv[1], h->rcv[2], h->rcv[3], h->rcv[4], h->rcv[5],
                   h->snd[0], h->snd[1], h->snd[2], h->barrier_word };
  for (size_t i = 0; i < sizeof(ptrs) / sizeof(ptrs[0]); i++) if (ptrs[i]) cudaFree(ptrs[i]);
  for (int i = 0; i < chan_handle::EV_MAX; i++) if (h->ev0[i]) { cudaEventDestroy(h->ev0[i]); cudaEventDestroy(h->ev1[i]); }
  for (int i = 0; i < 64; i++) { if (h->ev_batch[i]) cudaEventDestroy(h->ev_batch[i]); if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]); }
  if (h->ev_xs) cudaEventDestroy(h->ev_xs);
  if (h->ev_cf) cudaEventDestroy(h->ev_cf);
  for (int i = 0; i < 4; i++) { if (h->ev_cj[i]) cudaEventDestroy(h->ev_cj[i]); if (h->xc[i]) cudaStreamDestroy(h->xc[i]); }
  if (h->ev_pre) cudaEventDestroy(h->ev_pre);
  if (h->xs) cudaStreamDestroy(h->xs);
  if (h->xo) cudaStreamDestroy(h->xo);
  if (h->ev_sub) cudaEventDestroy(h->ev_sub);
  if (h->ev_piv) cudaEventDestroy(h->ev_piv);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->aux) cudaStreamDestroy(h->aux);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

// ---- peer-copy exchange set-up ----------------------------------------------------------------------
extern "C" int chan_ipc_export(chan_handle* h, void* out) {
  if (!h || !out) return CHAN_ERR_ARG;
  if (h->cfg.size < 2) return set_err(h, CHAN_ERR_STATE, "chan_ipc_export needs a multi-GPU handle");
  void* bufs[9] = { h->rcv[0], h->rcv[1], h->rcv[2], h->rcv[3], h->rcv[4], h->rcv[5], h->ox, h->oy, h->oz };
  for (int i = 0; i < 9; i++) {
    cudaIpcMemHandle_t mh;
    CHAN_CUDA(h, cudaIpcGetMemHandle(&mh, bufs[i]));
    static_assert(sizeof(mh) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy((char*)out + 64 * i, &mh, 64);
  }
  return CHAN_OK;
}
extern "C" int chan_ipc_import(chan_handle* h, int peer, const void* in) {
  if (!h || !in || peer < 0 || peer >= h->cfg.size) return CHAN_ERR_ARG;
  if (peer == h->cfg.rank) return CHAN_OK;
  for (int i = 0; i < 9; i++) {
    cudaIpcMemHandle_t mh; memcpy(&mh, (const char*)in + 64 * i, 64);
    void* p = 0;
    CHAN_CUDA(h, cudaIpcOpenMemHandle(&p, mh, cudaIpcMemLazyEnablePeerAccess));
    if (i < 6) h->peer_rcv[peer][i] = p; else h->peer_H[peer][i - 6] = p;
  }
  int have = 0;
  for (int r = 0; r < h->cfg.size; r++) have += (r == h->cfg.rank || h->peer_rcv[r][0]) ? 1 : 0;
  if (have == h->cfg.size) {
    for (int f = 0; f < 6; f++) h->peer_rcv[h->cfg.rank][f] = h->rcv[f];
    h->peer_H[h->cfg.rank][0] = h->ox; h->peer_H[h->cfg.rank][1] = h->oy; h->peer_H[h->cfg.rank][2] = h->oz;
    h->peer_ready = 1;      // the two exchanges of a substep become bulk peer copies (kernels_pencil.cu: launch_peer_scatter*)
  }
  return CHAN_OK;
}

extern "C" const char* chan_last_error(const chan_handle* h) { return h ? h->err : "null handle"; }
extern "C" long chan_launch_count(const chan_handle* h) { return h ? h->launches : 0; }
extern "C" int chan_sync(chan_handle* h) { if (!h) return CHAN_ERR_ARG; CHAN_CUDA(h, cudaStreamSynchronize(h->stream)); return CHAN_OK; }

static void* field_ptr(chan_handle* h, int which) {
  void* t[CHAN_F_COUNT] = { h->phi, h->g, h->v, h->dv, h->Rphi, h->Rg, h->u, h->w, h->ox, h->oy, h->oz };
  return (which >= 0 && which < CHAN_F_COUNT) ? t[which] : 0;
}

// The rank's active-kx slab [ikx0, ikx0+Kxl) is at most two runs of rows i of the reference array: kx >= 0 (i = kx)
// and kx < 0 (i = NX + kx).  run r covers slab rows [s0, s0+n) = reference rows [i0, i0+n).
struct IoRun { int s0, i0, n; };
static int io_runs(const chan_handle* h, IoRun* r) {
  const Dims& d = h->d;
  const int lo = d.ikx0, hi = d.ikx0 + d.Kxl;
  int nr = 0;
  const int p1 = hi < d.kxmax + 1 ? hi : d.kxmax + 1;
  if (lo < p1) { r[nr].s0 = 0; r[nr].i0 = lo; r[nr].n = p1 - lo; nr++; }
  const int n0 = lo > d.kxmax + 1 ? lo : d.kxmax + 1;
  if (n0 < hi) { r[nr].s0 = n0 - lo; r[nr].i0 = d.nx - d.Kx + n0; r[nr].n = hi - n0; nr++; }
  return nr;
}
static size_t io_row_bytes(const chan_handle* h) { return (size_t)h->d.nz * h->d.ny * csize(h); }

extern "C" int chan_set_field(chan_handle* h, int which, const void* host) {
  if (!h || !host) return CHAN_ERR_ARG;
  void* f = field_ptr(h, which); if (!f) return set_err(h, CHAN_ERR_ARG, "bad field selector %d", which);
  // a rank reads only the rows of its own kx slab out of the caller's full array
  IoRun r[2]; const int nr = io_runs(h, r); const size_t row = io_row_bytes(h);
  for (int q = 0; q < nr; q++)
    CHAN_CUDA(h, cudaMemcpyAsync((char*)h->io_buf + r[q].s0 * row, (const char*)host + r[q].i0 * row, (size_t)r[q].n * row, cudaMemcpyHostToDevice, h->stream));
  int rc = launch_pack(h, h->io_buf, f, 1); if (rc) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_get_field(chan_handle* h, int which, void* host) {
  if (!h || !host) return CHAN_ERR_ARG;
  void* f = field_ptr(h, which); if (!f) return set_err(h, CHAN_ERR_ARG, "bad field selector %d", which);
  int rc = launch_unpack(h, f, h->io_buf, 1); if (rc) return rc;
  // a rank writes only the rows of its own kx slab into the caller's full array; rank 0 also clears the rows of the
  // truncated kx (kxmax < kx < NX - kxmax), which belong to nobody's slab
  IoRun r[2]; const int nr = io_runs(h, r); const size_t row = io_row_bytes(h);
  for (int q = 0; q < nr; q++)
    CHAN_CUDA(h, cudaMemcpyAsync((char*)host + r[q].i0 * row, (const char*)h->io_buf + r[q].s0 * row, (size_t)r[q].n * row, cudaMemcpyDeviceToHost, h->stream));
  if (h->cfg.rank == 0) memset((char*)host + (size_t)(h->d.kxmax + 1) * row, 0, (size_t)(h->d.nx - h->d.Kx) * row);
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_set_fields(chan_handle* h, const void* ddv, const void* g) {
  int rc = chan_set_field(h, CHAN_F_DDV, ddv); if (rc) return rc;
  return chan_set_field(h, CHAN_F_G, g);
}
extern "C" int chan_get_fields(chan_handle* h, void* ddv, void* g) {
  int rc = chan_get_field(h, CHAN_F_DDV, ddv); if (rc) return rc;
  return chan_get_field(h, CHAN_F_G, g);
}
// slab variants: host array [Kxl][NZ][NY], rows in active-kx order ikx0 .. ikx0+Kxl-1 (chan_partition(chan_active_kx(NX), size, rank))
extern "C" int chan_set_field_slab(chan_handle* h, int which, const void* host_slab) {
  if (!h || !host_slab) return CHAN_ERR_ARG;
  void* f = field_ptr(h, which); if (!f) return set_err(h, CHAN_ERR_ARG, "bad field selector %d", which);
  CHAN_CUDA(h, cudaMemcpyAsync(h->io_buf, host_slab, h->io_bytes, cudaMemcpyHostToDevice, h->stream));
  int rc = launch_pack(h, h->io_buf, f, 1); if (rc) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_get_field_slab(chan_handle* h, int which, void* host_slab) {
  if (!h || !host_slab) return CHAN_ERR_ARG;
  void* f = field_ptr(h, which); if (!f) return set_err(h, CHAN_ERR_ARG, "bad field selector %d", which);
  int rc = launch_unpack(h, f, h->io_buf, 1); if (rc) return rc;
  CHAN_CUDA(h, cudaMemcpyAsync(host_slab, h->io_buf, h->io_bytes, cudaMemcpyDeviceToHost, h->stream));
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
// device variants: the caller's full reference-layout DEVICE arrays, no host staging (the coarse RKstep seam)
extern "C" int chan_set_field_dev(chan_handle* h, int which, const void* dev) {
  if (!h || !dev) return CHAN_ERR_ARG;
  void* f = field_ptr(h, which); if (!f) return set_err(h, CHAN_ERR_ARG, "bad field selector %d", which);
  int rc = launch_pack(h, dev, f, 0); if (rc) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_get_field_dev(chan_handle* h, int which, void* dev) {
  if (!h || !dev) return CHAN_ERR_ARG;
  void* f = field_ptr(h, which); if (!f) return set_err(h, CHAN_ERR_ARG, "bad field selector %d", which);
  if (h->cfg.size > 1) return set_err(h, CHAN_ERR_STATE, "chan_get_field_dev clears the whole array: single-GPU handles only");
  int rc = launch_unpack(h, f, dev, 0); if (rc) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_set_fields_dev(chan_handle* h, const void* ddv, const void* g) {
  int rc = chan_set_field_dev(h, CHAN_F_DDV, ddv); if (rc) return rc;
  return chan_set_field_dev(h, CHAN_F_G, g);
}
extern "C" int chan_get_fields_dev(chan_handle* h, void* ddv, void* g) {
  int rc = chan_get_field_dev(h, CHAN_F_DDV, ddv); if (rc) return rc;
  return chan_get_field_dev(h, CHAN_F_G, g);
}

// ---- checkpoint / restart (no counterpart in the reference, which restarts from its end-of-run field files and
// loses R_phi, R_g, dt: main.c:139, RK3.c:114) ------------------------------------------------------------
// Payload of one rank: header | Scalars | host_counter | mean_u, mean_N (T[NY]) | phi g v dv R_phi R_g (compact slabs)
struct CkptHeader { char magic[8]; int version, prec, nx, ny, nz, rank, size, ikx0, Kxl, Kz; long M; size_t bytes; };
static const char kCkptMagic[8] = { 'C', 'H', 'B', '2', 'C', 'K', 'P', 'T' };
static size_t ckpt_bytes(const chan_handle* h) {
  return sizeof(CkptHeader) + sizeof(Scalars) + sizeof(long) + 2 * (size_t)h->d.ny * h->prec + 6 * field_bytes(h);
}
extern "C" size_t chan_checkpoint_bytes(const chan_handle* h) { return h ? ckpt_bytes(h) : 0; }
extern "C" int chan_checkpoint_save(chan_handle* h, void* host_buf) {
  if (!h || !host_buf) return CHAN_ERR_ARG;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  char* p = (char*)host_buf;
  CkptHeader hd; memset(&hd, 0, sizeof(hd));
  memcpy(hd.magic, kCkptMagic, 8); hd.version = 1; hd.prec = h->prec; hd.nx = h->d.nx; hd.ny = h->d.ny; hd.nz = h->d.nz;
  hd.rank = h->cfg.rank; hd.size = h->cfg.size; hd.ikx0 = h->d.ikx0; hd.Kxl = h->d.Kxl; hd.Kz = h->d.Kz; hd.M = h->d.M; hd.bytes = ckpt_bytes(h);
  memcpy(p, &hd, sizeof(hd)); p += sizeof(hd);
  CHAN_CUDA(h, cudaMemcpy(p, h->sc, sizeof(Scalars), cudaMemcpyDeviceToHost)); p += sizeof(Scalars);
  memcpy(p, &h->host_counter, sizeof(long)); p += sizeof(long);
  const size_t rb = (size_t)h->d.ny * h->prec, fb = field_bytes(h);
  CHAN_CUDA(h, cudaMemcpy(p, h->mean_u, rb, cudaMemcpyDeviceToHost)); p += rb;
  CHAN_CUDA(h, cudaMemcpy(p, h->mean_N, rb, cudaMemcpyDeviceToHost)); p += rb;
  void* fl[6] = { h->phi, h->g, h->v, h->dv, h->Rphi, h->Rg };
  for (int i = 0; i < 6; i++) { CHAN_CUDA(h, cudaMemcpy(p, fl[i], fb, cudaMemcpyDeviceToHost)); p += fb; }
  return CHAN_OK;
}
extern "C" int chan_checkpoint_load(chan_handle* h, const void* host_buf) {
  if (!h || !host_buf) return CHAN_ERR_ARG;
  const char* p = (const char*)host_buf;
  CkptHeader hd; memcpy(&hd, p, sizeof(hd)); p += sizeof(hd);
  if (memcmp(hd.magic, kCkptMagic, 8) || hd.version != 1) return set_err(h, CHAN_ERR_ARG, "not a version-1 checkpoint");
  if (hd.prec != h->prec || hd.nx != h->d.nx || hd.ny != h->d.ny || hd.nz != h->d.nz || hd.rank != h->cfg.rank || hd.size != h->cfg.size ||
      hd.ikx0 != h->d.ikx0 || hd.Kxl != h->d.Kxl || hd.bytes != ckpt_bytes(h))
    return set_err(h, CHAN_ERR_ARG, "checkpoint of grid %dx%dx%d precision %d rank %d/%d does not match this handle", hd.nx, hd.ny, hd.nz, hd.prec, hd.rank, hd.size);
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  CHAN_CUDA(h, cudaMemcpy(h->sc, p, sizeof(Scalars), cudaMemcpyHostToDevice)); p += sizeof(Scalars);
  memcpy(&h->host_counter, p, sizeof(long)); p += sizeof(long);
  const size_t rb = (size_t)h->d.ny * h->prec, fb = field_bytes(h);
  CHAN_CUDA(h, cudaMemcpy(h->mean_u, p, rb, cudaMemcpyHostToDevice)); p += rb;
  CHAN_CUDA(h, cudaMemcpy(h->mean_N, p, rb, cudaMemcpyHostToDevice)); p += rb;
  void* fl[6] = { h->phi, h->g, h->v, h->dv, h->Rphi, h->Rg };
  for (int i = 0; i < 6; i++) { CHAN_CUDA(h, cudaMemcpy(fl[i], p, fb, cudaMemcpyHostToDevice)); p += fb; }
  return CHAN_OK;
}

extern "C" int chan_set_umean(chan_handle* h, const void* um) {
  if (!h || !um) return CHAN_ERR_ARG;
  const int ny = h->d.ny;
  if (h->prec == 4) { float* t = (float*)malloc(sizeof(float) * ny); for (int j = 0; j < ny; j++) t[j] = ((const float*)um)[2 * j];
    cudaError_t e = cudaMemcpy(h->mean_u, t, sizeof(float) * ny, cudaMemcpyHostToDevice); free(t); CHAN_CUDA(h, e); }
  else { double* t = (double*)malloc(sizeof(double) * ny); for (int j = 0; j < ny; j++) t[j] = ((const double*)um)[2 * j];
    cudaError_t e = cudaMemcpy(h->mean_u, t, sizeof(double) * ny, cudaMemcpyHostToDevice); free(t); CHAN_CUDA(h, e); }
  return CHAN_OK;
}
extern "C" int chan_get_umean(chan_handle* h, void* um) {
  if (!h || !um) return CHAN_ERR_ARG;
  const int ny = h->d.ny;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  if (h->prec == 4) { float* t = (float*)malloc(sizeof(float) * ny);
    cudaError_t e = cudaMemcpy(t, h->mean_u, sizeof(float) * ny, cudaMemcpyDeviceToHost);
    for (int j = 0; j < ny; j++) { ((float*)um)[2 * j] = t[j]; ((float*)um)[2 * j + 1] = 0.0f; } free(t); CHAN_CUDA(h, e); }
  else { double* t = (double*)malloc(sizeof(double) * ny);
    cudaError_t e = cudaMemcpy(t, h->mean_u, sizeof(double) * ny, cudaMemcpyDeviceToHost);
    for (int j = 0; j < ny; j++) { ((double*)um)[2 * j] = t[j]; ((double*)um)[2 * j + 1] = 0.0; } free(t); CHAN_CUDA(h, e); }
  return CHAN_OK;
}
extern "C" int chan_get_nmean(chan_handle* h, double* out) {
  if (!h || !out) return CHAN_ERR_ARG;
  const int ny = h->d.ny;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  if (h->prec == 4) { float* t = (float*)malloc(sizeof(float) * ny);
    cudaError_t e = cudaMemcpy(t, h->mean_N, sizeof(float) * ny, cudaMemcpyDeviceToHost);
    for (int j = 0; j < ny; j++) out[j] = t[j]; free(t); CHAN_CUDA(h, e); }
  else CHAN_CUDA(h, cudaMemcpy(out, h->mean_N, sizeof(double) * ny, cudaMemcpyDeviceToHost));
  return CHAN_OK;
}

extern "C" int chan_get_scalars(chan_handle* h, chan_scalars* out) {
  if (!h || !out) return CHAN_ERR_ARG;
  Scalars s;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  CHAN_CUDA(h, cudaMemcpy(&s, h->sc, sizeof(s), cudaMemcpyDeviceToHost));
  out->dt = s.dt; out->dt_next = s.dt_next; out->dt_c = s.dt_c; out->dt_v = s.dt_v;
  for (int c = 0; c < 3; c++) out->umax[c] = s.umax[c];
  out->utau[0] = s.utau[0]; out->utau[1] = s.utau[1];
  out->counter = s.counter; out->time = s.time;
  return CHAN_OK;
}
static int poke_scalar(chan_handle* h, size_t off, double v) {
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  CHAN_CUDA(h, cudaMemcpy((char*)h->sc + off, &v, sizeof(double), cudaMemcpyHostToDevice));
  return CHAN_OK;
}
extern "C" int chan_set_fixed_dt(chan_handle* h, double dt) { if (!h) return CHAN_ERR_ARG; return poke_scalar(h, offsetof(Scalars, fixed_dt), dt); }
extern "C" int chan_set_dt(chan_handle* h, double dt) {
  if (!h) return CHAN_ERR_ARG;
  double r = h->prec == 4 ? (double)(float)dt : dt;
  return poke_scalar(h, offsetof(Scalars, dt), r);
}

// ---- hot path ---------------------------------------------------------------------
extern "C" int chan_rk_init(chan_handle* h) {
  if (!h) return CHAN_ERR_ARG;
  // dealias x2 (RK3.c:120-121) is structural; imposeSymetry (RK3.c:122); calcVdV (RK3.c:123)
  int rc = launch_symmetrize(h, h->phi, h->g); if (rc) return rc;
  return launch_op(h, OP_VDV, h->phi, h->v, h->dv, 0.0, 0.0);
}

// ---- optional stage timing ----------------------------------------------------------
namespace chan {
int timing_begin(chan_handle* h, int stage) {
  if (!h->timing) return -1;
  if (h->n_ev >= chan_handle::EV_MAX) { h->n_ev_dropped++; return -1; }
  int i = h->n_ev++;
  if (!h->ev0[i]) { cudaEventCreate(&h->ev0[i]); cudaEventCreate(&h->ev1[i]); }
  h->ev_stage[i] = stage;
  cudaEventRecord(h->ev0[i], h->stream);
  return i;
}
void timing_end(chan_handle* h, int i) { if (i >= 0) cudaEventRecord(h->ev1[i], h->stream); }
}  // namespace chan
#define t_begin timing_begin
#define t_end timing_end

extern "C" int chan_set_timing(chan_handle* h, int on) { if (!h) return CHAN_ERR_ARG; h->timing = on; return CHAN_OK; }
extern "C" int chan_get_stage_ms(chan_handle* h, double* ms, long* cnt) {
  if (!h || !ms || !cnt) return CHAN_ERR_ARG;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  for (int s = 0; s < CHAN_STAGE_COUNT; s++) { ms[s] = 0; cnt[s] = 0; }
  for (int i = 0; i < h->n_ev; i++) {
    float t = 0; CHAN_CUDA(h, cudaEventElapsedTime(&t, h->ev0[i], h->ev1[i]));
    ms[h->ev_stage[i]] += t; cnt[h->ev_stage[i]]++;
  }
  const int dropped = h->n_ev_dropped;
  h->n_ev = 0; h->n_ev_dropped = 0;
  if (dropped) return set_err(h, CHAN_ERR_STATE, "%d stage instances were not timed: more than %d between two chan_get_stage_ms calls", dropped, (int)chan_handle::EV_MAX);
  return CHAN_OK;
}

extern "C" int chan_rk_substep(chan_handle* h, int n) {
  if (!h || n < 0 || n > 2) return CHAN_ERR_ARG;
  int rc, t;
  if (h->tune.lanes & 1) {      // pivot tables of this substep's implicit matrix: on the side stream, beside pre and the plane stage
    cudaStream_t main_stream = h->stream;
    CHAN_CUDA(h, cudaEventRecord(h->ev_sub, main_stream));
    CHAN_CUDA(h, cudaStreamWaitEvent(h->aux, h->ev_sub, 0));
    h->stream = h->aux;
    rc = launch_lanes_pivots(h, n);
    h->stream = main_stream;
    if (rc) return rc;
    CHAN_CUDA(h, cudaEventRecord(h->ev_piv, h->aux));
  }
  t = t_begin(h, CHAN_STAGE_PRE);
  if ((rc = (h->tune.lanes & 2) ? launch_pencil_pre_lanes(h)
          : ((h->tune.pencil_seg & 1) ? launch_pencil_pre_seg(h) : launch_pencil_pre(h)))) return rc;   // calcUW + writeUmean + d/dy + calcOmega
  t_end(h, t);
  if (h->peer_ready && !h->tune.pre_direct && (h->tune.fwd_direct || !h->tune.xchg_pipe)) {          // all rows at once, the plane stage waits for everything
    int t2 = t_begin(h, CHAN_STAGE_XCHG_OUT);
    if ((rc = launch_peer_scatter(h, -1))) return rc;
    t_end(h, t2);
  } else if (h->peer_ready && !h->tune.pre_direct) {
    // CHAN_TUNE_XCHG_PIPE=1, off by default: measured slower than the all-at-once copy on 2 and 8 GPUs (Re_tau=1000: 22.9 vs
    // 18.3 ms per step on 8 GPUs, profiles/r02_n8_notes.txt) - the copies and the transforms slow each other down and the
    // extra barriers cost more than the overlap hides.
    // pencil -> plane transpose over NVLink, one bulk copy kernel per plane batch of the destinations on a side stream, each
    // followed by a stream-ordered barrier of all ranks: the plane stage of batch b starts as soon as batch b has arrived
    // everywhere, while the rows of the later batches are still travelling
    if (h->nbatch > 64) return set_err(h, CHAN_ERR_ARG, "more than 64 plane batches per rank");
    cudaStream_t main_stream = h->stream;
    CHAN_CUDA(h, cudaEventRecord(h->ev_pre, main_stream));
    CHAN_CUDA(h, cudaStreamWaitEvent(h->xo, h->ev_pre, 0));
    h->stream = h->xo;
    for (int b = 0; b < h->nbatch && !rc; b++) {
      int t2 = t_begin(h, CHAN_STAGE_XCHG_OUT);
      rc = launch_peer_scatter(h, b);
      t_end(h, t2);
      if (!rc) rc = stream_barrier(h);
      if (!rc && cudaEventRecord(h->ev_in[b], h->xo) != cudaSuccess) rc = set_err(h, CHAN_ERR_CUDA, "event record failed");
    }
    h->stream = main_stream;
    if (rc) return rc;
  }
  t = t_begin(h, CHAN_STAGE_PLANES);
  if ((rc = launch_nonlinear_planes(h))) return rc;       // 6 C2R, u x omega, 3 R2C, truncation
  t_end(h, t);
  // The mean-flow / dt kernel is one block of serial work and shares nothing with pencil_post (it reads
  // the mean row of H1, the wall shear and the maxima; post reads sc->dt, which only end_step changes):
  // it runs on the side stream while post fills the GPU, and the main stream joins it afterwards.
  CHAN_CUDA(h, cudaEventRecord(h->ev_fork, h->stream));
  CHAN_CUDA(h, cudaStreamWaitEvent(h->aux, h->ev_fork, 0));
  {
    cudaStream_t main_stream = h->stream;
    h->stream = h->aux;
    t = t_begin(h, CHAN_STAGE_SCALAR);
    rc = launch_scalar_stage(h, n);                      // mean flow, calcDt
    t_end(h, t);
    h->stream = main_stream;
    if (rc) return rc;
  }
  CHAN_CUDA(h, cudaEventRecord(h->ev_join, h->aux));
  if (h->tune.lanes & 1) CHAN_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_piv, 0));
  t = t_begin(h, CHAN_STAGE_POST);
  rc = (h->tune.lanes & 1) ? launch_pencil_post_lanes(h, n)
     : ((h->tune.pencil_seg & 2) ? launch_pencil_post_seg(h, n) : launch_pencil_post(h, n));   // RK explicit + h_v,h_g + implicit + bi-Laplacian
  t_end(h, t);
  CHAN_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));
  return rc;
}

extern "C" int chan_rk_end_step(chan_handle* h) {
  if (!h) return CHAN_ERR_ARG;
  int rc;
  if (h->host_counter % 1000 == 0) { if ((rc = launch_symmetrize(h, h->phi, h->g))) return rc; }   // RK3.c:174-176
  h->host_counter++;
  return launch_end_step(h);
}

extern "C" int chan_rk_step(chan_handle* h, int nsteps) {
  if (!h || nsteps < 0) return CHAN_ERR_ARG;
  for (int s = 0; s < nsteps; s++) {
    int t = t_begin(h, CHAN_STAGE_STEP);
    for (int n = 0; n < 3; n++) { int rc = chan_rk_substep(h, n); if (rc) return rc; }
    int rc = chan_rk_end_step(h); if (rc) return rc;
    t_end(h, t);
  }
  return CHAN_OK;
}

extern "C" int chan_statistics(chan_handle* h, double* out) {
  if (!h || !out) return CHAN_ERR_ARG;
  double* d = h->stat_buf; const size_t b = sizeof(double) * 4 * h->d.ny;
  int rc = launch_stats(h, d);
  if (!rc) rc = allreduce_sum_host(h, d, 4 * h->d.ny);
  if (rc) return rc;
  CHAN_CUDA(h, cudaMemcpyAsync(out, d, b, cudaMemcpyDeviceToHost, h->stream));
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}

// ---- operator level: device pointers, reference layout, in place ----------------------
// work fields used as compact staging: u, w, ox, oy, oz (never live across API calls)
static int op_inplace(chan_handle* h, int op, void* a_dev, void* b_dev, void* c_dev, int nin, int nout, double betha, double dt) {
  void* st[3] = { h->ox, h->oy, h->oz };
  void* ext[3] = { a_dev, b_dev, c_dev };
  int rc;
  for (int i = 0; i < nin; i++) if ((rc = launch_pack(h, ext[i], st[i]))) return rc;
  if ((rc = launch_op(h, op, st[0], st[1], st[2], betha, dt))) return rc;
  for (int i = 0; i < nout; i++) if ((rc = launch_unpack(h, st[i], ext[i]))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}

extern "C" int chan_dealias(chan_handle* h, void* u) {
  if (!h || !u) return CHAN_ERR_ARG;
  int rc = launch_pack(h, u, h->ox); if (rc) return rc;
  if ((rc = launch_unpack(h, h->ox, u))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_deriv_y(chan_handle* h, void* u) { if (!h || !u) return CHAN_ERR_ARG; return op_inplace(h, OP_DERIV_Y, u, 0, 0, 1, 1, 0, 0); }
extern "C" int chan_deriv_yy(chan_handle* h, void* u) { if (!h || !u) return CHAN_ERR_ARG; return op_inplace(h, OP_DERIV_YY, u, 0, 0, 1, 1, 0, 0); }
extern "C" int chan_helmholtz(chan_handle* h, void* u) { if (!h || !u) return CHAN_ERR_ARG; return op_inplace(h, OP_HELM, u, 0, 0, 1, 1, 0, 0); }
extern "C" int chan_implicit(chan_handle* h, void* u, double betha, double dt) { if (!h || !u) return CHAN_ERR_ARG; return op_inplace(h, OP_IMPLICIT, u, 0, 0, 1, 1, betha, dt); }
extern "C" int chan_bilaplacian(chan_handle* h, void* ddv, void* v, void* dv, double betha, double dt) {
  if (!h || !ddv || !v || !dv) return CHAN_ERR_ARG;
  void* ext[3] = { ddv, v, dv }; void* st[3] = { h->ox, h->oy, h->oz };
  int rc = launch_pack(h, ddv, st[0]); if (rc) return rc;
  if ((rc = launch_op(h, OP_BILAP, st[0], st[1], st[2], betha, dt))) return rc;
  for (int i = 0; i < 3; i++) if ((rc = launch_unpack(h, st[i], ext[i]))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_calc_uw(chan_handle* h, void* ux, void* uz, const void* f, const void* g) {
  if (!h || !ux || !uz || !f || !g) return CHAN_ERR_ARG;
  int rc = launch_pack(h, f, h->ox); if (rc) return rc;
  if ((rc = launch_pack(h, g, h->oy))) return rc;
  if ((rc = launch_op(h, OP_UW, h->ox, h->oy, 0, 0, 0))) return rc;       // -> h->u, h->w
  if ((rc = launch_unpack(h, h->u, ux))) return rc;
  if ((rc = launch_unpack(h, h->w, uz))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_convolution(chan_handle* h, void* ux, void* uy, void* uz, void* wx, void* wy, void* wz) {
  if (!h || !ux || !uy || !uz || !wx || !wy || !wz) return CHAN_ERR_ARG;
  // uses the state slot v as staging for uy: save and restore it through the scratch field sC
  int rc;
  const size_t fb = field_bytes(h);
  CHAN_CUDA(h, cudaMemcpyAsync(h->sC, h->v, fb, cudaMemcpyDeviceToDevice, h->stream));
  if ((rc = launch_pack(h, ux, h->u))) return rc;
  if ((rc = launch_pack(h, uy, h->v))) return rc;
  if ((rc = launch_pack(h, uz, h->w))) return rc;
  if ((rc = launch_op(h, OP_OMEGA, 0, 0, 0, 0, 0))) return rc;            // d/dy u, d/dy w, omega
  if ((rc = launch_nonlinear_planes(h, true))) return rc;
  CHAN_CUDA(h, cudaMemcpyAsync(h->v, h->sC, fb, cudaMemcpyDeviceToDevice, h->stream));
  if ((rc = launch_unpack(h, h->ox, wx))) return rc;
  if ((rc = launch_unpack(h, h->oy, wy))) return rc;
  if ((rc = launch_unpack(h, h->oz, wz))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_calc_hvg(chan_handle* h, void* nlx, void* nly, void* nlz) {
  if (!h || !nlx || !nly || !nlz) return CHAN_ERR_ARG;
  void* ext[3] = { nlx, nly, nlz }; void* st[3] = { h->ox, h->oy, h->oz };
  int rc;
  for (int i = 0; i < 3; i++) if ((rc = launch_pack(h, ext[i], st[i]))) return rc;
  if ((rc = launch_op(h, OP_HVG, st[0], st[1], st[2], 0, 0))) return rc;
  if ((rc = launch_unpack(h, st[0], nlx))) return rc;
  if ((rc = launch_unpack(h, st[2], nlz))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_impose_symmetry(chan_handle* h, void* u, void* v) {
  if (!h || !u || !v) return CHAN_ERR_ARG;
  int rc = launch_pack(h, u, h->ox); if (rc) return rc;
  if ((rc = launch_pack(h, v, h->oy))) return rc;
  if ((rc = launch_symmetrize(h, h->ox, h->oy))) return rc;
  if ((rc = launch_unpack(h, h->ox, u))) return rc;
  if ((rc = launch_unpack(h, h->oy, v))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
extern "C" int chan_fft_backward_transpose(chan_handle* h, const void* u, void* phys) {
  if (!h || !u || !phys) return CHAN_ERR_ARG;
  int rc = launch_pack(h, u, h->ox); if (rc) return rc;
  if ((rc = launch_to_physical(h, h->ox, phys))) return rc;
  CHAN_CUDA(h, cudaStreamSynchronize(h->stream));
  return CHAN_OK;
}
