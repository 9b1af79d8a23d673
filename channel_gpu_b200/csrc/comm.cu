// Multi-GPU exchange: the y-pencil <-> xz-plane global transpose of the reference
// (transposeXYZ2YZX / transposeYZX2XYZ, channel_cuda_mpi.c:64-128: five cuBLAS geam passes, a
// full-field D2H, MPI_Alltoall through pageable host memory, a full-field H2D) as ONE grouped
// NCCL send/recv per field over NVLink, device to device.
//
// Thanks to the plane-major layout there is nothing to pack:
//   pencil side   F[j][ikx_local][kz]   rows j0_p .. j0_p+nyl_p-1 are what peer p needs: contiguous;
//   plane side    per-source blocks [jl][ikx_local(src)][kz], addressed in the FFT kernels through
//                 SegMap (plane_core.cuh) — as received / ready to send, no reorder pass.
// Only active modes cross the wire: 9 * (P/S) * (S-1)/S bytes per GPU per substep each way.
// Slabs may be uneven (all BASELINE Kx and NY are odd).
#include <nccl.h>
#include "chan_internal.h"

namespace chan {

#define CHAN_NCCL(h, call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
  return set_err(h, CHAN_ERR_CUDA, "%s: %s (%s:%d)", #call, ncclGetErrorString(r_), __FILE__, __LINE__); } while (0)

static ncclDataType_t real_type(const chan_handle* h) { return h->prec == 4 ? ncclFloat : ncclDouble; }

// pencil fields -> h->rcv[f]: peer p receives my rows [j0_p, j0_p + nyl_p) of every field
static int exchange(chan_handle* h, int nf, void* const* pencil, void* const* plane, bool to_planes) {
  ncclComm_t comm = (ncclComm_t)h->nccl_comm;
  const int S = h->cfg.size, me = h->cfg.rank;
  const size_t cb = 2 * (size_t)h->prec;
  // one group for all fields: NCCL fuses the nf*(S-1) send/recv pairs into a single launch
  CHAN_NCCL(h, ncclGroupStart());
  for (int f = 0; f < nf; f++) {
    for (int p = 0; p < S; p++) {
      long L[8];
      chan_exchange_layout(h->d.nx, h->d.ny, h->d.nz, S, me, p, L);   // the layout the CPU tests check
      char* pen = (char*)pencil[f] + (size_t)L[0] * cb; char* pla = (char*)plane[f] + (size_t)L[2] * cb;
      const char* sbuf = to_planes ? pen : pla; char* rbuf = to_planes ? pla : pen;
      size_t scount = (size_t)(to_planes ? L[1] : L[3]) * 2, rcount = (size_t)(to_planes ? L[3] : L[1]) * 2;
      if (p == me) { CHAN_CUDA(h, cudaMemcpyAsync(rbuf, sbuf, scount * h->prec, cudaMemcpyDeviceToDevice, h->stream)); continue; }
      CHAN_NCCL(h, ncclSend(sbuf, scount, real_type(h), p, comm, h->stream));
      CHAN_NCCL(h, ncclRecv(rbuf, rcount, real_type(h), p, comm, h->stream));
    }
  }
  CHAN_NCCL(h, ncclGroupEnd());
  return CHAN_OK;
}
// pencil fields -> h->rcv[f]: peer p receives my rows [j0_p, j0_p + nyl_p) of every field
int exchange_to_planes(chan_handle* h, int nf, void* const* fields) { return exchange(h, nf, fields, h->rcv, true); }
// h->snd[f] -> pencil fields: the mirror image
int exchange_to_pencils(chan_handle* h, int nf, void* const* fields) { return exchange(h, nf, fields, h->snd, false); }

// velocity maxima of calcUmax: replaces the three MPI_Allreduce(MAX) of reduceMAX (hit_mpi.c:427-447).
// The slots hold non-negative doubles as ordered bit patterns, so an unsigned max is a float max.
int allreduce_umax(chan_handle* h) {
  CHAN_NCCL(h, ncclAllReduce(h->sc->umax_bits, h->sc->umax_bits, 3, ncclUint64, ncclMax, (ncclComm_t)h->nccl_comm, h->stream));
  return CHAN_OK;
}

// Stream-ordered barrier across the ranks: a one-word all-reduce.  It completes on a rank only after
// every rank has reached it in its own stream, i.e. after everything issued before it there (the
// kernels that store into this rank's buffers over NVLink) has finished.
int stream_barrier(chan_handle* h) {
  CHAN_NCCL(h, ncclAllReduce(h->barrier_word, h->barrier_word, 1, ncclInt, ncclMax, (ncclComm_t)h->nccl_comm, h->stream));
  return CHAN_OK;
}

// statistics profiles: replaces MPI_Allreduce(SUM) of fft.c:260
int allreduce_sum_host(chan_handle* h, double* dev_buf, int n) {
  if (h->cfg.size > 1) CHAN_NCCL(h, ncclAllReduce(dev_buf, dev_buf, n, ncclDouble, ncclSum, (ncclComm_t)h->nccl_comm, h->stream));
  return CHAN_OK;
}

}  // namespace chan
