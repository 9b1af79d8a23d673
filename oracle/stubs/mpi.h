/* Single-rank stand-in for the 4 MPI entry points the reference's hot path
 * calls (channel_cuda_mpi.c:78,111  fft.c:260).  TEST INFRASTRUCTURE ONLY:
 * lets oracle/build_ref.sh link the reference's own sources on a box that has
 * no MPI.  Not part of the product. */
#ifndef ORACLE_STUB_MPI_H
#define ORACLE_STUB_MPI_H
#include <string.h>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
/* one MPI_DOUBLE carries one complex of the state type (float2 as shipped;
 * double2 in the retyped ref64 build), one MPI_FLOAT one real (ref_shim.h) */
#define MPI_DOUBLE (2*REF_REAL_BYTES)
#define MPI_FLOAT REF_REAL_BYTES
#define MPI_INT 4
#define MPI_CHAR 1
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_STATUS_IGNORE 0
static inline int MPI_Alltoall(const void* s, int n, MPI_Datatype t, void* r, int, MPI_Datatype, MPI_Comm) {
  memcpy(r, s, (size_t)n * (size_t)t); return 0; }
static inline int MPI_Allreduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op, MPI_Comm) {
  memcpy(r, s, (size_t)n * (size_t)t); return 0; }
#endif
