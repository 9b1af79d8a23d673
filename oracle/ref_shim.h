/* Force-included (-include) in front of every reference translation unit by
 * oracle/build_ref.sh.  TEST INFRASTRUCTURE ONLY — never linked into the
 * product library.
 *
 * 1. The reference calls cusparse{Z}gtsvStridedBatch
 *    (derivatives_nu_double.cu:406, hemholzt_nu_double.cu:242,
 *    implicitStep_nu_double.cu:240).  That legacy entry point was removed from
 *    cuSPARSE 11+; it is declared here and implemented in ref_harness.cu on top
 *    of cusparseZgtsv2StridedBatch, the closest surviving routine.
 * 2. `real` / `real2` name the state type of the build: float for the
 *    reference as shipped, double for the mechanically retyped "ref64" twin.
 * 3. Run-time knobs the patched RK3.c reads (step count, injected dt).
 */
#ifndef ORACLE_REF_SHIM_H
#define ORACLE_REF_SHIM_H
#include <time.h>
#include <stdio.h>
#include <cuda_runtime.h>
#include <cusparse.h>

#ifdef REF_DOUBLE
typedef double  real;
typedef double2 real2;
#else
typedef float  real;
typedef float2 real2;
#endif
#define REF_REAL_BYTES ((int)sizeof(real))

#ifndef REF_REYNOLDS
#define REF_REYNOLDS 3250.0
#endif

extern int    REF_NSTEPS;    /* replaces the literal 30000 of RK3.c:124          */
extern double REF_FIXED_DT;  /* >0: overrides calcDt's result (RK3.c:143-145)    */
extern double REF_LAST_DT;   /* dt_2 as computed by the reference's calcDt       */

cusparseStatus_t cusparseZgtsvStridedBatch(cusparseHandle_t handle, int m,
                                           const cuDoubleComplex* dl,
                                           const cuDoubleComplex* d,
                                           const cuDoubleComplex* du,
                                           cuDoubleComplex* x, int batchCount,
                                           int batchStride);
#endif
