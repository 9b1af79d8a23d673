#!/usr/bin/env bash
# Build the reference's own hot-path sources, where they lie under
# /root/reference/src, into oracle/_ref/ (git-ignored; travels to the GPU box).
# TEST INFRASTRUCTURE ONLY: the binary is the parity oracle and the timed
# baseline ("--impl reference"); nothing here is linked into the product.
#
#   oracle/build_ref.sh NX NY NZ [32|64] [REYNOLDS]
#   SHIM=1 oracle/build_ref.sh NX NY NZ [32|64]      the reference's DRIVER (RK3.c, RK3_kernels.cu, nonLinear.c,
#        meanUevol.c, statistics.cu, fft.c, ...) linked against host/channel_shim.cu + libchannel_b200.so instead of its
#        own operator files -> oracle/_ref/refshim_<grid>_f<prec>: the drop-in boundary as a program
#
# Grid sizes are compile-time macros in the reference (channel.h:42-44,
# Makefile:8-11), hence one binary per grid.  No reference source is copied
# into the repository: patched text lives only in a mktemp directory that is
# removed on exit.
#
# Patches (each is the minimum needed to run the reference's
# algorithm on this toolchain and on grids with NX != NY):
#  P1 channel_cuda_mpi.c:71,72,123,124  macro NY -> parameter Ny (only right when NX==NY)
#  P2 cusparse{Z}gtsvStridedBatch       legacy symbol gone -> shim (ref_shim.h / ref_harness.cu)
#  P3 -arch sm_100a
#  P5 channel_cuda_mpi.c:37,54          null beta / null B passed to cublasCgeam -> real zero, valid B
#  P6 RK3.c:124                         literal 30000 steps -> REF_NSTEPS; optional injected dt; per-step hook
#  P7 channel.h:61                      REYNOLDS -> -DREF_REYNOLDS (default 3250.0, the shipped value)
#  P8 14 pointwise launch sites         floor-divided grids -> ceil (kernels are already bounds-guarded);
#                                       without it matrix rows of the last k stay uninitialised when NZ*NY%16 != 0
#  P9 setUp.c:8                         drop setDerivatives_HO (dead fp32 twins, 6 field-sized allocations)
#  64: "ref64" = the same text with the state type widened (float->double,
#      cuFFT R2C/C2R -> D2Z/Z2D, Isamax/Cgeam -> Idamax/Zgeam); literals keep
#      their f suffix, i.e. constants stay float-valued.
set -euo pipefail
NXv=${1:?NX}; NYv=${2:?NY}; NZv=${3:?NZ}; PREC=${4:-32}; RE=${5:-3250.0}
HERE=$(cd "$(dirname "$0")" && pwd)
REF=${REF_SRC:-/root/reference/src}
OUT=$HERE/_ref
[ -d "$REF" ] || { echo "build_ref: $REF absent (expected on the GPU box): keeping prebuilt files"; exit 0; }
mkdir -p "$OUT"
TMP=$(mktemp -d /tmp/refbuild.XXXXXX); trap 'rm -rf "$TMP"' EXIT

FILES="RK3.c RK3_kernels.cu nonLinear.c nonLinear_kernels.cu convolution.c convolution_kernels.cu
dealias.cu fft.c channel_cuda_mpi.c derivatives_nu_double.cu hemholzt_nu_double.cu
implicitStep_nu_double.cu bilplacSolver_double.cu meanUevol.c statistics.cu imposeSymetry.c
setUp.c check.cu"
SHIM=${SHIM:-0}
# the hybrid keeps the reference's driver and plumbing; its nine operator files are replaced by host/channel_shim.cu
[ "$SHIM" = 1 ] && FILES="RK3.c RK3_kernels.cu nonLinear.c fft.c channel_cuda_mpi.c meanUevol.c statistics.cu setUp.c check.cu"

patch_common() {
  sed -E \
    -e 's#=NZ\*NY/(threadsPerBlock(_B)?\.y);#=(NZ*NY+\1-1)/\1;#' \
    -e 's#=NXSIZE/(threadsPerBlock\.x);#=(NXSIZE+\1-1)/\1;#' \
    -e 's#=NXSIZE/NSTEPS/(threadsPerBlock_B\.x);#=(NXSIZE/NSTEPS+\1-1)/\1;#'
}
retype() {
  if [ "$PREC" = 64 ]; then
    sed -E -e 's/float2/double2/g' -e 's/\bfloat\b/double/g' \
      -e 's/cufftExecR2C/cufftExecD2Z/g' -e 's/cufftExecC2R/cufftExecZ2D/g' \
      -e 's/CUFFT_R2C/CUFFT_D2Z/g' -e 's/CUFFT_C2R/CUFFT_Z2D/g' \
      -e 's/cublasIsamax/cublasIdamax/g' -e 's/cublasCgeam/cublasZgeam/g'
  else cat; fi
}

sed -E -e 's#^const float REYNOLDS=3250.0;#const float REYNOLDS=REF_REYNOLDS;#' "$REF/channel.h" | retype > "$TMP/channel.h"
grep -q REF_REYNOLDS "$TMP/channel.h" || { echo "P7 did not apply"; exit 1; }

for f in $FILES; do
  case $f in
    channel_cuda_mpi.c)
      sed -E \
        -e 's#u1,Nz,NY,myNx,domain\)#u1,Nz,Ny,myNx,domain)#' \
        -e 's#AUX,NY,myNx\*Nz,domain\)#AUX,Ny,myNx*Nz,domain)#' \
        -e 's#u1,myNx\*Nz,NY,domain\)#u1,myNx*Nz,Ny,domain)#' \
        -e 's#AUX,NY,Nz,myNx,domain\)#AUX,Ny,Nz,myNx,domain)#' \
        -e 's#u_1,Nx,0,0,Nx,#u_1,Nx,\&REF_ZERO,(const float2*)u_1,Nx,#' \
        -e 's#(const float2\*\)u_1\+stride,Nx,)0,0,Nx,#\1\&REF_ZERO,(const float2*)u_1+stride,Nx,#' \
        -e 's#^static float2 alpha\[1\];#static float2 alpha[1]; static float2 REF_ZERO;#' \
        "$REF/$f" > "$TMP/$f.p"
      [ "$(grep -c 'REF_ZERO' "$TMP/$f.p")" = 3 ] || { echo "P5 did not apply"; exit 1; }
      [ "$(grep -c '\bNY\b' "$TMP/$f.p")" = "$(( $(grep -c '\bNY\b' "$REF/$f") - 4 ))" ] || { echo "P1 did not apply"; exit 1; }
      ;;
    RK3.c)
      sed -E \
        -e 's#int Nsteps=30000;#int Nsteps=REF_NSTEPS;#' \
        -e 's#^( *)while\(counter<Nsteps\)\{#\1void ref_step_hook(int); while(counter<Nsteps){ ref_step_hook(counter);#' \
        -e 's#(dt_2=calcDt\(u,v,w,ddv,g,domain\);)#\1 REF_LAST_DT=dt_2; if(REF_FIXED_DT>0) dt_2=REF_FIXED_DT;#' \
        "$REF/$f" > "$TMP/$f.p"
      [ "$(grep -c 'REF_' "$TMP/$f.p")" -ge 2 ] || { echo "P6 did not apply"; exit 1; }
      grep -q ref_step_hook "$TMP/$f.p" || { echo "P6 hook did not apply"; exit 1; }
      ;;
    setUp.c)
      sed -E -e 's#^\s*setDerivatives_HO\(domain\);##' "$REF/$f" > "$TMP/$f.p"
      ! grep -q 'setDerivatives_HO' "$TMP/$f.p" || { echo "P9 did not apply"; exit 1; }
      ;;
    *) cp "$REF/$f" "$TMP/$f.p" ;;
  esac
  patch_common < "$TMP/$f.p" | retype > "$TMP/${f%.*}.cu"
  rm "$TMP/$f.p"
done
# P8 must have hit the 14 live launch sites at least (the hybrid keeps 3 of them)
n8=$(cat "$TMP"/*.cu | grep -c -- '-1)/threadsPerBlock' || true)
[ "$n8" -ge $([ "$SHIM" = 1 ] && echo 6 || echo 28) ] || { echo "P8 applied to $n8 lines only"; exit 1; }

DEFS="-DNX=$NXv -DNY=$NYv -DNZ=$NZv -DREF_REYNOLDS=$RE"
[ "$PREC" = 64 ] && DEFS="$DEFS -DREF_DOUBLE"
NVCC=${NVCC:-nvcc}
OBJS=""
for f in $FILES; do
  b=${f%.*}
  $NVCC -O3 -w -gencode arch=compute_100a,code=sm_100a $DEFS -I"$HERE/stubs" -I"$TMP" \
        -include "$HERE/ref_shim.h" -c "$TMP/$b.cu" -o "$TMP/$b.o" &
  OBJS="$OBJS $TMP/$b.o"
done
wait
$NVCC -O3 -w -gencode arch=compute_100a,code=sm_100a $DEFS -I"$HERE/stubs" -I"$TMP" \
      -include "$HERE/ref_shim.h" -c "$HERE/ref_harness.cu" -o "$TMP/ref_harness.o"
if [ "$SHIM" = 1 ]; then
  LIBDIR=$HERE/../channel_gpu_b200/lib
  [ -f "$LIBDIR/libchannel_b200.so" ] || { echo "build libchannel_b200.so first (python -m channel_gpu_b200.build)"; exit 1; }
  retype < "$HERE/../host/channel_shim.cu" > "$TMP/channel_shim.cu"
  $NVCC -O3 -w -gencode arch=compute_100a,code=sm_100a $DEFS -I"$HERE/stubs" -I"$TMP" -I"$HERE/../include" \
        -include "$HERE/ref_shim.h" -c "$TMP/channel_shim.cu" -o "$TMP/channel_shim.o"
  BIN="$OUT/refshim_${NXv}x${NYv}x${NZv}_f${PREC}"
  # the binary finds the library relative to itself: the repository travels to the GPU box under another path
  $NVCC -gencode arch=compute_100a,code=sm_100a -o "$BIN" $OBJS "$TMP/ref_harness.o" "$TMP/channel_shim.o" \
        -L"$LIBDIR" -lchannel_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../../channel_gpu_b200/lib' -lcufft -lcusparse -lcublas
else
  BIN="$OUT/ref_${NXv}x${NYv}x${NZv}_f${PREC}"
  $NVCC -gencode arch=compute_100a,code=sm_100a -o "$BIN" $OBJS "$TMP/ref_harness.o" -lcufft -lcusparse -lcublas
fi
echo "built $BIN"
