/* Host-side run driver above the C ABI: what main.c / io.c / statistics.cu / meanUevol.c do around
 * the hot path (file formats and cadences), in plain C. */
#ifndef CHAN_HOST_DRIVER_H
#define CHAN_HOST_DRIVER_H
#include <stddef.h>
#include "conf.h"
#include "../include/channel_b200.h"

/* genRandData (io.c:74-98): (rand()%NM)/NM is an integer division, so every entry is 0.5*F*(0-1) */
void gen_rand_data(void* ddv, void* g, double F, size_t n_complex, int precision);
/* field payload = the [NX][NZ][NY] complex array verbatim, the bytes the reference stores in its HDF5
 * dataset "u" (hit_mpi.c:272-328); here as a raw file (HDF5 is absent on the box) */
int read_field_raw(const char* name, void* dst, size_t bytes);
int write_field_raw(const char* name, const void* src, size_t bytes);
/* the same payload in the reference's container: HDF5 dataset "u", {NX, NY, 2NZ} of float / double (h5min.h).
 * read_field / write_field pick the container by the file name: "*.h5" = HDF5, anything else = raw */
int read_field(const char* name, void* dst, int nx, int ny, int nz, int precision);
int write_field(const char* name, const void* src, int nx, int ny, int nz, int precision);
/* checkpoint.c: one rank's complete state (chan_checkpoint_save / _load) to / from a file */
int write_checkpoint_file(chan_handle* h, const char* name);
int read_checkpoint_file(chan_handle* h, const char* name);
/* readU / writeU (meanUevol.c:153-176): NY complex values, raw */
int read_umean(const char* name, void* dst, int ny, int precision);
int write_umean(const char* name, const void* src, int ny, int precision);
/* calcSt files (statistics.cu:173-232): append one line to RSTRSS/URMS/VRMS/WRMS.dat */
int append_calcst_files(const char* path, const double* sums_4ny, int ny, int precision);
/* per-step files of meanURKstep_2 (meanUevol.c:489-557) + the stdout block */
int append_mean_files(const char* path, const void* umean, const double* nmean, const chan_scalars* sc,
                      int nx, int ny, int nz, double reynolds, double qvelocity, int precision, int verbose);
/* RKstep (RK3.c:111-193) + the main.c:107-147 frame around it; returns 0 on success */
int run_channel(const domain_t* d, const paths_t* p, const run_opts_t* o, int verbose);
#endif
