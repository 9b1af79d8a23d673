#include "driver.h"
#include "h5min.h"
#include <stdint.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

void gen_rand_data(void* ddv, void* g, double F, size_t n, int precision) {
  const double v = 0.5 * F * (0.0 - 1.0);
  for (size_t i = 0; i < 2 * n; i++) {
    if (precision == 4) { ((float*)ddv)[i] = (float)v; ((float*)g)[i] = (float)v; }
    else { ((double*)ddv)[i] = v; ((double*)g)[i] = v; }
  }
}

int read_field_raw(const char* name, void* dst, size_t bytes) {
  FILE* f = fopen(name, "rb");
  if (!f) { fprintf(stderr, "\nerror lectura: %s", name); return -1; }
  size_t got = fread(dst, 1, bytes, f); fclose(f);
  if (got != bytes) { fprintf(stderr, "\nreading error: %s,%zu", name, got); return -1; }
  return 0;
}
int write_field_raw(const char* name, const void* src, size_t bytes) {
  FILE* f = fopen(name, "wb");
  if (!f) { fprintf(stderr, "\nerror escritura: %s", name); return -1; }
  size_t put = fwrite(src, 1, bytes, f); fclose(f);
  if (put != bytes) { fprintf(stderr, "\nwriting error: %s", name); return -1; }
  return 0;
}
static int is_h5(const char* name) { size_t n = strlen(name); return n > 3 && !strcmp(name + n - 3, ".h5"); }
int read_field(const char* name, void* dst, int nx, int ny, int nz, int precision) {
  if (!is_h5(name)) return read_field_raw(name, dst, (size_t)nx * ny * nz * 2 * precision);
  const uint64_t dims[3] = { (uint64_t)nx, (uint64_t)ny, (uint64_t)(2 * nz) };      /* hit_mpi.c:272-279, io.c:48-54 */
  int rc = h5min_read(name, dst, precision, dims);
  if (rc) fprintf(stderr, "\nerror lectura: %s (%s)", name, rc == -2 ? "dataset u has another type or shape" : "not a readable HDF5 file");
  return rc;
}
int write_field(const char* name, const void* src, int nx, int ny, int nz, int precision) {
  if (!is_h5(name)) return write_field_raw(name, src, (size_t)nx * ny * nz * 2 * precision);
  const uint64_t dims[3] = { (uint64_t)nx, (uint64_t)ny, (uint64_t)(2 * nz) };
  int rc = h5min_write(name, src, precision, dims);
  if (rc) fprintf(stderr, "\nerror escritura: %s", name);
  return rc;
}
int read_umean(const char* name, void* dst, int ny, int precision) { return read_field_raw(name, dst, (size_t)ny * 2 * precision); }
int write_umean(const char* name, const void* src, int ny, int precision) { return write_field_raw(name, src, (size_t)ny * 2 * precision); }

static FILE* open_append(const char* path, const char* leaf) {
  char full[256];
  snprintf(full, sizeof(full), "%s%s", path, leaf);      /* strcpy + strcat of statistics.cu:176-190 */
  return fopen(full, "a");
}

int append_calcst_files(const char* path, const double* s, int ny, int precision) {
  const char* names[4] = { "RSTRSS.dat", "URMS.dat", "VRMS.dat", "WRMS.dat" };
  for (int q = 0; q < 4; q++) {
    FILE* f = open_append(path, names[q]);
    if (!f) return -1;
    for (int j = 0; j < ny; j++) {
      double v = precision == 4 ? (double)(float)s[q * ny + j] : s[q * ny + j];     /* float sum[NY] (statistics.cu:159) */
      fprintf(f, " %f", sqrt(v));                                                  /* sqrt of a negative stress prints -nan (Q14) */
    }
    fprintf(f, " \n");
    fclose(f);
  }
  return 0;
}

static double fmesh(double x) { return tanh(2.0 * x) / tanh(2.0); }   /* channel.h:21 */

int append_mean_files(const char* path, const void* umean, const double* nmean, const chan_scalars* sc,
                      int nx, int ny, int nz, double reynolds, double qvelocity, int precision, int verbose) {
  /* float arithmetic as in meanUevol.c:489-509 (double for the double-state build) */
#define R(x) (precision == 4 ? (double)(float)(x) : (double)(x))
  const double LY = 2.0, PI2 = R(2.0f * 3.14159265f), LX = R(1.0 * PI2), LZ = R(0.5 * PI2);
  const double DELTA_Y = R(2.0f / (float)(ny - 1));
  const double Umean = R(R(qvelocity) / LY), nu = R(1.0 / R(reynolds)), N2 = (double)nx * (2 * nz - 2);
  const double ut1 = sc->utau[0], ut2 = sc->utau[1];
  const double u_tau = R(sqrt(R(0.5 * R(R(ut1 * ut1) + R(ut2 * ut2)))));
  const double uc = precision == 4 ? (double)((const float*)umean)[2 * (ny / 2)] : ((const double*)umean)[2 * (ny / 2)];
  const double dy_max = fmesh(R((ny / 2 + 1) * DELTA_Y - 1.0)) - fmesh(R((ny / 2) * DELTA_Y - 1.0));
  const double dy_min = fmesh(R(1 * DELTA_Y - 1.0)) - fmesh(R(0 * DELTA_Y - 1.0));
  if (verbose) {
    printf("\n(tau_1,tau_2)=(%e,%e)", ut1, ut2);
    printf("\n****MEAN_PROFILE_STATISTICS****");
    printf("\n(RE_t,RE_c,RE_m)=(%e,%e,%e)", R(u_tau * LY * 0.5 / nu), R(uc * 0.5 * LY / (nu * N2)), R(1.0 / nu));
    printf("\n(Dx+,Dz+)=(%e,%e)", R(1.5 * u_tau * LX / (nu * nx)), R(1.5 * u_tau * LZ / (nu * (2 * nz - 2))));
    printf("\nDy+(max,min)=(%f,%f)", R(u_tau * dy_max / nu), R(u_tau * dy_min / nu));
    printf("\nC_f=%e", R(2.0 * u_tau / (Umean * Umean)));
    printf("\n(Um+,Ux+,Um/Uc)=(%f,%f,%f)\n", R(Umean / u_tau), R(uc / (u_tau * N2)), R(Umean * N2 / uc));
  }
  FILE* f;
  if (!(f = open_append(path, "MEANPROFILE.dat"))) return -1;
  for (int j = 0; j < ny; j++) fprintf(f, " %f", precision == 4 ? (double)((const float*)umean)[2 * j] : ((const double*)umean)[2 * j]);
  fprintf(f, " \n"); fclose(f);
  if (!(f = open_append(path, "MEANREAYNOLDS.dat"))) return -1;
  for (int j = 0; j < ny; j++) fprintf(f, " %f", nmean[j]);
  fprintf(f, " \n"); fclose(f);
  if (!(f = open_append(path, "UTAU.dat"))) return -1;
  fprintf(f, " %f", u_tau); fclose(f);
  if (!(f = open_append(path, "STATISTICS.dat"))) return -1;
  fprintf(f, "%e %e %e %e %e %e %e %e ", R(u_tau * LY * 0.5 / nu), R(uc * 0.5 * LY / (nu * N2)), R(1.0 / nu), R(Umean / u_tau),
          R(uc / (u_tau * N2)), R(Umean * N2 / uc), R(2.0 * u_tau * u_tau / (Umean * Umean)), R(2.0 * u_tau * u_tau / (uc * uc)));
  fprintf(f, "\n"); fclose(f);
  if (!(f = open_append(path, "RESOLUTION.dat"))) return -1;
  fprintf(f, "%f %f %f %f", R(1.5 * u_tau * LX / (nu * nx)), R(1.5 * u_tau * LZ / (nu * nz)), R(u_tau * dy_max / nu), R(u_tau * dy_min / nu));
  fprintf(f, "\n"); fclose(f);
#undef R
  return 0;
}

#define CK(call) do { int rc_ = (call); if (rc_) { fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, chan_last_error(h)); chan_destroy(h); free(ddv); free(g); free(um); free(nm); free(st); return 1; } } while (0)

int run_channel(const domain_t* d, const paths_t* p, const run_opts_t* o, int verbose) {
  const size_t n = (size_t)d->nx * d->nz * d->ny, bytes = n * 2 * (size_t)o->precision;
  void *ddv = malloc(bytes), *g = malloc(bytes), *um = calloc((size_t)d->ny * 2, (size_t)o->precision);
  double *nm = (double*)calloc((size_t)d->ny, sizeof(double)), *st = (double*)calloc((size_t)4 * d->ny, sizeof(double));
  chan_handle* h = 0;
  if (!ddv || !g || !um || !nm || !st) { fprintf(stderr, "out of host memory\n"); return 1; }
  chan_config c; chan_config_default(&c);
  c.nx = d->nx; c.ny = d->ny; c.nz = d->nz; c.precision = o->precision;
  c.reynolds = o->reynolds; c.qvelocity = o->qvelocity; c.cfl = o->cfl;
  CK(chan_create(&c, &h));
  if (verbose) printf("Allocation...\n");
  if (!strcmp(p->ginput, "-")) {                                        /* main.c:107-114 */
    if (verbose) printf("No input files specified. Creating empty files\n");
    gen_rand_data(ddv, g, (double)d->nx, n, o->precision);
  } else {
    if (verbose) printf("Reading Data...\nReading: %s, %s\n", p->ginput, p->ddvinput);
    if (read_field(p->ginput, g, d->nx, d->ny, d->nz, o->precision) || read_field(p->ddvinput, ddv, d->nx, d->ny, d->nz, o->precision)) { chan_destroy(h); return 1; }
  }
  CK(chan_set_fields(h, ddv, g));
  if (strcmp(p->umeaninput, "-")) {                                     /* main.c:119-123 */
    if (read_umean(p->umeaninput, um, d->ny, o->precision)) { chan_destroy(h); return 1; }
    CK(chan_set_umean(h, um));
  }
  if (o->fixed_dt > 0) CK(chan_set_fixed_dt(h, o->fixed_dt));
  if (verbose) printf("Starting RK iterations...\n");
  CK(chan_rk_init(h));                                                  /* RK3.c:120-123 */
  double time_elapsed = 0.0;
  int first = 0;
  if (o->restart[0] && strcmp(o->restart, "-")) {                       /* continue a checkpointed run bit for bit */
    if (read_checkpoint_file(h, o->restart)) { fprintf(stderr, "cannot restart from %s: %s\n", o->restart, chan_last_error(h)); chan_destroy(h); return 1; }
    chan_scalars sc0; CK(chan_get_scalars(h, &sc0));
    first = sc0.counter; time_elapsed = sc0.time;
    if (verbose) printf("Restarted from %s at (time,counter)=(%f,%d)\n", o->restart, time_elapsed, first);
  }
  for (int counter = first; counter < o->nsteps; counter++) {           /* RK3.c:127 */
    if (counter % 10 == 0) {                                            /* FREC_STATS, nonLinear.c:12-15 */
      CK(chan_statistics(h, st));
      if (append_calcst_files(p->path, st, d->ny, o->precision)) fprintf(stderr, "cannot append statistics under %s\n", p->path);
    }
    CK(chan_rk_step(h, 1));
    chan_scalars sc; CK(chan_get_scalars(h, &sc));
    CK(chan_get_umean(h, um)); CK(chan_get_nmean(h, nm));
    if (verbose) {                                                      /* RK3.c:91-97 */
      double N2 = (double)d->nx * (2 * d->nz - 2);
      printf("*****RK_STATISTICS****\nmax_V=(%e,%e,%e)\n(dt_c,dt_v)=(%f,%f)\n", sc.umax[0] / N2, sc.umax[1] / N2, sc.umax[2] / N2, sc.dt_c, sc.dt_v);
    }
    if (append_mean_files(p->path, um, nm, &sc, d->nx, d->ny, d->nz, o->reynolds, o->qvelocity, o->precision, verbose))
      fprintf(stderr, "cannot append mean-flow files under %s\n", p->path);
    time_elapsed += sc.dt;                                              /* dt of the step just taken (RK3.c:179-180) */
    if (verbose) printf("\n(time,counter)=(%f,%d)", time_elapsed, counter + 1);
    if (o->checkpoint_every > 0 && o->checkpoint[0] && strcmp(o->checkpoint, "-") && (counter + 1) % o->checkpoint_every == 0)
      if (write_checkpoint_file(h, o->checkpoint)) fprintf(stderr, "cannot write checkpoint %s\n", o->checkpoint);
  }
  CK(chan_get_fields(h, ddv, g));                                       /* writeData (main.c:139) */
  int rc = 0;
  if (p->goutput[0] && strcmp(p->goutput, "-")) rc |= write_field(p->goutput, g, d->nx, d->ny, d->nz, o->precision);
  if (p->ddvoutput[0] && strcmp(p->ddvoutput, "-")) rc |= write_field(p->ddvoutput, ddv, d->nx, d->ny, d->nz, o->precision);
  CK(chan_get_umean(h, um));
  if (p->umeanoutput[0] && strcmp(p->umeanoutput, "-")) rc |= write_umean(p->umeanoutput, um, d->ny, o->precision);   /* writeU (main.c:142) */
  chan_destroy(h); free(ddv); free(g); free(um); free(nm); free(st);
  return rc ? 1 : 0;
}
