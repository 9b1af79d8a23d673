/* run.conf reader: the libconfig subset the reference's run.conf uses (run.conf:1-23), with the
 * entry points of config.c (read_config_file / read_domain_from_config / read_filenames_from_config,
 * config.c:4-42).  libconfig itself is not needed.  Plain C. */
#ifndef CHAN_HOST_CONF_H
#define CHAN_HOST_CONF_H

#define CONF_MAX_ENTRIES 128

typedef struct conf_entry { char path[128]; char value[256]; int is_string; } conf_entry;
typedef struct config_t { conf_entry e[CONF_MAX_ENTRIES]; int n; char error[256]; int error_line; } config_t;

/* domain_t / paths_t as in channel.h:23-40 */
typedef struct domain_t { int nx, ny, nz, rank, size, iglobal; } domain_t;
typedef struct paths_t {
  char ginput[100], goutput[100], ddvinput[100], ddvoutput[100], umeaninput[100], umeanoutput[100], path[100];
} paths_t;

/* extra run-time keys (all optional; defaults = the reference's compile-time constants) */
typedef struct run_opts_t {
  double reynolds;   /* application.REYNOLDS   channel.h:61  3250 */
  double qvelocity;  /* application.QVELOCITY  channel.h:62  1.8  */
  double cfl;        /* application.CFL        RK3.c:68      0.5  */
  double fixed_dt;   /* application.FIXED_DT   0 = calcDt          */
  int nsteps;        /* application.NSTEPS     RK3.c:124     30000 */
  int precision;     /* application.PRECISION  4 | 8         4     */
  int checkpoint_every;    /* application.checkpoint.EVERY    steps between checkpoints, 0 = none (the reference has none) */
  char checkpoint[100];    /* application.checkpoint.FILE     written every EVERY steps (one rank's full state)      */
  char restart[100];       /* application.checkpoint.RESTART  "-" or absent = start from the input fields           */
} run_opts_t;

int  conf_parse_text(config_t* c, const char* text);          /* 0 = ok; else c->error / c->error_line */
int  read_config_file(config_t* c, const char* name);         /* config.c:4-15 (returns instead of exit(1)) */
int  conf_lookup_int(const config_t* c, const char* path, int* out);          /* 1 = found, like libconfig */
int  conf_lookup_float(const config_t* c, const char* path, double* out);
int  conf_lookup_string(const config_t* c, const char* path, const char** out);
void read_domain_from_config(domain_t* d, const config_t* c);                 /* config.c:18-23 */
int  read_filenames_from_config(paths_t* p, const config_t* c);               /* config.c:25-42; 0 = all present */
void read_run_opts_from_config(run_opts_t* o, const config_t* c);
#endif
