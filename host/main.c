/* channel_b200: the reference's program frame (main.c:10-150) on one GPU, above libchannel_b200.
 *   channel_b200 [run.conf]
 * Reads the same run.conf (application.NX/NY/NZ, input/output file names, path; "-" = no input),
 * plus optional keys REYNOLDS, QVELOCITY, CFL, NSTEPS, FIXED_DT, PRECISION (conf.h). */
#include <stdio.h>
#include <string.h>
#include "conf.h"
#include "driver.h"

int main(int argc, char** argv) {
  const char* conf = argc > 1 ? argv[1] : "run.conf";       /* main.c:34 hard-codes "run.conf" */
  static config_t config;
  domain_t domain = { 0, 0, 0, 0, 1, 0 };
  paths_t path; run_opts_t opts;
  if (read_config_file(&config, conf)) return 1;
  printf("Reading configuration\n");
  read_domain_from_config(&domain, &config);
  printf("Reading input file names\n");
  if (read_filenames_from_config(&path, &config)) { fprintf(stderr, "%s: missing application.input/output/path entries\n", conf); return 1; }
  read_run_opts_from_config(&opts, &config);
  printf("rank, size, iglobal, nx, ny ,nz\n===============================\n%d, %d, %d, %d, %d, %d\n",
         domain.rank, domain.size, domain.iglobal, domain.nx, domain.ny, domain.nz);
  int rc = run_channel(&domain, &path, &opts, 1);
  printf("\n");
  return rc;
}
