/* Periodic checkpoint / restart files above chan_checkpoint_save / chan_checkpoint_load.
 * The reference has none: it writes its fields once at the end (main.c:139) and a restarted run starts with
 * dt = 0 and R = 0 (RK3.c:114).  One file per rank: the library's versioned payload verbatim. */
#include "driver.h"
#include <stdio.h>
#include <stdlib.h>

int write_checkpoint_file(chan_handle* h, const char* name) {
  const size_t n = chan_checkpoint_bytes(h);
  void* buf = malloc(n);
  if (!buf) return -1;
  int rc = chan_checkpoint_save(h, buf);
  if (!rc) {
    /* write to a temporary name first: a crash while writing must not destroy the previous checkpoint */
    char tmp[512]; snprintf(tmp, sizeof(tmp), "%s.tmp", name);
    FILE* f = fopen(tmp, "wb");
    if (!f) rc = -1;
    else { rc = fwrite(buf, 1, n, f) == n ? 0 : -1; if (fclose(f)) rc = -1; if (!rc && rename(tmp, name)) rc = -1; }
  }
  free(buf);
  return rc;
}

int read_checkpoint_file(chan_handle* h, const char* name) {
  const size_t n = chan_checkpoint_bytes(h);
  FILE* f = fopen(name, "rb");
  if (!f) { fprintf(stderr, "\nerror lectura: %s", name); return -1; }
  void* buf = malloc(n);
  if (!buf) { fclose(f); return -1; }
  int rc = fread(buf, 1, n, f) == n ? 0 : -1;
  if (!rc && fgetc(f) != EOF) rc = -1;                    /* a longer file belongs to another grid */
  fclose(f);
  if (!rc) rc = chan_checkpoint_load(h, buf);
  free(buf);
  return rc;
}
