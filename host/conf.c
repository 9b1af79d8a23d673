/* See conf.h.  Grammar handled (libconfig subset):  setting := name (':'|'=') value [';'|','] ;
 * value := integer | float | "string" | '{' setting* '}' ;  comments: '#', '//' to end of line and C style. */
#include "conf.h"
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct { const char* s; int line; config_t* c; } P;

static void skip_ws(P* p) {
  for (;;) {
    while (*p->s && isspace((unsigned char)*p->s)) { if (*p->s == '\n') p->line++; p->s++; }
    if (*p->s == '#' || (p->s[0] == '/' && p->s[1] == '/')) { while (*p->s && *p->s != '\n') p->s++; continue; }
    if (p->s[0] == '/' && p->s[1] == '*') {
      p->s += 2;
      while (*p->s && !(p->s[0] == '*' && p->s[1] == '/')) { if (*p->s == '\n') p->line++; p->s++; }
      if (*p->s) p->s += 2;
      continue;
    }
    return;
  }
}
static int fail(P* p, const char* msg) {
  snprintf(p->c->error, sizeof(p->c->error), "%s", msg); p->c->error_line = p->line; return -1;
}
static int parse_settings(P* p, const char* prefix, int in_group);

static int parse_value(P* p, const char* path) {
  skip_ws(p);
  if (*p->s == '{') { p->s++; if (parse_settings(p, path, 1)) return -1; return 0; }
  if (p->c->n >= CONF_MAX_ENTRIES) return fail(p, "too many settings");
  conf_entry* e = &p->c->e[p->c->n];
  snprintf(e->path, sizeof(e->path), "%s", path);
  if (*p->s == '"') {
    size_t k = 0; p->s++;
    while (*p->s && *p->s != '"') {
      char ch = *p->s++;
      if (ch == '\\' && *p->s) { ch = *p->s++; if (ch == 'n') ch = '\n'; else if (ch == 't') ch = '\t'; }
      if (ch == '\n') p->line++;
      if (k + 1 < sizeof(e->value)) e->value[k++] = ch;
    }
    if (*p->s != '"') return fail(p, "unterminated string");
    p->s++; e->value[k] = 0; e->is_string = 1;
  } else {
    size_t k = 0;
    while (*p->s && (isalnum((unsigned char)*p->s) || strchr("+-.", *p->s))) { if (k + 1 < sizeof(e->value)) e->value[k++] = *p->s; p->s++; }
    if (!k) return fail(p, "syntax error: value expected");
    e->value[k] = 0; e->is_string = 0;
  }
  p->c->n++;
  return 0;
}

static int parse_settings(P* p, const char* prefix, int in_group) {
  for (;;) {
    skip_ws(p);
    if (!*p->s) return in_group ? fail(p, "syntax error: '}' expected") : 0;
    if (*p->s == '}') { if (!in_group) return fail(p, "syntax error: unexpected '}'"); p->s++; return 0; }
    char name[64]; size_t k = 0;
    while (*p->s && (isalnum((unsigned char)*p->s) || *p->s == '_' || *p->s == '-' || *p->s == '*')) { if (k + 1 < sizeof(name)) name[k++] = *p->s; p->s++; }
    name[k] = 0;
    if (!k) return fail(p, "syntax error: setting name expected");
    skip_ws(p);
    if (*p->s != ':' && *p->s != '=') return fail(p, "syntax error: ':' or '=' expected");
    p->s++;
    char path[128];
    if (prefix[0]) snprintf(path, sizeof(path), "%s.%s", prefix, name); else snprintf(path, sizeof(path), "%s", name);
    if (parse_value(p, path)) return -1;
    skip_ws(p);
    if (*p->s == ';' || *p->s == ',') p->s++;
  }
}

int conf_parse_text(config_t* c, const char* text) {
  P p; p.s = text; p.line = 1; p.c = c;
  c->n = 0; c->error[0] = 0; c->error_line = 0;
  return parse_settings(&p, "", 0);
}

int read_config_file(config_t* c, const char* name) {
  FILE* f = fopen(name, "rb");
  if (!f) { c->n = 0; snprintf(c->error, sizeof(c->error), "file I/O error"); c->error_line = 0; return -1; }
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  char* buf = (char*)malloc((size_t)n + 1);
  if (!buf) { fclose(f); return -1; }
  size_t got = fread(buf, 1, (size_t)n, f); buf[got] = 0; fclose(f);
  int rc = conf_parse_text(c, buf);
  free(buf);
  if (rc) fprintf(stderr, "%s:%d - %s\n", name, c->error_line, c->error);    /* config.c:9-10 */
  return rc;
}

static const conf_entry* find(const config_t* c, const char* path) {
  for (int i = 0; i < c->n; i++) if (!strcmp(c->e[i].path, path)) return &c->e[i];
  return 0;
}
int conf_lookup_int(const config_t* c, const char* path, int* out) {
  const conf_entry* e = find(c, path); if (!e || e->is_string) return 0;
  char* end; long v = strtol(e->value, &end, 0); if (*end) return 0;
  *out = (int)v; return 1;
}
int conf_lookup_float(const config_t* c, const char* path, double* out) {
  const conf_entry* e = find(c, path); if (!e || e->is_string) return 0;
  char* end; double v = strtod(e->value, &end); if (*end) return 0;
  *out = v; return 1;
}
int conf_lookup_string(const config_t* c, const char* path, const char** out) {
  const conf_entry* e = find(c, path); if (!e || !e->is_string) return 0;
  *out = e->value; return 1;
}

void read_domain_from_config(domain_t* d, const config_t* c) {
  conf_lookup_int(c, "application.NX", &d->nx);
  conf_lookup_int(c, "application.NY", &d->ny);
  conf_lookup_int(c, "application.NZ", &d->nz);
}

int read_filenames_from_config(paths_t* p, const config_t* c) {
  struct { const char* key; char* dst; } m[] = {
    { "application.input.G", p->ginput }, { "application.output.G", p->goutput },
    { "application.input.DDV", p->ddvinput }, { "application.output.DDV", p->ddvoutput },
    { "application.input.UMEAN", p->umeaninput }, { "application.output.UMEAN", p->umeanoutput },
    { "application.path", p->path } };
  int missing = 0;
  for (unsigned i = 0; i < sizeof(m) / sizeof(m[0]); i++) {
    const char* v = 0;
    if (conf_lookup_string(c, m[i].key, &v)) { strncpy(m[i].dst, v, 99); m[i].dst[99] = 0; }
    else { m[i].dst[0] = 0; missing++; }     /* the reference would strcpy from a dangling pointer here */
  }
  return missing;
}

void read_run_opts_from_config(run_opts_t* o, const config_t* c) {
  o->reynolds = 3250.0; o->qvelocity = 1.8; o->cfl = 0.5; o->fixed_dt = 0.0; o->nsteps = 30000; o->precision = 4;
  conf_lookup_float(c, "application.REYNOLDS", &o->reynolds);
  conf_lookup_float(c, "application.QVELOCITY", &o->qvelocity);
  conf_lookup_float(c, "application.CFL", &o->cfl);
  conf_lookup_float(c, "application.FIXED_DT", &o->fixed_dt);
  conf_lookup_int(c, "application.NSTEPS", &o->nsteps);
  conf_lookup_int(c, "application.PRECISION", &o->precision);
  int iv;
  if (conf_lookup_int(c, "application.REYNOLDS", &iv)) o->reynolds = iv;
  if (conf_lookup_int(c, "application.FIXED_DT", &iv)) o->fixed_dt = iv;
  o->checkpoint_every = 0; strcpy(o->checkpoint, "-"); strcpy(o->restart, "-");
  conf_lookup_int(c, "application.checkpoint.EVERY", &o->checkpoint_every);
  const char* sv;
  if (conf_lookup_string(c, "application.checkpoint.FILE", &sv)) { strncpy(o->checkpoint, sv, sizeof(o->checkpoint) - 1); o->checkpoint[sizeof(o->checkpoint) - 1] = 0; }
  if (conf_lookup_string(c, "application.checkpoint.RESTART", &sv)) { strncpy(o->restart, sv, sizeof(o->restart) - 1); o->restart[sizeof(o->restart) - 1] = 0; }
}
