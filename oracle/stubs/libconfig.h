/* Stand-in for libconfig: channel.h:103-105 only needs the type name. */
#ifndef ORACLE_STUB_LIBCONFIG_H
#define ORACLE_STUB_LIBCONFIG_H
typedef struct config_t { int unused; } config_t;
#endif
