/* Minimal HDF5 reader/writer for the reference's field files (one contiguous dataset "u", hit_mpi.c:257-422).
 * See h5min.c for the file map.  Plain C, no libhdf5. */
#ifndef CHAN_HOST_H5MIN_H
#define CHAN_HOST_H5MIN_H
#include <stdint.h>
#define H5MIN_DATA_OFFSET 2048
typedef struct h5min_info_t {
  uint64_t dims[3];          /* dataspace of "u": {NX, NY, 2*NZ} (hit_mpi.c:272-279) */
  int elem_bytes;            /* 4 = H5T_NATIVE_FLOAT, 8 = H5T_NATIVE_DOUBLE */
  uint64_t data_address, data_bytes, dataset_header, eof_address;
} h5min_info_t;
/* wrte_parallel_float (hit_mpi.c:257-339): creates `name`, dataset "u", dims, data verbatim.  0 = ok */
int h5min_write(const char* name, const void* data, int elem_bytes, const uint64_t dims[3]);
/* read_parallel_float (hit_mpi.c:341-422): 0 = ok, -2 = the file's type / dims differ from what was asked for */
int h5min_read(const char* name, void* dst, int elem_bytes, const uint64_t dims[3]);
int h5min_info(const char* name, h5min_info_t* info);
#endif
