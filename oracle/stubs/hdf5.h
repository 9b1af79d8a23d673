/* Empty stand-in: the hot-path sources include hdf5.h through channel.h but
 * call nothing from it (only hit_mpi.c/io.c do, and they are not built). */
#ifndef ORACLE_STUB_HDF5_H
#define ORACLE_STUB_HDF5_H
#endif
