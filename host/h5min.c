/* Minimal self-contained HDF5 container for the reference's field files.
 *
 * The reference stores every field as ONE contiguous dataset named "u" of H5T_NATIVE_FLOAT with dataspace
 * {NX, NY, 2*NZ} in a file of its own (hit_mpi.c:257-339 writer, :341-422 reader, dims at :272-279, called with
 * Nz = 2*domain.nz at io.c:48-54); the bytes are the device array verbatim, i.e. [i][k][j][re,im].
 * libhdf5 is not available here, so this file writes and reads exactly that shape of file from the published HDF5
 * File Format Specification (version 0 superblock, version 1 object headers, old-style group = B-tree node +
 * symbol-table node + local heap, contiguous layout) — the layout libhdf5 1.8 itself produces for such a file.
 * The double-state build uses H5T_NATIVE_DOUBLE in the same place (hit_mpi.c:92-255 are the reference's double twins).
 *
 * File map written by h5min_write (all addresses absolute, 8-byte offsets and lengths):
 *      0  superblock (96 bytes) incl. the root group's symbol-table entry
 *     96  root group object header: one Symbol Table message (B-tree, heap)
 *    136  B-tree node "TREE" (group node, K = 16: 544 bytes), one child
 *    680  local heap "HEAP" header (32 bytes)
 *    712  heap data segment (32 bytes): "" | "u" | free block
 *    744  symbol-table node "SNOD" (leaf K = 4: 328 bytes), one entry: "u"
 *   1072  dataset object header: Dataspace, Datatype, Fill Value, Data Layout messages
 *   2048  raw data (H5MIN_DATA_OFFSET)
 */
#include "h5min.h"
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define UNDEF 0xFFFFFFFFFFFFFFFFull
static const unsigned char kSig[8] = { 0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n' };

static void put(unsigned char* p, uint64_t v, int n) { for (int i = 0; i < n; i++) p[i] = (unsigned char)(v >> (8 * i)); }
static uint64_t get(const unsigned char* p, int n) { uint64_t v = 0; for (int i = 0; i < n; i++) v |= (uint64_t)p[i] << (8 * i); return v; }

enum { A_ROOT = 96, A_TREE = 136, A_HEAP = 680, A_HDATA = 712, A_SNOD = 744, A_DSET = 1072 };

/* one version-1 header message: type, data size (multiple of 8), flags; returns the position after it */
static unsigned char* msg(unsigned char* p, int type, int size) { put(p, (uint64_t)type, 2); put(p + 2, (uint64_t)size, 2); p[4] = 0; return p + 8; }

int h5min_write(const char* name, const void* data, int elem_bytes, const uint64_t dims[3]) {
  if (elem_bytes != 4 && elem_bytes != 8) return -1;
  unsigned char* hd = (unsigned char*)calloc(1, H5MIN_DATA_OFFSET);
  if (!hd) return -1;
  const uint64_t nbytes = dims[0] * dims[1] * dims[2] * (uint64_t)elem_bytes;
  unsigned char* p = hd;
  /* superblock, version 0 */
  memcpy(p, kSig, 8); p[8] = 0; p[9] = 0; p[10] = 0; p[11] = 0; p[12] = 0; p[13] = 8; p[14] = 8; p[15] = 0;
  put(p + 16, 4, 2); put(p + 18, 16, 2); put(p + 20, 0, 4);
  put(p + 24, 0, 8); put(p + 32, UNDEF, 8); put(p + 40, H5MIN_DATA_OFFSET + nbytes, 8); put(p + 48, UNDEF, 8);
  /* root group symbol-table entry: name offset 0, header address, cache type 1 (B-tree + heap in the scratch pad) */
  put(p + 56, 0, 8); put(p + 64, A_ROOT, 8); put(p + 72, 1, 4); put(p + 76, 0, 4); put(p + 80, A_TREE, 8); put(p + 88, A_HEAP, 8);
  /* root group object header (version 1): prefix 16 bytes, then one Symbol Table message (0x0011) of 16 bytes */
  p = hd + A_ROOT; p[0] = 1; put(p + 2, 1, 2); put(p + 4, 1, 4); put(p + 8, 24, 4);
  { unsigned char* q = msg(p + 16, 0x0011, 16); put(q, A_TREE, 8); put(q + 8, A_HEAP, 8); }
  /* B-tree node of the group: level 0, one entry; key0 = heap offset 0 (""), child = SNOD, key1 = heap offset of "u" */
  p = hd + A_TREE; memcpy(p, "TREE", 4); p[4] = 0; p[5] = 0; put(p + 6, 1, 2); put(p + 8, UNDEF, 8); put(p + 16, UNDEF, 8);
  put(p + 24, 0, 8); put(p + 32, A_SNOD, 8); put(p + 40, 8, 8);
  /* local heap */
  p = hd + A_HEAP; memcpy(p, "HEAP", 4); p[4] = 0; put(p + 8, 32, 8); put(p + 16, 16, 8); put(p + 24, A_HDATA, 8);
  p = hd + A_HDATA; p[8] = 'u'; put(p + 16, 1, 8); put(p + 24, 16, 8);      /* free block: next = 1 (none), size 16 */
  /* symbol-table node with the one link */
  p = hd + A_SNOD; memcpy(p, "SNOD", 4); p[4] = 1; put(p + 6, 1, 2);
  put(p + 8, 8, 8); put(p + 16, A_DSET, 8); put(p + 24, 0, 4);
  /* dataset object header */
  p = hd + A_DSET; p[0] = 1; put(p + 2, 4, 2); put(p + 4, 1, 4);
  unsigned char* q = p + 16;
  { unsigned char* d = msg(q, 0x0001, 32); d[0] = 1; d[1] = 3; d[2] = 0; for (int i = 0; i < 3; i++) put(d + 8 + 8 * i, dims[i], 8); q = d + 32; }
  { unsigned char* d = msg(q, 0x0003, 24);     /* IEEE little-endian floating point, version 1, class 1 */
    d[0] = 0x11; d[1] = 0x20; d[2] = (unsigned char)(elem_bytes == 4 ? 31 : 63); d[3] = 0; put(d + 4, (uint64_t)elem_bytes, 4);
    put(d + 8, 0, 2); put(d + 10, (uint64_t)(8 * elem_bytes), 2);
    d[12] = (unsigned char)(elem_bytes == 4 ? 23 : 52); d[13] = (unsigned char)(elem_bytes == 4 ? 8 : 11);
    d[14] = 0; d[15] = (unsigned char)(elem_bytes == 4 ? 23 : 52); put(d + 16, elem_bytes == 4 ? 127 : 1023, 4);
    q = d + 24; }
  { unsigned char* d = msg(q, 0x0005, 8); d[0] = 2; d[1] = 1; d[2] = 2; d[3] = 0; q = d + 8; }        /* fill value v2: early allocation, undefined */
  { unsigned char* d = msg(q, 0x0008, 24); d[0] = 3; d[1] = 1; put(d + 2, H5MIN_DATA_OFFSET, 8); put(d + 10, nbytes, 8); q = d + 24; }
  put(p + 8, (uint64_t)(q - (p + 16)), 4);
  FILE* f = fopen(name, "wb");
  if (!f) { free(hd); return -1; }
  int ok = fwrite(hd, 1, H5MIN_DATA_OFFSET, f) == H5MIN_DATA_OFFSET && fwrite(data, 1, nbytes, f) == nbytes;
  free(hd);
  return (fclose(f) == 0 && ok) ? 0 : -1;
}

/* ---- reader: follows the addresses in the file (nothing is assumed about where the writer put things) ---- */
typedef struct { FILE* f; uint64_t base; } rd_t;
static int rd_at(rd_t* r, uint64_t addr, void* dst, size_t n) {
  if (fseek(r->f, (long)(r->base + addr), SEEK_SET)) return -1;
  return fread(dst, 1, n, r->f) == n ? 0 : -1;
}
/* finds header message `type` in a version-1 object header (continuation blocks followed); copies up to cap bytes */
static int find_msg(rd_t* r, uint64_t hdr, int type, unsigned char* out, int cap) {
  unsigned char pre[16];
  if (rd_at(r, hdr, pre, 16) || pre[0] != 1) return -1;
  int nmsg = (int)get(pre + 2, 2);
  uint64_t pos = hdr + 16, end = pos + get(pre + 8, 4);
  uint64_t cont_a[8], cont_n[8]; int ncont = 0, icont = 0;
  for (int m = 0; m < nmsg; m++) {
    while (pos + 8 > end) { if (icont >= ncont) return -1; pos = cont_a[icont]; end = pos + cont_n[icont]; icont++; }
    unsigned char mh[8];
    if (rd_at(r, pos, mh, 8)) return -1;
    int t = (int)get(mh, 2), sz = (int)get(mh + 2, 2);
    if (t == type) { int n = sz < cap ? sz : cap; return rd_at(r, pos + 8, out, (size_t)n) ? -1 : n; }
    if (t == 0x0010 && ncont < 8) { unsigned char c[16]; if (rd_at(r, pos + 8, c, 16)) return -1; cont_a[ncont] = get(c, 8); cont_n[ncont] = get(c + 8, 8); ncont++; }
    pos += 8 + (uint64_t)sz;
  }
  return -1;
}

int h5min_info(const char* name, h5min_info_t* info) {
  memset(info, 0, sizeof(*info));
  rd_t r; r.f = fopen(name, "rb"); r.base = 0;
  if (!r.f) return -1;
  int rc = -1;
  unsigned char sb[96];
  do {
    if (rd_at(&r, 0, sb, 96) || memcmp(sb, kSig, 8)) break;
    if (sb[8] != 0 || sb[13] != 8 || sb[14] != 8) break;             /* version-0 superblock, 8-byte offsets/lengths */
    r.base = get(sb + 24, 8);
    info->eof_address = get(sb + 40, 8);
    const uint64_t root = get(sb + 64, 8);
    unsigned char st[16];
    if (find_msg(&r, root, 0x0011, st, 16) != 16) break;
    const uint64_t tree = get(st, 8), heap = get(st + 8, 8);
    unsigned char hh[32];
    if (rd_at(&r, heap, hh, 32) || memcmp(hh, "HEAP", 4)) break;
    const uint64_t hdata = get(hh + 24, 8);
    /* walk down the leftmost path of the B-tree to level 0, then scan the leaves left to right for "u" */
    uint64_t node = tree; unsigned char th[24];
    for (;;) { if (rd_at(&r, node, th, 24) || memcmp(th, "TREE", 4) || th[4] != 0) { node = UNDEF; break; }
      if (th[5] == 0) break;
      unsigned char c[8]; if (rd_at(&r, node + 32, c, 8)) { node = UNDEF; break; } node = get(c, 8); }
    uint64_t dset = UNDEF;
    while (node != UNDEF && dset == UNDEF) {
      if (rd_at(&r, node, th, 24)) break;
      int used = (int)get(th + 6, 2);
      for (int e = 0; e < used && dset == UNDEF; e++) {
        unsigned char c[8]; if (rd_at(&r, node + 24 + 8 + 16 * (uint64_t)e, c, 8)) break;
        unsigned char sn[8]; uint64_t snod = get(c, 8);
        if (rd_at(&r, snod, sn, 8) || memcmp(sn, "SNOD", 4)) break;
        int nsym = (int)get(sn + 6, 2);
        for (int s = 0; s < nsym; s++) {
          unsigned char en[40], nm[2];
          if (rd_at(&r, snod + 8 + 40 * (uint64_t)s, en, 40) || rd_at(&r, hdata + get(en, 8), nm, 2)) break;
          if (nm[0] == 'u' && nm[1] == 0) { dset = get(en + 8, 8); break; }
        }
      }
      node = get(th + 16, 8);       /* right sibling */
    }
    if (dset == UNDEF) break;
    info->dataset_header = dset;
    unsigned char m[64];
    int n = find_msg(&r, dset, 0x0001, m, 64);
    if (n < 8 || m[0] != 1 || m[1] != 3) break;
    for (int i = 0; i < 3; i++) info->dims[i] = get(m + 8 + 8 * i, 8);
    n = find_msg(&r, dset, 0x0003, m, 64);
    if (n < 20 || (m[0] & 0x0f) != 1 || (m[1] & 1)) break;             /* floating point, little endian */
    info->elem_bytes = (int)get(m + 4, 4);
    n = find_msg(&r, dset, 0x0008, m, 64);
    if (n < 18 || m[0] != 3 || m[1] != 1) break;                      /* version 3, contiguous */
    info->data_address = get(m + 2, 8); info->data_bytes = get(m + 10, 8);
    if (info->data_bytes != info->dims[0] * info->dims[1] * info->dims[2] * (uint64_t)info->elem_bytes) break;
    info->data_address += r.base;
    rc = 0;
  } while (0);
  fclose(r.f);
  return rc;
}

int h5min_read(const char* name, void* dst, int elem_bytes, const uint64_t dims[3]) {
  h5min_info_t in;
  if (h5min_info(name, &in)) return -1;
  if (in.elem_bytes != elem_bytes || in.dims[0] != dims[0] || in.dims[1] != dims[1] || in.dims[2] != dims[2]) return -2;
  FILE* f = fopen(name, "rb");
  if (!f) return -1;
  int ok = !fseek(f, (long)in.data_address, SEEK_SET) && fread(dst, 1, in.data_bytes, f) == in.data_bytes;
  fclose(f);
  return ok ? 0 : -1;
}
