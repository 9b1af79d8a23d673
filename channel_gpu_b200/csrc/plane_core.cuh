// Phases of the plane-stage kernels (kernels_plane.cu), written as __host__ __device__ functions
// of (tid, nthr) so tests/host/fft_host_test.cu can execute them on the CPU.
//
// Stage order.  Inverse transforms run DIT (digit-reversed in, natural out), forward transforms DIF
// (natural in, digit-reversed out).  Consequences:
//  * the m = 1 stage of an inverse transform is fused with the LOAD from global memory (the loader
//    gathers the r_0 inputs of a group through inv[], builds zero padding / Hermitian extension on
//    the fly, does the butterfly in registers and only then touches shared memory);
//  * the m = 1 stage of a forward x-transform is fused with the STORE (truncation = which outputs
//    are written);
//  * in z, the last inverse stage, the pointwise product u x omega and the first forward stage
//    share one pass: a thread holds all 6 lines of r_L positions k + q*m in registers.  These are
//    the large-span stages, whose shared-memory accesses are contiguous across threads.
// Physical space is in NATURAL order (x and z), so nothing downstream needs an index table.
#pragma once
#include "fft_core.cuh"
#include "chan_internal.h"

namespace chan {
namespace fft {

// scalar view of an FftPlan: the radix/span arrays are read from the FftPlan itself (a kernel parameter,
// i.e. constant bank) so that no per-thread local copy of them is made
template <class T2> struct PlanT { int n, nstage, r_first, r_last; FastDiv fsk; const T2* tw; const int* inv; const int* pos; };   // fsk: FastDiv of r_first

template <class T2> FFT_HD PlanT<T2> plan_of(const FftPlan& p) {
  PlanT<T2> q; q.n = p.n; q.nstage = p.nstage; q.r_first = p.r_first; q.r_last = p.r_last; q.fsk.d = (unsigned)p.r_first; q.fsk.M = p.rfirst_magic;
  q.tw = (const T2*)p.tw; q.inv = p.inv; q.pos = p.pos;
  return q;
}

// Offset of (plane jl, active-kx row ikx) in a plane-side buffer made of per-source blocks (SegMap).  The
// SegMap lives in the kernel parameters: it is only ever indexed with compile-time constants (a select
// chain over the segments) - a per-thread index would turn every lookup into a long-latency LDC in the
// middle of the address computation of the loads (18 % of xfft_inv's stall samples in the ncu source view).
FFT_HD size_t seg_offset(const SegMap& sm, int jl, int ikx, int Kz, int* seg_out = 0) {
  if (sm.closed) {
    const int big = sm.pr * (sm.pq + 1);                 // rows covered by the pr wider slabs
    FastDiv f1, f0; f1.d = (unsigned)(sm.pq + 1); f1.M = sm.mq1; f0.d = (unsigned)sm.pq; f0.M = sm.mq;
    const int sg = ikx < big ? fd_div(ikx, f1) : sm.pr + fd_div(ikx - big, f0);
    const int start = sg * sm.pq + (sg < sm.pr ? sg : sm.pr), width = sm.pq + (sg < sm.pr ? 1 : 0);
    if (seg_out) *seg_out = sg;
    return (size_t)(sm.bms * start + sm.bmw * width + ((long)jl * width + ikx - start) * Kz);
  }
  long base = sm.base[0]; int width = sm.width[0], start = sm.start[0], sg = 0;
  if (sm.nseg > 1) {
#pragma unroll
    for (int i = 1; i < 8; i++)
      if (i < sm.nseg && ikx >= sm.start[i]) { base = sm.base[i]; width = sm.width[i]; start = sm.start[i]; sg = i; }
  }
  if (seg_out) *seg_out = sg;
  return (size_t)base + ((size_t)jl * width + (ikx - start)) * Kz;
}
template <class T2> struct SegDst;
template <class T2> FFT_HD T2* seg_ptr(const SegDst<T2>& d, int nseg, int sg);
// destination of the forward store: one base pointer per kx segment (= per destination rank when the
// store goes straight into the peers' pencil-side fields over NVLink; all equal otherwise)
template <class T2> struct SegDst { T2* p[8]; };
template <class T2> FFT_HD T2* seg_ptr(const SegDst<T2>& d, int nseg, int sg) {
  T2* p = d.p[0];
  if (nseg > 1) {
#pragma unroll
    for (int i = 1; i < 8; i++) if (sg == i) p = d.p[i];
  }
  return p;
}

// =================================== x direction =================================================
// shared layout: s[pos*TK + t], t = kz within the tile (batch-fastest: conflict-free at every stride)

// inverse, stage 0 (m = 1) fused with the zero-padding load of the compact spectral rows
template <class T2, int R>
FFT_HD void x_inv_load_stage0(T2* s, const T2* src, const SegMap& sm, int jl, const PlanT<T2>& pl, int TK, int kz0, int Kz, int Kx, int kxmax, int tid, int nthr) {
  const int n = pl.n, ng = n / R;
  for (int b = tid; b < ng * TK; b += nthr) {
    int t = b % TK, g = b / TK, kz = kz0 + t;
    // all R loads are issued back to back (clamped address + select), never behind a branch
    T2 a[R]; int ik[R];
#pragma unroll
    const int nq0 = FFT_LDG(pl.inv + g * R);       // inv[g*R + q] = inv[g*R] + q*(n/R)  (perm_pos, kernels_tables.cu)
#pragma unroll
    for (int q = 0; q < R; q++) {
      int nq = nq0 + q * ng;
      ik[q] = nq <= kxmax ? nq : (nq >= n - kxmax ? nq - n + Kx : -1);
    }
    const bool kz_ok = kz < Kz;
#pragma unroll
    for (int q = 0; q < R; q++) a[q] = src[seg_offset(sm, jl, ik[q] < 0 ? 0 : ik[q], Kz) + (kz_ok ? kz : 0)];
#pragma unroll
    for (int q = 0; q < R; q++) if (ik[q] < 0 || !kz_ok) a[q] = mk2<T2>(0, 0);
    Dft<R, +1, T2>::run(a);
#pragma unroll
    for (int p = 0; p < R; p++) s[(size_t)(g * R + p) * TK + t] = a[p];
  }
}
// inverse, last stage fused with the store of the natural-order x line
template <class T2, int R>
FFT_HD void x_inv_last_store(const T2* s, T2* q, const PlanT<T2>& pl, int TK, int kz0, int Kz, int tid, int nthr) {
  const int n = pl.n, m = n / R;
  for (int b = tid; b < m * TK; b += nthr) {
    int t = b % TK, k = b / TK, kz = kz0 + t;
    T2 a[R];
#pragma unroll
    for (int i = 0; i < R; i++) a[i] = s[(size_t)(k + i * m) * TK + t];
    apply_twiddles<R, +1, T2>(a, pl.tw, k, 1);
    Dft<R, +1, T2>::run(a);
    if (kz < Kz) {
#pragma unroll
      for (int p = 0; p < R; p++) q[(size_t)(k + p * m) * Kz + kz] = a[p];
    }
  }
}
// forward, first stage (largest span) fused with the load of the natural-order x line
template <class T2, int R>
FFT_HD void x_fwd_load_first(T2* s, const T2* q, const PlanT<T2>& pl, int TK, int kz0, int Kz, int tid, int nthr) {
  const int n = pl.n, m = n / R;
  for (int b = tid; b < m * TK; b += nthr) {
    int t = b % TK, k = b / TK, kz = kz0 + t;
    T2 a[R];
#pragma unroll
    for (int p = 0; p < R; p++) a[p] = kz < Kz ? q[(size_t)(k + p * m) * Kz + kz] : mk2<T2>(0, 0);
    Dft<R, -1, T2>::run(a);
    apply_twiddles<R, -1, T2>(a, pl.tw, k, 1);
#pragma unroll
    for (int i = 0; i < R; i++) s[(size_t)(k + i * m) * TK + t] = a[i];
  }
}
// forward, stage 0 (m = 1) fused with the truncating store into the compact spectral rows
template <class T2, int R>
FFT_HD void x_fwd_stage0_store(const T2* s, const SegDst<T2>& dst, const SegMap& sm, int jl, const PlanT<T2>& pl, int TK, int kz0, int Kz, int Kx, int kxmax,
                               typename RealOf<T2>::type scale, int tid, int nthr) {
  const int n = pl.n, ng = n / R;
  for (int b = tid; b < ng * TK; b += nthr) {
    int t = b % TK, g = b / TK, kz = kz0 + t;
    T2 a[R];
#pragma unroll
    for (int i = 0; i < R; i++) a[i] = s[(size_t)(g * R + i) * TK + t];
    Dft<R, -1, T2>::run(a);
    if (kz < Kz) {
      const int np0 = FFT_LDG(pl.inv + g * R);
#pragma unroll
      for (int p = 0; p < R; p++) {
        int np = np0 + p * ng;
        int ikx = np <= kxmax ? np : (np >= n - kxmax ? np - n + Kx : -1);
        if (ikx >= 0) {
          T2 v = a[p];
          if (scale != 0) { v.x /= scale; v.y /= scale; }
          int sg; const size_t off = seg_offset(sm, jl, ikx, Kz, &sg);
          seg_ptr<T2>(dst, sm.nseg, sg)[off + kz] = v;
        }
      }
    }
  }
}
// one shared-memory stage of plan stage t over `nlines` participating lines (see smem_stage)
// W5: the length has a factor 5 (FFT_DISPATCH_RADIX, fft_core.cuh)
template <class T2, int SIGN, bool DIT, bool W5>
FFT_HD void plan_stage(T2* s, const PlanT<T2>& pl, int r, FastDiv fm, int tws, int st, FastDiv flines, int nact, int nper, int tid, int nthr, int skm = 1, int sha = -1) {
  FFT_DISPATCH_RADIX(W5, r, (smem_stage<T2, RR, SIGN, DIT>(s, pl.n, fm, st, flines, nact, nper, pl.tw, tid, nthr, pl.fsk, tws, skm, sha)));
}

template <class T2, int SIGN, bool DIT, int TK, bool W5>
FFT_HD void plan_stage_x(T2* s, const PlanT<T2>& pl, int r, int m, int tid, int nthr) {
  FFT_DISPATCH_RADIX(W5, r, (smem_stage_x<T2, RR, SIGN, DIT, TK>(s, pl.n, m, pl.tw, tid, nthr)));
}

// =================================== z direction =================================================
// One block works on XP x-pairs; shared layout s[zrow(pos) + slot], zrow(pos) = pos*SS + pos/r_0 (SS odd,
// >= XP*NIN; the pos/r_0 skew is explained at z_unpack_store).  Inverse side: slot = xpl*NIN + f.  Forward
// side: slot = xpl*NOUT + f, i.e. the NOUT output lines of all x-pairs are packed next to each other so
// that the forward stages touch contiguous elements (z_mid_fused makes the switch behind a barrier).
// Inverse side with an even number L of lines (always: 6 per x-pair): rows of exactly L elements and an extra
// skew row >> sha (sha = 2 for L = 2, 6 mod 8; 1 for 4; 0 for 0) instead of a one-element pad per row: the
// shared stages cross a hole only every 2^sha rows, the lanes of the middle stage (consecutive rows) stay apart.
// When the number L of forward lines is odd and the fused middle stage is a single round, the forward side
// uses rows of exactly L elements with skew L*(pos/r_0): the shared stages then walk one contiguous
// array (pitch SS left a hole per row that cost every second quarter-warp a bank conflict), L odd keeps
// the lanes of the middle stage (consecutive rows) apart, and r_0*L + L = L (mod 8) keeps the groups of
// the m = 1 stage apart.
// Two-for-one: line (xpl, f) holds field f at x0 = 2*xp (real part) and x1 = 2*xp+1 (imaginary part).
#ifdef __CUDA_ARCH__
#define FFT_BLOCK_SYNC() __syncthreads()
#else
#define FFT_BLOCK_SYNC()
#endif

// value of the Hermitian-extended, zero-padded packed spectrum at natural index nz
// index and kind of the Hermitian-extended, zero-padded packed spectrum at natural index nz:
// kind 0 -> zero, 1 -> A + iB (A,B at k = nz), 2 -> conj(A) + i conj(B) (k = n - nz)
FFT_HD void z_packed_index(int nz, int n, int Kz, int& k, int& kind) {
  kind = nz < Kz ? 1 : (nz > n - Kz ? 2 : 0);
  k = kind == 1 ? nz : (kind == 2 ? n - nz : 0);
}
template <class T2>
FFT_HD T2 z_packed_combine(T2 A, T2 B, int k, int kind) {
  typedef typename RealOf<T2>::type T;
  if (kind == 0) return mk2<T2>(0, 0);
  if (kind == 1 && k == 0) return mk2<T2>(A.x, B.x);            // imag of the kz=0 bin dropped, as C2R does
  T sg = kind == 1 ? (T)1 : (T)-1;
  return mk2<T2>(A.x - sg * B.y, sg * A.y + B.x);
}

// inverse, stage 0 fused with the packing load.  q0 -> Q[field 0][plane][x = 2*xp0][0]; fields are
// fstride apart, x rows Kz apart.
template <class T2, int R, int NIN>
FFT_HD void z_inv_load_stage0(T2* s, const T2* q0, size_t fstride, const PlanT<T2>& pl, FastDiv fng, int XP, int SS, int sha, int nxp_valid, int Kz, int tid, int nthr) {
  const int n = pl.n, ng = (int)fng.d, ST = XP * NIN;
  // lanes run over the groups g: for a fixed input q the natural indices inv[g*R] + q*ng of the ng groups
  // are a permutation of ng CONSECUTIVE kz, so a warp's load is one contiguous piece of a spectral row
  // (lanes over the lines instead would touch one 32-byte sector per lane); the shared-memory stores are
  // then R*SS + 1 elements apart, which the row skew makes conflict-free
  for (int b = tid; b < ng * ST; b += nthr) {
    int g, li; fd_divmod(b, fng, li, g);
    const int xpl = li / NIN, f = li - xpl * NIN;
    T2 a[R];
    if (xpl < nxp_valid) {
      const T2* qa = q0 + f * fstride + (size_t)(2 * xpl) * Kz;
      // loads are issued in independent batches of <= 8 pairs (clamped address + select), never behind a branch
      constexpr int HB = R == 16 ? 8 : (R == 12 ? 6 : R);
      const int nz0 = FFT_LDG(pl.inv + g * R);     // inv[g*R + q] = inv[g*R] + q*(n/R)
#pragma unroll
      for (int h0 = 0; h0 < R; h0 += HB) {
        int kk[HB], kind[HB]; T2 B[HB];
#pragma unroll
        for (int i = 0; i < HB; i++) z_packed_index(nz0 + (h0 + i) * ng, n, Kz, kk[i], kind[i]);
#pragma unroll
        for (int i = 0; i < HB; i++) { a[h0 + i] = qa[kk[i]]; B[i] = qa[Kz + kk[i]]; }
#pragma unroll
        for (int i = 0; i < HB; i++) a[h0 + i] = z_packed_combine<T2>(a[h0 + i], B[i], kk[i], kind[i]);
      }
      Dft<R, +1, T2>::run(a);
    } else {
#pragma unroll
      for (int i = 0; i < R; i++) a[i] = mk2<T2>(0, 0);
    }
#pragma unroll
    for (int p = 0; p < R; p++) s[(size_t)(g * R + p) * SS + g + (sha >= 0 ? (g * R + p) >> sha : 0) + li] = a[p];     // (g*R + p) / r_0 = g
  }
}

enum { ZM_ROTOR = 0, ZM_IDENTITY = 1, ZM_EXPORT = 2 };

// last inverse stage + pointwise + first forward stage, all 'NIN' lines of one x-pair in registers.
// umax3: per-thread running maxima of |u|,|v|,|w| (ROTOR).  phys0/phys1: output rows (EXPORT).
// Reads the inverse-side slots of s_in, writes the forward-side slots of s_out.  On the device
// s_out == s_in: the packed output slots of x-pair xpl overlap input slots of x-pairs <= xpl, which are
// read in the same or an earlier round of the loop, so one barrier per round between the reads and the
// writes is enough.  The host emulation (threads run one after the other) passes a separate s_out.
template <class T2, int R, int MODE>
FFT_HD void z_mid_fused(const T2* s_in, T2* s_out, const PlanT<T2>& pl, FastDiv fmid, int XP, int SS, int sha, int SSo, int SKo, int nxp_valid, typename RealOf<T2>::type n2,
                        typename RealOf<T2>::type* umax3, typename RealOf<T2>::type* phys, size_t phys_xstride, int tid, int nthr) {
  typedef typename RealOf<T2>::type T;
  constexpr int NIN = MODE == ZM_ROTOR ? 6 : (MODE == ZM_IDENTITY ? 2 : 1);
  constexpr int NOUT = MODE == ZM_ROTOR ? 3 : (MODE == ZM_IDENTITY ? 2 : 0);
  const int m = (int)fmid.d;                             // m = n / R
  const int ist = m * SS + skew_of(m, pl.fsk) + (sha >= 0 ? m >> sha : 0);   // r_0 (and 2^sha) divide m = n / r_last (nstage >= 2)
  const int osto = m * SSo + SKo * skew_of(m, pl.fsk);   // forward-side rows: pitch SSo, skew multiplier SKo
  const T rn2 = ((T)1 / n2) * ((T)1 / n2);
  const int nround = (m * XP + nthr - 1) / nthr;
  for (int it = 0; it < nround; it++) {
    const int b = tid + it * nthr;
    int k, xpl; fd_divmod(b, fmid, xpl, k);
    const bool work = b < m * XP && xpl < nxp_valid;
    const size_t row = (size_t)k * SS + skew_of(k, pl.fsk) + (sha >= 0 ? k >> sha : 0);
    const size_t rowo = (size_t)k * SSo + SKo * skew_of(k, pl.fsk);
    T2 a[NIN][R];
    if (work) {
#pragma unroll
      for (int f = 0; f < NIN; f++) {
#pragma unroll
        for (int i = 0; i < R; i++) a[f][i] = s_in[row + (size_t)i * ist + xpl * NIN + f];
        apply_twiddles<R, +1, T2>(a[f], pl.tw, k, 1);
        Dft<R, +1, T2>::run(a[f]);
      }
      if (MODE == ZM_EXPORT) {
        T* o0 = phys + (size_t)(2 * xpl) * phys_xstride; T* o1 = o0 + phys_xstride;
#pragma unroll
        for (int p = 0; p < R; p++) { o0[k + p * m] = a[0][p].x; o1[k + p * m] = a[0][p].y; }
      }
      if (MODE == ZM_ROTOR) {
        // rotorkernel (convolution_kernels.cu:107-131): every factor divided by N2, m = u x omega;
        // .x is the point (x0, z), .y the point (x1, z).  Maxima are of the un-normalised velocities.
#pragma unroll
        for (int p = 0; p < R; p++) {
          T2 u = a[0][p], v = a[1][p], w = a[2][p], ox = a[3][p], oy = a[4][p], oz = a[5][p];
          umax3[0] = fmax(umax3[0], fmax(fabs(u.x), fabs(u.y)));
          umax3[1] = fmax(umax3[1], fmax(fabs(v.x), fabs(v.y)));
          umax3[2] = fmax(umax3[2], fmax(fabs(w.x), fabs(w.y)));
          // (v/N2)(oz/N2) - (w/N2)(oy/N2) as one scaling by 1/N2^2: 12 divisions per point were 15 % of the
          // kernel's instructions; the result differs from the divide-first form in the last bit only
          a[0][p] = mk2<T2>((v.x * oz.x - w.x * oy.x) * rn2, (v.y * oz.y - w.y * oy.y) * rn2);
          a[1][p] = mk2<T2>((w.x * ox.x - u.x * oz.x) * rn2, (w.y * ox.y - u.y * oz.y) * rn2);
          a[2][p] = mk2<T2>((u.x * oy.x - v.x * ox.x) * rn2, (u.y * oy.y - v.y * ox.y) * rn2);
        }
      }
      if (MODE != ZM_EXPORT) {
#pragma unroll
        for (int f = 0; f < NOUT; f++) {
          Dft<R, -1, T2>::run(a[f]);
          apply_twiddles<R, -1, T2>(a[f], pl.tw, k, 1);
        }
      }
    }
    if (MODE == ZM_EXPORT) continue;
    if (NOUT != NIN) FFT_BLOCK_SYNC();
    if (work) {
#pragma unroll
      for (int f = 0; f < NOUT; f++) {
#pragma unroll
        for (int i = 0; i < R; i++) s_out[rowo + (size_t)i * osto + xpl * NOUT + f] = a[f][i];
      }
    }
  }
}

// Hermitian unpacking of the two-for-one pair + truncation to kz < Kz, from the digit-reversed
// output of the forward transform
// Threads run along k for coalesced stores, and consecutive k are digit-reversed positions that
// differ in their HIGH digits, i.e. rows a multiple of r_0 apart: with a plain row pitch they would
// all fall into the same banks (7x the ideal number of wavefronts in the ncu source view).  The
// pos/r_0 skew of the layout spreads exactly those rows over the banks.
template <class T2, int NIN, int NOUT>
FFT_HD void z_unpack_store(const T2* s, T2* q0, size_t fstride, const PlanT<T2>& pl, FastDiv fkz, int XP, int SS, int SK, int nxp_valid, int tid, int nthr) {
  typedef typename RealOf<T2>::type T;
  const int n = pl.n, Kz = (int)fkz.d;
  for (int b = tid; b < Kz * NOUT * XP; b += nthr) {
    int k, l; fd_divmod(b, fkz, l, k);
    const int xpl = l / NOUT, f = l - xpl * NOUT;
    if (xpl >= nxp_valid) continue;
    const int pk = FFT_LDG(pl.pos + k), pn = FFT_LDG(pl.pos + (k ? n - k : 0));
    T2 Yk = s[(size_t)pk * SS + SK * skew_of(pk, pl.fsk) + xpl * NOUT + f];
    T2 Yn = s[(size_t)pn * SS + SK * skew_of(pn, pl.fsk) + xpl * NOUT + f];
    const T h = (T)0.5;
    T2* qa = q0 + f * fstride + (size_t)(2 * xpl) * Kz;
    qa[k] = mk2<T2>(h * (Yk.x + Yn.x), h * (Yk.y - Yn.y));
    qa[Kz + k] = mk2<T2>(h * (Yk.y + Yn.y), h * (Yn.x - Yk.x));
  }
}

}  // namespace fft
}  // namespace chan
