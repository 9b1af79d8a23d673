// Hand-written mixed-radix FFT building blocks for the plane stage (kernels_plane.cu).
// Everything here is __host__ __device__ so that tests/host/fft_host_test.cu can run the very
// same code on the CPU, emulating the threads of a block one after the other (within a phase
// the threads touch disjoint shared-memory elements).
//
// Radices 2, 3, 4, 5, 7 are written out; 6, 8, 10, 12, 14, 16 are composites done entirely in registers with
// compile-time twiddles (fft_twiddle_consts.h, generated with
//   python -c "import math; ..."   see channel_gpu_b200/csrc/gen note in DESIGN.md section 4).
//
// Index conventions.  A plan is a radix list r_0..r_{s-1} with spans m_0 = 1, m_{t+1} = m_t r_t.
// Stage t works on groups g of m_t*r_t consecutive positions: elements g*m_t*r_t + q*m_t + k.
//   DIT stage:  a_q *= W_{m r}^{q k};  y = DFT_r(a);  store y_p at ... + p*m + k
//   DIF stage:  b = DFT_r(a);  b_q *= W_{m r}^{q k};  store          (the transposed graph)
// DIT stages 0..s-1 turn digit-reversed input into natural output; DIF stages s-1..0 do the
// opposite.  `inv[pos]` is the natural index held at digit-reversed position `pos`.
#pragma once
#include "fft_twiddle_consts.h"

namespace chan {
namespace fft {

template <class T2> struct RealOf { typedef decltype(T2().x) type; };
template <class T2> FFT_HD T2 mk2(typename RealOf<T2>::type a, typename RealOf<T2>::type b) { T2 r; r.x = a; r.y = b; return r; }
template <class T2> FFT_HD T2 cadd(T2 a, T2 b) { return mk2<T2>(a.x + b.x, a.y + b.y); }
template <class T2> FFT_HD T2 csub(T2 a, T2 b) { return mk2<T2>(a.x - b.x, a.y - b.y); }
template <class T2> FFT_HD T2 cmul(T2 a, T2 b) { return mk2<T2>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
// a * w for the forward transform (SIGN < 0), a * conj(w) for the inverse (SIGN > 0); w = exp(-i theta)
template <int SIGN, class T2> FFT_HD T2 twmul(T2 a, T2 w) {
  return SIGN < 0 ? mk2<T2>(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x) : mk2<T2>(a.x * w.x + a.y * w.y, a.y * w.x - a.x * w.y);
}
template <int SIGN, class T2> FFT_HD T2 muli(T2 a) { return SIGN > 0 ? mk2<T2>(-a.y, a.x) : mk2<T2>(a.y, -a.x); }   // SIGN * i * a

// multiply by the compile-time twiddle W_R^j (forward) or its conjugate (inverse)
template <int R, int SIGN, class T2> FFT_HD T2 tw_const(T2 a, int j) {
  typedef typename RealOf<T2>::type T;
  j = j % R;
  if (j == 0) return a;
  if ((4 * j) % R == 0) {                       // quarter turns: W = (-i)^(4j/R) forward
    int qt = (4 * j) / R;
    if (qt == 2) return mk2<T2>(-a.x, -a.y);
    T2 t = muli<SIGN, T2>(a);                    // SIGN*i*a ; forward W^(R/4) = -i = SIGN*i for SIGN=-1
    return qt == 1 ? t : mk2<T2>(-t.x, -t.y);
  }
  T2 w = mk2<T2>((T)fft_tw_cos(R, j), (T)(-fft_tw_sin(R, j)));
  return twmul<SIGN, T2>(a, w);
}

// ---- in-register DFTs, natural order in and out:  y_p = sum_q a_q exp(SIGN 2 pi i p q / R) -------
template <int R, int SIGN, class T2> struct Dft;
template <int SIGN, class T2> struct Dft<1, SIGN, T2> { FFT_HD static void run(T2*) {} };
template <int SIGN, class T2> struct Dft<2, SIGN, T2> {
  FFT_HD static void run(T2* a) { T2 t = a[0]; a[0] = cadd(t, a[1]); a[1] = csub(t, a[1]); }
};
template <int SIGN, class T2> struct Dft<3, SIGN, T2> {
  FFT_HD static void run(T2* a) {
    typedef typename RealOf<T2>::type T;
    const T h = (T)0.86602540378443864676372317075294;
    T2 t1 = cadd(a[1], a[2]);
    T2 t2 = mk2<T2>(a[0].x - (T)0.5 * t1.x, a[0].y - (T)0.5 * t1.y);
    T2 d = csub(a[1], a[2]);
    T2 t3 = muli<SIGN, T2>(mk2<T2>(h * d.x, h * d.y));
    a[0] = cadd(a[0], t1); a[1] = cadd(t2, t3); a[2] = csub(t2, t3);
  }
};
template <int SIGN, class T2> struct Dft<4, SIGN, T2> {
  FFT_HD static void run(T2* a) {
    T2 t0 = cadd(a[0], a[2]), t1 = csub(a[0], a[2]), t2 = cadd(a[1], a[3]), t3 = muli<SIGN, T2>(csub(a[1], a[3]));
    a[0] = cadd(t0, t2); a[2] = csub(t0, t2); a[1] = cadd(t1, t3); a[3] = csub(t1, t3);
  }
};
// 5 points: y_1,4 = a_0 - t/4 + (sqrt5/4)(t1 - t2) +- SIGN i (s1 d1 + s2 d2), y_2,3 = a_0 - t/4 - (sqrt5/4)(t1 - t2) +- SIGN i (s2 d1 - s1 d2)
// with t1 = a_1 + a_4, t2 = a_2 + a_3, t = t1 + t2, d1 = a_1 - a_4, d2 = a_2 - a_3, s1 = sin 72, s2 = sin 144
template <int SIGN, class T2> struct Dft<5, SIGN, T2> {
  FFT_HD static void run(T2* a) {
    typedef typename RealOf<T2>::type T;
    const T c5 = (T)0.55901699437494742410229341718282, s1 = (T)0.95105651629515357211643933337938, s2 = (T)0.58778525229247312916870595463907;
    T2 t1 = cadd(a[1], a[4]), t2 = cadd(a[2], a[3]), d1 = csub(a[1], a[4]), d2 = csub(a[2], a[3]);
    T2 t = cadd(t1, t2);
    T2 m1 = mk2<T2>(a[0].x - (T)0.25 * t.x, a[0].y - (T)0.25 * t.y);
    T2 m2 = mk2<T2>(c5 * (t1.x - t2.x), c5 * (t1.y - t2.y));
    T2 e1 = cadd(m1, m2), e2 = csub(m1, m2);
    T2 o1 = muli<SIGN, T2>(mk2<T2>(s1 * d1.x + s2 * d2.x, s1 * d1.y + s2 * d2.y));
    T2 o2 = muli<SIGN, T2>(mk2<T2>(s2 * d1.x - s1 * d2.x, s2 * d1.y - s1 * d2.y));
    a[0] = cadd(a[0], t); a[1] = cadd(e1, o1); a[4] = csub(e1, o1); a[2] = cadd(e2, o2); a[3] = csub(e2, o2);
  }
};
// 7 points: with t_j = a_j + a_(7-j), d_j = a_j - a_(7-j), c_k = cos(2 pi k/7), s_k = sin(2 pi k/7):
// y_p, y_(7-p) = a_0 + sum_j t_j c_(pj mod 7) +- SIGN i sum_j d_j s_(pj mod 7), p, j = 1..3
template <int SIGN, class T2> struct Dft<7, SIGN, T2> {
  FFT_HD static void run(T2* a) {
    typedef typename RealOf<T2>::type T;
    const T c1 = (T)0.62348980185873353052500488400424, c2 = (T)-0.22252093395631440428890256449679, c3 = (T)-0.90096886790241912623610231950745;
    const T s1 = (T)0.78183148246802980870844452667406, s2 = (T)0.97492791218182360701813168299393, s3 = (T)0.43388373911755812047576833284836;
    T2 t1 = cadd(a[1], a[6]), t2 = cadd(a[2], a[5]), t3 = cadd(a[3], a[4]);
    T2 d1 = csub(a[1], a[6]), d2 = csub(a[2], a[5]), d3 = csub(a[3], a[4]);
    T2 e1 = mk2<T2>(a[0].x + c1 * t1.x + c2 * t2.x + c3 * t3.x, a[0].y + c1 * t1.y + c2 * t2.y + c3 * t3.y);
    T2 e2 = mk2<T2>(a[0].x + c2 * t1.x + c3 * t2.x + c1 * t3.x, a[0].y + c2 * t1.y + c3 * t2.y + c1 * t3.y);
    T2 e3 = mk2<T2>(a[0].x + c3 * t1.x + c1 * t2.x + c2 * t3.x, a[0].y + c3 * t1.y + c1 * t2.y + c2 * t3.y);
    T2 o1 = muli<SIGN, T2>(mk2<T2>(s1 * d1.x + s2 * d2.x + s3 * d3.x, s1 * d1.y + s2 * d2.y + s3 * d3.y));
    T2 o2 = muli<SIGN, T2>(mk2<T2>(s2 * d1.x - s3 * d2.x - s1 * d3.x, s2 * d1.y - s3 * d2.y - s1 * d3.y));
    T2 o3 = muli<SIGN, T2>(mk2<T2>(s3 * d1.x - s1 * d2.x + s2 * d3.x, s3 * d1.y - s1 * d2.y + s2 * d3.y));
    a[0] = cadd(a[0], cadd(t1, cadd(t2, t3)));
    a[1] = cadd(e1, o1); a[6] = csub(e1, o1); a[2] = cadd(e2, o2); a[5] = csub(e2, o2); a[3] = cadd(e3, o3); a[4] = csub(e3, o3);
  }
};
// R = R1*R2: n = R1*n2 + n1, k = k2 + R2*k1
template <int R1, int R2, int SIGN, class T2> FFT_HD void dft_composite(T2* a) {
  constexpr int R = R1 * R2;
  T2 s[R1][R2];
#pragma unroll
  for (int n1 = 0; n1 < R1; n1++) {
#pragma unroll
    for (int n2 = 0; n2 < R2; n2++) s[n1][n2] = a[R1 * n2 + n1];
    Dft<R2, SIGN, T2>::run(s[n1]);
#pragma unroll
    for (int k2 = 1; k2 < R2; k2++) s[n1][k2] = tw_const<R, SIGN, T2>(s[n1][k2], n1 * k2);
  }
#pragma unroll
  for (int k2 = 0; k2 < R2; k2++) {
    T2 c[R1];
#pragma unroll
    for (int n1 = 0; n1 < R1; n1++) c[n1] = s[n1][k2];
    Dft<R1, SIGN, T2>::run(c);
#pragma unroll
    for (int k1 = 0; k1 < R1; k1++) a[k2 + R2 * k1] = c[k1];
  }
}
template <int SIGN, class T2> struct Dft<6, SIGN, T2> { FFT_HD static void run(T2* a) { dft_composite<2, 3, SIGN, T2>(a); } };
template <int SIGN, class T2> struct Dft<8, SIGN, T2> { FFT_HD static void run(T2* a) { dft_composite<2, 4, SIGN, T2>(a); } };
template <int SIGN, class T2> struct Dft<10, SIGN, T2> { FFT_HD static void run(T2* a) { dft_composite<2, 5, SIGN, T2>(a); } };
template <int SIGN, class T2> struct Dft<12, SIGN, T2> { FFT_HD static void run(T2* a) { dft_composite<3, 4, SIGN, T2>(a); } };
template <int SIGN, class T2> struct Dft<14, SIGN, T2> { FFT_HD static void run(T2* a) { dft_composite<2, 7, SIGN, T2>(a); } };
template <int SIGN, class T2> struct Dft<16, SIGN, T2> { FFT_HD static void run(T2* a) { dft_composite<4, 4, SIGN, T2>(a); } };

// ---- stage twiddles W_{mR}^{q k}, q = 1..R-1 -----------------------------------------------------
// fp64: one table load, the other powers by squaring/multiplying (log depth): the stages are bound
// by the shared-memory / L1 pipe, fp64 FMAs are spare.  fp32: direct table loads (accuracy).
template <class T2> struct UsePowers { static const bool value = sizeof(typename RealOf<T2>::type) == 8; };

#ifdef __CUDA_ARCH__
#define FFT_LDG(p) __ldg(p)
#else
#define FFT_LDG(p) (*(p))
#endif

template <int R, int SIGN, class T2> FFT_HD void apply_twiddles(T2* a, const T2* tw, int k, int tws) {
  if (R == 1) return;
  if (UsePowers<T2>::value) {
    T2 pw[R > 1 ? R : 2];
    pw[1] = FFT_LDG(tw + k * tws);
#pragma unroll
    for (int q = 2; q < R; q++) pw[q] = (q & 1) ? cmul(pw[q - 1], pw[1]) : cmul(pw[q / 2], pw[q / 2]);
#pragma unroll
    for (int q = 1; q < R; q++) a[q] = twmul<SIGN, T2>(a[q], pw[q]);
  } else {
#pragma unroll
    for (int q = 1; q < R; q++) a[q] = twmul<SIGN, T2>(a[q], FFT_LDG(tw + q * k * tws));
  }
}

// ---- a whole shared-memory stage ----------------------------------------------------------------
// element (row, line) lives at s[row*st + skew(row) + line_index(l)], line_index(l) = (l / nact) * nper + l % nact
// (nact of every nper interleaved lines take part; nlines = number of participating lines).
// skew(row) = row / sk for sk > 0 (z transforms: one extra element every r_0 rows, so that rows that differ
// in the HIGH digits of the digit-reversed position fall into different banks, see z_unpack_store), 0 for sk = 0.
// sk divides the span m of every stage but the m = 1 stage, where all rows of a butterfly share one skew.

// Division of a work-item index by a launch-time constant.  The generic integer divisions of the index
// arithmetic were ~15 % of the z kernel's instructions (ncu source view); with M = floor(2^32/d) + 1 the
// quotient (b*M) >> 32 is exact for 0 <= b < 2^16 and 1 < d < 2^16 (M = 0 encodes d = 1).  All numerators
// here are element counts of one shared-memory tile (< 14080 elements, see launch_z / launch_x).
struct FastDiv { unsigned d, M; };
inline FastDiv make_fastdiv(int d) { FastDiv f; f.d = (unsigned)d; f.M = d > 1 ? (unsigned)(0x100000000ull / (unsigned)d) + 1u : 0u; return f; }
FFT_HD int fd_div(int b, FastDiv f) {
#ifdef __CUDA_ARCH__
  return f.M ? (int)__umulhi((unsigned)b, f.M) : b;
#else
  return f.M ? (int)(((unsigned long long)(unsigned)b * f.M) >> 32) : b;
#endif
}
FFT_HD void fd_divmod(int b, FastDiv f, int& q, int& r) { q = fd_div(b, f); r = b - q * (int)f.d; }
// skew(row) = row / r_0 for the z transforms (fsk = FastDiv of r_0), 0 for fsk.d = 0 (x transforms)
FFT_HD int skew_of(int row, FastDiv fsk) { return fsk.d ? fd_div(row, fsk) : 0; }

template <class T2, int R, int SIGN, bool DIT>
FFT_HD void smem_stage(T2* s, int n, FastDiv fm, int st, FastDiv flines, int nact, int nper, const T2* tw, int tid, int nthr, FastDiv fsk, int tws, int skm = 1, int sha = -1) {
  // tws = n / (m R), the twiddle stride of this stage; row address = row*st + skm*(row/r_0) + (row >> sha) [sha >= 0]
  const int m = (int)fm.d, nlines = (int)flines.d;
  const int nb = n / R;
  const int qst = m * st + ((fsk.d && m >= (int)fsk.d) ? skm * skew_of(m, fsk) + (sha >= 0 ? m >> sha : 0) : 0);
  for (int b = tid; b < nb * nlines; b += nthr) {
    int l, bf; fd_divmod(b, flines, bf, l);
    int li = nact == nper ? l : (l / nact) * nper + l % nact;
    int g, k; fd_divmod(bf, fm, g, k);
    const int row0 = g * m * R + k;
    T2* p = s + (size_t)row0 * st + skm * skew_of(row0, fsk) + (sha >= 0 ? row0 >> sha : 0) + li;
    T2 a[R];
#pragma unroll
    for (int q = 0; q < R; q++) a[q] = p[(size_t)q * qst];
    if (DIT) apply_twiddles<R, SIGN, T2>(a, tw, k, tws);
    Dft<R, SIGN, T2>::run(a);
    if (!DIT) apply_twiddles<R, SIGN, T2>(a, tw, k, tws);
#pragma unroll
    for (int q = 0; q < R; q++) p[(size_t)q * qst] = a[q];
  }
}

// x transforms: the lines of a tile are a compile-time count (TK), the rows carry no skew
template <class T2, int R, int SIGN, bool DIT, int TK>
FFT_HD void smem_stage_x(T2* s, int n, int m, const T2* tw, int tid, int nthr) {
  const int nb = n / R, tws = n / (m * R);
  for (int b = tid; b < nb * TK; b += nthr) {
    int l = b % TK, bf = b / TK;
    int g = bf / m, k = bf - g * m;
    T2* p = s + (size_t)(g * m * R + k) * TK + l;
    T2 a[R];
#pragma unroll
    for (int q = 0; q < R; q++) a[q] = p[(size_t)q * m * TK];
    if (DIT) apply_twiddles<R, SIGN, T2>(a, tw, k, tws);
    Dft<R, SIGN, T2>::run(a);
    if (!DIT) apply_twiddles<R, SIGN, T2>(a, tw, k, tws);
#pragma unroll
    for (int q = 0; q < R; q++) p[(size_t)q * m * TK] = a[q];
  }
}

// Radix dispatch.  W5 (a compile-time flag of the calling kernel) says whether the length has a factor 5 or 7: the kernels
// are instantiated twice, and the instances for the 2^a 3^b lengths carry no radix-5/7/10/14 code at all (with the two extra
// cases in ONE switch the z stage and the inverse x transform of the 1024 x 385 x 768 grid came out 1.5 % slower - same registers,
// other code layout; same-box A/B, profiles/r02_radix5_ab.txt).
#define FFT_DISPATCH_RADIX_23(r, CALL)                                                 \
  switch (r) {                                                                         \
    case 2: { constexpr int RR = 2; CALL; } break;                                     \
    case 3: { constexpr int RR = 3; CALL; } break;                                     \
    case 4: { constexpr int RR = 4; CALL; } break;                                     \
    case 6: { constexpr int RR = 6; CALL; } break;                                     \
    case 8: { constexpr int RR = 8; CALL; } break;                                     \
    case 12: { constexpr int RR = 12; CALL; } break;                                   \
    default: { constexpr int RR = 16; CALL; } break;                                   \
  }
#define FFT_DISPATCH_RADIX_2357(r, CALL)                                                \
  switch (r) {                                                                         \
    case 2: { constexpr int RR = 2; CALL; } break;                                     \
    case 3: { constexpr int RR = 3; CALL; } break;                                     \
    case 4: { constexpr int RR = 4; CALL; } break;                                     \
    case 5: { constexpr int RR = 5; CALL; } break;                                     \
    case 6: { constexpr int RR = 6; CALL; } break;                                     \
    case 7: { constexpr int RR = 7; CALL; } break;                                     \
    case 8: { constexpr int RR = 8; CALL; } break;                                     \
    case 10: { constexpr int RR = 10; CALL; } break;                                   \
    case 12: { constexpr int RR = 12; CALL; } break;                                   \
    case 14: { constexpr int RR = 14; CALL; } break;                                   \
    default: { constexpr int RR = 16; CALL; } break;                                   \
  }
#define FFT_DISPATCH_RADIX(W5, r, CALL)                                                \
  do {                                                                                 \
    if constexpr (W5) { FFT_DISPATCH_RADIX_2357(r, CALL) } else { FFT_DISPATCH_RADIX_23(r, CALL) } \
  } while (0)

}  // namespace fft
}  // namespace chan
