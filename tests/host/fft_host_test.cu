// CPU execution of the plane-stage FFT phases (channel_gpu_b200/csrc/plane_core.cuh): the threads
// of a block are emulated one after the other, phase by phase, and the results are compared with
// O(N^2) DFTs.  Built and run by tests/test_fft_host.py (no GPU needed).
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <complex>
#include "../../channel_gpu_b200/csrc/plane_core.cuh"

using namespace chan;
using namespace chan::fft;
typedef std::complex<double> cd;
static const double PI = 3.14159265358979323846;

template <class T2> struct Host {
  FftPlan p; std::vector<int> inv, pos; std::vector<T2> tw;
  void build(int n, bool small_last) {
    make_plan(&p, n, small_last);
    if (p.n < 0) { printf("no plan for %d\n", n); exit(1); }
    inv.resize(n); pos.resize(n); tw.resize(n);
    for (int i = 0; i < n; i++) { pos[i] = perm_pos_of(p, i); inv[pos[i]] = i; }
    for (int k = 0; k < n; k++) { tw[k].x = (decltype(tw[k].x))cos(2 * PI * k / n); tw[k].y = (decltype(tw[k].x))(-sin(2 * PI * k / n)); }
    p.tw = tw.data(); p.inv = inv.data(); p.pos = pos.data();
  }
};
static double frand() { return 2.0 * rand() / RAND_MAX - 1.0; }
#define EACH_THREAD for (int tid = 0; tid < NTHR; tid++)
static const int NTHR = 96;

#define XSTAGE(SIGN, DIT)                                                                                         \
  switch (TK) {                                                                                                   \
    case 2: plan_stage_x<T2, SIGN, DIT, 2, true>(s.data(), pl, H.p.radix[t], H.p.span[t], tid, NTHR); break;            \
    case 4: plan_stage_x<T2, SIGN, DIT, 4, true>(s.data(), pl, H.p.radix[t], H.p.span[t], tid, NTHR); break;            \
    default: plan_stage_x<T2, SIGN, DIT, 8, true>(s.data(), pl, H.p.radix[t], H.p.span[t], tid, NTHR); break;           \
  }
template <class T2> double test_x(int nx, int Kz, int TK) {
  typedef typename RealOf<T2>::type T;
  Host<T2> H; H.build(nx, false);
  PlanT<T2> pl = plan_of<T2>(H.p);
  int kxmax = nx / 3, Kx = 2 * kxmax + 1;
  // two kx segments, as a 2-rank plane stage sees its receive buffer: rows of both at plane jl = 1
  SegMap sm; sm.nseg = 2; sm.closed = 0; int w0 = Kx / 2 + 1, w1 = Kx - w0;
  sm.start[0] = 0; sm.start[1] = w0; sm.start[2] = Kx; sm.width[0] = w0; sm.width[1] = w1; sm.base[0] = 0; sm.base[1] = (long)3 * w0 * Kz;
  const int jl = 1; const size_t segtot = (size_t)3 * Kx * Kz;
  auto at = [&](int i, int kz) { return seg_offset(sm, jl, i, Kz) + kz; };
  std::vector<T2> src(segtot), q((size_t)nx * Kz), dst(segtot), s((size_t)nx * TK);
  for (auto& v : src) { v.x = (T)frand(); v.y = (T)frand(); }
  double err = 0, nrm = 0;
  for (int kz0 = 0; kz0 < Kz; kz0 += TK) {
    EACH_THREAD { FFT_DISPATCH_RADIX(true, pl.r_first, (x_inv_load_stage0<T2, RR>(s.data(), src.data(), sm, jl, pl, TK, kz0, Kz, Kx, kxmax, tid, NTHR))); }
    for (int t = 1; t + 1 < pl.nstage; t++) EACH_THREAD XSTAGE(+1, true)
    EACH_THREAD { FFT_DISPATCH_RADIX(true, pl.r_last, (x_inv_last_store<T2, RR>(s.data(), q.data(), pl, TK, kz0, Kz, tid, NTHR))); }
  }
  for (int kz = 0; kz < Kz; kz++) for (int x = 0; x < nx; x++) {
    cd acc = 0;
    for (int i = 0; i < Kx; i++) { int kx = i <= kxmax ? i : i - Kx; acc += cd(src[at(i, kz)].x, src[at(i, kz)].y) * std::polar(1.0, 2 * PI * kx * x / nx); }
    cd got(q[(size_t)x * Kz + kz].x, q[(size_t)x * Kz + kz].y);
    err += std::norm(got - acc); nrm += std::norm(acc);
  }
  double e1 = sqrt(err / nrm);
  // forward of the physical lines must give back nx * src on the active modes
  SegDst<T2> dsts; for (int i = 0; i < 8; i++) dsts.p[i] = dst.data();
  for (int kz0 = 0; kz0 < Kz; kz0 += TK) {
    EACH_THREAD { FFT_DISPATCH_RADIX(true, pl.r_last, (x_fwd_load_first<T2, RR>(s.data(), q.data(), pl, TK, kz0, Kz, tid, NTHR))); }
    for (int t = pl.nstage - 2; t >= 1; t--) EACH_THREAD XSTAGE(-1, false)
    EACH_THREAD { FFT_DISPATCH_RADIX(true, pl.r_first, (x_fwd_stage0_store<T2, RR>(s.data(), dsts, sm, jl, pl, TK, kz0, Kz, Kx, kxmax, (T)nx, tid, NTHR))); }
  }
  err = nrm = 0;
  for (int i = 0; i < Kx; i++) for (int kz = 0; kz < Kz; kz++) { size_t o = at(i, kz); err += std::norm(cd(dst[o].x - src[o].x, dst[o].y - src[o].y)); nrm += std::norm(cd(src[o].x, src[o].y)); }
  double e2 = sqrt(err / nrm);
  printf("  x n=%4d plan", nx); for (int i = 0; i < pl.nstage; i++) printf(" %d", H.p.radix[i]);
  printf("  inv err %.2e  roundtrip err %.2e\n", e1, e2);
  return e1 > e2 ? e1 : e2;
}

static std::vector<double> c2r(const std::vector<cd>& A, int n, int Kz) {   // un-normalised, imag of bin 0 ignored
  std::vector<double> r(n);
  for (int z = 0; z < n; z++) {
    double acc = A[0].real();
    for (int k = 1; k < Kz; k++) acc += 2.0 * (A[k] * std::polar(1.0, 2 * PI * k * z / n)).real();
    r[z] = acc;
  }
  return r;
}
static std::vector<cd> r2c(const std::vector<double>& r, int n, int Kz) {
  std::vector<cd> A(Kz);
  for (int k = 0; k < Kz; k++) { cd acc = 0; for (int z = 0; z < n; z++) acc += r[z] * std::polar(1.0, -2 * PI * k * z / n); A[k] = acc; }
  return A;
}

template <class T2, int MODE> double test_z(int n, int XP, int nxp_valid) {
  typedef typename RealOf<T2>::type T;
  constexpr int NIN = MODE == ZM_ROTOR ? 6 : (MODE == ZM_IDENTITY ? 2 : 1);
  constexpr int NOUT = MODE == ZM_ROTOR ? 3 : (MODE == ZM_IDENTITY ? 2 : 0);
  Host<T2> H; H.build(n, true);
  PlanT<T2> pl = plan_of<T2>(H.p);
  int Kz = n / 3 + 1, nxrows = 2 * XP, ST = XP * NIN, SS = ST | 1, sha = -1;
  // inverse side as launch_z lays it out: packed rows + extra skew when the line count is even and 16 | n / r_last
  if (MODE == ZM_ROTOR && !(ST & 1) && (n / pl.r_last) % 16 == 0) { SS = ST; sha = (ST % 8 == 0) ? 0 : ((ST % 8 == 4) ? 1 : 2); }
  // forward side: packed rows of L = XP*NOUT elements with skew multiplier L when L is odd (as launch_z chooses)
  const int LO = XP * NOUT; const bool pack = (LO & 1) && MODE == ZM_ROTOR;
  const int SSo = pack ? LO : (SS | 1), SKo = pack ? LO : 1;
  size_t fstride = (size_t)nxrows * Kz;
  std::vector<T2> Q((size_t)NIN * fstride), Q0, s((size_t)n * (SS + 2) + n), s2((size_t)n * (SS + 2) + n);
  for (auto& v : Q) { v.x = (T)frand(); v.y = (T)frand(); }
  Q0 = Q;
  T n2 = (T)7.0;
  std::vector<T> phys((size_t)nxrows * (n + 2), (T)0);
  T um[3] = { 0, 0, 0 };
  EACH_THREAD { FFT_DISPATCH_RADIX(true, pl.r_first, (z_inv_load_stage0<T2, RR, NIN>(s.data(), Q.data(), fstride, pl, make_fastdiv(n / pl.r_first), XP, SS, sha, nxp_valid, Kz, tid, NTHR))); }
  for (int t = 1; t + 1 < pl.nstage; t++) EACH_THREAD plan_stage<T2, +1, true, true>(s.data(), pl, H.p.radix[t], make_fastdiv(H.p.span[t]), H.p.tws[t], SS, make_fastdiv(ST), ST, ST, tid, NTHR, 1, sha);
  EACH_THREAD {
    switch (pl.r_last) {
      case 2: z_mid_fused<T2, 2, MODE>(s.data(), s2.data(), pl, make_fastdiv(n / pl.r_last), XP, SS, sha, SSo, SKo, nxp_valid, n2, um, phys.data(), (size_t)(n + 2), tid, NTHR); break;
      case 3: z_mid_fused<T2, 3, MODE>(s.data(), s2.data(), pl, make_fastdiv(n / pl.r_last), XP, SS, sha, SSo, SKo, nxp_valid, n2, um, phys.data(), (size_t)(n + 2), tid, NTHR); break;
      default: z_mid_fused<T2, 4, MODE>(s.data(), s2.data(), pl, make_fastdiv(n / pl.r_last), XP, SS, sha, SSo, SKo, nxp_valid, n2, um, phys.data(), (size_t)(n + 2), tid, NTHR); break;
    }
  }
  if (MODE != ZM_EXPORT) {
    for (int t = pl.nstage - 2; t >= 0; t--) EACH_THREAD plan_stage<T2, -1, false, true>(s2.data(), pl, H.p.radix[t], make_fastdiv(H.p.span[t]), H.p.tws[t], SSo, make_fastdiv(XP * NOUT), NOUT, NOUT, tid, NTHR, SKo);
    EACH_THREAD z_unpack_store<T2, NIN, NOUT>(s2.data(), Q.data(), fstride, pl, make_fastdiv(Kz), XP, SSo, SKo, nxp_valid, tid, NTHR);
  }
  // reference
  double err = 0, nrm = 0, umref[3] = { 0, 0, 0 };
  for (int xr = 0; xr < 2 * nxp_valid; xr++) {
    std::vector<std::vector<double>> ph(NIN);
    for (int f = 0; f < NIN; f++) {
      std::vector<cd> A(Kz); for (int k = 0; k < Kz; k++) A[k] = cd(Q0[f * fstride + (size_t)xr * Kz + k].x, Q0[f * fstride + (size_t)xr * Kz + k].y);
      ph[f] = c2r(A, n, Kz);
    }
    if (MODE == ZM_EXPORT) {
      for (int z = 0; z < n; z++) { double d = phys[(size_t)xr * (n + 2) + z] - ph[0][z]; err += d * d; nrm += ph[0][z] * ph[0][z]; }
      continue;
    }
    std::vector<std::vector<double>> out(NOUT, std::vector<double>(n));
    for (int z = 0; z < n; z++) {
      if (MODE == ZM_ROTOR) {
        double u = ph[0][z], v = ph[1][z], w = ph[2][z], ox = ph[3][z], oy = ph[4][z], oz = ph[5][z];
        umref[0] = fmax(umref[0], fabs(u)); umref[1] = fmax(umref[1], fabs(v)); umref[2] = fmax(umref[2], fabs(w));
        u /= n2; v /= n2; w /= n2; ox /= n2; oy /= n2; oz /= n2;
        out[0][z] = v * oz - w * oy; out[1][z] = w * ox - u * oz; out[2][z] = u * oy - v * ox;
      } else { for (int f = 0; f < NOUT; f++) out[f][z] = ph[f][z]; }
    }
    for (int f = 0; f < NOUT; f++) {
      std::vector<cd> A = r2c(out[f], n, Kz);
      for (int k = 0; k < Kz; k++) { cd got(Q[f * fstride + (size_t)xr * Kz + k].x, Q[f * fstride + (size_t)xr * Kz + k].y); err += std::norm(got - A[k]); nrm += std::norm(A[k]); }
    }
  }
  double e = sqrt(err / nrm);
  if (MODE == ZM_ROTOR) for (int c = 0; c < 3; c++) e = fmax(e, fabs(um[c] - umref[c]) / umref[c]);
  // rows of invalid pairs must be untouched
  for (int f = 0; f < NOUT; f++) for (int xr = 2 * nxp_valid; xr < nxrows; xr++) for (int k = 0; k < Kz; k++)
    if (Q[f * fstride + (size_t)xr * Kz + k].x != Q0[f * fstride + (size_t)xr * Kz + k].x) e = 1.0;
  printf("  z n=%4d mode %d XP=%d/%d plan", n, MODE, nxp_valid, XP); for (int i = 0; i < pl.nstage; i++) printf(" %d", H.p.radix[i]);
  printf("  err %.2e\n", e);
  return e;
}

// the closed-form plane-side addressing (even partition of the active kx over the ranks) against the select chain
static int test_segment_closed_form() {
  int bad = 0;
  for (int S = 2; S <= 8; S++)
    for (int Kx : { 9, 21, 33, 85, 257, 683, 1365 }) {
      if (Kx < S) continue;
      const int Kz = 7, nyl = 5, q = Kx / S, r = Kx % S;
      for (int mode = 0; mode < 2; mode++) {            // plane-side blocks | peers' pencil-side fields (j0 = 11)
        SegMap a; a.nseg = S; a.closed = 0;
        long off = 0; int st = 0;
        for (int g = 0; g < S; g++) { int w = q + (g < r ? 1 : 0); a.start[g] = st; a.width[g] = w; a.base[g] = mode ? (long)11 * w * Kz : off; off += (long)nyl * w * Kz; st += w; }
        a.start[S] = Kx;
        SegMap b = a; b.closed = 1; b.pq = q; b.pr = r;
        b.mq = q > 1 ? (unsigned)(0x100000000ull / (unsigned)q) + 1u : 0u; b.mq1 = (unsigned)(0x100000000ull / (unsigned)(q + 1)) + 1u;
        b.bms = mode ? 0 : (long)nyl * Kz; b.bmw = mode ? (long)11 * Kz : 0;
        for (int jl = 0; jl < nyl; jl++) for (int i = 0; i < Kx; i++) {
          int sa = -1, sb = -1;
          if (seg_offset(a, jl, i, Kz, &sa) != seg_offset(b, jl, i, Kz, &sb) || sa != sb) bad++;
        }
      }
    }
  printf("  segment addressing: closed form vs select chain, %d mismatches\n", bad);
  return bad;
}

int main() {
  srand(7);
  if (test_segment_closed_form()) return 1;
  double worst_d = 0, worst_f = 0;
  const int xs[] = { 8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 384, 1024, 10, 20, 40, 50, 60, 80, 120, 160, 240, 250, 320, 640, 960, 1280, 14, 28, 56, 70, 112, 196, 224, 448, 896 };   // from 10 on: lengths with a factor 5 or 7
  for (int n : xs) worst_d = fmax(worst_d, test_x<double2>(n, n <= 64 ? 5 : 3, 4));
  worst_d = fmax(worst_d, test_x<double2>(2048, 2, 2));
  worst_f = fmax(worst_f, test_x<float2>(384, 3, 8));
  worst_f = fmax(worst_f, test_x<float2>(128, 9, 8));
  const int zs[] = { 8, 12, 16, 24, 32, 48, 96, 128, 384, 10, 20, 30, 40, 50, 80, 160, 240, 320, 14, 28, 42, 56, 98, 112, 140, 224 };
  for (int n : zs) {
    worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(n, 2, 2));
    worst_d = fmax(worst_d, test_z<double2, ZM_IDENTITY>(n, 1, 1));
    worst_d = fmax(worst_d, test_z<double2, ZM_EXPORT>(n, 2, 1));
  }
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(768, 1, 1));
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(48, 3, 2));
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(640, 1, 1));
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(960, 1, 1));
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(100, 3, 3));
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(448, 1, 1));
  worst_d = fmax(worst_d, test_z<double2, ZM_ROTOR>(896, 1, 1));
  worst_f = fmax(worst_f, test_z<float2, ZM_ROTOR>(224, 2, 2));
  worst_f = fmax(worst_f, test_z<float2, ZM_ROTOR>(320, 1, 1));
  worst_f = fmax(worst_f, test_x<float2>(640, 3, 8));
  worst_f = fmax(worst_f, test_z<float2, ZM_ROTOR>(384, 1, 1));
  worst_f = fmax(worst_f, test_z<float2, ZM_ROTOR>(128, 2, 2));
  printf("WORST double %.3e float %.3e\n", worst_d, worst_f);
  return (worst_d < 1e-12 && worst_f < 2e-5) ? 0 : 1;
}
