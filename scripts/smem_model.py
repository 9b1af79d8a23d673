"""Bank-conflict model of the z-stage shared-memory layouts (plane_core.cuh).

Elements are 16 bytes (double2): a 128-bit access is served per quarter-warp (8 lanes), conflict-free when the
8 element addresses fall into 8 different 16-byte bank groups (address mod 8).  For every phase of zstage_kernel
the model walks the thread -> element mapping of the code and counts wavefronts against the ideal number.
"""
import sys


def wavefronts(addrs):
    """addrs: element address per lane of one warp-wide 128-bit access (None = inactive lane)."""
    tot = ideal = 0
    for q in range(0, len(addrs), 8):
        grp = [a for a in addrs[q:q + 8] if a is not None]
        if not grp:
            continue
        cnt = {}
        for a in set(grp):
            cnt[a % 8] = cnt.get(a % 8, 0) + 1
        tot += max(cnt.values()); ideal += 1
    return tot, ideal


class Layout:
    def __init__(self, pitch, r0, skm=1, sha=-1):
        self.pitch, self.r0, self.skm, self.sha = pitch, r0, skm, sha

    def row(self, r):
        return r * self.pitch + self.skm * (r // self.r0) + ((r >> self.sha) if self.sha >= 0 else 0)


def phase_counts(n, radices, lines, nthr, lay, phase):
    """phase: 'load' (m=1 stage, lanes over groups), ('stage', t) shared stage t, 'mid' (lanes over rows),
    ('unpack', Kz).  Returns (wavefronts, ideal) summed over the accesses of one block."""
    r0, rl = radices[0], radices[-1]
    span = [1]
    for r in radices[:-1]:
        span.append(span[-1] * r)
    tot = ideal = 0

    def warp_access(addr_of_item, nitems, per_item):
        nonlocal tot, ideal
        for base in range(0, nitems, nthr):
            for w in range(0, nthr, 32):
                lanes = [b if b < nitems else None for b in range(base + w, base + w + 32)]
                for e in range(per_item):
                    t, i = wavefronts([None if b is None else addr_of_item(b, e) for b in lanes])
                    tot += t; ideal += i

    if phase == 'load':
        ng = n // r0
        warp_access(lambda b, p: lay.row((b % ng) * r0 + p) + b // ng, ng * lines, r0)
    elif phase[0] == 'stage':
        t = phase[1]; R, m = radices[t], span[t]
        nb = n // R
        def addr(b, q):
            l, bf = b % lines, b // lines
            g, k = bf // m, bf % m
            return lay.row(g * m * R + k + q * m) + l
        warp_access(addr, nb * lines, R)
    elif phase == 'mid':
        m = n // rl
        per_pair = lines_per_pair = phase_lines = None
        # lanes over k, one access per (line, i)
        xp = 1
        def addr(b, e):
            k, xpl = b % m, b // m
            f, i = e % lines, e // lines
            return lay.row(k + i * m) + xpl * lines + f
        warp_access(addr, m * xp, lines * rl)
    elif phase[0] == 'unpack':
        Kz = phase[1]
        # digit-reversed position of natural index k
        def pos(k):
            p, s = 0, n
            kk = k
            res = 0
            # perm_pos (kernels_tables.cu): least significant natural digit (mod r_last) -> most significant position digit
            sp = list(span)
            for t in range(len(radices) - 1, -1, -1):
                r = radices[t]
                res += (kk % r) * sp[t]; kk //= r
            return res
        def addr(b, e):
            k, l = b % Kz, b // Kz
            kk = k if e == 0 else ((n - k) if k else 0)
            return lay.row(pos(kk)) + l
        warp_access(addr, Kz * lines, 2)
    return tot, ideal


def report(n, radices, xp, nthr, Kz):
    r0 = radices[0]
    lin, lout = 6 * xp, 3 * xp
    cases = {
        "padded (old)": (Layout(lin | 1, r0), Layout(lin | 1, r0)),
        "packed (new)": (Layout(lin, r0, 1, {0: 0, 4: 1}.get(lin % 8, 2)) if lin % 2 == 0 else Layout(lin, r0),
                         Layout(lout, r0, lout) if lout % 2 else Layout(lin | 1, r0)),
    }
    out = {}
    for name, (li, lo) in cases.items():
        rows = []
        rows.append(("load stage stores", phase_counts(n, radices, lin, nthr, li, 'load')))
        for t in range(1, len(radices) - 1):
            rows.append(("inverse stage %d" % t, phase_counts(n, radices, lin, nthr, li, ('stage', t))))
        rows.append(("middle stage loads", phase_counts(n, radices, lin, nthr, li, 'mid')))
        rows.append(("middle stage stores", phase_counts(n, radices, lout, nthr, lo, 'mid')))
        for t in range(len(radices) - 2, -1, -1):
            rows.append(("forward stage %d" % t, phase_counts(n, radices, lout, nthr, lo, ('stage', t))))
        rows.append(("unpack loads", phase_counts(n, radices, lout, nthr, lo, ('unpack', Kz))))
        out[name] = rows
    return out


if __name__ == "__main__":
    for n, radices, xp, nthr, Kz in ((384, (16, 6, 4), 1, 128, 129), (768, (16, 12, 4), 1, 192, 257), (128, (8, 4, 4), 3, 128, 43)):
        print("n=%d radices=%s xp=%d threads=%d" % (n, radices, xp, nthr))
        rep = report(n, radices, xp, nthr, Kz)
        for name, rows in rep.items():
            t = sum(r[1][0] for r in rows); i = sum(r[1][1] for r in rows)
            print("  %-13s total %6d wavefronts, ideal %6d (x%.2f)" % (name, t, i, t / i))
            for label, (a, b) in rows:
                print("      %-22s %6d / %6d  x%.2f" % (label, a, b, a / b))
