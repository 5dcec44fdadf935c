"""Exact count of shared-memory bank conflicts of a row-form tile kernel (csrc/tile_kernel.cuh TileFFT, MODE_ROW) for a
given schedule and padding period -- the access pattern is fully determined at compile time, so it can be enumerated on
the CPU.  Calibrated against ncu (profiles/r2_ncu_full_mix.txt: n = 1000, E = 10, radices 10 10 10, C = 2, PER = 10:
9 395 260 load requests with 16 299 039 conflicts, 9 395 260 store requests with 8 176 947 conflicts over 67 109 CTAs).

  python scripts/analysis/smem_conflicts.py            # calibration against ncu + one padding period per schedule
  python scripts/analysis/smem_conflicts.py search     # unrestricted per-exchange search (what is possible)
  python scripts/analysis/smem_conflicts.py emit       # per-exchange layouts with an addressing-cost charge -> csrc/mix_layouts.h body

Model (fp32 complex = 8-byte accesses): a warp request is served as two half-warp requests; within a half-warp, threads that
touch different 8-byte words in the same one of 16 word-banks serialise; conflicts = wavefronts - (number of active halves)."""
import sys
from functools import lru_cache


def pad(o, per):
    return o if per == 0 else o + o // per


def conflicts(N, E, radices, C, per, esz=8):
    """(load requests, load conflicts, store requests, store conflicts) per CTA"""
    T = N // E
    LS = N if per == 0 else N + N // per
    nthreads = C * T
    banks = 128 // esz  # 8-byte words per 128-byte wavefront
    half = 16 if esz == 8 else 8  # threads served together
    NP = len(radices)
    ns = [1]
    for r in radices[:-1]:
        ns.append(ns[-1] * r)

    def count(addr_of):
        req = conf = 0
        for w0 in range(0, nthreads, 32):
            req += 1
            wf = halves = 0
            for h0 in range(w0, min(w0 + 32, nthreads), half):
                per_bank = {}
                for tid in range(h0, min(h0 + half, nthreads)):
                    a = addr_of(tid)
                    per_bank.setdefault(a % banks, set()).add(a)
                if per_bank:
                    halves += 1
                    wf += max(len(s) for s in per_bank.values())
            conf += wf - halves
        return req, conf

    lr = lc = sr = sc = 0
    for P in range(NP):
        R, NS = radices[P], ns[P]
        NB = E // R
        if P < NP - 1:  # stores of pass P
            for m in range(NB):
                for k in range(R):
                    def addr(tid, m=m, k=k):
                        c, t = divmod(tid, T)
                        j = t + m * T
                        jm = j % NS
                        o = (j - jm) * R + jm
                        return c * LS + pad(o + k * NS, per)
                    r, cf = count(addr)
                    sr += r
                    sc += cf
        if P >= 1:  # loads feeding pass P
            for q in range(E):
                def addr(tid, q=q):
                    c, t = divmod(tid, T)
                    return c * LS + pad(t + q * T, per)
                r, cf = count(addr)
                lr += r
                lc += cf
    return lr, lc, sr, sc


SCHEDULES = [  # (N, E, CROW, radices) as in csrc/mixtile.cu (fp32 row forms)
    (96, 24, 64, (12, 8)), (192, 24, 32, (24, 8)), (384, 24, 16, (8, 8, 6)), (768, 24, 8, (12, 8, 8)), (1536, 24, 4, (24, 8, 8)),
    (3072, 24, 2, (8, 8, 8, 6)), (6144, 24, 1, (12, 8, 8, 8)), (100, 10, 24, (10, 10)), (200, 20, 24, (20, 10)),
    (250, 10, 10, (10, 5, 5)), (400, 20, 12, (20, 20)), (500, 10, 5, (10, 10, 5)), (1000, 10, 2, (10, 10, 10)),
    (2000, 20, 2, (20, 10, 10)), (3000, 30, 2, (30, 10, 10)), (4000, 20, 1, (20, 20, 10)), (5000, 20, 1, (10, 10, 10, 5)),
    (81, 9, 28, (9, 9)), (243, 9, 9, (9, 9, 3)), (729, 9, 3, (9, 9, 9)), (2187, 9, 1, (9, 9, 9, 3)), (6561, 27, 1, (9, 9, 9, 9)),
    (125, 5, 10, (5, 5, 5)), (625, 5, 2, (5, 5, 5, 5)), (3125, 25, 2, (5, 5, 5, 5, 5)), (49, 7, 32, (7, 7)), (343, 7, 5, (7, 7, 7)),
    (2401, 7, 1, (7, 7, 7, 7)), (1080, 30, 7, (30, 6, 6)),
]


def current_per(radices):
    return radices[0] if radices[0] % 2 == 0 else 0


if __name__ == "__main__" and len(sys.argv) == 1:
    lr, lc, sr, sc = conflicts(1000, 10, (10, 10, 10), 2, 10)
    ctas = 67109
    print(f"calibration n=1000: loads {lr * ctas} requests {lc * ctas} conflicts (ncu 9395260 / 16299039); "
          f"stores {sr * ctas} / {sc * ctas} (ncu 9395260 / 8176947)")
    lr, lc, sr, sc = conflicts(1536, 24, (24, 8, 8), 4, 24)
    print(f"calibration n=1536: loads {lr * 21846} requests {lc * 21846} conflicts (ncu 8388864 / 5642337)")
    print("\nN E C radices | current PER: conflicts per request (loads, stores) | best PER")
    for N, E, C, rad in SCHEDULES:
        cur = current_per(rad)
        res = {}
        for per in sorted({0, 8, 16, 32, 64, cur, rad[0], 2 * rad[0], N // E} - {1}):
            if per and N // per > N // 4:
                continue
            lr, lc, sr, sc = conflicts(N, E, rad, C, per)
            res[per] = (lc + sc, lc / max(lr, 1), sc / max(sr, 1), lr + sr)
        best = min(res, key=lambda p: res[p][0])
        c = res[cur]
        b = res[best]
        print(f"{N} {E} {C} {rad} | PER={cur}: {c[1]:.2f} {c[2]:.2f} total {c[0]} of {c[3]} | best PER={best}: {b[1]:.2f} {b[2]:.2f} total {b[0]}"
              + "   " + " ".join(f"{p}:{res[p][0]}" for p in sorted(res)))


def exchange_conflicts(N, E, radices, C, X, P, a, LS, esz=8):
    """conflicts (stores of pass X + loads of pass X+1) per CTA when exchange X uses f(o) = o + (o // P) * a (P = 0: none)"""
    T = N // E
    nthreads = C * T
    banks = 128 // esz
    half = 16 if esz == 8 else 8
    ns = [1]
    for r in radices[:-1]:
        ns.append(ns[-1] * r)
    R, NS = radices[X], ns[X]
    NB = E // R

    def f(o):
        return o if P == 0 else o + (o // P) * a

    def count(addr_of):
        conf = 0
        for w0 in range(0, nthreads, 32):
            wf = halves = 0
            for h0 in range(w0, min(w0 + 32, nthreads), half):
                per_bank = {}
                for tid in range(h0, min(h0 + half, nthreads)):
                    ad = addr_of(tid)
                    per_bank.setdefault(ad % banks, set()).add(ad)
                halves += 1
                wf += max(len(s) for s in per_bank.values())
            conf += wf - halves
        return conf

    sc = lc = 0
    for m in range(NB):
        for k in range(R):
            def addr(tid, m=m, k=k):
                c, t = divmod(tid, T)
                j = t + m * T
                jm = j % NS
                return c * LS + f((j - jm) * R + jm + k * NS)
            sc += count(addr)
    for q in range(E):
        def addr(tid, q=q):
            c, t = divmod(tid, T)
            return c * LS + f(t + q * T)
        lc += count(addr)
    return sc, lc


def search(N, E, radices, C, esz=8, verbose=False):
    """best (P, a) per exchange; LS is fixed afterwards as the largest footprint (rounded so that LS - T is a multiple of the
    bank count when C > 1 would help is NOT attempted: LS = max footprint)"""
    T = N // E
    NP = len(radices)
    ns = [1]
    for r in radices[:-1]:
        ns.append(ns[-1] * r)
    cands = [(0, 0)]
    base = {8, 16, 32, 64, T, 2 * T}
    for X in range(NP - 1):
        base |= {radices[X], ns[X] * radices[X], ns[X]} - {1}
    for P in sorted(base):
        if P < 4 or P > N:
            continue
        for a in (1, 2, 3, 4, 5, 6, 7, 8):
            if (N // P) * a <= N // 2:
                cands.append((P, a))
    # LS depends on the chosen layouts; iterate twice: first with a generous LS guess per candidate (its own footprint)
    out = []
    for X in range(NP - 1):
        best = None
        for P, a in cands:
            foot = N if P == 0 else N + ((N - 1) // P) * a
            for LS in sorted({foot, foot + (-(foot - T)) % 16 if C > 1 else foot}):
                sc, lc = exchange_conflicts(N, E, radices, C, X, P, a, LS, esz)
                key = (sc + lc, foot)
                if best is None or key < best[0]:
                    best = (key, P, a, LS, sc, lc)
        out.append(best)
    return out


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "search":
    for N, E, C, rad in SCHEDULES:
        cur = current_per(rad)
        tot_cur = 0
        LS = N if cur == 0 else N + N // cur
        for X in range(len(rad) - 1):
            sc, lc = exchange_conflicts(N, E, rad, C, X, cur, 1 if cur else 0, LS)
            tot_cur += sc + lc
        res = search(N, E, rad, C)
        tot_new = sum(b[0][0] for b in res)
        print(f"{N} E={E} C={C} {rad}: now {tot_cur} -> {tot_new}  " + " | ".join(f"X{X}: P={b[1]} a={b[2]} LS={b[3]} st={b[4]} ld={b[5]}" for X, b in enumerate(res)), flush=True)


# ---- search per exchange: conflicts + a charge for accesses the kernel cannot address as base + immediate ----
FALLBACK_COST = 1.5  # issue slots per warp access that recomputes its padded address (tile_kernel.cuh physx_k fallback)


def fallback_accesses(N, E, radices, C, X, P):
    """warp-level access instructions of exchange X whose address is not base + immediate with padding period P"""
    if P == 0:
        return 0
    T = N // E
    warps = (C * T + 31) // 32
    ns = [1]
    for r in radices[:-1]:
        ns.append(ns[-1] * r)
    R, NS = radices[X], ns[X]
    reads_ok = T % P == 0
    writes_ok = (NS % P == 0) or (P % (NS * R) == 0)
    return (0 if reads_ok else E * warps) + (0 if writes_ok else E * warps)


def candidates(N, E, radices, X):
    T = N // E
    ns = [1]
    for r in radices[:-1]:
        ns.append(ns[-1] * r)
    R, NS = radices[X], ns[X]
    Ps = {8, 16, 32, 64, 128, R, 2 * R, radices[0]}
    Ps |= {d for d in range(2, T + 1) if T % d == 0}
    Ps |= {d for d in range(2, NS + 1) if NS % d == 0}
    Ps |= {m * NS * R for m in range(1, N // (NS * R) + 1)}
    return sorted(P for P in Ps if 2 <= P <= N)


def search_cost(N, E, radices, C, esz):
    T = N // E
    NP = len(radices)
    banks = 128 // esz
    out = []
    for X in range(NP - 1):
        best = None
        for P in [0] + candidates(N, E, radices, X):
            for a in ((0,) if P == 0 else (1, 2, 3, 4, 5, 6, 7, 8)):
                foot = N if P == 0 else N + ((N - 1) // P) * a
                if foot > N + N // 4 + 16:
                    continue
                LS = foot + ((T - foot) % banks if C > 1 else 0)  # (LS - T) % banks == 0: a half-warp that straddles two lines stays conflict free
                sc, lc = exchange_conflicts(N, E, radices, C, X, P, a, LS, esz)
                cost = sc + lc + FALLBACK_COST * fallback_accesses(N, E, radices, C, X, P)
                key = (cost, LS)
                if best is None or key < best[0]:
                    best = (key, P, a, LS, sc, lc)
        out.append(best)
    return out


# searched layouts that measured SLOWER than the default rule on B200 (profiles/r2_mix_probe.md: the cost model overrates the
# price of a recomputed address for these): (element size, N, E) keep the default
KEEP_DEFAULT = {(8, 1536, 24), (8, 200, 20), (8, 500, 10), (8, 729, 9), (8, 2187, 9), (8, 1080, 30),
                (16, 1536, 24), (16, 4000, 20), (16, 500, 10), (16, 625, 5)}

ALL_ROW_SHAPES = SCHEDULES + [(5000, 10, 1, (10, 10, 10, 5)), (6561, 9, 1, (9, 9, 9, 9)), (625, 25, 10, (5, 5, 5, 5)), (3125, 5, 1, (5, 5, 5, 5, 5))]

if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "emit":
    lines = []
    for esz, guard in ((8, "#if defined(FFTW_SINGLE)"), (16, "#else")):
        lines.append(guard)
        for N, E, C, rad in ALL_ROW_SHAPES:
            if esz == 16 and N > 4096:
                continue
            cur = current_per(rad)
            cost_cur = conf_cur = 0
            LS0 = N if cur == 0 else N + N // cur
            for X in range(len(rad) - 1):
                sc, lc = exchange_conflicts(N, E, rad, C, X, cur, 1 if cur else 0, LS0, esz)
                conf_cur += sc + lc
                cost_cur += sc + lc + FALLBACK_COST * fallback_accesses(N, E, rad, C, X, cur)
            res = search_cost(N, E, rad, C, esz)
            cost_new = sum(b[0][0] for b in res)
            conf_new = sum(b[4] + b[5] for b in res)
            body = ", ".join("{%d, %d, %d}" % (b[1], b[2], b[3]) for b in res)
            if (esz, N, E) in KEEP_DEFAULT:
                lines.append(f"// {N} E={E}: the default layout stays (the searched one measured slower on B200, profiles/r2_mix_probe.md): B2_MIX_LAYOUT({N}, {E}, {len(res)}, {body})")
            elif cost_new < cost_cur:
                lines.append(f"B2_MIX_LAYOUT({N}, {E}, {len(res)}, {body})  // bank conflicts per CTA {conf_cur} -> {conf_new} (cost {cost_cur:.0f} -> {cost_new:.0f})")
            else:
                lines.append(f"// {N} E={E}: the default layout stays (conflicts per CTA {conf_cur}, cost {cost_cur:.0f}; best found {cost_new:.0f})")
            print(lines[-1], file=sys.stderr, flush=True)
    lines.append("#endif")
    print("\n".join(lines))


MORE_ROW_SHAPES = [  # csrc/mixtile2.cu (not individually measured on the GPU)
    (144, 12, 21, (12, 12)), (288, 24, 21, (24, 12)), (576, 24, 10, (24, 24)), (1152, 24, 5, (24, 8, 6)), (2304, 24, 2, (24, 12, 8)),
    (4608, 24, 1, (24, 24, 8)), (1296, 12, 2, (6, 6, 6, 6)), (1728, 24, 3, (12, 12, 12)), (216, 12, 14, (6, 6, 6)), (432, 24, 14, (24, 6, 3)),
    (864, 24, 7, (24, 12, 3)), (320, 20, 16, (20, 4, 4)), (640, 20, 8, (20, 4, 4, 2)), (1280, 20, 4, (20, 4, 4, 4)), (800, 20, 6, (20, 10, 4)),
    (1600, 20, 3, (20, 20, 4)), (3200, 20, 1, (20, 20, 4, 2)), (8000, 20, 1, (20, 20, 20)), (162, 18, 28, (6, 9, 3)), (486, 18, 9, (6, 9, 9)),
    (1458, 18, 3, (6, 9, 9, 3)), (150, 30, 50, (30, 5)), (300, 30, 25, (30, 10)), (450, 30, 17, (30, 15)), (600, 30, 12, (10, 10, 6)),
    (750, 30, 10, (30, 5, 5)), (900, 30, 8, (30, 30)), (1500, 30, 5, (30, 10, 5)), (1800, 30, 4, (30, 10, 6)), (2700, 30, 2, (30, 30, 3)),
    (4500, 30, 1, (30, 30, 5)), (6000, 30, 1, (30, 10, 10, 2)), (7500, 30, 1, (30, 10, 5, 5)), (1250, 10, 2, (10, 5, 5, 5)), (2500, 10, 1, (10, 10, 5, 5)),
    (6250, 10, 1, (10, 5, 5, 5, 5)), (360, 30, 21, (6, 6, 10)), (540, 30, 14, (30, 6, 3)), (720, 30, 10, (6, 6, 10, 2)), (196, 14, 18, (14, 14)),
    (392, 14, 9, (14, 14, 2)), (686, 14, 5, (14, 7, 7)), (1372, 14, 2, (14, 14, 7)), (2744, 14, 1, (14, 14, 14)), (448, 28, 16, (28, 4, 4)),
    (896, 28, 8, (28, 4, 4, 2)), (1792, 28, 4, (28, 4, 4, 4)), (441, 21, 12, (21, 21)), (121, 11, 23, (11, 11)), (1331, 11, 2, (11, 11, 11)),
    (169, 13, 19, (13, 13)), (2197, 13, 1, (13, 13, 13)),
]


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "emit_more":
    # Unmeasured shapes: a searched layout is adopted only where the measured first list says it pays -- the default rule
    # has at least 1.2 conflicts per request and the search removes at least 60 % of them.
    fp32_only = {4608, 8000, 4500, 6000, 7500, 6250}
    lines = []
    for esz, guard in ((8, "#if defined(FFTW_SINGLE)"), (16, "#else")):
        lines.append(guard)
        for N, E, C, rad in MORE_ROW_SHAPES:
            if esz == 16 and N in fp32_only:
                continue
            cur = current_per(rad)
            conf_cur = req = 0
            LS0 = N if cur == 0 else N + N // cur
            warps = (C * (N // E) + 31) // 32
            for X in range(len(rad) - 1):
                sc, lc = exchange_conflicts(N, E, rad, C, X, cur, 1 if cur else 0, LS0, esz)
                conf_cur += sc + lc
                req += 2 * E * warps
            if conf_cur < 1.2 * req:
                lines.append(f"// {N} E={E}: the default layout stays ({conf_cur} conflicts on {req} requests per CTA)")
            else:
                res = search_cost(N, E, rad, C, esz)
                conf_new = sum(b[4] + b[5] for b in res)
                body = ", ".join("{%d, %d, %d}" % (b[1], b[2], b[3]) for b in res)
                if conf_new <= 0.4 * conf_cur:
                    lines.append(f"B2_MIX_LAYOUT({N}, {E}, {len(res)}, {body})  // bank conflicts per CTA {conf_cur} -> {conf_new} on {req} requests")
                else:
                    lines.append(f"// {N} E={E}: the default layout stays ({conf_cur} conflicts on {req} requests per CTA; best found {conf_new})")
            print(lines[-1], file=sys.stderr, flush=True)
    lines.append("#endif")
    print("\n".join(lines))


E60_ROW_SHAPES = [  # csrc/mixtile2.cu, fp32 only
    (240, 60, 32, (20, 12)), (480, 60, 16, (30, 4, 4)), (960, 60, 8, (20, 12, 4)), (1920, 60, 4, (30, 4, 4, 4)), (3840, 60, 2, (20, 12, 4, 4)),
    (1200, 60, 6, (30, 10, 4)), (2400, 60, 3, (30, 20, 4)), (1440, 60, 5, (30, 12, 4)), (2880, 60, 2, (20, 12, 12)), (3600, 60, 2, (30, 30, 4)),
    (4800, 60, 1, (30, 10, 4, 4)), (5760, 60, 1, (20, 12, 12, 2)), (7200, 60, 1, (30, 20, 12)), (7680, 60, 1, (20, 12, 4, 4, 2)),
]

if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "emit_e60":
    lines = ["#if defined(FFTW_SINGLE)"]
    for N, E, C, rad in E60_ROW_SHAPES:
        cur = current_per(rad)
        conf_cur = req = 0
        LS0 = N if cur == 0 else N + N // cur
        warps = (C * (N // E) + 31) // 32
        for X in range(len(rad) - 1):
            sc, lc = exchange_conflicts(N, E, rad, C, X, cur, 1 if cur else 0, LS0, 8)
            conf_cur += sc + lc
            req += 2 * E * warps
        if conf_cur < 1.2 * req:
            lines.append(f"// {N} E={E}: the default layout stays ({conf_cur} conflicts on {req} requests per CTA)")
        else:
            res = search_cost(N, E, rad, C, 8)
            conf_new = sum(b[4] + b[5] for b in res)
            body = ", ".join("{%d, %d, %d}" % (b[1], b[2], b[3]) for b in res)
            if conf_new <= 0.4 * conf_cur:
                lines.append(f"B2_MIX_LAYOUT({N}, {E}, {len(res)}, {body})  // bank conflicts per CTA {conf_cur} -> {conf_new} on {req} requests")
            else:
                lines.append(f"// {N} E={E}: the default layout stays ({conf_cur} conflicts on {req} requests per CTA; best found {conf_new})")
        print(lines[-1], file=sys.stderr, flush=True)
    lines.append("#endif")
    print("\n".join(lines))
