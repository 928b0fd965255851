"""Symmetric DUST (blastn's default query filter, "-dust 20 64 1") in the oracle: the definitional mask of
oracle/rcq_oracle.c (union of perfect intervals) against an independent restatement of the published
streaming algorithm (Morgulis et al. 2006, "A fast and symmetric DUST implementation"; window deque,
running pair counts, list of perfect intervals), plus the properties the name promises."""
import numpy as np
import pytest

from oracle import oracle as O

LET = np.frombuffer(b"ACGT", dtype=np.uint8)
CODE = np.full(256, 4, dtype=np.uint8)
for _k, _ch in enumerate(b"ACGT"):
    CODE[_ch] = _k


def dust_of(ascii_):
    n = len(ascii_)
    s = O.Sample(np.asarray(ascii_, dtype=np.uint8), np.array([0, n], dtype=np.int64), np.array([1], dtype=np.int32),
                 np.array([1], dtype=np.int32))
    return s.dust()


def sdust_stream(codes, T=20, W=64):
    """The streaming algorithm: triplets enter a window of W - 2 triplets; rw / rv count equal-triplet pairs in
    the window / in its suffix of length L whose counts stay <= 2T/10; when rw/L exceeds T/10 the suffixes are
    scanned and those scoring above T/10 and no lower than every perfect interval they contain join the list P;
    intervals leaving the window are written out merged."""
    n = len(codes)
    out = np.zeros(n, dtype=np.uint8)
    res, w, P = [], [], []
    cv, cw = [0] * 64, [0] * 64
    rv = rw = L = 0
    run = t = 0

    def save(start):
        if not P or P[-1][0] >= start:
            return
        ps, pf = P[-1][0], P[-1][1]
        if res and ps <= res[-1][1]:
            res[-1] = (res[-1][0], max(res[-1][1], pf))
        else:
            res.append((ps, pf))
        i = len(P) - 1
        while i >= 0 and P[i][0] < start:
            i -= 1
        del P[i + 1:]

    for i in range(n + 1):
        b = int(codes[i]) if i < n else 4
        if b < 4:
            run += 1
            t = ((t << 2) | b) & 63
            if run < 3:
                continue
            start = max(run - W, 0) + (i + 1 - run)
            save(start)
            if len(w) >= W - 2:
                s = w.pop(0)
                cw[s] -= 1
                rw -= cw[s]
                if L > len(w):
                    L -= 1
                    cv[s] -= 1
                    rv -= cv[s]
            w.append(t)
            L += 1
            rw += cw[t]; cw[t] += 1
            rv += cv[t]; cv[t] += 1
            if cv[t] * 10 > 2 * T:
                while True:
                    s = w[len(w) - L]
                    cv[s] -= 1
                    rv -= cv[s]
                    L -= 1
                    if s == t:
                        break
            if rw * 10 > L * T:
                c, r, max_r, max_l = cv[:], rv, 0, 0
                for ii in range(len(w) - L - 1, -1, -1):
                    r += c[w[ii]]
                    c[w[ii]] += 1
                    new_r, new_l = r, len(w) - ii - 1
                    if new_r * 10 > T * new_l:
                        j = 0
                        while j < len(P) and P[j][0] >= ii + start:
                            if max_r == 0 or P[j][2] * max_l > max_r * P[j][3]:
                                max_r, max_l = P[j][2], P[j][3]
                            j += 1
                        if max_r == 0 or new_r * max_l >= max_r * new_l:
                            max_r, max_l = new_r, new_l
                            P.insert(j, (ii + start, len(w) + 2 + start, new_r, new_l))
        else:
            start = max(run - W + 1, 0) + (i + 1 - run)
            while P:
                save(start)
                start += 1
            run = t = rv = rw = L = 0
            w, cv, cw = [], [0] * 64, [0] * 64
    for a, b in res:
        out[a:b] = 1
    return out


def low_complexity_mix(rng, n):
    parts = []
    while sum(map(len, parts)) < n:
        k = rng.integers(0, 6)
        if k == 0:
            parts.append(rng.integers(0, 4, size=rng.integers(20, 200)))
        elif k == 1:
            parts.append(np.full(rng.integers(5, 40), rng.integers(0, 4)))
        elif k == 2:
            parts.append(np.tile(rng.integers(0, 4, size=2), rng.integers(4, 30)))
        elif k == 3:
            parts.append(np.tile(rng.integers(0, 4, size=3), rng.integers(3, 20)))
        elif k == 4:
            parts.append(rng.choice([0, 3], size=rng.integers(10, 80)))
        else:
            parts.append(np.tile(rng.integers(0, 4, size=rng.integers(4, 9)), rng.integers(2, 10)))
    return np.concatenate(parts)[:n].astype(np.uint8)


def test_definition_equals_the_streaming_algorithm():
    rng = np.random.default_rng(0)
    masked = 0
    for it in range(150):
        n = int(rng.integers(3, 500))
        codes = low_complexity_mix(rng, n) if it % 3 else rng.integers(0, 4, size=n).astype(np.uint8)
        asc = LET[codes].copy()
        if it % 7 == 0 and n > 20:
            asc[rng.integers(0, n, size=3)] = ord("N")
        got = dust_of(asc)
        assert np.array_equal(got, sdust_stream(CODE[asc])), asc.tobytes()
        masked += int(got.sum())
    assert masked > 5000


def test_known_cases():
    assert dust_of(np.frombuffer(b"A" * 6, dtype=np.uint8)).sum() == 0            # 4 equal triplets: 6/3 = 2, not above
    assert dust_of(np.frombuffer(b"A" * 7, dtype=np.uint8)).all()                 # 5 equal triplets: 10/4 = 2.5
    rng = np.random.default_rng(1)
    body = LET[rng.integers(0, 4, size=300)]
    seq = np.concatenate([body, np.frombuffer(b"A" * 40, dtype=np.uint8), body[::-1].copy(), np.frombuffer(b"CA" * 20, dtype=np.uint8)])
    d = dust_of(seq)
    assert d[300:340].all() and d[640:680].all()                                   # the poly-A tail and the (CA)n repeat
    assert d[:290].mean() < 0.1                                                    # random sequence stays open


def test_symmetric_under_reverse_complement():
    rng = np.random.default_rng(2)
    comp = np.zeros(256, dtype=np.uint8)
    for x, y in zip(b"ACGTN", b"TGCAN"):
        comp[x] = y
    for _ in range(40):
        asc = LET[low_complexity_mix(rng, int(rng.integers(30, 400)))]
        assert np.array_equal(dust_of(asc)[::-1], dust_of(comp[asc[::-1]].copy()))


def test_masked_tails_do_not_seed():
    """Transcripts that share nothing but a poly-A tail and a microsatellite produce no seed; with the filter
    off every tail seeds against every other one."""
    rng = np.random.default_rng(3)
    n = 40

    def sample():
        seqs = [np.concatenate([LET[rng.integers(0, 4, size=200)], np.frombuffer(b"A" * 45, dtype=np.uint8),
                                LET[rng.integers(0, 4, size=50)], np.frombuffer(b"CA" * 22, dtype=np.uint8)]) for _ in range(n)]
        offs = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
        return np.concatenate(seqs), offs, np.arange(n, dtype=np.int32), np.ones(n, dtype=np.int32)

    a, b = sample(), sample()
    r = O.search_pair(O.Sample(*a), O.Sample(*b))
    assert r.counts["seeds"] == 0 and r.counts["rows"] == 0
    try:
        O.set_dust(False)
        r = O.search_pair(O.Sample(*a), O.Sample(*b))
        assert r.counts["seeds"] >= n * n
    finally:
        O.set_dust(True)
