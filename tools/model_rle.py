"""CPU model of rl_encode.cu's per-tile arithmetic (rows of 128 bytes, warp-level and tile-level closed-form
255-splits, positions mod 256 staged instead of counts), checked against the oracle.  Development aid:
    python tools/model_rle.py
Mirrors the kernel's names; nothing here is on the product path."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as o

ROW = 128
ROWS = 128
TILE = ROW * ROWS
WARPS = 4
WSPAN = TILE // WARPS
BIAS = 255
NONE = TILE


def encode_model(data, n=None, prev_run=0):
    data = np.asarray(data, dtype=np.uint8)
    n_avail = data.size
    if n is None:
        n = n_avail
    if n == 0:
        return np.zeros(0, np.uint8), np.zeros(0, np.uint8)
    n_scan = min(n_avail, n + 255)
    tiles = n_scan // TILE + 1
    prev_phase = prev_run % 255
    has_prev = prev_run != 0
    counts = {}
    values = {}
    g = 0                       # scanner prefix
    sbB = (BIAS - prev_phase) if prev_phase else 0
    for tile in range(tiles):
        tbase = tile * TILE
        rel_emit = min(max(n - tbase, 0), TILE)
        rel_scan = min(n_scan - tbase, TILE)
        ragged = rel_emit < TILE or rel_scan < TILE
        # ---- step A per warp
        M = [[0, 0, 0, 0] for _ in range(ROWS)]   # four 32-bit words per row
        F = [NONE] * WARPS
        L = [-1] * WARPS
        C = [0] * WARPS
        Ce = [0] * WARPS
        excl = [0] * ROWS
        excl_e = [0] * ROWS
        for r in range(ROWS):
            for i in range(ROW):
                pos = tbase + ROW * r + i
                if pos >= n_avail:
                    break
                if pos == 0:
                    st = not has_prev
                else:
                    st = data[pos] != data[pos - 1]
                if st:
                    M[r][i >> 5] |= 1 << (i & 31)
            if ragged:
                for gq in range(4):
                    kk = rel_scan - (ROW * r + 32 * gq)
                    if kk < 0:
                        M[r][gq] = 0
                    elif kk < 32:
                        M[r][gq] = (M[r][gq] & ((1 << kk) - 1)) | (1 << kk)
        for w in range(WARPS):
            rows = range(32 * w, 32 * w + 32)
            nz = [any(M[r]) for r in rows]
            my_last = {}
            lead = {}
            for r in rows:
                m128 = M[r][0] | (M[r][1] << 32) | (M[r][2] << 64) | (M[r][3] << 96)
                lead[r] = (m128 & -m128).bit_length() - 1 if m128 else ROW
                my_last[r] = ROW * r + m128.bit_length() - 1 if m128 else None
            for lane, r in enumerate(rows):
                lower = [q for q in range(lane) if nz[q]]
                rowstart = ROW * r
                if lower and not (ragged and rowstart > rel_scan):
                    S = my_last[32 * w + lower[-1]]
                    if (rowstart - S) + lead[r] > 255:
                        rr = (rowstart - 1 - S) % 255
                        j = 254 - rr
                        if j < lead[r]:
                            M[r][j >> 5] |= 1 << (j & 31)
            run = 0
            rune = 0
            for r in rows:
                c = sum(bin(x).count("1") for x in M[r])
                ce = c
                if ragged:
                    ce = 0
                    for gq in range(4):
                        kk = rel_emit - (ROW * r + 32 * gq)
                        em = 0 if kk <= 0 else (0xFFFFFFFF if kk >= 32 else (1 << kk) - 1)
                        ce += bin(M[r][gq] & em).count("1")
                excl[r] = run
                excl_e[r] = rune
                run += c
                rune += ce
            C[w] = run
            Ce[w] = rune
            if any(nz):
                first_lane = nz.index(True)
                last_lane = 31 - nz[::-1].index(True)
                F[w] = ROW * (32 * w + first_lane) + lead[32 * w + first_lane]
                L[w] = my_last[32 * w + last_lane]
        # ---- step B (every warp computes the same)
        base = 0
        base_e = 0
        Srun = -1
        wb = [0] * WARPS
        wnl = [0] * WARPS
        wfirst = [0] * WARPS
        for i in range(WARPS):
            nl_i = 0
            nle_i = 0
            first_i = 0
            ws = WSPAN * i
            if Srun >= 0 and not (ragged and ws > rel_scan):
                end_i = min(F[i], ws + WSPAN)
                q0 = (ws - 1 - Srun) // 255
                q1 = (end_i - 1 - Srun) // 255
                nl_i = q1 - q0
                first_i = Srun + 255 * (q0 + 1)
                nle_i = nl_i
                if ragged:
                    end_e = min(end_i, rel_emit)
                    nle_i = ((end_e - 1 - Srun) // 255 - q0) if end_e > ws else 0
            wb[i] = base
            wnl[i] = nl_i
            wfirst[i] = first_i
            base += nl_i + C[i]
            base_e += nle_i + Ce[i]
            if L[i] >= 0:
                Srun = L[i]
        total_all = base
        total_emit = base_e
        F_t = min(F)
        L_t = max(L)
        # ---- compaction: staged (value, pos & 255) in boundary order
        sval = [0] * total_all
        spos = [0] * total_all
        for w in range(WARPS):
            for j in range(wnl[w]):
                p = wfirst[w] + 255 * j
                sval[wb[w] + j] = int(data[tbase + WSPAN * w]) if tbase + WSPAN * w < n_avail else 0
                spos[wb[w] + j] = p & 255
            for r in range(32 * w, 32 * w + 32):
                kk = wb[w] + wnl[w] + excl[r]
                for gq in range(4):
                    bits = M[r][gq]
                    while bits:
                        bi = (bits & -bits).bit_length() - 1
                        bits &= bits - 1
                        p = ROW * r + 32 * gq + bi
                        sval[kk] = int(data[tbase + p]) if tbase + p < n_avail else 0
                        spos[kk] = p & 255
                        kk += 1
        # ---- service warp (closed form over the run that enters the tile)
        tB = tbase + BIAS
        FB = tB + F_t
        q0 = (tB - 1 - sbB) // 255
        BpB = sbB + 255 * q0
        limB = min(FB, tB + rel_emit)
        nl_all = (FB - 1 - sbB) // 255 - q0
        nl = ((limB - 1 - sbB) // 255 - q0) if limB > tB else 0
        lagged = 255 if nl_all else FB - BpB
        tot = nl + total_emit
        lead_first = BpB + 255 - tB
        e_lagged = lagged if (tile > 0 and BpB < n + BIAS) else 0
        first_byte = int(data[tbase]) if tbase < n_avail else 0
        g0 = g
        g += tot
        if e_lagged and g0 >= 1:
            counts[g0 - 1] = e_lagged
        # ---- store_tile
        lead_lag = nl_all == nl and total_all == 0
        for j in range(nl):
            p = lead_first + 255 * j
            values[g0 + j] = first_byte
            if not (j == nl - 1 and lead_lag):
                counts[g0 + j] = min(255, F_t - p)
        for k in range(total_emit):
            values[g0 + nl + k] = sval[k]
        ncounts = total_emit if total_all > total_emit else total_emit - 1
        for k in range(max(ncounts, 0)):
            counts[g0 + nl + k] = (spos[k + 1] - spos[k]) & 255
        # ---- d1
        if L_t >= 0:
            sbB = tbase + L_t + BIAS
    P = g
    cs = np.array([counts.get(k, 0) for k in range(P)], dtype=np.uint8)
    vs = np.array([values.get(k, 0) for k in range(P)], dtype=np.uint8)
    return cs, vs


def runs(rng, n, mean, alphabet=256):
    lens = rng.geometric(1.0 / mean, size=int(n / mean * 1.2) + 64)
    while lens.sum() < n:
        lens = np.concatenate([lens, rng.geometric(1.0 / mean, size=lens.size)])
    vals = rng.integers(0, alphabet, size=lens.size).astype(np.uint8)
    return np.repeat(vals, lens)[:n].copy()


def check(data, tag=""):
    c_ref, v_ref = o.rl_compress(data)
    c, v = encode_model(data)
    assert c.size == c_ref.size, (tag, data.size, c.size, c_ref.size)
    if not np.array_equal(c, c_ref):
        bad = int(np.flatnonzero(c != c_ref)[0])
        raise AssertionError(f"{tag}: counts differ at pair {bad}: {c[bad:bad+4]} vs {c_ref[bad:bad+4]} (n={data.size})")
    assert np.array_equal(v, v_ref), tag


def main():
    rng = np.random.default_rng(0)
    for run in [1, 2, 254, 255, 256, 509, 510, 511, 765, 766, 4095, 4096, 4097, 16383, 16384, 16385, 16384 + 255, 100000]:
        check(np.frombuffer(bytes([9]) * run + b"\x01", dtype=np.uint8), f"run{run}a")
        check(np.frombuffer(b"\x07\x07\x03" + bytes([9]) * run, dtype=np.uint8), f"run{run}b")
    for n in [1, 2, 15, 16, 17, 127, 128, 129, 255, 256, 257, 4095, 4096, 4097, 16383, 16384, 16385, 16384 + 256, 16384 * 2 - 1,
              16384 * 3 + 77, 65537, 100001]:
        for mean in [1.3, 4, 64, 700, 3000]:
            check(runs(np.random.default_rng(n + int(mean * 10)), n, mean), f"n{n}m{mean}")
            check(runs(np.random.default_rng(n + int(mean * 10)), n, mean, alphabet=2), f"n{n}m{mean}a2")
    for n in (16384 * 5, 16384 * 3 + 3, 255 * 512, 255 * 512 + 1):
        check(np.zeros(n, dtype=np.uint8), f"zeros{n}")
    d = np.zeros(16384 * 5 + 100, dtype=np.uint8)
    d[16384 * 2 + 17] = 1
    check(d, "interrupt")
    # shards
    for mean in [3, 500, 5000]:
        rng = np.random.default_rng(mean)
        data = runs(rng, 120_000, mean, alphabet=3)
        c_ref, v_ref = o.rl_compress(data)
        cuts = [0, 1, 77, 16384, 50_001, 90_000, 119_999, data.size]
        cs, vs = [], []
        for a, b in zip(cuts[:-1], cuts[1:]):
            prev = 0
            while a - prev - 1 >= 0 and data[a - prev - 1] == data[a]:
                prev += 1
            ahead = min(data.size, b + 255)
            c, v = encode_model(data[a:ahead], n=b - a, prev_run=prev)
            cs.append(c)
            vs.append(v)
        assert np.array_equal(np.concatenate(cs), c_ref), mean
        assert np.array_equal(np.concatenate(vs), v_ref), mean
    print("model ok")


if __name__ == "__main__":
    main()
