"""CPU model of rl_decode.cu's short-run expand (start-bit mask + PRMT selector table), checked against np.repeat.
Development aid: python tools/model_rld.py"""
import numpy as np

def entry_table():
    t = []
    for n in range(16):
        b1, b2, b3 = n & 1, (n >> 1) & 1, (n >> 2) & 1
        s1, s2, s3 = b1, b1 + b2, b1 + b2 + b3
        sel = 0 | (s1 << 4) | (s2 << 8) | (s3 << 12)
        t.append(sel | (bin(n).count("1") << 16))
    return t

ENT = entry_table()

def prmt(a, b, sel):
    by = [(a >> (8 * i)) & 255 for i in range(4)] + [(b >> (8 * i)) & 255 for i in range(4)]
    return sum(by[(sel >> (4 * i)) & 7] << (8 * i) for i in range(4))

def decode_tile(counts, vals, mis):
    counts = np.asarray(counts, dtype=np.int64)
    npairs = counts.size
    offs = np.concatenate([[0], np.cumsum(counts)])
    tot = int(offs[-1])
    nwords = (tot + 64) // 32 + 3
    bm = [0] * nwords
    for j in range(npairs):
        pos = int(offs[j])
        if pos:  # position 0 is not marked: rank(0) = 0
            bm[pos >> 5] |= 1 << (pos & 31)
    nblk = (tot + 15) // 16 + 1
    tbl = [0] * (nblk + 8)
    for b in range(nblk):
        x = min(16 * b, tot - 1)
        tbl[b] = int(np.searchsorted(offs, x, side="right") - 1)
    v = [int(x) for x in vals] + [0xEE] * 16
    out = {}
    nunits = (mis + tot + 63) >> 6
    for q in range(nunits):
        p = 64 * q - mis
        s = max(p, 0)
        blk = s >> 4
        B = blk << 4
        # starts in B+1 .. s
        w = (bm[B >> 5] >> (B & 31)) & 0xFFFFFFFF if (B & 31) == 0 else ((bm[B >> 5] >> 16) | (bm[(B >> 5) + 1] << 16)) & 0xFFFFFFFF
        d = s - B
        cur = tbl[blk] + bin(w & ((2 << d) - 2)).count("1")
        # the kernel's way: binary search over the 16-pair groups' first offsets, then start bits in (base, s]
        nthr = (npairs + 15) // 16
        base = [int(offs[min(16 * t, npairs)]) for t in range(256)] + [tot]
        lo_t, hi_t = 0, 256
        while hi_t - lo_t > 1:
            mid = (lo_t + hi_t) >> 1
            if base[mid] <= s:
                lo_t = mid
            else:
                hi_t = mid
        a, b = base[lo_t] + 1, s
        cnt = 0
        if a <= b:
            for wi in range(a >> 5, (b >> 5) + 1):
                m = bm[wi]
                if wi == (a >> 5):
                    m &= (0xFFFFFFFF << (a & 31)) & 0xFFFFFFFF
                if wi == (b >> 5):
                    m &= ((2 << (b & 31)) - 1) & 0xFFFFFFFF
                cnt += bin(m).count("1")
        assert 16 * lo_t + cnt == cur, (q, s, lo_t, cnt, cur)
        # 64 start bits for positions p+1 .. p+64
        if p >= 0:
            w0, sft = (p + 1) >> 5, (p + 1) & 31
            raw = bm[w0] | (bm[w0 + 1] << 32) | (bm[w0 + 2] << 64)
            bits = (raw >> sft) & ((1 << 64) - 1)
        else:
            k = -(p + 1)
            raw = bm[0] | (bm[1] << 32)
            bits = (raw << k) & ((1 << 64) - 1)
        for wq in range(16):
            n = (bits >> (4 * wq)) & 15
            e = ENT[n]
            a = cur & ~3
            lo = v[a] | (v[a + 1] << 8) | (v[a + 2] << 16) | (v[a + 3] << 24)
            hi = v[a + 4] | (v[a + 5] << 8) | (v[a + 6] << 16) | (v[a + 7] << 24)
            v4 = prmt(lo, hi, 0x3210 + 0x1111 * (cur & 3))
            word = prmt(v4, 0, e & 0xFFFF)
            for i in range(4):
                x = p + 4 * wq + i
                if 0 <= x < tot:
                    out[x] = (word >> (8 * i)) & 255
            if p + 4 * wq + 3 >= 0:  # the kernel gates the advance the same way: bits at positions <= 0 are zero anyway
                cur += e >> 16
    return np.array([out[x] for x in range(tot)], dtype=np.uint8)

def main():
    rng = np.random.default_rng(0)
    for trial in range(300):
        npairs = int(rng.integers(1, 400))
        mean = rng.choice([1.2, 2, 4, 8])
        counts = np.minimum(rng.geometric(1 / mean, size=npairs), 255)
        vals = rng.integers(0, 256, size=npairs).astype(np.uint8)
        mis = int(rng.integers(0, 32))
        want = np.repeat(vals, counts)
        got = decode_tile(counts, vals, mis)
        assert np.array_equal(got, want), (trial, npairs, mean, mis)
    print("model ok")

if __name__ == "__main__":
    main()
