"""Dev tool / design check: the segment-parallel interval coder as exact integer arithmetic (no GPU).

The arithmetic coder's `low` is an affine recurrence low' = (low << s) + a, so the final code value is
X = sum_i a_i << (T - S_i) (S_i = shifts up to and including word i, T = total shift); the byte stream the
reference writes (arith_codec.cpp:339-357, 524-546) is the bit string of X >> 8 on exactly T + 1 bits, then
the stop bit and zero padding.  The range recurrence is the only serial part; a segment that starts at an LPS
word depends on the range before it only through (range >> 5), i.e. 8 classes.  This script checks both
statements against the CPU oracle on random bin streams."""
import sys

import numpy as np

sys.path.insert(0, ".")
from oracle.checker import Checker, mk_ctx, mk_ep  # noqa: E402

MASK0, MASK1 = 0x7FE0, 0x7FFE


def resolve(cmds, qp, init_id, chk):
    ctx, rates = chk.ctx_init(qp, init_id)
    s0 = ctx["state0"].astype(np.int64).copy()
    s1 = ctx["state1"].astype(np.int64).copy()
    out = []
    for w in cmds:
        w = int(w)
        kind = w >> 29
        if kind == 0:
            c, b = w & 511, (w >> 9) & 1
            q = ((s0[c] + s1[c]) >> 8) & 0xFF
            mps = q >> 7
            qq = q ^ 0xFF if mps else q
            r0, r1 = int(rates[c]) >> 4, int(rates[c]) & 15
            s0[c] -= (s0[c] >> r0) & MASK0
            s1[c] -= (s1[c] >> r1) & MASK1
            if b:
                s0[c] += (0x7FFF >> r0) & MASK0
                s1[c] += (0x7FFF >> r1) & MASK1
            out.append(("ctx", int(qq >> 2), bool(b != mps)))
        elif kind == 1:
            out.append(("ep", (w >> 24) & 31, w & 0xFFFFFF))
    out.append(("trm",))
    return out


def step(word, r):
    """-> (shift, addend, new range) of one word at range r."""
    if word[0] == "ctx":
        _, k, is_lps = word
        lps = ((k * (r >> 5)) >> 1) + 4
        rmps = r - lps
        if is_lps:
            n = 0
            while (lps << n) < 256:
                n += 1
            return n, rmps << n, lps << n
        n = 1 if rmps < 256 else 0
        return n, 0, rmps << n
    if word[0] == "ep":
        return word[1], r * word[2], r
    r -= 2
    return 7, r << 7, 256  # encodeBinTrm(1)


def serial_stream(words):
    r, x, t = 510, 0, 0
    for w in words:
        s, a, r = step(w, r)
        x = (x << s) + a
        t += s
    return x, t


def to_bytes(x, t):
    bits = format(x >> 8, f"0{t + 1}b")
    assert len(bits) == t + 1
    bits += "1"
    bits += "0" * (-len(bits) % 8)
    return bytes(int(bits[i:i + 8], 2) for i in range(0, len(bits), 8))


def segmented_stream(words, seg=64):
    n = len(words)
    bounds = [0]
    for j in range(1, (n + seg - 1) // seg + 1):
        i = j * seg
        while i < n and not (words[i][0] == "ctx" and words[i][2]):
            i += 1
        bounds.append(min(i, n))
    bounds = sorted(set(bounds + [n]))
    # class maps: range after the segment for every class of the range before it
    maps = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        m = {}
        for cls in range(8, 16):
            r = cls << 5  # any representative of the class: the first word is an LPS word
            t = 0
            for w in words[a:b]:
                s, _, r = step(w, r)
                t += s
            m[cls] = (r, t)
        maps.append(m)
    # serial composition (segment 0 starts at the known range 510: simulated directly)
    r, t0 = 510, 0
    starts = []
    for j, (a, b) in enumerate(zip(bounds[:-1], bounds[1:])):
        starts.append((r, t0))
        if j == 0:
            for w in words[a:b]:
                s, _, r = step(w, r)
                t0 += s
        else:
            r, t = maps[j][r >> 5]
            t0 += t
    t_total = t0
    # every segment on its own from low = 0, then the sum of the shifted pieces
    x = 0
    for (a, b), (r, ts) in zip(zip(bounds[:-1], bounds[1:]), starts):
        v, t = 0, 0
        for w in words[a:b]:
            s, ad, r = step(w, r)
            v = (v << s) + ad
            t += s
        assert v < (1 << (t + 10))
        x += v << (t_total - ts - t)
    return x, t_total, len(bounds) - 1


if __name__ == "__main__":
    chk = Checker("oracle")
    rng = np.random.default_rng(5)
    for it in range(60):
        n = int(rng.integers(1, 3000))
        p1 = rng.random() * 0.5
        cmds = []
        for _ in range(n):
            if rng.random() < 0.75:
                cmds.append(mk_ctx(int(rng.integers(0, 40)), int(rng.random() < p1)))
            else:
                nb = int(rng.integers(0, 25))
                cmds.append(mk_ep(int(rng.integers(0, 1 << nb)) if nb else 0, nb))
        qp = int(rng.choice([0, 32, 51]))
        want = chk.encode_bins(np.asarray(cmds, np.uint32), qp, 2)[0].tobytes()
        words = resolve(cmds, qp, 2, chk)
        x, t = serial_stream(words)
        assert to_bytes(x, t) == want, (it, "serial characterisation")
        xs, ts, nseg = segmented_stream(words, seg=int(rng.choice([16, 64, 200])))
        assert (xs, ts) == (x, t), (it, "segmented sum")
    print("segment model ok")
