"""TEST INFRASTRUCTURE: seeded generator of expressions in the NFA writer's dialect (csrc/nfadb_writer.cpp) and an
independent ground truth for them: every end offset of every non-empty match, found with Python's `re`
(all-ends semantics of hs_scan, src/hs_runtime.h:479; dedupe of identical (id, to) as src/report.h:54-117)."""
import re

HS_FLAG_CASELESS = 1
HS_FLAG_DOTALL = 2

_WORDS = [b"GET", b"POST", b"HTTP", b"host", b"php", b"admin", b"select", b"union", b"passwd", b"cmd", b"exe",
          b"login", b"user", b"id", b"script", b"alert", b"eval", b"base64", b"etc", b"bin", b"sh", b"www"]
_CLASSES = [rb"[a-z]", rb"[A-Za-z0-9_]", rb"\d", rb"\w", rb"\s", rb"[^\r\n]", rb"[^ ]", rb".", rb"[0-9a-f]",
            rb"[\x00-\x1f]", rb"\S", rb"[^a-z]"]


def _lit(rng, lo=1, hi=6):
    if rng.random() < 0.5:
        return re.escape(rng.choice(_WORDS))
    n = rng.randint(lo, hi)
    return re.escape(bytes(rng.choice(b"abcdefgh/=?&. ") for _ in range(n)))


def _quant(rng):
    r = rng.random()
    if r < 0.25:
        return b"*"
    if r < 0.5:
        return b"+"
    if r < 0.6:
        return b"?"
    m = rng.randint(0, 4)
    if r < 0.75:
        return b"{%d}" % (m + 1)
    if r < 0.85:
        return b"{%d,}" % m
    return b"{%d,%d}" % (m, m + rng.randint(0, 5))


def random_expr(rng, depth=0):
    """One expression that cannot match the empty string (it always starts with a literal)."""
    parts = [_lit(rng)]
    for _ in range(rng.randint(1, 3)):
        r = rng.random()
        if r < 0.4:
            parts.append(rng.choice(_CLASSES) + _quant(rng))
        elif r < 0.6 and depth < 2:
            alts = [random_expr(rng, depth + 1) for _ in range(rng.randint(2, 3))]
            g = b"(" + b"|".join(alts) + b")"
            if rng.random() < 0.3:
                g += rng.choice([b"?", b"*", b"+", b"{1,2}"])
            parts.append(g)
        else:
            parts.append(_lit(rng))
    parts.append(_lit(rng, 1, 3))
    e = b"".join(parts)
    if depth == 0 and rng.random() < 0.15:
        e = b"^" + e
    return e


def random_set(rng, n):
    exprs, flags = [], []
    for _ in range(n):
        exprs.append(random_expr(rng))
        f = 0
        if rng.random() < 0.3:
            f |= HS_FLAG_CASELESS
        if rng.random() < 0.3:
            f |= HS_FLAG_DOTALL
        flags.append(f)
    return exprs, flags, list(range(100, 100 + n))


def random_text(rng, exprs, n):
    """Text over a small alphabet with fragments of the expressions' literal words planted."""
    alpha = b"abcdefgh/=?&. \r\n0123456789_AZ"
    t = bytearray(rng.choice(alpha) for _ in range(n))
    i = 0
    while i < n:
        w = rng.choice(_WORDS)
        if rng.random() < 0.5:
            w = w.swapcase()
        t[i:i + len(w)] = w
        i += len(w) + rng.randint(0, 12)
    return bytes(t[:n])


def brute_force(exprs, flags, ids, text):
    """[(id, to)] sorted by (to, id): to = end offset (exclusive) of any non-empty match of expr anywhere."""
    out = set()
    n = len(text)
    for e, f, i in zip(exprs, flags, ids):
        fl = 0
        if f & HS_FLAG_CASELESS:
            fl |= re.IGNORECASE
        if f & HS_FLAG_DOTALL:
            fl |= re.DOTALL
        anchored = e.startswith(b"^")
        body = e[1:] if anchored else e
        # all end offsets: reverse is awkward with re; instead test every (start) with a pattern that forces
        # each possible end via fullmatch on prefixes -- O(n^2) per pattern, so keep texts small.
        rx = re.compile(b"(?:" + body + b")", fl)
        starts = [0] if anchored else range(n)
        for s in starts:
            for t in range(s + 1, n + 1):
                if (i, t) in out:
                    continue
                if rx.fullmatch(text, s, t):
                    out.add((i, t))
    return sorted(out, key=lambda x: (x[1], x[0]))
