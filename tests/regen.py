"""TEST INFRASTRUCTURE: seeded generator of expressions in the NFA writer's dialect (csrc/nfadb_writer.cpp) and an
independent ground truth for them: every end offset of every non-empty match, found with Python's `re`
(all-ends semantics of hs_scan, src/hs_runtime.h:479; dedupe of identical (id, to) as src/report.h:54-117)."""
import re

HS_FLAG_CASELESS = 1
HS_FLAG_DOTALL = 2

_WORDS = [b"GET", b"POST", b"HTTP", b"host", b"php", b"admin", b"select", b"union", b"passwd", b"cmd", b"exe",
          b"login", b"user", b"id", b"script", b"alert", b"eval", b"base64", b"etc", b"bin", b"sh", b"www"]
_CLASSES = [(rb"[a-z]", b"qz"), (rb"[A-Za-z0-9_]", b"Q7_"), (rb"\d", b"07"), (rb"\w", b"k_"), (rb"\s", b" \t"),
            (rb"[^\r\n]", b"x;"), (rb"[^ ]", b"y-"), (rb".", b"./"), (rb"[0-9a-f]", b"3e"), (rb"[\x00-\x1f]", b"\x01\x1f"),
            (rb"\S", b"#s"), (rb"[^a-z]", b"Z5")]


def _lit(rng, lo=1, hi=6):
    if rng.random() < 0.5:
        w = rng.choice(_WORDS)
    else:
        w = bytes(rng.choice(b"abcdefgh/=?&. ") for _ in range(rng.randint(lo, hi)))
    return re.escape(w), w


def _quant(rng):
    """(quantifier, a repeat count it allows)"""
    r = rng.random()
    if r < 0.25:
        return b"*", rng.randint(0, 3)
    if r < 0.5:
        return b"+", rng.randint(1, 3)
    if r < 0.6:
        return b"?", rng.randint(0, 1)
    m = rng.randint(0, 4)
    if r < 0.75:
        return b"{%d}" % (m + 1), m + 1
    if r < 0.85:
        return b"{%d,}" % m, m + rng.randint(0, 2)
    k = rng.randint(0, 5)
    return b"{%d,%d}" % (m, m + k), m + rng.randint(0, k)


def random_expr(rng, depth=0):
    """(expression, a string it matches).  Never matches the empty string (it starts with a literal)."""
    e, x = _lit(rng)
    parts, sample = [e], [x]
    for _ in range(rng.randint(1, 3)):
        r = rng.random()
        if r < 0.4:
            cls, chars = rng.choice(_CLASSES)
            q, n = _quant(rng)
            parts.append(cls + q)
            sample.append(bytes(rng.choice(chars) for _ in range(n)))
        elif r < 0.6 and depth < 2:
            alts = [random_expr(rng, depth + 1) for _ in range(rng.randint(2, 3))]
            g = b"(" + b"|".join(a[0] for a in alts) + b")"
            reps = 1
            if rng.random() < 0.3:
                q, reps = rng.choice([(b"?", 0), (b"?", 1), (b"*", 2), (b"+", 1), (b"{1,2}", 2)])
                g += q
            parts.append(g)
            sample.append(b"".join(rng.choice(alts)[1] for _ in range(reps)))
        else:
            e, x = _lit(rng)
            parts.append(e)
            sample.append(x)
    e, x = _lit(rng, 1, 3)
    parts.append(e)
    sample.append(x)
    return b"".join(parts), b"".join(sample)


def random_set(rng, n, with_samples=False):
    exprs, flags, samples = [], [], []
    for _ in range(n):
        e, x = random_expr(rng)
        if rng.random() < 0.15:
            e = b"^" + e
        exprs.append(e)
        samples.append(x)
        f = 0
        if rng.random() < 0.3:
            f |= HS_FLAG_CASELESS
        if rng.random() < 0.3:
            f |= HS_FLAG_DOTALL
        flags.append(f)
    ids = list(range(100, 100 + n))
    return (exprs, flags, ids, samples) if with_samples else (exprs, flags, ids)


def random_text(rng, exprs, n, samples=None):
    """Text over a small alphabet with fragments of the expressions' literal words planted, and -- when given --
    strings known to match (an anchored pattern's sample goes to the front)."""
    alpha = b"abcdefgh/=?&. \r\n0123456789_AZ"
    t = bytearray(rng.choice(alpha) for _ in range(n))
    i = 0
    while i < n:
        w = rng.choice(_WORDS)
        if rng.random() < 0.5:
            w = w.swapcase()
        t[i:i + len(w)] = w
        i += len(w) + rng.randint(0, 12)
    if samples:
        i = 0
        for e, x in zip(exprs, samples):
            if e.startswith(b"^") and rng.random() < 0.5 and x:
                t[0:len(x)] = x
                i = len(x)
                break
        while i < n:
            x = rng.choice(samples)
            if rng.random() < 0.2:
                x = x.swapcase()
            t[i:i + len(x)] = x
            i += len(x) + rng.randint(0, 40)
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
        eod = body.endswith(b"\\z") and not body.endswith(b"\\\\z")   # writer dialect: match must end at end of data
        if eod:
            body = body[:-2]
        # all end offsets: reverse is awkward with re; instead test every (start) with a pattern that forces
        # each possible end via fullmatch on prefixes -- O(n^2) per pattern, so keep texts small.
        rx = re.compile(b"(?:" + body + b")", fl)
        starts = [0] if anchored else range(n)
        for s in starts:
            for t in ([n] if eod else range(s + 1, n + 1)):
                if t <= s or (i, t) in out:
                    continue
                if rx.fullmatch(text, s, t):
                    out.add((i, t))
    return sorted(out, key=lambda x: (x[1], x[0]))


def with_eod(rng, exprs, rate=0.5):
    """Append \\z (end-of-data anchor of the writer's dialect) to a random subset of the expressions."""
    return [e + b"\\z" if rng.random() < rate else e for e in exprs]
