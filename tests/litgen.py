"""Shared helpers for tests: seeded literal/text generators and the brute-force ground truth
(SURVEY.md section 8c "independent ground truth")."""
import random


def fold(b):
    """ASCII case fold as the reference does for caseless literals (clear 0x20 on alphabetic bytes only)."""
    return bytes((c | 0x20) if 65 <= c <= 90 else c for c in b)


def brute_force(lits, flags, ids, text):
    """Sorted unique (id, to) pairs: every occurrence of every literal, `to` one past the last byte,
    HS_FLAG_SINGLEMATCH ids keep only their earliest match."""
    hits = set()
    ftext = None
    for lit, f, i in zip(lits, flags, ids):
        if f & 1:
            if ftext is None:
                ftext = fold(text)
            t, l = ftext, fold(lit)
        else:
            t, l = text, lit
        pos = t.find(l)
        while pos != -1:
            hits.add((i, pos + len(l)))
            pos = t.find(l, pos + 1)
    single = {i for f, i in zip(flags, ids) if f & 8}
    out = {(i, to) for (i, to) in hits if i not in single}
    first = {}
    for i, to in hits:
        if i in single and (i not in first or to < first[i]):
            first[i] = to
    out |= set(first.items())
    return sorted(out)


def random_case(rng, lits, n, alpha, tl):
    lo = 1 if rng.random() < 0.3 else rng.choice([2, 3, 4])
    s = set()
    guard = 0
    while len(s) < n and guard < 20 * n + 100:
        guard += 1
        s.add(bytes(rng.choice(alpha) for _ in range(rng.randint(lo, rng.choice([4, 8, 12, 20])))))
    lits = sorted(s)
    flags = [(1 if rng.random() < 0.3 else 0) | (8 if rng.random() < 0.15 else 0) for _ in lits]
    ids = [rng.randrange(max(1, len(lits) // 2 + 1)) if rng.random() < 0.2 else k + 1000 for k in range(len(lits))]
    by_id = {}
    for k, (f, i) in enumerate(zip(flags, ids)):
        if i in by_id:
            flags[k] = (flags[k] & 1) | (by_id[i] & 8)
        else:
            by_id[i] = flags[k]
    text = bytearray(rng.choice(alpha) for _ in range(tl))
    for _ in range(tl // 50 + 1):
        l = rng.choice(lits)
        if rng.random() < 0.5:
            l = bytes((c ^ 0x20) if (65 <= c <= 90 or 97 <= c <= 122) and rng.random() < 0.5 else c for c in l)
        if len(l) <= tl:
            p = rng.randrange(0, tl - len(l) + 1)
            text[p:p + len(l)] = l
    if rng.random() < 0.2 and tl > 300:
        c = rng.choice(alpha)
        p = rng.randrange(0, tl - 260)
        text[p:p + 260] = bytes([c]) * 260
    return lits, flags, ids, bytes(text)
