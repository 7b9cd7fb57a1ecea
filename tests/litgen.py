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


def look_admits(entry, c):
    """One lookaround entry (offset, and, cmp, negate) or (offset, class_bitmap32) against byte c."""
    if len(entry) == 2:
        return bool((entry[1][c >> 3] >> (c & 7)) & 1)
    return ((c & entry[1]) == entry[2]) != bool(entry[3])


def brute_force_ext(lits, flags, ids, ext, text, groups_off=0):
    """Ground truth for the extended literal form (include/hs_litdb.h hsb_lit_ext): a literal matches at `to` when its
    text does, every lookaround entry is inside the buffer and admits its byte, and its group is initially on; quiet
    ids deliver nothing; SINGLEMATCH ids keep their earliest match (a quiet SINGLEMATCH id has none to keep)."""
    hits = set()
    ftext = fold(text)
    for lit, f, i, e in zip(lits, flags, ids, ext):
        e = e or {}
        if (groups_off >> e.get("group", 0)) & 1:
            continue
        t, l = (ftext, fold(lit)) if f & 1 else (text, lit)
        pos = t.find(l)
        while pos != -1:
            to = pos + len(l)
            if all(to + le[0] >= 0 and look_admits(le, text[to + le[0]]) for le in e.get("look", [])):
                hits.add((i, to))
            pos = t.find(l, pos + 1)
    quiet = {i for i, e in zip(ids, ext) if e and e.get("flags", 0) & 4}
    single = {i for f, i in zip(flags, ids) if f & 8}
    out = {(i, to) for (i, to) in hits if i not in single and i not in quiet}
    first = {}
    for i, to in hits:
        if i in single and i not in quiet and (i not in first or to < first[i]):
            first[i] = to
    out |= set(first.items())
    return sorted(out)


def random_class(rng):
    """A byte class as a 32-byte bitmap: a few nibble-product blocks (so that it fits the shufti buckets)."""
    bm = bytearray(32)
    for _ in range(rng.randint(1, 2)):
        his = rng.sample(range(16), rng.randint(1, 4))
        los = rng.sample(range(16), rng.randint(1, 8))
        for h in his:
            for l in los:
                c = h * 16 + l
                bm[c >> 3] |= 1 << (c & 7)
    if rng.random() < 0.3:
        bm = bytearray(b ^ 0xff for b in bm)
    return bytes(bm)


def random_ext_case(rng, n, alpha, tl, kind):
    """Literal set + ext records exercising one instruction family.  kind: 'byte', 'mask', 'mask32', 'mask64',
    'shufti16', 'shufti32', 'shufti64', 'groups', 'quiet'."""
    lits, flags, ids, text = random_case(rng, None, n, alpha, tl)
    ids = [k + 1 for k in range(len(lits))]          # one report per literal keeps the ground truth simple
    flags = [f & 1 | (8 if rng.random() < 0.2 else 0) for f in flags]
    text = bytearray(text)
    ext = []
    for k, lit in enumerate(lits):
        e = {}
        if kind in ("byte", "mask", "mask32", "mask64", "shufti16", "shufti32", "shufti64") and rng.random() < 0.8:
            span = {"byte": 1, "mask": 8, "mask32": 32, "mask64": 64, "shufti16": 16, "shufti32": 32, "shufti64": 64}[kind]
            lo = -rng.randint(max(1, span // 2) if span > 8 else 1, span) if kind != "byte" else -rng.randint(1, 12)
            hi = min(-1, lo + span - 1)
            cnt = 1 if kind == "byte" else rng.randint(2, 4)
            offs = sorted(set([lo] + [rng.randint(lo, hi) for _ in range(cnt - 1)]))
            if kind != "byte" and len(offs) == 1:
                offs = sorted(set(offs + [min(-1, lo + 1), lo]))
                if len(offs) == 1:
                    offs = [lo - 1, lo]
            look = []
            for o in offs:
                if kind.startswith("shufti"):
                    look.append((o, random_class(rng)))
                else:
                    c = rng.choice(alpha)
                    am = rng.choice([0xff, 0xdf, 0xf0, 0x0f, 0x7f])
                    look.append((o, am, c & am, rng.random() < 0.25))
            e["look"] = look
        if kind == "groups":
            # small sets: a group per literal, so that SINGLEMATCH ones may squash it; large sets share six groups
            e["group"] = k if len(lits) <= 40 else rng.randrange(6)
            e["flags"] = (1 if rng.random() < 0.7 else 0) | (2 if rng.random() < 0.7 else 0)
            if rng.random() < 0.4:
                flags[k] |= 8
        if kind == "quiet" and rng.random() < 0.4:
            e["flags"] = 4
        ext.append(e)
    # plant occurrences whose lookaround is satisfied, so that both outcomes are seen
    for _ in range(tl // 40 + 1):
        k = rng.randrange(len(lits))
        l, e = lits[k], ext[k]
        need = max([len(l)] + [-le[0] for le in e.get("look", [])])
        if need > tl:
            continue
        to = rng.randrange(need, tl + 1)
        text[to - len(l):to] = l
        for le in e.get("look", []):
            if -le[0] > len(l) or rng.random() < 0.5:
                ok = [c for c in range(256) if look_admits(le, c)]
                if ok and -le[0] > len(l):
                    text[to + le[0]] = rng.choice(ok)
    return lits, flags, ids, ext, bytes(text)
