"""Seeded synthetic workloads for BASELINE.json's configs (definitions: SURVEY.md section 8d).

  c1: 1000 literals, lengths 4..16 over [a-z], 20 % caseless; one buffer of space-separated pseudo-words
      (Zipf over a 50k-word vocabulary) with literals planted about once per 4 KiB.      [hs_scan]
  c2: 10 000 literals, lengths 3..12 over printable ASCII; batches of 1500-byte packets, payload = random
      printable bytes with HTTP-ish tokens, one planted literal in 10 % of the packets.  [hs_scan_vector shape]
Everything is numpy-vectorised so that 64 MiB .. 1 GiB corpora generate in seconds.
"""
import numpy as np

PRINTABLE = np.arange(32, 127, dtype=np.uint8)
LOWER = np.arange(97, 123, dtype=np.uint8)
HTTP_TOKENS = [b"GET /index.html HTTP/1.1\r\n", b"POST /api/v1/items HTTP/1.1\r\n", b"Host: www.example.com\r\n",
               b"User-Agent: Mozilla/5.0 (X11; Linux x86_64)\r\n", b"Accept: */*\r\n", b"Content-Length: 348\r\n",
               b"Content-Type: application/json\r\n", b"Connection: keep-alive\r\n", b"HTTP/1.1 200 OK\r\n",
               b"Cookie: session=", b"Accept-Encoding: gzip, deflate\r\n", b"\r\n\r\n"]


def _unique_literals(rng, n, lo, hi, alphabet):
    seen, out = set(), []
    while len(out) < n:
        ln = int(rng.integers(lo, hi + 1))
        s = bytes(alphabet[rng.integers(0, len(alphabet), ln)])
        if s not in seen:
            seen.add(s)
            out.append(s)
    return out


def c1_patterns(n=1000, seed=20260101):
    rng = np.random.default_rng(seed)
    lits = _unique_literals(rng, n, 4, 16, LOWER)
    flags = [1 if rng.random() < 0.2 else 0 for _ in lits]
    # caseless literals get a mixed-case spelling
    lits = [bytes((c - 32) if f and rng.random() < 0.5 else c for c in l) for l, f in zip(lits, flags)]
    return lits, flags, list(range(n))


def c1_text(nbytes, lits, seed=20260101):
    rng = np.random.default_rng(seed + 1)
    vocab_n = 50000
    wl = rng.integers(2, 11, vocab_n)
    starts = np.concatenate([[0], np.cumsum(wl + 1)])  # word + trailing space
    pool = LOWER[rng.integers(0, 26, int(starts[-1]))]
    pool[starts[1:] - 1] = 32
    ranks = np.arange(1, vocab_n + 1, dtype=np.float64)
    p = 1.0 / ranks
    p /= p.sum()
    out = np.empty(nbytes + 64, dtype=np.uint8)
    pos = 0
    while pos < nbytes:
        k = min(4_000_000, (nbytes - pos) // 4 + 16)
        w = rng.choice(vocab_n, size=k, p=p)
        lens = wl[w] + 1
        ends = np.cumsum(lens)
        total = int(ends[-1])
        idx = np.arange(total) - np.repeat(ends - lens, lens) + np.repeat(starts[w], lens)
        chunk = pool[idx]
        take = min(total, nbytes + 64 - pos)
        out[pos:pos + take] = chunk[:take]
        pos += take
    out = out[:nbytes]
    nplant = max(1, nbytes // 4096)
    where = rng.integers(0, max(1, nbytes - 32), nplant)
    which = rng.integers(0, len(lits), nplant)
    for w_, k_ in zip(where, which):
        l = lits[k_]
        if w_ + len(l) <= nbytes:
            out[w_:w_ + len(l)] = np.frombuffer(l, dtype=np.uint8)
    return out


def c2_patterns(n=10000, seed=20260102):
    rng = np.random.default_rng(seed)
    lits = _unique_literals(rng, n, 3, 12, PRINTABLE)
    return lits, [0] * n, list(range(n))


def c2_batch(nbytes, lits, seed=20260102, packet=1500):
    """Returns (flat uint8 array, uint64 offsets[count+1])."""
    rng = np.random.default_rng(seed + 7)
    count = (nbytes + packet - 1) // packet
    flat = PRINTABLE[rng.integers(0, len(PRINTABLE), nbytes)]
    offsets = np.minimum(np.arange(count + 1, dtype=np.uint64) * packet, nbytes).astype(np.uint64)
    # HTTP-ish tokens: ~30 % of the payload, written at random packet-relative positions
    ntok = int(nbytes * 0.3 / 24)
    tok_idx = rng.integers(0, len(HTTP_TOKENS), ntok)
    tok_pos = rng.integers(0, max(1, nbytes - 64), ntok)
    for t, tok in enumerate(HTTP_TOKENS):
        sel = tok_pos[tok_idx == t]
        tb = np.frombuffer(tok, dtype=np.uint8)
        idx = sel[:, None] + np.arange(len(tb))[None, :]
        flat[idx] = tb[None, :]
    # planted literals: one in 10 % of the packets
    npl = count // 10
    pk = rng.choice(count, size=npl, replace=False)
    which = rng.integers(0, len(lits), npl)
    inner = rng.integers(0, packet - 16, npl)
    for p_, k_, o_ in zip(pk, which, inner):
        l = lits[k_]
        s = int(offsets[p_]) + int(o_)
        if s + len(l) <= int(offsets[p_ + 1]):
            flat[s:s + len(l)] = np.frombuffer(l, dtype=np.uint8)
    return flat, offsets


def sweep_patterns(n, seed):
    rng = np.random.default_rng(seed)
    return _unique_literals(rng, n, 4, 12, PRINTABLE), [0] * n, list(range(n))


def sweep_batch(nbytes, bufsize, lits, seed):
    flat, offsets = c2_batch(nbytes, lits, seed=seed, packet=bufsize)
    return flat, offsets


# ---- c3: Snort-like regex set over synthetic HTTP payloads (SURVEY.md 8d "C3 concrete input") ------------------
# No pattern compiler exists in this image, so the set is what ONE LimEx NFA can hold:
# templates of the C3 description are drawn until the next one would not fit 512 states (csrc/nfadb_writer.cpp).
_C3_WORDS = [b"admin", b"login", b"passwd", b"select", b"union", b"script", b"eval", b"cmd", b"shell", b"upload",
             b"config", b"backup", b"token", b"session", b"debug", b"root", b"etc", b"proc", b"wp-", b"cgi-bin"]
_C3_EXT = [b"php", b"asp", b"jsp", b"cgi", b"exe", b"dll", b"sh"]


def c3_patterns(seed=20260103, max_patterns=2000, state_budget=500, templates=(0, 1, 2, 3, 4, 5)):
    """(expressions, flags, ids): expressions in the LimEx writer's dialect.  `templates`: which of the six C3
    templates to draw from (the counted-gap templates 0 and 3 determinise badly; the DFA configuration leaves
    them to the NFA engine, as the reference compiler does when a DFA exceeds its state limit)."""
    rng = np.random.default_rng(seed)
    ex, fl = [], []
    used = 2
    while len(ex) < max_patterns:
        k = int(rng.integers(0, 6))
        if k not in templates:
            continue
        w = _C3_WORDS[int(rng.integers(0, len(_C3_WORDS)))]
        w2 = _C3_WORDS[int(rng.integers(0, len(_C3_WORDS)))]
        if k == 0:
            m, n = int(rng.integers(1, 4)), int(rng.integers(4, 9))
            e = w + b"[0-9a-f]{%d,%d}" % (m, n) + w2
            cost = len(w) + n + len(w2)
        elif k == 1:
            n = int(rng.integers(4, 13))
            e = w + b"\\s+[A-Za-z0-9_]{4,%d}=" % n
            cost = len(w) + 1 + n + 1
        elif k == 2:
            e = b"^(GET|POST) /[^ ]*\\." + _C3_EXT[int(rng.integers(0, len(_C3_EXT)))] + b"\\?"
            cost = 7 + 2 + 1 + 1 + 3 + 1
        elif k == 3:
            g = int(rng.integers(2, 9))
            e = w + b"=[^\\r\\n]{%d}" % g + w2
            cost = len(w) + 1 + g + len(w2)
        elif k == 4:
            e = b"/" + w + b"/[^ ?]*\\.\\./"
            cost = len(w) + 7
        else:
            e = b"(" + w + b"|" + w2 + b")[^&]*&" + _C3_WORDS[int(rng.integers(0, len(_C3_WORDS)))] + b"="
            cost = len(w) + len(w2) + 12
        if used + cost > state_budget:
            break
        used += cost
        ex.append(e)
        fl.append(1 if rng.random() < 0.3 else 0)  # HS_FLAG_CASELESS
    return ex, fl, list(range(len(ex)))


def c3_corpus(nbytes, seed=20260103):
    """Synthetic HTTP requests / responses, 200..1400 bytes each: (flat uint8 array, uint64 offsets)."""
    rng = np.random.default_rng(seed + 11)
    lens = []
    tot = 0
    while tot < nbytes:
        chunk = rng.integers(200, 1401, 65536)
        lens.append(chunk)
        tot += int(chunk.sum())
    lens = np.concatenate(lens)
    ends = np.cumsum(lens)
    count = int(np.searchsorted(ends, nbytes, side="left")) + 1
    offsets = np.concatenate([[0], np.minimum(ends[:count], nbytes)]).astype(np.uint64)
    # body: URL-ish / header-ish bytes
    alpha = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz0123456789/=&?._-%+ :;", dtype=np.uint8)
    flat = alpha[rng.integers(0, len(alpha), nbytes)]
    # line structure: CRLF about every 60 bytes
    nl = rng.integers(0, max(1, nbytes - 2), nbytes // 60)
    flat[nl] = 13
    flat[nl + 1] = 10
    # request lines at the start of 60 % of the buffers, status lines on the rest
    starts = offsets[:-1].astype(np.int64)
    heads = [b"GET /", b"POST /", b"HTTP/1.1 200 OK\r\n"]
    pick = rng.choice(3, size=len(starts), p=[0.4, 0.2, 0.4])
    for t, h in enumerate(heads):
        sel = starts[(pick == t) & (np.diff(offsets).astype(np.int64) > 64)]
        hb = np.frombuffer(h, dtype=np.uint8)
        flat[sel[:, None] + np.arange(len(hb))[None, :]] = hb[None, :]
    # header tokens and attack-ish words sprinkled through the payload (~25 % of the bytes)
    toks = HTTP_TOKENS + [b"/" + w + b"/" for w in _C3_WORDS] + [w + b"=" for w in _C3_WORDS] + \
           [b"." + e + b"?" for e in _C3_EXT] + [b"/../"]
    ntok = int(nbytes * 0.25 / 14)
    tok_idx = rng.integers(0, len(toks), ntok)
    tok_pos = rng.integers(8, max(9, nbytes - 64), ntok)
    for t, tok in enumerate(toks):
        sel = tok_pos[tok_idx == t]
        tb = np.frombuffer(tok, dtype=np.uint8)
        flat[sel[:, None] + np.arange(len(tb))[None, :]] = tb[None, :]
    return flat, offsets


# ---- c4: pcapscan-shaped packet corpus, literal part of the 20k mixed set (SURVEY.md 8d "C4 concrete input") ----
def c4_patterns(n=16000, seed=20260104):
    """The 16 000 literals of C4's 20 000 mixed patterns.  The 4 000 regexes need Rose's decomposition into many
    small engines (full-Rose database), which no writer in this image produces: DESIGN.md section 6."""
    rng = np.random.default_rng(seed)
    return _unique_literals(rng, n, 4, 12, PRINTABLE), [0] * n, list(range(n))


def c4_corpus(nbytes, lits, seed=20260104):
    """Packets with the length mix {64 B: 40 %, 576 B: 20 %, 1500 B: 40 %} (by packet count), payload as c2;
    one block-mode scan per packet, as tools/.. pcapscan.cc:305-316 does.  (flat uint8 array, uint64 offsets)."""
    rng = np.random.default_rng(seed + 13)
    mean = 0.4 * 64 + 0.2 * 576 + 0.4 * 1500
    n_guess = int(nbytes / mean * 1.05) + 16
    lens = rng.choice(np.array([64, 576, 1500]), size=n_guess, p=[0.4, 0.2, 0.4])
    ends = np.cumsum(lens)
    count = int(np.searchsorted(ends, nbytes, side="left")) + 1
    offsets = np.concatenate([[0], np.minimum(ends[:count], nbytes)]).astype(np.uint64)
    flat, _ = c2_batch(nbytes, lits, seed=seed, packet=1500)   # payload bytes + tokens + planted literals
    return flat, offsets
