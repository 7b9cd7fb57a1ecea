"""tools/hsbench_gpu.py: the hsbench-shaped command line (pattern files in the reference's ID:/EXPR/FLAGS format, its
SQLite corpus layout).  CPU: the loaders; GPU: one run end to end, match count against the reference runtime."""
import importlib.util
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("hsbench_gpu", os.path.join(ROOT, "tools", "hsbench_gpu.py"))
cli = importlib.util.module_from_spec(spec)
spec.loader.exec_module(cli)

PATTERNS = b"""# comment
1:/needle/
2:/HayStack/i
3:/abc[0-9]{2,4}x/
7:/(login|admin)[^&]*&id=/i
9:/once/H
"""


def make_files(tmp_path):
    pf = tmp_path / "pats.txt"
    pf.write_bytes(PATTERNS)
    blocks = [b"a needle in a haystack once once", b"abc123x login=1&id=2", b"", b"ADMIN&ID= abc1x abc12345x needle"] * 50
    cf = tmp_path / "corpus.db"
    cli.write_corpus(str(cf), blocks)
    return pf, cf, blocks


def test_loaders(tmp_path):
    pf, cf, blocks = make_files(tmp_path)
    lits, regs = cli.read_patterns(str(pf))
    assert [p for p, _, _ in lits] == [1, 2, 9] and [p for p, _, _ in regs] == [3, 7]
    assert lits[1][2] == 1 and lits[2][2] == 8 and regs[1][2] == 1
    assert cli.read_corpus(str(cf), 1500) == blocks
    plain = tmp_path / "plain.bin"
    plain.write_bytes(b"x" * 4000)
    assert [len(b) for b in cli.read_corpus(str(plain), 1500)] == [1500, 1500, 1000]
    bad = tmp_path / "bad.txt"
    bad.write_bytes(b"1:/abc/8\n")
    with pytest.raises(SystemExit):
        cli.read_patterns(str(bad))


@pytest.mark.gpu
def test_cli_end_to_end(tmp_path, ref):
    pf, cf, blocks = make_files(tmp_path)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "hsbench_gpu.py"), "-e", str(pf), "-c", str(cf), "-n", "3", "-N"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    line = [l for l in r.stdout.splitlines() if l.startswith("Matches per iteration")][0]
    got = int(line.split(":")[1].split("(")[0].replace(",", ""))
    # the same database on the reference runtime
    from vectorscan_b200 import litdb
    lits, regs = cli.read_patterns(str(pf))
    db = litdb.compile_lit_multi([e for _, e, _ in lits], [f for _, _, f in lits], [p for p, _, _ in lits],
                                 regexes=[e for _, e, _ in regs], regex_flags=[f for _, _, f in regs], regex_ids=[p for p, _, _ in regs])
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    exp = sum(len(set(ref.scan(rdb, rs, b)[1])) for b in blocks)
    assert got == exp and exp > 300
    assert "Mean throughput (overall)" in r.stdout


def test_corpus_loader_keeps_stream_ids(tmp_path):
    cf = tmp_path / "streams.db"
    blocks = [b"ab", b"cd", b"ef", b"gh"]
    cli.write_corpus(str(cf), blocks, stream_ids=[5, 9, 5, 9])
    assert cli.read_corpus(str(cf), 1500, with_streams=True) == [(5, b"ab"), (9, b"cd"), (5, b"ef"), (9, b"gh")]
    assert cli.read_corpus(str(cf), 1500) == blocks


@pytest.mark.gpu
def test_cli_vectored_mode(tmp_path, ref):
    """-V: the blocks of one stream_id are one hs_scan_vector call on an HS_MODE_VECTORED database."""
    pf = tmp_path / "lits.txt"
    pf.write_bytes(b"1:/needle/\n2:/HayStack/i\n9:/once/H\n")
    blocks = [b"a nee", b"x hay", b"dle once on", b"STACK", b"ce once", b" needle"] * 20
    sids = [k % 2 + 10 * (k // 6) for k in range(len(blocks))]        # two interleaved streams per group of six blocks
    cf = tmp_path / "corpus.db"
    cli.write_corpus(str(cf), blocks, stream_ids=sids)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "hsbench_gpu.py"), "-e", str(pf), "-c", str(cf), "-n", "2", "-V"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    line = [l for l in r.stdout.splitlines() if l.startswith("Matches per iteration")][0]
    got = int(line.split(":")[1].split("(")[0].replace(",", ""))
    from vectorscan_b200 import litdb
    lits, _ = cli.read_patterns(str(pf))
    db = litdb.compile_lit_multi([e for _, e, _ in lits], [f for _, _, f in lits], [p for p, _, _ in lits], mode=litdb.HS_MODE_VECTORED)
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    streams = {}
    for sid, b in zip(sids, blocks):
        streams.setdefault(sid, []).append(b)
    exp = sum(len(ref.scan_vector(rdb, rs, bl)[1]) for bl in streams.values())
    assert got == exp and exp >= 60
    assert "vectored" in r.stdout
