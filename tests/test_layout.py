"""CPU test: every byte-layout constant the product declares (vectorscan_b200/csrc/hsb_layout.h) equals the
sizeof/offsetof/enum value the reference's own headers give (oracle/_ref/layout_probe, built from /root/reference)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _kv(text):
    out = {}
    for line in text.splitlines():
        k, _, v = line.rpartition(" ")
        if k:
            out[k] = int(v)
    return out


def test_layout_matches_reference_headers(tmp_path):
    probe = os.path.join(ROOT, "oracle", "_ref", "layout_probe")
    if not os.path.exists(probe):
        pytest.skip("oracle/_ref/layout_probe not built")
    exe = str(tmp_path / "layout_dump")
    subprocess.run(["g++", "-std=c++17", "-I" + os.path.join(ROOT, "vectorscan_b200", "csrc"),
                    os.path.join(ROOT, "tests", "layout_dump.cpp"), "-o", exe], check=True)
    ours = _kv(subprocess.run([exe], capture_output=True, text=True, check=True).stdout)
    theirs = _kv(subprocess.run([probe], capture_output=True, text=True, check=True).stdout)
    assert len(ours) > 300
    missing = [k for k in ours if k not in theirs]
    assert not missing, f"keys the probe does not print: {missing[:5]}"
    wrong = {k: (ours[k], theirs[k]) for k in ours if ours[k] != theirs[k]}
    assert not wrong, wrong
