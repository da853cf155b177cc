"""The 8-bytes-per-load container skip and span compare (device code of
csrc/nbrm_json.cuh) against the byte-at-a-time ones.

tests/hostplan/wide_fuzz.cc compiles the same source with g++ and drives it over random
JSON-like documents at every alignment, truncated documents included.  On the GPU the
same scanner is covered by the parity tests of the generators that use it (nbc, odh,
status, mutate).
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
HP = os.path.join(HERE, "hostplan")


def test_wide_scanners_match_scalar():
    subprocess.run(["make", "-s", "-C", HP, "wide_fuzz"], check=True)
    out = subprocess.run([os.path.join(HP, "wide_fuzz")], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.startswith("ok "), out.stdout
    assert int(out.stdout.split()[1]) > 100000
