"""A C program linked against BOTH libraries (tests/clink/link_both.c): each library's
release function frees only what that library allocated (include/nbrm.h nbrm_host_free =
pinned memory; include/nbrm_host.h nbrm_store_free_string = malloc'ed strings)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kubeflow-rm_b200")


def build(tmp_path):
    exe = os.path.join(str(tmp_path), "link_both")
    subprocess.run(["gcc", "-O1", "-Wall", "-Werror", os.path.join(ROOT, "tests", "clink", "link_both.c"), "-o", exe,
                    "-L" + PKG, "-lnbrm_host", "-lnbrm", "-Wl,-rpath," + PKG], check=True)
    return exe


def test_no_symbol_is_exported_by_both_libraries():
    def exported(lib):
        out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(PKG, lib)], capture_output=True, text=True,
                             check=True).stdout
        return {line.split()[-1] for line in out.splitlines() if " T " in line}
    both = exported("libnbrm.so") & exported("libnbrm_host.so")
    assert not {s for s in both if s.startswith("nbrm_")}, both


def test_store_half_links_and_runs(tmp_path):
    exe = build(tmp_path)
    r = subprocess.run([exe, "store"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "ok", (r.returncode, r.stdout, r.stderr)


@pytest.mark.gpu
def test_both_halves_on_a_gpu(tmp_path):
    exe = build(tmp_path)
    r = subprocess.run([exe, "all"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "ok", (r.returncode, r.stdout, r.stderr)
