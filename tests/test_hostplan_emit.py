"""The emit kernel's per-block routine (csrc/nbrm_emit.cuh) on the CPU: gather of one 16-byte
block of the output from the segments it spans, against a plain concatenation.  The warp-level
parts of emit_warp (the scan of the lengths, the ballot that counts a round's segments) are
restated serially in tests/hostplan; the GPU parity tests run the kernel itself."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    d = os.path.join(ROOT, "tests", "hostplan")
    subprocess.run(["make", "-C", d], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return C.CDLL(os.path.join(d, "libhostplan.so"))


def emit(lib, srcs, segs, head, win):
    n_src = len(srcs)
    arr = (C.c_char_p * n_src)(*srcs)
    lens = np.array([len(s) for s in srcs], dtype=np.uint64)
    so = np.array([s[1] for s in segs], dtype=np.uint32)
    sl = np.array([s[2] for s in segs], dtype=np.uint32)
    ss = np.array([s[0] for s in segs], dtype=np.uint32)
    total = int(sl.sum())
    out = bytearray(b"\xaa" * (head + total + 16))
    buf = (C.c_char * len(out)).from_buffer(out)
    rc = lib.hostplan_emit_blocks(arr, C.c_void_p(lens.ctypes.data), C.c_uint32(n_src),
                                  C.c_void_p(so.ctypes.data), C.c_void_p(sl.ctypes.data),
                                  C.c_void_p(ss.ctypes.data), C.c_uint32(len(segs)), C.c_uint32(head),
                                  C.c_uint32(win), buf)
    assert rc == 0
    return bytes(out), total


def check(lib, srcs, segs, head, win=128):
    got, total = emit(lib, srcs, segs, head, win)
    want = b"".join(srcs[k][o:o + n] for k, o, n in segs)
    assert got[:head] == b"\xaa" * head, "bytes before the object were touched"
    assert got[head + total:] == b"\xaa" * 16, "bytes after the object were touched"
    assert got[head:head + total] == want


def test_single_segment_every_alignment(lib):
    src = bytes(range(256)) * 3
    for head in range(16):
        for s_off in range(17):
            for n in (1, 2, 7, 8, 9, 15, 16, 17, 31, 32, 33, 100, 511, 512, 513):
                check(lib, [src], [(0, s_off, n)], head)


def test_short_segments(lib):
    rng = random.Random(5)
    srcs = [bytes(rng.randrange(256) for _ in range(4000)) for _ in range(5)]
    for trial in range(400):
        ns = rng.randrange(1, 60)
        segs = []
        for _ in range(ns):
            k = rng.randrange(5)
            n = rng.choice([1, 1, 2, 3, 5, 8, 13, 16, 17, 40])
            segs.append((k, rng.randrange(0, 4000 - n), n))
        check(lib, srcs, segs, rng.randrange(16))


def test_mixed_long_and_short(lib):
    rng = random.Random(6)
    srcs = [bytes(rng.randrange(256) for _ in range(20000)) for _ in range(3)]
    for trial in range(200):
        segs = []
        for _ in range(rng.randrange(1, 120)):
            k = rng.randrange(3)
            n = rng.choice([1, 4, 9, 16, 30, 64, 200, 512, 1500, 5000])
            segs.append((k, rng.randrange(0, 20000 - n), n))
        check(lib, srcs, segs, rng.randrange(16))


def test_more_segments_than_a_window(lib):
    rng = random.Random(7)
    srcs = [bytes(rng.randrange(256) for _ in range(3000))]
    for win in (1, 2, 5, 32, 128):
        for trial in range(30):
            segs = []
            for _ in range(rng.randrange(1, 700)):
                n = rng.choice([1, 2, 3, 10, 16, 33, 600])
                segs.append((0, rng.randrange(0, 3000 - n), n))
            check(lib, srcs, segs, rng.randrange(16), win)


def test_many_segments_start_in_one_round(lib):
    # more than 32 segment starts inside one 512-byte round: the count loop goes round twice
    src = bytes(range(256)) * 2
    segs = [(0, i % 200, 1 + (i % 3)) for i in range(300)]
    for head in (0, 5, 15):
        check(lib, [src], segs, head)
