"""GPU: the warp's index pass (nbrm_tape.cuh, build_tape_warp) against the scalar builder,
entry by entry — positions, kinds, depths and both links — on documents that exercise
what the pass has to get right: strings that hold structural characters, escaped quotes
and runs of backslashes across lane and chunk borders, whitespace, every alignment of the
document start, documents shorter than a lane's 16 bytes and longer than a warp's 512."""
import ctypes as C
import json
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CAP = 6000


@pytest.fixture(scope="module")
def scalar():
    import subprocess
    d = os.path.join(ROOT, "tests", "hostplan")
    subprocess.run(["make", "-C", d], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    lib = C.CDLL(os.path.join(d, "libhostplan_tape.so"))
    lib.hostplan_build_tape.restype = C.c_uint32

    def build(doc):
        out = np.zeros(CAP, dtype=np.uint64)
        fl = C.c_uint32(0)
        n = lib.hostplan_build_tape(C.c_char_p(doc), C.c_uint32(len(doc)), C.c_void_p(out.ctypes.data),
                                    C.c_uint32(CAP), C.byref(fl))
        return out[:n], fl.value
    return build


def nasty_strings(rng):
    pool = ['{', '}', '[', ']', ':', ',', '"', '\\', ' ', 'a', 'b', '\n', 'é', '\\\\', '\\"', '{"x":[1,2]}']
    return "".join(rng.choice(pool) for _ in range(rng.randrange(0, 40)))


def random_tree(rng, depth=0):
    r = rng.random()
    if depth > 6 or r < 0.3:
        return rng.choice([nasty_strings(rng), rng.randrange(-5, 10**6), True, None, "", 1.5])
    if r < 0.65:
        return {nasty_strings(rng) or "k%d" % i: random_tree(rng, depth + 1) for i in range(rng.randrange(0, 6))}
    return [random_tree(rng, depth + 1) for _ in range(rng.randrange(0, 6))]


def check(ctx, pkg, scalar, docs):
    ctx.upload(0, pkg.Docs(docs))
    ent, cnt, flg = ctx.tape_dump(0, len(docs), CAP)
    for i, d in enumerate(docs):
        exp, efl = scalar(d)
        assert cnt[i] == len(exp), (i, cnt[i], len(exp), d[:200])
        if len(exp):
            assert flg[i] == efl, (i, d[:200])
        got = ent[i, :cnt[i]]
        if not np.array_equal(got, exp):
            k = int(np.nonzero(got != exp)[0][0])
            raise AssertionError("doc %d entry %d: got %#x expected %#x\n%r" % (i, k, got[k], exp[k], d[:300]))


def test_random_trees(ctx, pkg, scalar):
    rng = random.Random(11)
    docs = []
    for i in range(1500):
        t = random_tree(rng)
        style = rng.randrange(3)
        if style == 0:
            s = json.dumps(t, separators=(",", ":"), ensure_ascii=False)
        elif style == 1:
            s = json.dumps(t, indent=rng.choice([1, 2, None]), ensure_ascii=bool(rng.randrange(2)))
        else:
            s = json.dumps(t)
        docs.append(s.encode())
    check(ctx, pkg, scalar, docs)


def test_alignment_and_borders(ctx, pkg, scalar):
    """Every start alignment (documents of every length mod 16 in front), quotes, escapes and
    backslash runs placed on lane (16 B) and chunk (512 B) borders."""
    docs = []
    for pad in range(0, 40):
        docs.append(b'{"p":"' + b"x" * pad + b'"}')
    for run in range(0, 40):
        for lead in (0, 5, 9, 15, 16, 17, 500, 509, 510, 511, 512, 513):
            s = b'{"k":"' + b"y" * lead + b"\\" * (2 * (run // 2)) + (b'\\"' if run & 1 else b"") + b'{,}:[","]","z":[1,{"q":2}]}'
            docs.append(s)
    docs += [b"", b"{}", b"[]", b"x", b'"', b"{", b"}", b'{"a":1', b'{"a":"\\', b"[" * 70 + b"]" * 70,
             b'{"a":' + b"[" * 63 + b"]" * 63 + b"}", b"[" * 62 + b"]" * 62, b"[1,2,3]" * 3, b'  {  "a" : [ 1 , 2 ] }  ']
    check(ctx, pkg, scalar, docs)


def test_workload_documents(ctx, pkg, wl, scalar):
    from oracle import nbc
    rng = random.Random(5)
    docs = []
    for nb in wl.corpus("fuzz", 400, start=3) + wl.corpus("c2", 20):
        docs.append(wl.dumps(nb, rng))
        try:
            docs.append(wl.dumps(nbc.generate_statefulset(nb, {}), rng))
        except nbc.ReferencePanic:
            pass
    big = wl.notebook_c2(7)
    big["metadata"]["annotations"]["blob"] = '{"x":[' + "1," * 20000 + '1]}'
    docs.append(wl.dumps(big))          # over the tape capacity in entries? no: the blob is a string
    docs.append(json.dumps({"a": [1] * 7000}).encode())   # 7001 commas: over CAP entries -> no tape
    check(ctx, pkg, scalar, docs)
