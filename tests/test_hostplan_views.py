"""CPU: the views the notebook-controller generators record of the objects they plan
(kubeflow-rm_b200/csrc/nbrm_plan.cuh) are exactly the spans view_object finds by walking
the emitted text — for every field the compare stage asks a desired object for."""
import ctypes as C
import random

import numpy as np

from test_hostplan_nbc import ENVS

# fields plan_copy_fields reads from the desired object, per kind (ViewField order:
# root kind meta labels annotations spec replicas tlabels tpod selector ports subjects)
USED = {"StatefulSet": [0, 1, 3, 4, 6, 7, 8], "Service": [0, 1, 3, 4, 9, 10], "VirtualService": [0, 1, 3, 4, 5]}
KINDS = ("StatefulSet", "Service", "VirtualService")


def test_recorded_views_match_a_walk(hostplan, wl):
    lib = hostplan.lib
    lib.hostplan_views.restype = C.c_void_p
    n_checked = 0
    for envi, env in enumerate(ENVS[:2]):
        rng = random.Random(envi)
        docs = [wl.dumps(nb, rng) for nb in wl.corpus("fuzz", 1500, start=envi * 1500) + wl.corpus("c2", 20)]
        for flags in (7, 7 | 8, 7 | 8 | 16):
            outs, st = hostplan.nbc_generate(docs, env, flags)
            nf = C.c_uint32()
            ptr = lib.hostplan_views(C.byref(nf))
            F = nf.value
            rec = np.frombuffer((C.c_uint32 * (len(outs) * 2 * F)).from_address(ptr), dtype=np.uint32).reshape(len(outs), 2 * F).copy()
            walk = np.zeros(2 * F, dtype=np.uint32)
            for i, o in enumerate(outs):
                if not o:
                    continue
                lib.hostplan_view_object(C.c_char_p(o), C.c_uint32(len(o)), C.c_void_p(walk.ctypes.data))
                for f in USED[KINDS[i % 3]]:
                    got = (int(rec[i, f]), int(rec[i, F + f]))
                    want = (int(walk[f]), int(walk[F + f]))
                    if want[1] == 0:
                        assert got[1] == 0, (i, f, got, want, o)
                    else:
                        assert got == want, (i, KINDS[i % 3], f, got, want, o[got[0]:got[0] + got[1]], o[want[0]:want[0] + want[1]])
                n_checked += 1
    assert n_checked > 20000


def test_recorded_odh_views_match_a_walk(hostplan, wl):
    """The same for the seven ODH objects and what the write decision reads: root, kind,
    metadata.labels, spec, subjects."""
    lib = hostplan.lib
    lib.hostplan_views.restype = C.c_void_p
    rng = random.Random(3)
    docs = [wl.dumps(nb, rng) for nb in wl.corpus("fuzz", 1500, start=7) + wl.corpus("c2", 20)]
    n_checked = 0
    for flags in (8 | 16, 8 | 16 | 32, 0):
        outs, st = hostplan.odh_generate(docs, None, {"CONTROLLER_NAMESPACE": "opendatahub"}, flags)
        nf = C.c_uint32()
        ptr = lib.hostplan_views(C.byref(nf))
        F = nf.value
        rec = np.frombuffer((C.c_uint32 * (len(outs) * 2 * F)).from_address(ptr), dtype=np.uint32).reshape(len(outs), 2 * F).copy()
        walk = np.zeros(2 * F, dtype=np.uint32)
        for i, o in enumerate(outs):
            if not o:
                continue
            lib.hostplan_view_object(C.c_char_p(o), C.c_uint32(len(o)), C.c_void_p(walk.ctypes.data))
            kind = o[int(walk[1]):int(walk[1]) + int(walk[F + 1])]
            fields = [0, 1] + ([3, 5, 11] if kind in (b'"Route"', b'"NetworkPolicy"', b'"RoleBinding"') else [])
            for f in fields:
                got = (int(rec[i, f]), int(rec[i, F + f]))
                want = (int(walk[f]), int(walk[F + f]))
                if want[1] == 0:
                    assert got[1] == 0, (i, f, got, want, o)
                else:
                    assert got == want, (i, kind, f, got, want, o)
            n_checked += 1
    assert n_checked > 10000


def test_mutated_notebook_views(hostplan, wl):
    """The webhook records where metadata, labels, annotations, the pod spec, its containers and the
    first of them are in the Notebook it writes; the generators that read that Notebook build
    their view from those extents.  That view equals the one the full walk finds."""
    lib = hostplan.lib
    lib.hostplan_views.restype = C.c_void_p
    rng = random.Random(9)
    nbs = wl.corpus("fuzz", 2500, start=77) + wl.corpus("c2", 200) + wl.corpus("c1", 100) + wl.corpus("loadtest", 50)
    docs = [wl.dumps(nb, rng) for nb in nbs]
    n_short = n_full = 0
    for flags in (1, 1 | 2, 0, 2):      # MUT_CREATE, + MUT_CA_BUNDLE, update, update + CA bundle
        outs, st = hostplan.mutate(docs, {"USE_ISTIO": "true"}, flags)
        nf = C.c_uint32()
        ptr = lib.hostplan_views(C.byref(nf))
        F = nf.value
        rec = np.frombuffer((C.c_uint32 * (len(outs) * 2 * F)).from_address(ptr), dtype=np.uint32).reshape(len(outs), 2 * F).copy()
        for i, o in enumerate(outs):
            if not o:
                continue
            r = np.ascontiguousarray(rec[i])
            rc = lib.hostplan_check_notebook_view(C.c_char_p(o), C.c_uint32(len(o)), C.c_void_p(r.ctypes.data))
            assert rc in (0, 100), (rc, flags, o, list(r))
            n_short += rc == 100
            n_full += rc == 0
    assert n_short > 2000, (n_short, n_full)
