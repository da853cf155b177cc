"""Shared parity checkers: library output (GPU, or the CPU build of the
planners) against the oracle, object by object, on canonical JSON."""
import json

from oracle import culling, gojson, nbc, odh, poddefault
from oracle import profile as oprofile


def check_nbc_generate(nbs, outs, status, env, flags, owner_api="kubeflow.org/v1beta1"):
    """nbs: Notebook trees; outs: 3 byte strings per Notebook; status: per Notebook."""
    assert len(outs) == 3 * len(nbs)
    n_checked = 0
    for i, nb in enumerate(nbs):
        gens = (nbc.generate_statefulset, nbc.generate_service, nbc.generate_virtualservice)
        exp, panic = [], False
        for k, g in enumerate(gens):
            if not flags & (1 << k):
                exp.append(None)
                continue
            try:
                e = g(nb, env)
            except nbc.ReferencePanic:
                panic = True
                break
            if flags & 8:
                nbc.set_controller_reference(nb, e, owner_api)
            exp.append(e)
        if panic:
            assert status[i] != 0, "doc %d: reference panics, library accepted it" % i
            assert all(len(outs[3 * i + k]) == 0 for k in range(3))
            continue
        assert status[i] == 0, "doc %d: status %d" % (i, status[i])
        for k in range(3):
            got = outs[3 * i + k]
            if exp[k] is None:
                assert got == b"", (i, k)
                continue
            tree = json.loads(got)
            assert gojson.canon(tree) == gojson.canon(exp[k]), \
                "doc %d object %d\n got %s\n exp %s" % (i, k, gojson.canon(tree), gojson.canon(exp[k]))
            n_checked += 1
    return n_checked


def check_odh_generate(nbs, outs, status, ctrl_ns, seeds, flags):
    """outs: 7 byte strings per Notebook (oracle.odh.ODH_SLOTS order)."""
    assert len(outs) == 7 * len(nbs)
    n_checked = 0
    for i, nb in enumerate(nbs):
        cookie = odh.cookie_secret_from_seed(seeds[16 * i:16 * i + 16] if seeds else b"\0" * 16)
        if not isinstance(nb, dict):
            continue
        exp = odh.reconcile_objects(nb, ctrl_ns, rbac=bool(flags & 32), cookie_secret=cookie,
                                    owner_ref=bool(flags & 8))
        assert status[i] == 0, (i, status[i])
        for k in range(7):
            got = outs[7 * i + k]
            if exp[k] is None:
                assert got == b"", (i, k, got)
                continue
            tree = json.loads(got)
            assert gojson.canon(tree) == gojson.canon(exp[k]), \
                "doc %d object %d\n got %s\n exp %s" % (i, k, gojson.canon(tree), gojson.canon(exp[k]))
            n_checked += 1
    return n_checked


def check_copy_fields(desired, found, outs, status, rejected=False):
    """desired/found: trees (found None = NotFound, desired None = nothing desired).
    rejected=True: a missing desired object means the generators rejected the Notebook (the
    reference panics), and the fused pass reports that on the pair (nbrm_nbc_reconcile)."""
    n_update = n_create = n_none = 0
    for i, (d, f) in enumerate(zip(desired, found)):
        if d is None:
            assert outs[i] == b"" and (status[i] != 0 if rejected else status[i] == 0), (i, status[i])
            continue
        try:
            exp = nbc.create_or_update(d, f)
        except nbc.ReferencePanic:
            assert status[i] == 5 and outs[i] == b"", i
            continue
        assert status[i] == 0, (i, status[i])
        if exp is None:
            assert outs[i] == b"", "pair %d: update not expected, got %s" % (i, outs[i])
            n_none += 1
            continue
        assert outs[i] != b"", "pair %d: expected %s" % (i, gojson.canon(exp))
        tree = json.loads(outs[i])
        assert gojson.canon(tree) == gojson.canon(exp), \
            "pair %d\n got %s\n exp %s" % (i, gojson.canon(tree), gojson.canon(exp))
        if f is None:
            n_create += 1
        else:
            n_update += 1
    return n_create, n_update, n_none


def check_odh_decide(desired, found, outs, status, result):
    """desired/found: trees (found None = NotFound, desired None = object not built)."""
    from oracle import odh
    n_create = n_update = n_none = 0
    for i, (d, f) in enumerate(zip(desired, found)):
        if d is None:
            assert outs[i] == b"" and status[i] == 0 and result[i] == 0, i
            continue
        what, exp = odh.write_decision(d, f)
        assert status[i] == 0, (i, status[i])
        assert result[i] == {None: 0, "create": 1, "update": 2}[what], (i, what, result[i])
        if what is None:
            assert outs[i] == b"", "pair %d: nothing expected, got %s" % (i, outs[i])
            n_none += 1
            continue
        tree = json.loads(outs[i])
        assert gojson.canon(tree) == gojson.canon(exp), \
            "pair %d (%s)\n got %s\n exp %s" % (i, what, gojson.canon(tree), gojson.canon(exp))
        if what == "create":
            n_create += 1
        else:
            n_update += 1
    return n_create, n_update, n_none


def check_status(nbs, stss, pods, outs, status, now_str):
    n = 0
    for i, (nb, sts, pod) in enumerate(zip(nbs, stss, pods)):
        assert status[i] == 0, (i, status[i])
        exp = nbc.create_notebook_status(nb, sts or {}, pod or {}, now_str)
        got = json.loads(outs[i])
        assert gojson.canon(got) == gojson.canon(exp), \
            "case %d\n got %s\n exp %s" % (i, gojson.canon(got), gojson.canon(exp))
        n += 1
    return n


def check_cull(cases, outs, status, result, now, idle_min, period_min):
    """cases: (notebook, pod_exists, kernels, terminals) tuples."""
    import collections
    seen = collections.Counter()
    cfg = {"CULL_IDLE_TIME": idle_min, "IDLENESS_CHECK_PERIOD": period_min}
    for i, (nb, pod_exists, kernels, terminals) in enumerate(cases):
        exp = json.loads(json.dumps(nb))
        outcome, updates, _ = culling.culling_step(exp, pod_exists, kernels, terminals,
                                                   now * 10**9, cfg)
        assert status[i] == 0, (i, status[i])
        assert int(result[i]) == outcome | (updates << 8), (i, int(result[i]), outcome, updates)
        if updates == 0:
            assert outs[i] == b"", i
        else:
            got = json.loads(outs[i])
            assert gojson.canon(got) == gojson.canon(exp), \
                "case %d outcome %d\n got %s\n exp %s" % (i, outcome, gojson.canon(got), gojson.canon(exp))
        seen[outcome] += 1
    return seen


PENDING = "notebooks.opendatahub.io/update-pending"


def _mask_pending(tree):
    a = (tree.get("metadata") or {}).get("annotations")
    if isinstance(a, dict) and PENDING in a:
        a[PENDING] = "<reason>"       # the reference's text is a go-cmp diff: presence is what is compared
    return tree


def _cert_env_as_set(tree, nb_name):
    """InjectCertConfig appends its env vars in Go map order (random in the
    reference): compare the five names as a set, at the tail of the list."""
    for c in (((tree.get("spec") or {}).get("template") or {}).get("spec") or {}).get("containers") or []:
        env = c.get("env")
        if c.get("name") == nb_name and isinstance(env, list):
            tail = [e for e in env if e.get("name") in odh.CERT_ENV_VARS and e.get("value") == odh.CERT_MOUNT_PATH]
            head = [e for e in env if e not in tail]
            c["env"] = head + sorted(tail, key=lambda e: e["name"])
    return tree


def check_mutate(requests, olds, outs, status, result, image, ca_bundle):
    """requests: Notebook trees; olds: None (CREATE) or trees (UPDATE)."""
    n_denied = n_blocked = n_ok = 0
    for i, nb in enumerate(requests):
        old = olds[i] if olds is not None else None
        try:
            exp, reason = odh.mutate_notebook(nb, "CREATE" if olds is None else "UPDATE", old=old,
                                              image=image, ca_bundle=ca_bundle)
        except odh.Denied:
            assert status[i] == 6 and outs[i] == b"", (i, status[i])
            n_denied += 1
            continue
        except (KeyError, TypeError, AttributeError):
            assert status[i] != 0, i          # a shape the Go decoder would have rejected
            continue
        assert status[i] == 0, (i, status[i])
        got = json.loads(outs[i])
        name = (nb.get("metadata") or {}).get("name")
        a = gojson.canon(_cert_env_as_set(_mask_pending(got), name))
        b = gojson.canon(_cert_env_as_set(_mask_pending(exp), name))
        assert a == b, "request %d\n got %s\n exp %s" % (i, a, b)
        assert bool(result[i]) == bool(reason), (i, result[i], reason)
        n_blocked += bool(reason)
        n_ok += 1
    return n_ok, n_denied, n_blocked


def check_poddefault(pods, pds, ranges, outs, status, result):
    """Patches compared by effect: apply(patch, pod) == the oracle's mutated pod."""
    import collections
    seen = collections.Counter()
    for i, pod in enumerate(pods):
        lst = pds[ranges[i][0]:ranges[i][1]] if ranges is not None else pds
        allowed, mutated, msg = poddefault.mutate_pod(pod, lst)
        if not allowed:
            exp_status = 7 if msg.startswith("conflict") else 9
            assert status[i] == exp_status and outs[i] == b"", (i, status[i], msg)
            seen["conflict" if exp_status == 7 else "selector"] += 1
            continue
        assert status[i] == 0, (i, status[i])
        if mutated is None:
            assert outs[i] == b"", (i, outs[i])
            seen["none"] += 1
            continue
        assert outs[i] != b"", i
        patch = json.loads(outs[i])
        got = poddefault.apply_json_patch(pod, patch)
        assert gojson.canon(got) == gojson.canon(mutated), \
            "pod %d\n got %s\n exp %s" % (i, gojson.canon(got), gojson.canon(mutated))
        n_applied = len([k for k in mutated["metadata"]["annotations"]
                         if k.startswith("poddefault.admission.kubeflow.org/poddefault-")])
        assert int(result[i]) >= 1
        seen["patched"] += 1
    return seen


def check_profile(profiles, outs, status, env, default_labels):
    n = 0
    for i, p in enumerate(profiles):
        exp = oprofile.profile_objects(p, env, default_labels)
        assert status[i] == 0, (i, status[i])
        for k in range(8):
            got = outs[8 * i + k]
            if exp[k] is None:
                assert got == b"", (i, k, got)
                continue
            assert gojson.canon(json.loads(got)) == gojson.canon(exp[k]), \
                "profile %d object %d\n got %s\n exp %s" % (i, k, gojson.canon(json.loads(got)), gojson.canon(exp[k]))
            n += 1
    return n
