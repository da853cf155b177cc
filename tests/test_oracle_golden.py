"""CPU: pin the oracle to the reference's own known-answer tables
(tests/golden/*.json, transcribed by tests/golden/make_golden.py)."""
import json
import os

import pytest

from oracle import culling, gojson, nbc, odh, poddefault, profile

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return json.load(open(os.path.join(G, name)))


def test_golden_files_are_current():
    """The committed JSON equals what make_golden.py writes."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(G, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    for name, obj in (("nbc_status.json", mg.status_cases), ("culling.json", mg.culling),
                      ("odh.json", mg.odh), ("adm.json", mg.adm),
                      ("nbc_involved_object.json", mg.involved_cases), ("profile.json", mg.profile)):
        assert load(name) == json.loads(json.dumps(obj)), name


# ---- notebook-controller ---------------------------------------------------


@pytest.mark.parametrize("case", load("nbc_status.json"), ids=lambda c: c["name"])
def test_create_notebook_status(case):
    got = nbc.create_notebook_status(case["notebook"], case["sts"], case["pod"], now="1970-01-01T00:00:00Z")
    assert gojson.canon(got) == gojson.canon(case["expected"])


def test_nb_name_from_involved_object():
    g = load("nbc_involved_object.json")
    pods = {(p["metadata"]["namespace"], p["metadata"]["name"]): p for p in g["pods"]}
    for c in g["cases"]:
        name, err = nbc.nb_name_from_involved_object(lambda ns, n: pods.get((ns, n)), c["involvedObject"])
        assert err is None and name == c["expected"], c["name"]
    assert nbc.nb_name_from_involved_object(lambda ns, n: None, {"kind": "Node"})[1]


# ---- culling -----------------------------------------------------------------


def _resolve(meta, now_ns):
    if meta is None:
        return None
    meta = json.loads(json.dumps(meta))
    for k, v in (meta.get("annotations") or {}).items():
        if v.startswith("now"):
            meta["annotations"][k] = gojson.format_rfc3339(now_ns + int(v[3:] or 0) * 10**9)
    return meta


def test_culling_tables():
    g = load("culling.json")
    now_ns = gojson.parse_rfc3339(g["now"])[0]
    for c in g["set_stop_annotation"]["cases"]:
        m = _resolve(c["meta"], now_ns)
        culling.set_stop_annotation(m, now_ns)
        if m is not None:
            assert culling.STOP_ANNOTATION in m["annotations"], c["name"]
    for c in g["stop_annotation_is_set"]["cases"]:
        assert culling.stop_annotation_is_set(_resolve(c["meta"], now_ns)) == c["result"], c["name"]
    for c in g["all_kernels_are_idle"]["cases"]:
        assert culling.all_kernels_are_idle(c["kernels"]) == c["result"], c["name"]
    env = {}
    for c in g["notebook_is_idle"]["cases"]:
        env.update(c["env"])                 # os.Setenv persists across the Go subtests
        cfg = culling.init_global_vars(env)
        assert culling.notebook_is_idle(_resolve(c["meta"], now_ns), now_ns, cfg) == c["result"], c["name"]


def test_rfc3339():
    p = gojson.parse_rfc3339
    assert p("should-fail") is None and p("2021-08-30 15:37:36Z") is None
    assert p("2021-08-30T15:37:36.990063Z")[0] == 1630337856990063000
    assert p("1900-08-30T15:37:36.990063Z")[0] < 0
    assert p("2021-08-30T17:37:36+02:00")[0] == p("2021-08-30T15:37:36Z")[0]
    assert p("2021-02-30T00:00:00Z") is None and p("2024-02-29T00:00:00Z") is not None
    assert gojson.format_rfc3339(1630337856990063000) == "2021-08-30T15:37:36Z"
    assert gojson.format_rfc3339(p("2021-08-30T17:37:36+02:00")[0], 7200) == "2021-08-30T17:37:36+02:00"
    for s in ("0001-01-01T00:00:00Z", "9999-12-31T23:59:59Z", "1969-12-31T23:59:59Z"):
        assert gojson.format_rfc3339(p(s)[0]) == s


# ---- ODH ---------------------------------------------------------------------


def _labels_spec(o):
    return gojson.canon([(o.get("metadata") or {}).get("labels"), o.get("spec")])


def test_odh_objects():
    g = load("odh.json")
    r = g["route"]
    assert odh.compare_notebook_routes(odh.new_notebook_route(r["notebook"]), r["expected"])
    assert gojson.canon(odh.new_notebook_route(r["notebook"])["status"]) == gojson.canon(r["expected"]["status"])
    n = g["network_policies"]
    got = odh.new_notebook_network_policy(n["notebook"], g["controller_namespace"])
    assert odh.compare_notebook_network_policies(got, n["expected_ctrl"])
    assert got["metadata"] == n["expected_ctrl"]["metadata"]
    got = odh.new_oauth_network_policy(n["notebook"])
    assert odh.compare_notebook_network_policies(got, n["expected_oauth"])
    assert got["metadata"] == n["expected_oauth"]["metadata"]

    o = g["oauth"]
    mutated, reason = odh.mutate_notebook(o["notebook"], "CREATE")
    assert reason == ""
    assert odh.compare_notebooks(mutated, o["expected_notebook"]), \
        (gojson.canon(mutated), gojson.canon(o["expected_notebook"]))
    sa = odh.new_notebook_service_account(o["notebook"])
    assert sa["metadata"] == o["expected_service_account"]["metadata"]
    svc = odh.new_notebook_oauth_service(o["notebook"])
    assert svc["metadata"] == o["expected_service"]["metadata"]
    assert svc["spec"]["ports"] == o["expected_service"]["spec"]["ports"]
    sec = odh.new_notebook_oauth_secret(o["notebook"], odh.cookie_secret_from_seed(b"\x01" * 16))
    assert len(sec["stringData"]["cookie_secret"]) == o["expected_secret_cookie_len"]
    assert odh.compare_notebook_routes(odh.new_notebook_oauth_route(o["notebook"]), o["expected_route"])

    ref = nbc.set_controller_reference(r["notebook"], {}, "kubeflow.org/v1")["metadata"]["ownerReferences"][0]
    ref.pop("uid")
    assert ref == g["owner_reference"]["expected"]

    nb = json.loads(json.dumps(o["notebook"]))
    nb["metadata"]["annotations"].update(g["service_mesh_denied"]["annotations"])
    with pytest.raises(odh.Denied):
        odh.mutate_notebook(nb, "CREATE")

    d = g["diff_string"]
    s = odh.first_difference(d["a"], d["b"], "{v1.Pod}")
    assert s.endswith(": node1 != node2") and s.startswith("{v1.Pod}")   # shape of expected_go


def test_odh_update_blocking():
    """maybeRestartRunningNotebook (notebook_webhook.go:314-370) on a running notebook."""
    o = load("odh.json")["oauth"]
    created, _ = odh.mutate_notebook(o["notebook"], "CREATE")
    running = json.loads(json.dumps(created))
    del running["metadata"]["annotations"]["kubeflow-resource-stopped"]
    # same object sent again, same image: nothing changes, nothing pending
    m, reason = odh.mutate_notebook(running, "UPDATE", old=running)
    assert reason == "" and odh.compare_notebooks(m, running)
    # webhook now configured with a new proxy image: blocked, marked pending, spec kept
    m, reason = odh.mutate_notebook(running, "UPDATE", old=running, image="new-image")
    assert reason and m["metadata"]["annotations"][odh.ANNOTATION_UPDATE_PENDING] == reason
    assert gojson.canon(m["spec"]) == gojson.canon(running["spec"])
    # the user changes the pod template in the same update: everything goes through
    upd = json.loads(json.dumps(running))
    upd["spec"]["template"]["spec"]["containers"][0]["image"] = "other"
    m, reason = odh.mutate_notebook(upd, "UPDATE", old=running, image="new-image")
    assert reason == "" and m["spec"]["template"]["spec"]["containers"][1]["image"] == "new-image"


# ---- PodDefault --------------------------------------------------------------


def test_adm_tables():
    g = load("adm.json")
    for c in g["merge_map_bad"]["cases"]:
        with pytest.raises(poddefault.Conflict):
            poddefault.merge_map(c["existing"], [c["defaults"]])
    for c in g["merge_map_good"]["cases"]:
        assert poddefault.merge_map(c["existing"], [c["defaults"]]) == c["out"], c["name"]
    for c in g["apply_no_conflict"]["cases"]:
        pod = json.loads(json.dumps(c["pod"]))
        poddefault.safe_to_apply_pod_defaults_on_pod(pod, c["poddefaults"])
        poddefault.apply_pod_defaults_on_pod(pod, c["poddefaults"])
        assert "poddefault.admission.kubeflow.org/poddefault-" in pod["metadata"]["annotations"]
    for c in g["set_command_and_args"]["cases"]:
        ctr = json.loads(json.dumps(c["in"]))
        poddefault.set_command_and_args(ctr, c["poddefaults"])
        assert ctr == c["out"], c["name"]


def test_selector_semantics():
    m = poddefault.selector_matches
    assert m({}, {}) and m({}, {"a": "b"}) and m(None, {"a": "b"})
    assert m({"matchLabels": {"a": "b"}}, {"a": "b", "c": "d"})
    assert not m({"matchLabels": {"a": "b"}}, {"a": "x"}) and not m({"matchLabels": {"a": "b"}}, {})
    ex = lambda op, vals=None: {"matchExpressions": [{"key": "k", "operator": op, "values": vals or []}]}
    assert m(ex("In", ["v"]), {"k": "v"}) and not m(ex("In", ["v"]), {"k": "w"}) and not m(ex("In", ["v"]), {})
    assert m(ex("NotIn", ["v"]), {}) and m(ex("NotIn", ["v"]), {"k": "w"}) and not m(ex("NotIn", ["v"]), {"k": "v"})
    assert m(ex("Exists"), {"k": ""}) and not m(ex("Exists"), {})
    assert m(ex("DoesNotExist"), {}) and not m(ex("DoesNotExist"), {"k": "v"})
    for bad in (ex("In"), ex("Exists", ["v"]), ex("Bogus")):
        with pytest.raises(poddefault.SelectorError):
            m(bad, {"k": "v"})


def test_profile_namespace_labels():
    for c in load("profile.json")["set_namespace_labels"]["cases"]:
        assert profile.set_namespace_labels(c["current"], c["labels"]) == c["expected"]
