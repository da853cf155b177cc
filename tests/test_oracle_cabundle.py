"""The trusted-CA bundle oracle (oracle/cabundle.py) against what the reference's own tests hold:
the two ConfigMap fixtures of odh-notebook-controller/controllers/notebook_controller_test.go
(:262-263 and :366-367, four distinct ed25519 certificates) must give a bundle in which
checkCertConfigMap (:1134-1159) counts 2 certificates (:315, :420)."""
import copy

import pytest

from oracle import cabundle

PEM = {
    "a": "-----BEGIN CERTIFICATE-----\nMIGrMF+gAwIBAgIBATAFBgMrZXAwADAeFw0yNDExMTMyMzI3MzdaFw0yNTExMTMy\nMzI3MzdaMAAwKjAFBgMrZXADIQDEMMlJ1P0gyxEV7A8PgpNosvKZgE4ttDDpu/w9\n35BHzjAFBgMrZXADQQDHT8ulalOcI6P5lGpoRcwLzpa4S/5pyqtbqw2zuj7dIJPI\ndNb1AkbARd82zc9bF+7yDkCNmLIHSlDORUYgTNEL\n-----END CERTIFICATE-----",
    "b": "-----BEGIN CERTIFICATE-----\nMIGrMF+gAwIBAgIBATAFBgMrZXAwADAeFw0yNDExMTMyMzI2NTlaFw0yNTExMTMy\nMzI2NTlaMAAwKjAFBgMrZXADIQB/v02zcoIIcuan/8bd7cvrBuCGTuVZBrYr1RdA\n0k58yzAFBgMrZXADQQBKsL1tkpOZ6NW+zEX3mD7bhmhxtODQHnANMXEXs0aljWrm\nAxDrLdmzsRRYFYxe23OdXhWqPs8SfO8EZWEvXoME\n-----END CERTIFICATE-----",
    "c": "-----BEGIN CERTIFICATE-----\nMIGrMF+gAwIBAgIBATAFBgMrZXAwADAeFw0yNDExMTMyMzI4MjZaFw0yNTExMTMy\nMzI4MjZaMAAwKjAFBgMrZXADIQD77pLvWIX0WmlkYthRZ79oIf7qrGO7yECf668T\nSB42vTAFBgMrZXADQQDs76j81LPh+lgnnf4L0ROUqB66YiBx9SyDTjm83Ya4KC+2\nLEP6Mw1//X2DX89f1chy7RxCpFS3eXb7U/p+GPwA\n-----END CERTIFICATE-----",
    "d": "-----BEGIN CERTIFICATE-----\nMIGrMF+gAwIBAgIBATAFBgMrZXAwADAeFw0yNDExMTMyMzI4NDJaFw0yNTExMTMy\nMzI4NDJaMAAwKjAFBgMrZXADIQAw01381TUVSxaCvjQckcw3RTcg+bsVMgNZU8eF\nXa/f3jAFBgMrZXADQQBeJZHSiMOYqa/tXUrQTfNIcklHuvieGyBRVSrX3bVUV2uM\nDBkZLsZt65rCk1A8NG+xkA6j3eIMAA9vBKJ0ht8F\n-----END CERTIFICATE-----",
}
CASES = [  # (odh-trusted-ca-bundle data or None, kube-root-ca.crt data or None, certificates expected / None = no ConfigMap)
    ({"ca-bundle.crt": PEM["a"], "odh-ca-bundle.crt": PEM["b"]}, None, 2),                       # test :255-315
    ({"ca-bundle.crt": PEM["c"], "odh-ca-bundle.crt": PEM["d"]}, None, 2),                       # test :360-420
    ({"ca-bundle.crt": PEM["a"] + "\n", "odh-ca-bundle.crt": "\n"}, {"ca.crt": PEM["d"]}, 2),    # RHOAIENG-15743: trailing newlines
    (None, {"ca.crt": PEM["a"]}, None),                                                          # :277-279
    ({"odh-ca-bundle.crt": PEM["b"]}, None, None),                                               # no ca-bundle.crt key :293
    ({"ca-bundle.crt": "  \n", "odh-ca-bundle.crt": PEM["b"]}, None, None),                      # empty ca-bundle.crt :293
    ({"ca-bundle.crt": PEM["a"]}, None, None),                                                   # odh-ca-bundle.crt key missing: !ok -> return nil
    ({"ca-bundle.crt": "not a pem", "odh-ca-bundle.crt": ""}, {"ca.crt": PEM["c"] + "\n" + PEM["d"]}, 2),  # whole file goes in
    ({"ca-bundle.crt": PEM["a"].replace("CERTIFICATE", "PRIVATE KEY"), "odh-ca-bundle.crt": ""}, {}, None),
    ({"ca-bundle.crt": PEM["a"][:120] + PEM["a"][124:], "odh-ca-bundle.crt": PEM["b"]}, None, 1),  # first block does not parse
]


def bundle_cases():
    return copy.deepcopy(CASES)


@pytest.mark.parametrize("case", range(len(CASES)))
def test_bundle(case):
    odh_cm, kube_cm, want = CASES[case]
    b = cabundle.cert_bundle(odh_cm, kube_cm)
    if want is None:
        assert b is None
    else:
        assert cabundle.count_certificates(b) == want
        cm = cabundle.trusted_ca_configmap("ns", b)
        assert cm["metadata"]["labels"] == {"opendatahub.io/managed-by": "workbenches"}


def notebook_with_certs(name="nb", dup_last=False):
    env = [{"name": "A", "value": "1"}] + [{"name": k, "value": "/etc/pki/tls/custom-certs/ca-bundle.crt"} for k in cabundle.ENV_VARS] + \
          [{"name": "Z", "value": "2"}]
    if dup_last:
        env.append({"name": "PIP_CERT", "value": "again"})
    return {"metadata": {"name": name, "namespace": "ns"},
            "spec": {"template": {"spec": {
                "containers": [{"name": "side", "env": [{"name": "PIP_CERT", "value": "kept"}]},
                               {"name": name, "env": env,
                                "volumeMounts": [{"name": "data", "mountPath": "/d"}, {"name": "trusted-ca", "mountPath": "/etc/pki"}]}],
                "volumes": [{"name": "data"}, {"name": "trusted-ca", "configMap": {"name": "workbench-trusted-ca-bundle"}}]}}}}


def test_unset_cert_config_and_its_loop_quirk():
    nb = notebook_with_certs()
    assert cabundle.is_configmap_deleted(nb, False) and not cabundle.is_configmap_deleted(nb, True)
    out, changed = cabundle.unset_notebook_cert_config(nb)
    ps = out["spec"]["template"]["spec"]
    assert changed and [e["name"] for e in ps["containers"][1]["env"]] == ["A", "Z"]
    assert [m["name"] for m in ps["containers"][1]["volumeMounts"]] == ["data"]
    assert ps["volumes"] == [{"name": "data"}] and ps["containers"][0]["env"] == [{"name": "PIP_CERT", "value": "kept"}]
    # the same variable twice with the second one last: the range reads a stale copy of the last
    # element past the shortened slice and slices beyond its length — the reference panics (:411-415)
    with pytest.raises(cabundle.ReferencePanic):
        cabundle.unset_notebook_cert_config(notebook_with_certs(dup_last=True))
    out, changed = cabundle.unset_notebook_cert_config({"metadata": {"name": "x"}, "spec": {"template": {"spec": {"containers": [{"name": "y"}]}}}})
    assert not changed


def test_host_library_matches_the_oracle(pkg, wl):
    """kubeflow-rm_b200/host/ca_bundle.cc against the oracle on the same cases."""
    import importlib
    import json
    import random
    import __graft_entry__ as graft
    host = importlib.import_module(graft.PKG_NAME + ".host")
    for odh_cm, kube_cm, want in bundle_cases():
        got = host.ca_bundle(odh_cm, kube_cm)
        assert got == cabundle.cert_bundle(odh_cm, kube_cm)
        if want is not None:
            assert host.pem_count_certificates(got) == want
    odh_cm = {"ca-bundle.crt": PEM["a"], "odh-ca-bundle.crt": PEM["b"]}
    ok = {"ca-bundle.crt": False, "odh-ca-bundle.crt": True}       # the caller's x509 parser rejects the first
    assert host.ca_bundle(odh_cm, None, ok) == cabundle.cert_bundle(odh_cm, None, ok) == PEM["b"]
    nb = notebook_with_certs()
    rc, out = host.unset_notebook_cert_config(wl.dumps(nb))
    assert rc == 1 and json.loads(out) == cabundle.unset_notebook_cert_config(nb)[0]
    assert host.unset_notebook_cert_config(wl.dumps(notebook_with_certs(dup_last=True)))[0] == -1
    assert host.unset_notebook_cert_config(wl.dumps({"metadata": {"name": "x"}, "spec": {"template": {"spec": {"containers": [{"name": "y"}]}}}}))[0] == 0
    rng = random.Random(3)
    for _ in range(300):                      # shuffled / duplicated variables: same result or same panic
        nb = notebook_with_certs()
        env = nb["spec"]["template"]["spec"]["containers"][1]["env"]
        env += [dict(rng.choice(env)) for _ in range(rng.randrange(3))]
        rng.shuffle(env)
        rc, out = host.unset_notebook_cert_config(wl.dumps(nb))
        try:
            want, _ = cabundle.unset_notebook_cert_config(nb)
            assert rc == 1 and json.loads(out) == want
        except cabundle.ReferencePanic:
            assert rc == -1
