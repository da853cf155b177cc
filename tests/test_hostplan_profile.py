"""CPU: the Profile reconcile planner (compiled by g++) against the oracle."""
import json
import random

from parity import check_profile

LABELS = {"katib.kubeflow.org/metrics-collector-injection": "enabled",
          "serving.kubeflow.org/inferenceservice": "enabled", "pipelines.kubeflow.org/enabled": "true",
          "app.kubernetes.io/part-of": "kubeflow-profile"}   # profile-controller/config/base/namespace-labels.yaml
ENVS = [
    ({}, {}),
    ({"NAMESPACE_LABELS_JSON": json.dumps(LABELS)}, LABELS),
    ({"NAMESPACE_LABELS_JSON": json.dumps(dict(LABELS, **{"istio-injection": "", "x": "y"})),
      "USERID_HEADER": "kubeflow-userid", "USERID_PREFIX": "", "KFP_UI_PRINCIPAL": "spiffe://a/b\"c"},
     dict(LABELS, **{"istio-injection": "", "x": "y"})),
]


def test_profiles(hostplan, wl):
    rng = random.Random(4)
    profiles = [wl.profile_cr(rng, i) for i in range(500)]
    docs = [wl.dumps(p, rng) for p in profiles]
    for env, labels in ENVS:
        outs, st = hostplan.profile(docs, env)
        oenv = {k: v for k, v in env.items() if k != "NAMESPACE_LABELS_JSON"}
        assert check_profile(profiles, outs, st, oenv, labels) > 3500
