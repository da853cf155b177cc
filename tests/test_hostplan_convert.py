"""CPU: Notebook version conversion (GenConvert, nbrm_convert.cuh) against the oracle's
restatement of ConvertTo / ConvertFrom (notebook-controller/api/v1/notebook_conversion.go:25-69)
on the reference's own sample CRs and on fuzzed Notebooks with statuses."""
import json
import os
import random

import yaml

from oracle import gojson, nbc

SAMPLES = "/root/reference/components/notebook-controller/config/samples"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nbc_samples.json")


def samples():
    """The three sample CRs of the reference (config/samples/_v1*_notebook.yaml), kept as a fixture
    because /root/reference does not travel to the GPU box."""
    if os.path.isdir(SAMPLES) and not os.path.exists(GOLDEN):
        out = {f: yaml.safe_load(open(os.path.join(SAMPLES, f))) for f in sorted(os.listdir(SAMPLES)) if f.endswith("notebook.yaml")}
        json.dump(out, open(GOLDEN, "w"), indent=1, sort_keys=True)
    return json.load(open(GOLDEN))


def with_status(rng, nb):
    nb = json.loads(json.dumps(nb))
    r = rng.random()
    if r < 0.8:
        nb["status"] = {"readyReplicas": rng.randint(0, 1),
                        "containerState": rng.choice([{}, {"running": {"startedAt": "2026-11-06T00:00:00Z"}},
                                                      {"waiting": {"reason": "PodInitializing"}}]),
                        "conditions": [{"type": t, "status": rng.choice(["True", "False"]),
                                        **({"lastProbeTime": "2026-11-06T10:00:00Z"} if rng.random() < 0.7 else {}),
                                        "lastTransitionTime": "2026-11-06T09:00:00Z",
                                        **({"reason": "r\"q"} if rng.random() < 0.5 else {}),
                                        **({"message": rng.choice(["m", ""])} if rng.random() < 0.5 else {})}
                                       for t in rng.sample(["Running", "Waiting", "Terminated", "Ready"], rng.randint(0, 3))]}
    return nb


def test_samples_and_fuzz(hostplan, wl):
    rng = random.Random(4)
    nbs = [with_status(rng, nb) for nb in list(samples().values()) + wl.corpus("fuzz", 800, start=3) + wl.corpus("c1", 3)]
    docs = [wl.dumps(nb, rng) for nb in nbs]
    for to_version in (0, 1, 2):
        for flags in (0, 1):
            outs, st = hostplan.convert(docs, to_version, flags)
            for nb, o, s in zip(nbs, outs, st):
                assert s == 0
                want = nbc.convert_notebook(nb, nbc.API_VERSIONS[to_version], keep_metadata=bool(flags))
                assert gojson.canon(json.loads(o)) == gojson.canon(want), (o, want)
    # what the reference's functions lose: a condition's status and lastTransitionTime
    nb = with_status(random.Random(1), samples()["_v1_notebook.yaml"])
    nb["status"]["conditions"] = [{"type": "Running", "status": "True", "lastTransitionTime": "2026-11-06T09:00:00Z"}]
    out = json.loads(hostplan.convert([wl.dumps(nb)], 0)[0][0])
    assert out["apiVersion"] == "kubeflow.org/v1beta1" and out["status"]["conditions"] == [{"type": "Running", "status": ""}]
    assert out["metadata"] == {}
