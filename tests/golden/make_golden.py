#!/usr/bin/env python
"""Writes tests/golden/*.json: the reference's own known-answer tables for the
hot path, transcribed by hand from its Go test files into JSON trees.

The reference cannot run in this image (no Go toolchain), so these are not
"outputs of the reference run here" but its tests' literal inputs and expected
values; each block cites the file:line it was read from (paths relative to
/root/reference/components).  Go struct literals are written as the JSON the
Kubernetes types serialise to, zero-valued fields left out.

Run:  python tests/golden/make_golden.py      (rewrites the .json files)
"""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def write(name, obj):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(obj, f, indent=1, sort_keys=True)
        f.write("\n")


# ---------------------------------------------------------------------------
# notebook-controller/controllers/notebook_controller_test.go:94-291 TestCreateNotebookStatus
T = "2022-08-30T01:10:30Z"    # v1.Date(2022, 8, 30, 1, 10, 30, 0, time.UTC)  (:194)
T2 = "2022-04-21T01:10:30Z"   # v1.Date(2022, 4, 21, 1, 10, 30, 0, time.UTC)  (:244)
NB = {"metadata": {"name": "test", "namespace": "kubeflow-user"}, "status": {}}
STS1 = {"metadata": {"name": "test", "namespace": "kubeflow-user"}, "status": {"readyReplicas": 1}}
status_cases = [
    {"name": "NotebookStatusInitialization", "src": "notebook_controller_test.go:103-119",
     "notebook": NB, "pod": {}, "sts": {},
     "expected": {"conditions": [], "readyReplicas": 0, "containerState": {}}},
    {"name": "NotebookStatusReadyReplicas", "src": "notebook_controller_test.go:120-144",
     "notebook": NB, "pod": {}, "sts": STS1,
     "expected": {"conditions": [], "readyReplicas": 1, "containerState": {}}},
    {"name": "NotebookContainerState", "src": "notebook_controller_test.go:145-182",
     "notebook": NB,
     "pod": {"metadata": {"name": "test", "namespace": "kubeflow-user"},
             "status": {"containerStatuses": [{"name": "test", "state": {"running": {"startedAt": None}}}]}},
     "sts": {},
     "expected": {"conditions": [], "readyReplicas": 0,
                  "containerState": {"running": {"startedAt": None}}}},
    {"name": "mirroringPodConditions", "src": "notebook_controller_test.go:183-232",
     "notebook": {},
     "pod": {"metadata": {"name": "test", "namespace": "kubeflow-user"},
             "status": {"conditions": [
                 {"type": "Running", "lastProbeTime": T, "lastTransitionTime": T},
                 {"type": "Waiting", "lastProbeTime": T, "lastTransitionTime": T,
                  "reason": "PodInitializing"}]}},
     "sts": STS1,
     "expected": {"conditions": [
         {"type": "Running", "status": "", "lastProbeTime": T, "lastTransitionTime": T},
         {"type": "Waiting", "status": "", "lastProbeTime": T, "lastTransitionTime": T,
          "reason": "PodInitializing"}],
         "readyReplicas": 1, "containerState": {}}},
    {"name": "unschedulablePod", "src": "notebook_controller_test.go:233-274",
     "notebook": {},
     "pod": {"metadata": {"name": "test", "namespace": "kubeflow-user"},
             "status": {"conditions": [
                 {"type": "PodScheduled", "lastProbeTime": T2, "lastTransitionTime": T2,
                  "message": "0/1 nodes are available: 1 Insufficient cpu.", "status": "false",
                  "reason": "Unschedulable"}]}},
     "sts": {"metadata": {"name": "test", "namespace": "kubeflow-user"}, "status": {}},
     "expected": {"conditions": [
         {"type": "PodScheduled", "lastProbeTime": T2, "lastTransitionTime": T2,
          "message": "0/1 nodes are available: 1 Insufficient cpu.", "status": "false",
          "reason": "Unschedulable"}],
         "readyReplicas": 0, "containerState": {}}},
]

# notebook-controller/controllers/notebook_controller_test.go:22-92 TestNbNameFromInvolvedObject
involved_cases = {
    "src": "notebook_controller_test.go:22-92",
    "pods": [{"metadata": {"name": "test-notebook-0", "namespace": "test-namespace",
                           "labels": {"notebook-name": "test-notebook"}}}],
    "cases": [
        {"name": "pod event", "involvedObject": {"kind": "Pod", "name": "test-notebook-0",
                                                 "namespace": "test-namespace"},
         "expected": "test-notebook"},
        {"name": "statefulset event", "involvedObject": {"kind": "StatefulSet", "name": "test-notebook",
                                                         "namespace": "test-namespace"},
         "expected": "test-notebook"},
    ],
}

# ---------------------------------------------------------------------------
# notebook-controller/controllers/culling_controller_test.go
# time.Now() in the tables is replaced by a fixed instant NOW; "now+N" strings
# are resolved by the test with the oracle's own RFC 3339 formatter.
STOP = "kubeflow-resource-stopped"
LAST = "notebooks.kubeflow.org/last-activity"
culling = {
    "now": "2026-11-06T12:00:00Z",
    "set_stop_annotation": {  # :14-56 — after the call the annotation must be present
        "src": "culling_controller_test.go:14-56",
        "cases": [
            {"name": "Nil Metadata", "meta": None},
            {"name": "No existing Annotations", "meta": {}},
            {"name": "Basic case", "meta": {"annotations": {}}},
            {"name": "Annotation is already set", "meta": {"annotations": {STOP: "now+0"}}},
        ]},
    "stop_annotation_is_set": {  # :58-94
        "src": "culling_controller_test.go:58-94",
        "cases": [
            {"name": "No existing Annotations", "meta": {}, "result": False},
            {"name": "Stop Annotation not Set", "meta": {"annotations": {}}, "result": False},
            {"name": "Annotation is already set", "meta": {"annotations": {STOP: "now+0"}}, "result": True},
        ]},
    "all_kernels_are_idle": {  # :96-141
        "src": "culling_controller_test.go:96-141",
        "cases": [
            {"name": "empty list", "kernels": [], "result": True},
            {"name": "idle kernels", "kernels": [{"execution_state": "idle"}, {"execution_state": "idle"}],
             "result": True},
            {"name": "one busy", "kernels": [{"execution_state": "idle"}, {"execution_state": "busy"}],
             "result": False},
        ]},
    "notebook_is_idle": {  # :143-263
        "src": "culling_controller_test.go:143-263",
        "cases": [
            {"name": "No existing Annotations", "meta": {}, "env": {}, "result": False},
            {"name": "Basic case", "meta": {"annotations": {}}, "env": {}, "result": False},
            {"name": "Stop Annotation already set", "meta": {"annotations": {STOP: "now+0"}}, "env": {},
             "result": False},
            {"name": "LAST_ACTIVITY_ANNOTATION is not set", "meta": {"annotations": {}}, "env": {},
             "result": False},
            {"name": "LAST_ACTIVITY_ANNOTATION is not RF3339",
             "meta": {"annotations": {LAST: "should-fail"}}, "env": {}, "result": False},
            {"name": "LAST_ACTIVITY_ANNOTATION is old",
             "meta": {"annotations": {LAST: "2021-08-30T15:37:36.990063Z"}}, "env": {}, "result": True},
            {"name": "LAST_ACTIVITY_ANNOTATION is too old",
             "meta": {"annotations": {LAST: "1900-08-30T15:37:36.990063Z"}}, "env": {}, "result": True},
            {"name": "LAST_ACTIVITY_ANNOTATION is the current time",
             "meta": {"annotations": {LAST: "now+0"}}, "env": {"CULL_IDLE_TIME": "5"}, "result": False},
            {"name": "1 minute MORE than the deadline",
             "meta": {"annotations": {LAST: "now-360"}}, "env": {"CULL_IDLE_TIME": "5"}, "result": True},
            {"name": "1 minute LESS than the deadline",
             "meta": {"annotations": {LAST: "now-180"}}, "env": {"CULL_IDLE_TIME": "5"}, "result": False},
        ]},
}

# ---------------------------------------------------------------------------
# odh-notebook-controller/controllers/notebook_controller_test.go
CTRL_NS = "redhat-ods-applications"     # suite_test.go:74


def create_notebook(name, ns):          # :970-984 createNotebook
    return {"apiVersion": "kubeflow.org/v1", "kind": "Notebook",
            "metadata": {"name": name, "namespace": ns},
            "spec": {"template": {"spec": {"containers": [
                {"name": name, "image": "registry.redhat.io/ubi8/ubi:latest"}]}}}}


def expected_route(name, ns, to, port, term):   # :61-87 and :798-823
    return {"metadata": {"name": name, "namespace": ns, "labels": {"notebook-name": name}},
            "spec": {"to": {"kind": "Service", "name": to, "weight": 100},
                     "port": {"targetPort": port},
                     "tls": {"termination": term, "insecureEdgeTerminationPolicy": "Redirect"},
                     "wildcardPolicy": "None"},
            "status": {"ingress": []}}


OAUTH_IMAGE = ("registry.redhat.io/openshift4/ose-oauth-proxy@sha256:"
               "4f8d66597feeb32bb18699326029f9a71a5aca4a57679d636b876377c2e95695")  # notebook_oauth.go:41


def create_oauth_container(name, ns):   # :1002-1086 createOAuthContainer
    probe = {"httpGet": {"path": "/oauth/healthz", "port": "oauth-proxy", "scheme": "HTTPS"},
             "timeoutSeconds": 1, "periodSeconds": 5, "successThreshold": 1, "failureThreshold": 3}
    return {
        "name": "oauth-proxy", "image": OAUTH_IMAGE, "imagePullPolicy": "Always",
        "env": [{"name": "NAMESPACE", "valueFrom": {"fieldRef": {"fieldPath": "metadata.namespace"}}}],
        "args": [
            "--provider=openshift", "--https-address=:8443", "--http-address=",
            "--openshift-service-account=" + name,
            "--cookie-secret-file=/etc/oauth/config/cookie_secret", "--cookie-expire=24h0m0s",
            "--tls-cert=/etc/tls/private/tls.crt", "--tls-key=/etc/tls/private/tls.key",
            "--upstream=http://localhost:8888",
            "--upstream-ca=/var/run/secrets/kubernetes.io/serviceaccount/ca.crt",
            "--email-domain=*", "--skip-provider-button",
            '--openshift-sar={"verb":"get","resource":"notebooks","resourceAPIGroup":"kubeflow.org",'
            '"resourceName":"' + name + '","namespace":"$(NAMESPACE)"}',
            "--logout-url=https://example.notebook-url/notebook/" + ns + "/" + name],
        "ports": [{"name": "oauth-proxy", "containerPort": 8443, "protocol": "TCP"}],
        "livenessProbe": dict(probe, initialDelaySeconds=30),
        "readinessProbe": dict(probe, initialDelaySeconds=5),
        "resources": {"requests": {"cpu": "100m", "memory": "64Mi"},
                      "limits": {"cpu": "100m", "memory": "64Mi"}},
        "volumeMounts": [{"name": "oauth-config", "mountPath": "/etc/oauth/config"},
                         {"name": "tls-certificates", "mountPath": "/etc/tls/private"}],
    }


def oauth_notebook(name, ns):           # :573-600
    nb = create_notebook(name, ns)
    nb["metadata"]["labels"] = {"app.kubernetes.io/instance": name}
    nb["metadata"]["annotations"] = {
        "notebooks.opendatahub.io/inject-oauth": "true",
        "notebooks.opendatahub.io/foo": "bar",
        "notebooks.opendatahub.io/oauth-logout-url": "https://example.notebook-url/notebook/" + ns + "/" + name}
    nb["spec"]["template"]["spec"]["volumes"] = [
        {"name": "notebook-data", "persistentVolumeClaim": {"claimName": name + "-data"}}]
    return nb


def expected_oauth_notebook(name, ns):  # :602-658
    return {
        "metadata": {"name": name, "namespace": ns,
                     "labels": {"app.kubernetes.io/instance": name},
                     "annotations": {
                         "notebooks.opendatahub.io/inject-oauth": "true",
                         "notebooks.opendatahub.io/foo": "bar",
                         "notebooks.opendatahub.io/oauth-logout-url":
                             "https://example.notebook-url/notebook/" + ns + "/" + name,
                         "kubeflow-resource-stopped": "odh-notebook-controller-lock"}},
        "spec": {"template": {"spec": {
            "serviceAccountName": name,
            "containers": [{"name": name, "image": "registry.redhat.io/ubi8/ubi:latest"},
                           create_oauth_container(name, ns)],
            "volumes": [
                {"name": "notebook-data", "persistentVolumeClaim": {"claimName": name + "-data"}},
                {"name": "oauth-config", "secret": {"secretName": name + "-oauth-config", "defaultMode": 420}},
                {"name": "tls-certificates", "secret": {"secretName": name + "-tls", "defaultMode": 420}}]}}}}


def expected_np(name, ns, suffix, port, from_ns=None):   # :436-477 and :1088-1119
    rule = {"ports": [{"protocol": "TCP", "port": port}]}
    if from_ns:
        rule["from"] = [{"namespaceSelector": {"matchLabels": {"kubernetes.io/metadata.name": from_ns}}}]
    return {"metadata": {"name": name + suffix, "namespace": ns},
            "spec": {"podSelector": {"matchLabels": {"notebook-name": name}},
                     "ingress": [rule], "policyTypes": ["Ingress"]}}


odh = {
    "controller_namespace": CTRL_NS,
    "route": {"src": "odh notebook_controller_test.go:53-87",
              "notebook": create_notebook("test-notebook", "default"),
              "expected": expected_route("test-notebook", "default", "test-notebook",
                                         "http-test-notebook", "edge")},
    "owner_reference": {"src": "odh notebook_controller_test.go:142-149",
                        "expected": {"apiVersion": "kubeflow.org/v1", "kind": "Notebook",
                                     "name": "test-notebook", "controller": True,
                                     "blockOwnerDeletion": True}},
    "network_policies": {"src": "odh notebook_controller_test.go:425-477,1088-1119",
                         "notebook": create_notebook("test-notebook-np", "default"),
                         "expected_ctrl": expected_np("test-notebook-np", "default", "-ctrl-np", 8888, CTRL_NS),
                         "expected_oauth": expected_np("test-notebook-np", "default", "-oauth-np", 8443)},
    "oauth": {"src": "odh notebook_controller_test.go:566-823,986-1086",
              "notebook": oauth_notebook("test-notebook-oauth", "default"),
              "expected_notebook": expected_oauth_notebook("test-notebook-oauth", "default"),
              "expected_service_account": {
                  "metadata": {"name": "test-notebook-oauth", "namespace": "default",
                               "labels": {"notebook-name": "test-notebook-oauth"},
                               "annotations": {
                                   "serviceaccounts.openshift.io/oauth-redirectreference.first":
                                       '{"kind":"OAuthRedirectReference","apiVersion":"v1",'
                                       '"reference":{"kind":"Route","name":"test-notebook-oauth"}}'}}},
              "expected_service": {
                  "metadata": {"name": "test-notebook-oauth-tls", "namespace": "default",
                               "labels": {"notebook-name": "test-notebook-oauth"},
                               "annotations": {"service.beta.openshift.io/serving-cert-secret-name":
                                               "test-notebook-oauth-tls"}},
                  "spec": {"ports": [{"name": "oauth-proxy", "port": 443, "targetPort": "oauth-proxy",
                                      "protocol": "TCP"}]}},
              "expected_secret_cookie_len": 32,   # :782
              "expected_route": expected_route("test-notebook-oauth", "default", "test-notebook-oauth-tls",
                                               "oauth-proxy", "reencrypt")},
    "service_mesh_denied": {"src": "odh notebook_controller_test.go:903-967",
                            "annotations": {"notebooks.opendatahub.io/inject-oauth": "true",
                                            "opendatahub.io/service-mesh": "true"}},
    "diff_string": {"src": "odh notebook_webhook_utils_test.go:28-64",
                    "expected_go": "{v1.Pod}.Spec.NodeName: node1 != node2",
                    "a": {"spec": {"nodeName": "node1"}}, "b": {"spec": {"nodeName": "node2"}}},
}

# ---------------------------------------------------------------------------
# admission-webhook/main_test.go
adm = {
    "merge_map_bad": {"src": "admission-webhook/main_test.go:12-28",
                      "cases": [{"existing": {"foo": "bar"}, "defaults": {"foo": "buz"}}]},
    "merge_map_good": {"src": "admission-webhook/main_test.go:30-73", "cases": [
        {"name": "Add annotation", "existing": {"foo": "bar"}, "defaults": {"baz": "bux"},
         "out": {"foo": "bar", "baz": "bux"}},
        {"name": "Add nothing", "existing": {"foo": "bar"}, "defaults": {}, "out": {"foo": "bar"}},
        {"name": "Same k/v in annotations", "existing": {"foo": "bar"}, "defaults": {"foo": "bar"},
         "out": {"foo": "bar"}}]},
    # :80-252 — the reference's assertion is inverted (fails only when equal), so it
    # pins "no conflict error" only; the expected trees below are what the source
    # text produces and are checked as such.
    "apply_no_conflict": {"src": "admission-webhook/main_test.go:80-252", "cases": [
        {"name": "Add Annotations", "pod": {"metadata": {"annotations": {"foo": "bar"}}},
         "poddefaults": [{"spec": {"annotations": {"baz": "bux"}, "serviceAccountName": "some-service-account",
                                   "automountServiceAccountToken": True}}]},
        {"name": "Same k/v", "pod": {"metadata": {"annotations": {"foo": "bar"}}},
         "poddefaults": [{"spec": {"annotations": {"foo": "bar"}}}]},
        {"name": "Add tolerations",
         "pod": {"spec": {"tolerations": [{"key": "oldToleration", "operator": "Exists", "effect": "NoSchedule"}]}},
         "poddefaults": [{"spec": {"tolerations": [
             {"key": "newToleration", "operator": "Equal", "value": "foo", "effect": "NoSchedule"}]}}]},
        {"name": "Add init containers", "pod": {"spec": {"containers": []}},
         "poddefaults": [{"spec": {"initContainers": [
             {"command": ["cmd1"], "args": ["arg1", "arg2"], "image": "nginx",
              "name": "test initcontainer and sidecar"}]}}]}]},
    "set_command_and_args": {"src": "admission-webhook/main_test.go:254-324", "cases": [
        {"name": "Set command and args", "in": {},
         "poddefaults": [{"metadata": {"name": "test-pod-default"},
                          "spec": {"command": ["cmd1"], "args": ["arg1", "arg2"]}}],
         "out": {"command": ["cmd1"], "args": ["arg1", "arg2"]}},
        {"name": "Do not overwrite", "in": {"command": ["cmd2"], "args": ["arg3", "arg4"]},
         "poddefaults": [{"metadata": {"name": "test-pod-default"},
                          "spec": {"command": ["cmd1"], "args": ["arg1", "arg2"]}}],
         "out": {"command": ["cmd2"], "args": ["arg3", "arg4"]}}]},
}

# ---------------------------------------------------------------------------
# profile-controller/controllers/profile_controller_test.go:25-111 TestEnforceNamespaceLabelsFromConfig
_DEF = {"katib.kubeflow.org/metrics-collector-injection": "enabled",
        "serving.kubeflow.org/inferenceservice": "enabled", "pipelines.kubeflow.org/enabled": "true",
        "app.kubernetes.io/part-of": "kubeflow-profile"}
profile = {"set_namespace_labels": {"src": "profile-controller/controllers/profile_controller_test.go:25-111", "cases": [
    {"current": None, "labels": _DEF, "expected": dict(_DEF)},
    {"current": {"user-name": "Jim", "serving.kubeflow.org/inferenceservice": "disabled"}, "labels": _DEF,
     "expected": dict(_DEF, **{"user-name": "Jim", "serving.kubeflow.org/inferenceservice": "disabled"})},
    {"current": {"user-name": "Jim", "removal-label": "enabled"}, "labels": dict(_DEF, **{"removal-label": ""}),
     "expected": dict(_DEF, **{"user-name": "Jim"})}]}}

if __name__ == "__main__":
    write("profile.json", profile)
    write("nbc_status.json", status_cases)
    write("nbc_involved_object.json", involved_cases)
    write("culling.json", culling)
    write("odh.json", odh)
    write("adm.json", adm)
    print("wrote golden fixtures to", HERE)
