"""Oracle (test infrastructure): the wire formats around the two webhooks.

Reference:
  components/admission-webhook/main.go:599-704 mutatePods, :708-749 serve,
  :64-70 toAdmissionResponse;
  components/odh-notebook-controller/controllers/notebook_webhook.go:234-302 Handle, whose
  return values are controller-runtime's admission.Errored / Denied / PatchResponseFromRaw
  (sigs.k8s.io/controller-runtime v0.17.0, not vendored: shapes recalled, unpinned).
Patches are compared by effect (the two jsonpatch libraries order their operations by Go map
iteration): `check_patch_effect` applies a patch to the original and compares with the
mutated tree.
"""
import base64
import json

from . import gojson, odh, poddefault


def review_request(obj, uid, operation="CREATE", old=None, namespace=None, resource=None, version="v1"):
    """An AdmissionReview request as the API server sends it."""
    req = {"uid": uid, "kind": {"group": "", "version": "v1", "kind": obj.get("kind", "Pod")},
           "resource": resource or {"group": "", "version": "v1", "resource": "pods"},
           "operation": operation, "object": obj}
    md = obj.get("metadata") or {}
    if md.get("name"):
        req["name"] = md["name"]
    if namespace is not None:
        req["namespace"] = namespace
    if old is not None:
        req["oldObject"] = old
    return {"apiVersion": "admission.k8s.io/" + version, "kind": "AdmissionReview", "request": req}


def check_patch_effect(original, patch, mutated):
    got = poddefault.apply_json_patch(original, patch)
    assert gojson.canon(got) == gojson.canon(mutated), \
        "patched %s\nexpected %s" % (gojson.canon(got), gojson.canon(mutated))


def poddefault_response(review, poddefaults):
    """serve(mutatePods) for one decoded review: what the response must say.  Returns
    ("empty",) for a resource that is not core/v1 pods (a review without a response),
    ("allowed", mutated pod or None, the pod the patch applies to) or ("error", message)."""
    req = review["request"]
    if req.get("resource") != {"group": "", "version": "v1", "resource": "pods"}:
        return ("empty",)                                                                # main.go:602-605
    allowed, mutated, msg = poddefault.mutate_pod(req["object"], poddefaults, req.get("namespace", ""))
    if not allowed:
        return ("error", msg)
    return ("allowed", mutated, req["object"])


def notebook_webhook_response(review, image=odh.OAUTH_PROXY_IMAGE, ca_bundle=False):
    """NotebookWebhook.Handle: ("denied", message) | ("allowed", mutated notebook)."""
    req = review["request"]
    try:
        mutated, _ = odh.mutate_notebook(req["object"], req.get("operation", "CREATE"), old=req.get("oldObject"),
                                         image=image, ca_bundle=ca_bundle)
    except odh.Denied as e:
        return ("denied", str(e))
    return ("allowed", mutated)


def decode_response(raw):
    """(uid, allowed, patch list or None, status dict or None, has_response) of a response body."""
    r = json.loads(raw)
    resp = r.get("response")
    if resp is None:
        return None, None, None, None, False
    patch = None
    if "patch" in resp:
        assert resp.get("patchType") == "JSONPatch"
        patch = json.loads(base64.b64decode(resp["patch"]))
    else:
        assert "patchType" not in resp
    return resp.get("uid"), resp.get("allowed"), patch, resp.get("status"), True
