"""Oracle: PodDefault admission mutation (test infrastructure only).

Follows components/admission-webhook/main.go:64-704 on JSON trees.  Golden:
admission-webhook/main_test.go:12-73 (mergeMap), :254-324 (setCommandAndArgs);
:80-252 pins only "no conflict" (its assertion is inverted).  filterPodDefaults
has no test in the reference: label-selector semantics are recalled from
k8s.io/apimachinery v0.28.1 -> "parity unpinned by
reference tests" for that function.
"""
from .gojson import deepcopy, get_map, meta
from .nbc import json_equal

ANNOTATION_PREFIX = "poddefault.admission.kubeflow.org"     # main.go:46
ISTIO_PROXY_CONTAINER_NAME = "istio-proxy"                   # :47
MIRROR_POD_ANNOTATION_KEY = "kubernetes.io/config.mirror"    # corev1.MirrorPodAnnotationKey


class SelectorError(Exception):
    pass


class Conflict(Exception):
    """utilerrors.NewAggregate(errs) with at least one error; .keys lists what
    conflicted.  The reference's message text uses Go %#v dumps (main.go:199...)
    and is not reproducible outside Go; compare on the keys."""

    def __init__(self, keys):
        super().__init__("conflict on: " + ", ".join(keys))
        self.keys = list(keys)


# ---------------------------------------------------------------- label selectors


def selector_matches(sel, labels):
    """metav1.LabelSelectorAsSelector(&sel).Matches(labels.Set(labels)), main.go:76-82.

    Empty selector matches everything; In/NotIn need >= 1 value, Exists /
    DoesNotExist need none; NotIn and DoesNotExist match when the key is absent;
    an unknown operator is an error that aborts the request (:77-79)."""
    sel = sel or {}
    labels = labels or {}
    # LabelSelectorAsSelector converts (and validates) every requirement before
    # anything is matched, so an invalid one is an error even when an earlier
    # requirement would not match
    for req in sel.get("matchExpressions") or []:
        op, vals = req.get("operator", ""), req.get("values") or []
        if op in ("In", "NotIn"):
            if not vals:
                raise SelectorError("values: Invalid value: for 'in', 'notin' operators, values set can't be empty")
        elif op in ("Exists", "DoesNotExist"):
            if vals:
                raise SelectorError("values: Invalid value: values set must be empty for exists and does not exist")
        else:
            raise SelectorError("%r is not a valid label selector operator" % op)
    for k, v in (sel.get("matchLabels") or {}).items():
        if k not in labels or labels[k] != v:
            return False
    for req in sel.get("matchExpressions") or []:
        key, op, vals = req.get("key", ""), req.get("operator", ""), req.get("values") or []
        if op == "In":
            if key not in labels or labels[key] not in vals:
                return False
        elif op == "NotIn":
            if key in labels and labels[key] in vals:
                return False
        elif op == "Exists":
            if key not in labels:
                return False
        elif key in labels:
            return False
    return True


def filter_pod_defaults(pds, pod):
    """filterPodDefaults, main.go:72-97.  The selector error is raised before
    the namespace test, as in the reference."""
    out = []
    for pd in pds:
        if not selector_matches((pd.get("spec") or {}).get("selector"), get_map(meta(pod), "labels")):
            continue
        if meta(pd).get("namespace", "") != meta(pod).get("namespace", ""):
            continue
        out.append(pd)
    return out


# ---------------------------------------------------------------- merges


def _spec(pd):
    return pd.get("spec") or {}


def _merge_keyed(existing, pds, field, key, what, nil_when_empty, errs):
    """Shared shape of mergeImagePullSecrets/Env/Volumes/Containers/Tolerations
    (main.go:170-450): append if the key is new, identical duplicate allowed,
    differing duplicate is a conflict."""
    orig = {}
    for x in existing or []:
        orig[(x or {}).get(key, "")] = x
    merged = list(existing or [])
    bad = []
    for pd in pds:
        for v in _spec(pd).get(field) or []:
            k = (v or {}).get(key, "")
            if k not in orig:
                orig[k] = v
                merged.append(v)
                continue
            if not json_equal(orig[k], v):
                bad.append("%s %s" % (what, k))
    if bad:
        errs.extend(bad)
        return None
    if nil_when_empty and not merged:
        return None                       # :355,:402,:445 return nil when empty
    return merged


def merge_image_pull_secrets(ips, pds, errs):
    return _merge_keyed(ips, pds, "imagePullSecrets", "name", "imagePullSecret", False, errs)   # :170-213


def merge_env(env, pds, errs):
    return _merge_keyed(env, pds, "env", "name", "env", False, errs)                            # :217-252


def merge_env_from(env_from, pds):
    """mergeEnvFrom, :254-262: plain concatenation, never conflicts; nil when nothing."""
    out = list(env_from or [])
    for pd in pds:
        out.extend(_spec(pd).get("envFrom") or [])
    return out or None


def merge_volume_mounts(mounts, pds, errs):
    """mergeVolumeMounts, :266-317: keyed by name AND by mountPath."""
    by_name, by_path = {}, {}
    for v in mounts or []:
        by_name[(v or {}).get("name", "")] = v
        by_path[(v or {}).get("mountPath", "")] = v
    merged = list(mounts or [])
    bad = []
    for pd in pds:
        for v in _spec(pd).get("volumeMounts") or []:
            n, p = (v or {}).get("name", ""), (v or {}).get("mountPath", "")
            if n not in by_name:
                by_name[n] = v
                merged.append(v)
            elif not json_equal(by_name[n], v):
                bad.append("volumeMount %s" % n)
            if p not in by_path:
                by_path[p] = v
            elif not json_equal(by_path[p], v):
                bad.append("mountPath %s" % p)
    if bad:
        errs.extend(bad)
        return None
    return merged


def merge_volumes(vols, pds, errs):
    return _merge_keyed(vols, pds, "volumes", "name", "volume", True, errs)                     # :321-360


def merge_containers(ctrs, pds, is_sidecar, errs):
    return _merge_keyed(ctrs, pds, "sidecars" if is_sidecar else "initContainers", "name",
                        "container", True, errs)                                                # :364-407


def merge_tolerations(tols, pds, errs):
    return _merge_keyed(tols, pds, "tolerations", "key", "toleration", True, errs)              # :411-450


def merge_map(existing, defaults, errs=None):
    """mergeMap, :454-475.  Returns the merged map (always non-nil); conflicts
    are appended to errs (or raised when errs is None)."""
    out = dict(existing or {})
    bad = []
    for d in defaults:
        for k, v in (d or {}).items():
            if k not in out:
                out[k] = v
            elif out[k] != v:
                bad.append("map key %s" % k)
    if bad:
        if errs is None:
            raise Conflict(bad)
        errs.extend(bad)
    return out


def safe_to_apply_pod_defaults_on_pod(pod, pds):
    """safeToApplyPodDefaultsOnPod + ...OnContainer, :101-166.  Raises Conflict."""
    errs = []
    spec = pod.get("spec") or {}
    merge_volumes(spec.get("volumes"), pds, errs)
    merge_tolerations(spec.get("tolerations"), pds, errs)
    merge_image_pull_secrets(spec.get("imagePullSecrets"), pds, errs)
    for c in spec.get("containers") or []:
        merge_env(c.get("env"), pds, errs)
        merge_volume_mounts(c.get("volumeMounts"), pds, errs)
    merge_map(get_map(meta(pod), "annotations"), [_spec(p).get("annotations") for p in pds], errs)
    merge_map(get_map(meta(pod), "labels"), [_spec(p).get("labels") for p in pds], errs)
    merge_containers(spec.get("initContainers"), pds, False, errs)
    merge_containers(spec.get("containers"), pds, True, errs)
    if errs:
        raise Conflict(errs)


def set_command_and_args(ctr, pds):
    """setCommandAndArgs, :582-597.  nil test: absent or null."""
    if ctr.get("name", "") == ISTIO_PROXY_CONTAINER_NAME:
        return
    for pd in pds:
        s = _spec(pd)
        if ctr.get("command") is None and s.get("command") is not None:
            ctr["command"] = deepcopy(s["command"])
        if ctr.get("args") is None and s.get("args") is not None:
            ctr["args"] = deepcopy(s["args"])


def _set_or_drop(obj, key, val):
    if val is None:
        obj.pop(key, None)
    else:
        obj[key] = deepcopy(val)


def apply_pod_defaults_on_pod(pod, pds):
    """applyPodDefaultsOnPod + ...OnContainer, :480-578.  Mutates pod."""
    if not pds:
        return
    errs = []
    spec = pod.setdefault("spec", {})
    m = pod.setdefault("metadata", {})
    _set_or_drop(spec, "volumes", merge_volumes(spec.get("volumes"), pds, errs))
    _set_or_drop(spec, "tolerations", merge_tolerations(spec.get("tolerations"), pds, errs))
    ips = merge_image_pull_secrets(spec.get("imagePullSecrets"), pds, errs)
    _set_or_drop(spec, "imagePullSecrets", ips or None)            # empty non-nil -> omitempty
    for pd in pds:                                                 # :509-518 last one wins
        s = _spec(pd)
        if s.get("automountServiceAccountToken") is not None:
            spec["automountServiceAccountToken"] = s["automountServiceAccountToken"]
        if s.get("serviceAccountName", "") != "":
            spec["serviceAccountName"] = s["serviceAccountName"]
    m["annotations"] = merge_map(get_map(m, "annotations"), [_spec(p).get("annotations") for p in pds], errs)
    labels = merge_map(get_map(m, "labels"), [_spec(p).get("labels") for p in pds], errs)
    _set_or_drop(m, "labels", labels or None)                      # empty map -> omitempty
    for c in spec.get("containers") or []:                         # :531-534
        _set_or_drop(c, "env", merge_env(c.get("env"), pds, errs) or None)
        _set_or_drop(c, "volumeMounts", merge_volume_mounts(c.get("volumeMounts"), pds, errs) or None)
        _set_or_drop(c, "envFrom", merge_env_from(c.get("envFrom"), pds))
        set_command_and_args(c, pds)
    # sidecars are appended AFTER the per-container injection (:540-550)
    _set_or_drop(spec, "initContainers", merge_containers(spec.get("initContainers"), pds, False, errs))
    _set_or_drop(spec, "containers", merge_containers(spec.get("containers"), pds, True, errs))
    for pd in pds:                                                 # :553-555
        m["annotations"]["%s/poddefault-%s" % (ANNOTATION_PREFIX, meta(pd).get("name", ""))] = \
            meta(pd).get("resourceVersion", "")


def mutate_pod(pod, pds, request_namespace=""):
    """mutatePods, :599-704, from the decoded Pod onwards.

    Returns (allowed, mutated_pod_or_None, message).  mutated is None when no
    patch is produced (:628,:631,:649,:660)."""
    pod = deepcopy(pod)
    m = pod.setdefault("metadata", {})
    if m.get("namespace", "") == "" and request_namespace != "":
        m["namespace"] = request_namespace                          # :616-619
    ann = m.get("annotations")
    if ann is not None:                                             # :625-633
        if ann.get(ANNOTATION_PREFIX + "/exclude", "") == "true":
            return True, None, ""
        if MIRROR_POD_ANNOTATION_KEY in ann:
            return True, None, ""
    if not pds:
        return True, None, ""
    try:
        matching = filter_pod_defaults(pds, pod)
    except SelectorError as e:
        return False, None, str(e)                                  # :653-656
    if not matching:
        return True, None, ""
    try:
        safe_to_apply_pod_defaults_on_pod(pod, matching)
    except Conflict as e:
        return False, None, str(e)                                  # :671-679
    apply_pod_defaults_on_pod(pod, matching)
    return True, pod, ""


def apply_json_patch(doc, patch):
    """Minimal RFC 6902 applier (add / replace / remove) used to check patches
    by effect: apply(patch, original) must equal the oracle's mutated pod."""
    doc = deepcopy(doc)
    for op in patch:
        parts = [p.replace("~1", "/").replace("~0", "~") for p in op["path"].split("/")[1:]]
        cur = doc
        for p in parts[:-1]:
            cur = cur[int(p)] if isinstance(cur, list) else cur[p]
        last = parts[-1]
        if isinstance(cur, list):
            if op["op"] == "add":
                if last == "-":
                    cur.append(deepcopy(op["value"]))
                else:
                    cur.insert(int(last), deepcopy(op["value"]))
            elif op["op"] == "replace":
                cur[int(last)] = deepcopy(op["value"])
            elif op["op"] == "remove":
                del cur[int(last)]
        else:
            if op["op"] in ("add", "replace"):
                cur[last] = deepcopy(op["value"])
            elif op["op"] == "remove":
                del cur[last]
    return doc
