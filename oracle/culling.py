"""Oracle: idle-culling predicates and annotation updates (test infrastructure).

Follows components/notebook-controller/controllers/culling_controller.go.
Time is injected: `now_ns` is time.Now() as Unix nanoseconds (UTC), so every
function is deterministic.  `cfg` holds the two integers initGlobalVars reads
(:511-544): CULL_IDLE_TIME and IDLENESS_CHECK_PERIOD, in minutes.

Golden: culling_controller_test.go:16-41 (setStopAnnotation), :59-85
(StopAnnotationIsSet), :97-131 (allKernelsAreIdle), :145-249 (notebookIsIdle).
"""
from .gojson import ZERO_TIME_NS, format_rfc3339, get_map, parse_rfc3339

STOP_ANNOTATION = "kubeflow-resource-stopped"                                       # :50
LAST_ACTIVITY_ANNOTATION = "notebooks.kubeflow.org/last-activity"                   # :51
LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION = "notebooks.kubeflow.org/last_activity_check_timestamp"  # :52
KERNEL_EXECUTION_STATE_IDLE = "idle"                                                # :55

MINUTE_NS = 60 * 1_000_000_000


def init_global_vars(env):
    """initGlobalVars, :511-544.  Returns cfg or raises on a bad check period."""
    def getenv_default(k, d):
        v = env.get(k, "")
        return v if v else d
    try:
        idle = int(getenv_default("CULL_IDLE_TIME", "1440"))       # strconv.Atoi
    except ValueError:
        idle = 1440                                                 # :521-526
    period = int(getenv_default("IDLENESS_CHECK_PERIOD", "1"))      # error returned, :537-540
    return {"CULL_IDLE_TIME": idle, "IDLENESS_CHECK_PERIOD": period}


def _ann(meta):
    return get_map(meta, "annotations")


def stop_annotation_is_set(meta):
    """StopAnnotationIsSet, :480-488."""
    return STOP_ANNOTATION in _ann(meta)


def set_stop_annotation(meta, now_ns):
    """setStopAnnotation, :461-478 (metrics left out).  Mutates meta."""
    if meta is None:
        return
    if not meta.get("annotations"):
        meta["annotations"] = {}
    meta["annotations"][STOP_ANNOTATION] = format_rfc3339(now_ns)


def annotations_exist(meta):
    """annotationsExist, :426-433."""
    a = _ann(meta)
    return LAST_ACTIVITY_ANNOTATION in a and LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION in a


def initialize_annotations(meta, now_ns):
    """initializeAnnotations, :435-442."""
    if not meta.get("annotations"):
        meta["annotations"] = {}
    t = format_rfc3339(now_ns)
    meta["annotations"][LAST_ACTIVITY_ANNOTATION] = t
    meta["annotations"][LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION] = t


def remove_annotations(meta):
    """removeAnnotations, :444-458."""
    if meta is None:
        return
    a = meta.get("annotations")
    if isinstance(a, dict):
        a.pop(LAST_ACTIVITY_ANNOTATION, None)
        a.pop(LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION, None)


def culling_check_period_has_passed(meta, now_ns, cfg):
    """cullingCheckPeriodHasPassed, :173-183.  A parse error is ignored (:178):
    the zero time plus the period is then compared with now."""
    a = _ann(meta)
    if LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION not in a:
        return False
    p = parse_rfc3339(a[LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION])
    stored = p[0] if p else ZERO_TIME_NS
    return stored + cfg["IDLENESS_CHECK_PERIOD"] * MINUTE_NS < now_ns


def notebook_is_idle(meta, now_ns, cfg):
    """notebookIsIdle, :186-207."""
    if meta.get("annotations") is not None:
        if stop_annotation_is_set(meta):
            return False
        p = parse_rfc3339(_ann(meta).get(LAST_ACTIVITY_ANNOTATION, ""))
        if p is None:
            return False
        if now_ns > p[0] + cfg["CULL_IDLE_TIME"] * MINUTE_NS:
            return True
    return False


def all_kernels_are_idle(kernels):
    """allKernelsAreIdle, :281-293."""
    return all((k.get("execution_state") or "") == KERNEL_EXECUTION_STATE_IDLE for k in kernels)


def get_notebook_recent_time(times):
    """getNotebookRecentTime, :295-315: most recent of the stamps, formatted
    back with whole seconds in the stamp's own zone; "" on any parse error."""
    best = parse_rfc3339(times[0])
    if best is None:
        return ""
    for t in times[1:]:
        p = parse_rfc3339(t)
        if p is None:
            return ""
        if p[0] > best[0]:
            best = p
    return format_rfc3339(best[0], best[1])


def compare_annotation_time_to_resource(meta, resource_time):
    """compareAnnotationTimeToResource, :337-355: never move backwards."""
    a = parse_rfc3339(_ann(meta).get(LAST_ACTIVITY_ANNOTATION, ""))
    if a is None:
        return False
    t = parse_rfc3339(resource_time)
    if t is None:
        return False
    return not a[0] > t[0]


def update_timestamp_from_kernels_activity(meta, kernels, now_ns):
    """updateTimestampFromKernelsActivity, :357-388.  A busy kernel sets
    last-activity to now and returns False (:364-372)."""
    if not kernels:
        return False
    if not all_kernels_are_idle(kernels):
        meta["annotations"][LAST_ACTIVITY_ANNOTATION] = format_rfc3339(now_ns)
        return False
    t = get_notebook_recent_time([k.get("last_activity") or "" for k in kernels])
    if t == "" or not compare_annotation_time_to_resource(meta, t):
        return False
    meta["annotations"][LAST_ACTIVITY_ANNOTATION] = t
    return True


def update_timestamp_from_terminals_activity(meta, terminals):
    """updateTimestampFromTerminalsActivity, :390-414."""
    if not terminals:
        return False
    t = get_notebook_recent_time([x.get("last_activity") or "" for x in terminals])
    if t == "" or not compare_annotation_time_to_resource(meta, t):
        return False
    meta["annotations"][LAST_ACTIVITY_ANNOTATION] = t
    return True


def update_notebook_last_activity_annotation(meta, kernels, terminals, now_ns):
    """updateNotebookLastActivityAnnotation, :318-335; kernels/terminals are the
    decoded /api/kernels and /api/terminals bodies, None when the GET failed."""
    if kernels is None and terminals is None:
        return False
    k = update_timestamp_from_kernels_activity(meta, kernels, now_ns)
    t = update_timestamp_from_terminals_activity(meta, terminals)
    return k or t


def update_last_culling_check_timestamp_annotation(meta, now_ns):
    """updateLastCullingCheckTimestampAnnotation, :416-424."""
    meta["annotations"][LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION] = format_rfc3339(now_ns)


# Outcome codes of one culling pass (what CullingReconciler.Reconcile does
# between its client calls, :85-169).
CULL_REMOVED_STOPPED = 1     # stop annotation set: culling annotations removed, Update (:103-111)
CULL_REMOVED_NO_POD = 2      # pod missing: culling annotations removed, Update (:116-125)
CULL_WAIT = 3                # period not passed: RequeueAfter (:141-144)
CULL_CHECKED = 4             # annotations refreshed, Update, not idle (:147-153,:168)
CULL_CULLED = 5              # as CHECKED, then stop annotation set, second Update (:156-167)


def culling_step(nb, pod_exists, kernels, terminals, now_ns, cfg):
    """The pure part of CullingReconciler.Reconcile (:85-169) for one Notebook.

    Mutates nb["metadata"]; returns (outcome, n_updates, requeue_after_minutes or None).
    """
    meta = nb.setdefault("metadata", {})
    if stop_annotation_is_set(meta):
        remove_annotations(meta)
        return CULL_REMOVED_STOPPED, 1, None
    if not pod_exists:
        remove_annotations(meta)
        return CULL_REMOVED_NO_POD, 1, None
    updates = 0
    if not annotations_exist(meta):
        initialize_annotations(meta, now_ns)
        updates += 1
    if not culling_check_period_has_passed(meta, now_ns, cfg):
        return CULL_WAIT, updates, cfg["IDLENESS_CHECK_PERIOD"]
    update_notebook_last_activity_annotation(meta, kernels, terminals, now_ns)
    update_last_culling_check_timestamp_annotation(meta, now_ns)
    updates += 1
    if notebook_is_idle(meta, now_ns, cfg):
        set_stop_annotation(meta, now_ns)
        return CULL_CULLED, updates + 1, cfg["IDLENESS_CHECK_PERIOD"]
    return CULL_CHECKED, updates, cfg["IDLENESS_CHECK_PERIOD"]
