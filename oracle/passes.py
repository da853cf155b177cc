"""Oracle: the full reconcile pass kubeflow-rm_b200/reconcile.py composes, on
JSON trees (test infrastructure; also the CPU baseline bench.py times)."""
from . import nbc, odh

CTRL_NS = "redhat-ods-applications"


def full_pass(request, found, sts, pod, env, now_str, cookie="", image=odh.OAUTH_PROXY_IMAGE,
              ctrl_ns=CTRL_NS, odh_found=None):
    """request: the Notebook as submitted; found: [StatefulSet, Service,
    VirtualService] trees or None each.  Returns (mutated notebook, 7 ODH objects,
    3 create-or-update results, status).  With odh_found (7 trees or None each: what the
    ODH reconcilers Get) the second element is instead the 7 objects those reconcilers
    write (None = nothing to write), odh.write_decision."""
    mutated, _ = odh.mutate_notebook(request, "CREATE", image=image)
    odh_objs = odh.reconcile_objects(mutated, ctrl_ns, rbac=False, cookie_secret=cookie)
    if odh_found is not None:
        odh_objs = [odh.write_decision(d, f)[1] for d, f in zip(odh_objs, odh_found)]
    desired = [nbc.generate_statefulset(mutated, env), nbc.generate_service(mutated, env),
               nbc.generate_virtualservice(mutated, env)]
    writes = []
    for d, f in zip(desired, found):
        nbc.set_controller_reference(mutated, d)
        writes.append(nbc.create_or_update(d, f))
    status = nbc.create_notebook_status(mutated, sts or {}, pod or {}, now_str)
    return mutated, odh_objs, writes, status
