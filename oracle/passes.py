"""Oracle: the full reconcile pass kubeflow-rm_b200/reconcile.py composes, on
JSON trees (test infrastructure; also the CPU baseline bench.py times)."""
from . import nbc, odh

CTRL_NS = "redhat-ods-applications"


def full_pass(request, found, sts, pod, env, now_str, cookie="", image=odh.OAUTH_PROXY_IMAGE,
              ctrl_ns=CTRL_NS):
    """request: the Notebook as submitted; found: [StatefulSet, Service,
    VirtualService] trees or None each.  Returns (mutated notebook, 7 ODH objects,
    3 create-or-update results, status)."""
    mutated, _ = odh.mutate_notebook(request, "CREATE", image=image)
    odh_objs = odh.reconcile_objects(mutated, ctrl_ns, rbac=False, cookie_secret=cookie)
    desired = [nbc.generate_statefulset(mutated, env), nbc.generate_service(mutated, env),
               nbc.generate_virtualservice(mutated, env)]
    writes = []
    for d, f in zip(desired, found):
        nbc.set_controller_reference(mutated, d)
        writes.append(nbc.create_or_update(d, f))
    status = nbc.create_notebook_status(mutated, sts or {}, pod or {}, now_str)
    return mutated, odh_objs, writes, status
