"""ctypes binding of include/nbrm_host.h (libnbrm_host.so): the C++ object store
and NotebookReconciler."""
import ctypes as C
import json
import os

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
ERR = {0: "Ok", 1: "NotFound", 2: "AlreadyExists", 3: "Conflict", 4: "Invalid"}
_lib = None


def host_lib_path():
    return os.path.join(_HERE, "libnbrm_host.so")


def load_host_library():
    global _lib
    if _lib is not None:
        return _lib
    capi.load_library()
    p = host_lib_path()
    if not os.path.exists(p):
        raise capi.NbrmError("%s is not built (there is no CPU fallback)" % p)
    lib = C.CDLL(p)
    lib.nbrm_store_new.restype = C.c_void_p
    lib.nbrm_store_free.argtypes = [C.c_void_p]
    lib.nbrm_store_create.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
    lib.nbrm_store_update.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
    lib.nbrm_store_delete.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
    lib.nbrm_store_get.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_uint64)]
    lib.nbrm_store_counters.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int]
    lib.nbrm_store_size.argtypes = [C.c_void_p]
    lib.nbrm_store_size.restype = C.c_uint64
    lib.nbrm_store_free_string.argtypes = [C.c_void_p]
    lib.nbrm_reconciler_new.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.nbrm_reconciler_new.restype = C.c_void_p
    lib.nbrm_reconciler_free.argtypes = [C.c_void_p]
    lib.nbrm_reconciler_reconcile.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.nbrm_store_patch.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64, C.c_int]
    lib.nbrm_odh_reconciler_new.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    lib.nbrm_odh_reconciler_new.restype = C.c_void_p
    lib.nbrm_odh_reconciler_free.argtypes = [C.c_void_p]
    lib.nbrm_odh_reconciler_free.restype = None
    lib.nbrm_odh_reconciler_reconcile.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
    lib.nbrm_ca_bundle.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_void_p,
                                   C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.nbrm_pem_count_certificates.argtypes = [C.c_char_p, C.c_uint64]
    lib.nbrm_pem_count_certificates.restype = C.c_uint64
    lib.nbrm_unset_notebook_cert_config.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.nbrm_kfam_binding_name.argtypes = [C.c_char_p] * 4
    lib.nbrm_kfam_binding_name.restype = C.c_void_p
    lib.nbrm_workqueue_new.restype = C.c_void_p
    for fn in ("free", "add", "forget", "done"):
        getattr(lib, "nbrm_workqueue_" + fn).restype = None
    lib.nbrm_workqueue_free.argtypes = [C.c_void_p]
    lib.nbrm_workqueue_add.argtypes = lib.nbrm_workqueue_forget.argtypes = lib.nbrm_workqueue_done.argtypes = \
        [C.c_void_p, C.c_char_p, C.c_char_p]
    lib.nbrm_workqueue_add_after.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_double, C.c_double]
    lib.nbrm_workqueue_add_after.restype = None
    lib.nbrm_workqueue_add_rate_limited.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_double]
    lib.nbrm_workqueue_add_rate_limited.restype = C.c_double
    lib.nbrm_workqueue_len.argtypes = [C.c_void_p]
    lib.nbrm_workqueue_len.restype = C.c_uint64
    lib.nbrm_workqueue_next_ready.argtypes = [C.c_void_p]
    lib.nbrm_workqueue_next_ready.restype = C.c_double
    lib.nbrm_workqueue_take.argtypes = [C.c_void_p, C.c_uint64, C.c_double, C.POINTER(C.c_void_p)]
    lib.nbrm_workqueue_take.restype = C.c_uint64
    _lib = lib
    return lib


class Store:
    def __init__(self):
        self.lib = load_host_library()
        self.h = self.lib.nbrm_store_new()

    def create(self, json_bytes):
        return self.lib.nbrm_store_create(self.h, json_bytes, len(json_bytes))

    def update(self, json_bytes):
        return self.lib.nbrm_store_update(self.h, json_bytes, len(json_bytes))

    def delete(self, kind, ns, name):
        return self.lib.nbrm_store_delete(self.h, kind.encode(), ns.encode(), name.encode())

    def get(self, kind, ns, name):
        """Returns the object's JSON bytes or None (NotFound)."""
        out, n = C.c_void_p(), C.c_uint64()
        rc = self.lib.nbrm_store_get(self.h, kind.encode(), ns.encode(), name.encode(), C.byref(out), C.byref(n))
        if rc != 0:
            return None
        b = C.string_at(out.value, n.value)
        self.lib.nbrm_store_free_string(out)
        return b

    def patch(self, kind, ns, name, patch_bytes, json_patch=False):
        """RFC 7386 merge patch (default) or RFC 6902 JSON patch on a stored object."""
        return self.lib.nbrm_store_patch(self.h, kind.encode(), ns.encode(), name.encode(), patch_bytes,
                                         len(patch_bytes), 1 if json_patch else 0)

    def counters(self, reset=False):
        c = (C.c_uint64 * 6)()
        self.lib.nbrm_store_counters(self.h, c, 1 if reset else 0)
        return dict(zip(("gets", "creates", "updates", "status_updates", "deletes", "conflicts"), list(c)))

    def size(self):
        return int(self.lib.nbrm_store_size(self.h))

    def close(self):
        if self.h:
            self.lib.nbrm_store_free(self.h)
            self.h = None


STAT_NAMES = ("keys", "reconciled", "not_found", "terminating", "creates", "updates", "status_updates",
              "status_unchanged", "pod_restarts", "errors", "events_reissued")


class NotebookReconciler:
    """NotebookReconciler (notebook-controller/controllers/notebook_controller.go:75-81,90-272)."""

    def __init__(self, store, ctx, use_istio=True, status_write_if_changed=False):
        self.lib = load_host_library()
        self.store, self.ctx = store, ctx
        self.h = self.lib.nbrm_reconciler_new(store.h, ctx.h, 1 if use_istio else 0,
                                              1 if status_write_if_changed else 0)

    def reconcile(self, keys, now):
        """keys: iterable of (namespace, name).  Returns (n_errors, stats dict)."""
        text = "\n".join("%s/%s" % k for k in keys).encode()
        st = (C.c_uint64 * 11)()
        rc = self.lib.nbrm_reconciler_reconcile(self.h, text, now, st)
        return rc, dict(zip(STAT_NAMES, list(st)))

    def close(self):
        if self.h:
            self.lib.nbrm_reconciler_free(self.h)
            self.h = None


ODH_STAT_NAMES = ("keys", "reconciled", "not_found", "creates", "updates", "conflicts_retried", "locks_removed",
                  "pull_secret_waits", "errors")


class OpenshiftNotebookReconciler:
    """OpenshiftNotebookReconciler (odh-notebook-controller/controllers/notebook_controller.go:150-247)."""

    def __init__(self, store, ctx, rbac=False):
        self.lib = load_host_library()
        self.store, self.ctx = store, ctx
        self.h = self.lib.nbrm_odh_reconciler_new(store.h, ctx.h, 1 if rbac else 0)

    def reconcile(self, keys):
        """keys: iterable of (namespace, name).  Returns (n_errors, stats dict, simulated seconds slept)."""
        text = "\n".join("%s/%s" % k for k in keys).encode()
        st = (C.c_uint64 * 9)()
        slept = C.c_double()
        rc = self.lib.nbrm_odh_reconciler_reconcile(self.h, text, st, C.byref(slept))
        return rc, dict(zip(ODH_STAT_NAMES, list(st))), slept.value

    def close(self):
        if self.h:
            self.lib.nbrm_odh_reconciler_free(self.h)
            self.h = None


PROFILE_STAT_NAMES = ("keys", "reconciled", "not_found", "creates", "updates", "deletes", "rejected", "profile_writes",
                      "finalizers_added", "finalizers_removed", "plugins_applied", "plugins_revoked", "errors")


class ProfileReconciler:
    """ProfileReconciler (profile-controller/controllers/profile_controller.go:105-334)."""

    def __init__(self, store, ctx, workload_identity="", default_labels=None):
        self.lib = load_host_library()
        self.store, self.ctx = store, ctx
        self.lib.nbrm_profile_reconciler_new.restype = C.c_void_p
        self.lib.nbrm_profile_reconciler_new.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_char_p]
        self.lib.nbrm_profile_reconciler_reconcile.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        self.lib.nbrm_profile_reconciler_free.argtypes = [C.c_void_p]
        labels = None if default_labels is None else json.dumps(default_labels).encode()
        self.h = self.lib.nbrm_profile_reconciler_new(store.h, ctx.h, workload_identity.encode(), labels)

    def reconcile(self, names):
        st = (C.c_uint64 * 13)()
        rc = self.lib.nbrm_profile_reconciler_reconcile(self.h, "\n".join(names).encode(), st)
        return rc, dict(zip(PROFILE_STAT_NAMES, list(st)))

    def close(self):
        if self.h:
            self.lib.nbrm_profile_reconciler_free(self.h)
            self.h = None


class WorkQueue:
    """controller-runtime's workqueue semantics (kubeflow-rm_b200/host/workqueue.h)."""

    def __init__(self):
        self.lib = load_host_library()
        self.h = self.lib.nbrm_workqueue_new()

    def add(self, ns, name):
        self.lib.nbrm_workqueue_add(self.h, ns.encode(), name.encode())

    def add_after(self, ns, name, now, seconds):
        self.lib.nbrm_workqueue_add_after(self.h, ns.encode(), name.encode(), now, seconds)

    def add_rate_limited(self, ns, name, now):
        return self.lib.nbrm_workqueue_add_rate_limited(self.h, ns.encode(), name.encode(), now)

    def forget(self, ns, name):
        self.lib.nbrm_workqueue_forget(self.h, ns.encode(), name.encode())

    def done(self, ns, name):
        self.lib.nbrm_workqueue_done(self.h, ns.encode(), name.encode())

    def take(self, max_keys, now):
        out = C.c_void_p()
        n = self.lib.nbrm_workqueue_take(self.h, max_keys, now, C.byref(out))
        text = C.string_at(out.value).decode()
        self.lib.nbrm_store_free_string(out)
        keys = [tuple(line.split("/", 1)) for line in text.splitlines() if line]
        assert len(keys) == n
        return keys

    def __len__(self):
        return int(self.lib.nbrm_workqueue_len(self.h))

    def next_ready(self):
        return self.lib.nbrm_workqueue_next_ready(self.h)

    def close(self):
        if self.h:
            self.lib.nbrm_workqueue_free(self.h)
            self.h = None


def ca_bundle(odh_cm, kube_cm, x509_ok=None):
    """CreateNotebookCertConfigMap's bundle (nbrm_ca_bundle): odh_cm / kube_cm are the ConfigMaps'
    data dicts or None (NotFound).  Returns the bundle text or None."""
    lib = load_host_library()
    enc = lambda cm, k: None if cm is None or k not in cm else cm[k].encode("latin-1")
    ok = None
    if x509_ok is not None:
        ok = (C.c_int8 * 3)(*[(-1 if x509_ok.get(k) is None else int(x509_ok[k]))
                              for k in ("ca-bundle.crt", "odh-ca-bundle.crt", "ca.crt")])
    out, n = C.c_void_p(), C.c_uint64()
    rc = lib.nbrm_ca_bundle(enc(odh_cm, "ca-bundle.crt"), enc(odh_cm, "odh-ca-bundle.crt"), 0 if odh_cm is None else 1,
                            enc(kube_cm, "ca.crt"), 0 if kube_cm is None else 1, ok, C.byref(out), C.byref(n))
    if rc != 1:
        return None
    b = C.string_at(out.value, n.value).decode("latin-1")
    lib.nbrm_store_free_string(out)
    return b


def pem_count_certificates(bundle):
    b = bundle.encode("latin-1")
    return int(load_host_library().nbrm_pem_count_certificates(b, len(b)))


def unset_notebook_cert_config(nb_json):
    """(rc, bytes or None): nbrm_unset_notebook_cert_config."""
    lib = load_host_library()
    out, n = C.c_void_p(), C.c_uint64()
    rc = lib.nbrm_unset_notebook_cert_config(nb_json, len(nb_json), C.byref(out), C.byref(n))
    if rc != 1:
        return rc, None
    b = C.string_at(out.value, n.value)
    lib.nbrm_store_free_string(out)
    return rc, b


def kfam_binding_name(user_kind, user_name, roleref_kind, roleref_name):
    lib = load_host_library()
    p = lib.nbrm_kfam_binding_name(user_kind.encode(), user_name.encode(), roleref_kind.encode(), roleref_name.encode())
    out = C.string_at(p).decode()
    lib.nbrm_store_free_string(C.c_void_p(p))
    return out
