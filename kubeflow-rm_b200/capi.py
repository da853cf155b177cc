"""ctypes binding of include/nbrm.h (libnbrm.so)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

GEN_STS, GEN_SVC, GEN_VS, GEN_OWNER_REF, GEN_OWNER_V1 = 1, 2, 4, 8, 16
GEN_ALL = GEN_STS | GEN_SVC | GEN_VS
ODH_RBAC = 32
MUT_CREATE, MUT_CA_BUNDLE = 1, 2
(G_NBC, G_ODH, G_COPY, G_STATUS, G_CULL, G_MUTATE, G_GUARD, G_PODDEFAULT) = range(8)
G_ODH_DECIDE = 9

STATUS_NAMES = {0: "OK", 1: "E_PARSE", 2: "E_NO_CONTAINERS", 3: "E_SEGS", 4: "E_PORTS_EMPTY",
                5: "E_NIL_REPLICAS", 6: "E_DENIED", 7: "E_CONFLICT", 8: "E_LIMIT", 9: "E_SELECTOR", 10: "E_ENV"}


class NbrmError(RuntimeError):
    pass


class _Env(C.Structure):
    _fields_ = [(k, C.c_char_p) for k in (
        "use_istio", "istio_gateway", "istio_host", "cluster_domain", "add_fsgroup",
        "controller_namespace", "oauth_proxy_image", "userid_header", "userid_prefix",
        "notebook_controller_principal", "istio_ingress_gateway_principal", "kfp_ui_principal",
        "namespace_labels_json", "tensorboard_image", "rwo_pvc_scheduling", "pvcviewer_default_podspec_json")]


class _Docs(C.Structure):
    _fields_ = [("data", C.c_void_p), ("off", C.c_void_p), ("n", C.c_uint64)]


class _Out(C.Structure):
    _fields_ = [("data", C.c_void_p), ("off", C.c_void_p), ("n", C.c_uint64),
                ("status", C.c_void_p), ("n_status", C.c_uint64), ("result", C.c_void_p)]


class Profile(C.Structure):
    _fields_ = [("plan_ms", C.c_double), ("scan_ms", C.c_double), ("emit_ms", C.c_double),
                ("plan_launches", C.c_uint64), ("scan_launches", C.c_uint64),
                ("emit_launches", C.c_uint64), ("emit_bytes_read", C.c_uint64),
                ("emit_bytes_written", C.c_uint64), ("plan_bytes_read", C.c_uint64),
                ("plan_ms_gen", C.c_double * 32), ("emit_ms_gen", C.c_double * 32),
                ("plan_bytes_gen", C.c_uint64 * 32), ("emit_bytes_gen", C.c_uint64 * 32),
                ("runs_gen", C.c_uint64 * 32)]


# reference process environment (nbrm_env) key -> struct field
OPT_GROUP_BY_LENGTH = 1

ENV_FIELDS = {"USE_ISTIO": "use_istio", "ISTIO_GATEWAY": "istio_gateway", "ISTIO_HOST": "istio_host",
              "CLUSTER_DOMAIN": "cluster_domain", "ADD_FSGROUP": "add_fsgroup",
              "CONTROLLER_NAMESPACE": "controller_namespace", "OAUTH_PROXY_IMAGE": "oauth_proxy_image",
              "USERID_HEADER": "userid_header", "USERID_PREFIX": "userid_prefix",
              "NOTEBOOK_CONTROLLER_PRINCIPAL": "notebook_controller_principal",
              "ISTIO_INGRESS_GATEWAY_PRINCIPAL": "istio_ingress_gateway_principal",
              "KFP_UI_PRINCIPAL": "kfp_ui_principal", "NAMESPACE_LABELS_JSON": "namespace_labels_json",
              "TENSORBOARD_IMAGE": "tensorboard_image", "RWO_PVC_SCHEDULING": "rwo_pvc_scheduling",
              "PVCVIEWER_DEFAULT_PODSPEC_JSON": "pvcviewer_default_podspec_json"}


def lib_path():
    return os.path.join(_HERE, "libnbrm.so")


_lib = None


def load_library():
    """Load libnbrm.so; raise (never fall back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise NbrmError("%s is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)" % p)
    lib = C.CDLL(p)
    lib.nbrm_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.nbrm_ctx_destroy.argtypes = [C.c_void_p]
    lib.nbrm_ctx_destroy.restype = None
    lib.nbrm_last_error.argtypes = [C.c_void_p]
    lib.nbrm_last_error.restype = C.c_char_p
    lib.nbrm_ctx_set_env.argtypes = [C.c_void_p, C.POINTER(_Env)]
    lib.nbrm_ctx_set_option.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.nbrm_sync.argtypes = [C.c_void_p]
    lib.nbrm_host_alloc.argtypes = [C.c_size_t]
    lib.nbrm_host_alloc.restype = C.c_void_p
    lib.nbrm_host_free.argtypes = [C.c_void_p]
    lib.nbrm_host_free.restype = None
    lib.nbrm_timer_start.argtypes = [C.c_void_p]
    lib.nbrm_timer_stop.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    lib.nbrm_profile_enable.argtypes = [C.c_void_p, C.c_int]
    lib.nbrm_profile_read.argtypes = [C.c_void_p, C.POINTER(Profile)]
    lib.nbrm_launch_count.argtypes = [C.c_void_p]
    lib.nbrm_launch_count.restype = C.c_uint64
    lib.nbrm_upload.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Docs)]
    lib.nbrm_slot_index.argtypes = [C.c_void_p, C.c_int]
    lib.nbrm_notebook_convert.argtypes = [C.c_void_p, C.POINTER(_Docs), C.c_int, C.c_uint32, C.POINTER(_Out)]
    lib.nbrm_json_diff.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_notebook_webhook_review.argtypes = [C.c_void_p, C.POINTER(_Docs), C.c_uint32, C.c_int, C.POINTER(_Out)]
    lib.nbrm_poddefault_review.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.c_void_p, C.POINTER(_Out)]
    lib.nbrm_tape_dump.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.nbrm_download.argtypes = [C.c_void_p, C.POINTER(_Out)]
    lib.nbrm_nbc_generate.argtypes = [C.c_void_p, C.POINTER(_Docs), C.c_uint32, C.POINTER(_Out)]
    lib.nbrm_nbc_generate_resident.argtypes = [C.c_void_p, C.c_uint32]
    lib.nbrm_upload_bytes.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    lib.nbrm_odh_generate.argtypes = [C.c_void_p, C.POINTER(_Docs), C.c_char_p, C.c_uint32, C.POINTER(_Out)]
    lib.nbrm_odh_generate_resident.argtypes = [C.c_void_p, C.c_uint32]
    lib.nbrm_result_to_slot.argtypes = [C.c_void_p, C.c_int]
    lib.nbrm_copy_fields.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_copy_fields_resident.argtypes = [C.c_void_p]
    lib.nbrm_nbc_reconcile.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.c_uint32,
                                       C.POINTER(_Out)]
    lib.nbrm_nbc_reconcile_resident.argtypes = [C.c_void_p, C.c_uint32]
    lib.nbrm_notebook_status.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Docs),
                                         C.c_uint64, C.POINTER(_Out)]
    lib.nbrm_notebook_status_resident.argtypes = [C.c_void_p, C.c_uint64]
    lib.nbrm_culling_pass.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Docs),
                                      C.c_char_p, C.c_uint64, C.c_int32, C.c_int32, C.POINTER(_Out)]
    lib.nbrm_culling_pass_resident.argtypes = [C.c_void_p, C.c_uint64, C.c_int32, C.c_int32]
    lib.nbrm_odh_mutate.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.c_uint32, C.POINTER(_Out)]
    lib.nbrm_odh_mutate_resident.argtypes = [C.c_void_p, C.c_uint32, C.c_int]
    lib.nbrm_poddefault_mutate.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.c_void_p,
                                           C.POINTER(_Out)]
    lib.nbrm_poddefault_mutate_resident.argtypes = [C.c_void_p]
    lib.nbrm_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint64,
                             C.c_uint32, C.c_uint32, C.c_int]
    lib.nbrm_download_slot.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Out)]
    lib.nbrm_upload_async.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Docs)]
    lib.nbrm_download_begin.argtypes = [C.c_void_p, C.c_int]
    lib.nbrm_download_slot_begin.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.nbrm_download_wait.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Out)]
    lib.nbrm_profile_generate.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_profile_generate_resident.argtypes = [C.c_void_p]
    lib.nbrm_tensorboard_update.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_pvcviewer_status.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_pvcviewer_webhook.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_pvcviewer_reconcile.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Docs), C.POINTER(_Out)]
    lib.nbrm_tensorboard_generate.argtypes = [C.c_void_p, C.POINTER(_Docs), C.POINTER(_Docs), C.c_uint32, C.POINTER(_Out)]
    lib.nbrm_scan_lengths_noalloc.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                              C.c_char_p, C.c_size_t]
    _lib = lib
    return lib


SCAN_NOALLOC_MAX_N = 4194304


def scan_lengths_noalloc(d_lens, n, d_off, stream=None):
    """nbrm_scan_lengths_noalloc on raw DEVICE pointers (ints).  Returns (rc, message)."""
    lib = load_library()
    err = C.create_string_buffer(256)
    rc = lib.nbrm_scan_lengths_noalloc(d_lens, n, d_off, stream, err, len(err))
    return rc, err.value.decode("utf-8", "replace")


class Docs:
    """A packed batch of JSON documents in host memory (optionally pinned)."""

    def __init__(self, docs=None, pinned=False, data=None, off=None):
        self._pin = []
        if docs is not None:
            lens = np.fromiter((len(d) for d in docs), dtype=np.uint64, count=len(docs))
            off = np.zeros(len(docs) + 1, dtype=np.uint64)
            np.cumsum(lens, out=off[1:])
            data = np.frombuffer(b"".join(docs), dtype=np.uint8)
        self.n = len(off) - 1
        self.nbytes = int(off[-1]) if self.n else 0
        if pinned:
            lib = load_library()
            self.data = self._pinned_copy(lib, np.ascontiguousarray(data, dtype=np.uint8))
            self.off = self._pinned_copy(lib, np.ascontiguousarray(off, dtype=np.uint64))
        else:
            self.data = np.ascontiguousarray(data, dtype=np.uint8)
            self.off = np.ascontiguousarray(off, dtype=np.uint64)

    def _pinned_copy(self, lib, a):
        nbytes = max(a.nbytes, 1)
        p = lib.nbrm_host_alloc(nbytes)
        if not p:
            raise NbrmError("nbrm_host_alloc(%d) failed" % nbytes)
        self._pin.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p)
        out = np.frombuffer(buf, dtype=a.dtype, count=a.size)
        out[...] = a
        return out

    def c_struct(self):
        return _Docs(self.data.ctypes.data, self.off.ctypes.data, self.n)

    def __del__(self):
        if self._pin and _lib is not None:
            for p in self._pin:
                _lib.nbrm_host_free(p)
            self._pin = []


class Result:
    """Copy-out view of an nbrm_out."""

    def __init__(self, out, copy=True):
        n = int(out.n)
        off = np.frombuffer((C.c_uint64 * (n + 1)).from_address(out.off), dtype=np.uint64)
        total = int(off[n]) if n else 0
        data = np.frombuffer((C.c_uint8 * max(total, 1)).from_address(out.data), dtype=np.uint8,
                             count=total)
        ns = int(out.n_status) if out.status else 0
        st = np.frombuffer((C.c_uint32 * max(ns, 1)).from_address(out.status), dtype=np.uint32,
                           count=ns) if out.status else np.zeros(0, dtype=np.uint32)
        rs = np.frombuffer((C.c_uint32 * max(ns, 1)).from_address(out.result), dtype=np.uint32,
                           count=ns) if out.result else np.zeros(ns, dtype=np.uint32)
        self.result = rs.copy() if copy else rs
        self.off = off.copy() if copy else off
        self.data = data.copy() if copy else data
        self.status = st.copy() if copy else st
        self.n = n

    def doc(self, i):
        return self.data[int(self.off[i]):int(self.off[i + 1])].tobytes()

    def docs(self):
        b = self.data.tobytes()
        o = self.off
        return [b[int(o[i]):int(o[i + 1])] for i in range(self.n)]


class Context:
    """One nbrm_ctx: a CUDA stream plus its device buffers on one GPU."""

    def __init__(self, device=0, env=None):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.nbrm_ctx_create(device, C.byref(h))
        if rc != 0 or not h:
            raise NbrmError("nbrm_ctx_create(device=%d) failed (rc=%d): libnbrm needs a CUDA "
                            "device, there is no CPU fallback" % (device, rc))
        self.h = h
        if env is not None:
            self.set_env(env)

    def close(self):
        if getattr(self, "h", None):
            self.lib.nbrm_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def _check(self, rc):
        if rc < 0:
            raise NbrmError(self.lib.nbrm_last_error(self.h).decode("utf-8", "replace"))
        return rc

    def set_env(self, env):
        e = _Env()
        for k, v in env.items():
            setattr(e, ENV_FIELDS[k], None if v is None else v.encode())
        self._check(self.lib.nbrm_ctx_set_env(self.h, C.byref(e)))

    def set_option(self, option, value):
        """nbrm_ctx_set_option; option: OPT_GROUP_BY_LENGTH."""
        self._check(self.lib.nbrm_ctx_set_option(self.h, int(option), int(value)))

    def sync(self):
        self._check(self.lib.nbrm_sync(self.h))

    def upload(self, slot, docs):
        d = docs.c_struct()
        self._check(self.lib.nbrm_upload(self.h, slot, C.byref(d)))

    def upload_async(self, slot, docs):
        """nbrm_upload on the upload stream; `docs` must be pinned (Docs(..., pinned=True))
        and alive until a generator that reads the slot has run and been synchronised."""
        d = docs.c_struct()
        self._check(self.lib.nbrm_upload_async(self.h, slot, C.byref(d)))

    def download(self, copy=True):
        o = _Out()
        self._check(self.lib.nbrm_download(self.h, C.byref(o)))
        return Result(o, copy)

    def timer_start(self):
        self._check(self.lib.nbrm_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self._check(self.lib.nbrm_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def profile_enable(self, on=True):
        self._check(self.lib.nbrm_profile_enable(self.h, 1 if on else 0))

    def profile_read(self):
        p = Profile()
        self._check(self.lib.nbrm_profile_read(self.h, C.byref(p)))
        return p

    def launch_count(self):
        return int(self.lib.nbrm_launch_count(self.h))

    # --- notebook-controller generators
    def nbc_generate(self, docs, flags=GEN_ALL, copy=True):
        d = docs.c_struct()
        o = _Out()
        self._check(self.lib.nbrm_nbc_generate(self.h, C.byref(d), flags, C.byref(o)))
        return Result(o, copy)

    def nbc_generate_resident(self, flags=GEN_ALL):
        self._check(self.lib.nbrm_nbc_generate_resident(self.h, flags))

    # --- ODH reconcile generators
    def odh_generate(self, docs, seeds=None, flags=GEN_OWNER_REF | GEN_OWNER_V1, copy=True):
        d = docs.c_struct()
        o = _Out()
        self._check(self.lib.nbrm_odh_generate(self.h, C.byref(d), seeds, flags, C.byref(o)))
        return Result(o, copy)

    def upload_bytes(self, data):
        self._check(self.lib.nbrm_upload_bytes(self.h, data, len(data) if data else 0))

    def odh_generate_resident(self, flags=GEN_OWNER_REF | GEN_OWNER_V1):
        self._check(self.lib.nbrm_odh_generate_resident(self.h, flags))

    # --- create-or-update decision / fused reconcile
    def result_to_slot(self, slot):
        self._check(self.lib.nbrm_result_to_slot(self.h, slot))

    def copy_fields(self, desired, found, copy=True):
        d, f, o = desired.c_struct(), found.c_struct(), _Out()
        self._check(self.lib.nbrm_copy_fields(self.h, C.byref(d), C.byref(f), C.byref(o)))
        return Result(o, copy)

    def copy_fields_resident(self):
        self._check(self.lib.nbrm_copy_fields_resident(self.h))

    def nbc_reconcile(self, notebooks, found, flags=GEN_ALL | GEN_OWNER_REF, copy=True):
        d, f, o = notebooks.c_struct(), found.c_struct(), _Out()
        self._check(self.lib.nbrm_nbc_reconcile(self.h, C.byref(d), C.byref(f), flags, C.byref(o)))
        return Result(o, copy)

    def nbc_reconcile_resident(self, flags=GEN_ALL | GEN_OWNER_REF):
        self._check(self.lib.nbrm_nbc_reconcile_resident(self.h, flags))

    # --- status mirroring / culling
    def notebook_status(self, notebooks, stss, pods, now, copy=True):
        a, b, d, o = notebooks.c_struct(), stss.c_struct(), pods.c_struct(), _Out()
        self._check(self.lib.nbrm_notebook_status(self.h, C.byref(a), C.byref(b), C.byref(d), now, C.byref(o)))
        return Result(o, copy)

    def culling_pass(self, notebooks, kernels, terminals, pod_exists, now, idle_min, period_min, copy=True):
        a, b, d, o = notebooks.c_struct(), kernels.c_struct(), terminals.c_struct(), _Out()
        self._check(self.lib.nbrm_culling_pass(self.h, C.byref(a), C.byref(b), C.byref(d), pod_exists, now,
                                               idle_min, period_min, C.byref(o)))
        return Result(o, copy)

    # --- ODH mutating webhook
    def odh_mutate(self, notebooks, old=None, flags=0, copy=True):
        d, o = notebooks.c_struct(), _Out()
        od = old.c_struct() if old is not None else None
        self._check(self.lib.nbrm_odh_mutate(self.h, C.byref(d), C.byref(od) if od is not None else None,
                                             flags if old is not None else flags | MUT_CREATE, C.byref(o)))
        return Result(o, copy)

    # --- PodDefault admission
    def poddefault_mutate(self, pods, poddefaults, ranges=None, copy=True):
        a, b, o = pods.c_struct(), poddefaults.c_struct(), _Out()
        r = None
        if ranges is not None:
            r = np.ascontiguousarray(np.asarray(ranges, dtype=np.uint32).reshape(-1))
        self._check(self.lib.nbrm_poddefault_mutate(self.h, C.byref(a), C.byref(b),
                                                    r.ctypes.data if r is not None else None, C.byref(o)))
        return Result(o, copy)

    def notebook_convert(self, notebooks, to_version, flags=0, copy=True):
        """ConvertTo / ConvertFrom (0 v1beta1, 1 v1, 2 v1alpha1)."""
        a, o = notebooks.c_struct(), _Out()
        self._check(self.lib.nbrm_notebook_convert(self.h, C.byref(a), to_version, flags, C.byref(o)))
        return Result(o, copy)

    # --- wire formats: JSON patches and AdmissionReview responses
    def json_diff(self, originals, mutated, copy=True):
        a, b, o = originals.c_struct(), mutated.c_struct(), _Out()
        self._check(self.lib.nbrm_json_diff(self.h, C.byref(a), C.byref(b), C.byref(o)))
        return Result(o, copy)

    def notebook_webhook_review(self, reviews, flags=MUT_CREATE, is_update=False, copy=True):
        a, o = reviews.c_struct(), _Out()
        self._check(self.lib.nbrm_notebook_webhook_review(self.h, C.byref(a), flags, 1 if is_update else 0, C.byref(o)))
        return Result(o, copy)

    def poddefault_review(self, reviews, poddefaults, ranges=None, copy=True):
        a, b, o = reviews.c_struct(), poddefaults.c_struct(), _Out()
        r = None
        if ranges is not None:
            r = np.ascontiguousarray(np.asarray(ranges, dtype=np.uint32).reshape(-1))
        self._check(self.lib.nbrm_poddefault_review(self.h, C.byref(a), C.byref(b),
                                                    r.ctypes.data if r is not None else None, C.byref(o)))
        return Result(o, copy)

    def poddefault_mutate_resident(self):
        self._check(self.lib.nbrm_poddefault_mutate_resident(self.h))

    # --- composing generators over arbitrary slots
    def run(self, gen, a, b=-1, c=-1, flags=0, now=0, arg0=0, arg1=0, side=False):
        self._check(self.lib.nbrm_run(self.h, gen, a, b, c, flags, now, arg0, arg1, 1 if side else 0))

    def download_slot(self, slot, copy=True):
        o = _Out()
        self._check(self.lib.nbrm_download_slot(self.h, slot, C.byref(o)))
        return Result(o, copy)

    def slot_index(self, slot):
        """Index the slot's documents once (nbrm_slot_index): for documents that stay resident
        while many batches are planned against them."""
        self._check(self.lib.nbrm_slot_index(self.h, slot))

    def tape_dump(self, slot, n, cap):
        """The structural index of every document of `slot` as the plan kernels build it
        (nbrm_tape_dump; n = documents in the slot): (entries uint64[n, cap], counts uint32[n],
        flags uint32[n])."""
        import numpy as np
        ent = np.zeros((n, cap), dtype=np.uint64)
        cnt = np.zeros(n, dtype=np.uint32)
        flg = np.zeros(n, dtype=np.uint32)
        self._check(self.lib.nbrm_tape_dump(self.h, slot, cap, ent.ctypes.data, cnt.ctypes.data, flg.ctypes.data))
        return ent, cnt, flg

    def odh_decide(self, desired, found, copy=True):
        """ODH write decision for (desired, found) pairs: result word 0 nothing, 1 create, 2 update."""
        self.upload(0, desired)
        self.upload(1, found)
        self.run(G_ODH_DECIDE, 0, 1)
        return self.download(copy)

    # --- downloads that overlap the next generator (one at a time per lane, 0..3)
    def download_begin(self, lane):
        self._check(self.lib.nbrm_download_begin(self.h, lane))

    def download_slot_begin(self, slot, lane):
        self._check(self.lib.nbrm_download_slot_begin(self.h, slot, lane))

    def download_wait(self, lane, copy=True):
        o = _Out()
        self._check(self.lib.nbrm_download_wait(self.h, lane, C.byref(o)))
        return Result(o, copy)

    # --- pvcviewer-controller: the bodies a reconcile writes
    def pvcviewer_reconcile(self, viewers, found=None, node_names=None, copy=True):
        """Deployment, Service, VirtualService per PVCViewer (include/nbrm.h); found: 3 per viewer."""
        d, o = viewers.c_struct(), _Out()
        f = found.c_struct() if found is not None else None
        nn = node_names.c_struct() if node_names is not None else None
        self._check(self.lib.nbrm_pvcviewer_reconcile(self.h, C.byref(d), C.byref(f) if f is not None else None,
                                                      C.byref(nn) if nn is not None else None, C.byref(o)))
        return Result(o, copy)

    def tensorboard_update(self, tensorboards, desired, found, copy=True):
        """Per Tensorboard: the Deployment Update body (or b"") and the status (include/nbrm.h)."""
        a, b, d, o = tensorboards.c_struct(), desired.c_struct(), found.c_struct(), _Out()
        self._check(self.lib.nbrm_tensorboard_update(self.h, C.byref(a), C.byref(b), C.byref(d), C.byref(o)))
        return Result(o, copy)

    def pvcviewer_status(self, viewers, deployments, copy=True):
        a, b, o = viewers.c_struct(), deployments.c_struct(), _Out()
        self._check(self.lib.nbrm_pvcviewer_status(self.h, C.byref(a), C.byref(b), C.byref(o)))
        return Result(o, copy)

    def pvcviewer_webhook(self, viewers, copy=True):
        """Default + validate of the viewer's webhook: defaulted viewers, result = validation code."""
        d, o = viewers.c_struct(), _Out()
        self._check(self.lib.nbrm_pvcviewer_webhook(self.h, C.byref(d), C.byref(o)))
        return Result(o, copy)

    # --- tensorboard-controller objects
    def tensorboard_generate(self, tensorboards, node_names=None, flags=0, copy=True):
        """Deployment, Service, VirtualService per Tensorboard CR (include/nbrm.h)."""
        d, o = tensorboards.c_struct(), _Out()
        nn = node_names.c_struct() if node_names is not None else None
        self._check(self.lib.nbrm_tensorboard_generate(self.h, C.byref(d), C.byref(nn) if nn is not None else None,
                                                       flags, C.byref(o)))
        return Result(o, copy)

    # --- profile-controller objects
    def profile_generate(self, profiles, copy=True):
        d, o = profiles.c_struct(), _Out()
        self._check(self.lib.nbrm_profile_generate(self.h, C.byref(d), C.byref(o)))
        return Result(o, copy)
