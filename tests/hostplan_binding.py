"""ctypes binding of tests/hostplan/libhostplan.so (TEST INFRASTRUCTURE: the
planner headers compiled for the CPU; see hostplan.cc)."""
import ctypes as C

import numpy as np


class HostPlan:
    def __init__(self, path):
        self.lib = C.CDLL(path)

    def _call(self, fn, docs_list, env, flags, cap, out_per_doc, extra=()):
        import __graft_entry__ as graft
        capi = __import__(graft.PKG_NAME + ".capi", fromlist=["x"])
        data = b"".join(docs_list)
        off = np.zeros(len(docs_list) + 1, dtype=np.uint64)
        if docs_list:
            off[1:] = np.cumsum([len(d) for d in docs_list])
        e = None
        if env is not None:
            e = capi._Env()
            for k, v in env.items():
                setattr(e, capi.ENV_FIELDS[k], None if v is None else v.encode())
        pd, po, ps = C.c_void_p(), C.c_void_p(), C.c_void_p()
        fn.restype = C.c_int
        rc = fn(C.c_char_p(data), C.c_void_p(off.ctypes.data), C.c_uint64(len(docs_list)),
                C.byref(e) if e is not None else None, C.c_uint32(flags), C.c_uint32(cap), *extra,
                C.byref(pd), C.byref(po), C.byref(ps))
        assert rc == 0, rc
        n = len(docs_list) * out_per_doc
        o = np.frombuffer((C.c_uint64 * (n + 1)).from_address(po.value), dtype=np.uint64).copy()
        buf = C.string_at(pd.value, int(o[n])) if int(o[n]) else b""
        st = np.frombuffer((C.c_uint32 * max(len(docs_list), 1)).from_address(ps.value),
                           dtype=np.uint32, count=len(docs_list)).copy()
        return [buf[int(o[i]):int(o[i + 1])] for i in range(n)], st

    def nbc_generate(self, docs_list, env=None, flags=7, cap=96):
        return self._call(self.lib.hostplan_nbc_generate, docs_list, env, flags, cap, 3)
