"""ctypes binding of tests/hostplan/libhostplan.so (TEST INFRASTRUCTURE: the
planner headers compiled for the CPU; see hostplan.cc)."""
import ctypes as C

import numpy as np

GEN_NBC, GEN_ODH, GEN_COPY, GEN_STATUS, GEN_CULL, GEN_MUTATE, GEN_GUARD, GEN_PD, GEN_PROFILE = 0, 1, 2, 3, 4, 5, 6, 7, 8
GEN_ODH_DECIDE = 9
GEN_DIFF, GEN_REVIEW_OPEN, GEN_REVIEW_CLOSE = 10, 11, 12
GEN_TENSORBOARD = 14
GEN_PVCVIEWER = 15
GEN_PVCVIEWER_DEFAULT = 16
GEN_TB_UPDATE, GEN_PV_STATUS = 17, 18


def _pack(docs_list):
    data = b"".join(docs_list)
    off = np.zeros(len(docs_list) + 1, dtype=np.uint64)
    if docs_list:
        off[1:] = np.cumsum([len(d) for d in docs_list])
    return data, off


class HostPlan:
    def __init__(self, path):
        self.lib = C.CDLL(path)
        self.lib.hostplan_run.restype = C.c_int

    def run(self, gen, slots, env=None, flags=0, extra=None, now=0, arg0=0, arg1=0, want_result=False, shared1=False):
        """slots: up to three lists of byte strings (None = slot unused)."""
        import __graft_entry__ as graft
        capi = __import__(graft.PKG_NAME + ".capi", fromlist=["x"])
        slots = list(slots) + [None] * (3 - len(slots))
        n = len(next(s for s in slots if s is not None))
        shared_n = len(slots[1]) if shared1 else 0
        args, keep = [], []
        for k, s in enumerate(slots):
            if s is None:
                args += [None, None]
            else:
                assert len(s) == n or (shared1 and k == 1)
                data, off = _pack(s)
                keep.append((data, off))
                args += [C.c_char_p(data), C.c_void_p(off.ctypes.data)]
        e = None
        if env is not None:
            e = capi._Env()
            for k, v in env.items():
                setattr(e, capi.ENV_FIELDS[k], None if v is None else v.encode())
        pd, po, ps, pr, nobj = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_uint32()
        rc = self.lib.hostplan_run(C.c_int(gen), *args, C.c_uint64(n),
                                   C.byref(e) if e is not None else None, C.c_uint32(flags),
                                   C.c_char_p(extra) if extra is not None else None, C.c_uint64(now),
                                   C.c_uint32(arg0), C.c_uint32(arg1), C.c_uint32(shared_n),
                                   C.byref(nobj), C.byref(pd), C.byref(po), C.byref(ps), C.byref(pr))
        assert rc == 0, rc
        m = n * nobj.value
        o = np.frombuffer((C.c_uint64 * (m + 1)).from_address(po.value), dtype=np.uint64).copy()
        buf = C.string_at(pd.value, int(o[m])) if int(o[m]) else b""
        st = np.frombuffer((C.c_uint32 * max(n, 1)).from_address(ps.value), dtype=np.uint32,
                           count=n).copy()
        outs = [buf[int(o[i]):int(o[i + 1])] for i in range(m)]
        if want_result:
            res = np.frombuffer((C.c_uint32 * max(n, 1)).from_address(pr.value), dtype=np.uint32,
                                count=n).copy()
            return outs, st, res
        return outs, st

    def nbc_generate(self, docs_list, env=None, flags=7):
        if not docs_list:
            return [], np.zeros(0, dtype=np.uint32)
        return self.run(GEN_NBC, [docs_list], env, flags)

    def odh_generate(self, docs_list, seeds=None, env=None, flags=0):
        if not docs_list:
            return [], np.zeros(0, dtype=np.uint32)
        return self.run(GEN_ODH, [docs_list], env, flags, extra=seeds)

    def copy_fields(self, desired_list, found_list):
        """found_list entries may be b"" (NotFound)."""
        if not desired_list:
            return [], np.zeros(0, dtype=np.uint32)
        return self.run(GEN_COPY, [desired_list, found_list])

    def status(self, notebooks, stss, pods, now=0):
        return self.run(GEN_STATUS, [notebooks, stss, pods], now=now)

    def cull(self, notebooks, kernels, terminals, pod_exists, now, idle_min, period_min):
        return self.run(GEN_CULL, [notebooks, kernels, terminals], extra=bytes(pod_exists), now=now,
                        arg0=idle_min, arg1=period_min, want_result=True)

    def mutate(self, notebooks, env=None, flags=1):
        return self.run(GEN_MUTATE, [notebooks], env, flags)

    def restart_guard(self, mutated, updated, old, env=None):
        return self.run(GEN_GUARD, [mutated, updated, old], env, want_result=True)

    def poddefault(self, pods, pds, ranges=None):
        extra = None
        if ranges is not None:
            extra = np.asarray(ranges, dtype=np.uint32).reshape(-1).tobytes()
        return self.run(GEN_PD, [pods, pds], extra=extra, want_result=True, shared1=True)

    def odh_decide(self, desired_list, found_list):
        if not desired_list:
            return [], np.zeros(0, dtype=np.uint32), np.zeros(0, dtype=np.uint32)
        return self.run(GEN_ODH_DECIDE, [desired_list, found_list], want_result=True)

    def pvcviewer(self, viewers, found=None, node_names=None, env=None):
        return self.run(GEN_PVCVIEWER, [viewers, found if found is not None else [], node_names], env, shared1=True)

    def tensorboard_update(self, tensorboards, desired, found):
        return self.run(GEN_TB_UPDATE, [tensorboards, desired, found], want_result=True)

    def pvcviewer_status(self, viewers, deployments):
        return self.run(GEN_PV_STATUS, [viewers, deployments])

    def pvcviewer_webhook(self, viewers, env=None):
        return self.run(GEN_PVCVIEWER_DEFAULT, [viewers], env, want_result=True)

    def tensorboard(self, tensorboards, node_names=None, env=None, flags=0):
        return self.run(GEN_TENSORBOARD, [tensorboards, node_names], env, flags=flags)

    def profile(self, profiles, env=None):
        return self.run(GEN_PROFILE, [profiles], env)

    def json_diff(self, originals, mutated):
        return self.run(GEN_DIFF, [originals, mutated])

    def review_open(self, reviews, flags=0):
        return self.run(GEN_REVIEW_OPEN, [reviews], flags=flags, want_result=True)

    def review_close(self, reviews, patches, statuses, flags=0):
        import numpy as np
        return self.run(GEN_REVIEW_CLOSE, [reviews, patches], flags=flags,
                        extra=np.asarray(statuses, dtype=np.uint32).tobytes())

    def convert(self, notebooks, to_version, flags=0):
        return self.run(13, [notebooks], flags=flags, arg0=to_version)
