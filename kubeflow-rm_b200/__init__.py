"""kubeflow-rm_b200 — B200 batch engine for the kubeflow-rm Notebook reconcile
and PodDefault admission hot path.

The compute lives in libnbrm.so (csrc/, CUDA for sm_100a) behind the C ABI in
include/nbrm.h.  This package is the thin Python binding used by the tests and
bench.py, plus the synthetic workloads of SURVEY.md section 8(d).  There is no
CPU fallback: importing works anywhere, but every call needs the built library
and a CUDA device.
"""
from .capi import (G_NBC, G_ODH, G_COPY, G_STATUS, G_CULL, G_MUTATE, G_GUARD, G_PODDEFAULT, G_ODH_DECIDE,  # noqa: F401
                   Context, Docs, Result, NbrmError, lib_path, load_library,  # noqa: F401
                   GEN_STS, GEN_SVC, GEN_VS, GEN_OWNER_REF, GEN_OWNER_V1, GEN_ALL, ODH_RBAC, MUT_CREATE, MUT_CA_BUNDLE,
                   scan_lengths_noalloc, SCAN_NOALLOC_MAX_N)
