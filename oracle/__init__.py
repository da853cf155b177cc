"""CPU oracle for the kubeflow-rm hot path.  TEST INFRASTRUCTURE ONLY.

This package is a plain-Python restatement of the reference's pure functions
on the Notebook reconcile path and the PodDefault admission path
(SURVEY.md section 8a).  It works on parsed JSON trees (dict / list / str /
int / bool / None) the way the Go code works on typed structs, and keeps the
Go nil-vs-empty distinctions that matter:

    absent key or JSON null  <->  Go nil slice / nil map / nil pointer
    [] or {}                 <->  Go empty, non-nil slice / map

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import it.  The product path (libnbrm.so and the
host code in kubeflow-rm_b200/) never does: it fails loudly without the CUDA
library.

Parity pinning.  The reference is Go and cannot be built in this image (no
Go toolchain, un-vendored modules).  The oracle is pinned against the
reference's own known-answer tables, transcribed into tests/golden/*.json by
tests/golden/make_golden.py:
  - status mirroring   notebook-controller/controllers/notebook_controller_test.go:103-275
  - culling predicates notebook-controller/controllers/culling_controller_test.go:16-249
  - ODH objects        odh-notebook-controller/controllers/notebook_controller_test.go:61-87,436-477,602-658,728-747,798-823,986-1119
  - mergeMap, setCommandAndArgs  admission-webhook/main_test.go:12-73,254-324
generateStatefulSet / generateService / generateVirtualService /
Copy*Fields / filterPodDefaults have no asserting test in the reference
(SURVEY.md 8c): for those the oracle follows the source text only and says
"parity unpinned by reference tests" where it matters.
"""
