"""GPU: Profile reconcile objects through the C ABI against the oracle (BASELINE config 4: 500 Profiles)."""
import random

import pytest

from parity import check_profile
from test_hostplan_profile import ENVS

pytestmark = pytest.mark.gpu


def test_500_profiles(ctx, pkg, wl, hostplan):
    rng = random.Random(4)
    profiles = [wl.profile_cr(rng, i) for i in range(500)]
    docs = [wl.dumps(p, rng) for p in profiles]
    for env, labels in ENVS:
        ctx.set_env(env)
        res = ctx.profile_generate(pkg.Docs(docs))
        oenv = {k: v for k, v in env.items() if k != "NAMESPACE_LABELS_JSON"}
        assert check_profile(profiles, res.docs(), res.status, oenv, labels) > 3500
        outs, st = hostplan.profile(docs, env)
        assert outs == res.docs()
    ctx.set_env({})
