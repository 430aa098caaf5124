"""ptb_raycast_noalloc (include/ptb.h; no counterpart in the reference): the
ray-cast query through the library's file-scope static device scratch instead of per-call cudaMalloc'd staging.  Checked
through the C ABI against the existing entry point ptb_raycast - same kernel, so the hit records must be byte-identical -
on two calls in a row that share the scratch (the second smaller, so stale contents of the first would show), in both
traversal backends, plus the size limit.  Collected last (test_zz_*) so that a failure here cannot hide other tests."""
import numpy as np
import pytest

from util import random_rays

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def I():
    from path_tracer_b200 import integrator
    return integrator


def test_noalloc_raycast_equals_raycast_on_two_calls_in_a_row(I, gpu, oracle, scenes):
    limit = I.lib().ptb_raycast_noalloc_max_rays()
    assert limit == 65536
    for name, params in (("cornell", ()), ("terrain", (64,))):
        sc = scenes(name, *params)
        d = I.DeviceScene(gpu, sc)
        first = random_rays(sc, limit, 41, oracle)             # fills the whole scratch
        second = random_rays(sc, 5000, 42, oracle)             # a shorter batch over the first one's leftovers
        for rays in (first, second):
            want = d.raycast(rays)
            got = d.raycast_noalloc(rays)
            assert got.tobytes() == want.tobytes()
            assert (want["shape"] != 0xFFFFFFFF).mean() > 0.2    # an actual test: plenty of hits
        # the reference-layout traversal (one thread per ray, no fetch counter) through the same scratch
        want = d.raycast(second, I.TRAVERSAL_BVH2_REFERENCE, I.TIE_FAITHFUL)
        got = d.raycast_noalloc(second, I.TRAVERSAL_BVH2_REFERENCE, I.TIE_FAITHFUL)
        assert got.tobytes() == want.tobytes()
        # and against the oracle directly, so the check does not rest on ptb_raycast alone
        bf = oracle.raycast_bruteforce(sc.view, second[:500])
        got = d.raycast_noalloc(second[:500])
        assert np.array_equal(got["shape"], bf["shape"]) and np.array_equal(got["prim"], bf["prim"])
        assert np.array_equal(got["t"].view(np.uint32), bf["t"].view(np.uint32))
        assert len(d.raycast_noalloc(second[:0])) == 0           # empty batch
        with pytest.raises(I.PtbError) as e:                     # one ray too many: refused, with the limit in the message
            d.raycast_noalloc(np.concatenate([first, second[:1]]))
        assert e.value.code == -1 and "65536" in str(e.value)
        with pytest.raises(I.PtbError):
            d.raycast_noalloc(second, I.TRAVERSAL_WIDE, I.TIE_FAITHFUL)
