"""atn_trace_closest_static -- the closest-hit query for hosts that must not allocate device memory
(a device-memory budget that is already spent, or a frame loop that cannot stall on cudaMalloc).
It stages rays and hits in an array compiled into the library instead of buffers grown on demand.

Checked here, through the C ABI: two calls in a row give exactly what atn_trace_closest gives for the
same rays (the shared array keeps the previous call's contents, so the second call uses fewer and
different rays and a t_max the first one did not have), and a batch beyond the fixed capacity is
refused with a clear error instead of being truncated.

This file is named so that pytest collects it last: a failure here cannot hide another test's result
in a run that stops at the first failure.  It also passes when run alone in a fresh process."""
import numpy as np
import pytest

from aten_b200 import AtnError, PathTracing

pytestmark = pytest.mark.gpu


def _same(a, b):
    return a.tobytes() == b.tobytes()


def test_static_query_matches_allocating_query_on_two_calls(golden):
    scene = golden.scene("sponza_atrium")
    rays, tmax = golden.rays("sponza_atrium")
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    try:
        # call 1: all rays, no distance bound
        want1 = r.trace_closest(rays)
        got1 = r.trace_closest_static(rays)
        assert _same(got1, want1)
        assert (want1["objid"] >= 0).mean() > 0.5
        # call 2, straight after: fewer rays, reversed order, with per-ray bounds
        rays2 = np.ascontiguousarray(rays[::-1][:1000])
        tmax2 = np.ascontiguousarray(tmax[::-1][:1000])
        want2 = r.trace_closest(rays2, tmax2)
        got2 = r.trace_closest_static(rays2, tmax2)
        assert _same(got2, want2)
        # and the answers are the reference's (golden hits of the same rays, same bounds)
        ref = golden.hits("sponza_atrium")[::-1][:1000]
        agree = (got2["primid"] == np.where(ref["objid"] >= 0, ref["primid"], -1)).mean()
        assert agree > 0.999
    finally:
        r.close()


def test_static_query_refuses_batches_beyond_its_capacity(golden):
    scene = golden.scene("cornell")
    rays, _ = golden.rays("cornell")
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    try:
        too_many = np.resize(rays, PathTracing.STATIC_MAX_RAYS + 1)
        with pytest.raises(AtnError, match="exceed the static staging area"):
            r.trace_closest_static(too_many)
        full = np.resize(rays, PathTracing.STATIC_MAX_RAYS)          # exactly the capacity is fine
        assert _same(r.trace_closest_static(full), r.trace_closest(full))
        assert len(r.trace_closest_static(rays[:0])) == 0            # empty batch
    finally:
        r.close()
