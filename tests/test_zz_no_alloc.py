"""paths_trace_closest_static: the f32 per-ray seam with all of its device scratch in a static device array (run on the B200 box, -m gpu).

The entry point has no counterpart in the reference; it exists for callers that may not call cudaMalloc (include/paths_b200.h).  It runs
the very kernels `paths_trace_closest(precision=32)` runs, so the bar is the one the production-ray tests hold the f32 kernels to among
themselves: primitive ids, distances and barycentrics BIT-IDENTICAL to the allocating entry point's, on two calls in a row that share
the scratch array (the second one shorter and with different rays, so that anything left over from the first would show), then a
full-size batch, an oversize batch (refused, limit in the message) and a scene that was never staged (refused, nothing allocated).

Collected last on purpose; self-contained, so it also passes alone in a fresh process.
"""
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path")
    return paths_b200


def ray_sets(pb, gsc):
    """Two different batches: 64 000 coherent camera rays, then 40 000 incoherent rays leaving the surfaces those hit."""
    o, d, _ = pb.camera_rays(gsc.camera, 11, 0, precision=pb.PRECISION_F32)
    assert len(o) == 320 * 200
    prim, t = gsc.find_intersection(o, d, precision=pb.PRECISION_F32)  # also stages the scene on the device (staging allocates)
    rng = np.random.default_rng(23)
    hit = np.nonzero(prim != 0xFFFFFFFF)[0]
    assert len(hit) > 1000
    sel = rng.choice(hit, size=40_000, replace=True)
    d2 = rng.normal(size=(len(sel), 3))
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o2 = o[sel] + d[sel] * t[sel, None] + d2 * 1e-3
    return (o, d), (o2, d2)


def same_answers(pb, gsc, o, d):
    p0, t0, b0 = gsc.find_intersection(o, d, precision=pb.PRECISION_F32, want_bary=True)
    p1, t1, b1 = gsc.find_intersection_static(o, d, want_bary=True)
    assert (p0 != 0xFFFFFFFF).any()
    assert np.array_equal(p1, p0), f"{int((p1 != p0).sum())} of {len(p0)} primitive ids differ"
    assert np.array_equal(t1, t0), "distances differ"   # inf on a miss in both
    assert np.array_equal(b1, b0), "barycentrics differ"
    # without the optional output
    p2, t2 = gsc.find_intersection_static(o, d)
    assert np.array_equal(p2, p0) and np.array_equal(t2, t0)


@pytest.mark.parametrize("name", ["teapot", "dragon"])  # two-wide lane-refill kernel / four-wide parked-leaf kernel
def test_static_scratch_entry_point_matches_the_allocating_one(pb, name):
    if name == "dragon":
        sys.path.insert(0, str(ROOT / "tools"))
        import make_assets

        make_assets.ensure("dragon")
    gsc = pb.Scene.load(f"scenes/{name}.yml", width=320, height=200)
    first, second = ray_sets(pb, gsc)
    # two calls in a row over the shared array: a long batch, then a shorter one with other rays
    same_answers(pb, gsc, *first)
    same_answers(pb, gsc, *second)
    # and the first again: nothing of the second may survive
    same_answers(pb, gsc, *first)
    # exactly the maximum
    n_max = pb.Scene.STATIC_MAX_RAYS
    reps = -(-n_max // len(first[0]))
    o = np.tile(first[0], (reps, 1))[:n_max]
    d = np.tile(first[1], (reps, 1))[:n_max]
    same_answers(pb, gsc, o, d)
    # one more than the maximum: refused with the limit spelled out, outputs untouched
    o1, d1 = np.vstack([o, o[:1]]), np.vstack([d, d[:1]])
    with pytest.raises(pb.PathsError) as e:
        gsc.find_intersection_static(o1, d1)
    assert e.value.status == pb._capi.ERR_INVALID_ARG and str(n_max) in str(e.value)
    # an empty batch is fine, and the entry point still works after the refusal
    p, t = gsc.find_intersection_static(np.zeros((0, 3)), np.zeros((0, 3)))
    assert len(p) == 0 and len(t) == 0
    same_answers(pb, gsc, *second)


def test_static_scratch_entry_point_refuses_an_unstaged_scene(pb):
    gsc = pb.Scene.load("scenes/teapot.yml", width=320, height=200)  # a fresh handle: staged on no device yet
    o = np.array([[0.0, 1.0, 5.0]])
    d = np.array([[0.0, 0.0, -1.0]])
    with pytest.raises(pb.PathsError) as e:
        gsc.find_intersection_static(o, d)
    assert e.value.status == pb._capi.ERR_UNSUPPORTED and "staged" in str(e.value)
    gsc.find_intersection(o, d, precision=pb.PRECISION_F32)  # stages it
    p0, t0 = gsc.find_intersection(o, d, precision=pb.PRECISION_F32)
    p1, t1 = gsc.find_intersection_static(o, d)
    assert np.array_equal(p0, p1) and np.array_equal(t0, t1)
