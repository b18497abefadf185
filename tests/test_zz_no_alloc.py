"""-m gpu: rt_cast_rays_static, the entry point that allocates nothing (no counterpart in the reference; added
for callers that cannot call cudaMalloc). Its device scratch is a static array inside
librtb200.so whose address the library looks up with cudaGetSymbolAddress; the results must equal rt_cast_rays' bit for
bit (the bar tests/test_gpu_parity.py::test_random_rays_vs_oracle sets for ids and t), on calls in a row that share the
array. Collected last on purpose."""
import numpy as np
import pytest

from conftest import scene_path

pytestmark = pytest.mark.gpu


def rays(seed, n):
    rng = np.random.default_rng(seed)
    org = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    org[:, 1] = rng.uniform(0, 14, n)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d[: n // 4] = np.array([0.0, 4.0, 0.0], np.float32) - org[: n // 4]  # a quarter aimed at the terrain
    return org, d


def same(a, b):
    return all((x.view(np.uint32) == y.view(np.uint32)).all() for x, y in zip(a, b))


def test_cast_rays_static_equals_cast_rays(pkg, gpu_ok):
    w = pkg.World(scene_path("world8"))
    s = pkg.Scene(w)
    # builds nothing: without a BVH the call is refused, not served by an allocation
    o0, d0 = rays(1, 64)
    with pytest.raises(RuntimeError, match="status 5"):
        s.cast_rays_static(o0, d0, use_bvh=1)
    s.build_bvh()
    # two calls in a row on the shared array; the second is smaller, so stale results of the first lie behind it
    o1, d1 = rays(2, 50000)
    o2, d2 = rays(3, 1237)
    want1 = s.cast_rays(o1, d1, use_bvh=1)
    want2 = s.cast_rays(o2, d2, use_bvh=1)
    got1 = s.cast_rays_static(o1, d1, use_bvh=1)
    got2 = s.cast_rays_static(o2, d2, use_bvh=1)
    assert (want1[0] >= 0).sum() > 5000 and (want2[0] >= 0).sum() > 100
    assert same(got1, want1)
    assert same(got2, want2)
    # again, and through the other tracer modes (brute force needs no BVH, only the instance boxes)
    assert same(s.cast_rays_static(o1, d1, use_bvh=1), want1)
    for mode in (0, 2, 3):
        assert same(s.cast_rays_static(o2, d2, use_bvh=mode), s.cast_rays(o2, d2, use_bvh=mode)), mode
    # the stated maximum: exactly at it works, one beyond is refused with a message naming the limit
    nmax = pkg.RT_CAST_STATIC_MAX_RAYS
    o3, d3 = rays(4, nmax)
    assert same(s.cast_rays_static(o3, d3, use_bvh=1), s.cast_rays(o3, d3, use_bvh=1))
    o4, d4 = rays(5, nmax + 1)
    with pytest.raises(RuntimeError, match="status 6.*RT_CAST_STATIC_MAX_RAYS"):
        s.cast_rays_static(o4, d4, use_bvh=1)
    assert same(s.cast_rays_static(o2, d2, use_bvh=1), want2)  # still usable after the refusal
    assert s.cast_rays_static(o2[:0], d2[:0], use_bvh=1)[0].size == 0
    # a second scene shares the same array
    w2 = pkg.World(scene_path("world4"))
    s2 = pkg.Scene(w2)
    s2.build_bvh()
    assert same(s2.cast_rays_static(o2, d2, use_bvh=1), s2.cast_rays(o2, d2, use_bvh=1))
    assert same(s.cast_rays_static(o2, d2, use_bvh=1), want2)
    s2.close()
    s.close()
