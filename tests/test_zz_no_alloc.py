"""b200reg_point_to_plane_residual_noalloc — the allocation-free variant of b200reg_point_to_plane_residual (static device scratch,
include/b200reg.h; no counterpart in the reference).

The file is named so that pytest collects it LAST: a failure here must not hide other tests.
It needs nothing from the tests before it (own clouds, own matches from the library's kNN) and passes alone in a fresh process.
Bar: the value of the existing entry point, bit for bit (same kernels, same grid) — asserted at the 1e-9 relative tolerance of
tests/test_gpu_round2.py::test_point_to_plane_residual_matches_oracle and, stricter, for equality — on two calls in a row, with a
different input in between so that stale scratch contents would show."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(ctx, seed, n, m, k):
    """A noisy plane patch + clutter as reference (with normals), a displaced sampling of it as reading; matches from b200reg_knn."""
    rng = np.random.default_rng(seed)
    ref = np.zeros((m, 4), np.float32)
    ref[:, 0:2] = rng.uniform(-5, 5, (m, 2))
    ref[:, 2] = 0.2 * np.sin(ref[:, 0]) + rng.normal(0, 0.01, m)
    ref[:, 3] = 1.0
    nrm = np.zeros((m, 3), np.float32)
    nrm[:, 0] = -0.2 * np.cos(ref[:, 0])
    nrm[:, 2] = 1.0
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    q = np.zeros((n, 4), np.float32)
    q[:, 0:2] = rng.uniform(-5.5, 5.5, (n, 2))          # some readings fall outside: no match within max_dist
    q[:, 2] = 0.2 * np.sin(q[:, 0]) + 0.05 + rng.normal(0, 0.02, n)
    q[:, 3] = 1.0
    d_ref, d_q = ctx.upload(ref, nrm), ctx.upload(q)
    ix = ctx.build_index(d_ref)
    ids, d2, _ = ctx.knn(ix, d_q, k=k, max_dist=0.3)
    ix.free()
    w = rng.choice(np.array([0.0, 0.25, 1.0], np.float32), size=d2.shape)
    return d_q, d_ref, ids, d2, w


def test_noalloc_residual_equals_the_allocating_entry_point_twice_in_a_row(ctx):
    from open3d_slam_advanced_rss_2024_public_b200 import capi
    cases = [_case(ctx, 1, 60_000, 90_000, 1), _case(ctx, 2, 9_000, 20_000, 3), _case(ctx, 3, 300, 500, 1)]
    want = [[ctx.point_to_plane_residual(q, r, ids, d2, w, force2D=f) for f in (False, True)] for q, r, ids, d2, w in cases]
    assert all((np.isinf(d2)).any() and (~np.isinf(d2)).any() for _, _, _, d2, _ in cases)      # missing AND present matches
    for rep in range(2):                                   # two calls in a row per input; a smaller input follows a larger one
        for (q, r, ids, d2, w), exp in zip(cases, want):
            for f, (res_e, kept_e) in zip((False, True), exp):
                for again in range(2):
                    res, kept = ctx.point_to_plane_residual_noalloc(q, r, ids, d2, w, force2D=f)
                    assert kept == kept_e == int(((d2 != np.inf) & (w != 0)).sum())
                    assert abs(res - res_e) <= 1e-9 * max(1.0, abs(res_e))
                    assert res == res_e, (rep, again, res, res_e)
    # all weights zero after a non-trivial call: nothing left over in the scratch
    q, r, ids, d2, w = cases[0]
    assert ctx.point_to_plane_residual_noalloc(q, r, ids, d2, np.zeros_like(w)) == (0.0, 0)
    # a reference without normals and an over-sized request are refused with a clear message
    with pytest.raises(capi.InvalidField):
        ctx.point_to_plane_residual_noalloc(q, ctx.upload(np.zeros((10, 4), np.float32)), ids, d2, w)
    big_n = capi.NOALLOC_MAX_PAIRS // 4 + 1
    big = ctx.upload(np.zeros((big_n, 4), np.float32))
    zi = np.zeros((big_n, 4), np.int32)
    zf = np.zeros((big_n, 4), np.float32)
    with pytest.raises(capi.InvalidParameter, match="B200REG_NOALLOC_MAX_PAIRS"):
        ctx.point_to_plane_residual_noalloc(big, r, zi, zf, zf)
    ok_n = capi.NOALLOC_MAX_PAIRS // 4                      # exactly the maximum is accepted
    res, kept = ctx.point_to_plane_residual_noalloc(ctx.upload(np.zeros((ok_n, 4), np.float32)), r, zi[:ok_n], zf[:ok_n], zf[:ok_n])
    assert (res, kept) == (0.0, 0)
    res, kept = ctx.point_to_plane_residual_noalloc(q, r, ids, d2, w)      # and the array is as good as new afterwards
    assert (res, kept) == want[0][0]
