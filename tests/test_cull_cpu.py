"""The warp-footprint culling of the compositors (guava_renderer_b200/csrc/internal.cuh: splat_may_touch) restated
in numpy float32, and the property every bit-exact claim of the compositors rests on: a splat the test rejects
fails the reference's per-pixel alpha test (forward.cu:350-360: power > 0 or alpha < 1/255) at EVERY pixel centre
of the 8x4 block, so skipping it changes no output of the reference's loop.  Host-side arithmetic only."""
import numpy as np

f32 = np.float32


def splat_may_touch(mx, my, A, B, C, lthr, x0, x1, y0, y1):
    """internal.cuh: splat_may_touch, operation for operation in float32 (no FMA contraction here; the margin of the
    test covers that)."""
    with np.errstate(all="ignore"):
        dxlo, dxhi = mx - x1, mx - x0
        dylo, dyhi = my - y1, my - y0
        ex = np.minimum(np.maximum(f32(0), dxlo), dxhi)
        ey = np.minimum(np.maximum(f32(0), dylo), dyhi)
        dy1 = np.minimum(np.maximum((-B * ex) / C, dylo), dyhi)
        dx2 = np.minimum(np.maximum((-B * ey) / A, dxlo), dxhi)
        q1 = f32(0.5) * (A * ex * ex + C * dy1 * dy1) + B * ex * dy1
        q2 = f32(0.5) * (A * dx2 * dx2 + C * ey * ey) + B * dx2 * ey
        qmin = np.minimum(q1, q2)
        dxm = np.maximum(np.abs(dxlo), np.abs(dxhi))
        dym = np.maximum(np.abs(dylo), np.abs(dyhi))
        M = f32(0.5) * (A * dxm * dxm + C * dym * dym) + np.abs(B) * dxm * dym
        thr = lthr + f32(1e-5) * M
        pd = (A > 0) & (C > 0) & (A * C > f32(1.0001) * B * B)
        return ~(pd & (qmin > thr))


def _random_splats(n, rng):
    # 2D covariances as the forward preprocess makes them (forward.cu:215-235): random SPD + 0.3 on the diagonal
    sx = np.exp(rng.normal(0.3, 1.0, n))
    sy = np.exp(rng.normal(0.3, 1.0, n))
    th = rng.uniform(0, np.pi, n)
    c, s = np.cos(th), np.sin(th)
    a = (c * c * sx * sx + s * s * sy * sy + 0.3).astype(f32)
    b = (c * s * (sx * sx - sy * sy)).astype(f32)
    d = (s * s * sx * sx + c * c * sy * sy + 0.3).astype(f32)
    det = a * d - b * b
    A, B, C = d / det, -b / det, a / det          # conic (x, y, z)
    op = np.clip(1.0 / (1.0 + np.exp(-rng.normal(1.0, 2.0, n))), 0.002, 0.999).astype(f32)
    return A.astype(f32), B.astype(f32), C.astype(f32), op


def test_rejected_splats_fail_the_alpha_test_at_every_pixel_of_the_block():
    rng = np.random.default_rng(20260101)
    n = 400_000
    A, B, C, op = _random_splats(n, rng)
    bx = (rng.integers(0, 64, n) * 8).astype(f32)
    by = (rng.integers(0, 128, n) * 4).astype(f32)
    # means spread around the block: inside, near the edges, and up to a few radii away
    reach = 3.0 * np.sqrt(np.maximum(1.0 / np.minimum(A, C), 1.0))
    mx = (bx + 3.5 + rng.normal(0, 1, n) * (4 + reach)).astype(f32)
    my = (by + 1.5 + rng.normal(0, 1, n) * (2 + reach)).astype(f32)
    lthr = (np.log(f32(255) * op) + f32(0.01)).astype(f32)   # SplatRec::lthr

    keep = splat_may_touch(mx, my, A, B, C, lthr, bx, bx + f32(7), by, by + f32(3))
    rejected = ~keep
    assert 0.2 < rejected.mean() < 0.95, rejected.mean()      # the property below is not vacuous

    worst = np.full(n, -np.inf, dtype=np.float64)
    touched = np.zeros(n, dtype=bool)
    for j in range(4):
        for i in range(8):
            dx = mx - (bx + f32(i))
            dy = my - (by + f32(j))
            power = f32(-0.5) * (A * dx * dx + C * dy * dy) - B * dx * dy     # forward.cu:350-352
            alpha = np.minimum(f32(0.99), op * np.exp(power))                 # :356
            ok = ~(power > 0) & ~(alpha < f32(1.0 / 255.0))                   # :353, :357
            touched |= ok
            worst = np.maximum(worst, np.where(power > 0, -np.inf, alpha))
    bad = rejected & touched
    assert not bad.any(), (int(bad.sum()), float(worst[bad].max()))
    # and the test is not hopelessly loose: most splats that reach no pixel of the block are rejected
    assert (rejected & ~touched).sum() > 0.85 * (~touched).sum()


def test_degenerate_conics_and_nans_are_kept():
    z = f32(0)
    one = np.array([1.0], dtype=f32)
    far = np.array([1e6], dtype=f32)
    lthr = np.array([1.0], dtype=f32)
    args = (z, f32(7), z, f32(3))
    # not positive definite, zero, negative, NaN: never rejected, however far away
    for A, B, C in ((one, 2 * one, one), (0 * one, 0 * one, one), (-one, 0 * one, one),
                    (one * np.nan, 0 * one, one), (one, one * np.nan, one)):
        assert splat_may_touch(far, far, A, B, C, lthr, *args).all()
    assert splat_may_touch(far * np.nan, far, one, 0 * one, one, lthr, *args).all()
    # a proper far-away splat is rejected, the same one on top of the block is kept
    assert not splat_may_touch(far, far, one, 0 * one, one, lthr, *args).any()
    assert splat_may_touch(one * 3, one * 2, one, 0 * one, one, lthr, *args).all()
