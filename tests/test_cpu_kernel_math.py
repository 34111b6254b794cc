"""CPU checks of the algebra the compositing kernels rely on (numpy restatements of the device code,
no GPU): the places where vegs_b200 does NOT follow the reference's operation order step by step.

* backward: the reference's five "accumulated behind the hit" recurrences (colour x3, depth, alpha;
  cuda_rasterizer/backward.cu:586-623) collapse, under the pixel's cotangent, into one scalar
  recurrence, and a working set can be cut into segments that are reduced to affine maps and
  composed (composite_bwd.cu, S2 pass 1 / pass 2);
* forward: the real part of glm::mix(identity, q, w) is cos(w theta) when q.w = cos(theta), and the
  sin / cos pair on [0, pi] used for it (composite_fwd.cu, sincos_0_pi);
* forward: the per-row interval that narrows the exact contribution test is conservative
  (composite_common.cuh, entry_hits_fwd).
"""
import numpy as np
import pytest


# ------------------------------------------------------------------ backward recurrences
def _reference_dl_dalpha(alpha, col, depth, dpix, dpd, dpa, T_final, bg):
    """dL/dalpha_j for j = n-1 .. 0 exactly as the reference's loop states it (float64)."""
    n = len(alpha)
    T = T_final
    accum_rec = np.zeros(3); accum_red = 0.0; accum_rea = 0.0
    last_alpha = 0.0; last_c = np.zeros(3); last_d = 0.0
    bg_dot = float(bg @ dpix)
    out = np.zeros(n); w = np.zeros(n)
    for j in range(n - 1, -1, -1):
        T = T / (1.0 - alpha[j])
        w[j] = alpha[j] * T
        accum_rec = last_alpha * last_c + (1.0 - last_alpha) * accum_rec
        last_c = col[j]
        dL = float((col[j] - accum_rec) @ dpix)
        accum_red = last_alpha * last_d + (1.0 - last_alpha) * accum_red
        last_d = depth[j]
        dL += (depth[j] - accum_red) * dpd
        accum_rea = last_alpha + (1.0 - last_alpha) * accum_rea
        dL += (1.0 - accum_rea) * dpa
        dL *= T
        last_alpha = alpha[j]
        dL += (-T_final / (1.0 - alpha[j])) * bg_dot
        out[j] = dL
    return w, out


def _collapsed_segmented(alpha, col, depth, dpix, dpd, dpa, T_final, bg, cuts):
    """The kernel's formulation: v = <c, depth, 1>.<dL/dC, dL/dD, dL/dalpha_map>, one recurrence
    A' = alpha v + (1 - alpha) A, processed back to front in segments that are first reduced to
    (prod 1/(1-alpha), a, b) and then re-walked from the composed incoming state."""
    n = len(alpha)
    v = col @ dpix + depth * dpd + dpa
    bg_term = -T_final * float(bg @ dpix)
    order = list(range(n - 1, -1, -1))                      # back to front
    segs = [order[a:b] for a, b in zip([0] + cuts, cuts + [n]) if b > a]
    agg = []
    for seg in segs:                                        # pass 1
        P, a, b = 1.0, 0.0, 1.0
        for j in seg:
            om = 1.0 - alpha[j]
            P *= 1.0 / om
            a = alpha[j] * v[j] + om * a
            b *= om
        agg.append((P, a, b))
    w = np.zeros(n); out = np.zeros(n)
    T_in, A_in = T_final, 0.0
    for seg, (P, a, b) in zip(segs, agg):                   # pass 2
        T, A = T_in, A_in
        for j in seg:
            r = 1.0 / (1.0 - alpha[j])
            T *= r
            w[j] = alpha[j] * T
            out[j] = T * (v[j] - A) + bg_term * r
            A = alpha[j] * v[j] + (1.0 - alpha[j]) * A
        T_in, A_in = T_in * P, a + b * A_in
        assert np.isclose(T, T_in, rtol=1e-12) and np.isclose(A, A_in, rtol=1e-10, atol=1e-12)
    return w, out


@pytest.mark.parametrize("n,seed", [(1, 0), (7, 1), (64, 2), (300, 3)])
def test_collapsed_and_segmented_backward_recurrence_equals_the_reference_loop(n, seed):
    rng = np.random.default_rng(seed)
    alpha = rng.uniform(1 / 255, 0.99, n)
    col, depth = rng.uniform(0, 1, (n, 3)), rng.uniform(0.2, 200, n)
    dpix, dpd, dpa = rng.normal(size=3), rng.normal(), rng.normal()
    bg = rng.uniform(0, 1, 3)
    T_final = float(np.prod(1.0 - alpha))
    w_ref, g_ref = _reference_dl_dalpha(alpha, col, depth, dpix, dpd, dpa, T_final, bg)
    cuts = sorted(rng.choice(np.arange(1, n), size=min(5, max(n - 1, 0)), replace=False).tolist()) if n > 1 else []
    w, g = _collapsed_segmented(alpha, col, depth, dpix, dpd, dpa, T_final, bg, cuts)
    assert np.allclose(w, w_ref, rtol=1e-10)
    assert np.allclose(g, g_ref, rtol=1e-8, atol=1e-10 * np.abs(g_ref).max())


# ------------------------------------------------------------------ slerp from the identity
def test_real_part_of_mix_from_identity_is_cos_w_theta():
    rng = np.random.default_rng(0)
    qw = rng.uniform(-0.999, 0.999, 10000)
    w = rng.uniform(0, 1, 10000)
    theta = np.arccos(qw)
    mix_real = (np.sin((1 - w) * theta) + qw * np.sin(w * theta)) / np.sin(theta)   # glm::mix, real part
    assert np.abs(mix_real - np.cos(w * theta)).max() < 1e-12


def _sincos_0_pi(x):
    """float32 restatement of sincos_0_pi (composite_fwd.cu)."""
    f = np.float32
    x = x.astype(f)
    fma = lambda a, b, c: (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f)
    k = lambda v: np.full_like(x, v, dtype=f)
    j = np.rint(x * f(0.636619772)).astype(f)
    r = fma(j, k(-1.5703125), x)
    r = fma(j, k(-4.837512969970703125e-4), r)
    r = fma(j, k(-7.54978995489188e-8), r)
    r2 = (r * r).astype(f)
    sp = fma(fma(fma(k(-1.9515295891e-4), r2, k(8.3321608736e-3)), r2, k(-1.6666654611e-1)), (r2 * r).astype(f), r)
    cp = fma(fma(fma(k(2.443315711809948e-5), r2, k(-1.388731625493765e-3)), r2, k(4.166664568298827e-2)),
             (r2 * r2).astype(f), fma(k(-0.5), r2, k(1.0)))
    q = j.astype(int)
    return np.where(q == 1, cp, np.where(q == 2, -sp, sp)), np.where(q == 1, -sp, np.where(q == 2, -cp, cp))


def test_sincos_on_zero_to_pi_is_accurate_to_an_ulp_or_so():
    x = np.linspace(0.0, np.pi, 400001)
    s, c = _sincos_0_pi(x)
    x32 = x.astype(np.float32).astype(np.float64)
    assert np.abs(s - np.sin(x32)).max() < 2e-7
    assert np.abs(c - np.cos(x32)).max() < 2e-7


# ------------------------------------------------------------------ row interval of the hit test
def test_row_interval_never_excludes_a_contributing_pixel():
    """For random conics / opacities / positions: every pixel of a row that passes the reference's
    test (power <= 0 and min(0.99, opacity * exp(power)) >= 1/255) lies inside the interval
    entry_hits_fwd computes for that row (float32 restatement)."""
    rng = np.random.default_rng(5)
    f = np.float32
    n = 20000
    # covariance = R diag(s1, s2) R^T + 0.3 I, conic = inverse (what the preprocess stage produces)
    ang = rng.uniform(0, np.pi, n)
    s1, s2 = 10 ** rng.uniform(-1.5, 2.5, n), 10 ** rng.uniform(-1.5, 2.5, n)
    ca, sa = np.cos(ang), np.sin(ang)
    cxx = ca * ca * s1 + sa * sa * s2 + 0.3
    cyy = sa * sa * s1 + ca * ca * s2 + 0.3
    cxy = ca * sa * (s1 - s2)
    det = cxx * cyy - cxy * cxy
    A, B, C = (cyy / det).astype(f), (-cxy / det).astype(f), (cxx / det).astype(f)
    op = rng.uniform(0.004, 1.0, n).astype(f)
    mx, my = rng.uniform(-4, 12, n).astype(f), rng.uniform(-4, 8, n).astype(f)
    missed = 0
    checked = 0
    for y in range(4):
        dy = (my - f(y)).astype(f)
        tau = (np.log(f(255.0) * op) + f(0.01)).astype(f)
        bb = (B * B).astype(f)
        dt = (A.astype(np.float64) * C - bb).astype(f) - (B.astype(np.float64) * B - bb).astype(f)
        disc = (-dt.astype(np.float64) * (dy * dy).astype(f) + (f(2.0) * tau * A).astype(f)).astype(f)
        root = (np.sqrt(np.maximum(disc, 0)).astype(f) * f(1.0005)).astype(f)
        mid = (-B * dy).astype(f)
        inv_a = (f(1.0) / A).astype(f)
        x_lo = (mx - ((mid + root) * inv_a).astype(f) - f(0.02)).astype(f)
        x_hi = (mx - ((mid - root) * inv_a).astype(f) + f(0.02)).astype(f)
        for x in range(8):
            dx = (mx - f(x)).astype(f)
            power = (-0.5 * (A.astype(np.float64) * dx * dx + C.astype(np.float64) * dy * dy)
                     - B.astype(np.float64) * dx * dy)
            alpha = np.minimum(0.99, op.astype(np.float64) * np.exp(power))
            hit = (power <= 0) & (alpha >= 1.0 / 255.0) & (tau >= 0)
            inside = (disc >= 0) & (x >= np.ceil(x_lo)) & (x <= np.floor(x_hi))
            missed += int((hit & ~inside).sum())
            checked += int(hit.sum())
    assert checked > 10000
    assert missed == 0


def test_box_extents_contain_every_contributing_pixel():
    """The per-Gaussian half extents (ex, ey) the preprocess stage stores as fp16 (preprocess.cu) bound
    the pixels that can pass the reference's test, also for strongly anisotropic splats and after
    the rounding to half precision."""
    rng = np.random.default_rng(9)
    f = np.float32
    n = 200000
    ang = rng.uniform(0, np.pi, n)
    s1, s2 = 10 ** rng.uniform(-3, 4, n), 10 ** rng.uniform(-3, 1, n)
    ca, sa = np.cos(ang), np.sin(ang)
    cxx = (ca * ca * s1 + sa * sa * s2 + 0.3).astype(f)
    cyy = (sa * sa * s1 + ca * ca * s2 + 0.3).astype(f)
    cxy = (ca * sa * (s1 - s2)).astype(f)
    det = (cxx.astype(np.float64) * cyy - (cxy * cxy).astype(f)).astype(f)
    keep = det > 0
    det_inv = (f(1) / det).astype(f)
    A, B, C = (cyy * det_inv).astype(f), (-cxy * det_inv).astype(f), (cxx * det_inv).astype(f)
    op = rng.uniform(0.004, 1.0, n).astype(f)
    tau = (np.log(f(255.0) * op) + f(0.01)).astype(f)
    cdet = (A * C - B * B).astype(f)
    ok = keep & (cdet > 0) & (tau >= 0)
    with np.errstate(invalid="ignore", divide="ignore"):
        s2t = (f(2.0) * tau / cdet).astype(f)
        ex = (np.sqrt(s2t * C).astype(f) * f(1.002) + f(0.01)).astype(np.float16).astype(f)
        ey = (np.sqrt(s2t * A).astype(f) * f(1.002) + f(0.01)).astype(np.float16).astype(f)
    # sample offsets around the splat and apply the reference's float32 test
    missed = checked = 0
    for _ in range(6):
        scale = np.maximum(ex, ey) * 1.5 + 1.0
        dx = (rng.uniform(-1, 1, n) * scale).astype(f)
        dy = (rng.uniform(-1, 1, n) * scale).astype(f)
        p1 = ((A * dx).astype(f) * dx).astype(f)
        p2 = ((C * dy).astype(f) * dy).astype(f)
        power = ((p1.astype(np.float64) + p2).astype(f).astype(np.float64) * -0.5
                 - ((B * dx).astype(f) * dy).astype(f)).astype(f)
        alpha = np.minimum(f(0.99), (op * np.exp(power.astype(np.float64)).astype(f)).astype(f))
        hit = ok & (power <= 0) & (alpha >= f(1.0 / 255.0))
        inside = (np.abs(dx) <= ex) & (np.abs(dy) <= ey)
        missed += int((hit & ~inside & np.isfinite(ex) & np.isfinite(ey)).sum())
        checked += int(hit.sum())
    assert checked > 50000
    assert missed == 0


def test_unit_masks_from_pixel_boxes_equal_the_per_entry_box_test():
    """binning.cu derives, per tile instance, the mask of compositing units the Gaussian reaches from
    the integer pixel box the preprocess stage stores; the scan of P >= 2^24 scenes instead clips
    the float box per entry (unit_box).  Both must keep exactly the same entries, including for
    off-screen, huge, infinite and NaN boxes."""
    rng = np.random.default_rng(3)
    f = np.float32
    W, H = 1408, 376
    n = 200000
    x = rng.uniform(-300, W + 300, n).astype(f)
    y = rng.uniform(-300, H + 300, n).astype(f)
    ex = (10 ** rng.uniform(-2, 3, n)).astype(np.float16).astype(f)
    ey = (10 ** rng.uniform(-2, 3, n)).astype(np.float16).astype(f)
    special = rng.integers(0, 40, n)
    ex[special == 0] = np.inf; ey[special == 1] = np.inf
    ex[special == 2] = -1.0; ey[special == 2] = -1.0
    x[special == 3] = np.nan; y[special == 4] = np.nan
    x[special == 5] = 3e9; y[special == 6] = -3e9

    def to_int(v):                       # cvt.rzi.s32.f32 semantics: saturate, NaN -> 0
        v = np.where(np.isnan(v), 0.0, v.astype(np.float64))
        return np.clip(np.trunc(v), -2 ** 31, 2 ** 31 - 1).astype(np.int64)

    def fmax(a, b):                      # fmaxf / fminf return the non-NaN operand
        return np.where(np.isnan(a), b, np.where(np.isnan(b), a, np.maximum(a, b)))

    def fmin(a, b):
        return np.where(np.isnan(a), b, np.where(np.isnan(b), a, np.minimum(a, b)))

    with np.errstate(invalid="ignore", over="ignore"):
        lo_x, hi_x, lo_y, hi_y = (x - ex).astype(f), (x + ex).astype(f), (y - ey).astype(f), (y + ey).astype(f)
        # preprocess.cu: the box clipped to the image, in whole pixels
        gx0 = np.minimum(to_int(np.ceil(fmax(lo_x, f(0)))), W)
        gx1 = np.maximum(to_int(np.floor(fmin(hi_x, f(W - 1)))), -1)
        gy0 = np.minimum(to_int(np.ceil(fmax(lo_y, f(0)))), H)
        gy1 = np.maximum(to_int(np.floor(fmin(hi_y, f(H - 1)))), -1)
        tiles = rng.integers(0, [(W + 15) // 16, (H + 15) // 16], (n, 2))
        for blk in range(8):
            bx0 = tiles[:, 0] * 16 + (blk & 1) * 8
            by0 = tiles[:, 1] * 16 + (blk >> 1) * 4
            bx1 = np.minimum(bx0 + 7, W - 1)
            by1 = np.minimum(by0 + 3, H - 1)
            on_image = (bx0 < W) & (by0 < H)
            # composite_common.cuh unit_box
            ux0 = to_int(np.ceil(fmax(lo_x, bx0.astype(f))))
            ux1 = to_int(np.floor(fmin(hi_x, bx1.astype(f))))
            uy0 = to_int(np.ceil(fmax(lo_y, by0.astype(f))))
            uy1 = to_int(np.floor(fmin(hi_y, by1.astype(f))))
            keep_box = on_image & (ux0 <= ux1) & (uy0 <= uy1)
            # binning.cu mask bit
            keep_mask = on_image & (np.maximum(gx0, bx0) <= np.minimum(gx1, bx1)) \
                & (np.maximum(gy0, by0) <= np.minimum(gy1, by1))
            assert np.array_equal(keep_box, keep_mask), f"unit {blk}: {int((keep_box != keep_mask).sum())} differ"
    assert keep_box.sum() > 100


def test_margins_scale_with_the_cancelling_terms_for_long_thin_splats():
    """Long thin splats sampled far from their centre along the major axis: the terms
    A dx^2, C dy^2, 2 B dx dy reach 1e5 - 1e7 and cancel to a power of a few units, so the float32
    evaluation of the reference is off by far more than a fixed 0.01.  The box extents of
    preprocess.cu widen tau by 4e-6 of those terms; with that every pixel that passes the
    reference's float32 test lies inside the box, while the fixed margin alone misses some."""
    rng = np.random.default_rng(21)
    f = np.float32
    n = 400000
    ang = rng.uniform(0.2, np.pi / 2 - 0.2, n)                  # well away from the axes: all three terms large
    s1, s2 = 10 ** rng.uniform(4.5, 6.5, n), 10 ** rng.uniform(-2, 0, n)   # sigma 180 - 1800 px by ~0.1 - 1 px
    ca, sa = np.cos(ang), np.sin(ang)
    cxx = (ca * ca * s1 + sa * sa * s2 + 0.3).astype(f)
    cyy = (sa * sa * s1 + ca * ca * s2 + 0.3).astype(f)
    cxy = (ca * sa * (s1 - s2)).astype(f)
    det = (cxx.astype(np.float64) * cyy - (cxy * cxy).astype(f).astype(np.float64)).astype(f)
    det_inv = (f(1) / det).astype(f)
    A, B, C = (cyy * det_inv).astype(f), (-cxy * det_inv).astype(f), (cxx * det_inv).astype(f)
    op = rng.uniform(0.3, 1.0, n).astype(f)
    tau = (np.log(f(255.0) * op) + f(0.01)).astype(f)
    cdet = (A * C - B * B).astype(f)
    ok = (det > 0) & (cdet > 0) & np.isfinite(A) & np.isfinite(C)

    def extents(t):
        with np.errstate(invalid="ignore", divide="ignore"):
            s2t = (f(2.0) * t / cdet).astype(f)
            return np.sqrt(s2t * C).astype(f), np.sqrt(s2t * A).astype(f)

    tx, ty = extents(tau)
    mag = (A * tx * tx + C * ty * ty + f(2.0) * np.abs(B) * tx * ty).astype(f)
    tx2, ty2 = extents((tau + f(4e-6) * mag).astype(f))
    ex_new = (tx2 * f(1.002) + f(0.01)).astype(np.float16).astype(f)
    ey_new = (ty2 * f(1.002) + f(0.01)).astype(np.float16).astype(f)
    ex_old = (tx * f(1.002) + f(0.01)).astype(np.float16).astype(f)
    ey_old = (ty * f(1.002) + f(0.01)).astype(np.float16).astype(f)
    ok &= np.isfinite(ex_new) & np.isfinite(ey_new)               # inf = "test the whole tile", always safe
    missed_new = missed_old = checked = 0
    for _ in range(8):
        # points on the major axis near the end of the box, jittered across it by a pixel or so
        r = (rng.uniform(0.9, 1.02, n) * np.hypot(tx, ty)).astype(np.float64) * rng.choice([-1.0, 1.0], n)
        jit = rng.uniform(-1.5, 1.5, n)
        dx = np.rint(ca * r - sa * jit).astype(f)               # whole-pixel offsets, as pixel centres are
        dy = np.rint(sa * r + ca * jit).astype(f)
        p1 = ((A * dx).astype(f) * dx).astype(f)
        p2 = ((C * dy).astype(f) * dy).astype(f)
        power = ((p1.astype(np.float64) + p2).astype(f).astype(np.float64) * -0.5
                 - ((B * dx).astype(f) * dy).astype(f)).astype(f)
        alpha = np.minimum(f(0.99), (op * np.exp(np.minimum(power, 0).astype(np.float64)).astype(f)).astype(f))
        hit = ok & (power <= 0) & (alpha >= f(1.0 / 255.0))
        missed_new += int((hit & ~((np.abs(dx) <= ex_new) & (np.abs(dy) <= ey_new))).sum())
        missed_old += int((hit & ~((np.abs(dx) <= ex_old) & (np.abs(dy) <= ey_old))).sum())
        checked += int(hit.sum())
    assert checked > 1000
    assert missed_new == 0, missed_new
    print("fixed-margin box would have missed", missed_old, "of", checked)
