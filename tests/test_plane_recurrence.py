"""The algebra behind the FP32 stepper (csrc/rrt_trace.cuh fast_init / fast_step), checked on the CPU
against the reference's own chords (tests/golden/kat_geodesic.npz, produced by BlackHole::next_micro_ray): the recurrence
the reference evaluates from positions every step is, in the orbital plane, a scalar recurrence in (u, u') plus a frame
rotation with a possible flip — including after a ray has escaped (u+ < 0) and bounces between two far points.  Evaluated
in double here, the plane form must follow the reference's chords to rounding; in float32 (what the device carries), to
well inside BASELINE.json's 1e-4.  CPU only; the device kernel itself is tested in test_gpu_parity.py."""
import numpy as np
import pytest

from util import golden


def plane_chords(o, d, c, r, h, nsteps, dt):
    """numpy restatement of fast_init + fast_step for n rays; returns [n, nsteps, 8] like rrt_geodesic"""
    o = np.asarray(o, np.float64); d = np.asarray(d, np.float64); c = np.asarray(c, np.float64)
    with np.errstate(all="ignore"):
        x = o - c
        u = 1.0 / np.sqrt((x * x).sum(1)); x = x * u[:, None]
        dx = (d * x).sum(1); y = d - dx[:, None] * x
        rdy = 1.0 / np.sqrt((y * y).sum(1)); y = y * rdy[:, None]
        up = -u * dx * rdy
        X = x.astype(dt); Y = y.astype(dt); u = u.astype(dt); up = up.astype(dt); cc = c.astype(dt)
        k = dt(2 * np.sin(h / 2) ** 2); sh = dt(np.sin(h)); ch = dt(np.cos(h)); cot = dt(np.cos(h) / np.sin(h)); tan2 = dt(np.tan(h / 2))
        r15 = dt(1.5 * r); hh = dt(h); h2 = dt(h / 2); h24 = dt(h * h / 4); h26 = dt(h * h / 6); r2 = dt(r * r)
        f = lambda v: r15 * v * v - v
        out = np.zeros((len(o), nsteps, 8))
        for j in range(nsteps):
            rho = dt(1) / u
            org = cc + rho[:, None] * X
            f1 = f(u); v2 = u + up * h2; f2 = f(v2); v3 = v2 + f1 * h24; f3 = f(v3)
            G = up * hh + (f1 + f2 + f3) * h26
            un = u + G
            rhon = dt(1) / un
            A = -(u * k + G) * rho * rhon; B = sh * rhon
            q = A[:, None] * X + B[:, None] * Y
            ln = np.sqrt((q * q).sum(1)); dirn = q / ln[:, None]
            perp = rho * B / ln; disc = r2 - perp * perp
            sq = np.sqrt(np.maximum(disc, 0)); tc = -rho * A / ln
            cap = (disc >= 0) & (((0 <= tc - sq) & (tc - sq <= ln)) | ((0 <= tc + sq) & (tc + sq <= ln)))
            out[:, j, 0:3] = org; out[:, j, 3:6] = dirn; out[:, j, 6] = ln; out[:, j, 7] = cap
            sg = np.where(un < 0, dt(-1), dt(1))
            E = ch * X + sh * Y
            Y = ch * Y - sh * X
            X = sg[:, None] * E
            up = G * cot - u * tan2
            u = np.abs(un)
    return out


def errors(ref, got, c):
    end_r = ref[..., 0:3] + ref[..., 3:6] * ref[..., 6:7]; end_g = got[..., 0:3] + got[..., 3:6] * got[..., 6:7]
    with np.errstate(all="ignore"):
        perr = np.linalg.norm(end_r - end_g, axis=-1) / np.linalg.norm(end_r - c, axis=-1)
        derr = np.linalg.norm(ref[..., 3:6] - got[..., 3:6], axis=-1)
    captured_before = np.cumsum(ref[..., 7] > 0, axis=1) - (ref[..., 7] > 0) > 0   # chords after a capture are never used
    ok = np.isfinite(perr) & np.isfinite(derr) & ~captured_before
    return perr, derr, ok


@pytest.mark.parametrize("tag", ["default", "alt"])
def test_plane_form_is_the_reference_recurrence(tag):
    g = golden("kat_geodesic")
    bh = g["bh_" + tag]; ref = g["seg_" + tag]
    got = plane_chords(g["o"], g["d"], bh[:3], bh[3], bh[4], ref.shape[1], np.float64)
    perr, derr, ok = errors(ref, got, bh[:3])
    assert ok.sum() > 0.5 * ok.size
    # same recurrence, different operation order: agreement to double rounding amplified by the recurrence (escaped rays
    # jump to radii of 1e4..1e7 when u+ passes close to 0; errors are relative to the radius)
    assert np.median(perr[ok]) < 1e-12 and np.percentile(perr[ok], 99) < 1e-10 and perr[ok].max() < 1e-8
    assert derr[ok].max() < 1e-11
    assert np.array_equal((ref[..., 7] > 0)[ok], (got[..., 7] > 0)[ok])          # horizon verdicts
    # the set does exercise the escaped regime (a chord whose end lies on the far side of the hole from its start)
    v0 = ref[..., 0:3] - bh[:3]; v1 = ref[..., 0:3] + ref[..., 3:6] * ref[..., 6:7] - bh[:3]
    assert ((v0 * v1).sum(-1)[ok] < 0).sum() > 100


@pytest.mark.parametrize("tag", ["default", "alt"])
def test_plane_form_in_float32_is_within_tolerance(tag):
    g = golden("kat_geodesic")
    bh = g["bh_" + tag]; ref = g["seg_" + tag]
    got = plane_chords(g["o"], g["d"], bh[:3], bh[3], bh[4], ref.shape[1], np.float32)
    perr, derr, ok = errors(ref, got, bh[:3])
    # BASELINE.json: end points and directions within 1e-4 relative.  Directions: always.  End points: 99 % of the chords;
    # the rest are far points of escaped rays, r = 1 / |u+| with u+ = u + G a difference of nearly equal numbers: the
    # radius loses the digits the subtraction cancels, the chord's direction and its near end do not
    assert np.median(perr[ok]) < 2e-6 and np.percentile(perr[ok], 99) < 1e-4 and perr[ok].max() < 1e-3
    assert np.median(derr[ok]) < 1e-6 and derr[ok].max() < 1e-5
    assert np.array_equal((ref[..., 7] > 0)[ok], (got[..., 7] > 0)[ok])
