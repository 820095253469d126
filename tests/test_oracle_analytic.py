"""The oracle's Poisson and Neumann estimators against closed-form solutions (rows A13, A14).

The walker oracle is pinned walk-for-walk against the reference's CUDA build (test_oracle_b1_pin.py); this file checks that
what both compute is the PDE solution: source term through the harmonic Green's function of the ball (util/green.h:14-119,
integrator/uniform/integrator.cu:234-316) in 2D and 3D, and the Neumann-flux estimator with reflection (:318-445, :447-526)
with non-zero, spatially varying flux in 2D and 3D. Standard errors are MEASURED: the oracle is run once per sample (one
sample per pixel, seed = sample index), so the per-pixel sample standard deviation is exact (255 degrees of freedom)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from oracle_lib import Oracle  # noqa: E402

# name -> (frame, spp, depth): the 3D cube needs a deep walk budget -- with maxWalkingDepth 64 about 9 % of its walks are cut
# before they reach the Dirichlet faces and the estimate is biased low by 0.04 (the reference truncates the same way)
CASES = {"poisson_disk_2d": (24, 256, 64), "poisson_ball_3d": (10, 256, 64), "neumann_square_2d": (24, 256, 128), "neumann_cube_3d": (12, 256, 512)}


def oracle_samples(orc, spec, W, spp, depth, eps=2e-3):
    dim = spec["dim"]
    kw = {k: spec[k] for k in ("color_dirichlet", "color_neumann", "source", "source_origin", "source_voxel", "grid_scale", "grid_pos", "grid_up",
                               "grid_right", "aabb_min", "aabb_max") if spec.get(k) is not None}
    prob = dict(dim=dim, dirichlet=orc.scene(dim, spec["dverts"], spec["dprims"]),
                neumann=orc.scene(dim, spec["nverts"], spec["nprims"]) if spec.get("nverts") is not None else None, **kw)
    vals = []
    for s in range(spp):
        st = dict(frame_w=W, frame_h=W, max_walking_depth=depth, samples_per_pixel=1, epsilon_shell=eps, seed=1000 + s)
        sol, _ = orc.solve_uniform(prob, st)
        vals.append(sol[:, 0].astype(np.float64))
    pts = orc.eval_points(prob, st, np.arange(W * W)).astype(np.float64)
    return np.asarray(vals), pts


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_closed_form(name):
    W, spp, depth = CASES[name]
    spec, u, inside = helpers.ANALYTIC[name]()
    vals, pts = oracle_samples(Oracle(), spec, W, spp, depth)
    mean, se = vals.mean(0), vals.std(0, ddof=1) / np.sqrt(spp)
    m = inside(pts)
    ref = u(pts)
    z = (mean[m] - ref[m]) / se[m]
    # |delta| <= 3 measured standard errors (nominal coverage 99.7 %; the count over a few hundred pixels is itself noisy)
    # measured on these seeds: 0.992 / 1.000 / 1.000 / 0.986 (2 of 144 pixels)
    assert np.mean(np.abs(z) <= 3.0) >= 0.98, (name, float(np.mean(np.abs(z) <= 3.0)))
    assert np.abs(z).max() < 4.5, (name, float(np.abs(z).max()))
    # the standard errors are calibrated, not just large: about 68 % of the pixels lie within one of them
    assert 0.58 <= np.mean(np.abs(z) <= 1.0) <= 0.80, (name, float(np.mean(np.abs(z) <= 1.0)))
    # no systematic bias beyond the mesh discretisation (inscribed polygons / triangles make the domain slightly smaller)
    assert abs(z.mean()) < 0.5, (name, float(z.mean()))
    relmse = np.mean((mean[m] - ref[m]) ** 2 / (ref[m] ** 2 + 1e-2))
    expected = np.mean(se[m] ** 2 / (ref[m] ** 2 + 1e-2))
    assert relmse < 1.5 * expected, (name, float(relmse), float(expected))


def test_walk_depth_truncation_is_the_only_bias_of_the_3d_neumann_problem():
    """documents the effect the depth budget has (and that the estimator itself is unbiased): same problem, depth 64 vs 512"""
    spec, u, inside = helpers.ANALYTIC["neumann_cube_3d"]()
    orc = Oracle()
    v64, pts = oracle_samples(orc, spec, 8, 256, 64)
    v512, _ = oracle_samples(orc, spec, 8, 256, 512)
    ref = u(pts)
    assert (v64.mean(0) - ref).mean() < -0.04          # truncated walks contribute nothing: biased low (measured -0.067)
    assert abs((v512.mean(0) - ref).mean()) < 0.03     # measured -0.018 with a standard error of 0.009 on this frame mean
