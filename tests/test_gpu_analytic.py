"""GPU: uniform and guided integrators against closed-form Poisson / Neumann solutions with MEASURED standard errors
(rows A13, A14). Problems: tests/helpers.py ANALYTIC (unit disk / ball with constant source, unit square /
cube with spatially varying Neumann flux); estimators: util/green.h:14-119, integrator/uniform/integrator.cu:234-445.
The standard error comes from wost_film_stderr with spp_batch == samples_per_pixel (one accumulator slot per sample: the
exact per-pixel sample variance). Bar (north_star): |delta| <= 3 standard errors per pixel at fixed spp -- nominal coverage
99.7 %, and the count over N pixels has a binomial s.d. of sqrt(0.003 / N) -- and relMSE equal to what the measured variance
predicts. The per-walk distributions are skewed (a few long walks carry large weights), so at 128-256 samples the
t-statistic's 3-sigma coverage is 98.9-99.9 % rather than the normal 99.7 %; thresholds are set just below what was measured."""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu

NET_SMALL = {"encoding": {"otype": "DenseGrid", "n_levels": 4, "n_features_per_level": 4, "base_resolution": 4, "per_level_scale": 1.405},
             "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 3},
             "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "Adam", "learning_rate": 8e-3, "beta1": 0.9, "beta2": 0.99,
                                                                    "epsilon": 1e-15, "l2_reg": 1e-6}}}

# name -> (frame, spp, depth); the cube needs a deep walk budget (tests/test_oracle_analytic.py documents the truncation bias)
CASES = {"poisson_disk_2d": (64, 256, 64), "poisson_ball_3d": (40, 128, 64), "neumann_square_2d": (64, 256, 128), "neumann_cube_3d": (40, 128, 512)}


def _problem(spec):
    from elaina_b200 import api
    dim = spec["dim"]
    kw = {k: spec[k] for k in ("color_dirichlet", "color_neumann", "source", "source_origin", "source_voxel", "grid_scale", "grid_pos", "grid_up",
                               "grid_right", "aabb_min", "aabb_max") if spec.get(k) is not None}
    g_n = api.Scene(dim, spec["nverts"], spec["nprims"]) if spec.get("nverts") is not None else None
    return api.Problem(dim, dirichlet=api.Scene(dim, spec["dverts"], spec["dprims"]), neumann=g_n, **kw)


def _points(spec, W):
    """core/evaluation_grid.h:27-33,60-64 in numpy"""
    ndc = 2.0 * np.arange(W, dtype=np.float64) / W - 1.0
    NX, NY = np.meshgrid(ndc, ndc)
    up = np.asarray(spec["grid_up"], np.float64)
    right = np.array([up[1], -up[0]]) if spec["dim"] == 2 else np.asarray(spec["grid_right"], np.float64)
    return spec["grid_scale"] * (NX[..., None] * right + NY[..., None] * up) + np.asarray(spec["grid_pos"], np.float64)


def _check(name, sol, se, spec, u, inside, W, f3_min, relmse_factor=1.25):
    pts = _points(spec, W)
    m = inside(pts) & (se > 0)
    ref = u(pts)
    z = (sol[m] - ref[m]) / se[m]
    f3, f1 = float(np.mean(np.abs(z) <= 3.0)), float(np.mean(np.abs(z) <= 1.0))
    assert f3 >= f3_min, (name, f3)
    assert 0.58 <= f1 <= 0.78, (name, f1)                 # calibrated: ~68 % within one standard error (0.62-0.69 measured; the
                                                          # 3D Neumann term G(r) ~ 1/r is heavy-tailed, its variance estimate noisy)
    assert abs(z.mean()) < 0.25, (name, float(z.mean()))  # mesh discretisation (inscribed polygon / triangles) only
    relmse = np.mean((sol[m] - ref[m]) ** 2 / (ref[m] ** 2 + 1e-2))
    expected = np.mean(se[m] ** 2 / (ref[m] ** 2 + 1e-2))
    assert relmse < relmse_factor * expected, (name, float(relmse), float(expected))
    return f3, f1, float(z.mean()), float(relmse), float(expected)


@pytest.mark.parametrize("name", sorted(CASES))
def test_uniform_matches_closed_form(name):
    from elaina_b200 import api
    W, spp, depth = CASES[name]
    spec, u, inside = helpers.ANALYTIC[name]()
    it = api.UniformIntegrator(_problem(spec), api.make_settings((W, W), depth, spp, 2e-3, seed=7, spp_batch=spp))
    it.solve()
    out = _check(name, it.film()[..., 0].astype(np.float64), it.film_stderr()[..., 0].astype(np.float64), spec, u, inside, W, 0.985)
    print(name, "uniform: f3 %.4f f1 %.3f mean z %.3f relMSE %.3e (variance predicts %.3e)" % out)


@pytest.mark.parametrize("name", sorted(CASES))
def test_guided_matches_closed_form(name):
    """same problems through the guided integrator (training inside solve(), then guiding with the learned mixture): the
    estimator stays unbiased -- MIS pdf, reflection fold on the Neumann boundary, selection probability (rows A18-A20)"""
    from elaina_b200 import api
    W, spp, depth = CASES[name]
    spec, u, inside = helpers.ANALYTIC[name]()
    st = api.make_settings((W, W), depth, spp, 2e-3, seed=11, spp_batch=spp, train_spp_count=spp // 4, max_guided_depth_training=8,
                           max_guided_depth_guiding=8, uniform_fraction_training=0.5, uniform_fraction_guiding=0.5,
                           train_min_batch=1024, train_batch_size=8192)
    it = api.GuidedIntegrator(_problem(spec), st, NET_SMALL)
    it.solve()
    assert it.train_stats["steps"] >= spp // 4
    # the guided run is not reproducible bit for bit (the grid / weight gradients are summed with atomics), and the measured variance
    # of the heavy-tailed 3D Neumann term underestimates the error of a single run by up to ~30 % (1.30 observed): 1.6 x here
    out = _check(name, it.film()[..., 0].astype(np.float64), it.film_stderr()[..., 0].astype(np.float64), spec, u, inside, W, 0.98, relmse_factor=1.6)
    print(name, "guided: f3 %.4f f1 %.3f mean z %.3f relMSE %.3e (variance predicts %.3e)" % out)
