"""GPU: the guided integrator through the C ABI.

Layers of evidence, bottom up:
  * VMM / vMF arithmetic on the device == the oracle restatement (which is pinned by the reference's known answers);
  * DenseGrid + MLP forward: SIMT kernel == oracle (fp16 operands, wide accumulation), tcgen05 tile kernel == SIMT;
  * one training step (forward, KL gradient, backward on tcgen05, Adam + EMA) == numpy on the oracle's forward;
  * guided walks with FIXED weights: field within 3 standard errors of the oracle's guided walker and of the analytic
    harmonic solution, fused tcgen05 mode == validation mode (SIMT MLP through HBM);
  * training inside solve(): optimiser steps are taken, the loss goes down, the field stays unbiased.
Tolerances: fp16 tensor-core operands -> 2e-3 absolute on O(1) network outputs; fields: |delta| <= 3 s.e. per pixel."""
import numpy as np
import pytest

from helpers import build_both, mixed_problem_2d

pytestmark = pytest.mark.gpu

NET_SMALL = {"encoding": {"otype": "DenseGrid", "n_levels": 4, "n_features_per_level": 4, "base_resolution": 4, "per_level_scale": 1.405},
             "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 3},
             "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "Adam", "learning_rate": 8e-3, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}}
NET_LADYBUG = {"encoding": {"otype": "DenseGrid", "n_levels": 8, "n_features_per_level": 4, "base_resolution": 8, "per_level_scale": 1.4049999713897705},
               "network": NET_SMALL["network"], "optimizer": NET_SMALL["optimizer"]}


def _rand_dirs(rng, n, dim):
    d = rng.randn(n, dim).astype(np.float32)
    return d / np.linalg.norm(d, axis=1, keepdims=True)


@pytest.mark.parametrize("dim", [2, 3])
def test_vmm_device_matches_oracle(oracle, dim):
    from elaina_b200 import api
    rng = np.random.RandomState(dim)
    n, nl = 4096, dim + 2
    raw = (rng.randn(n, 8 * nl + 1) * 1.5).astype(np.float32)
    raw[:, 1:8 * nl:nl] += rng.rand(n, 8).astype(np.float32) * 6  # kappa up to ~e^9
    d, nn = _rand_dirs(rng, n, dim), _rand_dirs(rng, n, dim)
    on = (rng.rand(n) < 0.5).astype(np.uint8)
    pg, gg, sg = api.vmm_eval(dim, raw, d, nn, on, seed=9)
    po, go, so = oracle.vmm_eval(dim, raw, d, nn, on, seed=9)
    # The device evaluates the densities with the hardware exp / log / reciprocal (what the reference's --use_fast_math build runs)
    # and forms the von Mises log-normaliser without the cancellation of the reference's expression, which the oracle keeps:
    # kappa cos - log 2 pi - log I0(kappa) loses ~kappa ulps, i.e. up to 1e-3 relative at kappa ~ e^9. Rows whose lobes all stay
    # below kappa = 50 agree at the 2e-4 level, all rows within the cancellation noise.
    kap = np.exp(np.clip(raw[:, 1:8 * nl:nl], -10, 15)).max(axis=1)
    low = kap < 50.0
    assert low.sum() > 50
    np.testing.assert_allclose(pg[low], po[low], rtol=2e-4, atol=1e-7)
    np.testing.assert_allclose(pg, po, rtol=3e-3, atol=1e-7)
    assert np.median(np.abs(pg - po) / (np.abs(po) + 1e-12)) < 2e-5
    scale = np.abs(go).max(axis=1, keepdims=True) + 1e-12
    assert np.max(np.abs(gg - go) / scale) < 3e-3
    # same Philox draws, same rejection loop (double for the proposal parameter, candidate and acceptance variable) -> the same
    # direction up to float rounding, except where a 1e-7 perturbation of the acceptance test flips a decision
    assert np.mean(np.linalg.norm(sg - so, axis=1) < 1e-3) > 0.998
    np.testing.assert_allclose(np.linalg.norm(sg, axis=1), 1.0, atol=1e-4)


def _guided(problem, frame=(32, 32), depth=24, spp=4, eps=5e-3, seed=1, network=NET_SMALL, **kw):
    from elaina_b200 import api
    return api.GuidedIntegrator(problem, api.make_settings(frame, depth, spp, eps, seed=seed, **kw), network)


def _oracle_net(oracle, it, cfg, aabb_min, aabb_max, params):
    e = cfg["encoding"]
    net = oracle.net(it.dim, e["n_levels"], e["base_resolution"], e["per_level_scale"], aabb_min, aabb_max)
    assert net.n_params == it.n_params and net.n_mlp == it.n_mlp_params and net.k0 == it.encoded_width
    net.set_params(params)
    return net


@pytest.mark.parametrize("cfg", [NET_SMALL, NET_LADYBUG], ids=["k16", "k32"])
def test_network_forward_simt_and_tcgen05(oracle, cfg):
    spec = mixed_problem_2d(200)
    oprob, gprob = build_both(oracle, spec)
    it = _guided(gprob, network=cfg)
    rng = np.random.RandomState(4)
    params = it.get_params()
    assert np.abs(params[:it.n_mlp_params]).max() <= np.sqrt(6.0 / (16 + 64)) + 1e-6  # Xavier bound of the widest-range matrix
    assert np.abs(params[it.n_mlp_params:]).max() <= 1e-4 + 1e-9                       # grid init range (grid.h:1361)
    params = (rng.randn(it.n_params) * 0.25).astype(np.float32)
    it.set_params(params)
    np.testing.assert_array_equal(it.get_params(), params)
    net = _oracle_net(oracle, it, cfg, spec["aabb_min"], spec["aabb_max"], params)
    pts = (rng.rand(1000, 2) * 3.2 - 1.6).astype(np.float32)  # includes points outside the AABB (wrap-around indexing)
    want, _, _, _ = net.forward(pts)
    simt = it.infer(pts, impl=1)
    kout = 8 * 4 + 1
    np.testing.assert_allclose(simt[:, :kout], want[:, :kout], rtol=2e-3, atol=2e-3)
    tc = it.infer(pts, impl=0)
    np.testing.assert_allclose(tc[:, :kout], simt[:, :kout], rtol=1e-3, atol=1e-3)
    # query_network == one row of the same function
    np.testing.assert_allclose(it.queryNetwork(pts[3]), simt[3, :kout], rtol=1e-6, atol=1e-6)


def test_network_forward_3d_tcgen05(oracle):
    from elaina_b200 import api, scenes
    dv, dp = scenes.icosphere(3, 1.0, 0.05)
    prob = api.Problem(3, dirichlet=api.Scene(3, dv, dp), color_dirichlet=scenes.gray_pairs(scenes.harmonic_3d(dv)),
                       grid_scale=0.8, grid_pos=(0, 0, 0), grid_up=(0, 1, 0), grid_right=(1, 0, 0), aabb_min=(-1.2, -1.2, -1.2), aabb_max=(1.2, 1.2, 1.2))
    it = _guided(prob, frame=(16, 16), network=NET_SMALL)
    rng = np.random.RandomState(8)
    params = (rng.randn(it.n_params) * 0.25).astype(np.float32)
    it.set_params(params)
    net = _oracle_net(oracle, it, NET_SMALL, (-1.2, -1.2, -1.2), (1.2, 1.2, 1.2), params)
    pts = (rng.rand(777, 3) * 2.4 - 1.2).astype(np.float32)
    want, _, _, _ = net.forward(pts)
    tc = it.infer(pts, impl=0)
    np.testing.assert_allclose(tc[:, :41], want[:, :41], rtol=3e-3, atol=3e-3)


def _numpy_train_step(oracle, net, cfg, samples, loss_scale=128.0):
    """forward (oracle) -> KL gradient (oracle) -> backward + Adam + EMA (numpy, fp64), mirroring tcnn's order
    (fully_fused_mlp.cu:150-260, grid.h:395-540, adam.h:48-118, ema.h:44-60). -> (grad slab, new params, new ema)"""
    pos, d, nn, pdf, li, on = samples
    n = len(pos)
    out, acts, ci, cw = net.forward(pos, want_acts=True, want_cells=True)
    dl, loss = oracle.kl_grad(net.dim, out, d, nn, on, pdf, li, loss_scale / n)
    dz = dl.astype(np.float16).astype(np.float64)  # padded to 48 columns
    k0, P = net.k0, net.params.astype(np.float64)
    grad = np.zeros(net.n_params)
    dims = [(k0, 64), (64, 64), (64, 64), (64, 48)]
    offs = np.cumsum([0] + [k * m for k, m in dims])
    for l in (3, 2, 1, 0):
        K, N = dims[l]
        a = acts[:, l, :K].astype(np.float64)
        W = P[offs[l]:offs[l] + K * N].reshape(N, K).astype(np.float16).astype(np.float64)
        grad[offs[l]:offs[l] + K * N] = (dz.T @ a).reshape(-1)
        da = dz @ W
        if l > 0:
            dz = (da * (a > 0)).astype(np.float16).astype(np.float64)
        else:
            g = np.zeros((net.n_params - net.n_mlp) // 4 * 4).reshape(-1, 4)
            de = da.reshape(n, net.n_levels, 4) if k0 == net.n_levels * 4 else da[:, :net.n_levels * 4].reshape(n, net.n_levels, 4)
            np.add.at(g, ci.reshape(-1), (cw[..., None] * de[:, :, None, :]).reshape(-1, 4))
            grad[net.n_mlp:] = g.reshape(-1)
    o = cfg["optimizer"]["nested"]
    g = grad / loss_scale
    m = np.zeros_like(g)
    v = np.zeros_like(g)
    upd = np.ones(len(g), bool)
    upd[net.n_mlp:] = g[net.n_mlp:] != 0
    g[:net.n_mlp] += o["l2_reg"] * P[:net.n_mlp]
    m = (1 - o["beta1"]) * g
    v = (1 - o["beta2"]) * g * g
    lr = o["learning_rate"] * np.sqrt(1 - o["beta2"]) / (1 - o["beta1"])
    newp = np.where(upd, P - lr / (np.sqrt(v) + o["epsilon"]) * m, P)
    return grad, newp, newp.copy(), float(loss.mean())  # first EMA step: debias makes ema == weights


def _random_samples(rng, n, dim, lo, hi):
    pos = (lo + rng.rand(n, dim) * (hi - lo)).astype(np.float32)
    d, nn = _rand_dirs(rng, n, dim), _rand_dirs(rng, n, dim)
    pdf = (0.05 + rng.rand(n) * 0.4).astype(np.float32)
    li = (rng.rand(n) * 2.0).astype(np.float32)
    on = (rng.rand(n) < 0.3).astype(np.uint8)
    return pos, d, nn, pdf, li, on


@pytest.mark.parametrize("cfg", [NET_SMALL, NET_LADYBUG], ids=["k16", "k32"])
def test_training_step_matches_numpy(oracle, cfg):
    spec = mixed_problem_2d(200)
    oprob, gprob = build_both(oracle, spec)
    it = _guided(gprob, network=cfg)
    rng = np.random.RandomState(12)
    params = (rng.randn(it.n_params) * 0.2).astype(np.float32)
    it.set_params(params)
    net = _oracle_net(oracle, it, cfg, spec["aabb_min"], spec["aabb_max"], params)
    samples = _random_samples(rng, 128 * 24, 2, -1.4, 1.4)
    want_grad, want_p, want_ema, want_loss = _numpy_train_step(oracle, net, cfg, samples)
    loss, grad = it.train_batch(*samples)
    assert loss == pytest.approx(want_loss, rel=2e-3, abs=1e-4)
    # gradient slab: fp16 dZ operands on both sides, different summation order -> compare per block against the block's scale
    nm = it.n_mlp_params
    for lo_, hi_ in [(0, nm), (nm, it.n_params)]:
        scale = np.abs(want_grad[lo_:hi_]).max() + 1e-12
        assert np.max(np.abs(grad[lo_:hi_] - want_grad[lo_:hi_])) / scale < 2e-2, (lo_, hi_)
        cos = np.dot(grad[lo_:hi_], want_grad[lo_:hi_]) / (np.linalg.norm(grad[lo_:hi_]) * np.linalg.norm(want_grad[lo_:hi_]) + 1e-30)
        assert cos > 0.9995
    # Adam's first step moves every touched weight by ~lr against the sign of its gradient
    newp = it.get_params()
    moved = np.abs(want_p - params) > 1e-4
    agree = np.sign(newp - params)[moved] == np.sign(want_p - params)[moved]
    strong = np.abs(want_grad[moved]) > 1e-3 * np.abs(want_grad).max()
    assert np.mean(agree[strong]) > 0.999
    np.testing.assert_allclose(newp[moved][strong], want_p[moved][strong], atol=2e-3)
    np.testing.assert_allclose(it.get_params(inference=True), newp, atol=1e-6)  # first EMA step == weights
    # a second step keeps the EMA between old and new weights
    it.train_batch(*samples, want_grad=False)
    p2, e2 = it.get_params(), it.get_params(inference=True)
    lo2, hi2 = np.minimum(newp, p2) - 1e-6, np.maximum(newp, p2) + 1e-6
    assert np.all((e2 >= lo2) & (e2 <= hi2))


def _stderr_band(sol_a, sol_b, se):
    """|a - b| <= 3 standard errors of the difference of two estimates with the MEASURED per-pixel standard error `se`
    (wost_film_stderr, one accumulator slot per sample). At the handful of samples these same-stream comparisons run, a pixel
    whose samples all returned the same value has se = 0 although one flipped decision changes it, so the per-pixel value is
    floored by the frame median of the measured errors."""
    se = np.asarray(se, np.float64).reshape(sol_a.shape)
    floor = np.median(se[se > 0]) if np.any(se > 0) else 0.0
    return np.abs(sol_a - sol_b) <= 3.0 * np.sqrt(2.0) * np.sqrt(se ** 2 + floor ** 2) + 1e-4


@pytest.mark.parametrize("mlp_impl", [1, 0], ids=["simt", "tcgen05"])
def test_guided_walk_fixed_weights_vs_oracle(oracle, mlp_impl):
    """Guiding phase only (trainSppCount = 0), network weights set by hand to a smooth, clearly non-uniform mixture:
    the device walks use the same Philox draws as the oracle's guided walker, so the two fields agree far inside the
    3-s.e. band (walks only part where an fp16-level difference of the selection probability flips a decision)."""
    from elaina_b200 import api
    spec = mixed_problem_2d(300)
    oprob, gprob = build_both(oracle, spec)
    W, spp, depth = 24, 8, 16
    it = _guided(gprob, frame=(W, W), depth=depth, spp=spp, eps=5e-3, seed=5, train_spp_count=0, max_guided_depth_guiding=6,
                 uniform_fraction_guiding=0.5, mlp_impl=mlp_impl, spp_batch=spp)
    rng = np.random.RandomState(21)
    params = np.zeros(it.n_params, np.float32)
    nm = it.n_mlp_params
    params[:nm] = (rng.randn(nm) * 0.18).astype(np.float32)
    params[nm:] = (rng.randn(it.n_params - nm) * 0.5).astype(np.float32)
    it.set_params(params)
    it.solve()
    sol_g = it.film()[..., :3].reshape(-1, 3)
    net = _oracle_net(oracle, it, NET_SMALL, spec["aabb_min"], spec["aabb_max"], params)
    st = dict(frame_w=W, frame_h=W, max_walking_depth=depth, samples_per_pixel=spp, epsilon_shell=5e-3, seed=5)
    sol_o, steps_o, _ = oracle.solve_guided(oprob, st, net, max_guided_depth=6, uniform_fraction=0.5)
    assert abs(it.last_walk_steps - steps_o) <= 0.02 * steps_o
    close = np.isclose(sol_g, sol_o, rtol=2e-3, atol=2e-3)
    assert close.mean() > 0.9, close.mean()          # most pixels: identical decisions
    assert np.all(_stderr_band(sol_g, sol_o, it.film_stderr().reshape(-1, 3)))
    assert abs(sol_g.mean() - sol_o.mean()) < 0.02


def test_guided_fused_equals_validation_mode(oracle):
    spec = mixed_problem_2d(300)
    _, gprob = build_both(oracle, spec)
    sols = []
    for impl in (1, 0):
        it = _guided(gprob, frame=(32, 32), depth=16, spp=6, seed=9, train_spp_count=0, max_guided_depth_guiding=16,
                     uniform_fraction_guiding=0.0, mlp_impl=impl, network=NET_LADYBUG, spp_batch=6)
        rng = np.random.RandomState(3)
        p = (rng.randn(it.n_params) * 0.2).astype(np.float32)
        it.set_params(p)
        it.solve()
        sols.append(it.film()[..., :3].copy())
        se = it.film_stderr()
    # 16 guided bounces, always guided: an fp16-level difference of the raw outputs (tcgen05 vs SIMT summation order)
    # perturbs every sampled direction a little, and ~1e-3 of the ~100 discrete decisions per pixel (lobe pick, shell
    # test) flip, after which a walk is a different walk: most pixels agree closely, all agree statistically
    close = np.isclose(sols[0], sols[1], rtol=2e-2, atol=2e-2)
    assert close.mean() > 0.7, close.mean()
    assert np.all(_stderr_band(sols[0], sols[1], se))
    assert abs(sols[0].mean() - sols[1].mean()) < 0.01


def test_guided_training_learns_and_stays_unbiased():
    """Laplace on the star polygon with g = Re(z^3) (analytic solution known), Dirichlet only: training runs inside
    solve(); afterwards the field is within the 3-s.e. band of the analytic solution, optimiser steps were taken and
    the training loss of the last batch is finite."""
    from elaina_b200 import api, scenes
    dv, dp = scenes.star_polygon(1500)
    prob = api.Problem(2, dirichlet=api.Scene(2, dv, dp), color_dirichlet=scenes.gray_pairs(scenes.harmonic_2d(dv, 3)),
                       grid_scale=0.6, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0), aabb_min=(-1.4, -1.4), aabb_max=(1.4, 1.4))
    W, spp = 64, 96
    it = _guided(prob, frame=(W, W), depth=48, spp=spp, eps=1e-3, seed=42, train_spp_count=32, max_guided_depth_training=8,
                 max_guided_depth_guiding=8, uniform_fraction_training=0.5, uniform_fraction_guiding=0.5,
                 train_min_batch=2048, train_batch_size=8192, network=NET_SMALL, spp_batch=spp)
    p0 = it.get_params()
    it.solve()
    stats = it.train_stats
    assert stats["steps"] >= 32 and stats["samples"] > 2048 and np.isfinite(stats["loss"])
    p1 = it.get_params()
    assert np.isfinite(p1).all() and np.abs(p1 - p0).max() > 1e-3
    sol = it.film()[..., 0]
    g = 2.0 * np.arange(W, dtype=np.float32) / W - 1.0
    X, Y = np.meshgrid(0.6 * g, 0.6 * g)
    ref = scenes.harmonic_2d(np.stack([X, Y], -1), 3)
    # the corners of the frame lie outside the star-shaped curve (exterior problem, different solution): interior pixels only
    inside = np.hypot(X, Y) < 1.0 + 0.3 * np.sin(5.0 * np.arctan2(Y, X)) - 0.02
    assert inside.mean() > 0.9
    err = sol - ref
    se = it.film_stderr()[..., 0]                      # measured: one accumulator slot per sample (spp_batch == spp)
    z = (err / np.maximum(se, 1e-9))[inside]
    assert np.mean(np.abs(z) <= 3.0) >= 0.985, float(np.mean(np.abs(z) <= 3.0))  # nominal 99.7 %; 96 samples of a skewed distribution
    assert 0.6 <= np.mean(np.abs(z) <= 1.0) <= 0.78                           # calibrated errors: ~68 % within one
    assert abs(z.mean()) < 0.2
    relmse = np.mean((err ** 2 / (ref ** 2 + 1e-2))[inside])
    assert relmse < 1.3 * np.mean((se ** 2 / (ref ** 2 + 1e-2))[inside])


def test_guided_training_samples_match_oracle(oracle):
    """One training frame with the network frozen (learning rate 0): the compacted training set produced from the walk
    records (generate_training_data) has the oracle's size and statistics."""
    import copy
    spec = mixed_problem_2d(300)
    oprob, gprob = build_both(oracle, spec)
    cfg = copy.deepcopy(NET_SMALL)
    cfg["optimizer"]["nested"]["learning_rate"] = 0.0
    cfg["optimizer"]["nested"]["l2_reg"] = 0.0
    W, depth = 32, 16
    it = _guided(gprob, frame=(W, W), depth=depth, spp=1, eps=5e-3, seed=13, train_spp_count=1, max_guided_depth_training=6,
                 uniform_fraction_training=0.5, train_min_batch=128, train_batch_size=1024, network=cfg)
    rng = np.random.RandomState(2)
    params = (rng.randn(it.n_params) * 0.2).astype(np.float32)
    it.set_params(params)
    it.solve()
    net = _oracle_net(oracle, it, cfg, spec["aabb_min"], spec["aabb_max"], params)
    st = dict(frame_w=W, frame_h=W, max_walking_depth=depth, samples_per_pixel=1, epsilon_shell=5e-3, seed=13)
    sol_o, _, smp = oracle.solve_guided(oprob, st, net, max_guided_depth=6, uniform_fraction=0.5, emit_samples=True)
    stats = it.train_stats
    assert abs(stats["samples"] - len(smp["li"])) <= 0.02 * len(smp["li"]) + 2
    # the samples themselves (generate_training_data, train.h:422-471): same Philox draws and, with a fixed network, the same walks
    # except where an fp16-level difference of a selection probability flips a decision -- every device sample must have a
    # twin in the oracle's set (position, direction, normal, pdf, target, boundary flag), order aside
    got = it.last_training_samples()
    assert len(got["li"]) == stats["samples"]
    key_o = np.concatenate([smp["pos"], smp["dir"]], 1).astype(np.float64)
    key_g = np.concatenate([got["pos"], got["dir"]], 1).astype(np.float64)
    d2 = ((key_g[:, None, :] - key_o[None, :, :]) ** 2).sum(-1)
    j = d2.argmin(1)
    twin = np.sqrt(d2[np.arange(len(j)), j]) < 1e-4
    assert twin.mean() > 0.97, twin.mean()
    jj = j[twin]
    np.testing.assert_allclose(got["nrm"][twin], smp["nrm"][jj], atol=1e-5)
    assert np.array_equal(got["onN"][twin], smp["onN"][jj] > 0)
    # pdf and target depend on the network output (fp16 tensor-core forward vs the oracle's fp32 network): 2 % of the row's scale
    assert np.mean(np.abs(got["pdf"][twin] - smp["pdf"][jj]) <= 0.02 * np.abs(smp["pdf"][jj]) + 1e-4) > 0.98
    assert np.mean(np.abs(got["li"][twin] - smp["li"][jj]) <= 0.02 * np.abs(smp["li"][jj]) + 1e-3) > 0.95
    np.testing.assert_allclose(it.get_params(), params, atol=1e-7)  # lr = 0: weights untouched
    sol_g = it.film()[..., :3].reshape(-1, 3)
    assert np.isclose(sol_g, sol_o, rtol=2e-3, atol=2e-3).mean() > 0.9


def test_guided_3d_vmf_vs_oracle_and_training(oracle):
    """3D: vMF lobes, Poisson source term, triangles. Fixed weights vs the oracle's guided walker, then a short
    training run (records -> samples -> tcgen05 training kernel with K0 = 16, D = 3) that must stay finite."""
    from elaina_b200 import api, scenes
    dv, dp = scenes.icosphere(3, 1.0, 0.05)
    src = np.ones((8, 8, 8, 3), np.float32) * 0.5
    spec = dict(dim=3, dverts=dv, dprims=dp, color_dirichlet=scenes.gray_pairs(scenes.harmonic_3d(dv)),
                source=src, source_origin=(-1.6, -1.6, -1.6), source_voxel=0.4, source_intensity=1.0,
                grid_scale=0.7, grid_pos=(0.0, 0.0, 0.1), grid_up=(0.0, 1.0, 0.0), grid_right=(1.0, 0.0, 0.0),
                aabb_min=(-1.2, -1.2, -1.2), aabb_max=(1.2, 1.2, 1.2))
    oprob, gprob = build_both(oracle, spec)
    W, spp, depth = 20, 6, 24
    it = _guided(gprob, frame=(W, W), depth=depth, spp=spp, eps=4e-3, seed=17, train_spp_count=0, max_guided_depth_guiding=5,
                 uniform_fraction_guiding=0.5, spp_batch=spp)
    rng = np.random.RandomState(6)
    params = (rng.randn(it.n_params) * 0.2).astype(np.float32)
    it.set_params(params)
    it.solve()
    sol_g = it.film()[..., :3].reshape(-1, 3)
    net = _oracle_net(oracle, it, NET_SMALL, spec["aabb_min"], spec["aabb_max"], params)
    st = dict(frame_w=W, frame_h=W, max_walking_depth=depth, samples_per_pixel=spp, epsilon_shell=4e-3, seed=17)
    sol_o, steps_o, _ = oracle.solve_guided(oprob, st, net, max_guided_depth=5, uniform_fraction=0.5)
    assert abs(it.last_walk_steps - steps_o) <= 0.02 * steps_o
    assert np.isclose(sol_g, sol_o, rtol=5e-3, atol=5e-3).mean() > 0.85
    assert np.all(_stderr_band(sol_g, sol_o, it.film_stderr().reshape(-1, 3)))
    it2 = _guided(gprob, frame=(32, 32), depth=depth, spp=12, eps=4e-3, seed=3, train_spp_count=8, max_guided_depth_training=5,
                  max_guided_depth_guiding=5, uniform_fraction_training=0.5, uniform_fraction_guiding=0.5, train_min_batch=512, train_batch_size=2048)
    it2.solve()
    stats = it2.train_stats
    assert stats["steps"] >= 8 and np.isfinite(stats["loss"])
    assert np.isfinite(it2.get_params()).all() and np.isfinite(it2.film()).all()
