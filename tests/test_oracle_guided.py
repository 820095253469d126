"""CPU tests pinning the ORACLE's guided arithmetic (oracle/wost_oracle_guided.inl):
known answers lifted from the reference's own (commented-out) tests test/vonmises_test.cu:7,13-14,57-59,133,146,
sampler moments (:72-122), finite-difference checks of the mixture gradients, and the DenseGrid/MLP bookkeeping
(parameter counts of data/ladybug/n.json and data/fille/n.json: SURVEY.md 8(a) row A22)."""
import numpy as np
import pytest


def test_reference_known_answers(oracle):
    oracle._bind_guided()
    L = oracle.lib
    for mode in (1, 0):  # libm and detmath
        oracle.set_math_mode(mode)
        assert L.orc_eval_poly_large0(1.14514) == pytest.approx(0.4184690292340133, rel=1e-5)
        for x, want in zip([1.0, 2.0, 3.0, 4.0], [0.23591432, 0.82399356, 1.58530772, 2.42497277]):
            assert L.orc_log_bessel(x, 0) == pytest.approx(want, rel=1e-5)
        for th, lp, p in zip([-2.0, -1.0, 0.0, 1.0, 2.0], [-6.18411160, -2.16702533, -0.23629522, -2.16702533, -6.18411160],
                             [0.00206193, 0.11451776, 0.78954756, 0.11451776, 0.00206193]):
            c = float(np.float32(np.cos(th)))
            assert L.orc_vonmises_log_pdf(4.2, c) == pytest.approx(lp, rel=2e-5)
            assert L.orc_vonmises_pdf(4.2, c) == pytest.approx(p, rel=2e-5)
        c = float(np.float32(np.cos(0.5)))
        assert L.orc_vonmises_dlog_dkappa(1.45, c) == pytest.approx(0.29405486583709717, rel=1e-4)
        assert L.orc_vonmises_dlog_dkappa(14.5, c) == pytest.approx(-0.08729398250579834, rel=1e-4)
    oracle.set_math_mode(0)


def test_pdfs_integrate_to_one(oracle):
    oracle._bind_guided()
    L = oracle.lib
    th = np.linspace(-np.pi, np.pi, 20001)
    for kappa in (0.5, 4.2, 60.0, 2000.0):
        p = np.array([L.orc_vonmises_pdf(kappa, float(np.cos(t))) for t in th])
        assert np.trapezoid(p, th) == pytest.approx(1.0, rel=2e-3)
    ct = np.linspace(-1, 1, 200001)
    for kappa in (0.5, 5.0, 300.0):
        p = np.array([L.orc_vmf_pdf(kappa, float(c)) for c in ct])
        assert 2 * np.pi * np.trapezoid(p, ct) == pytest.approx(1.0, rel=2e-3)


def _single_lobe_raw(dim, kappa, mu, n):
    kout = 8 * (dim + 2) + 1
    raw = np.zeros((n, kout), np.float32)
    raw[:, 0::1] = 0.0
    for i in range(8):
        raw[:, i * (dim + 2)] = -10.0  # lambda = e^-10
        raw[:, i * (dim + 2) + 2] = 1.0
    raw[:, 0] = 10.0                    # lobe 0 dominates
    raw[:, 1] = np.log(kappa)
    raw[:, 2:2 + dim] = mu
    return raw


@pytest.mark.parametrize("kappa,tol", [(145.0, 0.05), (1.45, 0.02)])
def test_von_mises_sampler_moments(oracle, kappa, tol):
    # test/vonmises_test.cu:72-122: |mean| < 0.1, circular variance == 1 - I1/I0 within 5 % / 1 %
    n = 200000
    raw = _single_lobe_raw(2, kappa, (1.0, 0.0), n)
    d = np.tile(np.array([1.0, 0.0], np.float32), (n, 1))
    _, _, smp = oracle.vmm_eval(2, raw, d, d, np.zeros(n, np.uint8), seed=7)
    ang = np.arctan2(smp[:, 1], smp[:, 0])
    assert abs(ang.mean()) < 0.1
    R = np.hypot(np.cos(ang).mean(), np.sin(ang).mean())
    L = oracle.lib
    want = 1.0 - np.exp(L.orc_log_bessel(kappa, 1) - L.orc_log_bessel(kappa, 0))
    assert (1 - R) == pytest.approx(want, rel=tol)


def test_vmf_sampler_mean_resultant(oracle):
    n = 200000
    mu = np.array([0.0, 0.6, 0.8], np.float32)
    for kappa in (2.0, 50.0):
        raw = _single_lobe_raw(3, kappa, mu, n)
        d = np.tile(mu, (n, 1))
        _, _, smp = oracle.vmm_eval(3, raw, d, d, np.zeros(n, np.uint8), seed=3)
        assert np.allclose(np.linalg.norm(smp, axis=1), 1.0, atol=1e-4)
        want = 1.0 / np.tanh(kappa) - 1.0 / kappa  # E[mu . w] of a vMF lobe
        assert float((smp @ mu).mean()) == pytest.approx(want, abs=5e-3)


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("on_neumann", [0, 1])
def test_mixture_gradient_matches_finite_differences(oracle, dim, on_neumann):
    """VMM::gradients_probability (distribution.h:200-264 / 347-431) against central differences of the pdf in the
    ACTIVATED parameters (lambda, kappa, mu_original)."""
    rng = np.random.RandomState(5 + dim)
    n, nl = 16, dim + 2
    raw = (rng.randn(n, 8 * nl + 1) * 0.7).astype(np.float32)
    raw[:, 1::nl][:, :8] += 1.0  # kappa ~ e^1
    d = rng.randn(n, dim).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    nn = rng.randn(n, dim).astype(np.float32)
    nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    on = np.full(n, on_neumann, np.uint8)
    oracle.set_math_mode(1)
    pdf, grad, _ = oracle.vmm_eval(dim, raw, d, nn, on)
    worst = 0.0
    for sg in range(8):
        for j in range(nl):
            col = sg * nl + j
            h = 2e-3
            rp, rm = raw.copy(), raw.copy()
            if j < 2:  # exponential activation: perturb log-parameter, d/d(param) = d/d(raw) / param
                rp[:, col] += h
                rm[:, col] -= h
                scale = 1.0 / np.exp(raw[:, col])
            else:
                rp[:, col] += h
                rm[:, col] -= h
                scale = 1.0
            pp, _, _ = oracle.vmm_eval(dim, rp, d, nn, on)
            pm, _, _ = oracle.vmm_eval(dim, rm, d, nn, on)
            fd = (pp.astype(np.float64) - pm) / (2 * h) * scale
            err = np.abs(fd - grad[:, col]) / (np.abs(grad).max(axis=1) + 1e-6)  # second-order FD error scales with the row, not the entry
            if dim == 3 and j == 1:
                continue  # the reference's vMF d/dkappa uses a fitted polynomial for kappa < 1 (distribution.h:389-397): not an exact derivative
            worst = max(worst, float(err.max()))
    oracle.set_math_mode(0)
    assert worst < 2e-2, worst


def test_kl_gradient_is_the_derivative_of_the_loss(oracle):
    dim, n = 2, 12
    rng = np.random.RandomState(11)
    raw = np.zeros((n, 48), np.float32)
    raw[:, :33] = rng.randn(n, 33) * 0.5
    raw[:, 2:32:4] += 1.0  # keep |mu_original| away from 0, where normalisation is singular and central differences are meaningless
    d = rng.randn(n, dim).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    on = np.zeros(n, np.uint8)
    pdf = (rng.rand(n) * 0.3 + 0.05).astype(np.float32)
    li = (rng.rand(n) * 2).astype(np.float32)
    oracle.set_math_mode(1)
    dl, loss = oracle.kl_grad(dim, raw, d, d, on, pdf, li, 1.0)
    for col in range(32):
        h = 5e-3
        rp, rm = raw.copy(), raw.copy()
        rp[:, col] += h
        rm[:, col] -= h
        lp = oracle.kl_grad(dim, rp, d, d, on, pdf, li, 1.0)[1].astype(np.float64)
        lm = oracle.kl_grad(dim, rm, d, d, on, pdf, li, 1.0)[1].astype(np.float64)
        fd = (lp - lm) / (2 * h)
        assert np.allclose(fd, dl[:, col], rtol=5e-2, atol=2e-3 * np.abs(dl).max()), col
    oracle.set_math_mode(0)
    assert np.all(dl[:, 33:] == 0)


def test_network_parameter_counts_match_the_reference_configs(oracle):
    # ladybug 61 536 grid + 13 312 MLP parameters, fille 1 056 + 12 288
    lady = oracle.net(2, 8, 8, 1.4049999713897705, (-100.0, -100.0), (600.0, 600.0))
    assert (lady.n_params - lady.n_mlp, lady.n_mlp, lady.k0) == (61536, 13312, 32)
    fille = oracle.net(2, 4, 4, 1.4049999713897705, (-1.0, -1.0), (1.0, 1.0))
    assert (fille.n_params - fille.n_mlp, fille.n_mlp, fille.k0) == (1056, 12288, 16)


def test_network_forward_against_numpy(oracle):
    net = oracle.net(2, 4, 4, 1.405, (-1.0, -1.0), (1.0, 1.0))
    rng = np.random.RandomState(0)
    params = (rng.randn(net.n_params) * 0.2).astype(np.float32)
    net.set_params(params)
    pts = (rng.rand(64, 2) * 2 - 1).astype(np.float32)
    out, acts, ci, cw = net.forward(pts, want_acts=True, want_cells=True)
    grid = params[net.n_mlp:].reshape(-1, 4).astype(np.float16).astype(np.float64)
    enc = np.einsum("nlc,nlcf->nlf", cw.astype(np.float64), grid[ci]).reshape(64, -1)
    a = enc.astype(np.float16).astype(np.float64)
    assert np.allclose(a, acts[:, 0, :16], atol=2e-3)  # fp32 vs fp64 interpolation sums may round to neighbouring halfs
    a = acts[:, 0, :16].astype(np.float64)
    off = 0
    for l, (K, N) in enumerate([(16, 64), (64, 64), (64, 64), (64, 48)]):
        W = params[off:off + K * N].reshape(N, K).astype(np.float16).astype(np.float64)
        off += K * N
        z = a @ W.T
        if l < 3:
            a = np.maximum(z, 0).astype(np.float16).astype(np.float64)
            assert np.allclose(a, acts[:, l + 1, :], atol=1e-2, rtol=2e-3)
            a = acts[:, l + 1, :].astype(np.float64)
        else:
            assert np.allclose(z, out, atol=1e-4, rtol=1e-4)
    assert np.allclose(cw.sum(axis=2), 1.0, atol=1e-5)
