"""Statistical tests of the reference for the phase functions and BSDFs on the path (chi^2 of `sample`
against `eval` / `pdf` in src/phase/tests/test_rayleigh.py, test_hg.py (g = 0.6), test_isotropic.py,
src/bsdfs/tests/test_diffuse.py, test_rpv.py), restated against the oracle's functions (CPU only; the device functions
are held bit-exact to these by the parity suite).  The reference's ChiSquareTest bins the sphere in (cos theta, phi),
integrates the pdf over each bin and compares with a histogram of samples at significance 0.01; so does this file."""
import ctypes as C

import numpy as np
import pytest
from scipy import stats

N_SAMPLES = 400000
RNG_SEED = 7


def chi2_sphere(dirs, density, n_mu=24, n_phi=12, axis=(0.0, 0.0, 1.0), mu_range=(-1.0, 1.0), sub=16):
    """Pearson chi^2 of `dirs` (n x 3 unit vectors) against `density(mu, phi)` (per steradian, in polar coordinates
    around `axis`).  Bins with an expected count below 5 are pooled, as in the reference's ChiSquareTest.  Returns the
    p-value and the integral of the density over the binned domain."""
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    s = np.cross(axis, [1.0, 0.0, 0.0] if abs(axis[0]) < 0.9 else [0.0, 1.0, 0.0]); s /= np.linalg.norm(s)
    t = np.cross(axis, s)
    mu = np.clip(dirs @ axis, -1, 1)
    phi = np.arctan2(dirs @ t, dirs @ s) % (2 * np.pi)
    mu_edges = np.linspace(mu_range[0], mu_range[1], n_mu + 1)
    phi_edges = np.linspace(0, 2 * np.pi, n_phi + 1)
    inside = (mu >= mu_range[0]) & (mu <= mu_range[1])
    observed, _, _ = np.histogram2d(mu[inside], phi[inside], bins=[mu_edges, phi_edges])
    # composite midpoint rule inside each bin
    mu_f = (np.arange(n_mu * sub) + 0.5) / (n_mu * sub) * (mu_range[1] - mu_range[0]) + mu_range[0]
    phi_f = (np.arange(n_phi * sub) + 0.5) / (n_phi * sub) * 2 * np.pi
    dens = density(mu_f[:, None], phi_f[None, :]) * np.ones((mu_f.size, phi_f.size))
    cell = (mu_range[1] - mu_range[0]) / (n_mu * sub) * 2 * np.pi / (n_phi * sub)
    prob = dens.reshape(n_mu, sub, n_phi, sub).sum(axis=(1, 3)) * cell
    expected = prob * len(dirs)
    o, e = observed.ravel(), expected.ravel()
    order = np.argsort(e)
    o, e = o[order], e[order]
    small = np.cumsum(e) < 5
    if small.any():
        o = np.concatenate([[o[small].sum()], o[~small]]); e = np.concatenate([[e[small].sum()], e[~small]])
    keep = e > 0
    chi2 = ((o[keep] - e[keep]) ** 2 / e[keep]).sum()
    dof = keep.sum() - 1
    return 1.0 - stats.chi2.cdf(chi2, dof), prob.sum()


def _phase(pkg, kind, g=0.0, values=None):
    abi = pkg.abi
    ph = abi.Phase()
    ph.type = {"isotropic": abi.PHASE_ISOTROPIC, "rayleigh": abi.PHASE_RAYLEIGH, "hg": abi.PHASE_HG,
               "tabphase": abi.PHASE_TABULATED}[kind]
    ph.g = g
    keep = None
    if values is not None:
        keep = (C.c_double * len(values))(*values)
        ph.n_values, ph.values = len(values), keep
    return ph, keep


PHASES = [("isotropic", {}), ("rayleigh", {}), ("hg", {"g": 0.6}), ("hg", {"g": -0.3}), ("hg", {"g": 0.0}),
          ("tabphase", {"values": [0.5, 1.0, 1.5, 0.75, 2.5, 4.0]})]


@pytest.mark.parametrize("wi", [[0.0, 0.0, 1.0], [0.48, -0.6, 0.64]])
@pytest.mark.parametrize("kind,kw", PHASES)
def test_phase_sample_follows_eval(pkg, oracle, kind, kw, wi):
    """test_rayleigh.py / test_hg.py / test_isotropic.py: chi^2 of sample() against eval(); eval() integrates to 1"""
    ph, keep = _phase(pkg, kind, **kw)
    wi = np.asarray(wi) / np.linalg.norm(wi)
    u = np.random.default_rng(RNG_SEED).random((N_SAMPLES, 3))
    wo, value = oracle.phase_samples(pkg, ph, wi, u)
    assert np.allclose(np.linalg.norm(wo, axis=1), 1, atol=1e-9)

    def density(mu, phi):   # phase functions depend on dot(wo, wi) only
        mu = np.broadcast_to(mu, np.broadcast(mu, phi).shape)
        flat = mu[:, 0]
        ref = np.array([oracle.phase_eval(pkg, ph, wi, _dir_at(wi, m)) for m in flat])
        return ref[:, None]
    p, total = chi2_sphere(wo, density, axis=wi)
    assert abs(total - 1) < 2e-3, "eval() integrates to %.5f" % total
    assert p > 0.01, "chi^2 p-value %.2e for %s %s" % (p, kind, kw)
    # the value returned with a sample is eval() in that direction
    for i in (0, 17, 4711):
        assert np.isclose(value[i], oracle.phase_eval(pkg, ph, wi, wo[i]), rtol=1e-12)


def _dir_at(axis, mu):
    axis = np.asarray(axis, float)
    s = np.cross(axis, [1.0, 0.0, 0.0] if abs(axis[0]) < 0.9 else [0.0, 1.0, 0.0]); s /= np.linalg.norm(s)
    return mu * axis + np.sqrt(max(0.0, 1 - mu * mu)) * s


def _bsdf(pkg, kind, *p):
    abi = pkg.abi
    b = abi.Bsdf()
    b.type = {"diffuse": abi.BSDF_DIFFUSE, "rpv": abi.BSDF_RPV, "bilambertian": abi.BSDF_BILAMBERTIAN}[kind]
    for i, v in enumerate(p):
        b.p[i] = v
    return b


BSDFS = [("diffuse", (0.5,)), ("rpv", (0.1, 0.7, -0.1, 0.1)), ("rpv", (0.3, 1.2, 0.25, 0.6)), ("bilambertian", (0.5, 0.25)),
         ("bilambertian", (0.0, 0.6))]


@pytest.mark.parametrize("wi", [[0.0, 0.0, 1.0], [0.6, 0.0, 0.8], [-0.3, 0.5, 0.2]])
@pytest.mark.parametrize("kind,params", BSDFS)
def test_bsdf_sample_follows_pdf(pkg, oracle, kind, params, wi):
    """test_diffuse.py / test_rpv.py: chi^2 of sample() against pdf(); weight = eval / pdf; pdf integrates to 1"""
    b = _bsdf(pkg, kind, *params)
    wi = np.asarray(wi) / np.linalg.norm(wi)
    u = np.random.default_rng(RNG_SEED + 1).random((N_SAMPLES, 3))
    wo, value, pdf, weight = oracle.bsdf_samples(pkg, b, wi, u)
    ok = pdf > 0
    assert ok.mean() > 0.99
    if kind == "rpv":
        # SURVEY F8, reproduced on purpose: rpv::sample returns the BRDF value f (incl. 1 / pi) as the path weight,
        # not f cos / pdf (src/bsdfs/rpv.cpp:97-105); eval() = f |cos theta_o|
        assert np.allclose(weight[ok], value[ok] / wo[ok, 2], rtol=1e-9, atol=1e-12)
    else:
        assert np.allclose(weight[ok], value[ok] / pdf[ok], rtol=1e-9, atol=1e-12)

    def density(mu, phi):
        mu, phi = np.broadcast_arrays(mu, phi)
        st = np.sqrt(np.maximum(0, 1 - mu * mu))
        d = np.stack([st * np.cos(phi), st * np.sin(phi), mu], axis=-1).reshape(-1, 3)
        return oracle.bsdf_eval(pkg, b, wi, d)[1].reshape(mu.shape)
    p, total = chi2_sphere(wo, density, axis=(0, 0, 1))
    assert abs(total - 1) < 3e-3, "pdf() integrates to %.5f" % total
    assert p > 0.01, "chi^2 p-value %.2e for %s %s" % (p, kind, params)


def test_diffuse_values(pkg, oracle):
    """test_diffuse.py: eval = rho / pi * cos(theta_o) on the upper hemisphere, 0 below; pdf = cos / pi"""
    b = _bsdf(pkg, "diffuse", 0.5)
    wi = np.array([0.0, 0.0, 1.0])
    th = np.linspace(0, np.pi, 20)
    wo = np.stack([np.sin(th), np.zeros_like(th), np.cos(th)], axis=1)
    value, pdf = oracle.bsdf_eval(pkg, b, wi, wo)
    up = np.cos(th) > 0
    assert np.allclose(value[up], 0.5 / np.pi * np.cos(th[up]), rtol=1e-12) and np.all(value[~up] == 0)
    assert np.allclose(pdf[up], np.cos(th[up]) / np.pi, rtol=1e-12) and np.all(pdf[~up] == 0)
