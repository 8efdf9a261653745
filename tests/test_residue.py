"""WHERE the parity residue lies (SURVEY 8d "parity statistic"; north_star: "the residue confined to fp32-vs-fp64 edge and
refraction pixels"), and the random scenes a plain 99.9 % threshold cannot judge.

CPU half: the host-compiled copy of the per-ray device functions (tests/hostsim) against the oracle.  The GPU half
(tests/test_gpu_parity.py) runs the same checks on the CUDA path.  The classifier is `conftest.residue_report`: every pixel
that differs by more than 1/255 must be reference-unstable (the oracle's own value moves under a 1e-6 relative perturbation
of its rays and hit points), or lie on a silhouette / material edge / shadow edge, or in a mirror / refraction region."""
import numpy as np
import pytest

import hostsim
import oracle
import python_raytracer_b200 as rt
from conftest import (ALL_CASES, FUZZ_CHAOTIC_SEEDS, PARITY_FRACTION, load_bundled, load_case, load_fuzz, residue_report,
                      u8_cr, unstable_fraction)

abi = rt._abi


def test_stability_probe_is_deterministic_and_off_by_default():
    sc = load_case("MirrorRefraction", {"resolution": [90, 51]})
    fs = sc.flatten()
    cols, rows = np.arange(0, 90, 3), np.arange(0, 90, 3) % 51
    base, _ = oracle.render_pixels(fs, abi, cols, rows)
    assert np.array_equal(oracle.render_pixels_perturbed(fs, cols, rows, eps=0.0), base)       # eps 0: the reference itself
    a = oracle.render_pixels_perturbed(fs, cols, rows, eps=1e-6, seed=5)
    assert np.array_equal(a, oracle.render_pixels_perturbed(fs, cols, rows, eps=1e-6, seed=5))
    assert np.abs(a - base).max() < 0.5 and not np.array_equal(a, base)


@pytest.mark.parametrize("seed", FUZZ_CHAOTIC_SEEDS)
def test_chaotic_random_scenes_host_twin_matches_on_reference_stable_pixels(seed):
    """The 13 seeds round 1 screened out (6 of them miss the plain 99.9 % threshold, worst 99.15 %): every out-of-tolerance
    pixel is classified; the statistic over the reference-STABLE pixels must meet the tolerance, nothing may be unexplained,
    and the probe itself must call only a small part of the frame unstable."""
    sc = load_fuzz(seed)
    fs = sc.flatten()
    ref, _ = oracle.render(fs, abi)
    img, _ = hostsim.render(fs, abi)
    rep = residue_report(fs, ref, img)
    print(f"seed {seed}: {rep}")
    assert rep["frac_stable"] >= PARITY_FRACTION, rep
    assert rep["unexplained"] == 0, rep
    # the probe must not call everything unstable: 0.5-6 % of these frames are (textured planes sampled with the nearest
    # texel, simple_geometry.py:165-169, are unstable along every texel boundary), against 100 % of the residue
    assert unstable_fraction(fs, ref) <= 0.10


@pytest.mark.parametrize("case", ALL_CASES)
def test_golden_residue_is_on_edges_and_refraction_regions(case):
    sc = load_case(case)
    fs = sc.flatten()
    ref, _ = oracle.render(fs, abi)
    img, _ = hostsim.render(fs, abi)
    rep = residue_report(fs, ref, img)
    print(f"{case}: {rep}")
    assert rep["frac"] >= PARITY_FRACTION and rep["unexplained"] == 0, rep


@pytest.mark.parametrize("name,n_px", [("NovelScene1", 6000), ("NovelScene2", 1500)])
def test_bundled_csg_scenes_at_shipped_settings_sampled_host_twin(name, n_px):
    """scenes/NovelScene1.json:2-3 (2048x1024 x 32 spp) and NovelScene2.json:2-5 (1024x512 x 2*15*16 samples) at their
    SHIPPED settings: random pixels, all samples each, host twin vs oracle (the GPU test renders the full frame)."""
    sc = load_bundled(name)
    fs = sc.flatten()
    W, H = sc.vc.width, sc.vc.height
    assert (W, H) == ((2048, 1024) if name == "NovelScene1" else (1024, 512))
    rng = np.random.default_rng(3)
    cols, rows = rng.integers(0, W, n_px), rng.integers(0, H, n_px)
    ref, _ = oracle.render_pixels(fs, abi, cols, rows)
    img, _ = hostsim.render_pixels(fs, abi, cols, rows)
    d = np.abs(u8_cr(ref) - u8_cr(img)).max(axis=1)
    assert (d <= 1).mean() >= PARITY_FRACTION, ((d <= 1).mean(), d.max())
