"""CPU: the oracle against the committed golden manifest (tests/golden/manifest.json, made by
tests/golden/make_manifest.py in the container that holds /root/reference).  Pins the oracle: any edit to
oracle/ or a library upgrade that changes a result shows up as a digest mismatch."""
import json
import os

import numpy as np
import pytest

import golden_cases as G


@pytest.fixture(scope="module")
def manifest():
    with open(G.MANIFEST) as f:
        return json.load(f)


def test_manifest_covers_every_case(manifest):
    ids = [cid for cid, _, _, _ in G.cases()]
    assert len(ids) == len(set(ids)) == len(manifest["cases"]) and set(ids) == set(manifest["cases"])
    assert set(G.inputs()) == set(manifest["inputs"]) and "natural_120x160" in manifest["inputs"]
    for name, img in G.inputs().items():
        assert G.digest(img) == manifest["inputs"][name], name
    # every noise of SURVEY 8(a) a2-a11 + a13 and every denoiser a16-a20 appears
    specs = {cid.split("|")[1] for cid in ids}
    for noise in ("gaussian", "poisson", "sap", "speckle", "quant", "uniform", "brownian", "periodic", "gamma", "rayleigh", "shader"):
        assert any(s.startswith(noise) for s in specs), noise
    for den in ("gaus_blur", "mean", "median", "bilateral", "wavelet"):
        assert any(("_" + den) in s for s in specs), den
    # the whole ladder, no cell skipped: noise x {none, 5 denoisers} x its three levels
    from oracle import closures as C
    for name in G.inputs():
        have = {cid.split("|")[1] for cid in ids if cid.startswith(name + "|")}
        for noise in C.NOISE_ORDER[:10]:
            for lvl in C.LEVELS.get(noise, [None]):
                for den in ("", "_wavelet", "_gaus_blur", "_mean", "_median", "_bilateral"):
                    assert noise + den + ("_var" + lvl if lvl else "") in have, (name, noise, den, lvl)
    assert len(ids) >= 330


def test_oracle_reproduces_manifest(manifest):
    if G.versions() != manifest["versions"]:
        pytest.skip("library versions differ from the manifest's: %r" % (G.versions(),))
    bad = []
    for cid, img, spec, fields in G.cases():
        if G.digest(G.evaluate(img, spec, fields)) != manifest["cases"][cid]["sha256"]:
            bad.append(cid)
    assert not bad, bad[:10]


def test_digest_is_shape_sensitive():
    a = np.zeros((2, 3, 3), np.uint8)
    assert G.digest(a) != G.digest(a.reshape(3, 2, 3))
