"""CPU tier: the oracle is pinned -- (1) the C restatement (oracle/oracle_rt.c) is bit-identical to the
reference's own kernels compiled as C++ (oracle/_ref) wherever that library is present, (2) both reproduce the
committed golden vectors (KAT-1 / KAT-2 of SURVEY.md section 4, tests/golden/kat.json)."""
import json
import os

import numpy as np
import pytest

from helpers import frame_rays
from pyrt import scenes as S
from pyrt import types as T

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "kat.json")))


def _scene(name):
    return S.kat2_scene() if name == "kat2" else S.load_reference_scene(name)


@pytest.mark.parametrize("kat", KAT["cases"], ids=lambda k: k["name"])
def test_port_reproduces_golden_vectors(oracle_mod, kat):
    objs, lights = _scene(kat["scene"])
    W, H = kat["width"], kat["height"]
    z, rays = frame_rays(W, H)
    assert abs(float(z) - kat["zdir"]) <= 1e-3
    rgba, hid, ht, st = oracle_mod.port_render(kat["mode"], kat["max_bounces"], objs, lights, rays)
    for px in kat["pixels"]:
        got = rgba[px["jj"] * W + px["ii"], :3]
        assert np.allclose(got, px["rgb"], rtol=0, atol=2e-7), (px, got)
    if "hit_pixels" in kat:
        assert int((hid >= 0).sum()) == kat["hit_pixels"] == st.hit_pixels
    if "sum_rgb" in kat:
        assert np.allclose(rgba[:, :3].astype(np.float64).sum(axis=0), kat["sum_rgb"], rtol=1e-9)
    if "rays" in kat:
        assert [st.rays_primary, st.rays_shadow, st.rays_reflect] == kat["rays"]


@pytest.mark.parametrize("name", list(S.REFERENCE_SCENES) + ["simpleScene3L", "kat2"])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_port_equals_reference_kernels(oracle_mod, name, mode):
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    objs, lights = _scene(name)
    _, rays = frame_rays(160, 90)
    a = oracle_mod.ref_render(mode, 30, objs, lights, rays)
    b, hid, ht, st = oracle_mod.port_render(mode, 30, objs, lights, rays)
    assert np.array_equal(a, b)
    assert np.array_equal(hid >= 0, ht < T.MAX_FLOAT)


def test_port_equals_reference_kernels_synthetic(oracle_mod):
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref not built")
    objs, lights = S.synthetic_scene(1500, 200)
    _, rays = frame_rays(96, 54)
    for mode in (1, 2):
        a = oracle_mod.ref_render(mode, 4, objs, lights, rays)
        b, *_ = oracle_mod.port_render(mode, 4, objs, lights, rays)
        assert np.array_equal(a, b)


def test_ray_count_formula(oracle_mod):
    """a fully absorbing hit pixel with one light costs 3 rays (primary, shadow, wasted reflection)."""
    objs, lights = S.load_reference_scene("simpleSphere")
    _, rays = frame_rays(160, 90)
    _, hid, _, st = oracle_mod.port_render(2, 30, objs, lights, rays)
    hits = int((hid >= 0).sum())
    assert (st.rays_primary, st.rays_shadow, st.rays_reflect) == (len(rays), hits, hits)


def test_quantize_matches_ppm_exporter_rule():
    c = np.array([[0.0, 0.3, 1.0, 1.0], [1.5, -0.2, 0.999, 1.0], [0.00392157, 0.5, 0.25, 1.0]], dtype=np.float32)
    q = T.quantize_ppm(c)
    assert q.tolist() == [[0, 76, 255], [255, -51, 254], [1, 127, 63]]
