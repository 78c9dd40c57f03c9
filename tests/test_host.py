"""CPU: the host-side C++ mirror (path-tracing_b200/host) — loader, Triangle constructor, Config parser — and that the
C-ABI library loads and exports every symbol include/pt_render.h declares.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import ptb200
import scenes
from util import GOLDEN, ROOT, assert_bits_equal

G = np.load(GOLDEN + "/golden.npz")


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "pt_render.h")).read()
    declared = set(re.findall(r"\b(pt_[a-z_]+)\s*\(", hdr))
    assert declared == set(ptb200.ABI_SYMBOLS), declared ^ set(ptb200.ABI_SYMBOLS)
    for s in declared:
        assert hasattr(lib.lib, s), s
    assert b"sm_100a" in lib.lib.pt_version()


def test_struct_layouts_match_header():
    assert C.sizeof(ptb200.PtConfig) == 4 * 18
    assert C.sizeof(ptb200.PtScene) == 80
    assert C.sizeof(ptb200.PtStats) == 4 * 8 + 3 * 4 + 4 * 4 + 8 * 4 + 4 + 17 * 8 + 8  # 8-byte aligned


def test_no_device_fails_loudly(lib):
    if lib.device_count() > 0:
        pytest.skip("a GPU is present")
    sc = scenes.cornell()
    with pytest.raises(ptb200.PtError) as e:
        lib.render(sc, ptb200.config(8, 8, 1))
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)


def test_argument_errors(lib):
    sc = scenes.cornell()
    bad = dict(sc)
    bad["mat_ids"] = sc["mat_ids"].copy()
    bad["mat_ids"][3] = 99
    with pytest.raises(ptb200.PtError) as e:
        lib.render(bad, ptb200.config(8, 8, 1))
    assert e.value.code == 1
    with pytest.raises(ptb200.PtError) as e:
        lib.render(sc, ptb200.config(0, 8, 1))
    assert e.value.code == 1


def test_loader_matches_reference_loader(tor):
    assert tor["verts"].shape[0] == int(G["tor_n"]) == 270
    assert_bits_equal(tor["verts"], G["tor_verts"], "vertices")
    assert_bits_equal(tor["planes"], G["tor_planes"], "planes (vn override)")
    assert_bits_equal(tor["squares"], G["tor_squares"], "squares")
    assert tor["mats"].shape == (5, 10)
    assert list(tor["mat_ids"][:256]) == [4] * 256
    assert list(tor["mat_ids"][256:]) == [2, 2, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 3, 3]
    np.testing.assert_allclose(tor["mats"][0], [2, 2, 2, 1, 1, 1, .5, .5, .5, 96.078431], rtol=1e-7)
    np.testing.assert_allclose(tor["mats"][3][3:6], [1, .01, .01], rtol=1e-7)


def test_loader_missing_file(loader):
    with pytest.raises(FileNotFoundError):
        loader.load_obj("/nonexistent", "x.obj")


def test_host_triangle_constructor(loader):
    sc = scenes.cornell()
    planes, squares = loader.triangle_setup(sc["verts"])
    assert_bits_equal(planes, G["cornell_planes"])
    assert_bits_equal(squares, G["cornell_squares"])


def test_config_parser(loader):
    c = loader.parse_config(["prog", "--W", "1920", "--H", "1080", "-RPP", "256", "-MRR", "8", "-ERR", "-1", "-SEED", "7",
                             "-EPS"])  # trailing unpaired flag is ignored (config.h:36)
    assert (c.width, c.height, c.rays_per_pixel, c.max_ray_reflections, c.seed) == (1920, 1080, 256, 8, 7)
    assert c.error == -1.0 and abs(c.eps - 1e-4) < 1e-10
