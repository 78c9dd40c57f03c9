"""CPU: the oracle (oracle/pt_oracle.c, plain-C restatement) against the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py) and, where /root/reference was compiled, against the reference live."""
import numpy as np
import pytest

import scenes
from util import GOLDEN, MODELS, assert_bits_equal, with_planes

G = np.load(GOLDEN + "/golden.npz")
GB = np.load(GOLDEN + "/golden_big.npz")  # BASELINE configs[2] / configs[3] pins, tests/golden/make_golden_big.py


def tor_scene():
    # geometry from the golden dump; materials from Tor.mtl (Tor.mtl:4-52) in file order
    ns = 96.078431
    mats = np.array([[2, 2, 2, 1, 1, 1, .5, .5, .5, ns], [0, 0, 0, 1, 1, 1, .5, .5, .5, ns],
                     [0, 0, 0, .01, 1, .01, .5, .5, .5, ns], [0, 0, 0, 1, .01, .01, .5, .5, .5, ns],
                     [0, 0, 0, .5, 1, .01, .5, .5, .5, ns]], np.float32)
    ids = np.concatenate([np.full(256, 4), [2, 2, 0, 0], np.full(8, 1), [3, 3]]).astype(np.int32)
    return {"verts": G["tor_verts"], "planes": G["tor_planes"], "squares": G["tor_squares"], "mat_ids": ids, "mats": mats}


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32 10 rounds
    assert oracle.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert oracle.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert oracle.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    assert_bits_equal(G["philox_kat"][0], np.array([0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8], np.uint32))


def test_triangle_setup_matches_reference_dump(oracle):
    sc = scenes.cornell()
    planes, squares = oracle.triangle_setup(sc["verts"])
    assert_bits_equal(planes, G["cornell_planes"], "cornell planes")
    assert_bits_equal(squares, G["cornell_squares"], "cornell squares")


@pytest.mark.parametrize("name", ["tor_cam", "tor_rnd"])
def test_nearest_hit_tor(oracle, name):
    ids, t = oracle.intersect(tor_scene(), G[name + "_rays"])
    assert_bits_equal(ids, G[name + "_ids"], name + " ids")
    assert_bits_equal(t, G[name + "_t"], name + " t")
    assert (ids >= 0).mean() > 0.9


@pytest.mark.parametrize("name,make", [("cornell", scenes.cornell), ("glossy", scenes.glossy_mix),
                                       ("torus40", lambda: scenes.torus_room(40, 20))])
def test_nearest_hit_synthetic(oracle, name, make):
    sc = with_planes(make(), oracle)
    ids, t = oracle.intersect(sc, G[name + "_rays"])
    assert_bits_equal(ids, G[name + "_ids"], name + " ids")
    assert_bits_equal(t, G[name + "_t"], name + " t")


def test_primary_rays(oracle):
    cfg = oracle.config(64, 64, 256)
    for p, want in zip(G["prim_xys"], G["prim_rays"]):
        o, d = oracle.primary_ray(cfg, *map(int, p))
        assert_bits_equal(np.concatenate([o, d]), want, f"primary {p}")


def render_case(oracle, sc, name):
    W, H, spp, depth, seed, sb, ss = map(int, G[name + "_cfg"])
    cfg = oracle.config(W, H, spp, depth, 1e-4, float(G[name + "_error"]), seed, sb, ss)
    sm, s2, cnt, segs, _ = oracle.render_accum(sc, cfg)
    assert_bits_equal(cnt, G[name + "_count"], name + " count")
    assert_bits_equal(sm, G[name + "_sum"], name + " sum")
    assert_bits_equal(s2, G[name + "_sum2"], name + " sum2")
    assert segs == int(G[name + "_segments"])
    return cnt


@pytest.mark.parametrize("name", ["tor_flat", "tor_adapt", "tor_part"])
def test_render_tor(oracle, name):
    cnt = render_case(oracle, tor_scene(), name)
    assert cnt.sum() > 0


@pytest.mark.parametrize("name,make", [("cornell", scenes.cornell), ("glossy", scenes.glossy_mix),
                                       ("torus40", lambda: scenes.torus_room(40, 20))])
def test_render_synthetic(oracle, name, make):
    render_case(oracle, with_planes(make(), oracle), name)


def composite_scene(loader):
    """BASELINE configs[3] (C4): Tor.obj as the product's loader reads it + the two Cornell boxes, 290 triangles."""
    return scenes.composite(loader.load_obj(MODELS, "Tor.obj"))


def render_case_big(oracle, sc, name):
    W, H, spp, depth, seed, sb, ss = map(int, GB[name + "_cfg"])
    sm, s2, cnt, segs, _ = oracle.render_accum(sc, oracle.config(W, H, spp, depth, 1e-4, float(GB[name + "_error"]), seed, sb, ss))
    assert_bits_equal(cnt, GB[name + "_count"], name + " count")
    assert_bits_equal(sm, GB[name + "_sum"], name + " sum")
    assert_bits_equal(s2, GB[name + "_sum2"], name + " sum2")
    assert segs == int(GB[name + "_segments"])


def test_composite_c4_golden(oracle, loader):
    """C4's scene: ray set, a 32x32x16 frame and one 1-of-4 sample partition of it, against the unmodified reference."""
    sc = with_planes(composite_scene(loader), oracle)
    ids, t = oracle.intersect(sc, GB["composite_rays"])
    assert_bits_equal(ids, GB["composite_ids"], "composite ids")
    assert_bits_equal(t, GB["composite_t"], "composite t")
    render_case_big(oracle, sc, "composite")
    render_case_big(oracle, sc, "composite_part")


def test_torus1m_c3_golden_subset(oracle):
    """C3's own scene (1,000,014 triangles): every 16th ray of the golden set (the oracle is brute force like the
    reference: the whole set takes the host cores half a minute; the GPU test runs all of it)."""
    sc = scenes.torus_room(1000, 500)
    import ptb200
    planes, squares = ptb200.HostLoader().triangle_setup_bulk(sc["verts"])
    sc = dict(sc, planes=planes, squares=squares)
    sel = slice(5, None, 16)
    ids, t = oracle.intersect(sc, GB["torus1m_rays"][sel])
    assert_bits_equal(ids, GB["torus1m_ids"][sel], "torus1m ids")
    assert_bits_equal(t, GB["torus1m_t"][sel], "torus1m t")
    assert ((ids >= 0) & (ids < 1000000)).mean() > 0.5  # most of these rays end on the torus itself


def test_vn_planes_golden(oracle):
    """Triangle ctor + the `vn` override of scene.cpp:102-104 from Tor.obj's raw vn values: bit-exact planes."""
    vn = [v if h else None for v, h in zip(GB["torvn_vn"], GB["torvn_has"])]
    planes, squares = oracle.triangle_setup(GB["torvn_verts"], vn)
    assert GB["torvn_has"].all()
    assert_bits_equal(planes, GB["torvn_planes"], "vn planes")
    assert_bits_equal(squares, GB["torvn_squares"], "squares")
    assert_bits_equal(GB["torvn_planes"], G["tor_planes"])


CONVERGED_RMSE_TOL = 1e-3  # north_star: converged images at 4096 spp agree within RMSE 1e-3


def converged_image(name):
    """Σc/n of the unmodified reference at 4096 spp (tests/golden/converged.npz, made by make_converged.py)."""
    CV = np.load(GOLDEN + "/converged.npz")
    sm, cnt = CV[name + "_sum"], CV[name + "_count"]
    img = np.where(cnt[..., None] > 0, sm / np.maximum(cnt, 1)[..., None].astype(np.float32), np.float32(0))
    return CV, img.astype(np.float32)


def test_converged_4096spp_cornell(oracle):
    """The restatement against the reference's converged Cornell image (the Tor.obj image is left to the GPU test:
    brute force over 270 triangles takes the host cores most of a minute)."""
    CV, img = converged_image("cornell")
    W, H, spp, depth, seed = map(int, CV["cornell_cfg"])
    sc = with_planes(scenes.cornell(), oracle)
    sm, _, cnt, segs, _ = oracle.render_accum(sc, oracle.config(W, H, spp, depth, 1e-4, -1.0, seed))
    rgb = oracle.resolve(sm, cnt)
    assert np.sqrt(np.mean((rgb.astype(np.float64) - img) ** 2)) <= CONVERGED_RMSE_TOL
    assert_bits_equal(sm, CV["cornell_sum"])  # in fact bit for bit
    assert_bits_equal(cnt, CV["cornell_count"])
    assert segs == int(CV["cornell_segments"])


def sky_scene(oracle):
    sc = with_planes(scenes.open_scene(), oracle)
    sc["sky"] = G["sky_bgr"]
    return sc


def test_render_skybox(oracle):
    """miss -> equirect lookup with the reference's inverted lerp weights, accumulated without the throughput"""
    cnt = render_case(oracle, sky_scene(oracle), "sky")
    assert cnt.sum() > 32 * 32 * 4  # most paths leave the open room


def test_adaptive_rule_changes_the_result():
    # the adaptive fixture must actually exercise main.cpp:118-125 (otherwise it pins nothing)
    assert int(G["tor_adapt_segments"]) < 32 * 32 * 24 * 8
    flat_rate = int(G["tor_flat_segments"]) / (32 * 32 * 16)
    assert int(G["tor_adapt_segments"]) / (32 * 32 * 24) < flat_rate


def test_oracle_vs_live_reference(oracle, reference):
    """Fresh seeds and sizes, not in the fixtures: restatement == reference, bit for bit."""
    reference.set_config(24, 20, 12, 8, 1e-4, 0.001)
    assert reference.load_obj(MODELS, "Tor.obj") == 270
    planes, verts, squares = reference.dump()
    sc = tor_scene()
    assert_bits_equal(planes, sc["planes"])
    rays = scenes.random_rays(3000, seed=77)
    ri, rt = reference.intersect(rays)
    oi, ot = oracle.intersect(sc, rays)
    assert_bits_equal(oi, ri)
    assert_bits_equal(ot, rt)
    rs, rs2, rc, rseg, _ = reference.render_accum(seed=7)
    os_, os2, oc, oseg, _ = oracle.render_accum(sc, oracle.config(24, 20, 12, 8, 1e-4, 0.001, 7))
    assert_bits_equal(oc, rc)
    assert_bits_equal(os_, rs)
    assert_bits_equal(os2, rs2)
    assert oseg == rseg


def test_extensions_default_to_the_reference(oracle):
    """rr_depth = 0 and an all-zero materials_ext must not change a single bit (C5's extensions are opt-in)."""
    sc = with_planes(scenes.glossy_mix(), oracle)
    base = oracle.render_accum(sc, oracle.config(24, 16, 8, 12, 1e-4, 0.001, 7))
    sc2 = dict(sc, mats_ext=np.zeros((sc["mats"].shape[0], 2), np.float32))
    got = oracle.render_accum_w(sc2, oracle.config(24, 16, 8, 12, 1e-4, 0.001, 7, rr_depth=0))
    for a, b in zip(base[:3], got[:3]):
        assert_bits_equal(a, b)
    assert np.array_equal(got[3], got[2].astype(np.float32))    # weight == count without roulette
    assert base[3] == got[4]


def test_russian_roulette_keeps_the_expectation(oracle):
    """Extension (no reference semantics): the compensated throughput keeps E[sum of contributions per path]; the
    weight sum is an unbiased (if heavy-tailed) estimate of the no-roulette contributing-path count."""
    sc = with_planes(scenes.cornell(), oracle)
    n_paths = 32 * 18 * 256
    sm0, _, cnt0, wt0, segs0 = oracle.render_accum_w(sc, oracle.config(32, 18, 256, 16, 1e-4, -1.0, 42))
    sm1, _, cnt1, wt1, segs1 = oracle.render_accum_w(sc, oracle.config(32, 18, 256, 16, 1e-4, -1.0, 42, rr_depth=6))
    assert segs1 < 0.5 * segs0                                   # what roulette is for
    assert abs(sm1.sum() / sm0.sum() - 1) < 0.02, (sm1.sum() / n_paths, sm0.sum() / n_paths)
    # weights reach 20^k (p >= 0.05 per bounce): the sum converges slowly and sits below its mean at this sample count
    assert 0.5 < wt1.sum() / wt0.sum() < 1.5
    assert cnt1.sum() < cnt0.sum() and np.all(wt1 >= cnt1 - 1e-3)


def test_refraction_lobe_runs_and_is_selected(oracle):
    sc = with_planes(scenes.glass_mix(), oracle)
    a = oracle.render_accum_w(sc, oracle.config(32, 18, 16, 16, 1e-4, -1.0, 42))
    b = oracle.render_accum_w(dict(sc, mats_ext=None), oracle.config(32, 18, 16, 16, 1e-4, -1.0, 42))
    assert np.isfinite(a[0]).all() and a[2].sum() > 0
    assert not np.array_equal(a[0], b[0])                       # the glass boxes changed the image


def test_refract_matches_the_references_dead_code(oracle, reference):
    """The refractive lobe's direction arithmetic against Refract() of the reference's own main.cpp (:35-47, never
    called upstream), both branches, bit for bit."""
    rng = np.random.default_rng(5)
    n_tir = 0
    for _ in range(4000):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        if np.dot(n, d) > 0:
            n = -n
        n, d = n.astype(np.float32), d.astype(np.float32)
        eta = float(np.float32(rng.choice([1 / 1.5, 1.5, 1.33, 1 / 1.33, 2.4])))
        got, tir = oracle.refract(d, n, eta)
        n_tir += tir
        assert_bits_equal(got, reference.refract(d, n, eta))
    assert 0 < n_tir < 4000


def test_powf_restatement_matches_host_libm(oracle):
    """glibc powf restated (oracle/pt_oracle.c or_powf_restated = what csrc/pt_post.cuh ref_powf implements) against
    the host's powf: every k * 2^-24 in [0, 1) plus random bit patterns over the whole positive range."""
    rng = np.random.default_rng(3)
    unit = np.arange(1 << 24, dtype=np.float32) * np.float32(2.0 ** -24)
    wide = rng.integers(0, 0x7f800000, 4_000_000, dtype=np.uint32).view(np.float32)
    edge = np.array([0.0, 1.0, np.inf, 1e-45, 1.17e-38, 3.4e38, 255.0, 1.0000001, 0.99999994], np.float32)
    for gamma in (1 / 2.2, 0.5, 1.0, 2.2):
        for x in (unit, wide, edge):
            bad, first = oracle.powf_check(x, gamma)
            assert bad == 0, (gamma, first)


def test_post_goldens_are_the_references(oracle, reference):
    """The filter goldens come from the reference's own GaussBlur / MedianFilter (main.cpp:11-33, 49-80)."""
    img = G["post_img"]
    assert_bits_equal(reference.gauss_blur(img, 2), G["post_img_gauss2"])
    assert_bits_equal(reference.median_filter(img, 1), G["post_img_median1"])
    color, rgb8, st = oracle.resolve_display(G["cornell_sum"], G["cornell_sum2"], G["cornell_count"], 1 / 2.2)
    assert np.array_equal(rgb8, G["post_cornell_rgb8"])
    assert np.array_equal(oracle.to_rgb8(reference.gauss_blur(color, 2), G["cornell_count"]), G["post_cornell_gauss2"])
    assert np.array_equal(oracle.to_rgb8(reference.median_filter(color, 3), G["cornell_count"]), G["post_cornell_median3"])
