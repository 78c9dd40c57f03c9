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
    assert C.sizeof(ptb200.PtConfig) == 4 * 20
    assert C.sizeof(ptb200.PtScene) == 80
    assert C.sizeof(ptb200.PtStats) == 4 * 8 + 3 * 4 + 4 * 4 + 8 * 4 + 4 + 17 * 8 + 8 + 8  # 8-byte aligned


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


def _write_awkward_obj(d, n_quads=4000, seed=2):
    """An OBJ that exercises the tokeniser: CRLF and tabs, scientific notation, explicit signs, vt / vn records, the
    f forms a, a/b, a//c and a/b/c, several usemtl switches, comments-as-unknown-tokens; plus an MTL with the lines
    the reference parses, the extension lines (Ni, d) and lines it ignores."""
    rng = np.random.default_rng(seed)
    with open(os.path.join(d, "awk.mtl"), "w") as f:
        for k in range(3):
            f.write(f"newmtl m{k}\nNs {96.078431 * k}\nKa 1 1 1\nKd {0.25 * (k + 1)} 0.5 1e-2\nKs .5 .5 .5\n"
                    f"Ke {2 if k == 0 else 0} 0 0\nNi {1.0 + 0.25 * k}\nd {1.0 - 0.4 * k}\nillum 2\n\n")
    verts = rng.uniform(-9, 9, (n_quads * 4, 3)).astype(np.float32)
    fmt = ["{:.9g}", "{:+.8e}", "{:.7f}", "{:E}"]
    lines = ["mtllib awk.mtl", "o thing"]
    for i, v in enumerate(verts):
        txt = [fmt[(i + k) % 4].format(float(c)) for k, c in enumerate(v)]
        lines.append("v\t" + "  ".join(txt))
        verts[i] = [np.float32(float(t)) for t in txt]  # what the text says (the short formats drop digits)
    lines += ["vt 0.5 0.25", "vn 0 1 0", "vn 0.6 0 0.8", "s off"]
    tris, mats, has_vn = [], [], []
    forms = ["{a}", "{a}/1", "{a}//2", "{a}/1/1"]
    for q in range(n_quads):
        if q % 1000 == 0:
            lines.append(f"usemtl {(q // 1000) % 3}")
        a, b, c, e = 4 * q + 1, 4 * q + 2, 4 * q + 3, 4 * q + 4
        for t in ((a, b, c), (a, c, e)):
            form = forms[(q + t[1]) % 4]
            lines.append("f " + " ".join(form.format(a=i) for i in t))
            tris.append([verts[i - 1] for i in t])
            mats.append((q // 1000) % 3)
            has_vn.append({"{a}": 0, "{a}/1": 0, "{a}//2": 2, "{a}/1/1": 1}[form])
    with open(os.path.join(d, "awk.obj"), "w", newline="") as f:
        f.write("\r\n".join(lines) + "\r\n")
    return np.asarray(tris, np.float32).reshape(-1, 9), np.asarray(mats, np.int32), np.asarray(has_vn)


def test_fast_tokeniser_keeps_the_references_loader_semantics(loader, tmp_path):
    """host/scene.cpp reads the OBJ from one buffer instead of `file >> token`; it must build what
    the reference's Scene::LoadModel builds from the same file — compared live where oracle/_ref exists."""
    verts, mats, has_vn = _write_awkward_obj(str(tmp_path))
    sc = loader.load_obj(str(tmp_path), "awk.obj")
    assert_bits_equal(sc["verts"], verts, "vertices")
    assert np.array_equal(sc["mat_ids"], mats)
    assert sc["mats"].shape[0] == 3 and sc["mats"][2][9] == np.float32(96.078431 * 2)
    np.testing.assert_allclose(sc["mats_ext_raw"], [[1.0, 0.0], [1.25, 0.4], [1.5, 0.8]], rtol=1e-6)
    vn = {1: (0, 1, 0), 2: (0.6, 0, 0.8)}
    for i in np.flatnonzero(has_vn)[:50]:
        np.testing.assert_allclose(sc["planes"][i][:3], vn[has_vn[i]], rtol=1e-6)
    from oracle_py import Reference
    if Reference.available():
        ref = Reference()
        ref.set_config(8, 8, 1, 1)
        assert ref.load_obj(str(tmp_path), "awk.obj") == verts.shape[0]
        planes, rverts, squares = ref.dump()
        assert_bits_equal(sc["verts"], rverts, "vertices vs reference loader")
        assert_bits_equal(sc["planes"], planes, "planes vs reference loader")
        assert_bits_equal(sc["squares"], squares, "squares vs reference loader")


LOADER_KEYS = ("verts", "planes", "squares", "mat_ids", "mats", "mats_ext_raw")


def _load_with(loader, d, name, threads, min_bytes=1):
    os.environ["PT_LOADER_THREADS"], os.environ["PT_LOADER_MIN_BYTES"] = str(threads), str(min_bytes)
    try:
        return loader.load_obj(d, name)
    finally:
        del os.environ["PT_LOADER_THREADS"], os.environ["PT_LOADER_MIN_BYTES"]


def test_parallel_parse_is_the_sequential_parse(loader, tmp_path):
    """A large OBJ is cut after line breaks and parsed by several threads (host/scene.cpp LoadModelParallel); the
    result must be the sequential token machine's, bit for bit — here forced onto a small file, for several cuts."""
    d = str(tmp_path)
    _write_awkward_obj(d, n_quads=3000)
    seq = _load_with(loader, d, "awk.obj", 1)
    assert seq["parse_pieces"] == 0
    for threads in (2, 3, 7, 16):
        par = _load_with(loader, d, "awk.obj", threads)
        assert par["parse_pieces"] == threads
        for k in LOADER_KEYS:
            assert_bits_equal(np.asarray(par[k]), np.asarray(seq[k]), f"{k} with {threads} threads")
    assert loader.load_obj(d, "awk.obj")["parse_pieces"] == 0  # default: small files are not cut up


def test_parallel_parse_falls_back_when_the_cuts_do_not_fit(loader, tmp_path):
    """The loader's grammar is token based: statements may span lines and keywords inside "comments" count
    (scene.cpp:37-38).  Where a cut after a line break lands inside a statement, or an index points at a vertex the
    sequential loader would not have seen yet in another piece's range, the pieces are discarded and the sequential
    loop runs."""
    d = str(tmp_path)
    rng = np.random.default_rng(4)
    v = rng.uniform(-5, 5, (600, 3)).astype(np.float32)
    with open(os.path.join(d, "m.mtl"), "w") as f:
        f.write("newmtl a\nKd 1 0 0\nnewmtl b\nKd 0 1 0\n")
    # every vertex statement spans three lines: any cut after a line break is inside one with probability 2/3
    lines = ["mtllib m.mtl", "# v 9 9 9 is a vertex too: the reference reads tokens, not lines"]
    for x, y, z in v:
        lines += [f"v {x:.9g}", f"{y:.9g}", f"{z:.9g}"]
    lines += ["usemtl 1"] + [f"f {3 * k + 2} {3 * k + 3}\n{3 * k + 4}" for k in range(199)]
    open(os.path.join(d, "span.obj"), "w").write("\n".join(lines) + "\n")
    seq = _load_with(loader, d, "span.obj", 1)
    assert seq["verts"].shape[0] == 199 and seq["verts"][0][0] == v[0][0]  # vertex 1 is the comment's (9, 9, 9)
    hit_fallback = False
    for threads in (2, 3, 5, 8, 13):
        par = _load_with(loader, d, "span.obj", threads)
        hit_fallback |= par["parse_pieces"] == 0
        for k in LOADER_KEYS:
            assert_bits_equal(np.asarray(par[k]), np.asarray(seq[k]), f"{k} with {threads} threads")
    assert hit_fallback
    # one statement per line, but the faces come BEFORE their vertices in file order only for the last one: the
    # sequential loader throws there (positions.at), so this file keeps every face behind its vertices
    lines = [f"v {x:.9g} {y:.9g} {z:.9g}" for x, y, z in v] + [f"f {k + 1} {k + 2} {k + 3}" for k in range(0, 597, 3)]
    open(os.path.join(d, "plain.obj"), "w").write("\n".join(lines) + "\n")
    seq = _load_with(loader, d, "plain.obj", 1)
    par = _load_with(loader, d, "plain.obj", 6)
    assert par["parse_pieces"] == 6 and par["mats"].shape[0] == 0
    for k in ("verts", "planes", "squares", "mat_ids"):
        assert_bits_equal(np.asarray(par[k]), np.asarray(seq[k]), k)


def test_fast_float_conversion_is_strtof(loader):
    """host/scene.cpp fast_strtof converts short decimals through one double operation and must return strtof's float
    bit for bit, ending at the same character; everything it cannot guarantee goes to strtof (hex, inf / nan, more
    than 15 digits, large exponents, doubles that sit on a float midpoint)."""
    rng = np.random.default_rng(11)
    toks = ["0", "-0", "+0.0", ".5", "5.", "-.5e-3", "1e", "1e+", "1.e5", "12abc", "1.5e3x", "0x10", "0x1p3", "inf", "-inf",
            "nan", "+nan", "infinity", ".", "-", "+.e5", "1e5000", "1e-5000", "123456789012345", "1234567890123456",
            "0.000000000000000000001", "1e22", "1e23", "1e-22", "1e-23", "3.4028235e38", "1.17549435e-38", "1e-45",
            "00000000000000000000001.5", "9007199254740993", "16777217", "16777217.0", "33554434.0000001",
            "0.1", "0.2", "0.30000001", "1.00000005960464477539", "1.000000059604645", "1.00000005960464"]
    x = np.concatenate([rng.uniform(-30, 30, 200000), rng.normal(size=100000) * 1e-3, rng.uniform(-1, 1, 100000) * 1e6,
                        10.0 ** rng.uniform(-30, 30, 100000)])
    fmts = ["%.9g", "%.6f", "%.8e", "%.7g", "%+.12g", "%.15g", "%.17g", "%.3f", "%E", "%.10f"]
    for i, v in enumerate(x):
        toks.append(fmts[i % len(fmts)] % v)
    # decimals at and next to float midpoints: the double-rounding hazard the guard is there for
    f = np.abs(rng.normal(size=50000).astype(np.float32) * np.float32(10.0)) + np.float32(1e-3)
    mid = (f.astype(np.float64) + np.nextafter(f, np.float32(np.inf)).astype(np.float64)) / 2
    for m in mid:
        toks += ["%.15g" % m, "%.14g" % m, "%.13g" % m, repr(float(m)), "%.15g" % np.nextafter(m, np.inf), "%.15g" % np.nextafter(m, -np.inf)]
    n, fast, bad = loader.check_fast_strtof(" ".join(toks))
    assert n == len(toks)
    assert bad == 0
    assert fast > 0.6 * n  # the fast path is what usually runs


def test_binary_scene_cache(loader, tmp_path):
    """-CACHE <dir>: the second load takes the binary image and yields the same arrays bit for
    bit; a changed OBJ or MTL, a truncated or foreign cache file fall back to parsing."""
    d, cache = str(tmp_path), str(tmp_path / "cache")
    os.mkdir(cache)
    _write_awkward_obj(d, n_quads=500)
    plain = loader.load_obj(d, "awk.obj")
    first = loader.load_obj(d, "awk.obj", cache_dir=cache)
    assert not plain["from_cache"] and not first["from_cache"]
    img = os.path.join(cache, "awk.obj.ptscene")
    assert os.path.exists(img) and not os.path.exists(img + ".tmp")
    again = loader.load_obj(d, "awk.obj", cache_dir=cache)
    assert again["from_cache"]
    for k in ("verts", "planes", "squares", "mat_ids", "mats", "mats_ext_raw"):
        assert_bits_equal(np.asarray(again[k]), np.asarray(plain[k]), k)
    assert (again["mats_ext"] is None) == (plain["mats_ext"] is None)
    # truncated image: ignored, rewritten
    blob = open(img, "rb").read()
    open(img, "wb").write(blob[: len(blob) // 2])
    assert not loader.load_obj(d, "awk.obj", cache_dir=cache)["from_cache"]
    assert open(img, "rb").read() == blob
    # trailing garbage / foreign file
    open(img, "wb").write(blob + b"x")
    assert not loader.load_obj(d, "awk.obj", cache_dir=cache)["from_cache"]
    open(img, "wb").write(b"not a cache")
    assert not loader.load_obj(d, "awk.obj", cache_dir=cache)["from_cache"]
    # right magic and stamps, absurd triangle count (header: magic[8], then int64 n_triangles): the payload size is
    # checked against the file before anything is allocated from the header
    open(img, "wb").write(blob[:8] + (2**31 - 1).to_bytes(8, "little") + blob[16:])
    assert not loader.load_obj(d, "awk.obj", cache_dir=cache)["from_cache"]
    assert open(img, "rb").read() == blob
    assert loader.load_obj(d, "awk.obj", cache_dir=cache)["from_cache"]
    # the MTL changes: stale
    with open(os.path.join(d, "awk.mtl"), "a") as f:
        f.write("\nnewmtl extra\nKd 0.1 0.2 0.3\n")
    changed = loader.load_obj(d, "awk.obj", cache_dir=cache)
    assert not changed["from_cache"] and changed["mats"].shape[0] == plain["mats"].shape[0] + 1
    assert loader.load_obj(d, "awk.obj", cache_dir=cache)["from_cache"]
    # the OBJ changes: stale
    with open(os.path.join(d, "awk.obj"), "a") as f:
        f.write("f 1 2 3\n")
    grown = loader.load_obj(d, "awk.obj", cache_dir=cache)
    assert not grown["from_cache"] and grown["verts"].shape[0] == plain["verts"].shape[0] + 1
    # an unwritable cache directory is not an error
    assert not loader.load_obj(d, "awk.obj", cache_dir="/nonexistent/dir")["from_cache"]


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


def test_host_ray_reflect(loader):
    """host/ray.h Ray::Reflect = ray.h:45-50: new origin, glm::normalize(vec4) of the direction ((x+y)+(z+w) dot,
    v * (1 / sqrt)), throughput times the tint, depth + 1."""
    f = np.float32
    rng = np.random.default_rng(8)
    for _ in range(200):
        nd = rng.normal(size=4).astype(np.float32)
        nd[3] = 0
        nb = np.append(rng.normal(size=3), 1).astype(np.float32)
        col, tint = rng.random(3).astype(np.float32), rng.random(3).astype(np.float32)
        b, d, c, depth = loader.ray_reflect([0, 0, -20, 1], [0, 0, 1, 0], col, 3, nb, nd, tint)
        inv = f(1) / np.sqrt(f(f(nd[0] * nd[0]) + f(nd[1] * nd[1])) + f(f(nd[2] * nd[2]) + f(nd[3] * nd[3])))
        assert_bits_equal(d, (nd * inv).astype(np.float32), "direction")
        assert_bits_equal(b, nb, "begin")
        assert_bits_equal(c, col * tint, "colour")
        assert depth == 4


def test_host_triangle_constructor_bulk(loader):
    GB = np.load(GOLDEN + "/golden_big.npz")
    p, s = loader.triangle_setup_bulk(GB["torvn_verts"], GB["torvn_vn"], GB["torvn_has"])
    assert_bits_equal(p, GB["torvn_planes"], "vn planes")
    assert_bits_equal(s, GB["torvn_squares"], "squares")
