"""Synthetic scenes of BASELINE.json's configs, as the arrays the C-ABI takes.

All scenes live in the frame the reference's fixed camera looks at (main.cpp:126-129: eye (0,0,-20), looking +Z,
tan(half-fov) = 0.5 on both axes): a closed room x,y in [-10,10], z in [-21,10], like the `Main` object of
models/Tor.obj (Tor.obj:456-508).  Triangles are wound so that the constructor's geometric normal
(triangles.h:31-34) points where the reference needs it: INTO the room for walls, OUT of objects, DOWN for the
ceiling light — the reference never flips a normal toward the ray (scene.cpp:123, material.h:84,99).
Material rows are (Ke[3], Kd[3], Ks[3], Ns) = MaterialCharacteristics (material.h:22-28).
"""
import numpy as np

WHITE, RED, GREEN = (1, 1, 1), (1, .01, .01), (.01, 1, .01)


def mat(Kd, Ke=(0, 0, 0), Ks=(0, 0, 0), Ns=0.0):
    return [*Ke, *Kd, *Ks, Ns]


def _quad(a, b, c, d, want):
    """Two triangles (a,b,c), (a,c,d) of the planar quad a-b-c-d, wound so that cross(AB, AC) points along `want`."""
    a, b, c, d = (np.asarray(p, np.float64) for p in (a, b, c, d))
    if np.dot(np.cross(b - a, c - a), np.asarray(want, np.float64)) < 0:
        b, d = d, b
    return [np.concatenate([a, b, c]), np.concatenate([a, c, d])]


def _box(lo, hi, outward=True, skip_bottom=False):
    (x0, y0, z0), (x1, y1, z1) = lo, hi
    s = 1.0 if outward else -1.0
    faces = [
        ([x0, y0, z0], [x0, y1, z0], [x0, y1, z1], [x0, y0, z1], (-s, 0, 0)),
        ([x1, y0, z0], [x1, y1, z0], [x1, y1, z1], [x1, y0, z1], (s, 0, 0)),
        ([x0, y1, z0], [x1, y1, z0], [x1, y1, z1], [x0, y1, z1], (0, s, 0)),
        ([x0, y0, z0], [x1, y0, z0], [x1, y1, z0], [x0, y1, z0], (0, 0, -s)),
        ([x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1], (0, 0, s)),
    ]
    if not skip_bottom:
        faces.insert(2, ([x0, y0, z0], [x1, y0, z0], [x1, y0, z1], [x0, y0, z1], (0, -s, 0)))
    tris = []
    for a, b, c, d, n in faces:
        tris += _quad(a, b, c, d, n)
    return tris


def _rotate_y(tris, centre, deg):
    t = np.asarray(tris, np.float64).reshape(-1, 3, 3)
    c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
    R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    ctr = np.asarray(centre, np.float64)
    return list(((t - ctr) @ R.T + ctr).reshape(-1, 9))


def _scene(tris, mat_ids, mats):
    return {"verts": np.asarray(tris, np.float64).astype(np.float32).reshape(-1, 9),
            "planes": None, "squares": None,
            "mat_ids": np.asarray(mat_ids, np.int32), "mats": np.asarray(mats, np.float32).reshape(-1, 10)}


def room(light_half=3.0, light_y=9.99, wall_Ns=0.0, wall_Ks=(0, 0, 0)):
    """12 wall triangles (inward normals; left red, right green, rest white) + a 2-triangle ceiling light."""
    mats = [mat(WHITE, Ks=wall_Ks, Ns=wall_Ns), mat(RED, Ks=wall_Ks, Ns=wall_Ns), mat(GREEN, Ks=wall_Ks, Ns=wall_Ns),
            mat(WHITE, Ke=(1, 1, 1))]
    tris = _box((-10, -10, -21), (10, 10, 10), outward=False)  # faces: -x, +x, -y, +y, -z, +z
    ids = [1, 1, 2, 2, 0, 0, 0, 0, 0, 0, 0, 0]                 # left (x=-10) red, right (x=+10) green
    h = light_half
    tris += _quad([-h, light_y, -h], [h, light_y, -h], [h, light_y, h], [-h, light_y, h], (0, -1, 0))
    ids += [3, 3]
    return tris, ids, mats


def cornell():
    """C2: ~34 triangles, diffuse + emissive."""
    tris, ids, mats = room()
    short = _rotate_y(_box((1.5, -10, -4), (7.5, -4, 2), skip_bottom=True), (4.5, -7, -1), -18)
    tall = _rotate_y(_box((-7.5, -10, 1), (-1.5, 2, 7), skip_bottom=True), (-4.5, -4, 4), 20)
    tris += short + tall
    ids += [0] * 20
    return _scene(tris, ids, mats)


def torus_tris(nu, nv, centre=(-4, 0, 4), axis=(-0.35355, -0.70711, 0.61237), R=3.0421, r=0.8663):
    """Parametric torus with Tor.obj's measured placement, nu x nv quads -> 2*nu*nv triangles, outward normals."""
    ax = np.asarray(axis, np.float64)
    ax /= np.linalg.norm(ax)
    e1 = np.cross(ax, [1.0, 0.0, 0.0])
    e1 /= np.linalg.norm(e1)
    e2 = np.cross(ax, e1)
    u = np.arange(nu + 1) * (2 * np.pi / nu)
    v = np.arange(nv + 1) * (2 * np.pi / nv)
    U, V = np.meshgrid(u, v, indexing="ij")
    ring = np.cos(U)[..., None] * e1 + np.sin(U)[..., None] * e2
    P = np.asarray(centre, np.float64) + (R + r * np.cos(V))[..., None] * ring + (r * np.sin(V))[..., None] * ax
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    # orientation: make cross(AB, AC) point away from the tube centre line
    tube_c = np.asarray(centre, np.float64) + R * ring[:-1, :-1]
    n = np.cross(b - a, c - a)
    flip = (np.einsum("ijk,ijk->ij", n, a - tube_c) < 0)[..., None]
    b2 = np.where(flip, d, b)
    d2 = np.where(flip, b, d)
    t1 = np.concatenate([a, b2, c], -1).reshape(-1, 9)
    t2 = np.concatenate([a, c, d2], -1).reshape(-1, 9)
    return np.stack([t1, t2], 1).reshape(-1, 9)


def torus_room(nu=1000, nv=500):
    """C3: subdivided torus (2*nu*nv triangles, glossy+diffuse material 4 of Tor.mtl:44-52) in the lit room."""
    tris, ids, mats = room(light_half=1.0, light_y=9.0)
    mats = mats + [mat((0.5, 1, 0.01), Ks=(0.5, 0.5, 0.5), Ns=96.078431)]
    tt = torus_tris(nu, nv)
    verts = np.concatenate([tt, np.asarray(tris, np.float64)], 0)
    mat_ids = np.concatenate([np.full(len(tt), 4, np.int32), np.asarray(ids, np.int32)])
    return _scene(verts, mat_ids, mats)


def composite(tor_scene):
    """C4: Tor.obj (as loaded by the host loader, planes included) + the two Cornell boxes."""
    c = cornell()
    boxes = c["verts"][14:]
    n0 = tor_scene["verts"].shape[0]
    white = tor_scene["mats"].shape[0]
    mats = np.concatenate([tor_scene["mats"], np.asarray([mat(WHITE)], np.float32)], 0)
    verts = np.concatenate([tor_scene["verts"], boxes], 0)
    ids = np.concatenate([tor_scene["mat_ids"], np.full(len(boxes), white, np.int32)])
    return {"verts": verts, "planes": None, "squares": None, "mat_ids": ids, "mats": mats, "n_loaded": n0}


def glossy_mix():
    """C5: Cornell geometry with per-object Ns in {0, 96.08, 500, 900, 1000} and Ks in {0.5, 0.9} (depth 32)."""
    tris, ids, mats = room(wall_Ns=96.078431, wall_Ks=(0.5, 0.5, 0.5))
    mats += [mat(WHITE, Ks=(0.9, 0.9, 0.9), Ns=500.0), mat((0.8, 0.8, 1), Ks=(0.9, 0.9, 0.9), Ns=900.0),
             mat(WHITE, Ks=(0.9, 0.9, 0.9), Ns=1000.0)]
    short = _rotate_y(_box((1.5, -10, -4), (7.5, -4, 2), skip_bottom=True), (4.5, -7, -1), -18)
    tall = _rotate_y(_box((-7.5, -10, 1), (-1.5, 2, 7), skip_bottom=True), (-4.5, -4, 4), 20)
    tris += short + tall
    ids += [4] * 10 + [5] * 8 + [6] * 2
    return _scene(tris, ids, mats)


def glass_mix():
    """C5 with the refractive extension: glossy_mix() where the short box is glass (Ni 1.5, Tr 0.8, tint Ks) and the
    tall box a mirror/diffuse/refractive mix, so that all three lobes and every lobe-count case are exercised.
    `mats_ext` rows are (Ni, Tr) per material (PtScene.materials_ext)."""
    sc = glossy_mix()
    sc["mats"][4] = mat((0.9, 0.9, 0.9), Ks=(0.95, 0.97, 0.95), Ns=200.0)   # mirror + diffuse + refract
    sc["mats"][5] = mat((0.8, 0.8, 1), Ks=(0.9, 0.9, 1.0), Ns=1000.0)       # mirror + refract (no diffuse lobe)
    ext = np.zeros((sc["mats"].shape[0], 2), np.float32)
    ext[4] = (1.5, 0.8)
    ext[5] = (1.33, 0.5)
    ext[6] = (1.5, 1.0)                                                     # Ns 1000, Tr 1: chances (0, -, 1)
    sc["mats_ext"] = ext
    return sc


def open_scene():
    """Skybox test scene (scene.cpp:125-156): the Cornell room without ceiling, back wall and light, so that most
    paths leave it; plus the two boxes.  No emitter: all radiance comes from the sky lookup."""
    c = cornell()
    keep = [i for i in range(34) if i not in (6, 7, 10, 11, 12, 13)]  # drop +y wall, +z wall, light
    return {"verts": c["verts"][keep], "planes": None, "squares": None, "mat_ids": c["mat_ids"][keep], "mats": c["mats"]}


def sky_image(width=64, height=32, seed=7):
    """Synthetic equirect skybox: smooth gradients plus noise, uint8 (H, W, 3) in blue-green-red order."""
    rng = np.random.default_rng(seed)
    y, x = np.meshgrid(np.linspace(0, 1, height), np.linspace(0, 1, width), indexing="ij")
    img = np.stack([200 * (1 - y) + 30, 120 + 100 * np.sin(6.283 * x), 80 + 150 * y], -1)
    return np.clip(img + rng.integers(-20, 20, img.shape), 0, 255).astype(np.uint8)


def write_bmp(path, bgr):
    """24-bit bottom-up BMP of a (H, W, 3) blue-green-red image (for the reference's bitmap_image loader)."""
    h, w = bgr.shape[:2]
    stride = (w * 3 + 3) & ~3
    rows = np.zeros((h, stride), np.uint8)
    rows[:, :w * 3] = bgr[::-1].reshape(h, w * 3)
    hdr = bytearray(54)
    hdr[0:2] = b"BM"
    hdr[2:6] = (54 + stride * h).to_bytes(4, "little")
    hdr[10:14] = (54).to_bytes(4, "little")
    hdr[14:18] = (40).to_bytes(4, "little")
    hdr[18:22] = w.to_bytes(4, "little")
    hdr[22:26] = h.to_bytes(4, "little")
    hdr[26:28] = (1).to_bytes(2, "little")
    hdr[28:30] = (24).to_bytes(2, "little")
    hdr[34:38] = (stride * h).to_bytes(4, "little")
    open(path, "wb").write(bytes(hdr) + rows.tobytes())


def random_rays(n, seed=0, inside=True):
    """Seeded ray set: origins in the room, uniform directions (unnormalised: Ray's constructor normalises)."""
    rng = np.random.default_rng(seed)
    o = rng.uniform([-9.5, -9.5, -20.5], [9.5, 9.5, 9.5], (n, 3)) if inside else rng.uniform(-30, 30, (n, 3))
    d = rng.normal(size=(n, 3))
    return np.concatenate([o, d], 1).astype(np.float32)


def camera_rays(width, height):
    """Primary rays through pixel centres (no jitter), main.cpp:126-129 with the jitter at 0."""
    x, y = np.meshgrid(np.arange(width), np.arange(height))
    d = np.stack([(x / width - 0.5), (-(y) / height + 0.5), np.ones_like(x, np.float64)], -1).reshape(-1, 3)
    o = np.broadcast_to(np.array([0.0, 0.0, -20.0]), d.shape)
    return np.concatenate([o, d], 1).astype(np.float32)
