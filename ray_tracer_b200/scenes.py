"""Procedural scenes for the bench and the large-size tests.

BASELINE.json's C3/C4/C5 name assets that are not in the reference checkout (sponza.obj, dragon.obj:
/root/reference/.MISSING_LARGE_BLOBS) and there is no network, so the bench uses labelled STAND-INS
with the same structure: one RenderObject + BVH2 per material group, every group scattered through
the whole atrium (as Sponza's usemtl groups are), ~262k triangles in 25 groups, identity transform,
environment light on. A user-supplied sponza.obj goes through SceneBuilder.read_obj instead.

World conventions of the reference: -y is up, the hard-wired NEE light is the quad y = -1.5,
x,z in +-1/3 (shaders/raytrace.comp:368-387), the default camera sits at (0,-0.5,-3.5) looking +z.
"""
import numpy as np

from . import scene as _scene


GPU_BVH = None     # (device, min_triangles): objects at least that large get their BVH2 from b200pt_build_bvh (same bytes)


def _builder(make_builder=None):
    """make_builder: another builder class with the same methods (bench.py's reference arm passes the oracle's, so that it
    loads no library of the product)."""
    if make_builder is not None:
        return make_builder()
    b = _scene.SceneBuilder()
    if GPU_BVH is not None:
        b.use_gpu_bvh(*GPU_BVH)
    return b


def _pack(tri_pos, tri_nrm, uv_scale=None):
    """(n,3,3) positions + (n,3,3) normals -> (n,3,8) TrianglePoints. uv = 0, or with uv_scale a planar mapping: the two
    world axes across the face normal's dominant axis, times uv_scale (texture repeats per world unit; sampler 0 wraps)."""
    n = tri_pos.shape[0]
    out = np.zeros((n, 3, 8), np.float32)
    out[:, :, 0:3] = tri_pos
    out[:, :, 4:7] = tri_nrm
    if uv_scale is not None:
        fn = np.cross(tri_pos[:, 1] - tri_pos[:, 0], tri_pos[:, 2] - tri_pos[:, 0])
        dom = np.argmax(np.abs(fn), axis=1)
        ua = np.where(dom == 0, 2, 0)                 # x-facing: (z, y); y-facing: (x, z); z-facing: (x, y)
        va = np.where(dom == 1, 2, 1)
        idx = np.arange(n)
        out[:, :, 3] = tri_pos[idx, :, ua] * uv_scale
        out[:, :, 7] = tri_pos[idx, :, va] * uv_scale
    return out


def _grid(origin, du, dv, nu, nv, height=None):
    """Tessellated parallelogram: 2*nu*nv triangles, optional displacement height(u,v) along the normal."""
    origin, du, dv = (np.asarray(a, np.float64) for a in (origin, du, dv))
    u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
    nrm = np.cross(du, dv)
    nrm /= np.linalg.norm(nrm)
    p = origin + u[..., None] * du + v[..., None] * dv
    if height is not None:
        p = p + height(u, v)[..., None] * nrm
    a, b, c, d = p[:-1, :-1], p[1:, :-1], p[1:, 1:], p[:-1, 1:]
    tris = np.concatenate([np.stack([a, b, c], 2).reshape(-1, 3, 3), np.stack([a, c, d], 2).reshape(-1, 3, 3)])
    fn = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
    fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-30)
    return tris, np.repeat(fn[:, None, :], 3, 1)


def _revolve(center, profile_r, profile_y, segments):
    """Surface of revolution around the y axis through `center`: column shafts, vases."""
    k = len(profile_r) - 1
    ang = np.linspace(0, 2 * np.pi, segments + 1)
    r = np.asarray(profile_r, np.float64)[:, None]
    y = np.asarray(profile_y, np.float64)[:, None]
    p = np.stack([r * np.cos(ang), np.broadcast_to(y, (k + 1, segments + 1)), r * np.sin(ang)], -1) + np.asarray(center)
    a, b, c, d = p[:-1, :-1], p[1:, :-1], p[1:, 1:], p[:-1, 1:]
    tris = np.concatenate([np.stack([a, b, c], 2).reshape(-1, 3, 3), np.stack([a, c, d], 2).reshape(-1, 3, 3)])
    # smooth normals: radial direction corrected by the profile slope
    def nrm_of(q):
        rel = q - np.asarray(center)
        rad = rel * [1, 0, 1]
        ln = np.maximum(np.linalg.norm(rad, axis=-1, keepdims=True), 1e-30)
        return rad / ln
    return tris, nrm_of(tris)


def _blob(center, radius, rings, segments, bump, seed, outward=False):
    """Bumpy sphere (plants, lion heads stand-in). The default winding faces inwards (the atrium workload was
    measured that way and is kept); outward=True winds the triangles so that the outside is the front face."""
    rng = np.random.default_rng(seed)
    th = np.linspace(0, np.pi, rings + 1)[:, None]
    ph = np.linspace(0, 2 * np.pi, segments + 1)[None, :]
    f = 1 + bump * (np.sin(5 * th + rng.uniform(0, 6)) * np.cos(7 * ph + rng.uniform(0, 6)))
    d = np.stack([np.sin(th) * np.cos(ph), np.cos(th) * np.ones_like(ph), np.sin(th) * np.sin(ph)], -1)
    p = np.asarray(center) + radius * f[..., None] * d
    a, b, c, dd = p[:-1, :-1], p[1:, :-1], p[1:, 1:], p[:-1, 1:]
    tris = np.concatenate([np.stack([a, b, c], 2).reshape(-1, 3, 3), np.stack([a, c, dd], 2).reshape(-1, 3, 3)])
    area = np.linalg.norm(np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0]), axis=1)
    tris = tris[area > 1e-12]                              # pole triangles degenerate
    if outward:
        tris = tris[:, ::-1].copy()
    nrm = tris - np.asarray(center)
    nrm /= np.maximum(np.linalg.norm(nrm, axis=-1, keepdims=True), 1e-30)
    return tris, nrm


def _box(lo, hi, skip=()):
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    parts = []
    faces = {"x0": (lo, [0, 0, 1], [0, 1, 0]), "x1": ([hi[0], lo[1], lo[2]], [0, 1, 0], [0, 0, 1]),
             "y0": (lo, [1, 0, 0], [0, 0, 1]), "y1": ([lo[0], hi[1], lo[2]], [0, 0, 1], [1, 0, 0]),
             "z0": (lo, [0, 1, 0], [1, 0, 0]), "z1": ([lo[0], lo[1], hi[2]], [1, 0, 0], [0, 1, 0])}
    ext = hi - lo
    for name, (o, du, dv) in faces.items():
        if name in skip:
            continue
        parts.append(_grid(o, np.asarray(du) * ext, np.asarray(dv) * ext, 1, 1))
    return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])


def atrium_parts(detail=1.0):
    """List of (group name, tris (n,3,3), normals (n,3,3)) for the Sponza stand-in."""
    s = lambda n: max(2, int(round(n * detail)))
    parts = []
    add = lambda name, tn: parts.append((name, tn[0], tn[1]))
    X, Z0, Z1, FLOOR, ROOF = 3.0, -5.0, 9.0, 0.5, -5.0
    L = Z1 - Z0
    tile = lambda u, v: 0.004 * np.sin(u * 60 * np.pi) * np.sin(v * 140 * np.pi)
    add("floor", _grid([-X, FLOOR, Z0], [2 * X, 0, 0], [0, 0, L], s(56), s(130), tile))
    brick = lambda u, v: 0.006 * np.sign(np.sin(u * 40 * np.pi)) * np.sign(np.sin(v * 90 * np.pi))
    add("bricks_left", _grid([-X, FLOOR, Z0], [0, 0, L], [0, ROOF - FLOOR, 0], s(96), s(40), brick))
    add("bricks_right", _grid([X, FLOOR, Z1], [0, 0, -L], [0, ROOF - FLOOR, 0], s(96), s(40), brick))
    add("wall_front", _grid([X, FLOOR, Z0], [-2 * X, 0, 0], [0, ROOF - FLOOR, 0], s(30), s(40), brick))
    add("wall_back", _grid([-X, FLOOR, Z1], [2 * X, 0, 0], [0, ROOF - FLOOR, 0], s(30), s(40), brick))
    # roof: two side strips, the middle is open to the sky
    add("roof", _grid([-X, ROOF, Z0], [0, 0, L], [1.5, 0, 0], s(60), s(8)))
    add("roof", _grid([X, ROOF, Z1], [0, 0, -L], [-1.5, 0, 0], s(60), s(8)))
    # gallery walkway at mid height on both sides + its ceiling
    for side in (-1, 1):
        x0 = side * X
        add("gallery", _grid([x0, -2.0, Z0 if side < 0 else Z1], [0, 0, side * -L], [-side * 1.2, 0, 0], s(90), s(8)))
        add("gallery", _grid([x0 - side * 1.2, -2.0, Z0 if side > 0 else Z1], [0, 0, side * L], [0, 0.15, 0], s(90), 1))
    # columns: two levels, both sides
    prof_y = np.concatenate([[0, 0.04, 0.08, 0.12], np.linspace(0.16, 2.1, s(16)), [2.16, 2.22, 2.3, 2.35]])
    prof_r = np.concatenate([[0.17, 0.17, 0.14, 0.12], 0.1 + 0.008 * np.cos(np.linspace(0, 3, s(16))), [0.12, 0.15, 0.17, 0.17]])
    zs = np.arange(Z0 + 1.0, Z1 - 0.5, 1.5)
    for level, base in enumerate((FLOOR, -2.0)):
        for side in (-1, 1):
            for i, z in enumerate(zs):
                name = "column_%s_%d" % ("a" if (i + level) % 2 else "b", level)
                add(name, _revolve([side * 1.8, base, z], prof_r, -prof_y, s(56)))
        # lintels on top of the columns
        for side in (-1, 1):
            add("lintel_%d" % level, _box([side * 1.8 - 0.18, base - 2.5, Z0 + 0.5], [side * 1.8 + 0.18, base - 2.35, Z1 - 0.2]))
    # curtains hanging between upper columns
    for k, z in enumerate(zs[:-1]):
        for side in (-1, 1):
            ph = 0.7 * k + side
            wave = lambda u, v, ph=ph: 0.06 * np.sin(u * 9 * np.pi + ph) * (0.3 + v) + 0.02 * np.sin(v * 5 * np.pi + 2 * ph)
            add("curtain_%s" % ("red", "green", "blue")[k % 3],
                _grid([side * 1.8, -4.3, z + 0.15], [0, 0, 1.2], [0, 1.9, 0], s(40), s(56), wave))
    # vases and plants along the nave, reliefs on the end walls
    vr = [0.02, 0.1, 0.16, 0.2, 0.21, 0.19, 0.15, 0.1, 0.08, 0.1, 0.13]
    vy = [0.0, 0.0, 0.05, 0.14, 0.25, 0.36, 0.45, 0.52, 0.58, 0.63, 0.66]
    vy_f = np.interp(np.linspace(0, 1, s(30)), np.linspace(0, 1, len(vy)), vy)
    vr_f = np.interp(np.linspace(0, 1, s(30)), np.linspace(0, 1, len(vr)), vr)
    for k, z in enumerate(zs[::2]):
        for side in (-1, 1):
            add("vase", _revolve([side * 1.1, FLOOR, z + 0.75], vr_f, -vy_f, s(64)))
            add("plant", _blob([side * 1.1, FLOOR - 0.95, z + 0.75], 0.3, s(44), s(64), 0.25, seed=100 + 17 * k + side))
    for z, sgn in ((Z0 + 0.12, 1), (Z1 - 0.12, -1)):
        add("lion", _blob([0.0, -1.2, z], 0.45, s(72), s(100), 0.18, seed=50 + sgn))
    return parts


# 25 diffuse materials (all MTL materials come out diffuse in the reference: Ni/d are ignored, A-Q18)
ATRIUM_GROUPS = ["floor", "bricks_left", "bricks_right", "wall_front", "wall_back", "roof", "gallery", "column_a_0", "column_b_0",
                 "column_a_1", "column_b_1", "lintel_0", "lintel_1", "curtain_red", "curtain_green", "curtain_blue", "vase",
                 "plant", "lion", "trim_a", "trim_b", "trim_c", "trim_d", "trim_e", "trim_f"]


def procedural_texture(kind, seed, size=1024):
    """size x size x 4 uint8, sRGB-encoded, in the role of one file of assets/sponza_textures (absent geometry, so the
    maps are stand-ins too): brick / tile / fabric / stone patterns with per-texel noise."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:size, 0:size].astype(np.float32) / size
    base = rng.uniform(0.35, 0.95, 3).astype(np.float32)
    noise = rng.uniform(0.85, 1.0, (size, size, 1)).astype(np.float32)
    if kind == "brick":
        row = np.floor(y * 16)
        mortar = (np.abs((y * 16) % 1 - 0.5) > 0.44) | (np.abs(((x * 8 + 0.5 * (row % 2)) % 1) - 0.5) > 0.47)
        shade = np.where(mortar, 0.55, 1.0)[..., None] * (0.8 + 0.2 * np.sin(row * 12.9898)[..., None] ** 2)
    elif kind == "tile":
        shade = np.where(((np.floor(x * 8) + np.floor(y * 8)) % 2) == 0, 1.0, 0.6)[..., None]
    elif kind == "fabric":
        shade = (0.75 + 0.25 * np.sin(x * 2 * np.pi * 48) * np.sin(y * 2 * np.pi * 6))[..., None]
    else:
        shade = (0.7 + 0.3 * np.sin(x * 37.0 + 5 * np.sin(y * 11.0)) * np.cos(y * 29.0))[..., None]
    rgb = np.clip(base * shade * noise, 0.0, 1.0)
    out = np.empty((size, size, 4), np.uint8)
    out[..., :3] = np.round(rgb * 255.0)
    out[..., 3] = 255
    return out


def metal_mask_texture(seed, size=512):
    """metalness map (map_Ks): red channel >= 0.5 marks mirror texels — inlaid discs on an otherwise diffuse surface."""
    y, x = np.mgrid[0:size, 0:size].astype(np.float32) / size
    disc = ((x * 4 % 1 - 0.5) ** 2 + (y * 4 % 1 - 0.5) ** 2) < 0.04
    out = np.zeros((size, size, 4), np.uint8)
    out[..., 0] = np.where(disc, 255, 0)
    out[..., 3] = 255
    return out


ATRIUM_TEXTURE_KIND = {"floor": "tile", "bricks_left": "brick", "bricks_right": "brick", "wall_front": "brick", "wall_back": "brick",
                       "curtain_red": "fabric", "curtain_green": "fabric", "curtain_blue": "fabric"}


def atrium(detail=0.89, with_light=True, seed=1, make_builder=None, textured=False, extra=None):
    """Sponza stand-in: returns (blobs, info). Materials 0..24 diffuse, 25 = emissive light box at the NEE quad.
    textured: every material group gets an albedo map (TextureBuffer slot = its material index; `vase` and `lintel_*` a
    metalness map as well) and planar texture coordinates — the role of assets/sponza.mtl + sponza_textures; the images
    are in info["texture_images"] {slot: (h, w, 4) uint8} for b200pt_upload_texture. extra: list of
    (tris, normals, material64) appended as further objects (the C5 statue)."""
    rng = np.random.default_rng(seed)
    groups = {g: [] for g in ATRIUM_GROUPS}
    for name, tris, nrm in atrium_parts(detail):
        groups[name].append((tris, nrm))
    # trim_*: small boxes scattered on walls/floor so that every one of the 25 groups has geometry everywhere
    for i, g in enumerate(ATRIUM_GROUPS[19:]):
        for k in range(40):
            c = np.array([rng.uniform(-2.8, 2.8), rng.uniform(-4.5, 0.3), rng.uniform(-4.8, 8.8)])
            h = rng.uniform(0.03, 0.12, 3)
            groups[g].append(_box(c - h, c + h))
    b = _builder(make_builder)
    b.set_spheres(np.zeros(320, np.uint8))      # spheres.resize(10): ten zero spheres, always (src/vk_engine.cpp:682)
    albedo = rng.uniform(0.25, 0.9, (25, 3))
    albedo[13], albedo[14], albedo[15] = (0.8, 0.15, 0.12), (0.15, 0.7, 0.2), (0.15, 0.2, 0.8)
    images = {}
    for gi, a in enumerate(albedo):
        if textured:
            g = ATRIUM_GROUPS[gi]
            images[gi] = procedural_texture(ATRIUM_TEXTURE_KIND.get(g, "stone"), 1000 + gi)
            metal = -1
            if g in ("vase", "lintel_0", "lintel_1"):
                metal = 32 + len([k for k in images if k >= 32])
                images[metal] = metal_mask_texture(gi)
            b.add_material(_scene.material(albedo=a, albedo_index=gi, metalness_index=metal))
        else:
            b.add_material(_scene.material(albedo=a))
    b.add_material(_scene.material(albedo=(0, 0, 0), emission=(1, 1, 1), strength=2.4))     # 25: the reference's light material
    counts = {}
    for gi, g in enumerate(ATRIUM_GROUPS):
        tris = np.concatenate([t for t, _ in groups[g]])
        nrm = np.concatenate([n for _, n in groups[g]])
        counts[g] = len(tris)
        b.add_mesh(_pack(tris, nrm, uv_scale=0.5 if textured else None), material=gi)
    if with_light:
        t, n = _box([-1 / 3, -1.52, -1 / 3], [1 / 3, -1.5, 1 / 3])
        b.add_mesh(_pack(t, n), material=25)
        counts["light"] = len(t)
    for k, (t, n, mat64) in enumerate(extra or []):
        b.add_material(mat64)
        b.add_mesh(_pack(t, n), material=26 + k)
        counts["extra_%d" % k] = len(t)
    blobs = b.blobs()
    info = dict(triangles=int(sum(counts.values())), objects=len(counts), groups=counts, stand_in_for="sponza.obj")
    if textured:
        info["texture_images"] = images
        info["textures"] = "%d procedural albedo maps 1024x1024 + %d metalness maps (stand-ins for assets/sponza_textures), planar uv, sampler 0 (REPEAT)" % (
            len([k for k in images if k < 32]), len([k for k in images if k >= 32]))
    return blobs, info


def atrium_with_statue(triangles=800_000, make_builder=None, seed=7):
    """C5 stand-in ("Sponza + dragon", BASELINE.json configs[4]): the atrium with one high-polygon mirror statue (the C3
    stand-in's bumpy blob, ~`triangles` triangles, the reference's mirror material) standing in the nave."""
    n = max(8, int(round((triangles / 2) ** 0.5)))
    st, sn = _blob([0.0, -0.45, 2.2], 0.9, n, n, 0.22, seed=seed, outward=True)
    blobs, info = atrium(make_builder=make_builder, extra=[(st, sn, _scene.material(albedo=(1, 0.84, 0.3), reflectance=1.0))])
    info["stand_in_for"] = "sponza.obj + dragon.obj"
    return blobs, info


def statue_room(triangles=800_000, statue_material=4, seed=7, make_builder=None):
    """C3 stand-in (dragon.obj is not in the reference checkout): a Cornell-style room - white floor/ceiling/back,
    red and green side walls, the reference's light material on a box at the hard-wired NEE quad - around one
    high-polygon statue (bumpy blob, ~`triangles` triangles) with the reference's mirror material 4 (gold-metal
    dragon preset) or dielectric 5. Materials 0..5 are the reference's defaults (src/vk_engine.cpp:689-722)."""
    b = _builder(make_builder)
    b.set_spheres(np.zeros(320, np.uint8))
    b.add_material(_scene.material(albedo=(1, 1, 1)))
    b.add_material(_scene.material(albedo=(1, 0, 0)))
    b.add_material(_scene.material(albedo=(0, 1, 0)))
    b.add_material(_scene.material(albedo=(0, 0, 0), emission=(1, 1, 1), strength=2.4))
    b.add_material(_scene.material(albedo=(1, 0.84, 0.3), reflectance=1.0))
    b.add_material(_scene.material(albedo=(1, 1, 1), ior=2.0))
    lo, hi = np.array([-1.5, -1.52, -4.0]), np.array([1.5, 0.5, 1.5])          # closed: the default camera is inside
    white = [_box(lo, hi, skip=("x0", "x1"))]
    b.add_mesh(_pack(*[np.concatenate(x) for x in zip(*white)]), material=0)
    b.add_mesh(_pack(*_grid([lo[0], hi[1], lo[2]], [0, 0, hi[2] - lo[2]], [0, lo[1] - hi[1], 0], 1, 1)), material=1)
    b.add_mesh(_pack(*_grid([hi[0], hi[1], hi[2]], [0, 0, lo[2] - hi[2]], [0, lo[1] - hi[1], 0], 1, 1)), material=2)
    b.add_mesh(_pack(*_box([-1 / 3, -1.5, -1 / 3], [1 / 3, -1.48, 1 / 3])), material=3)
    n = max(8, int(round((triangles / 2) ** 0.5)))
    st, sn = _blob([0.0, -0.25, 0.2], 0.62, n, n, 0.22, seed=seed, outward=True)
    b.add_mesh(_pack(st, sn), material=statue_material)
    info = dict(triangles=int(len(st)) + 8 + 2 + 2 + 12, objects=5, stand_in_for="dragon.obj")
    return b.blobs(), info


def soup(n, seed=0x5EED, edge=0.01, make_builder=None):
    """n random triangles, centres uniform in [-1,1]^3, one object, material 0 (diffuse white)."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1, 1, (n, 1, 3))
    p = c + rng.uniform(-edge, edge, (n, 3, 3))
    fn = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
    fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-30)
    b = _builder(make_builder)
    b.add_material(_scene.material())
    b.add_mesh(_pack(p, np.repeat(fn[:, None, :], 3, 1)), material=0)
    return b.blobs(), dict(triangles=n, objects=1)


def atrium_camera(pc):
    """Camera inside the nave looking down its length (the reference's default pose already is: (0,-0.5,-3.5), +z)."""
    pc["lightDir"][3] = 1.0        # environment on, as the reference's Sponza preset
    pc.set_camera_angles(-6.0, 8.0, 0.0)
    return pc
