"""Deterministic procedural inputs of the benchmark configurations: the ~10 M-triangle terrain +
spheres scene and its HDR environment map (config C3), the many-light room (config C4). Everything is a pure function of
its seed, so the CUDA path, the oracle and every rank of a multi-GPU job build identical arrays."""
import numpy as np

from . import scene as sc


# ---- environment map ---------------------------------------------------------------------------------------------
def procedural_envmap(width=2048, height=1024, seed=1234, sun_radiance=5.0e4, sun_half_angle_deg=0.5):
    """RGBA32F lat-long sky: vertical gradient + a sun disc of `sun_radiance` within `sun_half_angle_deg` + seeded
    low-amplitude noise (keeps the alias table non-trivial). Row 0 is the first row in memory; envmap_sample maps
    texel (x, y) to theta = y / height * pi, phi = x / width * 2 pi (Envmap.h:95-109)."""
    rng = np.random.default_rng(seed)
    v = (np.arange(height, dtype=np.float64) + 0.5) / height
    u = (np.arange(width, dtype=np.float64) + 0.5) / width
    theta = v[:, None] * np.pi                      # 0 = one pole
    phi = u[None, :] * 2.0 * np.pi
    # direction of the texel, as envmap_sample builds it (Y up)
    d = np.stack([-np.sin(theta) * np.cos(phi), -np.cos(theta) * np.ones_like(phi), -np.sin(theta) * np.sin(phi)], axis=-1)
    up = d[..., 1]
    horizon = np.exp(-np.abs(up) * 4.0)
    sky = np.stack([0.25 + 0.55 * horizon, 0.45 + 0.45 * horizon, 0.95 + 0.05 * horizon], axis=-1) * np.where(up[..., None] > 0, 1.0, 0.15)
    sun_dir = np.array([0.45, 0.55, -0.35])
    sun_dir /= np.linalg.norm(sun_dir)
    cos_sun = d @ sun_dir
    sun = cos_sun > np.cos(np.radians(sun_half_angle_deg))
    noise = rng.uniform(0.9, 1.1, size=(height, width, 1))
    rgb = sky * noise
    rgb[sun] = sun_radiance
    out = np.ones((height, width, 4), np.float32)
    out[..., :3] = rgb.astype(np.float32)
    return out


# ---- geometry ----------------------------------------------------------------------------------------------------
def _fbm(x, z, rng, octaves=5):
    h = np.zeros_like(x)
    amp, freq = 1.0, 1.0
    for _ in range(octaves):
        px, pz = rng.uniform(0, 2 * np.pi, 2)
        ax, az = rng.uniform(0.7, 1.3, 2)
        h += amp * (np.sin(freq * ax * x + px) * np.cos(freq * az * z + pz) + 0.5 * np.sin(freq * 0.7 * (x + z) + px * pz))
        amp *= 0.5
        freq *= 2.1
    return h


def heightfield(quads, extent, amplitude, rng):
    """(quads+1)^2 shared vertices, 2*quads^2 triangles, fBm displacement, no vertex normals (flat shading)."""
    n = quads + 1
    g = np.linspace(-extent, extent, n, dtype=np.float64)
    x, z = np.meshgrid(g, g, indexing="xy")
    y = amplitude * _fbm(x * (6.0 / extent), z * (6.0 / extent), rng)
    verts = np.stack([x, y, z], axis=-1).reshape(-1, 3).astype(np.float32)
    i = np.arange(quads, dtype=np.int64)
    a = (i[:, None] * n + i[None, :]).reshape(-1)          # row-major quad corners
    tri = np.stack([a, a + n, a + 1, a + 1, a + n, a + n + 1], axis=-1).reshape(-1)
    return verts, tri.astype(np.int32)


def icosphere(subdivisions):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    for _ in range(subdivisions):
        edges = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        key = np.sort(edges, axis=1)
        uniq, inv = np.unique(key, axis=0, return_inverse=True)
        mid = v[uniq[:, 0]] + v[uniq[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        m = inv.reshape(3, -1).T + len(v)                  # midpoints of edges (01, 12, 20) per face
        v = np.concatenate([v, mid])
        a, b, c = f[:, 0], f[:, 1], f[:, 2]
        ab, bc, ca = m[:, 0], m[:, 1], m[:, 2]
        f = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    return v, f


def random_materials(count, rng):
    """The fixed distributions of config C3: metallic in {0,1} (p = 0.3), roughness U[0.05,1], coat U[0,1] on 25 %,
    sheen on 25 %, transmission 1 with ior U[1.1,2] on 10 %, anisotropy U[0,0.8]; no textures, no emission."""
    mats = []
    for _ in range(count):
        m = sc.default_material()
        b = m.base
        b.base_color.set(*rng.uniform(0.2, 0.95, 3))
        b.metallic = 1.0 if rng.uniform() < 0.3 else 0.0
        b.roughness = rng.uniform(0.05, 1.0)
        b.anisotropy = rng.uniform(0.0, 0.8)
        if rng.uniform() < 0.25:
            b.coat = rng.uniform(0.0, 1.0)
            b.coat_roughness = rng.uniform(0.0, 0.3)
        if rng.uniform() < 0.25:
            b.sheen = rng.uniform(0.2, 1.0)
        if rng.uniform() < 0.10:
            b.specular_transmission = 1.0
            b.metallic = 0.0
            b.ior = rng.uniform(1.1, 2.0)
            b.dielectric_priority = 1
        mats.append(sc.make_safe(m))
    return mats


def terrain_and_spheres(quads=1581, spheres=3906, subdivisions=3, materials=64, seed=1234):
    """Config C3: height-field of quads x quads quads + `spheres` flattened icospheres (1280 triangles each at 3
    subdivisions). Defaults give 2*1581^2 + 3906*1280 = 9 998 802 triangles."""
    rng = np.random.default_rng(seed)
    s = sc.Scene()
    s.name = f"terrain_and_spheres(quads={quads}, spheres={spheres}, subdiv={subdivisions}, seed={seed})"
    s.materials = random_materials(materials, rng)
    extent, amplitude = 50.0, 2.5
    hv, hi = heightfield(quads, extent, amplitude, rng)
    nverts = [hv]
    nidx = [hi]
    nnorm = [np.zeros_like(hv)]
    nhas = [np.zeros(len(hv), np.uint8)]
    # terrain material per triangle: bands of the first 8 materials (coherent regions, several materials)
    tri_centroid_x = hv[hi.reshape(-1, 3)[:, 0], 0]
    nmat = [np.clip(((tri_centroid_x + extent) / (2 * extent) * 8).astype(np.int32), 0, 7) % max(1, materials)]
    sv, sf = icosphere(subdivisions)
    offset = len(hv)
    if spheres > 0:
        centers = np.stack([rng.uniform(-extent * 0.9, extent * 0.9, spheres), rng.uniform(1.0, 9.0, spheres), rng.uniform(-extent * 0.9, extent * 0.9, spheres)], axis=-1)
        radii = rng.uniform(0.25, 1.2, spheres)
        mat_of = rng.integers(0, materials, spheres)
        allv = (sv[None, :, :] * radii[:, None, None] + centers[:, None, :]).astype(np.float32)
        alln = np.broadcast_to(sv[None, :, :], allv.shape).astype(np.float32)
        alli = (sf.reshape(-1)[None, :] + (offset + np.arange(spheres, dtype=np.int64) * len(sv))[:, None]).astype(np.int32)
        nverts.append(allv.reshape(-1, 3))
        nnorm.append(alln.reshape(-1, 3))
        nhas.append(np.ones(spheres * len(sv), np.uint8))
        nidx.append(alli.reshape(-1))
        nmat.append(np.repeat(mat_of.astype(np.int32), len(sf)))
    s.vertices = np.concatenate(nverts)
    s.normals = np.concatenate(nnorm)
    s.has_normals = np.concatenate(nhas)
    s.indices = np.concatenate(nidx)
    s.material_indices = np.concatenate(nmat)
    s.texcoords = np.zeros((len(s.vertices), 2), np.float32)
    s.compute_emissive_triangles()
    s.camera = dict(position=np.array([0.0, 12.0, 62.0]), target=np.array([0.0, 2.0, 0.0]), up=np.array([0.0, 1.0, 0.0]), hfov=0.6, near=0.1, far=1000.0)
    return s


def _box(center, half, rot):
    c = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], np.float64) * half
    f = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4], [2, 3, 7], [2, 7, 6], [1, 2, 6], [1, 6, 5], [0, 4, 7], [0, 7, 3]], np.int64)
    return c @ rot.T + center, f


def _random_rotations(n, rng):
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    return np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                     np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                     np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], axis=1)


def many_lights_room(boxes=2000, lights=100_000, emissive_materials=256, seed=4321):
    """Config C4: closed room 20 x 10 x 20 (12 triangles) + `boxes` diffuse boxes (12 triangles each) + `lights` emissive
    triangles of edge 2-5 cm, random pose, random hue x strength 50, spread over `emissive_materials` materials."""
    rng = np.random.default_rng(seed)
    s = sc.Scene()
    s.name = f"many_lights_room(boxes={boxes}, lights={lights}, seed={seed})"
    wall = sc.default_material()
    wall.base.base_color.set(0.75, 0.75, 0.75)
    wall.base.roughness = 0.8
    s.materials.append(sc.make_safe(wall))
    for _ in range(7):
        m = sc.default_material()
        m.base.base_color.set(*rng.uniform(0.2, 0.9, 3))
        m.base.roughness = rng.uniform(0.2, 1.0)
        m.base.metallic = 1.0 if rng.uniform() < 0.2 else 0.0
        s.materials.append(sc.make_safe(m))
    first_emissive = len(s.materials)
    for _ in range(emissive_materials):
        m = sc.default_material()
        hue = rng.uniform(0.0, 1.0, 3)
        hue /= hue.max()
        m.base.emission.set(*hue)
        m.base.emission_strength = 50.0
        s.materials.append(sc.make_safe(m))
    # room: inward-facing box
    rv, rf = _box(np.array([0.0, 5.0, 0.0]), np.array([10.0, 5.0, 10.0]), np.eye(3))
    s.append_mesh(rv, rf[:, ::-1].reshape(-1), None, 0)
    if boxes > 0:
        rots = _random_rotations(boxes, rng)
        centers = np.stack([rng.uniform(-9, 9, boxes), rng.uniform(0.3, 6.0, boxes), rng.uniform(-9, 9, boxes)], -1)
        halves = rng.uniform(0.1, 0.45, (boxes, 3))
        vs, fs, ms = [], [], []
        for k in range(boxes):
            bv, bf = _box(centers[k], halves[k], rots[k])
            vs.append(bv)
            fs.append(bf.reshape(-1) + 8 * k)
            ms.append(np.full(12, 1 + k % 7, np.int32))
        off = len(s.vertices)
        s.vertices = np.concatenate([s.vertices, np.concatenate(vs).astype(np.float32)])
        s.normals = np.concatenate([s.normals, np.zeros((8 * boxes, 3), np.float32)])
        s.has_normals = np.concatenate([s.has_normals, np.zeros(8 * boxes, np.uint8)])
        s.texcoords = np.concatenate([s.texcoords, np.zeros((8 * boxes, 2), np.float32)])
        s.indices = np.concatenate([s.indices, (np.concatenate(fs) + off).astype(np.int32)])
        s.material_indices = np.concatenate([s.material_indices, np.concatenate(ms)])
    if lights > 0:
        centers = np.stack([rng.uniform(-9.5, 9.5, lights), rng.uniform(0.2, 9.8, lights), rng.uniform(-9.5, 9.5, lights)], -1)
        rots = _random_rotations(lights, rng)
        edge = rng.uniform(0.02, 0.05, lights)
        local = np.array([[-0.5, -0.2887, 0.0], [0.5, -0.2887, 0.0], [0.0, 0.5774, 0.0]])
        tv = np.einsum("nij,kj->nki", rots, local) * edge[:, None, None] + centers[:, None, :]
        off = len(s.vertices)
        s.vertices = np.concatenate([s.vertices, tv.reshape(-1, 3).astype(np.float32)])
        s.normals = np.concatenate([s.normals, np.zeros((3 * lights, 3), np.float32)])
        s.has_normals = np.concatenate([s.has_normals, np.zeros(3 * lights, np.uint8)])
        s.texcoords = np.concatenate([s.texcoords, np.zeros((3 * lights, 2), np.float32)])
        s.indices = np.concatenate([s.indices, (np.arange(3 * lights, dtype=np.int64) + off).astype(np.int32)])
        s.material_indices = np.concatenate([s.material_indices, (first_emissive + rng.integers(0, emissive_materials, lights)).astype(np.int32)])
    s.compute_emissive_triangles()
    s.camera = dict(position=np.array([0.0, 5.0, 9.5]), target=np.array([0.0, 4.5, 0.0]), up=np.array([0.0, 1.0, 0.0]), hfov=0.9, near=0.05, far=100.0)
    return s
