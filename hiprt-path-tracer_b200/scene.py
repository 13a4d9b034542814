"""Scene inputs for the hot path: the reference's `Scene` arrays (src/Scene/SceneParser.h:80-131), a loader shim for the
glTF files the reference ships, the camera maths of src/Scene/Camera.cpp, and the procedural scenes of SURVEY.md section 8(d).

ASSIMP (the reference's importer, pinned TomClabault/assimp@c867ddbc) is absent from /root/reference, so the glTF shim
below restates what SceneParser::parse_scene_file (src/Scene/SceneParser.cpp:22-253, 362-407) obtains from it: one mesh
per material with node transforms pre-applied (aiProcess_PreTransformVertices), material factors mapped like ASSIMP's
glTF2 importer maps them, emissive triangles listed per ThreadFunctions.cpp:116-143. Import parity is "unpinned" (no
reference test pins it); the SAME arrays feed the oracle and the CUDA path, so import differences cancel in parity.

Textures: the 8-bit texture pipeline of SceneParser::prepare_textures / assign_material_texture_indices
(SceneParser.cpp:277-353, 409-471) and ThreadFunctions::load_scene_texture (src/Threads/ThreadFunctions.cpp:30-114) is restated below:
per material the reference's slot order and global index offsets, its channel count per slot, constant emissive textures folded into
the material's emission. Files are decoded with Pillow in place of stb_image and converted between channel counts by stb_image's rules;
tests/test_scene_import.py checks the decode against the reference's own loader (oracle/ref_image.cpp) on the textures it ships."""
import ctypes as C
import json
import os

import numpy as np

from . import structs as S


def default_material():
    """RendererMaterial with the reference's default member initialisers (src/HostDeviceCommon/Material.h:155-268)."""
    m = S.RendererMaterial()
    b = m.base
    b.emission_strength = 1.0
    b.base_color.set(1.0)
    b.roughness = 0.3
    b.oren_nayar_sigma = 0.34906585039886591538
    b.metallic_F90_falloff_exponent = 5.0
    b.metallic_F82.set(1.0)
    b.metallic_F90.set(1.0)
    b.second_roughness = 0.5
    b.specular = 1.0
    b.specular_tint = 1.0
    b.specular_color.set(1.0)
    b.coat_medium_absorption.set(1.0)
    b.coat_medium_thickness = 5.0
    b.coat_roughening = 1.0
    b.coat_darkening = 1.0
    b.coat_ior = 1.5
    b.sheen_roughness = 0.5
    b.sheen_color.set(1.0)
    b.ior = 1.40
    b.absorption_at_distance = 1.0
    b.absorption_color.set(1.0)
    b.dispersion_abbe_number = 20.0
    b.thin_film_ior = 1.3
    b.thin_film_thickness = 500.0
    b.thin_film_base_ior_override = 1.0
    b.srgb = 1
    b.alpha_opacity = 1.0
    b.energy_preservation_monte_carlo_samples = 12
    for name, _ in S.RendererMaterial._fields_[1:]:
        setattr(m, name, -1)
    return m


def make_safe(m):
    """SimplifiedRendererMaterial::make_safe + precompute_properties (Material.h:45-71), called by
    SceneParser::read_material_properties (SceneParser.cpp:405-406)."""
    b = m.base
    f = np.float32
    b.roughness = max(f(1.0e-4), f(b.roughness))
    b.coat_roughness = max(f(1.0e-4), f(b.coat_roughness))
    b.sheen_roughness = max(f(1.0e-4), f(b.sheen_roughness))
    b.emission.set(max(0.0, b.emission.r), max(0.0, b.emission.g), max(0.0, b.emission.b))
    b.emission_strength = max(0.0, b.emission_strength)
    b.absorption_at_distance = max(b.absorption_at_distance, f(1.0e-4))
    b.absorption_color.set(*[max(c, f(1.0 / 512.0)) for c in (b.absorption_color.r, b.absorption_color.g, b.absorption_color.b)])
    b.thin_film_ior = max(f(1.0005), f(b.thin_film_ior))
    if b.specular_transmission == 0.0:
        b.dielectric_priority = (1 << 15) - 1
    return m


def material_is_emissive(m):
    b = m.base
    z = lambda x: -1.0e-10 < x < 1.0e-10
    return (not z(b.emission.r * b.emission_strength)) or (not z(b.emission.g * b.emission_strength)) or (not z(b.emission.b * b.emission_strength)) or bool(b.emissive_texture_used)


class Scene:
    def __init__(self):
        self.vertices = np.zeros((0, 3), np.float32)
        self.indices = np.zeros((0,), np.int32)          # 3 per triangle
        self.normals = np.zeros((0, 3), np.float32)
        self.has_normals = np.zeros((0,), np.uint8)
        self.texcoords = np.zeros((0, 2), np.float32)
        self.material_indices = np.zeros((0,), np.int32)  # per triangle
        self.materials = []                               # list of S.RendererMaterial
        self.emissive = np.zeros((0,), np.int32)
        self.textures = []                                # list of (h, w, c) uint8 arrays
        self.camera = None                                # dict(position, target, up, hfov, near, far) or None
        self.name = ""

    @property
    def triangle_count(self):
        return len(self.indices) // 3

    @property
    def material_count(self):
        return len(self.materials)

    def materials_bytes(self):
        buf = (S.RendererMaterial * max(1, len(self.materials)))()
        for i, m in enumerate(self.materials):
            buf[i] = m
        return np.frombuffer(buf, dtype=np.uint8).reshape(-1, C.sizeof(S.RendererMaterial))[:max(1, len(self.materials))].copy()

    def arrays(self):
        return {
            "vertices": np.ascontiguousarray(self.vertices, np.float32), "indices": np.ascontiguousarray(self.indices, np.int32),
            "normals": np.ascontiguousarray(self.normals, np.float32), "has_normals": np.ascontiguousarray(self.has_normals, np.uint8),
            "texcoords": np.ascontiguousarray(self.texcoords, np.float32), "material_indices": np.ascontiguousarray(self.material_indices, np.int32),
            "materials": self.materials_bytes(), "emissive": np.ascontiguousarray(self.emissive, np.int32),
        }

    def compute_emissive_triangles(self):
        """ThreadFunctions::load_scene_parse_emissive_triangles (src/Threads/ThreadFunctions.cpp:116-143)."""
        emissive_mat = np.array([material_is_emissive(m) and not m.base.emissive_texture_used for m in self.materials], dtype=bool)
        self.emissive = np.nonzero(emissive_mat[self.material_indices])[0].astype(np.int32) if len(self.materials) else np.zeros((0,), np.int32)

    def append_mesh(self, vertices, indices, normals, material_index, texcoords=None):
        vertices = np.asarray(vertices, np.float32).reshape(-1, 3)
        indices = np.asarray(indices, np.int64).reshape(-1)
        offset = len(self.vertices)
        self.vertices = np.concatenate([self.vertices, vertices])
        if normals is not None:
            self.normals = np.concatenate([self.normals, np.asarray(normals, np.float32).reshape(-1, 3)])
            self.has_normals = np.concatenate([self.has_normals, np.ones(len(vertices), np.uint8)])
        else:
            self.normals = np.concatenate([self.normals, np.zeros((len(vertices), 3), np.float32)])
            self.has_normals = np.concatenate([self.has_normals, np.zeros(len(vertices), np.uint8)])
        tc = np.zeros((len(vertices), 2), np.float32) if texcoords is None else np.asarray(texcoords, np.float32).reshape(-1, 2)
        self.texcoords = np.concatenate([self.texcoords, tc])
        self.indices = np.concatenate([self.indices, (indices + offset).astype(np.int32)])
        self.material_indices = np.concatenate([self.material_indices, np.full(len(indices) // 3, material_index, np.int32)])


# ---- 8-bit textures (src/Image/Image.cpp, src/Threads/ThreadFunctions.cpp) ----------------------------
NO_TEXTURE = -1                    # RendererMaterial::NO_TEXTURE (src/HostDeviceCommon/Material.h:231)
CONSTANT_EMISSIVE_TEXTURE = -2     # RendererMaterial::CONSTANT_EMISSIVE_TEXTURE (Material.h:237)


def convert_channels(texels, output_channels):
    """stb_image's stbi__convert_format: what Image8Bit::read_image(path, output_channels) gets when the file has another channel count.
    Grey from colour is the integer luma (77 r + 150 g + 29 b) >> 8, a missing alpha is 255, grey replicates into r, g, b."""
    texels = np.asarray(texels, np.uint8)
    if texels.ndim == 2:
        texels = texels[:, :, None]
    n = texels.shape[2]
    if n == output_channels:
        return np.ascontiguousarray(texels)
    h, w = texels.shape[:2]
    alpha = texels[:, :, n - 1] if n in (2, 4) else np.full((h, w), 255, np.uint8)
    if n >= 3:
        rgb = texels[:, :, :3]
        grey = ((rgb[:, :, 0].astype(np.uint32) * 77 + rgb[:, :, 1].astype(np.uint32) * 150 + rgb[:, :, 2].astype(np.uint32) * 29) >> 8).astype(np.uint8)
    else:
        grey = texels[:, :, 0]
        rgb = np.repeat(grey[:, :, None], 3, axis=2)
    if output_channels == 1:
        out = grey[:, :, None]
    elif output_channels == 2:
        out = np.stack([grey, alpha], axis=2)
    elif output_channels == 3:
        out = rgb
    elif output_channels == 4:
        out = np.concatenate([rgb, alpha[:, :, None]], axis=2)
    else:
        raise ValueError("output_channels must be 1..4")
    return np.ascontiguousarray(out)


def read_image_8bit(path, output_channels):
    """Image8Bit::read_image(path, output_channels, flipY = false) (src/Image/Image.cpp:33-61): rows top to bottom, 8 bits per channel."""
    from PIL import Image
    with Image.open(path) as im:
        if im.mode == "1":
            texels = np.asarray(im.convert("L"))
        elif im.mode in ("I", "I;16", "I;16B", "I;16L", "F"):
            # 16-bit files: stb_image keeps the high byte
            texels = (np.clip(np.asarray(im.convert("I"), np.int64), 0, 65535) >> 8).astype(np.uint8)
        elif im.mode == "P":
            texels = np.asarray(im.convert("RGBA" if "transparency" in im.info else "RGB"))
        elif im.mode in ("L", "LA", "RGB", "RGBA"):
            texels = np.asarray(im)
        else:   # CMYK, YCbCr, ...
            texels = np.asarray(im.convert("RGB"))
    return convert_channels(texels, output_channels)


def image_is_constant_color(texels, threshold=5):
    """Image8Bit::is_constant_color (Image.cpp:250-271): every channel of every texel within `threshold` of the first texel."""
    if texels.size == 0:
        return False
    first = texels[0, 0].astype(np.int32)
    return bool((np.abs(texels.astype(np.int32) - first) <= threshold).all())


def image_sample_rgba32f(texels, u, v):
    """Image8Bit::sample_rgba32f (Image.cpp:161-195): repeat addressing, v flipped, x = int(u (w - 1)), nearest texel, / 255."""
    u, v = np.float32(u), np.float32(v)
    if u != 1.0:
        u = np.float32(u - int(u))
    if v != 1.0:
        v = np.float32(v - int(v))
    u = np.float32(1.0 + u) if u < 0 else u
    v = np.float32(1.0 + v) if v < 0 else v
    v = np.float32(1.0) - v
    h, w = texels.shape[:2]
    x, y = int(u * np.float32(w - 1)), int(v * np.float32(h - 1))
    out = np.zeros(4, np.float32)
    out[:texels.shape[2]] = texels[y, x].astype(np.float32) / np.float32(255.0)
    return out


# Texture slots in the order SceneParser::get_textures_paths_and_indices queries them (SceneParser.cpp:409-447), with the channel count
# ThreadFunctions::load_scene_texture asks for (ThreadFunctions.cpp:43-89). "roughness_metallic" is the packed glTF texture: ASSIMP reports
# pbrMetallicRoughness.metallicRoughnessTexture as DIFFUSE_ROUGHNESS and as METALNESS with one path, which the parser folds into one 3-channel load.
TEXTURE_SLOTS = (("base_color", 4), ("emission", 4), ("roughness_metallic", 3), ("roughness", 1), ("metallic", 1), ("specular", 1), ("coat", 1), ("sheen", 1),
                 ("specular_transmission", 1), ("normal_map", 4))


def _gltf_texture_uri(g, info):
    if not info or "index" not in info:
        return None
    tex = g["textures"][info["index"]]
    if "source" not in tex:
        return None
    uri = g["images"][tex["source"]].get("uri")
    if not uri or uri.startswith("data:"):
        return None
    return uri.replace("%20", " ")          # SceneParser::normalize_texture_paths (SceneParser.cpp:472-490)


def _gltf_material_textures(g, gm):
    """(slot, uri) of one material, in the order the reference pushes them into its texture list."""
    pbr = gm.get("pbrMetallicRoughness", {})
    ext = gm.get("extensions", {})
    found = {
        "base_color": _gltf_texture_uri(g, pbr.get("baseColorTexture")),
        "emission": _gltf_texture_uri(g, gm.get("emissiveTexture")),
        "roughness_metallic": _gltf_texture_uri(g, pbr.get("metallicRoughnessTexture")),
        "specular": _gltf_texture_uri(g, ext.get("KHR_materials_specular", {}).get("specularTexture")),
        "coat": _gltf_texture_uri(g, ext.get("KHR_materials_clearcoat", {}).get("clearcoatTexture")),
        "sheen": _gltf_texture_uri(g, ext.get("KHR_materials_sheen", {}).get("sheenColorTexture")),
        "specular_transmission": _gltf_texture_uri(g, ext.get("KHR_materials_transmission", {}).get("transmissionTexture")),
        "normal_map": _gltf_texture_uri(g, gm.get("normalTexture")),
    }
    return [(slot, found[slot]) for slot, _ in TEXTURE_SLOTS if found.get(slot)]


def import_textures(scene, material_textures, base_directory, read_image=read_image_8bit):
    """prepare_textures + assign_material_texture_indices + load_scene_texture for `material_textures[i]` = [(slot, path)] of material i.
    Fills scene.textures (one (h, w, c) uint8 array per listed texture; a constant emissive texture leaves an empty 1x1 placeholder, like the
    default-constructed Image8Bit the reference leaves there) and the *_texture_index members of the materials."""
    channels = dict(TEXTURE_SLOTS)
    offset = 0
    scene.textures = []
    scene.texture_paths = []
    for material_index, listed in enumerate(material_textures):
        m = scene.materials[material_index]
        for local, (slot, path) in enumerate(listed):
            setattr(m, slot + "_texture_index", local + offset)
            full = os.path.join(base_directory, path)
            texels = read_image(full, channels[slot])
            if slot == "emission" and image_is_constant_color(texels, 5):
                # ThreadFunctions.cpp:91-101: the texture is not kept; its colour becomes the material's emission
                m.emission_texture_index = CONSTANT_EMISSIVE_TEXTURE
                c = image_sample_rgba32f(texels, 0.0, 0.0)
                m.base.emission.set(float(c[0]), float(c[1]), float(c[2]))
                texels = np.zeros((1, 1, channels[slot]), np.uint8)
            scene.textures.append(texels)
            scene.texture_paths.append(full)
        offset += len(listed)
    # emissive_texture_used stays false on the host (only get_intersection_material sets it, Device/includes/Material.h:82), so
    # load_scene_parse_emissive_triangles lists the triangles of a material with a non-zero emission factor whatever its textures
    return scene


# ---- glTF 2.0 loader shim ---------------------------------------------------------------------------
_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _quat_to_mat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]], dtype=np.float64)


def _node_matrix(node):
    if "matrix" in node:
        return np.array(node["matrix"], dtype=np.float64).reshape(4, 4).T
    m = np.eye(4)
    r = _quat_to_mat(node.get("rotation", [0, 0, 0, 1]))
    s = np.array(node.get("scale", [1, 1, 1]), dtype=np.float64)
    m[:3, :3] = r * s[None, :]
    m[:3, 3] = node.get("translation", [0, 0, 0])
    return m


def _gltf_material(gm):
    """glTF material -> RendererMaterial through the keys read_material_properties reads (SceneParser.cpp:362-407),
    with ASSIMP glTF2-importer defaults (metallicFactor and roughnessFactor default to 1, baseColorFactor to white)."""
    m = default_material()
    b = m.base
    pbr = gm.get("pbrMetallicRoughness", {})
    bc = pbr.get("baseColorFactor", [1, 1, 1, 1])
    b.base_color.set(bc[0], bc[1], bc[2])
    em = gm.get("emissiveFactor", [0, 0, 0])
    b.emission.set(em[0], em[1], em[2])
    ext = gm.get("extensions", {})
    if "KHR_materials_emissive_strength" in ext:
        b.emission_strength = ext["KHR_materials_emissive_strength"].get("emissiveStrength", 1.0)
    b.metallic = pbr.get("metallicFactor", 1.0)
    b.roughness = pbr.get("roughnessFactor", 1.0)
    if "KHR_materials_sheen" in ext:
        sc = ext["KHR_materials_sheen"].get("sheenColorFactor", [0, 0, 0])
        b.sheen_color.set(sc[0], sc[1], sc[2])
        b.sheen_roughness = ext["KHR_materials_sheen"].get("sheenRoughnessFactor", 0.0)
        b.sheen = 1.0
    if "KHR_materials_specular" in ext:
        b.specular = ext["KHR_materials_specular"].get("specularFactor", 1.0)
        b.specular_tint = 1.0
        b.specular_color.set(1.0)
    if "KHR_materials_clearcoat" in ext:
        b.coat = ext["KHR_materials_clearcoat"].get("clearcoatFactor", 0.0)
        b.coat_roughness = ext["KHR_materials_clearcoat"].get("clearcoatRoughnessFactor", 0.0)
    if "KHR_materials_ior" in ext:
        b.ior = ext["KHR_materials_ior"].get("ior", 1.5)
    if "KHR_materials_transmission" in ext:
        b.specular_transmission = ext["KHR_materials_transmission"].get("transmissionFactor", 0.0)
    if "KHR_materials_volume" in ext:
        ac = ext["KHR_materials_volume"].get("attenuationColor", [1, 1, 1])
        b.absorption_color.set(ac[0], ac[1], ac[2])
        if "attenuationDistance" in ext["KHR_materials_volume"]:
            b.absorption_at_distance = ext["KHR_materials_volume"]["attenuationDistance"]
    b.alpha_opacity = bc[3]
    return make_safe(m)


def load_gltf(path, load_textures=True):
    g = json.load(open(path))
    base = os.path.dirname(path)
    buffers = [np.fromfile(os.path.join(base, b["uri"]), dtype=np.uint8) for b in g["buffers"]]

    def accessor(i):
        a = g["accessors"][i]
        bv = g["bufferViews"][a["bufferView"]]
        dt = np.dtype(_COMPONENT[a["componentType"]])
        nc = _NCOMP[a["type"]]
        start = bv.get("byteOffset", 0) + a.get("byteOffset", 0)
        stride = bv.get("byteStride", 0) or dt.itemsize * nc
        raw = buffers[bv["buffer"]]
        if stride == dt.itemsize * nc:
            return np.frombuffer(raw, dtype=dt, count=a["count"] * nc, offset=start).reshape(a["count"], nc)
        out = np.empty((a["count"], nc), dtype=dt)
        for k in range(a["count"]):
            out[k] = np.frombuffer(raw, dtype=dt, count=nc, offset=start + k * stride)
        return out

    scene = Scene()
    scene.name = os.path.basename(path)
    scene.materials = [_gltf_material(m) for m in g.get("materials", [])]
    material_textures = [_gltf_material_textures(g, m) if load_textures else [] for m in g.get("materials", [])]

    # Walk the node hierarchy; collect (primitive, world matrix) instances and the first camera
    instances = []
    camera = None

    def walk(ni, parent):
        nonlocal camera
        node = g["nodes"][ni]
        world = parent @ _node_matrix(node)
        if "mesh" in node:
            for prim in g["meshes"][node["mesh"]]["primitives"]:
                instances.append((prim, world))
        if "camera" in node and camera is None:
            camera = (g["cameras"][node["camera"]], world)
        for c in node.get("children", []):
            walk(c, world)

    for root in g["scenes"][g.get("scene", 0)]["nodes"]:
        walk(root, np.eye(4))

    # aiProcess_PreTransformVertices emits one mesh per material, in material order
    instances.sort(key=lambda pw: pw[0].get("material", 0))
    for prim, world in instances:
        pos = accessor(prim["attributes"]["POSITION"]).astype(np.float64)
        pos = (pos @ world[:3, :3].T + world[:3, 3]).astype(np.float32)
        nrm = None
        if "NORMAL" in prim["attributes"]:
            nm = np.linalg.inv(world[:3, :3]).T
            nrm = accessor(prim["attributes"]["NORMAL"]).astype(np.float64) @ nm.T
            nrm = (nrm / np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-30)).astype(np.float32)
        idx = accessor(prim["indices"]).reshape(-1) if "indices" in prim else np.arange(len(pos))
        # texcoords are only kept when the material has textures (SceneParser.cpp:136-140); ASSIMP's glTF2 importer flips v (y = 1 - y)
        tc = None
        mi = prim.get("material", 0)
        if "TEXCOORD_0" in prim["attributes"] and mi < len(material_textures) and len(material_textures[mi]) > 0:
            tc = accessor(prim["attributes"]["TEXCOORD_0"]).astype(np.float32)
            tc = np.stack([tc[:, 0], np.float32(1.0) - tc[:, 1]], axis=1)
        scene.append_mesh(pos, idx, nrm, mi, tc)

    if any(material_textures):
        import_textures(scene, material_textures, base)
    scene.compute_emissive_triangles()

    if camera is not None:
        cam, world = camera
        persp = cam["perspective"]
        aspect = persp.get("aspectRatio", 1.0)
        scene.camera = {
            "position": world[:3, 3].copy(),
            # ASSIMP leaves mLookAt a direction (0,0,-1) rotated by the node; SceneParser.cpp:230-236 feeds it to glm::lookAt as a target POINT
            "target": world[:3, :3] @ np.array([0.0, 0.0, -1.0]),
            "up": world[:3, :3] @ np.array([0.0, 1.0, 0.0]),
            "hfov": 2.0 * np.arctan(np.tan(persp["yfov"] * 0.5) * (aspect if aspect else 1.0)),
            "near": persp.get("znear", 0.1), "far": persp.get("zfar", 1000.0),
        }
    return scene


# ---- camera (src/Scene/Camera.cpp:9-53, SceneParser.cpp:222-253) --------------------------------------
def _look_at(eye, target, up):
    f = target - eye
    f = f / np.linalg.norm(f)
    s = np.cross(f, up)
    s = s / np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[0, 3], m[1, 3], m[2, 3] = -s @ eye, -u @ eye, f @ eye
    return m


def _perspective(fovy, aspect, near, far):
    t = np.tan(fovy / 2.0)
    m = np.zeros((4, 4))
    m[0, 0] = 1.0 / (aspect * t)
    m[1, 1] = 1.0 / t
    m[2, 2] = -(far + near) / (far - near)
    m[3, 2] = -1.0
    m[2, 3] = -(2.0 * far * near) / (far - near)
    return m


def make_camera(position, target, up, hfov, width, height, near=0.1, far=1000.0, jitter=True, vertical_fov=None):
    """HIPRTCamera as Camera::to_hiprt builds it. Matrices are the glm (column-major) matrices reinterpreted as row-major
    float4x4, i.e. transposed: inverse_view = V^-1, inverse_projection = (P^-1)^T, view_projection = P^T V -- kept exactly
    that way because the kernels' matrix_X_point / reprojection arithmetic depends on it (Camera.cpp:13-23)."""
    aspect = width / height
    if vertical_fov is None:
        vertical_fov = 2.0 * np.arctan(np.tan(hfov * 0.5) * aspect) + 0.425   # SceneParser.cpp:247-249
    V = _look_at(np.asarray(position, np.float64), np.asarray(target, np.float64), np.asarray(up, np.float64))
    P = _perspective(vertical_fov, aspect, near, far)
    cam = S.HIPRTCamera()
    cam.inverse_view.set(np.linalg.inv(V).astype(np.float32))
    cam.inverse_projection.set(np.linalg.inv(P).T.astype(np.float32))
    cam.view_projection.set((P.T @ V).astype(np.float32))
    cam.do_jittering = 1 if jitter else 0
    return cam


def scene_camera(scene, width, height, jitter=True):
    c = scene.camera
    if c is None:
        # default camera of SceneParser.cpp:255-275: at the origin looking down -z, 45 degree vertical fov
        return make_camera([0, 0, 0], [0, 0, -1], [0, 1, 0], 0.0, width, height, jitter=jitter, vertical_fov=np.pi / 4)
    return make_camera(c["position"], c["target"], c["up"], c["hfov"], width, height, c["near"], c["far"], jitter)


# ---- scene fixtures (npz) -----------------------------------------------------------------------------
def save_scene(scene, path):
    a = scene.arrays()
    cam = scene.camera or {}
    textures = {f"texture_{i}": np.ascontiguousarray(t, np.uint8) for i, t in enumerate(scene.textures)}
    np.savez_compressed(path, material_count=scene.material_count, name=scene.name, texture_count=len(scene.textures),
                        camera_keys=np.array(sorted(cam.keys())), camera_values=np.array([np.asarray(cam[k], np.float64).ravel() for k in sorted(cam.keys())], dtype=object),
                        **a, **textures)


def load_scene(path):
    z = np.load(path, allow_pickle=True)
    s = Scene()
    s.name = str(z["name"])
    s.vertices, s.indices, s.normals = z["vertices"], z["indices"], z["normals"]
    s.has_normals, s.texcoords, s.material_indices, s.emissive = z["has_normals"], z["texcoords"], z["material_indices"], z["emissive"]
    raw = np.ascontiguousarray(z["materials"])
    n = int(z["material_count"])
    s.materials = [S.RendererMaterial.from_buffer_copy(raw[i].tobytes()) for i in range(n)]
    s.textures = [z[f"texture_{i}"] for i in range(int(z["texture_count"]) if "texture_count" in z.files else 0)]
    keys = [str(k) for k in z["camera_keys"]]
    if keys:
        vals = z["camera_values"]
        s.camera = {k: (np.asarray(v, np.float64) if np.size(v) > 1 else float(np.asarray(v).ravel()[0])) for k, v in zip(keys, vals)}
    return s
