"""Texture import (hiprt-path-tracer_b200/scene.py) against the reference's 8-bit texture pipeline:
SceneParser::prepare_textures / assign_material_texture_indices (src/Scene/SceneParser.cpp:277-353, 409-471), ThreadFunctions::load_scene_texture
(src/Threads/ThreadFunctions.cpp:30-114) and Image8Bit::read_image / is_constant_color / sample_rgba32f (src/Image/Image.cpp:33-61, 161-195, 250-271).
The decode is checked against the reference's own loader (oracle/ref_image.cpp over stb_image); the imported scene is rendered by the CUDA path
and by the reference CPU kernels from the same arrays.

The only textured scene the reference ships, data/GLTFs/the-white-room-low.gltf, comes without its .bin buffer, so the scene here is a
generated glTF (a lit room, one material per texture case); the-white-room's 16 JPEG files serve as decoder inputs where /root/reference exists."""
import glob
import json
import os
import struct

import numpy as np
import pytest

WHITE_ROOM = "/root/reference/data/GLTFs/the-white-room"


@pytest.fixture(scope="module")
def sc():
    return __import__("importlib").import_module("hiprt-path-tracer_b200.scene")


def _image_oracle():
    from oracle import ref
    if not os.path.exists(ref.image_lib_path()):
        pytest.skip("oracle/_ref image library is not built")
    return ref


# ---- a generated textured glTF ------------------------------------------------------------------------------------------------------------------
def _quad(p0, ex, ey, n, uv_scale=1.0):
    p0, ex, ey = (np.asarray(v, np.float32) for v in (p0, ex, ey))
    pos = np.stack([p0, p0 + ex, p0 + ex + ey, p0 + ey])
    nrm = np.tile(np.asarray(n, np.float32), (4, 1))
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32) * uv_scale - 0.25 * (uv_scale > 1)          # leaves [0, 1] when scaled: repeat addressing
    return pos, nrm, uv, np.array([0, 1, 2, 0, 2, 3], np.uint32)


def _grid(p0, ex, ey, n, cells, uv_scale=1.0):
    """A quad cut into cells x cells quads (more triangles, texcoords interpolate across them)."""
    p0, ex, ey = (np.asarray(v, np.float32) for v in (p0, ex, ey))
    k = np.arange(cells + 1, dtype=np.float32) / cells
    u, v = np.meshgrid(k, k, indexing="xy")
    pos = p0 + u[..., None] * ex + v[..., None] * ey
    uv = np.stack([u, v], axis=-1) * uv_scale
    idx = []
    for j in range(cells):
        for i in range(cells):
            a = j * (cells + 1) + i
            idx += [a, a + 1, a + cells + 2, a, a + cells + 2, a + cells + 1]
    return pos.reshape(-1, 3).astype(np.float32), np.tile(np.asarray(n, np.float32), ((cells + 1) ** 2, 1)), uv.reshape(-1, 2).astype(np.float32), np.array(idx, np.uint32)


def write_textured_gltf(directory):
    """Room 2 x 2 x 2 around the origin, open towards +z where the camera sits. Returns (path, expected texture list)."""
    from PIL import Image
    rng = np.random.default_rng(5)
    os.makedirs(os.path.join(directory, "tex dir"), exist_ok=True)

    def save(name, arr, **kw):
        Image.fromarray(arr).save(os.path.join(directory, name), **kw)

    yy, xx = np.mgrid[0:64, 0:64]
    checker = (((xx // 8 + yy // 8) % 2) * 150 + 60).astype(np.uint8)
    floor = np.stack([checker, np.roll(checker, 3, 1), 255 - checker], axis=2)
    save("floor.png", floor)
    save("floor_orm.png", rng.integers(30, 226, size=(16, 16, 3), dtype=np.uint8))
    smooth = np.stack([(xx * 3 + 40), (yy * 2 + 60), ((xx + yy) + 80)], axis=2).astype(np.uint8)
    save("tex dir/back wall.jpg", smooth, quality=92)
    rgba = rng.integers(50, 256, size=(32, 16, 4), dtype=np.uint8)
    save("left.png", rgba)
    nm = rng.integers(110, 146, size=(8, 8, 3), dtype=np.uint8)
    nm[..., 2] = 255
    save("left_normal.png", nm)
    save("light.png", np.full((4, 4, 3), (250, 247, 249), np.uint8) + rng.integers(0, 4, size=(4, 4, 3), dtype=np.uint8))      # constant within threshold 5
    glow = np.zeros((8, 8, 3), np.uint8)
    glow[2:6, 2:6] = (255, 180, 90)
    save("ceiling_glow.png", glow)
    save("grey.png", rng.integers(0, 256, size=(8, 8), dtype=np.uint8))                                                        # 1 channel -> RGBA request

    images = ["floor.png", "floor_orm.png", "tex%20dir/back%20wall.jpg", "left.png", "left_normal.png", "light.png", "ceiling_glow.png", "grey.png"]
    tex = {name: i for i, name in enumerate(images)}
    T = lambda name: {"index": tex[name]}   # noqa: E731
    materials = [
        {"name": "floor", "pbrMetallicRoughness": {"baseColorTexture": T("floor.png"), "metallicRoughnessTexture": T("floor_orm.png"), "baseColorFactor": [1, 1, 1, 1]}},
        {"name": "back", "pbrMetallicRoughness": {"baseColorTexture": T("tex%20dir/back%20wall.jpg"), "metallicFactor": 0.0, "roughnessFactor": 0.8}},
        {"name": "left", "pbrMetallicRoughness": {"baseColorTexture": T("left.png"), "metallicFactor": 0.0, "roughnessFactor": 0.5}, "normalTexture": T("left_normal.png")},
        {"name": "right", "pbrMetallicRoughness": {"baseColorFactor": [0.2, 0.7, 0.3, 1.0], "metallicFactor": 0.0, "roughnessFactor": 0.6}},
        {"name": "light", "emissiveTexture": T("light.png"), "emissiveFactor": [1, 1, 1], "extensions": {"KHR_materials_emissive_strength": {"emissiveStrength": 12.0}},
         "pbrMetallicRoughness": {"baseColorFactor": [0, 0, 0, 1]}},
        {"name": "ceiling", "emissiveTexture": T("ceiling_glow.png"), "pbrMetallicRoughness": {"baseColorFactor": [0.7, 0.7, 0.7, 1], "metallicFactor": 0.0, "roughnessFactor": 0.9}},
        {"name": "box", "pbrMetallicRoughness": {"baseColorTexture": T("grey.png"), "metallicFactor": 0.0, "roughnessFactor": 0.4}},
    ]
    meshes_geo = [
        _grid([-1, -1, -1], [2, 0, 0], [0, 0, 2], [0, 1, 0], 4, uv_scale=2.0),        # floor (material 0), texcoords up to 2: wraps
        _grid([-1, -1, -1], [2, 0, 0], [0, 2, 0], [0, 0, 1], 3),                        # back wall
        _grid([-1, -1, 1], [0, 0, -2], [0, 2, 0], [1, 0, 0], 3, uv_scale=1.5),         # left wall
        _quad([1, -1, -1], [0, 0, 2], [0, 2, 0], [-1, 0, 0]),                           # right wall: texcoords present in the file, dropped on import
        _quad([-0.3, 0.98, -0.3], [0.6, 0, 0], [0, 0, 0.6], [0, -1, 0]),                # light
        _grid([-1, 1, -1], [2, 0, 0], [0, 0, 2], [0, -1, 0], 2),                        # ceiling
        _grid([-0.5, -1, -0.4], [0.6, 0, 0.2], [0, 0.9, 0], [-0.316, 0, 0.949], 2),     # a slanted panel standing on the floor
    ]
    blob = bytearray()
    views, accessors, meshes, nodes = [], [], [], []

    def add(arr, type_, component, target=None, minmax=False):
        nonlocal blob
        while len(blob) % 4:
            blob += b"\0"
        view = {"buffer": 0, "byteOffset": len(blob), "byteLength": arr.nbytes}
        if target:
            view["target"] = target
        blob += arr.tobytes()
        views.append(view)
        acc = {"bufferView": len(views) - 1, "componentType": component, "count": len(arr), "type": type_}
        if minmax:
            acc["min"], acc["max"] = arr.min(0).tolist(), arr.max(0).tolist()
        accessors.append(acc)
        return len(accessors) - 1

    for mi, (pos, nrm, uv, idx) in enumerate(meshes_geo):
        prim = {"attributes": {"POSITION": add(pos, "VEC3", 5126, 34962, True), "NORMAL": add(nrm, "VEC3", 5126, 34962), "TEXCOORD_0": add(uv, "VEC2", 5126, 34962)},
                "indices": add(idx if mi % 2 else idx.astype(np.uint16), "SCALAR", 5125 if mi % 2 else 5123, 34963), "material": mi}
        meshes.append({"primitives": [prim]})
        nodes.append({"mesh": mi})
    nodes.append({"camera": 0, "translation": [0.0, 0.0, 3.4]})
    g = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": list(range(len(nodes)))}], "nodes": nodes, "meshes": meshes, "materials": materials,
         "cameras": [{"type": "perspective", "perspective": {"yfov": 0.7, "aspectRatio": 1.5, "znear": 0.1, "zfar": 100.0}}],
         "textures": [{"source": i} for i in range(len(images))], "images": [{"uri": u} for u in images],
         "accessors": accessors, "bufferViews": views, "buffers": [{"uri": "room.bin", "byteLength": len(blob)}],
         "extensionsUsed": ["KHR_materials_emissive_strength"]}
    open(os.path.join(directory, "room.bin"), "wb").write(bytes(blob))
    path = os.path.join(directory, "room.gltf")
    json.dump(g, open(path, "w"))
    # (material, slot, file, channels) in the order the reference lists textures: material by material, base colour, emission, packed roughness-metallic, ..., normals
    expected = [(0, "base_color", "floor.png", 4), (0, "roughness_metallic", "floor_orm.png", 3), (1, "base_color", "tex dir/back wall.jpg", 4),
                (2, "base_color", "left.png", 4), (2, "normal_map", "left_normal.png", 4), (4, "emission", "light.png", 4), (5, "emission", "ceiling_glow.png", 4),
                (6, "base_color", "grey.png", 4)]
    return path, expected


# ---- host logic (no GPU) ----------------------------------------------------------------------------------------------------------------------------
def test_convert_channels_follows_stb_image(sc):
    rng = np.random.default_rng(1)
    rgb = rng.integers(0, 256, size=(5, 7, 3), dtype=np.uint8)
    luma = ((rgb[..., 0].astype(int) * 77 + rgb[..., 1].astype(int) * 150 + rgb[..., 2].astype(int) * 29) >> 8).astype(np.uint8)
    assert np.array_equal(sc.convert_channels(rgb, 1)[..., 0], luma)
    assert np.array_equal(sc.convert_channels(rgb, 2), np.stack([luma, np.full_like(luma, 255)], axis=2))
    assert np.array_equal(sc.convert_channels(rgb, 4), np.concatenate([rgb, np.full((5, 7, 1), 255, np.uint8)], axis=2))
    grey_alpha = rng.integers(0, 256, size=(5, 7, 2), dtype=np.uint8)
    assert np.array_equal(sc.convert_channels(grey_alpha, 4), np.concatenate([np.repeat(grey_alpha[..., :1], 3, axis=2), grey_alpha[..., 1:]], axis=2))
    assert np.array_equal(sc.convert_channels(grey_alpha, 3), np.repeat(grey_alpha[..., :1], 3, axis=2))
    rgba = rng.integers(0, 256, size=(5, 7, 4), dtype=np.uint8)
    assert np.array_equal(sc.convert_channels(rgba, 3), rgba[..., :3]) and np.array_equal(sc.convert_channels(rgba, 2)[..., 1], rgba[..., 3])
    assert sc.convert_channels(rgb[..., 0], 3).shape == (5, 7, 3)


def test_decoder_matches_the_reference_loader(sc, tmp_path):
    """Image8Bit::read_image (stb_image) against read_image_8bit (Pillow + stb's conversion rules): PNG is lossless, so every mode x requested
    channel count is byte-equal, as are is_constant_color and the texel sample_rgba32f((0, 0)) returns; baseline JPEG decoders may round the
    inverse DCT and the chroma upsampling differently, so JPEG is bounded: mean difference under half a level, no texel off by more than 16."""
    ref = _image_oracle()
    from PIL import Image
    rng = np.random.default_rng(2)
    files = {}
    for mode, c in (("L", 1), ("LA", 2), ("RGB", 3), ("RGBA", 4)):
        arr = rng.integers(0, 256, size=(9, 13, c), dtype=np.uint8)
        files[mode] = str(tmp_path / f"{mode}.png")
        Image.fromarray(arr[..., 0] if c == 1 else arr, mode).save(files[mode])
    files["P"] = str(tmp_path / "P.png")
    Image.fromarray(rng.integers(0, 256, size=(9, 13, 3), dtype=np.uint8)).quantize(16).save(files["P"])
    files["I16"] = str(tmp_path / "I16.png")
    Image.fromarray(rng.integers(0, 65536, size=(9, 13), dtype=np.uint16)).save(files["I16"])
    files["const"] = str(tmp_path / "const.png")
    Image.fromarray(np.full((6, 6, 3), (200, 100, 50), np.uint8) + rng.integers(0, 5, size=(6, 6, 3), dtype=np.uint8)).save(files["const"])
    for name, path in files.items():
        for channels in (1, 2, 3, 4):
            want, constant, at00 = ref.read_image_8bit(path, channels)
            got = sc.read_image_8bit(path, channels)
            assert got.shape == want.shape and np.array_equal(got, want), (name, channels)
            assert sc.image_is_constant_color(got, 5) == constant, (name, channels)
            assert np.array_equal(sc.image_sample_rgba32f(got, 0.0, 0.0)[:channels], at00[:channels]), (name, channels)
    assert sc.image_is_constant_color(sc.read_image_8bit(files["const"], 4), 5) and not sc.image_is_constant_color(sc.read_image_8bit(files["RGB"], 4), 5)
    jpegs = sorted(glob.glob(os.path.join(WHITE_ROOM, "*.jpg")))
    if not jpegs:
        jpegs = [str(tmp_path / "own.jpg")]
        yy, xx = np.mgrid[0:96, 0:128]
        Image.fromarray(np.stack([xx * 2, yy * 2, xx + yy], axis=2).astype(np.uint8)).save(jpegs[0], quality=90)
    worst = (0.0, 0)
    for path in jpegs:
        want, _, _ = ref.read_image_8bit(path, 4)
        got = sc.read_image_8bit(path, 4)
        assert got.shape == want.shape, path
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        worst = (max(worst[0], float(d.mean())), max(worst[1], int(d.max())))
        assert d.mean() < 0.5 and d.max() <= 16 and np.array_equal(got[..., 3], want[..., 3]), (path, d.mean(), d.max())
    print(f"{len(jpegs)} JPEG files: worst mean |difference| {worst[0]:.3f} levels, worst texel {worst[1]}")


def test_gltf_texture_import_follows_the_reference_parser(sc, tmp_path):
    path, expected = write_textured_gltf(str(tmp_path))
    scene = sc.load_gltf(path)
    assert len(scene.textures) == len(expected) == len(scene.texture_paths)
    for i, (material, slot, name, channels) in enumerate(expected):
        assert getattr(scene.materials[material], slot + "_texture_index") == (sc.CONSTANT_EMISSIVE_TEXTURE if name == "light.png" else i), (i, slot)
        assert scene.texture_paths[i].endswith(name)
        if name != "light.png":
            assert scene.textures[i].shape[2] == channels and scene.textures[i].dtype == np.uint8
            assert np.array_equal(scene.textures[i], sc.read_image_8bit(os.path.join(str(tmp_path), name), channels))
    # every other slot of every material stays NO_TEXTURE
    listed = {(m, s) for m, s, _, _ in expected}
    for mi, m in enumerate(scene.materials):
        for slot, _ in sc.TEXTURE_SLOTS:
            if (mi, slot) not in listed:
                assert getattr(m, slot + "_texture_index") == sc.NO_TEXTURE, (mi, slot)
    # the packed glTF texture is ONE 3-channel load; the single-channel file was widened to RGBA (grey, grey, grey, 255)
    assert scene.textures[1].shape == (16, 16, 3)
    grey = scene.textures[7]
    assert (grey[..., 0] == grey[..., 1]).all() and (grey[..., 3] == 255).all()
    # constant emissive texture: folded into the material (emission = texel / 255, strength from the extension), placeholder left in the list
    light = scene.materials[4]
    assert scene.textures[5].shape == (1, 1, 4)
    e = np.array([light.base.emission.r, light.base.emission.g, light.base.emission.b])
    assert np.all(np.abs(e * 255.0 - np.array([250, 247, 249])) <= 4.0) and light.base.emission_strength == 12.0
    ceiling = scene.materials[5]
    assert ceiling.emission_texture_index == 6 and ceiling.base.emission.r == 0.0
    # emissive triangles: the light's two (constant texture folded); the ceiling's emission lives in a texture and is not importance-sampled
    light_triangles = np.nonzero(scene.material_indices == 4)[0]
    assert np.array_equal(scene.emissive, light_triangles.astype(np.int32)) and len(light_triangles) == 2
    # texcoords: kept (v flipped, as ASSIMP's glTF2 importer hands them over) only for meshes whose material lists textures
    tri = np.nonzero(scene.material_indices == 3)[0]
    assert not scene.texcoords[scene.indices.reshape(-1, 3)[tri].ravel()].any()
    floor_vertices = scene.indices.reshape(-1, 3)[np.nonzero(scene.material_indices == 0)[0]].ravel()
    uv = scene.texcoords[floor_vertices]
    assert uv[:, 0].max() == 2.0 and uv[:, 1].min() == -1.0 and uv[:, 1].max() == 1.0
    assert scene.triangle_count == 2 * (16 + 9 + 9 + 1 + 1 + 4 + 4) and scene.camera is not None
    # the fixture format keeps the textures
    sc.save_scene(scene, str(tmp_path / "room.npz"))
    again = sc.load_scene(str(tmp_path / "room.npz"))
    assert len(again.textures) == len(scene.textures) and all(np.array_equal(a, b) for a, b in zip(again.textures, scene.textures))
    assert again.materials[0].base_color_texture_index == 0 and np.array_equal(again.texcoords, scene.texcoords)
    # load_textures = False: geometry and factors only
    plain = sc.load_gltf(path, load_textures=False)
    assert not plain.textures and not plain.texcoords.any() and plain.materials[0].base_color_texture_index == sc.NO_TEXTURE


# ---- GPU: the imported scene through both renderers -------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_imported_textured_gltf_renders_like_the_reference(pt, wl, ctx, ref, sc, tmp_path):
    """load_gltf -> Scene (textures, texcoords, folded emission) -> CUDA renderer and reference CPU kernels, matched RNG streams: G-buffer
    materials byte-equal except the sRGB base colour (one ulp of powf), shading normals (normal map) byte-equal, radiance within 1e-3."""
    path, _ = write_textured_gltf(str(tmp_path))
    scene = sc.load_gltf(path)
    W, H, spp = 168, 112, 4
    cfg = {"name": "generated textured room", "scene": scene, "camera": sc.scene_camera(scene, W, H), "width": W, "height": H,
           "options": dict(lss="mis", ess="none", ggx=0, later=3), "oracle_variant": "mis", "envmap": None,
           "settings": dict(nb_bounces=6, do_alpha_testing=0, enable_adaptive_sampling=0, samples_per_frame=1)}
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    try:
        r.render(spp)
        for _ in range(spp):
            g.render_frame()
        g.synchronize()
        hit = r.buffer("camera_ray_hit") == 1
        assert hit.mean() > 0.9
        assert np.array_equal(r.buffer("first_hit_prim_index"), g.download("first_hit_prim_index"))
        ma, mb = r.buffer("materials")[hit][:, 4:244].copy(), g.download("materials")[hit][:, 4:244].copy()
        o = pt.structs.SimplifiedRendererMaterial.base_color.offset - 4
        ca, cb = ma[:, o:o + 12].copy().view(np.float32), mb[:, o:o + 12].copy().view(np.float32)
        assert np.allclose(ca, cb, rtol=1.5e-7, atol=0.0)
        ma[:, o:o + 12] = 0
        mb[:, o:o + 12] = 0
        assert np.array_equal(ma, mb)
        assert np.array_equal(r.buffer("shading_normals")[hit], g.download("shading_normals")[hit])
        a, b = r.pixels() / spp, g.download("pixels") / spp
        assert a.mean() > 0.02                                  # the folded emissive texture does light the room
        err = float((np.abs(a - b) / np.maximum(a, 1e-3)).mean())
        assert err < 1e-3, err
    finally:
        g.close()
