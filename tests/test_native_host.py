"""The native host (rtb200_cli: C++ TOML reader, scene loader, camera, PNG writer — rust-ray-tracing_b200/csrc/host/)
against Python's tomllib and the Python mirror of the reference's loaders (src/main.rs:27-74, src/materials.rs:12-33,
src/objects.rs:38-52,292-299, src/geometry.rs:190-211, src/color.rs:99-123, src/toml_utils.rs:2-11)."""
import json
import math
import os
import struct
import subprocess
import tomllib
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "rust-ray-tracing_b200", "rtb200_cli")


def run(args, check=True):
    r = subprocess.run([CLI] + args, capture_output=True, text=True)
    if check and r.returncode != 0:
        raise AssertionError(f"rtb200_cli {args} failed: {r.stderr}")
    return r


def same(a, b):
    if isinstance(a, float) or isinstance(b, float):
        if isinstance(a, bool) or isinstance(b, bool):
            return False
        if isinstance(a, float) != isinstance(b, float):
            return False            # TOML keeps integer and float apart
        return (math.isnan(a) and math.isnan(b)) or a == b
    if isinstance(a, dict):
        return isinstance(b, dict) and a.keys() == b.keys() and all(same(a[k], b[k]) for k in a)
    if isinstance(a, list):
        return isinstance(b, list) and len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
    return type(a) is type(b) and a == b


VALID = [
    'a = 1\nb = "x"\nc = 1.5\nd = true\n',
    '# comment\n[materials.ground]\ntype = "lambertian" # trailing\nalbedo = [0.5, 0.5, 0.5]\n\n[[hitables]]\ntype = "sphere"\ncenter = { x = 0, y = -1000.0, z = 0 }\nradius = 1000\nmaterial = "ground"\n',
    'x = [1, 2,\n  3, # c\n  4,\n]\ny = [[1, 2], ["a", \'b\']]\nz = { a = { b = 1 }, c = [1.0, 2e3, -3E-2, 1_000.5] }\n',
    'i = [0x1F, 0o17, 0b101, 1_000, -17, +5, 0]\nf = [inf, -inf, +inf, 1e10, 6.02e+23, -0.0, 3.14]\nn = nan\n',
    's1 = "tab\\tquote\\" uni\\u00e9 \\U0001F600"\ns2 = \'C:\\raw\\path\'\ns3 = """\nmulti\nline \\\n   joined"""\ns4 = \'\'\'\nlit\nraw\'\'\'\n',
    'a.b.c = 1\na.b.d = 2\n"quoted key" = 3\n\'lit key\'.x = 4\n[t."sub.table"]\nk = 1\n',
    '[[fruit]]\nname = "apple"\n[fruit.physical]\ncolor = "red"\n[[fruit.variety]]\nname = "red delicious"\n[[fruit.variety]]\nname = "granny"\n[[fruit]]\nname = "banana"\n',
    '[a]\n[a.b]\nx = 1\n[c]\ny = {}\nz = []\n',
    'title = "crlf"\r\nv = 1\r\n[t]\r\nw = 2\r\n',
]

INVALID = [
    'a = 1\na = 2\n',
    'a = \n',
    'a = [1, 2\n',
    '[t]\n[t]\n',
    'a = 01\n',
    'a = 1__0\n',
    'a = "unterminated\n',
    'a = {b = 1}\na.c = 2\n',
    'a = 1 b = 2\n',
    '[[t]]\n[t]\n',
]


@pytest.mark.parametrize("i", range(len(VALID)))
def test_toml_reader_matches_tomllib(tmp_path, i):
    p = tmp_path / "doc.toml"
    p.write_bytes(VALID[i].encode("utf-8"))
    got = json.loads(run([str(p), "--dump-toml"]).stdout)
    want = tomllib.loads(VALID[i])
    assert same(got, want), (got, want)


@pytest.mark.parametrize("i", range(len(INVALID)))
def test_toml_reader_rejects_what_tomllib_rejects(tmp_path, i):
    with pytest.raises(tomllib.TOMLDecodeError):
        tomllib.loads(INVALID[i])
    p = tmp_path / "doc.toml"
    p.write_bytes(INVALID[i].encode("utf-8"))
    r = run([str(p), "--dump-toml"], check=False)
    assert r.returncode == 1 and "Failed parsing scene from file" in r.stderr     # src/main.rs:36-39


def test_toml_dates_are_reported_as_unsupported(tmp_path):
    p = tmp_path / "doc.toml"
    p.write_text("d = 1979-05-27\n")
    r = run([str(p), "--dump-toml"], check=False)
    assert r.returncode == 1 and "dates and times are not supported" in r.stderr


SCENE = '''
[materials.ground]
type = "Lambertian"
albedo = { red = 0.5, green = 0.5, blue = 1 }
[materials.steel]
type = "METAL"
albedo = [0.7, 0.6, 0.5, 99]
fuzzy_factor = 3
[materials.glass]
type = "dielectric"
index_of_refraction = 1.5
hollow = true
[materials.odd]
type = "lambertian"
albedo = { red = "x", green = 0.25, blue = 0.75 }

[[hitables]]
type = "Sphere"
center = [0, -1000, 0]
radius = 1000
material = "ground"
[[hitables]]
type = "sphere"
center = { x = 4.0, y = 1, z = 0.5 }
radius = 1.0
material = "steel"
[[hitables]]
type = "SPHERE"
center = [0.0, 1.0, 0.0, 7.0]
radius = 0.5
material = "glass"
[[hitables]]
type = "sphere"
center = [1, 2, 3]
radius = 2
material = "odd"
'''


def test_scene_loader_matches_python_host(rtb, tmp_path):
    p = tmp_path / "scene.toml"
    p.write_text(SCENE)
    got = json.loads(run([str(p), "--dump-scene"]).stdout)
    scene = rtb.Scene.from_toml(SCENE)
    spheres, mat_index, kind, hollow, albedo, param = scene.arrays()
    assert len(got["spheres"]) == len(scene) == 4 and got["n_materials"] == 4
    for i, s in enumerate(got["spheres"]):
        m = mat_index[i]
        assert s["center"] == list(spheres[i, :3]) and s["radius"] == spheres[i, 3]
        assert s["material"] == scene.objects[i].material
        assert s["kind"] == kind[m] and s["hollow"] == hollow[m]
        assert s["albedo"] == list(albedo[m]) and s["param"] == param[m]
    assert got["spheres"][3]["albedo"] == [0.0, 0.25, 0.75]            # wrongly typed channel -> unwrap_or(0.0), src/color.rs:102
    assert got["spheres"][1]["param"] == 1.0                           # fuzzy_factor = 3 clamped by Metal::new, src/materials.rs:79-88


def test_generated_scenes_round_trip(rtb, tmp_path):
    for name in ("three_spheres", "random_spheres", "dielectric_heavy"):
        scene = getattr(rtb.scenes, name)()
        p = tmp_path / f"{name}.toml"
        p.write_text(scene.to_toml())
        got = json.loads(run([str(p), "--dump-scene"]).stdout)
        spheres, mat_index, kind, hollow, albedo, param = scene.arrays()
        assert len(got["spheres"]) == len(scene)
        c = np.array([s["center"] + [s["radius"]] for s in got["spheres"]])
        assert np.array_equal(c, spheres)
        assert [s["kind"] for s in got["spheres"]] == list(kind[mat_index])
        assert np.array_equal(np.array([s["albedo"] for s in got["spheres"]]), albedo[mat_index])
        assert np.array_equal(np.array([s["param"] for s in got["spheres"]]), param[mat_index])


@pytest.mark.parametrize("mutation,message", [
    (('type = "METAL"', 'type = "plastic"'), "Unknown material type plastic!"),          # src/materials.rs:31
    (('type = "SPHERE"', 'type = "cube"'), "Unknown object type cube"),                  # src/objects.rs:50
    (('material = "glass"', 'material = "nope"'), "does not exist"),                     # src/objects.rs:296
    (('hollow = true', 'hollow = 1'), "hollow must be a boolean"),                       # src/materials.rs:151
    (('radius = 1.0', 'radius = "big"'), "expected a TOML float or integer"),            # src/toml_utils.rs:2-11
    (('center = [1, 2, 3]', 'center = [1, 2]'), "array.len() >= 3"),                     # src/geometry.rs:200
    (('fuzzy_factor = 3', ''), 'missing key "fuzzy_factor"'),
    (('[materials.ground]', '[stuff.ground]'), "does not exist"),
])
def test_scene_loader_errors_follow_the_reference(rtb, tmp_path, mutation, message):
    text = SCENE.replace(*mutation)
    p = tmp_path / "scene.toml"
    p.write_text(text)
    r = run([str(p), "--dump-scene"], check=False)
    assert r.returncode == 1 and message in r.stderr, r.stderr
    with pytest.raises(Exception):
        rtb.Scene.from_toml(text)


def test_missing_scene_file_message():
    r = run(["/nonexistent/scene.toml", "--dump-scene"], check=False)
    assert r.returncode == 1 and "Can't read scene from file /nonexistent/scene.toml" in r.stderr   # src/main.rs:31-34


def test_camera_matches_python_host_and_oracle(rtb, oracle, tmp_path):
    p = tmp_path / "scene.toml"
    p.write_text(SCENE + '\n[camera]\nimage_width = 320\nimage_height = 200\nfocal_length = 3.5\nview_angle = 40.0\n'
                         'center = [1, 2, 3]\nlook_at = { x = 0, y = 0.5, z = -1 }\nup = [0, 1, 0]\ndefocus_angle = 0.6\n')
    got = json.loads(run([str(p), "--dump-camera", "--camera-from-scene"]).stdout)
    cam = rtb.Camera(320, 200, 3.5, 40.0, (1, 2, 3), (0, 0.5, -1), (0, 1, 0), 0.6)
    vec = np.asarray(cam.vectors()).reshape(6, 3)
    keys = ["center", "viewport_upper_left_corner", "viewport_u", "viewport_v", "defocus_u", "defocus_v"]
    for k, v in zip(keys, vec):
        assert got[k] == list(v)
    ref = np.asarray(oracle.camera_new(320, 200, 3.5, 40.0, (1, 2, 3), (0, 0.5, -1), (0, 1, 0), 0.6)).reshape(6, 3)
    assert np.allclose(vec, ref, rtol=1e-15, atol=1e-15)
    # defaults = the camera main hard-codes (src/main.rs:45-58)
    d = json.loads(run([str(p), "--dump-camera"]).stdout)
    assert (d["image_width"], d["image_height"]) == (3840, 2160) and d["center"] == [16.0, 2.0, 18.5]
    # the reference's Camera::from_toml insists on TOML floats for focal_length / view_angle (src/ray_tracing.rs:67-68)
    p.write_text(SCENE + '\n[camera]\nimage_width = 320\nimage_height = 200\nfocal_length = 3\nview_angle = 40.0\n'
                         'center = [1, 2, 3]\nlook_at = [0, 0, 0]\nup = [0, 1, 0]\n')
    r = run([str(p), "--dump-camera", "--camera-from-scene"], check=False)
    assert r.returncode == 1 and "must be TOML floats" in r.stderr


def read_png(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert crc == zlib.crc32(tag + body) & 0xFFFFFFFF
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 2)
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert np.all(raw[:, 0] == 0)
    return raw[:, 1:].reshape(h, w, 3)


@pytest.mark.gpu
def test_cli_renders_the_same_image_as_the_python_host(rtb, renderer, tmp_path):
    """scene.toml -> output.png through the native binary == Renderer.render through ctypes, byte for byte."""
    scene = rtb.scenes.three_spheres()
    p = tmp_path / "scene.toml"
    p.write_text(scene.to_toml())
    out = tmp_path / "output.png"
    r = run([str(p), "-o", str(out), "--width", "160", "--height", "90", "--spp", "8", "--depth", "50", "--seed", "5", "--stats-json"])
    assert "Image Size: 160 x 90" in r.stdout and "Total Pixels: 14400" in r.stdout and "Average Pixel Rate:" in r.stdout
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    renderer.seed = 5
    img, stat = renderer.render(50, 8, scene, rtb.Camera.default(160, 90))
    assert np.array_equal(read_png(out), img)
    assert stats["paths"] == 160 * 90 * 8 and stats["segments"] == stat.segments
    ppm = tmp_path / "o.ppm"
    run([str(p), "-o", str(ppm), "--width", "160", "--height", "90", "--spp", "8", "--seed", "5", "--quiet"])
    body = open(ppm, "rb").read()
    assert body.startswith(b"P6\n160 90\n255\n") and np.array_equal(np.frombuffer(body[len(b"P6\n160 90\n255\n"):], np.uint8).reshape(90, 160, 3), img)


@pytest.mark.gpu
def test_cli_progress_and_both_roots(rtb, renderer, tmp_path):
    """--progress draws on stderr without changing the image; --both-roots == the Python host with set_hit_mode(True)."""
    scene = rtb.scenes.three_spheres()
    p = tmp_path / "scene.toml"
    p.write_text(scene.to_toml())
    a, b = tmp_path / "a.png", tmp_path / "b.png"
    r = run([str(p), "-o", str(a), "--width", "320", "--height", "180", "--spp", "32", "--seed", "3", "--quiet", "--progress"])
    assert "100.0 %" in r.stderr and r.stdout.strip() == ""
    renderer.seed = 3
    img, _ = renderer.render(50, 32, scene, rtb.Camera.default(320, 180))
    assert np.array_equal(read_png(a), img)
    run([str(p), "-o", str(b), "--width", "320", "--height", "180", "--spp", "32", "--seed", "3", "--quiet", "--both-roots"])
    renderer.set_hit_mode(True)
    try:
        img2, _ = renderer.render(50, 32, scene, rtb.Camera.default(320, 180))
    finally:
        renderer.set_hit_mode(False)
    assert np.array_equal(read_png(b), img2) and not np.array_equal(img2, img)
