"""Host-side logic (no GPU): the TOML scene schema and loaders mirror the reference's, including its
error behaviour — what panics there raises here."""
import numpy as np
import pytest

SCENE = """
[materials.ground]
type = "Lambertian"
albedo = [0.5, 0.5, 0.5]

[materials.mirror]
type = "metal"
albedo = {red = 0.7, green = 0.6, blue = 0.5}
fuzzy_factor = 3

[materials.glass]
type = "DIELECTRIC"
index_of_refraction = 1.5
hollow = false

[[hitables]]
type = "Sphere"
center = [0, -1000, 0]
radius = 1000
material = "ground"

[[hitables]]
type = "sphere"
center = {x = 4.0, y = 1, z = 0.5}
radius = 1.0
material = "mirror"

[[hitables]]
type = "sphere"
center = [0.0, 1.0, 0.0, 99.0]
radius = 1
material = "glass"
"""


def test_scene_from_toml(rtb):
    s = rtb.Scene.from_toml(SCENE)
    assert len(s) == 3
    assert s.materials["ground"].kind == 0 and s.materials["ground"].albedo == (0.5, 0.5, 0.5)
    m = s.materials["mirror"]
    assert m.kind == 1 and m.albedo == (0.7, 0.6, 0.5)
    assert m.param == 1.0                                   # Metal::new clamps fuzzy_factor to <= 1 (materials.rs:79-88)
    g = s.materials["glass"]
    assert g.kind == 2 and g.param == 1.5 and g.hollow is False
    assert s.objects[0].center == (0.0, -1000.0, 0.0) and s.objects[0].radius == 1000.0   # integers accepted (toml_utils.rs)
    assert s.objects[1].center == (4.0, 1.0, 0.5)
    assert s.objects[2].center == (0.0, 1.0, 0.0)           # arrays longer than 3 are accepted (geometry.rs:200)
    assert [o.material for o in s.objects] == ["ground", "mirror", "glass"]


def test_scene_toml_roundtrip(rtb):
    for gen in ("three_spheres", "random_spheres", "dielectric_heavy"):
        s = getattr(rtb.scenes, gen)()
        t = rtb.Scene.from_toml(s.to_toml())
        assert len(t) == len(s)
        for a, b in zip(s.arrays(), t.arrays()):
            assert np.array_equal(a, b)


@pytest.mark.parametrize("bad,exc", [
    ('[materials.m]\ntype="plastic"\n[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius=1\nmaterial="m"\n', ValueError),   # materials.rs:31
    ('[materials.m]\ntype="lambertian"\nalbedo=[1,1,1]\n[[hitables]]\ntype="cube"\ncenter=[0,0,0]\nradius=1\nmaterial="m"\n', ValueError),  # objects.rs:50
    ('[materials.m]\ntype="lambertian"\nalbedo=[1,1,1]\n[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius=1\nmaterial="x"\n', KeyError),   # objects.rs:296
    ('[materials.m]\ntype="lambertian"\nalbedo=[1,1]\n[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius=1\nmaterial="m"\n', ValueError),   # color.rs:110
    ('[materials.m]\ntype="dielectric"\nindex_of_refraction=1.5\n[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius=1\nmaterial="m"\n', KeyError),  # hollow required, materials.rs:151
    ('[materials.m]\ntype="metal"\nalbedo=[1,1,1]\n[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius=1\nmaterial="m"\n', KeyError),        # fuzzy_factor required
    ('[materials.m]\ntype="lambertian"\nalbedo=[1,1,1]\n[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius="big"\nmaterial="m"\n', TypeError),
    ('[[hitables]]\ntype="sphere"\n', KeyError),                                                                                      # scene["materials"], main.rs:41
])
def test_scene_errors_follow_the_reference(rtb, bad, exc):
    with pytest.raises(exc):
        rtb.Scene.from_toml(bad)


def test_color_table_wrong_type_is_zero(rtb):
    # to_float(...).unwrap_or(0.0): a wrongly typed channel becomes 0.0 (src/color.rs:101-105)
    s = rtb.Scene.from_toml('[materials.m]\ntype="lambertian"\nalbedo={red="x", green=0.5, blue=1}\n'
                            '[[hitables]]\ntype="sphere"\ncenter=[0,0,0]\nradius=1\nmaterial="m"\n')
    assert s.materials["m"].albedo == (0.0, 0.5, 1.0)


def test_generators_match_survey(rtb):
    assert len(rtb.scenes.three_spheres()) == 4
    n = len(rtb.scenes.random_spheres())
    assert 470 <= n <= 490                                   # "~485 spheres"
    assert len(rtb.scenes.ten_k_spheres()) == 10001
    s4 = rtb.scenes.dielectric_heavy()
    kinds = np.array([s4.materials[o.material].kind for o in s4.objects])
    assert (kinds == 2).mean() > 0.7
    # generators are seeded: same scene every time
    a, b = rtb.scenes.random_spheres(), rtb.scenes.random_spheres()
    assert all(np.array_equal(x, y) for x, y in zip(a.arrays(), b.arrays()))


def test_camera_matches_oracle_camera_new(rtb, oracle):
    cam = rtb.Camera(1920, 1080, 10.0, 30.0, (16.0, 2.0, 18.5), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 0.6)
    ref = oracle.camera_new(1920, 1080, 10.0, 30.0, (16.0, 2.0, 18.5), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 0.6)
    assert np.array_equal(np.array(cam.vectors()), ref)      # same f64 arithmetic, bit for bit
    assert cam.image_width() == 1920 and cam.image_height() == 1080


def test_camera_from_toml(rtb):
    t = {"image_width": 64, "image_height": 36, "focal_length": 10.0, "view_angle": 30.0, "center": [16, 2, 18.5],
         "look_at": {"x": 0, "y": 0, "z": 0}, "up": [0, 1, 0]}
    cam = rtb.Camera.from_toml(t)
    assert cam.vectors() == rtb.Camera.default(64, 36).vectors()
    with pytest.raises(TypeError):
        rtb.Camera.from_toml({**t, "focal_length": 10})      # as_float().unwrap() on an integer (ray_tracing.rs:67)


def test_image_writers(rtb, tmp_path):
    import struct, zlib
    img = (np.arange(5 * 7 * 3) % 256).astype(np.uint8).reshape(5, 7, 3)
    rtb.write_ppm(str(tmp_path / "a.ppm"), img)
    raw = (tmp_path / "a.ppm").read_bytes()
    assert raw.startswith(b"P6\n7 5\n255\n") and raw.endswith(img.tobytes())
    rtb.write_png(str(tmp_path / "a.png"), img)
    png = (tmp_path / "a.png").read_bytes()
    assert png[:8] == b"\x89PNG\r\n\x1a\n"
    w, h = struct.unpack(">II", png[16:24])
    assert (w, h) == (7, 5)
    idat = png[png.index(b"IDAT") + 4: png.index(b"IEND") - 8]
    rows = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(5, 1 + 7 * 3)
    assert np.array_equal(rows[:, 1:].reshape(5, 7, 3), img) and np.all(rows[:, 0] == 0)
