"""Malformed mesh assets through paths_scene_load_yaml (CPU, no device needed): the reference panics on these (`unwrap`/`expect` in
src/obj.rs:9-40, src/ply.rs:13-70, src/bvh.rs:201 "Must have at least one cluster"); the C ABI must return a status and a message."""
import os
import struct
import textwrap

import pytest

SCENE = textwrap.dedent("""\
    camera:
      image_width: 32
      image_height: 24
      location: {{ x: 0.0, y: 1.0, z: -5.0 }}
      orientation: {{ pitch: 0.0, yaw: 0.0, roll: 0.0 }}
      sensor_width: 0.036
      sensor_height: 0.024
      focal_length: 0.05
      focus_distance: 5.0
      aperture: 8.0
    skybox:
      type: Gradient
      overhead_colour: {{ r: 0.5, g: 0.8, b: 0.9 }}
      horizon_colour: {{ r: 0.7, g: 0.85, b: 0.9 }}
    models:
      m:
        file: "{mesh}"
    lights: []
    objects:
      - shape:
          type: Mesh
          model: m
          translation: {{ x: 0.0, y: 0.0, z: 0.0 }}
          rotation: {{ pitch: 0.0, yaw: 0.0, roll: 0.0 }}
          scale: 1.0
        material:
          type: Lambertian
          albedo: {{ type: Rgb, r: 0.5, g: 0.5, b: 0.5 }}
    """)

PLY = (b"ply\nformat binary_little_endian 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\n"
       b"element face 1\nproperty list uchar int vertex_indices\nend_header\n" + struct.pack("<9f", 0, 0, 0, 1, 0, 0, 0, 1, 0) + struct.pack("<B3i", 3, 0, 1, 2))
OBJ = b"v 0 0 0\nv 1 0 0\nv 0 1 0\n"

# name -> (bytes, expected: number of primitives, or a fragment of the error message)
CASES = {
    "good.ply": (PLY, 1),
    "trunc_body.ply": (PLY[:-7], "truncated"),
    "trunc_header.ply": (PLY[:60], "not a PLY"),
    "bad_index.ply": (PLY[:-12] + struct.pack("<3i", 0, 1, 99), "out of range"),
    "neg_index.ply": (PLY[:-12] + struct.pack("<3i", 0, -1, 2), "out of range"),
    "huge_count.ply": (PLY.replace(b"element vertex 3", b"element vertex 4000000000"), "more rows than the file can hold"),
    "endless_element.ply": (PLY.replace(b"end_header", b"element junk 18446744073709551615\nend_header"), "more rows than the file can hold"),
    "empty.ply": (b"", "not a PLY"),
    "garbage.ply": (bytes(range(256)), "not a PLY"),
    "good.obj": (OBJ + b"f 1 2 3\n", 1),
    "relative_index.obj": (OBJ + b"f -1 -2 -3\n", 1),
    "bad_index.obj": (OBJ + b"f 1 2 9\n", "out of range"),
    "zero_index.obj": (OBJ + b"f 0 1 2\n", "out of range"),
    "junk.obj": (b"v 0 0\nf\nf 1\n", "fewer than 3"),
    "nan_vertex.obj": (b"v nan 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n", "at least one cluster"),  # NaN face normal: dropped (model.rs:174-249)
    "empty.obj": (b"", "at least one cluster"),
    "no_faces.obj": (OBJ, "at least one cluster"),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_malformed_assets_return_a_status(tmp_path, name):
    import paths_b200 as pb

    data, expect = CASES[name]
    mesh = tmp_path / name
    mesh.write_bytes(data)
    yml = tmp_path / (name + ".yml")
    yml.write_text(SCENE.format(mesh=str(mesh)))
    if isinstance(expect, int):
        assert pb.Scene.load(str(yml)).info["n_prims"] == expect
    else:
        with pytest.raises(pb.PathsError) as e:
            pb.Scene.load(str(yml))
        assert expect in str(e.value), str(e.value)


def test_missing_asset_and_missing_scene(tmp_path):
    import paths_b200 as pb

    yml = tmp_path / "s.yml"
    yml.write_text(SCENE.format(mesh=str(tmp_path / "nope.obj")))
    with pytest.raises(pb.PathsError):
        pb.Scene.load(str(yml))
    with pytest.raises(pb.PathsError):
        pb.Scene.load(str(tmp_path / "nope.yml"))
    assert os.path.exists(yml)


HDR_SCENE = textwrap.dedent("""\
    camera:
      image_width: 16
      image_height: 8
      location: {{ x: 0.0, y: 1.0, z: -5.0 }}
      orientation: {{ pitch: 0.0, yaw: 0.0, roll: 0.0 }}
      sensor_width: 0.036
      sensor_height: 0.024
      focal_length: 0.05
      focus_distance: 5.0
      aperture: 8.0
    skybox:
      type: Hdri
      filename: "{hdr}"
    models: {{}}
    lights: []
    objects:
      - shape:
          type: Sphere
          radius: 1.0
          center: {{ x: 0.0, y: 1.0, z: 0.0 }}
        material:
          type: Mirror
    """)
HDR_HEAD = b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 2 +X 8\n"
HDR_CASES = {
    "good.hdr": (HDR_HEAD + bytes([128, 128, 128, 129]) * 16, True),
    "short_flat.hdr": (HDR_HEAD + bytes([128, 128, 128, 129]) * 9, False),
    "rle_overrun.hdr": (HDR_HEAD + bytes([2, 2, 0, 8]) + bytes([128 + 100, 7]) * 4, False),  # a run longer than the scanline
    "rle_truncated.hdr": (HDR_HEAD + bytes([2, 2, 0, 8]) + bytes([128 + 8, 7]), False),
    "no_resolution.hdr": (b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n", False),
    "bad_magic.hdr": (b"P6\n8 2\n255\n" + bytes(48), False),
    "zero_size.hdr": (b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 0 +X 0\n", False),
    "empty.hdr": (b"", False),
    "huge_resolution.hdr": (b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 4000000000 +X 4000000000\n" + bytes(64), False),  # refused before allocating
}


@pytest.mark.parametrize("name", sorted(HDR_CASES))
def test_malformed_hdr_returns_a_status(tmp_path, name):
    import paths_b200 as pb

    data, ok = HDR_CASES[name]
    hdr = tmp_path / name
    hdr.write_bytes(data)
    yml = tmp_path / (name + ".yml")
    yml.write_text(HDR_SCENE.format(hdr=str(hdr)))
    if ok:
        assert pb.Scene.load(str(yml)).info["n_prims"] == 1
    else:
        with pytest.raises(pb.PathsError):
            pb.Scene.load(str(yml))
