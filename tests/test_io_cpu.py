"""Host-side formats (guava_renderer_b200/io.py): the Gaussian PLY of ubody_gaussian.py:350-420 and the
avatar tensor dictionary.  plyfile is not installed, so the expected file is spelled out byte by byte."""
import numpy as np
import pytest
import torch

from guava_renderer_b200 import io as gio
from guava_renderer_b200 import synthetic


def _cloud(P=257, seed=3):
    c = synthetic.guava_cloud(P, seed)
    return c["xyz"], c["features_color"], c["opacity"], c["scaling"], c["rotation"]


def test_ply_bytes_are_what_the_reference_writes(tmp_path):
    xyz, feat, op, sc, rot = _cloud(5)
    p = str(tmp_path / "GS_canonical.ply")
    gio.write_gaussian_ply(p, xyz, feat, op, sc, rot)
    raw = open(p, "rb").read()
    header = ("ply\nformat binary_little_endian 1.0\nelement vertex 5\n" +
              "".join(f"property float {a}\n" for a in
                      ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2", "opacity",
                       "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]) + "end_header\n").encode()
    assert raw.startswith(header) and len(raw) == len(header) + 5 * 17 * 4
    body = np.frombuffer(raw[len(header):], "<f4").reshape(5, 17)
    # ubody_gaussian.py:355-372: xyz, zero normals, rgb / C0, inverse_sigmoid(opacity), log(scale), rotation
    assert np.array_equal(body[:, 0:3], xyz.numpy()) and not body[:, 3:6].any()
    assert np.allclose(body[:, 6:9], feat.numpy()[:, :3] / 0.28209479177387814, rtol=1e-6)
    o = op.numpy()
    assert np.allclose(body[:, 9:10], np.log(o / (1 - o)), rtol=1e-5, atol=1e-6)
    assert np.allclose(body[:, 10:13], np.log(sc.numpy()), rtol=1e-6)
    assert np.array_equal(body[:, 13:17], rot.numpy())


@pytest.mark.parametrize("P", [0, 1, 257])
def test_ply_round_trip(tmp_path, P):
    xyz, feat, op, sc, rot = _cloud(P) if P else [torch.zeros(0, w) for w in (3, 32, 1, 3, 4)]
    p = str(tmp_path / "a.ply")
    gio.write_gaussian_ply(p, xyz, feat, op, sc, rot)
    g = gio.read_gaussian_ply(p)
    assert g["xyz"].shape == (1, P, 3) and g["features_color"].shape == (1, P, 32)
    assert torch.equal(g["xyz"][0], xyz)
    assert torch.allclose(g["features_color"][0, :, :3], feat[:, :3], rtol=1e-5, atol=1e-7)
    assert not g["features_color"][0, :, 3:].any()
    assert torch.allclose(g["opacity"][0], op, rtol=1e-5, atol=1e-6)
    assert torch.allclose(g["scaling"][0], sc, rtol=1e-5)
    assert torch.allclose(g["rotation"][0], rot, atol=1e-6)


def test_reads_ascii_big_endian_and_extra_elements(tmp_path):
    vals = np.arange(2 * 17, dtype=np.float32).reshape(2, 17) / 7
    names = list(gio.PLY_ATTRIBUTES)
    p1 = tmp_path / "ascii.ply"
    p1.write_text("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\n" +
                  "".join(f"property float {n}\n" for n in names) + "element face 0\nproperty list uchar int vertex_indices\nend_header\n" +
                  "\n".join(" ".join(repr(float(x)) for x in r) for r in vals) + "\n")
    p2 = tmp_path / "be.ply"
    hdr = ("ply\nformat binary_big_endian 1.0\nelement vertex 2\n" + "".join(f"property double {n}\n" for n in names) +
           "property uchar red\nend_header\n").encode()
    dt = np.dtype([(n, ">f8") for n in names] + [("red", "u1")])
    body = np.zeros(2, dt)
    for i, n in enumerate(names):
        body[n] = vals[:, i]
    p2.write_bytes(hdr + body.tobytes())
    a, b = gio.read_ply_vertices(str(p1)), gio.read_ply_vertices(str(p2))
    for i, n in enumerate(names):
        assert np.allclose(a[n], vals[:, i]) and np.allclose(b[n], vals[:, i])
    assert b["red"].dtype == np.uint8
    with pytest.raises(ValueError):
        (tmp_path / "bad.ply").write_bytes(hdr + body.tobytes()[:-3])
        gio.read_ply_vertices(str(tmp_path / "bad.ply"))
    with pytest.raises(ValueError):
        (tmp_path / "no.ply").write_text("plx\n")
        gio.read_ply_vertices(str(tmp_path / "no.ply"))


def test_avatar_state_round_trip(tmp_path):
    from oracle import deform_oracle as do
    a = do.synthetic_avatar(1, 20, 30, 50, seed=1)
    per_avatar = {k: v for k, v in a.items() if k not in ("vertices", "ver_transform_mat", "faces")}
    p = str(tmp_path / "avatar.pt")
    gio.save_avatar_state(p, **per_avatar)
    st = torch.load(p, weights_only=True)
    assert "_uv_local_xyz" in st and "_smplx_features_color" in st          # the reference's attribute names
    st["ehm.some_buffer"] = torch.zeros(3)                                    # foreign keys are ignored
    torch.save({("module." + k): v for k, v in st.items()}, p)
    back = gio.load_avatar_state(p)
    assert set(back) == set(per_avatar)
    for k in per_avatar:
        assert torch.equal(back[k], per_avatar[k]), k
    torch.save({"_uv_rotation": torch.zeros(1, 2, 4)}, p)
    with pytest.raises(KeyError):
        gio.load_avatar_state(p)
