"""Row N3: NanoVDB Vec3f source files. The fixture tests/golden/source_vec3f.nvdb and its samples were produced by the
reference's own NanoVDB headers (tests/golden/make_nvdb_golden.cpp: tools::createNanoGrid + io::writeGrid, sampled with
worldToIndex + SampleFromVoxels<Vec3fTree, 1, false> exactly like integrator/uniform/integrator.cu:309-311).
CPU: the reader (csrc/nvdb.h through the C ABI) + the dense trilinear lookup reproduce those samples.
GPU: a configuration with scene.source_path runs through elaina-exec's code path and equals the dense-grid run."""
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVDB = os.path.join(ROOT, "tests", "golden", "source_vec3f.nvdb")
SAMPLES = os.path.join(ROOT, "tests", "golden", "source_vec3f_samples.txt")


def _trilinear(dense, origin, voxel, pts):
    """the walker's source_lookup (csrc/walk.cuh) in numpy: zero outside the grid"""
    nz, ny, nx, _ = dense.shape
    x = (pts - np.asarray(origin, np.float32)) * np.float32(1.0 / voxel)
    fl = np.floor(x)
    f = (x - fl).astype(np.float32)
    i0 = fl.astype(np.int64)
    out = np.zeros((len(pts), 3), np.float32)
    for a in (0, 1):
        for b in (0, 1):
            for c in (0, 1):
                i, j, k = i0[:, 0] + a, i0[:, 1] + b, i0[:, 2] + c
                ok = (i >= 0) & (j >= 0) & (k >= 0) & (i < nx) & (j < ny) & (k < nz)
                v = np.zeros((len(pts), 3), np.float32)
                v[ok] = dense[k[ok], j[ok], i[ok]]
                w = (f[:, 0] if a else 1 - f[:, 0]) * (f[:, 1] if b else 1 - f[:, 1]) * (f[:, 2] if c else 1 - f[:, 2])
                out += v * w[:, None]
    return out


def test_reader_reproduces_reference_samples():
    from elaina_b200 import scenes
    dense, origin, voxel = scenes.load_nvdb_source(NVDB)
    assert dense.shape == (9, 11, 14, 3) and voxel == 0.25            # index bbox [3 122 2]-[16 132 10]
    np.testing.assert_allclose(origin, (-1.5 + 3 * 0.25, -1.0 + 122 * 0.25, -0.75 + 2 * 0.25))
    # voxel values: f(i, j, k) of the generator, the hole is an inactive voxel holding the background
    i, j, k = 5, 130, 7
    np.testing.assert_allclose(dense[k - 2, j - 122, i - 3], (0.1 * i + 0.01 * j, np.sin(np.float32(0.3 * j)) + 0.2 * k, 0.05 * i * k - 0.3), rtol=1e-6)
    assert np.all(dense[5 - 2, 127 - 122, 9 - 3] == 0.0)
    g = np.loadtxt(SAMPLES, dtype=np.float64)
    pts, want = g[:, :3].astype(np.float32), g[:, 3:].astype(np.float32)
    got = _trilinear(dense, origin, voxel, pts)
    assert (np.abs(want).sum(1) > 0).sum() > 150                        # the samples do cover the grid
    np.testing.assert_allclose(got, want, rtol=0, atol=3e-5)            # fp32 index arithmetic differs by rounding only


def test_zip_codec_file_equals_uncompressed():
    """Codec::ZIP (io/IO.h:221-236,279-292), fixture written by the reference's own NanoVDB headers built with NANOVDB_USE_ZIP
    (tests/golden/make_nvdb_golden.cpp, third argument): the reader inflates it to the same dense grid as the Codec::NONE file."""
    from elaina_b200 import scenes
    a, oa, va = scenes.load_nvdb_source(NVDB)
    b, ob, vb = scenes.load_nvdb_source(os.path.join(ROOT, "tests", "golden", "source_vec3f_zip.nvdb"))
    assert a.shape == b.shape and oa == ob and va == vb
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_reader_errors_are_status_codes(tmp_path):
    from elaina_b200 import _lib, scenes
    with pytest.raises(_lib.WostError) as e:
        scenes.load_nvdb_source(tmp_path / "missing.nvdb")
    assert e.value.status == 3                                           # WOST_ERR_IO
    bad = tmp_path / "bad.nvdb"
    bad.write_bytes(b"not a nanovdb file at all....")
    with pytest.raises(_lib.WostError) as e:
        scenes.load_nvdb_source(bad)
    assert "bad magic" in str(e.value)
    raw = bytearray(open(NVDB, "rb").read())
    raw[14] = 2                                                          # FileHeader.codec = BLOSC (no blosc library in this image)
    z = tmp_path / "blosc.nvdb"
    z.write_bytes(bytes(raw))
    with pytest.raises(_lib.WostError) as e:
        scenes.load_nvdb_source(z)
    assert "BLOSC" in str(e.value)
    t = tmp_path / "trunc.nvdb"
    t.write_bytes(bytes(open(NVDB, "rb").read()[:5000]))
    with pytest.raises(_lib.WostError):
        scenes.load_nvdb_source(t)


@pytest.mark.gpu
def test_source_path_config_equals_dense_run(tmp_path):
    from elaina_b200 import _lib, api, scenes
    dense, origin, voxel = scenes.load_nvdb_source(NVDB)
    # a sphere around the centre of the source box
    c = np.array(origin) + 0.5 * voxel * np.array([dense.shape[2], dense.shape[1], dense.shape[0]])
    dv, dp = scenes.icosphere(3, 1.6, 0.05)
    dv = (dv + c.astype(np.float32)).astype(np.float32)
    scenes.save_obj(tmp_path / "model.obj", dv, dp)
    scenes.save_vertex_colors(tmp_path / "color.json", scenes.gray_pairs(scenes.harmonic_3d(dv - c.astype(np.float32))))
    setting = {"frameSize": [24, 24], "maxWalkingDepth": 24, "samplesPerPixel": 6, "epsilonShell": 0.01, "seed": 11}
    conf = {"dimensionality": 3, "base_path": str(tmp_path / "exp") + "/", "exp_name": "s",
            "integrator": {"setting": setting, "type": "uniform", "channels": ["SOLUTION"]},
            "export": [{"type": "image", "channel": "SOLUTION", "file_name": "solution"}],
            "scene": {"aabb": {"min": [float(x) for x in c - 2.0], "max": [float(x) for x in c + 2.0]},
                      "evaluation_grid": {"mData": {"pos": [float(x) for x in c], "scale": 1.2, "up": [0.0, 1.0, 0.0], "right": [1.0, 0.0, 0.0]}},
                      "source_path": NVDB, "source_intensity": 0.8,
                      "mesh": {"dirichlet_path": str(tmp_path / "model.obj"), "vertex_color_dirichlet_path": str(tmp_path / "color.json")}}}
    p = tmp_path / "s.json"
    p.write_text(json.dumps(conf))
    L = _lib.load()
    assert L.wost_run_config(str(p).encode()) == 0, L.wost_last_error()
    res = json.load(open(tmp_path / "exp" / "s" / "result.json"))
    prob = api.Problem(3, dirichlet=api.Scene(3, dv, dp), color_dirichlet=scenes.gray_pairs(scenes.harmonic_3d(dv - c.astype(np.float32))),
                       source=dense, source_origin=origin, source_voxel=voxel, source_intensity=0.8,
                       grid_scale=1.2, grid_pos=tuple(float(x) for x in c), grid_up=(0.0, 1.0, 0.0), grid_right=(1.0, 0.0, 0.0),
                       aabb_min=tuple(float(x) for x in c - 2.0), aabb_max=tuple(float(x) for x in c + 2.0))
    it = api.UniformIntegrator(prob, api.settings_from_json(setting, seed=11))
    it.solve()
    assert it.last_walk_steps == res["walk_steps"]
    it0 = api.UniformIntegrator(api.Problem(3, dirichlet=api.Scene(3, dv, dp), color_dirichlet=scenes.gray_pairs(scenes.harmonic_3d(dv - c.astype(np.float32))),
                                            grid_scale=1.2, grid_pos=tuple(float(x) for x in c), grid_up=(0.0, 1.0, 0.0), grid_right=(1.0, 0.0, 0.0)),
                                api.settings_from_json(setting, seed=11))
    it0.solve()
    assert np.abs(it.film()[..., :3] - it0.film()[..., :3]).max() > 1e-3   # the source term does contribute
