"""The elaina-exec configuration surface (exec.cu:39-221 -> wost_run_config / elaina_b200/elaina-exec).
CPU: error behaviour that needs no GPU (missing file / key, unsupported values -> status codes, never exit or throw).
GPU: full runs of a uniform and a guided configuration written in the reference's JSON format; outputs are read back
with independent decoders (zlib PNG, hand-parsed uncompressed EXR) and compared with the film from the API."""
import json
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(path):
    from elaina_b200 import _lib
    L = _lib.load()
    rc = L.wost_run_config(str(path).encode())
    return rc, (L.wost_last_error() or b"").decode()


def test_missing_file_and_keys_are_status_codes(tmp_path):
    rc, msg = _run(tmp_path / "nope.json")
    assert rc == 3 and "does not exist" in msg                      # WOST_ERR_IO
    p = tmp_path / "c.json"
    p.write_text("{ not json")
    assert _run(p)[0] == 4                                          # WOST_ERR_CONFIG (parse error)
    p.write_text(json.dumps({"base_path": str(tmp_path), "exp_name": "x"}))
    rc, msg = _run(p)
    assert rc == 4 and "Key not found: dimensionality" in msg       # json_get_or_throw (core/common.h:167-170)
    conf = {"dimensionality": 4, "base_path": str(tmp_path), "exp_name": "x", "scene": {}, "integrator": {"type": "uniform", "setting": {}}}
    p.write_text(json.dumps(conf))
    rc, msg = _run(p)
    assert rc == 5 and "Unsupported dimensionality" in msg          # reference: exit(1), exec.cu:125-129
    conf["dimensionality"] = 2
    conf["integrator"]["type"] = "fancy"
    p.write_text(json.dumps(conf))
    rc, msg = _run(p)
    assert rc == 5 and "Unrecognized integrator type" in msg        # exec.cu:96-100
    assert os.path.exists(tmp_path / "x" / "conf.json")             # the configuration copy is written first (exec.cu:69-72)


def _read_png(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        assert zlib.crc32(typ + data) == struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])[0]
        if typ == b"IHDR":
            w, h, depth, color = struct.unpack(">IIBB", data[:10])
            assert (depth, color) == (8, 6)
        elif typ == b"IDAT":
            idat += data
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 4 * w)
    assert np.all(raw[:, 0] == 0)
    return raw[:, 1:].reshape(h, w, 4)


def _read_exr(path):
    b = open(path, "rb").read()
    assert struct.unpack("<II", b[:8]) == (20000630, 2)
    pos, attrs = 8, {}
    while b[pos] != 0:
        e = b.index(b"\0", pos)
        name = b[pos:e].decode()
        pos = e + 1
        e = b.index(b"\0", pos)
        typ = b[pos:e].decode()
        pos = e + 1
        size = struct.unpack("<i", b[pos:pos + 4])[0]
        attrs[name] = (typ, b[pos + 4:pos + 4 + size])
        pos += 4 + size
    pos += 1
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    assert attrs["compression"][1] == b"\0"
    names = [c[:1].decode() for c in attrs["channels"][1].split(b"\0") if len(c) == 1 and c.isalpha()]
    offs = struct.unpack("<%dQ" % h, b[pos:pos + 8 * h])
    img = np.zeros((h, w, 4), np.float32)
    for y in range(h):
        yy, n = struct.unpack("<ii", b[offs[y]:offs[y] + 8])
        line = np.frombuffer(b[offs[y] + 8:offs[y] + 8 + n], np.float16).reshape(4, w)  # A B G R
        img[yy, :, 3], img[yy, :, 2], img[yy, :, 1], img[yy, :, 0] = line[0], line[1], line[2], line[3]
    return img


def _write_scene(tmp_path, guided):
    from elaina_b200 import scenes
    dv, dp = scenes.star_polygon(400, 100.0, 30.0, 5)
    dv = dv + np.float32(250.0)
    nv = np.array([[50, 50], [450, 50], [450, 450], [50, 450]], np.float32)
    npr = np.array([[0, 1], [1, 2], [2, 3], [3, 0]], np.int32)
    scenes.save_obj(tmp_path / "model.obj", dv, dp)
    scenes.save_obj(tmp_path / "boundary.obj", nv, npr)
    cols = scenes.synth_stroke_colors(len(dv), dp, seed=3)
    scenes.save_vertex_colors(tmp_path / "color.json", cols)
    W = 48
    setting = {"debugPixel": 0, "frameSize": [W, W], "maxWalkingDepth": 24, "samplesPerPixel": 8, "saveSppMetricsDuration": -1,
               "saveSppMetricsUntil": -1, "saveTimeMetricsDuration": -1, "epsilonShell": 1, "seed": 7}
    if guided:
        setting.update({"trainSppCount": 4, "uniformFractionInGuidingPhase": 0.5, "uniformFractionInTrainingPhase": 0.5,
                        "maxGuidedDepthInGuidingPhase": 6, "maxGuidedDepthInTrainingPhase": 6})
    from elaina_b200 import api
    conf = {"dimensionality": 2, "base_path": str(tmp_path / "exp") + "/", "exp_name": "g" if guided else "u", "print_network": False,
            "integrator": {"setting": setting, "type": "guided" if guided else "uniform", "channels": ["SOLUTION", "DIRICHLET_SDF"]},
            "export": [{"type": "image", "channel": "SOLUTION", "file_name": "solution"},
                       {"type": "image", "channel": "DIRICHLET_SDF", "file_name": "sdf"},
                       {"type": "energy", "tone": "MATLAB_JET", "channel": "SOLUTION", "file_name": "solution_energy_jet"}],
            "network": api.DEFAULT_NETWORK,
            "scene": {"aabb": {"max": [500.0, 500.0], "min": [0.0, 0.0]},
                      "evaluation_grid": {"mData": {"pos": [250.0, 250.0], "scale": 200, "up": [-1.0, 0.0]}},
                      "mesh": {"dirichlet_path": str(tmp_path / "model.obj"), "vertex_color_dirichlet_path": str(tmp_path / "color.json"),
                               "neumann_path": str(tmp_path / "boundary.obj")}}}
    p = tmp_path / ("g.json" if guided else "u.json")
    p.write_text(json.dumps(conf, indent=4))
    return p, conf, (dv, dp, nv, npr, cols)


@pytest.mark.gpu
def test_uniform_config_run_matches_api(tmp_path):
    from elaina_b200 import api
    p, conf, (dv, dp, nv, npr, cols) = _write_scene(tmp_path, guided=False)
    exe = os.path.join(ROOT, "elaina_b200", "elaina-exec")
    r = subprocess.run([exe, str(p)], capture_output=True, text=True, timeout=300)  # the command line of the reference (main.cpp:10-18)
    assert r.returncode == 0, r.stderr
    out = tmp_path / "exp" / "u"
    res = json.load(open(out / "result.json"))
    assert "duration" in res and "timestamp" in res and res["walk_steps"] > 0
    assert json.load(open(out / "conf.json")) == conf
    prob = api.Problem(2, dirichlet=api.Scene(2, dv, dp), neumann=api.Scene(2, nv, npr), color_dirichlet=cols,
                       grid_scale=200.0, grid_pos=(250.0, 250.0), grid_up=(-1.0, 0.0), aabb_min=(0.0, 0.0), aabb_max=(500.0, 500.0))
    it = api.UniformIntegrator(prob, api.settings_from_json(conf["integrator"]["setting"], seed=7))
    it.solve()
    assert it.last_walk_steps == res["walk_steps"]
    film = it.film()
    want = film[..., :4] / film[..., 4:5]
    exr = _read_exr(out / "solution.exr")
    np.testing.assert_allclose(exr, want[::-1].astype(np.float16).astype(np.float32), rtol=0, atol=0)  # flipped, HALF
    png = _read_png(out / "solution.png")
    np.testing.assert_array_equal(png, np.clip((want[::-1] * 255).astype(np.int64), 0, 255).astype(np.uint8))
    it.renderDirichletSDF()
    sdf = it.film(api.CH_DIRICHLET_SDF)
    np.testing.assert_array_equal(_read_exr(out / "sdf.exr")[..., 0], sdf[::-1, :, 0].astype(np.float16).astype(np.float32))
    jet = _read_png(out / "solution_energy_jet.png")
    assert jet.shape == (48, 48, 4) and np.all(jet[..., 3] == 255) and jet[..., :3].std() > 10


@pytest.mark.gpu
def test_guided_config_run(tmp_path):
    from elaina_b200 import _lib
    p, conf, _ = _write_scene(tmp_path, guided=True)
    rc, msg = _run(p)
    assert rc == 0, msg
    out = tmp_path / "exp" / "g"
    res = json.load(open(out / "result.json"))
    assert res["walk_steps"] > 0 and res["walk_steps_per_sec"] > 0
    exr = _read_exr(out / "solution.exr")
    assert np.isfinite(exr).all() and exr[..., 3].min() == 1.0
    # same scene through the uniform integrator: the guided field is an unbiased estimate of the same solution
    pu, _, _ = _write_scene(tmp_path, guided=False)
    assert _run(pu)[0] == 0
    ref = _read_exr(tmp_path / "exp" / "u" / "solution.exr")
    assert abs(float(exr[..., :3].mean() - ref[..., :3].mean())) < 0.03


@pytest.mark.gpu
@pytest.mark.parametrize("guided", [False, True])
def test_progressive_frames(tmp_path, guided):
    """saveSppMetricsDuration / saveSppMetricsUntil / saveTimeMetricsDuration (uniform/integrator.cu:577-609):
    frames/<sampleId>.{exr,png} hold solution / (sampleId + 1) for sampleId % duration == 0 and sampleId < until;
    frames_time/<ms>.* likewise for the time metric. The last frame of a run equals a shorter run's solution."""
    from elaina_b200 import api
    p, conf, (dv, dp, nv, npr, cols) = _write_scene(tmp_path, guided=guided)
    conf["integrator"]["setting"].update({"saveSppMetricsDuration": 3, "saveSppMetricsUntil": 7, "saveTimeMetricsDuration": 4})
    p.write_text(json.dumps(conf, indent=4))
    rc, msg = _run(p)
    assert rc == 0, msg
    out = tmp_path / "exp" / ("g" if guided else "u")
    assert sorted(os.listdir(out / "frames")) == ["0.exr", "0.png", "3.exr", "3.png", "6.exr", "6.png"]   # 8 spp: ids 0, 3, 6 (< 7)
    assert len([f for f in os.listdir(out / "frames_time") if f.endswith(".exr")]) in (1, 2)         # ids 0 and 4 (same ms -> same name)
    f6 = _read_exr(out / "frames" / "6.exr")
    assert np.isfinite(f6).all() and f6[..., 3].min() == 1.0
    if not guided:
        # the frame after sample 6 is the 7-spp estimate of the same Philox streams
        prob = api.Problem(2, dirichlet=api.Scene(2, dv, dp), neumann=api.Scene(2, nv, npr), color_dirichlet=cols,
                           grid_scale=200.0, grid_pos=(250.0, 250.0), grid_up=(-1.0, 0.0), aabb_min=(0.0, 0.0), aabb_max=(500.0, 500.0))
        st = dict(conf["integrator"]["setting"], samplesPerPixel=7, saveSppMetricsDuration=-1, saveTimeMetricsDuration=-1)
        it = api.UniformIntegrator(prob, api.settings_from_json(st, seed=7))
        it.solve()
        film = it.film()
        want = (film[..., :4] / film[..., 4:5])[::-1]
        np.testing.assert_allclose(f6, want, rtol=2e-3, atol=2e-3)
        # and the callback surface of the API delivers the same frames
        got = []
        it2 = api.UniformIntegrator(prob, api.settings_from_json(conf["integrator"]["setting"], seed=7))
        it2.set_frame_callback(lambda kind, sid, ms, film: got.append((kind, sid, film)))
        it2.solve()
        assert [(k, s) for k, s, _ in got] == [(0, 0), (1, 0), (0, 3), (1, 4), (0, 6)]
        np.testing.assert_allclose((got[-1][2][..., :4] / got[-1][2][..., 4:5])[::-1], f6, rtol=2e-3, atol=2e-3)
