"""The BASELINE.json configurations as host-side workload specs, and their elaina-exec case directories.

`workload(name)` -> dict with the geometry / boundary data / evaluation grid / integrator settings of one configuration:

  c1        configs[0]: data/ladybug on its REAL geometry (61 476 Dirichlet segments + the 4-segment Neumann square,
            elaina_b200/data/scene_ladybug.npz), settings of data/ladybug/u.json (uniform) / n.json (guided): 1024^2, 1024 spp,
            depth 64, epsilonShell 1. The shipped color.json is missing from the reference snapshot: colours are
            synthesised per connected stroke (scenes.synth_stroke_colors, seeded).
  c2        configs[1]: data/fille real geometry (66 367 segments), depth 128, 4-level DenseGrid network (data/fille/n.json)
  c3        configs[2]: synthetic 2D Laplace on a 10 000-segment star polygon, g = Re(z^3), 1024^2, 256 spp
  c4        configs[3]: synthetic 3D Poisson on a 1 310 720-triangle displaced icosphere (c4-small: 327 680)
  c5        configs[4]: c3's geometry at 4096^2, 1024 spp (strong-scaling sweep)
  c1-synth  round-1 stand-in for c1 (seeded stroke soup of the same size)

`write_case(wl, out_dir, ...)` writes the same workload as an elaina-exec case (OBJ meshes, color.json, mask.pgm, conf.json
with the reference's keys, exec.cu:65-76 / core/problem.cu:158-181), runnable unchanged by the reference's elaina-exec and by
elaina_b200/elaina-exec.
"""
import json
import math
import os

import numpy as np

from . import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")  # only the NanoVDB test fixture of the "nvdb" demo workload below
DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")  # the package owns its inputs (data/make_scene_fixtures.py wrote them)

NETWORK_LADYBUG = {  # data/ladybug/n.json:49-81
    "encoding": {"otype": "DenseGrid", "n_levels": 8, "n_features_per_level": 4, "base_resolution": 8,
                 "per_level_scale": 1.4049999713897705, "interpolation": "Linear"},
    "loss": {"otype": "L2"},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 3},
    "optimizer": {"otype": "Ema", "decay": 0.949999988079071,
                  "nested": {"otype": "Adam", "learning_rate": 0.00800000037997961, "beta1": 0.8999999761581421, "beta2": 0.9900000095367432,
                             "epsilon": 1.0000000036274937e-15, "l2_reg": 9.999999974752427e-07, "adabound": False}},
}
NETWORK_FILLE = json.loads(json.dumps(NETWORK_LADYBUG))  # data/fille/n.json: 4 levels from base resolution 4
NETWORK_FILLE["encoding"].update(n_levels=4, base_resolution=4)


def _shipped(name):
    z = np.load(os.path.join(DATA, "scene_%s.npz" % name))
    return z["dverts"], z["dprims"], z["nverts"], z["nprims"]


def workload(name):
    if name in ("c3", "c3-small", "c5"):
        nseg = 10000
        dv, dp = scenes.star_polygon(nseg)
        W = {"c3": 1024, "c3-small": 256, "c5": 4096}[name]
        spp = {"c3": 256, "c3-small": 16, "c5": 1024}[name]
        diag = float(np.linalg.norm(np.array([3.0, 3.0])))
        # evaluate the interior problem only (that is where Re(z^3) is the solution): mask = pixels inside the curve
        g = 1.5 * (2.0 * np.arange(W, dtype=np.float64) / W - 1.0)
        X, Y = np.meshgrid(g, g)
        mask = (np.hypot(X, Y) < 1.0 + 0.3 * np.sin(5.0 * np.arctan2(Y, X)) - 4e-3).astype(np.uint8)
        return dict(name=name, dim=2, dverts=dv, dprims=dp, nverts=None, nprims=None, mask=mask,
                    color_dirichlet=scenes.gray_pairs(scenes.harmonic_2d(dv, 3)), color_neumann=None,
                    grid_scale=1.5, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0), aabb_min=(-1.5, -1.5), aabb_max=(1.5, 1.5),
                    frame=(W, W), depth=64, spp=spp, eps=1e-3 * diag, seed=42, network=NETWORK_LADYBUG, bytes_per_step=938,
                    desc="synthetic 2D Laplace, %d-segment star polygon, g=Re(z^3), %dx%d grid masked to the interior (%d active pixels), %d spp, depth 64" % (nseg, W, W, int(mask.sum()), spp))
    if name in ("c1", "c2"):
        scene = "ladybug" if name == "c1" else "fille"
        dv, dp, nv, npr = _shipped(scene)
        depth = 64 if name == "c1" else 128
        scale = 250.0 if name == "c1" else 300.0
        return dict(name=name, dim=2, dverts=dv, dprims=dp, nverts=nv, nprims=npr, mask=None,
                    color_dirichlet=scenes.synth_stroke_colors(len(dv), dp), color_neumann=None,
                    grid_scale=scale, grid_pos=(250.0, 250.0), grid_up=(-1.0, 0.0), aabb_min=(-100.0, -100.0), aabb_max=(600.0, 600.0),
                    frame=(1024, 1024), depth=depth, spp=1024, eps=1.0, seed=42, network=NETWORK_LADYBUG if name == "c1" else NETWORK_FILLE,
                    bytes_per_step=1850,
                    desc="data/%s real geometry (%d Dirichlet segments, 64-bit Morton path) inside the shipped 4-segment Neumann square, "
                         "synthesised per-stroke colours (color.json is missing from the reference snapshot), 1024x1024 grid, 1024 spp, depth %d, epsilonShell 1"
                         % (scene, len(dp), depth))
    if name == "c1-synth":
        dv, dp = scenes.stroke_soup(60, 1024)
        nv = np.array([[-90, -90], [590, -90], [590, 590], [-90, 590]], np.float32)
        npr = np.array([[0, 1], [1, 2], [2, 3], [3, 0]], np.int32)
        return dict(name=name, dim=2, dverts=dv, dprims=dp, nverts=nv, nprims=npr, color_dirichlet=scenes.stroke_colors(60, 1024), color_neumann=None,
                    grid_scale=250.0, grid_pos=(250.0, 250.0), grid_up=(-1.0, 0.0), aabb_min=(-100.0, -100.0), aabb_max=(600.0, 600.0),
                    frame=(1024, 1024), depth=64, spp=256, eps=1.0, seed=42, network=NETWORK_LADYBUG, bytes_per_step=1850,
                    desc="synthetic ladybug-shaped 2D mixed problem: 61440-segment Dirichlet stroke soup in a 4-segment Neumann square [-90,590]^2, 1024x1024 grid, 256 spp, depth 64, epsilonShell 1")
    if name in ("c4", "c4-small"):
        dv, dp = scenes.icosphere(8 if name == "c4" else 7, 1.0, 0.1)
        src = np.ones((16, 16, 16, 3), np.float32)
        g = 2.0 * np.arange(512, dtype=np.float64) / 512 - 1.0
        X, Y = np.meshgrid(g, g)
        mask = (np.hypot(X, Y) < 0.88).astype(np.uint8)
        return dict(name=name, dim=3, dverts=dv, dprims=dp, nverts=None, nprims=None, mask=mask,
                    color_dirichlet=scenes.gray_pairs(scenes.harmonic_3d(dv)), color_neumann=None,
                    source=src, source_origin=(-1.6, -1.6, -1.6), source_voxel=0.2,
                    grid_scale=1.0, grid_pos=(0.0, 0.0, 0.0), grid_up=(0.0, 1.0, 0.0), grid_right=(1.0, 0.0, 0.0),
                    aabb_min=(-1.5, -1.5, -1.5), aabb_max=(1.5, 1.5, 1.5),
                    frame=(512, 512), depth=64, spp=64, eps=2e-3, seed=42, network=NETWORK_LADYBUG, bytes_per_step=1734,
                    desc="synthetic 3D Poisson, %d-triangle displaced icosphere, 512x512 slice z=0 masked to the interior (%d active pixels), 64 spp, depth 64" % (len(dp), int(mask.sum())))
    if name in ("c4-slices", "c4-small-slices"):
        # configs[3] as BASELINE.json words it: a SET of z-slices of the 512^3 evaluation volume. The bench runs `nslices` of the 512
        # planes per step (evenly spaced over the part of the volume that cuts the mesh), each a 512^2 frame masked to the interior.
        wl = workload("c4" if name == "c4-slices" else "c4-small")
        nslices = 8
        g = 2.0 * np.arange(512, dtype=np.float64) / 512 - 1.0
        X, Y = np.meshgrid(g, g)
        planes = (np.arange(nslices) + 0.5) / nslices * 2.0 - 1.0  # z in (-1, 1)
        zs = 0.8 * planes
        slices = []
        for z in zs:
            r = math.sqrt(max(0.0, 0.88 * 0.88 - z * z))
            slices.append(dict(grid_pos=(0.0, 0.0, float(z)), mask=(np.hypot(X, Y) < r).astype(np.uint8)))
        active = int(sum(int(sl["mask"].sum()) for sl in slices))
        wl.update(name=name, slices=slices,
                  desc=wl["desc"].replace("512x512 slice z=0 masked to the interior (%d active pixels)" % int(wl["mask"].sum()),
                                          "%d of the 512 z-slices of the 512^3 evaluation volume per step (z = %s), each 512x512 masked to the interior (%d active pixels in total)"
                                          % (nslices, ", ".join("%.2f" % z for z in zs), active)))
        return wl
    if name.startswith("pin-"):
        return _pin_workload(name)
    raise SystemExit("unknown workload " + name)


NVDB_FIXTURE = os.path.join(GOLDEN, "source_vec3f.nvdb")  # world box x [-0.75, 2.75], y [29.5, 32.25], z [-0.25, 2.0], voxel 0.25
PIN_CASES = ("pin-dirichlet2d", "pin-mixed2d", "pin-ladybug", "pin-poisson2d", "pin-poisson3d", "pin-mixed3d")


def _box_mesh(lo, hi):
    """12 triangles of an axis-aligned box, outward orientation."""
    (x0, y0, z0), (x1, y1, z1) = lo, hi
    v = np.array([[x0, y0, z0], [x1, y0, z0], [x1, y1, z0], [x0, y1, z0], [x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1]], np.float32)
    f = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4], [2, 3, 7], [2, 7, 6], [1, 2, 6], [1, 6, 5], [3, 0, 4], [3, 4, 7]], np.int32)
    return v, f


def _pin_workload(name):
    """Small cases run through the reference's own CUDA build to pin the walker oracle (tests/golden/b1_<name>.npz,
    tools/b1_run.py --pin): every estimator of the uniform integrator is exercised with non-trivial data."""
    base = dict(name=name, nverts=None, nprims=None, mask=None, color_neumann=None, seed=0, network=NETWORK_LADYBUG, bytes_per_step=938)
    if name == "pin-dirichlet2d":
        wl = workload("c3-small")
        wl.update(name=name, frame=(128, 128), spp=8, mask=wl["mask"][::2, ::2].copy(), desc="pin: c3 geometry, 128x128, 8 spp")
        return wl
    if name == "pin-ladybug":
        wl = workload("c1")
        wl.update(name=name, frame=(256, 256), spp=2, desc="pin: data/ladybug real geometry, 256x256, 2 spp")
        return wl
    if name == "pin-mixed2d":
        dv, dp = scenes.star_polygon(400, 0.8, 0.25, 5)
        nv = np.array([[-1.4, -1.4], [1.4, -1.4], [1.4, 1.4], [-1.4, 1.4]], np.float32)
        npr = np.array([[0, 1], [1, 2], [2, 3], [3, 0]], np.int32)
        cn = np.random.RandomState(3).rand(4, 6).astype(np.float32) * 0.2
        return dict(base, dim=2, dverts=dv, dprims=dp, nverts=nv, nprims=npr, color_dirichlet=scenes.gray_pairs(scenes.harmonic_2d(dv, 2)), color_neumann=cn,
                    grid_scale=1.3, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0), aabb_min=(-1.5, -1.5), aabb_max=(1.5, 1.5),
                    frame=(128, 128), depth=32, spp=8, eps=2e-3, bytes_per_step=1850,
                    desc="pin: star-shaped Dirichlet curve inside a square Neumann box with non-zero flux, 128x128, 8 spp")
    if name == "pin-poisson2d":
        dv, dp = scenes.star_polygon(256, 0.8, 0.0, 5)
        dv = (dv + np.array([1.0, 30.8], np.float32)).astype(np.float32)
        return dict(base, dim=2, dverts=dv, dprims=dp, color_dirichlet=scenes.gray_pairs(0.05 * (dv[:, 0] - 1.0) + 0.1),
                    source_nvdb=NVDB_FIXTURE, grid_scale=0.85, grid_pos=(1.0, 30.8), grid_up=(0.0, 1.0), aabb_min=(0.0, 29.8), aabb_max=(2.0, 31.8),
                    frame=(128, 128), depth=32, spp=8, eps=2e-3,
                    desc="pin: 2D Poisson in a disk with the NanoVDB Vec3f source fixture, 128x128, 8 spp")
    if name in ("pin-poisson3d", "pin-mixed3d"):
        c = np.array([1.0, 30.8, 0.9], np.float32)
        dv, dp = scenes.icosphere(3, 0.8 if name == "pin-poisson3d" else 0.35, 0.0)
        dv = (dv + c).astype(np.float32)
        wl = dict(base, dim=3, dverts=dv, dprims=dp, color_dirichlet=scenes.gray_pairs(0.05 * (dv[:, 0] - 1.0) + 0.02 * (dv[:, 2] - 0.9) + 0.1),
                  source_nvdb=NVDB_FIXTURE, grid_scale=0.85, grid_pos=tuple(float(x) for x in c), grid_up=(0.0, 1.0, 0.0), grid_right=(1.0, 0.0, 0.0),
                  aabb_min=(0.0, 29.8, -0.1), aabb_max=(2.0, 31.8, 1.9), frame=(96, 96), depth=32, spp=4, eps=2e-3, bytes_per_step=1734,
                  desc="pin: 3D Poisson in a ball (1280 triangles) with the NanoVDB Vec3f source fixture, 96x96 slice, 4 spp")
        if name == "pin-mixed3d":
            nv, npr = _box_mesh(c - 0.9, c + 0.9)
            cn = np.random.RandomState(5).rand(len(nv), 6).astype(np.float32) * 0.2
            wl.update(nverts=nv, nprims=npr, color_neumann=cn, desc="pin: 3D Poisson between a Dirichlet sphere and a Neumann box with non-zero flux + source, 96x96 slice, 4 spp")
        return wl
    raise SystemExit("unknown workload " + name)


def problem_kwargs(wl):
    keys = ("color_dirichlet", "color_neumann", "mask", "source", "source_origin", "source_voxel", "grid_scale", "grid_pos", "grid_up",
            "grid_right", "aabb_min", "aabb_max")
    kw = {k: wl[k] for k in keys if wl.get(k) is not None}
    if wl.get("source_nvdb") and "source" not in kw:
        # core/problem.cu:135-149: the .nvdb Vec3f grid, flattened by the product's reader (csrc/nvdb.h) into the dense grid the ABI takes
        kw["source"], kw["source_origin"], kw["source_voxel"] = scenes.load_nvdb_source(wl["source_nvdb"])
    return kw


def guided_settings(wl):
    """data/{ladybug,fille}/n.json: trainSppCount = spp/4, guided depth 10 in both phases, uniformFraction 0.5 (learned selection)."""
    return dict(train_spp_count=max(1, wl["spp"] // 4), max_guided_depth_training=10, max_guided_depth_guiding=10,
                uniform_fraction_training=0.5, uniform_fraction_guiding=0.5)


def write_pgm(path, mask):
    """binary PGM of a (H, W) 0/1 mask, written so that after the reference's flipped load (core/problem.cu:222,
    loadImage(path, flip = true)) mask[pixelId = x + y*W] is the array's [y, x]."""
    m = (np.asarray(mask) != 0).astype(np.uint8) * 255
    h, w = m.shape
    with open(path, "wb") as f:
        f.write(b"P5\n%d %d\n255\n" % (w, h))
        f.write(m[::-1].tobytes())


def write_case(wl, out_dir, integrator="uniform", exp_name="run", spp=None, extra_setting=None, exports=("image", "raw")):
    """-> path of conf.json. Files: dirichlet.obj [neumann.obj] color_d.json [color_n.json] [mask.pgm] conf.json; outputs go to
    <out_dir>/<exp_name>/ (created here: the reference's run_expr opens conf.json in it without creating it, exec.cu:69-72).
    3D source grids are only understood by our runner (`source_dense` extension) and are skipped for the reference."""
    os.makedirs(os.path.join(out_dir, exp_name, "frames"), exist_ok=True)
    os.makedirs(os.path.join(out_dir, exp_name, "frames_time"), exist_ok=True)
    dim = wl["dim"]
    mesh = {}
    scenes.save_obj(os.path.join(out_dir, "dirichlet.obj"), wl["dverts"], wl["dprims"])
    mesh["dirichlet_path"] = os.path.join(out_dir, "dirichlet.obj")
    if wl.get("color_dirichlet") is not None:
        scenes.save_vertex_colors(os.path.join(out_dir, "color_d.json"), wl["color_dirichlet"])
        mesh["vertex_color_dirichlet_path"] = os.path.join(out_dir, "color_d.json")
    if wl.get("nverts") is not None:
        scenes.save_obj(os.path.join(out_dir, "neumann.obj"), wl["nverts"], wl["nprims"])
        mesh["neumann_path"] = os.path.join(out_dir, "neumann.obj")
        if wl.get("color_neumann") is not None:
            scenes.save_vertex_colors(os.path.join(out_dir, "color_n.json"), wl["color_neumann"])
            mesh["vertex_color_neumann_path"] = os.path.join(out_dir, "color_n.json")
    grid = {"pos": [float(x) for x in wl["grid_pos"]], "scale": float(wl["grid_scale"]), "up": [float(x) for x in wl["grid_up"]]}
    if dim == 3:
        grid["right"] = [float(x) for x in wl["grid_right"]]
    scene = {"aabb": {"min": [float(x) for x in wl["aabb_min"]], "max": [float(x) for x in wl["aabb_max"]]},
             "evaluation_grid": {"mData": grid}, "mesh": mesh}
    if wl.get("mask") is not None:
        write_pgm(os.path.join(out_dir, "mask.pgm"), wl["mask"])
        scene["mask_path"] = os.path.join(out_dir, "mask.pgm")
    if wl.get("source_nvdb"):
        scene["source_path"] = wl["source_nvdb"]
    elif wl.get("source") is not None:
        src = np.ascontiguousarray(wl["source"], np.float32)
        src.tofile(os.path.join(out_dir, "source.f32"))
        nz, ny, nx = src.shape[:3]
        scene["source_dense"] = {"path": os.path.join(out_dir, "source.f32"), "dims": [nx, ny, nz],
                                 "origin": [float(x) for x in wl["source_origin"]], "voxel": float(wl["source_voxel"])}
    W, H = wl["frame"]
    setting = {"debugPixel": 0, "frameSize": [W, H], "maxWalkingDepth": wl["depth"], "samplesPerPixel": int(spp or wl["spp"]),
               "saveSppMetricsDuration": -1, "saveSppMetricsUntil": -1, "saveTimeMetricsDuration": -1, "epsilonShell": float(wl["eps"]),
               "seed": int(wl["seed"])}
    conf = {"dimensionality": dim, "base_path": out_dir, "exp_name": exp_name,
            "integrator": {"setting": setting, "type": integrator, "channels": ["SOLUTION"]},
            "export": [{"type": t, "channel": "SOLUTION", "file_name": "solution"} for t in exports], "scene": scene}
    if integrator == "guided":
        g = guided_settings(dict(wl, spp=int(spp or wl["spp"])))
        setting.update(trainSppCount=g["train_spp_count"], maxGuidedDepthInTrainingPhase=g["max_guided_depth_training"],
                       maxGuidedDepthInGuidingPhase=g["max_guided_depth_guiding"], uniformFractionInTrainingPhase=g["uniform_fraction_training"],
                       uniformFractionInGuidingPhase=g["uniform_fraction_guiding"])
        conf["network"] = wl["network"]
    if extra_setting:
        setting.update(extra_setting)
    path = os.path.join(out_dir, "conf.json")
    with open(path, "w") as f:
        json.dump(conf, f, indent=1)
    return path
