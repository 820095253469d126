"""Scene ingest timing: wost_scene_create on a 1.3M-triangle icosphere, device vs host silhouette adjacency (row N1)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elaina_b200 import api, scenes
dv, dp = scenes.icosphere(8, 1.0, 0.1)
for mode in ("device", "host", "device", "host"):
    if mode == "host":
        os.environ["WOST_HOST_SILHOUETTES"] = "1"
    else:
        os.environ.pop("WOST_HOST_SILHOUETTES", None)
    t = time.perf_counter()
    s = api.Scene(3, dv, dp)
    print("%s adjacency: wost_scene_create of %d triangles %.1f ms" % (mode, len(dp), 1e3 * (time.perf_counter() - t)), flush=True)
    s.close() if hasattr(s, "close") else None
