"""Audit of the conservative f32 triangle prefilter (traverse.cuh:tri_prefilter) with a -DRT_PREFILTER_CHECK=1 build of the library
(RT_LIBRTCORE=.../librtcore_check.so): renders every built-in scene that holds triangles and reads the device-side counters.  In that build the exact
f64 test runs for EVERY triangle slot; a violation is a slot the prefilter would have skipped whose exact hit the traversal then used.
Prints one JSON object; `violations` must be 0."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rustraytracer_b200 import Film, IntType, render_samples, scenes  # noqa: E402
from rustraytracer_b200._native import check, lib  # noqa: E402


def stats():
    s = (C.c_uint64 * 4)()
    check(lib().rt_debug_prefilter_stats(s))
    return [int(x) for x in s]


rows, prev = [], stats()
for name, kw, depth, size, spp in (("plastic_dragon", {}, 8, 512, 8), ("plastic_dragon", {"variant": 1}, 8, 512, 4), ("two_dragons", {"variant": 1}, 16, 512, 4),
                                   ("cornell_box_statue", {}, 25, 512, 4), ("material_hdr", {"variant": 3}, 12, 256, 4), ("teapot_hdr", {}, 8, 256, 4)):
    hs = scenes.HostScene(name, size, size, **kw)
    for build in ("sah", "lbvh"):
        gpu = hs.upload(build=build)
        film = Film(size, size)
        render_samples(gpu, hs.camera(), IntType.Path(depth, shade_f32=True), film, size, size, 0, spp, seed=9)
        now = stats()
        d = [a - b for a, b in zip(now, prev)]; prev = now
        rows.append({"scene": name, **kw, "tree": build, "triangle_slot_tests": d[0], "skipped_by_prefilter": d[1], "exact_hits": d[2], "violations": d[3],
                     "skip_rate": d[1] / max(d[0], 1)})
tot = stats()
print(json.dumps({"build": os.environ.get("RT_LIBRTCORE", "default"), "triangle_slot_tests": tot[0], "skipped_by_prefilter": tot[1], "exact_hits": tot[2], "violations": tot[3],
                  "skip_rate": tot[1] / max(tot[0], 1), "rows": rows}, indent=1))
sys.exit(0 if tot[3] == 0 and tot[0] > 0 else 1)
