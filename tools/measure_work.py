"""Offline measurement of the algorithmic work per extend ray with the ORACLE's ordered reference
traversal, on a bounded sample of the bench workload.  Test-side tooling: writes profiles/algorithmic_work_r01.json,
which bench.py reads (bench.py itself never calls the oracle for the roofline numerator)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc  # noqa: E402
from rustraytracer_b200 import scenes  # noqa: E402

out = {}
for name, variant, depth in [("plastic_dragon", 0, 8), ("plastic_dragon", 1, 8)]:
    hs = scenes.HostScene(name, 1024, 1024, variant=variant)
    o = orc.OracleScene(hs.tables())
    film, c, secs = o.render_ex(hs.camera(), 1024, 1024, 0, 1, seed=0, max_depth=depth, threads=os.cpu_count(), pixel_stride=17, count_ordered=True)
    c = [int(x) for x in c]
    out[f"{name}_v{variant}"] = dict(paths=int(film[..., 3].sum()), extend=c[0], shadow=c[1], probe=c[2], box_tests=c[3], prim_tests=c[4],
                                     n_box=c[3] / c[0], n_tri=c[4] / c[0], secs=secs)
    print(name, variant, out[f"{name}_v{variant}"])
json.dump(out, open(os.path.join(ROOT, "profiles", "algorithmic_work_r01.json"), "w"), indent=1)
