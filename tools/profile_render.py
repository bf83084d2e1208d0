"""Small render used under ncu / for quick timing: python tools/profile_render.py <scene> <variant> <size> <spp> <depth>"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rustraytracer_b200 import Film, IntType, render_samples, scenes  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "plastic_dragon"
variant, size, spp, depth, f32 = (int(x) for x in (sys.argv[2:7] + ["0", "512", "4", "8", "0"][len(sys.argv) - 2:]))
hs = scenes.HostScene(name, size, size, variant=variant)
gpu = hs.upload()
if os.environ.get("RT_WAVEFRONT"):
    gpu.set_wavefront_size(int(os.environ["RT_WAVEFRONT"]))
cam = hs.camera()
film = Film(size, size)
for rep in range(2):
    film.clear()
    render_samples(gpu, cam, IntType.Path(depth, shade_f32=bool(f32)), film, size, size, 0, spp, seed=rep)
    c = gpu.counters()
    rays = c["extend_rays"] + c["shadow_rays"] + c["probe_rays"]
    print(f"f32={f32} {name} v{variant} {size}^2 x{spp}: {c['render_ms']:.2f} ms {rays / c['render_ms'] / 1e3:.1f} Mrays/s  extend {c['extend_ms']:.2f} shade {c['shade_ms']:.2f} connect {c['connect_ms']:.2f} iters {c['iterations']}", flush=True)
