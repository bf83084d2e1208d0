"""python tools/film_dump.py <scene> <variant> <size> <spp> <depth> <f32> <out.npy> — render and save the film (A/B of two builds: RT_LIBRTCORE selects the library)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from rustraytracer_b200 import Film, IntType, render_samples, scenes  # noqa: E402

name = sys.argv[1]
variant, size, spp, depth, f32 = (int(x) for x in sys.argv[2:7])
hs = scenes.HostScene(name, size, size, variant=variant)
gpu = hs.upload()
film = Film(size, size)
render_samples(gpu, hs.camera(), IntType.Path(depth, shade_f32=bool(f32)), film, size, size, 0, spp, seed=5)
np.save(sys.argv[7], film.read())
