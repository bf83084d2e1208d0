"""Multi-GPU host logic on CPU: the sample-range partition bench.py uses and a world-size-2 gloo all-reduce of
the (r,g,b,n) film — the only collective on the path.  The per-rank "render" here is the oracle
(test infrastructure), which shares the RNG keying with the CUDA path: the union of the ranks' sample ranges must
equal one render of all samples."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
    import numpy as np, torch, torch.distributed as dist
    import orc
    from rustraytracer_b200 import scenes
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    w = h = 24; spp = 2
    hs = scenes.HostScene("sphere_roughness", w, h)
    oracle = orc.OracleScene(hs.tables())
    cam = hs.camera()
    step = 0
    base = (step * world + rank) * spp                      # bench.py's partition of the global sample space
    film, _, _ = oracle.render(cam, w, h, base, base + spp, seed=9, threads=2)
    t = torch.from_numpy(film.reshape(-1).copy())
    dist.all_reduce(t)                                      # sum of the accumulation buffers
    if rank == 0:
        full, _, _ = oracle.render(cam, w, h, 0, world * spp, seed=9, threads=2)
        got = t.numpy().reshape(h, w, 4)
        assert np.array_equal(got[..., 3], full[..., 3]) and np.all(full[..., 3] == world * spp)
        assert np.allclose(got, full, rtol=1e-12, atol=1e-12)
        print("PARTITION_OK")
    dist.destroy_process_group()
""") % (ROOT, ROOT)


def test_sample_range_partition_allreduce_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29517", str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    assert "PARTITION_OK" in out.stdout
