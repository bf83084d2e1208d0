"""Per-ray constants of a traversal kernel from an ncu metrics capture of EVERY launch of that kernel in one render, so that the ray mix (camera rays,
all bounces) is the bench's own:

    ncu --metrics <METRICS> --clock-control none -k regex:k_extend --launch-skip N -c N --csv --page raw --log-file ext.csv python tools/profile_render.py ...
    python tools/extend_work.py ext.csv <rays traced by those launches> k_extend profiles/extend_work_r02.json

rays = the render's extend_rays counter (printed by tools/profile_render.py).  Output: bytes the kernel's global loads request per ray (L1 sectors x 32 B),
L2->L1 sectors, DRAM bytes per ray, warp instructions and thread instructions per ray, lanes per instruction, and the captured launches' time.
bench.py multiplies these by its own live rays/s to get achieved GB/s and instruction issue rates against the measured peaks."""
import csv
import json
import sys

METRICS = ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,lts__t_sectors_srcunit_tex_op_read.sum,dram__bytes_read.sum,dram__bytes_write.sum,"
           "smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum")


def main():
    path, rays, kernel, out = sys.argv[1], float(sys.argv[2]), sys.argv[3], sys.argv[4]
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    data = [r for r in data if kernel in r[ki]]
    tot = {}
    for m in METRICS.split(","):
        if m in hdr:
            i = hdr.index(m)
            scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(units[i], 1.0)
            tot[m] = sum(float(r[i].replace(",", "")) for r in data) * scale
    ld = tot["l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"] * 32
    res = {"kernel": kernel, "launches_captured": len(data), "rays": rays, "ms_under_ncu": tot["gpu__time_duration.sum"],
           "load_bytes_per_ray": ld / rays, "l2_to_l1_bytes_per_ray": tot.get("lts__t_sectors_srcunit_tex_op_read.sum", 0.0) * 32 / rays,
           "dram_bytes_per_ray": (tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]) / rays,
           "warp_inst_per_ray": tot["smsp__inst_executed.sum"] / rays, "thread_inst_per_ray": tot["smsp__thread_inst_executed.sum"] / rays,
           "lanes_per_inst": tot["smsp__thread_inst_executed.sum"] / tot["smsp__inst_executed.sum"],
           "local_bytes_per_ray": (tot.get("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", 0) + tot.get("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum", 0)) * 32 / rays,
           "how": "ncu --metrics (tools/extend_work.py METRICS) --clock-control none over every launch of the kernel in one render of the bench's scene "
                  "(python tools/profile_render.py plastic_dragon 0 1024 8 8 1, second render); sums over the launches / the render's ray counter"}
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    if len(sys.argv) == 2 and sys.argv[1] == "--metrics":
        print(METRICS)
    else:
        main()
