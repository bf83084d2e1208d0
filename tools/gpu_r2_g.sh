#!/bin/bash
# round 2, GPU pass G: per-ray constants of k_extend / k_connect (ncu metrics over every launch of one render), full-set captures of one launch each of
# extend / connect / shade, launch list of the bench command, dragon batch rates, soup HBM roofline at 10 M triangles
set -x
M=$(python tools/extend_work.py --metrics)
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
timeout 200 $R > gpurun_out/prof_plain.log 2>&1; cat gpurun_out/prof_plain.log
RT_TRACE_ITER=1 timeout 200 $R > gpurun_out/trace_iter.log 2>&1
# the second render's launches: every render here is 10 iterations -> 10 k_extend / k_connect launches
timeout 900 ncu --metrics $M --clock-control none -k regex:k_extend --launch-skip 10 -c 10 --csv --page raw --log-file gpurun_out/extend_metrics.csv $R > gpurun_out/ncu_em.log 2>&1
timeout 900 ncu --metrics $M --clock-control none -k regex:k_connect --launch-skip 10 -c 10 --csv --page raw --log-file gpurun_out/connect_metrics.csv $R > gpurun_out/ncu_cm.log 2>&1
timeout 900 ncu --metrics $M --clock-control none -k regex:k_shade --launch-skip 20 -c 20 --csv --page raw --log-file gpurun_out/shade_metrics.csv $R > gpurun_out/ncu_sm.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 11 -c 1 -o gpurun_out/extend_full_r02 -f $R > gpurun_out/ncu_e.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 10 -c 1 -o gpurun_out/connect_full_r02 -f $R > gpurun_out/ncu_c.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 20 -c 2 -o gpurun_out/shade_full_r02 -f $R > gpurun_out/ncu_s.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
timeout 600 python tools/dragon_batches.py --out gpurun_out/dragon_batches.json > gpurun_out/dragon_batches.log 2>&1; cat gpurun_out/dragon_batches.log
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_closest_persistent --csv --page raw --log-file gpurun_out/soup10_metrics.csv python tools/soup_sweep.py --sizes 10 > gpurun_out/soup10_ncu.log 2>&1
RT_BUILD=lbvh timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_closest_persistent --csv --page raw --log-file gpurun_out/soup50_metrics.csv python tools/soup_sweep.py --sizes 50 > gpurun_out/soup50_ncu.log 2>&1
ls -la gpurun_out | head -40; du -sh gpurun_out
