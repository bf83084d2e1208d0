#!/bin/bash
set -x
timeout 900 python tools/config_runs.py --out gpurun_out/configs.json > gpurun_out/configs.log 2>&1
timeout 1500 python tools/soup_sweep.py --sizes 1,2,5,10,20 --out gpurun_out/soup_sweep.json > gpurun_out/soup.log 2>&1
tail -n 12 gpurun_out/configs.log | cut -c1-400; tail -n 8 gpurun_out/soup.log | cut -c1-400
