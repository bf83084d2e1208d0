#!/bin/bash
set -x
RT_TRACE_ITER=1 timeout 100 python tools/profile_render.py sphere_roughness 0 64 4 8 1 > gpurun_out/dbg_trace.log 2>&1
RT_TRACE_ITER=1 timeout 100 python tools/profile_render.py material_hdr 0 64 4 8 1 > gpurun_out/dbg_trace_inf.log 2>&1
timeout 300 compute-sanitizer --tool memcheck python tools/profile_render.py sphere_roughness 0 32 2 8 1 > gpurun_out/dbg_san.log 2>&1
tail -n 30 gpurun_out/dbg_trace.log; tail -n 12 gpurun_out/dbg_trace_inf.log; tail -n 30 gpurun_out/dbg_san.log
