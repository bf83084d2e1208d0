#!/bin/bash
# SAH leaf-cost sweep (host builder knob RT_SAH_ISECT): smaller leaves = fewer f64 leaf tests, more f32 node steps
for c in 1.5 3 6 12 24; do
  echo "RT_SAH_ISECT=$c"
  RT_SAH_ISECT=$c RT_TRACE_BUILD=1 timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 2>&1 | tail -n 2
done > gpurun_out/sah_sweep.log 2>&1
cat gpurun_out/sah_sweep.log
