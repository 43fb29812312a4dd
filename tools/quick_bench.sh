#!/bin/bash
# bench.py without the CPU baseline, printing the headline numbers and the per-group kernel times
python bench.py --no-cpu-baseline --configs humanoid "$@" 2>gpurun_out/qb.err | python -c "
import json,sys
o=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.3fM  ms/step %.4f  e2e %.3fM' % (o['value']/1e6, o['ms_per_step'], o['e2e']['value']/1e6))
print({k:round(v['ms']*1e3,1) for k,v in o['roofline']['groups'].items()}, 'sum', round(o['roofline']['step']['kernels_ms']*1e3,1))
"
