for cfg in "2 3 0" "3 3 0" "2 3 4" "2 3 5" "2 3 6" "2 3 7" "3 3 6" "3 3 5"; do set -- $cfg
 CVR_PATH_H_FWD=$1 CVR_PATH_H_REV=$2 CVR_H3_OCC=$3 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-side-shapes > gpurun_out/sw.json 2> gpurun_out/sw.err
 python -c "
import json
d=json.load(open('gpurun_out/sw.json')); s=d['stage_us_per_frame']
print('fwd=$1 rev=$2 occ=$3', round(d['value']), 'h_fwd', round(s['path_h_fwd'],1), 'h_rev', round(s['path_h_rev_wta'],1))"
done
