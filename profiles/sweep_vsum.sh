# tuning sweep of k_vsum: CVR_VSUM_OCC (blocks per SM, 0 = no cap) x CVR_VSUM_P (rows prefetched ahead)
for cfg in "0 16" "0 24" "0 32" "4 24" "3 32"; do set -- $cfg
 CVR_VSUM_OCC=$1 CVR_VSUM_P=$2 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-side-shapes > gpurun_out/sw.json 2> gpurun_out/sw.err
 python -c "
import json
d=json.load(open('gpurun_out/sw.json')); s=d['stage_us_per_frame']
print('occ=$1 P=$2', round(d['value']), 'vsum', round(s['vsum'],1))"
done
