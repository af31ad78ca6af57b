# cluster size x batch sweep for the vertical path kernel (tuning record; CVR_PATHV_CL forces the cluster size)
for cl in 8 6 5; do
 for b in 60 66 75 78 88 104; do
  CVR_PATHV_CL=$cl timeout 300 python bench.py --exact-batch --batch $b --steps 5 --warmup 3 --no-cpu-baseline 2> /dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l)
    except Exception: continue
    st=j['stage_ms_per_step']; B=$b
    print('CL=$cl B=$b fps=%.0f e2e=%.0f  us/frame: '%(j['value'], j['e2e']['value']) + ' '.join('%s=%.1f'%(k,v*1000/B) for k,v in st.items()))
"
 done
done
