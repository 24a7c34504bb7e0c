for LIB in libhoomd_b200.so libhoomd_b200_b128.so; do
echo "== $LIB"
HOOMD_B200_LIB=/root/repo/hoomd_b200/$LIB HOOMD_B200_PAIR_MODE=0 TPPS=1,2,4 python scratch/sweep.py lj64k 2000 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(d['workload'], 'tpp', d['tpp'], 'tps', round(d['tps']), 'us/step', round(d['us_per_step'],2), 'pair', round(d['pair_us'],2), 'rebuild', round(d['rebuild_us'],1))"
HOOMD_B200_LIB=/root/repo/hoomd_b200/$LIB HOOMD_B200_PAIR_MODE=0 TPPS=1,2 python scratch/sweep.py lj1m 400 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(d['workload'], 'tpp', d['tpp'], 'tps', round(d['tps']), 'us/step', round(d['us_per_step'],2), 'pair', round(d['pair_us'],2), 'rebuild', round(d['rebuild_us'],1))"
done
