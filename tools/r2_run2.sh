# Round 2, run 2: parity of the new search (tournament + deferral), A/B timing, instruction counts.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2b_pytest.txt 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2b_pytest.txt
B="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-model --no-adjacent"
summ() { python - "$1" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'ms/step %.3f'%d['ms_per_step'], {k:(round(v['ms'],4), round(v.get('frac',0),3)) for k,v in d['kernels'].items()}, d.get('verified'), d.get('neighbour_set_mismatches'))
except Exception as e:
    print(sys.argv[1], 'parse failed', e)
P
}
for v in tour notour; do
  for d in defer nodefer; do
    L=""; [ $v = notour ] && L="MVT_LIB=$PWD/build/libmvt_notour.so"
    D=""; [ $d = nodefer ] && D="MVT_NO_DEFER=1"
    env $L $D timeout 300 $B > gpurun_out/r2b_${v}_${d}.json 2> gpurun_out/r2b_${v}_${d}.err; summ gpurun_out/r2b_${v}_${d}.json
  done
done
for wl in droid dexycb; do timeout 300 $B --workload $wl > gpurun_out/r2b_$wl.json 2>gpurun_out/r2b_$wl.err; summ gpurun_out/r2b_$wl.json; done
timeout 300 $B --dtype bf16 > gpurun_out/r2b_bf16.json 2>gpurun_out/r2b_bf16.err; summ gpurun_out/r2b_bf16.json
# instruction counts of the fused kernel (and the hard kernel)
M=smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
Q="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-model --no-adjacent --no-distributions --no-verify"
timeout 300 ncu --metrics $M --clock-control none -k regex:knn_corr -s 8 -c 4 --csv --log-file gpurun_out/r2b_inst_tour.csv $Q > /dev/null 2>&1
MVT_LIB=$PWD/build/libmvt_notour.so MVT_NO_DEFER=1 timeout 300 ncu --metrics $M --clock-control none -k regex:knn_corr -s 8 -c 2 --csv --log-file gpurun_out/r2b_inst_notour.csv $Q > /dev/null 2>&1
grep -h "knn_corr" gpurun_out/r2b_inst_tour.csv | cut -d, -f5,13- | head -20
grep -h "knn_corr" gpurun_out/r2b_inst_notour.csv | cut -d, -f5,13- | head -8
