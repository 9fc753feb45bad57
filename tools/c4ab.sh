# A/B runs of the config-4 search (51 PRN x 141 bins x 327 360 lags x 10 steps): each argument is an environment assignment list
i=0
for e in "$@"; do
env $e timeout 500 python bench.py --no-cpu --no-trk --c5-streams 0 --no-c3 --no-dropin --no-sweep --steps 3 > gpurun_out/c4_$i.json 2> gpurun_out/c4_$i.err; echo "rc=$? [$e]"; tail -1 gpurun_out/c4_$i.err
python - <<P
import json
d=json.loads([x for x in open('gpurun_out/c4_$i.json') if x.startswith('{')][-1])
c=d["strong_scaling_c4"]; print(c.get("kernel_share_this_rank")); print('  C4 %.1f G cells/s %.1f ms per search, found bins %s' % (c['cells_per_s']/1e9, c['ms_per_search'], c['found_bins'] == {k:[v, c['found_bins'][k][1]] for k,v in c['true_bins'].items()}))
P
i=$((i+1))
done
