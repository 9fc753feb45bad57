for c in 4 2 1 8; do
BENCH_C5_CLUSTER=$c timeout 300 python bench.py --no-cpu --no-trk --no-c3 --strong-prns 0 --no-dropin --no-sweep --steps 3 > gpurun_out/c5_$c.json 2> gpurun_out/c5_$c.err; echo "rc=$? cluster $c"
python - <<P
import json
d=json.loads([x for x in open('gpurun_out/c5_$c.json') if x.startswith('{')][-1])
c=d['streams_c5']; print('  full %.1f ms  trk %.0f channel.s/s (%.2f us/period) cluster %s' % (c['full_job']['ms'], c['tracking']['value'], c['tracking']['us_per_code_period_per_channel'], c['tracking']['cluster_size']))
P
done
