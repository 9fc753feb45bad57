# A/B runs of the GLONASS (configs[2]) object: each argument is an environment assignment list
i=0
for e in "$@"; do
env $e timeout 300 python bench.py --no-cpu --no-trk --c5-streams 0 --strong-prns 0 --no-dropin --no-sweep --steps 5 > gpurun_out/c3_$i.json 2> gpurun_out/c3_$i.err; echo "rc=$? [$e]"; tail -2 gpurun_out/c3_$i.err
python - <<P
import json
d=json.loads([x for x in open('gpurun_out/c3_$i.json') if x.startswith('{')][-1])
c=d['glonass_c3']; print('  C3 %.1f G cells/s %.2f ms' % (c['cells_per_s']/1e9, c['ms_per_search']), c['kernel_ms_per_search'])
P
i=$((i+1))
done
