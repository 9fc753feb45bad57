for e in "A=1" "SDRB200_FWD_WALK=1"; do
env $e timeout 300 python bench.py --no-cpu --no-trk --c5-streams 0 --strong-prns 0 --no-dropin --no-sweep --steps 5 > gpurun_out/c3.json 2> gpurun_out/c3.err; echo "rc=$? [$e]"
python - <<P
import json
d=json.loads([x for x in open('gpurun_out/c3.json') if x.startswith('{')][-1])
c=d['glonass_c3']; print('  C3 %.1f G cells/s %.2f ms' % (c['cells_per_s']/1e9, c['ms_per_search']), ' C1', round(d['value']/1e9,1), {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
P
done
