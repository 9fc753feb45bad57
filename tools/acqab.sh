# A/B runs of the acquisition search kernel: each argument is an environment assignment list (quoted), e.g.
#   bash tools/acqab.sh "SDRB200_ACQ_ROWC=1" "SDRB200_ROWP_PAD=32768"
BA="--no-cpu --no-trk --c5-streams 0 --no-c3 --strong-prns 0 --no-dropin --no-sweep"
i=0
for e in "$@"; do
  env $e SDRB200_VERBOSE=1 python bench.py $BA > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err; echo "rc=$? [$e]"
  grep -h "k_rowp" gpurun_out/ab_$i.err | head -1
  python - <<P
import json
d=json.loads([x for x in open('gpurun_out/ab_$i.json') if x.startswith('{')][-1])
print('  value %.1f G cells/s  %.4f ms/step  e2e %.1f G  ' % (d['value']/1e9, d['ms_per_step'], d['e2e']['value']/1e9), {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
P
  i=$((i+1))
done
