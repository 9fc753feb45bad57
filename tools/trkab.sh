# A/B runs of the tracking kernel: each argument is an environment assignment list, e.g.  bash tools/trkab.sh "A=1" "SDRB200_TRK_SMALL=1"
BA="--no-cpu --c5-streams 0 --no-c3 --strong-prns 0 --no-dropin --trk-steps 1 --steps 3"
i=0
for e in "$@"; do
  env $e SDRB200_VERBOSE=1 python bench.py $BA > gpurun_out/tab_$i.json 2> gpurun_out/tab_$i.err; echo "rc=$? [$e]"
  grep -h "k_trk_cl" gpurun_out/tab_$i.err | sort | uniq -c | head -8
  python - <<P
import json
d=json.loads([x for x in open('gpurun_out/tab_$i.json') if x.startswith('{')][-1])
t=d['tracking']
print('  C2 %.0f channel.s/s  %.2f us/period' % (t['value'], t['us_per_code_period']), [(p['channels'], round(p['channel_s_per_s'])) for p in t.get('channel_sweep',{}).get('points',[])])
P
  i=$((i+1))
done
