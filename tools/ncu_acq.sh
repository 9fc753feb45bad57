# ncu --set full capture of the acquisition search kernel (one launch), after a plain run of the same command
BA="--no-cpu --no-trk --c5-streams 0 --no-c3 --strong-prns 0 --no-dropin --no-sweep --steps 2 --warmup 1"
PAT=${1:-k_rowp}
python bench.py $BA > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$PAT -s 2 -c 1 -o gpurun_out/prof_acq -f python bench.py $BA > gpurun_out/ncu.log 2>&1
echo rc=$?; tail -3 gpurun_out/ncu.log
