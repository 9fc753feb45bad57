# ncu --set full capture of the tracking kernel (one launch of 12 channels x 2 s), after a plain run of the same command
BA="--no-cpu --c5-streams 0 --no-c3 --strong-prns 0 --no-dropin --no-sweep --trk-seconds 2 --trk-steps 1 --steps 2 --warmup 1"
python bench.py $BA > gpurun_out/plain_trk.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_trk_cl -s 1 -c 1 -o gpurun_out/prof_trk -f python bench.py $BA > gpurun_out/ncu_trk.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_trk.log
