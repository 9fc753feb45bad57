# ncu --set full capture of one launch of the two kernels of the two-pass path at config-4 size, after a plain run
BA="--no-cpu --no-trk --c5-streams 0 --no-c3 --no-dropin --no-sweep --steps 2 --warmup 1"
python bench.py $BA > gpurun_out/plain_c4.log 2>&1 || { echo plain run failed; tail -3 gpurun_out/plain_c4.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:k_comb_fft -s 2 -c 1 -o gpurun_out/prof_c4_comb -f python bench.py $BA > gpurun_out/ncu_c4a.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_c4a.log
ncu --set full --clock-control none --import-source on -k k_row -s 2 -c 1 -o gpurun_out/prof_c4_row -f python bench.py $BA > gpurun_out/ncu_c4b.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_c4b.log
