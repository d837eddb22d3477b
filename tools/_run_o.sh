mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_golden.py -m gpu -x -q > gpurun_out/r02o_pytest_nb.log 2>&1; echo "parity rc=$?"; tail -4 gpurun_out/r02o_pytest_nb.log
for wl in C2 C1; do
timeout 600 python tools/tune.py --stage frame --workload $wl --frames 40 --reps 30 --sweep ray_nb=0,1 --sweep ray_ctas=8,9,12 > gpurun_out/r02o_tune_frame_nb_$wl.txt 2> gpurun_out/r02o_tune_$wl.err; echo "tune $wl rc=$?"; cat gpurun_out/r02o_tune_frame_nb_$wl.txt
done
