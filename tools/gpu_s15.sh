cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "inactive" 2>&1 | grep "^E  " | head -5
run() { env $2 timeout 600 python bench.py --scene c5 --settle 120 --warmup 5 --steps 20 --no-cpu-baseline --no-also --scale $3 > gpurun_out/s15_$1.json 2> gpurun_out/s15_$1.err; }
run h_bins X=1 0.5
run h_nobins BPC_NO_BINS=1 0.5
run q_bins X=1 0.25
run q_nobins BPC_NO_BINS=1 0.25
