cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider -k "reproducible or free_run or 64_dynamic or multi_space or sleep or each_space or engine or set_state or joint" > gpurun_out/s24_tests.log 2>&1
tail -8 gpurun_out/s24_tests.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/s24_bench.json 2> gpurun_out/s24_bench.err
tail -c 300 gpurun_out/s24_bench.err
