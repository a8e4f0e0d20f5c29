cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --timeout 2400 -p no:cacheprovider > gpurun_out/s37_tests.log 2>&1
tail -15 gpurun_out/s37_tests.log | cut -c1-300
python bench.py --steps 20 --warmup 5 > gpurun_out/s37_bench.json 2> gpurun_out/s37_bench.err
python tools/show_bench.py gpurun_out/s37_bench.json 2>/dev/null || tail -c 1500 gpurun_out/s37_bench.err
