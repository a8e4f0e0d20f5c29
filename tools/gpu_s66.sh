cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --timeout 2400 -p no:cacheprovider > gpurun_out/s66_tests.log 2>&1
tail -8 gpurun_out/s66_tests.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s66_smoke.log 2>&1; tail -2 gpurun_out/s66_smoke.log
python bench.py > gpurun_out/s66_bench.json 2> gpurun_out/s66_bench.err
python tools/show_bench.py gpurun_out/s66_bench.json 2>/dev/null || tail -c 1500 gpurun_out/s66_bench.err
