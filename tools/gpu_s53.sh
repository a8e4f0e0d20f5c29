cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --timeout 2400 -p no:cacheprovider > gpurun_out/s53_tests.log 2>&1
tail -15 gpurun_out/s53_tests.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s53_smoke.log 2>&1; tail -3 gpurun_out/s53_smoke.log
python bench.py > gpurun_out/s53_bench.json 2> gpurun_out/s53_bench.err
python tools/show_bench.py gpurun_out/s53_bench.json 2>/dev/null || tail -c 1500 gpurun_out/s53_bench.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/s53_ref.json 2> gpurun_out/s53_ref.err; tail -c 600 gpurun_out/s53_ref.json
