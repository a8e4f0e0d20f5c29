cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider > gpurun_out/s27_engine.log 2>&1
tail -25 gpurun_out/s27_engine.log
python -m pytest tests -q -m gpu --timeout 2400 -p no:cacheprovider --deselect tests/test_gpu_engine_module.py > gpurun_out/s27_tests.log 2>&1
tail -15 gpurun_out/s27_tests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/s27_bench.json 2> gpurun_out/s27_bench.err
python tools/show_bench.py gpurun_out/s27_bench.json 2>/dev/null || tail -c 1500 gpurun_out/s27_bench.json
