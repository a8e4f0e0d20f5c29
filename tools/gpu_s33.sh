cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_engine_module.py tests/test_gpu_queries.py -q -m gpu --timeout 1500 -p no:cacheprovider > gpurun_out/s33_tests.log 2>&1
tail -30 gpurun_out/s33_tests.log | cut -c1-300
