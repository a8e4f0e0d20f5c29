cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "with_areas or shape or direct_c_abi or queries" > gpurun_out/s70_tests.log 2>&1
tail -15 gpurun_out/s70_tests.log | cut -c1-600
