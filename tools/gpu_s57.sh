cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "pick_ray or mixed_pile or direct_c_abi" > gpurun_out/s57_tests.log 2>&1
tail -12 gpurun_out/s57_tests.log | cut -c1-400
