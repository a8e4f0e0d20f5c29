cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "areas or several_shapes" > gpurun_out/s69_tests.log 2>&1
tail -15 gpurun_out/s69_tests.log | cut -c1-600
