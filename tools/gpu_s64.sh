cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "disabled" > gpurun_out/s64_tests.log 2>&1
tail -25 gpurun_out/s64_tests.log | cut -c1-500
