cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_engine_module.py tests/test_gpu_features.py -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider > gpurun_out/s16_tests.log 2>&1
tail -30 gpurun_out/s16_tests.log
