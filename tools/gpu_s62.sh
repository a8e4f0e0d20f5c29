cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider > gpurun_out/s62_tests.log 2>&1
tail -25 gpurun_out/s62_tests.log | cut -c1-500
