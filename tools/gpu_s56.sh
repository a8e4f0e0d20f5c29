cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "simple_hinges" > gpurun_out/s56_tests.log 2>&1
tail -12 gpurun_out/s56_tests.log | cut -c1-400
