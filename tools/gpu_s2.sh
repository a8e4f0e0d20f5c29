cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py tests/test_gpu_scale_parity.py -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider > gpurun_out/s2_new_tests.log 2>&1
tail -40 gpurun_out/s2_new_tests.log
python -m pytest tests -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider --deselect tests/test_gpu_features.py --deselect tests/test_gpu_scale_parity.py > gpurun_out/s2_old_tests.log 2>&1
tail -15 gpurun_out/s2_old_tests.log
