cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "areas" > gpurun_out/s35_tests.log 2>&1
tail -40 gpurun_out/s35_tests.log | cut -c1-400
