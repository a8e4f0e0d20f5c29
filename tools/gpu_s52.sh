cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "continuous" > gpurun_out/s52_tests.log 2>&1
tail -30 gpurun_out/s52_tests.log | cut -c1-600
