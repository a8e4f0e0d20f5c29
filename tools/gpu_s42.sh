cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "convex or hull or chain" > gpurun_out/s42_tests.log 2>&1
tail -8 gpurun_out/s42_tests.log | cut -c1-300
