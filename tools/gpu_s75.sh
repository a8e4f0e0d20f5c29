cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 600 -p no:cacheprovider -k "committed_golden" > gpurun_out/s75_tests.log 2>&1
tail -8 gpurun_out/s75_tests.log | cut -c1-500
