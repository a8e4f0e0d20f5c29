cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "body_test_motion" > gpurun_out/s39_tests.log 2>&1
tail -40 gpurun_out/s39_tests.log | cut -c1-400
