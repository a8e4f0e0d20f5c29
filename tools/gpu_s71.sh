cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "closest_point" > gpurun_out/s71_tests.log 2>&1
tail -15 gpurun_out/s71_tests.log | cut -c1-600
