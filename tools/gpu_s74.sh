cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py -q -m gpu --timeout 600 -p no:cacheprovider -k "intersect_rays or with_areas or pick" > gpurun_out/s74_tests.log 2>&1
tail -6 gpurun_out/s74_tests.log | cut -c1-400
