cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "height_map or triangle_mesh" > gpurun_out/s51_tests.log 2>&1
tail -30 gpurun_out/s51_tests.log | cut -c1-400
