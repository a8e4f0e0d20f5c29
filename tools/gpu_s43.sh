cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "triangle_mesh" > gpurun_out/s43_tests.log 2>&1
tail -30 gpurun_out/s43_tests.log | cut -c1-400
