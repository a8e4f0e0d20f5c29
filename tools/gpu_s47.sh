cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_edges_and_scale.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "round_2_entry" > gpurun_out/s47_tests.log 2>&1
tail -30 gpurun_out/s47_tests.log | cut -c1-300
