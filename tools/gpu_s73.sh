cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "triangle_mesh or mixed_pile_boxes or matches_direct_c_abi or edge_cases" > gpurun_out/s73_tests.log 2>&1
tail -15 gpurun_out/s73_tests.log | cut -c1-600
