cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_queries.py tests/test_gpu_engine_module.py tests/test_gpu_features.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "with_areas or pick_ray or areas or mixed_pile" > gpurun_out/s58_tests.log 2>&1
tail -25 gpurun_out/s58_tests.log | cut -c1-400
