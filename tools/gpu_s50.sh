cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "slider or cone_twist or joint or sixdof or 6dof" > gpurun_out/s50_tests.log 2>&1
tail -30 gpurun_out/s50_tests.log | cut -c1-400
