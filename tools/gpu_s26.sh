cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "contact_reporting" > gpurun_out/s26_tests.log 2>&1
tail -25 gpurun_out/s26_tests.log
