cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "triangle_meshes or height_maps" > gpurun_out/s41_tests.log 2>&1
tail -40 gpurun_out/s41_tests.log | cut -c1-400
