cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_edges_and_scale.py tests/test_gpu_queries.py tests/test_gpu_parity.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "pipelined or with_areas or bodies_added or applied_forces or forces_and_ccd" > gpurun_out/s59_tests.log 2>&1
tail -12 gpurun_out/s59_tests.log | cut -c1-400
timeout 900 python bench.py --scene c5 --warmup 5 --steps 60 --no-cpu-baseline --no-also > gpurun_out/s59_c5.json 2> gpurun_out/s59_c5.err
python tools/show_bench.py gpurun_out/s59_c5.json 2>/dev/null || tail -c 800 gpurun_out/s59_c5.err
