cd $GRAFT_REPO_ROOT
timeout 900 python bench.py --scene c5 --warmup 5 --steps 60 --no-cpu-baseline --no-also > gpurun_out/s54_c5.json 2> gpurun_out/s54_c5.err
python tools/show_bench.py gpurun_out/s54_c5.json 2>/dev/null || tail -c 800 gpurun_out/s54_c5.err
python -m pytest tests/test_gpu_features.py tests/test_gpu_engine_module.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "slider or cone_twist or joint or sixdof or 6dof or custom_solver_bias or forces_and_ccd" > gpurun_out/s54_tests.log 2>&1
tail -12 gpurun_out/s54_tests.log | cut -c1-400
