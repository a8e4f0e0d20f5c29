cd $GRAFT_REPO_ROOT
run() { BPC_LIBRARY=$PWD/bradot_b200/$2 timeout 600 python bench.py --scene c5 --warmup 5 --steps 30 --no-cpu-baseline --no-also > gpurun_out/s55_$1.json 2> gpurun_out/s55_$1.err; python tools/show_bench.py gpurun_out/s55_$1.json 2>/dev/null | grep -o "ms/step [0-9.]*\|'solve': [0-9.]*" | tr '\n' ' '; echo " $1"; }
run full libbradot_physics_cuda.so
run A_noext libbpc_A.so
run B_noext_nobias libbpc_B.so
run full2 libbradot_physics_cuda.so
run A2 libbpc_A.so
run B2 libbpc_B.so
python -m pytest tests/test_gpu_features.py -q -m gpu --timeout 1500 -p no:cacheprovider -k "custom_solver_bias" > gpurun_out/s55_tests.log 2>&1
tail -5 gpurun_out/s55_tests.log | cut -c1-400
