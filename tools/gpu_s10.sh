cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges_and_scale.py tests/test_gpu_features.py -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider > gpurun_out/s10_tests.log 2>&1
tail -12 gpurun_out/s10_tests.log
run() { # name lib env scene warm
  BPC_LIBRARY=$PWD/bradot_b200/$2 env $3 timeout 600 python bench.py --scene $4 --warmup $5 --steps 30 --no-cpu-baseline > gpurun_out/s10_$1.json 2> gpurun_out/s10_$1.err
}
run c5_bins libbradot_physics_cuda.so X=1 c5 120
run c5_nobins libbradot_physics_cuda.so BPC_NO_BINS=1 c5 120
run c5_A libbpc_A.so BPC_NO_BINS=1 c5 120
run c5_B libbpc_B.so BPC_NO_BINS=1 c5 120
run c2_def libbradot_physics_cuda.so BPC_NO_BINS=1 c2 150
run c2_A libbpc_A.so BPC_NO_BINS=1 c2 150
run c2_B libbpc_B.so BPC_NO_BINS=1 c2 150
cat gpurun_out/s10*.err | tail -5
