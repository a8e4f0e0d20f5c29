cd $GRAFT_REPO_ROOT
run() { # name lib env scene warm
  BPC_LIBRARY=$PWD/bradot_b200/$2 env $3 timeout 600 python bench.py --scene $4 --warmup $5 --steps 30 --no-cpu-baseline > gpurun_out/s8_$1.json 2> gpurun_out/s8_$1.err
}
run c5_bins64 libbradot_physics_cuda.so X=1 c5 120
run c5_bins128 libbpc_C.so X=1 c5 120
run c5_A libbpc_A.so BPC_NO_BINS=1 c5 120
run c5_B libbpc_B.so BPC_NO_BINS=1 c5 120
run c2_A libbpc_A.so BPC_NO_BINS=1 c2 150
run c2_B libbpc_B.so BPC_NO_BINS=1 c2 150
run c4_bins64 libbradot_physics_cuda.so X=1 c4 120
run c4_bins128 libbpc_C.so X=1 c4 120
cat gpurun_out/s8*.err | tail -5
