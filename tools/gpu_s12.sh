cd $GRAFT_REPO_ROOT
run() { # name env scene warm extra
  env $2 timeout 600 python bench.py --scene $3 --warmup $4 --steps 30 --no-cpu-baseline $5 > gpurun_out/s12_$1.json 2> gpurun_out/s12_$1.err
}
run c5_bins X=1 c5 120
run c5_nobins BPC_NO_BINS=1 c5 120
run c5s_bins X=1 c5 120 "--scale 0.125"
run c5s_nobins BPC_NO_BINS=1 c5 120 "--scale 0.125"
run c2_nobins BPC_NO_BINS=1 c2 150
cat gpurun_out/s12*.err | tail -5
