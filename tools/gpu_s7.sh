cd $GRAFT_REPO_ROOT
BPC_NO_BINS=1 timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_solve -s 120 -c 1 -o gpurun_out/s7_c5_solve_nobins python tools/profile_step.py --scene c5 --warmup 120 --steps 2 > gpurun_out/s7_ncu.log 2>&1
BPC_NO_BINS=1 timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_solve -s 150 -c 1 -o gpurun_out/s7_c2_solve_nobins python tools/profile_step.py --scene c2 --warmup 150 --steps 2 >> gpurun_out/s7_ncu.log 2>&1
tail -5 gpurun_out/s7_ncu.log
