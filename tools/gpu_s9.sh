cd $GRAFT_REPO_ROOT
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_solve_bins -s 120 -c 1 -o gpurun_out/s9_c5_bins64 python tools/profile_step.py --scene c5 --warmup 120 --steps 2 > gpurun_out/s9_ncu.log 2>&1
tail -3 gpurun_out/s9_ncu.log
