cd $GRAFT_REPO_ROOT
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_solve_bins -s 122 -c 1 -o gpurun_out/s31_bins python tools/profile_step.py --scene c5 --scale 0.125 --warmup 121 --steps 3 > gpurun_out/s31_ncu.log 2>&1
tail -2 gpurun_out/s31_ncu.log
