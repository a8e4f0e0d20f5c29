cd $GRAFT_REPO_ROOT
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_solve -s 240 -c 2 -o gpurun_out/s6_c5_solve python tools/profile_step.py --scene c5 --warmup 120 --steps 2 > gpurun_out/s6_ncu.log 2>&1
tail -5 gpurun_out/s6_ncu.log
ls -la gpurun_out/*.ncu-rep
