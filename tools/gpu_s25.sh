cd $GRAFT_REPO_ROOT
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8700 --launch-count 150 --csv --log-file gpurun_out/s25_c5s_launches.csv python tools/profile_step.py --scene c5 --scale 0.125 --warmup 125 --steps 4 > gpurun_out/s25.log 2>&1
tail -2 gpurun_out/s25.log
