cd $GRAFT_REPO_ROOT
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8700 --launch-count 160 --csv --log-file gpurun_out/s17_c5_launches.csv python tools/profile_step.py --scene c5 --warmup 125 --steps 4 > gpurun_out/s17_c5.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 9000 --launch-count 150 --csv --log-file gpurun_out/s17_c2_launches.csv python tools/profile_step.py --scene c2 --warmup 155 --steps 4 > gpurun_out/s17_c2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 26000 --launch-count 150 --csv --log-file gpurun_out/s17_c3_launches.csv python tools/profile_step.py --scene c3 --warmup 400 --steps 4 > gpurun_out/s17_c3.log 2>&1
tail -2 gpurun_out/s17_c5.log gpurun_out/s17_c2.log gpurun_out/s17_c3.log
