cd $GRAFT_REPO_ROOT
P=gpurun_out/r02c
mkdir -p $P
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 9000 --launch-count 150 --csv --log-file $P/r02c_c2_settled_launches.csv python tools/profile_step.py --scene c2 --warmup 155 --steps 4 > $P/c2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 20000 --launch-count 200 --csv --log-file $P/r02c_c3_settled_launches.csv python tools/profile_step.py --scene c3 --warmup 400 --steps 4 > $P/c3.log 2>&1
ls -la $P
