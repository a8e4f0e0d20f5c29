cd $GRAFT_REPO_ROOT
P=gpurun_out/r02
mkdir -p $P
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8700 --launch-count 150 --csv --log-file $P/r02_c5_settled_launches.csv python tools/profile_step.py --scene c5 --warmup 125 --steps 4 > $P/c5.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 9000 --launch-count 150 --csv --log-file $P/r02_c2_settled_launches.csv python tools/profile_step.py --scene c2 --warmup 155 --steps 4 > $P/c2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 20000 --launch-count 200 --csv --log-file $P/r02_c3_settled_launches.csv python tools/profile_step.py --scene c3 --warmup 400 --steps 4 > $P/c3.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_sweeps|k_presolve" -s 250 -c 2 -o $P/r02_c5_solve python tools/profile_step.py --scene c5 --warmup 125 --steps 2 > $P/c5_solve.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_sweeps|k_presolve" -s 310 -c 2 -o $P/r02_c2_solve python tools/profile_step.py --scene c2 --warmup 155 --steps 2 > $P/c2_solve.log 2>&1
ls -la $P
