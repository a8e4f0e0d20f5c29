cd $GRAFT_REPO_ROOT
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8700 --launch-count 160 --csv --log-file gpurun_out/s20_c5_launches.csv python tools/profile_step.py --scene c5 --warmup 125 --steps 4 > gpurun_out/s20_c5.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_traverse" -s 125 -c 1 -o gpurun_out/s20_c5_trav python tools/profile_step.py --scene c5 --warmup 121 --steps 2 > gpurun_out/s20_ncu.log 2>&1
