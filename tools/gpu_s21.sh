cd $GRAFT_REPO_ROOT
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_traverse -s 121 -c 1 -o gpurun_out/s21_c5_trav python tools/profile_step.py --scene c5 --warmup 121 --steps 2 > gpurun_out/s21_ncu.log 2>&1
tail -2 gpurun_out/s21_ncu.log
