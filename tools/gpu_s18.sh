cd $GRAFT_REPO_ROOT
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_pair_setup|k_pair_test|k_traverse|k_color|k_match" -s 605 -c 5 -o gpurun_out/s18_c5_np python tools/profile_step.py --scene c5 --warmup 121 --steps 2 > gpurun_out/s18_ncu.log 2>&1
tail -3 gpurun_out/s18_ncu.log
