set -x
nvidia-smi -L; nproc
cd $GRAFT_REPO_ROOT
timeout 600 python bench.py --scene c5 --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s1_c5_w120.json 2> gpurun_out/s1_c5_w120.err
timeout 600 python bench.py --scene c5 --warmup 400 --steps 30 --no-cpu-baseline > gpurun_out/s1_c5_w400.json 2> gpurun_out/s1_c5_w400.err
timeout 300 python bench.py --scene c2 --warmup 150 --steps 30 --no-cpu-baseline > gpurun_out/s1_c2_w150.json 2> gpurun_out/s1_c2_w150.err
timeout 600 python bench.py --scene c3 --warmup 400 --steps 10 --no-cpu-baseline > gpurun_out/s1_c3_w400.json 2> gpurun_out/s1_c3_w400.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s1_c5_launches.csv python tools/profile_step.py --scene c5 --warmup 120 --steps 3 > gpurun_out/s1_ncu.log 2>&1
ls -la gpurun_out
