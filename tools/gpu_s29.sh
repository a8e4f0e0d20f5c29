cd $GRAFT_REPO_ROOT
python bench.py --scene c5 --scale 0.125 --steps 30 --warmup 5 --no-cpu-baseline --no-also > gpurun_out/s29_c5s.json 2> gpurun_out/s29_c5s.err
python tools/show_bench.py gpurun_out/s29_c5s.json
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8000 -c 400 --csv --log-file gpurun_out/s29_c5s_launches.csv python bench.py --scene c5 --scale 0.125 --steps 8 --warmup 2 --no-cpu-baseline --no-also > gpurun_out/s29_ncu.log 2>&1
tail -2 gpurun_out/s29_ncu.log
