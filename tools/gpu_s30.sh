cd $GRAFT_REPO_ROOT
python bench.py --scene c5 --scale 0.125 --steps 30 --warmup 5 --no-cpu-baseline --no-also > gpurun_out/s30_c5s.json 2> gpurun_out/s30_c5s.err
python tools/show_bench.py gpurun_out/s30_c5s.json
python bench.py --scene c2 --steps 30 --warmup 5 --no-cpu-baseline --no-also > gpurun_out/s30_c2.json 2> gpurun_out/s30_c2.err
python tools/show_bench.py gpurun_out/s30_c2.json
