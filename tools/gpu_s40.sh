cd $GRAFT_REPO_ROOT
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-also > gpurun_out/s40_bench.json 2> gpurun_out/s40_bench.err
python tools/show_bench.py gpurun_out/s40_bench.json 2>/dev/null || tail -c 1500 gpurun_out/s40_bench.err
