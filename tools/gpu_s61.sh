cd $GRAFT_REPO_ROOT
for s in 0.125 0.25 0.5; do
timeout 600 python bench.py --scene c5 --scale $s --warmup 5 --steps 60 --no-cpu-baseline --no-also > gpurun_out/s61_$s.json 2> gpurun_out/s61_$s.err
python tools/show_bench.py gpurun_out/s61_$s.json 2>/dev/null || tail -c 500 gpurun_out/s61_$s.err
done
