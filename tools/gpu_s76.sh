cd $GRAFT_REPO_ROOT
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-also > gpurun_out/s76_bench.json 2> gpurun_out/s76_bench.err
python tools/show_bench.py gpurun_out/s76_bench.json 2>/dev/null | head -2 || tail -c 400 gpurun_out/s76_bench.err
