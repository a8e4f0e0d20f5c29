cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --timeout 2400 -p no:cacheprovider > gpurun_out/s60_tests.log 2>&1
tail -15 gpurun_out/s60_tests.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s60_smoke.log 2>&1; tail -3 gpurun_out/s60_smoke.log
python bench.py > gpurun_out/s60_bench.json 2> gpurun_out/s60_bench.err
python tools/show_bench.py gpurun_out/s60_bench.json 2>/dev/null || tail -c 1500 gpurun_out/s60_bench.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/s60_ref.json 2> gpurun_out/s60_ref.err; tail -c 300 gpurun_out/s60_ref.json
P=gpurun_out/r02c
mkdir -p $P
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8700 --launch-count 150 --csv --log-file $P/r02c_c5_settled_launches.csv python tools/profile_step.py --scene c5 --warmup 125 --steps 4 > $P/c5.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"k_presolve|k_sweeps" -s 240 -c 2 -o $P/r02c_c5_solve python tools/profile_step.py --scene c5 --warmup 120 --steps 2 > $P/c5_full.log 2>&1
tail -3 $P/c5_full.log
ls -la $P
