cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider > gpurun_out/s13_tests.log 2>&1
tail -15 gpurun_out/s13_tests.log
( time python bench.py --steps 20 --warmup 5 > gpurun_out/s13_bench.json 2> gpurun_out/s13_bench.err ) 2> gpurun_out/s13_bench.time
( time python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/s13_ref.json 2> gpurun_out/s13_ref.err ) 2> gpurun_out/s13_ref.time
cat gpurun_out/s13_bench.time gpurun_out/s13_ref.time; tail -3 gpurun_out/s13_bench.err gpurun_out/s13_ref.err
