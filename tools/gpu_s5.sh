cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider -k "reproducible or free_run or 64_dynamic or multi_space or sleep or each_space" > gpurun_out/s5_tests.log 2>&1
tail -5 gpurun_out/s5_tests.log
for sc in c5 c2; do
timeout 600 python bench.py --scene $sc --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s5_${sc}_bins.json 2> gpurun_out/s5_${sc}_bins.err
BPC_NO_BINS=1 timeout 600 python bench.py --scene $sc --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s5_${sc}_nobins.json 2> gpurun_out/s5_${sc}_nobins.err
done
timeout 600 python bench.py --scene c3 --warmup 400 --steps 10 --no-cpu-baseline > gpurun_out/s5_c3_bins.json 2> gpurun_out/s5_c3_bins.err
timeout 600 python bench.py --scene c4 --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s5_c4_bins.json 2> gpurun_out/s5_c4_bins.err
cat gpurun_out/s5*.err | tail -5
