cd $GRAFT_REPO_ROOT
python -m pytest tests -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider -k "reproducible or free_run or 64_dynamic or multi_space or sleep or each_space" > gpurun_out/s4_tests.log 2>&1
tail -30 gpurun_out/s4_tests.log
for sc in c5 c4; do
timeout 600 python bench.py --scene $sc --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s4_${sc}_bins.json 2> gpurun_out/s4_${sc}_bins.err
BPC_LIBRARY=$PWD/bradot_b200/libbpc_massshared.so timeout 600 python bench.py --scene $sc --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s4_${sc}_massshared.json 2> gpurun_out/s4_${sc}_massshared.err
done
timeout 600 python bench.py --scene c2 --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s4_c2_bins.json 2> gpurun_out/s4_c2_bins.err
cat gpurun_out/s4*.err | tail -5
