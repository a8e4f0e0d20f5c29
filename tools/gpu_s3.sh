cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py tests/test_gpu_scale_parity.py -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider -k "set_state or inactive or c4_layout or c5_layout or regrows" > gpurun_out/s3_tests.log 2>&1
tail -30 gpurun_out/s3_tests.log
python -m pytest tests -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider -k "reproducible or free_run or 64_dynamic or engine or multi_space or sleep" > gpurun_out/s3_tests2.log 2>&1
tail -30 gpurun_out/s3_tests2.log
for sc in c5 c2 c4; do
timeout 600 python bench.py --scene $sc --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s3_${sc}_bins.json 2> gpurun_out/s3_${sc}_bins.err
BPC_NO_BINS=1 timeout 600 python bench.py --scene $sc --warmup 120 --steps 30 --no-cpu-baseline > gpurun_out/s3_${sc}_nobins.json 2> gpurun_out/s3_${sc}_nobins.err
done
tail -3 gpurun_out/*.err
