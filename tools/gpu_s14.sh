cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_features.py tests/test_gpu_scale_parity.py -q -m gpu --maxfail=30 --timeout 1500 -p no:cacheprovider -k "set_state or inactive or c4_layout or c5_layout" > gpurun_out/s14_tests.log 2>&1
tail -8 gpurun_out/s14_tests.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/s14_bench2.json 2> gpurun_out/s14_bench2.err
tail -c 1500 gpurun_out/s14_bench2.json
grep -i "error\|Traceback" gpurun_out/s14_bench2.err | head
