cd $GRAFT_REPO_ROOT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/s63_n2.json 2> gpurun_out/s63_n2.err
python tools/show_bench.py gpurun_out/s63_n2.json 2>/dev/null || tail -c 1500 gpurun_out/s63_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/s63_n2_ref.json 2> gpurun_out/s63_n2_ref.err
tail -c 300 gpurun_out/s63_n2_ref.json
