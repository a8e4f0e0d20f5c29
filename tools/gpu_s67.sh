cd $GRAFT_REPO_ROOT
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/s67_n$n.json 2> gpurun_out/s67_n$n.err
python tools/show_bench.py gpurun_out/s67_n$n.json 2>/dev/null || tail -c 800 gpurun_out/s67_n$n.err
done
