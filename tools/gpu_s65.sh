cd $GRAFT_REPO_ROOT
run() { BPC_LIBRARY=$PWD/bradot_b200/$2 timeout 600 python bench.py --scene c5 --scale $3 --warmup 5 --steps 30 --no-cpu-baseline --no-also > gpurun_out/s65_$1_$3.json 2> gpurun_out/s65_$1_$3.err; python tools/show_bench.py gpurun_out/s65_$1_$3.json 2>/dev/null | grep -o "ms/step [0-9.]*\|'solve': [0-9.]*\|checksum [0-9.]*" | tr '\n' ' '; echo " $1 scale $3"; }
for s in 1.0 0.5 0.25; do
run default libbradot_physics_cuda.so $s
run B_bins_always libbpc_B.so $s
run A_small_smem libbpc_A.so $s
done
