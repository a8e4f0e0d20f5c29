cd $GRAFT_REPO_ROOT
for sc in c5 c2; do
timeout 600 python bench.py --scene $sc --warmup 5 --steps 20 --no-cpu-baseline --no-also > gpurun_out/s22_${sc}.json 2> gpurun_out/s22_${sc}.err
done
timeout 600 python bench.py --scene c3 --warmup 3 --steps 10 --no-cpu-baseline --no-also > gpurun_out/s22_c3.json 2> gpurun_out/s22_c3.err
