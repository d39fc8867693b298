mkdir -p gpurun_out
for rep in 1 2 3; do for pdl in 1 0; do for B in 1024 2048 8192; do
echo "K2_PDL=$pdl B=$B: $(B200RL_K2_PDL=$pdl timeout 600 python bench.py --workload ppo --B $B --steps 50 --warmup 3 --repeats 10 --no-cpu-baseline --no-e2e --others= 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.5f' % d['ms_per_step'])")"
done; done; done
