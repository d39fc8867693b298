mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for cfg in "6 0 0" "3 0 0" "2 0 0" "1 0 0" "6 5 0" "3 5 0" "2 5 0" "1 5 0" "6 0 1" "3 0 1"; do
  set -- $cfg
  export B200RL_K1_G=$2; if [ "$2" = "0" ]; then unset B200RL_K1_G; fi
  export B200RL_K1_TILE=$3
  echo "streams=$1 G=$2 tile=$3: $(timeout 300 python bench.py --steps 200 --warmup 5 --repeats 20 --streams $1 --pipeline independent --no-cpu-baseline --no-e2e --others '' 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.5f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']), 'serial', {k: round(v,5) if isinstance(v,float) else None for k,v in d['config']['serial_single_stream'].items() if k in ('ms_per_step','frac')}, d['config']['timing']['min_ms'], d['config']['timing']['max_ms'])")"
done
