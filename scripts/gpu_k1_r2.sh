mkdir -p gpurun_out
python - <<'PY'
import torch, time
h = torch.empty(11468800, dtype=torch.uint8).pin_memory(); d = torch.empty_like(h, device="cuda")
for _ in range(5): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(50): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 50
print("H2D 11.47 MB pinned: %.1f us -> %.1f GB/s" % (dt * 1e6, 11468800 / dt / 1e9))
x = torch.zeros(1, device="cuda"); torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(2000): x.add_(1)
torch.cuda.synchronize(); print("eager tiny-kernel launch rate: %.2f us per launch" % ((time.perf_counter() - t0) / 2000 * 1e6))
PY
nproc; uptime
for i in 1 2; do
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --others= 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('e2e %.3e' % d['e2e']['value'], d['e2e']['timing'])"
done
