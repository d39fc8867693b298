"""K8 check under torchrun (one process per GPU): the fused peer-memory allreduce + clip + Adam against an explicit
rank-ordered sum followed by K6, bit for bit, plus device timings of both routes.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 scripts/peer_adam_check.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from impala_b200 import optim, parallel  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    results = {}
    timed_sizes, check_only = (630_352, 1001, 3_000_000), (9, 4099, 70_001)
    for route, n in [(r, n) for r in ("one_shot", "two_shot", "nvls") for n in timed_sizes + check_only]:
        # the reference model, a ragged tiny one, a streaming-path one (timed); fewer elements than one CTA / than the
        # ranks' two-shot slices, sizes that are no multiple of 4 (checked only); every exchange route
        os.environ["B200RL_K8_TWO_SHOT"] = "0" if route == "one_shot" else "1"
        os.environ["B200RL_K8_NVLS"] = "1" if route == "nvls" else "0"
        os.environ["B200RL_PEER_TRANSPORT"] = "auto" if route == "nvls" else "ipc"
        torch.manual_seed(0)
        pa = [torch.nn.Parameter(torch.randn(n - 7, device=dev)), torch.nn.Parameter(torch.randn(7, device=dev))]
        pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
        opt_a = optim.FusedClipAdam(pa, lr=1e-2, eps=1e-5)
        opt_b = optim.FusedClipAdam(pb, lr=1e-2, eps=1e-5)
        bucket = parallel.PeerGradBucket(pa, n_extra=8)
        nvls = route == "nvls" and bucket.mc_ptr != 0          # the switch's summation order is its own: compare to rounding
        if route == "nvls" and not nvls and rank == 0:
            print("no multicast address (transport %s): the nvls rows below are the unicast two-shot route" % bucket.transport)
        ref = parallel.GradBucket(pb, n_extra=8)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        for step in range(4):
            bucket.attach()
            ref.attach()
            local = torch.randn(n + 8, device=dev, generator=g) * (0.01 if step % 2 else 1.0)   # clipped and unclipped steps
            bucket.flat.copy_(local)
            gathered = [torch.empty_like(local) for _ in range(world)]
            dist.all_gather(gathered, local)
            want = gathered[0].clone()
            for r in range(1, world):
                want += gathered[r]                                  # rank order, as the kernel adds
            ref.flat.copy_(want)
            norm_b = opt_b.step_flat(ref.grads, 0.5).clone()
            norm_a = opt_a.step_peer(bucket, 0.5).clone()
            torch.cuda.synchronize()
            if nvls:
                torch.testing.assert_close(norm_a, norm_b, rtol=1e-5, atol=1e-6)
                torch.testing.assert_close(bucket.reduced, ref.grads, rtol=1e-5, atol=1e-6)
                torch.testing.assert_close(opt_a.flat, opt_b.flat, rtol=1e-5, atol=1e-5)
                opt_b.flat.copy_(opt_a.flat); opt_b.exp_avg.copy_(opt_a.exp_avg); opt_b.exp_avg_sq.copy_(opt_a.exp_avg_sq)
            else:
                assert torch.equal(norm_a, norm_b), (n, step, norm_a, norm_b)
                assert torch.equal(opt_a.flat, opt_b.flat), (n, step, "params")
                assert torch.equal(opt_a.exp_avg, opt_b.exp_avg) and torch.equal(opt_a.exp_avg_sq, opt_b.exp_avg_sq)
                assert torch.equal(bucket.reduced, ref.grads), (n, step, "reduced clipped gradient")
            assert torch.equal(bucket.extra_reduced, want[n:]), (n, step, "scalar tail")
            everyone = [torch.empty_like(opt_a.flat) for _ in range(world)]
            dist.all_gather(everyone, opt_a.flat)
            assert all(torch.equal(everyone[0], e) for e in everyone), "ranks diverged"
        # timing: K8 vs NCCL allreduce + K6, 50 steps each, CUDA events, max over ranks
        def timed(fn, iters=50):
            for _ in range(5):
                fn()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(iters):
                fn()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b) / iters * 1e3], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()

        def nccl_route():
            dist.all_reduce(ref.flat)
            opt_b.step_flat(ref.grads, 0.5)

        if n in check_only:
            dist.barrier()
            bucket.close()
            continue
        results["%s n=%d" % (route, n)] = {"k8_us": round(timed(lambda: opt_a.step_peer(bucket, 0.5)), 2), "nccl_plus_k6_us": round(timed(nccl_route), 2)}
        dist.barrier()
        bucket.close()
    if rank == 0:
        print(json.dumps({"peer_adam_check": "ok", "world": world, "us_per_step": results}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
