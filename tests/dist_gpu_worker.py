"""Worker of tests/test_gpu_exchange.py::test_two_gpu_step_matches_whole_batch (launched by torch.distributed.run, one
process per GPU): the sharded SpectralStep -- eager, then CUDA-graph replays on fresh data -- must reproduce the
whole-batch step of one GPU (bench.check_parity_n: loss, selection, projected latents, loss gradient)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200"))

import torch
import torch.distributed as dist

import bench
from tac_spectral.step import SpectralStep


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    w = dict(batch=64, views=2, T=16000, d=128, dim_p=24)
    rows = w["batch"] * w["views"]
    step = SpectralStep(rows, w["T"], w["views"], w["batch"], w["d"], w["dim_p"], bench.N_FFTS, bench.HOPS, device=dev,
                        group=dist.group.WORLD)
    sets = [bench.synth_host(w["batch"], w["views"], w["T"], w["d"], seed=100 * rank + i) for i in range(3)]
    bufs = (torch.empty(rows, w["T"], device=dev), torch.empty(rows, w["T"], device=dev),
            [torch.empty(w["batch"], w["d"], device=dev) for _ in range(w["views"])])

    def load(i):
        hx, hy, hz = sets[i]
        bufs[0].copy_(hx); bufs[1].copy_(hy)
        for a, b in zip(bufs[2], hz):
            a.copy_(b)

    verdicts = []
    load(0)
    verdicts.append(bench.check_parity_n(step, bufs, w, dev, rank, world))            # eager
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step.run(*bufs)
    for i in (1, 2, 0, 1):                                                               # replays on fresh data, both parities
        load(i)
        g.replay()
        verdicts.append(bench.check_parity_n(step, bufs, w, dev, rank, world, rerun=False))
    if rank == 0:
        for v in verdicts:
            print(v, flush=True)
        if all(v and v["verdict"] == "ok" for v in verdicts):
            print(f"PARITY_N_OK exchange={step.exchange} backend={step.xchg.backend if step.xchg else None}", flush=True)
    g = None
    torch.cuda.synchronize()
    dist.barrier(device_ids=[local])
    step.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
