"""The data-parallel training block of bench.py alone (config 4 through model.train_on_batch, bucketed NCCL exchange),
for A/B runs of NCCL settings under torchrun: prints one JSON line with ms per step, the step without exchange and the
exposed all-reduce. Usage: torchrun --nproc-per-node N tools/bench_dp.py [--steps 20] [--tag text]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import bench
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    args = argparse.Namespace(train_batch=a.batch, train_steps=a.steps)
    out = bench.train_block(args, world, rank, dev, barrier)
    if rank == 0:
        keep = {k: out[k] for k in ("value", "ms_per_step", "split_ms", "exchange")}
        keep["e2e_ms_per_step"] = out["e2e"]["ms_per_step"]
        keep["dp_check_ok"] = (out.get("dp_check") or {}).get("ok")
        keep["tag"] = a.tag
        keep["env"] = {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}
        print(json.dumps(keep))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
