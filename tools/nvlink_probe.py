"""torchrun --nproc-per-node 2 tools/nvlink_probe.py: raw peer-write rates over NVLink through torch symmetric memory
(copy engine copy_, an elementwise kernel's stores), the yardsticks for the in-kernel peer stores of the depth-slab mode."""
import json
import os

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    n = 141557760 // 4                                  # bytes a rank sends to its peer at cfg4, V = 10, 2 GPUs
    buf = symm.empty((n,), dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(buf, dist.group.WORLD)
    peer = hdl.get_buffer((rank + 1) % world, (n,), torch.float32)
    local = torch.randn(n, device=dev)
    out = {"rank": rank, "world": world, "mbytes": n * 4 / 1e6}

    def timed(fn, steps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    ms = timed(lambda: peer.copy_(local))
    out["copy_ms"], out["copy_gbs"] = ms, n * 4 / ms / 1e6
    ms = timed(lambda: torch.add(local, 1.0, out=peer))          # elementwise kernel: loads local, stores to the peer
    out["kernel_store_ms"], out["kernel_store_gbs"] = ms, n * 4 / ms / 1e6
    chunks = 20
    c = n // chunks

    def chunked():
        for i in range(chunks):
            peer[i * c:(i + 1) * c].copy_(local[i * c:(i + 1) * c], non_blocking=True)
    ms = timed(chunked)
    out["copy_20_chunks_ms"] = ms
    ms = timed(lambda: local.add_(1.0))
    out["local_kernel_ms"] = ms
    print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
