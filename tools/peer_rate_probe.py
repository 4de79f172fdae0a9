"""torchrun --nproc-per-node 2 tools/peer_rate_probe.py: peer-write rate over NVLink against the number of SMs that write
(st.global.v4 from the SMs' threads, or cp.async.bulk through shared memory).  Yardstick for the depth-slab "push" transport.
Build first (here, no GPU needed):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -shared -Xcompiler -fPIC -o tools/_peer_rate_probe.so tools/peer_rate_probe.cu"""
import ctypes
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
    lib = ctypes.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_peer_rate_probe.so"))
    lib.probe_launch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    n = 32 * 1024 * 1024                                 # 128 MB
    buf = symm.empty((n,), dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(buf, dist.group.WORLD)
    peer = hdl.get_buffer((rank + 1) % world, (n,), torch.float32)
    local = torch.randn(n, device=dev)
    own = torch.empty(n, device=dev)

    def timed(dst, ctas, mode, chunk, steps=5):
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):
            rc = lib.probe_launch(dst.data_ptr(), local.data_ptr(), n * 4, ctas, mode, chunk, st)
            assert rc == 0, rc
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            lib.probe_launch(dst.data_ptr(), local.data_ptr(), n * 4, ctas, mode, chunk, st)
        e1.record()
        torch.cuda.synchronize()
        return n * 4 / (e0.elapsed_time(e1) / steps) / 1e6

    for ctas in (1, 4, 8, 16, 32, 64, 148):
        rec = {"rank": rank, "ctas": ctas, "st_v4_peer_gbs": timed(peer, ctas, 0, 0), "st_v4_local_gbs": timed(own, ctas, 0, 0)}
        for chunk in (128, 512, 4096):
            rec[f"bulk{chunk}_peer_gbs"] = timed(peer, ctas, 1, chunk)
        rec["bulk4096_local_gbs"] = timed(own, ctas, 1, 4096)
        if rank == 0:
            print(json.dumps(rec), flush=True)
    dist.barrier()
    ok = torch.equal(hdl.get_buffer(rank, (n,), torch.float32), buf)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
