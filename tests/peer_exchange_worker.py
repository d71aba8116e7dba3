"""torchrun worker of tests/test_multigpu.py::test_peer_exchange_matches_nccl: every rank fills its
half of the symmetric buffer with seeded counts, the peer-memory sum kernel (gfe_peer_sum_u32) must
equal the NCCL all-reduce of the same records, over many epochs (both halves, skewed rank timing)."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gfalign_b200.sharded import PeerExchange   # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n_cap = 5000
    px = PeerExchange(n_cap, dev)
    out = torch.zeros(n_cap, dtype=torch.int32, device=dev)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    bad = 0
    for epoch in range(1, 41):
        n = [1, 2, 63, 2048, 2049, 5000][epoch % 6]
        rec = torch.randint(0, 1 << 20, (n,), generator=g, dtype=torch.int32).to(dev)
        px.half_tensor(n).copy_(rec)
        if epoch % 5 == rank % 5:
            time.sleep(0.01)                 # ranks arrive at the exchange at different times
        px.sum_into(out.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
        want = rec.clone()
        dist.all_reduce(want)
        torch.cuda.synchronize()
        if not torch.equal(out[:n], want):
            bad += 1
    t = torch.tensor([bad], device=dev)
    dist.all_reduce(t)
    if rank == 0:
        sys.stderr.write(f"peer exchange: world {world}, mismatching epochs {int(t.item())}\n")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()
