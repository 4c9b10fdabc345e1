"""Worker of tests/test_multi_gpu.py: one process per GPU (torchrun, NCCL).  Every rank routes its own
seeded read batch through the chosen transport and checks its shard of the table against the oracle run
over ALL ranks' batches (filtered by ownership)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import kreeq_b200 as kb                                   # noqa: E402
import kreeq_oracle as ko                                 # noqa: E402
from helpers import random_genome, sample_reads           # noqa: E402
from kreeq_b200 import shard                              # noqa: E402


def main():
    transport = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    k = 21
    genome = random_genome(np.random.default_rng(11), 400_000)
    # ranks bring different amounts of reads: the round counts differ and the short ranks join empty rounds
    batches = [sample_reads(np.random.default_rng(100 + r), genome, 6000 + 3000 * r, 150, err=0.01) +
               (genome[1000 * r:1000 * r + 400].tobytes() + b"N") * 200 for r in range(world)]      # + a high-copy stretch each
    ref = ko.Table(k)
    for b in batches:
        ref.count_batch(b)
    rk, rr = ref.export()
    mine = shard.owner_of(rk, world) == rank
    g = kb.DBG(k=k, device=local, rank=rank, nranks=world)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    g.set_stream(stream.cuda_stream)
    round_pos = 500_000
    if transport == "peer":
        router = shard.PeerRoutedCounter(g, rank, world, round_pos, dev, slack=1.6, stream=stream)
    else:
        router = shard.RoutedCounter(g, rank, world, round_pos, dev, slack=1.6)
    bases, inv = kb.pack_host(batches[rank])
    d_b = torch.from_numpy(bases.view(np.int64)).to(dev)
    d_i = torch.from_numpy(inv.view(np.int64)).to(dev)
    if transport == "peer":
        router.count(d_b, d_i, len(batches[rank]))
    else:                                                  # the NCCL transport needs equal round counts
        rounds = max((len(b) - k + 1 + round_pos - 1) // round_pos for b in batches)
        router.count(d_b, d_i, len(batches[rank]), rounds=rounds)
    g.finalize()
    kk, rec = g.export()
    assert np.array_equal(kk, rk[mine]), "rank %d: keys differ" % rank
    assert rec.tobytes() == rr[mine].tobytes(), "rank %d: records differ" % rank
    t = torch.tensor([g.instances()], dtype=torch.int64, device=dev)
    dist.all_reduce(t)
    assert int(t.item()) == int(ref.stats()["total"])
    # validate: every rank takes its own slice of the assembly and probes the owners' shards over peer memory
    span = len(genome) // world
    seg = genome[rank * span:(rank + 1) * span].tobytes()
    _, rout, racc = ref.validate_segment(seg, cutoff=2)
    with shard.PeerTables(g, world, dev):
        out = g.evaluateSegment(seg, cov_cutoff=2)
    assert out.tobytes() == rout.tobytes(), "rank %d: per-base validate output differs" % rank
    assert (g.tot_missing, g.tot_kcount, g.tot_edge_missing) == racc
    # the handle is usable for counting again after the block
    g.traverseInReads(batches[rank][:1510]); g.finalize()
    g.close()
    dist.barrier()
    if rank == 0:
        print("mp_route_worker ok: transport=%s world=%d distinct=%d" % (transport, world, len(rk)))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
